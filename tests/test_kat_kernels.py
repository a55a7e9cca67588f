"""RTCD-kernel known-answer tests.
  * CPU (-m "not gpu"): oracle kernels vs the reference's C and SIMD kernels (needs oracle/_ref), and vs
    the committed outputs of the reference's C kernels (tests/golden/kat_rtcd.npz) where it is absent.
  * GPU (-m gpu): the CUDA C-ABI entry points vs the oracle kernels."""
import os

import numpy as np
import pytest

import kat_cases as kc
from kat_runner import Impl, same
from oracle import pyoracle as po

import golden_lib as gl

HAVE_REF = os.path.isdir(po.REFERENCE_ROOT) or po.ref_available()
KAT_GOLDEN = os.path.join(gl.GOLDEN_DIR, "kat_rtcd.npz")
GOLDEN_CTX = {"peer": None}  # Recorder (make_golden.py) or Replayer (test_oracle_kernels_match_golden)


def impls(kind):
    """-> (candidate, [references...])"""
    orc = Impl(po.oracle_lib(), "orc_")
    if kind == "cpu":
        return orc, [Impl(po.ref_lib(), "ref_", simd) for simd in (0, 1, 2)]
    if kind == "golden":
        return orc, [GOLDEN_CTX["peer"]]
    import svtme_b200 as me
    return Impl(me.lib(), "svtme_"), [orc]


def run_sad_loop(kind):
    cand, refs = impls(kind)
    for i, (bw, bh, sw, sh, skip, pat) in enumerate(kc.sad_loop_cases()):
        doubled = (i % 3) == 0 and pat != "UNALIGN"  # the SUB_SAD_SEARCH calling convention (strides x2)
        src, ss, ref, rs, raw = kc.sad_loop_buffers(bw, bh, sw, sh, pat, 1000 + i, doubled)
        got = cand.sad_loop(src, ss, ref, rs, bh, bw, raw, skip, sw, sh)
        for r in refs:
            if r.simd and (rs != raw or bw % 4 or bw > 64):
                continue  # the SIMD kernels are only defined for the widths/strides the encoder uses
            same(got, r.sad_loop(src, ss, ref, rs, bh, bw, raw, skip, sw, sh), "sad_loop %s" % ((bw, bh, sw, sh, skip, pat),))
    # empty search areas: outputs other than best_sad stay untouched
    src, ss, ref, rs, raw = kc.sad_loop_buffers(16, 8, 8, 8, "RANDOM", 1)
    assert cand.sad_loop(src, ss, ref, rs, 8, 16, raw, 0, 0, 4) == (0xffffff, -7, 9)
    assert cand.sad_loop(src, ss, ref, rs, 8, 16, raw, 1, 8, 1) == (0xffffff, -7, 9)  # only an even row, skipped


def run_all_sad(kind):
    cand, refs = impls(kind)
    n = 0
    for pat in kc.PATTERNS:
        for sub in (0, 1):
            for init in ("max", "small"):
                src, ss, ref, rs = kc.me_block_buffers(pat, 77 + n)
                rng = np.random.default_rng(n)
                if init == "max":
                    b8, b16 = np.full(64, 0x7FFFFFFF, np.uint32), np.full(16, 0x7FFFFFFF, np.uint32)
                else:  # some running bests already lower than anything in this group: must be kept
                    b8 = rng.integers(0, 3000, 64).astype(np.uint32)
                    b16 = rng.integers(0, 12000, 16).astype(np.uint32)
                m8 = rng.integers(0, 1 << 32, 64, dtype=np.uint64).astype(np.uint32)
                m16 = rng.integers(0, 1 << 32, 16, dtype=np.uint64).astype(np.uint32)
                mv = int(((np.uint32(-5 & 0xFFFF) << 16) | np.uint32(0xFFFD)))  # y = -5, x = -3: x crosses zero
                got = cand.all_sad(src, ss, ref, rs, mv, b8, b16, m8, m16, sub)
                for r in refs:
                    same(got, r.all_sad(src, ss, ref, rs, mv, b8, b16, m8, m16, sub), "all_sad %s sub=%d %s" % (pat, sub, init))
                n += 1


def run_eight_sad(kind):
    cand, refs = impls(kind)
    for i, pat in enumerate(["BUF_MAX", "BUF_MIN", "BUF_SMALL", "BUF_RANDOM"]):
        for j, bpat in enumerate(["BUF_MAX", "BUF_SMALL"]):
            if pat in ("BUF_MAX", "BUF_RANDOM"):
                continue  # sums overflow uint32: implementation-defined lanes in the SIMD versions
            s16 = kc.sad_buffer(pat, (16, 8), 5 + i)
            b32, b64 = kc.sad_buffer(bpat, 4, 9 + j), kc.sad_buffer(bpat, 1, 11 + j)
            if bpat == "BUF_SMALL":
                b32, b64 = b32 * 3, b64 * 20
            else:  # the SSE4.1 kernel compares signed: keep the initial bests below 2^31 like the encoder does
                b32, b64 = b32 >> 1, b64 >> 1
            m32, m64 = kc.sad_buffer("BUF_RANDOM", 4, 13), kc.sad_buffer("BUF_RANDOM", 1, 14)
            mv = (7 << 16) | 0x7FFC  # x + index wraps in int16 exactly as in the reference
            got = cand.eight_sad(s16, b32, b64, m32, m64, mv)
            for r in refs:
                same(got, r.eight_sad(s16, b32, b64, m32, m64, mv), "eight_sad %s %s" % (pat, bpat))
            got = cand.ext_sad_32_64(s16[:, 0], b32, b64, m32, m64, mv)
            for r in refs:
                same(got, r.ext_sad_32_64(s16[:, 0], b32, b64, m32, m64, mv), "ext_sad_32_64 %s %s" % (pat, bpat))


def run_ext_sad(kind):
    cand, refs = impls(kind)
    for n, pat in enumerate(kc.PATTERNS):
        for sub in (0, 1):
            src, ss, ref, rs = kc.me_block_buffers(pat, 300 + n, ref_extra=1)
            b8, b16 = np.full(4, 0x7FFFFFFF, np.uint32), np.full(1, 2000 if pat == "RANDOM" else 0x7FFFFFFF, np.uint32)
            m8, m16 = np.arange(4, dtype=np.uint32), np.full(1, 99, np.uint32)
            got = cand.ext_sad(src, ss, ref[:, 3:], rs, b8, b16, m8, m16, 0x00030002, sub)
            for r in refs:
                same(got, r.ext_sad(src, ss, ref[:, 3:], rs, b8, b16, m8, m16, 0x00030002, sub), "ext_sad %s sub=%d" % (pat, sub))


def run_nxm_init_downsample(kind):
    cand, refs = impls(kind)
    rng = np.random.default_rng(8)
    for (bw, bh) in kc.BLOCK_SIZES + [(64, 28), (56, 32), (8, 4)]:
        src = rng.integers(0, 256, (bh, 131), dtype=np.uint8)
        ref = rng.integers(0, 256, (bh, 197), dtype=np.uint8)
        got = cand.nxm(src[:, 1:], 131, ref[:, 2:], 197, bh, bw)
        for r in refs:
            if r.simd and (bw % 8 or bw > 64):
                continue
            assert got == r.nxm(src[:, 1:], 131, ref[:, 2:], 197, bh, bw), (bw, bh)
    for c128, c32, v in ((21, 1, 128 * 128 * 255), (0, 3, 7), (4, 0, 0xFFFFFFFF)):
        got = cand.init32(c128, c32, v)
        for r in refs:
            same(got, r.init32(c128, c32, v), "init32")
    for (w, h) in kc.DOWNSAMPLE_SIZES:
        img = rng.integers(0, 256, (h, w), dtype=np.uint8)
        for step in (2, 4, 8):
            if w // step < 2:
                continue
            got = cand.downsample(img, step)
            for r in refs:
                if r.simd:
                    continue
                same(got, r.downsample(img, step), "downsample %dx%d /%d" % (w, h, step))


RUNNERS = [run_sad_loop, run_all_sad, run_eight_sad, run_ext_sad, run_nxm_init_downsample]


@pytest.mark.skipif(not HAVE_REF, reason="reference sources not present")
@pytest.mark.parametrize("runner", RUNNERS, ids=lambda f: f.__name__[4:])
def test_oracle_kernels_match_reference(orc, runner):
    runner("cpu")


@pytest.mark.parametrize("runner", RUNNERS, ids=lambda f: f.__name__[4:])
def test_oracle_kernels_match_golden(orc, runner):
    """same runners, the reference's recorded outputs instead of the reference (tests/golden/make_golden.py)"""
    if not os.path.exists(KAT_GOLDEN):
        pytest.skip("tests/golden/kat_rtcd.npz not generated")
    peer = GOLDEN_CTX["peer"] = gl.Replayer(np.load(KAT_GOLDEN), runner.__name__)
    runner("golden")
    assert peer.pos == len(peer.calls) and peer.pos > 0


@pytest.mark.gpu
@pytest.mark.parametrize("runner", RUNNERS, ids=lambda f: f.__name__[4:])
def test_cuda_kernels_match_oracle(orc, runner):
    runner("gpu")
