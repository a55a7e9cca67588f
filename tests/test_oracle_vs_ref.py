"""Pins the oracle (oracle/me_oracle.c) against the reference's own code (oracle/_ref, built from
/root/reference by oracle/Makefile). Skipped where the reference sources are absent (GPU box) —
there the committed fixtures of tests/golden/ take over (test_oracle_golden.py)."""
import ctypes as C

import numpy as np
import pytest

from oracle import pyoracle as po
import svtme_b200  # noqa: F401
from svt_av1_psy_b200 import abi
from svt_av1_psy_b200.synth import PanningNoise

pytestmark = pytest.mark.skipif(not po.ref_available() and not __import__("os").path.isdir(po.REFERENCE_ROOT),
                                reason="reference sources not present")

from util import CASES, EXTRA, RES_FIELDS, SMALL, case_id, case_inputs, compare, extra_inputs, orc_me  # noqa: E402


def run_both(src, refs, pocs, poc, enc_mode, **kw):
    pin = po.make_pic_in(src, refs, pocs, poc, enc_mode, **kw)
    params = po.ref_derive_params(pin)
    ref_res, ref_dbg, _ = po.ref_me_picture(pin, params, debug=True)
    orc_res, orc_dbg, ops = orc_me(src, refs, params)
    return params, ref_res, ref_dbg, orc_res, orc_dbg


@pytest.mark.parametrize("case", CASES, ids=case_id)
def test_picture_me_matches_reference(orc, case):
    src, refs, pocs, poc, enc_mode, kw = case_inputs(case)
    _, ref_res, ref_dbg, orc_res, orc_dbg = run_both(src, refs, pocs, poc, enc_mode, **kw)
    compare(ref_res, orc_res, ref_dbg, orc_dbg)


@pytest.mark.parametrize("case", SMALL, ids=case_id)
def test_smallest_pictures_match_reference(orc, case):
    src, refs, pocs, poc, enc_mode, kw = case_inputs(case)
    _, ref_res, ref_dbg, orc_res, orc_dbg = run_both(src, refs, pocs, poc, enc_mode, **kw)
    compare(ref_res, orc_res, ref_dbg, orc_dbg)


@pytest.mark.parametrize("name,geo,kw", EXTRA, ids=[e[0] for e in EXTRA])
def test_picture_me_variants(orc, name, geo, kw):
    src, refs, pocs, poc, enc_mode, kw = extra_inputs(geo, kw)
    _, ref_res, ref_dbg, orc_res, orc_dbg = run_both(src, refs, pocs, poc, enc_mode, **kw)
    compare(ref_res, orc_res, ref_dbg, orc_dbg)


@pytest.mark.parametrize("enc_mode", [4, 8, -2])
def test_uncorrelated_pictures(orc, enc_mode):
    """references unrelated to the source: exercises ref pruning, large SADs, ties at window edges"""
    rng = np.random.default_rng(5)
    w, h = 256, 192
    src = rng.integers(0, 256, (h, w), dtype=np.uint8)
    flat = np.full((h, w), 77, np.uint8)  # every SAD ties: first raster position must win
    refs = [[rng.integers(0, 256, (h, w), dtype=np.uint8), flat, src.copy()], [flat, np.roll(src, (3, -5), (0, 1))]]
    _, ref_res, ref_dbg, orc_res, orc_dbg = run_both(src, refs, [[2, 0, 1], [8, 6]], 4, enc_mode, temporal_layer=2)
    compare(ref_res, orc_res, ref_dbg, orc_dbg)


@pytest.mark.parametrize("simd", [1, 2])
def test_reference_simd_tables_equal_c(orc, simd):
    """the reference's AVX2 / AVX-512 ME kernels give the same stage results as its C kernels"""
    w, h = 320, 192
    clip = PanningNoise(w, h, seed=3, vx=3, vy=2, max_frames=40)
    l0, l1 = [0, 8, 4, 2], [32, 24, 20]
    src = clip.frame(16)
    refs = [[clip.frame(p) for p in l0], [clip.frame(p) for p in l1]]
    lib = po.ref_lib()
    a, _, _ = po.ref_me_picture(po.make_pic_in(src, refs, [l0, l1], 16, 4, simd=0))
    assert lib.ref_rtcd_table_at_last_b64() == 0
    b, _, _ = po.ref_me_picture(po.make_pic_in(src, refs, [l0, l1], 16, 4, simd=simd))
    # the table the picture's b64 calls really ran with (round 1 compared C with C here: the plane builder reset it)
    assert lib.ref_rtcd_table_at_last_b64() == simd
    compare(a, b)


def test_reference_simd_path_is_the_fast_one(orc):
    """The CPU baseline bench.py quotes is the reference's x86 production table (aom_dsp_rtcd.c:499,507-514); a
    picture-level run with simd=2 must execute it (probe) and be an order of magnitude faster than the C kernels."""
    if "avx512bw" not in open("/proc/cpuinfo").read():
        pytest.skip("host without AVX-512")
    w, h = 1920, 1080
    clip = PanningNoise(w, h, seed=7, vx=2, vy=1, max_frames=40)
    l0, l1 = [0, 8, 4, 2], [32, 24, 20]
    src = clip.frame(16)
    refs = [[clip.frame(p) for p in l0], [clip.frame(p) for p in l1]]
    lib = po.ref_lib()
    b, _, t_simd = po.ref_me_picture(po.make_pic_in(src, refs, [l0, l1], 16, 4, simd=2), debug=False)
    assert lib.ref_rtcd_table_at_last_b64() == 2
    a, _, t_c = po.ref_me_picture(po.make_pic_in(src, refs, [l0, l1], 16, 4, simd=0), debug=False)
    assert lib.ref_rtcd_table_at_last_b64() == 0
    compare(a, b)
    assert t_c > 10 * t_simd, (t_c, t_simd)
