"""-m gpu: the CUDA picture pipeline (through the C ABI) against the oracle, bit-exact."""
import numpy as np
import pytest

import svtme_b200 as me
from oracle import pyoracle as po
from util import (CASES, EXTRA, SMALL, case_id, case_inputs, compare, compare_hme_state, extra_inputs, gpu_me, orc_me,
                  params_from_encoder_inputs)

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module", autouse=True)
def _built(orc):
    me.lib()  # fails loudly if the CUDA library is missing


def planes_oracle(luma):
    h, w = luma.shape
    full = np.zeros((h + 144, w + 144), np.uint8)
    quarter = np.zeros((h // 2 + 64, w // 2 + 64), np.uint8)
    sixteenth = np.zeros((h // 4 + 32, w // 4 + 32), np.uint8)
    po.oracle_lib().orc_build_planes(po.ptr(np.ascontiguousarray(luma)), w, h, po.ptr(full), po.ptr(quarter), po.ptr(sixteenth))
    return full, quarter, sixteenth


@pytest.mark.parametrize("w,h", [(320, 192), (200, 72), (1920, 1080), (88, 72)])
def test_pyramid_matches_oracle(w, h):
    rng = np.random.default_rng(w * 7 + h)
    luma = rng.integers(0, 256, (h, w), dtype=np.uint8)
    s = me.Session(w, h, max_pictures=2)
    s.upload(1, luma)
    for lvl, want in enumerate(planes_oracle(luma)):
        got = s.download_plane(1, lvl)
        np.testing.assert_array_equal(got, want, err_msg="level %d" % lvl)
    s.close()


def test_pyramid_10bit_upload():
    rng = np.random.default_rng(3)
    w, h = 320, 192
    luma10 = rng.integers(0, 1024, (h, w), dtype=np.uint16)
    s = me.Session(w, h, max_pictures=2)
    s.upload(0, luma10)
    want = planes_oracle((luma10 >> 2).astype(np.uint8))
    for lvl in range(3):
        np.testing.assert_array_equal(s.download_plane(0, lvl), want[lvl])
    s.close()


@pytest.mark.parametrize("case", CASES, ids=case_id)
def test_picture_me_matches_oracle(case):
    src, refs, pocs, poc, enc_mode, kw = case_inputs(case)
    params = params_from_encoder_inputs(src, refs, pocs, poc, enc_mode, **kw)
    want, dbg, ops = orc_me(src, refs, params)
    got, st = gpu_me(src, refs, params)
    compare_hme_state(st, dbg, params)
    compare(want, got, fields=list(got.FIELDS))
    assert got.sad_ops == ops  # the device's own count of the algorithmic work equals the oracle's


@pytest.mark.parametrize("case", SMALL, ids=case_id)
def test_smallest_pictures_match_oracle(case):
    src, refs, pocs, poc, enc_mode, kw = case_inputs(case)
    params = params_from_encoder_inputs(src, refs, pocs, poc, enc_mode, **kw)
    want, dbg, ops = orc_me(src, refs, params)
    got, st = gpu_me(src, refs, params)
    compare_hme_state(st, dbg, params)
    compare(want, got, fields=list(got.FIELDS))
    assert got.sad_ops == ops


@pytest.mark.parametrize("name,geo,kw", EXTRA, ids=[e[0] for e in EXTRA])
def test_picture_me_variants(name, geo, kw):
    src, refs, pocs, poc, enc_mode, kw = extra_inputs(geo, kw)
    params = params_from_encoder_inputs(src, refs, pocs, poc, enc_mode, **kw)
    want, dbg, _ = orc_me(src, refs, params)
    got, st = gpu_me(src, refs, params)
    compare_hme_state(st, dbg, params)
    compare(want, got, fields=list(got.FIELDS))


@pytest.mark.parametrize("enc_mode", [4, 8, -2])
def test_uncorrelated_and_flat_references(enc_mode):
    rng = np.random.default_rng(5)
    w, h = 256, 192
    src = rng.integers(0, 256, (h, w), dtype=np.uint8)
    flat = np.full((h, w), 77, np.uint8)  # every SAD ties: the first raster position must win
    refs = [[rng.integers(0, 256, (h, w), dtype=np.uint8), flat, src.copy()], [flat, np.roll(src, (3, -5), (0, 1))]]
    params = params_from_encoder_inputs(src, refs, [[2, 0, 1], [8, 6]], 4, enc_mode, temporal_layer=2)
    want, dbg, _ = orc_me(src, refs, params)
    got, st = gpu_me(src, refs, params)
    compare_hme_state(st, dbg, params)
    compare(want, got, fields=list(got.FIELDS))


def test_session_rejects_bad_arguments():
    s = me.Session(320, 192, max_pictures=2)
    src = np.zeros((192, 320), np.uint8)
    params = params_from_encoder_inputs(src, [[src], []], [[0], []], 1, 8)
    with pytest.raises(me.binding.SvtMeError):
        s.me(0, params)  # nothing uploaded
    s.upload(0, src)
    params.ref_pic_id[0][0] = 1
    with pytest.raises(me.binding.SvtMeError):
        s.me(0, params)  # reference picture 1 not uploaded
    with pytest.raises(me.binding.SvtMeError):
        me.Session(321, 192)  # width not a multiple of 8
    s.close()


def test_batch_equals_single_picture_calls():
    """a mini-GOP submitted as one batch gives, picture by picture, what single-picture calls give"""
    import copy
    from svt_av1_psy_b200.synth import PanningNoise
    w, h = 320, 192
    clip = PanningNoise(w, h, seed=21, vx=2, vy=1, max_frames=40)
    s = me.Session(w, h, max_pictures=20)
    for t in range(17):
        s.upload(t, clip.frame(t))
    plans = [(8, 1, [0, 4], [16, 12]), (4, 2, [0], [8, 16]), (2, 3, [0], [4, 8]), (16, 0, [0, 8], [0, 8]), (5, 4, [4, 0], [6, 8])]
    ids, prms, singles = [], [], []
    for poc, tl, l0, l1 in plans:
        src = clip.frame(poc)
        refs = [[clip.frame(p) for p in l0], [clip.frame(p) for p in l1]]
        p = params_from_encoder_inputs(src, refs, [l0, l1], poc, 4, temporal_layer=tl, hier_levels=5, is_ref=1 if tl < 5 else 0)
        for l, lst in enumerate((l0, l1)):
            for r, rp in enumerate(lst):
                p.ref_pic_id[l][r] = rp
        ids.append(poc)
        prms.append(p)
        singles.append(s.me(poc, copy.copy(p)))
    from svt_av1_psy_b200 import abi
    batch = [abi.PicResultArrays(w, h, p) for p in prms]
    s.me_batch(ids, prms, batch)
    for a, b in zip(singles, batch):
        compare(a, b, fields=list(a.FIELDS))
    s.close()


def test_pipelined_batches_equal_synchronous_ones():
    """batches queued back to back (sync=0, uploads of later pictures in between, two sets of host arrays) give what
    the same batches give one at a time; a slot re-uploaded while an earlier batch may still read it is ordered"""
    from svt_av1_psy_b200 import abi
    from svt_av1_psy_b200.synth import PanningNoise
    w, h = 320, 192
    clip = PanningNoise(w, h, seed=33, vx=3, vy=1, max_frames=64)
    n_slots = 24

    def plan(poc):
        return [p for p in (poc - 1, poc - 2) if p >= 0], [p for p in (poc + 1,) if p < 40]

    def run(pipelined):
        s = me.Session(w, h, max_pictures=n_slots)
        out, sets, seqs = [], [None, None], [0, 0]
        for t in range(10):
            s.upload(t % n_slots, clip.frame(t))
        for k in range(6):  # batch k: pictures 4k+1 .. 4k+4; picture 4k+5 (list 1 of the last one) is uploaded first
            for t in range(10 + 4 * k, 10 + 4 * (k + 1)):
                s.upload(t % n_slots, clip.frame(t))  # wraps onto slots that earlier batches read
            ids, prms = [], []
            for poc in range(4 * k + 1, 4 * k + 5):
                l0, l1 = plan(poc)
                refs = [[clip.frame(p) for p in l0], [clip.frame(p) for p in l1]]
                p = params_from_encoder_inputs(clip.frame(poc), refs, [l0, l1], poc, 6, temporal_layer=2, hier_levels=5)
                for l, lst in enumerate((l0, l1)):
                    for r, rp in enumerate(lst):
                        p.ref_pic_id[l][r] = rp % n_slots
                ids.append(poc % n_slots)
                prms.append(p)
            if not pipelined:
                out.append(s.me_batch(ids, prms, [abi.PicResultArrays(w, h, p) for p in prms], sync=True))
                continue
            j = k & 1
            if sets[j] is not None:
                s.wait_batch(seqs[j])
                out.append([r.snapshot() for r in sets[j]])
            sets[j] = [abi.PicResultArrays(w, h, p, pinned=True) for p in prms]
            s.me_batch(ids, prms, sets[j], sync=False)
            seqs[j] = s.batch_seq()
        if pipelined:
            s.sync()
            tail = sorted(((seqs[j], [r.snapshot() for r in sets[j]]) for j in range(2)), key=lambda t: t[0])
            out += [t[1] for t in tail]
        s.close()
        return out

    a, b = run(False), run(True)
    assert len(a) == len(b) == 6
    for ba, bb in zip(a, b):
        for ra, rb in zip(ba, bb):
            compare(ra, rb, fields=list(ra.FIELDS))


def test_batched_uploads_build_the_same_pyramids():
    """svtme_picture_upload_batch / _10bit_batch / _device_batch: two launches for all pictures, same planes"""
    import torch
    rng = np.random.default_rng(12)
    w, h, n = 200, 72, 5
    frames = [rng.integers(0, 256, (h, w), dtype=np.uint8) for _ in range(n)]
    frames10 = [(f.astype(np.uint16) << 2) | rng.integers(0, 4, f.shape, dtype=np.uint16) for f in frames]
    s = me.Session(w, h, max_pictures=3 * n)
    s.upload_batch(list(range(n)), [f.ctypes.data for f in frames], w, "host8")
    s.upload_batch(list(range(n, 2 * n)), [f.ctypes.data for f in frames10], w, "host10")
    dev = [torch.from_numpy(f).cuda() for f in frames]
    torch.cuda.synchronize()
    s.upload_batch(list(range(2 * n, 3 * n)), [d.data_ptr() for d in dev], w, "device")
    for i, f in enumerate(frames):
        want = planes_oracle(f)
        for lvl in range(3):
            for slot in (i, n + i, 2 * n + i):
                np.testing.assert_array_equal(s.download_plane(slot, lvl), want[lvl], err_msg="picture %d slot %d level %d" % (i, slot, lvl))
    s.close()


def test_two_compute_streams_give_the_same_results():
    """svtme_session_set_concurrency(2): consecutive batches on alternating compute streams, uploads in between (host and
    device), slots re-used while the other stream may still read them"""
    import torch
    from svt_av1_psy_b200 import abi
    from svt_av1_psy_b200.synth import PanningNoise
    w, h, n_slots = 320, 192, 20
    clip = PanningNoise(w, h, seed=77, vx=2, vy=2, max_frames=64)

    def run(concurrency):
        s = me.Session(w, h, max_pictures=n_slots)
        s.set_concurrency(concurrency)
        outs, pend = [], []
        for t in range(8):
            s.upload(t % n_slots, clip.frame(t))
        dev = {}
        for k in range(7):
            new = list(range(8 + 4 * k, 12 + 4 * k))
            if k % 2:  # device uploads every other round
                for t in new:
                    dev[t] = torch.from_numpy(clip.frame(t)).cuda()
                torch.cuda.synchronize()
                s.upload_batch([t % n_slots for t in new], [dev[t].data_ptr() for t in new], w, "device")
            else:
                frames = [clip.frame(t) for t in new]
                s.upload_batch([t % n_slots for t in new], [f.ctypes.data for f in frames], w, "host8")
            ids, prms = [], []
            for poc in range(4 * k + 2, 4 * k + 6):
                l0, l1 = [poc - 1, poc - 2], [poc + 1]
                refs = [[clip.frame(p) for p in l0], [clip.frame(p) for p in l1]]
                p = params_from_encoder_inputs(clip.frame(poc), refs, [l0, l1], poc, 5, temporal_layer=2, hier_levels=5)
                for l, lst in enumerate((l0, l1)):
                    for r, rp in enumerate(lst):
                        p.ref_pic_id[l][r] = rp % n_slots
                ids.append(poc % n_slots)
                prms.append(p)
            res = [abi.PicResultArrays(w, h, p, pinned=True) for p in prms]
            s.me_batch(ids, prms, res, sync=False)
            pend.append((s.batch_seq(), res))
            if len(pend) > 2:
                seq, r = pend.pop(0)
                s.wait_batch(seq)
                outs.append([x.snapshot() for x in r])
        s.sync()
        outs += [[x.snapshot() for x in r] for _, r in pend]
        s.close()
        return outs

    a, b = run(1), run(2)
    assert len(a) == len(b) == 7
    for ba, bb in zip(a, b):
        for ra, rb in zip(ba, bb):
            compare(ra, rb, fields=list(ra.FIELDS))


@pytest.mark.parametrize("case", [CASES[0], CASES[2], CASES[3], CASES[4], CASES[8]], ids=case_id)
@pytest.mark.parametrize("mode", [0, 1])
def test_unit_granularity_does_not_change_results(case, mode):
    """k_search16 walks (b64, list, ref) groups in large batches and single jobs in small ones: both against the oracle"""
    src, refs, pocs, poc, enc_mode, kw = case_inputs(case)
    params = params_from_encoder_inputs(src, refs, pocs, poc, enc_mode, **kw)
    want, dbg, ops = orc_me(src, refs, params)
    h, w = src.shape
    s = me.Session(w, h, max_pictures=12)
    s.set_unit_granularity(mode)
    got, st = gpu_me(src, refs, params, session=s)
    s.close()
    compare_hme_state(st, dbg, params)
    compare(want, got, fields=list(got.FIELDS))
    assert got.sad_ops == ops
