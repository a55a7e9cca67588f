"""BASELINE.json's other configurations as parity cases at their full sizes (-m gpu).
  configs[1]  1920x1080 8-bit preset 4              one picture against the oracle (SADs, MVs, candidates, work count)
  configs[2]  3840x2160 10-bit preset 4             one picture, 10-bit upload path, against the oracle
  configs[3]  7680x4320 8-bit preset 2              the oracle needs a minute per picture here: pure-translation and
                                                    determinism properties at full size + the 8K controls against the oracle
                                                    on a 2048x256 crop
  configs[4]  preset -2 at 1080p, 16 streams        widest windows: properties at 1080p, oracle at 640x360; 16 sessions
                                                    interleaved give what one session gives
All through the C ABI (svt-av1-psy_b200/binding.py)."""
import copy

import numpy as np
import pytest

import svtme_b200 as me
from svt_av1_psy_b200.synth import PanningNoise
from util import compare, compare_hme_state, gpu_me, orc_me, params_from_encoder_inputs

pytestmark = pytest.mark.gpu


def setup(w, h, preset, l0, l1, poc, tl=2, seed=1, sigma=2.0, vx=3, vy=2, **kw):
    clip = PanningNoise(w, h, seed=seed, vx=vx, vy=vy, max_frames=40, sigma=sigma)
    src = clip.frame(poc)
    refs = [[clip.frame(p) for p in l0], [clip.frame(p) for p in l1]]
    prm = params_from_encoder_inputs(src, refs, [l0, l1], poc, preset, temporal_layer=tl, hier_levels=5, **kw)
    return clip, src, refs, prm


def test_1080p_preset4_matches_oracle():
    _, src, refs, prm = setup(1920, 1080, 4, [4, 0], [16], 8)
    want, dbg, ops = orc_me(src, refs, prm)
    got, st = gpu_me(src, refs, prm)
    compare_hme_state(st, dbg, prm)
    compare(want, got, fields=list(got.FIELDS))
    assert got.sad_ops == ops


def test_4k_10bit_preset4_matches_oracle():
    w, h = 3840, 2160
    clip, src, refs, prm = setup(w, h, 4, [4], [16], 8)
    want, dbg, _ = orc_me(src, refs, prm)  # the ME works on the 8 MSBs (pack_unpack_c.c:77-116)
    s = me.Session(w, h, max_pictures=4)
    s.upload(0, clip.frame10(8))
    s.upload(1, clip.frame10(4))
    s.upload(2, clip.frame10(16))
    p = copy.copy(prm)
    p.ref_pic_id[0][0], p.ref_pic_id[1][0] = 1, 2
    got = s.me(0, p)
    compare_hme_state(s.hme_state(), dbg, prm)
    compare(want, got, fields=list(got.FIELDS))
    s.close()


def translation_properties(w, h, preset, vx, vy):
    """reference = source moved by a whole-sample vector inside the HME range: away from the picture edges every PU of
    every b64 must find SAD 0 at one and the same vector; two runs must agree bit for bit"""
    clip, src, refs, prm = setup(w, h, preset, [7], [9], 8, sigma=0.0, vx=vx, vy=vy)
    a, _ = gpu_me(src, refs, prm)
    b, _ = gpu_me(src, refs, prm)
    compare(a, b, fields=list(a.FIELDS))
    cols, rows = (w + 63) // 64, (h + 63) // 64
    interior = [r * cols + c for r in range(2, rows - 2) for c in range(2, cols - 2)]
    for l in (0, 1):
        sad = a.p_sb_best_sad[interior, l, 0, :]
        assert int(sad.max()) == 0, "list %d: a displaced copy must be found exactly" % l
        mv = a.p_sb_best_mv[interior, l, 0, :]
        assert len(np.unique(mv)) == 1, "one translation -> one vector for every PU"
    assert not np.array_equal(a.p_sb_best_mv[interior[0], 0, 0, 0], a.p_sb_best_mv[interior[0], 1, 0, 0])  # past vs future
    return a


def test_8k_preset2_properties():
    translation_properties(7680, 4320, 2, 5, 3)


def test_8k_controls_match_oracle_on_a_crop():
    _, _, _, prm8k = setup(7680, 4320, 2, [4, 0], [16], 8)
    w, h = 2048, 256
    clip = PanningNoise(w, h, seed=4, vx=6, vy=2, max_frames=40)
    src, refs = clip.frame(8), [[clip.frame(4), clip.frame(0)], [clip.frame(16)]]
    want, dbg, ops = orc_me(src, refs, prm8k)
    got, st = gpu_me(src, refs, prm8k)
    compare_hme_state(st, dbg, prm8k)
    compare(want, got, fields=list(got.FIELDS))
    assert got.sad_ops == ops


def test_preset_m2_1080p_properties():
    translation_properties(1920, 1080, -2, 4, 3)


def test_preset_m2_matches_oracle_640x360():
    _, src, refs, prm = setup(640, 360, -2, [4, 0, 6, 2], [16, 12, 10], 8)
    want, dbg, ops = orc_me(src, refs, prm)
    got, st = gpu_me(src, refs, prm)
    compare_hme_state(st, dbg, prm)
    compare(want, got, fields=list(got.FIELDS))
    assert got.sad_ops == ops


def test_16_concurrent_streams():
    """16 sessions (one per stream) used in an interleaved order give, stream by stream, what a session used alone gives"""
    w, h, n = 640, 368, 16
    clips = [PanningNoise(w, h, seed=50 + i, vx=1 + i % 4, vy=i % 3, max_frames=40) for i in range(n)]
    l0, l1 = [4, 0], [16]
    prms = []
    for c in clips:
        refs = [[c.frame(p) for p in l0], [c.frame(p) for p in l1]]
        p = params_from_encoder_inputs(c.frame(8), refs, [l0, l1], 8, 6, temporal_layer=2, hier_levels=5)
        p.ref_pic_id[0][0], p.ref_pic_id[0][1], p.ref_pic_id[1][0] = 1, 2, 3
        prms.append(p)
    sessions = [me.Session(w, h, max_pictures=4) for _ in range(n)]
    for slot, poc in enumerate((8, 4, 0, 16)):  # upload picture by picture, stream by stream
        for s, c in zip(sessions, clips):
            s.upload(slot, c.frame(poc))
    from svt_av1_psy_b200 import abi
    outs = [abi.PicResultArrays(w, h, p, pinned=True) for p in prms]
    for s, p, o in zip(sessions, prms, outs):
        s.me_batch([0], [p], [o], sync=False)  # all 16 in flight
    for s in sessions:
        s.sync()
    for i in (0, 5, 15):  # alone, on a fresh session
        c = clips[i]
        alone, _ = gpu_me(c.frame(8), [[c.frame(p) for p in l0], [c.frame(p) for p in l1]], prms[i])
        compare(alone, outs[i], fields=list(alone.FIELDS))
    want, _, _ = orc_me(clips[3].frame(8), [[clips[3].frame(p) for p in l0], [clips[3].frame(p) for p in l1]], prms[3])
    compare(want, outs[3], fields=list(outs[3].FIELDS))
    for s in sessions:
        s.close()
