"""Seeded random configurations, CUDA against the oracle: picture size, preset, prediction structure, screen-content
class, QP, temporal layer, number of references and the motion / noise of the content all vary together. The fixed
case lists of util.py pin known corners; this net catches interactions between them."""
import numpy as np
import pytest

import svtme_b200  # noqa: F401  (registers the svt_av1_psy_b200 package)
from svt_av1_psy_b200 import gop
from svt_av1_psy_b200.synth import PanningNoise
from util import compare, compare_hme_state, gpu_me, orc_me, params_from_encoder_inputs

pytestmark = pytest.mark.gpu
N_CASES = 48


def draw(i):
    rng = np.random.default_rng(9000 + i)
    w = int(rng.integers(8, 49)) * 8
    h = int(rng.integers(8, 33)) * 8
    enc_mode = int(rng.integers(-3, 14))
    if enc_mode <= 0:  # the widest windows: keep the oracle's run time in seconds
        w, h = min(w, 192), min(h, 128)
    rtc = enc_mode >= 7 and rng.random() < 0.3
    sc = rng.random() < 0.2
    mrp = gop.mrp_for_preset(enc_mode)
    n0 = int(rng.integers(1, mrp["l0"] + 1))
    n1 = int(rng.integers(0, mrp["l1"] + 1))
    if rtc:
        n0, n1 = min(n0, 2), min(n1, 2)
    hier = 2 if rtc else 5
    tl = int(rng.integers(0, hier + 1))
    poc = 16
    past = [poc - d for d in sorted(rng.choice(np.arange(1, 17), size=n0, replace=False))]
    future = [poc + d for d in sorted(rng.choice(np.arange(1, 17), size=n1, replace=False))]
    if tl == 0 or rtc:  # base layer / low delay: list 1 holds past pictures too
        future = [poc - d for d in sorted(rng.choice(np.arange(1, 17), size=n1, replace=False))]
    vx, vy = int(rng.integers(0, 13)), int(rng.integers(0, 7))
    clip = PanningNoise(w, h, seed=int(rng.integers(1, 1 << 30)), vx=vx, vy=vy, max_frames=40, sigma=float(rng.choice([0.0, 1.5, 4.0, 12.0])),
                        scale=int(rng.choice([2, 4, 8])))
    kw = dict(temporal_layer=tl, hier_levels=hier, is_ref=0 if tl == hier else 1, sc_class1=int(sc), pred_structure=1 if rtc else 2,
              qp=int(rng.integers(10, 64)))
    return clip, poc, past, future, enc_mode, kw


@pytest.mark.parametrize("i", range(N_CASES))
def test_random_configuration(i):
    clip, poc, l0, l1, enc_mode, kw = draw(i)
    src = clip.frame(poc)
    refs = [[clip.frame(p) for p in l0], [clip.frame(p) for p in l1]]
    params = params_from_encoder_inputs(src, refs, [l0, l1], poc, enc_mode, **kw)
    want, dbg, ops = orc_me(src, refs, params)
    got, st = gpu_me(src, refs, params)
    what = "case %d: %dx%d M%d %s l0=%s l1=%s" % (i, clip.w, clip.h, enc_mode, kw, l0, l1)
    try:
        compare_hme_state(st, dbg, params)
        compare(want, got, fields=list(got.FIELDS))
    except AssertionError as e:
        raise AssertionError(what + "\n" + str(e)[:1500])
    assert got.sad_ops == ops, what
