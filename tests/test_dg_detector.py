"""Dynamic-GOP detector HME (dg_detector_hme_level0, pd_process.c:392-589).
  * CPU: the oracle's restatement against the reference's own function (oracle/_ref) and against the committed
    metrics (tests/golden/dg_detector.json) where the reference is absent;
  * GPU: svtme_dg_detector_hme against the oracle, metrics and per-b64 results."""
import json
import os

import numpy as np
import pytest

from oracle import pyoracle as po
import svtme_b200 as me
from svt_av1_psy_b200.synth import PanningNoise
import golden_lib as gl

GOLDEN = os.path.join(gl.GOLDEN_DIR, "dg_detector.json")
# (w, h, vx, vy, frame distance): the three window classes (16 / 64 / 128), incomplete b64s, zoom-like and static pairs
CASES = [(320, 192, 2, 1, 4), (640, 360, 3, 2, 6), (864, 480, 5, 1, 3), (1280, 720, 4, 3, 5), (1920, 1080, 2, 1, 8), (200, 72, 1, 1, 2),
         (1280, 720, 0, 0, 1)]


def pair(case):
    w, h, vx, vy, d = case
    clip = PanningNoise(w, h, seed=w + h, vx=max(vx, 0), vy=max(vy, 0), max_frames=16, sigma=3.0 if vx or vy else 0.0)
    return clip.frame(d), clip.frame(0)


def key(case):
    return "x".join(str(v) for v in case)


@pytest.mark.skipif(not os.path.isdir(po.REFERENCE_ROOT), reason="reference sources not present")
@pytest.mark.parametrize("case", CASES, ids=key)
def test_oracle_matches_reference(orc, case):
    src, ref = pair(case)
    want, _, _ = po.dg_detector(po.ref_lib(), "ref_", src, ref)
    got, _, _ = po.dg_detector(po.oracle_lib(), "orc_", src, ref)
    assert got.as_tuple() == want.as_tuple()


@pytest.mark.parametrize("case", CASES, ids=key)
def test_oracle_matches_golden(orc, case):
    if not os.path.exists(GOLDEN):
        pytest.skip("tests/golden/dg_detector.json not generated")
    src, ref = pair(case)
    got, _, _ = po.dg_detector(po.oracle_lib(), "orc_", src, ref)
    assert list(got.as_tuple()) == json.load(open(GOLDEN))[key(case)]


@pytest.mark.gpu
@pytest.mark.parametrize("case", CASES, ids=key)
def test_cuda_matches_oracle(case):
    src, ref = pair(case)
    want, wsad, wmv = po.dg_detector(po.oracle_lib(), "orc_", src, ref)
    h, w = src.shape
    s = me.Session(w, h, max_pictures=2)
    s.upload(0, src)
    s.upload(1, ref)
    got, sad, mv = s.dg_detector(0, 1)
    np.testing.assert_array_equal(sad, wsad)
    np.testing.assert_array_equal(mv, wmv)
    assert got.as_tuple() == want.as_tuple()
    if os.path.exists(GOLDEN):
        assert list(got.as_tuple()) == json.load(open(GOLDEN))[key(case)]  # the reference's own numbers
    again, _, _ = s.dg_detector(0, 1, per_b64=False)
    assert again.as_tuple() == got.as_tuple()
    s.close()
