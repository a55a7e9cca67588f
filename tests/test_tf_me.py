"""Temporal-filter ME (ME_MCTF): svt_aom_motion_estimation_b64 as temporal_filtering.c:3095-3146
drives it — one list, one reference (the neighbouring picture), FULL_SAD HME, no pruning, no candidate lists, and the
HME early exit of motion_estimation.c:3179-3183.
  * CPU: the host controls (svtme_sig_deriv_me_tf) and the oracle against the reference's own code (oracle/_ref), and
    against the committed fixture (tests/golden/tf_me.npz) where the reference is absent;
  * GPU: the CUDA path against the oracle and the fixture."""
import os

import numpy as np
import pytest

from oracle import pyoracle as po
import svtme_b200 as me
from svt_av1_psy_b200 import binding
from svt_av1_psy_b200.synth import PanningNoise
import golden_lib as gl
from util import compare_hme_state, gpu_me, orc_me

GOLDEN = os.path.join(gl.GOLDEN_DIR, "tf_me.npz")
TF_FIELDS = ("p_sb_best_sad", "p_sb_best_mv", "search_results")
# (w, h, enc_mode, hme_me_level, qp_opt, me_exit_th, qp, centre poc, ref poc)
CASES = [
    (320, 192, 4, 0, 0, 0, 35, 8, 7),
    (320, 192, 4, 1, 1, 0, 35, 8, 10),
    (320, 192, 8, 2, 0, 0, 20, 8, 9),
    (256, 136, 8, 3, 1, 0, 55, 8, 6),        # SUB_SAD full-pel under FULL_SAD HME; incomplete last b64 row
    (320, 192, 10, 4, 0, 0, 35, 8, 7),
    (320, 192, 4, 1, 0, 7564, 35, 8, 7),     # the early exit takes some b64s and leaves others
    (864, 480, 6, 2, 1, 7385, 35, 8, 5),
    (200, 72, 9, 4, 0, 0, 35, 8, 7),
]


def key(c):
    return "_".join(str(v) for v in c)


def inputs(case):
    w, h, m, lvl, qo, th, qp, cp, rp = case
    clip = PanningNoise(w, h, seed=w + lvl, vx=3, vy=2, max_frames=20, sigma=2.5)
    prm = binding.sig_deriv_me_tf(m, w, h, lvl, qp_opt=qo, me_exit_th=th, qp=qp, centre_poc=cp, ref_poc=rp)
    return clip.frame(cp), clip.frame(rp), prm


def same(a, b, what):
    for f in TF_FIELDS:
        x, y = getattr(a, f) if not isinstance(a, dict) else a[f], getattr(b, f)
        if x.dtype.names:
            for n in x.dtype.names:
                np.testing.assert_array_equal(x[n], y[n], err_msg="%s %s.%s" % (what, f, n))
        else:
            np.testing.assert_array_equal(x, y, err_msg="%s %s" % (what, f))


@pytest.mark.skipif(not os.path.isdir(po.REFERENCE_ROOT), reason="reference sources not present")
@pytest.mark.parametrize("case", CASES, ids=key)
def test_oracle_and_host_controls_match_reference(orc, case):
    w, h, m, lvl, qo, th, qp, cp, rp = case
    centre, ref, prm = inputs(case)
    rprm, want, wdbg = po.ref_tf_me_picture(centre, ref, m, lvl, qp_opt=qo, me_exit_th=th, qp=qp, centre_poc=cp, ref_poc=rp)
    assert gl.struct_bytes(prm) == gl.struct_bytes(rprm), "svtme_sig_deriv_me_tf differs from svt_aom_sig_deriv_me_tf"
    got, gdbg, _ = orc_me(centre, [[ref]], prm)
    same(want, got, "oracle vs reference")
    for n in ("x_hme_l0", "y_hme_l0", "hme_l0_sad", "x_hme_l1", "y_hme_l1", "hme_l1_sad", "x_hme_l2", "y_hme_l2", "hme_l2_sad"):
        np.testing.assert_array_equal(wdbg[n][:, 0, 0], gdbg[n][:, 0, 0], err_msg=n)
    if th:
        exits = (want.search_results["hme_sad"][:, 0, 0] < th).sum()
        assert 0 < exits < want.n_b64, "the case is meant to mix exited and searched b64s (%d of %d)" % (exits, want.n_b64)


@pytest.mark.parametrize("case", CASES, ids=key)
def test_oracle_matches_golden(orc, case):
    if not os.path.exists(GOLDEN):
        pytest.skip("tests/golden/tf_me.npz not generated")
    z = np.load(GOLDEN)
    centre, ref, prm = inputs(case)
    assert gl.struct_bytes(prm) == z[key(case) + "/params"].tobytes()
    got, _, _ = orc_me(centre, [[ref]], prm)
    same({f: z[key(case) + "/" + f] for f in TF_FIELDS}, got, "oracle vs fixture")


@pytest.mark.gpu
@pytest.mark.parametrize("case", CASES, ids=key)
def test_cuda_matches_oracle(case):
    centre, ref, prm = inputs(case)
    want, dbg, ops = orc_me(centre, [[ref]], prm)
    got, st = gpu_me(centre, [[ref]], prm)
    compare_hme_state(st, dbg, prm)
    same(want, got, "CUDA vs oracle")
    for f in ("total_me_candidate_index", "me_mv_array", "me_candidate_array", "me_64x64_distortion", "rc_me_allow_gm"):
        assert not getattr(got, f).any(), f + " must stay empty for the temporal filter"
    assert got.sad_ops == ops
    if os.path.exists(GOLDEN):
        z = np.load(GOLDEN)
        same({f: z[key(case) + "/" + f] for f in TF_FIELDS}, got, "CUDA vs the reference's fixture")
