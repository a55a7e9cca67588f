"""Picture-level parity against the committed outputs of the REFERENCE (tests/golden/me_*.npz, written by
tests/golden/make_golden.py from oracle/_ref). These run where /root/reference does not exist:
  * CPU: the oracle (and the host parameter derivation) against the fixtures = the oracle stays pinned;
  * GPU: the CUDA path, through the C ABI, against the reference's outputs directly."""
import numpy as np
import pytest

import golden_lib as gl
from util import (CASES, EXTRA, RES_FIELDS, case_id, case_inputs, compare_hme_state, extra_inputs, gpu_me, orc_me,
                  params_from_encoder_inputs)

ALL = [(case_id(c), lambda c=c: case_inputs(c)) for c in CASES] + [(n, lambda g=g, k=k: extra_inputs(g, k)) for n, g, k in EXTRA]


def load(name, make):
    if not gl.have_picture_fixture(name):
        pytest.skip("fixture me_%s.npz not generated" % name)
    fx = gl.PictureFixture(name)
    src, refs, pocs, poc, enc_mode, kw = make()
    fx.check_inputs(src, refs)
    return fx, src, refs, pocs, poc, enc_mode, kw


@pytest.mark.parametrize("name,make", ALL, ids=[a[0] for a in ALL])
def test_oracle_matches_golden(orc, name, make):
    fx, src, refs, pocs, poc, enc_mode, kw = load(name, make)
    host = params_from_encoder_inputs(src, refs, pocs, poc, enc_mode, **kw)
    assert gl.struct_bytes(host) == gl.struct_bytes(fx.params), "host svtme_sig_deriv_me differs from the reference's derivation"
    res, dbg, _ = orc_me(src, refs, fx.params)
    fx.compare(res, RES_FIELDS, "oracle vs reference fixture")
    for n in fx.debug.dtype.names:
        np.testing.assert_array_equal(fx.debug[n], dbg[n], err_msg="debug field " + n)


@pytest.mark.gpu
@pytest.mark.parametrize("name,make", ALL, ids=[a[0] for a in ALL])
def test_cuda_matches_golden(name, make):
    fx, src, refs, pocs, poc, enc_mode, kw = load(name, make)
    res, st = gpu_me(src, refs, fx.params)
    fx.compare(res, RES_FIELDS, "CUDA vs reference fixture")
    compare_hme_state(st, fx.debug, fx.params)
