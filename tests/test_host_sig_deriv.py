"""Host C logic (svt-av1-psy_b200/host/me_sig_deriv.c) against the reference's own
svt_aom_sig_deriv_me, driven through oracle/_ref (needs /root/reference), and against the golden
parameter fixture where the reference is absent."""
import ctypes as C
import gzip
import json
import os

import numpy as np
import pytest

from oracle import pyoracle as po
import svtme_b200 as me
from svt_av1_psy_b200 import abi
import golden_lib as gl
from util import params_from_encoder_inputs

GOLDEN = os.path.join(gl.GOLDEN_DIR, "sig_deriv_params.json.gz")


def derive_host(enc_mode, w, h, tl, sc, ps, qp, frame_rate=30 << 16):
    src = np.zeros((h, w), np.uint8)
    refs = [[src, src], [src]]
    return params_from_encoder_inputs(src, refs, [[0, 1], [9]], 5, enc_mode, temporal_layer=tl, sc_class1=sc,
                                      pred_structure=ps, qp=qp, hier_levels=2 if ps == 1 else 5, frame_rate=frame_rate)


@pytest.mark.skipif(not os.path.isdir(po.REFERENCE_ROOT), reason="reference sources not present")
def test_sig_deriv_matches_reference(orc):
    n = 0
    for key in gl.sig_deriv_full_grid():
        ref = po.ref_derive_params(gl.sig_deriv_pic_in(po, *key))
        host = derive_host(*key)
        assert gl.fields_of(ref) == gl.fields_of(host), key
        assert gl.struct_bytes(ref) == gl.struct_bytes(host), key
        n += 1
    assert n > 1000


def test_sig_deriv_low_frame_rate_and_errors():
    a = derive_host(5, 640, 360, 1, 0, 2, 35, frame_rate=30 << 16)
    assert a.ctrls.me_sa.sa_min.width == 24  # (16*3)>>1: scs->frame_rate >> 16 is non-zero for any Q16 rate
    bad = me.binding.SigDerivIn()
    p = abi.MePicParams()
    assert me.host_lib().svtme_sig_deriv_me(C.byref(bad), C.byref(p)) == abi.ERR_BAD_PARAMETER


def test_sig_deriv_golden():
    """committed fixture generated from the reference (tests/golden/make_golden.py)"""
    want = json.loads(gzip.open(GOLDEN).read())
    assert len(want) > 400
    for key, fields in want.items():
        got = gl.fields_of(derive_host(*[int(v) for v in key.split(",")]))
        assert got == fields, key


def test_shard_owner():
    f = me.host_lib().svtme_shard_owner
    f.argtypes = [C.c_int64, C.c_int32, C.c_int32, C.c_int32]
    owners = [f(p, 32, 0, 4) for p in range(0, 129)]
    assert owners[0] == 0 and owners[1] == 0 and owners[32] == 0 and owners[33] == 1 and owners[128] == 3
    assert f(5, 32, 3, 4) == 3 and f(5, 32, 0, 1) == 0
