"""Host C logic (svt-av1-psy_b200/host/me_sig_deriv.c) against the reference's own
svt_aom_sig_deriv_me, driven through oracle/_ref (needs /root/reference), and against the golden
parameter fixture where the reference is absent."""
import ctypes as C
import itertools
import json
import os

import numpy as np
import pytest

from oracle import pyoracle as po
import svtme_b200 as me
from svt_av1_psy_b200 import abi
from util import params_from_encoder_inputs

GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "sig_deriv_params.json")
RES = [(320, 192), (640, 360), (864, 480), (1280, 720), (1920, 1080), (3840, 2160), (7680, 4320)]


def grid():
    for enc_mode, (w, h), tl, sc, ps, qp in itertools.product(range(-3, 14), RES, (0, 2), (0, 1), (2, 1), (20, 35, 55)):
        if ps == 1 and enc_mode < 7:
            continue  # low delay forces enc_mode >= 7 (enc_handle.c:4780-4785)
        yield enc_mode, w, h, tl, sc, ps, qp


def struct_bytes(p):
    return bytes(memoryview(p).cast("B")) if False else C.string_at(C.addressof(p), C.sizeof(p))


def derive_host(enc_mode, w, h, tl, sc, ps, qp, frame_rate=30 << 16):
    src = np.zeros((h, w), np.uint8)
    refs = [[src, src], [src]]
    return params_from_encoder_inputs(src, refs, [[0, 1], [9]], 5, enc_mode, temporal_layer=tl, sc_class1=sc,
                                      pred_structure=ps, qp=qp, hier_levels=2 if ps == 1 else 5, frame_rate=frame_rate)


def fields_of(p):
    out = {}
    for name, _ in abi.MeCtrls._fields_:
        v = getattr(p.ctrls, name)
        if isinstance(v, abi.SearchAreaMinMax):
            v = [v.sa_min.width, v.sa_min.height, v.sa_max.width, v.sa_max.height]
        elif isinstance(v, abi.SearchArea):
            v = [v.width, v.height]
        elif hasattr(v, "__len__"):
            v = [[a.sa_min.width, a.sa_min.height, a.sa_max.width, a.sa_max.height] for a in v]
        out[name] = v
    for name in ("enable_me_8x8", "enable_me_16x16", "max_cand", "max_refs", "max_l0", "only_l_bwd", "input_resolution",
                 "gm_enabled", "gm_use_distance_based_active_th"):
        out[name] = getattr(p, name)
    return out


@pytest.mark.skipif(not os.path.isdir(po.REFERENCE_ROOT), reason="reference sources not present")
def test_sig_deriv_matches_reference(orc):
    n = 0
    for enc_mode, w, h, tl, sc, ps, qp in grid():
        src = np.zeros((8, 8), np.uint8)
        pin = po.make_pic_in(np.zeros((h, w), np.uint8)[:8, :8].copy() if False else np.zeros((h, w), np.uint8),
                             [[np.zeros((h, w), np.uint8)] * 2, [np.zeros((h, w), np.uint8)]], [[0, 1], [9]], 5, enc_mode,
                             temporal_layer=tl, sc_class1=sc, pred_structure=ps, qp=qp, hier_levels=2 if ps == 1 else 5) \
            if (w, h) == (320, 192) else None
        if pin is None:
            # only the picture size matters for the derivation: avoid allocating 8K planes
            pin = po.RefMePicIn()
            mrp = po.mrp_for_preset(enc_mode)
            pin.enc_mode, pin.width, pin.height, pin.qp, pin.pred_structure = enc_mode, w, h, qp, ps
            pin.hierarchical_levels, pin.temporal_layer_index, pin.is_ref, pin.sc_class1 = (2 if ps == 1 else 5), tl, 1, sc
            pin.frame_rate, pin.gm_level = 30 << 16, -1
            pin.only_l_bwd, pin.safe_limit_nref, pin.safe_limit_zz_th = mrp["only_l_bwd"], mrp["safe_limit_nref"], mrp["safe_limit_zz_th"]
            pin.max_l0, pin.max_l1, pin.picture_number, pin.num_lists = mrp["l0"], mrp["l1"], 5, 2
            pin.ref_count[0], pin.ref_count[1] = 2, 1
            pin.ref_picture_number[0][0], pin.ref_picture_number[0][1], pin.ref_picture_number[1][0] = 0, 1, 9
        ref = po.ref_derive_params(pin)
        host = derive_host(enc_mode, w, h, tl, sc, ps, qp)
        assert fields_of(ref) == fields_of(host), (enc_mode, w, h, tl, sc, ps, qp)
        assert struct_bytes(ref) == struct_bytes(host), (enc_mode, w, h, tl, sc, ps, qp)
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
    if not os.path.exists(GOLDEN):
        pytest.skip("golden fixture not generated yet")
    want = json.load(open(GOLDEN))
    for key, fields in want.items():
        enc_mode, w, h, tl, sc, ps, qp = [int(v) for v in key.split(",")]
        got = fields_of(derive_host(enc_mode, w, h, tl, sc, ps, qp))
        assert got == fields, key


def test_shard_owner():
    f = me.host_lib().svtme_shard_owner
    f.argtypes = [C.c_int64, C.c_int32, C.c_int32, C.c_int32]
    owners = [f(p, 32, 0, 4) for p in range(0, 129)]
    assert owners[0] == 0 and owners[1] == 0 and owners[32] == 0 and owners[33] == 1 and owners[128] == 3
    assert f(5, 32, 3, 4) == 3 and f(5, 32, 0, 1) == 0
