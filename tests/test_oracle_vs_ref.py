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

RES_FIELDS = [f for f in abi.PicResultArrays.FIELDS if f != "me_windows"]


def run_both(src, refs, pocs, poc, enc_mode, **kw):
    pin = po.make_pic_in(src, refs, pocs, poc, enc_mode, **kw)
    params = po.ref_derive_params(pin)
    ref_res, ref_dbg, _ = po.ref_me_picture(pin, params, debug=True)
    orc_res, orc_dbg, ops = orc_me(src, refs, params)
    return params, ref_res, ref_dbg, orc_res, orc_dbg


def orc_me(src, refs, params, debug=True):
    h, w = src.shape
    res = abi.PicResultArrays(w, h, params)
    dbg = np.zeros(res.n_b64, dtype=po.DEBUG_DTYPE) if debug else None
    arr = ((C.c_void_p * 4) * 2)()
    keep = []
    for l in range(len(refs)):
        for r, p in enumerate(refs[l]):
            a = np.ascontiguousarray(p)
            keep.append(a)
            arr[l][r] = a.ctypes.data
    ops = C.c_uint64(0)
    s = np.ascontiguousarray(src)
    rc = po.oracle_lib().orc_me_picture(C.c_void_p(s.ctypes.data), arr, w, h, C.byref(params), C.byref(res.struct),
                                        po.ptr(dbg) if debug else None, C.byref(ops))
    assert rc == 0
    return res, dbg, ops.value


def compare(a, b, dbg_a=None, dbg_b=None, fields=RES_FIELDS):
    if dbg_a is not None:
        for name in po.DEBUG_DTYPE.names:
            np.testing.assert_array_equal(dbg_a[name], dbg_b[name], err_msg="debug field " + name)
    for f in fields:
        x, y = getattr(a, f), getattr(b, f)
        if x.dtype.names:
            for n in x.dtype.names:
                np.testing.assert_array_equal(x[n], y[n], err_msg="%s.%s" % (f, n))
        else:
            np.testing.assert_array_equal(x, y, err_msg=f)


CASES = [
    # (w, h, enc_mode, temporal_layer, poc, l0 pocs, l1 pocs)
    (320, 192, 4, 1, 16, [0, 8, 4, 2], [32, 24, 20]),
    (320, 192, 8, 3, 6, [4, 0, 2], [8, 16]),
    (320, 192, 4, 0, 32, [0, 16, 8, 4], [0, 16, 8]),     # base layer: list 1 mirrors list 0
    (256, 136, 2, 2, 12, [8, 0, 4, 2], [16, 24, 32]),    # incomplete last b64 row (136 = 2*64 + 8)
    (192, 128, -2, 1, 8, [0, 4, 2, 6], [16, 12, 10]),    # MRP: widest windows
    (320, 192, 10, 4, 3, [2, 0], [4, 8]),
    (320, 192, 6, 2, 4, [0], [8]),                       # one ref per list: mrp_off builder
    (320, 192, 12, 1, 4, [0], []),                       # P picture, single ref builder
]


@pytest.mark.parametrize("case", CASES, ids=lambda c: "%dx%d_M%d_tl%d" % (c[0], c[1], c[2], c[3]))
def test_picture_me_matches_reference(orc, case):
    w, h, enc_mode, tl, poc, l0, l1 = case
    clip = PanningNoise(w, h, seed=11 + enc_mode, vx=3, vy=2, max_frames=40)
    src = clip.frame(poc)
    refs = [[clip.frame(p) for p in l0], [clip.frame(p) for p in l1]]
    hier = 5 if enc_mode <= 8 else 4
    _, ref_res, ref_dbg, orc_res, orc_dbg = run_both(src, refs, [l0, l1], poc, enc_mode, temporal_layer=tl,
                                                     hier_levels=hier, is_ref=1 if tl < hier else 0)
    compare(ref_res, orc_res, ref_dbg, orc_dbg)


EXTRA = [
    # name, dict(w,h,enc_mode,tl,poc,l0,l1), make_pic_in kwargs
    ("sc_M4", dict(w=320, h=192, enc_mode=4, tl=1, poc=16, l0=[0, 8], l1=[32, 24]), dict(sc_class1=1)),
    ("sc_M9", dict(w=320, h=192, enc_mode=9, tl=2, poc=6, l0=[4, 0], l1=[8, 16]), dict(sc_class1=1)),
    ("rtc_M9", dict(w=320, h=192, enc_mode=9, tl=1, poc=6, l0=[5, 4, 2], l1=[3, 1]), dict(pred_structure=1, hier_levels=2)),
    ("rtc_sc_M10", dict(w=320, h=192, enc_mode=10, tl=0, poc=6, l0=[5, 4], l1=[5, 4]), dict(pred_structure=1, sc_class1=1, hier_levels=2)),
    ("rtc_M12", dict(w=320, h=192, enc_mode=12, tl=1, poc=6, l0=[5], l1=[4]), dict(pred_structure=1, hier_levels=2)),
    ("gm_M0", dict(w=320, h=192, enc_mode=0, tl=1, poc=8, l0=[0, 4, 2, 6], l1=[16, 12, 10]), dict()),
    ("gm_MRS_720", dict(w=1280, h=128, enc_mode=-3, tl=1, poc=4, l0=[0, 2], l1=[8, 6]), dict()),
    ("qp20_M8", dict(w=320, h=192, enc_mode=8, tl=3, poc=6, l0=[4, 0, 2], l1=[8, 16]), dict(qp=20)),
    ("qp63_M8", dict(w=320, h=192, enc_mode=8, tl=3, poc=6, l0=[4, 0, 2], l1=[8, 16]), dict(qp=63)),
    ("lowfps_M5", dict(w=320, h=192, enc_mode=5, tl=3, poc=6, l0=[4, 0, 2, 3], l1=[8, 16, 7]), dict(frame_rate=10 << 16)),
    ("safe_limit_M8", dict(w=320, h=192, enc_mode=8, tl=5, poc=5, l0=[4, 0, 2], l1=[6, 8]),
     dict(similar_brightness=1, mrp=dict(only_l_bwd=1, safe_limit_nref=1, safe_limit_zz_th=60000, l0=4, l1=3))),
]


@pytest.mark.parametrize("name,geo,kw", EXTRA, ids=[e[0] for e in EXTRA])
def test_picture_me_variants(orc, name, geo, kw):
    w, h = geo["w"], geo["h"]
    clip = PanningNoise(w, h, seed=101, vx=5, vy=-3 if False else 3, max_frames=40, sigma=4.0)
    src = clip.frame(geo["poc"])
    refs = [[clip.frame(p) for p in geo["l0"]], [clip.frame(p) for p in geo["l1"]]]
    kw = dict(kw)
    kw.setdefault("hier_levels", 5)
    _, ref_res, ref_dbg, orc_res, orc_dbg = run_both(src, refs, [geo["l0"], geo["l1"]], geo["poc"], geo["enc_mode"],
                                                     temporal_layer=geo["tl"], **kw)
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
    a, _, _ = po.ref_me_picture(po.make_pic_in(src, refs, [l0, l1], 16, 4, simd=0))
    b, _, _ = po.ref_me_picture(po.make_pic_in(src, refs, [l0, l1], 16, 4, simd=simd))
    compare(a, b)
