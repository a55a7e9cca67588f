"""Shared helpers of the test-suite: running the oracle, comparing result sets, the case lists."""
import ctypes as C

import numpy as np

from oracle import pyoracle as po
import svtme_b200  # noqa: F401
from svt_av1_psy_b200 import abi
from svt_av1_psy_b200.synth import PanningNoise

RES_FIELDS = [f for f in abi.PicResultArrays.FIELDS if f != "me_windows"]

# (w, h, enc_mode, temporal_layer, poc, l0 pocs, l1 pocs)
CASES = [
    (320, 192, 4, 1, 16, [0, 8, 4, 2], [32, 24, 20]),
    (320, 192, 8, 3, 6, [4, 0, 2], [8, 16]),
    (320, 192, 4, 0, 32, [0, 16, 8, 4], [0, 16, 8]),     # base layer: list 1 mirrors list 0
    (256, 136, 2, 2, 12, [8, 0, 4, 2], [16, 24, 32]),    # incomplete last b64 row (136 = 2*64 + 8)
    (192, 128, -2, 1, 8, [0, 4, 2, 6], [16, 12, 10]),    # MRP: widest windows
    (320, 192, 10, 4, 3, [2, 0], [4, 8]),
    (320, 192, 6, 2, 4, [0], [8]),                       # one ref per list: mrp_off builder
    (320, 192, 12, 1, 4, [0], []),                       # P picture, single ref builder
    (200, 72, 4, 1, 8, [0, 4], [16, 12]),                # widths/heights that are not multiples of 64
]
# smallest pictures: one b64, one partial b64 (search windows larger than the picture, every clamp active), a column of b64s
SMALL = [
    (64, 64, 4, 2, 8, [4, 0], [12]),
    (72, 40, 8, 2, 8, [4, 0], [12]),
    (136, 64, 2, 2, 8, [4, 0], [12]),
    (64, 200, 6, 2, 8, [4, 0], [12]),
]

EXTRA = [
    ("sc_M4", dict(w=320, h=192, enc_mode=4, tl=1, poc=16, l0=[0, 8], l1=[32, 24]), dict(sc_class1=1)),
    ("sc_M9", dict(w=320, h=192, enc_mode=9, tl=2, poc=6, l0=[4, 0], l1=[8, 16]), dict(sc_class1=1)),
    ("rtc_M9", dict(w=320, h=192, enc_mode=9, tl=1, poc=6, l0=[5, 4], l1=[3, 1]), dict(pred_structure=1, hier_levels=2)),
    ("rtc_sc_M10", dict(w=320, h=192, enc_mode=10, tl=0, poc=6, l0=[5, 4], l1=[5, 4]), dict(pred_structure=1, sc_class1=1, hier_levels=2)),
    ("rtc_M12", dict(w=320, h=192, enc_mode=12, tl=1, poc=6, l0=[5], l1=[4]), dict(pred_structure=1, hier_levels=2)),
    ("gm_M0", dict(w=320, h=192, enc_mode=0, tl=1, poc=8, l0=[0, 4, 2, 6], l1=[16, 12, 10]), dict()),
    ("gm_MRS_720", dict(w=1280, h=128, enc_mode=-3, tl=1, poc=4, l0=[0, 2], l1=[8, 6]), dict()),
    ("qp20_M8", dict(w=320, h=192, enc_mode=8, tl=3, poc=6, l0=[4, 0, 2], l1=[8, 16]), dict(qp=20)),
    ("qp63_M8", dict(w=320, h=192, enc_mode=8, tl=3, poc=6, l0=[4, 0, 2], l1=[8, 16]), dict(qp=63)),
    ("lowfps_M5", dict(w=320, h=192, enc_mode=5, tl=3, poc=6, l0=[4, 0, 2], l1=[8, 16]), dict(frame_rate=10 << 16)),
    ("safe_limit_M8", dict(w=320, h=192, enc_mode=8, tl=5, poc=5, l0=[4, 0, 2], l1=[6, 8]),
     dict(similar_brightness=1, mrp=dict(only_l_bwd=1, safe_limit_nref=1, safe_limit_zz_th=60000, l0=4, l1=3))),
]


def case_id(c):
    return "%dx%d_M%d_tl%d" % (c[0], c[1], c[2], c[3])


def case_inputs(case):
    w, h, enc_mode, tl, poc, l0, l1 = case
    clip = PanningNoise(w, h, seed=11 + enc_mode, vx=3, vy=2, max_frames=40)
    src = clip.frame(poc)
    refs = [[clip.frame(p) for p in l0], [clip.frame(p) for p in l1]]
    hier = 5 if enc_mode <= 8 else 4
    kw = dict(temporal_layer=tl, hier_levels=hier, is_ref=1 if tl < hier else 0)
    return src, refs, [l0, l1], poc, enc_mode, kw


def extra_inputs(geo, kw):
    w, h = geo["w"], geo["h"]
    clip = PanningNoise(w, h, seed=101, vx=5, vy=3, max_frames=40, sigma=4.0)
    src = clip.frame(geo["poc"])
    refs = [[clip.frame(p) for p in geo["l0"]], [clip.frame(p) for p in geo["l1"]]]
    kw = dict(kw)
    kw.setdefault("hier_levels", 5)
    kw["temporal_layer"] = geo["tl"]
    return src, refs, [geo["l0"], geo["l1"]], geo["poc"], geo["enc_mode"], kw


def orc_me(src, refs, params, debug=True):
    """run oracle/me_oracle.c on one picture -> (PicResultArrays, debug array, sad_ops)"""
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


def params_from_encoder_inputs(src, refs, pocs, poc, enc_mode, **kw):
    """SvtMePicParams through the product's host C logic (svtme_sig_deriv_me)."""
    import svtme_b200 as me
    h, w = src.shape
    mrp = kw.get("mrp") or po.mrp_for_preset(enc_mode)
    sin = me.binding.SigDerivIn()
    sin.enc_mode, sin.width, sin.height, sin.qp = enc_mode, w, h, kw.get("qp", 35)
    sin.pred_structure, sin.hierarchical_levels = kw.get("pred_structure", 2), kw.get("hier_levels", 5)
    sin.temporal_layer_index, sin.is_ref, sin.sc_class1 = kw.get("temporal_layer", 1), kw.get("is_ref", 1), kw.get("sc_class1", 0)
    sin.frame_rate, sin.similar_brightness_refs, sin.gm_level = kw.get("frame_rate", 30 << 16), kw.get("similar_brightness", 0), kw.get("gm_level", -1)
    sin.only_l_bwd, sin.safe_limit_nref, sin.safe_limit_zz_th = mrp["only_l_bwd"], mrp["safe_limit_nref"], mrp["safe_limit_zz_th"]
    sin.max_l0, sin.max_l1 = mrp["l0"], mrp["l1"]
    sin.picture_number = poc
    sin.num_lists = 2 if len(refs) > 1 and len(refs[1]) > 0 else 1
    for l in range(sin.num_lists):
        sin.ref_count[l] = len(refs[l])
        for r in range(len(refs[l])):
            sin.ref_picture_number[l][r] = pocs[l][r]
    return me.sig_deriv_me(sin)


def gpu_me(src, refs, params, session=None):
    """upload src + refs into a session and run the CUDA picture ME -> (PicResultArrays, hme state)"""
    import copy
    import svtme_b200 as me
    h, w = src.shape
    own = session is None
    s = session or me.Session(w, h, max_pictures=12)
    p = copy.copy(params)
    s.upload(0, src)
    pid = 1
    for l in range(len(refs)):
        for r, plane in enumerate(refs[l]):
            s.upload(pid, plane)
            p.ref_pic_id[l][r] = pid
            pid += 1
    s.sad_ops(reset=True)
    res = s.me(0, p)
    st = s.hme_state()
    res.sad_ops = sum(s.sad_ops())
    if own:
        s.close()
    return res, st


def compare_hme_state(st, dbg, params):
    """device HME state vs the oracle's debug dump, for the refs that were searched"""
    lv = [("x_hme_l0", "y_hme_l0", "hme_l0_sad"), ("x_hme_l1", "y_hme_l1", "hme_l1_sad"), ("x_hme_l2", "y_hme_l2", "hme_l2_sad")]
    c = params.ctrls
    enabled = [c.enable_hme_level0_flag, c.enable_hme_level1_flag, c.enable_hme_level2_flag]
    for l in range(params.num_of_list_to_search):
        if l == 1 and params.temporal_layer_index == 0:
            continue
        for r in range(params.num_of_ref_pic_to_search[l]):
            for k in range(3):
                if not (c.enable_hme_flag and enabled[k]):
                    continue
                np.testing.assert_array_equal(st["x_hme"][:, k, l, r], dbg[lv[k][0]][:, l, r], err_msg="x_hme L%d l%d r%d" % (k, l, r))
                np.testing.assert_array_equal(st["y_hme"][:, k, l, r], dbg[lv[k][1]][:, l, r], err_msg="y_hme L%d l%d r%d" % (k, l, r))
                np.testing.assert_array_equal(st["hme_sad"][:, k, l, r], dbg[lv[k][2]][:, l, r], err_msg="hme_sad L%d l%d r%d" % (k, l, r))
            if c.prehme_enable:
                np.testing.assert_array_equal(st["prehme_sad"][:, l, r], dbg["prehme_sad"][:, l, r], err_msg="prehme_sad l%d r%d" % (l, r))
                np.testing.assert_array_equal(st["prehme_mv_x"][:, l, r], dbg["prehme_mv_x"][:, l, r], err_msg="prehme_mv_x")
                np.testing.assert_array_equal(st["prehme_mv_y"][:, l, r], dbg["prehme_mv_y"][:, l, r], err_msg="prehme_mv_y")
            np.testing.assert_array_equal(st["reduce_me_sr_divisor"][:, l, r], dbg["reduce_me_sr_divisor"][:, l, r], err_msg="divisor")
