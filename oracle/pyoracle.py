"""TEST INFRASTRUCTURE ONLY — ctypes access to the two CPU checkers:

  * oracle/libme_oracle.so    : our plain-C restatement of the reference's ME path (me_oracle.c)
  * oracle/_ref/libsvtref.so  : the reference's own sources compiled by oracle/Makefile

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module. The product (svt-av1-psy_b200/) never does."""
import ctypes as C
import os
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
import svtme_b200  # noqa: E402,F401  (registers svt_av1_psy_b200)
from svt_av1_psy_b200 import abi  # noqa: E402

REFERENCE_ROOT = "/root/reference"
u8p = C.POINTER(C.c_uint8)


class RefMePicIn(C.Structure):
    _fields_ = [
        ("enc_mode", C.c_int32),
        ("width", C.c_int32),
        ("height", C.c_int32),
        ("qp", C.c_int32),
        ("pred_structure", C.c_int32),
        ("hierarchical_levels", C.c_int32),
        ("temporal_layer_index", C.c_int32),
        ("is_ref", C.c_int32),
        ("sc_class1", C.c_int32),
        ("frame_rate", C.c_uint32),
        ("similar_brightness_refs", C.c_int32),
        ("gm_level", C.c_int32),
        ("only_l_bwd", C.c_int32),
        ("safe_limit_nref", C.c_int32),
        ("safe_limit_zz_th", C.c_int32),
        ("max_l0", C.c_int32),
        ("max_l1", C.c_int32),
        ("picture_number", C.c_uint64),
        ("num_lists", C.c_int32),
        ("ref_count", C.c_int32 * 2),
        ("ref_picture_number", (C.c_uint64 * 4) * 2),
        ("src_luma", C.c_void_p),
        ("ref_luma", (C.c_void_p * 4) * 2),
        ("simd", C.c_int32),
    ]


class RefMeDebug(C.Structure):
    _q = (C.c_int16 * 2) * 2
    _fields_ = (
        [(n, ((C.c_int16 * 2) * 2 * 4) * 2) for n in
         ("x_hme_l0", "y_hme_l0", "x_hme_l1", "y_hme_l1", "x_hme_l2", "y_hme_l2")]
        + [(n, ((C.c_uint64 * 2) * 2 * 4) * 2) for n in ("hme_l0_sad", "hme_l1_sad", "hme_l2_sad")]
        + [("prehme_mv_x", ((C.c_int16 * 2) * 4) * 2), ("prehme_mv_y", ((C.c_int16 * 2) * 4) * 2),
           ("prehme_sad", ((C.c_uint64 * 2) * 4) * 2),
           ("zz_sad", (C.c_uint32 * 4) * 2), ("reduce_me_sr_divisor", (C.c_uint32 * 4) * 2)]
    )


DEBUG_DTYPE = np.dtype([
    ("x_hme_l0", "<i2", (2, 4, 2, 2)), ("y_hme_l0", "<i2", (2, 4, 2, 2)),
    ("x_hme_l1", "<i2", (2, 4, 2, 2)), ("y_hme_l1", "<i2", (2, 4, 2, 2)),
    ("x_hme_l2", "<i2", (2, 4, 2, 2)), ("y_hme_l2", "<i2", (2, 4, 2, 2)),
    ("hme_l0_sad", "<u8", (2, 4, 2, 2)), ("hme_l1_sad", "<u8", (2, 4, 2, 2)), ("hme_l2_sad", "<u8", (2, 4, 2, 2)),
    ("prehme_mv_x", "<i2", (2, 4, 2)), ("prehme_mv_y", "<i2", (2, 4, 2)), ("prehme_sad", "<u8", (2, 4, 2)),
    ("zz_sad", "<u4", (2, 4)), ("reduce_me_sr_divisor", "<u4", (2, 4)),
])
assert DEBUG_DTYPE.itemsize == C.sizeof(RefMeDebug), (DEBUG_DTYPE.itemsize, C.sizeof(RefMeDebug))


def build(ref=True, quiet=True):
    """(Re)build the checkers. Building the checker is not using it."""
    targets = ["oracle"]
    if ref and os.path.isdir(REFERENCE_ROOT):
        targets.append("ref")
    out = subprocess.run(["make", "-C", HERE] + targets, capture_output=True, text=True)
    if out.returncode != 0:
        raise RuntimeError("oracle build failed:\n" + out.stdout + out.stderr)
    if not quiet:
        print(out.stdout)


_ref = None
_orc = None


def ref_available():
    return os.path.exists(os.path.join(HERE, "_ref", "libsvtref.so"))


def ref_lib():
    global _ref
    if _ref is None:
        path = os.path.join(HERE, "_ref", "libsvtref.so")
        if not os.path.exists(path):
            raise RuntimeError("oracle/_ref/libsvtref.so not built (needs /root/reference): run `make -C oracle ref`")
        _ref = C.CDLL(path)
        _ref.ref_me_picture.argtypes = [C.POINTER(RefMePicIn), C.POINTER(abi.MePicResults), C.c_void_p,
                                        C.POINTER(C.c_double)]
        _ref.ref_me_picture.restype = C.c_int
        _ref.ref_derive_params.argtypes = [C.POINTER(RefMePicIn), C.POINTER(abi.MePicParams)]
        _ref.ref_nxm_sad_kernel.restype = C.c_uint32
    return _ref


_wrap = None


def wrap_available():
    return os.path.exists(os.path.join(HERE, "_ref", "libsvtref_wrap.so"))


def wrap_lib():
    """oracle/_ref/libsvtref_wrap.so: the reference's ME callers linked with -Wl,--wrap and the product's
    reference-side binding (svt-av1-psy_b200/shim/svtme_wrap.c): svt_aom_motion_estimation_b64 and
    dg_detector_hme_level0 are served by the GPU library. Same harness entry points as ref_lib()."""
    global _wrap
    if _wrap is None:
        path = os.path.join(HERE, "_ref", "libsvtref_wrap.so")
        if not os.path.exists(path):
            raise RuntimeError("oracle/_ref/libsvtref_wrap.so not built (needs /root/reference): run `make -C oracle ref`")
        _wrap = C.CDLL(path)
        _wrap.ref_me_picture.argtypes = [C.POINTER(RefMePicIn), C.POINTER(abi.MePicResults), C.c_void_p,
                                         C.POINTER(C.c_double)]
        _wrap.ref_me_picture.restype = C.c_int
        _wrap.ref_derive_params.argtypes = [C.POINTER(RefMePicIn), C.POINTER(abi.MePicParams)]
        _wrap.ref_derive_params_shim.argtypes = [C.POINTER(RefMePicIn), C.POINTER(abi.MePicParams)]
        _wrap.svtme_shim_stats.argtypes = [C.POINTER(C.c_uint64)] * 3
        _wrap.svtme_shim_stats.restype = None
        _wrap.svtme_shim_batches.restype = C.c_uint64
    return _wrap


def shim_stats():
    """(pictures, temporal-filter pairs, detector pairs) the wrapped library has run on the GPU"""
    a, b, c = C.c_uint64(0), C.c_uint64(0), C.c_uint64(0)
    wrap_lib().svtme_shim_stats(C.byref(a), C.byref(b), C.byref(c))
    return a.value, b.value, c.value


def oracle_lib():
    global _orc
    if _orc is None:
        path = os.path.join(HERE, "libme_oracle.so")
        if not os.path.exists(path):
            build(ref=False)
        _orc = C.CDLL(path)
        _orc.orc_me_picture.restype = C.c_int
        _orc.orc_nxm_sad_kernel.restype = C.c_uint32
    return _orc


def ptr(a):
    return a.ctypes.data_as(C.c_void_p)


from svt_av1_psy_b200.gop import mrp_for_preset  # noqa: E402,F401


def make_pic_in(src, refs, ref_pocs, poc, enc_mode, qp=35, temporal_layer=1, hier_levels=5, is_ref=1,
                simd=0, sc_class1=0, pred_structure=2, frame_rate=30 << 16, similar_brightness=0,
                gm_level=-1, mrp=None):
    """src: (H,W) u8; refs: [[l0 planes...],[l1 planes...]]; ref_pocs same shape."""
    h, w = src.shape
    mrp = mrp or mrp_for_preset(enc_mode)
    pin = RefMePicIn()
    pin.enc_mode, pin.width, pin.height, pin.qp = enc_mode, w, h, qp
    pin.pred_structure, pin.hierarchical_levels = pred_structure, hier_levels
    pin.temporal_layer_index, pin.is_ref, pin.sc_class1 = temporal_layer, is_ref, sc_class1
    pin.frame_rate, pin.similar_brightness_refs, pin.gm_level = frame_rate, similar_brightness, gm_level
    pin.only_l_bwd, pin.safe_limit_nref, pin.safe_limit_zz_th = mrp["only_l_bwd"], mrp["safe_limit_nref"], mrp["safe_limit_zz_th"]
    pin.max_l0, pin.max_l1 = mrp["l0"], mrp["l1"]
    pin.picture_number = poc
    pin.num_lists = 2 if len(refs) > 1 and len(refs[1]) > 0 else 1
    keep = [np.ascontiguousarray(src)]
    pin.src_luma = keep[0].ctypes.data
    for l in range(pin.num_lists):
        pin.ref_count[l] = len(refs[l])
        for r, plane in enumerate(refs[l]):
            a = np.ascontiguousarray(plane)
            assert a.shape == (h, w) and a.dtype == np.uint8
            keep.append(a)
            pin.ref_luma[l][r] = a.ctypes.data
            pin.ref_picture_number[l][r] = ref_pocs[l][r]
    pin.simd = simd
    pin._keep = keep
    return pin


def ref_derive_params(pin, lib=None):
    p = abi.MePicParams()
    (lib or ref_lib()).ref_derive_params(C.byref(pin), C.byref(p))
    return p


def ref_derive_params_shim(pin):
    """the parameters as the product's shim flattens them from the reference's MeContext / PCS"""
    p = abi.MePicParams()
    rc = wrap_lib().ref_derive_params_shim(C.byref(pin), C.byref(p))
    assert rc == 0
    return p


def ref_me_picture(pin, params=None, debug=False, lib=None):
    """run the reference's own ME on one picture -> (PicResultArrays, debug array or None, seconds).
    lib=wrap_lib(): the same caller code, with the b64 function served by the GPU library through the shim."""
    params = params or ref_derive_params(pin, lib)
    res = abi.PicResultArrays(pin.width, pin.height, params)
    dbg = np.zeros(res.n_b64, dtype=DEBUG_DTYPE) if debug else None
    sec = C.c_double(0)
    rc = (lib or ref_lib()).ref_me_picture(C.byref(pin), C.byref(res.struct), ptr(dbg) if debug else None, C.byref(sec))
    if rc != 0:
        raise RuntimeError("ref_me_picture failed rc=%d" % rc)
    return res, dbg, sec.value


def ref_me_picture_mt(pin, n_threads, params=None, lib=None):
    """the same picture with its b64s dealt out to n_threads threads, one MeContext each (the ME segment threads)"""
    lib = lib or ref_lib()
    params = params or ref_derive_params(pin, lib)
    res = abi.PicResultArrays(pin.width, pin.height, params)
    f = lib.ref_me_picture_mt
    f.argtypes = [C.POINTER(RefMePicIn), C.POINTER(abi.MePicResults), C.c_int, C.POINTER(C.c_double)]
    f.restype = C.c_int
    sec = C.c_double(0)
    rc = f(C.byref(pin), C.byref(res.struct), n_threads, C.byref(sec))
    if rc != 0:
        raise RuntimeError("ref_me_picture_mt failed rc=%d" % rc)
    return res, sec.value


def ref_me_minigop(pins, n_threads=4, submit=True, lib=None):
    """several pictures released together (a mini-GOP), their b64s dealt out to n_threads ME threads; with
    lib=wrap_lib() and submit=True they are first queued as ONE GPU batch by the shim -> [PicResultArrays]"""
    lib = lib or ref_lib()
    n = len(pins)
    arr = (RefMePicIn * n)()
    for i, p in enumerate(pins):
        C.memmove(C.addressof(arr[i]), C.addressof(p), C.sizeof(RefMePicIn))
    results = [abi.PicResultArrays(p.width, p.height, ref_derive_params(p, lib)) for p in pins]
    rs = (abi.MePicResults * n)()
    for i, r in enumerate(results):
        C.memmove(C.addressof(rs[i]), C.addressof(r.struct), C.sizeof(abi.MePicResults))
    f = lib.ref_me_minigop
    f.argtypes = [C.POINTER(RefMePicIn), C.c_int, C.POINTER(abi.MePicResults), C.c_int, C.c_int]
    f.restype = C.c_int
    rc = f(arr, n, rs, n_threads, 1 if submit else 0)
    if rc != 0:
        raise RuntimeError("ref_me_minigop failed rc=%d" % rc)
    return results


def dg_detector(lib, prefix, src, ref, per_b64=False):
    """dynamic-GOP detector metrics through `orc_dg_detector` (oracle) or `ref_dg_detector` (the reference's own code)"""
    h, w = src.shape
    m = abi.DgMetrics()
    m.input_resolution = abi.input_resolution_of(w, h)
    s, r = np.ascontiguousarray(src), np.ascontiguousarray(ref)
    if prefix == "ref_":
        rc = lib.ref_dg_detector(C.c_void_p(s.ctypes.data), C.c_void_p(r.ctypes.data), w, h, C.byref(m))
        assert rc == 0
        return m, None, None
    n = abi.n_b64_of(w, h)
    sad, mv = np.zeros(n, np.uint64), np.zeros((n, 2), np.int16)
    rc = lib.orc_dg_detector(C.c_void_p(s.ctypes.data), C.c_void_p(r.ctypes.data), w, h, C.byref(m),
                             C.c_void_p(sad.ctypes.data), C.c_void_p(mv.ctypes.data))
    assert rc == 0
    return m, sad, mv


class RefTfIn(C.Structure):
    """RefTfIn of oracle/ref_harness.h"""
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("enc_mode", C.c_int32), ("qp", C.c_int32),
                ("tf_hme_me_level", C.c_int32), ("tf_qp_opt", C.c_int32), ("tf_me_exit_th", C.c_uint32),
                ("temporal_layer_index", C.c_int32), ("is_ref", C.c_int32),
                ("centre_picture_number", C.c_uint64), ("ref_picture_number", C.c_uint64),
                ("centre_luma", C.c_void_p), ("ref_luma", C.c_void_p), ("simd", C.c_int32)]


def ref_tf_me_picture(centre, ref, enc_mode, hme_me_level, qp_opt=0, me_exit_th=0, qp=35, centre_poc=8, ref_poc=7,
                      temporal_layer=0, is_ref=1, simd=0, debug=True, lib=None):
    """temporal-filter ME of `centre` against `ref` through the reference's own code
    -> (params, PicResultArrays, debug array)"""
    h, w = centre.shape
    tin = RefTfIn()
    tin.width, tin.height, tin.enc_mode, tin.qp = w, h, enc_mode, qp
    tin.tf_hme_me_level, tin.tf_qp_opt, tin.tf_me_exit_th = hme_me_level, qp_opt, me_exit_th
    tin.temporal_layer_index, tin.is_ref = temporal_layer, is_ref
    tin.centre_picture_number, tin.ref_picture_number = centre_poc, ref_poc
    c8, r8 = np.ascontiguousarray(centre), np.ascontiguousarray(ref)
    tin.centre_luma, tin.ref_luma, tin.simd = c8.ctypes.data, r8.ctypes.data, simd
    params = abi.MePicParams()
    tmp = abi.MePicParams()
    tmp.enable_me_8x8, tmp.max_refs, tmp.max_cand = 1, 2, 3
    res = abi.PicResultArrays(w, h, tmp)
    dbg = np.zeros(res.n_b64, dtype=DEBUG_DTYPE) if debug else None
    f = (lib or ref_lib()).ref_tf_me_picture
    f.argtypes = [C.POINTER(RefTfIn), C.POINTER(abi.MePicParams), C.POINTER(abi.MePicResults), C.c_void_p]
    f.restype = C.c_int
    rc = f(C.byref(tin), C.byref(params), C.byref(res.struct), ptr(dbg) if debug else None)
    if rc != 0:
        raise RuntimeError("ref_tf_me_picture failed rc=%d" % rc)
    return params, res, dbg
