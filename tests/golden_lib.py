"""Helpers shared by tests/golden/make_golden.py (which writes the fixtures from the reference) and the
tests that read them back where /root/reference does not exist (the GPU box)."""
import ctypes as C
import itertools
import os

import numpy as np

import svtme_b200  # noqa: F401
from svt_av1_psy_b200 import abi

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
RES = [(320, 192), (640, 360), (864, 480), (1280, 720), (1920, 1080), (3840, 2160), (7680, 4320)]


def sig_deriv_full_grid():
    for enc_mode, (w, h), tl, sc, ps, qp in itertools.product(range(-3, 14), RES, (0, 2), (0, 1), (2, 1), (20, 35, 55)):
        if ps == 1 and enc_mode < 7:
            continue  # low delay forces enc_mode >= 7 (enc_handle.c:4780-4785)
        yield enc_mode, w, h, tl, sc, ps, qp


def sig_deriv_grid():
    """the committed fixture keeps every 5th combination (+ all of 1080p random access at qp 35)"""
    for i, key in enumerate(sig_deriv_full_grid()):
        enc_mode, w, h, tl, sc, ps, qp = key
        if i % 5 == 0 or ((w, h) == (1920, 1080) and ps == 2 and qp == 35):
            yield key


def sig_deriv_pic_in(po, enc_mode, w, h, tl, sc, ps, qp):
    """RefMePicIn without planes: only the picture size matters for the derivation"""
    pin = po.RefMePicIn()
    mrp = po.mrp_for_preset(enc_mode)
    pin.enc_mode, pin.width, pin.height, pin.qp, pin.pred_structure = enc_mode, w, h, qp, ps
    pin.hierarchical_levels, pin.temporal_layer_index, pin.is_ref, pin.sc_class1 = (2 if ps == 1 else 5), tl, 1, sc
    pin.frame_rate, pin.gm_level = 30 << 16, -1
    pin.only_l_bwd, pin.safe_limit_nref, pin.safe_limit_zz_th = mrp["only_l_bwd"], mrp["safe_limit_nref"], mrp["safe_limit_zz_th"]
    pin.max_l0, pin.max_l1, pin.picture_number, pin.num_lists = mrp["l0"], mrp["l1"], 5, 2
    pin.ref_count[0], pin.ref_count[1] = 2, 1
    pin.ref_picture_number[0][0], pin.ref_picture_number[0][1], pin.ref_picture_number[1][0] = 0, 1, 9
    return pin


def struct_bytes(p):
    return C.string_at(C.addressof(p), C.sizeof(p))


def fields_of(p):
    """SvtMePicParams -> plain dict (JSON-able)"""
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


def params_from_bytes(raw):
    p = abi.MePicParams()
    assert len(raw) == C.sizeof(p)
    C.memmove(C.addressof(p), bytes(raw), C.sizeof(p))
    return p


class PictureFixture:
    """me_<name>.npz: the reference's outputs for one picture of tests/util.py's case lists"""

    def __init__(self, name):
        self.path = os.path.join(GOLDEN_DIR, "me_%s.npz" % name)
        self.z = np.load(self.path)
        self.params = params_from_bytes(self.z["params"].tobytes())
        self.debug = self.z["debug"]
        self.input_crc = self.z["input_crc"]

    def field(self, f):
        return self.z["res_" + f]

    def check_inputs(self, src, refs):
        import zlib
        crc = [zlib.crc32(np.ascontiguousarray(p).tobytes()) for p in [src] + refs[0] + (refs[1] if len(refs) > 1 else [])]
        assert list(self.input_crc) == crc, "synthetic input generator drifted: regenerate tests/golden with make_golden.py"

    def compare(self, res, fields, what):
        for f in fields:
            x, y = self.field(f), getattr(res, f)
            if x.dtype.names:
                for n in x.dtype.names:
                    np.testing.assert_array_equal(x[n], y[n], err_msg="%s %s.%s" % (what, f, n))
            else:
                np.testing.assert_array_equal(x, y, err_msg="%s %s" % (what, f))


def have_picture_fixture(name):
    return os.path.exists(os.path.join(GOLDEN_DIR, "me_%s.npz" % name))


# ---- record / replay of RTCD-kernel outputs ------------------------------------------------------
def _flatten(out):
    if isinstance(out, tuple):
        return [np.asarray(v) for v in out]
    return [np.asarray(out)]


class Recorder:
    """wraps an Impl; stores every output in call order (make_golden.py)"""
    simd = None

    def __init__(self, impl):
        self.impl, self.log = impl, []

    def __getattr__(self, name):
        fn = getattr(self.impl, name)

        def call(*a, **k):
            out = fn(*a, **k)
            self.log.append((name, _flatten(out)))
            return out
        return call

    def arrays(self, prefix):
        out = {}
        for i, (name, vals) in enumerate(self.log):
            for j, v in enumerate(vals):
                out["%s/%04d/%s/%d" % (prefix, i, name, j)] = v
        return out


class Replayer:
    """returns the recorded outputs in the same call order (the runners are deterministic)"""
    simd = None

    def __init__(self, z, prefix):
        calls = {}
        for key in z.files:
            p, i, name, j = key.split("/")
            if p == prefix:
                calls.setdefault(int(i), (name, {}))[1][int(j)] = z[key]
        self.calls = [calls[i] for i in sorted(calls)]
        self.pos = 0

    def __getattr__(self, name):
        def call(*a, **k):
            want, vals = self.calls[self.pos]
            assert want == name, "golden KAT log out of step: %s vs %s at call %d" % (want, name, self.pos)
            self.pos += 1
            out = [vals[j] for j in sorted(vals)]
            conv = [v if v.ndim else (int(v) if v.dtype.kind in "iu" else v.item()) for v in out]
            return tuple(conv) if len(conv) > 1 else conv[0]
        return call
