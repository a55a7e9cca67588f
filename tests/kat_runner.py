"""Calls one RTCD kernel through a given implementation (oracle C, reference C/SIMD, CUDA C-ABI) with
identical arguments and returns every output, so that tests can compare implementations pairwise."""
import ctypes as C

import numpy as np


class Impl:
    """prefix: 'orc_' (oracle/libme_oracle.so), 'ref_' (oracle/_ref, extra leading simd arg), 'svtme_' (CUDA C ABI)"""

    def __init__(self, lib, prefix, simd=None):
        self.lib, self.prefix, self.simd = lib, prefix, simd

    def _fn(self, name, restype=None):
        f = getattr(self.lib, self.prefix + name)
        f.restype = restype
        return f

    def _args(self, *a):
        return ((self.simd,) if self.simd is not None else ()) + a

    def sad_loop(self, src, src_stride, ref, ref_stride, bh, bw, raw, skip, sa_w, sa_h, init=(12345, -7, 9)):
        best = C.c_uint64(init[0])
        x, y = C.c_int16(init[1]), C.c_int16(init[2])
        self._fn("sad_loop_kernel")(*self._args(
            C.c_void_p(src.ctypes.data), C.c_uint32(src_stride), C.c_void_p(ref.ctypes.data), C.c_uint32(ref_stride),
            C.c_uint32(bh), C.c_uint32(bw), C.byref(best), C.byref(x), C.byref(y), C.c_uint32(raw), C.c_uint8(skip),
            C.c_int16(sa_w), C.c_int16(sa_h)))
        return best.value, x.value, y.value

    def all_sad(self, src, src_stride, ref, ref_stride, mv, best8, best16, mv8, mv16, sub):
        best8, best16, mv8, mv16 = best8.copy(), best16.copy(), mv8.copy(), mv16.copy()
        e16 = np.zeros((16, 8), np.uint32)
        e8 = np.zeros((64, 8), np.uint32)
        self._fn("ext_all_sad_calculation_8x8_16x16")(*self._args(
            C.c_void_p(src.ctypes.data), C.c_uint32(src_stride), C.c_void_p(ref.ctypes.data), C.c_uint32(ref_stride),
            C.c_uint32(mv), C.c_void_p(best8.ctypes.data), C.c_void_p(best16.ctypes.data), C.c_void_p(mv8.ctypes.data),
            C.c_void_p(mv16.ctypes.data), C.c_void_p(e16.ctypes.data), C.c_void_p(e8.ctypes.data), C.c_uint8(sub)))
        return best8, best16, mv8, mv16, e16

    def eight_sad(self, sad16, best32, best64, mv32, mv64, mv):
        best32, best64, mv32, mv64 = best32.copy(), best64.copy(), mv32.copy(), mv64.copy()
        s32 = np.zeros((4, 8), np.uint32)
        sad16 = np.ascontiguousarray(sad16)
        self._fn("ext_eight_sad_calculation_32x32_64x64")(*self._args(
            C.c_void_p(sad16.ctypes.data), C.c_void_p(best32.ctypes.data), C.c_void_p(best64.ctypes.data),
            C.c_void_p(mv32.ctypes.data), C.c_void_p(mv64.ctypes.data), C.c_uint32(mv), C.c_void_p(s32.ctypes.data)))
        return best32, best64, mv32, mv64, s32

    def ext_sad(self, src, src_stride, ref, ref_stride, best8, best16, mv8, mv16, mv, sub):
        best8, best16, mv8, mv16 = best8.copy(), best16.copy(), mv8.copy(), mv16.copy()
        s16 = np.zeros(1, np.uint32)
        s8 = np.zeros(4, np.uint32)
        self._fn("ext_sad_calculation_8x8_16x16")(*self._args(
            C.c_void_p(src.ctypes.data), C.c_uint32(src_stride), C.c_void_p(ref.ctypes.data), C.c_uint32(ref_stride),
            C.c_void_p(best8.ctypes.data), C.c_void_p(best16.ctypes.data), C.c_void_p(mv8.ctypes.data),
            C.c_void_p(mv16.ctypes.data), C.c_uint32(mv), C.c_void_p(s16.ctypes.data), C.c_void_p(s8.ctypes.data), C.c_uint8(sub)))
        return best8, best16, mv8, mv16, s16, s8

    def ext_sad_32_64(self, sad16, best32, best64, mv32, mv64, mv):
        best32, best64, mv32, mv64 = best32.copy(), best64.copy(), mv32.copy(), mv64.copy()
        s32 = np.zeros(4, np.uint32)
        sad16 = np.ascontiguousarray(sad16)
        self._fn("ext_sad_calculation_32x32_64x64")(*self._args(
            C.c_void_p(sad16.ctypes.data), C.c_void_p(best32.ctypes.data), C.c_void_p(best64.ctypes.data),
            C.c_void_p(mv32.ctypes.data), C.c_void_p(mv64.ctypes.data), C.c_uint32(mv), C.c_void_p(s32.ctypes.data)))
        return best32, best64, mv32, mv64, s32

    def nxm(self, src, src_stride, ref, ref_stride, h, w):
        return self._fn("nxm_sad_kernel", C.c_uint32)(*self._args(
            C.c_void_p(src.ctypes.data), C.c_uint32(src_stride), C.c_void_p(ref.ctypes.data), C.c_uint32(ref_stride),
            C.c_uint32(h), C.c_uint32(w)))

    def init32(self, count128, count32, value):
        buf = np.full(count128 * 4 + count32 + 4, 0xDEADBEEF, np.uint32)
        self._fn("initialize_buffer_32bits")(*self._args(C.c_void_p(buf.ctypes.data), C.c_uint32(count128),
                                                         C.c_uint32(count32), C.c_uint32(value)))
        return buf

    def downsample(self, img, step):
        h, w = img.shape
        out = np.zeros((h // step + 1, w // step + 3), np.uint8)
        getattr(self.lib, self.prefix + "downsample_2d")(
            C.c_void_p(img.ctypes.data), C.c_uint32(img.strides[0]), C.c_uint32(w), C.c_uint32(h), C.c_void_p(out.ctypes.data),
            C.c_uint32(out.strides[0]), C.c_uint32(step))
        return out


def same(a, b, what):
    if isinstance(a, tuple):
        assert len(a) == len(b)
        for i, (x, y) in enumerate(zip(a, b)):
            same(x, y, "%s[%d]" % (what, i))
    elif isinstance(a, np.ndarray):
        np.testing.assert_array_equal(a, b, err_msg=what)
    else:
        assert a == b, (what, a, b)
