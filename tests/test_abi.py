"""The C ABI without a GPU: the library loads, exports every function include/svtme_b200.h and
svt-av1-psy_b200/host/me_sig_deriv.h declare, the ctypes mirror of the structs has the sizes the C compiler gave
them, the header is valid C, and the product path fails loudly (no CPU fallback) where there is no CUDA device."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

import svtme_b200 as me
from svt_av1_psy_b200 import abi, binding

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "svtme_b200.h")
HOST_HEADER = os.path.join(ROOT, "svt-av1-psy_b200", "host", "me_sig_deriv.h")


def declared_functions(path):
    text = open(path).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)  # comments mention functions too
    return sorted(set(re.findall(r"\b(svtme_[a-z0-9_]+)\s*\(", text)))


def test_every_declared_entry_point_is_exported():
    names = declared_functions(HEADER)
    assert len(names) >= 35, names
    lib = me.lib()
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, "declared in include/svtme_b200.h but not exported by libsvtme_b200.so: %s" % missing
    for must in ("svtme_sad_loop_kernel", "svtme_ext_all_sad_calculation_8x8_16x16", "svtme_ext_eight_sad_calculation_32x32_64x64",
                 "svtme_picture_me_batch", "svtme_dg_detector_hme", "svtme_sad_loop_search_noalloc"):
        assert must in names


def test_host_library_exports_its_header():
    names = declared_functions(HOST_HEADER)
    assert {"svtme_sig_deriv_me", "svtme_sig_deriv_me_tf", "svtme_shard_owner"} <= set(names)
    lib = me.host_lib()
    assert not [n for n in names if not hasattr(lib, n)]


def test_struct_sizes_match_the_compiler():
    lib = me.lib()
    lib.svtme_abi_sizeof.argtypes = [C.c_int]
    lib.svtme_abi_sizeof.restype = C.c_uint32
    for what, ty in enumerate((abi.MeCtrls, abi.MePicParams, abi.MeSearchResult, abi.MeWindow, abi.MePicResults)):
        assert lib.svtme_abi_sizeof(what) == C.sizeof(ty), ty.__name__
    assert lib.svtme_abi_sizeof(5) == abi.HME_STATE_DTYPE.itemsize
    assert C.sizeof(abi.MeCtrls) == 128 and C.sizeof(abi.MePicParams) == 248  # the fixtures store these bytes
    assert C.sizeof(abi.DgMetrics) == 24


def test_header_is_valid_c(tmp_path):
    src = tmp_path / "use_header.c"
    src.write_text('#include "svtme_b200.h"\n#include "me_sig_deriv.h"\n'
                   "int main(void) { SvtMePicParams p; SvtMeDgMetrics m; (void)p; (void)m; return (int)sizeof(SvtMeCtrls) - 128; }\n")
    out = subprocess.run(["gcc", "-std=c99", "-Wall", "-Werror", "-pedantic", "-I", os.path.dirname(HEADER), "-I", os.path.dirname(HOST_HEADER),
                          "-o", str(tmp_path / "use_header"), str(src)], capture_output=True, text=True)
    assert out.returncode == 0, out.stderr
    assert subprocess.run([str(tmp_path / "use_header")]).returncode == 0


def test_version_and_error_codes():
    lib = me.lib()
    lib.svtme_version.restype = C.c_char_p
    assert b"sm_100a" in lib.svtme_version()
    assert abi.OK == 0 and abi.ERR_BAD_PARAMETER == 0x80001005  # EB_ErrorNone / EB_ErrorBadParameter (EbSvtAv1.h:127-142)


def test_no_cpu_fallback_without_a_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    with pytest.raises(binding.SvtMeError):
        me.Session(64, 64, max_pictures=2)
    # the RTCD-signature entry points report the failure too instead of computing on the host
    src = np.zeros((16, 16), np.uint8)
    best = C.c_uint64(0)
    x, y = C.c_int16(0), C.c_int16(0)
    lib = me.lib()
    lib.svtme_last_error.restype = C.c_uint32
    lib.svtme_sad_loop_kernel(C.c_void_p(src.ctypes.data), C.c_uint32(16), C.c_void_p(src.ctypes.data), C.c_uint32(16), C.c_uint32(8),
                              C.c_uint32(8), C.byref(best), C.byref(x), C.byref(y), C.c_uint32(16), C.c_uint8(0), C.c_int16(2), C.c_int16(2))
    assert lib.svtme_last_error() != abi.OK
