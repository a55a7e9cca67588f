"""B200-native open-loop motion estimation for SVT-AV1-PSY — Python side of the C ABI.

The product is csrc/ (CUDA kernels + the extern "C" library, include/svtme_b200.h) and host/
(C host logic). This module only binds the library with ctypes for tests and bench.py.
There is no CPU fallback: if the CUDA library is missing, lib() raises."""
import os

from . import abi  # noqa: F401

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(PKG_DIR, "libsvtme_b200.so")
HOST_LIB_PATH = os.path.join(PKG_DIR, "libsvtme_host.so")

from .binding import lib, host_lib, Session, sig_deriv_me, last_error  # noqa: E402,F401
