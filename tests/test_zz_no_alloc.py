"""GPU test of svtme_sad_loop_search_noalloc, the entry point for callers that may not allocate
device memory (real-time encoder threads, plug-ins inside a host that owns the GPU memory pool, code
inside a stream capture). It must return exactly what the allocating entry point
svtme_sad_loop_kernel returns for the same search. The new entry point keeps its scratch in one
array inside the library, which every call reuses, so the check is made on two calls in a row: a large
search first, then a small one that would pick up stale slots of the first if they leaked.

Collected last on purpose (file name): a failure here cannot hide the result of any other test."""
import ctypes as C

import numpy as np
import pytest

import kat_cases as kc
from oracle import pyoracle as po

pytestmark = pytest.mark.gpu


def host_search(lib, src, ss, ref, rs, bh, bw, raw, skip, sw, sh):
    best, x, y = C.c_uint64(1), C.c_int16(-7), C.c_int16(9)
    lib.svtme_sad_loop_kernel(C.c_void_p(src.ctypes.data), C.c_uint32(ss), C.c_void_p(ref.ctypes.data), C.c_uint32(rs),
                              C.c_uint32(bh), C.c_uint32(bw), C.byref(best), C.byref(x), C.byref(y), C.c_uint32(raw),
                              C.c_uint8(skip), C.c_int16(sw), C.c_int16(sh))
    return best.value, x.value, y.value


def noalloc_search(lib, torch, src, ss, ref, rs, bh, bw, raw, skip, sw, sh, pad=0, poison=False):
    """inputs and outputs live in device memory owned by the caller (here: torch tensors). pad = 0: the buffers begin and
    end exactly where the caller's data does; poison: bytes around them are 0xFF, so a read outside the bytes
    svt_sad_loop_kernel touches would change a SAD"""
    d_src = torch.full((src.size + 2 * pad,), 255 if poison else 0, dtype=torch.uint8, device="cuda")
    d_ref = torch.full((ref.size + 2 * pad,), 255 if poison else 0, dtype=torch.uint8, device="cuda")
    d_src[pad:pad + src.size] = torch.from_numpy(src.reshape(-1)).cuda()
    d_ref[pad:pad + ref.size] = torch.from_numpy(ref.reshape(-1)).cuda()
    out_sad = torch.tensor([1], dtype=torch.int64, device="cuda")
    out_xy = torch.tensor([-7, 9], dtype=torch.int16, device="cuda")
    torch.cuda.synchronize()
    rc = lib.svtme_sad_loop_search_noalloc(
        C.c_void_p(d_src.data_ptr() + pad), C.c_uint32(ss), C.c_void_p(d_ref.data_ptr() + pad), C.c_uint32(rs),
        C.c_uint32(bh), C.c_uint32(bw), C.c_void_p(out_sad.data_ptr()), C.c_void_p(out_xy.data_ptr()),
        C.c_void_p(out_xy.data_ptr() + 2), C.c_uint32(raw), C.c_uint8(skip), C.c_int16(sw), C.c_int16(sh), C.c_void_p(0))
    torch.cuda.synchronize()
    xy = out_xy.cpu().numpy()
    return rc, (int(out_sad.item()), int(xy[0]), int(xy[1]))


def test_noalloc_search_matches_allocating_entry_point_twice_in_a_row():
    torch = pytest.importorskip("torch")
    import svtme_b200 as me
    lib = me.lib()
    lib.svtme_sad_loop_search_noalloc.restype = C.c_uint32
    # (block w, h, search area w, h, skip_search_line, pattern, doubled strides)
    calls = [
        (16, 8, 640, 400, 0, "RANDOM", True),    # 13 bands of 32 rows: fills many scratch slots
        (16, 8, 15, 6, 0, "RANDOM", False),      # 1 band: must not see the 12 stale slots
        (16, 16, 96, 80, 1, "UNALIGN", False),   # skip_search_line across band boundaries
        (8, 8, 48, 40, 0, "REF_MAX", False),     # every position ties: the first one in raster order wins
        (12, 5, 33, 65, 0, "RANDOM", False),     # block narrower than 16, odd sizes
    ]
    for i, (bw, bh, sw, sh, skip, pat, doubled) in enumerate(calls):
        src, ss, ref, rs, raw = kc.sad_loop_buffers(bw, bh, sw, sh, pat, 4242 + i, doubled)
        want = host_search(lib, src, ss, ref, rs, bh, bw, raw, skip, sw, sh)
        rc, got = noalloc_search(lib, torch, src, ss, ref, rs, bh, bw, raw, skip, sw, sh)
        assert rc == 0
        assert got == want, (i, got, want)  # bit-exact: integer SADs and positions
        # ... and what the oracle (the plain-C restatement of svt_sad_loop_kernel) says
        best, x, y = C.c_uint64(1), C.c_int16(-7), C.c_int16(9)
        po.oracle_lib().orc_sad_loop_kernel(C.c_void_p(src.ctypes.data), C.c_uint32(ss), C.c_void_p(ref.ctypes.data), C.c_uint32(rs),
                                            C.c_uint32(bh), C.c_uint32(bw), C.byref(best), C.byref(x), C.byref(y), C.c_uint32(raw),
                                            C.c_uint8(skip), C.c_int16(sw), C.c_int16(sh))
        assert got == (best.value, x.value, y.value), (i, got)
    # an empty search leaves the centres alone, like the allocating entry point does
    src, ss, ref, rs, raw = kc.sad_loop_buffers(16, 8, 8, 8, "RANDOM", 1)
    rc, got = noalloc_search(lib, torch, src, ss, ref, rs, 8, 16, raw, 0, 0, 4)
    assert rc == 0 and got == (0xffffff, -7, 9)


def test_noalloc_search_stays_inside_the_callers_buffers():
    """The source and reference buffers end exactly where the last byte svt_sad_loop_kernel reads lies (a plane that ends
    flush with its allocation), first with nothing around them, then with 0xFF around them: the result must not move."""
    torch = pytest.importorskip("torch")
    import svtme_b200 as me
    lib = me.lib()
    lib.svtme_sad_loop_search_noalloc.restype = C.c_uint32
    for i, (bw, bh, sw, sh, skip) in enumerate([(16, 8, 40, 33, 0), (13, 7, 9, 5, 0), (16, 16, 64, 64, 1), (4, 4, 3, 2, 0)]):
        src, ss, ref, rs, raw = kc.sad_loop_buffers(bw, bh, sw, sh, "RANDOM", 777 + i)
        # cut the buffers to the touched bytes: (bh-1) rows + bw bytes, (sh-1 + bh-1) rows + sw+bw-1 bytes
        src_c = np.ascontiguousarray(src.reshape(-1)[:(bh - 1) * ss + bw])
        ref_c = np.ascontiguousarray(ref.reshape(-1)[:(sh - 1) * raw + (bh - 1) * rs + sw + bw - 1])
        want = host_search(lib, src, ss, ref, rs, bh, bw, raw, skip, sw, sh)
        for pad, poison in ((0, False), (64, True), (3, True)):
            rc, got = noalloc_search(lib, torch, src_c, ss, ref_c, rs, bh, bw, raw, skip, sw, sh, pad=pad, poison=poison)
            assert rc == 0 and got == want, (i, pad, got, want)


def test_noalloc_search_refuses_what_its_scratch_cannot_hold():
    torch = pytest.importorskip("torch")
    import svtme_b200 as me
    from svt_av1_psy_b200 import abi
    lib = me.lib()
    lib.svtme_sad_loop_search_noalloc.restype = C.c_uint32
    src, ss, ref, rs, raw = kc.sad_loop_buffers(16, 8, 16, 8, "RANDOM", 2)
    rc, _ = noalloc_search(lib, torch, src, ss, ref, rs, 8, 16, raw, 0, 1025, 4)  # wider than SVTME_NOALLOC_MAX_WIDTH
    assert rc == abi.ERR_BAD_PARAMETER
    rc, _ = noalloc_search(lib, torch, src, ss, ref, rs, 24, 16, raw, 0, 8, 8)  # block taller than 16 rows
    assert rc == abi.ERR_BAD_PARAMETER
