"""Input generators for the RTCD-kernel known-answer tests. They follow the parameter space of the
reference's own C-vs-SIMD unit tests (test/SadTest.cc:62-110 block sizes and patterns, :507-518 search
areas, :266-294 SAD-buffer patterns; test/PictureOperatorTest.cc:224-335 decimation sizes); the
cross products are thinned so that the CPU oracle finishes in seconds."""
import numpy as np

BLOCK_SIZES = [(4, 10), (4, 12), (8, 10), (8, 12), (24, 10), (40, 10), (48, 10), (56, 10), (24, 14), (40, 14), (48, 14),
               (56, 14), (16, 10), (16, 5), (32, 10), (32, 20), (64, 20), (64, 64), (64, 32), (32, 64), (32, 32), (32, 16),
               (16, 32), (16, 16), (16, 8), (8, 16), (8, 8), (8, 4), (4, 4), (4, 8), (4, 16), (16, 4), (8, 32), (32, 8),
               (16, 64), (64, 16), (24, 24), (24, 16), (16, 24), (24, 8), (8, 24), (64, 24), (48, 24), (32, 24), (24, 32),
               (48, 48), (48, 16), (48, 32), (16, 48), (32, 48), (48, 64), (64, 48), (56, 32), (40, 32)]
BLOCK_SIZES_SMALL = [(6, 2), (6, 16), (12, 5), (12, 33), (31, 1), (31, 8), (15, 10), (7, 11), (5, 10), (17, 10), (19, 11),
                     (16, 17), (16, 33), (39, 1), (3, 40), (43, 5), (55, 3), (37, 37), (63, 21), (63, 33), (4, 2), (4, 33),
                     (8, 15), (8, 31), (16, 2), (16, 15), (24, 3), (24, 31), (32, 9), (32, 33)]
LOOP_AREAS = [(8, 15), (16, 31), (12, 31), (64, 125), (192, 75), (128, 50), (64, 25), (240, 200), (144, 120), (96, 80),
              (48, 40), (240, 120), (144, 72), (96, 48), (48, 24), (560, 320), (336, 192), (224, 128), (112, 64), (640, 400),
              (384, 240), (256, 160), (128, 80), (480, 120), (288, 72), (192, 48), (96, 24), (160, 60), (96, 36), (64, 24),
              (32, 12), (15, 6)]
PATTERNS = ["REF_MAX", "SRC_MAX", "RANDOM", "UNALIGN"]


def sad_loop_cases():
    """(bw, bh, sa_w, sa_h, skip, pattern): every block size on two small areas, every area on the
    production block shapes (16x8 sub-sampled HME-L0, 32x16 L1, 64x32 L2) capped for oracle time."""
    cases = []
    for i, (bw, bh) in enumerate(BLOCK_SIZES + BLOCK_SIZES_SMALL):
        for (sw, sh) in ((15, 6), (32, 12)):
            cases.append((bw, bh, sw, sh, i & 1, PATTERNS[i % 4]))
    for i, (sw, sh) in enumerate(LOOP_AREAS):
        cases.append((16, 8, sw, sh, i & 1, PATTERNS[i % 4]))
        if sw * sh <= 160 * 60:
            cases.append((32, 16, sw, sh, 0, PATTERNS[(i + 1) % 4]))
        if sw * sh <= 96 * 48:
            cases.append((64, 32, sw, sh, 0, PATTERNS[(i + 2) % 4]))
    return cases


def sad_loop_buffers(bw, bh, sa_w, sa_h, pattern, seed, doubled=False):
    """-> src (2D view), src_stride, ref (2D), ref_stride; doubled: strides as the SUB_SAD callers pass them"""
    rng = np.random.default_rng(seed)
    step = 2 if doubled else 1
    src_stride = 128 - (1 if pattern == "UNALIGN" else 0)
    ref_stride_raw = sa_w + bw + 9 + (1 if pattern == "UNALIGN" else 0)
    src = np.zeros(((bh * step) + 1, src_stride), np.uint8)
    ref = np.zeros((sa_h + bh * step + 1, ref_stride_raw), np.uint8)
    if pattern == "REF_MAX":
        ref[:] = 255
    elif pattern == "SRC_MAX":
        src[:] = 255
    else:
        src[:] = rng.integers(0, 256, src.shape, dtype=np.uint8)
        ref[:] = rng.integers(0, 256, ref.shape, dtype=np.uint8)
        # plant an exact copy so that several positions tie at SAD 0 is NOT the case, but one does exist
        py, px = sa_h // 2, sa_w // 3
        ref[py:py + bh * step:step, px:px + bw] = src[0:bh * step:step, :bw]
    return src, src_stride * step, ref, ref_stride_raw * step, ref_stride_raw


def me_block_buffers(pattern, seed, ref_extra=8):
    """64x64 source and a (64+ref_extra-1)-wide reference for the all-SAD kernels"""
    rng = np.random.default_rng(seed)
    src_stride = 64 + (0 if pattern != "UNALIGN" else 3)
    ref_stride = 64 + ref_extra + 16 + (0 if pattern != "UNALIGN" else 1)
    src = np.zeros((64, src_stride), np.uint8)
    ref = np.zeros((64, ref_stride), np.uint8)
    if pattern == "REF_MAX":
        ref[:] = 255
    elif pattern == "SRC_MAX":
        src[:] = 255
    else:
        src[:] = rng.integers(0, 256, src.shape, dtype=np.uint8)
        ref[:] = rng.integers(0, 256, ref.shape, dtype=np.uint8)
        ref[:, 3:67] = (src[:, :64].astype(np.int16) + rng.integers(-3, 4, (64, 64))).clip(0, 255).astype(np.uint8)
    return src, src_stride, ref, ref_stride


def sad_buffer(pattern, shape, seed):
    """SADPattern BUF_MAX / BUF_MIN / BUF_SMALL / BUF_RANDOM (test/SadTest.cc:266-294)"""
    rng = np.random.default_rng(seed)
    if pattern == "BUF_MAX":
        return np.full(shape, 0xFFFFFFFF, np.uint32)
    if pattern == "BUF_MIN":
        return np.zeros(shape, np.uint32)
    if pattern == "BUF_SMALL":
        return rng.integers(0, 1 << 12, shape, dtype=np.uint32)
    return rng.integers(0, 1 << 32, shape, dtype=np.uint64).astype(np.uint32)


DOWNSAMPLE_SIZES = [(1920, 1080), (960, 540), (176, 144), (88, 72), (64, 64), (16, 16)]
