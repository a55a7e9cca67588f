"""BASELINE.json's configurations at FULL size with the reference counts the benchmark runs (4 + 3), as parity cases whose
expected results come from the REFERENCE itself: tests/golden/make_full_size.py runs oracle/_ref (the reference's own
sources, SIMD table) on each picture and stores the CRC32 of every result array in tests/golden/full_size_crc.json; the
GPU test recomputes the picture through the C ABI and compares CRCs. The widest windows (<= 256 x 256 at 8K preset 2,
<= 512 x 512 at preset -2: enc_mode_config.c:228-355, motion_estimation.c:1296-1366) at the picture sizes that stress the
full-pel tile queue, the key packing and the unit-queue capacities are covered this way."""
import zlib

import numpy as np

from svt_av1_psy_b200.synth import PanningNoise

# name -> (w, h, bits, preset, hierarchical levels, temporal layer, poc, l0, l1, clip velocity)
FULL_SIZE = {
    "1080p_8bit_M4_4+3": (1920, 1080, 8, 4, 5, 2, 8, [4, 0, 6, 2], [16, 12, 10], (3, 2)),
    "2160p_10bit_M4_4+3": (3840, 2160, 10, 4, 5, 2, 8, [4, 0, 6, 2], [16, 12, 10], (5, 3)),
    "4320p_8bit_M2_4+3": (7680, 4320, 8, 2, 4, 2, 4, [2, 0, 3, 1], [8, 6, 5], (9, 4)),
    "1080p_8bit_M-2_4+3": (1920, 1080, 8, -2, 5, 2, 8, [4, 0, 6, 2], [16, 12, 10], (7, 3)),
}


def inputs(name):
    """-> (clip, src8, refs8, pocs, poc, preset, kw); src8 / refs8 are the 8-bit planes ME sees (the 8 MSBs of 10-bit input)"""
    w, h, bits, preset, hier, tl, poc, l0, l1, (vx, vy) = FULL_SIZE[name]
    clip = PanningNoise(w, h, seed=900 + preset, vx=vx, vy=vy, max_frames=20, sigma=2.0)
    src = clip.frame(poc)
    refs = [[clip.frame(p) for p in l0], [clip.frame(p) for p in l1]]
    kw = dict(temporal_layer=tl, hier_levels=hier, is_ref=1)
    return clip, src, refs, [l0, l1], poc, preset, kw


def crc_of(arr):
    return zlib.crc32(np.ascontiguousarray(arr).tobytes()) & 0xffffffff


def result_crcs(res, fields):
    return {f: crc_of(getattr(res, f)) for f in fields}
