"""not gpu: the tile plan of the 1/16-plane searches (k_hme_ctl makes it on the device with the same functions,
svt-av1-psy_b200/csrc/me_ctl.cuh s16_tile_geometry + me_search.cuh s16_eval_choice): tiles cover the window exactly, tile
sizes are the ones the TMA boxes hold, the evaluators are the ones measured best for the production shapes. Window shapes: the ones the bench workload produces (profiles/r02_summary.md) and a
sweep."""
import ctypes as C

import pytest

import svtme_b200 as me

GEN, VB2, VB3, POS = 0, 1, 2, 3


def plan(sa_w, sa_h, block16x8=1, level0=1):
    lib = me.lib()
    lib.svtme_s16_tile_plan.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_int)]
    lib.svtme_s16_tile_plan.restype = None
    out = (C.c_int * 7)()
    lib.svtme_s16_tile_plan(sa_w, sa_h, block16x8, level0, out)
    return dict(tw=out[0], th=out[1], ntx=out[2], nty=out[3], full=out[4], last=out[5], cost=out[6])


@pytest.mark.parametrize("sa_w,sa_h", [(8, 285), (8, 200), (8, 100), (384, 3), (192, 3), (96, 3), (64, 48), (32, 16), (32, 32), (96, 80),
                                       (96, 96), (32, 28), (32, 20), (16, 8), (1, 1), (200, 200), (480, 480), (17, 5)])
@pytest.mark.parametrize("level0", [0, 1])
def test_tiles_cover_the_window(sa_w, sa_h, level0):
    p = plan(sa_w, sa_h, 1, level0)
    if sa_w <= 16:
        assert (p["tw"], p["th"]) == (16, 96)  # the tall box: 16 positions x 96 rows
    else:
        assert p["tw"] == 128 and p["th"] in (18, 16, 12)  # the wide box holds 18 search rows
    assert (p["ntx"] - 1) * p["tw"] < sa_w <= p["ntx"] * p["tw"]
    assert (p["nty"] - 1) * p["th"] < sa_h <= p["nty"] * p["th"]
    assert p["ntx"] < 128 and p["nty"] < 256 and p["ntx"] * p["nty"] < 65536  # the cursor's bit fields (S16Cursor)
    assert p["full"] in (GEN, VB2, VB3, POS) and p["last"] in (GEN, VB2, VB3, POS)
    if not level0:
        assert VB2 not in (p["full"], p["last"])  # the pre-HME instantiation has no two-row evaluator


def test_other_block_shapes_keep_the_fixed_tiles():
    for sa_w, sa_h in [(64, 48), (32, 32), (128, 128)]:
        p = plan(sa_w, sa_h, 0, 1)  # the detector's 16x16 full-SAD blocks
        assert (p["tw"], p["th"]) == (128, 18)
        assert p["full"] == GEN and p["last"] == GEN


def test_the_production_shapes_get_the_evaluators_measured_best():
    assert plan(8, 285, 1, 0)["full"] == VB3        # tall pre-HME strip: 32 (octet, 3-row) tasks per 96-row tile
    assert plan(8, 100, 1, 0)["last"] == POS        # its 4-row remainder: one position per lane
    assert plan(384, 3, 1, 0)["full"] == GEN        # wide pre-HME strip
    p = plan(64, 48, 1, 1)
    assert p["th"] == 12 and p["full"] == VB3        # four tiles of 12 rows: one full pass each
    p = plan(32, 16, 1, 1)
    assert p["nty"] == 1 and p["full"] == VB2        # 32 (octet, 2-row) tasks
    p = plan(96, 80, 1, 1)
    assert p["th"] == 16 and p["full"] == VB2        # five tiles of 16 rows, three full passes each
