"""-m gpu: BASELINE.json's configurations at full size, 4 + 3 references, against the REFERENCE's own results
(tests/golden/full_size_crc.json, written by tests/golden/make_full_size.py from oracle/_ref): CRC32 of every result array
of the picture. CPU part: the fixture is complete and the inputs regenerate to the recorded CRCs."""
import copy
import json
import os

import numpy as np
import pytest

import svtme_b200 as me
import full_size_cases as fs
from util import RES_FIELDS, params_from_encoder_inputs

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "full_size_crc.json")


def golden():
    if not os.path.exists(GOLDEN):
        pytest.skip("tests/golden/full_size_crc.json not generated")
    return json.load(open(GOLDEN))


def test_fixture_covers_every_case():
    g = golden()
    assert set(g) == set(fs.FULL_SIZE)
    for name, rec in g.items():
        assert set(rec["result_crc"]) == set(RES_FIELDS) and rec["reference_table"] in (1, 2)


def test_inputs_regenerate(orc):
    g = golden()
    name = "1080p_8bit_M4_4+3"
    _, src, refs, _, _, _, _ = fs.inputs(name)
    assert [fs.crc_of(p) for p in [src] + refs[0] + refs[1]] == g[name]["input_crc"]


@pytest.mark.gpu
@pytest.mark.parametrize("name", list(fs.FULL_SIZE))
def test_full_size_matches_reference(name):
    g = golden()[name]
    w, h, bits = fs.FULL_SIZE[name][:3]
    clip, src, refs, pocs, poc, preset, kw = fs.inputs(name)
    assert [fs.crc_of(p) for p in [src] + refs[0] + refs[1]] == g["input_crc"], "the input generator drifted"
    prm = params_from_encoder_inputs(src, refs, pocs, poc, preset, **kw)
    s = me.Session(w, h, max_pictures=8)
    planes = [(0, poc)] + [(1 + i, p) for i, p in enumerate(pocs[0] + pocs[1])]
    for slot, p in planes:
        s.upload(slot, clip.frame10(p) if bits == 10 else clip.frame(p))  # 10-bit: the device keeps the 8 MSBs
    p = copy.copy(prm)
    pid = 1
    for l in range(2):
        for r in range(len(pocs[l])):
            p.ref_pic_id[l][r] = pid
            pid += 1
    got = s.me(0, p)
    again = s.me(0, p)
    s.close()
    crc = fs.result_crcs(got, RES_FIELDS)
    bad = [f for f in RES_FIELDS if crc[f] != g["result_crc"][f]]
    assert not bad, "%s: arrays that differ from the reference's: %s" % (name, bad)
    assert fs.result_crcs(again, RES_FIELDS) == crc  # and run to run
