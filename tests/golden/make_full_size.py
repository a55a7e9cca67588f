#!/usr/bin/env python
"""Writes tests/golden/full_size_crc.json from the REFERENCE (oracle/_ref, its AVX2 / AVX-512 table: the C and SIMD
kernels give identical results, tests/test_oracle_vs_ref.py): per full-size case the CRC32 of every result array of the
picture, of its inputs, and the wall time the reference took. Development container only (needs /root/reference):

    python tests/golden/make_full_size.py [case ...]
"""
import json
import os
import sys
import time

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from oracle import pyoracle as po  # noqa: E402
import svtme_b200  # noqa: E402,F401
from util import RES_FIELDS  # noqa: E402
import full_size_cases as fs  # noqa: E402

OUT = os.path.join(HERE, "full_size_crc.json")


def main():
    assert po.ref_available(), "oracle/_ref missing: run `make -C oracle ref` where /root/reference exists"
    simd = 2 if "avx512bw" in open("/proc/cpuinfo").read() else 1
    out = json.load(open(OUT)) if os.path.exists(OUT) else {}
    for name in (sys.argv[1:] or list(fs.FULL_SIZE)):
        _, src, refs, pocs, poc, preset, kw = fs.inputs(name)
        pin = po.make_pic_in(src, refs, pocs, poc, preset, simd=simd, **kw)
        params = po.ref_derive_params(pin)
        t0 = time.time()
        res, _ = po.ref_me_picture_mt(pin, os.cpu_count() or 1, params)
        sec = time.time() - t0
        assert po.ref_lib().ref_rtcd_table_at_last_b64() == simd
        out[name] = {"input_crc": [fs.crc_of(p) for p in [src] + refs[0] + refs[1]], "result_crc": fs.result_crcs(res, RES_FIELDS),
                     "reference_seconds": round(sec, 2), "reference_threads": os.cpu_count() or 1, "reference_table": simd}
        print(name, "%.1f s" % sec, flush=True)
        with open(OUT, "w") as f:
            json.dump(out, f, indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
