#!/usr/bin/env python
"""Generates the committed fixtures of tests/golden/ from the REFERENCE itself (oracle/_ref, built by
oracle/Makefile from /root/reference). Run in the development container only:

    python tests/golden/make_golden.py

  sig_deriv_params.json.gz  svt_aom_sig_deriv_me outputs (enc_mode_config.c:694-843) for a grid of
                          (enc_mode, width, height, temporal layer, sc_class1, pred structure, qp)
  me_<case>.npz           svt_aom_motion_estimation_b64 outputs of every b64 of one picture (MeSbResults,
                          PCS statistics, MeContext bests, HME state) for the cases of tests/util.py
  kat_rtcd.npz            outputs of the reference's C RTCD kernels for a thinned list of kat_cases.py
  tf_me.npz               temporal-filter ME (ME_MCTF) controls and results of the cases of test_tf_me.py
  dg_detector.json        DGDetectorMetrics of dg_detector_hme_level0 (pd_process.c:490) for the pairs of test_dg_detector.py

Inputs are NOT stored: they are regenerated from their seeds (numpy Generator streams are stable); the
CRC32 of every input picture is stored so that a drifting generator fails loudly instead of silently.
"""
import gzip
import json
import os
import sys
import zlib

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from oracle import pyoracle as po  # noqa: E402
import svtme_b200  # noqa: E402,F401
from util import CASES, EXTRA, RES_FIELDS, case_id, case_inputs, extra_inputs  # noqa: E402
import golden_lib as gl  # noqa: E402


def sig_deriv_fixture():
    out = {}
    for key in gl.sig_deriv_grid():
        enc_mode, w, h, tl, sc, ps, qp = key
        pin = gl.sig_deriv_pic_in(po, enc_mode, w, h, tl, sc, ps, qp)
        out[",".join(str(v) for v in key)] = gl.fields_of(po.ref_derive_params(pin))
    with gzip.GzipFile(os.path.join(HERE, "sig_deriv_params.json.gz"), "wb", mtime=0) as f:
        f.write(json.dumps(out, separators=(",", ":"), sort_keys=True).encode())
    print("sig_deriv_params.json.gz:", len(out), "combinations")


def picture_fixture(name, src, refs, pocs, poc, enc_mode, kw):
    pin = po.make_pic_in(src, refs, pocs, poc, enc_mode, **kw)
    params = po.ref_derive_params(pin)
    res, dbg, _ = po.ref_me_picture(pin, params, debug=True)
    arrays = {"params": np.frombuffer(gl.struct_bytes(params), np.uint8).copy(), "debug": dbg,
              "input_crc": np.array([zlib.crc32(np.ascontiguousarray(p).tobytes()) for p in [src] + refs[0] + (refs[1] if len(refs) > 1 else [])],
                                    np.uint32)}
    for f in RES_FIELDS:
        arrays["res_" + f] = getattr(res, f)
    np.savez_compressed(os.path.join(HERE, "me_%s.npz" % name), **arrays)
    print("me_%s.npz" % name)


def kat_fixture():
    """replays the runners of tests/test_kat_kernels.py with the reference's C kernels as the recorded peer"""
    import test_kat_kernels as tk
    from kat_runner import Impl
    out = {}
    for runner in tk.RUNNERS:
        rec = tk.GOLDEN_CTX["peer"] = gl.Recorder(Impl(po.ref_lib(), "ref_", 0))
        runner("golden")  # candidate = oracle, compared against the reference while recording
        out.update(rec.arrays(runner.__name__))
    np.savez_compressed(os.path.join(HERE, "kat_rtcd.npz"), **out)
    print("kat_rtcd.npz:", len(out), "arrays")


def dg_fixture():
    import test_dg_detector as td
    out = {}
    for case in td.CASES:
        src, ref = td.pair(case)
        m, _, _ = po.dg_detector(po.ref_lib(), "ref_", src, ref)
        out[td.key(case)] = list(m.as_tuple())
    with open(os.path.join(HERE, "dg_detector.json"), "w") as f:
        json.dump(out, f, indent=1, sort_keys=True)
    print("dg_detector.json:", len(out), "pairs")


def tf_fixture():
    import test_tf_me as tt
    out = {}
    for case in tt.CASES:
        w, h, m, lvl, qo, th, qp, cp, rp = case
        centre, ref, _ = tt.inputs(case)
        prm, res, _ = po.ref_tf_me_picture(centre, ref, m, lvl, qp_opt=qo, me_exit_th=th, qp=qp, centre_poc=cp, ref_poc=rp)
        out[tt.key(case) + "/params"] = np.frombuffer(gl.struct_bytes(prm), np.uint8).copy()
        for f in tt.TF_FIELDS:
            out[tt.key(case) + "/" + f] = getattr(res, f)
    np.savez_compressed(os.path.join(HERE, "tf_me.npz"), **out)
    print("tf_me.npz:", len(tt.CASES), "cases")


def main():
    if not po.ref_available():
        po.build(ref=True)
    assert po.ref_available(), "oracle/_ref missing: run `make -C oracle ref` where /root/reference exists"
    sig_deriv_fixture()
    for case in CASES:
        picture_fixture(case_id(case), *case_inputs(case))
    for name, geo, kw in EXTRA:
        picture_fixture(name, *extra_inputs(geo, kw))
    kat_fixture()
    dg_fixture()
    tf_fixture()


if __name__ == "__main__":
    main()
