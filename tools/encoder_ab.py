#!/usr/bin/env python
"""Informational: wall time of the reference encoder (C-only build of oracle/encoder.mk) with its own CPU motion
estimation against the same encoder re-linked with the GPU library (shim/svtme_wrap.c), same clip, same arguments.
Not a bench.py number: the C-only build has no SIMD kernels, so the CPU side is slower than a production build.
usage: python tools/encoder_ab.py [W H FRAMES PRESET LP]"""
import os
import re
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from test_encoder_dropin import GPU_APP, REF_APP, write_y4m  # noqa: E402


def main():
    w, h, n, preset, lp = (int(v) for v in (sys.argv[1:6] if len(sys.argv) >= 6 else (1920, 1080, 33, 8, 8)))
    with tempfile.TemporaryDirectory() as d:
        y4m = os.path.join(d, "in.y4m")
        write_y4m(y4m, w, h, n, seed=3)
        out = {}
        for name, app in (("ref", REF_APP), ("gpu", GPU_APP)):
            ivf = os.path.join(d, name + ".ivf")
            env = dict(os.environ, SVTME_SHIM_STATS="1")
            t0 = time.time()
            r = subprocess.run([app, "-i", y4m, "-b", ivf, "--preset", str(preset), "--lp", str(lp)], capture_output=True, text=True, env=env)
            dt = time.time() - t0
            assert r.returncode == 0, r.stderr[-2000:]
            enc = re.search(r"Total Encoding Time:\s+(\d+) ms", r.stdout + r.stderr)
            shim = re.search(r"svtme_wrap: .*", r.stderr)
            out[name] = (dt, int(enc.group(1)) if enc else None, open(ivf, "rb").read(), shim.group(0) if shim else "")
        same = out["ref"][2] == out["gpu"][2]
        print("%dx%d %d frames preset %d lp %d: CPU-ME encoder %.1f s (encode %s ms), GPU-ME encoder %.1f s (encode %s ms), bitstreams %s (%d bytes)"
              % (w, h, n, preset, lp, out["ref"][0], out["ref"][1], out["gpu"][0], out["gpu"][1], "identical" if same else "DIFFER", len(out["ref"][2])))
        print(out["gpu"][3])


if __name__ == "__main__":
    main()
