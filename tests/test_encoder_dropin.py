"""Whole-encoder drop-in check: the reference ENCODER, re-linked with the GPU library, must write the same bitstream.

oracle/encoder.mk compiles the reference's encoder library and command-line application from their own sources
(C only, plain gcc, no cmake) and links them twice: SvtAv1EncApp_ref (unmodified) and SvtAv1EncApp_gpu
(-Wl,--wrap=svt_aom_motion_estimation_b64 -Wl,--wrap=dg_detector_hme_level0 + svt-av1-psy_b200/shim/svtme_wrap.c +
libsvtme_b200.so). In the second one every open-loop ME picture, every temporal-filter ME pair and every dynamic-GOP
detector search of the encode runs on the GPU; everything downstream of motion estimation (TPL, rate control, mode
decision, entropy coding) consumes those results, so an identical .ivf means the stage is a drop-in for the whole
encoder, not only for the arrays the parity tests look at. This is the reference's own end-to-end criterion (its CI
compares bitstreams, .gitlab/workflows/linux/.gitlab-ci.yml:338-391).

The binaries are built in the development container (they need the reference's sources) and travel to the GPU box as
prebuilt files; nothing here reads /root/reference."""
import os
import re
import subprocess

import numpy as np
import pytest

import svtme_b200  # noqa: F401
from svt_av1_psy_b200.synth import PanningNoise

ENC_DIR = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "_ref", "enc")
REF_APP, GPU_APP = os.path.join(ENC_DIR, "SvtAv1EncApp_ref"), os.path.join(ENC_DIR, "SvtAv1EncApp_gpu")
BATCHED_APP = os.path.join(ENC_DIR, "SvtAv1EncApp_gpu_batched")  # + oracle/pd_process_submit.patch: mini-GOPs submitted as GPU batches

pytestmark = pytest.mark.skipif(not (os.path.exists(REF_APP) and os.path.exists(GPU_APP)),
                                reason="oracle/_ref/enc not built (make -f oracle/encoder.mk, needs the reference's sources)")

# (name, width, height, frames, extra encoder arguments)
ENCODES = [
    ("M8_default", 320, 192, 40, ["--preset", "8"]),                       # temporal filter + dynamic-GOP detector on (defaults)
    ("M4_lp4", 320, 192, 34, ["--preset", "4", "--lp", "4"]),              # pre-HME, 4+3 references, several ME threads
    ("M10_360p", 640, 360, 24, ["--preset", "10"]),
    ("M7_low_delay", 320, 192, 24, ["--preset", "7", "--pred-struct", "1", "--tune", "1"]),
    ("M9_low_delay_cbr", 320, 192, 24, ["--preset", "9", "--pred-struct", "1", "--tune", "1", "--rc", "2", "--tbr", "500"]),
    ("M8_screen_content", 320, 192, 20, ["--preset", "8", "--scm", "1"]),
    ("M5_720p", 1280, 720, 18, ["--preset", "5", "--lp", "6"]),
    ("M2_odd_size", 200, 136, 20, ["--preset", "2"]),                       # width / height not multiples of 64
    ("M8_10bit", 320, 192, 20, ["--preset", "8"]),                         # 10-bit input: ME reads the 8 MSBs
    ("M12_no_8x8", 640, 360, 20, ["--preset", "12"]),                       # fastest presets: 8x8 ME off at larger sizes (21 PUs per b64)
    ("M0_gm_mrp", 192, 128, 12, ["--preset", "0"]),                         # global-motion flags, widest windows
    ("M6_three_layers", 320, 192, 20, ["--preset", "6", "--hierarchical-levels", "3"]),
    ("M8_no_tf", 320, 192, 20, ["--preset", "8", "--enable-tf", "0"]),
    ("M13_1080p", 1920, 1080, 6, ["--preset", "13", "--lp", "8"]),
]


def write_y4m(path, w, h, n, seed, ten_bit=False):
    clip = PanningNoise(w, h, seed=seed, vx=2, vy=1, max_frames=n)
    with open(path, "wb") as f:
        f.write(b"YUV4MPEG2 W%d H%d F30:1 Ip A1:1 %s\n" % (w, h, b"C420p10" if ten_bit else b"C420jpeg"))
        for i in range(n):
            y = clip.frame(i)
            f.write(b"FRAME\n")
            if ten_bit:
                rng = np.random.default_rng(seed * 1000 + i)
                y16 = (y.astype(np.uint16) << 2) | rng.integers(0, 4, y.shape, dtype=np.uint16)
                f.write(y16.astype("<u2").tobytes())
                f.write(np.full((h // 2) * (w // 2) * 2, 512, "<u2").tobytes())
            else:
                f.write(y.tobytes())
                f.write(np.full((h // 2) * (w // 2) * 2, 128, np.uint8).tobytes())


def encode(app, y4m, out, args, env=None):
    e = dict(os.environ)
    e.update(env or {})
    r = subprocess.run([app, "-i", y4m, "-b", out] + args, capture_output=True, text=True, env=e, timeout=900)
    assert r.returncode == 0, "%s failed:\n%s\n%s" % (os.path.basename(app), r.stdout[-2000:], r.stderr[-2000:])
    return r.stderr


def test_binaries_are_linked_as_described():
    """CPU: the GPU encoder's two call sites of the b64 function and the detector's caller reach the shim"""
    dis = subprocess.run(["objdump", "-d", "--no-show-raw-insn", GPU_APP], capture_output=True, text=True).stdout
    assert len(re.findall(r"call\s+\w+ <__wrap_svt_aom_motion_estimation_b64>", dis)) == 2  # me_process.c:266, temporal_filtering.c:3141
    assert len(re.findall(r"call\s+\w+ <__wrap_dg_detector_hme_level0>", dis)) == 1         # me_process.c:329
    ref = subprocess.run(["objdump", "-d", "--no-show-raw-insn", REF_APP], capture_output=True, text=True).stdout
    assert "__wrap_" not in ref and "svtme_" not in ref


@pytest.mark.gpu
@pytest.mark.parametrize("name,w,h,n,args", ENCODES, ids=[e[0] for e in ENCODES])
def test_gpu_encoder_writes_the_reference_bitstream(tmp_path, name, w, h, n, args):
    y4m = str(tmp_path / "in.y4m")
    write_y4m(y4m, w, h, n, seed=len(name) + w, ten_bit=name.endswith("10bit"))
    ref_ivf, gpu_ivf = str(tmp_path / "ref.ivf"), str(tmp_path / "gpu.ivf")
    encode(REF_APP, y4m, ref_ivf, args)
    log = encode(GPU_APP, y4m, gpu_ivf, args, env={"SVTME_SHIM_STATS": "1"})
    m = re.search(r"svtme_wrap: on the GPU: (\d+) open-loop pictures \((\d+) batches\), (\d+) temporal-filter pairs, (\d+) detector pairs", log)
    assert m, "no report from the shim:\n" + log[-1500:]
    pictures, _, tf_pairs, dg_pairs = (int(v) for v in m.groups())
    assert pictures >= n - 1 - n // 16, "open-loop ME of (almost) every picture must have run on the GPU, got %d of %d" % (pictures, n)
    a, b = open(ref_ivf, "rb").read(), open(gpu_ivf, "rb").read()
    assert len(a) > 32 + 12 * n
    assert a == b, "bitstreams differ (%d vs %d bytes); GPU: %d pictures, %d TF pairs, %d DG pairs" % (len(a), len(b), pictures, tf_pairs, dg_pairs)
    print("%s: %d bytes identical; on the GPU: %d pictures, %d TF pairs, %d DG pairs" % (name, len(a), pictures, tf_pairs, dg_pairs))


# the throughput path inside the real encoder: Picture Decision (patched copy of pd_process.c, oracle/pd_process_submit.patch)
# hands every mini-GOP to svtme_shim_submit_pictures, which queues ONE GPU batch and ONE packed result download for it
BATCHED = [e for e in ENCODES if e[0] in ("M8_default", "M4_lp4", "M5_720p", "M2_odd_size", "M8_10bit", "M6_three_layers")] + [
    ("M4_1080p_65", 1920, 1080, 40, ["--preset", "4", "--lp", "8"]),
]


@pytest.mark.gpu
@pytest.mark.skipif(not os.path.exists(BATCHED_APP), reason="SvtAv1EncApp_gpu_batched not built")
@pytest.mark.parametrize("name,w,h,n,args", BATCHED, ids=[e[0] for e in BATCHED])
def test_batched_gpu_encoder_writes_the_reference_bitstream(tmp_path, name, w, h, n, args):
    y4m = str(tmp_path / "in.y4m")
    write_y4m(y4m, w, h, n, seed=len(name) + w, ten_bit=name.endswith("10bit"))
    ref_ivf, gpu_ivf = str(tmp_path / "ref.ivf"), str(tmp_path / "gpu.ivf")
    encode(REF_APP, y4m, ref_ivf, args)
    log = encode(BATCHED_APP, y4m, gpu_ivf, args, env={"SVTME_SHIM_STATS": "1"})
    m = re.search(r"svtme_wrap: on the GPU: (\d+) open-loop pictures \((\d+) batches\), (\d+) temporal-filter pairs, (\d+) detector pairs", log)
    assert m, "no report from the shim:\n" + log[-1500:]
    pictures, batches, tf_pairs, dg_pairs = (int(v) for v in m.groups())
    assert batches > 0, "no mini-GOP went through svtme_shim_submit_pictures"
    assert pictures >= n - 1 - n // 16
    a, b = open(ref_ivf, "rb").read(), open(gpu_ivf, "rb").read()
    assert a == b, "bitstreams differ (%d vs %d bytes); GPU: %d pictures in %d batches" % (len(a), len(b), pictures, batches)
    print("%s: %d bytes identical; on the GPU: %d pictures (%d batches), %d TF pairs, %d DG pairs" % (name, len(a), pictures, batches, tf_pairs, dg_pairs))
