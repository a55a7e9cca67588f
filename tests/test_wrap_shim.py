"""The drop-in boundary exercised from the REFERENCE's side.

oracle/_ref/libsvtref_wrap.so is the reference's own ME sources and the harness that plays me_process.c /
temporal_filtering.c / the detector's caller, linked with
    -Wl,--wrap=svt_aom_motion_estimation_b64 -Wl,--wrap=dg_detector_hme_level0
and the product's binding svt-av1-psy_b200/shim/svtme_wrap.c. The harness code is unchanged: it fills a MeContext
the way the encoder does, calls svt_aom_motion_estimation_b64 per b64 and reads the results out of the MeContext,
pcs->pa_me_data->me_results[] and the PCS statistic arrays. Here those calls land in the GPU library; what the
harness reads back must equal what the unmodified reference produced (the committed fixtures).

The library is built in the development container (it needs the reference's headers) and travels to the GPU box
as a prebuilt file; nothing here reads /root/reference at run time."""
import json
import os

import numpy as np
import pytest

from oracle import pyoracle as po
import golden_lib as gl
from util import CASES, EXTRA, RES_FIELDS, SMALL, case_id, case_inputs, extra_inputs
import test_dg_detector as td
import test_tf_me as tt

pytestmark = pytest.mark.skipif(not po.wrap_available(), reason="oracle/_ref/libsvtref_wrap.so not built (needs the reference's headers)")

PICTURES = [(case_id(c), lambda c=c: case_inputs(c)) for c in CASES] + [(n, lambda g=g, k=k: extra_inputs(g, k)) for n, g, k in EXTRA]


def test_wrapped_library_exports_the_binding():
    lib = po.wrap_lib()
    for sym in ("__wrap_svt_aom_motion_estimation_b64", "__wrap_dg_detector_hme_level0", "svtme_shim_flatten", "svtme_shim_set_device",
                "svtme_shim_release_picture", "svtme_shim_forget_pictures", "svtme_shim_shutdown", "svtme_shim_stats",
                "svtme_shim_submit_pictures", "svtme_shim_batches"):
        assert hasattr(lib, sym), sym


@pytest.mark.parametrize("name,make", PICTURES, ids=[a[0] for a in PICTURES])
def test_shim_flattens_the_reference_context_like_the_fixture(name, make):
    """CPU: MeContext / PCS -> SvtMePicParams by the product's svtme_shim_flatten == the parameters of the fixture
    (which were exported by the test harness's own, independent field copy)"""
    if not gl.have_picture_fixture(name):
        pytest.skip("fixture me_%s.npz not generated" % name)
    fx = gl.PictureFixture(name)
    src, refs, pocs, poc, enc_mode, kw = make()
    pin = po.make_pic_in(src, refs, pocs, poc, enc_mode, **kw)
    assert gl.struct_bytes(po.ref_derive_params_shim(pin)) == gl.struct_bytes(fx.params)


def test_shim_flatten_over_the_parameter_grid():
    """CPU: the same over the (preset, resolution, layer, screen content, prediction structure, qp) grid of the fixture"""
    import gzip
    with gzip.open(os.path.join(gl.GOLDEN_DIR, "sig_deriv_params.json.gz")) as f:
        want = json.loads(f.read().decode())
    n = 0
    for key in gl.sig_deriv_grid():
        pin = gl.sig_deriv_pic_in(po, *key)
        assert gl.fields_of(po.ref_derive_params_shim(pin)) == want[",".join(str(v) for v in key)], key
        n += 1
    assert n == len(want)


@pytest.mark.gpu
@pytest.mark.parametrize("name,make", PICTURES, ids=[a[0] for a in PICTURES])
def test_reference_caller_gets_gpu_results_open_loop(name, make):
    if not gl.have_picture_fixture(name):
        pytest.skip("fixture me_%s.npz not generated" % name)
    fx = gl.PictureFixture(name)
    src, refs, pocs, poc, enc_mode, kw = make()
    fx.check_inputs(src, refs)
    before = po.shim_stats()
    pin = po.make_pic_in(src, refs, pocs, poc, enc_mode, **kw)
    res, _, _ = po.ref_me_picture(pin, lib=po.wrap_lib())
    assert po.shim_stats()[0] == before[0] + 1, "the picture did not run on the GPU"
    fx.compare(res, RES_FIELDS, "reference caller + GPU shim vs reference fixture")


@pytest.mark.gpu
@pytest.mark.parametrize("n_threads", [2, 7, 20])
def test_concurrent_me_threads_share_one_gpu_launch(n_threads):
    """the encoder's ME threads (up to 20-25, enc_handle.c:1066) call the b64 function concurrently, each with its own
    MeContext, on different b64s of the same picture: one of them launches the picture, all of them read their rows"""
    for c in (CASES[0], CASES[4]):
        fx = gl.PictureFixture(case_id(c))
        src, refs, pocs, poc, enc_mode, kw = case_inputs(c)
        before = po.shim_stats()
        pin = po.make_pic_in(src, refs, pocs, poc, enc_mode, **kw)
        res, _ = po.ref_me_picture_mt(pin, n_threads, lib=po.wrap_lib())
        assert po.shim_stats()[0] == before[0] + 1, "exactly one GPU launch per picture"
        fx.compare(res, RES_FIELDS, "%d reference ME threads + GPU shim vs reference fixture" % n_threads)


# a 320x192 preset-4 mini-GOP out of one clip: (temporal layer, poc, list 0, list 1)
MINIGOP = [(1, 16, [0, 8, 4, 2], [32, 24, 20]), (0, 32, [0, 16, 8, 4], [0, 16, 8]), (2, 8, [0, 4, 2], [16, 32]),
           (3, 4, [0, 2], [8, 16]), (3, 12, [8, 4], [16, 32]), (4, 2, [0], [4, 8])]


@pytest.mark.gpu
@pytest.mark.parametrize("n_threads", [1, 6])
def test_minigop_is_one_gpu_batch(orc, n_threads):
    """svtme_shim_submit_pictures: the pictures Picture Decision releases together are queued as ONE batch before the ME
    threads start; the b64 calls wait for it and copy. A picture whose stage structure differs (preset 12 here) is left
    to the per-picture launch. Peer: the oracle (pinned by the fixtures) on the parameters the reference derived."""
    from util import compare, orc_me
    lib = po.wrap_lib()
    pics = [(320, 192, 4, tl, poc, l0, l1) for tl, poc, l0, l1 in MINIGOP] + [CASES[7]]
    ins = [case_inputs(c) for c in pics]
    pins = [po.make_pic_in(src, refs, pocs, poc, enc_mode, **kw) for src, refs, pocs, poc, enc_mode, kw in ins]
    before, batches = po.shim_stats(), lib.svtme_shim_batches()
    got = po.ref_me_minigop(pins, n_threads=n_threads, submit=True, lib=lib)
    assert lib.svtme_shim_batches() == batches + 1
    assert po.shim_stats()[0] == before[0] + len(pics), "6 pictures in the batch + 1 launched on its first b64"
    for (src, refs, pocs, poc, enc_mode, kw), pin, g in zip(ins, pins, got):
        want, _, _ = orc_me(src, refs, po.ref_derive_params(pin, lib), debug=False)
        compare(want, g, fields=RES_FIELDS)
    # the two pictures that have fixtures: against the reference's own output
    gl.PictureFixture(case_id(CASES[0])).compare(got[0], RES_FIELDS, "batched mini-GOP vs reference fixture")
    gl.PictureFixture(case_id(CASES[2])).compare(got[1], RES_FIELDS, "batched mini-GOP vs reference fixture")
    gl.PictureFixture(case_id(CASES[7])).compare(got[6], RES_FIELDS, "picture outside the batch vs reference fixture")


@pytest.mark.gpu
@pytest.mark.parametrize("case", SMALL, ids=case_id)
def test_reference_caller_gets_gpu_results_small_pictures(orc, case):
    """no fixture for these: the oracle (pinned by the fixtures) is the peer"""
    from util import compare, orc_me
    src, refs, pocs, poc, enc_mode, kw = case_inputs(case)
    pin = po.make_pic_in(src, refs, pocs, poc, enc_mode, **kw)
    lib = po.wrap_lib()
    params = po.ref_derive_params(pin, lib)
    got, _, _ = po.ref_me_picture(pin, params, lib=lib)
    want, _, _ = orc_me(src, refs, params, debug=False)
    compare(want, got, fields=RES_FIELDS)


@pytest.mark.gpu
@pytest.mark.parametrize("case", tt.CASES, ids=tt.key)
def test_reference_caller_gets_gpu_results_temporal_filter(case):
    if not os.path.exists(tt.GOLDEN):
        pytest.skip("tests/golden/tf_me.npz not generated")
    z = np.load(tt.GOLDEN)
    w, h, m, lvl, qo, th, qp, cp, rp = case
    centre, ref, _ = tt.inputs(case)
    before = po.shim_stats()
    prm, res, _ = po.ref_tf_me_picture(centre, ref, m, lvl, qp_opt=qo, me_exit_th=th, qp=qp, centre_poc=cp, ref_poc=rp, debug=False,
                                       lib=po.wrap_lib())
    assert po.shim_stats()[1] == before[1] + 1, "the pair did not run on the GPU"
    assert gl.struct_bytes(prm) == z[tt.key(case) + "/params"].tobytes()
    tt.same({f: z[tt.key(case) + "/" + f] for f in tt.TF_FIELDS}, res, "reference TF caller + GPU shim vs reference fixture")


@pytest.mark.gpu
@pytest.mark.parametrize("case", td.CASES, ids=td.key)
def test_reference_caller_gets_gpu_results_dg_detector(case):
    if not os.path.exists(td.GOLDEN):
        pytest.skip("tests/golden/dg_detector.json not generated")
    src, ref = td.pair(case)
    before = po.shim_stats()
    got, _, _ = po.dg_detector(po.wrap_lib(), "ref_", src, ref)
    assert po.shim_stats()[2] == before[2] + 1, "the pair did not run on the GPU"
    assert list(got.as_tuple()) == json.load(open(td.GOLDEN))[td.key(case)]
