"""-m gpu: the throughput forms of the C ABI.
  * svtme_picture_me_batch_packed: one device-to-host copy for a whole batch, any subset of the result groups, against the
    oracle and against the per-array form (svtme_picture_me_batch);
  * merged host uploads: frames adjacent in host memory, named in any order, mixed with frames that are not adjacent;
  * the device-resident view: a later stage reads planes and ME vectors in HBM, no second upload.
CPU part (no GPU): svtme_packed_layout is a pure function of the geometry."""
import copy
import ctypes as C

import numpy as np
import pytest

import svtme_b200 as me
from svt_av1_psy_b200 import abi
from util import CASES, case_inputs, compare, orc_me, params_from_encoder_inputs

GROUPS = {"total_me_candidate_index": abi.F_SB_RESULTS, "me_mv_array": abi.F_SB_RESULTS, "me_candidate_array": abi.F_SB_RESULTS,
          "search_results": abi.F_SEARCH_RESULTS, "me_windows": abi.F_WINDOWS, "p_sb_best_sad": abi.F_BEST, "p_sb_best_mv": abi.F_BEST}


def group_of(field):
    return GROUPS.get(field, abi.F_STATS)


def test_packed_layout_is_dense_and_ordered():
    p = abi.MePicParams()
    p.enable_me_8x8, p.max_refs, p.max_cand = 1, 7, 23
    w, h = 1920, 1080
    n, n_pus, shapes = abi.PicResultArrays.shapes_of(w, h, p)
    for fields in (abi.F_SB_RESULTS | abi.F_STATS, abi.F_ALL, abi.F_BEST | abi.F_SEARCH_RESULTS, abi.F_STATS):
        lay = abi.PackedLayout()
        assert me.lib().svtme_packed_layout(w, h, C.byref(p), fields, C.byref(lay)) == abi.OK
        spans = []
        for i, name in enumerate(abi.PicResultArrays.FIELDS):
            shape, dt = shapes[name]
            nbytes = int(np.prod(shape)) * np.dtype(dt).itemsize
            if group_of(name) & fields:
                assert lay.off[i] != abi.ABSENT and lay.off[i] % 64 == 0
                spans.append((lay.off[i], lay.off[i] + nbytes))
            else:
                assert lay.off[i] == abi.ABSENT
        spans.sort()
        for (a0, a1), (b0, b1) in zip(spans, spans[1:]):
            assert a1 <= b0
        assert spans[0][0] == 0 and spans[-1][1] <= lay.bytes < spans[-1][1] + 256 and lay.bytes % 256 == 0
    # what an open-loop consumer downloads per 1080p picture (preset 4: 85 PUs, 7 references, 23 candidates)
    lay = abi.PackedLayout()
    me.lib().svtme_packed_layout(w, h, C.byref(p), abi.F_SB_RESULTS | abi.F_STATS, C.byref(lay))
    assert lay.bytes < 2.3e6
    assert me.lib().svtme_packed_layout(w, h, C.byref(p), 64, C.byref(lay)) == abi.ERR_BAD_PARAMETER


def test_compact_strides():
    """the picture's own strides: n0 + n1 vectors and the candidates construct_me_candidate_array* can emit"""
    p = abi.MePicParams()
    p.enable_me_8x8, p.max_refs, p.max_cand, p.max_l0 = 1, 7, 23, 4
    p.num_of_list_to_search = 2
    for n0, n1, only_l_bwd, want in [(4, 3, 1, (7, 4, 8)), (4, 3, 0, (7, 4, 23)), (2, 1, 1, (3, 2, 4)), (2, 2, 0, (4, 2, 9)), (1, 1, 1, (2, 1, 3))]:
        p.num_of_ref_pic_to_search[0], p.num_of_ref_pic_to_search[1], p.only_l_bwd = n0, n1, only_l_bwd
        st = me.Session.sb_strides(p, abi.F_SB_RESULTS | abi.F_SB_COMPACT)
        assert (st.mv_stride, st.mv_l1_offset, st.cand_stride) == want and st.n_pus == 85
        st = me.Session.sb_strides(p, abi.F_SB_RESULTS)
        assert (st.mv_stride, st.mv_l1_offset, st.cand_stride) == (7, 4, 23)
    p.num_of_list_to_search, p.num_of_ref_pic_to_search[0], p.num_of_ref_pic_to_search[1] = 1, 1, 0
    st = me.Session.sb_strides(p, abi.F_SB_COMPACT)
    assert (st.mv_stride, st.mv_l1_offset, st.cand_stride) == (1, 1, 1)


def _batch(n_pics=5):
    """n_pics pictures of one stream (one session): sources, references, parameters with slot ids"""
    src0, refs0, pocs, poc, enc_mode, kw = case_inputs(CASES[0])  # 320x192 preset 4, 4+3 references
    h, w = src0.shape
    rng = np.random.default_rng(11)
    planes = [src0] + [r for lst in refs0 for r in lst]
    pics = []
    for k in range(n_pics):
        src = np.roll(src0, (k, 2 * k), (0, 1)) if k else src0
        src = np.clip(src.astype(np.int16) + rng.integers(-2, 3, src.shape), 0, 255).astype(np.uint8)
        params = params_from_encoder_inputs(src, refs0, pocs, poc, enc_mode, **kw)
        pics.append((src, params))
    return w, h, planes, refs0, pics


@pytest.mark.gpu
@pytest.mark.parametrize("fields", [abi.F_SB_RESULTS | abi.F_STATS, abi.F_ALL, abi.F_BEST | abi.F_SEARCH_RESULTS | abi.F_WINDOWS])
def test_packed_batch_matches_oracle(orc, fields):
    import torch
    w, h, planes, refs, pics = _batch()
    s = me.Session(w, h, max_pictures=16)
    for i, pl in enumerate(planes[1:]):
        s.upload(1 + i, pl)
    ids, prms = [], []
    for k, (src, params) in enumerate(pics):
        s.upload(8 + k, src)
        p = copy.copy(params)
        pid = 1
        for l in range(len(refs)):
            for r in range(len(refs[l])):
                p.ref_pic_id[l][r] = pid
                pid += 1
        ids.append(8 + k)
        prms.append(p)
    n = len(ids)
    lay = s.packed_layout(prms[0], fields)
    slab = torch.zeros(lay.bytes * n, dtype=torch.uint8).pin_memory()
    slab.fill_(0xA5)
    ids_a, prm_a = (C.c_int32 * n)(*ids), (abi.MePicParams * n)(*prms)
    for rep in range(2):  # twice: the two result sets of the session
        s.me_batch_packed(ids_a, prm_a, n, fields, slab.data_ptr(), slab.numel(), sync=False)
        s.wait_batch(s.batch_seq())
        block = slab.numpy()
        for k, (src, params) in enumerate(pics):
            want, _, _ = orc_me(src, refs, params, debug=False)
            got = abi.PicResultArrays(w, h, prms[k], block=block[k * lay.bytes:(k + 1) * lay.bytes], layout=lay)
            names = [f for f in got.FIELDS if getattr(got, f) is not None]
            assert set(names) == {f for f in got.FIELDS if group_of(f) & fields}
            compare(want, got, fields=names)
    # the per-array form of the same batch gives the same bytes
    res = [abi.PicResultArrays(w, h, p) for p in prms]
    s.me_batch(ids, prms, res, sync=True)
    block = slab.numpy()
    for k in range(n):
        got = abi.PicResultArrays(w, h, prms[k], block=block[k * lay.bytes:(k + 1) * lay.bytes], layout=lay)
        for f in got.FIELDS:
            if getattr(got, f) is not None:
                np.testing.assert_array_equal(getattr(got, f), getattr(res[k], f), err_msg=f)
    # a slab that is too small is refused before anything is queued
    with pytest.raises(me.binding.SvtMeError):
        s.me_batch_packed(ids_a, prm_a, n, fields, slab.data_ptr(), slab.numel() - 256)
    # a field the batch was not asked to produce cannot be fetched afterwards
    if not fields & abi.F_BEST:
        s.me_batch_packed(ids_a, prm_a, n, fields, slab.data_ptr(), slab.numel(), sync=True)
        with pytest.raises(me.binding.SvtMeError):
            s.fetch_batch(0, abi.PicResultArrays(w, h, prms[0]))
    s.close()


@pytest.mark.gpu
@pytest.mark.parametrize("concurrency", [1, 2])
def test_three_slabs_in_rotation(orc, concurrency):
    """The bench's end-to-end pipeline: batches queued without waiting into three host slabs in turn, two compute streams;
    the library alone orders the reuse of its two device-side result regions (the first writing kernel of batch k waits
    for the download of batch k - 2), svtme_session_wait_batch(seq) waits for that batch's OWN download."""
    import torch
    w, h, planes, refs, pics = _batch(6)
    s = me.Session(w, h, max_pictures=16)
    s.set_concurrency(concurrency)
    for i, pl in enumerate(planes[1:]):
        s.upload(1 + i, pl)
    fields = abi.F_SB_RESULTS | abi.F_SB_COMPACT | abi.F_STATS
    prms = []
    for k, (src, params) in enumerate(pics):
        s.upload(8 + k, src)
        p = copy.copy(params)
        pid = 1
        for l in range(len(refs)):
            for r in range(len(refs[l])):
                p.ref_pic_id[l][r] = pid
                pid += 1
        prms.append(p)
    lay = s.packed_layout(prms[0], fields)
    strides = me.Session.sb_strides(prms[0], fields)
    slabs = [torch.zeros(lay.bytes * 2, dtype=torch.uint8).pin_memory() for _ in range(3)]
    s.reserve(prms[0], 2)
    # batch b = pictures (b, (b + 1) % 6): six batches, each slab used twice
    seqs, want = [], {}
    checked = 0
    for b in range(6):
        k = b % 3
        if b >= 3:  # the slab's previous batch: wait, check, hand out again
            s.wait_batch(seqs[b - 3])
            checked += _check_slab(slabs[k], lay, strides, w, h, prms, refs, pics, [b - 3, (b - 2) % 6], want)
            slabs[k].fill_(0x5A)
        pair = [b, (b + 1) % 6]
        ids_a, prm_a = (C.c_int32 * 2)(*[8 + i for i in pair]), (abi.MePicParams * 2)(*[prms[i] for i in pair])
        s.me_batch_packed(ids_a, prm_a, 2, fields, slabs[k].data_ptr(), slabs[k].numel(), sync=False)
        seqs.append(s.batch_seq())
    for b in (5, 3, 4):  # out of order
        s.wait_batch(seqs[b])
        checked += _check_slab(slabs[b % 3], lay, strides, w, h, prms, refs, pics, [b, (b + 1) % 6], want)
    assert checked == 12
    s.sync()
    s.close()


@pytest.mark.gpu
def test_wait_for_a_batch_older_than_the_event_ring(orc):
    """ten batches queued into ten slabs without a single wait: waiting for the first one (its event slots were handed
    on to batch 9) still returns only when its results are in host memory"""
    import torch
    w, h, planes, refs, pics = _batch(6)
    s = me.Session(w, h, max_pictures=16)
    s.set_concurrency(2)
    for i, pl in enumerate(planes[1:]):
        s.upload(1 + i, pl)
    fields = abi.F_SB_RESULTS | abi.F_SB_COMPACT | abi.F_STATS
    prms = []
    for k, (src, params) in enumerate(pics):
        s.upload(8 + k, src)
        p = copy.copy(params)
        pid = 1
        for l in range(len(refs)):
            for r in range(len(refs[l])):
                p.ref_pic_id[l][r] = pid
                pid += 1
        prms.append(p)
    lay = s.packed_layout(prms[0], fields)
    strides = me.Session.sb_strides(prms[0], fields)
    slabs = [torch.zeros(lay.bytes * 2, dtype=torch.uint8).pin_memory() for _ in range(10)]
    s.reserve(prms[0], 2)
    seqs, want = [], {}
    for b in range(10):
        pair = [b % 6, (b + 1) % 6]
        ids_a, prm_a = (C.c_int32 * 2)(*[8 + i for i in pair]), (abi.MePicParams * 2)(*[prms[i] for i in pair])
        s.me_batch_packed(ids_a, prm_a, 2, fields, slabs[b].data_ptr(), slabs[b].numel(), sync=False)
        seqs.append(s.batch_seq())
    s.wait_batch(seqs[0])
    _check_slab(slabs[0], lay, strides, w, h, prms, refs, pics, [0, 1], want)
    s.wait_batch(seqs[9])
    _check_slab(slabs[9], lay, strides, w, h, prms, refs, pics, [9 % 6, 10 % 6], want)
    s.sync()
    s.close()


def _check_slab(slab, lay, strides, w, h, prms, refs, pics, pair, want):
    block = slab.numpy()
    for j, i in enumerate(pair):
        if i not in want:
            want[i] = orc_me(pics[i][0], refs, pics[i][1], debug=False)[0]
        got_c = abi.PicResultArrays(w, h, prms[i], block=block[j * lay.bytes:(j + 1) * lay.bytes], layout=lay, strides=strides)
        got = me.Session.sb_expand(prms[i], abi.F_SB_RESULTS | abi.F_SB_COMPACT | abi.F_STATS, got_c, abi.PicResultArrays(w, h, prms[i]))
        for f in ("total_me_candidate_index", "me_mv_array", "me_candidate_array"):
            np.testing.assert_array_equal(getattr(got, f), getattr(want[i], f), err_msg="picture %d %s" % (i, f))
        for f in got_c.FIELDS:
            if group_of(f) == abi.F_STATS:
                np.testing.assert_array_equal(getattr(got_c, f), getattr(want[i], f), err_msg="picture %d %s" % (i, f))
    return len(pair)


@pytest.mark.gpu
def test_merged_uploads(orc):
    """frames adjacent in host memory cross as one copy whatever the order they are named in; pitched and scattered
    frames take the per-frame path; the planes on the device are the same either way"""
    import torch
    w, h = 320, 192
    rng = np.random.default_rng(5)
    clip = torch.from_numpy(rng.integers(0, 256, (7, h, w), dtype=np.uint8)).pin_memory()
    lone = torch.from_numpy(rng.integers(0, 256, (h, w), dtype=np.uint8)).pin_memory()
    s = me.Session(w, h, max_pictures=16)
    order = [4, 0, 6, 5, 1]  # frames 0-1 and 4-6 are two runs; plus one frame elsewhere
    ids = [10, 11, 12, 13, 14, 15]
    ptrs = [clip[f].data_ptr() for f in order] + [lone.data_ptr()]
    for rep in range(40):  # several laps of the staging ring
        s.upload_batch(ids, ptrs, w, "host8")
    want = [clip[f].numpy() for f in order] + [lone.numpy()]
    ref = me.Session(w, h, max_pictures=2)
    for pid, luma in zip(ids, want):
        ref.upload(0, luma)
        for lvl in range(3):
            np.testing.assert_array_equal(s.download_plane(pid, lvl), ref.download_plane(0, lvl), err_msg="picture %d level %d" % (pid, lvl))
    # 10-bit frames through the same ring
    clip16 = torch.from_numpy(rng.integers(0, 1024, (3, h, w), dtype=np.uint16).view(np.int16)).pin_memory()
    s.upload_batch([1, 2, 3], [clip16[f].data_ptr() for f in (2, 0, 1)], w, "host10")
    for pid, f in zip((1, 2, 3), (2, 0, 1)):
        ref.upload(0, (clip16[f].numpy().view(np.uint16) >> 2).astype(np.uint8))
        np.testing.assert_array_equal(s.download_plane(pid, 2), ref.download_plane(0, 2))
    s.close()
    ref.close()


@pytest.mark.gpu
def test_device_resident_view(orc):
    """row f4: GM pre-processing / TPL-style consumers read the planes and the ME vectors where they are"""
    import torch
    w, h, planes, refs, pics = _batch(3)
    s = me.Session(w, h, max_pictures=16)
    for i, pl in enumerate(planes[1:]):
        s.upload(1 + i, pl)
    ids, prms = [], []
    for k, (src, params) in enumerate(pics):
        s.upload(8 + k, src)
        p = copy.copy(params)
        pid = 1
        for l in range(len(refs)):
            for r in range(len(refs[l])):
                p.ref_pic_id[l][r] = pid
                pid += 1
        ids.append(8 + k)
        prms.append(p)
    # planes: the device view of a picture = what the oracle builds from the same luma
    s.sync()
    dp = s.device_planes(8)
    for lvl in range(3):
        pl = dp[lvl]
        assert (pl.width, pl.height) == (w >> lvl, h >> lvl) and pl.stride >= pl.width + 2 * pl.pad
        rows = s.read_device(pl.origin - pl.pad * pl.stride - pl.pad, (pl.height + 2 * pl.pad) * pl.stride).reshape(pl.height + 2 * pl.pad, pl.stride)
        exp = s.download_plane(8, lvl)  # padded by 72 / 32 / 16 as the reference's buffers
        ep = (72, 32, 16)[lvl]
        got = rows[pl.pad - ep:pl.pad + pl.height + ep, pl.pad - ep:pl.pad + pl.width + ep]
        np.testing.assert_array_equal(got, exp)
    # results: read the MeSbResults arrays of every picture of the batch in place, compare with the host copy
    fields = abi.F_SB_RESULTS | abi.F_STATS
    n = len(ids)
    lay = s.packed_layout(prms[0], fields)
    slab = torch.zeros(lay.bytes * n, dtype=torch.uint8).pin_memory()
    ids_a, prm_a = (C.c_int32 * n)(*ids), (abi.MePicParams * n)(*prms)
    s.me_batch_packed(ids_a, prm_a, n, fields, slab.data_ptr(), slab.numel(), sync=False)
    seq = s.batch_seq()
    consumer = torch.cuda.Stream()
    views = [s.acquire_results(seq, k, consumer.cuda_stream) for k in range(n)]
    consumer.synchronize()  # the consumer's stream was made to wait for the batch's kernels
    s.wait_batch(seq)
    for k in range(n):
        host = abi.PicResultArrays(w, h, prms[k], block=slab.numpy()[k * lay.bytes:(k + 1) * lay.bytes], layout=lay)
        dev = views[k]
        assert dev.p_sb_best_sad is None and dev.search_results is None  # not produced
        for f in ("total_me_candidate_index", "me_mv_array", "me_candidate_array", "me_8x8_distortion", "rc_me_allow_gm"):
            a = getattr(host, f)
            got = s.read_device(getattr(dev, f), a.nbytes).view(a.dtype).reshape(a.shape)
            np.testing.assert_array_equal(got, a, err_msg="picture %d %s" % (k, f))
    s.release_results(seq, consumer.cuda_stream)
    # two batches later the set is reused: the view is refused
    s.me_batch_packed(ids_a, prm_a, n, fields, slab.data_ptr(), slab.numel(), sync=True)
    s.me_batch_packed(ids_a, prm_a, n, fields, slab.data_ptr(), slab.numel(), sync=True)
    with pytest.raises(me.binding.SvtMeError):
        s.acquire_results(seq, 0, consumer.cuda_stream)
    s.close()


@pytest.mark.gpu
@pytest.mark.parametrize("case", [CASES[0], CASES[3], CASES[4], CASES[6], CASES[7]], ids=lambda c: "%dx%d_M%d" % (c[0], c[1], c[2]))
def test_compact_wire_format_expands_to_the_oracle(orc, case):
    """SVTME_F_SB_COMPACT: fewer bytes cross PCIe, svtme_sb_expand_b64 gives the reference's arrays back (general, MRP,
    mrp-off and single-reference candidate builders, only_l_bwd on and off)"""
    import torch
    src, refs, pocs, poc, enc_mode, kw = case_inputs(case)
    h, w = src.shape
    params = params_from_encoder_inputs(src, refs, pocs, poc, enc_mode, **kw)
    want, _, _ = orc_me(src, refs, params, debug=False)
    s = me.Session(w, h, max_pictures=12)
    s.upload(0, src)
    p = copy.copy(params)
    pid = 1
    for l in range(len(refs)):
        for r, plane in enumerate(refs[l]):
            s.upload(pid, plane)
            p.ref_pic_id[l][r] = pid
            pid += 1
    fields = abi.F_SB_RESULTS | abi.F_SB_COMPACT | abi.F_STATS
    lay, full = s.packed_layout(p, fields), s.packed_layout(p, abi.F_SB_RESULTS | abi.F_STATS)
    st = me.Session.sb_strides(p, fields)
    assert lay.bytes <= full.bytes
    if st.cand_stride < p.max_cand or st.mv_stride < p.max_refs:
        assert lay.bytes < full.bytes
    slab = torch.zeros(lay.bytes, dtype=torch.uint8).pin_memory()
    s.me_batch_packed((C.c_int32 * 1)(0), (abi.MePicParams * 1)(p), 1, fields, slab.data_ptr(), slab.numel(), sync=True)
    got_c = abi.PicResultArrays(w, h, p, block=slab.numpy(), layout=lay, strides=st)
    got = me.Session.sb_expand(p, fields, got_c, abi.PicResultArrays(w, h, p))
    for f in ("total_me_candidate_index", "me_mv_array", "me_candidate_array"):
        np.testing.assert_array_equal(getattr(got, f), getattr(want, f), err_msg=f)
    for f in got_c.FIELDS:
        if group_of(f) == abi.F_STATS:
            np.testing.assert_array_equal(getattr(got_c, f), getattr(want, f), err_msg=f)
    s.close()
