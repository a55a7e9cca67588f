"""ctypes binding of libsvtme_b200.so (CUDA, include/svtme_b200.h) and libsvtme_host.so (host C)."""
import ctypes as C
import os

import numpy as np

from . import abi

_PKG = os.path.dirname(os.path.abspath(__file__))
_LIB = None
_HOST = None

u8p = C.c_void_p


class SvtMeError(RuntimeError):
    def __init__(self, code, what):
        super().__init__("%s failed: error 0x%08x" % (what, code))
        self.code = code


def lib():
    """The CUDA C-ABI library. Raises (loudly) if it has not been built: there is no fallback."""
    global _LIB
    if _LIB is None:
        path = os.path.join(_PKG, "libsvtme_b200.so")
        if not os.path.exists(path):
            raise RuntimeError(
                "libsvtme_b200.so is missing (%s). Build it with __graft_entry__.build() or "
                "`make -C svt-av1-psy_b200/csrc`. There is no CPU fallback." % path)
        L = C.CDLL(path)
        L.svtme_version.restype = C.c_char_p
        L.svtme_last_error.restype = C.c_uint32
        L.svtme_nxm_sad_kernel.restype = C.c_uint32
        L.svtme_session_create.argtypes = [C.POINTER(C.c_void_p), C.c_int, C.c_uint32, C.c_uint32, C.c_uint32]
        L.svtme_session_create.restype = C.c_uint32
        L.svtme_session_destroy.argtypes = [C.c_void_p]
        L.svtme_session_destroy.restype = None
        L.svtme_picture_upload.argtypes = [C.c_void_p, C.c_int32, C.c_void_p, C.c_uint32]
        L.svtme_picture_upload.restype = C.c_uint32
        L.svtme_picture_upload_10bit.argtypes = [C.c_void_p, C.c_int32, C.c_void_p, C.c_uint32]
        L.svtme_picture_upload_10bit.restype = C.c_uint32
        L.svtme_picture_upload_device.argtypes = [C.c_void_p, C.c_int32, C.c_void_p, C.c_uint32]
        L.svtme_picture_upload_device.restype = C.c_uint32
        L.svtme_picture_me.argtypes = [C.c_void_p, C.c_int32, C.POINTER(abi.MePicParams),
                                       C.POINTER(abi.MePicResults), C.c_int]
        L.svtme_picture_me.restype = C.c_uint32
        L.svtme_picture_me_device.argtypes = [C.c_void_p, C.c_int32, C.POINTER(abi.MePicParams)]
        L.svtme_picture_me_device.restype = C.c_uint32
        L.svtme_fetch_results.argtypes = [C.c_void_p, C.POINTER(abi.MePicResults), C.c_int]
        L.svtme_fetch_results.restype = C.c_uint32
        L.svtme_session_sync.argtypes = [C.c_void_p]
        L.svtme_session_sync.restype = C.c_uint32
        L.svtme_dg_detector_hme.argtypes = [C.c_void_p, C.c_int32, C.c_int32, C.POINTER(abi.DgMetrics), C.c_void_p, C.c_void_p]
        L.svtme_dg_detector_hme.restype = C.c_uint32
        L.svtme_session_batch_seq.argtypes = [C.c_void_p]
        L.svtme_session_batch_seq.restype = C.c_uint64
        L.svtme_session_wait_batch.argtypes = [C.c_void_p, C.c_uint64]
        L.svtme_session_wait_batch.restype = C.c_uint32
        L.svtme_picture_download_plane.argtypes = [C.c_void_p, C.c_int32, C.c_int, C.c_void_p, C.c_size_t,
                                                   C.POINTER(C.c_uint32), C.POINTER(C.c_uint32),
                                                   C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]
        L.svtme_picture_download_plane.restype = C.c_uint32
        L.svtme_session_stream.argtypes = [C.c_void_p]
        L.svtme_session_stream.restype = C.c_void_p
        L.svtme_session_set_stream.argtypes = [C.c_void_p, C.c_void_p]
        L.svtme_session_set_stream.restype = C.c_uint32
        L.svtme_session_launch_count.argtypes = [C.c_void_p]
        L.svtme_session_launch_count.restype = C.c_uint64
        L.svtme_session_kernel_time.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_double),
                                                C.POINTER(C.c_double), C.POINTER(C.c_uint64)]
        L.svtme_session_kernel_time.restype = C.c_uint32
        L.svtme_picture_me_batch.argtypes = [C.c_void_p, C.c_uint32, C.POINTER(C.c_int32), C.POINTER(abi.MePicParams),
                                             C.POINTER(abi.MePicResults), C.c_int]
        L.svtme_picture_me_batch.restype = C.c_uint32
        L.svtme_fetch_results_batch.argtypes = [C.c_void_p, C.c_uint32, C.POINTER(abi.MePicResults), C.c_int]
        L.svtme_fetch_results_batch.restype = C.c_uint32
        L.svtme_session_enable_timing.argtypes = [C.c_void_p, C.c_int]
        L.svtme_session_enable_timing.restype = C.c_uint32
        L.svtme_fetch_hme_state.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t]
        L.svtme_fetch_hme_state.restype = C.c_uint32
        L.svtme_session_sad_ops.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
        L.svtme_session_sad_ops.restype = C.c_uint32
        for fn in (L.svtme_picture_upload_batch, L.svtme_picture_upload_10bit_batch, L.svtme_picture_upload_device_batch):
            fn.argtypes = [C.c_void_p, C.c_uint32, C.POINTER(C.c_int32), C.POINTER(C.c_void_p), C.c_uint32]
            fn.restype = C.c_uint32
        L.svtme_kernel_name.argtypes = [C.c_int]
        L.svtme_kernel_name.restype = C.c_char_p
        L.svtme_session_kernel_times.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_double), C.c_int]
        L.svtme_session_kernel_times.restype = C.c_uint32
        L.svtme_session_sad_ops_by_stage.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_uint64)]
        L.svtme_session_sad_ops_by_stage.restype = C.c_uint32
        L.svtme_packed_layout.argtypes = [C.c_uint32, C.c_uint32, C.POINTER(abi.MePicParams), C.c_uint32, C.POINTER(abi.PackedLayout)]
        L.svtme_packed_layout.restype = C.c_uint32
        L.svtme_picture_me_batch_packed.argtypes = [C.c_void_p, C.c_uint32, C.POINTER(C.c_int32), C.POINTER(abi.MePicParams), C.c_uint32,
                                                    C.c_void_p, C.c_size_t, C.c_int]
        L.svtme_picture_me_batch_packed.restype = C.c_uint32
        L.svtme_session_set_result_fields.argtypes = [C.c_void_p, C.c_uint32]
        L.svtme_session_set_result_fields.restype = C.c_uint32
        L.svtme_session_set_unit_granularity.argtypes = [C.c_void_p, C.c_int]
        L.svtme_session_set_unit_granularity.restype = C.c_uint32
        L.svtme_session_reserve.argtypes = [C.c_void_p, C.POINTER(abi.MePicParams), C.c_uint32]
        L.svtme_session_reserve.restype = C.c_uint32
        L.svtme_picture_device_planes.argtypes = [C.c_void_p, C.c_int32, C.POINTER(abi.DevicePlane)]
        L.svtme_picture_device_planes.restype = C.c_uint32
        L.svtme_batch_results_acquire.argtypes = [C.c_void_p, C.c_uint64, C.c_uint32, C.POINTER(abi.MePicResults), C.c_void_p]
        L.svtme_batch_results_acquire.restype = C.c_uint32
        L.svtme_batch_results_release.argtypes = [C.c_void_p, C.c_uint64, C.c_void_p]
        L.svtme_batch_results_release.restype = C.c_uint32
        L.svtme_copy_from_device.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t]
        L.svtme_copy_from_device.restype = C.c_uint32
        L.svtme_sb_strides.argtypes = [C.POINTER(abi.MePicParams), C.c_uint32, C.POINTER(abi.SbStrides)]
        L.svtme_sb_strides.restype = None
        L.svtme_sb_expand_b64.argtypes = [C.POINTER(abi.MePicParams), C.c_uint32, C.POINTER(abi.MePicResults), C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p]
        L.svtme_sb_expand_b64.restype = None
        L.svtme_host_alloc.argtypes = [C.c_size_t]
        L.svtme_host_alloc.restype = C.c_void_p
        L.svtme_host_free.argtypes = [C.c_void_p]
        L.svtme_host_free.restype = None
        L.svtme_abi_sizeof.argtypes = [C.c_int]
        L.svtme_abi_sizeof.restype = C.c_uint32
        _LIB = L
    return _LIB


def host_lib():
    """Host-side C logic (signal derivation, prediction structure): no CUDA inside."""
    global _HOST
    if _HOST is None:
        path = os.path.join(_PKG, "libsvtme_host.so")
        if not os.path.exists(path):
            raise RuntimeError("libsvtme_host.so is missing (%s): run __graft_entry__.build()" % path)
        H = C.CDLL(path)
        H.svtme_sig_deriv_me.restype = C.c_uint32
        _HOST = H
    return _HOST


def last_error():
    return lib().svtme_last_error()


def _check(code, what):
    if code != abi.OK:
        raise SvtMeError(code, what)


class SigDerivTfIn(C.Structure):
    """SvtMeSigDerivTfIn (svt-av1-psy_b200/host/me_sig_deriv.h)"""
    _fields_ = [("enc_mode", C.c_int32), ("width", C.c_int32), ("height", C.c_int32), ("qp", C.c_int32),
                ("tf_hme_me_level", C.c_int32), ("tf_qp_opt", C.c_int32), ("tf_me_exit_th", C.c_uint32),
                ("temporal_layer_index", C.c_int32), ("is_ref", C.c_int32),
                ("centre_picture_number", C.c_uint64), ("ref_picture_number", C.c_uint64)]


def sig_deriv_me_tf(enc_mode, width, height, hme_me_level, qp_opt=0, me_exit_th=0, qp=35, centre_poc=8, ref_poc=7,
                    temporal_layer=0, is_ref=1):
    """temporal-filter ME controls through the product's host C logic (svtme_sig_deriv_me_tf)"""
    t = SigDerivTfIn(enc_mode, width, height, qp, hme_me_level, qp_opt, me_exit_th, temporal_layer, is_ref, centre_poc, ref_poc)
    p = abi.MePicParams()
    f = host_lib().svtme_sig_deriv_me_tf
    f.argtypes = [C.POINTER(SigDerivTfIn), C.POINTER(abi.MePicParams)]
    f.restype = C.c_uint32
    _check(f(C.byref(t), C.byref(p)), "svtme_sig_deriv_me_tf")
    return p


class SigDerivIn(C.Structure):
    """SvtMeSigDerivIn of host/me_sig_deriv.h: the encoder-level inputs of svt_aom_sig_deriv_me."""
    _fields_ = [
        ("enc_mode", C.c_int32), ("width", C.c_int32), ("height", C.c_int32), ("qp", C.c_int32),
        ("pred_structure", C.c_int32), ("hierarchical_levels", C.c_int32), ("temporal_layer_index", C.c_int32),
        ("is_ref", C.c_int32), ("sc_class1", C.c_int32), ("frame_rate", C.c_uint32),
        ("similar_brightness_refs", C.c_int32), ("gm_level", C.c_int32), ("only_l_bwd", C.c_int32),
        ("safe_limit_nref", C.c_int32), ("safe_limit_zz_th", C.c_int32), ("max_l0", C.c_int32), ("max_l1", C.c_int32),
        ("picture_number", C.c_uint64), ("num_lists", C.c_int32), ("ref_count", C.c_int32 * 2),
        ("ref_picture_number", (C.c_uint64 * 4) * 2),
    ]


def sig_deriv_me(sin):
    """host C restatement of svt_aom_sig_deriv_me + PCS-level ME fields -> abi.MePicParams"""
    p = abi.MePicParams()
    _check(host_lib().svtme_sig_deriv_me(C.byref(sin), C.byref(p)), "svtme_sig_deriv_me")
    return p


class Session:
    """One GPU's picture store + ME launcher (SvtMeSession)."""

    def __init__(self, width, height, max_pictures=16, device=0):
        self.width, self.height = width, height
        self._h = C.c_void_p()
        _check(lib().svtme_session_create(C.byref(self._h), device, width, height, max_pictures), "svtme_session_create")

    def close(self):
        if self._h:
            lib().svtme_session_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def upload(self, pic_id, luma):
        a = np.ascontiguousarray(luma)
        if a.dtype == np.uint16:
            assert a.shape == (self.height, self.width)
            _check(lib().svtme_picture_upload_10bit(self._h, pic_id, a.ctypes.data, a.strides[0] // 2), "svtme_picture_upload_10bit")
        else:
            assert a.dtype == np.uint8 and a.shape == (self.height, self.width)
            _check(lib().svtme_picture_upload(self._h, pic_id, a.ctypes.data, a.strides[0]), "svtme_picture_upload")

    def upload_ptr(self, pic_id, host_ptr, stride):
        _check(lib().svtme_picture_upload(self._h, pic_id, host_ptr, stride), "svtme_picture_upload")

    def upload_ptr16(self, pic_id, host_ptr, stride_in_samples):
        _check(lib().svtme_picture_upload_10bit(self._h, pic_id, host_ptr, stride_in_samples), "svtme_picture_upload_10bit")

    def upload_device(self, pic_id, dev_ptr, stride):
        _check(lib().svtme_picture_upload_device(self._h, pic_id, dev_ptr, stride), "svtme_picture_upload_device")

    def upload_batch(self, pic_ids, ptrs, stride, kind="host8"):
        """several pictures at once (two kernel launches for all their pyramids). kind: host8 / host10 / device;
        ptrs: raw addresses (pinned host memory or device memory)"""
        n = len(pic_ids)
        ids = (C.c_int32 * n)(*pic_ids)
        pp = (C.c_void_p * n)(*ptrs)
        fn = {"host8": lib().svtme_picture_upload_batch, "host10": lib().svtme_picture_upload_10bit_batch,
              "device": lib().svtme_picture_upload_device_batch}[kind]
        _check(fn(self._h, n, ids, pp, stride), "svtme_picture_upload_*_batch")

    def me(self, pic_id, params, results=None, sync=True):
        results = results if results is not None else abi.PicResultArrays(self.width, self.height, params)
        _check(lib().svtme_picture_me(self._h, pic_id, C.byref(params), C.byref(results.struct), 1 if sync else 0),
               "svtme_picture_me")
        return results

    def me_batch(self, pic_ids, params_list, results_list=None, sync=True):
        """ME of several pictures in one launch sequence. results_list: PicResultArrays per picture or None."""
        n = len(pic_ids)
        ids = (C.c_int32 * n)(*pic_ids)
        prm = (abi.MePicParams * n)(*params_list)
        res = None
        if results_list is not None:
            res = (abi.MePicResults * n)(*[r.struct for r in results_list])
        _check(lib().svtme_picture_me_batch(self._h, n, ids, prm, res, 1 if sync else 0), "svtme_picture_me_batch")
        return results_list

    def packed_layout(self, params, fields):
        lay = abi.PackedLayout()
        _check(lib().svtme_packed_layout(self.width, self.height, C.byref(params), fields, C.byref(lay)), "svtme_packed_layout")
        return lay

    def me_batch_packed(self, ids_array, params_array, n, fields, slab_ptr, slab_bytes, sync=False):
        """one launch sequence + ONE device-to-host copy for the whole batch; ids_array / params_array: ctypes arrays the
        caller keeps (no per-call marshalling), slab_ptr: host memory of slab_bytes bytes (pinned for an asynchronous copy)"""
        _check(lib().svtme_picture_me_batch_packed(self._h, n, ids_array, params_array, fields, slab_ptr, slab_bytes, 1 if sync else 0),
               "svtme_picture_me_batch_packed")

    @staticmethod
    def sb_strides(params, fields):
        st = abi.SbStrides()
        lib().svtme_sb_strides(C.byref(params), fields, C.byref(st))
        return st

    @staticmethod
    def sb_expand(params, fields, res, out):
        """compact block arrays `res` -> `out` (PicResultArrays with the reference's strides), b64 by b64 through the C helper"""
        n_pus = 85 if params.enable_me_8x8 else 21
        for b in range(out.n_b64):
            lib().svtme_sb_expand_b64(C.byref(params), fields, C.byref(res.struct), b, out.total_me_candidate_index[b].ctypes.data,
                                      out.me_mv_array[b].ctypes.data, out.me_candidate_array[b].ctypes.data)
        return out

    def set_result_fields(self, fields):
        _check(lib().svtme_session_set_result_fields(self._h, fields), "svtme_session_set_result_fields")

    def set_unit_granularity(self, mode):
        """-1 auto, 0 one k_search16 unit per (b64, list, ref), 1 one per job"""
        _check(lib().svtme_session_set_unit_granularity(self._h, mode), "svtme_session_set_unit_granularity")

    def reserve(self, params, n_pictures):
        _check(lib().svtme_session_reserve(self._h, C.byref(params), n_pictures), "svtme_session_reserve")

    def device_planes(self, pic_id):
        planes = (abi.DevicePlane * 3)()
        _check(lib().svtme_picture_device_planes(self._h, pic_id, planes), "svtme_picture_device_planes")
        return planes

    def acquire_results(self, seq, index, consumer_stream=None):
        """device pointers (abi.MePicResults) to picture `index` of batch `seq`; consumer_stream waits for the batch"""
        r = abi.MePicResults()
        _check(lib().svtme_batch_results_acquire(self._h, seq, index, C.byref(r), C.c_void_p(consumer_stream)), "svtme_batch_results_acquire")
        return r

    def release_results(self, seq, consumer_stream=None):
        _check(lib().svtme_batch_results_release(self._h, seq, C.c_void_p(consumer_stream)), "svtme_batch_results_release")

    def read_device(self, dev_ptr, nbytes):
        out = np.zeros(nbytes, np.uint8)
        _check(lib().svtme_copy_from_device(self._h, out.ctypes.data, C.c_void_p(dev_ptr), nbytes), "svtme_copy_from_device")
        return out

    def fetch_batch(self, index, results, sync=True):
        _check(lib().svtme_fetch_results_batch(self._h, index, C.byref(results.struct), 1 if sync else 0), "svtme_fetch_results_batch")
        return results

    def me_device(self, pic_id, params):
        _check(lib().svtme_picture_me_device(self._h, pic_id, C.byref(params)), "svtme_picture_me_device")

    def fetch(self, results, sync=True):
        _check(lib().svtme_fetch_results(self._h, C.byref(results.struct), 1 if sync else 0), "svtme_fetch_results")
        return results

    def sync(self):
        _check(lib().svtme_session_sync(self._h), "svtme_session_sync")

    def dg_detector(self, src_pic_id, ref_pic_id, per_b64=True):
        """dynamic-GOP detector HME of an uploaded (source, reference) pair -> (DgMetrics, sad[n_b64], mv[n_b64][2])"""
        m = abi.DgMetrics()
        m.input_resolution = abi.input_resolution_of(self.width, self.height)
        n = abi.n_b64_of(self.width, self.height)
        sad = np.zeros(n, np.uint64) if per_b64 else None
        mv = np.zeros((n, 2), np.int16) if per_b64 else None
        _check(lib().svtme_dg_detector_hme(self._h, src_pic_id, ref_pic_id, C.byref(m), sad.ctypes.data if per_b64 else None,
                                           mv.ctypes.data if per_b64 else None), "svtme_dg_detector_hme")
        return m, sad, mv

    def set_concurrency(self, n):
        """2: consecutive batches alternate between two compute streams (await results with wait_batch / sync)"""
        lib().svtme_session_set_concurrency.argtypes = [C.c_void_p, C.c_int]
        lib().svtme_session_set_concurrency.restype = C.c_uint32
        _check(lib().svtme_session_set_concurrency(self._h, n), "svtme_session_set_concurrency")

    def batch_seq(self):
        """sequence number (1, 2, ...) of the batch queued last"""
        return int(lib().svtme_session_batch_seq(self._h))

    def wait_batch(self, seq):
        """block until the results of batch `seq` are in the host arrays it was given"""
        _check(lib().svtme_session_wait_batch(self._h, seq), "svtme_session_wait_batch")

    def download_plane(self, pic_id, level):
        pad = (72, 32, 16)[level]
        w, h = self.width >> level, self.height >> level
        buf = np.zeros(((h + 2 * pad), (w + 2 * pad)), dtype=np.uint8)
        st, rows, ox, oy = C.c_uint32(), C.c_uint32(), C.c_uint32(), C.c_uint32()
        _check(lib().svtme_picture_download_plane(self._h, pic_id, level, buf.ctypes.data, buf.nbytes, C.byref(st),
                                                  C.byref(rows), C.byref(ox), C.byref(oy)), "svtme_picture_download_plane")
        assert st.value == buf.shape[1] and rows.value == buf.shape[0], (st.value, rows.value, buf.shape)
        return buf

    @property
    def stream(self):
        return lib().svtme_session_stream(self._h)

    def set_stream(self, cuda_stream_ptr):
        _check(lib().svtme_session_set_stream(self._h, C.c_void_p(cuda_stream_ptr)), "svtme_session_set_stream")

    def enable_timing(self, on=True):
        _check(lib().svtme_session_enable_timing(self._h, 1 if on else 0), "svtme_session_enable_timing")

    def hme_state(self):
        n = abi.n_b64_of(self.width, self.height)
        out = np.zeros(n, dtype=abi.HME_STATE_DTYPE)
        _check(lib().svtme_fetch_hme_state(self._h, out.ctypes.data, out.nbytes), "svtme_fetch_hme_state")
        return out

    def sad_ops(self, reset=False):
        a, b = C.c_uint64(), C.c_uint64()
        _check(lib().svtme_session_sad_ops(self._h, 1 if reset else 0, C.byref(a), C.byref(b)), "svtme_session_sad_ops")
        return int(a.value), int(b.value)

    def sad_ops_by_stage(self, reset=False):
        """[HME stage total, full-pel, pre-HME, HME level 0, level 1, level 2]"""
        v = (C.c_uint64 * 6)()
        _check(lib().svtme_session_sad_ops_by_stage(self._h, 1 if reset else 0, v), "svtme_session_sad_ops_by_stage")
        return [int(x) for x in v]

    def kernel_times(self, reset=False):
        """{kernel name: device ms accumulated by the timed batches}"""
        ms = (C.c_double * 16)()
        _check(lib().svtme_session_kernel_times(self._h, 1 if reset else 0, ms, 16), "svtme_session_kernel_times")
        out = {}
        for tag in range(1, 16):
            name = lib().svtme_kernel_name(tag)
            if not name:
                break
            out[name.decode()] = ms[tag]
        return out

    def launch_count(self):
        return int(lib().svtme_session_launch_count(self._h))

    def kernel_time(self, reset=False):
        f, h, n = C.c_double(), C.c_double(), C.c_uint64()
        _check(lib().svtme_session_kernel_time(self._h, 1 if reset else 0, C.byref(f), C.byref(h), C.byref(n)),
               "svtme_session_kernel_time")
        return f.value, h.value, int(n.value)
