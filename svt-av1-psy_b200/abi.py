"""ctypes mirror of include/svtme_b200.h (the C ABI of libsvtme_b200.so).

Field order and types follow the header exactly; tests/test_abi.py checks sizeof() of every
struct against the values the C compiler reports (svtme_abi_sizeof_*)."""
import ctypes as C

MAX_LIST = 2
MAX_REF = 4
PU_COUNT = 85
SUB_SAD_SEARCH = 0
FULL_SAD_SEARCH = 1

OK = 0
ERR_INSUFFICIENT_RESOURCES = 0x80001000
ERR_BAD_PARAMETER = 0x80001005
ERR_CUDA = 0x80002000


class SearchArea(C.Structure):
    _fields_ = [("width", C.c_uint16), ("height", C.c_uint16)]


class SearchAreaMinMax(C.Structure):
    _fields_ = [("sa_min", SearchArea), ("sa_max", SearchArea)]


class MeCtrls(C.Structure):
    _fields_ = [
        ("enable_hme_flag", C.c_uint8),
        ("enable_hme_level0_flag", C.c_uint8),
        ("enable_hme_level1_flag", C.c_uint8),
        ("enable_hme_level2_flag", C.c_uint8),
        ("hme_search_method", C.c_uint8),
        ("me_search_method", C.c_uint8),
        ("num_hme_sa_w", C.c_uint16),
        ("num_hme_sa_h", C.c_uint16),
        ("hme_l0_sa", SearchAreaMinMax),
        ("hme_l1_sa", SearchArea),
        ("hme_l2_sa", SearchArea),
        ("me_sa", SearchAreaMinMax),
        ("prehme_enable", C.c_uint8),
        ("prehme_skip_search_line", C.c_uint8),
        ("prehme_l1_early_exit", C.c_uint8),
        ("prehme_sa_cfg", SearchAreaMinMax * 2),
        ("enable_me_hme_ref_pruning", C.c_uint8),
        ("prune_ref_if_hme_sad_dev_bigger_than_th", C.c_uint16),
        ("prune_ref_if_me_sad_dev_bigger_than_th", C.c_uint16),
        ("zz_sad_th", C.c_uint32),
        ("zz_sad_pct", C.c_uint16),
        ("phme_sad_th", C.c_uint32),
        ("phme_sad_pct", C.c_uint16),
        ("enable_me_sr_adjustment", C.c_uint8),
        ("reduce_me_sr_based_on_mv_length_th", C.c_uint16),
        ("stationary_hme_sad_abs_th", C.c_uint16),
        ("stationary_me_sr_divisor", C.c_uint16),
        ("reduce_me_sr_based_on_hme_sad_abs_th", C.c_uint16),
        ("me_sr_divisor_for_low_hme_sad", C.c_uint16),
        ("distance_based_hme_resizing", C.c_uint8),
        ("me_8x8_var_enabled", C.c_uint8),
        ("me_sr_div4_th", C.c_uint32),
        ("me_sr_div2_th", C.c_uint32),
        ("me_sr_mult2_th", C.c_uint32),
        ("mv_sa_adj_enabled", C.c_uint8),
        ("mv_sa_adj_nearest_ref_only", C.c_uint8),
        ("mv_sa_adj_mv_size_th", C.c_uint16),
        ("mv_sa_adj_sa_multiplier", C.c_uint16),
        ("reduce_hme_l0_sr_th_min", C.c_uint8),
        ("reduce_hme_l0_sr_th_max", C.c_uint8),
        ("me_early_exit_th", C.c_uint32),
        ("me_safe_limit_zz_th", C.c_uint32),
        ("prev_me_stage_based_exit_th", C.c_uint32),
        ("prune_me_candidates_th", C.c_int32),
        ("use_best_unipred_cand_only", C.c_uint8),
        ("me_type", C.c_uint8),
        ("tf_me_exit_th", C.c_uint16),
    ]


class MePicParams(C.Structure):
    _fields_ = [
        ("ctrls", MeCtrls),
        ("picture_number", C.c_uint64),
        ("num_of_list_to_search", C.c_uint8),
        ("num_of_ref_pic_to_search", C.c_uint8 * 2),
        ("temporal_layer_index", C.c_uint8),
        ("is_ref", C.c_uint8),
        ("hierarchical_levels", C.c_uint8),
        ("similar_brightness_refs", C.c_uint8),
        ("enable_me_8x8", C.c_uint8),
        ("enable_me_16x16", C.c_uint8),
        ("max_cand", C.c_uint8),
        ("max_refs", C.c_uint8),
        ("max_l0", C.c_uint8),
        ("only_l_bwd", C.c_uint8),
        ("input_resolution", C.c_uint8),
        ("gm_enabled", C.c_uint8),
        ("gm_use_distance_based_active_th", C.c_uint8),
        ("ref_pic_id", (C.c_int32 * MAX_REF) * MAX_LIST),
        ("ref_picture_number", (C.c_uint64 * MAX_REF) * MAX_LIST),
    ]


class MeSearchResult(C.Structure):
    _fields_ = [
        ("hme_sc_x", C.c_int16),
        ("hme_sc_y", C.c_int16),
        ("do_ref", C.c_uint8),
        ("pad_", C.c_uint8 * 3),
        ("hme_sad", C.c_uint64),
    ]


class MeWindow(C.Structure):
    _fields_ = [("x_origin", C.c_int16), ("y_origin", C.c_int16), ("width", C.c_int16), ("height", C.c_int16)]


class DgMetrics(C.Structure):
    """SvtMeDgMetrics (include/svtme_b200.h): DGDetectorMetrics of pcs.h:696-702 + the resolution class as input"""
    _fields_ = [("tot_dist", C.c_uint64), ("tot_cplx", C.c_uint32), ("tot_active", C.c_uint32),
                ("sum_in_vectors", C.c_int32), ("input_resolution", C.c_uint32)]

    def as_tuple(self):
        return (self.tot_dist, self.tot_cplx, self.tot_active, self.sum_in_vectors)


def input_resolution_of(width, height):
    """EbInputResolution class (definitions.h:2088-2094, thresholds :2060-2066)"""
    n = width * height
    for cls, th in enumerate((0x28500, 0x4CE00, 0xA1400, 0x16DA00, 0x535200, 0x140A000)):
        if n < th:
            return cls
    return 6


class MePicResults(C.Structure):
    _fields_ = [
        ("p_sb_best_sad", C.c_void_p),
        ("p_sb_best_mv", C.c_void_p),
        ("search_results", C.c_void_p),
        ("me_windows", C.c_void_p),
        ("total_me_candidate_index", C.c_void_p),
        ("me_mv_array", C.c_void_p),
        ("me_candidate_array", C.c_void_p),
        ("me_64x64_distortion", C.c_void_p),
        ("me_32x32_distortion", C.c_void_p),
        ("me_16x16_distortion", C.c_void_p),
        ("me_8x8_distortion", C.c_void_p),
        ("rc_me_distortion", C.c_void_p),
        ("me_8x8_cost_variance", C.c_void_p),
        ("stationary_block_present_sb", C.c_void_p),
        ("rc_me_allow_gm", C.c_void_p),
    ]


F_SB_RESULTS, F_STATS, F_SEARCH_RESULTS, F_BEST, F_WINDOWS, F_ALL = 1, 2, 4, 8, 16, 31
F_SB_COMPACT = 32  # me_mv_array / me_candidate_array rows at the picture's own strides (SbStrides)
ABSENT = (1 << (8 * C.sizeof(C.c_size_t))) - 1


class PackedLayout(C.Structure):
    """SvtMePackedLayout: block size and the offset of each SvtMePicResults member (ABSENT = not part of the block)"""
    _fields_ = [("bytes", C.c_size_t), ("off", C.c_size_t * 15)]


class SbStrides(C.Structure):
    """SvtMeSbStrides"""
    _fields_ = [("mv_stride", C.c_uint8), ("mv_l1_offset", C.c_uint8), ("cand_stride", C.c_uint8), ("n_pus", C.c_uint8)]


class DevicePlane(C.Structure):
    """SvtMeDevicePlane"""
    _fields_ = [("origin", C.c_void_p), ("stride", C.c_uint32), ("width", C.c_uint32), ("height", C.c_uint32), ("pad", C.c_uint32)]


# numpy dtypes of the packed result records
import numpy as np  # noqa: E402

SEARCH_RESULT_DTYPE = np.dtype(
    [("hme_sc_x", "<i2"), ("hme_sc_y", "<i2"), ("do_ref", "u1"), ("pad_", "u1", (3,)), ("hme_sad", "<u8")]
)
WINDOW_DTYPE = np.dtype([("x_origin", "<i2"), ("y_origin", "<i2"), ("width", "<i2"), ("height", "<i2")])
assert SEARCH_RESULT_DTYPE.itemsize == C.sizeof(MeSearchResult) == 16
assert WINDOW_DTYPE.itemsize == C.sizeof(MeWindow) == 8


HME_STATE_DTYPE = np.dtype([
    ("x_hme", "<i2", (3, 2, 4, 2, 2)), ("y_hme", "<i2", (3, 2, 4, 2, 2)), ("hme_sad", "<u8", (3, 2, 4, 2, 2)),
    ("prehme_mv_x", "<i2", (2, 4, 2)), ("prehme_mv_y", "<i2", (2, 4, 2)), ("prehme_sad", "<u8", (2, 4, 2)),
    ("zz_sad", "<u4", (2, 4)), ("reduce_me_sr_divisor", "<u4", (2, 4)),
    ("hme_sc_x", "<i2", (2, 4)), ("hme_sc_y", "<i2", (2, 4)), ("sr_hme_sad", "<u8", (2, 4)), ("do_ref", "u1", (2, 4)),
    ("performed_phme", "u1", (2, 4, 2)), ("prehme_valid", "u1", (2, 4, 2)), ("tf_exit", "u1"), ("pad_", "u1", (7,)),
], align=True)


def n_b64_of(width, height):
    return ((width + 63) // 64) * ((height + 63) // 64)


class PicResultArrays:
    """Owns numpy arrays laid out as SvtMePicResults expects and the ctypes struct pointing at them."""

    FIELDS = (
        "p_sb_best_sad", "p_sb_best_mv", "search_results", "me_windows", "total_me_candidate_index",
        "me_mv_array", "me_candidate_array", "me_64x64_distortion", "me_32x32_distortion",
        "me_16x16_distortion", "me_8x8_distortion", "rc_me_distortion", "me_8x8_cost_variance",
        "stationary_block_present_sb", "rc_me_allow_gm",
    )

    @staticmethod
    def shapes_of(width, height, params, strides=None):
        """strides: SbStrides of a compact block (rows of the two MeSbResults arrays at the picture's own strides)"""
        n = n_b64_of(width, height)
        n_pus = PU_COUNT if params.enable_me_8x8 else 21
        mv_stride = strides.mv_stride if strides is not None else max(1, params.max_refs)
        cand_stride = strides.cand_stride if strides is not None else max(1, params.max_cand)
        return n, n_pus, {
            "p_sb_best_sad": ((n, 2, 4, 85), np.uint32),
            "p_sb_best_mv": ((n, 2, 4, 85), np.uint32),
            "search_results": ((n, 2, 4), SEARCH_RESULT_DTYPE),
            "me_windows": ((n, 2, 4), WINDOW_DTYPE),
            "total_me_candidate_index": ((n, n_pus), np.uint8),
            "me_mv_array": ((n, n_pus * mv_stride), np.uint32),
            "me_candidate_array": ((n, n_pus * cand_stride), np.uint8),
            "me_64x64_distortion": ((n,), np.uint32),
            "me_32x32_distortion": ((n,), np.uint32),
            "me_16x16_distortion": ((n,), np.uint32),
            "me_8x8_distortion": ((n,), np.uint32),
            "rc_me_distortion": ((n,), np.uint32),
            "me_8x8_cost_variance": ((n,), np.uint32),
            "stationary_block_present_sb": ((n,), np.uint8),
            "rc_me_allow_gm": ((n,), np.uint8),
        }

    def __init__(self, width, height, params, pinned=False, block=None, layout=None, strides=None):
        """block + layout: views into one picture's block of a packed result slab (svtme_picture_me_batch_packed) instead
        of arrays of their own; members the layout does not hold are None; strides: the block is compact"""
        n, n_pus, shapes = self.shapes_of(width, height, params, strides)
        self.n_b64, self.n_pus = n, n_pus
        self.struct = MePicResults()
        self._keep = []
        if block is not None:
            for i, name in enumerate(self.FIELDS):
                shape, dt = shapes[name]
                if layout.off[i] == ABSENT:
                    setattr(self, name, None)
                    continue
                nbytes = int(np.prod(shape)) * np.dtype(dt).itemsize
                arr = block[layout.off[i]:layout.off[i] + nbytes].view(dt).reshape(shape)
                setattr(self, name, arr)
                setattr(self.struct, name, arr.ctypes.data)
            return
        self._keep = []
        for name, (shape, dt) in shapes.items():
            if pinned:
                import torch
                nbytes = int(np.prod(shape)) * np.dtype(dt).itemsize
                t = torch.zeros(nbytes, dtype=torch.uint8).pin_memory()
                self._keep.append(t)
                arr = t.numpy().view(dt).reshape(shape)
            else:
                arr = np.zeros(shape, dtype=dt)
            setattr(self, name, arr)
            setattr(self.struct, name, arr.ctypes.data)

    def nbytes(self):
        return sum(getattr(self, f).nbytes for f in self.FIELDS)

    def snapshot(self):
        """plain copies of the arrays (the originals may be reused by a later batch)"""
        import types
        out = types.SimpleNamespace(FIELDS=self.FIELDS, n_b64=self.n_b64, n_pus=self.n_pus)
        for f in self.FIELDS:
            setattr(out, f, getattr(self, f).copy())
        return out
