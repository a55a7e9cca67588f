/*
 * svtme_b200.h — C ABI of the B200 (sm_100a) open-loop motion-estimation library.
 *
 * This is the drop-in boundary for ONE stage of SVT-AV1-PSY 2.3.0-A: the open-loop
 * motion-estimation stage (pre-HME, HME level 0/1/2, full-pel 64x64 search, ME candidate
 * construction) that svt_aom_motion_estimation_kernel drives
 * (Source/Lib/Codec/me_process.c:97, Source/Lib/Codec/motion_estimation.c:3146).
 *
 * Everything here is plain C: pointers, sizes, PODs. No torch / C++ types cross the
 * boundary. All "file:line" citations are relative to the reference tree
 * (Source/Lib/... of gianni-rosato/svt-av1-psy 2.3.0-A).
 *
 * There is NO CPU fallback: every entry point returns SVTME_ERR_CUDA (and logs to stderr)
 * if the device is missing or a CUDA call fails.
 */
#ifndef SVTME_B200_H
#define SVTME_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- error convention: mirrors EbErrorType (Source/API/EbSvtAv1.h:127-142) ---- */
#define SVTME_OK 0                                  /* EB_ErrorNone */
#define SVTME_ERR_INSUFFICIENT_RESOURCES 0x80001000 /* EB_ErrorInsufficientResources */
#define SVTME_ERR_BAD_PARAMETER 0x80001005          /* EB_ErrorBadParameter */
#define SVTME_ERR_CUDA 0x80002000                   /* no reference counterpart: device failure */

/* ---- fixed dimensions of the MeContext result layout ---- */
#define SVTME_MAX_LIST 2   /* MAX_NUM_OF_REF_PIC_LIST */
#define SVTME_MAX_REF 4    /* REF_LIST_MAX_DEPTH / MAX_REF_IDX (definitions.h:2365) */
#define SVTME_PU_COUNT 85  /* SQUARE_PU_COUNT (me_sb_results.h:24) */
#define SVTME_SUB_SAD_SEARCH 0  /* definitions.h:2080 */
#define SVTME_FULL_SAD_SEARCH 1 /* definitions.h:2081 */

typedef struct SvtMeSearchArea { /* SearchArea, me_context.h:320-323 */
    uint16_t width;
    uint16_t height;
} SvtMeSearchArea;

typedef struct SvtMeSearchAreaMinMax { /* SearchAreaMinMax, me_context.h:324-327 */
    SvtMeSearchArea sa_min;
    SvtMeSearchArea sa_max;
} SvtMeSearchAreaMinMax;

/*
 * The ME controls that svt_aom_sig_deriv_me (enc_mode_config.c:694-843) writes into
 * MeContext (me_context.h:366-509), flattened. The host derives them exactly as the
 * reference does (the reference's own function can be used unchanged; svtme_sig_deriv_me()
 * in the host library is our C restatement of it) and hands them to the device as launch
 * parameters.
 */
typedef struct SvtMeCtrls {
    /* HME flags, me_context.h:427-432 */
    uint8_t enable_hme_flag;
    uint8_t enable_hme_level0_flag;
    uint8_t enable_hme_level1_flag;
    uint8_t enable_hme_level2_flag;
    uint8_t hme_search_method; /* SVTME_SUB_SAD_SEARCH / SVTME_FULL_SAD_SEARCH */
    uint8_t me_search_method;
    uint16_t num_hme_sa_w; /* always 2 (enc_mode_config.c:146) */
    uint16_t num_hme_sa_h;
    SvtMeSearchAreaMinMax hme_l0_sa; /* total HME-L0 area */
    SvtMeSearchArea       hme_l1_sa;
    SvtMeSearchArea       hme_l2_sa;
    SvtMeSearchAreaMinMax me_sa; /* full-pel search area min/max */
    /* PreHmeCtrls, me_context.h:336-341 */
    uint8_t               prehme_enable;
    uint8_t               prehme_skip_search_line;
    uint8_t               prehme_l1_early_exit;
    SvtMeSearchAreaMinMax prehme_sa_cfg[2];
    /* MeHmeRefPruneCtrls, me_context.h:280-290 */
    uint8_t  enable_me_hme_ref_pruning;
    uint16_t prune_ref_if_hme_sad_dev_bigger_than_th;
    uint16_t prune_ref_if_me_sad_dev_bigger_than_th;
    uint32_t zz_sad_th;
    uint16_t zz_sad_pct;
    uint32_t phme_sad_th;
    uint16_t phme_sad_pct;
    /* MeSrCtrls, me_context.h:292-305 */
    uint8_t  enable_me_sr_adjustment;
    uint16_t reduce_me_sr_based_on_mv_length_th;
    uint16_t stationary_hme_sad_abs_th;
    uint16_t stationary_me_sr_divisor;
    uint16_t reduce_me_sr_based_on_hme_sad_abs_th;
    uint16_t me_sr_divisor_for_low_hme_sad;
    uint8_t  distance_based_hme_resizing;
    /* Me8x8VarCtrls, me_context.h:310-318 */
    uint8_t  me_8x8_var_enabled;
    uint32_t me_sr_div4_th;
    uint32_t me_sr_div2_th;
    uint32_t me_sr_mult2_th;
    /* MvBasedSearchAdj, me_context.h:356-364 */
    uint8_t  mv_sa_adj_enabled;
    uint8_t  mv_sa_adj_nearest_ref_only;
    uint16_t mv_sa_adj_mv_size_th;
    uint16_t mv_sa_adj_sa_multiplier;
    /* scalars, me_context.h:489-508 */
    uint8_t  reduce_hme_l0_sr_th_min;
    uint8_t  reduce_hme_l0_sr_th_max;
    uint32_t me_early_exit_th;
    uint32_t me_safe_limit_zz_th;
    uint32_t prev_me_stage_based_exit_th;
    int32_t  prune_me_candidates_th;
    uint8_t  use_best_unipred_cand_only;
    /* the three bytes that used to be tail padding (sizeof stays 128) */
    uint8_t  me_type;        /* SVTME_ME_OPEN_LOOP / SVTME_ME_MCTF (EbMeType, me_context.h): temporal-filter ME, SURVEY 8f row f1 */
    uint16_t tf_me_exit_th;  /* me_ctx->tf_me_exit_th (uint16_t, me_context.h:495): MCTF skips the full-pel search of a b64 whose
                                HME SAD is below it (motion_estimation.c:3179-3183) */
} SvtMeCtrls;
#define SVTME_ME_OPEN_LOOP 0
#define SVTME_ME_MCTF 1

/*
 * Per-picture launch parameters: the PictureParentControlSet / SequenceControlSet fields the
 * ME path reads (motion_estimation.c, me_process.c:219-262) plus the reference-picture list.
 */
typedef struct SvtMePicParams {
    SvtMeCtrls ctrls;
    uint64_t picture_number; /* pcs->picture_number */
    uint8_t  num_of_list_to_search;       /* 1 for P, 2 for B (me_process.c:219-221) */
    uint8_t  num_of_ref_pic_to_search[2]; /* pcs->ref_list{0,1}_count_try */
    uint8_t  temporal_layer_index;
    uint8_t  is_ref;
    uint8_t  hierarchical_levels;
    uint8_t  similar_brightness_refs; /* pd_process.c:3444-3461 */
    uint8_t  enable_me_8x8;           /* pcs.c:1423 */
    uint8_t  enable_me_16x16;
    uint8_t  max_cand;   /* pa_me_data->max_cand (pd_process.c:3509) */
    uint8_t  max_refs;   /* pa_me_data->max_refs */
    uint8_t  max_l0;     /* pa_me_data->max_l0 */
    uint8_t  only_l_bwd; /* scs->mrp_ctrls.only_l_bwd */
    uint8_t  input_resolution; /* EbInputResolution, definitions.h:2088-2094 */
    uint8_t  gm_enabled;       /* pcs->gm_ctrls.enabled */
    uint8_t  gm_use_distance_based_active_th;
    /* reference pictures: ids of pictures previously uploaded with svtme_picture_upload(),
       [list][ref_idx]; and their display-order numbers (me_ds_ref_array[..].picture_number) */
    int32_t  ref_pic_id[SVTME_MAX_LIST][SVTME_MAX_REF];
    uint64_t ref_picture_number[SVTME_MAX_LIST][SVTME_MAX_REF];
} SvtMePicParams;

/* SearchResults (me_context.h:348-355), one per (b64, list, ref). */
typedef struct SvtMeSearchResult {
    int16_t  hme_sc_x;
    int16_t  hme_sc_y;
    uint8_t  do_ref;
    uint8_t  pad_[3];
    uint64_t hme_sad;
} SvtMeSearchResult;

/* Full-pel window actually searched for (b64, list, ref): used for the SAD-op count of
   SURVEY.md §8(d); all zero when the ref was pruned before the integer search. */
typedef struct SvtMeWindow {
    int16_t x_origin, y_origin; /* relative to the b64 origin */
    int16_t width, height;      /* search points */
} SvtMeWindow;

/*
 * Result arrays of one picture. Every pointer is caller-owned HOST memory (NULL = not wanted).
 * n_b64 = ceil(W/64)*ceil(H/64), raster order (me_process.c:177).
 *   n_pus   = enable_me_8x8 ? 85 : 21 (pcs.c:108-117)
 */
typedef struct SvtMePicResults {
    /* MeContext mirror, me_context.h:395-396: [n_b64][2][4][85], ME z-order PU index */
    uint32_t          *p_sb_best_sad;
    uint32_t          *p_sb_best_mv; /* (uint16 y << 16) | uint16 x, full-pel */
    SvtMeSearchResult *search_results; /* [n_b64][2][4] */
    SvtMeWindow       *me_windows;     /* [n_b64][2][4] */
    /* MeSbResults mirror, me_sb_results.h:44-51 */
    uint8_t  *total_me_candidate_index; /* [n_b64][n_pus] */
    uint32_t *me_mv_array;              /* [n_b64][n_pus * max_refs]  MvCandidate.as_int */
    uint8_t  *me_candidate_array;       /* [n_b64][n_pus * max_cand]  MeCandidate bitfield byte */
    /* per-b64 PCS statistics, pcs.h:835-845 */
    uint32_t *me_64x64_distortion;
    uint32_t *me_32x32_distortion;
    uint32_t *me_16x16_distortion;
    uint32_t *me_8x8_distortion;
    uint32_t *rc_me_distortion;
    uint32_t *me_8x8_cost_variance;
    uint8_t  *stationary_block_present_sb;
    uint8_t  *rc_me_allow_gm;
} SvtMePicResults;

typedef struct SvtMeSession SvtMeSession; /* opaque */

/* ------------------------------------------------------------------------------------------
 * 1. RTCD-signature kernels (known-answer parity path; one launch per call).
 *    Pointers are HOST pointers exactly as the reference's callers pass them; the library
 *    stages the touched bytes to the device, runs the sm_100a kernel and copies the outputs
 *    back. These can be stored straight into the reference's RTCD pointers.
 * ---------------------------------------------------------------------------------------- */

/* replaces svt_sad_loop_kernel (aom_dsp_rtcd.h:776; C: C_DEFAULT/compute_sad_c.c:58-101) */
void svtme_sad_loop_kernel(uint8_t *src, uint32_t src_stride, uint8_t *ref, uint32_t ref_stride,
                           uint32_t block_height, uint32_t block_width, uint64_t *best_sad,
                           int16_t *x_search_center, int16_t *y_search_center,
                           uint32_t src_stride_raw, uint8_t skip_search_line,
                           int16_t search_area_width, int16_t search_area_height);

/* replaces svt_ext_all_sad_calculation_8x8_16x16 (aom_dsp_rtcd.h:850; C: motion_estimation.c:335-362) */
void svtme_ext_all_sad_calculation_8x8_16x16(uint8_t *src, uint32_t src_stride, uint8_t *ref,
                                             uint32_t ref_stride, uint32_t mv,
                                             uint32_t *p_best_sad_8x8, uint32_t *p_best_sad_16x16,
                                             uint32_t *p_best_mv8x8, uint32_t *p_best_mv16x16,
                                             uint32_t p_eight_sad16x16[16][8],
                                             uint32_t p_eight_sad8x8[64][8], uint8_t sub_sad);

/* replaces svt_ext_eight_sad_calculation_32x32_64x64 (aom_dsp_rtcd.h:851; C: motion_estimation.c:369-425) */
void svtme_ext_eight_sad_calculation_32x32_64x64(uint32_t p_sad16x16[16][8],
                                                 uint32_t *p_best_sad_32x32,
                                                 uint32_t *p_best_sad_64x64,
                                                 uint32_t *p_best_mv32x32, uint32_t *p_best_mv64x64,
                                                 uint32_t mv, uint32_t p_sad32x32[4][8]);

/* replaces svt_ext_sad_calculation_8x8_16x16 (aom_dsp_rtcd.h:839; C: motion_estimation.c:98-164) */
void svtme_ext_sad_calculation_8x8_16x16(uint8_t *src, uint32_t src_stride, uint8_t *ref,
                                         uint32_t ref_stride, uint32_t *p_best_sad_8x8,
                                         uint32_t *p_best_sad_16x16, uint32_t *p_best_mv8x8,
                                         uint32_t *p_best_mv16x16, uint32_t mv,
                                         uint32_t *p_sad16x16, uint32_t *p_sad8x8, uint8_t sub_sad);

/* replaces svt_ext_sad_calculation_32x32_64x64 (aom_dsp_rtcd.h:845; C: motion_estimation.c:171-205) */
void svtme_ext_sad_calculation_32x32_64x64(uint32_t *p_sad16x16, uint32_t *p_best_sad_32x32,
                                           uint32_t *p_best_sad_64x64, uint32_t *p_best_mv32x32,
                                           uint32_t *p_best_mv64x64, uint32_t mv,
                                           uint32_t *p_sad32x32);

/* replaces svt_nxm_sad_kernel (aom_dsp_rtcd.h:854; C: C_DEFAULT/compute_sad_c.c:20-37,209-212) */
uint32_t svtme_nxm_sad_kernel(const uint8_t *src, uint32_t src_stride, const uint8_t *ref,
                              uint32_t ref_stride, uint32_t height, uint32_t width);

/* replaces svt_initialize_buffer_32bits (aom_dsp_rtcd.h:852; C: me_sad_calculation.c:14-17) */
void svtme_initialize_buffer_32bits(uint32_t *pointer, uint32_t count128, uint32_t count32,
                                    uint32_t value);

/* replaces svt_aom_downsample_2d (C: pic_analysis_process.c:131-161): 2x2 box decimation */
void svtme_downsample_2d(uint8_t *input_samples, uint32_t input_stride, uint32_t input_area_width,
                         uint32_t input_area_height, uint8_t *decim_samples, uint32_t decim_stride,
                         uint32_t decim_step);

/* No counterpart in the reference: the same search as svt_sad_loop_kernel, for callers that must
   not allocate GPU memory. Some users of this library run where cudaMalloc is forbidden or too
   costly: a real-time encoder thread (cudaMalloc locks the device and can stall other streams), a
   plug-in inside a host application that has reserved all device memory for itself, or code
   between cudaStreamBeginCapture and cudaStreamEndCapture. They keep their pictures in device
   memory already, so every pointer here is a DEVICE pointer (src, ref, and the three outputs),
   and the kernels are queued on `cuda_stream` (a cudaStream_t, NULL = default stream); the call
   returns without waiting. The scratch the search needs between its two kernels lives in an 8 KB
   array compiled into the library (no cudaMalloc, cudaFree or other allocating call is made).
   Limits: block at most 16x16, ref_stride equal to src_stride_raw or twice it, search area at
   most SVTME_NOALLOC_MAX_WIDTH wide and SVTME_NOALLOC_MAX_BANDS*32 rows; anything larger is
   refused with SVTME_ERR_BAD_PARAMETER and a message on stderr. All calls share the one scratch
   array: do not let two of them run on the GPU at the same time (other streams or threads). The
   first call on a device fetches the array's address from the CUDA runtime; issue that first call
   before starting any stream capture. Only the bytes svt_sad_loop_kernel itself reads are touched
   (block_width bytes per source row, search_area_width + block_width - 1 per reference row): the
   buffers may end flush with their allocation and need no alignment. Output semantics are those of svt_sad_loop_kernel:
   *d_best_sad = 0xffffff and the centres untouched when no position was evaluated. */
#define SVTME_NOALLOC_MAX_WIDTH 1024
#define SVTME_NOALLOC_MAX_BANDS 1024
uint32_t svtme_sad_loop_search_noalloc(const uint8_t *d_src, uint32_t src_stride, const uint8_t *d_ref,
                                       uint32_t ref_stride, uint32_t block_height, uint32_t block_width,
                                       uint64_t *d_best_sad, int16_t *d_x_search_center,
                                       int16_t *d_y_search_center, uint32_t src_stride_raw,
                                       uint8_t skip_search_line, int16_t search_area_width,
                                       int16_t search_area_height, void *cuda_stream);

/* Last error of the calling thread for the void-returning entry points above (SVTME_OK if none). */
uint32_t svtme_last_error(void);

/* ------------------------------------------------------------------------------------------
 * 2. Picture-batched path (throughput path): one session per GPU; source/reference luma
 *    planes and their 1/4 and 1/16 pyramids live in HBM; one call runs the whole ME stage
 *    of a picture (all b64, all lists/refs) as a fixed kernel sequence on a CUDA stream.
 *    Replaces the b64 loop of svt_aom_motion_estimation_kernel (me_process.c:174-291) and
 *    svt_aom_motion_estimation_b64 (motion_estimation.c:3146-3223) for ME_OPEN_LOOP.
 * ---------------------------------------------------------------------------------------- */

/* Create a session on CUDA device `device`. width/height: luma size (multiples of 8, as the
   reference guarantees after svt_aom_pad_picture_to_multiple_of_min_blk_size_dimensions);
   max_pictures: capacity of the device picture store (source + reference pictures alive). */
uint32_t svtme_session_create(SvtMeSession **out, int device, uint32_t width, uint32_t height,
                              uint32_t max_pictures);
void     svtme_session_destroy(SvtMeSession *s);

/* Upload the 8-bit luma plane of one picture (interior only, `stride` bytes per row, HOST
   memory; pinned memory makes the copy asynchronous). The device builds the replicate-padded
   full-resolution plane (pad 72, enc_handle.c:4485-4497), the 1/4 (pad 32) and 1/16 (pad 16)
   planes (pic_analysis_process.c:2139-2196, pic_operators.c:338-384). pic_id is the caller's
   handle (0 <= pic_id < max_pictures); re-uploading an id replaces its planes (this is how the
   post-temporal-filter refresh, temporal_filtering.c:3842-3882, is expressed). */
uint32_t svtme_picture_upload(SvtMeSession *s, int32_t pic_id, const uint8_t *luma,
                              uint32_t stride);

/* 10-bit input: `luma16` holds 10-bit samples in uint16 (y4m C420p10 layout). ME reads only the
   8 MSBs (reference_object.c:327, pack_unpack_c.c:77-116); the split is done on the device. */
uint32_t svtme_picture_upload_10bit(SvtMeSession *s, int32_t pic_id, const uint16_t *luma16,
                                    uint32_t stride_in_samples);

/* Same as svtme_picture_upload but the luma plane is already in DEVICE memory. */
uint32_t svtme_picture_upload_device(SvtMeSession *s, int32_t pic_id, const uint8_t *d_luma,
                                     uint32_t stride);

/* The same uploads for several pictures at once (n <= 40): the pyramids of all of them are built by two kernel
   launches. Host frames may have any pitch; device frames must be dense (stride == width). */
uint32_t svtme_picture_upload_batch(SvtMeSession *s, uint32_t n, const int32_t *pic_ids, const uint8_t *const *luma,
                                    uint32_t stride);
uint32_t svtme_picture_upload_10bit_batch(SvtMeSession *s, uint32_t n, const int32_t *pic_ids,
                                          const uint16_t *const *luma16, uint32_t stride_in_samples);
uint32_t svtme_picture_upload_device_batch(SvtMeSession *s, uint32_t n, const int32_t *pic_ids,
                                           const uint8_t *const *d_luma, uint32_t stride);

/* Run open-loop ME of picture pic_id against params->ref_pic_id[][] and copy the requested
   result arrays to host (asynchronously on the session stream; call svtme_session_sync, or
   pass sync=1). */
uint32_t svtme_picture_me(SvtMeSession *s, int32_t pic_id, const SvtMePicParams *params,
                          SvtMePicResults *results, int sync);

/* The same for n pictures at once: every kernel of the stage is launched once for the whole batch
   (grid.y = picture), which is how a mini-GOP is meant to be submitted — Picture Decision releases
   all pictures of a mini-GOP together (pd_process.c:4373-4400) and open-loop ME pictures do not
   depend on each other. pic_ids[i] / params[i] / results[i] belong to picture i (results may be NULL).
   All pictures of a batch must share the HME stage structure (same enable_hme_* flags, pre-HME on/off,
   search methods), which holds inside one encode. */
uint32_t svtme_picture_me_batch(SvtMeSession *s, uint32_t n, const int32_t *pic_ids, const SvtMePicParams *params,
                                SvtMePicResults *results, int sync);
/* Copy the device results of picture `index` of the last batch. */
uint32_t svtme_fetch_results_batch(SvtMeSession *s, uint32_t index, SvtMePicResults *results, int sync);

/* ---- packed results: one download per batch ------------------------------------------------------------------------
   The arrays of SvtMePicResults are what the rest of the encoder reads after svt_aom_motion_estimation_b64 returns:
     SVTME_F_SB_RESULTS      total_me_candidate_index, me_mv_array, me_candidate_array = pcs->pa_me_data->me_results[b64]
                             (me_sb_results.h:44-51; consumers: src_ops_process.c (TPL), mode decision)
     SVTME_F_STATS           the eight per-b64 PCS arrays compute_distortion / perform_gm_detection write
                             (motion_estimation.c:3034-3077, :2908-3031)
     SVTME_F_SEARCH_RESULTS  me_ctx->search_results (me_context.h:348-355)
     SVTME_F_BEST            me_ctx->p_sb_best_sad / p_sb_best_mv (me_context.h:395-396)
     SVTME_F_WINDOWS         the full-pel windows searched (no reference counterpart; SAD-op accounting)
   Outside motion_estimation.c the MeContext members (SEARCH_RESULTS, BEST) are read only by the temporal filter
   (temporal_filtering.c:1860-2226, :2672: ME_MCTF); open-loop ME's caller (me_process.c:266-291) reads none of them, so
   an open-loop consumer needs SB_RESULTS | STATS only.
   svtme_picture_me_batch_packed runs the batch like svtme_picture_me_batch and lands the requested fields of all its
   pictures in ONE caller-owned host slab with ONE device-to-host copy (pinned memory makes it asynchronous): picture i's
   block starts at the sum of the block sizes of pictures 0..i-1 (svtme_packed_layout of each picture's params; every
   block size is a multiple of 256 bytes) and holds the arrays at the offsets the layout gives, each array exactly as
   SvtMePicResults documents it (the host's own strides: n_pus * max_refs, n_pus * max_cand). Fields that are not
   requested are not written by the kernels at all. svtme_packed_view fills a SvtMePicResults with pointers into a block. */
#define SVTME_F_SB_RESULTS 1u
#define SVTME_F_STATS 2u
#define SVTME_F_SEARCH_RESULTS 4u
#define SVTME_F_BEST 8u
#define SVTME_F_WINDOWS 16u
#define SVTME_F_ALL 31u
/* With SVTME_F_SB_RESULTS: me_mv_array and me_candidate_array cross PCIe at the strides THIS picture can fill instead of
   the sequence-wide max_refs / max_cand the reference allocates (pcs.c:91-96): a row of me_mv_array holds the picture's own
   n0 + n1 references (list 1 starts at n0), a row of me_candidate_array the most candidates construct_me_candidate_array*
   can emit for its reference counts (n0 + n1 unipred + the bipred pairs only_l_bwd allows, motion_estimation.c:2602-2905):
   8 instead of 23 bytes per PU at presets 0-4. svtme_sb_strides reports the strides; svtme_sb_expand_b64 writes one
   b64's rows at the reference's strides (what a consumer does anyway when it copies the rows into pa_me_data). */
#define SVTME_F_SB_COMPACT 32u
typedef struct SvtMeSbStrides {
    uint8_t mv_stride;    /* entries per PU row of me_mv_array */
    uint8_t mv_l1_offset; /* first list-1 entry inside a row */
    uint8_t cand_stride;  /* bytes per PU row of me_candidate_array */
    uint8_t n_pus;
} SvtMeSbStrides;
void svtme_sb_strides(const SvtMePicParams *params, uint32_t fields, SvtMeSbStrides *out);
/* dst_* : one b64's arrays at the reference's strides (n_pus, n_pus * max_refs, n_pus * max_cand entries; zero filled
   first); src: the picture-level arrays of a block produced with `fields` */
void svtme_sb_expand_b64(const SvtMePicParams *params, uint32_t fields, const SvtMePicResults *src, uint32_t b64_index,
                         uint8_t *dst_total, uint32_t *dst_mv, uint8_t *dst_cand);
typedef struct SvtMePackedLayout {
    size_t bytes;   /* size of one picture's block */
    size_t off[15]; /* byte offset of each SvtMePicResults member, in member order; (size_t)-1 = not part of the block */
} SvtMePackedLayout;
uint32_t svtme_packed_layout(uint32_t width, uint32_t height, const SvtMePicParams *params, uint32_t fields,
                             SvtMePackedLayout *out); /* fields may include SVTME_F_SB_COMPACT */
void     svtme_packed_view(const SvtMePackedLayout *layout, void *block, SvtMePicResults *out);
uint32_t svtme_picture_me_batch_packed(SvtMeSession *s, uint32_t n, const int32_t *pic_ids, const SvtMePicParams *params,
                                       uint32_t fields, void *host_slab, size_t host_slab_bytes, int sync);
/* Fields produced by calls that name no host arrays (svtme_picture_me_device, results == NULL): default SVTME_F_ALL. */
uint32_t svtme_session_set_result_fields(SvtMeSession *s, uint32_t fields);
/* Work granularity of the 1/16-plane searches (pre-HME, HME level 0): 0 = one queue unit per (b64, list, ref), its jobs and
   tiles searched by one warp (large batches: least per-unit work); 1 = one unit per job (small batches: a single picture
   has too few (b64, list, ref) groups to occupy every warp of the GPU); -1 = chosen per batch from its size (default).
   Results do not depend on it. */
uint32_t svtme_session_set_unit_granularity(SvtMeSession *s, int mode);
/* Size the unit queues, result slabs and job arrays for batches of up to n_pictures pictures with these controls now,
   instead of growing them (with a pipeline drain) when the first such batch arrives. */
uint32_t svtme_session_reserve(SvtMeSession *s, const SvtMePicParams *params, uint32_t n_pictures);

/* ---- device-resident view ---------------------------------------------------------------------
   Later stages of the encoder re-read what already sits in HBM: global-motion pre-processing reads the three planes of a
   picture and of its references (global_me.c:62-110, :309), the TPL dispenser reads the ME vectors and candidate lists
   of pa_me_data (src_ops_process.c:205-299, :605-616). These two calls hand such a consumer DEVICE pointers instead of a
   second upload / a round trip through the host.
   svtme_picture_device_planes: the padded planes of an uploaded picture (level 0 full, 1 quarter, 2 sixteenth):
   `origin` = interior sample (0,0), rows `stride` bytes apart, `pad` replicated samples on every side. Valid until pic_id
   is uploaded again or the session is destroyed; work reading them must be ordered after the upload (the session's
   stream, or after svtme_session_sync).
   svtme_batch_results_acquire: device pointers (same arrays, same layout as SvtMePicResults; NULL = field not produced)
   to the results of picture `index` of batch `seq`, and makes `consumer_stream` (a cudaStream_t) wait for the batch's
   kernels. The session keeps two result sets: the arrays stay intact until batch seq + 2 is launched;
   svtme_batch_results_release(seq, consumer_stream) makes that later batch wait for the work queued on consumer_stream
   so far. SVTME_ERR_BAD_PARAMETER if the batch's set was already reused. */
typedef struct SvtMeDevicePlane {
    const uint8_t *origin;
    uint32_t       stride, width, height, pad;
} SvtMeDevicePlane;
uint32_t svtme_picture_device_planes(SvtMeSession *s, int32_t pic_id, SvtMeDevicePlane planes[3]);
uint32_t svtme_batch_results_acquire(SvtMeSession *s, uint64_t seq, uint32_t index, SvtMePicResults *device_ptrs,
                                     void *consumer_stream);
uint32_t svtme_batch_results_release(SvtMeSession *s, uint64_t seq, void *consumer_stream);
/* Page-locked host memory for frames and result slabs (copies from / to it are asynchronous and run at full PCIe rate);
   for callers that do not link the CUDA runtime themselves. NULL on failure. A CUDA device must exist. */
void *svtme_host_alloc(size_t bytes);
void  svtme_host_free(void *p);
/* Convenience for callers without the CUDA runtime (tests): synchronous copy of device memory to the host. */
uint32_t svtme_copy_from_device(SvtMeSession *s, void *dst_host, const void *src_device, size_t bytes);

/* Same, but results stay on the device (read them back with svtme_fetch_results, or in place through
   svtme_batch_results_acquire). Used to time the kernels with inputs and outputs resident in HBM. */
uint32_t svtme_picture_me_device(SvtMeSession *s, int32_t pic_id, const SvtMePicParams *params);

/* Copy the device result arrays of the last svtme_picture_me_device call into `results`. */
uint32_t svtme_fetch_results(SvtMeSession *s, SvtMePicResults *results, int sync);

/* Wait for everything queued on the session: host uploads, ME kernels, result downloads. */
uint32_t svtme_session_sync(SvtMeSession *s);

/* Pipelining across batches. The session runs host uploads, the ME kernels and the result downloads on three
   streams: with sync = 0, svtme_picture_me_batch returns once the batch is queued, and the uploads of the next
   pictures and the next batch's kernels overlap its downloads. svtme_session_batch_seq returns the sequence number
   (1, 2, ...) of the batch queued last; svtme_session_wait_batch blocks until the results of batch `seq` are in the
   host arrays it was given (the batch's own download, not a later batch's). The caller rotates its host arrays: with
   k sets it waits for batch seq - k before it hands a set out again. Three sets keep both compute streams of
   svtme_session_set_concurrency(2) busy: the device-side result region of batch seq - 2 is reused by batch seq, and the
   library makes the first kernel that writes it wait for that download on the device, so the host need not. (With two
   sets the host itself waits for the download of batch seq - 2 before it can queue batch seq.) No reference counterpart:
   the reference's ME threads synchronise through the picture-decision / ME FIFOs (me_process.c:75-330). */
uint64_t svtme_session_batch_seq(SvtMeSession *s);
/* n = 2: consecutive batches run on two compute streams (the session's stream and an internal one), so that the
   latency-bound control kernels of one batch overlap the searches of the other. Results must then be awaited with
   svtme_session_wait_batch / svtme_session_sync (work queued on the session's stream is no longer ordered after every
   batch), and device buffers given to svtme_picture_upload_device* must be complete when the call is made. n = 1 (the
   default) keeps everything on the session's stream. While kernel timing is enabled one stream is used. */
uint32_t svtme_session_set_concurrency(SvtMeSession *s, int n);
uint32_t svtme_session_wait_batch(SvtMeSession *s, uint64_t seq);

/* Read back a device plane (level 0 = full padded, 1 = quarter, 2 = sixteenth) into host memory,
   including padding; *stride / *rows receive the padded geometry. For parity tests. */
uint32_t svtme_picture_download_plane(SvtMeSession *s, int32_t pic_id, int level, uint8_t *dst,
                                      size_t dst_capacity, uint32_t *stride, uint32_t *rows,
                                      uint32_t *org_x, uint32_t *org_y);

/* The CUDA stream the session launches on (cudaStream_t as void*), so a caller can record
   events on it; svtme_session_set_stream lets the caller provide one. */
void    *svtme_session_stream(SvtMeSession *s);
uint32_t svtme_session_set_stream(SvtMeSession *s, void *cuda_stream);

/* Number of kernel launches issued by this session since creation (bench "gpu_launches"). */
uint64_t svtme_session_launch_count(SvtMeSession *s);

/* Switch per-stage CUDA-event timing of svtme_picture_me* on or off (off by default: each timed
   call records three events on the session stream). */
uint32_t svtme_session_enable_timing(SvtMeSession *s, int on);

/* Device time (ms, CUDA events on the session stream) of the dominant full-pel search kernel
   accumulated since the last call with reset=1, and the number of launches it covers. */
uint32_t svtme_session_kernel_time(SvtMeSession *s, int reset, double *fullpel_ms,
                                   double *hme_ms, uint64_t *n_launch);

/* Per-kernel device time (ms) accumulated by the timed batches since the last reset: ms[tag] for tag = 1 .. the first
   tag whose svtme_kernel_name is NULL (CUDA events behind every launch of the sequence, on the launching stream). */
const char *svtme_kernel_name(int tag);
uint32_t    svtme_session_kernel_times(SvtMeSession *s, int reset, double *ms, int n_ms);

/* How k_search16 cuts a 1/16-plane search window of sa_w x sa_h positions into tiles and which evaluator takes each kind
   of tile: the plan k_hme_ctl makes on the device, computed on the host (no GPU needed; tests, diagnostics). block16x8: the
   block is 16 wide and 8 rows two apart (the production shape of the sub-sampled SAD search); level0: the launch that also
   has the two-row blocked evaluator. out[0..3] = tile width, tile height, tiles per row, tile rows; out[4], out[5] =
   evaluator of a full tile and of a tile of the last tile row (0 general, 1 two-row blocked, 2 three-row blocked, 3 one
   position per lane); out[6] = the cost model's warp instructions for the whole window. No reference counterpart. */
void svtme_s16_tile_plan(int sa_w, int sa_h, int block16x8, int level0, int out[7]);

/* Algorithmic work done by the kernels since the last reset: "SAD operations" = search positions x
   block pixels compared (|a-b| and accumulate), the unit of SURVEY.md 8(d) / BASELINE.md. Counted on
   the device by the thread that sets up each search. */
uint32_t svtme_session_sad_ops(SvtMeSession *s, int reset, uint64_t *hme_ops, uint64_t *fullpel_ops);
/* The same split by stage: [0] HME stage total, [1] full-pel, [2] pre-HME, [3] HME level 0, [4] level 1, [5] level 2. */
uint32_t svtme_session_sad_ops_by_stage(SvtMeSession *s, int reset, uint64_t ops[6]);

/* Per-b64 HME-stage state of the last svtme_picture_me* call (parity debugging): n_b64 records of
   SvtMeHmeState. No reference counterpart as an array; the fields mirror MeContext members
   (me_context.h:435-452, :329-334, :348-355). */
typedef struct SvtMeHmeState {
    int16_t  x_hme[3][2][4][2][2]; /* [level][list][ref][sr_w][sr_h] x_hme_levelN_search_center */
    int16_t  y_hme[3][2][4][2][2];
    uint64_t hme_sad[3][2][4][2][2];
    int16_t  prehme_mv_x[2][4][2], prehme_mv_y[2][4][2];
    uint64_t prehme_sad[2][4][2];
    uint32_t zz_sad[2][4];
    uint32_t reduce_me_sr_divisor[2][4];
    int16_t  hme_sc_x[2][4], hme_sc_y[2][4]; /* before me_prune_ref */
    uint64_t sr_hme_sad[2][4];
    uint8_t  do_ref[2][4];
    uint8_t  performed_phme[2][4][2];
    uint8_t  prehme_valid[2][4][2];
    uint8_t  tf_exit;  /* ME_MCTF: the HME SAD of (0,0) was below tf_me_exit_th, no full-pel search (motion_estimation.c:3179-3183) */
    uint8_t  pad_[7];
} SvtMeHmeState;
uint32_t svtme_fetch_hme_state(SvtMeSession *s, void *dst, size_t dst_bytes);

/* --- 3b. Dynamic-GOP detector HME ---------------------------------------------------------
   Replaces dg_detector_hme_level0 + early_hme_b64 (Codec/pd_process.c:392-589): for every b64 one svt_sad_loop_kernel
   search of its 16x16 block of the 1/16 picture (FULL_SAD_SEARCH, window 16 / 64 / 128 square by input resolution,
   :494-495) in the 1/16 picture of `ref_pic_id`, and the DGDetectorMetrics sums (Codec/pcs.h:696-702) over the picture.
   Both pictures must have been uploaded. b64_sad / b64_mv (col, row per b64, 1/4-pel units as in the reference:
   x4) may be NULL. Synchronous. */
typedef struct SvtMeDgMetrics {
    uint64_t tot_dist;        /* sum of the best SADs */
    uint32_t tot_cplx;        /* b64s with SAD > 16*16*30 */
    uint32_t tot_active;      /* b64s with a non-zero vector */
    int32_t  sum_in_vectors;  /* inward minus outward pointing vector components */
    uint32_t input_resolution; /* in: EbInputResolution class (0 = <=240p .. 6 = 8K); svtme_sig_deriv_me derives it */
} SvtMeDgMetrics;
uint32_t svtme_dg_detector_hme(SvtMeSession *s, int32_t src_pic_id, int32_t ref_pic_id, SvtMeDgMetrics *metrics,
                               uint64_t *b64_sad, int16_t *b64_mv);

/* sizeof() of the ABI structs as the library was compiled (0 SvtMeCtrls, 1 SvtMePicParams,
   2 SvtMeSearchResult, 3 SvtMeWindow, 4 SvtMePicResults, 5 SvtMeHmeState): binding self-check. */
uint32_t svtme_abi_sizeof(int what);

/* Library/build identification string. */
const char *svtme_version(void);

#ifdef __cplusplus
}
#endif
#endif /* SVTME_B200_H */
