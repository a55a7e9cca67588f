/*
 * ref_harness.h — TEST INFRASTRUCTURE ONLY (never linked into the product).
 *
 * C interface of oracle/_ref/libsvtref.so: the UNMODIFIED reference sources of the ME path
 * (compiled where they lie under /root/reference by oracle/Makefile) plus ref_harness.c, a
 * driver that fakes the few PictureParentControlSet / SequenceControlSet fields the path reads
 * and then calls the reference's own svt_aom_sig_deriv_me and svt_aom_motion_estimation_b64.
 */
#ifndef REF_HARNESS_H
#define REF_HARNESS_H
#include <stdint.h>
#include "svtme_b200.h"

#ifdef __cplusplus
extern "C" {
#endif

/* "encoder level" description of one picture's ME job: what Picture Decision hands to ME */
typedef struct RefMePicIn {
    int32_t  enc_mode;            /* EncMode: -3 (MRS) .. 13 */
    int32_t  width, height;       /* luma size, multiples of 8 */
    int32_t  qp;                  /* scs->static_config.qp */
    int32_t  pred_structure;      /* 1 low-delay-B, 2 random access */
    int32_t  hierarchical_levels; /* pcs->hierarchical_levels */
    int32_t  temporal_layer_index;
    int32_t  is_ref;
    int32_t  sc_class1;
    uint32_t frame_rate;          /* scs->frame_rate (Q16) */
    int32_t  similar_brightness_refs;
    int32_t  gm_level;            /* -1: derive like the reference (svt_aom_derive_gm_level) */
    /* scs->mrp_ctrls fields the path reads */
    int32_t  only_l_bwd, safe_limit_nref, safe_limit_zz_th;
    int32_t  max_l0, max_l1;      /* ref_count_used_list{0,1} (pd_process.c:3495-3511) */
    uint64_t picture_number;
    int32_t  num_lists;           /* 1 = P, 2 = B */
    int32_t  ref_count[2];
    uint64_t ref_picture_number[2][4];
    const uint8_t *src_luma;      /* interior plane, stride = width */
    const uint8_t *ref_luma[2][4];
    int32_t  simd;                /* 0 = C kernels, 1 = AVX2 table, 2 = AVX-512 table */
} RefMePicIn;

/* extra per-b64 state dumped for debugging parity of the HME stages */
typedef struct RefMeDebug {
    int16_t  x_hme_l0[2][4][2][2], y_hme_l0[2][4][2][2];
    int16_t  x_hme_l1[2][4][2][2], y_hme_l1[2][4][2][2];
    int16_t  x_hme_l2[2][4][2][2], y_hme_l2[2][4][2][2];
    uint64_t hme_l0_sad[2][4][2][2], hme_l1_sad[2][4][2][2], hme_l2_sad[2][4][2][2];
    int16_t  prehme_mv_x[2][4][2], prehme_mv_y[2][4][2];
    uint64_t prehme_sad[2][4][2];
    uint32_t zz_sad[2][4];
    uint32_t reduce_me_sr_divisor[2][4];
} RefMeDebug;

/* Fill `out` with the controls the reference derives for this picture (svt_aom_sig_deriv_me)
   and the PCS-level fields (enable_me_8x8, max_cand, gm flags ...). */
int ref_derive_params(const RefMePicIn *in, SvtMePicParams *out);
/* libsvtref_wrap.so only: the same through the product's svtme_shim_flatten (-2 in libsvtref.so) */
int ref_derive_params_shim(const RefMePicIn *in, SvtMePicParams *out);

/* Run the reference ME stage on one picture. `res` arrays follow SvtMePicResults (host memory,
   NULL = skip). `dbg` is [n_b64] or NULL. Returns 0 on success. *seconds (optional) receives the
   wall time spent inside svt_aom_motion_estimation_b64 only. */
int ref_me_picture(const RefMePicIn *in, SvtMePicResults *res, RefMeDebug *dbg, double *seconds);
/* the same with the b64s dealt out to n_threads threads, one MeContext each (the encoder's ME segment threads) */
int ref_me_picture_mt(const RefMePicIn *in, SvtMePicResults *res, int n_threads, double *seconds);
/* n_pics pictures released together (a mini-GOP): res[k] belongs to ins[k]; submit != 0 queues them as one GPU batch first
   when the library was linked with the GPU shim (libsvtref_wrap.so) */
int ref_me_minigop(const RefMePicIn *ins, int n_pics, SvtMePicResults *res, int n_threads, int submit);

/* Which ME table is installed now / was installed when the most recent svt_aom_motion_estimation_b64 call of a
   picture-level run started: 0 = C, 1 = AVX2, 2 = AVX-512 SAD loop + AVX2, -1 = mixed */
int ref_rtcd_table(void);
int ref_rtcd_table_at_last_b64(void);

/* Plane cache (bench.py's CPU baseline only): with it on, ref_me_picture* build the pyramid of a picture once per luma
   pointer, as the encoder does once per picture in Picture Analysis, instead of once per use as a reference. */
void ref_plane_cache_enable(int on);
int  ref_plane_cache_build(const uint8_t *luma, int width, int height);
void ref_plane_cache_drop(const uint8_t *luma); /* NULL = everything */

/* Build the three padded planes of a picture with the reference's own decimation/padding.
   Buffers must hold (w+2*72)*(h+2*72), (w/2+64)*(h/2+64), (w/4+32)*(h/4+32) bytes. */
int ref_build_planes(const uint8_t *luma, int width, int height, uint8_t *full, uint8_t *quarter,
                     uint8_t *sixteenth);

/* RTCD kernels; simd: 0 = C, 1 = AVX2 (SSE4.1 where the reference's x86 table has no AVX2
   entry), 2 = AVX-512 for the SAD loop */
void ref_sad_loop_kernel(int simd, uint8_t *src, uint32_t src_stride, uint8_t *ref,
                         uint32_t ref_stride, uint32_t block_height, uint32_t block_width,
                         uint64_t *best_sad, int16_t *x_search_center, int16_t *y_search_center,
                         uint32_t src_stride_raw, uint8_t skip_search_line,
                         int16_t search_area_width, int16_t search_area_height);
void ref_ext_all_sad_calculation_8x8_16x16(int simd, uint8_t *src, uint32_t src_stride,
                                           uint8_t *ref, uint32_t ref_stride, uint32_t mv,
                                           uint32_t *p_best_sad_8x8, uint32_t *p_best_sad_16x16,
                                           uint32_t *p_best_mv8x8, uint32_t *p_best_mv16x16,
                                           uint32_t p_eight_sad16x16[16][8],
                                           uint32_t p_eight_sad8x8[64][8], uint8_t sub_sad);
void ref_ext_eight_sad_calculation_32x32_64x64(int simd, uint32_t p_sad16x16[16][8],
                                               uint32_t *p_best_sad_32x32,
                                               uint32_t *p_best_sad_64x64,
                                               uint32_t *p_best_mv32x32, uint32_t *p_best_mv64x64,
                                               uint32_t mv, uint32_t p_sad32x32[4][8]);
void ref_ext_sad_calculation_8x8_16x16(int simd, uint8_t *src, uint32_t src_stride, uint8_t *ref,
                                       uint32_t ref_stride, uint32_t *p_best_sad_8x8,
                                       uint32_t *p_best_sad_16x16, uint32_t *p_best_mv8x8,
                                       uint32_t *p_best_mv16x16, uint32_t mv, uint32_t *p_sad16x16,
                                       uint32_t *p_sad8x8, uint8_t sub_sad);
void ref_ext_sad_calculation_32x32_64x64(int simd, uint32_t *p_sad16x16, uint32_t *p_best_sad_32x32,
                                         uint32_t *p_best_sad_64x64, uint32_t *p_best_mv32x32,
                                         uint32_t *p_best_mv64x64, uint32_t mv,
                                         uint32_t *p_sad32x32);
uint32_t ref_nxm_sad_kernel(int simd, const uint8_t *src, uint32_t src_stride, const uint8_t *ref,
                            uint32_t ref_stride, uint32_t height, uint32_t width);
void ref_initialize_buffer_32bits(int simd, uint32_t *pointer, uint32_t count128, uint32_t count32,
                                  uint32_t value);
void ref_downsample_2d(uint8_t *input_samples, uint32_t input_stride, uint32_t input_area_width,
                       uint32_t input_area_height, uint8_t *decim_samples, uint32_t decim_stride,
                       uint32_t decim_step);
void ref_generate_padding(uint8_t *pic, uint32_t stride, uint32_t width, uint32_t height,
                          uint32_t pad_w, uint32_t pad_h);
void ref_unpack_10bit_msb(const uint16_t *in16, uint32_t in_stride, uint8_t *out8,
                          uint32_t out_stride, uint32_t width, uint32_t height);

/* Temporal-filter ME (ME_MCTF, temporal_filtering.c:3095-3146) of the centre picture against ONE neighbouring picture:
   svt_aom_sig_deriv_me_tf + the per-block set-up of create_me_context_and_picture_control + svt_aom_motion_estimation_b64.
   params_out (optional) receives the controls as the product's SvtMePicParams; res: p_sb_best_sad / p_sb_best_mv /
   search_results (entries of b64s that took the HME early exit are zero). */
typedef struct RefTfIn {
    int32_t  width, height, enc_mode, qp;
    int32_t  tf_hme_me_level, tf_qp_opt; /* pcs->tf_ctrls.hme_me_level (0..4), .qp_opt */
    uint32_t tf_me_exit_th;              /* pcs->tf_ctrls.me_exit_th */
    int32_t  temporal_layer_index, is_ref;
    uint64_t centre_picture_number, ref_picture_number;
    const uint8_t *centre_luma, *ref_luma;
    int32_t  simd;
} RefTfIn;
int ref_tf_me_picture(const RefTfIn *in, SvtMePicParams *params_out, SvtMePicResults *res, RefMeDebug *dbg);

/* dynamic-GOP detector metrics of one (source, reference) pair through the reference's dg_detector_hme_level0 */
int ref_dg_detector(const uint8_t *src_luma, const uint8_t *ref_luma, int width, int height, SvtMeDgMetrics *out);

#ifdef __cplusplus
}
#endif
#endif
