/*
 * me_oracle.h — TEST INFRASTRUCTURE ONLY (never linked into, imported by or executed from the
 * product). Plain-C restatement of the reference's open-loop ME path; see me_oracle.c.
 * Parity status: PINNED — checked bit-for-bit against the reference's own sources compiled into
 * oracle/_ref/libsvtref.so (tests/test_oracle_vs_ref.py) and against the committed fixtures under
 * tests/golden/ that were generated from that library (tests/golden/make_golden.py).
 */
#ifndef ME_ORACLE_H
#define ME_ORACLE_H
#include <stdint.h>
#include "svtme_b200.h"

#ifdef __cplusplus
extern "C" {
#endif

/* One padded luma plane: EbPictureBufferDesc fields the path reads (pic_buffer_desc.h:33-74) */
typedef struct OrcPlane {
    const uint8_t *buffer_y;
    int32_t stride_y, org_x, org_y, width, height;
} OrcPlane;

/* full / quarter / sixteenth planes of one picture (EbPaReferenceObject, reference_object.h:73-89) */
typedef struct OrcPic {
    OrcPlane lvl[3];
} OrcPic;

/* extra per-b64 HME state, same layout as RefMeDebug in ref_harness.h */
typedef struct OrcDebug {
    int16_t  x_hme_l0[2][4][2][2], y_hme_l0[2][4][2][2];
    int16_t  x_hme_l1[2][4][2][2], y_hme_l1[2][4][2][2];
    int16_t  x_hme_l2[2][4][2][2], y_hme_l2[2][4][2][2];
    uint64_t hme_l0_sad[2][4][2][2], hme_l1_sad[2][4][2][2], hme_l2_sad[2][4][2][2];
    int16_t  prehme_mv_x[2][4][2], prehme_mv_y[2][4][2];
    uint64_t prehme_sad[2][4][2];
    uint32_t zz_sad[2][4];
    uint32_t reduce_me_sr_divisor[2][4];
} OrcDebug;

/* padded-plane sizes for a WxH picture */
#define ORC_FULL_PAD 72
#define ORC_QUARTER_PAD 32
#define ORC_SIXTEENTH_PAD 16

/* pyramid construction: pad + 2x2 decimation (pic_analysis_process.c:131-161, :2139-2196;
   pic_operators.c:338-384) */
void orc_downsample_2d(const uint8_t *in, uint32_t in_stride, uint32_t in_w, uint32_t in_h, uint8_t *out,
                       uint32_t out_stride, uint32_t step);
void orc_generate_padding(uint8_t *pic, uint32_t stride, uint32_t w, uint32_t h, uint32_t pad_w, uint32_t pad_h);
int  orc_build_planes(const uint8_t *luma, int w, int h, uint8_t *full, uint8_t *quarter, uint8_t *sixteenth);
void orc_unpack_10bit_msb(const uint16_t *in16, uint32_t in_stride, uint8_t *out8, uint32_t out_stride,
                          uint32_t w, uint32_t h);

/* RTCD kernels */
void orc_sad_loop_kernel(const uint8_t *src, uint32_t src_stride, const uint8_t *ref, uint32_t ref_stride,
                         uint32_t block_height, uint32_t block_width, uint64_t *best_sad,
                         int16_t *x_search_center, int16_t *y_search_center, uint32_t src_stride_raw,
                         uint8_t skip_search_line, int16_t search_area_width, int16_t search_area_height);
void orc_ext_all_sad_calculation_8x8_16x16(const uint8_t *src, uint32_t src_stride, const uint8_t *ref,
                                           uint32_t ref_stride, uint32_t mv, uint32_t *p_best_sad_8x8,
                                           uint32_t *p_best_sad_16x16, uint32_t *p_best_mv8x8,
                                           uint32_t *p_best_mv16x16, uint32_t p_eight_sad16x16[16][8],
                                           uint32_t p_eight_sad8x8[64][8], uint8_t sub_sad);
void orc_ext_eight_sad_calculation_32x32_64x64(uint32_t p_sad16x16[16][8], uint32_t *p_best_sad_32x32,
                                               uint32_t *p_best_sad_64x64, uint32_t *p_best_mv32x32,
                                               uint32_t *p_best_mv64x64, uint32_t mv,
                                               uint32_t p_sad32x32[4][8]);
void orc_ext_sad_calculation_8x8_16x16(const uint8_t *src, uint32_t src_stride, const uint8_t *ref,
                                       uint32_t ref_stride, uint32_t *p_best_sad_8x8,
                                       uint32_t *p_best_sad_16x16, uint32_t *p_best_mv8x8,
                                       uint32_t *p_best_mv16x16, uint32_t mv, uint32_t *p_sad16x16,
                                       uint32_t *p_sad8x8, uint8_t sub_sad);
void orc_ext_sad_calculation_32x32_64x64(uint32_t *p_sad16x16, uint32_t *p_best_sad_32x32,
                                         uint32_t *p_best_sad_64x64, uint32_t *p_best_mv32x32,
                                         uint32_t *p_best_mv64x64, uint32_t mv, uint32_t *p_sad32x32);
uint32_t orc_nxm_sad_kernel(const uint8_t *src, uint32_t src_stride, const uint8_t *ref, uint32_t ref_stride,
                            uint32_t height, uint32_t width);
void orc_initialize_buffer_32bits(uint32_t *pointer, uint32_t count128, uint32_t count32, uint32_t value);

/* Whole ME stage of one picture: b64 loop of me_process.c:174-271 around
   svt_aom_motion_estimation_b64 (motion_estimation.c:3146-3223), sequential, one context.
   luma planes are interior (stride = width); the pyramids are built internally.
   Returns 0, or -1 on bad arguments. sad_ops (optional) receives the SURVEY §8(d) op count. */
int orc_me_picture(const uint8_t *src_luma, const uint8_t *const ref_luma[2][4], int width, int height,
                   const SvtMePicParams *params, SvtMePicResults *res, OrcDebug *dbg, uint64_t *sad_ops);

/* dg_detector_hme_level0, Codec/pd_process.c:392-589: metrics of one (source, reference) pair; m->input_resolution is
   an input. b64_sad [n_b64] and b64_mv [n_b64][2] (col, row) are optional. */
int orc_dg_detector(const uint8_t *src_luma, const uint8_t *ref_luma, int width, int height, SvtMeDgMetrics *m,
                    uint64_t *b64_sad, int16_t *b64_mv);

#ifdef __cplusplus
}
#endif
#endif
