/*
 * me_oracle.c — TEST INFRASTRUCTURE ONLY.
 *
 * CPU restatement, in plain C, of the open-loop motion-estimation path of SVT-AV1-PSY 2.3.0-A.
 * It is the checker for the CUDA kernels on machines where /root/reference does not exist (the
 * GPU box). It is never imported, linked or executed by the product (svt-av1-psy_b200/);
 * only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs use it.
 *
 * Parity status: PINNED. Every function below is checked bit-for-bit against the reference's own
 * code compiled into oracle/_ref/libsvtref.so (tests/test_oracle_vs_ref.py, run where the
 * reference is present) and against fixtures generated from that library (tests/golden/).
 *
 * Each function cites the reference lines it follows (paths relative to Source/Lib/). The code is
 * organised differently from the reference (one window-clamp helper instead of five inlined
 * copies, per-point SAD pyramid instead of 8-position groups) but must keep every arithmetic quirk:
 * int16 wrap, the no-op left/top width corrections, strict '<' raster-order arg-min, stale hmeMvSad.
 */
#include "me_oracle.h"

#include <stdlib.h>
#include <string.h>

#define ORC_MIN(a, b) ((a) < (b) ? (a) : (b))
#define ORC_MAX(a, b) ((a) > (b) ? (a) : (b))
#define ORC_ABS(a) ((a) < 0 ? -(a) : (a))
#define MAX_SAD_VALUE (128 * 128 * 255) /* motion_estimation.h:85 */
#define MAX_U32 0xFFFFFFFFu

/* ====================================================================================== */
/* pyramid                                                                                 */
/* ====================================================================================== */

/* svt_aom_downsample_2d_c, Codec/pic_analysis_process.c:131-161: average of the 2x2 samples at
   (h-1,h) x (v-1,v) with h,v starting at step/2 and advancing by step, rounded (+2)>>2 */
void orc_downsample_2d(const uint8_t *in, uint32_t in_stride, uint32_t in_w, uint32_t in_h, uint8_t *out,
                       uint32_t out_stride, uint32_t step) {
    const uint32_t half = step >> 1;
    uint32_t       oy   = 0;
    for (uint32_t v = half; v < in_h; v += step, oy++) {
        const uint8_t *cur  = in + (size_t)v * in_stride;
        const uint8_t *prev = cur - in_stride;
        uint32_t       ox   = 0;
        for (uint32_t h = half; h < in_w; h += step, ox++) {
            uint32_t sum = (uint32_t)prev[h - 1] + prev[h] + cur[h - 1] + cur[h];
            out[(size_t)oy * out_stride + ox] = (uint8_t)((sum + 2) >> 2);
        }
    }
}

/* svt_aom_generate_padding, Codec/pic_operators.c:338-384: replicate left/right per row, then copy
   the first/last padded row up/down pad_h times */
void orc_generate_padding(uint8_t *pic, uint32_t stride, uint32_t w, uint32_t h, uint32_t pad_w, uint32_t pad_h) {
    for (uint32_t y = 0; y < h; y++) {
        uint8_t *row = pic + (size_t)(pad_h + y) * stride + pad_w;
        memset(row - pad_w, row[0], pad_w);
        memset(row + w, row[w - 1], pad_w);
    }
    uint8_t *top = pic + (size_t)pad_h * stride;
    uint8_t *bot = pic + (size_t)(pad_h + h - 1) * stride;
    for (uint32_t i = 1; i <= pad_h; i++) {
        memcpy(top - (size_t)i * stride, top, stride);
        memcpy(bot + (size_t)i * stride, bot, stride);
    }
}

/* svt_aom_pad_input_pictures + svt_aom_downsample_filtering_input_picture,
   Codec/pic_analysis_process.c:2139-2211; geometry Codec/reference_object.c:320-380 and
   Globals/enc_handle.c:4485-4497. The 1/16 picture is decimated from the 1/4 picture. */
int orc_build_planes(const uint8_t *luma, int w, int h, uint8_t *full, uint8_t *quarter, uint8_t *sixteenth) {
    const int fs = w + 2 * ORC_FULL_PAD, qs = (w >> 1) + 2 * ORC_QUARTER_PAD, ss = (w >> 2) + 2 * ORC_SIXTEENTH_PAD;
    for (int y = 0; y < h; y++) memcpy(full + (size_t)(y + ORC_FULL_PAD) * fs + ORC_FULL_PAD, luma + (size_t)y * w, (size_t)w);
    orc_generate_padding(full, (uint32_t)fs, (uint32_t)w, (uint32_t)h, ORC_FULL_PAD, ORC_FULL_PAD);
    orc_downsample_2d(full + (size_t)ORC_FULL_PAD * fs + ORC_FULL_PAD, (uint32_t)fs, (uint32_t)w, (uint32_t)h,
                      quarter + (size_t)ORC_QUARTER_PAD * qs + ORC_QUARTER_PAD, (uint32_t)qs, 2);
    orc_generate_padding(quarter, (uint32_t)qs, (uint32_t)(w >> 1), (uint32_t)(h >> 1), ORC_QUARTER_PAD, ORC_QUARTER_PAD);
    orc_downsample_2d(quarter + (size_t)ORC_QUARTER_PAD * qs + ORC_QUARTER_PAD, (uint32_t)qs, (uint32_t)(w >> 1),
                      (uint32_t)(h >> 1), sixteenth + (size_t)ORC_SIXTEENTH_PAD * ss + ORC_SIXTEENTH_PAD, (uint32_t)ss, 2);
    orc_generate_padding(sixteenth, (uint32_t)ss, (uint32_t)(w >> 2), (uint32_t)(h >> 2), ORC_SIXTEENTH_PAD, ORC_SIXTEENTH_PAD);
    return 0;
}

/* 8 MSBs of a 10-bit sample: svt_unpack_and_2bcompress_c, C_DEFAULT/pack_unpack_c.c:77-116
   (ME never reads the 2-bit plane: Codec/reference_object.c:327) */
void orc_unpack_10bit_msb(const uint16_t *in16, uint32_t in_stride, uint8_t *out8, uint32_t out_stride,
                          uint32_t w, uint32_t h) {
    for (uint32_t y = 0; y < h; y++)
        for (uint32_t x = 0; x < w; x++) out8[(size_t)y * out_stride + x] = (uint8_t)(in16[(size_t)y * in_stride + x] >> 2);
}

/* ====================================================================================== */
/* RTCD kernels                                                                            */
/* ====================================================================================== */

static inline uint32_t absdiff(uint8_t a, uint8_t b) { return a > b ? (uint32_t)(a - b) : (uint32_t)(b - a); }

/* svt_nxm_sad_kernel_helper_c / svt_fast_loop_nxm_sad_kernel, C_DEFAULT/compute_sad_c.c:20-37,209-212 */
uint32_t orc_nxm_sad_kernel(const uint8_t *src, uint32_t src_stride, const uint8_t *ref, uint32_t ref_stride,
                            uint32_t height, uint32_t width) {
    uint32_t sad = 0;
    for (uint32_t y = 0; y < height; y++, src += src_stride, ref += ref_stride)
        for (uint32_t x = 0; x < width; x++) sad += absdiff(src[x], ref[x]);
    return sad;
}

/* svt_sad_loop_kernel_c, C_DEFAULT/compute_sad_c.c:58-101. Raster scan, strict '<', best starts at
   0xffffff, outputs untouched when nothing beats it; even search rows skipped when
   skip_search_line && width == 16 && height <= 16; the ref row advance uses src_stride_raw. */
void orc_sad_loop_kernel(const uint8_t *src, uint32_t src_stride, const uint8_t *ref, uint32_t ref_stride,
                         uint32_t block_height, uint32_t block_width, uint64_t *best_sad,
                         int16_t *x_search_center, int16_t *y_search_center, uint32_t src_stride_raw,
                         uint8_t skip_search_line, int16_t search_area_width, int16_t search_area_height) {
    const int skip = skip_search_line && block_width == 16 && block_height <= 16;
    *best_sad      = 0xffffff;
    for (int16_t ys = 0; ys < search_area_height; ys++, ref += src_stride_raw) {
        if (skip && (ys & 1) == 0) continue;
        for (int16_t xs = 0; xs < search_area_width; xs++) {
            uint32_t sad = orc_nxm_sad_kernel(src, src_stride, ref + xs, ref_stride, block_height, block_width);
            if (sad < *best_sad) {
                *best_sad        = sad;
                *x_search_center = xs;
                *y_search_center = ys;
            }
        }
    }
}

/* svt_initialize_buffer_32bits_c, Codec/me_sad_calculation.c:14-17 */
void orc_initialize_buffer_32bits(uint32_t *pointer, uint32_t count128, uint32_t count32, uint32_t value) {
    for (uint32_t i = 0; i < count128 * 4 + count32; i++) pointer[i] = value;
}

/* 8x8 SAD; sub_sad: rows 0,2,4,6 only, doubled (Codec/motion_estimation.c:42-63,105-135,224-231) */
static uint32_t sad8x8(const uint8_t *src, uint32_t src_stride, const uint8_t *ref, uint32_t ref_stride, int sub_sad) {
    if (sub_sad) return orc_nxm_sad_kernel(src, 2 * src_stride, ref, 2 * ref_stride, 4, 8) << 1;
    return orc_nxm_sad_kernel(src, src_stride, ref, ref_stride, 8, 8);
}

static inline uint32_t pack_mv(int16_t x, int16_t y) { return ((uint32_t)(uint16_t)y << 16) | (uint16_t)x; }

/* svt_ext_sad_calculation_8x8_16x16_c, Codec/motion_estimation.c:98-164: one 16x16 block, one
   search position; p_sad8x8[4] / p_sad16x16[1] written; bests updated with strict '<' */
void orc_ext_sad_calculation_8x8_16x16(const uint8_t *src, uint32_t src_stride, const uint8_t *ref,
                                       uint32_t ref_stride, uint32_t *p_best_sad_8x8,
                                       uint32_t *p_best_sad_16x16, uint32_t *p_best_mv8x8,
                                       uint32_t *p_best_mv16x16, uint32_t mv, uint32_t *p_sad16x16,
                                       uint32_t *p_sad8x8, uint8_t sub_sad) {
    uint32_t sum = 0;
    for (int k = 0; k < 4; k++) {
        const uint32_t off_s = (uint32_t)(k >> 1) * 8 * src_stride + (uint32_t)(k & 1) * 8;
        const uint32_t off_r = (uint32_t)(k >> 1) * 8 * ref_stride + (uint32_t)(k & 1) * 8;
        p_sad8x8[k]          = sad8x8(src + off_s, src_stride, ref + off_r, ref_stride, sub_sad);
        if (p_sad8x8[k] < p_best_sad_8x8[k]) {
            p_best_sad_8x8[k] = p_sad8x8[k];
            p_best_mv8x8[k]   = mv;
        }
        sum += p_sad8x8[k];
    }
    if (sum < p_best_sad_16x16[0]) {
        p_best_sad_16x16[0] = sum;
        p_best_mv16x16[0]   = mv;
    }
    *p_sad16x16 = sum;
}

/* svt_ext_sad_calculation_32x32_64x64_c, Codec/motion_estimation.c:171-205 */
void orc_ext_sad_calculation_32x32_64x64(uint32_t *p_sad16x16, uint32_t *p_best_sad_32x32,
                                         uint32_t *p_best_sad_64x64, uint32_t *p_best_mv32x32,
                                         uint32_t *p_best_mv64x64, uint32_t mv, uint32_t *p_sad32x32) {
    uint32_t total = 0;
    for (int k = 0; k < 4; k++) {
        uint32_t s = p_sad16x16[4 * k] + p_sad16x16[4 * k + 1] + p_sad16x16[4 * k + 2] + p_sad16x16[4 * k + 3];
        p_sad32x32[k] = s;
        if (s < p_best_sad_32x32[k]) {
            p_best_sad_32x32[k] = s;
            p_best_mv32x32[k]   = mv;
        }
        total += s;
    }
    if (total < p_best_sad_64x64[0]) {
        p_best_sad_64x64[0] = total;
        p_best_mv64x64[0]   = mv;
    }
}

/* z-order index of the 16x16 block at raster (y,x): offsets[] of Codec/motion_estimation.c:341 */
static const uint8_t k_z16[16] = {0, 1, 4, 5, 2, 3, 6, 7, 8, 9, 12, 13, 10, 11, 14, 15};

/* svt_ext_all_sad_calculation_8x8_16x16_c, Codec/motion_estimation.c:210-362: 8 consecutive
   horizontal positions; for each 16x16 block (raster), positions 0..7 in order; the MV of
   position i is (x + i, y) with the x add done in int16 (:233-235). p_eight_sad8x8 is unused. */
void orc_ext_all_sad_calculation_8x8_16x16(const uint8_t *src, uint32_t src_stride, const uint8_t *ref,
                                           uint32_t ref_stride, uint32_t mv, uint32_t *p_best_sad_8x8,
                                           uint32_t *p_best_sad_16x16, uint32_t *p_best_mv8x8,
                                           uint32_t *p_best_mv16x16, uint32_t p_eight_sad16x16[16][8],
                                           uint32_t p_eight_sad8x8[64][8], uint8_t sub_sad) {
    (void)p_eight_sad8x8;
    const int16_t mvx = (int16_t)(mv & 0xFFFF), mvy = (int16_t)(mv >> 16);
    for (int by = 0; by < 4; by++)
        for (int bx = 0; bx < 4; bx++) {
            const int      z16 = k_z16[4 * by + bx];
            const uint8_t *s   = src + (size_t)16 * by * src_stride + 16 * bx;
            const uint8_t *r   = ref + (size_t)16 * by * ref_stride + 16 * bx;
            for (int i = 0; i < 8; i++) {
                const uint32_t mvi = pack_mv((int16_t)(mvx + (int16_t)i), mvy);
                uint32_t       sum = 0;
                for (int k = 0; k < 4; k++) {
                    const uint32_t off_s = (uint32_t)(k >> 1) * 8 * src_stride + (uint32_t)(k & 1) * 8;
                    const uint32_t off_r = (uint32_t)(k >> 1) * 8 * ref_stride + (uint32_t)(k & 1) * 8;
                    const uint32_t v     = sad8x8(s + off_s, src_stride, r + off_r + i, ref_stride, sub_sad);
                    if (v < p_best_sad_8x8[4 * z16 + k]) {
                        p_best_sad_8x8[4 * z16 + k] = v;
                        p_best_mv8x8[4 * z16 + k]   = mvi;
                    }
                    sum += v;
                }
                p_eight_sad16x16[z16][i] = sum;
                if (sum < p_best_sad_16x16[z16]) {
                    p_best_sad_16x16[z16] = sum;
                    p_best_mv16x16[z16]   = mvi;
                }
            }
        }
}

/* svt_ext_eight_sad_calculation_32x32_64x64_c, Codec/motion_estimation.c:369-425 */
void orc_ext_eight_sad_calculation_32x32_64x64(uint32_t p_sad16x16[16][8], uint32_t *p_best_sad_32x32,
                                               uint32_t *p_best_sad_64x64, uint32_t *p_best_mv32x32,
                                               uint32_t *p_best_mv64x64, uint32_t mv,
                                               uint32_t p_sad32x32[4][8]) {
    const int16_t mvx = (int16_t)(mv & 0xFFFF), mvy = (int16_t)(mv >> 16);
    for (int i = 0; i < 8; i++) {
        const uint32_t mvi   = pack_mv((int16_t)(mvx + (int16_t)i), mvy);
        uint32_t       total = 0;
        for (int k = 0; k < 4; k++) {
            uint32_t s = p_sad16x16[4 * k][i] + p_sad16x16[4 * k + 1][i] + p_sad16x16[4 * k + 2][i] + p_sad16x16[4 * k + 3][i];
            p_sad32x32[k][i] = s;
            if (s < p_best_sad_32x32[k]) {
                p_best_sad_32x32[k] = s;
                p_best_mv32x32[k]   = mvi;
            }
            total += s;
        }
        if (total < p_best_sad_64x64[0]) {
            p_best_sad_64x64[0] = total;
            p_best_mv64x64[0]   = mvi;
        }
    }
}

/* ====================================================================================== */
/* per-b64 context (the MeContext fields the path uses, Codec/me_context.h:366-509)        */
/* ====================================================================================== */

typedef struct SearchInfo { /* me_context.h:329-334 */
    uint16_t sa_w, sa_h;
    int16_t  mv_x, mv_y; /* best_mv.as_mv.col / .row */
    uint64_t sad;
    uint8_t  valid;
} SearchInfo;

typedef struct SearchRes { /* me_context.h:348-355 */
    int16_t  hme_sc_x, hme_sc_y;
    uint64_t hme_sad;
    uint8_t  do_ref;
} SearchRes;

typedef struct Ctx {
    const SvtMePicParams *pp;
    const SvtMeCtrls     *c;
    int                   pic_w, pic_h;
    OrcPic                src;
    OrcPic                ref[2][4];
    /* current b64 */
    uint32_t       b64_index, org_x, org_y, b64_w, b64_h, b64_geom_w, b64_geom_h;
    const uint8_t *b64_src, *quarter_src, *sixteenth_src;
    uint32_t       b64_src_stride, quarter_stride, sixteenth_stride;
    /* state that persists from one b64 to the next exactly as in MeContext */
    SvtMeSearchAreaMinMax hme_l0_sa;
    int16_t    xc[3][2][4][2][2], yc[3][2][4][2][2]; /* [level][list][ref][sr_w][sr_h] */
    uint64_t   hsad[3][2][4][2][2];
    SearchInfo prehme[2][4][2];
    uint8_t    performed_phme[2][4][2];
    SearchRes  sr[2][4];
    uint32_t   reduce_me_sr_divisor[2][4];
    uint32_t   zz_sad[2][4];
    uint32_t   best_sad[2][4][85], best_mv[2][4][85];
    uint32_t   me_distortion[85];
    SvtMeWindow win[2][4];
    uint64_t   sad_ops;
} Ctx;

/* get_me_reference distance, Codec/motion_estimation.c:1232-1234 */
static uint16_t ref_dist(const Ctx *x, int l, int r) {
    int64_t d = (int64_t)x->pp->picture_number - (int64_t)x->pp->ref_picture_number[l][r];
    return (uint16_t)(int16_t)ORC_ABS(d);
}

/* svt_aom_get_scaled_picture_distance, Codec/motion_estimation.c:1239-1243 */
static uint16_t scaled_dist(uint16_t dist) {
    uint8_t round_up = ((dist % 8) == 0) ? 0 : 1;
    return (uint16_t)(((dist * 5) / 8) + round_up);
}

/* The edge correction every search stage applies to one axis (Codec/motion_estimation.c:848-880,
   949-982, 1050-1078, 1481-1500, 1542-1560, 1655-1699). All int16. The left/top branch moves the
   origin first, so the size correction that follows in the reference subtracts zero: kept as is. */
static void clamp_axis(int16_t org, int16_t *origin, int16_t *size, int16_t pad, int16_t dim) {
    if ((org + *origin) < -pad) {
        *origin = (int16_t)(-pad - org);
        *size   = (int16_t)(*size - (-pad - (org + *origin)));
    }
    if ((org + *origin) > dim - 1) *origin = (int16_t)(*origin - ((org + *origin) - (dim - 1)));
    if ((org + *origin + *size) > dim) *size = (int16_t)ORC_MAX(1, *size - ((org + *origin + *size) - dim));
}

/* Body shared by prehme_core (:1638-1736) and hme_level_0/1/2 (:820-1113): clamp the window, run
   svt_sad_loop_kernel on the plane, scale the result back to full resolution. */
static void hme_window_search(Ctx *x, const OrcPlane *ref, const uint8_t *src, uint32_t src_stride, int16_t org_x,
                              int16_t org_y, uint32_t bw, uint32_t bh, int16_t origin_x, int16_t origin_y,
                              int16_t sa_w, int16_t sa_h, int16_t pad_w, int16_t pad_h, int round_w8, uint8_t skip,
                              int scale, uint64_t *sad, int16_t *cx, int16_t *cy) {
    const int full = x->c->hme_search_method == SVTME_FULL_SAD_SEARCH;
    clamp_axis(org_x, &origin_x, &sa_w, pad_w, (int16_t)ref->width);
    if (round_w8) sa_w = (int16_t)((sa_w < 8) ? sa_w : sa_w & ~0x07);
    clamp_axis(org_y, &origin_y, &sa_h, pad_h, (int16_t)ref->height);
    int16_t  x_tl  = (int16_t)(((int16_t)ref->org_x + org_x) + origin_x);
    int16_t  y_tl  = (int16_t)(((int16_t)ref->org_y + org_y) + origin_y);
    uint32_t index = (uint32_t)(x_tl + y_tl * ref->stride_y);
    orc_sad_loop_kernel(src, full ? src_stride : src_stride * 2, ref->buffer_y + index,
                        full ? (uint32_t)ref->stride_y : (uint32_t)ref->stride_y * 2, full ? bh : bh >> 1, bw, sad, cx,
                        cy, (uint32_t)ref->stride_y, skip, sa_w, sa_h);
    if (sa_w > 0 && sa_h > 0) x->sad_ops += (uint64_t)sa_w * (uint64_t)(skip && bw == 16 && (full ? bh : bh >> 1) <= 16 ? sa_h / 2 : sa_h) * bw * (full ? bh : bh >> 1);
    if (!full) *sad = *sad * 2;
    *cx = (int16_t)(*cx + origin_x);
    *cx = (int16_t)(*cx * scale);
    *cy = (int16_t)(*cy + origin_y);
    *cy = (int16_t)(*cy * scale);
}

/* get_zz_sad, Codec/motion_estimation.c:1737-1759 */
static uint32_t get_zz_sad(Ctx *x, const OrcPlane *ref) {
    uint32_t index = (uint32_t)((int16_t)ref->org_x + (int16_t)x->org_x + ((int16_t)ref->org_y + (int16_t)x->org_y) * ref->stride_y);
    uint32_t s = orc_nxm_sad_kernel(x->b64_src, x->b64_src_stride << 1, ref->buffer_y + index, (uint32_t)ref->stride_y << 1,
                                    x->b64_h >> 1, x->b64_w);
    x->sad_ops += (uint64_t)(x->b64_h >> 1) * x->b64_w;
    return s << 1;
}

/* init_zz_sad, Codec/motion_estimation.c:2452-2507 */
static void init_zz_sad(Ctx *x) {
    const SvtMePicParams *pp = x->pp;
    uint32_t best_zz_sad = MAX_U32;
    for (int l = 0; l < pp->num_of_list_to_search; l++)
        for (int r = 0; r < pp->num_of_ref_pic_to_search[l]; r++)
            if (pp->temporal_layer_index > 0 || l == 0) {
                uint32_t zz = get_zz_sad(x, &x->ref[l][r].lvl[0]);
                zz          = (zz * 64 * 64) / (x->b64_w * x->b64_h);
                x->zz_sad[l][r] = zz;
                best_zz_sad     = ORC_MIN(best_zz_sad, zz);
            }
    const uint32_t zz_th = x->c->zz_sad_th;
    if (pp->temporal_layer_index > 0 && best_zz_sad < zz_th)
        for (int l = 0; l < pp->num_of_list_to_search; l++)
            for (int r = 1; r < pp->num_of_ref_pic_to_search[l]; r++) {
                const uint32_t pct = x->c->zz_sad_pct;
                if ((x->zz_sad[l][r] - best_zz_sad) * 100 > (pct * best_zz_sad)) x->sr[l][r].do_ref = 0;
            }
    const uint32_t safe_limit_zz_th = x->c->me_safe_limit_zz_th;
    if (safe_limit_zz_th) {
        int limit = pp->hierarchical_levels > 0 && pp->num_of_list_to_search == 2 &&
            pp->temporal_layer_index >= pp->hierarchical_levels && pp->similar_brightness_refs &&
            x->zz_sad[0][0] < safe_limit_zz_th && x->zz_sad[1][0] < safe_limit_zz_th;
        for (int l = 0; l < pp->num_of_list_to_search; l++)
            for (int r = 0; r < pp->num_of_ref_pic_to_search[l]; r++)
                if (limit && r > 0) x->sr[l][r].do_ref = 0;
    }
}

/* check_prehme_early_exit, Codec/motion_estimation.c:1763-1789 */
static int prehme_early_exit(Ctx *x, int l, int r, int s) {
    SearchInfo *d = &x->prehme[l][r][s];
    if (x->c->me_early_exit_th && x->zz_sad[l][r] < x->c->me_early_exit_th) {
        d->mv_x = d->mv_y = 0;
        d->sad   = 0;
        d->valid = 1;
        return 1;
    }
    if (x->c->prehme_l1_early_exit) {
        const SearchInfo *p = &x->prehme[0][r][s];
        if (l == 1 && p->valid && ((p->sad < (32 * 32)) || ((ORC_ABS(p->mv_x) < 16) && (ORC_ABS(p->mv_y) < 16)))) {
            d->mv_x  = (int16_t)-p->mv_x;
            d->mv_y  = (int16_t)-p->mv_y;
            d->sad   = p->sad;
            d->valid = 1;
            return 1;
        }
    }
    return 0;
}

/* prehme_b64 + prehme_core, Codec/motion_estimation.c:1638-1736, 1792-1866 */
static void prehme_b64(Ctx *x) {
    const SvtMePicParams *pp = x->pp;
    const SvtMeCtrls     *c  = x->c;
    uint32_t best_sad = MAX_U32;
    for (int l = 0; l < pp->num_of_list_to_search; l++)
        for (int r = 0; r < pp->num_of_ref_pic_to_search[l]; r++) {
            const OrcPlane *ref = &x->ref[l][r].lvl[2];
            if (pp->temporal_layer_index > 0 || l == 0) {
                uint32_t factor = scaled_dist(ref_dist(x, l, r));
                for (int s = 0; s < 2; s++) {
                    if (prehme_early_exit(x, l, r, s)) continue;
                    SearchInfo *d = &x->prehme[l][r][s];
                    if (!x->sr[l][r].do_ref) {
                        d->mv_x = d->mv_y = 0;
                        d->sad = MAX_U32;
                        continue;
                    }
                    d->sa_w = (uint16_t)ORC_MIN(c->prehme_sa_cfg[s].sa_min.width * factor, (uint32_t)c->prehme_sa_cfg[s].sa_max.width);
                    d->sa_h = (uint16_t)ORC_MIN(c->prehme_sa_cfg[s].sa_min.height * factor, (uint32_t)c->prehme_sa_cfg[s].sa_max.height);
                    int16_t sa_w = (int16_t)d->sa_w, sa_h = (int16_t)d->sa_h;
                    hme_window_search(x, ref, x->sixteenth_src, x->sixteenth_stride, (int16_t)(((int16_t)x->org_x) >> 2),
                                      (int16_t)(((int16_t)x->org_y) >> 2), x->b64_w >> 2, x->b64_h >> 2,
                                      (int16_t)(-(int16_t)(sa_w >> 1)), (int16_t)(-(int16_t)(sa_h >> 1)), sa_w, sa_h,
                                      (int16_t)(ref->org_x - 1), (int16_t)(ref->org_y - 1), 0, c->prehme_skip_search_line, 4,
                                      &d->sad, &d->mv_x, &d->mv_y);
                    d->valid                    = 1;
                    x->performed_phme[l][r][s] = 1;
                }
                uint32_t min_sad = (uint32_t)ORC_MIN(x->prehme[l][r][0].sad, x->prehme[l][r][1].sad);
                best_sad         = ORC_MIN(best_sad, min_sad);
            } else {
                for (int s = 0; s < 2; s++) {
                    x->prehme[1][r][s].mv_x = (int16_t)-x->prehme[0][r][s].mv_x;
                    x->prehme[1][r][s].mv_y = (int16_t)-x->prehme[0][r][s].mv_y;
                    x->prehme[1][r][s].sad  = x->prehme[0][r][s].sad;
                }
            }
        }
    if (pp->temporal_layer_index > 0 && best_sad < c->phme_sad_th)
        for (int l = 0; l < pp->num_of_list_to_search; l++)
            for (int r = 1; r < pp->num_of_ref_pic_to_search[l]; r++) {
                if (!x->sr[l][r].do_ref) continue;
                const uint32_t th  = c->phme_sad_pct;
                uint32_t       sad = (uint32_t)ORC_MIN(x->prehme[l][r][0].sad, x->prehme[l][r][1].sad);
                if ((sad - best_sad) * 100 > (th * best_sad)) x->sr[l][r].do_ref = 0;
            }
}

/* get_hme_l0_search_area, Codec/motion_estimation.c:1870-1937 (mutates x->hme_l0_sa; the caller
   restores it, :2069-2073) */
static void hme_l0_search_area(Ctx *x, int l, int r, uint16_t dist, int16_t *sa_w, int16_t *sa_h) {
    const SvtMeCtrls *c = x->c;
    if (c->enable_me_sr_adjustment && c->distance_based_hme_resizing) {
        uint8_t is_hor = 1, is_ver = 1, is_still = 0;
        if (c->reduce_hme_l0_sr_th_min && c->reduce_hme_l0_sr_th_max && (l || r)) {
            int16_t mvx = x->xc[0][0][0][0][0], mvy = x->yc[0][0][0][0][0];
            is_ver   = (ORC_ABS(mvx) < c->reduce_hme_l0_sr_th_min) && (ORC_ABS(mvy) > c->reduce_hme_l0_sr_th_max);
            is_hor   = (ORC_ABS(mvx) > c->reduce_hme_l0_sr_th_max) && (ORC_ABS(mvy) < c->reduce_hme_l0_sr_th_min);
            is_still = (ORC_ABS(mvx) < (c->reduce_hme_l0_sr_th_min * 3)) && (ORC_ABS(mvy) < (c->reduce_hme_l0_sr_th_min * 3));
        }
        uint8_t x_offset = is_hor ? 1 : 2, y_offset = is_ver ? 1 : 2;
        if (c->enable_me_sr_adjustment == 2 && is_still) x_offset = y_offset = 4;
        x->hme_l0_sa.sa_min.width  = (uint16_t)(x->hme_l0_sa.sa_min.width / (x_offset + r));
        x->hme_l0_sa.sa_min.height = (uint16_t)(x->hme_l0_sa.sa_min.height / (y_offset + r));
        x->hme_l0_sa.sa_max.width  = (uint16_t)(x->hme_l0_sa.sa_max.width / (x_offset + r));
        x->hme_l0_sa.sa_max.height = (uint16_t)(x->hme_l0_sa.sa_max.height / (y_offset + r));
    }
    int32_t factor = scaled_dist(dist);
    int16_t w      = (int16_t)(x->hme_l0_sa.sa_min.width / c->num_hme_sa_w);
    w = (int16_t)ORC_MIN((((w * factor) + 15) & ~0x0F), (((x->hme_l0_sa.sa_max.width / c->num_hme_sa_w) + 15) & ~0x0F));
    int16_t h = (int16_t)(x->hme_l0_sa.sa_min.height / c->num_hme_sa_h);
    h         = (int16_t)ORC_MIN((h * factor), x->hme_l0_sa.sa_max.height / c->num_hme_sa_h);
    *sa_w = w;
    *sa_h = h;
}

static void set_quadrants(Ctx *x, int lvl, int l, int r, int16_t cx, int16_t cy, uint64_t sad) {
    for (int sh = 0; sh < x->c->num_hme_sa_h; sh++)
        for (int sw = 0; sw < x->c->num_hme_sa_w; sw++) {
            x->xc[lvl][l][r][sw][sh]   = cx;
            x->yc[lvl][l][r][sw][sh]   = cy;
            x->hsad[lvl][l][r][sw][sh] = sad;
        }
}

/* hme_level0_b64 + hme_level_0 + get_worst_quadrant, Codec/motion_estimation.c:820-920, 1942-2106 */
static void hme_level0_b64(Ctx *x) {
    const SvtMePicParams *pp   = x->pp;
    const SvtMeCtrls     *c    = x->c;
    const SvtMeSearchAreaMinMax base = x->hme_l0_sa;
    for (int l = 0; l < pp->num_of_list_to_search; l++)
        for (int r = 0; r < pp->num_of_ref_pic_to_search[l]; r++) {
            if (c->me_early_exit_th && x->zz_sad[l][r] < (c->me_early_exit_th >> 2)) {
                set_quadrants(x, 0, l, r, 0, 0, 0);
                continue;
            }
            if (c->prev_me_stage_based_exit_th) {
                int s = x->prehme[l][r][0].sad <= x->prehme[l][r][1].sad ? 0 : 1;
                if (x->performed_phme[l][r][s] && x->prehme[l][r][s].sad < (c->prev_me_stage_based_exit_th >> 4)) {
                    set_quadrants(x, 0, l, r, x->prehme[l][r][s].mv_x, x->prehme[l][r][s].mv_y, x->prehme[l][r][s].sad);
                    continue;
                }
            }
            if (!x->sr[l][r].do_ref) {
                set_quadrants(x, 0, l, r, 0, 0, MAX_U32);
                continue;
            }
            if (!(pp->temporal_layer_index > 0 || l == 0)) continue;
            const OrcPlane *ref = &x->ref[l][r].lvl[2];
            int16_t sa_w = 0, sa_h = 0;
            hme_l0_search_area(x, l, r, ref_dist(x, l, r), &sa_w, &sa_h);
            for (int sh = 0; sh < c->num_hme_sa_h; sh++)
                for (int sw = 0; sw < c->num_hme_sa_w; sw++) {
                    /* hme_level_0 :837-846: width rounded up to 8 first, origin of quadrant (sw,sh) */
                    int16_t w  = (int16_t)((sa_w + 7) & ~0x07);
                    int16_t ox = (int16_t)(-(int16_t)((w * c->num_hme_sa_w) >> 1) + (int16_t)(w * sw));
                    int16_t oy = (int16_t)(-(int16_t)((sa_h * c->num_hme_sa_h) >> 1) + (int16_t)(sa_h * sh));
                    hme_window_search(x, ref, x->sixteenth_src, x->sixteenth_stride, (int16_t)(((int16_t)x->org_x) >> 2),
                                      (int16_t)(((int16_t)x->org_y) >> 2), x->b64_w >> 2, x->b64_h >> 2, ox, oy, w, sa_h,
                                      (int16_t)(ref->org_x - 1), (int16_t)(ref->org_y - 1), 1, 0, 4,
                                      &x->hsad[0][l][r][sw][sh], &x->xc[0][l][r][sw][sh], &x->yc[0][l][r][sw][sh]);
                }
            if (c->enable_me_sr_adjustment && c->distance_based_hme_resizing) x->hme_l0_sa = base;
            if (c->prehme_enable) {
                /* get_worst_quadrant :1942-1971: strict '>' from 0, order (0,0),(1,0),(0,1),(1,1);
                   the last test does not update max_sad */
                uint8_t  bw = 0, bh = 0;
                uint64_t max_sad = 0;
                if (x->hsad[0][l][r][0][0] > max_sad) { max_sad = x->hsad[0][l][r][0][0]; bw = 0; bh = 0; }
                if (x->hsad[0][l][r][1][0] > max_sad) { max_sad = x->hsad[0][l][r][1][0]; bw = 1; bh = 0; }
                if (x->hsad[0][l][r][0][1] > max_sad) { max_sad = x->hsad[0][l][r][0][1]; bw = 0; bh = 1; }
                if (x->hsad[0][l][r][1][1] > max_sad) { bw = 1; bh = 1; }
                int s = x->prehme[l][r][0].sad <= x->prehme[l][r][1].sad ? 0 : 1;
                if (x->prehme[l][r][s].sad < x->hsad[0][l][r][bw][bh]) {
                    x->hsad[0][l][r][bw][bh] = x->prehme[l][r][s].sad;
                    x->xc[0][l][r][bw][bh]   = x->prehme[l][r][s].mv_x;
                    x->yc[0][l][r][bw][bh]   = x->prehme[l][r][s].mv_y;
                }
            }
        }
}

/* hme_level1_b64 + hme_level_1, Codec/motion_estimation.c:923-1022, 2111-2192 */
static void hme_level1_b64(Ctx *x) {
    const SvtMePicParams *pp = x->pp;
    const SvtMeCtrls     *c  = x->c;
    for (int l = 0; l < pp->num_of_list_to_search; l++)
        for (int r = 0; r < pp->num_of_ref_pic_to_search[l]; r++) {
            if (!(pp->temporal_layer_index > 0 || l == 0)) continue;
            if (c->me_early_exit_th && x->zz_sad[l][r] < (c->me_early_exit_th >> 2)) {
                set_quadrants(x, 1, l, r, 0, 0, 0);
                continue;
            }
            if (!x->sr[l][r].do_ref) {
                set_quadrants(x, 1, l, r, 0, 0, MAX_U32);
                continue;
            }
            const OrcPlane *ref = &x->ref[l][r].lvl[1];
            for (int sh = 0; sh < c->num_hme_sa_h; sh++)
                for (int sw = 0; sw < c->num_hme_sa_w; sw++) {
                    if (c->prev_me_stage_based_exit_th && x->hsad[0][l][r][sw][sh] < (c->prev_me_stage_based_exit_th >> 5)) {
                        x->xc[1][l][r][sw][sh]   = x->xc[0][l][r][sw][sh];
                        x->yc[1][l][r][sw][sh]   = x->yc[0][l][r][sw][sh];
                        x->hsad[1][l][r][sw][sh] = x->hsad[0][l][r][sw][sh];
                        continue;
                    }
                    int16_t w  = (int16_t)(((int16_t)c->hme_l1_sa.width + 7) & ~0x07);
                    int16_t h  = (int16_t)c->hme_l1_sa.height;
                    int16_t ox = (int16_t)(-(w >> 1) + (int16_t)(x->xc[0][l][r][sw][sh] >> 1));
                    int16_t oy = (int16_t)(-(h >> 1) + (int16_t)(x->yc[0][l][r][sw][sh] >> 1));
                    hme_window_search(x, ref, x->quarter_src, x->quarter_stride, (int16_t)(((int16_t)x->org_x) >> 1),
                                      (int16_t)(((int16_t)x->org_y) >> 1), x->b64_w >> 1, x->b64_h >> 1, ox, oy, w, h,
                                      (int16_t)(ref->org_x - 1), (int16_t)(ref->org_y - 1), 1, 0, 2,
                                      &x->hsad[1][l][r][sw][sh], &x->xc[1][l][r][sw][sh], &x->yc[1][l][r][sw][sh]);
                }
        }
}

/* hme_level2_b64 + hme_level_2, Codec/motion_estimation.c:1025-1113, 2197-2247 (no do_ref test) */
static void hme_level2_b64(Ctx *x) {
    const SvtMePicParams *pp = x->pp;
    const SvtMeCtrls     *c  = x->c;
    for (int l = 0; l < pp->num_of_list_to_search; l++)
        for (int r = 0; r < pp->num_of_ref_pic_to_search[l]; r++) {
            if (!(pp->temporal_layer_index > 0 || l == 0)) continue;
            const OrcPlane *ref = &x->ref[l][r].lvl[0];
            for (int sh = 0; sh < c->num_hme_sa_h; sh++)
                for (int sw = 0; sw < c->num_hme_sa_w; sw++) {
                    if (c->prev_me_stage_based_exit_th && x->hsad[1][l][r][sw][sh] < (c->prev_me_stage_based_exit_th >> 2)) {
                        x->xc[2][l][r][sw][sh]   = x->xc[1][l][r][sw][sh];
                        x->yc[2][l][r][sw][sh]   = x->yc[1][l][r][sw][sh];
                        x->hsad[2][l][r][sw][sh] = x->hsad[1][l][r][sw][sh];
                        continue;
                    }
                    int16_t w  = (int16_t)(((int16_t)c->hme_l2_sa.width + 7) & ~0x07);
                    int16_t h  = (int16_t)c->hme_l2_sa.height;
                    int16_t ox = (int16_t)(-(w >> 1) + x->xc[1][l][r][sw][sh]);
                    int16_t oy = (int16_t)(-(h >> 1) + x->yc[1][l][r][sw][sh]);
                    hme_window_search(x, ref, x->b64_src, x->b64_src_stride, (int16_t)x->org_x, (int16_t)x->org_y, x->b64_w,
                                      x->b64_h, ox, oy, w, h, 63, 63, 1, 0, 1, &x->hsad[2][l][r][sw][sh],
                                      &x->xc[2][l][r][sw][sh], &x->yc[2][l][r][sw][sh]);
                }
        }
}

/* set_final_seach_centre_sb, Codec/motion_estimation.c:2252-2450. Best of the quadrants of the
   deepest enabled level (first wins ties; scan sw fastest starting at (1,0)). hme_sad is carried
   across refs when a ref skips HME (list 1 of a base-layer picture): hmeMvSad lives outside the
   loops (:2263). */
static void set_final_search_centre(Ctx *x) {
    const SvtMePicParams *pp = x->pp;
    const SvtMeCtrls     *c  = x->c;
    int16_t  hx = 0, hy = 0, scx = 0, scy = 0;
    uint64_t hsad = 0;
    for (int l = 0; l < pp->num_of_list_to_search; l++)
        for (int r = 0; r < pp->num_of_ref_pic_to_search[l]; r++) {
            if (pp->temporal_layer_index > 0 || l == 0) {
                if (c->enable_hme_flag) {
                    int lvl = -1;
                    /* the three blocks :2286, :2334, :2381 run in this order; the last that applies wins */
                    if (c->enable_hme_level0_flag && !c->enable_hme_level1_flag && !c->enable_hme_level2_flag) lvl = 0;
                    if (c->enable_hme_level1_flag && !c->enable_hme_level2_flag) lvl = 1;
                    if (c->enable_hme_level2_flag) lvl = 2;
                    if (lvl >= 0) {
                        hx   = x->xc[lvl][l][r][0][0];
                        hy   = x->yc[lvl][l][r][0][0];
                        hsad = x->hsad[lvl][l][r][0][0];
                        for (int sh = 0; sh < c->num_hme_sa_h; sh++)
                            for (int sw = (sh == 0 ? 1 : 0); sw < c->num_hme_sa_w; sw++)
                                if (x->hsad[lvl][l][r][sw][sh] < hsad) {
                                    hx   = x->xc[lvl][l][r][sw][sh];
                                    hy   = x->yc[lvl][l][r][sw][sh];
                                    hsad = x->hsad[lvl][l][r][sw][sh];
                                }
                    }
                    scx = hx;
                    scy = hy;
                }
            } else {
                scx = 0;
                scy = 0;
            }
            x->sr[l][r].hme_sc_x = scx;
            x->sr[l][r].hme_sc_y = scy;
            x->sr[l][r].hme_sad  = hsad;
        }
}

/* hme_prune_ref_and_adjust_sr, Codec/motion_estimation.c:2547-2588 (loops cover all [2][4]) */
static void hme_prune_ref_and_adjust_sr(Ctx *x) {
    const SvtMeCtrls *c  = x->c;
    uint16_t          th = c->prune_ref_if_hme_sad_dev_bigger_than_th;
    if (c->enable_me_hme_ref_pruning && th != (uint16_t)~0) {
        uint64_t best = (uint64_t)~0;
        for (int l = 0; l < 2; l++)
            for (int r = 0; r < 4; r++)
                if (x->sr[l][r].hme_sad < best) best = x->sr[l][r].hme_sad;
        for (int l = 0; l < 2; l++)
            for (int r = 1; r < 4; r++)
                if ((x->sr[l][r].hme_sad - best) * 100 > (th * best)) x->sr[l][r].do_ref = 0;
    }
    if (c->enable_me_sr_adjustment) {
        for (int l = 0; l < 2; l++)
            for (int r = 0; r < 4; r++) {
                if (ORC_ABS(x->sr[l][r].hme_sc_x) <= c->reduce_me_sr_based_on_mv_length_th &&
                    ORC_ABS(x->sr[l][r].hme_sc_y) <= c->reduce_me_sr_based_on_mv_length_th &&
                    x->sr[l][r].hme_sad < c->stationary_hme_sad_abs_th)
                    x->reduce_me_sr_divisor[l][r] = c->stationary_me_sr_divisor;
                else if (x->sr[l][r].hme_sad < c->reduce_me_sr_based_on_hme_sad_abs_th)
                    x->reduce_me_sr_divisor[l][r] = c->me_sr_divisor_for_low_hme_sad;
            }
    }
}

/* ====================================================================================== */
/* full-pel search                                                                         */
/* ====================================================================================== */

/* One search point of open_loop_me_fullpel_search_sblock (Codec/motion_estimation.c:429-817): the
   64 8x8 SADs, summed to 16x16 / 32x32 / 64x64, each compared (strict '<') with the running best
   of its PU. The reference visits points in raster order (8-position groups, then the remainder,
   :790-815), and for one PU that is plain raster order, so per-point evaluation gives the same
   winners. PU numbering inside p_sb_best_*: 0 = 64x64, 1..4 = 32x32, 5..20 = 16x16 (z-order),
   21..84 = 8x8 (z-order of the 16x16 parent * 4 + {TL,TR,BL,BR}) (me_context.h:52-138). */
static void fullpel_point(Ctx *x, int l, int r, const uint8_t *ref, uint32_t ref_stride, int16_t mvx, int16_t mvy) {
    const int      sub = x->c->me_search_method == SVTME_SUB_SAD_SEARCH;
    const uint32_t mv  = pack_mv(mvx, mvy);
    uint32_t      *bs = x->best_sad[l][r], *bm = x->best_mv[l][r];
    uint32_t       s16[16], s32[4] = {0, 0, 0, 0}, s64 = 0;
    for (int by = 0; by < 4; by++)
        for (int bx = 0; bx < 4; bx++) {
            const int z = k_z16[4 * by + bx];
            uint32_t  sum = 0;
            for (int k = 0; k < 4; k++) {
                const uint32_t px = (uint32_t)(16 * bx + 8 * (k & 1)), py = (uint32_t)(16 * by + 8 * (k >> 1));
                const uint32_t v  = sad8x8(x->b64_src + (size_t)py * x->b64_src_stride + px, x->b64_src_stride,
                                           ref + (size_t)py * ref_stride + px, ref_stride, sub);
                if (v < bs[21 + 4 * z + k]) {
                    bs[21 + 4 * z + k] = v;
                    bm[21 + 4 * z + k] = mv;
                }
                sum += v;
            }
            s16[z] = sum;
            if (sum < bs[5 + z]) {
                bs[5 + z] = sum;
                bm[5 + z] = mv;
            }
        }
    for (int k = 0; k < 4; k++) {
        s32[k] = s16[4 * k] + s16[4 * k + 1] + s16[4 * k + 2] + s16[4 * k + 3];
        if (s32[k] < bs[1 + k]) {
            bs[1 + k] = s32[k];
            bm[1 + k] = mv;
        }
        s64 += s32[k];
    }
    if (s64 < bs[0]) {
        bs[0] = s64;
        bm[0] = mv;
    }
}

/* open_loop_me_fullpel_search_sblock, Codec/motion_estimation.c:781-817; the ref pointer of point
   (xs,ys) is the window origin + xs + ys*stride (:443-446, :1563-1570 net to that) */
static void fullpel_search(Ctx *x, int l, int r, const OrcPlane *ref, int16_t x_origin, int16_t y_origin, uint32_t sa_w,
                           uint32_t sa_h) {
    int16_t x_tl = (int16_t)((int16_t)(ref->org_x + x->org_x) + x_origin);
    int16_t y_tl = (int16_t)((int16_t)(ref->org_y + x->org_y) + y_origin);
    const uint8_t *base = ref->buffer_y + (x_tl + y_tl * ref->stride_y);
    for (uint32_t ys = 0; ys < sa_h; ys++)
        for (uint32_t xs = 0; xs < sa_w; xs++)
            fullpel_point(x, l, r, base + xs + (size_t)ys * ref->stride_y, (uint32_t)ref->stride_y,
                          (int16_t)((int32_t)xs + x_origin), (int16_t)((int32_t)ys + y_origin));
    x->sad_ops += (uint64_t)sa_w * sa_h * 64 * (x->c->me_search_method == SVTME_SUB_SAD_SEARCH ? 32 : 64);
}

/* check_00_center, Codec/motion_estimation.c:1139-1206 (COST_PRECISION shift cancels in the MIN) */
static uint32_t check_00_center(Ctx *x, const OrcPlane *ref, int16_t *xc, int16_t *yc, uint32_t zz_sad) {
    const int16_t org_x = (int16_t)x->org_x, org_y = (int16_t)x->org_y, pad = 63;
    uint32_t index = (uint32_t)((int16_t)ref->org_x + org_x + ((int16_t)ref->org_y + org_y) * ref->stride_y);
    uint32_t zero_mv_sad;
    if (x->c->me_early_exit_th)
        zero_mv_sad = zz_sad;
    else {
        zero_mv_sad = orc_nxm_sad_kernel(x->b64_src, x->b64_src_stride << 1, ref->buffer_y + index,
                                         (uint32_t)ref->stride_y << 1, x->b64_h >> 1, x->b64_w);
        x->sad_ops += (uint64_t)(x->b64_h >> 1) * x->b64_w;
    }
    zero_mv_sad = zero_mv_sad << 1;
    *xc = ((org_x + *xc) < -pad) ? (int16_t)(-pad - org_x) : *xc;
    *xc = ((org_x + *xc) > (int16_t)ref->width - 1) ? (int16_t)(*xc - ((org_x + *xc) - ((int16_t)ref->width - 1))) : *xc;
    *yc = ((org_y + *yc) < -pad) ? (int16_t)(-pad - org_y) : *yc;
    *yc = ((org_y + *yc) > (int16_t)ref->height - 1) ? (int16_t)(*yc - ((org_y + *yc) - ((int16_t)ref->height - 1))) : *yc;
    uint64_t zero_mv_cost = (uint64_t)(zero_mv_sad << 8);
    index = (uint32_t)((int16_t)(ref->org_x + org_x) + *xc + ((int16_t)(ref->org_y + org_y) + *yc) * ref->stride_y);
    uint32_t hme_mv_sad = orc_nxm_sad_kernel(x->b64_src, x->b64_src_stride << 1, ref->buffer_y + index,
                                             (uint32_t)ref->stride_y << 1, x->b64_h >> 1, x->b64_w);
    x->sad_ops += (uint64_t)(x->b64_h >> 1) * x->b64_w;
    hme_mv_sad = hme_mv_sad << 1;
    uint64_t hme_mv_cost = (uint64_t)(hme_mv_sad << 8);
    uint64_t centre_cost = ORC_MIN(zero_mv_cost, hme_mv_cost);
    *xc = (centre_cost == zero_mv_cost) ? 0 : *xc;
    *yc = (centre_cost == zero_mv_cost) ? 0 : *yc;
    return hme_mv_sad;
}

/* integer_search_b64, Codec/motion_estimation.c:1249-1586 (restricted_motion_vector == 0) */
static void integer_search_b64(Ctx *x) {
    const SvtMePicParams *pp = x->pp;
    const SvtMeCtrls     *c  = x->c;
    const int16_t pic_w = (int16_t)x->pic_w, pic_h = (int16_t)x->pic_h, pad = 63;
    const int16_t org_x = (int16_t)x->org_x, org_y = (int16_t)x->org_y;
    for (int l = 0; l < pp->num_of_list_to_search; l++)
        for (int r = 0; r < pp->num_of_ref_pic_to_search[l]; r++) {
            const OrcPlane *ref = &x->ref[l][r].lvl[0];
            uint16_t dist = ref_dist(x, l, r);
            memset(&x->win[l][r], 0, sizeof(SvtMeWindow));
            if (x->sr[l][r].do_ref == 0) continue;
            int16_t xc = x->sr[l][r].hme_sc_x, yc = x->sr[l][r].hme_sc_y;
            int16_t sa_w = (int16_t)c->me_sa.sa_min.width, sa_h = (int16_t)c->me_sa.sa_min.height;
            if (c->me_type != SVTME_ME_MCTF) dist = scaled_dist(dist); /* :1302-1304 */
            sa_w = (int16_t)ORC_MIN((sa_w * dist), c->me_sa.sa_max.width);
            sa_h = (int16_t)ORC_MIN((sa_h * dist), c->me_sa.sa_max.height);
            if (c->mv_sa_adj_enabled && (!c->mv_sa_adj_nearest_ref_only || r == 0)) {
                if (ORC_ABS(xc) > c->mv_sa_adj_mv_size_th) sa_w = (int16_t)(sa_w * c->mv_sa_adj_sa_multiplier);
                if (ORC_ABS(yc) > c->mv_sa_adj_mv_size_th) sa_h = (int16_t)(sa_h * c->mv_sa_adj_sa_multiplier);
            }
            const uint32_t div = x->reduce_me_sr_divisor[l][r];
            sa_w = (int16_t)((ORC_MAX(1, ((uint32_t)sa_w / div)) + 7) & ~0x07);
            sa_h = (int16_t)ORC_MAX(3, ((uint32_t)sa_h / div));
            const int16_t sa_h0 = sa_h, sa_w0 = sa_w;
            uint64_t best_hme_sad = (uint64_t)~0;
            if (c->me_early_exit_th) {
                if (x->zz_sad[l][r] < (c->me_early_exit_th / 6)) sa_w = sa_h = 1;
            } else {
                uint8_t hme_is_accurate = 1;
                if ((xc != 0 || yc != 0) && pp->is_ref) {
                    best_hme_sad = check_00_center(x, ref, &xc, &yc, x->zz_sad[l][r]);
                    if (xc == 0 && yc == 0) hme_is_accurate = 0;
                }
                if (c->enable_me_sr_adjustment == 2) {
                    if ((hme_is_accurate && (best_hme_sad < (24 * 24))) || (pp->is_ref && x->sr[l][r].hme_sad < (24 * 24)))
                        sa_h = (int16_t)(sa_h / 2);
                    if ((l || r) && x->best_sad[0][0][0] < 5000 && sa_h == sa_h0 && sa_w == sa_w0) {
                        sa_h = (int16_t)(sa_h >> 1);
                        sa_w = (int16_t)(sa_w >> 1);
                    }
                }
            }
            orc_initialize_buffer_32bits(x->best_sad[l][r], 21, 1, MAX_SAD_VALUE);
            if (c->me_8x8_var_enabled && (sa_w * sa_h > 24)) {
                /* 1x1 probe at the search centre, then the variance of its 64 8x8 SADs (:1393-1441) */
                fullpel_search(x, l, r, ref, xc, yc, 1, 1);
                const uint32_t mean = x->best_sad[l][r][0] / 64;
                uint32_t       sumsq = 0;
                for (int i = 0; i < 64; i++) {
                    const int32_t d = (int32_t)x->best_sad[l][r][21 + i] - (int32_t)mean;
                    sumsq += (uint32_t)(d * d);
                }
                const uint32_t var = sumsq / 64;
                if (var > c->me_sr_mult2_th) {
                    sa_w = (int16_t)((ORC_MAX(1, sa_w * 3 / 2) + 7) & ~0x7);
                    sa_h = (int16_t)ORC_MAX(1, sa_h * 3 / 2);
                }
                if (var < c->me_sr_div4_th) {
                    sa_w = (int16_t)((ORC_MAX(1, sa_w >> 2) + 7) & ~0x7);
                    sa_h = (int16_t)ORC_MAX(1, sa_h >> 2);
                    sa_h = (int16_t)ORC_MAX(3, sa_h);
                } else if (var < c->me_sr_div2_th) {
                    sa_w = (int16_t)((ORC_MIN(sa_w, sa_w >> 1) + 7) & ~0x7);
                    sa_h = (int16_t)ORC_MIN(sa_h, sa_h >> 1);
                    sa_h = (int16_t)ORC_MAX(3, sa_h);
                }
            }
            int16_t x_origin = (int16_t)(xc - (sa_w >> 1));
            int16_t y_origin = (int16_t)(yc - (sa_h >> 1));
            clamp_axis(org_x, &x_origin, &sa_w, pad, pic_w);
            sa_w = (int16_t)((sa_w < 8) ? sa_w : sa_w & ~0x07);
            clamp_axis(org_y, &y_origin, &sa_h, pad, pic_h);
            x->win[l][r].x_origin = x_origin;
            x->win[l][r].y_origin = y_origin;
            x->win[l][r].width    = sa_w;
            x->win[l][r].height   = sa_h;
            fullpel_search(x, l, r, ref, x_origin, y_origin, (uint32_t)sa_w, (uint32_t)sa_h);
        }
}

static const uint32_t k_tab8x8[64] = {/* motion_estimation.h:107-116 */
    0,  1,  4,  5,  16, 17, 20, 21, 2,  3,  6,  7,  18, 19, 22, 23, 8,  9,  12, 13, 24, 25,
    28, 29, 10, 11, 14, 15, 26, 27, 30, 31, 32, 33, 36, 37, 48, 49, 52, 53, 34, 35, 38, 39,
    50, 51, 54, 55, 40, 41, 44, 45, 56, 57, 60, 61, 42, 43, 46, 47, 58, 59, 62, 63};

/* me_prune_ref, Codec/motion_estimation.c:1592-1635 */
static void me_prune_ref(Ctx *x) {
    const SvtMePicParams *pp = x->pp;
    const SvtMeCtrls     *c  = x->c;
    for (int l = 0; l < pp->num_of_list_to_search; l++)
        for (int r = 0; r < pp->num_of_ref_pic_to_search[l]; r++) {
            x->sr[l][r].hme_sad = 0;
            if (x->sr[l][r].do_ref == 0) {
                x->sr[l][r].hme_sad = (uint64_t)(MAX_SAD_VALUE * 64);
                continue;
            }
            for (int p = 0; p < 64; p++) x->sr[l][r].hme_sad += x->best_sad[l][r][21 + k_tab8x8[p]];
        }
    uint16_t th = c->prune_ref_if_me_sad_dev_bigger_than_th;
    if (c->enable_me_hme_ref_pruning && th != (uint16_t)~0) {
        uint64_t best = (uint64_t)~0;
        for (int l = 0; l < 2; l++)
            for (int r = 0; r < 4; r++)
                if (x->sr[l][r].hme_sad < best) best = x->sr[l][r].hme_sad;
        for (int l = 0; l < 2; l++)
            for (int r = 1; r < 4; r++)
                if ((x->sr[l][r].hme_sad - best) * 100 > (th * best)) x->sr[l][r].do_ref = 0;
    }
}

/* ====================================================================================== */
/* candidate lists, distortion statistics, GM detection                                    */
/* ====================================================================================== */

static const uint8_t k_z_to_raster[85] = {/* Codec/motion_estimation.c:2590-2601 */
    0,  1,  2,  3,  4,  5,  6,  9,  10, 7,  8,  11, 12, 13, 14, 17, 18, 15, 16, 19, 20, 21, 22, 29, 30, 23, 24, 31, 32,
    37, 38, 45, 46, 39, 40, 47, 48, 25, 26, 33, 34, 27, 28, 35, 36, 41, 42, 49, 50, 43, 44, 51, 52, 53, 54, 61, 62, 55,
    56, 63, 64, 69, 70, 77, 78, 71, 72, 79, 80, 57, 58, 65, 66, 59, 60, 67, 68, 73, 74, 81, 82, 75, 76, 83, 84};

/* MeCandidate is a 1-byte bit-field (me_sb_results.h:28-34): direction:2 | ref_idx_l0:2 |
   ref_idx_l1:2 | ref0_list:1 | ref1_list:1, LSB first (GCC, little-endian). The reference stores
   24 into the 1-bit list fields for "none"; that truncates to 0. */
static uint8_t cand_byte(unsigned dir, unsigned i0, unsigned i1, unsigned l0, unsigned l1) {
    return (uint8_t)((dir & 3) | ((i0 & 3) << 2) | ((i1 & 3) << 4) | ((l0 & 1) << 6) | ((l1 & 1) << 7));
}

typedef struct SbOut { /* MeSbResults of the current b64 */
    uint8_t  *total;
    uint32_t *mv;
    uint8_t  *cand;
} SbOut;

static int use_me_pu(const SvtMePicParams *pp, uint32_t n_idx) {
    return pp->enable_me_16x16 ? (pp->enable_me_8x8 || n_idx < 21) : (n_idx < 5);
}
static uint32_t number_of_pus(const SvtMePicParams *pp) { return pp->enable_me_16x16 ? (pp->enable_me_8x8 ? 85u : 21u) : 5u; }

/* construct_me_candidate_array_single_ref, Codec/motion_estimation.c:2716-2767 */
static void cand_single_ref(Ctx *x, SbOut *o) {
    const SvtMePicParams *pp = x->pp;
    memset(o->total, 1, number_of_pus(pp));
    for (uint32_t n = 0; n < 85; n++) {
        const uint8_t pu       = k_z_to_raster[n];
        x->me_distortion[pu] = x->best_sad[0][0][n];
        if (x->sr[0][0].do_ref == 0) continue;
        if (use_me_pu(pp, n)) {
            o->cand[pu * pp->max_cand] = cand_byte(0, 0, 0, 0, 0);
            o->mv[pu * pp->max_refs]   = x->best_mv[0][0][n];
        }
    }
}

/* construct_me_candidate_array_mrp_off, Codec/motion_estimation.c:2602-2715 */
static void cand_mrp_off(Ctx *x, SbOut *o, uint32_t n_lists) {
    const SvtMePicParams *pp = x->pp;
    uint8_t org[2];
    org[0] = x->sr[0][0].do_ref;
    org[1] = (n_lists == 1) ? 0 : x->sr[1][0].do_ref;
    if (n_lists < 2 || !x->sr[1][0].do_ref) n_lists = 1;
    const uint32_t th = (org[0] && org[1]) ? (uint32_t)x->c->prune_me_candidates_th : 0;
    memset(o->total, 1, number_of_pus(pp));
    for (uint32_t n = 0; n < 85; n++) {
        const uint8_t pu  = k_z_to_raster[n];
        const int     use = use_me_pu(pp, n);
        uint8_t       off = 0;
        uint8_t       doref[2] = {org[0], org[1]};
        const uint32_t best = (org[0] && org[1]) ? ORC_MIN(x->best_sad[0][0][n], x->best_sad[1][0][n])
            : org[0]                             ? x->best_sad[0][0][n]
                                                 : x->best_sad[1][0][n];
        x->me_distortion[pu] = best;
        int min_list = -1;
        if (x->c->use_best_unipred_cand_only && doref[0] && doref[1]) min_list = x->best_sad[0][0][n] < x->best_sad[1][0][n] ? 0 : 1;
        for (uint32_t l = 0; l < n_lists && (use || off == 0); l++) {
            if (doref[l] == 0) continue;
            if (th > 0) {
                uint32_t dev = (x->best_sad[l][0][n] - best) * 100;
                if (dev > (best * th)) {
                    doref[l] = 0;
                    continue;
                }
            }
            if (min_list != -1 && min_list != (int)l) {
                if (use) o->mv[pu * pp->max_refs + (l ? pp->max_l0 : 0)] = x->best_mv[l][0][n];
                continue;
            }
            if (use) {
                o->cand[pu * pp->max_cand + off] = cand_byte(l, 0, 0, l == 0 ? l : 24, l == 1 ? l : 24);
                o->mv[pu * pp->max_refs + (l ? pp->max_l0 : 0)] = x->best_mv[l][0][n];
            }
            off++;
        }
        if (doref[0] && doref[1] && use) {
            o->cand[pu * pp->max_cand + off] = cand_byte(2 /* BI_PRED */, 0, 0, 0, 1);
            o->total[pu]                     = (uint8_t)(off + 1);
        }
    }
}

/* construct_me_candidate_array, Codec/motion_estimation.c:2768-2905 */
static void cand_general(Ctx *x, SbOut *o, uint32_t n_lists) {
    const SvtMePicParams *pp  = x->pp;
    const uint8_t        *nref = pp->num_of_ref_pic_to_search;
    for (uint32_t n = 0; n < 85; n++) {
        const uint8_t  pu  = (n > 4) ? k_z_to_raster[n] : (uint8_t)n;
        const int      use = use_me_pu(pp, n);
        uint8_t        off = 0;
        uint8_t        doref[2][4];
        const uint32_t th   = (uint32_t)x->c->prune_me_candidates_th;
        uint32_t       best = MAX_U32;
        for (uint32_t l = 0; l < n_lists; l++)
            for (uint32_t r = 0; r < nref[l]; r++) {
                doref[l][r] = x->sr[l][r].do_ref;
                if (doref[l][r] == 0) continue;
                best = x->best_sad[l][r][n] < best ? x->best_sad[l][r][n] : best;
            }
        x->me_distortion[pu] = best;
        for (uint32_t l = 0; l < n_lists && (use || off == 0); l++)
            for (uint32_t r = 0; r < nref[l] && (use || off == 0); r++) {
                if (doref[l][r] == 0) continue;
                if (th > 0) {
                    uint32_t dev = (x->best_sad[l][r][n] - best) * 100;
                    if (dev > (best * th)) {
                        doref[l][r] = 0;
                        continue;
                    }
                }
                if (use) {
                    o->cand[pu * pp->max_cand + off]                    = cand_byte(l, r, r, l == 0 ? l : 24, l == 1 ? l : 24);
                    o->mv[pu * pp->max_refs + (l ? pp->max_l0 : 0) + r] = x->best_mv[l][r][n];
                }
                off++;
            }
        if (n_lists == 2 && use) {
            for (uint32_t a = 0; a < nref[0]; a++)
                for (uint32_t b = 0; b < nref[1]; b++) {
                    if (pp->only_l_bwd && (a > 0 || b > 0)) continue;
                    if (doref[0][a] && doref[1][b]) o->cand[pu * pp->max_cand + off++] = cand_byte(2, a, b, 0, 1);
                }
            if (!pp->only_l_bwd) {
                for (uint32_t a = 1; a < nref[0]; a++)
                    if (doref[0][0] && doref[0][a]) o->cand[pu * pp->max_cand + off++] = cand_byte(2, 0, a, 0, 0);
                if (nref[1] == 3 && doref[1][0] && doref[1][2]) o->cand[pu * pp->max_cand + off++] = cand_byte(2, 0, 2, 1, 1);
            }
        }
        if (use) o->total[pu] = off;
    }
}

/* compute_distortion, Codec/motion_estimation.c:3034-3077 */
static void compute_distortion(Ctx *x, SvtMePicResults *res) {
    const uint32_t b = x->b64_index;
    uint32_t d64 = x->me_distortion[0], d32 = 0, d16 = 0, d8 = 0;
    for (int i = 0; i < 4; i++) d32 += x->me_distortion[1 + i];
    for (int i = 0; i < 16; i++) d16 += x->me_distortion[5 + i];
    for (int i = 0; i < 64; i++) d8 += x->me_distortion[21 + i];
    uint64_t mean = d8 / 64, sumsq = 0;
    for (int i = 0; i < 64; i++) {
        const int64_t diff = (int64_t)x->me_distortion[21 + i] - (int64_t)mean;
        sumsq += (uint64_t)(diff * diff);
    }
    const uint32_t pix = x->b64_geom_w * x->b64_geom_h, size = 64 * 64;
    if (res->me_8x8_cost_variance) res->me_8x8_cost_variance[b] = (uint32_t)(sumsq / 64);
    if (res->rc_me_distortion) res->rc_me_distortion[b] = (x->pp->input_resolution <= 2 /* 480p range */) ? d8 : d16;
    if (res->me_64x64_distortion) res->me_64x64_distortion[b] = (d64 * size) / pix;
    if (res->me_32x32_distortion) res->me_32x32_distortion[b] = (d32 * size) / pix;
    if (res->me_16x16_distortion) res->me_16x16_distortion[b] = (d16 * size) / pix;
    if (res->me_8x8_distortion) res->me_8x8_distortion[b] = (d8 * size) / pix;
}

/* perform_gm_detection, Codec/motion_estimation.c:2908-3031 (enable_me_16x16 == 1). The candidate
   is fetched with the z-order index n_idx, as the reference does. */
static void gm_detection(Ctx *x, const SbOut *o, uint8_t *stationary, uint8_t *allow_gm) {
    const SvtMePicParams *pp = x->pp;
    uint64_t cnt[2][4][2][2];
    uint64_t stationary_cnt = 0, tot = 0;
    memset(cnt, 0, sizeof(cnt));
    const int low_res = pp->input_resolution <= 2;
    const unsigned n_blk = low_res ? 64 : 16;
    for (unsigned i = 0; i < n_blk; i++) {
        uint8_t n = (uint8_t)((low_res ? 21 : 5) + i);
        if (low_res && !pp->enable_me_8x8 && n >= 21) n = (uint8_t)(5 + ((i >> 3) >> 1) * 4 + ((i & 7) >> 1)); /* definitions.h:2626-2638 */
        const uint8_t cb  = o->cand[n * pp->max_cand];
        const unsigned dir = cb & 3;
        const uint32_t l = (dir == 0 || dir == 2) ? ((cb >> 6) & 1) : ((cb >> 7) & 1);
        const uint32_t r = (dir == 0 || dir == 2) ? ((cb >> 2) & 3) : ((cb >> 4) & 3);
        int active_th;
        if (low_res) {
            uint64_t a = pp->picture_number, b = pp->ref_picture_number[l][r];
            uint16_t dist = (uint16_t)ORC_ABS((int16_t)(ORC_MAX(a, b) - ORC_MIN(a, b)));
            active_th = pp->gm_use_distance_based_active_th ? ORC_MAX(dist >> 1, 4) : 4;
        } else {
            uint16_t dist = (uint16_t)ORC_ABS((int16_t)(pp->picture_number - pp->ref_picture_number[l][r]));
            active_th = pp->gm_use_distance_based_active_th ? ORC_MAX(dist * 16, 32) : 32;
        }
        const int mx = ((int16_t)(x->best_mv[l][r][n] & 0xFFFF)) << 2;
        const int my = ((int16_t)(x->best_mv[l][r][n] >> 16)) << 2;
        if (mx < -active_th) cnt[l][r][0][0]++;
        else if (mx > active_th) cnt[l][r][0][1]++;
        if (my < -active_th) cnt[l][r][1][0]++;
        else if (my > active_th) cnt[l][r][1][1]++;
        const int st = low_res ? 0 : 4;
        if (abs(mx) <= st && abs(my) <= st) stationary_cnt++;
        tot++;
    }
    if (stationary_cnt > ((tot * 5) / 100)) *stationary = 1;
    for (int l = 0; l < 2; l++)
        for (int r = 0; r < 4; r++)
            for (int cc = 0; cc < 2; cc++)
                for (int s = 0; s < 2; s++)
                    if (cnt[l][r][cc][s] > (tot / 2)) *allow_gm = 1;
}

/* ====================================================================================== */
/* per-b64 orchestration and the picture loop                                              */
/* ====================================================================================== */

/* init_me_hme_data, Codec/motion_estimation.c:3080-3140. Not reset (on purpose, as in the
   reference): hsad[], prehme mv/sad, hme_sc_x/y, best_sad. */
static void init_me_hme_data(Ctx *x) {
    if (x->c->enable_hme_flag) {
        memset(x->xc, 0, sizeof(x->xc));
        memset(x->yc, 0, sizeof(x->yc));
    }
    memset(x->best_mv, 0, sizeof(x->best_mv));
    for (int l = 0; l < 2; l++)
        for (int r = 0; r < 4; r++) {
            x->sr[l][r].do_ref            = 1;
            x->sr[l][r].hme_sad           = MAX_U32;
            x->reduce_me_sr_divisor[l][r] = 1;
            x->zz_sad[l][r]               = MAX_U32;
            x->prehme[l][r][0].valid      = 0;
            x->prehme[l][r][1].valid      = 0;
        }
    memset(x->performed_phme, 0, sizeof(x->performed_phme));
}

/* svt_aom_motion_estimation_b64, Codec/motion_estimation.c:3146-3223 (ME_OPEN_LOOP and ME_MCTF). Returns 1 when the
   temporal-filter early exit skipped the full-pel search (:3179-3183): p_sb_best_* of this b64 are then not produced. */
static int motion_estimation_b64(Ctx *x, SvtMePicResults *res, SbOut *o) {
    const SvtMePicParams *pp = x->pp;
    const SvtMeCtrls     *c  = x->c;
    const uint32_t n_lists = pp->num_of_list_to_search;
    const int      mctf = c->me_type == SVTME_ME_MCTF;
    const uint8_t  prune_ref = c->enable_hme_flag && !mctf; /* :3173 */
    init_me_hme_data(x);
    /* hme_b64 :2511-2545 */
    if (c->me_early_exit_th || c->me_safe_limit_zz_th) init_zz_sad(x);
    if (c->prehme_enable) prehme_b64(x);
    if (c->enable_hme_flag) {
        if (c->enable_hme_level0_flag) hme_level0_b64(x);
        if (c->enable_hme_level1_flag) hme_level1_b64(x);
        if (c->enable_hme_level2_flag) hme_level2_b64(x);
    }
    set_final_search_centre(x);
    if (mctf && x->sr[0][0].hme_sad < c->tf_me_exit_th) return 1;
    if (prune_ref) hme_prune_ref_and_adjust_sr(x);
    integer_search_b64(x);
    if (prune_ref && c->enable_me_hme_ref_pruning) me_prune_ref(x);
    if (mctf) return 0; /* :3196: no candidate lists, statistics or GM flags for the temporal filter */
    if (pp->num_of_ref_pic_to_search[0] == 1 && pp->num_of_ref_pic_to_search[1] == 0)
        cand_single_ref(x, o);
    else if (pp->num_of_ref_pic_to_search[0] == 1 && pp->num_of_ref_pic_to_search[1] == 1)
        cand_mrp_off(x, o, n_lists);
    else
        cand_general(x, o, n_lists);
    compute_distortion(x, res);
    uint8_t stationary = 0, allow_gm = 0;
    if (pp->gm_enabled) gm_detection(x, o, &stationary, &allow_gm);
    if (res->stationary_block_present_sb) res->stationary_block_present_sb[x->b64_index] = stationary;
    if (res->rc_me_allow_gm) res->rc_me_allow_gm[x->b64_index] = allow_gm;
    return 0;
}

static void pic_alloc(OrcPic *p, const uint8_t *luma, int w, int h, uint8_t *mem[3]) {
    const int dims[3][3] = {{w, h, ORC_FULL_PAD}, {w >> 1, h >> 1, ORC_QUARTER_PAD}, {w >> 2, h >> 2, ORC_SIXTEENTH_PAD}};
    for (int i = 0; i < 3; i++) {
        const int s = dims[i][0] + 2 * dims[i][2];
        mem[i]      = (uint8_t *)calloc((size_t)s * (size_t)(dims[i][1] + 2 * dims[i][2]), 1);
        p->lvl[i].buffer_y = mem[i];
        p->lvl[i].stride_y = s;
        p->lvl[i].org_x = p->lvl[i].org_y = dims[i][2];
        p->lvl[i].width  = dims[i][0];
        p->lvl[i].height = dims[i][1];
    }
    orc_build_planes(luma, w, h, mem[0], mem[1], mem[2]);
}

/* b64 loop of svt_aom_motion_estimation_kernel, Codec/me_process.c:174-271 */
int orc_me_picture(const uint8_t *src_luma, const uint8_t *const ref_luma[2][4], int width, int height,
                   const SvtMePicParams *params, SvtMePicResults *res_in, OrcDebug *dbg, uint64_t *sad_ops) {
    if (!src_luma || !params || width <= 0 || height <= 0 || (width & 7) || (height & 7)) return -1;
    if (params->num_of_list_to_search < 1 || params->num_of_list_to_search > 2) return -1;
    if (!params->enable_me_16x16) return -1;
    SvtMePicResults none;
    memset(&none, 0, sizeof(none));
    SvtMePicResults *res = res_in ? res_in : &none;
    Ctx *x = (Ctx *)calloc(1, sizeof(Ctx));
    x->pp        = params;
    x->c         = &params->ctrls;
    x->pic_w     = width;
    x->pic_h     = height;
    x->hme_l0_sa = params->ctrls.hme_l0_sa;
    uint8_t *src_mem[3], *ref_mem[2][4][3];
    pic_alloc(&x->src, src_luma, width, height, src_mem);
    for (int l = 0; l < params->num_of_list_to_search; l++)
        for (int r = 0; r < params->num_of_ref_pic_to_search[l]; r++) {
            if (!ref_luma[l][r]) return -1;
            pic_alloc(&x->ref[l][r], ref_luma[l][r], width, height, ref_mem[l][r]);
        }
    const uint32_t bw = (uint32_t)(width + 63) / 64, bh = (uint32_t)(height + 63) / 64;
    const uint32_t n_pus = params->enable_me_8x8 ? 85u : 21u;
    uint8_t  *tmp_total = (uint8_t *)calloc(n_pus, 1);
    uint32_t *tmp_mv    = (uint32_t *)calloc((size_t)n_pus * ORC_MAX(1, params->max_refs), 4);
    uint8_t  *tmp_cand  = (uint8_t *)calloc((size_t)n_pus * ORC_MAX(1, params->max_cand), 1);
    for (uint32_t by = 0; by < bh; by++)
        for (uint32_t bx = 0; bx < bw; bx++) {
            const uint32_t b = (uint16_t)(bx + by * bw);
            x->b64_index = b;
            x->org_x     = bx * 64;
            x->org_y     = by * 64;
            /* motion_estimation.c:3163-3170 (8-aligned picture) and pcs.c:1573-1579 */
            x->b64_w = x->b64_geom_w = (uint32_t)ORC_MIN(64, width - (int)x->org_x);
            x->b64_h = x->b64_geom_h = (uint32_t)ORC_MIN(64, height - (int)x->org_y);
            const OrcPlane *f = &x->src.lvl[0], *q = &x->src.lvl[1], *s = &x->src.lvl[2];
            x->b64_src          = f->buffer_y + (size_t)(f->org_y + x->org_y) * f->stride_y + f->org_x + x->org_x;
            x->b64_src_stride   = (uint32_t)f->stride_y;
            x->quarter_src      = q->buffer_y + (size_t)(q->org_y + (x->org_y >> 1)) * q->stride_y + q->org_x + (x->org_x >> 1);
            x->quarter_stride   = (uint32_t)q->stride_y;
            x->sixteenth_src    = s->buffer_y + (size_t)(s->org_y + (x->org_y >> 2)) * s->stride_y + s->org_x + (x->org_x >> 2);
            x->sixteenth_stride = (uint32_t)s->stride_y;
            SbOut o;
            o.total = res->total_me_candidate_index ? res->total_me_candidate_index + (size_t)b * n_pus : tmp_total;
            o.mv    = res->me_mv_array ? res->me_mv_array + (size_t)b * n_pus * params->max_refs : tmp_mv;
            o.cand  = res->me_candidate_array ? res->me_candidate_array + (size_t)b * n_pus * params->max_cand : tmp_cand;
            memset(o.total, 0, n_pus);
            memset(o.mv, 0, (size_t)n_pus * params->max_refs * 4);
            memset(o.cand, 0, (size_t)n_pus * params->max_cand);
            const int tf_exit = motion_estimation_b64(x, res, &o);
            for (int l = 0; l < 2; l++)
                for (int r = 0; r < 4; r++) {
                    const size_t k        = ((size_t)b * 2 + (size_t)l) * 4 + (size_t)r;
                    const int    searched = l < params->num_of_list_to_search && r < params->num_of_ref_pic_to_search[l];
                    const int    valid    = searched && x->sr[l][r].do_ref && !tf_exit;
                    for (int p = 0; p < 85; p++) {
                        if (res->p_sb_best_sad) res->p_sb_best_sad[k * 85 + (size_t)p] = valid ? x->best_sad[l][r][p] : 0;
                        if (res->p_sb_best_mv) res->p_sb_best_mv[k * 85 + (size_t)p] = valid ? x->best_mv[l][r][p] : 0;
                    }
                    if (res->search_results) {
                        SvtMeSearchResult *o2 = &res->search_results[k];
                        memset(o2, 0, sizeof(*o2));
                        if (searched) {
                            o2->hme_sc_x = x->sr[l][r].hme_sc_x;
                            o2->hme_sc_y = x->sr[l][r].hme_sc_y;
                            o2->do_ref   = x->sr[l][r].do_ref;
                            o2->hme_sad  = x->sr[l][r].hme_sad;
                        }
                    }
                    if (res->me_windows) {
                        memset(&res->me_windows[k], 0, sizeof(SvtMeWindow));
                        if (searched && !tf_exit) res->me_windows[k] = x->win[l][r];
                    }
                }
            if (dbg) {
                OrcDebug *d = &dbg[b];
                memcpy(d->x_hme_l0, x->xc[0], sizeof(d->x_hme_l0));
                memcpy(d->y_hme_l0, x->yc[0], sizeof(d->y_hme_l0));
                memcpy(d->x_hme_l1, x->xc[1], sizeof(d->x_hme_l1));
                memcpy(d->y_hme_l1, x->yc[1], sizeof(d->y_hme_l1));
                memcpy(d->x_hme_l2, x->xc[2], sizeof(d->x_hme_l2));
                memcpy(d->y_hme_l2, x->yc[2], sizeof(d->y_hme_l2));
                memcpy(d->hme_l0_sad, x->hsad[0], sizeof(d->hme_l0_sad));
                memcpy(d->hme_l1_sad, x->hsad[1], sizeof(d->hme_l1_sad));
                memcpy(d->hme_l2_sad, x->hsad[2], sizeof(d->hme_l2_sad));
                for (int l = 0; l < 2; l++)
                    for (int r = 0; r < 4; r++) {
                        for (int s2 = 0; s2 < 2; s2++) {
                            d->prehme_mv_x[l][r][s2] = x->prehme[l][r][s2].mv_x;
                            d->prehme_mv_y[l][r][s2] = x->prehme[l][r][s2].mv_y;
                            d->prehme_sad[l][r][s2]  = x->prehme[l][r][s2].sad;
                        }
                        d->zz_sad[l][r]               = x->zz_sad[l][r];
                        d->reduce_me_sr_divisor[l][r] = x->reduce_me_sr_divisor[l][r];
                    }
            }
        }
    if (sad_ops) *sad_ops = x->sad_ops;
    free(tmp_total);
    free(tmp_mv);
    free(tmp_cand);
    for (int i = 0; i < 3; i++) free(src_mem[i]);
    for (int l = 0; l < params->num_of_list_to_search; l++)
        for (int r = 0; r < params->num_of_ref_pic_to_search[l]; r++)
            for (int i = 0; i < 3; i++) free(ref_mem[l][r][i]);
    free(x);
    return 0;
}


/* ====================================================================================== */
/* Dynamic-GOP detector HME: dg_detector_hme_level0 + early_hme_b64, Codec/pd_process.c:392-589 */
/* ====================================================================================== */
int orc_dg_detector(const uint8_t *src_luma, const uint8_t *ref_luma, int width, int height, SvtMeDgMetrics *m,
                    uint64_t *b64_sad, int16_t *b64_mv) {
    if (!src_luma || !ref_luma || !m || width <= 0 || height <= 0 || (width & 7) || (height & 7)) return -1;
    const int sw = width >> 2, sh = height >> 2, pad = ORC_SIXTEENTH_PAD, stride = sw + 2 * pad;
    const size_t fb = (size_t)(width + 2 * ORC_FULL_PAD) * (height + 2 * ORC_FULL_PAD);
    const size_t qb = (size_t)((width >> 1) + 2 * ORC_QUARTER_PAD) * ((height >> 1) + 2 * ORC_QUARTER_PAD);
    const size_t sb = (size_t)stride * (sh + 2 * pad);
    uint8_t *full = (uint8_t *)calloc(fb, 1), *quarter = (uint8_t *)calloc(qb, 1);
    uint8_t *s16 = (uint8_t *)calloc(sb, 1), *r16 = (uint8_t *)calloc(sb, 1);
    orc_build_planes(src_luma, width, height, full, quarter, s16);
    orc_build_planes(ref_luma, width, height, full, quarter, r16);
    /* :494-495: INPUT_SIZE_360p_RANGE = 1, INPUT_SIZE_480p_RANGE = 2 (definitions.h:2088-2090) */
    const int16_t sa0 = m->input_resolution <= 1 ? 16 : m->input_resolution <= 2 ? 64 : 128;
    const uint32_t cols = (uint32_t)(width + 63) / 64, rows = (uint32_t)(height + 63) / 64;
    m->tot_dist = 0; m->tot_cplx = 0; m->tot_active = 0; m->sum_in_vectors = 0;
    /* declared outside the b64 loops as in the reference (:499-500): svt_sad_loop_kernel resets best_sad but leaves the
       centre untouched when nothing beats 0xffffff (cannot happen for a 16x16 8-bit block) */
    uint64_t sad = ~0ull;
    int16_t  col = 0, row = 0;
    for (uint32_t by = 0; by < rows; by++) {
        for (uint32_t bx = 0; bx < cols; bx++) {
            /* early_hme_b64 :392-487, all in int16_t as there */
            const int16_t org_x = (int16_t)(((int16_t)(bx * 64)) >> 2), org_y = (int16_t)(((int16_t)(by * 64)) >> 2);
            int16_t sa_w = (int16_t)((sa0 + 7) & ~0x07), sa_h = sa0;
            const int16_t pad_w = (int16_t)pad - 1, pad_h = (int16_t)pad - 1;
            int16_t ox = (int16_t)(-(sa_w >> 1)), oy = (int16_t)(-(sa_h >> 1));
            if ((org_x + ox) < -pad_w) { /* :424-427: origin reassigned first: the width is not reduced */
                ox   = (int16_t)(-pad_w - org_x);
                sa_w = (int16_t)(sa_w - (-pad_w - (org_x + ox)));
            }
            if ((org_x + ox) > (int16_t)sw - 1) ox = (int16_t)(ox - ((org_x + ox) - ((int16_t)sw - 1)));
            if ((org_x + ox + sa_w) > (int16_t)sw) {
                const int v = sa_w - ((org_x + ox + sa_w) - (int16_t)sw);
                sa_w = (int16_t)(v > 1 ? v : 1);
            }
            sa_w = (int16_t)((sa_w < 8) ? sa_w : sa_w & ~0x07);
            if ((org_y + oy) < -pad_h) {
                oy   = (int16_t)(-pad_h - org_y);
                sa_h = (int16_t)(sa_h - (-pad_h - (org_y + oy)));
            }
            if ((org_y + oy) > (int16_t)sh - 1) oy = (int16_t)(oy - ((org_y + oy) - ((int16_t)sh - 1)));
            if ((org_y + oy + sa_h) > (int16_t)sh) {
                const int v = sa_h - ((org_y + oy + sa_h) - (int16_t)sh);
                sa_h = (int16_t)(v > 1 ? v : 1);
            }
            const int16_t x_tl = (int16_t)(((int16_t)pad + org_x) + ox), y_tl = (int16_t)(((int16_t)pad + org_y) + oy);
            const uint8_t *src = s16 + (size_t)(pad + (by * 64 >> 2)) * stride + pad + (bx * 64 >> 2);
            orc_sad_loop_kernel(src, (uint32_t)stride, r16 + (x_tl + y_tl * stride), (uint32_t)stride, 16, 16, &sad, &col, &row,
                                (uint32_t)stride, 0, sa_w, sa_h);
            col = (int16_t)(col + ox); col = (int16_t)(col * 4);
            row = (int16_t)(row + oy); row = (int16_t)(row * 4);
            m->tot_dist += sad;
            m->tot_cplx += (sad > (16 * 16 * 30));
            m->tot_active += ((abs(col) > 0) || (abs(row) > 0));
            if (by < rows / 2) { if (row > 0) --m->sum_in_vectors; else if (row < 0) ++m->sum_in_vectors; }
            else if (by > rows / 2) { if (row > 0) ++m->sum_in_vectors; else if (row < 0) --m->sum_in_vectors; }
            if (bx < cols / 2) { if (col > 0) --m->sum_in_vectors; else if (col < 0) ++m->sum_in_vectors; }
            else if (bx > cols / 2) { if (col > 0) ++m->sum_in_vectors; else if (col < 0) --m->sum_in_vectors; }
            if (b64_sad) b64_sad[by * cols + bx] = sad;
            if (b64_mv) { b64_mv[2 * (by * cols + bx)] = col; b64_mv[2 * (by * cols + bx) + 1] = row; }
        }
    }
    free(full); free(quarter); free(s16); free(r16);
    return 0;
}
