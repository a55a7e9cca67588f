/*
 * ref_harness.c — TEST INFRASTRUCTURE ONLY. Driver around the UNMODIFIED reference sources.
 *
 * Compiled by oracle/Makefile together with these reference files (from /root/reference, not
 * copied): Codec/motion_estimation.c, Codec/enc_mode_config.c, Codec/me_sad_calculation.c,
 * Codec/pic_analysis_process.c, Codec/pic_operators.c, Codec/svt_log.c, C_DEFAULT/compute_sad_c.c,
 * C_DEFAULT/pack_unpack_c.c and the intrinsics-only SIMD kernels ASM_AVX2/compute_sad_intrin_avx2.c,
 * ASM_AVX512/compute_sad_intrin_avx512.c, ASM_SSE4_1/compute_sad_intrin_sse4_1.c,
 * ASM_SSE2/me_sad_calc_intrin_sse2.c. Symbols those objects reference but the ME path never
 * calls are resolved by generated abort() stubs (oracle/gen_stubs.py).
 *
 * What this file restates (because the reference's own code for it lives in files that need the
 * whole encoder): the b64 loop and pointer set-up of svt_aom_motion_estimation_kernel
 * (me_process.c:174-271), the PA-reference buffer geometry (reference_object.c:320-380,
 * enc_handle.c:4485-4497), the HME enable flags (enc_mode_config.c:1820-1831),
 * use_best_me_unipred_cand_only (enc_mode_config.c:1984-1987), b64_geom (pcs.c:1552-1606) and
 * the ME result sizing (pcs.c:91-96, pd_process.c:3495-3511). Everything else — signal
 * derivation, HME, integer search, candidate construction, statistics — is the reference's code.
 */
#define RTCD_C
#define AOM_RTCD_C
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#include <pthread.h>

#include "definitions.h"
#include "common_dsp_rtcd.h"
#include "aom_dsp_rtcd.h"
#include "pcs.h"
#include "sequence_control_set.h"
#include "me_context.h"
#include "motion_estimation.h"
#include "enc_mode_config.h"
#include "pic_operators.h"
#include "compute_sad_c.h"

#include "ref_harness.h"

/* SIMD kernels of the reference (declared in its ASM_* headers; declared here to avoid pulling
   every ASM header) */
void svt_sad_loop_kernel_avx2_intrin(uint8_t *src, uint32_t src_stride, uint8_t *ref, uint32_t ref_stride,
                                     uint32_t block_height, uint32_t block_width, uint64_t *best_sad,
                                     int16_t *x_search_center, int16_t *y_search_center, uint32_t src_stride_raw,
                                     uint8_t skip_search_line, int16_t search_area_width,
                                     int16_t search_area_height);
void svt_sad_loop_kernel_avx512_intrin(uint8_t *src, uint32_t src_stride, uint8_t *ref, uint32_t ref_stride,
                                       uint32_t block_height, uint32_t block_width, uint64_t *best_sad,
                                       int16_t *x_search_center, int16_t *y_search_center,
                                       uint32_t src_stride_raw, uint8_t skip_search_line,
                                       int16_t search_area_width, int16_t search_area_height);
void svt_ext_all_sad_calculation_8x8_16x16_avx2(uint8_t *src, uint32_t src_stride, uint8_t *ref,
                                                uint32_t ref_stride, uint32_t mv, uint32_t *p_best_sad_8x8,
                                                uint32_t *p_best_sad_16x16, uint32_t *p_best_mv8x8,
                                                uint32_t *p_best_mv16x16, uint32_t p_eight_sad16x16[16][8],
                                                uint32_t p_eight_sad8x8[64][8], Bool sub_sad);
void svt_ext_eight_sad_calculation_32x32_64x64_avx2(uint32_t p_sad16x16[16][8], uint32_t *p_best_sad_32x32,
                                                    uint32_t *p_best_sad_64x64, uint32_t *p_best_mv32x32,
                                                    uint32_t *p_best_mv64x64, uint32_t mv,
                                                    uint32_t p_sad32x32[4][8]);
void svt_ext_sad_calculation_8x8_16x16_avx2_intrin(uint8_t *src, uint32_t src_stride, uint8_t *ref,
                                                   uint32_t ref_stride, uint32_t *p_best_sad_8x8,
                                                   uint32_t *p_best_sad_16x16, uint32_t *p_best_mv8x8,
                                                   uint32_t *p_best_mv16x16, uint32_t mv, uint32_t *p_sad16x16,
                                                   uint32_t *p_sad8x8, Bool sub_sad);
void svt_ext_sad_calculation_32x32_64x64_sse4_intrin(uint32_t *p_sad16x16, uint32_t *p_best_sad_32x32,
                                                     uint32_t *p_best_sad_64x64, uint32_t *p_best_mv32x32,
                                                     uint32_t *p_best_mv64x64, uint32_t mv,
                                                     uint32_t *p_sad32x32);
uint32_t svt_nxm_sad_kernel_helper_avx2(const uint8_t *src, uint32_t src_stride, const uint8_t *ref,
                                        uint32_t ref_stride, uint32_t height, uint32_t width);
uint32_t svt_nxm_sad_kernel_sub_sampled_helper_avx2(const uint8_t *src, uint32_t src_stride,
                                                    const uint8_t *ref, uint32_t ref_stride, uint32_t height,
                                                    uint32_t width);
void svt_initialize_buffer_32bits_sse2_intrin(uint32_t *pointer, uint32_t count128, uint32_t count32,
                                              uint32_t value);
void svt_initialize_buffer_32bits_c(uint32_t *pointer, uint32_t count128, uint32_t count32, uint32_t value);
void svt_aom_downsample_2d_c(uint8_t *input_samples, uint32_t input_stride, uint32_t input_area_width,
                             uint32_t input_area_height, uint8_t *decim_samples, uint32_t decim_stride,
                             uint32_t decim_step);
void svt_unpack_and_2bcompress_c(uint16_t *in16b_buffer, uint32_t in16b_stride, uint8_t *out8b_buffer,
                                 uint32_t out8b_stride, uint8_t *out2b_buffer, uint32_t out2b_stride,
                                 uint32_t width, uint32_t height);

static void memcpy_wrap(void *dst, void const *src, size_t n) { memcpy(dst, src, n); }

/* The ME rows of svt_aom_setup_rtcd_internal (aom_dsp_rtcd.c:499,507-514 for x86, :1199-1214 C) */
static void set_rtcd_planes(void) {
    /* the two pointers the pyramid builder reads; the ME rows are left as the caller selected them */
    svt_memcpy    = memcpy_wrap;
    downsample_2d = svt_aom_downsample_2d_c;
}
static void set_rtcd(int simd) {
    set_rtcd_planes();
    svt_sad_loop_kernel                       = svt_sad_loop_kernel_c;
    svt_ext_sad_calculation_8x8_16x16         = svt_ext_sad_calculation_8x8_16x16_c;
    svt_ext_sad_calculation_32x32_64x64       = svt_ext_sad_calculation_32x32_64x64_c;
    svt_ext_all_sad_calculation_8x8_16x16     = svt_ext_all_sad_calculation_8x8_16x16_c;
    svt_ext_eight_sad_calculation_32x32_64x64 = svt_ext_eight_sad_calculation_32x32_64x64_c;
    svt_initialize_buffer_32bits              = svt_initialize_buffer_32bits_c;
    svt_nxm_sad_kernel_sub_sampled            = svt_nxm_sad_kernel_helper_c;
    svt_nxm_sad_kernel                        = svt_nxm_sad_kernel_helper_c;
    if (simd >= 1) {
        svt_sad_loop_kernel                       = svt_sad_loop_kernel_avx2_intrin;
        svt_ext_sad_calculation_8x8_16x16         = svt_ext_sad_calculation_8x8_16x16_avx2_intrin;
        svt_ext_sad_calculation_32x32_64x64       = svt_ext_sad_calculation_32x32_64x64_sse4_intrin;
        svt_ext_all_sad_calculation_8x8_16x16     = svt_ext_all_sad_calculation_8x8_16x16_avx2;
        svt_ext_eight_sad_calculation_32x32_64x64 = svt_ext_eight_sad_calculation_32x32_64x64_avx2;
        svt_initialize_buffer_32bits              = svt_initialize_buffer_32bits_sse2_intrin;
        svt_nxm_sad_kernel_sub_sampled            = svt_nxm_sad_kernel_sub_sampled_helper_avx2;
        svt_nxm_sad_kernel                        = svt_nxm_sad_kernel_helper_avx2;
    }
    if (simd >= 2)
        svt_sad_loop_kernel = svt_sad_loop_kernel_avx512_intrin;
}

/* Which ME table is installed right now: 0 = C, 1 = AVX2, 2 = AVX-512 SAD loop + AVX2 (the x86 production table,
   aom_dsp_rtcd.c:499,507-514), -1 = a mixture. ref_rtcd_table_at_last_b64() is the same probe taken by run_b64 just
   before the most recent svt_aom_motion_estimation_b64 call, so a test can see what a picture-level run executed. */
static int g_table_at_last_b64 = -1;
int ref_rtcd_table(void) {
    const int loop = svt_sad_loop_kernel == svt_sad_loop_kernel_avx512_intrin ? 2
        : svt_sad_loop_kernel == svt_sad_loop_kernel_avx2_intrin              ? 1
        : svt_sad_loop_kernel == svt_sad_loop_kernel_c                        ? 0 : -1;
    const int all8 = svt_ext_all_sad_calculation_8x8_16x16 == svt_ext_all_sad_calculation_8x8_16x16_avx2 ? 1
        : svt_ext_all_sad_calculation_8x8_16x16 == svt_ext_all_sad_calculation_8x8_16x16_c               ? 0 : -1;
    const int e32 = svt_ext_eight_sad_calculation_32x32_64x64 == svt_ext_eight_sad_calculation_32x32_64x64_avx2 ? 1
        : svt_ext_eight_sad_calculation_32x32_64x64 == svt_ext_eight_sad_calculation_32x32_64x64_c              ? 0 : -1;
    const int nxm = svt_nxm_sad_kernel == svt_nxm_sad_kernel_helper_avx2 ? 1
        : svt_nxm_sad_kernel == svt_nxm_sad_kernel_helper_c              ? 0 : -1;
    if (loop < 0 || all8 < 0 || e32 < 0 || nxm < 0) return -1;
    if (all8 != e32 || all8 != nxm) return -1;
    if (loop == 0) return all8 == 0 ? 0 : -1;
    return all8 == 1 ? loop : -1;
}
int ref_rtcd_table_at_last_b64(void) { return g_table_at_last_b64; }

/* ---------------- RTCD kernel pass-throughs ---------------- */
void ref_sad_loop_kernel(int simd, uint8_t *src, uint32_t src_stride, uint8_t *ref, uint32_t ref_stride,
                         uint32_t block_height, uint32_t block_width, uint64_t *best_sad,
                         int16_t *x_search_center, int16_t *y_search_center, uint32_t src_stride_raw,
                         uint8_t skip_search_line, int16_t search_area_width, int16_t search_area_height) {
    set_rtcd(simd);
    svt_sad_loop_kernel(src, src_stride, ref, ref_stride, block_height, block_width, best_sad,
                        x_search_center, y_search_center, src_stride_raw, skip_search_line,
                        search_area_width, search_area_height);
}
void ref_ext_all_sad_calculation_8x8_16x16(int simd, uint8_t *src, uint32_t src_stride, uint8_t *ref,
                                           uint32_t ref_stride, uint32_t mv, uint32_t *p_best_sad_8x8,
                                           uint32_t *p_best_sad_16x16, uint32_t *p_best_mv8x8,
                                           uint32_t *p_best_mv16x16, uint32_t p_eight_sad16x16[16][8],
                                           uint32_t p_eight_sad8x8[64][8], uint8_t sub_sad) {
    set_rtcd(simd);
    svt_ext_all_sad_calculation_8x8_16x16(src, src_stride, ref, ref_stride, mv, p_best_sad_8x8,
                                          p_best_sad_16x16, p_best_mv8x8, p_best_mv16x16,
                                          p_eight_sad16x16, p_eight_sad8x8, sub_sad);
}
void ref_ext_eight_sad_calculation_32x32_64x64(int simd, uint32_t p_sad16x16[16][8],
                                               uint32_t *p_best_sad_32x32, uint32_t *p_best_sad_64x64,
                                               uint32_t *p_best_mv32x32, uint32_t *p_best_mv64x64,
                                               uint32_t mv, uint32_t p_sad32x32[4][8]) {
    set_rtcd(simd);
    svt_ext_eight_sad_calculation_32x32_64x64(p_sad16x16, p_best_sad_32x32, p_best_sad_64x64,
                                              p_best_mv32x32, p_best_mv64x64, mv, p_sad32x32);
}
void ref_ext_sad_calculation_8x8_16x16(int simd, uint8_t *src, uint32_t src_stride, uint8_t *ref,
                                       uint32_t ref_stride, uint32_t *p_best_sad_8x8,
                                       uint32_t *p_best_sad_16x16, uint32_t *p_best_mv8x8,
                                       uint32_t *p_best_mv16x16, uint32_t mv, uint32_t *p_sad16x16,
                                       uint32_t *p_sad8x8, uint8_t sub_sad) {
    set_rtcd(simd);
    svt_ext_sad_calculation_8x8_16x16(src, src_stride, ref, ref_stride, p_best_sad_8x8, p_best_sad_16x16,
                                      p_best_mv8x8, p_best_mv16x16, mv, p_sad16x16, p_sad8x8, sub_sad);
}
void ref_ext_sad_calculation_32x32_64x64(int simd, uint32_t *p_sad16x16, uint32_t *p_best_sad_32x32,
                                         uint32_t *p_best_sad_64x64, uint32_t *p_best_mv32x32,
                                         uint32_t *p_best_mv64x64, uint32_t mv, uint32_t *p_sad32x32) {
    set_rtcd(simd);
    svt_ext_sad_calculation_32x32_64x64(p_sad16x16, p_best_sad_32x32, p_best_sad_64x64, p_best_mv32x32,
                                        p_best_mv64x64, mv, p_sad32x32);
}
uint32_t ref_nxm_sad_kernel(int simd, const uint8_t *src, uint32_t src_stride, const uint8_t *ref,
                            uint32_t ref_stride, uint32_t height, uint32_t width) {
    set_rtcd(simd);
    return svt_nxm_sad_kernel(src, src_stride, ref, ref_stride, height, width);
}
void ref_initialize_buffer_32bits(int simd, uint32_t *pointer, uint32_t count128, uint32_t count32,
                                  uint32_t value) {
    set_rtcd(simd);
    svt_initialize_buffer_32bits(pointer, count128, count32, value);
}
void ref_downsample_2d(uint8_t *input_samples, uint32_t input_stride, uint32_t input_area_width,
                       uint32_t input_area_height, uint8_t *decim_samples, uint32_t decim_stride,
                       uint32_t decim_step) {
    svt_aom_downsample_2d_c(input_samples, input_stride, input_area_width, input_area_height,
                            decim_samples, decim_stride, decim_step);
}
void ref_generate_padding(uint8_t *pic, uint32_t stride, uint32_t width, uint32_t height, uint32_t pad_w,
                          uint32_t pad_h) {
    set_rtcd(0);
    svt_aom_generate_padding(pic, stride, width, height, pad_w, pad_h);
}
void ref_unpack_10bit_msb(const uint16_t *in16, uint32_t in_stride, uint8_t *out8, uint32_t out_stride,
                          uint32_t width, uint32_t height) {
    uint32_t s2  = (width + 3) / 4;
    uint8_t *lsb = (uint8_t *)malloc((size_t)s2 * height);
    svt_unpack_and_2bcompress_c((uint16_t *)in16, in_stride, out8, out_stride, lsb, s2, width, height);
    free(lsb);
}

/* libsvtref_wrap.so only (the same harness linked with -Wl,--wrap and svt-av1-psy_b200/shim/svtme_wrap.c): the shim
   keeps device copies of host pictures keyed by (picture number, plane address); this harness frees and reallocates its
   planes for every call, so the copies are dropped around each one. Absent (NULL) in libsvtref.so. */
extern void svtme_shim_forget_pictures(void) __attribute__((weak));
static void shim_forget(void) {
    if (svtme_shim_forget_pictures) svtme_shim_forget_pictures();
}

/* ---------------- picture planes (PA reference object geometry) ---------------- */
#define FULL_PAD 72 /* enc_handle.c:4485-4497: BLOCK_SIZE_64 + 4 + 4 */
#define QUARTER_PAD 32 /* reference_object.c:348-351: b64_size >> 1 */
#define SIXTEENTH_PAD 16 /* reference_object.c:362-365: b64_size >> 2 */

typedef struct Planes {
    EbPictureBufferDesc full, quarter, sixteenth;
    uint8_t            *mem[3];
    int                 cached; /* owned by the plane cache: planes_free leaves it alone */
} Planes;

static void desc_init(EbPictureBufferDesc *d, uint8_t *buf, int w, int h, int pad) {
    memset(d, 0, sizeof(*d));
    d->buffer_y   = buf;
    d->stride_y   = (uint16_t)(w + 2 * pad);
    d->org_x      = (uint16_t)pad;
    d->org_y      = (uint16_t)pad;
    d->origin_bot_y = (uint16_t)pad;
    d->width      = (uint16_t)w;
    d->height     = (uint16_t)h;
    d->max_width  = (uint16_t)w;
    d->max_height = (uint16_t)h;
    d->bit_depth  = EB_EIGHT_BIT;
}

/* mirrors svt_aom_pad_input_pictures (pic_analysis_process.c:2198-2211) followed by
   svt_aom_downsample_filtering_input_picture (pic_analysis_process.c:2139-2196), calling the
   reference's own svt_aom_generate_padding and downsample_2d */
static void build_planes_into(const uint8_t *luma, int w, int h, uint8_t *full, uint8_t *quarter,
                              uint8_t *sixteenth, Planes *p) {
    set_rtcd_planes();
    desc_init(&p->full, full, w, h, FULL_PAD);
    desc_init(&p->quarter, quarter, w >> 1, h >> 1, QUARTER_PAD);
    desc_init(&p->sixteenth, sixteenth, w >> 2, h >> 2, SIXTEENTH_PAD);
    for (int y = 0; y < h; y++)
        memcpy(full + (size_t)(y + FULL_PAD) * p->full.stride_y + FULL_PAD, luma + (size_t)y * w, w);
    svt_aom_generate_padding(full, p->full.stride_y, w, h, FULL_PAD, FULL_PAD);
    PictureParentControlSet *pcs = (PictureParentControlSet *)calloc(1, sizeof(*pcs));
    pcs->enable_hme_flag         = 1;
    pcs->enable_hme_level0_flag  = 1;
    pcs->enable_hme_level1_flag  = 1;
    svt_aom_downsample_filtering_input_picture(pcs, &p->full, &p->quarter, &p->sixteenth);
    free(pcs);
}

int ref_build_planes(const uint8_t *luma, int width, int height, uint8_t *full, uint8_t *quarter,
                     uint8_t *sixteenth) {
    Planes p;
    build_planes_into(luma, width, height, full, quarter, sixteenth, &p);
    return 0;
}

static void planes_alloc_fresh(Planes *p, const uint8_t *luma, int w, int h) {
    p->mem[0] = (uint8_t *)calloc((size_t)(w + 2 * FULL_PAD) * (h + 2 * FULL_PAD), 1);
    p->mem[1] = (uint8_t *)calloc((size_t)((w >> 1) + 2 * QUARTER_PAD) * ((h >> 1) + 2 * QUARTER_PAD), 1);
    p->mem[2] = (uint8_t *)calloc((size_t)((w >> 2) + 2 * SIXTEENTH_PAD) * ((h >> 2) + 2 * SIXTEENTH_PAD), 1);
    p->cached = 0;
    build_planes_into(luma, w, h, p->mem[0], p->mem[1], p->mem[2], p);
}

/* Plane cache for the timed CPU baseline (bench.py): the encoder builds a picture's pyramid ONCE (Picture Analysis) and
   every picture that references it reads the same PA reference object; without the cache this harness would rebuild the
   pyramids of all references for every picture searched and the baseline would be charged work the reference never does.
   Keyed by the caller's luma pointer. Off by default: the parity tests build their planes per call. */
#define PLANE_CACHE_MAX 1024
static struct {
    const uint8_t *key[PLANE_CACHE_MAX];
    int            w[PLANE_CACHE_MAX], h[PLANE_CACHE_MAX];
    Planes         planes[PLANE_CACHE_MAX];
    int            n, on;
    pthread_mutex_t mu;
} g_cache = {.mu = PTHREAD_MUTEX_INITIALIZER};

static int cache_find(const uint8_t *luma, int w, int h) {
    for (int i = 0; i < g_cache.n; i++)
        if (g_cache.key[i] == luma && g_cache.w[i] == w && g_cache.h[i] == h) return i;
    return -1;
}
void ref_plane_cache_enable(int on) { g_cache.on = on; }
/* build (once) and keep the pyramid of the picture whose interior luma plane is at `luma`; 0 = built, 1 = was present */
int ref_plane_cache_build(const uint8_t *luma, int width, int height) {
    pthread_mutex_lock(&g_cache.mu);
    int i = cache_find(luma, width, height);
    pthread_mutex_unlock(&g_cache.mu);
    if (i >= 0) return 1;
    Planes p;
    planes_alloc_fresh(&p, luma, width, height); /* outside the lock: callers build distinct pictures in parallel */
    p.cached = 1;
    pthread_mutex_lock(&g_cache.mu);
    if (g_cache.n >= PLANE_CACHE_MAX) {
        pthread_mutex_unlock(&g_cache.mu);
        for (int k = 0; k < 3; k++) free(p.mem[k]);
        return -1;
    }
    i = g_cache.n++;
    g_cache.key[i] = luma; g_cache.w[i] = width; g_cache.h[i] = height; g_cache.planes[i] = p;
    pthread_mutex_unlock(&g_cache.mu);
    return 0;
}
void ref_plane_cache_drop(const uint8_t *luma) {
    pthread_mutex_lock(&g_cache.mu);
    for (int i = 0; i < g_cache.n; i++)
        if (g_cache.key[i] == luma || !luma) {
            for (int k = 0; k < 3; k++) free(g_cache.planes[i].mem[k]);
            g_cache.n--;
            g_cache.key[i] = g_cache.key[g_cache.n]; g_cache.w[i] = g_cache.w[g_cache.n]; g_cache.h[i] = g_cache.h[g_cache.n];
            g_cache.planes[i] = g_cache.planes[g_cache.n];
            i--;
        }
    pthread_mutex_unlock(&g_cache.mu);
}

static void planes_alloc(Planes *p, const uint8_t *luma, int w, int h) {
    if (g_cache.on) {
        ref_plane_cache_build(luma, w, h);
        pthread_mutex_lock(&g_cache.mu);
        int i = cache_find(luma, w, h);
        if (i >= 0) *p = g_cache.planes[i];
        pthread_mutex_unlock(&g_cache.mu);
        if (i >= 0) return;
    }
    planes_alloc_fresh(p, luma, w, h);
}
static void planes_free(Planes *p) {
    if (p->cached) return;
    for (int i = 0; i < 3; i++) free(p->mem[i]);
}

/* ---------------- fake sequence / picture control sets ---------------- */
typedef struct Fake {
    SequenceControlSet      *scs;
    PictureParentControlSet *pcs;
    MotionEstimationData    *me_data;
    MeSbResults            **me_results;
    MeSbResults             *me_results_store;
    uint32_t                 n_b64, b64_w, b64_h, n_pus;
} Fake;

static EbInputResolution derive_input_resolution(uint32_t input_size) {
    /* sequence_control_set.c:113-131 */
    if (input_size < INPUT_SIZE_240p_TH) return INPUT_SIZE_240p_RANGE;
    if (input_size < INPUT_SIZE_360p_TH) return INPUT_SIZE_360p_RANGE;
    if (input_size < INPUT_SIZE_480p_TH) return INPUT_SIZE_480p_RANGE;
    if (input_size < INPUT_SIZE_720p_TH) return INPUT_SIZE_720p_RANGE;
    if (input_size < INPUT_SIZE_1080p_TH) return INPUT_SIZE_1080p_RANGE;
    if (input_size < INPUT_SIZE_4K_TH) return INPUT_SIZE_4K_RANGE;
    return INPUT_SIZE_8K_RANGE;
}

static void fake_init(Fake *f, const RefMePicIn *in) {
    memset(f, 0, sizeof(*f));
    SequenceControlSet      *scs = f->scs = (SequenceControlSet *)calloc(1, sizeof(SequenceControlSet));
    PictureParentControlSet *pcs = f->pcs = (PictureParentControlSet *)calloc(1, sizeof(PictureParentControlSet));
    const bool rtc_tune = in->pred_structure == SVT_AV1_PRED_LOW_DELAY_B;

    scs->static_config.qp             = (uint32_t)in->qp;
    scs->static_config.pred_structure = (uint8_t)in->pred_structure;
    scs->static_config.enc_mode       = (int8_t)in->enc_mode;
    scs->static_config.restricted_motion_vector = 0;
    scs->frame_rate                   = in->frame_rate;
    scs->input_resolution             = derive_input_resolution((uint32_t)in->width * (uint32_t)in->height);
    scs->b64_size                     = 64;
    scs->max_input_luma_width         = (uint16_t)in->width;
    scs->max_input_luma_height        = (uint16_t)in->height;
    scs->mrp_ctrls.only_l_bwd         = (uint8_t)in->only_l_bwd;
    scs->mrp_ctrls.safe_limit_nref    = (uint8_t)in->safe_limit_nref;
    scs->mrp_ctrls.safe_limit_zz_th   = (uint32_t)in->safe_limit_zz_th;

    pcs->scs                  = scs;
    pcs->enc_mode             = (EncMode)in->enc_mode;
    pcs->sc_class1            = (uint8_t)in->sc_class1;
    pcs->picture_number       = in->picture_number;
    pcs->aligned_width        = (uint16_t)in->width;
    pcs->aligned_height       = (uint16_t)in->height;
    pcs->hierarchical_levels  = (uint8_t)in->hierarchical_levels;
    pcs->temporal_layer_index = (uint8_t)in->temporal_layer_index;
    pcs->is_ref               = (uint8_t)in->is_ref;
    pcs->slice_type           = in->num_lists == 2 ? B_SLICE : P_SLICE;
    pcs->ref_list0_count_try  = (uint8_t)in->ref_count[0];
    pcs->ref_list1_count_try  = (uint8_t)(in->num_lists == 2 ? in->ref_count[1] : 0);
    pcs->similar_brightness_refs = in->similar_brightness_refs != 0;
    /* enc_mode_config.c:1820-1831 (svt_aom_sig_deriv_multi_processes) */
    pcs->enable_hme_flag        = 1;
    pcs->enable_hme_level0_flag = 1;
    pcs->enable_hme_level1_flag = 1;
    pcs->enable_hme_level2_flag = (pcs->sc_class1 || pcs->enc_mode <= ENC_M6) ? 1 : 0;
    /* enc_mode_config.c:1984-1987 */
    pcs->use_best_me_unipred_cand_only = pcs->enc_mode <= ENC_M4 ? 0 : 1;
    /* pcs.c:1420-1426, resource_coordination_process.c:425 */
    pcs->enable_me_16x16          = 1;
    pcs->enable_me_8x8            = svt_aom_get_enable_me_8x8(pcs->enc_mode, rtc_tune, scs->input_resolution);
    pcs->max_number_of_pus_per_sb = SQUARE_PU_COUNT;
    /* pd_process.c:3517-3519 */
    svt_aom_set_gm_controls(pcs, in->gm_level >= 0 ? (uint8_t)in->gm_level : svt_aom_derive_gm_level(pcs, true));

    f->b64_w = (uint32_t)(in->width + 63) / 64;
    f->b64_h = (uint32_t)(in->height + 63) / 64;
    f->n_b64 = f->b64_w * f->b64_h;
    f->n_pus = pcs->enable_me_8x8 ? SQUARE_PU_COUNT : MAX_SB64_PU_COUNT_NO_8X8;
    pcs->b64_total_count = (uint16_t)f->n_b64;
    /* pcs.c:1552-1606 */
    pcs->b64_geom = (B64Geom *)calloc(f->n_b64, sizeof(B64Geom));
    for (uint32_t i = 0; i < f->n_b64; i++) {
        B64Geom *g          = &pcs->b64_geom[i];
        g->horizontal_index = (uint8_t)(i % f->b64_w);
        g->vertical_index   = (uint8_t)(i / f->b64_w);
        g->org_x            = (uint16_t)(g->horizontal_index * 64);
        g->org_y            = (uint16_t)(g->vertical_index * 64);
        g->width            = (uint8_t)MIN(64, in->width - g->org_x);
        g->height           = (uint8_t)MIN(64, in->height - g->org_y);
        g->tile_end_x       = (uint32_t)in->width;
        g->tile_end_y       = (uint32_t)in->height;
    }
    /* ME result storage: pcs.c:91-117, pd_process.c:3495-3511 */
    MotionEstimationData *md = f->me_data = (MotionEstimationData *)calloc(1, sizeof(*md));
    uint8_t max_refs, max_cand;
    /* svt_aom_get_max_allocated_me_refs, pcs.c:91-96 */
    max_refs = (uint8_t)(in->max_l0 + in->max_l1);
    max_cand = (uint8_t)(in->max_l0 + in->max_l1 + in->max_l0 * in->max_l1 + (in->max_l0 - 1) + (in->max_l1 == 3 ? 1 : 0));
    md->max_cand       = max_cand;
    md->max_refs       = max_refs;
    md->max_l0         = (uint8_t)in->max_l0;
    f->me_results      = (MeSbResults **)calloc(f->n_b64, sizeof(MeSbResults *));
    f->me_results_store = (MeSbResults *)calloc(f->n_b64, sizeof(MeSbResults));
    for (uint32_t i = 0; i < f->n_b64; i++) {
        MeSbResults *r              = &f->me_results_store[i];
        r->me_mv_array              = (MvCandidate *)calloc((size_t)f->n_pus * max_refs, sizeof(MvCandidate));
        r->me_candidate_array       = (MeCandidate *)calloc((size_t)f->n_pus * max_cand, sizeof(MeCandidate));
        r->total_me_candidate_index = (uint8_t *)calloc(f->n_pus, 1);
        f->me_results[i]            = r;
    }
    md->me_results  = f->me_results;
    pcs->pa_me_data = md;
    pcs->rc_me_distortion            = (uint32_t *)calloc(f->n_b64, 4);
    pcs->me_8x8_cost_variance        = (uint32_t *)calloc(f->n_b64, 4);
    pcs->me_64x64_distortion         = (uint32_t *)calloc(f->n_b64, 4);
    pcs->me_32x32_distortion         = (uint32_t *)calloc(f->n_b64, 4);
    pcs->me_16x16_distortion         = (uint32_t *)calloc(f->n_b64, 4);
    pcs->me_8x8_distortion           = (uint32_t *)calloc(f->n_b64, 4);
    pcs->stationary_block_present_sb = (uint8_t *)calloc(f->n_b64, 1);
    pcs->rc_me_allow_gm              = (uint8_t *)calloc(f->n_b64, 1);
}

static void fake_free(Fake *f) {
    PictureParentControlSet *pcs = f->pcs;
    for (uint32_t i = 0; i < f->n_b64; i++) {
        free(f->me_results_store[i].me_mv_array);
        free(f->me_results_store[i].me_candidate_array);
        free(f->me_results_store[i].total_me_candidate_index);
    }
    free(f->me_results_store);
    free(f->me_results);
    free(f->me_data);
    free(pcs->b64_geom);
    free(pcs->rc_me_distortion);
    free(pcs->me_8x8_cost_variance);
    free(pcs->me_64x64_distortion);
    free(pcs->me_32x32_distortion);
    free(pcs->me_16x16_distortion);
    free(pcs->me_8x8_distortion);
    free(pcs->stationary_block_present_sb);
    free(pcs->rc_me_allow_gm);
    free(pcs);
    free(f->scs);
}

/* ---------------- export of the derived controls ---------------- */
static SvtMeSearchArea sa_of(SearchArea a) {
    SvtMeSearchArea r = {a.width, a.height};
    return r;
}
static SvtMeSearchAreaMinMax samm_of(SearchAreaMinMax a) {
    SvtMeSearchAreaMinMax r = {sa_of(a.sa_min), sa_of(a.sa_max)};
    return r;
}

static void export_params(const RefMePicIn *in, const Fake *f, const MeContext *m, SvtMePicParams *o) {
    memset(o, 0, sizeof(*o));
    SvtMeCtrls *c = &o->ctrls;
    c->enable_hme_flag        = m->enable_hme_flag;
    c->enable_hme_level0_flag = m->enable_hme_level0_flag;
    c->enable_hme_level1_flag = m->enable_hme_level1_flag;
    c->enable_hme_level2_flag = m->enable_hme_level2_flag;
    c->hme_search_method      = m->hme_search_method;
    c->me_search_method       = m->me_search_method;
    c->num_hme_sa_w           = m->num_hme_sa_w;
    c->num_hme_sa_h           = m->num_hme_sa_h;
    c->hme_l0_sa              = samm_of(m->hme_l0_sa);
    c->hme_l1_sa              = sa_of(m->hme_l1_sa);
    c->hme_l2_sa              = sa_of(m->hme_l2_sa);
    c->me_sa                  = samm_of(m->me_sa);
    c->prehme_enable           = m->prehme_ctrl.enable;
    c->prehme_skip_search_line = m->prehme_ctrl.enable ? m->prehme_ctrl.skip_search_line : 0;
    c->prehme_l1_early_exit    = m->prehme_ctrl.enable ? m->prehme_ctrl.l1_early_exit : 0;
    for (int i = 0; i < 2; i++)
        if (m->prehme_ctrl.enable) c->prehme_sa_cfg[i] = samm_of(m->prehme_ctrl.prehme_sa_cfg[i]);
    c->enable_me_hme_ref_pruning = m->me_hme_prune_ctrls.enable_me_hme_ref_pruning;
    c->prune_ref_if_hme_sad_dev_bigger_than_th = m->me_hme_prune_ctrls.prune_ref_if_hme_sad_dev_bigger_than_th;
    c->prune_ref_if_me_sad_dev_bigger_than_th  = m->me_hme_prune_ctrls.prune_ref_if_me_sad_dev_bigger_than_th;
    c->zz_sad_th    = m->me_hme_prune_ctrls.zz_sad_th;
    c->zz_sad_pct   = m->me_hme_prune_ctrls.zz_sad_pct;
    c->phme_sad_th  = m->me_hme_prune_ctrls.phme_sad_th;
    c->phme_sad_pct = m->me_hme_prune_ctrls.phme_sad_pct;
    const MeSrCtrls *s = &m->me_sr_adjustment_ctrls;
    c->enable_me_sr_adjustment = s->enable_me_sr_adjustment;
    if (s->enable_me_sr_adjustment) {
        c->reduce_me_sr_based_on_mv_length_th   = s->reduce_me_sr_based_on_mv_length_th;
        c->stationary_hme_sad_abs_th            = s->stationary_hme_sad_abs_th;
        c->stationary_me_sr_divisor             = s->stationary_me_sr_divisor;
        c->reduce_me_sr_based_on_hme_sad_abs_th = s->reduce_me_sr_based_on_hme_sad_abs_th;
        c->me_sr_divisor_for_low_hme_sad        = s->me_sr_divisor_for_low_hme_sad;
        c->distance_based_hme_resizing          = s->distance_based_hme_resizing;
    }
    c->me_8x8_var_enabled = m->me_8x8_var_ctrls.enabled;
    if (c->me_8x8_var_enabled) {
        c->me_sr_div4_th  = m->me_8x8_var_ctrls.me_sr_div4_th;
        c->me_sr_div2_th  = m->me_8x8_var_ctrls.me_sr_div2_th;
        c->me_sr_mult2_th = m->me_8x8_var_ctrls.me_sr_mult2_th;
    }
    c->mv_sa_adj_enabled = m->mv_based_sa_adj.enabled;
    if (c->mv_sa_adj_enabled) {
        c->mv_sa_adj_nearest_ref_only = m->mv_based_sa_adj.nearest_ref_only;
        c->mv_sa_adj_mv_size_th       = m->mv_based_sa_adj.mv_size_th;
        c->mv_sa_adj_sa_multiplier    = m->mv_based_sa_adj.sa_multiplier;
    }
    c->reduce_hme_l0_sr_th_min     = m->reduce_hme_l0_sr_th_min;
    c->reduce_hme_l0_sr_th_max     = m->reduce_hme_l0_sr_th_max;
    c->me_early_exit_th            = m->me_early_exit_th;
    c->me_safe_limit_zz_th         = m->me_safe_limit_zz_th;
    c->prev_me_stage_based_exit_th = m->prev_me_stage_based_exit_th;
    c->prune_me_candidates_th      = m->prune_me_candidates_th;
    c->use_best_unipred_cand_only  = m->use_best_unipred_cand_only;

    const PictureParentControlSet *pcs = f->pcs;
    o->picture_number              = pcs->picture_number;
    o->num_of_list_to_search       = (uint8_t)in->num_lists;
    o->num_of_ref_pic_to_search[0] = pcs->ref_list0_count_try;
    o->num_of_ref_pic_to_search[1] = pcs->ref_list1_count_try;
    o->temporal_layer_index        = pcs->temporal_layer_index;
    o->is_ref                      = pcs->is_ref;
    o->hierarchical_levels         = pcs->hierarchical_levels;
    o->similar_brightness_refs     = pcs->similar_brightness_refs;
    o->enable_me_8x8               = pcs->enable_me_8x8;
    o->enable_me_16x16             = pcs->enable_me_16x16;
    o->max_cand                    = f->me_data->max_cand;
    o->max_refs                    = f->me_data->max_refs;
    o->max_l0                      = f->me_data->max_l0;
    o->only_l_bwd                  = pcs->scs->mrp_ctrls.only_l_bwd;
    o->input_resolution            = (uint8_t)pcs->scs->input_resolution;
    o->gm_enabled                  = pcs->gm_ctrls.enabled;
    o->gm_use_distance_based_active_th = pcs->gm_ctrls.use_distance_based_active_th;
    for (int l = 0; l < 2; l++)
        for (int r = 0; r < 4; r++) {
            o->ref_pic_id[l][r]         = -1;
            o->ref_picture_number[l][r] = in->ref_picture_number[l][r];
        }
}

int ref_derive_params(const RefMePicIn *in, SvtMePicParams *out) {
    Fake f;
    fake_init(&f, in);
    MeContext *m = (MeContext *)calloc(1, sizeof(MeContext));
    m->me_type   = ME_OPEN_LOOP;
    svt_aom_sig_deriv_me(f.scs, f.pcs, m); /* enc_mode_config.c:694 — the reference's own */
    export_params(in, &f, m, out);
    free(m);
    fake_free(&f);
    return 0;
}

/* The same derivation, exported through the PRODUCT's reference-side binding (svt-av1-psy_b200/shim/svtme_wrap.c,
   svtme_shim_flatten) instead of export_params above: libsvtref_wrap.so only. Lets the CPU tests check the shim's
   MeContext/PCS -> SvtMePicParams flattening against the golden parameter fixture without a GPU. */
extern void svtme_shim_flatten(const PictureParentControlSet *pcs, const MeContext *m, SvtMePicParams *o) __attribute__((weak));
int ref_derive_params_shim(const RefMePicIn *in, SvtMePicParams *out) {
    if (!svtme_shim_flatten) return -2;
    Fake f;
    fake_init(&f, in);
    PictureParentControlSet *pcs = f.pcs;
    MeContext *m = (MeContext *)calloc(1, sizeof(MeContext));
    m->me_type   = ME_OPEN_LOOP;
    svt_aom_sig_deriv_me(f.scs, pcs, m);
    /* what me_process.c:219-262 sets before every b64 call */
    m->num_of_list_to_search       = (pcs->slice_type == P_SLICE) ? 1 : 2;
    m->num_of_ref_pic_to_search[0] = pcs->ref_list0_count_try;
    if (pcs->slice_type == B_SLICE) m->num_of_ref_pic_to_search[1] = pcs->ref_list1_count_try;
    m->temporal_layer_index = pcs->temporal_layer_index;
    m->is_ref               = pcs->is_ref;
    for (int i = 0; i < m->num_of_list_to_search; i++)
        for (int j = 0; j < m->num_of_ref_pic_to_search[i]; j++)
            m->me_ds_ref_array[i][j].picture_number = in->ref_picture_number[i][j];
    svtme_shim_flatten(pcs, m, out);
    free(m);
    fake_free(&f);
    return 0;
}

static double now_s(void) {
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec;
}

/* ---------------- whole-picture ME through the reference's own b64 function ---------------- */
static void dump_debug(RefMeDebug *d, const MeContext *me);
typedef struct PicRun {
    const RefMePicIn *in;
    Fake              f;
    Planes            src, refs[2][4];
    SvtMePicResults  *res;
    RefMeDebug       *dbg;
    /* the PA reference objects Picture Decision would have attached to the PCS (pd_process.c:4176): read by the GPU
       shim's svtme_shim_submit_pictures only, the reference's b64 function gets its planes through the MeContext */
    EbObjectWrapper     src_w, ref_w[2][4];
    EbPaReferenceObject src_o, ref_o[2][4];
} PicRun;

/* one b64 the way svt_aom_motion_estimation_kernel does it (me_process.c:174-271), then the dump of what the call left
   in the MeContext and in the PCS; returns the seconds spent inside svt_aom_motion_estimation_b64 */
static double run_b64(PicRun *run, MeContext *me, uint32_t x_b64, uint32_t y_b64) {
    const RefMePicIn        *in  = run->in;
    Fake                    *f   = &run->f;
    PictureParentControlSet *pcs = f->pcs;
    SvtMePicResults         *res = run->res;
    const uint32_t n_pus = f->n_pus;
    const uint8_t  max_refs = f->me_data->max_refs, max_cand = f->me_data->max_cand;

    uint32_t b64_index    = (uint16_t)(x_b64 + y_b64 * f->b64_w);
    uint32_t b64_origin_x = x_b64 * 64, b64_origin_y = y_b64 * 64;
    EbPictureBufferDesc *ip = &run->src.full, *qp = &run->src.quarter, *sp = &run->src.sixteenth;
    uint32_t buffer_index = (ip->org_y + b64_origin_y) * ip->stride_y + ip->org_x + b64_origin_x;
    me->b64_src_ptr    = &ip->buffer_y[buffer_index];
    me->b64_src_stride = ip->stride_y;
    if (me->enable_hme_level1_flag) {
        buffer_index = (qp->org_y + (b64_origin_y >> 1)) * qp->stride_y + qp->org_x + (b64_origin_x >> 1);
        me->quarter_b64_buffer        = &qp->buffer_y[buffer_index];
        me->quarter_b64_buffer_stride = qp->stride_y;
    }
    if (me->enable_hme_level0_flag) {
        buffer_index = (sp->org_y + (b64_origin_y >> 2)) * sp->stride_y + sp->org_x + (b64_origin_x >> 2);
        me->sixteenth_b64_buffer        = &sp->buffer_y[buffer_index];
        me->sixteenth_b64_buffer_stride = sp->stride_y;
    }
    me->me_type                     = ME_OPEN_LOOP;
    me->num_of_list_to_search       = (pcs->slice_type == P_SLICE) ? 1 : 2;
    me->num_of_ref_pic_to_search[0] = pcs->ref_list0_count_try;
    if (pcs->slice_type == B_SLICE) me->num_of_ref_pic_to_search[1] = pcs->ref_list1_count_try;
    me->temporal_layer_index = pcs->temporal_layer_index;
    me->is_ref               = pcs->is_ref;
    for (int i = 0; i < me->num_of_list_to_search; i++)
        for (int j = 0; j < me->num_of_ref_pic_to_search[i]; j++) {
            me->me_ds_ref_array[i][j].picture_ptr           = &run->refs[i][j].full;
            me->me_ds_ref_array[i][j].quarter_picture_ptr   = &run->refs[i][j].quarter;
            me->me_ds_ref_array[i][j].sixteenth_picture_ptr = &run->refs[i][j].sixteenth;
            me->me_ds_ref_array[i][j].picture_number        = in->ref_picture_number[i][j];
        }

    g_table_at_last_b64 = ref_rtcd_table();
    double t0 = now_s();
    svt_aom_motion_estimation_b64(pcs, b64_index, b64_origin_x, b64_origin_y, me, ip);
    const double t_me = now_s() - t0;

    /* ---- dump ---- */
    if (res) {
        for (int l = 0; l < 2; l++)
            for (int r = 0; r < 4; r++) {
                size_t o = ((size_t)b64_index * 2 + l) * 4 + r;
                int searched = l < me->num_of_list_to_search && r < me->num_of_ref_pic_to_search[l];
                /* entries of refs with do_ref == 0 are stale: masked */
                int valid = searched && me->search_results[l][r].do_ref;
                if (res->p_sb_best_sad)
                    for (int p = 0; p < 85; p++)
                        res->p_sb_best_sad[o * 85 + p] = valid ? me->p_sb_best_sad[l][r][p] : 0;
                if (res->p_sb_best_mv)
                    for (int p = 0; p < 85; p++)
                        res->p_sb_best_mv[o * 85 + p] = valid ? me->p_sb_best_mv[l][r][p] : 0;
                if (res->search_results) {
                    SvtMeSearchResult *sr = &res->search_results[o];
                    memset(sr, 0, sizeof(*sr));
                    if (searched) {
                        sr->hme_sc_x = me->search_results[l][r].hme_sc_x;
                        sr->hme_sc_y = me->search_results[l][r].hme_sc_y;
                        sr->do_ref   = me->search_results[l][r].do_ref;
                        sr->hme_sad  = me->search_results[l][r].hme_sad;
                    }
                }
            }
        MeSbResults *r = f->me_results[b64_index];
        if (res->total_me_candidate_index)
            memcpy(res->total_me_candidate_index + (size_t)b64_index * n_pus, r->total_me_candidate_index, n_pus);
        if (res->me_mv_array)
            memcpy(res->me_mv_array + (size_t)b64_index * n_pus * max_refs, r->me_mv_array,
                   (size_t)n_pus * max_refs * 4);
        if (res->me_candidate_array)
            memcpy(res->me_candidate_array + (size_t)b64_index * n_pus * max_cand, r->me_candidate_array,
                   (size_t)n_pus * max_cand);
    }
    if (run->dbg) dump_debug(&run->dbg[b64_index], me);
    return t_me;
}

static int pic_run_init(PicRun *run, const RefMePicIn *in, SvtMePicResults *res, RefMeDebug *dbg) {
    if (in->width % 8 || in->height % 8 || in->num_lists < 1 || in->num_lists > 2) return -1;
    set_rtcd(in->simd);
    run->in = in; run->res = res; run->dbg = dbg;
    fake_init(&run->f, in);
    planes_alloc(&run->src, in->src_luma, in->width, in->height);
    for (int l = 0; l < in->num_lists; l++)
        for (int r = 0; r < in->ref_count[l]; r++) planes_alloc(&run->refs[l][r], in->ref_luma[l][r], in->width, in->height);
    memset(&run->src_w, 0, sizeof(run->src_w)); memset(run->ref_w, 0, sizeof(run->ref_w));
    memset(&run->src_o, 0, sizeof(run->src_o)); memset(run->ref_o, 0, sizeof(run->ref_o));
    run->src_o.input_padded_pic = &run->src.full;
    run->src_o.quarter_downsampled_picture_ptr = &run->src.quarter;
    run->src_o.sixteenth_downsampled_picture_ptr = &run->src.sixteenth;
    run->src_o.picture_number = in->picture_number;
    run->src_w.object_ptr = &run->src_o;
    run->f.pcs->pa_ref_pic_wrapper = &run->src_w;
    for (int l = 0; l < in->num_lists; l++)
        for (int r = 0; r < in->ref_count[l]; r++) {
            run->ref_o[l][r].input_padded_pic = &run->refs[l][r].full;
            run->ref_o[l][r].quarter_downsampled_picture_ptr = &run->refs[l][r].quarter;
            run->ref_o[l][r].sixteenth_downsampled_picture_ptr = &run->refs[l][r].sixteenth;
            run->ref_o[l][r].picture_number = in->ref_picture_number[l][r];
            run->ref_w[l][r].object_ptr = &run->ref_o[l][r];
            run->f.pcs->ref_pa_pic_ptr_array[l][r] = &run->ref_w[l][r];
        }
    return 0;
}

static void pic_run_finish(PicRun *run) {
    const RefMePicIn        *in  = run->in;
    PictureParentControlSet *pcs = run->f.pcs;
    SvtMePicResults         *res = run->res;
    if (res) {
#define CP(field, bytes) if (res->field) memcpy(res->field, pcs->field, (size_t)run->f.n_b64 * (bytes))
        CP(me_64x64_distortion, 4);
        CP(me_32x32_distortion, 4);
        CP(me_16x16_distortion, 4);
        CP(me_8x8_distortion, 4);
        CP(rc_me_distortion, 4);
        CP(me_8x8_cost_variance, 4);
        CP(stationary_block_present_sb, 1);
        CP(rc_me_allow_gm, 1);
#undef CP
    }
    planes_free(&run->src);
    for (int l = 0; l < in->num_lists; l++)
        for (int r = 0; r < in->ref_count[l]; r++) planes_free(&run->refs[l][r]);
    fake_free(&run->f);
}

int ref_me_picture(const RefMePicIn *in, SvtMePicResults *res, RefMeDebug *dbg, double *seconds) {
    PicRun run;
    shim_forget();
    if (pic_run_init(&run, in, res, dbg)) return -1;
    MeContext *me = (MeContext *)calloc(1, sizeof(MeContext));
    me->me_type   = ME_OPEN_LOOP;
    svt_aom_sig_deriv_me(run.f.scs, run.f.pcs, me); /* me_process.c:121 */
    double t_me = 0.0;
    for (uint32_t y_b64 = 0; y_b64 < run.f.b64_h; y_b64++)
        for (uint32_t x_b64 = 0; x_b64 < run.f.b64_w; x_b64++) t_me += run_b64(&run, me, x_b64, y_b64);
    if (seconds) *seconds = t_me;
    free(me);
    shim_forget();
    pic_run_finish(&run);
    return 0;
}

/* The same picture with the b64s dealt out to n_threads threads, each with its own MeContext, as the encoder's ME
   threads work on the segments of one picture (me_process.c:86, :148-172; up to 20-25 threads, enc_handle.c:1066). */
typedef struct MtArg {
    PicRun  *run;
    int      tid, n_threads;
    double   seconds;
} MtArg;
static void *mt_worker(void *p) {
    MtArg     *a   = (MtArg *)p;
    PicRun    *run = a->run;
    MeContext *me  = (MeContext *)calloc(1, sizeof(MeContext));
    me->me_type    = ME_OPEN_LOOP;
    svt_aom_sig_deriv_me(run->f.scs, run->f.pcs, me);
    for (uint32_t i = (uint32_t)a->tid; i < run->f.n_b64; i += (uint32_t)a->n_threads)
        a->seconds += run_b64(run, me, i % run->f.b64_w, i / run->f.b64_w);
    free(me);
    return NULL;
}
int ref_me_picture_mt(const RefMePicIn *in, SvtMePicResults *res, int n_threads, double *seconds) {
    if (n_threads < 1 || n_threads > 64) return -1;
    PicRun run;
    shim_forget();
    if (pic_run_init(&run, in, res, NULL)) return -1;
    pthread_t th[64];
    MtArg     arg[64];
    for (int t = 0; t < n_threads; t++) {
        arg[t] = (MtArg){&run, t, n_threads, 0.0};
        pthread_create(&th[t], NULL, mt_worker, &arg[t]);
    }
    double tot = 0.0;
    for (int t = 0; t < n_threads; t++) {
        pthread_join(th[t], NULL);
        tot += arg[t].seconds;
    }
    if (seconds) *seconds = tot;
    shim_forget();
    pic_run_finish(&run);
    return 0;
}

/* n pictures released together, as Picture Decision releases a mini-GOP (pd_process.c:4373-4400): with the GPU shim they
   are queued as one batch (svtme_shim_submit_pictures) before the ME threads start; the threads then take (picture,
   b64) pairs in the order the ME FIFO would hand them out. Without the shim (libsvtref.so) the submit step is absent. */
extern void svtme_shim_submit_pictures(PictureParentControlSet **pcs_list, uint32_t n) __attribute__((weak));
typedef struct GopArg {
    PicRun *runs;
    int     n_pics, tid, n_threads;
} GopArg;
static void *gop_worker(void *p) {
    GopArg *a = (GopArg *)p;
    for (int k = 0; k < a->n_pics; k++) {
        PicRun    *run = &a->runs[k];
        MeContext *me  = (MeContext *)calloc(1, sizeof(MeContext));
        me->me_type    = ME_OPEN_LOOP;
        svt_aom_sig_deriv_me(run->f.scs, run->f.pcs, me);
        for (uint32_t i = (uint32_t)a->tid; i < run->f.n_b64; i += (uint32_t)a->n_threads)
            run_b64(run, me, i % run->f.b64_w, i / run->f.b64_w);
        free(me);
    }
    return NULL;
}
int ref_me_minigop(const RefMePicIn *ins, int n_pics, SvtMePicResults *res, int n_threads, int submit) {
    if (n_pics < 1 || n_pics > 32 || n_threads < 1 || n_threads > 64) return -1;
    PicRun *runs = (PicRun *)calloc((size_t)n_pics, sizeof(PicRun));
    shim_forget();
    for (int k = 0; k < n_pics; k++)
        if (pic_run_init(&runs[k], &ins[k], &res[k], NULL)) return -1;
    if (submit && svtme_shim_submit_pictures) {
        PictureParentControlSet *list[32];
        for (int k = 0; k < n_pics; k++) list[k] = runs[k].f.pcs;
        svtme_shim_submit_pictures(list, (uint32_t)n_pics);
    }
    pthread_t th[64];
    GopArg    arg[64];
    for (int t = 0; t < n_threads; t++) {
        arg[t] = (GopArg){runs, n_pics, t, n_threads};
        pthread_create(&th[t], NULL, gop_worker, &arg[t]);
    }
    for (int t = 0; t < n_threads; t++) pthread_join(th[t], NULL);
    shim_forget();
    for (int k = 0; k < n_pics; k++) pic_run_finish(&runs[k]);
    free(runs);
    return 0;
}


/* ---------------- dynamic-GOP detector HME (pd_process.c:392-589) ---------------- */
void dg_detector_hme_level0(struct PictureParentControlSet *ppcs, uint32_t seg_idx); /* pd_process.c:490 */
/* the detector takes a mutex per b64 and posts a semaphore at the end: single-threaded here */
EbErrorType svt_block_on_mutex(EbHandle h) { (void)h; return EB_ErrorNone; }
EbErrorType svt_release_mutex(EbHandle h) { (void)h; return EB_ErrorNone; }
EbErrorType svt_post_semaphore(EbHandle h) { (void)h; return EB_ErrorNone; }

int ref_dg_detector(const uint8_t *src_luma, const uint8_t *ref_luma, int width, int height, SvtMeDgMetrics *out) {
    Planes ps, pr;
    planes_alloc(&ps, src_luma, width, height);
    planes_alloc(&pr, ref_luma, width, height);
    set_rtcd(0);
    SequenceControlSet      *scs  = (SequenceControlSet *)calloc(1, sizeof(*scs));
    PictureParentControlSet *src  = (PictureParentControlSet *)calloc(1, sizeof(*src));
    PictureParentControlSet *refp = (PictureParentControlSet *)calloc(1, sizeof(*refp));
    EbObjectWrapper          ws, wr;
    EbPaReferenceObject      os, orf;
    DGDetectorSeg            dg;
    memset(&ws, 0, sizeof(ws)); memset(&wr, 0, sizeof(wr)); memset(&os, 0, sizeof(os)); memset(&orf, 0, sizeof(orf));
    memset(&dg, 0, sizeof(dg));
    scs->b64_size = 64;
    os.sixteenth_downsampled_picture_ptr  = &ps.sixteenth;
    orf.sixteenth_downsampled_picture_ptr = &pr.sixteenth;
    os.input_padded_pic = &ps.full; /* read by the GPU shim only: the reference's detector needs just the 1/16 pictures */
    orf.input_padded_pic = &pr.full;
    src->picture_number = 1; refp->picture_number = 0;
    shim_forget();
    ws.object_ptr = &os; wr.object_ptr = &orf;
    src->pa_ref_pic_wrapper = &ws; refp->pa_ref_pic_wrapper = &wr;
    src->scs = scs; src->aligned_width = (uint16_t)width; src->aligned_height = (uint16_t)height;
    src->input_resolution = derive_input_resolution((uint32_t)width * (uint32_t)height);
    src->me_segments_column_count = 1; src->me_segments_row_count = 1; src->me_segments_total_count = 1;
    src->dg_detector = &dg;
    dg.ref_pic = refp;
    dg_detector_hme_level0(src, 0);
    out->tot_dist = dg.metrics.tot_dist; out->tot_cplx = dg.metrics.tot_cplx; out->tot_active = dg.metrics.tot_active;
    out->sum_in_vectors = dg.metrics.sum_in_vectors; out->input_resolution = (uint32_t)src->input_resolution;
    shim_forget();
    free(scs); free(src); free(refp);
    planes_free(&ps); planes_free(&pr);
    return 0;
}


/* ---------------- temporal-filter ME (ME_MCTF) ---------------- */
void svt_aom_sig_deriv_me_tf(PictureParentControlSet *pcs, MeContext *me_ctx); /* enc_mode_config.c:849 */

static void dump_debug(RefMeDebug *d, const MeContext *me) {
    memcpy(d->x_hme_l0, me->x_hme_level0_search_center, sizeof(d->x_hme_l0));
    memcpy(d->y_hme_l0, me->y_hme_level0_search_center, sizeof(d->y_hme_l0));
    memcpy(d->x_hme_l1, me->x_hme_level1_search_center, sizeof(d->x_hme_l1));
    memcpy(d->y_hme_l1, me->y_hme_level1_search_center, sizeof(d->y_hme_l1));
    memcpy(d->x_hme_l2, me->x_hme_level2_search_center, sizeof(d->x_hme_l2));
    memcpy(d->y_hme_l2, me->y_hme_level2_search_center, sizeof(d->y_hme_l2));
    memcpy(d->hme_l0_sad, me->hme_level0_sad, sizeof(d->hme_l0_sad));
    memcpy(d->hme_l1_sad, me->hme_level1_sad, sizeof(d->hme_l1_sad));
    memcpy(d->hme_l2_sad, me->hme_level2_sad, sizeof(d->hme_l2_sad));
    for (int l = 0; l < 2; l++)
        for (int r = 0; r < 4; r++) {
            for (int s = 0; s < 2; s++) {
                d->prehme_mv_x[l][r][s] = me->prehme_data[l][r][s].best_mv.as_mv.col;
                d->prehme_mv_y[l][r][s] = me->prehme_data[l][r][s].best_mv.as_mv.row;
                d->prehme_sad[l][r][s]  = me->prehme_data[l][r][s].sad;
            }
            d->zz_sad[l][r]               = me->zz_sad[l][r];
            d->reduce_me_sr_divisor[l][r] = me->reduce_me_sr_divisor[l][r];
        }
}

int ref_tf_me_picture(const RefTfIn *in, SvtMePicParams *params_out, SvtMePicResults *res, RefMeDebug *dbg) {
    if (in->width % 8 || in->height % 8 || in->tf_hme_me_level < 0 || in->tf_hme_me_level > 4) return -1;
    set_rtcd(in->simd);
    RefMePicIn pin;
    memset(&pin, 0, sizeof(pin));
    pin.enc_mode = in->enc_mode; pin.width = in->width; pin.height = in->height; pin.qp = in->qp;
    pin.pred_structure = 2; pin.hierarchical_levels = 5;
    pin.temporal_layer_index = in->temporal_layer_index; pin.is_ref = in->is_ref;
    pin.frame_rate = 30 << 16; pin.gm_level = -1;
    pin.max_l0 = 1; pin.max_l1 = 1; pin.picture_number = in->centre_picture_number;
    pin.num_lists = 1; pin.ref_count[0] = 1; pin.ref_picture_number[0][0] = in->ref_picture_number;
    Fake f;
    fake_init(&f, &pin);
    PictureParentControlSet *pcs = f.pcs;
    pcs->tf_ctrls.hme_me_level = (uint8_t)in->tf_hme_me_level;
    pcs->tf_ctrls.qp_opt       = (uint8_t)in->tf_qp_opt;
    pcs->tf_ctrls.me_exit_th   = in->tf_me_exit_th;
    /* enc_mode_config.c:1832-1856 (svt_aom_sig_deriv_multi_processes) */
    pcs->tf_enable_hme_flag        = 1;
    pcs->tf_enable_hme_level0_flag = 1;
    pcs->tf_enable_hme_level1_flag = in->tf_hme_me_level <= 2;
    pcs->tf_enable_hme_level2_flag = in->tf_hme_me_level == 0;

    Planes src, ref;
    shim_forget();
    planes_alloc(&src, in->centre_luma, in->width, in->height);
    planes_alloc(&ref, in->ref_luma, in->width, in->height);
    MeContext *me = (MeContext *)calloc(1, sizeof(MeContext));
    me->me_type   = ME_MCTF;
    svt_aom_sig_deriv_me_tf(pcs, me); /* me_process.c:124 */

    for (uint32_t y_b64 = 0; y_b64 < f.b64_h; y_b64++) {
        for (uint32_t x_b64 = 0; x_b64 < f.b64_w; x_b64++) {
            const uint32_t b64_index = x_b64 + y_b64 * f.b64_w, ox = x_b64 * 64, oy = y_b64 * 64;
            EbPictureBufferDesc *ip = &src.full, *qp = &src.quarter, *sp = &src.sixteenth;
            /* create_me_context_and_picture_control, temporal_filtering.c:288-343 */
            me->me_type           = ME_MCTF;
            me->hme_search_method = FULL_SAD_SEARCH;
            me->b64_src_ptr       = &ip->buffer_y[(ip->org_y + oy) * ip->stride_y + ip->org_x + ox];
            me->b64_src_stride    = ip->stride_y;
            me->quarter_b64_buffer          = &qp->buffer_y[(qp->org_y + (oy >> 1)) * qp->stride_y + qp->org_x + (ox >> 1)];
            me->quarter_b64_buffer_stride   = qp->stride_y;
            me->sixteenth_b64_buffer        = &sp->buffer_y[(sp->org_y + (oy >> 2)) * sp->stride_y + sp->org_x + (ox >> 2)];
            me->sixteenth_b64_buffer_stride = sp->stride_y;
            /* temporal_filtering.c:3112-3140 */
            me->num_of_list_to_search       = 1;
            me->num_of_ref_pic_to_search[0] = 1;
            me->num_of_ref_pic_to_search[1] = 0;
            me->temporal_layer_index        = pcs->temporal_layer_index;
            me->is_ref                      = pcs->is_ref;
            me->me_ds_ref_array[0][0].picture_ptr           = &ref.full;
            me->me_ds_ref_array[0][0].quarter_picture_ptr   = &ref.quarter;
            me->me_ds_ref_array[0][0].sixteenth_picture_ptr = &ref.sixteenth;
            me->me_ds_ref_array[0][0].picture_number        = in->ref_picture_number;
            me->tf_me_exit_th             = (uint16_t)pcs->tf_ctrls.me_exit_th;
            me->tf_use_pred_64x64_only_th = 0;
            me->hme_l0_sa = me->hme_l0_sa_default_tf; /* set_hme_search_params_mctf(ctx, 0), :2752-2760 */
            if (params_out && b64_index == 0) {
                export_params(&pin, &f, me, params_out);
                params_out->ctrls.me_type       = SVTME_ME_MCTF;
                params_out->ctrls.tf_me_exit_th = me->tf_me_exit_th;
                /* PCS-level fields of the open-loop stage that MCTF never reads (candidate lists, GM): the values
                   svtme_sig_deriv_me_tf writes */
                params_out->enable_me_8x8 = params_out->enable_me_16x16 = 1;
                params_out->max_cand = 3; params_out->max_refs = 2; params_out->max_l0 = 1;
                params_out->only_l_bwd = 0; params_out->similar_brightness_refs = 0;
                params_out->gm_enabled = 0; params_out->gm_use_distance_based_active_th = 0;
                params_out->hierarchical_levels = 5;
            }
            g_table_at_last_b64 = ref_rtcd_table();
            svt_aom_motion_estimation_b64(pcs, b64_index, ox, oy, me, ip);
            const int exited = me->tf_use_pred_64x64_only_th == (uint8_t)~0; /* motion_estimation.c:3179-3183 */
            if (res) {
                const size_t o = (size_t)b64_index * 8;
                for (int p = 0; p < 85; p++) {
                    if (res->p_sb_best_sad) res->p_sb_best_sad[o * 85 + p] = exited ? 0 : me->p_sb_best_sad[0][0][p];
                    if (res->p_sb_best_mv) res->p_sb_best_mv[o * 85 + p] = exited ? 0 : me->p_sb_best_mv[0][0][p];
                }
                if (res->search_results) {
                    SvtMeSearchResult *sr = &res->search_results[o];
                    memset(sr, 0, sizeof(*sr));
                    sr->hme_sc_x = me->search_results[0][0].hme_sc_x;
                    sr->hme_sc_y = me->search_results[0][0].hme_sc_y;
                    sr->do_ref   = me->search_results[0][0].do_ref;
                    sr->hme_sad  = me->search_results[0][0].hme_sad;
                }
            }
            if (dbg) dump_debug(&dbg[b64_index], me);
        }
    }
    shim_forget();
    free(me);
    planes_free(&src);
    planes_free(&ref);
    fake_free(&f);
    return 0;
}
