// me_device.cuh — device-side data layout and the inner byte-SAD routines (sm_100a).
//
// Data layout in HBM
//   * one slab per picture: three replicate-padded luma planes (full: pad 80, 1/4: pad 32,
//     1/16: pad 16; the reference pads 72/32/16, reference_object.c:320-380, enc_handle.c:4485-4497),
//     row stride rounded up to 16 bytes, 512 bytes of slack before/after so that 16-byte aligned
//     TMA boxes and 4-byte aligned staging reads never leave the allocation.
//   * per-b64 HME state (SbState) written by k_hme_ctl and read by k_fp / k_post.
//   * per-(b64,list,ref) best SAD / MV arrays [85] written by k_fp / k_fp_big (MeContext
//     p_sb_best_sad / p_sb_best_mv, me_context.h:395-396).
//
// The arithmetic unit of the whole path is VABSDIFF4.U8.ACC (sad4_acc below): 4 pixel |a-b|
// accumulated per lane-instruction. The integer-ALU pipe it shares with shifts, adds and compares
// issues 64 lane-instr/clk/SM on B200 (profiles/r01_vabs_peak_microbench.txt) =
// the integer roofline of this library.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "svtme_b200.h"

#define SVTME_FULL_PAD 80 /* >= 64 + 4 (the reference pads b64 + ME_FILTER_TAP); 80 keeps b64 origins 16-byte aligned */
#define SVTME_QUARTER_PAD 32
#define SVTME_SIXTEENTH_PAD 16
#define SVTME_SLAB_SLACK 512

#define MAX_SAD_VALUE_U32 (128u * 128u * 255u) /* motion_estimation.h:85 */
#define MAX_U32_ 0xFFFFFFFFu

struct DevPlane {
    const uint8_t *buf; // address of padded sample (0,0)
    int stride, org_x, org_y, width, height;
};
struct DevPic {
    DevPlane    lvl[3]; // 0 full, 1 quarter, 2 sixteenth
    const void *tmap;   // CUtensorMap[TM_KINDS] of this picture's planes (me_tma.cuh), device memory
};

// per-b64 state produced by the HME stage (the MeContext fields later stages read) + the
// intermediate HME arrays kept for parity debugging (same layout as OrcDebug / RefMeDebug).
struct SbState {
    int16_t  x_hme[3][2][4][2][2]; // [level][list][ref][sr_w][sr_h]
    int16_t  y_hme[3][2][4][2][2];
    uint64_t hme_sad[3][2][4][2][2];
    int16_t  prehme_mv_x[2][4][2], prehme_mv_y[2][4][2];
    uint64_t prehme_sad[2][4][2];
    uint32_t zz_sad[2][4];
    uint32_t reduce_me_sr_divisor[2][4];
    // SearchResults
    int16_t  hme_sc_x[2][4], hme_sc_y[2][4];
    uint64_t sr_hme_sad[2][4];
    uint8_t  do_ref[2][4];
    uint8_t  performed_phme[2][4][2];
    uint8_t  prehme_valid[2][4][2];
    uint8_t  tf_exit; // ME_MCTF: skip the full-pel search of this b64 (motion_estimation.c:3179-3183)
    uint8_t  pad_[7];
};

// One window search of the HME stage (pre-HME region or HME level-N quadrant), set up by k_hme_ctl
// and executed by k_search16 (1/16 plane) / k_search_blk (1/4 and full planes). 32 jobs per b64: slot = (list*4+ref)*4 + quadrant
// (pre-HME uses quadrant 0/1 for its two regions).
struct HmeJobG {
    unsigned long long key;  // (sad << 32) | (y*sa_w + x): minimum over the job's tiles (k_search16: one warp, plain store; detector bands: k_dg_reduce)
    int16_t            sa_w, sa_h, origin_x, origin_y;
    uint16_t           tile_w, tile_h; // positions per tile (tile_w multiple of 8)
    uint8_t            n_tx, n_ty, active, wide; // wide: 1 = 128x18-position tiles, 0 = 16x96 (1/16-plane searches)
    int16_t            px, py;                   // plane coordinates of the window's top-left sample
    int32_t            src_off;                  // offset of the block's top-left sample in the source plane of the level
    uint8_t            bw, rows;                 // block width, block rows compared (after row sub-sampling)
    uint8_t            merged, pad_;             // 1: this job's unit covers the 2 x 2 union of the windows of slots +0..+3
                                                 //    (m_w x m_h, split at this job's own sa_w / sa_h); 2: covered by that unit
    int16_t            m_w, m_h;
};
#define HME_JOBS_PER_B64 32
#define HME_KEY_NONE ((0xffffffull << 32) | 0xffffffffull) /* best_sad = 0xffffff, nothing found */

// Full-pel search of one (b64, list, ref), set up by k_fp; windows of more than FP_INLINE_TILES tiles are finished by k_fp_big.
struct FpJobG {
    int16_t  xc, yc;             // search centre (variance-probe position, scan index 0)
    int16_t  x_origin, y_origin; // window origin relative to the b64
    int16_t  sa_w, sa_h;
    uint16_t n_tx, n_ty;         // tiles of TM_FP_TILE_W x TM_FP_TILE_H positions (me_tma.cuh)
    uint8_t  active, probe, pad_[2];
};

// everything one picture's kernels need; passed by value as a __grid_constant__ parameter
struct alignas(16) PicJob { // a multiple of 16 bytes: the job array is fetched by a kernel in 16-byte words
    SvtMePicParams pp;
    DevPic         src;
    DevPic         ref[2][4];
    int            pic_w, pic_h, b64_cols, b64_rows, n_b64, n_pus;
    int            mv_stride, mv_l1_off, cand_stride; // row geometry of me_mv_array / me_candidate_array as they are written (svtme_sb_strides)
    // device buffers
    SbState       *sb_state;   // [n_b64]
    uint32_t      *best_sad;   // [n_b64][2][4][85]
    uint32_t      *best_mv;    // [n_b64][2][4][85]
    SvtMeWindow   *windows;    // [n_b64][2][4]
    SvtMeSearchResult *search_results; // [n_b64][2][4]
    uint8_t       *total_me_candidate_index;
    uint32_t      *me_mv_array;
    uint8_t       *me_candidate_array;
    uint32_t      *me_64x64_distortion, *me_32x32_distortion, *me_16x16_distortion, *me_8x8_distortion;
    uint32_t      *rc_me_distortion, *me_8x8_cost_variance;
    uint8_t       *stationary_block_present_sb, *rc_me_allow_gm;
    unsigned long long *sad_ops; // [0] HME stage, [1] full-pel stage, [2] pre-HME, [3..5] HME level 0..2: algorithmic |a-b| operations
    // v2 pipeline scratch
    HmeJobG            *hme_jobs;  // [n_b64][HME_JOBS_PER_B64]
    FpJobG             *fp_jobs;   // [n_b64][8]
    unsigned long long *fp_keys;   // [n_b64][8][85]: (sad << 32) | global scan index (0 = probe)
    uint2              *fp_units;  // unit queue shared by the batch: x = picture in batch, y = (job << 16) | tile
    uint32_t           *fp_counters; // [0] number of units, [1] queue head
    uint32_t            fp_units_cap;
    uint2              *s16_units;    // unit queue of k_search16: x = picture in batch, y = (job slot in picture << 8) | tile
    uint32_t           *s16_counters; // [0] number of units, [1] queue head
    uint32_t            s16_units_cap;
    int                 s16_split;    // 1: k_hme_ctl makes every 1/16-plane job a queue unit of its own (small batches)
    uint32_t            inv_cols, inv_refs; // ceil(2^32 / b64_cols), ceil(2^32 / searched references): div_by(); 0 = divide
};

// n / d without the ~20-instruction integer division every warp of the per-b64 kernels started with: inv = ceil(2^32 / d) from
// the host (picjob_set_inverses), exact while n * d < 2^32; inv == 0 (d == 1, or a job record filled elsewhere): divide
__device__ __forceinline__ int div_by(int n, int d, uint32_t inv) { return inv ? (int)__umulhi((uint32_t)n, inv) : n / d; }

static_assert(sizeof(PicJob) % 16 == 0, "PicJob is copied in 16-byte words");

// ------------------------------------------------------------------------------------------
// small helpers
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ int imin(int a, int b) { return a < b ? a : b; }
__device__ __forceinline__ int imax(int a, int b) { return a > b ? a : b; }
__device__ __forceinline__ int iabs(int a) { return a < 0 ? -a : a; }

// 4 bytes starting at an arbitrary global address (two aligned loads + funnel shift). Reads at
// most 3 bytes before and 4..7 bytes after p: callers stay inside the slab slack.
__device__ __forceinline__ uint32_t ldg_u32_unaligned(const uint8_t *p) {
    const uintptr_t a  = (uintptr_t)p;
    const uint32_t *q  = (const uint32_t *)(a & ~(uintptr_t)3);
    const uint32_t  lo = __ldg(q), hi = __ldg(q + 1);
    return __funnelshift_r(lo, hi, (uint32_t)(a & 3) * 8);
}

// c + sum of |a.byte[i] - b.byte[i]|: one VABSDIFF4.U8.ACC. Written as PTX so that the accumulate operand is
// used: from `acc += __vsadu4(a, b)` the compiler builds VABSDIFF4(.., RZ) pairs joined by IADD3 (one
// extra ALU-pipe instruction per two SADs on the pipe that bounds these kernels).
__device__ __forceinline__ uint32_t sad4_acc(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t d;
    asm("vabsdiff4.u32.u32.u32.add %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}

// (hi:lo) >> 8*NB for a compile-time NB = 1..3 (bytes): one SHF on the integer-ALU pipe, the pipe that bounds every
// search kernel. SVTME_FSR_IMAD=1 builds the shift from IMAD.HI + IMAD instead (multiplier from constant memory so that
// ptxas cannot turn it back into shifts) to use the FMA pipe these kernels leave idle. MEASURED SLOWER on B200 and
// therefore off: IMAD.HI issues at 32 lane-instr/clk/SM, half the IMAD rate, and slows the ALU pipe down with it;
// 7090 against 7628 frames/s on the 1080p bench (profiles/r01_fsr_imad_experiment.md, tools/microbench/fsr_pipes.cu).
#ifndef SVTME_FSR_IMAD
#define SVTME_FSR_IMAD 0
#endif
static __constant__ uint32_t c_fsr_mul[4] = {0u, 1u << 24, 1u << 16, 1u << 8};
template <int NB>
__device__ __forceinline__ uint32_t fsr_bytes(uint32_t lo, uint32_t hi) {
    if (NB == 0) return lo;
#if SVTME_FSR_IMAD
    uint32_t h, r;
    const uint32_t m = c_fsr_mul[NB];
    asm("mul.hi.u32 %0, %1, %2;" : "=r"(h) : "r"(lo), "r"(m));
    asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(hi), "r"(m), "r"(h));
    return r;
#else
    return __funnelshift_r(lo, hi, 8 * NB);
#endif
}

// sh[j][i] = the word that starts j+1 bytes into w[i] (i = 0..4): the 15 shifted words the 8 positions of a strip share
__device__ __forceinline__ void strip_shifted_words(const uint32_t (&w)[6], uint32_t (&sh)[3][5]) {
#pragma unroll
    for (int i = 0; i < 5; i++) {
        sh[0][i] = fsr_bytes<1>(w[i], w[i + 1]);
        sh[1][i] = fsr_bytes<2>(w[i], w[i + 1]);
        sh[2][i] = fsr_bytes<3>(w[i], w[i + 1]);
    }
}

// acc[p] += SAD of the 16-byte strip s[0..3] against the strip starting p bytes into w[0..5]
// for the 8 consecutive search positions p = 0..7. 32 VABSDIFF4.U8.ACC + 15 funnel shifts.
__device__ __forceinline__ void strip16_sad8(const uint32_t (&w)[6], const uint32_t (&s)[4], uint32_t (&acc)[8]) {
    uint32_t sh[3][5];
    strip_shifted_words(w, sh);
#pragma unroll
    for (int p = 0; p < 8; p++) {
        const int o = p >> 2, j = p & 3;
#pragma unroll
        for (int k = 0; k < 4; k++) acc[p] = sad4_acc(j ? sh[j - 1][o + k] : w[o + k], s[k], acc[p]);
    }
}

// same, but the strip is split into its left and right 8 pixels (two 8x8 blocks side by side)
__device__ __forceinline__ void strip16_sad8_split(const uint32_t (&w)[6], const uint32_t (&s)[4],
                                                   uint32_t (&accl)[8], uint32_t (&accr)[8]) {
    uint32_t sh[3][5];
    strip_shifted_words(w, sh);
#pragma unroll
    for (int p = 0; p < 8; p++) {
        const int o = p >> 2, j = p & 3;
#pragma unroll
        for (int k = 0; k < 4; k++) {
            const uint32_t v = j ? sh[j - 1][o + k] : w[o + k];
            if (k < 2)
                accl[p] = sad4_acc(v, s[k], accl[p]);
            else
                accr[p] = sad4_acc(v, s[k], accr[p]);
        }
    }
}

// byte mask keeping the first n (0..4) bytes of a little-endian word
__device__ __forceinline__ uint32_t lo_bytes_mask(int n) {
    return n <= 0 ? 0u : (n >= 4 ? 0xFFFFFFFFu : ((1u << (8 * n)) - 1u));
}

// Stage a rectangle of bytes from global memory (arbitrary alignment) into shared memory whose
// rows are 4-byte aligned: smem word (r, c) = global bytes [g + r*gstride + 4c, +4).
// Plane strides are multiples of 4, so the byte misalignment is the same for every row: one funnel
// shift amount per call, (row, column) advanced incrementally (no division per word).
__device__ __forceinline__ void stage_rect(uint32_t *smem, int smem_stride_words, const uint8_t *g, int gstride,
                                           int rows, int words_per_row, int tid, int nthreads) {
    if ((gstride & 3) == 0) {
        const uint32_t  sa  = (uint32_t)((uintptr_t)g & 3) * 8;
        const uint32_t *q   = (const uint32_t *)((uintptr_t)g & ~(uintptr_t)3);
        const int       gs4 = gstride >> 2;
        int r = tid / words_per_row, c = tid - r * words_per_row;
        const int dr = nthreads / words_per_row, dc = nthreads - dr * words_per_row;
        // 4 words (8 loads) in flight per lane: the copy is latency bound, not bandwidth bound
        while (r < rows) {
            int      rk[4], ck[4];
            uint32_t lo[4], hi[4];
#pragma unroll
            for (int k = 0; k < 4; k++) {
                rk[k] = r; ck[k] = c;
                c += dc; r += dr;
                if (c >= words_per_row) { c -= words_per_row; r++; }
            }
#pragma unroll
            for (int k = 0; k < 4; k++) {
                if (rk[k] < rows) {
                    const uint32_t *p = q + (size_t)rk[k] * gs4 + ck[k];
                    lo[k] = __ldg(p);
                    hi[k] = __ldg(p + 1);
                }
            }
#pragma unroll
            for (int k = 0; k < 4; k++)
                if (rk[k] < rows) smem[rk[k] * smem_stride_words + ck[k]] = __funnelshift_r(lo[k], hi[k], sa);
        }
    } else { // odd strides only occur in the RTCD-signature test entry points
        const int total = rows * words_per_row;
        for (int i = tid; i < total; i += nthreads) {
            const int r = i / words_per_row, c = i - r * words_per_row;
            smem[r * smem_stride_words + c] = ldg_u32_unaligned(g + (size_t)r * gstride + 4 * c);
        }
    }
}

// svt_aom_get_scaled_picture_distance, motion_estimation.c:1239-1243
__device__ __forceinline__ uint16_t scaled_picture_distance(uint16_t dist) {
    const uint8_t round_up = ((dist % 8) == 0) ? 0 : 1;
    return (uint16_t)(((dist * 5) / 8) + round_up);
}

// |poc - ref_poc| as get_me_reference computes it, motion_estimation.c:1232-1234
__device__ __forceinline__ uint16_t ref_distance(const SvtMePicParams &pp, int l, int r) {
    long long d = (long long)pp.picture_number - (long long)pp.ref_picture_number[l][r];
    if (d < 0) d = -d;
    return (uint16_t)(int16_t)d;
}

// The edge correction every search stage applies to one axis, in int16 like the reference
// (motion_estimation.c:848-880, 949-982, 1050-1078, 1481-1500, 1542-1560, 1655-1699). The left/top
// branch moves the origin first, so the size correction that follows it subtracts zero.
__device__ __forceinline__ void clamp_axis(int16_t org, int16_t &origin, int16_t &size, int16_t pad, int16_t dim) {
    if ((org + origin) < -pad) {
        origin = (int16_t)(-pad - org);
        size   = (int16_t)(size - (-pad - (org + origin)));
    }
    if ((org + origin) > dim - 1) origin = (int16_t)(origin - ((org + origin) - (dim - 1)));
    if ((org + origin + size) > dim) size = (int16_t)imax(1, size - ((org + origin + size) - dim));
}
