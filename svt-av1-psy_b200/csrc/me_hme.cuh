// me_hme.cuh — k_hme: the whole hierarchical-ME stage of one 64x64 block per CTA.
//
// Replaces hme_b64 + hme_prune_ref_and_adjust_sr (motion_estimation.c:2511-2588) and everything
// they call: init_zz_sad (:2452), prehme_b64 (:1792), hme_level0/1/2_b64 (:1976/:2111/:2197),
// set_final_seach_centre_sb (:2252). The searches are svt_sad_loop_kernel (compute_sad_c.c:58-101)
// semantics: raster scan, strict '<', best starts at 0xffffff.
//
// Work split inside the CTA
//   * scalar control (window clamping, early exits, pruning): thread 0, literal int16 arithmetic
//   * pre-HME / level 0 (16-wide blocks on the 1/16 plane, windows of 1000s of points): the window
//     of one search is staged in shared memory (4-byte aligned rows) and every thread evaluates
//     8 consecutive positions per task with strip16_sad8 (32 VABSDIFF4 per block row)
//   * level 1 / level 2 (32- and 64-wide blocks, 8x3..16x16 windows): all searches of all refs form
//     one flat task list (search, row, octet, 16-pixel strip); reference rows come straight from
//     global memory (L1-resident), strips are summed with warp shuffles
//   * arg-min: 64-bit keys (sad << 32 | y*sa_w + x), in-thread strict '<' in scan order, then
//     warp shuffle + shared-memory atomicMin — the smallest key is the first minimum in raster order.
#pragma once
#include "me_device.cuh"

#define HME_THREADS 256
#define HME_MAX_JOBS 32
#define HME_KEY_INIT ((0xffffffull << 32) | 0xffffffffull) /* best_sad = 0xffffff, nothing found */

struct HmeJob {
    const uint8_t *ref;        // window top-left sample in the reference plane
    int            ref_stride; // plane stride (bytes)
    int            sa_w, sa_h;
    int16_t        origin_x, origin_y;
    uint8_t        l, r, a, b; // a,b: (sr_w, sr_h) for HME levels, (region, 0) for pre-HME
};

struct HmeSh {
    SbState  st;
    uint8_t  performed_phme[2][4][2];
    uint8_t  prehme_valid[2][4][2];
    HmeJob   jobs[HME_MAX_JOBS];
    unsigned long long job_key[HME_MAX_JOBS];
    int      task_prefix[HME_MAX_JOBS + 1];
    int      n_jobs;
    // source block of this b64 at the three pyramid levels, rows already sub-sampled when
    // hme_search_method == SUB_SAD_SEARCH; bytes beyond the block width are zero
    alignas(16) uint32_t src16[16 * 4];  // 1/16 plane: <= 16 rows x 16 bytes
    alignas(16) uint32_t src4[32 * 8];   // 1/4 plane : <= 32 rows x 32 bytes
    alignas(16) uint32_t src1[64 * 16];  // full plane: <= 64 rows x 64 bytes
};

// ---- scalar helpers (thread 0) --------------------------------------------------------------

// window clamp + plane address shared by prehme_core and hme_level_0/1/2
__device__ inline void hme_make_job(HmeJob &j, const DevPlane &ref, int16_t org_x, int16_t org_y, int16_t origin_x,
                                    int16_t origin_y, int16_t sa_w, int16_t sa_h, int16_t pad_w, int16_t pad_h,
                                    bool round_w8, uint8_t l, uint8_t r, uint8_t a, uint8_t b) {
    clamp_axis(org_x, origin_x, sa_w, pad_w, (int16_t)ref.width);
    if (round_w8) sa_w = (int16_t)((sa_w < 8) ? sa_w : sa_w & ~0x07);
    clamp_axis(org_y, origin_y, sa_h, pad_h, (int16_t)ref.height);
    const int16_t x_tl = (int16_t)(((int16_t)ref.org_x + org_x) + origin_x);
    const int16_t y_tl = (int16_t)(((int16_t)ref.org_y + org_y) + origin_y);
    j.ref        = ref.buf + (x_tl + y_tl * ref.stride);
    j.ref_stride = ref.stride;
    j.sa_w       = sa_w;
    j.sa_h       = sa_h;
    j.origin_x   = origin_x;
    j.origin_y   = origin_y;
    j.l = l; j.r = r; j.a = a; j.b = b;
}

// result of one search: sad (doubled when rows were sub-sampled), centre scaled to full resolution.
// cx/cy keep their previous value when no position was evaluated (svt_sad_loop_kernel contract).
__device__ inline void hme_job_result(const HmeJob &j, unsigned long long key, bool sub, int scale, uint64_t &sad,
                                      int16_t &cx, int16_t &cy) {
    uint64_t s = key >> 32;
    if ((uint32_t)key != 0xffffffffu) {
        const int idx = (int)(uint32_t)key;
        cx = (int16_t)(idx % j.sa_w);
        cy = (int16_t)(idx / j.sa_w);
    }
    if (sub) s = s * 2;
    sad = s;
    cx  = (int16_t)(cx + j.origin_x);
    cx  = (int16_t)(cx * scale);
    cy  = (int16_t)(cy + j.origin_y);
    cy  = (int16_t)(cy * scale);
}

__device__ inline void hme_set_quadrants(HmeSh &sh, const SvtMeCtrls &c, int lvl, int l, int r, int16_t cx, int16_t cy,
                                         uint64_t sad) {
    for (int b = 0; b < c.num_hme_sa_h; b++)
        for (int a = 0; a < c.num_hme_sa_w; a++) {
            sh.st.x_hme[lvl][l][r][a][b]   = cx;
            sh.st.y_hme[lvl][l][r][a][b]   = cy;
            sh.st.hme_sad[lvl][l][r][a][b] = sad;
        }
}

// ---- cooperative pieces ----------------------------------------------------------------------

// Stage the b64's source block of one pyramid level into shared memory: n_rows rows of
// `words` 32-bit words, row r taken from src + r*row_step*stride; bytes >= bw are zeroed.
__device__ inline void hme_stage_src(uint32_t *dst, int words, const uint8_t *src, int stride, int row_step, int n_rows,
                                     int bw, int tid) {
    for (int i = tid; i < n_rows * words; i += HME_THREADS) {
        const int r = i / words, k = i - r * words;
        uint32_t  v = ldg_u32_unaligned(src + (size_t)r * row_step * stride + 4 * k);
        dst[i]      = v & lo_bytes_mask(bw - 4 * k);
    }
}

// 64-bit min across the CTA into *slot
__device__ __forceinline__ void hme_commit_key(unsigned long long key, unsigned long long *slot) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const unsigned long long other = __shfl_xor_sync(0xffffffffu, key, o);
        key = other < key ? other : key;
    }
    if ((threadIdx.x & 31) == 0 && key != HME_KEY_INIT) atomicMin(slot, key);
}

// Search A: one search with a 16-wide (or narrower) block; window staged in shared memory.
// svt_sad_loop_kernel semantics incl. skip_search_line (even search rows skipped when the block is
// 16 wide and at most 16 rows, compute_sad_c.c:74-79).
__device__ inline void hme_search_staged(HmeSh &sh, const HmeJob &j, int job_index, uint32_t *win, int win_words,
                                         const uint32_t *src /*[n_rows][4]*/, int bw, int n_rows, int row_step,
                                         bool skip_lines) {
    const int tid     = threadIdx.x;
    const int n_oct   = (j.sa_w + 7) >> 3;
    const int stride  = 2 * n_oct + 4; // words per staged row: 8*n_oct + 16 bytes
    const int span    = row_step * (n_rows - 1);
    int       tile_h  = win_words / stride - span;
    if (tile_h < 1) tile_h = 1; // host sizes the buffer so that this never triggers
    const bool     partial = bw < 16;
    const uint32_t m0 = lo_bytes_mask(bw), m1 = lo_bytes_mask(bw - 4), m2 = lo_bytes_mask(bw - 8), m3 = lo_bytes_mask(bw - 12);
    uint32_t best_sad = 0xffffff, best_idx = 0xffffffffu;
    for (int y0 = 0; y0 < j.sa_h; y0 += tile_h) {
        const int th = imin(tile_h, j.sa_h - y0);
        __syncthreads(); // previous tile fully consumed
        stage_rect(win, stride, j.ref + (size_t)y0 * j.ref_stride, j.ref_stride, th + span, stride, tid, HME_THREADS);
        __syncthreads();
        const int n_task = n_oct * th;
        for (int t = tid; t < n_task; t += HME_THREADS) {
            const int yy = t / n_oct, oct = t - yy * n_oct;
            const int y  = y0 + yy;
            if (skip_lines && (y & 1) == 0) continue;
            uint32_t acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
            const uint32_t *row = win + yy * stride + 2 * oct;
            for (int rr = 0; rr < n_rows; rr++, row += row_step * stride) {
                uint32_t w[6], s[4];
                const uint2 a = *(const uint2 *)(row), b = *(const uint2 *)(row + 2), cc = *(const uint2 *)(row + 4);
                w[0] = a.x; w[1] = a.y; w[2] = b.x; w[3] = b.y; w[4] = cc.x; w[5] = cc.y;
                const uint4 sv = *(const uint4 *)(src + 4 * rr);
                s[0] = sv.x; s[1] = sv.y; s[2] = sv.z; s[3] = sv.w;
                if (!partial) {
                    strip16_sad8(w, s, acc);
                } else {
#pragma unroll
                    for (int p = 0; p < 8; p++) {
                        const int o = p >> 2, shf = (p & 3) * 8;
                        const uint32_t v0 = shf ? __funnelshift_r(w[o], w[o + 1], shf) : w[o];
                        const uint32_t v1 = shf ? __funnelshift_r(w[o + 1], w[o + 2], shf) : w[o + 1];
                        const uint32_t v2 = shf ? __funnelshift_r(w[o + 2], w[o + 3], shf) : w[o + 2];
                        const uint32_t v3 = shf ? __funnelshift_r(w[o + 3], w[o + 4], shf) : w[o + 3];
                        acc[p] += __vsadu4(v0 & m0, s[0]) + __vsadu4(v1 & m1, s[1]) + __vsadu4(v2 & m2, s[2]) +
                            __vsadu4(v3 & m3, s[3]);
                    }
                }
            }
#pragma unroll
            for (int p = 0; p < 8; p++) {
                const int x = 8 * oct + p;
                if (x < j.sa_w && acc[p] < best_sad) {
                    best_sad = acc[p];
                    best_idx = (uint32_t)(y * j.sa_w + x);
                }
            }
        }
    }
    hme_commit_key(((unsigned long long)best_sad << 32) | best_idx, &sh.job_key[job_index]);
}

// Search B: all jobs in sh.jobs[0..n_jobs) at once, blocks of `strips` 16-pixel strips, reference
// rows read from global memory. One task = (job, y, octet, strip).
__device__ inline void hme_search_direct(HmeSh &sh, const uint32_t *src /*[n_rows][4*strips]*/, int strips, int bw,
                                         int n_rows, int row_step) {
    const int tid   = threadIdx.x;
    const int total = sh.task_prefix[sh.n_jobs];
    const int sbits = strips == 4 ? 2 : (strips == 2 ? 1 : 0);
    for (int base = (tid & ~31); base < total; base += HME_THREADS) {
        const int  t      = base + (tid & 31);
        const bool active = t < total;
        int ji = 0;
        if (active)
            while (t >= sh.task_prefix[ji + 1]) ji++;
        const HmeJob &j     = sh.jobs[ji];
        const int     n_oct = (j.sa_w + 7) >> 3;
        const int     local = active ? t - sh.task_prefix[ji] : 0;
        const int     strip = local & (strips - 1);
        const int     oy    = local >> sbits;
        const int     y = oy / n_oct, oct = oy - y * n_oct;
        uint32_t      acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        if (active) {
            const uint8_t *g    = j.ref + (size_t)y * j.ref_stride + 8 * oct + 16 * strip;
            const int      rstr = row_step * j.ref_stride;
            const uint32_t m0 = lo_bytes_mask(bw - 16 * strip), m1 = lo_bytes_mask(bw - 16 * strip - 4),
                           m2 = lo_bytes_mask(bw - 16 * strip - 8), m3 = lo_bytes_mask(bw - 16 * strip - 12);
            const bool partial = bw < 16 * strip + 16;
            for (int rr = 0; rr < n_rows; rr++, g += rstr) {
                const uintptr_t a  = (uintptr_t)g;
                const uint32_t *q  = (const uint32_t *)(a & ~(uintptr_t)3);
                const uint32_t  sa = (uint32_t)(a & 3) * 8;
                uint32_t        gw[7], w[6], s[4];
#pragma unroll
                for (int k = 0; k < 7; k++) gw[k] = __ldg(q + k);
#pragma unroll
                for (int k = 0; k < 6; k++) w[k] = __funnelshift_r(gw[k], gw[k + 1], sa);
                const uint4 sv = *(const uint4 *)(src + (size_t)rr * 4 * strips + 4 * strip);
                s[0] = sv.x; s[1] = sv.y; s[2] = sv.z; s[3] = sv.w;
                if (!partial) {
                    strip16_sad8(w, s, acc);
                } else {
#pragma unroll
                    for (int p = 0; p < 8; p++) {
                        const int o = p >> 2, shf = (p & 3) * 8;
                        const uint32_t v0 = shf ? __funnelshift_r(w[o], w[o + 1], shf) : w[o];
                        const uint32_t v1 = shf ? __funnelshift_r(w[o + 1], w[o + 2], shf) : w[o + 1];
                        const uint32_t v2 = shf ? __funnelshift_r(w[o + 2], w[o + 3], shf) : w[o + 2];
                        const uint32_t v3 = shf ? __funnelshift_r(w[o + 3], w[o + 4], shf) : w[o + 3];
                        acc[p] += __vsadu4(v0 & m0, s[0]) + __vsadu4(v1 & m1, s[1]) + __vsadu4(v2 & m2, s[2]) +
                            __vsadu4(v3 & m3, s[3]);
                    }
                }
            }
        }
        // sum the strips of one (job, y, octet): they sit in adjacent lanes
#pragma unroll
        for (int p = 0; p < 8; p++) {
            if (strips >= 2) acc[p] += __shfl_xor_sync(0xffffffffu, acc[p], 1);
            if (strips >= 4) acc[p] += __shfl_xor_sync(0xffffffffu, acc[p], 2);
        }
        if (active && strip == 0) {
            uint32_t best_sad = 0xffffff, best_idx = 0xffffffffu;
#pragma unroll
            for (int p = 0; p < 8; p++) {
                const int x = 8 * oct + p;
                if (x < j.sa_w && acc[p] < best_sad) {
                    best_sad = acc[p];
                    best_idx = (uint32_t)(y * j.sa_w + x);
                }
            }
            if (best_idx != 0xffffffffu) atomicMin(&sh.job_key[ji], ((unsigned long long)best_sad << 32) | best_idx);
        }
    }
}

// plain w x h SAD at one position, computed by one warp (svt_nxm_sad_kernel, compute_sad_c.c:20-37)
__device__ inline uint32_t warp_nxm_sad(const uint8_t *src, int src_stride, const uint8_t *ref, int ref_stride,
                                        int height, int width) {
    const int lane = threadIdx.x & 31;
    uint32_t  sad  = 0;
    const int wpr  = (width + 3) >> 2; // words per row
    for (int i = lane; i < height * wpr; i += 32) {
        const int      r = i / wpr, k = i - r * wpr;
        const uint32_t m = lo_bytes_mask(width - 4 * k);
        const uint32_t a = ldg_u32_unaligned(src + (size_t)r * src_stride + 4 * k) & m;
        const uint32_t b = ldg_u32_unaligned(ref + (size_t)r * ref_stride + 4 * k) & m;
        sad += __vsadu4(a, b);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) sad += __shfl_xor_sync(0xffffffffu, sad, o);
    return sad;
}
