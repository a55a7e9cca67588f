// me_noalloc_search.cuh — the CTA-wide staged window search used by the no-allocation entry point
// (me_noalloc.cu). svt_sad_loop_kernel semantics (compute_sad_c.c:58-101): raster scan, strict '<', best starts at
// 0xffffff; result = 64-bit key (sad << 32 | y*sa_w + x) whose minimum is the first minimum in raster order.
// The picture pipeline uses the TMA tile searches of me_search.cuh instead; this one needs no tensor map and no
// device memory besides the caller's buffers, which is the point of that entry.
#pragma once
#include "me_device.cuh"

#define HME_THREADS 256
#define HME_KEY_INIT ((0xffffffull << 32) | 0xffffffffull) /* best_sad = 0xffffff, nothing found */

struct HmeJob {
    const uint8_t *ref;        // window top-left sample in the reference plane
    int            ref_stride; // plane stride (bytes)
    int            sa_w, sa_h;
    int16_t        origin_x, origin_y;
};

struct HmeSh {
    HmeJob             jobs[1];
    unsigned long long job_key[1];
    alignas(16) uint32_t src16[16 * 4]; // the block: <= 16 rows x 16 bytes, bytes beyond its width zeroed
};

// ---- cooperative pieces ----------------------------------------------------------------------

// Stage the b64's source block of one pyramid level into shared memory: n_rows rows of
// `words` 32-bit words, row r taken from src + r*row_step*stride; bytes >= bw are zeroed.
// The caller's buffers may end flush with their allocation, so only the bytes svt_sad_loop_kernel itself would read are
// touched: bw bytes per source row, sa_w + bw - 1 bytes per reference row.
__device__ inline void hme_stage_src(uint32_t *dst, int words, const uint8_t *src, int stride, int row_step, int n_rows,
                                     int bw, int tid) {
    for (int i = tid; i < n_rows * words; i += HME_THREADS) {
        const int r = i / words, k = i - r * words;
        const uint8_t *p = src + (size_t)r * row_step * stride + 4 * k;
        uint32_t v = 0;
        for (int b = 0; b < 4 && 4 * k + b < bw; b++) v |= (uint32_t)__ldg(p + b) << (8 * b);
        dst[i] = v;
    }
}

// smem word (r, c) = bytes [4c, 4c + 4) of row r of the rectangle at g (row pitch gstride), bytes at or beyond `need`
// read as zero and are not loaded. Whole aligned words are used where both aligned words under a staged word lie inside
// the row's needed bytes, single bytes at the two ends of a row.
__device__ inline void stage_rect_exact(uint32_t *smem, int smem_stride_words, const uint8_t *g, int gstride, int rows,
                                        int words_per_row, int need, int tid, int nthreads) {
    const int total = rows * words_per_row;
    for (int i = tid; i < total; i += nthreads) {
        const int r = i / words_per_row, c = i - r * words_per_row;
        const uint8_t *row = g + (size_t)r * gstride, *p = row + 4 * c;
        const int n = need - 4 * c; // bytes of this word that exist
        uint32_t v = 0;
        if (n > 0) {
            const uintptr_t a = (uintptr_t)p, lo = a & ~(uintptr_t)3;
            if (n >= 4 && lo >= (uintptr_t)row && lo + 8 <= (uintptr_t)row + (uintptr_t)need) {
                const uint32_t *q = (const uint32_t *)lo;
                v = __funnelshift_r(__ldg(q), __ldg(q + 1), (uint32_t)(a & 3) * 8);
            } else if (n >= 4 && (a & 3) == 0) {
                v = __ldg((const uint32_t *)p);
            } else {
                for (int b = 0; b < 4 && b < n; b++) v |= (uint32_t)__ldg(p + b) << (8 * b);
            }
        }
        smem[r * smem_stride_words + c] = v;
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
        stage_rect_exact(win, stride, j.ref + (size_t)y0 * j.ref_stride, j.ref_stride, th + span, stride, j.sa_w + bw - 1, tid, HME_THREADS);
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
                        acc[p] = sad4_acc(v0 & m0, s[0], acc[p]);
                        acc[p] = sad4_acc(v1 & m1, s[1], acc[p]);
                        acc[p] = sad4_acc(v2 & m2, s[2], acc[p]);
                        acc[p] = sad4_acc(v3 & m3, s[3], acc[p]);
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
