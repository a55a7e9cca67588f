// me_search.cuh — the HME-stage search kernels of the v2 pipeline: every window search of every
// b64 / reference of a picture is an independent unit of work, one warp each, no CTA barriers.
//
//   k_zz          zero-MV SAD per (b64, ref)                      init_zz_sad        motion_estimation.c:2452
//   k_search16    pre-HME regions and HME level-0 quadrants       prehme_core :1638, hme_level_0 :820
//                 (blocks <= 16 wide on the 1/16 plane): unit = (job, tile of the window); the warp
//                 stages the tile (+ the rows/cols the block reaches) in its own 4.5 KB of shared
//                 memory with 4-byte aligned rows, then each lane evaluates 8 consecutive positions
//                 per task (strip16_sad8: 6 words -> 24 SHF + 32 VABSDIFF4 per block row)
//   k_search_wide HME level 1 / 2 (32- / 64-wide blocks, windows of 8x3 .. 16x16 positions): unit =
//                 32 consecutive tasks (job, y, octet, 16-pixel strip) of a flat list with a fixed
//                 number of task slots per job; reference rows straight from global memory
//
// All follow svt_sad_loop_kernel (compute_sad_c.c:58-101): raster order, strict '<', best starts at
// 0xffffff. Result = 64-bit key (sad << 32 | y*sa_w + x); atomicMin over the tiles of a job keeps
// the first minimum in raster order.
#pragma once
#include "me_device.cuh"

#define SEARCH_WARPS 4

struct B64Geo {
    int org_x, org_y, b64_w, b64_h;
};
__device__ __forceinline__ B64Geo b64_geo(const PicJob &job, int b) {
    B64Geo g;
    g.org_x = (b % job.b64_cols) * 64;
    g.org_y = (b / job.b64_cols) * 64;
    g.b64_w = imin(64, job.pic_w - g.org_x); // motion_estimation.c:3163-3170, picture already 8-aligned
    g.b64_h = imin(64, job.pic_h - g.org_y);
    return g;
}

__device__ __forceinline__ unsigned long long warp_min_u64(unsigned long long key) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const unsigned long long other = __shfl_xor_sync(0xffffffffu, key, o);
        key = other < key ? other : key;
    }
    return key;
}

// plain w x h SAD at one position by one warp (svt_nxm_sad_kernel, compute_sad_c.c:20-37)
__device__ inline uint32_t warp_nxm_sad2(const uint8_t *src, int src_stride, const uint8_t *ref, int ref_stride,
                                         int height, int width) {
    const int lane = threadIdx.x & 31;
    uint32_t  sad  = 0;
    const int wpr  = (width + 3) >> 2;
    for (int i = lane; i < height * wpr; i += 32) {
        const int      r = i / wpr, k = i - r * wpr;
        const uint32_t m = lo_bytes_mask(width - 4 * k);
        const uint32_t a = ldg_u32_unaligned(src + (size_t)r * src_stride + 4 * k) & m;
        const uint32_t b = ldg_u32_unaligned(ref + (size_t)r * ref_stride + 4 * k) & m;
        sad = sad4_acc(a, b, sad);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) sad += __shfl_xor_sync(0xffffffffu, sad, o);
    return sad;
}

// ---- k_zz ------------------------------------------------------------------------------------
__global__ void __launch_bounds__(32 * SEARCH_WARPS) k_zz(const PicJob *__restrict__ jobs) {
    const PicJob &job = jobs[blockIdx.y];
    const int u = blockIdx.x * SEARCH_WARPS + (threadIdx.x >> 5);
    if (u >= job.n_b64 * 8) return;
    const int b = u >> 3, l = (u >> 2) & 1, r = u & 3;
    const SvtMePicParams &pp = job.pp;
    if (l >= pp.num_of_list_to_search || r >= pp.num_of_ref_pic_to_search[l]) return;
    if (!(pp.temporal_layer_index > 0 || l == 0)) return;
    const B64Geo    g  = b64_geo(job, b);
    const DevPlane &sf = job.src.lvl[0], &rf = job.ref[l][r].lvl[0];
    const uint8_t *sp = sf.buf + (size_t)(sf.org_y + g.org_y) * sf.stride + sf.org_x + g.org_x;
    const uint8_t *rp = rf.buf + ((int16_t)rf.org_x + (int16_t)g.org_x + ((int16_t)rf.org_y + (int16_t)g.org_y) * rf.stride);
    uint32_t zz = warp_nxm_sad2(sp, sf.stride << 1, rp, rf.stride << 1, g.b64_h >> 1, g.b64_w) << 1;
    zz = (zz * 64 * 64) / (uint32_t)(g.b64_w * g.b64_h); // :2465
    if ((threadIdx.x & 31) == 0) {
        job.sb_state[b].zz_sad[l][r] = zz;
        if (job.sad_ops) atomicAdd(&job.sad_ops[0], (unsigned long long)(g.b64_h >> 1) * g.b64_w);
    }
}

// ---- warp-level staged search of one tile, block <= 16 wide -----------------------------------
// win: this warp's staging buffer (S16_WIN_WORDS words, 16-byte aligned); src: n_rows x 4 words,
// bytes >= bw zeroed. Returns the warp-wide best key of the tile (positions indexed in the window).
__device__ inline unsigned long long warp_search16_tile(uint32_t *win, const uint32_t *src, const uint8_t *ref,
                                                        int ref_stride, int sa_w, int x0, int y0, int tw, int th, int bw,
                                                        int n_rows, int row_step, bool skip_lines) {
    const int lane   = threadIdx.x & 31;
    const int n_oct  = (tw + 7) >> 3;
    const int wpr    = 2 * n_oct + 4;                      // words per staged row
    const int stride = (wpr & 15) == 0 ? wpr + 2 : wpr;    // a row pitch of 16k words would put every row on the same banks
    const int span   = row_step * (n_rows - 1);
    stage_rect(win, stride, ref + (size_t)y0 * ref_stride + x0, ref_stride, th + span, wpr, lane, 32);
    __syncwarp();
    const bool     partial = bw < 16;
    const uint32_t m0 = lo_bytes_mask(bw), m1 = lo_bytes_mask(bw - 4), m2 = lo_bytes_mask(bw - 8), m3 = lo_bytes_mask(bw - 12);
    uint32_t best_sad = 0xffffff, best_idx = 0xffffffffu;
    int yy = lane / n_oct, oct = lane - yy * n_oct;
    const int dyy = 32 / n_oct, doct = 32 - dyy * n_oct;
    for (; yy < th; yy += dyy, oct += doct) {
        if (oct >= n_oct) { oct -= n_oct; yy++; if (yy >= th) break; }
        const int y  = y0 + yy;
        if (skip_lines && (y & 1) == 0) continue;
        uint32_t acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        const uint32_t *row = win + yy * stride + 2 * oct;
#pragma unroll 2
        for (int rr = 0; rr < n_rows; rr++, row += row_step * stride) {
            const uint2 a = *(const uint2 *)(row), b = *(const uint2 *)(row + 2), cc = *(const uint2 *)(row + 4);
            const uint4 sv = *(const uint4 *)(src + 4 * rr);
            const uint32_t w[6] = {a.x, a.y, b.x, b.y, cc.x, cc.y};
            const uint32_t s[4] = {sv.x, sv.y, sv.z, sv.w};
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
                    acc[p] += __vsadu4(v0 & m0, s[0]) + __vsadu4(v1 & m1, s[1]) + __vsadu4(v2 & m2, s[2]) + __vsadu4(v3 & m3, s[3]);
                }
            }
        }
#pragma unroll
        for (int p = 0; p < 8; p++) {
            const int x = 8 * oct + p;
            if (x < tw && acc[p] < best_sad) {
                best_sad = acc[p];
                best_idx = (uint32_t)(y * sa_w + x0 + x);
            }
        }
    }
    __syncwarp();
    return warp_min_u64(((unsigned long long)best_sad << 32) | best_idx);
}

// stage the source block rows of one pyramid level for one warp: n_rows x `words` words
__device__ inline void warp_stage_src(uint32_t *dst, int words, const uint8_t *src, int stride, int row_step, int n_rows,
                                      int bw) {
    const int lane = threadIdx.x & 31;
    for (int i = lane; i < n_rows * words; i += 32) {
        const int r = i / words, k = i - r * words;
        dst[i] = ldg_u32_unaligned(src + (size_t)r * row_step * stride + 4 * k) & lo_bytes_mask(bw - 4 * k);
    }
}

// ---- k_search16 ------------------------------------------------------------------------------
// persistent warps pull (picture, job, tile) units from the queue k_hme_ctl filled
__global__ void __launch_bounds__(32 * SEARCH_WARPS, 8) k_search16(const PicJob *__restrict__ jobs, int skip_cfg) {
    __shared__ __align__(16) uint32_t s_win[SEARCH_WARPS][S16_WIN_WORDS];
    __shared__ __align__(16) uint32_t s_src[SEARCH_WARPS][16 * 4];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const PicJob &job0 = jobs[0];
    const uint32_t n_units = min(job0.s16_counters[0], job0.s16_units_cap);
    int cur_b = -1, cur_pic = -1;
    const bool sub = job0.pp.ctrls.hme_search_method == SVTME_SUB_SAD_SEARCH;
    const int row_step = sub ? 2 : 1;
    int bw = 0, rows = 0;
    for (;;) {
        uint32_t ui = 0;
        if (lane == 0) ui = atomicAdd(&job0.s16_counters[1], 1u);
        ui = __shfl_sync(0xffffffffu, ui, 0);
        if (ui >= n_units) break;
        const uint2 unit = job0.s16_units[ui];
        const PicJob &job = jobs[unit.x];
        const int slot = (int)(unit.y >> 8), tile = (int)(unit.y & 0xff);
        HmeJobG &j = job.hme_jobs[slot];
        const int b = slot / HME_JOBS_PER_B64;
        if (b != cur_b || (int)unit.x != cur_pic) { // source block of this b64 on the 1/16 plane
            const B64Geo g = b64_geo(job, b);
            bw = g.b64_w >> 2;
            rows = sub ? (g.b64_h >> 2) >> 1 : (g.b64_h >> 2);
            const DevPlane &ss = job.src.lvl[2];
            __syncwarp();
            warp_stage_src(s_src[warp], 4, ss.buf + (size_t)(ss.org_y + (g.org_y >> 2)) * ss.stride + ss.org_x + (g.org_x >> 2),
                           ss.stride, row_step, rows, bw);
            cur_b = b; cur_pic = (int)unit.x;
        }
        const int ty = tile / j.n_tx, tx = tile - ty * j.n_tx;
        const int x0 = tx * j.tile_w, y0 = ty * j.tile_h;
        const int tw = imin(j.tile_w, j.sa_w - x0), th = imin(j.tile_h, j.sa_h - y0);
        const bool skip = skip_cfg && bw == 16 && rows <= 16;
        const unsigned long long key = warp_search16_tile(s_win[warp], s_src[warp], j.ref, j.ref_stride, j.sa_w, x0, y0, tw, th,
                                                         bw, rows, row_step, skip);
        if (lane == 0 && key != HME_KEY_NONE) atomicMin(&j.key, key);
        __syncwarp();
    }
}

// ---- wide blocks: one task = (y, octet, 16-pixel strip) of one search, rows from global memory ----
// Returns in acc[] the strip-summed SADs of the 8 positions (valid in every lane of the strip group).
// search rows advance by search_stride, block rows by ref_blk_stride / src_blk_stride.
__device__ inline void warp_search_wide_task(bool active, const uint8_t *ref, int search_stride, int ref_blk_stride,
                                             const uint8_t *src, int src_blk_stride, int bw, int rows, int y, int oct,
                                             int strip, int strips, uint32_t (&acc)[8]) {
#pragma unroll
    for (int p = 0; p < 8; p++) acc[p] = 0;
    if (active) {
        const uint8_t *s  = src + 16 * strip;
        const uint8_t *gr = ref + (size_t)y * search_stride + 8 * oct + 16 * strip;
        const int rstr = ref_blk_stride, sstr = src_blk_stride;
        const int rem = bw - 16 * strip; // valid bytes of this strip
        const uint32_t m0 = lo_bytes_mask(rem), m1 = lo_bytes_mask(rem - 4), m2 = lo_bytes_mask(rem - 8), m3 = lo_bytes_mask(rem - 12);
#pragma unroll 2
        for (int rr = 0; rr < rows; rr++, gr += rstr, s += sstr) {
            const uintptr_t a  = (uintptr_t)gr;
            const uint32_t *q  = (const uint32_t *)(a & ~(uintptr_t)3);
            const uint32_t  sa = (uint32_t)(a & 3) * 8;
            uint32_t gw[7], w[6];
#pragma unroll
            for (int k = 0; k < 7; k++) gw[k] = __ldg(q + k);
#pragma unroll
            for (int k = 0; k < 6; k++) w[k] = __funnelshift_r(gw[k], gw[k + 1], sa);
            const uint32_t sv[4] = {ldg_u32_unaligned(s) & m0, ldg_u32_unaligned(s + 4) & m1, ldg_u32_unaligned(s + 8) & m2,
                                    ldg_u32_unaligned(s + 12) & m3};
            if (rem >= 16) {
                strip16_sad8(w, sv, acc);
            } else {
#pragma unroll
                for (int p = 0; p < 8; p++) {
                    const int o = p >> 2, shf = (p & 3) * 8;
                    const uint32_t v0 = shf ? __funnelshift_r(w[o], w[o + 1], shf) : w[o];
                    const uint32_t v1 = shf ? __funnelshift_r(w[o + 1], w[o + 2], shf) : w[o + 1];
                    const uint32_t v2 = shf ? __funnelshift_r(w[o + 2], w[o + 3], shf) : w[o + 2];
                    const uint32_t v3 = shf ? __funnelshift_r(w[o + 3], w[o + 4], shf) : w[o + 3];
                    acc[p] += __vsadu4(v0 & m0, sv[0]) + __vsadu4(v1 & m1, sv[1]) + __vsadu4(v2 & m2, sv[2]) + __vsadu4(v3 & m3, sv[3]);
                }
            }
        }
    }
#pragma unroll
    for (int p = 0; p < 8; p++) {
        if (strips >= 2) acc[p] += __shfl_xor_sync(0xffffffffu, acc[p], 1);
        if (strips >= 4) acc[p] += __shfl_xor_sync(0xffffffffu, acc[p], 2);
    }
}

__device__ __forceinline__ unsigned long long octet_best_key(const uint32_t (&acc)[8], int y, int oct, int sa_w) {
    uint32_t best_sad = 0xffffff, best_idx = 0xffffffffu;
#pragma unroll
    for (int p = 0; p < 8; p++) {
        const int x = 8 * oct + p;
        if (x < sa_w && acc[p] < best_sad) {
            best_sad = acc[p];
            best_idx = (uint32_t)(y * sa_w + x);
        }
    }
    return ((unsigned long long)best_sad << 32) | best_idx;
}

// ---- block-parallel search for wide blocks (HME level 1 / 2) -----------------------------------
// Windows here are tiny (8x3 .. 16x16 positions) and blocks large (32x16 .. 64x64 samples), so the
// lanes are spread over the BLOCK instead of over positions: lane = (ypos, band, strip) with
// NY*NB*NS = 32; a lane sums RB rows of its 16-pixel strip for 8 consecutive positions of search row
// y0+ypos, straight from global memory (7 aligned words per row + one funnel shift per word), and
// shfl_xor adds bands and strips. No shared memory, no staging latency chain, all loads of a pass
// in flight at once. Leaders (band 0, strip 0) keep the running best of their search rows.
template <int RB>
struct BlkLane {
    uint32_t src[RB][4];
    uint32_t m[4];     // byte masks of this strip (block narrower than 16*NS)
    int      strip, band, ypos, row0, n_rows;
    bool     full;
};

template <int RB>
__device__ __forceinline__ void blk_lane_init(BlkLane<RB> &L, const uint8_t *src, int src_blk_stride, int bw, int n_rows,
                                              int ns_bits, int nb_bits) {
    const int lane = threadIdx.x & 31;
    L.strip = lane & ((1 << ns_bits) - 1);
    L.band  = (lane >> ns_bits) & ((1 << nb_bits) - 1);
    L.ypos  = lane >> (ns_bits + nb_bits);
    L.row0  = L.band * RB;
    L.n_rows = n_rows;
    const int rem = bw - 16 * L.strip;
    L.full = rem >= 16;
    L.m[0] = lo_bytes_mask(rem); L.m[1] = lo_bytes_mask(rem - 4); L.m[2] = lo_bytes_mask(rem - 8); L.m[3] = lo_bytes_mask(rem - 12);
#pragma unroll
    for (int q = 0; q < RB; q++)
#pragma unroll
        for (int k = 0; k < 4; k++)
            L.src[q][k] = (L.row0 + q < n_rows) ? (ldg_u32_unaligned(src + (size_t)(L.row0 + q) * src_blk_stride + 16 * L.strip + 4 * k) & L.m[k]) : 0u;
}

// acc[p] = SAD of the whole block at positions (x+p, y) for the lanes of that ypos group
template <int RB>
__device__ __forceinline__ void blk_eval_pass(const BlkLane<RB> &L, const uint8_t *ref /* sample aligned with block (0,0) at position x,y */,
                                              int ref_blk_stride, bool active, int ns_bits, int nb_bits, uint32_t (&acc)[8]) {
#pragma unroll
    for (int p = 0; p < 8; p++) acc[p] = 0;
    if (active) {
        const uint8_t *g = ref + (size_t)L.row0 * ref_blk_stride + 16 * L.strip;
        uint32_t gw[RB][7];
        uint32_t sa[RB];
#pragma unroll
        for (int q = 0; q < RB; q++) {
            const uintptr_t a = (uintptr_t)(g + (size_t)q * ref_blk_stride);
            const uint32_t *qp = (const uint32_t *)(a & ~(uintptr_t)3);
            sa[q] = (uint32_t)(a & 3) * 8;
            const bool ok = L.row0 + q < L.n_rows;
#pragma unroll
            for (int k = 0; k < 7; k++) gw[q][k] = ok ? __ldg(qp + k) : 0u;
        }
#pragma unroll
        for (int q = 0; q < RB; q++) {
            uint32_t w[6];
#pragma unroll
            for (int k = 0; k < 6; k++) w[k] = __funnelshift_r(gw[q][k], gw[q][k + 1], sa[q]);
            if (L.full) {
                strip16_sad8(w, L.src[q], acc);
            } else {
#pragma unroll
                for (int p = 0; p < 8; p++) {
                    const int o = p >> 2, shf = (p & 3) * 8;
                    const uint32_t v0 = shf ? __funnelshift_r(w[o], w[o + 1], shf) : w[o];
                    const uint32_t v1 = shf ? __funnelshift_r(w[o + 1], w[o + 2], shf) : w[o + 1];
                    const uint32_t v2 = shf ? __funnelshift_r(w[o + 2], w[o + 3], shf) : w[o + 2];
                    const uint32_t v3 = shf ? __funnelshift_r(w[o + 3], w[o + 4], shf) : w[o + 3];
                    acc[p] += __vsadu4(v0 & L.m[0], L.src[q][0]) + __vsadu4(v1 & L.m[1], L.src[q][1]) +
                        __vsadu4(v2 & L.m[2], L.src[q][2]) + __vsadu4(v3 & L.m[3], L.src[q][3]);
                }
            }
        }
    }
    const int gbits = ns_bits + nb_bits;
#pragma unroll
    for (int p = 0; p < 8; p++) {
        if (gbits >= 1) acc[p] += __shfl_xor_sync(0xffffffffu, acc[p], 1);
        if (gbits >= 2) acc[p] += __shfl_xor_sync(0xffffffffu, acc[p], 2);
        if (gbits >= 3) acc[p] += __shfl_xor_sync(0xffffffffu, acc[p], 4);
        if (gbits >= 4) acc[p] += __shfl_xor_sync(0xffffffffu, acc[p], 8);
        if (gbits >= 5) acc[p] += __shfl_xor_sync(0xffffffffu, acc[p], 16);
    }
}

// whole search of one job by one warp; returns the best key (all lanes)
template <int RB>
__device__ inline unsigned long long warp_search_blk(const uint8_t *src, int src_blk_stride, const uint8_t *ref, int search_stride,
                                                     int ref_blk_stride, int bw, int n_rows, int sa_w, int sa_h, int ns_bits,
                                                     int nb_bits) {
    BlkLane<RB> L;
    blk_lane_init<RB>(L, src, src_blk_stride, bw, n_rows, ns_bits, nb_bits);
    const int ny = 32 >> (ns_bits + nb_bits); // search rows per pass
    const int n_oct = (sa_w + 7) >> 3;
    uint32_t best_sad = 0xffffff, best_idx = 0xffffffffu;
    for (int y0 = 0; y0 < sa_h; y0 += ny) {
        const int  y      = y0 + L.ypos;
        const bool active = y < sa_h;
        for (int oct = 0; oct < n_oct; oct++) {
            uint32_t acc[8];
            blk_eval_pass<RB>(L, ref + (size_t)y * search_stride + 8 * oct, ref_blk_stride, active, ns_bits, nb_bits, acc);
            if (active && L.band == 0 && L.strip == 0) {
#pragma unroll
                for (int p = 0; p < 8; p++) {
                    const int x = 8 * oct + p;
                    if (x < sa_w && acc[p] < best_sad) {
                        best_sad = acc[p];
                        best_idx = (uint32_t)(y * sa_w + x);
                    }
                }
            }
        }
    }
    return warp_min_u64(((unsigned long long)best_sad << 32) | best_idx);
}

// ---- k_search_blk ----------------------------------------------------------------------------
// one warp per job slot; level 1: 1/4 plane (32-wide blocks: 2 strips x 4 bands x 4 search rows),
// level 2: full plane (64-wide blocks: 4 strips x 8 bands x 1 search row)
template <bool SUB>
__global__ void __launch_bounds__(32 * SEARCH_WARPS) k_search_blk(const PicJob *__restrict__ jobs, int level) {
    const PicJob &job = jobs[blockIdx.y];
    const int lane = threadIdx.x & 31;
    const int slot = blockIdx.x * SEARCH_WARPS + (threadIdx.x >> 5);
    if (slot >= job.n_b64 * HME_JOBS_PER_B64) return;
    HmeJobG &j = job.hme_jobs[slot];
    if (!j.active || j.sa_w <= 0 || j.sa_h <= 0) return;
    const int b = slot / HME_JOBS_PER_B64;
    const B64Geo g = b64_geo(job, b);
    const int row_step = SUB ? 2 : 1, shift = level == 1 ? 1 : 0;
    const int bw = g.b64_w >> shift, rows = SUB ? (g.b64_h >> shift) >> 1 : (g.b64_h >> shift);
    const DevPlane &sp = job.src.lvl[level == 1 ? 1 : 0];
    const uint8_t *src = sp.buf + (size_t)(sp.org_y + (g.org_y >> shift)) * sp.stride + sp.org_x + (g.org_x >> shift);
    constexpr int RB = SUB ? 4 : 8;
    const unsigned long long key = level == 1
        ? warp_search_blk<RB>(src, row_step * sp.stride, j.ref, j.ref_stride, row_step * j.ref_stride, bw, rows, j.sa_w, j.sa_h, 1, 2)
        : warp_search_blk<RB>(src, row_step * sp.stride, j.ref, j.ref_stride, row_step * j.ref_stride, bw, rows, j.sa_w, j.sa_h, 2, 3);
    if (lane == 0) j.key = key;
}
