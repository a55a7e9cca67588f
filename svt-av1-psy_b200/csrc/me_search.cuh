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
#include "me_tma.cuh"

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
// lanes are spread over the BLOCK instead of over positions: lane = band * NS + strip; a lane sums NR
// rows of its 16-pixel strip for 8 consecutive positions per pass, from a tile that one TMA load put in
// shared memory (me_tma.cuh); 9 shuffles add the 32 lanes. Tiles are 16 positions x 4 search rows and
// double-buffered per warp, across the quadrant searches of one (b64, list, ref).
#define BLK_IDX_BITS 12 /* y * sa_w + x < 4096; SAD < 2^20 */

// block rows of lane `band` for the HME levels: used row j = band + NB * q, sample row = row_step * j
template <int NS, int NR>
__device__ __forceinline__ void blk_rows(int (&row_of)[NR], int row_step) {
    const int band = (threadIdx.x & 31) / NS;
#pragma unroll
    for (int q = 0; q < NR; q++) row_of[q] = row_step * (band + (32 / NS) * q);
}

// every position of tile (tx, ty) of a window whose position (0,0) is plane sample (px, py); buf = shared address of
// the tile, loaded from (px + 16 tx) & ~15. best: per-lane running key (sad << BLK_IDX_BITS | y * sa_w + x).
template <int NS, int NR>
__device__ __forceinline__ void blk_tile_eval(const TLane<NS, NR> &L, uint32_t buf, int pitch, int px, int tx, int ty, int sa_w,
                                              int sa_h, uint32_t &best) {
    const uint32_t a = (uint32_t)(px + TM_BLK_TILE_W * tx) & 15u;
    const int n_x = imin(TM_BLK_TILE_W, sa_w - TM_BLK_TILE_W * tx), n_y = imin(TM_BLK_TILE_H, sa_h - TM_BLK_TILE_H * ty);
    const int pos = reduce8_pos();
    for (int yy = 0; yy < n_y; yy++) {
        for (int o = 0; 8 * o < n_x; o++) {
            uint32_t acc[8], unused[8];
            tile_pass<NS, NR, false>(L, buf + (uint32_t)(yy * pitch), a + 8u * o, acc, unused);
            const uint32_t sad = reduce8_all(acc);
            const int x = 8 * o + pos;
            if (x < n_x) {
                const uint32_t idx = (uint32_t)((TM_BLK_TILE_H * ty + yy) * sa_w + TM_BLK_TILE_W * tx + x);
                best = min(best, (sad << BLK_IDX_BITS) | idx);
            }
        }
    }
}

__device__ __forceinline__ unsigned long long blk_key64(uint32_t best) {
    best = __reduce_min_sync(0xffffffffu, best);
    return best == 0xffffffffu ? HME_KEY_NONE
                               : (((unsigned long long)(best >> BLK_IDX_BITS) << 32) | (best & ((1u << BLK_IDX_BITS) - 1)));
}

// ---- k_search_blk ----------------------------------------------------------------------------
// one warp per (b64, list, ref): its (up to) 4 quadrant searches run back to back through one tile pipeline.
// level 1: 1/4 plane, 32-wide blocks (NS = 2); level 2: full plane, 64-wide blocks (NS = 4).
template <bool SUB, int LEVEL>
__global__ void __launch_bounds__(32 * SEARCH_WARPS) k_search_blk(const PicJob *__restrict__ jobs) {
    constexpr int NS = LEVEL == 1 ? 2 : 4;
    constexpr int NR = (LEVEL == 1 ? 1 : 4) * (SUB ? 1 : 2);
    constexpr int PITCH = LEVEL == 1 ? TM_L1_BOX_W : TM_L2_BOX_W;
    constexpr int BOX_BYTES = LEVEL == 1 ? TM_L1_BOX_W * TM_L1_BOX_H : TM_L2_BOX_W * TM_L2_BOX_H;
    constexpr int BUF = LEVEL == 1 ? TM_L1_BUF : TM_L2_BUF;
    extern __shared__ __align__(128) uint8_t s_dyn[];
    __shared__ __align__(8) uint64_t s_bar[SEARCH_WARPS][2];
    const PicJob &job = jobs[blockIdx.y];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int u = blockIdx.x * SEARCH_WARPS + warp; // (b64, list, ref)
    if (u >= job.n_b64 * 8) return;
    HmeJobG *jq = job.hme_jobs + (size_t)u * 4;
    // lanes 0..3 hold the four quadrant jobs
    int my_act = 0, my_px = 0, my_py = 0, my_sw = 0, my_sh = 0;
    const int b = u >> 3, l = (u >> 2) & 1, r = u & 3;
    const DevPic   &rp = job.ref[l][r];
    const DevPlane &pl = rp.lvl[LEVEL == 1 ? 1 : 0];
    if (lane < 4) {
        const HmeJobG &j = jq[lane];
        if (j.active && j.sa_w > 0 && j.sa_h > 0) {
            const long long off = j.ref - pl.buf;
            my_act = 1; my_py = (int)(off / pl.stride); my_px = (int)(off - (long long)my_py * pl.stride);
            my_sw = j.sa_w; my_sh = j.sa_h;
        }
    }
    const uint32_t act_mask = __ballot_sync(0xffffffffu, my_act) & 0xfu;
    if (!act_mask) return;
    const CUtensorMap *map = (const CUtensorMap *)rp.tmap + (LEVEL == 1 ? TM_L1 : TM_L2);
    uint8_t *sbase = s_dyn + ((128u - (smem_u32(s_dyn) & 127u)) & 127u);
    TilePipe P;
    pipe_init(P, sbase + (size_t)warp * 2 * BUF, BUF, s_bar[warp], BOX_BYTES);

    const B64Geo g = b64_geo(job, b);
    constexpr int shift = LEVEL == 1 ? 1 : 0;
    const int bw = g.b64_w >> shift, bh = g.b64_h >> shift;
    const DevPlane &sp = job.src.lvl[LEVEL == 1 ? 1 : 0];
    const uint8_t *src = sp.buf + (size_t)(sp.org_y + (g.org_y >> shift)) * sp.stride + sp.org_x + (g.org_x >> shift);
    TLane<NS, NR> L;
    int row_of[NR];
    blk_rows<NS, NR>(row_of, SUB ? 2 : 1);

    int q = __ffs(act_mask) - 1, t = 0;
    int px = __shfl_sync(0xffffffffu, my_px, q), py = __shfl_sync(0xffffffffu, my_py, q);
    int sw = __shfl_sync(0xffffffffu, my_sw, q), sh = __shfl_sync(0xffffffffu, my_sh, q);
    int ntx = (sw + TM_BLK_TILE_W - 1) / TM_BLK_TILE_W, nt = ntx * ((sh + TM_BLK_TILE_H - 1) / TM_BLK_TILE_H);
    pipe_issue(P, 0, map, (px) & ~15, py);
    tlane_init<NS, NR>(L, src, sp.stride, row_of, bw, bh, PITCH); // overlaps the first tile load
    int stage = 0;
    uint32_t best = 0xffffffffu;
    while (q < 4) {
        // the tile after this one: next tile of the job, else first tile of the next active job
        int nq = q, ntile = t + 1, npx = px, npy = py, nsw = sw, nsh = sh, nntx = ntx, nnt = nt;
        if (ntile >= nt) {
            const uint32_t rest = act_mask & ~((2u << q) - 1u);
            nq = rest ? __ffs(rest) - 1 : 4;
            ntile = 0;
            if (nq < 4) {
                npx = __shfl_sync(0xffffffffu, my_px, nq); npy = __shfl_sync(0xffffffffu, my_py, nq);
                nsw = __shfl_sync(0xffffffffu, my_sw, nq); nsh = __shfl_sync(0xffffffffu, my_sh, nq);
                nntx = (nsw + TM_BLK_TILE_W - 1) / TM_BLK_TILE_W; nnt = nntx * ((nsh + TM_BLK_TILE_H - 1) / TM_BLK_TILE_H);
            }
        }
        if (nq < 4) {
            const int ty = ntile / nntx, tx = ntile - ty * nntx;
            pipe_issue(P, stage ^ 1, map, (npx + TM_BLK_TILE_W * tx) & ~15, npy + TM_BLK_TILE_H * ty);
        }
        pipe_wait(P, stage);
        {
            const int ty = t / ntx, tx = t - ty * ntx;
            blk_tile_eval<NS, NR>(L, P.buf[stage], PITCH, px, tx, ty, sw, sh, best);
        }
        __syncwarp();
        if (nq != q) {
            const unsigned long long key = blk_key64(best);
            if (lane == 0) jq[q].key = key;
            best = 0xffffffffu;
        }
        q = nq; t = ntile; px = npx; py = npy; sw = nsw; sh = nsh; ntx = nntx; nt = nnt;
        stage ^= 1;
    }
}
