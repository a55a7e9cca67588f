// me_search.cuh — the HME-stage search kernels of the v2 pipeline: every window search of every
// b64 / reference of a picture is an independent unit of work, one warp each, no CTA barriers.
//
//   k_zz          zero-MV SAD per (b64, ref)                      init_zz_sad        motion_estimation.c:2452
//   k_search16    pre-HME regions and HME level-0 quadrants       prehme_core :1638, hme_level_0 :820
//                 (blocks <= 16 wide on the 1/16 plane): a warp walks the tiles of the windows of one (b64, list,
//                 ref); one TMA load brings a tile (+ the rows/cols the block reaches) into the warp's shared
//                 memory; one of four evaluators, chosen per tile by a cost model, searches it (s16_eval_choice)
//   k_search_blk  HME level 1 / 2 (32- / 64-wide blocks, windows of 8x3 .. 16x16 positions): lanes
//                 spread over the block, TMA tiles (see below)
//
// All follow svt_sad_loop_kernel (compute_sad_c.c:58-101): raster order, strict '<', best starts at
// 0xffffff. Result = 64-bit key (sad << 32 | y*sa_w + x); the minimum over the tiles of a job keeps the
// first minimum in raster order (one warp searches all tiles of a job: a running key in registers, a plain store).
#pragma once
#include <type_traits>
#include "me_device.cuh"
#include "me_tma.cuh"

#define SEARCH_WARPS 4

struct B64Geo {
    int org_x, org_y, b64_w, b64_h;
};
__device__ __forceinline__ B64Geo b64_geo(const PicJob &job, int b) {
    B64Geo g;
    const int by = div_by(b, job.b64_cols, job.inv_cols);
    g.org_x = (b - by * job.b64_cols) * 64;
    g.org_y = by * 64;
    g.b64_w = imin(64, job.pic_w - g.org_x); // motion_estimation.c:3163-3170, picture already 8-aligned
    g.b64_h = imin(64, job.pic_h - g.org_y);
    return g;
}

// Warps are handed only the (b64, list, ref) combinations the picture has: compact unit uc -> b64 and slot
// u = b64 * 8 + list * 4 + ref. (Mapping warps to all 8 slots left half of every CTA idle while it kept its shared memory.)
__device__ __forceinline__ bool compact_unit(const PicJob &job, int uc, int &b, int &l, int &r) {
    const SvtMePicParams &pp = job.pp;
    const int n0 = pp.num_of_ref_pic_to_search[0], n1 = pp.num_of_list_to_search > 1 ? pp.num_of_ref_pic_to_search[1] : 0;
    const int na = n0 + n1;
    if (na <= 0 || uc >= job.n_b64 * na) return false;
    b = div_by(uc, na, job.inv_refs);
    const int k = uc - b * na;
    l = k >= n0 ? 1 : 0;
    r = l ? k - n0 : k;
    return true;
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
    int b, l, r;
    if (!compact_unit(job, blockIdx.x * SEARCH_WARPS + (threadIdx.x >> 5), b, l, r)) return;
    const SvtMePicParams &pp = job.pp;
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

// ---- 1/16-plane searches (blocks <= 16 wide): tiles through TMA ---------------------------------
// One tile = TM_S16T (16 positions x 96 rows) or TM_S16W (128 x 18 / 16 / 12: s16_tile_geometry) of a window, fetched from the 16-byte boundary at
// or before its first position. src: n_rows x 16 bytes in shared memory, bytes >= bw zeroed.
//
// A window may be the union of the four quadrant windows of an HME level-0 search (k_hme_ctl merges them when they
// form an exact 2 x 2 grid: one unit instead of four, lanes filled). Every routine therefore keeps FOUR running bests,
// one per quadrant: quadrant = (x >= split_x) + 2 * (y >= split_y) in window coordinates; an ordinary window has its
// splits beyond its size. Keys: (sad << 16 | yy << 8 | xx) with tile-local coordinates (sad <= 16*16*255, yy < 128,
// xx < 128); min() = first minimum in raster order, inside a quadrant too.
struct S16Split {
    int x, y; // window coordinates of the first position of the right / bottom quadrants
};

__device__ __forceinline__ void s16_merge_best(uint32_t (&best)[4], uint32_t key, int qd) {
#pragma unroll
    for (int k = 0; k < 4; k++) best[k] = (qd == k) ? min(best[k], key) : best[k];
}
// a lane walks its tasks in raster order, so the quadrant changes a few times per tile at most: the running best of
// the current quadrant lives in one register and is folded into best[] only on a change
struct S16Run {
    uint32_t key;
    int      qd;
};
__device__ __forceinline__ void s16_run_update(S16Run &run, uint32_t (&best)[4], uint32_t key, int qd) {
    if (qd != run.qd) {
        s16_merge_best(best, run.key, run.qd);
        run.key = 0xffffffffu;
        run.qd  = qd;
    }
    run.key = min(run.key, key);
}

// x / n for 1 <= n <= 16 and 0 <= x < 4096 without the ~20-instruction integer division (n = octets of a tile row):
// ceil(2^16 / n) * x >> 16 is exact while x * (ceil(2^16/n) - 2^16/n) < 2^16 / n, i.e. for x < 2^16 / n
__constant__ uint16_t c_inv16[17] = {0, 0, 32768, 21846, 16384, 13108, 10923, 9363, 8192, 7282, 6554, 5958, 5462, 5042, 4682, 4370, 4096};
__device__ __forceinline__ int div_small(int x, int n) { return n == 1 ? x : (int)(((uint32_t)x * c_inv16[n]) >> 16); }

// lane = (search row, octet of 8 positions); a lane reads 28 bytes per block row from the 4-byte boundary at or before
// its octet (7 LDS.32, consecutive lanes 8 bytes apart); the remaining 0..3 bytes of misalignment are the same for
// every octet of the job: one funnel shift per word unless ALIGNED (no misalignment, nearly every window). PARTIAL: blocks
// narrower than 16 (right picture edge). Those and the unaligned windows are kept out of line (s16_tile_tasks_slow) so
// that the code every other tile runs stays small. The row
// loops of this and the per-position evaluator are NOT unrolled: the pre-HME launch was stalling on instruction fetch
// (ncu: 1.2 warps per issue cycle in no_instruction); 2928 -> 1936 SASS instructions, 37.4 -> 35.2 us per picture.
template <bool PARTIAL, bool ALIGNED>
__device__ __forceinline__ void s16_tile_tasks(uint32_t buf, int pitch, uint32_t src, uint32_t a4, uint32_t t8, int x0, int y0, int tw,
                                               int th, int bw, int n_rows, int row_step, bool skip_lines, S16Split sp,
                                               uint32_t (&best)[4]) {
    const int lane  = threadIdx.x & 31;
    const int n_oct = (tw + 7) >> 3;
    constexpr bool partial = PARTIAL;
    uint32_t m0 = 0xffffffffu, m1 = 0xffffffffu, m2 = 0xffffffffu, m3 = 0xffffffffu;
    if (partial) { m0 = lo_bytes_mask(bw); m1 = lo_bytes_mask(bw - 4); m2 = lo_bytes_mask(bw - 8); m3 = lo_bytes_mask(bw - 12); }
    int yy = div_small(lane, n_oct), oct = lane - yy * n_oct;
    const int dyy = div_small(32, n_oct), doct = 32 - dyy * n_oct;
    const uint32_t rstep = (uint32_t)(row_step * pitch);
    S16Run run = {0xffffffffu, 0};
    for (; yy < th; yy += dyy, oct += doct) {
        if (oct >= n_oct) { oct -= n_oct; yy++; if (yy >= th) break; }
        const int y = y0 + yy;
        if (skip_lines && (y & 1) == 0) continue;
        uint32_t acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        uint32_t row = buf + (uint32_t)(yy * pitch) + a4 + 8u * (uint32_t)oct; // a4: the window's byte offset in the tile rows, rounded down to a word
#pragma unroll 1
        for (int rr = 0; rr < n_rows; rr++, row += rstep) {
            uint32_t w7[7], w[6];
#pragma unroll
            for (int i = 0; i < 7; i++) asm volatile("ld.shared.u32 %0, [%1];" : "=r"(w7[i]) : "r"(row + 4u * i));
            const uint4 sv = lds128(src + 16u * rr);
            if (!ALIGNED) { // see s16_tile_tasks_vb: nearly every window starts on a word boundary
#pragma unroll
                for (int k = 0; k < 6; k++) w7[k] = __funnelshift_r(w7[k], w7[k + 1], t8);
            }
#pragma unroll
            for (int k = 0; k < 6; k++) w[k] = w7[k];
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
                    acc[p] = sad4_acc(v0 & m0, s[0], acc[p]);
                    acc[p] = sad4_acc(v1 & m1, s[1], acc[p]);
                    acc[p] = sad4_acc(v2 & m2, s[2], acc[p]);
                    acc[p] = sad4_acc(v3 & m3, s[3], acc[p]);
                }
            }
        }
        uint32_t tb = 0xffffffffu;
        const uint32_t idx = ((uint32_t)yy << 8) | (uint32_t)(8 * oct);
        if (8 * oct + 8 <= tw) {
#pragma unroll
            for (int p = 0; p < 8; p++) tb = min(tb, (acc[p] << 16) + idx + p);
        } else {
#pragma unroll
            for (int p = 0; p < 8; p++)
                if (8 * oct + p < tw) tb = min(tb, (acc[p] << 16) + idx + p);
        }
        s16_run_update(run, best, tb, (x0 + 8 * oct >= sp.x ? 1 : 0) + (y >= sp.y ? 2 : 0));
    }
    s16_merge_best(best, run.key, run.qd);
}

__device__ __noinline__ uint4 s16_tile_tasks_slow(uint32_t buf, int pitch, uint32_t src, uint32_t a4, uint32_t t8, int x0, int y0, int tw, int th,
                                                    int bw, int n_rows, int row_step, bool skip_lines, int spx, int spy) {
    uint32_t best[4] = {0xffffffffu, 0xffffffffu, 0xffffffffu, 0xffffffffu};
    const S16Split sp = {spx, spy};
    if (bw < 16) s16_tile_tasks<true, false>(buf, pitch, src, a4, t8, x0, y0, tw, th, bw, n_rows, row_step, skip_lines, sp, best);
    else s16_tile_tasks<false, false>(buf, pitch, src, a4, t8, x0, y0, tw, th, bw, n_rows, row_step, skip_lines, sp, best);
    return make_uint4(best[0], best[1], best[2], best[3]);
}

// Vertically blocked form for the common shape (16x16 block, rows two apart: 8 block rows): a lane takes G search
// rows of the same parity (yy0, yy0+2, ...) of one octet. Search rows two apart read the same reference rows against
// different source rows, so each reference row is loaded and funnel-shifted ONCE (7 LDS + 21 SHF) and feeds up to G
// x 32 VABSDIFF4: 21 SHF per 96 SADs instead of per 32. par0 / npar: first parity and number of parities searched
// (skip_search_line searches odd rows only).
// G = 3: blocks of 6 search rows. Lanes then sit 6 rows apart (72 words at the 48-byte pitch of the tall tile = 8
// banks): 4-way conflicts on the LDS.32, 28 cycles per reference row against ~58 of ALU work. With G = 4 lanes sit 8
// rows apart = a multiple of 32 words for ANY pitch the TMA allows (multiples of 16 bytes): 16-way conflicts, measured
// 30 % slower than no blocking at all.
#define S16_VB_G 3
template <int G, bool ALIGNED>
__device__ __forceinline__ void s16_tile_tasks_vb(uint32_t buf, int pitch, uint32_t src, uint32_t a, int x0, int y0, int tw, int th,
                                                  int par0, int npar, S16Split sp, uint32_t (&best)[4]) {
    const int lane  = threadIdx.x & 31;
    const int n_oct = (tw + 7) >> 3;
    const int n_tasks = n_oct * ((th + 2 * G - 1) / (2 * G)) * npar;
    const uint32_t t8 = (a & 3u) * 8u;
    S16Run run = {0xffffffffu, 0};
    for (int t = lane; t < n_tasks; t += 32) {
        const int ch = div_small(t, n_oct), oct = t - ch * n_oct;
        const int yy0 = npar == 2 ? (ch >> 1) * 2 * G + (ch & 1) : ch * 2 * G + par0;
        if (yy0 >= th) continue;
        uint32_t acc[G][8];
#pragma unroll
        for (int g = 0; g < G; g++)
#pragma unroll
            for (int p = 0; p < 8; p++) acc[g][p] = 0;
        // word address of the octet's first byte (LDS.32: lanes 8 bytes apart, 2-way bank conflict, far below the ALU time)
        const uint32_t row0 = buf + (uint32_t)(yy0 * pitch) + ((a + 8u * (uint32_t)oct) & ~3u);
        // real loops over the reference rows (the source row each one meets is read from shared memory, a broadcast):
        // unrolled it is ~1100 instructions and the kernel stalls on instruction fetch. Reference row k meets block row
        // k - g of search row yy0 + 2g: rows G-1 .. 7 meet ALL G search rows (steady loop, no conditions), the first
        // and last G-1 rows only some of them (one short loop with the range checks).
        auto ref_row = [&](int k, auto all_rows) {
            const uint32_t row = row0 + (uint32_t)(2 * k * pitch);
            uint32_t w[7], s1[5], s2[5], s3[5];
#pragma unroll
            for (int i = 0; i < 7; i++) asm volatile("ld.shared.u32 %0, [%1];" : "=r"(w[i]) : "r"(row + 4u * i));
            // windows start on a word boundary of the plane unless the picture edge clamped them (block origins are multiples
            // of 16 samples, window origins multiples of 4): ALIGNED has no alignment shifts (6 of the 21 per reference row;
            // written as "if (t8)" ptxas predicates them and they still issue: no gain; without them pre-HME 34.4 -> 32.2 us,
            // level 0 15.1 -> 14.4). The other windows take the same code with the shifts, out of line (s16_vb_unaligned).
            if (!ALIGNED) {
#pragma unroll
                for (int i = 0; i < 6; i++) w[i] = __funnelshift_r(w[i], w[i + 1], t8);
            }
#pragma unroll
            for (int i = 0; i < 5; i++) {
                s1[i] = fsr_bytes<1>(w[i], w[i + 1]);
                s2[i] = fsr_bytes<2>(w[i], w[i + 1]);
                s3[i] = fsr_bytes<3>(w[i], w[i + 1]);
            }
#pragma unroll
            for (int g = 0; g < G; g++) {
                const int rr = k - g; // warp-uniform
                if (decltype(all_rows)::value || (unsigned)rr < 8u) {
                    const uint4 sv = lds128(src + 16u * (uint32_t)rr);
                    const uint32_t s[4] = {sv.x, sv.y, sv.z, sv.w};
#pragma unroll
                    for (int i = 0; i < 4; i++) {
                        acc[g][0] = sad4_acc(w[i], s[i], acc[g][0]);
                        acc[g][4] = sad4_acc(w[i + 1], s[i], acc[g][4]);
                        acc[g][1] = sad4_acc(s1[i], s[i], acc[g][1]);
                        acc[g][5] = sad4_acc(s1[i + 1], s[i], acc[g][5]);
                        acc[g][2] = sad4_acc(s2[i], s[i], acc[g][2]);
                        acc[g][6] = sad4_acc(s2[i + 1], s[i], acc[g][6]);
                        acc[g][3] = sad4_acc(s3[i], s[i], acc[g][3]);
                        acc[g][7] = sad4_acc(s3[i + 1], s[i], acc[g][7]);
                    }
                }
            }
        };
#pragma unroll 1
        for (int j = 0; j < 2 * (G - 1); j++) ref_row(j < G - 1 ? j : j + 9 - G, std::false_type{});
#pragma unroll 1
        for (int k = G - 1; k < 8; k++) ref_row(k, std::true_type{});
        const int right = x0 + 8 * oct >= sp.x ? 1 : 0;
#pragma unroll
        for (int g = 0; g < G; g++) {
            const int yy = yy0 + 2 * g;
            if (yy < th) {
                const uint32_t idx = ((uint32_t)yy << 8) | (uint32_t)(8 * oct);
                uint32_t tb = 0xffffffffu;
                if (8 * oct + 8 <= tw) { // all but the last octet of a row: no per-position range checks
#pragma unroll
                    for (int p = 0; p < 8; p++) tb = min(tb, (acc[g][p] << 16) + idx + p);
                } else {
#pragma unroll
                    for (int p = 0; p < 8; p++)
                        if (8 * oct + p < tw) tb = min(tb, (acc[g][p] << 16) + idx + p);
                }
                s16_run_update(run, best, tb, right + (y0 + yy >= sp.y ? 2 : 0));
            }
        }
    }
    s16_merge_best(best, run.key, run.qd);
}

template <bool VB2>
__device__ __noinline__ uint4 s16_vb_unaligned(int g, uint32_t buf, int pitch, uint32_t src, uint32_t a, int x0, int y0, int tw, int th, int par0,
                                               int npar, int spx, int spy) {
    uint32_t best[4] = {0xffffffffu, 0xffffffffu, 0xffffffffu, 0xffffffffu};
    const S16Split sp = {spx, spy};
    if (VB2 && g == 2) s16_tile_tasks_vb<2, false>(buf, pitch, src, a, x0, y0, tw, th, par0, npar, sp, best);
    else s16_tile_tasks_vb<3, false>(buf, pitch, src, a, x0, y0, tw, th, par0, npar, sp, best);
    return make_uint4(best[0], best[1], best[2], best[3]);
}

// Tiles with few positions (the 4 rows a 100-row pre-HME window leaves after its 96-row tile, windows cut by the picture
// edge): lane = ONE position. The octet evaluator would run its 8 block rows with 4 of 32 lanes busy (~490 warp
// instructions for 32 positions); here a lane reads the 5 words its position touches per block row, shifts them by its own
// byte offset and issues 4 VABSDIFF4: ~110 warp instructions per 32 positions.
__device__ __forceinline__ void s16_tile_tasks_pos(uint32_t buf, int pitch, uint32_t src, uint32_t a, int x0, int y0, int tw, int th,
                                                   int n_rows, int row_step, bool skip_lines, S16Split sp, uint32_t (&best)[4]) {
    const int lane = threadIdx.x & 31;
    S16Run run = {0xffffffffu, 0};
    for (int t = lane; t < tw * th; t += 32) {
        const int yy = div_small(t, tw), x = t - yy * tw;
        if (skip_lines && ((y0 + yy) & 1) == 0) continue;
        const uint32_t off = a + (uint32_t)x, sh = (off & 3u) * 8u;
        uint32_t row = buf + (uint32_t)(yy * pitch) + (off & ~3u);
        const uint32_t rstep = (uint32_t)(row_step * pitch);
        uint32_t acc = 0;
#pragma unroll 1
        for (int rr = 0; rr < n_rows; rr++, row += rstep) {
            uint32_t c[5];
#pragma unroll
            for (int i = 0; i < 5; i++) asm volatile("ld.shared.u32 %0, [%1];" : "=r"(c[i]) : "r"(row + 4u * i));
            const uint4 sv = lds128(src + 16u * (uint32_t)rr);
            acc = sad4_acc(__funnelshift_r(c[0], c[1], sh), sv.x, acc);
            acc = sad4_acc(__funnelshift_r(c[1], c[2], sh), sv.y, acc);
            acc = sad4_acc(__funnelshift_r(c[2], c[3], sh), sv.z, acc);
            acc = sad4_acc(__funnelshift_r(c[3], c[4], sh), sv.w, acc);
        }
        s16_run_update(run, best, (acc << 16) + ((uint32_t)yy << 8) + (uint32_t)x, (x0 + x >= sp.x ? 1 : 0) + (y0 + yy >= sp.y ? 2 : 0));
    }
    s16_merge_best(best, run.key, run.qd);
}

// Which evaluator searches a tile of tw x th positions: the one with the fewest warp instructions = passes over the 32
// lanes x instructions per pass (measured loop bodies: a reference row costs 28 load / shift instructions, a block row of
// 8 positions 33).
//   general : task = (search row, octet)                   8 x 61 per pass
//   vb<G>   : task = (G search rows of one parity, octet)  (7 + G) x 28 + 8 G x 33 per pass
//   pos     : task = one position                          8 x 14 per pass
// Shared with the tile geometry (k_hme_ctl picks the tile height of a window by the same model).
enum { S16_EVAL_GEN = 0, S16_EVAL_VB2 = 1, S16_EVAL_VB3 = 2, S16_EVAL_POS = 3 };
__host__ __device__ inline int s16_eval_choice(int tw, int th, int bw, int n_rows, int row_step, bool skip_lines, bool wide_pitch, bool vb2,
                                               int *cost_out) {
    const int n_oct = (tw + 7) >> 3, npar = skip_lines ? 1 : 2, rows_eff = skip_lines ? (th + 1) >> 1 : th;
    int cost = ((n_oct * rows_eff + 31) >> 5) * 500, which = S16_EVAL_GEN;
    if (bw == 16 && tw <= 16) {
        const int c = ((tw * th + 31) >> 5) * (14 * n_rows + 10);
        if (c <= cost) { cost = c; which = S16_EVAL_POS; }
    }
    if (bw == 16 && n_rows == 8 && row_step == 2) {
        // G = 2 only in the wide tile (the 48-byte pitch of the tall one puts lanes 4 rows apart on 4 banks) and not for the 3-row
        // pre-HME strips (measured: one pass with half-empty chains 39.8 us per picture, two general passes 36.9)
        if (vb2 && wide_pitch && th >= 8) {
            const int c = ((n_oct * ((th + 3) >> 2) * npar + 31) >> 5) * 840;
            if (c <= cost) { cost = c; which = S16_EVAL_VB2; }
        }
        const int c = ((n_oct * ((th + 5) / 6) * npar + 31) >> 5) * 1140;
        if (c <= cost) { cost = c; which = S16_EVAL_VB3; }
    }
    if (cost_out) *cost_out = cost;
    return which;
}

// a: byte offset of the tile's position 0 inside its rows ((plane x of the position) & 15). On return best[k] is the
// warp-wide best of quadrant k in this tile, or 0xffffffff.
template <bool VB2>
__device__ inline void warp_search16_tile(uint32_t buf, int pitch, uint32_t src, uint32_t a, int x0, int y0, int tw, int th, int bw,
                                          int n_rows, int row_step, bool skip_lines, S16Split sp, uint32_t (&best)[4]) {
    const bool wide_pitch = pitch != TM_S16T_BOX_W;
#pragma unroll
    for (int k = 0; k < 4; k++) best[k] = 0xffffffffu;
    const int which = s16_eval_choice(tw, th, bw, n_rows, row_step, skip_lines, wide_pitch, VB2, nullptr);
    const int par0 = skip_lines ? ((y0 & 1) ? 0 : 1) : 0; // skip_search_line drops the even rows of the WINDOW: parity is relative to y0
    const bool aligned = (a & 3u) == 0u;
    if (which == S16_EVAL_VB3 || (VB2 && which == S16_EVAL_VB2)) {
        if (!aligned) {
            const uint4 b4 = s16_vb_unaligned<VB2>(which == S16_EVAL_VB3 ? 3 : 2, buf, pitch, src, a, x0, y0, tw, th, par0, skip_lines ? 1 : 2, sp.x, sp.y);
            best[0] = b4.x; best[1] = b4.y; best[2] = b4.z; best[3] = b4.w;
        } else if (!VB2 || which == S16_EVAL_VB3) {
            s16_tile_tasks_vb<3, true>(buf, pitch, src, a, x0, y0, tw, th, par0, skip_lines ? 1 : 2, sp, best);
        } else {
            s16_tile_tasks_vb<2, true>(buf, pitch, src, a, x0, y0, tw, th, par0, skip_lines ? 1 : 2, sp, best);
        }
    } else if (which == S16_EVAL_POS) {
        s16_tile_tasks_pos(buf, pitch, src, a, x0, y0, tw, th, n_rows, row_step, skip_lines, sp, best);
    } else if (bw < 16 || !aligned) {
        const uint4 b4 = s16_tile_tasks_slow(buf, pitch, src, a & ~3u, (a & 3u) * 8u, x0, y0, tw, th, bw, n_rows, row_step, skip_lines, sp.x, sp.y);
        best[0] = b4.x; best[1] = b4.y; best[2] = b4.z; best[3] = b4.w;
    } else {
        s16_tile_tasks<false, true>(buf, pitch, src, a, 0u, x0, y0, tw, th, bw, n_rows, row_step, skip_lines, sp, best);
    }
#pragma unroll
    for (int k = 0; k < 4; k++) best[k] = __reduce_min_sync(0xffffffffu, best[k]);
}

// tile-local key of quadrant qd -> the job's 64-bit key (sad << 32 | index inside the quadrant's own window of width wq)
__device__ __forceinline__ unsigned long long s16_key64(uint32_t key, int x0, int y0, int qd, S16Split sp, int wq) {
    const int x = x0 + (int)(key & 0xffu) - ((qd & 1) ? sp.x : 0), y = y0 + (int)((key >> 8) & 0xffu) - ((qd & 2) ? sp.y : 0);
    return ((unsigned long long)(key >> 16) << 32) | (uint32_t)(y * wq + x);
}

// source block rows of the 1/16 plane -> shared memory (n_rows x 16 bytes, bytes >= bw zeroed); block origins are 16-byte aligned
__device__ __forceinline__ void warp_stage_src16(uint32_t dst, const uint8_t *src, int stride, int row_step, int n_rows, int bw) {
    const int lane = threadIdx.x & 31;
    if (lane < n_rows) {
        uint4 v = __ldg((const uint4 *)(src + (size_t)lane * row_step * stride));
        if (bw < 16) { v.x &= lo_bytes_mask(bw); v.y &= lo_bytes_mask(bw - 4); v.z &= lo_bytes_mask(bw - 8); v.w &= lo_bytes_mask(bw - 12); }
        asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(dst + 16u * lane), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
    }
}

// ---- k_search16 ------------------------------------------------------------------------------
// Persistent warps pull UNITS from the queue k_hme_ctl filled. A unit = one (picture, b64, list, ref): the jobs of its
// four slots (two pre-HME regions, or four level-0 quadrants / their merged window) are searched by this warp tile after
// tile. Per unit, once: the queue ticket, the unit record, the source block rows; per job, once: the job record and the
// store of its best position (kept in registers over its tiles: no atomics); per tile: one TMA load, in flight while the
// previous tile is searched, and the search. (Round 1 made every tile a queue unit: ~650 of the ~1590 instructions a
// unit executed were this per-unit work, profiles/r02_k_search16_units.md.)
// Where a warp is in its walk (unit -> job -> tile), packed: the cursor is alive across every tile search and the
// evaluators need the registers.
struct S16Cursor {
    int      pic, slot0, src_off;       // unit record
    uint32_t tile;                      // tile inside the current job: tx | ty << 8 | index << 16 (stepped, never divided)
    uint32_t u;                         // bw | rows << 8 | remaining-jobs mask << 16 | current job q << 20 | new_unit << 24 | valid << 25
    uint32_t pxy, wh, split, misc;      // job: px | py << 16; sa_w | sa_h << 16; w0 | h0 << 16; n_tiles | n_tx << 16 | wide << 24 | merged << 25 | tile_h / 2 << 26
};
#define S16C_VALID(c) (((c).u >> 25) & 1u)
#define S16C_NEW_UNIT(c) (((c).u >> 24) & 1u)

__device__ __forceinline__ void s16_load_job(const PicJob *__restrict__ jobs, S16Cursor &c) {
    const HmeJobG &j = jobs[c.pic].hme_jobs[c.slot0 + (int)((c.u >> 20) & 3u)];
    int sa_w = j.sa_w, sa_h = j.sa_h, w0 = 32767, h0 = 32767, merged = 0;
    if (j.merged == 1) { merged = 1; w0 = sa_w; h0 = sa_h; sa_w = j.m_w; sa_h = j.m_h; }
    c.pxy   = (uint32_t)(uint16_t)j.px | ((uint32_t)(uint16_t)j.py << 16);
    c.wh    = (uint32_t)sa_w | ((uint32_t)sa_h << 16);
    c.split = (uint32_t)w0 | ((uint32_t)h0 << 16);
    c.misc  = ((uint32_t)j.n_tx * (uint32_t)j.n_ty) | ((uint32_t)j.n_tx << 16) | ((uint32_t)j.wide << 24) | ((uint32_t)merged << 25) |
              ((uint32_t)(j.tile_h >> 1) << 26); // tile heights are even (96; 18 / 16 / 12)
    c.tile  = 0;
}
// next unit of the queue: ticket, record, which of its four jobs have tiles
__device__ __forceinline__ void s16_next_unit(const PicJob *__restrict__ jobs, uint32_t n_units, S16Cursor &c) {
    const int lane = threadIdx.x & 31;
    uint32_t ui = 0;
    if (lane == 0) ui = atomicAdd(&jobs[0].s16_counters[1], 1u);
    ui = __shfl_sync(0xffffffffu, ui, 0);
    c.u = 0;
    if (ui >= n_units) return;
    const uint2 rec = jobs[0].s16_units[ui];
    const bool single = (rec.y >> 31) != 0; // the unit is ONE job (small batches), not the four slots of a (list, ref)
    const int  slot   = (int)(rec.y & 0x7fffffffu);
    c.pic = (int)rec.x; c.slot0 = slot & ~3;
    bool has = false;
    if (lane < 4 && (!single || lane == (slot & 3))) {
        const HmeJobG &j = jobs[c.pic].hme_jobs[c.slot0 + lane];
        has = j.active && j.merged != 2 && j.n_tx && j.n_ty && j.sa_w > 0 && j.sa_h > 0;
    }
    uint32_t todo = __ballot_sync(0xffffffffu, has) & 0xfu;
    if (!todo) return; // cannot happen: k_hme_ctl enqueues only groups with work
    const uint32_t q = (uint32_t)__ffs(todo) - 1u;
    todo &= todo - 1u;
    const HmeJobG &j0 = jobs[c.pic].hme_jobs[c.slot0 + (int)q];
    c.src_off = j0.src_off; // the jobs of a unit compare the same block
    c.u = (uint32_t)j0.bw | ((uint32_t)j0.rows << 8) | (todo << 16) | (q << 20) | (1u << 24) | (1u << 25);
    s16_load_job(jobs, c);
}
// step to the next tile: same job, next job of the unit, or next unit
__device__ __forceinline__ void s16_advance(const PicJob *__restrict__ jobs, uint32_t n_units, S16Cursor &c) {
    c.u &= ~(1u << 24);
    uint32_t t = c.tile + 0x10001u; // next index, next column
    if ((t & 0xffu) == ((c.misc >> 16) & 0xffu)) t = (t & ~0xffu) + 0x100u; // first column of the next tile row
    c.tile = t;
    if ((t >> 16) < (c.misc & 0xffffu)) return;
    const uint32_t todo = (c.u >> 16) & 0xfu;
    if (todo) {
        const uint32_t q = (uint32_t)__ffs(todo) - 1u;
        c.u = (c.u & ~(0xffu << 16)) | ((todo & (todo - 1u)) << 16) | (q << 20);
        s16_load_job(jobs, c);
        return;
    }
    s16_next_unit(jobs, n_units, c);
}
struct S16Tile { // tile coordinates inside the job's window
    int x0, y0, tw, th;
};
__device__ __forceinline__ S16Tile s16_tile_of(const S16Cursor &c) {
    S16Tile t;
    const int n_tx = (int)((c.misc >> 16) & 0xffu), wide = (int)((c.misc >> 24) & 1u);
    const int tile_w = wide ? S16W_TILE_W : S16T_TILE_W, tile_h = (int)(c.misc >> 26) << 1;
    const int tx = (int)(c.tile & 0xffu), ty = (int)((c.tile >> 8) & 0xffu);
    t.x0 = tx * tile_w; t.y0 = ty * tile_h;
    t.tw = imin(tile_w, (int)(c.wh & 0xffffu) - t.x0); t.th = imin(tile_h, (int)(c.wh >> 16) - t.y0);
    return t;
}

// VB2: with the two-row blocked evaluator (the level-0 launch: windows of 16 .. 96 rows; the pre-HME launch leaves it out,
// its code only costs instruction-cache space there: 36.9 -> 38.5 us per picture with it compiled in and never chosen)
template <bool VB2>
__global__ void __launch_bounds__(32 * SEARCH_WARPS, 5) k_search16(const PicJob *__restrict__ jobs, int skip_cfg, int row_step) {
    extern __shared__ __align__(128) uint8_t s_dyn[];
    __shared__ __align__(16) uint32_t s_src[SEARCH_WARPS][2][16 * 4];
    __shared__ __align__(8) uint64_t s_bar[SEARCH_WARPS][2];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const PicJob &job0 = jobs[0];
    const uint32_t n_units = min(job0.s16_counters[0], job0.s16_units_cap);
    if (blockIdx.x == 0 && threadIdx.x == 0 && job0.s16_counters[0] > job0.s16_units_cap) job0.fp_counters[8] = 1; // queue overflow: reported by the host
    if (n_units == 0) return;
    // row_step (2: SAD over every other block row, hme_search_method == SUB_SAD_SEARCH) is a launch argument: read from the job
    // record it was re-read from global memory for every tile
    uint8_t *sbase = s_dyn + ((128u - (smem_u32(s_dyn) & 127u)) & 127u);
    TilePipe P;
    pipe_init(P, sbase + (size_t)warp * 2 * TM_S16_BUF, TM_S16_BUF, s_bar[warp], 0);
    const uint32_t src32 = smem_u32(s_src[warp][0]); // source rows of the unit with parity sb at + 256 sb bytes
    S16Cursor c;
    s16_next_unit(jobs, n_units, c);
    if (!S16C_VALID(c)) return;
    // issue: the tile load into stage `st`; the source rows of a NEW unit's block go to source buffer `sb`
    auto issue = [&](const S16Cursor &cc, const S16Tile &t, int st, int sb) {
        const PicJob &job = jobs[cc.pic];
        const int lr = (cc.slot0 & (HME_JOBS_PER_B64 - 1)) >> 2;
        const int wide = (int)((cc.misc >> 24) & 1u);
        const CUtensorMap *map = (const CUtensorMap *)job.ref[lr >> 2][lr & 3].tmap + (wide ? TM_S16W : TM_S16T);
        const int px = (int)(int16_t)(cc.pxy & 0xffffu), py = (int)(int16_t)(cc.pxy >> 16);
        pipe_issue_bytes(P, st, map, (px + t.x0) & ~15, py + t.y0, wide ? TM_S16W_BOX_W * TM_S16W_BOX_H : TM_S16T_BOX_W * TM_S16T_BOX_H);
        if (S16C_NEW_UNIT(cc)) {
            const DevPlane &ss = job.src.lvl[2];
            warp_stage_src16(src32 + 256u * (uint32_t)sb, ss.buf + cc.src_off, ss.stride, row_step, (int)((cc.u >> 8) & 0xffu), (int)(cc.u & 0xffu));
        }
    };
    int stage = 0, sbuf = 0; // sbuf: source buffer of the unit being searched
    issue(c, s16_tile_of(c), stage, sbuf);
    unsigned long long run_key = HME_KEY_NONE; // lanes 0..3: best of quadrant `lane` of the current job over its tiles so far
    while (true) {
        // everything the search of the current tile needs, before the cursor moves on
        const S16Tile t = s16_tile_of(c);
        const int bw = (int)(c.u & 0xffu), rows = (int)((c.u >> 8) & 0xffu), wide = (int)((c.misc >> 24) & 1u), merged = (int)((c.misc >> 25) & 1u);
        const int px = (int)(int16_t)(c.pxy & 0xffffu), sa_w = (int)(c.wh & 0xffffu);
        const S16Split sp = {(int)(c.split & 0xffffu), (int)(c.split >> 16)};
        const bool last_of_job = (c.tile >> 16) + 1u == (c.misc & 0xffffu);
        unsigned long long *key_out = &jobs[c.pic].hme_jobs[c.slot0 + (int)((c.u >> 20) & 3u) + lane].key;
        const int cur_sbuf = sbuf;
        // the tile after this one: its load overlaps this tile's search
        s16_advance(jobs, n_units, c);
        if (S16C_VALID(c)) {
            if (S16C_NEW_UNIT(c)) sbuf ^= 1;
            issue(c, s16_tile_of(c), stage ^ 1, sbuf);
        }
        const bool skip = skip_cfg && bw == 16 && rows <= 16;
        pipe_wait(P, stage);
        __syncwarp(); // the source rows written by other lanes
        uint32_t best[4];
        warp_search16_tile<VB2>(P.buf[stage], wide ? TM_S16W_BOX_W : TM_S16T_BOX_W, src32 + 256u * (uint32_t)cur_sbuf,
                           (uint32_t)(px + t.x0) & 15u, t.x0, t.y0, t.tw, t.th, bw, rows, row_step, skip, sp, best);
        if (lane < 4 && (lane == 0 || merged)) {
            const uint32_t k = lane == 0 ? best[0] : (lane == 1 ? best[1] : (lane == 2 ? best[2] : best[3]));
            if (k != 0xffffffffu) {
                const int wq = merged ? ((lane & 1) ? sa_w - sp.x : sp.x) : sa_w;
                const unsigned long long k64 = s16_key64(k, t.x0, t.y0, lane, sp, wq);
                run_key = k64 < run_key ? k64 : run_key;
            }
            if (last_of_job) { // this warp alone searched the job: plain store
                *key_out = run_key;
                run_key = HME_KEY_NONE;
            }
        }
        __syncwarp();
        if (!S16C_VALID(c)) break;
        stage ^= 1;
    }
}

// ---- block-parallel search for wide blocks (HME level 1 / 2) -----------------------------------
// Windows here are tiny (8x3 .. 16x16 positions) and blocks large (32x16 .. 64x64 samples), so the
// lanes are spread over the BLOCK instead of over positions: lane = band * NS + strip; a lane sums NR
// rows of its 16-pixel strip for 8 consecutive positions per pass, from a tile that one TMA load put in
// shared memory (me_tma.cuh); REDUX.SUM adds the 32 lanes. Tiles are 16 positions x 4 search rows and
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
    for (int yy = 0; yy < n_y; yy++) {
        for (int o = 0; 8 * o < n_x; o++) {
            uint32_t acc[8], unused[8];
            tile_pass<NS, NR, false>(L, buf + (uint32_t)(yy * pitch), a + 8u * o, acc, unused);
            // warp-wide sums on the reduction unit (REDUX.SUM), one per position, instead of a shuffle tree on the
            // ALU pipe that bounds this kernel; every lane then keeps the running minimum (the values are uniform)
            const uint32_t idx0 = (uint32_t)((TM_BLK_TILE_H * ty + yy) * sa_w + TM_BLK_TILE_W * tx + 8 * o);
            if (8 * o + 8 <= n_x) { // a whole octet (every window of 8 or more positions per row): no per-position range test;
                                    // key = sad * 2^12 + idx0 as a multiply-add (FMA pipe), "+ p, min" as one add-min
#pragma unroll
                for (int p = 0; p < 8; p++) {
                    const uint32_t sad = __reduce_add_sync(0xffffffffu, acc[p]);
                    best = min(best, sad * (1u << BLK_IDX_BITS) + idx0 + (uint32_t)p);
                }
            } else {
#pragma unroll
                for (int p = 0; p < 8; p++) {
                    const uint32_t sad = __reduce_add_sync(0xffffffffu, acc[p]);
                    if (8 * o + p < n_x) best = min(best, (sad << BLK_IDX_BITS) | (idx0 + (uint32_t)p));
                }
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
    int b, l, r;
    if (!compact_unit(job, blockIdx.x * SEARCH_WARPS + warp, b, l, r)) return;
    const int u = b * 8 + l * 4 + r; // (b64, list, ref)
    HmeJobG *jq = job.hme_jobs + (size_t)u * 4;
    // lanes 0..3 hold the four quadrant jobs
    int my_act = 0, my_px = 0, my_py = 0, my_sw = 0, my_sh = 0;
    const DevPic   &rp = job.ref[l][r];
    if (lane < 4) {
        const HmeJobG &j = jq[lane];
        if (j.active && j.sa_w > 0 && j.sa_h > 0) {
            my_act = 1; my_px = j.px; my_py = j.py;
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
