// me_tma.cuh — search windows through TMA tile loads, and the block-parallel SAD evaluator that
// reads them from shared memory.
//
// The searches with large blocks and small windows (HME level 1 / 2, full-pel) read a
// (block + window) rectangle of a reference plane at an arbitrary byte position. Reading it with
// per-lane LDGs costs one L1 wavefront per row touched per instruction and made those kernels
// LSU-bound (profiles/r01_v3_summary.md). Here one lane issues ONE cp.async.bulk.tensor.2d per tile
// (measured: ~2.5 cycles per box row per SM, asynchronous, tools/microbench/tma_box.cu) and the warp
// reads the tile back with conflict-free LDS.128.
//
// TMA constraint (measured, traps otherwise): the innermost start coordinate must be 16-byte
// aligned. A tile is therefore fetched from x & ~15 and the evaluator starts `beta` bytes into each
// row: chunk select by address, word select by a warp-uniform 4-way switch, byte select by one
// funnel shift per word.
#pragma once
#include <cuda.h>
#include "me_device.cuh"

// tensor-map kinds of one picture (DevPic::tmap[kind]); box = bytes per row x rows
enum { TM_L1 = 0, TM_L2 = 1, TM_FP = 2, TM_S16T = 3, TM_S16W = 4, TM_KINDS = 5 };
#define TM_L1_BOX_W 80  /* 32 + 15 (positions) + 15 (alignment) = 62; 80 = 5 chunks makes rows 2 apart hit different banks */
#define TM_L1_BOX_H 35  /* 32 block rows + 4 search rows - 1 */
#define TM_L2_BOX_W 96  /* 64 + 15 + 15 = 94 */
#define TM_L2_BOX_H 67
#define TM_FP_BOX_W 96
#define TM_FP_BOX_H 71  /* 64 + 8 search rows - 1 */
/* 1/16 plane, 16-wide blocks (pre-HME, HME level 0): a tall tile (16 x 96 positions) for windows up to 16 positions wide, a wide one otherwise.
   Rows are read with LDS.64 from the 8-byte boundary at or before a position: 8 * 17 + 32 = 168 > 160 never happens
   because the last octet of a wide tile starts at most 8 * 16 bytes in (8 * (1 + 15) + 32 = 160). */
#define TM_S16T_BOX_W 48
#define TM_S16T_BOX_H 111 /* 96 search rows + 15 block rows */
#define TM_S16W_BOX_W 160
#define TM_S16W_BOX_H 33  /* 18 search rows + 15 */
#define S16T_TILE_W 16
#define S16T_TILE_H 96
#define S16W_TILE_W 128
#define S16W_TILE_H 18
#define TM_BLK_TILE_W 16
#define TM_BLK_TILE_H 4
#define TM_FP_TILE_W 16
#define TM_FP_TILE_H 8
#define TM_ALIGN128(x) (((x) + 127) & ~127)
#define TM_L1_BUF TM_ALIGN128(TM_L1_BOX_W * TM_L1_BOX_H)
#define TM_L2_BUF TM_ALIGN128(TM_L2_BOX_W * TM_L2_BOX_H)
#define TM_FP_BUF TM_ALIGN128(TM_FP_BOX_W * TM_FP_BOX_H)
#define TM_S16_BUF TM_ALIGN128(TM_S16T_BOX_W * TM_S16T_BOX_H) /* 5376 >= the wide box (5280) */

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
// bar32: shared-window address of the mbarrier (smem_u32 once, at pipe_init: a generic->shared conversion per wait was
// 2 % of k_search16's instructions)
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar32, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar32), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar32, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(bar32), "r"(parity)
        : "memory");
}
__device__ __forceinline__ uint2 lds64(uint32_t addr) {
    uint2 v;
    asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(addr));
    return v;
}
__device__ __forceinline__ uint4 lds128(uint32_t addr) {
    uint4 v;
    asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr));
    return v;
}

// one warp's double-buffered tile pipeline
struct TilePipe {
    uint32_t  buf[2]; // shared addresses of the two stages
    uint32_t  bar[2]; // shared addresses of the two mbarriers
    uint32_t  phase;  // bit s = parity the next wait on stage s expects
    uint32_t  bytes;
};
__device__ __forceinline__ void pipe_init(TilePipe &P, uint8_t *stage0, uint32_t stage_bytes, uint64_t *bars, uint32_t box_bytes) {
    P.buf[0] = smem_u32(stage0);
    P.buf[1] = P.buf[0] + stage_bytes;
    P.bar[0] = smem_u32(&bars[0]);
    P.bar[1] = P.bar[0] + 8u;
    P.phase = 0;
    P.bytes = box_bytes;
    if ((threadIdx.x & 31) == 0) { mbar_init(&bars[0], 1); mbar_init(&bars[1], 1); }
    mbar_fence_init();
    __syncwarp();
}
// x_aligned must be a multiple of 16 (bytes); coordinates are relative to the plane buffer and may lie outside it (zero
// fill). The caller guarantees that every lane has finished reading stage s (a __syncwarp after the reads).
__device__ __forceinline__ void pipe_issue(TilePipe &P, int s, const CUtensorMap *map, int x_aligned, int y) {
    if ((threadIdx.x & 31) == 0) {
        mbar_expect_tx(P.bar[s], P.bytes);
        asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(
                         P.buf[s]),
                     "l"(map), "r"(x_aligned), "r"(y), "r"(P.bar[s])
                     : "memory");
    }
}
__device__ __forceinline__ void pipe_issue_bytes(TilePipe &P, int s, const CUtensorMap *map, int x_aligned, int y, uint32_t bytes) {
    if ((threadIdx.x & 31) == 0) {
        mbar_expect_tx(P.bar[s], bytes);
        asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(
                         P.buf[s]),
                     "l"(map), "r"(x_aligned), "r"(y), "r"(P.bar[s])
                     : "memory");
    }
}
__device__ __forceinline__ void pipe_wait(TilePipe &P, int s) {
    mbar_wait(P.bar[s], (P.phase >> s) & 1u);
    P.phase ^= 1u << s;
}

// ---- the evaluator ---------------------------------------------------------------------------
// lane = band * NS + strip: strip = 16-pixel column of the block, band = group of NR block rows.
// roff[q]: byte offset inside the tile of the q-th row this lane sums, for search row 0.
template <int NS, int NR>
struct TLane {
    uint32_t src[NR][4];
    uint32_t roff[NR];
    uint32_t m[4];   // byte masks of a strip that is cut by the block's right edge
    uint32_t rowok;  // bit q: row q exists (blocks cut by the picture's bottom edge)
    int      strip, band;
    bool     full;   // warp-uniform: the block has all its columns and rows
};

// rows: lane-specific block rows row_of[q] (block coordinates, already multiplied by the row step); pitch: tile row pitch
template <int NS, int NR>
__device__ __forceinline__ void tlane_init(TLane<NS, NR> &L, const uint8_t *src, int src_stride, const int (&row_of)[NR], int bw, int bh,
                                           int pitch) {
    const int lane = threadIdx.x & 31;
    L.strip = lane % NS;
    L.band  = lane / NS;
    const int rem = bw - 16 * L.strip;
    L.m[0] = lo_bytes_mask(rem); L.m[1] = lo_bytes_mask(rem - 4); L.m[2] = lo_bytes_mask(rem - 8); L.m[3] = lo_bytes_mask(rem - 12);
    L.rowok = 0;
    bool all = rem >= 16;
#pragma unroll
    for (int q = 0; q < NR; q++) {
        const bool ok = row_of[q] < bh;
        all = all && ok;
        L.rowok |= ok ? (1u << q) : 0u;
        L.roff[q] = (uint32_t)(row_of[q] * pitch + 16 * L.strip);
        uint4 v = make_uint4(0, 0, 0, 0);
        if (ok && rem > 0) v = __ldg((const uint4 *)(src + (size_t)row_of[q] * src_stride + 16 * L.strip)); // block origins are 16-byte aligned
        L.src[q][0] = v.x & L.m[0]; L.src[q][1] = v.y & L.m[1]; L.src[q][2] = v.z & L.m[2]; L.src[q][3] = v.w & L.m[3];
    }
    L.full = __all_sync(0xffffffffu, all);
}

// w[q][0..5] = the 24 bytes that start 4*J + t8/8 bytes into the 16-byte chunk at base + roff[q], for every row q of the
// lane. Only this extraction depends on the alignment: it is the body of a warp-uniform 4-way switch, the SAD code
// after it exists once (instruction-cache footprint: the first version unrolled everything four times and stalled
// on instruction fetch, profiles/r01_v4_summary.md).
template <int NR, int J>
__device__ __forceinline__ void tile_rows_words(const uint32_t (&roff)[NR], uint32_t base, uint32_t t8, uint32_t (&w)[NR][6]) {
#pragma unroll
    for (int q = 0; q < NR; q++) {
        const uint32_t addr = base + roff[q];
        uint32_t c[12];
        const uint4 a = lds128(addr), b = lds128(addr + 16);
        c[0] = a.x; c[1] = a.y; c[2] = a.z; c[3] = a.w; c[4] = b.x; c[5] = b.y; c[6] = b.z; c[7] = b.w;
        if (J >= 2) {
            const uint4 d = lds128(addr + 32);
            c[8] = d.x; c[9] = d.y; c[10] = d.z; c[11] = d.w;
        } else {
            c[8] = c[9] = c[10] = c[11] = 0; // J <= 1 stops at c[J+6] <= c[7]
        }
#pragma unroll
        for (int k = 0; k < 6; k++) w[q][k] = __funnelshift_r(c[J + k], c[J + k + 1], t8);
    }
}

// one pass: the 8 positions whose first byte lies beta bytes into tile row `yy` (row_base = shared address of that row).
// SPLIT: left / right 8 pixels of the strip kept apart (accl, accr), else summed into accl.
template <int NS, int NR, bool SPLIT>
__device__ __forceinline__ void tile_pass(const TLane<NS, NR> &L, uint32_t row_base, uint32_t beta, uint32_t (&accl)[8], uint32_t (&accr)[8]) {
#pragma unroll
    for (int p = 0; p < 8; p++) accl[p] = accr[p] = 0;
    const uint32_t base = row_base + (beta & ~15u), t8 = (beta & 3u) * 8u;
    uint32_t w[NR][6];
    switch ((beta & 15u) >> 2) { // warp-uniform
    case 0: tile_rows_words<NR, 0>(L.roff, base, t8, w); break;
    case 1: tile_rows_words<NR, 1>(L.roff, base, t8, w); break;
    case 2: tile_rows_words<NR, 2>(L.roff, base, t8, w); break;
    default: tile_rows_words<NR, 3>(L.roff, base, t8, w); break;
    }
    if (L.full) {
#pragma unroll
        for (int q = 0; q < NR; q++) {
            if (SPLIT) strip16_sad8_split(w[q], L.src[q], accl, accr);
            else strip16_sad8(w[q], L.src[q], accl);
        }
    } else {
        // a strip cut by the block's right edge or a row below its bottom edge: mask what the source masks
#pragma unroll 1
        for (int q = 0; q < NR; q++) {
            const uint32_t rm = (L.rowok >> q) & 1u ? 0xffffffffu : 0u;
            uint32_t wq[6], sq[4];
#pragma unroll
            for (int qq = 0; qq < NR; qq++)
                if (qq == q) {
#pragma unroll
                    for (int k = 0; k < 6; k++) wq[k] = w[qq][k];
#pragma unroll
                    for (int k = 0; k < 4; k++) sq[k] = L.src[qq][k];
                }
#pragma unroll
            for (int p = 0; p < 8; p++) {
                const int o = p >> 2, sh = (p & 3) * 8;
#pragma unroll
                for (int k = 0; k < 4; k++) {
                    const uint32_t v = (sh ? __funnelshift_r(wq[o + k], wq[o + k + 1], sh) : wq[o + k]) & L.m[k] & rm;
                    if (SPLIT && k >= 2) accr[p] = sad4_acc(v, sq[k], accr[p]);
                    else accl[p] = sad4_acc(v, sq[k], accl[p]);
                }
            }
        }
    }
}

// SAD of the lane's rows of its strip at ONE position whose first byte lies beta bytes into tile row `row_base` (any
// alignment: five aligned words per row, one funnel shift and one VABSDIFF4 per word). Full blocks only (L.full).
template <int NS, int NR>
__device__ __forceinline__ uint32_t tile_sad1(const TLane<NS, NR> &L, uint32_t row_base, uint32_t beta) {
    const uint32_t base = row_base + (beta & ~3u), sh = (beta & 3u) * 8u;
    uint32_t acc = 0;
#pragma unroll
    for (int q = 0; q < NR; q++) {
        const uint32_t addr = base + L.roff[q];
        uint32_t c[5];
#pragma unroll
        for (int i = 0; i < 5; i++) asm volatile("ld.shared.u32 %0, [%1];" : "=r"(c[i]) : "r"(addr + 4u * i));
#pragma unroll
        for (int i = 0; i < 4; i++) acc = sad4_acc(__funnelshift_r(c[i], c[i + 1], sh), L.src[q][i], acc);
    }
    return acc;
}
