// me_fp3.cuh — full-pel (integer) search, block-parallel form over TMA tiles.
//
// One warp evaluates 8 consecutive search positions ("octet") of the 64x64 block per pass, with
// the 32 lanes spread over the BLOCK: lane = band*4 + strip, strip = 16-pixel column (0..3), band =
// 8-pixel row band (0..7). A lane keeps its 16x8 piece of the source block in registers for the
// whole job and reads the matching reference rows from a shared-memory tile that one TMA load
// fetched (me_tma.cuh): 16 positions x 8 search rows per tile. It accumulates the two 8x8 blocks
// of its piece for the 8 positions (strip16_sad8_split); 16x16 / 32x32 / 64x64 SADs are completed with
// halving exchanges (4 + 2 + 1 + 2 shuffles per pass). Each lane owns exactly two 8x8 PUs (and some
// lanes a 16x16, 32x32 or the 64x64 PU): 5 packed keys per lane instead of 85 per thread.
//
//   k_fp       one warp per (b64, list, ref): window sizing (integer_search_b64,
//              motion_estimation.c:1296-1561, check_00_center :1139, 8x8-variance probe :1393-1441) and,
//              when the window is small, the whole scan (open_loop_me_fullpel_search_sblock :781,
//              svt_ext_all_sad_calculation_8x8_16x16 :335, svt_ext_eight_sad_calculation_32x32_64x64
//              :369, tails :98-205). Larger windows are cut into tiles and queued.
//   k_fp_big   persistent warps pull (picture, job, tile) units and merge with 64-bit atomicMin.
//
// Keys: 32-bit (sad << 12 | index), index 0 = the variance probe (evaluated first in the reference and
// never reset, so it wins ties), index 1 + y*sa_w + x for window positions; min() = first minimum in
// raster order = the reference's strict '<'. SADs stay un-doubled (SUB_SAD_SEARCH doubles all alike).
#pragma once
#include "me_device.cuh"
#include "me_tma.cuh"
#include "me_search.cuh"

#define FP_WARPS 4
#define FP_IDX_BITS 12 /* 20-bit SAD (64x64x255) + 12-bit index = 32 bits */
#define FP_INLINE_TILES 3 /* windows of up to this many tiles are scanned by the planning warp itself */

__constant__ uint8_t c_z16b[16] = {0, 1, 4, 5, 2, 3, 6, 7, 8, 9, 12, 13, 10, 11, 14, 15}; // motion_estimation.c:341

template <bool SUB>
struct FpLane {
    static constexpr int NR = SUB ? 4 : 8; // rows of the band this lane touches
    TLane<4, NR> T;
    uint32_t kl, kr, k16, k32, k64; // running bests (packed keys) of the PUs it may own
};

// rows of a band are visited in a rotated order by the odd bands: two lanes of a quarter-warp then read tile rows
// 2 or 10 apart, which a 96-byte pitch maps to different banks (rows 8 apart would collide)
template <bool SUB>
__device__ __forceinline__ void fp_lane_init(FpLane<SUB> &L, const uint8_t *src_full, int src_stride, int bw, int bh) {
    constexpr int NR = FpLane<SUB>::NR;
    const int band = (threadIdx.x & 31) >> 2;
    int row_of[NR];
#pragma unroll
    for (int q = 0; q < NR; q++) row_of[q] = 8 * band + (SUB ? 2 : 1) * ((q + (band & 1) * (SUB ? 1 : 2)) % NR);
    tlane_init<4, NR>(L.T, src_full, src_stride, row_of, bw, bh, TM_FP_BOX_W);
    L.kl = L.kr = L.k16 = L.k32 = L.k64 = 0xffffffffu;
}

// one pass: 8 positions starting beta bytes into tile row `row_base`; idx0 = key index of the first, nv = how many exist
// raw (optional): this lane's left / right 8x8 sums and the 64x64 sum of the FIRST position (the variance probe)
template <bool SUB>
__device__ __forceinline__ void fp_eval_pass(FpLane<SUB> &L, uint32_t row_base, uint32_t beta, uint32_t idx0, int nv,
                                             uint32_t *raw = nullptr) {
    constexpr int NR = FpLane<SUB>::NR;
    uint32_t accl[8], accr[8];
    tile_pass<4, NR, true>(L.T, row_base, beta, accl, accr);
    const int  lane = threadIdx.x & 31;
    const bool h4 = lane & 4, h1 = lane & 1, h8 = lane & 8;
    uint32_t s8[8];
    const bool all8 = nv >= 8; // warp-uniform; the usual case (windows are multiples of 8 positions wide): no range tests
#pragma unroll
    for (int p = 0; p < 8; p++) {
        s8[p] = accl[p] + accr[p];
        if (all8 || p < nv) {
            L.kl = min(L.kl, accl[p] * (1u << FP_IDX_BITS) + idx0 + p);
            L.kr = min(L.kr, accr[p] * (1u << FP_IDX_BITS) + idx0 + p);
        }
    }
    uint32_t t16[4], u[2];
    const int p16 = h4 ? 4 : 0;
#pragma unroll
    for (int i = 0; i < 4; i++) { // the other 8-row band of the 16x16: it keeps the other four positions
        const uint32_t mine = h4 ? s8[4 + i] : s8[i], theirs = h4 ? s8[i] : s8[4 + i];
        t16[i] = mine + __shfl_xor_sync(0xffffffffu, theirs, 4);
        if (all8 || p16 + i < nv) L.k16 = min(L.k16, t16[i] * (1u << FP_IDX_BITS) + idx0 + p16 + i);
    }
#pragma unroll
    for (int i = 0; i < 2; i++) { // the other strip of the 32x32
        const uint32_t mine = h1 ? t16[2 + i] : t16[i], theirs = h1 ? t16[i] : t16[2 + i];
        u[i] = mine + __shfl_xor_sync(0xffffffffu, theirs, 1);
    }
    const uint32_t mine = h8 ? u[1] : u[0], theirs = h8 ? u[0] : u[1];
    const uint32_t t32 = mine + __shfl_xor_sync(0xffffffffu, theirs, 8); // the other 16-row band
    const int p32 = p16 + (h1 ? 2 : 0) + (h8 ? 1 : 0);
    uint32_t t64 = t32 + __shfl_xor_sync(0xffffffffu, t32, 2);
    t64 += __shfl_xor_sync(0xffffffffu, t64, 16);
    if (p32 < nv) {
        L.k32 = min(L.k32, (t32 << FP_IDX_BITS) + idx0 + p32);
        L.k64 = min(L.k64, (t64 << FP_IDX_BITS) + idx0 + p32);
    }
    if (raw) {
        raw[0] = accl[0];
        raw[1] = accr[0];
        raw[2] = __shfl_sync(0xffffffffu, t64, 0); // lane 0 holds position 0 after the halving exchanges
    }
}

// one position (the variance probe / the RTCD known-answer path): raw = sums of this lane's left 8x8, right 8x8, its
// 16x16, its 32x32 and the 64x64; keys updated with index idx
template <bool SUB>
__device__ __forceinline__ void fp_eval_one(FpLane<SUB> &L, uint32_t row_base, uint32_t beta, uint32_t idx, uint32_t (&raw)[5]) {
    constexpr int NR = FpLane<SUB>::NR;
    uint32_t accl[8], accr[8];
    tile_pass<4, NR, true>(L.T, row_base, beta, accl, accr); // 8 positions computed, the first one used
    const uint32_t s8  = accl[0] + accr[0];
    const uint32_t t16 = s8 + __shfl_xor_sync(0xffffffffu, s8, 4);
    uint32_t       t32 = t16 + __shfl_xor_sync(0xffffffffu, t16, 1);
    t32 += __shfl_xor_sync(0xffffffffu, t32, 8);
    uint32_t t64 = t32 + __shfl_xor_sync(0xffffffffu, t32, 2);
    t64 += __shfl_xor_sync(0xffffffffu, t64, 16);
    raw[0] = accl[0]; raw[1] = accr[0]; raw[2] = t16; raw[3] = t32; raw[4] = t64;
    L.kl  = min(L.kl, (accl[0] << FP_IDX_BITS) + idx);
    L.kr  = min(L.kr, (accr[0] << FP_IDX_BITS) + idx);
    L.k16 = min(L.k16, (t16 << FP_IDX_BITS) + idx);
    L.k32 = min(L.k32, (t32 << FP_IDX_BITS) + idx);
    L.k64 = min(L.k64, (t64 << FP_IDX_BITS) + idx);
}

// the lanes of a PU hold the bests of different positions: combine, then visit the PUs this lane owns: f(key, pu)
template <bool SUB, typename F>
__device__ __forceinline__ void fp_for_owned(const FpLane<SUB> &L, F f) {
    const int lane = threadIdx.x & 31, strip = lane & 3, band = lane >> 2;
    const uint32_t k16 = min(L.k16, __shfl_xor_sync(0xffffffffu, L.k16, 4));
    uint32_t k32 = min(L.k32, __shfl_xor_sync(0xffffffffu, L.k32, 4));
    k32 = min(k32, __shfl_xor_sync(0xffffffffu, k32, 1));
    k32 = min(k32, __shfl_xor_sync(0xffffffffu, k32, 8));
    const uint32_t k64 = __reduce_min_sync(0xffffffffu, L.k64);
    const int z = c_z16b[4 * (band >> 1) + strip];
    const int k = 2 * (band & 1);
    f(L.kl, 21 + 4 * z + k);
    f(L.kr, 21 + 4 * z + k + 1);
    if ((band & 1) == 0) f(k16, 5 + z);
    if ((band & 3) == 0 && (strip & 1) == 0) f(k32, 1 + 2 * (band >> 2) + (strip >> 1));
    if (band == 0 && strip == 0) f(k64, 0);
}

// positions [ox, ox+n_x) x [oy, oy+n_y) of a tile whose position (0,0) lies `a` bytes into its first row;
// key index of the rectangle's position (0,0) = idx_base, of the next row = idx_base + idx_row
template <bool SUB>
__device__ __forceinline__ void fp_scan_rect(FpLane<SUB> &L, uint32_t buf, uint32_t a, int ox, int oy, int n_x, int n_y, uint32_t idx_base,
                                             uint32_t idx_row) {
    for (int yy = 0; yy < n_y; yy++) {
        const uint32_t row_base = buf + (uint32_t)((oy + yy) * TM_FP_BOX_W);
        for (int x = 0; x < n_x; x += 8)
            fp_eval_pass<SUB>(L, row_base, a + (uint32_t)(ox + x), idx_base + (uint32_t)yy * idx_row + (uint32_t)x, imin(8, n_x - x));
    }
}

// ---- k_fp ------------------------------------------------------------------------------------
// phase: 0 all refs, 1 only (0,0), 2 all but (0,0) (enable_me_sr_adjustment == 2, :1357-1366)
template <bool SUB>
__global__ void __launch_bounds__(32 * FP_WARPS) k_fp(const PicJob *__restrict__ jobs, int phase) {
    extern __shared__ __align__(128) uint8_t s_dyn[];
    __shared__ __align__(8) uint64_t s_bar[FP_WARPS][2];
    const PicJob &job = jobs[blockIdx.y];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    int b, l, r;
    if (!compact_unit(job, blockIdx.x * FP_WARPS + warp, b, l, r)) return; // slots the picture does not have: k_post skips them
    const int u = b * 8 + l * 4 + r;
    const SvtMePicParams &pp = job.pp;
    const SvtMeCtrls     &c  = pp.ctrls;
    FpJobG &fj = job.fp_jobs[u];
    if (phase == 1 && (l || r)) return;
    if (phase == 2 && !(l || r)) return;
    const SbState &st = job.sb_state[b];
    unsigned long long *gk = job.fp_keys + (size_t)u * 85;
    // everything this warp reads of the b64's state and of other warps' results, requested together, before the first
    // branch that depends on any of it (the loads are one round trip instead of four dependent ones)
    const uint8_t st_do_ref = st.do_ref[l][r], st_tf_exit = st.tf_exit;
    const int16_t st_xc = st.hme_sc_x[l][r], st_yc = st.hme_sc_y[l][r];
    const uint32_t st_div = st.reduce_me_sr_divisor[l][r], st_zz = st.zz_sad[l][r];
    const auto st_sr_hme_sad = st.sr_hme_sad[l][r];
    const unsigned long long key00 = phase == 2 ? job.fp_keys[(size_t)b * 8 * 85] : 0ull; // p_sb_best_sad[0][0][0]: written by the phase-1 launch
    if (!st_do_ref || st_tf_exit) { // :1294-1295; temporal filter: HME was good enough (:3179-3183)
        if (lane == 0) { fj.active = 0; if (job.windows) job.windows[u] = SvtMeWindow{0, 0, 0, 0}; }
        return;
    }
    const B64Geo    g   = b64_geo(job, b);
    const DevPic   &rp  = job.ref[l][r];
    const DevPlane &ref = rp.lvl[0];
    const DevPlane &sf  = job.src.lvl[0];
    const uint8_t  *src_full = sf.buf + (size_t)(sf.org_y + g.org_y) * sf.stride + sf.org_x + g.org_x;
    const CUtensorMap *map = (const CUtensorMap *)rp.tmap + TM_FP;
    uint8_t *sbase = s_dyn + ((128u - (smem_u32(s_dyn) & 127u)) & 127u);
    TilePipe P;
    pipe_init(P, sbase + (size_t)warp * 2 * TM_FP_BUF, TM_FP_BUF, s_bar[warp], TM_FP_BOX_W * TM_FP_BOX_H);
    unsigned long long ops = 0;

    // window sizing :1296-1366 (all lanes compute the same scalars; SADs are warp-cooperative)
    int16_t xc = st_xc, yc = st_yc;
    int16_t sa_w = (int16_t)c.me_sa.sa_min.width, sa_h = (int16_t)c.me_sa.sa_min.height;
    const uint16_t dist = c.me_type == SVTME_ME_MCTF ? ref_distance(pp, l, r) : scaled_picture_distance(ref_distance(pp, l, r)); // :1302-1304
    sa_w = (int16_t)imin((sa_w * dist), c.me_sa.sa_max.width);
    sa_h = (int16_t)imin((sa_h * dist), c.me_sa.sa_max.height);
    if (c.mv_sa_adj_enabled && (!c.mv_sa_adj_nearest_ref_only || r == 0)) {
        if (iabs(xc) > c.mv_sa_adj_mv_size_th) sa_w = (int16_t)(sa_w * c.mv_sa_adj_sa_multiplier);
        if (iabs(yc) > c.mv_sa_adj_mv_size_th) sa_h = (int16_t)(sa_h * c.mv_sa_adj_sa_multiplier);
    }
    const uint32_t div = st_div;
    sa_w = (int16_t)((max(1u, ((uint32_t)sa_w / div)) + 7) & ~0x07u);
    sa_h = (int16_t)max(3u, ((uint32_t)sa_h / div));
    const int16_t sa_h0 = sa_h, sa_w0 = sa_w;
    uint64_t best_hme_sad = ~0ull;
    if (c.me_early_exit_th) {
        if (st_zz < (c.me_early_exit_th / 6)) sa_w = sa_h = 1;
    }
    // plane coordinates of the reference sample under the block's top-left sample at MV (0,0)
    const int bx = ref.org_x + g.org_x, by = ref.org_y + g.org_y;
    FpLane<SUB> L;
    bool lane_ready = false, probe_loaded = false; // probe_loaded: stages 0 / 1 hold the probe tiles of centre (0,0) / the HME centre
    if (!c.me_early_exit_th) {
        uint8_t hme_is_accurate = 1;
        if ((xc != 0 || yc != 0) && pp.is_ref) {
            // check_00_center :1139-1206 (the COST_PRECISION shift is monotonic)
            const int16_t ox = (int16_t)g.org_x, oy = (int16_t)g.org_y, pad = 63;
            int16_t xh = xc, yh = yc;
            xh = ((ox + xh) < -pad) ? (int16_t)(-pad - ox) : xh;
            xh = ((ox + xh) > (int16_t)ref.width - 1) ? (int16_t)(xh - ((ox + xh) - ((int16_t)ref.width - 1))) : xh;
            yh = ((oy + yh) < -pad) ? (int16_t)(-pad - oy) : yh;
            yh = ((oy + yh) > (int16_t)ref.height - 1) ? (int16_t)(yh - ((oy + yh) - ((int16_t)ref.height - 1))) : yh;
            uint32_t zero_mv_sad, hme_mv_sad;
            if (SUB && g.b64_w == 64 && g.b64_h == 64) {
                // both SADs (64 wide, the 32 even rows: svt_nxm_sad_kernel with doubled strides) from TMA tiles and the
                // lane's source rows, which the scan needs anyway: two tile loads in flight together, ~60 instructions per
                // position, no per-lane global loads (the unaligned-LDG form below was a quarter of this kernel's
                // instructions and its largest stall, profiles/r02_ncu_notes.md)
                // The two tiles are fetched where the PROBE tile of either outcome lies (8 columns and 4 rows before the
                // position): whichever centre wins, its probe tile is already in shared memory and the probe below does
                // not wait for a third load (the set-up of a warp is a chain of dependent latencies: job record, b64 state,
                // these tiles, the probe tile; ncu: long-scoreboard and barrier waits are the kernel's largest stalls).
                pipe_issue(P, 0, map, (bx - 8) & ~15, by - 4);
                pipe_issue(P, 1, map, (bx + xh - 8) & ~15, by + yh - 4);
                fp_lane_init<SUB>(L, src_full, sf.stride, 64, 64);
                lane_ready = true;
                pipe_wait(P, 0);
                zero_mv_sad = __reduce_add_sync(0xffffffffu, tile_sad1<4, FpLane<SUB>::NR>(L.T, P.buf[0] + 4u * TM_FP_BOX_W, (uint32_t)((bx - 8) & 15) + 8u)) << 1;
                pipe_wait(P, 1);
                hme_mv_sad = __reduce_add_sync(0xffffffffu, tile_sad1<4, FpLane<SUB>::NR>(L.T, P.buf[1] + 4u * TM_FP_BOX_W, (uint32_t)((bx + xh - 8) & 15) + 8u)) << 1;
                probe_loaded = true;
            } else {
                const uint8_t *p0 = ref.buf + ((int16_t)ref.org_x + ox + ((int16_t)ref.org_y + oy) * ref.stride);
                zero_mv_sad = warp_nxm_sad2(src_full, sf.stride << 1, p0, ref.stride << 1, g.b64_h >> 1, g.b64_w) << 1;
                const uint8_t *p1 = ref.buf + ((int16_t)(ref.org_x + ox) + xh + ((int16_t)(ref.org_y + oy) + yh) * ref.stride);
                hme_mv_sad = warp_nxm_sad2(src_full, sf.stride << 1, p1, ref.stride << 1, g.b64_h >> 1, g.b64_w) << 1;
            }
            xc = xh; yc = yh;
            const uint64_t zero_cost = (uint64_t)(zero_mv_sad << 8), hme_cost = (uint64_t)(hme_mv_sad << 8);
            if ((zero_cost < hme_cost ? zero_cost : hme_cost) == zero_cost) xc = yc = 0;
            best_hme_sad = hme_mv_sad;
            if (xc == 0 && yc == 0) hme_is_accurate = 0;
            ops += 2ull * (unsigned long long)(g.b64_h >> 1) * g.b64_w;
        }
        if (c.enable_me_sr_adjustment == 2) {
            if ((hme_is_accurate && (best_hme_sad < (24 * 24))) || (pp.is_ref && st_sr_hme_sad < (24 * 24)))
                sa_h = (int16_t)(sa_h / 2);
            const uint32_t best00 = (uint32_t)(key00 >> 32) << (SUB ? 1 : 0); // p_sb_best_sad[0][0][0]; only references other than (0,0) use it: phase 2
            if ((l || r) && best00 < 5000 && sa_h == sa_h0 && sa_w == sa_w0) {
                sa_h = (int16_t)(sa_h >> 1);
                sa_w = (int16_t)(sa_w >> 1);
            }
        }
    }
    const bool probe = c.me_8x8_var_enabled && (sa_w * sa_h > 24);
    const int  ptx = bx + xc - 8, pty = by + yc - 4; // the probe tile: 16 x 8 positions around the search centre
    const int  pst = probe_loaded && (xc != 0 || yc != 0) ? 1 : 0; // the stage that holds it
    if (probe && !probe_loaded) pipe_issue(P, 0, map, ptx & ~15, pty);
    if (!lane_ready) fp_lane_init<SUB>(L, src_full, sf.stride, 64, 64); // the full-pel stage always works on 64x64 (padded source)
    if (probe) { // :1393-1441: one position at the search centre, then the variance of its 64 8x8 SADs
        if (!probe_loaded) pipe_wait(P, 0);
        uint32_t raw[3];
        fp_eval_pass<SUB>(L, P.buf[pst] + 4u * TM_FP_BOX_W, (uint32_t)(ptx & 15) + 8u, 0u, 1, raw);
        const uint32_t dbl  = SUB ? 1 : 0;
        const uint32_t mean = (raw[2] << dbl) / 64;
        const int32_t  dl = (int32_t)(raw[0] << dbl) - (int32_t)mean, dr = (int32_t)(raw[1] << dbl) - (int32_t)mean;
        uint32_t sumsq = (uint32_t)(dl * dl) + (uint32_t)(dr * dr);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) sumsq += __shfl_xor_sync(0xffffffffu, sumsq, o);
        const uint32_t var = sumsq / 64;
        if (var > c.me_sr_mult2_th) {
            sa_w = (int16_t)((imax(1, sa_w * 3 / 2) + 7) & ~0x7);
            sa_h = (int16_t)imax(1, sa_h * 3 / 2);
        }
        if (var < c.me_sr_div4_th) {
            sa_w = (int16_t)((imax(1, sa_w >> 2) + 7) & ~0x7);
            sa_h = (int16_t)imax(1, sa_h >> 2);
            sa_h = (int16_t)imax(3, sa_h);
        } else if (var < c.me_sr_div2_th) {
            sa_w = (int16_t)((imin(sa_w, sa_w >> 1) + 7) & ~0x7);
            sa_h = (int16_t)imin(sa_h, sa_h >> 1);
            sa_h = (int16_t)imax(3, sa_h);
        }
    }
    // final window :1442-1561 (restricted_motion_vector == 0)
    int16_t x_origin = (int16_t)(xc - (sa_w >> 1));
    int16_t y_origin = (int16_t)(yc - (sa_h >> 1));
    clamp_axis((int16_t)g.org_x, x_origin, sa_w, 63, (int16_t)job.pic_w);
    sa_w = (int16_t)((sa_w < 8) ? sa_w : sa_w & ~0x07);
    clamp_axis((int16_t)g.org_y, y_origin, sa_h, 63, (int16_t)job.pic_h);
    const int ntx = (sa_w + TM_FP_TILE_W - 1) / TM_FP_TILE_W, nty = (sa_h + TM_FP_TILE_H - 1) / TM_FP_TILE_H, n_tiles = ntx * nty;
    const int wx = bx + x_origin, wy = by + y_origin; // plane coordinates of window position (0,0)
    const int rx = wx - ptx, ry = wy - pty;           // ... relative to the probe tile
    const bool in_probe_tile = probe && rx >= 0 && ry >= 0 && rx + sa_w <= TM_FP_TILE_W && ry + sa_h <= TM_FP_TILE_H;
    const bool inline_scan = in_probe_tile || n_tiles <= FP_INLINE_TILES;
    // one loop, one call site of the scan (code size): the probe tile when it covers the window, else the window's tiles
    if (inline_scan) {
        __syncwarp();
        int stage = in_probe_tile ? pst : pst ^ 1;
        if (!in_probe_tile) pipe_issue(P, stage, map, wx & ~15, wy);
        const int n_iter = in_probe_tile ? 1 : n_tiles;
        for (int t = 0; t < n_iter; t++) {
            const int ty = t / ntx, tx = t - ty * ntx;
            uint32_t a = (uint32_t)(ptx & 15), idx_base = 1u;
            int ox = rx, oy = ry, nx = sa_w, ny = sa_h;
            if (!in_probe_tile) {
                if (t + 1 < n_tiles) {
                    const int ty1 = (t + 1) / ntx, tx1 = (t + 1) - ty1 * ntx;
                    pipe_issue(P, stage ^ 1, map, (wx + TM_FP_TILE_W * tx1) & ~15, wy + TM_FP_TILE_H * ty1);
                }
                pipe_wait(P, stage);
                a = (uint32_t)((wx + TM_FP_TILE_W * tx) & 15);
                ox = oy = 0;
                nx = imin(TM_FP_TILE_W, sa_w - TM_FP_TILE_W * tx);
                ny = imin(TM_FP_TILE_H, sa_h - TM_FP_TILE_H * ty);
                idx_base = 1u + (uint32_t)(TM_FP_TILE_H * ty * sa_w + TM_FP_TILE_W * tx);
            }
            fp_scan_rect<SUB>(L, P.buf[stage], a, ox, oy, nx, ny, idx_base, (uint32_t)sa_w);
            __syncwarp();
            stage ^= 1;
        }
    }
    // running bests -> 64-bit keys (sad << 32 | scan index); ~0 = untouched (MAX_SAD_VALUE, :1368)
    fp_for_owned<SUB>(L, [&](uint32_t k, int pu) {
        gk[pu] = k == 0xffffffffu ? ~0ull : (((unsigned long long)(k >> FP_IDX_BITS) << 32) | (k & ((1u << FP_IDX_BITS) - 1)));
    });
    if (lane == 0) {
        const int n_units = inline_scan ? 0 : n_tiles;
        fj.xc = xc; fj.yc = yc; fj.x_origin = x_origin; fj.y_origin = y_origin; fj.sa_w = sa_w; fj.sa_h = sa_h;
        fj.n_tx = (uint16_t)ntx; fj.n_ty = (uint16_t)nty; fj.active = 1; fj.probe = probe ? 1 : 0;
        if (job.windows) job.windows[u] = SvtMeWindow{x_origin, y_origin, sa_w, sa_h};
        if (n_units) {
            const uint32_t base = atomicAdd(&job.fp_counters[0], (uint32_t)n_units);
            for (int i = 0; i < n_units && base + i < job.fp_units_cap; i++)
                job.fp_units[base + i] = make_uint2((uint32_t)u, (blockIdx.y << 16) | (uint32_t)i);
        }
        ops += ((unsigned long long)sa_w * sa_h + (probe ? 1 : 0)) * 64ull * (SUB ? 32 : 64);
        if (job.sad_ops) atomicAdd(&job.sad_ops[1], ops);
    }
}

// ---- k_fp_big --------------------------------------------------------------------------------
struct FpUnit {
    int pic, u, tile, wx, wy, sa_w, sa_h, ntx;
    bool valid;
};
__device__ __forceinline__ FpUnit fp_fetch_unit(const PicJob *__restrict__ jobs, uint32_t n_units) {
    FpUnit un;
    uint32_t ui = 0;
    if ((threadIdx.x & 31) == 0) ui = atomicAdd(&jobs[0].fp_counters[1], 1u);
    ui = __shfl_sync(0xffffffffu, ui, 0);
    un.valid = ui < n_units;
    un.pic = un.u = un.tile = un.wx = un.wy = un.sa_w = un.sa_h = 0; un.ntx = 1;
    if (un.valid) {
        const uint2 q = jobs[0].fp_units[ui];
        un.pic = (int)(q.y >> 16); un.u = (int)q.x; un.tile = (int)(q.y & 0xffff);
        const PicJob &job = jobs[un.pic];
        const FpJobG &fj  = job.fp_jobs[un.u];
        const int b = un.u >> 3, l = (un.u >> 2) & 1, r = un.u & 3;
        const B64Geo    g   = b64_geo(job, b);
        const DevPlane &ref = job.ref[l][r].lvl[0];
        un.wx = ref.org_x + g.org_x + fj.x_origin; un.wy = ref.org_y + g.org_y + fj.y_origin;
        un.sa_w = fj.sa_w; un.sa_h = fj.sa_h; un.ntx = fj.n_tx;
    }
    return un;
}
__device__ __forceinline__ void fp_issue_unit(TilePipe &P, int stage, const PicJob *__restrict__ jobs, const FpUnit &un) {
    const int l = (un.u >> 2) & 1, r = un.u & 3;
    const int ty = un.tile / un.ntx, tx = un.tile - ty * un.ntx;
    const CUtensorMap *map = (const CUtensorMap *)jobs[un.pic].ref[l][r].tmap + TM_FP;
    pipe_issue(P, stage, map, (un.wx + TM_FP_TILE_W * tx) & ~15, un.wy + TM_FP_TILE_H * ty);
}

template <bool SUB>
__global__ void __launch_bounds__(32 * FP_WARPS) k_fp_big(const PicJob *__restrict__ jobs) {
    extern __shared__ __align__(128) uint8_t s_dyn[];
    __shared__ __align__(8) uint64_t s_bar[FP_WARPS][2];
    const int warp = threadIdx.x >> 5;
    const uint32_t n_units = min(jobs[0].fp_counters[0], jobs[0].fp_units_cap); // the unit queue is shared by the batch
    if (blockIdx.x == 0 && threadIdx.x == 0 && jobs[0].fp_counters[0] > jobs[0].fp_units_cap) jobs[0].fp_counters[8] = 1; // queue overflow: reported by the host
    if (n_units == 0) return;
    uint8_t *sbase = s_dyn + ((128u - (smem_u32(s_dyn) & 127u)) & 127u);
    TilePipe P;
    pipe_init(P, sbase + (size_t)warp * 2 * TM_FP_BUF, TM_FP_BUF, s_bar[warp], TM_FP_BOX_W * TM_FP_BOX_H);
    FpUnit cur = fp_fetch_unit(jobs, n_units);
    if (!cur.valid) return;
    int stage = 0;
    fp_issue_unit(P, stage, jobs, cur);
    FpLane<SUB> L;
    int have_pic = -1, have_u = -1;
    while (cur.valid) {
        const FpUnit nxt = fp_fetch_unit(jobs, n_units);
        if (nxt.valid) fp_issue_unit(P, stage ^ 1, jobs, nxt);
        const PicJob &job = jobs[cur.pic];
        if (cur.pic != have_pic || cur.u != have_u) {
            const B64Geo    g  = b64_geo(job, cur.u >> 3);
            const DevPlane &sf = job.src.lvl[0];
            fp_lane_init<SUB>(L, sf.buf + (size_t)(sf.org_y + g.org_y) * sf.stride + sf.org_x + g.org_x, sf.stride, 64, 64);
            have_pic = cur.pic; have_u = cur.u;
        }
        L.kl = L.kr = L.k16 = L.k32 = L.k64 = 0xffffffffu;
        const int ty = cur.tile / cur.ntx, tx = cur.tile - ty * cur.ntx;
        pipe_wait(P, stage);
        // key indices are local to the tile (row pitch 16) so that they fit the 12-bit field
        fp_scan_rect<SUB>(L, P.buf[stage], (uint32_t)((cur.wx + TM_FP_TILE_W * tx) & 15), 0, 0, imin(TM_FP_TILE_W, cur.sa_w - TM_FP_TILE_W * tx),
                          imin(TM_FP_TILE_H, cur.sa_h - TM_FP_TILE_H * ty), 0u, (uint32_t)TM_FP_TILE_W);
        __syncwarp();
        unsigned long long *gk = job.fp_keys + (size_t)cur.u * 85;
        const uint32_t gbase = 1u + (uint32_t)(TM_FP_TILE_H * ty * cur.sa_w + TM_FP_TILE_W * tx);
        const uint32_t sa_w = (uint32_t)cur.sa_w;
        fp_for_owned<SUB>(L, [&](uint32_t k, int pu) {
            if (k != 0xffffffffu) {
                const uint32_t li = k & ((1u << FP_IDX_BITS) - 1);
                atomicMin(&gk[pu], ((unsigned long long)(k >> FP_IDX_BITS) << 32) | (gbase + (li >> 4) * sa_w + (li & 15u)));
            }
        });
        cur = nxt;
        stage ^= 1;
    }
}
