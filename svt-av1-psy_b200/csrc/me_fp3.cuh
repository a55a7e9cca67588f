// me_fp3.cuh — full-pel (integer) search, block-parallel form.
//
// One warp evaluates 8 consecutive search positions ("octet") of the 64x64 block per pass, with
// the 32 lanes spread over the BLOCK: lane = band*4 + strip, strip = 16-pixel column (0..3), band =
// 8-pixel row band (0..7). A lane keeps its 16x8 piece of the source block in registers for the
// whole job, loads the matching reference rows straight from global memory (7 aligned words per
// row, funnel-shifted once to the window's byte alignment), and accumulates the two 8x8 blocks of
// its piece for the 8 positions (strip16_sad8_split). 16x16 / 32x32 / 64x64 SADs are completed with
// 5 shfl_xor per position. Each lane therefore owns exactly two 8x8 PUs (and some lanes a 16x16,
// 32x32 or the 64x64 PU): 5 packed keys per lane instead of 85 per thread, no shared memory, no
// barriers, all lanes busy even for the 8x3 windows that dominate on easy content.
//
//   k_fp       one warp per (b64, list, ref): window sizing (integer_search_b64,
//              motion_estimation.c:1296-1561, check_00_center :1139, 8x8-variance probe :1393-1441) and,
//              when the window is small, the whole scan (open_loop_me_fullpel_search_sblock :781,
//              svt_ext_all_sad_calculation_8x8_16x16 :335, svt_ext_eight_sad_calculation_32x32_64x64
//              :369, tails :98-205). Larger windows are cut into chunks of passes and queued.
//   k_fp_big   persistent warps pull (picture, job, chunk) units and merge with 64-bit atomicMin.
//
// Keys: 32-bit (sad << 12 | index), index 0 = the variance probe (evaluated first in the reference and
// never reset, so it wins ties), index 1 + y*sa_w + x for window positions; min() = first minimum in
// raster order = the reference's strict '<'. SADs stay un-doubled (SUB_SAD_SEARCH doubles all alike).
#pragma once
#include "me_device.cuh"
#include "me_search.cuh"

#define FP_WARPS 4
#define FP_IDX_BITS 12 /* 20-bit SAD (64x64x255) + 12-bit index = 32 bits; chunk-local indices stay below 256 + 2*sa_w */
#define FP_INLINE_PASSES 48 /* windows up to this many octet passes are scanned by the planning warp itself */
#define FP_CHUNK_PASSES 32

__constant__ uint8_t c_z16b[16] = {0, 1, 4, 5, 2, 3, 6, 7, 8, 9, 12, 13, 10, 11, 14, 15}; // motion_estimation.c:341

template <bool SUB>
struct FpLane {
    static constexpr int NR = SUB ? 4 : 8; // rows of the band this lane touches
    uint32_t src[NR][4];                   // its 16 x NR piece of the source block
    uint32_t kl, kr, k16, k32, k64;        // running bests (packed keys) of the PUs it may own
    int      strip, band;
};

template <bool SUB>
__device__ __forceinline__ void fp_lane_init(FpLane<SUB> &L, const uint8_t *src_full, int src_stride) {
    const int lane = threadIdx.x & 31;
    L.strip = lane & 3;
    L.band  = lane >> 2;
    const uint8_t *p = src_full + (size_t)(8 * L.band) * src_stride + 16 * L.strip;
#pragma unroll
    for (int q = 0; q < FpLane<SUB>::NR; q++)
#pragma unroll
        for (int k = 0; k < 4; k++) L.src[q][k] = ldg_u32_unaligned(p + (size_t)(SUB ? 2 * q : q) * src_stride + 4 * k);
    L.kl = L.kr = L.k16 = L.k32 = L.k64 = 0xffffffffu;
}

// One pass: positions (x .. x+7, y) given by `ref` = address of the reference sample that aligns with
// the block's top-left sample for position x. pmask: which of the 8 positions exist. idx0: key index
// of position x. Outputs the raw sums of position 0 (for the probe / KAT) when wanted.
template <bool SUB>
__device__ __forceinline__ void fp_eval_pass(FpLane<SUB> &L, const uint8_t *ref, int ref_stride, uint32_t pmask, uint32_t idx0,
                                             uint32_t *raw /* [5]: l, r, 16, 32, 64 of position 0, or nullptr */) {
    constexpr int NR = FpLane<SUB>::NR;
    const uint8_t  *g  = ref + (size_t)(8 * L.band) * ref_stride + 16 * L.strip;
    const uintptr_t a0 = (uintptr_t)g;
    const uint32_t  sa = (uint32_t)(a0 & 3) * 8; // strides are multiples of 4: same misalignment on every row
    const uint32_t *q0 = (const uint32_t *)(a0 & ~(uintptr_t)3);
    const int       rs4 = (SUB ? 2 : 1) * (ref_stride >> 2);
    uint32_t gw[NR][7];
#pragma unroll
    for (int q = 0; q < NR; q++)
#pragma unroll
        for (int k = 0; k < 7; k++) gw[q][k] = __ldg(q0 + (size_t)q * rs4 + k);
    uint32_t accl[8] = {0, 0, 0, 0, 0, 0, 0, 0}, accr[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#pragma unroll
    for (int q = 0; q < NR; q++) {
        uint32_t w[6];
#pragma unroll
        for (int k = 0; k < 6; k++) w[k] = __funnelshift_r(gw[q][k], gw[q][k + 1], sa);
        strip16_sad8_split(w, L.src[q], accl, accr);
    }
    const uint32_t mul = 1u << FP_IDX_BITS;
#pragma unroll
    for (int p = 0; p < 8; p++) {
        const bool     ok  = (pmask >> p) & 1;
        const uint32_t id  = idx0 + p;
        const uint32_t s8  = accl[p] + accr[p];
        const uint32_t t16 = s8 + __shfl_xor_sync(0xffffffffu, s8, 4);       // the other 8-row band of the 16x16
        uint32_t       t32 = t16 + __shfl_xor_sync(0xffffffffu, t16, 1);     // the other strip of the 32x32
        t32 += __shfl_xor_sync(0xffffffffu, t32, 8);                         // the other 16-row band
        uint32_t t64 = t32 + __shfl_xor_sync(0xffffffffu, t32, 2);
        t64 += __shfl_xor_sync(0xffffffffu, t64, 16);
        if (ok) {
            L.kl  = min(L.kl, accl[p] * mul + id);
            L.kr  = min(L.kr, accr[p] * mul + id);
            L.k16 = min(L.k16, t16 * mul + id);
            L.k32 = min(L.k32, t32 * mul + id);
            L.k64 = min(L.k64, t64 * mul + id);
        }
        if (p == 0 && raw) { raw[0] = accl[0]; raw[1] = accr[0]; raw[2] = t16; raw[3] = t32; raw[4] = t64; }
    }
}

// visit the PUs this lane owns: f(key, pu)
template <bool SUB, typename F>
__device__ __forceinline__ void fp_for_owned(const FpLane<SUB> &L, F f) {
    const int z = c_z16b[4 * (L.band >> 1) + L.strip];
    const int k = 2 * (L.band & 1);
    f(L.kl, 21 + 4 * z + k);
    f(L.kr, 21 + 4 * z + k + 1);
    if ((L.band & 1) == 0) f(L.k16, 5 + z);
    if ((L.band & 3) == 0 && (L.strip & 1) == 0) f(L.k32, 1 + 2 * (L.band >> 2) + (L.strip >> 1));
    if (L.band == 0 && L.strip == 0) f(L.k64, 0);
}

template <bool SUB>
__device__ __forceinline__ void fp_scan_passes(FpLane<SUB> &L, const DevPlane &ref, int org_x, int org_y, int x_origin,
                                               int y_origin, int sa_w, int sa_h, int pass0, int pass1, uint32_t idx_base) {
    const int n_oct = (sa_w + 7) >> 3;
    int y = pass0 / n_oct, oct = pass0 - y * n_oct;
    const uint8_t *row = ref.buf + (size_t)(ref.org_y + org_y + y_origin + y) * ref.stride + (ref.org_x + org_x + x_origin);
    for (int ps = pass0; ps < pass1; ps++) {
        const int      x     = 8 * oct;
        const int      left  = sa_w - x;
        const uint32_t pmask = left >= 8 ? 0xffu : ((1u << left) - 1u);
        fp_eval_pass<SUB>(L, row + x, ref.stride, pmask, idx_base + (uint32_t)(y * sa_w + x), nullptr);
        if (++oct == n_oct) { oct = 0; y++; row += ref.stride; }
    }
    (void)sa_h;
}

// ---- k_fp ------------------------------------------------------------------------------------
// phase: 0 all refs, 1 only (0,0), 2 all but (0,0) (enable_me_sr_adjustment == 2, :1357-1366)
template <bool SUB>
__global__ void __launch_bounds__(32 * FP_WARPS) k_fp(const PicJob *__restrict__ jobs, int phase) {
    const PicJob &job = jobs[blockIdx.y];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int u = blockIdx.x * FP_WARPS + warp;
    if (u >= job.n_b64 * 8) return;
    const int b = u >> 3, l = (u >> 2) & 1, r = u & 3;
    const SvtMePicParams &pp = job.pp;
    const SvtMeCtrls     &c  = pp.ctrls;
    FpJobG &fj = job.fp_jobs[u];
    if (l >= pp.num_of_list_to_search || r >= pp.num_of_ref_pic_to_search[l]) {
        if (lane == 0 && phase != 2) { fj.active = 0; job.windows[u] = SvtMeWindow{0, 0, 0, 0}; }
        return;
    }
    if (phase == 1 && (l || r)) return;
    if (phase == 2 && !(l || r)) return;
    const SbState &st = job.sb_state[b];
    unsigned long long *gk = job.fp_keys + (size_t)u * 85;
    if (!st.do_ref[l][r]) { // :1294-1295
        if (lane == 0) { fj.active = 0; job.windows[u] = SvtMeWindow{0, 0, 0, 0}; }
        return;
    }
    const B64Geo    g   = b64_geo(job, b);
    const DevPlane &ref = job.ref[l][r].lvl[0];
    const DevPlane &sf  = job.src.lvl[0];
    const uint8_t  *src_full = sf.buf + (size_t)(sf.org_y + g.org_y) * sf.stride + sf.org_x + g.org_x;
    unsigned long long ops = 0;

    // window sizing :1296-1366 (all lanes compute the same scalars; SADs are warp-cooperative)
    int16_t xc = st.hme_sc_x[l][r], yc = st.hme_sc_y[l][r];
    int16_t sa_w = (int16_t)c.me_sa.sa_min.width, sa_h = (int16_t)c.me_sa.sa_min.height;
    const uint16_t dist = scaled_picture_distance(ref_distance(pp, l, r));
    sa_w = (int16_t)imin((sa_w * dist), c.me_sa.sa_max.width);
    sa_h = (int16_t)imin((sa_h * dist), c.me_sa.sa_max.height);
    if (c.mv_sa_adj_enabled && (!c.mv_sa_adj_nearest_ref_only || r == 0)) {
        if (iabs(xc) > c.mv_sa_adj_mv_size_th) sa_w = (int16_t)(sa_w * c.mv_sa_adj_sa_multiplier);
        if (iabs(yc) > c.mv_sa_adj_mv_size_th) sa_h = (int16_t)(sa_h * c.mv_sa_adj_sa_multiplier);
    }
    const uint32_t div = st.reduce_me_sr_divisor[l][r];
    sa_w = (int16_t)((max(1u, ((uint32_t)sa_w / div)) + 7) & ~0x07u);
    sa_h = (int16_t)max(3u, ((uint32_t)sa_h / div));
    const int16_t sa_h0 = sa_h, sa_w0 = sa_w;
    uint64_t best_hme_sad = ~0ull;
    if (c.me_early_exit_th) {
        if (st.zz_sad[l][r] < (c.me_early_exit_th / 6)) sa_w = sa_h = 1;
    } else {
        uint8_t hme_is_accurate = 1;
        if ((xc != 0 || yc != 0) && pp.is_ref) {
            // check_00_center :1139-1206 (the COST_PRECISION shift is monotonic)
            const int16_t ox = (int16_t)g.org_x, oy = (int16_t)g.org_y, pad = 63;
            const uint8_t *p0 = ref.buf + ((int16_t)ref.org_x + ox + ((int16_t)ref.org_y + oy) * ref.stride);
            const uint32_t zero_mv_sad = warp_nxm_sad2(src_full, sf.stride << 1, p0, ref.stride << 1, g.b64_h >> 1, g.b64_w) << 1;
            xc = ((ox + xc) < -pad) ? (int16_t)(-pad - ox) : xc;
            xc = ((ox + xc) > (int16_t)ref.width - 1) ? (int16_t)(xc - ((ox + xc) - ((int16_t)ref.width - 1))) : xc;
            yc = ((oy + yc) < -pad) ? (int16_t)(-pad - oy) : yc;
            yc = ((oy + yc) > (int16_t)ref.height - 1) ? (int16_t)(yc - ((oy + yc) - ((int16_t)ref.height - 1))) : yc;
            const uint8_t *p1 = ref.buf + ((int16_t)(ref.org_x + ox) + xc + ((int16_t)(ref.org_y + oy) + yc) * ref.stride);
            const uint32_t hme_mv_sad = warp_nxm_sad2(src_full, sf.stride << 1, p1, ref.stride << 1, g.b64_h >> 1, g.b64_w) << 1;
            const uint64_t zero_cost = (uint64_t)(zero_mv_sad << 8), hme_cost = (uint64_t)(hme_mv_sad << 8);
            if ((zero_cost < hme_cost ? zero_cost : hme_cost) == zero_cost) xc = yc = 0;
            best_hme_sad = hme_mv_sad;
            if (xc == 0 && yc == 0) hme_is_accurate = 0;
            ops += 2ull * (unsigned long long)(g.b64_h >> 1) * g.b64_w;
        }
        if (c.enable_me_sr_adjustment == 2) {
            if ((hme_is_accurate && (best_hme_sad < (24 * 24))) || (pp.is_ref && st.sr_hme_sad[l][r] < (24 * 24)))
                sa_h = (int16_t)(sa_h / 2);
            const uint32_t best00 = (uint32_t)(job.fp_keys[(size_t)b * 8 * 85] >> 32) << (SUB ? 1 : 0); // p_sb_best_sad[0][0][0]
            if ((l || r) && best00 < 5000 && sa_h == sa_h0 && sa_w == sa_w0) {
                sa_h = (int16_t)(sa_h >> 1);
                sa_w = (int16_t)(sa_w >> 1);
            }
        }
    }
    FpLane<SUB> L;
    fp_lane_init<SUB>(L, src_full, sf.stride);
    const bool probe = c.me_8x8_var_enabled && (sa_w * sa_h > 24);
    if (probe) { // :1393-1441: one position at the search centre, then the variance of its 64 8x8 SADs
        uint32_t raw[5];
        const uint8_t *pc = ref.buf + (size_t)(ref.org_y + g.org_y + yc) * ref.stride + (ref.org_x + g.org_x + xc);
        fp_eval_pass<SUB>(L, pc, ref.stride, 1u, 0u, raw);
        const uint32_t dbl  = SUB ? 1 : 0;
        const uint32_t mean = (raw[4] << dbl) / 64;
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
    const int n_oct = (sa_w + 7) >> 3, passes = n_oct * sa_h;
    const bool inline_scan = passes <= FP_INLINE_PASSES;
    if (inline_scan) fp_scan_passes<SUB>(L, ref, g.org_x, g.org_y, x_origin, y_origin, sa_w, sa_h, 0, passes, 1u);
    // running bests -> 64-bit keys (sad << 32 | scan index); ~0 = untouched (MAX_SAD_VALUE, :1368)
    fp_for_owned<SUB>(L, [&](uint32_t k, int pu) {
        gk[pu] = k == 0xffffffffu ? ~0ull : (((unsigned long long)(k >> FP_IDX_BITS) << 32) | (k & ((1u << FP_IDX_BITS) - 1)));
    });
    if (lane == 0) {
        const int n_chunks = inline_scan ? 0 : (passes + FP_CHUNK_PASSES - 1) / FP_CHUNK_PASSES;
        fj.xc = xc; fj.yc = yc; fj.x_origin = x_origin; fj.y_origin = y_origin; fj.sa_w = sa_w; fj.sa_h = sa_h;
        fj.n_tx = (uint16_t)n_chunks; fj.n_ty = 1; fj.active = 1; fj.probe = probe ? 1 : 0;
        job.windows[u] = SvtMeWindow{x_origin, y_origin, sa_w, sa_h};
        if (n_chunks) {
            const uint32_t base = atomicAdd(&job.fp_counters[0], (uint32_t)n_chunks);
            for (int i = 0; i < n_chunks && base + i < job.fp_units_cap; i++)
                job.fp_units[base + i] = make_uint2((uint32_t)u, (blockIdx.y << 16) | (uint32_t)i);
        }
        ops += ((unsigned long long)sa_w * sa_h + (probe ? 1 : 0)) * 64ull * (SUB ? 32 : 64);
        if (job.sad_ops) atomicAdd(&job.sad_ops[1], ops);
    }
}

// ---- k_fp_big --------------------------------------------------------------------------------
template <bool SUB>
__global__ void __launch_bounds__(32 * FP_WARPS) k_fp_big(const PicJob *__restrict__ jobs) {
    const PicJob &job0 = jobs[0]; // the unit queue is shared by the batch
    const int lane = threadIdx.x & 31;
    const uint32_t n_units = min(job0.fp_counters[0], job0.fp_units_cap);
    for (;;) {
        uint32_t ui = 0;
        if (lane == 0) ui = atomicAdd(&job0.fp_counters[1], 1u);
        ui = __shfl_sync(0xffffffffu, ui, 0);
        if (ui >= n_units) break;
        const uint2 unit = job0.fp_units[ui];
        const PicJob &job = jobs[unit.y >> 16];
        const int u = (int)unit.x, chunk = (int)(unit.y & 0xffff);
        const FpJobG &fj = job.fp_jobs[u];
        const int b = u >> 3, l = (u >> 2) & 1, r = u & 3;
        const B64Geo    g   = b64_geo(job, b);
        const DevPlane &ref = job.ref[l][r].lvl[0];
        const DevPlane &sf  = job.src.lvl[0];
        FpLane<SUB> L;
        fp_lane_init<SUB>(L, sf.buf + (size_t)(sf.org_y + g.org_y) * sf.stride + sf.org_x + g.org_x, sf.stride);
        const int n_oct = (fj.sa_w + 7) >> 3, passes = n_oct * fj.sa_h;
        const int p0 = chunk * FP_CHUNK_PASSES, p1 = imin(passes, p0 + FP_CHUNK_PASSES);
        // indices inside the chunk start at the chunk's first position so that they fit the 12-bit field
        const int y_first = p0 / n_oct;
        const uint32_t gbase = 1u + (uint32_t)(y_first * fj.sa_w);
        fp_scan_passes<SUB>(L, ref, g.org_x, g.org_y, fj.x_origin, fj.y_origin + y_first, fj.sa_w, fj.sa_h - y_first,
                            p0 - y_first * n_oct, p1 - y_first * n_oct, 0u);
        unsigned long long *gk = job.fp_keys + (size_t)u * 85;
        fp_for_owned<SUB>(L, [&](uint32_t k, int pu) {
            if (k != 0xffffffffu)
                atomicMin(&gk[pu], ((unsigned long long)(k >> FP_IDX_BITS) << 32) | (gbase + (k & ((1u << FP_IDX_BITS) - 1))));
        });
    }
}
