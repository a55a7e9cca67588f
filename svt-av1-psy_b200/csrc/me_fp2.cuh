// me_fp2.cuh — full-pel (integer) search of the v2 pipeline.
//
//   k_fp_plan    one warp per (b64, list, ref): everything integer_search_b64 does before the window
//                scan (motion_estimation.c:1296-1561): window size from the HME results,
//                check_00_center (:1139), the 8x8-SAD-variance probe at the search centre (:1393-1441),
//                clamping. Emits the job, seeds the 85 running bests with the probe (scan index 0) and
//                appends the window's tiles to a unit queue.
//   k_fp_search  persistent warps pull (job, tile) units. A tile is FP_TILE_W x FP_TILE_H positions;
//                the warp stages the tile (+63 rows / +64 bytes the 64x64 block reaches) in its own
//                shared memory, 4-byte aligned rows. Task = 8 consecutive positions x one 16-pixel,
//                64-row strip of the block; the 4 strips of a position octet sit in adjacent lanes.
//                A strip yields 16 8x8 + 4 16x16 SADs, 32x32 / 64x64 are completed by shfl_xor across
//                the strip lanes (open_loop_me_fullpel_search_sblock :781, svt_ext_all_sad_calculation_
//                8x8_16x16 :335, svt_ext_eight_sad_calculation_32x32_64x64 :369 and the tails :98-205).
//                Each PU keeps a packed 32-bit key (sad << 7 | position inside the tile); min() is the
//                first minimum in raster order. Tiles merge with 64-bit atomicMin on
//                (sad << 32 | 1 + y*sa_w + x); the probe is index 0 and so wins ties, as in the reference
//                where it is evaluated first and the bests are not reset afterwards.
// SADs stay un-doubled while searching (SUB_SAD_SEARCH doubles all alike) and are doubled by k_post.
#pragma once
#include "me_device.cuh"
#include "me_search.cuh"

#define FP_WARPS 4
#define FP_WIN_ROWS (FP_TILE_H + 63)
#define FP_WIN_STRIDE 18 /* words: 8 positions + 64 bytes */
#define FP_IDX_BITS 7

__constant__ uint8_t c_z16b[16] = {0, 1, 4, 5, 2, 3, 6, 7, 8, 9, 12, 13, 10, 11, 14, 15}; // motion_estimation.c:341

struct FpKeys {
    uint32_t k8[4][2][2], k16[4], k32[2], k64;
};

// stage the 64x64 source block: row `row` at src_sm[(row / row_step) * 16 ..]
__device__ inline void warp_stage_src64(uint32_t *src_sm, const uint8_t *src, int stride, int row_step) {
    const int lane = threadIdx.x & 31;
    const int n = (64 / row_step) * 16;
    for (int i = lane; i < n; i += 32) src_sm[i] = ldg_u32_unaligned(src + (size_t)(i >> 4) * row_step * stride + 4 * (i & 15));
}

// One tile: positions (x, yy), x in [0,tw) (tw <= 8), yy in [0,th) (th <= FP_TILE_H). win rows are the
// staged reference rows starting at the tile's first position; src_sm as above.
__device__ inline void warp_fp_tile(const uint32_t *win, const uint32_t *src_sm, int tw, int th, bool sub, FpKeys &K) {
    const int lane  = threadIdx.x & 31;
    const int strip = lane & 3;
    const uint32_t mul = 1u << FP_IDX_BITS;
#pragma unroll
    for (int a = 0; a < 4; a++) {
        K.k16[a] = 0xffffffffu;
#pragma unroll
        for (int h = 0; h < 2; h++) K.k8[a][h][0] = K.k8[a][h][1] = 0xffffffffu;
    }
    K.k32[0] = K.k32[1] = K.k64 = 0xffffffffu;
    const int n_task = th * 4;
    for (int base = 0; base < n_task; base += 32) {
        const int  t     = base + lane;
        const bool valid = t < n_task;
        const int  yy    = valid ? (t >> 2) : 0;
        const uint32_t idx0 = (uint32_t)(yy * 8);
        const uint32_t pmask = valid ? ((1u << tw) - 1u) : 0u;
        const uint32_t *wbase = win + yy * FP_WIN_STRIDE + 4 * strip;
        uint32_t s64[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        uint32_t s32[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#pragma unroll
        for (int b16 = 0; b16 < 4; b16++) {
            uint32_t s16[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#pragma unroll
            for (int half = 0; half < 2; half++) {
                uint32_t accl[8] = {0, 0, 0, 0, 0, 0, 0, 0}, accr[8] = {0, 0, 0, 0, 0, 0, 0, 0};
                const int row0 = 16 * b16 + 8 * half;
                const int nq = sub ? 4 : 8, rs = sub ? 2 : 1;
#pragma unroll
                for (int q = 0; q < 8; q++) {
                    if (q < nq) {
                        const int       row = row0 + rs * q;
                        const uint32_t *rp  = wbase + row * FP_WIN_STRIDE;
                        const uint2 a = *(const uint2 *)(rp), b = *(const uint2 *)(rp + 2), cc = *(const uint2 *)(rp + 4);
                        const uint4 sv = *(const uint4 *)(src_sm + (row / rs) * 16 + 4 * strip);
                        const uint32_t w[6] = {a.x, a.y, b.x, b.y, cc.x, cc.y};
                        const uint32_t s[4] = {sv.x, sv.y, sv.z, sv.w};
                        strip16_sad8_split(w, s, accl, accr);
                    }
                }
#pragma unroll
                for (int p = 0; p < 8; p++) {
                    const bool ok = (pmask >> p) & 1;
                    K.k8[b16][half][0] = min(K.k8[b16][half][0], ok ? accl[p] * mul + (idx0 + p) : 0xffffffffu);
                    K.k8[b16][half][1] = min(K.k8[b16][half][1], ok ? accr[p] * mul + (idx0 + p) : 0xffffffffu);
                    s16[p] += accl[p] + accr[p];
                }
            }
#pragma unroll
            for (int p = 0; p < 8; p++) {
                const bool ok = (pmask >> p) & 1;
                K.k16[b16] = min(K.k16[b16], ok ? s16[p] * mul + (idx0 + p) : 0xffffffffu);
                s32[p] += s16[p];
            }
            if (b16 & 1) {
#pragma unroll
                for (int p = 0; p < 8; p++) {
                    const bool     ok = (pmask >> p) & 1;
                    const uint32_t v  = s32[p] + __shfl_xor_sync(0xffffffffu, s32[p], 1);
                    K.k32[b16 >> 1]   = min(K.k32[b16 >> 1], ok ? v * mul + (idx0 + p) : 0xffffffffu);
                    s64[p] += v;
                    s32[p] = 0;
                }
            }
        }
#pragma unroll
        for (int p = 0; p < 8; p++) {
            const bool     ok = (pmask >> p) & 1;
            const uint32_t v  = s64[p] + __shfl_xor_sync(0xffffffffu, s64[p], 2);
            K.k64 = min(K.k64, ok ? v * mul + (idx0 + p) : 0xffffffffu);
        }
    }
}

// visit every (key, PU index) this lane's strip owns, after reducing over the lanes of the same strip
template <typename F>
__device__ __forceinline__ void fp_for_each_key(const FpKeys &K, F f) {
    const int      lane = threadIdx.x & 31, strip = lane & 3;
    const uint32_t mask = 0x11111111u << strip;
#pragma unroll
    for (int b16 = 0; b16 < 4; b16++) {
        const int z = c_z16b[4 * b16 + strip];
#pragma unroll
        for (int half = 0; half < 2; half++) {
            f(__reduce_min_sync(mask, K.k8[b16][half][0]), 21 + 4 * z + 2 * half + 0);
            f(__reduce_min_sync(mask, K.k8[b16][half][1]), 21 + 4 * z + 2 * half + 1);
        }
        f(__reduce_min_sync(mask, K.k16[b16]), 5 + z);
    }
    f(__reduce_min_sync(mask, K.k32[0]), 1 + (strip >> 1));
    f(__reduce_min_sync(mask, K.k32[1]), 3 + (strip >> 1));
    f(__reduce_min_sync(mask, K.k64), 0);
}

// ---- k_fp_plan -------------------------------------------------------------------------------
// phase: 0 all refs, 1 only (0,0), 2 all but (0,0) (enable_me_sr_adjustment == 2, :1357-1366)
__global__ void __launch_bounds__(32 * FP_WARPS) k_fp_plan(const PicJob *__restrict__ jobs, int phase) {
    const PicJob &job = jobs[blockIdx.y];
    __shared__ __align__(16) uint32_t s_src[FP_WARPS][64 * 16];
    __shared__ __align__(16) uint32_t s_win[FP_WARPS][64 * FP_WIN_STRIDE];
    __shared__ uint32_t s_probe[FP_WARPS][85];
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
    const bool      sub = c.me_search_method == SVTME_SUB_SAD_SEARCH;
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
            const uint32_t best00 = (uint32_t)(job.fp_keys[(size_t)b * 8 * 85] >> 32) << (sub ? 1 : 0); // p_sb_best_sad[0][0][0]
            if ((l || r) && best00 < 5000 && sa_h == sa_h0 && sa_w == sa_w0) {
                sa_h = (int16_t)(sa_h >> 1);
                sa_w = (int16_t)(sa_w >> 1);
            }
        }
    }
    // the running bests start at MAX_SAD_VALUE (:1368); ~0 keys mean "nothing yet"
    for (int i = lane; i < 85; i += 32) s_probe[warp][i] = 0xffffffffu;
    const bool probe = c.me_8x8_var_enabled && (sa_w * sa_h > 24);
    if (probe) { // :1393-1441
        warp_stage_src64(s_src[warp], src_full, sf.stride, sub ? 2 : 1);
        const uint8_t *pc = ref.buf + (size_t)(ref.org_y + g.org_y + yc) * ref.stride + (ref.org_x + g.org_x + xc);
        stage_rect(s_win[warp], FP_WIN_STRIDE, pc, ref.stride, 64, FP_WIN_STRIDE, lane, 32);
        __syncwarp();
        FpKeys K;
        warp_fp_tile(s_win[warp], s_src[warp], 1, 1, sub, K);
        fp_for_each_key(K, [&](uint32_t m, int pu) {
            if (lane < 4) s_probe[warp][pu] = m >> FP_IDX_BITS; // un-doubled SAD of the probe position
        });
        __syncwarp();
        const uint32_t dbl  = sub ? 1 : 0;
        const uint32_t mean = (s_probe[warp][0] << dbl) / 64;
        uint32_t sumsq = 0;
        for (int i = 0; i < 64; i++) {
            const int32_t d = (int32_t)(s_probe[warp][21 + i] << dbl) - (int32_t)mean;
            sumsq += (uint32_t)(d * d);
        }
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
    for (int i = lane; i < 85; i += 32) {
        const uint32_t v = s_probe[warp][i];
        gk[i] = v == 0xffffffffu ? ~0ull : ((unsigned long long)v << 32);
    }
    // final window :1442-1561 (restricted_motion_vector == 0)
    int16_t x_origin = (int16_t)(xc - (sa_w >> 1));
    int16_t y_origin = (int16_t)(yc - (sa_h >> 1));
    clamp_axis((int16_t)g.org_x, x_origin, sa_w, 63, (int16_t)job.pic_w);
    sa_w = (int16_t)((sa_w < 8) ? sa_w : sa_w & ~0x07);
    clamp_axis((int16_t)g.org_y, y_origin, sa_h, 63, (int16_t)job.pic_h);
    if (lane == 0) {
        const int n_tx = (sa_w + FP_TILE_W - 1) / FP_TILE_W, n_ty = (sa_h + FP_TILE_H - 1) / FP_TILE_H;
        fj.xc = xc; fj.yc = yc; fj.x_origin = x_origin; fj.y_origin = y_origin; fj.sa_w = sa_w; fj.sa_h = sa_h;
        fj.n_tx = (uint16_t)n_tx; fj.n_ty = (uint16_t)n_ty; fj.active = 1; fj.probe = probe ? 1 : 0;
        job.windows[u] = SvtMeWindow{x_origin, y_origin, sa_w, sa_h};
        const uint32_t n = (uint32_t)(n_tx * n_ty);
        const uint32_t base = atomicAdd(&job.fp_counters[0], n);
        for (uint32_t i = 0; i < n && base + i < job.fp_units_cap; i++) job.fp_units[base + i] = make_uint2(blockIdx.y, ((uint32_t)u << 16) | i);
        ops += ((unsigned long long)sa_w * sa_h + (probe ? 1 : 0)) * 64ull * (sub ? 32 : 64);
        if (job.sad_ops) atomicAdd(&job.sad_ops[1], ops);
    }
}

// ---- k_fp_search -----------------------------------------------------------------------------
__global__ void __launch_bounds__(32 * FP_WARPS, 4) k_fp_search(const PicJob *__restrict__ jobs) {
    const PicJob &job0 = jobs[0]; // the unit queue is shared by the batch
    __shared__ __align__(16) uint32_t s_src[FP_WARPS][64 * 16];
    __shared__ __align__(16) uint32_t s_win[FP_WARPS][FP_WIN_ROWS * FP_WIN_STRIDE];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const bool sub = job0.pp.ctrls.me_search_method == SVTME_SUB_SAD_SEARCH;
    const uint32_t n_units = min(job0.fp_counters[0], job0.fp_units_cap);
    int cur_job = -1, cur_pic = -1;
    for (;;) {
        uint32_t ui = 0;
        if (lane == 0) ui = atomicAdd(&job0.fp_counters[1], 1u);
        ui = __shfl_sync(0xffffffffu, ui, 0);
        if (ui >= n_units) break;
        const uint2 unit = job0.fp_units[ui];
        const PicJob &job = jobs[unit.x];
        const int u = (int)(unit.y >> 16), tile = (int)(unit.y & 0xffff);
        const FpJobG &fj = job.fp_jobs[u];
        const int b = u >> 3, l = (u >> 2) & 1, r = u & 3;
        const B64Geo    g   = b64_geo(job, b);
        const DevPlane &ref = job.ref[l][r].lvl[0];
        if (u != cur_job || (int)unit.x != cur_pic) {
            cur_pic = (int)unit.x;
            const DevPlane &sf = job.src.lvl[0];
            __syncwarp();
            warp_stage_src64(s_src[warp], sf.buf + (size_t)(sf.org_y + g.org_y) * sf.stride + sf.org_x + g.org_x, sf.stride, sub ? 2 : 1);
            cur_job = u;
        }
        const int ty = tile / fj.n_tx, tx = tile - ty * fj.n_tx;
        const int x0 = tx * FP_TILE_W, y0 = ty * FP_TILE_H;
        const int tw = imin(FP_TILE_W, fj.sa_w - x0), th = imin(FP_TILE_H, fj.sa_h - y0);
        const uint8_t *g0 = ref.buf + (size_t)(ref.org_y + g.org_y + fj.y_origin + y0) * ref.stride + (ref.org_x + g.org_x + fj.x_origin + x0);
        stage_rect(s_win[warp], FP_WIN_STRIDE, g0, ref.stride, th + 63, FP_WIN_STRIDE, lane, 32);
        __syncwarp();
        FpKeys K;
        warp_fp_tile(s_win[warp], s_src[warp], tw, th, sub, K);
        unsigned long long *gk = job.fp_keys + (size_t)u * 85;
        const int sa_w = fj.sa_w;
        fp_for_each_key(K, [&](uint32_t m, int pu) {
            if (lane < 4 && m != 0xffffffffu) {
                const uint32_t loc = m & ((1u << FP_IDX_BITS) - 1);
                const uint32_t gidx = 1u + (uint32_t)((y0 + (int)(loc >> 3)) * sa_w + x0 + (int)(loc & 7));
                atomicMin(&gk[pu], ((unsigned long long)(m >> FP_IDX_BITS) << 32) | gidx);
            }
        });
        __syncwarp();
    }
}
