// me_fullpel.cuh — k_fullpel: integer (full-pel) search of one (b64, list, ref) per CTA.
//
// Replaces integer_search_b64 for one reference (motion_estimation.c:1249-1586), i.e. the window
// sizing logic, check_00_center (:1139-1206), the 8x8-SAD-variance probe (:1393-1441) and
// open_loop_me_fullpel_search_sblock (:781-817) with the RTCD kernels it drives:
// svt_ext_all_sad_calculation_8x8_16x16 (:335-362), svt_ext_eight_sad_calculation_32x32_64x64
// (:369-425) and their one-position tail variants (:98-205).
//
// Mapping: the search window (sa_w x sa_h positions, plus the 63 extra rows/columns a 64x64 block
// reaches) is staged in shared memory with 4-byte aligned rows. One task = 8 consecutive
// positions ("octet") x one 16-pixel-wide, 64-row strip of the block; the 4 strips of an octet sit
// in adjacent lanes. A strip yields the 16 8x8 and 4 16x16 SADs of its column of blocks; 32x32 and
// 64x64 SADs are completed with warp shuffles across the strip lanes. Every PU keeps a packed
// 32-bit key (sad << idx_bits | scan index inside the tile): min() over keys in any order returns
// the first minimum of the raster scan, which is the reference's strict '<' rule. Tiles (y ranges
// of at most 2^idx_bits positions) are merged through 64-bit keys in shared memory; the variance
// probe at the search centre is scan index 0, so it wins ties exactly like in the reference,
// where it is evaluated first and the running bests are not reset afterwards.
//
// SADs are kept un-doubled while searching (row sub-sampling doubles every SAD alike, which
// cannot change an arg-min) and doubled once on output.
#pragma once
#include "me_device.cuh"
#include "me_hme.cuh" // warp_nxm_sad

#define FP_THREADS 128

__constant__ uint8_t c_z16[16] = {0, 1, 4, 5, 2, 3, 6, 7, 8, 9, 12, 13, 10, 11, 14, 15}; // motion_estimation.c:341

struct FpSh {
    unsigned long long gkey[85]; // (sad << 32) | global scan index; ~0 = nothing yet
    int16_t  xc, yc;             // search centre
    int16_t  sa_w, sa_h, x_origin, y_origin;
    int      active;
    alignas(16) uint32_t src[64 * 16]; // the 64x64 source block, row-major, 16 words per row
};

// One tile of a window: positions (x, y0+yy), x in [0,sa_w), yy in [0,tile_h). win holds rows
// y0 .. y0+tile_h+62 of the staged region. gidx_base: global scan index of position (0, y0).
__device__ inline void fp_search_tile(FpSh &sh, const uint32_t *win, int stride, int sa_w, int tile_h, bool sub,
                                      int idx_bits, uint32_t gidx_base) {
    const int tid    = threadIdx.x;
    const int strip  = tid & 3;
    const int n_oct  = (sa_w + 7) >> 3;
    const int n_task = n_oct * tile_h * 4;
    const uint32_t mul = 1u << idx_bits;
    uint32_t k8[4][2][2], k16[4], k32[2], k64;
#pragma unroll
    for (int a = 0; a < 4; a++) {
        k16[a] = 0xffffffffu;
#pragma unroll
        for (int h = 0; h < 2; h++) k8[a][h][0] = k8[a][h][1] = 0xffffffffu;
    }
    k32[0] = k32[1] = k64 = 0xffffffffu;

    for (int base = 0; base < n_task; base += FP_THREADS) {
        const int  t     = base + tid;
        const bool valid = t < n_task;
        const int  oy    = valid ? (t >> 2) : 0;
        const int  yy = oy / n_oct, oct = oy - yy * n_oct;
        const uint32_t idx0 = (uint32_t)(yy * sa_w + 8 * oct);
        uint32_t pmask = 0; // bit p set: position p of the octet exists
#pragma unroll
        for (int p = 0; p < 8; p++) pmask |= (valid && (8 * oct + p) < sa_w) ? (1u << p) : 0u;
        const uint32_t *wbase = win + yy * stride + 2 * oct + 4 * strip;
        uint32_t s64[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        uint32_t s32[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#pragma unroll
        for (int b16 = 0; b16 < 4; b16++) {
            uint32_t s16[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#pragma unroll
            for (int half = 0; half < 2; half++) {
                uint32_t accl[8] = {0, 0, 0, 0, 0, 0, 0, 0}, accr[8] = {0, 0, 0, 0, 0, 0, 0, 0};
                const int row0 = 16 * b16 + 8 * half;
                if (sub) {
#pragma unroll
                    for (int q = 0; q < 4; q++) {
                        const int       row = row0 + 2 * q;
                        const uint32_t *rp  = wbase + row * stride;
                        const uint2 a = *(const uint2 *)(rp), b = *(const uint2 *)(rp + 2), cc = *(const uint2 *)(rp + 4);
                        const uint4 sv = *(const uint4 *)(sh.src + row * 16 + 4 * strip);
                        const uint32_t w[6] = {a.x, a.y, b.x, b.y, cc.x, cc.y};
                        const uint32_t s[4] = {sv.x, sv.y, sv.z, sv.w};
                        strip16_sad8_split(w, s, accl, accr);
                    }
                } else {
#pragma unroll
                    for (int q = 0; q < 8; q++) {
                        const int       row = row0 + q;
                        const uint32_t *rp  = wbase + row * stride;
                        const uint2 a = *(const uint2 *)(rp), b = *(const uint2 *)(rp + 2), cc = *(const uint2 *)(rp + 4);
                        const uint4 sv = *(const uint4 *)(sh.src + row * 16 + 4 * strip);
                        const uint32_t w[6] = {a.x, a.y, b.x, b.y, cc.x, cc.y};
                        const uint32_t s[4] = {sv.x, sv.y, sv.z, sv.w};
                        strip16_sad8_split(w, s, accl, accr);
                    }
                }
#pragma unroll
                for (int p = 0; p < 8; p++) {
                    const bool ok = (pmask >> p) & 1;
                    const uint32_t kl = ok ? accl[p] * mul + (idx0 + p) : 0xffffffffu;
                    const uint32_t kr = ok ? accr[p] * mul + (idx0 + p) : 0xffffffffu;
                    k8[b16][half][0]  = min(k8[b16][half][0], kl);
                    k8[b16][half][1]  = min(k8[b16][half][1], kr);
                    s16[p] += accl[p] + accr[p];
                }
            }
#pragma unroll
            for (int p = 0; p < 8; p++) {
                const bool ok = (pmask >> p) & 1;
                k16[b16]      = min(k16[b16], ok ? s16[p] * mul + (idx0 + p) : 0xffffffffu);
                s32[p] += s16[p];
            }
            if (b16 & 1) {
#pragma unroll
                for (int p = 0; p < 8; p++) {
                    const bool     ok = (pmask >> p) & 1;
                    const uint32_t v  = s32[p] + __shfl_xor_sync(0xffffffffu, s32[p], 1);
                    k32[b16 >> 1]     = min(k32[b16 >> 1], ok ? v * mul + (idx0 + p) : 0xffffffffu);
                    s64[p] += v;
                    s32[p] = 0;
                }
            }
        }
#pragma unroll
        for (int p = 0; p < 8; p++) {
            const bool     ok = (pmask >> p) & 1;
            const uint32_t v  = s64[p] + __shfl_xor_sync(0xffffffffu, s64[p], 2);
            k64               = min(k64, ok ? v * mul + (idx0 + p) : 0xffffffffu);
        }
    }
    // reduce over the lanes that hold the same strip, then merge into the window-global keys
    const uint32_t lane_mask = 0x11111111u << strip;
    const uint32_t imask     = mul - 1;
    const bool     leader    = (tid & 31) < 4;
#define FP_COMMIT(key, pu)                                                                                        \
    {                                                                                                             \
        const uint32_t m_ = __reduce_min_sync(lane_mask, key);                                                    \
        if (leader && m_ != 0xffffffffu)                                                                          \
            atomicMin(&sh.gkey[pu], ((unsigned long long)(m_ >> idx_bits) << 32) | (gidx_base + (m_ & imask)));   \
    }
#pragma unroll
    for (int b16 = 0; b16 < 4; b16++) {
        const int z = c_z16[4 * b16 + strip];
#pragma unroll
        for (int half = 0; half < 2; half++) {
            FP_COMMIT(k8[b16][half][0], 21 + 4 * z + 2 * half + 0);
            FP_COMMIT(k8[b16][half][1], 21 + 4 * z + 2 * half + 1);
        }
        FP_COMMIT(k16[b16], 5 + z);
    }
    FP_COMMIT(k32[0], 1 + (strip >> 1));
    FP_COMMIT(k32[1], 3 + (strip >> 1));
    FP_COMMIT(k64, 0);
#undef FP_COMMIT
}

// Search a whole window (tiles in y). Positions get global scan indices gidx0 + y*sa_w + x.
__device__ inline void fp_search_window(FpSh &sh, uint32_t *win, int win_words, const DevPlane &ref, int org_x, int org_y,
                                        int x_origin, int y_origin, int sa_w, int sa_h, bool sub, uint32_t gidx0) {
    const int n_oct    = (sa_w + 7) >> 3;
    const int stride   = 2 * n_oct + 16; // words: 8*n_oct + 64 bytes per staged row
    const int idx_bits = sub ? 13 : 12;
    int tile_h = win_words / stride - 63;
    tile_h     = imin(tile_h, (1 << idx_bits) / imax(1, sa_w));
    if (tile_h < 1) tile_h = 1;
    const uint8_t *g0 = ref.buf + (size_t)(ref.org_y + org_y + y_origin) * ref.stride + (ref.org_x + org_x + x_origin);
    for (int y0 = 0; y0 < sa_h; y0 += tile_h) {
        const int th = imin(tile_h, sa_h - y0);
        __syncthreads();
        stage_rect(win, stride, g0 + (size_t)y0 * ref.stride, ref.stride, th + 63, stride, threadIdx.x, FP_THREADS);
        __syncthreads();
        fp_search_tile(sh, win, stride, sa_w, th, sub, idx_bits, gidx0 + (uint32_t)(y0 * sa_w));
    }
    __syncthreads();
}

__global__ void __launch_bounds__(FP_THREADS) k_fullpel(const __grid_constant__ PicJob job) {
    extern __shared__ __align__(16) uint32_t fp_win[];
    __shared__ __align__(16) FpSh sh;
    const SvtMePicParams &pp  = job.pp;
    const SvtMeCtrls     &c   = pp.ctrls;
    const int             tid = threadIdx.x;
    const int             l = blockIdx.x >> 2, r = blockIdx.x & 3, b = blockIdx.y;
    if (l >= pp.num_of_list_to_search || r >= pp.num_of_ref_pic_to_search[l]) return;
    if (job.fullpel_phase == 1 && (l || r)) return;
    if (job.fullpel_phase == 2 && !(l || r)) return;
    const int org_x = (b % job.b64_cols) * 64, org_y = (b / job.b64_cols) * 64;
    const int b64_w = imin(64, job.pic_w - org_x), b64_h = imin(64, job.pic_h - org_y);
    const bool sub = c.me_search_method == SVTME_SUB_SAD_SEARCH;
    const SbState  &st  = job.sb_state[b];
    const DevPlane &ref = job.ref[l][r].lvl[0];
    const DevPlane &sf  = job.src.lvl[0];
    const size_t    out = ((size_t)b * 2 + l) * 4 + r;
    const uint8_t  *src_full = sf.buf + (size_t)(sf.org_y + org_y) * sf.stride + sf.org_x + org_x;
    const int       win_words = job.fp_smem_bytes / 4;

    if (!st.do_ref[l][r]) { // motion_estimation.c:1294-1295
        if (tid == 0) job.windows[out] = SvtMeWindow{0, 0, 0, 0};
        return;
    }
    for (int i = tid; i < 85; i += FP_THREADS) sh.gkey[i] = ~0ull;
    for (int i = tid; i < 64 * 16; i += FP_THREADS)
        sh.src[i] = ldg_u32_unaligned(src_full + (size_t)(i >> 4) * sf.stride + 4 * (i & 15));

    // ---- window sizing: warp 0, scalar logic on lane 0 (motion_estimation.c:1296-1366) ----
    int16_t sa_w = 0, sa_h = 0, xc = 0, yc = 0;
    bool    probe = false;
    unsigned long long ops = 0; // thread 0
    if (tid < 32) {
        xc   = st.hme_sc_x[l][r];
        yc   = st.hme_sc_y[l][r];
        sa_w = (int16_t)c.me_sa.sa_min.width;
        sa_h = (int16_t)c.me_sa.sa_min.height;
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
                // check_00_center, :1139-1206 (COST_PRECISION shift is monotonic: compare the SADs)
                const int16_t ox = (int16_t)org_x, oy = (int16_t)org_y, pad = 63;
                const uint8_t *p0 = ref.buf + ((int16_t)ref.org_x + ox + ((int16_t)ref.org_y + oy) * ref.stride);
                uint32_t zero_mv_sad = warp_nxm_sad(src_full, sf.stride << 1, p0, ref.stride << 1, b64_h >> 1, b64_w) << 1;
                xc = ((ox + xc) < -pad) ? (int16_t)(-pad - ox) : xc;
                xc = ((ox + xc) > (int16_t)ref.width - 1) ? (int16_t)(xc - ((ox + xc) - ((int16_t)ref.width - 1))) : xc;
                yc = ((oy + yc) < -pad) ? (int16_t)(-pad - oy) : yc;
                yc = ((oy + yc) > (int16_t)ref.height - 1) ? (int16_t)(yc - ((oy + yc) - ((int16_t)ref.height - 1))) : yc;
                const uint8_t *p1 = ref.buf + ((int16_t)(ref.org_x + ox) + xc + ((int16_t)(ref.org_y + oy) + yc) * ref.stride);
                const uint32_t hme_mv_sad = warp_nxm_sad(src_full, sf.stride << 1, p1, ref.stride << 1, b64_h >> 1, b64_w) << 1;
                const uint64_t zero_cost = (uint64_t)(zero_mv_sad << 8), hme_cost = (uint64_t)(hme_mv_sad << 8);
                if ((zero_cost < hme_cost ? zero_cost : hme_cost) == zero_cost) xc = yc = 0;
                best_hme_sad = hme_mv_sad;
                ops += 2ull * (unsigned long long)(b64_h >> 1) * b64_w;
                if (xc == 0 && yc == 0) hme_is_accurate = 0;
            }
            if (c.enable_me_sr_adjustment == 2) {
                if ((hme_is_accurate && (best_hme_sad < (24 * 24))) || (pp.is_ref && st.sr_hme_sad[l][r] < (24 * 24)))
                    sa_h = (int16_t)(sa_h / 2);
                if ((l || r) && job.best_sad[(size_t)b * 2 * 4 * 85] < 5000 && sa_h == sa_h0 && sa_w == sa_w0) {
                    sa_h = (int16_t)(sa_h >> 1);
                    sa_w = (int16_t)(sa_w >> 1);
                }
            }
        }
        probe = c.me_8x8_var_enabled && (sa_w * sa_h > 24);
        if (tid == 0) {
            sh.xc = xc; sh.yc = yc; sh.sa_w = sa_w; sh.sa_h = sa_h;
            sh.active = probe ? 1 : 0;
        }
    }
    __syncthreads();
    xc = sh.xc; yc = sh.yc; sa_w = sh.sa_w; sa_h = sh.sa_h;
    probe = sh.active != 0;

    // ---- 8x8 SAD variance probe, :1393-1441 ----
    if (probe) {
        fp_search_window(sh, fp_win, win_words, ref, org_x, org_y, xc, yc, 1, 1, sub, 0u);
        if (tid == 0) {
            const uint32_t dbl  = sub ? 1 : 0;
            const uint32_t mean = ((uint32_t)(sh.gkey[0] >> 32) << dbl) / 64;
            uint32_t sumsq = 0;
            for (int i = 0; i < 64; i++) {
                const int32_t d = (int32_t)((uint32_t)(sh.gkey[21 + i] >> 32) << dbl) - (int32_t)mean;
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
            sh.sa_w = sa_w; sh.sa_h = sa_h;
        }
        __syncthreads();
        sa_w = sh.sa_w; sa_h = sh.sa_h;
    }

    // ---- final window, :1442-1561 (restricted_motion_vector == 0) ----
    int16_t x_origin = (int16_t)(xc - (sa_w >> 1));
    int16_t y_origin = (int16_t)(yc - (sa_h >> 1));
    clamp_axis((int16_t)org_x, x_origin, sa_w, 63, (int16_t)job.pic_w);
    sa_w = (int16_t)((sa_w < 8) ? sa_w : sa_w & ~0x07);
    clamp_axis((int16_t)org_y, y_origin, sa_h, 63, (int16_t)job.pic_h);
    if (tid == 0) {
        job.windows[out] = SvtMeWindow{x_origin, y_origin, sa_w, sa_h};
        ops += ((unsigned long long)sa_w * sa_h + (probe ? 1 : 0)) * 64ull * (sub ? 32 : 64);
        if (job.sad_ops) atomicAdd(&job.sad_ops[1], ops);
    }

    fp_search_window(sh, fp_win, win_words, ref, org_x, org_y, x_origin, y_origin, sa_w, sa_h, sub, 1u);

    // ---- outputs: p_sb_best_sad / p_sb_best_mv of this (b64, list, ref), me_context.h:395-396 ----
    for (int pu = tid; pu < 85; pu += FP_THREADS) {
        const unsigned long long k = sh.gkey[pu];
        uint32_t sad = MAX_SAD_VALUE_U32, mv = 0;
        if (k != ~0ull) {
            sad = (uint32_t)(k >> 32) << (sub ? 1 : 0);
            const uint32_t gi = (uint32_t)k;
            int16_t mx, my;
            if (gi == 0) {
                mx = xc; my = yc;
            } else {
                mx = (int16_t)((int32_t)((gi - 1) % (uint32_t)sa_w) + x_origin);
                my = (int16_t)((int32_t)((gi - 1) / (uint32_t)sa_w) + y_origin);
            }
            mv = ((uint32_t)(uint16_t)my << 16) | (uint16_t)mx;
        }
        job.best_sad[out * 85 + pu] = sad;
        job.best_mv[out * 85 + pu]  = mv;
    }
}
