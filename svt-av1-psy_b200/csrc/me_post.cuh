// me_post.cuh — k_post: per-b64 epilogue of svt_aom_motion_estimation_b64
// (motion_estimation.c:3191-3221): me_prune_ref (:1592-1635), construct_me_candidate_array and its
// two special cases (:2602-2905), compute_distortion (:3034-3077), perform_gm_detection
// (:2908-3031). One CTA per b64, one thread per PU (ME z-order index n_idx).
#pragma once
#include "me_device.cuh"

#define POST_THREADS 96

__constant__ uint8_t c_z_to_raster[85] = {/* motion_estimation.c:2590-2601 */
    0,  1,  2,  3,  4,  5,  6,  9,  10, 7,  8,  11, 12, 13, 14, 17, 18, 15, 16, 19, 20, 21, 22, 29, 30, 23, 24, 31, 32,
    37, 38, 45, 46, 39, 40, 47, 48, 25, 26, 33, 34, 27, 28, 35, 36, 41, 42, 49, 50, 43, 44, 51, 52, 53, 54, 61, 62, 55,
    56, 63, 64, 69, 70, 77, 78, 71, 72, 79, 80, 57, 58, 65, 66, 59, 60, 67, 68, 73, 74, 81, 82, 75, 76, 83, 84};

// MeCandidate bit-field byte (me_sb_results.h:28-34); list fields are 1 bit wide, so the
// reference's "24" (no list) is stored as 0
__device__ __forceinline__ uint8_t cand_byte(unsigned dir, unsigned i0, unsigned i1, unsigned l0, unsigned l1) {
    return (uint8_t)((dir & 3) | ((i0 & 3) << 2) | ((i1 & 3) << 4) | ((l0 & 1) << 6) | ((l1 & 1) << 7));
}

struct PostSh {
    uint64_t hme_sad[2][4];
    uint8_t  do_ref[2][4];
    uint32_t me_distortion[85];
    uint32_t best_sad[8 * 85], best_mv[8 * 85]; // this b64's p_sb_best_sad / p_sb_best_mv: built here, read by every later step
    // this b64's MeSbResults arrays (me_sb_results.h:44-51), built here and written to HBM once, as whole words
    uint32_t mv[85 * 8];
    uint8_t  cand[85 * 32];
    uint8_t  total[88];
};

// n bytes from shared memory to global memory at any alignment: whole aligned words where the destination has them
__device__ __forceinline__ void post_store_bytes(uint8_t *dst, const uint8_t *src, uint32_t n, int tid) {
    const uint32_t head = (uint32_t)((4u - ((uintptr_t)dst & 3u)) & 3u);
    const uint32_t h = head < n ? head : n, n_words = (n - h) >> 2, tail = h + 4u * n_words;
    for (uint32_t i = tid; i < n_words; i += POST_THREADS) {
        const uint8_t *q = src + h + 4u * i;
        *(uint32_t *)(dst + h + 4u * i) = (uint32_t)q[0] | ((uint32_t)q[1] << 8) | ((uint32_t)q[2] << 16) | ((uint32_t)q[3] << 24);
    }
    if ((uint32_t)tid < h) dst[tid] = src[tid];
    if ((uint32_t)tid < n - tail) dst[tail + tid] = src[tail + tid];
}

__global__ void __launch_bounds__(POST_THREADS) k_post(const PicJob *__restrict__ jobs) {
    const PicJob &job = jobs[blockIdx.y];
    __shared__ PostSh sh;
    const SvtMePicParams &pp  = job.pp;
    const SvtMeCtrls     &c   = pp.ctrls;
    const int             tid = threadIdx.x, b = blockIdx.x;
    const SbState        &st  = job.sb_state[b];
    // result arrays the caller did not ask for are NULL (svtme_picture_me_batch_packed / SvtMePicResults members): nothing is written
    uint32_t *gbs = job.best_sad ? job.best_sad + (size_t)b * 2 * 4 * 85 : nullptr;
    uint32_t *gbm = job.best_mv ? job.best_mv + (size_t)b * 2 * 4 * 85 : nullptr;
#define BS(l, r, n) sh.best_sad[((l) * 4 + (r)) * 85 + (n)]
#define BM(l, r, n) sh.best_mv[((l) * 4 + (r)) * 85 + (n)]
    const uint8_t *nref = pp.num_of_ref_pic_to_search;
    const int n_lists_cfg = pp.num_of_list_to_search;

    // ---- full-pel keys -> p_sb_best_sad / p_sb_best_mv (me_context.h:395-396), kept in shared memory; global memory
    //      gets them once, at the end, with the entries of references that end with do_ref == 0 zeroed ----
    {
        const bool sub = c.me_search_method == SVTME_SUB_SAD_SEARCH;
        for (int i = tid; i < 8 * 85; i += POST_THREADS) {
            const int lr = i / 85, pu = i - lr * 85;
            const bool have = job.fp_keys && (lr >> 2) < n_lists_cfg && (lr & 3) < nref[lr >> 2] && job.fp_jobs[(size_t)b * 8 + lr].active;
            uint32_t sad = 0, mv = 0;
            // !have: a slot the picture does not have, or a reference that was not searched (do_ref == 0 before the full-pel
            // stage): stale in the reference, never consumed (a reference is only used while its do_ref is set), zero here
            if (have) {
                const FpJobG &fj = job.fp_jobs[(size_t)b * 8 + lr];
                const unsigned long long k = job.fp_keys[((size_t)b * 8 + lr) * 85 + pu];
                sad = MAX_SAD_VALUE_U32; // initial bests, :1368 and :3119
                mv  = 0;
                if (k != ~0ull) {
                    sad = (uint32_t)(k >> 32) << (sub ? 1 : 0);
                    const uint32_t gi = (uint32_t)k;
                    int16_t mx, my;
                    if (gi == 0) { // the variance probe at the search centre
                        mx = fj.xc;
                        my = fj.yc;
                    } else {
                        mx = (int16_t)((int32_t)((gi - 1) % (uint32_t)fj.sa_w) + fj.x_origin);
                        my = (int16_t)((int32_t)((gi - 1) / (uint32_t)fj.sa_w) + fj.y_origin);
                    }
                    mv = ((uint32_t)(uint16_t)my << 16) | (uint16_t)mx;
                }
            }
            sh.best_sad[i] = sad;
            sh.best_mv[i]  = mv;
        }
        __syncthreads();
    }
    // zero the MeContext mirrors of references that end with do_ref == 0, and of a b64 the temporal filter left early
    // (their content is stale in the reference), so that dumps are deterministic; everything
    // else is written as built. Called by all threads once sh.do_ref is final and nobody reads the mirrors any more.
    auto write_mirrors = [&]() {
        if (!gbs) return;
        for (int i = tid; i < 8 * 85; i += POST_THREADS) {
            const int lr = i / 85, l = lr >> 2, r = lr & 3;
            const bool keep = l < n_lists_cfg && r < nref[l] && sh.do_ref[l][r] && !st.tf_exit;
            gbs[i] = keep ? sh.best_sad[i] : 0u;
            gbm[i] = keep ? sh.best_mv[i] : 0u;
        }
    };

    const bool mctf = c.me_type == SVTME_ME_MCTF; // no pruning, candidate lists, statistics or GM flags (:3173, :3196)
    // ---- me_prune_ref ----
    if (tid < 8) {
        const int l = tid >> 2, r = tid & 3;
        uint64_t  h = st.sr_hme_sad[l][r];
        sh.do_ref[l][r] = st.do_ref[l][r];
        if (!mctf && c.enable_hme_flag && c.enable_me_hme_ref_pruning && l < n_lists_cfg && r < nref[l]) {
            h = 0;
            if (st.do_ref[l][r] == 0)
                h = (uint64_t)(MAX_SAD_VALUE_U32 * 64u);
            else
                for (int p = 0; p < 64; p++) h += BS(l, r, 21 + p); // tab8x8 is a permutation: plain sum
        }
        sh.hme_sad[l][r] = h;
    }
    __syncthreads();
    if (tid == 0 && !mctf && c.enable_hme_flag && c.enable_me_hme_ref_pruning) {
        const uint16_t th = c.prune_ref_if_me_sad_dev_bigger_than_th;
        if (th != (uint16_t)~0) {
            uint64_t best = ~0ull;
            for (int l = 0; l < 2; l++)
                for (int r = 0; r < 4; r++) best = sh.hme_sad[l][r] < best ? sh.hme_sad[l][r] : best;
            for (int l = 0; l < 2; l++)
                for (int r = 1; r < 4; r++)
                    if ((sh.hme_sad[l][r] - best) * 100 > ((uint64_t)th * best)) sh.do_ref[l][r] = 0;
        }
    }
    __syncthreads();

    // ---- candidate construction: thread n = ME z-order PU index ----
    const uint32_t n_pus  = (uint32_t)job.n_pus;
    const bool want_sb = job.total_me_candidate_index != nullptr, want_stats = job.me_64x64_distortion != nullptr;
    // row geometry of the two MeSbResults arrays: the reference's (max_refs, max_l0, max_cand) or this picture's own (SVTME_F_SB_COMPACT)
    const uint32_t mv_stride = (uint32_t)job.mv_stride, mv_l1 = (uint32_t)job.mv_l1_off, cand_stride = (uint32_t)job.cand_stride;
    uint8_t  *o_total = sh.total; // built in shared memory, stored at the end
    uint32_t *o_mv    = sh.mv;
    uint8_t  *o_cand  = sh.cand;
    auto write_sb_results = [&]() {
        if (!want_sb) return;
        post_store_bytes(job.total_me_candidate_index + (size_t)b * n_pus, sh.total, n_pus, tid);
        uint32_t *g_mv = job.me_mv_array + (size_t)b * n_pus * mv_stride;
        for (uint32_t i = tid; i < n_pus * mv_stride; i += POST_THREADS) g_mv[i] = sh.mv[i];
        post_store_bytes(job.me_candidate_array + (size_t)b * n_pus * cand_stride, sh.cand, n_pus * cand_stride, tid);
    };
    auto write_search_results = [&]() {
        if (tid >= 8) return;
        const int l = tid >> 2, r = tid & 3;
        SvtMeSearchResult sr;
        sr.hme_sc_x = 0; sr.hme_sc_y = 0; sr.do_ref = 0; sr.pad_[0] = sr.pad_[1] = sr.pad_[2] = 0; sr.hme_sad = 0;
        if (l < n_lists_cfg && r < nref[l]) {
            sr.hme_sc_x = st.hme_sc_x[l][r];
            sr.hme_sc_y = st.hme_sc_y[l][r];
            sr.do_ref   = sh.do_ref[l][r];
            sr.hme_sad  = sh.hme_sad[l][r];
        } else if (job.windows) {
            job.windows[(size_t)b * 8 + tid] = SvtMeWindow{0, 0, 0, 0};
        }
        if (job.search_results) job.search_results[(size_t)b * 8 + tid] = sr;
    };
    if (mctf) { // the temporal filter consumes p_sb_best_sad / p_sb_best_mv and SearchResults only
        for (uint32_t i = tid; i < n_pus; i += POST_THREADS) o_total[i] = 0;
        for (uint32_t i = tid; i < n_pus * mv_stride; i += POST_THREADS) o_mv[i] = 0;
        for (uint32_t i = tid; i < n_pus * cand_stride; i += POST_THREADS) o_cand[i] = 0;
        if (tid == 0 && want_stats) {
            job.me_64x64_distortion[b] = job.me_32x32_distortion[b] = job.me_16x16_distortion[b] = job.me_8x8_distortion[b] = 0;
            job.rc_me_distortion[b] = job.me_8x8_cost_variance[b] = 0;
            job.stationary_block_present_sb[b] = 0;
            job.rc_me_allow_gm[b]              = 0;
        }
        __syncthreads();
        write_sb_results();
        write_search_results();
        write_mirrors();
        return;
    }
    if (!want_sb && !want_stats) { // only the MeContext members were asked for
        write_search_results();
        write_mirrors();
        return;
    }
    // the two special builders start from memset(total_me_candidate_index, 1, number_of_pus) (:2625, :2734)
    const uint8_t total_init = (nref[0] == 1 && nref[1] <= 1) ? 1 : 0;
    for (uint32_t i = tid; i < n_pus; i += POST_THREADS) o_total[i] = total_init;
    for (uint32_t i = tid; i < n_pus * mv_stride; i += POST_THREADS) o_mv[i] = 0;
    for (uint32_t i = tid; i < n_pus * cand_stride; i += POST_THREADS) o_cand[i] = 0;
    __syncthreads();
    if (tid < 85) {
        const uint32_t n   = (uint32_t)tid;
        const bool     use = pp.enable_me_8x8 || n < 21; // enable_me_16x16 == 1
        if (nref[0] == 1 && nref[1] == 0) {
            // construct_me_candidate_array_single_ref
            const uint8_t pu = c_z_to_raster[n];
            sh.me_distortion[pu] = BS(0, 0, n);
            if (sh.do_ref[0][0] && use) {
                o_cand[pu * cand_stride] = cand_byte(0, 0, 0, 0, 0);
                o_mv[pu * mv_stride]   = BM(0, 0, n);
            }
        } else if (nref[0] == 1 && nref[1] == 1) {
            // construct_me_candidate_array_mrp_off
            const uint8_t pu = c_z_to_raster[n];
            uint32_t n_lists = (uint32_t)n_lists_cfg;
            uint8_t  org[2];
            org[0] = sh.do_ref[0][0];
            org[1] = (n_lists == 1) ? 0 : sh.do_ref[1][0];
            if (n_lists < 2 || !sh.do_ref[1][0]) n_lists = 1;
            const uint32_t th = (org[0] && org[1]) ? (uint32_t)c.prune_me_candidates_th : 0;
            uint8_t  off = 0, doref[2] = {org[0], org[1]};
            const uint32_t best = (org[0] && org[1]) ? min(BS(0, 0, n), BS(1, 0, n)) : org[0] ? BS(0, 0, n) : BS(1, 0, n);
            sh.me_distortion[pu] = best;
            int min_list = -1;
            if (c.use_best_unipred_cand_only && doref[0] && doref[1]) min_list = BS(0, 0, n) < BS(1, 0, n) ? 0 : 1;
            for (uint32_t l = 0; l < n_lists && (use || off == 0); l++) {
                if (doref[l] == 0) continue;
                if (th > 0) {
                    const uint32_t dev = (BS(l, 0, n) - best) * 100;
                    if (dev > (best * th)) {
                        doref[l] = 0;
                        continue;
                    }
                }
                if (min_list != -1 && min_list != (int)l) {
                    if (use) o_mv[pu * mv_stride + (l ? mv_l1 : 0)] = BM(l, 0, n);
                    continue;
                }
                if (use) {
                    o_cand[pu * cand_stride + off]                  = cand_byte(l, 0, 0, l == 0 ? l : 24, l == 1 ? l : 24);
                    o_mv[pu * mv_stride + (l ? mv_l1 : 0)] = BM(l, 0, n);
                }
                off++;
            }
            if (doref[0] && doref[1] && use) {
                o_cand[pu * cand_stride + off] = cand_byte(2, 0, 0, 0, 1);
                o_total[pu]                    = (uint8_t)(off + 1);
            }
        } else {
            // construct_me_candidate_array
            const uint8_t  pu      = (n > 4) ? c_z_to_raster[n] : (uint8_t)n;
            const uint32_t n_lists = (uint32_t)n_lists_cfg;
            const uint32_t th      = (uint32_t)c.prune_me_candidates_th;
            uint8_t  off = 0, doref[2][4];
            uint32_t best = MAX_U32_;
            for (uint32_t l = 0; l < n_lists; l++)
                for (uint32_t r = 0; r < nref[l]; r++) {
                    doref[l][r] = sh.do_ref[l][r];
                    if (doref[l][r] == 0) continue;
                    best = BS(l, r, n) < best ? BS(l, r, n) : best;
                }
            sh.me_distortion[pu] = best;
            for (uint32_t l = 0; l < n_lists && (use || off == 0); l++)
                for (uint32_t r = 0; r < nref[l] && (use || off == 0); r++) {
                    if (doref[l][r] == 0) continue;
                    if (th > 0) {
                        const uint32_t dev = (BS(l, r, n) - best) * 100;
                        if (dev > (best * th)) {
                            doref[l][r] = 0;
                            continue;
                        }
                    }
                    if (use) {
                        o_cand[pu * cand_stride + off]                       = cand_byte(l, r, r, l == 0 ? l : 24, l == 1 ? l : 24);
                        o_mv[pu * mv_stride + (l ? mv_l1 : 0) + r] = BM(l, r, n);
                    }
                    off++;
                }
            if (n_lists == 2 && use) {
                for (uint32_t a = 0; a < nref[0]; a++)
                    for (uint32_t bb = 0; bb < nref[1]; bb++) {
                        if (pp.only_l_bwd && (a > 0 || bb > 0)) continue;
                        if (doref[0][a] && doref[1][bb]) o_cand[pu * cand_stride + off++] = cand_byte(2, a, bb, 0, 1);
                    }
                if (!pp.only_l_bwd) {
                    for (uint32_t a = 1; a < nref[0]; a++)
                        if (doref[0][0] && doref[0][a]) o_cand[pu * cand_stride + off++] = cand_byte(2, 0, a, 0, 0);
                    if (nref[1] == 3 && doref[1][0] && doref[1][2]) o_cand[pu * cand_stride + off++] = cand_byte(2, 0, 2, 1, 1);
                }
            }
            if (use) o_total[pu] = off;
        }
    }
    __syncthreads();

    // ---- compute_distortion + perform_gm_detection + SearchResults ----
    uint32_t d8 = 0, d16 = 0, d32 = 0;
    uint64_t sumsq = 0;
    if (tid < 32) { // warp 0: the sums over the 64 / 16 / 4 PUs of a size as lane-parallel reductions
        d8  = sh.me_distortion[21 + tid] + sh.me_distortion[21 + 32 + tid];
        d16 = tid < 16 ? sh.me_distortion[5 + tid] : 0u;
        d32 = tid < 4 ? sh.me_distortion[1 + tid] : 0u;
        d8  = __reduce_add_sync(0xffffffffu, d8);
        d16 = __reduce_add_sync(0xffffffffu, d16);
        d32 = __reduce_add_sync(0xffffffffu, d32);
        const long long mean = (long long)(d8 / 64);
        const long long e0 = (long long)sh.me_distortion[21 + tid] - mean, e1 = (long long)sh.me_distortion[21 + 32 + tid] - mean;
        sumsq = (uint64_t)(e0 * e0) + (uint64_t)(e1 * e1);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) sumsq += __shfl_xor_sync(0xffffffffu, sumsq, o);
    }
    if (tid == 0 && want_stats) {
        const int b_row = div_by(b, job.b64_cols, job.inv_cols);
        const int org_x = (b - b_row * job.b64_cols) * 64, org_y = b_row * 64;
        const uint32_t gw = (uint32_t)imin(64, job.pic_w - org_x), gh = (uint32_t)imin(64, job.pic_h - org_y);
        const uint32_t d64 = sh.me_distortion[0];
        const uint32_t pix = gw * gh, size = 64 * 64;
        job.me_8x8_cost_variance[b] = (uint32_t)(sumsq / 64);
        job.rc_me_distortion[b]     = (pp.input_resolution <= 2) ? d8 : d16;
        job.me_64x64_distortion[b]  = (d64 * size) / pix;
        job.me_32x32_distortion[b]  = (d32 * size) / pix;
        job.me_16x16_distortion[b]  = (d16 * size) / pix;
        job.me_8x8_distortion[b]    = (d8 * size) / pix;

        uint8_t stationary = 0, allow_gm = 0;
        if (pp.gm_enabled) {
            uint32_t cnt[2][4][2][2];
            for (int i = 0; i < 32; i++) ((uint32_t *)cnt)[i] = 0;
            uint32_t stationary_cnt = 0, tot = 0;
            const bool     low_res = pp.input_resolution <= 2;
            const unsigned n_blk   = low_res ? 64 : 16;
            for (unsigned i = 0; i < n_blk; i++) {
                uint8_t n = (uint8_t)((low_res ? 21 : 5) + i);
                if (low_res && !pp.enable_me_8x8) n = (uint8_t)(5 + ((i >> 3) >> 1) * 4 + ((i & 7) >> 1));
                const uint8_t  cb  = o_cand[n * cand_stride];
                const unsigned dir = cb & 3;
                const uint32_t l = (dir == 0 || dir == 2) ? ((cb >> 6) & 1) : ((cb >> 7) & 1);
                const uint32_t r = (dir == 0 || dir == 2) ? ((cb >> 2) & 3) : ((cb >> 4) & 3);
                int active_th;
                if (low_res) {
                    const uint64_t pa = pp.picture_number, pb = pp.ref_picture_number[l][r];
                    const uint16_t dist = (uint16_t)iabs((int16_t)((pa > pb ? pa : pb) - (pa > pb ? pb : pa)));
                    active_th = pp.gm_use_distance_based_active_th ? imax(dist >> 1, 4) : 4;
                } else {
                    const uint16_t dist = (uint16_t)iabs((int16_t)(pp.picture_number - pp.ref_picture_number[l][r]));
                    active_th = pp.gm_use_distance_based_active_th ? imax(dist * 16, 32) : 32;
                }
                const uint32_t mv = BM(l, r, n);
                const int mx = ((int16_t)(mv & 0xFFFF)) * 4, my = ((int16_t)(mv >> 16)) * 4;
                if (mx < -active_th) cnt[l][r][0][0]++;
                else if (mx > active_th) cnt[l][r][0][1]++;
                if (my < -active_th) cnt[l][r][1][0]++;
                else if (my > active_th) cnt[l][r][1][1]++;
                const int sth = low_res ? 0 : 4;
                if (iabs(mx) <= sth && iabs(my) <= sth) stationary_cnt++;
                tot++;
            }
            if (stationary_cnt > ((tot * 5) / 100)) stationary = 1;
            for (int i = 0; i < 32; i++)
                if (((uint32_t *)cnt)[i] > (tot / 2)) allow_gm = 1;
        }
        job.stationary_block_present_sb[b] = stationary;
        job.rc_me_allow_gm[b]              = allow_gm;
    }
    write_sb_results();
    write_search_results();
    write_mirrors();
#undef BS
#undef BM
}

