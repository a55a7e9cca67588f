// me_ctl.cuh — k_hme_ctl: the scalar control of the HME stage between two search kernels.
//
// One warp per b64; lane = job slot = (list*4 + ref)*4 + quadrant. The per-b64 state (SbState, the
// MeContext fields of me_context.h:435-452,329-355) is worked on in place, in global memory. One launch does
//   1. POST     : turn the keys of the searches that just ran into SADs / search centres
//                 (tail of prehme_core :1727-1734, hme_level_0/1/2 :911-917 / :1013-1019 / :1106-1110,
//                 worst-quadrant replacement :2075-2102)
//   2. FINALIZE : the cross-reference step that closes a stage (pre-HME mirroring and pruning
//                 :1839-1865; set_final_seach_centre_sb :2252 + hme_prune_ref_and_adjust_sr :2547)
//   3. BUILD    : early-exit tests and window clamping of the next stage's searches
//                 (prehme_b64 :1792, hme_level0_b64 :1976, hme_level1_b64 :2111, hme_level2_b64 :2197)
// on the subset of references given by bit masks, so that the few control settings with a
// dependency between references (pre-HME l1_early_exit, reduce_hme_l0_sr_th_*) run as two rounds.
// Arithmetic is the reference's, int16 where the reference uses int16.
#pragma once
#include "me_device.cuh"
#include "me_search.cuh"

#define CTL_WARPS 4
enum { ST_NONE = -1, ST_PREHME = 0, ST_L0 = 1, ST_L1 = 2, ST_L2 = 3 };

struct CtlOp {
    int      init;        // init_me_hme_data (:3080) + zz pruning (:2471-2504)
    int      post_stage;  // ST_*
    uint32_t post_mask;   // bit (list*4+ref)
    int      finalize;    // ST_PREHME: mirror + prune; ST_L2: final centre + prune + sr adjust (any last level)
    int      build_stage; // ST_*
    uint32_t build_mask;
};

// tile geometry of a 1/16-plane search: shared by the builder (device) and the detector / known-answer launchers (host).
// vb_shape: the block is 16 wide, 8 rows two apart (the production shape). The wide tile's box holds 18 search rows; a
// window is cut into tiles of 18, 16 or 12 rows, whichever leaves the evaluators (s16_eval_choice) the fewest warp
// instructions: 96 x 80 positions are 5 tiles of 16 rows (3 full passes of the two-row blocked evaluator each) instead
// of 4 x 18 + 8 (3 passes of the three-row one at 3/4 occupancy each), 64 x 48 are 4 x 12 instead of 18 + 18 + 12.
__host__ __device__ inline void s16_tile_geometry(int sa_w, int sa_h, int *tile_w, int *tile_h, int *n_tx, int *n_ty, int *wide,
                                                  bool vb_shape = false, bool vb2 = false) {
    *wide = sa_w > S16T_TILE_W;
    *tile_w = *wide ? S16W_TILE_W : S16T_TILE_W;
    *tile_h = *wide ? S16W_TILE_H : S16T_TILE_H;
    if (sa_w <= 0 || sa_h <= 0) { *n_tx = 0; *n_ty = 0; return; }
    *n_tx = (sa_w + *tile_w - 1) / *tile_w;
    if (*wide && vb_shape && sa_h > 12) {
        int best_cost = 1 << 30;
        for (int h = S16W_TILE_H; h >= 12; h = h == 18 ? 16 : h - 4) { // 18, 16, 12
            int cost = 0;
            for (int tx = 0; tx < *n_tx; tx++) {
                const int tw = (sa_w - tx * *tile_w) < *tile_w ? (sa_w - tx * *tile_w) : *tile_w;
                int c_full = 0, c_rem = 0;
                s16_eval_choice(tw, h, 16, 8, 2, false, true, vb2, &c_full);
                if (sa_h % h) s16_eval_choice(tw, sa_h % h, 16, 8, 2, false, true, vb2, &c_rem);
                cost += (sa_h / h) * (c_full + 150) + (sa_h % h ? c_rem + 150 : 0); // + the per-tile work (load, cursor, reduction)
            }
            if (cost < best_cost) { best_cost = cost; *tile_h = h; }
        }
    }
    *n_ty = (sa_h + *tile_h - 1) / *tile_h;
}

// window clamp + plane address shared by prehme_core and hme_level_0/1/2
__device__ inline void ctl_make_job(HmeJobG &j, const DevPlane &ref, int16_t org_x, int16_t org_y, int16_t origin_x,
                                    int16_t origin_y, int16_t sa_w, int16_t sa_h, int16_t pad_w, int16_t pad_h,
                                    bool round_w8, int rows, int bw, bool vb_rows = false, bool vb2 = false) {
    clamp_axis(org_x, origin_x, sa_w, pad_w, (int16_t)ref.width);
    if (round_w8) sa_w = (int16_t)((sa_w < 8) ? sa_w : sa_w & ~0x07);
    clamp_axis(org_y, origin_y, sa_h, pad_h, (int16_t)ref.height);
    const int16_t x_tl = (int16_t)(((int16_t)ref.org_x + org_x) + origin_x);
    const int16_t y_tl = (int16_t)(((int16_t)ref.org_y + org_y) + origin_y);
    j.sa_w = sa_w; j.sa_h = sa_h; j.origin_x = origin_x; j.origin_y = origin_y;
    int tw, th, ntx, nty, wide;
    s16_tile_geometry(sa_w, sa_h, &tw, &th, &ntx, &nty, &wide, bw == 16 && rows == 8 && vb_rows, vb2);
    j.tile_w = (uint16_t)tw; j.tile_h = (uint16_t)th; j.n_tx = (uint8_t)ntx; j.n_ty = (uint8_t)nty; j.wide = (uint8_t)wide;
    j.px = x_tl; j.py = y_tl;
    // the block the job compares: offset of its top-left sample in the SOURCE plane of the same level (all planes of a
    // session share one geometry), width, and rows after sub-sampling
    j.src_off = (int32_t)((ref.org_y + org_y) * ref.stride + ref.org_x + org_x);
    j.bw = (uint8_t)bw; j.rows = (uint8_t)rows;
    j.merged = 0; j.m_w = j.m_h = 0;
    j.key    = HME_KEY_NONE;
    j.active = 1;
}

// Append the searches a warp just built to the batch-wide unit queue of k_search16. One queue entry = one (b64, list,
// ref): the up to four jobs of slots 4 lr .. 4 lr + 3 (the two pre-HME regions, or the four level-0 quadrants / their
// merged window) are searched by ONE warp, tile after tile, so that what a tile search needs besides its tile (the unit
// record, the job records, the source block rows) is fetched once per (b64, list, ref) and not once per tile, and a
// job's best position is kept in registers until its last tile. Called by all 32 lanes; has_work: this lane's job was
// built in this launch and has at least one tile to search.
__device__ inline void ctl_enqueue_units(const PicJob &job, bool has_work, int b) {
    const int      lane  = threadIdx.x & 31;
    const uint32_t work  = __ballot_sync(0xffffffffu, has_work);
    // small batches (one picture, a temporal-filter pair): every job is a unit of its own (bit 31 of the record), so that
    // a few hundred b64s still spread over all warps of the GPU
    const bool     lead  = job.s16_split ? has_work : ((lane & 3) == 0 && ((work >> lane) & 0xfu) != 0); // else: first lane of a (list, ref) group with work
    const uint32_t leads = __ballot_sync(0xffffffffu, lead);
    if (!leads) return;
    uint32_t base = 0;
    const int first = __ffs(leads) - 1;
    if (lane == first) base = atomicAdd(&job.s16_counters[0], (uint32_t)__popc(leads)); // one atomic per warp (the counter is shared by the batch)
    base = __shfl_sync(0xffffffffu, base, first) + (uint32_t)__popc(leads & ((1u << lane) - 1u));
    if (lead && base < job.s16_units_cap)
        job.s16_units[base] = make_uint2(blockIdx.y, (uint32_t)(b * HME_JOBS_PER_B64 + lane) | (job.s16_split ? 0x80000000u : 0u));
}

// key -> (sad, centre); centre keeps its value when no position was evaluated
__device__ inline void ctl_job_result(const HmeJobG &j, bool sub, int scale, uint64_t &sad, int16_t &cx, int16_t &cy) {
    const unsigned long long key = j.key;
    uint64_t s = key >> 32;
    if ((uint32_t)key != 0xffffffffu) {
        const int idx = (int)(uint32_t)key;
        cx = (int16_t)(idx % j.sa_w);
        cy = (int16_t)(idx / j.sa_w);
    }
    if (sub) s = s * 2;
    sad = s;
    cx  = (int16_t)(cx + j.origin_x);
    cx  = (int16_t)(cx * scale);
    cy  = (int16_t)(cy + j.origin_y);
    cy  = (int16_t)(cy * scale);
}

__device__ inline void ctl_set_quadrants(SbState &st, int lvl, int l, int r, int16_t cx, int16_t cy, uint64_t sad) {
    for (int b = 0; b < 2; b++)
        for (int a = 0; a < 2; a++) {
            st.x_hme[lvl][l][r][a][b]   = cx;
            st.y_hme[lvl][l][r][a][b]   = cy;
            st.hme_sad[lvl][l][r][a][b] = sad;
        }
}

__global__ void __launch_bounds__(32 * CTL_WARPS) k_hme_ctl(const PicJob *__restrict__ jobs, const CtlOp op) {
    const PicJob &job = jobs[blockIdx.y];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int b = blockIdx.x * CTL_WARPS + warp;
    if (b >= job.n_b64) return;
    const SvtMePicParams &pp = job.pp;
    const SvtMeCtrls     &c  = pp.ctrls;
    // The per-b64 state is worked on where it lives (global memory, through L1/L2): a launch touches a few hundred of its
    // 1552 bytes (the levels and references of the stage it posts / builds). Round 1 staged the whole record into shared
    // memory and wrote it back in every launch: 5 x ~44 MB of DRAM traffic per mini-GOP, more than the searches read.
    // Lanes exchange values through it; every such hand-over is separated by __syncwarp(), which orders memory accesses
    // among the lanes of the warp.
    SbState &st = job.sb_state[b];
    const B64Geo g = b64_geo(job, b);
    const bool sub = c.hme_search_method == SVTME_SUB_SAD_SEARCH;
    const int  row_step = sub ? 2 : 1;
    const int  lr = lane >> 2, l = lr >> 2, r = lr & 3, q = lane & 3, qa = q & 1, qb = q >> 1;
    const bool searched = l < pp.num_of_list_to_search && r < pp.num_of_ref_pic_to_search[l];
    const bool hme_ref  = pp.temporal_layer_index > 0 || l == 0; // refs that go through HME (:1806, :2044, ...)
    HmeJobG &j = job.hme_jobs[(size_t)b * HME_JOBS_PER_B64 + lane];
    unsigned long long ops = 0;

    // ---- state in ----
    uint32_t *gw = (uint32_t *)&job.sb_state[b];
    if (op.init) {
        uint32_t zz = 0;
        if (lane < 8) zz = gw[offsetof(SbState, zz_sad) / 4 + lane]; // written by k_zz for the refs it covers
        __syncwarp();
        for (int i = lane; i < (int)(sizeof(SbState) / 4); i += 32) gw[i] = 0u;
        __syncwarp();
        if (lane < 8) {
            const int l8 = lane >> 2, r8 = lane & 3;
            const bool has_zz = (c.me_early_exit_th || c.me_safe_limit_zz_th) && l8 < pp.num_of_list_to_search &&
                r8 < pp.num_of_ref_pic_to_search[l8] && (pp.temporal_layer_index > 0 || l8 == 0);
            st.do_ref[l8][r8]               = 1;
            st.sr_hme_sad[l8][r8]           = MAX_U32_;
            st.reduce_me_sr_divisor[l8][r8] = 1;
            st.zz_sad[l8][r8]               = has_zz ? zz : MAX_U32_;
        }
        j.active = 0;
        __syncwarp();
        if (lane == 0 && (c.me_early_exit_th || c.me_safe_limit_zz_th)) { // init_zz_sad :2455-2504
            uint32_t best = MAX_U32_;
            for (int ll = 0; ll < pp.num_of_list_to_search; ll++)
                for (int rr = 0; rr < pp.num_of_ref_pic_to_search[ll]; rr++) {
                    if (pp.temporal_layer_index > 0 || ll == 0) best = min(best, st.zz_sad[ll][rr]);
                }
            if (pp.temporal_layer_index > 0 && best < c.zz_sad_th)
                for (int ll = 0; ll < pp.num_of_list_to_search; ll++)
                    for (int rr = 1; rr < pp.num_of_ref_pic_to_search[ll]; rr++)
                        if ((st.zz_sad[ll][rr] - best) * 100 > ((uint32_t)c.zz_sad_pct * best)) st.do_ref[ll][rr] = 0;
            if (c.me_safe_limit_zz_th) {
                const bool limit = pp.hierarchical_levels > 0 && pp.num_of_list_to_search == 2 &&
                    pp.temporal_layer_index >= pp.hierarchical_levels && pp.similar_brightness_refs &&
                    st.zz_sad[0][0] < c.me_safe_limit_zz_th && st.zz_sad[1][0] < c.me_safe_limit_zz_th;
                if (limit)
                    for (int ll = 0; ll < pp.num_of_list_to_search; ll++)
                        for (int rr = 1; rr < pp.num_of_ref_pic_to_search[ll]; rr++) st.do_ref[ll][rr] = 0;
            }
        }
    }
    __syncwarp();

    // ---- POST ----
    const bool in_post = searched && ((op.post_mask >> lr) & 1) && j.active;
    if (op.post_stage == ST_PREHME) {
        if (in_post && q < 2) {
            ctl_job_result(j, sub, 4, st.prehme_sad[l][r][q], st.prehme_mv_x[l][r][q], st.prehme_mv_y[l][r][q]);
            st.prehme_valid[l][r][q]   = 1;
            st.performed_phme[l][r][q] = 1;
            j.active = 0;
        }
    } else if (op.post_stage >= ST_L0) {
        const int lvl = op.post_stage - ST_L0;
        if (in_post) {
            ctl_job_result(j, sub, lvl == 0 ? 4 : (lvl == 1 ? 2 : 1), st.hme_sad[lvl][l][r][qa][qb], st.x_hme[lvl][l][r][qa][qb],
                           st.y_hme[lvl][l][r][qa][qb]);
            j.active = 0;
        }
        __syncwarp();
        if (lvl == 0 && in_post && q == 0 && c.prehme_enable) {
            // get_worst_quadrant :1942-1971, then replace it by the better pre-HME result :2084-2101
            uint8_t  wa = 0, wb = 0;
            uint64_t mx = 0;
            const uint64_t(*hs)[2] = st.hme_sad[0][l][r];
            if (hs[0][0] > mx) { mx = hs[0][0]; wa = 0; wb = 0; }
            if (hs[1][0] > mx) { mx = hs[1][0]; wa = 1; wb = 0; }
            if (hs[0][1] > mx) { mx = hs[0][1]; wa = 0; wb = 1; }
            if (hs[1][1] > mx) { wa = 1; wb = 1; }
            const int s = st.prehme_sad[l][r][0] <= st.prehme_sad[l][r][1] ? 0 : 1;
            if (st.prehme_sad[l][r][s] < st.hme_sad[0][l][r][wa][wb]) {
                st.hme_sad[0][l][r][wa][wb] = st.prehme_sad[l][r][s];
                st.x_hme[0][l][r][wa][wb]   = st.prehme_mv_x[l][r][s];
                st.y_hme[0][l][r][wa][wb]   = st.prehme_mv_y[l][r][s];
            }
        }
    }
    __syncwarp();

    // ---- FINALIZE ----
    if (op.finalize == ST_PREHME) {
        if (searched && !hme_ref && q < 2) { // list 1 of a base-layer picture mirrors list 0, :1841-1847
            st.prehme_mv_x[1][r][q] = (int16_t)-st.prehme_mv_x[0][r][q];
            st.prehme_mv_y[1][r][q] = (int16_t)-st.prehme_mv_y[0][r][q];
            st.prehme_sad[1][r][q]  = st.prehme_sad[0][r][q];
        }
        __syncwarp();
        if (lane == 0) { // :1837-1838, :1851-1865
            uint32_t best_sad = MAX_U32_;
            for (int ll = 0; ll < pp.num_of_list_to_search; ll++)
                for (int rr = 0; rr < pp.num_of_ref_pic_to_search[ll]; rr++)
                    if (pp.temporal_layer_index > 0 || ll == 0)
                        best_sad = min(best_sad, (uint32_t)min(st.prehme_sad[ll][rr][0], st.prehme_sad[ll][rr][1]));
            if (pp.temporal_layer_index > 0 && best_sad < c.phme_sad_th)
                for (int ll = 0; ll < pp.num_of_list_to_search; ll++)
                    for (int rr = 1; rr < pp.num_of_ref_pic_to_search[ll]; rr++) {
                        if (!st.do_ref[ll][rr]) continue;
                        const uint32_t sad = (uint32_t)min(st.prehme_sad[ll][rr][0], st.prehme_sad[ll][rr][1]);
                        if ((sad - best_sad) * 100 > ((uint32_t)c.phme_sad_pct * best_sad)) st.do_ref[ll][rr] = 0;
                    }
        }
    } else if (op.finalize == ST_L2 && lane == 0) {
        // set_final_seach_centre_sb :2252-2450: hmeMvSad is carried from one reference to the next
        int16_t  hx = 0, hy = 0, scx = 0, scy = 0;
        uint64_t hsad = 0;
        for (int ll = 0; ll < pp.num_of_list_to_search; ll++)
            for (int rr = 0; rr < pp.num_of_ref_pic_to_search[ll]; rr++) {
                if (pp.temporal_layer_index > 0 || ll == 0) {
                    if (c.enable_hme_flag) {
                        int lvl = -1;
                        if (c.enable_hme_level0_flag && !c.enable_hme_level1_flag && !c.enable_hme_level2_flag) lvl = 0;
                        if (c.enable_hme_level1_flag && !c.enable_hme_level2_flag) lvl = 1;
                        if (c.enable_hme_level2_flag) lvl = 2;
                        if (lvl >= 0) {
                            hx   = st.x_hme[lvl][ll][rr][0][0];
                            hy   = st.y_hme[lvl][ll][rr][0][0];
                            hsad = st.hme_sad[lvl][ll][rr][0][0];
                            for (int bb = 0; bb < 2; bb++)
                                for (int aa = (bb == 0 ? 1 : 0); aa < 2; aa++)
                                    if (st.hme_sad[lvl][ll][rr][aa][bb] < hsad) {
                                        hx   = st.x_hme[lvl][ll][rr][aa][bb];
                                        hy   = st.y_hme[lvl][ll][rr][aa][bb];
                                        hsad = st.hme_sad[lvl][ll][rr][aa][bb];
                                    }
                        }
                        scx = hx;
                        scy = hy;
                    }
                } else {
                    scx = 0;
                    scy = 0;
                }
                st.hme_sc_x[ll][rr]   = scx;
                st.hme_sc_y[ll][rr]   = scy;
                st.sr_hme_sad[ll][rr] = hsad;
            }
        const bool mctf = c.me_type == SVTME_ME_MCTF;
        st.tf_exit = mctf && st.sr_hme_sad[0][0] < c.tf_me_exit_th; // :3179-3183
        if (c.enable_hme_flag && !mctf) { // hme_prune_ref_and_adjust_sr :2547-2588 (loops over all [2][4]); never for MCTF (:3173)
            const uint16_t th = c.prune_ref_if_hme_sad_dev_bigger_than_th;
            if (c.enable_me_hme_ref_pruning && th != (uint16_t)~0) {
                uint64_t best = ~0ull;
                for (int i = 0; i < 8; i++) best = st.sr_hme_sad[i >> 2][i & 3] < best ? st.sr_hme_sad[i >> 2][i & 3] : best;
                for (int ll = 0; ll < 2; ll++)
                    for (int rr = 1; rr < 4; rr++)
                        if ((st.sr_hme_sad[ll][rr] - best) * 100 > ((uint64_t)th * best)) st.do_ref[ll][rr] = 0;
            }
            if (c.enable_me_sr_adjustment)
                for (int ll = 0; ll < 2; ll++)
                    for (int rr = 0; rr < 4; rr++) {
                        if (iabs(st.hme_sc_x[ll][rr]) <= c.reduce_me_sr_based_on_mv_length_th &&
                            iabs(st.hme_sc_y[ll][rr]) <= c.reduce_me_sr_based_on_mv_length_th &&
                            st.sr_hme_sad[ll][rr] < c.stationary_hme_sad_abs_th)
                            st.reduce_me_sr_divisor[ll][rr] = c.stationary_me_sr_divisor;
                        else if (st.sr_hme_sad[ll][rr] < c.reduce_me_sr_based_on_hme_sad_abs_th)
                            st.reduce_me_sr_divisor[ll][rr] = c.me_sr_divisor_for_low_hme_sad;
                    }
        }
    }
    __syncwarp();

    // ---- BUILD ----
    bool l0_built = false, s16_work = false; // s16_work: this lane built a 1/16-plane search that k_search16 must run
    const bool in_build = searched && ((op.build_mask >> lr) & 1);
    const int16_t ox16 = (int16_t)g.org_x, oy16 = (int16_t)g.org_y;
    if (op.build_stage == ST_PREHME && in_build && q < 2 && hme_ref) {
        const int s = q;
        const int bw0 = g.b64_w >> 2, rows0 = sub ? (g.b64_h >> 2) >> 1 : (g.b64_h >> 2);
        // check_prehme_early_exit :1763-1789
        if (c.me_early_exit_th && st.zz_sad[l][r] < c.me_early_exit_th) {
            st.prehme_mv_x[l][r][s] = st.prehme_mv_y[l][r][s] = 0;
            st.prehme_sad[l][r][s]   = 0;
            st.prehme_valid[l][r][s] = 1;
        } else if (c.prehme_l1_early_exit && l == 1 && st.prehme_valid[0][r][s] &&
                   ((st.prehme_sad[0][r][s] < (32 * 32)) ||
                    ((iabs(st.prehme_mv_x[0][r][s]) < 16) && (iabs(st.prehme_mv_y[0][r][s]) < 16)))) {
            st.prehme_mv_x[l][r][s]  = (int16_t)-st.prehme_mv_x[0][r][s];
            st.prehme_mv_y[l][r][s]  = (int16_t)-st.prehme_mv_y[0][r][s];
            st.prehme_sad[l][r][s]   = st.prehme_sad[0][r][s];
            st.prehme_valid[l][r][s] = 1;
        } else if (!st.do_ref[l][r]) {
            st.prehme_mv_x[l][r][s] = st.prehme_mv_y[l][r][s] = 0;
            st.prehme_sad[l][r][s] = MAX_U32_;
        } else {
            const uint32_t factor = scaled_picture_distance(ref_distance(pp, l, r));
            const int16_t w = (int16_t)(uint16_t)min((uint32_t)c.prehme_sa_cfg[s].sa_min.width * factor, (uint32_t)c.prehme_sa_cfg[s].sa_max.width);
            const int16_t h = (int16_t)(uint16_t)min((uint32_t)c.prehme_sa_cfg[s].sa_min.height * factor, (uint32_t)c.prehme_sa_cfg[s].sa_max.height);
            const DevPlane &rf = job.ref[l][r].lvl[2];
            ctl_make_job(j, rf, (int16_t)(ox16 >> 2), (int16_t)(oy16 >> 2), (int16_t)(-(int16_t)(w >> 1)), (int16_t)(-(int16_t)(h >> 1)),
                         w, h, (int16_t)(rf.org_x - 1), (int16_t)(rf.org_y - 1), false, rows0, bw0, sub, false);
            const bool sk = c.prehme_skip_search_line && bw0 == 16 && rows0 <= 16;
            if (j.sa_w > 0 && j.sa_h > 0) ops += (unsigned long long)j.sa_w * (sk ? j.sa_h / 2 : j.sa_h) * bw0 * rows0;
            s16_work = j.n_tx && j.n_ty;
        }
    } else if (op.build_stage == ST_L0 && in_build) {
        const int bw0 = g.b64_w >> 2, rows0 = sub ? (g.b64_h >> 2) >> 1 : (g.b64_h >> 2);
        bool done = false; // the three early exits of hme_level0_b64 :1993-2038, taken before the layer test
        int16_t ex = 0, ey = 0;
        uint64_t esad = 0;
        if (c.me_early_exit_th && st.zz_sad[l][r] < (c.me_early_exit_th >> 2)) done = true;
        if (!done && c.prev_me_stage_based_exit_th) {
            const int s = st.prehme_sad[l][r][0] <= st.prehme_sad[l][r][1] ? 0 : 1;
            if (st.performed_phme[l][r][s] && st.prehme_sad[l][r][s] < (c.prev_me_stage_based_exit_th >> 4)) {
                done = true; ex = st.prehme_mv_x[l][r][s]; ey = st.prehme_mv_y[l][r][s]; esad = st.prehme_sad[l][r][s];
            }
        }
        if (!done && !st.do_ref[l][r]) { done = true; esad = MAX_U32_; }
        if (done) {
            st.x_hme[0][l][r][qa][qb] = ex; st.y_hme[0][l][r][qa][qb] = ey; st.hme_sad[0][l][r][qa][qb] = esad;
        } else if (hme_ref) {
            // get_hme_l0_search_area :1870-1937
            SvtMeSearchAreaMinMax sa = c.hme_l0_sa;
            if (c.enable_me_sr_adjustment && c.distance_based_hme_resizing) {
                uint8_t is_hor = 1, is_ver = 1, is_still = 0;
                if (c.reduce_hme_l0_sr_th_min && c.reduce_hme_l0_sr_th_max && (l || r)) {
                    const int mvx = st.x_hme[0][0][0][0][0], mvy = st.y_hme[0][0][0][0][0];
                    is_ver   = (iabs(mvx) < c.reduce_hme_l0_sr_th_min) && (iabs(mvy) > c.reduce_hme_l0_sr_th_max);
                    is_hor   = (iabs(mvx) > c.reduce_hme_l0_sr_th_max) && (iabs(mvy) < c.reduce_hme_l0_sr_th_min);
                    is_still = (iabs(mvx) < (c.reduce_hme_l0_sr_th_min * 3)) && (iabs(mvy) < (c.reduce_hme_l0_sr_th_min * 3));
                }
                uint8_t x_off = is_hor ? 1 : 2, y_off = is_ver ? 1 : 2;
                if (c.enable_me_sr_adjustment == 2 && is_still) x_off = y_off = 4;
                sa.sa_min.width  = (uint16_t)(sa.sa_min.width / (x_off + r));
                sa.sa_min.height = (uint16_t)(sa.sa_min.height / (y_off + r));
                sa.sa_max.width  = (uint16_t)(sa.sa_max.width / (x_off + r));
                sa.sa_max.height = (uint16_t)(sa.sa_max.height / (y_off + r));
            }
            const int32_t factor = scaled_picture_distance(ref_distance(pp, l, r));
            int16_t sw0 = (int16_t)(sa.sa_min.width / c.num_hme_sa_w);
            sw0 = (int16_t)imin((((sw0 * factor) + 15) & ~0x0F), (((sa.sa_max.width / c.num_hme_sa_w) + 15) & ~0x0F));
            int16_t sh0 = (int16_t)(sa.sa_min.height / c.num_hme_sa_h);
            sh0 = (int16_t)imin((sh0 * factor), sa.sa_max.height / c.num_hme_sa_h);
            // hme_level_0 :837-846
            const int16_t w  = (int16_t)((sw0 + 7) & ~0x07);
            const int16_t ox = (int16_t)(-(int16_t)((w * c.num_hme_sa_w) >> 1) + (int16_t)(w * qa));
            const int16_t oy = (int16_t)(-(int16_t)((sh0 * c.num_hme_sa_h) >> 1) + (int16_t)(sh0 * qb));
            const DevPlane &rf = job.ref[l][r].lvl[2];
            ctl_make_job(j, rf, (int16_t)(ox16 >> 2), (int16_t)(oy16 >> 2), ox, oy, w, sh0, (int16_t)(rf.org_x - 1),
                         (int16_t)(rf.org_y - 1), true, rows0, bw0, sub, true);
            if (j.sa_w > 0 && j.sa_h > 0) ops += (unsigned long long)j.sa_w * j.sa_h * bw0 * rows0;
            l0_built = true; // enqueued below, once the four quadrants of a reference know whether they merge
        }
    } else if ((op.build_stage == ST_L1 || op.build_stage == ST_L2) && in_build && hme_ref) {
        const int lvl = op.build_stage - ST_L0;
        bool done = false;
        if (lvl == 1) { // hme_level1_b64 :2128-2153
            if (c.me_early_exit_th && st.zz_sad[l][r] < (c.me_early_exit_th >> 2)) {
                st.x_hme[1][l][r][qa][qb] = 0; st.y_hme[1][l][r][qa][qb] = 0; st.hme_sad[1][l][r][qa][qb] = 0;
                done = true;
            } else if (!st.do_ref[l][r]) {
                st.x_hme[1][l][r][qa][qb] = 0; st.y_hme[1][l][r][qa][qb] = 0; st.hme_sad[1][l][r][qa][qb] = MAX_U32_;
                done = true;
            }
        }
        const uint32_t exit_th = lvl == 1 ? (c.prev_me_stage_based_exit_th >> 5) : (c.prev_me_stage_based_exit_th >> 2);
        if (!done && c.prev_me_stage_based_exit_th && st.hme_sad[lvl - 1][l][r][qa][qb] < exit_th) { // :2156-2166, :2214-2224
            st.x_hme[lvl][l][r][qa][qb]   = st.x_hme[lvl - 1][l][r][qa][qb];
            st.y_hme[lvl][l][r][qa][qb]   = st.y_hme[lvl - 1][l][r][qa][qb];
            st.hme_sad[lvl][l][r][qa][qb] = st.hme_sad[lvl - 1][l][r][qa][qb];
            done = true;
        }
        if (!done) {
            const SvtMeSearchArea area = lvl == 1 ? c.hme_l1_sa : c.hme_l2_sa;
            const int16_t w  = (int16_t)(((int16_t)area.width + 7) & ~0x07);
            const int16_t h  = (int16_t)area.height;
            const int16_t px = st.x_hme[lvl - 1][l][r][qa][qb], py = st.y_hme[lvl - 1][l][r][qa][qb];
            const int16_t ox = (int16_t)(-(w >> 1) + (lvl == 1 ? (int16_t)(px >> 1) : px));
            const int16_t oy = (int16_t)(-(h >> 1) + (lvl == 1 ? (int16_t)(py >> 1) : py));
            const int shift = lvl == 1 ? 1 : 0;
            const int bw = g.b64_w >> shift, rows = sub ? (g.b64_h >> shift) >> 1 : (g.b64_h >> shift);
            const DevPlane &rf = job.ref[l][r].lvl[lvl == 1 ? 1 : 0];
            if (lvl == 1)
                ctl_make_job(j, rf, (int16_t)(ox16 >> 1), (int16_t)(oy16 >> 1), ox, oy, w, h, (int16_t)(rf.org_x - 1),
                             (int16_t)(rf.org_y - 1), true, rows, bw);
            else
                ctl_make_job(j, rf, ox16, oy16, ox, oy, w, h, 63, 63, true, rows, bw);
            if (j.sa_w > 0 && j.sa_h > 0) ops += (unsigned long long)j.sa_w * j.sa_h * bw * rows;

        }
    }
    if (op.build_stage == ST_L0) { // warp-uniform: all lanes are here
        // The four quadrant windows of one reference (lanes 4k..4k+3: q = qa + 2 qb) are adjacent unless the picture edge
        // clamped them differently. When they form an exact 2 x 2 grid, ONE unit searches their union (k_search16 keeps a
        // best per quadrant): a quarter of the units, and lanes that a 16 x 8 quadrant alone leaves idle are filled.
        const int act = l0_built && j.sa_w > 0 && j.sa_h > 0;
        const int base = lane & ~3;
        int v_act[4], v_px[4], v_py[4], v_w[4], v_h[4];
#pragma unroll
        for (int k = 0; k < 4; k++) {
            v_act[k] = __shfl_sync(0xffffffffu, act, base + k);
            v_px[k]  = __shfl_sync(0xffffffffu, (int)j.px, base + k);
            v_py[k]  = __shfl_sync(0xffffffffu, (int)j.py, base + k);
            v_w[k]   = __shfl_sync(0xffffffffu, (int)j.sa_w, base + k);
            v_h[k]   = __shfl_sync(0xffffffffu, (int)j.sa_h, base + k);
        }
        const bool grid = v_act[0] && v_act[1] && v_act[2] && v_act[3] && v_w[0] == v_w[2] && v_w[1] == v_w[3] && v_h[0] == v_h[1] &&
            v_h[2] == v_h[3] && v_px[1] == v_px[0] + v_w[0] && v_px[2] == v_px[0] && v_px[3] == v_px[1] && v_py[1] == v_py[0] &&
            v_py[2] == v_py[0] + v_h[0] && v_py[3] == v_py[2] && (v_w[0] & 7) == 0;
        if (grid) {
            if (q == 0) {
                int tw, th, ntx, nty, wide;
                s16_tile_geometry(v_w[0] + v_w[1], v_h[0] + v_h[2], &tw, &th, &ntx, &nty, &wide, j.bw == 16 && j.rows == 8 && sub, true);
                j.merged = 1; j.m_w = (int16_t)(v_w[0] + v_w[1]); j.m_h = (int16_t)(v_h[0] + v_h[2]);
                j.tile_w = (uint16_t)tw; j.tile_h = (uint16_t)th; j.n_tx = (uint8_t)ntx; j.n_ty = (uint8_t)nty; j.wide = (uint8_t)wide;
                s16_work = ntx && nty;
            } else {
                j.merged = 2;
            }
        } else if (l0_built) {
            s16_work = j.n_tx && j.n_ty;
        }
    }
    if (op.build_stage == ST_PREHME || op.build_stage == ST_L0) ctl_enqueue_units(job, s16_work, b); // warp-uniform
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) ops += __shfl_xor_sync(0xffffffffu, ops, o);
    if (lane == 0 && ops && job.sad_ops) {
        atomicAdd(&job.sad_ops[0], ops);
        if (op.build_stage >= 0) atomicAdd(&job.sad_ops[2 + op.build_stage], ops); // [2] pre-HME, [3..5] HME level 0..2 (ops are counted when a stage is built)
    }
}
