// me_hme_kernel.cuh — body of k_hme (see me_hme.cuh for the work split).
#pragma once
#include "me_hme.cuh"

__global__ void __launch_bounds__(HME_THREADS) k_hme(const __grid_constant__ PicJob job) {
    extern __shared__ __align__(16) uint32_t hme_win[];
    __shared__ __align__(16) HmeSh sh;

    const SvtMePicParams &pp  = job.pp;
    const SvtMeCtrls     &c   = pp.ctrls;
    const int             tid = threadIdx.x;
    const int             b   = blockIdx.x;
    const int             org_x = (b % job.b64_cols) * 64, org_y = (b / job.b64_cols) * 64;
    // svt_aom_motion_estimation_b64, motion_estimation.c:3163-3170 (picture already 8-aligned)
    const int  b64_w = imin(64, job.pic_w - org_x), b64_h = imin(64, job.pic_h - org_y);
    const bool sub      = c.hme_search_method == SVTME_SUB_SAD_SEARCH;
    const int  row_step = sub ? 2 : 1;
    const int  win_words = job.hme_smem_bytes / 4;
    const bool list1_searched_tl0 = pp.temporal_layer_index > 0; // (tl > 0 || l == 0) tests below
    unsigned long long ops = 0; // thread 0: positions x block pixels of every search it set up

    // ---- init_me_hme_data, motion_estimation.c:3080-3140 ----
    for (int i = tid; i < (int)(sizeof(SbState) / 4); i += HME_THREADS) ((uint32_t *)&sh.st)[i] = 0;
    if (tid < 16) {
        ((uint8_t *)sh.performed_phme)[tid] = 0;
        ((uint8_t *)sh.prehme_valid)[tid]   = 0;
    }
    __syncthreads();
    if (tid < 8) {
        const int l = tid >> 2, r = tid & 3;
        sh.st.do_ref[l][r]               = 1;
        sh.st.sr_hme_sad[l][r]           = MAX_U32_;
        sh.st.reduce_me_sr_divisor[l][r] = 1;
        sh.st.zz_sad[l][r]               = MAX_U32_;
    }
    // source blocks (me_process.c:183-214)
    const DevPlane &sf = job.src.lvl[0], &sq = job.src.lvl[1], &ss = job.src.lvl[2];
    const uint8_t *src_full = sf.buf + (size_t)(sf.org_y + org_y) * sf.stride + sf.org_x + org_x;
    const uint8_t *src_q    = sq.buf + (size_t)(sq.org_y + (org_y >> 1)) * sq.stride + sq.org_x + (org_x >> 1);
    const uint8_t *src_s    = ss.buf + (size_t)(ss.org_y + (org_y >> 2)) * ss.stride + ss.org_x + (org_x >> 2);
    const int bw0 = b64_w >> 2, bh0 = b64_h >> 2, rows0 = sub ? bh0 >> 1 : bh0;
    const int bw1 = b64_w >> 1, bh1 = b64_h >> 1, rows1 = sub ? bh1 >> 1 : bh1;
    const int bw2 = b64_w, bh2 = b64_h, rows2 = sub ? bh2 >> 1 : bh2;
    hme_stage_src(sh.src16, 4, src_s, ss.stride, row_step, rows0, bw0, tid);
    hme_stage_src(sh.src4, 8, src_q, sq.stride, row_step, rows1, bw1, tid);
    hme_stage_src(sh.src1, 16, src_full, sf.stride, row_step, rows2, bw2, tid);
    __syncthreads();

    // ---- init_zz_sad, motion_estimation.c:2452-2507 ----
    if (c.me_early_exit_th || c.me_safe_limit_zz_th) {
        const int warp = tid >> 5;
        for (int i = warp; i < 8; i += HME_THREADS / 32) {
            const int l = i >> 2, r = i & 3;
            if (l < pp.num_of_list_to_search && r < pp.num_of_ref_pic_to_search[l] && (list1_searched_tl0 || l == 0)) {
                const DevPlane &rf  = job.ref[l][r].lvl[0];
                const uint8_t  *rp  = rf.buf + ((int16_t)rf.org_x + (int16_t)org_x + ((int16_t)rf.org_y + (int16_t)org_y) * rf.stride);
                uint32_t        zz  = warp_nxm_sad(src_full, sf.stride << 1, rp, rf.stride << 1, b64_h >> 1, b64_w) << 1;
                zz                  = (zz * 64 * 64) / (uint32_t)(b64_w * b64_h);
                if ((tid & 31) == 0) sh.st.zz_sad[l][r] = zz;
            }
        }
        __syncthreads();
        if (tid == 0) {
            uint32_t best = MAX_U32_;
            for (int l = 0; l < pp.num_of_list_to_search; l++)
                for (int r = 0; r < pp.num_of_ref_pic_to_search[l]; r++)
                    if (list1_searched_tl0 || l == 0) {
                        best = min(best, sh.st.zz_sad[l][r]);
                        ops += (unsigned long long)(b64_h >> 1) * b64_w;
                    }
            if (pp.temporal_layer_index > 0 && best < c.zz_sad_th)
                for (int l = 0; l < pp.num_of_list_to_search; l++)
                    for (int r = 1; r < pp.num_of_ref_pic_to_search[l]; r++)
                        if ((sh.st.zz_sad[l][r] - best) * 100 > ((uint32_t)c.zz_sad_pct * best)) sh.st.do_ref[l][r] = 0;
            if (c.me_safe_limit_zz_th) {
                const bool limit = pp.hierarchical_levels > 0 && pp.num_of_list_to_search == 2 &&
                    pp.temporal_layer_index >= pp.hierarchical_levels && pp.similar_brightness_refs &&
                    sh.st.zz_sad[0][0] < c.me_safe_limit_zz_th && sh.st.zz_sad[1][0] < c.me_safe_limit_zz_th;
                if (limit)
                    for (int l = 0; l < pp.num_of_list_to_search; l++)
                        for (int r = 1; r < pp.num_of_ref_pic_to_search[l]; r++) sh.st.do_ref[l][r] = 0;
            }
        }
        __syncthreads();
    }

    // ---- prehme_b64, motion_estimation.c:1792-1866 ----
    if (c.prehme_enable) {
        uint32_t best_sad = MAX_U32_; // thread 0 only
        for (int l = 0; l < pp.num_of_list_to_search; l++)
            for (int r = 0; r < pp.num_of_ref_pic_to_search[l]; r++) {
                const DevPlane &rf = job.ref[l][r].lvl[2];
                if (tid == 0) {
                    sh.n_jobs = 0;
                    if (list1_searched_tl0 || l == 0) {
                        const uint32_t factor = scaled_picture_distance(ref_distance(pp, l, r));
                        for (int s = 0; s < 2; s++) {
                            // check_prehme_early_exit, :1763-1789
                            if (c.me_early_exit_th && sh.st.zz_sad[l][r] < c.me_early_exit_th) {
                                sh.st.prehme_mv_x[l][r][s] = sh.st.prehme_mv_y[l][r][s] = 0;
                                sh.st.prehme_sad[l][r][s]                              = 0;
                                sh.prehme_valid[l][r][s]                               = 1;
                                continue;
                            }
                            if (c.prehme_l1_early_exit && l == 1 && sh.prehme_valid[0][r][s] &&
                                ((sh.st.prehme_sad[0][r][s] < (32 * 32)) ||
                                 ((iabs(sh.st.prehme_mv_x[0][r][s]) < 16) && (iabs(sh.st.prehme_mv_y[0][r][s]) < 16)))) {
                                sh.st.prehme_mv_x[l][r][s] = (int16_t)-sh.st.prehme_mv_x[0][r][s];
                                sh.st.prehme_mv_y[l][r][s] = (int16_t)-sh.st.prehme_mv_y[0][r][s];
                                sh.st.prehme_sad[l][r][s]  = sh.st.prehme_sad[0][r][s];
                                sh.prehme_valid[l][r][s]   = 1;
                                continue;
                            }
                            if (!sh.st.do_ref[l][r]) {
                                sh.st.prehme_mv_x[l][r][s] = sh.st.prehme_mv_y[l][r][s] = 0;
                                sh.st.prehme_sad[l][r][s]                              = MAX_U32_;
                                continue;
                            }
                            const uint16_t saw = (uint16_t)min((uint32_t)c.prehme_sa_cfg[s].sa_min.width * factor,
                                                               (uint32_t)c.prehme_sa_cfg[s].sa_max.width);
                            const uint16_t sah = (uint16_t)min((uint32_t)c.prehme_sa_cfg[s].sa_min.height * factor,
                                                               (uint32_t)c.prehme_sa_cfg[s].sa_max.height);
                            const int16_t w = (int16_t)saw, h = (int16_t)sah;
                            const int     ji = sh.n_jobs++;
                            hme_make_job(sh.jobs[ji], rf, (int16_t)(((int16_t)org_x) >> 2), (int16_t)(((int16_t)org_y) >> 2),
                                         (int16_t)(-(int16_t)(w >> 1)), (int16_t)(-(int16_t)(h >> 1)), w, h,
                                         (int16_t)(rf.org_x - 1), (int16_t)(rf.org_y - 1), false, (uint8_t)l, (uint8_t)r,
                                         (uint8_t)s, 0);
                            sh.job_key[ji] = HME_KEY_INIT;
                            {
                                const HmeJob &j = sh.jobs[ji];
                                const bool sk = c.prehme_skip_search_line && bw0 == 16 && rows0 <= 16;
                                if (j.sa_w > 0 && j.sa_h > 0) ops += (unsigned long long)j.sa_w * (sk ? j.sa_h / 2 : j.sa_h) * bw0 * rows0;
                            }
                        }
                    } else {
                        for (int s = 0; s < 2; s++) {
                            sh.st.prehme_mv_x[1][r][s] = (int16_t)-sh.st.prehme_mv_x[0][r][s];
                            sh.st.prehme_mv_y[1][r][s] = (int16_t)-sh.st.prehme_mv_y[0][r][s];
                            sh.st.prehme_sad[1][r][s]  = sh.st.prehme_sad[0][r][s];
                        }
                    }
                }
                __syncthreads();
                const int nj = sh.n_jobs;
                for (int ji = 0; ji < nj; ji++)
                    hme_search_staged(sh, sh.jobs[ji], ji, hme_win, win_words, sh.src16, bw0, rows0, row_step,
                                      c.prehme_skip_search_line && bw0 == 16 && rows0 <= 16);
                __syncthreads();
                if (tid == 0) {
                    for (int ji = 0; ji < nj; ji++) {
                        const HmeJob &j = sh.jobs[ji];
                        hme_job_result(j, sh.job_key[ji], sub, 4, sh.st.prehme_sad[l][r][j.a], sh.st.prehme_mv_x[l][r][j.a],
                                       sh.st.prehme_mv_y[l][r][j.a]);
                        sh.prehme_valid[l][r][j.a]   = 1;
                        sh.performed_phme[l][r][j.a] = 1;
                    }
                    if (list1_searched_tl0 || l == 0) {
                        const uint32_t m = (uint32_t)min(sh.st.prehme_sad[l][r][0], sh.st.prehme_sad[l][r][1]);
                        best_sad         = min(best_sad, m);
                    }
                }
            }
        if (tid == 0 && pp.temporal_layer_index > 0 && best_sad < c.phme_sad_th)
            for (int l = 0; l < pp.num_of_list_to_search; l++)
                for (int r = 1; r < pp.num_of_ref_pic_to_search[l]; r++) {
                    if (!sh.st.do_ref[l][r]) continue;
                    const uint32_t sad = (uint32_t)min(sh.st.prehme_sad[l][r][0], sh.st.prehme_sad[l][r][1]);
                    if ((sad - best_sad) * 100 > ((uint32_t)c.phme_sad_pct * best_sad)) sh.st.do_ref[l][r] = 0;
                }
        __syncthreads();
    }

    if (c.enable_hme_flag) {
        // ---- hme_level0_b64, motion_estimation.c:1976-2106 ----
        if (c.enable_hme_level0_flag) {
            for (int l = 0; l < pp.num_of_list_to_search; l++)
                for (int r = 0; r < pp.num_of_ref_pic_to_search[l]; r++) {
                    const DevPlane &rf = job.ref[l][r].lvl[2];
                    if (tid == 0) {
                        sh.n_jobs = 0;
                        bool done = false;
                        if (c.me_early_exit_th && sh.st.zz_sad[l][r] < (c.me_early_exit_th >> 2)) {
                            hme_set_quadrants(sh, c, 0, l, r, 0, 0, 0);
                            done = true;
                        }
                        if (!done && c.prev_me_stage_based_exit_th) {
                            const int s = sh.st.prehme_sad[l][r][0] <= sh.st.prehme_sad[l][r][1] ? 0 : 1;
                            if (sh.performed_phme[l][r][s] && sh.st.prehme_sad[l][r][s] < (c.prev_me_stage_based_exit_th >> 4)) {
                                hme_set_quadrants(sh, c, 0, l, r, sh.st.prehme_mv_x[l][r][s], sh.st.prehme_mv_y[l][r][s],
                                                  sh.st.prehme_sad[l][r][s]);
                                done = true;
                            }
                        }
                        if (!done && !sh.st.do_ref[l][r]) {
                            hme_set_quadrants(sh, c, 0, l, r, 0, 0, MAX_U32_);
                            done = true;
                        }
                        if (!done && (list1_searched_tl0 || l == 0)) {
                            // get_hme_l0_search_area, :1870-1937
                            SvtMeSearchAreaMinMax sa = c.hme_l0_sa;
                            if (c.enable_me_sr_adjustment && c.distance_based_hme_resizing) {
                                uint8_t is_hor = 1, is_ver = 1, is_still = 0;
                                if (c.reduce_hme_l0_sr_th_min && c.reduce_hme_l0_sr_th_max && (l || r)) {
                                    const int mvx = sh.st.x_hme[0][0][0][0][0], mvy = sh.st.y_hme[0][0][0][0][0];
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
                            int16_t       sw0    = (int16_t)(sa.sa_min.width / c.num_hme_sa_w);
                            sw0 = (int16_t)imin((((sw0 * factor) + 15) & ~0x0F), (((sa.sa_max.width / c.num_hme_sa_w) + 15) & ~0x0F));
                            int16_t sh0 = (int16_t)(sa.sa_min.height / c.num_hme_sa_h);
                            sh0         = (int16_t)imin((sh0 * factor), sa.sa_max.height / c.num_hme_sa_h);
                            for (int qb = 0; qb < c.num_hme_sa_h; qb++)
                                for (int qa = 0; qa < c.num_hme_sa_w; qa++) {
                                    // hme_level_0, :837-846
                                    const int16_t w  = (int16_t)((sw0 + 7) & ~0x07);
                                    const int16_t ox = (int16_t)(-(int16_t)((w * c.num_hme_sa_w) >> 1) + (int16_t)(w * qa));
                                    const int16_t oy = (int16_t)(-(int16_t)((sh0 * c.num_hme_sa_h) >> 1) + (int16_t)(sh0 * qb));
                                    const int     ji = sh.n_jobs++;
                                    hme_make_job(sh.jobs[ji], rf, (int16_t)(((int16_t)org_x) >> 2), (int16_t)(((int16_t)org_y) >> 2),
                                                 ox, oy, w, sh0, (int16_t)(rf.org_x - 1), (int16_t)(rf.org_y - 1), true,
                                                 (uint8_t)l, (uint8_t)r, (uint8_t)qa, (uint8_t)qb);
                                    sh.job_key[ji] = HME_KEY_INIT;
                                    if (sh.jobs[ji].sa_w > 0 && sh.jobs[ji].sa_h > 0)
                                        ops += (unsigned long long)sh.jobs[ji].sa_w * sh.jobs[ji].sa_h * bw0 * rows0;
                                }
                        }
                    }
                    __syncthreads();
                    const int nj = sh.n_jobs;
                    for (int ji = 0; ji < nj; ji++)
                        hme_search_staged(sh, sh.jobs[ji], ji, hme_win, win_words, sh.src16, bw0, rows0, row_step, false);
                    __syncthreads();
                    if (tid == 0 && nj > 0) {
                        for (int ji = 0; ji < nj; ji++) {
                            const HmeJob &j = sh.jobs[ji];
                            hme_job_result(j, sh.job_key[ji], sub, 4, sh.st.hme_sad[0][l][r][j.a][j.b],
                                           sh.st.x_hme[0][l][r][j.a][j.b], sh.st.y_hme[0][l][r][j.a][j.b]);
                        }
                        if (c.prehme_enable) {
                            // get_worst_quadrant, :1942-1971, then replace it by the better pre-HME result
                            uint8_t  qa = 0, qb = 0;
                            uint64_t mx = 0;
                            const uint64_t(*hs)[2] = sh.st.hme_sad[0][l][r];
                            if (hs[0][0] > mx) { mx = hs[0][0]; qa = 0; qb = 0; }
                            if (hs[1][0] > mx) { mx = hs[1][0]; qa = 1; qb = 0; }
                            if (hs[0][1] > mx) { mx = hs[0][1]; qa = 0; qb = 1; }
                            if (hs[1][1] > mx) { qa = 1; qb = 1; }
                            const int s = sh.st.prehme_sad[l][r][0] <= sh.st.prehme_sad[l][r][1] ? 0 : 1;
                            if (sh.st.prehme_sad[l][r][s] < sh.st.hme_sad[0][l][r][qa][qb]) {
                                sh.st.hme_sad[0][l][r][qa][qb] = sh.st.prehme_sad[l][r][s];
                                sh.st.x_hme[0][l][r][qa][qb]   = sh.st.prehme_mv_x[l][r][s];
                                sh.st.y_hme[0][l][r][qa][qb]   = sh.st.prehme_mv_y[l][r][s];
                            }
                        }
                    }
                }
            __syncthreads();
        }

        // ---- hme_level1_b64 (:2111-2192) and hme_level2_b64 (:2197-2247): one flat batch each ----
        for (int lvl = 1; lvl <= 2; lvl++) {
            if (lvl == 1 ? !c.enable_hme_level1_flag : !c.enable_hme_level2_flag) continue;
            if (tid == 0) {
                sh.n_jobs         = 0;
                sh.task_prefix[0] = 0;
                const int strips  = lvl == 1 ? 2 : 4;
                for (int l = 0; l < pp.num_of_list_to_search; l++)
                    for (int r = 0; r < pp.num_of_ref_pic_to_search[l]; r++) {
                        if (!(list1_searched_tl0 || l == 0)) continue;
                        if (lvl == 1) {
                            if (c.me_early_exit_th && sh.st.zz_sad[l][r] < (c.me_early_exit_th >> 2)) {
                                hme_set_quadrants(sh, c, 1, l, r, 0, 0, 0);
                                continue;
                            }
                            if (!sh.st.do_ref[l][r]) {
                                hme_set_quadrants(sh, c, 1, l, r, 0, 0, MAX_U32_);
                                continue;
                            }
                        }
                        const DevPlane &rf = job.ref[l][r].lvl[lvl == 1 ? 1 : 0];
                        for (int qb = 0; qb < c.num_hme_sa_h; qb++)
                            for (int qa = 0; qa < c.num_hme_sa_w; qa++) {
                                const uint32_t exit_th = lvl == 1 ? (c.prev_me_stage_based_exit_th >> 5) : (c.prev_me_stage_based_exit_th >> 2);
                                if (c.prev_me_stage_based_exit_th && sh.st.hme_sad[lvl - 1][l][r][qa][qb] < exit_th) {
                                    sh.st.x_hme[lvl][l][r][qa][qb]   = sh.st.x_hme[lvl - 1][l][r][qa][qb];
                                    sh.st.y_hme[lvl][l][r][qa][qb]   = sh.st.y_hme[lvl - 1][l][r][qa][qb];
                                    sh.st.hme_sad[lvl][l][r][qa][qb] = sh.st.hme_sad[lvl - 1][l][r][qa][qb];
                                    continue;
                                }
                                const SvtMeSearchArea area = lvl == 1 ? c.hme_l1_sa : c.hme_l2_sa;
                                const int16_t w  = (int16_t)(((int16_t)area.width + 7) & ~0x07);
                                const int16_t h  = (int16_t)area.height;
                                const int16_t px = sh.st.x_hme[lvl - 1][l][r][qa][qb], py = sh.st.y_hme[lvl - 1][l][r][qa][qb];
                                const int16_t ox = (int16_t)(-(w >> 1) + (lvl == 1 ? (int16_t)(px >> 1) : px));
                                const int16_t oy = (int16_t)(-(h >> 1) + (lvl == 1 ? (int16_t)(py >> 1) : py));
                                const int     ji = sh.n_jobs++;
                                if (lvl == 1)
                                    hme_make_job(sh.jobs[ji], rf, (int16_t)(((int16_t)org_x) >> 1), (int16_t)(((int16_t)org_y) >> 1),
                                                 ox, oy, w, h, (int16_t)(rf.org_x - 1), (int16_t)(rf.org_y - 1), true, (uint8_t)l,
                                                 (uint8_t)r, (uint8_t)qa, (uint8_t)qb);
                                else
                                    hme_make_job(sh.jobs[ji], rf, (int16_t)org_x, (int16_t)org_y, ox, oy, w, h, 63, 63, true,
                                                 (uint8_t)l, (uint8_t)r, (uint8_t)qa, (uint8_t)qb);
                                sh.job_key[ji] = HME_KEY_INIT;
                                const HmeJob &j = sh.jobs[ji];
                                const int tasks = (j.sa_w > 0 && j.sa_h > 0) ? j.sa_h * ((j.sa_w + 7) >> 3) * strips : 0;
                                sh.task_prefix[ji + 1] = sh.task_prefix[ji] + tasks;
                                if (j.sa_w > 0 && j.sa_h > 0)
                                    ops += (unsigned long long)j.sa_w * j.sa_h * (lvl == 1 ? bw1 * rows1 : bw2 * rows2);
                            }
                    }
            }
            __syncthreads();
            if (lvl == 1)
                hme_search_direct(sh, sh.src4, 2, bw1, rows1, row_step);
            else
                hme_search_direct(sh, sh.src1, 4, bw2, rows2, row_step);
            __syncthreads();
            if (tid == 0)
                for (int ji = 0; ji < sh.n_jobs; ji++) {
                    const HmeJob &j = sh.jobs[ji];
                    hme_job_result(j, sh.job_key[ji], sub, lvl == 1 ? 2 : 1, sh.st.hme_sad[lvl][j.l][j.r][j.a][j.b],
                                   sh.st.x_hme[lvl][j.l][j.r][j.a][j.b], sh.st.y_hme[lvl][j.l][j.r][j.a][j.b]);
                }
            __syncthreads();
        }
    }

    if (tid == 0) {
        // ---- set_final_seach_centre_sb, motion_estimation.c:2252-2450 (hmeMvSad carried across refs) ----
        int16_t  hx = 0, hy = 0, scx = 0, scy = 0;
        uint64_t hsad = 0;
        for (int l = 0; l < pp.num_of_list_to_search; l++)
            for (int r = 0; r < pp.num_of_ref_pic_to_search[l]; r++) {
                if (list1_searched_tl0 || l == 0) {
                    if (c.enable_hme_flag) {
                        int lvl = -1;
                        if (c.enable_hme_level0_flag && !c.enable_hme_level1_flag && !c.enable_hme_level2_flag) lvl = 0;
                        if (c.enable_hme_level1_flag && !c.enable_hme_level2_flag) lvl = 1;
                        if (c.enable_hme_level2_flag) lvl = 2;
                        if (lvl >= 0) {
                            hx   = sh.st.x_hme[lvl][l][r][0][0];
                            hy   = sh.st.y_hme[lvl][l][r][0][0];
                            hsad = sh.st.hme_sad[lvl][l][r][0][0];
                            for (int qb = 0; qb < c.num_hme_sa_h; qb++)
                                for (int qa = (qb == 0 ? 1 : 0); qa < c.num_hme_sa_w; qa++)
                                    if (sh.st.hme_sad[lvl][l][r][qa][qb] < hsad) {
                                        hx   = sh.st.x_hme[lvl][l][r][qa][qb];
                                        hy   = sh.st.y_hme[lvl][l][r][qa][qb];
                                        hsad = sh.st.hme_sad[lvl][l][r][qa][qb];
                                    }
                        }
                        scx = hx;
                        scy = hy;
                    }
                } else {
                    scx = 0;
                    scy = 0;
                }
                sh.st.hme_sc_x[l][r]   = scx;
                sh.st.hme_sc_y[l][r]   = scy;
                sh.st.sr_hme_sad[l][r] = hsad;
            }
        // ---- hme_prune_ref_and_adjust_sr, :2547-2588 (prune_ref = enable_hme_flag for open loop) ----
        if (c.enable_hme_flag) {
            const uint16_t th = c.prune_ref_if_hme_sad_dev_bigger_than_th;
            if (c.enable_me_hme_ref_pruning && th != (uint16_t)~0) {
                uint64_t best = ~0ull;
                for (int l = 0; l < 2; l++)
                    for (int r = 0; r < 4; r++) best = sh.st.sr_hme_sad[l][r] < best ? sh.st.sr_hme_sad[l][r] : best;
                for (int l = 0; l < 2; l++)
                    for (int r = 1; r < 4; r++)
                        if ((sh.st.sr_hme_sad[l][r] - best) * 100 > ((uint64_t)th * best)) sh.st.do_ref[l][r] = 0;
            }
            if (c.enable_me_sr_adjustment)
                for (int l = 0; l < 2; l++)
                    for (int r = 0; r < 4; r++) {
                        if (iabs(sh.st.hme_sc_x[l][r]) <= c.reduce_me_sr_based_on_mv_length_th &&
                            iabs(sh.st.hme_sc_y[l][r]) <= c.reduce_me_sr_based_on_mv_length_th &&
                            sh.st.sr_hme_sad[l][r] < c.stationary_hme_sad_abs_th)
                            sh.st.reduce_me_sr_divisor[l][r] = c.stationary_me_sr_divisor;
                        else if (sh.st.sr_hme_sad[l][r] < c.reduce_me_sr_based_on_hme_sad_abs_th)
                            sh.st.reduce_me_sr_divisor[l][r] = c.me_sr_divisor_for_low_hme_sad;
                    }
        }
    }
    if (tid == 0 && job.sad_ops) atomicAdd(&job.sad_ops[0], ops);
    __syncthreads();
    for (int i = tid; i < (int)(sizeof(SbState) / 4); i += HME_THREADS)
        ((uint32_t *)&job.sb_state[b])[i] = ((const uint32_t *)&sh.st)[i];
}
