/* me_sig_deriv.c — see me_sig_deriv.h. Citations: Source/Lib/Codec/enc_mode_config.c unless noted. */
#include "me_sig_deriv.h"

#include <string.h>

#define CLIP3(lo, hi, v) ((v) < (lo) ? (lo) : ((v) > (hi) ? (hi) : (v)))
#define MAXV(a, b) ((a) > (b) ? (a) : (b))
enum { MRS = -3, MRP = -2, MR = -1 };
enum { RES_240 = 0, RES_360, RES_480, RES_720, RES_1080, RES_4K, RES_8K }; /* definitions.h:2088-2094 */

static SvtMeSearchArea SA(int w, int h) {
    SvtMeSearchArea a;
    a.width  = (uint16_t)w;
    a.height = (uint16_t)h;
    return a;
}

/* svt_aom_derive_input_resolution, sequence_control_set.c:113-131; thresholds definitions.h:2060-2066 */
static int input_resolution(uint32_t px) {
    if (px < 0x28500) return RES_240;
    if (px < 0x4CE00) return RES_360;
    if (px < 0xA1400) return RES_480;
    if (px < 0x16DA00) return RES_720;
    if (px < 0x535200) return RES_1080;
    if (px < 0x140A000) return RES_4K;
    return RES_8K;
}

/* set_hme_search_params, :143-223 */
static void hme_search_params(const SvtMeSigDerivIn *in, int res, int rtc, SvtMeCtrls *c) {
    const int m = in->enc_mode;
    int q_mult = 0;
    c->num_hme_sa_w = c->num_hme_sa_h = 2;
    SvtMeSearchArea mn = SA(32, 32), mx = SA(192, 192);
    if (m <= MRS) {
        if (res < RES_4K) { mn = SA(128, 128); mx = SA(256, 256); } else { mn = SA(240, 240); mx = SA(480, 480); }
    } else if (m <= 1) {
        if (res >= RES_4K) { mn = SA(240, 240); mx = SA(480, 480); }
    } else if (m <= 3) {
    } else if (m <= 7) {
        q_mult = 3;
    } else if (!rtc && m <= 8) {
        if (!(in->sc_class1 || res >= RES_4K)) mn = SA(16, 16);
        q_mult = 3;
    } else if ((!rtc && m <= 10) || (rtc && m <= 8)) {
        if (!in->sc_class1) mn = SA(16, 16);
        q_mult = 3;
    } else {
        if (!in->sc_class1) {
            if (res < RES_4K) { mn = SA(8, 8); mx = SA(96, 96); } else { mn = SA(16, 16); mx = SA(96, 96); }
        }
        q_mult = 3;
    }
    if (q_mult) {
        const uint16_t qw = (uint16_t)CLIP3(500, 1000, (int)(q_mult * ((8 * in->qp) - 125)));
        mn.width  = (uint16_t)MAXV(8, (mn.width * qw) / 1000);
        mn.height = (uint16_t)MAXV(8, (mn.height * qw) / 1000);
        mx.width  = (uint16_t)MAXV(96, (mx.width * qw) / 1000);
        mx.height = (uint16_t)MAXV(96, (mx.height * qw) / 1000);
    }
    c->hme_l0_sa.sa_min = mn;
    c->hme_l0_sa.sa_max = mx;
    c->hme_l1_sa = c->hme_l2_sa = (m <= 0) ? SA(16, 16) : SA(8, 3);
}

/* set_me_search_params, :228-363 */
static void me_search_params(const SvtMeSigDerivIn *in, int res, int rtc, SvtMeCtrls *c) {
    const int m = in->enc_mode, sc = in->sc_class1;
    int q_mult = 0;
    SvtMeSearchArea mn, mx;
    if (rtc) {
        if (sc) {
            if (m <= 8) { mn = SA(32, 32); mx = SA(96, 96); }
            else if (m <= 10) { if (res < RES_1080) { mn = SA(16, 16); mx = SA(32, 16); } else { mn = SA(24, 24); mx = SA(24, 24); } }
            else { if (res < RES_1080) { mn = SA(16, 16); mx = SA(32, 16); } else { mn = SA(16, 6); mx = SA(16, 9); } }
        } else {
            if (m <= 9) { if (res < RES_1080) { mn = SA(16, 16); mx = SA(32, 16); } else { mn = SA(16, 6); mx = SA(16, 9); } }
            else if (res < RES_720) { mn = SA(8, 3); mx = SA(16, 9); }
            else if (res < RES_1080) { mn = SA(8, 1); mx = SA(16, 7); }
            else if (res < RES_4K) { mn = SA(8, 1); mx = SA(8, 7); }
            else { mn = SA(8, 1); mx = SA(8, 1); }
        }
    } else if (sc) {
        if (m <= 1) { mn = SA(175, 175); mx = SA(750, 750); }
        else if (m <= 7) { mn = SA(48, 48); mx = SA(224, 224); }
        else if (m <= 8) { mn = SA(32, 32); mx = SA(164, 164); }
        else if (m <= 9) { mn = SA(32, 32); mx = SA(96, 96); }
        else if (m <= 10) { mn = SA(16, 16); mx = SA(96, 96); }
        else { mn = SA(8, 8); mx = SA(32, 32); }
    } else if (m <= 1) { mn = SA(64, 64); mx = SA(256, 256); }
    else if (m <= 2) { mn = SA(32, 32); mx = SA(128, 128); }
    else if (m <= 3) { mn = SA(24, 24); mx = SA(104, 104); }
    else if (m <= 6) { mn = SA(16, 16); mx = SA(64, 32); }
    else if (m <= 8) {
        if (in->hierarchical_levels <= 3) {
            if (res < RES_4K) { mn = SA(8, 5); mx = SA(16, 9); } else { mn = SA(8, 1); mx = SA(8, 1); }
        } else if (res < RES_1080) { mn = SA(16, 16); mx = SA(32, 16); }
        else { mn = SA(16, 6); mx = SA(16, 9); }
        q_mult = 4;
    } else if (m <= 10) {
        if (res < RES_1080) { mn = SA(16, 16); mx = SA(32, 16); } else { mn = SA(16, 6); mx = SA(16, 9); }
        q_mult = 4;
    } else { mn = SA(8, 3); mx = SA(8, 3); q_mult = 2; }
    if (q_mult) {
        const uint16_t qw = (uint16_t)CLIP3(500, 1000, (int)(q_mult * ((8 * in->qp) - 125)));
        mn.width  = (uint16_t)MAXV(8, (mn.width * qw) / 1000);
        mn.height = (uint16_t)MAXV(3, (mn.height * qw) / 1000);
        mx.width  = (uint16_t)MAXV(8, (mx.width * qw) / 1000);
        mx.height = (uint16_t)MAXV(3, (mx.height * qw) / 1000);
    }
    if (in->frame_rate >> 16) { /* low_frame_rate_flag as written in the reference (:358) */
        mn.width  = (uint16_t)((mn.width * 3) >> 1);
        mn.height = (uint16_t)((mn.height * 3) >> 1);
    }
    c->me_sa.sa_min = mn;
    c->me_sa.sa_max = mx;
}

/* svt_aom_set_prehme_ctrls, :555-606 */
static void prehme_ctrls(SvtMeCtrls *c, int level) {
    static const int cfg[5][10] = {/* v.min w,h, v.max w,h, h.min w,h, h.max w,h, skip, l1_exit */
        {0, 0, 0, 0, 0, 0, 0, 0, 0, 0},       {8, 144, 8, 496, 144, 3, 496, 3, 0, 0}, {8, 100, 8, 400, 96, 3, 384, 3, 0, 0},
        {8, 100, 8, 350, 32, 7, 200, 7, 1, 0}, {8, 100, 8, 350, 32, 7, 128, 7, 1, 1}};
    c->prehme_enable = level ? 1 : 0;
    if (!level) return;
    const int *k = cfg[level];
    c->prehme_sa_cfg[0].sa_min = SA(k[0], k[1]);
    c->prehme_sa_cfg[0].sa_max = SA(k[2], k[3]);
    c->prehme_sa_cfg[1].sa_min = SA(k[4], k[5]);
    c->prehme_sa_cfg[1].sa_max = SA(k[6], k[7]);
    c->prehme_skip_search_line = (uint8_t)k[8];
    c->prehme_l1_early_exit    = (uint8_t)k[9];
}

/* svt_aom_set_me_hme_ref_prune_ctrls, :364-443 */
static void ref_prune_ctrls(SvtMeCtrls *c, int level) {
    static const uint16_t hme_th[7] = {0xFFFF, 80, 50, 30, 15, 5, 5};
    static const uint16_t me_th[7]  = {0xFFFF, 0xFFFF, 0xFFFF, 0xFFFF, 60, 60, 60};
    c->enable_me_hme_ref_pruning               = level ? 1 : 0;
    c->prune_ref_if_hme_sad_dev_bigger_than_th = hme_th[level];
    c->prune_ref_if_me_sad_dev_bigger_than_th  = me_th[level];
    c->zz_sad_th    = level == 6 ? 20 * 64 * 64 : 0;
    c->zz_sad_pct   = level == 6 ? 5 : 0;
    c->phme_sad_th  = level == 6 ? 10 * 64 * 64 : 0;
    c->phme_sad_pct = level == 6 ? 5 : 0;
}

/* svt_aom_set_me_sr_adjustment_ctrls, :466-533 */
static void sr_adjustment_ctrls(SvtMeCtrls *c, int level) {
    static const int k[6][7] = {/* enable, mv_len_th, stat_abs_th, stat_div, low_abs_th, low_div, dist_resize */
        {0, 0, 0, 0, 0, 0, 0},          {1, 4, 12000, 8, 6000, 8, 0},   {1, 4, 12000, 8, 6000, 8, 1},
        {1, 4, 12000, 8, 12000, 8, 1},  {2, 16, 20000, 8, 20000, 8, 1}, {2, 20, 24000, 8, 24000, 8, 1}};
    c->enable_me_sr_adjustment = (uint8_t)k[level][0];
    if (!level) return; /* the other fields keep their previous value in the reference: unused when disabled */
    c->reduce_me_sr_based_on_mv_length_th   = (uint16_t)k[level][1];
    c->stationary_hme_sad_abs_th            = (uint16_t)k[level][2];
    c->stationary_me_sr_divisor             = (uint16_t)k[level][3];
    c->reduce_me_sr_based_on_hme_sad_abs_th = (uint16_t)k[level][4];
    c->me_sr_divisor_for_low_hme_sad        = (uint16_t)k[level][5];
    c->distance_based_hme_resizing          = (uint8_t)k[level][6];
    if (c->enable_hme_level2_flag == 0) {
        const int d = c->enable_hme_level1_flag == 1 ? 4 : 16;
        c->stationary_hme_sad_abs_th            = (uint16_t)(c->stationary_hme_sad_abs_th / d);
        c->reduce_me_sr_based_on_hme_sad_abs_th = (uint16_t)(c->reduce_me_sr_based_on_hme_sad_abs_th / d);
    }
}

/* svt_aom_get_enable_me_8x8, :76-95 */
static uint8_t enable_me_8x8(int m, int rtc, int res) {
    if (rtc) return m <= 9;
    if (m <= 5) return 1;
    if (m <= 10) return res <= RES_720;
    return 0;
}

uint32_t svtme_sig_deriv_me(const SvtMeSigDerivIn *in, SvtMePicParams *out) {
    if (!in || !out || in->width <= 0 || in->height <= 0 || (in->width & 7) || (in->height & 7) || in->num_lists < 1 ||
        in->num_lists > 2 || in->ref_count[0] < 1 || in->ref_count[0] > 4 || in->ref_count[1] < 0 || in->ref_count[1] > 4 ||
        in->max_l0 < 1)
        return SVTME_ERR_BAD_PARAMETER;
    memset(out, 0, sizeof(*out));
    SvtMeCtrls *c   = &out->ctrls;
    const int   m   = in->enc_mode, sc = in->sc_class1 != 0;
    const int   rtc = in->pred_structure == 1; /* SVT_AV1_PRED_LOW_DELAY_B */
    const int   res = input_resolution((uint32_t)in->width * (uint32_t)in->height);
    const int   is_base = in->temporal_layer_index == 0;

    me_search_params(in, res, rtc, c);
    hme_search_params(in, res, rtc, c);
    /* svt_aom_sig_deriv_multi_processes, :1820-1831 */
    c->enable_hme_flag = c->enable_hme_level0_flag = c->enable_hme_level1_flag = 1;
    c->enable_hme_level2_flag = (sc || m <= 6) ? 1 : 0;
    c->hme_search_method = c->me_search_method = SVTME_SUB_SAD_SEARCH; /* :712-713 */
    /* :715-731 */
    if (rtc && sc) { c->reduce_hme_l0_sr_th_min = 8; c->reduce_hme_l0_sr_th_max = 100; }
    else if (rtc && m > 8) { c->reduce_hme_l0_sr_th_min = 8; c->reduce_hme_l0_sr_th_max = 200; }
    /* :733-754 */
    int prehme_level;
    if (m <= MRS) prehme_level = 1;
    else if (sc) prehme_level = rtc ? 1 : 2;
    else if (rtc) prehme_level = m <= 8 ? 2 : (m <= 10 ? 4 : 0);
    else prehme_level = m <= 8 ? 2 : 4;
    if (c->enable_hme_level1_flag == 0) prehme_level = 0;
    prehme_ctrls(c, prehme_level);
    /* :758-786 */
    int prune;
    if (sc) prune = m <= MRS ? 0 : (m <= 2 ? 1 : (m <= 8 ? 3 : 6));
    else if (m <= MRS) prune = 0;
    else if (m <= MR) prune = 1;
    else if (m <= 0) prune = is_base ? 1 : 2;
    else if (m <= 1) prune = is_base ? 1 : 4;
    else if (m <= 5) prune = is_base ? 1 : 5;
    else if (m <= 10) prune = is_base ? 1 : 6;
    else prune = 6;
    ref_prune_ctrls(c, prune);
    /* :789-801 */
    sr_adjustment_ctrls(c, sc ? (m <= 8 ? 4 : 5) : (m <= MR ? 0 : (m <= 0 ? 1 : 3)));
    /* :803-810, :445-464 */
    const int mv_adj = m <= MRS ? 1 : (m <= 3 ? 2 : 0);
    c->mv_sa_adj_enabled = mv_adj ? 1 : 0;
    if (mv_adj) {
        c->mv_sa_adj_nearest_ref_only = mv_adj == 2;
        c->mv_sa_adj_mv_size_th       = 25;
        c->mv_sa_adj_sa_multiplier    = 2;
    }
    /* :812-813, :534-553 (level 2) */
    c->me_8x8_var_enabled = 1;
    c->me_sr_div4_th      = 80000;
    c->me_sr_div2_th      = 150000;
    c->me_sr_mult2_th     = 0xFFFFFFFFu;
    c->prune_me_candidates_th     = m <= 5 ? 0 : 65; /* :814-817 */
    c->use_best_unipred_cand_only = m <= 4 ? 0 : 1;  /* :819, :1984-1987 */
    /* :820-832 */
    if (rtc) c->me_early_exit_th = sc ? 64 * 64 : (m <= 10 ? 64 * 64 * 8 : 64 * 64 * 9);
    else c->me_early_exit_th = m <= 4 ? 0 : 64 * 64 * 8;
    c->me_safe_limit_zz_th         = in->safe_limit_nref == 1 ? (uint32_t)in->safe_limit_zz_th : 0; /* :834 */
    c->prev_me_stage_based_exit_th = (rtc && sc) ? 64 * 64 * 4 : 0;                                 /* :837-842 */

    /* PCS-level fields */
    out->picture_number              = in->picture_number;
    out->num_of_list_to_search       = (uint8_t)in->num_lists;
    out->num_of_ref_pic_to_search[0] = (uint8_t)in->ref_count[0];
    out->num_of_ref_pic_to_search[1] = (uint8_t)(in->num_lists == 2 ? in->ref_count[1] : 0);
    out->temporal_layer_index        = (uint8_t)in->temporal_layer_index;
    out->is_ref                      = (uint8_t)in->is_ref;
    out->hierarchical_levels         = (uint8_t)in->hierarchical_levels;
    out->similar_brightness_refs     = in->similar_brightness_refs != 0;
    out->enable_me_16x16             = 1; /* svt_aom_get_enable_me_16x16, :97-101 */
    out->enable_me_8x8               = enable_me_8x8(m, rtc, res);
    /* svt_aom_get_max_allocated_me_refs, pcs.c:91-96 */
    out->max_refs = (uint8_t)(in->max_l0 + in->max_l1);
    out->max_cand = (uint8_t)(in->max_l0 + in->max_l1 + in->max_l0 * in->max_l1 + (in->max_l0 - 1) + (in->max_l1 == 3 ? 1 : 0));
    out->max_l0   = (uint8_t)in->max_l0;
    out->only_l_bwd       = (uint8_t)in->only_l_bwd;
    out->input_resolution = (uint8_t)res;
    /* svt_aom_get_gm_core_level :103-116 (super-res off, non-I picture), svt_aom_set_gm_controls :2001-2130 */
    int gm = in->gm_level >= 0 ? in->gm_level : (m <= MRP ? 1 : (m <= MR ? 2 : (m <= 2 ? 4 : 0)));
    out->gm_enabled = gm ? 1 : 0;
    out->gm_use_distance_based_active_th = (gm >= 7) ? 1 : 0; /* :2019-2160 */
    for (int l = 0; l < 2; l++)
        for (int r = 0; r < 4; r++) {
            out->ref_pic_id[l][r]         = -1;
            out->ref_picture_number[l][r] = in->ref_picture_number[l][r];
        }
    return SVTME_OK;
}

/* svt_aom_sig_deriv_me_tf, enc_mode_config.c:849-895 with tf_set_me_hme_params_oq :611-688 */
uint32_t svtme_sig_deriv_me_tf(const SvtMeSigDerivTfIn *in, SvtMePicParams *out) {
    if (!in || !out || in->width <= 0 || in->height <= 0 || (in->width & 7) || (in->height & 7) || in->tf_hme_me_level < 0 ||
        in->tf_hme_me_level > 4)
        return SVTME_ERR_BAD_PARAMETER;
    memset(out, 0, sizeof(*out));
    SvtMeCtrls *c   = &out->ctrls;
    const int   m   = in->enc_mode, lvl = in->tf_hme_me_level;
    const int   res = input_resolution((uint32_t)in->width * (uint32_t)in->height);
    c->num_hme_sa_w = c->num_hme_sa_h = 2;
    SvtMeSearchArea l0mn, l0mx, l1, l2, memn, memx;
    switch (lvl) {
    case 0: l0mn = SA(30, 30); l0mx = SA(60, 60); l1 = l2 = SA(16, 16); memn = SA(60, 60); memx = SA(120, 120); break;
    case 1: l0mn = SA(16, 16); l0mx = SA(32, 32); l1 = l2 = SA(16, 16); memn = SA(16, 16); memx = SA(32, 32); break;
    case 2:
        if (res <= RES_360) { l0mn = SA(8, 8); l0mx = SA(8, 8); l1 = SA(8, 8); }
        else if (res <= RES_480) { l0mn = SA(8, 8); l0mx = SA(16, 16); l1 = SA(8, 8); }
        else { l0mn = SA(16, 16); l0mx = SA(32, 32); l1 = SA(16, 16); }
        l2 = SA(16, 16); memn = memx = SA(8, 8);
        break;
    case 3: l0mn = l0mx = SA(8, 8); l1 = l2 = SA(8, 8); memn = memx = SA(8, 8); break;
    default: l0mn = l0mx = SA(4, 4); l1 = l2 = SA(8, 8); memn = memx = SA(8, 8); break;
    }
    if (in->tf_qp_opt) { /* :679-687 */
        const uint16_t qw = (uint16_t)CLIP3(250, 1000, (int)((8 * in->qp) - 125));
        memn.width  = (uint16_t)MAXV(8, (memn.width * qw) / 1000);
        memn.height = (uint16_t)MAXV(8, (memn.height * qw) / 1000);
        memx.width  = (uint16_t)MAXV(8, (memx.width * qw) / 1000);
        memx.height = (uint16_t)MAXV(8, (memx.height * qw) / 1000);
    }
    c->hme_l0_sa.sa_min = l0mn; c->hme_l0_sa.sa_max = l0mx; /* set_hme_search_params_mctf(ctx, 0): the defaults as they are */
    c->hme_l1_sa = l1; c->hme_l2_sa = l2;
    c->me_sa.sa_min = memn; c->me_sa.sa_max = memx;
    /* :1832-1856 */
    c->enable_hme_flag = c->enable_hme_level0_flag = 1;
    c->enable_hme_level1_flag = lvl <= 2;
    c->enable_hme_level2_flag = lvl == 0;
    c->hme_search_method = SVTME_FULL_SAD_SEARCH; /* temporal_filtering.c:313, whatever the level */
    c->me_search_method  = lvl <= 2 ? SVTME_FULL_SAD_SEARCH : SVTME_SUB_SAD_SEARCH; /* :859-869 */
    prehme_ctrls(c, 0);
    ref_prune_ctrls(c, 0);
    sr_adjustment_ctrls(c, 0);
    /* svt_aom_set_mv_based_sa_ctrls(0), svt_aom_set_me_8x8_var_ctrls(0): disabled = all zero */
    c->me_early_exit_th            = (m <= 6 || res <= RES_720) ? 0 : 64 * 64 * 4; /* :884-886 */
    c->prev_me_stage_based_exit_th = (m <= 7 || res <= RES_720) ? 0 : 64 * 64 * 4; /* :891-893 */
    c->me_type       = SVTME_ME_MCTF;
    c->tf_me_exit_th = (uint16_t)in->tf_me_exit_th; /* me_context.h:495 */
    out->picture_number              = in->centre_picture_number;
    out->num_of_list_to_search       = 1;
    out->num_of_ref_pic_to_search[0] = 1;
    out->temporal_layer_index        = (uint8_t)in->temporal_layer_index;
    out->is_ref                      = (uint8_t)in->is_ref;
    out->hierarchical_levels         = 5;
    out->enable_me_8x8 = out->enable_me_16x16 = 1;
    out->max_cand = 3; out->max_refs = 2; out->max_l0 = 1; /* unused by MCTF: sizes of the (empty) candidate arrays */
    out->input_resolution = (uint8_t)res;
    for (int l = 0; l < 2; l++)
        for (int r = 0; r < 4; r++) out->ref_pic_id[l][r] = -1;
    out->ref_picture_number[0][0] = in->ref_picture_number;
    return SVTME_OK;
}

int32_t svtme_shard_owner(int64_t picture_number, int32_t mini_gop_size, int32_t stream, int32_t world_size) {
    if (world_size <= 1 || mini_gop_size <= 0) return 0;
    /* picture 0 (key frame) belongs to the first mini-GOP; pictures 1..N form mini-GOP 0, and so on */
    const int64_t g = picture_number <= 0 ? 0 : (picture_number - 1) / mini_gop_size;
    return (int32_t)((g + stream) % world_size);
}
