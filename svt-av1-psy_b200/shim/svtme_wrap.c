/*
 * svtme_wrap.c — the reference-side binding of libsvtme_b200.so.
 *
 * Link the encoder (static, non-LTO libSvtAv1Enc.a) with
 *     -Wl,--wrap=svt_aom_motion_estimation_b64 -Wl,--wrap=dg_detector_hme_level0  svtme_wrap.o  -lsvtme_b200
 * and every caller of those two functions (me_process.c:266 open-loop ME, temporal_filtering.c:3141 temporal-filter ME,
 * me_process.c:329 dynamic-GOP detector) reaches this file instead. The first b64 of a picture that arrives launches the
 * WHOLE picture on the GPU; every call then copies its own b64's rows from the picture-level result arrays into the memory
 * the reference owns (MeContext, pcs->pa_me_data->me_results[b64], the PCS statistic arrays), so the code downstream of
 * the call sees exactly what svt_aom_motion_estimation_b64 (motion_estimation.c:3146-3223) would have left there.
 *
 * It compiles against the reference's own headers (pcs.h, me_context.h ...), which are not part of this repository:
 * oracle/Makefile builds it where /root/reference exists, together with the reference's ME sources and the test harness
 * that plays the role of me_process.c, into oracle/_ref/libsvtref_wrap.so (tests/test_wrap_shim.py).
 *
 * There is no CPU fallback: a failing GPU call is fatal. __real_svt_aom_motion_estimation_b64 is reached only for ME
 * types this library does not implement (anything but ME_OPEN_LOOP and ME_MCTF; the reference has no such caller).
 */
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "definitions.h"
#include "pcs.h"
#include "sequence_control_set.h"
#include "me_context.h"
#include "motion_estimation.h"
#include "reference_object.h"
#include "enc_mode_config.h"

#include "svtme_b200.h"
#include "svtme_wrap.h"

#define SHIM_SLOTS 64 /* device picture store: sources + references alive (look-ahead + mini-GOP) */
#define SHIM_JOBS 64  /* pictures / temporal-filter pairs whose results are being handed out */

typedef struct ShimSlot {
    int            valid;
    uint64_t       picture_number;
    const uint8_t *buf; /* identity of the host plane */
    const void    *scope; /* who may reuse the copy, see shim_slot_of_plane */
    uint64_t       stamp;
} ShimSlot;

typedef struct ShimJob {
    int             used;
    const void     *pcs;
    int             me_type;
    uint64_t        ref_number; /* ME_MCTF: the neighbouring picture of this (centre, neighbour) pair */
    uint32_t        n_b64, n_pus, delivered;
    uint64_t        stamp;
    int             readers; /* threads copying rows out right now: never evicted */
    int             pending; /* queued by svtme_shim_submit_pictures: wait for batch `seq` before the first read */
    uint64_t        seq;
    uint8_t         max_refs, max_cand;
    uint32_t        fields;  /* result groups of `res` (SVTME_F_SB_COMPACT: rows at the picture's own strides) */
    SvtMePicParams  params;
    SvtMePicResults res;
    SvtMeHmeState  *hme; /* ME_MCTF only (tf_exit) */
    void           *mem;
    struct ShimBatch *batch; /* queued by svtme_shim_submit_pictures: `res` points into the batch's slab */
} ShimJob;

/* the host slab one svtme_picture_me_batch_packed call fills (one device-to-host copy for the whole mini-GOP); freed when
   the last of its pictures has handed out its last b64 */
typedef struct ShimBatch {
    void *slab;
    int   live;
} ShimBatch;

static struct {
    pthread_mutex_t mu;
    SvtMeSession   *sess;
    uint32_t        w, h;
    int             device;
    uint64_t        clock;
    ShimSlot        slot[SHIM_SLOTS];
    ShimJob         job[SHIM_JOBS];
    uint64_t        n_pictures, n_tf_pairs, n_dg, n_batches;
} g = {.mu = PTHREAD_MUTEX_INITIALIZER};

static void shim_job_free(ShimJob *j);
static void shim_fatal(const char *what, uint32_t rc) {
    fprintf(stderr, "svtme_wrap: %s failed (0x%x): the GPU motion-estimation path has no CPU fallback\n", what, rc);
    abort();
}
#define SHIM_CHECK(call)                         \
    do {                                         \
        const uint32_t rc_ = (call);             \
        if (rc_ != SVTME_OK) shim_fatal(#call, rc_); \
    } while (0)

/* ---- session and picture store -------------------------------------------------------------- */
static void shim_session(uint32_t w, uint32_t h) {
    if (g.sess && (g.w != w || g.h != h)) {
        svtme_session_destroy(g.sess);
        g.sess = NULL;
    }
    if (!g.sess) {
        SHIM_CHECK(svtme_session_create(&g.sess, g.device, w, h, SHIM_SLOTS));
        g.w = w;
        g.h = h;
        memset(g.slot, 0, sizeof(g.slot));
    }
}

/* device slot of a host picture, uploading it when the store does not hold it. Slots touched during the current launch
   (stamp == g.clock) are never evicted.
   `scope`: the samples behind one (picture number, plane address) change during an encode: the temporal filter rewrites a
   base picture in place after its neighbours were searched against it (temporal_filtering.c:3842-3882), and the
   dynamic-GOP detector looks at pictures before they are filtered. A device copy is therefore only reused inside the
   scope it was uploaded for: open-loop ME (scope NULL: a picture is final once it is a source or a reference there), the
   temporal filtering of ONE centre picture (scope = its PCS), the detector (scope = &g). */
static int32_t shim_slot_of_plane(const uint8_t *luma, uint32_t stride, uint64_t picture_number, const void *scope) {
    int victim = -1;
    for (int i = 0; i < SHIM_SLOTS; i++) {
        ShimSlot *s = &g.slot[i];
        if (s->valid && s->picture_number == picture_number && s->buf == luma && s->scope == scope) {
            s->stamp = g.clock;
            return i;
        }
        if (!s->valid) {
            if (victim < 0 || g.slot[victim].valid) victim = i;
        } else if (s->stamp != g.clock && (victim < 0 || (g.slot[victim].valid && s->stamp < g.slot[victim].stamp)))
            victim = i;
    }
    if (victim < 0) shim_fatal("picture store full", 0);
    SHIM_CHECK(svtme_picture_upload(g.sess, victim, luma, stride));
    g.slot[victim] = (ShimSlot){1, picture_number, luma, scope, g.clock};
    return victim;
}
static int32_t shim_slot_of(const EbPictureBufferDesc *d, uint64_t picture_number, const void *scope) {
    /* scaled references (super-resolution / resize, me_process.c:232-249) have other dimensions: not supported */
    if ((uint32_t)ALIGN_POWER_OF_TWO(d->width, 3) != g.w || (uint32_t)ALIGN_POWER_OF_TWO(d->height, 3) != g.h)
        shim_fatal("reference picture size differs from the source (super-res / resize are not supported)", d->width);
    return shim_slot_of_plane(d->buffer_y + (size_t)d->org_y * d->stride_y + d->org_x, d->stride_y, picture_number, scope);
}

void svtme_shim_set_device(int device) { g.device = device; }

void svtme_shim_release_picture(uint64_t picture_number) {
    pthread_mutex_lock(&g.mu);
    for (int i = 0; i < SHIM_SLOTS; i++)
        if (g.slot[i].valid && g.slot[i].picture_number == picture_number) g.slot[i].valid = 0;
    pthread_mutex_unlock(&g.mu);
}

void svtme_shim_forget_pictures(void) {
    pthread_mutex_lock(&g.mu);
    for (int i = 0; i < SHIM_SLOTS; i++) g.slot[i].valid = 0;
    pthread_mutex_unlock(&g.mu);
}

void svtme_shim_shutdown(void) {
    pthread_mutex_lock(&g.mu);
    if (g.sess) svtme_session_destroy(g.sess);
    g.sess = NULL;
    for (int i = 0; i < SHIM_JOBS; i++) shim_job_free(&g.job[i]);
    memset(g.slot, 0, sizeof(g.slot));
    pthread_mutex_unlock(&g.mu);
}

uint64_t svtme_shim_batches(void) { return g.n_batches; }

void svtme_shim_stats(uint64_t *pictures, uint64_t *tf_pairs, uint64_t *dg_pairs) {
    pthread_mutex_lock(&g.mu);
    if (pictures) *pictures = g.n_pictures;
    if (tf_pairs) *tf_pairs = g.n_tf_pairs;
    if (dg_pairs) *dg_pairs = g.n_dg;
    pthread_mutex_unlock(&g.mu);
}

/* SVTME_SHIM_STATS=1 in the environment: one line on stderr at process exit (the whole-encoder test reads it) */
__attribute__((destructor)) static void shim_report(void) {
    const char *e = getenv("SVTME_SHIM_STATS");
    if (e && e[0] == '1')
        fprintf(stderr, "svtme_wrap: on the GPU: %llu open-loop pictures (%llu batches), %llu temporal-filter pairs, %llu detector pairs\n",
                (unsigned long long)g.n_pictures, (unsigned long long)g.n_batches, (unsigned long long)g.n_tf_pairs,
                (unsigned long long)g.n_dg);
}

/* ---- MeContext / PCS -> SvtMePicParams ------------------------------------------------------ */
static SvtMeSearchArea sa_of(SearchArea a) {
    SvtMeSearchArea r = {a.width, a.height};
    return r;
}
static SvtMeSearchAreaMinMax samm_of(SearchAreaMinMax a) {
    SvtMeSearchAreaMinMax r = {sa_of(a.sa_min), sa_of(a.sa_max)};
    return r;
}

/* Field by field what motion_estimation.c reads from the MeContext (filled by svt_aom_sig_deriv_me /
   svt_aom_sig_deriv_me_tf, enc_mode_config.c:694-895, and by the caller, me_process.c:219-262 /
   temporal_filtering.c:3112-3140) and from the PCS. Members of a disabled feature are left 0: the reference leaves them
   unset, and they are never read. */
void svtme_shim_flatten(const PictureParentControlSet *pcs, const MeContext *m, SvtMePicParams *o) {
    /* --rmv 1 clamps the full-pel window to the b64's tile instead of the padded picture (motion_estimation.c:1445-1470,
       :1507-1533). The library has no tile geometry and would search other windows than the reference: refuse, like every
       other mode it does not implement. */
    if (pcs->scs->static_config.restricted_motion_vector)
        shim_fatal("restricted_motion_vector (--rmv 1: tile-restricted search windows are not implemented)", 1);
    memset(o, 0, sizeof(*o));
    SvtMeCtrls *c = &o->ctrls;
    c->enable_hme_flag        = m->enable_hme_flag;
    c->enable_hme_level0_flag = m->enable_hme_level0_flag;
    c->enable_hme_level1_flag = m->enable_hme_level1_flag;
    c->enable_hme_level2_flag = m->enable_hme_level2_flag;
    c->hme_search_method      = m->hme_search_method;
    c->me_search_method       = m->me_search_method;
    c->num_hme_sa_w           = m->num_hme_sa_w;
    c->num_hme_sa_h           = m->num_hme_sa_h;
    c->hme_l0_sa              = samm_of(m->hme_l0_sa);
    c->hme_l1_sa              = sa_of(m->hme_l1_sa);
    c->hme_l2_sa              = sa_of(m->hme_l2_sa);
    c->me_sa                  = samm_of(m->me_sa);
    if (m->prehme_ctrl.enable) {
        c->prehme_enable           = m->prehme_ctrl.enable;
        c->prehme_skip_search_line = m->prehme_ctrl.skip_search_line;
        c->prehme_l1_early_exit    = m->prehme_ctrl.l1_early_exit;
        c->prehme_sa_cfg[0]        = samm_of(m->prehme_ctrl.prehme_sa_cfg[0]);
        c->prehme_sa_cfg[1]        = samm_of(m->prehme_ctrl.prehme_sa_cfg[1]);
    }
    const MeHmeRefPruneCtrls *p = &m->me_hme_prune_ctrls;
    c->enable_me_hme_ref_pruning               = p->enable_me_hme_ref_pruning;
    c->prune_ref_if_hme_sad_dev_bigger_than_th = p->prune_ref_if_hme_sad_dev_bigger_than_th;
    c->prune_ref_if_me_sad_dev_bigger_than_th  = p->prune_ref_if_me_sad_dev_bigger_than_th;
    c->zz_sad_th    = p->zz_sad_th;
    c->zz_sad_pct   = p->zz_sad_pct;
    c->phme_sad_th  = p->phme_sad_th;
    c->phme_sad_pct = p->phme_sad_pct;
    const MeSrCtrls *s = &m->me_sr_adjustment_ctrls;
    if (s->enable_me_sr_adjustment) {
        c->enable_me_sr_adjustment              = s->enable_me_sr_adjustment; /* 1 or 2 (enc_mode_config.c:470-499) */
        c->reduce_me_sr_based_on_mv_length_th   = s->reduce_me_sr_based_on_mv_length_th;
        c->stationary_hme_sad_abs_th            = s->stationary_hme_sad_abs_th;
        c->stationary_me_sr_divisor             = s->stationary_me_sr_divisor;
        c->reduce_me_sr_based_on_hme_sad_abs_th = s->reduce_me_sr_based_on_hme_sad_abs_th;
        c->me_sr_divisor_for_low_hme_sad        = s->me_sr_divisor_for_low_hme_sad;
        c->distance_based_hme_resizing          = s->distance_based_hme_resizing;
    }
    if (m->me_8x8_var_ctrls.enabled) {
        c->me_8x8_var_enabled = m->me_8x8_var_ctrls.enabled;
        c->me_sr_div4_th      = m->me_8x8_var_ctrls.me_sr_div4_th;
        c->me_sr_div2_th      = m->me_8x8_var_ctrls.me_sr_div2_th;
        c->me_sr_mult2_th     = m->me_8x8_var_ctrls.me_sr_mult2_th;
    }
    if (m->mv_based_sa_adj.enabled) {
        c->mv_sa_adj_enabled          = m->mv_based_sa_adj.enabled;
        c->mv_sa_adj_nearest_ref_only = m->mv_based_sa_adj.nearest_ref_only;
        c->mv_sa_adj_mv_size_th       = m->mv_based_sa_adj.mv_size_th;
        c->mv_sa_adj_sa_multiplier    = m->mv_based_sa_adj.sa_multiplier;
    }
    c->reduce_hme_l0_sr_th_min     = m->reduce_hme_l0_sr_th_min;
    c->reduce_hme_l0_sr_th_max     = m->reduce_hme_l0_sr_th_max;
    c->me_early_exit_th            = m->me_early_exit_th;
    c->me_safe_limit_zz_th         = m->me_safe_limit_zz_th;
    c->prev_me_stage_based_exit_th = m->prev_me_stage_based_exit_th;
    c->prune_me_candidates_th      = m->prune_me_candidates_th;
    c->use_best_unipred_cand_only  = m->use_best_unipred_cand_only;

    o->picture_number              = pcs->picture_number;
    o->num_of_list_to_search       = m->num_of_list_to_search;
    o->num_of_ref_pic_to_search[0] = m->num_of_ref_pic_to_search[0];
    o->num_of_ref_pic_to_search[1] = m->num_of_list_to_search > 1 ? m->num_of_ref_pic_to_search[1] : 0;
    o->temporal_layer_index        = m->temporal_layer_index;
    o->is_ref                      = m->is_ref;
    o->hierarchical_levels         = pcs->hierarchical_levels;
    o->input_resolution            = (uint8_t)pcs->scs->input_resolution;
    if (m->me_type == ME_MCTF) {
        c->me_type       = SVTME_ME_MCTF;
        c->tf_me_exit_th = m->tf_me_exit_th;
        /* the candidate-list / GM outputs are not produced for ME_MCTF (motion_estimation.c:3195): neutral geometry */
        o->enable_me_8x8 = o->enable_me_16x16 = 1;
        o->max_cand = 3; o->max_refs = 2; o->max_l0 = 1;
    } else {
        c->me_type                     = SVTME_ME_OPEN_LOOP;
        o->similar_brightness_refs     = pcs->similar_brightness_refs;
        o->enable_me_8x8               = pcs->enable_me_8x8;
        o->enable_me_16x16             = pcs->enable_me_16x16;
        o->max_cand                    = pcs->pa_me_data->max_cand;
        o->max_refs                    = pcs->pa_me_data->max_refs;
        o->max_l0                      = pcs->pa_me_data->max_l0;
        o->only_l_bwd                  = pcs->scs->mrp_ctrls.only_l_bwd;
        o->gm_enabled                  = pcs->gm_ctrls.enabled;
        o->gm_use_distance_based_active_th = pcs->gm_ctrls.use_distance_based_active_th;
    }
    for (int l = 0; l < 2; l++)
        for (int r = 0; r < 4; r++) {
            o->ref_pic_id[l][r] = -1;
            if (l < o->num_of_list_to_search && r < o->num_of_ref_pic_to_search[l])
                o->ref_picture_number[l][r] = m->me_ds_ref_array[l][r].picture_number;
        }
}

/* ---- jobs: one per picture (open loop) or per (centre, neighbour) pair (temporal filter) ----- */
static void shim_job_free(ShimJob *j) {
    free(j->mem);
    if (j->batch && --j->batch->live == 0) {
        svtme_host_free(j->batch->slab);
        free(j->batch);
    }
    memset(j, 0, sizeof(*j));
}

static ShimJob *shim_find_job(const void *pcs, int me_type, uint64_t ref_number) {
    for (int i = 0; i < SHIM_JOBS; i++) {
        ShimJob *j = &g.job[i];
        if (j->used && j->pcs == pcs && j->me_type == me_type && j->ref_number == ref_number) return j;
    }
    return NULL;
}

static ShimJob *shim_new_job(const void *pcs, int me_type, uint64_t ref_number, uint32_t n_b64, const SvtMePicParams *p, int own_arrays) {
    ShimJob *j = NULL;
    for (int i = 0; i < SHIM_JOBS && !j; i++)
        if (!g.job[i].used) j = &g.job[i];
    if (!j) {
        /* results nobody came back for: the temporal filter skips the ME call of some blocks (temporal_filtering.c:3080-3094),
           so a (centre, neighbour) pair may never see all its b64s. The oldest set goes; if one of its b64s is still asked
           for, the pair is simply launched again. */
        for (int i = 0; i < SHIM_JOBS; i++)
            if (!g.job[i].pending && !g.job[i].readers && (!j || g.job[i].stamp < j->stamp)) j = &g.job[i];
        if (!j) shim_fatal("job table full", 0);
        shim_job_free(j);
    }
    memset(j, 0, sizeof(*j));
    j->stamp = g.clock;
    j->used = 1; j->pcs = pcs; j->me_type = me_type; j->ref_number = ref_number;
    j->n_b64 = n_b64; j->n_pus = p->enable_me_8x8 ? 85 : 21; j->max_refs = p->max_refs; j->max_cand = p->max_cand;
    j->params = *p;
    j->fields = SVTME_F_SB_RESULTS | SVTME_F_STATS | SVTME_F_SEARCH_RESULTS | SVTME_F_BEST;
    if (!own_arrays) return j; /* the caller points j->res into a batch slab */
    const size_t n = n_b64, npu = j->n_pus;
    const size_t sz[] = {n * 680 * 4, n * 680 * 4, n * 8 * sizeof(SvtMeSearchResult), n * npu, n * npu * j->max_refs * 4,
                         n * npu * j->max_cand, n * 4, n * 4, n * 4, n * 4, n * 4, n * 4, n, n,
                         me_type == ME_MCTF ? n * sizeof(SvtMeHmeState) : 0};
    size_t total = 0, off[15];
    for (int i = 0; i < 15; i++) { off[i] = total; total += (sz[i] + 63) & ~(size_t)63; }
    uint8_t *mem = (uint8_t *)calloc(1, total ? total : 64);
    if (!mem) shim_fatal("result allocation", 0);
    j->mem = mem;
    SvtMePicResults *r = &j->res;
    r->p_sb_best_sad  = (uint32_t *)(mem + off[0]);
    r->p_sb_best_mv   = (uint32_t *)(mem + off[1]);
    r->search_results = (SvtMeSearchResult *)(mem + off[2]);
    if (me_type != ME_MCTF) {
        r->total_me_candidate_index    = mem + off[3];
        r->me_mv_array                 = (uint32_t *)(mem + off[4]);
        r->me_candidate_array          = mem + off[5];
        r->me_64x64_distortion         = (uint32_t *)(mem + off[6]);
        r->me_32x32_distortion         = (uint32_t *)(mem + off[7]);
        r->me_16x16_distortion         = (uint32_t *)(mem + off[8]);
        r->me_8x8_distortion           = (uint32_t *)(mem + off[9]);
        r->rc_me_distortion            = (uint32_t *)(mem + off[10]);
        r->me_8x8_cost_variance        = (uint32_t *)(mem + off[11]);
        r->stationary_block_present_sb = mem + off[12];
        r->rc_me_allow_gm              = mem + off[13];
    } else
        j->hme = (SvtMeHmeState *)(mem + off[14]);
    return j;
}

/* launch the whole picture: uploads (source + every reference the store does not hold), one svtme_picture_me */
static ShimJob *shim_launch(PictureParentControlSet *pcs, MeContext *me_ctx, EbPictureBufferDesc *input, uint32_t b64_origin_x,
                            uint32_t b64_origin_y, uint64_t ref_number) {
    const uint32_t w = (uint32_t)ALIGN_POWER_OF_TWO(input->width, 3), h = (uint32_t)ALIGN_POWER_OF_TWO(input->height, 3);
    shim_session(w, h);
    g.clock++;
    SvtMePicParams p;
    svtme_shim_flatten(pcs, me_ctx, &p);
    /* the source samples are the ones the reference would read: me_ctx->b64_src_ptr points at this b64 inside the padded
       source plane (me_process.c:196-197: the PA reference object's input_padded_pic, not `input`) */
    const uint8_t *src_luma = me_ctx->b64_src_ptr - (size_t)b64_origin_y * me_ctx->b64_src_stride - b64_origin_x;
    const void    *scope    = me_ctx->me_type == ME_MCTF ? (const void *)pcs : NULL;
    const int32_t  src_slot = shim_slot_of_plane(src_luma, me_ctx->b64_src_stride, pcs->picture_number, scope);
    for (int l = 0; l < p.num_of_list_to_search; l++)
        for (int r = 0; r < p.num_of_ref_pic_to_search[l]; r++)
            p.ref_pic_id[l][r] = shim_slot_of(me_ctx->me_ds_ref_array[l][r].picture_ptr, me_ctx->me_ds_ref_array[l][r].picture_number, scope);
    ShimJob *j = shim_new_job(pcs, me_ctx->me_type, ref_number, ((w + 63) / 64) * ((h + 63) / 64), &p, 1);
    SHIM_CHECK(svtme_picture_me(g.sess, src_slot, &p, &j->res, /*sync*/ 1));
    if (j->hme) SHIM_CHECK(svtme_fetch_hme_state(g.sess, j->hme, (size_t)j->n_b64 * sizeof(SvtMeHmeState)));
    if (me_ctx->me_type == ME_MCTF) g.n_tf_pairs++;
    else g.n_pictures++;
    return j;
}

EbErrorType __real_svt_aom_motion_estimation_b64(PictureParentControlSet *pcs, uint32_t b64_index, uint32_t b64_origin_x,
                                                 uint32_t b64_origin_y, MeContext *me_ctx, EbPictureBufferDesc *input_ptr);

EbErrorType __wrap_svt_aom_motion_estimation_b64(PictureParentControlSet *pcs, uint32_t b64_index, uint32_t b64_origin_x,
                                                 uint32_t b64_origin_y, MeContext *me_ctx, EbPictureBufferDesc *input_ptr) {
    if (me_ctx->me_type != ME_OPEN_LOOP && me_ctx->me_type != ME_MCTF)
        return __real_svt_aom_motion_estimation_b64(pcs, b64_index, b64_origin_x, b64_origin_y, me_ctx, input_ptr);
    const int      mctf       = me_ctx->me_type == ME_MCTF;
    const uint64_t ref_number = mctf ? me_ctx->me_ds_ref_array[0][0].picture_number : 0;
    pthread_mutex_lock(&g.mu);
    ShimJob *j = shim_find_job(pcs, me_ctx->me_type, ref_number);
    if (!j) j = shim_launch(pcs, me_ctx, input_ptr, b64_origin_x, b64_origin_y, ref_number); /* whichever thread arrives first launches */
    if (j->pending) { /* queued ahead of time as part of a batch: the first reader waits for it */
        SHIM_CHECK(svtme_session_wait_batch(g.sess, j->seq));
        j->pending = 0;
    }
    j->readers++;
    pthread_mutex_unlock(&g.mu);
    if (b64_index >= j->n_b64) shim_fatal("b64 index out of range", b64_index);

    /* motion_estimation.c:3163-3170 */
    const uint16_t aw = (uint16_t)ALIGN_POWER_OF_TWO(input_ptr->width, 3), ah = (uint16_t)ALIGN_POWER_OF_TWO(input_ptr->height, 3);
    me_ctx->b64_width  = (aw - b64_origin_x) < BLOCK_SIZE_64 ? aw - b64_origin_x : BLOCK_SIZE_64;
    me_ctx->b64_height = (ah - b64_origin_y) < BLOCK_SIZE_64 ? ah - b64_origin_y : BLOCK_SIZE_64;

    const SvtMePicResults *r = &j->res;
    for (int l = 0; l < 2; l++)
        for (int k = 0; k < 4; k++) {
            const size_t             o  = ((size_t)b64_index * 2 + l) * 4 + k;
            const SvtMeSearchResult *sr = &r->search_results[o];
            me_ctx->search_results[l][k].hme_sc_x = sr->hme_sc_x;
            me_ctx->search_results[l][k].hme_sc_y = sr->hme_sc_y;
            me_ctx->search_results[l][k].hme_sad  = sr->hme_sad;
            me_ctx->search_results[l][k].do_ref   = sr->do_ref;
            memcpy(me_ctx->p_sb_best_sad[l][k], r->p_sb_best_sad + o * 85, 85 * 4);
            memcpy(me_ctx->p_sb_best_mv[l][k], r->p_sb_best_mv + o * 85, 85 * 4);
        }
    /* the per-size views of the best arrays, as integer_search_b64 leaves them (motion_estimation.c:1370-1386): on the last
       (list, reference) it searched. Read outside motion_estimation.c only by the temporal filter, whose single
       reference is [0][0] (temporal_filtering.c:1860-2226, :2672). */
    {
        const int ll = me_ctx->num_of_list_to_search > 1 && me_ctx->num_of_ref_pic_to_search[1] ? 1 : 0;
        const int lr = me_ctx->num_of_ref_pic_to_search[ll] ? me_ctx->num_of_ref_pic_to_search[ll] - 1 : 0;
        me_ctx->p_best_sad_64x64 = &me_ctx->p_sb_best_sad[ll][lr][ME_TIER_ZERO_PU_64x64];
        me_ctx->p_best_sad_32x32 = &me_ctx->p_sb_best_sad[ll][lr][ME_TIER_ZERO_PU_32x32_0];
        me_ctx->p_best_sad_16x16 = &me_ctx->p_sb_best_sad[ll][lr][ME_TIER_ZERO_PU_16x16_0];
        me_ctx->p_best_sad_8x8   = &me_ctx->p_sb_best_sad[ll][lr][ME_TIER_ZERO_PU_8x8_0];
        me_ctx->p_best_mv64x64   = &me_ctx->p_sb_best_mv[ll][lr][ME_TIER_ZERO_PU_64x64];
        me_ctx->p_best_mv32x32   = &me_ctx->p_sb_best_mv[ll][lr][ME_TIER_ZERO_PU_32x32_0];
        me_ctx->p_best_mv16x16   = &me_ctx->p_sb_best_mv[ll][lr][ME_TIER_ZERO_PU_16x16_0];
        me_ctx->p_best_mv8x8     = &me_ctx->p_sb_best_mv[ll][lr][ME_TIER_ZERO_PU_8x8_0];
    }
    if (mctf) {
        /* motion_estimation.c:2539-2544, :3179-3183 */
        if (ABS(me_ctx->search_results[0][0].hme_sc_x) > ABS(me_ctx->search_results[0][0].hme_sc_y)) me_ctx->tf_tot_horz_blks++;
        else me_ctx->tf_tot_vert_blks++;
        if (j->hme[b64_index].tf_exit) me_ctx->tf_use_pred_64x64_only_th = (uint8_t)~0;
    } else {
        MeSbResults *sb = pcs->pa_me_data->me_results[b64_index];
        /* this b64's rows at the reference's strides (a batch downloads them at the picture's own, narrower strides) */
        svtme_sb_expand_b64(&j->params, j->fields, r, b64_index, sb->total_me_candidate_index, (uint32_t *)sb->me_mv_array,
                            (uint8_t *)sb->me_candidate_array);
        pcs->me_64x64_distortion[b64_index]         = r->me_64x64_distortion[b64_index];
        pcs->me_32x32_distortion[b64_index]         = r->me_32x32_distortion[b64_index];
        pcs->me_16x16_distortion[b64_index]         = r->me_16x16_distortion[b64_index];
        pcs->me_8x8_distortion[b64_index]           = r->me_8x8_distortion[b64_index];
        pcs->rc_me_distortion[b64_index]            = r->rc_me_distortion[b64_index];
        pcs->me_8x8_cost_variance[b64_index]        = r->me_8x8_cost_variance[b64_index];
        pcs->stationary_block_present_sb[b64_index] = r->stationary_block_present_sb[b64_index];
        pcs->rc_me_allow_gm[b64_index]              = r->rc_me_allow_gm[b64_index];
    }

    pthread_mutex_lock(&g.mu);
    j->readers--;
    if (++j->delivered == j->n_b64) shim_job_free(j); /* every b64 is asked for exactly once (me_process.c:174-271) */
    pthread_mutex_unlock(&g.mu);
    return EB_ErrorNone;
}

/* ---- a mini-GOP at once ---------------------------------------------------------------------------------------------
   Picture Decision releases the pictures of a mini-GOP together (pd_process.c:4373-4400) and open-loop ME pictures do
   not depend on each other, so their ME can be ONE GPU batch (every kernel launched once for all of them) queued before
   the first TASK_PAME task is consumed; the b64 calls then only wait for the batch and copy. For every picture this does
   what svt_aom_motion_estimation_kernel does before its b64 loop (me_process.c:121, :131-146, :219-262). */
static int same_stage_structure(const SvtMeCtrls *a, const SvtMeCtrls *b) {
    return a->enable_hme_flag == b->enable_hme_flag && a->enable_hme_level0_flag == b->enable_hme_level0_flag &&
        a->enable_hme_level1_flag == b->enable_hme_level1_flag && a->enable_hme_level2_flag == b->enable_hme_level2_flag &&
        a->prehme_enable == b->prehme_enable && a->hme_search_method == b->hme_search_method &&
        a->me_search_method == b->me_search_method && a->prehme_skip_search_line == b->prehme_skip_search_line &&
        a->hme_l1_sa.width == b->hme_l1_sa.width && a->hme_l1_sa.height == b->hme_l1_sa.height &&
        a->hme_l2_sa.width == b->hme_l2_sa.width && a->hme_l2_sa.height == b->hme_l2_sa.height;
}

void svtme_shim_submit_pictures(PictureParentControlSet **pcs_list, uint32_t n) {
    if (n == 0) return;
    if (n > SHIM_JOBS / 2) shim_fatal("svtme_shim_submit_pictures: too many pictures", n);
    MeContext      *m      = (MeContext *)calloc(1, sizeof(MeContext));
    SvtMePicParams *params = (SvtMePicParams *)calloc(n, sizeof(SvtMePicParams));
    int32_t        *ids    = (int32_t *)calloc(n, sizeof(int32_t));
    ShimJob       **jobs   = (ShimJob **)calloc(n, sizeof(ShimJob *));
    if (!m || !params || !ids || !jobs) shim_fatal("allocation", 0);
    pthread_mutex_lock(&g.mu);
    uint32_t k = 0;
    for (uint32_t i = 0; i < n; i++) {
        PictureParentControlSet *pcs = pcs_list[i];
        if (pcs->slice_type == I_SLICE || shim_find_job(pcs, ME_OPEN_LOOP, 0)) continue; /* me_process.c:165 */
        const EbPaReferenceObject *po  = (const EbPaReferenceObject *)pcs->pa_ref_pic_wrapper->object_ptr;
        const EbPictureBufferDesc *src = po->input_padded_pic;
        if (k == 0) {
            shim_session((uint32_t)ALIGN_POWER_OF_TWO(src->width, 3), (uint32_t)ALIGN_POWER_OF_TWO(src->height, 3));
            g.clock++;
        }
        memset(m, 0, sizeof(*m));
        m->me_type = ME_OPEN_LOOP;
        svt_aom_sig_deriv_me(pcs->scs, pcs, m);
        m->num_of_list_to_search       = (pcs->slice_type == P_SLICE) ? 1 : 2;
        m->num_of_ref_pic_to_search[0] = pcs->ref_list0_count_try;
        if (pcs->slice_type == B_SLICE) m->num_of_ref_pic_to_search[1] = pcs->ref_list1_count_try;
        m->temporal_layer_index = pcs->temporal_layer_index;
        m->is_ref               = pcs->is_ref;
        for (int l = 0; l < m->num_of_list_to_search; l++)
            for (int r = 0; r < m->num_of_ref_pic_to_search[l]; r++) {
                const EbPaReferenceObject *ro = (const EbPaReferenceObject *)pcs->ref_pa_pic_ptr_array[l][r]->object_ptr;
                m->me_ds_ref_array[l][r].picture_ptr    = ro->input_padded_pic;
                m->me_ds_ref_array[l][r].picture_number = ro->picture_number;
            }
        SvtMePicParams *p = &params[k];
        svtme_shim_flatten(pcs, m, p);
        if (k && !same_stage_structure(&params[0].ctrls, &p->ctrls)) continue; /* left to the per-picture launch */
        ids[k] = shim_slot_of(src, pcs->picture_number, NULL);
        for (int l = 0; l < p->num_of_list_to_search; l++)
            for (int r = 0; r < p->num_of_ref_pic_to_search[l]; r++)
                p->ref_pic_id[l][r] = shim_slot_of(m->me_ds_ref_array[l][r].picture_ptr, m->me_ds_ref_array[l][r].picture_number, NULL);
        jobs[k] = shim_new_job(pcs, ME_OPEN_LOOP, 0, pcs->b64_total_count, p, 0);
        k++;
    }
    if (k) {
        /* what the b64 calls hand out: the MeSbResults arrays and PCS statistics, plus the MeContext members */
        const uint32_t   fields = SVTME_F_SB_RESULTS | SVTME_F_SB_COMPACT | SVTME_F_STATS | SVTME_F_SEARCH_RESULTS | SVTME_F_BEST;
        SvtMePackedLayout *lay  = (SvtMePackedLayout *)calloc(k, sizeof(SvtMePackedLayout));
        ShimBatch         *bt   = (ShimBatch *)calloc(1, sizeof(ShimBatch));
        if (!lay || !bt) shim_fatal("allocation", 0);
        size_t total = 0;
        for (uint32_t i = 0; i < k; i++) {
            SHIM_CHECK(svtme_packed_layout(g.w, g.h, &params[i], fields, &lay[i]));
            total += lay[i].bytes;
        }
        bt->slab = svtme_host_alloc(total);
        if (!bt->slab) shim_fatal("pinned result slab", 0);
        bt->live = (int)k;
        size_t at = 0;
        for (uint32_t i = 0; i < k; i++) {
            svtme_packed_view(&lay[i], (uint8_t *)bt->slab + at, &jobs[i]->res);
            jobs[i]->batch  = bt;
            jobs[i]->fields = fields;
            at += lay[i].bytes;
        }
        SHIM_CHECK(svtme_picture_me_batch_packed(g.sess, k, ids, params, fields, bt->slab, total, /*sync*/ 0));
        const uint64_t seq = svtme_session_batch_seq(g.sess);
        for (uint32_t i = 0; i < k; i++) {
            jobs[i]->pending = 1;
            jobs[i]->seq     = seq;
        }
        g.n_pictures += k;
        g.n_batches++;
        free(lay);
    }
    pthread_mutex_unlock(&g.mu);
    free(m); free(params); free(ids); free(jobs);
}

/* ---- dynamic-GOP detector (pd_process.c:490-589; one call per ME segment, me_process.c:329) ----------------------- */
void __wrap_dg_detector_hme_level0(struct PictureParentControlSet *ppcs, uint32_t seg_idx) {
    (void)seg_idx;
    DGDetectorSeg *dg = ppcs->dg_detector;
    svt_block_on_mutex(dg->metrics_mutex);
    if (dg->metrics.seg_completed == 0) { /* the first segment does the whole picture, the others only count down */
        const EbPaReferenceObject *so = (const EbPaReferenceObject *)ppcs->pa_ref_pic_wrapper->object_ptr;
        const EbPaReferenceObject *ro = (const EbPaReferenceObject *)dg->ref_pic->pa_ref_pic_wrapper->object_ptr;
        pthread_mutex_lock(&g.mu);
        shim_session(ppcs->aligned_width, ppcs->aligned_height);
        g.clock++;
        const int32_t  s = shim_slot_of(so->input_padded_pic, ppcs->picture_number, &g);
        const int32_t  r = shim_slot_of(ro->input_padded_pic, dg->ref_pic->picture_number, &g);
        SvtMeDgMetrics m = {.input_resolution = (uint32_t)ppcs->input_resolution};
        SHIM_CHECK(svtme_dg_detector_hme(g.sess, s, r, &m, NULL, NULL));
        g.n_dg++;
        pthread_mutex_unlock(&g.mu);
        dg->metrics.tot_dist += m.tot_dist;
        dg->metrics.tot_cplx += m.tot_cplx;
        dg->metrics.tot_active += m.tot_active;
        dg->metrics.sum_in_vectors += m.sum_in_vectors;
    }
    dg->metrics.seg_completed++;
    if (dg->metrics.seg_completed == (ppcs->me_segments_column_count * ppcs->me_segments_row_count))
        svt_post_semaphore(dg->frame_done_sem); /* pd_process.c:583-586 */
    svt_release_mutex(dg->metrics_mutex);
}
