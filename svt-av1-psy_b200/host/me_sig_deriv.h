/*
 * me_sig_deriv.h — host-side (plain C) derivation of the ME launch parameters.
 * Restates svt_aom_sig_deriv_me (Source/Lib/Codec/enc_mode_config.c:694-843) and the PCS-level
 * fields Picture Decision / multi-process signal derivation set for the ME stage, so that a caller
 * without the reference encoder (bench.py, tests) obtains exactly the controls the encoder would
 * use. Inside the encoder the reference's own function runs unchanged and its MeContext is
 * flattened into SvtMeCtrls.
 */
#ifndef SVTME_SIG_DERIV_H
#define SVTME_SIG_DERIV_H
#include <stdint.h>
#include "svtme_b200.h"
#ifdef __cplusplus
extern "C" {
#endif

typedef struct SvtMeSigDerivIn {
    int32_t  enc_mode;            /* EncMode (EbSvtAv1Enc.h:46-63): -3 MRS, -2 MRP, -1 MR, 0..13 */
    int32_t  width, height;       /* luma size, multiples of 8 */
    int32_t  qp;                  /* static_config.qp */
    int32_t  pred_structure;      /* 1 low delay B, 2 random access */
    int32_t  hierarchical_levels;
    int32_t  temporal_layer_index;
    int32_t  is_ref;
    int32_t  sc_class1;
    uint32_t frame_rate;          /* Q16 */
    int32_t  similar_brightness_refs;
    int32_t  gm_level;            /* -1: derive (svt_aom_get_gm_core_level) */
    int32_t  only_l_bwd, safe_limit_nref, safe_limit_zz_th; /* scs->mrp_ctrls */
    int32_t  max_l0, max_l1;      /* ref_count_used_list{0,1} */
    uint64_t picture_number;
    int32_t  num_lists;
    int32_t  ref_count[2];
    uint64_t ref_picture_number[2][4];
} SvtMeSigDerivIn;

uint32_t svtme_sig_deriv_me(const SvtMeSigDerivIn *in, SvtMePicParams *out);

/* Temporal-filter ME controls (ME_MCTF): svt_aom_sig_deriv_me_tf + tf_set_me_hme_params_oq (enc_mode_config.c:611-688,
   :849-895), the tf_enable_hme_* switch (:1832-1856), and what create_me_context_and_picture_control /
   set_hme_search_params_mctf / the block loop set per block (temporal_filtering.c:288-343, :2752-2760, :3112-3140):
   one list, one reference (the neighbouring picture being aligned to the centre picture), FULL_SAD HME. */
typedef struct SvtMeSigDerivTfIn {
    int32_t  enc_mode, width, height, qp;
    int32_t  tf_hme_me_level; /* pcs->tf_ctrls.hme_me_level, 0..4 */
    int32_t  tf_qp_opt;       /* pcs->tf_ctrls.qp_opt */
    uint32_t tf_me_exit_th;   /* pcs->tf_ctrls.me_exit_th (stored in a uint16_t by the reference) */
    int32_t  temporal_layer_index, is_ref; /* of the centre picture */
    uint64_t centre_picture_number, ref_picture_number;
} SvtMeSigDerivTfIn;
uint32_t svtme_sig_deriv_me_tf(const SvtMeSigDerivTfIn *in, SvtMePicParams *out);

/* Picture-level sharding of the ME stage over GPUs: mini-GOP g of a stream goes to
   rank (g + stream) % world; pictures inside a mini-GOP stay together so most refs are local. */
int32_t svtme_shard_owner(int64_t picture_number, int32_t mini_gop_size, int32_t stream, int32_t world_size);

#ifdef __cplusplus
}
#endif
#endif
