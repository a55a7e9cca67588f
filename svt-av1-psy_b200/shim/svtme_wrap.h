/* svtme_wrap.h — what the encoder's glue code may call besides the two wrapped functions (svtme_wrap.c). */
#ifndef SVTME_WRAP_H
#define SVTME_WRAP_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif
struct PictureParentControlSet;

/* CUDA device of the session the shim creates on first use (default 0); call before the first picture. */
void svtme_shim_set_device(int device);

/* Queue the open-loop ME of n pictures as ONE GPU batch (every kernel launched once for all of them) and return without
   waiting. Call it where Picture Decision has released the pictures of a mini-GOP (pd_process.c:4373-4400: reference
   lists final, samples final, i.e. after the temporal filter of the base pictures), before their TASK_PAME tasks are
   consumed. The b64 calls of these pictures then wait for the batch instead of launching. I pictures, pictures already
   queued and pictures whose HME stage structure differs from the first one's are skipped (the latter are launched on
   their first b64 call as usual). n <= 32. */
void svtme_shim_submit_pictures(struct PictureParentControlSet **pcs_list, uint32_t n);
/* batches queued by svtme_shim_submit_pictures so far */
uint64_t svtme_shim_batches(void);

/* Drop the device copy of a picture. Hook: svt_aom_release_pa_reference_objects (initial_rc_process.c:721-736), and
   after the temporal filter rewrote a base picture in place (temporal_filtering.c:3842-3882) so that open-loop ME
   uploads the filtered samples. */
void svtme_shim_release_picture(uint64_t picture_number);
/* Drop every device copy (end of stream; the test harness calls it between unrelated pictures). */
void svtme_shim_forget_pictures(void);

/* Destroy the session and every pending result (svt_av1_enc_deinit). */
void svtme_shim_shutdown(void);

/* Pictures, temporal-filter pairs and detector pairs that were run on the GPU so far. */
void svtme_shim_stats(uint64_t *pictures, uint64_t *tf_pairs, uint64_t *dg_pairs);

#ifdef __cplusplus
}
#endif
#endif
