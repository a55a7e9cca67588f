/* svtme_wrap.h — what the encoder's glue code may call besides the two wrapped functions (svtme_wrap.c). */
#ifndef SVTME_WRAP_H
#define SVTME_WRAP_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* CUDA device of the session the shim creates on first use (default 0); call before the first picture. */
void svtme_shim_set_device(int device);

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
