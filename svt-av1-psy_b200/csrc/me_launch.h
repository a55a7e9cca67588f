// me_launch.h — host-callable launchers defined in me_kernels.cu (internal to the library).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "svtme_b200.h"

struct DevPic;
struct PicJob;

// every launcher returns the number of kernels it launched on `st`
int  svtme_launch_build_pyramid(const DevPic &pic, int mode, const void *d_in, int in_stride, cudaStream_t st);
int  svtme_launch_build_pyramids(const DevPic *pics, const void *const *d_in, int in_stride, int n, int mode, cudaStream_t st);
int  svtme_launch_downsample_generic(const uint8_t *d_in, int in_stride, int in_w, int in_h, uint8_t *d_out, int out_stride,
                                     int step, cudaStream_t st);
uint32_t svtme_dg_units_cap(int sa0, int n_b64);
int  svtme_launch_dg_detector(const PicJob *h_job, const PicJob *d_job, int sa0, void *d_out, unsigned long long *d_b64_sad,
                              int16_t *d_b64_mv, cudaStream_t st);
uint32_t svtme_fp_units_cap(const SvtMePicParams &pp, int n_b64);
uint32_t svtme_s16_units_cap(const SvtMePicParams &pp, int n_b64);
// per-kernel timeline of one launch sequence: event i was recorded behind the launch tagged tag[i]
enum { SVTME_TL_START = 0, SVTME_TL_ZZ, SVTME_TL_CTL, SVTME_TL_S16_PREHME, SVTME_TL_S16_L0, SVTME_TL_BLK_L1, SVTME_TL_BLK_L2, SVTME_TL_FP,
       SVTME_TL_FP_BIG, SVTME_TL_POST, SVTME_TL_TAGS };
#define SVTME_TL_MAX 32
struct SvtMeTimeline {
    cudaEvent_t ev[SVTME_TL_MAX];
    int         tag[SVTME_TL_MAX];
    int         n;
};
int  svtme_launch_picture_me(const PicJob *h_jobs, int n, const PicJob *d_jobs, cudaStream_t st, cudaEvent_t ev_hme0,
                             cudaEvent_t ev_hme1, cudaEvent_t ev_fp1, SvtMeTimeline *tl, const cudaEvent_t *out_free, int n_out_free);
int  svtme_launch_kat_sad_loop(const uint8_t *d_src, int src_stride, const uint8_t *d_ref, int ref_stride, int bh, int bw,
                               int src_stride_raw, int skip, int sa_w, int sa_h, unsigned long long *d_key, uint8_t *d_pitched,
                               size_t pitched_cap, cudaStream_t st);
int  svtme_launch_kat_octet(const uint8_t *d_src, int src_stride, const uint8_t *d_ref, int ref_pitch, int ref_rows, int sub,
                            int n_pos, uint32_t *d_out, cudaStream_t st);
// CUtensorMap (128 bytes at `out`) of a u8 plane: rows of `pitch` bytes (multiple of 16), base 16-byte aligned. 0 = ok.
int  svtme_encode_tmap(void *out, const void *base, size_t pitch, size_t rows, uint32_t box_w, uint32_t box_h);
int  svtme_launch_kat_nxm(const uint8_t *d_src, int src_stride, const uint8_t *d_ref, int ref_stride, int h, int w,
                          uint32_t *d_out, cudaStream_t st);
void svtme_launch_zero_words(void *d_p, int n_words, cudaStream_t st);
// dst: device memory, src: pinned host memory (device-visible under unified addressing); sizes rounded up to 16 bytes
void svtme_launch_copy_from_pinned(void *d_dst, const void *h_src, size_t bytes, cudaStream_t st);
int  svtme_launch_fill32(uint32_t *d_p, uint32_t n, uint32_t v, cudaStream_t st);
