// me_kernels.cu — kernel definitions and their host-side launchers (sm_100a only).
#include "me_launch.h"

#include "me_device.cuh"
#include "me_fullpel.cuh"
#include "me_hme_kernel.cuh"
#include "me_post.cuh"

// ------------------------------------------------------------------------------------------
// pyramid: replicate padding + 2x2 decimation (pic_operators.c:338-384,
// pic_analysis_process.c:131-161, :2139-2196), 10-bit -> 8 MSBs (pack_unpack_c.c:77-116)
// ------------------------------------------------------------------------------------------

// Fill the padding of a plane whose interior is already in place: every padded sample is the
// interior sample at the clamped coordinate (what horizontal-then-vertical replication produces).
__global__ void k_pad_plane(uint8_t *buf, int stride, int pad, int w, int h) {
    const int W = w + 2 * pad, H = h + 2 * pad;
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
    if (x >= W || y >= H) return;
    const int ix = x - pad, iy = y - pad;
    if (ix >= 0 && ix < w && iy >= 0 && iy < h) return;
    const int cx = min(max(ix, 0), w - 1), cy = min(max(iy, 0), h - 1);
    buf[(size_t)y * stride + x] = buf[(size_t)(cy + pad) * stride + cx + pad];
}

// dst (incl. its padding) <- 2x2 box decimation of the interior of src; step 2 only.
// out[x,y] = (in[2x,2y] + in[2x+1,2y] + in[2x,2y+1] + in[2x+1,2y+1] + 2) >> 2
__global__ void k_downsample2(const uint8_t *src, int sstride, int spad, uint8_t *dst, int dstride, int dpad, int dw,
                              int dh) {
    const int W = dw + 2 * dpad, H = dh + 2 * dpad;
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
    if (x >= W || y >= H) return;
    const int cx = min(max(x - dpad, 0), dw - 1), cy = min(max(y - dpad, 0), dh - 1);
    const uint8_t *p = src + (size_t)(2 * cy + spad) * sstride + 2 * cx + spad;
    const uint32_t s = (uint32_t)p[0] + p[1] + p[sstride] + p[sstride + 1];
    dst[(size_t)y * dstride + x] = (uint8_t)((s + 2) >> 2);
}

// generic svt_aom_downsample_2d_c semantics (any step), for the RTCD-signature wrapper
__global__ void k_downsample_generic(const uint8_t *in, int in_stride, int in_w, int in_h, uint8_t *out, int out_stride,
                                     int step) {
    const int half = step >> 1;
    const int ox = blockIdx.x * blockDim.x + threadIdx.x, oy = blockIdx.y;
    const int h = half + ox * step, v = half + oy * step;
    if (h >= in_w || v >= in_h) return;
    const uint8_t *cur = in + (size_t)v * in_stride, *prev = cur - in_stride;
    const uint32_t s = (uint32_t)prev[h - 1] + prev[h] + cur[h - 1] + cur[h];
    out[(size_t)oy * out_stride + ox] = (uint8_t)((s + 2) >> 2);
}

__global__ void k_unpack10(const uint16_t *in, int in_stride, uint8_t *out, int out_stride, int w, int h) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
    if (x < w && y < h) out[(size_t)y * out_stride + x] = (uint8_t)(in[(size_t)y * in_stride + x] >> 2);
}

__global__ void k_copy2d(const uint8_t *in, int in_stride, uint8_t *out, int out_stride, int w, int h) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
    if (x < w && y < h) out[(size_t)y * out_stride + x] = in[(size_t)y * in_stride + x];
}

static inline dim3 grid2d(int w, int h, int bx) { return dim3((unsigned)((w + bx - 1) / bx), (unsigned)h, 1); }

int svtme_launch_build_pyramid(const DevPic &pic, cudaStream_t st) {
    const DevPlane &f = pic.lvl[0], &q = pic.lvl[1], &s = pic.lvl[2];
    k_pad_plane<<<grid2d(f.width + 2 * f.org_x, f.height + 2 * f.org_y, 256), 256, 0, st>>>(
        const_cast<uint8_t *>(f.buf), f.stride, f.org_x, f.width, f.height);
    k_downsample2<<<grid2d(q.width + 2 * q.org_x, q.height + 2 * q.org_y, 256), 256, 0, st>>>(
        f.buf, f.stride, f.org_x, const_cast<uint8_t *>(q.buf), q.stride, q.org_x, q.width, q.height);
    k_downsample2<<<grid2d(s.width + 2 * s.org_x, s.height + 2 * s.org_y, 256), 256, 0, st>>>(
        q.buf, q.stride, q.org_x, const_cast<uint8_t *>(s.buf), s.stride, s.org_x, s.width, s.height);
    return 3;
}

int svtme_launch_unpack10(const uint16_t *d_in, int in_stride, uint8_t *d_out, int out_stride, int w, int h,
                          cudaStream_t st) {
    k_unpack10<<<grid2d(w, h, 256), 256, 0, st>>>(d_in, in_stride, d_out, out_stride, w, h);
    return 1;
}

int svtme_launch_copy2d(const uint8_t *d_in, int in_stride, uint8_t *d_out, int out_stride, int w, int h,
                        cudaStream_t st) {
    k_copy2d<<<grid2d(w, h, 256), 256, 0, st>>>(d_in, in_stride, d_out, out_stride, w, h);
    return 1;
}

int svtme_launch_downsample_generic(const uint8_t *d_in, int in_stride, int in_w, int in_h, uint8_t *d_out,
                                    int out_stride, int step, cudaStream_t st) {
    const int ow = (in_w - (step >> 1) + step - 1) / step, oh = (in_h - (step >> 1) + step - 1) / step;
    if (ow <= 0 || oh <= 0) return 0;
    k_downsample_generic<<<grid2d(ow, oh, 128), 128, 0, st>>>(d_in, in_stride, in_w, in_h, d_out, out_stride, step);
    return 1;
}

// ------------------------------------------------------------------------------------------
// picture pipeline
// ------------------------------------------------------------------------------------------
static bool g_attr_set = false;

void svtme_smem_plan(const SvtMePicParams &pp, int *hme_bytes, int *fp_bytes) {
    const SvtMeCtrls &c = pp.ctrls;
    // HME staged searches: a window row is 8*ceil(sa_w/8) + 16 bytes, rows = sa_h + 2*(rows-1) <= sa_h + 30.
    // Budget: fit the common windows whole, cap at 48 KB (larger windows are tiled in y).
    size_t need = 16 * 1024;
    auto   want = [&](int w, int h) {
        const size_t row = (size_t)(((w + 7) & ~7) + 16);
        const size_t b   = row * (size_t)(h + 31);
        if (b > need) need = b;
    };
    if (c.prehme_enable)
        for (int i = 0; i < 2; i++) want(c.prehme_sa_cfg[i].sa_max.width, c.prehme_sa_cfg[i].sa_max.height);
    want(((c.hme_l0_sa.sa_max.width / 2 + 15) & ~15), c.hme_l0_sa.sa_max.height / 2);
    if (need > 48 * 1024) need = 48 * 1024;
    *hme_bytes = (int)((need + 15) & ~(size_t)15);
    // full-pel: row = 8*ceil(sa_w/8) + 64 bytes, rows = sa_h + 63; widths can double (mv_based_sa_adj) and
    // grow x1.5 (8x8 variance). Cap at 64 KB; bigger windows are tiled in y.
    int mw = c.me_sa.sa_max.width, mh = c.me_sa.sa_max.height;
    if (c.mv_sa_adj_enabled) { mw *= c.mv_sa_adj_sa_multiplier; mh *= c.mv_sa_adj_sa_multiplier; }
    mw = ((mw * 3 / 2) + 7) & ~7;
    mh = mh * 3 / 2;
    size_t fp = (size_t)(mw + 64) * (size_t)(mh + 63);
    const size_t fp_min = (size_t)(mw + 64) * 64; // one search row per tile at least
    if (fp > 64 * 1024) fp = 64 * 1024;
    if (fp < fp_min) fp = fp_min;
    if (fp < 12 * 1024) fp = 12 * 1024;
    *fp_bytes = (int)((fp + 15) & ~(size_t)15);
}

int svtme_launch_picture_me(PicJob &job, cudaStream_t st, cudaEvent_t ev_hme0, cudaEvent_t ev_hme1, cudaEvent_t ev_fp1) {
    if (!g_attr_set) {
        cudaFuncSetAttribute(k_hme, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
        cudaFuncSetAttribute(k_fullpel, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024);
        g_attr_set = true;
    }
    int launches = 0;
    if (ev_hme0) cudaEventRecord(ev_hme0, st);
    k_hme<<<job.n_b64, HME_THREADS, job.hme_smem_bytes, st>>>(job);
    launches++;
    if (ev_hme1) cudaEventRecord(ev_hme1, st);
    const dim3 fgrid(8, (unsigned)job.n_b64, 1);
    if (job.pp.ctrls.enable_me_sr_adjustment == 2) {
        // the window of every ref but (0,0) depends on p_sb_best_sad[0][0][0] (motion_estimation.c:1357-1366)
        job.fullpel_phase = 1;
        k_fullpel<<<fgrid, FP_THREADS, job.fp_smem_bytes, st>>>(job);
        job.fullpel_phase = 2;
        k_fullpel<<<fgrid, FP_THREADS, job.fp_smem_bytes, st>>>(job);
        launches += 2;
    } else {
        job.fullpel_phase = 0;
        k_fullpel<<<fgrid, FP_THREADS, job.fp_smem_bytes, st>>>(job);
        launches++;
    }
    if (ev_fp1) cudaEventRecord(ev_fp1, st);
    k_post<<<job.n_b64, POST_THREADS, 0, st>>>(job);
    k_mask_unsearched<<<job.n_b64, 96, 0, st>>>(job);
    launches += 2;
    return launches;
}

// ------------------------------------------------------------------------------------------
// RTCD-signature kernels (known-answer path)
// ------------------------------------------------------------------------------------------

// svt_sad_loop_kernel semantics for ANY block shape: one thread per search position, byte loads.
// Used by the C-ABI wrapper when the shape is outside what the staged strip search covers.
__global__ void k_kat_sad_loop_generic(const uint8_t *src, int src_stride, const uint8_t *ref, int ref_stride, int bh,
                                       int bw, int src_stride_raw, int skip, int sa_w, int sa_h,
                                       unsigned long long *out_key) {
    unsigned long long best = HME_KEY_INIT;
    const bool sk = skip && bw == 16 && bh <= 16;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < sa_w * sa_h; i += gridDim.x * blockDim.x) {
        const int y = i / sa_w, x = i - y * sa_w;
        if (sk && (y & 1) == 0) continue;
        const uint8_t *r = ref + (size_t)y * src_stride_raw + x;
        uint32_t sad = 0;
        for (int yy = 0; yy < bh; yy++)
            for (int xx = 0; xx < bw; xx++) {
                const int a = src[(size_t)yy * src_stride + xx], b = r[(size_t)yy * ref_stride + xx];
                sad += (uint32_t)(a > b ? a - b : b - a);
            }
        const unsigned long long k = ((unsigned long long)sad << 32) | (uint32_t)i;
        if (sad < 0xffffff && k < best) best = k;
    }
    if (best != HME_KEY_INIT) atomicMin(out_key, best);
}

// svt_sad_loop_kernel through the production staged search (block <= 16 wide) or the direct
// strip search (17..64 wide); one CTA.
__global__ void __launch_bounds__(HME_THREADS) k_kat_sad_loop(const uint8_t *src, int src_stride, const uint8_t *ref,
                                                               int ref_stride, int bh, int bw, int src_stride_raw,
                                                               int skip, int sa_w, int sa_h, int win_bytes,
                                                               unsigned long long *out_key) {
    extern __shared__ __align__(16) uint32_t kat_win[];
    __shared__ __align__(16) HmeSh sh;
    const int tid = threadIdx.x;
    const int strips = (bw + 15) >> 4;
    // the reference advances the ref row by src_stride_raw per search row and by ref_stride per
    // block row; the production searches need ref_stride == row_step * src_stride_raw
    const int row_step = ref_stride / src_stride_raw;
    if (tid == 0) {
        HmeJob &j = sh.jobs[0];
        j.ref = ref; j.ref_stride = src_stride_raw; j.sa_w = sa_w; j.sa_h = sa_h; j.origin_x = j.origin_y = 0;
        sh.job_key[0] = HME_KEY_INIT;
        sh.n_jobs = 1;
        sh.task_prefix[0] = 0;
        sh.task_prefix[1] = sa_h * ((sa_w + 7) >> 3) * (strips == 3 ? 4 : strips);
    }
    uint32_t *srcw = strips == 1 ? sh.src16 : (strips == 2 ? sh.src4 : sh.src1);
    const int words = strips == 1 ? 4 : (strips == 2 ? 8 : 16);
    hme_stage_src(srcw, words, src, src_stride, 1, bh, bw, tid);
    __syncthreads();
    if (strips == 1)
        hme_search_staged(sh, sh.jobs[0], 0, kat_win, win_bytes / 4, sh.src16, bw, bh, row_step, skip && bw == 16 && bh <= 16);
    else
        hme_search_direct(sh, srcw, strips == 2 ? 2 : 4, bw, bh, row_step);
    __syncthreads();
    if (tid == 0) *out_key = sh.job_key[0];
}

int svtme_launch_kat_sad_loop(const uint8_t *d_src, int src_stride, const uint8_t *d_ref, int ref_stride, int bh, int bw,
                              int src_stride_raw, int skip, int sa_w, int sa_h, unsigned long long *d_key,
                              cudaStream_t st) {
    const bool fast = bw >= 1 && bw <= 64 && bh >= 1 && bh <= (bw <= 16 ? 16 : (bw <= 32 ? 32 : 64)) && src_stride_raw > 0 &&
        (ref_stride == src_stride_raw || ref_stride == 2 * src_stride_raw) && sa_w > 0 && sa_h > 0;
    if (fast) {
        const int win_bytes = 48 * 1024;
        if (!g_attr_set) {
            cudaFuncSetAttribute(k_hme, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
            cudaFuncSetAttribute(k_fullpel, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024);
            g_attr_set = true;
        }
        static bool kat_attr = false;
        if (!kat_attr) {
            cudaFuncSetAttribute(k_kat_sad_loop, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
            kat_attr = true;
        }
        k_kat_sad_loop<<<1, HME_THREADS, win_bytes, st>>>(d_src, src_stride, d_ref, ref_stride, bh, bw, src_stride_raw, skip,
                                                         sa_w, sa_h, win_bytes, d_key);
    } else {
        const unsigned long long init = HME_KEY_INIT;
        cudaMemcpyAsync(d_key, &init, sizeof(init), cudaMemcpyHostToDevice, st);
        if (sa_w > 0 && sa_h > 0) {
            const int n = sa_w * sa_h;
            k_kat_sad_loop_generic<<<(n + 127) / 128, 128, 0, st>>>(d_src, src_stride, d_ref, ref_stride, bh, bw,
                                                                      src_stride_raw, skip, sa_w, sa_h, d_key);
        }
    }
    return 1;
}

// svt_ext_all_sad_calculation_8x8_16x16 + svt_ext_eight_sad_calculation_32x32_64x64 for ONE group of
// 8 positions, through the production tile routine. d_ref points at position 0; rows: ref_stride.
// out: sad8[64][8] (z-order), sad16[16][8], sad32[4][8], sad64[8] (un-updated raw SADs, doubled when
// sub_sad) computed with single-position windows so that every position is observable.
__global__ void __launch_bounds__(FP_THREADS) k_kat_octet(const uint8_t *src, int src_stride, const uint8_t *ref,
                                                           int ref_stride, int sub, int n_pos, uint32_t *out /*[8][85]*/) {
    extern __shared__ __align__(16) uint32_t kat_fp_win[];
    __shared__ __align__(16) FpSh sh;
    const int tid = threadIdx.x;
    for (int i = tid; i < 64 * 16; i += FP_THREADS)
        sh.src[i] = ldg_u32_unaligned(src + (size_t)(i >> 4) * src_stride + 4 * (i & 15));
    const int stride = 2 + 16;
    for (int p = 0; p < n_pos; p++) {
        for (int i = tid; i < 85; i += FP_THREADS) sh.gkey[i] = ~0ull;
        __syncthreads();
        stage_rect(kat_fp_win, stride, ref + p, ref_stride, 64, stride, tid, FP_THREADS);
        __syncthreads();
        fp_search_tile(sh, kat_fp_win, stride, 1, 1, sub != 0, sub ? 13 : 12, 0u);
        __syncthreads();
        for (int i = tid; i < 85; i += FP_THREADS) out[p * 85 + i] = (uint32_t)(sh.gkey[i] >> 32) << (sub ? 1 : 0);
        __syncthreads();
    }
}

int svtme_launch_kat_octet(const uint8_t *d_src, int src_stride, const uint8_t *d_ref, int ref_stride, int sub, int n_pos,
                           uint32_t *d_out, cudaStream_t st) {
    k_kat_octet<<<1, FP_THREADS, 64 * 18 * 4, st>>>(d_src, src_stride, d_ref, ref_stride, sub, n_pos, d_out);
    return 1;
}

__global__ void k_kat_nxm(const uint8_t *src, int src_stride, const uint8_t *ref, int ref_stride, int h, int w,
                          uint32_t *out) {
    const uint32_t s = warp_nxm_sad(src, src_stride, ref, ref_stride, h, w);
    if (threadIdx.x == 0) *out = s;
}
int svtme_launch_kat_nxm(const uint8_t *d_src, int src_stride, const uint8_t *d_ref, int ref_stride, int h, int w,
                         uint32_t *d_out, cudaStream_t st) {
    k_kat_nxm<<<1, 32, 0, st>>>(d_src, src_stride, d_ref, ref_stride, h, w, d_out);
    return 1;
}

__global__ void k_fill32(uint32_t *p, uint32_t n, uint32_t v) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = v;
}
int svtme_launch_fill32(uint32_t *d_p, uint32_t n, uint32_t v, cudaStream_t st) {
    if (n) k_fill32<<<(n + 255) / 256, 256, 0, st>>>(d_p, n, v);
    return n ? 1 : 0;
}
