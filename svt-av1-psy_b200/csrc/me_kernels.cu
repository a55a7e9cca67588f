// me_kernels.cu — kernel definitions and their host-side launchers (sm_100a only).
#include <cuda.h>
#include <cstdlib>
#include <cstring>
#include "me_launch.h"

#include "me_device.cuh"
#include "me_search.cuh"
#include "me_ctl.cuh"
#include "me_fp3.cuh"
#include "me_post.cuh"

// ------------------------------------------------------------------------------------------
// pyramid: replicate padding + 2x2 decimation (pic_operators.c:338-384,
// pic_analysis_process.c:131-161, :2139-2196), 10-bit -> 8 MSBs (pack_unpack_c.c:77-116)
// ------------------------------------------------------------------------------------------

// One pass over the input builds all three interiors. A thread owns an 8x4 patch of the full-resolution
// picture: it reads 4 x 8 samples (8-bit, or 16-bit words holding 10-bit samples: 8 MSBs kept,
// pack_unpack_c.c:77-116), writes them into the full plane unless they already are there (host uploads
// land in the plane directly), and writes the 4x2 quarter and 2x1 sixteenth samples they decimate to:
// out = (a + b + c + d + 2) >> 2 per 2x2 box, the sixteenth plane from the ROUNDED quarter samples
// (pic_operators.c:338-384 applied twice, pic_analysis_process.c:131-161).
// MODE 0: 8-bit input in `in`; 1: interior already in place; 2: 16-bit input.
// Up to INGEST_BATCH pictures per launch (blockIdx.z / blockIdx.y = picture): the jobs travel as a kernel parameter.
#define INGEST_BATCH 32
struct IngestJob {
    DevPic      pic;
    const void *in;
    int         in_stride; // samples
    int         pad_;
};
struct IngestBatch {
    IngestJob j[INGEST_BATCH];
};

template <int MODE>
__global__ void __launch_bounds__(256) k_ingest(const __grid_constant__ IngestBatch batch) {
    const IngestJob &job = batch.j[blockIdx.z];
    const DevPic &pic = job.pic;
    const void *in = job.in;
    const int in_stride = job.in_stride;
    const DevPlane &f = pic.lvl[0], &q = pic.lvl[1], &s = pic.lvl[2];
    const int px = blockIdx.x * blockDim.x + threadIdx.x, py = blockIdx.y * blockDim.y + threadIdx.y;
    const int x = 8 * px, y = 4 * py;
    if (x >= f.width || y >= f.height) return;
    uint8_t *fi = const_cast<uint8_t *>(f.buf) + (size_t)(f.org_y + y) * f.stride + f.org_x + x; // 8-byte aligned
    uint2 r[4];
#pragma unroll
    for (int k = 0; k < 4; k++) {
        if (MODE == 1) {
            r[k] = *(const uint2 *)(fi + (size_t)k * f.stride);
        } else if (MODE == 0) {
            const uint8_t *p = (const uint8_t *)in + (size_t)(y + k) * in_stride + x;
            if ((((uintptr_t)p) & 7) == 0) {
                r[k] = __ldg((const uint2 *)p);
            } else {
                r[k].x = ldg_u32_unaligned(p);
                r[k].y = ldg_u32_unaligned(p + 4);
            }
        } else {
            const uint16_t *p = (const uint16_t *)in + (size_t)(y + k) * in_stride + x;
            uint32_t v[8];
#pragma unroll
            for (int i = 0; i < 8; i++) v[i] = (uint32_t)(__ldg(p + i) >> 2) & 0xffu;
            r[k].x = v[0] | (v[1] << 8) | (v[2] << 16) | (v[3] << 24);
            r[k].y = v[4] | (v[5] << 8) | (v[6] << 16) | (v[7] << 24);
        }
        if (MODE != 1) *(uint2 *)(fi + (size_t)k * f.stride) = r[k];
    }
    // quarter: 4 samples per row pair. Bytes of a word pair up as (0,1) and (2,3).
    uint32_t qrow[2];
#pragma unroll
    for (int k = 0; k < 2; k++) {
        const uint2 a = r[2 * k], b = r[2 * k + 1];
        const uint32_t w[2] = {a.x, a.y}, v[2] = {b.x, b.y};
        uint32_t out = 0;
#pragma unroll
        for (int i = 0; i < 4; i++) {
            const uint32_t wa = w[i >> 1] >> (16 * (i & 1)), wb = v[i >> 1] >> (16 * (i & 1));
            const uint32_t sum = (wa & 0xff) + ((wa >> 8) & 0xff) + (wb & 0xff) + ((wb >> 8) & 0xff);
            out |= ((sum + 2) >> 2) << (8 * i);
        }
        qrow[k] = out;
        *(uint32_t *)(const_cast<uint8_t *>(q.buf) + (size_t)(q.org_y + (y >> 1) + k) * q.stride + q.org_x + (x >> 1)) = out;
    }
    uint32_t s16 = 0;
#pragma unroll
    for (int i = 0; i < 2; i++) {
        const uint32_t wa = qrow[0] >> (16 * i), wb = qrow[1] >> (16 * i);
        const uint32_t sum = (wa & 0xff) + ((wa >> 8) & 0xff) + (wb & 0xff) + ((wb >> 8) & 0xff);
        s16 |= ((sum + 2) >> 2) << (8 * i);
    }
    *(uint16_t *)(const_cast<uint8_t *>(s.buf) + (size_t)(s.org_y + (y >> 2)) * s.stride + s.org_x + (x >> 2)) = (uint16_t)s16;
}

// Replicate padding of the three planes in one launch: every padded sample is the interior sample at the
// clamped coordinate. A thread writes one aligned 4-byte group; only groups that hold padding are enumerated
// (top band, bottom band, then the left and right bands of the interior rows), level after level.
__global__ void __launch_bounds__(256) k_pad_planes(const __grid_constant__ IngestBatch batch) {
    const DevPic &pic = batch.j[blockIdx.y].pic;
    int i = blockIdx.x * blockDim.x + threadIdx.x;
#pragma unroll
    for (int lv = 0; lv < 3; lv++) {
        const DevPlane &p = pic.lvl[lv];
        const int pad = p.org_x, w = p.width, h = p.height;
        const int W4 = (w + 2 * pad + 3) >> 2;
        const int lg = pad >> 2;          // groups wholly left of the interior (pads are multiples of 4)
        const int rg0 = (pad + w) >> 2;   // first group that touches the right padding
        const int side = lg + (W4 - rg0);
        const int n_band = pad * W4, n_mid = h * side;
        if (i >= 2 * n_band + n_mid) { i -= 2 * n_band + n_mid; continue; }
        int y, g;
        if (i < 2 * n_band) {
            const int r = i / W4;
            g = i - r * W4;
            y = r < pad ? r : h + r; // r in [pad, 2 pad) -> rows pad + h ..
        } else {
            const int k = i - 2 * n_band, r = k / side;
            g = k - r * side;
            if (g >= lg) g = rg0 + (g - lg);
            y = pad + r;
        }
        uint8_t *row = const_cast<uint8_t *>(p.buf) + (size_t)y * p.stride;
        const uint8_t *srow = p.buf + (size_t)(min(max(y - pad, 0), h - 1) + pad) * p.stride + pad;
        const int x0 = 4 * g - pad; // interior coordinate of the group's first byte
        uint32_t v;
        if (x0 >= 0 && x0 + 3 < w) {
            v = *(const uint32_t *)(srow + x0);
        } else {
            v = 0;
#pragma unroll
            for (int k = 0; k < 4; k++) v |= (uint32_t)srow[min(max(x0 + k, 0), w - 1)] << (8 * k);
        }
        *(uint32_t *)(row + 4 * g) = v;
        return;
    }
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

static inline dim3 grid2d(int w, int h, int bx) { return dim3((unsigned)((w + bx - 1) / bx), (unsigned)h, 1); }

// mode 0: 8-bit device input, 1: the full-resolution interior is already in the plane, 2: 16-bit device input.
// All pictures of a call share the mode and the geometry (one session).
int svtme_launch_build_pyramids(const DevPic *pics, const void *const *d_in, int in_stride, int n, int mode, cudaStream_t st) {
    int launches = 0;
    for (int base = 0; base < n; base += INGEST_BATCH) {
        const int m = n - base < INGEST_BATCH ? n - base : INGEST_BATCH;
        IngestBatch b;
        memset(&b, 0, sizeof(b));
        for (int i = 0; i < m; i++) {
            b.j[i].pic = pics[base + i];
            b.j[i].in = d_in ? d_in[base + i] : nullptr;
            b.j[i].in_stride = in_stride;
        }
        const DevPlane &f = pics[base].lvl[0];
        const dim3 blk(32, 8), grid((unsigned)((f.width / 8 + 31) / 32), (unsigned)((f.height / 4 + 7) / 8), (unsigned)m);
        if (mode == 0) k_ingest<0><<<grid, blk, 0, st>>>(b);
        else if (mode == 1) k_ingest<1><<<grid, blk, 0, st>>>(b);
        else k_ingest<2><<<grid, blk, 0, st>>>(b);
        long long groups = 0;
        for (int lv = 0; lv < 3; lv++) {
            const DevPlane &p = pics[base].lvl[lv];
            const int W4 = (p.width + 2 * p.org_x + 3) >> 2;
            groups += 2LL * p.org_x * W4 + (long long)p.height * ((p.org_x >> 2) + W4 - ((p.org_x + p.width) >> 2));
        }
        k_pad_planes<<<dim3((unsigned)((groups + 255) / 256), (unsigned)m), 256, 0, st>>>(b);
        launches += 2;
    }
    return launches;
}

int svtme_launch_build_pyramid(const DevPic &pic, int mode, const void *d_in, int in_stride, cudaStream_t st) {
    return svtme_launch_build_pyramids(&pic, &d_in, in_stride, 1, mode, st);
}

int svtme_launch_downsample_generic(const uint8_t *d_in, int in_stride, int in_w, int in_h, uint8_t *d_out,
                                    int out_stride, int step, cudaStream_t st) {
    const int ow = (in_w - (step >> 1) + step - 1) / step, oh = (in_h - (step >> 1) + step - 1) / step;
    if (ow <= 0 || oh <= 0) return 0;
    k_downsample_generic<<<grid2d(ow, oh, 128), 128, 0, st>>>(d_in, in_stride, in_w, in_h, d_out, out_stride, step);
    return 1;
}

// ------------------------------------------------------------------------------------------
// picture pipeline (v2): flat warp-per-tile searches, control in k_hme_ctl between them
// ------------------------------------------------------------------------------------------
// Inside a launch sequence nothing goes through a copy engine: cudaMemsetAsync / cudaMemcpyAsync of a few bytes on the
// compute stream share the engines' in-order queues with the session's large host uploads, so a counter reset that waits
// (in stream order) for the previous kernel holds back the next mini-GOP's upload behind it, and the uploads stop
// overlapping the kernels (measured at N = 8, where an upload takes 3 ms: 6.1 ms per step instead of 3.9). Counters are
// zeroed and the job array is fetched (from pinned, device-visible host memory) by small kernels instead.
__global__ void k_zero_words(uint32_t *p, int n) {
    if ((int)threadIdx.x < n) p[threadIdx.x] = 0u;
}
__global__ void __launch_bounds__(256) k_copy_words(uint4 *dst, const uint4 *src, int n16) {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += gridDim.x * blockDim.x) dst[i] = src[i];
}
void svtme_launch_zero_words(void *d_p, int n_words, cudaStream_t st) { k_zero_words<<<1, 32, 0, st>>>((uint32_t *)d_p, n_words); }
void svtme_launch_copy_from_pinned(void *d_dst, const void *h_src, size_t bytes, cudaStream_t st) {
    const int n16 = (int)((bytes + 15) / 16); // both buffers are allocated in multiples of 16 bytes
    k_copy_words<<<(n16 + 255) / 256 < 64 ? (n16 + 255) / 256 : 64, 256, 0, st>>>((uint4 *)d_dst, (const uint4 *)h_src, n16);
}

static int g_fp_search_ctas_per_sm = 0, g_s16_ctas_per_sm = 0;
// experiment knob (SVTME_PERSISTENT_SLACK=k): the persistent search kernels leave k CTA slots per SM free, so that with
// two compute streams (svtme_session_set_concurrency) the control / post kernels of one batch can run under the searches
// of the other. 0 = fill the SM (default).
static int persistent_slack() {
    static int v = -1;
    if (v < 0) {
        const char *e = getenv("SVTME_PERSISTENT_SLACK");
        v = e ? atoi(e) : 0;
        if (v < 0) v = 0;
    }
    return v;
}

// kernels whose dynamic shared memory exceeds the 48 KB default opt in once per device
static void opt_in_shared_memory() {
    static int done_for = -1;
    int dev = 0;
    cudaGetDevice(&dev);
    if (done_for == dev) return;
    const int blk2 = SEARCH_WARPS * 2 * TM_L2_BUF + 128, blk1 = SEARCH_WARPS * 2 * TM_L1_BUF + 128, fp = FP_WARPS * 2 * TM_FP_BUF + 128;
    cudaFuncSetAttribute(k_search16<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, SEARCH_WARPS * 2 * TM_S16_BUF + 128);
    cudaFuncSetAttribute(k_search16<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, SEARCH_WARPS * 2 * TM_S16_BUF + 128);
    cudaFuncSetAttribute(k_search_blk<true, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, blk1);
    cudaFuncSetAttribute(k_search_blk<false, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, blk1);
    cudaFuncSetAttribute(k_search_blk<true, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, blk2);
    cudaFuncSetAttribute(k_search_blk<false, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, blk2);
    cudaFuncSetAttribute(k_fp<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, fp);
    cudaFuncSetAttribute(k_fp<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, fp);
    cudaFuncSetAttribute(k_fp_big<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, fp);
    cudaFuncSetAttribute(k_fp_big<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, fp);
    done_for = dev;
}

static inline int imax_h(int a, int b) { return a > b ? a : b; }
static int next_pow2(int v) {
    int p = 1;
    while (p < v) p <<= 1;
    return p;
}

// ------------------------------------------------------------------------------------------
// dynamic-GOP detector HME (dg_detector_hme_level0 + early_hme_b64, pd_process.c:392-589)
// ------------------------------------------------------------------------------------------
struct DgOut { // device mirror of SvtMeDgMetrics' outputs
    unsigned long long tot_dist;
    unsigned int       tot_cplx, tot_active;
    int                sum_in_vectors;
};

// one lane per b64: window of its 16x16 block (:410-455). The window is cut into bands of one tile height, one job and one
// queue unit per band (slots 4k of the b64's eight (list, ref) groups, which this search does not use otherwise): a pair
// of pictures has only n_b64 windows, and a unit is searched by one warp, so the bands are what keeps every SM busy.
__global__ void __launch_bounds__(128) k_dg_build(const PicJob *__restrict__ jobs, int sa0) {
    const PicJob &job = jobs[0];
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= job.n_b64) return;
    const B64Geo    g   = b64_geo(job, b);
    const DevPlane &ref = job.ref[0][0].lvl[2];
    HmeJobG *jb = job.hme_jobs + (size_t)b * HME_JOBS_PER_B64;
    HmeJobG J;
    int16_t sa_w = (int16_t)((sa0 + 7) & ~0x07), sa_h = (int16_t)sa0;
    const int16_t ox = (int16_t)(-(sa_w >> 1)), oy = (int16_t)(-(sa_h >> 1));
    ctl_make_job(J, ref, (int16_t)(g.org_x >> 2), (int16_t)(g.org_y >> 2), ox, oy, sa_w, sa_h, (int16_t)(ref.org_x - 1),
                 (int16_t)(ref.org_y - 1), true, 16, 16);
    const int band_h = J.tile_h, n_bands = J.sa_w > 0 && J.sa_h > 0 ? imin(8, (J.sa_h + band_h - 1) / band_h) : 0;
    for (int k = 0; k < HME_JOBS_PER_B64; k++) jb[k].active = 0;
    if (!n_bands) { // an empty window (cannot happen with the detector's sizes): the record k_dg_reduce falls back to
        jb[0] = J;
        jb[0].active = 0;
    }
    uint32_t at = n_bands ? atomicAdd(&job.s16_counters[0], (uint32_t)n_bands) : 0u;
    for (int k = 0; k < n_bands; k++) {
        HmeJobG Jk = J;
        Jk.py = (int16_t)(J.py + k * band_h);
        Jk.origin_y = (int16_t)(J.origin_y + k * band_h);
        Jk.sa_h = (int16_t)(k + 1 == n_bands ? J.sa_h - k * band_h : band_h); // (8 bands of 18 rows hold the largest window: 128 rows)
        Jk.n_ty = (uint8_t)((Jk.sa_h + band_h - 1) / band_h);
        jb[4 * k] = Jk;
        if (at + k < job.s16_units_cap) job.s16_units[at + k] = make_uint2(0u, (uint32_t)(b * HME_JOBS_PER_B64 + 4 * k));
    }
}

// per-b64 result (:470-486) and the metrics (:540-575)
__global__ void __launch_bounds__(128) k_dg_reduce(const PicJob *__restrict__ jobs, DgOut *out, unsigned long long *b64_sad,
                                                   int16_t *b64_mv) {
    const PicJob &job = jobs[0];
    const int b = blockIdx.x * blockDim.x + threadIdx.x;
    if (b >= job.n_b64) return;
    const HmeJobG *jb = job.hme_jobs + (size_t)b * HME_JOBS_PER_B64;
    // first minimum in raster order over the bands: bands are consecutive row ranges, so a tie goes to the earlier band
    uint64_t sad = 0xffffff;
    int16_t  col = 0, row = 0;
    bool     found = false;
    for (int k = 0; k < 8; k++) {
        const HmeJobG &j = jb[4 * k];
        if (!j.active || (uint32_t)j.key == 0xffffffffu) continue;
        uint64_t s = 0;
        int16_t  c = 0, r = 0;
        ctl_job_result(j, false, 4, s, c, r);
        if (!found || s < sad) { sad = s; col = c; row = r; found = true; }
    }
    if (!found) { // no position was evaluated: the reference's initial values (pd_process.c:470-486 with svt_sad_loop_kernel's 0xffffff)
        ctl_job_result(jb[0], false, 4, sad, col, row);
    }
    b64_sad[b] = sad;
    b64_mv[2 * b] = col; b64_mv[2 * b + 1] = row;
    const int bx = b % job.b64_cols, by = b / job.b64_cols, cols = job.b64_cols, rows = job.b64_rows;
    int in = 0;
    if (by < rows / 2) in += row > 0 ? -1 : (row < 0 ? 1 : 0);
    else if (by > rows / 2) in += row > 0 ? 1 : (row < 0 ? -1 : 0);
    if (bx < cols / 2) in += col > 0 ? -1 : (col < 0 ? 1 : 0);
    else if (bx > cols / 2) in += col > 0 ? 1 : (col < 0 ? -1 : 0);
    atomicAdd(&out->tot_dist, (unsigned long long)sad);
    if (sad > 16 * 16 * 30) atomicAdd(&out->tot_cplx, 1u);
    if (col != 0 || row != 0) atomicAdd(&out->tot_active, 1u);
    if (in) atomicAdd(&out->sum_in_vectors, in);
}

uint32_t svtme_dg_units_cap(int sa0, int n_b64) {
    (void)sa0;
    return (uint32_t)(8 * n_b64); // at most eight bands per b64
}

int svtme_launch_dg_detector(const PicJob *h_job, const PicJob *d_job, int sa0, void *d_out, unsigned long long *d_b64_sad,
                             int16_t *d_b64_mv, cudaStream_t st) {
    opt_in_shared_memory();
    const int n_b64 = h_job->n_b64;
    svtme_launch_zero_words(h_job->s16_counters, 2, st);
    svtme_launch_zero_words(d_out, (int)(sizeof(DgOut) / 4), st);
    k_dg_build<<<(n_b64 + 127) / 128, 128, 0, st>>>(d_job, sa0);
    int n_sm = 148, dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev);
    const size_t s16_smem = (size_t)SEARCH_WARPS * 2 * TM_S16_BUF + 128;
    if (!g_s16_ctas_per_sm) {
        int k = 0;
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&k, k_search16<true>, 32 * SEARCH_WARPS, s16_smem);
        g_s16_ctas_per_sm = k > 0 ? k : 1;
    }
    k_search16<true><<<n_sm * g_s16_ctas_per_sm, 32 * SEARCH_WARPS, s16_smem, st>>>(d_job, 0, h_job->pp.ctrls.hme_search_method == SVTME_SUB_SAD_SEARCH ? 2 : 1);
    k_dg_reduce<<<(n_b64 + 127) / 128, 128, 0, st>>>(d_job, (DgOut *)d_out, d_b64_sad, d_b64_mv);
    return 3;
}

uint32_t svtme_s16_units_cap(const SvtMePicParams &pp, int n_b64) {
    (void)pp;
    return (uint32_t)(32 * n_b64); // one unit per (b64, list, ref) and stage, or one per job slot in small batches
}

uint32_t svtme_fp_units_cap(const SvtMePicParams &pp, int n_b64) {
    const SvtMeCtrls &c = pp.ctrls;
    long long mw = c.me_sa.sa_max.width, mh = c.me_sa.sa_max.height;
    if (c.mv_sa_adj_enabled) { mw *= c.mv_sa_adj_sa_multiplier; mh *= c.mv_sa_adj_sa_multiplier; }
    mw = ((mw * 3 / 2) + 7) & ~7LL; // the 8x8-variance rule can grow the window by 3/2 (:1426-1429)
    mh = mh * 3 / 2;
    if (mw < 8) mw = 8;
    if (mh < 3) mh = 3;
    long long chunks = ((mw + TM_FP_TILE_W - 1) / TM_FP_TILE_W) * ((mh + TM_FP_TILE_H - 1) / TM_FP_TILE_H);
    if (chunks > 65535) chunks = 65535;
    long long cap = chunks * n_b64 * 8;
    if (cap > 0x7fffffffLL) cap = 0x7fffffffLL;
    return (uint32_t)cap;
}

// Launch the whole ME stage for a batch of n pictures (h_jobs: host copy, d_jobs: the same array on
// the device). Every kernel is launched once for the batch (grid.y = picture); the control sequence
// is the union of what the pictures need (an extra round is a no-op for pictures without the
// dependency it serves).
int svtme_launch_picture_me(const PicJob *h_jobs, int n, const PicJob *d_jobs, cudaStream_t st, cudaEvent_t ev_hme0,
                            cudaEvent_t ev_hme1, cudaEvent_t ev_fp1, SvtMeTimeline *tl, const cudaEvent_t *out_free, int n_out_free) {
    // per-kernel timing (svtme_session_enable_timing): an event behind every launch, tagged with the kernel it closes
    auto mark = [&](int tag) {
        if (!tl || tl->n >= SVTME_TL_MAX) return;
        cudaEventRecord(tl->ev[tl->n], st);
        tl->tag[tl->n++] = tag;
    };
    const PicJob         &j0 = h_jobs[0];
    const SvtMeCtrls     &c  = j0.pp.ctrls; // HME flags / search methods are equal across a batch (validated by the caller)
    int launches = 0;
    const bool sub = c.hme_search_method == SVTME_SUB_SAD_SEARCH;
    const uint32_t ALL = 0xffu, LIST0 = 0x0fu, LIST1 = 0xf0u;
    const int n_b64 = j0.n_b64;
    const unsigned ny = (unsigned)n;
    const int ctl_grid = (n_b64 + CTL_WARPS - 1) / CTL_WARPS;
    bool any_zz = false, any_l1dep = false, any_dep = false, any_sr2 = false;
    int  max_act = 1; // most (list, ref) combinations of a picture of the batch: sizes the grids of the per-(b64, list, ref) kernels
    for (int i = 0; i < n; i++) {
        const SvtMePicParams &pp = h_jobs[i].pp;
        const SvtMeCtrls     &ci = pp.ctrls;
        any_zz |= ci.me_early_exit_th || ci.me_safe_limit_zz_th;
        any_l1dep |= ci.prehme_enable && ci.prehme_l1_early_exit && pp.num_of_list_to_search == 2 && pp.temporal_layer_index > 0;
        any_dep |= ci.enable_me_sr_adjustment && ci.distance_based_hme_resizing && ci.reduce_hme_l0_sr_th_min && ci.reduce_hme_l0_sr_th_max;
        any_sr2 |= ci.enable_me_sr_adjustment == 2;
        max_act = imax_h(max_act, pp.num_of_ref_pic_to_search[0] + (pp.num_of_list_to_search > 1 ? pp.num_of_ref_pic_to_search[1] : 0));
    }
    opt_in_shared_memory();
    if (ev_hme0) cudaEventRecord(ev_hme0, st);
    svtme_launch_zero_words(j0.fp_counters, 2, st);
    mark(SVTME_TL_START);
    if (any_zz) {
        k_zz<<<dim3((n_b64 * max_act + SEARCH_WARPS - 1) / SEARCH_WARPS, ny), 32 * SEARCH_WARPS, 0, st>>>(d_jobs);
        launches++;
        mark(SVTME_TL_ZZ);
    }
    bool first = true;
    int  post = ST_NONE;
    uint32_t pmask = 0;
    auto emit = [&](int fin, int build, uint32_t bmask) {
        CtlOp op;
        op.init = first ? 1 : 0; op.post_stage = post; op.post_mask = pmask; op.finalize = fin; op.build_stage = build; op.build_mask = bmask;
        if (build == ST_PREHME || build == ST_L0) svtme_launch_zero_words(j0.s16_counters, 2, st);
        k_hme_ctl<<<dim3(ctl_grid, ny), 32 * CTL_WARPS, 0, st>>>(d_jobs, op);
        launches++;
        mark(SVTME_TL_CTL);
        first = false;
        post  = ST_NONE;
    };
    int n_sm = 148;
    {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev);
    }
    const size_t s16_smem = (size_t)SEARCH_WARPS * 2 * TM_S16_BUF + 128;
    if (!g_s16_ctas_per_sm) {
        int k = 0;
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&k, k_search16<true>, 32 * SEARCH_WARPS, s16_smem);
        g_s16_ctas_per_sm = k > 0 ? k : 1;
        if (g_s16_ctas_per_sm > persistent_slack()) g_s16_ctas_per_sm -= persistent_slack();
    }
    auto search16 = [&](int skip, int tag) {
        if (tag == SVTME_TL_S16_PREHME) k_search16<false><<<n_sm * g_s16_ctas_per_sm, 32 * SEARCH_WARPS, s16_smem, st>>>(d_jobs, skip, sub ? 2 : 1);
        else k_search16<true><<<n_sm * g_s16_ctas_per_sm, 32 * SEARCH_WARPS, s16_smem, st>>>(d_jobs, skip, sub ? 2 : 1);
        launches++;
        mark(tag);
    };
    auto search_wide = [&](int level) {
        const dim3 grid((n_b64 * max_act + SEARCH_WARPS - 1) / SEARCH_WARPS, ny);
        const size_t smem = (size_t)SEARCH_WARPS * 2 * (level == 1 ? TM_L1_BUF : TM_L2_BUF) + 128;
        if (level == 1) {
            if (sub) k_search_blk<true, 1><<<grid, 32 * SEARCH_WARPS, smem, st>>>(d_jobs);
            else k_search_blk<false, 1><<<grid, 32 * SEARCH_WARPS, smem, st>>>(d_jobs);
        } else {
            if (sub) k_search_blk<true, 2><<<grid, 32 * SEARCH_WARPS, smem, st>>>(d_jobs);
            else k_search_blk<false, 2><<<grid, 32 * SEARCH_WARPS, smem, st>>>(d_jobs);
        }
        launches++;
        mark(level == 1 ? SVTME_TL_BLK_L1 : SVTME_TL_BLK_L2);
    };
    int fin_pending = ST_NONE;
    if (c.prehme_enable) {
        if (any_l1dep) {
            emit(ST_NONE, ST_PREHME, LIST0);
            search16(c.prehme_skip_search_line, SVTME_TL_S16_PREHME);
            post = ST_PREHME; pmask = LIST0;
            emit(ST_NONE, ST_PREHME, LIST1);
            search16(c.prehme_skip_search_line, SVTME_TL_S16_PREHME);
            post = ST_PREHME; pmask = LIST1;
        } else {
            emit(ST_NONE, ST_PREHME, ALL);
            search16(c.prehme_skip_search_line, SVTME_TL_S16_PREHME);
            post = ST_PREHME; pmask = ALL;
        }
        fin_pending = ST_PREHME;
    }
    if (c.enable_hme_flag) {
        if (c.enable_hme_level0_flag) {
            if (any_dep) {
                emit(fin_pending, ST_L0, 0x01u);
                search16(0, SVTME_TL_S16_L0);
                post = ST_L0; pmask = 0x01u;
                emit(ST_NONE, ST_L0, ALL & ~0x01u);
                search16(0, SVTME_TL_S16_L0);
                post = ST_L0; pmask = ALL & ~0x01u;
            } else {
                emit(fin_pending, ST_L0, ALL);
                search16(0, SVTME_TL_S16_L0);
                post = ST_L0; pmask = ALL;
            }
            fin_pending = ST_NONE;
        }
        if (c.enable_hme_level1_flag) {
            emit(fin_pending, ST_L1, ALL);
            fin_pending = ST_NONE;
            search_wide(1);
            post = ST_L1; pmask = ALL;
        }
        if (c.enable_hme_level2_flag) {
            emit(fin_pending, ST_L2, ALL);
            fin_pending = ST_NONE;
            search_wide(2);
            post = ST_L2; pmask = ALL;
        }
    }
    if (fin_pending == ST_PREHME) emit(ST_PREHME, ST_NONE, 0);
    emit(ST_L2, ST_NONE, 0); // posts the last searches, then set_final_seach_centre_sb + pruning
    if (ev_hme1) cudaEventRecord(ev_hme1, st);

    // ---- full-pel ----
    // k_fp and k_post are the first kernels that write the output arrays of the jobs (the batch's result region): only
    // here does the sequence wait for the readers of the region's previous contents (the download of the batch two
    // before this one, a device-side consumer), not at its start. The HME stage, two thirds of the sequence, then runs
    // under that download instead of behind it.
    for (int i = 0; i < n_out_free; i++) cudaStreamWaitEvent(st, out_free[i], 0);
    const bool fsub = c.me_search_method == SVTME_SUB_SAD_SEARCH;
    const size_t fp_smem = (size_t)FP_WARPS * 2 * TM_FP_BUF + 128;
    if (!g_fp_search_ctas_per_sm) {
        int k = 0;
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&k, k_fp_big<true>, 32 * FP_WARPS, fp_smem);
        g_fp_search_ctas_per_sm = k > 0 ? k : 1;
        if (g_fp_search_ctas_per_sm > persistent_slack()) g_fp_search_ctas_per_sm -= persistent_slack();
    }
    const dim3 plan_grid((n_b64 * max_act + FP_WARPS - 1) / FP_WARPS, ny);
    const int big_grid = n_sm * g_fp_search_ctas_per_sm;
    auto fp_round = [&](int phase) {
        if (fsub) k_fp<true><<<plan_grid, 32 * FP_WARPS, fp_smem, st>>>(d_jobs, phase);
        else k_fp<false><<<plan_grid, 32 * FP_WARPS, fp_smem, st>>>(d_jobs, phase);
        mark(SVTME_TL_FP);
        if (fsub) k_fp_big<true><<<big_grid, 32 * FP_WARPS, fp_smem, st>>>(d_jobs);
        else k_fp_big<false><<<big_grid, 32 * FP_WARPS, fp_smem, st>>>(d_jobs);
        mark(SVTME_TL_FP_BIG);
        launches += 2;
    };
    if (any_sr2) {
        // the window of every ref but (0,0) depends on p_sb_best_sad[0][0][0] (motion_estimation.c:1357-1366)
        fp_round(1);
        svtme_launch_zero_words(j0.fp_counters, 2, st);
        fp_round(2);
    } else {
        fp_round(0);
    }
    if (ev_fp1) cudaEventRecord(ev_fp1, st);
    k_post<<<dim3(n_b64, ny), POST_THREADS, 0, st>>>(d_jobs);
    mark(SVTME_TL_POST);
    launches += 1;
    return launches;
}

// ------------------------------------------------------------------------------------------
// RTCD-signature kernels (known-answer path): thin wrappers around the same warp-level routines
// ------------------------------------------------------------------------------------------

// svt_sad_loop_kernel semantics for ANY block shape: one thread per search position, byte loads.
// Only used for blocks wider than 64 pixels, which the encoder never passes.
__global__ void k_kat_sad_loop_generic(const uint8_t *src, int src_stride, const uint8_t *ref, int ref_stride, int bh,
                                       int bw, int src_stride_raw, int skip, int sa_w, int sa_h,
                                       unsigned long long *out_key) {
    unsigned long long best = HME_KEY_NONE;
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
    if (best != HME_KEY_NONE) atomicMin(out_key, best);
}

// blocks up to 16x16: the TMA tile search of k_search16, one warp per tile. The window was re-pitched by the caller
// (row pitch multiple of 16, `map` describes it with the tall or the wide box); position (0,0) is its sample (0,0).
__global__ void __launch_bounds__(32 * SEARCH_WARPS) k_kat_search16(const __grid_constant__ CUtensorMap map, const uint8_t *src,
                                                                     int src_stride, int row_step, int bh, int bw, int skip,
                                                                     int sa_w, int sa_h, int tile_w, int tile_h, int n_tx,
                                                                     int n_ty, int wide, unsigned long long *out_key) {
    __shared__ __align__(128) uint8_t s_buf[SEARCH_WARPS][TM_S16_BUF];
    __shared__ __align__(16) uint32_t s_src[SEARCH_WARPS][16 * 4];
    __shared__ __align__(8) uint64_t s_bar[SEARCH_WARPS][2];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int tile = blockIdx.x * SEARCH_WARPS + warp;
    if (tile >= n_tx * n_ty) return;
    TilePipe P;
    pipe_init(P, s_buf[warp], 0, s_bar[warp], 0);
    const int ty = tile / n_tx, tx = tile - ty * n_tx;
    const int x0 = tx * tile_w, y0 = ty * tile_h;
    pipe_issue_bytes(P, 0, &map, x0 & ~15, y0, wide ? TM_S16W_BOX_W * TM_S16W_BOX_H : TM_S16T_BOX_W * TM_S16T_BOX_H);
    warp_stage_src16(smem_u32(s_src[warp]), src, src_stride, 1, bh, bw);
    pipe_wait(P, 0);
    __syncwarp();
    uint32_t best[4];
    const S16Split sp = {32767, 32767};
    warp_search16_tile<false>(P.buf[0], wide ? TM_S16W_BOX_W : TM_S16T_BOX_W, smem_u32(s_src[warp]), (uint32_t)x0 & 15u, x0, y0,
                       imin(tile_w, sa_w - x0), imin(tile_h, sa_h - y0), bw, bh, row_step, skip && bw == 16 && bh <= 16, sp, best);
    const unsigned long long key = best[0] == 0xffffffffu ? HME_KEY_NONE : s16_key64(best[0], x0, y0, 0, sp, sa_w);
    if (lane == 0 && key != HME_KEY_NONE) atomicMin(out_key, key);
}

// blocks 17..64 wide or taller than 16 rows: the TMA tile search of k_search_blk, one warp. The reference window was
// re-pitched by the caller (row pitch multiple of 16, `map` describes it); position (0,0) is its sample (0,0).
template <int NS, int NR>
__global__ void __launch_bounds__(32) k_kat_search_blk(const __grid_constant__ CUtensorMap map, const uint8_t *src, int src_stride,
                                                       int bh, int bw, int row_step, int sa_w, int sa_h,
                                                       unsigned long long *out_key) {
    constexpr int PITCH = NS == 2 ? TM_L1_BOX_W : TM_L2_BOX_W;
    constexpr int BOX_BYTES = NS == 2 ? TM_L1_BOX_W * TM_L1_BOX_H : TM_L2_BOX_W * TM_L2_BOX_H;
    constexpr int BUF = NS == 2 ? TM_L1_BUF : TM_L2_BUF;
    __shared__ __align__(128) uint8_t s_buf[2 * BUF];
    __shared__ __align__(8) uint64_t s_bar[2];
    TilePipe P;
    pipe_init(P, s_buf, BUF, s_bar, BOX_BYTES);
    TLane<NS, NR> L;
    int row_of[NR];
    blk_rows<NS, NR>(row_of, row_step);
    const int ntx = (sa_w + TM_BLK_TILE_W - 1) / TM_BLK_TILE_W, nt = ntx * ((sa_h + TM_BLK_TILE_H - 1) / TM_BLK_TILE_H);
    pipe_issue(P, 0, &map, 0, 0);
    tlane_init<NS, NR>(L, src, src_stride, row_of, bw, bh * row_step, PITCH);
    uint32_t best = 0xffffffffu;
    int stage = 0;
    for (int t = 0; t < nt; t++) {
        const int ty = t / ntx, tx = t - ty * ntx;
        if (t + 1 < nt) {
            const int ty1 = (t + 1) / ntx, tx1 = (t + 1) - ty1 * ntx;
            pipe_issue(P, stage ^ 1, &map, TM_BLK_TILE_W * tx1, TM_BLK_TILE_H * ty1);
        }
        pipe_wait(P, stage);
        blk_tile_eval<NS, NR>(L, P.buf[stage], PITCH, 0, tx, ty, sa_w, sa_h, best);
        __syncwarp();
        stage ^= 1;
    }
    const unsigned long long key = blk_key64(best);
    if (threadIdx.x == 0) *out_key = key;
}

int svtme_launch_kat_sad_loop(const uint8_t *d_src, int src_stride, const uint8_t *d_ref, int ref_stride, int bh, int bw,
                              int src_stride_raw, int skip, int sa_w, int sa_h, unsigned long long *d_key,
                              uint8_t *d_pitched, size_t pitched_cap, cudaStream_t st) {
    const unsigned long long init = HME_KEY_NONE;
    cudaMemcpyAsync(d_key, &init, sizeof(init), cudaMemcpyHostToDevice, st);
    if (sa_w <= 0 || sa_h <= 0) return 0;
    const bool strided = src_stride_raw > 0 && (ref_stride == src_stride_raw || ref_stride == 2 * src_stride_raw);
    if (strided && bw <= 16 && bh <= 16 && d_pitched && (bh - 1) * (ref_stride / src_stride_raw) <= 15) {
        const int row_step = ref_stride / src_stride_raw; // the production shapes: 16 rows, or 8 rows two apart
        const int w_bytes = sa_w - 1 + bw, rows = sa_h - 1 + (bh - 1) * row_step + 1;
        const int pitch = (w_bytes + 15 + 16) & ~15;
        if ((size_t)pitch * rows > pitched_cap) return 0;
        cudaMemsetAsync(d_pitched, 0, (size_t)pitch * rows, st);
        cudaMemcpy2DAsync(d_pitched, pitch, d_ref, src_stride_raw, w_bytes, rows, cudaMemcpyDeviceToDevice, st);
        int tw, th, ntx, nty, wide;
        s16_tile_geometry(sa_w, sa_h, &tw, &th, &ntx, &nty, &wide);
        CUtensorMap map;
        if (svtme_encode_tmap(&map, d_pitched, pitch, rows, wide ? TM_S16W_BOX_W : TM_S16T_BOX_W, wide ? TM_S16W_BOX_H : TM_S16T_BOX_H)) return 0;
        k_kat_search16<<<(ntx * nty + SEARCH_WARPS - 1) / SEARCH_WARPS, 32 * SEARCH_WARPS, 0, st>>>(
            map, d_src, src_stride, row_step, bh, bw, skip, sa_w, sa_h, tw, th, ntx, nty, wide, d_key);
    } else if (strided && bw <= 64 && bh * (ref_stride / src_stride_raw) <= 64 && d_pitched && sa_w * sa_h <= 4096) {
        // re-pitch the window (rows 16-byte aligned, as the picture store keeps its planes) and describe it to the TMA
        const int row_step = ref_stride / src_stride_raw;
        const int w_bytes = sa_w - 1 + bw, rows = sa_h - 1 + (bh - 1) * row_step + 1;
        const int pitch = (w_bytes + 15 + 16) & ~15;
        if ((size_t)pitch * rows > pitched_cap) return 0;
        cudaMemsetAsync(d_pitched, 0, (size_t)pitch * rows, st);
        cudaMemcpy2DAsync(d_pitched, pitch, d_ref, src_stride_raw, w_bytes, rows, cudaMemcpyDeviceToDevice, st);
        const bool small = bw <= 32 && bh * row_step <= 32;
        CUtensorMap map;
        if (svtme_encode_tmap(&map, d_pitched, pitch, rows, small ? TM_L1_BOX_W : TM_L2_BOX_W, small ? TM_L1_BOX_H : TM_L2_BOX_H)) return 0;
        if (small) {
            if (bh <= 16) k_kat_search_blk<2, 1><<<1, 32, 0, st>>>(map, d_src, src_stride / row_step, bh, bw, row_step, sa_w, sa_h, d_key);
            else k_kat_search_blk<2, 2><<<1, 32, 0, st>>>(map, d_src, src_stride / row_step, bh, bw, row_step, sa_w, sa_h, d_key);
        } else {
            if (bh <= 32) k_kat_search_blk<4, 4><<<1, 32, 0, st>>>(map, d_src, src_stride / row_step, bh, bw, row_step, sa_w, sa_h, d_key);
            else k_kat_search_blk<4, 8><<<1, 32, 0, st>>>(map, d_src, src_stride / row_step, bh, bw, row_step, sa_w, sa_h, d_key);
        }
    } else {
        const int n = sa_w * sa_h;
        k_kat_sad_loop_generic<<<(n + 127) / 128, 128, 0, st>>>(d_src, src_stride, d_ref, ref_stride, bh, bw, src_stride_raw,
                                                                  skip, sa_w, sa_h, d_key);
    }
    return 1;
}

// svt_ext_all_sad_calculation_8x8_16x16 / svt_ext_sad_calculation_8x8_16x16: raw SADs of all 85 PUs at
// n_pos consecutive horizontal positions, through the tile routines of k_fp (one warp). d_ref: 96-byte pitch.
template <bool SUB>
__global__ void __launch_bounds__(32) k_kat_octet(const __grid_constant__ CUtensorMap map, const uint8_t *src, int src_stride, int n_pos,
                                                  uint32_t *out /*[n_pos][85]*/) {
    __shared__ __align__(128) uint8_t s_buf[2 * TM_FP_BUF];
    __shared__ __align__(8) uint64_t s_bar[2];
    TilePipe P;
    pipe_init(P, s_buf, TM_FP_BUF, s_bar, TM_FP_BOX_W * TM_FP_BOX_H);
    pipe_issue(P, 0, &map, 0, 0);
    FpLane<SUB> L;
    fp_lane_init<SUB>(L, src, src_stride, 64, 64);
    pipe_wait(P, 0);
    const int lane = threadIdx.x & 31, strip = lane & 3, band = lane >> 2;
    const int z = c_z16b[4 * (band >> 1) + strip], k = 2 * (band & 1);
    for (int p = 0; p < n_pos; p++) {
        uint32_t raw[5];
        fp_eval_one<SUB>(L, P.buf[0], (uint32_t)p, 0u, raw);
        uint32_t *o = out + p * 85;
        const int dbl = SUB ? 1 : 0;
        o[21 + 4 * z + k] = raw[0] << dbl;
        o[21 + 4 * z + k + 1] = raw[1] << dbl;
        if ((band & 1) == 0) o[5 + z] = raw[2] << dbl;
        if ((band & 3) == 0 && (strip & 1) == 0) o[1 + 2 * (band >> 2) + (strip >> 1)] = raw[3] << dbl;
        if (lane == 0) o[0] = raw[4] << dbl;
    }
}

int svtme_launch_kat_octet(const uint8_t *d_src, int src_stride, const uint8_t *d_ref, int ref_pitch, int ref_rows, int sub, int n_pos,
                           uint32_t *d_out, cudaStream_t st) {
    CUtensorMap map;
    if (svtme_encode_tmap(&map, d_ref, ref_pitch, ref_rows, TM_FP_BOX_W, TM_FP_BOX_H)) return 0;
    if (sub) k_kat_octet<true><<<1, 32, 0, st>>>(map, d_src, src_stride, n_pos, d_out);
    else k_kat_octet<false><<<1, 32, 0, st>>>(map, d_src, src_stride, n_pos, d_out);
    return 1;
}

__global__ void k_kat_nxm(const uint8_t *src, int src_stride, const uint8_t *ref, int ref_stride, int h, int w,
                          uint32_t *out) {
    const uint32_t s = warp_nxm_sad2(src, src_stride, ref, ref_stride, h, w);
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

// ------------------------------------------------------------------------------------------
// the tile plan of a 1/16-plane window, on the host (include/svtme_b200.h)
// ------------------------------------------------------------------------------------------
extern "C" void svtme_s16_tile_plan(int sa_w, int sa_h, int block16x8, int level0, int out[7]) {
    int tw = 0, th = 0, ntx = 0, nty = 0, wide = 0;
    s16_tile_geometry(sa_w, sa_h, &tw, &th, &ntx, &nty, &wide, block16x8 != 0, level0 != 0);
    out[0] = tw; out[1] = th; out[2] = ntx; out[3] = nty;
    const int bw = 16, rows = block16x8 ? 8 : 16, step = block16x8 ? 2 : 1;
    int total = 0;
    out[4] = out[5] = -1;
    for (int ty = 0; ty < nty; ty++)
        for (int tx = 0; tx < ntx; tx++) {
            const int w = sa_w - tx * tw < tw ? sa_w - tx * tw : tw, h = sa_h - ty * th < th ? sa_h - ty * th : th;
            int cost = 0;
            const int which = s16_eval_choice(w, h, bw, rows, step, false, wide != 0, level0 != 0, &cost);
            total += cost;
            if (tx == 0 && ty == 0) out[4] = which;
            if (tx == 0 && ty == nty - 1) out[5] = which;
        }
    out[6] = total;
}
