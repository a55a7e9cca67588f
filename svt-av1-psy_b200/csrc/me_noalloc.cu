// me_noalloc.cu — svtme_sad_loop_search_noalloc: window search on device-resident planes for
// callers that are not allowed to allocate GPU memory.
//
// Who needs this: code that runs where cudaMalloc is off limits — an encoder thread inside a
// frame-deadline loop (cudaMalloc takes a device-wide lock and may synchronise), a plug-in hosted
// by an application that owns the whole memory pool of the GPU, or a library called between
// cudaStreamBeginCapture / EndCapture, where allocation calls are illegal. Such callers already
// hold their pictures in device memory and only need the search itself.
//
// The search is split over thread blocks (one block per band of search rows). Each block leaves
// the best position of its band in a scratch slot; a second small kernel picks the first minimum
// in raster order among the slots. That scratch is NOT allocated: it is the file-scope array
// g_noalloc_band_keys below (8 KB, present in every GPU context that loads this library until the
// process exits). Consequences the caller must know:
//   * the array is shared by every call in the process: two calls whose kernels can overlap on
//     the GPU (different streams, different host threads) race on it. Serialise them.
//   * nothing in it has to be cleared between calls: every slot a call reads was written by the
//     same call. Slots beyond the current call's band count keep the previous call's contents and
//     are never read.
//   * the first call on each device looks the array's device address up (cudaGetSymbolAddress).
//     Make that first call outside any stream capture; later calls only launch kernels.
#include <cuda_runtime.h>

#include <cstdio>
#include <mutex>

#include "me_noalloc_search.cuh"
#include "svtme_b200.h"

#define NOALLOC_MAX_BANDS SVTME_NOALLOC_MAX_BANDS
#define NOALLOC_BAND_ROWS 32 /* even: keeps the parity the skip_search_line rule looks at */
#define NOALLOC_WIN_BYTES (36 * 1024) /* + the static HmeSh stays under the 48 KB a kernel gets without opting in */
#define NOALLOC_MAX_DEVICES 64

static __device__ __align__(256) unsigned long long g_noalloc_band_keys[NOALLOC_MAX_BANDS];

// In host code the name g_noalloc_band_keys is only a handle for the runtime, not a device
// address, so kernels get the address that cudaGetSymbolAddress returned for the current device.
static unsigned long long *g_band_keys_addr[NOALLOC_MAX_DEVICES];
static std::mutex          g_noalloc_mu;

__global__ void __launch_bounds__(HME_THREADS)
    k_noalloc_band_search(const uint8_t *src, int src_stride, const uint8_t *ref, int raw_stride, int row_step, int bh,
                          int bw, int skip, int sa_w, int sa_h, unsigned long long *band_keys) {
    extern __shared__ __align__(16) uint32_t na_win[];
    __shared__ __align__(16) HmeSh sh;
    const int tid = threadIdx.x, band = blockIdx.x;
    const int y0 = band * NOALLOC_BAND_ROWS;
    if (tid == 0) {
        HmeJob &j    = sh.jobs[0];
        j.ref        = ref + (size_t)y0 * raw_stride;
        j.ref_stride = raw_stride;
        j.sa_w       = sa_w;
        j.sa_h       = imin(NOALLOC_BAND_ROWS, sa_h - y0);
        j.origin_x = j.origin_y = 0;
        sh.job_key[0] = HME_KEY_INIT;
    }
    hme_stage_src(sh.src16, 4, src, src_stride, 1, bh, bw, tid);
    __syncthreads();
    hme_search_staged(sh, sh.jobs[0], 0, na_win, NOALLOC_WIN_BYTES / 4, sh.src16, bw, bh, row_step,
                      skip && bw == 16 && bh <= 16);
    __syncthreads();
    if (tid == 0) {
        unsigned long long k = sh.job_key[0];
        if ((uint32_t)k != 0xffffffffu) // position inside the band -> position inside the whole window
            k = (k & 0xffffffff00000000ull) | (uint32_t)((uint32_t)k + (uint32_t)(y0 * sa_w));
        band_keys[band] = k;
    }
}

__global__ void k_noalloc_pick(const unsigned long long *band_keys, int n_bands, int sa_w, unsigned long long *best_sad,
                               int16_t *x_center, int16_t *y_center) {
    __shared__ unsigned long long best;
    if (threadIdx.x == 0) best = HME_KEY_INIT;
    __syncthreads();
    unsigned long long k = HME_KEY_INIT;
    for (int i = threadIdx.x; i < n_bands; i += blockDim.x) k = band_keys[i] < k ? band_keys[i] : k;
    if (k != HME_KEY_INIT) atomicMin(&best, k);
    __syncthreads();
    if (threadIdx.x == 0) {
        *best_sad = best >> 32; // 0xffffff when no position was evaluated, as svt_sad_loop_kernel leaves it
        if ((uint32_t)best != 0xffffffffu) {
            *x_center = (int16_t)((uint32_t)best % (uint32_t)sa_w);
            *y_center = (int16_t)((uint32_t)best / (uint32_t)sa_w);
        }
    }
}

extern "C" uint32_t svtme_sad_loop_search_noalloc(const uint8_t *d_src, uint32_t src_stride, const uint8_t *d_ref,
                                                  uint32_t ref_stride, uint32_t block_height, uint32_t block_width,
                                                  uint64_t *d_best_sad, int16_t *d_x_search_center,
                                                  int16_t *d_y_search_center, uint32_t src_stride_raw,
                                                  uint8_t skip_search_line, int16_t search_area_width,
                                                  int16_t search_area_height, void *cuda_stream) {
    if (!d_src || !d_ref || !d_best_sad || !d_x_search_center || !d_y_search_center || src_stride_raw == 0)
        return SVTME_ERR_BAD_PARAMETER;
    if (block_width < 1 || block_width > 16 || block_height < 1 || block_height > 16 ||
        (ref_stride != src_stride_raw && ref_stride != 2 * src_stride_raw)) {
        fprintf(stderr, "[svtme_b200] svtme_sad_loop_search_noalloc: block %ux%u / strides %u,%u not supported "
                        "(block at most 16x16, ref_stride = 1x or 2x src_stride_raw)\n",
                block_width, block_height, ref_stride, src_stride_raw);
        return SVTME_ERR_BAD_PARAMETER;
    }
    const int n_bands = search_area_height > 0 ? (search_area_height + NOALLOC_BAND_ROWS - 1) / NOALLOC_BAND_ROWS : 0;
    if (search_area_width > SVTME_NOALLOC_MAX_WIDTH || n_bands > NOALLOC_MAX_BANDS) {
        fprintf(stderr, "[svtme_b200] svtme_sad_loop_search_noalloc: search area %dx%d exceeds the built-in scratch "
                        "(at most %d wide, %d rows)\n",
                search_area_width, search_area_height, SVTME_NOALLOC_MAX_WIDTH, NOALLOC_MAX_BANDS * NOALLOC_BAND_ROWS);
        return SVTME_ERR_BAD_PARAMETER;
    }
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= NOALLOC_MAX_DEVICES) return SVTME_ERR_CUDA;
    unsigned long long *keys = nullptr;
    {
        std::lock_guard<std::mutex> g(g_noalloc_mu);
        if (!g_band_keys_addr[dev]) { // first call on this device (it is the current one)
            void       *p = nullptr;
            cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_band_keys);
            if (e != cudaSuccess || !p) {
                fprintf(stderr, "[svtme_b200] cudaGetSymbolAddress(g_noalloc_band_keys): %s\n", cudaGetErrorString(e));
                return SVTME_ERR_CUDA;
            }
            g_band_keys_addr[dev] = (unsigned long long *)p;
        }
        keys = g_band_keys_addr[dev];
    }
    cudaStream_t st = (cudaStream_t)cuda_stream;
    const int    sa_w = search_area_width > 0 ? search_area_width : 0;
    if (n_bands > 0 && sa_w > 0)
        k_noalloc_band_search<<<n_bands, HME_THREADS, NOALLOC_WIN_BYTES, st>>>(
            d_src, (int)src_stride, d_ref, (int)src_stride_raw, (int)(ref_stride / src_stride_raw), (int)block_height,
            (int)block_width, skip_search_line, sa_w, search_area_height, keys);
    k_noalloc_pick<<<1, 256, 0, st>>>(keys, sa_w > 0 ? n_bands : 0, sa_w > 0 ? sa_w : 1, (unsigned long long *)d_best_sad,
                                      d_x_search_center, d_y_search_center);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        fprintf(stderr, "[svtme_b200] svtme_sad_loop_search_noalloc launch: %s\n", cudaGetErrorString(e));
        return SVTME_ERR_CUDA;
    }
    return SVTME_OK;
}
