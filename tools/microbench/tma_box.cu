// How fast can one warp pull small, arbitrarily aligned 2-D windows of a u8 plane into shared memory?
//   mode 0: TMA tile load (cp.async.bulk.tensor.2d) + mbarrier, double-buffered per warp
//   mode 1: LDG.128 (16-byte aligned chunks) -> STS.128 by the warp itself
// Each warp walks "jobs" = windows at pseudo-random byte offsets around its own 64x64 block, like the
// HME level-2 / full-pel searches do, and touches the window once (one LDS per lane) so that the load
// cannot be dropped. Reports windows/us and cycles per window per SM.
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <cuda.h>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); return 1; } } while (0)

typedef CUresult (*EncodeTiled)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void tma_load_2d(void *dst, const CUtensorMap *map, int x, int y, uint64_t *bar) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(smem_u32(dst)),
                 "l"(map), "r"(x), "r"(y), "r"(smem_u32(bar))
                 : "memory");
}

constexpr int WARPS = 4;

__device__ __forceinline__ void job_xy(int w, int it, int plane_w, int plane_h, int bw, int bh, int *x, int *y) {
    const int nbx = (plane_w - 256) / 64, nby = (plane_h - 256) / 64;
    const int b = (w * 7 + (it >> 4)) % (nbx * nby);
    uint32_t h = (uint32_t)(w * 2654435761u) ^ (uint32_t)(it * 40503u);
    h ^= h >> 13; h *= 0x5bd1e995u; h ^= h >> 15;
    *x = 96 + (b % nbx) * 64 + (int)(h & 63) - 32;
    if (bw < 0) *x &= ~15; // TMA: the innermost start coordinate must be 16-byte aligned (measured: traps otherwise)
    *y = 96 + (b / nbx) * 64 + (int)((h >> 8) & 31) - 16;
    (void)bw; (void)bh;
}

template <int MODE>
__global__ void __launch_bounds__(32 * WARPS) k(const __grid_constant__ CUtensorMap pmap, const CUtensorMap *gmap, const uint8_t *plane, int stride, int plane_w, int plane_h, int bw,
                                                 int bh, int iters, uint32_t *out, long long *cyc) {
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ __align__(8) uint64_t bars[WARPS][2];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int pitch_all = ((bw + 30) >> 4) * 16;
    const int box_bytes = pitch_all * bh, buf_bytes = (box_bytes + 127) & ~127;
    uint8_t *sbase = smem + ((128u - (smem_u32(smem) & 127u)) & 127u);
    uint8_t *buf0 = sbase + (size_t)warp * 2 * buf_bytes, *buf1 = buf0 + buf_bytes;
    if (lane == 0) { mbar_init(&bars[warp][0], 1); mbar_init(&bars[warp][1], 1); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    __syncthreads();
    const int gw = blockIdx.x * WARPS + warp;
    const CUtensorMap *map = MODE == 2 ? &pmap : gmap;
    if (MODE == 0) asm volatile("fence.proxy.tensormap::generic.acquire.gpu [%0], 128;" ::"l"(gmap) : "memory");
    uint32_t acc = 0;
    const long long t0 = clock64();
    if (MODE != 1) {
        int x, y;
        job_xy(gw, 0, plane_w, plane_h, -1, bh, &x, &y);
        if (lane == 0) { mbar_expect_tx(&bars[warp][0], box_bytes); tma_load_2d(buf0, map, x, y, &bars[warp][0]); }
        uint32_t ph0 = 0, ph1 = 0;
        for (int it = 0; it < iters; it++) {
            const int cur = it & 1;
            if (it + 1 < iters) {
                job_xy(gw, it + 1, plane_w, plane_h, -1, bh, &x, &y);
                if (lane == 0) { mbar_expect_tx(&bars[warp][cur ^ 1], box_bytes); tma_load_2d(cur ? buf0 : buf1, map, x, y, &bars[warp][cur ^ 1]); }
            }
            mbar_wait(&bars[warp][cur], cur ? ph1 : ph0);
            if (cur) ph1 ^= 1; else ph0 ^= 1;
            const uint32_t *w32 = (const uint32_t *)(cur ? buf1 : buf0);
            acc += w32[lane] + w32[(box_bytes >> 2) - 1 - lane];
            __syncwarp();
        }
    } else {
        const int chunks = (bw + 15 + 15) >> 4; // aligned 16-byte chunks that cover bw bytes at any alignment
        const int pitch = chunks * 16;
        for (int it = 0; it < iters; it++) {
            int x, y;
            job_xy(gw, it, plane_w, plane_h, bw, bh, &x, &y);
            const uint8_t *g = plane + (size_t)y * stride + (x & ~15);
            const int n = chunks * bh;
            uint4 v[12];
#pragma unroll
            for (int k = 0; k < 12; k++) {
                const int i = lane + 32 * k;
                if (i < n) { const int r = i / chunks, c = i - r * chunks; v[k] = __ldg((const uint4 *)(g + (size_t)r * stride + 16 * c)); }
            }
#pragma unroll
            for (int k = 0; k < 12; k++) {
                const int i = lane + 32 * k;
                if (i < n) { const int r = i / chunks, c = i - r * chunks; *(uint4 *)(buf0 + r * pitch + 16 * c) = v[k]; }
            }
            for (int i = lane + 32 * 12; i < n; i += 32) {
                const int r = i / chunks, c = i - r * chunks;
                *(uint4 *)(buf0 + r * pitch + 16 * c) = __ldg((const uint4 *)(g + (size_t)r * stride + 16 * c));
            }
            __syncwarp();
            const uint32_t *w32 = (const uint32_t *)buf0;
            acc += w32[lane] + w32[(n * 4) - 1 - lane];
            __syncwarp();
        }
    }
    const long long t1 = clock64();
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

int main() {
    cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
    const int nsm = p.multiProcessorCount;
    const int W = 2048 + 192, H = 1200 + 192;
    uint8_t *d_plane; CK(cudaMalloc(&d_plane, (size_t)W * H));
    CK(cudaMemset(d_plane, 7, (size_t)W * H));
    EncodeTiled encode = nullptr;
    cudaDriverEntryPointQueryResult qres;
    CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", (void **)&encode, cudaEnableDefault, &qres));
    if (!encode) { printf("no cuTensorMapEncodeTiled\n"); return 1; }
    uint32_t *d_out; long long *d_cyc; CUtensorMap *d_map;
    CK(cudaMalloc(&d_out, 4 * nsm * 8 * 128)); CK(cudaMalloc(&d_cyc, 8 * nsm * 8)); CK(cudaMalloc(&d_map, sizeof(CUtensorMap)));
    const int shapes[][2] = {{39, 34}, {71, 66}, {79, 72}, {71, 34}, {143, 40}, {23, 16}};
    for (auto &s : shapes) {
        const int bw = s[0], bh = s[1];
        CUtensorMap m;
        cuuint64_t dims[2] = {(cuuint64_t)W, (cuuint64_t)H}, strides[1] = {(cuuint64_t)W};
        cuuint32_t box[2] = {(cuuint32_t)(((bw + 30) >> 4) * 16), (cuuint32_t)bh}, es[2] = {1, 1};
        CUresult r = encode(&m, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, d_plane, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                            CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) { printf("encode failed %d for %dx%d\n", (int)r, bw, bh); continue; }
        CK(cudaMemcpy(d_map, &m, sizeof(m), cudaMemcpyHostToDevice));
        for (int mode = 2; mode >= 0; mode--) {
            const int pitch = ((bw + 30) >> 4) * 16;
            const int buf = ((pitch * bh + 127) & ~127) * 2 * WARPS + 128;
            for (int ctas : {1, 2, 4, 8}) {
                if ((size_t)buf * ctas > 200 * 1024) continue;
                const int iters = 2000;
                for (int rep = 0; rep < 2; rep++) {
                    if (mode == 0) {
                        CK(cudaFuncSetAttribute(k<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, buf));
                        k<0><<<nsm * ctas, 32 * WARPS, buf>>>(m, d_map, d_plane, W, W, H, bw, bh, iters, d_out, d_cyc);
                    } else if (mode == 1) {
                        CK(cudaFuncSetAttribute(k<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, buf));
                        k<1><<<nsm * ctas, 32 * WARPS, buf>>>(m, d_map, d_plane, W, W, H, bw, bh, iters, d_out, d_cyc);
                    } else {
                        CK(cudaFuncSetAttribute(k<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, buf));
                        k<2><<<nsm * ctas, 32 * WARPS, buf>>>(m, d_map, d_plane, W, W, H, bw, bh, iters, d_out, d_cyc);
                    }
                }
                {
                    cudaError_t e = cudaDeviceSynchronize();
                    if (e != cudaSuccess) { printf("mode %d box %dx%d ctas %d: %s\n", mode, bw, bh, ctas, cudaGetErrorString(e)); return 1; }
                }
                long long *h = new long long[nsm * ctas];
                CK(cudaMemcpy(h, d_cyc, sizeof(long long) * nsm * ctas, cudaMemcpyDeviceToHost));
                double avg = 0; for (int i = 0; i < nsm * ctas; i++) avg += (double)h[i]; avg /= nsm * ctas; delete[] h;
                const double per_sm = (double)iters * WARPS * ctas; // windows per SM
                printf("%s box %3dx%2d (%5d B) warps/SM=%2d : %7.1f cycles/window/SM  %6.1f B/clk/SM\n", mode == 1 ? "LDG->STS  " : (mode == 0 ? "TMA(gmap) " : "TMA(param)"), bw, bh,
                       bw * bh, WARPS * ctas, avg / per_sm, per_sm * bw * bh / avg);
            }
        }
    }
    printf("done\n");
    return 0;
}
