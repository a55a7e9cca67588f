// Micro-benchmark of the HME inner loop variants: how close to the VABSDIFF4 issue limit
// (41.5 lane-instr/clk/SM) does a (shared-memory loads + funnel shifts + VABSDIFF4) row loop get?
//   V0: 8 positions / task : 3 LDS.64 + LDS.128(src) + SHF + 32 VABS per row   (current kernels)
//   V1: 16 positions / task: 5 LDS.64 + LDS.128(src) + SHF + 64 VABS per row
//   V2: 8 positions, 4 pre-shifted window copies: 4 x (LDS.128 + LDS.32) + LDS.128(src) + 32 VABS, no SHF
//   V3: 32 VABS per row on registers only (upper bound of the loop structure)
//   V4: 8 positions, PRMT instead of SHF
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o loop_bench loop_bench.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); return 1; } } while (0)
constexpr int ROWS = 8, ITERS = 512;

template <int V>
__global__ void k_loop(uint32_t *out, long long *cycles, int stride) {
    extern __shared__ __align__(16) uint32_t sm[];
    uint32_t *win = sm, *src = sm + 4 * 4096;
    for (int i = threadIdx.x; i < 4 * 4096 + 64; i += blockDim.x) sm[i] = i * 2654435761u;
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t accv[16];
#pragma unroll
    for (int i = 0; i < 16; i++) accv[i] = 0;
    long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < ITERS; it++) {
        const uint32_t *row = win + ((it * 7 + warp * 13) & 63) * stride + 2 * (lane & 7) + ((lane >> 3) * stride);
#pragma unroll 2
        for (int rr = 0; rr < ROWS; rr++, row += 2 * stride) {
            const uint4 sv = *(const uint4 *)(src + 4 * rr);
            const uint32_t s[4] = {sv.x, sv.y, sv.z, sv.w};
            if (V == 0 || V == 4 || V == 5 || V == 6) {
                const uint2 a = *(const uint2 *)(row), b = *(const uint2 *)(row + 2), c = *(const uint2 *)(row + 4);
                const uint32_t w[6] = {a.x, a.y, b.x, b.y, c.x, c.y};
#pragma unroll
                for (int p = 0; p < 8; p++) {
                    const int o = p >> 2, sh = (p & 3) * 8;
#pragma unroll
                    for (int k = 0; k < 4; k++) {
                        uint32_t v;
                        if (V == 0) v = sh ? __funnelshift_r(w[o + k], w[o + k + 1], sh) : w[o + k];
                        else if (V == 5) v = sh ? __umulhi(w[o + k], 1u << (32 - sh)) + w[o + k + 1] * (1u << (32 - sh)) : w[o + k];
                        else if (V == 6) v = sh ? (uint32_t)(((unsigned long long)w[o + k] * (1u << (32 - sh))) >> 32) + w[o + k + 1] * (1u << (32 - sh)) : w[o + k];
                        else v = sh ? __byte_perm(w[o + k], w[o + k + 1], sh == 8 ? 0x4321 : (sh == 16 ? 0x5432 : 0x6543)) : w[o + k];
                        accv[p] += __vsadu4(v, s[k]);
                    }
                }
            } else if (V == 1) {
                uint32_t w[10];
#pragma unroll
                for (int k = 0; k < 5; k++) { const uint2 a = *(const uint2 *)(row + 2 * k); w[2 * k] = a.x; w[2 * k + 1] = a.y; }
#pragma unroll
                for (int p = 0; p < 16; p++) {
                    const int o = p >> 2, sh = (p & 3) * 8;
#pragma unroll
                    for (int k = 0; k < 4; k++) {
                        const uint32_t v = sh ? __funnelshift_r(w[o + k], w[o + k + 1], sh) : w[o + k];
                        accv[p] += __vsadu4(v, s[k]);
                    }
                }
            } else if (V == 2) {
#pragma unroll
                for (int c = 0; c < 4; c++) {
                    const uint32_t *cp = row + c * 4096 - 2 * (lane & 7) + 4 * (lane & 7);
                    const uint4 a = *(const uint4 *)((const uint32_t *)((uintptr_t)cp & ~(uintptr_t)15));
                    const uint32_t e = cp[4];
                    const uint32_t w[5] = {a.x, a.y, a.z, a.w, e};
#pragma unroll
                    for (int k = 0; k < 4; k++) {
                        accv[c] += __vsadu4(w[k], s[k]);
                        accv[c + 4] += __vsadu4(w[k + 1], s[k]);
                    }
                }
            } else {
#pragma unroll
                for (int p = 0; p < 8; p++)
#pragma unroll
                    for (int k = 0; k < 4; k++) accv[p] += __vsadu4(accv[(p + 1) & 7] ^ s[k], s[k]);
            }
        }
    }
    long long t1 = clock64();
    uint32_t r = 0;
#pragma unroll
    for (int i = 0; i < 16; i++) r ^= accv[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
    if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

template <int V>
int run(const char *name, int nsm, int nthr, int ctas_per_sm, double vabs_per_row, uint32_t *d_out, long long *d_cyc) {
    const int smem = (4 * 4096 + 64) * 4;
    CK(cudaFuncSetAttribute(k_loop<V>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    const int nblk = nsm * ctas_per_sm;
    for (int stride : {18}) {
        k_loop<V><<<nblk, nthr, smem>>>(d_out, d_cyc, stride);
        CK(cudaDeviceSynchronize());
        k_loop<V><<<nblk, nthr, smem>>>(d_out, d_cyc, stride);
        CK(cudaDeviceSynchronize());
        long long *h = new long long[nblk];
        CK(cudaMemcpy(h, d_cyc, sizeof(long long) * nblk, cudaMemcpyDeviceToHost));
        double avg = 0; for (int i = 0; i < nblk; i++) avg += (double)h[i]; avg /= nblk;
        delete[] h;
        printf("%-28s warps/SM=%2d stride=%2d words  VABS lane-instr/clk/SM = %6.2f\n", name, nthr / 32 * ctas_per_sm, stride,
               vabs_per_row * ROWS * ITERS * nthr * ctas_per_sm / avg);
    }
    return 0;
}

int main() {
    cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
    const int nsm = p.multiProcessorCount;
    uint32_t *d_out; long long *d_cyc;
    CK(cudaMalloc(&d_out, sizeof(uint32_t) * nsm * 8 * 1024));
    CK(cudaMalloc(&d_cyc, sizeof(long long) * nsm * 8));
    for (int cps : {2, 3, 4}) {
        const int nthr = 256;
        run<0>("V0 8pos SHF", nsm, nthr, cps, 32, d_out, d_cyc);
        run<5>("V5 8pos umulhi+imad shift", nsm, nthr, cps, 32, d_out, d_cyc);
        run<6>("V6 8pos wide-mul shift", nsm, nthr, cps, 32, d_out, d_cyc);
        run<1>("V1 16pos SHF", nsm, nthr, cps, 64, d_out, d_cyc);
        run<2>("V2 8pos preshifted copies", nsm, nthr, cps, 32, d_out, d_cyc);
        run<3>("V3 registers only", nsm, nthr, cps, 32, d_out, d_cyc);
    }
    printf("done\n");
    return 0;
}
