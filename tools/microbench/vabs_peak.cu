// What is the real issue-rate ceiling of VABSDIFF4.U8.ACC on sm_100a? Independent accumulators,
// operands refreshed by a few cheap ALU ops per iteration so that nothing can be hoisted.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); return 1; } } while (0)
constexpr int ITERS = 2048;
template <int NACC, int NW>
__global__ void k(uint32_t *out, long long *cyc, uint32_t seed) {
    uint32_t acc[NACC], w[NW], s[4];
#pragma unroll
    for (int i = 0; i < NACC; i++) acc[i] = 0;
#pragma unroll
    for (int i = 0; i < NW; i++) w[i] = seed * (i + 3) + threadIdx.x;
#pragma unroll
    for (int i = 0; i < 4; i++) s[i] = seed * (i + 11) ^ threadIdx.x;
    long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < ITERS; it++) {
#pragma unroll
        for (int i = 0; i < NW; i++) w[i] += 0x01020304u;
#pragma unroll
        for (int i = 0; i < NACC; i++) acc[i] += __vsadu4(w[i % NW], s[(i / NW) & 3]);
    }
    long long t1 = clock64();
    uint32_t r = 0;
#pragma unroll
    for (int i = 0; i < NACC; i++) r ^= acc[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}
template <int NACC, int NW>
int run(int nsm, uint32_t *d_out, long long *d_cyc) {
    for (int nthr : {128, 256, 512, 1024}) {
        k<NACC, NW><<<nsm, nthr>>>(d_out, d_cyc, 12345u); CK(cudaDeviceSynchronize());
        k<NACC, NW><<<nsm, nthr>>>(d_out, d_cyc, 12345u); CK(cudaDeviceSynchronize());
        long long *h = new long long[nsm];
        CK(cudaMemcpy(h, d_cyc, sizeof(long long) * nsm, cudaMemcpyDeviceToHost));
        double avg = 0; for (int i = 0; i < nsm; i++) avg += (double)h[i]; avg /= nsm; delete[] h;
        printf("NACC=%2d NW=%2d warps/SM=%2d  VABS lane-instr/clk/SM = %6.2f (+%d IADD per %d VABS)\n", NACC, NW, nthr / 32,
               (double)NACC * ITERS * nthr / avg, NW, NACC);
    }
    return 0;
}
int main() {
    cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
    const int nsm = p.multiProcessorCount;
    uint32_t *d_out; long long *d_cyc;
    CK(cudaMalloc(&d_out, 4 * nsm * 1024)); CK(cudaMalloc(&d_cyc, 8 * nsm));
    run<32, 8>(nsm, d_out, d_cyc);
    run<32, 4>(nsm, d_out, d_cyc);
    run<64, 8>(nsm, d_out, d_cyc);
    run<16, 2>(nsm, d_out, d_cyc);
    printf("done\n");
    return 0;
}
