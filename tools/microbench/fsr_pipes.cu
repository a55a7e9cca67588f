// Can the funnel shifts of the SAD kernels move off the integer-ALU pipe? One loop iteration = the work of one strip
// row of strip16_sad8: 6 words refreshed (IADD), 15 shifted words, 32 VABSDIFF4.U8.ACC.
//   MODE 0: shifts as SHF (ALU pipe)              MODE 1: shifts as IMAD.HI + IMAD (FMA pipe), multiplier from a kernel
//   MODE 2: IMAD.HI alone (throughput)            argument so that ptxas cannot strength-reduce it
//   MODE 3: IMAD alone                            MODE 4: 8 of 15 shifts on IMAD, 7 on SHF
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); return 1; } } while (0)
constexpr int ITERS = 2048;

__device__ __forceinline__ uint32_t sad4_acc(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t d;
    asm("vabsdiff4.u32.u32.u32.add %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}
__device__ __forceinline__ uint32_t fsr_imad(uint32_t lo, uint32_t hi, uint32_t m) {
    uint32_t h, r;
    asm("mul.hi.u32 %0, %1, %2;" : "=r"(h) : "r"(lo), "r"(m));
    asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(hi), "r"(m), "r"(h));
    return r;
}

template <int MODE>
__global__ void k(uint32_t *out, long long *cyc, uint32_t seed, uint32_t m1, uint32_t m2, uint32_t m3) {
    uint32_t acc[8], w[6], s[4];
#pragma unroll
    for (int i = 0; i < 8; i++) acc[i] = (MODE == 2 || MODE == 3) ? seed * (i + 5) + threadIdx.x : 0u;
#pragma unroll
    for (int i = 0; i < 6; i++) w[i] = seed * (i + 3) + threadIdx.x;
#pragma unroll
    for (int i = 0; i < 4; i++) s[i] = seed * (i + 11) ^ threadIdx.x;
    const uint32_t mm[3] = {m1, m2, m3};
    long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < ITERS; it++) {
        if (MODE == 2 || MODE == 3) {
#pragma unroll
            for (int r = 0; r < 4; r++)
#pragma unroll
                for (int i = 0; i < 8; i++) {
                    if (MODE == 2) asm("mul.hi.u32 %0, %1, %2;" : "=r"(acc[i]) : "r"(acc[i] + 0u), "r"(mm[i % 3]));
                    else asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(acc[i]) : "r"(acc[i] + 0u), "r"(mm[i % 3]), "r"(s[i & 3]));
                }
            continue;
        }
#pragma unroll
        for (int i = 0; i < 6; i++) w[i] += 0x01020304u;
        uint32_t sh[3][5];
#pragma unroll
        for (int j = 0; j < 3; j++)
#pragma unroll
            for (int i = 0; i < 5; i++) {
                const bool imad = MODE == 1 || (MODE == 4 && ((j * 5 + i) & 1) == 0);
                sh[j][i] = imad ? fsr_imad(w[i], w[i + 1], mm[j]) : __funnelshift_r(w[i], w[i + 1], 8 * (j + 1));
            }
#pragma unroll
        for (int p = 0; p < 8; p++) {
            const int o = p >> 2, j = p & 3;
#pragma unroll
            for (int q = 0; q < 4; q++) acc[p] = sad4_acc(j ? sh[j - 1][o + q] : w[o + q], s[q], acc[p]);
        }
    }
    long long t1 = clock64();
    uint32_t r = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) r ^= acc[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

template <int MODE>
int run(const char *what, int nsm, uint32_t *d_out, long long *d_cyc) {
    for (int nthr : {256, 512, 1024}) {
        for (int rep = 0; rep < 2; rep++) { k<MODE><<<nsm, nthr>>>(d_out, d_cyc, 12345u, 1u << 24, 1u << 16, 1u << 8); CK(cudaDeviceSynchronize()); }
        long long *h = new long long[nsm];
        CK(cudaMemcpy(h, d_cyc, sizeof(long long) * nsm, cudaMemcpyDeviceToHost));
        double avg = 0; for (int i = 0; i < nsm; i++) avg += (double)h[i]; avg /= nsm; delete[] h;
        printf("%-34s warps/SM=%2d  strip rows (32 SAD4)/clk/SM = %6.3f  -> %6.2f SAD4 lane-instr/clk/SM\n", what, nthr / 32,
               (double)ITERS * nthr / avg, 32.0 * ITERS * nthr / avg);
    }
    return 0;
}
int main() {
    cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
    const int nsm = p.multiProcessorCount;
    uint32_t *d_out; long long *d_cyc;
    CK(cudaMalloc(&d_out, 4 * nsm * 1024)); CK(cudaMalloc(&d_cyc, 8 * nsm));
    run<0>("32 VABS + 15 SHF + 6 IADD", nsm, d_out, d_cyc);
    run<1>("32 VABS + 15 (IMAD.HI+IMAD) + 6 IADD", nsm, d_out, d_cyc);
    run<4>("32 VABS + 8 IMAD pairs + 7 SHF", nsm, d_out, d_cyc);
    run<2>("32 IMAD.HI (as 32 'SAD4')", nsm, d_out, d_cyc);
    run<3>("32 IMAD (as 32 'SAD4')", nsm, d_out, d_cyc);
    printf("done\n");
    return 0;
}
