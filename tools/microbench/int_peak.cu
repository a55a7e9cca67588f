// int_peak.cu — which pipe issues VABSDIFF4.U8.ACC on sm_100a and at what rate, alone and next to the instructions
// that surround it in the search kernels (SHF, IADD3, LOP3, IMAD, VIMNMX, DP4A).
//
// Every mix is written as inline PTX on operands that change every iteration, so nothing is hoisted, fused or
// strength-reduced (the round-1 pipe_bench rows "SHF+VABS" and "VABSDIFF4+DP4A" counted instructions the compiler had
// removed, and vabs_peak's 111.42 row was a launch that failed unnoticed). Launch errors are checked. The SASS of each
// kernel was checked with cuobjdump -sass (opcode counts inside the loop, profiles/r02_int_peak.md).
//
// Run plain for the cycle-count rates, and under
//   ncu --metrics sm__inst_executed_pipe_alu.sum,sm__inst_executed_pipe_fma.sum,sm__inst_executed_pipe_fmaheavy.sum,\
//       sm__inst_executed_pipe_fmalite.sum,sm__inst_executed_pipe_uniform.sum,sm__inst_executed.sum,sm__cycles_elapsed.max
// for the pipe each opcode is counted on.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -o int_peak int_peak.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at line %d\n", cudaGetErrorString(e_), __LINE__); return 1; } } while (0)

constexpr int ITERS = 2048;
constexpr int N = 16; // independent chains per thread

__device__ __forceinline__ uint32_t vabs_acc(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t d;
    asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}
__device__ __forceinline__ uint32_t shf_r(uint32_t lo, uint32_t hi, uint32_t s) {
    uint32_t d;
    asm volatile("shf.r.wrap.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(lo), "r"(hi), "r"(s));
    return d;
}
__device__ __forceinline__ uint32_t iadd(uint32_t a, uint32_t b) {
    uint32_t d;
    asm volatile("add.u32 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b));
    return d;
}
__device__ __forceinline__ uint32_t lop_xor(uint32_t a, uint32_t b) {
    uint32_t d;
    asm volatile("xor.b32 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b));
    return d;
}
__device__ __forceinline__ uint32_t imad(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t d;
    asm volatile("mad.lo.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}
__device__ __forceinline__ uint32_t umin_(uint32_t a, uint32_t b) {
    uint32_t d;
    asm volatile("min.u32 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b));
    return d;
}
__device__ __forceinline__ uint32_t dp4a(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t d;
    asm volatile("dp4a.u32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}
__device__ __forceinline__ uint32_t vabs_noacc(uint32_t a, uint32_t b) {
    uint32_t d;
    const uint32_t z = 0;
    asm volatile("vabsdiff4.u32.u32.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(z));
    return d;
}

enum { M_VABS = 0, M_SHF, M_IADD, M_LOP, M_IMAD, M_MIN, M_DP4A, M_VABS_SHF, M_VABS_IADD, M_VABS_IMAD, M_VABS_DP4A, M_VABS_MIN,
       M_VABS2_SHF1, M_VABS4_SHF1_IMAD1, M_COUNT };
static const char *NAMES[M_COUNT] = {"VABSDIFF4.ACC", "SHF.R", "IADD", "LOP3(xor)", "IMAD", "VIMNMX(min.u32)", "IDP.4A", "VABS+SHF 1:1",
                                     "VABS+IADD 1:1", "VABS+IMAD 1:1", "VABSnoacc+DP4A 1:1", "VABS+MIN 1:1", "VABS+SHF 2:1", "VABS+SHF+IMAD 4:1:1"};
static const double OPS[M_COUNT] = {1, 1, 1, 1, 1, 1, 1, 2, 2, 2, 2, 2, 3, 6}; // instructions per chain step (see below)

template <int MODE>
__global__ void __launch_bounds__(1024) k(uint32_t *out, long long *cyc, uint32_t seed) {
    uint32_t a[N], b[N];
#pragma unroll
    for (int i = 0; i < N; i++) { a[i] = seed * (i + 3) + threadIdx.x; b[i] = seed * (i + 11) ^ threadIdx.x; }
    const uint32_t sh = (threadIdx.x & 3) * 8 + 8;
    long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < ITERS; it++) {
#pragma unroll
        for (int i = 0; i < N; i++) {
            // every step feeds its own result back: a dependent chain per i, N independent chains per thread
            if (MODE == M_VABS) a[i] = vabs_acc(a[i], b[i], a[i]);
            else if (MODE == M_SHF) a[i] = shf_r(a[i], b[i], sh);
            else if (MODE == M_IADD) a[i] = iadd(a[i], b[i]);
            else if (MODE == M_LOP) a[i] = lop_xor(a[i], b[i]);
            else if (MODE == M_IMAD) a[i] = imad(a[i], b[i], a[i]);
            else if (MODE == M_MIN) a[i] = umin_(a[i], b[i]); 
            else if (MODE == M_DP4A) a[i] = dp4a(a[i], b[i], a[i]);
            else if (MODE == M_VABS_SHF) { b[i] = shf_r(b[i], a[i], sh); a[i] = vabs_acc(a[i], b[i], a[i]); }
            else if (MODE == M_VABS_IADD) { b[i] = iadd(b[i], a[i]); a[i] = vabs_acc(a[i], b[i], a[i]); }
            else if (MODE == M_VABS_IMAD) { b[i] = imad(b[i], a[i], b[i]); a[i] = vabs_acc(a[i], b[i], a[i]); }
            else if (MODE == M_VABS_DP4A) { b[i] = vabs_noacc(a[i], b[i]); a[i] = dp4a(b[i], 0x01010101u, a[i]); }
            else if (MODE == M_VABS_MIN) { b[i] = umin_(b[i], a[i]); a[i] = vabs_acc(a[i], b[i], a[i]); }
            else if (MODE == M_VABS2_SHF1) { b[i] = shf_r(b[i], a[i], sh); a[i] = vabs_acc(a[i], b[i], a[i]); a[i] = vabs_acc(b[i], a[i], a[i]); }
            else if (MODE == M_VABS4_SHF1_IMAD1) {
                b[i] = shf_r(b[i], a[i], sh);
                a[i] = vabs_acc(a[i], b[i], a[i]); a[i] = vabs_acc(b[i], a[i], a[i]);
                b[i] = imad(b[i], a[i], b[i]);
                a[i] = vabs_acc(a[i], b[i], a[i]); a[i] = vabs_acc(b[i], a[i], a[i]);
            }
        }
    }
    long long t1 = clock64();
    uint32_t r = 0;
#pragma unroll
    for (int i = 0; i < N; i++) r ^= a[i] ^ b[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
    if (threadIdx.x == 0) cyc[blockIdx.x] = t1 - t0;
}

template <int MODE>
static int run(int nsm, uint32_t *d_out, long long *d_cyc) {
    for (int nthr : {256, 512, 1024}) {
        for (int rep = 0; rep < 2; rep++) {
            CK(cudaMemset(d_cyc, 0, 8 * nsm));
            k<MODE><<<nsm, nthr>>>(d_out, d_cyc, 12345u);
            CK(cudaGetLastError()); // a launch that fails (too many resources) would leave the previous row's cycles behind
            CK(cudaDeviceSynchronize());
        }
        long long *h = new long long[nsm];
        CK(cudaMemcpy(h, d_cyc, sizeof(long long) * nsm, cudaMemcpyDeviceToHost));
        double avg = 0;
        for (int i = 0; i < nsm; i++) avg += (double)h[i];
        avg /= nsm;
        delete[] h;
        printf("%-22s warps/SM=%2d  lane-instr/clk/SM = %7.2f  (%g instr per step)\n", NAMES[MODE], nthr / 32,
               OPS[MODE] * N * ITERS * nthr / avg, OPS[MODE]);
    }
    return 0;
}

int main() {
    cudaDeviceProp p;
    CK(cudaGetDeviceProperties(&p, 0));
    const int nsm = p.multiProcessorCount;
    printf("device %s, %d SMs\n", p.name, nsm);
    uint32_t *d_out;
    long long *d_cyc;
    CK(cudaMalloc(&d_out, 4 * nsm * 1024));
    CK(cudaMalloc(&d_cyc, 8 * nsm));
    if (run<M_VABS>(nsm, d_out, d_cyc)) return 1;
    if (run<M_SHF>(nsm, d_out, d_cyc)) return 1;
    if (run<M_IADD>(nsm, d_out, d_cyc)) return 1;
    if (run<M_LOP>(nsm, d_out, d_cyc)) return 1;
    if (run<M_IMAD>(nsm, d_out, d_cyc)) return 1;
    if (run<M_MIN>(nsm, d_out, d_cyc)) return 1;
    if (run<M_DP4A>(nsm, d_out, d_cyc)) return 1;
    if (run<M_VABS_SHF>(nsm, d_out, d_cyc)) return 1;
    if (run<M_VABS_IADD>(nsm, d_out, d_cyc)) return 1;
    if (run<M_VABS_IMAD>(nsm, d_out, d_cyc)) return 1;
    if (run<M_VABS_DP4A>(nsm, d_out, d_cyc)) return 1;
    if (run<M_VABS_MIN>(nsm, d_out, d_cyc)) return 1;
    if (run<M_VABS2_SHF1>(nsm, d_out, d_cyc)) return 1;
    if (run<M_VABS4_SHF1_IMAD1>(nsm, d_out, d_cyc)) return 1;
    printf("done\n");
    return 0;
}
