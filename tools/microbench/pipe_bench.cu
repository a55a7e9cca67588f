// Micro-benchmark: issue rates of the integer byte-SAD instruction (VABSDIFF4.U8.ACC)
// and the instructions that surround it in the full-pel search kernel on sm_100a.
// Output: lane-instructions per clock per SM for each mix, at several occupancies.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o pipe_bench pipe_bench.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); return 1; } } while (0)

constexpr int ITERS = 4096;

template <int MODE>
__global__ void k_pipe(uint32_t *out, uint32_t seed, long long *cycles) {
    uint32_t a[16], acc[16], k[16];
    uint32_t b = seed * (threadIdx.x + 1);
#pragma unroll
    for (int i = 0; i < 16; i++) { a[i] = seed + i * 0x01020304u + threadIdx.x; acc[i] = 0; k[i] = 0xffffffffu; }
    __syncthreads();
    long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < ITERS; it++) {
#pragma unroll
        for (int i = 0; i < 16; i++) {
            if (MODE == 0) {            // VABSDIFF4.ACC only
                acc[i] = (acc[i] + __vsadu4(a[i], b));
            } else if (MODE == 1) {     // IADD3-ish (alu pipe)
                acc[i] = acc[i] + a[i] + b;
            } else if (MODE == 2) {     // IMAD (fma pipe)
                acc[i] = acc[i] * b + a[i];
            } else if (MODE == 3) {     // VABS + IMAD pack + min  (8 vabs : 1 imad : 1 min)
                acc[i] = (acc[i] + __vsadu4(a[i], b));
                if ((i & 7) == 7) { uint32_t key = acc[i] * 2048u + it; k[i] = min(k[i], key); }
            } else if (MODE == 4) {     // VABS : IMAD 1:1
                acc[i] = (acc[i] + __vsadu4(a[i], b));
                k[i] = k[i] * b + a[i];
            } else if (MODE == 5) {     // VABS : IADD 1:1 (same pipe?)
                acc[i] = (acc[i] + __vsadu4(a[i], b));
                k[i] = k[i] + a[i] + b;
            } else if (MODE == 6) {     // umin only
                k[i] = min(k[i], a[i] ^ (uint32_t)it);
            } else if (MODE == 7) {     // vabsdiff4 without acc + dp4a
                acc[i] = __dp4a(__vabsdiffu4(a[i], b), 0x01010101u, acc[i]);
            } else if (MODE == 8) {     // funnel shift + vabs
                uint32_t s = __funnelshift_r(a[i], a[(i + 1) & 15], (threadIdx.x & 3) * 8);
                acc[i] = (acc[i] + __vsadu4(s, b));
            }
        }
        b += 0x01010101u;
    }
    long long t1 = clock64();
    uint32_t r = 0;
#pragma unroll
    for (int i = 0; i < 16; i++) r ^= acc[i] ^ k[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
    if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

// shared-memory read patterns: LDS.32 distinct, LDS.128 distinct, LDS.128 4-address broadcast
template <int MODE>
__global__ void k_lds(uint32_t *out, long long *cycles) {
    __shared__ __align__(16) uint32_t sm[8192];
    for (int i = threadIdx.x; i < 8192; i += blockDim.x) sm[i] = i * 2654435761u;
    __syncthreads();
    uint32_t r = 0;
    int lane = threadIdx.x & 31;
    long long t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < ITERS; it++) {
        int base = (it * 64) & 4095;
#pragma unroll
        for (int j = 0; j < 8; j++) {
            if (MODE == 0) {         // LDS.32, 32 distinct consecutive words
                r ^= sm[base + j * 32 + lane];
            } else if (MODE == 1) {  // LDS.128, 32 distinct 16B chunks (512 B)
                uint4 v = *reinterpret_cast<uint4 *>(&sm[base + j * 128 + lane * 4]);
                r ^= v.x ^ v.y ^ v.z ^ v.w;
            } else if (MODE == 2) {  // LDS.128, 4 distinct chunks (lane&3), broadcast to 8 lanes each
                uint4 v = *reinterpret_cast<uint4 *>(&sm[base + j * 16 + (lane & 3) * 4]);
                r ^= v.x ^ v.y ^ v.z ^ v.w;
            } else if (MODE == 3) {  // LDS.128, all lanes same chunk
                uint4 v = *reinterpret_cast<uint4 *>(&sm[base + j * 4]);
                r ^= v.x ^ v.y ^ v.z ^ v.w;
            } else if (MODE == 4) {  // LDS.128, 8 distinct chunks by lane>>2 (quarter-warp-friendly)
                uint4 v = *reinterpret_cast<uint4 *>(&sm[base + j * 32 + (lane >> 2) * 4]);
                r ^= v.x ^ v.y ^ v.z ^ v.w;
            } else if (MODE == 5) {  // LDS.64 distinct
                uint2 v = *reinterpret_cast<uint2 *>(&sm[base + j * 64 + lane * 2]);
                r ^= v.x ^ v.y;
            } else if (MODE == 6) {  // shfl
                r ^= __shfl_xor_sync(0xffffffffu, r + j, 1);
            }
        }
    }
    long long t1 = clock64();
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
    if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

template <typename F>
static int run(const char *name, F launch, int nblk, int nthr, double ops_per_thread, uint32_t *d_out, long long *d_cyc) {
    launch(nblk, nthr);
    CK(cudaDeviceSynchronize());
    launch(nblk, nthr);
    CK(cudaDeviceSynchronize());
    long long *h = new long long[nblk];
    CK(cudaMemcpy(h, d_cyc, sizeof(long long) * nblk, cudaMemcpyDeviceToHost));
    double avg = 0; for (int i = 0; i < nblk; i++) avg += (double)h[i]; avg /= nblk;
    delete[] h;
    // one block per SM: lane-instr per clk per SM
    printf("%-34s thr/SM=%4d  cycles=%10.0f  lane-ops/clk/SM=%7.2f\n", name, nthr, avg, ops_per_thread * nthr / avg);
    return 0;
}

int main() {
    cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
    int nsm = p.multiProcessorCount;
    printf("device %s, %d SMs, clock %d kHz\n", p.name, nsm, p.clockRate);
    uint32_t *d_out; long long *d_cyc;
    CK(cudaMalloc(&d_out, sizeof(uint32_t) * nsm * 1024));
    CK(cudaMalloc(&d_cyc, sizeof(long long) * nsm));
    const char *names[] = {"VABSDIFF4.ACC", "IADD3", "IMAD", "8VABS+IMAD+MIN", "VABS+IMAD 1:1", "VABS+IADD 1:1", "UMIN", "VABSDIFF4+DP4A", "SHF+VABS"};
    double ops[] = {16, 16, 16, 16 + 4, 32, 32, 16, 32, 32};
    int thr[] = {128, 256, 512, 1024};
    for (int t = 0; t < 4; t++) {
        int nt = thr[t];
#define RUNP(M) run(names[M], [&](int b, int n) { k_pipe<M><<<b, n>>>(d_out, 12345u, d_cyc); }, nsm, nt, ops[M] * ITERS, d_out, d_cyc)
        RUNP(0); RUNP(1); RUNP(2); RUNP(3); RUNP(4); RUNP(5); RUNP(6); RUNP(7); RUNP(8);
    }
    const char *lnames[] = {"LDS.32 distinct", "LDS.128 distinct", "LDS.128 4-addr bcast", "LDS.128 1-addr bcast", "LDS.128 8-addr by lane>>2", "LDS.64 distinct", "SHFL.BFLY"};
    for (int t = 0; t < 4; t++) {
        int nt = thr[t];
#define RUNL(M) run(lnames[M], [&](int b, int n) { k_lds<M><<<b, n>>>(d_out, d_cyc); }, nsm, nt, 8.0 * ITERS, d_out, d_cyc)
        RUNL(0); RUNL(1); RUNL(2); RUNL(3); RUNL(4); RUNL(5); RUNL(6);
    }
    printf("done\n");
    return 0;
}
