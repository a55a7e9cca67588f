// me_kat.cu — RTCD-signature entry points (include/svtme_b200.h, section 1). Each call stages the
// bytes the reference kernel would touch into device memory, runs the sm_100a kernel (the same
// device routines the picture pipeline uses) and copies the outputs back. Known-answer parity path:
// one launch per call is ~10^4x too fine-grained for throughput.
#include <cuda_runtime.h>

#include <cstdio>
#include <cstring>
#include <mutex>

#include "me_device.cuh"
#include "me_launch.h"
#include "svtme_b200.h"

static thread_local uint32_t t_last_error = SVTME_OK;
extern "C" uint32_t svtme_last_error(void) { return t_last_error; }

namespace {
struct Scratch {
    uint8_t *d = nullptr;
    size_t   cap = 0;
    std::mutex mu;
    // returns a device pointer with SVTME_SLAB_SLACK bytes readable before and after [p, p+bytes)
    uint8_t *need(size_t bytes) {
        const size_t want = bytes + 2 * SVTME_SLAB_SLACK;
        if (want > cap) {
            if (d) cudaFree(d);
            cap = (want + (1u << 20)) & ~((size_t)(1u << 20) - 1);
            if (cudaMalloc(&d, cap) != cudaSuccess) { d = nullptr; cap = 0; return nullptr; }
            cudaMemset(d, 0, cap);
        }
        return d + SVTME_SLAB_SLACK;
    }
};
Scratch g_src, g_ref, g_out, g_pitched;

bool fail(const char *what, cudaError_t e) {
    fprintf(stderr, "[svtme_b200] %s: %s (no CPU fallback)\n", what, cudaGetErrorString(e));
    t_last_error = SVTME_ERR_CUDA;
    return false;
}
bool ok(cudaError_t e, const char *what) { return e == cudaSuccess ? true : fail(what, e); }

// simple one-thread kernels for the SAD-aggregation entry points (inputs are 16x16 SADs, not pixels)
__global__ void k_eight_32_64(const uint32_t *sad16 /*[16][8]*/, uint32_t *best32, uint32_t *best64, uint32_t *mv32,
                              uint32_t *mv64, uint32_t mv, uint32_t *sad32 /*[4][8]*/, int n_pos) {
    // motion_estimation.c:369-425 (n_pos = 8) and :171-205 (n_pos = 1, arrays are [16] / [4])
    const int16_t mvx = (int16_t)(mv & 0xFFFF), mvy = (int16_t)(mv >> 16);
    for (int i = 0; i < n_pos; i++) {
        const uint32_t mvi = n_pos == 1 ? mv : (((uint32_t)(uint16_t)mvy << 16) | (uint16_t)(int16_t)(mvx + (int16_t)i));
        uint32_t total = 0;
        for (int k = 0; k < 4; k++) {
            uint32_t s = 0;
            for (int q = 0; q < 4; q++) s += sad16[(4 * k + q) * n_pos + i];
            sad32[k * n_pos + i] = s;
            if (s < best32[k]) { best32[k] = s; mv32[k] = mvi; }
            total += s;
        }
        if (total < best64[0]) { best64[0] = total; mv64[0] = mvi; }
    }
}
} // namespace

#define KAT_BEGIN()                      \
    t_last_error = SVTME_OK;             \
    std::lock_guard<std::mutex> lk_(g_src.mu)

extern "C" void svtme_sad_loop_kernel(uint8_t *src, uint32_t src_stride, uint8_t *ref, uint32_t ref_stride,
                                      uint32_t block_height, uint32_t block_width, uint64_t *best_sad,
                                      int16_t *x_search_center, int16_t *y_search_center, uint32_t src_stride_raw,
                                      uint8_t skip_search_line, int16_t search_area_width, int16_t search_area_height) {
    KAT_BEGIN();
    *best_sad = 0xffffff;
    if (search_area_width <= 0 || search_area_height <= 0 || block_height == 0 || block_width == 0) return;
    const size_t src_bytes = (size_t)(block_height - 1) * src_stride + block_width;
    const size_t ref_bytes = (size_t)(search_area_height - 1) * src_stride_raw + (size_t)(block_height - 1) * ref_stride +
        (size_t)(search_area_width - 1) + block_width;
    // blocks up to 64 wide: source rows re-pitched to 64 bytes (the picture store keeps block origins 16-byte aligned)
    const bool repitch = block_width <= 64;
    const size_t pitched_cap = ((size_t)search_area_width + block_width + 48) * ((size_t)search_area_height + 64 * 2 + 8);
    uint8_t *ds = g_src.need(repitch ? (size_t)64 * block_height : src_bytes), *dr = g_ref.need(ref_bytes), *dk = g_out.need(64);
    uint8_t *dp = g_pitched.need(pitched_cap);
    if (!ds || !dr || !dk || !dp) { t_last_error = SVTME_ERR_INSUFFICIENT_RESOURCES; return; }
    if (repitch) {
        if (!ok(cudaMemset(ds, 0, (size_t)64 * block_height), "memset src")) return;
        if (!ok(cudaMemcpy2D(ds, 64, src, src_stride, block_width, block_height, cudaMemcpyHostToDevice), "H2D src")) return;
    } else if (!ok(cudaMemcpy(ds, src, src_bytes, cudaMemcpyHostToDevice), "H2D src")) {
        return;
    }
    if (!ok(cudaMemcpy(dr, ref, ref_bytes, cudaMemcpyHostToDevice), "H2D ref")) return;
    svtme_launch_kat_sad_loop(ds, repitch ? 64 : (int)src_stride, dr, (int)ref_stride, (int)block_height, (int)block_width,
                              (int)src_stride_raw, skip_search_line, search_area_width, search_area_height,
                              (unsigned long long *)dk, dp, pitched_cap, 0);
    unsigned long long key = 0;
    if (!ok(cudaMemcpy(&key, dk, 8, cudaMemcpyDeviceToHost), "sad_loop kernel")) return;
    *best_sad = key >> 32;
    if ((uint32_t)key != 0xffffffffu) {
        *x_search_center = (int16_t)((uint32_t)key % (uint32_t)search_area_width);
        *y_search_center = (int16_t)((uint32_t)key / (uint32_t)search_area_width);
    }
}

// runs the production octet routine; h_out[0..8*85) raw SADs per position, then [8*85 .. 9*85) best
// SAD over the 8 positions and [9*85 .. 10*85) the index of its first occurrence
static bool run_octet(const uint8_t *src, uint32_t src_stride, const uint8_t *ref, uint32_t ref_stride, int sub, int n_pos,
                      int rows, int cols, uint32_t *h_out) {
    const size_t src_bytes = (size_t)(rows - 1) * src_stride + cols;
    const size_t ref_bytes = (size_t)(rows - 1) * ref_stride + cols + (n_pos - 1);
    // the device routine always evaluates the full 64x64 block: give it zero-padded copies, the reference one with the
    // 96-byte row pitch of a full-pel tile
    uint8_t *ds = g_src.need((size_t)64 * 64), *dr = g_ref.need((size_t)96 * 72), *dk = g_out.need((size_t)8 * 85 * 4);
    if (!ds || !dr || !dk) { t_last_error = SVTME_ERR_INSUFFICIENT_RESOURCES; return false; }
    (void)src_bytes; (void)ref_bytes;
    if (!ok(cudaMemset(ds, 0, 64 * 64), "memset") || !ok(cudaMemset(dr, 0, 96 * 72), "memset")) return false;
    if (!ok(cudaMemcpy2D(ds, 64, src, src_stride, (size_t)cols, (size_t)rows, cudaMemcpyHostToDevice), "H2D src")) return false;
    if (!ok(cudaMemcpy2D(dr, 96, ref, ref_stride, (size_t)cols + (size_t)(n_pos - 1), (size_t)rows, cudaMemcpyHostToDevice), "H2D ref")) return false;
    if (!svtme_launch_kat_octet(ds, 64, dr, 96, 72, sub, n_pos, (uint32_t *)dk, 0)) { t_last_error = SVTME_ERR_CUDA; return false; }
    return ok(cudaMemcpy(h_out, dk, (size_t)n_pos * 85 * 4, cudaMemcpyDeviceToHost), "octet kernel");
}

static const uint8_t k_z16[16] = {0, 1, 4, 5, 2, 3, 6, 7, 8, 9, 12, 13, 10, 11, 14, 15};

extern "C" void svtme_ext_all_sad_calculation_8x8_16x16(uint8_t *src, uint32_t src_stride, uint8_t *ref,
                                                        uint32_t ref_stride, uint32_t mv, uint32_t *p_best_sad_8x8,
                                                        uint32_t *p_best_sad_16x16, uint32_t *p_best_mv8x8,
                                                        uint32_t *p_best_mv16x16, uint32_t p_eight_sad16x16[16][8],
                                                        uint32_t p_eight_sad8x8[64][8], uint8_t sub_sad) {
    KAT_BEGIN();
    (void)p_eight_sad8x8; // unused by the reference too (motion_estimation.c:218)
    static uint32_t h[8 * 85];
    if (!run_octet(src, src_stride, ref, ref_stride, sub_sad ? 1 : 0, 8, 64, 64, h)) return;
    const int16_t mvx = (int16_t)(mv & 0xFFFF), mvy = (int16_t)(mv >> 16);
    // SADs come from the device; the running-best update in position order (strict '<') is the
    // reference's bookkeeping (:231-280) on 80 values
    for (int i = 0; i < 8; i++) {
        const uint32_t mvi = ((uint32_t)(uint16_t)mvy << 16) | (uint16_t)(int16_t)(mvx + (int16_t)i);
        for (int z = 0; z < 16; z++) {
            for (int k = 0; k < 4; k++) {
                const uint32_t v = h[i * 85 + 21 + 4 * z + k];
                if (v < p_best_sad_8x8[4 * z + k]) { p_best_sad_8x8[4 * z + k] = v; p_best_mv8x8[4 * z + k] = mvi; }
            }
            const uint32_t s = h[i * 85 + 5 + z];
            p_eight_sad16x16[z][i] = s;
            if (s < p_best_sad_16x16[z]) { p_best_sad_16x16[z] = s; p_best_mv16x16[z] = mvi; }
        }
    }
}

extern "C" void svtme_ext_sad_calculation_8x8_16x16(uint8_t *src, uint32_t src_stride, uint8_t *ref, uint32_t ref_stride,
                                                    uint32_t *p_best_sad_8x8, uint32_t *p_best_sad_16x16,
                                                    uint32_t *p_best_mv8x8, uint32_t *p_best_mv16x16, uint32_t mv,
                                                    uint32_t *p_sad16x16, uint32_t *p_sad8x8, uint8_t sub_sad) {
    KAT_BEGIN();
    static uint32_t h[85];
    // one 16x16 block = the top-left 16x16 of a zero-padded 64x64 evaluation (PU 5, 8x8 PUs 21..24)
    if (!run_octet(src, src_stride, ref, ref_stride, sub_sad ? 1 : 0, 1, 16, 16, h)) return;
    for (int k = 0; k < 4; k++) {
        p_sad8x8[k] = h[21 + k];
        if (p_sad8x8[k] < p_best_sad_8x8[k]) { p_best_sad_8x8[k] = p_sad8x8[k]; p_best_mv8x8[k] = mv; }
    }
    const uint32_t s = h[5];
    if (s < p_best_sad_16x16[0]) { p_best_sad_16x16[0] = s; p_best_mv16x16[0] = mv; }
    *p_sad16x16 = s;
}

static void run_32_64(uint32_t *sad16, uint32_t *best32, uint32_t *best64, uint32_t *mv32, uint32_t *mv64, uint32_t mv,
                      uint32_t *sad32, int n_pos) {
    uint32_t *d = (uint32_t *)g_out.need(4096);
    if (!d) { t_last_error = SVTME_ERR_INSUFFICIENT_RESOURCES; return; }
    uint32_t *d16 = d, *db32 = d + 128, *db64 = d + 132, *dm32 = d + 136, *dm64 = d + 140, *d32 = d + 144;
    if (!ok(cudaMemcpy(d16, sad16, (size_t)16 * n_pos * 4, cudaMemcpyHostToDevice), "H2D")) return;
    cudaMemcpy(db32, best32, 16, cudaMemcpyHostToDevice);
    cudaMemcpy(db64, best64, 4, cudaMemcpyHostToDevice);
    cudaMemcpy(dm32, mv32, 16, cudaMemcpyHostToDevice);
    cudaMemcpy(dm64, mv64, 4, cudaMemcpyHostToDevice);
    k_eight_32_64<<<1, 1>>>(d16, db32, db64, dm32, dm64, mv, d32, n_pos);
    if (!ok(cudaMemcpy(best32, db32, 16, cudaMemcpyDeviceToHost), "32x32/64x64 kernel")) return;
    cudaMemcpy(best64, db64, 4, cudaMemcpyDeviceToHost);
    cudaMemcpy(mv32, dm32, 16, cudaMemcpyDeviceToHost);
    cudaMemcpy(mv64, dm64, 4, cudaMemcpyDeviceToHost);
    cudaMemcpy(sad32, d32, (size_t)4 * n_pos * 4, cudaMemcpyDeviceToHost);
}

extern "C" void svtme_ext_eight_sad_calculation_32x32_64x64(uint32_t p_sad16x16[16][8], uint32_t *p_best_sad_32x32,
                                                            uint32_t *p_best_sad_64x64, uint32_t *p_best_mv32x32,
                                                            uint32_t *p_best_mv64x64, uint32_t mv,
                                                            uint32_t p_sad32x32[4][8]) {
    KAT_BEGIN();
    run_32_64(&p_sad16x16[0][0], p_best_sad_32x32, p_best_sad_64x64, p_best_mv32x32, p_best_mv64x64, mv, &p_sad32x32[0][0], 8);
}

extern "C" void svtme_ext_sad_calculation_32x32_64x64(uint32_t *p_sad16x16, uint32_t *p_best_sad_32x32,
                                                      uint32_t *p_best_sad_64x64, uint32_t *p_best_mv32x32,
                                                      uint32_t *p_best_mv64x64, uint32_t mv, uint32_t *p_sad32x32) {
    KAT_BEGIN();
    run_32_64(p_sad16x16, p_best_sad_32x32, p_best_sad_64x64, p_best_mv32x32, p_best_mv64x64, mv, p_sad32x32, 1);
}

extern "C" uint32_t svtme_nxm_sad_kernel(const uint8_t *src, uint32_t src_stride, const uint8_t *ref, uint32_t ref_stride,
                                         uint32_t height, uint32_t width) {
    KAT_BEGIN();
    if (height == 0 || width == 0) return 0;
    const size_t sb = (size_t)(height - 1) * src_stride + width, rb = (size_t)(height - 1) * ref_stride + width;
    uint8_t *ds = g_src.need(sb), *dr = g_ref.need(rb), *dk = g_out.need(64);
    if (!ds || !dr || !dk) { t_last_error = SVTME_ERR_INSUFFICIENT_RESOURCES; return 0; }
    if (!ok(cudaMemcpy(ds, src, sb, cudaMemcpyHostToDevice), "H2D src")) return 0;
    if (!ok(cudaMemcpy(dr, ref, rb, cudaMemcpyHostToDevice), "H2D ref")) return 0;
    svtme_launch_kat_nxm(ds, (int)src_stride, dr, (int)ref_stride, (int)height, (int)width, (uint32_t *)dk, 0);
    uint32_t out = 0;
    ok(cudaMemcpy(&out, dk, 4, cudaMemcpyDeviceToHost), "nxm kernel");
    return out;
}

extern "C" void svtme_initialize_buffer_32bits(uint32_t *pointer, uint32_t count128, uint32_t count32, uint32_t value) {
    KAT_BEGIN();
    const uint32_t n = count128 * 4 + count32;
    if (!n) return;
    uint32_t *d = (uint32_t *)g_out.need((size_t)n * 4);
    if (!d) { t_last_error = SVTME_ERR_INSUFFICIENT_RESOURCES; return; }
    svtme_launch_fill32(d, n, value, 0);
    ok(cudaMemcpy(pointer, d, (size_t)n * 4, cudaMemcpyDeviceToHost), "fill kernel");
}

extern "C" void svtme_downsample_2d(uint8_t *input_samples, uint32_t input_stride, uint32_t input_area_width,
                                    uint32_t input_area_height, uint8_t *decim_samples, uint32_t decim_stride,
                                    uint32_t decim_step) {
    KAT_BEGIN();
    if (decim_step < 2 || input_area_width == 0 || input_area_height == 0) return;
    const uint32_t half = decim_step >> 1;
    if (input_area_width <= half || input_area_height <= half) return;
    const uint32_t ow = (input_area_width - half + decim_step - 1) / decim_step, oh = (input_area_height - half + decim_step - 1) / decim_step;
    const size_t ib = (size_t)(input_area_height - 1) * input_stride + input_area_width;
    uint8_t *di = g_src.need(ib), *dout = g_ref.need((size_t)ow * oh);
    if (!di || !dout) { t_last_error = SVTME_ERR_INSUFFICIENT_RESOURCES; return; }
    if (!ok(cudaMemcpy(di, input_samples, ib, cudaMemcpyHostToDevice), "H2D")) return;
    svtme_launch_downsample_generic(di, (int)input_stride, (int)input_area_width, (int)input_area_height, dout, (int)ow,
                                    (int)decim_step, 0);
    ok(cudaMemcpy2D(decim_samples, decim_stride, dout, ow, ow, oh, cudaMemcpyDeviceToHost), "downsample kernel");
}
