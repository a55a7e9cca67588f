// me_capi.cu — the extern "C" boundary of libsvtme_b200.so (include/svtme_b200.h): session
// management, picture upload, per-picture ME launch, result fetch. No CPU fallback anywhere: a
// failing CUDA call is reported on stderr and returned as SVTME_ERR_CUDA.
#include <cuda.h>
#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <algorithm>
#include <deque>
#include <mutex>
#include <vector>

#include "me_device.cuh"
#include "me_tma.cuh"
#include "me_launch.h"
#include "svtme_b200.h"

#define CU_TRY(expr)                                                                                       \
    do {                                                                                                   \
        cudaError_t e_ = (expr);                                                                           \
        if (e_ != cudaSuccess) {                                                                           \
            fprintf(stderr, "[svtme_b200] CUDA error %s at %s:%d (%s)\n", cudaGetErrorString(e_), __FILE__, \
                    __LINE__, #expr);                                                                      \
            return SVTME_ERR_CUDA;                                                                         \
        }                                                                                                  \
    } while (0)

static inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

struct ResultLayout { // offsets into the per-picture device scratch slab (what crosses to the host lives in the download regions)
    size_t dg_sad, dg_mv, sb_state, ops, hme_jobs, fp_jobs, fp_keys, fp_counters, end;
};

// SvtMePicResults members in declaration order, and the request group each belongs to
enum { RF_BEST_SAD, RF_BEST_MV, RF_SEARCH, RF_WINDOWS, RF_TOTAL, RF_MV, RF_CAND, RF_D64, RF_D32, RF_D16, RF_D8, RF_RC, RF_VAR, RF_STAT, RF_GM, RF_COUNT };
static const uint32_t rf_group[RF_COUNT] = {SVTME_F_BEST, SVTME_F_BEST, SVTME_F_SEARCH_RESULTS, SVTME_F_WINDOWS, SVTME_F_SB_RESULTS,
                                            SVTME_F_SB_RESULTS, SVTME_F_SB_RESULTS, SVTME_F_STATS, SVTME_F_STATS, SVTME_F_STATS,
                                            SVTME_F_STATS, SVTME_F_STATS, SVTME_F_STATS, SVTME_F_STATS, SVTME_F_STATS};
// placement order inside a block: what every open-loop consumer wants first
static const int rf_order[RF_COUNT] = {RF_TOTAL, RF_MV, RF_CAND, RF_D64, RF_D32, RF_D16, RF_D8, RF_RC, RF_VAR, RF_STAT, RF_GM, RF_SEARCH, RF_WINDOWS,
                                       RF_BEST_SAD, RF_BEST_MV};
static size_t rf_bytes(int f, size_t n_b64, size_t n_pus, size_t max_refs, size_t max_cand) {
    switch (f) {
    case RF_BEST_SAD: case RF_BEST_MV: return n_b64 * 8 * 85 * 4;
    case RF_SEARCH: return n_b64 * 8 * sizeof(SvtMeSearchResult);
    case RF_WINDOWS: return n_b64 * 8 * sizeof(SvtMeWindow);
    case RF_TOTAL: return n_b64 * n_pus;
    case RF_MV: return n_b64 * n_pus * max_refs * 4;
    case RF_CAND: return n_b64 * n_pus * max_cand;
    case RF_STAT: case RF_GM: return n_b64;
    default: return n_b64 * 4;
    }
}
extern "C" void svtme_sb_strides(const SvtMePicParams *pp, uint32_t fields, SvtMeSbStrides *out) {
    out->n_pus = pp->enable_me_8x8 ? 85 : 21;
    out->mv_stride = pp->max_refs; out->mv_l1_offset = pp->max_l0; out->cand_stride = pp->max_cand;
    if (!(fields & SVTME_F_SB_COMPACT)) return;
    const uint32_t n0 = pp->num_of_ref_pic_to_search[0], n1 = pp->num_of_list_to_search > 1 ? pp->num_of_ref_pic_to_search[1] : 0;
    // construct_me_candidate_array* (motion_estimation.c:2602-2905): unipred of every reference, then the bipred pairs
    uint32_t cand = n0 + n1;
    if (n1) cand += pp->only_l_bwd ? 1u : n0 * n1 + (n0 - 1u) + (n1 == 3 ? 1u : 0u);
    if (cand < 1) cand = 1;
    if (cand < out->cand_stride) out->cand_stride = (uint8_t)cand;
    if (n0 + n1 >= 1 && n0 + n1 <= out->mv_stride) { out->mv_stride = (uint8_t)(n0 + n1); out->mv_l1_offset = (uint8_t)n0; }
}

extern "C" void svtme_sb_expand_b64(const SvtMePicParams *pp, uint32_t fields, const SvtMePicResults *src, uint32_t b, uint8_t *dst_total,
                                    uint32_t *dst_mv, uint8_t *dst_cand) {
    SvtMeSbStrides st;
    svtme_sb_strides(pp, fields, &st);
    const size_t np = st.n_pus;
    memcpy(dst_total, src->total_me_candidate_index + (size_t)b * np, np);
    if (st.mv_stride == pp->max_refs && st.mv_l1_offset == pp->max_l0 && st.cand_stride == pp->max_cand) {
        memcpy(dst_mv, src->me_mv_array + (size_t)b * np * pp->max_refs, np * pp->max_refs * 4);
        memcpy(dst_cand, src->me_candidate_array + (size_t)b * np * pp->max_cand, np * pp->max_cand);
        return;
    }
    memset(dst_mv, 0, np * pp->max_refs * 4);
    memset(dst_cand, 0, np * pp->max_cand);
    const uint32_t *mv = src->me_mv_array + (size_t)b * np * st.mv_stride;
    const uint8_t  *cd = src->me_candidate_array + (size_t)b * np * st.cand_stride;
    const uint32_t n0 = st.mv_l1_offset, n1 = st.mv_stride - st.mv_l1_offset;
    for (size_t pu = 0; pu < np; pu++) {
        memcpy(dst_mv + pu * pp->max_refs, mv + pu * st.mv_stride, n0 * 4);
        memcpy(dst_mv + pu * pp->max_refs + pp->max_l0, mv + pu * st.mv_stride + n0, n1 * 4);
        memcpy(dst_cand + pu * pp->max_cand, cd + pu * st.cand_stride, st.cand_stride);
    }
}

static void packed_layout(size_t n_b64, size_t n_pus, size_t max_refs, size_t max_cand, uint32_t fields, SvtMePackedLayout *L) {
    size_t o = 0;
    for (int f = 0; f < RF_COUNT; f++) L->off[f] = (size_t)-1;
    for (int k = 0; k < RF_COUNT; k++) {
        const int f = rf_order[k];
        if (!(fields & rf_group[f])) continue;
        L->off[f] = o;
        o = align_up(o + rf_bytes(f, n_b64, n_pus, max_refs, max_cand), 64);
    }
    L->bytes = align_up(o, 256);
}
static void **result_member(SvtMePicResults *r, int f) { return (void **)r + f; } // fifteen pointers, declaration order
static_assert(sizeof(SvtMePicResults) == RF_COUNT * sizeof(void *), "SvtMePicResults is fifteen pointers");

struct SvtMeSession {
    int          device = 0;
    int          w = 0, h = 0, max_pics = 0, n_b64 = 0, b64_cols = 0, b64_rows = 0;
    cudaStream_t stream = nullptr;
    bool         own_stream = false;
    uint8_t     *pic_slab = nullptr;
    CUtensorMap *d_tmaps = nullptr; // [max_pics][TM_KINDS]
    size_t       pic_bytes = 0;
    std::vector<DevPic> pics;
    std::vector<char>   pic_valid;
    // Result slabs come in two sets used by alternate batches: the D2H copies of batch k (download stream) run
    // while batch k+1 (compute stream) writes the other set. res_slab (set 0, slab 0) also holds the counters.
    std::vector<uint8_t *> res_sets[2]; // one result/scratch slab per picture of a batch
    uint8_t     *res_slab = nullptr;
    cudaStream_t up_stream = nullptr, dl_stream = nullptr; // host uploads (H2D + pyramid) / result downloads (D2H)
    // concurrency 2: batches alternate between `stream` and `stream2` so that the latency-bound control / post kernels
    // of one batch run under the ALU-bound searches of the other. Each set of result slabs then also owns its half of
    // the job array and of the unit queues.
    cudaStream_t stream2 = nullptr;
    int          concurrency = 1;
    cudaEvent_t  ev_built[2] = {};       // last pyramid build queued on compute stream 0 / 1
    bool         built_recorded[2] = {false, false};
    std::vector<uint64_t> slot_last_seq; // last batch that reads each picture slot
    uint32_t     d_jobs_n = 0;           // pictures per set the job array holds
    cudaEvent_t  ev_me[8] = {}, ev_dl[2] = {};
    bool         dl_recorded[2] = {false, false};
    cudaEvent_t  ev_dl_seq[8] = {};      // the last download of batch seq (slot seq % 8): what svtme_session_wait_batch waits for
    uint64_t     dl_seq[8] = {};
    uint64_t     batch_seq = 0; // sequence number of the last launched batch (1-based)
    PicJob      *d_jobs = nullptr;
    // host copies of the job arrays: a ring of pinned arrays (the H2D copy of one is asynchronous; it is rewritten only
    // after the event recorded behind that copy)
    PicJob      *h_jobs_ring = nullptr;
    cudaEvent_t  ev_jobs[4] = {};
    bool         jobs_used[4] = {false, false, false, false};
    uint64_t     jobs_head = 0;
    ResultLayout lay{};
    // What a batch hands to the host: per result set one contiguous device region, the blocks of the batch's pictures
    // back to back (svtme_packed_layout of each picture). BatchOut describes the batch that owns the set now.
    struct BatchOut {
        uint64_t seq = 0;
        bool     me = false; // false: the set was last used by the detector
        struct Geo { size_t n_pus, max_refs, max_cand; };
        std::vector<SvtMePackedLayout> lay;
        std::vector<size_t>            base;
        std::vector<Geo>               geo;
        size_t   total = 0;
    } out[2];
    uint8_t     *dl_region[2] = {nullptr, nullptr};
    size_t       dl_cap[2] = {0, 0};
    uint32_t     default_fields = SVTME_F_ALL;
    int          s16_unit_mode = -1; // svtme_session_set_unit_granularity
    uint32_t    *h_flags = nullptr; // pinned: [set] = the set's unit-queue overflow flag as of its last download
    cudaEvent_t  ev_release[2] = {}; // device-resident consumers of a set (svtme_batch_results_release)
    bool         release_recorded[2] = {false, false};
    // host uploads go through a byte ring of staging memory: one allocation per upload call
    struct StageAlloc { size_t begin, end; cudaEvent_t ev_ingested; };
    uint8_t     *stage_slab = nullptr;
    size_t       stage_cap = 0, stage_head = 0;
    std::deque<StageAlloc>   stage_live;
    std::vector<cudaEvent_t> stage_ev_pool;
    cudaEvent_t  ev_copied = nullptr;
    uint2       *fp_units = nullptr, *s16_units = nullptr;
    uint32_t     fp_units_cap = 0, s16_units_cap = 0;
    uint64_t     launches = 0;
    bool         timing = false;
    std::vector<cudaEvent_t> ev; // triples: before hme, after hme, after full-pel
    size_t       ev_used = 0;
    double       acc_hme_ms = 0, acc_fp_ms = 0;
    uint64_t     n_timed = 0;
    std::vector<SvtMeTimeline> tls; // per-kernel timelines of the timed batches not folded yet
    std::vector<cudaEvent_t>   tl_pool;
    double       acc_tag_ms[SVTME_TL_TAGS] = {};
    std::mutex   mu;
};

// ---- TMA descriptors -------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                  const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn encode_tiled_fn() {
    static EncodeTiledFn fn = [] {
        void *p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess)
            p = nullptr;
        return (EncodeTiledFn)p;
    }();
    return fn;
}
int svtme_encode_tmap(void *out, const void *base, size_t pitch, size_t rows, uint32_t box_w, uint32_t box_h) {
    EncodeTiledFn fn = encode_tiled_fn();
    if (!fn || (pitch & 15) || ((uintptr_t)base & 15)) return 1;
    cuuint64_t dims[2] = {(cuuint64_t)pitch, (cuuint64_t)rows}, strides[1] = {(cuuint64_t)pitch};
    cuuint32_t box[2] = {box_w, box_h}, es[2] = {1, 1};
    const CUresult r = fn((CUtensorMap *)out, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<void *>(base), dims, strides, box, es,
                          CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { fprintf(stderr, "[svtme_b200] cuTensorMapEncodeTiled failed (%d)\n", (int)r); return 1; }
    return 0;
}

static void plane_geom(int w, int h, int level, int *pw, int *ph, int *pad, int *stride, size_t *bytes) {
    static const int pads[3] = {SVTME_FULL_PAD, SVTME_QUARTER_PAD, SVTME_SIXTEENTH_PAD};
    *pw     = w >> level;
    *ph     = h >> level;
    *pad    = pads[level];
    *stride = (int)align_up((size_t)(*pw + 2 * *pad), 16);
    *bytes  = align_up((size_t)*stride * (size_t)(*ph + 2 * *pad) + 2 * SVTME_SLAB_SLACK, 256);
}

static void result_layout(ResultLayout &L, int n_b64) {
    size_t o = 0;
    auto take = [&](size_t bytes) { size_t at = o; o = align_up(o + bytes, 256); return at; };
    L.dg_sad   = take((size_t)n_b64 * 8); // dynamic-GOP detector: per-b64 SAD and vector
    L.dg_mv    = take((size_t)n_b64 * 4);
    L.sb_state = take((size_t)n_b64 * sizeof(SbState));
    L.ops = take(64);
    L.hme_jobs    = take((size_t)n_b64 * HME_JOBS_PER_B64 * sizeof(HmeJobG));
    L.fp_jobs     = take((size_t)n_b64 * 8 * sizeof(FpJobG));
    L.fp_keys     = take((size_t)n_b64 * 8 * 85 * sizeof(unsigned long long));
    L.fp_counters = take(64);
    L.end = o;
}

extern "C" const char *svtme_version(void) { return "svtme_b200 0.1 (sm_100a, open-loop ME of SVT-AV1-PSY 2.3.0-A)"; }

extern "C" uint32_t svtme_session_create(SvtMeSession **out, int device, uint32_t width, uint32_t height,
                                         uint32_t max_pictures) {
    if (!out || width == 0 || height == 0 || (width & 7) || (height & 7) || width > 16384 || height > 16384 ||
        max_pictures == 0 || max_pictures > 4096)
        return SVTME_ERR_BAD_PARAMETER;
    *out = nullptr;
    int n_dev = 0;
    CU_TRY(cudaGetDeviceCount(&n_dev));
    if (device < 0 || device >= n_dev) {
        fprintf(stderr, "[svtme_b200] no CUDA device %d (found %d); there is no CPU fallback\n", device, n_dev);
        return SVTME_ERR_CUDA;
    }
    CU_TRY(cudaSetDevice(device));
    SvtMeSession *s = new SvtMeSession();
    s->device = device; s->w = (int)width; s->h = (int)height; s->max_pics = (int)max_pictures;
    s->b64_cols = (s->w + 63) / 64; s->b64_rows = (s->h + 63) / 64; s->n_b64 = s->b64_cols * s->b64_rows;
    size_t off[3], total = 0;
    int pw[3], ph[3], pad[3], st[3];
    for (int i = 0; i < 3; i++) {
        size_t b;
        plane_geom(s->w, s->h, i, &pw[i], &ph[i], &pad[i], &st[i], &b);
        off[i] = total;
        total += b;
    }
    s->pic_bytes = total;
    cudaError_t e = cudaMalloc(&s->pic_slab, total * max_pictures);
    if (e != cudaSuccess) { delete s; fprintf(stderr, "[svtme_b200] cudaMalloc picture store: %s\n", cudaGetErrorString(e)); return SVTME_ERR_INSUFFICIENT_RESOURCES; }
    cudaMemset(s->pic_slab, 0, total * max_pictures);
    s->pics.resize(max_pictures);
    s->pic_valid.assign(max_pictures, 0);
    for (uint32_t p = 0; p < max_pictures; p++)
        for (int i = 0; i < 3; i++) {
            DevPlane &pl = s->pics[p].lvl[i];
            pl.buf = s->pic_slab + (size_t)p * total + off[i] + SVTME_SLAB_SLACK;
            pl.stride = st[i]; pl.org_x = pad[i]; pl.org_y = pad[i]; pl.width = pw[i]; pl.height = ph[i];
        }
    { // TMA descriptors of every plane the tile searches read (me_tma.cuh)
        std::vector<CUtensorMap> maps((size_t)max_pictures * TM_KINDS);
        e = cudaMalloc(&s->d_tmaps, maps.size() * sizeof(CUtensorMap));
        if (e != cudaSuccess) { cudaFree(s->pic_slab); delete s; return SVTME_ERR_INSUFFICIENT_RESOURCES; }
        static const int kind_level[TM_KINDS] = {1, 0, 0, 2, 2};
        static const uint32_t kind_box[TM_KINDS][2] = {{TM_L1_BOX_W, TM_L1_BOX_H}, {TM_L2_BOX_W, TM_L2_BOX_H}, {TM_FP_BOX_W, TM_FP_BOX_H},
                                                        {TM_S16T_BOX_W, TM_S16T_BOX_H}, {TM_S16W_BOX_W, TM_S16W_BOX_H}};
        for (uint32_t p = 0; p < max_pictures; p++) {
            for (int k = 0; k < TM_KINDS; k++) {
                const DevPlane &pl = s->pics[p].lvl[kind_level[k]];
                if (svtme_encode_tmap(&maps[(size_t)p * TM_KINDS + k], pl.buf, (size_t)pl.stride, (size_t)pl.height + 2 * pl.org_y,
                                      kind_box[k][0], kind_box[k][1])) {
                    cudaFree(s->pic_slab); cudaFree(s->d_tmaps); delete s;
                    return SVTME_ERR_CUDA;
                }
            }
            s->pics[p].tmap = s->d_tmaps + (size_t)p * TM_KINDS;
        }
        e = cudaMemcpy(s->d_tmaps, maps.data(), maps.size() * sizeof(CUtensorMap), cudaMemcpyHostToDevice);
        if (e != cudaSuccess) { cudaFree(s->pic_slab); cudaFree(s->d_tmaps); delete s; return SVTME_ERR_CUDA; }
    }
    result_layout(s->lay, s->n_b64);
    e = cudaMalloc(&s->res_slab, s->lay.end);
    if (e != cudaSuccess) { cudaFree(s->pic_slab); delete s; return SVTME_ERR_INSUFFICIENT_RESOURCES; }
    cudaMemset(s->res_slab, 0, s->lay.end);
    s->res_sets[0].push_back(s->res_slab);
    e = cudaMalloc(&s->d_jobs, sizeof(PicJob) * 2);
    if (e != cudaSuccess) { cudaFree(s->pic_slab); cudaFree(s->res_slab); delete s; return SVTME_ERR_INSUFFICIENT_RESOURCES; }
    CU_TRY(cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking));
    s->own_stream = true;
    CU_TRY(cudaStreamCreateWithFlags(&s->stream2, cudaStreamNonBlocking));
    for (auto &e2 : s->ev_built) CU_TRY(cudaEventCreateWithFlags(&e2, cudaEventDisableTiming));
    s->slot_last_seq.assign(max_pictures, 0);
    s->d_jobs_n = 1;
    CU_TRY(cudaStreamCreateWithFlags(&s->up_stream, cudaStreamNonBlocking));
    CU_TRY(cudaStreamCreateWithFlags(&s->dl_stream, cudaStreamNonBlocking));
    for (auto &e2 : s->ev_me) CU_TRY(cudaEventCreateWithFlags(&e2, cudaEventDisableTiming));
    for (auto &e2 : s->ev_dl) CU_TRY(cudaEventCreateWithFlags(&e2, cudaEventDisableTiming));
    for (auto &e2 : s->ev_dl_seq) CU_TRY(cudaEventCreateWithFlags(&e2, cudaEventDisableTiming));
    for (auto &e2 : s->ev_jobs) CU_TRY(cudaEventCreateWithFlags(&e2, cudaEventDisableTiming));
    for (auto &e2 : s->ev_release) CU_TRY(cudaEventCreateWithFlags(&e2, cudaEventDisableTiming));
    CU_TRY(cudaEventCreateWithFlags(&s->ev_copied, cudaEventDisableTiming));
    CU_TRY(cudaMallocHost(&s->h_flags, 2 * sizeof(uint32_t)));
    s->h_flags[0] = s->h_flags[1] = 0;
    CU_TRY(cudaMallocHost(&s->h_jobs_ring, sizeof(PicJob) * 4));
    CU_TRY(cudaDeviceSynchronize());
    *out = s;
    return SVTME_OK;
}

extern "C" void svtme_session_destroy(SvtMeSession *s) {
    if (!s) return;
    cudaSetDevice(s->device);
    cudaStreamSynchronize(s->up_stream);
    cudaStreamSynchronize(s->stream);
    cudaStreamSynchronize(s->stream2);
    cudaStreamSynchronize(s->dl_stream);
    for (auto &e : s->ev) cudaEventDestroy(e);
    for (auto &t : s->tls) for (auto &e : t.ev) cudaEventDestroy(e);
    for (auto &e : s->tl_pool) cudaEventDestroy(e);
    for (auto &e : s->ev_me) if (e) cudaEventDestroy(e);
    for (auto &e : s->ev_dl) if (e) cudaEventDestroy(e);
    for (auto &e : s->ev_dl_seq) if (e) cudaEventDestroy(e);
    if (s->up_stream) cudaStreamDestroy(s->up_stream);
    if (s->stream2) cudaStreamDestroy(s->stream2);
    for (auto &e : s->ev_built) if (e) cudaEventDestroy(e);
    if (s->dl_stream) cudaStreamDestroy(s->dl_stream);
    if (s->own_stream) cudaStreamDestroy(s->stream);
    cudaFree(s->pic_slab);
    if (s->d_tmaps) cudaFree(s->d_tmaps);
    for (auto &set : s->res_sets)
        for (uint8_t *p : set) cudaFree(p);
    if (s->d_jobs) cudaFree(s->d_jobs);
    if (s->stage_slab) cudaFree(s->stage_slab);
    if (s->ev_copied) cudaEventDestroy(s->ev_copied);
    for (auto &a : s->stage_live) cudaEventDestroy(a.ev_ingested);
    for (auto &e : s->stage_ev_pool) cudaEventDestroy(e);
    for (auto &e : s->ev_jobs) if (e) cudaEventDestroy(e);
    for (auto &e : s->ev_release) if (e) cudaEventDestroy(e);
    for (auto &p : s->dl_region) if (p) cudaFree(p);
    if (s->h_flags) cudaFreeHost(s->h_flags);
    if (s->h_jobs_ring) cudaFreeHost(s->h_jobs_ring);
    if (s->fp_units) cudaFree(s->fp_units);
    if (s->s16_units) cudaFree(s->s16_units);
    delete s;
}

extern "C" void *svtme_session_stream(SvtMeSession *s) { return s ? (void *)s->stream : nullptr; }

extern "C" uint32_t svtme_session_set_stream(SvtMeSession *s, void *cuda_stream) {
    if (!s) return SVTME_ERR_BAD_PARAMETER;
    std::lock_guard<std::mutex> g(s->mu);
    cudaSetDevice(s->device);
    CU_TRY(cudaStreamSynchronize(s->up_stream));
    CU_TRY(cudaStreamSynchronize(s->stream));
    CU_TRY(cudaStreamSynchronize(s->stream2));
    CU_TRY(cudaStreamSynchronize(s->dl_stream));
    if (s->own_stream) cudaStreamDestroy(s->stream);
    s->stream     = (cudaStream_t)cuda_stream;
    s->own_stream = false;
    return SVTME_OK;
}

extern "C" uint64_t svtme_session_launch_count(SvtMeSession *s) { return s ? s->launches : 0; }

static uint32_t sync_all(SvtMeSession *s) {
    CU_TRY(cudaStreamSynchronize(s->up_stream));
    CU_TRY(cudaStreamSynchronize(s->stream));
    CU_TRY(cudaStreamSynchronize(s->stream2));
    CU_TRY(cudaStreamSynchronize(s->dl_stream));
    return SVTME_OK;
}

// compute stream of the batches of result set `set` (kernel timing measures stages: one stream then)
static cudaStream_t compute_stream(const SvtMeSession *s, int set) {
    return (s->concurrency == 2 && !s->timing && set == 1) ? s->stream2 : s->stream;
}
static int stream_index(const SvtMeSession *s, int set) { return compute_stream(s, set) == s->stream2 ? 1 : 0; }

// The unit queues are sized from the largest windows the controls allow; a kernel that finds one too small raises a
// flag instead of dropping work silently. The flag of a result set travels to the host behind every download of the set
// (h_flags, checked by svtme_session_wait_batch and the sync = 1 paths) and is read directly at svtme_session_sync.
static uint32_t report_overflow() {
    fprintf(stderr, "[svtme_b200] a search-unit queue overflowed: results of the last batches are incomplete\n");
    return SVTME_ERR_INSUFFICIENT_RESOURCES;
}
static uint32_t check_queue_overflow(SvtMeSession *s) {
    uint32_t flag = 0;
    for (int set = 0; set < 2 && !flag; set++)
        if (!s->res_sets[set].empty()) {
            CU_TRY(cudaMemcpyAsync(&flag, (uint32_t *)(s->res_sets[set][0] + s->lay.fp_counters) + 8, sizeof(flag), cudaMemcpyDeviceToHost, s->stream));
            CU_TRY(cudaStreamSynchronize(s->stream));
            if (flag) cudaMemsetAsync((uint32_t *)(s->res_sets[set][0] + s->lay.fp_counters) + 8, 0, sizeof(flag), s->stream);
        }
    s->h_flags[0] = s->h_flags[1] = 0;
    return flag ? report_overflow() : SVTME_OK;
}
// after a download of `set` has completed on the host side
static uint32_t check_downloaded_flag(SvtMeSession *s, int set) {
    if (!s->h_flags[set]) return SVTME_OK;
    s->h_flags[set] = 0;
    cudaMemsetAsync((uint32_t *)(s->res_sets[set][0] + s->lay.fp_counters) + 8, 0, sizeof(uint32_t), s->dl_stream);
    return report_overflow();
}

extern "C" uint32_t svtme_session_sync(SvtMeSession *s) {
    if (!s) return SVTME_ERR_BAD_PARAMETER;
    cudaSetDevice(s->device);
    uint32_t rc = sync_all(s);
    return rc != SVTME_OK ? rc : check_queue_overflow(s);
}

extern "C" uint32_t svtme_session_set_concurrency(SvtMeSession *s, int n) {
    if (!s || n < 1 || n > 2) return SVTME_ERR_BAD_PARAMETER;
    std::lock_guard<std::mutex> g(s->mu);
    cudaSetDevice(s->device);
    uint32_t rc = sync_all(s);
    if (rc != SVTME_OK) return rc;
    s->concurrency = n;
    return SVTME_OK;
}

extern "C" uint64_t svtme_session_batch_seq(SvtMeSession *s) { return s ? s->batch_seq : 0; }

extern "C" uint32_t svtme_session_wait_batch(SvtMeSession *s, uint64_t seq) {
    if (!s) return SVTME_ERR_BAD_PARAMETER;
    cudaEvent_t ev[4];
    int n_ev = 0;
    bool downloaded = false;
    const int set = (int)(seq & 1);
    { // the events live as long as the session; only the bookkeeping needs the lock (other threads may be launching)
        std::lock_guard<std::mutex> g(s->mu);
        if (seq == 0 || seq > s->batch_seq) return SVTME_OK;
        if (seq + 8 <= s->batch_seq) {
            // eight or more batches were queued since (a caller with nine or more host buffers that never waited): the
            // batch's own event slots were handed on. Everything queued later on the same streams completes after it:
            // the newest batch of either compute stream, and the newest download of either result set.
            ev[n_ev++] = s->ev_me[s->batch_seq % 8];
            ev[n_ev++] = s->ev_me[(s->batch_seq - 1) % 8];
            for (int k = 0; k < 2; k++)
                if (s->dl_recorded[k]) { ev[n_ev++] = s->ev_dl[k]; downloaded = true; }
        } else {
            // the batch's own last download (not a later batch's of the same set: a caller that rotates three or more host
            // buffers waits for batch k - 2 while batch k's download may not even have started)
            downloaded = s->dl_seq[seq % 8] == seq;
            ev[n_ev++] = downloaded ? s->ev_dl_seq[seq % 8] : s->ev_me[seq % 8];
        }
    }
    cudaSetDevice(s->device);
    for (int i = 0; i < n_ev; i++) CU_TRY(cudaEventSynchronize(ev[i]));
    if (!downloaded) return SVTME_OK;
    std::lock_guard<std::mutex> g(s->mu);
    return check_downloaded_flag(s, set);
}

// Host uploads: the frame crosses PCIe as ONE linear copy into a ring of staging frames on the upload stream (pure
// copy-engine work, runs ahead of the kernels); the pyramid build reads the staging frame on the COMPUTE stream, so
// picture planes are only ever written in compute-stream order (no hazard with the ME kernels of earlier batches).
// Measured: building the pyramid on the upload stream gated every copy behind a kernel waiting for an SM slot next
// to the persistent search kernels, and a pitched 2-D copy into the padded plane reaches ~11 GB/s.
#define SVTME_STAGE_FRAMES 40
// a region of the staging ring; the regions it overlaps were consumed by pyramid builds queued earlier: the upload
// stream waits for those before the copy engine overwrites them
static uint32_t stage_alloc(SvtMeSession *s, size_t bytes, uint8_t **dst, cudaEvent_t *ev_ingested) {
    if (!s->stage_slab) {
        s->stage_cap = (size_t)s->w * s->h * 2 * SVTME_STAGE_FRAMES; // 40 frames of 16-bit samples
        cudaError_t e = cudaMalloc(&s->stage_slab, s->stage_cap);
        if (e != cudaSuccess) { s->stage_cap = 0; fprintf(stderr, "[svtme_b200] cudaMalloc staging ring: %s\n", cudaGetErrorString(e)); return SVTME_ERR_INSUFFICIENT_RESOURCES; }
    }
    bytes = align_up(bytes, 256);
    if (bytes > s->stage_cap) return SVTME_ERR_BAD_PARAMETER;
    if (s->stage_head + bytes > s->stage_cap) s->stage_head = 0;
    const size_t b = s->stage_head, e = b + bytes;
    s->stage_head = e;
    while (!s->stage_live.empty()) { // FIFO in ring order: the overlapping regions are the oldest ones
        SvtMeSession::StageAlloc &a = s->stage_live.front();
        if (a.end <= b || a.begin >= e) break;
        CU_TRY(cudaStreamWaitEvent(s->up_stream, a.ev_ingested, 0));
        s->stage_ev_pool.push_back(a.ev_ingested);
        s->stage_live.pop_front();
    }
    cudaEvent_t ev;
    if (!s->stage_ev_pool.empty()) { ev = s->stage_ev_pool.back(); s->stage_ev_pool.pop_back(); }
    else CU_TRY(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
    s->stage_live.push_back({b, e, ev});
    *dst = s->stage_slab + b;
    *ev_ingested = ev;
    return SVTME_OK;
}

// Pyramids are built on the compute stream of the NEXT batch (the one that will read them). A slot is not overwritten
// before the last batch that reads it is done (it may run on the other compute stream), and a batch waits for the
// builds queued on the other stream (after_builds / launch_me_batch).
static uint32_t before_build(SvtMeSession *s, cudaStream_t cs, int32_t pic_id) {
    if (s->concurrency == 2 && s->slot_last_seq[pic_id]) CU_TRY(cudaStreamWaitEvent(cs, s->ev_me[s->slot_last_seq[pic_id] % 8], 0));
    return SVTME_OK;
}
static uint32_t after_builds(SvtMeSession *s, cudaStream_t cs) {
    const int k = cs == s->stream2 ? 1 : 0;
    CU_TRY(cudaEventRecord(s->ev_built[k], cs));
    s->built_recorded[k] = true;
    return SVTME_OK;
}

// Uploads of n pictures: host frames are copied into ONE region of the staging ring, frames that are adjacent in host
// memory (a decoded clip, the frames of a mini-GOP in a capture buffer) as one copy, whatever the order they were named
// in; then the pyramids of all n are built by two launches.
static uint32_t upload_batch(SvtMeSession *s, uint32_t n, const int32_t *pic_ids, const void *const *src, size_t row_bytes,
                             size_t host_pitch, int mode, bool host) {
    if (!s || !pic_ids || !src || n == 0 || n > SVTME_STAGE_FRAMES) return SVTME_ERR_BAD_PARAMETER;
    for (uint32_t i = 0; i < n; i++)
        if (pic_ids[i] < 0 || pic_ids[i] >= s->max_pics || !src[i]) return SVTME_ERR_BAD_PARAMETER;
    std::lock_guard<std::mutex> g(s->mu);
    CU_TRY(cudaSetDevice(s->device));
    std::vector<DevPic>       pics(n);
    std::vector<const void *> in(n);
    cudaStream_t cs = compute_stream(s, (int)((s->batch_seq + 1) & 1));
    cudaEvent_t ev_ingested = nullptr;
    if (host) {
        const size_t frame_bytes = row_bytes * (size_t)s->h;
        uint8_t *region = nullptr;
        uint32_t rc = stage_alloc(s, frame_bytes * n, &region, &ev_ingested);
        if (rc != SVTME_OK) return rc;
        std::vector<uint32_t> order(n);
        for (uint32_t i = 0; i < n; i++) order[i] = i;
        std::sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) { return (uintptr_t)src[a] < (uintptr_t)src[b]; });
        for (uint32_t k = 0; k < n;) {
            uint32_t run = 1;
            if (host_pitch == row_bytes)
                while (k + run < n && (const uint8_t *)src[order[k + run]] == (const uint8_t *)src[order[k]] + frame_bytes * run) run++;
            uint8_t *dst = region + frame_bytes * k;
            if (host_pitch == row_bytes)
                CU_TRY(cudaMemcpyAsync(dst, src[order[k]], frame_bytes * run, cudaMemcpyHostToDevice, s->up_stream));
            else
                CU_TRY(cudaMemcpy2DAsync(dst, row_bytes, src[order[k]], host_pitch, row_bytes, (size_t)s->h, cudaMemcpyHostToDevice, s->up_stream));
            for (uint32_t j = 0; j < run; j++) in[order[k + j]] = region + frame_bytes * (k + j);
            k += run;
        }
        CU_TRY(cudaEventRecord(s->ev_copied, s->up_stream));
        CU_TRY(cudaStreamWaitEvent(cs, s->ev_copied, 0));
    }
    for (uint32_t i = 0; i < n; i++) {
        pics[i] = s->pics[pic_ids[i]];
        if (!host) in[i] = src[i];
        uint32_t rc = before_build(s, cs, pic_ids[i]);
        if (rc != SVTME_OK) return rc;
    }
    s->launches += (uint64_t)svtme_launch_build_pyramids(pics.data(), in.data(), s->w, (int)n, mode, cs);
    CU_TRY(cudaGetLastError());
    if (host) CU_TRY(cudaEventRecord(ev_ingested, cs));
    for (uint32_t i = 0; i < n; i++) s->pic_valid[pic_ids[i]] = 1;
    return after_builds(s, cs);
}

extern "C" uint32_t svtme_picture_upload(SvtMeSession *s, int32_t pic_id, const uint8_t *luma, uint32_t stride) {
    if (!s || !luma || stride < (uint32_t)s->w) return SVTME_ERR_BAD_PARAMETER;
    const void *p = luma;
    return upload_batch(s, 1, &pic_id, &p, (size_t)s->w, stride, 0, true);
}

extern "C" uint32_t svtme_picture_upload_device(SvtMeSession *s, int32_t pic_id, const uint8_t *d_luma, uint32_t stride) {
    if (!s || !d_luma || pic_id < 0 || pic_id >= s->max_pics || stride < (uint32_t)s->w) return SVTME_ERR_BAD_PARAMETER;
    std::lock_guard<std::mutex> g(s->mu);
    CU_TRY(cudaSetDevice(s->device));
    cudaStream_t cs = compute_stream(s, (int)((s->batch_seq + 1) & 1));
    uint32_t rc = before_build(s, cs, pic_id);
    if (rc != SVTME_OK) return rc;
    s->launches += (uint64_t)svtme_launch_build_pyramid(s->pics[pic_id], 0, d_luma, (int)stride, cs);
    CU_TRY(cudaGetLastError());
    s->pic_valid[pic_id] = 1;
    return after_builds(s, cs);
}

extern "C" uint32_t svtme_picture_upload_10bit(SvtMeSession *s, int32_t pic_id, const uint16_t *luma16, uint32_t stride_in_samples) {
    if (!s || !luma16 || stride_in_samples < (uint32_t)s->w) return SVTME_ERR_BAD_PARAMETER;
    const void *p = luma16;
    return upload_batch(s, 1, &pic_id, &p, (size_t)s->w * 2, (size_t)stride_in_samples * 2, 2, true);
}

extern "C" uint32_t svtme_picture_upload_batch(SvtMeSession *s, uint32_t n, const int32_t *pic_ids, const uint8_t *const *luma,
                                               uint32_t stride) {
    if (!s || stride < (uint32_t)s->w) return SVTME_ERR_BAD_PARAMETER;
    return upload_batch(s, n, pic_ids, (const void *const *)luma, (size_t)s->w, stride, 0, true);
}

extern "C" uint32_t svtme_picture_upload_10bit_batch(SvtMeSession *s, uint32_t n, const int32_t *pic_ids,
                                                     const uint16_t *const *luma16, uint32_t stride_in_samples) {
    if (!s || stride_in_samples < (uint32_t)s->w) return SVTME_ERR_BAD_PARAMETER;
    return upload_batch(s, n, pic_ids, (const void *const *)luma16, (size_t)s->w * 2, (size_t)stride_in_samples * 2, 2, true);
}

extern "C" uint32_t svtme_picture_upload_device_batch(SvtMeSession *s, uint32_t n, const int32_t *pic_ids,
                                                      const uint8_t *const *d_luma, uint32_t stride) {
    if (!s || stride != (uint32_t)s->w) return SVTME_ERR_BAD_PARAMETER; // dense device frames only
    return upload_batch(s, n, pic_ids, (const void *const *)d_luma, 0, 0, 0, false);
}

extern "C" uint32_t svtme_picture_download_plane(SvtMeSession *s, int32_t pic_id, int level, uint8_t *dst, size_t dst_capacity,
                                                 uint32_t *stride, uint32_t *rows, uint32_t *org_x, uint32_t *org_y) {
    if (!s || !dst || pic_id < 0 || pic_id >= s->max_pics || level < 0 || level > 2 || !s->pic_valid[pic_id]) return SVTME_ERR_BAD_PARAMETER;
    std::lock_guard<std::mutex> g(s->mu);
    CU_TRY(cudaSetDevice(s->device));
    const DevPlane &p = s->pics[pic_id].lvl[level];
    // exported with the padding of the reference's buffers (72 / 32 / 16); the store may pad wider for alignment
    static const int export_pad[3] = {72, SVTME_QUARTER_PAD, SVTME_SIXTEENTH_PAD};
    const int ep = export_pad[level], skip = p.org_x - ep;
    const size_t W = (size_t)p.width + 2 * ep, H = (size_t)p.height + 2 * ep;
    if (dst_capacity < W * H || skip < 0) return SVTME_ERR_BAD_PARAMETER;
    CU_TRY(cudaStreamSynchronize(s->up_stream));
    CU_TRY(cudaStreamSynchronize(s->stream2));
    CU_TRY(cudaMemcpy2DAsync(dst, W, p.buf + (size_t)skip * p.stride + skip, (size_t)p.stride, W, H, cudaMemcpyDeviceToHost, s->stream));
    CU_TRY(cudaStreamSynchronize(s->stream));
    if (stride) *stride = (uint32_t)W;
    if (rows) *rows = (uint32_t)H;
    if (org_x) *org_x = (uint32_t)ep;
    if (org_y) *org_y = (uint32_t)ep;
    return SVTME_OK;
}

static uint32_t validate_params(const SvtMeSession *s, int32_t pic_id, const SvtMePicParams *pp) {
    if (!s || !pp || pic_id < 0 || pic_id >= s->max_pics || !s->pic_valid[pic_id]) return SVTME_ERR_BAD_PARAMETER;
    if (pp->num_of_list_to_search < 1 || pp->num_of_list_to_search > 2) return SVTME_ERR_BAD_PARAMETER;
    if (!pp->enable_me_16x16) return SVTME_ERR_BAD_PARAMETER; // svt_aom_get_enable_me_16x16 is constant 1
    if (pp->ctrls.num_hme_sa_w != 2 || pp->ctrls.num_hme_sa_h != 2) return SVTME_ERR_BAD_PARAMETER; // enc_mode_config.c:146-147
    if (pp->max_refs > 8 || pp->max_cand > 32) return SVTME_ERR_BAD_PARAMETER;
    for (int l = 0; l < pp->num_of_list_to_search; l++) {
        if (pp->num_of_ref_pic_to_search[l] > SVTME_MAX_REF) return SVTME_ERR_BAD_PARAMETER;
        for (int r = 0; r < pp->num_of_ref_pic_to_search[l]; r++) {
            const int id = pp->ref_pic_id[l][r];
            if (id < 0 || id >= s->max_pics || !s->pic_valid[id]) return SVTME_ERR_BAD_PARAMETER;
        }
    }
    if (pp->num_of_ref_pic_to_search[0] < 1) return SVTME_ERR_BAD_PARAMETER;
    // the MeSbResults arrays are sized by max_l0 / max_refs (pcs.c:91-96): more refs than that would alias entries
    if (pp->num_of_ref_pic_to_search[0] > pp->max_l0 || pp->max_l0 > pp->max_refs) return SVTME_ERR_BAD_PARAMETER;
    if (pp->num_of_list_to_search == 2 && pp->num_of_ref_pic_to_search[1] > pp->max_refs - pp->max_l0) return SVTME_ERR_BAD_PARAMETER;
    { // window sizes are int16 throughout, as in the reference; a full-pel window is cut into at most 65535 tiles
        long long mw = pp->ctrls.me_sa.sa_max.width, mh = pp->ctrls.me_sa.sa_max.height;
        if (pp->ctrls.mv_sa_adj_enabled) { mw *= pp->ctrls.mv_sa_adj_sa_multiplier; mh *= pp->ctrls.mv_sa_adj_sa_multiplier; }
        mw = mw * 3 / 2 + 8; mh = mh * 3 / 2;
        if (mw > 16384 || mh > 16384 || ((mw + 15) / 16) * ((mh + 7) / 8) > 65535) return SVTME_ERR_BAD_PARAMETER;
    }
    return SVTME_OK;
}

static void fold_timelines(SvtMeSession *s) {
    for (SvtMeTimeline &t : s->tls) {
        for (int i = 1; i < t.n; i++) {
            float ms = 0;
            if (cudaEventElapsedTime(&ms, t.ev[i - 1], t.ev[i]) == cudaSuccess) s->acc_tag_ms[t.tag[i]] += ms;
        }
        for (int i = 0; i < SVTME_TL_MAX; i++) s->tl_pool.push_back(t.ev[i]);
    }
    s->tls.clear();
}
static void fold_events(SvtMeSession *s) {
    fold_timelines(s);
    for (size_t i = 0; i + 3 <= s->ev_used; i += 3) {
        float a = 0, b = 0;
        if (cudaEventElapsedTime(&a, s->ev[i], s->ev[i + 1]) == cudaSuccess &&
            cudaEventElapsedTime(&b, s->ev[i + 1], s->ev[i + 2]) == cudaSuccess) {
            s->acc_hme_ms += a;
            s->acc_fp_ms += b;
            s->n_timed++;
        }
    }
    s->ev_used = 0;
}

static uint32_t ensure_batch(SvtMeSession *s, uint32_t n) {
    if (n <= s->res_sets[0].size() && n <= s->res_sets[1].size() && n <= s->d_jobs_n) return SVTME_OK;
    uint32_t rc = sync_all(s);
    if (rc != SVTME_OK) return rc;
    for (auto &set : s->res_sets)
        while (set.size() < n) {
            uint8_t *p = nullptr;
            cudaError_t e = cudaMalloc(&p, s->lay.end);
            if (e != cudaSuccess) { fprintf(stderr, "[svtme_b200] cudaMalloc result slab: %s\n", cudaGetErrorString(e)); return SVTME_ERR_INSUFFICIENT_RESOURCES; }
            cudaMemset(p, 0, s->lay.end);
            set.push_back(p);
        }
    if (s->d_jobs) cudaFree(s->d_jobs);
    s->d_jobs = nullptr;
    cudaError_t e = cudaMalloc(&s->d_jobs, sizeof(PicJob) * n * 2); // one half per result set
    if (e != cudaSuccess) return SVTME_ERR_INSUFFICIENT_RESOURCES;
    if (s->h_jobs_ring) cudaFreeHost(s->h_jobs_ring);
    s->h_jobs_ring = nullptr;
    if (cudaMallocHost(&s->h_jobs_ring, sizeof(PicJob) * n * 4) != cudaSuccess) return SVTME_ERR_INSUFFICIENT_RESOURCES;
    for (auto &u : s->jobs_used) u = false;
    s->d_jobs_n = n;
    return SVTME_OK;
}

// a pinned job array nobody is copying from any more
static uint32_t next_job_array(SvtMeSession *s, PicJob **out, int *slot) {
    const int k = (int)(s->jobs_head++ & 3);
    if (s->jobs_used[k]) CU_TRY(cudaEventSynchronize(s->ev_jobs[k])); // four batches back: long done
    *out = s->h_jobs_ring + (size_t)k * s->d_jobs_n;
    *slot = k;
    return SVTME_OK;
}

static uint32_t ensure_units(SvtMeSession *s, uint64_t need_units, uint64_t need_s16) {
    if (need_units > 0x7fffffffull || need_s16 > 0x7fffffffull) return SVTME_ERR_INSUFFICIENT_RESOURCES;
    if (need_units > s->fp_units_cap) { // the unit queue grows with the largest search window of the controls in use
        uint32_t rcs = sync_all(s);
        if (rcs != SVTME_OK) return rcs;
        if (s->fp_units) cudaFree(s->fp_units);
        s->fp_units = nullptr; s->fp_units_cap = 0;
        cudaError_t e = cudaMalloc(&s->fp_units, (size_t)need_units * 2 * sizeof(uint2)); // one half per result set
        if (e != cudaSuccess) { fprintf(stderr, "[svtme_b200] cudaMalloc unit queue (%llu units): %s\n", (unsigned long long)need_units, cudaGetErrorString(e)); return SVTME_ERR_INSUFFICIENT_RESOURCES; }
        s->fp_units_cap = (uint32_t)need_units;
    }
    if (need_s16 > s->s16_units_cap) {
        uint32_t rcs = sync_all(s);
        if (rcs != SVTME_OK) return rcs;
        if (s->s16_units) cudaFree(s->s16_units);
        s->s16_units = nullptr; s->s16_units_cap = 0;
        cudaError_t e = cudaMalloc(&s->s16_units, (size_t)need_s16 * 2 * sizeof(uint2));
        if (e != cudaSuccess) { fprintf(stderr, "[svtme_b200] cudaMalloc HME unit queue (%llu units): %s\n", (unsigned long long)need_s16, cudaGetErrorString(e)); return SVTME_ERR_INSUFFICIENT_RESOURCES; }
        s->s16_units_cap = (uint32_t)need_s16;
    }
    return SVTME_OK;
}

static uint32_t ensure_download_region(SvtMeSession *s, int set, size_t bytes) {
    if (bytes <= s->dl_cap[set]) return SVTME_OK;
    uint32_t rc = sync_all(s);
    if (rc != SVTME_OK) return rc;
    if (s->dl_region[set]) cudaFree(s->dl_region[set]);
    s->dl_region[set] = nullptr; s->dl_cap[set] = 0;
    cudaError_t e = cudaMalloc(&s->dl_region[set], bytes);
    if (e != cudaSuccess) { fprintf(stderr, "[svtme_b200] cudaMalloc result region (%zu bytes): %s\n", bytes, cudaGetErrorString(e)); return SVTME_ERR_INSUFFICIENT_RESOURCES; }
    cudaMemset(s->dl_region[set], 0, bytes);
    s->dl_cap[set] = bytes;
    return SVTME_OK;
}

extern "C" uint32_t svtme_session_reserve(SvtMeSession *s, const SvtMePicParams *params, uint32_t n_pictures) {
    if (!s || !params || n_pictures == 0 || n_pictures > 4096) return SVTME_ERR_BAD_PARAMETER;
    std::lock_guard<std::mutex> g(s->mu);
    CU_TRY(cudaSetDevice(s->device));
    uint32_t rc = ensure_batch(s, n_pictures);
    if (rc != SVTME_OK) return rc;
    rc = ensure_units(s, (uint64_t)svtme_fp_units_cap(*params, s->n_b64) * n_pictures, (uint64_t)svtme_s16_units_cap(*params, s->n_b64) * n_pictures);
    if (rc != SVTME_OK) return rc;
    SvtMePackedLayout L;
    packed_layout((size_t)s->n_b64, 85, 8, 32, SVTME_F_ALL, &L); // the largest block any picture of this size can have
    for (int set = 0; set < 2; set++) {
        rc = ensure_download_region(s, set, L.bytes * n_pictures);
        if (rc != SVTME_OK) return rc;
    }
    return SVTME_OK;
}

extern "C" uint32_t svtme_session_set_unit_granularity(SvtMeSession *s, int mode) {
    if (!s || mode < -1 || mode > 1) return SVTME_ERR_BAD_PARAMETER;
    std::lock_guard<std::mutex> g(s->mu);
    s->s16_unit_mode = mode;
    return SVTME_OK;
}

extern "C" uint32_t svtme_session_set_result_fields(SvtMeSession *s, uint32_t fields) {
    if (!s || (fields & ~(SVTME_F_ALL | SVTME_F_SB_COMPACT))) return SVTME_ERR_BAD_PARAMETER;
    std::lock_guard<std::mutex> g(s->mu);
    s->default_fields = fields;
    return SVTME_OK;
}

extern "C" uint32_t svtme_packed_layout(uint32_t width, uint32_t height, const SvtMePicParams *params, uint32_t fields,
                                        SvtMePackedLayout *out) {
    if (!params || !out || !width || !height || (fields & ~(SVTME_F_ALL | SVTME_F_SB_COMPACT)) || params->max_refs > 8 || params->max_cand > 32) return SVTME_ERR_BAD_PARAMETER;
    const size_t n_b64 = (size_t)((width + 63) / 64) * ((height + 63) / 64);
    SvtMeSbStrides st;
    svtme_sb_strides(params, fields, &st);
    packed_layout(n_b64, st.n_pus, st.mv_stride, st.cand_stride, fields, out);
    return SVTME_OK;
}

extern "C" void svtme_packed_view(const SvtMePackedLayout *layout, void *block, SvtMePicResults *out) {
    for (int f = 0; f < RF_COUNT; f++)
        *result_member(out, f) = layout->off[f] == (size_t)-1 ? nullptr : (void *)((uint8_t *)block + layout->off[f]);
}

static uint32_t fields_of_results(const SvtMePicResults *r) {
    uint32_t m = 0;
    for (int f = 0; f < RF_COUNT; f++)
        if (*result_member(const_cast<SvtMePicResults *>(r), f)) m |= rf_group[f];
    return m;
}

// fields[i]: what picture i must produce (SVTME_F_*)
// reciprocals for div_by() (me_device.cuh): b64 index -> row, compact unit index -> b64
static void picjob_set_inverses(PicJob &job) {
    auto inv = [](int d) -> uint32_t { return d > 1 ? (uint32_t)(((1ull << 32) + (uint64_t)d - 1) / (uint64_t)d) : 0u; };
    const SvtMePicParams &pp = job.pp;
    const int na = pp.num_of_ref_pic_to_search[0] + (pp.num_of_list_to_search > 1 ? pp.num_of_ref_pic_to_search[1] : 0);
    job.inv_cols = inv(job.b64_cols);
    job.inv_refs = (uint64_t)job.n_b64 * 8 * (uint64_t)(na > 0 ? na : 1) < (1ull << 32) ? inv(na) : 0u;
}

static uint32_t launch_me_batch(SvtMeSession *s, uint32_t n, const int32_t *pic_ids, const SvtMePicParams *params, const uint32_t *fields) {
    if (!s || !pic_ids || !params || n == 0 || n > 4096) return SVTME_ERR_BAD_PARAMETER;
    CU_TRY(cudaSetDevice(s->device));
    uint64_t need_units = 0, need_s16 = 0;
    for (uint32_t i = 0; i < n; i++) {
        uint32_t rc = validate_params(s, pic_ids[i], &params[i]);
        if (rc != SVTME_OK) return rc;
        // one launch sequence serves the batch: the stage structure must agree (it does inside one encode)
        const SvtMeCtrls &a = params[0].ctrls, &b = params[i].ctrls;
        if (a.enable_hme_flag != b.enable_hme_flag || a.enable_hme_level0_flag != b.enable_hme_level0_flag ||
            a.enable_hme_level1_flag != b.enable_hme_level1_flag || a.enable_hme_level2_flag != b.enable_hme_level2_flag ||
            a.prehme_enable != b.prehme_enable || a.hme_search_method != b.hme_search_method ||
            a.me_search_method != b.me_search_method || a.prehme_skip_search_line != b.prehme_skip_search_line || a.me_type != b.me_type ||
            a.hme_l1_sa.width != b.hme_l1_sa.width || a.hme_l1_sa.height != b.hme_l1_sa.height ||
            a.hme_l2_sa.width != b.hme_l2_sa.width || a.hme_l2_sa.height != b.hme_l2_sa.height)
            return SVTME_ERR_BAD_PARAMETER;
        need_units += svtme_fp_units_cap(params[i], s->n_b64);
        need_s16 += svtme_s16_units_cap(params[i], s->n_b64);
    }
    uint32_t rc = ensure_batch(s, n);
    if (rc != SVTME_OK) return rc;
    rc = ensure_units(s, need_units, need_s16);
    if (rc != SVTME_OK) return rc;
    const uint64_t seq = s->batch_seq + 1;
    const int set = (int)(seq & 1);
    SvtMeSession::BatchOut &bo = s->out[set];
    std::vector<SvtMePackedLayout> lay(n);
    std::vector<size_t> base(n);
    size_t total = 0;
    std::vector<SvtMeSbStrides> strides(n);
    for (uint32_t i = 0; i < n; i++) {
        svtme_sb_strides(&params[i], fields[i], &strides[i]);
        packed_layout((size_t)s->n_b64, strides[i].n_pus, strides[i].mv_stride, strides[i].cand_stride, fields[i], &lay[i]);
        base[i] = total;
        total += lay[i].bytes;
    }
    rc = ensure_download_region(s, set, total);
    if (rc != SVTME_OK) return rc;
    s->batch_seq = seq;
    bo.seq = seq; bo.me = true; bo.lay.swap(lay); bo.base.swap(base); bo.total = total;
    bo.geo.resize(n);
    for (uint32_t i = 0; i < n; i++) bo.geo[i] = {(size_t)strides[i].n_pus, (size_t)strides[i].mv_stride, (size_t)strides[i].cand_stride};
    cudaStream_t cs = compute_stream(s, set);
    // the set's previous results have left for the host and its device-side readers are done: waited for inside the launch
    // sequence, before the first kernel that writes the region (svtme_launch_picture_me)
    cudaEvent_t out_free[2];
    int n_out_free = 0;
    if (s->dl_recorded[set]) out_free[n_out_free++] = s->ev_dl[set];
    if (s->release_recorded[set]) { out_free[n_out_free++] = s->ev_release[set]; s->release_recorded[set] = false; }
    if (s->concurrency == 2) { // pictures whose pyramids were queued on the other compute stream
        const int other = 1 - stream_index(s, set);
        if (s->built_recorded[other]) CU_TRY(cudaStreamWaitEvent(cs, s->ev_built[other], 0));
    }
    PicJob *d_jobs = s->d_jobs + (size_t)set * s->d_jobs_n;
    PicJob *h_jobs = nullptr;
    int jslot = 0;
    rc = next_job_array(s, &h_jobs, &jslot);
    if (rc != SVTME_OK) return rc;
    for (uint32_t i = 0; i < n; i++) {
        const SvtMePicParams *pp = &params[i];
        PicJob &job = h_jobs[i];
        memset(&job, 0, sizeof(job));
        job.pp  = *pp;
        job.src = s->pics[pic_ids[i]];
        s->slot_last_seq[pic_ids[i]] = seq;
        for (int l = 0; l < pp->num_of_list_to_search; l++)
            for (int r = 0; r < pp->num_of_ref_pic_to_search[l]; r++) {
                job.ref[l][r] = s->pics[pp->ref_pic_id[l][r]];
                s->slot_last_seq[pp->ref_pic_id[l][r]] = seq;
            }
        job.pic_w = s->w; job.pic_h = s->h; job.b64_cols = s->b64_cols; job.b64_rows = s->b64_rows; job.n_b64 = s->n_b64;
        job.n_pus = pp->enable_me_8x8 ? 85 : 21;
        picjob_set_inverses(job);
        job.mv_stride = strides[i].mv_stride; job.mv_l1_off = strides[i].mv_l1_offset; job.cand_stride = strides[i].cand_stride;
        uint8_t *R = s->res_sets[set][i];
        uint8_t *D = s->dl_region[set] + bo.base[i];
        const SvtMePackedLayout &L = bo.lay[i];
        auto at = [&](int f) -> uint8_t * { return L.off[f] == (size_t)-1 ? nullptr : D + L.off[f]; }; // nullptr: not produced
        job.sb_state       = (SbState *)(R + s->lay.sb_state);
        job.best_sad       = (uint32_t *)at(RF_BEST_SAD);
        job.best_mv        = (uint32_t *)at(RF_BEST_MV);
        job.windows        = (SvtMeWindow *)at(RF_WINDOWS);
        job.search_results = (SvtMeSearchResult *)at(RF_SEARCH);
        job.total_me_candidate_index = at(RF_TOTAL);
        job.me_mv_array              = (uint32_t *)at(RF_MV);
        job.me_candidate_array       = at(RF_CAND);
        job.me_64x64_distortion = (uint32_t *)at(RF_D64);
        job.me_32x32_distortion = (uint32_t *)at(RF_D32);
        job.me_16x16_distortion = (uint32_t *)at(RF_D16);
        job.me_8x8_distortion   = (uint32_t *)at(RF_D8);
        job.rc_me_distortion    = (uint32_t *)at(RF_RC);
        job.me_8x8_cost_variance = (uint32_t *)at(RF_VAR);
        job.stationary_block_present_sb = at(RF_STAT);
        job.rc_me_allow_gm              = at(RF_GM);
        job.hme_jobs     = (HmeJobG *)(R + s->lay.hme_jobs);
        job.fp_jobs      = (FpJobG *)(R + s->lay.fp_jobs);
        job.fp_keys      = (unsigned long long *)(R + s->lay.fp_keys);
        // shared by the batch (slab 0)
        job.sad_ops      = (unsigned long long *)(s->res_slab + s->lay.ops);
        // per result set: counters in the set's first slab, one half of each unit queue
        job.fp_counters  = (uint32_t *)(s->res_sets[set][0] + s->lay.fp_counters);
        job.fp_units     = s->fp_units + (size_t)set * s->fp_units_cap;
        job.fp_units_cap = s->fp_units_cap;
        job.s16_units     = s->s16_units + (size_t)set * s->s16_units_cap;
        job.s16_counters  = (uint32_t *)(s->res_sets[set][0] + s->lay.fp_counters) + 4;
        job.s16_units_cap = s->s16_units_cap;
        // units of k_search16 = (b64, list, ref) groups when the batch has enough of them to occupy the GPU several times over
        // (~3000 resident warps), single jobs otherwise
        job.s16_split = s->s16_unit_mode >= 0 ? s->s16_unit_mode : ((size_t)n * s->n_b64 * 8 < 24000 ? 1 : 0);
    }
    svtme_launch_copy_from_pinned(d_jobs, h_jobs, sizeof(PicJob) * n, cs); // a kernel, not a copy-engine transfer (me_kernels.cu)
    CU_TRY(cudaEventRecord(s->ev_jobs[jslot], cs));
    s->jobs_used[jslot] = true;
    cudaEvent_t e0 = nullptr, e1 = nullptr, e2 = nullptr;
    if (s->timing) {
        if (s->ev_used + 3 > s->ev.size()) {
            if (s->ev.size() >= 3 * 2048) {
                CU_TRY(cudaStreamSynchronize(s->stream));
                CU_TRY(cudaStreamSynchronize(s->stream2));
                fold_events(s);
            } else {
                for (int i = 0; i < 3; i++) {
                    cudaEvent_t e;
                    CU_TRY(cudaEventCreate(&e));
                    s->ev.push_back(e);
                }
            }
        }
        e0 = s->ev[s->ev_used]; e1 = s->ev[s->ev_used + 1]; e2 = s->ev[s->ev_used + 2];
        s->ev_used += 3;
    }
    SvtMeTimeline *tl = nullptr;
    if (s->timing) {
        if (s->tls.size() >= 256) {
            CU_TRY(cudaStreamSynchronize(s->stream));
            fold_timelines(s);
        }
        s->tls.emplace_back();
        tl = &s->tls.back();
        tl->n = 0;
        for (int i = 0; i < SVTME_TL_MAX; i++) {
            if (!s->tl_pool.empty()) { tl->ev[i] = s->tl_pool.back(); s->tl_pool.pop_back(); }
            else CU_TRY(cudaEventCreate(&tl->ev[i]));
        }
    }
    s->launches += (uint64_t)svtme_launch_picture_me(h_jobs, (int)n, d_jobs, cs, e0, e1, e2, tl, out_free, n_out_free);
    CU_TRY(cudaGetLastError());
    CU_TRY(cudaEventRecord(s->ev_me[seq % 8], cs));
    return SVTME_OK;
}

// the download stream waits for the kernels of the batch that owns `set`; afterwards: the set's overflow flag, the event
static uint32_t begin_download(SvtMeSession *s, int set) {
    CU_TRY(cudaStreamWaitEvent(s->dl_stream, s->ev_me[s->out[set].seq % 8], 0));
    return SVTME_OK;
}
static uint32_t end_download(SvtMeSession *s, int set) {
    CU_TRY(cudaMemcpyAsync(&s->h_flags[set], (uint32_t *)(s->res_sets[set][0] + s->lay.fp_counters) + 8, sizeof(uint32_t), cudaMemcpyDeviceToHost, s->dl_stream));
    CU_TRY(cudaEventRecord(s->ev_dl[set], s->dl_stream));
    s->dl_recorded[set] = true;
    const uint64_t seq = s->out[set].seq;
    CU_TRY(cudaEventRecord(s->ev_dl_seq[seq % 8], s->dl_stream));
    s->dl_seq[seq % 8] = seq;
    return SVTME_OK;
}
static uint32_t finish_sync(SvtMeSession *s, int set, bool downloaded) {
    if (downloaded) {
        CU_TRY(cudaStreamSynchronize(s->dl_stream)); // the download waited for the batch's kernels
        return check_downloaded_flag(s, set);
    }
    CU_TRY(cudaStreamSynchronize(compute_stream(s, set)));
    return SVTME_OK;
}

// per-field copies of picture `index` of the batch that owns `set` into the caller's arrays (between begin_/end_download)
static uint32_t fetch_fields(SvtMeSession *s, int set, uint32_t index, SvtMePicResults *res) {
    const SvtMeSession::BatchOut &bo = s->out[set];
    if (!bo.me || !res || index >= bo.lay.size()) return SVTME_ERR_BAD_PARAMETER;
    const SvtMePackedLayout &L = bo.lay[index];
    const uint8_t *D = s->dl_region[set] + bo.base[index];
    const SvtMeSession::BatchOut::Geo &g = bo.geo[index];
    for (int f = 0; f < RF_COUNT; f++)
        if (*result_member(res, f) && L.off[f] == (size_t)-1) return SVTME_ERR_BAD_PARAMETER; // the batch was not asked to produce this field
    for (int f = 0; f < RF_COUNT; f++) {
        void *dst = *result_member(res, f);
        if (dst) CU_TRY(cudaMemcpyAsync(dst, D + L.off[f], rf_bytes(f, (size_t)s->n_b64, g.n_pus, g.max_refs, g.max_cand), cudaMemcpyDeviceToHost, s->dl_stream));
    }
    return SVTME_OK;
}

extern "C" uint32_t svtme_picture_me_batch(SvtMeSession *s, uint32_t n, const int32_t *pic_ids, const SvtMePicParams *params,
                                           SvtMePicResults *results, int sync) {
    if (!s || n == 0 || n > 4096) return SVTME_ERR_BAD_PARAMETER;
    std::lock_guard<std::mutex> g(s->mu);
    std::vector<uint32_t> fields(n, s->default_fields);
    if (results)
        for (uint32_t i = 0; i < n; i++) fields[i] = fields_of_results(&results[i]);
    uint32_t rc = launch_me_batch(s, n, pic_ids, params, fields.data());
    if (rc != SVTME_OK) return rc;
    const int set = (int)(s->batch_seq & 1);
    if (results) {
        rc = begin_download(s, set);
        for (uint32_t i = 0; i < n && rc == SVTME_OK; i++) rc = fetch_fields(s, set, i, &results[i]);
        if (rc != SVTME_OK) return rc;
        rc = end_download(s, set);
        if (rc != SVTME_OK) return rc;
    }
    return sync ? finish_sync(s, set, results != nullptr) : SVTME_OK;
}

extern "C" uint32_t svtme_picture_me_batch_packed(SvtMeSession *s, uint32_t n, const int32_t *pic_ids, const SvtMePicParams *params,
                                                  uint32_t fields, void *host_slab, size_t host_slab_bytes, int sync) {
    if (!s || n == 0 || n > 4096 || !host_slab || !(fields & SVTME_F_ALL) || (fields & ~(SVTME_F_ALL | SVTME_F_SB_COMPACT))) return SVTME_ERR_BAD_PARAMETER;
    std::lock_guard<std::mutex> g(s->mu);
    { // the slab must hold the batch before anything is queued
        size_t total = 0;
        SvtMePackedLayout L;
        for (uint32_t i = 0; i < n; i++) {
            if (params[i].max_refs > 8 || params[i].max_cand > 32) return SVTME_ERR_BAD_PARAMETER;
            SvtMeSbStrides st;
            svtme_sb_strides(&params[i], fields, &st);
            packed_layout((size_t)s->n_b64, st.n_pus, st.mv_stride, st.cand_stride, fields, &L);
            total += L.bytes;
        }
        if (total > host_slab_bytes) return SVTME_ERR_BAD_PARAMETER;
    }
    std::vector<uint32_t> f(n, fields);
    uint32_t rc = launch_me_batch(s, n, pic_ids, params, f.data());
    if (rc != SVTME_OK) return rc;
    const int set = (int)(s->batch_seq & 1);
    rc = begin_download(s, set);
    if (rc != SVTME_OK) return rc;
    CU_TRY(cudaMemcpyAsync(host_slab, s->dl_region[set], s->out[set].total, cudaMemcpyDeviceToHost, s->dl_stream)); // the whole batch: ONE copy
    rc = end_download(s, set);
    if (rc != SVTME_OK) return rc;
    return sync ? finish_sync(s, set, true) : SVTME_OK;
}

extern "C" uint32_t svtme_fetch_results_batch(SvtMeSession *s, uint32_t index, SvtMePicResults *results, int sync) {
    if (!s) return SVTME_ERR_BAD_PARAMETER;
    std::lock_guard<std::mutex> g(s->mu);
    CU_TRY(cudaSetDevice(s->device));
    const int set = (int)(s->batch_seq & 1);
    if (s->out[set].seq != s->batch_seq) return SVTME_ERR_BAD_PARAMETER;
    uint32_t rc = begin_download(s, set);
    if (rc == SVTME_OK) rc = fetch_fields(s, set, index, results);
    if (rc == SVTME_OK) rc = end_download(s, set);
    if (rc != SVTME_OK) return rc;
    return sync ? finish_sync(s, set, true) : SVTME_OK;
}

extern "C" uint32_t svtme_picture_me_device(SvtMeSession *s, int32_t pic_id, const SvtMePicParams *params) {
    if (!s) return SVTME_ERR_BAD_PARAMETER;
    std::lock_guard<std::mutex> g(s->mu);
    return launch_me_batch(s, 1, &pic_id, params, &s->default_fields);
}

extern "C" uint32_t svtme_fetch_results(SvtMeSession *s, SvtMePicResults *results, int sync) {
    return svtme_fetch_results_batch(s, 0, results, sync);
}

extern "C" uint32_t svtme_picture_me(SvtMeSession *s, int32_t pic_id, const SvtMePicParams *params, SvtMePicResults *results,
                                     int sync) {
    return svtme_picture_me_batch(s, 1, &pic_id, params, results, sync);
}

// ---- device-resident view ----------------------------------------------------
extern "C" uint32_t svtme_picture_device_planes(SvtMeSession *s, int32_t pic_id, SvtMeDevicePlane planes[3]) {
    if (!s || !planes || pic_id < 0 || pic_id >= s->max_pics) return SVTME_ERR_BAD_PARAMETER;
    std::lock_guard<std::mutex> g(s->mu);
    if (!s->pic_valid[pic_id]) return SVTME_ERR_BAD_PARAMETER;
    for (int i = 0; i < 3; i++) {
        const DevPlane &p = s->pics[pic_id].lvl[i];
        planes[i].origin = p.buf + (size_t)p.org_y * p.stride + p.org_x;
        planes[i].stride = (uint32_t)p.stride; planes[i].width = (uint32_t)p.width; planes[i].height = (uint32_t)p.height;
        planes[i].pad = (uint32_t)p.org_x;
    }
    return SVTME_OK;
}

extern "C" uint32_t svtme_batch_results_acquire(SvtMeSession *s, uint64_t seq, uint32_t index, SvtMePicResults *device_ptrs,
                                                void *consumer_stream) {
    if (!s || !device_ptrs || seq == 0) return SVTME_ERR_BAD_PARAMETER;
    std::lock_guard<std::mutex> g(s->mu);
    CU_TRY(cudaSetDevice(s->device));
    const int set = (int)(seq & 1);
    const SvtMeSession::BatchOut &bo = s->out[set];
    if (bo.seq != seq || !bo.me || index >= bo.lay.size()) return SVTME_ERR_BAD_PARAMETER; // a later batch owns the set by now
    svtme_packed_view(&bo.lay[index], s->dl_region[set] + bo.base[index], device_ptrs);
    CU_TRY(cudaStreamWaitEvent((cudaStream_t)consumer_stream, s->ev_me[seq % 8], 0));
    return SVTME_OK;
}

extern "C" uint32_t svtme_batch_results_release(SvtMeSession *s, uint64_t seq, void *consumer_stream) {
    if (!s || seq == 0) return SVTME_ERR_BAD_PARAMETER;
    std::lock_guard<std::mutex> g(s->mu);
    CU_TRY(cudaSetDevice(s->device));
    const int set = (int)(seq & 1);
    if (s->out[set].seq != seq) return SVTME_ERR_BAD_PARAMETER;
    CU_TRY(cudaEventRecord(s->ev_release[set], (cudaStream_t)consumer_stream));
    s->release_recorded[set] = true;
    return SVTME_OK;
}

extern "C" void *svtme_host_alloc(size_t bytes) {
    void *p = nullptr;
    if (cudaMallocHost(&p, bytes ? bytes : 1) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    return p;
}
extern "C" void svtme_host_free(void *p) {
    if (p) cudaFreeHost(p);
}

extern "C" uint32_t svtme_copy_from_device(SvtMeSession *s, void *dst_host, const void *src_device, size_t bytes) {
    if (!s || !dst_host || !src_device) return SVTME_ERR_BAD_PARAMETER;
    CU_TRY(cudaSetDevice(s->device));
    CU_TRY(cudaMemcpy(dst_host, src_device, bytes, cudaMemcpyDeviceToHost));
    return SVTME_OK;
}

// dg_detector_hme_level0 (pd_process.c:490-589) for one (source, reference) pair of uploaded pictures
extern "C" uint32_t svtme_dg_detector_hme(SvtMeSession *s, int32_t src_pic_id, int32_t ref_pic_id, SvtMeDgMetrics *metrics,
                                          uint64_t *b64_sad, int16_t *b64_mv) {
    if (!s || !metrics || src_pic_id < 0 || src_pic_id >= s->max_pics || ref_pic_id < 0 || ref_pic_id >= s->max_pics ||
        !s->pic_valid[src_pic_id] || !s->pic_valid[ref_pic_id])
        return SVTME_ERR_BAD_PARAMETER;
    std::lock_guard<std::mutex> g(s->mu);
    CU_TRY(cudaSetDevice(s->device));
    const int sa0 = metrics->input_resolution <= 1 ? 16 : metrics->input_resolution <= 2 ? 64 : 128; // :494-495
    uint32_t rc = ensure_batch(s, 1);
    if (rc != SVTME_OK) return rc;
    const uint32_t need = svtme_dg_units_cap(sa0, s->n_b64);
    if (need > s->s16_units_cap) {
        rc = sync_all(s);
        if (rc != SVTME_OK) return rc;
        if (s->s16_units) cudaFree(s->s16_units);
        s->s16_units = nullptr; s->s16_units_cap = 0;
        if (cudaMalloc(&s->s16_units, (size_t)need * 2 * sizeof(uint2)) != cudaSuccess) return SVTME_ERR_INSUFFICIENT_RESOURCES;
        s->s16_units_cap = need;
    }
    const uint64_t seq = ++s->batch_seq;
    const int set = (int)(seq & 1);
    cudaStream_t cs = compute_stream(s, set);
    if (s->dl_recorded[set]) CU_TRY(cudaStreamWaitEvent(cs, s->ev_dl[set], 0));
    if (s->concurrency == 2) {
        const int other = 1 - stream_index(s, set);
        if (s->built_recorded[other]) CU_TRY(cudaStreamWaitEvent(cs, s->ev_built[other], 0));
    }
    if (s->release_recorded[set]) { CU_TRY(cudaStreamWaitEvent(cs, s->ev_release[set], 0)); s->release_recorded[set] = false; }
    s->out[set].seq = seq; s->out[set].me = false; // the set no longer holds ME results
    PicJob *d_job = s->d_jobs + (size_t)set * s->d_jobs_n;
    PicJob *h_job = nullptr;
    int jslot = 0;
    rc = next_job_array(s, &h_job, &jslot);
    if (rc != SVTME_OK) return rc;
    PicJob &job = *h_job;
    memset(&job, 0, sizeof(job));
    job.pp.ctrls.hme_search_method = SVTME_FULL_SAD_SEARCH; // :501
    job.src = s->pics[src_pic_id];
    for (int l = 0; l < 2; l++) // the bands of a window occupy the job slots of all eight (list, ref) groups: all of them search this reference
        for (int r = 0; r < 4; r++) job.ref[l][r] = s->pics[ref_pic_id];
    s->slot_last_seq[src_pic_id] = s->slot_last_seq[ref_pic_id] = seq;
    job.pic_w = s->w; job.pic_h = s->h; job.b64_cols = s->b64_cols; job.b64_rows = s->b64_rows; job.n_b64 = s->n_b64;
    uint8_t *R = s->res_sets[set][0];
    job.hme_jobs      = (HmeJobG *)(R + s->lay.hme_jobs);
    job.s16_units     = s->s16_units + (size_t)set * s->s16_units_cap;
    job.s16_counters  = (uint32_t *)(R + s->lay.fp_counters) + 4;
    job.fp_counters   = (uint32_t *)(R + s->lay.fp_counters);
    job.s16_units_cap = s->s16_units_cap;
    svtme_launch_copy_from_pinned(d_job, &job, sizeof(PicJob), cs);
    CU_TRY(cudaEventRecord(s->ev_jobs[jslot], cs));
    s->jobs_used[jslot] = true;
    uint8_t *d_out = R + s->lay.ops + 32;
    unsigned long long *d_sad = (unsigned long long *)(R + s->lay.dg_sad);
    int16_t *d_mv = (int16_t *)(R + s->lay.dg_mv);
    s->launches += (uint64_t)svtme_launch_dg_detector(&job, d_job, sa0, d_out, d_sad, d_mv, cs);
    CU_TRY(cudaGetLastError());
    CU_TRY(cudaEventRecord(s->ev_me[seq % 8], cs));
    struct { unsigned long long tot_dist; unsigned int tot_cplx, tot_active; int sum_in_vectors; } out;
    CU_TRY(cudaMemcpyAsync(&out, d_out, sizeof(out), cudaMemcpyDeviceToHost, cs));
    if (b64_sad) CU_TRY(cudaMemcpyAsync(b64_sad, d_sad, (size_t)s->n_b64 * 8, cudaMemcpyDeviceToHost, cs));
    if (b64_mv) CU_TRY(cudaMemcpyAsync(b64_mv, d_mv, (size_t)s->n_b64 * 4, cudaMemcpyDeviceToHost, cs));
    CU_TRY(cudaStreamSynchronize(cs));
    metrics->tot_dist = out.tot_dist; metrics->tot_cplx = out.tot_cplx; metrics->tot_active = out.tot_active;
    metrics->sum_in_vectors = out.sum_in_vectors;
    return SVTME_OK;
}

extern "C" uint32_t svtme_session_enable_timing(SvtMeSession *s, int on) {
    if (!s) return SVTME_ERR_BAD_PARAMETER;
    std::lock_guard<std::mutex> g(s->mu);
    s->timing = on != 0;
    return SVTME_OK;
}

extern "C" uint32_t svtme_session_kernel_time(SvtMeSession *s, int reset, double *fullpel_ms, double *hme_ms, uint64_t *n_launch) {
    if (!s) return SVTME_ERR_BAD_PARAMETER;
    std::lock_guard<std::mutex> g(s->mu);
    CU_TRY(cudaSetDevice(s->device));
    CU_TRY(cudaStreamSynchronize(s->stream));
    fold_events(s);
    if (fullpel_ms) *fullpel_ms = s->acc_fp_ms;
    if (hme_ms) *hme_ms = s->acc_hme_ms;
    if (n_launch) *n_launch = s->n_timed;
    if (reset) { s->acc_fp_ms = s->acc_hme_ms = 0; s->n_timed = 0; }
    return SVTME_OK;
}

static const char *const tl_names[SVTME_TL_TAGS] = {"", "k_zz", "k_hme_ctl", "k_search16 (pre-HME)", "k_search16 (HME level 0)", "k_search_blk (HME level 1)",
                                                    "k_search_blk (HME level 2)", "k_fp", "k_fp_big", "k_post"};
extern "C" const char *svtme_kernel_name(int tag) { return tag > 0 && tag < SVTME_TL_TAGS ? tl_names[tag] : nullptr; }

extern "C" uint32_t svtme_session_kernel_times(SvtMeSession *s, int reset, double *ms, int n_ms) {
    if (!s || !ms || n_ms < 0) return SVTME_ERR_BAD_PARAMETER;
    std::lock_guard<std::mutex> g(s->mu);
    CU_TRY(cudaSetDevice(s->device));
    CU_TRY(cudaStreamSynchronize(s->stream));
    fold_timelines(s);
    for (int i = 0; i < n_ms; i++) ms[i] = i < SVTME_TL_TAGS ? s->acc_tag_ms[i] : 0.0;
    if (reset)
        for (auto &v : s->acc_tag_ms) v = 0;
    return SVTME_OK;
}

extern "C" uint32_t svtme_session_sad_ops_by_stage(SvtMeSession *s, int reset, uint64_t ops[6]) {
    if (!s || !ops) return SVTME_ERR_BAD_PARAMETER;
    std::lock_guard<std::mutex> g(s->mu);
    CU_TRY(cudaSetDevice(s->device));
    unsigned long long v[6] = {0, 0, 0, 0, 0, 0};
    CU_TRY(cudaStreamSynchronize(s->stream2));
    CU_TRY(cudaMemcpyAsync(v, s->res_slab + s->lay.ops, sizeof(v), cudaMemcpyDeviceToHost, s->stream));
    CU_TRY(cudaStreamSynchronize(s->stream));
    for (int i = 0; i < 6; i++) ops[i] = v[i];
    if (reset) CU_TRY(cudaMemsetAsync(s->res_slab + s->lay.ops, 0, sizeof(v), s->stream));
    return SVTME_OK;
}

extern "C" uint32_t svtme_fetch_hme_state(SvtMeSession *s, void *dst, size_t dst_bytes) {
    if (!s || !dst || dst_bytes < (size_t)s->n_b64 * sizeof(SbState)) return SVTME_ERR_BAD_PARAMETER;
    std::lock_guard<std::mutex> g(s->mu);
    CU_TRY(cudaSetDevice(s->device));
    cudaStream_t cs = compute_stream(s, (int)(s->batch_seq & 1));
    CU_TRY(cudaMemcpyAsync(dst, s->res_sets[s->batch_seq & 1][0] + s->lay.sb_state, (size_t)s->n_b64 * sizeof(SbState), cudaMemcpyDeviceToHost, cs));
    CU_TRY(cudaStreamSynchronize(cs));
    return SVTME_OK;
}

extern "C" uint32_t svtme_session_sad_ops(SvtMeSession *s, int reset, uint64_t *hme_ops, uint64_t *fullpel_ops) {
    if (!s) return SVTME_ERR_BAD_PARAMETER;
    std::lock_guard<std::mutex> g(s->mu);
    CU_TRY(cudaSetDevice(s->device));
    unsigned long long v[2] = {0, 0};
    CU_TRY(cudaStreamSynchronize(s->stream2));
    CU_TRY(cudaMemcpyAsync(v, s->res_slab + s->lay.ops, sizeof(v), cudaMemcpyDeviceToHost, s->stream));
    CU_TRY(cudaStreamSynchronize(s->stream));
    if (hme_ops) *hme_ops = v[0];
    if (fullpel_ops) *fullpel_ops = v[1];
    if (reset) CU_TRY(cudaMemsetAsync(s->res_slab + s->lay.ops, 0, 6 * sizeof(unsigned long long), s->stream));
    return SVTME_OK;
}

extern "C" uint32_t svtme_abi_sizeof(int what) {
    switch (what) {
    case 0: return (uint32_t)sizeof(SvtMeCtrls);
    case 1: return (uint32_t)sizeof(SvtMePicParams);
    case 2: return (uint32_t)sizeof(SvtMeSearchResult);
    case 3: return (uint32_t)sizeof(SvtMeWindow);
    case 4: return (uint32_t)sizeof(SvtMePicResults);
    case 5: return (uint32_t)sizeof(SbState);
    default: return 0;
    }
}
