// SeRANN population engine: host side of the C ABI in include/serann_b200.h.
//
// One engine = one GPU = one shard of the population.  srn_build turns the flat per-individual layer
// lists into (a) packed arenas - fp16 parameters, fp16 Adam slots, fp32 gradient accumulators,
// fp16 activations and activation gradients - and (b) a launch plan: layers of all individuals are
// bucketed by (depth, kind, tile shape) and every bucket is ONE kernel launch over a device table of
// problems.  A training step is then ~3 launches per bucket regardless of the population size.
#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <functional>
#include <map>
#include <random>
#include <string>
#include <tuple>
#include <vector>

#include <cuda.h>          // CUtensorMap and the encode entry point (fetched through the runtime, libcuda is not linked)

#include "../../include/serann_b200.h"
#include "srn_gemm_simt.cuh"
#include "srn_gemm_tc.cuh"
#include "srn_conv_slab.cuh"
#include "srn_kernels.cuh"
#include "srn_host.h"
#include "srn_tma.h"

namespace srn {

thread_local std::string g_error;
std::atomic<unsigned long long> g_launches{0};
std::atomic<unsigned long long> g_allocs{0};      // cudaMalloc calls made by this library (srn_alloc_count)

cudaError_t dmalloc(void** p, size_t bytes) {
    g_allocs.fetch_add(1, std::memory_order_relaxed);
    return cudaMalloc(p, bytes);
}

int fail(int code, const char* fmt, ...) {
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    g_error = buf;
    return code;
}

static inline float h16(float x) { return __half2float(__float2half_rn(x)); }

enum Pass { P_FWD = 0, P_DGRAD, P_WGRAD, P_COL, P_BWD_EW, P_COUNT };

struct Tensor {
    long long rows = 0;        // per sample: product of the leading axes
    int C = 0, Cs = 0;         // channels and channel stride (C rounded up to 8; 1 for the model inputs)
    long long elems = 0;       // rows * C
    long long bs = 0;          // batch stride in elements = rows * Cs
    int root = -1;             // tensor whose storage this one views
    bool view = false;
    int inplace_of = -1;       // concat tensor whose buffer this tensor is written into directly (no copy)
    long long inplace_off = 0;
    // Padded concatenation (a (1, F) concatenate read only by Dense layers): every member sits at an offset that is a
    // multiple of 8 and a member produced in place keeps its padded [row][Cs] layout, so no member is compacted; the
    // readers' kernels get zero rows at the pad positions.  phys_C = physical row length, colmap[logical] = physical.
    long long phys_C = 0;
    std::vector<int> colmap;
    bool needs_grad = false;
    int depth = 0;
    int consumers = 0;         // consumers needing din (counted on the root)
    long long act_off = -1, grad_off = -1;   // offsets into the activation arena (elements)
    __half* act = nullptr;
    __half* grad = nullptr;
};

struct LayerRt {
    srn_layer_desc d;
    long long w_off = -1, w_cnt = 0, b_off = -1, b_cnt = 0;   // trainable arena (elements; w_cnt counts padded rows)
    int w_rows = 0, w_cols = 0, ldw = 0;                      // kernel as [w_rows][ldw], w_cols real columns
    long long mov_off = -1;                                   // moving-statistics arena
    long long save_off = -1;                                  // fp32 BN scratch arena
    bool copy = false, dead = false;                          // reshape that needs a compacting copy / is never read
    std::vector<long long> in_poff;                           // padded concatenate: physical offset of every member
    std::vector<int> rowmap;                                  // Dense on a padded concatenation: logical kernel row -> arena row
    int w_rows_log = 0;                                       // kernel rows as Keras counts them
};

struct Indiv {
    std::vector<LayerRt> layers;
    std::vector<Tensor> tensors;
    float lb = 0.f;
    long long pcount = 0;
    int lane = 0;
    long long act_cost = 0;        // activation elements per sample: what the lanes are balanced on
};

// Device memory owned by an engine handle is grow-only: a buffer is reallocated (with some headroom) only when a call
// needs more than it holds, so that a steady-state generation - same population size, similar architectures - makes
// no cudaMalloc / cudaFree at all (both synchronise the whole device).
// SRN_GUARD=1 (tests; compute-sanitizer is closed on the GPU pool): every buffer gets a 4 KiB band of 0xA5 right behind
// the bytes the current use asked for, and srn_debug_check_guards reports any band that a kernel has written into.
constexpr size_t GUARD_BYTES = 4096;
static bool guards_on() { static const bool on = getenv("SRN_GUARD") != nullptr; return on; }
struct DevBuf {
    void* p = nullptr;
    size_t cap = 0;
    size_t used = 0;         // bytes of the current use (where the guard band starts)
    template <class T> T* as() const { return static_cast<T*>(p); }
};
static cudaError_t buf_reserve_raw(DevBuf& b, size_t bytes, double headroom);
static cudaError_t buf_reserve(DevBuf& b, size_t bytes, double headroom = 1.0) {
    if (!guards_on()) return buf_reserve_raw(b, bytes, headroom);
    cudaError_t e = buf_reserve_raw(b, bytes + GUARD_BYTES, headroom);
    if (e != cudaSuccess) return e;
    b.used = bytes;
    return cudaMemset(static_cast<char*>(b.p) + bytes, 0xA5, GUARD_BYTES);
}
static bool buf_guard_ok(const DevBuf& b) {
    if (!guards_on() || !b.p) return true;
    unsigned char host[GUARD_BYTES];
    if (cudaMemcpy(host, static_cast<const char*>(b.p) + b.used, GUARD_BYTES, cudaMemcpyDeviceToHost) != cudaSuccess) return false;
    for (size_t i = 0; i < GUARD_BYTES; ++i) if (host[i] != 0xA5) return false;
    return true;
}
static cudaError_t buf_reserve_raw(DevBuf& b, size_t bytes, double headroom) {
    if (bytes <= b.cap && b.p) return cudaSuccess;
    if (b.p) { cudaFree(b.p); b.p = nullptr; b.cap = 0; }
    size_t want = (size_t)((double)bytes * headroom);
    want = (std::max(want, bytes) + 255) & ~(size_t)255;
    cudaError_t e = dmalloc(&b.p, want);
    if (e != cudaSuccess && want > bytes) { cudaGetLastError(); want = (bytes + 255) & ~(size_t)255; e = dmalloc(&b.p, want); }
    if (e != cudaSuccess) { b.p = nullptr; return e; }
    b.cap = want;
    return cudaSuccess;
}
static void buf_free(DevBuf& b) { if (b.p) cudaFree(b.p); b.p = nullptr; b.cap = 0; }

// Bump allocator for everything a build and its launch plans create (problem tables, prefix sums, tile tables, tensor
// maps, split-K scratch): nothing is freed one by one, the pool is rewound by the next build; if it ever needed a
// second block the next rewind replaces the blocks by one of their combined size.
struct DevPool {
    std::vector<DevBuf> blocks;
    size_t off = 0;          // offset in the last block
    void* alloc(size_t bytes, cudaError_t* err) {
        bytes = (bytes + 255) & ~(size_t)255;
        if (blocks.empty() || off + bytes > blocks.back().cap) {
            DevBuf b;
            *err = buf_reserve(b, std::max(bytes, (size_t)8 << 20));
            if (*err != cudaSuccess) return nullptr;
            blocks.push_back(b);
            off = 0;
        }
        void* r = static_cast<char*>(blocks.back().p) + off;
        off += bytes;
        *err = cudaSuccess;
        return r;
    }
    cudaError_t rewind() {
        if (blocks.size() > 1) {
            size_t total = 0;
            for (auto& b : blocks) { total += b.cap; buf_free(b); }
            blocks.clear();
            DevBuf b;
            cudaError_t e = buf_reserve(b, total + total / 4);
            if (e != cudaSuccess) return e;
            blocks.push_back(b);
        }
        off = 0;
        return cudaSuccess;
    }
    void release() { for (auto& b : blocks) buf_free(b); blocks.clear(); off = 0; }
};

struct Cached {
    int* prefix[P_COUNT] = {};
    int total[P_COUNT] = {};
    TcTile* tiles[P_COUNT] = {};   // tensor-core passes: one descriptor per tile, heaviest first
    std::vector<TcTile> host_fwd;  // forward tiles kept on the host for the per-level merged tables
    std::vector<SplitFin> host_fin;   // split-K forward problems of this bucket (k_splitk_finish), blk0 relative to the bucket
    SplitFin* fin = nullptr;
    int fin_blocks = 0;
    SlabProb* slab = nullptr;         // problems of the bucket that run forward on the slab kernel (srn_conv_slab.cuh)
    int* slab_prefix = nullptr;       // their tile prefix (device, slab_np + 1 entries)
    int slab_np = 0, slab_total = 0, slab_stride = 0, slab_wbytes = 0;
};

struct Group {
    int depth, kind, cfgF, cfgC, ordinal, tc;
    int lane = 0;                  // half of the shard this bucket's individuals belong to (srn_engine::LaneCtx)
    int mode = 3;                  // bit 0: runs when the inputs are a shared batch, bit 1: when they are per individual
    std::vector<Prob> h;
    Prob* d = nullptr;
    int* done = nullptr;           // BatchNorm buckets: per-problem chunk counters of the fused finalize
    std::map<int, Cached> cache;
};

}  // namespace srn

using namespace srn;

struct ProfRec { int cat; int group; cudaEvent_t a, b; };

struct srn_engine {
    bool profiling = false;
    int cur_group = -1;                   // launch bucket being issued (profiling detail)
    std::vector<ProfRec*> prof;
    srn_config cfg{};
    cudaStream_t stream = 0;
    cudaStream_t cur = 0;                 // stream the LAUNCH macros use (main stream or one of the side streams)
    static constexpr int NSIDE = 7;
    static constexpr int NWG = 3;
    static constexpr int NLANES = 2;
    // A lane = one half of the shard's individuals with its own set of streams.  Individuals are independent, so the two
    // halves walk their depth levels out of step: while one half is in its deep, latency-bound levels (a few small
    // launches in a dependency chain) the other half's wide first levels keep the SMs busy.  Lane 0 runs on the
    // engine's main stream; inference passes (one forward over all individuals) use lane 0 for everything.
    struct LaneCtx {
        cudaStream_t main = nullptr;
        cudaStream_t side[NSIDE] = {};        // independent launch buckets of one depth level overlap their tails on these
        cudaEvent_t ev_fork = nullptr, ev_join[NSIDE] = {};
        cudaStream_t wg_stream[NWG] = {};
        cudaEvent_t ev_wg = nullptr, ev_wg_done[NWG] = {};
        int wg_next = 0; bool wg_used[NWG] = {};
        cudaEvent_t ev_done = nullptr;
        Prob* d_loss = nullptr; int n_loss = 0;      // this lane's slice of the loss problems
    } lane[NLANES];
    LaneCtx* L = nullptr;                 // lane whose streams the launch helpers use
    int n_lanes = 1;                      // lanes the current population is split into
    int stagger_depth = 0;                // lane 1 starts once lane 0 has issued this forward depth level (0: together)
    cudaEvent_t ev_inputs = nullptr;
    // the optimiser pass runs on its own stream so that the next step's batch gather and im2col overlap it
    // weight-gradient launches float: they only need their layer's activation backward, nothing in the backward pass
    // needs them, so they run on their own streams, fill the idle SMs of the latency-bound deep levels and are joined
    // before the optimiser
    cudaStream_t adam_stream = nullptr;
    cudaEvent_t ev_adam_go = nullptr, ev_adam_done = nullptr;
    bool adam_pending = false;
    bool multi_stream = true;
    int num_sms = 148;
    std::vector<Im2colGeo> geos;         // shared im2col buffers of the first layers (one per window geometry)
    __half* im2col = nullptr;
    int xe = 0, G = 0, ncls = 0, NL = 0;
    // dataset
    __half* x[2] = {nullptr, nullptr};
    uint8_t* y[2] = {nullptr, nullptr};
    uint8_t* gb[2] = {nullptr, nullptr};
    int nrows[2] = {0, 0};
    // staging
    __half* xin = nullptr; __half* gin = nullptr; uint8_t* lab = nullptr;
    // population
    int n = 0;
    std::vector<Indiv> inds;
    std::vector<Group> groups;
    Prob* d_loss = nullptr; std::vector<Prob> h_loss;
    float* d_lb = nullptr;
    __half *P = nullptr, *M = nullptr, *V = nullptr, *S = nullptr, *act = nullptr;
    float *Gr = nullptr, *save = nullptr;
    long long p_elems = 0, s_elems = 0, act_elems = 0, save_elems = 0;
    long long device_bytes = 0;
    int adam_t = 0;
    bool built = false;
    // metric buffers
    int* met_correct = nullptr; float* met_ce = nullptr; float* met_mse = nullptr; long long met_slots = 0;
    // grow-only device memory (see DevBuf / DevPool)
    DevBuf bP, bM, bV, bGr, bS, bSave, bAct, bIm2col, bMetC, bMetCe, bMetMse;
    DevBuf bPerm, bLogits, bIdx, bRepX, bRepG, bRepPar, bRepProbs, bRepPacked, bRepMut;
    DevPool pool;
    cudaEvent_t ev_t0 = nullptr, ev_t1 = nullptr;      // srn_fit's timing events
    // forward launches of one depth level that run the same kernel (same N tile) share one tile table: key (nb, depth, BN, mode)
    struct MergedFwd { TcTile* dev = nullptr; int total = 0; int first_group = -1; SplitFin* fin = nullptr; int nfin = 0, fin_blocks = 0; };
    std::map<std::tuple<int, int, int, int>, MergedFwd> merged_fwd;
};

#define CK(call)                                                                                   \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess) return fail(SRN_ERR_CUDA, "%s: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
    } while (0)

// device memory from the handle's pool (rewound by the next build)
#define POOL_ALLOC(ptr, T, bytes)                                                                  \
    do {                                                                                           \
        cudaError_t pe_;                                                                           \
        (ptr) = static_cast<T*>(h->pool.alloc((bytes), &pe_));                                     \
        if (pe_ != cudaSuccess) return fail(SRN_ERR_CUDA, "device pool: %s (%s:%d)", cudaGetErrorString(pe_), __FILE__, __LINE__); \
    } while (0)

// every launch is counted and, when profiling is on, bracketed by CUDA events tagged with its category
#define LAUNCH(cat, kern, grid, block, ...)                                                        \
    do {                                                                                           \
        if (grid_nonzero(grid)) {                                                                          \
            ProfRec* pr_ = h->profiling ? prof_begin(h, (cat)) : nullptr;                          \
            kern<<<(grid), (block), 0, h->cur>>>(__VA_ARGS__);                                     \
            if (pr_) cudaEventRecord(pr_->b, h->cur);                                           \
            g_launches.fetch_add(1, std::memory_order_relaxed);                                    \
        }                                                                                          \
    } while (0)

static inline bool grid_nonzero(unsigned g) { return g > 0; }
static inline bool grid_nonzero(int g) { return g > 0; }
static inline bool grid_nonzero(const dim3& g) { return g.x > 0 && g.y > 0 && g.z > 0; }

#define LAUNCH_SMEM(cat, kern, grid, block, smem, ...)                                             \
    do {                                                                                           \
        if (grid_nonzero(grid)) {                                                                  \
            ProfRec* pr_ = h->profiling ? prof_begin(h, (cat)) : nullptr;                          \
            kern<<<(grid), (block), (smem), h->cur>>>(__VA_ARGS__);                                \
            if (pr_) cudaEventRecord(pr_->b, h->cur);                                           \
            g_launches.fetch_add(1, std::memory_order_relaxed);                                    \
        }                                                                                          \
    } while (0)

static ProfRec* prof_begin(srn_engine* h, int cat) {
    ProfRec* r = new ProfRec();
    r->cat = cat;
    r->group = h->cur_group;
    cudaEventCreate(&r->a);
    cudaEventCreate(&r->b);
    cudaEventRecord(r->a, h->cur);
    h->prof.push_back(r);
    return r;
}

extern "C" const char* srn_last_error(void) { return g_error.c_str(); }
extern "C" uint64_t srn_launch_count(void) { return g_launches.load(); }
extern "C" uint64_t srn_alloc_count(void) { return g_allocs.load(); }

// ------------------------------------------------------------------ create / destroy
// Forget the population.  The device buffers stay with the handle for the next build (grow-only); release = true
// (srn_destroy) gives them back.
static void free_population(srn_engine* h, bool release = false) {
    h->merged_fwd.clear();
    h->groups.clear();
    h->inds.clear();
    h->d_loss = nullptr; h->d_lb = nullptr;
    for (auto& L : h->lane) { L.d_loss = nullptr; L.n_loss = 0; }
    h->n_lanes = 1;
    h->im2col = nullptr; h->geos.clear();
    h->P = h->M = h->V = h->S = h->act = nullptr; h->Gr = h->save = nullptr;
    h->met_correct = nullptr; h->met_ce = h->met_mse = nullptr; h->met_slots = 0;
    h->built = false; h->n = 0; h->device_bytes = 0;
    if (release) {
        for (DevBuf* b : {&h->bP, &h->bM, &h->bV, &h->bGr, &h->bS, &h->bSave, &h->bAct, &h->bIm2col, &h->bMetC, &h->bMetCe, &h->bMetMse,
                          &h->bPerm, &h->bLogits, &h->bIdx, &h->bRepX, &h->bRepG, &h->bRepPar, &h->bRepProbs, &h->bRepPacked, &h->bRepMut})
            buf_free(*b);
        h->pool.release();
    }
}

static int create_resources(srn_engine* h);

extern "C" int srn_create(const srn_config* cfg, srn_engine** out) {
    if (!cfg || !out) return fail(SRN_ERR_ARG, "srn_create: null argument");
    if (cfg->image_h <= 0 || cfg->image_w <= 0 || cfg->genotype_size <= 0 || cfg->n_classes <= 0 || cfg->max_batch <= 0)
        return fail(SRN_ERR_ARG, "srn_create: non-positive dimension");
    if (cfg->n_classes > 255) return fail(SRN_ERR_LIMIT, "srn_create: n_classes > 255");
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0)
        return fail(SRN_ERR_CUDA, "srn_create: no CUDA device (%s); this engine has no CPU path", cudaGetErrorString(e));
    if (cfg->device < 0 || cfg->device >= count) return fail(SRN_ERR_ARG, "srn_create: device %d of %d", cfg->device, count);
    CK(cudaSetDevice(cfg->device));
    {
        const int max_smem = tc_smem_bytes(256);
#define TC_ATTR(MODE, SCALAR, BN_) CK(cudaFuncSetAttribute(k_gemm_tc<MODE, SCALAR, BN_>, cudaFuncAttributeMaxDynamicSharedMemorySize, tc_smem_bytes(BN_)))
#define TC_ATTR4(MODE, SCALAR) TC_ATTR(MODE, SCALAR, 32); TC_ATTR(MODE, SCALAR, 64); TC_ATTR(MODE, SCALAR, 128); TC_ATTR(MODE, SCALAR, 256)
        TC_ATTR4(TC_FWD, false); TC_ATTR4(TC_FWD, true); TC_ATTR4(TC_DGRAD, false); TC_ATTR4(TC_WGRAD, false); TC_ATTR4(TC_WGRAD, true);
#undef TC_ATTR4
#undef TC_ATTR
        CK(cudaFuncSetAttribute(k_conv_slab<32>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024));
        CK(cudaFuncSetAttribute(k_conv_slab<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024));
    }
    srn_engine* h = new srn_engine();
    h->cfg = *cfg;
    int rc = create_resources(h);
    if (rc) { srn_destroy(h); return rc; }      // streams, events and buffers made so far are given back
    *out = h;
    return SRN_OK;
}

static int create_resources(srn_engine* h) {
    const srn_config* cfg = &h->cfg;
    CK(cudaDeviceGetAttribute(&h->num_sms, cudaDevAttrMultiProcessorCount, cfg->device));
    for (int l = 0; l < srn_engine::NLANES; ++l) {
        srn_engine::LaneCtx& L = h->lane[l];
        if (l == 0) L.main = h->stream;
        else CK(cudaStreamCreateWithFlags(&L.main, cudaStreamNonBlocking));
        for (int i = 0; i < srn_engine::NSIDE; ++i) {
            CK(cudaStreamCreateWithFlags(&L.side[i], cudaStreamNonBlocking));
            CK(cudaEventCreateWithFlags(&L.ev_join[i], cudaEventDisableTiming));
        }
        CK(cudaEventCreateWithFlags(&L.ev_fork, cudaEventDisableTiming));
        CK(cudaEventCreateWithFlags(&L.ev_wg, cudaEventDisableTiming));
        CK(cudaEventCreateWithFlags(&L.ev_done, cudaEventDisableTiming));
        for (int i = 0; i < srn_engine::NWG; ++i) {
            CK(cudaStreamCreateWithFlags(&L.wg_stream[i], cudaStreamNonBlocking));
            CK(cudaEventCreateWithFlags(&L.ev_wg_done[i], cudaEventDisableTiming));
        }
    }
    h->L = &h->lane[0];
    CK(cudaEventCreateWithFlags(&h->ev_inputs, cudaEventDisableTiming));
    CK(cudaStreamCreateWithFlags(&h->adam_stream, cudaStreamNonBlocking));
    CK(cudaEventCreateWithFlags(&h->ev_adam_go, cudaEventDisableTiming));
    CK(cudaEventCreateWithFlags(&h->ev_adam_done, cudaEventDisableTiming));
    h->multi_stream = getenv("SRN_SINGLE_STREAM") == nullptr;
    h->cur = h->stream;
    h->xe = cfg->image_h * cfg->image_w;
    h->G = cfg->genotype_size;
    h->ncls = cfg->n_classes;
    h->NL = h->ncls + h->G;
    CK(cudaEventCreate(&h->ev_t0)); CK(cudaEventCreate(&h->ev_t1));
    CK(dmalloc((void**)&h->xin, (size_t)cfg->max_batch * h->xe * sizeof(__half)));
    CK(dmalloc((void**)&h->gin, (size_t)cfg->max_batch * h->G * sizeof(__half)));
    CK(dmalloc((void**)&h->lab, (size_t)cfg->max_batch));
    return SRN_OK;
}

extern "C" void srn_destroy(srn_engine* h) {
    if (!h) return;
    cudaSetDevice(h->cfg.device);
    cudaDeviceSynchronize();
    free_population(h, true);
    if (h->ev_t0) cudaEventDestroy(h->ev_t0);
    if (h->ev_t1) cudaEventDestroy(h->ev_t1);
    for (int l = 0; l < srn_engine::NLANES; ++l) {
        srn_engine::LaneCtx& L = h->lane[l];
        for (int i = 0; i < srn_engine::NSIDE; ++i) { if (L.side[i]) cudaStreamDestroy(L.side[i]); if (L.ev_join[i]) cudaEventDestroy(L.ev_join[i]); }
        if (L.ev_fork) cudaEventDestroy(L.ev_fork);
        if (L.ev_wg) cudaEventDestroy(L.ev_wg);
        if (L.ev_done) cudaEventDestroy(L.ev_done);
        for (int i = 0; i < srn_engine::NWG; ++i) {
            if (L.ev_wg_done[i]) cudaEventDestroy(L.ev_wg_done[i]);
            if (L.wg_stream[i]) cudaStreamDestroy(L.wg_stream[i]);
        }
        if (l > 0 && L.main) cudaStreamDestroy(L.main);
    }
    if (h->ev_inputs) cudaEventDestroy(h->ev_inputs);
    if (h->ev_adam_go) cudaEventDestroy(h->ev_adam_go);
    if (h->ev_adam_done) cudaEventDestroy(h->ev_adam_done);
    if (h->adam_stream) cudaStreamDestroy(h->adam_stream);
    for (int s = 0; s < 2; ++s) { cudaFree(h->x[s]); cudaFree(h->y[s]); cudaFree(h->gb[s]); }
    cudaFree(h->xin); cudaFree(h->gin); cudaFree(h->lab);
    delete h;
}

extern "C" int srn_set_dataset(srn_engine* h, const float* x_train, const uint8_t* y_train, const uint8_t* g_train,
                               int32_t n_train, const float* x_test, const uint8_t* y_test, const uint8_t* g_test,
                               int32_t n_test) {
    if (!h) return fail(SRN_ERR_ARG, "null engine");
    CK(cudaSetDevice(h->cfg.device));
    const float* xs[2] = {x_train, x_test};
    const uint8_t* ys[2] = {y_train, y_test};
    const uint8_t* gs[2] = {g_train, g_test};
    const int ns[2] = {n_train, n_test};
    for (int s = 0; s < 2; ++s) {
        cudaFree(h->x[s]); cudaFree(h->y[s]); cudaFree(h->gb[s]);
        h->x[s] = nullptr; h->y[s] = nullptr; h->gb[s] = nullptr; h->nrows[s] = 0;
        if (ns[s] <= 0) continue;
        if (!xs[s] || !gs[s]) return fail(SRN_ERR_ARG, "srn_set_dataset: null array for split %d", s);
        const long long ne = (long long)ns[s] * h->xe;
        float* tmp = nullptr;
        CK(dmalloc((void**)&tmp, ne * sizeof(float)));
        CK(cudaMemcpy(tmp, xs[s], ne * sizeof(float), cudaMemcpyHostToDevice));
        CK(dmalloc((void**)&h->x[s], ne * sizeof(__half)));
        LAUNCH(SRN_CAT_OTHER, k_f32_to_f16, (unsigned)((ne + 255) / 256), 256, tmp, h->x[s], ne);
        CK(cudaStreamSynchronize(h->stream));
        CK(cudaFree(tmp));
        CK(dmalloc((void**)&h->y[s], ns[s]));
        if (ys[s]) CK(cudaMemcpy(h->y[s], ys[s], ns[s], cudaMemcpyHostToDevice));
        else CK(cudaMemset(h->y[s], 0, ns[s]));
        CK(dmalloc((void**)&h->gb[s], (size_t)ns[s] * h->G));
        CK(cudaMemcpy(h->gb[s], gs[s], (size_t)ns[s] * h->G, cudaMemcpyHostToDevice));
        h->nrows[s] = ns[s];
    }
    return SRN_OK;
}

// ------------------------------------------------------------------ build
static inline long long align_up(long long v, long long a) { return (v + a - 1) / a * a; }

static int cfg_for(int n) { return n <= 16 ? 2 : (n <= 32 ? 1 : 0); }
// tensor-core N tile: 32 / 64 / 128 / 256 columns
static int tc_cfg_for(int n) { return n <= 32 ? 0 : (n <= 64 ? 1 : (n <= 128 ? 2 : 3)); }
static const int TC_BN[4] = {32, 64, 128, 256};
static int g_tensor_cores = 1;      // srn_debug_set_tensor_cores (test hook): 0 routes every GEMM through the SIMT kernels
static const int CFG_BM[3] = {64, 128, 256};
static const int CFG_BN[3] = {64, 32, 16};

static bool has_params(int kind) { return kind == SRN_KIND_DENSE || kind == SRN_KIND_CONV || kind == SRN_KIND_BN; }

static int validate_layer(const srn_layer_desc& d, const std::vector<Tensor>& ts, int li) {
    if (d.n_in < 1 || d.n_in > SRN_MAX_CONCAT_INPUTS) return fail(SRN_ERR_UNSUPPORTED, "layer %d: %d inputs", li, d.n_in);
    for (int j = 0; j < d.n_in; ++j)
        if (d.in[j] < 0 || d.in[j] >= (int)ts.size()) return fail(SRN_ERR_ARG, "layer %d: input id %d out of range", li, d.in[j]);
    if (d.kind != SRN_KIND_CONCAT && d.n_in != 1) return fail(SRN_ERR_ARG, "layer %d: needs exactly one input", li);
    const Tensor& t0 = ts[d.in[0]];
    const long long in_e = (long long)d.in_h * d.in_w * d.in_c, out_e = (long long)d.out_h * d.out_w * d.out_c;
    if (in_e <= 0 || out_e <= 0) return fail(SRN_ERR_ARG, "layer %d: empty tensor", li);
    if (d.kind != SRN_KIND_CONCAT && in_e != t0.elems) return fail(SRN_ERR_ARG, "layer %d: input has %lld elements, descriptor says %lld", li, t0.elems, in_e);
    if (d.kind != SRN_KIND_CONCAT && d.kind != SRN_KIND_RESHAPE && (d.in_c != t0.C || (long long)d.in_h * d.in_w != t0.rows))
        return fail(SRN_ERR_ARG, "layer %d: input is (%lld rows, %d channels), descriptor says (%d x %d, %d)", li, t0.rows, t0.C, d.in_h, d.in_w, d.in_c);
    switch (d.kind) {
        case SRN_KIND_DENSE:
            if (d.units <= 0 || d.out_c != d.units || d.out_h * d.out_w != d.in_h * d.in_w) return fail(SRN_ERR_ARG, "layer %d: bad Dense shape", li);
            break;
        case SRN_KIND_CONV:
            if (d.units <= 0 || d.out_c != d.units || d.kh <= 0 || d.kw <= 0 || d.sh <= 0 || d.sw <= 0 ||
                d.in_h < d.kh || d.in_w < d.kw || d.out_h != (d.in_h - d.kh) / d.sh + 1 || d.out_w != (d.in_w - d.kw) / d.sw + 1)
                return fail(SRN_ERR_ARG, "layer %d: bad convolution shape", li);
            break;
        case SRN_KIND_MAXPOOL:
            if (d.kh <= 0 || d.kw <= 0 || d.sh <= 0 || d.sw <= 0 || d.in_h < d.kh || d.in_w < d.kw || d.out_c != d.in_c ||
                d.out_h != (d.in_h - d.kh) / d.sh + 1 || d.out_w != (d.in_w - d.kw) / d.sw + 1)
                return fail(SRN_ERR_ARG, "layer %d: bad pooling shape", li);
            break;
        case SRN_KIND_BN:
        case SRN_KIND_RESHAPE:
            if (in_e != out_e) return fail(SRN_ERR_ARG, "layer %d: element count changes", li);
            if (d.kind == SRN_KIND_BN && d.in_c != d.out_c) return fail(SRN_ERR_ARG, "layer %d: BN channels", li);
            break;
        case SRN_KIND_CONCAT: {
            long long csum = 0;
            for (int j = 0; j < d.n_in; ++j) {
                const Tensor& t = ts[d.in[j]];
                const long long rows = (long long)d.out_h * d.out_w;
                if (t.elems % rows) return fail(SRN_ERR_ARG, "layer %d: concat rows mismatch", li);
                csum += t.elems / rows;
            }
            if (csum != d.out_c) return fail(SRN_ERR_ARG, "layer %d: concat channels %lld != %d", li, csum, d.out_c);
            break;
        }
        default:
            return fail(SRN_ERR_UNSUPPORTED, "layer %d: kind %d", li, d.kind);
    }
    return SRN_OK;
}

extern "C" int srn_build(srn_engine* h, const srn_layer_desc* layers, const int32_t* off, const float* loss_balance,
                         int32_t n, uint64_t seed) {
    if (n <= 0) return fail(SRN_ERR_ARG, "srn_build: bad argument");
    std::vector<uint64_t> seeds((size_t)n);
    for (int i = 0; i < n; ++i) seeds[(size_t)i] = seed * 0x9E3779B97F4A7C15ull + ((uint64_t)i << 20);
    return srn_build_seeded(h, layers, off, loss_balance, n, seeds.data());
}

extern "C" int srn_build_seeded(srn_engine* h, const srn_layer_desc* layers, const int32_t* off, const float* loss_balance,
                                int32_t n, const uint64_t* seeds) {
    if (!h || !layers || !off || !loss_balance || !seeds || n <= 0) return fail(SRN_ERR_ARG, "srn_build: bad argument");
    CK(cudaSetDevice(h->cfg.device));
    CK(cudaStreamSynchronize(h->stream));
    free_population(h);
    const int MB = h->cfg.max_batch;
    h->inds.resize(n);
    long long p_off = 0, s_off = 0, a_off = 0, sv_off = 0;
    // ---- pass 1: tensors (rows, C, channel stride), parameter offsets, sizes
    for (int i = 0; i < n; ++i) {
        Indiv& ind = h->inds[i];
        ind.lb = loss_balance[i];
        const int nl = off[i + 1] - off[i];
        if (nl <= 0) return fail(SRN_ERR_ARG, "individual %d has no layers", i);
        ind.layers.resize(nl);
        ind.tensors.resize(2 + nl);
        Tensor& tx = ind.tensors[0]; tx.rows = h->xe; tx.C = tx.Cs = 1; tx.elems = h->xe; tx.bs = h->xe; tx.root = 0;
        Tensor& tg = ind.tensors[1]; tg.rows = h->G; tg.C = tg.Cs = 1; tg.elems = h->G; tg.bs = h->G; tg.root = 1;
        // a concatenate reads through reshapes (flat order is preserved), so a reshape that would need a
        // compacting copy is only materialised when something else consumes it
        std::vector<int> used(2 + nl, 0);       // direct readers of a tensor (a concatenate counts on the chased source)
        std::vector<int> used_src(2 + nl, 0);   // readers of a tensor through any chain of reshapes
        auto chase = [&](int t) { while (t >= 2 && layers[off[i] + t - 2].kind == SRN_KIND_RESHAPE) t = layers[off[i] + t - 2].in[0]; return t; };
        for (int l = 0; l < nl; ++l) {
            const srn_layer_desc& d = layers[off[i] + l];
            for (int j = 0; j < d.n_in && j < SRN_MAX_CONCAT_INPUTS; ++j) {
                if (d.in[j] < 0 || d.in[j] >= 2 + l) return fail(SRN_ERR_ARG, "individual %d layer %d: input id %d out of range", i, l, d.in[j]);
                used[d.kind == SRN_KIND_CONCAT ? chase(d.in[j]) : d.in[j]] += 1;
                if (d.kind != SRN_KIND_RESHAPE) used_src[chase(d.in[j])] += 1;
            }
        }
        used[1 + nl] += 1;
        used_src[chase(1 + nl)] += 1;
        // readers of a tensor (through reshapes) that are not Dense layers on a one-row view: a concatenation with none
        // of them can take the padded layout
        std::vector<int> other_readers(2 + nl, 0);
        for (int l = 0; l < nl; ++l) {
            const srn_layer_desc& d = layers[off[i] + l];
            if (d.kind == SRN_KIND_RESHAPE) continue;
            for (int j = 0; j < d.n_in && j < SRN_MAX_CONCAT_INPUTS; ++j) {
                bool flat = d.kind == SRN_KIND_DENSE;
                int t = d.in[j];
                while (t >= 2 && layers[off[i] + t - 2].kind == SRN_KIND_RESHAPE) {
                    const srn_layer_desc& r = layers[off[i] + t - 2];
                    if ((long long)r.out_h * r.out_w != 1) flat = false;
                    t = r.in[0];
                }
                if (!flat) other_readers[t] += 1;
            }
        }
        other_readers[chase(1 + nl)] += 1;
        const bool allow_padded = getenv("SRN_NO_PADDED_CONCAT") == nullptr;
        for (int l = 0; l < nl; ++l) {
            LayerRt& L = ind.layers[l];
            L.d = layers[off[i] + l];
            const srn_layer_desc& d = L.d;
            std::vector<Tensor> visible(ind.tensors.begin(), ind.tensors.begin() + 2 + l);
            int rc = validate_layer(d, visible, l);
            if (rc) { std::string m = g_error; return fail(rc, "individual %d: %s", i, m.c_str()); }
            Tensor& to = ind.tensors[2 + l];
            const Tensor& ti = ind.tensors[d.in[0]];
            to.rows = (long long)d.out_h * d.out_w; to.C = d.out_c; to.elems = to.rows * to.C;
            bool ng = has_params(d.kind);
            int depth = 0;
            for (int j = 0; j < d.n_in; ++j) {
                const Tensor& tj = ind.tensors[d.kind == SRN_KIND_CONCAT ? chase(d.in[j]) : d.in[j]];
                ng = ng || tj.needs_grad; depth = std::max(depth, tj.depth);
            }
            to.needs_grad = ng;
            to.root = 2 + l; to.depth = depth + 1;
            to.Cs = (int)align_up(to.C, 8);
            if (d.kind == SRN_KIND_RESHAPE) {
                const bool src_flat = ti.Cs == ti.C || ti.rows == 1;
                if (src_flat && (to.rows == 1 || to.C % 8 == 0)) {       // pure view of the source's storage
                    to.root = ti.root; to.depth = depth; to.view = true;
                    to.Cs = to.rows == 1 ? (int)ti.bs : to.C;
                } else {
                    L.copy = true;                                       // compacting copy, skipped when unused
                    L.dead = used[2 + l] == 0;
                }
            }
            to.bs = to.view ? ti.bs : to.rows * to.Cs;
            if (to.view && ti.phys_C) { to.phys_C = ti.phys_C; to.colmap = ti.colmap; }
            if (d.kind == SRN_KIND_CONCAT && to.rows == 1 && allow_padded && other_readers[2 + l] == 0) {
                long long poff = 0;
                to.colmap.resize((size_t)to.C);
                long long e0 = 0;
                for (int j = 0; j < d.n_in; ++j) {
                    const int si = chase(d.in[j]);
                    Tensor& S = ind.tensors[si];
                    const bool cand = si >= 2 && !S.view && S.inplace_of < 0 && used_src[si] == 1 &&
                                      ind.layers[si - 2].d.kind != SRN_KIND_CONCAT && ind.layers[si - 2].d.kind != SRN_KIND_RESHAPE;
                    L.in_poff.push_back(poff);
                    long long plen = S.elems;
                    if (cand) {
                        plen = S.rows * S.Cs;
                        S.inplace_of = 2 + l; S.inplace_off = poff;
                        for (long long r = 0; r < S.rows; ++r)
                            for (int c = 0; c < S.C; ++c) to.colmap[(size_t)(e0 + r * S.C + c)] = (int)(poff + r * S.Cs + c);
                    } else {
                        for (long long e = 0; e < S.elems; ++e) to.colmap[(size_t)(e0 + e)] = (int)(poff + e);
                    }
                    e0 += S.elems;
                    poff += align_up(plen, 8);
                }
                to.phys_C = poff; to.Cs = (int)poff; to.bs = poff;
                for (int j = 0; j < d.n_in; ++j) {
                    Tensor& S = ind.tensors[chase(d.in[j])];
                    if (S.inplace_of == 2 + l) S.bs = poff;
                }
            }
            if (d.kind == SRN_KIND_DENSE || d.kind == SRN_KIND_CONV) {
                L.w_rows = d.kind == SRN_KIND_DENSE ? d.in_c : d.kh * d.kw * d.in_c;
                L.w_rows_log = L.w_rows;
                if (d.kind == SRN_KIND_DENSE && ti.phys_C) { L.w_rows = (int)ti.phys_C; L.rowmap = ti.colmap; }
                L.w_cols = d.units; L.ldw = (int)align_up(d.units, 8);
                L.w_cnt = (long long)L.w_rows * L.ldw; L.b_cnt = d.units;
                L.w_off = p_off; p_off = align_up(p_off + L.w_cnt, 8);
                L.b_off = p_off; p_off = align_up(p_off + L.b_cnt, 8);
                ind.pcount += (long long)L.w_rows_log * L.w_cols + L.b_cnt;
            } else if (d.kind == SRN_KIND_BN) {
                L.w_cnt = L.b_cnt = d.in_c; L.w_rows = 1; L.w_cols = L.ldw = d.in_c;
                L.w_off = p_off; p_off = align_up(p_off + L.w_cnt, 8);
                L.b_off = p_off; p_off = align_up(p_off + L.b_cnt, 8);
                L.mov_off = s_off; s_off = align_up(s_off + 2LL * d.in_c, 8);
                const long long rows_max = (long long)MB * ti.rows;
                L.save_off = sv_off; sv_off += 4LL * ti.Cs + 2LL * ti.Cs * col_chunks(rows_max, ti.Cs);
                ind.pcount += 4LL * d.in_c;
            }
        }
        if (ind.tensors.back().elems != h->NL || ind.tensors.back().rows != 1)
            return fail(SRN_ERR_ARG, "individual %d: last layer must emit one row of n_classes + genotype_size = %d values", i, h->NL);
        // in-place concatenation: an input that is the dense (C % 8 == 0) output of a layer, read by nothing else,
        // is produced straight into its slice of the (1, F) concatenation buffer - forward copy and backward
        // slice copy disappear; the producers and gradient readers honour the buffer's batch stride
        for (int l = 0; l < nl; ++l) {
            const srn_layer_desc& d = ind.layers[l].d;
            if (d.kind != SRN_KIND_CONCAT || ind.tensors[2 + l].rows != 1 || ind.tensors[2 + l].phys_C) continue;
            long long coff = 0;
            for (int j = 0; j < d.n_in; ++j) {
                const int si = chase(d.in[j]);
                Tensor& S = ind.tensors[si];
                if (si >= 2 && !S.view && S.inplace_of < 0 && S.Cs == S.C && coff % 8 == 0 && used_src[si] == 1 &&
                    ind.layers[si - 2].d.kind != SRN_KIND_CONCAT && ind.layers[si - 2].d.kind != SRN_KIND_RESHAPE) {
                    S.inplace_of = 2 + l; S.inplace_off = coff;
                    S.bs = ind.tensors[2 + l].bs;
                }
                coff += S.elems;
            }
        }
        // storage: every tensor that owns memory gets activations; gradients where needed
        for (size_t t = 2; t < ind.tensors.size(); ++t) {
            Tensor& T = ind.tensors[t];
            if (T.view || ind.layers[t - 2].dead || T.inplace_of >= 0) continue;
            a_off = align_up(a_off, 64);
            T.act_off = a_off; a_off += T.bs * MB;
            if (T.needs_grad) { a_off = align_up(a_off, 64); T.grad_off = a_off; a_off += T.bs * MB; }
        }
    }
    h->p_elems = align_up(std::max(p_off, 8LL), 8); h->s_elems = std::max(s_off, 8LL);
    h->act_elems = std::max(a_off, 64LL); h->save_elems = std::max(sv_off, 8LL);
    const size_t pb = (size_t)h->p_elems * sizeof(__half);
    // arenas are kept across builds and only grow (with headroom, so that the next generation's similar population fits)
    CK(h->pool.rewind());
    CK(buf_reserve(h->bP, pb, 1.25)); CK(buf_reserve(h->bM, pb, 1.25)); CK(buf_reserve(h->bV, pb, 1.25));
    CK(buf_reserve(h->bGr, (size_t)h->p_elems * sizeof(float), 1.25));
    CK(buf_reserve(h->bS, (size_t)h->s_elems * sizeof(__half), 1.25));
    CK(buf_reserve(h->bSave, (size_t)h->save_elems * sizeof(float), 1.25));
    {
        cudaError_t e = buf_reserve(h->bAct, (size_t)h->act_elems * sizeof(__half), 1.15);
        if (e != cudaSuccess) {
            cudaGetLastError();
            free_population(h);
            return fail(SRN_ERR_LIMIT, "activation arena of %.2f GB does not fit: %s", h->act_elems * 2.0 / 1e9, cudaGetErrorString(e));
        }
    }
    h->P = h->bP.as<__half>(); h->M = h->bM.as<__half>(); h->V = h->bV.as<__half>(); h->Gr = h->bGr.as<float>();
    h->S = h->bS.as<__half>(); h->save = h->bSave.as<float>(); h->act = h->bAct.as<__half>();
    h->device_bytes = 3LL * pb + h->p_elems * 4 + h->s_elems * 2 + h->save_elems * 4 + h->act_elems * 2;
    // everything starts at zero; the channel / row pads of tensors and kernels are never written afterwards
    CK(cudaMemsetAsync(h->P, 0, pb, h->stream)); CK(cudaMemsetAsync(h->M, 0, pb, h->stream)); CK(cudaMemsetAsync(h->V, 0, pb, h->stream));
    CK(cudaMemsetAsync(h->Gr, 0, (size_t)h->p_elems * sizeof(float), h->stream));
    CK(cudaMemsetAsync(h->S, 0, (size_t)h->s_elems * sizeof(__half), h->stream));
    CK(cudaMemsetAsync(h->save, 0, (size_t)h->save_elems * sizeof(float), h->stream));
    CK(cudaMemsetAsync(h->act, 0, (size_t)h->act_elems * sizeof(__half), h->stream));
    // ---- lanes: two halves of about equal activation volume (greedy on the sorted sizes); a small shard stays whole
    // (measured on the benchmark population: 9.7 ms per step with two lanes against 8.95 with one - the extra launches
    // and the halved buckets cost more than the overlap of the deep levels gains - so one lane is the default)
    h->n_lanes = (n >= 8 && getenv("SRN_TWO_LANES") != nullptr) ? srn_engine::NLANES : 1;
    {
        std::vector<int> order((size_t)n);
        for (int i = 0; i < n; ++i) {
            order[(size_t)i] = i;
            long long c = 0;
            for (const Tensor& T : h->inds[i].tensors) if (T.act_off >= 0) c += T.bs;
            h->inds[i].act_cost = c;
            h->inds[i].lane = 0;
        }
        if (h->n_lanes > 1) {
            std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return h->inds[a].act_cost > h->inds[b].act_cost; });
            long long load[srn_engine::NLANES] = {};
            for (int i : order) {
                int best = 0;
                for (int l = 1; l < h->n_lanes; ++l) if (load[l] < load[best]) best = l;
                h->inds[i].lane = best;
                load[best] += h->inds[i].act_cost;
            }
        }
    }
    // ---- pass 2: pointers, consumer ordinals, problem groups
    std::map<std::tuple<int, int, int, int, int, int>, Group> gmap;
    long long geo_elems = 0;
    h->h_loss.clear();
    for (int i = 0; i < n; ++i) {
        Indiv& ind = h->inds[i];
        for (auto& T : ind.tensors) {
            if (T.act_off >= 0) T.act = h->act + T.act_off;
            if (T.grad_off >= 0) T.grad = h->act + T.grad_off;
            T.consumers = 0;
        }
        for (auto& T : ind.tensors)              // tensors produced in place inside a concatenation buffer
            if (T.inplace_of >= 0) {
                const Tensor& Cn = ind.tensors[T.inplace_of];
                T.act = Cn.act + T.inplace_off;
                T.grad = (T.needs_grad && Cn.grad) ? Cn.grad + T.inplace_off : nullptr;
            }
        for (auto& T : ind.tensors)              // views share the root's storage (roots precede their views)
            if (T.view) { T.act = ind.tensors[T.root].act; T.grad = ind.tensors[T.root].grad; }
        auto chase = [&](int t) { while (t >= 2 && ind.layers[t - 2].d.kind == SRN_KIND_RESHAPE) t = ind.layers[t - 2].d.in[0]; return t; };
        for (size_t l = 0; l < ind.layers.size(); ++l) {
            const LayerRt& L = ind.layers[l];
            const srn_layer_desc& d = L.d;
            if (d.kind == SRN_KIND_RESHAPE && (!L.copy || L.dead)) continue;
            const bool is_copy = d.kind == SRN_KIND_CONCAT || d.kind == SRN_KIND_RESHAPE;
            const Tensor& to = ind.tensors[2 + l];
            int coff = 0;
            for (int j = 0; j < d.n_in; ++j) {
                const int tin = d.kind == SRN_KIND_CONCAT ? chase(d.in[j]) : d.in[j];
                Tensor& ti = ind.tensors[tin];
                if (d.kind == SRN_KIND_CONCAT && !L.in_poff.empty()) coff = (int)L.in_poff[j];     // padded concatenation: aligned physical offsets
                if (d.kind == SRN_KIND_CONCAT && ti.inplace_of == (int)(2 + l)) { coff += (int)ti.elems; continue; }   // already in place
                Tensor& troot = ind.tensors[ti.root];
                Prob p{};
                p.indiv = i;
                p.in_sel = ti.root == 0 ? 1 : (ti.root == 1 ? 2 : 0);
                p.in = ti.act;
                p.in_bs = (int)ti.bs;
                p.out = to.act; p.out_bs = (int)to.bs;
                p.gout = to.grad;
                p.din = ti.needs_grad ? ti.grad : nullptr;
                int ordinal = 0;
                if (p.din) ordinal = troot.consumers++;
                p.H = d.in_h; p.W = d.in_w; p.C = ti.C; p.Cs = ti.Cs;
                p.Ho = d.out_h; p.Wo = d.out_w; p.F = to.C; p.Fs = to.Cs;
                p.kh = 1; p.kw = 1; p.sh = 1; p.sw = 1;
                p.act = d.act;
                int cfgF = 0, cfgC = 0, tc = 0, kind = d.kind;
                if (d.kind == SRN_KIND_DENSE || d.kind == SRN_KIND_CONV) {
                    if (d.kind == SRN_KIND_CONV) { p.kh = d.kh; p.kw = d.kw; p.sh = d.sh; p.sw = d.sw; }
                    else { p.H = (int)ti.rows; p.W = 1; p.Ho = p.H; p.Wo = 1; }
                    if (d.kind == SRN_KIND_DENSE && ti.phys_C) p.C = (int)ti.phys_C;      // zero kernel rows at the pad positions
                    p.w = h->P + L.w_off; p.bias = h->P + L.b_off;
                    p.gw = h->Gr + L.w_off; p.gb = h->Gr + L.b_off;
                    p.ldw = L.ldw;
                    cfgF = cfg_for(p.F); cfgC = cfg_for(p.C);
                    if (g_tensor_cores) {
                        // 1: operands in 16-byte chunks (channel stride padded to 8); 2: first layers on a one-channel
                        // model input, im2col gathered element by element
                        if (p.Cs % 8 == 0 && p.Fs % 8 == 0) tc = 1;
                        else if (p.in_sel != 0 && p.Cs == 1 && p.C == 1 && p.Fs % 8 == 0) tc = 2;   // only the model inputs have Cs == 1
                        // 3: Dense / 1 x 1 convolution on a one-channel model input: an outer product, elementwise kernels
                        if (tc == 2 && p.kh * p.kw == 1) tc = 3;
                        if (tc) { cfgF = tc_cfg_for(p.F); cfgC = tc_cfg_for(p.C); }
                    }
                } else if (d.kind == SRN_KIND_MAXPOOL) {
                    p.kh = d.kh; p.kw = d.kw; p.sh = d.sh; p.sw = d.sw;
                } else if (d.kind == SRN_KIND_BN) {
                    p.H = (int)ti.rows; p.W = 1;
                    p.w = h->P + L.w_off; p.bias = h->P + L.b_off;
                    p.gw = h->Gr + L.w_off; p.gb = h->Gr + L.b_off;
                    p.mov = h->S + L.mov_off; p.save = h->save + L.save_off;
                } else if (is_copy) {
                    kind = SRN_KIND_CONCAT;
                    p.n_el = (int)ti.elems;
                    if (to.phys_C) p.F = (int)to.phys_C;
                    p.Cj = (int)(ti.elems / to.rows);
                    p.coff = coff;
                    coff += p.Cj;
                }
                int mode = 3;
                if (tc == 2) {
                    // first layer on a model input: with a shared batch it runs as a Dense over the shared im2col
                    // matrix of its window geometry (vector path); the element-wise gather variant below is kept
                    // for per-individual inputs (replication)
                    int gi = -1;
                    for (size_t q = 0; q < h->geos.size(); ++q) {
                        const Im2colGeo& e = h->geos[q];
                        if (e.in_bs == -p.in_sel && e.kh == p.kh && e.kw == p.kw && e.sh == p.sh && e.sw == p.sw) { gi = (int)q; break; }
                    }
                    if (gi < 0) {
                        Im2colGeo e{};
                        e.in_bs = -p.in_sel;                 // placeholder: which input (resolved per launch)
                        e.H = p.H; e.W = p.W; e.kh = p.kh; e.kw = p.kw; e.sh = p.sh; e.sw = p.sw; e.Ho = p.Ho; e.Wo = p.Wo;
                        e.KK = p.kh * p.kw; e.Kp = (int)align_up(e.KK, 8);
                        e.out = (__half*)nullptr + geo_elems;    // offset for now, fixed up after allocation
                        geo_elems += align_up((long long)MB * e.Ho * e.Wo * e.Kp, 64);
                        h->geos.push_back(e);
                        gi = (int)h->geos.size() - 1;
                    }
                    Prob s2 = p;
                    const Im2colGeo& e = h->geos[gi];
                    s2.in_sel = 0; s2.in = (const __half*)nullptr + (e.out - (__half*)nullptr);   // fixed up below
                    s2.in_bs = e.Ho * e.Wo * e.Kp;
                    s2.H = e.Ho * e.Wo; s2.W = 1; s2.C = e.KK; s2.Cs = e.Kp; s2.Ho = s2.H; s2.Wo = 1;
                    s2.kh = s2.kw = s2.sh = s2.sw = 1;
                    s2.coff = gi;                            // geometry index, used for the fix-up
                    s2.Cj = e.KK < e.Kp ? 1 : 0;             // the matrix has a ones column: bias gradient from the weight-gradient GEMM
                    const int cF = tc_cfg_for(s2.F), cC = tc_cfg_for(s2.C);
                    auto key2 = std::make_tuple(to.depth, kind, cF + 16 * 1 + 64 * 1, cC, ordinal, ind.lane);
                    Group& g2 = gmap[key2];
                    g2.lane = ind.lane;
                    g2.depth = to.depth; g2.kind = kind; g2.cfgF = cF; g2.cfgC = cC; g2.ordinal = ordinal; g2.tc = 1; g2.mode = 1;
                    g2.h.push_back(s2);
                    mode = 2;
                }
                auto key = std::make_tuple(to.depth, kind, cfgF + 16 * tc + 64 * mode, cfgC, ordinal, ind.lane);
                Group& g = gmap[key];
                g.lane = ind.lane;
                g.depth = to.depth; g.kind = kind; g.cfgF = cfgF; g.cfgC = cfgC; g.ordinal = ordinal; g.tc = tc; g.mode = mode;
                g.h.push_back(p);
            }
        }
        Prob lp{};
        lp.indiv = i;
        lp.out = ind.tensors.back().act; lp.gout = ind.tensors.back().grad; lp.out_bs = (int)ind.tensors.back().bs;
        h->h_loss.push_back(lp);
    }
    h->groups.clear();
    for (auto& kv : gmap) h->groups.push_back(std::move(kv.second));   // std::map iterates in key order
    if (geo_elems > 0) {
        CK(buf_reserve(h->bIm2col, (size_t)geo_elems * sizeof(__half), 1.25));
        h->im2col = h->bIm2col.as<__half>();
        CK(cudaMemsetAsync(h->im2col, 0, (size_t)geo_elems * sizeof(__half), h->stream));
        h->device_bytes += geo_elems * 2;
        for (auto& e : h->geos) e.out = h->im2col + (e.out - (__half*)nullptr);
        for (auto& g : h->groups)
            if (g.mode == 1)
                for (auto& q : g.h) { q.in = h->geos[q.coff].out; q.coff = 0; }
    }
    // backward walks the depth levels from the deepest; inside a level the buckets of first consumers (ordinal 0) run
    // first, concurrently, then the others one by one (see backward_pass).  In that order the first writer of each
    // input gradient overwrites and every later one adds.
    {
        std::map<const __half*, bool> written;
        auto visit = [&](Group& g) {
            for (auto& p : g.h) {
                if (!p.din) continue;
                bool& w = written[p.din];
                p.din_acc = w ? 1 : 0;
                w = true;
            }
        };
        int gi = (int)h->groups.size() - 1;
        while (gi >= 0) {
            int ge = gi;
            std::vector<Group*> later;
            while (ge >= 0 && h->groups[ge].depth == h->groups[gi].depth) {
                Group& g = h->groups[ge];
                if (g.mode & 1) { if (g.ordinal == 0) visit(g); else later.push_back(&g); }
                --ge;
            }
            for (Group* g : later) visit(*g);
            gi = ge;
        }
    }
    for (auto& g : h->groups) {
        if (g.kind == SRN_KIND_BN) {
            POOL_ALLOC(g.done, int, g.h.size() * sizeof(int));
            CK(cudaMemsetAsync(g.done, 0, g.h.size() * sizeof(int), h->stream));
        }
        POOL_ALLOC(g.d, Prob, g.h.size() * sizeof(Prob));
        CK(cudaMemcpyAsync(g.d, g.h.data(), g.h.size() * sizeof(Prob), cudaMemcpyHostToDevice, h->stream));
    }
    POOL_ALLOC(h->d_loss, Prob, n * sizeof(Prob));
    CK(cudaMemcpyAsync(h->d_loss, h->h_loss.data(), n * sizeof(Prob), cudaMemcpyHostToDevice, h->stream));
    {
        // the same problems once more, lane by lane (the loss kernel of a lane covers only its own individuals)
        std::vector<Prob> by_lane;
        Prob* d_by_lane = nullptr;
        POOL_ALLOC(d_by_lane, Prob, n * sizeof(Prob));
        for (int l = 0; l < srn_engine::NLANES; ++l) {
            h->lane[l].d_loss = d_by_lane + by_lane.size();
            for (int i = 0; i < n; ++i) if (h->inds[i].lane == l) by_lane.push_back(h->h_loss[(size_t)i]);
            h->lane[l].n_loss = (int)(d_by_lane + by_lane.size() - h->lane[l].d_loss);
        }
        CK(cudaMemcpyAsync(d_by_lane, by_lane.data(), n * sizeof(Prob), cudaMemcpyHostToDevice, h->stream));
        CK(cudaStreamSynchronize(h->stream));      // by_lane goes out of scope
    }
    POOL_ALLOC(h->d_lb, float, n * sizeof(float));
    CK(cudaMemcpyAsync(h->d_lb, loss_balance, n * sizeof(float), cudaMemcpyHostToDevice, h->stream));
    // ---- Keras default initialisation on the device
    std::vector<long long> pad_off; std::vector<int> pad_len;     // kernel rows to clear afterwards (padded concatenations)
    for (int i = 0; i < n; ++i) {
        Indiv& ind = h->inds[i];
        for (size_t l = 0; l < ind.layers.size(); ++l) {
            const LayerRt& L = ind.layers[l];
            const srn_layer_desc& d = L.d;
            const uint64_t tseed = seeds[i] + l;
            if (d.kind == SRN_KIND_DENSE || d.kind == SRN_KIND_CONV) {
                const int rf = d.kind == SRN_KIND_CONV ? d.kh * d.kw : 1;
                const double fan_in = (double)d.in_c * rf;
                const bool heads = (l + 1 == ind.layers.size());
                float la, lbm; int split = d.units;
                if (heads) {      // the fused head stands for Dense(n_classes) and Dense(genotype_size)
                    la = (float)std::sqrt(6.0 / (fan_in + h->ncls));
                    lbm = (float)std::sqrt(6.0 / (fan_in + h->G));
                    split = h->ncls;
                } else {
                    la = lbm = (float)std::sqrt(6.0 / (fan_in + (double)d.units * rf));
                }
                LAUNCH(SRN_CAT_OTHER, k_init_uniform, (unsigned)((L.w_cnt + 255) / 256), 256, h->P + L.w_off, L.w_cnt, d.units, L.ldw, split, la, lbm, tseed);
                if (!L.rowmap.empty()) {      // kernel rows of pad positions are zero and stay zero (zero activations -> zero gradient -> no update)
                    std::vector<char> live((size_t)L.w_rows, 0);
                    for (int r : L.rowmap) live[(size_t)r] = 1;
                    for (int r = 0; r < L.w_rows; ++r)
                        if (!live[(size_t)r]) { pad_off.push_back(L.w_off + (long long)r * L.ldw); pad_len.push_back(L.ldw); }
                }
            } else if (d.kind == SRN_KIND_BN) {
                LAUNCH(SRN_CAT_OTHER, k_fill_f16, (unsigned)((L.w_cnt + 255) / 256), 256, h->P + L.w_off, L.w_cnt, 1.0f);          // gamma
                LAUNCH(SRN_CAT_OTHER, k_fill_f16, (unsigned)((d.in_c + 255) / 256), 256, h->S + L.mov_off + d.in_c, (long long)d.in_c, 1.0f);  // moving variance
            }
        }
    }
    long long* d_pad_off = nullptr; int* d_pad_len = nullptr;
    if (!pad_off.empty()) {
        POOL_ALLOC(d_pad_off, long long, pad_off.size() * sizeof(long long));
        POOL_ALLOC(d_pad_len, int, pad_len.size() * sizeof(int));
        CK(cudaMemcpyAsync(d_pad_off, pad_off.data(), pad_off.size() * sizeof(long long), cudaMemcpyHostToDevice, h->stream));
        CK(cudaMemcpyAsync(d_pad_len, pad_len.data(), pad_len.size() * sizeof(int), cudaMemcpyHostToDevice, h->stream));
        LAUNCH(SRN_CAT_OTHER, k_zero_spans, (unsigned)pad_off.size(), 64, h->P, d_pad_off, d_pad_len);
    }
    CK(cudaStreamSynchronize(h->stream));
    CK(cudaGetLastError());
    h->n = n;
    h->adam_t = 0;
    h->built = true;
    return SRN_OK;
}

extern "C" int64_t srn_parameters_count(srn_engine* h, int32_t i) {
    if (!h || !h->built || i < 0 || i >= h->n) return -1;
    return h->inds[i].pcount;
}

extern "C" int64_t srn_device_bytes(srn_engine* h) { return h ? h->device_bytes : -1; }

// weight `which` of a layer as a [rows][cols] matrix with leading dimension ld inside the arena
static int locate_weights(srn_engine* h, int ind, int layer, int which, __half** ptr, int64_t* cnt, int* rows, int* cols, int* ld,
                          const std::vector<int>** rowmap = nullptr) {
    if (!h || !h->built) return fail(SRN_ERR_STATE, "weights: engine has no population");
    if (ind < 0 || ind >= h->n) return fail(SRN_ERR_ARG, "weights: individual %d", ind);
    Indiv& I = h->inds[ind];
    if (layer < 0 || layer >= (int)I.layers.size()) return fail(SRN_ERR_ARG, "weights: layer %d", layer);
    const LayerRt& L = I.layers[layer];
    const int kind = L.d.kind;
    *ptr = nullptr;
    if (rowmap) *rowmap = nullptr;
    if (kind == SRN_KIND_DENSE || kind == SRN_KIND_CONV) {
        // a Dense on a padded concatenation keeps zero rows at the pad positions: callers see the Keras rows through rowmap
        if (which == 0) { *ptr = h->P + L.w_off; *rows = L.w_rows_log; *cols = L.w_cols; *ld = L.ldw; if (rowmap && !L.rowmap.empty()) *rowmap = &L.rowmap; }
        if (which == 1) { *ptr = h->P + L.b_off; *rows = 1; *cols = (int)L.b_cnt; *ld = (int)L.b_cnt; }
    } else if (kind == SRN_KIND_BN) {
        const int C = L.d.in_c;
        *rows = 1; *cols = C; *ld = C;
        if (which == 0) *ptr = h->P + L.w_off;
        if (which == 1) *ptr = h->P + L.b_off;
        if (which == 2) *ptr = h->S + L.mov_off;
        if (which == 3) *ptr = h->S + L.mov_off + C;
    }
    if (!*ptr) return fail(SRN_ERR_ARG, "weights: layer %d (kind %d) has no weight %d", layer, kind, which);
    *cnt = (int64_t)*rows * *cols;
    return SRN_OK;
}

// rows of a [rows][cols] matrix between the host (dense) and the arena (leading dimension ld, optional row map)
static int copy_rows(void* host, char* dev, int rows, int cols, int ld, size_t esz, const std::vector<int>* rowmap, bool to_host) {
    if (!rowmap) {
        if (to_host) CK(cudaMemcpy2D(host, cols * esz, dev, ld * esz, cols * esz, rows, cudaMemcpyDeviceToHost));
        else CK(cudaMemcpy2D(dev, ld * esz, host, cols * esz, cols * esz, rows, cudaMemcpyHostToDevice));
        return SRN_OK;
    }
    int prows = 0;
    for (int r : *rowmap) prows = std::max(prows, r + 1);
    std::vector<char> tmp((size_t)prows * cols * esz, 0);
    if (to_host) {
        CK(cudaMemcpy2D(tmp.data(), cols * esz, dev, ld * esz, cols * esz, prows, cudaMemcpyDeviceToHost));
        for (int r = 0; r < rows; ++r) memcpy((char*)host + (size_t)r * cols * esz, tmp.data() + (size_t)(*rowmap)[r] * cols * esz, cols * esz);
    } else {
        // pad rows are rewritten with zeros, which is what they hold anyway
        for (int r = 0; r < rows; ++r) memcpy(tmp.data() + (size_t)(*rowmap)[r] * cols * esz, (const char*)host + (size_t)r * cols * esz, cols * esz);
        CK(cudaMemcpy2D(dev, ld * esz, tmp.data(), cols * esz, cols * esz, prows, cudaMemcpyHostToDevice));
    }
    return SRN_OK;
}

extern "C" int srn_get_weights(srn_engine* h, int32_t ind, int32_t layer, int32_t which, uint16_t* buf, int64_t* count) {
    __half* p; int64_t c; int rows, cols, ld; const std::vector<int>* rm;
    int rc = locate_weights(h, ind, layer, which, &p, &c, &rows, &cols, &ld, &rm);
    if (rc) return rc;
    if (count) *count = c;
    if (buf) {
        CK(cudaSetDevice(h->cfg.device));
        CK(cudaStreamSynchronize(h->stream));
        CK(cudaStreamSynchronize(h->adam_stream));
        return copy_rows(buf, reinterpret_cast<char*>(p), rows, cols, ld, sizeof(__half), rm, true);
    }
    return SRN_OK;
}

extern "C" int srn_set_weights(srn_engine* h, int32_t ind, int32_t layer, int32_t which, const uint16_t* buf, int64_t count) {
    __half* p; int64_t c; int rows, cols, ld; const std::vector<int>* rm;
    int rc = locate_weights(h, ind, layer, which, &p, &c, &rows, &cols, &ld, &rm);
    if (rc) return rc;
    if (!buf || count != c) return fail(SRN_ERR_ARG, "set_weights: expected %lld elements, got %lld", (long long)c, (long long)count);
    CK(cudaSetDevice(h->cfg.device));
    CK(cudaStreamSynchronize(h->stream));
    CK(cudaStreamSynchronize(h->adam_stream));
    return copy_rows(const_cast<uint16_t*>(buf), reinterpret_cast<char*>(p), rows, cols, ld, sizeof(__half), rm, false);
}

extern "C" int srn_debug_set_tensor_cores(int32_t on) {
    g_tensor_cores = on ? 1 : 0;
    return SRN_OK;
}

// ------------------------------------------------------------------ launch plan per batch size
// ------------------------------------------------------------------ TMA tensor maps of Dense problems
// The two maps (A operand, B operand) of one pass of a Dense problem whose operands are plain matrices; false = gather path
static bool dense_tensor_maps(const Prob& p, int nb, int mode, int BN, CUtensorMap* m) {
    static const int mask = getenv("SRN_TMA_MASK") ? atoi(getenv("SRN_TMA_MASK")) : 7;     // bit per pass (fwd, dgrad, wgrad): experiments
    if (!((mask >> mode) & 1)) return false;
    if (p.kh * p.kw != 1 || p.sh != 1 || p.sw != 1 || p.in_sel != 0) return false;
    const int HW = p.H * p.W, HoWo = p.Ho * p.Wo;
    if (HW != HoWo) return false;
    const bool a_ok = HW == 1 || p.in_bs == HW * p.Cs, g_ok = HoWo == 1 || p.out_bs == HoWo * p.Fs;
    const long long M = (long long)nb * HW;
    const long long a_pitch = HW == 1 ? p.in_bs : p.Cs, g_pitch = HoWo == 1 ? p.out_bs : p.Fs;
    if (mode == TC_FWD && tc_fwd_pack(p, nb, false) > 1) return false;      // packed short reductions: gathered rows, several sub-tiles per stage
    if (mode == TC_FWD)
        return a_ok && encode_matrix(&m[0], p.in, M, p.Cs, a_pitch, 128) && encode_matrix(&m[1], p.w, p.C, p.ldw, p.ldw, 64);
    if (mode == TC_DGRAD)
        return g_ok && encode_matrix(&m[0], p.gout, M, p.Fs, g_pitch, 128) && encode_matrix(&m[1], p.w, p.C, p.ldw, p.ldw, BN);
    return a_ok && g_ok && encode_matrix(&m[0], p.in, M, p.Cs, a_pitch, 64) && encode_matrix(&m[1], p.gout, M, p.Fs, g_pitch, 64);
}

static int tc_grid(srn_engine* h, int total, int BN);
// Forward by the slab convolution kernel (srn_conv_slab.cuh): stride-1 windows over a dense tensor whose pixels are one
// swizzle row, weights that fit the resident region.  Such problems get no tiles in the bucket's GEMM tile table.
// OFF by default (SRN_SLAB=1 turns it on; parity-tested in that mode): on the benchmark population the eligible layers
// cost 0.19 ms of the 3.0 ms forward GEMM time on the gather path and 0.36 ms here - with N <= 64 an MMA is 16-32 tensor
// cycles, one issuing warp per CTA does not keep up, and the resident weights (128-byte rows) leave room for one CTA per
// SM only.  The layers that dominate (7 x 7 x 64 windows, 8-channel inputs, strides) are not eligible yet.
static bool slab_ok(const Group& g, const Prob& p) {
    static const bool on = getenv("SRN_SLAB") != nullptr;
    return on && g.tc == 1 && g.kind == SRN_KIND_CONV && p.kh * p.kw > 1 && p.sh == 1 && p.sw == 1 && p.in_sel == 0 && p.C == p.Cs &&
           (p.C == 16 || p.C == 32 || p.C == 64) && p.F <= 64 && p.in_bs == p.H * p.W * p.Cs && p.kh * p.kw * p.C <= CS_KMAX &&
           p.Ho == p.H - p.kh + 1 && p.Wo == p.W - p.kw + 1 && (p.kh - 1) * p.W + p.kw - 1 <= 384 && encode_tiled_fn() != nullptr;
}
// Target tile length (k-blocks) of the forward launch this bucket runs in = half the mean load of a CTA over all the
// tensor-core buckets of its depth level with the same N tile (they are issued as one launch).  Long reductions are split
// down to it (tc_fwd_split).  (Every split costs an fp32 atomic pass over its output tile: "four tiles per SM" for every
// splittable problem measured 0.13 ms per step slower than one; no splits at all 0.6 ms more serialised time.  Unsplit,
// single tiles of 140-230 k-blocks left their CTA with 2.4-3.9 times the mean load of the launch.)
static int group_tstar(srn_engine* h, const Group& g, int nb) {
    if (g.tc != 1) return 0;
    static const int fixed = getenv("SRN_SPLIT_TSTAR") ? atoi(getenv("SRN_SPLIT_TSTAR")) : 0;      // A/B switch
    if (fixed) return fixed;
    const int BN = TC_BN[g.cfgF];
    long long kblocks = 0;
    for (const Group& o : h->groups) {
        if (o.depth != g.depth || o.tc != 1 || o.cfgF != g.cfgF || o.lane != g.lane || !(o.mode & g.mode)) continue;
        if (o.kind != SRN_KIND_DENSE && o.kind != SRN_KIND_CONV) continue;
        for (const Prob& p : o.h) {
            if (slab_ok(o, p)) continue;
            const long long mt = ((long long)nb * p.Ho * p.Wo + TC_BM - 1) / TC_BM, tn = (p.F + BN - 1) / BN, nk = (tc_k_fwd(p) + TC_BK - 1) / TC_BK;
            kblocks += mt * tn * nk;
        }
    }
    const int G = tc_grid(h, 1 << 30, BN);
    return (int)std::max(48LL, kblocks / G / 2);      // (fixed lengths 24 / 40 / 64 / 96 measured 8.22 / 8.17 / 8.14 / 8.13 ms per step: flat above 40)
}

static int pass_items(const Group& g, const Prob& p, int pass, int nb, int smax) {
    const long long in_rows = (long long)nb * p.H * p.W, out_rows = (long long)nb * p.Ho * p.Wo;
    switch (g.kind) {
        case SRN_KIND_DENSE:
        case SRN_KIND_CONV:
            if (pass == P_COL) return col_chunks(out_rows, p.Fs);
            if (g.tc == 3) return pass == P_FWD ? ew_blocks(out_rows, p.Fs) : 0;      // wgrad rides on the bias-gradient pass
            if (g.tc) {
                if (pass == P_FWD && slab_ok(g, p)) return 0;      // forward by the slab kernel (its own table in plan_for)
                if (pass == P_FWD) return tc_tiles_fwd(p, nb, TC_BN[g.cfgF], smax);
                if (pass == P_DGRAD) return p.din ? tc_tiles_dgrad(p, nb, TC_BN[g.cfgC]) : 0;
                if (pass == P_WGRAD) return tc_tiles_wgrad(p, nb, TC_BN[g.cfgF]);
                return 0;
            }
            if (pass == P_FWD) return fwd_tiles(nb * p.Ho * p.Wo, p.F, CFG_BM[g.cfgF], CFG_BN[g.cfgF]);
            if (pass == P_DGRAD) return p.din ? fwd_tiles(nb * p.H * p.W, p.C, CFG_BM[g.cfgC], CFG_BN[g.cfgC]) : 0;
            if (pass == P_WGRAD) return wgrad_tiles(p.kh * p.kw * p.C, p.F, nb * p.Ho * p.Wo);
            return 0;
        case SRN_KIND_MAXPOOL:
            if (pass == P_FWD) return ew_blocks(out_rows, p.Cs);
            if (pass == P_BWD_EW) return p.din ? ew_blocks(in_rows, p.Cs) : 0;
            return 0;
        case SRN_KIND_BN:
            if (pass == P_FWD) return ew_blocks(in_rows, p.Cs);
            if (pass == P_COL) return col_chunks(in_rows, p.Cs);
            if (pass == P_BWD_EW) return p.din ? ew_blocks(in_rows, p.Cs) : 0;
            return 0;
        case SRN_KIND_CONCAT:
            if (pass == P_FWD) return copy_blocks_fwd(p, nb);
            if (pass == P_BWD_EW) return p.din ? copy_blocks(p, nb) : 0;
            return 0;
    }
    return 0;
}

// heaviest tiles first, ties in table order: sorts 4-byte indices and moves every 320-byte descriptor once (sorting the
// descriptors themselves was most of the host time of a launch plan)
static void sort_tiles_by_cost(std::vector<TcTile>& tt) {
    std::vector<uint32_t> order(tt.size());
    for (size_t i = 0; i < order.size(); ++i) order[i] = (uint32_t)i;
    std::stable_sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) { return tt[a].nk * tt[a].nsub > tt[b].nk * tt[b].nsub; });
    std::vector<TcTile> sorted(tt.size());
    for (size_t i = 0; i < order.size(); ++i) sorted[i] = tt[order[i]];
    tt.swap(sorted);
}

static int plan_for(srn_engine* h, Group& g, int nb, Cached** out) {
    auto it = g.cache.find(nb);
    if (it != g.cache.end()) { *out = &it->second; return SRN_OK; }
    Cached c;
    const int np = (int)g.h.size();
    const int smax = group_tstar(h, g, nb);      // forward split-K: target tile length
    std::vector<int> host((size_t)P_COUNT * (np + 1));
    for (int pass = 0; pass < P_COUNT; ++pass) {
        int* pre = host.data() + (size_t)pass * (np + 1);
        long long acc = 0;
        for (int i = 0; i < np; ++i) { pre[i] = (int)acc; acc += pass_items(g, g.h[i], pass, nb, smax); }
        if (acc > 0x7fffffffLL) return fail(SRN_ERR_LIMIT, "launch grid too large");
        pre[np] = (int)acc;
        c.total[pass] = (int)acc;
    }
    int* dev = nullptr;
    POOL_ALLOC(dev, int, host.size() * sizeof(int));
    CK(cudaMemcpyAsync(dev, host.data(), host.size() * sizeof(int), cudaMemcpyHostToDevice, h->stream));
    CK(cudaStreamSynchronize(h->stream));      // host vector goes out of scope
    for (int pass = 0; pass < P_COUNT; ++pass) c.prefix[pass] = dev + (size_t)pass * (np + 1);
    if (g.tc == 1 || g.tc == 2) {
        if (g.tc == 1 && g.kind == SRN_KIND_CONV) {
            std::vector<SlabProb> sp;
            std::vector<int> spre(1, 0);
            const int MB = h->cfg.max_batch;
            for (int i = 0; i < np; ++i) {
                const Prob& p = g.h[i];
                if (!slab_ok(g, p)) continue;
                SlabProb q;
                memset(&q, 0, sizeof(q));
                q.bias = p.bias; q.out = p.out; q.H = p.H; q.W = p.W; q.Wo = p.Wo; q.P = (p.Ho - 1) * p.W + p.Wo; q.tps = (q.P + TC_BM - 1) / TC_BM;
                q.C = p.C; q.F = p.F; q.Fs = p.Fs; q.K = p.kh * p.kw * p.C; q.kw = p.kw; q.out_bs = p.out_bs; q.act = p.act;
                q.rowb = p.C * 2; q.layout = p.C == 16 ? 6 : (p.C == 32 ? 4 : 2); q.sbo = 8 * q.rowb;
                const int rows = TC_BM + (p.kh - 1) * p.W + p.kw - 1;
                q.nbox = rows <= 256 ? 1 : 2;
                q.box_rows = (int)align_up((rows + q.nbox - 1) / q.nbox, 8);
                q.slab_bytes = q.nbox * q.box_rows * q.rowb;
                c.slab_stride = std::max(c.slab_stride, (int)align_up(q.slab_bytes, 1024));
                c.slab_wbytes = std::max(c.slab_wbytes, ((q.K + 63) / 64) * CS_WCHUNK);
                if (!encode_slab(&q.amap, p.in, (long long)MB * p.H * p.W, p.C, q.box_rows) || !encode_matrix(&q.wmap, p.w, q.K, p.ldw, p.ldw, 64))
                    return fail(SRN_ERR_CUDA, "slab convolution: tensor map encoding failed");
                sp.push_back(q);
                spre.push_back(spre.back() + nb * q.tps);
            }
            if (!sp.empty()) {
                c.slab_np = (int)sp.size(); c.slab_total = spre.back();
                POOL_ALLOC(c.slab, SlabProb, sp.size() * sizeof(SlabProb));
                CK(cudaMemcpy(c.slab, sp.data(), sp.size() * sizeof(SlabProb), cudaMemcpyHostToDevice));
                POOL_ALLOC(c.slab_prefix, int, spre.size() * sizeof(int));
                CK(cudaMemcpy(c.slab_prefix, spre.data(), spre.size() * sizeof(int), cudaMemcpyHostToDevice));
            }
        }
        for (int pass : {P_FWD, P_DGRAD, P_WGRAD}) {
            if (!c.total[pass]) continue;
            const int* pre = host.data() + (size_t)pass * (np + 1);
            const int mode = pass == P_FWD ? TC_FWD : (pass == P_DGRAD ? TC_DGRAD : TC_WGRAD);
            const int BN = TC_BN[pass == P_DGRAD ? g.cfgC : g.cfgF];
            std::vector<TcTile> tt((size_t)c.total[pass]);
            for (int i = 0; i < np; ++i)
                for (int t = pre[i]; t < pre[i + 1]; ++t) {
                    memset(&tt[t], 0, sizeof(TcTile));
                    tt[t].p = g.h[i];
                    tc_describe_tile(tt[t], t - pre[i], nb, BN, mode, g.tc == 2, smax);
                }
            if (pass == P_FWD && g.tc == 1) {
                // split-K scratch: zeroed once, every launch leaves it zeroed (srn_gemm_tc.cuh)
                for (int i = 0; i < np; ++i) {
                    if (tc_fwd_split(g.h[i], nb, BN, smax) <= 1) continue;
                    const Prob& p = g.h[i];
                    const size_t rows = (size_t)nb * p.Ho * p.Wo;
                    const size_t acc_bytes = rows * p.Fs * sizeof(float);
                    char* buf = nullptr;
                    POOL_ALLOC(buf, char, acc_bytes);
                    CK(cudaMemset(buf, 0, acc_bytes));
                    h->device_bytes += (long long)acc_bytes;
                    for (int t = pre[i]; t < pre[i + 1]; ++t) tt[t].sk_acc = reinterpret_cast<float*>(buf);
                    SplitFin f;
                    f.acc = reinterpret_cast<float*>(buf); f.out = p.out; f.bias = p.bias; f.rows = (int)rows; f.Fs = p.Fs; f.F = p.F; f.act = p.act;
                    f.HoWo = p.Ho * p.Wo; f.out_bs = p.out_bs; f.blk0 = c.fin_blocks;
                    c.fin_blocks += (int)((rows * (size_t)(p.Fs / 8) + 255) / 256);
                    c.host_fin.push_back(f);
                }
                if (!c.host_fin.empty()) {
                    POOL_ALLOC(c.fin, SplitFin, c.host_fin.size() * sizeof(SplitFin));
                    CK(cudaMemcpy(c.fin, c.host_fin.data(), c.host_fin.size() * sizeof(SplitFin), cudaMemcpyHostToDevice));
                }
            }
            if (g.tc == 1) {        // any problem with a 1 x 1 window over row-major operands (Dense, first layers over the shared im2col matrix)
                std::vector<CUtensorMap> maps;
                std::vector<int> first(np, -1);
                CUtensorMap pair[2];
                for (int i = 0; i < np; ++i)
                    if (pre[i + 1] > pre[i] && dense_tensor_maps(g.h[i], nb, mode, BN, pair)) {
                        first[i] = (int)maps.size();
                        maps.push_back(pair[0]); maps.push_back(pair[1]);
                    }
                if (!maps.empty()) {
                    CUtensorMap* dm = nullptr;
                    POOL_ALLOC(dm, CUtensorMap, maps.size() * sizeof(CUtensorMap));
                    CK(cudaMemcpy(dm, maps.data(), maps.size() * sizeof(CUtensorMap), cudaMemcpyHostToDevice));
                    for (int i = 0; i < np; ++i)
                        if (first[i] >= 0)
                            for (int t = pre[i]; t < pre[i + 1]; ++t) tt[t].tm = dm + first[i];
                }
            }
            sort_tiles_by_cost(tt);
            TcTile* dt = nullptr;
            POOL_ALLOC(dt, TcTile, tt.size() * sizeof(TcTile));
            CK(cudaMemcpy(dt, tt.data(), tt.size() * sizeof(TcTile), cudaMemcpyHostToDevice));
            c.tiles[pass] = dt;
            if (pass == P_FWD && g.tc == 1) c.host_fwd = tt;
        }
    }
    g.cache[nb] = c;
    *out = &g.cache[nb];
    return SRN_OK;
}

// persistent tensor-core launch: as many CTAs as fit the machine (shared memory bound), each walking the tile table
static int tc_grid(srn_engine* h, int total, int BN) {
    const int per_sm = std::max(1, std::min(4, (228 * 1024) / (tc_smem_bytes(BN) + 1024)));      // 228 KB per SM, 1 KB reserved per CTA
    return std::min(total, h->num_sms * per_sm);
}
#define LAUNCH_TC1(cat, MODE, SCALAR, pass, BN_)                                                   \
    LAUNCH_SMEM(cat, (k_gemm_tc<MODE, SCALAR, BN_>), tc_grid(h, c->total[pass], BN_), TC_THREADS, tc_smem_bytes(BN_), \
                c->tiles[pass], c->total[pass], nb, I)
#define LAUNCH_TC(cat, MODE, SCALAR, pass, BN_)                                                    \
    do {                                                                                           \
        switch (BN_) {                                                                             \
            case 32: LAUNCH_TC1(cat, MODE, SCALAR, pass, 32); break;                               \
            case 64: LAUNCH_TC1(cat, MODE, SCALAR, pass, 64); break;                               \
            case 128: LAUNCH_TC1(cat, MODE, SCALAR, pass, 128); break;                             \
            default: LAUNCH_TC1(cat, MODE, SCALAR, pass, 256); break;                              \
        }                                                                                          \
    } while (0)

template <int CFG>
static void launch_fwd_cfg(srn_engine* h, Group& g, Cached* c, int nb, const Inputs& I) {
    LAUNCH(SRN_CAT_GEMM_FWD, (k_conv_fwd<(CFG == 0 ? 64 : CFG == 1 ? 128 : 256), (CFG == 0 ? 64 : CFG == 1 ? 32 : 16)>), c->total[P_FWD], SIMT_THREADS,
           g.d, c->prefix[P_FWD], (int)g.h.size(), nb, I);
}
template <int CFG>
static void launch_dgrad_cfg(srn_engine* h, Group& g, Cached* c, int nb) {
    LAUNCH(SRN_CAT_GEMM_DGRAD, (k_conv_dgrad<(CFG == 0 ? 64 : CFG == 1 ? 128 : 256), (CFG == 0 ? 64 : CFG == 1 ? 32 : 16)>), c->total[P_DGRAD], SIMT_THREADS,
           g.d, c->prefix[P_DGRAD], (int)g.h.size(), nb);
}

// Runs independent launch buckets: bucket 0 on the main stream, the others round-robin on the side streams, fenced
// by a fork event before and join events after.
template <class F>
static int run_level(srn_engine* h, std::vector<Group*>& level, F&& body) {
    if (level.empty()) return SRN_OK;
    const bool fan = h->multi_stream && level.size() > 1;
    if (fan) {
        cudaEventRecord(h->L->ev_fork, h->L->main);
        for (int i = 0; i < srn_engine::NSIDE; ++i) cudaStreamWaitEvent(h->L->side[i], h->L->ev_fork, 0);
    }
    bool used[srn_engine::NSIDE] = {};
    int rc = SRN_OK;
    for (size_t k = 0; k < level.size() && !rc; ++k) {
        const int lane = fan ? (int)(k % (srn_engine::NSIDE + 1)) : 0;
        h->cur = lane == 0 ? h->L->main : h->L->side[lane - 1];
        if (lane) used[lane - 1] = true;
        rc = body(*level[k]);
    }
    h->cur = h->L->main;
    for (int i = 0; i < srn_engine::NSIDE; ++i)
        if (used[i]) { cudaEventRecord(h->L->ev_join[i], h->L->side[i]); cudaStreamWaitEvent(h->L->main, h->L->ev_join[i], 0); }
    return rc;
}

// same fan-out for arbitrary launch closures
static int run_tasks(srn_engine* h, std::vector<std::function<int()>>& tasks) {
    if (tasks.empty()) return SRN_OK;
    const bool fan = h->multi_stream && tasks.size() > 1;
    if (fan) {
        cudaEventRecord(h->L->ev_fork, h->L->main);
        for (int i = 0; i < srn_engine::NSIDE; ++i) cudaStreamWaitEvent(h->L->side[i], h->L->ev_fork, 0);
    }
    bool used[srn_engine::NSIDE] = {};
    int rc = SRN_OK;
    for (size_t k = 0; k < tasks.size() && !rc; ++k) {
        const int lane = fan ? (int)(k % (srn_engine::NSIDE + 1)) : 0;
        h->cur = lane == 0 ? h->L->main : h->L->side[lane - 1];
        if (lane) used[lane - 1] = true;
        rc = tasks[k]();
    }
    h->cur = h->L->main;
    for (int i = 0; i < srn_engine::NSIDE; ++i)
        if (used[i]) { cudaEventRecord(h->L->ev_join[i], h->L->side[i]); cudaStreamWaitEvent(h->L->main, h->L->ev_join[i], 0); }
    return rc;
}

// the bucket's slab-eligible problems (forward only)
static void launch_slab(srn_engine* h, Group& g, Cached* c) {
    if (!c->slab_total) return;
    const int smem = cs_smem_bytes(c->slab_stride, c->slab_wbytes);
    const int per_sm = std::max(1, std::min(2, (228 * 1024) / (smem + 1024)));      // two CTAs per SM when slabs and weights are small
    const int grid = std::min(c->slab_total, h->num_sms * per_sm);
    if (TC_BN[g.cfgF] == 32)
        LAUNCH_SMEM(SRN_CAT_GEMM_FWD, (k_conv_slab<32>), grid, CS_THREADS, smem, c->slab, c->slab_prefix, c->slab_np, c->slab_total, c->slab_stride, c->slab_wbytes);
    else
        LAUNCH_SMEM(SRN_CAT_GEMM_FWD, (k_conv_slab<64>), grid, CS_THREADS, smem, c->slab, c->slab_prefix, c->slab_np, c->slab_total, c->slab_stride, c->slab_wbytes);
}

static int fwd_group(srn_engine* h, Group& g, int nb, const Inputs& I, bool training) {
    Cached* c;
    h->cur_group = (int)(&g - h->groups.data());
    int rc = plan_for(h, g, nb, &c);
    if (rc) return rc;
    const int np = (int)g.h.size();
    switch (g.kind) {
        case SRN_KIND_DENSE:
        case SRN_KIND_CONV:
            if (g.tc == 3) {
                LAUNCH(SRN_CAT_GEMM_FWD, k_dense_k1_fwd, c->total[P_FWD], EW_THREADS, g.d, c->prefix[P_FWD], np, nb, I);
                break;
            }
            if (g.tc) {
                const int BN = TC_BN[g.cfgF];
                if (g.tc == 1) {
                    launch_slab(h, g, c);
                    LAUNCH_TC(SRN_CAT_GEMM_FWD, TC_FWD, false, P_FWD, BN);
                    if (c->fin_blocks) LAUNCH(SRN_CAT_GEMM_FWD, k_splitk_finish, c->fin_blocks, 256, c->fin, (int)c->host_fin.size());
                }
                else LAUNCH_TC(SRN_CAT_GEMM_FWD, TC_FWD, true, P_FWD, BN);
                break;
            }
            if (g.cfgF == 0) launch_fwd_cfg<0>(h, g, c, nb, I);
            else if (g.cfgF == 1) launch_fwd_cfg<1>(h, g, c, nb, I);
            else launch_fwd_cfg<2>(h, g, c, nb, I);
            break;
        case SRN_KIND_MAXPOOL:
            LAUNCH(SRN_CAT_POOL, k_maxpool_fwd, c->total[P_FWD], EW_THREADS, g.d, c->prefix[P_FWD], np, nb, I);
            break;
        case SRN_KIND_BN:
            if (training) {
                LAUNCH(SRN_CAT_BN, k_bn_partial, c->total[P_COL], COL_THREADS, g.d, c->prefix[P_COL], np, nb, I, 0, g.done);
            } else {
                LAUNCH(SRN_CAT_BN, k_bn_prepare_infer, np, 128, g.d);
            }
            LAUNCH(SRN_CAT_BN, k_bn_apply, c->total[P_FWD], EW_THREADS, g.d, c->prefix[P_FWD], np, nb, I);
            break;
        case SRN_KIND_CONCAT:
            LAUNCH(SRN_CAT_CONCAT, k_copy, c->total[P_FWD], EW_THREADS, g.d, c->prefix[P_FWD], np, nb, I, 0);
            break;
    }
    return SRN_OK;
}

// the main stream must not touch parameters or gradient accumulators before the last optimiser pass is done
static void wait_adam(srn_engine* h) {
    if (h->adam_pending) { cudaStreamWaitEvent(h->stream, h->ev_adam_done, 0); h->adam_pending = false; }
}

// shared im2col matrices of the first layers (one per window geometry), written once per batch on lane 0's streams
static void shared_im2col(srn_engine* h, int nb, const Inputs& I) {
    const int mode = (I.x_istride == 0 && I.g_istride == 0) ? 1 : 2;
    if (mode == 1 && !h->geos.empty()) {
        // one small launch per window geometry: independent, so they go out on all streams at once
        const bool fan = h->multi_stream && h->geos.size() > 1;
        if (fan) {
            cudaEventRecord(h->L->ev_fork, h->L->main);
            for (int i = 0; i < srn_engine::NSIDE; ++i) cudaStreamWaitEvent(h->L->side[i], h->L->ev_fork, 0);
        }
        bool used[srn_engine::NSIDE] = {};
        for (size_t k = 0; k < h->geos.size(); ++k) {
            Im2colGeo e = h->geos[k];
            const bool img = e.in_bs == -1;
            e.in = img ? I.x : I.g; e.in_bs = img ? I.x_bs : I.g_bs;
            const long long total = (long long)nb * e.Ho * e.Wo * (e.Kp >> 3);
            const int lane = fan ? (int)(k % (srn_engine::NSIDE + 1)) : 0;
            h->cur = lane == 0 ? h->L->main : h->L->side[lane - 1];
            if (lane) used[lane - 1] = true;
            LAUNCH(SRN_CAT_GATHER, k_im2col, (unsigned)((total + 255) / 256), 256, e, nb);
        }
        h->cur = h->L->main;
        for (int i = 0; i < srn_engine::NSIDE; ++i)
            if (used[i]) { cudaEventRecord(h->L->ev_join[i], h->L->side[i]); cudaStreamWaitEvent(h->L->main, h->L->ev_join[i], 0); }
    }
}

// lane < 0: every bucket (inference passes); else the buckets of that lane's individuals, on the current lane context
static int forward_pass(srn_engine* h, int nb, const Inputs& I, bool training, int lane = -1) {
    const int mode = (I.x_istride == 0 && I.g_istride == 0) ? 1 : 2;
    if (lane < 0) { shared_im2col(h, nb, I); wait_adam(h); }
    // buckets of one depth level are independent: spread them over the side streams so that the tail of one
    // persistent launch overlaps the ramp-up of the next; levels are separated by events
    size_t gi = 0;
    while (gi < h->groups.size()) {
        size_t ge = gi;
        std::vector<Group*> level;
        while (ge < h->groups.size() && h->groups[ge].depth == h->groups[gi].depth) {
            if ((h->groups[ge].mode & mode) && (lane < 0 || h->groups[ge].lane == lane)) level.push_back(&h->groups[ge]);
            ++ge;
        }
        static const bool merge = getenv("SRN_NO_MERGE_FWD") == nullptr;
        int rc = SRN_OK;
        if (!merge) {
            rc = run_level(h, level, [&](Group& g) { return fwd_group(h, g, nb, I, training); });
        } else {
            // tensor-core GEMM buckets with the same N tile run the same kernel: one launch over their joined tile tables
            std::vector<std::function<int()>> tasks;
            std::vector<Group*> by_bn[4];
            for (Group* g : level) {
                if ((g->kind == SRN_KIND_DENSE || g->kind == SRN_KIND_CONV) && g->tc == 1) by_bn[g->cfgF].push_back(g);
                else tasks.push_back([=, &I]() { return fwd_group(h, *g, nb, I, training); });
            }
            for (int b = 0; b < 4; ++b) {
                if (by_bn[b].empty()) continue;
                if (by_bn[b].size() == 1) { Group* g = by_bn[b][0]; tasks.push_back([=, &I]() { return fwd_group(h, *g, nb, I, training); }); continue; }
                auto key = std::make_tuple(nb, by_bn[b][0]->depth, b, mode + 16 * (lane + 1));
                auto it = h->merged_fwd.find(key);
                if (it == h->merged_fwd.end()) {
                    std::vector<TcTile> all;
                    std::vector<SplitFin> fins;
                    int fin_blocks = 0;
                    for (Group* g : by_bn[b]) {
                        Cached* c; rc = plan_for(h, *g, nb, &c); if (rc) return rc;
                        all.insert(all.end(), c->host_fwd.begin(), c->host_fwd.end());
                        for (SplitFin f : c->host_fin) { f.blk0 += fin_blocks; fins.push_back(f); }
                        fin_blocks += c->fin_blocks;
                    }
                    sort_tiles_by_cost(all);
                    if (getenv("SRN_DUMP_FWD")) {      // debug: what a merged forward launch holds
                        const int BNd = TC_BN[b], G = tc_grid(h, (int)all.size(), BNd);
                        long long sum = 0, mx = 0, nsplit_tiles = 0; std::vector<long long> load((size_t)std::max(G, 1), 0);
                        for (size_t t = 0; t < all.size(); ++t) {
                            const long long c = (long long)all[t].nk * all[t].nsub;
                            sum += c; mx = std::max(mx, c); nsplit_tiles += all[t].nsplit > 1;
                            const int j = (int)(t / G), r = (int)(t % G);
                            load[(size_t)((j & 1) ? G - 1 - r : r)] += c;
                        }
                        long long lmax = 0; for (long long v : load) lmax = std::max(lmax, v);
                        fprintf(stderr, "[fwd launch] depth %d BN %d buckets %zu tiles %zu grid %d kblocks %lld max_tile %lld max_cta %lld mean_cta %.1f split_tiles %lld\n",
                                by_bn[b][0]->depth, BNd, by_bn[b].size(), all.size(), G, sum, mx, lmax, G ? (double)sum / G : 0.0, nsplit_tiles);
                        for (Group* g : by_bn[b])
                            for (const Prob& p : g->h)
                                fprintf(stderr, "    bucket %d: H %d W %d C %d -> Ho %d Wo %d F %d k %dx%d s %d  K %d rows %d\n", (int)(g - h->groups.data()), p.H, p.W, p.C, p.Ho, p.Wo, p.F,
                                        p.kh, p.kw, p.sh, p.kh * p.kw * p.Cs, nb * p.Ho * p.Wo);
                    }
                    srn_engine::MergedFwd m;
                    m.total = (int)all.size();
                    m.first_group = (int)(by_bn[b][0] - h->groups.data());
                    if (m.total) {
                        POOL_ALLOC(m.dev, TcTile, all.size() * sizeof(TcTile));
                        CK(cudaMemcpy(m.dev, all.data(), all.size() * sizeof(TcTile), cudaMemcpyHostToDevice));
                    }
                    if (!fins.empty()) {
                        POOL_ALLOC(m.fin, SplitFin, fins.size() * sizeof(SplitFin));
                        CK(cudaMemcpy(m.fin, fins.data(), fins.size() * sizeof(SplitFin), cudaMemcpyHostToDevice));
                        m.nfin = (int)fins.size(); m.fin_blocks = fin_blocks;
                    }
                    it = h->merged_fwd.emplace(key, m).first;
                }
                for (Group* g : by_bn[b]) {          // their slab-eligible problems: one launch per bucket, beside the merged GEMM launch
                    Cached* c; rc = plan_for(h, *g, nb, &c); if (rc) return rc;
                    if (c->slab_total) tasks.push_back([=]() { h->cur_group = (int)(g - h->groups.data()); launch_slab(h, *g, c); return (int)SRN_OK; });
                }
                const srn_engine::MergedFwd m = it->second;
                if (!m.total) continue;
                const int BN = TC_BN[b];
                tasks.push_back([=, &I]() {
                    h->cur_group = m.first_group;
                    switch (BN) {
                        case 32: LAUNCH_SMEM(SRN_CAT_GEMM_FWD, (k_gemm_tc<TC_FWD, false, 32>), tc_grid(h, m.total, 32), TC_THREADS, tc_smem_bytes(32), m.dev, m.total, nb, I); break;
                        case 64: LAUNCH_SMEM(SRN_CAT_GEMM_FWD, (k_gemm_tc<TC_FWD, false, 64>), tc_grid(h, m.total, 64), TC_THREADS, tc_smem_bytes(64), m.dev, m.total, nb, I); break;
                        case 128: LAUNCH_SMEM(SRN_CAT_GEMM_FWD, (k_gemm_tc<TC_FWD, false, 128>), tc_grid(h, m.total, 128), TC_THREADS, tc_smem_bytes(128), m.dev, m.total, nb, I); break;
                        default: LAUNCH_SMEM(SRN_CAT_GEMM_FWD, (k_gemm_tc<TC_FWD, false, 256>), tc_grid(h, m.total, 256), TC_THREADS, tc_smem_bytes(256), m.dev, m.total, nb, I); break;
                    }
                    if (m.fin_blocks) LAUNCH(SRN_CAT_GEMM_FWD, k_splitk_finish, m.fin_blocks, 256, m.fin, m.nfin);
                    return (int)SRN_OK;
                });
            }
            rc = run_tasks(h, tasks);
        }
        if (rc) return rc;
        if (lane == 0 && training && h->stagger_depth > 0 && h->groups[gi].depth == h->stagger_depth) cudaEventRecord(h->ev_inputs, h->L->main);
        gi = ge;
    }
    return SRN_OK;
}

static int bwd_group(srn_engine* h, Group& g, int nb, const Inputs& I) {
    Cached* c;
    h->cur_group = (int)(&g - h->groups.data());
    int rc = plan_for(h, g, nb, &c);
    if (rc) return rc;
    const int np = (int)g.h.size();
    switch (g.kind) {
        case SRN_KIND_DENSE:
        case SRN_KIND_CONV:
            if (g.tc == 3) {
                LAUNCH(SRN_CAT_ACT_BWD, k_act_bwd_bias<true>, c->total[P_COL], COL_THREADS, g.d, c->prefix[P_COL], np, nb, I);
                break;
            }
            LAUNCH(SRN_CAT_ACT_BWD, k_act_bwd_bias<false>, c->total[P_COL], COL_THREADS, g.d, c->prefix[P_COL], np, nb, I);
            {
                // the weight gradient goes to a floating stream behind this bucket's activation backward
                cudaStream_t lane = h->cur;
                if (h->multi_stream) {
                    const int w = h->L->wg_next; h->L->wg_next = (w + 1) % srn_engine::NWG; h->L->wg_used[w] = true;
                    cudaEventRecord(h->L->ev_wg, lane);
                    cudaStreamWaitEvent(h->L->wg_stream[w], h->L->ev_wg, 0);
                    h->cur = h->L->wg_stream[w];
                }
                if (g.tc) {
                    const int BNf = TC_BN[g.cfgF];
                    if (g.tc == 1) LAUNCH_TC(SRN_CAT_GEMM_WGRAD, TC_WGRAD, false, P_WGRAD, BNf);
                    else LAUNCH_TC(SRN_CAT_GEMM_WGRAD, TC_WGRAD, true, P_WGRAD, BNf);
                } else {
                    LAUNCH(SRN_CAT_GEMM_WGRAD, k_conv_wgrad, c->total[P_WGRAD], SIMT_THREADS, g.d, c->prefix[P_WGRAD], np, nb, I);
                }
                h->cur = lane;
            }
            if (g.tc) {
                LAUNCH_TC(SRN_CAT_GEMM_DGRAD, TC_DGRAD, false, P_DGRAD, TC_BN[g.cfgC]);
                break;
            }
            if (g.cfgC == 0) launch_dgrad_cfg<0>(h, g, c, nb);
            else if (g.cfgC == 1) launch_dgrad_cfg<1>(h, g, c, nb);
            else launch_dgrad_cfg<2>(h, g, c, nb);
            break;
        case SRN_KIND_MAXPOOL:
            LAUNCH(SRN_CAT_POOL, k_maxpool_bwd, c->total[P_BWD_EW], EW_THREADS, g.d, c->prefix[P_BWD_EW], np, nb, I);
            break;
        case SRN_KIND_BN:
            LAUNCH(SRN_CAT_BN, k_bn_partial, c->total[P_COL], COL_THREADS, g.d, c->prefix[P_COL], np, nb, I, 1, g.done);
            LAUNCH(SRN_CAT_BN, k_bn_bwd_apply, c->total[P_BWD_EW], EW_THREADS, g.d, c->prefix[P_BWD_EW], np, nb, I);
            break;
        case SRN_KIND_CONCAT:
            LAUNCH(SRN_CAT_CONCAT, k_copy, c->total[P_BWD_EW], EW_THREADS, g.d, c->prefix[P_BWD_EW], np, nb, I, 1);
            break;
    }
    return SRN_OK;
}

static int backward_pass(srn_engine* h, int nb, const Inputs& I, int lane = -1) {
    const int mode = (I.x_istride == 0 && I.g_istride == 0) ? 1 : 2;
    int gi = (int)h->groups.size() - 1;
    while (gi >= 0) {
        int ge = gi;
        std::vector<Group*> first, later;      // consumers that overwrite an input gradient / that add to it afterwards
        while (ge >= 0 && h->groups[ge].depth == h->groups[gi].depth) {
            Group& g = h->groups[ge];
            if ((g.mode & mode) && (lane < 0 || g.lane == lane)) (g.ordinal == 0 ? first : later).push_back(&g);
            --ge;
        }
        int rc = run_level(h, first, [&](Group& g) { return bwd_group(h, g, nb, I); });
        if (rc) return rc;
        for (Group* g : later) { rc = bwd_group(h, *g, nb, I); if (rc) return rc; }    // accumulation order matters
        gi = ge;
    }
    for (int w = 0; w < srn_engine::NWG; ++w)      // every weight gradient is in before the optimiser (and the next batch) starts
        if (h->L->wg_used[w]) {
            cudaEventRecord(h->L->ev_wg_done[w], h->L->wg_stream[w]);
            cudaStreamWaitEvent(h->L->main, h->L->ev_wg_done[w], 0);
            h->L->wg_used[w] = false;
        }
    return SRN_OK;
}

static int ensure_metrics(srn_engine* h, long long slots) {
    if (slots > h->met_slots || !h->met_correct) {
        CK(buf_reserve(h->bMetC, slots * h->n * sizeof(int), 1.25));
        CK(buf_reserve(h->bMetCe, slots * h->n * sizeof(float), 1.25));
        CK(buf_reserve(h->bMetMse, slots * h->n * sizeof(float), 1.25));
        h->met_correct = h->bMetC.as<int>(); h->met_ce = h->bMetCe.as<float>(); h->met_mse = h->bMetMse.as<float>();
        h->met_slots = slots;
    }
    CK(cudaMemsetAsync(h->met_correct, 0, slots * h->n * sizeof(int), h->stream));
    CK(cudaMemsetAsync(h->met_ce, 0, slots * h->n * sizeof(float), h->stream));
    CK(cudaMemsetAsync(h->met_mse, 0, slots * h->n * sizeof(float), h->stream));
    return SRN_OK;
}

static Inputs shared_inputs(srn_engine* h) {
    Inputs I;
    I.x = h->xin; I.x_istride = 0; I.x_bs = h->xe;
    I.g = h->gin; I.g_istride = 0; I.g_bs = h->G;
    return I;
}

static void launch_loss(srn_engine* h, int nb, const Inputs& I, int slot_base, int chunk, int row0, int training, int lane = -1) {
    LossArgs a;
    a.labels = h->lab;
    a.bits = I.g; a.bits_istride = I.g_istride; a.bits_bs = I.g_bs;
    a.loss_balance = h->d_lb;
    a.met_correct = h->met_correct; a.met_ce = h->met_ce; a.met_mse = h->met_mse;
    a.n = h->n; a.ncls = h->ncls; a.G = h->G;
    a.slot_base = slot_base; a.chunk = chunk; a.row0 = row0; a.training = training;
    const Prob* probs = lane < 0 ? h->d_loss : h->lane[lane].d_loss;
    const int np = lane < 0 ? h->n : h->lane[lane].n_loss;
    const int ctas = np * ((nb + 7) / 8);
    LAUNCH(SRN_CAT_LOSS, k_loss, ctas, 256, probs, np, nb, a);
}

static AdamScalars adam_scalars(int t) {
    // Keras Adam._prepare_local + TF ApplyAdam with every scalar held in fp16
    const float lr = h16(1e-3f), b1 = h16(0.9f), b2 = h16(0.999f), eps = h16(1e-4f);
    const float step = h16((float)t);
    const float b1p = h16(powf(b1, step)), b2p = h16(powf(b2, step));
    AdamScalars s;
    s.alpha = h16(h16(lr * h16(sqrtf(h16(1.f - b2p)))) / h16(1.f - b1p));
    s.omb1 = h16(1.f - b1);
    s.omb2 = h16(1.f - b2);
    s.eps = eps;
    return s;
}

// one training step on staged rows: gather -> forward -> loss -> backward -> Adam
static int train_step(srn_engine* h, const int* d_idx, int start, int nb, int slot, bool apply_adam) {
    LAUNCH(SRN_CAT_GATHER, k_gather_batch, nb, 256, h->x[0], h->gb[0], h->y[0], d_idx, start, nb, h->xe, h->G, h->xin, h->gin, h->lab);
    Inputs I = shared_inputs(h);
    int rc = SRN_OK;
    h->L = &h->lane[0];
    shared_im2col(h, nb, I);
    wait_adam(h);
    const int lanes = h->multi_stream ? h->n_lanes : 1;
    h->stagger_depth = getenv("SRN_STAGGER") ? atoi(getenv("SRN_STAGGER")) : 2;
    if (lanes > 1) cudaEventRecord(h->ev_inputs, h->stream);
    for (int l = 0; l < h->n_lanes && !rc; ++l) {
        // with one stream (profiling) the lanes simply follow each other on lane 0's streams
        srn_engine::LaneCtx* L = &h->lane[lanes > 1 ? l : 0];
        h->L = L;
        h->cur = L->main;
        if (lanes > 1 && l > 0) cudaStreamWaitEvent(L->main, h->ev_inputs, 0);
        rc = forward_pass(h, nb, I, true, l);
        if (!rc) {
            launch_loss(h, nb, I, slot, 1 << 30, 0, 1, l);
            rc = backward_pass(h, nb, I, l);
        }
        if (lanes > 1 && l > 0) cudaEventRecord(L->ev_done, L->main);
    }
    h->L = &h->lane[0];
    h->cur = h->stream;
    if (lanes > 1)
        for (int l = 1; l < h->n_lanes; ++l) cudaStreamWaitEvent(h->stream, h->lane[l].ev_done, 0);
    if (rc) return rc;
    if (apply_adam) {
        h->adam_t += 1;
        const long long n8 = h->p_elems / 8;
        if (h->multi_stream) {
            cudaEventRecord(h->ev_adam_go, h->stream);
            cudaStreamWaitEvent(h->adam_stream, h->ev_adam_go, 0);
            h->cur = h->adam_stream;
        }
        LAUNCH(SRN_CAT_ADAM, k_adam, (unsigned)((n8 + 255) / 256), 256, h->P, h->M, h->V, h->Gr, n8, adam_scalars(h->adam_t));
        if (h->multi_stream) {
            cudaEventRecord(h->ev_adam_done, h->adam_stream);
            h->cur = h->stream;
            h->adam_pending = true;
        }
    }
    return SRN_OK;
}

// inference over dataset rows [start, start+rows) with metrics per `chunk` rows into slots slot_base..
static int eval_rows(srn_engine* h, int split, int start, int rows, int chunk, int slot_base) {
    const int MB = h->cfg.max_batch;
    for (int r = 0; r < rows; r += MB) {
        const int nb = std::min(MB, rows - r);
        LAUNCH(SRN_CAT_GATHER, k_gather_batch, nb, 256, h->x[split], h->gb[split], h->y[split], (const int*)nullptr, start + r, nb, h->xe, h->G,
               h->xin, h->gin, h->lab);
        Inputs I = shared_inputs(h);
        int rc = forward_pass(h, nb, I, false);
        if (rc) return rc;
        launch_loss(h, nb, I, slot_base, chunk, r, 0);
    }
    return SRN_OK;
}

static int check_ready(srn_engine* h, const char* what) {
    if (!h) return fail(SRN_ERR_ARG, "%s: null engine", what);
    if (!h->built) return fail(SRN_ERR_STATE, "%s: call srn_build first", what);
    if (!h->x[0]) return fail(SRN_ERR_STATE, "%s: call srn_set_dataset first", what);
    return SRN_OK;
}

extern "C" int srn_fit(srn_engine* h, int32_t epochs, int32_t batch, int32_t n_val, const int32_t* perms,
                       uint64_t shuffle_seed, int32_t max_steps, int32_t* train_correct, float* train_loss,
                       int32_t* val_correct, float* val_mse_sum, float* device_ms) {
    int rc = check_ready(h, "srn_fit");
    if (rc) return rc;
    CK(cudaSetDevice(h->cfg.device));
    const int n_fit = h->nrows[0] - n_val;
    if (epochs <= 0 || batch <= 0 || batch > h->cfg.max_batch || n_val < 0 || n_fit <= 0)
        return fail(SRN_ERR_ARG, "srn_fit: epochs %d batch %d (max %d) n_val %d of %d rows", epochs, batch, h->cfg.max_batch, n_val, h->nrows[0]);
    const int S = (n_fit + batch - 1) / batch, Vc = n_val ? (n_val + batch - 1) / batch : 0;
    std::vector<int32_t> gen;
    if (!perms) {
        gen.resize((size_t)epochs * n_fit);
        std::mt19937_64 rng(shuffle_seed);
        for (int e = 0; e < epochs; ++e) {
            int32_t* p = gen.data() + (size_t)e * n_fit;
            for (int i = 0; i < n_fit; ++i) p[i] = i;
            for (int i = n_fit - 1; i > 0; --i) std::swap(p[i], p[rng() % (uint64_t)(i + 1)]);
        }
        perms = gen.data();
    } else {
        for (size_t i = 0; i < (size_t)epochs * n_fit; ++i)
            if (perms[i] < 0 || perms[i] >= n_fit) return fail(SRN_ERR_ARG, "srn_fit: permutation entry %d out of range", perms[i]);
    }
    CK(buf_reserve(h->bPerm, (size_t)epochs * n_fit * sizeof(int)));
    int* d_perm = h->bPerm.as<int>();
    CK(cudaMemcpy(d_perm, perms, (size_t)epochs * n_fit * sizeof(int), cudaMemcpyHostToDevice));
    const long long slots = (long long)epochs * (S + Vc);
    rc = ensure_metrics(h, slots);
    if (rc) return rc;
    cudaEvent_t ev0 = h->ev_t0, ev1 = h->ev_t1;
    float ms_total = 0.f;
    double host_enqueue_s = 0.0;          // host time spent issuing the steps' launches (SRN_HOST_TIMING prints it)
    int steps_done = 0;
    bool stop = false;
    // gradient accumulators start at zero (a debug step without the optimiser leaves its sums behind)
    wait_adam(h);
    CK(cudaMemsetAsync(h->Gr, 0, (size_t)h->p_elems * sizeof(float), h->stream));
    // warm the launch plans outside the timed region
    const auto tp0 = std::chrono::steady_clock::now();
    for (auto& g : h->groups) { Cached* c; rc = plan_for(h, g, batch, &c); if (rc) return rc; }
    const double plan_s = std::chrono::duration<double>(std::chrono::steady_clock::now() - tp0).count();
    for (int e = 0; e < epochs && !stop; ++e) {
        CK(cudaEventRecord(ev0, h->stream));
        for (int s = 0; s < S; ++s) {
            const int nb = std::min(batch, n_fit - s * batch);
            const auto th0 = std::chrono::steady_clock::now();
            rc = train_step(h, d_perm + (size_t)e * n_fit + (size_t)s * batch, 0, nb, e * S + s, true);
            host_enqueue_s += std::chrono::duration<double>(std::chrono::steady_clock::now() - th0).count();
            if (rc) return rc;
            if (max_steps > 0 && ++steps_done >= max_steps) { stop = true; break; }
        }
        wait_adam(h);
        CK(cudaEventRecord(ev1, h->stream));
        if (n_val && !stop) {
            rc = eval_rows(h, 0, n_fit, n_val, batch, epochs * S + e * Vc);
            if (rc) return rc;
        }
        CK(cudaEventSynchronize(ev1));
        float ms = 0.f;
        CK(cudaEventElapsedTime(&ms, ev0, ev1));
        ms_total += ms;
    }
    CK(cudaStreamSynchronize(h->stream));
    CK(cudaGetLastError());
    const size_t tn = (size_t)epochs * S * h->n, vn = (size_t)epochs * Vc * h->n;
    if (train_correct) CK(cudaMemcpy(train_correct, h->met_correct, tn * sizeof(int), cudaMemcpyDeviceToHost));
    if (train_loss) {
        std::vector<float> ce(tn), mse(tn);
        CK(cudaMemcpy(ce.data(), h->met_ce, tn * sizeof(float), cudaMemcpyDeviceToHost));
        CK(cudaMemcpy(mse.data(), h->met_mse, tn * sizeof(float), cudaMemcpyDeviceToHost));
        for (int e = 0; e < epochs; ++e)
            for (int s = 0; s < S; ++s) {
                const float nb = (float)std::min(batch, n_fit - s * batch);
                for (int i = 0; i < h->n; ++i) {
                    const size_t k = ((size_t)e * S + s) * h->n + i;
                    train_loss[2 * k] = ce[k] / nb;
                    train_loss[2 * k + 1] = mse[k] / nb;
                }
            }
    }
    if (val_correct && vn) CK(cudaMemcpy(val_correct, h->met_correct + tn, vn * sizeof(int), cudaMemcpyDeviceToHost));
    if (val_mse_sum && vn) CK(cudaMemcpy(val_mse_sum, h->met_mse + tn, vn * sizeof(float), cudaMemcpyDeviceToHost));
    if (device_ms) *device_ms = ms_total;
    if (getenv("SRN_HOST_TIMING"))
        fprintf(stderr, "srn_fit: %d steps, device %.3f ms, host enqueue %.3f ms, launch plans %.3f ms\n", max_steps > 0 ? steps_done : epochs * S, ms_total, 1e3 * host_enqueue_s, 1e3 * plan_s);
    return SRN_OK;
}

extern "C" int srn_evaluate(srn_engine* h, int32_t split, int32_t start, int32_t rows, int32_t chunk,
                            int32_t* correct, float* mse_sum) {
    int rc = check_ready(h, "srn_evaluate");
    if (rc) return rc;
    if (split < 0 || split > 1 || start < 0 || rows <= 0 || chunk <= 0 || start + rows > h->nrows[split])
        return fail(SRN_ERR_ARG, "srn_evaluate: rows [%d, %d) of split %d (%d rows), chunk %d", start, start + rows, split, split >= 0 && split <= 1 ? h->nrows[split] : -1, chunk);
    CK(cudaSetDevice(h->cfg.device));
    const int slots = (rows + chunk - 1) / chunk;
    rc = ensure_metrics(h, slots);
    if (rc) return rc;
    rc = eval_rows(h, split, start, rows, chunk, 0);
    if (rc) return rc;
    CK(cudaStreamSynchronize(h->stream));
    CK(cudaGetLastError());
    if (correct) CK(cudaMemcpy(correct, h->met_correct, (size_t)slots * h->n * sizeof(int), cudaMemcpyDeviceToHost));
    if (mse_sum) CK(cudaMemcpy(mse_sum, h->met_mse, (size_t)slots * h->n * sizeof(float), cudaMemcpyDeviceToHost));
    return SRN_OK;
}

// head logits of every individual -> [n][rows_total][NL] (rows r0.. of this chunk), optionally sigmoid of the
// replication part -> [n][R][G]
__global__ void k_collect_heads(const Prob* __restrict__ loss_probs, int n, int nb, int NL, int ncls, int r0, int rows_total,
                                __half* __restrict__ logits, __half* __restrict__ probs) {
    const int i = blockIdx.y;
    const int b = blockIdx.x;
    const Prob p = loss_probs[i];
    const __half* z = p.out + (long long)b * p.out_bs;
    for (int j = threadIdx.x; j < NL; j += blockDim.x) {
        if (logits) logits[((long long)i * rows_total + r0 + b) * NL + j] = z[j];
        if (probs && j >= ncls) {
            float s = rh(1.f / (1.f + expf(-__half2float(z[j]))));
            probs[((long long)i * rows_total + r0 + b) * (NL - ncls) + (j - ncls)] = __float2half_rn(s);
        }
    }
    (void)n; (void)nb;
}

extern "C" int srn_forward_logits(srn_engine* h, int32_t split, int32_t start, int32_t rows, int32_t training,
                                  uint16_t* logits) {
    int rc = check_ready(h, "srn_forward_logits");
    if (rc) return rc;
    if (split < 0 || split > 1 || start < 0 || rows <= 0 || start + rows > h->nrows[split] || !logits)
        return fail(SRN_ERR_ARG, "srn_forward_logits: bad range");
    if (training && rows > h->cfg.max_batch) return fail(SRN_ERR_ARG, "srn_forward_logits: training-mode batch > max_batch");
    CK(cudaSetDevice(h->cfg.device));
    CK(buf_reserve(h->bLogits, (size_t)h->n * rows * h->NL * sizeof(__half)));
    __half* d_out = h->bLogits.as<__half>();
    const int MB = h->cfg.max_batch;
    for (int r = 0; r < rows; r += MB) {
        const int nb = std::min(MB, rows - r);
        LAUNCH(SRN_CAT_GATHER, k_gather_batch, nb, 256, h->x[split], h->gb[split], h->y[split], (const int*)nullptr, start + r, nb, h->xe, h->G,
               h->xin, h->gin, h->lab);
        Inputs I = shared_inputs(h);
        rc = forward_pass(h, nb, I, training != 0);
        if (rc) return rc;
        dim3 grid(nb, h->n);
        LAUNCH(SRN_CAT_OTHER, k_collect_heads, grid, 128, h->d_loss, h->n, nb, h->NL, h->ncls, r, rows, d_out, (__half*)nullptr);
    }
    CK(cudaStreamSynchronize(h->stream));
    CK(cudaGetLastError());
    CK(cudaMemcpy(logits, d_out, (size_t)h->n * rows * h->NL * sizeof(__half), cudaMemcpyDeviceToHost));
    return SRN_OK;
}

extern "C" int srn_replicate(srn_engine* h, const uint8_t* parent_bits, int32_t R, const int32_t* image_index,
                             uint16_t* probs, uint32_t* packed_bits, int32_t* mutations) {
    int rc = check_ready(h, "srn_replicate");
    if (rc) return rc;
    if (!parent_bits || R <= 0) return fail(SRN_ERR_ARG, "srn_replicate: bad argument");
    if (!h->x[1]) return fail(SRN_ERR_STATE, "srn_replicate: no test split");
    CK(cudaSetDevice(h->cfg.device));
    const int n = h->n, G = h->G, xe = h->xe, words = (G + 31) / 32;
    const long long rows = (long long)n * R;
    std::vector<int32_t> idx((size_t)rows);
    if (image_index) {
        for (long long r = 0; r < rows; ++r) {
            if (image_index[r] < 0 || image_index[r] >= h->nrows[1]) return fail(SRN_ERR_ARG, "srn_replicate: image index out of range");
            idx[r] = image_index[r];
        }
    } else {      // np.repeat(x_test, ceil(n*R / n_test), axis=0)[:n*R]  (experiment_worker.py:139-150)
        const long long repeat = (rows + h->nrows[1] - 1) / h->nrows[1];
        for (long long r = 0; r < rows; ++r) idx[r] = (int32_t)(r / repeat);
    }
    // staging buffers belong to the handle and only grow: no cudaMalloc / cudaFree per generation
    auto cleanup = [&]() {};
#define CKR(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { cleanup(); return fail(SRN_ERR_CUDA, "%s: %s", #call, cudaGetErrorString(e_)); } } while (0)
    CKR(buf_reserve(h->bIdx, rows * sizeof(int), 1.25));
    CKR(buf_reserve(h->bRepX, rows * xe * sizeof(__half), 1.25));
    CKR(buf_reserve(h->bRepG, (size_t)n * G * sizeof(__half), 1.25));
    CKR(buf_reserve(h->bRepPar, (size_t)n * G, 1.25));
    CKR(buf_reserve(h->bRepProbs, rows * G * sizeof(__half), 1.25));
    CKR(buf_reserve(h->bRepPacked, rows * words * sizeof(uint32_t), 1.25));
    CKR(buf_reserve(h->bRepMut, rows * sizeof(int), 1.25));
    int* d_idx = h->bIdx.as<int>(); __half* d_x = h->bRepX.as<__half>(); __half* d_g = h->bRepG.as<__half>();
    uint8_t* d_par = h->bRepPar.as<uint8_t>(); __half* d_probs = h->bRepProbs.as<__half>();
    uint32_t* d_packed = h->bRepPacked.as<uint32_t>(); int* d_mut = h->bRepMut.as<int>();
    CKR(cudaMemcpy(d_idx, idx.data(), rows * sizeof(int), cudaMemcpyHostToDevice));
    CKR(cudaMemcpy(d_par, parent_bits, (size_t)n * G, cudaMemcpyHostToDevice));
    LAUNCH(SRN_CAT_GATHER, k_gather_batch, (unsigned)rows, 256, h->x[1], (const uint8_t*)nullptr, (const uint8_t*)nullptr, d_idx, 0, (int)rows, xe, G,
           d_x, (__half*)nullptr, (uint8_t*)nullptr);
    LAUNCH(SRN_CAT_OTHER, k_u8_to_f16, (unsigned)(((long long)n * G + 255) / 256), 256, d_par, d_g, (long long)n * G);
    const int MB = h->cfg.max_batch;
    for (int r0 = 0; r0 < R; r0 += MB) {
        const int nb = std::min(MB, R - r0);
        Inputs I;
        I.x = d_x + (long long)r0 * xe; I.x_istride = (long long)R * xe; I.x_bs = xe;
        I.g = d_g; I.g_istride = G; I.g_bs = 0;     // the parent's genotype, broadcast over its rows
        rc = forward_pass(h, nb, I, false);
        if (rc) { cleanup(); return rc; }
        dim3 grid(nb, n);
        LAUNCH(SRN_CAT_READOUT, k_collect_heads, grid, 128, h->d_loss, n, nb, h->NL, h->ncls, r0, R, (__half*)nullptr, d_probs);
    }
    LAUNCH(SRN_CAT_READOUT, k_readout, (unsigned)((rows + 7) / 8), 256, d_probs, d_par, (int)rows, R, G, d_packed, d_mut);
    CKR(cudaStreamSynchronize(h->stream));
    CKR(cudaGetLastError());
    if (probs) CKR(cudaMemcpy(probs, d_probs, rows * G * sizeof(__half), cudaMemcpyDeviceToHost));
    if (packed_bits) CKR(cudaMemcpy(packed_bits, d_packed, rows * words * sizeof(uint32_t), cudaMemcpyDeviceToHost));
    if (mutations) CKR(cudaMemcpy(mutations, d_mut, rows * sizeof(int), cudaMemcpyDeviceToHost));
    cleanup();
#undef CKR
    return SRN_OK;
}

// ------------------------------------------------------------------ debug / parity hooks (declared in srn_host.h)
extern "C" int srn_debug_step(srn_engine* h, const int32_t* rows, int32_t nb, int32_t apply_adam) {
    int rc = check_ready(h, "srn_debug_step");
    if (rc) return rc;
    if (!rows || nb <= 0 || nb > h->cfg.max_batch) return fail(SRN_ERR_ARG, "srn_debug_step: bad batch");
    CK(cudaSetDevice(h->cfg.device));
    CK(buf_reserve(h->bIdx, nb * sizeof(int), 1.25));
    int* d_idx = h->bIdx.as<int>();
    CK(cudaMemcpy(d_idx, rows, nb * sizeof(int), cudaMemcpyHostToDevice));
    rc = ensure_metrics(h, 1);
    cudaMemsetAsync(h->Gr, 0, (size_t)h->p_elems * sizeof(float), h->stream);
    if (!rc) rc = train_step(h, d_idx, 0, nb, 0, apply_adam != 0);
    wait_adam(h);
    cudaError_t e = cudaStreamSynchronize(h->stream);
    if (rc) return rc;
    if (e != cudaSuccess) return fail(SRN_ERR_CUDA, "srn_debug_step: %s", cudaGetErrorString(e));
    CK(cudaGetLastError());
    return SRN_OK;
}

extern "C" int srn_debug_get_tensor(srn_engine* h, int32_t ind, int32_t tensor, int32_t grad, int32_t nb, uint16_t* buf) {
    if (!h || !h->built || ind < 0 || ind >= h->n) return fail(SRN_ERR_ARG, "srn_debug_get_tensor: bad individual");
    Indiv& I = h->inds[ind];
    if (tensor < 2 || tensor >= (int)I.tensors.size() || nb <= 0 || nb > h->cfg.max_batch) return fail(SRN_ERR_ARG, "srn_debug_get_tensor: bad tensor/batch");
    const Tensor& T = I.tensors[tensor];
    const __half* src = grad ? T.grad : T.act;
    if (!src) return fail(SRN_ERR_ARG, "srn_debug_get_tensor: tensor %d has no %s buffer", tensor, grad ? "gradient" : "activation");
    CK(cudaSetDevice(h->cfg.device));
    CK(cudaStreamSynchronize(h->stream));
    if (T.phys_C) {             // padded concatenation (or a view of one): gather the logical columns
        std::vector<uint16_t> tmp((size_t)nb * T.phys_C);
        CK(cudaMemcpy2D(tmp.data(), (size_t)T.phys_C * 2, src, (size_t)T.bs * 2, (size_t)T.phys_C * 2, (size_t)nb, cudaMemcpyDeviceToHost));
        for (int b = 0; b < nb; ++b)
            for (long long e = 0; e < T.elems; ++e) buf[(size_t)b * T.elems + e] = tmp[(size_t)b * T.phys_C + T.colmap[(size_t)e]];
    } else if (T.bs == T.rows * T.Cs) {
        CK(cudaMemcpy2D(buf, (size_t)T.C * sizeof(__half), src, (size_t)T.Cs * sizeof(__half), (size_t)T.C * sizeof(__half), (size_t)nb * T.rows, cudaMemcpyDeviceToHost));
    } else {                    // produced in place inside a concatenation buffer: rows of Cs, the buffer's batch stride
        for (int b = 0; b < nb; ++b)
            CK(cudaMemcpy2D(buf + (size_t)b * T.elems, (size_t)T.C * sizeof(__half), src + (size_t)b * T.bs, (size_t)T.Cs * sizeof(__half),
                            (size_t)T.C * sizeof(__half), (size_t)T.rows, cudaMemcpyDeviceToHost));
    }
    return SRN_OK;
}

extern "C" int srn_debug_get_grad(srn_engine* h, int32_t ind, int32_t layer, int32_t which, float* buf, int64_t count) {
    __half* p; int64_t c; int rows, cols, ld; const std::vector<int>* rm;
    int rc = locate_weights(h, ind, layer, which, &p, &c, &rows, &cols, &ld, &rm);
    if (rc) return rc;
    if (which > 1 || count != c) return fail(SRN_ERR_ARG, "srn_debug_get_grad: bad which/count");
    CK(cudaSetDevice(h->cfg.device));
    CK(cudaStreamSynchronize(h->stream));
    return copy_rows(buf, reinterpret_cast<char*>(h->Gr + (p - h->P)), rows, cols, ld, sizeof(float), rm, true);
}

// SRN_GUARD=1: number of the first handle buffer whose guard band was overwritten (1-based), 0 = all intact, -1 = guards off
extern "C" int srn_debug_check_guards(srn_engine* h) {
    if (!h) return -1;
    if (!guards_on()) return -1;
    if (cudaSetDevice(h->cfg.device) != cudaSuccess || cudaDeviceSynchronize() != cudaSuccess) return -2;
    const DevBuf* bufs[] = {&h->bP, &h->bM, &h->bV, &h->bGr, &h->bS, &h->bSave, &h->bAct, &h->bIm2col, &h->bMetC, &h->bMetCe, &h->bMetMse,
                            &h->bPerm, &h->bLogits, &h->bIdx, &h->bRepX, &h->bRepG, &h->bRepPar, &h->bRepProbs, &h->bRepPacked, &h->bRepMut};
    int k = 0;
    for (const DevBuf* b : bufs) { ++k; if (!buf_guard_ok(*b)) return k; }
    return 0;
}

extern "C" int srn_debug_launch_groups(srn_engine* h) { return h && h->built ? (int)h->groups.size() : -1; }

// ------------------------------------------------------------------ per-category device timing (bench.py roofline)
extern "C" int srn_profile(srn_engine* h, int32_t on) {
    if (!h) return fail(SRN_ERR_ARG, "srn_profile: null engine");
    CK(cudaSetDevice(h->cfg.device));
    CK(cudaStreamSynchronize(h->stream));
    for (ProfRec* r : h->prof) { cudaEventDestroy(r->a); cudaEventDestroy(r->b); delete r; }
    h->prof.clear();
    h->profiling = on != 0;
    // per-launch event times only mean something when launches do not overlap: profile on one stream
    h->multi_stream = (!h->profiling || on == 2) && getenv("SRN_SINGLE_STREAM") == nullptr;      // on == 2: timeline of the overlapped schedule
    return SRN_OK;
}

extern "C" int srn_profile_read(srn_engine* h, float* ms, int32_t* launches) {
    if (!h || !ms || !launches) return fail(SRN_ERR_ARG, "srn_profile_read: null argument");
    CK(cudaSetDevice(h->cfg.device));
    CK(cudaStreamSynchronize(h->stream));
    for (int c = 0; c < SRN_CAT_COUNT; ++c) { ms[c] = 0.f; launches[c] = 0; }
    for (ProfRec* r : h->prof) {
        float t = 0.f;
        CK(cudaEventElapsedTime(&t, r->a, r->b));
        ms[r->cat] += t;
        launches[r->cat] += 1;
    }
    return SRN_OK;
}

// Per-launch detail of the last profiled region (debug hook): ms[i] = event time of launch i, info[8 i ..] =
// {category, bucket, kind, tc, BN forward, BN dgrad, problems, depth}, start[i] = offset of its start from the first
// launch (meaningful across streams: srn_profile(h, 2) keeps the side streams on), work[3 i ..] = {FLOPs of one GEMM pass over the
// bucket's real (unpadded) shapes, activation bytes read, activation bytes written} at batch nb.
extern "C" int srn_debug_profile_detail(srn_engine* h, int32_t nb, int32_t cap, float* ms, float* start, int32_t* info, double* work) {
    if (!h || !ms || !start || !info || !work) return -1;
    if (cudaSetDevice(h->cfg.device) != cudaSuccess || cudaStreamSynchronize(h->stream) != cudaSuccess) return -1;
    int n = 0;
    for (ProfRec* r : h->prof) {
        if (n >= cap) break;
        float t = 0.f;
        cudaEventElapsedTime(&t, r->a, r->b);
        ms[n] = t;
        start[n] = 0.f;
        cudaEventElapsedTime(&start[n], h->prof.front()->a, r->a);      // offset from the first launch of the region
        int32_t* o = info + 8 * (size_t)n;
        double* w = work + 3 * (size_t)n;
        o[0] = r->cat; o[1] = r->group;
        for (int k = 2; k < 8; ++k) o[k] = -1;
        w[0] = w[1] = w[2] = 0.0;
        if (r->group >= 0 && r->group < (int)h->groups.size()) {
            const Group& g = h->groups[r->group];
            o[2] = g.kind; o[3] = g.tc; o[4] = g.cfgF; o[5] = g.cfgC; o[6] = (int)g.h.size(); o[7] = g.depth;
            for (const Prob& p : g.h) {
                const double in_rows = (double)nb * p.H * p.W, out_rows = (double)nb * p.Ho * p.Wo;
                if (g.kind == SRN_KIND_DENSE || g.kind == SRN_KIND_CONV)
                    w[0] += 2.0 * out_rows * p.kh * p.kw * p.C * p.F;
                w[1] += in_rows * p.Cs * 2.0;
                w[2] += out_rows * (g.kind == SRN_KIND_DENSE || g.kind == SRN_KIND_CONV ? p.Fs : p.Cs) * 2.0;
            }
        }
        ++n;
    }
    return n;
}

// Shapes of the problems of one launch bucket (debug hook): 20 ints per problem
// {H, W, C, Cs, Ho, Wo, F, Fs, kh, kw, sh, sw, act, din_acc, individual, has din, coff, n_el, Cj, in_sel}.
extern "C" int srn_debug_group_probs(srn_engine* h, int32_t group, int32_t cap, int32_t* out) {
    if (!h || !out || group < 0 || group >= (int)h->groups.size()) return -1;
    int n = 0;
    for (const Prob& p : h->groups[group].h) {
        if (n >= cap) break;
        int32_t* o = out + 20 * (size_t)n++;
        o[0] = p.H; o[1] = p.W; o[2] = p.C; o[3] = p.Cs; o[4] = p.Ho; o[5] = p.Wo; o[6] = p.F; o[7] = p.Fs;
        o[8] = p.kh; o[9] = p.kw; o[10] = p.sh; o[11] = p.sw; o[12] = p.act; o[13] = p.din_acc; o[14] = p.indiv; o[15] = p.din != nullptr;
        o[16] = p.coff; o[17] = p.n_el; o[18] = p.Cj; o[19] = p.in_sel;
    }
    return n;
}
