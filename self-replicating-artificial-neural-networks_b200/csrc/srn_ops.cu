// Stand-alone entry points: edit distance, genotype readout, BatchNormalizationF16 forward (the autoencoder decode is srn_decode.cu).
#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstddef>
#include <vector>

#include "../../include/serann_b200.h"
#include "srn_kernels.cuh"
#include "srn_host.h"

using namespace srn;

#define CK(call)                                                                                   \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess) return fail(SRN_ERR_CUDA, "%s: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
    } while (0)

static int check_device(int device) {
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0) return fail(SRN_ERR_CUDA, "no CUDA device (%s); this library has no CPU path", cudaGetErrorString(e));
    if (device < 0 || device >= count) return fail(SRN_ERR_ARG, "device %d of %d", device, count);
    CK(cudaSetDevice(device));
    return SRN_OK;
}

// ------------------------------------------------------------------ edit distance (host)
// Levenshtein distance between two byte strings, two-row dynamic programme.  Replaces edlib.align(a, b)['editDistance']
// (reference evolutionary_experiment/logic/experiment.py:16-17), which is not installable here.
extern "C" int32_t srn_levenshtein(const char* a, int32_t la, const char* b, int32_t lb) {
    if (la < 0 || lb < 0 || (!a && la) || (!b && lb)) return -1;
    if (la < lb) { std::swap(a, b); std::swap(la, lb); }
    if (lb == 0) return la;
    std::vector<int32_t> prev(lb + 1), cur(lb + 1);
    for (int j = 0; j <= lb; ++j) prev[j] = j;
    for (int i = 1; i <= la; ++i) {
        cur[0] = i;
        const char ca = a[i - 1];
        for (int j = 1; j <= lb; ++j) {
            const int32_t sub = prev[j - 1] + (ca != b[j - 1]);
            cur[j] = std::min(sub, std::min(prev[j], cur[j - 1]) + 1);
        }
        prev.swap(cur);
    }
    return prev[lb];
}

// ------------------------------------------------------------------ stand-alone readout
extern "C" int srn_readout(int32_t device, const uint16_t* probs, const uint8_t* parent_bits, int32_t rows,
                           int32_t rows_per_parent, int32_t G, uint32_t* packed_bits, int32_t* mutations) {
    if (!probs || !parent_bits || rows < 0 || rows_per_parent <= 0 || G <= 0) return fail(SRN_ERR_ARG, "srn_readout: bad argument");
    if (rows == 0) return SRN_OK;
    int rc = check_device(device);
    if (rc) return rc;
    const int words = (G + 31) / 32, parents = (rows + rows_per_parent - 1) / rows_per_parent;
    __half* d_p = nullptr; uint8_t* d_par = nullptr; uint32_t* d_w = nullptr; int* d_m = nullptr;
    auto cleanup = [&]() { cudaFree(d_p); cudaFree(d_par); cudaFree(d_w); cudaFree(d_m); };
#define CKR(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { cleanup(); return fail(SRN_ERR_CUDA, "%s: %s", #call, cudaGetErrorString(e_)); } } while (0)
    CKR(cudaMalloc(&d_p, (size_t)rows * G * sizeof(__half)));
    CKR(cudaMalloc(&d_par, (size_t)parents * G));
    CKR(cudaMalloc(&d_w, (size_t)rows * words * sizeof(uint32_t)));
    CKR(cudaMalloc(&d_m, (size_t)rows * sizeof(int)));
    CKR(cudaMemcpy(d_p, probs, (size_t)rows * G * sizeof(__half), cudaMemcpyHostToDevice));
    CKR(cudaMemcpy(d_par, parent_bits, (size_t)parents * G, cudaMemcpyHostToDevice));
    k_readout<<<(rows + 7) / 8, 256>>>(d_p, d_par, rows, rows_per_parent, G, d_w, d_m);
    g_launches.fetch_add(1);
    CKR(cudaDeviceSynchronize());
    CKR(cudaGetLastError());
    if (packed_bits) CKR(cudaMemcpy(packed_bits, d_w, (size_t)rows * words * sizeof(uint32_t), cudaMemcpyDeviceToHost));
    if (mutations) CKR(cudaMemcpy(mutations, d_m, (size_t)rows * sizeof(int), cudaMemcpyDeviceToHost));
    cleanup();
#undef CKR
    return SRN_OK;
}

// ------------------------------------------------------------------ stand-alone BatchNormalizationF16 forward
extern "C" int srn_bn_train_forward(int32_t device, const uint16_t* x, int64_t rows, int32_t C,
                                    const uint16_t* gamma, const uint16_t* beta, uint16_t* moving_mean,
                                    uint16_t* moving_variance, uint16_t* y) {
    if (!x || !gamma || !beta || !moving_mean || !moving_variance || !y || rows <= 0 || C <= 0)
        return fail(SRN_ERR_ARG, "srn_bn_train_forward: bad argument");
    if (rows * (int64_t)C > 0x7fffffffLL) return fail(SRN_ERR_LIMIT, "srn_bn_train_forward: more than 2^31 elements");
    int rc = check_device(device);
    if (rc) return rc;
    const int chunks = col_chunks(rows, C);
    __half *d_x = nullptr, *d_y = nullptr, *d_par = nullptr; float* d_save = nullptr; Prob* d_prob = nullptr; int* d_prefix = nullptr;
    auto cleanup = [&]() { cudaFree(d_x); cudaFree(d_y); cudaFree(d_par); cudaFree(d_save); cudaFree(d_prob); cudaFree(d_prefix); };
#define CKR(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { cleanup(); return fail(SRN_ERR_CUDA, "%s: %s", #call, cudaGetErrorString(e_)); } } while (0)
    const size_t ne = (size_t)rows * C;
    CKR(cudaMalloc(&d_x, ne * sizeof(__half)));
    CKR(cudaMalloc(&d_y, ne * sizeof(__half)));
    CKR(cudaMalloc(&d_par, 4 * (size_t)C * sizeof(__half)));          // gamma, beta, moving mean, moving variance
    CKR(cudaMalloc(&d_save, (4 * (size_t)C + 2 * (size_t)C * chunks) * sizeof(float)));
    CKR(cudaMalloc(&d_prob, sizeof(Prob)));
    CKR(cudaMalloc(&d_prefix, 4 * sizeof(int)));
    CKR(cudaMemcpy(d_x, x, ne * sizeof(__half), cudaMemcpyHostToDevice));
    CKR(cudaMemcpy(d_par, gamma, C * sizeof(__half), cudaMemcpyHostToDevice));
    CKR(cudaMemcpy(d_par + C, beta, C * sizeof(__half), cudaMemcpyHostToDevice));
    CKR(cudaMemcpy(d_par + 2 * C, moving_mean, C * sizeof(__half), cudaMemcpyHostToDevice));
    CKR(cudaMemcpy(d_par + 3 * C, moving_variance, C * sizeof(__half), cudaMemcpyHostToDevice));
    Prob p{};
    p.in = d_x; p.out = d_y; p.w = d_par; p.bias = d_par + C; p.mov = d_par + 2 * C; p.save = d_save;
    p.H = 1; p.W = 1; p.C = C; p.Cs = C; p.Ho = 1; p.Wo = 1; p.F = C; p.Fs = C; p.in_bs = C; p.out_bs = C;    // nb = rows samples of one row each
    const int nb = (int)rows;
    const int pre[4] = {0, chunks, 0, ew_blocks(rows, C)};
    CKR(cudaMemcpy(d_prob, &p, sizeof(Prob), cudaMemcpyHostToDevice));
    CKR(cudaMemcpy(d_prefix, pre, sizeof pre, cudaMemcpyHostToDevice));
    Inputs I{};
    k_bn_partial<<<chunks, COL_THREADS>>>(d_prob, d_prefix, 1, nb, I, 0);
    k_bn_finalize_fwd<<<1, 128>>>(d_prob, nb, I);
    k_bn_apply<<<pre[3], EW_THREADS>>>(d_prob, d_prefix + 2, 1, nb, I);
    g_launches.fetch_add(3);
    CKR(cudaDeviceSynchronize());
    CKR(cudaGetLastError());
    CKR(cudaMemcpy(y, d_y, ne * sizeof(__half), cudaMemcpyDeviceToHost));
    CKR(cudaMemcpy(moving_mean, d_par + 2 * C, C * sizeof(__half), cudaMemcpyDeviceToHost));
    CKR(cudaMemcpy(moving_variance, d_par + 3 * C, C * sizeof(__half), cudaMemcpyDeviceToHost));
    cleanup();
#undef CKR
    return SRN_OK;
}

// ------------------------------------------------------------------ BatchNormalizationF16 forward without allocation
// Same three kernels as srn_bn_train_forward, but the per-channel statistics, the per-chunk partial sums and
// the problem descriptor live in the file-scope array below instead of cudaMalloc'ed scratch.  The array is part
// of the library image: every GPU that loads the library holds SRN_NOALLOC_SCRATCH_BYTES of it until exit.
constexpr size_t NOALLOC_HEADER_BYTES = 512;                       // Prob descriptor + launch prefix
constexpr size_t NOALLOC_FLOATS = (SRN_NOALLOC_SCRATCH_BYTES - NOALLOC_HEADER_BYTES) / sizeof(float);
static_assert(sizeof(Prob) + 8 * sizeof(int) <= NOALLOC_HEADER_BYTES, "descriptor does not fit the scratch header");

static __device__ __align__(256) unsigned char g_noalloc_scratch[SRN_NOALLOC_SCRATCH_BYTES];

constexpr int NOALLOC_MAX_DEVICES = 64;
// device address of the array, per device; written once per device (atomically: two host threads may make the first
// call at the same time and would store the same value).  cudaDeviceReset invalidates it - a process that resets a
// device must not call this entry point on that device afterwards.
static std::atomic<unsigned char*> g_noalloc_ptr[NOALLOC_MAX_DEVICES] = {};

extern "C" int srn_bn_train_forward_noalloc(int32_t device, void* stream, const uint16_t* d_x, int64_t rows, int32_t C,
                                            const uint16_t* d_gamma, const uint16_t* d_beta,
                                            uint16_t* d_moving_mean, uint16_t* d_moving_variance, uint16_t* d_y) {
    if (!d_x || !d_gamma || !d_beta || !d_moving_mean || !d_moving_variance || !d_y || rows <= 0 || C <= 0)
        return fail(SRN_ERR_ARG, "srn_bn_train_forward_noalloc: bad argument");
    if (rows * (int64_t)C > 0x7fffffffLL) return fail(SRN_ERR_LIMIT, "srn_bn_train_forward_noalloc: more than 2^31 elements");
    if (device < 0 || device >= NOALLOC_MAX_DEVICES) return fail(SRN_ERR_ARG, "srn_bn_train_forward_noalloc: device %d", device);
    if (d_moving_variance != d_moving_mean + C)
        return fail(SRN_ERR_ARG, "srn_bn_train_forward_noalloc: moving_variance must directly follow moving_mean (one [2][C] array)");
    // With C a multiple of 8 the kernels move 16-byte groups of 8 channels; a buffer that starts at an odd element
    // offset (a slice of a framework tensor) would fault there and poison the caller's CUDA context, so it is refused.
    if (C % 8 == 0 && ((reinterpret_cast<uintptr_t>(d_x) | reinterpret_cast<uintptr_t>(d_y)) & 15u) != 0)
        return fail(SRN_ERR_ARG, "srn_bn_train_forward_noalloc: with C a multiple of 8, d_x and d_y must be 16-byte aligned");
    if (((reinterpret_cast<uintptr_t>(d_x) | reinterpret_cast<uintptr_t>(d_y) | reinterpret_cast<uintptr_t>(d_gamma) |
          reinterpret_cast<uintptr_t>(d_beta) | reinterpret_cast<uintptr_t>(d_moving_mean)) & 1u) != 0)
        return fail(SRN_ERR_ARG, "srn_bn_train_forward_noalloc: fp16 buffers must be 2-byte aligned");
    const int chunks = col_chunks(rows, C);
    const size_t need = 4 * (size_t)C + 2 * (size_t)C * chunks;
    if (need > NOALLOC_FLOATS)
        return fail(SRN_ERR_LIMIT, "srn_bn_train_forward_noalloc: C * (4 + 2 * ceil(rows / %d)) = %zu floats exceeds the built-in "
                    "scratch of %zu floats (%zu bytes); use srn_bn_train_forward for this size", col_geom(C).rows_per_cta, need,
                    NOALLOC_FLOATS, (size_t)SRN_NOALLOC_SCRATCH_BYTES);
    // the caller's current device is put back before returning, whatever happens below
    struct DeviceGuard {
        int prev = -1;
        DeviceGuard() { if (cudaGetDevice(&prev) != cudaSuccess) prev = -1; }
        ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
    } guard;
    int rc = check_device(device);
    if (rc) return rc;
    if (!g_noalloc_ptr[device].load(std::memory_order_acquire)) {
        // In host code the array's name is only a handle for the runtime; the address kernels can use comes from
        // cudaGetSymbolAddress, asked once per device while that device is current.
        void* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess || !p) return fail(SRN_ERR_CUDA, "cudaGetSymbolAddress(g_noalloc_scratch) on device %d: %s", device, cudaGetErrorString(e));
        g_noalloc_ptr[device].store(static_cast<unsigned char*>(p), std::memory_order_release);
    }
    unsigned char* base = g_noalloc_ptr[device].load(std::memory_order_acquire);
    Prob* d_prob = reinterpret_cast<Prob*>(base);
    int* d_prefix = reinterpret_cast<int*>(base + sizeof(Prob));
    float* d_save = reinterpret_cast<float*>(base + NOALLOC_HEADER_BYTES);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    struct { Prob p; int pre[4]; } hdr{};
    Prob& p = hdr.p;
    p.in = reinterpret_cast<const __half*>(d_x); p.out = reinterpret_cast<__half*>(d_y);
    p.w = reinterpret_cast<const __half*>(d_gamma); p.bias = reinterpret_cast<const __half*>(d_beta);
    p.mov = reinterpret_cast<__half*>(d_moving_mean); p.save = d_save;
    p.H = 1; p.W = 1; p.C = C; p.Cs = C; p.Ho = 1; p.Wo = 1; p.F = C; p.Fs = C; p.in_bs = C; p.out_bs = C;
    const int nb = (int)rows;
    const int blocks = ew_blocks(rows, C);
    hdr.pre[0] = 0; hdr.pre[1] = chunks; hdr.pre[2] = 0; hdr.pre[3] = blocks;
    static_assert(offsetof(decltype(hdr), pre) == sizeof(Prob), "descriptor layout");
    CK(cudaMemcpyAsync(d_prob, &hdr, sizeof hdr, cudaMemcpyHostToDevice, st));
    // Nothing in the scratch has to start at zero: every statistic and every partial sum that is read below is
    // written first by this call (k_bn_partial covers all chunks x channels, k_bn_finalize_fwd all channels).
    Inputs I{};
    k_bn_partial<<<chunks, COL_THREADS, 0, st>>>(d_prob, d_prefix, 1, nb, I, 0);
    k_bn_finalize_fwd<<<1, 128, 0, st>>>(d_prob, nb, I);
    k_bn_apply<<<blocks, EW_THREADS, 0, st>>>(d_prob, d_prefix + 2, 1, nb, I);
    g_launches.fetch_add(3);
    CK(cudaGetLastError());
    CK(cudaStreamSynchronize(st));
    return SRN_OK;
}
