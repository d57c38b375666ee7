// Ribosomal-autoencoder decode on the 5th-generation tensor cores (tcgen05 + TMEM + TMA), sm_100a.
//
// generative_net in inference mode (reference ribosomal_autoencoder/model.py:35-48, called at
// evolutionary_experiment/logic/ribosomal_autoencoder.py:122-124):
//   one_hot(bits) (n, G, 2) -> Conv1D(32, 5) -> BN -> Flatten -> Dense(T * V) -> Reshape(T, V) -> BN -> argmax over V
//
// The input is binary, so the first two layers are a table: the 32 outputs at position l depend only on the five bits
// l .. l + 4, i.e. on a window value w_l in [0, 32).  Folding that table into the Dense kernel once, when the decoder is
// created, gives   z[r][j] = b[j] + sum_l Tab[l * 32 + w_l(r)][j],   Tab[l * 32 + w][j] = sum_f LUT[w][f] W[l * 32 + f][j]
// (accumulated in double).  As a GEMM this is  OneHot (n x 32 L) x Tab (32 L x T V)  whose left operand is EXACTLY
// representable in fp16 (a 0/1 matrix with one 1 per position), so only Tab needs splitting: Tab * 2^s = hi + lo with
// fp16 hi, lo (22 mantissa bits, products exact, fp32 accumulation in TMEM) - two kind::f16 passes instead of the
// three an fp16 split of both operands would need, and the left operand is not even read from memory: the CTA writes
// the two ones per row and 64-wide k-block straight into the swizzled shared-memory tile.
//
// k_dec_gemm: persistent, warp-specialised.  Tile = 256 genotypes x (tokens_per_tile * V <= 256) logits, two 128-row
// accumulators side by side in TMEM (512 columns).  Warps 0-3 maintain the one-hot A tile of each stage (clear the
// previous ones, set the new ones, generic-proxy stores + fence.proxy.async), warp 4 issues the TMA loads of the hi
// and lo table tiles (2-D tiled maps over this GPU's memory, swizzle-128B), warp 5 issues 16 tcgen05.mma per k-block,
// warps 6-9 run the epilogue: per token the second BN as one FMA per logit (y = acc * a[j] + c[j]), first-maximum
// argmax with the runner-up tracked; logits never reach HBM.  A token whose margin is below DEC_MARGIN is appended to
// a list and recomputed from the fp32 table in double by k_dec_recheck, so the result is the argmax of the exact
// arithmetic wherever the fp16 split could matter at all.
#include <algorithm>
#include <cmath>
#include <cstddef>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "../../include/serann_b200.h"
#include "srn_gemm_tc.cuh"
#include "srn_host.h"
#include "srn_tma.h"

using namespace srn;

#define CK(call)                                                                                   \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess) return fail(SRN_ERR_CUDA, "%s: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
    } while (0)

namespace {

constexpr int DEC_F = 32, DEC_KW = 5, DEC_WIN = 32;      // conv filters, taps, window values (2^taps)
constexpr float KERAS_BN_EPS = 1e-3f;
constexpr int DEC_M = 256;                                // genotypes per tile (two 128-row accumulators)
constexpr int DEC_BK = 64;                                // k per stage = two positions
constexpr int DEC_STAGES = 2;
constexpr int DEC_A_BYTES = DEC_M * DEC_BK * 2;           // 32 KB
constexpr int DEC_B_BYTES = 4 * 8192;                     // up to 256 columns of one table half, 64 k rows
constexpr int DEC_STAGE_BYTES = DEC_A_BYTES + 2 * DEC_B_BYTES;
constexpr int DEC_MAX_L = 100;                            // positions whose window bytes fit the shared-memory budget
constexpr int DEC_WIN_BYTES = DEC_M * DEC_MAX_L;
constexpr int DEC_SMEM = DEC_STAGES * DEC_STAGE_BYTES + DEC_WIN_BYTES + 2 * 256 * 4 + 256;
constexpr int DEC_THREADS = 320;                          // 4 writer warps, TMA warp, MMA warp, 4 epilogue warps
constexpr float DEC_MARGIN = 5e-4f;                       // relative top-1 / top-2 margin below which a token is rechecked

struct DecGeom {
    int n, L, N, T, V, tpt, BN, nkb, m_tiles, n_tiles;
};

__device__ __forceinline__ void tmem_ld8(uint32_t taddr, float (&v)[8]) {
    uint32_t r[8];
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];\n"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
                 : "r"(taddr) : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;\n" ::: "memory");
#pragma unroll
    for (int i = 0; i < 8; ++i) v[i] = __uint_as_float(r[i]);
}
__device__ __forceinline__ void named_bar(int id, int threads) { asm volatile("bar.sync %0, %1;\n" ::"r"(id), "r"(threads) : "memory"); }
__device__ __forceinline__ void st_shared_u16(uint32_t addr, uint16_t v) { asm volatile("st.shared.u16 [%0], %1;\n" ::"r"(addr), "h"(v) : "memory"); }

// window values w_l = sum_t bit[l + t] << t of every genotype: [n][L] bytes
__global__ void k_dec_windows(const uint8_t* __restrict__ bits, int n, int G, int L, uint8_t* __restrict__ win) {
    const long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= (long long)n * L) return;
    const int l = (int)(e % L);
    const long long r = e / L;
    int w = 0;
#pragma unroll
    for (int t = 0; t < DEC_KW; ++t) w |= (bits[r * G + l + t] ? 1 : 0) << t;
    win[e] = (uint8_t)w;
}

// Tab[l * 32 + w][j] = sum_f LUT[w][f] * W[(l * 32 + f) * N + j], accumulated in double
__global__ void k_dec_table(const double* __restrict__ lut, const float* __restrict__ W, int L, int N, float* __restrict__ tab, unsigned* __restrict__ maxbits) {
    const long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long total = (long long)L * DEC_WIN * N;
    float out = 0.f;
    if (e < total) {
        const int j = (int)(e % N);
        const int lw = (int)(e / N);
        const int l = lw / DEC_WIN, w = lw % DEC_WIN;
        double acc = 0.0;
#pragma unroll 8
        for (int f = 0; f < DEC_F; ++f) acc += lut[w * DEC_F + f] * (double)W[((long long)l * DEC_F + f) * N + j];
        out = (float)acc;
        tab[e] = out;
    }
    // largest magnitude, for the power-of-two scale of the fp16 split
    unsigned m = __float_as_uint(fabsf(out));
    for (int o = 16; o > 0; o >>= 1) m = max(m, __shfl_xor_sync(0xffffffffu, m, o));
    if ((threadIdx.x & 31) == 0 && m) atomicMax(maxbits, m);
}

// Tab * scale = hi + lo, both fp16 (hi = round(x), lo = round(x - hi): exact to 22 bits)
__global__ void k_dec_split(const float* __restrict__ tab, long long total, float scale, __half* __restrict__ hi, __half* __restrict__ lo) {
    const long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= total) return;
    const float x = tab[e] * scale;
    const __half h = __float2half_rn(x);
    hi[e] = h;
    lo[e] = __float2half_rn(x - __half2float(h));
}

__global__ void __launch_bounds__(DEC_THREADS, 1)
k_dec_gemm(const CUtensorMap* __restrict__ maps, const uint8_t* __restrict__ win, const float* __restrict__ acol,
           const float* __restrict__ ccol, DecGeom g, int* __restrict__ tokens, int* __restrict__ flagged, int flag_cap,
           int* __restrict__ flag_count) {
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const uint32_t sbase = smem_u32(smem_raw);
    if (sbase & 1023u) __trap();
    const uint32_t sWin = sbase + DEC_STAGES * DEC_STAGE_BYTES;
    const uint32_t sCoef = sWin + DEC_WIN_BYTES;                 // a[256] then c[256]
    const uint32_t sBar = sCoef + 2 * 256 * 4;                   // full[2], empty[2], tmem_full, tmem_empty, slot
    const uint32_t sFull = sBar, sEmpty = sBar + 16, sTFull = sBar + 32, sTEmpty = sBar + 40, sSlot = sBar + 48;
    unsigned char* win_s = smem_raw + (sWin - sbase);
    float* coef_s = reinterpret_cast<float*>(smem_raw + (sCoef - sbase));
    volatile uint32_t* slot_ptr = reinterpret_cast<volatile uint32_t*>(smem_raw + (sSlot - sbase));

    if (tid == 0) {
        for (int i = 0; i < DEC_STAGES; ++i) { mbar_init(sFull + i * 8, 128 + 1); mbar_init(sEmpty + i * 8, 1); }
        mbar_init(sTFull, 1); mbar_init(sTEmpty, 128);
        asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
    }
    if (warp == 5) tmem_alloc(sSlot, 512u);
    // the one-hot tiles start as zeros
    for (int i = tid; i < DEC_STAGES * DEC_A_BYTES / 16; i += DEC_THREADS) {
        const int st = i / (DEC_A_BYTES / 16), o = i % (DEC_A_BYTES / 16);
        *reinterpret_cast<uint4*>(smem_raw + st * DEC_STAGE_BYTES + o * 16) = make_uint4(0, 0, 0, 0);
    }
    fence_proxy_async();
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *slot_ptr;
    const int total_tiles = g.m_tiles * g.n_tiles;

    if (warp < 4) {
        // ================================================================ one-hot A tiles (128 threads, two rows each)
        uint32_t old_addr[DEC_STAGES][2][2];
#pragma unroll
        for (int s = 0; s < DEC_STAGES; ++s)
#pragma unroll
            for (int j = 0; j < 2; ++j) { old_addr[s][j][0] = 0xFFFFFFFFu; old_addr[s][j][1] = 0xFFFFFFFFu; }
        int st = 0; uint32_t ph = 1;
        for (int t = blockIdx.x; t < total_tiles; t += gridDim.x) {
            const int m0 = (t % g.m_tiles) * DEC_M;
            named_bar(1, 128);                                    // nobody still reads the previous tile's windows
            const int rows = min(DEC_M, g.n - m0);
            const long long bytes = (long long)rows * g.L;
            const uint8_t* src = win + (long long)m0 * g.L;
            if (((g.L & 15) == 0)) {
                for (long long i = tid; i < bytes / 16; i += 128)
                    reinterpret_cast<uint4*>(win_s)[i] = __ldg(reinterpret_cast<const uint4*>(src) + i);
            } else {
                for (long long i = tid; i < bytes; i += 128) win_s[i] = src[i];
            }
            named_bar(1, 128);
            for (int kb = 0; kb < g.nkb; ++kb) {
                mbar_wait(sEmpty + st * 8, ph);
                const uint32_t aS = sbase + st * DEC_STAGE_BYTES;
#pragma unroll
                for (int j = 0; j < 2; ++j) {
                    const int row = tid + j * 128;
#pragma unroll
                    for (int q = 0; q < 2; ++q) {
                        if (old_addr[st][j][q] != 0xFFFFFFFFu) st_shared_u16(old_addr[st][j][q], 0);
                        const int l = kb * 2 + q;
                        uint32_t addr = 0xFFFFFFFFu;
                        if (row < rows && l < g.L) {
                            const int k = q * DEC_WIN + win_s[row * g.L + l];
                            addr = aS + j * (DEC_A_BYTES / 2) + sw_k_off(tid, k >> 3) + (uint32_t)((k & 7) * 2);
                            st_shared_u16(addr, 0x3C00);          // fp16 1.0
                        }
                        old_addr[st][j][q] = addr;
                    }
                }
                fence_proxy_async();                              // generic-proxy stores -> visible to the tensor core's reads
                mbar_arrive(sFull + st * 8);
                if (++st == DEC_STAGES) { st = 0; ph ^= 1u; }
            }
        }
    } else if (warp == 4) {
        // ================================================================ TMA producer: hi and lo table tiles
        if (lane == 0) {
            const int nbox = (g.BN + 63) / 64;
            int st = 0; uint32_t ph = 1;
            for (int t = blockIdx.x; t < total_tiles; t += gridDim.x) {
                const int n0 = (t / g.m_tiles) * g.BN;
                for (int kb = 0; kb < g.nkb; ++kb) {
                    mbar_wait(sEmpty + st * 8, ph);
                    const uint32_t bar = sFull + st * 8;
                    const uint32_t bHi = sbase + st * DEC_STAGE_BYTES + DEC_A_BYTES, bLo = bHi + DEC_B_BYTES;
                    mbar_arrive_expect_tx(bar, (uint32_t)(2 * nbox * 8192));
                    for (int b = 0; b < nbox; ++b) {
                        tma_load_2d(bHi + b * 8192, maps, n0 + b * 64, kb * DEC_BK, bar);
                        tma_load_2d(bLo + b * 8192, reinterpret_cast<const char*>(maps) + 128, n0 + b * 64, kb * DEC_BK, bar);
                    }
                    if (++st == DEC_STAGES) { st = 0; ph ^= 1u; }
                }
            }
        }
    } else if (warp == 5) {
        // ================================================================ MMA issuer
        if (lane == 0) {
            const uint32_t idesc = umma_idesc(0, 1, g.BN);        // A K-major, B MN-major, N = BN
            int st = 0; uint32_t fph = 0;
            int it = 0;
            for (int t = blockIdx.x; t < total_tiles; t += gridDim.x, ++it) {
                if (it > 0) mbar_wait(sTEmpty, (uint32_t)((it - 1) & 1));      // the epilogue has drained both accumulators
                tc_fence_after();
                for (int kb = 0; kb < g.nkb; ++kb) {
                    mbar_wait(sFull + st * 8, fph);
                    tc_fence_after();
                    const uint32_t aS = sbase + st * DEC_STAGE_BYTES, bS = aS + DEC_A_BYTES;
#pragma unroll
                    for (int half = 0; half < 2; ++half) {
                        const uint64_t dA = umma_desc_sw128(aS + half * (DEC_A_BYTES / 2), 16, 1024);
                        const uint32_t tacc = tmem + (uint32_t)(half * 256);
#pragma unroll
                        for (int part = 0; part < 2; ++part) {
                            const uint64_t dB = umma_desc_sw128(bS + part * DEC_B_BYTES, 8192, 1024);
#pragma unroll
                            for (int ks = 0; ks < DEC_BK / 16; ++ks)
                                umma_f16(tacc, dA + (uint64_t)((32 >> 4) * ks), dB + (uint64_t)((2048 >> 4) * ks), idesc,
                                         (kb | part | ks) ? 1u : 0u);
                        }
                    }
                    umma_commit(sEmpty + st * 8);
                    if (++st == DEC_STAGES) { st = 0; fph ^= 1u; }
                }
                umma_commit(sTFull);
            }
        }
    } else {
        // ================================================================ epilogue (4 warps): BN2 affine + argmax per token
        const int ewarp = warp & 3, et = (warp - 6) * 32 + lane;             // TMEM lane quarter / thread index 0..127
        int it = 0;
        for (int t = blockIdx.x; t < total_tiles; t += gridDim.x, ++it) {
            const int m0 = (t % g.m_tiles) * DEC_M, nt = t / g.m_tiles, n0 = nt * g.BN, tok0 = nt * g.tpt;
            named_bar(2, 128);                                    // everybody is done with the previous tile's coefficients
            for (int c = et; c < g.BN; c += 128) {
                const bool ok = n0 + c < g.N;
                coef_s[c] = ok ? acol[n0 + c] : 0.f;
                coef_s[256 + c] = ok ? ccol[n0 + c] : 0.f;
            }
            named_bar(2, 128);
            mbar_wait(sTFull, (uint32_t)(it & 1));
            tc_fence_after();
            for (int half = 0; half < 2; ++half) {
                const int row = m0 + half * 128 + ewarp * 32 + lane;
                const uint32_t trow = tmem + ((uint32_t)(ewarp * 32) << 16) + (uint32_t)(half * 256);
                for (int k = 0; k < g.tpt; ++k) {
                    float best = -INFINITY, second = -INFINITY;
                    int arg = 0;
                    for (int c8 = 0; c8 < g.V; c8 += 8) {
                        float v[8];
                        tmem_ld8(trow + (uint32_t)(k * g.V + c8), v);
#pragma unroll
                        for (int i = 0; i < 8; ++i) {
                            const int c = k * g.V + c8 + i;
                            const float y = fmaf(v[i], coef_s[c], coef_s[256 + c]);
                            if (y > best) { second = best; best = y; arg = c8 + i; }
                            else if (y > second) second = y;
                        }
                    }
                    const int tok = tok0 + k;
                    if (row < g.n && tok < g.T) {
                        tokens[(long long)row * g.T + tok] = arg;
                        if (!(best - second > DEC_MARGIN * fmaxf(1.f, fabsf(best)))) {      // also catches NaN
                            const int slot = atomicAdd(flag_count, 1);
                            if (slot < flag_cap) flagged[slot] = row * g.T + tok;
                        }
                    }
                }
            }
            tc_fence_before();
            mbar_arrive(sTEmpty);
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 5) tmem_dealloc(tmem, 512u);
}

// One warp per token to recompute: z[j] = b[j] + sum_l Tab[l * 32 + w_l][j] in double from the fp32 table, second BN in
// double, first maximum wins.  list == nullptr: every (genotype, token) pair (shapes the tensor-core kernel does not serve).
__global__ void __launch_bounds__(256)
k_dec_recheck(const int* __restrict__ list, long long count, const uint8_t* __restrict__ win, const float* __restrict__ tab,
              const double* __restrict__ inv2, const double* __restrict__ shift2, const float* __restrict__ bias,
              int L, int N, int T, int V, int* __restrict__ tokens) {
    const long long wi = (long long)blockIdx.x * 8 + (threadIdx.x >> 5);
    if (wi >= count) return;
    const int lane = threadIdx.x & 31;
    const long long id = list ? list[wi] : wi;
    const long long row = id / T;
    const int tok = (int)(id % T);
    const uint8_t* w = win + row * L;
    double best = -INFINITY;
    int arg = 0x7fffffff;
    for (int v0 = 0; v0 < V; v0 += 32) {
        const int v = v0 + lane;
        double y = -INFINITY;
        if (v < V) {
            const int j = tok * V + v;
            double acc = 0.0;
            for (int l = 0; l < L; ++l) acc += (double)tab[((long long)l * DEC_WIN + w[l]) * N + j];
            y = (acc + (double)bias[j]) * inv2[v] + shift2[v];
        }
        if (y > best || (y == best && v < arg)) { best = y; arg = v; }
    }
    for (int o = 16; o > 0; o >>= 1) {
        const double ob = __shfl_xor_sync(0xffffffffu, best, o);
        const int oa = __shfl_xor_sync(0xffffffffu, arg, o);
        if (ob > best || (ob == best && oa < arg)) { best = ob; arg = oa; }
    }
    if (lane == 0) tokens[row * T + tok] = arg == 0x7fffffff ? 0 : arg;
}

}  // namespace

struct srn_decoder {
    int device = 0, G = 0, T = 0, V = 0, L = 0, N = 0;
    bool tensor_path = false;
    int tpt = 0, BN = 0;
    float scale = 1.f;
    float* tab = nullptr;                 // fp32 table [L * 32][N]
    __half *hi = nullptr, *lo = nullptr;  // its fp16 split
    float *acol = nullptr, *ccol = nullptr, *bias = nullptr;
    double *inv2 = nullptr, *shift2 = nullptr;
    CUtensorMap* maps = nullptr;          // device copy of the two tensor maps (hi, lo)
    // grow-only working buffers
    uint8_t* bits = nullptr; uint8_t* win = nullptr; int* tokens = nullptr; int* flagged = nullptr; int* flag_count = nullptr;
    long long cap_rows = 0; int flag_cap = 0;
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    int num_sms = 148;
    // statistics of the last decode
    long long last_flagged = 0;
};

extern "C" void srn_decoder_destroy(srn_decoder* d) {
    if (!d) return;
    cudaSetDevice(d->device);
    cudaFree(d->tab); cudaFree(d->hi); cudaFree(d->lo); cudaFree(d->acol); cudaFree(d->ccol); cudaFree(d->bias);
    cudaFree(d->inv2); cudaFree(d->shift2); cudaFree(d->maps);
    cudaFree(d->bits); cudaFree(d->win); cudaFree(d->tokens); cudaFree(d->flagged); cudaFree(d->flag_count);
    if (d->e0) cudaEventDestroy(d->e0);
    if (d->e1) cudaEventDestroy(d->e1);
    delete d;
}

template <class T>
static int upload(T** dst, const T* src, size_t n) {
    CK(dmalloc(reinterpret_cast<void**>(dst), n * sizeof(T)));
    CK(cudaMemcpy(*dst, src, n * sizeof(T), cudaMemcpyHostToDevice));
    return SRN_OK;
}

static int decoder_setup(srn_decoder* d, const float* conv_kernel, const float* conv_bias, const float* bn1,
                         const float* dense_kernel, const float* dense_bias, const float* bn2) {
    const int L = d->L, N = d->N, V = d->V;
    CK(cudaDeviceGetAttribute(&d->num_sms, cudaDevAttrMultiProcessorCount, d->device));
    CK(cudaEventCreate(&d->e0)); CK(cudaEventCreate(&d->e1));
    // first two layers as a table over the five-bit window: LUT[w][f] = BN1(bias[f] + sum_t K[t][bit_t(w)][f])
    std::vector<double> lut((size_t)DEC_WIN * DEC_F);
    for (int w = 0; w < DEC_WIN; ++w)
        for (int f = 0; f < DEC_F; ++f) {
            float acc = 0.f;                              // fp32 like the reference's convolution
            for (int t = 0; t < DEC_KW; ++t) acc += conv_kernel[(t * 2 + ((w >> t) & 1)) * DEC_F + f];      // Keras layout (5, 2, 32)
            acc += conv_bias[f];
            const float inv = (1.0f / sqrtf(bn1[3 * DEC_F + f] + KERAS_BN_EPS)) * bn1[f];
            lut[(size_t)w * DEC_F + f] = (double)(acc * inv + (bn1[DEC_F + f] - bn1[2 * DEC_F + f] * inv));
        }
    // second BN per vocabulary index: y = (z + b) * inv2 + shift2
    std::vector<double> inv2((size_t)V), shift2((size_t)V);
    for (int v = 0; v < V; ++v) {
        const float inv = (1.0f / sqrtf(bn2[3 * V + v] + KERAS_BN_EPS)) * bn2[v];
        inv2[(size_t)v] = inv;
        shift2[(size_t)v] = (double)bn2[V + v] - (double)bn2[2 * V + v] * inv;
    }
    int rc;
    if ((rc = upload(&d->inv2, inv2.data(), (size_t)V))) return rc;
    if ((rc = upload(&d->shift2, shift2.data(), (size_t)V))) return rc;
    if ((rc = upload(&d->bias, dense_bias, (size_t)N))) return rc;
    // the table, on the device
    double* d_lut = nullptr; float* d_W = nullptr; unsigned* d_max = nullptr;
    auto cleanup = [&]() { cudaFree(d_lut); cudaFree(d_W); cudaFree(d_max); };
    const long long total = (long long)L * DEC_WIN * N;
#define CKS(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { cleanup(); return fail(SRN_ERR_CUDA, "%s: %s", #call, cudaGetErrorString(e_)); } } while (0)
    CKS(dmalloc((void**)&d_lut, lut.size() * sizeof(double)));
    CKS(cudaMemcpy(d_lut, lut.data(), lut.size() * sizeof(double), cudaMemcpyHostToDevice));
    CKS(dmalloc((void**)&d_W, (size_t)L * DEC_F * N * sizeof(float)));
    CKS(cudaMemcpy(d_W, dense_kernel, (size_t)L * DEC_F * N * sizeof(float), cudaMemcpyHostToDevice));
    CKS(dmalloc((void**)&d_max, sizeof(unsigned)));
    CKS(cudaMemset(d_max, 0, sizeof(unsigned)));
    CKS(dmalloc((void**)&d->tab, (size_t)total * sizeof(float)));
    k_dec_table<<<(unsigned)((total + 255) / 256), 256>>>(d_lut, d_W, L, N, d->tab, d_max);
    g_launches.fetch_add(1);
    unsigned maxbits = 0;
    CKS(cudaMemcpy(&maxbits, d_max, sizeof(unsigned), cudaMemcpyDeviceToHost));
    CKS(cudaGetLastError());
    float maxabs;
    memcpy(&maxabs, &maxbits, sizeof(float));
    cleanup();
    d_lut = nullptr; d_W = nullptr; d_max = nullptr;
    if (!std::isfinite(maxabs)) return fail(SRN_ERR_ARG, "srn_decoder_create: non-finite decoder weights");
    // tensor-core path: whole tokens per tile, 8-aligned vocabulary, window bytes in shared memory
    d->tpt = 0;
    if (V % 8 == 0 && V <= 256 && L <= DEC_MAX_L && encode_tiled_fn() != nullptr && (N * 2) % 16 == 0)
        for (int t = 256 / V; t >= 1; --t)
            if ((t * V) % 16 == 0) { d->tpt = t; break; }
    d->tensor_path = d->tpt > 0 && getenv("SRN_DEC_EXACT") == nullptr;
    if (d->tensor_path) {
        d->BN = d->tpt * V;
        int e = 0;                                                    // scale so that the largest entry lies in [8192, 16384)
        if (maxabs > 0.f) { frexpf(maxabs, &e); e = 14 - e; }
        d->scale = ldexpf(1.f, e);
        CK(dmalloc((void**)&d->hi, (size_t)total * sizeof(__half)));
        CK(dmalloc((void**)&d->lo, (size_t)total * sizeof(__half)));
        k_dec_split<<<(unsigned)((total + 255) / 256), 256>>>(d->tab, total, d->scale, d->hi, d->lo);
        g_launches.fetch_add(1);
        std::vector<float> acol((size_t)N), ccol((size_t)N);
        for (int j = 0; j < N; ++j) {
            const int v = j % V;
            acol[(size_t)j] = (float)(inv2[(size_t)v] / (double)d->scale);
            ccol[(size_t)j] = (float)((double)dense_bias[j] * inv2[(size_t)v] + shift2[(size_t)v]);
        }
        if ((rc = upload(&d->acol, acol.data(), (size_t)N))) return rc;
        if ((rc = upload(&d->ccol, ccol.data(), (size_t)N))) return rc;
        CUtensorMap maps[2];
        const long long K = (long long)L * DEC_WIN;
        if (!encode_matrix(&maps[0], d->hi, K, N, N, DEC_BK) || !encode_matrix(&maps[1], d->lo, K, N, N, DEC_BK))
            return fail(SRN_ERR_CUDA, "srn_decoder_create: tensor map encoding failed");
        if ((rc = upload(&d->maps, maps, 2))) return rc;
        CK(cudaFuncSetAttribute(k_dec_gemm, cudaFuncAttributeMaxDynamicSharedMemorySize, DEC_SMEM));
    }
    CK(dmalloc((void**)&d->flag_count, sizeof(int)));
    CK(cudaDeviceSynchronize());
    CK(cudaGetLastError());
    return SRN_OK;
}

extern "C" int srn_decoder_create(int32_t device, int32_t genotype_size, int32_t max_tokens, int32_t vocab,
                                  const float* conv_kernel, const float* conv_bias, const float* bn1,
                                  const float* dense_kernel, const float* dense_bias, const float* bn2,
                                  srn_decoder** out) {
    if (!out || !conv_kernel || !conv_bias || !bn1 || !dense_kernel || !dense_bias || !bn2) return fail(SRN_ERR_ARG, "srn_decoder_create: null argument");
    if (genotype_size < DEC_KW || max_tokens <= 0 || vocab <= 0) return fail(SRN_ERR_ARG, "srn_decoder_create: bad dimensions");
    if ((long long)max_tokens * vocab > 0x3fffffffLL) return fail(SRN_ERR_LIMIT, "srn_decoder_create: max_tokens * vocab too large");
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0) return fail(SRN_ERR_CUDA, "no CUDA device (%s); this library has no CPU path", cudaGetErrorString(e));
    if (device < 0 || device >= count) return fail(SRN_ERR_ARG, "device %d of %d", device, count);
    CK(cudaSetDevice(device));
    srn_decoder* d = new srn_decoder();
    d->device = device; d->G = genotype_size; d->T = max_tokens; d->V = vocab;
    d->L = genotype_size - DEC_KW + 1; d->N = max_tokens * vocab;
    const int rc = decoder_setup(d, conv_kernel, conv_bias, bn1, dense_kernel, dense_bias, bn2);
    if (rc) { srn_decoder_destroy(d); return rc; }       // every upload made so far is given back
    *out = d;
    return SRN_OK;
}

extern "C" int srn_decoder_decode(srn_decoder* d, const uint8_t* bits, int32_t n, int32_t* tokens, float* device_ms) {
    if (!d || !bits || !tokens || n < 0) return fail(SRN_ERR_ARG, "srn_decoder_decode: bad argument");
    if (device_ms) *device_ms = 0.f;
    if (n == 0) return SRN_OK;
    CK(cudaSetDevice(d->device));
    const int CH = 1 << 18;                    // rows per pass: bounds the token buffer to CH * T ints
    const long long cap = std::min<long long>(n, CH);
    if (cap > d->cap_rows) {                   // grow-only working buffers
        cudaFree(d->bits); cudaFree(d->win); cudaFree(d->tokens); cudaFree(d->flagged);
        d->bits = nullptr; d->win = nullptr; d->tokens = nullptr; d->flagged = nullptr; d->cap_rows = 0;
        const long long rows = std::min<long long>(CH, cap + cap / 4);
        CK(dmalloc((void**)&d->bits, (size_t)rows * d->G));
        CK(dmalloc((void**)&d->win, (size_t)rows * d->L + 64));
        CK(dmalloc((void**)&d->tokens, (size_t)rows * d->T * sizeof(int)));
        d->flag_cap = (int)std::min<long long>(rows * d->T, 1 << 22);
        CK(dmalloc((void**)&d->flagged, (size_t)d->flag_cap * sizeof(int)));
        d->cap_rows = rows;
    }
    float total_ms = 0.f;
    d->last_flagged = 0;
    for (long long s = 0; s < n; s += d->cap_rows) {
        const int m = (int)std::min<long long>(d->cap_rows, n - s);
        CK(cudaMemcpy(d->bits, bits + (size_t)s * d->G, (size_t)m * d->G, cudaMemcpyHostToDevice));
        CK(cudaEventRecord(d->e0, 0));
        const long long we = (long long)m * d->L;
        k_dec_windows<<<(unsigned)((we + 255) / 256), 256>>>(d->bits, m, d->G, d->L, d->win);
        g_launches.fetch_add(1);
        const long long all = (long long)m * d->T;
        int flagged = 0;
        bool dense = !d->tensor_path;
        if (d->tensor_path) {
            CK(cudaMemsetAsync(d->flag_count, 0, sizeof(int), 0));
            DecGeom g;
            g.n = m; g.L = d->L; g.N = d->N; g.T = d->T; g.V = d->V; g.tpt = d->tpt; g.BN = d->BN;
            g.nkb = (d->L + 1) / 2;
            g.m_tiles = (m + DEC_M - 1) / DEC_M;
            g.n_tiles = (d->T + d->tpt - 1) / d->tpt;
            const int grid = std::min(g.m_tiles * g.n_tiles, d->num_sms);
            k_dec_gemm<<<grid, DEC_THREADS, DEC_SMEM>>>(d->maps, d->win, d->acol, d->ccol, g, d->tokens, d->flagged, d->flag_cap, d->flag_count);
            g_launches.fetch_add(1);
            CK(cudaMemcpy(&flagged, d->flag_count, sizeof(int), cudaMemcpyDeviceToHost));      // waits for the kernel
            CK(cudaGetLastError());
            if (flagged > d->flag_cap) dense = true;       // list overflow (degenerate weights): recompute everything exactly
            else if (flagged > 0) {
                k_dec_recheck<<<(unsigned)((flagged + 7) / 8), 256>>>(d->flagged, flagged, d->win, d->tab, d->inv2, d->shift2, d->bias,
                                                                     d->L, d->N, d->T, d->V, d->tokens);
                g_launches.fetch_add(1);
            }
            d->last_flagged += flagged;
        }
        if (dense) {
            k_dec_recheck<<<(unsigned)((all + 7) / 8), 256>>>(nullptr, all, d->win, d->tab, d->inv2, d->shift2, d->bias,
                                                             d->L, d->N, d->T, d->V, d->tokens);
            g_launches.fetch_add(1);
        }
        CK(cudaEventRecord(d->e1, 0));
        CK(cudaEventSynchronize(d->e1));
        CK(cudaGetLastError());
        float ms = 0.f;
        CK(cudaEventElapsedTime(&ms, d->e0, d->e1));
        total_ms += ms;
        CK(cudaMemcpy(tokens + (size_t)s * d->T, d->tokens, (size_t)m * d->T * sizeof(int), cudaMemcpyDeviceToHost));
    }
    if (device_ms) *device_ms = total_ms;
    return SRN_OK;
}

// Tokens sent to the exact recheck by the last srn_decoder_decode call, and whether the tensor-core kernel served it
extern "C" int srn_decoder_stats(srn_decoder* d, int64_t* rechecked, int32_t* tensor_path) {
    if (!d) return fail(SRN_ERR_ARG, "srn_decoder_stats: null decoder");
    if (rechecked) *rechecked = d->last_flagged;
    if (tensor_path) *tensor_path = d->tensor_path ? 1 : 0;
    return SRN_OK;
}
