// Micro-benchmark: how fast can one SM's copy engine fill shared memory with the im2col operand of a small-channel
// convolution, against the quarter-warp LDGSTS gather the GEMM kernel uses?  ("measure before
// arguing".)  Shape: NHWC fp16 activations [750][26][26][C], 3 x 3 window, stride 1, no padding -> 750 * 24 * 24 =
// 432 000 output pixels = 3375 tiles of 128 pixels; one ring stage = one window tap of one tile = 128 pixels x C
// channels (C * 2 = 16, 32, 64 or 128 bytes per pixel).  Nothing consumes the data: a consumer warp hands the stage
// back as soon as it has landed, so the figure is the fill rate of the load path alone.
//   tma     cp.async.bulk.tensor.4d ... im2col, one elected thread, box {C, 128 pixels}, swizzle = C * 2 bytes
//   ldgsts  256 threads, 16-byte cp.async, 4 rows x 8 chunks per warp instruction where the row allows it,
//           completion by cp.async.mbarrier.arrive.noinc (the kernel's scheme)
// The tensor maps cover this GPU's memory only.  The first tile of every variant is copied back and checked against
// a host im2col (swizzle undone), so the coordinate / corner conventions of the im2col mode are pinned by the run.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tma_small_box tma_small_box.cu && ./tma_small_box
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cstring>
#include <vector>
#include <cuda.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1); } } while (0)

constexpr int NB = 750, H = 26, W = 26, KH = 3, KW = 3, HO = H - KH + 1, WO = W - KW + 1;
constexpr int M = NB * HO * WO, TILES = M / 128;          // 432 000 / 128 = 3375 exactly
constexpr int STAGES = 8;
constexpr int ROUNDS = 16;          // passes over the tile set per timed launch (the activations stay L2-resident, as they do between the taps of the real kernel)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;\n" ::"r"(bar), "r"(count) : "memory"); }
__device__ __forceinline__ void mbar_arrive(uint32_t bar) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];\n" ::"r"(bar) : "memory"); }
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" ::"r"(bar), "r"(bytes) : "memory"); }
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    asm volatile("{\n.reg .pred P1;\nWAIT_LOOP:\nmbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1, 20000;\n@P1 bra WAIT_DONE;\nbra WAIT_LOOP;\nWAIT_DONE:\n}\n" ::"r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ void tma_im2col_4d(uint32_t dst, const void* tmap, int c, int w, int h, int n, uint16_t woff, uint16_t hoff, uint32_t bar) {
    asm volatile("cp.async.bulk.tensor.4d.shared::cluster.global.im2col.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2], {%7, %8};\n"
                 ::"r"(dst), "l"(tmap), "r"(bar), "r"(c), "r"(w), "r"(h), "r"(n), "h"(woff), "h"(hoff) : "memory");
}

// stage s: [128 pixels][C] (C * 2 bytes per row, swizzled by the copy engine / by the gather)
template <int C, bool TMA>
__global__ void __launch_bounds__(320, 2) k_fill(const __grid_constant__ CUtensorMap tm, const __half* __restrict__ x, int tiles, __half* dump) {
    extern __shared__ __align__(1024) unsigned char sm[];
    constexpr uint32_t STAGE = 128 * C * 2;
    const uint32_t sbase = smem_u32(sm), sFull = sbase + STAGES * STAGE, sEmpty = sFull + STAGES * 8;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    if (tid == 0) {
        for (int i = 0; i < STAGES; ++i) { mbar_init(sFull + i * 8, TMA ? 1 : 256); mbar_init(sEmpty + i * 8, 1); }
        asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
    }
    __syncthreads();
    if (warp < 8) {
        int st = 0; uint32_t ph = 1;
        for (int tt = blockIdx.x; tt < tiles; tt += gridDim.x) {
            const int t = tt % TILES, m0 = t * 128;
            for (int tap = 0; tap < KH * KW; ++tap) {
                const int ki = tap / KW, kj = tap % KW;
                if (TMA) {
                    if (tid == 0) {
                        mbar_wait(sEmpty + st * 8, ph);
                        const int n = m0 / (HO * WO), rem = m0 % (HO * WO), ho = rem / WO, wo = rem % WO;
                        mbar_arrive_expect_tx(sFull + st * 8, STAGE);
                        tma_im2col_4d(sbase + st * STAGE, &tm, 0, wo, ho, n, (uint16_t)kj, (uint16_t)ki, sFull + st * 8);
                    }
                } else {
                    mbar_wait(sEmpty + st * 8, ph);
                    // chunks per row CPR = C / 8; a warp instruction covers 32 / CPR rows; 128 * CPR chunks per stage over 256 threads
                    constexpr int CPR = C / 8, PER = (128 * CPR + 255) / 256;
#pragma unroll
                    for (int i = 0; i < PER; ++i) {
                        const int idx = i * 256 + tid;
                        if (128 * CPR >= 256 || idx < 128 * CPR) {
                            const int row = idx / CPR, ch = idx % CPR, m = m0 + row;
                            const int n = m / (HO * WO), rem = m % (HO * WO), ho = rem / WO, wo = rem % WO;
                            const __half* src = x + (((size_t)n * H + ho + ki) * W + wo + kj) * C + ch * 8;
                            // row pitch C * 2 bytes; 16-byte chunk index XOR-ed with the row index inside the swizzle span, as the copy engine does
                            constexpr int SPAN = CPR;            // chunks per swizzle span (span = row for C <= 64)
                            const int rsw = SPAN == 8 ? (row & 7) : (SPAN == 4 ? ((row >> 1) & 3) : (SPAN == 2 ? ((row >> 2) & 1) : 0));
                            const uint32_t dst = sbase + st * STAGE + row * (C * 2) + ((ch ^ rsw) << 4);
                            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(dst), "l"(src) : "memory");
                        }
                    }
                    asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];\n" ::"r"(sFull + st * 8) : "memory");
                }
                if (++st == STAGES) { st = 0; ph ^= 1u; }
            }
        }
    } else if (warp == 8) {
        int st = 0; uint32_t ph = 0;
        bool first = blockIdx.x == 0 && dump != nullptr;
        int taps_dumped = 0;
        for (int tt = blockIdx.x; tt < tiles; tt += gridDim.x) {
            for (int tap = 0; tap < KH * KW; ++tap) {
                mbar_wait(sFull + st * 8, ph);
                if (first && taps_dumped < KH * KW) {          // copy the first tile's taps back for the host check
                    asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");
                    const __half* s = reinterpret_cast<const __half*>(sm + st * STAGE);
                    for (int i = lane; i < 128 * C; i += 32) dump[(size_t)taps_dumped * 128 * C + i] = s[i];
                    __syncwarp();
                    ++taps_dumped;
                }
                if (lane == 0) mbar_arrive(sEmpty + st * 8);
                if (++st == STAGES) { st = 0; ph ^= 1u; }
            }
        }
    }
}

typedef CUresult (*EncodeIm2colFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const int*, const int*,
                                   cuuint32_t, cuuint32_t, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

template <int C>
static void run(const std::vector<__half>& hx, const __half* dx, EncodeIm2colFn enc) {
    CUtensorMap tm;
    memset(&tm, 0, sizeof(tm));
    const cuuint64_t dims[4] = {(cuuint64_t)C, W, H, NB};
    const cuuint64_t strides[3] = {(cuuint64_t)C * 2, (cuuint64_t)W * C * 2, (cuuint64_t)H * W * C * 2};
    const int lower[2] = {0, 0}, upper[2] = {-(KW - 1), -(KH - 1)};      // {W, H}: window base positions 0 .. W - KW, 0 .. H - KH
    const cuuint32_t estr[4] = {1, 1, 1, 1};
    const CUtensorMapSwizzle sw = C == 8 ? CU_TENSOR_MAP_SWIZZLE_NONE : (C == 16 ? CU_TENSOR_MAP_SWIZZLE_32B : (C == 32 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_128B));
    const CUresult r = enc(&tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 4, const_cast<__half*>(dx), dims, strides, lower, upper, C, 128, estr,
                           CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { printf("C=%d: cuTensorMapEncodeIm2col failed (%d)\n", C, (int)r); return; }
    const size_t smem = (size_t)STAGES * 128 * C * 2 + 2 * STAGES * 8 + 64;
    __half* dump = nullptr;
    CK(cudaMalloc(&dump, (size_t)KH * KW * 128 * C * 2));
    int sms = 0;
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    for (int variant = 0; variant < 2; ++variant) {
        auto kern = variant == 0 ? k_fill<C, true> : k_fill<C, false>;
        CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        const int grid = 2 * sms;
        // check run
        CK(cudaMemset(dump, 0xFF, (size_t)KH * KW * 128 * C * 2));
        kern<<<grid, 320, smem>>>(tm, dx, TILES, dump);
        CK(cudaDeviceSynchronize());
        std::vector<__half> hd((size_t)KH * KW * 128 * C);
        CK(cudaMemcpy(hd.data(), dump, hd.size() * 2, cudaMemcpyDeviceToHost));
        long long bad = 0;
        constexpr int CPR = C / 8;
        for (int tap = 0; tap < KH * KW; ++tap)
            for (int row = 0; row < 128; ++row)
                for (int c = 0; c < C; ++c) {
                    const int ki = tap / KW, kj = tap % KW, n = row / (HO * WO), rem = row % (HO * WO), ho = rem / WO, wo = rem % WO;
                    const __half want = hx[(((size_t)n * H + ho + ki) * W + wo + kj) * C + c];
                    const int ch = c / 8;
                    const int rsw = CPR == 8 ? (row & 7) : (CPR == 4 ? ((row >> 1) & 3) : (CPR == 2 ? ((row >> 2) & 1) : 0));
                    const __half got = hd[(size_t)tap * 128 * C + (size_t)row * C + ((ch ^ rsw) * 8) + (c & 7)];
                    if (__half2float(want) != __half2float(got)) ++bad;
                }
        // timed runs
        cudaEvent_t e0, e1;
        CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
        for (int i = 0; i < 3; ++i) kern<<<grid, 320, smem>>>(tm, dx, TILES * ROUNDS, nullptr);
        CK(cudaEventRecord(e0));
        const int reps = 10;
        for (int i = 0; i < reps; ++i) kern<<<grid, 320, smem>>>(tm, dx, TILES * ROUNDS, nullptr);
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        float ms = 0.f;
        CK(cudaEventElapsedTime(&ms, e0, e1));
        ms /= reps;
        const double bytes = (double)TILES * ROUNDS * KH * KW * 128 * C * 2;
        printf("C=%3d (%3d B/pixel) %-6s: %8.3f ms  %7.1f GB/s into shared memory (%.2f GB), %.0f k-tap stages/s/SM, first tile %s (%lld wrong)\n", C, C * 2,
               variant == 0 ? "tma" : "ldgsts", ms, bytes / ms / 1e6, bytes / 1e9, (double)TILES * ROUNDS * KH * KW / (ms * 1e-3) / sms, bad ? "MISMATCH" : "ok", bad);
        CK(cudaEventDestroy(e0)); CK(cudaEventDestroy(e1));
    }
    CK(cudaFree(dump));
}

int main() {
    void* f = nullptr;
    cudaDriverEntryPointQueryResult q;
    CK(cudaFree(0));
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeIm2col", &f, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) { printf("no cuTensorMapEncodeIm2col\n"); return 1; }
    EncodeIm2colFn enc = reinterpret_cast<EncodeIm2colFn>(f);
    const size_t nmax = (size_t)NB * H * W * 64;
    std::vector<__half> hx(nmax);
    uint32_t s = 12345u;
    for (size_t i = 0; i < nmax; ++i) { s = s * 1664525u + 1013904223u; hx[i] = __float2half((float)((s >> 16) & 1023) / 64.f - 8.f); }
    __half* dx = nullptr;
    CK(cudaMalloc(&dx, nmax * 2));
    CK(cudaMemcpy(dx, hx.data(), nmax * 2, cudaMemcpyHostToDevice));
    printf("im2col operand of a 3x3 window over [750][26][26][C] fp16, 3375 tiles x 9 taps, %d stages of 128 x C, grid 2 CTAs/SM\n", STAGES);
    run<8>(hx, dx, enc);
    run<16>(hx, dx, enc);
    run<32>(hx, dx, enc);
    run<64>(hx, dx, enc);
    CK(cudaFree(dx));
    return 0;
}
