// Micro-benchmark / feasibility check: a stride-1 convolution on tcgen05 WITHOUT re-fetching every input pixel kh*kw
// times.  The GEMM kernel's loaders fill shared memory with the im2col operand tap by tap (9x - 49x the input bytes,
// LSU-bound: DESIGN.md section 4).  Here a tile is 128 consecutive positions p = ho * W + wo of one sample's INPUT grid;
// the tile's input rows [p0, p0 + 128 + (kh-1) W + kw-1) are loaded ONCE (one 2-D TMA, swizzled K-major rows of C
// channels) and the A operand of tap (ki, kj) is the same slab read through a matrix descriptor whose start address is
// advanced by (ki W + kj) rows - a row offset that is not a multiple of the 8-row swizzle atom.  Positions with
// wo >= Wo are computed and dropped (Wo / W of the MMA work is useful).  The weights stay resident in shared memory.
// Shape: [750][26][26][32] fp16, 3 x 3, 32 filters (K = 288), ReLU-free plain store; output checked against a host
// convolution for the first samples.  Variant flag: 0 = descriptor base_offset 0, 1 = base_offset = (start >> 7) & 7.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o conv_slab conv_slab.cu && ./conv_slab [variant]
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cstring>
#include <cmath>
#include <vector>
#include <cuda.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1); } } while (0)

constexpr int NB = 750, H = 26, W = 26, C = 32, KH = 3, KW = 3, F = 32, HO = H - KH + 1, WO = W - KW + 1;
constexpr int P = (HO - 1) * W + WO;                    // input-grid positions that can hold a window's top-left corner
constexpr int TPS = (P + 127) / 128;                    // tiles per sample
constexpr int HALO = (KH - 1) * W + (KW - 1);
constexpr int SLAB_ROWS = 128 + HALO;                   // 182
constexpr int SLAB_BOX = (SLAB_ROWS + 7) / 8 * 8;       // 184 rows per TMA box (<= 256)
constexpr int ROWB = C * 2;                             // 64 bytes per pixel: swizzle-64B rows
constexpr int SLAB_BYTES = SLAB_BOX * ROWB;             // 11 776
constexpr int SLAB_STRIDE = (SLAB_BYTES + 1023) / 1024 * 1024;
constexpr int K = KH * KW * C;                          // 288
constexpr int WB_BYTES = K * 128;                       // weights, MN-major swizzle-128B rows of 64 n (32 used)
constexpr int NSLAB = 3, NACC = 2;
constexpr int THREADS = 192;                            // warp 0: TMA, warp 1: MMA issuer, warps 2-5: epilogue

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;\n" ::"r"(bar), "r"(count) : "memory"); }
__device__ __forceinline__ void mbar_arrive(uint32_t bar) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];\n" ::"r"(bar) : "memory"); }
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" ::"r"(bar), "r"(bytes) : "memory"); }
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    asm volatile("{\n.reg .pred P1;\nWAIT_LOOP:\nmbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1, 20000;\n@P1 bra WAIT_DONE;\nbra WAIT_LOOP;\nWAIT_DONE:\n}\n" ::"r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const void* tmap, int c0, int c1, uint32_t bar) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];\n"
                 ::"r"(dst), "l"(tmap), "r"(bar), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;\n" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;\n" ::: "memory"); }
__device__ __forceinline__ bool elect_one_sync() {
    uint32_t pred = 0;
    asm volatile("{\n.reg .pred P1;\nelect.sync _|P1, 0xffffffff;\nselp.u32 %0, 1, 0, P1;\n}\n" : "=r"(pred));
    return pred != 0;
}
// shared-memory matrix descriptor: start >> 4 | LBO >> 4 << 16 | SBO >> 4 << 32 | version 1 << 46 | base_offset << 49 | layout << 61
__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr, uint32_t lbo, uint32_t sbo, uint32_t layout, uint32_t base_off) {
    return (uint64_t)((saddr & 0x3FFFFu) >> 4) | ((uint64_t)(lbo >> 4) << 16) | ((uint64_t)(sbo >> 4) << 32) | (1ull << 46) |
           ((uint64_t)(base_off & 7) << 49) | ((uint64_t)layout << 61);
}
__device__ __forceinline__ void umma_f16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n}\n"
                 ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) { asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n" ::"r"(bar) : "memory"); }
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];\n"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
          "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
          "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr) : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;\n" ::: "memory");
}

// x: [NB * H * W][C], w: [K][F] (k = (ki, kj, c)), out: [NB][HO][WO][F]
__global__ void __launch_bounds__(THREADS, 2) k_conv_slab(const __grid_constant__ CUtensorMap tm, const __half* __restrict__ wgt, __half* __restrict__ out, int tiles, int variant) {
    extern __shared__ __align__(1024) unsigned char sm[];
    const uint32_t sbase = smem_u32(sm);
    const uint32_t sW = sbase, sSlab = sbase + (WB_BYTES + 1023) / 1024 * 1024;
    const uint32_t sBar = sSlab + NSLAB * SLAB_STRIDE;          // full[NSLAB], empty[NSLAB], tfull[NACC], tempty[NACC], slot
    const uint32_t sFull = sBar, sEmpty = sBar + 8 * NSLAB, sTFull = sEmpty + 8 * NSLAB, sTEmpty = sTFull + 8 * NACC, sSlot = sTEmpty + 8 * NACC;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    if (tid == 0) {
        for (int i = 0; i < NSLAB; ++i) { mbar_init(sFull + 8 * i, 1); mbar_init(sEmpty + 8 * i, 1); }
        for (int i = 0; i < NACC; ++i) { mbar_init(sTFull + 8 * i, 1); mbar_init(sTEmpty + 8 * i, 4); }
        asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;\n" ::"r"(sSlot), "r"(64u) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;\n" ::: "memory");
    }
    // weights -> MN-major swizzle-128B rows: k row kl at (kl >> 3) * 1024 + (kl & 7) * 128, 16-byte chunk u at (u ^ (kl & 7)) << 4
    for (int i = tid; i < K * 4; i += THREADS) {
        const int kl = i >> 2, u = i & 3;
        const uint4 v = *reinterpret_cast<const uint4*>(wgt + (size_t)kl * F + u * 8);
        *reinterpret_cast<uint4*>(sm + (kl >> 3) * 1024 + (kl & 7) * 128 + ((u ^ (kl & 7)) << 4)) = v;
    }
    asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *reinterpret_cast<volatile uint32_t*>(sm + (sSlot - sbase));
    const int G = gridDim.x;
    if (warp == 0) {
        if (lane == 0) {
            int st = 0; uint32_t ph = 1;
            for (int t = blockIdx.x; t < tiles; t += G) {
                const int b = t / TPS, p0 = (t - b * TPS) * 128;
                mbar_wait(sEmpty + 8 * st, ph);
                mbar_arrive_expect_tx(sFull + 8 * st, SLAB_BYTES);
                tma_load_2d(sSlab + st * SLAB_STRIDE, &tm, 0, b * H * W + p0, sFull + 8 * st);
                if (++st == NSLAB) { st = 0; ph ^= 1u; }
            }
        }
    } else if (warp == 1) {
        int st = 0; uint32_t fph = 0; int j = 0;
        // instruction descriptor kind::f16: D f32, A = B = f16, A K-major, B MN-major, N = 32, M = 128
        const uint32_t idesc = (1u << 4) | (0u << 15) | (1u << 16) | ((uint32_t)(F >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
        for (int t = blockIdx.x; t < tiles; t += G, ++j) {
            const int acc = j % NACC;
            if (j >= NACC) { mbar_wait(sTEmpty + 8 * acc, (uint32_t)((j / NACC - 1) & 1)); }
            mbar_wait(sFull + 8 * st, fph);
            tc_fence_after();
            const uint32_t slab = sSlab + st * SLAB_STRIDE, tacc = tmem + (uint32_t)(acc * 32);
            if (elect_one_sync()) {
#pragma unroll
                for (int tap = 0; tap < KH * KW; ++tap) {
                    const int off = (tap / KW) * W + (tap % KW);
#pragma unroll
                    for (int s = 0; s < C / 16; ++s) {
                        const uint32_t a0 = slab + off * ROWB + s * 32;
                        const uint64_t dA = umma_desc(a0, 16, 8 * ROWB, 4 /* swizzle 64B */, variant == 1 ? (a0 >> 7) & 7 : 0);
                        const uint64_t dB = umma_desc(sW + (tap * (C / 16) + s) * 2048, 8192, 1024, 2 /* swizzle 128B */, 0);
                        umma_f16(tacc, dA, dB, idesc, (tap | s) ? 1u : 0u);
                    }
                }
                umma_commit(sEmpty + 8 * st);
                umma_commit(sTFull + 8 * acc);
            }
            __syncwarp();
            if (++st == NSLAB) { st = 0; fph ^= 1u; }
        }
    } else {
        const int q = warp & 3;                     // TMEM lanes 32 q .. 32 q + 31
        int j = 0;
        for (int t = blockIdx.x; t < tiles; t += G, ++j) {
            const int acc = j % NACC;
            const int b = t / TPS, p = (t - b * TPS) * 128 + q * 32 + lane;
            mbar_wait(sTFull + 8 * acc, (uint32_t)((j / NACC) & 1));
            tc_fence_after();
            uint32_t r[32];
            tmem_ld32(tmem + (uint32_t)(acc * 32) + ((uint32_t)(q * 32) << 16), r);
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(sTEmpty + 8 * acc);
            const int ho = p / W, wo = p - ho * W;
            if (p < P && wo < WO) {
                __half* o = out + (((size_t)b * HO + ho) * WO + wo) * F;
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    __half2 h[4];
#pragma unroll
                    for (int i = 0; i < 4; ++i) h[i] = __floats2half2_rn(__uint_as_float(r[c * 8 + 2 * i]), __uint_as_float(r[c * 8 + 2 * i + 1]));
                    *reinterpret_cast<uint4*>(o + c * 8) = *reinterpret_cast<const uint4*>(h);
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;\n" ::"r"(tmem), "r"(64u) : "memory");
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                  CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int main(int argc, char** argv) {
    const int variant = argc > 1 ? atoi(argv[1]) : 0;
    CK(cudaFree(0));
    void* f = nullptr;
    cudaDriverEntryPointQueryResult qr;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &qr) != cudaSuccess || qr != cudaDriverEntryPointSuccess) { printf("no cuTensorMapEncodeTiled\n"); return 1; }
    EncodeTiledFn enc = reinterpret_cast<EncodeTiledFn>(f);
    const size_t nx = (size_t)NB * H * W * C, nw = (size_t)K * F, no = (size_t)NB * HO * WO * F;
    std::vector<__half> hx(nx), hw(nw);
    uint32_t s = 777u;
    for (auto& v : hx) { s = s * 1664525u + 1013904223u; v = __float2half((float)((s >> 16) & 255) / 128.f - 1.f); }
    for (auto& v : hw) { s = s * 1664525u + 1013904223u; v = __float2half((float)((s >> 16) & 255) / 2048.f - 0.0625f); }
    __half *dx, *dw, *dout;
    CK(cudaMalloc(&dx, nx * 2)); CK(cudaMalloc(&dw, nw * 2)); CK(cudaMalloc(&dout, no * 2));
    CK(cudaMemcpy(dx, hx.data(), nx * 2, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(dw, hw.data(), nw * 2, cudaMemcpyHostToDevice));
    CK(cudaMemset(dout, 0, no * 2));
    CUtensorMap tm;
    const cuuint64_t dims[2] = {C, (cuuint64_t)NB * H * W};
    const cuuint64_t strides[1] = {(cuuint64_t)C * 2};
    const cuuint32_t box[2] = {C, SLAB_BOX};
    const cuuint32_t estr[2] = {1, 1};
    if (enc(&tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, dx, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_64B,
            CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS) { printf("tensor map encode failed\n"); return 1; }
    const int smem = (WB_BYTES + 1023) / 1024 * 1024 + NSLAB * SLAB_STRIDE + 256;
    CK(cudaFuncSetAttribute(k_conv_slab, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    int sms = 0;
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    const int tiles = NB * TPS, grid = 2 * sms;
    k_conv_slab<<<grid, THREADS, smem>>>(tm, dw, dout, tiles, variant);
    CK(cudaDeviceSynchronize());
    // check the first three samples against a host convolution (fp32 accumulation, rounded once)
    std::vector<__half> ho_((size_t)3 * HO * WO * F);
    CK(cudaMemcpy(ho_.data(), dout, ho_.size() * 2, cudaMemcpyDeviceToHost));
    long long bad = 0; double maxerr = 0;
    for (int b = 0; b < 3; ++b)
        for (int y = 0; y < HO; ++y)
            for (int x = 0; x < WO; ++x)
                for (int fo = 0; fo < F; ++fo) {
                    float acc = 0.f;
                    for (int ki = 0; ki < KH; ++ki)
                        for (int kj = 0; kj < KW; ++kj)
                            for (int c = 0; c < C; ++c)
                                acc += __half2float(hx[(((size_t)b * H + y + ki) * W + x + kj) * C + c]) * __half2float(hw[(size_t)((ki * KW + kj) * C + c) * F + fo]);
                    const float got = __half2float(ho_[(((size_t)b * HO + y) * WO + x) * F + fo]);
                    const double err = fabs((double)got - (double)acc);
                    if (err > maxerr) maxerr = err;
                    if (err > 2e-2 + 4e-3 * fabs(acc)) ++bad;
                }
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    for (int i = 0; i < 3; ++i) k_conv_slab<<<grid, THREADS, smem>>>(tm, dw, dout, tiles, variant);
    CK(cudaEventRecord(e0));
    const int reps = 20;
    for (int i = 0; i < reps; ++i) k_conv_slab<<<grid, THREADS, smem>>>(tm, dw, dout, tiles, variant);
    CK(cudaEventRecord(e1));
    CK(cudaEventSynchronize(e1));
    float ms = 0.f;
    CK(cudaEventElapsedTime(&ms, e0, e1));
    ms /= reps;
    const double flop = 2.0 * NB * HO * WO * (double)K * F;
    printf("variant %d: conv 3x3 32->32 over [750][26][26][32]: %.3f ms, %.1f TFLOP/s algorithmic (%.2f GFLOP), %.1f GB/s in+out, output check %s (%lld wrong of %zu, max abs err %.4f)\n",
           variant, ms, flop / ms / 1e9, flop / 1e9, (nx * 2.0 + no * 2.0) / ms / 1e6, bad ? "MISMATCH" : "ok", bad, ho_.size(), maxerr);
    return 0;
}
