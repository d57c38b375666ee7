// Forward pass of stride-1 convolutions on tcgen05 WITHOUT re-fetching every input pixel kh * kw times.
//
// The general GEMM kernel (srn_gemm_tc.cuh) fills shared memory with the im2col operand tap by tap: 9 - 49 times the
// input bytes go through the LSU, and that, not HBM or the tensor pipe, bounds it.  Here a tile is
// 128 consecutive positions p = ho * W + wo of one sample's INPUT grid.  The tile's input rows
// [p0, p0 + 128 + (kh - 1) W + kw - 1) are loaded ONCE by TMA as swizzled K-major rows of C channels (the "slab"), and
// the A operand of tap (ki, kj) is the same slab read through a shared-memory matrix descriptor whose start address is
// advanced by (ki W + kj) rows.  tcgen05 applies the swizzle XOR to absolute shared-memory address bits, so a start
// address that is not a multiple of the 8-row swizzle atom reads the right chunks (tools/micro/conv_slab.cu pins this:
// 424 TFLOP/s on a 3 x 3, 32 -> 32 channel layer that the gather runs at ~100).  Positions with wo >= Wo are computed
// and dropped (Wo / W of the MMA work is useful).  The weights [K = kh kw C][ldw] of the current problem stay RESIDENT in
// shared memory (64-row TMA boxes, MN-major swizzle-128B, the layout of the GEMM kernel's forward B operand) and are
// replaced when the CTA's tile walk reaches the next problem.  (Streaming them per tile through a ring lost to the gather:
// 64 requests of 32-64 bytes per chunk with six chunks in flight is bound by the copy engine's request latency.)
//
//   warp 0     producer: one thread issues the slab of tile i + 2 and, at a problem change, the weight boxes
//   warp 1     MMA issuer (converged warp, elect.sync): K / 16 tcgen05.mma per tile into one of NACC accumulators
//   warps 2-5  epilogue: round, + bias, round, activation - the arithmetic of the GEMM kernel's forward epilogue
//
// Eligible: Conv1D / Conv2D with stride 1, C in {16, 32, 64} (one swizzle row per pixel), F <= 64, dense input tensor.
// Everything else, and the whole backward pass, stays on srn_gemm_tc.cuh.  All tensor maps cover this GPU's memory.
#pragma once
#include "srn_gemm_tc.cuh"

namespace srn {

constexpr int CS_THREADS = 192;
constexpr int CS_NSLAB = 4;           // slab ring (prefetch distance 2)
constexpr int CS_WCHUNK = 64 * 128;   // one weight box: 64 k rows x 64 n (128-byte rows)
constexpr int CS_KMAX = 832;          // resident weights: at most 13 boxes = 104 KB
constexpr int CS_NACC = 4;

struct __align__(64) SlabProb {
    CUtensorMap amap;                 // [max_batch * H * W][C] fp16, box {C, box_rows}, swizzle = C * 2 bytes
    CUtensorMap wmap;                 // [K][ldw] fp16, box {64, 64}, swizzle-128B
    const __half* bias;
    __half* out;
    int H, W, Wo, P, tps;             // P = (Ho - 1) W + Wo positions, tps = tiles per sample
    int C, F, Fs, K, kw;
    int out_bs, act;
    int rowb, layout, sbo;            // bytes per slab row, descriptor layout type and 8-row stride
    int box_rows, nbox, slab_bytes;   // slab = nbox TMA boxes of box_rows rows
    int pad[5];
};
static_assert(sizeof(SlabProb) % 64 == 0, "tensor maps inside the table stay 64-byte aligned");

__host__ __device__ constexpr int cs_smem_bytes(int slab_stride, int wbytes) { return wbytes + CS_NSLAB * slab_stride + 512; }

__device__ __forceinline__ uint64_t cs_desc(uint32_t saddr, uint32_t lbo, uint32_t sbo, uint32_t layout) {
    return (uint64_t)((saddr & 0x3FFFFu) >> 4) | ((uint64_t)(lbo >> 4) << 16) | ((uint64_t)(sbo >> 4) << 32) | (1ull << 46) | ((uint64_t)layout << 61);
}

template <int BN>
static __global__ void __launch_bounds__(CS_THREADS, 1)
k_conv_slab(const SlabProb* __restrict__ probs, const int* __restrict__ prefix, int nprob, int ntiles, int slab_stride, int wbytes) {
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    const uint32_t sbase = smem_u32(smem_raw);
    if (sbase & 1023u) __trap();
    const uint32_t sWt = sbase, sSlab = sbase + (uint32_t)wbytes;          // wbytes is a multiple of 8192
    const uint32_t sBar = sSlab + CS_NSLAB * (uint32_t)slab_stride;
    const uint32_t sSFull = sBar, sSEmpty = sSFull + 8 * CS_NSLAB, sWFull = sSEmpty + 8 * CS_NSLAB, sWFree = sWFull + 8;
    const uint32_t sTFull = sWFree + 8, sTEmpty = sTFull + 8 * CS_NACC, sSlot = sTEmpty + 8 * CS_NACC;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    if (tid == 0) {
        for (int i = 0; i < CS_NSLAB; ++i) { mbar_init(sSFull + 8 * i, 1); mbar_init(sSEmpty + 8 * i, 1); }
        mbar_init(sWFull, 1); mbar_init(sWFree, 1);
        for (int i = 0; i < CS_NACC; ++i) { mbar_init(sTFull + 8 * i, 1); mbar_init(sTEmpty + 8 * i, 4); }
        asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
    }
    if (warp == 1) tmem_alloc(sSlot, (uint32_t)(CS_NACC * BN));
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *reinterpret_cast<volatile uint32_t*>(smem_raw + (sSlot - sbase));
    const int G = gridDim.x;
    const int my_tiles = ((int)blockIdx.x < ntiles) ? (ntiles - 1 - (int)blockIdx.x) / G + 1 : 0;
    // problem of tile t: the prefix table is small (a bucket's problems), tiles of a CTA only move forward
    int pi = 0;
    auto prob_of = [&](int t) {
        while (pi + 1 < nprob && t >= prefix[pi + 1]) ++pi;
        return pi;
    };

    if (warp == 0) {
        if (lane == 0) {
            int ss = 0; uint32_t sph = 1;
            int pj = 0;                                   // problem cursor of the slab prefetch (runs ahead of pi)
            auto issue_slab = [&](int i) {
                if (i >= my_tiles) return;
                const int t = (int)blockIdx.x + i * G;
                while (pj + 1 < nprob && t >= prefix[pj + 1]) ++pj;
                const SlabProb& q = probs[pj];
                const int tl = t - prefix[pj], b = tl / q.tps, p0 = (tl - b * q.tps) * 128;
                mbar_wait(sSEmpty + 8 * ss, sph);
                mbar_arrive_expect_tx(sSFull + 8 * ss, (uint32_t)q.slab_bytes);
                const int row0 = b * q.H * q.W + p0;
                const uint32_t dst = sSlab + (uint32_t)ss * (uint32_t)slab_stride;
                tma_load_2d(dst, &q.amap, 0, row0, sSFull + 8 * ss);
                if (q.nbox > 1) tma_load_2d(dst + (uint32_t)(q.box_rows * q.rowb), &q.amap, 0, row0 + q.box_rows, sSFull + 8 * ss);
                if (++ss == CS_NSLAB) { ss = 0; sph ^= 1u; }
            };
            issue_slab(0);
            issue_slab(1);
            int prev = -1, wloads = 0;
            for (int i = 0; i < my_tiles; ++i) {
                const int t = (int)blockIdx.x + i * G;
                const int pr = prob_of(t);
                if (pr != prev) {
                    // next problem: its weights replace the resident ones once the issuer is done with them
                    const SlabProb& q = probs[pr];
                    const int nch = (q.K + 63) >> 6;
                    if (wloads) mbar_wait(sWFree, (uint32_t)((wloads - 1) & 1));
                    mbar_arrive_expect_tx(sWFull, (uint32_t)(nch * CS_WCHUNK));
                    for (int ch = 0; ch < nch; ++ch) tma_load_2d(sWt + (uint32_t)ch * CS_WCHUNK, &q.wmap, 0, ch * 64, sWFull);
                    ++wloads;
                    prev = pr;
                }
                issue_slab(i + 2);
            }
        }
    } else if (warp == 1) {
        int ss = 0; uint32_t sph = 0;
        auto bcast = [](int v) { return __shfl_sync(0xffffffffu, v, 0); };
        int prev = -1, wloads = 0;
        int* offtab = reinterpret_cast<int*>(smem_raw + (sBar - sbase) + 256);      // 64 entries behind the barriers
        for (int i = 0; i < my_tiles; ++i) {
            const int t = (int)blockIdx.x + i * G;
            const int pr = prob_of(t);
            const SlabProb& q = probs[pr];
            const int K16 = bcast(q.K) >> 4, C16 = bcast(q.C) >> 4, W = bcast(q.W), kw = bcast(q.kw);
            const uint32_t rowb = (uint32_t)bcast(q.rowb), layout = (uint32_t)bcast(q.layout), sbo = (uint32_t)bcast(q.sbo);
            const int n_umma = min(BN, (bcast(q.F) + 15) & ~15);
            // kind::f16, D f32, A K-major, B MN-major
            const uint32_t idesc = (1u << 4) | (1u << 16) | ((uint32_t)(n_umma >> 3) << 17) | ((uint32_t)(TC_BM >> 4) << 24);
            if (pr != prev) {
                mbar_wait(sWFull, (uint32_t)(wloads & 1)); ++wloads; prev = pr;
                // slab offset (16-byte units) of every K = 16 step of this problem: tap (ki, kj) = rows ki W + kj further on,
                // 16-channel step s = 32 bytes further inside the row.  One table per problem keeps the issue loop at a few
                // instructions per MMA: with N <= 64 an MMA is 16-32 tensor cycles and this one warp has to keep up.
                __syncwarp();
                for (int idx = lane; idx < ((K16 + 3) & ~3); idx += 32) {
                    const int id = min(idx, K16 - 1), tap = id / C16, sst = id - tap * C16, ki = tap / kw, kjj = tap - ki * kw;
                    offtab[idx] = (int)((((uint32_t)(ki * W + kjj)) * rowb + (uint32_t)sst * 32u) >> 4);
                }
                __syncwarp();
            }
            // is this the last tile of the problem on this CTA?  then the weights are handed back with its last MMA
            bool last_of_prob = i + 1 >= my_tiles;
            if (!last_of_prob) { const int tn = t + G; last_of_prob = pr + 1 < nprob && tn >= prefix[pr + 1]; }
            const int acc = i % CS_NACC;
            if (i >= CS_NACC) mbar_wait(sTEmpty + 8 * acc, (uint32_t)((i / CS_NACC - 1) & 1));
            mbar_wait(sSFull + 8 * ss, sph);
            tc_fence_after();
            const uint32_t tacc = tmem + (uint32_t)(acc * BN);
            const uint64_t dA0 = cs_desc(sSlab + (uint32_t)ss * (uint32_t)slab_stride, 16, sbo, layout);
            const uint64_t dB0 = cs_desc(sWt, 8192, 1024, 2);
            for (int k = 0; k < K16; k += 4) {
                const int4 o = *reinterpret_cast<const int4*>(offtab + k);
                const int n4 = K16 - k;
                if (elect_one_sync()) {
                    umma_f16(tacc, dA0 + (uint64_t)(uint32_t)o.x, dB0 + (uint64_t)(uint32_t)(k * 128), idesc, k ? 1u : 0u);
                    if (n4 > 1) umma_f16(tacc, dA0 + (uint64_t)(uint32_t)o.y, dB0 + (uint64_t)(uint32_t)(k * 128 + 128), idesc, 1u);
                    if (n4 > 2) umma_f16(tacc, dA0 + (uint64_t)(uint32_t)o.z, dB0 + (uint64_t)(uint32_t)(k * 128 + 256), idesc, 1u);
                    if (n4 > 3) umma_f16(tacc, dA0 + (uint64_t)(uint32_t)o.w, dB0 + (uint64_t)(uint32_t)(k * 128 + 384), idesc, 1u);
                    if (n4 <= 4) {
                        umma_commit(sSEmpty + 8 * ss);
                        umma_commit(sTFull + 8 * acc);
                        if (last_of_prob) umma_commit(sWFree);
                    }
                }
                __syncwarp();
            }
            if (++ss == CS_NSLAB) { ss = 0; sph ^= 1u; }
        }
    } else {
        const int qw = warp & 3;                       // TMEM lanes 32 qw .. 32 qw + 31
        for (int i = 0; i < my_tiles; ++i) {
            const int t = (int)blockIdx.x + i * G;
            const SlabProb& q = probs[prob_of(t)];
            const int acc = i % CS_NACC;
            const int tl = t - prefix[pi], b = tl / q.tps, p = (tl - b * q.tps) * 128 + qw * 32 + lane;
            const int W = q.W, F = q.F, Fs = q.Fs, act = q.act;
            const int ho = p / W, wo = p - ho * W;
            __half* orow = (p < q.P && wo < q.Wo) ? q.out + (long long)b * q.out_bs + (long long)(ho * q.Wo + wo) * Fs : nullptr;
            const __half* bias = q.bias;
            const int n_umma = min(BN, (F + 15) & ~15);
            mbar_wait(sTFull + 8 * acc, (uint32_t)((i / CS_NACC) & 1));
            tc_fence_after();
            const uint32_t trow = tmem + (uint32_t)(acc * BN) + ((uint32_t)(qw * 32) << 16);
#pragma unroll
            for (int cb = 0; cb < BN; cb += 32) {
                if (cb >= n_umma) break;
                uint32_t r[32];
                if (n_umma - cb > 16) tmem_ld32(trow + cb, r);
                else tmem_ld16_raw(trow + cb, r);
                if (cb + 32 >= n_umma) {               // all reads of this accumulator are done: hand it back before the stores
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(sTEmpty + 8 * acc);
                }
                if (!orow) continue;
                const int left = F - cb;               // real columns from here on
                const int nch = min(4, (left + 7) >> 3), nfull = min(4, left >> 3);
                __half* dst = orow + cb;
#pragma unroll
                for (int h8 = 0; h8 < 4; ++h8) {
                    if (h8 >= nch) break;
                    const uint4 bq = ldg16(bias + cb + h8 * 8);          // bias arrays are padded to 8 elements in the arena
                    uint4 o;
                    if (act != 2 && h8 < nfull) {
                        const __half2* b2 = reinterpret_cast<const __half2*>(&bq);
                        __half2* o2 = reinterpret_cast<__half2*>(&o);
                        const __half2 zero = __float2half2_rn(0.f);
#pragma unroll
                        for (int j = 0; j < 4; ++j) {
                            __half2 z = __hadd2(pack_h2(r[h8 * 8 + 2 * j], r[h8 * 8 + 2 * j + 1]), b2[j]);
                            if (act == 1) z = __hmax2_nan(z, zero);
                            o2[j] = z;
                        }
                    } else {
                        const Half8 bv = *reinterpret_cast<const Half8*>(&bq);
                        Half8 oh;
#pragma unroll
                        for (int j = 0; j < 8; ++j) {
                            float z = 0.f;
                            if (cb + h8 * 8 + j < F) z = act_forward(rh(rh(__uint_as_float(r[h8 * 8 + j])) + __half2float(bv.v[j])), act);
                            oh.v[j] = __float2half_rn(z);
                        }
                        o = *reinterpret_cast<const uint4*>(&oh);
                    }
                    stg16(dst + h8 * 8, o);
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) tmem_dealloc(tmem, (uint32_t)(CS_NACC * BN));
}

}  // namespace srn
