// Population-grouped implicit-GEMM kernels on the CUDA cores (fp16 operands, fp32 accumulate).
//
// These are the general kernels: any channel count, any stride, unaligned shapes, K < 16 first layers.
// Every Dense / Conv1D / Conv2D of every individual maps onto three GEMMs:
//   forward   Out[m][f]  = sum_k A[m][k] W[k][f]      m = (b, ho, wo), k = (i, j, c)
//   dgrad     dIn[m][c]  = sum_k dOut'[m][k] W[k'][c] m = (b, h, w),   k = (i, j, f)
//   wgrad     dW[k][f]  += sum_m A[m][k] dOut[m][f]   reduction split across CTAs, fp32 atomics
// A is never materialised: tiles are gathered straight from the NHWC activation.
#pragma once
#include "srn_common.cuh"

namespace srn {

constexpr int SIMT_BK = 16;
constexpr int SIMT_THREADS = 256;
constexpr int WGRAD_ROWS_PER_SPLIT = 2048;

template <int BM, int BN>
struct SimtTile {
    static constexpr int TM = 4, TN = 4;
    static constexpr int TX = BN / TN;            // threads along n
    static constexpr int TY = BM / TM;            // threads along m
    static_assert(TX * TY == SIMT_THREADS, "tile shape must use 256 threads");
};

template <int BM, int BN>
__device__ __forceinline__ void simt_mma(const float (&As)[SIMT_BK][BM + 4], const float (&Bs)[SIMT_BK][BN + 4],
                                         float (&acc)[4][4], int ty, int tx) {
#pragma unroll
    for (int k = 0; k < SIMT_BK; ++k) {
        float4 a = *reinterpret_cast<const float4*>(&As[k][ty * 4]);
        float4 b = *reinterpret_cast<const float4*>(&Bs[k][tx * 4]);
        float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
    }
}

__host__ __device__ inline int fwd_tiles(int M, int N, int BM, int BN) {
    return ((M + BM - 1) / BM) * ((N + BN - 1) / BN);
}

// ------------------------------------------------------------------ forward
template <int BM, int BN>
static __global__ void __launch_bounds__(SIMT_THREADS)
k_conv_fwd(const Prob* __restrict__ probs, const int* __restrict__ prefix, int nprob, int nb, Inputs I) {
    using T = SimtTile<BM, BN>;
    __shared__ __align__(16) float As[SIMT_BK][BM + 4];
    __shared__ __align__(16) float Bs[SIMT_BK][BN + 4];
    const int pi = find_prob(prefix, nprob, blockIdx.x);
    const Prob p = probs[pi];
    const int tile = blockIdx.x - prefix[pi];
    const int HoWo = p.Ho * p.Wo;
    const int M = nb * HoWo, N = p.F, K = p.kh * p.kw * p.C;
    const int tilesN = (N + BN - 1) / BN;
    const int m0 = (tile / tilesN) * BM, n0 = (tile % tilesN) * BN;
    int in_bs;
    const __half* in = resolve_in(p, I, in_bs);
    const int tid = threadIdx.x;

    // A loader: lane along k (contiguous (j, c) runs), rows tid/16 + 16 r
    constexpr int AR = BM / 16;
    const int ak = tid % SIMT_BK;
    long long arow_off[AR];
#pragma unroll
    for (int r = 0; r < AR; ++r) {
        int m = m0 + tid / SIMT_BK + 16 * r;
        if (m < M) {
            int b = m / HoWo, rem = m - b * HoWo;
            int ho = rem / p.Wo, wo = rem - ho * p.Wo;
            arow_off[r] = (long long)b * in_bs + ((long long)(ho * p.sh) * p.W + wo * p.sw) * p.Cs;
        } else {
            arow_off[r] = -1;
        }
    }
    // B loader: lane along n
    constexpr int BR = SIMT_BK * BN / SIMT_THREADS;        // k rows per thread
    constexpr int BKSTEP = SIMT_THREADS / BN;
    const int bn = tid % BN, bk0 = tid / BN;

    const int ty = tid / T::TX, tx = tid % T::TX;
    float acc[4][4] = {};
    const int kwC = p.kw * p.C;

    for (int k0 = 0; k0 < K; k0 += SIMT_BK) {
        {
            int k = k0 + ak;
            int koff = -1;
            if (k < K) {                      // k = (i, j, c) over real channels; memory steps by the channel stride
                int i = k / kwC, rem = k - i * kwC;
                int j = rem / p.C, c = rem - j * p.C;
                koff = (i * p.W + j) * p.Cs + c;
            }
#pragma unroll
            for (int r = 0; r < AR; ++r) {
                float v = 0.f;
                if (koff >= 0 && arow_off[r] >= 0) v = __half2float(in[arow_off[r] + koff]);
                As[ak][tid / SIMT_BK + 16 * r] = v;
            }
        }
#pragma unroll
        for (int r = 0; r < BR; ++r) {
            int k = k0 + bk0 + BKSTEP * r, n = n0 + bn;
            float v = 0.f;
            if (k < K && n < N) v = __half2float(p.w[(long long)k * p.ldw + n]);
            Bs[bk0 + BKSTEP * r][bn] = v;
        }
        __syncthreads();
        simt_mma<BM, BN>(As, Bs, acc, ty, tx);
        __syncthreads();
    }
    // epilogue: conv result -> fp16, + bias -> fp16, activation -> fp16   (three TF ops)
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        int m = m0 + ty * 4 + i;
        if (m >= M) continue;
        int b = m / HoWo, rem = m - b * HoWo;
        __half* orow = p.out + (long long)b * p.out_bs + (long long)rem * p.Fs;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            int n = n0 + tx * 4 + j;
            if (n >= N) continue;
            float z = rh(rh(acc[i][j]) + __half2float(p.bias[n]));
            orow[n] = __float2half_rn(act_forward(z, p.act));
        }
    }
}

// ------------------------------------------------------------------ dgrad
// M = nb*H*W input positions, N = C input channels, K = kh*kw*F.
template <int BM, int BN>
static __global__ void __launch_bounds__(SIMT_THREADS)
k_conv_dgrad(const Prob* __restrict__ probs, const int* __restrict__ prefix, int nprob, int nb) {
    using T = SimtTile<BM, BN>;
    __shared__ __align__(16) float As[SIMT_BK][BM + 4];
    __shared__ __align__(16) float Bs[SIMT_BK][BN + 4];
    const int pi = find_prob(prefix, nprob, blockIdx.x);
    const Prob p = probs[pi];
    const int tile = blockIdx.x - prefix[pi];
    const int HW = p.H * p.W, HoWo = p.Ho * p.Wo;
    const int M = nb * HW, N = p.C, F = p.F, K = p.kh * p.kw * F;
    const int tilesN = (N + BN - 1) / BN;
    const int m0 = (tile / tilesN) * BM, n0 = (tile % tilesN) * BN;
    const int tid = threadIdx.x;

    constexpr int AR = BM / 16;
    const int ak = tid % SIMT_BK;
    int ab[AR], ah[AR], aw[AR];
#pragma unroll
    for (int r = 0; r < AR; ++r) {
        int m = m0 + tid / SIMT_BK + 16 * r;
        if (m < M) {
            ab[r] = m / HW;
            int rem = m - ab[r] * HW;
            ah[r] = rem / p.W;
            aw[r] = rem - ah[r] * p.W;
        } else {
            ab[r] = -1; ah[r] = 0; aw[r] = 0;
        }
    }
    constexpr int BR = SIMT_BK * BN / SIMT_THREADS;
    constexpr int BKSTEP = SIMT_THREADS / BN;
    const int bn = tid % BN, bk0 = tid / BN;
    const int ty = tid / T::TX, tx = tid % T::TX;
    float acc[4][4] = {};

    for (int k0 = 0; k0 < K; k0 += SIMT_BK) {
        {
            int k = k0 + ak;
            int ij = k / F, f = k - ij * F;
            int i = ij / p.kw, j = ij - i * p.kw;
#pragma unroll
            for (int r = 0; r < AR; ++r) {
                float v = 0.f;
                if (k < K && ab[r] >= 0) {
                    int hh = ah[r] - i, ww = aw[r] - j;
                    if (hh >= 0 && ww >= 0) {
                        int ho = hh / p.sh, wo = ww / p.sw;
                        if (ho * p.sh == hh && wo * p.sw == ww && ho < p.Ho && wo < p.Wo)
                            v = __half2float(p.gout[(long long)ab[r] * p.out_bs + ((long long)ho * p.Wo + wo) * p.Fs + f]);
                    }
                }
                As[ak][tid / SIMT_BK + 16 * r] = v;
            }
        }
#pragma unroll
        for (int r = 0; r < BR; ++r) {
            int k = k0 + bk0 + BKSTEP * r, c = n0 + bn;
            float v = 0.f;
            if (k < K && c < N) {
                int ij = k / F, f = k - ij * F;
                v = __half2float(p.w[((long long)ij * N + c) * p.ldw + f]);
            }
            Bs[bk0 + BKSTEP * r][bn] = v;
        }
        __syncthreads();
        simt_mma<BM, BN>(As, Bs, acc, ty, tx);
        __syncthreads();
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        int m = m0 + ty * 4 + i;
        if (m >= M) continue;
        int b = m / HW, rem = m - b * HW;
        __half* drow = p.din + (long long)b * p.in_bs + (long long)rem * p.Cs;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            int c = n0 + tx * 4 + j;
            if (c >= N) continue;
            float v = acc[i][j];
            if (p.din_acc) v = rh(v) + __half2float(drow[c]);
            drow[c] = __float2half_rn(v);
        }
    }
}

__host__ __device__ inline int wgrad_tiles(int Kf, int F, int M) {
    return ((Kf + 63) / 64) * ((F + 63) / 64) * ((M + WGRAD_ROWS_PER_SPLIT - 1) / WGRAD_ROWS_PER_SPLIT);
}

// ------------------------------------------------------------------ wgrad
// dW[kk][f] += sum over a slab of rows m of A[m][kk] * dOut[m][f]; 64x64 tile, fp32 atomics into gw.
static __global__ void __launch_bounds__(SIMT_THREADS)
k_conv_wgrad(const Prob* __restrict__ probs, const int* __restrict__ prefix, int nprob, int nb, Inputs I) {
    constexpr int BM = 64, BN = 64;
    using T = SimtTile<BM, BN>;
    __shared__ __align__(16) float As[SIMT_BK][BM + 4];
    __shared__ __align__(16) float Bs[SIMT_BK][BN + 4];
    const int pi = find_prob(prefix, nprob, blockIdx.x);
    const Prob p = probs[pi];
    int tile = blockIdx.x - prefix[pi];
    const int HoWo = p.Ho * p.Wo;
    const int M = nb * HoWo, F = p.F, Kf = p.kh * p.kw * p.C;
    const int tilesK = (Kf + 63) / 64, tilesF = (F + 63) / 64;
    const int split = tile / (tilesK * tilesF);
    tile -= split * tilesK * tilesF;
    const int kk0 = (tile / tilesF) * BM, f0 = (tile % tilesF) * BN;
    const int r0 = split * WGRAD_ROWS_PER_SPLIT;
    const int r1 = min(M, r0 + WGRAD_ROWS_PER_SPLIT);
    int in_bs;
    const __half* in = resolve_in(p, I, in_bs);
    const int tid = threadIdx.x;

    // A' loader: lane along kk (contiguous in memory), 4 rows of m per thread per slab
    const int akk = kk0 + (tid % 64);
    long long akoff = -1;
    if (akk < Kf) {
        int kwC = p.kw * p.C;
        int i = akk / kwC, rem = akk - i * kwC;
        int j = rem / p.C, c = rem - j * p.C;
        akoff = ((long long)i * p.W + j) * p.Cs + c;
    }
    const int bf = f0 + (tid % 64);
    const int mr0 = tid / 64;                 // 0..3
    const int ty = tid / T::TX, tx = tid % T::TX;
    float acc[4][4] = {};

    for (int ms = r0; ms < r1; ms += SIMT_BK) {
#pragma unroll
        for (int r = 0; r < 4; ++r) {
            int ml = mr0 + 4 * r;
            int m = ms + ml;
            float av = 0.f, bv = 0.f;
            if (m < r1) {
                int b = m / HoWo, rem = m - b * HoWo;
                if (akoff >= 0) {
                    int ho = rem / p.Wo, wo = rem - ho * p.Wo;
                    av = __half2float(in[(long long)b * in_bs + ((long long)(ho * p.sh) * p.W + wo * p.sw) * p.Cs + akoff]);
                }
                if (bf < F) bv = __half2float(p.gout[(long long)b * p.out_bs + (long long)rem * p.Fs + bf]);
            }
            As[ml][tid % 64] = av;
            Bs[ml][tid % 64] = bv;
        }
        __syncthreads();
        simt_mma<BM, BN>(As, Bs, acc, ty, tx);
        __syncthreads();
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        int kk = kk0 + ty * 4 + i;
        if (kk >= Kf) continue;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            int f = f0 + tx * 4 + j;
            if (f < F && acc[i][j] != 0.f) atomicAdd(&p.gw[(long long)kk * p.ldw + f], acc[i][j]);
        }
    }
}

}  // namespace srn
