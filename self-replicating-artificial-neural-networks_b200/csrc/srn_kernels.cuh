// HBM-bound kernels of the SeRANN engine: activation backward + bias gradient, BatchNormalizationF16,
// max pooling, copy (concatenate / reshape), fused losses, multi-tensor Adam, batch gather, readout, init.
// All are population-grouped (see Prob in srn_common.cuh) except Adam/readout, which run flat.
// Tensors are [row][Cs] with Cs a multiple of 8 (srn_common.cuh), so the elementwise kernels move 16-byte
// chunks of 8 channels; the scalar branch (Cs = 1) only serves layers applied directly to a model input.
#pragma once
#include "srn_common.cuh"

namespace srn {

// Packed-half arithmetic for the op-by-op fp16 parts: the product or sum of two fp16 values is exact in fp32, so
// rh(float(a) * float(b)) == __hmul(a, b) bit for bit, and likewise for add / sub.  Eight elements cost four packed
// instructions instead of eight convert-operate-convert chains; these passes were bound by instruction issue, not HBM.
struct __align__(16) Half2x4 { __half2 v[4]; };
__device__ __forceinline__ Half2x4 as_h2(const Half8& h) { return *reinterpret_cast<const Half2x4*>(&h); }
__device__ __forceinline__ Half8 as_h8(const Half2x4& h) { return *reinterpret_cast<const Half8*>(&h); }
// gradient of an activation, eight elements: ReLU keeps g where y > 0 (anything else, NaN included, gives +0),
// sigmoid is TF's SigmoidGrad dy * y * (1 - y) with one rounding per op
__device__ __forceinline__ Half8 act_grad8(const Half8& g, const Half8& y, int act) {
    Half2x4 g2 = as_h2(g);
    const Half2x4 y2 = as_h2(y);
    if (act == 1) {
        const __half2 zero = __float2half2_rn(0.f);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const unsigned m = __hgt2_mask(y2.v[k], zero);
            unsigned bits = *reinterpret_cast<const unsigned*>(&g2.v[k]) & m;
            g2.v[k] = *reinterpret_cast<__half2*>(&bits);
        }
    } else if (act == 2) {
        const __half2 one = __float2half2_rn(1.f);
#pragma unroll
        for (int k = 0; k < 4; ++k) g2.v[k] = __hmul2_rn(__hmul2_rn(g2.v[k], y2.v[k]), __hsub2_rn(one, y2.v[k]));      // _rn: never contracted into an fma
    }
    return as_h8(g2);
}

constexpr int COL_THREADS = 256;
constexpr int COL_ITERS = 16;         // rows per row lane and CTA in the column-reduction kernels

// Thread layout of the column reductions: channel chunk cq = tid % nq, row lane rl = tid / nq.
struct ColGeom { int cw, nq, RL, rows_per_cta; };
__host__ __device__ inline ColGeom col_geom(int Cs) {
    ColGeom g;
    g.cw = (Cs % 8 == 0) ? 8 : 1;
    g.nq = Cs / g.cw;
    g.RL = g.nq <= COL_THREADS ? COL_THREADS / g.nq : 1;
    g.rows_per_cta = g.RL * COL_ITERS;
    return g;
}
__host__ __device__ inline int col_chunks(long long rows, int Cs) {
    const int rpc = col_geom(Cs).rows_per_cta;
    return (int)((rows + rpc - 1) / rpc);
}

// offset of row r (over all samples) of an input tensor; model inputs may be strided or broadcast per sample
__device__ __forceinline__ long long in_row_off(long long r, int rps, int bs, int Cs) {
    if (bs == rps * Cs) return r * Cs;                       // samples back to back: no division
    const unsigned b = (unsigned)r / (unsigned)rps;          // rows of a batch fit 32 bits
    return (long long)b * bs + (long long)((unsigned)r - b * (unsigned)rps) * Cs;
}

// offset of row r (over all samples) of an output / output-gradient tensor.  Normally samples are contiguous
// (bs == rps * Fs); a tensor written in place into a concatenation buffer has the buffer's batch stride instead
__device__ __forceinline__ long long out_row_off(long long r, int rps, int bs, int Fs) {
    if (bs == rps * Fs) return r * Fs;
    const unsigned b = (unsigned)r / (unsigned)rps;
    return (long long)b * bs + (long long)((unsigned)r - b * (unsigned)rps) * Fs;
}

// sums over the row lanes of a CTA: sm holds one value per (thread, j); thread `ch` (< nq * cw) gets channel ch
__device__ __forceinline__ float col_block_sum(const float* sm, const ColGeom& g, int ch) {
    const int cq = ch / g.cw, j = ch - cq * g.cw;
    float s = 0.f;
    for (int rl = 0; rl < g.RL; ++rl) s += sm[(rl * g.nq + cq) * 8 + j];
    return s;
}

// ------------------------------------------------------------------ activation backward + bias gradient
// gout <- fp16(gout * act'(out)) in place (TF's ReluGrad / SigmoidGrad); gb[f] += column sums of the result.
// K1: Dense (or 1 x 1 convolution) on a one-channel model input - the kernel is a single row, its gradient
// gw[f] = sum_r x[r] * dZ[r][f] is one more column sum of the same pass (no GEMM launch for these layers).
template <bool K1>
static __global__ void __launch_bounds__(COL_THREADS, 4)
k_act_bwd_bias(const Prob* __restrict__ probs, const int* __restrict__ prefix, int nprob, int nb, Inputs I) {
    __shared__ float sm[COL_THREADS * 8];
    const int pi = find_prob(prefix, nprob, blockIdx.x);
    const Prob p = probs[pi];
    // first layers over the shared im2col matrix get their bias gradient from the weight-gradient GEMM (Cj == 1 marks
    // the ones column): a linear layer has nothing to do here, an activation still needs its mask applied
    const bool bias_here = K1 || p.Cj != 1;
    if (!bias_here && p.act == 0) return;
    const int chunk = blockIdx.x - prefix[pi];
    const int F = p.F, Fs = p.Fs;
    const ColGeom g = col_geom(Fs);
    const long long rows = (long long)nb * p.Ho * p.Wo;
    const long long r0 = (long long)chunk * g.rows_per_cta, r1 = min(rows, r0 + g.rows_per_cta);
    const int tid = threadIdx.x;
    for (int cq0 = 0; cq0 < g.nq; cq0 += COL_THREADS) {
        const int cq = g.nq <= COL_THREADS ? tid % g.nq : cq0 + tid;
        const int rl = g.nq <= COL_THREADS ? tid / g.nq : 0;
        const bool active = rl < g.RL && cq < g.nq;
        float s[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        float sx[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        int in_bs = 0;
        const __half* xin = K1 ? resolve_in(p, I, in_bs) : nullptr;
        if (active) {
            long long r = r0 + rl;
            if (g.cw == 8) {
                // four rows per trip: all loads of a trip are issued before the first dependent instruction, so a thread
                // keeps 64 - 128 bytes in flight instead of 16 - 32 (the pass is latency-bound otherwise: 1.6 TB/s)
                for (; r + 3LL * g.RL < r1; r += 4LL * g.RL) {
                    long long o[4]; Half8 gv[4], y[4]; float xr[4];
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
                        const long long rr = r + (long long)u * g.RL;
                        o[u] = out_row_off(rr, p.Ho * p.Wo, p.out_bs, Fs) + (long long)cq * 8;
                        gv[u] = ld8(p.gout + o[u]);
                        if (p.act) y[u] = ld8(p.out + o[u]);
                        xr[u] = K1 ? __half2float(xin[in_row_off(rr, p.H * p.W, in_bs, 1)]) : 0.f;
                    }
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
                        if (p.act) {
                            gv[u] = act_grad8(gv[u], y[u], p.act);
                            st8(p.gout + o[u], gv[u]);
                        }
                        const Half2x4 g2 = as_h2(gv[u]);
#pragma unroll
                        for (int k = 0; k < 4; ++k) {
                            const float2 f = __half22float2(g2.v[k]);
                            s[2 * k] += f.x; s[2 * k + 1] += f.y;
                            if (K1) { sx[2 * k] += f.x * xr[u]; sx[2 * k + 1] += f.y * xr[u]; }
                        }
                    }
                }
            }
            for (; r < r1; r += g.RL) {
                const long long o = out_row_off(r, p.Ho * p.Wo, p.out_bs, Fs) + (long long)cq * g.cw;
                const float xr = K1 ? __half2float(xin[in_row_off(r, p.H * p.W, in_bs, 1)]) : 0.f;
                if (g.cw == 8) {
                    Half8 gv = ld8(p.gout + o);
                    if (p.act) {
                        gv = act_grad8(gv, ld8(p.out + o), p.act);
                        st8(p.gout + o, gv);
                    }
#pragma unroll
                    for (int j = 0; j < 8; ++j) { const float gj = __half2float(gv.v[j]); s[j] += gj; if (K1) sx[j] += gj * xr; }
                } else {
                    float gg = __half2float(p.gout[o]);
                    if (p.act == 1) { gg = __half2float(p.out[o]) > 0.f ? gg : 0.f; p.gout[o] = __float2half_rn(gg); }
                    else if (p.act == 2) { const float yy = __half2float(p.out[o]); gg = rh(rh(gg * yy) * rh(1.f - yy)); p.gout[o] = __float2half_rn(gg); }
                    s[0] += gg;
                    if (K1) sx[0] += gg * xr;
                }
            }
        }
        if (g.RL == 1) {
            if (active)
                for (int j = 0; j < g.cw; ++j) {
                    const int f = cq * g.cw + j;
                    if (bias_here && f < F && s[j] != 0.f) atomicAdd(&p.gb[f], s[j]);
                    if (K1 && f < F && sx[j] != 0.f) atomicAdd(&p.gw[f], sx[j]);
                }
        } else {
#pragma unroll
            for (int j = 0; j < 8; ++j) sm[tid * 8 + j] = active ? s[j] : 0.f;
            __syncthreads();
            for (int ch = tid; ch < g.nq * g.cw; ch += COL_THREADS) {
                const float t = col_block_sum(sm, g, ch);
                if (bias_here && ch < F && t != 0.f) atomicAdd(&p.gb[ch], t);
            }
            __syncthreads();
            if (K1) {
#pragma unroll
                for (int j = 0; j < 8; ++j) sm[tid * 8 + j] = active ? sx[j] : 0.f;
                __syncthreads();
                for (int ch = tid; ch < g.nq * g.cw; ch += COL_THREADS) {
                    const float t = col_block_sum(sm, g, ch);
                    if (ch < F && t != 0.f) atomicAdd(&p.gw[ch], t);
                }
                __syncthreads();
            }
        }
    }
}

// ------------------------------------------------------------------ BatchNormalizationF16
// save layout per problem: [0][Cs] mean32 (the fp32 batch mean; mean16 = its fp16 rounding), [1][Cs] inv_raw (rsqrt(var16 + eps), fp16-valued), [2][Cs] inv
// (times gamma), [3][Cs] shift; then partial sums [chunks][Cs][2].
constexpr float BN_EPS16 = 0.0010004043579101562f;       // fp16(1e-3)
constexpr float BN_MOM16 = 0.990234375f;                  // fp16(0.99): tf.cast(momentum, fp16)
constexpr float BN_ONE_MINUS_MOM16 = 0.009765625f;        // 1 - fp16(0.99), exact in fp16
constexpr float BN_ONE_EPS16 = 1.0009765625f;             // fp16(1 + 1e-3)

// forward: shifted sums (x - x0), (x - x0)^2 around the first row; backward: sum dy, sum dy * (x - mean16)
static __device__ void bn_finalize_fwd_dev(const Prob& p, int nb, const Inputs& I);
static __device__ void bn_finalize_bwd_dev(const Prob& p, int nb);
// done != nullptr: the CTA that finishes a problem's last chunk also combines the partials (no separate finalize launch);
// done[problem] counts finished chunks and is left at zero
static __global__ void __launch_bounds__(COL_THREADS, 4)
k_bn_partial(const Prob* __restrict__ probs, const int* __restrict__ prefix, int nprob, int nb, Inputs I, int bwd, int* done = nullptr) {
    __shared__ float sm1[COL_THREADS * 8];
    __shared__ float sm2[COL_THREADS * 8];
    const int pi = find_prob(prefix, nprob, blockIdx.x);
    const Prob p = probs[pi];
    const int chunk = blockIdx.x - prefix[pi];
    const int rps = p.H * p.W, Cs = p.Cs;
    const ColGeom g = col_geom(Cs);
    const long long rows = (long long)nb * rps;
    const long long r0 = (long long)chunk * g.rows_per_cta, r1 = min(rows, r0 + g.rows_per_cta);
    int in_bs;
    const __half* in = resolve_in(p, I, in_bs);
    float* part = p.save + 4 * Cs + (long long)chunk * Cs * 2;
    const int tid = threadIdx.x;
    for (int cq0 = 0; cq0 < g.nq; cq0 += COL_THREADS) {
        const int cq = g.nq <= COL_THREADS ? tid % g.nq : cq0 + tid;
        const int rl = g.nq <= COL_THREADS ? tid / g.nq : 0;
        const bool active = rl < g.RL && cq < g.nq;
        float s1[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f}, s2[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        if (active) {
            float pivot[8];
            for (int j = 0; j < g.cw; ++j) pivot[j] = bwd ? rh(p.save[cq * g.cw + j]) : __half2float(in[cq * g.cw + j]);      // backward: mean16 = fp16(mean32)
            long long r = r0 + rl;
            if (g.cw == 8) {
                // four rows per trip: the loads of a trip are independent, so four (eight backward) are in flight per thread
                for (; r + 3LL * g.RL < r1; r += 4LL * g.RL) {
                    Half8 x[4], dy[4];
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
                        const long long rr = r + (long long)u * g.RL;
                        x[u] = ld8(in + in_row_off(rr, rps, in_bs, Cs) + (long long)cq * 8);
                        if (bwd) dy[u] = ld8(p.gout + out_row_off(rr, rps, p.out_bs, p.Fs) + (long long)cq * 8);
                    }
#pragma unroll
                    for (int u = 0; u < 4; ++u) {
#pragma unroll
                        for (int j = 0; j < 8; ++j) {
                            if (bwd) {
                                const float d = __half2float(dy[u].v[j]);
                                s1[j] += d;
                                s2[j] += d * (__half2float(x[u].v[j]) - pivot[j]);
                            } else {
                                const float xv = __half2float(x[u].v[j]) - pivot[j];
                                s1[j] += xv;
                                s2[j] += xv * xv;
                            }
                        }
                    }
                }
            }
            for (; r < r1; r += g.RL) {
                const long long io = in_row_off(r, rps, in_bs, Cs) + (long long)cq * g.cw;
                if (g.cw == 8) {
                    const Half8 x = ld8(in + io);
                    if (bwd) {
                        const Half8 dy = ld8(p.gout + out_row_off(r, rps, p.out_bs, p.Fs) + (long long)cq * 8);
#pragma unroll
                        for (int j = 0; j < 8; ++j) {
                            const float d = __half2float(dy.v[j]);
                            s1[j] += d;
                            s2[j] += d * (__half2float(x.v[j]) - pivot[j]);
                        }
                    } else {
#pragma unroll
                        for (int j = 0; j < 8; ++j) {
                            const float xv = __half2float(x.v[j]) - pivot[j];
                            s1[j] += xv;
                            s2[j] += xv * xv;
                        }
                    }
                } else {
                    const float xv = __half2float(in[io]) - pivot[0];
                    if (bwd) { const float d = __half2float(p.gout[out_row_off(r, rps, p.out_bs, p.Fs) + cq]); s1[0] += d; s2[0] += d * xv; }
                    else { s1[0] += xv; s2[0] += xv * xv; }
                }
            }
        }
        if (g.RL == 1) {
            if (active)
                for (int j = 0; j < g.cw; ++j) { part[(cq * g.cw + j) * 2] = s1[j]; part[(cq * g.cw + j) * 2 + 1] = s2[j]; }
        } else {
#pragma unroll
            for (int j = 0; j < 8; ++j) { sm1[tid * 8 + j] = active ? s1[j] : 0.f; sm2[tid * 8 + j] = active ? s2[j] : 0.f; }
            __syncthreads();
            for (int ch = tid; ch < g.nq * g.cw; ch += COL_THREADS) {
                part[ch * 2] = col_block_sum(sm1, g, ch);
                part[ch * 2 + 1] = col_block_sum(sm2, g, ch);
            }
            __syncthreads();
        }
    }
    if (done) {
        __shared__ int last;
        __threadfence();
        __syncthreads();
        if (tid == 0) last = atomicAdd(&done[pi], 1) == prefix[pi + 1] - prefix[pi] - 1;
        __syncthreads();
        if (last) {
            __threadfence();
            if (bwd) bn_finalize_bwd_dev(p, nb); else bn_finalize_fwd_dev(p, nb, I);
            if (tid == 0) done[pi] = 0;
        }
    }
}

__device__ __forceinline__ void bn_scale_shift(float mean16, float var16, float gamma, float beta,
                                               float& inv_raw, float& inv, float& shift) {
    inv_raw = rh(1.0f / sqrtf(rh(var16 + BN_EPS16)));
    inv = rh(inv_raw * gamma);
    shift = rh(beta - rh(mean16 * inv));
}

// Sum of the per-chunk partials of channel c in chunk order (double accumulators); eight chunks are loaded at a time
// so that the L2 latencies overlap.
__device__ __forceinline__ void bn_chunk_sums(const float* part, int chunks, int Cs, int c, double& s1, double& s2) {
    s1 = 0.0; s2 = 0.0;
    const float2* pp = reinterpret_cast<const float2*>(part) + c;
    int k = 0;
    for (; k + 8 <= chunks; k += 8) {
        float2 v[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) v[j] = __ldcg(pp + (long long)(k + j) * Cs);
#pragma unroll
        for (int j = 0; j < 8; ++j) { s1 += v[j].x; s2 += v[j].y; }
    }
    for (; k < chunks; ++k) { const float2 v = __ldcg(pp + (long long)k * Cs); s1 += v.x; s2 += v.y; }
}

// one CTA per problem: combine partials, derive scale/shift, update the moving statistics (fp16 op by op)
static __global__ void __launch_bounds__(128)
k_bn_finalize_fwd(const Prob* __restrict__ probs, int nb, Inputs I) {
    const Prob p = probs[blockIdx.x];
    bn_finalize_fwd_dev(p, nb, I);
}
static __device__ void bn_finalize_fwd_dev(const Prob& p, int nb, const Inputs& I) {
    const int C = p.C, Cs = p.Cs;
    const long long rows = (long long)nb * p.H * p.W;
    const int chunks = col_chunks(rows, Cs);
    int in_bs;
    const __half* in = resolve_in(p, I, in_bs);
    const float* part = p.save + 4 * Cs;
    for (int c = threadIdx.x; c < Cs; c += blockDim.x) {
        if (c >= C) {           // pad channels: identity scale of zero so that they stay exactly zero
            p.save[c] = 0.f; p.save[Cs + c] = 0.f; p.save[2 * Cs + c] = 0.f; p.save[3 * Cs + c] = 0.f;
            continue;
        }
        double s1, s2;
        bn_chunk_sums(part, chunks, Cs, c, s1, s2);
        const double n = (double)rows;
        const double dm = s1 / n;
        const float mean32 = (float)((double)__half2float(in[c]) + dm);
        const float var32 = (float)fmax(s2 / n - dm * dm, 0.0);
        const float mean16 = rh(mean32), var16 = rh(var32);
        float inv_raw, inv, shift;
        bn_scale_shift(mean16, var16, __half2float(p.w[c]), __half2float(p.bias[c]), inv_raw, inv, shift);
        p.save[c] = mean32; p.save[Cs + c] = inv_raw; p.save[2 * Cs + c] = inv; p.save[3 * Cs + c] = shift;
        // moving statistics: variance * n16 / (n16 - (1 + eps)) with n cast to fp16 (inf/inf = NaN above 65504)
        const float n16 = rh((float)rows);
        const float bessel = rh(n16 / rh(n16 - BN_ONE_EPS16));
        const float var_unb = rh(var16 * bessel);
        float mm = __half2float(p.mov[c]), mv = __half2float(p.mov[C + c]);
        // keras 2.6 backend.moving_average_update (TF2 branch): x.assign(x * momentum + value * (1 - momentum)), fp16 op by op
        mm = rh(rh(mm * BN_MOM16) + rh(mean16 * BN_ONE_MINUS_MOM16));
        mv = rh(rh(mv * BN_MOM16) + rh(var_unb * BN_ONE_MINUS_MOM16));
        p.mov[c] = __float2half_rn(mm);
        p.mov[C + c] = __float2half_rn(mv);
    }
}

// inference: scale/shift from the moving statistics
static __global__ void __launch_bounds__(128)
k_bn_prepare_infer(const Prob* __restrict__ probs) {
    const Prob p = probs[blockIdx.x];
    const int C = p.C, Cs = p.Cs;
    for (int c = threadIdx.x; c < Cs; c += blockDim.x) {
        float inv_raw, inv = 0.f, shift = 0.f;
        if (c < C)
            bn_scale_shift(__half2float(p.mov[c]), __half2float(p.mov[C + c]), __half2float(p.w[c]),
                           __half2float(p.bias[c]), inv_raw, inv, shift);
        p.save[2 * Cs + c] = inv; p.save[3 * Cs + c] = shift;
    }
}

constexpr int EW_THREADS = 256;
constexpr int EW_PER_THREAD = 4;      // 16-byte chunks (or scalars) per thread
constexpr int EW_PER_CTA = EW_THREADS * EW_PER_THREAD;
// work items of an elementwise pass over `rows` rows of channel stride Cs
__host__ __device__ inline int ew_blocks(long long rows, int Cs) {
    const long long units = rows * (Cs % 8 == 0 ? Cs / 8 : Cs);
    return (int)((units + EW_PER_CTA - 1) / EW_PER_CTA);
}

// y = fp16(fp16(x * inv) + shift)
static __global__ void __launch_bounds__(EW_THREADS)
k_bn_apply(const Prob* __restrict__ probs, const int* __restrict__ prefix, int nprob, int nb, Inputs I) {
    const int pi = find_prob(prefix, nprob, blockIdx.x);
    const Prob p = probs[pi];
    const int blk = blockIdx.x - prefix[pi];
    const int Cs = p.Cs, rps = p.H * p.W;
    const int cw = (Cs % 8 == 0) ? 8 : 1, nq = Cs / cw;
    const long long total = (long long)nb * rps * nq;
    int in_bs;
    const __half* in = resolve_in(p, I, in_bs);
    const float* inv = p.save + 2 * Cs;
    const float* shift = p.save + 3 * Cs;
    long long q = (long long)blk * EW_PER_CTA + threadIdx.x;
#pragma unroll
    for (int i = 0; i < EW_PER_THREAD; ++i, q += EW_THREADS) {
        if (q >= total) break;
        const unsigned ru = (unsigned)q / (unsigned)nq; const long long r = ru; const int c0 = (int)((unsigned)q - ru * (unsigned)nq) * cw;
        const long long io = in_row_off(r, rps, in_bs, Cs) + c0, oo = out_row_off(r, rps, p.out_bs, p.Fs) + c0;
        if (cw == 8) {
            // inv and shift hold fp16 values: y = fp16(fp16(x * inv) + shift) is two packed instructions per pair (the _rn forms
            // are never contracted into one fma, which would round once instead of twice)
            const Half2x4 x2 = as_h2(ld8(in + io));
            const float4 i0 = *reinterpret_cast<const float4*>(inv + c0), i1 = *reinterpret_cast<const float4*>(inv + c0 + 4);
            const float4 s0 = *reinterpret_cast<const float4*>(shift + c0), s1 = *reinterpret_cast<const float4*>(shift + c0 + 4);
            Half2x4 y2;
            y2.v[0] = __hadd2_rn(__hmul2_rn(x2.v[0], __floats2half2_rn(i0.x, i0.y)), __floats2half2_rn(s0.x, s0.y));
            y2.v[1] = __hadd2_rn(__hmul2_rn(x2.v[1], __floats2half2_rn(i0.z, i0.w)), __floats2half2_rn(s0.z, s0.w));
            y2.v[2] = __hadd2_rn(__hmul2_rn(x2.v[2], __floats2half2_rn(i1.x, i1.y)), __floats2half2_rn(s1.x, s1.y));
            y2.v[3] = __hadd2_rn(__hmul2_rn(x2.v[3], __floats2half2_rn(i1.z, i1.w)), __floats2half2_rn(s1.z, s1.w));
            st8(p.out + oo, as_h8(y2));
        } else {
            p.out[oo] = __float2half_rn(rh(__half2float(in[io]) * inv[c0]) + shift[c0]);
        }
    }
}

// one CTA per problem: dgamma, dbeta into the fp32 gradient arena; coefficients for the dx pass are left in
// chunk 0's partial slots: [c][0] = mean(dy), [c][1] = gamma * inv_raw^3 * mean(dy * (x - mean16))
static __global__ void __launch_bounds__(128)
k_bn_finalize_bwd(const Prob* __restrict__ probs, int nb) {
    const Prob p = probs[blockIdx.x];
    bn_finalize_bwd_dev(p, nb);
}
static __device__ void bn_finalize_bwd_dev(const Prob& p, int nb) {
    const int C = p.C, Cs = p.Cs;
    const long long rows = (long long)nb * p.H * p.W;
    const int chunks = col_chunks(rows, Cs);
    float* part = p.save + 4 * Cs;
    for (int c = threadIdx.x; c < Cs; c += blockDim.x) {
        if (c >= C) { part[c * 2] = 0.f; part[c * 2 + 1] = 0.f; continue; }
        double s1, s2;
        bn_chunk_sums(part, chunks, Cs, c, s1, s2);
        const float inv_raw = p.save[Cs + c];
        const float gamma = __half2float(p.w[c]);
        atomicAdd(&p.gb[c], (float)s1);                       // dbeta
        atomicAdd(&p.gw[c], (float)(s2 * inv_raw));           // dgamma
        part[c * 2] = (float)(s1 / (double)rows);
        part[c * 2 + 1] = (float)((double)gamma * inv_raw * inv_raw * inv_raw * s2 / (double)rows);
    }
}

// dx = inv * (dy - mean(dy)) - (x - mean32) * gamma * inv_raw^3 * mean(dy * (x - mean16)).
// The two means differ on purpose: TF's moments() takes the variance around the fp32 mean (its gradient carries
// x - mean32), while the normalisation y = x * inv + (beta - mean16 * inv) uses the fp16 mean (d inv = sum dy * (x - mean16)).
// On channels whose spread is small against the fp16 spacing of their mean the difference is a visible share of dx.
static __global__ void __launch_bounds__(EW_THREADS)
k_bn_bwd_apply(const Prob* __restrict__ probs, const int* __restrict__ prefix, int nprob, int nb, Inputs I) {
    const int pi = find_prob(prefix, nprob, blockIdx.x);
    const Prob p = probs[pi];
    if (p.din == nullptr) return;
    const int blk = blockIdx.x - prefix[pi];
    const int Cs = p.Cs, rps = p.H * p.W;
    const int cw = (Cs % 8 == 0) ? 8 : 1, nq = Cs / cw;
    const long long total = (long long)nb * rps * nq;
    int in_bs;
    const __half* in = resolve_in(p, I, in_bs);
    const float* mean32 = p.save;
    const float* inv = p.save + 2 * Cs;
    const float* coef = p.save + 4 * Cs;
    long long q = (long long)blk * EW_PER_CTA + threadIdx.x;
#pragma unroll
    for (int i = 0; i < EW_PER_THREAD; ++i, q += EW_THREADS) {
        if (q >= total) break;
        const unsigned ru = (unsigned)q / (unsigned)nq; const long long r = ru; const int c0 = (int)((unsigned)q - ru * (unsigned)nq) * cw;
        const long long io = in_row_off(r, rps, in_bs, Cs) + c0, oo = out_row_off(r, rps, p.out_bs, p.Fs) + c0;
        if (cw == 8) {
            const Half2x4 x2 = as_h2(ld8(in + io)), dy2 = as_h2(ld8(p.gout + oo));
            Half2x4 d2;
            if (p.din_acc) d2 = as_h2(ld8(p.din + io));
            float dx[8];
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const float2 xf = __half22float2(x2.v[k]), df = __half22float2(dy2.v[k]);
                const int c = c0 + 2 * k;
                const float4 cf = *reinterpret_cast<const float4*>(coef + c * 2);       // {mean dy, coefficient} of channels c, c + 1
                dx[2 * k] = inv[c] * (df.x - cf.x) - (xf.x - mean32[c]) * cf.y;
                dx[2 * k + 1] = inv[c + 1] * (df.y - cf.z) - (xf.y - mean32[c + 1]) * cf.w;
            }
#pragma unroll
            for (int j = 0; j < 8; ++j) if (c0 + j >= p.C) dx[j] = 0.f;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                __half2 h = __floats2half2_rn(dx[2 * k], dx[2 * k + 1]);
                if (p.din_acc) h = __hadd2_rn(h, d2.v[k]);                                 // fp16(dx) + fp16 accumulator, one rounding
                d2.v[k] = h;
            }
            st8(p.din + io, as_h8(d2));
        } else {
            float dx = inv[c0] * (__half2float(p.gout[oo]) - coef[c0 * 2]) - (__half2float(in[io]) - mean32[c0]) * coef[c0 * 2 + 1];
            if (p.din_acc) dx = rh(dx) + __half2float(p.din[io]);
            p.din[io] = __float2half_rn(dx);
        }
    }
}

// ------------------------------------------------------------------ max pooling (C == F, Cs == Fs)
__device__ __forceinline__ bool better_max(float v, float best) { return v > best || (v != v && best == best); }

static __global__ void __launch_bounds__(EW_THREADS)
k_maxpool_fwd(const Prob* __restrict__ probs, const int* __restrict__ prefix, int nprob, int nb, Inputs I) {
    const int pi = find_prob(prefix, nprob, blockIdx.x);
    const Prob p = probs[pi];
    const int blk = blockIdx.x - prefix[pi];
    const int Cs = p.Cs, HoWo = p.Ho * p.Wo;
    const int cw = (Cs % 8 == 0) ? 8 : 1, nq = Cs / cw;
    const long long total = (long long)nb * HoWo * nq;
    int in_bs;
    const __half* in = resolve_in(p, I, in_bs);
    long long q = (long long)blk * EW_PER_CTA + threadIdx.x;
#pragma unroll 2
    for (int it = 0; it < EW_PER_THREAD; ++it, q += EW_THREADS) {
        if (q >= total) break;
        const unsigned pix = (unsigned)q / (unsigned)nq; const int c0 = (int)((unsigned)q - pix * (unsigned)nq) * cw;
        const unsigned bu = pix / (unsigned)HoWo; const long long b = bu; const int rem = (int)(pix - bu * (unsigned)HoWo);
        const int ho = rem / p.Wo, wo = rem - ho * p.Wo;
        const __half* base = in + b * in_bs + ((long long)(ho * p.sh) * p.W + wo * p.sw) * Cs + c0;
        __half* o = p.out + out_row_off(pix, HoWo, p.out_bs, p.Fs) + c0;
        if (cw == 8) {
            // the maximum of fp16 values is one of them: packed max, NaN wins once seen (as better_max does)
            Half2x4 best = as_h2(ld8(base));
            for (int i = 0; i < p.kh; ++i)
                for (int jj = (i == 0 ? 1 : 0); jj < p.kw; ++jj) {
                    const Half2x4 x = as_h2(ld8(base + ((long long)i * p.W + jj) * Cs));
#pragma unroll
                    for (int k = 0; k < 4; ++k) best.v[k] = __hmax2_nan(best.v[k], x.v[k]);
                }
            st8(o, as_h8(best));
        } else {
            float best = -INFINITY;
            for (int i = 0; i < p.kh; ++i)
                for (int jj = 0; jj < p.kw; ++jj) {
                    const float v = __half2float(base[((long long)i * p.W + jj) * Cs]);
                    if (better_max(v, best)) best = v;
                }
            *o = __float2half_rn(best);
        }
    }
}

// gradient to the first maximum of each window (row-major scan), gathered per input element
static __global__ void __launch_bounds__(EW_THREADS)
k_maxpool_bwd(const Prob* __restrict__ probs, const int* __restrict__ prefix, int nprob, int nb, Inputs I) {
    const int pi = find_prob(prefix, nprob, blockIdx.x);
    const Prob p = probs[pi];
    if (p.din == nullptr) return;
    const int blk = blockIdx.x - prefix[pi];
    const int Cs = p.Cs, HW = p.H * p.W, HoWo = p.Ho * p.Wo;
    const int cw = (Cs % 8 == 0) ? 8 : 1, nq = Cs / cw;
    const long long total = (long long)nb * HW * nq;
    int in_bs;
    const __half* in = resolve_in(p, I, in_bs);
    long long q = (long long)blk * EW_PER_CTA + threadIdx.x;
    if (p.sh == p.kh && p.sw == p.kw) {
        // windows do not overlap (Keras' default stride): one thread per window and channel chunk finds the first
        // maximum and writes the whole window's gradients; the threads of the last window column / row also clear
        // the input pixels that no window covers
        const long long wtotal = (long long)nb * HoWo * nq;
#pragma unroll 1
        for (int it = 0; it < EW_PER_THREAD; ++it, q += EW_THREADS) {
            if (q >= wtotal) break;
            const unsigned pix = (unsigned)q / (unsigned)nq; const int c0 = (int)((unsigned)q - pix * (unsigned)nq) * cw;
            const unsigned bu = pix / (unsigned)HoWo; const long long b = bu; const int rem = (int)(pix - bu * (unsigned)HoWo);
            const int ho = rem / p.Wo, wo = rem - ho * p.Wo;
            const long long oo = out_row_off(pix, HoWo, p.out_bs, p.Fs) + c0;
            const int h_end = (ho == p.Ho - 1) ? p.H - ho * p.kh : p.kh;      // also the uncovered rows / columns
            const int w_end = (wo == p.Wo - 1) ? p.W - wo * p.kw : p.kw;
            const __half* ib = in + b * in_bs + ((long long)(ho * p.kh) * p.W + wo * p.kw) * Cs + c0;
            __half* db = p.din + b * p.in_bs + ((long long)(ho * p.kh) * p.W + wo * p.kw) * Cs + c0;
            if (cw == 8) {
                // packed: an element takes the window's gradient iff it equals the maximum (or both are NaN) and no earlier
                // element of the window did; everything is bit masks on half pairs, no conversions (pad channels carry a
                // zero gradient, so they need no mask of their own)
                const Half2x4 m8 = as_h2(ld8(p.out + oo)), g8 = as_h2(ld8(p.gout + oo));
                unsigned found[4] = {0u, 0u, 0u, 0u}, nan_m[4];
#pragma unroll
                for (int k = 0; k < 4; ++k) nan_m[k] = __hneu2_mask(m8.v[k], m8.v[k]);
                for (int i = 0; i < h_end; ++i)
                    for (int jj = 0; jj < w_end; ++jj) {
                        __half* d = db + ((long long)i * p.W + jj) * Cs;
                        Half2x4 o;
                        if (i < p.kh && jj < p.kw) {
                            const Half2x4 e = as_h2(ld8(ib + ((long long)i * p.W + jj) * Cs));
#pragma unroll
                            for (int k = 0; k < 4; ++k) {
                                const unsigned eq = __heq2_mask(e.v[k], m8.v[k]) | (__hneu2_mask(e.v[k], e.v[k]) & nan_m[k]);
                                const unsigned take = eq & ~found[k];
                                found[k] |= eq;
                                unsigned bits = *reinterpret_cast<const unsigned*>(&g8.v[k]) & take;
                                o.v[k] = *reinterpret_cast<__half2*>(&bits);
                            }
                        } else {
#pragma unroll
                            for (int k = 0; k < 4; ++k) o.v[k] = __float2half2_rn(0.f);
                        }
                        if (p.din_acc) {
                            const Half2x4 prev = as_h2(ld8(d));
#pragma unroll
                            for (int k = 0; k < 4; ++k) o.v[k] = __hadd2_rn(o.v[k], prev.v[k]);
                        }
                        st8(d, as_h8(o));
                    }
                continue;
            }
            float mx = __half2float(p.out[oo]), go = __half2float(p.gout[oo]);
            int arg = -1;
            for (int i = 0; i < p.kh; ++i)
                for (int jj = 0; jj < p.kw; ++jj) {
                    const float v = __half2float(ib[((long long)i * p.W + jj) * Cs]);
                    if (arg < 0 && (v == mx || (v != v && mx != mx))) arg = i * p.kw + jj;
                }
            for (int i = 0; i < h_end; ++i)
                for (int jj = 0; jj < w_end; ++jj) {
                    __half* d = db + ((long long)i * p.W + jj) * Cs;
                    const int id = (i < p.kh && jj < p.kw) ? i * p.kw + jj : -2;
                    float v = (arg == id) ? go : 0.f;
                    if (p.din_acc) v = rh(v) + __half2float(*d);
                    *d = __float2half_rn(v);
                }
        }
        return;
    }
#pragma unroll 1
    for (int it = 0; it < EW_PER_THREAD; ++it, q += EW_THREADS) {
        if (q >= total) break;
        const unsigned pix = (unsigned)q / (unsigned)nq; const int c0 = (int)((unsigned)q - pix * (unsigned)nq) * cw;
        const unsigned bu = pix / (unsigned)HW; const long long b = bu; const int rem = (int)(pix - bu * (unsigned)HW);
        const int h = rem / p.W, w = rem - h * p.W;
        const __half* ib = in + b * in_bs + c0;
        float xv[8], g[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) { xv[j] = 0.f; g[j] = 0.f; }
        if (cw == 8) { const Half8 x = ld8(ib + ((long long)h * p.W + w) * Cs);
#pragma unroll
            for (int j = 0; j < 8; ++j) xv[j] = __half2float(x.v[j]);
        } else xv[0] = __half2float(ib[((long long)h * p.W + w) * Cs]);
        // windows (ho, wo) that contain (h, w)
        int ho_lo = h - p.kh + 1; ho_lo = ho_lo <= 0 ? 0 : (ho_lo + p.sh - 1) / p.sh;
        int wo_lo = w - p.kw + 1; wo_lo = wo_lo <= 0 ? 0 : (wo_lo + p.sw - 1) / p.sw;
        const int ho_hi = min(p.Ho - 1, h / p.sh), wo_hi = min(p.Wo - 1, w / p.sw);
        for (int ho = ho_lo; ho <= ho_hi; ++ho)
            for (int wo = wo_lo; wo <= wo_hi; ++wo) {
                const long long oo = b * p.out_bs + ((long long)ho * p.Wo + wo) * p.Fs + c0;
                float mx[8], go[8];
                if (cw == 8) { const Half8 m8 = ld8(p.out + oo), g8 = ld8(p.gout + oo);
#pragma unroll
                    for (int j = 0; j < 8; ++j) { mx[j] = __half2float(m8.v[j]); go[j] = __half2float(g8.v[j]); }
                } else { mx[0] = __half2float(p.out[oo]); go[0] = __half2float(p.gout[oo]); }
                unsigned cand = 0;          // lanes where this element equals the window maximum
                for (int j = 0; j < cw; ++j) if (xv[j] == mx[j] || (xv[j] != xv[j] && mx[j] != mx[j])) cand |= 1u << j;
                if (!cand) continue;
                // first maximum wins: drop lanes where an earlier element of the window equals the maximum
                const int hi0 = ho * p.sh, wi0 = wo * p.sw;
                for (int i = 0; i < p.kh && cand; ++i)
                    for (int jj = 0; jj < p.kw && cand; ++jj) {
                        const int hh = hi0 + i, ww = wi0 + jj;
                        if (hh == h && ww == w) { i = p.kh; break; }
                        const __half* s = ib + ((long long)hh * p.W + ww) * Cs;
                        if (cw == 8) { const Half8 e = ld8(s);
#pragma unroll
                            for (int j = 0; j < 8; ++j) { const float v = __half2float(e.v[j]); if (v == mx[j] || (v != v && mx[j] != mx[j])) cand &= ~(1u << j); }
                        } else { const float v = __half2float(*s); if (v == mx[0] || (v != v && mx[0] != mx[0])) cand = 0; }
                    }
                for (int j = 0; j < cw; ++j) if (cand & (1u << j)) g[j] += go[j];
            }
        __half* d = p.din + b * p.in_bs + (long long)rem * Cs + c0;
        if (cw == 8) {
            Half8 o;
            if (p.din_acc) o = ld8(d);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                float v = g[j];
                if (c0 + j >= p.C) v = 0.f;
                if (p.din_acc) v = rh(v) + __half2float(o.v[j]);
                o.v[j] = __float2half_rn(v);
            }
            st8(d, o);
        } else {
            float v = g[0];
            if (p.din_acc) v = rh(v) + __half2float(*d);
            *d = __float2half_rn(v);
        }
    }
}

// ------------------------------------------------------------------ Dense on a one-channel model input (K = 1)
// out[r][f] = act(fp16(fp16(x[r] * w[f]) + b[f])): an outer product, bound by the output write.  One thread per row and
// chunk of 8 units; bit-identical to the GEMM path (one exact fp32 product, then the same roundings).
static __global__ void __launch_bounds__(EW_THREADS)
k_dense_k1_fwd(const Prob* __restrict__ probs, const int* __restrict__ prefix, int nprob, int nb, Inputs I) {
    const int pi = find_prob(prefix, nprob, blockIdx.x);
    const Prob p = probs[pi];
    const int blk = blockIdx.x - prefix[pi];
    const int Fs = p.Fs, rps = p.Ho * p.Wo, nq = Fs >> 3;
    const unsigned total = (unsigned)nb * (unsigned)rps * (unsigned)nq;
    int in_bs;
    const __half* in = resolve_in(p, I, in_bs);
    unsigned q = (unsigned)blk * EW_PER_CTA + threadIdx.x;
#pragma unroll
    for (int i = 0; i < EW_PER_THREAD; ++i, q += EW_THREADS) {
        if (q >= total) break;
        const unsigned r = q / (unsigned)nq; const int c0 = (int)(q - r * (unsigned)nq) * 8;
        const float x = __half2float(in[in_row_off(r, p.H * p.W, in_bs, 1)]);
        const Half8 w = ld8(p.w + c0), b = ld8(p.bias + c0);
        Half8 o;
        if (p.act != 2 && c0 + 8 <= p.F) {
            // linear / ReLU: fp16(fp16(x * w) + b) and the clamp as packed halves (pad columns of w and b are zero)
            const Half2x4 w2 = as_h2(w), b2 = as_h2(b);
            const __half2 x2 = __float2half2_rn(x), zero = __float2half2_rn(0.f);
            Half2x4 o2;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                __half2 z = __hadd2_rn(__hmul2_rn(x2, w2.v[k]), b2.v[k]);
                if (p.act == 1) z = __hmax2_nan(z, zero);
                o2.v[k] = z;
            }
            o = as_h8(o2);
        } else {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                float z = 0.f;
                if (c0 + j < p.F) z = act_forward(rh(rh(x * __half2float(w.v[j])) + __half2float(b.v[j])), p.act);
                o.v[j] = __float2half_rn(z);
            }
        }
        st8(p.out + out_row_off(r, rps, p.out_bs, Fs) + c0, o);
    }
}

// ------------------------------------------------------------------ copy: concatenate (last axis) and non-view reshape
// The source's n_el real elements per sample, in row-major (row, channel) order, fill rows of Cj channels at
// channel offset coff of the destination.  fwd: out <- in;  bwd: din (+)= gout.
// Work unit = one 16-byte chunk of 8 source channels when the source has a padded channel stride (every tensor but
// the model inputs), else one element.  mode 2: source and destination chunks are both aligned (C, Cj, coff
// multiples of 8): 16-byte load and store.  mode 1: aligned source chunk, compacted (unaligned) destination:
// 16-byte access on the source side, 2-byte accesses on the destination side.  mode 0: element by element.
__host__ __device__ inline int copy_mode(const Prob& p) {
    if (p.Cs % 8 != 0) return 0;
    return (p.C % 8 == 0 && p.Cj % 8 == 0 && p.coff % 8 == 0 && p.Fs % 8 == 0) ? 2 : 1;
}
__host__ __device__ inline int copy_blocks(const Prob& p, int nb) {
    const long long per = copy_mode(p) ? (long long)(p.n_el / p.C) * (p.Cs / 8) : p.n_el;
    return (int)(((long long)nb * per + EW_PER_CTA - 1) / EW_PER_CTA);
}
// Forward compaction into one destination row (flatten + concatenate of a tensor whose channel count is not a multiple
// of 8, or at an unaligned offset): the work unit is an aligned 16-byte chunk of the DESTINATION, gathered element by
// element from the padded source, so that the stores are full chunks instead of eight 2-byte stores per source chunk.
__host__ __device__ inline bool copy_dst_major(const Prob& p) { return copy_mode(p) == 1 && p.Cj == p.n_el; }
__host__ __device__ inline int copy_blocks_fwd(const Prob& p, int nb) {
    if (!copy_dst_major(p)) return copy_blocks(p, nb);
    const long long per = (p.coff + p.n_el + 7) / 8 - p.coff / 8;
    return (int)(((long long)nb * per + EW_PER_CTA - 1) / EW_PER_CTA);
}

static __global__ void __launch_bounds__(EW_THREADS)
k_copy(const Prob* __restrict__ probs, const int* __restrict__ prefix, int nprob, int nb, Inputs I, int bwd) {
    const int pi = find_prob(prefix, nprob, blockIdx.x);
    const Prob p = probs[pi];
    if (bwd && p.din == nullptr) return;
    const int blk = blockIdx.x - prefix[pi];
    const int mode = copy_mode(p);
    int in_bs;
    const __half* in = resolve_in(p, I, in_bs);
    if (!bwd && copy_dst_major(p)) {
        const int q0 = p.coff >> 3, nqd = ((p.coff + p.n_el + 7) >> 3) - q0;
        const unsigned total = (unsigned)nb * (unsigned)nqd;
        unsigned u = (unsigned)blk * EW_PER_CTA + threadIdx.x;
#pragma unroll 1
        for (int i = 0; i < EW_PER_THREAD; ++i, u += EW_THREADS) {
            if (u >= total) break;
            const unsigned b = u / (unsigned)nqd; const int q = q0 + (int)(u - b * (unsigned)nqd);
            const int e0 = q * 8 - p.coff;                       // flat source element of the chunk's first slot (may be < 0)
            const int ec = max(e0, 0);
            int sr = (int)((unsigned)ec / (unsigned)p.C), sc = ec - sr * p.C;
            const __half* src = in + (long long)b * in_bs;
            __half* dst = p.out + (long long)b * p.out_bs + q * 8;
            Half8 v;
            bool full = e0 >= 0 && e0 + 8 <= p.n_el;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const int e = e0 + j;
                if (e >= 0 && e < p.n_el) {
                    v.v[j] = src[(long long)sr * p.Cs + sc];
                    if (++sc == p.C) { sc = 0; ++sr; }
                    if (!full) dst[j] = v.v[j];
                }
            }
            if (full) st8(dst, v);
        }
        return;
    }
    const int nq = mode ? p.Cs / 8 : 1;
    const int per = mode ? (p.n_el / p.C) * nq : p.n_el;      // units per sample
    const unsigned total = (unsigned)nb * (unsigned)per;      // units of the whole batch fit 32 bits (srn_build checks tensor sizes)
    unsigned u = (unsigned)blk * EW_PER_CTA + threadIdx.x;
#pragma unroll 1
    for (int i = 0; i < EW_PER_THREAD; ++i, u += EW_THREADS) {
        if (u >= total) break;
        const unsigned bb = u / (unsigned)per; const int uu = (int)(u - bb * (unsigned)per);
        const long long b = bb;
        int sr, sc;
        if (mode) { sr = (int)((unsigned)uu / (unsigned)nq); sc = (uu - sr * nq) * 8; }
        else if (p.C == 1) { sr = uu; sc = 0; }
        else { sr = (int)((unsigned)uu / (unsigned)p.C); sc = uu - sr * p.C; }
        if (sc >= p.C) continue;                                // a chunk that is pad only
        const long long so = b * in_bs + (long long)sr * p.Cs + sc;
        const int ee = sr * p.C + sc;                           // flat index of the first element of the unit
        int dr = (int)((unsigned)ee / (unsigned)p.Cj), dc = ee - dr * p.Cj;
        if (mode == 2) {
            const long long dd = b * p.out_bs + (long long)dr * p.Fs + p.coff + dc;
            if (!bwd) st8(p.out + dd, ld8(in + so));
            else {
                const long long go = b * p.in_bs + (long long)sr * p.Cs + sc;
                Half8 g = ld8(p.gout + dd);
                if (p.din_acc) {
                    const Half8 o = ld8(p.din + go);
#pragma unroll
                    for (int j = 0; j < 8; ++j) g.v[j] = __float2half_rn(__half2float(g.v[j]) + __half2float(o.v[j]));
                }
                st8(p.din + go, g);
            }
        } else {
            const int n = mode ? min(8, p.C - sc) : 1;          // real channels in this unit
            Half8 v;
            const long long go = b * p.in_bs + (long long)sr * p.Cs + sc;
            if (!bwd) { if (mode) v = ld8(in + so); else v.v[0] = in[so]; }
            else if (p.din_acc) { if (mode) v = ld8(p.din + go); else v.v[0] = p.din[go]; }
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                if (j < n) {
                    const long long dd = b * p.out_bs + (long long)dr * p.Fs + p.coff + dc;
                    if (!bwd) p.out[dd] = v.v[j];
                    else {
                        float g = __half2float(p.gout[dd]);
                        if (p.din_acc) g += __half2float(v.v[j]);
                        v.v[j] = __float2half_rn(g);
                    }
                    if (++dc == p.Cj) { dc = 0; ++dr; }
                } else if (bwd) {
                    v.v[j] = __float2half_rn(0.f);                 // pad channels of the source gradient stay zero
                }
            }
            if (bwd) { if (mode) st8(p.din + go, v); else p.din[go] = v.v[0]; }
        }
    }
}

// clears spans of the parameter arena (one block per span): kernel rows that stand for pad positions of a padded
// concatenation
static __global__ void k_zero_spans(__half* __restrict__ P, const long long* __restrict__ off, const int* __restrict__ len) {
    __half* p = P + off[blockIdx.x];
    for (int i = threadIdx.x; i < len[blockIdx.x]; i += blockDim.x) p[i] = __float2half_rn(0.f);
}

// ------------------------------------------------------------------ shared im2col of a one-channel model input
// In training and evaluation every individual sees the same image / genotype batch, so the patch matrix of a
// first layer depends only on its window geometry: it is written once per batch and geometry as
// [sample][position][Kp] (Kp = kh*kw rounded up to 8, pad zero) and every first layer with that geometry becomes a
// plain Dense over it on the vectorised tensor-core path.
struct Im2colGeo {
    const __half* in; __half* out;
    int in_bs, H, W, kh, kw, sh, sw, Ho, Wo, KK, Kp;
};

static __global__ void __launch_bounds__(256)
k_im2col(Im2colGeo g, int nb) {
    const int nq = g.Kp >> 3, HoWo = g.Ho * g.Wo;
    const long long total = (long long)nb * HoWo * nq;
    long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= total) return;
    const unsigned pos = (unsigned)q / (unsigned)nq; const int k0 = (int)((unsigned)q - pos * (unsigned)nq) * 8;
    const unsigned bu = pos / (unsigned)HoWo; const long long b = bu; const int rem = (int)(pos - bu * (unsigned)HoWo);
    const int ho = rem / g.Wo, wo = rem - ho * g.Wo;
    const __half* src = g.in + b * g.in_bs + (long long)(ho * g.sh) * g.W + wo * g.sw;
    Half8 v;
    int ki = k0 / g.kw, kj = k0 - ki * g.kw;
#pragma unroll
    for (int e = 0; e < 8; ++e) {
        // column KK (the first pad column, when there is one) holds 1: the weight-gradient GEMM over this matrix then
        // yields the bias gradient in its row KK, and the forward pass is unaffected (there is no kernel row KK)
        v.v[e] = (k0 + e < g.KK) ? src[ki * g.W + kj] : __float2half_rn(k0 + e == g.KK ? 1.f : 0.f);
        if (++kj == g.kw) { kj = 0; ++ki; }
    }
    st8(g.out + pos * g.Kp + k0, v);
}

// ------------------------------------------------------------------ fused losses
// One warp per (individual, row).  logits [nb][ncls + G] fp16 (head output), targets: labels and the
// genotype bits of the same rows.  Writes dlogits (training) and per-slot metric sums.
struct LossArgs {
    const uint8_t* labels;       // [nb]
    const __half* bits;          // genotype targets, row stride bits_bs, per-individual stride bits_istride
    long long bits_istride; int bits_bs;
    const float* loss_balance;   // [n]
    int* met_correct; float* met_ce; float* met_mse;   // [slots][n]
    int n, ncls, G;
    int slot_base, chunk, row0;  // metric slot of row b = slot_base + (row0 + b) / chunk
    int training;
};

static __global__ void __launch_bounds__(256)
k_loss(const Prob* __restrict__ probs, int nprob, int nb, LossArgs a) {
    const int warps_per_cta = 8;
    const int ctas_per_prob = (nb + warps_per_cta - 1) / warps_per_cta;
    const int pi = blockIdx.x / ctas_per_prob;
    if (pi >= nprob) return;
    const Prob p = probs[pi];
    const int b_raw = (blockIdx.x - pi * ctas_per_prob) * warps_per_cta + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    // a warp past the end of the batch recomputes the last sample (its gradient stores repeat the same values) and is left
    // out of the metrics: every warp reaches the block-wide combination of the metric sums below
    const bool counted = b_raw < nb;
    const int b = counted ? b_raw : nb - 1;
    const int NL = a.ncls + a.G;
    const __half* z = p.out + (long long)b * p.out_bs;
    const int label = a.labels[b];
    // ---- classification: log-sum-exp in fp32 from the fp16 logits
    float lse; int arg = 0;
    if (a.ncls <= 32) {
        // lane j holds class j: one expf per lane, and the sums run over the classes in index order through shuffles, so
        // every lane gets the bits the sequential loop below gives
        const float zl = lane < a.ncls ? __half2float(z[lane]) : -INFINITY;
        float zmax = -INFINITY;
        for (int j = 0; j < a.ncls; ++j) {
            const float v = __shfl_sync(0xffffffffu, zl, j);
            if (v > zmax) zmax = v;
        }
        const float el = expf(zl - zmax);
        float se = 0.f;
        for (int j = 0; j < a.ncls; ++j) se += __shfl_sync(0xffffffffu, el, j);
        lse = zmax + logf(se);
        // categorical_accuracy takes argmax over the model's fp16 softmax OUTPUT: probabilities that round to the same
        // half tie and the first index wins (an all-NaN row keeps index 0, as tf.argmax does)
        const float pl = rh(el / se);
        float pbest = __shfl_sync(0xffffffffu, pl, 0);
        for (int j = 1; j < a.ncls; ++j) {
            const float pj = __shfl_sync(0xffffffffu, pl, j);
            if (pj > pbest) { pbest = pj; arg = j; }
        }
    } else {
        float zmax = -INFINITY;
        for (int j = 0; j < a.ncls; ++j) zmax = fmaxf(zmax, __half2float(z[j]));       // (NaN logits: se is NaN either way)
        float se = 0.f;
        for (int j = 0; j < a.ncls; ++j) se += expf(__half2float(z[j]) - zmax);
        lse = zmax + logf(se);
        float pbest = rh(expf(__half2float(z[0]) - zmax) / se);
        for (int j = 1; j < a.ncls; ++j) {
            const float pj = rh(expf(__half2float(z[j]) - zmax) / se);
            if (pj > pbest) { pbest = pj; arg = j; }
        }
    }
    const float ce = rh(lse - __half2float(z[label]));
    // ---- replication: sigmoid + squared error
    const float lb16 = rh(a.loss_balance[p.indiv]);
    const float wr16 = rh(1.0f - lb16);
    const float gc = rh(lb16 / (float)nb);                 // d total / d ce_b
    const float g2 = rh(rh(wr16 / (float)nb) / (float)a.G); // d total / d sq_j
    const __half* t = a.bits + (long long)p.indiv * a.bits_istride + (long long)b * a.bits_bs;
    float sq_sum = 0.f;
    for (int j = lane; j < a.G; j += 32) {
        float s = rh(1.f / (1.f + expf(-__half2float(z[a.ncls + j]))));
        float d = s - __half2float(t[j]);
        sq_sum += rh(d * d);
        if (a.training) {
            // SquaredDifference's gradient scalar_mul(2, grad) * (x - y) and SigmoidGrad, one rounding per op
            float dy = rh(rh(2.f * g2) * rh(d));
            float ga = rh(dy * s), gb = rh(1.f - s);
            p.gout[(long long)b * p.out_bs + a.ncls + j] = __float2half_rn(ga * gb);
        }
    }
    if (a.training) {
        for (int j = lane; j < a.ncls; j += 32) {
            float pj = expf(__half2float(z[j]) - lse);
            p.gout[(long long)b * p.out_bs + j] = __float2half_rn(gc * (pj - (j == label ? 1.f : 0.f)));
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) sq_sum += __shfl_xor_sync(0xffffffffu, sq_sum, o);
    // The 8 samples of a block belong to one individual: their metric sums are combined in shared memory and added with
    // one set of atomics per block (750 same-address atomics per individual and metric serialised at the L2 otherwise).
    __shared__ float s_ce[8], s_mse[8];
    __shared__ int s_ok[8], s_slot[8];
    const int w = threadIdx.x >> 5;
    if (lane == 0) {
        s_mse[w] = counted ? rh(sq_sum / (float)a.G) : 0.f;
        s_ce[w] = counted ? ce : 0.f;
        s_ok[w] = counted && arg == label ? 1 : 0;
        s_slot[w] = counted ? a.slot_base + (a.row0 + b) / a.chunk : -1;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        int k = 0;
        while (k < warps_per_cta && s_slot[k] >= 0) {          // runs of samples in the same metric slot (a chunk boundary can cut the block)
            const int slot = s_slot[k];
            float ce_sum = 0.f, mse_sum = 0.f; int ok = 0;
            for (; k < warps_per_cta && s_slot[k] == slot; ++k) { ce_sum += s_ce[k]; mse_sum += s_mse[k]; ok += s_ok[k]; }
            const long long mi = (long long)slot * a.n + p.indiv;
            if (ok) atomicAdd(&a.met_correct[mi], ok);
            atomicAdd(&a.met_ce[mi], ce_sum);
            atomicAdd(&a.met_mse[mi], mse_sum);
        }
    }
    (void)NL;
}

// ------------------------------------------------------------------ multi-tensor Adam over the packed arena
// TF's ApplyAdam with T = half: every intermediate rounded to fp16.  Gradients are fp32 accumulators,
// rounded to fp16 on read and cleared for the next step.
struct AdamScalars { float alpha, omb1, omb2, eps; };

static __global__ void __launch_bounds__(256)
k_adam(__half* __restrict__ p, __half* __restrict__ m, __half* __restrict__ v, float* __restrict__ g,
       long long n8, AdamScalars s) {
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n8) return;
    uint4 pv = reinterpret_cast<uint4*>(p)[i], mv = reinterpret_cast<uint4*>(m)[i], vv = reinterpret_cast<uint4*>(v)[i];
    float4 g0 = reinterpret_cast<float4*>(g)[2 * i], g1 = reinterpret_cast<float4*>(g)[2 * i + 1];
    float gf[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
    __half* ph = reinterpret_cast<__half*>(&pv);
    __half* mh = reinterpret_cast<__half*>(&mv);
    __half* vh = reinterpret_cast<__half*>(&vv);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        float gi = rh(gf[k]);
        float mi = __half2float(mh[k]), vi = __half2float(vh[k]), pi = __half2float(ph[k]);
        mi = rh(mi + rh(s.omb1 * rh(gi - mi)));
        vi = rh(vi + rh(s.omb2 * rh(rh(gi * gi) - vi)));
        float upd = rh(rh(s.alpha * mi) / rh(s.eps + rh(sqrtf(vi))));
        pi = rh(pi - upd);
        mh[k] = __float2half_rn(mi); vh[k] = __float2half_rn(vi); ph[k] = __float2half_rn(pi);
    }
    reinterpret_cast<uint4*>(p)[i] = pv;
    reinterpret_cast<uint4*>(m)[i] = mv;
    reinterpret_cast<uint4*>(v)[i] = vv;
    const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
    reinterpret_cast<float4*>(g)[2 * i] = z;
    reinterpret_cast<float4*>(g)[2 * i + 1] = z;
}

// ------------------------------------------------------------------ batch gather from the resident dataset
// rows idx[b] (or start + b when idx == nullptr) of the image / genotype / label arrays -> staged batch
static __global__ void __launch_bounds__(256)
k_gather_batch(const __half* __restrict__ x, const uint8_t* __restrict__ gbits, const uint8_t* __restrict__ y,
               const int* __restrict__ idx, int start, int nb, int xe, int G,
               __half* __restrict__ xo, __half* __restrict__ go, uint8_t* __restrict__ yo) {
    const int b = blockIdx.x;
    if (b >= nb) return;
    const long long src = idx ? idx[b] : start + b;
    for (int e = threadIdx.x; e < xe; e += blockDim.x) xo[(long long)b * xe + e] = x[src * xe + e];
    if (go) for (int e = threadIdx.x; e < G; e += blockDim.x) go[(long long)b * G + e] = __float2half_rn((float)gbits[src * G + e]);
    if (yo && threadIdx.x == 0) yo[b] = y[src];
}

static __global__ void k_f32_to_f16(const float* __restrict__ in, __half* __restrict__ out, long long n) {
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = __float2half_rn(in[i]);
}

static __global__ void k_u8_to_f16(const uint8_t* __restrict__ in, __half* __restrict__ out, long long n) {
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = __float2half_rn((float)in[i]);
}

// ------------------------------------------------------------------ genotype readout
// One warp per offspring row: sigmoid probabilities (fp16) -> bit = p > 0.5 (== np.round(np.clip(p, 0, 1)),
// half-to-even), ballot-packed 32 loci per word, and the Hamming distance to the parent.
// NaN probabilities give bit 0 and count as a mutation (nan != parent), as NumPy does.
static __global__ void __launch_bounds__(256)
k_readout(const __half* __restrict__ probs, const uint8_t* __restrict__ parent, int rows, int rows_per_parent, int G,
          uint32_t* __restrict__ packed, int* __restrict__ mutations) {
    const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (row >= rows) return;
    const int words = (G + 31) / 32;
    const __half* p = probs + (long long)row * G;
    const uint8_t* par = parent + (long long)(row / rows_per_parent) * G;
    int mut = 0;
    for (int w = 0; w < words; ++w) {
        int j = w * 32 + lane;
        bool bit = false, diff = false;
        if (j < G) {
            float v = __half2float(p[j]);
            bit = v > 0.5f;
            diff = (v != v) || (bit != (par[j] != 0));
        }
        uint32_t word = __ballot_sync(0xffffffffu, bit);
        mut += __popc(__ballot_sync(0xffffffffu, diff));
        if (lane == 0) packed[(long long)row * words + w] = word;
    }
    if (lane == 0) mutations[row] = mut;
}

// ------------------------------------------------------------------ parameter initialisation
__device__ __forceinline__ uint64_t splitmix64(uint64_t x) {
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}

// uniform(-limit, limit) per element; columns >= split_col use limit_b (the fused head: two Dense layers)
// rows are `ld` elements apart; the pad columns [ncols, ld) stay zero
static __global__ void k_init_uniform(__half* w, long long count, int ncols, int ld, int split_col, float limit_a, float limit_b,
                               uint64_t seed) {
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    const int col = (int)(i % ld);
    if (col >= ncols) { w[i] = __float2half_rn(0.f); return; }
    uint64_t r = splitmix64(seed ^ splitmix64((uint64_t)i));
    float u = (float)(r >> 40) * (1.0f / 16777216.0f);          // [0, 1)
    float limit = (col >= split_col) ? limit_b : limit_a;
    w[i] = __float2half_rn((2.f * u - 1.f) * limit);
}

static __global__ void k_fill_f16(__half* w, long long count, float value) {
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < count) w[i] = __float2half_rn(value);
}

}  // namespace srn
