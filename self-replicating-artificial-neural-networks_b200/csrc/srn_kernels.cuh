// HBM-bound kernels of the SeRANN engine: activation backward + bias gradient, BatchNormalizationF16,
// max pooling, concat, fused losses, multi-tensor Adam, batch gather, readout, initialisation.
// All are population-grouped (see Prob in srn_common.cuh) except Adam/readout, which run flat.
#pragma once
#include "srn_common.cuh"

namespace srn {

constexpr int COL_THREADS = 256;      // 32 channel lanes x 8 row lanes
constexpr int COL_ROWS = 1024;        // rows per CTA of the column-reduction kernels

__host__ __device__ inline int col_chunks(long long rows) { return (int)((rows + COL_ROWS - 1) / COL_ROWS); }

// rows x C view of a per-sample-strided tensor: row r -> sample r / rps, inner row r % rps
__device__ __forceinline__ long long row_off(long long r, int rps, int bs, int C) {
    long long b = r / rps;
    return b * bs + (r - b * rps) * (long long)C;
}

// block-level sum over the 8 row lanes; result valid in row lane 0
__device__ __forceinline__ float rowlane_sum(float v, float (*sm)[33], int cl, int rl) {
    sm[rl][cl] = v;
    __syncthreads();
    float s = 0.f;
    if (rl == 0) {
#pragma unroll
        for (int i = 0; i < 8; ++i) s += sm[i][cl];
    }
    __syncthreads();
    return s;
}

// ------------------------------------------------------------------ activation backward + bias gradient
// gout <- fp16(gout * act'(out)) in place; gb[f] += column sums of the result.
static __global__ void __launch_bounds__(COL_THREADS)
k_act_bwd_bias(const Prob* __restrict__ probs, const int* __restrict__ prefix, int nprob, int nb) {
    __shared__ float sm[8][33];
    const int pi = find_prob(prefix, nprob, blockIdx.x);
    const Prob p = probs[pi];
    const int chunk = blockIdx.x - prefix[pi];
    const int rps = p.Ho * p.Wo, F = p.F;
    const long long rows = (long long)nb * rps;
    const long long r0 = (long long)chunk * COL_ROWS, r1 = min(rows, r0 + COL_ROWS);
    const int cl = threadIdx.x & 31, rl = threadIdx.x >> 5;
    for (int f0 = 0; f0 < F; f0 += 32) {
        const int f = f0 + cl;
        float s = 0.f;
        if (f < F) {
            for (long long r = r0 + rl; r < r1; r += 8) {
                long long o = row_off(r, rps, p.out_bs, F) + f;
                float g = __half2float(p.gout[o]);
                if (p.act == 1) {
                    g = __half2float(p.out[o]) > 0.f ? g : 0.f;
                    p.gout[o] = __float2half_rn(g);
                } else if (p.act == 2) {
                    float y = __half2float(p.out[o]);
                    g = rh(rh(g * y) * rh(1.f - y));
                    p.gout[o] = __float2half_rn(g);
                }
                s += g;
            }
        }
        s = rowlane_sum(s, sm, cl, rl);
        if (rl == 0 && f < F && s != 0.f) atomicAdd(&p.gb[f], s);
    }
}

// ------------------------------------------------------------------ BatchNormalizationF16
// save layout per problem: [0][C] mean16, [1][C] inv_raw (rsqrt(var16 + eps), fp16-valued),
// [2][C] inv (times gamma), [3][C] shift; then partial sums [chunks][C][2].
constexpr float BN_EPS16 = 0.0010004043579101562f;       // fp16(1e-3)
constexpr float BN_DECAY16 = 0.01000213623046875f;        // fp16(1 - 0.99)
constexpr float BN_ONE_EPS16 = 1.0009765625f;             // fp16(1 + 1e-3)

static __global__ void __launch_bounds__(COL_THREADS)
k_bn_partial(const Prob* __restrict__ probs, const int* __restrict__ prefix, int nprob, int nb, Inputs I, int bwd) {
    __shared__ float sm[8][33];
    const int pi = find_prob(prefix, nprob, blockIdx.x);
    const Prob p = probs[pi];
    const int chunk = blockIdx.x - prefix[pi];
    const int rps = p.H * p.W, C = p.C;
    const long long rows = (long long)nb * rps;
    const long long r0 = (long long)chunk * COL_ROWS, r1 = min(rows, r0 + COL_ROWS);
    int in_bs;
    const __half* in = resolve_in(p, I, in_bs);
    float* part = p.save + 4 * C + (long long)chunk * C * 2;
    const int cl = threadIdx.x & 31, rl = threadIdx.x >> 5;
    for (int c0 = 0; c0 < C; c0 += 32) {
        const int c = c0 + cl;
        float s1 = 0.f, s2 = 0.f;
        if (c < C) {
            // forward: shifted sums around the first row's value; backward: sum dy, sum dy * (x - mean16)
            const float pivot = bwd ? p.save[c] : __half2float(in[c]);
            for (long long r = r0 + rl; r < r1; r += 8) {
                float x = __half2float(in[row_off(r, rps, in_bs, C) + c]) - pivot;
                if (bwd) {
                    float dy = __half2float(p.gout[row_off(r, rps, p.out_bs, C) + c]);
                    s1 += dy;
                    s2 += dy * x;
                } else {
                    s1 += x;
                    s2 += x * x;
                }
            }
        }
        s1 = rowlane_sum(s1, sm, cl, rl);
        s2 = rowlane_sum(s2, sm, cl, rl);
        if (rl == 0 && c < C) { part[c * 2] = s1; part[c * 2 + 1] = s2; }
    }
}

__device__ __forceinline__ void bn_scale_shift(float mean16, float var16, float gamma, float beta,
                                               float& inv_raw, float& inv, float& shift) {
    inv_raw = rh(1.0f / sqrtf(rh(var16 + BN_EPS16)));
    inv = rh(inv_raw * gamma);
    shift = rh(beta - rh(mean16 * inv));
}

// one CTA per problem: combine partials, derive scale/shift, update the moving statistics (fp16 op by op)
static __global__ void __launch_bounds__(128)
k_bn_finalize_fwd(const Prob* __restrict__ probs, int nb, Inputs I) {
    const Prob p = probs[blockIdx.x];
    const int C = p.C;
    const long long rows = (long long)nb * p.H * p.W;
    const int chunks = col_chunks(rows);
    int in_bs;
    const __half* in = resolve_in(p, I, in_bs);
    const float* part = p.save + 4 * C;
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
        double s1 = 0.0, s2 = 0.0;
        for (int k = 0; k < chunks; ++k) { s1 += part[((long long)k * C + c) * 2]; s2 += part[((long long)k * C + c) * 2 + 1]; }
        const double n = (double)rows;
        const double dm = s1 / n;
        const float mean32 = (float)((double)__half2float(in[c]) + dm);
        const float var32 = (float)fmax(s2 / n - dm * dm, 0.0);
        const float mean16 = rh(mean32), var16 = rh(var32);
        float inv_raw, inv, shift;
        bn_scale_shift(mean16, var16, __half2float(p.w[c]), __half2float(p.bias[c]), inv_raw, inv, shift);
        p.save[c] = mean16; p.save[C + c] = inv_raw; p.save[2 * C + c] = inv; p.save[3 * C + c] = shift;
        // moving statistics: variance * n16 / (n16 - (1 + eps)) with n cast to fp16 (inf/inf = NaN above 65504)
        const float n16 = rh((float)rows);
        const float bessel = rh(n16 / rh(n16 - BN_ONE_EPS16));
        const float var_unb = rh(var16 * bessel);
        float mm = __half2float(p.mov[c]), mv = __half2float(p.mov[C + c]);
        mm = rh(mm - rh(rh(mm - mean16) * BN_DECAY16));
        mv = rh(mv - rh(rh(mv - var_unb) * BN_DECAY16));
        p.mov[c] = __float2half_rn(mm);
        p.mov[C + c] = __float2half_rn(mv);
    }
}

// inference: scale/shift from the moving statistics
static __global__ void __launch_bounds__(128)
k_bn_prepare_infer(const Prob* __restrict__ probs) {
    const Prob p = probs[blockIdx.x];
    const int C = p.C;
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
        float inv_raw, inv, shift;
        bn_scale_shift(__half2float(p.mov[c]), __half2float(p.mov[C + c]), __half2float(p.w[c]),
                       __half2float(p.bias[c]), inv_raw, inv, shift);
        p.save[2 * C + c] = inv; p.save[3 * C + c] = shift;
    }
}

constexpr int EW_THREADS = 256;
constexpr int EW_PER_THREAD = 8;
constexpr int EW_PER_CTA = EW_THREADS * EW_PER_THREAD;
__host__ __device__ inline int ew_blocks(long long elems) { return (int)((elems + EW_PER_CTA - 1) / EW_PER_CTA); }

// y = fp16(fp16(x * inv) + shift)
static __global__ void __launch_bounds__(EW_THREADS)
k_bn_apply(const Prob* __restrict__ probs, const int* __restrict__ prefix, int nprob, int nb, Inputs I) {
    const int pi = find_prob(prefix, nprob, blockIdx.x);
    const Prob p = probs[pi];
    const int blk = blockIdx.x - prefix[pi];
    const int C = p.C, per = p.H * p.W * C;
    const long long total = (long long)nb * per;
    int in_bs;
    const __half* in = resolve_in(p, I, in_bs);
    const float* inv = p.save + 2 * C;
    const float* shift = p.save + 3 * C;
    long long e = (long long)blk * EW_PER_CTA + threadIdx.x;
#pragma unroll
    for (int i = 0; i < EW_PER_THREAD; ++i, e += EW_THREADS) {
        if (e >= total) break;
        long long b = e / per; int r = (int)(e - b * per); int c = r % C;
        float x = __half2float(in[b * in_bs + r]);
        p.out[b * p.out_bs + r] = __float2half_rn(rh(x * inv[c]) + shift[c]);
    }
}

// one CTA per problem: dgamma, dbeta into the fp32 gradient arena; coefficients for the dx pass
// save[3][C] is reused: [3][c] <- inv * s1 / n ... kept simple: store s1/n and k2 = gamma*inv_raw^3*s2/n
static __global__ void __launch_bounds__(128)
k_bn_finalize_bwd(const Prob* __restrict__ probs, int nb) {
    const Prob p = probs[blockIdx.x];
    const int C = p.C;
    const long long rows = (long long)nb * p.H * p.W;
    const int chunks = col_chunks(rows);
    float* part = p.save + 4 * C;
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
        double s1 = 0.0, s2 = 0.0;
        for (int k = 0; k < chunks; ++k) { s1 += part[((long long)k * C + c) * 2]; s2 += part[((long long)k * C + c) * 2 + 1]; }
        const float inv_raw = p.save[C + c];
        const float gamma = __half2float(p.w[c]);
        atomicAdd(&p.gb[c], (float)s1);                       // dbeta
        atomicAdd(&p.gw[c], (float)(s2 * inv_raw));           // dgamma
        // coefficients overwrite chunk 0's slot of this channel (only this thread reads it); dx pass reads them
        part[c * 2] = (float)(s1 / (double)rows);
        part[c * 2 + 1] = (float)((double)gamma * inv_raw * inv_raw * inv_raw * s2 / (double)rows);
    }
}

// dx = inv * (dy - mean(dy)) - (x - mean16) * gamma * inv_raw^3 * mean(dy * (x - mean16))
static __global__ void __launch_bounds__(EW_THREADS)
k_bn_bwd_apply(const Prob* __restrict__ probs, const int* __restrict__ prefix, int nprob, int nb, Inputs I) {
    const int pi = find_prob(prefix, nprob, blockIdx.x);
    const Prob p = probs[pi];
    if (p.din == nullptr) return;
    const int blk = blockIdx.x - prefix[pi];
    const int C = p.C, per = p.H * p.W * C;
    const long long total = (long long)nb * per;
    int in_bs;
    const __half* in = resolve_in(p, I, in_bs);
    const float* mean16 = p.save;
    const float* inv = p.save + 2 * C;
    const float* coef = p.save + 4 * C;
    long long e = (long long)blk * EW_PER_CTA + threadIdx.x;
#pragma unroll
    for (int i = 0; i < EW_PER_THREAD; ++i, e += EW_THREADS) {
        if (e >= total) break;
        long long b = e / per; int r = (int)(e - b * per); int c = r % C;
        float x = __half2float(in[b * in_bs + r]) - mean16[c];
        float dy = __half2float(p.gout[b * p.out_bs + r]);
        float dx = inv[c] * (dy - coef[c * 2]) - x * coef[c * 2 + 1];
        __half* d = p.din + b * p.in_bs + r;
        if (p.din_acc) dx = rh(dx) + __half2float(*d);
        *d = __float2half_rn(dx);
    }
}

// ------------------------------------------------------------------ max pooling
static __global__ void __launch_bounds__(EW_THREADS)
k_maxpool_fwd(const Prob* __restrict__ probs, const int* __restrict__ prefix, int nprob, int nb, Inputs I) {
    const int pi = find_prob(prefix, nprob, blockIdx.x);
    const Prob p = probs[pi];
    const int blk = blockIdx.x - prefix[pi];
    const int C = p.C, per = p.Ho * p.Wo * C;
    const long long total = (long long)nb * per;
    int in_bs;
    const __half* in = resolve_in(p, I, in_bs);
    long long e = (long long)blk * EW_PER_CTA + threadIdx.x;
#pragma unroll
    for (int it = 0; it < EW_PER_THREAD; ++it, e += EW_THREADS) {
        if (e >= total) break;
        long long b = e / per; int r = (int)(e - b * per);
        int c = r % C; int t = r / C; int wo = t % p.Wo, ho = t / p.Wo;
        const __half* base = in + b * in_bs + ((long long)(ho * p.sh) * p.W + wo * p.sw) * C + c;
        float best = -INFINITY;
        for (int i = 0; i < p.kh; ++i)
            for (int j = 0; j < p.kw; ++j) {
                float v = __half2float(base[((long long)i * p.W + j) * C]);
                if (v > best || (v != v && best == best)) best = v;
            }
        p.out[b * p.out_bs + r] = __float2half_rn(best);
    }
}

// gradient to the first maximum of each window (row-major scan), gathered per input element
static __global__ void __launch_bounds__(EW_THREADS)
k_maxpool_bwd(const Prob* __restrict__ probs, const int* __restrict__ prefix, int nprob, int nb, Inputs I) {
    const int pi = find_prob(prefix, nprob, blockIdx.x);
    const Prob p = probs[pi];
    if (p.din == nullptr) return;
    const int blk = blockIdx.x - prefix[pi];
    const int C = p.C, per = p.H * p.W * C;
    const long long total = (long long)nb * per;
    int in_bs;
    const __half* in = resolve_in(p, I, in_bs);
    long long e = (long long)blk * EW_PER_CTA + threadIdx.x;
#pragma unroll 1
    for (int it = 0; it < EW_PER_THREAD; ++it, e += EW_THREADS) {
        if (e >= total) break;
        long long b = e / per; int r = (int)(e - b * per);
        int c = r % C; int t = r / C; int w = t % p.W, h = t / p.W;
        const __half* ib = in + b * in_bs + c;
        const float xv = __half2float(ib[((long long)h * p.W + w) * C]);
        float g = 0.f;
        // windows (ho, wo) that contain (h, w)
        int ho_lo = h - p.kh + 1; ho_lo = ho_lo <= 0 ? 0 : (ho_lo + p.sh - 1) / p.sh;
        int wo_lo = w - p.kw + 1; wo_lo = wo_lo <= 0 ? 0 : (wo_lo + p.sw - 1) / p.sw;
        int ho_hi = min(p.Ho - 1, h / p.sh), wo_hi = min(p.Wo - 1, w / p.sw);
        for (int ho = ho_lo; ho <= ho_hi; ++ho)
            for (int wo = wo_lo; wo <= wo_hi; ++wo) {
                long long oo = b * p.out_bs + ((long long)ho * p.Wo + wo) * C + c;
                float mx = __half2float(p.out[oo]);
                bool is_max = (xv == mx) || (xv != xv && mx != mx);
                if (!is_max) continue;
                // first maximum wins: no earlier element of the window may equal the maximum
                bool first = true;
                int hi0 = ho * p.sh, wi0 = wo * p.sw;
                for (int i = 0; i < p.kh && first; ++i)
                    for (int j = 0; j < p.kw; ++j) {
                        int hh = hi0 + i, ww = wi0 + j;
                        if (hh == h && ww == w) { i = p.kh; break; }
                        float v = __half2float(ib[((long long)hh * p.W + ww) * C]);
                        if (v == mx || (v != v && mx != mx)) { first = false; break; }
                    }
                if (first) g += __half2float(p.gout[oo]);
            }
        __half* d = p.din + b * p.in_bs + r;
        if (p.din_acc) g = rh(g) + __half2float(*d);
        *d = __float2half_rn(g);
    }
}

// ------------------------------------------------------------------ concat (last axis)
// fwd: out[b][r][coff + c] = in[b][r][c];  bwd: din[b][r][c] (+)= gout[b][r][coff + c]
static __global__ void __launch_bounds__(EW_THREADS)
k_concat(const Prob* __restrict__ probs, const int* __restrict__ prefix, int nprob, int nb, Inputs I, int bwd) {
    const int pi = find_prob(prefix, nprob, blockIdx.x);
    const Prob p = probs[pi];
    if (bwd && p.din == nullptr) return;
    const int blk = blockIdx.x - prefix[pi];
    const int C = p.C, per = p.H * p.W * C;
    const long long total = (long long)nb * per;
    int in_bs;
    const __half* in = resolve_in(p, I, in_bs);
    long long e = (long long)blk * EW_PER_CTA + threadIdx.x;
#pragma unroll
    for (int i = 0; i < EW_PER_THREAD; ++i, e += EW_THREADS) {
        if (e >= total) break;
        long long b = e / per; int r = (int)(e - b * per);
        int c = r % C; int row = r / C;
        long long oo = b * p.out_bs + (long long)row * p.F + p.coff + c;
        if (!bwd) {
            p.out[oo] = in[b * in_bs + r];
        } else {
            __half* d = p.din + b * p.in_bs + r;
            float g = __half2float(p.gout[oo]);
            if (p.din_acc) g += __half2float(*d);
            *d = __float2half_rn(g);
        }
    }
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
    const int b = (blockIdx.x - pi * ctas_per_prob) * warps_per_cta + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (b >= nb) return;
    const int NL = a.ncls + a.G;
    const __half* z = p.out + (long long)b * p.out_bs;
    const int label = a.labels[b];
    // ---- classification: log-sum-exp in fp32 from the fp16 logits (every lane computes it redundantly)
    float zmax = -INFINITY; int arg = 0;
    for (int j = 0; j < a.ncls; ++j) {
        float v = __half2float(z[j]);
        if (v > zmax) { zmax = v; arg = j; }
    }
    float se = 0.f;
    for (int j = 0; j < a.ncls; ++j) se += expf(__half2float(z[j]) - zmax);
    const float lse = zmax + logf(se);
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
            float dy = 2.f * d * g2;
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
    if (lane == 0) {
        const float mse = rh(sq_sum / (float)a.G);
        const int slot = a.slot_base + (a.row0 + b) / a.chunk;
        const long long mi = (long long)slot * a.n + p.indiv;
        if (arg == label) atomicAdd(&a.met_correct[mi], 1);
        atomicAdd(&a.met_ce[mi], ce);
        atomicAdd(&a.met_mse[mi], mse);
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
