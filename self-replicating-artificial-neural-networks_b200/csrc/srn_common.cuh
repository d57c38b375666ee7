// Shared device-side definitions for the SeRANN population engine (sm_100a).
//
// Tensor layout: every activation / activation-gradient tensor is [sample][row][Cs] fp16, channel last, where
// rows = product of the leading axes, C real channels and Cs = C rounded up to 8 the channel stride.  The pad
// channels are zero when the arena is created and no kernel ever writes them, so 16-byte chunks of 8 channels
// are always aligned, never straddle a pixel, and contribute exact zeros to every contraction.  The two model
// inputs (image, genotype) have a single channel and are the only tensors with Cs = C = 1.
#pragma once
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace srn {

// One (individual, layer) problem of a population-grouped launch.  A launch carries an array of these
// plus an exclusive prefix sum of per-problem work items; a CTA finds its problem by binary search.
struct Prob {
    const __half* in;      // input activations (ignored when in_sel != 0)
    __half* out;           // output activations
    const __half* w;       // kernel rows [K][ldw] (Keras layout, rows padded to ldw) or BN gamma
    const __half* bias;    // bias [F] or BN beta
    __half* din;           // gradient wrt input (nullptr: not needed)
    __half* gout;          // gradient wrt output (activation backward rewrites it in place)
    float* gw;             // fp32 gradient accumulators (kernel / gamma)
    float* gb;             // (bias / beta)
    __half* mov;           // BN: moving_mean [C] then moving_variance [C]
    float* save;           // BN: [4][Cs] saved mean16, inv_raw, inv, shift; then partial sums [chunks][Cs][2]
    int H, W, C, Cs;       // input  (per sample): C channels, channel stride Cs
    int Ho, Wo, F, Fs;     // output (per sample)
    int kh, kw, sh, sw;
    int act;
    int in_bs, out_bs;     // batch strides (elements)
    int in_sel;            // 0: `in`; 1: image input; 2: genotype input
    int din_acc;           // 1: add into din instead of overwriting
    int indiv;             // index of the individual in the shard
    int ldw;               // leading dimension of the kernel rows (F rounded up to 8; the pad stays zero)
    // copy problems (concatenate / non-view reshape): n_el real elements of the source per sample are laid
    // out row-major into rows of Cj channels at channel offset coff of the destination rows
    int coff, n_el, Cj;
};

// Where tensors 0 (image) and 1 (genotype) come from in this launch.
struct Inputs {
    const __half* x; long long x_istride; int x_bs;   // per-individual stride (0 = shared batch), batch stride
    const __half* g; long long g_istride; int g_bs;
};

__device__ __forceinline__ const __half* resolve_in(const Prob& p, const Inputs& I, int& bs) {
    if (p.in_sel == 1) { bs = I.x_bs; return I.x + (long long)p.indiv * I.x_istride; }
    if (p.in_sel == 2) { bs = I.g_bs; return I.g + (long long)p.indiv * I.g_istride; }
    bs = p.in_bs;
    return p.in;
}

// round-to-fp16 and back: the value an all-fp16 graph would hold after this op
__device__ __forceinline__ float rh(float x) { return __half2float(__float2half_rn(x)); }

__device__ __forceinline__ int find_prob(const int* __restrict__ prefix, int nprob, int item) {
    int lo = 0, hi = nprob;            // prefix has nprob + 1 entries; find last p with prefix[p] <= item
    while (hi - lo > 1) {
        int mid = (lo + hi) >> 1;
        if (prefix[mid] <= item) lo = mid; else hi = mid;
    }
    return lo;
}

__device__ __forceinline__ float act_forward(float z, int act) {
    if (act == 1) return z < 0.f ? 0.f : z;          // NaN propagates, as Eigen's max does
    if (act == 2) return rh(1.f / (1.f + expf(-z)));
    return z;
}

// 8 consecutive fp16 values as one 16-byte access
struct __align__(16) Half8 { __half v[8]; };
__device__ __forceinline__ Half8 ld8(const __half* p) { return *reinterpret_cast<const Half8*>(p); }
__device__ __forceinline__ void st8(__half* p, const Half8& h) { *reinterpret_cast<Half8*>(p) = h; }

}  // namespace srn
