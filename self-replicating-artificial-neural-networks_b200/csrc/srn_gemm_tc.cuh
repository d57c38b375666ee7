// Population-grouped implicit GEMM on the 5th-generation tensor cores (tcgen05 + TMEM), sm_100a.
//
// One CTA (128 threads) owns a 128 x BN output tile of one (individual, layer) problem.  Operands are
// gathered from the NHWC fp16 activations / padded weight rows with 16-byte cp.async into shared memory in
// the *canonical no-swizzle* UMMA layout: 8x8-element core matrices of 128 contiguous bytes, core (mn_group,
// k_group) at ((mn_group * 8 + k_group) * 128) bytes.  Both K-major operands (8 rows x 16 B of k) and MN-major
// operands (8 k-rows x 16 B of m/n) use that arrangement, so every descriptor has LBO = 128 B (next k group)
// and SBO = 1024 B (next m/n group), and only the major bits of the instruction descriptor differ:
//   forward  Out = A(K-major, im2col rows) x W(MN-major, [k][n] rows)
//   dgrad    dIn = dOut'(K-major)         x W(K-major, rows (ij, c), contiguous f)
//   wgrad    dW  = A(MN-major, [row][kk]) x dOut(MN-major, [row][f])      (reduction over rows)
// A 4-stage ring of (A, B) tiles is filled two k-blocks ahead; one thread issues four K=16 tcgen05.mma per
// k-block and commits to the ring's mbarriers; the fp32 accumulator lives in TMEM and is read back with
// tcgen05.ld by the four warps (warp w owns TMEM lanes 32w..32w+31) for the same epilogues as the SIMT kernels.
//
// Every activation tensor has its channel stride padded to 8 with zero pad (srn_common.cuh), so a 16-byte
// chunk is always aligned and inside one pixel; the k index of a pass runs over the padded (pixel, channel)
// space and weight rows of pad channels are zero-filled.  Layers reading a model input (one channel, no pad)
// use the SCALAR_A variant: the im2col operand is gathered element by element and stored to shared memory
// with ordinary 16-byte stores; everything else is identical.
#pragma once
#include "srn_common.cuh"

namespace srn {

constexpr int TC_BM = 128;            // UMMA M
constexpr int TC_BK = 64;             // k elements per stage (4 MMAs of K = 16)
constexpr int TC_STAGES = 4;
constexpr int TC_PF = 3;              // prefetch distance in k-blocks
constexpr int TC_THREADS = 128;
constexpr int TC_A_BYTES = TC_BM * TC_BK * 2;
constexpr int TC_WGRAD_ROWS = 4096;   // reduction rows per wgrad CTA (64 k-blocks)

enum { TC_FWD = 0, TC_DGRAD = 1, TC_WGRAD = 2 };

__host__ __device__ inline int tc_smem_bytes(int BN) { return TC_STAGES * (TC_A_BYTES + BN * TC_BK * 2) + 256; }
__host__ __device__ inline int tc_tmem_cols(int BN) { return BN <= 32 ? 32 : (BN <= 64 ? 64 : (BN <= 128 ? 128 : 256)); }
__host__ __device__ inline int round_up8(int v) { return (v + 7) & ~7; }

// reduction length of each pass as the kernel sees it: over padded channels
__host__ __device__ inline int tc_k_fwd(const Prob& p) { return p.kh * p.kw * p.Cs; }
__host__ __device__ inline int tc_k_dgrad(const Prob& p) { return p.kh * p.kw * p.Fs; }

__host__ __device__ inline int tc_tiles_fwd(const Prob& p, int nb, int BN) {
    return ((nb * p.Ho * p.Wo + TC_BM - 1) / TC_BM) * ((p.F + BN - 1) / BN);
}
__host__ __device__ inline int tc_tiles_dgrad(const Prob& p, int nb, int BN) {
    return ((nb * p.H * p.W + TC_BM - 1) / TC_BM) * ((p.C + BN - 1) / BN);
}
__host__ __device__ inline int tc_tiles_wgrad(const Prob& p, int nb, int BN) {
    return ((tc_k_fwd(p) + TC_BM - 1) / TC_BM) * ((p.F + BN - 1) / BN) * ((nb * p.Ho * p.Wo + TC_WGRAD_ROWS - 1) / TC_WGRAD_ROWS);
}

// ------------------------------------------------------------------ PTX wrappers
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, bool valid) {
    const int sz = valid ? 16 : 0;                      // src-size 0: the 16 bytes are zero-filled
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(dst), "l"(src), "r"(sz) : "memory");
}
__device__ __forceinline__ void cp_async16_ca(uint32_t dst, const void* src, bool valid) {   // keeps the line in L1 (im2col reuse)
    const int sz = valid ? 16 : 0;
    asm volatile("cp.async.ca.shared.global [%0], [%1], 16, %2;\n" ::"r"(dst), "l"(src), "r"(sz) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;\n" ::"n"(N) : "memory"); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory"); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;\n" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred P1;\n"
        "WAIT_LOOP:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n"
        "@P1 bra WAIT_DONE;\n"
        "bra WAIT_LOOP;\n"
        "WAIT_DONE:\n"
        "}\n" ::"r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;\n" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;\n" ::: "memory"); }

__device__ __forceinline__ void tmem_alloc(uint32_t slot_smem, uint32_t cols) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;\n" ::"r"(slot_smem), "r"(cols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;\n" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t cols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;\n" ::"r"(taddr), "r"(cols) : "memory");
}

// shared-memory matrix descriptor, no swizzle: start >> 4 | LBO >> 4 << 16 | SBO >> 4 << 32 | version 1 << 46
__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr) {
    return (uint64_t)((saddr & 0x3FFFFu) >> 4) | ((uint64_t)(128u >> 4) << 16) | ((uint64_t)(1024u >> 4) << 32) | (1ull << 46);
}
// instruction descriptor, kind::f16: D = f32 (bits 4-5 = 1), A = B = f16 (0), major bits 15/16, N >> 3 at 17, M >> 4 at 24
__device__ __forceinline__ uint32_t umma_idesc(int a_mn_major, int b_mn_major, int n) {
    return (1u << 4) | ((uint32_t)a_mn_major << 15) | ((uint32_t)b_mn_major << 16) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(TC_BM >> 4) << 24);
}
__device__ __forceinline__ void umma_f16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n"
        "}\n" ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n" ::"r"(bar) : "memory");
}
// 16 consecutive fp32 columns of this thread's TMEM lane
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float (&v)[16]) {
    uint32_t r[16];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];\n"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr)
        : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;\n" ::: "memory");
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}

// byte offset of the 16-byte unit (mn_group, k_group, row-in-core) inside a tile
__device__ __forceinline__ uint32_t core_off(int mn_group, int k_group, int r8) {
    return (uint32_t)((mn_group * 8 + k_group) * 128 + r8 * 16);
}

// ------------------------------------------------------------------ the kernel
template <int MODE, bool SCALAR_A>
static __global__ void __launch_bounds__(TC_THREADS)
k_gemm_tc(const Prob* __restrict__ probs, const int* __restrict__ prefix, int nprob, int nb, int BN, Inputs I) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int l8 = lane & 7, lq = lane >> 3;              // row inside a core matrix, chunk inside a quad
    const uint32_t sraw = smem_u32(smem_raw);
    const uint32_t sbase = (sraw + 127u) & ~127u;
    unsigned char* sgen = smem_raw + (sbase - sraw);      // generic pointer to the same place as sbase
    const uint32_t B_BYTES = (uint32_t)BN * TC_BK * 2;
    const uint32_t sA = sbase, sB = sbase + TC_STAGES * TC_A_BYTES;
    const uint32_t sBar = sB + TC_STAGES * B_BYTES;      // TC_STAGES "empty" barriers + 1 "done" barrier
    const uint32_t sSlot = sBar + (TC_STAGES + 1) * 8;
    volatile uint32_t* slot_ptr = reinterpret_cast<volatile uint32_t*>(sgen + (sSlot - sbase));

    const int pi = find_prob(prefix, nprob, blockIdx.x);
    const Prob p = probs[pi];
    int tile = blockIdx.x - prefix[pi];
    const int tmem_cols = tc_tmem_cols(BN);

    if (tid == 0) {
        for (int i = 0; i <= TC_STAGES; ++i) mbar_init(sBar + i * 8, 1);
        asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
    }
    if (warp == 0) tmem_alloc(sSlot, (uint32_t)tmem_cols);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *slot_ptr;

    // ---- problem geometry
    int in_bs;
    const __half* in = resolve_in(p, I, in_bs);
    const int HoWo = p.Ho * p.Wo, HW = p.H * p.W;
    const int Cs = p.Cs, Fs = p.Fs;
    const int KK = p.kh * p.kw;
    const int Kf = SCALAR_A ? KK : KK * Cs;               // forward reduction length (padded channels)
    const int Kd = KK * Fs;
    int m0 = 0, n0 = 0, Mrows = 0, Ncols = 0, nk = 0, r0 = 0, r1 = 0;
    if (MODE == TC_FWD) {
        Mrows = nb * HoWo; Ncols = p.F;
        const int tn = (Ncols + BN - 1) / BN;
        m0 = (tile / tn) * TC_BM; n0 = (tile % tn) * BN;
        nk = (Kf + TC_BK - 1) / TC_BK;
    } else if (MODE == TC_DGRAD) {
        Mrows = nb * HW; Ncols = p.C;
        const int tn = (Ncols + BN - 1) / BN;
        m0 = (tile / tn) * TC_BM; n0 = (tile % tn) * BN;
        nk = (Kd + TC_BK - 1) / TC_BK;
    } else {
        Mrows = Kf; Ncols = p.F;                          // output rows are kk = (i, j, padded c)
        const int tm = (Mrows + TC_BM - 1) / TC_BM, tn = (Ncols + BN - 1) / BN;
        const int split = tile / (tm * tn);
        tile -= split * tm * tn;
        m0 = (tile / tn) * TC_BM; n0 = (tile % tn) * BN;
        r0 = split * TC_WGRAD_ROWS; r1 = min(nb * HoWo, r0 + TC_WGRAD_ROWS);
        nk = (r1 - r0 + TC_BK - 1) / TC_BK;
    }
    const int n_umma = min(BN, ((Ncols - n0 + 15) / 16) * 16);

    // ---- per-thread loader state that does not change over the k loop
    const __half* a_row[4];          // FWD: base of the im2col row; null = out of range
    int a_b[4], a_h[4], a_w[4];      // DGRAD: input position of the row
    long long w_koff[4]; bool w_kok[4];   // WGRAD (vector): offset of this thread's four kk chunks inside the window
    if (MODE == TC_FWD) {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int m = m0 + (warp * 4 + j) * 8 + l8;
            a_row[j] = nullptr;
            if (m < Mrows) {
                const int b = m / HoWo, rem = m - b * HoWo;
                const int ho = rem / p.Wo, wo = rem - ho * p.Wo;
                a_row[j] = in + (long long)b * in_bs + ((long long)(ho * p.sh) * p.W + wo * p.sw) * Cs;
            }
        }
    } else if (MODE == TC_DGRAD) {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int m = m0 + (warp * 4 + j) * 8 + l8;
            a_b[j] = -1; a_h[j] = 0; a_w[j] = 0;
            if (m < Mrows) {
                a_b[j] = m / HW;
                const int rem = m - a_b[j] * HW;
                a_h[j] = rem / p.W; a_w[j] = rem - a_h[j] * p.W;
            }
        }
    } else if (!SCALAR_A) {
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const int kk = m0 + (q * 4 + lq) * 8;
            w_kok[q] = kk < Kf;
            const int ij = kk / Cs, cp = kk - ij * Cs;
            const int i = ij / p.kw, j = ij - i * p.kw;
            w_koff[q] = ((long long)i * p.W + j) * Cs + cp;
        }
    }

    auto load_stage = [&](int stage, int kb) {
        const uint32_t aS = sA + stage * TC_A_BYTES, bS = sB + stage * B_BYTES;
        const int items_b = BN >> 4;                     // B items per warp
        if (MODE == TC_FWD) {
            // A, K-major: 16 row blocks x 2 quads of k chunks
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const int q = warp * 8 + i, blk = q >> 1, ch = (q & 1) * 4 + lq;
                const int k = kb * TC_BK + ch * 8;
                const __half* row = a_row[i >> 1];
                if (SCALAR_A) {
                    Half8 v;
                    int ki = k / p.kw, kj = k - ki * p.kw;
#pragma unroll
                    for (int e = 0; e < 8; ++e) {
                        v.v[e] = (row != nullptr && k + e < Kf) ? row[ki * p.W + kj] : __float2half_rn(0.f);
                        if (++kj == p.kw) { kj = 0; ++ki; }
                    }
                    *reinterpret_cast<Half8*>(sgen + (aS - sbase) + core_off(blk, ch, l8)) = v;
                } else {
                    const bool ok = row != nullptr && k < Kf;
                    const int ij = k / Cs, cp = k - ij * Cs;
                    const int ki = ij / p.kw, kj = ij - ki * p.kw;
                    cp_async16_ca(aS + core_off(blk, ch, l8),
                                  ok ? (const void*)(row + ((long long)ki * p.W + kj) * Cs + cp) : (const void*)p.w, ok);
                }
            }
            // B, MN-major: 8 k blocks x BN/32 quads of n chunks; W rows are [(ij, c)][ldw], pad channels have no row
            const int Q = BN >> 5;
            for (int i = 0; i < items_b; ++i) {
                const int q = warp * items_b + i, blk = q / Q, nch = (q - blk * Q) * 4 + lq;
                const int k = kb * TC_BK + blk * 8 + l8, n = n0 + nch * 8;
                long long wrow = k;
                bool ok = k < Kf && n < p.ldw;
                if (!SCALAR_A) { const int ij = k / Cs, cp = k - ij * Cs; ok = ok && cp < p.C; wrow = (long long)ij * p.C + cp; }
                cp_async16(bS + core_off(nch, blk, l8), ok ? (const void*)(p.w + wrow * p.ldw + n) : (const void*)p.w, ok);
            }
        } else if (MODE == TC_DGRAD) {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const int q = warp * 8 + i, blk = q >> 1, ch = (q & 1) * 4 + lq;
                const int k = kb * TC_BK + ch * 8;
                const int j = i >> 1;
                bool ok = a_b[j] >= 0 && k < Kd;
                const __half* src = p.gout;
                if (ok) {
                    const int ij = k / Fs, f = k - ij * Fs;
                    const int ki = ij / p.kw, kj = ij - ki * p.kw;
                    const int hh = a_h[j] - ki, ww = a_w[j] - kj;
                    ok = hh >= 0 && ww >= 0;
                    if (ok) {
                        const int ho = hh / p.sh, wo = ww / p.sw;
                        ok = ho * p.sh == hh && wo * p.sw == ww && ho < p.Ho && wo < p.Wo;
                        if (ok) src = p.gout + (long long)a_b[j] * p.out_bs + ((long long)ho * p.Wo + wo) * Fs + f;
                    }
                }
                cp_async16_ca(aS + core_off(blk, ch, l8), src, ok);
            }
            // B, K-major: rows n = input channel c, k = (ij, f) contiguous in f
            for (int i = 0; i < items_b; ++i) {
                const int q = warp * items_b + i, blk = q >> 1, ch = (q & 1) * 4 + lq;
                const int c = n0 + blk * 8 + l8, k = kb * TC_BK + ch * 8;
                const bool ok = c < p.C && k < Kd;
                const int ij = k / Fs, f = k - ij * Fs;
                cp_async16(bS + core_off(blk, ch, l8),
                           ok ? (const void*)(p.w + ((long long)ij * p.C + c) * p.ldw + f) : (const void*)p.w, ok);
            }
        } else {
            // rows of this k-block: r = r0 + kb*64 + blk*8 + l8, two blocks per thread
            const __half* in_row[2]; const __half* go_row[2];
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                const int r = r0 + kb * TC_BK + (warp * 2 + j) * 8 + l8;
                in_row[j] = nullptr; go_row[j] = nullptr;
                if (r < r1) {
                    const int b = r / HoWo, rem = r - b * HoWo;
                    const int ho = rem / p.Wo, wo = rem - ho * p.Wo;
                    in_row[j] = in + (long long)b * in_bs + ((long long)(ho * p.sh) * p.W + wo * p.sw) * Cs;
                    go_row[j] = p.gout + (long long)b * p.out_bs + (long long)rem * Fs;
                }
            }
            // A, MN-major: 8 row blocks x 4 quads of kk chunks
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const int q = warp * 8 + i, blk = q >> 2, qd = q & 3, cc = qd * 4 + lq;
                const __half* row = (i < 4) ? in_row[0] : in_row[1];
                if (SCALAR_A) {
                    Half8 v;
                    const int kk = m0 + cc * 8;
                    int ki = kk / p.kw, kj = kk - ki * p.kw;
#pragma unroll
                    for (int e = 0; e < 8; ++e) {
                        v.v[e] = (row != nullptr && kk + e < Kf) ? row[ki * p.W + kj] : __float2half_rn(0.f);
                        if (++kj == p.kw) { kj = 0; ++ki; }
                    }
                    *reinterpret_cast<Half8*>(sgen + (aS - sbase) + core_off(cc, blk, l8)) = v;
                } else {
                    const bool ok = row != nullptr && w_kok[qd];
                    cp_async16_ca(aS + core_off(cc, blk, l8), ok ? (const void*)(row + w_koff[qd]) : (const void*)p.w, ok);
                }
            }
            // B, MN-major: 8 row blocks x BN/32 quads of f chunks
            const int Q = BN >> 5;
            for (int i = 0; i < items_b; ++i) {
                const int q = warp * items_b + i, blk = q / Q, nch = (q - blk * Q) * 4 + lq;
                const __half* row = (blk - warp * 2) ? go_row[1] : go_row[0];
                const int f = n0 + nch * 8;
                const bool ok = row != nullptr && f < Fs;
                cp_async16(bS + core_off(nch, blk, l8), ok ? (const void*)(row + f) : (const void*)p.w, ok);
            }
        }
    };

    const uint32_t idesc = umma_idesc(MODE == TC_WGRAD ? 1 : 0, MODE == TC_DGRAD ? 0 : 1, n_umma);

    // ---- main loop: fill TC_PF k-blocks ahead, one thread issues the MMAs
    for (int s = 0; s < TC_PF; ++s) {
        if (s < nk) load_stage(s, s);
        cp_async_commit();
    }
    for (int kb = 0; kb < nk; ++kb) {
        const int nxt = kb + TC_PF;
        if (nxt < nk) {
            const int sb = nxt % TC_STAGES;
            if (nxt >= TC_STAGES) mbar_wait(sBar + sb * 8, (uint32_t)((nxt / TC_STAGES - 1) & 1));
            load_stage(sb, nxt);
        }
        cp_async_commit();
        cp_async_wait<TC_PF>();
        fence_proxy_async();
        __syncthreads();
        if (tid == 0) {
            tc_fence_after();
            const int st = kb % TC_STAGES;
            const uint32_t aS = sA + st * TC_A_BYTES, bS = sB + st * B_BYTES;
#pragma unroll
            for (int ks = 0; ks < TC_BK / 16; ++ks)
                umma_f16(tmem, umma_desc(aS + ks * 256), umma_desc(bS + ks * 256), idesc, (kb | ks) ? 1u : 0u);
            umma_commit(sBar + st * 8);
            if (kb == nk - 1) umma_commit(sBar + TC_STAGES * 8);
        }
    }
    mbar_wait(sBar + TC_STAGES * 8, 0);
    tc_fence_after();

    // ---- epilogue: this thread owns accumulator row (warp * 32 + lane); whole 8-channel chunks are written,
    //      pad channels as zero
    const int mrow = m0 + warp * 32 + lane;
    const uint32_t trow = tmem + ((uint32_t)(warp * 32) << 16);
    if (MODE == TC_FWD) {
        __half* orow = nullptr;
        if (mrow < Mrows) {
            const int b = mrow / HoWo, rem = mrow - b * HoWo;
            orow = p.out + (long long)b * p.out_bs + (long long)rem * Fs;
        }
        for (int cb = 0; cb < n_umma; cb += 16) {
            float v[16];
            tmem_ld16(trow + cb, v);
            if (!orow) continue;
#pragma unroll
            for (int h8 = 0; h8 < 2; ++h8) {
                const int n = n0 + cb + h8 * 8;
                if (n >= Ncols) break;
                Half8 o;
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    float z = 0.f;
                    if (n + j < Ncols) z = act_forward(rh(rh(v[h8 * 8 + j]) + __half2float(p.bias[n + j])), p.act);
                    o.v[j] = __float2half_rn(z);
                }
                st8(orow + n, o);
            }
        }
    } else if (MODE == TC_DGRAD) {
        __half* drow = nullptr;
        if (mrow < Mrows) {
            const int b = mrow / HW, rem = mrow - b * HW;
            drow = p.din + (long long)b * p.in_bs + (long long)rem * Cs;
        }
        for (int cb = 0; cb < n_umma; cb += 16) {
            float v[16];
            tmem_ld16(trow + cb, v);
            if (!drow) continue;
#pragma unroll
            for (int h8 = 0; h8 < 2; ++h8) {
                const int c = n0 + cb + h8 * 8;
                if (c >= Ncols) break;
                Half8 o;
                if (p.din_acc) o = ld8(drow + c);
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    float x = (c + j < Ncols) ? v[h8 * 8 + j] : 0.f;
                    if (p.din_acc) x = rh(x) + __half2float(o.v[j]);
                    o.v[j] = __float2half_rn(x);
                }
                st8(drow + c, o);
            }
        }
    } else {
        // accumulator row = kk = (ij, padded c): rows of pad channels have no weight row
        float* grow = nullptr;
        if (mrow < Kf) {
            if (SCALAR_A) grow = p.gw + (long long)mrow * p.ldw;
            else { const int ij = mrow / Cs, cp = mrow - ij * Cs; if (cp < p.C) grow = p.gw + ((long long)ij * p.C + cp) * p.ldw; }
        }
        const bool single = (nb * HoWo) <= TC_WGRAD_ROWS;     // one split: plain stores, else fp32 atomics
        for (int cb = 0; cb < n_umma; cb += 16) {
            float v[16];
            tmem_ld16(trow + cb, v);
            if (!grow) continue;
#pragma unroll
            for (int j = 0; j < 16; ++j) {
                const int f = n0 + cb + j;
                if (f < Ncols) {
                    if (single) grow[f] = v[j];
                    else if (v[j] != 0.f) atomicAdd(&grow[f], v[j]);
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem, (uint32_t)tmem_cols);
}

}  // namespace srn
