// Population-grouped implicit GEMM on the 5th-generation tensor cores (tcgen05 + TMEM), sm_100a.
//
// One CTA (416 threads) owns 128 x BN output tiles of (individual, layer) problems.  Operands are gathered from the
// NHWC fp16 activations / padded weight rows with 16-byte cp.async into shared memory in the swizzle-128B canonical
// UMMA layouts (rows of 128 bytes in 1024-byte atoms of 8 rows, 16-byte chunk index XOR row index):
//   forward  Out = A(K-major, im2col rows) x W(MN-major, [k][n] rows)
//   dgrad    dIn = dOut'(K-major)         x W(K-major, rows (ij, c), contiguous f)
//   wgrad    dW  = A(MN-major, [row][kk]) x dOut(MN-major, [row][f])      (reduction over rows)
// K-major tiles hold 64 k per row, MN-major tiles 64 m/n of one k per row (LBO = 8192 B between groups of 64 m/n,
// SBO = 1024 B between groups of 8 rows / 8 k); only the major bits of the instruction descriptor and the per-step
// advance of the matrix descriptors differ between the passes.
// The kernel is persistent and warp-specialised.  A CTA walks its share of a host-built table of tile descriptors
// (sorted by cost, dealt out in snake order, prefetched four tiles ahead) and treats the k-blocks of its tiles as
// one stream:
//   warps 0-7  loaders: gather k-block p into ring stage p % S with cp.async, 4 rows x 8 chunks per warp-wide copy
//              (a quarter warp reads 128 contiguous bytes of one row: one L1 wavefront per line); each thread's
//              copies arrive on the stage's "full" barrier by themselves when they land (cp.async.mbarrier.arrive.noinc);
//   warp 8     MMA issuer: the converged warp waits "full", one lane elected by elect.sync issues up to four K=16
//              tcgen05.mma into one of the TMEM accumulators and commits to the stage's "empty" barrier (and to
//              "tmem_full" after a tile's last k-block);
//   warps 9-12 epilogue: wait "tmem_full", read the accumulator with tcgen05.ld (warp w owns TMEM lanes
//              32 (w % 4) ..), apply bias / activation / gradient stores, arrive on "tmem_empty".
// Loads, MMAs and epilogues of neighbouring tiles therefore overlap; nothing in the loop is a CTA-wide barrier.
//
// Every activation tensor has its channel stride padded to 8 with zero pad (srn_common.cuh), so a 16-byte
// chunk is always aligned and inside one pixel; the k index of a pass runs over the padded (pixel, channel)
// space and weight rows of pad channels are zero-filled.  Layers reading a model input (one channel, no pad)
// use the SCALAR_A variant: the im2col operand is gathered element by element and stored to shared memory
// with ordinary 16-byte stores; everything else is identical.
#pragma once
#include <cstdlib>
#include <type_traits>
#include "srn_common.cuh"

namespace srn {

constexpr int TC_BM = 128;            // UMMA M
constexpr int TC_BK = 64;             // k elements per stage (4 MMAs of K = 16)
constexpr int TC_MAX_STAGES = 8;
constexpr int TC_LW = 8;               // loader warps
constexpr int TC_LOADERS = TC_LW * 32;
constexpr int TC_THREADS = TC_LOADERS + 32 + 128;   // loaders + 1 MMA warp + 4 epilogue warps
constexpr int TC_A_BYTES = TC_BM * TC_BK * 2;
constexpr int TC_WGRAD_ROWS = 4096;   // reduction rows per wgrad CTA (64 k-blocks)

enum { TC_FWD = 0, TC_DGRAD = 1, TC_WGRAD = 2 };

constexpr int TC_RING = 24;           // tile descriptors in shared memory: epilogue <= 10 tiles behind the loader, 4 ahead
constexpr int TC_TILE_BYTES = 320;    // one tile descriptor (table entry / ring slot)
constexpr int TC_DESC_AHEAD = 4;
constexpr int TC_LUT = 512;            // 16-byte k chunks per lookup table: reductions up to 4096 elements, longer ones compute on the fly
// Pipeline depth per N tile: BN <= 64 keeps two CTAs per SM (4 stages each), wider tiles take the SM alone with as
// many stages as fit.  Loaders never wait for their own copies: completion is signalled to the stage's "full"
// barrier by the copy engine itself (cp.async.mbarrier.arrive.noinc), so up to `stages` k-blocks are in flight.
__host__ __device__ constexpr int tc_stages(int BN) { return BN <= 64 ? 4 : (BN <= 128 ? 3 : 4); }
// B tiles are stored with 128-byte rows (swizzle-128B atoms), so a 32-wide MN-major tile takes the room of a 64-wide one
__host__ __device__ constexpr int tc_b_bytes(int BN) { return (BN < 64 ? 64 : BN) * TC_BK * 2; }
__host__ __device__ constexpr int tc_bias_bytes(int BN) { return 4 * BN * 2; }      // the tile's bias chunk, one copy per epilogue warp
__host__ __device__ constexpr int tc_smem_bytes(int BN) { return tc_stages(BN) * (TC_A_BYTES + tc_b_bytes(BN)) + 384 + TC_RING * TC_TILE_BYTES + 2 * TC_LUT * 8 + tc_bias_bytes(BN); }
__host__ __device__ constexpr int tc_acc_cols(int BN) { return BN <= 32 ? 32 : (BN <= 64 ? 64 : (BN <= 128 ? 128 : 256)); }
// TMEM accumulators per CTA: 256 columns (two CTAs share the SM's 512) hold 8, 4 or 2 of them; a CTA that has the SM
// alone (BN = 256) takes all 512 columns for two.  Short tiles (one or two k-blocks) are bound by the round trip
// MMA -> epilogue -> MMA over an accumulator, so more accumulators mean more tiles in flight.
__host__ __device__ constexpr int tc_num_acc(int BN) { return BN <= 128 ? 256 / tc_acc_cols(BN) : 2; }
__host__ __device__ constexpr int tc_tmem_cols(int BN) { return tc_num_acc(BN) * tc_acc_cols(BN); }
__host__ __device__ inline int round_up8(int v) { return (v + 7) & ~7; }

// reduction length of each pass as the kernel sees it: over padded channels
__host__ __device__ inline int tc_k_fwd(const Prob& p) { return p.kh * p.kw * p.Cs; }
__host__ __device__ inline int tc_k_dgrad(const Prob& p) { return p.kh * p.kw * p.Fs; }

// Forward problems whose reduction fits one k-block and that have many row tiles (first layers, per-pixel Dense) are cut
// into tiles of TC_SUB consecutive 128-row sub-tiles: one descriptor, one loader/issuer/epilogue set-up for all of them.
constexpr int TC_SUB = 8;
__host__ __device__ inline int tc_fwd_sub(const Prob& p, int nb) {
    const int mt = (nb * p.Ho * p.Wo + TC_BM - 1) / TC_BM;
    // (any reduction length: what a descriptor costs the loaders - fetching it, the geometry set-up, the k table - is
    // about as much as five k-blocks of copies, so a problem with many row tiles shares it among TC_SUB of them)
#ifndef __CUDA_ARCH__
    static const bool only_short = getenv("SRN_SUB_K64") != nullptr;       // A/B: sub-tiles only for one-k-block problems
    if (only_short && tc_k_fwd(p) > TC_BK) return 1;
#endif
    if (mt < 8 * TC_SUB) return 1;
    // a descriptor should not carry much more than 64 k-blocks either: the tile table is dealt out by cost, and a
    // descriptor of 8 sub-tiles x 29 k-blocks left its CTA with 2.4 times the mean load of the launch
    const int nk = (tc_k_fwd(p) + TC_BK - 1) / TC_BK;
    const int sub = (TC_SUB * 8) / (nk < 1 ? 1 : nk);
    return sub < 1 ? 1 : (sub > TC_SUB ? TC_SUB : sub);
}
// Sub-tiles of such a problem whose reduction is at most 16 (32) elements are packed four (two) to a ring stage: row r of
// the stage's 128-byte rows holds the k chunks of row r of every packed sub-tile side by side, the MMA of sub-tile q
// reads them through the matrix descriptor advanced by q * Kpad elements (exactly like a K = 16 step), and the weight
// tile is loaded once per stage.  One barrier round trip then carries four sub-tiles: with one k-block per tile the
// pipeline is bound by how many bytes a CTA keeps in flight, not by bandwidth.
__host__ __device__ inline int tc_fwd_pack(const Prob& p, int nb, bool scalar_a) {
    if (scalar_a || tc_fwd_sub(p, nb) == 1) return 1;
    const int K = tc_k_fwd(p);
    return K <= 16 ? 4 : (K <= 32 ? 2 : 1);
}
// Forward problems with few output tiles and a long reduction (Dense on a flattened feature map: 750 rows, K up to
// 25 000) are split along k so that the reduction runs on many SMs; the partial sums meet in an fp32 scratch tile.
__host__ __device__ inline int tc_fwd_split_tiles(const Prob& p, int nb, int BN) {     // output tiles of a splittable problem, else 0
    const int mt = (nb * p.Ho * p.Wo + TC_BM - 1) / TC_BM, tn = (p.F + BN - 1) / BN, nk = (tc_k_fwd(p) + TC_BK - 1) / TC_BK;
    return (mt >= 8 * TC_SUB || mt * tn > 256 || nk < 16) ? 0 : mt * tn;
}
// tstar = the tile length (k-blocks) the launch this problem runs in aims at: about half the mean load of a CTA over all
// the buckets merged into that launch (group_tstar in srn_engine.cu), so that no tile unbalances the launch; every split
// costs an fp32 atomic pass over its output tile, so splits keep at least 8 k-blocks
__host__ __device__ inline int tc_fwd_split(const Prob& p, int nb, int BN, int tstar) {
    if (tstar <= 0 || tc_fwd_split_tiles(p, nb, BN) == 0) return 1;
    const int nk = (tc_k_fwd(p) + TC_BK - 1) / TC_BK;
    int s = (nk + tstar - 1) / tstar;
    if (s > nk / 8) s = nk / 8;
    if (s > 32) s = 32;
    return s < 1 ? 1 : s;
}
__host__ __device__ inline int tc_tiles_fwd(const Prob& p, int nb, int BN, int smax) {
    const int mt = (nb * p.Ho * p.Wo + TC_BM - 1) / TC_BM, sub = tc_fwd_sub(p, nb);
    return ((mt + sub - 1) / sub) * ((p.F + BN - 1) / BN) * tc_fwd_split(p, nb, BN, smax);
}
// The input gradient of a strided convolution is computed per residue class (ph, pw) of the input position: a pixel
// h = ph + sh h' only receives the taps ki = ph + sh ki', so each class is a stride-1 problem over Hc x Wc pixels with
// khp x kwp taps (a class with no tap still gets tiles: its pixels receive zeros).  Stride 1 is the single class (0, 0).
struct DgClass { int ph, pw, Hc, Wc, khp, kwp; };
__host__ __device__ inline DgClass tc_dgrad_class(const Prob& p, int cls) {
    DgClass d;
    d.ph = cls / p.sw; d.pw = cls - d.ph * p.sw;
    d.Hc = p.H > d.ph ? (p.H - d.ph + p.sh - 1) / p.sh : 0;
    d.Wc = p.W > d.pw ? (p.W - d.pw + p.sw - 1) / p.sw : 0;
    d.khp = p.kh > d.ph ? (p.kh - d.ph + p.sh - 1) / p.sh : 0;
    d.kwp = p.kw > d.pw ? (p.kw - d.pw + p.sw - 1) / p.sw : 0;
    return d;
}
__host__ __device__ inline int tc_tiles_dgrad_class(const Prob& p, const DgClass& d, int nb, int BN) {
    return ((nb * d.Hc * d.Wc + TC_BM - 1) / TC_BM) * ((p.C + BN - 1) / BN);
}
__host__ __device__ inline int tc_tiles_dgrad(const Prob& p, int nb, int BN) {
    int n = 0;
    for (int cls = 0; cls < p.sh * p.sw; ++cls) n += tc_tiles_dgrad_class(p, tc_dgrad_class(p, cls), nb, BN);
    return n;
}
__host__ __device__ inline int tc_tiles_wgrad(const Prob& p, int nb, int BN) {
    return ((tc_k_fwd(p) + TC_BM - 1) / TC_BM) * ((p.F + BN - 1) / BN) * ((nb * p.Ho * p.Wo + TC_WGRAD_ROWS - 1) / TC_WGRAD_ROWS);
}

// ------------------------------------------------------------------ PTX wrappers
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// 16-byte asynchronous copy; when !valid the destination is zero-filled (src-size 0) and src is not read, but it
// must still be a mapped address: callers pass offset 0 of the tensor then.  ptxas folds the size select into the
// zero-fill predicate operand of one LDGSTS.
__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, bool valid) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(dst), "l"(src), "r"(valid ? 16u : 0u) : "memory");
}
__device__ __forceinline__ void cp_async16_ca(uint32_t dst, const void* src, bool valid) {   // keeps the line in L1 (im2col reuse)
    asm volatile("cp.async.ca.shared.global [%0], [%1], 16, %2;\n" ::"r"(dst), "l"(src), "r"(valid ? 16u : 0u) : "memory");
}
// arrive on an mbarrier once all cp.async issued so far by this thread have landed (no wait in the thread)
__device__ __forceinline__ void cp_async_arrive_noinc(uint32_t bar) {
    asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];\n" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;\n" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;\n" ::"n"(N) : "memory"); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory"); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;\n" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];\n" ::"r"(bar) : "memory");
}
// TMA: the issuing thread arrives on the stage's "full" barrier with the byte count of its copies, the copies
// complete the transaction count when they land
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.expect_tx.relaxed.cta.shared::cta.b64 [%0], %1;\n" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const void* tmap, int c0, int c1, uint32_t bar) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];\n"
                 ::"r"(dst), "l"(tmap), "r"(bar), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void loaders_sync() { asm volatile("bar.sync 1, 256;\n" ::: "memory"); }   // the loader warps only
// try_wait suspends the thread until the phase completes or the time hint (ns) runs out, so a long hint means few
// polls: every poll is a shared-memory access that competes with the copies and the operand reads
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    const uint32_t hint = 20000u;
    asm volatile(
        "{\n"
        ".reg .pred P1;\n"
        "WAIT_LOOP:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1, %2;\n"
        "@P1 bra WAIT_DONE;\n"
        "bra WAIT_LOOP;\n"
        "WAIT_DONE:\n"
        "}\n" ::"r"(bar), "r"(parity), "r"(hint) : "memory");
}
// same, for waits that last a whole tile (epilogue, accumulator hand-back): back off so that the spinning warps
// do not take issue slots from the loaders
__device__ __forceinline__ void mbar_wait_relaxed(uint32_t bar, uint32_t parity) {
    uint32_t done = 0;
    while (true) {
        asm volatile(
            "{\n"
            ".reg .pred P1;\n"
            "mbarrier.try_wait.parity.shared::cta.b64 P1, [%1], %2;\n"
            "selp.u32 %0, 1, 0, P1;\n"
            "}\n" : "=r"(done) : "r"(bar), "r"(parity) : "memory");
        if (done) break;
        __nanosleep(200);
    }
}
// one lane of the (converged) warp
__device__ __forceinline__ bool elect_one_sync() {
    uint32_t pred = 0;
    asm volatile("{\n.reg .pred P1;\nelect.sync _|P1, 0xffffffff;\nselp.u32 %0, 1, 0, P1;\n}\n" : "=r"(pred));
    return pred != 0;
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

// shared-memory matrix descriptor, swizzle-128B (layout type 2 at bits 61-63): start >> 4 | LBO >> 4 << 16 |
// SBO >> 4 << 32 | version 1 << 46.  Canonical layouts in 16-byte units (cute/atom/mma_traits_sm100.hpp):
//   K-major   ((8, n), 2)      : ((8, SBO), 1)           rows of 128 B = 64 k, 8 rows per 1024-byte atom, LBO unused
//   MN-major  ((8, n), (8, k)) : ((1, LBO), (8, SBO))    rows of 128 B = 64 m/n of one k, 8 k per atom
// with the 16-byte chunk index XOR-ed with the row index inside the atom (address bits 4-6 ^= bits 7-9).
__device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
    return (uint64_t)((saddr & 0x3FFFFu) >> 4) | ((uint64_t)(lbo_bytes >> 4) << 16) | ((uint64_t)(sbo_bytes >> 4) << 32) | (1ull << 46) | (2ull << 61);
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

// same, raw bits into the first 16 of 32 registers (narrow tiles)
__device__ __forceinline__ void tmem_ld16_raw(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];\n"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr)
        : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;\n" ::: "memory");
}
// 32 consecutive fp32 columns of this thread's TMEM lane (one wait for the whole batch)
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];\n"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
          "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
          "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr)
        : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;\n" ::: "memory");
}
// epilogue traffic goes straight to the global window (tile descriptors hold generic pointers, which would compile
// to generic LD / ST): read-only 16-byte loads, 16- and 32-byte stores
__device__ __forceinline__ uint4 ldg16(const __half* p) { return __ldg(reinterpret_cast<const uint4*>(p)); }
__device__ __forceinline__ void stg16(__half* p, const uint4& v) {
    asm volatile("st.global.v4.b32 [%0], {%1, %2, %3, %4};\n" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ void stg32(__half* p, const uint4& a, const uint4& b) {      // p is 32-byte aligned
    asm volatile("st.global.v8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};\n" ::"l"(p), "r"(a.x), "r"(a.y), "r"(a.z), "r"(a.w),
                 "r"(b.x), "r"(b.y), "r"(b.z), "r"(b.w) : "memory");
}
// two fp32 accumulators -> packed fp16 (round to nearest even), low half = first
__device__ __forceinline__ __half2 pack_h2(uint32_t lo, uint32_t hi) {
    return __floats2half2_rn(__uint_as_float(lo), __uint_as_float(hi));
}

// byte offset of a 16-byte chunk inside an operand tile (tile bases are 1024-byte aligned)
__device__ __forceinline__ uint32_t sw_k_off(int row, int chunk) {    // K-major: row = m/n index, chunk = k chunk 0..7
    return (uint32_t)((row >> 3) * 1024 + (row & 7) * 128 + ((chunk ^ (row & 7)) << 4));
}
__device__ __forceinline__ uint32_t sw_mn_off(int u, int kk) {        // MN-major: u = m/n chunk, kk = k row 0..63; 64 m/n per 8192-byte group
    return (uint32_t)((u >> 3) * 8192 + (kk >> 3) * 1024 + (kk & 7) * 128 + (((u & 7) ^ (kk & 7)) << 4));
}

// ------------------------------------------------------------------ the kernel
// what the loader, the MMA issuer and the epilogue need to know about one tile (lives in a shared-memory ring)
// division by a runtime constant as multiply-high + shift (valid for numerators < 2^31); the loaders do
// several index decompositions per 16-byte chunk and a hardware-less integer division costs ~20 dependent instructions
struct FastDiv {
    uint32_t mul, shr;
    int d;
};
inline FastDiv make_fastdiv(int d) {
    FastDiv f; f.d = d; f.mul = 0; f.shr = 0;
    if (d > 1) {
        int lg = 31 - __builtin_clz((unsigned)d);
        if (d & (d - 1)) ++lg;
        const int p = 31 + lg;
        f.mul = (uint32_t)((((unsigned long long)1 << p) + (unsigned)d - 1) / (unsigned)d);
        f.shr = (uint32_t)(p - 32);
    }
    return f;
}
__device__ __forceinline__ int fdiv(int n, const FastDiv& f) { return f.d == 1 ? n : (int)(__umulhi((uint32_t)n, f.mul) >> f.shr); }

struct __align__(16) TcTile {
    Prob p;
    int m0, n0, nk, n_umma, r0, r1, Mrows, Ncols;
    FastDiv fd_rows;   // rows per sample of the M index: Ho*Wo (fwd, wgrad) or H*W (dgrad)
    FastDiv fd_w;      // Wo (fwd, wgrad) or W (dgrad)
    FastDiv fd_c;      // channel stride of the k index: Cs (fwd, wgrad) or Fs (dgrad)
    FastDiv fd_kw;
    int klast;         // K = 16 steps of the last k-block (1..4): the tail of the reduction is neither loaded nor multiplied
    int nsub;          // 128-row sub-tiles m0, m0 + 128, .. covered by this descriptor (forward only, else 1)
    // split-K (forward, few output tiles and a long reduction): this tile covers k-blocks kb0 .. kb0 + nk - 1 of nsplit
    // splits; partial sums meet in the fp32 scratch sk_acc [Mrows][Fs] (k_splitk_finish turns them into the output)
    int kb0, nsplit, sk_tile;
    int pack;          // forward sub-tiles per ring stage (tc_fwd_pack), else 1
    float* sk_acc;
    int* reserved;     // (was the arrival counter of the in-kernel split-K finish)
    int ph, pw, Hc, Wc, khp, kwp;     // dgrad: residue class of this tile (tc_dgrad_class)
    // Dense problems whose operands are plain row-major matrices are loaded by TMA (2-D tiled mode, swizzle-128B):
    // tm -> two 128-byte tensor maps (A operand, B operand) of this pass, nullptr = cp.async gather path
    const void* tm;
    unsigned char pad[TC_TILE_BYTES - sizeof(Prob) - 20 * sizeof(int) - 4 * sizeof(FastDiv) - 3 * sizeof(void*)];
};
static_assert(sizeof(TcTile) == TC_TILE_BYTES, "TcTile is one table entry / ring slot");

// index in the table of the j-th tile of CTA c (G CTAs): rounds alternate direction so that costs even out
__host__ __device__ inline int tc_tile_index(int c, int j, int G) { return j * G + ((j & 1) ? G - 1 - c : c); }

inline void tc_describe_tile(TcTile& t, int tile, int nb, int BN, int MODE, bool SCALAR_A, int smax) {
    const Prob& p = t.p;
    const int HoWo = p.Ho * p.Wo;
    const int KK = p.kh * p.kw;
    const int Kf = SCALAR_A ? KK : KK * p.Cs;
    t.tm = nullptr;
    t.pack = 1;
    t.r0 = 0; t.r1 = 0; t.nsub = 1; t.kb0 = 0; t.nsplit = 1; t.sk_tile = 0; t.sk_acc = nullptr; t.reserved = nullptr;
    int Kend = 0;                                           // end of this tile's part of the reduction (elements)
    if (MODE == TC_FWD) {
        t.Mrows = nb * HoWo; t.Ncols = p.F;
        const int tn = (t.Ncols + BN - 1) / BN, sub = tc_fwd_sub(p, nb), mt = (t.Mrows + TC_BM - 1) / TC_BM;
        const int per_split = ((mt + sub - 1) / sub) * tn, split = tile / per_split;
        tile -= split * per_split;
        t.m0 = (tile / tn) * sub * TC_BM; t.n0 = (tile % tn) * BN;
        t.nsub = min(sub, mt - (tile / tn) * sub);
        t.pack = tc_fwd_pack(p, nb, SCALAR_A);
        t.nk = (Kf + TC_BK - 1) / TC_BK;
        t.nsplit = SCALAR_A ? 1 : tc_fwd_split(p, nb, BN, smax);
        if (t.nsplit > 1) {
            const int base = t.nk / t.nsplit, rem = t.nk % t.nsplit;
            t.kb0 = split * base + min(split, rem);
            t.nk = base + (split < rem ? 1 : 0);
            t.sk_tile = tile;
        }
        Kend = min(Kf, (t.kb0 + t.nk) * TC_BK);
    } else if (MODE == TC_DGRAD) {
        DgClass d = tc_dgrad_class(p, 0);
        for (int cls = 0; cls < p.sh * p.sw; ++cls) {
            d = tc_dgrad_class(p, cls);
            const int n = tc_tiles_dgrad_class(p, d, nb, BN);
            if (tile < n) break;
            tile -= n;
        }
        t.ph = d.ph; t.pw = d.pw; t.Hc = d.Hc; t.Wc = d.Wc; t.khp = d.khp; t.kwp = d.kwp;
        t.Mrows = nb * d.Hc * d.Wc; t.Ncols = p.C;
        const int tn = (t.Ncols + BN - 1) / BN;
        t.m0 = (tile / tn) * TC_BM; t.n0 = (tile % tn) * BN;
        t.nk = max(1, (d.khp * d.kwp * p.Fs + TC_BK - 1) / TC_BK);
    } else {
        t.Mrows = Kf; t.Ncols = p.F;                      // output rows are kk = (i, j, padded c)
        const int tm = (t.Mrows + TC_BM - 1) / TC_BM, tn = (t.Ncols + BN - 1) / BN;
        const int split = tile / (tm * tn);
        tile -= split * tm * tn;
        t.m0 = (tile / tn) * TC_BM; t.n0 = (tile % tn) * BN;
        t.r0 = split * TC_WGRAD_ROWS; t.r1 = min(nb * HoWo, t.r0 + TC_WGRAD_ROWS);
        t.nk = (t.r1 - t.r0 + TC_BK - 1) / TC_BK;
    }
    {
        const int Ktot = MODE == TC_FWD ? Kend - t.kb0 * TC_BK : (MODE == TC_DGRAD ? t.khp * t.kwp * p.Fs : t.r1 - t.r0);
        t.klast = max(1, (Ktot - (t.nk - 1) * TC_BK + 15) / 16);
    }
    t.n_umma = min(BN, ((t.Ncols - t.n0 + 15) / 16) * 16);
    t.fd_rows = make_fastdiv(MODE == TC_DGRAD ? max(1, t.Hc * t.Wc) : HoWo);
    t.fd_w = make_fastdiv(MODE == TC_DGRAD ? max(1, t.Wc) : p.Wo);
    t.fd_c = make_fastdiv(MODE == TC_DGRAD ? p.Fs : p.Cs);
    t.fd_kw = make_fastdiv(MODE == TC_DGRAD ? max(1, t.kwp) : p.kw);
}

// What the k index of a 16-byte chunk means for a problem: (window position, channel) of the padded reduction space.
struct KGeo { int cs, kw, ww, C, klim, bias; FastDiv fdc, fdkw; bool dense; };
// Table entry of chunk `chunk` (k = 8 * chunk .. + 7).  y: bit 31 = inside the reduction; x = byte offset to add to the
// row's base.  FWD: x = ((ki * W + kj) * Cs + c) * 2, y bits 0-19 = weight row ij * C + c, bits 20-23 = real channels in
// the chunk.  DGRAD: x = (f - (ki * Wo + kj) * Fs + bias) * 2, y bits 0-7 = ki, 8-15 = kj, 16-30 = f / 8.
template <int MODE>
__device__ __forceinline__ int2 tc_k_entry(const KGeo& g, int chunk) {
    const int k = chunk * 8;
    if (k >= g.klim) return make_int2(0, 0);
    int ij = 0, cp = k, ki = 0, kj = 0;
    if (!g.dense) { ij = fdiv(k, g.fdc); cp = k - ij * g.cs; ki = fdiv(ij, g.fdkw); kj = ij - ki * g.kw; }
    if (MODE == TC_DGRAD)
        return make_int2(2 * (cp - (ki * g.ww + kj) * g.cs + g.bias),
                         (int)(0x80000000u | (uint32_t)ki | ((uint32_t)kj << 8) | ((uint32_t)(cp >> 3) << 16)));
    const int nvalid = min(8, max(0, g.C - cp));
    return make_int2(2 * ((ki * g.ww + kj) * g.cs + cp), (int)(0x80000000u | (uint32_t)(ij * g.C + cp) | ((uint32_t)nvalid << 20)));
}

// Split-K forward problems: the GEMM kernel adds the partial sums of every split into an fp32 scratch matrix [rows][Fs];
// this kernel (launched behind it on the same stream) adds the bias, applies the activation, stores fp16 and leaves the
// scratch zeroed for the next launch.  One thread per 8 columns, consecutive threads on consecutive chunks of a row.
// (As the last-arriving epilogue warp's job, this pass was a chain of dependent L2 round trips per thread that outlasted
// the tile's whole reduction, and its registers counted against the 72 of the main loops.)
struct SplitFin { float* acc; __half* out; const __half* bias; int rows, Fs, F, act, HoWo, out_bs, blk0; };
static __global__ void __launch_bounds__(256) k_splitk_finish(const SplitFin* __restrict__ fins, int nfin) {
    int i = 0;
    while (i + 1 < nfin && (int)blockIdx.x >= fins[i + 1].blk0) ++i;
    const SplitFin f = fins[i];
    const int cpr = f.Fs >> 3;                                     // 8-column chunks per row
    const long long chunk = (long long)(blockIdx.x - f.blk0) * 256 + threadIdx.x;
    if (chunk >= (long long)f.rows * cpr) return;
    const int row = (int)(chunk / cpr), n = (int)(chunk - (long long)row * cpr) * 8;
    float* a = f.acc + (long long)row * f.Fs + n;
    const float4 x0 = __ldcg(reinterpret_cast<const float4*>(a)), x1 = __ldcg(reinterpret_cast<const float4*>(a + 4));
    const float v[8] = {x0.x, x0.y, x0.z, x0.w, x1.x, x1.y, x1.z, x1.w};
    const Half8 bv = ld8(f.bias + n);                              // bias arrays are padded to 8 elements in the arena
    Half8 o;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        float z = 0.f;
        if (n + j < f.F) z = act_forward(rh(rh(v[j]) + __half2float(bv.v[j])), f.act);
        o.v[j] = __float2half_rn(z);
    }
    const int b = row / f.HoWo, rem = row - b * f.HoWo;
    st8(f.out + (long long)b * f.out_bs + (long long)rem * f.Fs + n, o);
    __stcg(reinterpret_cast<float4*>(a), make_float4(0.f, 0.f, 0.f, 0.f));
    __stcg(reinterpret_cast<float4*>(a + 4), make_float4(0.f, 0.f, 0.f, 0.f));
}

template <int MODE, bool SCALAR_A, int BN>
static __global__ void __launch_bounds__(TC_THREADS, (BN <= 128 ? 2 : 1))
k_gemm_tc(const TcTile* __restrict__ table, int ntiles, int nb, Inputs I) {
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const uint32_t sraw = smem_u32(smem_raw);
    const uint32_t sbase = sraw;
    if (sraw & 1023u) __trap();                           // swizzle-128B atoms need 1024-byte alignment (the dynamic window starts on one)
    unsigned char* sgen = smem_raw + (sbase - sraw);      // generic pointer to the same place as sbase
    constexpr uint32_t B_BYTES = (uint32_t)tc_b_bytes(BN);
    constexpr int STAGES = tc_stages(BN);
    const uint32_t sA = sbase, sB = sbase + STAGES * TC_A_BYTES;
    const uint32_t sRing = sB + STAGES * B_BYTES;         // TC_RING tile descriptors
    const uint32_t sFull = sRing + TC_RING * TC_TILE_BYTES;         // barriers: full[S], empty[S], tmem_full[2], tmem_empty[2]
    const uint32_t sEmpty = sFull + TC_MAX_STAGES * 8;
    const uint32_t sTFull = sEmpty + TC_MAX_STAGES * 8;
    const uint32_t sTEmpty = sTFull + 8 * 8;
    const uint32_t sSlot = sTEmpty + 8 * 8;
    const uint32_t sLut = sFull + 384;                    // two k-chunk tables of TC_LUT entries (this tile, next tile)
    const uint32_t sBias = sLut + 2 * TC_LUT * 8;         // forward: bias of the current tile's columns, per epilogue warp
    TcTile* ring = reinterpret_cast<TcTile*>(sgen + (sRing - sbase));
    volatile uint32_t* slot_ptr = reinterpret_cast<volatile uint32_t*>(sgen + (sSlot - sbase));
    int2* lut = reinterpret_cast<int2*>(sgen + (sLut - sbase));
    constexpr int acc_cols = tc_acc_cols(BN), tmem_cols = tc_tmem_cols(BN), NACC = tc_num_acc(BN);

    const int G = gridDim.x, cta = blockIdx.x;
    int my_tiles = ntiles / G;                                      // full rounds, plus maybe one of the last round
    if (tc_tile_index(cta, my_tiles, G) < ntiles) ++my_tiles;
    // tile descriptors travel global -> registers (one tile early) -> ring slot j % TC_RING, 16 bytes per thread
    // of the first half-warp; the load of the next one is in flight while the current tile is being gathered
    uint4 dreg = make_uint4(0, 0, 0, 0);
    auto desc_load = [&](int j) {
        if (j < my_tiles && tid < TC_TILE_BYTES / 16)
            dreg = __ldg(reinterpret_cast<const uint4*>(table + tc_tile_index(cta, j, G)) + tid);
    };
    auto desc_store = [&](int j) {
        if (j < my_tiles && tid < TC_TILE_BYTES / 16) reinterpret_cast<uint4*>(ring + (j % TC_RING))[tid] = dreg;
    };

    if (tid == 0) {
        for (int i = 0; i < STAGES; ++i) { mbar_init(sFull + i * 8, TC_LOADERS); mbar_init(sEmpty + i * 8, 1); }
        for (int i = 0; i < NACC; ++i) { mbar_init(sTFull + i * 8, 1); mbar_init(sTEmpty + i * 8, 4); }
        asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory");
    }
    if (warp == TC_LW) tmem_alloc(sSlot, (uint32_t)tmem_cols);
    for (int j = 0; j < TC_DESC_AHEAD; ++j) { desc_load(j); desc_store(j); }
    desc_load(TC_DESC_AHEAD);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    // (every role reads the TMEM base from the slot when it needs it: held in a register across the roles it was spilled
    // to local memory and reloaded once per accumulator hand-over)

    if (warp < TC_LW) {
        // =================================================================== loaders (8 warps)
        // Operand tiles are stored in the swizzle-128B canonical layouts.  One warp-wide copy instruction moves
        // 4 rows x 8 chunks: the 8 lanes of a quarter warp read 128 contiguous bytes of one row (one L1 line when
        // the row is contiguous in memory - a cached 16-byte LDGSTS costs one L1 wavefront per line a quarter warp
        // touches, tools/micro/ldgsts_wavefronts.cu) and write the 8 chunk slots of one 128-byte shared-memory
        // row (all 32 banks).  A thread therefore owns one chunk column (lane & 7) and rows 4 * warp + lane / 8
        // + 32 i, so what depends only on k (window offset, weight row, validity of a 16-byte k chunk) is read once
        // per k-block from a per-tile lookup table in shared memory that the loaders fill one tile ahead.  All
        // gather offsets are 32-bit byte offsets inside one tensor; an out-of-range chunk gets offset 0 and a
        // zero-filling copy.
        const int ch = lane & 7, sub = lane >> 3, rlo = warp * 4 + sub;          // chunk column, row inside a block of 32
        const uint32_t k_dst = sw_k_off(rlo, ch);                                 // K-major tile: rows rlo + 32 i at + 4096 i
        const char* l_in = nullptr; int l_in_bs = 0;
        const char* t_w = nullptr; const char* t_gout = nullptr;
        uint32_t a_off[4] = {0, 0, 0, 0}; bool a_ok[4] = {false, false, false, false};   // FWD: this thread's (up to) four im2col rows
        uint32_t a_dst[2] = {0, 0};
        // Short reductions (first layers: K = 16 or 32) use 2 or 4 chunks per row, so a copy instruction covers 16 or 8
        // rows and no lane copies zeros: t_cs = log2(chunks per row), t_items = rows per thread, t_ch = this thread's chunk
        int t_cs = 3, t_items = 4, t_ch = ch, t_ga = 2;
        int a_h[4] = {0, 0, 0, 0}, a_w[4] = {0, 0, 0, 0}, a_base[4] = {0, 0, 0, 0};       // DGRAD: input position of its rows, biased offset
        int w_koff[2] = {0, 0}; bool w_kok[2] = {false, false};    // WGRAD (vector): offsets of this thread's two kk chunks inside the window
        int t_ldw = 0, t_Cs = 1, t_Fs = 1, t_kw = 1, t_W = 1, t_Wo = 1, t_Ho = 1, t_sh = 1, t_sw = 1;
        int t_Kf = 0, t_n0 = 0, t_out_bs = 0, t_HoWo = 1, t_r0 = 0, t_r1 = 0, t_m0 = 0, t_Cldw = 0, t_C8 = 0, t_C = 0;
        bool t_lut = false, t_s1 = true, t_flat = false, t_fast = false;
        int t_tap0 = 0, t_tapi = 0, t_tapj = 0, t_dW = 1;        // DGRAD: weight-row offsets of the class's taps, class width
        KGeo geo{};
        FastDiv t_fdrows{}, t_fdw{};
        int lj = 0, lnk = 0, lsub = 1, t_Mrows = 0, lkb0 = 0, lpack = 1;
        const char* t_tm = nullptr;
        auto geo_of = [&](const TcTile& T) {
            const Prob& p = T.p;
            KGeo g;
            const int KK = p.kh * p.kw;
            g.dense = KK == 1; g.kw = p.kw; g.C = p.C; g.fdc = T.fd_c; g.fdkw = T.fd_kw;
            if (MODE == TC_DGRAD) {      // taps of the tile's residue class
                g.dense = T.khp * T.kwp == 1; g.kw = max(1, T.kwp);
                g.cs = p.Fs; g.ww = p.Wo; g.klim = T.khp * T.kwp * p.Fs; g.bias = max(0, (T.khp - 1) * p.Wo + (T.kwp - 1)) * p.Fs;
            }
            else { g.cs = p.Cs; g.ww = p.W; g.klim = SCALAR_A ? KK : KK * p.Cs; g.bias = 0; }
            return g;
        };
        auto fill_lut = [&](int j) {       // table of tile j -> buffer j & 1 (FWD and DGRAD only)
            if (MODE == TC_WGRAD || SCALAR_A || j >= my_tiles) return;
            const TcTile& T = ring[j % TC_RING];
            const int chunks = (T.kb0 + T.nk) * 8;
            if (chunks > TC_LUT || (MODE == TC_DGRAD ? T.khp * T.kwp : T.p.kh * T.p.kw) == 1) return;      // Dense entries are cheaper to compute than to look up
            const KGeo g = geo_of(T);
            int2* L = lut + (j & 1) * TC_LUT;
            for (int c = tid; c < chunks; c += TC_LOADERS) L[c] = tc_k_entry<MODE>(g, c);
        };
        auto loader_enter_tile = [&]() {
            loaders_sync();                 // descriptor lj and table lj were stored before the previous loaders_sync
            const TcTile& T = ring[lj % TC_RING];
            desc_store(lj + TC_DESC_AHEAD);
            desc_load(lj + TC_DESC_AHEAD + 1);
            fill_lut(lj + 1);
            const Prob& p = T.p;
            { const __half* in = resolve_in(p, I, l_in_bs); l_in = reinterpret_cast<const char*>(in); }
            lnk = T.nk; lsub = T.nsub; t_Mrows = T.Mrows; lkb0 = T.kb0; t_tm = reinterpret_cast<const char*>(T.tm);
            lpack = MODE == TC_FWD ? T.pack : 1;
            t_w = reinterpret_cast<const char*>(p.w); t_gout = reinterpret_cast<const char*>(p.gout);
            t_ldw = p.ldw; t_Cs = p.Cs; t_Fs = p.Fs; t_kw = p.kw; t_W = p.W;
            t_Wo = p.Wo; t_Ho = p.Ho; t_sh = p.sh; t_sw = p.sw; t_n0 = T.n0; t_out_bs = p.out_bs; t_HoWo = p.Ho * p.Wo;
            t_r0 = T.r0; t_r1 = T.r1; t_m0 = T.m0;
            t_Cldw = p.C * p.ldw; t_C8 = p.C - T.n0; t_C = p.C;
            const int KK = p.kh * p.kw;
            t_Kf = SCALAR_A ? KK : KK * p.Cs;
            t_lut = (T.kb0 + T.nk) * 8 <= TC_LUT && (MODE == TC_DGRAD ? T.khp * T.kwp : KK) != 1;
            t_s1 = true;
            t_flat = MODE == TC_FWD && KK == 1 && p.sh == 1 && p.sw == 1 && l_in_bs == p.H * p.W * p.Cs;
            t_tap0 = (T.ph * p.kw + T.pw) * t_Cldw; t_tapi = p.sh * p.kw * t_Cldw; t_tapj = p.sw * t_Cldw;
            geo = geo_of(T);
            t_fdrows = T.fd_rows; t_fdw = T.fd_w;
            const int HW = MODE == TC_DGRAD ? T.Hc * T.Wc : p.H * p.W;
            t_dW = T.Wc;
            if (MODE == TC_FWD) {
                t_cs = (T.nk == 1 && t_Kf <= 32 && T.pack == 1) ? (t_Kf <= 16 ? 1 : 2) : 3;      // packed stages use whole 128-byte rows
                t_items = 1 << (t_cs - 1);
                t_ch = lane & ((1 << t_cs) - 1);
                const int rpw = 32 >> t_cs, subr = lane >> t_cs;
#pragma unroll
                for (int i = 0; i < 2; ++i) a_dst[i] = sw_k_off((i * rpw * 8 + warp * rpw + subr) & 127, t_ch);
            } else if (MODE == TC_DGRAD) {
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const int m = T.m0 + i * 32 + rlo;
                    a_h[i] = -0x40000000; a_w[i] = 0; a_base[i] = 0;        // out of range: every window test fails
                    if (m < T.Mrows) {
                        const int b = fdiv(m, t_fdrows);
                        const int rem = m - b * HW;
                        a_h[i] = fdiv(rem, t_fdw); a_w[i] = rem - a_h[i] * t_dW;
                        // stride 1: source element = b * out_bs + ((h - ki) * Wo + (w - kj)) * Fs + kf; the table holds
                        // kf - (ki * Wo + kj) * Fs + bias >= 0
                        a_base[i] = 2 * (b * t_out_bs + (a_h[i] * t_Wo + a_w[i]) * t_Fs - geo.bias);
                    }
                }
            } else {
                // valid kk chunks of this tile: 2 or 4 (first layers) use the narrow lane mappings, up to 8 one group, else two
                const int mu = min(16, (T.Mrows - T.m0 + 7) >> 3);
                t_cs = mu <= 2 ? 1 : (mu <= 4 ? 2 : 3);
                t_ga = mu > 8 ? 2 : 1;
                t_ch = lane & ((1 << t_cs) - 1);
                if (!SCALAR_A) {
#pragma unroll
                    for (int g = 0; g < 2; ++g) {
                        const int kk = T.m0 + (g * 8 + t_ch) * 8;
                        w_kok[g] = kk < T.Mrows;
                        const int kc = w_kok[g] ? kk : 0;
                        const int ij = fdiv(kc, geo.fdc), cp = kc - ij * t_Cs;
                        const int i = fdiv(ij, geo.fdkw), j = ij - i * t_kw;
                        w_koff[g] = (i * t_W + j) * t_Cs + cp;
                    }
                }
            }
            t_fast = MODE == TC_FWD ? (!SCALAR_A && t_lut && t_cs == 3) : (MODE == TC_DGRAD ? t_lut : (!SCALAR_A && t_cs == 3));
        };
        auto fwd_rows = [&](int m0) {        // FWD: this thread's rows of the sub-tile starting at row m0
            const int rpw = 32 >> t_cs, subr = lane >> t_cs;
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                if (i >= t_items) { a_ok[i] = false; continue; }
                const int m = m0 + i * rpw * 8 + warp * rpw + subr;
                a_ok[i] = m < t_Mrows;
                const int mc = a_ok[i] ? m : 0;
                if (t_flat) { a_off[i] = 2u * (uint32_t)mc * (uint32_t)t_Cs; continue; }      // rows back to back: no decomposition
                const int b = fdiv(mc, t_fdrows), rem = mc - b * t_HoWo;
                const int ho = fdiv(rem, t_fdw), wo = rem - ho * t_Wo;
                a_off[i] = 2u * ((uint32_t)b * (uint32_t)l_in_bs + (uint32_t)(((ho * t_sh) * t_W + wo * t_sw) * t_Cs));
            }
        };
        // FAST = the tile's k table is in shared memory and the tile uses the standard 4 rows x 8 chunks mapping: the k
        // loop of such tiles (all convolutions of the benchmark) is compiled without the fallback branches
        auto k_entry = [&](int chunk, auto FAST) -> int2 {
            if (decltype(FAST)::value || t_lut) return lut[(lj & 1) * TC_LUT + chunk];
            return tc_k_entry<MODE>(geo, chunk);
        };
        // MN-major B tile (k rows x n chunks): BN = 32 has four chunks per row, so a quarter warp takes rows r and
        // r + 4 (their swizzled slots never share a bank); wider tiles take whole 128-byte row segments
        auto load_b_mn = [&](uint32_t bS, int krows, auto&& row_src) {   // rows >= krows are not read by the MMAs of this k-block
            if (BN == 32) {
                const int q = lane >> 2, kl = warp * 8 + ((q & 1) * 4 + (q >> 1)), u = lane & 3;
                if (kl < krows) {
                    int left; const char* src = row_src(kl, left);
                    const bool ok = u * 16 < left;
                    cp_async16(bS + sw_mn_off(u, kl), src + (ok ? u * 16 : 0), ok);
                }
            } else {
                const int G = BN >> 6;
                const uint32_t d0 = bS + sw_mn_off(ch, rlo);           // row rlo + 32 h at + 4096 h, group g at + 8192 g
                if (krows >= 64) {
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        int left; const char* src = row_src(h * 32 + rlo, left);
#pragma unroll 4
                        for (int g = 0; g < G; ++g) {
                            const int ub = g * 128 + ch * 16;
                            const bool ok = ub < left;
                            cp_async16(d0 + h * 4096 + g * 8192, src + (ok ? ub : 0), ok);
                        }
                    }
                } else if (rlo < krows) {                                 // short reductions: krows is 16, 32 or 48
                    int left; const char* src = row_src(rlo, left);
                    for (int g = 0; g < G; ++g) {
                        const int ub = g * 128 + ch * 16;
                        const bool ok = ub < left;
                        cp_async16(d0 + g * 8192, src + (ok ? ub : 0), ok);
                    }
                    if (krows > 32 && rlo + 32 < krows) {
                        src = row_src(rlo + 32, left);
                        for (int g = 0; g < G; ++g) {
                            const int ub = g * 128 + ch * 16;
                            const bool ok = ub < left;
                            cp_async16(d0 + 4096 + g * 8192, src + (ok ? ub : 0), ok);
                        }
                    }
                }
            }
        };
        auto load_stage = [&](int stage, int kb, auto FAST) {
            constexpr bool fast = decltype(FAST)::value;
            const uint32_t aS = sA + stage * TC_A_BYTES, bS = sB + stage * B_BYTES;
            if (MODE == TC_FWD) {
                // A, K-major: four rows, one k chunk
                if (!SCALAR_A) {
                    const int2 e = k_entry(kb * 8 + (fast ? ch : t_ch), FAST);
                    if (fast || t_cs == 3) {
#pragma unroll
                        for (int i = 0; i < 4; ++i) {
                            const bool ok = a_ok[i] && e.y < 0;
                            cp_async16_ca(aS + k_dst + i * 4096, l_in + (ok ? a_off[i] + (uint32_t)e.x : 0u), ok);
                        }
                    } else {
#pragma unroll
                        for (int i = 0; i < 2; ++i) {
                            if (i < t_items) {
                                const bool ok = a_ok[i] && e.y < 0;
                                cp_async16_ca(aS + a_dst[i], l_in + (ok ? a_off[i] + (uint32_t)e.x : 0u), ok);
                            }
                        }
                    }
                } else {
                    // first layer on a one-channel model input with per-individual inputs: element-wise gather
                    const __half* in_h = reinterpret_cast<const __half*>(l_in);
                    const int k = kb * TC_BK + t_ch * 8;
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        if (i < t_items) {
                            const __half* row = in_h + (a_off[i] >> 1);
                            Half8 v;
                            int ki = fdiv(k, geo.fdkw), kj = k - ki * t_kw;
#pragma unroll
                            for (int e = 0; e < 8; ++e) {
                                v.v[e] = (a_ok[i] && k + e < t_Kf) ? row[ki * t_W + kj] : __float2half_rn(0.f);
                                if (++kj == t_kw) { kj = 0; ++ki; }
                            }
                            *reinterpret_cast<Half8*>(sgen + (aS - sbase) + (t_cs == 3 ? k_dst + i * 4096 : a_dst[i & 1])) = v;
                        }
                    }
                }
                // B, MN-major: k rows kb*64 .. +63, n chunks from n0.  W rows are [(ij, c)][ldw]; pad channels have no row.
                load_b_mn(bS, min(64, (t_Kf - kb * TC_BK + 15) & ~15), [&](int kl, int& left) -> const char* {
                    int wrow; bool kok;
                    if (SCALAR_A) { wrow = kb * TC_BK + kl; kok = wrow < t_Kf; }
                    else { const int2 eb = k_entry(kb * 8 + (kl >> 3), FAST); wrow = (eb.y & 0xFFFFF) + (kl & 7); kok = (kl & 7) < ((eb.y >> 20) & 15); }
                    left = kok ? 2 * (t_ldw - t_n0) : 0;
                    return t_w + (kok ? 2u * (uint32_t)(wrow * t_ldw + t_n0) : 0u);
                });
            } else if (MODE == TC_DGRAD) {
                // k = (ij, f) over the padded filters: one chunk per thread and k-block, four rows
                const int2 e = k_entry(kb * 8 + ch, FAST);
                const int ki = e.y & 255, kj = (e.y >> 8) & 255, kf = ((e.y >> 16) & 0x7FFF) * 8;
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const int hh = a_h[i] - ki, ww = a_w[i] - kj;
                    const bool ok = e.y < 0 && (unsigned)hh < (unsigned)t_Ho && (unsigned)ww < (unsigned)t_Wo;
                    const uint32_t off = (uint32_t)(a_base[i] + e.x);
                    cp_async16_ca(aS + k_dst + i * 4096, t_gout + (ok ? off : 0u), ok);
                }
                // B, K-major: rows n = input channel c, k = (ij, f) contiguous in f
                const uint32_t wb = 2u * (uint32_t)(t_tap0 + ki * t_tapi + kj * t_tapj + kf + t_n0 * t_ldw);
#pragma unroll 4
                for (int i = 0; i < (BN >> 5); ++i) {
                    const int c = i * 32 + rlo;
                    const bool ok = c < t_C8 && e.y < 0;
                    cp_async16(bS + k_dst + i * 4096, t_w + (ok ? wb + 2u * (uint32_t)(c * t_ldw) : 0u), ok);
                }
            } else {
                // reduction rows r0 + kb*64 .. +63; A, MN-major (rows x kk chunks); B, MN-major: dOut rows, f chunks from n0
                auto in_row_of = [&](int kl, bool& rok, int& b, int& rem) -> const __half* {
                    const int r = t_r0 + kb * TC_BK + kl;
                    rok = r < t_r1;
                    const int rc = rok ? r : t_r0;
                    b = fdiv(rc, t_fdrows); rem = rc - b * t_HoWo;
                    const int ho = fdiv(rem, t_fdw), wo = rem - ho * t_Wo;
                    return reinterpret_cast<const __half*>(l_in) + (long long)b * l_in_bs + ((long long)(ho * t_sh) * t_W + wo * t_sw) * t_Cs;
                };
                auto copy_a = [&](const __half* in_row, bool rok, int kl, int g) {
                    const int u = g * 8 + t_ch;
                    if (SCALAR_A) {
                        Half8 v;
                        const int kk = t_m0 + u * 8;
                        int ki = fdiv(kk, geo.fdkw), kj = kk - ki * t_kw;
#pragma unroll
                        for (int e = 0; e < 8; ++e) {
                            v.v[e] = (rok && kk + e < t_Kf) ? in_row[ki * t_W + kj] : __float2half_rn(0.f);
                            if (++kj == t_kw) { kj = 0; ++ki; }
                        }
                        *reinterpret_cast<Half8*>(sgen + (aS - sbase) + sw_mn_off(u, kl)) = v;
                    } else {
                        cp_async16_ca(aS + sw_mn_off(u, kl), in_row + w_koff[g], rok && w_kok[g]);
                    }
                };
                auto go_of = [&](int kl, int& left) -> const char* {
                    const int r = t_r0 + kb * TC_BK + kl;
                    const bool rok = r < t_r1;
                    const int rc = rok ? r : t_r0;
                    const int b = fdiv(rc, t_fdrows), rem = rc - b * t_HoWo;
                    left = rok ? 2 * (t_Fs - t_n0) : 0;
                    return t_gout + 2 * ((long long)b * t_out_bs + (long long)rem * t_Fs + t_n0);
                };
                if (fast || t_cs == 3) {
                    const char* go_row[2]; int go_left[2];
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        const int kl = h * 32 + rlo;
                        bool rok; int b, rem;
                        const __half* in_row = in_row_of(kl, rok, b, rem);
                        go_row[h] = t_gout + 2 * ((long long)b * t_out_bs + (long long)rem * t_Fs + t_n0);
                        go_left[h] = rok ? 2 * (t_Fs - t_n0) : 0;
                        copy_a(in_row, rok, kl, 0);
                        if (t_ga == 2) copy_a(in_row, rok, kl, 1);
                    }
                    if (BN == 32) load_b_mn(bS, 64, go_of);
                    else load_b_mn(bS, 64, [&](int kl, int& left) -> const char* { left = go_left[kl >> 5]; return go_row[kl >> 5]; });
                } else {
                    const int rpw = 32 >> t_cs, kl = warp * rpw + (lane >> t_cs);      // 16 or 8 rows per warp
                    if (kl < 64) {
                        bool rok; int b, rem;
                        const __half* in_row = in_row_of(kl, rok, b, rem);
                        copy_a(in_row, rok, kl, 0);
                    }
                    load_b_mn(bS, 64, go_of);
                }
            }
        };

        fill_lut(0);
        int sb = 0; uint32_t ph = 1;     // ring stage and the parity a fresh "empty" barrier passes immediately
        auto stage_issued = [&]() {
            cp_async_arrive_noinc(sFull + sb * 8);       // fires when this thread's copies for the stage have landed
            if (++sb == STAGES) { sb = 0; ph ^= 1u; }
        };
        for (lj = 0; lj < my_tiles; ++lj) {
            loader_enter_tile();
            if (MODE == TC_FWD && lpack > 1) {
                // packed short reductions: chunk column ch belongs to sub-tile s + ch / cps, k chunk ch % cps
                const int cps = 8 / lpack, q = ch / cps, kc = ch - q * cps;
                const bool kvalid = kc * 8 < t_Kf;
                for (int s = 0; s < lsub; s += lpack) {
                    fwd_rows(t_m0 + (s + q) * TC_BM);
                    const bool sub_ok = s + q < lsub && kvalid;
                    mbar_wait(sEmpty + sb * 8, ph);
                    const uint32_t aS = sA + sb * TC_A_BYTES, bS = sB + sb * B_BYTES;
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        const bool ok = sub_ok && a_ok[i];
                        cp_async16_ca(aS + k_dst + i * 4096, l_in + (ok ? a_off[i] + (uint32_t)(kc * 16) : 0u), ok);
                    }
                    load_b_mn(bS, (t_Kf + 15) & ~15, [&](int kl, int& left) -> const char* {
                        const bool kok = kl < t_C;
                        left = kok ? 2 * (t_ldw - t_n0) : 0;
                        return t_w + (kok ? 2u * (uint32_t)(kl * t_ldw + t_n0) : 0u);
                    });
                    stage_issued();
                }
                continue;
            }
            if (!t_tm) {
                // gathered tiles: the k loop is instantiated once per path so that it carries no per-k-block path tests
                auto k_loops = [&](auto FAST) {
                    for (int s = 0; s < lsub; ++s) {
                        if (MODE == TC_FWD) fwd_rows(t_m0 + s * TC_BM);
                        for (int kb = lkb0; kb < lkb0 + lnk; ++kb) {
                            mbar_wait(sEmpty + sb * 8, ph);
                            load_stage(sb, kb, FAST);
                            if (SCALAR_A) fence_proxy_async();       // the element-wise gather used ordinary shared-memory stores
                            stage_issued();
                        }
                    }
                };
                if (t_fast) k_loops(std::true_type{});
                else k_loops(std::false_type{});
                continue;
            }
            for (int s = 0; s < lsub; ++s) {
                for (int kb = lkb0; kb < lkb0 + lnk; ++kb) {
                    mbar_wait(sEmpty + sb * 8, ph);
                    {
                        // row-major operands: one thread issues the tile copies
                        const uint32_t bar = sFull + sb * 8;
                        if (tid == 0) {
                            const uint32_t aS = sA + sb * TC_A_BYTES, bS = sB + sb * B_BYTES;
                            constexpr int G = BN >= 64 ? BN / 64 : 1;
                            if (MODE == TC_FWD) {
                                mbar_arrive_expect_tx(bar, TC_A_BYTES + G * 8192);
                                tma_load_2d(aS, t_tm, kb * TC_BK, t_m0 + s * TC_BM, bar);
#pragma unroll
                                for (int g = 0; g < G; ++g) tma_load_2d(bS + g * 8192, t_tm + 128, t_n0 + g * 64, kb * TC_BK, bar);
                            } else if (MODE == TC_DGRAD) {
                                mbar_arrive_expect_tx(bar, TC_A_BYTES + BN * 128);
                                tma_load_2d(aS, t_tm, kb * TC_BK, t_m0, bar);
                                tma_load_2d(bS, t_tm + 128, kb * TC_BK, t_n0, bar);
                            } else {
                                // accumulator rows 64..127 are only loaded when the tile has them (their garbage is never stored)
                                mbar_arrive_expect_tx(bar, t_ga * 8192 + G * 8192);
                                const int r = t_r0 + kb * TC_BK;
                                tma_load_2d(aS, t_tm, t_m0, r, bar);
                                if (t_ga == 2) tma_load_2d(aS + 8192, t_tm, t_m0 + 64, r, bar);
#pragma unroll
                                for (int g = 0; g < G; ++g) tma_load_2d(bS + g * 8192, t_tm + 128, t_n0 + g * 64, r, bar);
                            }
                        } else {
                            mbar_arrive(bar);
                        }
                        if (++sb == STAGES) { sb = 0; ph ^= 1u; }
                    }
                }
            }
        }
        cp_async_wait_all();
    } else if (warp == TC_LW) {
        // =================================================================== MMA issuer
        // The whole warp walks the loop in lock step and one elected lane issues: every operand of the tensor-core
        // instructions is then a warp-uniform value that the compiler keeps in uniform registers.  Run under
        // `if (lane == 0)` the same loop cost ~145 instructions per k-block (a uniformity "waterfall" around every
        // UTCHMMA / UTCBAR), and with tiles of one to five k-blocks of a 128 x 32 tile (32 tensor-pipe cycles each)
        // that single thread's instruction stream bounded the whole kernel (ncu: the issuer warp never stalled).
        {
            const uint32_t tmem_u = __shfl_sync(0xffffffffu, *slot_ptr, 0);
            auto bcast = [](int v) { return __shfl_sync(0xffffffffu, v, 0); };
            int st = 0; uint32_t fph = 0;     // ring stage and the parity of its "full" barrier
            // K-major operands advance 32 bytes per K = 16 step inside the 128-byte rows, MN-major ones two 8-k atoms
            const uint64_t dA0 = MODE == TC_WGRAD ? umma_desc_sw128(sA, 8192, 1024) : umma_desc_sw128(sA, 16, 1024);
            const uint64_t dB0 = MODE == TC_DGRAD ? umma_desc_sw128(sB, 16, 1024) : umma_desc_sw128(sB, 8192, 1024);
            constexpr uint64_t a_step = MODE == TC_WGRAD ? (2048 >> 4) : (32 >> 4), b_step = MODE == TC_DGRAD ? (32 >> 4) : (2048 >> 4);
            int jj = 0;                       // accumulator hand-overs so far (one per sub-tile)
            for (int j = 0; j < my_tiles; ++j) {
                // descriptor j was stored by the loaders before they published the tile's first stage
                mbar_wait(sFull + st * 8, fph);
                const TcTile& T = ring[j % TC_RING];
                const int nk = bcast(T.nk), klast = bcast(T.klast), nsub = bcast(T.nsub);
                const int pack = MODE == TC_FWD ? bcast(T.pack) : 1;
                const uint32_t idesc = umma_idesc(MODE == TC_WGRAD ? 1 : 0, MODE == TC_DGRAD ? 0 : 1, bcast(T.n_umma));
                if (MODE == TC_FWD && pack > 1) {
                    // packed short reductions (tc_fwd_pack): one stage feeds `pack` sub-tiles, each into its own accumulator
                    const int kper = 4 / pack;             // K = 16 steps per sub-tile inside the stage
                    for (int s = 0; s < nsub; s += pack) {
                        if (s) mbar_wait(sFull + st * 8, fph);
                        fence_proxy_async();
                        tc_fence_after();
                        const uint64_t dA = dA0 + (uint64_t)((uint32_t)(st * TC_A_BYTES) >> 4), dB = dB0 + (uint64_t)(((uint32_t)st * B_BYTES) >> 4);
                        for (int q = 0; q < pack && s + q < nsub; ++q, ++jj) {
                            const int acc = jj % NACC;
                            if (jj >= NACC) { mbar_wait(sTEmpty + acc * 8, (uint32_t)((jj / NACC - 1) & 1)); tc_fence_after(); }
                            const uint32_t tacc = tmem_u + (uint32_t)(acc * acc_cols);
                            if (elect_one_sync()) {
                                umma_f16(tacc, dA + a_step * (uint64_t)(q * kper), dB, idesc, 0u);
                                if (klast > 1) umma_f16(tacc, dA + a_step * (uint64_t)(q * kper + 1), dB + b_step, idesc, 1u);
                                umma_commit(sTFull + acc * 8);
                            }
                            __syncwarp();
                        }
                        if (elect_one_sync()) umma_commit(sEmpty + st * 8);
                        __syncwarp();
                        if (++st == STAGES) { st = 0; fph ^= 1u; }
                    }
                    continue;
                }
                for (int s = 0; s < nsub; ++s, ++jj) {
                    const int acc = jj % NACC;
                    if (jj >= NACC) mbar_wait(sTEmpty + acc * 8, (uint32_t)((jj / NACC - 1) & 1));   // epilogue drained this accumulator
                    const uint32_t tacc = tmem_u + (uint32_t)(acc * acc_cols);
                    for (int kb = 0; kb < nk; ++kb) {
                        if (s | kb) mbar_wait(sFull + st * 8, fph);
                        fence_proxy_async();     // the stage was written through the generic proxy (cp.async), UMMA reads it through the async proxy
                        tc_fence_after();
                        const uint64_t dA = dA0 + (uint64_t)((uint32_t)(st * TC_A_BYTES) >> 4), dB = dB0 + (uint64_t)(((uint32_t)st * B_BYTES) >> 4);
                        const int nks = kb == nk - 1 ? klast : TC_BK / 16;
                        if (elect_one_sync()) {
                            umma_f16(tacc, dA, dB, idesc, kb ? 1u : 0u);
                            if (nks > 1) umma_f16(tacc, dA + a_step, dB + b_step, idesc, 1u);
                            if (nks > 2) umma_f16(tacc, dA + a_step * 2, dB + b_step * 2, idesc, 1u);
                            if (nks > 3) umma_f16(tacc, dA + a_step * 3, dB + b_step * 3, idesc, 1u);
                            umma_commit(sEmpty + st * 8);
                            if (kb == nk - 1) umma_commit(sTFull + acc * 8);
                        }
                        __syncwarp();
                        if (++st == STAGES) { st = 0; fph ^= 1u; }
                    }
                }
            }
        }
    } else {
        // =================================================================== epilogue (4 warps)
        const int ewarp = warp & 3;      // TMEM lanes 32 * ewarp .. + 31
        const uint4* wbias = reinterpret_cast<const uint4*>(sgen + (sBias - sbase)) + ewarp * (BN / 8);
        // Forward tiles without k-split.  (Per-tile values are read from the ring slot per sub-tile on purpose: hoisted, they
        // are spilled to local memory under the 72-register cap, and local loads miss the L1 that the cp.async traffic churns.)
        auto epilogue_fwd = [&](const TcTile& T, int sub, uint32_t tacc) {
            const Prob& p = T.p;
            const int e_m0 = T.m0, e_n0 = T.n0, e_Ncols = T.Ncols, e_numma = T.n_umma, e_Fs = p.Fs, e_act = p.act, e_Mrows = T.Mrows, e_HoWo = p.Ho * p.Wo;
            const int e_out_bs = p.out_bs; __half* const e_out = p.out; const bool e_flat = e_out_bs == e_HoWo * e_Fs;
            const int mrow = e_m0 + sub * TC_BM + ewarp * 32 + lane, n0 = e_n0, Ncols = e_Ncols, act = e_act;
            const uint32_t trow = tacc + ((uint32_t)(ewarp * 32) << 16);
            __half* orow = nullptr;
            if (mrow < e_Mrows) {
                if (e_flat) orow = e_out + (long long)mrow * e_Fs;      // samples back to back
                else {
                    const int b = fdiv(mrow, T.fd_rows), rem = mrow - b * e_HoWo;
                    orow = e_out + (long long)b * e_out_bs + (long long)rem * e_Fs;
                }
            }
            // 32 accumulator columns per TMEM read (16 when that is all the tile has).  Linear and ReLU layers round, add
            // the bias and clamp as packed halves (one rounding per TensorFlow op, as the float path does); sigmoid goes
            // through fp32.  Rows are stored 32 bytes at a time when aligned.
            for (int cb = 0; cb < e_numma; cb += 32) {
                const int left = Ncols - n0 - cb;                       // real columns from here on
                const int nch = min(4, (left + 7) >> 3), nfull = min(4, left >> 3);
                uint32_t r[32];
                if (e_numma - cb > 16) tmem_ld32(trow + cb, r);
                else tmem_ld16_raw(trow + cb, r);
                if (!orow) continue;
                __half* dst = orow + n0 + cb;
                if (act != 2) {
                    const __half2 zero = __float2half2_rn(0.f);
                    const bool wide = nfull == 4 && (reinterpret_cast<uintptr_t>(dst) & 31) == 0;
#pragma unroll
                    for (int h16 = 0; h16 < 2; ++h16) {
                        if (h16 * 2 >= nfull) break;
                        uint4 o[2];
#pragma unroll
                        for (int q = 0; q < 2; ++q) {
                            const int h8 = h16 * 2 + q;
                            const uint4 bq = wbias[(cb >> 3) + h8];
                            const __half2* b2 = reinterpret_cast<const __half2*>(&bq);
                            __half2* o2 = reinterpret_cast<__half2*>(&o[q]);
#pragma unroll
                            for (int j = 0; j < 4; ++j) {
                                __half2 z = __hadd2(pack_h2(r[h8 * 8 + 2 * j], r[h8 * 8 + 2 * j + 1]), b2[j]);
                                if (act == 1) z = __hmax2_nan(z, zero);
                                o2[j] = z;
                            }
                        }
                        if (wide || (nfull >= h16 * 2 + 2 && (reinterpret_cast<uintptr_t>(dst + h16 * 16) & 31) == 0)) stg32(dst + h16 * 16, o[0], o[1]);
                        else {
                            stg16(dst + h16 * 16, o[0]);
                            if (h16 * 2 + 1 < nfull) stg16(dst + h16 * 16 + 8, o[1]);
                        }
                    }
                    if (nch == nfull) continue;
                }
                // sigmoid layers, and the chunk that holds the last real columns when F is not a multiple of 8
#pragma unroll
                for (int h8 = 0; h8 < 4; ++h8) {
                    if (h8 >= nch || (act != 2 && h8 < nfull)) continue;
                    const int n = n0 + cb + h8 * 8;
                    const uint4 bq = wbias[(cb >> 3) + h8];
                    const Half8 bv = *reinterpret_cast<const Half8*>(&bq);
                    Half8 o;
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        float z = 0.f;
                        if (n + j < Ncols) z = act_forward(rh(rh(__uint_as_float(r[h8 * 8 + j])) + __half2float(bv.v[j])), act);
                        o.v[j] = __float2half_rn(z);
                    }
                    stg16(dst + h8 * 8, *reinterpret_cast<const uint4*>(&o));
                }
            }
        };
        uint32_t rel_bar = 0;      // "tmem_empty" barrier of the accumulator in hand; an epilogue that releases it early (all TMEM reads done) clears this
        auto release_acc = [&]() {
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(rel_bar);
            rel_bar = 0;
        };
        auto epilogue = [&](const TcTile& T, int sub, uint32_t tacc, int ewarp) {
            const Prob& p = T.p;
            const int Cs = p.Cs, Fs = p.Fs;
            const int mrow = T.m0 + sub * TC_BM + ewarp * 32 + lane, n0 = T.n0, Ncols = T.Ncols;
            const uint32_t trow = tacc + ((uint32_t)(ewarp * 32) << 16);
            if (MODE == TC_FWD) {
                const int HoWo = p.Ho * p.Wo;
                const __half* bias = p.bias; const int act = p.act;
                __half* orow = nullptr;
                if (mrow < T.Mrows) {
                    if (p.out_bs == HoWo * Fs) orow = p.out + (long long)mrow * Fs;      // samples back to back
                    else {
                        const int b = fdiv(mrow, T.fd_rows), rem = mrow - b * HoWo;
                        orow = p.out + (long long)b * p.out_bs + (long long)rem * Fs;
                    }
                }
                if (T.nsplit > 1) {
                    // split-K: add this part of the reduction into the fp32 scratch row
                    float* arow = mrow < T.Mrows ? T.sk_acc + (long long)mrow * Fs : nullptr;
                    for (int cb = 0; cb < T.n_umma; cb += 32) {
                        uint32_t r[32];
                        tmem_ld32(trow + cb, r);
                        if (!arow) continue;
#pragma unroll
                        for (int q = 0; q < 8; ++q) {
                            const int n = n0 + cb + q * 4;
                            if (n < Ncols && cb + q * 4 < T.n_umma)
                                atomicAdd(reinterpret_cast<float4*>(arow + n), make_float4(__uint_as_float(r[q * 4]), __uint_as_float(r[q * 4 + 1]),
                                                                                           __uint_as_float(r[q * 4 + 2]), __uint_as_float(r[q * 4 + 3])));
                        }
                    }
                    return;      // bias, activation and the fp16 store happen in k_splitk_finish, launched behind this kernel
                }
            } else if (MODE == TC_DGRAD) {
                __half* drow = nullptr;
                if (mrow < T.Mrows) {
                    const int b = fdiv(mrow, T.fd_rows), rem = mrow - b * (T.Hc * T.Wc);
                    const int hc = fdiv(rem, T.fd_w), wc = rem - hc * T.Wc;
                    drow = p.din + (long long)b * p.in_bs + (long long)((T.ph + p.sh * hc) * p.W + T.pw + p.sw * wc) * Cs;
                }
                const bool accum = p.din_acc != 0;
                for (int cb = 0; cb < T.n_umma; cb += 32) {
                    uint32_t r[32];
                    tmem_ld32(trow + cb, r);
                    if (!drow) continue;
                    Half8 prev[4];
                    if (accum) {
#pragma unroll
                        for (int h8 = 0; h8 < 4; ++h8)
                            if (n0 + cb + h8 * 8 < Ncols) prev[h8] = ld8(drow + n0 + cb + h8 * 8);
                    }
#pragma unroll
                    for (int h8 = 0; h8 < 4; ++h8) {
                        const int c = n0 + cb + h8 * 8;
                        if (c >= Ncols) break;
                        Half8 o;
                        __half2* o2 = reinterpret_cast<__half2*>(&o);
#pragma unroll
                        for (int j = 0; j < 4; ++j) {
                            __half2 z = pack_h2(r[h8 * 8 + 2 * j], r[h8 * 8 + 2 * j + 1]);
                            if (accum) z = __hadd2(z, reinterpret_cast<const __half2*>(&prev[h8])[j]);
                            o2[j] = z;
                        }
                        if (c + 8 > Ncols) {
#pragma unroll
                            for (int j = 0; j < 8; ++j) if (c + j >= Ncols) o.v[j] = __float2half_rn(0.f);
                        }
                        st8(drow + c, o);
                    }
                }
            } else {
                // accumulator row = kk = (ij, padded c): rows of pad channels have no weight row
                float* grow = nullptr;
                if (mrow < T.Mrows) {
                    if (SCALAR_A) grow = p.gw + (long long)mrow * p.ldw;
                    else {
                        const int ij = fdiv(mrow, T.fd_c), cp = mrow - ij * Cs;
                        if (cp < p.C) grow = p.gw + ((long long)ij * p.C + cp) * p.ldw;
                        else if (p.Cj == 1 && mrow == p.C) grow = p.gb;       // ones column of the shared im2col matrix: the bias gradient
                    }
                }
                const bool single = (nb * p.Ho * p.Wo) <= TC_WGRAD_ROWS;     // one split: plain stores, else fp32 atomics
                for (int cb = 0; cb < T.n_umma; cb += 16) {
                    float v[16];
                    tmem_ld16(trow + cb, v);
                    if (!grow) continue;
                    if (single && n0 + cb + 16 <= p.ldw) {   // whole chunk inside the padded row: pad columns get exact zeros
#pragma unroll
                        for (int j = 0; j < 16; j += 4)
                            *reinterpret_cast<float4*>(grow + n0 + cb + j) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
                    } else if (!single && n0 + cb + 16 <= p.ldw) {   // split reduction: four-wide fp32 atomics (pad columns add zeros)
#pragma unroll
                        for (int j = 0; j < 16; j += 4)
                            if (n0 + cb + j < Ncols)
                                atomicAdd(reinterpret_cast<float4*>(grow + n0 + cb + j), make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]));
                    } else {
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
            }
        };
        int jj = 0;
        for (int j = 0; j < my_tiles; ++j) {
            const TcTile& T = ring[j % TC_RING];
            const int nsub = T.nsub;
            if (MODE == TC_FWD) {
                // this warp's copy of the tile's bias columns (bias arrays are padded to 8 elements in the arena)
                __syncwarp();
                if (lane * 8 < T.n_umma && T.n0 + lane * 8 < T.Ncols)
                    const_cast<uint4*>(wbias)[lane] = ldg16(T.p.bias + T.n0 + lane * 8);
                __syncwarp();
            }
            for (int s = 0; s < nsub; ++s, ++jj) {
                const int acc = jj % NACC;
                // long tiles: back off while waiting so that the spin does not take issue slots from the loaders
                if (T.nk > 32) mbar_wait_relaxed(sTFull + acc * 8, (uint32_t)((jj / NACC) & 1));
                else mbar_wait(sTFull + acc * 8, (uint32_t)((jj / NACC) & 1));
                tc_fence_after();
                rel_bar = sTEmpty + acc * 8;
                if (MODE == TC_FWD && T.nsplit == 1) epilogue_fwd(T, s, *slot_ptr + (uint32_t)(acc * acc_cols));
                else epilogue(T, s, *slot_ptr + (uint32_t)(acc * acc_cols), ewarp);
                if (rel_bar) release_acc();
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == TC_LW) tmem_dealloc(*slot_ptr, (uint32_t)tmem_cols);
}

}  // namespace srn
