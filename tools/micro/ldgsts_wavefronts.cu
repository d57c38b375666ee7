// Micro-benchmark: shared-memory wavefronts and time of 16-byte cp.async (LDGSTS.128) for several lane -> global address
// patterns.  nvcc -gencode arch=compute_100a,code=sm_100a -o ldgsts ldgsts_wavefronts.cu ; ncu --metrics
// l1tex__data_pipe_lsu_wavefronts_mem_shared.sum,l1tex__t_requests_pipe_lsu_mem_global_op_ldgsts.sum ./ldgsts
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

template <int V>
__global__ void __launch_bounds__(256) k(const char* __restrict__ g, int iters, size_t span) {
    extern __shared__ __align__(128) char sm[];
    const uint32_t s = (uint32_t)__cvta_generic_to_shared(sm);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, l8 = lane & 7, lq = lane >> 3;
    size_t base = ((size_t)blockIdx.x * 8 + warp) * 65536;
    for (int it = 0; it < iters; ++it) {
        size_t off;
        if (V == 0) off = (size_t)lane * 16;                               // 512 contiguous bytes, 512-aligned
        else if (V == 1) off = (size_t)l8 * 4096 + lq * 16;                // 8 rows x 64 B, rows far apart
        else if (V == 2) off = (size_t)l8 * 64 + lq * 16 + 64;             // 8 rows x 64 B adjacent: contiguous, 64-aligned only
        else if (V == 3) off = (size_t)lane * 256;                         // 32 different lines
        else if (V == 4) off = (size_t)l8 * 16 + lq * 4096;                // 4 rows x 128 B, rows far apart
        else off = (size_t)(lane >> 1) * 4096 + (lane & 1) * 16;           // 16 sectors of 32 B
        const char* src = g + (base + (size_t)it * 8192 + off) % span;
        uint32_t dst = s + warp * 2048 + (it & 3) * 512 + lane * 16;
        if (V == 6) {      // global: 4 rows x 128 B far apart; shared: swizzle-128B rows (row r at r * 128, chunk c at c ^ (r & 7))
            const int r = (it & 3) * 4 + (lane >> 3), c = lane & 7;
            src = g + (base + (size_t)it * 8192 + (size_t)(lane >> 3) * 4096 + c * 16) % span;
            dst = s + warp * 2048 + r * 128 + ((c ^ (r & 7)) << 4);
        }
        if (V == 7) {      // the loader's old mapping: 8 rows x 4 chunks, lane = chunk * 8 + row, core-matrix destination
            src = g + (base + (size_t)it * 8192 + (size_t)l8 * 64 + lq * 16) % span;
        }
        if (V >= 8) {      // .ca (allocate in L1) on a 16 KB window per warp: L1 hits after the first pass
            const int v = V - 8;
            size_t o = v == 0 ? (size_t)lane * 16 : (v == 1 ? (size_t)l8 * 64 + lq * 16 : (size_t)(lane >> 3) * 128 + (lane & 7) * 16 + 64);
            src = g + ((size_t)blockIdx.x * 8 + warp) * 16384 + ((size_t)(it & 15) * 1024 + o) % 16384;
            asm volatile("cp.async.ca.shared.global [%0], [%1], 16;\n" ::"r"(dst), "l"(src) : "memory");
        } else
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(dst), "l"(src) : "memory");
        if ((it & 3) == 3) asm volatile("cp.async.wait_all;\n" ::: "memory");
    }
    asm volatile("cp.async.wait_all;\n" ::: "memory");
}

template <int V>
void run(const char* g, size_t span, const char* name) {
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    const int iters = 4096, grid = 148 * 2;
    k<V><<<grid, 256, 16384>>>(g, 64, span);
    cudaEventRecord(a);
    k<V><<<grid, 256, 16384>>>(g, iters, span);
    cudaEventRecord(b); cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms, a, b);
    const double ops = (double)grid * 8 * iters;
    printf("%-44s %.3f ms  %.1f cycles/op/SM (at 1.9 GHz)  %.0f GB/s\n", name, ms, ms * 1e-3 * 1.9e9 / (ops / 148), ops * 512 / ms / 1e6);
}

int main() {
    const size_t span = (size_t)64 << 20;          // 64 MB: L2-resident
    char* g; cudaMalloc(&g, span + (1 << 20)); cudaMemset(g, 1, span + (1 << 20));
    run<0>(g, span, "V0 512 B contiguous, aligned");
    run<1>(g, span, "V1 8 rows x 64 B, far apart");
    run<2>(g, span, "V2 8 rows x 64 B adjacent (+64 B)");
    run<3>(g, span, "V3 32 lines");
    run<4>(g, span, "V4 4 rows x 128 B, far apart");
    run<5>(g, span, "V5 16 rows x 32 B");
    run<6>(g, span, "V6 4 rows x 128 B -> swizzled rows");
    run<7>(g, span, "V7 8 adjacent rows x 64 B, row-fastest lanes");
    run<8>(g, span, "V8 .ca L1-resident, 512 B contiguous");
    run<9>(g, span, "V9 .ca L1-resident, 8 rows x 64 B row-fastest");
    run<10>(g, span, "V10 .ca L1-resident, 4 rows x 128 B +64");
    cudaError_t e = cudaDeviceSynchronize();
    printf("%s\n", cudaGetErrorString(e));
    return e != cudaSuccess;
}
