// Host-side helpers shared by the translation units of libserann_b200.so, and the debug hooks
// the parity tests use (not part of the product ABI in include/serann_b200.h).
#pragma once
#include <atomic>
#include <cuda_runtime.h>
#include <stdint.h>
#include <string>

struct srn_engine;

namespace srn {
extern thread_local std::string g_error;
extern std::atomic<unsigned long long> g_launches;
extern std::atomic<unsigned long long> g_allocs;
cudaError_t dmalloc(void** p, size_t bytes);      // counted cudaMalloc
int fail(int code, const char* fmt, ...);
}  // namespace srn

extern "C" {
// Test hook, process-wide, read by srn_build: 1 (default) runs aligned Dense/Conv problems on the tcgen05 kernels, 0 routes
// every GEMM through the general CUDA-core kernels that otherwise only serve odd-stride views (A/B parity tests).
int srn_debug_set_tensor_cores(int32_t on);
// one training step on the given rows of the training split; apply_adam = 0 leaves the fp32 gradients in place
int srn_debug_step(srn_engine* h, const int32_t* rows, int32_t nb, int32_t apply_adam);
// activation (grad = 0) or activation gradient (grad = 1) of tensor `tensor` of one individual, first nb rows
int srn_debug_get_tensor(srn_engine* h, int32_t individual, int32_t tensor, int32_t grad, int32_t nb, uint16_t* buf);
// fp32 gradient accumulator of a weight (which = 0 kernel/gamma, 1 bias/beta)
int srn_debug_get_grad(srn_engine* h, int32_t individual, int32_t layer, int32_t which, float* buf, int64_t count);
int srn_debug_launch_groups(srn_engine* h);
// with SRN_GUARD=1 in the environment every handle buffer carries a guard band behind its live bytes: 0 = all intact,
// k > 0 = buffer k was overrun, -1 = guards are off (the pool's compute-sanitizer is closed; this is the stand-in)
int srn_debug_check_guards(srn_engine* h);
// per-launch records of the last profiled region; returns the number written (see srn_engine.cu)
int srn_debug_group_probs(srn_engine* h, int32_t group, int32_t cap, int32_t* out);
int srn_debug_profile_detail(srn_engine* h, int32_t nb, int32_t cap, float* ms, float* start, int32_t* info, double* work);
}
