// Host-side helper shared by the translation units: 2-D tiled, swizzle-128B tensor maps over row-major fp16 matrices in
// THIS GPU's memory (cuTensorMapEncodeTiled fetched through the runtime, libcuda is not linked).
#pragma once
#include <cstdlib>
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace srn {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
inline EncodeTiledFn encode_tiled_fn() {
    static EncodeTiledFn fn = nullptr;
    static bool tried = false;
    if (!tried) {
        tried = true;
        void* f = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (getenv("SRN_NO_TMA") == nullptr &&
            cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &f, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(f);
    }
    return fn;
}
// row-major fp16 matrix [rows][cols] with a row pitch (elements), box = 64 columns x box_rows rows, swizzle-128B;
// rows and columns outside the matrix are zero-filled by the copy engine
inline bool encode_matrix(CUtensorMap* m, const void* base, long long rows, long long cols, long long pitch, int box_rows) {
    EncodeTiledFn fn = encode_tiled_fn();
    if (!fn || !base || rows <= 0 || cols <= 0 || (pitch * 2) % 16 != 0 || ((uintptr_t)base & 15) != 0) return false;
    const cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
    const cuuint64_t strides[1] = {(cuuint64_t)pitch * 2};
    const cuuint32_t box[2] = {64u, (cuuint32_t)box_rows};
    const cuuint32_t estr[2] = {1u, 1u};
    return fn(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
              CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// input slab of the stride-1 convolution kernel (srn_conv_slab.cuh): rows of C channels (C * 2 = 32, 64 or 128 bytes, one
// swizzle row per pixel) of a dense [rows][C] fp16 tensor, box = C x box_rows, swizzle span = the row
inline bool encode_slab(CUtensorMap* m, const void* base, long long rows, int C, int box_rows) {
    EncodeTiledFn fn = encode_tiled_fn();
    if (!fn || !base || rows <= 0 || (C != 16 && C != 32 && C != 64) || box_rows <= 0 || box_rows > 256 || ((uintptr_t)base & 15) != 0) return false;
    const cuuint64_t dims[2] = {(cuuint64_t)C, (cuuint64_t)rows};
    const cuuint64_t strides[1] = {(cuuint64_t)C * 2};
    const cuuint32_t box[2] = {(cuuint32_t)C, (cuuint32_t)box_rows};
    const cuuint32_t estr[2] = {1u, 1u};
    const CUtensorMapSwizzle sw = C == 16 ? CU_TENSOR_MAP_SWIZZLE_32B : (C == 32 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_128B);
    return fn(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, sw,
              CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

}  // namespace srn
