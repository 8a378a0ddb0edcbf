// Host-side internals shared by the .cu files: error reporting, launch counting, TMA descriptor encoding.
#pragma once
#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <cuda.h>
#include <cuda_runtime.h>

#include "../../include/b200sd.h"

namespace b200sd {

int32_t set_error(int32_t code, const char* fmt, ...);
void count_launch(int n = 1);
int num_sms();

// Encode a bf16 tiled tensor map with SWIZZLE_128B (inner box = 64 elements = 128 bytes).
// dims/strides innermost first; strides[i] is the byte stride of dim i+1 (dim 0 is contiguous).
int32_t encode_tmap_bf16(CUtensorMap* map, const void* base, int rank, const uint64_t* dims,
                         const uint64_t* strides_bytes, const uint32_t* box);
int32_t encode_tmap_2d(CUtensorMap* map, const void* base, int elem_bytes, uint64_t cols, uint64_t rows,
                       uint64_t row_stride_bytes, uint32_t box_cols, uint32_t box_rows, int swizzle_bytes);

#define B200SD_CHECK_ARG(cond, ...)                                  \
    do {                                                             \
        if (!(cond)) return set_error(B200SD_ERR_BAD_ARG, __VA_ARGS__); \
    } while (0)

#define B200SD_CHECK_LAUNCH(name)                                                                       \
    do {                                                                                                \
        cudaError_t e__ = cudaGetLastError();                                                           \
        if (e__ != cudaSuccess) return set_error(B200SD_ERR_CUDA, "%s: %s", name, cudaGetErrorString(e__)); \
    } while (0)

static inline cudaStream_t as_stream(b200sd_stream_t s) { return reinterpret_cast<cudaStream_t>(s); }

}  // namespace b200sd
