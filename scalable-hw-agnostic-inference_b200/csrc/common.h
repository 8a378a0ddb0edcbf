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
// elem_strides (optional): traversal stride per dimension -- a box of extent b with stride s delivers ceil(b / s) elements.
int32_t encode_tmap_bf16(CUtensorMap* map, const void* base, int rank, const uint64_t* dims,
                         const uint64_t* strides_bytes, const uint32_t* box, const uint32_t* elem_strides = nullptr);
int32_t encode_tmap_2d(CUtensorMap* map, const void* base, int elem_bytes, uint64_t cols, uint64_t rows,
                       uint64_t row_stride_bytes, uint32_t box_cols, uint32_t box_rows, int swizzle_bytes,
                       int dtype = -1);  // dtype: b200sd_dtype of 2-byte elements (needed by reduce-adds), -1 = opaque

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

// B200SD_PDL=0 launches every kernel fully serialised (no programmatic dependent launch)
bool pdl_enabled();

// One launch path for every kernel of the library: optional thread-block cluster, and the programmatic-dependent-launch
// attribute (ptx.cuh, pdl_sync) so that consecutive kernels of a stream -- and of a captured CUDA graph -- overlap
// one kernel's set-up with the previous kernel's tail.
template <typename... KArgs, typename... Args>
static inline cudaError_t launch_kernel(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                                        int cluster_x, Args&&... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid, cfg.blockDim = block, cfg.dynamicSmemBytes = smem, cfg.stream = stream;
    cudaLaunchAttribute attr[2];
    unsigned n = 0;
    if (cluster_x > 1) {
        attr[n].id = cudaLaunchAttributeClusterDimension;
        attr[n].val.clusterDim.x = (unsigned)cluster_x, attr[n].val.clusterDim.y = 1, attr[n].val.clusterDim.z = 1;
        ++n;
    }
    if (pdl_enabled()) {
        attr[n].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[n].val.programmaticStreamSerializationAllowed = 1;
        ++n;
    }
    cfg.attrs = attr, cfg.numAttrs = n;
    return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

#define B200SD_LAUNCH(name, kernel, grid, block, smem, stream, ...)                                              \
    do {                                                                                                         \
        cudaError_t le__ = launch_kernel(kernel, grid, block, smem, stream, 1, __VA_ARGS__);                       \
        if (le__ != cudaSuccess) return set_error(B200SD_ERR_CUDA, "%s launch: %s", name, cudaGetErrorString(le__)); \
    } while (0)

}  // namespace b200sd
