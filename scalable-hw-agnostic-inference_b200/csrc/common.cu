// Library plumbing: status / error string, device check, launch counter, TMA descriptor encoder.
#include <atomic>
#include <cstdlib>
#include <cstring>
#include <mutex>

#include "common.h"

namespace b200sd {

static thread_local char g_err[512] = "";
static std::atomic<int64_t> g_launches{0};

int32_t set_error(int32_t code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}

void count_launch(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }

int num_sms() {
    static int cached[64] = {0};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= 64) dev = 0;
    if (cached[dev] == 0) {
        int n = 0;
        cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
        cached[dev] = n > 0 ? n : 148;
    }
    return cached[dev];
}

bool pdl_enabled() {
    static const bool on = [] {
        const char* e = getenv("B200SD_PDL");
        return !(e && e[0] == '0');
    }();
    return on;
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
    static EncodeTiledFn fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult qres;
        // resolved through the runtime so the library has no link-time dependency on libcuda
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
    });
    return fn;
}

int32_t encode_tmap_bf16(CUtensorMap* map, const void* base, int rank, const uint64_t* dims,
                         const uint64_t* strides_bytes, const uint32_t* box, const uint32_t* elem_strides) {
    EncodeTiledFn fn = get_encode_fn();
    if (!fn) return set_error(B200SD_ERR_DEVICE, "cuTensorMapEncodeTiled not available from the driver");
    cuuint64_t gdim[5];
    cuuint64_t gstr[4];
    cuuint32_t bx[5];
    cuuint32_t es[5];
    for (int i = 0; i < rank; ++i) {
        gdim[i] = dims[i];
        bx[i] = box[i];
        es[i] = elem_strides ? elem_strides[i] : 1;
        if (i > 0) gstr[i - 1] = strides_bytes[i - 1];
    }
    CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank, const_cast<void*>(base), gdim, gstr, bx,
                    es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        return set_error(B200SD_ERR_BAD_ARG,
                         "cuTensorMapEncodeTiled failed (%d): base=%p rank=%d dims=[%llu,%llu,%llu,%llu] "
                         "stride0=%llu box=[%u,%u,%u,%u]",
                         (int)r, base, rank, (unsigned long long)dims[0], (unsigned long long)(rank > 1 ? dims[1] : 0),
                         (unsigned long long)(rank > 2 ? dims[2] : 0), (unsigned long long)(rank > 3 ? dims[3] : 0),
                         (unsigned long long)(rank > 1 ? strides_bytes[0] : 0), box[0], rank > 1 ? box[1] : 0,
                         rank > 2 ? box[2] : 0, rank > 3 ? box[3] : 0);
    }
    return B200SD_OK;
}

// 2-D row-major tensor [rows][cols] of 2- or 4-byte elements, box = box_cols x box_rows, swizzle 64 or 128 bytes
// (= box_cols * elem_bytes).  Used for the GEMM epilogue's TMA stores.
int32_t encode_tmap_2d(CUtensorMap* map, const void* base, int elem_bytes, uint64_t cols, uint64_t rows,
                       uint64_t row_stride_bytes, uint32_t box_cols, uint32_t box_rows, int swizzle_bytes, int dtype) {
    EncodeTiledFn fn = get_encode_fn();
    if (!fn) return set_error(B200SD_ERR_DEVICE, "cuTensorMapEncodeTiled not available from the driver");
    cuuint64_t gdim[2] = {cols, rows};
    cuuint64_t gstr[1] = {row_stride_bytes};
    cuuint32_t bx[2] = {box_cols, box_rows};
    cuuint32_t es[2] = {1, 1};
    // a TMA reduce-add does arithmetic on the elements, so 16-bit maps name their floating-point format
    const CUtensorMapDataType dt = elem_bytes == 4 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32
                                   : dtype == B200SD_F16 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16
                                   : dtype == B200SD_BF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_UINT16;
    CUresult r = fn(map, dt, 2,
                    const_cast<void*>(base), gdim, gstr, bx, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                    swizzle_bytes == 128 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B,
                    CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS)
        return set_error(B200SD_ERR_BAD_ARG, "cuTensorMapEncodeTiled (2-D, %d-byte) failed (%d): base=%p cols=%llu rows=%llu "
                         "stride=%llu", elem_bytes, (int)r, base, (unsigned long long)cols, (unsigned long long)rows,
                         (unsigned long long)row_stride_bytes);
    return B200SD_OK;
}

}  // namespace b200sd

using namespace b200sd;

extern "C" {

int32_t b200sd_abi_version(void) { return B200SD_ABI_VERSION; }

const char* b200sd_last_error(void) { return g_err; }

int64_t b200sd_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }

int32_t b200sd_check_device(int32_t dev) {
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0)
        return set_error(B200SD_ERR_DEVICE, "no CUDA device: %s", cudaGetErrorString(e));
    if (dev < 0 || dev >= n) return set_error(B200SD_ERR_BAD_ARG, "device %d out of range (%d devices)", dev, n);
    int major = 0, minor = 0;
    cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev);
    cudaDeviceGetAttribute(&minor, cudaDevAttrComputeCapabilityMinor, dev);
    if (major != 10)
        return set_error(B200SD_ERR_DEVICE, "device %d is sm_%d%d; libb200sd is built for sm_100a only", dev, major,
                         minor);
    return B200SD_OK;
}

}  // extern "C"
