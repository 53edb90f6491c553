// Bucket-major index materialisation.  The reference keeps the bucket layout implicit in
// data_prediction + the DataFrame index and re-derives it every round with a groupby and a
// per-bucket .loc copy (search/li/LearnedIndex.py:101-104, :350, :357); here the rows are
// permuted once into a bucket-major fp16 matrix that stays resident in HBM.
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/lmi_b200.h"

namespace lmi {

// one warp per source row; 8-byte (4 x fp16) stores
template <typename T>
__global__ void scatter_rows_kernel(const T* __restrict__ src, const int64_t* __restrict__ dest_row, int64_t m, int d,
                                    __half* __restrict__ dst, int ld) {
    const int warps = blockDim.x >> 5;
    const int lane = threadIdx.x & 31;
    for (int64_t i = (int64_t)blockIdx.x * warps + (threadIdx.x >> 5); i < m; i += (int64_t)gridDim.x * warps) {
        const int64_t to = dest_row ? dest_row[i] : i;
        const T* s = src + (size_t)i * d;
        __half* o = dst + (size_t)to * ld;
        for (int c = lane; c < d; c += 32) {
            if constexpr (sizeof(T) == 2) o[c] = s[c];
            else o[c] = __float2half_rn(static_cast<float>(s[c]));
        }
        for (int c = d + lane; c < ld; c += 32) o[c] = __float2half_rn(0.f);  // zero padding: inner products unchanged
    }
}

__global__ void cast_f16_kernel(const float* __restrict__ src, __half* __restrict__ dst, int64_t count) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += stride)
        dst[i] = __float2half_rn(src[i]);
}

}  // namespace lmi

extern "C" int lmi_scatter_rows(const void* src, int src_is_f16, const int64_t* dest_row, int64_t m, int d,
                                void* dst_f16, int dst_ld, void* stream) {
    if (!src || !dst_f16 || m < 0 || d <= 0 || dst_ld < d) return LMI_E_BADARG;
    if (m == 0) return 0;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int warps = 8;
    int64_t grid = (m + warps - 1) / warps;
    if (grid > 148 * 64) grid = 148 * 64;
    if (src_is_f16)
        lmi::scatter_rows_kernel<__half><<<(unsigned)grid, warps * 32, 0, st>>>(static_cast<const __half*>(src),
                                                                               dest_row, m, d,
                                                                               static_cast<__half*>(dst_f16), dst_ld);
    else
        lmi::scatter_rows_kernel<float><<<(unsigned)grid, warps * 32, 0, st>>>(static_cast<const float*>(src),
                                                                              dest_row, m, d,
                                                                              static_cast<__half*>(dst_f16), dst_ld);
    return static_cast<int>(cudaGetLastError());
}

extern "C" int lmi_cast_f16(const float* src, void* dst_f16, int64_t count, void* stream) {
    if (!src || !dst_f16 || count < 0) return LMI_E_BADARG;
    if (count == 0) return 0;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    int64_t grid = (count + 255) / 256;
    if (grid > 148 * 32) grid = 148 * 32;
    lmi::cast_f16_kernel<<<(unsigned)grid, 256, 0, st>>>(src, static_cast<__half*>(dst_f16), count);
    return static_cast<int>(cudaGetLastError());
}

extern "C" int lmi_version(void) { return LMI_B200_ABI_VERSION; }

extern "C" const char* lmi_error_string(int code) {
    switch (code) {
        case 0: return "success";
        case LMI_E_BADARG: return "lmi: bad argument";
        case LMI_E_UNSUPPORTED: return "lmi: shape not supported by the sm_100a kernels";
        case LMI_E_DRIVER: return "lmi: CUDA driver entry point unavailable / tensor map encode failed";
        default: break;
    }
    if (code > 0) return cudaGetErrorString(static_cast<cudaError_t>(code));
    return "lmi: unknown error";
}
