// Bucket-major index materialisation.  The reference keeps the bucket layout implicit in
// data_prediction + the DataFrame index and re-derives it every round with a groupby and a
// per-bucket .loc copy (search/li/LearnedIndex.py:101-104, :350, :357); here the rows are
// permuted once into a bucket-major fp16 matrix that stays resident in HBM.
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/lmi_b200.h"

namespace lmi {

// one warp per source row.  Rows with dest_row[i] < 0 are dropped (objects without a bucket).
// `inexact` (optional) counts the values the fp32 -> fp16 conversion changed: the scan computes
// f32 arithmetic on fp16-representable values (what SISAP's fp16 `emb` gives the reference after
// its float32 casts), so a non-zero count means the distances are those of ROUNDED vectors.
template <typename T>
__global__ void scatter_rows_kernel(const T* __restrict__ src, const int64_t* __restrict__ dest_row, int64_t m, int d,
                                    __half* __restrict__ dst, int ld, unsigned long long* __restrict__ inexact) {
    const int warps = blockDim.x >> 5;
    const int lane = threadIdx.x & 31;
    int bad = 0;
    for (int64_t i = (int64_t)blockIdx.x * warps + (threadIdx.x >> 5); i < m; i += (int64_t)gridDim.x * warps) {
        const int64_t to = dest_row ? dest_row[i] : i;
        if (to < 0) continue;
        const T* s = src + (size_t)i * d;
        __half* o = dst + (size_t)to * ld;
        for (int c = lane; c < d; c += 32) {
            if constexpr (sizeof(T) == 2) {
                o[c] = s[c];
            } else {
                const float v = static_cast<float>(s[c]);
                const __half h = __float2half_rn(v);
                o[c] = h;
                bad += (__half2float(h) != v && v == v) ? 1 : 0;
            }
        }
        for (int c = d + lane; c < ld; c += 32) o[c] = __float2half_rn(0.f);  // zero padding: inner products unchanged
    }
    if (inexact) {
        bad = __reduce_add_sync(0xffffffffu, bad);
        if (lane == 0 && bad) atomicAdd(inexact, (unsigned long long)bad);
    }
}

__global__ void cast_f16_kernel(const float* __restrict__ src, __half* __restrict__ dst, int64_t count,
                                unsigned long long* __restrict__ inexact) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    int bad = 0;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += stride) {
        const float v = src[i];
        const __half h = __float2half_rn(v);
        dst[i] = h;
        bad += (__half2float(h) != v && v == v) ? 1 : 0;
    }
    if (inexact) {
        bad = __reduce_add_sync(0xffffffffu, bad);
        if ((threadIdx.x & 31) == 0 && bad) atomicAdd(inexact, (unsigned long long)bad);
    }
}

}  // namespace lmi

extern "C" int lmi_scatter_rows(const void* src, int src_is_f16, const int64_t* dest_row, int64_t m, int d,
                                void* dst_f16, int dst_ld, int64_t* inexact, void* stream) {
    if (!src || !dst_f16 || m < 0 || d <= 0 || dst_ld < d) return LMI_E_BADARG;
    if (m == 0) return 0;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int warps = 8;
    int64_t grid = (m + warps - 1) / warps;
    if (grid > 148 * 64) grid = 148 * 64;
    if (src_is_f16)
        lmi::scatter_rows_kernel<__half><<<(unsigned)grid, warps * 32, 0, st>>>(static_cast<const __half*>(src),
                                                                               dest_row, m, d,
                                                                               static_cast<__half*>(dst_f16), dst_ld,
                                                                               nullptr);
    else
        lmi::scatter_rows_kernel<float><<<(unsigned)grid, warps * 32, 0, st>>>(static_cast<const float*>(src),
                                                                              dest_row, m, d,
                                                                              static_cast<__half*>(dst_f16), dst_ld,
                                                                              reinterpret_cast<unsigned long long*>(inexact));
    return static_cast<int>(cudaGetLastError());
}

extern "C" int lmi_cast_f16(const float* src, void* dst_f16, int64_t count, int64_t* inexact, void* stream) {
    if (!src || !dst_f16 || count < 0) return LMI_E_BADARG;
    if (count == 0) return 0;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    int64_t grid = (count + 255) / 256;
    if (grid > 148 * 32) grid = 148 * 32;
    lmi::cast_f16_kernel<<<(unsigned)grid, 256, 0, st>>>(src, static_cast<__half*>(dst_f16), count,
                                                         reinterpret_cast<unsigned long long*>(inexact));
    return static_cast<int>(cudaGetLastError());
}

extern "C" int lmi_version(void) { return LMI_B200_ABI_VERSION; }

extern "C" const char* lmi_error_string(int code) {
    switch (code) {
        case 0: return "success";
        case LMI_E_BADARG: return "lmi: bad argument";
        case LMI_E_UNSUPPORTED: return "lmi: shape not supported by the sm_100a kernels";
        case LMI_E_DRIVER: return "lmi: CUDA driver entry point unavailable / tensor map encode failed";
        case LMI_E_SCRATCH_LIMIT:
            return "lmi: batch needs more scratch than the library's built-in array holds "
                   "(lmi_invert_static: lmi_invert_workspace_ints > LMI_INVERT_STATIC_MAX_INTS); use lmi_invert";
        default: break;
    }
    if (code > 0) return cudaGetErrorString(static_cast<cudaError_t>(code));
    return "lmi: unknown error";
}
