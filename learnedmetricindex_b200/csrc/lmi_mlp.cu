// MLP inference for the learned navigation step:
//   lmi_linear        y = act(x . W^T + b)            (model.py:29-79, :97-99, :232)
//   lmi_softmax_topn  softmax + top-B classes          (model.py:238-239; LearnedIndex.py:209-214)
//   lmi_argmax        first maximum per row            (model.py:223)
//
// lmi_linear is the fp32 CUDA-core form of a Linear layer: the verification kernel for the
// tensor-core path (lmi_mlp_tc.cu, split-precision tcgen05) and the fallback for input
// widths that are not a multiple of 4.  Parity demands probed-bucket lists identical to the
// reference's fp32 torch path except where probabilities tie within 1e-6, so both keep
// fp32 quality; a single-pass fp16/tf32 product would perturb logits by ~1e-3 relative.
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <float.h>
#include <stdint.h>

#include "../../include/lmi_b200.h"

namespace lmi {

constexpr int LIN_BM = 128;  // rows of x per block
constexpr int LIN_BN = 64;   // output features per block
constexpr int LIN_BK = 16;
constexpr int LIN_THREADS = 256;  // 16 x 16 threads, each 8 x 4 outputs

template <typename T>
__device__ __forceinline__ float4 load4(const T* p);
template <>
__device__ __forceinline__ float4 load4<float>(const float* p) {
    return *reinterpret_cast<const float4*>(p);
}
template <>
__device__ __forceinline__ float4 load4<__half>(const __half* p) {
    const uint2 raw = *reinterpret_cast<const uint2*>(p);
    const float2 a = __half22float2(*reinterpret_cast<const __half2*>(&raw.x));
    const float2 b = __half22float2(*reinterpret_cast<const __half2*>(&raw.y));
    return make_float4(a.x, a.y, b.x, b.y);
}
template <typename T>
__device__ __forceinline__ float load1(const T* p);
template <>
__device__ __forceinline__ float load1<float>(const float* p) {
    return *p;
}
template <>
__device__ __forceinline__ float load1<__half>(const __half* p) {
    return __half2float(*p);
}

// y[M,N] = act(x[M,K] . W[N,K]^T + b[N]).  Requires K % 4 == 0 for the vector path
// (VEC) -- otherwise scalar loads.  Accumulation runs over k in ascending order.
template <typename TX, bool VEC>
__global__ void __launch_bounds__(LIN_THREADS)
linear_kernel(const TX* __restrict__ x, const float* __restrict__ W, const float* __restrict__ bias,
              float* __restrict__ y, int M, int K, int N, int relu) {
    __shared__ __align__(16) float As[2][LIN_BK][LIN_BM + 4];
    __shared__ __align__(16) float Bs[2][LIN_BK][LIN_BN + 4];

    const int tid = threadIdx.x;
    const int tx = tid & 15;   // output column group (4 columns)
    const int ty = tid >> 4;   // output row group (8 rows)
    const int m0 = blockIdx.y * LIN_BM;
    const int n0 = blockIdx.x * LIN_BN;

    // global -> smem assignment: A tile 128 x 16 = 512 float4 (2 per thread), B tile 64 x 16 = 256 float4
    const int a_row = tid >> 2;          // 0..63 (+64 for the second)
    const int a_k4 = (tid & 3) * 4;      // 0,4,8,12
    const int b_row = tid >> 2;          // 0..63
    const int b_k4 = (tid & 3) * 4;

    float acc[8][4];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

    auto fetchA = [&](int k0, int which) -> float4 {
        const int r = m0 + a_row + which * 64;
        const int k = k0 + a_k4;
        if (r >= M) return make_float4(0.f, 0.f, 0.f, 0.f);
        const TX* p = x + (size_t)r * K + k;
        if (VEC) {
            if (k + 3 < K) return load4<TX>(p);
        }
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (k < K) v.x = load1<TX>(p);
        if (k + 1 < K) v.y = load1<TX>(p + 1);
        if (k + 2 < K) v.z = load1<TX>(p + 2);
        if (k + 3 < K) v.w = load1<TX>(p + 3);
        return v;
    };
    auto fetchB = [&](int k0) -> float4 {
        const int r = n0 + b_row;
        const int k = k0 + b_k4;
        if (r >= N) return make_float4(0.f, 0.f, 0.f, 0.f);
        const float* p = W + (size_t)r * K + k;
        if (VEC) {
            if (k + 3 < K) return *reinterpret_cast<const float4*>(p);
        }
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (k < K) v.x = p[0];
        if (k + 1 < K) v.y = p[1];
        if (k + 2 < K) v.z = p[2];
        if (k + 3 < K) v.w = p[3];
        return v;
    };
    auto stash = [&](int buf, const float4& a0, const float4& a1, const float4& b0) {
        As[buf][a_k4 + 0][a_row] = a0.x; As[buf][a_k4 + 1][a_row] = a0.y;
        As[buf][a_k4 + 2][a_row] = a0.z; As[buf][a_k4 + 3][a_row] = a0.w;
        As[buf][a_k4 + 0][a_row + 64] = a1.x; As[buf][a_k4 + 1][a_row + 64] = a1.y;
        As[buf][a_k4 + 2][a_row + 64] = a1.z; As[buf][a_k4 + 3][a_row + 64] = a1.w;
        Bs[buf][b_k4 + 0][b_row] = b0.x; Bs[buf][b_k4 + 1][b_row] = b0.y;
        Bs[buf][b_k4 + 2][b_row] = b0.z; Bs[buf][b_k4 + 3][b_row] = b0.w;
    };

    const int nk = (K + LIN_BK - 1) / LIN_BK;
    float4 ra0 = fetchA(0, 0), ra1 = fetchA(0, 1), rb0 = fetchB(0);
    stash(0, ra0, ra1, rb0);
    __syncthreads();

    for (int kt = 0; kt < nk; ++kt) {
        const int buf = kt & 1;
        if (kt + 1 < nk) {
            ra0 = fetchA((kt + 1) * LIN_BK, 0);
            ra1 = fetchA((kt + 1) * LIN_BK, 1);
            rb0 = fetchB((kt + 1) * LIN_BK);
        }
#pragma unroll
        for (int kk = 0; kk < LIN_BK; ++kk) {
            const float4 a_lo = *reinterpret_cast<const float4*>(&As[buf][kk][ty * 8]);
            const float4 a_hi = *reinterpret_cast<const float4*>(&As[buf][kk][ty * 8 + 4]);
            const float4 b4 = *reinterpret_cast<const float4*>(&Bs[buf][kk][tx * 4]);
            const float a[8] = {a_lo.x, a_lo.y, a_lo.z, a_lo.w, a_hi.x, a_hi.y, a_hi.z, a_hi.w};
            const float b[4] = {b4.x, b4.y, b4.z, b4.w};
#pragma unroll
            for (int i = 0; i < 8; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
        }
        if (kt + 1 < nk) {
            stash(buf ^ 1, ra0, ra1, rb0);
            __syncthreads();
        }
    }

#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int r = m0 + ty * 8 + i;
        if (r >= M) continue;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int c = n0 + tx * 4 + j;
            if (c >= N) continue;
            float v = acc[i][j] + bias[c];
            if (relu) v = fmaxf(v, 0.f);
            y[(size_t)r * N + c] = v;
        }
    }
}

__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// One warp per query row; the row's probabilities live in shared memory.
__global__ void softmax_topn_kernel(const float* __restrict__ logits, int nq, int C, int B,
                                    int32_t* __restrict__ order, float* __restrict__ probs_out) {
    extern __shared__ float sm[];
    const int warps = blockDim.x >> 5;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int q = blockIdx.x * warps + warp;
    if (q >= nq) return;
    float* p = sm + (size_t)warp * C;
    const float* src = logits + (size_t)q * C;

    float mx = -INFINITY;
    for (int c = lane; c < C; c += 32) {
        const float v = src[c];
        p[c] = v;
        mx = fmaxf(mx, v);
    }
    mx = warp_max(mx);
    float sum = 0.f;
    for (int c = lane; c < C; c += 32) {
        const float e = expf(p[c] - mx);
        p[c] = e;
        sum += e;
    }
    sum = warp_sum(sum);
    for (int c = lane; c < C; c += 32) p[c] = p[c] / sum;
    __syncwarp();

    for (int r = 0; r < B; ++r) {
        float best = -1.f;
        int arg = 0x7fffffff;
        for (int c = lane; c < C; c += 32) {
            const float v = p[c];
            if (v > best) {  // strided ascending scan: first maximum wins inside a lane
                best = v;
                arg = c;
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float ob = __shfl_xor_sync(0xffffffffu, best, o);
            const int oa = __shfl_xor_sync(0xffffffffu, arg, o);
            if (ob > best || (ob == best && oa < arg)) {
                best = ob;
                arg = oa;
            }
        }
        if (lane == 0) {
            order[(size_t)q * B + r] = (arg == 0x7fffffff) ? -1 : arg;
            probs_out[(size_t)q * B + r] = best;
            if (arg != 0x7fffffff) p[arg] = -2.f;  // taken
        }
        __syncwarp();
    }
}

__global__ void argmax_kernel(const float* __restrict__ logits, int64_t n, int C, int64_t* __restrict__ out) {
    const int warps = blockDim.x >> 5;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t row = (int64_t)blockIdx.x * warps + warp;
    if (row >= n) return;
    const float* src = logits + (size_t)row * C;
    float best = -INFINITY;
    int arg = 0x7fffffff;
    for (int c = lane; c < C; c += 32) {
        const float v = src[c];
        if (v > best || arg == 0x7fffffff) {
            best = v;
            arg = c;
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float ob = __shfl_xor_sync(0xffffffffu, best, o);
        const int oa = __shfl_xor_sync(0xffffffffu, arg, o);
        if (ob > best || (ob == best && oa < arg)) {
            best = ob;
            arg = oa;
        }
    }
    if (lane == 0) out[row] = arg;
}

}  // namespace lmi

extern "C" int lmi_linear(const void* x, int x_is_f16, const float* W, const float* b, float* y, int nq,
                          int in_dim, int out_dim, int relu, void* stream) {
    if (!x || !W || !b || !y || nq < 0 || in_dim <= 0 || out_dim <= 0) return LMI_E_BADARG;
    if (nq == 0) return 0;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    dim3 grid((out_dim + lmi::LIN_BN - 1) / lmi::LIN_BN, (nq + lmi::LIN_BM - 1) / lmi::LIN_BM);
    const bool vec = (in_dim % 4 == 0);
    if (x_is_f16) {
        const __half* xx = static_cast<const __half*>(x);
        if (vec)
            lmi::linear_kernel<__half, true><<<grid, lmi::LIN_THREADS, 0, st>>>(xx, W, b, y, nq, in_dim, out_dim, relu);
        else
            lmi::linear_kernel<__half, false><<<grid, lmi::LIN_THREADS, 0, st>>>(xx, W, b, y, nq, in_dim, out_dim, relu);
    } else {
        const float* xx = static_cast<const float*>(x);
        if (vec)
            lmi::linear_kernel<float, true><<<grid, lmi::LIN_THREADS, 0, st>>>(xx, W, b, y, nq, in_dim, out_dim, relu);
        else
            lmi::linear_kernel<float, false><<<grid, lmi::LIN_THREADS, 0, st>>>(xx, W, b, y, nq, in_dim, out_dim, relu);
    }
    return static_cast<int>(cudaGetLastError());
}

extern "C" int lmi_softmax_topn(const float* logits, int nq, int C, int B, int32_t* bucket_order, float* probs,
                                void* stream) {
    if (!logits || !bucket_order || !probs || nq < 0 || C <= 0 || B <= 0 || B > C) return LMI_E_BADARG;
    if (nq == 0) return 0;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    int warps = 4;
    while (warps > 1 && (size_t)warps * C * sizeof(float) > 96 * 1024) warps >>= 1;
    const size_t smem = (size_t)warps * C * sizeof(float);
    if (smem > 200 * 1024) return LMI_E_UNSUPPORTED;
    if (smem > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(lmi::softmax_topn_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             (int)smem);
        if (e != cudaSuccess) return (int)e;
    }
    const int grid = (nq + warps - 1) / warps;
    lmi::softmax_topn_kernel<<<grid, warps * 32, smem, st>>>(logits, nq, C, B, bucket_order, probs);
    return static_cast<int>(cudaGetLastError());
}

extern "C" int lmi_argmax(const float* logits, int64_t n, int C, int64_t* out, void* stream) {
    if (!logits || !out || n < 0 || C <= 0) return LMI_E_BADARG;
    if (n == 0) return 0;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int warps = 8;
    const int64_t grid = (n + warps - 1) / warps;
    if (grid > 0x7fffffff) return LMI_E_UNSUPPORTED;
    lmi::argmax_kernel<<<(unsigned)grid, warps * 32, 0, st>>>(logits, n, C, out);
    return static_cast<int>(cudaGetLastError());
}
