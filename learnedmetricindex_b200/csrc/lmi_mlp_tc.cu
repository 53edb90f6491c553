// Tensor-core form of the navigation MLP's Linear layers (search/li/model.py:29-79, :97-99,
// :232): y = act(x . W^T + b) as tcgen05.mma kind::tf32 with SPLIT operands, so that the
// result keeps fp32 quality -- parity needs probed-bucket lists identical to the reference's
// fp32 torch path except where probabilities tie within 1e-6.
//
//   * every fp32 operand is written as hi + mid + lo, each part exactly representable in tf32
//     (11 significant bits); queries stored as fp16 are one exact part.  A product of two
//     parts is exact in fp32 (22 bits), and all part pairs whose weight is >= 2^-22 of the
//     leading one are issued: 3 MMAs per k-step for a 1 x 3 split (fp16 x, fp32 W), 6 for 3 x 3;
//   * the tensor core's accumulator adds with truncation, so sums are kept SHORT: one K-block
//     of 32 is accumulated in TMEM (smallest terms first, so the truncations of the leading
//     term are the only ones that matter) and the epilogue warps add each K-block's partial
//     tile into fp32 registers with round-to-nearest;
//   * x and W parts stream by TMA (SWIZZLE_128B boxes of 32 fp32 x rows) through a 2-3 stage
//     ring; the partial tiles are double buffered in TMEM (2 x 64 columns);
//   * the epilogue adds the bias, applies ReLU and writes either fp32 logits or the hi/mid/lo
//     parts the next layer reads.
// Tile: 128 rows of x (TMEM lanes; thread == row) x 64 output features per CTA.
// Roles: warps 0-3 epilogue, warp 4 TMA producer, warp 5 MMA issuer.
#include <cuda.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/lmi_b200.h"
#include "lmi_ptx.cuh"

namespace lmi {

constexpr int TC_BM = 128;
constexpr int TC_BN = 64;
constexpr int TC_BK = 32;  // fp32 per 128-byte shared-memory row == one K-block == one accumulation chunk
constexpr int TC_A_BYTES = TC_BM * 128;
constexpr int TC_B_BYTES = TC_BN * 128;
constexpr int TC_THREADS = 192;
constexpr int TC_MAX_PARTS = 3;
constexpr int TC_MAX_STAGES = 4;

__device__ __forceinline__ float tf32_trunc(float v) { return __uint_as_float(__float_as_uint(v) & 0xffffe000u); }

// hi = tf32(v), mid = tf32(v - hi), lo = v - hi - mid: all three exact, |lo| < 2^-22 |v|
__device__ __forceinline__ void split3(float v, float& hi, float& mid, float& lo) {
    hi = tf32_trunc(v);
    const float r = v - hi;
    mid = tf32_trunc(r);
    lo = r - mid;
}

template <typename T>
__global__ void split_tf32_kernel(const T* __restrict__ x, int64_t count, float* __restrict__ hi,
                                  float* __restrict__ mid, float* __restrict__ lo) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    float v;
    if constexpr (sizeof(T) == 2) v = __half2float(x[i]);
    else v = x[i];
    float h, m, l;
    split3(v, h, m, l);
    hi[i] = h;
    if (mid) mid[i] = m;
    if (lo) lo[i] = l;
}

struct TcMaps {
    CUtensorMap a[TC_MAX_PARTS];
    CUtensorMap b[TC_MAX_PARTS];
};

struct TcBarriers {
    uint64_t full[TC_MAX_STAGES];
    uint64_t empty[TC_MAX_STAGES];
    uint64_t acc_full[2];
    uint64_t acc_empty[2];
    uint32_t tmem_base;
};

// terms: up to 6 (a part, b part) pairs, 4 bits each (a << 2 | b), issued in the order given
// (least significant pair first).
__global__ void __launch_bounds__(TC_THREADS)
linear_tc_kernel(const __grid_constant__ TcMaps maps, int na, int nb, uint32_t terms, int nterms, int stages,
                 const float* __restrict__ bias, int M, int K, int N, int relu, float* __restrict__ y,
                 float* __restrict__ y_hi, float* __restrict__ y_mid, float* __restrict__ y_lo, int ldy, int ycol0) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* ring = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    const int stage_bytes = na * TC_A_BYTES + nb * TC_B_BYTES;
    TcBarriers* bars = reinterpret_cast<TcBarriers*>(ring + (size_t)stages * stage_bytes);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int n0 = blockIdx.x * TC_BN, m0 = blockIdx.y * TC_BM;
    const int grp = blockIdx.z;  // grouped form: model `grp` of a tree level (x columns, W rows, y columns of its own)
    const int KB = (K + TC_BK - 1) / TC_BK;

    if (warp == 4 && lane == 0) {
        for (int i = 0; i < na; ++i) ptx::tma_prefetch_desc(&maps.a[i]);
        for (int i = 0; i < nb; ++i) ptx::tma_prefetch_desc(&maps.b[i]);
        for (int s = 0; s < stages; ++s) {
            ptx::mbar_init(&bars->full[s], 1);
            ptx::mbar_init(&bars->empty[s], 1);
        }
        for (int b = 0; b < 2; ++b) {
            ptx::mbar_init(&bars->acc_full[b], 1);
            ptx::mbar_init(&bars->acc_empty[b], 4);
        }
        ptx::fence_barrier_init();
    }
    if (warp == 5) {
        ptx::tmem_alloc(&bars->tmem_base, 2 * TC_BN);
        ptx::tmem_relinquish();
    }
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem = bars->tmem_base;

    if (warp == 4) {
        // ===================== TMA producer =====================
        const uint64_t pol_x = ptx::policy_evict_first();   // x tile: read by the N/64 CTAs of this row block
        const uint64_t pol_w = ptx::policy_evict_normal();  // W: re-read by every row block
        uint32_t stage = 0, phase = 0;
        for (int kb = 0; kb < KB; ++kb) {
            ptx::mbar_wait(&bars->empty[stage], phase ^ 1);
            if (ptx::elect_one()) {
                uint8_t* dst = ring + (size_t)stage * stage_bytes;
                ptx::mbar_arrive_expect_tx(&bars->full[stage], stage_bytes);
                for (int i = 0; i < na; ++i)
                    ptx::tma_load_3d(dst + i * TC_A_BYTES, &maps.a[i], kb * TC_BK, m0, grp, &bars->full[stage], pol_x);
                for (int i = 0; i < nb; ++i)
                    ptx::tma_load_3d(dst + na * TC_A_BYTES + i * TC_B_BYTES, &maps.b[i], kb * TC_BK, n0, grp,
                                     &bars->full[stage], pol_w);
            }
            __syncwarp();
            if (++stage == (uint32_t)stages) {
                stage = 0;
                phase ^= 1;
            }
        }
    } else if (warp == 5) {
        // ===================== MMA issuer =====================
        constexpr uint32_t idesc = ptx::umma_idesc_tf32(TC_BM, TC_BN);
        const uint64_t desc0 = ptx::umma_desc_sw128_kmajor(ptx::smem_u32(ring));
        const uint32_t desc0_lo = (uint32_t)desc0;
        const uint64_t desc0_hi = desc0 >> 32;
        uint32_t stage = 0, phase = 0;
        for (int kb = 0; kb < KB; ++kb) {
            const uint32_t buf = kb & 1;
            ptx::mbar_wait(&bars->acc_empty[buf], ((kb >> 1) & 1) ^ 1);
            ptx::mbar_wait(&bars->full[stage], phase);
            ptx::tc_fence_after();
            if (ptx::elect_one()) {
                const uint32_t slo = desc0_lo + ((stage * stage_bytes) >> 4);
                const uint32_t d_tmem = tmem + buf * TC_BN;
                for (int t = 0; t < nterms; ++t) {
                    const uint32_t pair = (terms >> (4 * t)) & 15u;
                    const uint32_t alo = slo + (((pair >> 2) * TC_A_BYTES) >> 4);
                    const uint32_t blo = slo + ((na * TC_A_BYTES + (pair & 3u) * TC_B_BYTES) >> 4);
#pragma unroll
                    for (int k = 0; k < TC_BK / 8; ++k) {
                        const uint64_t a_desc = (desc0_hi << 32) | (uint64_t)(alo + 2 * k);
                        const uint64_t b_desc = (desc0_hi << 32) | (uint64_t)(blo + 2 * k);
                        ptx::mma_tf32_ss(d_tmem, a_desc, b_desc, idesc, (t | k) != 0 ? 1u : 0u);
                    }
                }
                ptx::mma_commit(&bars->empty[stage]);
                ptx::mma_commit(&bars->acc_full[buf]);
            }
            __syncwarp();
            if (++stage == (uint32_t)stages) {
                stage = 0;
                phase ^= 1;
            }
        }
    } else {
        // ===================== epilogue warps (thread == row of x) =====================
        const uint32_t lane_field = static_cast<uint32_t>(warp * 32) << 16;
        float acc[TC_BN];
#pragma unroll
        for (int j = 0; j < TC_BN; ++j) acc[j] = 0.f;
        for (int kb = 0; kb < KB; ++kb) {
            const uint32_t buf = kb & 1;
            ptx::mbar_wait(&bars->acc_full[buf], (kb >> 1) & 1);
            ptx::tc_fence_after();
            uint32_t v0[32], v1[32];
            ptx::tmem_ld32(tmem + lane_field + buf * TC_BN, v0);
            ptx::tmem_ld32(tmem + lane_field + buf * TC_BN + 32, v1);
            ptx::tmem_ld_wait();
            ptx::tc_fence_before();
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive(&bars->acc_empty[buf]);
#pragma unroll
            for (int j = 0; j < 32; ++j) {
                acc[j] += __uint_as_float(v0[j]);  // round-to-nearest fp32 adds of the K-block partials
                acc[32 + j] += __uint_as_float(v1[j]);
            }
        }
        const int r = m0 + warp * 32 + lane;
        if (r < M) {
            const int gcol0 = ycol0 + grp * N;  // first output column of this group
            const float* gbias = bias + (size_t)grp * N;
            const bool vec = (N % 4 == 0) && (ldy % 4 == 0) && (gcol0 % 4 == 0);
#pragma unroll
            for (int j = 0; j < TC_BN; j += 4) {
                float o[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const int c = n0 + j + u;
                    float t = acc[j + u] + (c < N ? gbias[c] : 0.f);
                    if (relu) t = fmaxf(t, 0.f);
                    o[u] = t;
                }
                const size_t off = (size_t)r * ldy + gcol0 + n0 + j;
                if (vec && n0 + j + 3 < N) {
                    if (y) *reinterpret_cast<float4*>(y + off) = make_float4(o[0], o[1], o[2], o[3]);
                    if (y_hi) {
                        float h[4], m[4], l[4];
#pragma unroll
                        for (int u = 0; u < 4; ++u) split3(o[u], h[u], m[u], l[u]);
                        *reinterpret_cast<float4*>(y_hi + off) = make_float4(h[0], h[1], h[2], h[3]);
                        *reinterpret_cast<float4*>(y_mid + off) = make_float4(m[0], m[1], m[2], m[3]);
                        *reinterpret_cast<float4*>(y_lo + off) = make_float4(l[0], l[1], l[2], l[3]);
                    }
                } else {
                    for (int u = 0; u < 4; ++u) {
                        if (n0 + j + u >= N) break;
                        if (y) y[off + u] = o[u];
                        if (y_hi) {
                            float h, m, l;
                            split3(o[u], h, m, l);
                            y_hi[off + u] = h;
                            y_mid[off + u] = m;
                            y_lo[off + u] = l;
                        }
                    }
                }
            }
        }
    }

    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 5) {
        ptx::tc_fence_after();
        ptx::tmem_dealloc(tmem, 2 * TC_BN);
    }
}

typedef CUresult (*TcEncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static TcEncodeTiledFn tc_encode_fn() {
    static TcEncodeTiledFn fn = nullptr;
    if (fn) return fn;
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) != cudaSuccess ||
        qres != cudaDriverEntryPointSuccess)
        return nullptr;
    fn = reinterpret_cast<TcEncodeTiledFn>(p);
    return fn;
}

// `groups` fp32 matrices [rows, K] -- matrix g starts `group_stride` elements after matrix g-1, rows are `ld`
// elements apart -> boxes of 32 fp32 x box_rows rows x 1 group, SWIZZLE_128B, zero fill out of bounds
static bool tc_make_map(CUtensorMap* map, const float* base, int64_t rows, int K, int box_rows, int64_t ld,
                        int groups, int64_t group_stride) {
    TcEncodeTiledFn encode = tc_encode_fn();
    if (!encode) return false;
    const cuuint64_t gdim[3] = {(cuuint64_t)K, (cuuint64_t)rows, (cuuint64_t)groups};
    const cuuint64_t gstride[2] = {(cuuint64_t)ld * sizeof(float),
                                   (cuuint64_t)(groups > 1 ? group_stride : ld * rows) * sizeof(float)};
    const cuuint32_t box[3] = {(cuuint32_t)TC_BK, (cuuint32_t)box_rows, 1};
    const cuuint32_t estride[3] = {1, 1, 1};
    return encode(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<float*>(base), gdim, gstride, box, estride,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

}  // namespace lmi

extern "C" int lmi_split_tf32(const void* x, int x_is_f16, int64_t count, float* hi, float* mid, float* lo,
                              void* stream) {
    if (!x || !hi || count < 0 || ((mid == nullptr) != (lo == nullptr))) return LMI_E_BADARG;
    if (count == 0) return 0;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int64_t blocks = (count + 255) / 256;
    if (blocks > 0x7fffffff) return LMI_E_UNSUPPORTED;
    if (x_is_f16)
        lmi::split_tf32_kernel<__half><<<(unsigned)blocks, 256, 0, st>>>(static_cast<const __half*>(x), count, hi, mid, lo);
    else
        lmi::split_tf32_kernel<float><<<(unsigned)blocks, 256, 0, st>>>(static_cast<const float*>(x), count, hi, mid, lo);
    return static_cast<int>(cudaGetLastError());
}

extern "C" int lmi_linear_tc(const float* x_hi, const float* x_mid, const float* x_lo, const float* w_hi,
                             const float* w_mid, const float* w_lo, const float* bias, int nq, int in_dim,
                             int out_dim, int relu, float* y, float* y_hi, float* y_mid, float* y_lo, void* stream) {
    return lmi_linear_tc_grouped(x_hi, x_mid, x_lo, in_dim, w_hi, w_mid, w_lo, bias, nq, in_dim, out_dim, 1, relu, y,
                                 y_hi, y_mid, y_lo, out_dim, 0, stream);
}

extern "C" int lmi_linear_tc_grouped(const float* x_hi, const float* x_mid, const float* x_lo, int ldx,
                                     const float* w_hi, const float* w_mid, const float* w_lo, const float* bias,
                                     int nq, int in_dim, int out_dim, int groups, int relu, float* y, float* y_hi,
                                     float* y_mid, float* y_lo, int ldy, int y_col0, void* stream) {
    if (!x_hi || !w_hi || !bias || nq < 0 || in_dim <= 0 || out_dim <= 0) return LMI_E_BADARG;
    if (groups <= 0 || groups > 65535 || ldx < groups * in_dim || y_col0 < 0 || ldy < y_col0 + groups * out_dim)
        return LMI_E_BADARG;
    if (ldx % 4 != 0) return LMI_E_UNSUPPORTED;
    if ((x_mid == nullptr) != (x_lo == nullptr) || (w_mid == nullptr) != (w_lo == nullptr)) return LMI_E_BADARG;
    if (!y && !y_hi) return LMI_E_BADARG;
    if (y_hi && (!y_mid || !y_lo)) return LMI_E_BADARG;
    if (in_dim % 4 != 0) return LMI_E_UNSUPPORTED;  // TMA row pitch must be a multiple of 16 bytes
    if (nq == 0) return 0;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const float* xa[3] = {x_hi, x_mid, x_lo};
    const float* wb[3] = {w_hi, w_mid, w_lo};
    const int na = x_mid ? 3 : 1, nb = w_mid ? 3 : 1;
    for (int i = 0; i < na; ++i)
        if (reinterpret_cast<uintptr_t>(xa[i]) & 15) return LMI_E_BADARG;
    for (int i = 0; i < nb; ++i)
        if (reinterpret_cast<uintptr_t>(wb[i]) & 15) return LMI_E_BADARG;

    lmi::TcMaps maps;
    for (int i = 0; i < na; ++i)
        if (!lmi::tc_make_map(&maps.a[i], xa[i], nq, in_dim, lmi::TC_BM, ldx, groups, in_dim)) return LMI_E_DRIVER;
    for (int i = 0; i < nb; ++i)
        if (!lmi::tc_make_map(&maps.b[i], wb[i], out_dim, in_dim, lmi::TC_BN, in_dim, groups,
                              (int64_t)out_dim * in_dim))
            return LMI_E_DRIVER;
    for (int i = na; i < 3; ++i) maps.a[i] = maps.a[0];
    for (int i = nb; i < 3; ++i) maps.b[i] = maps.b[0];

    // part pairs (a, b) with a + b <= 2 (weight >= 2^-22 of the leading product), least significant first
    uint32_t terms = 0;
    int nterms = 0;
    for (int w = 2; w >= 0; --w)
        for (int a = 0; a <= w; ++a) {
            const int b = w - a;
            if (a < na && b < nb) terms |= (uint32_t)((a << 2) | b) << (4 * nterms++);
        }

    const int stage_bytes = na * lmi::TC_A_BYTES + nb * lmi::TC_B_BYTES;
    int stages = (200 * 1024) / stage_bytes;
    if (stages > lmi::TC_MAX_STAGES) stages = lmi::TC_MAX_STAGES;
    if (stage_bytes <= 48 * 1024 && stages > 2) stages = 2;  // small stages: leave room for two CTAs per SM
    if (stages < 2) return LMI_E_UNSUPPORTED;
    const size_t smem = (size_t)stages * stage_bytes + 1024 + sizeof(lmi::TcBarriers);
    cudaError_t e = cudaFuncSetAttribute(lmi::linear_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
    dim3 grid((out_dim + lmi::TC_BN - 1) / lmi::TC_BN, (nq + lmi::TC_BM - 1) / lmi::TC_BM, groups);
    lmi::linear_tc_kernel<<<grid, lmi::TC_THREADS, smem, st>>>(maps, na, nb, terms, nterms, stages, bias, nq, in_dim,
                                                               out_dim, relu, y, y_hi, y_mid, y_lo, ldy, y_col0);
    return static_cast<int>(cudaGetLastError());
}
