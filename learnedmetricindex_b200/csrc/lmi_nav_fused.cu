// ONE launch for the 1-level navigation step -- NeuralNetwork.predict_proba (search/li/model.py:226-241:
// Linear -> ReLU -> Linear -> softmax -> topk) and the slice bucket_order = classes[:, :B]
// (search/li/LearnedIndex.py:209-214) -- for the 2-Linear-layer models of the reference's zoo
// (search/li/model.py:29-79: MLP, MLP-2, MLP-3, MLP-4).
//
// Same arithmetic as the 4-launch path (lmi_split_tf32 -> lmi_linear_tc x 2 -> lmi_softmax_topn, which stays
// as the check and as the path for every other shape): tcgen05.mma kind::tf32 on SPLIT operands
// (hi + mid + lo, each part exact in tf32; all part pairs of weight >= 2^-22), one K-block of 32 summed in
// TMEM, K-blocks added in fp32 registers with round-to-nearest -- fp32-quality logits.  What the fusion
// removes: the fp32 copy of the queries (31 MB at 10k x 768), the three parts of the hidden activations
// (3 x 20 MB written and re-read), the logits round trip, and three launches.
//
// Grid = (H / 64 hidden chunks, ceil(nq / 128) query tiles).  CTA (chunk j, tile m):
//   phase 1  hid[128, 64] = relu(x[m] . W1[64 j .. 64 j + 63]^T + b1): queries come straight from the
//            fp16 matrix -- four converter warps widen each 128 x 32 block into the SWIZZLE_128B A tile of
//            the pipeline stage (fp16 is exact in one tf32 part) while the producer warp streams the three
//            parts of W1 by TMA; 3 MMAs per k-step; the epilogue warps (thread == query row) add the K-block
//            partials, apply bias + ReLU, split hid into hi/mid/lo and write the parts into shared memory
//            in the A-operand layout of phase 2 -- the hidden activations never leave the SM;
//   phase 2  partial logits[128, C] = hid . W2[:, 64 j .. 64 j + 63]^T (K = 64: two K-blocks, 6 part pairs),
//            W2's parts for this chunk pre-loaded by TMA at kernel start; written to an L2-resident scratch;
//   finish   the LAST CTA of the tile to arrive (atomic ticket) sums the H/64 partials in chunk order
//            (deterministic), adds b2, and does softmax + top-B from shared memory, one thread per query row.
// Roles: warps 0-3 epilogue, warp 4 TMA producer, warp 5 MMA issuer, warps 6-9 query converters.
#include <cuda.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

#include "../../include/lmi_b200.h"
#include "lmi_ptx.cuh"

namespace lmi {

constexpr int NF_BM = 128;                 // queries per tile (TMEM lanes)
constexpr int NF_HC = 64;                  // hidden units per CTA
constexpr int NF_BK = 32;                  // fp32 per 128-byte shared-memory row == one K-block
constexpr int NF_A_BYTES = NF_BM * 128;    // one 128-row A tile (queries / hidden part): 16 KB
constexpr int NF_W1_BYTES = NF_HC * 128;   // one part of a W1 K-block: 8 KB
constexpr int NF_STAGE_BYTES = NF_A_BYTES + 3 * NF_W1_BYTES;  // 40 KB
constexpr int NF_STAGES = 3;
constexpr int NF_RING_BYTES = NF_STAGES * NF_STAGE_BYTES;     // 120 KB; phase 2 reuses it for the hidden parts
constexpr int NF_CP_MAX = 128;             // classes, padded to a multiple of 16 (MMA N)
constexpr int NF_W2_TILE_BYTES = NF_CP_MAX * 128;             // one (part, K-block) tile of W2: 16 KB
constexpr int NF_W2_BYTES = 6 * NF_W2_TILE_BYTES;             // 96 KB
constexpr int NF_THREADS = 320;
constexpr int NF_EPI_THREADS = 128;
constexpr int NF_CONV_WARP0 = 6, NF_CONV_WARPS = 4;
constexpr uint32_t NF_TMEM_COLS = 512;
constexpr uint32_t NF_ACC2_COL = 128;      // phase-1 partial tiles: [0,64) [64,128); phase-2: [128,256) [256,384)

struct NfBarriers {
    uint64_t full[NF_STAGES];    // W1 parts of the stage have landed (TMA)
    uint64_t conv[NF_STAGES];    // the converters have written the stage's A tile
    uint64_t empty[NF_STAGES];   // the MMAs reading the stage are done
    uint64_t acc1_full[2];
    uint64_t acc1_empty[2];
    uint64_t w2_full;
    uint64_t h_ready;
    uint64_t acc2_full[2];
    uint32_t tmem_base;
    int32_t ticket;
};
constexpr size_t NF_SMEM_BYTES = (size_t)NF_RING_BYTES + NF_W2_BYTES + 1024 /*align*/ + 256 /*barriers*/;
static_assert(sizeof(NfBarriers) <= 256, "barrier block too large");
static_assert(2 * 3 * NF_A_BYTES <= NF_RING_BYTES, "hidden parts must fit the ring they alias");
static_assert((size_t)NF_BM * (NF_CP_MAX + 1) * 4 <= NF_RING_BYTES, "logit rows must fit the ring they alias");

struct NfMaps {
    CUtensorMap w1[3];  // [H, d_in] fp32, box 32 x 64 rows
    CUtensorMap w2[3];  // [C, H] fp32, box 32 x Cp rows
};

__device__ __forceinline__ float nf_tf32_trunc(float v) { return __uint_as_float(__float_as_uint(v) & 0xffffe000u); }

// byte offset of the 16-byte chunk `c4` (0..7) of row r inside a 128-row SWIZZLE_128B K-major tile
__device__ __forceinline__ uint32_t nf_swz(int r, int c4) {
    return (uint32_t)((r >> 3) * 1024 + (r & 7) * 128 + ((c4 ^ (r & 7)) << 4));
}
__device__ __forceinline__ void nf_sts128(uint32_t addr, float a, float b, float c, float d) {
    asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}
__device__ __forceinline__ void nf_named_barrier(int id, int threads) {
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(threads) : "memory");
}

__global__ void __launch_bounds__(NF_THREADS, 1)
nav_fused_kernel(const __grid_constant__ NfMaps maps, const __half* __restrict__ x, int nq, int d_in, int H, int C,
                 int Cp, const float* __restrict__ b1, const float* __restrict__ b2, float* __restrict__ scratch,
                 int32_t* __restrict__ tile_counter, int B, int32_t* __restrict__ order, float* __restrict__ probs,
                 float* __restrict__ logits_out) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* ring = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* w2s = ring + NF_RING_BYTES;
    NfBarriers* bars = reinterpret_cast<NfBarriers*>(w2s + NF_W2_BYTES);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int chunk = blockIdx.x, tile = blockIdx.y, chunks = gridDim.x;
    const int m0 = tile * NF_BM;
    const int KB1 = (d_in + NF_BK - 1) / NF_BK;
    const uint32_t w2_tile_bytes = (uint32_t)Cp * 128u;

    if (warp == 4 && lane == 0) {
        for (int i = 0; i < 3; ++i) {
            ptx::tma_prefetch_desc(&maps.w1[i]);
            ptx::tma_prefetch_desc(&maps.w2[i]);
        }
        for (int s = 0; s < NF_STAGES; ++s) {
            ptx::mbar_init(&bars->full[s], 1);
            ptx::mbar_init(&bars->conv[s], NF_CONV_WARPS);
            ptx::mbar_init(&bars->empty[s], 1);
        }
        for (int b = 0; b < 2; ++b) {
            ptx::mbar_init(&bars->acc1_full[b], 1);
            ptx::mbar_init(&bars->acc1_empty[b], 4);
            ptx::mbar_init(&bars->acc2_full[b], 1);
        }
        ptx::mbar_init(&bars->w2_full, 1);
        ptx::mbar_init(&bars->h_ready, 4);
        ptx::fence_barrier_init();
    }
    if (warp == 5) {
        ptx::tmem_alloc(&bars->tmem_base, NF_TMEM_COLS);
        ptx::tmem_relinquish();
    }
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem = bars->tmem_base;

    if (warp == 4) {
        // ===================== TMA producer: W2's parts for this chunk (once), W1's parts per K-block ==========
        const uint64_t pol = ptx::policy_evict_normal();  // every query tile re-reads the weights: keep them in L2
        if (ptx::elect_one()) {
            ptx::mbar_arrive_expect_tx(&bars->w2_full, 6u * w2_tile_bytes);
            for (int p = 0; p < 3; ++p)
                for (int kk = 0; kk < 2; ++kk)
                    ptx::tma_load_2d(w2s + (size_t)(p * 2 + kk) * NF_W2_TILE_BYTES, &maps.w2[p],
                                     chunk * NF_HC + kk * NF_BK, 0, &bars->w2_full, pol);
        }
        __syncwarp();
        uint32_t stage = 0, phase = 0;
        for (int kb = 0; kb < KB1; ++kb) {
            ptx::mbar_wait(&bars->empty[stage], phase ^ 1);
            if (ptx::elect_one()) {
                uint8_t* dst = ring + (size_t)stage * NF_STAGE_BYTES + NF_A_BYTES;
                ptx::mbar_arrive_expect_tx(&bars->full[stage], 3 * NF_W1_BYTES);
                for (int p = 0; p < 3; ++p)
                    ptx::tma_load_2d(dst + p * NF_W1_BYTES, &maps.w1[p], kb * NF_BK, chunk * NF_HC, &bars->full[stage],
                                     pol);
            }
            __syncwarp();
            if (++stage == NF_STAGES) {
                stage = 0;
                phase ^= 1;
            }
        }
    } else if (warp >= NF_CONV_WARP0) {
        // ===================== query converters: fp16 rows -> fp32 (one exact tf32 part) A tiles ================
        // thread t: 16-byte chunk (t & 3) of rows (t >> 2) + 32 i, i = 0..3, of every K-block; the next K-block's
        // four loads are in flight while this one is converted and stored
        const int t = threadIdx.x - NF_CONV_WARP0 * 32;
        const int ch = t & 3, r0 = t >> 2;
        auto load_block = [&](uint4 (&v)[4], int kb) {
            const int col = kb * NF_BK + ch * 8;
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int r = m0 + r0 + 32 * i;
                v[i] = (r < nq && col < d_in) ? __ldg(reinterpret_cast<const uint4*>(x + (size_t)r * d_in + col))
                                               : make_uint4(0u, 0u, 0u, 0u);
            }
        };
        // four K-blocks of loads in flight per thread (an L2 round trip is ~3 MMA K-block times)
        uint4 v0[4], v1[4], v2[4], v3[4];
        load_block(v0, 0);
        load_block(v1, 1);
        load_block(v2, 2);
        load_block(v3, 3);
        uint32_t stage = 0, phase = 0;
        auto convert_block = [&](const uint4 (&cur)[4]) {
            ptx::mbar_wait(&bars->empty[stage], phase ^ 1);
            const uint32_t a_tile = ptx::smem_u32(ring + (size_t)stage * NF_STAGE_BYTES);
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int r = r0 + 32 * i;
                const __half2* hp = reinterpret_cast<const __half2*>(&cur[i]);
                const float2 f0 = __half22float2(hp[0]), f1 = __half22float2(hp[1]);
                const float2 f2 = __half22float2(hp[2]), f3 = __half22float2(hp[3]);
                nf_sts128(a_tile + nf_swz(r, ch * 2), f0.x, f0.y, f1.x, f1.y);
                nf_sts128(a_tile + nf_swz(r, ch * 2 + 1), f2.x, f2.y, f3.x, f3.y);
            }
            ptx::fence_proxy_async();  // generic-proxy stores -> visible to the tensor core's async-proxy reads
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive(&bars->conv[stage]);
            if (++stage == NF_STAGES) {
                stage = 0;
                phase ^= 1;
            }
        };
        for (int kb = 0; kb < KB1; kb += 4) {
            convert_block(v0);
            load_block(v0, kb + 4);
            if (kb + 1 < KB1) {
                convert_block(v1);
                load_block(v1, kb + 5);
            }
            if (kb + 2 < KB1) {
                convert_block(v2);
                load_block(v2, kb + 6);
            }
            if (kb + 3 < KB1) {
                convert_block(v3);
                load_block(v3, kb + 7);
            }
        }
    } else if (warp == 5) {
        // ===================== MMA issuer =====================
        const uint64_t desc0 = ptx::umma_desc_sw128_kmajor(ptx::smem_u32(ring));
        const uint32_t desc0_lo = (uint32_t)desc0;
        const uint64_t desc0_hi = desc0 >> 32;
        {   // phase 1: x (one part) x W1 (three parts), least significant part first
            constexpr uint32_t idesc = ptx::umma_idesc_tf32(NF_BM, NF_HC);
            uint32_t stage = 0, phase = 0;
            for (int kb = 0; kb < KB1; ++kb) {
                const uint32_t buf = kb & 1;
                ptx::mbar_wait(&bars->acc1_empty[buf], ((kb >> 1) & 1) ^ 1);
                ptx::mbar_wait(&bars->full[stage], phase);
                ptx::mbar_wait(&bars->conv[stage], phase);
                ptx::tc_fence_after();
                if (ptx::elect_one()) {
                    const uint32_t slo = desc0_lo + ((stage * NF_STAGE_BYTES) >> 4);
                    const uint32_t d_tmem = tmem + buf * NF_HC;
#pragma unroll
                    for (int p = 2; p >= 0; --p) {
                        const uint32_t blo = slo + ((NF_A_BYTES + p * NF_W1_BYTES) >> 4);
#pragma unroll
                        for (int k = 0; k < NF_BK / 8; ++k) {
                            const uint64_t a_desc = (desc0_hi << 32) | (uint64_t)(slo + 2 * k);
                            const uint64_t b_desc = (desc0_hi << 32) | (uint64_t)(blo + 2 * k);
                            ptx::mma_tf32_ss(d_tmem, a_desc, b_desc, idesc, (p != 2 || k != 0) ? 1u : 0u);
                        }
                    }
                    ptx::mma_commit(&bars->empty[stage]);
                    ptx::mma_commit(&bars->acc1_full[buf]);
                }
                __syncwarp();
                if (++stage == NF_STAGES) {
                    stage = 0;
                    phase ^= 1;
                }
            }
        }
        {   // phase 2: hid (three parts, written by the epilogue warps over the ring) x W2 (three parts)
            const uint32_t idesc = ptx::umma_idesc_tf32(NF_BM, Cp);
            const uint32_t wlo = desc0_lo + (NF_RING_BYTES >> 4);
            ptx::mbar_wait(&bars->w2_full, 0);
            ptx::mbar_wait(&bars->h_ready, 0);
            ptx::tc_fence_after();
            if (ptx::elect_one()) {
                for (int kk = 0; kk < 2; ++kk) {
                    const uint32_t d_tmem = tmem + NF_ACC2_COL + kk * NF_CP_MAX;
                    bool first = true;
                    for (int w = 2; w >= 0; --w) {          // part pairs (a, b) with a + b == w, smallest first
                        for (int a = 0; a <= w; ++a) {
                            const int b = w - a;
                            const uint32_t alo = desc0_lo + (((a * 2 + kk) * NF_A_BYTES) >> 4);
                            const uint32_t blo = wlo + (((b * 2 + kk) * NF_W2_TILE_BYTES) >> 4);
#pragma unroll
                            for (int k = 0; k < NF_BK / 8; ++k) {
                                const uint64_t a_desc = (desc0_hi << 32) | (uint64_t)(alo + 2 * k);
                                const uint64_t b_desc = (desc0_hi << 32) | (uint64_t)(blo + 2 * k);
                                ptx::mma_tf32_ss(d_tmem, a_desc, b_desc, idesc, first ? 0u : 1u);
                                first = false;
                            }
                        }
                    }
                    ptx::mma_commit(&bars->acc2_full[kk]);
                }
            }
            __syncwarp();
        }
    } else {
        // ===================== epilogue warps (thread == query row) =====================
        const uint32_t lane_field = static_cast<uint32_t>(warp * 32) << 16;
        const int row = warp * 32 + lane;
        {   // phase 1: add the K-block partials (round-to-nearest fp32), then bias + ReLU + split into the h tiles
            float acc[NF_HC];
#pragma unroll
            for (int j = 0; j < NF_HC; ++j) acc[j] = 0.f;
            for (int kb = 0; kb < KB1; ++kb) {
                const uint32_t buf = kb & 1;
                ptx::mbar_wait(&bars->acc1_full[buf], (kb >> 1) & 1);
                ptx::tc_fence_after();
                uint32_t v0[32], v1[32];
                ptx::tmem_ld32(tmem + lane_field + buf * NF_HC, v0);
                ptx::tmem_ld32(tmem + lane_field + buf * NF_HC + 32, v1);
                ptx::tmem_ld_wait();
                ptx::tc_fence_before();
                __syncwarp();
                if (lane == 0) ptx::mbar_arrive(&bars->acc1_empty[buf]);
#pragma unroll
                for (int j = 0; j < 32; ++j) {
                    acc[j] += __uint_as_float(v0[j]);
                    acc[32 + j] += __uint_as_float(v1[j]);
                }
            }
            // every MMA of phase 1 has completed (its last accumulator was just read), so the ring is free:
            // hid part p, K-block kk lives at ring + (p * 2 + kk) * 16 KB in the A-operand layout
            const uint32_t ring_u32 = ptx::smem_u32(ring);
            const float* bias = b1 + chunk * NF_HC;
#pragma unroll
            for (int kk = 0; kk < 2; ++kk) {
#pragma unroll
                for (int c4 = 0; c4 < 8; ++c4) {
                    float hi[4], mid[4], lo[4];
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        const int j = kk * 32 + c4 * 4 + e;
                        const float h = fmaxf(acc[j] + __ldg(bias + j), 0.f);
                        hi[e] = nf_tf32_trunc(h);
                        const float r = h - hi[e];
                        mid[e] = nf_tf32_trunc(r);
                        lo[e] = r - mid[e];
                    }
                    const uint32_t off = nf_swz(row, c4);
                    nf_sts128(ring_u32 + (0 * 2 + kk) * NF_A_BYTES + off, hi[0], hi[1], hi[2], hi[3]);
                    nf_sts128(ring_u32 + (1 * 2 + kk) * NF_A_BYTES + off, mid[0], mid[1], mid[2], mid[3]);
                    nf_sts128(ring_u32 + (2 * 2 + kk) * NF_A_BYTES + off, lo[0], lo[1], lo[2], lo[3]);
                }
            }
            ptx::fence_proxy_async();
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive(&bars->h_ready);
        }
        // phase 2: partial logits of this hidden chunk = sum of its two K-block partials -> scratch (row fastest)
        float* part = scratch + ((size_t)tile * chunks + chunk) * (size_t)Cp * NF_BM;
        for (int c0 = 0; c0 < Cp; c0 += 32) {
            uint32_t u0[32], u1[32];
            if (c0 == 0) {
                ptx::mbar_wait(&bars->acc2_full[0], 0);
                ptx::mbar_wait(&bars->acc2_full[1], 0);
                ptx::tc_fence_after();
            }
            ptx::tmem_ld32(tmem + lane_field + NF_ACC2_COL + c0, u0);
            ptx::tmem_ld32(tmem + lane_field + NF_ACC2_COL + NF_CP_MAX + c0, u1);
            ptx::tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 32; ++j)
                if (c0 + j < Cp) part[(size_t)(c0 + j) * NF_BM + row] = __uint_as_float(u0[j]) + __uint_as_float(u1[j]);
        }
        ptx::tc_fence_before();
        // ticket: the last CTA of this query tile to arrive finishes the job (below, with every warp)
        __threadfence();
        nf_named_barrier(1, NF_EPI_THREADS);
        if (threadIdx.x == 0) {
            // release / acquire in the thread that owns the atomic: the barrier above made every epilogue
            // thread's partials visible to this thread, the fence before the atomic publishes them (fences are
            // cumulative) together with the ticket, and the fence after it orders the finishing CTA's reads of
            // the OTHER CTAs' partials behind the ticket value that proved them written.  The __syncthreads
            // below hands that ordering to the rest of the CTA.
            __threadfence();
            const int32_t t = atomicAdd(tile_counter + tile, 1);
            __threadfence();
            bars->ticket = t;
        }
    }

    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 5) {
        ptx::tc_fence_after();
        ptx::tmem_dealloc(tmem, NF_TMEM_COLS);
    }
    if (bars->ticket != chunks - 1) return;

    // ===================== finish (last CTA of the tile; all MMAs are done: the ring is free) ==================
    __threadfence();
    float* lall = reinterpret_cast<float*>(ring);  // logits[row][c], row pitch NF_CP_MAX + 1 (conflict-free rows)
    {
        // logits = b2 + sum over hidden chunks IN CHUNK ORDER (the result does not depend on which CTA came last);
        // element e = c * 128 + row of a partial; 8 elements x all chunks of loads in flight per thread
        const float* tile_parts = scratch + (size_t)tile * chunks * (size_t)Cp * NF_BM;
        const int total = C * NF_BM;
        const size_t part_stride = (size_t)Cp * NF_BM;
        constexpr int U = 8;
        for (int base = threadIdx.x; base < total; base += U * NF_THREADS) {
            float v[U];
#pragma unroll
            for (int u = 0; u < U; ++u) v[u] = 0.f;
#pragma unroll 4
            for (int j = 0; j < chunks; ++j) {
                float t[U];
#pragma unroll
                for (int u = 0; u < U; ++u) {
                    const int e = base + u * NF_THREADS;
                    t[u] = e < total ? __ldcg(tile_parts + (size_t)j * part_stride + e) : 0.f;
                }
#pragma unroll
                for (int u = 0; u < U; ++u) v[u] += t[u];
            }
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const int e = base + u * NF_THREADS;
                if (e < total) {
                    const int c = e >> 7, r = e & (NF_BM - 1);
                    lall[(size_t)r * (NF_CP_MAX + 1) + c] = v[u] + __ldg(b2 + c);
                }
            }
        }
    }
    __syncthreads();
    if (logits_out) {
        const int rows_here = min(NF_BM, nq - m0);
        for (int idx = threadIdx.x; idx < rows_here * C; idx += NF_THREADS) {
            const int r = idx / C, c = idx - r * C;
            logits_out[(size_t)(m0 + r) * C + c] = lall[(size_t)r * (NF_CP_MAX + 1) + c];
        }
    }
    // ready for the next launch.  The counters reset themselves, which is only sound with ONE launch in flight per
    // workspace: launches that share a workspace must be ordered on a stream (the host side keeps one workspace
    // per batch size and issues every search of an index on the caller's current stream).
    if (threadIdx.x == 0) tile_counter[tile] = 0;
    if (order && threadIdx.x < NF_BM && m0 + (int)threadIdx.x < nq) {
        // softmax (model.py:238) and the B most probable classes, ties to the lower class index -- the same
        // formulas as softmax_topn_kernel; one thread per query row
        const int q = m0 + threadIdx.x;
        float* lrow = lall + (size_t)threadIdx.x * (NF_CP_MAX + 1);
        float mx = -INFINITY;
        for (int c = 0; c < C; ++c) mx = fmaxf(mx, lrow[c]);
        float sum = 0.f;
        for (int c = 0; c < C; ++c) {
            const float e = expf(lrow[c] - mx);
            lrow[c] = e;
            sum += e;
        }
        for (int c = 0; c < C; ++c) lrow[c] = lrow[c] / sum;
        for (int r = 0; r < B; ++r) {
            float best = -1.f;
            int arg = -1;
            for (int c = 0; c < C; ++c) {
                const float v = lrow[c];
                if (v > best) {
                    best = v;
                    arg = c;
                }
            }
            order[(size_t)q * B + r] = arg;
            probs[(size_t)q * B + r] = best;
            if (arg >= 0) lrow[arg] = -2.f;
        }
    }
}
typedef CUresult (*NfEncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static NfEncodeTiledFn nf_encode_fn() {
    static NfEncodeTiledFn fn = nullptr;
    if (fn) return fn;
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) != cudaSuccess ||
        qres != cudaDriverEntryPointSuccess)
        return nullptr;
    fn = reinterpret_cast<NfEncodeTiledFn>(p);
    return fn;
}

// fp32 matrix [rows, cols] row-major -> boxes of 32 fp32 x box_rows rows, SWIZZLE_128B, zero fill out of bounds
static bool nf_make_map(CUtensorMap* map, const float* base, int rows, int cols, int box_rows) {
    NfEncodeTiledFn encode = nf_encode_fn();
    if (!encode) return false;
    const cuuint64_t gdim[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
    const cuuint64_t gstride[1] = {(cuuint64_t)cols * sizeof(float)};
    const cuuint32_t box[2] = {(cuuint32_t)NF_BK, (cuuint32_t)box_rows};
    const cuuint32_t estride[2] = {1, 1};
    return encode(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(base), gdim, gstride, box, estride,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

}  // namespace lmi

extern "C" int lmi_mlp_topn_supported(int x_is_f16, int in_dim, int hidden, int C, int B) {
    return x_is_f16 && in_dim > 0 && in_dim % 8 == 0 && hidden >= lmi::NF_HC && hidden % lmi::NF_HC == 0 &&
           hidden <= 1024 && C > 0 && C <= lmi::NF_CP_MAX && B >= 0 && B <= C;
}

extern "C" int64_t lmi_mlp_topn_workspace_bytes(int nq, int hidden, int C) {
    if (nq <= 0 || hidden <= 0 || C <= 0) return 0;
    const int64_t tiles = (nq + lmi::NF_BM - 1) / lmi::NF_BM;
    const int64_t Cp = (C + 15) / 16 * 16;
    const int64_t scratch = tiles * (hidden / lmi::NF_HC) * Cp * lmi::NF_BM * 4;
    return (scratch + 255) / 256 * 256 + tiles * 4;
}

extern "C" int lmi_mlp_topn(const void* x_f16, int nq, int in_dim, const float* w1_hi, const float* w1_mid,
                            const float* w1_lo, const float* b1, int hidden, const float* w2_hi, const float* w2_mid,
                            const float* w2_lo, const float* b2, int C, int B, void* workspace, int32_t* bucket_order,
                            float* probs, float* logits, void* stream) {
    if (!x_f16 || !w1_hi || !w1_mid || !w1_lo || !b1 || !w2_hi || !w2_mid || !w2_lo || !b2 || !workspace)
        return LMI_E_BADARG;
    if (nq < 0 || B < 0 || (B > 0 && (!bucket_order || !probs)) || (B == 0 && !logits)) return LMI_E_BADARG;
    if (!lmi_mlp_topn_supported(1, in_dim, hidden, C, B)) return LMI_E_UNSUPPORTED;
    if ((reinterpret_cast<uintptr_t>(x_f16) & 15) || (reinterpret_cast<uintptr_t>(workspace) & 15)) return LMI_E_BADARG;
    if (nq == 0) return 0;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int Cp = (C + 15) / 16 * 16;
    const int chunks = hidden / lmi::NF_HC;
    const int tiles = (nq + lmi::NF_BM - 1) / lmi::NF_BM;
    if (tiles > 65535) return LMI_E_UNSUPPORTED;

    lmi::NfMaps maps;
    const float* w1[3] = {w1_hi, w1_mid, w1_lo};
    const float* w2[3] = {w2_hi, w2_mid, w2_lo};
    for (int i = 0; i < 3; ++i) {
        if ((reinterpret_cast<uintptr_t>(w1[i]) & 15) || (reinterpret_cast<uintptr_t>(w2[i]) & 15)) return LMI_E_BADARG;
        if (!lmi::nf_make_map(&maps.w1[i], w1[i], hidden, in_dim, lmi::NF_HC)) return LMI_E_DRIVER;
        if (!lmi::nf_make_map(&maps.w2[i], w2[i], C, hidden, Cp)) return LMI_E_DRIVER;
    }
    int dev = 0;
    cudaGetDevice(&dev);
    static bool attr_set[64] = {};
    if (dev < 0 || dev >= 64 || !attr_set[dev]) {
        cudaError_t e = cudaFuncSetAttribute(lmi::nav_fused_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             (int)lmi::NF_SMEM_BYTES);
        if (e != cudaSuccess) return (int)e;
        if (dev >= 0 && dev < 64) attr_set[dev] = true;
    }
    float* scratch = static_cast<float*>(workspace);
    const int64_t scratch_bytes = ((int64_t)tiles * chunks * Cp * lmi::NF_BM * 4 + 255) / 256 * 256;
    int32_t* counters = reinterpret_cast<int32_t*>(static_cast<char*>(workspace) + scratch_bytes);
    dim3 grid(chunks, tiles);
    lmi::nav_fused_kernel<<<grid, lmi::NF_THREADS, lmi::NF_SMEM_BYTES, st>>>(
        maps, static_cast<const __half*>(x_f16), nq, in_dim, hidden, C, Cp, b1, b2, scratch, counters, B, bucket_order,
        probs, logits);
    return static_cast<int>(cudaGetLastError());
}
