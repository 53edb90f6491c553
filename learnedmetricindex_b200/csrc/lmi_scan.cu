// Candidate distance scan with a fused per-query top-10 (replaces groupby + .loc gather +
// faiss.knn(..., METRIC_INNER_PRODUCT), search/li/LearnedIndex.py:350-365).
//
// Work item = (bucket, row range of <= range_len rows, tile of <= 128 probing queries).
// Persistent CTAs (one per SM) draw items IN ORDER from a global counter; the items of one
// (bucket, range) -- one per query tile -- are adjacent in the list, so they run at the same
// time on different SMs and the range is read from HBM once (the other tiles hit L2).
// Per item:
//   * the 128 query rows are gathered straight from the fp16 query matrix into TENSOR
//     MEMORY (tcgen05.st), where they stay as the A operand for the whole range
//     (K = d <= 768 fp16 -> <= 384 TMEM columns); the queries are never re-streamed;
//   * the bucket's rows stream HBM/L2 -> shared memory by TMA (64 rows x 128 B boxes,
//     SWIZZLE_128B, 4 boxes = 32 KB per pipeline stage, ring of 6 stages) as the B operand;
//   * tcgen05.mma kind::f16 (fp16 x fp16 -> fp32) M=128, N=64, K=16 accumulates a
//     128 x 64 score tile in TMEM, double buffered (2 x 64 columns).  MMA issue turned out
//     to be synchronous with the tensor pipe, so two issuer warps alternate by tile and
//     hide each other's barrier round trips;
//   * eight epilogue warps read the tile (tcgen05.ld, thread == query row, warp == lane
//     quarter x column half) and keep a threshold-gated sorted top-10 in registers; the
//     score matrix never leaves the SM.
// Roles: warps 0-7 epilogue (+ query gather), warp 8 TMA producer + item scheduler,
// warps 9-10 MMA issuers.
//
// Kernels in this file (lmi_scan_topk's `impl`; all write bit-identical merged results):
//   scan_topk_tc_kernel<64, false>   impl 0, the product path described above, plus the gate the ranges of a
//                                    bucket share through global memory (see the epilogue);
//   scan_topk_tc_kernel<128, false>  impl 3: N = 128 score tiles, ONE accumulator (TMEM: 384 query columns + 128);
//   scan_topk_tc_kernel<128, true>   impl 4: N = 128 tiles, TWO accumulators -- the queries' k-blocks 8..11 live in
//                                    shared memory and enter their MMAs through a descriptor;
//   scan_topk_pair_kernel            impl 2: clusters of two CTAs, bucket rows delivered by TMA multicast;
//   scan_topk_simt_kernel            impl 1: CUDA-core verification kernel (any d; the path for d > 768).
// Measured against each other in profiles/r02_scan_lab.md; DESIGN.md section 3.3 says why impl 0 stays the default.
#include <cuda.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <stdlib.h>

#include "../../include/lmi_b200.h"
#include "lmi_ptx.cuh"

namespace lmi {

constexpr int SCAN_M = 128;        // queries per item (TMEM lanes)
constexpr int SCAN_N = 64;         // bucket rows per MMA tile
constexpr int SCAN_KBLK = 64;      // fp16 per 128-byte shared-memory row
constexpr int SCAN_BOX_BYTES = SCAN_N * 128;  // one TMA box: 64 rows x 64 fp16
constexpr int SCAN_BOXES = 4;                  // boxes (k-blocks) per pipeline stage: one barrier round-trip per 16 MMAs
constexpr int SCAN_STAGE_BYTES = SCAN_BOXES * SCAN_BOX_BYTES;
constexpr int SCAN_STAGES = 6;
constexpr int SCAN_EPI_WARPS = 8;
constexpr int TMA_WARP = SCAN_EPI_WARPS, MMA_WARP = SCAN_EPI_WARPS + 1;  // MMA issuers: warps MMA_WARP, MMA_WARP + 1
constexpr int SCAN_MMA_WARPS = 2;
constexpr int SCAN_THREADS = (SCAN_EPI_WARPS + 1 + SCAN_MMA_WARPS) * 32;
constexpr int SCAN_MAX_D = 768;
constexpr uint32_t TMEM_COLS = 512;
constexpr uint32_t TMEM_ACC_COL = SCAN_MAX_D / 2;  // A: columns [0,384); accumulators: [384,448), [448,512)
constexpr int TOPK = LMI_PER_BUCKET_K;
constexpr int SCAN_SPLIT_KT = 8;       // SPLIT arrangement: k-blocks of the queries in TMEM (the rest in shared memory)
constexpr int SCAN_SPLIT_STAGES = 5;
constexpr int SCAN_SPLIT_A_BYTES = (SCAN_MAX_D / SCAN_KBLK - SCAN_SPLIT_KT) * SCAN_M * 128;  // 64 KB
constexpr int SCAN_SCHED_SLOTS = 4;   // item indices the producer may run ahead of the consumers
constexpr size_t SCAN_SMEM_BYTES = (size_t)SCAN_STAGES * SCAN_STAGE_BYTES + 1024 /*align*/ + 512 /*barriers*/ +
                                   2 * SCAN_M * sizeof(float) /*gates*/;
constexpr size_t SCAN_SPLIT_SMEM_BYTES = SCAN_SMEM_BYTES - (size_t)SCAN_STAGES * SCAN_STAGE_BYTES +
                                         (size_t)SCAN_SPLIT_STAGES * SCAN_STAGE_BYTES + SCAN_SPLIT_A_BYTES;
static_assert(SCAN_SPLIT_SMEM_BYTES <= 232448, "exceeds the 227 KB of dynamic shared memory per CTA");

// Scan tuning flags (lmi_scan_topk's `impl` >> 8): results are identical for every value.
constexpr int SCAN_FLAG_L2_SIBLINGS = 1;  // L2 eviction hints by sibling order (see the producer)
// (tried and dropped: requesting tile t + 2..4 into L2 with cp.async.bulk.prefetch.tensor while tile t is loaded --
// +14 % scan time at 10M, profiles/r02_scan_lab.md)

#ifdef LMI_PROFILE
// profiling build only (tools/): `dbg` ablations give WRONG results and are compiled out of the product
// (dbg & 1: no TMA loads; & 2: no candidate processing; & 4: half the operand traffic; & 8: MMA-warp
// wait cycles per CTA into g_scan_prof; & 16: stages pre-filled with pseudo-random fp16, so that MMAs
// without loads multiply real data -- zeros draw far less power; & 32: every load reads a 50 MB window,
// i.e. L2 -> SM traffic as usual but nothing from HBM)
__device__ long long g_scan_prof[148 * 8];
#define LMI_DBG(x) (x)
#else
#define LMI_DBG(x) 0
#endif

struct ScanBarriers {
    uint64_t full[SCAN_STAGES];
    uint64_t full_b[SCAN_STAGES];  // SPLIT: the second issuer's "stage filled" barriers
    uint64_t empty[SCAN_STAGES];
    uint64_t acc_full[2];
    uint64_t acc_empty[2];
    uint64_t q_ready;
    uint64_t sched_full[SCAN_SCHED_SLOTS];
    uint64_t sched_empty[SCAN_SCHED_SLOTS];
    int32_t sched_item[SCAN_SCHED_SLOTS];
    uint32_t tmem_base;
};
static_assert(sizeof(ScanBarriers) <= 512, "barrier block too large");

// sorted (descending) top-10 kept in registers; strict '>' keeps the earlier row on ties
struct TopK {
    float s[TOPK];
    int32_t i[TOPK];
    __device__ __forceinline__ void reset() {
#pragma unroll
        for (int t = 0; t < TOPK; ++t) {
            s[t] = -INFINITY;
            i[t] = -1;
        }
    }
    __device__ __forceinline__ void insert(float v, int32_t idx) {
        s[TOPK - 1] = v;
        i[TOPK - 1] = idx;
#pragma unroll
        for (int t = TOPK - 1; t > 0; --t) {
            const bool sw = s[t] > s[t - 1];
            const float fs = sw ? s[t - 1] : s[t];
            const float bs = sw ? s[t] : s[t - 1];
            const int32_t fi = sw ? i[t - 1] : i[t];
            const int32_t bi = sw ? i[t] : i[t - 1];
            s[t] = fs;
            s[t - 1] = bs;
            i[t] = fi;
            i[t - 1] = bi;
        }
    }
    __device__ __forceinline__ void offer(float v, int32_t idx) {
        if (v > s[TOPK - 1]) insert(v, idx);
    }
};

__device__ __forceinline__ void store_topk(const TopK& tk, float* __restrict__ part_sim,
                                           int32_t* __restrict__ part_row, size_t out) {
#pragma unroll
    for (int t = 0; t < TOPK; t += 2) {
        *reinterpret_cast<float2*>(part_sim + out + t) = make_float2(tk.s[t], tk.s[t + 1]);
        *reinterpret_cast<int2*>(part_row + out + t) = make_int2(tk.i[t], tk.i[t + 1]);
    }
}

// NT = bucket rows per score tile (= N of the MMA).  NT = 64: two accumulator buffers of 64 columns, two issuer
// warps that alternate by tile (the layout described above).  NT = 128: K = 768 of queries fills 384 of the 512
// tensor-memory columns, so ONE accumulator of 128 columns is left -- the tensor pipe pauses while the epilogue
// warps pull a finished tile out of TMEM (the candidate processing itself overlaps the next tile) -- but every
// MMA reads the A operand (4 KB of queries out of TMEM) for twice as many outputs, and one issuer warp suffices
// (an N = 128 instruction lasts 64 cycles).  Same work items, same partial lists (column half h of a tile ->
// list h), bit-identical results.
// order-preserving float <-> uint32 (0 is below every float), so that bounds can be compared as integers
__device__ __forceinline__ uint32_t gate_encode(float f) {
    const uint32_t b = __float_as_uint(f);
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ float gate_decode(uint32_t u) {
    return __uint_as_float((u & 0x80000000u) ? (u & 0x7fffffffu) : ~u);
}
// A value strictly below `bound` to gate with (`score > gate_below(bound)`), so that a score EQUAL to a bound that
// another list published still reaches the (similarity, row)-keyed merge, where the lower row wins the tie.  Two
// ulps relative for ordinary values; an absolute step of FLT_MIN where the relative one vanishes -- a bound of
// exactly 0.0 (all-zero queries or rows, padding) or a denormal-sized one -- and -inf stays -inf.
__device__ __forceinline__ float gate_below(float bound) {
    return bound - fmaxf(fabsf(bound) * 2.4e-7f, 1.17549435e-38f);
}
__device__ __forceinline__ uint32_t ld_relaxed_u32(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed_u32(uint32_t* p, uint32_t v) {
    asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

template <int NT, bool SPLIT>
__global__ void __launch_bounds__(SCAN_THREADS, 1)
scan_topk_tc_kernel(const __grid_constant__ CUtensorMap tmap, const __half* __restrict__ Q,
                    const int32_t* __restrict__ pair_list, int B, const int32_t* __restrict__ items,
                    const int32_t* __restrict__ n_items_ptr, int64_t max_items, int d, int lmax,
                    float* __restrict__ part_sim, int32_t* __restrict__ part_row, uint32_t* __restrict__ gate_g,
                    size_t gate_stride, int flags, int dbg) {
    constexpr int SCAN_N = NT;
    constexpr int SCAN_BOX_BYTES = NT * 128;                    // one TMA box: NT rows x 64 fp16
    constexpr int SCAN_BOXES = SCAN_STAGE_BYTES / SCAN_BOX_BYTES;  // k-blocks per 32 KB stage: 4 (NT = 64) / 2 (NT = 128)
    // SPLIT (NT = 128 only): the queries' first KT = 8 k-blocks (512 dimensions, 256 columns) live in tensor memory
    // and the remaining <= 4 k-blocks in shared memory (64 KB, the UMMA SWIZZLE_128B K-major layout, written by the
    // gather) -- their MMAs take A from a shared-memory descriptor -- which leaves 256 columns for TWO accumulators
    // of 128, so N = 128 tiles are double buffered like the N = 64 ones; the ring shrinks to 5 stages.
    static_assert(!SPLIT || NT == 128, "SPLIT is the NT = 128 arrangement");
    constexpr int NACC = (NT == 64 || SPLIT) ? 2 : 1;           // accumulator buffers that fit beside the queries
    constexpr int ISSUERS = (NT == 64 || SPLIT) ? 2 : 1;        // issuer warps in use (they alternate by tile)
    constexpr int HCOLS = NT / 2;                               // columns of a tile one epilogue warp owns
    constexpr int KT = SPLIT ? SCAN_SPLIT_KT : SCAN_MAX_D / SCAN_KBLK;  // k-blocks of the queries held in TMEM
    constexpr uint32_t ACC_COL = KT * (SCAN_KBLK / 2);          // first accumulator column
    constexpr int NSTAGES = SPLIT ? SCAN_SPLIT_STAGES : SCAN_STAGES;
    constexpr int A_S_BYTES = SPLIT ? SCAN_SPLIT_A_BYTES : 0;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* a_s = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    uint8_t* stages = a_s + A_S_BYTES;
    ScanBarriers* bars = reinterpret_cast<ScanBarriers*>(stages + (size_t)NSTAGES * SCAN_STAGE_BYTES);
    float* gates = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(bars) + 512);  // [2 halves][128 rows]

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int KB = d / SCAN_KBLK;
    int n_items = *n_items_ptr;
    if ((int64_t)n_items > max_items) n_items = (int)max_items;
    // Dynamic in-order scheduling: the producer warp draws item indices from a global counter and
    // hands them to the MMA and epilogue warps through a small shared-memory ring.  Items of one
    // (bucket, row range) -- one per tile of 128 probing queries -- are adjacent in the list, so
    // they are drawn back to back by CTAs that stream the same rows at the same time: the
    // range is read from HBM once and from L2 by the other query tiles.
    int* sched_counter = const_cast<int*>(n_items_ptr) + 1;
    auto next_item = [&](uint32_t& slot, uint32_t& phase) -> int {
        ptx::mbar_wait(&bars->sched_full[slot], phase);
        const int it = *reinterpret_cast<volatile int32_t*>(&bars->sched_item[slot]);
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(&bars->sched_empty[slot]);
        if (++slot == SCAN_SCHED_SLOTS) {
            slot = 0;
            phase ^= 1;
        }
        return it;
    };

    if (warp == TMA_WARP && lane == 0) {
        ptx::tma_prefetch_desc(&tmap);
        for (int s = 0; s < NSTAGES; ++s) {
            ptx::mbar_init(&bars->full[s], 1);
            ptx::mbar_init(&bars->full_b[s], 1);
            ptx::mbar_init(&bars->empty[s], 1);
        }
        for (int b = 0; b < 2; ++b) {
            ptx::mbar_init(&bars->acc_full[b], 1);
            ptx::mbar_init(&bars->acc_empty[b], SCAN_EPI_WARPS);
        }
        ptx::mbar_init(&bars->q_ready, SCAN_EPI_WARPS * 32);
        for (int b = 0; b < SCAN_SCHED_SLOTS; ++b) {
            ptx::mbar_init(&bars->sched_full[b], 1);
            ptx::mbar_init(&bars->sched_empty[b], SCAN_EPI_WARPS + ISSUERS);
        }
        ptx::fence_barrier_init();
    }
    if (warp == MMA_WARP) {
        ptx::tmem_alloc(&bars->tmem_base, TMEM_COLS);
        ptx::tmem_relinquish();
    }
#ifdef LMI_PROFILE
    if (dbg & 16) {  // ablation support: pseudo-random fp16 (|x| < 0.06) in every stage, so MMAs without loads see real data
        for (int i = threadIdx.x; i < NSTAGES * SCAN_STAGE_BYTES / 4; i += SCAN_THREADS) {
            uint32_t x = (uint32_t)i * 2654435761u + blockIdx.x * 40503u;
            x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
            reinterpret_cast<uint32_t*>(stages)[i] = (x & 0x83ff83ffu) | 0x28002800u;
        }
        ptx::fence_proxy_async();
    }
#endif
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem = bars->tmem_base;

    if (warp == TMA_WARP) {
        // ===================== TMA producer =====================
        // The whole warp walks the loop (warp-uniform control flow); one elected lane issues, so
        // ptxas emits the TMA / mbarrier instructions straight, without a per-thread waterfall loop.
        // The query tiles of one (bucket, row range) are adjacent items, drawn back to back by different CTAs:
        // the first reader pulls the rows from HBM, its siblings should find them in L2.  With
        // SCAN_FLAG_L2_SIBLINGS every tile but the last of a range loads with evict_last (keep the rows for
        // the siblings still to come) and the last one with evict_first (nobody reads them again this launch).
        const uint64_t pol_normal = ptx::policy_evict_normal();
        const uint64_t pol_keep = ptx::policy_evict_last();
        const uint64_t pol_drop = ptx::policy_evict_first();
        const int SPT = (KB + SCAN_BOXES - 1) / SCAN_BOXES;  // stages per 64-row tile
        uint32_t stage = 0, phase = 0, sslot = 0, sphase = 0, ptile = 0;
        while (true) {
            ptx::mbar_wait(&bars->sched_empty[sslot], sphase ^ 1);
            int it = 0;
            if (lane == 0) it = atomicAdd(sched_counter, 1);
            it = __shfl_sync(0xffffffffu, it, 0);
            if (it >= n_items) it = -1;
            if (lane == 0) {
                *reinterpret_cast<volatile int32_t*>(&bars->sched_item[sslot]) = it;
                ptx::mbar_arrive(&bars->sched_full[sslot]);
            }
            if (++sslot == SCAN_SCHED_SLOTS) {
                sslot = 0;
                sphase ^= 1;
            }
            if (it < 0) break;
            const int4 w0 = *reinterpret_cast<const int4*>(items + (size_t)it * LMI_ITEM_WORDS);
            const int4 w1 = *reinterpret_cast<const int4*>(items + (size_t)it * LMI_ITEM_WORDS + 4);
            const int row0 = w0.x, nrows = w0.y;
            const int ntiles = (nrows + SCAN_N - 1) / SCAN_N;
            uint64_t policy = pol_normal;
            if ((flags & SCAN_FLAG_L2_SIBLINGS) && w1.w > 1) policy = (w1.z + 1 == w1.w) ? pol_drop : pol_keep;
            for (int t = 0; t < ntiles; ++t, ++ptile) {
                // SPLIT: a tile spans more stages than the ring holds, so a stage serves both issuers in turn; each
                // issuer has its own "filled" barrier per stage (it then sees every phase of the barriers it waits on)
                uint64_t* full_bars = (SPLIT && (ptile & 1)) ? bars->full_b : bars->full;
                for (int s = 0; s < SPT; ++s) {
                    const int kb0 = s * SCAN_BOXES;
                    int nb = min(SCAN_BOXES, KB - kb0);
                    ptx::mbar_wait(&bars->empty[stage], phase ^ 1);
                    if (ptx::elect_one()) {
                        uint8_t* dst = stages + (size_t)stage * SCAN_STAGE_BYTES;
                        if (LMI_DBG(dbg & 1)) {  // ablation: no loads, the MMAs run on whatever the stage holds
                            ptx::mbar_arrive(&full_bars[stage]);
                            nb = 0;
                        } else {
                            if (LMI_DBG(dbg & 4)) nb = (nb + 1) / 2;  // ablation: half the operand traffic
                            ptx::mbar_arrive_expect_tx(&full_bars[stage], nb * SCAN_BOX_BYTES);
                        }
                        int trow = row0 + t * SCAN_N;
                        if (LMI_DBG(dbg & 32)) trow &= 32767;  // ablation: every load hits a 50 MB window (L2, no HBM)
                        for (int b = 0; b < nb; ++b)
                            ptx::tma_load_2d(dst + b * SCAN_BOX_BYTES, &tmap, (kb0 + b) * SCAN_KBLK, trow,
                                             &full_bars[stage], policy);
                    }
                    __syncwarp();
                    if (++stage == NSTAGES) {
                        stage = 0;
                        phase ^= 1;
                    }
                }
            }
        }
    } else if (warp >= MMA_WARP) {
        // ===================== MMA issuers =====================
        // tcgen05.mma issue is synchronous with the tensor pipe (measured: the issuing thread
        // advances one MMA per ~36 cycles and the pipe drains within one or two MMAs once it
        // stops), so every barrier round trip of a single issuer is tensor-pipe idle time.  Two
        // issuer warps therefore alternate by score tile: warp w owns the tiles with
        // (tile counter & 1) == w, i.e. accumulator buffer w and every other group of SPT
        // pipeline stages, and does its barrier waits while the other warp is issuing.
        const uint32_t mw = warp - MMA_WARP;
        if (mw >= (uint32_t)ISSUERS) goto done;  // NT = 128: one issuer
        constexpr uint32_t idesc = ptx::umma_idesc_f16(SCAN_M, SCAN_N);
        const int SPT = (KB + SCAN_BOXES - 1) / SCAN_BOXES;
        const uint64_t desc0 = ptx::umma_desc_sw128_kmajor(ptx::smem_u32(stages));
        const uint32_t desc0_lo = (uint32_t)desc0;
        const uint64_t desc0_hi = desc0 >> 32;  // the same for every SWIZZLE_128B K-major operand
        const uint32_t descA_lo = (uint32_t)ptx::umma_desc_sw128_kmajor(ptx::smem_u32(a_s));
        (void)descA_lo;
        uint32_t stage = 0, phase = 0, tcount = 0, item_parity = 0, sslot = 0, sphase = 0;
        uint32_t use_bits = 0;  // SPLIT: bit s = parity of this issuer's uses of stage s so far
        uint64_t* full_bars = (SPLIT && mw == 1) ? bars->full_b : bars->full;
        auto skip_tile_stages = [&]() {
            for (int s = 0; s < SPT; ++s)
                if (++stage == NSTAGES) {
                    stage = 0;
                    phase ^= 1;
                }
        };
        if (ISSUERS == 2 && mw == 1) skip_tile_stages();
#ifdef LMI_PROFILE
        const bool prof = (dbg & 8) != 0 && mw == 0;
        long long p_sched = 0, p_q = 0, p_acc = 0, p_full = 0, p_issue = 0, p_n = 0, p_t0 = clock64(), p_c;
#define PROF_BEGIN() if (prof) p_c = clock64()
#define PROF_END(acc) if (prof) acc += clock64() - p_c
#define PROF_COUNT() ++p_n
#else
#define PROF_BEGIN()
#define PROF_END(acc)
#define PROF_COUNT()
#endif
        while (true) {
            PROF_BEGIN();
            const int it = next_item(sslot, sphase);
            PROF_END(p_sched);
            if (it < 0) break;
            const int4 w0 = *reinterpret_cast<const int4*>(items + (size_t)it * LMI_ITEM_WORDS);
            const int nrows = w0.y;
            const int ntiles = (nrows + SCAN_N - 1) / SCAN_N;
            PROF_BEGIN();
            ptx::mbar_wait(&bars->q_ready, item_parity);
            PROF_END(p_q);
            item_parity ^= 1;
            ptx::tc_fence_after();
            for (int t = 0; t < ntiles; ++t, ++tcount) {
                const uint32_t buf = tcount & (NACC - 1);
                if (ISSUERS == 2 && buf != mw) continue;
                PROF_BEGIN();
                ptx::mbar_wait(&bars->acc_empty[buf], ((tcount / NACC) & 1) ^ 1);
                PROF_END(p_acc);
                ptx::tc_fence_after();
                const uint32_t d_tmem = tmem + ACC_COL + buf * SCAN_N;
                for (int s = 0; s < SPT; ++s) {
                    const int kb0 = s * SCAN_BOXES;
                    const int nb = min(SCAN_BOXES, KB - kb0);
                    PROF_BEGIN();
                    ptx::mbar_wait(&full_bars[stage], SPLIT ? ((use_bits >> stage) & 1u) : phase);
                    PROF_END(p_full);
                    use_bits ^= 1u << stage;
                    ptx::tc_fence_after();
                    PROF_BEGIN();
                    PROF_COUNT();
                    if (ptx::elect_one()) {
                        // The descriptor's start-address field counts 16-byte units in its low word,
                        // so stage / box / k offsets are plain adds on that word.  A full stage is
                        // issued as 16 straight-line MMAs whose operands differ from the stage base
                        // by compile-time constants (no dependent address arithmetic between them).
                        const uint32_t dlo = desc0_lo + ((stage * SCAN_STAGE_BYTES) >> 4);
                        const uint32_t a0 = tmem + kb0 * (SCAN_KBLK / 2);
                        if (SPLIT && kb0 >= KT) {
                            // the k-blocks of the queries that live in shared memory (KT is a multiple of the
                            // boxes per stage, so a stage is all-TMEM or all-shared): A by descriptor
                            const uint32_t alo = descA_lo + (((kb0 - KT) * (SCAN_M * 128)) >> 4);
                            for (int b = 0; b < nb; ++b) {
#pragma unroll
                                for (int k = 0; k < SCAN_KBLK / 16; ++k) {
                                    const uint64_t b_desc =
                                        (desc0_hi << 32) | (uint64_t)(dlo + ((b * SCAN_BOX_BYTES + k * 32) >> 4));
                                    const uint64_t a_desc =
                                        (desc0_hi << 32) | (uint64_t)(alo + ((b * (SCAN_M * 128) + k * 32) >> 4));
                                    ptx::mma_f16_ss(d_tmem, a_desc, b_desc, idesc, 1u);  // never a tile's first MMA (KT > 0)
                                }
                            }
                        } else if (nb == SCAN_BOXES) {
#pragma unroll
                            for (int b = 0; b < SCAN_BOXES; ++b) {
#pragma unroll
                                for (int k = 0; k < SCAN_KBLK / 16; ++k) {
                                    const uint64_t b_desc =
                                        (desc0_hi << 32) | (uint64_t)(dlo + ((b * SCAN_BOX_BYTES + k * 32) >> 4));
                                    ptx::mma_f16_ts(d_tmem, a0 + b * (SCAN_KBLK / 2) + k * 8, b_desc, idesc,
                                                    (b | k) != 0 ? 1u : (s != 0 ? 1u : 0u));
                                }
                            }
                        } else {
                            for (int b = 0; b < nb; ++b) {
#pragma unroll
                                for (int k = 0; k < SCAN_KBLK / 16; ++k) {
                                    const uint64_t b_desc =
                                        (desc0_hi << 32) | (uint64_t)(dlo + ((b * SCAN_BOX_BYTES + k * 32) >> 4));
                                    ptx::mma_f16_ts(d_tmem, a0 + b * (SCAN_KBLK / 2) + k * 8, b_desc, idesc,
                                                    (s | b | k) != 0 ? 1u : 0u);
                                }
                            }
                        }
                        ptx::mma_commit(&bars->empty[stage]);
                        if (s == SPT - 1) ptx::mma_commit(&bars->acc_full[buf]);
                    }
                    __syncwarp();
                    PROF_END(p_issue);
                    if (++stage == NSTAGES) {
                        stage = 0;
                        phase ^= 1;
                    }
                }
                if (ISSUERS == 2) skip_tile_stages();  // the other issuer's tile
            }
        }
#ifdef LMI_PROFILE
        if (prof && lane == 0 && blockIdx.x < 148) {
            long long* o = g_scan_prof + blockIdx.x * 8;
            o[0] = clock64() - p_t0; o[1] = p_sched; o[2] = p_q; o[3] = p_acc; o[4] = p_full; o[5] = p_issue; o[6] = p_n;
        }
#endif
#undef PROF_BEGIN
#undef PROF_END
#undef PROF_COUNT
    } else {
        // ===================== epilogue warps (thread == query row) =====================
        // 8 warps: warp w reads TMEM lanes [32*(w&3), +32) (the hardware's lane quarter of warp
        // w%4) and the 32-column half h = w>>2 of every 64-column score tile, so two warps per
        // scheduler hide each other's TMEM-load and branch latencies.  Each (row, half) keeps its
        // own top-10 list; the two halves gate on the larger of their 10th-best (exchanged through
        // shared memory) and the merge kernel folds the lists with a (similarity, row) key.
        const int q4 = warp & 3, h = warp >> 2;
        const int row = q4 * 32 + lane;
        const uint32_t lane_field = static_cast<uint32_t>(q4 * 32) << 16;
        float* gate_mine = gates + h * SCAN_M + row;
        const float* gate_other = gates + (h ^ 1) * SCAN_M + row;
        uint32_t tcount = 0, sslot = 0, sphase = 0;
        while (true) {
            const int it = next_item(sslot, sphase);
            if (it < 0) break;
            const int4 w0 = *reinterpret_cast<const int4*>(items + (size_t)it * LMI_ITEM_WORDS);
            const int4 w1 = *reinterpret_cast<const int4*>(items + (size_t)it * LMI_ITEM_WORDS + 4);
            const int row0 = w0.x, nrows = w0.y, slot0 = w0.z, mvalid = w0.w, list0 = w1.x & 0xffff;
            const int ntiles = (nrows + SCAN_N - 1) / SCAN_N;
            const bool valid = row < mvalid;

            // gather this item's query rows into TMEM; half h takes the k-blocks c with c % 2 == h
            // (all MMAs of the previous item have completed: every epilogue warp consumed its
            // last accumulator tile before getting here)
            const int q = valid ? (pair_list[slot0 + row] / B) : 0;
            const uint4* qsrc = reinterpret_cast<const uint4*>(Q + (size_t)q * d);
            // two register buffers: the loads of k-block c + 2 are in flight while c is stored to TMEM
            auto load_kblock = [&](uint32_t (&r)[32], int c) {
                if (valid) {
#pragma unroll
                    for (int v = 0; v < 8; ++v) {
                        const uint4 u = __ldg(qsrc + c * 8 + v);
                        r[4 * v + 0] = u.x;
                        r[4 * v + 1] = u.y;
                        r[4 * v + 2] = u.z;
                        r[4 * v + 3] = u.w;
                    }
                } else {
#pragma unroll
                    for (int v = 0; v < 32; ++v) r[v] = 0u;
                }
            };
            // k-block c of the row: into TMEM columns [32c, 32c + 32) of the row's lane, or (SPLIT, c >= KT) into the
            // shared-memory A tile of k-block c - KT: row r of a 128-row SWIZZLE_128B K-major tile lies at
            // (r / 8) * 1024 + (r % 8) * 128 and its 16-byte chunk j at position j ^ (r % 8)
            auto store_kblock = [&](const uint32_t (&r)[32], int c) {
                if (!SPLIT || c < KT) {
                    ptx::tmem_st32(tmem + lane_field + c * (SCAN_KBLK / 2), r);
                } else {
                    uint8_t* dst = a_s + (size_t)(c - KT) * (SCAN_M * 128) + (row >> 3) * 1024 + (row & 7) * 128;
#pragma unroll
                    for (int j = 0; j < 8; ++j)
                        *reinterpret_cast<uint4*>(dst + ((j ^ (row & 7)) << 4)) =
                            make_uint4(r[4 * j + 0], r[4 * j + 1], r[4 * j + 2], r[4 * j + 3]);
                }
            };
            {
                uint32_t ra[32], rb[32];
                int c = h;
                if (c < KB) load_kblock(ra, c);
                while (c < KB) {
                    if (c + 2 < KB) load_kblock(rb, c + 2);
                    store_kblock(ra, c);
                    c += 2;
                    if (c >= KB) break;
                    if (c + 2 < KB) load_kblock(ra, c + 2);
                    store_kblock(rb, c);
                    c += 2;
                }
            }
            ptx::tmem_st_wait();
            if (SPLIT) ptx::fence_proxy_async();  // the shared-memory part is read by the tensor core (async proxy)
            *gate_mine = -INFINITY;
            ptx::tc_fence_before();
            ptx::mbar_arrive(&bars->q_ready);

            TopK tk;
            tk.reset();
            float tau = -INFINITY;  // gate: a lower bound of this (row, range)'s final 10th best
            // The (query, bucket) pair's gate shared with the CTAs scanning the bucket's other row ranges: list l of
            // the pair publishes its GJ-th best so far in gate_g[l * pairs + slot]; GM lists with GJ entries above
            // a value v (GM * GJ >= 10) mean ten rows of the bucket beat v, so the GM-th largest published value
            // is a lower bound of the bucket's final 10th best.
            const int n_lists = (w1.x >> 16) & 0xffff;
            const bool wide = n_lists >= 5;  // (GJ, GM) = (2, 5), else (5, 2)
            uint32_t* gate_pair = (gate_g != nullptr && valid && n_lists >= 2) ? gate_g + slot0 + row : nullptr;
            uint32_t gate_pub = 0;  // what this list last published
            for (int t = 0; t < ntiles; ++t, ++tcount) {
                const uint32_t buf = tcount & (NACC - 1);
                // refresh the bound at tiles 1, 3, 7, 15, 23, ...: the loads are in flight while the tile is awaited
                uint32_t gq = 0;
                if (gate_pair != nullptr && (t == 1 || t == 3 || (t & 7) == 7)) {
                    uint32_t top[5] = {0u, 0u, 0u, 0u, 0u};  // descending
                    for (int l0 = 0; l0 < n_lists; l0 += 8) {
                        uint32_t u[8];  // eight independent loads in flight, then the insertion network
#pragma unroll
                        for (int i = 0; i < 8; ++i)
                            u[i] = l0 + i < n_lists ? ld_relaxed_u32(gate_pair + (size_t)(l0 + i) * gate_stride) : 0u;
#pragma unroll
                        for (int i = 0; i < 8; ++i) {
#pragma unroll
                            for (int e = 0; e < 5; ++e) {
                                const uint32_t hi = max(top[e], u[i]);
                                u[i] = min(top[e], u[i]);
                                top[e] = hi;
                            }
                        }
                    }
                    gq = wide ? top[4] : top[1];
                }
                ptx::mbar_wait(&bars->acc_full[buf], (tcount / NACC) & 1);
                ptx::tc_fence_after();
                uint32_t v[HCOLS];
                {
                    uint32_t(&v0)[32] = *reinterpret_cast<uint32_t(*)[32]>(&v[0]);
                    ptx::tmem_ld32(tmem + lane_field + ACC_COL + buf * SCAN_N + h * HCOLS, v0);
                    if (HCOLS == 64) {
                        uint32_t(&v1)[32] = *reinterpret_cast<uint32_t(*)[32]>(&v[HCOLS - 32]);
                        ptx::tmem_ld32(tmem + lane_field + ACC_COL + buf * SCAN_N + h * HCOLS + 32, v1);
                    }
                }
                // the other half's 10th best, nudged down so that an exact tie with it still
                // reaches the (similarity, row)-keyed merge
                const float og = *gate_other;
                tau = fmaxf(tau, gate_below(og));
                if (gq != 0u) {  // same nudge: an exact tie with the bound still reaches the merge
                    const float gg = gate_decode(gq);
                    tau = fmaxf(tau, gate_below(gg));
                }
                ptx::tmem_ld_wait();
                ptx::tc_fence_before();
                __syncwarp();
                if (lane == 0) ptx::mbar_arrive(&bars->acc_empty[buf]);  // scores are in registers now

                const int base_row = row0 + t * SCAN_N + h * HCOLS;
                const int lim = nrows - t * SCAN_N - h * HCOLS;  // columns of this half inside the range
                if (lim < HCOLS) {
#pragma unroll
                    for (int j = 0; j < HCOLS; ++j)
                        if (j >= lim) v[j] = 0xff800000u;  // -inf
                }
                constexpr int NG = HCOLS / 8;  // groups of 8 columns
                float g[NG];
#pragma unroll
                for (int gi = 0; gi < NG; ++gi) {
                    float m = __uint_as_float(v[gi * 8]);
#pragma unroll
                    for (int j = 1; j < 8; ++j) m = fmaxf(m, __uint_as_float(v[gi * 8 + j]));
                    g[gi] = m;
                }
                float mall = g[0];
#pragma unroll
                for (int gi = 1; gi < NG; ++gi) mall = fmaxf(mall, g[gi]);
                if (!LMI_DBG(dbg & 2) && __any_sync(0xffffffffu, mall > tau)) {
#pragma unroll
                    for (int gi = 0; gi < NG; ++gi) {
                        if (__any_sync(0xffffffffu, g[gi] > tau)) {
                            uint32_t mask = 0;
#pragma unroll
                            for (int j = 0; j < 8; ++j)
                                mask |= (__uint_as_float(v[gi * 8 + j]) > tau) ? (1u << j) : 0u;
                            // every lane pops its own lowest candidate column per trip, so lanes with
                            // candidates in different columns share one pass through the insertion code
                            while (__any_sync(0xffffffffu, mask != 0)) {
                                if (mask != 0) {
                                    const int j = __ffs(mask) - 1;
                                    mask &= mask - 1;
                                    const uint32_t a0 = (j & 1) ? v[gi * 8 + 1] : v[gi * 8 + 0];
                                    const uint32_t a1 = (j & 1) ? v[gi * 8 + 3] : v[gi * 8 + 2];
                                    const uint32_t a2 = (j & 1) ? v[gi * 8 + 5] : v[gi * 8 + 4];
                                    const uint32_t a3 = (j & 1) ? v[gi * 8 + 7] : v[gi * 8 + 6];
                                    const uint32_t b0 = (j & 2) ? a1 : a0;
                                    const uint32_t b1 = (j & 2) ? a3 : a2;
                                    const float f = __uint_as_float((j & 4) ? b1 : b0);
                                    if (f > tau) {
                                        tk.insert(f, base_row + gi * 8 + j);
                                        tau = fmaxf(tau, tk.s[TOPK - 1]);
                                    }
                                }
                            }
                        }
                    }
                    *gate_mine = tk.s[TOPK - 1];
                    if (gate_pair != nullptr) {
                        const float jth = wide ? tk.s[1] : tk.s[4];
                        const uint32_t mine = jth > -INFINITY ? gate_encode(jth) : 0u;
                        if (mine > gate_pub) {
                            gate_pub = mine;
                            st_relaxed_u32(gate_pair + (size_t)(list0 + h) * gate_stride, mine);
                        }
                    }
                }
            }
            if (valid)
                store_topk(tk, part_sim, part_row, ((size_t)(slot0 + row) * lmax + list0 + h) * TOPK);
        }
    }

done:
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == MMA_WARP) {
        ptx::tc_fence_after();
        ptx::tmem_dealloc(tmem, TMEM_COLS);
    }
}

// ------------------------------------------------------------------------------------------------
// The same scan with the bucket rows shared between TWO CTAs (a thread-block cluster of 2 = the two SMs
// of a TPC) by TMA MULTICAST.  The cluster draws PAIRS of consecutive work items; CTA `rank` of the pair
// takes item 2p + rank.  Items of one (bucket, row range) -- one per tile of 128 probing queries -- are
// adjacent in the list, so most pairs are two query tiles over the SAME rows ("shared" pair): each CTA then
// issues the TMA loads of half of every pipeline stage's boxes with .multicast::cluster, the box lands in
// BOTH CTAs' shared memory and completes bytes on both CTAs' full barriers -- the rows cross L2 -> SM once
// for the two tiles instead of twice.  A pair of unrelated items (a bucket's odd last tile next to the
// first tile of the following range) runs the two CTAs independently, exactly like the single-CTA kernel.
// Everything per item -- query gather into TMEM, tcgen05.mma with A in TMEM, the fused top-10 epilogue,
// the partial lists written -- is identical to scan_topk_tc_kernel, so results are bit-identical.
//
// Protocol.  One cluster barrier per pair (all warps of both CTAs): when it completes every MMA of the
// previous pair has finished in both CTAs (the epilogue warps only arrive after draining the last
// accumulator tile), so all pipeline stages of both CTAs are free, the pipeline restarts at stage 0 in
// lock step, and the next pair index -- drawn by the leader CTA's producer lane from the global counter and
// written into both CTAs' shared memory before it arrives -- is visible.  `empty` barriers count 2
// arrivals: in a shared pair a stage of CTA x is written by both producers, so it is free when BOTH issuers
// are done with their copy (each tcgen05.commit is multicast to the two CTAs' barriers); in an independent
// pair the CTA's own issuer commits twice.  Stage parities are tracked per stage (a bit mask), because a
// pair may leave the ring anywhere and the next one starts at stage 0 again.
struct PairBarriers {
    uint64_t full[SCAN_STAGES];
    uint64_t empty[SCAN_STAGES];
    uint64_t acc_full[2];
    uint64_t acc_empty[2];
    uint64_t q_ready;
    int32_t sched_pair[2];  // pair index of cluster round r is sched_pair[r & 1]
    uint32_t tmem_base;
};
static_assert(sizeof(PairBarriers) <= 512, "barrier block too large");

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(SCAN_THREADS, 1)
scan_topk_pair_kernel(const __grid_constant__ CUtensorMap tmap, const __half* __restrict__ Q,
                      const int32_t* __restrict__ pair_list, int B, const int32_t* __restrict__ items,
                      const int32_t* __restrict__ n_items_ptr, int64_t max_items, int d, int lmax,
                      float* __restrict__ part_sim, int32_t* __restrict__ part_row) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* stages = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    PairBarriers* bars = reinterpret_cast<PairBarriers*>(stages + (size_t)SCAN_STAGES * SCAN_STAGE_BYTES);
    float* gates = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(bars) + 512);  // [2 halves][128 rows]

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int KB = d / SCAN_KBLK;
    const int SPT = (KB + SCAN_BOXES - 1) / SCAN_BOXES;  // stages per 64-row tile
    int n_items = *n_items_ptr;
    if ((int64_t)n_items > max_items) n_items = (int)max_items;
    const int n_pairs = (n_items + 1) >> 1;
    const uint32_t rank = ptx::cluster_ctarank();
    int* sched_counter = const_cast<int*>(n_items_ptr) + 1;
    const bool scheduler = rank == 0 && warp == TMA_WARP && lane == 0;  // draws the pair indices for the cluster
    auto publish_pair = [&](int slot, int p) {
        *reinterpret_cast<volatile int32_t*>(&bars->sched_pair[slot]) = p;
        ptx::st_cluster_u32(ptx::mapa(ptx::smem_u32(&bars->sched_pair[slot]), 1), (uint32_t)p);
    };

    if (warp == TMA_WARP && lane == 0) {
        ptx::tma_prefetch_desc(&tmap);
        for (int s = 0; s < SCAN_STAGES; ++s) {
            ptx::mbar_init(&bars->full[s], 1);
            ptx::mbar_init(&bars->empty[s], 2);
        }
        for (int b = 0; b < 2; ++b) {
            ptx::mbar_init(&bars->acc_full[b], 1);
            ptx::mbar_init(&bars->acc_empty[b], SCAN_EPI_WARPS);
        }
        ptx::mbar_init(&bars->q_ready, SCAN_EPI_WARPS * 32);
        ptx::fence_barrier_init();
    }
    if (warp == MMA_WARP) {
        ptx::tmem_alloc(&bars->tmem_base, TMEM_COLS);
        ptx::tmem_relinquish();
    }
    ptx::tc_fence_before();
    ptx::cluster_sync();  // barrier inits of BOTH CTAs are in place before anything is multicast
    ptx::tc_fence_after();
    const uint32_t tmem = bars->tmem_base;
    int next_p = 0;
    if (scheduler) {
        publish_pair(0, atomicAdd(sched_counter, 1));
        next_p = atomicAdd(sched_counter, 1);
    }
    ptx::cluster_sync();

    // per-role running state (identical bookkeeping in the producer and both issuers)
    uint32_t use_bits = 0;      // bit s: parity of how often stage s has been used so far in this CTA
    uint32_t tcount = 0;        // score tiles so far (accumulator buffer = tcount & 1)
    uint32_t item_parity = 0;   // items this CTA has processed so far (q_ready phase)
    const uint32_t mw = warp - MMA_WARP;

    for (uint32_t round = 0;; ++round) {
        const int p = *reinterpret_cast<volatile int32_t*>(&bars->sched_pair[round & 1]);
        if (p >= n_pairs) break;
        const int it = 2 * p + (int)rank, it_peer = 2 * p + (int)(rank ^ 1);
        const bool has = it < n_items;
        int4 w0 = make_int4(0, 0, 0, 0), w1 = make_int4(0, 0, 0, 0);
        if (has) {
            w0 = *reinterpret_cast<const int4*>(items + (size_t)it * LMI_ITEM_WORDS);
            w1 = *reinterpret_cast<const int4*>(items + (size_t)it * LMI_ITEM_WORDS + 4);
        }
        bool shared = false;
        if (has && it_peer < n_items) {
            const int2 pw = *reinterpret_cast<const int2*>(items + (size_t)it_peer * LMI_ITEM_WORDS);
            shared = pw.x == w0.x && pw.y == w0.y;  // same rows: two query tiles of one (bucket, range)
        }
        const int row0 = w0.x, nrows = w0.y, slot0 = w0.z, mvalid = w0.w, list0 = w1.x & 0xffff;
        const int ntiles = has ? (nrows + SCAN_N - 1) / SCAN_N : 0;

        if (warp == TMA_WARP) {
            // ===================== TMA producer =====================
            const uint64_t policy = ptx::policy_evict_normal();
            uint32_t stage = 0;
            for (int t = 0; t < ntiles; ++t) {
                for (int s = 0; s < SPT; ++s) {
                    const int kb0 = s * SCAN_BOXES;
                    const int nb = min(SCAN_BOXES, KB - kb0);
                    ptx::mbar_wait(&bars->empty[stage], ((use_bits >> stage) & 1u) ^ 1u);
                    if (ptx::elect_one()) {
                        uint8_t* dst = stages + (size_t)stage * SCAN_STAGE_BYTES;
                        ptx::mbar_arrive_expect_tx(&bars->full[stage], nb * SCAN_BOX_BYTES);
                        if (shared) {
                            // this CTA's half of the stage (boxes b with b % 2 == rank), delivered to both CTAs
                            for (int b = (int)rank; b < nb; b += 2)
                                ptx::tma_load_2d_mcast(dst + b * SCAN_BOX_BYTES, &tmap, (kb0 + b) * SCAN_KBLK,
                                                       row0 + t * SCAN_N, &bars->full[stage], (uint16_t)3, policy);
                        } else {
                            for (int b = 0; b < nb; ++b)
                                ptx::tma_load_2d(dst + b * SCAN_BOX_BYTES, &tmap, (kb0 + b) * SCAN_KBLK, row0 + t * SCAN_N,
                                                 &bars->full[stage], policy);
                        }
                    }
                    __syncwarp();
                    use_bits ^= 1u << stage;
                    if (++stage == SCAN_STAGES) stage = 0;
                }
            }
        } else if (warp >= MMA_WARP) {
            // ===================== MMA issuers (alternate by score tile, as in scan_topk_tc_kernel) ========
            constexpr uint32_t idesc = ptx::umma_idesc_f16(SCAN_M, SCAN_N);
            const uint64_t desc0 = ptx::umma_desc_sw128_kmajor(ptx::smem_u32(stages));
            const uint32_t desc0_lo = (uint32_t)desc0;
            const uint64_t desc0_hi = desc0 >> 32;
            if (has) {
                ptx::mbar_wait(&bars->q_ready, item_parity);
                item_parity ^= 1;
                ptx::tc_fence_after();
            }
            uint32_t stage = 0;
            for (int t = 0; t < ntiles; ++t, ++tcount) {
                const uint32_t buf = tcount & 1;
                if (buf != mw) {  // the other issuer's tile: account for its stages
                    for (int s = 0; s < SPT; ++s) {
                        use_bits ^= 1u << stage;
                        if (++stage == SCAN_STAGES) stage = 0;
                    }
                    continue;
                }
                ptx::mbar_wait(&bars->acc_empty[buf], ((tcount >> 1) & 1) ^ 1);
                ptx::tc_fence_after();
                const uint32_t d_tmem = tmem + TMEM_ACC_COL + buf * SCAN_N;
                for (int s = 0; s < SPT; ++s) {
                    const int kb0 = s * SCAN_BOXES;
                    const int nb = min(SCAN_BOXES, KB - kb0);
                    ptx::mbar_wait(&bars->full[stage], (use_bits >> stage) & 1u);
                    ptx::tc_fence_after();
                    if (ptx::elect_one()) {
                        const uint32_t dlo = desc0_lo + ((stage * SCAN_STAGE_BYTES) >> 4);
                        const uint32_t a0 = tmem + kb0 * (SCAN_KBLK / 2);
                        if (nb == SCAN_BOXES) {
#pragma unroll
                            for (int b = 0; b < SCAN_BOXES; ++b) {
#pragma unroll
                                for (int k = 0; k < SCAN_KBLK / 16; ++k) {
                                    const uint64_t b_desc =
                                        (desc0_hi << 32) | (uint64_t)(dlo + ((b * SCAN_BOX_BYTES + k * 32) >> 4));
                                    ptx::mma_f16_ts(d_tmem, a0 + b * (SCAN_KBLK / 2) + k * 8, b_desc, idesc,
                                                    (b | k) != 0 ? 1u : (s != 0 ? 1u : 0u));
                                }
                            }
                        } else {
                            for (int b = 0; b < nb; ++b) {
#pragma unroll
                                for (int k = 0; k < SCAN_KBLK / 16; ++k) {
                                    const uint64_t b_desc =
                                        (desc0_hi << 32) | (uint64_t)(dlo + ((b * SCAN_BOX_BYTES + k * 32) >> 4));
                                    ptx::mma_f16_ts(d_tmem, a0 + b * (SCAN_KBLK / 2) + k * 8, b_desc, idesc,
                                                    (s | b | k) != 0 ? 1u : 0u);
                                }
                            }
                        }
                        if (shared) {
                            ptx::mma_commit_mcast(&bars->empty[stage], (uint16_t)3);  // frees the stage in BOTH CTAs
                        } else {
                            ptx::mma_commit(&bars->empty[stage]);
                            ptx::mma_commit(&bars->empty[stage]);
                        }
                        if (s == SPT - 1) ptx::mma_commit(&bars->acc_full[buf]);
                    }
                    __syncwarp();
                    use_bits ^= 1u << stage;
                    if (++stage == SCAN_STAGES) stage = 0;
                }
            }
        } else if (has) {
            // ===================== epilogue warps (thread == query row), as in scan_topk_tc_kernel ==========
            const int q4 = warp & 3, h = warp >> 2;
            const int row = q4 * 32 + lane;
            const uint32_t lane_field = static_cast<uint32_t>(q4 * 32) << 16;
            float* gate_mine = gates + h * SCAN_M + row;
            const float* gate_other = gates + (h ^ 1) * SCAN_M + row;
            const bool valid = row < mvalid;
            const int q = valid ? (pair_list[slot0 + row] / B) : 0;
            const uint4* qsrc = reinterpret_cast<const uint4*>(Q + (size_t)q * d);
            auto load_kblock = [&](uint32_t (&r)[32], int c) {
                if (valid) {
#pragma unroll
                    for (int v = 0; v < 8; ++v) {
                        const uint4 u = __ldg(qsrc + c * 8 + v);
                        r[4 * v + 0] = u.x;
                        r[4 * v + 1] = u.y;
                        r[4 * v + 2] = u.z;
                        r[4 * v + 3] = u.w;
                    }
                } else {
#pragma unroll
                    for (int v = 0; v < 32; ++v) r[v] = 0u;
                }
            };
            {
                uint32_t ra[32], rb[32];
                int c = h;
                if (c < KB) load_kblock(ra, c);
                while (c < KB) {
                    if (c + 2 < KB) load_kblock(rb, c + 2);
                    ptx::tmem_st32(tmem + lane_field + c * (SCAN_KBLK / 2), ra);
                    c += 2;
                    if (c >= KB) break;
                    if (c + 2 < KB) load_kblock(ra, c + 2);
                    ptx::tmem_st32(tmem + lane_field + c * (SCAN_KBLK / 2), rb);
                    c += 2;
                }
            }
            ptx::tmem_st_wait();
            *gate_mine = -INFINITY;
            ptx::tc_fence_before();
            ptx::mbar_arrive(&bars->q_ready);

            TopK tk;
            tk.reset();
            float tau = -INFINITY;
            for (int t = 0; t < ntiles; ++t, ++tcount) {
                const uint32_t buf = tcount & 1;
                ptx::mbar_wait(&bars->acc_full[buf], (tcount >> 1) & 1);
                ptx::tc_fence_after();
                uint32_t v[32];
                ptx::tmem_ld32(tmem + lane_field + TMEM_ACC_COL + buf * SCAN_N + h * 32, v);
                const float og = *gate_other;
                tau = fmaxf(tau, gate_below(og));
                ptx::tmem_ld_wait();
                ptx::tc_fence_before();
                __syncwarp();
                if (lane == 0) ptx::mbar_arrive(&bars->acc_empty[buf]);

                const int base_row = row0 + t * SCAN_N + h * 32;
                const int lim = nrows - t * SCAN_N - h * 32;
                if (lim < 32) {
#pragma unroll
                    for (int j = 0; j < 32; ++j)
                        if (j >= lim) v[j] = 0xff800000u;  // -inf
                }
                float g[4];
#pragma unroll
                for (int gi = 0; gi < 4; ++gi) {
                    float m = __uint_as_float(v[gi * 8]);
#pragma unroll
                    for (int j = 1; j < 8; ++j) m = fmaxf(m, __uint_as_float(v[gi * 8 + j]));
                    g[gi] = m;
                }
                const float mall = fmaxf(fmaxf(g[0], g[1]), fmaxf(g[2], g[3]));
                if (__any_sync(0xffffffffu, mall > tau)) {
#pragma unroll
                    for (int gi = 0; gi < 4; ++gi) {
                        if (__any_sync(0xffffffffu, g[gi] > tau)) {
                            uint32_t mask = 0;
#pragma unroll
                            for (int j = 0; j < 8; ++j)
                                mask |= (__uint_as_float(v[gi * 8 + j]) > tau) ? (1u << j) : 0u;
                            while (__any_sync(0xffffffffu, mask != 0)) {
                                if (mask != 0) {
                                    const int j = __ffs(mask) - 1;
                                    mask &= mask - 1;
                                    const uint32_t a0 = (j & 1) ? v[gi * 8 + 1] : v[gi * 8 + 0];
                                    const uint32_t a1 = (j & 1) ? v[gi * 8 + 3] : v[gi * 8 + 2];
                                    const uint32_t a2 = (j & 1) ? v[gi * 8 + 5] : v[gi * 8 + 4];
                                    const uint32_t a3 = (j & 1) ? v[gi * 8 + 7] : v[gi * 8 + 6];
                                    const uint32_t b0 = (j & 2) ? a1 : a0;
                                    const uint32_t b1 = (j & 2) ? a3 : a2;
                                    const float f = __uint_as_float((j & 4) ? b1 : b0);
                                    if (f > tau) {
                                        tk.insert(f, base_row + gi * 8 + j);
                                        tau = fmaxf(tau, tk.s[TOPK - 1]);
                                    }
                                }
                            }
                        }
                    }
                    *gate_mine = tk.s[TOPK - 1];
                }
            }
            if (valid)
                store_topk(tk, part_sim, part_row, ((size_t)(slot0 + row) * lmax + list0 + h) * TOPK);
        }

        // ---- end of the pair: publish the next pair index, then the cluster barrier ----
        if (scheduler) {
            publish_pair((round + 1) & 1, next_p);
            next_p = atomicAdd(sched_counter, 1);
        }
        __syncwarp();
        ptx::cluster_arrive();
        ptx::cluster_wait();
    }

    ptx::tc_fence_before();
    ptx::cluster_sync();
    if (warp == MMA_WARP) {
        ptx::tc_fence_after();
        ptx::tmem_dealloc(tmem, TMEM_COLS);
    }
}

// Verification kernel (tests only): same item list, same top-k bookkeeping, plain fp32
// FMA dot products on CUDA cores.  One block per item, thread == query row.
__global__ void __launch_bounds__(SCAN_M)
scan_topk_simt_kernel(const __half* __restrict__ X, const __half* __restrict__ Q,
                      const int32_t* __restrict__ pair_list, int B, const int32_t* __restrict__ items,
                      const int32_t* __restrict__ n_items_ptr, int64_t max_items, int d, int lmax,
                      float* __restrict__ part_sim, int32_t* __restrict__ part_row) {
    extern __shared__ __half xs[];  // one row of X
    int n_items = *n_items_ptr;
    if ((int64_t)n_items > max_items) n_items = (int)max_items;
    for (int it = blockIdx.x; it < n_items; it += gridDim.x) {
        const int4 w0 = *reinterpret_cast<const int4*>(items + (size_t)it * LMI_ITEM_WORDS);
        const int4 w1 = *reinterpret_cast<const int4*>(items + (size_t)it * LMI_ITEM_WORDS + 4);
        const int row0 = w0.x, nrows = w0.y, slot0 = w0.z, mvalid = w0.w, list0 = w1.x & 0xffff;
        const int row = threadIdx.x;
        const bool valid = row < mvalid;
        const int q = valid ? (pair_list[slot0 + row] / B) : 0;
        const __half* qrow = Q + (size_t)q * d;
        TopK tk;
        tk.reset();
        for (int r = 0; r < nrows; ++r) {
            __syncthreads();
            const __half* xrow = X + (size_t)(row0 + r) * d;
            for (int c = threadIdx.x; c < d; c += blockDim.x) xs[c] = xrow[c];
            __syncthreads();
            if (valid) {
                float acc = 0.f;
                for (int c = 0; c < d; ++c) acc = fmaf(__half2float(qrow[c]), __half2float(xs[c]), acc);
                tk.offer(acc, row0 + r);
            }
        }
        if (valid) {
            store_topk(tk, part_sim, part_row, ((size_t)(slot0 + row) * lmax + list0) * TOPK);
            tk.reset();  // the item's second list (the tcgen05 kernel's other column half) stays empty
            store_topk(tk, part_sim, part_row, ((size_t)(slot0 + row) * lmax + list0 + 1) * TOPK);
        }
    }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
    static EncodeTiledFn fn = nullptr;
    if (fn) return fn;
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) != cudaSuccess ||
        qres != cudaDriverEntryPointSuccess)
        return nullptr;
    fn = reinterpret_cast<EncodeTiledFn>(p);
    return fn;
}

}  // namespace lmi

#ifdef LMI_PROFILE
extern "C" int lmi_debug_scan_profile(long long* host_out /* [148*8] */) {
    return (int)cudaMemcpyFromSymbol(host_out, lmi::g_scan_prof, sizeof(long long) * 148 * 8);
}
#endif

extern "C" int lmi_scan_topk(const void* X, int64_t n, int d, const void* Q, int nq, int B,
                             const int32_t* pair_list, const int32_t* items, const int32_t* n_items,
                             int64_t max_items, int lmax, float* part_sim, int32_t* part_row, int impl,
                             uint32_t* gate, void* stream) {
    if (!X || !Q || !pair_list || !items || !n_items || !part_sim || !part_row) return LMI_E_BADARG;
    if (n <= 0 || d <= 0 || nq <= 0 || B <= 0 || lmax <= 0 || max_items <= 0) return LMI_E_BADARG;
    if (n > 0x7fffffffLL) return LMI_E_UNSUPPORTED;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    int dev = 0, sms = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (sms <= 0) sms = 148;

    const int flags = impl >> 8;
    impl &= 0xff;
    if (impl == 1) {
        const int grid = (int)(max_items < (int64_t)sms * 8 ? max_items : (int64_t)sms * 8);
        lmi::scan_topk_simt_kernel<<<grid, lmi::SCAN_M, (size_t)d * sizeof(__half), st>>>(
            static_cast<const __half*>(X), static_cast<const __half*>(Q), pair_list, B, items, n_items, max_items, d,
            lmax, part_sim, part_row);
        return static_cast<int>(cudaGetLastError());
    }
    if (impl != 0 && impl != 2 && impl != 3 && impl != 4) return LMI_E_BADARG;  // 3 / 4: 128-row score tiles
    if (d % lmi::SCAN_KBLK != 0 || d > lmi::SCAN_MAX_D) return LMI_E_UNSUPPORTED;
    if ((reinterpret_cast<uintptr_t>(X) & 15) != 0 || (reinterpret_cast<uintptr_t>(Q) & 15) != 0) return LMI_E_BADARG;

    lmi::EncodeTiledFn encode = lmi::get_encode_fn();
    if (!encode) return LMI_E_DRIVER;
    CUtensorMap tmap;
    const cuuint64_t gdim[2] = {(cuuint64_t)d, (cuuint64_t)n};
    const cuuint64_t gstride[1] = {(cuuint64_t)d * sizeof(__half)};
    const cuuint32_t box[2] = {(cuuint32_t)lmi::SCAN_KBLK, (cuuint32_t)(impl >= 3 ? 128 : lmi::SCAN_N)};
    const cuuint32_t estride[2] = {1, 1};
    CUresult cr = encode(&tmap, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<void*>(X), gdim, gstride, box, estride,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (cr != CUDA_SUCCESS) return LMI_E_DRIVER;

    // the opt-in to > 48 KB of dynamic shared memory is per device (context): remember it per device id
    static bool attr_set[64] = {};
    if (dev < 0 || dev >= 64 || !attr_set[dev]) {
        cudaError_t e = cudaFuncSetAttribute(lmi::scan_topk_tc_kernel<64, false>,
                                             cudaFuncAttributeMaxDynamicSharedMemorySize, (int)lmi::SCAN_SMEM_BYTES);
        if (e == cudaSuccess)
            e = cudaFuncSetAttribute(lmi::scan_topk_tc_kernel<128, false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     (int)lmi::SCAN_SMEM_BYTES);
        if (e == cudaSuccess)
            e = cudaFuncSetAttribute(lmi::scan_topk_tc_kernel<128, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     (int)lmi::SCAN_SPLIT_SMEM_BYTES);
        if (e != cudaSuccess) return (int)e;
        if (dev >= 0 && dev < 64) attr_set[dev] = true;
    }
    if (impl == 2) {
        // cluster-of-2 kernel: TMA multicast between the query tiles of one (bucket, row range)
        static bool pair_attr_set[64] = {};
        if (dev < 0 || dev >= 64 || !pair_attr_set[dev]) {
            cudaError_t e = cudaFuncSetAttribute(lmi::scan_topk_pair_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                 (int)lmi::SCAN_SMEM_BYTES);
            if (e != cudaSuccess) return (int)e;
            if (dev >= 0 && dev < 64) pair_attr_set[dev] = true;
        }
        int64_t clusters = (max_items + 1) / 2;
        if (clusters > sms / 2) clusters = sms / 2;
        if (clusters < 1) clusters = 1;
        cudaError_t me = cudaMemsetAsync(const_cast<int32_t*>(n_items) + 1, 0, sizeof(int32_t), st);
        if (me != cudaSuccess) return (int)me;
        lmi::scan_topk_pair_kernel<<<(unsigned)(2 * clusters), lmi::SCAN_THREADS, lmi::SCAN_SMEM_BYTES, st>>>(
            tmap, static_cast<const __half*>(Q), pair_list, B, items, n_items, max_items, d, lmax, part_sim, part_row);
        return static_cast<int>(cudaGetLastError());
    }
    const int grid = (int)(max_items < (int64_t)sms ? max_items : (int64_t)sms);
    int dbg = 0;
#ifdef LMI_PROFILE
    const char* dbg_env = getenv("LMI_SCAN_ABLATE");  // profiling ablations only; results are wrong when set
    dbg = dbg_env ? atoi(dbg_env) : 0;
#endif
    cudaError_t me = cudaMemsetAsync(const_cast<int32_t*>(n_items) + 1, 0, sizeof(int32_t), st);  // scheduler counter
    if (me != cudaSuccess) return (int)me;
    if (gate != nullptr) {
        me = cudaMemsetAsync(gate, 0, sizeof(uint32_t) * (size_t)nq * B * lmax, st);  // 0 = below every similarity
        if (me != cudaSuccess) return (int)me;
    }
    if (impl == 4)
        lmi::scan_topk_tc_kernel<128, true><<<grid, lmi::SCAN_THREADS, lmi::SCAN_SPLIT_SMEM_BYTES, st>>>(
            tmap, static_cast<const __half*>(Q), pair_list, B, items, n_items, max_items, d, lmax, part_sim, part_row,
            gate, (size_t)nq * B, flags, dbg);
    else if (impl == 3)
        lmi::scan_topk_tc_kernel<128, false><<<grid, lmi::SCAN_THREADS, lmi::SCAN_SMEM_BYTES, st>>>(
            tmap, static_cast<const __half*>(Q), pair_list, B, items, n_items, max_items, d, lmax, part_sim, part_row,
            gate, (size_t)nq * B, flags, dbg);
    else
        lmi::scan_topk_tc_kernel<64, false><<<grid, lmi::SCAN_THREADS, lmi::SCAN_SMEM_BYTES, st>>>(
            tmap, static_cast<const __half*>(Q), pair_list, B, items, n_items, max_items, d, lmax, part_sim, part_row,
            gate, (size_t)nq * B, flags, dbg);
    return static_cast<int>(cudaGetLastError());
}
