// Query-to-bucket inversion: the device-side replacement for the B*C calls of
// filter_path_idxs (search/li/utils.py:61-65; LearnedIndex.py:353).  A stable counting
// sort of the nq*B (query, probe-rank) pairs by probed bucket, so inside a bucket the
// pairs keep ascending flat index p = q*B + r (the order np.where yields), followed by
// the cut of the scan into work items.  No atomics: every warp owns one contiguous
// chunk of pairs and one private histogram row, so the result is deterministic.
#include <cuda_runtime.h>
#include <stdint.h>

#include <mutex>

#include "../../include/lmi_b200.h"

namespace lmi {

constexpr int INV_WARPS_PER_BLOCK = 8;
constexpr int SCAN_M_TILE = 128;  // queries per scan work item (TMEM lanes)

__host__ __device__ inline int inv_num_chunks(int64_t P) {
    int64_t g = P / 256;
    if (g < 64) g = 64;
    if (g > 4096) g = 4096;
    return (int)g;
}
__host__ __device__ inline int64_t inv_chunk_len(int64_t P, int G) {
    int64_t c = (P + G - 1) / G;
    return (c + 31) / 32 * 32;
}

// pass 1: hist[w][c] = number of pairs of chunk w that probe bucket c
// a probe counts only if its bucket id is valid AND the bucket holds rows in THIS layout: on a rank that
// owns 1/world of the buckets the other probes are dropped right here (they need no slot, no work item and
// no partial lists; the merge gives them the reference's empty-bucket filler)
__device__ __forceinline__ int live_key(int key, int C, const int64_t* __restrict__ bucket_off) {
    if (key < 0 || key >= C) return -1;
    return bucket_off[key + 1] > bucket_off[key] ? key : -1;
}

__global__ void invert_hist_kernel(const int32_t* __restrict__ order, int64_t P, int C, int G, int64_t chunk,
                                   const int64_t* __restrict__ bucket_off, int32_t* __restrict__ hist) {
    const int w = blockIdx.x * INV_WARPS_PER_BLOCK + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (w >= G) return;
    int32_t* row = hist + (size_t)w * C;
    for (int c = lane; c < C; c += 32) row[c] = 0;
    __syncwarp();
    const int64_t lo = (int64_t)w * chunk;
    const int64_t hi = min(P, lo + chunk);
    for (int64_t base = lo; base < hi; base += 32) {
        const int64_t p = base + lane;
        int key = -1;
        if (p < hi) key = live_key(order[p], C, bucket_off);
        const unsigned peers = __match_any_sync(0xffffffffu, key);
        if (key >= 0 && (peers & ((1u << lane) - 1)) == 0) row[key] += __popc(peers);
        __syncwarp();
    }
}

template <typename T>
__device__ __forceinline__ T block_exclusive_scan(T v, T* warp_tot, T& total) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    T incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const T t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
    }
    if (lane == 31) warp_tot[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        T t = (lane < (int)(blockDim.x >> 5)) ? warp_tot[lane] : T(0);
        T ti = t;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const T u = __shfl_up_sync(0xffffffffu, ti, o);
            if (lane >= o) ti += u;
        }
        warp_tot[lane] = ti - t;           // exclusive warp base
        if (lane == 31) warp_tot[32] = ti;  // grand total
    }
    __syncthreads();
    const T base = warp_tot[warp];
    total = warp_tot[32];
    __syncthreads();
    return base + incl - v;
}

// pass 2 (one block): column scan of hist over chunks, bucket offsets, work items.
// Work items are listed in bucket order and drawn in that order by the scan's persistent CTAs, so the
// LAST items decide how ragged the end of the launch is.  Buckets are therefore cut with `range_len` rows
// per item until (1000 - tail_permille)/1000 of the launch's work (rows x query tiles) has been listed and
// with the shorter `tail_range_len` after that: long items while there is plenty of work (the per-item
// query gather and top-10 warm-up amortise), short ones where the CTAs run dry.
__global__ void __launch_bounds__(1024)
invert_scan_kernel(int32_t* __restrict__ hist, int C, int G, const int64_t* __restrict__ bucket_off, int range_len,
                   int tail_range_len, int tail_permille, int32_t* __restrict__ counts, int32_t* __restrict__ offsets,
                   int32_t* __restrict__ bucket_lists, int32_t* __restrict__ items, int32_t* __restrict__ n_items,
                   int64_t max_items) {
    __shared__ int warp_tot[33];
    __shared__ long long warp_tot_ll[33];
    __shared__ int carry_pairs, carry_items;
    __shared__ long long carry_work, total_work;
    if (threadIdx.x == 0) {
        carry_pairs = 0;
        carry_items = 0;
        carry_work = 0;
    }
    __syncthreads();
    // column scan of the per-chunk histograms: one warp per bucket, 32 chunks per trip (independent
    // loads), exclusive scan across the lanes; the bucket totals land in shared memory
    for (int c = threadIdx.x >> 5; c < C; c += (int)(blockDim.x >> 5)) {
        const int lane = threadIdx.x & 31;
        int run = 0;
        for (int g0 = 0; g0 < G; g0 += 32) {
            const int g = g0 + lane;
            const int t = (g < G) ? hist[(size_t)g * C + c] : 0;
            int incl = t;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const int u = __shfl_up_sync(0xffffffffu, incl, o);
                if (lane >= o) incl += u;
            }
            if (g < G) hist[(size_t)g * C + c] = run + incl - t;
            run += __shfl_sync(0xffffffffu, incl, 31);
        }
        if (lane == 0) counts[c] = run;
    }
    __threadfence_block();
    __syncthreads();
    // total work of the launch, in (row, query tile) units
    {
        long long w = 0;
        for (int c = threadIdx.x; c < C; c += blockDim.x) {
            const int m = counts[c];
            if (m > 0) w += (bucket_off[c + 1] - bucket_off[c]) * (long long)((m + SCAN_M_TILE - 1) / SCAN_M_TILE);
        }
        long long tot;
        block_exclusive_scan<long long>(w, warp_tot_ll, tot);
        if (threadIdx.x == 0) total_work = tot;
        __syncthreads();
    }
    const bool two_level = tail_permille > 0 && tail_range_len > 0 && tail_range_len < range_len;
    const long long tail_from = two_level ? total_work / 1000 * (1000 - tail_permille) : total_work + 1;
    for (int c0 = 0; c0 < C; c0 += blockDim.x) {
        const int c = c0 + threadIdx.x;
        int m = 0;
        if (c < C) m = counts[c];
        int tot_pairs;
        const int off = block_exclusive_scan<int>(m, warp_tot, tot_pairs) + carry_pairs;
        if (c < C) offsets[c] = off;

        int nt = 0;
        int64_t lo = 0, hi = 0;
        if (c < C && m > 0) {
            lo = bucket_off[c];
            hi = bucket_off[c + 1];
            nt = (m + SCAN_M_TILE - 1) / SCAN_M_TILE;
        }
        long long tot_w;
        const long long before = block_exclusive_scan<long long>((hi - lo) * (long long)nt, warp_tot_ll, tot_w) + carry_work;
        const int rl = (before >= tail_from) ? tail_range_len : range_len;
        int nr = 0;
        int64_t len = 0;
        if (hi > lo) {
            // ranges of equal length (a multiple of 64 rows, <= rl): the items of a bucket then take
            // equally long, which keeps the scan's in-order dynamic schedule balanced and the query
            // tiles of one range in step
            nr = (int)((hi - lo + rl - 1) / rl);
            len = ((hi - lo + nr - 1) / nr + 63) / 64 * 64;
        }
        if (c < C) bucket_lists[c] = nr * LMI_LISTS_PER_ITEM;
        int tot_items;
        const int ibase = block_exclusive_scan<int>(nr * nt, warp_tot, tot_items) + carry_items;
        for (int r = 0; r < nr; ++r) {
            const int64_t row0 = lo + (int64_t)r * len;
            const int nrows = (int)min(len, hi - row0);
            for (int t = 0; t < nt; ++t) {
                const int64_t idx = (int64_t)ibase + (int64_t)r * nt + t;
                if (idx >= max_items) continue;
                int4* dst = reinterpret_cast<int4*>(items + idx * LMI_ITEM_WORDS);
                dst[0] = make_int4((int)row0, nrows, off + t * SCAN_M_TILE, min(SCAN_M_TILE, m - t * SCAN_M_TILE));
                dst[1] = make_int4(r * LMI_LISTS_PER_ITEM | (nr * LMI_LISTS_PER_ITEM) << 16, c, t, nt);
            }
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            carry_pairs += tot_pairs;
            carry_items += tot_items;
            carry_work += tot_w;
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        offsets[C] = carry_pairs;
        n_items[0] = (int)min((int64_t)carry_items, max_items);
        n_items[1] = 0;  // the scan's item counter (lmi_scan_topk draws items from it)
    }
}

// pass 3: stable scatter, same chunking as pass 1
__global__ void invert_scatter_kernel(const int32_t* __restrict__ order, int64_t P, int C, int G, int64_t chunk,
                                      const int64_t* __restrict__ bucket_off, int32_t* __restrict__ hist,
                                      const int32_t* __restrict__ offsets,
                                      int32_t* __restrict__ pair_list, int32_t* __restrict__ slot_of_pair) {
    const int w = blockIdx.x * INV_WARPS_PER_BLOCK + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (w >= G) return;
    int32_t* row = hist + (size_t)w * C;
    const int64_t lo = (int64_t)w * chunk;
    const int64_t hi = min(P, lo + chunk);
    for (int64_t base = lo; base < hi; base += 32) {
        const int64_t p = base + lane;
        int key = -1;
        if (p < hi) key = live_key(order[p], C, bucket_off);
        const unsigned peers = __match_any_sync(0xffffffffu, key);
        const int rank = __popc(peers & ((1u << lane) - 1));
        const int leader = __ffs(peers) - 1;
        int cur = 0;
        if (key >= 0 && rank == 0) {
            cur = row[key];
            row[key] = cur + __popc(peers);
        }
        cur = __shfl_sync(0xffffffffu, cur, leader);
        if (p < hi) {
            if (key >= 0) {
                const int slot = offsets[key] + cur + rank;
                pair_list[slot] = (int32_t)p;
                slot_of_pair[p] = slot;
            } else {
                slot_of_pair[p] = -1;
            }
        }
        __syncwarp();
    }
}

}  // namespace lmi

extern "C" int64_t lmi_max_items(int nq, int B, int C, int rmax) {
    const int64_t P = (int64_t)nq * B;
    const int64_t touched = P < C ? P : C;
    return (int64_t)rmax * (P / lmi::SCAN_M_TILE + touched + 1);
}

extern "C" int64_t lmi_invert_workspace_ints(int nq, int B, int C) {
    const int64_t P = (int64_t)nq * B;
    return (int64_t)lmi::inv_num_chunks(P) * C;
}

extern "C" int64_t lmi_workspace_layout(int nq, int B, int C, int rmax, int d_pad, int64_t* offsets_host) {
    if (nq <= 0 || B <= 0 || C <= 0 || rmax <= 0 || d_pad < 0) return LMI_E_BADARG;
    const int64_t P = (int64_t)nq * B;
    const int64_t lists = P * rmax * LMI_LISTS_PER_ITEM * LMI_PER_BUCKET_K;
    const int64_t bytes[LMI_WS_ARRAYS] = {
        4LL * C,                                            // LMI_WS_COUNTS
        4LL * (C + 1),                                      // LMI_WS_OFFSETS
        4LL * C,                                            // LMI_WS_BUCKET_LISTS
        4LL * P,                                            // LMI_WS_PAIR_LIST
        4LL * P,                                            // LMI_WS_SLOT_OF_PAIR
        4LL * LMI_ITEM_WORDS * lmi_max_items(nq, B, C, rmax),  // LMI_WS_ITEMS
        4LL * 2,                                            // LMI_WS_N_ITEMS
        4LL * lmi_invert_workspace_ints(nq, B, C),          // LMI_WS_INVERT_SCRATCH
        4LL * lists,                                        // LMI_WS_PART_SIM
        4LL * lists,                                        // LMI_WS_PART_ROW
        2LL * nq * d_pad,                                   // LMI_WS_Q16
        4LL * P * rmax * LMI_LISTS_PER_ITEM,                // LMI_WS_GATE
    };
    int64_t at = 0;
    for (int i = 0; i < LMI_WS_ARRAYS; ++i) {
        if (offsets_host) offsets_host[i] = at;
        at += (bytes[i] + 255) / 256 * 256;
    }
    if (offsets_host) offsets_host[LMI_WS_ARRAYS] = at;
    return at;
}

extern "C" int64_t lmi_workspace_bytes(int nq, int B, int C, int rmax) {
    return lmi_workspace_layout(nq, B, C, rmax, 0, nullptr);
}

extern "C" int lmi_invert(const int32_t* bucket_order, int nq, int B, int C, const int64_t* bucket_off,
                          int range_len, int tail_range_len, int tail_permille, int32_t* counts, int32_t* offsets,
                          int32_t* bucket_lists, int32_t* pair_list, int32_t* slot_of_pair, int32_t* items,
                          int32_t* n_items, int64_t max_items, int32_t* workspace, void* stream) {
    if (!bucket_order || !bucket_off || !counts || !offsets || !bucket_lists || !pair_list || !slot_of_pair ||
        !items || !n_items || !workspace)
        return LMI_E_BADARG;
    if (nq <= 0 || B <= 0 || C <= 0 || range_len <= 0 || range_len % 64 != 0 || max_items <= 0) return LMI_E_BADARG;
    if (tail_permille < 0 || tail_permille > 1000 || tail_range_len < 0 || tail_range_len % 64 != 0) return LMI_E_BADARG;
    const int64_t P = (int64_t)nq * B;
    if (P > 0x7fffffffLL) return LMI_E_UNSUPPORTED;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int G = lmi::inv_num_chunks(P);
    const int64_t chunk = lmi::inv_chunk_len(P, G);
    const int blocks = (G + lmi::INV_WARPS_PER_BLOCK - 1) / lmi::INV_WARPS_PER_BLOCK;
    lmi::invert_hist_kernel<<<blocks, lmi::INV_WARPS_PER_BLOCK * 32, 0, st>>>(bucket_order, P, C, G, chunk, bucket_off,
                                                                              workspace);
    lmi::invert_scan_kernel<<<1, 1024, 0, st>>>(workspace, C, G, bucket_off, range_len, tail_range_len, tail_permille,
                                                counts, offsets, bucket_lists, items, n_items, max_items);
    lmi::invert_scatter_kernel<<<blocks, lmi::INV_WARPS_PER_BLOCK * 32, 0, st>>>(bucket_order, P, C, G, chunk,
                                                                                 bucket_off, workspace, offsets,
                                                                                 pair_list, slot_of_pair);
    return static_cast<int>(cudaGetLastError());
}

// ---- lmi_invert_static: the same three passes with the histogram scratch inside the library ----------------
//
// The scratch lives in the module image instead of in a caller allocation: LMI_INVERT_STATIC_MAX_INTS int32
// (4 MiB) that every GPU which loads this library carries until the process exits.  `g_invert_scratch` below is
// what HOST code sees of it -- a stand-in the runtime uses as a lookup key, NOT an address a kernel can
// dereference -- so the host asks the runtime where the array sits on the CURRENT device
// (cudaGetSymbolAddress) and hands that pointer to the kernels like any other workspace.
namespace lmi {

static __device__ __align__(256) int32_t g_invert_scratch[LMI_INVERT_STATIC_MAX_INTS];

constexpr int STATIC_MAX_DEVICES = 64;
static std::mutex g_static_mu;
static int32_t* g_static_ptr[STATIC_MAX_DEVICES];  // per-device address of g_invert_scratch, filled on first use

// Address of the scratch array on the device that is current on the calling thread.  The first call per
// device resolves the symbol (this may load the module on that device, which is why it has to happen
// outside stream capture); later calls read the cached pointer.
static int static_scratch(int32_t** out) {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return static_cast<int>(e);
    if (dev < 0 || dev >= STATIC_MAX_DEVICES) return LMI_E_UNSUPPORTED;
    std::lock_guard<std::mutex> lock(g_static_mu);
    if (!g_static_ptr[dev]) {
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_invert_scratch);
        if (e != cudaSuccess) return static_cast<int>(e);
        if (!p) return LMI_E_DRIVER;
        g_static_ptr[dev] = static_cast<int32_t*>(p);
    }
    *out = g_static_ptr[dev];
    return 0;
}

}  // namespace lmi

extern "C" int lmi_invert_static(const int32_t* bucket_order, int nq, int B, int C, const int64_t* bucket_off,
                                 int range_len, int tail_range_len, int tail_permille, int32_t* counts,
                                 int32_t* offsets, int32_t* bucket_lists, int32_t* pair_list, int32_t* slot_of_pair,
                                 int32_t* items, int32_t* n_items, int64_t max_items, void* stream) {
    if (!bucket_order || !bucket_off || !counts || !offsets || !bucket_lists || !pair_list || !slot_of_pair ||
        !items || !n_items)
        return LMI_E_BADARG;
    if (nq <= 0 || B <= 0 || C <= 0 || range_len <= 0 || range_len % 64 != 0 || max_items <= 0) return LMI_E_BADARG;
    if (tail_permille < 0 || tail_permille > 1000 || tail_range_len < 0 || tail_range_len % 64 != 0) return LMI_E_BADARG;
    const int64_t P = (int64_t)nq * B;
    if (P > 0x7fffffffLL) return LMI_E_UNSUPPORTED;
    const int G = lmi::inv_num_chunks(P);
    // the whole scratch need of this shape, checked BEFORE anything is launched: chunks x buckets histogram cells
    if ((int64_t)G * C > (int64_t)LMI_INVERT_STATIC_MAX_INTS) return LMI_E_SCRATCH_LIMIT;
    int32_t* scratch = nullptr;
    const int rc = lmi::static_scratch(&scratch);
    if (rc != 0) return rc;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int64_t chunk = lmi::inv_chunk_len(P, G);
    const int blocks = (G + lmi::INV_WARPS_PER_BLOCK - 1) / lmi::INV_WARPS_PER_BLOCK;
    // nothing to clear by hand: pass 1 zeroes every histogram row it is about to count into, so what the
    // previous call (of any shape) left in the array is never read
    lmi::invert_hist_kernel<<<blocks, lmi::INV_WARPS_PER_BLOCK * 32, 0, st>>>(bucket_order, P, C, G, chunk, bucket_off,
                                                                              scratch);
    lmi::invert_scan_kernel<<<1, 1024, 0, st>>>(scratch, C, G, bucket_off, range_len, tail_range_len, tail_permille,
                                                counts, offsets, bucket_lists, items, n_items, max_items);
    lmi::invert_scatter_kernel<<<blocks, lmi::INV_WARPS_PER_BLOCK * 32, 0, st>>>(bucket_order, P, C, G, chunk,
                                                                                 bucket_off, scratch, offsets,
                                                                                 pair_list, slot_of_pair);
    return static_cast<int>(cudaGetLastError());
}
