// Query-to-bucket inversion: the device-side replacement for the B*C calls of
// filter_path_idxs (search/li/utils.py:61-65; LearnedIndex.py:353).  A stable counting
// sort of the nq*B (query, probe-rank) pairs by probed bucket, so inside a bucket the
// pairs keep ascending flat index p = q*B + r (the order np.where yields), followed by
// the cut of the scan into work items.  No atomics: every warp owns one contiguous
// chunk of pairs and one private histogram row, so the result is deterministic.
#include <cuda_runtime.h>
#include <stdint.h>

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
__global__ void invert_hist_kernel(const int32_t* __restrict__ order, int64_t P, int C, int G, int64_t chunk,
                                   int32_t* __restrict__ hist) {
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
        if (p < hi) {
            key = order[p];
            if (key < 0 || key >= C) key = -1;
        }
        const unsigned peers = __match_any_sync(0xffffffffu, key);
        if (key >= 0 && (peers & ((1u << lane) - 1)) == 0) row[key] += __popc(peers);
        __syncwarp();
    }
}

__device__ __forceinline__ int block_exclusive_scan(int v, int* warp_tot, int& total) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
    }
    if (lane == 31) warp_tot[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        int t = (lane < (int)(blockDim.x >> 5)) ? warp_tot[lane] : 0;
        int ti = t;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const int u = __shfl_up_sync(0xffffffffu, ti, o);
            if (lane >= o) ti += u;
        }
        warp_tot[lane] = ti - t;           // exclusive warp base
        if (lane == 31) warp_tot[32] = ti;  // grand total
    }
    __syncthreads();
    const int base = warp_tot[warp];
    total = warp_tot[32];
    __syncthreads();
    return base + incl - v;
}

// pass 2 (one block): column scan of hist over chunks, bucket offsets, work items.
__global__ void __launch_bounds__(1024)
invert_scan_kernel(int32_t* __restrict__ hist, int C, int G, const int64_t* __restrict__ bucket_off, int range_len,
                   int32_t* __restrict__ counts, int32_t* __restrict__ offsets, int32_t* __restrict__ bucket_lists,
                   int32_t* __restrict__ items, int32_t* __restrict__ n_items, int64_t max_items) {
    __shared__ int warp_tot[33];
    __shared__ int carry_pairs, carry_items;
    if (threadIdx.x == 0) {
        carry_pairs = 0;
        carry_items = 0;
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
    for (int c0 = 0; c0 < C; c0 += blockDim.x) {
        const int c = c0 + threadIdx.x;
        int m = 0;
        if (c < C) m = counts[c];
        int tot_pairs;
        const int off = block_exclusive_scan(m, warp_tot, tot_pairs) + carry_pairs;
        if (c < C) offsets[c] = off;

        int nr = 0, nt = 0;
        int64_t lo = 0, hi = 0, len = 0;
        if (c < C && m > 0) {
            lo = bucket_off[c];
            hi = bucket_off[c + 1];
            if (hi > lo) {
                // ranges of equal length (a multiple of 64 rows, <= range_len): the items of a
                // bucket then take equally long, which keeps the scan's in-order dynamic schedule
                // balanced and the query tiles of one range in step
                nr = (int)((hi - lo + range_len - 1) / range_len);
                len = ((hi - lo + nr - 1) / nr + 63) / 64 * 64;
                nt = (m + SCAN_M_TILE - 1) / SCAN_M_TILE;
            }
        }
        if (c < C) bucket_lists[c] = nr * LMI_LISTS_PER_ITEM;
        int tot_items;
        const int ibase = block_exclusive_scan(nr * nt, warp_tot, tot_items) + carry_items;
        for (int r = 0; r < nr; ++r) {
            const int64_t row0 = lo + (int64_t)r * len;
            const int nrows = (int)min(len, hi - row0);
            for (int t = 0; t < nt; ++t) {
                const int64_t idx = (int64_t)ibase + (int64_t)r * nt + t;
                if (idx >= max_items) continue;
                int4* dst = reinterpret_cast<int4*>(items + idx * LMI_ITEM_WORDS);
                dst[0] = make_int4((int)row0, nrows, off + t * SCAN_M_TILE, min(SCAN_M_TILE, m - t * SCAN_M_TILE));
                dst[1] = make_int4(r * LMI_LISTS_PER_ITEM, c, 0, 0);
            }
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            carry_pairs += tot_pairs;
            carry_items += tot_items;
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        offsets[C] = carry_pairs;
        n_items[0] = (int)min((int64_t)carry_items, max_items);
    }
}

// pass 3: stable scatter, same chunking as pass 1
__global__ void invert_scatter_kernel(const int32_t* __restrict__ order, int64_t P, int C, int G, int64_t chunk,
                                      int32_t* __restrict__ hist, const int32_t* __restrict__ offsets,
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
        if (p < hi) {
            key = order[p];
            if (key < 0 || key >= C) key = -1;
        }
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

extern "C" int lmi_invert(const int32_t* bucket_order, int nq, int B, int C, const int64_t* bucket_off,
                          int range_len, int32_t* counts, int32_t* offsets, int32_t* bucket_lists,
                          int32_t* pair_list, int32_t* slot_of_pair, int32_t* items, int32_t* n_items,
                          int64_t max_items, int32_t* workspace, void* stream) {
    if (!bucket_order || !bucket_off || !counts || !offsets || !bucket_lists || !pair_list || !slot_of_pair ||
        !items || !n_items || !workspace)
        return LMI_E_BADARG;
    if (nq <= 0 || B <= 0 || C <= 0 || range_len <= 0 || range_len % 64 != 0 || max_items <= 0) return LMI_E_BADARG;
    const int64_t P = (int64_t)nq * B;
    if (P > 0x7fffffffLL) return LMI_E_UNSUPPORTED;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int G = lmi::inv_num_chunks(P);
    const int64_t chunk = lmi::inv_chunk_len(P, G);
    const int blocks = (G + lmi::INV_WARPS_PER_BLOCK - 1) / lmi::INV_WARPS_PER_BLOCK;
    lmi::invert_hist_kernel<<<blocks, lmi::INV_WARPS_PER_BLOCK * 32, 0, st>>>(bucket_order, P, C, G, chunk, workspace);
    lmi::invert_scan_kernel<<<1, 1024, 0, st>>>(workspace, C, G, bucket_off, range_len, counts, offsets,
                                                bucket_lists, items, n_items, max_items);
    lmi::invert_scatter_kernel<<<blocks, lmi::INV_WARPS_PER_BLOCK * 32, 0, st>>>(bucket_order, P, C, G, chunk,
                                                                                 workspace, offsets, pair_list,
                                                                                 slot_of_pair);
    return static_cast<int>(cudaGetLastError());
}
