// Multi-level navigation: the batched best-first search of
// LearnedIndex._precompute_bucket_order (search/li/LearnedIndex.py:216-252) with
// _visit_internal_nodes (:254-301), _visit_buckets (:303-325) and PriorityQueue.py:7-94.
//
// The reference pops, per query and per iteration, the most probable entry of a sorted
// per-query array, runs the popped node's model on the queries that popped it, and pushes
// the children with their RAW child probabilities.  A node's probability does not depend
// on when it is evaluated, so here every internal model is evaluated for the whole query
// batch up front (dense grouped MLP -> `node_prob[nq, n_nodes]`, one column per tree node)
// and the queue becomes a per-query FRONTIER over that table: one warp per query keeps the
// frontier in shared memory (prob >= 0 while pushed-and-not-popped, -1 otherwise), finds its
// maximum with a strided scan + shuffle reduction, then either opens the node's children
// (internal node) or appends the node's bucket to the query's bucket order.
//
// Nodes are numbered breadth first: depth-1 nodes 0..C1-1 (id == class of the root model),
// then all depth-2 nodes parent-major, ...; the children of a node are contiguous.
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/lmi_b200.h"

namespace lmi {

__device__ __forceinline__ float nav_warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
__device__ __forceinline__ float nav_warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// softmax(dim=1) of logits[nq, C] (model.py:238) written to probs[q * stride + col0 + c].
// Same arithmetic as softmax_topn_kernel: max-subtracted expf, sum by warp tree, division.
__global__ void softmax_probs_kernel(const float* __restrict__ logits, int nq, int C, float* __restrict__ probs,
                                     int64_t stride, int64_t col0) {
    const int warps = blockDim.x >> 5;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int q = blockIdx.x * warps + warp;
    if (q >= nq) return;
    const float* src = logits + (size_t)q * C;
    float* dst = probs + (size_t)q * stride + col0;
    float mx = -INFINITY;
    for (int c = lane; c < C; c += 32) mx = fmaxf(mx, src[c]);
    mx = nav_warp_max(mx);
    float sum = 0.f;
    for (int c = lane; c < C; c += 32) sum += expf(src[c] - mx);
    sum = nav_warp_sum(sum);
    for (int c = lane; c < C; c += 32) dst[c] = expf(src[c] - mx) / sum;
}

// In-place softmax of column segments (one per model) of the node table; one warp per (query, segment).
__global__ void softmax_segments_kernel(float* __restrict__ table, int nq, int64_t ld,
                                        const int32_t* __restrict__ seg_start, const int32_t* __restrict__ seg_len,
                                        int n_seg) {
    const int warps = blockDim.x >> 5;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int64_t w = (int64_t)blockIdx.x * warps + warp;
    if (w >= (int64_t)nq * n_seg) return;
    const int q = (int)(w / n_seg), s = (int)(w % n_seg);
    const int C = seg_len[s];
    if (C <= 0) return;
    float* row = table + (size_t)q * ld + seg_start[s];
    float mx = -INFINITY;
    for (int c = lane; c < C; c += 32) mx = fmaxf(mx, row[c]);
    mx = nav_warp_max(mx);
    float sum = 0.f;
    for (int c = lane; c < C; c += 32) sum += expf(row[c] - mx);
    sum = nav_warp_sum(sum);
    for (int c = lane; c < C; c += 32) row[c] = expf(row[c] - mx) / sum;
}

// One warp per query.  smem: frontier[warps][n_nodes].
__global__ void navigate_tree_kernel(const float* __restrict__ node_prob, int nq, int n_nodes,
                                     const int32_t* __restrict__ first_child, const int32_t* __restrict__ n_children,
                                     const int32_t* __restrict__ node_bucket, int n_roots, int B,
                                     int32_t* __restrict__ order) {
    extern __shared__ float nav_sm[];
    const int warps = blockDim.x >> 5;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int q = blockIdx.x * warps + warp;
    if (q >= nq) return;
    float* fr = nav_sm + (size_t)warp * n_nodes;
    const float* prob = node_prob + (size_t)q * n_nodes;
    // seed: the level-1 nodes with the root model's probabilities (:222-227)
    for (int v = lane; v < n_nodes; v += 32) fr[v] = v < n_roots ? prob[v] : -1.f;
    __syncwarp();

    int found = 0;
    while (found < B) {
        // pop: the most probable frontier node (ties: lowest node id)
        float best = -1.f;
        int arg = 0x7fffffff;
        for (int v = lane; v < n_nodes; v += 32) {
            const float p = fr[v];
            if (p > best) {
                best = p;
                arg = v;
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
        if (arg == 0x7fffffff) break;  // frontier exhausted (the host rejects such budgets up front)
        const int nc = n_children[arg];
        const int fc = first_child[arg];
        const int bucket = node_bucket[arg];
        if (lane == 0) fr[arg] = -1.f;
        // internal node: its children enter the frontier with their own model's probabilities (:283-299)
        for (int j = lane; j < nc; j += 32) fr[fc + j] = prob[fc + j];
        // bucket: next slot of this query's order (:321-325)
        if (bucket >= 0) {
            if (lane == 0) order[(size_t)q * B + found] = bucket;
            ++found;
        }
        __syncwarp();
    }
    for (int r = found + lane; r < B; r += 32) order[(size_t)q * B + r] = -1;  // EMPTY_VALUE
}

}  // namespace lmi

extern "C" int lmi_softmax_probs(const float* logits, int nq, int C, float* probs, int64_t out_stride,
                                 int64_t out_col0, void* stream) {
    if (!logits || !probs || nq < 0 || C <= 0 || out_col0 < 0 || out_stride < out_col0 + C) return LMI_E_BADARG;
    if (nq == 0) return 0;
    const int warps = 8;
    lmi::softmax_probs_kernel<<<(nq + warps - 1) / warps, warps * 32, 0, static_cast<cudaStream_t>(stream)>>>(
        logits, nq, C, probs, out_stride, out_col0);
    return static_cast<int>(cudaGetLastError());
}

extern "C" int lmi_softmax_segments(float* table, int nq, int64_t ld, const int32_t* seg_start,
                                    const int32_t* seg_len, int n_seg, void* stream) {
    if (!table || !seg_start || !seg_len || nq < 0 || ld <= 0 || n_seg <= 0) return LMI_E_BADARG;
    if (nq == 0) return 0;
    const int warps = 8;
    const int64_t blocks = ((int64_t)nq * n_seg + warps - 1) / warps;
    if (blocks > 0x7fffffff) return LMI_E_UNSUPPORTED;
    lmi::softmax_segments_kernel<<<(unsigned)blocks, warps * 32, 0, static_cast<cudaStream_t>(stream)>>>(
        table, nq, ld, seg_start, seg_len, n_seg);
    return static_cast<int>(cudaGetLastError());
}

extern "C" int lmi_navigate_tree(const float* node_prob, int nq, int n_nodes, const int32_t* first_child,
                                 const int32_t* n_children, const int32_t* node_bucket, int n_roots, int B,
                                 int32_t* bucket_order, void* stream) {
    if (!node_prob || !first_child || !n_children || !node_bucket || !bucket_order) return LMI_E_BADARG;
    if (nq < 0 || n_nodes <= 0 || n_roots <= 0 || n_roots > n_nodes || B <= 0) return LMI_E_BADARG;
    if (nq == 0) return 0;
    int warps = 8;
    while (warps > 1 && (size_t)warps * n_nodes * sizeof(float) > 64 * 1024) warps >>= 1;
    const size_t smem = (size_t)warps * n_nodes * sizeof(float);
    if (smem > 200 * 1024) return LMI_E_UNSUPPORTED;
    if (smem > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(lmi::navigate_tree_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             (int)smem);
        if (e != cudaSuccess) return (int)e;
    }
    lmi::navigate_tree_kernel<<<(nq + warps - 1) / warps, warps * 32, smem, static_cast<cudaStream_t>(stream)>>>(
        node_prob, nq, n_nodes, first_child, n_children, node_bucket, n_roots, B, bucket_order);
    return static_cast<int>(cudaGetLastError());
}
