// Per-query merge of the scan's partial lists (replaces distances = 1 - similarity, the
// local-row -> DataFrame-label mapping and the cross-round stable-argsort merge,
// search/li/LearnedIndex.py:368-371 and :125-146), plus the cross-GPU merge that follows
// the all-gather of per-rank results.
#include <cuda_runtime.h>
#include <float.h>
#include <math.h>
#include <stdint.h>

#include "../../include/lmi_b200.h"

namespace lmi {

constexpr int MK = LMI_PER_BUCKET_K;

struct Key {
    float d;
    int e;
};
__device__ __forceinline__ bool key_less(float d0, int e0, float d1, int e1) {
    return d0 < d1 || (d0 == d1 && e0 < e1);
}
__device__ __forceinline__ void warp_argmin(float& d, int& e) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float od = __shfl_xor_sync(0xffffffffu, d, o);
        const int oe = __shfl_xor_sync(0xffffffffu, e, o);
        if (key_less(od, oe, d, e)) {
            d = od;
            e = oe;
        }
    }
}

constexpr int MERGE_LPL = 8;  // lists per lane: up to 256 partial lists per (query, probe)

// One warp per query.  Shared memory per warp: B*10 (dist f32, id u32) pairs.
__global__ void merge_topk_kernel(const float* __restrict__ part_sim, const int32_t* __restrict__ part_row,
                                  const int32_t* __restrict__ slot_of_pair, const int32_t* __restrict__ order,
                                  const int64_t* __restrict__ bucket_off, const uint32_t* __restrict__ ids,
                                  const int32_t* __restrict__ bucket_lists, int nq, int B, int C, int lmax,
                                  int k_out, double* __restrict__ dists, uint32_t* __restrict__ nns,
                                  uint16_t* __restrict__ probe_out) {
    extern __shared__ float sm[];
    const int warps = blockDim.x >> 5;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int q = blockIdx.x * warps + warp;
    if (q >= nq) return;
    const int E = B * MK;
    float* e_dist = sm + (size_t)warp * 2 * E;
    uint32_t* e_id = reinterpret_cast<uint32_t*>(e_dist + E);

    // stage 1: every probe -> that bucket's top-10 as (distance, id), padded like the reference
    for (int r = 0; r < B; ++r) {
        const int p = q * B + r;
        const int b = order[p];
        int64_t lo = 0, hi = 0;
        if (b >= 0 && b < C) {
            lo = bucket_off[b];
            hi = bucket_off[b + 1];
        }
        if (hi <= lo) {
            // bucket without objects: the reference's per-round buffers stay (id 0, +inf) (:340-341)
            if (lane < MK) {
                e_dist[r * MK + lane] = INFINITY;
                e_id[r * MK + lane] = 0u;
            }
            continue;
        }
        const int nl = min(bucket_lists[b], lmax);
        const int slot = slot_of_pair[p];
        if (nl <= 32) {
            // common case: one list per lane, held entirely in registers; the winner's list shifts down
            const size_t lbase = ((size_t)slot * lmax + lane) * MK;
            float ls[MK];
            int lr[MK];
#pragma unroll
            for (int t = 0; t < MK; t += 2) {
                float2 sv = make_float2(-INFINITY, -INFINITY);
                int2 rv = make_int2(-1, -1);
                if (lane < nl) {
                    sv = *reinterpret_cast<const float2*>(part_sim + lbase + t);
                    rv = *reinterpret_cast<const int2*>(part_row + lbase + t);
                }
                ls[t] = sv.x;
                ls[t + 1] = sv.y;
                lr[t] = rv.x;
                lr[t + 1] = rv.y;
            }
            float res_d = 0.f;
            int res_row = 0x7fffffff;
#pragma unroll 1
            for (int t = 0; t < MK; ++t) {
                float wd = (lr[0] >= 0) ? -ls[0] : INFINITY;
                int wrow = (lr[0] >= 0) ? lr[0] : 0x7fffffff;
                warp_argmin(wd, wrow);
                if (lane == t) {
                    res_d = wd;
                    res_row = wrow;
                }
                if (wrow != 0x7fffffff && lr[0] == wrow) {
#pragma unroll
                    for (int u = 0; u + 1 < MK; ++u) {
                        ls[u] = ls[u + 1];
                        lr[u] = lr[u + 1];
                    }
                    lr[MK - 1] = -1;
                }
            }
            if (lane < MK) {
                if (res_row == 0x7fffffff) {
                    // fewer than 10 objects: faiss pads with I = -1 / D = -FLT_MAX, the reference maps
                    // -1 to the bucket's last label and 1 - (-FLT_MAX) to FLT_MAX (:368-370)
                    e_dist[r * MK + lane] = 1.0f - (-FLT_MAX);
                    e_id[r * MK + lane] = ids[hi - 1];
                } else {
                    e_dist[r * MK + lane] = 1.0f + res_d;
                    e_id[r * MK + lane] = ids[res_row];
                }
            }
            continue;
        }
        // lane walks the (descending) lists lane, lane+32, ...; a taken head is replaced by the
        // list's next entry
        const size_t base = (size_t)slot * lmax * MK;
        float hv[MERGE_LPL];
        int hrow[MERGE_LPL], head[MERGE_LPL];
#pragma unroll
        for (int u = 0; u < MERGE_LPL; ++u) {
            const int l = lane + 32 * u;
            hv[u] = -INFINITY;
            hrow[u] = -1;
            head[u] = 0;
            if (l < nl) {
                hv[u] = part_sim[base + (size_t)l * MK];
                hrow[u] = part_row[base + (size_t)l * MK];
            }
        }
        const uint32_t last_id = ids[hi - 1];
        for (int t = 0; t < MK; ++t) {
            // arg max by (similarity desc, row asc): the lower row wins a tie, as in one pass
            // over the bucket in row order
            float kd = INFINITY;
            int ke = 0x7fffffff, ku = 0;
#pragma unroll
            for (int u = 0; u < MERGE_LPL; ++u) {
                if (hrow[u] >= 0 && key_less(-hv[u], hrow[u], kd, ke)) {
                    kd = -hv[u];
                    ke = hrow[u];
                    ku = u;
                }
            }
            float wd = kd;
            int wrow = ke;
            warp_argmin(wd, wrow);
            const bool none = wrow == 0x7fffffff;
            if (lane == 0) {
                if (none) {
                    // fewer than 10 objects: faiss pads with I = -1 / D = -FLT_MAX, the reference maps
                    // -1 to the bucket's last label and 1 - (-FLT_MAX) to FLT_MAX (:368-370)
                    e_dist[r * MK + t] = 1.0f - (-FLT_MAX);
                    e_id[r * MK + t] = last_id;
                } else {
                    e_dist[r * MK + t] = 1.0f + wd;
                    e_id[r * MK + t] = ids[wrow];
                }
            }
            if (!none && ke == wrow) {  // rows are unique across lists: exactly one lane owns the winner
#pragma unroll
                for (int u = 0; u < MERGE_LPL; ++u) {
                    if (u == ku) {
                        const int l = lane + 32 * u;
                        ++head[u];
                        if (head[u] < MK) {
                            hv[u] = part_sim[base + (size_t)l * MK + head[u]];
                            hrow[u] = part_row[base + (size_t)l * MK + head[u]];
                        } else {
                            hrow[u] = -1;
                        }
                    }
                }
            }
        }
    }
    __syncwarp();

    // stage 2: k_out smallest by (distance, probe rank, position) == stable argsort over the rounds
    for (int t = 0; t < k_out; ++t) {
        float bd = INFINITY;
        int be = 0x7fffffff;
        for (int e = lane; e < E; e += 32) {
            const float dd = e_dist[e];
            if (!(dd != dd) && key_less(dd, e, bd, be)) {  // NaN marks a taken slot
                bd = dd;
                be = e;
            }
        }
        // +inf entries are legal candidates (empty-bucket fillers): order them by index
        float kd = bd;
        int ke = be;
        warp_argmin(kd, ke);
        if (lane == 0) {
            if (ke == 0x7fffffff) {
                dists[(size_t)q * k_out + t] = INFINITY;
                nns[(size_t)q * k_out + t] = 0u;
                if (probe_out) probe_out[(size_t)q * k_out + t] = 0xffffu;
            } else {
                dists[(size_t)q * k_out + t] = (double)e_dist[ke];
                nns[(size_t)q * k_out + t] = e_id[ke];
                if (probe_out) probe_out[(size_t)q * k_out + t] = (uint16_t)(ke / MK);  // probe rank of the entry
                e_dist[ke] = __int_as_float(0x7fc00000);
            }
        }
        __syncwarp();
    }
}

// Cross-rank key.  The unsharded search orders equal distances by (probe rank, position in the bucket's
// top-10) -- the stable argsort over the rounds, LearnedIndex.py:131-134.  Inside one rank's list that order is
// already in place; ACROSS ranks an equal distance is resolved by the entry's probe rank (carried per entry when
// the caller passes it), then by the rank: whole-bucket sharding puts different probes on different ranks, and
// row-slice sharding gives the lower rows of a bucket to the lower rank.  Without probe ranks: (distance, rank).
__device__ __forceinline__ bool rank_key_less(double d0, int p0, int e0, double d1, int p1, int e1) {
    return d0 < d1 || (d0 == d1 && (p0 < p1 || (p0 == p1 && e0 < e1)));
}

// rank w's ascending lists live at d_all + w*stride_d / i_all + w*stride_i (/ p_all + w*stride_p), each [nq, k];
// output [nq, k]
__global__ void merge_ranks_kernel(const double* __restrict__ d_all, const uint32_t* __restrict__ i_all,
                                   const uint16_t* __restrict__ p_all, int64_t stride_d, int64_t stride_i,
                                   int64_t stride_p, int world, int nq, int k, double* __restrict__ dists,
                                   uint32_t* __restrict__ nns) {
    const int warps = blockDim.x >> 5;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int q = blockIdx.x * warps + warp;
    if (q >= nq) return;
    // lane w (< world) walks rank w's ascending list; world <= 32
    int head = 0;
    const size_t base_d = (size_t)lane * stride_d + (size_t)q * k;
    const size_t base_i = (size_t)lane * stride_i + (size_t)q * k;
    const size_t base_p = (size_t)lane * stride_p + (size_t)q * k;
    bool live = lane < world;
    double hv = live ? d_all[base_d] : INFINITY;
    int hp = (live && p_all) ? (int)p_all[base_p] : 0;
    for (int t = 0; t < k; ++t) {
        double kd = hv;
        int kp = hp;
        int ke = live ? lane : 0x7fffffff;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const double od = __shfl_xor_sync(0xffffffffu, kd, o);
            const int op = __shfl_xor_sync(0xffffffffu, kp, o);
            const int oe = __shfl_xor_sync(0xffffffffu, ke, o);
            if (oe != 0x7fffffff && (ke == 0x7fffffff || rank_key_less(od, op, oe, kd, kp, ke))) {
                kd = od;
                kp = op;
                ke = oe;
            }
        }
        if (ke == 0x7fffffff) {
            if (lane == 0) {
                dists[(size_t)q * k + t] = INFINITY;
                nns[(size_t)q * k + t] = 0u;
            }
            continue;
        }
        if (lane == ke) {
            dists[(size_t)q * k + t] = hv;
            nns[(size_t)q * k + t] = i_all[base_i + head];
            ++head;
            if (head < k) {
                hv = d_all[base_d + head];
                if (p_all) hp = (int)p_all[base_p + head];
            } else {
                live = false;
            }
        }
    }
}

// The exchange step and the merge in ONE launch: rank w's packed block [dists f64[nq,k] | nns u32[nq,k] |
// probe u16[nq,k]] is read straight out of PEER memory (rank_base[w] is that rank's symmetric-memory buffer mapped
// into this process: NVLink P2P loads through NVSwitch), staged per warp in shared memory with every lane
// issuing independent remote loads, then merged as in merge_ranks_kernel.
__global__ void merge_ranks_peer_kernel(const void* const* __restrict__ rank_base, int64_t nns_off, int64_t probe_off,
                                        int world, int nq, int k, double* __restrict__ dists,
                                        uint32_t* __restrict__ nns) {
    extern __shared__ double peer_sm[];
    const int warps = blockDim.x >> 5;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int q = blockIdx.x * warps + warp;
    if (q >= nq) return;
    const int per = world * k;
    double* sd = peer_sm + (size_t)warp * per;
    uint32_t* si = reinterpret_cast<uint32_t*>(peer_sm + (size_t)warps * per) + (size_t)warp * per;
    uint16_t* sp = reinterpret_cast<uint16_t*>(reinterpret_cast<uint32_t*>(peer_sm + (size_t)warps * per) +
                                               (size_t)warps * per) + (size_t)warp * per;
    for (int idx = lane; idx < per; idx += 32) {
        const int w = idx / k, j = idx - w * k;
        const char* base = static_cast<const char*>(rank_base[w]);
        sd[idx] = reinterpret_cast<const double*>(base)[(size_t)q * k + j];
        si[idx] = reinterpret_cast<const uint32_t*>(base + nns_off)[(size_t)q * k + j];
        sp[idx] = probe_off >= 0 ? reinterpret_cast<const uint16_t*>(base + probe_off)[(size_t)q * k + j] : (uint16_t)0;
    }
    __syncwarp();
    int head = 0;
    bool live = lane < world;
    double hv = live ? sd[lane * k] : INFINITY;
    int hp = live ? (int)sp[lane * k] : 0;
    for (int t = 0; t < k; ++t) {
        double kd = hv;
        int kp = hp;
        int ke = live ? lane : 0x7fffffff;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const double od = __shfl_xor_sync(0xffffffffu, kd, o);
            const int op = __shfl_xor_sync(0xffffffffu, kp, o);
            const int oe = __shfl_xor_sync(0xffffffffu, ke, o);
            if (oe != 0x7fffffff && (ke == 0x7fffffff || rank_key_less(od, op, oe, kd, kp, ke))) {
                kd = od;
                kp = op;
                ke = oe;
            }
        }
        if (ke == 0x7fffffff) {
            if (lane == 0) {
                dists[(size_t)q * k + t] = INFINITY;
                nns[(size_t)q * k + t] = 0u;
            }
            continue;
        }
        if (lane == ke) {
            dists[(size_t)q * k + t] = hv;
            nns[(size_t)q * k + t] = si[lane * k + head];
            ++head;
            if (head < k) {
                hv = sd[lane * k + head];
                hp = (int)sp[lane * k + head];
            } else {
                live = false;
            }
        }
    }
}

// All-gather by loads from peer memory: dst[w * bytes + i] = rank_base[w][i] (16-byte words).
__global__ void gather_peer_kernel(const void* const* __restrict__ rank_base, int64_t words_per_rank, int world,
                                   uint4* __restrict__ dst) {
    const int64_t total = words_per_rank * world;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int w = (int)(i / words_per_rank);
        const int64_t j = i - (int64_t)w * words_per_rank;
        dst[i] = static_cast<const uint4*>(rank_base[w])[j];
    }
}

}  // namespace lmi

extern "C" int lmi_merge_topk(const float* part_sim, const int32_t* part_row, const int32_t* slot_of_pair,
                              const int32_t* bucket_order, const int64_t* bucket_off, const uint32_t* ids,
                              const int32_t* bucket_lists, int nq, int B, int C, int lmax, int k_out, double* dists,
                              uint32_t* nns, uint16_t* probe_rank, void* stream) {
    if (!part_sim || !part_row || !slot_of_pair || !bucket_order || !bucket_off || !ids || !bucket_lists || !dists ||
        !nns)
        return LMI_E_BADARG;
    if (nq < 0 || B <= 0 || C <= 0 || lmax <= 0 || k_out <= 0) return LMI_E_BADARG;
    if (k_out > B * lmi::MK) return LMI_E_BADARG;
    if (lmax > 32 * lmi::MERGE_LPL || B >= 0xffff) return LMI_E_UNSUPPORTED;
    if (nq == 0) return 0;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const size_t per_warp = (size_t)2 * B * lmi::MK * sizeof(float);
    int warps = 4;
    while (warps > 1 && per_warp * warps > 160 * 1024) warps >>= 1;
    const size_t smem = per_warp * warps;
    if (smem > 200 * 1024) return LMI_E_UNSUPPORTED;
    if (smem > 48 * 1024) {
        cudaError_t e =
            cudaFuncSetAttribute(lmi::merge_topk_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return (int)e;
    }
    const int grid = (nq + warps - 1) / warps;
    lmi::merge_topk_kernel<<<grid, warps * 32, smem, st>>>(part_sim, part_row, slot_of_pair, bucket_order, bucket_off,
                                                           ids, bucket_lists, nq, B, C, lmax, k_out, dists, nns,
                                                           probe_rank);
    return static_cast<int>(cudaGetLastError());
}

extern "C" int lmi_merge_ranks(const double* dists_all, const uint32_t* nns_all, const uint16_t* probe_all,
                               int64_t rank_stride_d, int64_t rank_stride_i, int64_t rank_stride_p, int world, int nq,
                               int k, double* dists, uint32_t* nns, void* stream) {
    if (!dists_all || !nns_all || !dists || !nns || world <= 0 || nq < 0 || k <= 0) return LMI_E_BADARG;
    if (world > 32) return LMI_E_UNSUPPORTED;
    if (nq == 0) return 0;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int warps = 4;
    lmi::merge_ranks_kernel<<<(nq + warps - 1) / warps, warps * 32, 0, st>>>(
        dists_all, nns_all, probe_all, rank_stride_d, rank_stride_i, rank_stride_p, world, nq, k, dists, nns);
    return static_cast<int>(cudaGetLastError());
}

extern "C" int lmi_merge_ranks_peer(const void* const* rank_base, int64_t nns_offset_bytes, int64_t probe_offset_bytes,
                                    int world, int nq, int k, double* dists, uint32_t* nns, void* stream) {
    if (!rank_base || !dists || !nns || world <= 0 || nq < 0 || k <= 0 || nns_offset_bytes < 0) return LMI_E_BADARG;
    if (world > 32 || (nns_offset_bytes & 3) || (probe_offset_bytes >= 0 && (probe_offset_bytes & 1)))
        return LMI_E_UNSUPPORTED;
    if (nq == 0) return 0;
    const int warps = 4;
    const size_t smem = (size_t)warps * world * k * (sizeof(double) + sizeof(uint32_t) + sizeof(uint16_t));
    if (smem > 48 * 1024) return LMI_E_UNSUPPORTED;
    lmi::merge_ranks_peer_kernel<<<(nq + warps - 1) / warps, warps * 32, smem, static_cast<cudaStream_t>(stream)>>>(
        rank_base, nns_offset_bytes, probe_offset_bytes, world, nq, k, dists, nns);
    return static_cast<int>(cudaGetLastError());
}

extern "C" int lmi_gather_peer(const void* const* rank_base, int64_t bytes_per_rank, int world, void* dst,
                               void* stream) {
    if (!rank_base || !dst || world <= 0 || bytes_per_rank < 0) return LMI_E_BADARG;
    if ((bytes_per_rank & 15) || (reinterpret_cast<uintptr_t>(dst) & 15)) return LMI_E_BADARG;
    if (bytes_per_rank == 0) return 0;
    const int64_t words = bytes_per_rank / 16;
    int64_t blocks = (words * world + 255) / 256;
    if (blocks > 148 * 8) blocks = 148 * 8;
    lmi::gather_peer_kernel<<<(unsigned)blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(
        rank_base, words, world, static_cast<uint4*>(dst));
    return static_cast<int>(cudaGetLastError());
}
