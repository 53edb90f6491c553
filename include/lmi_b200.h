/*
 * lmi_b200.h -- C ABI of the B200-native LMI query path.
 *
 * Drop-in boundary for the learned-navigation search behind
 * LearnedIndex.search() of Coda-Research-Group/LearnedMetricIndex.  The reference
 * is pure Python; the arithmetic on its hot path lives in four library calls.
 * Each entry point below replaces one of them (file:line relative to the
 * reference checkout):
 *
 *   lmi_mlp_topn (ONE launch) or
 *   lmi_linear_tc (+ lmi_split_tf32) / lmi_linear / lmi_softmax_topn
 *                                  <- NeuralNetwork.predict_proba
 *                                     search/li/model.py:226-241 (Linear/ReLU
 *                                     stack :29-79, softmax :238, topk :239) and
 *                                     the 1-level slice bucket_order = classes[:, :B]
 *                                     search/li/LearnedIndex.py:209-214
 *   lmi_argmax                     <- NeuralNetwork.predict  search/li/model.py:213-224
 *   lmi_softmax_probs / lmi_navigate_tree
 *                                  <- the multi-level branch of _precompute_bucket_order
 *                                     search/li/LearnedIndex.py:216-252 with
 *                                     _visit_internal_nodes :254-301, _visit_buckets
 *                                     :303-325 and PriorityQueue (search/li/PriorityQueue.py:7-94)
 *   lmi_invert                     <- filter_path_idxs       search/li/utils.py:61-65,
 *                                     as called per bucket per round at
 *                                     search/li/LearnedIndex.py:353
 *   lmi_scan_topk                  <- groupby + .loc gather + faiss.knn(...,
 *                                     METRIC_INNER_PRODUCT)
 *                                     search/li/LearnedIndex.py:350-365
 *   lmi_merge_topk                 <- distances = 1 - similarity, local->label id
 *                                     mapping and the cross-round stable-argsort
 *                                     merge  search/li/LearnedIndex.py:368-371, :125-146
 *   lmi_scatter_rows / lmi_cast_f16<- the bucket layout the reference re-derives
 *                                     every round from data_prediction + the
 *                                     DataFrame index (:101-104, :350, :357),
 *                                     materialised once as a bucket-major fp16 matrix
 *
 * Conventions
 *   - extern "C", plain pointers and sizes; no C++/torch types.
 *   - every pointer is a DEVICE pointer owned by the caller unless the name ends
 *     in _host; nothing is allocated or freed inside, nothing is retained.  Scratch is passed in by the
 *     caller too -- or, for callers that cannot allocate any, see lmi_invert_static.
 *   - `stream` is a cudaStream_t passed as void*; calls only enqueue work.
 *   - return value: 0 on success, a positive cudaError_t, or a negative LMI_E_*.
 *     lmi_error_string() renders either.  No exceptions cross the boundary.
 *   - thread-safety: stateless; concurrent calls on different streams are fine (one exception, documented at
 *     its declaration: lmi_invert_static, whose scratch is one array per GPU shared by all of its calls).
 *   - there is no CPU fallback: every entry point launches sm_100a kernels.
 */
#ifndef LMI_B200_H_
#define LMI_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define LMI_B200_ABI_VERSION 7

#define LMI_E_BADARG (-1)      /* null pointer / non-positive size / unsupported shape */
#define LMI_E_UNSUPPORTED (-2) /* shape outside what the sm_100a kernels implement */
#define LMI_E_DRIVER (-3)      /* CUDA driver entry point (tensor map encode) unavailable */
#define LMI_E_SCRATCH_LIMIT (-4) /* lmi_invert_static: the batch needs more than the built-in scratch array */

#define LMI_PER_BUCKET_K 10 /* `k` is not forwarded to _search_single_bucket (LearnedIndex.py:111-117,:334) */
#define LMI_ITEM_WORDS 8    /* int32 words per scan work item, see lmi_invert */
#define LMI_LISTS_PER_ITEM 2 /* partial top-10 lists a scan work item writes per query (one per column half) */

int lmi_version(void);
const char *lmi_error_string(int code);

/* ---- model.py ---------------------------------------------------------------- */

/* y[nq,out] = act(x[nq,in] . W[out,in]^T + b), fp32 FMA.  x is fp32, or fp16 when
 * x_is_f16 != 0 (widened exactly, as data_X_to_torch does, model.py:102-105). */
int lmi_linear(const void *x, int x_is_f16, const float *W, const float *b, float *y,
               int nq, int in_dim, int out_dim, int relu, void *stream);

/* Tensor-core form of lmi_linear (tcgen05 kind::tf32, TMA-fed) on SPLIT operands: every fp32
 * value v is passed as hi + mid + lo with each part exact in tf32 (lmi_split_tf32), all part
 * products of weight >= 2^-22 are accumulated, K-blocks of 32 in the tensor core and across
 * K-blocks in fp32 registers -- fp32-quality logits, as parity with model.py:232 needs.
 * x_mid/x_lo (w_mid/w_lo) are both NULL when the operand is exact in one part (fp16 data).
 * Output: fp32 y and/or the three parts of y (the next layer's x operand). in_dim % 4 == 0. */
int lmi_split_tf32(const void *x, int x_is_f16, int64_t count, float *hi, float *mid, float *lo,
                   void *stream);
int lmi_linear_tc(const float *x_hi, const float *x_mid, const float *x_lo,
                  const float *w_hi, const float *w_mid, const float *w_lo, const float *bias,
                  int nq, int in_dim, int out_dim, int relu,
                  float *y, float *y_hi, float *y_mid, float *y_lo, void *stream);

/* Grouped form -- every internal model of one tree level in ONE launch (blockIdx.z = model g):
 *   x parts  [nq, ldx]: model g reads columns [g*in_dim, (g+1)*in_dim)  (the previous grouped layer's
 *                       output; for a first layer stack the W of all models and call lmi_linear_tc)
 *   w parts  [groups, out_dim, in_dim] contiguous, bias [groups, out_dim]
 *   y / y parts: row pitch ldy, model g writes columns [y_col0 + g*out_dim, y_col0 + (g+1)*out_dim)
 * so the last layer can write its logits straight into the node table of lmi_navigate_tree. */
int lmi_linear_tc_grouped(const float *x_hi, const float *x_mid, const float *x_lo, int ldx,
                          const float *w_hi, const float *w_mid, const float *w_lo, const float *bias,
                          int nq, int in_dim, int out_dim, int groups, int relu,
                          float *y, float *y_hi, float *y_mid, float *y_lo, int ldy, int y_col0,
                          void *stream);

/* softmax over the C logits of each row, then the B most probable classes in
 * descending probability (ties: lower class index first).
 * bucket_order int32[nq,B] (== the reference's bucket_order[:, :, 0]), probs f32[nq,B]. */
int lmi_softmax_topn(const float *logits, int nq, int C, int B,
                     int32_t *bucket_order, float *probs, void *stream);

/* NeuralNetwork.predict_proba (model.py:226-241) of a 2-Linear-layer model (MLP, MLP-2, MLP-3, MLP-4,
 * model.py:29-79) in ONE launch, with the slice bucket_order = classes[:, :B] (LearnedIndex.py:209-214):
 * Linear + bias + ReLU + Linear + bias + softmax + top-B, the hidden activations never leaving the SM.
 * Same split-tf32 arithmetic as lmi_linear_tc (fp32-quality logits); x is the fp16 query matrix (exact
 * in one tf32 part), w1/w2 parts from lmi_split_tf32 of the [hidden,in_dim] / [C,hidden] weights.
 *   lmi_mlp_topn_supported  1 when the shape is implemented: fp16 x, in_dim % 8 == 0, hidden % 64 == 0,
 *                           hidden <= 1024, C <= 128, B <= C; otherwise use the four-call path above
 *   workspace   lmi_mlp_topn_workspace_bytes(nq, hidden, C) bytes, 16-byte aligned, ZEROED ONCE by the
 *               caller before the first use (the kernel leaves its tile counters at zero again)
 *   B > 0:  bucket_order int32[nq,B], probs f32[nq,B] as lmi_softmax_topn;  logits f32[nq,C] optional
 *   B == 0: only logits are written (predict / multi-level tables) */
int lmi_mlp_topn_supported(int x_is_f16, int in_dim, int hidden, int C, int B);
int64_t lmi_mlp_topn_workspace_bytes(int nq, int hidden, int C);
int lmi_mlp_topn(const void *x_f16, int nq, int in_dim,
                 const float *w1_hi, const float *w1_mid, const float *w1_lo, const float *b1, int hidden,
                 const float *w2_hi, const float *w2_mid, const float *w2_lo, const float *b2, int C, int B,
                 void *workspace, int32_t *bucket_order, float *probs, float *logits, void *stream);

/* argmax over the C logits of each row (first maximum), int64 out. */
int lmi_argmax(const float *logits, int64_t n, int C, int64_t *out, void *stream);

/* ---- multi-level navigation (n_levels > 1) --------------------------------------
 * Tree nodes are numbered breadth first: depth-1 nodes 0..C1-1 (node id == class of the root
 * model), then every depth-2 node parent-major, and so on; the children of a node are
 * contiguous.  node_prob[q, v] is the probability the model of v's PARENT gives v's class for
 * query q (raw, not multiplied by the parent's own probability -- LearnedIndex.py:295-299). */

/* probs[q*out_stride + out_col0 + c] = softmax(logits[q, :])[c]  (model.py:238), C columns. */
int lmi_softmax_probs(const float *logits, int nq, int C, float *probs, int64_t out_stride,
                      int64_t out_col0, void *stream);

/* In-place softmax of n_seg column segments of table[nq, ld]: segment s covers columns
 * [seg_start[s], seg_start[s] + seg_len[s]); seg_len[s] <= 0 is skipped.  One segment per
 * model: the root's [0, C1) and, for every node with a model, its children's columns. */
int lmi_softmax_segments(float *table, int nq, int64_t ld, const int32_t *seg_start,
                         const int32_t *seg_len, int n_seg, void *stream);

/* Per query: best-first search over the tree.  The frontier starts as the n_roots depth-1
 * nodes; the most probable frontier node is popped (ties: lowest node id; the reference's
 * unstable argsort leaves ties undefined, PriorityQueue.py:80); a popped node with
 * n_children[v] > 0 (it has an internal model) opens children first_child[v] .. +n_children[v];
 * a popped node with node_bucket[v] >= 0 (its path is in bucket_paths) appends that flat
 * bucket id to the query's order; a node that is neither is dropped, as in the reference.
 * Stops after B buckets.  bucket_order int32[nq,B]; -1 where the tree ran out (the host
 * rejects budgets beyond the reachable buckets before launching). */
int lmi_navigate_tree(const float *node_prob, int nq, int n_nodes, const int32_t *first_child,
                      const int32_t *n_children, const int32_t *node_bucket, int n_roots, int B,
                      int32_t *bucket_order, void *stream);

/* ---- bucket layout ------------------------------------------------------------ */

/* dst[dest_row[i], 0:d] = (fp16) src[i, :] for i < m, and dst[dest_row[i], d:dst_ld] = 0; src is
 * fp32 or fp16 [m,d], dst has a row pitch of dst_ld >= d elements (the scan kernel wants a
 * multiple of 64: zero columns leave every inner product unchanged).  dest_row == NULL is the
 * identity (cast + pad of a query matrix); a row with dest_row[i] < 0 is dropped (an object without
 * a bucket).  inexact (optional, device int64, caller-zeroed): incremented by the number of fp32
 * values the conversion to fp16 changed -- zero for data that is fp16 on disk and was widened, as
 * SISAP's `emb` is after the reference's float32 casts (model.py:104, faiss.knn); non-zero means
 * the scan runs on ROUNDED vectors (relative error 2^-11 per component), which the caller must
 * surface (learnedmetricindex_b200 warns, or refuses with allow_fp16_rounding=False). */
int lmi_scatter_rows(const void *src, int src_is_f16, const int64_t *dest_row, int64_t m, int d,
                     void *dst_f16, int dst_ld, int64_t *inexact, void *stream);

/* dst[i] = (fp16) src[i], count elements; inexact as above. */
int lmi_cast_f16(const float *src, void *dst_f16, int64_t count, int64_t *inexact, void *stream);

/* ---- query-to-bucket inversion ------------------------------------------------ */

/* Upper bound on the number of scan work items lmi_invert can emit (rmax = the most ranges any bucket
 * is cut into, i.e. ceil(largest bucket / the SHORTER of range_len and tail_range_len)). */
int64_t lmi_max_items(int nq, int B, int C, int rmax);
/* int32 words of scratch lmi_invert needs (per-chunk histograms). */
int64_t lmi_invert_workspace_ints(int nq, int B, int C);

/* One allocation for everything a query batch needs between navigation and the final merge: byte offsets
 * (256-byte aligned) of the arrays below inside a workspace of the returned size; offsets_host (HOST
 * int64[LMI_WS_ARRAYS + 1], may be NULL) receives them, entry LMI_WS_ARRAYS the total.  Shapes depend only
 * on (nq, B, C, rmax, d_pad), so a caller keeps one workspace per batch shape and a search allocates
 * nothing.  d_pad = row pitch of the fp16 query copy (0: queries are fp16 with the layout's pitch already). */
#define LMI_WS_COUNTS 0         /* int32[C]               lmi_invert counts        */
#define LMI_WS_OFFSETS 1        /* int32[C+1]             lmi_invert offsets       */
#define LMI_WS_BUCKET_LISTS 2   /* int32[C]               lmi_invert bucket_lists  */
#define LMI_WS_PAIR_LIST 3      /* int32[nq*B]            lmi_invert pair_list     */
#define LMI_WS_SLOT_OF_PAIR 4   /* int32[nq*B]            lmi_invert slot_of_pair  */
#define LMI_WS_ITEMS 5          /* int32[max_items*8]     lmi_invert items         */
#define LMI_WS_N_ITEMS 6        /* int32[2]               lmi_invert n_items       */
#define LMI_WS_INVERT_SCRATCH 7 /* int32[lmi_invert_workspace_ints]                */
#define LMI_WS_PART_SIM 8       /* f32[nq*B, 2*rmax, 10]  lmi_scan_topk part_sim   */
#define LMI_WS_PART_ROW 9       /* int32[nq*B, 2*rmax, 10] lmi_scan_topk part_row  */
#define LMI_WS_Q16 10           /* fp16[nq, d_pad]        cast / padded queries    */
#define LMI_WS_GATE 11          /* uint32[2*rmax, nq*B]   lmi_scan_topk gate       */
#define LMI_WS_ARRAYS 12
int64_t lmi_workspace_layout(int nq, int B, int C, int rmax, int d_pad, int64_t *offsets_host);
int64_t lmi_workspace_bytes(int nq, int B, int C, int rmax);

/* Groups the nq*B (query, probe-rank) pairs by probed bucket -- ascending flat pair
 * index p = q*B + r inside a bucket, i.e. the order filter_path_idxs yields -- and
 * cuts the scan into work items.
 *   bucket_order int32[nq,B]   in   entries outside [0,C) are ignored, and so are probes of a bucket
 *                                   that holds no rows in THIS layout (hi == lo in bucket_off: an empty
 *                                   bucket, or one that lives on another GPU): slot -1, no item, the
 *                                   merge pads them as the reference pads an empty bucket
 *   bucket_off   int64[C+1]    in   row range of bucket b in the bucket-major matrix
 *   range_len                  in   upper bound on the rows of a work item (multiple of 64); a
 *                                   bucket of s rows is cut into ceil(s / range_len) ranges of
 *                                   equal length, rounded up to a multiple of 64
 *   tail_range_len, tail_permille   two-level cut: once (1000 - tail_permille)/1000 of the launch's work
 *                                   (rows x query tiles, buckets in id order = the order the scan draws
 *                                   items in) has been listed, buckets are cut with tail_range_len
 *                                   (multiple of 64, < range_len) instead, so the persistent CTAs run dry
 *                                   within a short item of each other; 0 / 0 = one level
 *   counts       int32[C]      out  m_b = pairs probing bucket b (0 for a bucket without rows)
 *   offsets      int32[C+1]    out  exclusive scan of counts
 *   bucket_lists int32[C]      out  partial lists the scan writes per (query, bucket b) pair
 *                                   (= ranges of b x LMI_LISTS_PER_ITEM; 0 for an unprobed bucket)
 *   pair_list    int32[nq*B]   out  flat pair index per sorted slot
 *   slot_of_pair int32[nq*B]   out  inverse permutation (-1 for ignored pairs)
 *   items        int32[max_items*8] out  {row0, nrows, slot0, mvalid, first list | lists of the bucket << 16,
 *                                   bucket, query tile index within (bucket, range), query tiles of the bucket}
 *   n_items      int32[2]      out  [0] number of items; [1] = 0, the scan's item counter
 *   workspace    int32[lmi_invert_workspace_ints(nq,B,C)]  scratch
 */
int lmi_invert(const int32_t *bucket_order, int nq, int B, int C, const int64_t *bucket_off,
               int range_len, int tail_range_len, int tail_permille,
               int32_t *counts, int32_t *offsets, int32_t *bucket_lists,
               int32_t *pair_list, int32_t *slot_of_pair, int32_t *items, int32_t *n_items,
               int64_t max_items, int32_t *workspace, void *stream);

/* lmi_invert for callers that have no device memory to give it as scratch.  (No counterpart in the reference:
 * it is the same replacement of filter_path_idxs as lmi_invert; only where the scratch comes from differs.)
 *
 * Who needs it: code that runs where allocating device memory is off the table.  A plugin living inside a
 * host application (a database operator, a serving backend) is typically handed fixed input/output buffers
 * and a stream, and must not touch the allocator: the host has already claimed the GPU's memory for its own
 * pool, so a cudaMalloc from the plugin either fails or breaks the footprint the host budgeted.  A
 * latency-critical serving thread cannot take the device-wide synchronisation that cudaMalloc / cudaFree
 * imply.  And a caller written in another language over FFI may simply have no device allocator to call.
 * Such callers can still own the OUTPUT arrays below (they are the result, sized by the host), but the
 * per-chunk histograms of lmi_invert are pure scratch -- so this variant keeps them in an array that is part
 * of the library image itself: LMI_INVERT_STATIC_MAX_INTS int32 = 4 MiB on every GPU that loads the library,
 * held until the process exits.  The call performs no cudaMalloc / cudaFree and calls nothing that allocates.
 *
 * Arguments and outputs: exactly lmi_invert's, minus `workspace`; the outputs are bit-identical to lmi_invert's.
 *
 * Limit: lmi_invert_workspace_ints(nq, B, C) <= LMI_INVERT_STATIC_MAX_INTS, checked before anything is
 * launched; beyond it the call returns LMI_E_SCRATCH_LIMIT and writes nothing (use lmi_invert with your own
 * scratch).  The headline batch (10 000 queries x 4 probes, 120 buckets) needs 18 720 ints; 100 000 queries x
 * 120 probes over 120 buckets needs 491 520; the limit is first hit at 4096 chunks x more than 256 buckets.
 *
 * First call on each device: the library looks up where its array lives on the device that is current on the
 * calling thread (cudaGetSymbolAddress, result cached per device, return code checked).  That look-up can
 * make the runtime load the module, so make the first call on a device with that device current, on the
 * stream's own device, and BEFORE you begin any stream capture; later calls only enqueue the three kernels.
 *
 * Sharing: there is ONE array per GPU and every call on that GPU uses it.  Nothing has to be cleared between
 * calls (the first pass zeroes the cells it is about to count into, so the previous call's leftovers are never
 * read), but two calls whose kernels can be in flight at the same time -- different streams, or different
 * host threads on different streams -- overwrite each other's histograms and produce garbage.  Keep all
 * calls for one GPU on one stream, or order them with events.  lmi_invert has no such restriction. */
#define LMI_INVERT_STATIC_MAX_INTS (1 << 20)
int lmi_invert_static(const int32_t *bucket_order, int nq, int B, int C, const int64_t *bucket_off,
                      int range_len, int tail_range_len, int tail_permille,
                      int32_t *counts, int32_t *offsets, int32_t *bucket_lists,
                      int32_t *pair_list, int32_t *slot_of_pair, int32_t *items, int32_t *n_items,
                      int64_t max_items, void *stream);

/* ---- candidate scan + fused per-range top-10 ---------------------------------- */

/* For every work item: inner products of its <=128 queries with its rows of X
 * (tcgen05 fp16 MMA, fp32 accumulation in TMEM, X streamed by TMA), keeping the
 * 10 largest per query.  The full similarity matrix is never written.
 *   X         fp16[n,d] bucket-major, 16-byte aligned, d % 64 == 0, 64 <= d <= 768
 *   Q         fp16[nq,d]
 *   lmax      lists allocated per (query, probe) pair: >= max_b bucket_lists[b]
 *   part_sim  f32 [nq*B, lmax, 10] out  descending; -inf where fewer than 10 rows
 *   part_row  i32 [nq*B, lmax, 10] out  row index into X; -1 where empty
 * An item writes lists {first list, first list + 1} of each of its queries.
 *   gate      uint32[lmax, nq*B] scratch, optional (NULL: off).  The row ranges of one bucket are scanned by
 *             different CTAs at the same time, each keeping its own top-10 lists.  Through this array every list
 *             publishes, per (query, bucket) pair, its j-th best similarity so far; when m lists hold j entries
 *             above a value v and m*j >= 10, ten rows of the bucket beat v, so every list may skip scores below
 *             the m-th largest published value ((j, m) = (2, 5) for a bucket of >= 5 lists, else (5, 2)).  The
 *             partial lists then hold fewer entries below the final 10th best (and vary run to run); the merged
 *             result is unchanged.  The call resets the array itself.
 * impl (low 8 bits; every value gives bit-identical lists):
 *   0 = tcgen05/TMA kernel, one CTA per work item, 64-row score tiles, two accumulators (product path)
 *   1 = plain CUDA-core kernel, any d: the verification path of the tests, and what a search with d > 768 runs
 *       on (correct, far from the roofline; learnedmetricindex_b200 warns when it takes it)
 *   2 = the impl-0 kernel over thread-block clusters of two CTAs: two query tiles of one (bucket, range)
 *       receive the bucket rows by ONE multicast TMA load (measured: not faster)
 *   3 = 128-row score tiles (N = 128 MMAs, half the A-operand reads per flop), one accumulator
 *   4 = 128-row score tiles with the queries split between tensor memory (K-blocks 0-7) and shared memory
 *       (K-blocks 8-11), which leaves room for two accumulators (measured: not faster than 0)
 * impl >> 8: tuning flags (bit 0: L2 eviction hints by sibling order); results do not depend on them.
 */
int lmi_scan_topk(const void *X, int64_t n, int d, const void *Q, int nq, int B,
                  const int32_t *pair_list, const int32_t *items, const int32_t *n_items,
                  int64_t max_items, int lmax, float *part_sim, int32_t *part_row,
                  int impl, uint32_t *gate, void *stream);

/* ---- merge ------------------------------------------------------------------- */

/* Per query: reduce each probe's partial lists (bucket_lists[b] of them, ties between
 * equal similarities go to the lower row) to that bucket's top-10, turn
 * similarities into distances 1 - sim (f32), map rows to ids, pad as the reference
 * does (undersized bucket: last id of the bucket, 1 + FLT_MAX; empty or unknown
 * bucket: id 0, +inf), then keep the k_out smallest over all probes -- ties keep the
 * earlier probe first (stable argsort over [previous | new], LearnedIndex.py:131-134).
 *   ids        uint32[n]   label of each bucket-major row (the DataFrame index)
 *   dists      f64[nq,k_out] out, nns uint32[nq,k_out] out
 *   probe_rank uint16[nq,k_out] out, optional (NULL: not written): the probe rank r of each returned entry
 *              (0xffff for a pad) -- what the cross-GPU merge needs to order equal distances exactly as
 *              the single-GPU search does
 *   k_out <= 10*B.
 */
int lmi_merge_topk(const float *part_sim, const int32_t *part_row, const int32_t *slot_of_pair,
                   const int32_t *bucket_order, const int64_t *bucket_off, const uint32_t *ids,
                   const int32_t *bucket_lists, int nq, int B, int C, int lmax, int k_out,
                   double *dists, uint32_t *nns, uint16_t *probe_rank, void *stream);

/* Cross-GPU merge after the all-gather of the per-rank results: rank w's ascending
 * lists are dists_all + w*rank_stride_d (f64[nq,k]), nns_all + w*rank_stride_i
 * (uint32[nq,k]) and probe_all + w*rank_stride_p (uint16[nq,k], lmi_merge_topk's probe_rank;
 * NULL: absent) -- strides in elements, so one packed gather buffer works -- and the
 * output is the k smallest per query.  Equal distances are ordered by (probe rank, GPU rank):
 * with whole buckets per GPU, or with every bucket cut into row slices in GPU order, that is
 * the single-GPU search's stable order (LearnedIndex.py:131-134), so the sharded result equals
 * the unsharded one bit for bit.  Without probe ranks ties keep the lower GPU rank. world <= 32. */
int lmi_merge_ranks(const double *dists_all, const uint32_t *nns_all, const uint16_t *probe_all,
                    int64_t rank_stride_d, int64_t rank_stride_i, int64_t rank_stride_p,
                    int world, int nq, int k, double *dists, uint32_t *nns, void *stream);

/* The exchange step fused with the merge: instead of an all-gather followed by lmi_merge_ranks,
 * every rank reads its peers' result blocks straight out of peer memory (NVLink P2P through
 * NVSwitch) inside the merge kernel.  rank_base: DEVICE array of `world` pointers, entry w = the
 * address IN THIS PROCESS of rank w's packed block [dists f64[nq,k] | nns u32[nq,k] | probe u16[nq,k]]
 * (symmetric memory / CUDA IPC mapping); nns_offset_bytes / probe_offset_bytes = offsets of those
 * parts in a block (probe_offset_bytes < 0: no probe ranks, ties keep the lower GPU rank).
 * The caller orders the launch after a cross-rank barrier (all blocks written) and must not
 * overwrite its own block until every peer has passed the next barrier (double buffering). */
int lmi_merge_ranks_peer(const void *const *rank_base, int64_t nns_offset_bytes,
                         int64_t probe_offset_bytes, int world, int nq, int k, double *dists,
                         uint32_t *nns, void *stream);

/* All-gather by loads from peer memory (no collective library): dst[w*bytes_per_rank + i] =
 * rank_base[w][i] for every rank w; rank_base as in lmi_merge_ranks_peer, bytes_per_rank a multiple
 * of 16, dst 16-byte aligned.  Used for the query-sharded navigation's bucket order and for the
 * query upload (each rank copies 1/world of the host batch over PCIe, the rest arrives over NVLink). */
int lmi_gather_peer(const void *const *rank_base, int64_t bytes_per_rank, int world, void *dst,
                    void *stream);

#ifdef __cplusplus
}
#endif
#endif /* LMI_B200_H_ */
