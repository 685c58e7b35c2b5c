/* b200ann — C ABI of the B200-native hot path of Cookiecoolkid/hnsw-nsg.
 *
 * Everything behind this header is hand-written sm_100a CUDA (hnsw-nsg_b200/csrc). There is no CPU
 * fallback: every entry point fails with B200ANN_ERR_CUDA when no CUDA device is usable.
 *
 * Conventions
 *   - every function returns 0 on success or a negative B200ANN_ERR_* code; the message is kept per
 *     handle and read with b200ann_last_error() (the C++ shims re-throw it as std::runtime_error, the
 *     reference's error convention: hnswlib/hnswlib/hnswalg.h:1178, :720).
 *   - pointers marked [h|d] may be host or device memory (resolved with cudaPointerGetAttributes);
 *     results are written to wherever the output pointer lives.
 *   - vectors are fp32 row-major [n, stride] (stride in floats, >= dim); ids are u32; an index copies
 *     the rows to HBM with stride rounded up to 4 floats (16 B) and zero padding.
 *   - adjacency is the flat layout u32 [n, gstride], row i = out-neighbours in list order, padded
 *     with B200ANN_INVALID.
 *   - one handle per GPU and per host thread; calls on different handles run concurrently.
 */
#ifndef B200ANN_H
#define B200ANN_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif
#pragma GCC visibility push(default)

#define B200ANN_INVALID 0xFFFFFFFFu

enum {
  B200ANN_L2 = 0,   /* hnswlib::L2Space */
  B200ANN_IP = 1,   /* hnswlib::InnerProductSpace (1 - <a,b>) */
  B200ANN_L2_U8 = 2 /* hnswlib::L2SpaceI: unsigned-byte rows, int squared L2 (b200ann_index_create_u8 only) */
};

enum {
  B200ANN_OK = 0,
  B200ANN_ERR_ARG = -1,   /* bad argument (message says which) */
  B200ANN_ERR_CUDA = -2,  /* CUDA runtime error / no device */
  B200ANN_ERR_STATE = -3, /* index has no graph of the requested kind */
  B200ANN_ERR_FEWER = -4, /* fewer than k results for some query (bindings.cpp:646-648) */
  B200ANN_ERR_CAPACITY = -5 /* filtered search: candidate set overflow on some query (see b200ann_hnsw_search_filtered) */
};

typedef struct b200ann_handle *b200ann_handle_t;
typedef struct b200ann_index *b200ann_index_t;

/* ---- handle: device + stream + workspace + last error ------------------------------------ */
int b200ann_create(int device, b200ann_handle_t *out);
int b200ann_destroy(b200ann_handle_t h);
const char *b200ann_last_error(b200ann_handle_t h);
/* cudaStream_t the handle launches on (for CUDA-event timing by the caller) */
void *b200ann_stream(b200ann_handle_t h);
/* number of kernels this handle has launched since creation */
uint64_t b200ann_launch_count(b200ann_handle_t h);
/* milliseconds the kernels of the last search/build call took on the device (CUDA events) */
float b200ann_last_kernel_ms(b200ann_handle_t h);

/* ---- index: vectors (+ graphs) resident in HBM -------------------------------------------- */
/* replaces: the data ownership of hnswlib::HierarchicalNSW (data_level0_memory_, hnswalg.h:130) and
 * the borrowed `data_` of efanna2e::IndexNSG (CNNS/src/nsg/index_nsg.cpp:469,509) */
int b200ann_index_create(b200ann_handle_t h, const float *data /*[h|d]*/, size_t n, size_t dim, size_t stride,
                         int metric, b200ann_index_t *out);
/* replaces the data ownership of hnswlib::HierarchicalNSW<int> over hnswlib::L2SpaceI (space_l2.h:255-323: elements
 * are dim unsigned bytes, the distance is the int sum of (a-b)^2, L2SqrI4x / L2SqrI): data = n rows of dim bytes,
 * stride BYTES apart; rows live in HBM padded to 16 bytes - a quarter of the float index's bytes per hop of a
 * bandwidth-bound walk. dim <= 8192. Such an index is taken by the HNSW entry points only (b200ann_hnsw_build,
 * b200ann_hnsw_insert_batch with byte rows and a byte stride, b200ann_index_set_hnsw / get_hnsw,
 * b200ann_hnsw_search, _filtered, _epsilon); there `q` points at unsigned-byte query rows, `qstride` is in BYTES,
 * and the `dists` array holds int32 values (the reference's dist_t = int; unfilled slots INT32_MAX), bit-exact -
 * integer sums do not depend on the summation order. epsilon of the range search is compared as an integer.
 * Every other entry point returns B200ANN_ERR_ARG for it. */
int b200ann_index_create_u8(b200ann_handle_t h, const uint8_t *data /*[h|d]*/, size_t n, size_t dim, size_t stride,
                            b200ann_index_t *out);
int b200ann_index_destroy(b200ann_index_t ix);
size_t b200ann_index_size(b200ann_index_t ix);
size_t b200ann_index_dim(b200ann_index_t ix);

/* NSG / kNN flat graph: final_graph_ + ep_ (+ width) of efanna2e::IndexNSG (index_nsg.h:62-84);
 * what IndexNSG::Load (index_nsg.cpp:51-68) or OptimizeGraph (:659-682) produce on the CPU */
int b200ann_index_set_graph(b200ann_index_t ix, const uint32_t *graph /*[h|d]*/, size_t gstride, uint32_t ep);
/* copy the graph back: graph [n, gstride_out] */
int b200ann_index_get_graph(b200ann_index_t ix, uint32_t *graph /*[h|d]*/, size_t gstride_out, uint32_t *ep,
                            uint32_t *width);

/* HNSW graph in the flat layout: what HierarchicalNSW::loadIndex (hnswalg.h:716-822)
 * holds in data_level0_memory_/linkLists_/element_levels_. upper_links [upper_rows, maxM]. */
int b200ann_index_set_hnsw(b200ann_index_t ix, const uint32_t *links0, size_t maxM0, const int32_t *levels,
                           const uint32_t *upper_ptr, const uint32_t *upper_links, size_t upper_rows,
                           size_t maxM, const uint64_t *labels /* NULL = identity */, uint32_t ep,
                           int maxlevel);

/* replaces the hnswlib construction loop: HierarchicalNSW(space, n, M, ef_construction, random_seed) followed by
 * addPoint(row i, label i) for i = 0..n-1 (hnswlib/hnswlib/hnswalg.h:89-144 ctor, :954-992 / :1153-1267 addPoint,
 * :225-305 searchBaseLayer, :443-483 getNeighborsByHeuristic2, :506-630 mutuallyConnectNewElement; driven like
 * hnswlib/tests/cpp/sift_1m.cpp:321-336). Levels are drawn exactly like the reference (:206-210, same generator
 * and seed); rows are inserted in batches (DESIGN.md "HNSW build"): ratio = rows already inserted per new row of a
 * batch (0 = 32; at most 8 for the first min(2048, n/16) rows), bmax = largest batch (0 = 16384). labels [n] (NULL = row number) are what addPoint's `label` is.
 * On success the index holds the HNSW graph (b200ann_hnsw_search, b200ann_index_get_hnsw, hybrid search).
 * info (optional, 8 x u64): {beam distance evaluations, expanded nodes, heuristic distance evaluations, batches,
 * upper rows, maxlevel, entry point, largest reverse-link group}. */
int b200ann_hnsw_build(b200ann_handle_t h, b200ann_index_t ix, size_t M, size_t ef_construction, uint32_t random_seed,
                       size_t ratio, size_t bmax, const uint64_t *labels /*[h|d] or NULL*/, uint64_t *info /*[h]*/);
/* replaces nb addPoint calls on a LIVE hnswlib index (hnswalg.h:954-992 / :1153-1267): appends nb rows (data [nb,
 * stride], labels [nb] or NULL = row numbers) to the index and links them into the HNSW graph it already holds, in
 * place - the device arrays grow with doubling capacity, the rows already linked are not touched except through the
 * reverse links the new points add, and the level generator continues where the build / the previous append left it
 * (a graph installed with b200ann_index_set_hnsw starts a fresh generator with seed 100 like the reference's load
 * ctor). The new rows are inserted with the batch schedule of b200ann_hnsw_build (one batch for nb <= n/32).
 * ef_construction: 0 = the value of the build (200 for an installed graph). Any NSG graph of the index is dropped.
 * b200ann_index_size grows by nb. info (optional, 8 x u64): as b200ann_hnsw_build, counters of this call. */
int b200ann_hnsw_insert_batch(b200ann_handle_t h, b200ann_index_t ix, const float *data /*[h|d]*/, size_t stride,
                              const uint64_t *labels /*[h|d] or NULL*/, size_t nb, size_t ef_construction,
                              uint64_t *info /*[h]*/);
/* replaces HierarchicalNSW::updatePoint(dataPoint, internalId, 1.0) + repairConnectionsForUpdate (hnswalg.h:993-1140),
 * i.e. what addPoint does for a label that already exists (:1160-1174), for nb elements at once and IN PLACE: ids [nb]
 * = distinct internal ids (rows of the index), data [nb, stride] = their new vectors (bytes for a u8 index). The
 * vectors are replaced; every one-hop neighbour of an updated element gets a new list from {element, one-hop,
 * two-hop} through the heuristic (phase A, on the graph as it was before the call; a list claimed by several
 * updated elements is computed for the last of them); then the elements are re-linked like an insertion with
 * themselves filtered out of their candidates, sources already present in a neighbour's list skipped (phase B).
 * Levels, entry point and labels do not change; no rebuild, cost proportional to nb. Any NSG graph of the index is
 * dropped. The batch form is restated by oracle.c orc_hnsw_update variant 1 and equals it link for link; its
 * variant 0 is the reference's one-at-a-time form, pinned to the compiled reference. One pass moves at most n/32
 * elements (the build's batch ratio: elements of a pass do not see each other's new positions); a larger call is a
 * sequence of passes, each on the graph the previous one left, and gives the same result as that sequence of calls.
 * info (optional, 8 x u64): {distance evaluations, expanded nodes, heuristic evaluations,
 * chunks, lists rewritten by phase A, maxlevel, entry point, largest reverse-link group}. */
int b200ann_hnsw_update_batch(b200ann_handle_t h, b200ann_index_t ix, const uint32_t *ids /*[h|d]*/,
                              const float *data /*[h|d]*/, size_t stride, size_t nb, uint64_t *info /*[h]*/);
/* new labels for single elements of an index that holds an HNSW graph: what addPoint(.., replace_deleted = true) does
 * to the slot it reuses (hnswalg.h:982-987 setExternalLabel). ids [nb] internal ids, labels [nb]. */
int b200ann_index_set_labels(b200ann_index_t ix, const uint32_t *ids /*[h|d]*/, const uint64_t *labels /*[h|d]*/, size_t nb);
/* sizes of the index's HNSW graph: {maxM0, maxM, upper_rows, maxlevel, entry point} */
int b200ann_index_hnsw_shape(b200ann_index_t ix, uint64_t *out5 /*[h]*/);
/* copy the HNSW graph back in the layout b200ann_index_set_hnsw takes (what HierarchicalNSW::saveIndex
 * serialises, hnswalg.h:685-713): links0 [n,maxM0], levels [n], upper_ptr [n], upper_links [upper_rows,maxM],
 * labels [n]; any pointer may be NULL. */
int b200ann_index_get_hnsw(b200ann_index_t ix, uint32_t *links0 /*[h|d]*/, int32_t *levels, uint32_t *upper_ptr,
                           uint32_t *upper_links, uint64_t *labels);

/* ---- exact brute-force kNN ---------------------------------------------------------------- */
/* replaces hnswlib::BruteforceSearch::searchKnn (bruteforce.h:106-135) and
 * IndexGraph::generate_control_set (efanna_graph/src/index_graph.cpp:138-153), batched.
 * ids/dists [nq,k] ascending by (distance,id); ties -> smaller id. Unfilled slots (k > n):
 * B200ANN_INVALID / +inf. */
int b200ann_bf_knn(b200ann_handle_t h, b200ann_index_t ix, const float *q /*[h|d]*/, size_t nq, size_t qstride,
                   size_t k, uint32_t *ids /*[h|d]*/, float *dists /*[h|d] or NULL*/);
/* No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it). b200ann_bf_knn on the
 * fp32 SIMT kernel pair (csrc/bf_knn.cu) with NO allocation of any kind: the per-split partial lists, which
 * b200ann_bf_knn takes from the stream-ordered allocator, live in one static device array of
 * B200ANN_NOALLOC_SCRATCH_BYTES that every GPU loading the library holds until exit. q, ids and dists must be DEVICE
 * memory (host buffers would need staging allocations -> B200ANN_ERR_ARG). Needs nq * k * 16 + nq bytes <=
 * B200ANN_NOALLOC_SCRATCH_BYTES (e.g. 1000 queries x k = 1000, or 100000 x k = 10), else B200ANN_ERR_ARG with the
 * sizes in the message; within the limit the base range is split over as many CTAs as the array holds lists for.
 * Results are those of b200ann_bf_knn, bit for bit. Every call shares the array (the part a call uses is cleared
 * first): calls that overlap on the GPU - two streams, two handles, two host threads - race; serialise them. Runs on
 * the handle's stream and synchronises it before returning unless the stream is being captured or b200ann_set_async
 * is on. */
#define B200ANN_NOALLOC_SCRATCH_BYTES ((size_t)16 << 20)
int b200ann_bf_knn_noalloc(b200ann_handle_t h, b200ann_index_t ix, const float *q /*[d]*/, size_t nq, size_t qstride,
                           size_t k, uint32_t *ids /*[d]*/, float *dists /*[d] or NULL*/);
/* replaces hnswlib::BruteforceSearch<int>::searchKnn (bruteforce.h:106-135) over hnswlib::L2SpaceI
 * (space_l2.h:255-323, L2SqrI4x / L2SqrI: int sum of (a-b)^2 over unsigned bytes), batched. base [n] rows of dim
 * bytes at bstride bytes, q [nq] rows at qstride bytes (host or device; copied per call when on the host).
 * ids/dists [nq,k] ascending by (distance,id); integer arithmetic, so both are bit-exact. Unfilled slots (k > n):
 * B200ANN_INVALID / INT32_MAX. dim <= 32768 (beyond it the reference's int accumulator can overflow). */
int b200ann_bf_knn_u8(b200ann_handle_t h, const uint8_t *base /*[h|d]*/, size_t n, size_t dim, size_t bstride,
                      const uint8_t *q /*[h|d]*/, size_t nq, size_t qstride, size_t k, uint32_t *ids /*[h|d]*/,
                      int32_t *dists /*[h|d] or NULL*/);
/* a device-resident copy of host bytes for b200ann_bf_knn_u8's `base` (so repeated searches do not re-upload the rows);
 * release with b200ann_free_device. */
int b200ann_upload_u8(b200ann_handle_t h, const uint8_t *src /*[h]*/, size_t bytes, uint8_t **out_dev);
int b200ann_free_device(void *dev);
/* exact kNN graph of the index's own rows, self excluded: knn [n,K] ascending (distance,id).
 * The exact counterpart of IndexGraph::Build's output (index_graph.cpp:275-292). */
int b200ann_knn_graph(b200ann_handle_t h, b200ann_index_t ix, size_t K, uint32_t *knn /*[h|d]*/);

/* rows [row0, row0 + nrows) of the same graph: knn [nrows,K]. The row-block partition of the kNN-graph build across
 * GPUs (BASELINE north_star "graph build is partitioned by row blocks"): every GPU holds the base, computes the rows
 * of its block and the blocks are exchanged with one all-gather; the union equals b200ann_knn_graph's output bit for
 * bit. */
int b200ann_knn_graph_rows(b200ann_handle_t h, b200ann_index_t ix, size_t row0, size_t nrows, size_t K,
                           uint32_t *knn /*[h|d]*/);

/* replaces efanna2e::IndexGraph::Build with the IndexRandom initializer (CNNS/efanna_graph/src/index_graph.cpp:
 * 265-294: InitializeGraph :175-201, NNDescent :122-136 = iter x {join :22-33, update :34-120}, top-K extraction
 * :275-291; index_random.cpp:24-27). K, L, iter, S, R are the reference's parameters of the same names. L2 indexes
 * only (the reference's Index base always constructs DistanceL2, CNNS/src/nsg/index.cpp:8-16).
 * knn [n,K] = the K best pool entries per node, ascending (distance,id), B200ANN_INVALID past the pool;
 * dists [n,K] optional. The approximate kNN graph b200ann_nsg_build / IndexNSG::Load_nn_graph consume.
 * Every phase is computed in an execution-order independent "batch form" (DESIGN.md): rand() -> counter hash of
 * (seed, node, draw), pools = sets of the L smallest admitted offers, reverse samples = R smallest hashes.
 * info (optional, 8 x u64): {distance evaluations, offers queued, pool inserts, iterations, join us, update us, 0, 0}.
 * Needs 2 <= S <= 64, S < L <= 512, K <= L, 0 < R <= 512, n > 2S+1. */
int b200ann_nndescent(b200ann_handle_t h, b200ann_index_t ix, size_t K, size_t L, size_t iter, size_t S, size_t R,
                      uint32_t seed, uint32_t *knn /*[h|d]*/, float *dists /*[h|d] or NULL*/, uint64_t *info /*[h]*/);

/* replaces efanna2e::IndexKDtree::Build (CNNS/efanna_graph/src/index_kdtree.cpp:279-435; driver
 * efanna_graph/tests/test_kdtree_graph.cpp): an initial approximate kNN graph from nTrees randomized KD-trees
 * (meanSplit over 101 sampled points, a random one of the 5 highest-variance dimensions, leaves of <= 10 points) whose
 * subtrees below depth mLevel are "conquered" - every point of an internal node's smaller child walks the other
 * child to a leaf and the distances to that leaf's points are offered to both ends' K-lists. "nTrees", "mLevel", "K"
 * are the reference's parameters of those names. The reference is not reproducible (std::random_shuffle, generators
 * seeded with the OpenMP thread number); this is the batch form oracle.c orc_kdtree_graph restates, bit for bit: the
 * shuffle of tree t is an affine permutation, the random pick a counter hash of (seed, tree, node), partitions are
 * stable, a K-list is the set of the K smallest (distance, id) offered. knn [n,K] ascending (distance, id); rows with
 * fewer than K entries are filled with distinct pseudo-random ids (dists +inf) like :413-424. L2 only. Feed it to
 * b200ann_nndescent_refine (IndexGraph::RefineGraph). info (optional, 8 x u64): {distance evaluations, tree nodes,
 * leaves, deepest level, levels, 0, 0, 0}. */
int b200ann_kdtree_graph(b200ann_handle_t h, b200ann_index_t ix, size_t K, size_t nTrees, size_t mLevel, uint32_t seed,
                         uint32_t *knn /*[h|d]*/, float *dists /*[h|d] or NULL*/, uint64_t *info /*[h]*/);
/* replaces efanna2e::IndexGraph::RefineGraph (CNNS/efanna_graph/src/index_graph.cpp:235-262 = InitializeGraph_Refine
 * :203-232 + NNDescent :122-136): NN-descent whose pools start from the rows of an initializer's graph - init_graph
 * [n, init_k], B200ANN_INVALID padding allowed, self / duplicate / out-of-range ids dropped, every entry flagged new,
 * init_k <= L - instead of from S random ids (what the KD-tree initializer of the reference feeds it, index_kdtree.cpp;
 * here any graph: b200ann_kdtree_graph, a previous b200ann_nndescent result, a graph file). Everything else is
 * b200ann_nndescent. Bit-exact against oracle.c orc_nndescent_refine. */
int b200ann_nndescent_refine(b200ann_handle_t h, b200ann_index_t ix, size_t K, size_t L, size_t iter, size_t S, size_t R,
                             uint32_t seed, const uint32_t *init_graph /*[h|d]*/, size_t init_k, uint32_t *knn /*[h|d]*/,
                             float *dists /*[h|d] or NULL*/, uint64_t *info /*[h]*/);

/* Which kernel b200ann_bf_knn / b200ann_knn_graph run on: 0 = auto (tensor cores when n >= 4096 and
 * nq*n >= 2^24, else fp32 SIMT), 1 = fp32 SIMT tiles (csrc/bf_knn.cu), 2 = tcgen05 tensor cores with fp32
 * re-ranking (csrc/knn_tc.cu). Both give the same bits. slack = extra shortlist entries per row kept by the
 * tensor path beyond k (< 0 = default max(32, k/4)); rows whose shortlist cannot be PROVEN to hold the exact
 * top k are recomputed by the SIMT kernel, so slack only moves work between the two. */
int b200ann_set_knn_mode(b200ann_handle_t h, int mode, int slack);
/* statistics of the last bf_knn / knn_graph call on this handle, 8 x u64: {query rows, rows recomputed by the
 * SIMT fallback, shortlisted candidates re-ranked in fp32, tensor passes (1 = bf16-exact data, 3 = bf16x3),
 * base-range splits, shortlist size, 128x256 tiles, A-resident}; all 0 when the SIMT kernel ran. */
int b200ann_knn_last_stats(b200ann_handle_t h, uint64_t *out8);

/* ---- graph search ------------------------------------------------------------------------- */
/* replaces efanna2e::IndexNSG::Search / SearchWithOptGraph (index_nsg.cpp:506-568, :570-657) for a
 * batch of queries. K results per query from a sorted pool of L (= "L_search"), seeds = first L
 * out-neighbours of ep then pseudo-random ids (counter hash of (seed, query index, draw)).
 * ids [nq,K]; dists [nq,K] or NULL (the reference returns ids only); stats [nq,2] u32 =
 * {distance evaluations, expanded nodes} or NULL. */
int b200ann_nsg_search(b200ann_handle_t h, b200ann_index_t ix, const float *q /*[h|d]*/, size_t nq,
                       size_t qstride, size_t K, size_t L, uint32_t seed, uint32_t *ids /*[h|d]*/,
                       float *dists /*[h|d]*/, uint32_t *stats /*[h|d]*/);

/* replaces hnswlib::HierarchicalNSW::searchKnn (hnswalg.h:1270-1324: upper-layer greedy descent +
 * searchBaseLayerST<true> :309-440) for a batch of queries, in the [nq,k] row shape of
 * python_bindings/bindings.cpp:612-695. labels/dists ascending by (distance, internal id);
 * beam = max(ef,k). Returns B200ANN_ERR_FEWER if some row has fewer than k results (rows are then
 * padded with ~0 / +inf). */
int b200ann_hnsw_search(b200ann_handle_t h, b200ann_index_t ix, const float *q /*[h|d]*/, size_t nq,
                        size_t qstride, size_t k, size_t ef, uint64_t *labels /*[h|d]*/,
                        float *dists /*[h|d]*/, uint32_t *stats /*[h|d]*/);

/* replaces hnswlib::HierarchicalNSW::searchKnn(query, k, isIdAllowed) for a batch of queries when a filter functor
 * is given and / or elements are marked deleted (hnswalg.h:1305-1313 -> searchBaseLayerST<false>, :309-440, no stop
 * condition; markDelete / unmarkDelete :853-911). The host functor and the delete marks become ONE bitmap over
 * internal ids (= rows of the index): bit (i & 31) of allow[i >> 5] set <=> element i is not deleted and its label
 * passes the filter; allow holds ceil(n/32) words. Elements that are filtered out are still walked through, exactly
 * like the reference; rows with fewer than k allowed results are padded with ~0 / +inf (no error: the reference
 * returns short results too). The reference's candidate_set is unbounded; the first pass bounds it per query
 * (shared memory) and PROVES per query whether the bound mattered; queries for which it could have are searched again
 * with global-memory heaps that hold all n elements and cannot overflow, so every answer is the reference's bit for
 * bit. inexact [nq] (host, optional) reports queries that stayed unproven (all 0 today; B200ANN_ERR_CAPACITY when
 * inexact is NULL and one did). */
int b200ann_hnsw_search_filtered(b200ann_handle_t h, b200ann_index_t ix, const float *q /*[h|d]*/, size_t nq,
                                 size_t qstride, size_t k, size_t ef, const uint32_t *allow /*[h|d]*/,
                                 uint64_t *labels /*[h|d]*/, float *dists /*[h|d]*/, uint32_t *stats /*[h|d]*/,
                                 uint8_t *inexact /*[h] or NULL*/);

/* replaces hnswlib::HierarchicalNSW::searchStopConditionClosest(query, EpsilonSearchStopCondition(epsilon, min, max),
 * isIdAllowed) for a batch of queries (hnswalg.h:1327-1378 -> searchBaseLayerST<false>(.., ef = 0, filter, stop
 * condition) :309-440; stop_condition.h:219-275): a range search that walks like a search with ef =
 * max_num_candidates and also stops once the closest open candidate lies outside the epsilon ball and at least
 * min_num_candidates results are held; afterwards results farther than epsilon are dropped (filter_results).
 * epsilon is in the index's distance (squared L2, or 1 - inner product). allow = the bitmap of
 * b200ann_hnsw_search_filtered, or NULL (no filter functor, nothing deleted). Output rows have max_num_candidates
 * slots, ascending by distance, counts[q] of them valid, the rest ~0 / +inf. internal_ids != 0 returns internal ids
 * instead of labels - what the reference does (it copies the (dist, tableint) pairs of top_candidates into the
 * result without getExternalLabel, hnswalg.h:1368-1371); 0 returns labels. A query the bounded first pass could not
 * prove exact is searched again like in b200ann_hnsw_search_filtered; B200ANN_ERR_CAPACITY if one stayed unproven. */
int b200ann_hnsw_search_epsilon(b200ann_handle_t h, b200ann_index_t ix, const float *q /*[h|d]*/, size_t nq,
                                size_t qstride, float epsilon, size_t min_num_candidates, size_t max_num_candidates,
                                const uint32_t *allow /*[h|d] or NULL*/, int internal_ids,
                                uint64_t *labels /*[h|d] [nq,max]*/, float *dists /*[h|d] [nq,max]*/,
                                uint32_t *counts /*[h|d] [nq]*/, uint32_t *stats /*[h|d] or NULL*/);

/* replaces hnswlib::HierarchicalNSW::searchStopConditionClosest(query, MultiVectorSearchStopCondition(space,
 * num_docs_to_search, ef_collection), isIdAllowed) for a batch of queries over a multi-vector space
 * (stop_condition.h:9-215: every element carries a document id behind its vector, several elements per document;
 * hnswalg.h:1327-1378, :309-440). doc_ids [n] = the document id of every element (internal id order; what
 * BaseMultiVectorSpace::get_doc_id reads from the row). The walk bounds its result set by the number of DISTINCT
 * documents in it (ef_collection, raised to num_docs_to_search like the reference) instead of by its size: a
 * candidate is considered while fewer documents are held or it is closer than the farthest result; the farthest
 * results are dropped while more than ef_collection documents are held; afterwards the farthest results go until
 * num_docs_to_search documents are left (filter_results). Rows of max_results slots, ascending by distance,
 * counts[q] valid; a query whose result list would need more slots -> B200ANN_ERR_CAPACITY (raise max_results; at
 * most about 28000). allow / internal_ids as in b200ann_hnsw_search_epsilon. */
int b200ann_hnsw_search_multivector(b200ann_handle_t h, b200ann_index_t ix, const float *q /*[h|d]*/, size_t nq,
                                    size_t qstride, const uint32_t *doc_ids /*[h|d] [n]*/, size_t num_docs_to_search,
                                    size_t ef_collection, size_t max_results, const uint32_t *allow /*[h|d] or NULL*/,
                                    int internal_ids, uint64_t *labels /*[h|d] [nq,max_results]*/,
                                    float *dists /*[h|d] [nq,max_results]*/, uint32_t *counts /*[h|d] [nq]*/,
                                    uint32_t *stats /*[h|d] or NULL*/);

/* replaces hnsw_nsg::IndexNSG::SearchFromEnterpoint (hnsw_nsg/src/index_nsg.cpp:703-782) for a batch:
 * like b200ann_nsg_search, but query i starts from entry[i] instead of the navigating node. The
 * reference's seed rules are kept: the entry point itself is marked visited and never
 * evaluated; seeds = its first L out-neighbours, then neighbours of those seeds (a slot whose
 * neighbour was already visited holds node 0), then pseudo-random ids. Needs K <= L < n. */
int b200ann_nsg_search_from(b200ann_handle_t h, b200ann_index_t ix, const float *q /*[h|d]*/, size_t nq,
                            size_t qstride, size_t K, size_t L, uint32_t seed, const uint32_t *entry /*[h|d] [nq]*/,
                            uint32_t *ids /*[h|d]*/, float *dists /*[h|d] or NULL*/, uint32_t *stats /*[h|d] or NULL*/);

/* replaces hnsw_nsg::HNSW_NSG::searchKnn (hnsw_nsg/include/index_hnsw_nsg.h:107-151) for a batch:
 * greedy descent on the HNSW levels >= 1 of the index (b200ann_index_set_hnsw), external label of the
 * node it ends on = NSG entry point, then SearchFromEnterpoint over the NSG graph
 * (b200ann_index_set_graph / b200ann_nsg_build) in ONE kernel. K = the reference's max(k, ef_), L = its
 * "L_search". ids [nq,K]; entry_out [nq] (optional) = the entry labels; stats count both phases. */
int b200ann_hybrid_search(b200ann_handle_t h, b200ann_index_t ix, const float *q /*[h|d]*/, size_t nq,
                          size_t qstride, size_t K, size_t L, uint32_t seed, uint32_t *ids /*[h|d]*/,
                          float *dists /*[h|d] or NULL*/, uint32_t *entry_out /*[h|d] or NULL*/,
                          uint32_t *stats /*[h|d] or NULL*/);

/* replaces the per-cluster `nsg_index->Search(query, cluster_data, search_K, paras, tmp)` loop of the reference's
 * cluster drivers (CNNS/tests/cluster_hnsw_nsg_search.cpp:189-215, ..._pipeline.cpp:165-270) for a batch: the index
 * holds ALL clusters back to back (rows of cluster c = [base_c, base_c + size_c)) under one block-diagonal graph
 * with global ids (each cluster's NSG shifted by base_c); query i searches the cluster given by ep[i] (that cluster's
 * navigating node, global id), base[i], size[i] exactly like IndexNSG::Search would search the cluster alone
 * (random seed fill inside the cluster; a cluster smaller than L is searched with a pool of its size).
 * size[i] == 0 skips query i (its row is B200ANN_INVALID / +inf). ids are global rows of this index. */
int b200ann_nsg_search_clustered(b200ann_handle_t h, b200ann_index_t ix, const float *q /*[h|d]*/, size_t nq,
                                 size_t qstride, size_t K, size_t L, uint32_t seed, const uint32_t *ep /*[h|d] [nq]*/,
                                 const uint32_t *base /*[h|d] [nq]*/, const uint32_t *size /*[h|d] [nq]*/,
                                 uint32_t *ids /*[h|d]*/, float *dists /*[h|d] or NULL*/, uint32_t *stats /*[h|d] or NULL*/);

/* the whole per-query loop of the reference's cluster search driver (CNNS/tests/cluster_hnsw_nsg_search.cpp:138-270)
 * for a batch, as one stream of launches with no host round trip: the nprobe nearest routing points by an exact scan
 * of route_ix (:143-160; the reference asks a faiss IndexHNSWFlat) -> clusters ranked by how many of their routing
 * points were hit, ties to the smaller cluster id (:162-187) -> b200ann_nsg_search_clustered in "every query's r-th
 * cluster" for r = 0..nprobe-1 with the early-stop rule between ranks when early_stop != 0 (:249-263: a query stops
 * after the first cluster whose best distance is not below the k-th smallest distance found so far) -> the k smallest
 * of everything found (:217-247, :265-270), local rows mapped through perm (mapping_<id>, NULL = rows).
 * ix: all clusters back to back + block-diagonal graph (see b200ann_nsg_search_clustered); route_ix: n_clusters *
 * per_cluster rows, cluster c's routing points (centroid + m samples: per_cluster = m + 1) at [c*per_cluster, ...);
 * cl_ep / cl_base / cl_size [n_clusters]: navigating node, first row, row count of every cluster.
 * ids / dists [nq,k] (B200ANN_INVALID / +inf padded); searched [nq] or NULL: clusters searched per query. */
int b200ann_cluster_search(b200ann_handle_t h, b200ann_index_t ix, b200ann_index_t route_ix, size_t per_cluster,
                           const uint32_t *cl_ep /*[h|d]*/, const uint32_t *cl_base /*[h|d]*/,
                           const uint32_t *cl_size /*[h|d]*/, size_t n_clusters, const uint32_t *perm /*[h|d] [n] or NULL*/,
                           const float *q /*[h|d]*/, size_t nq, size_t qstride, size_t k, size_t nprobe, size_t search_K,
                           size_t L, uint32_t seed, int early_stop, uint32_t *ids /*[h|d]*/, float *dists /*[h|d] or NULL*/,
                           uint32_t *searched /*[h|d] or NULL*/);

/* ---- shard-result merge (multi-GPU search) ---------------------------------------------------- */
/* d/ids [parts, nq, k]: per-shard result lists, each row ascending by distance (what every search of this library
 * writes), ids global, unfilled slots +inf / B200ANN_INVALID. out_d (may be NULL) / out_i [nq, kout] = the kout
 * smallest of the union, ascending by distance; equal distances keep (part, position) order - with row shards and
 * (distance,id)-sorted lists that is (distance,id) order. GPU counterpart of the reference's per-cluster result merge
 * (CNNS/tests/cluster_hnsw_nsg_search.cpp:189-270; merge_topk_queue, ..._pipeline.cpp:195-212);
 * the lists come from one NCCL all-gather of every rank's b200ann_nsg_search / _hnsw_search output. */
int b200ann_merge_topk(b200ann_handle_t h, const float *d /*[h|d]*/, const uint32_t *ids /*[h|d]*/, size_t parts,
                       size_t nq, size_t k, size_t kout, float *out_d /*[h|d]*/, uint32_t *out_i /*[h|d]*/);

/* the same merge over the buffer ONE all-gather fills: buf = world blocks, block r = what rank r sent:
 * [distances: lp x nq x k floats][ids: lp x nq x k u32] (lp lists per rank: the shards it owns). The searches write
 * straight into the rank's block (ids already global through b200ann_set_id_base), so a multi-GPU step is
 * lp searches -> one ncclAllGather -> this kernel. Device memory only. */
int b200ann_merge_topk_gathered(b200ann_handle_t h, const void *buf /*[d]*/, size_t world, size_t lp, size_t nq, size_t k,
                                size_t kout, float *out_d /*[d] or NULL*/, uint32_t *out_i /*[d]*/);

/* Stream-ordered mode for pipelines that keep everything on the device (multi-GPU search: search -> all-gather ->
 * merge): when on, b200ann_nsg_search, _nsg_search_from, _hybrid_search and b200ann_merge_topk(_gathered) calls whose inputs AND outputs are all device memory
 * return as soon as their kernels are enqueued on the handle's stream (b200ann_stream); the caller orders later work
 * on that stream or synchronizes it. b200ann_last_kernel_ms is not updated by such calls; a kernel fault surfaces at
 * the next synchronizing call. Calls with any host buffer still synchronize. Default off. */
int b200ann_set_async(b200ann_handle_t h, int on);

/* Scheduling / tuning knobs of the graph-search kernels; none of them changes a result (-1 = leave as is).
 * prefetch: while a node is expanded, pull the adjacency row of the next unexpanded pool entry into L2 (the reference's
 *   _mm_prefetch hints, index_nsg.cpp:627-632, hnswalg.h:370-383); default on.
 * visited_compact: 16-bit visited-set entries (twice the slots in the same shared memory) when n allows; default off
 *   (a probe costs ~3x the instructions of a 32-bit atomicCAS; measured slower despite fewer re-evaluations).
 * ctas_per_sm: cap on resident blocks per SM of the persistent kernel (0 = what occupancy allows). */
int b200ann_set_search_options(b200ann_handle_t h, int prefetch, int visited_compact, int ctas_per_sm);
/* Which graph-search kernel runs (results are bit-identical): version 2 = rows through the TMA bulk-copy
 * engine into a per-warp shared-memory ring of `ring_stages` stages (0 = default 1) with the pool in registers for
 * L <= 64 (csrc/beam2.cuh); version 1 = the round-1 register-load kernel (csrc/beam.cuh); version 0 (default) =
 * version 2 for beams of L <= 64, version 1 beyond (where it is the faster one) and for byte-row indexes.
 * mode (version 2): 0 (default) = full rounds of the resident warps run one query per warp, a remainder of at most one
 * round of blocks runs one query per block of 4 warps; 1 = one query per warp only; 2 = one query per block only.
 * -1 leaves a value as it is. */
int b200ann_set_search_kernel(b200ann_handle_t h, int version, int ring_stages, int mode);
/* Row shards (multi-GPU search): id_base is added to every valid id the NSG-family searches of this handle write
 * (b200ann_nsg_search, _nsg_search_from, _hybrid_search, _nsg_search_clustered), i.e. local row -> global row number
 * in the kernel's own epilogue. Default 0. */
int b200ann_set_id_base(b200ann_handle_t h, uint32_t id_base);
/* Work order of the following b200ann_*_search calls on this handle whose nq == n: work item w processes query
 * order[w] (a permutation of 0..n-1 in DEVICE memory, borrowed until reset with NULL). Queries are independent, so the
 * order only decides which queries run at the same time (queries that walk the same region of the graph share L2). */
int b200ann_set_query_order(b200ann_handle_t h, const uint32_t *order /*[d] or NULL*/, size_t n);
/* Measurement aid: reads `rows` whole rows of the index at pseudo-random positions with the search kernels' own loads
 * (16 B per lane, a team of lanes per row, `inflight` in {1,2,4} rows per team in flight) and returns the device time.
 * rows * dim*4 / ms = what HBM delivers for graph search's access pattern on this index (the random-gather ceiling
 * bench.py reports next to the streaming-copy peak). */
int b200ann_probe_gather(b200ann_handle_t h, b200ann_index_t ix, uint64_t rows, int inflight, int warps_per_block,
                         int blocks_per_sm, uint32_t seed, float *ms /*[h]*/);

/* tuning knob: log2 of the per-query visited-set size in shared-memory words. 0 (default) = auto: search kernel v2
 * takes 2^9 .. 2^11, whatever keeps the most warps resident; v1 and the build kernels use 2^10. Results never depend
 * on it (a forgotten node is re-evaluated and rejected again). */
int b200ann_set_visited_bits(b200ann_handle_t h, int bits);
/* same for the per-node searches of b200ann_nsg_link / b200ann_hnsw_build. 0 (default) = auto: 9 (dim <= 128) or 10 for
 * NSG Link, which is bound by shared-memory occupancy, 11 for the HNSW build. Results do not depend on it. */
int b200ann_set_link_visited_bits(b200ann_handle_t h, int bits);

/* ---- NSG construction --------------------------------------------------------------------- */
/* replaces efanna2e::IndexNSG::Build (CNNS/src/nsg/index_nsg.cpp:465-504): init_graph, Link
 * (get_neighbors + sync_prune per node, then InterInsert), compaction, tree_grow. knn [n,kstride] is
 * the kNN graph IndexNSG::Load_nn_graph (:128-148) reads (rows may be padded with B200ANN_INVALID),
 * e.g. the output of b200ann_knn_graph. L, R, C are the reference's "L", "R", "C" parameters.
 * ep0 = the node the reference draws with rand() % nd_ (:299); B200ANN_INVALID = derive it from seed.
 * On success the index holds the NSG graph (b200ann_nsg_search / b200ann_index_get_graph).
 * InterInsert runs in its order-independent batch form; rows have R + slack slots
 * (slack = 14) for the edges tree_grow appends.
 * info (optional, 8 x u64): {ep, width, tree_grow edges added, beam distance evaluations, expanded
 * nodes, prune distance evaluations, InterInsert distance evaluations, 0}. */
int b200ann_nsg_build(b200ann_handle_t h, b200ann_index_t ix, const uint32_t *knn /*[h|d]*/, size_t kstride,
                      size_t L, size_t R, size_t C, uint32_t seed, uint32_t ep0, uint64_t *info /*[h]*/);

/* the phases of b200ann_nsg_build one by one (parity tests compare each with the oracle) */
/* init_graph (:287-303): centroid + beam search from ep0; center_out [dim] may be NULL */
int b200ann_nsg_init_ep(b200ann_handle_t h, b200ann_index_t ix, const uint32_t *knn /*[h|d]*/, size_t kstride,
                        size_t L, uint32_t seed, uint32_t ep0, uint32_t *ep_out /*[h]*/, float *center_out /*[h|d]*/);
/* Link's first loop (:437-457): cut_ids/cut_d [n,R], unused tail = B200ANN_INVALID / -1;
 * stats [n,3] u32 = {beam distance evaluations, expanded nodes, prune distance evaluations} or NULL */
int b200ann_nsg_link(b200ann_handle_t h, b200ann_index_t ix, const uint32_t *knn /*[h|d]*/, size_t kstride,
                     uint32_t ep, size_t L, size_t R, size_t C, uint32_t seed, uint32_t *cut_ids /*[h|d]*/,
                     float *cut_d /*[h|d]*/, uint32_t *stats /*[h|d]*/);
/* InterInsert (:357-424, batch form), in place on cut_ids/cut_d [n,R]; n_eval (optional) = distance evaluations */
int b200ann_nsg_interinsert(b200ann_handle_t h, b200ann_index_t ix, size_t R, uint32_t *cut_ids /*[h|d]*/,
                            float *cut_d /*[h|d]*/, uint64_t *n_eval /*[h]*/);
/* compaction (:473-487) + tree_grow (:684-764): installs the graph in the index; rows get R+slack slots */
int b200ann_nsg_finish(b200ann_handle_t h, b200ann_index_t ix, const uint32_t *cut_ids /*[h|d]*/,
                       const float *cut_d /*[h|d]*/, size_t R, uint32_t ep, size_t L, uint32_t seed, size_t slack,
                       uint32_t *added /*[h]*/);

#pragma GCC visibility pop
#ifdef __cplusplus
}
#endif
#endif
