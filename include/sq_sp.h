/*
 * sq_sp.h -- C ABI of the B200 shortest-path library (HP-2).
 *
 * Part 1 is the drop-in surface: the ten symbols the reference's `sp` library exports from
 * projects/bolinas_civic/sp/src/py.cc:6-88 and that projects/bolinas_civic/sp/interface.py:12-67 binds with
 * ctypes.  Names, argument order, argument meaning and return conventions are the reference's; the
 * work behind `dijkstra` runs on the GPU (one near-far SSSP tree per CTA, f64).
 *
 * Part 2 is additive: batched trees (one tree per origin), device-resident variants that take
 * plain device pointers (no torch types), weight upload, and tuning knobs.
 *
 * All functions are `extern "C"`, take plain pointers and sizes, and never throw across the
 * boundary.  There is no CPU fallback: every compute entry point returns an error (or aborts with a
 * message for the reference-shaped calls that have no error channel) when no CUDA device is usable.
 */
#ifndef SQ_SP_H_
#define SQ_SP_H_

#include <stdbool.h>
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct SqGraph Graph;               /* opaque; reference: class Graph, sp/include/graph.h:57 */
typedef struct SqShortestPath ShortestPath; /* opaque; reference: struct ShortestPath, sp/include/graph.h:20 */

/* ---- Part 1: reference-compatible surface (sp/src/py.cc) ------------------------------------ */

/* py.cc:11-15  toy graph of graph.cc:121-135 (7 vertices, 11 edges). */
Graph* simplegraph(bool directed);
/* py.cc:21-25  MatrixMarket reader, graph.cc:55-97 (first data line is the header; nvertices = n+1). */
Graph* readgraph(char* filename, bool directed);
/* py.cc:36-41  COO edge list; `nvertices` is the maximum vertex id (graph.cc:141 adds one). */
Graph* creategraph(int* edge_starts, int* edge_ends, double* edge_wghs, int nedges, int nvertices,
                   bool directed);
/* py.cc:46-48  MatrixMarket writer, graph.cc:100-119 (unique (v1,v2) keys in key order). */
bool writegraph(Graph* graph, char* filename);
/* py.cc:54-56  one tree from `origin`; destination -1 = all.  The GPU always settles the whole
 * tree, which is a superset of the reference's early-exit result (graph.cc:196).  Invalid ids give
 * a tree with no distances (graph.cc:161-163). */
ShortestPath* dijkstra(Graph* graph, int origin, int destination);
/* py.cc:63-65  set the weight of edge (vertex1, vertex2).  A missing edge aborts the reference
 * (std::out_of_range across the C boundary); here it is ignored and counted (sp_last_error). */
void update_edge(Graph* graph, int vertex1, int vertex2, double weight);
/* py.cc:69 */
int origin(ShortestPath* sp);
/* py.cc:74-77  DBL_MAX when the tree is empty or the vertex is unreachable. */
double distance(ShortestPath* sp, int destination);
/* py.cc:80 */
void clear(ShortestPath* sp);
/* py.cc:85-87  -1 for the origin and for unreachable vertices.  Out-of-range ids return -1
 * (the reference throws). */
int parent(ShortestPath* sp, int destination);

/* ---- Part 2: additive surface ------------------------------------------------------------------ */

/* Graph::remove_edge (graph.cc:40-52) is public C++ in the reference but not exported; exported
 * here so the reference's own known-answer test (sp/tests/graph_test.cc:47-64, 95-103) can be
 * replayed through the C ABI.  Removes the edge currently mapped at key (v1,v2) (after the
 * undirected swap).  Returns 0 on success, -1 if absent. */
int sp_remove_edge(Graph* graph, int vertex1, int vertex2);
void sp_free_graph(Graph* graph);
int sp_nvertices(Graph* graph);
int sp_nedges(Graph* graph); /* adjacency entries (duplicates included) */
int sp_nedges_logged(Graph* graph); /* edges ever inserted = length of the sp_set_weights arrays */

/* One full tree per origin.  Host buffers: dist_out[n][V] (f64, DBL_MAX = unreachable) and
 * parent_out[n][V] (i32, -1 = none); either may be NULL.  Predecessors follow the canonical rule
 * parent[v] = min{u : fl(dist[u] + w(u,v)) == dist[v]}.  Returns 0, or a negative error code. */
int sp_sssp_batch(Graph* graph, const int* origins, int n, double* dist_out, int* parent_out);
/* reverse != 0: trees on the reversed graph (distances to, and next vertex towards, each origin). */
int sp_sssp_batch_ex(Graph* graph, const int* origins, int n, double* dist_out, int* parent_out, int reverse);

/* sp_sssp_batch_ex with the predecessors in the compact form of sp_set_parent_format 1: codes_out[n][V], one byte per
 * vertex = the index of the in-edge (creation order at that vertex; out-edge for reverse trees) that attains the
 * distance, 255 = none; dist_out may be NULL.  Every consumer that holds the graph can follow the codes directly
 * (sp_decode_parents_host turns them into vertex ids: parent_out[n][V]).  On config 5 a tree is 50 KB over PCIe instead
 * of 602 KB, so the call is bound by the tree kernel, not by the link. */
int sp_sssp_batch_codes(Graph* graph, const int* origins, int n, double* dist_out, unsigned char* codes_out, int reverse);
int sp_decode_parents_host(Graph* graph, const unsigned char* codes, int n, int* parent_out, int reverse);
/* The same with two vertices per byte: codes_out[n][(V + 1) / 2], low nibble = even vertex, 15 = none.  For graphs whose
 * largest in-degree (out-degree for reverse trees) is at most 15 -- road networks -- else SQ_ERR_LIMIT.  25 KB per tree on
 * config 5: with eight GPUs behind one host the byte form is bound by what the host can take in (83 GB/s on the
 * measured box), this form is not. */
int sp_sssp_batch_codes4(Graph* graph, const int* origins, int n, double* dist_out, unsigned char* codes_out, int reverse);
int sp_decode_parents_host4(Graph* graph, const unsigned char* codes, int n, int* parent_out, int reverse);

/* Agent.get_path for a batch of agents (queue_class.py:491-512 / model/queue_class.py:491-512: g.dijkstra(origin,
 * destin), sp.distance(destin), sp.route(destin)): for every pair k the distance origins[k] -> dests[k] (DBL_MAX
 * when there is no path, like distance()) and the path as a vertex sequence origins[k], ..., dests[k] -- what
 * route() returns as (start, end) pairs; empty when there is no path or origin == destination.  One full tree per
 * DISTINCT origin is computed on the device and walked there, so only distances and paths cross PCIe (the full
 * tables of sp_sssp_batch are 12 bytes per vertex per tree).  Host arrays: dist_out[n], route_off_out[n] (offset
 * of path k in the pool), route_len_out[n] (vertices in path k).  The pool is page-locked memory owned by the graph,
 * valid until the next sp_route_batch or sp_free_graph on it. */
int sp_route_batch(Graph* graph, const int* origins, const int* dests, int n, double* dist_out,
                   long long* route_off_out, int* route_len_out);
const int* sp_route_pool(Graph* graph);

/* Page-locked host memory for sp_sssp_batch outputs: with both destinations allocated here the device-to-host
 * copies overlap the computation of the next chunk of trees (PCIe-bound instead of staging-bound). */
void* sp_alloc_host(size_t bytes);
void sp_free_host(void* ptr);

/* Same with device pointers (CUDA device memory owned by the caller, e.g. a torch tensor's
 * data_ptr()); `stream` is a cudaStream_t passed as void* (NULL = default stream).  Asynchronous
 * with respect to the host.  Either output may be NULL (the tree is then kept in library scratch
 * only and just timed). */
int sp_sssp_batch_device(Graph* graph, const int* origins_dev, int n, double* dist_dev,
                         int* parent_dev, void* stream);

/* No counterpart in the reference: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same result as
 * sp_sssp_batch_device (bit-identical distances and canonical predecessors; dist_dev is required, parent_dev may be
 * null) computed by the bitmap tree kernel with its scratch in a file-scope device array of 8 MiB + 256 bytes that every
 * GPU loading the library holds until exit.  Calls no cudaMalloc / cudaFree / allocating library routine.  Limits: at
 * most 65,536 vertices (SQ_ERR_LIMIT beyond); the graph must already be on the device (sp_sync after the last edit,
 * else SQ_ERR_STATE).  Every call shares the one array: calls that overlap on a GPU (two streams, two threads) race. */
int sp_sssp_batch_device_noalloc(Graph* graph, const int* origins_dev, int n, double* dist_dev, int* parent_dev,
                                 void* stream);
/* sp_sssp_batch_device with a direction flag: reverse != 0 runs the trees on the reversed graph, i.e.
 * dist[i][v] = shortest distance from v TO origins[i] and parent[i][v] = the next vertex on that path (canonical:
 * smallest id among exact ties).  One reverse tree per destination routes every agent heading there
 * (Agent.get_path, queue_class.py:491-512, calls dijkstra once per agent instead). */
int sp_sssp_batch_device_ex(Graph* graph, const int* origins_dev, int n, double* dist_dev, int* parent_dev,
                            void* stream, int reverse);

/* Predecessor output format of the device batches: 0 (default) = int32 vertex ids; 1 = one byte per vertex, the
 * index of the in-edge (creation order at that vertex) that attains the distance, 255 = none.  The compact form
 * is what the multi-GPU path all-gathers (4x less NVLink traffic); sp_decode_parents_device expands it with the
 * graph replica every rank holds.  Host batches (sp_sssp_batch) always return vertex ids. */
int sp_set_parent_format(Graph* graph, int format);
int sp_decode_parents_device(Graph* graph, const unsigned char* codes_dev, int n, int* parent_dev, void* stream,
                             int reverse);

/* Device-side sp_set_weights: weights_dev[nedges] in creategraph order (device memory), NaN = keep the current
 * weight.  Rewrites the device graph in place with one kernel; the host edge log is kept in step. */
int sp_set_weights_device(Graph* graph, const double* weights_dev, int nedges, void* stream);

/* Bulk weight refresh, one weight per edge in insertion order (creategraph order); replaces n calls
 * to update_edge (queue_class.py:463-466 is called once per link every 100 ticks). */
int sp_set_weights(Graph* graph, const double* weights, int nedges);

/* 1 if update_edge / sp_set_weights / structural edits have not reached the device yet. */
int sp_has_pending_edits(Graph* graph);

/* Push pending host-side edits (update_edge / sp_set_weights / structure) to the device now.
 * Called implicitly by every compute entry point. */
int sp_sync(Graph* graph);

/* Tuning knobs for the tree kernels: threads per CTA = per tree (ELL kernel: 32, 64, 128 or 256; bitmap
 * kernel: 128..1024), CTAs per SM (0 = default), delta = delta_factor x mean finite edge weight (<=0 keeps
 * current), shared-memory queue slots per queue of the bitmap kernel (0 = default). */
int sp_set_tuning(Graph* graph, int threads, int ctas_per_sm, double delta_factor, int smem_queue);

/* Kernel choice: 3 (default) = auto: the warp kernel for batches of at least 13 trees per SM (the group path, 4, only
 * if sp_set_group_tuning sets a batch-size threshold), the ELL kernel
 * below that (a lone tree finishes ~10x sooner on a whole CTA); if more than 5 % of a batch's trees overflow the
 * warp kernel's ring, the next batches use a 2048-entry ring and, failing that, the ELL kernel; 2 = warp kernel (one tree per warp, four-byte
 * hashed queue entries in a shared-memory ring), 0 = ELL kernel (one tree per CTA, four-edge vertex records, tagged queue entries, queues of
 * queue_factor x V entries), 1 = bitmap kernel only (CSR, per-vertex shared-memory bitmaps; also the fallback
 * that reruns trees whose queues overflowed in the other two), 4 = group kernel (eight trees per warp on interleaved
 * distance rows, no atomics, no overflow path).  All produce identical results.
 * queue_factor <= 0 keeps the current value. */
int sp_set_algo(Graph* graph, int algo, double queue_factor);

/* Group-kernel knobs (algo 4, and auto mode on dense batches): eight trees of neighbouring origins share one
 * frontier and one warp (csrc/sp_group.cuh).  warps_per_sm: resident groups per SM (multiple of 4, 0 keeps; default
 * 16); batch: pops decided per batch, 8 / 16 / 32 (mode 2: pops in flight, 4 / 8 / 16) (0 keeps; default 16); prefetch: 1 = prefetch a batch's rows
 * into L1 (< 0 keeps; default 1); auto mode picks the group path for batches of at least min_trees origins
 * (0 keeps; default: never -- on config 5 the warp kernel is the faster one) that hold at least
 * V / min_density origins (0 keeps; default 8). */
int sp_set_group_tuning(Graph* graph, int warps_per_sm, int batch, int prefetch, int min_trees, int min_density);
/* Which tree kernel works on the group layout: 0 (default) = the warp kernel, a CTA of eight warps per group and one
 * tree per warp (ring_cap entries per tree, 0 keeps; default 1024; trees that outgrow it are rerun by the bitmap
 * kernel); 1 = sssp_group_kernel, one warp per group, no atomics and no capacity limits (also what mode 0 falls back
 * to when the graph is too large for the bitmap kernel); 2 = sssp_union_kernel, one warp per group with a shared
 * frontier, 64-bit atomics on the interleaved rows and `batch` (sp_set_group_tuning: 4, 8 or 16) pops in flight; two
 * bits per vertex per warp in shared memory bound it to about 400k vertices (larger graphs run mode 1). */
int sp_set_group_mode(Graph* graph, int mode, int ring_cap);
/* Auto mode, batches of at most one tree per SM (the device router's reverse tree per destination): keep each tree's
 * distance row in shared memory while it is computed (graphs of up to ~25.7k vertices on B200); 1 = on (default), 0 = off. */
int sp_set_smem_rows(Graph* graph, int on);

/* Warp-kernel knobs: ring capacity in entries (rounded up to a power of two; 0 keeps; default 1024), frontier
 * vertices per step = 8 x ilp (1, 2 or 4; 0 keeps; default 4), CTAs of four warps per SM (0 = default 4, at most
 * what shared memory allows; < 0 keeps). */
int sp_set_warp_tuning(Graph* graph, int ring_cap, int ilp, int ctas_per_sm);

/* Statistics of the last batch: [0] trees, [1] rounds, [2] relaxations, [3] bucket transitions,
 * [4] trees the ELL kernel handed to the bitmap kernel (queue overflow).  Filled only when stats were enabled at build time
 * (SQ_SSSP_STATS); otherwise zeros. */
int sp_last_stats(Graph* graph, long long* out5);

/* Last error text of the calling thread ("" if none). */
const char* sp_last_error(void);

/* Number of CUDA kernels this library has launched in this process (for bench gpu_launches). */
long long sp_kernel_launches(void);

#ifdef __cplusplus
}
#endif
#endif /* SQ_SP_H_ */
