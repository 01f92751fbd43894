// sp_graph.cu -- B200 shortest-path library behind the reference's `sp` C ABI (include/sq_sp.h).
//
// What it replaces: projects/bolinas_civic/sp/src/graph.cc (Graph, dijkstra_priority_queue :153-218)
// and sp/src/py.cc (the extern "C" surface).  Nothing here is derived from that code: the host side
// is a COO edge log + CSR/CSC build, the device side is a near-far (delta-stepping) label-correcting
// SSSP, one tree per CTA, with queues staged in shared memory and distances updated by 64-bit L2
// atomicMin.  Distances are bit-identical to a binary-heap Dijkstra in f64 because both compute the
// least fixed point of d[v] = min_u fl(d[u] + w(u,v)) (fl(+) is monotone, weights are >= 0).
//
// Reference behaviours kept on purpose (each is exercised by tests against oracle/_ref/liblsp.so):
//  * nvertices = max id + 1 (graph.cc:141; readgraph: header + 1, graph.cc:80).
//  * "undirected" graphs store every edge as (min,max) and Dijkstra only ever relaxes edge.second
//    (graph.cc:7-8, 202), so an undirected graph is a directed graph oriented low id -> high id.
//  * duplicate (v1,v2) inserts stay in the adjacency (all are relaxed) while update_edge reaches only
//    the last one inserted (graph.cc:13, 16-19, 34-36).
#include <algorithm>
#include <cfloat>
#include <climits>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <fstream>
#include <iostream>
#include <map>
#include <mutex>
#include <sstream>
#include <string>
#include <unordered_map>
#include <numeric>
#include <cstring>
#include <vector>

#include "../../include/sq_sp.h"
#include "sq_common.cuh"

using sq::SQ_ERR_ARG;
using sq::SQ_ERR_CUDA;
using sq::SQ_ERR_LIMIT;
using sq::SQ_OK;

typedef unsigned long long u64;

#define SQ_MAX_DEVICES 64
static std::mutex g_attr_mutex;   // kernel attributes are set once per device

// =================================================================================================
// Device side
// =================================================================================================

struct __align__(16) EdgeRec {  // one adjacency entry: 16 B so a relaxation needs a single LDG.128
  double w;
  int v;    // forward CSR: head vertex; reverse CSC: tail vertex
  int pad;
};

struct TreeParams {
  int V;
  int n_trees;
  const int* row_ptr;   // [V+1] forward CSR
  const EdgeRec* adj;   // [E]
  const int* rrow_ptr;  // [V+1] reverse CSC
  const EdgeRec* radj;  // [E]
  const int* origins;   // [n_trees] device
  u64* dist_out;        // [n_trees][V] or nullptr
  int* parent_out;      // [n_trees][V] or nullptr
  u64* dist_scratch;    // [grid][V], used when dist_out == nullptr
  int* qscratch;        // [grid][4][V] overflow parts of the four queues
  int* tree_counter;    // dynamic tree scheduler
  long long* stats;     // [5] or nullptr
  double delta;
  int nq;               // shared-memory slots per queue
  int words;            // ceil(V/32)
  // bitmap kernel as the fallback of the ELL kernel: trees to run come from a device list
  const int* tree_list;   // nullptr: trees 0..n_trees-1
  const int* n_list;      // device count of tree_list entries
  // ELL kernel
  const EdgeRec* ell;     // [V][4] forward; .pad = back-edge slot (3 bits, 7 = none) | (slot 0: overflow start, slot 1: count) << 3
  const EdgeRec* ovf;     // forward edges beyond the fourth of a vertex
  const EdgeRec* rell;    // [V][4] reverse
  const EdgeRec* rovf;
  int2* q2scratch;        // [grid][4*near_cap + 2*far_cap] u64 distances of the global queue parts
  int* q2v;               // same shape, vertex ids
  int near_cap, far_cap;
  int* fail_list;         // trees whose queues overflowed (rerun by the bitmap kernel)
  int* fail_count;
  // warp kernel
  unsigned* wfar;         // [warps in grid][2][wfar_cap] far piles
  int wfar_cap;
  int ring_cap;           // entries per shared-memory ring (power of two)
  int vbits;              // bits of a queue entry that hold the vertex id
  // grouped variant of the warp kernel (eight sibling trees per CTA on interleaved rows)
  u64* gdist;             // [groups][V][8]
  const int* tord;        // [groups * 8] tree at each sorted position, -1 = padding
  const int* perm;        // caller id -> internal id
  int parent_codes;       // bitmap kernel's own predecessor pass: 1 = one-byte in-edge codes
};

#define DIST_INF_BITS 0x7FEFFFFFFFFFFFFFull  // DBL_MAX, what the reference leaves in unreached slots

// Bucket thresholds must move: with delta below half an ulp of T (distances of 1e8 behind a closed link next to
// sub-nanosecond weights) T + delta == T and a bucket transition would find the same empty window for ever.
__device__ __forceinline__ double next_up(double x) { return __longlong_as_double(__double_as_longlong(x) + 1); }
__device__ __forceinline__ double advance(double T, double delta) {
  const double n = T + delta;
  return n > T ? n : next_up(T);
}

// A queue whose first `cap_s` slots live in shared memory and the rest in a per-CTA global scratch
// row of V ints.  Entries are unique (guarded by a shared-memory bit per vertex), so V slots suffice.
struct HQ {
  int* s;
  int* g;
  int cap_s;
  __device__ __forceinline__ void put(int pos, int v) const {
    if (pos < cap_s) s[pos] = v; else g[pos - cap_s] = v;
  }
  __device__ __forceinline__ int get(int pos) const {
    return pos < cap_s ? s[pos] : g[pos - cap_s];
  }
};

// Read-only 16-byte adjacency load through the non-coherent path (the graph is immutable in a launch).
__device__ __forceinline__ EdgeRec ld_edge(const EdgeRec* p) {
  const int4 raw = __ldg(reinterpret_cast<const int4*>(p));
  EdgeRec er;
  er.w = __hiloint2double(raw.y, raw.x);
  er.v = raw.z;
  er.pad = 0;
  return er;
}

// Warp-aggregated append: one shared-memory atomicAdd per warp, lanes take consecutive slots.
// Must be reached by all 32 lanes.
__device__ __forceinline__ void warp_push(bool want, int v, const HQ& q, int* cnt) {
  const unsigned m = __ballot_sync(0xffffffffu, want);
  if (m == 0) return;
  const int leader = __ffs(m) - 1;
  int base = 0;
  if ((int)sq::lane_id() == leader) base = atomicAdd(cnt, __popc(m));
  base = __shfl_sync(0xffffffffu, base, leader);
  if (want) q.put(base + __popc(m & sq::lanemask_lt()), v);
}

__device__ __forceinline__ bool test_and_set(unsigned* bits, int v) {
  const unsigned b = 1u << (v & 31);
  return (atomicOr(&bits[v >> 5], b) & b) != 0;
}
__device__ __forceinline__ void clear_bit(unsigned* bits, int v) {
  atomicAnd(&bits[v >> 5], ~(1u << (v & 31)));
}

template <int THREADS>
__global__ void __launch_bounds__(THREADS) sssp_nearfar_kernel(const TreeParams p) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  // layout: 3 bitmaps (near parity 0/1, far) | 4 queues (near0, near1, far0, far1)
  unsigned* bits_near0 = reinterpret_cast<unsigned*>(smem_raw);
  unsigned* bits_near1 = bits_near0 + p.words;
  unsigned* bits_far = bits_near1 + p.words;
  int* qs = reinterpret_cast<int*>(bits_far + p.words);

  __shared__ int s_cnt[4];  // near0, near1, far0, far1
  __shared__ int s_tree;
  __shared__ u64 s_min;

  const int tid = threadIdx.x;
  const int V = p.V;
  int* gq = p.qscratch + (size_t)blockIdx.x * 4 * (size_t)V;
  // queue i in {near0, near1, far0, far1}; built on the fly so nothing is indexed dynamically in
  // local memory
  auto Q = [&](int i) -> HQ { return HQ{qs + i * p.nq, gq + (size_t)i * V, p.nq}; };
  auto NB = [&](int parity) -> unsigned* { return bits_near0 + parity * p.words; };

  long long st_rounds = 0, st_relax = 0, st_trans = 0;

  for (;;) {
    if (tid == 0) {
      int k = atomicAdd(p.tree_counter, 1);
      if (p.tree_list) k = (k < *p.n_list) ? p.tree_list[k] : p.n_trees;
      s_tree = k;
    }
    __syncthreads();
    const int tree = s_tree;
    if (tree >= p.n_trees) break;

    u64* dist = p.dist_out ? p.dist_out + (size_t)tree * V : p.dist_scratch + (size_t)blockIdx.x * V;
    const int src = p.origins[tree];

    for (int i = tid; i < V; i += THREADS) dist[i] = DIST_INF_BITS;
    for (int i = tid; i < 3 * p.words; i += THREADS) bits_near0[i] = 0u;
    if (tid < 4) s_cnt[tid] = 0;
    __syncthreads();

    const bool valid = (src >= 0 && src < V);
    if (tid == 0 && valid) {
      dist[src] = 0ull;
      Q(0).put(0, src);
      s_cnt[0] = 1;
      bits_near0[src >> 5] |= 1u << (src & 31);
    }
    __syncthreads();

    double T = p.delta;
    int cur = 0, fcur = 0;

    while (valid) {
      const int n = s_cnt[cur];
      if (n == 0) {
        // ---- bucket transition: refill the near queue from the far pile ----------------------
        int nf = s_cnt[2 + fcur];
        if (nf == 0) break;
        for (;;) {
          __syncthreads();
          if (tid == 0) { s_min = DIST_INF_BITS; s_cnt[2 + (fcur ^ 1)] = 0; }
          __syncthreads();
          const double Told = T;
          T = advance(Told, p.delta);
          u64 lmin = DIST_INF_BITS;
          const HQ qfar_in = Q(2 + fcur), qfar_out = Q(2 + (fcur ^ 1)), qnear_out = Q(cur);
          for (int i0 = 0; i0 < nf; i0 += THREADS) {
            const int i = i0 + tid;
            bool to_near = false, to_far = false;
            int v = 0;
            if (i < nf) {
              v = qfar_in.get(i);
              const u64 db = sq::ld_cg_u64(&dist[v]);
              const double d = __longlong_as_double((long long)db);
              if (d < Told) {
                clear_bit(bits_far, v);  // already settled through the near queue
              } else if (d < T) {
                clear_bit(bits_far, v);
                to_near = !test_and_set(NB(cur), v);
              } else {
                to_far = true;  // far bit stays set
                lmin = db < lmin ? db : lmin;
              }
            }
            warp_push(to_near, v, qnear_out, &s_cnt[cur]);
            warp_push(to_far, v, qfar_out, &s_cnt[2 + (fcur ^ 1)]);
          }
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) {
            const u64 other = __shfl_xor_sync(0xffffffffu, lmin, o);
            lmin = other < lmin ? other : lmin;
          }
          if (sq::lane_id() == 0 && lmin != DIST_INF_BITS) atomicMin(&s_min, lmin);
          __syncthreads();
          fcur ^= 1;
          nf = s_cnt[2 + fcur];
          ++st_trans;
          if (s_cnt[cur] > 0 || nf == 0) break;
          // nothing fell into [Told, T): jump the threshold to the smallest far distance
          T = __longlong_as_double((long long)s_min);
        }
        continue;
      }

      // ---- one relaxation round over the near queue ------------------------------------------
      const HQ qin = Q(cur), qnear = Q(cur ^ 1), qfar = Q(2 + fcur);
      unsigned* bcur = NB(cur);
      unsigned* bnext = NB(cur ^ 1);
      for (int i0 = 0; i0 < n; i0 += THREADS) {
        const int i = i0 + tid;
        int b = 0, e = 0;
        double du = 0.0;
        if (i < n) {
          const int u = qin.get(i);
          clear_bit(bcur, u);
          b = __ldg(&p.row_ptr[u]);
          e = __ldg(&p.row_ptr[u + 1]);
          du = __longlong_as_double((long long)sq::ld_cg_u64(&dist[u]));
        }
        const int maxdeg = __reduce_max_sync(0xffffffffu, e - b);
        for (int k0 = 0; k0 < maxdeg; k0 += 4) {
          int v[4];
          u64 ndb[4], old[4];
          bool has[4];
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const int idx = b + k0 + j;
            has[j] = idx < e;
            v[j] = 0;
            ndb[j] = DIST_INF_BITS;
            if (has[j]) {
              const EdgeRec er = ld_edge(&p.adj[idx]);
              v[j] = er.v;
              ndb[j] = (u64)__double_as_longlong(du + er.w);
            }
          }
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            old[j] = 0ull;
            if (has[j]) old[j] = atomicMin(&dist[v[j]], ndb[j]);
          }
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const bool improved = has[j] && ndb[j] < old[j];
            const bool is_near = __longlong_as_double((long long)ndb[j]) < T;
            bool to_near = false, to_far = false;
            if (improved) {
              if (is_near) to_near = !test_and_set(bnext, v[j]);
              else to_far = !test_and_set(bits_far, v[j]);
            }
            warp_push(to_near, v[j], qnear, &s_cnt[cur ^ 1]);
            warp_push(to_far, v[j], qfar, &s_cnt[2 + fcur]);
            st_relax += has[j];
          }
        }
      }
      __syncthreads();
      if (tid == 0) s_cnt[cur] = 0;
      cur ^= 1;
      ++st_rounds;
      __syncthreads();
    }
    __syncthreads();

    // ---- canonical predecessor pass: parent[v] = min{u : fl(d[u] + w(u,v)) == d[v]} -----------
    if (p.parent_out) {
      int* parent = p.parent_out + (size_t)tree * V;
      unsigned char* pcode = reinterpret_cast<unsigned char*>(p.parent_out) + (size_t)tree * V;
      for (int v = tid; v < V; v += THREADS) {
        int best = -1, bslot = 255;
        if (valid && v != src) {
          const u64 dvb = sq::ld_cg_u64(&dist[v]);
          if (dvb != DIST_INF_BITS) {
            const int b = __ldg(&p.rrow_ptr[v]);
            const int e = __ldg(&p.rrow_ptr[v + 1]);
            best = INT_MAX;
            for (int idx = b; idx < e; ++idx) {
              const EdgeRec er = ld_edge(&p.radj[idx]);
              const u64 dub = sq::ld_cg_u64(&dist[er.v]);
              if (dub == DIST_INF_BITS) continue;
              const double cand = __longlong_as_double((long long)dub) + er.w;
              if ((u64)__double_as_longlong(cand) == dvb && er.v < best) { best = er.v; bslot = min(idx - b, 254); }
            }
            if (best == INT_MAX) { best = -1; bslot = 255; }
          }
        }
        if (p.parent_codes) pcode[v] = (unsigned char)bslot;
        else parent[v] = best;
      }
    }
    __syncthreads();
  }

  if (p.stats) {
    // per-thread relaxation counts -> one atomic per warp
    long long r = st_relax;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) r += __shfl_xor_sync(0xffffffffu, r, o);
    if (sq::lane_id() == 0) atomicAdd((u64*)&p.stats[2], (u64)r);
    if (tid == 0) {
      atomicAdd((u64*)&p.stats[1], (u64)st_rounds);
      atomicAdd((u64*)&p.stats[3], (u64)st_trans);
    }
  }
}


// -------------------------------------------------------------------------------------------------
// ELL kernel: the near-far schedule without per-vertex bitmaps, row pointers or far-pile rescans.
//
//  * adjacency = one 64-byte record of four 16-byte edges per vertex (road networks rarely exceed
//    degree 4; further edges sit in an overflow list).  Four adjacent lanes own one frontier vertex, so a
//    warp step touches 8 lines for its 32 edges instead of 32, and there is no row_ptr load in the chain.
//  * queue entries are (vertex, distance when pushed).  An entry whose distance no longer equals
//    dist[vertex] is stale: the thread that lowered the distance pushed a newer entry.  This replaces the
//    three shared-memory bitmaps (and their ATOMS traffic); duplicates are possible, so queues are
//    bounded and a tree that overflows is rerun by the bitmap kernel.
//  * three buckets instead of two: improvements below T go to the next round of the current bucket,
//    those in [T, T+delta) straight into the queue of the next bucket (which simply becomes the current
//    queue at the transition: no rescan, no dist reload), and only those beyond T+delta into a far pile
//    (empty on road networks whose link times stay below delta; closed links at 1e8 end up there).
//  * queues are hybrids: the first QS entries in shared memory, the rest in a per-CTA global scratch row;
//    four of them rotate (current / next round / spare being reset / next bucket) so that a round needs a
//    single __syncthreads.
//  * predecessors are a separate kernel (sssp_parent_kernel) at full occupancy.
// -------------------------------------------------------------------------------------------------
__device__ __forceinline__ int2 ld_cg_int2(const int2* p) {
  int2 v;
  asm volatile("ld.global.cg.v2.s32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "l"(p));
  return v;
}

constexpr int QS = 384;  // shared-memory entries per queue

// Queue entry = (vertex, vertex it was reached from, distance when pushed).  The 'from' vertex lets the
// relaxation skip the edge straight back (it can never improve: weights are >= 0), a quarter of all atomics
// on a grid.  Stored as int2 {v, from} + u64.
// The global parts are reached through a small table in shared memory (six pointers per array, filled once per
// CTA): building them per round cost ~100 dependent 64-bit multiply-adds on the path of every relaxation round,
// and they are only needed when a queue outgrows its shared part.
struct HQ2 {
  int2* sv;  // shared part
  u64* sd;
  int2* const* gvt;  // tables of the global parts, indexed by qi
  u64* const* gdt;
  int qi;
  __device__ __forceinline__ void get(int pos, int2& v, u64& d) const {
    if (pos < QS) { v = sv[pos]; d = sd[pos]; }
    else { v = ld_cg_int2(&gvt[qi][pos - QS]); d = sq::ld_cg_u64(&gdt[qi][pos - QS]); }
  }
  __device__ __forceinline__ void put(int pos, int2 v, u64 d, int gcap, int* fail) const {
    if (pos < QS) { sv[pos] = v; sd[pos] = d; }
    else if (pos - QS < gcap) { gvt[qi][pos - QS] = v; gdt[qi][pos - QS] = d; }
    else *fail = 1;
  }
};

// Append up to two entries per lane to one CTA-wide queue: one shared atomicAdd per warp.
__device__ __forceinline__ void cta_push2(bool w0, int2 v0, u64 d0, bool w1, int2 v1, u64 d1, const HQ2& q, int* cnt,
                                          int gcap, int* fail) {
  const unsigned m0 = __ballot_sync(0xffffffffu, w0);
  const unsigned m1 = __ballot_sync(0xffffffffu, w1);
  if ((m0 | m1) == 0) return;
  const int c0 = __popc(m0);
  int base = 0;
  if (sq::lane_id() == 0) base = atomicAdd(cnt, c0 + __popc(m1));
  base = __shfl_sync(0xffffffffu, base, 0);
  const unsigned lt = sq::lanemask_lt();
  if (w0) q.put(base + __popc(m0 & lt), v0, d0, gcap, fail);
  if (w1) q.put(base + c0 + __popc(m1 & lt), v1, d1, gcap, fail);
}

// SD: the tree's distance row lives in SHARED memory for the whole tree (V x 8 bytes next to the queues: up to ~25.7k
// vertices in B200's 227 KB) and is copied to the caller's row at the end.  A lone tree per SM -- the device router's
// one reverse tree per destination at every weight refresh -- is bound by its ~330 dependent rounds, and half of a
// round was the two L2 round trips for dist[u] and the atomicMin; with the row in shared memory a
// round waits for the ELL record only.
template <int NW, bool SD>
__global__ void __launch_bounds__(NW * 32, SD ? 1 : 32 / NW) sssp_ell_kernel(const TreeParams p) {
  constexpr int THREADS = NW * 32;
  extern __shared__ __align__(16) u64 s_dist_row[];
  __shared__ int2 s_qv[4][QS];
  __shared__ u64 s_qd[4][QS];
  __shared__ int s_cnt[4];   // one counter per physical queue
  __shared__ int s_fcnt[2];  // far piles
  __shared__ int s_tree, s_fail;
  __shared__ u64 s_min;

  const int tid = threadIdx.x;
  const int lane = tid & 31, warp = tid >> 5;
  const int slot = lane & 3, grp = lane >> 2;
  const int V = p.V;
  const int gcap = p.near_cap, fcap = p.far_cap;
  // per-CTA global scratch: 4 queue tails + 2 far piles, vertex ids then distances
  const size_t per_cta = 4 * (size_t)gcap + 2 * (size_t)fcap;
  u64* gd_base = reinterpret_cast<u64*>(p.q2scratch) + (size_t)blockIdx.x * per_cta;
  int2* gv_base = reinterpret_cast<int2*>(p.q2v) + (size_t)blockIdx.x * per_cta;
  __shared__ int2* s_gv[6];   // global parts: queues 0..3, far piles 4..5
  __shared__ u64* s_gd[6];
  if (tid < 6) {
    const size_t off = tid < 4 ? (size_t)tid * gcap : 4 * (size_t)gcap + (size_t)(tid - 4) * fcap;
    s_gv[tid] = gv_base + off;
    s_gd[tid] = gd_base + off;
  }
  __syncthreads();
  auto Q = [&](int i) -> HQ2 { return HQ2{s_qv[i], s_qd[i], s_gv, s_gd, i}; };
  auto FARQ = [&](int i) -> HQ2 { return HQ2{nullptr, nullptr, s_gv, s_gd, 4 + i}; };  // global only: positions start at QS

  int st_rounds = 0, st_relax = 0, st_trans = 0;  // per launch; 32 bits are plenty per thread

  for (;;) {
    if (tid == 0) s_tree = atomicAdd(p.tree_counter, 1);
    __syncthreads();
    const int tree = s_tree;
    if (tree >= p.n_trees) break;

    u64* dist_g = p.dist_out ? p.dist_out + (size_t)tree * V : p.dist_scratch + (size_t)blockIdx.x * V;
    u64* dist = SD ? s_dist_row : dist_g;
    auto LD = [&](const u64* q) -> u64 { return SD ? *reinterpret_cast<const volatile u64*>(q) : sq::ld_cg_u64(q); };
    const int src = p.origins[tree];
    const bool valid = (src >= 0 && src < V);

    {  // dist rows are 8-byte aligned; 16-byte stores on the aligned middle
      const int head = (int)(((uintptr_t)dist >> 3) & 1);
      if (tid == 0 && head && V > 0) dist[0] = DIST_INF_BITS;
      const int pairs = (V - head) >> 1;
      ulonglong2* d2 = reinterpret_cast<ulonglong2*>(dist + head);
      const ulonglong2 inf2 = make_ulonglong2(DIST_INF_BITS, DIST_INF_BITS);
      for (int i = tid; i < pairs; i += THREADS) d2[i] = inf2;
      if (tid == 0 && ((V - head) & 1)) dist[V - 1] = DIST_INF_BITS;
    }
    if (tid < 4) s_cnt[tid] = 0;
    if (tid < 2) s_fcnt[tid] = 0;
    if (tid == 0) s_fail = 0;
    __syncthreads();
    if (tid == 0 && valid) {
      dist[src] = 0ull;
      s_qv[0][0] = make_int2(src, -1);
      s_qd[0][0] = 0ull;
      s_cnt[0] = 1;
    }

    double T = p.delta;          // current bucket: d < T; next bucket: T <= d < T + delta
    int cur = 0, nxt = 1, spare = 2, bkt = 3, fcur = 0;

    while (valid) {
      __syncthreads();
      const int n = s_cnt[cur];
      if (s_fail) break;
      if (tid == 0) s_cnt[spare] = 0;
      if (n == 0) {
        // ---- bucket transition ---------------------------------------------------------------------
        bool done = false;
        for (;;) {
          const int nb = s_cnt[bkt];
          const int nf = min(s_fcnt[fcur], fcap);
          if (nb == 0 && nf == 0) { done = true; break; }
          bool advanced = false;
          if (nb > 0) {  // the next bucket becomes the current queue as it is
            T = advance(T, p.delta);
            const int t = cur; cur = bkt; bkt = t;
            advanced = true;
          }
          if (nf > 0) {
            // split the far pile against the new next-bucket range [T, T + delta)
            __syncthreads();
            if (tid == 0) { s_min = DIST_INF_BITS; s_fcnt[fcur ^ 1] = 0; }
            __syncthreads();
            const double T2 = advance(T, p.delta);
            u64 lmin = DIST_INF_BITS;
            const int2* fiv = s_gv[4 + fcur];
            const u64* fid = s_gd[4 + fcur];
            const HQ2 qb = Q(bkt);
            const HQ2 qf = FARQ(fcur ^ 1);
            for (int i0 = 0; i0 < nf; i0 += THREADS) {
              const int i = i0 + tid;
              bool to_b = false, to_f = false;
              int2 v = make_int2(0, -1);
              u64 d = 0ull;
              if (i < nf) {
                v = ld_cg_int2(&fiv[i]);
                d = sq::ld_cg_u64(&fid[i]);
                if (LD(&dist[v.x]) == d) {  // else stale: a newer entry exists
                  if (__longlong_as_double((long long)d) < T2) to_b = true;
                  else { to_f = true; lmin = d < lmin ? d : lmin; }
                }
              }
              cta_push2(to_b, v, d, false, v, 0ull, qb, &s_cnt[bkt], gcap, &s_fail);
              // far pile: positions start at QS in HQ2 terms so that put() takes its global branch
              {
                const unsigned m = __ballot_sync(0xffffffffu, to_f);
                if (m) {
                  int base = 0;
                  if (lane == 0) base = atomicAdd(&s_fcnt[fcur ^ 1], __popc(m));
                  base = __shfl_sync(0xffffffffu, base, 0);
                  if (to_f) qf.put(QS + base + __popc(m & sq::lanemask_lt()), v, d, fcap, &s_fail);
                }
              }
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
              const u64 other = __shfl_xor_sync(0xffffffffu, lmin, o);
              lmin = other < lmin ? other : lmin;
            }
            if (lane == 0 && lmin != DIST_INF_BITS) atomicMin(&s_min, lmin);
            __syncthreads();
            fcur ^= 1;
            ++st_trans;
            if (!advanced && s_cnt[bkt] == 0 && s_fcnt[fcur] > 0 && !s_fail) {
              // nothing within reach: move the window so that the smallest far distance opens the next bucket
              T = __longlong_as_double((long long)s_min);
              continue;
            }
          }
          if (advanced || s_fail) break;
        }
        if (done) break;
        continue;
      }

      // ---- one relaxation round: 4 lanes per frontier vertex, 16 vertices per warp iteration -------
      const HQ2 qin = Q(cur), qn = Q(nxt), qb = Q(bkt);
      const HQ2 qf = FARQ(fcur);
      const double T2 = advance(T, p.delta);
      // SD (a lone tree on an SM: most of the 32 warps have nothing to do in a round): eight vertices per warp instead
      // of sixteen, so that a wavefront of ~150 vertices is spread over twice as many warps and the round's critical
      // path -- one warp's instruction stream -- is half as long (J = 1: the j = 1 halves below compile away).
      constexpr int J = SD ? 1 : 2;
      for (int base = warp * 8 * J; base < n; base += NW * 8 * J) {
        int u[2], from[2];
        u64 dub[2];
        bool act[2];
        EdgeRec rec[2];
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          const int idx = base + j * 8 + grp;
          act[j] = j < J && idx < n;
          int2 ent = make_int2(0, -1);
          dub[j] = 0ull;
          if (act[j]) qin.get(idx, ent, dub[j]);
          u[j] = ent.x; from[j] = ent.y;
        }
        u64 dnow[2];
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          rec[j].w = 0.0; rec[j].v = -1; rec[j].pad = 0;
          dnow[j] = 0ull;
          if (act[j]) {
            dnow[j] = LD(&dist[u[j]]);
            const int4 raw = __ldg(reinterpret_cast<const int4*>(&p.ell[(size_t)u[j] * 4 + slot]));
            rec[j].w = __hiloint2double(raw.y, raw.x);
            rec[j].v = raw.z;
            rec[j].pad = raw.w;
          }
        }
        u64 ndb[2], old[2];
        bool has[2];
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          act[j] = act[j] && dnow[j] == dub[j];
          has[j] = act[j] && rec[j].v >= 0 && rec[j].v != from[j];
          ndb[j] = (u64)__double_as_longlong(__longlong_as_double((long long)dub[j]) + rec[j].w);
          old[j] = 0ull;
        }
#pragma unroll
        for (int j = 0; j < 2; ++j)
          if (has[j]) old[j] = atomicMin(&dist[rec[j].v], ndb[j]);
        bool imp[2];
        double nd[2];
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          imp[j] = has[j] && ndb[j] < old[j];
          nd[j] = __longlong_as_double((long long)ndb[j]);
          st_relax += has[j];
        }
        const int2 e0 = make_int2(rec[0].v, u[0]), e1 = make_int2(rec[1].v, u[1]);
        if (SD) {
          // one vertex per lane group: both queues' appends in one go, their two counters bumped by two lanes at once
          const bool pn = imp[0] && nd[0] < T, pb = imp[0] && nd[0] >= T && nd[0] < T2;
          const unsigned mn = __ballot_sync(0xffffffffu, pn), mb = __ballot_sync(0xffffffffu, pb);
          if (mn | mb) {
            int base = 0;
            if (lane == 0 && mn) base = atomicAdd(&s_cnt[nxt], __popc(mn));
            if (lane == 1 && mb) base = atomicAdd(&s_cnt[bkt], __popc(mb));
            const int bn = __shfl_sync(0xffffffffu, base, 0), bb = __shfl_sync(0xffffffffu, base, 1);
            const unsigned lt = sq::lanemask_lt();
            if (pn) qn.put(bn + __popc(mn & lt), e0, ndb[0], gcap, &s_fail);
            if (pb) qb.put(bb + __popc(mb & lt), e0, ndb[0], gcap, &s_fail);
          }
        } else {
          cta_push2(imp[0] && nd[0] < T, e0, ndb[0], imp[1] && nd[1] < T, e1, ndb[1], qn, &s_cnt[nxt], gcap, &s_fail);
          cta_push2(imp[0] && nd[0] >= T && nd[0] < T2, e0, ndb[0], imp[1] && nd[1] >= T && nd[1] < T2, e1, ndb[1], qb,
                    &s_cnt[bkt], gcap, &s_fail);
        }
        if (__any_sync(0xffffffffu, (imp[0] && nd[0] >= T2) || (imp[1] && nd[1] >= T2))) {
          const bool f0 = imp[0] && nd[0] >= T2, f1 = imp[1] && nd[1] >= T2;
          const unsigned m0 = __ballot_sync(0xffffffffu, f0), m1 = __ballot_sync(0xffffffffu, f1);
          int fb = 0;
          if (lane == 0) fb = atomicAdd(&s_fcnt[fcur], __popc(m0) + __popc(m1));
          fb = __shfl_sync(0xffffffffu, fb, 0);
          if (f0) qf.put(QS + fb + __popc(m0 & sq::lanemask_lt()), e0, ndb[0], fcap, &s_fail);
          if (f1) qf.put(QS + fb + __popc(m0) + __popc(m1 & sq::lanemask_lt()), e1, ndb[1], fcap, &s_fail);
        }
        // edges beyond the fourth of a vertex (rare on road networks)
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          const int my_pad = act[j] ? (rec[j].pad >> 3) : 0;
          if (!__any_sync(0xffffffffu, my_pad != 0)) continue;
          const int ovf_b = __shfl_sync(0xffffffffu, my_pad, lane & ~3);
          const int ovf_n = __shfl_sync(0xffffffffu, my_pad, (lane & ~3) + 1);
          const int max_o = __reduce_max_sync(0xffffffffu, ovf_n);
          const double du = __longlong_as_double((long long)dub[j]);
          for (int k = slot; k - slot < max_o; k += 4) {
            const bool h2 = act[j] && k < ovf_n;
            int2 v2 = make_int2(0, u[j]);
            u64 nb2 = DIST_INF_BITS, old2 = 0ull;
            if (h2) {
              const EdgeRec er = ld_edge(&p.ovf[ovf_b + k]);
              v2.x = er.v;
              nb2 = (u64)__double_as_longlong(du + er.w);
              old2 = atomicMin(&dist[v2.x], nb2);
            }
            const bool imp2 = h2 && nb2 < old2;
            const double nd2 = __longlong_as_double((long long)nb2);
            cta_push2(imp2 && nd2 < T, v2, nb2, false, v2, 0ull, qn, &s_cnt[nxt], gcap, &s_fail);
            cta_push2(imp2 && nd2 >= T && nd2 < T2, v2, nb2, false, v2, 0ull, qb, &s_cnt[bkt], gcap, &s_fail);
            const bool f2 = imp2 && nd2 >= T2;
            const unsigned m2 = __ballot_sync(0xffffffffu, f2);
            if (m2) {
              int fb = 0;
              if (lane == 0) fb = atomicAdd(&s_fcnt[fcur], __popc(m2));
              fb = __shfl_sync(0xffffffffu, fb, 0);
              if (f2) qf.put(QS + fb + __popc(m2 & sq::lanemask_lt()), v2, nb2, fcap, &s_fail);
            }
            st_relax += h2;
          }
        }
      }
      { const int t = cur; cur = nxt; nxt = spare; spare = t; }
      ++st_rounds;
    }
    __syncthreads();
    if (SD && (p.dist_out || p.dist_scratch)) {
      // the finished row goes out in one coalesced pass (the predecessor kernel and the caller read it from there)
      for (int i = tid; i < V; i += THREADS) dist_g[i] = dist[i];
    }

    if (s_fail) {
      // queue capacity exceeded: hand the tree to the bitmap kernel (its queues cannot overflow)
      if (tid == 0) p.fail_list[atomicAdd(p.fail_count, 1)] = tree;
    }
    __syncthreads();
  }

  if (p.stats) {
    long long r = st_relax;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) r += __shfl_xor_sync(0xffffffffu, r, o);
    if (lane == 0) atomicAdd((u64*)&p.stats[2], (u64)r);
    if (tid == 0) {
      atomicAdd((u64*)&p.stats[1], (u64)st_rounds);
      atomicAdd((u64*)&p.stats[3], (u64)st_trans);
    }
  }
}

// -------------------------------------------------------------------------------------------------
// Warp kernel: one tree per WARP, no CTA barriers, no shared-memory atomics.
//
// The CTA-per-tree kernels spend their time on per-round overhead at low lane fill: a road-network wavefront
// holds a few hundred vertices, so 256 threads find 1-2 passes of work between barriers (ncu, r1b: 27 %
// barrier stalls, 40 % issue utilisation, 1.33 M warp instructions per tree).  Here a warp owns the tree:
//  * queue state (head / tail / bucket count) lives in registers and is warp-uniform; pushes are a ballot
//    and a popc, visibility is a __syncwarp;
//  * the current bucket is a FIFO ring in shared memory (pops 8 x ILP vertices per step, four lanes per
//    vertex = one 64-byte ELL record), improvements below T are appended to the same ring, those in
//    [T, T + delta) to the next-bucket buffer, which becomes the ring at the transition; the rest go to a far
//    pile in global memory (closed links at 1e8);
//  * a queue entry is 4 bytes: vertex id | slot of the edge straight back (skipped: it cannot improve) | hash
//    of the distance the vertex had when pushed.  An entry whose hash no longer matches dist[vertex] is stale
//    (a newer entry exists).  A hash collision only makes a stale entry look valid, and valid entries are
//    relaxed with the distance read from memory, so collisions cost a duplicate pass, never a wrong result;
//  * trees whose wavefront outgrows the ring are handed to the fallback kernel through fail_list.
// Latency (queue -> dist/ELL loads -> L2 atomicMin) is hidden by 16 resident warps per SM, each on its own tree,
// and by ILP independent groups of eight vertices per step.  What bounds the kernel is not HBM but the L1TEX
// wavefront pipe: every relaxation is an 8-byte atomic on its own cache line (~2 wavefronts per relaxation all
// told), see DESIGN.md section 2.
// -------------------------------------------------------------------------------------------------
__device__ __forceinline__ unsigned dist_hash(u64 d) {
  return ((unsigned)d * 0x9E3779B1u) ^ ((unsigned)(d >> 32) * 0x85EBCA6Bu);
}

// GROUPED: the eight warps of a CTA run the eight trees of one group of neighbouring origins (sorted by the host along
// the locality order) on INTERLEAVED distance rows, dist[v][8] in internal numbering (p.gdist), so that the sectors
// and lines the sibling trees touch are the same ones: a quarter of the L2 footprint per tree, and the two rows of a
// 128-byte line belong to adjacent vertices.  group_finalize_kernel un-interleaves the rows afterwards.  The ELL is the
// group form (sp_group.cuh): pad bit 1 = empty slot, bits 3..5 = slot of the edge straight back, bits 6.. = overflow.
template <int ILP, bool GROUPED>
__global__ void __launch_bounds__(GROUPED ? 256 : 128) sssp_warp_kernel(const TreeParams p) {
  extern __shared__ unsigned s_rings[];
  __shared__ int s_grp;
  const unsigned FULL = 0xffffffffu;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int slot = lane & 3, grp = lane >> 2;
  const unsigned lt = sq::lanemask_lt();
  const int V = p.V;
  const int C = p.ring_cap, Cm = C - 1;
  const int vbits = p.vbits;
  const unsigned vmask = (1u << vbits) - 1u;
  const unsigned himask = ~((1u << (vbits + 3)) - 1u);
  const int fcap = p.wfar_cap;
  const int gwarp = blockIdx.x * (blockDim.x >> 5) + warp;
  unsigned* ring_a = s_rings + (size_t)warp * 2 * C;
  unsigned* far_base = p.wfar + (size_t)gwarp * 2 * fcap;
  const double delta = p.delta;

  int st_steps = 0, st_relax = 0, st_trans = 0;

  for (;;) {
    int tree = 0, src = -1;
    u64* dist;
    if (GROUPED) {
      __syncthreads();                       // every warp is done with the previous group
      if (threadIdx.x == 0) s_grp = atomicAdd(p.tree_counter, 1);
      __syncthreads();
      const int gidx = s_grp;
      if (gidx >= p.n_trees) break;          // n_trees = number of groups here
      dist = p.gdist + (size_t)gidx * V * 8;
      {  // rows = DBL_MAX (graph.cc:178), all eight warps together
        ulonglong2* d2 = reinterpret_cast<ulonglong2*>(dist);
        const ulonglong2 inf2 = make_ulonglong2(DIST_INF_BITS, DIST_INF_BITS);
        const int n2 = V * 4;
#pragma unroll 4
        for (int i = threadIdx.x; i < n2; i += 256) d2[i] = inf2;
      }
      tree = __ldg(&p.tord[gidx * 8 + warp]);
      if (tree >= 0) {
        const int o = __ldg(&p.origins[tree]);
        if (o >= 0 && o < V) src = __ldg(&p.perm[o]);
      }
      dist += warp;                          // this warp's column
      __syncthreads();
    } else {
      if (lane == 0) tree = atomicAdd(p.tree_counter, 1);
      tree = __shfl_sync(FULL, tree, 0);
      if (tree >= p.n_trees) break;
      dist = p.dist_out ? p.dist_out + (size_t)tree * V : p.dist_scratch + (size_t)gwarp * V;
      src = p.origins[tree];
      {  // distance row = DBL_MAX (graph.cc:178); rows are 8-byte aligned, 16-byte stores on the aligned middle
        const int head = (int)(((uintptr_t)dist >> 3) & 1);
        if (lane == 0 && head && V > 0) dist[0] = DIST_INF_BITS;
        const int pairs = (V - head) >> 1;
        ulonglong2* d2 = reinterpret_cast<ulonglong2*>(dist + head);
        const ulonglong2 inf2 = make_ulonglong2(DIST_INF_BITS, DIST_INF_BITS);
#pragma unroll 4
        for (int i = lane; i < pairs; i += 32) d2[i] = inf2;
        if (lane == 0 && ((V - head) & 1)) dist[V - 1] = DIST_INF_BITS;
      }
      __syncwarp();
    }
    const bool valid_src = (src >= 0 && src < V);
    if (!valid_src) continue;
    // element v of this tree's distance column
    auto DA = [&](unsigned v) -> u64* { return GROUPED ? dist + ((size_t)v << 3) : dist + v; };

    unsigned* cur = ring_a;
    unsigned* bkt = ring_a + C;
    int ch = 0, ct = 1, bt = 0;    // ring head / tail (monotonic), next-bucket count
    int nf = 0, fcur = 0;          // far pile
    bool fail = false;
    if (lane == 0) {
      *DA((unsigned)src) = 0ull;
      cur[0] = (unsigned)src | (7u << vbits) | (dist_hash(0ull) & himask);
    }
    __syncwarp();
    double T = delta, T2 = advance(delta, delta);
    // Classifies one candidate per lane and appends it to the ring / next bucket / far pile.  Convergent.
    auto push = [&](bool imp, u64 ndb, unsigned ent) {
      const double nd = __longlong_as_double((long long)ndb);
      const bool nr = imp && nd < T;
      const bool md = imp && !nr && nd < T2;
      const bool fr = imp && !nr && !md;
      const unsigned mn = __ballot_sync(FULL, nr), mm = __ballot_sync(FULL, md), mf = __ballot_sync(FULL, fr);
      if (nr) cur[(ct + __popc(mn & lt)) & Cm] = ent;
      ct += __popc(mn);
      if (md) {
        const int pos = bt + __popc(mm & lt);
        if (pos < C) bkt[pos] = ent;
      }
      bt += __popc(mm);
      if (mf) {
        if (fr) {
          const int pos = nf + __popc(mf & lt);
          if (pos < fcap) far_base[(size_t)fcur * fcap + pos] = ent;
        }
        nf += __popc(mf);
      }
    };

    for (;;) {
      if (ch == ct) {
        // ---- bucket transition (warp-uniform) ---------------------------------------------------------
        if (bt == 0) {
          if (nf == 0) break;  // tree finished
          // nothing within reach: open the next bucket at the smallest live far distance
          u64 lmin = DIST_INF_BITS;
          const unsigned* fp = far_base + (size_t)fcur * fcap;
          for (int i = lane; i < nf; i += 32) {
            const unsigned e = (unsigned)sq::ld_cg_s32(reinterpret_cast<const int*>(fp + i));
            const u64 d = sq::ld_cg_u64(DA(e & vmask));
            if (((dist_hash(d) ^ e) & himask) == 0 && d < lmin) lmin = d;
          }
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) {
            const u64 other = __shfl_xor_sync(FULL, lmin, o);
            lmin = other < lmin ? other : lmin;
          }
          if (lmin == DIST_INF_BITS) break;  // only stale entries were left
          T2 = __longlong_as_double((long long)lmin);
        }
        { unsigned* t = cur; cur = bkt; bkt = t; }
        ch = 0; ct = bt; bt = 0;
        T = T2; T2 = advance(T, delta);
        ++st_trans;
        if (nf > 0) {
          // split the far pile against the new next-bucket range [T, T2)
          const unsigned* fp = far_base + (size_t)fcur * fcap;
          unsigned* fq = far_base + (size_t)(fcur ^ 1) * fcap;
          int nf2 = 0;
          for (int i0 = 0; i0 < nf; i0 += 32) {
            const int i = i0 + lane;
            unsigned e = 0u;
            bool live = false;
            double d = 0.0;
            if (i < nf) {
              e = (unsigned)sq::ld_cg_s32(reinterpret_cast<const int*>(fp + i));
              const u64 db = sq::ld_cg_u64(DA(e & vmask));
              live = ((dist_hash(db) ^ e) & himask) == 0;
              d = __longlong_as_double((long long)db);
            }
            const bool to_b = live && d < T2, to_f = live && !to_b;
            const unsigned mb = __ballot_sync(FULL, to_b), mf = __ballot_sync(FULL, to_f);
            if (to_b) {
              const int pos = bt + __popc(mb & lt);
              if (pos < C) bkt[pos] = e;
            }
            bt += __popc(mb);
            if (to_f) fq[nf2 + __popc(mf & lt)] = e;  // nf2 <= i0 <= fcap: cannot overflow
            nf2 += __popc(mf);
          }
          nf = nf2;
          fcur ^= 1;
          if (bt > C) { fail = true; break; }
        }
        __syncwarp();
        continue;
      }

      // ---- one step: 8 x ILP frontier vertices, four lanes (= one 64-byte ELL record) per vertex -------
      ++st_steps;
      unsigned ent[ILP];
      bool act[ILP];
#pragma unroll
      for (int j = 0; j < ILP; ++j) {
        const int idx = ch + j * 8 + grp;
        act[j] = idx < ct;
        ent[j] = act[j] ? cur[idx & Cm] : 0u;
      }
      ch = min(ct, ch + 8 * ILP);
      u64 dnow[ILP];
      int4 raw[ILP];
#pragma unroll
      for (int j = 0; j < ILP; ++j) {
        dnow[j] = 0ull;
        raw[j] = make_int4(0, 0, -1, 7);
        if (act[j]) {
          const unsigned u = ent[j] & vmask;
          dnow[j] = sq::ld_cg_u64(DA(u));
          raw[j] = __ldg(reinterpret_cast<const int4*>(&p.ell[(size_t)u * 4 + slot]));   // GROUPED: p.ell = the group ELL
        }
      }
      bool ok[ILP], has[ILP];
      u64 ndb[ILP], old[ILP];
#pragma unroll
      for (int j = 0; j < ILP; ++j) {
        ok[j] = act[j] && ((dist_hash(dnow[j]) ^ ent[j]) & himask) == 0;
        has[j] = ok[j] && (GROUPED ? !(raw[j].w & 2) : raw[j].z >= 0) && slot != (int)((ent[j] >> vbits) & 7u);
        ndb[j] = (u64)__double_as_longlong(__longlong_as_double((long long)dnow[j]) + __hiloint2double(raw[j].y, raw[j].x));
        old[j] = 0ull;
      }
      // (measured and dropped, profiles/r2c_group_sweep.jsonl: reading the head's distance first and sending a RED.MIN
      //  without return value only where the candidate is smaller -- 46 % of the relaxations -- is 14 % SLOWER than the
      //  atomic with return; read + atomicMin where smaller is 5 % slower.  The kernel is not bound by the L2 atomic units.)
#pragma unroll
      for (int j = 0; j < ILP; ++j)
        if (has[j]) old[j] = atomicMin(DA((unsigned)raw[j].z), ndb[j]);
#pragma unroll
      for (int j = 0; j < ILP; ++j) {
        st_relax += has[j];
        const unsigned back = GROUPED ? (unsigned)(raw[j].w >> 3) & 7u : (unsigned)(raw[j].w & 7);
        const unsigned ne = (unsigned)raw[j].z | (back << vbits) | (dist_hash(ndb[j]) & himask);
        push(has[j] && ndb[j] < old[j], ndb[j], ne);
      }
      // edges beyond the fourth of a vertex (rare on road networks)
#pragma unroll
      for (int j = 0; j < ILP; ++j) {
        const int my = ok[j] ? (raw[j].w >> (GROUPED ? 6 : 3)) : 0;
        if (!__any_sync(FULL, slot == 1 && my != 0)) continue;
        const int ovf_b = __shfl_sync(FULL, my, lane & ~3);
        const int ovf_n = __shfl_sync(FULL, my, (lane & ~3) + 1);
        const int max_o = __reduce_max_sync(FULL, ovf_n);
        for (int k = slot; k - slot < max_o; k += 4) {
          const bool h2 = ok[j] && k < ovf_n;
          unsigned ne = 0u;
          u64 nb2 = DIST_INF_BITS, old2 = 0ull;
          if (h2) {
            const EdgeRec er = ld_edge(&p.ovf[ovf_b + k]);
            nb2 = (u64)__double_as_longlong(__longlong_as_double((long long)dnow[j]) + er.w);
            old2 = atomicMin(DA((unsigned)er.v), nb2);
            ne = (unsigned)er.v | (7u << vbits) | (dist_hash(nb2) & himask);
          }
          st_relax += h2;
          push(h2 && nb2 < old2, nb2, ne);
        }
      }
      if (ct - ch > C || bt > C || nf > fcap) { fail = true; break; }
      __syncwarp();
    }
    __syncwarp();
    if (fail && lane == 0) p.fail_list[atomicAdd(p.fail_count, 1)] = tree;  // rerun by the fallback kernel
  }

  if (p.stats) {
    long long r = st_relax;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) r += __shfl_xor_sync(FULL, r, o);
    if (lane == 0) {
      atomicAdd((u64*)&p.stats[2], (u64)r);
      atomicAdd((u64*)&p.stats[1], (u64)st_steps);
      atomicAdd((u64*)&p.stats[3], (u64)st_trans);
    }
  }
}

#include "sp_group.cuh"

// Canonical predecessors for a batch of finished trees: parent[v] = min{u : fl(d[u] + w(u,v)) == d[v]}, -1 for
// the origin and for unreached vertices.  One thread per vertex; its (up to four) in-edges come from a
// slot-major copy of the reverse ELL (coalesced) and stay in registers while blockIdx.y strides over trees, so
// the graph is read once per ~16 trees and every dist / parent access of a warp is a run of 32 vertices.
__global__ void __launch_bounds__(256) sssp_parent_kernel(int V, int n_trees, const int* __restrict__ rs_v,
                                                          const double* __restrict__ rs_w, const EdgeRec* __restrict__ rell,
                                                          const EdgeRec* __restrict__ rovf, int has_ovf,
                                                          const int* __restrict__ origins, const u64* __restrict__ dist_all,
                                                          int* __restrict__ parent_all, int codes) {
  const int v = blockIdx.x * 256 + threadIdx.x;
  if (v >= V) return;
  int nv[4];
  double nw[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    nv[j] = __ldg(&rs_v[(size_t)j * V + v]);
    nw[j] = __ldg(&rs_w[(size_t)j * V + v]);
  }
  int ovf_b = 0, ovf_n = 0;
  if (has_ovf) {
    ovf_b = __ldg(&rell[(size_t)v * 4 + 0].pad) >> 3;
    ovf_n = __ldg(&rell[(size_t)v * 4 + 1].pad) >> 3;
  }
  for (int tree = blockIdx.y; tree < n_trees; tree += gridDim.y) {
    const u64* dist = dist_all + (size_t)tree * V;
    const u64 dvb = __ldg(&dist[v]);
    int best = INT_MAX, slot = 255;
    if (dvb != DIST_INF_BITS) {
      u64 dub[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) dub[j] = nv[j] >= 0 ? __ldg(&dist[nv[j]]) : DIST_INF_BITS;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        if (dub[j] != DIST_INF_BITS) {
          const double cand = __longlong_as_double((long long)dub[j]) + nw[j];
          if ((u64)__double_as_longlong(cand) == dvb && nv[j] < best) { best = nv[j]; slot = j; }
        }
      }
      for (int k = 0; k < ovf_n; ++k) {
        const EdgeRec er = ld_edge(&rovf[ovf_b + k]);
        const u64 du = __ldg(&dist[er.v]);
        if (du != DIST_INF_BITS) {
          const double cand = __longlong_as_double((long long)du) + er.w;
          if ((u64)__double_as_longlong(cand) == dvb && er.v < best) { best = er.v; slot = min(4 + k, 254); }
        }
      }
    }
    const int src = __ldg(&origins[tree]);
    const bool none = best == INT_MAX || v == src;
    if (codes) reinterpret_cast<unsigned char*>(parent_all)[(size_t)tree * V + v] = none ? (unsigned char)255 : (unsigned char)slot;
    else parent_all[(size_t)tree * V + v] = none ? -1 : best;
  }
}

// Expands one-byte predecessor codes (sp_set_parent_format 1: index of the in-edge that attains the distance,
// 255 = none) back into vertex ids.
__global__ void __launch_bounds__(256) decode_parents_kernel(int V, int n_trees, const int* __restrict__ rs_v,
                                                             const EdgeRec* __restrict__ rell, const EdgeRec* __restrict__ rovf,
                                                             const unsigned char* __restrict__ codes, int* __restrict__ parent_all) {
  const int v = blockIdx.x * 256 + threadIdx.x;
  if (v >= V) return;
  int nv[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) nv[j] = __ldg(&rs_v[(size_t)j * V + v]);
  for (int tree = blockIdx.y; tree < n_trees; tree += gridDim.y) {
    const int c = codes[(size_t)tree * V + v];
    int par = -1;
    if (c < 4) par = c == 0 ? nv[0] : c == 1 ? nv[1] : c == 2 ? nv[2] : nv[3];
    else if (c != 255) par = rovf[(__ldg(&rell[(size_t)v * 4].pad) >> 3) + (c - 4)].v;
    parent_all[(size_t)tree * V + v] = par;
  }
}

// One-byte codes -> two vertices per byte (low nibble = even vertex), 255 -> 15.  The caller has checked that every
// code is below 15 (largest in-degree <= 15).  One thread per output byte.
__global__ void __launch_bounds__(256) pack_codes4_kernel(int V, int n_trees, const unsigned char* __restrict__ codes,
                                                          unsigned char* __restrict__ packed) {
  const size_t vh = ((size_t)V + 1) / 2;
  const size_t i = (size_t)blockIdx.x * 256 + threadIdx.x;
  if (i >= vh * (size_t)n_trees) return;
  const size_t tree = i / vh, b = i - tree * vh;
  const unsigned char* row = codes + tree * (size_t)V;
  const unsigned lo = row[2 * b];
  const unsigned hi = 2 * b + 1 < (size_t)V ? row[2 * b + 1] : 255u;
  packed[i] = (unsigned char)((lo >= 15u ? 15u : lo) | ((hi >= 15u ? 15u : hi) << 4));
}

// Device-side weight refresh: weights[i] belongs to edge i of the insertion log; NaN keeps the current weight.
// Every copy of the weight (CSR/CSC records, ELL + overflow records, slot-major arrays) is rewritten in place.
struct WeightSlots {
  const int* adj; const int* radj;    // CSR / CSC record index, -1 = edge not in the graph
  const int* ell; const int* rell;    // >= 0: index into the ELL array; < 0: -(overflow index + 1)
  const int* gell; const int* grell;  // same for the group kernel's ELL pair (internal numbering)
};
__global__ void __launch_bounds__(256) scatter_weights_kernel(int n, int V, const double* __restrict__ w, WeightSlots ws,
                                                              EdgeRec* adj, EdgeRec* radj, EdgeRec* ell, EdgeRec* ovf,
                                                              EdgeRec* rell, EdgeRec* rovf, double* fs_w, double* rs_w,
                                                              EdgeRec* gell, EdgeRec* govf, EdgeRec* grell, EdgeRec* grovf,
                                                              double* __restrict__ wlog) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i >= n) return;
  const double x = w[i];
  if (x != x) return;
  wlog[i] = x;   // the edge log in insertion order; the host copy catches up on demand (g_pull_host_log)
  const int a = ws.adj[i];
  if (a < 0) return;
  adj[a].w = x;
  radj[ws.radj[i]].w = x;
  const int e = ws.ell[i];
  if (e >= 0) { ell[e].w = x; fs_w[(size_t)(e & 3) * V + (e >> 2)] = x; }
  else ovf[-(e + 1)].w = x;
  const int r = ws.rell[i];
  if (r >= 0) { rell[r].w = x; rs_w[(size_t)(r & 3) * V + (r >> 2)] = x; }
  else rovf[-(r + 1)].w = x;
  const int ge = ws.gell[i];
  if (ge >= 0) gell[ge].w = x; else govf[-(ge + 1)].w = x;
  const int gr = ws.grell[i];
  if (gr >= 0) grell[gr].w = x; else grovf[-(gr + 1)].w = x;
}

// =================================================================================================
// Host side
// =================================================================================================

struct SqGraph {
  bool directed = true;
  unsigned nvertices = UINT_MAX;  // graph.h:127 default
  // edge log in insertion order (after the undirected swap)
  std::vector<int> es, ee;
  std::vector<double> ew;
  std::vector<char> alive;
  std::map<std::pair<int, int>, int> key2edge;  // ordered like the reference's std::map (writegraph order)
  long long ignored_updates = 0;

  // host CSR bookkeeping
  bool topo_dirty = true, w_dirty = true;
  std::vector<int> edge_slot, edge_rslot;  // edge id -> position in adj / radj
  std::vector<EdgeRec> h_adj, h_radj;
  std::vector<int> h_rp, h_rrp;
  std::vector<double> w_stage;
  int V = 0, E = 0;

  // device
  int device = -1;
  int* d_row_ptr = nullptr;
  int* d_rrow_ptr = nullptr;
  EdgeRec* d_adj = nullptr;
  EdgeRec* d_radj = nullptr;
  int* d_counter = nullptr;
  long long* d_stats = nullptr;
  u64* d_dist_scratch = nullptr;
  int* d_qscratch = nullptr;
  int scratch_grid = 0, scratch_V = 0;
  int* d_origins = nullptr;
  int origins_cap = 0;
  u64* d_dist_buf = nullptr;
  int* d_parent_buf = nullptr;
  size_t dist_buf_elems = 0, parent_buf_elems = 0;
  // ELL form (four edges per vertex + overflow), forward and reverse
  std::vector<EdgeRec> h_ell, h_ovf, h_rell, h_rovf;
  EdgeRec* d_ell = nullptr; EdgeRec* d_ovf = nullptr; EdgeRec* d_rell = nullptr; EdgeRec* d_rovf = nullptr;
  size_t ovf_cap = 0, rovf_cap = 0;
  int* d_rs_v = nullptr;       // slot-major reverse ELL for the predecessor kernel
  double* d_rs_w = nullptr;
  int* d_fs_v = nullptr;       // slot-major forward ELL (predecessors of trees on the reversed graph)
  double* d_fs_w = nullptr;
  int* d_wslots = nullptr;     // 4 x n_log: where the weight of edge i lives (scatter_weights_kernel)
  // sp_set_weights_device leaves the newest weights on the device only (no host round trip inside a simulation's
  // refresh): d_wlog is the edge log's weight column while dev_leads; every host reader / editor of the log pulls first
  unsigned char* d_pack_buf = nullptr;   // sp_sssp_batch_codes4: two chunks of packed codes
  size_t pack_buf_bytes = 0;
  double* d_wlog = nullptr;
  size_t wlog_cap = 0;
  bool dev_leads = false;
  cudaEvent_t ev_wlog = nullptr;   // recorded behind the last device-side refresh (outlives the caller's stream)
  size_t wslots_n = 0;
  int2* d_q2scratch = nullptr;
  int* d_q2v = nullptr;
  u64* d_dist_tmp = nullptr;   // distances of parent-only batches
  size_t dist_tmp_elems = 0;
  int q2_grid = 0, q2_near = 0, q2_far = 0;
  int* d_fail = nullptr;       // [0] count, [1] fallback tree counter, [2..] list
  int fail_cap = 0;
  cudaStream_t s_compute = nullptr, s_copy = nullptr;
  cudaEvent_t ev_done[2] = {nullptr, nullptr}, ev_copied[2] = {nullptr, nullptr};
  int parent_format = 0;       // 0 = int32 vertex ids, 1 = one-byte in-edge codes (device batches only)
  int algo = 3;                // 3 = auto (default), 2 = warp kernel, 0 = ELL kernel, 1 = bitmap kernel (also everyone's fallback)
  int warp_ring = 1024, warp_ilp = 4, warp_ctas_per_sm = 0;
  // auto mode learns from the overflow count of its last warp-kernel batch (read at the start of the next one)
  int auto_ring = 0;           // ring capacity auto mode uses (0 = warp_ring)
  bool auto_warp_unfit = false; // wavefronts of this graph do not fit the largest ring: auto stays on the ELL kernel
  int auto_checks = 0;
  int pending_check_n = 0;      // trees of the last auto warp batch whose overflow count has not been looked at yet
  unsigned* d_wfar = nullptr;
  int wfar_warps = 0, wfar_cap = 0;
  // sp_route_batch: per-pair device arrays, path pool on both sides (the host pool is page-locked)
  int* d_pair = nullptr;        // 3 x pair_cap: destination, tree, path length
  long long* d_pair_off = nullptr;
  double* d_pair_dist = nullptr;
  size_t pair_cap = 0;
  int* d_pool = nullptr; size_t d_pool_cap = 0;
  int* h_pool = nullptr; size_t h_pool_cap = 0;
  double queue_factor = 1.0;   // queue capacities of the ELL kernel: near = factor*V, far = 2*factor*V
  // group kernel (sp_group.cuh): internal numbering along a Hilbert curve over landmark pseudo-coordinates
  std::vector<int> h_perm, h_iperm;      // caller id -> internal id and back
  int* d_perm = nullptr; int* d_iperm = nullptr;
  unsigned char* d_tile_ord = nullptr;   // per 256-vertex tile of internal ids: its vertices in ascending caller id
  std::vector<EdgeRec> h_gell[2], h_govf[2];   // [0] forward, [1] reverse; internal numbering
  EdgeRec* d_gell[2] = {nullptr, nullptr}; EdgeRec* d_govf[2] = {nullptr, nullptr};
  size_t govf_cap[2] = {0, 0};
  int4* d_gheads[2] = {nullptr, nullptr};
  int* d_gkey = nullptr; int* d_tord = nullptr; size_t gkey_cap = 0;
  int* d_ghist = nullptr; size_t ghist_cap = 0;
  u64* d_gdist = nullptr; size_t gdist_groups = 0;
  unsigned* d_gring = nullptr; unsigned* d_gpend = nullptr; unsigned* d_ginp = nullptr;
  int g_warps = 0, g_ring_cap = 0, g_pend_cap = 0, g_words = 0;
  int group_warps_per_sm = 16, group_pb = 16, group_pf = 1;
  cudaDeviceProp prop; int prop_dev = -1;
  int ell_smem_row = 1;                  // auto mode: lone trees keep their distance row in shared memory when it fits
  int group_mode = 0;                    // 0 = grouped warp kernel (a CTA of eight warps per group), 1 = sssp_group_kernel
  int group_ring = 1024;                 // ring entries per tree of the grouped warp kernel
  // Auto mode (algo 3) does NOT take the group path by default: all three kernels on the group layout measured slower
  // than sssp_warp_kernel on config 5 (profiles/r2*_group_sweep.jsonl).  sp_set_group_tuning can
  // lower the threshold; algo 4 forces the path.
  int group_min_trees = INT_MAX;         // auto mode: batches at least this large ...
  int group_min_density = 8;             // ... with at least V / this many origins go to the group path
  size_t group_scratch_bytes = (size_t)24 << 30;   // upper bound of the interleaved distance scratch per chunk

  // tuning
  int threads = 256;
  int ctas_per_sm = 0;
  double delta_factor = 4.0;
  double delta = 1.0;
  int smem_queue = 0;
  long long last_stats[5] = {0, 0, 0, 0, 0};
};

struct SqShortestPath {
  int source = -1;
  std::vector<double> distances;
  std::vector<int> parent;
};

// -------------------------------------------------------------------------------------------------
// Route extraction (sp_route_batch): Agent.get_path needs distance(destin) and route(destin) of its tree, not the
// tree.  One thread per (origin, destination) pair walks the predecessor row from the destination back to the
// origin, once to count and once to write the path in forward order, so only paths cross PCIe.
// -------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128) route_len_kernel(const u64* __restrict__ dist, const int* __restrict__ parent,
                                                        const int* __restrict__ origins, int V, int tree0, int n,
                                                        const int* __restrict__ pdest, const int* __restrict__ ptree,
                                                        int* __restrict__ plen, double* __restrict__ pdist) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n) return;
  const int tl = ptree[k] - tree0;
  const int dest = pdest[k];
  const int src = origins[ptree[k]];
  const size_t row = (size_t)tl * V;
  const u64 d = (src >= 0 && src < V) ? dist[row + dest] : DIST_INF_BITS;
  pdist[k] = __longlong_as_double((long long)d);
  int len = 0;
  if (d != DIST_INF_BITS && dest != src) {
    int v = dest;
    len = 1;
    while (v != src && len <= V) { v = parent[row + v]; ++len; if (v < 0) { len = 0; break; } }
    if (len > V) len = 0;
  }
  plen[k] = len;
}

__global__ void __launch_bounds__(128) route_write_kernel(const int* __restrict__ parent, int V, int tree0, int n,
                                                          const int* __restrict__ pdest, const int* __restrict__ ptree,
                                                          const int* __restrict__ plen, const long long* __restrict__ poff,
                                                          long long pool0, int* __restrict__ pool) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n) return;
  const int len = plen[k];
  if (len == 0) return;
  const size_t row = (size_t)(ptree[k] - tree0) * V;
  int* out = pool + (poff[k] - pool0);
  int v = pdest[k];
  for (int i = len - 1; i >= 0; --i) { out[i] = v; if (i) v = parent[row + v]; }
}

// The device holds newer weights than the host log (sp_set_weights_device): bring the log and the host CSR up to date.
static int g_pull_host_log(SqGraph* g) {
  if (!g->dev_leads) return SQ_OK;
  const size_t n = g->ew.size();
  if (n == 0) { g->dev_leads = false; return SQ_OK; }
  SQ_CUDA_TRY(cudaEventSynchronize(g->ev_wlog));
  SQ_CUDA_TRY(cudaMemcpy(g->ew.data(), g->d_wlog, sizeof(double) * n, cudaMemcpyDeviceToHost));
  g->dev_leads = false;
  for (size_t i = 0; i < n && i < g->edge_slot.size(); ++i) {
    const int a = g->edge_slot[i];
    if (a >= 0) { g->h_adj[a].w = g->ew[i]; g->h_radj[g->edge_rslot[i]].w = g->ew[i]; }
  }
  return SQ_OK;
}

static void g_add_edge(SqGraph* g, int v1, int v2, double w) {
  g_pull_host_log(g);
  if (!g->directed && v1 > v2) std::swap(v1, v2);
  const int id = (int)g->es.size();
  g->es.push_back(v1);
  g->ee.push_back(v2);
  g->ew.push_back(w);
  g->alive.push_back(1);
  g->key2edge[std::make_pair(v1, v2)] = id;  // last insert owns the key
  g->topo_dirty = true;
}

static void g_free_device(SqGraph* g) {
  cudaFree(g->d_row_ptr); cudaFree(g->d_rrow_ptr); cudaFree(g->d_adj); cudaFree(g->d_radj);
  cudaFree(g->d_ell); cudaFree(g->d_rell); cudaFree(g->d_ovf); cudaFree(g->d_rovf);
  cudaFree(g->d_rs_v); cudaFree(g->d_rs_w); g->d_rs_v = nullptr; g->d_rs_w = nullptr;
  cudaFree(g->d_fs_v); cudaFree(g->d_fs_w); g->d_fs_v = nullptr; g->d_fs_w = nullptr;
  cudaFree(g->d_wslots); g->d_wslots = nullptr; g->wslots_n = 0;
  cudaFree(g->d_pack_buf); g->d_pack_buf = nullptr; g->pack_buf_bytes = 0;
  cudaFree(g->d_wlog); g->d_wlog = nullptr; g->wlog_cap = 0; g->dev_leads = false;
  g->d_row_ptr = g->d_rrow_ptr = nullptr; g->d_adj = g->d_radj = nullptr;
  g->d_ell = g->d_rell = g->d_ovf = g->d_rovf = nullptr;
  g->ovf_cap = g->rovf_cap = 0;
  cudaFree(g->d_perm); cudaFree(g->d_iperm); g->d_perm = g->d_iperm = nullptr;
  cudaFree(g->d_tile_ord); g->d_tile_ord = nullptr;
  for (int d = 0; d < 2; ++d) {
    cudaFree(g->d_gell[d]); cudaFree(g->d_govf[d]); cudaFree(g->d_gheads[d]);
    g->d_gell[d] = g->d_govf[d] = nullptr; g->d_gheads[d] = nullptr; g->govf_cap[d] = 0;
  }
}

// Locality order for the group kernel.  Road networks come without usable coordinates at this level (the ABI takes
// an edge list), so the order is built from the graph alone: hop distances from two landmarks a, c (a = a peripheral
// vertex, c = the vertex farthest from both a and the vertex b farthest from a) are pseudo-coordinates; their sum and
// difference are roughly orthogonal on a planar network; vertices are sorted along a Hilbert curve over those two.
static unsigned hilbert_key(unsigned n, unsigned x, unsigned y) {
  unsigned d = 0;
  for (unsigned s = n / 2; s > 0; s /= 2) {
    const unsigned rx = (x & s) ? 1u : 0u, ry = (y & s) ? 1u : 0u;
    d += s * s * ((3u * rx) ^ ry);
    if (ry == 0) {
      if (rx == 1) { x = s - 1 - x; y = s - 1 - y; }
      std::swap(x, y);
    }
  }
  return d;
}

static void locality_order(int V, const std::vector<int>& rp, const std::vector<EdgeRec>& adj, const std::vector<int>& rrp,
                           const std::vector<EdgeRec>& radj, std::vector<int>& perm, std::vector<int>& iperm) {
  perm.assign(V, 0); iperm.assign(V, 0);
  if (V == 0) return;
  std::vector<int> q(V), da(V), db(V), dc(V);
  auto bfs = [&](int src, std::vector<int>& d) {   // hops on the undirected graph; unreached vertices keep -1
    std::fill(d.begin(), d.end(), -1);
    int h = 0, tl = 0;
    q[tl++] = src; d[src] = 0;
    while (h < tl) {
      const int x = q[h++];
      for (int i = rp[x]; i < rp[x + 1]; ++i) { const int y = adj[i].v; if (d[y] < 0) { d[y] = d[x] + 1; q[tl++] = y; } }
      for (int i = rrp[x]; i < rrp[x + 1]; ++i) { const int y = radj[i].v; if (d[y] < 0) { d[y] = d[x] + 1; q[tl++] = y; } }
    }
  };
  // start from the vertex of largest degree's component: vertex 0 may be isolated (ids need not be dense)
  int start = 0;
  for (int v = 0; v < V; ++v)
    if (rp[v + 1] - rp[v] + rrp[v + 1] - rrp[v] > rp[start + 1] - rp[start] + rrp[start + 1] - rrp[start]) start = v;
  bfs(start, db);
  int a = start;
  for (int v = 0; v < V; ++v) if (db[v] > db[a]) a = v;
  bfs(a, da);
  int b = a;
  for (int v = 0; v < V; ++v) if (da[v] > da[b]) b = v;
  bfs(b, db);
  int c = a; long best = -1;
  for (int v = 0; v < V; ++v) { const long m = std::min(da[v], db[v]); if (m > best) { best = m; c = v; } }
  bfs(c, dc);
  int pmin = INT_MAX, qmin = INT_MAX, pmax = INT_MIN, qmax = INT_MIN;
  for (int v = 0; v < V; ++v) {
    if (da[v] < 0 || dc[v] < 0) continue;
    const int pp = da[v] + dc[v], qq = da[v] - dc[v];
    pmin = std::min(pmin, pp); pmax = std::max(pmax, pp); qmin = std::min(qmin, qq); qmax = std::max(qmax, qq);
  }
  if (pmin > pmax) { pmin = pmax = qmin = qmax = 0; }
  unsigned side = 2;
  while ((int)side <= std::max(pmax - pmin, qmax - qmin) / 2 + 1) side <<= 1;
  std::vector<std::pair<unsigned long long, int>> keyed(V);
  for (int v = 0; v < V; ++v) {
    unsigned long long kx = ~0ull >> 1;   // other components: after the main one, in id order
    if (da[v] >= 0 && dc[v] >= 0) kx = hilbert_key(side, (unsigned)(da[v] + dc[v] - pmin) / 2, (unsigned)(da[v] - dc[v] - qmin) / 2);
    keyed[v] = std::make_pair(kx, v);
  }
  std::sort(keyed.begin(), keyed.end());
  for (int x = 0; x < V; ++x) { iperm[x] = keyed[x].second; perm[keyed[x].second] = x; }
}

// CSR -> four-wide ELL + overflow list.  .pad of every record: low 3 bits = the slot, in the head vertex's own
// record, of an edge leading straight back to this vertex (7 = none; relaxing it can never improve anything
// since weights are >= 0); slot 0 / slot 1 of a vertex carry the overflow start / count above those bits.
static void build_ell(const std::vector<int>& rp, const std::vector<EdgeRec>& adj, int V, std::vector<EdgeRec>& ell,
                      std::vector<EdgeRec>& ovf) {
  ell.assign((size_t)4 * std::max(V, 1), EdgeRec{0.0, -1, 7});
  ovf.clear();
  for (int v = 0; v < V; ++v) {
    const int b = rp[v], deg = rp[v + 1] - rp[v];
    for (int j = 0; j < std::min(deg, 4); ++j) {
      EdgeRec r = adj[b + j];
      r.pad = 7;
      const int h = r.v, hb = rp[h], hdeg = std::min(rp[h + 1] - rp[h], 4);
      for (int k = 0; k < hdeg; ++k)
        if (adj[hb + k].v == v) { r.pad = k; break; }
      ell[(size_t)4 * v + j] = r;
    }
    if (deg > 4) {
      ell[(size_t)4 * v + 0].pad |= (int)ovf.size() << 3;
      ell[(size_t)4 * v + 1].pad |= (deg - 4) << 3;
      for (int j = 4; j < deg; ++j) ovf.push_back(adj[b + j]);
    }
  }
}

// The group kernel's ELL pair: the CSR / CSC in internal numbering (per-vertex edge order kept, so slot numbers and
// one-byte predecessor codes mean the same as in the caller's numbering), four edges per record + overflow.  The
// record's slot-0 pad carries the "sequential path" flag in bit 0.
static void build_group_ell(int V, const std::vector<int>& rp, const std::vector<EdgeRec>& adj, const std::vector<int>& perm,
                            const std::vector<int>& iperm, std::vector<EdgeRec>& gell, std::vector<EdgeRec>& govf,
                            std::vector<int4>& heads) {
  std::vector<int> rpn(V + 1, 0);
  std::vector<EdgeRec> adjn(adj.size());
  for (int x = 0; x < V; ++x) rpn[x + 1] = rpn[x] + (rp[iperm[x] + 1] - rp[iperm[x]]);
  for (int x = 0; x < V; ++x) {
    const int v = iperm[x];
    for (int i = rp[v], o = rpn[x]; i < rp[v + 1]; ++i, ++o) adjn[o] = EdgeRec{adj[i].w, perm[adj[i].v], 0};
  }
  build_ell(rpn, adjn, V, gell, govf);
  heads.assign(std::max(V, 1), make_int4(-1, -1, -1, -1));
  for (int x = 0; x < V; ++x) {
    EdgeRec* r = &gell[(size_t)4 * x];
    bool slow = rpn[x + 1] - rpn[x] > 4;
    for (int a = 0; a < 4; ++a) {
      if (r[a].v == x) slow = true;
      for (int b = a + 1; b < 4; ++b) if (r[a].v >= 0 && r[a].v == r[b].v) slow = true;
    }
    heads[x] = make_int4(r[0].v, r[1].v, r[2].v, r[3].v);
    for (int a = 0; a < 4; ++a) {
      const bool empty = r[a].v < 0;
      // bit 0: sequential path, bit 1: empty slot, bits 3..5: slot of the edge straight back (7 = none), bits 6..: overflow
      r[a].pad = (slow ? 1 : 0) | (empty ? 2 : 0) | ((r[a].pad & 7) << 3) | ((r[a].pad >> 3) << 6);
      if (empty) { r[a].v = x; r[a].w = 0.0; }   // never improves: the fast path needs no "has edge" test
    }
  }
}

static int g_build_and_upload(SqGraph* g) {
  if (int rc = sq::require_device()) return rc;
  if (g->nvertices == UINT_MAX) return sq::set_error(SQ_ERR_ARG, "graph has no vertex count");
  if (g->topo_dirty) {
    const int V = (int)g->nvertices;
    std::vector<int> rp(V + 1, 0), rrp(V + 1, 0);
    int E = 0;
    for (size_t i = 0; i < g->es.size(); ++i) {
      if (!g->alive[i]) continue;
      const int a = g->es[i], b = g->ee[i];
      if (a < 0 || a >= V || b < 0 || b >= V) continue;  // the reference would relax into .at() and abort
      ++rp[a + 1]; ++rrp[b + 1]; ++E;
    }
    for (int v = 0; v < V; ++v) { rp[v + 1] += rp[v]; rrp[v + 1] += rrp[v]; }
    g->h_adj.assign(E, EdgeRec{0.0, 0, 0});
    g->h_radj.assign(E, EdgeRec{0.0, 0, 0});
    g->edge_slot.assign(g->es.size(), -1);
    g->edge_rslot.assign(g->es.size(), -1);
    std::vector<int> fill(rp.begin(), rp.end() - 1), rfill(rrp.begin(), rrp.end() - 1);
    for (size_t i = 0; i < g->es.size(); ++i) {  // insertion order within a vertex, like the reference
      if (!g->alive[i]) continue;
      const int a = g->es[i], b = g->ee[i];
      if (a < 0 || a >= V || b < 0 || b >= V) continue;
      const int s = fill[a]++, r = rfill[b]++;
      g->edge_slot[i] = s; g->edge_rslot[i] = r;
      g->h_adj[s] = EdgeRec{g->ew[i], b, 0};
      g->h_radj[r] = EdgeRec{g->ew[i], a, 0};
    }
    g_free_device(g);
    g->V = V; g->E = E;
    g->h_rp = rp; g->h_rrp = rrp;
    SQ_CUDA_TRY(cudaMalloc(&g->d_ell, sizeof(EdgeRec) * 4 * (size_t)std::max(V, 1)));
    SQ_CUDA_TRY(cudaMalloc(&g->d_rell, sizeof(EdgeRec) * 4 * (size_t)std::max(V, 1)));
    SQ_CUDA_TRY(cudaMalloc(&g->d_rs_v, sizeof(int) * 4 * (size_t)std::max(V, 1)));
    SQ_CUDA_TRY(cudaMalloc(&g->d_rs_w, sizeof(double) * 4 * (size_t)std::max(V, 1)));
    SQ_CUDA_TRY(cudaMalloc(&g->d_fs_v, sizeof(int) * 4 * (size_t)std::max(V, 1)));
    SQ_CUDA_TRY(cudaMalloc(&g->d_fs_w, sizeof(double) * 4 * (size_t)std::max(V, 1)));
    SQ_CUDA_TRY(cudaMalloc(&g->d_row_ptr, sizeof(int) * (size_t)(V + 1)));
    SQ_CUDA_TRY(cudaMalloc(&g->d_rrow_ptr, sizeof(int) * (size_t)(V + 1)));
    SQ_CUDA_TRY(cudaMalloc(&g->d_adj, sizeof(EdgeRec) * (size_t)std::max(E, 1)));
    SQ_CUDA_TRY(cudaMalloc(&g->d_radj, sizeof(EdgeRec) * (size_t)std::max(E, 1)));
    SQ_CUDA_TRY(cudaMemcpy(g->d_row_ptr, rp.data(), sizeof(int) * (size_t)(V + 1), cudaMemcpyHostToDevice));
    SQ_CUDA_TRY(cudaMemcpy(g->d_rrow_ptr, rrp.data(), sizeof(int) * (size_t)(V + 1), cudaMemcpyHostToDevice));
    if (!g->d_counter) SQ_CUDA_TRY(cudaMalloc(&g->d_counter, sizeof(int) * 1024));
    if (!g->d_stats) SQ_CUDA_TRY(cudaMalloc(&g->d_stats, sizeof(long long) * 5));
    // group kernel: internal numbering and its device arrays
    locality_order(V, rp, g->h_adj, rrp, g->h_radj, g->h_perm, g->h_iperm);
    SQ_CUDA_TRY(cudaMalloc(&g->d_perm, sizeof(int) * (size_t)std::max(V, 1)));
    SQ_CUDA_TRY(cudaMalloc(&g->d_iperm, sizeof(int) * (size_t)std::max(V, 1)));
    SQ_CUDA_TRY(cudaMalloc(&g->d_tile_ord, (size_t)std::max(V, 1)));
    if (V > 0) {
      SQ_CUDA_TRY(cudaMemcpy(g->d_perm, g->h_perm.data(), sizeof(int) * (size_t)V, cudaMemcpyHostToDevice));
      SQ_CUDA_TRY(cudaMemcpy(g->d_iperm, g->h_iperm.data(), sizeof(int) * (size_t)V, cudaMemcpyHostToDevice));
      // group_finalize_kernel: thread i of a tile takes the tile's i-th vertex in caller-id order
      std::vector<unsigned char> ord((size_t)V);
      std::vector<int> idx;
      for (int b0 = 0; b0 < V; b0 += FIN_TILE) {
        const int m = std::min(FIN_TILE, V - b0);
        idx.resize(m);
        std::iota(idx.begin(), idx.end(), 0);
        std::sort(idx.begin(), idx.end(), [&](int a, int b) { return g->h_iperm[b0 + a] < g->h_iperm[b0 + b]; });
        for (int i = 0; i < m; ++i) ord[(size_t)b0 + i] = (unsigned char)idx[i];
      }
      SQ_CUDA_TRY(cudaMemcpy(g->d_tile_ord, ord.data(), (size_t)V, cudaMemcpyHostToDevice));
    }
    for (int d = 0; d < 2; ++d) {
      SQ_CUDA_TRY(cudaMalloc(&g->d_gell[d], sizeof(EdgeRec) * 4 * (size_t)std::max(V, 1)));
      SQ_CUDA_TRY(cudaMalloc(&g->d_gheads[d], sizeof(int4) * (size_t)std::max(V, 1)));
    }
    g->topo_dirty = false;
    g->auto_ring = 0; g->auto_warp_unfit = false; g->pending_check_n = 0; g->auto_checks = 0;
    g->w_dirty = true;
  } else if (g->w_dirty) {
    for (size_t i = 0; i < g->es.size(); ++i) {
      const int s = g->edge_slot[i];
      if (s < 0) continue;
      g->h_adj[s].w = g->ew[i];
      g->h_radj[g->edge_rslot[i]].w = g->ew[i];
    }
  }
  if (g->w_dirty) {
    if (g->E > 0) {
      SQ_CUDA_TRY(cudaMemcpy(g->d_adj, g->h_adj.data(), sizeof(EdgeRec) * (size_t)g->E, cudaMemcpyHostToDevice));
      SQ_CUDA_TRY(cudaMemcpy(g->d_radj, g->h_radj.data(), sizeof(EdgeRec) * (size_t)g->E, cudaMemcpyHostToDevice));
    }
    build_ell(g->h_rp, g->h_adj, g->V, g->h_ell, g->h_ovf);
    build_ell(g->h_rrp, g->h_radj, g->V, g->h_rell, g->h_rovf);
    if (g->h_ovf.size() + 1 > g->ovf_cap) {
      cudaFree(g->d_ovf); g->d_ovf = nullptr;
      g->ovf_cap = g->h_ovf.size() + 1;
      SQ_CUDA_TRY(cudaMalloc(&g->d_ovf, sizeof(EdgeRec) * g->ovf_cap));
    }
    if (g->h_rovf.size() + 1 > g->rovf_cap) {
      cudaFree(g->d_rovf); g->d_rovf = nullptr;
      g->rovf_cap = g->h_rovf.size() + 1;
      SQ_CUDA_TRY(cudaMalloc(&g->d_rovf, sizeof(EdgeRec) * g->rovf_cap));
    }
    SQ_CUDA_TRY(cudaMemcpy(g->d_ell, g->h_ell.data(), sizeof(EdgeRec) * g->h_ell.size(), cudaMemcpyHostToDevice));
    SQ_CUDA_TRY(cudaMemcpy(g->d_rell, g->h_rell.data(), sizeof(EdgeRec) * g->h_rell.size(), cudaMemcpyHostToDevice));
    {
      const size_t Vs = (size_t)std::max(g->V, 1);
      std::vector<int> sv(4 * Vs, -1);
      std::vector<double> sw(4 * Vs, 0.0);
      for (int dir = 0; dir < 2; ++dir) {
        const std::vector<EdgeRec>& src = dir ? g->h_ell : g->h_rell;
        for (int v = 0; v < g->V; ++v)
          for (int j = 0; j < 4; ++j) {
            sv[(size_t)j * Vs + v] = src[(size_t)4 * v + j].v;
            sw[(size_t)j * Vs + v] = src[(size_t)4 * v + j].w;
          }
        SQ_CUDA_TRY(cudaMemcpy(dir ? g->d_fs_v : g->d_rs_v, sv.data(), sizeof(int) * sv.size(), cudaMemcpyHostToDevice));
        SQ_CUDA_TRY(cudaMemcpy(dir ? g->d_fs_w : g->d_rs_w, sw.data(), sizeof(double) * sw.size(), cudaMemcpyHostToDevice));
      }
    }
    for (int d = 0; d < 2; ++d) {
      std::vector<int4> heads;
      build_group_ell(g->V, d ? g->h_rrp : g->h_rp, d ? g->h_radj : g->h_adj, g->h_perm, g->h_iperm, g->h_gell[d], g->h_govf[d], heads);
      if (g->h_govf[d].size() + 1 > g->govf_cap[d]) {
        cudaFree(g->d_govf[d]); g->d_govf[d] = nullptr;
        g->govf_cap[d] = g->h_govf[d].size() + 1;
        SQ_CUDA_TRY(cudaMalloc(&g->d_govf[d], sizeof(EdgeRec) * g->govf_cap[d]));
      }
      SQ_CUDA_TRY(cudaMemcpy(g->d_gell[d], g->h_gell[d].data(), sizeof(EdgeRec) * g->h_gell[d].size(), cudaMemcpyHostToDevice));
      SQ_CUDA_TRY(cudaMemcpy(g->d_gheads[d], heads.data(), sizeof(int4) * heads.size(), cudaMemcpyHostToDevice));
      if (!g->h_govf[d].empty())
        SQ_CUDA_TRY(cudaMemcpy(g->d_govf[d], g->h_govf[d].data(), sizeof(EdgeRec) * g->h_govf[d].size(), cudaMemcpyHostToDevice));
    }
    if (g->wslots_n != g->es.size()) {
      // where each logged edge's weight lives on the device (for sp_set_weights_device)
      const size_t n = g->es.size();
      std::vector<int> slots(6 * std::max<size_t>(n, 1), -1);
      for (size_t i = 0; i < n; ++i) {
        const int a = g->edge_slot[i], r = g->edge_rslot[i];
        if (a < 0) continue;
        slots[i] = a;
        slots[n + i] = r;
        const int v = g->es[i], ja = a - g->h_rp[v];
        slots[2 * n + i] = ja < 4 ? 4 * v + ja : -((g->h_ell[(size_t)4 * v].pad >> 3) + (ja - 4) + 1);
        const int u = g->ee[i], jr = r - g->h_rrp[u];
        slots[3 * n + i] = jr < 4 ? 4 * u + jr : -((g->h_rell[(size_t)4 * u].pad >> 3) + (jr - 4) + 1);
        const int xv = g->h_perm[v], xu = g->h_perm[u];
        slots[4 * n + i] = ja < 4 ? 4 * xv + ja : -((g->h_gell[0][(size_t)4 * xv].pad >> 6) + (ja - 4) + 1);
        slots[5 * n + i] = jr < 4 ? 4 * xu + jr : -((g->h_gell[1][(size_t)4 * xu].pad >> 6) + (jr - 4) + 1);
      }
      cudaFree(g->d_wslots); g->d_wslots = nullptr;
      SQ_CUDA_TRY(cudaMalloc(&g->d_wslots, sizeof(int) * slots.size()));
      SQ_CUDA_TRY(cudaMemcpy(g->d_wslots, slots.data(), sizeof(int) * slots.size(), cudaMemcpyHostToDevice));
      g->wslots_n = n;
    }
    if (!g->h_ovf.empty())
      SQ_CUDA_TRY(cudaMemcpy(g->d_ovf, g->h_ovf.data(), sizeof(EdgeRec) * g->h_ovf.size(), cudaMemcpyHostToDevice));
    if (!g->h_rovf.empty())
      SQ_CUDA_TRY(cudaMemcpy(g->d_rovf, g->h_rovf.data(), sizeof(EdgeRec) * g->h_rovf.size(), cudaMemcpyHostToDevice));
    // delta = factor x mean weight of the "open" edges (closed links carry 1e8, queue_class.py:469)
    double sum = 0.0; long long cnt = 0;
    for (int i = 0; i < g->E; ++i) {
      const double w = g->h_adj[i].w;
      if (w < 1e7 && w > 0.0) { sum += w; ++cnt; }
    }
    const double mean = cnt ? sum / (double)cnt : 1.0;
    g->delta = std::max(g->delta_factor * mean, 1e-300);
    g->w_dirty = false;
  }
  return SQ_OK;
}

template <int THREADS>
static int launch_trees(SqGraph* g, TreeParams& p, int grid, size_t smem, cudaStream_t st) {
  // function attributes belong to a device's context: one flag per device, set under a lock
  static bool attr_set[SQ_MAX_DEVICES] = {false};
  int dev = 0;
  SQ_CUDA_TRY(cudaGetDevice(&dev));
  {
    std::lock_guard<std::mutex> lock(g_attr_mutex);
    if (dev < 0 || dev >= SQ_MAX_DEVICES || !attr_set[dev]) {
      SQ_CUDA_TRY(cudaFuncSetAttribute(sssp_nearfar_kernel<THREADS>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
      if (dev >= 0 && dev < SQ_MAX_DEVICES) attr_set[dev] = true;
    }
  }
  sssp_nearfar_kernel<THREADS><<<grid, THREADS, smem, st>>>(p);
  sq::count_launch();
  SQ_CUDA_TRY(cudaGetLastError());
  return SQ_OK;
}

static int launch_bitmap_kernel(SqGraph* g, TreeParams& p, int grid, size_t smem, int threads, cudaStream_t st) {
  switch (threads) {
    case 128: return launch_trees<128>(g, p, grid, smem, st);
    case 256: return launch_trees<256>(g, p, grid, smem, st);
    case 512: return launch_trees<512>(g, p, grid, smem, st);
    case 1024: return launch_trees<1024>(g, p, grid, smem, st);
    default: return sq::set_error(SQ_ERR_ARG, "bitmap kernel: threads must be 128, 256, 512 or 1024 (got %d)", threads);
  }
}

// Largest V whose distance row fits next to the ELL kernel's static queues in one CTA's shared memory.
static size_t ell_smem_row_bytes(const cudaDeviceProp& prop) {
  cudaFuncAttributes fa;
  if (cudaFuncGetAttributes(&fa, sssp_ell_kernel<32, true>) != cudaSuccess) { cudaGetLastError(); return 0; }
  const size_t have = prop.sharedMemPerBlockOptin;
  return have > fa.sharedSizeBytes + 256 ? (have - fa.sharedSizeBytes - 256) & ~(size_t)15 : 0;
}
static size_t ell_smem_row_limit(const cudaDeviceProp& prop) { return ell_smem_row_bytes(prop) / sizeof(u64); }

static int launch_ell_kernel(TreeParams& p, int grid, int threads, cudaStream_t st, bool smem_row = false,
                             const cudaDeviceProp* prop = nullptr) {
  if (smem_row) {
    // a lone tree per SM with its distance row in shared memory (1024 threads)
    const size_t bytes = sizeof(u64) * (size_t)p.V;
    static bool attr_set[SQ_MAX_DEVICES] = {false};
    int dev = 0;
    SQ_CUDA_TRY(cudaGetDevice(&dev));
    {
      std::lock_guard<std::mutex> lock(g_attr_mutex);
      if (dev < 0 || dev >= SQ_MAX_DEVICES || !attr_set[dev]) {
        SQ_CUDA_TRY(cudaFuncSetAttribute(sssp_ell_kernel<32, true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)ell_smem_row_bytes(*prop)));
        if (dev >= 0 && dev < SQ_MAX_DEVICES) attr_set[dev] = true;
      }
    }
    sssp_ell_kernel<32, true><<<grid, 1024, bytes, st>>>(p);
    sq::count_launch();
    SQ_CUDA_TRY(cudaGetLastError());
    return SQ_OK;
  }
  switch (threads) {
    case 32: sssp_ell_kernel<1, false><<<grid, 32, 0, st>>>(p); break;
    case 64: sssp_ell_kernel<2, false><<<grid, 64, 0, st>>>(p); break;
    case 128: sssp_ell_kernel<4, false><<<grid, 128, 0, st>>>(p); break;
    case 256: sssp_ell_kernel<8, false><<<grid, 256, 0, st>>>(p); break;
    case 512: sssp_ell_kernel<16, false><<<grid, 512, 0, st>>>(p); break;
    case 1024: sssp_ell_kernel<32, false><<<grid, 1024, 0, st>>>(p); break;
    default: return sq::set_error(SQ_ERR_ARG, "ELL kernel: threads must be 32 ... 1024, a power of two (got %d)", threads);
  }
  sq::count_launch();
  SQ_CUDA_TRY(cudaGetLastError());
  return SQ_OK;
}

// Group path: the trees are sorted by the internal id of their origin (a Hilbert order over the network) and cut into
// groups of eight; a group's distances live interleaved, dist[v][8], in library scratch.  Two tree kernels work on that
// layout: the grouped variant of sssp_warp_kernel (group_mode 0: one CTA of eight warps per group, one tree per warp)
// and sssp_group_kernel (group_mode 1: one warp per group, no atomics, no capacity limits).  group_finalize_kernel then
// un-interleaves the rows into the caller's tables and writes the canonical predecessors in the same pass.
struct BitmapGeom { int grid, threads; size_t smem; bool fits; };

// host_list != nullptr: run only the trees listed there (the rerun of trees another kernel gave up on, on graphs too
// large for the bitmap kernel), in list order, on sssp_group_kernel.
static int g_run_group(SqGraph* g, const TreeParams& base, const BitmapGeom& bm, const cudaDeviceProp& prop, const int* origins_dev,
                       int n, u64* dist_dev, int* parent_dev, cudaStream_t st, bool reverse, const int* host_list = nullptr,
                       int n_list = 0) {
  const int V = g->V;
  const int n_sm = prop.multiProcessorCount;
  const int dir = reverse ? 1 : 0;
  const int n_groups = ((host_list ? n_list : n) + 7) / 8, n_pad = n_groups * 8;
  if (parent_dev && g->parent_format) {
    size_t worst = 0;
    const std::vector<int>& rp = reverse ? g->h_rp : g->h_rrp;
    for (int v = 0; v < V; ++v) worst = std::max<size_t>(worst, (size_t)(rp[v + 1] - rp[v]));
    if (worst > 254) return sq::set_error(SQ_ERR_LIMIT, "one-byte predecessor codes need in-degree <= 254 (found %zu)", worst);
  }
  int vbits = 1;
  while ((1ll << vbits) < (long long)V) ++vbits;
  // the warp kernel needs a fallback for trees whose wavefront outgrows its ring; without one (graph too large for the
  // bitmap kernel, or ids too wide for the hashed queue entries) the group runs on the kernel that cannot overflow
  const bool warp_mode = g->group_mode == 0 && vbits <= 24 && bm.fits && !host_list;
  // union kernel: two bits per vertex per warp in shared memory
  const size_t union_smem = (size_t)4 * 2 * ((V + 31) / 32) * sizeof(unsigned);
  const bool union_mode = g->group_mode == 2 && !host_list && bm.fits && union_smem <= 200 * 1024;

  // ---- sort the trees by origin (counting sort on the internal ids) ----
  if (g->gkey_cap < (size_t)n_pad) {
    cudaFree(g->d_gkey); cudaFree(g->d_tord); g->d_gkey = g->d_tord = nullptr;
    SQ_CUDA_TRY(cudaMalloc(&g->d_gkey, sizeof(int) * (size_t)n_pad));
    SQ_CUDA_TRY(cudaMalloc(&g->d_tord, sizeof(int) * (size_t)n_pad));
    g->gkey_cap = (size_t)n_pad;
  }
  if (g->ghist_cap < (size_t)V + 1) {
    cudaFree(g->d_ghist); g->d_ghist = nullptr;
    SQ_CUDA_TRY(cudaMalloc(&g->d_ghist, sizeof(int) * ((size_t)V + 1)));
    g->ghist_cap = (size_t)V + 1;
  }
  if (host_list) {
    std::vector<int> tl((size_t)n_pad, -1);
    std::copy(host_list, host_list + n_list, tl.begin());
    SQ_CUDA_TRY(cudaMemcpyAsync(g->d_tord, tl.data(), sizeof(int) * (size_t)n_pad, cudaMemcpyHostToDevice, st));
    SQ_CUDA_TRY(cudaStreamSynchronize(st));   // tl goes out of scope
    SQ_CUDA_TRY(cudaMemsetAsync(g->d_counter + 8, 0, sizeof(int) * 128, st));
  } else {
    SQ_CUDA_TRY(cudaMemsetAsync(g->d_ghist, 0, sizeof(int) * ((size_t)V + 1), st));
    group_hist_kernel<<<(n + 255) / 256, 256, 0, st>>>(n, V, origins_dev, g->d_perm, g->d_gkey, g->d_ghist);
    group_scan_kernel<<<1, 1024, 0, st>>>(V + 1, g->d_ghist);
    group_scatter_kernel<<<(n_pad + 255) / 256, 256, 0, st>>>(n, n_pad, g->d_gkey, g->d_ghist, g->d_tord);
    sq::count_launch(3);
    SQ_CUDA_TRY(cudaGetLastError());
  }

  // ---- scratch: interleaved distance rows for a chunk of groups ----
  const size_t row_bytes = (size_t)V * 64;
  size_t chunk = std::max<size_t>(1, std::min<size_t>((size_t)n_groups, g->group_scratch_bytes / std::max<size_t>(row_bytes, 1)));
  if (g->gdist_groups < chunk) {
    cudaFree(g->d_gdist); g->d_gdist = nullptr; g->gdist_groups = 0;
    SQ_CUDA_TRY(cudaMalloc(&g->d_gdist, row_bytes * chunk));
    g->gdist_groups = chunk;
  }

  // ---- kernel geometry ----
  int grid = 0;
  size_t warp_smem = 0;
  int C = 64, fcap = 0;
  int ring_cap = 64;
  const int pend_cap = V + 32, words = (V + 31) / 32;
  typedef void (*WarpKernel)(const TreeParams);
  typedef void (*GroupKernel)(const GroupParams);
  static const WarpKernel wkernels[4] = {sssp_warp_kernel<1, true>, sssp_warp_kernel<2, true>, sssp_warp_kernel<4, true>,
                                         sssp_warp_kernel<8, true>};
  static const GroupKernel ukernels[3] = {sssp_union_kernel<4>, sssp_union_kernel<8>, sssp_union_kernel<16>};
  static const GroupKernel gkernels[3][2] = {{sssp_group_kernel<8, 0>, sssp_group_kernel<8, 1>},
                                             {sssp_group_kernel<16, 0>, sssp_group_kernel<16, 1>},
                                             {sssp_group_kernel<32, 0>, sssp_group_kernel<32, 1>}};
  int dev = 0;
  SQ_CUDA_TRY(cudaGetDevice(&dev));
  if (warp_mode) {
    while (C < g->group_ring) C <<= 1;
    warp_smem = (size_t)8 * 2 * C * sizeof(unsigned);
    if (warp_smem > 200 * 1024) return sq::set_error(SQ_ERR_ARG, "group path: ring capacity %d does not fit shared memory", C);
    int per_sm = (int)((size_t)prop.sharedMemPerMultiprocessor / (warp_smem + 1024));
    per_sm = std::max(1, std::min(per_sm, std::max(1, g->group_warps_per_sm / 8)));
    per_sm = std::min(per_sm, 2048 / 256);
    grid = (int)std::min<size_t>(chunk, (size_t)n_sm * per_sm);
    fcap = std::min(std::max(V, 32), std::max(1024, V / 8));
    if (g->wfar_warps < grid * 8 || g->wfar_cap != fcap) {
      cudaFree(g->d_wfar); g->d_wfar = nullptr;
      SQ_CUDA_TRY(cudaMalloc(&g->d_wfar, sizeof(unsigned) * (size_t)grid * 8 * 2 * fcap));
      g->wfar_warps = grid * 8; g->wfar_cap = fcap;
    }
    static bool attr_set[SQ_MAX_DEVICES] = {false};
    static int last_pct[SQ_MAX_DEVICES];
    std::lock_guard<std::mutex> lock(g_attr_mutex);
    if (dev >= 0 && dev < SQ_MAX_DEVICES) {
      if (!attr_set[dev]) {
        for (int i = 0; i < 4; ++i)
          SQ_CUDA_TRY(cudaFuncSetAttribute(wkernels[i], cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        attr_set[dev] = true;
        last_pct[dev] = -1;
      }
      int pct = (int)((100 * (size_t)per_sm * (warp_smem + 1024) + prop.sharedMemPerMultiprocessor - 1) / prop.sharedMemPerMultiprocessor);
      pct = std::max(0, std::min(100, pct));
      if (pct != last_pct[dev]) {
        for (int i = 0; i < 4; ++i)
          SQ_CUDA_TRY(cudaFuncSetAttribute(wkernels[i], cudaFuncAttributePreferredSharedMemoryCarveout, pct));
        last_pct[dev] = pct;
      }
    }
  } else {
    int wps = std::max(4, std::min(g->group_warps_per_sm, 32)) / 4 * 4;
    if (union_mode) wps = std::max(4, std::min(wps, 4 * (int)(prop.sharedMemPerMultiprocessor / (union_smem + 1024))));
    const int warps = (int)std::min<size_t>(chunk, (size_t)n_sm * wps);
    grid = (warps + 3) / 4;
    while (ring_cap < V) ring_cap <<= 1;
    if (g->g_warps < grid * 4 || g->g_ring_cap != ring_cap || g->g_pend_cap != pend_cap) {
      cudaFree(g->d_gring); cudaFree(g->d_gpend); cudaFree(g->d_ginp);
      g->d_gring = g->d_gpend = g->d_ginp = nullptr; g->g_warps = 0;
      SQ_CUDA_TRY(cudaMalloc(&g->d_gring, sizeof(unsigned) * (size_t)grid * 4 * ring_cap));
      SQ_CUDA_TRY(cudaMalloc(&g->d_gpend, sizeof(unsigned) * (size_t)grid * 4 * 2 * pend_cap));
      SQ_CUDA_TRY(cudaMalloc(&g->d_ginp, sizeof(unsigned) * (size_t)grid * 4 * words));
      g->g_warps = grid * 4; g->g_ring_cap = ring_cap; g->g_pend_cap = pend_cap; g->g_words = words;
    }
    // sssp_group_kernel: no shared memory at all, give every byte of the SM's array to L1; union kernel: its bitmaps
    static bool attr_set[SQ_MAX_DEVICES] = {false};
    static int u_pct[SQ_MAX_DEVICES];
    std::lock_guard<std::mutex> lock(g_attr_mutex);
    if (dev >= 0 && dev < SQ_MAX_DEVICES) {
      if (!attr_set[dev]) {
        for (int a2 = 0; a2 < 3; ++a2) {
          for (int b2 = 0; b2 < 2; ++b2)
            SQ_CUDA_TRY(cudaFuncSetAttribute(gkernels[a2][b2], cudaFuncAttributePreferredSharedMemoryCarveout, 0));
          SQ_CUDA_TRY(cudaFuncSetAttribute(ukernels[a2], cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        }
        attr_set[dev] = true;
        u_pct[dev] = -1;
      }
      if (union_mode) {
        int pct = (int)((100 * (size_t)(wps / 4) * (union_smem + 1024) + prop.sharedMemPerMultiprocessor - 1) / prop.sharedMemPerMultiprocessor);
        pct = std::max(0, std::min(100, pct));
        if (pct != u_pct[dev]) {
          for (int a2 = 0; a2 < 3; ++a2)
            SQ_CUDA_TRY(cudaFuncSetAttribute(ukernels[a2], cudaFuncAttributePreferredSharedMemoryCarveout, pct));
          u_pct[dev] = pct;
        }
      }
    }
  }

  for (size_t g0 = 0; g0 < (size_t)n_groups; g0 += chunk) {
    const int m = (int)std::min(chunk, (size_t)n_groups - g0);
    int* counter = g->d_counter + 8 + (int)((g0 / chunk) & 127);   // [0..7] belong to the other kernels; zeroed per batch
    if (warp_mode) {
      TreeParams p = base;
      p.n_trees = m;                      // groups
      p.origins = origins_dev; p.dist_out = nullptr; p.parent_out = nullptr;
      p.tord = g->d_tord + g0 * 8; p.perm = g->d_perm; p.gdist = g->d_gdist;
      p.ell = g->d_gell[dir]; p.ovf = g->d_govf[dir];
      p.tree_counter = counter;
      p.wfar = g->d_wfar; p.wfar_cap = fcap; p.ring_cap = C; p.vbits = vbits;
      wkernels[g->warp_ilp == 1 ? 0 : g->warp_ilp == 2 ? 1 : g->warp_ilp == 8 ? 3 : 2]<<<std::min(m, grid), 256, warp_smem, st>>>(p);
    } else {
      GroupParams gp;
      memset(&gp, 0, sizeof(gp));
      gp.V = V; gp.n_groups = m;
      gp.gell = g->d_gell[dir]; gp.govf = g->d_govf[dir]; gp.gheads = g->d_gheads[dir];
      gp.perm = g->d_perm; gp.origins = origins_dev; gp.tord = g->d_tord + g0 * 8;
      gp.gdist = g->d_gdist; gp.ring = g->d_gring; gp.pend = g->d_gpend; gp.inp = g->d_ginp;
      gp.ring_cap = ring_cap; gp.pend_cap = pend_cap; gp.words = words;
      gp.counter = counter;
      gp.delta = g->delta;
      gp.stats = base.stats;
      gp.fail_list = base.fail_list; gp.fail_count = base.fail_count;
      if (union_mode)
        ukernels[g->group_pb <= 4 ? 0 : g->group_pb <= 8 ? 1 : 2]<<<(std::min(m, grid * 4) + 3) / 4, 128, union_smem, st>>>(gp);
      else
        gkernels[g->group_pb <= 8 ? 0 : g->group_pb <= 16 ? 1 : 2][g->group_pf ? 1 : 0]<<<(std::min(m, grid * 4) + 3) / 4, 128, 0, st>>>(gp);
    }
    sq::count_launch();
    SQ_CUDA_TRY(cudaGetLastError());
    if (dist_dev || parent_dev) {
      dim3 fgrid((unsigned)(((size_t)V + 255) / 256), (unsigned)std::max(1, std::min(m, 2048)));
      group_finalize_kernel<<<fgrid, FIN_TILE, 0, st>>>(V, m, g->d_gell[dir ^ 1], g->d_govf[dir ^ 1], g->d_iperm, g->d_tile_ord,
                                                       origins_dev, g->d_tord + g0 * 8, g->d_gdist, dist_dev, parent_dev,
                                                       g->parent_format);
      sq::count_launch();
      SQ_CUDA_TRY(cudaGetLastError());
    }
  }
  if (warp_mode || union_mode) {
    // trees whose wavefront outgrew the ring (normally none: these CTAs exit at once): the bitmap kernel recomputes
    // them in the caller's numbering, straight into the caller's tables, predecessors included
    TreeParams pf = base;
    pf.origins = origins_dev; pf.dist_out = dist_dev; pf.parent_out = parent_dev; pf.parent_codes = g->parent_format;
    pf.tree_counter = g->d_fail + 1;
    pf.tree_list = g->d_fail + 2;
    pf.n_list = g->d_fail;
    pf.stats = nullptr;
    if (int rc = launch_bitmap_kernel(g, pf, bm.grid, bm.smem, bm.threads, st)) return rc;
  }
  if (!host_list) g->last_stats[0] = n;
  return SQ_OK;
}

// Runs n trees; origins_dev / outputs are device pointers (outputs may be null).
static int g_run_batch_device(SqGraph* g, const int* origins_dev, int n, u64* dist_dev, int* parent_dev,
                              cudaStream_t st, bool reverse = false) {
  if (n <= 0) return SQ_OK;
  if (int rc = g_build_and_upload(g)) return rc;
  const int V = g->V;
  if (V <= 0) return sq::set_error(SQ_ERR_ARG, "graph has no vertices");
  int dev = 0;
  SQ_CUDA_TRY(cudaGetDevice(&dev));
  if (g->prop_dev != dev) {     // cached per graph (a graph lives on one device), not per process
    SQ_CUDA_TRY(cudaGetDeviceProperties(&g->prop, dev));
    g->prop_dev = dev;
  }
  const cudaDeviceProp& prop = g->prop;
  const int n_sm = prop.multiProcessorCount;

  // ---- bitmap kernel geometry (main kernel when algo == 1, else the fallback of the ELL kernel) ----
  const int words = (V + 31) / 32;
  const size_t bits_bytes = (size_t)3 * words * sizeof(unsigned);
  const bool bitmap_fits = bits_bytes <= 160 * 1024;
  if (g->algo == 1 && !bitmap_fits)
    return sq::set_error(SQ_ERR_LIMIT, "graph too large for the shared-memory vertex bitmaps (V=%d, limit ~436k vertices)", V);
  int nq = g->smem_queue > 0 ? g->smem_queue : 2048;
  nq = std::min(nq, std::max(32, V));
  while (bits_bytes + (size_t)4 * nq * sizeof(int) > 196 * 1024 && nq > 32) nq /= 2;
  const size_t smem = bits_bytes + (size_t)4 * nq * sizeof(int);
  const int bm_threads = (g->algo == 1 && g->threads >= 128) ? g->threads : 256;
  int bm_per_sm = (int)((size_t)prop.sharedMemPerMultiprocessor / (smem + 1024 + 64));
  bm_per_sm = std::max(1, std::min(bm_per_sm, 2048 / bm_threads));
  bm_per_sm = std::min(bm_per_sm, 32);
  if (g->algo == 1 && g->ctas_per_sm > 0) bm_per_sm = std::min(bm_per_sm, g->ctas_per_sm);
  if (g->algo != 1) bm_per_sm = 1;  // fallback use only: keep its scratch small
  const int bm_grid = std::min(n, n_sm * bm_per_sm);

  if (bitmap_fits && (g->scratch_grid < bm_grid || g->scratch_V < V)) {
    cudaFree(g->d_dist_scratch); cudaFree(g->d_qscratch);
    g->d_dist_scratch = nullptr; g->d_qscratch = nullptr;
    SQ_CUDA_TRY(cudaMalloc(&g->d_dist_scratch, sizeof(u64) * (size_t)bm_grid * V));
    SQ_CUDA_TRY(cudaMalloc(&g->d_qscratch, sizeof(int) * (size_t)bm_grid * 4 * V));
    g->scratch_grid = bm_grid; g->scratch_V = V;
  }
  if (g->pending_check_n > 0 && g->d_fail) {
    // how did the wavefronts of this graph fit the warp kernel's ring last time?  More than 5 % of the trees handed
    // to the fallback kernel: double the ring once (2048 entries, 12 warps per SM), then give the graph to the ELL
    // kernel, whose queues spill to global memory
    int nfail = 0;
    SQ_CUDA_TRY(cudaMemcpyAsync(&nfail, g->d_fail, sizeof(int), cudaMemcpyDeviceToHost, st));
    SQ_CUDA_TRY(cudaStreamSynchronize(st));
    if ((long long)nfail * 20 > g->pending_check_n) {
      const int cur = g->auto_ring > 0 ? g->auto_ring : g->warp_ring;
      if (cur < 2048) g->auto_ring = 2048; else g->auto_warp_unfit = true;
    }
    g->pending_check_n = 0;
    g->auto_checks += 1;
  }
  if (g->fail_cap < n + 2) {
    cudaFree(g->d_fail); g->d_fail = nullptr;
    SQ_CUDA_TRY(cudaMalloc(&g->d_fail, sizeof(int) * (size_t)(n + 2)));
    g->fail_cap = n + 2;
  }
  SQ_CUDA_TRY(cudaMemsetAsync(g->d_counter, 0, sizeof(int) * 1024, st));
  SQ_CUDA_TRY(cudaMemsetAsync(g->d_stats, 0, sizeof(long long) * 5, st));
  SQ_CUDA_TRY(cudaMemsetAsync(g->d_fail, 0, sizeof(int) * 2, st));

  TreeParams p;
  memset(&p, 0, sizeof(p));
  p.V = V; p.n_trees = n;
  p.row_ptr = g->d_row_ptr; p.adj = g->d_adj; p.rrow_ptr = g->d_rrow_ptr; p.radj = g->d_radj;
  p.origins = origins_dev; p.dist_out = dist_dev; p.parent_out = parent_dev;
  p.dist_scratch = g->d_dist_scratch; p.qscratch = g->d_qscratch;
  p.tree_counter = g->d_counter;
#ifdef SQ_SSSP_STATS
  p.stats = g->d_stats;
#else
  p.stats = nullptr;
#endif
  p.delta = g->delta; p.nq = nq; p.words = words;
  p.ell = g->d_ell; p.ovf = g->d_ovf; p.rell = g->d_rell; p.rovf = g->d_rovf;
  p.fail_count = g->d_fail; p.fail_list = g->d_fail + 2;
  if (reverse) {  // trees on the reversed graph: distances *to* the origin, parent = next vertex towards it
    std::swap(p.row_ptr, p.rrow_ptr); std::swap(p.adj, p.radj);
    std::swap(p.ell, p.rell); std::swap(p.ovf, p.rovf);
  }

  if (g->algo == 1) {
    if (int rc = launch_bitmap_kernel(g, p, bm_grid, smem, bm_threads, st)) return rc;
    g->last_stats[0] = n;
    return SQ_OK;
  }

  // ---- group kernel: large batches whose origins lie close together (algo 4, or auto on dense batches) ----
  if (g->algo == 4 || (g->algo == 3 && n >= g->group_min_trees && (long long)n * g->group_min_density >= (long long)V)) {
    const BitmapGeom bm{bm_grid, bm_threads, smem, bitmap_fits};
    return g_run_group(g, p, bm, prop, origins_dev, n, dist_dev, parent_dev, st, reverse);
  }

  // ---- warp kernel (algo 2, default) or ELL kernel (algo 0) -------------------------------------------
  int vbits = 1;
  while ((1ll << vbits) < (long long)V) ++vbits;
  // auto (3): a warp per tree wins on throughput once its 16 warp slots per SM are mostly filled; below ~13 trees per SM
  // the CTA-per-tree ELL kernel is faster (r2 sweep, config 5: 512 trees 286 k/s and full batches 345 k/s on the ELL
  // kernel; the warp kernel 220 k/s at 1,024 trees, 353 k/s at 2,048, 390 k/s from 2,368 on) and finishes a lone tree
  // ~10x sooner (the router's 8 trees per weight refresh)
  const bool want_warp = g->algo == 2 || (g->algo == 3 && n >= 13 * n_sm && !g->auto_warp_unfit);
  const bool use_warp = want_warp && vbits <= 24;   // >= 5 hash bits per queue entry
  // a handful of trees (the device router's one per destination): the GPU is empty, so give each tree a whole
  // SM's worth of threads; a round is then one or two passes over the wavefront instead of five
  int threads = use_warp ? 128 : std::min(g->threads, 1024);
  if (!use_warp && g->algo == 3) threads = n <= n_sm ? 1024 : n <= 2 * n_sm ? 512 : 256;
  // a lone tree per SM is bound by the number of rounds, not by work: four times wider buckets save the rounds spent
  // on bucket transitions (8 trees on the config-4 graph: 0.57 -> 0.52 ms, 7 % more relaxations)
  if (!use_warp && g->algo == 3 && n <= n_sm) p.delta = g->delta * 4.0;
  int per_sm, grid, rows;   // rows = trees in flight (scratch distance rows in timing-only mode)
  size_t warp_smem = 0;
  if (use_warp) {
    int C = 64;
    while (C < (g->algo == 3 && g->auto_ring > 0 ? g->auto_ring : g->warp_ring)) C <<= 1;
    if (g->algo == 3 && g->auto_checks < 3) g->pending_check_n = n;   // the first batches after a topology change
    warp_smem = (size_t)4 * 2 * C * sizeof(unsigned);
    if (warp_smem > 200 * 1024) return sq::set_error(SQ_ERR_ARG, "warp kernel: ring capacity %d does not fit shared memory", C);
    per_sm = (int)((size_t)prop.sharedMemPerMultiprocessor / (warp_smem + 1024));
    // 16 warps per SM measured best on B200 (tools/sssp_sweep.py, profiles/): beyond that the L1TEX wavefront
    // pipe queues up and every step's two dependent memory waits grow faster than the warp count
    per_sm = std::max(1, std::min(per_sm, g->warp_ctas_per_sm > 0 ? std::min(g->warp_ctas_per_sm, 16) : 4));
    {
      // all trees of a graph take about the same time, so a batch finishes in whole waves: spread the trees
      // evenly over the waves instead of leaving the last one mostly empty
      const int max_rows = n_sm * per_sm * 4;
      const int waves = (n + max_rows - 1) / max_rows;
      const int even = (n + waves - 1) / waves;
      grid = std::min((even + 3) / 4, n_sm * per_sm);
    }
    rows = grid * 4;
    const int fcap = std::min(std::max(V, 32), std::max(1024, V / 8));
    if (g->wfar_warps < rows || g->wfar_cap != fcap) {
      cudaFree(g->d_wfar); g->d_wfar = nullptr;
      SQ_CUDA_TRY(cudaMalloc(&g->d_wfar, sizeof(unsigned) * (size_t)rows * 2 * fcap));
      g->wfar_warps = rows; g->wfar_cap = fcap;
    }
    p.wfar = g->d_wfar; p.wfar_cap = fcap; p.ring_cap = C; p.vbits = vbits;
  } else {
    per_sm = g->ctas_per_sm > 0 ? g->ctas_per_sm : std::min(9, 1024 / threads);
    per_sm = std::max(1, std::min(per_sm, std::min(9, 2048 / threads)));
    grid = std::min(n, n_sm * per_sm);
    rows = grid;
    const int near_cap = std::max(32, (int)(g->queue_factor * V));
    const int far_cap = 2 * near_cap;
    const size_t per_cta = 4 * (size_t)near_cap + 2 * (size_t)far_cap;
    if (g->q2_grid < grid || g->q2_near != near_cap) {
      cudaFree(g->d_q2scratch); g->d_q2scratch = nullptr;
      cudaFree(g->d_q2v); g->d_q2v = nullptr;
      SQ_CUDA_TRY(cudaMalloc(&g->d_q2scratch, sizeof(u64) * (size_t)grid * per_cta));
      SQ_CUDA_TRY(cudaMalloc(&g->d_q2v, sizeof(int2) * (size_t)grid * per_cta));
      g->q2_grid = grid; g->q2_near = near_cap; g->q2_far = far_cap;
    }
    p.q2scratch = g->d_q2scratch; p.q2v = g->d_q2v; p.near_cap = near_cap; p.far_cap = far_cap;
  }
  if (parent_dev && !dist_dev) {
    // predecessors need the finished distances of every tree: keep them in a library buffer
    if (g->dist_tmp_elems < (size_t)n * V) {
      cudaFree(g->d_dist_tmp); g->d_dist_tmp = nullptr;
      SQ_CUDA_TRY(cudaMalloc(&g->d_dist_tmp, sizeof(u64) * (size_t)n * V));
      g->dist_tmp_elems = (size_t)n * V;
    }
    p.dist_out = g->d_dist_tmp;
  }
  if (!p.dist_out) {
    // timing-only mode keeps distances in a scratch row per tree in flight
    if (g->scratch_grid < rows || g->scratch_V < V || !g->d_dist_scratch) {
      cudaFree(g->d_dist_scratch); cudaFree(g->d_qscratch);
      g->d_dist_scratch = nullptr; g->d_qscratch = nullptr;
      SQ_CUDA_TRY(cudaMalloc(&g->d_dist_scratch, sizeof(u64) * (size_t)rows * V));
      SQ_CUDA_TRY(cudaMalloc(&g->d_qscratch, sizeof(int) * (size_t)std::max(bm_grid, 1) * 4 * V));
      g->scratch_grid = rows; g->scratch_V = V;
    }
    p.dist_scratch = g->d_dist_scratch; p.qscratch = g->d_qscratch;
  }
  p.parent_out = nullptr;  // predecessors come from sssp_parent_kernel below
  if (use_warp) {
    typedef void (*WarpKernel)(const TreeParams);
    static const WarpKernel kernels[4] = {sssp_warp_kernel<1, false>, sssp_warp_kernel<2, false>, sssp_warp_kernel<4, false>,
                                          sssp_warp_kernel<8, false>};
    static bool attr_set[SQ_MAX_DEVICES] = {false};
    static int last_pct[SQ_MAX_DEVICES];
    std::lock_guard<std::mutex> lock(g_attr_mutex);
    const int di = (dev >= 0 && dev < SQ_MAX_DEVICES) ? dev : 0;
    if (!attr_set[di]) {
      for (int i = 0; i < 4; ++i)
        SQ_CUDA_TRY(cudaFuncSetAttribute(kernels[i], cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
      attr_set[di] = true;
      last_pct[di] = -1;
    }
    {
      // shared-memory carve-out = what the resident CTAs need and no more: with the default (largest) carve-out
      // L1 shrinks to ~28 KB and the kernel loses a quarter of its throughput (measured)
      int pct = (int)((100 * (size_t)per_sm * (warp_smem + 1024) + prop.sharedMemPerMultiprocessor - 1) / prop.sharedMemPerMultiprocessor);
      pct = std::max(0, std::min(100, pct));
      if (pct != last_pct[di]) {
        for (int i = 0; i < 4; ++i)
          SQ_CUDA_TRY(cudaFuncSetAttribute(kernels[i], cudaFuncAttributePreferredSharedMemoryCarveout, pct));
        last_pct[di] = pct;
      }
    }
    kernels[g->warp_ilp == 1 ? 0 : g->warp_ilp == 2 ? 1 : g->warp_ilp == 8 ? 3 : 2]<<<grid, 128, warp_smem, st>>>(p);
    sq::count_launch();
    SQ_CUDA_TRY(cudaGetLastError());
  } else {
    // a handful of trees on an empty GPU whose vertex count fits: distance rows in shared memory
    const bool smem_row = g->algo == 3 && g->ell_smem_row && n <= n_sm && threads == 1024 && (size_t)V <= ell_smem_row_limit(prop);
    if (int rc = launch_ell_kernel(p, grid, threads, st, smem_row, &prop)) return rc;
  }
  // fallback pass over the trees whose queues overflowed (normally none: its CTAs exit at once)
  if (bitmap_fits) {
    TreeParams pf = p;
    pf.tree_counter = g->d_fail + 1;
    pf.tree_list = g->d_fail + 2;
    pf.n_list = g->d_fail;
    pf.stats = nullptr;
    if (int rc = launch_bitmap_kernel(g, pf, bm_grid, smem, bm_threads, st)) return rc;
  }
  std::vector<int> redo;
  if (!bitmap_fits) {
    // the graph is too large for the bitmap kernel's shared-memory bitmaps (V > ~436k): look at the overflow list now
    // and rerun those trees on the group kernel, whose queues are bounded by V and cannot overflow
    int nfail = 0;
    SQ_CUDA_TRY(cudaMemcpyAsync(&nfail, g->d_fail, sizeof(int), cudaMemcpyDeviceToHost, st));
    SQ_CUDA_TRY(cudaStreamSynchronize(st));
    if (nfail > 0) {
      redo.resize((size_t)nfail);
      SQ_CUDA_TRY(cudaMemcpy(redo.data(), g->d_fail + 2, sizeof(int) * (size_t)nfail, cudaMemcpyDeviceToHost));
    }
  }
  if (parent_dev) {
    const int codes = g->parent_format;
    if (codes) {
      size_t worst = 0;
      const std::vector<int>& rp = reverse ? g->h_rp : g->h_rrp;
      for (int v = 0; v < V; ++v) worst = std::max<size_t>(worst, (size_t)(rp[v + 1] - rp[v]));
      if (worst > 254) return sq::set_error(SQ_ERR_LIMIT, "one-byte predecessor codes need in-degree <= 254 (found %zu)", worst);
    }
    // each thread keeps its in-edge record in registers and walks over ~16 trees
    dim3 pgrid((unsigned)(((size_t)V + 255) / 256), (unsigned)std::max(1, std::min((n + 15) / 16, 32768)));
    if (reverse)
      sssp_parent_kernel<<<pgrid, 256, 0, st>>>(V, n, g->d_fs_v, g->d_fs_w, g->d_ell, g->d_ovf, g->h_ovf.empty() ? 0 : 1,
                                                origins_dev, p.dist_out, parent_dev, codes);
    else
      sssp_parent_kernel<<<pgrid, 256, 0, st>>>(V, n, g->d_rs_v, g->d_rs_w, g->d_rell, g->d_rovf, g->h_rovf.empty() ? 0 : 1,
                                                origins_dev, p.dist_out, parent_dev, codes);
    sq::count_launch();
    SQ_CUDA_TRY(cudaGetLastError());
  }
  if (!redo.empty()) {
    const BitmapGeom bm{bm_grid, bm_threads, smem, bitmap_fits};
    if (int rc = g_run_group(g, p, bm, prop, origins_dev, n, dist_dev, parent_dev, st, reverse, redo.data(), (int)redo.size()))
      return rc;
  }
  g->last_stats[0] = n;
  return SQ_OK;
}

static int g_fetch_stats(SqGraph* g) {
#ifdef SQ_SSSP_STATS
  long long tmp[5];
  SQ_CUDA_TRY(cudaMemcpy(tmp, g->d_stats, sizeof(tmp), cudaMemcpyDeviceToHost));
  for (int i = 1; i < 5; ++i) g->last_stats[i] = tmp[i];
#endif
  if (g->d_fail) {
    int nfail = 0;
    SQ_CUDA_TRY(cudaMemcpy(&nfail, g->d_fail, sizeof(int), cudaMemcpyDeviceToHost));
    g->last_stats[4] = nfail;
  }
  return SQ_OK;
}

// =================================================================================================
// C ABI
// =================================================================================================
extern "C" {

Graph* simplegraph(bool directed) {
  SqGraph* g = new SqGraph();
  g->directed = directed;
  g->nvertices = 7;
  static const int a[11] = {1, 1, 1, 2, 2, 3, 3, 3, 4, 4, 5};
  static const int b[11] = {2, 3, 6, 3, 4, 1, 4, 6, 3, 5, 6};
  static const double w[11] = {1.5, 9.1, 14.3, 15.9, 5.5, 5.6, 11.6, 2.4, 0.2, 6.2, 9.7};
  for (int i = 0; i < 11; ++i) g_add_edge(g, a[i], b[i], w[i]);
  return g;
}

Graph* readgraph(char* filename, bool directed) {
  SqGraph* g = new SqGraph();
  g->directed = directed;
  std::ifstream file(filename ? filename : "");
  if (!file.is_open() || !file.good()) {
    // the reference prints and returns a graph with no vertex count (graph.cc:88-95)
    std::cout << "Read matrix market file: Input file not found\n";
    sq::set_error(SQ_ERR_ARG, "readgraph: cannot open %s", filename ? filename : "(null)");
    return g;
  }
  std::string line;
  bool header = true;
  while (std::getline(file, line)) {
    if (line.find('#') != std::string::npos || line.find('%') != std::string::npos || line.empty()) continue;
    std::istringstream is(line);
    if (header) {
      unsigned nv = 0;
      is >> nv;
      g->nvertices = nv + 1;
      header = false;
      continue;
    }
    int v1, v2;
    double w;
    while (is >> v1 >> v2 >> w) g_add_edge(g, v1, v2, w);
  }
  std::cout << "Graph summary #edges: " << g->key2edge.size() << " #vertices: " << g->nvertices << "\n";
  return g;
}

Graph* creategraph(int* edge_starts, int* edge_ends, double* edge_wghs, int nedges, int nvertices, bool directed) {
  SqGraph* g = new SqGraph();
  g->directed = directed;
  g->nvertices = (unsigned)(nvertices + 1);
  g->es.reserve(nedges); g->ee.reserve(nedges); g->ew.reserve(nedges); g->alive.reserve(nedges);
  for (int i = 0; i < nedges; ++i) g_add_edge(g, edge_starts[i], edge_ends[i], edge_wghs[i]);
  return g;
}

bool writegraph(Graph* g, char* filename) {
  if (!g || !filename) return false;
  std::ofstream file(filename);
  if (!file.is_open()) {
    std::cout << "Write matrix market file: Output file not found\n";
    return false;
  }
  g_pull_host_log(g);
  file << "%%MatrixMarket matrix coordinate real general\n%\n";
  file << g->nvertices << " " << g->nvertices << " " << g->key2edge.size() << "\n";
  for (const auto& kv : g->key2edge)
    file << kv.first.first << " " << kv.first.second << " " << g->ew[kv.second] << "\n";
  return true;
}

void update_edge(Graph* g, int v1, int v2, double w) {
  if (!g) return;
  auto it = g->key2edge.find(std::make_pair(v1, v2));
  if (it == g->key2edge.end()) {
    ++g->ignored_updates;
    sq::set_error(SQ_ERR_ARG, "update_edge: no edge (%d,%d); ignored", v1, v2);
    return;
  }
  g_pull_host_log(g);
  g->ew[it->second] = w;
  g->w_dirty = true;
}

int sp_remove_edge(Graph* g, int v1, int v2) {
  if (!g) return SQ_ERR_ARG;
  if (!g->directed && v1 > v2) std::swap(v1, v2);
  auto it = g->key2edge.find(std::make_pair(v1, v2));
  if (it == g->key2edge.end()) return sq::set_error(SQ_ERR_ARG, "remove_edge: no edge (%d,%d)", v1, v2);
  if (int rc = g_pull_host_log(g)) return rc;
  g->alive[it->second] = 0;
  g->key2edge.erase(it);
  g->topo_dirty = true;
  return SQ_OK;
}

int sp_set_weights(Graph* g, const double* w, int nedges) {
  if (!g || !w) return sq::set_error(SQ_ERR_ARG, "sp_set_weights: null argument");
  if (nedges != (int)g->ew.size())
    return sq::set_error(SQ_ERR_ARG, "sp_set_weights: %d weights for %d edges", nedges, (int)g->ew.size());
  if (int rc = g_pull_host_log(g)) return rc;   // keeps the host CSR consistent; every weight is replaced below
  std::copy(w, w + nedges, g->ew.begin());
  g->w_dirty = true;
  return SQ_OK;
}

int sp_sync(Graph* g) {
  if (!g) return sq::set_error(SQ_ERR_ARG, "sp_sync: null graph");
  return g_build_and_upload(g);
}

int sp_nvertices(Graph* g) { return g ? (int)g->nvertices : 0; }
int sp_nedges(Graph* g) {
  if (!g) return 0;
  int n = 0;
  for (char a : g->alive) n += a;
  return n;
}

int sp_has_pending_edits(Graph* g) { return (g && (g->topo_dirty || g->w_dirty)) ? 1 : 0; }

int sp_nedges_logged(Graph* g) { return g ? (int)g->es.size() : 0; }

void sp_free_graph(Graph* g) {
  if (!g) return;
  g_free_device(g);
  if (g->ev_wlog) cudaEventDestroy(g->ev_wlog);
  cudaFree(g->d_counter); cudaFree(g->d_stats); cudaFree(g->d_dist_scratch); cudaFree(g->d_qscratch);
  cudaFree(g->d_origins); cudaFree(g->d_dist_buf); cudaFree(g->d_parent_buf);
  cudaFree(g->d_q2scratch); cudaFree(g->d_q2v); cudaFree(g->d_dist_tmp); cudaFree(g->d_fail); cudaFree(g->d_wfar);
  cudaFree(g->d_pair); cudaFree(g->d_pair_off); cudaFree(g->d_pair_dist); cudaFree(g->d_pool);
  cudaFree(g->d_gkey); cudaFree(g->d_tord); cudaFree(g->d_ghist); cudaFree(g->d_gdist);
  cudaFree(g->d_gring); cudaFree(g->d_gpend); cudaFree(g->d_ginp);
  if (g->h_pool) cudaFreeHost(g->h_pool);
  if (g->s_compute) {
    cudaStreamDestroy(g->s_compute); cudaStreamDestroy(g->s_copy);
    for (int i = 0; i < 2; ++i) { cudaEventDestroy(g->ev_done[i]); cudaEventDestroy(g->ev_copied[i]); }
  }
  delete g;
}

int sp_set_tuning(Graph* g, int threads, int ctas_per_sm, double delta_factor, int smem_queue) {
  if (!g) return sq::set_error(SQ_ERR_ARG, "sp_set_tuning: null graph");
  if (threads > 0) {
    if (threads != 32 && threads != 64 && threads != 128 && threads != 256 && threads != 512 && threads != 1024)
      return sq::set_error(SQ_ERR_ARG, "threads must be 32, 64, 128, 256, 512 or 1024");
    g->threads = threads;
  }
  if (ctas_per_sm >= 0) g->ctas_per_sm = ctas_per_sm;
  if (delta_factor > 0.0) { g->delta_factor = delta_factor; g->w_dirty = true; }
  if (smem_queue >= 0) g->smem_queue = smem_queue;
  return SQ_OK;
}

int sp_set_algo(Graph* g, int algo, double queue_factor) {
  if (!g) return sq::set_error(SQ_ERR_ARG, "sp_set_algo: null graph");
  if (algo < 0 || algo > 4)
    return sq::set_error(SQ_ERR_ARG, "sp_set_algo: algo must be 0 (ELL), 1 (bitmap), 2 (warp), 3 (auto) or 4 (group)");
  g->algo = algo;
  if (algo == 1 && g->threads < 128) g->threads = 256;
  if (algo == 3 && g->threads > 256) g->threads = 256;
  if (queue_factor > 0.0) g->queue_factor = queue_factor;
  return SQ_OK;
}

int sp_set_warp_tuning(Graph* g, int ring_cap, int ilp, int ctas_per_sm) {
  if (!g) return sq::set_error(SQ_ERR_ARG, "sp_set_warp_tuning: null graph");
  if (ilp != 0 && ilp != 1 && ilp != 2 && ilp != 4 && ilp != 8) return sq::set_error(SQ_ERR_ARG, "sp_set_warp_tuning: ilp must be 1, 2, 4 or 8");
  if (ring_cap > 0) g->warp_ring = ring_cap;
  if (ilp > 0) g->warp_ilp = ilp;
  if (ctas_per_sm >= 0) g->warp_ctas_per_sm = ctas_per_sm;
  return SQ_OK;
}

int sp_set_smem_rows(Graph* g, int on) {
  if (!g) return sq::set_error(SQ_ERR_ARG, "sp_set_smem_rows: null graph");
  g->ell_smem_row = on ? 1 : 0;
  return SQ_OK;
}

int sp_set_group_mode(Graph* g, int mode, int ring_cap) {
  if (!g || mode < 0 || mode > 2) return sq::set_error(SQ_ERR_ARG, "sp_set_group_mode: mode must be 0, 1 or 2");
  g->group_mode = mode;
  if (ring_cap > 0) g->group_ring = ring_cap;
  return SQ_OK;
}

int sp_set_group_tuning(Graph* g, int warps_per_sm, int batch, int prefetch, int min_trees, int min_density) {
  if (!g) return sq::set_error(SQ_ERR_ARG, "sp_set_group_tuning: null graph");
  if (batch != 0 && batch != 4 && batch != 8 && batch != 16 && batch != 32)
    return sq::set_error(SQ_ERR_ARG, "sp_set_group_tuning: batch must be 4, 8, 16 or 32");
  if (warps_per_sm > 0) g->group_warps_per_sm = warps_per_sm;
  if (batch > 0) g->group_pb = batch;
  if (prefetch >= 0) g->group_pf = prefetch;
  if (min_trees > 0) g->group_min_trees = min_trees;
  if (min_density > 0) g->group_min_density = min_density;
  return SQ_OK;
}

int sp_set_parent_format(Graph* g, int format) {
  if (!g || (format != 0 && format != 1)) return sq::set_error(SQ_ERR_ARG, "sp_set_parent_format: format must be 0 or 1");
  g->parent_format = format;
  return SQ_OK;
}

int sp_decode_parents_device(Graph* g, const unsigned char* codes_dev, int n, int* parent_dev, void* stream, int reverse) {
  if (!g || !codes_dev || !parent_dev) return sq::set_error(SQ_ERR_ARG, "sp_decode_parents_device: null argument");
  if (n <= 0) return SQ_OK;
  if (int rc = g_build_and_upload(g)) return rc;
  dim3 grid((unsigned)(((size_t)g->V + 255) / 256), (unsigned)std::max(1, std::min((n + 15) / 16, 32768)));
  if (reverse)
    decode_parents_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(g->V, n, g->d_fs_v, g->d_ell, g->d_ovf, codes_dev, parent_dev);
  else
    decode_parents_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(g->V, n, g->d_rs_v, g->d_rell, g->d_rovf, codes_dev, parent_dev);
  sq::count_launch();
  SQ_CUDA_TRY(cudaGetLastError());
  return SQ_OK;
}

int sp_last_stats(Graph* g, long long* out5) {
  if (!g || !out5) return SQ_ERR_ARG;
  for (int i = 0; i < 5; ++i) out5[i] = g->last_stats[i];
  return SQ_OK;
}

const char* sp_last_error(void) { return sq::err_buf(); }
long long sp_kernel_launches(void) { return sq::launch_counter().load(); }

int sp_sssp_batch_device_ex(Graph* g, const int* origins_dev, int n, double* dist_dev, int* parent_dev, void* stream,
                            int reverse) {
  if (!g || (!origins_dev && n > 0)) return sq::set_error(SQ_ERR_ARG, "sp_sssp_batch_device_ex: null argument");
  return g_run_batch_device(g, origins_dev, n, reinterpret_cast<u64*>(dist_dev), parent_dev, (cudaStream_t)stream, reverse != 0);
}

int sp_set_weights_device(Graph* g, const double* weights_dev, int nedges, void* stream) {
  if (!g || !weights_dev) return sq::set_error(SQ_ERR_ARG, "sp_set_weights_device: null argument");
  if (nedges != (int)g->ew.size())
    return sq::set_error(SQ_ERR_ARG, "sp_set_weights_device: %d weights for %d edges", nedges, (int)g->ew.size());
  if (int rc = g_build_and_upload(g)) return rc;  // pending host edits first: the device copy is about to lead
  if (nedges == 0) return SQ_OK;
  cudaStream_t st = (cudaStream_t)stream;
  const size_t n = (size_t)nedges;
  WeightSlots ws{g->d_wslots, g->d_wslots + n, g->d_wslots + 2 * n, g->d_wslots + 3 * n, g->d_wslots + 4 * n, g->d_wslots + 5 * n};
  if (!g->dev_leads) {   // first device-side refresh since the host last led: seed the device log with the host's
    if (g->wlog_cap < n) {
      cudaFree(g->d_wlog); g->d_wlog = nullptr;
      SQ_CUDA_TRY(cudaMalloc(&g->d_wlog, sizeof(double) * n));
      g->wlog_cap = n;
    }
    SQ_CUDA_TRY(cudaMemcpyAsync(g->d_wlog, g->ew.data(), sizeof(double) * n, cudaMemcpyHostToDevice, st));
  }
  scatter_weights_kernel<<<(nedges + 255) / 256, 256, 0, st>>>(nedges, g->V, weights_dev, ws, g->d_adj, g->d_radj, g->d_ell,
                                                             g->d_ovf, g->d_rell, g->d_rovf, g->d_fs_w, g->d_rs_w,
                                                             g->d_gell[0], g->d_govf[0], g->d_gell[1], g->d_govf[1], g->d_wlog);
  sq::count_launch();
  SQ_CUDA_TRY(cudaGetLastError());
  // The host log (what update_edge / writegraph / the next structural rebuild start from) is not touched here: a
  // device-to-host copy and a stream synchronisation per refresh leave the GPU idle while the host catches up, every
  // weight_freq ticks of a simulation.  The log catches up when somebody on the host looks (g_pull_host_log).
  if (!g->ev_wlog) SQ_CUDA_TRY(cudaEventCreateWithFlags(&g->ev_wlog, cudaEventDisableTiming));
  SQ_CUDA_TRY(cudaEventRecord(g->ev_wlog, st));
  g->dev_leads = true;
  return SQ_OK;
}

int sp_sssp_batch_device(Graph* g, const int* origins_dev, int n, double* dist_dev, int* parent_dev, void* stream) {
  if (!g || (!origins_dev && n > 0)) return sq::set_error(SQ_ERR_ARG, "sp_sssp_batch_device: null argument");
  return g_run_batch_device(g, origins_dev, n, reinterpret_cast<u64*>(dist_dev), parent_dev, (cudaStream_t)stream);
}

// ---- allocation-free variant (no counterpart in the reference) -----
// For callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops): the bitmap tree kernel with its
// scratch -- the tree scheduler's counter and the global overflow parts of the four queues of every CTA -- in a
// file-scope device array instead of per-graph allocations.  Every GPU that loads the library holds the array until
// exit; every call shares it, so calls that overlap on one GPU race.
#define SQ_NOALLOC_MAX_VERTICES 65536
#define SQ_NOALLOC_HEADER_BYTES 256                 /* [0] tree counter, cleared by every call */
#define SQ_NOALLOC_QUEUE_BYTES (8u << 20)           /* [CTAs][4][V] ints: 8 CTAs at the largest V, more on smaller graphs */
static __device__ __align__(256) unsigned char g_noalloc_scratch[SQ_NOALLOC_HEADER_BYTES + SQ_NOALLOC_QUEUE_BYTES];
static unsigned char* g_noalloc_addr[SQ_MAX_DEVICES];   // the array's device address on each device (host code cannot
                                                        // use the symbol's name as a pointer)

// Set-up, once per device, with that device current: the device address of the scratch array.
static int noalloc_scratch_address(unsigned char** out) {
  int dev = -1;
  SQ_CUDA_TRY(cudaGetDevice(&dev));
  if (dev < 0 || dev >= SQ_MAX_DEVICES) return sq::set_error(SQ_ERR_LIMIT, "sp_sssp_batch_device_noalloc: device %d out of range", dev);
  std::lock_guard<std::mutex> lock(g_attr_mutex);
  if (!g_noalloc_addr[dev]) {
    void* ptr = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&ptr, g_noalloc_scratch);
    if (e != cudaSuccess || !ptr) {
      cudaGetLastError();
      return sq::set_error(SQ_ERR_CUDA, "sp_sssp_batch_device_noalloc: cudaGetSymbolAddress failed on device %d: %s", dev,
                           cudaGetErrorString(e));
    }
    g_noalloc_addr[dev] = static_cast<unsigned char*>(ptr);
  }
  *out = g_noalloc_addr[dev];
  return SQ_OK;
}

int sp_sssp_batch_device_noalloc(Graph* g, const int* origins_dev, int n, double* dist_dev, int* parent_dev, void* stream) {
  if (!g || !dist_dev || (!origins_dev && n > 0))
    return sq::set_error(SQ_ERR_ARG, "sp_sssp_batch_device_noalloc: null argument (the distance table is required)");
  if (n <= 0) return SQ_OK;
  if (g->topo_dirty || g->w_dirty || !g->d_row_ptr)
    return sq::set_error(sq::SQ_ERR_STATE, "sp_sssp_batch_device_noalloc: the graph has edits that are not on the device yet; "
                                       "call sp_sync first (it allocates, this call must not)");
  const int V = g->V;
  if (V <= 0) return sq::set_error(SQ_ERR_ARG, "graph has no vertices");
  if (V > SQ_NOALLOC_MAX_VERTICES)
    return sq::set_error(SQ_ERR_LIMIT, "sp_sssp_batch_device_noalloc: %d vertices, the static scratch is sized for at most %d",
                         V, SQ_NOALLOC_MAX_VERTICES);
  unsigned char* scratch = nullptr;
  if (int rc = noalloc_scratch_address(&scratch)) return rc;
  int dev = 0, n_sm = 0;
  SQ_CUDA_TRY(cudaGetDevice(&dev));
  SQ_CUDA_TRY(cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev));
  cudaStream_t st = (cudaStream_t)stream;

  const int words = (V + 31) / 32;
  const size_t bits_bytes = (size_t)3 * words * sizeof(unsigned);
  int nq = std::min(2048, std::max(32, V));
  const size_t smem = bits_bytes + (size_t)4 * nq * sizeof(int);      // <= 24 KB + 32 KB at the largest V
  const int threads = 256;
  const size_t per_cta = sizeof(int) * 4 * (size_t)V;
  const int grid = (int)std::min<size_t>({(size_t)n, (size_t)SQ_NOALLOC_QUEUE_BYTES / per_cta, (size_t)n_sm * 4});

  TreeParams p;
  memset(&p, 0, sizeof(p));
  p.V = V; p.n_trees = n;
  p.row_ptr = g->d_row_ptr; p.adj = g->d_adj; p.rrow_ptr = g->d_rrow_ptr; p.radj = g->d_radj;
  p.origins = origins_dev;
  p.dist_out = reinterpret_cast<u64*>(dist_dev);
  p.parent_out = parent_dev;
  p.tree_counter = reinterpret_cast<int*>(scratch);
  p.qscratch = reinterpret_cast<int*>(scratch + SQ_NOALLOC_HEADER_BYTES);
  p.delta = g->delta; p.nq = nq; p.words = words;
  // the array keeps the last call's contents: the counter must start at zero, the queue parts are written before read
  SQ_CUDA_TRY(cudaMemsetAsync(scratch, 0, SQ_NOALLOC_HEADER_BYTES, st));
  return launch_bitmap_kernel(g, p, grid, smem, threads, st);
}

static bool is_pinned_host(const void* ptr) {
  if (!ptr) return true;
  cudaPointerAttributes at;
  if (cudaPointerGetAttributes(&at, ptr) != cudaSuccess) {
    cudaGetLastError();
    return false;
  }
  return at.type == cudaMemoryTypeHost;
}

void* sp_alloc_host(size_t bytes) {
  void* ptr = nullptr;
  if (sq::require_device()) return nullptr;
  if (cudaHostAlloc(&ptr, std::max<size_t>(bytes, 1), cudaHostAllocDefault) != cudaSuccess) {
    sq::set_error(SQ_ERR_CUDA, "sp_alloc_host: cudaHostAlloc(%zu) failed: %s", bytes, cudaGetErrorString(cudaGetLastError()));
    return nullptr;
  }
  return ptr;
}

void sp_free_host(void* ptr) {
  if (ptr) cudaFreeHost(ptr);
}

static int g_ensure_streams(SqGraph* g) {
  if (g->s_compute) return SQ_OK;
  SQ_CUDA_TRY(cudaStreamCreateWithFlags(&g->s_compute, cudaStreamNonBlocking));
  SQ_CUDA_TRY(cudaStreamCreateWithFlags(&g->s_copy, cudaStreamNonBlocking));
  for (int i = 0; i < 2; ++i) {
    SQ_CUDA_TRY(cudaEventCreateWithFlags(&g->ev_done[i], cudaEventDisableTiming));
    SQ_CUDA_TRY(cudaEventCreateWithFlags(&g->ev_copied[i], cudaEventDisableTiming));
  }
  return SQ_OK;
}

int sp_sssp_batch(Graph* g, const int* origins, int n, double* dist_out, int* parent_out) {
  return sp_sssp_batch_ex(g, origins, n, dist_out, parent_out, 0);
}

// Host-buffer batches: trees are produced in chunks into two device buffers; while chunk i is computed on one stream,
// chunk i-1 travels to the host on another.  With page-locked destinations (sp_alloc_host) the copies are asynchronous
// and the call runs at PCIe speed; pageable destinations still work, the driver then stages every copy.
// codes != 0: predecessors leave as one-byte in-edge codes (parent_out is then unsigned char[n][V]): 50 KB per tree on
// config 5 instead of 200 KB, and with dist_out == NULL nothing else crosses PCIe.
static int g_batch_host(SqGraph* g, const int* origins, int n, double* dist_out, void* parent_out, int reverse, int codes) {
  if (!g || (!origins && n > 0)) return sq::set_error(SQ_ERR_ARG, "sp_sssp_batch: null argument");
  if (n <= 0) return SQ_OK;
  if (int rc = g_build_and_upload(g)) return rc;
  const size_t V = (size_t)g->V;
  if (g->origins_cap < n) {
    cudaFree(g->d_origins); g->d_origins = nullptr;
    SQ_CUDA_TRY(cudaMalloc(&g->d_origins, sizeof(int) * (size_t)n));
    g->origins_cap = n;
  }
  SQ_CUDA_TRY(cudaMemcpy(g->d_origins, origins, sizeof(int) * (size_t)n, cudaMemcpyHostToDevice));
  const size_t psize = codes ? 1 : sizeof(int);
  const size_t vh = (V + 1) / 2;   // bytes of a tree's packed codes (codes == 2)
  if (codes == 2) {
    size_t worst = 0;
    const std::vector<int>& rp = reverse ? g->h_rp : g->h_rrp;
    for (size_t v = 0; v < V; ++v) worst = std::max(worst, (size_t)(rp[v + 1] - rp[v]));
    if (worst > 15)
      return sq::set_error(SQ_ERR_LIMIT, "four-bit predecessor codes need in-degree <= 15 (found %zu); use sp_sssp_batch_codes", worst);
  }
  const size_t per_tree = V * ((dist_out ? sizeof(u64) : 0) + (parent_out ? psize : 0));
  size_t chunk = std::max<size_t>(1, std::min<size_t>((size_t)n, ((size_t)1 << 29) / std::max<size_t>(per_tree, 1)));
  // at least four chunks on large batches (the copy of one hides behind the trees of the next), but never so small that
  // a chunk leaves SMs idle
  chunk = std::min(chunk, std::max<size_t>(((size_t)n + 3) / 4, 2368));
  if ((size_t)n > chunk) chunk = std::max<size_t>(chunk, 64);
  if (dist_out && g->dist_buf_elems < 2 * chunk * V) {
    cudaFree(g->d_dist_buf); g->d_dist_buf = nullptr;
    SQ_CUDA_TRY(cudaMalloc(&g->d_dist_buf, sizeof(u64) * 2 * chunk * V));
    g->dist_buf_elems = 2 * chunk * V;
  }
  const size_t pelems = (2 * chunk * V * psize + sizeof(int) - 1) / sizeof(int);   // in ints
  if (parent_out && g->parent_buf_elems < pelems) {
    cudaFree(g->d_parent_buf); g->d_parent_buf = nullptr;
    SQ_CUDA_TRY(cudaMalloc(&g->d_parent_buf, sizeof(int) * pelems));
    g->parent_buf_elems = pelems;
  }
  if (codes == 2 && g->pack_buf_bytes < 2 * chunk * vh) {
    cudaFree(g->d_pack_buf); g->d_pack_buf = nullptr;
    SQ_CUDA_TRY(cudaMalloc(&g->d_pack_buf, 2 * chunk * vh));
    g->pack_buf_bytes = 2 * chunk * vh;
  }
  if (int rc = g_ensure_streams(g)) return rc;
  const bool pinned = is_pinned_host(dist_out) && is_pinned_host(parent_out);
  int k = 0;
  for (size_t off = 0; off < (size_t)n; off += chunk, ++k) {
    const int m = (int)std::min(chunk, (size_t)n - off);
    const int b = k & 1;
    u64* dbuf = dist_out ? g->d_dist_buf + (size_t)b * chunk * V : nullptr;
    unsigned char* pbuf = parent_out ? reinterpret_cast<unsigned char*>(g->d_parent_buf) + (size_t)b * chunk * V * psize : nullptr;
    if (k >= 2) SQ_CUDA_TRY(cudaStreamWaitEvent(g->s_compute, g->ev_copied[b], 0));  // buffer b is free again
    const int fmt = g->parent_format;
    g->parent_format = codes ? 1 : 0;
    const int rc_run = g_run_batch_device(g, g->d_origins + off, m, dbuf, reinterpret_cast<int*>(pbuf), g->s_compute, reverse != 0);
    g->parent_format = fmt;
    if (rc_run) return rc_run;
    unsigned char* qbuf = nullptr;
    if (codes == 2) {
      qbuf = g->d_pack_buf + (size_t)b * chunk * vh;
      const size_t total = vh * (size_t)m;
      pack_codes4_kernel<<<(unsigned)((total + 255) / 256), 256, 0, g->s_compute>>>((int)V, m, pbuf, qbuf);
      sq::count_launch();
      SQ_CUDA_TRY(cudaGetLastError());
    }
    SQ_CUDA_TRY(cudaEventRecord(g->ev_done[b], g->s_compute));
    SQ_CUDA_TRY(cudaStreamWaitEvent(g->s_copy, g->ev_done[b], 0));
    if (!pinned) SQ_CUDA_TRY(cudaStreamSynchronize(g->s_compute));  // pageable copies block the host anyway
    if (dist_out)
      SQ_CUDA_TRY(cudaMemcpyAsync(dist_out + off * V, dbuf, sizeof(u64) * (size_t)m * V, cudaMemcpyDeviceToHost, g->s_copy));
    if (parent_out && codes == 2)
      SQ_CUDA_TRY(cudaMemcpyAsync(reinterpret_cast<unsigned char*>(parent_out) + off * vh, qbuf, (size_t)m * vh,
                                  cudaMemcpyDeviceToHost, g->s_copy));
    else if (parent_out)
      SQ_CUDA_TRY(cudaMemcpyAsync(reinterpret_cast<unsigned char*>(parent_out) + off * V * psize, pbuf, psize * (size_t)m * V,
                                  cudaMemcpyDeviceToHost, g->s_copy));
    SQ_CUDA_TRY(cudaEventRecord(g->ev_copied[b], g->s_copy));
  }
  SQ_CUDA_TRY(cudaStreamSynchronize(g->s_compute));
  SQ_CUDA_TRY(cudaStreamSynchronize(g->s_copy));
  g->last_stats[0] = n;
  return g_fetch_stats(g);
}

int sp_sssp_batch_ex(Graph* g, const int* origins, int n, double* dist_out, int* parent_out, int reverse) {
  return g_batch_host(g, origins, n, dist_out, parent_out, reverse, 0);
}

int sp_sssp_batch_codes(Graph* g, const int* origins, int n, double* dist_out, unsigned char* codes_out, int reverse) {
  if (!codes_out) return sq::set_error(SQ_ERR_ARG, "sp_sssp_batch_codes: null codes_out");
  return g_batch_host(g, origins, n, dist_out, codes_out, reverse, 1);
}

int sp_sssp_batch_codes4(Graph* g, const int* origins, int n, double* dist_out, unsigned char* codes_out, int reverse) {
  if (!codes_out) return sq::set_error(SQ_ERR_ARG, "sp_sssp_batch_codes4: null codes_out");
  return g_batch_host(g, origins, n, dist_out, codes_out, reverse, 2);
}

// Four-bit codes -> vertex ids on the host: codes[n][(V + 1) / 2], low nibble = even vertex, 15 = none.
int sp_decode_parents_host4(Graph* g, const unsigned char* codes, int n, int* parent_out, int reverse) {
  if (!g || !codes || !parent_out) return sq::set_error(SQ_ERR_ARG, "sp_decode_parents_host4: null argument");
  if (int rc = g_build_and_upload(g)) return rc;
  const std::vector<int>& rp = reverse ? g->h_rp : g->h_rrp;
  const std::vector<EdgeRec>& adj = reverse ? g->h_adj : g->h_radj;
  const size_t V = (size_t)g->V, vh = (V + 1) / 2;
  for (int i = 0; i < n; ++i) {
    const unsigned char* c = codes + (size_t)i * vh;
    int* p = parent_out + (size_t)i * V;
    for (size_t v = 0; v < V; ++v) {
      const int code = (c[v >> 1] >> ((v & 1) * 4)) & 15;
      p[v] = (code == 15 || rp[v] + code >= rp[v + 1]) ? -1 : adj[(size_t)rp[v] + code].v;
    }
  }
  return SQ_OK;
}

// One-byte in-edge codes -> vertex ids on the host (what sp_decode_parents_device does on the GPU): code c of vertex v
// is the c-th in-edge of v in creation order (for reverse trees: the c-th out-edge), 255 = none.
int sp_decode_parents_host(Graph* g, const unsigned char* codes, int n, int* parent_out, int reverse) {
  if (!g || !codes || !parent_out) return sq::set_error(SQ_ERR_ARG, "sp_decode_parents_host: null argument");
  if (int rc = g_build_and_upload(g)) return rc;
  const std::vector<int>& rp = reverse ? g->h_rp : g->h_rrp;
  const std::vector<EdgeRec>& adj = reverse ? g->h_adj : g->h_radj;
  const size_t V = (size_t)g->V;
  for (int i = 0; i < n; ++i) {
    const unsigned char* c = codes + (size_t)i * V;
    int* p = parent_out + (size_t)i * V;
    for (size_t v = 0; v < V; ++v) {
      const int code = c[v];
      p[v] = (code == 255 || rp[v] + code >= rp[v + 1]) ? -1 : adj[(size_t)rp[v] + code].v;
    }
  }
  return SQ_OK;
}

int sp_route_batch(Graph* g, const int* origins, const int* dests, int n, double* dist_out, long long* route_off_out,
                   int* route_len_out) {
  if (!g || ((!origins || !dests || !dist_out || !route_off_out || !route_len_out) && n > 0))
    return sq::set_error(SQ_ERR_ARG, "sp_route_batch: null argument");
  if (n <= 0) return SQ_OK;
  if (int rc = g_build_and_upload(g)) return rc;
  const int Vi = g->V;
  const size_t V = (size_t)Vi;
  for (int k = 0; k < n; ++k)
    if (dests[k] < 0 || dests[k] >= Vi) return sq::set_error(SQ_ERR_ARG, "sp_route_batch: destination out of range");
  // one tree per distinct origin; pairs sorted by tree so that a chunk of trees serves a contiguous run of pairs
  std::vector<int> uo(origins, origins + n);
  std::sort(uo.begin(), uo.end());
  uo.erase(std::unique(uo.begin(), uo.end()), uo.end());
  const int nt = (int)uo.size();
  std::vector<int> order((size_t)n), ptree((size_t)n), pdest((size_t)n);
  {
    std::vector<int> tree_of((size_t)n);
    for (int k = 0; k < n; ++k) tree_of[k] = (int)(std::lower_bound(uo.begin(), uo.end(), origins[k]) - uo.begin());
    std::iota(order.begin(), order.end(), 0);
    std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return tree_of[a] < tree_of[b]; });
    for (int i = 0; i < n; ++i) { ptree[i] = tree_of[order[i]]; pdest[i] = dests[order[i]]; }
  }
  if (int rc = g_ensure_streams(g)) return rc;
  cudaStream_t st = g->s_compute;
  if (g->origins_cap < nt) {
    cudaFree(g->d_origins); g->d_origins = nullptr;
    SQ_CUDA_TRY(cudaMalloc(&g->d_origins, sizeof(int) * (size_t)nt));
    g->origins_cap = nt;
  }
  if (g->pair_cap < (size_t)n) {
    cudaFree(g->d_pair); cudaFree(g->d_pair_off); cudaFree(g->d_pair_dist);
    g->d_pair = nullptr; g->d_pair_off = nullptr; g->d_pair_dist = nullptr; g->pair_cap = 0;
    SQ_CUDA_TRY(cudaMalloc(&g->d_pair, sizeof(int) * 3 * (size_t)n));
    SQ_CUDA_TRY(cudaMalloc(&g->d_pair_off, sizeof(long long) * (size_t)n));
    SQ_CUDA_TRY(cudaMalloc(&g->d_pair_dist, sizeof(double) * (size_t)n));
    g->pair_cap = (size_t)n;
  }
  int* d_pdest = g->d_pair;
  int* d_ptree = g->d_pair + g->pair_cap;
  int* d_plen = g->d_pair + 2 * g->pair_cap;
  SQ_CUDA_TRY(cudaMemcpyAsync(g->d_origins, uo.data(), sizeof(int) * (size_t)nt, cudaMemcpyHostToDevice, st));
  SQ_CUDA_TRY(cudaMemcpyAsync(d_pdest, pdest.data(), sizeof(int) * (size_t)n, cudaMemcpyHostToDevice, st));
  SQ_CUDA_TRY(cudaMemcpyAsync(d_ptree, ptree.data(), sizeof(int) * (size_t)n, cudaMemcpyHostToDevice, st));
  // trees per chunk: as many as a third of the free memory holds (the tree kernels want thousands per launch)
  const size_t per_tree = V * (sizeof(u64) + sizeof(int));
  size_t free_b = 0, total_b = 0;
  SQ_CUDA_TRY(cudaMemGetInfo(&free_b, &total_b));
  const size_t have = std::min(g->dist_buf_elems, g->parent_buf_elems) / std::max<size_t>(V, 1);   // trees the buffers hold already
  size_t chunk = std::min<size_t>((size_t)nt, std::max(have, std::max<size_t>(64, (free_b / 3) / std::max<size_t>(per_tree, 1))));
  if (const char* cap = getenv("SQB200_ROUTE_CHUNK")) chunk = std::min(chunk, (size_t)std::max(1, atoi(cap)));   // tests: force several chunks
  if (g->dist_buf_elems < chunk * V) {
    cudaFree(g->d_dist_buf); g->d_dist_buf = nullptr; g->dist_buf_elems = 0;
    SQ_CUDA_TRY(cudaMalloc(&g->d_dist_buf, sizeof(u64) * chunk * V));
    g->dist_buf_elems = chunk * V;
  }
  if (g->parent_buf_elems < chunk * V) {
    cudaFree(g->d_parent_buf); g->d_parent_buf = nullptr; g->parent_buf_elems = 0;
    SQ_CUDA_TRY(cudaMalloc(&g->d_parent_buf, sizeof(int) * chunk * V));
    g->parent_buf_elems = chunk * V;
  }
  std::vector<int> len((size_t)n);
  std::vector<long long> off((size_t)n);
  std::vector<double> dist((size_t)n);
  long long pool_used = 0;
  int p0 = 0;
  for (size_t t0 = 0; t0 < (size_t)nt; t0 += chunk) {
    const int m = (int)std::min(chunk, (size_t)nt - t0);
    int p1 = p0;
    while (p1 < n && ptree[p1] < (int)t0 + m) ++p1;
    const int np = p1 - p0;
    const int fmt = g->parent_format;
    g->parent_format = 0;
    const int rc_run = g_run_batch_device(g, g->d_origins + t0, m, g->d_dist_buf, g->d_parent_buf, st, false);
    g->parent_format = fmt;
    if (rc_run) return rc_run;
    if (np > 0) {
      const int blocks = (np + 127) / 128;
      route_len_kernel<<<blocks, 128, 0, st>>>(g->d_dist_buf, g->d_parent_buf, g->d_origins, Vi, (int)t0, np, d_pdest + p0,
                                              d_ptree + p0, d_plen + p0, g->d_pair_dist + p0);
      sq::count_launch();
      SQ_CUDA_TRY(cudaMemcpyAsync(len.data() + p0, d_plen + p0, sizeof(int) * (size_t)np, cudaMemcpyDeviceToHost, st));
      SQ_CUDA_TRY(cudaMemcpyAsync(dist.data() + p0, g->d_pair_dist + p0, sizeof(double) * (size_t)np, cudaMemcpyDeviceToHost, st));
      SQ_CUDA_TRY(cudaStreamSynchronize(st));
      const long long pool0 = pool_used;
      for (int i = p0; i < p1; ++i) { off[i] = pool_used; pool_used += len[i]; }
      const size_t need = (size_t)(pool_used - pool0);
      if (need > 0) {
        if (g->d_pool_cap < need) {
          cudaFree(g->d_pool); g->d_pool = nullptr; g->d_pool_cap = 0;
          SQ_CUDA_TRY(cudaMalloc(&g->d_pool, sizeof(int) * need));
          g->d_pool_cap = need;
        }
        if (g->h_pool_cap < (size_t)pool_used) {
          const size_t cap = std::max<size_t>((size_t)pool_used, std::max<size_t>(2 * g->h_pool_cap, 1 << 16));
          int* nb = nullptr;
          SQ_CUDA_TRY(cudaMallocHost(&nb, sizeof(int) * cap));
          if (g->h_pool) { memcpy(nb, g->h_pool, sizeof(int) * (size_t)pool0); cudaFreeHost(g->h_pool); }
          g->h_pool = nb; g->h_pool_cap = cap;
        }
        SQ_CUDA_TRY(cudaMemcpyAsync(g->d_pair_off + p0, off.data() + p0, sizeof(long long) * (size_t)np, cudaMemcpyHostToDevice, st));
        route_write_kernel<<<blocks, 128, 0, st>>>(g->d_parent_buf, Vi, (int)t0, np, d_pdest + p0, d_ptree + p0, d_plen + p0,
                                                  g->d_pair_off + p0, pool0, g->d_pool);
        sq::count_launch();
        SQ_CUDA_TRY(cudaMemcpyAsync(g->h_pool + pool0, g->d_pool, sizeof(int) * need, cudaMemcpyDeviceToHost, st));
        SQ_CUDA_TRY(cudaStreamSynchronize(st));
      }
    }
    p0 = p1;
  }
  SQ_CUDA_TRY(cudaStreamSynchronize(st));
  for (int i = 0; i < n; ++i) {
    const int k = order[i];
    dist_out[k] = dist[i]; route_off_out[k] = off[i]; route_len_out[k] = len[i];
  }
  g->last_stats[0] = nt;
  return g_fetch_stats(g);
}

const int* sp_route_pool(Graph* g) { return g ? g->h_pool : nullptr; }

ShortestPath* dijkstra(Graph* g, int origin_v, int destination) {
  SqShortestPath* sp = new SqShortestPath();
  sp->source = origin_v;
  if (!g) return sp;
  const long long nv = (long long)g->nvertices;
  // graph.cc:161-163 (the reference's guard uses '>' and then aborts in .at(); '>=' here)
  if (origin_v < 0 || origin_v >= nv || ((destination < 0 || destination >= nv) && destination != -1)) return sp;
  const size_t V = (size_t)nv;
  sp->distances.resize(V);
  sp->parent.resize(V);
  const int rc = sp_sssp_batch(g, &origin_v, 1, sp->distances.data(), sp->parent.data());
  if (rc != SQ_OK) {
    // no error channel in the reference signature and no CPU fallback: fail loudly
    fprintf(stderr, "spatial_queue_b200: dijkstra() failed (%d): %s\n", rc, sq::err_buf());
    abort();
  }
  return sp;
}

int origin(ShortestPath* sp) { return sp ? sp->source : -1; }

double distance(ShortestPath* sp, int destination) {
  if (!sp || sp->distances.empty() || destination < 0 || (size_t)destination >= sp->distances.size()) return DBL_MAX;
  return sp->distances[(size_t)destination];
}

void clear(ShortestPath* sp) { delete sp; }

int parent(ShortestPath* sp, int destination) {
  if (!sp || destination < 0 || (size_t)destination >= sp->parent.size()) return -1;
  return sp->parent[(size_t)destination];
}

}  // extern "C"
