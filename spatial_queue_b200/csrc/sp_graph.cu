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
#include <sstream>
#include <string>
#include <unordered_map>
#include <vector>

#include "../../include/sq_sp.h"
#include "sq_common.cuh"

using sq::SQ_ERR_ARG;
using sq::SQ_ERR_CUDA;
using sq::SQ_ERR_LIMIT;
using sq::SQ_OK;

typedef unsigned long long u64;

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
  const EdgeRec* ell;     // [V][4] forward, slot 0/1 .pad = overflow start / count
  const EdgeRec* ovf;     // forward edges beyond the fourth of a vertex
  const EdgeRec* rell;    // [V][4] reverse
  const EdgeRec* rovf;
  int2* q2scratch;        // [grid][2*near_cap + 2*far_cap] (vertex, low 32 bits of its distance when pushed)
  int near_cap, far_cap;
  int* fail_list;         // trees whose queues overflowed (rerun by the bitmap kernel)
  int* fail_count;
};

#define DIST_INF_BITS 0x7FEFFFFFFFFFFFFFull  // DBL_MAX, what the reference leaves in unreached slots

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
          T = Told + p.delta;
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
      for (int v = tid; v < V; v += THREADS) {
        int best = -1;
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
              if ((u64)__double_as_longlong(cand) == dvb && er.v < best) best = er.v;
            }
            if (best == INT_MAX) best = -1;
          }
        }
        parent[v] = best;
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
// ELL kernel: the same near-far schedule without per-vertex bitmaps and without row pointers.
//
//  * adjacency = one 64-byte record of four 16-byte edges per vertex (road networks rarely exceed
//    degree 4; further edges sit in an overflow list).  Four adjacent lanes own one frontier vertex, so a
//    warp step touches 8 lines for its 32 edges instead of 32, and there is no row_ptr load in the chain.
//  * queue entries are (vertex, low 32 bits of its distance when pushed).  An entry whose tag no longer
//    matches dist[vertex] is stale: the thread that lowered the distance pushed a newer entry.  This
//    replaces the three shared-memory bitmaps (and their ATOMS traffic); duplicates are possible, so
//    queues are bounded by a capacity and a tree that overflows is rerun by the bitmap kernel.
//  * queues live in a per-CTA global scratch row (L2-resident, read with ld.cg, written coalesced);
//    shared memory only holds the four counters, so many trees fit on an SM.
// -------------------------------------------------------------------------------------------------
__device__ __forceinline__ int2 ld_cg_int2(const int2* p) {
  int2 v;
  asm volatile("ld.global.cg.v2.s32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "l"(p));
  return v;
}

// Append from any subset of a warp's lanes to a CTA-wide queue: one shared atomicAdd per warp.
__device__ __forceinline__ void cta_push(bool want, int v, unsigned tag, int2* q, int* cnt, int cap, int* fail) {
  const unsigned m = __ballot_sync(0xffffffffu, want);
  if (m == 0) return;
  const int leader = __ffs(m) - 1;
  int base = 0;
  if ((int)sq::lane_id() == leader) base = atomicAdd(cnt, __popc(m));
  base = __shfl_sync(0xffffffffu, base, leader);
  if (want) {
    const int pos = base + __popc(m & sq::lanemask_lt());
    if (pos < cap) q[pos] = make_int2(v, (int)tag);
    else *fail = 1;
  }
}

template <int NW>
__global__ void __launch_bounds__(NW * 32) sssp_ell_kernel(const TreeParams p) {
  constexpr int THREADS = NW * 32;
  __shared__ int s_cnt[4];  // near0, near1, far0, far1
  __shared__ int s_tree, s_fail;
  __shared__ u64 s_min;

  const int tid = threadIdx.x;
  const int lane = tid & 31, warp = tid >> 5;
  const int slot = lane & 3, grp = lane >> 2;
  const int V = p.V;
  int2* qbase = p.q2scratch + (size_t)blockIdx.x * (2 * (size_t)p.near_cap + 2 * (size_t)p.far_cap);
  auto NEAR = [&](int i) -> int2* { return qbase + (size_t)i * p.near_cap; };
  auto FAR = [&](int i) -> int2* { return qbase + 2 * (size_t)p.near_cap + (size_t)i * p.far_cap; };

  long long st_rounds = 0, st_relax = 0, st_trans = 0;

  for (;;) {
    if (tid == 0) s_tree = atomicAdd(p.tree_counter, 1);
    __syncthreads();
    const int tree = s_tree;
    if (tree >= p.n_trees) break;

    u64* dist = p.dist_out ? p.dist_out + (size_t)tree * V : p.dist_scratch + (size_t)blockIdx.x * V;
    const int src = p.origins[tree];
    const bool valid = (src >= 0 && src < V);

    // dist rows are 8-byte aligned; use 16-byte stores on the aligned middle
    {
      const int head = (int)(((uintptr_t)dist >> 3) & 1);  // 1 if the row starts on an odd 8-byte slot
      if (tid == 0 && head && V > 0) dist[0] = DIST_INF_BITS;
      const int pairs = (V - head) >> 1;
      ulonglong2* d2 = reinterpret_cast<ulonglong2*>(dist + head);
      const ulonglong2 inf2 = make_ulonglong2(DIST_INF_BITS, DIST_INF_BITS);
      for (int i = tid; i < pairs; i += THREADS) d2[i] = inf2;
      if (tid == 0 && ((V - head) & 1)) dist[V - 1] = DIST_INF_BITS;
    }
    if (tid < 4) s_cnt[tid] = 0;
    if (tid == 0) s_fail = 0;
    __syncthreads();
    if (tid == 0 && valid) {
      dist[src] = 0ull;
      NEAR(0)[0] = make_int2(src, 0);
      s_cnt[0] = 1;
    }
    __syncthreads();

    double T = p.delta;
    int cur = 0, fcur = 0;

    while (valid) {
      const int n = min(s_cnt[cur], p.near_cap);
      if (s_fail) break;
      if (n == 0) {
        // ---- bucket transition: split the far pile against the next threshold --------------------
        int nf = min(s_cnt[2 + fcur], p.far_cap);
        if (nf == 0) break;
        for (;;) {
          __syncthreads();
          if (tid == 0) { s_min = DIST_INF_BITS; s_cnt[2 + (fcur ^ 1)] = 0; }
          __syncthreads();
          const double Told = T;
          T = Told + p.delta;
          u64 lmin = DIST_INF_BITS;
          const int2* fin = FAR(fcur);
          int2* fout = FAR(fcur ^ 1);
          int2* nout = NEAR(cur);
          for (int i0 = 0; i0 < nf; i0 += THREADS) {
            const int i = i0 + tid;
            bool to_near = false, to_far = false;
            int2 ent = make_int2(0, 0);
            if (i < nf) {
              ent = ld_cg_int2(&fin[i]);
              const u64 db = sq::ld_cg_u64(&dist[ent.x]);
              if ((unsigned)db == (unsigned)ent.y) {  // else stale: a newer entry exists
                if (__longlong_as_double((long long)db) < T) to_near = true;
                else { to_far = true; lmin = db < lmin ? db : lmin; }
              }
            }
            cta_push(to_near, ent.x, (unsigned)ent.y, nout, &s_cnt[cur], p.near_cap, &s_fail);
            cta_push(to_far, ent.x, (unsigned)ent.y, fout, &s_cnt[2 + (fcur ^ 1)], p.far_cap, &s_fail);
          }
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) {
            const u64 other = __shfl_xor_sync(0xffffffffu, lmin, o);
            lmin = other < lmin ? other : lmin;
          }
          if (lane == 0 && lmin != DIST_INF_BITS) atomicMin(&s_min, lmin);
          __syncthreads();
          fcur ^= 1;
          nf = min(s_cnt[2 + fcur], p.far_cap);
          ++st_trans;
          if (s_cnt[cur] > 0 || nf == 0 || s_fail) break;
          T = __longlong_as_double((long long)s_min);  // nothing in range: jump to the smallest far distance
        }
        continue;
      }

      // ---- one relaxation round: 4 lanes per frontier vertex, 8 vertices per warp step -----------
      const int2* qin = NEAR(cur);
      int2* qnear = NEAR(cur ^ 1);
      int2* qfar = FAR(fcur);
      int* cnt_near = &s_cnt[cur ^ 1];
      int* cnt_far = &s_cnt[2 + fcur];
      for (int base = warp * 8; base < n; base += NW * 8) {
        const int idx = base + grp;
        bool act = idx < n;
        int u = 0;
        u64 dub = 0ull;
        if (act) {
          const int2 ent = ld_cg_int2(&qin[idx]);
          u = ent.x;
          dub = sq::ld_cg_u64(&dist[u]);
          act = ((unsigned)dub == (unsigned)ent.y);
        }
        EdgeRec rec;
        rec.w = 0.0; rec.v = -1; rec.pad = 0;
        if (act) {
          const int4 raw = __ldg(reinterpret_cast<const int4*>(&p.ell[(size_t)u * 4 + slot]));
          rec.w = __hiloint2double(raw.y, raw.x);
          rec.v = raw.z;
          rec.pad = raw.w;
        }
        const double du = __longlong_as_double((long long)dub);
        {
          const bool has = rec.v >= 0;
          const u64 ndb = (u64)__double_as_longlong(du + rec.w);
          u64 old = 0ull;
          if (has) old = atomicMin(&dist[rec.v], ndb);
          const bool imp = has && ndb < old;
          const bool is_near = __longlong_as_double((long long)ndb) < T;
          cta_push(imp && is_near, rec.v, (unsigned)ndb, qnear, cnt_near, p.near_cap, &s_fail);
          cta_push(imp && !is_near, rec.v, (unsigned)ndb, qfar, cnt_far, p.far_cap, &s_fail);
          st_relax += has;
        }
        // edges beyond the fourth (rare on road networks)
        const int ovf_b = __shfl_sync(0xffffffffu, rec.pad, lane & ~3);
        const int ovf_n = __shfl_sync(0xffffffffu, rec.pad, (lane & ~3) + 1);
        const int max_o = __reduce_max_sync(0xffffffffu, act ? ovf_n : 0);
        for (int k = slot; k - slot < max_o; k += 4) {
          const bool has = act && k < ovf_n;
          int v2 = 0;
          u64 ndb = DIST_INF_BITS, old = 0ull;
          if (has) {
            const EdgeRec er = ld_edge(&p.ovf[ovf_b + k]);
            v2 = er.v;
            ndb = (u64)__double_as_longlong(du + er.w);
            old = atomicMin(&dist[v2], ndb);
          }
          const bool imp = has && ndb < old;
          const bool is_near = __longlong_as_double((long long)ndb) < T;
          cta_push(imp && is_near, v2, (unsigned)ndb, qnear, cnt_near, p.near_cap, &s_fail);
          cta_push(imp && !is_near, v2, (unsigned)ndb, qfar, cnt_far, p.far_cap, &s_fail);
          st_relax += has;
        }
      }
      __syncthreads();
      if (tid == 0) s_cnt[cur] = 0;
      cur ^= 1;
      ++st_rounds;
      __syncthreads();
    }
    __syncthreads();

    if (s_fail) {
      // queue capacity exceeded: hand the tree to the bitmap kernel (its queues cannot overflow)
      if (tid == 0) p.fail_list[atomicAdd(p.fail_count, 1)] = tree;
      __syncthreads();
      continue;
    }

    // ---- canonical predecessor pass: parent[v] = min{u : fl(d[u] + w(u,v)) == d[v]} -------------
    if (p.parent_out) {
      int* parent = p.parent_out + (size_t)tree * V;
      for (int v0 = 0; v0 < V; v0 += THREADS / 4) {
        const int v = v0 + (tid >> 2);
        const bool ok = valid && v < V && v != src;
        u64 dvb = DIST_INF_BITS;
        EdgeRec rec;
        rec.w = 0.0; rec.v = -1; rec.pad = 0;
        if (ok) {
          dvb = sq::ld_cg_u64(&dist[v]);
          if (dvb != DIST_INF_BITS) {
            const int4 raw = __ldg(reinterpret_cast<const int4*>(&p.rell[(size_t)v * 4 + slot]));
            rec.w = __hiloint2double(raw.y, raw.x);
            rec.v = raw.z;
            rec.pad = raw.w;
          }
        }
        int best = INT_MAX;
        if (rec.v >= 0) {
          const u64 dub = sq::ld_cg_u64(&dist[rec.v]);
          if (dub != DIST_INF_BITS) {
            const double cand = __longlong_as_double((long long)dub) + rec.w;
            if ((u64)__double_as_longlong(cand) == dvb) best = rec.v;
          }
        }
        const int ovf_b = __shfl_sync(0xffffffffu, rec.pad, lane & ~3);
        const int ovf_n = __shfl_sync(0xffffffffu, rec.pad, (lane & ~3) + 1);
        const int max_o = __reduce_max_sync(0xffffffffu, ovf_n);
        for (int k = slot; k - slot < max_o; k += 4) {
          if (k < ovf_n) {
            const EdgeRec er = ld_edge(&p.rovf[ovf_b + k]);
            const u64 dub = sq::ld_cg_u64(&dist[er.v]);
            if (dub != DIST_INF_BITS) {
              const double cand = __longlong_as_double((long long)dub) + er.w;
              if ((u64)__double_as_longlong(cand) == dvb && er.v < best) best = er.v;
            }
          }
        }
        best = min(best, __shfl_xor_sync(0xffffffffu, best, 1));
        best = min(best, __shfl_xor_sync(0xffffffffu, best, 2));
        if (slot == 0 && v < V) parent[v] = (best == INT_MAX) ? -1 : best;
      }
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
  int2* d_q2scratch = nullptr;
  int q2_grid = 0, q2_near = 0, q2_far = 0;
  int* d_fail = nullptr;       // [0] count, [1] fallback tree counter, [2..] list
  int fail_cap = 0;
  int algo = 0;                // 0 = ELL kernel (bitmap kernel only as its fallback), 1 = bitmap kernel
  double queue_factor = 1.0;   // queue capacities of the ELL kernel: near = factor*V, far = 2*factor*V

  // tuning
  int threads = 128;
  int ctas_per_sm = 0;
  double delta_factor = 2.0;
  double delta = 1.0;
  int smem_queue = 0;
  long long last_stats[5] = {0, 0, 0, 0, 0};
};

struct SqShortestPath {
  int source = -1;
  std::vector<double> distances;
  std::vector<int> parent;
};

static void g_add_edge(SqGraph* g, int v1, int v2, double w) {
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
  g->d_row_ptr = g->d_rrow_ptr = nullptr; g->d_adj = g->d_radj = nullptr;
  g->d_ell = g->d_rell = g->d_ovf = g->d_rovf = nullptr;
  g->ovf_cap = g->rovf_cap = 0;
}

// CSR -> four-wide ELL + overflow list.  Slot 0 / slot 1 of a vertex carry the overflow start / count in .pad.
static void build_ell(const std::vector<int>& rp, const std::vector<EdgeRec>& adj, int V, std::vector<EdgeRec>& ell,
                      std::vector<EdgeRec>& ovf) {
  ell.assign((size_t)4 * std::max(V, 1), EdgeRec{0.0, -1, 0});
  ovf.clear();
  for (int v = 0; v < V; ++v) {
    const int b = rp[v], deg = rp[v + 1] - rp[v];
    for (int j = 0; j < std::min(deg, 4); ++j) {
      EdgeRec r = adj[b + j];
      r.pad = 0;
      ell[(size_t)4 * v + j] = r;
    }
    if (deg > 4) {
      ell[(size_t)4 * v + 0].pad = (int)ovf.size();
      ell[(size_t)4 * v + 1].pad = deg - 4;
      for (int j = 4; j < deg; ++j) ovf.push_back(adj[b + j]);
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
    SQ_CUDA_TRY(cudaMalloc(&g->d_row_ptr, sizeof(int) * (size_t)(V + 1)));
    SQ_CUDA_TRY(cudaMalloc(&g->d_rrow_ptr, sizeof(int) * (size_t)(V + 1)));
    SQ_CUDA_TRY(cudaMalloc(&g->d_adj, sizeof(EdgeRec) * (size_t)std::max(E, 1)));
    SQ_CUDA_TRY(cudaMalloc(&g->d_radj, sizeof(EdgeRec) * (size_t)std::max(E, 1)));
    SQ_CUDA_TRY(cudaMemcpy(g->d_row_ptr, rp.data(), sizeof(int) * (size_t)(V + 1), cudaMemcpyHostToDevice));
    SQ_CUDA_TRY(cudaMemcpy(g->d_rrow_ptr, rrp.data(), sizeof(int) * (size_t)(V + 1), cudaMemcpyHostToDevice));
    if (!g->d_counter) SQ_CUDA_TRY(cudaMalloc(&g->d_counter, sizeof(int)));
    if (!g->d_stats) SQ_CUDA_TRY(cudaMalloc(&g->d_stats, sizeof(long long) * 5));
    g->topo_dirty = false;
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
  static bool attr_set = false;
  if (!attr_set) {
    SQ_CUDA_TRY(cudaFuncSetAttribute(sssp_nearfar_kernel<THREADS>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    attr_set = true;
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

static int launch_ell_kernel(TreeParams& p, int grid, int threads, cudaStream_t st) {
  switch (threads) {
    case 32: sssp_ell_kernel<1><<<grid, 32, 0, st>>>(p); break;
    case 64: sssp_ell_kernel<2><<<grid, 64, 0, st>>>(p); break;
    case 128: sssp_ell_kernel<4><<<grid, 128, 0, st>>>(p); break;
    case 256: sssp_ell_kernel<8><<<grid, 256, 0, st>>>(p); break;
    default: return sq::set_error(SQ_ERR_ARG, "ELL kernel: threads must be 32, 64, 128 or 256 (got %d)", threads);
  }
  sq::count_launch();
  SQ_CUDA_TRY(cudaGetLastError());
  return SQ_OK;
}

// Runs n trees; origins_dev / outputs are device pointers (outputs may be null).
static int g_run_batch_device(SqGraph* g, const int* origins_dev, int n, u64* dist_dev, int* parent_dev,
                              cudaStream_t st) {
  if (n <= 0) return SQ_OK;
  if (int rc = g_build_and_upload(g)) return rc;
  const int V = g->V;
  if (V <= 0) return sq::set_error(SQ_ERR_ARG, "graph has no vertices");
  static cudaDeviceProp prop;
  static int prop_dev = -1;
  int dev = 0;
  SQ_CUDA_TRY(cudaGetDevice(&dev));
  if (prop_dev != dev) {
    SQ_CUDA_TRY(cudaGetDeviceProperties(&prop, dev));
    prop_dev = dev;
  }
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
  if (g->fail_cap < n + 2) {
    cudaFree(g->d_fail); g->d_fail = nullptr;
    SQ_CUDA_TRY(cudaMalloc(&g->d_fail, sizeof(int) * (size_t)(n + 2)));
    g->fail_cap = n + 2;
  }
  SQ_CUDA_TRY(cudaMemsetAsync(g->d_counter, 0, sizeof(int), st));
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

  if (g->algo == 1) {
    if (int rc = launch_bitmap_kernel(g, p, bm_grid, smem, bm_threads, st)) return rc;
    g->last_stats[0] = n;
    return SQ_OK;
  }

  // ---- ELL kernel -------------------------------------------------------------------------------
  const int threads = std::min(g->threads, 256);
  int per_sm = g->ctas_per_sm > 0 ? g->ctas_per_sm : 12;
  per_sm = std::max(1, std::min(per_sm, std::min(32, 2048 / threads)));
  const int grid = std::min(n, n_sm * per_sm);
  const int near_cap = std::max(32, (int)(g->queue_factor * V));
  const int far_cap = 2 * near_cap;
  if (g->q2_grid < grid || g->q2_near != near_cap) {
    cudaFree(g->d_q2scratch); g->d_q2scratch = nullptr;
    SQ_CUDA_TRY(cudaMalloc(&g->d_q2scratch, sizeof(int2) * (size_t)grid * (2 * (size_t)near_cap + 2 * (size_t)far_cap)));
    g->q2_grid = grid; g->q2_near = near_cap; g->q2_far = far_cap;
  }
  p.q2scratch = g->d_q2scratch; p.near_cap = near_cap; p.far_cap = far_cap;
  if (!dist_dev) {
    // timing-only mode keeps distances in a per-CTA scratch row
    if (g->scratch_grid < grid || g->scratch_V < V || !g->d_dist_scratch) {
      cudaFree(g->d_dist_scratch); cudaFree(g->d_qscratch);
      g->d_dist_scratch = nullptr; g->d_qscratch = nullptr;
      SQ_CUDA_TRY(cudaMalloc(&g->d_dist_scratch, sizeof(u64) * (size_t)grid * V));
      SQ_CUDA_TRY(cudaMalloc(&g->d_qscratch, sizeof(int) * (size_t)grid * 4 * V));
      g->scratch_grid = grid; g->scratch_V = V;
      p.dist_scratch = g->d_dist_scratch; p.qscratch = g->d_qscratch;
    }
  }
  if (int rc = launch_ell_kernel(p, grid, threads, st)) return rc;
  // fallback pass over the trees whose queues overflowed (normally none: its CTAs exit at once)
  if (bitmap_fits) {
    TreeParams pf = p;
    pf.tree_counter = g->d_fail + 1;
    pf.tree_list = g->d_fail + 2;
    pf.n_list = g->d_fail;
    pf.stats = nullptr;
    if (int rc = launch_bitmap_kernel(g, pf, bm_grid, smem, bm_threads, st)) return rc;
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
  g->ew[it->second] = w;
  g->w_dirty = true;
}

int sp_remove_edge(Graph* g, int v1, int v2) {
  if (!g) return SQ_ERR_ARG;
  if (!g->directed && v1 > v2) std::swap(v1, v2);
  auto it = g->key2edge.find(std::make_pair(v1, v2));
  if (it == g->key2edge.end()) return sq::set_error(SQ_ERR_ARG, "remove_edge: no edge (%d,%d)", v1, v2);
  g->alive[it->second] = 0;
  g->key2edge.erase(it);
  g->topo_dirty = true;
  return SQ_OK;
}

int sp_set_weights(Graph* g, const double* w, int nedges) {
  if (!g || !w) return sq::set_error(SQ_ERR_ARG, "sp_set_weights: null argument");
  if (nedges != (int)g->ew.size())
    return sq::set_error(SQ_ERR_ARG, "sp_set_weights: %d weights for %d edges", nedges, (int)g->ew.size());
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

void sp_free_graph(Graph* g) {
  if (!g) return;
  g_free_device(g);
  cudaFree(g->d_counter); cudaFree(g->d_stats); cudaFree(g->d_dist_scratch); cudaFree(g->d_qscratch);
  cudaFree(g->d_origins); cudaFree(g->d_dist_buf); cudaFree(g->d_parent_buf);
  cudaFree(g->d_q2scratch); cudaFree(g->d_fail);
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
  if (algo != 0 && algo != 1) return sq::set_error(SQ_ERR_ARG, "sp_set_algo: algo must be 0 (ELL) or 1 (bitmap)");
  g->algo = algo;
  if (algo == 1 && g->threads < 128) g->threads = 256;
  if (algo == 0 && g->threads > 256) g->threads = 128;
  if (queue_factor > 0.0) g->queue_factor = queue_factor;
  return SQ_OK;
}

int sp_last_stats(Graph* g, long long* out5) {
  if (!g || !out5) return SQ_ERR_ARG;
  for (int i = 0; i < 5; ++i) out5[i] = g->last_stats[i];
  return SQ_OK;
}

const char* sp_last_error(void) { return sq::err_buf(); }
long long sp_kernel_launches(void) { return sq::launch_counter().load(); }

int sp_sssp_batch_device(Graph* g, const int* origins_dev, int n, double* dist_dev, int* parent_dev, void* stream) {
  if (!g || (!origins_dev && n > 0)) return sq::set_error(SQ_ERR_ARG, "sp_sssp_batch_device: null argument");
  return g_run_batch_device(g, origins_dev, n, reinterpret_cast<u64*>(dist_dev), parent_dev, (cudaStream_t)stream);
}

int sp_sssp_batch(Graph* g, const int* origins, int n, double* dist_out, int* parent_out) {
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
  // chunk so that the staging buffers stay below ~2 GiB
  const size_t per_tree = V * (sizeof(u64) + sizeof(int));
  size_t chunk = std::max<size_t>(1, std::min<size_t>((size_t)n, ((size_t)2 << 30) / std::max<size_t>(per_tree, 1)));
  if (dist_out && g->dist_buf_elems < chunk * V) {
    cudaFree(g->d_dist_buf); g->d_dist_buf = nullptr;
    SQ_CUDA_TRY(cudaMalloc(&g->d_dist_buf, sizeof(u64) * chunk * V));
    g->dist_buf_elems = chunk * V;
  }
  if (parent_out && g->parent_buf_elems < chunk * V) {
    cudaFree(g->d_parent_buf); g->d_parent_buf = nullptr;
    SQ_CUDA_TRY(cudaMalloc(&g->d_parent_buf, sizeof(int) * chunk * V));
    g->parent_buf_elems = chunk * V;
  }
  for (size_t off = 0; off < (size_t)n; off += chunk) {
    const int m = (int)std::min(chunk, (size_t)n - off);
    if (int rc = g_run_batch_device(g, g->d_origins + off, m, dist_out ? g->d_dist_buf : nullptr,
                                    parent_out ? g->d_parent_buf : nullptr, 0))
      return rc;
    if (dist_out)
      SQ_CUDA_TRY(cudaMemcpy(dist_out + off * V, g->d_dist_buf, sizeof(u64) * (size_t)m * V, cudaMemcpyDeviceToHost));
    if (parent_out)
      SQ_CUDA_TRY(cudaMemcpy(parent_out + off * V, g->d_parent_buf, sizeof(int) * (size_t)m * V, cudaMemcpyDeviceToHost));
    SQ_CUDA_TRY(cudaDeviceSynchronize());
  }
  g->last_stats[0] = n;
  return g_fetch_stats(g);
}

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
