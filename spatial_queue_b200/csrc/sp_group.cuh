// sp_group.cuh -- grouped tree kernel: EIGHT trees of neighbouring origins per warp, one shared frontier.
//
// Why: a batch of reroute trees has thousands of origins on one road network, so most
// origins have seven others within a few hops.  Their wavefronts sweep the graph almost together.  sssp_warp_kernel
// gives every tree its own queue and its own 8-byte distance words, so each relaxation is a 64-bit L2 atomic on a
// sector of its own (563 k L2 sectors and 315 k L1TEX wavefronts per tree, profiles/r1g_sssp_warp_full_ncu.txt).
// Here the distances of the eight trees of a group are interleaved, dist[v][8] = one 64-byte row per vertex, and the
// group shares ONE queue: a pop relaxes the four out-edges of a vertex for all eight trees at once with the warp's
// 32 lanes = 4 edge slots x 8 trees.  A row is touched as two full sectors, one ELL record serves eight trees, and
// -- because a warp owns its rows outright and pops run in order -- there are no atomics on distances at all:
// load, compare, store.  Per tree that is ~17 k pops and ~255 k sectors instead of 69 k pops and 563 k
// (tools/sim_group2.c models this schedule on the config-5 graph).
//
// Schedule (label-correcting near-far, same fixed point as graph.cc:153-218, hence bit-identical distances):
//  * the sign bit of dist[v][k] marks "improved since v was last popped" (distances are >= 0, so the bit is free).
//    A vertex with a marked lane below the threshold T sits in the FIFO ring exactly once; one whose marked lanes
//    are all >= T sits in the pending list exactly once (a bitmap in global memory makes that exact).  So ring and
//    pending list hold at most V entries each and nothing can overflow: this kernel has no failure path.
//  * a pop clears the marks of its vertex, relaxes all eight lanes, and records four ballots; whether a head goes
//    to the ring or to the pending list is decided for a whole batch of pops at once, one pop per lane.
//  * when the ring runs dry the pending list is scanned: vertices whose smallest marked distance is below the new
//    threshold move to the ring, the rest stay; if none qualifies the threshold jumps to the smallest (closed links
//    at 1e8, queue_class.py:469).
//  * rows, ELL records and the rows of the heads of a batch are prefetched into L1 when the batch is read, so the
//    in-order pops mostly hit L1 (the data is private to the warp; L1 is coherent within an SM).
// Vertices are renumbered along a Hilbert curve over landmark pseudo-coordinates (host side), which puts the two
// rows of a 128-byte line, and the eight origins of a group, next to each other in the network.
#pragma once

#define SQ_SIGN64 0x8000000000000000ull
#define SQ_MASK64 0x7FFFFFFFFFFFFFFFull

struct GroupParams {
  int V;
  int n_groups;
  const EdgeRec* gell;   // [V][4], internal numbering.  .pad bit 0 (all four slots) = record needs the sequential path
                         // (more than four edges, two edges to one head, or a self loop); bit 1 = empty slot (its
                         // head is the vertex itself with weight 0, so the fast path needs no "has edge" test);
                         // bits 3..5 = slot of the edge straight back (warp kernel); slot 0 bits 6.. = overflow
                         // start; slot 1 bits 6.. = overflow count
  const EdgeRec* govf;
  const int4* gheads;    // [V] the four heads of a record (-1 = none), for prefetching
  const int* perm;       // caller id -> internal id
  const int* origins;    // caller ids, one per tree
  const int* tord;       // [n_groups * 8] tree index at each sorted position, -1 = padding
  u64* gdist;            // [n_groups][V][8]
  unsigned* ring;        // [warps][ring_cap]
  unsigned* pend;        // [warps][2][pend_cap]
  unsigned* inp;         // [warps][words] "listed in the pending list"
  int ring_cap;          // power of two >= V
  int pend_cap;          // >= V
  int words;
  int* counter;
  long long* stats;      // [1] pops, [2] edge relaxations (per group, not per tree), [3] transitions
  double delta;
  int* fail_list;        // union kernel: trees of groups whose pending list overflowed (rerun by the bitmap kernel)
  int* fail_count;
};

__device__ __forceinline__ u64 ld_row(const u64* p) {
  u64 v;
  asm volatile("ld.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_row(u64* p, u64 v) { asm volatile("st.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory"); }
__device__ __forceinline__ unsigned ld_u32(const unsigned* p) {
  unsigned v;
  asm volatile("ld.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_u32(unsigned* p, unsigned v) { asm volatile("st.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }
__device__ __forceinline__ void prefetch_l1(const void* p) { asm volatile("prefetch.global.L1 [%0];" ::"l"(p)); }

// PB = pops per batch (8, 16 or 32); PF = 1: prefetch a batch's rows into L1 when it is read
template <int PB, int PF>
__global__ void __launch_bounds__(128) sssp_group_kernel(const GroupParams p) {
  const unsigned FULL = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  const int j = lane >> 3, k = lane & 7;
  const unsigned lt = sq::lanemask_lt();
  const int V = p.V;
  const int gwarp = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const unsigned rmask = (unsigned)p.ring_cap - 1u;
  unsigned* ring = p.ring + (size_t)gwarp * p.ring_cap;
  unsigned* pend0 = p.pend + (size_t)gwarp * 2 * p.pend_cap;
  unsigned* inp = p.inp + (size_t)gwarp * p.words;
  const double delta = p.delta;
  const u64 INF = DIST_INF_BITS;

  int st_pops = 0, st_slow = 0, st_trans = 0;

  for (;;) {
    int grp = 0;
    if (lane == 0) grp = atomicAdd(p.counter, 1);
    grp = __shfl_sync(FULL, grp, 0);
    if (grp >= p.n_groups) break;
    u64* D = p.gdist + (size_t)grp * V * 8;

    int src = -1;
    {
      const int t = __ldg(&p.tord[grp * 8 + k]);
      if (t >= 0) {
        const int o = __ldg(&p.origins[t]);
        if (o >= 0 && o < V) src = __ldg(&p.perm[o]);
      }
    }
    {  // rows = DBL_MAX, unmarked (graph.cc:178); the pending bitmap = 0
      ulonglong2* d2 = reinterpret_cast<ulonglong2*>(D);
      const ulonglong2 inf2 = make_ulonglong2(INF, INF);
      const int n2 = V * 4;
#pragma unroll 4
      for (int i = lane; i < n2; i += 32) d2[i] = inf2;
      for (int i = lane; i < p.words; i += 32) inp[i] = 0u;
    }
    __syncwarp();
    int rh = 0, rt = 0, np = 0, pcur = 0;
    {  // seeds: distance 0, marked; each distinct origin enters the ring once
      const bool mine = (j == 0 && src >= 0);
      if (mine) st_row(D + (size_t)src * 8 + k, SQ_SIGN64);
      bool first = mine;
#pragma unroll
      for (int q = 0; q < 8; ++q) {
        const int other = __shfl_sync(FULL, src, q);
        if (q < k && other == src) first = false;
      }
      const unsigned m = __ballot_sync(FULL, first);
      if (first) st_u32(&ring[__popc(m & lt)], (unsigned)src);
      rt = __popc(m);
    }
    __syncwarp();
    double T = delta;
    // per-lane base addresses: row of vertex x for this lane = Dk + 64 x, record slot = gk + 64 x
    const char* Dk = reinterpret_cast<const char*>(D + k);
    const char* gk = reinterpret_cast<const char*>(p.gell + j);
    asm volatile("" : "+l"(Dk), "+l"(gk));

    for (;;) {
      if (rh == rt) {
        // ---- bucket transition: scan the pending list, four entries per step (eight lanes read one row) ----------
        if (np == 0) break;
        ++st_trans;
        double T_new = advance(T, delta);
        for (;;) {
          const unsigned* P = pend0 + (size_t)pcur * p.pend_cap;
          unsigned* Pn = pend0 + (size_t)(pcur ^ 1) * p.pend_cap;
          int nn = 0;
          rh = 0; rt = 0;
          u64 gmin = INF;
          for (int i0 = 0; i0 < np; i0 += 4) {
            const int idx = i0 + j;
            const bool valid = idx < np;
            const unsigned v = valid ? ld_u32(&P[idx]) : 0u;
            const u64 raw = valid ? ld_row(D + (size_t)v * 8 + k) : INF;
            u64 m = ((long long)raw < 0) ? (raw & SQ_MASK64) : INF;
#pragma unroll
            for (int o = 1; o < 8; o <<= 1) {
              const u64 other = __shfl_xor_sync(FULL, m, o);
              m = other < m ? other : m;
            }
            const bool dirty = m != INF;
            const bool to_ring = valid && dirty && __longlong_as_double((long long)m) < T_new;
            const bool keep = valid && dirty && !to_ring;
            const bool leader = k == 0;
            const unsigned mr = __ballot_sync(FULL, to_ring && leader), mk = __ballot_sync(FULL, keep && leader);
            if (leader && valid) {
              if (to_ring) st_u32(&ring[(unsigned)(rt + __popc(mr & lt)) & rmask], v);
              else if (keep) st_u32(&Pn[nn + __popc(mk & lt)], v);
              if (!keep) atomicAnd(&inp[v >> 5], ~(1u << (v & 31)));
            }
            rt += __popc(mr);
            nn += __popc(mk);
            if (keep && m < gmin) gmin = m;
          }
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) {
            const u64 other = __shfl_xor_sync(FULL, gmin, o);
            gmin = other < gmin ? other : gmin;
          }
          __syncwarp();
          pcur ^= 1;
          np = nn;
          if (rt > 0 || nn == 0) break;
          // nothing within reach: open the next bucket at the smallest marked distance
          const double Tg = __longlong_as_double((long long)gmin);
          T_new = advance(Tg, delta);
        }
        T = T_new;
        continue;
      }

      // ---- one batch of pops ---------------------------------------------------------------------------------
      const int nb = min(PB, rt - rh);
      unsigned ent = 0u;
      int4 hd = make_int4(-1, -1, -1, -1);
      if (lane < nb) {
        ent = ld_u32(&ring[(unsigned)(rh + lane) & rmask]);
        if (PF) {
          prefetch_l1(D + (size_t)ent * 8);
          prefetch_l1(p.gell + (size_t)ent * 4);
        }
        hd = __ldg(&p.gheads[ent]);
        if (PF) {
          if (hd.x >= 0) prefetch_l1(D + (size_t)hd.x * 8);
          if (hd.y >= 0) prefetch_l1(D + (size_t)hd.y * 8);
          if (hd.z >= 0) prefetch_l1(D + (size_t)hd.z * 8);
          if (hd.w >= 0) prefetch_l1(D + (size_t)hd.w * 8);
        }
      }
      rh += nb;
      st_pops += nb;
      unsigned m1 = 0u, m2 = 0u, m3 = 0u, m4 = 0u;   // lane b keeps the ballots of pop b

      for (int b = 0; b < nb; ++b) {
        const unsigned u = __shfl_sync(FULL, ent, b);
        u64* ru = reinterpret_cast<u64*>(const_cast<char*>(Dk) + ((size_t)u << 6));
        const u64 du_raw = ld_row(ru);
        const int4 rc = __ldg(reinterpret_cast<const int4*>(gk + ((size_t)u << 6)));
        const int du_hi = (int)(du_raw >> 32);
        const u64 duc = du_raw & SQ_MASK64;
        if (j == 0 && du_hi < 0) st_row(ru, duc);   // the pop clears the marks of u
        const double dud = __longlong_as_double((long long)duc);
        if (rc.w & 1) {
          // ---- sequential path: parallel edges, self loops, more than four out-edges.  Lanes 0..7 relax one edge at
          // a time and push at once (the batch decode sees no ballots for this pop).
          __syncwarp();
          const int ovf_b = __shfl_sync(FULL, rc.w, 0) >> 6;
          const int ovf_n = __shfl_sync(FULL, rc.w, 8) >> 6;
          for (int e = 0; e < 4 + ovf_n; ++e) {
            int v, pd = 0;
            double w;
            if (e < 4) {
              v = __shfl_sync(FULL, rc.z, e * 8);
              pd = __shfl_sync(FULL, rc.w, e * 8);
              const int wlo = __shfl_sync(FULL, rc.x, e * 8), whi = __shfl_sync(FULL, rc.y, e * 8);
              w = __hiloint2double(whi, wlo);
            } else {
              const EdgeRec er = ld_edge(&p.govf[ovf_b + (e - 4)]);
              v = er.v; w = er.w;
            }
            if ((pd & 2) || v == (int)u) continue;   // warp-uniform: empty slot or self loop
            ++st_slow;
            bool imp = false, near = false, wasd = false, wasn = false;
            if (j == 0) {
              u64* rv = reinterpret_cast<u64*>(const_cast<char*>(Dk) + ((size_t)(unsigned)v << 6));
              const u64 old_raw = ld_row(rv);
              const double oldd = __longlong_as_double((long long)(old_raw & SQ_MASK64));
              const double nd = dud + w;
              imp = nd < oldd;
              near = imp && nd < T;
              wasd = (int)(old_raw >> 32) < 0;
              wasn = wasd && oldd < T;
              if (imp) st_row(rv, (u64)__double_as_longlong(nd) | SQ_SIGN64);
            }
            const bool a_near = __any_sync(FULL, near), a_imp = __any_sync(FULL, imp);
            const bool a_wasn = __any_sync(FULL, wasn), a_wasd = __any_sync(FULL, wasd);
            if (a_near && !a_wasn) {
              if (lane == 0) st_u32(&ring[(unsigned)rt & rmask], (unsigned)v);
              ++rt;
            } else if (a_imp && !a_near && !a_wasd) {
              unsigned old = 0u;
              if (lane == 0) old = atomicOr(&inp[v >> 5], 1u << (v & 31));
              old = __shfl_sync(FULL, old, 0);
              if (!(old & (1u << (v & 31)))) {
                if (lane == 0) st_u32(&pend0[(size_t)pcur * p.pend_cap + np], (unsigned)v);
                ++np;
              }
            }
            __syncwarp();
          }
          continue;
        }
        // fast path: the four heads are distinct and none is u (empty slots point at u with weight 0: never improve)
        u64* rv = reinterpret_cast<u64*>(const_cast<char*>(Dk) + ((size_t)(unsigned)rc.z << 6));
        const u64 old_raw = ld_row(rv);
        const double oldd = __longlong_as_double((long long)(old_raw & SQ_MASK64));
        const double nd = dud + __hiloint2double(rc.y, rc.x);
        const bool imp = nd < oldd;
        const bool near = imp && nd < T;
        const bool wasd = (int)(old_raw >> 32) < 0;
        const bool wasn = wasd && oldd < T;
        const unsigned b1 = __ballot_sync(FULL, near), b2 = __ballot_sync(FULL, imp);
        const unsigned b3 = __ballot_sync(FULL, wasn), b4 = __ballot_sync(FULL, wasd);
        if (imp) st_row(rv, (u64)__double_as_longlong(nd) | SQ_SIGN64);
        if (lane == b) { m1 = b1; m2 = b2; m3 = b3; m4 = b4; }
        __syncwarp();
      }

      // ---- batch decode: lane b turns the ballots of pop b into ring / pending pushes of its four heads ----------
      {
        const int hv[4] = {hd.x, hd.y, hd.z, hd.w};
        bool tr[4], tp[4];
        int cr = 0, cp = 0;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const unsigned n1 = (m1 >> (8 * q)) & 0xffu, n2 = (m2 >> (8 * q)) & 0xffu;
          const unsigned n3 = (m3 >> (8 * q)) & 0xffu, n4 = (m4 >> (8 * q)) & 0xffu;
          tr[q] = n1 != 0u && n3 == 0u;                 // became near-marked: enters the ring
          tp[q] = n2 != 0u && n1 == 0u && n4 == 0u;     // became marked, all beyond T: pending list, if not listed yet
        }
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          if (tp[q]) {
            const unsigned bit = 1u << (hv[q] & 31);
            tp[q] = (atomicOr(&inp[hv[q] >> 5], bit) & bit) == 0u;
          }
          cr += tr[q];
          cp += tp[q];
        }
        const int packed = cr | (cp << 16);
        int incl = packed;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
          const int other = __shfl_up_sync(FULL, incl, o);
          if (lane >= o) incl += other;
        }
        const int total = __shfl_sync(FULL, incl, 31);
        int r0 = rt + ((incl - packed) & 0xffff), p0 = np + ((incl - packed) >> 16);
        unsigned* Pc = pend0 + (size_t)pcur * p.pend_cap;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          if (tr[q]) st_u32(&ring[(unsigned)(r0++) & rmask], (unsigned)hv[q]);
          if (tp[q]) st_u32(&Pc[p0++], (unsigned)hv[q]);
        }
        rt += total & 0xffff;
        np += total >> 16;
      }
      __syncwarp();
    }
    __syncwarp();
  }

  if (p.stats && lane == 0) {
    atomicAdd((u64*)&p.stats[1], (u64)st_pops);
    atomicAdd((u64*)&p.stats[2], (u64)st_pops * 4ull + (u64)st_slow);   // edge slots visited (each for 8 trees)
    atomicAdd((u64*)&p.stats[3], (u64)st_trans);
  }
}

// Un-interleaves a chunk of finished groups into the caller's row-major tables and computes the canonical
// predecessors, parent[v] = min{u : fl(d[u] + w(u,v)) == d[v]} (caller ids), in the same pass.  A block owns a TILE of
// 256 consecutive internal ids = a compact patch of the network (Hilbert order): per group it stages the tile's 256
// rows (16 KB, one coalesced read) in shared memory, so a vertex's own row and those of its in-neighbours inside the
// tile -- three quarters of them -- cost no L2 traffic; neighbours outside the tile are read from L2.  Threads are
// assigned to the tile's vertices in ascending CALLER id (tile_ord), so the eight rows a warp writes per group go out
// as runs of consecutive addresses.  gin = the group ELL of the opposite direction (in-edges).
constexpr int FIN_TILE = 256;
__global__ void __launch_bounds__(FIN_TILE) group_finalize_kernel(int V, int n_groups, const EdgeRec* __restrict__ gin,
                                                                  const EdgeRec* __restrict__ ginovf, const int* __restrict__ iperm,
                                                                  const unsigned char* __restrict__ tile_ord,
                                                                  const int* __restrict__ origins, const int* __restrict__ tord,
                                                                  const u64* __restrict__ gdist, u64* __restrict__ dist_out,
                                                                  int* __restrict__ parent_out, int codes) {
  __shared__ u64 s_rows[FIN_TILE][9];   // 72-byte pitch: the column reads of a warp spread over the banks
  const int base = blockIdx.x * FIN_TILE;
  const int nrows = min(FIN_TILE, V - base);
  const int tid = threadIdx.x;
  const bool live = tid < nrows;
  const int xl = live ? (int)tile_ord[base + tid] : 0;   // my vertex, as an offset into the tile
  const int x = base + xl;
  const int v = live ? __ldg(&iperm[x]) : -1;             // caller id
  int nx[4], nv[4];
  double nw[4];
  int ovf_b = 0, ovf_n = 0;
#pragma unroll
  for (int q = 0; q < 4; ++q) { nx[q] = -1; nv[q] = -1; nw[q] = 0.0; }
  if (parent_out && live) {
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int4 raw = __ldg(reinterpret_cast<const int4*>(gin + (size_t)x * 4 + q));
      nw[q] = __hiloint2double(raw.y, raw.x);
      nx[q] = (raw.w & 2) ? -1 : raw.z;            // empty slots point at the vertex itself
      nv[q] = nx[q] >= 0 ? __ldg(&iperm[nx[q]]) : -1;
      if (q == 0) ovf_b = raw.w >> 6;
      if (q == 1) ovf_n = raw.w >> 6;
    }
  }
  for (int g = blockIdx.y; g < n_groups; g += gridDim.y) {
    const u64* G = gdist + (size_t)g * V * 8;
    {  // stage the tile: nrows x 64 bytes, contiguous in the group's table
      const ulonglong2* src = reinterpret_cast<const ulonglong2*>(G + (size_t)base * 8);
      for (int e = tid; e < nrows * 4; e += FIN_TILE) {
        const ulonglong2 t2 = __ldg(&src[e]);
        s_rows[e >> 2][(e & 3) * 2] = t2.x & SQ_MASK64;
        s_rows[e >> 2][(e & 3) * 2 + 1] = t2.y & SQ_MASK64;
      }
    }
    __syncthreads();
    if (live) {
      u64 dv[8];
#pragma unroll
      for (int kk = 0; kk < 8; ++kk) dv[kk] = s_rows[xl][kk];
      int best[8], slot[8];
#pragma unroll
      for (int kk = 0; kk < 8; ++kk) { best[kk] = INT_MAX; slot[kk] = 255; }
      if (parent_out) {
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          if (nx[q] < 0) continue;
          u64 du[8];
          const int off = nx[q] - base;
          if (off >= 0 && off < nrows) {
#pragma unroll
            for (int kk = 0; kk < 8; ++kk) du[kk] = s_rows[off][kk];
          } else {
            const ulonglong2* r2 = reinterpret_cast<const ulonglong2*>(G + (size_t)nx[q] * 8);
#pragma unroll
            for (int h = 0; h < 4; ++h) {
              const ulonglong2 t2 = __ldg(&r2[h]);
              du[2 * h] = t2.x & SQ_MASK64;
              du[2 * h + 1] = t2.y & SQ_MASK64;
            }
          }
#pragma unroll
          for (int kk = 0; kk < 8; ++kk) {
            if (du[kk] != DIST_INF_BITS && dv[kk] != DIST_INF_BITS) {
              const double cand = __longlong_as_double((long long)du[kk]) + nw[q];
              if ((u64)__double_as_longlong(cand) == dv[kk] && nv[q] < best[kk]) { best[kk] = nv[q]; slot[kk] = q; }
            }
          }
        }
        for (int e = 0; e < ovf_n; ++e) {
          const EdgeRec er = ld_edge(&ginovf[ovf_b + e]);
          const int uv = __ldg(&iperm[er.v]);
          const u64* r = G + (size_t)er.v * 8;
#pragma unroll
          for (int kk = 0; kk < 8; ++kk) {
            const u64 du = __ldg(&r[kk]) & SQ_MASK64;
            if (du != DIST_INF_BITS && dv[kk] != DIST_INF_BITS) {
              const double cand = __longlong_as_double((long long)du) + er.w;
              if ((u64)__double_as_longlong(cand) == dv[kk] && uv < best[kk]) { best[kk] = uv; slot[kk] = min(4 + e, 254); }
            }
          }
        }
      }
#pragma unroll
      for (int kk = 0; kk < 8; ++kk) {
        const int t = __ldg(&tord[g * 8 + kk]);
        if (t < 0) continue;
        if (dist_out) dist_out[(size_t)t * V + v] = dv[kk];
        if (parent_out) {
          const bool none = best[kk] == INT_MAX || v == __ldg(&origins[t]);
          if (codes) reinterpret_cast<unsigned char*>(parent_out)[(size_t)t * V + v] = none ? (unsigned char)255 : (unsigned char)slot[kk];
          else parent_out[(size_t)t * V + v] = none ? -1 : best[kk];
        }
      }
    }
    __syncthreads();
  }
}

// ---- origins -> groups: counting sort of the trees by the internal id of their origin ---------------------------------
__global__ void __launch_bounds__(256) group_hist_kernel(int n, int V, const int* __restrict__ origins,
                                                         const int* __restrict__ perm, int* __restrict__ key, int* __restrict__ hist) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i >= n) return;
  const int o = origins[i];
  const int kx = (o >= 0 && o < V) ? perm[o] : V;   // invalid origins sort last
  key[i] = kx;
  atomicAdd(&hist[kx], 1);
}
// exclusive prefix sum of hist[0..m) in place, one block
__global__ void __launch_bounds__(1024) group_scan_kernel(int m, int* __restrict__ hist) {
  __shared__ int part[1024];
  const int tid = threadIdx.x;
  const int per = (m + 1023) / 1024;
  const int b = tid * per, e = min(m, b + per);
  int s = 0;
  for (int i = b; i < e; ++i) s += hist[i];
  part[tid] = s;
  __syncthreads();
  for (int o = 1; o < 1024; o <<= 1) {
    const int t = tid >= o ? part[tid - o] : 0;
    __syncthreads();
    part[tid] += t;
    __syncthreads();
  }
  int run = part[tid] - s;
  for (int i = b; i < e; ++i) { const int c = hist[i]; hist[i] = run; run += c; }
}
__global__ void __launch_bounds__(256) group_scatter_kernel(int n, int n_pad, const int* __restrict__ key, int* __restrict__ cursor,
                                                            int* __restrict__ tord) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i < n) tord[atomicAdd(&cursor[key[i]], 1)] = i;
  else if (i < n_pad) tord[i] = -1;
}

// -------------------------------------------------------------------------------------------------------------------
// Union kernel: the group layout and the shared frontier of sssp_group_kernel, but relaxations are 64-bit L2 atomicMin
// (eight lanes of a slot hit one 64-byte row: two sectors per edge for eight trees) and P pops are in flight at once.
// sssp_group_kernel showed what in-order pops cost: two dependent DRAM-latency round trips per pop, ~2,600 cycles, with
// nothing to overlap them.  Atomics need no ordering between pops, so a step issues the row and record loads of P pops
// together, then their 4 x P edge atomics, then does the queue bookkeeping -- the structure of sssp_warp_kernel, with
// the requests of eight trees folded into one.
//  * no marks in the distance words (they are atomicMin operands); a bit per vertex in shared memory says "in the ring"
//    (set by the push that wins atomicOr, cleared when the batch holding the vertex is read);
//  * improvements beyond the threshold go to the pending list unless the vertex is in the ring anyway; duplicates are
//    possible there, the scan at the bucket transition removes them with the same bit;
//  * the scan reads a row and takes the smallest distance >= T: below T + delta the vertex enters the ring.
// The bitmap bounds V by shared memory (4 warps x V/8 bytes per CTA); larger graphs use the other group kernels.
// -------------------------------------------------------------------------------------------------------------------
template <int P>
__global__ void __launch_bounds__(128) sssp_union_kernel(const GroupParams p) {
  extern __shared__ unsigned s_union_bits[];
  const unsigned FULL = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  const int j = lane >> 3, k = lane & 7;
  const unsigned lt = sq::lanemask_lt();
  const int V = p.V;
  const int gwarp = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const unsigned rmask = (unsigned)p.ring_cap - 1u;
  unsigned* ring = p.ring + (size_t)gwarp * p.ring_cap;
  unsigned* pend0 = p.pend + (size_t)gwarp * 2 * p.pend_cap;
  unsigned* bits = s_union_bits + (size_t)(threadIdx.x >> 5) * 2 * p.words;   // "in the ring"
  unsigned* pbits = bits + p.words;                                            // "listed in the pending list"
  const double delta = p.delta;
  const u64 INF = DIST_INF_BITS;
  const bool leader = k == 0;
  const int sh = 8 * j;

  int st_pops = 0, st_trans = 0;

  for (;;) {
    int grp = 0;
    if (lane == 0) grp = atomicAdd(p.counter, 1);
    grp = __shfl_sync(FULL, grp, 0);
    if (grp >= p.n_groups) break;
    u64* D = p.gdist + (size_t)grp * V * 8;
    int src = -1;
    {
      const int t = __ldg(&p.tord[grp * 8 + k]);
      if (t >= 0) {
        const int o = __ldg(&p.origins[t]);
        if (o >= 0 && o < V) src = __ldg(&p.perm[o]);
      }
    }
    {
      ulonglong2* d2 = reinterpret_cast<ulonglong2*>(D);
      const ulonglong2 inf2 = make_ulonglong2(INF, INF);
      const int n2 = V * 4;
#pragma unroll 4
      for (int i = lane; i < n2; i += 32) d2[i] = inf2;
      for (int i = lane; i < 2 * p.words; i += 32) bits[i] = 0u;
    }
    __syncwarp();
    int rh = 0, rt = 0, np = 0, pcur = 0;
    bool fail = false;
    {
      const bool mine = (j == 0 && src >= 0);
      if (mine) st_row(D + (size_t)src * 8 + k, 0ull);
      bool first = mine;
#pragma unroll
      for (int q = 0; q < 8; ++q) {
        const int other = __shfl_sync(FULL, src, q);
        if (q < k && other == src) first = false;
      }
      const unsigned m = __ballot_sync(FULL, first);
      if (first) {
        st_u32(&ring[__popc(m & lt)], (unsigned)src);
        atomicOr(&bits[src >> 5], 1u << (src & 31));
      }
      rt = __popc(m);
    }
    __syncwarp();
    double T = delta;
    u64* Dk = D + k;

    while (!fail) {
      if (rh == rt) {
        // ---- bucket transition: four pending entries per step, eight lanes read one row ------------------------------
        if (np == 0) break;
        ++st_trans;
        double T_lo = T;                       // everything below is settled and was relaxed onward
        double T_new = advance(T, delta);
        for (;;) {
          const unsigned* Pc = pend0 + (size_t)pcur * p.pend_cap;
          unsigned* Pn = pend0 + (size_t)(pcur ^ 1) * p.pend_cap;
          int nn = 0;
          rh = 0; rt = 0;
          u64 gmin = INF;
          for (int i0 = 0; i0 < np; i0 += 4) {
            const int idx = i0 + j;
            const bool valid = idx < np;
            const unsigned v = valid ? ld_u32(&Pc[idx]) : 0u;
            const u64 raw = valid ? sq::ld_cg_u64(Dk + ((size_t)v << 3)) : INF;
            u64 m = (__longlong_as_double((long long)raw) >= T_lo) ? raw : INF;
#pragma unroll
            for (int o = 1; o < 8; o <<= 1) {
              const u64 other = __shfl_xor_sync(FULL, m, o);
              m = other < m ? other : m;
            }
            const bool live = valid && m != INF;
            bool to_ring = live && __longlong_as_double((long long)m) < T_new;
            const bool keep = live && !to_ring;
            if (leader && valid) {
              const unsigned bit = 1u << (v & 31);
              if (to_ring) to_ring = (atomicOr(&bits[v >> 5], bit) & bit) == 0u;   // may be in the ring already
              if (!keep) atomicAnd(&pbits[v >> 5], ~bit);                           // leaves the list
            }
            const unsigned mr = __ballot_sync(FULL, to_ring && leader), mk = __ballot_sync(FULL, keep && leader);
            if (leader) {
              if (to_ring) st_u32(&ring[(unsigned)(rt + __popc(mr & lt)) & rmask], v);
              else if (keep) st_u32(&Pn[nn + __popc(mk & lt)], v);
            }
            rt += __popc(mr);
            nn += __popc(mk);
            if (keep && m < gmin) gmin = m;
          }
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) {
            const u64 other = __shfl_xor_sync(FULL, gmin, o);
            gmin = other < gmin ? other : gmin;
          }
          __syncwarp();
          pcur ^= 1;
          np = nn;
          if (rt > 0 || nn == 0) break;
          const double Tg = __longlong_as_double((long long)gmin);   // nothing within reach: jump
          T_lo = Tg;
          T_new = advance(Tg, delta);
        }
        T = T_new;
        continue;
      }

      // ---- one batch of up to 32 pops, P at a time ----------------------------------------------------------------------
      const int nb = min(32, rt - rh);
      unsigned ent = 0u;
      if (lane < nb) ent = ld_u32(&ring[(unsigned)(rh + lane) & rmask]);
      rh += nb;
      st_pops += nb;
      unsigned* Pc = pend0 + (size_t)pcur * p.pend_cap;
      for (int b0 = 0; b0 < nb; b0 += P) {
        {  // the P vertices of this step leave the ring now: an improvement from here on re-queues them
          const unsigned e = __shfl_sync(FULL, ent, (b0 + lane) & 31);
          if (lane < P && b0 + lane < nb) atomicAnd(&bits[e >> 5], ~(1u << (e & 31)));
        }
        u64 du[P];
        int4 rc[P];
        bool act[P];
#pragma unroll
        for (int q = 0; q < P; ++q) {
          act[q] = b0 + q < nb;
          const unsigned u = __shfl_sync(FULL, ent, (b0 + q) & 31);
          du[q] = INF;
          rc[q] = make_int4(0, 0, 0, 2);
          if (act[q]) {
            du[q] = sq::ld_cg_u64(Dk + ((size_t)u << 3));
            rc[q] = __ldg(reinterpret_cast<const int4*>(p.gell + (size_t)u * 4 + j));
          }
        }
        u64 ndb[P], old[P];
        bool has[P];
#pragma unroll
        for (int q = 0; q < P; ++q) {
          has[q] = act[q] && !(rc[q].w & 2) && du[q] != INF;
          ndb[q] = (u64)__double_as_longlong(__longlong_as_double((long long)du[q]) + __hiloint2double(rc[q].y, rc[q].x));
          old[q] = 0ull;
        }
#pragma unroll
        for (int q = 0; q < P; ++q)
          if (has[q]) old[q] = atomicMin(Dk + ((size_t)(unsigned)rc[q].z << 3), ndb[q]);
#pragma unroll
        for (int q = 0; q < P; ++q) {
          const bool imp = has[q] && ndb[q] < old[q];
          const bool near = imp && __longlong_as_double((long long)ndb[q]) < T;
          const unsigned b1 = __ballot_sync(FULL, near), b2 = __ballot_sync(FULL, imp);
          if (b2 == 0u) continue;                                   // warp-uniform
          const bool s_near = ((b1 >> sh) & 0xffu) != 0u, s_imp = ((b2 >> sh) & 0xffu) != 0u;
          bool pr = false, pp = false;
          const unsigned v = (unsigned)rc[q].z;
          if (leader && s_imp) {
            const unsigned bit = 1u << (v & 31);
            if (s_near) pr = (atomicOr(&bits[v >> 5], bit) & bit) == 0u;
            else if (!(bits[v >> 5] & bit)) pp = (atomicOr(&pbits[v >> 5], bit) & bit) == 0u;
          }
          const unsigned mr = __ballot_sync(FULL, pr), mp = __ballot_sync(FULL, pp);
          if (pr) st_u32(&ring[(unsigned)(rt + __popc(mr & lt)) & rmask], v);
          if (pp) {
            const int pos = np + __popc(mp & lt);
            if (pos < p.pend_cap) st_u32(&Pc[pos], v);
          }
          rt += __popc(mr);
          np += __popc(mp);
        }
        // edges beyond the fourth of a vertex (rare on road networks): one edge at a time, lanes 0..7
#pragma unroll
        for (int q = 0; q < P; ++q) {
          const int w0 = __shfl_sync(FULL, rc[q].w, 0), w1 = __shfl_sync(FULL, rc[q].w, 8);
          const int ovf_n = act[q] ? (w1 >> 6) : 0;
          if (ovf_n == 0) continue;                                 // warp-uniform
          const int ovf_b = w0 >> 6;
          const u64 du0 = __shfl_sync(FULL, du[q], k);              // lanes 0..7 hold the row of u
          for (int e = 0; e < ovf_n; ++e) {
            const EdgeRec er = ld_edge(&p.govf[ovf_b + e]);
            bool imp = false, near = false;
            if (j == 0 && du0 != INF) {
              const u64 nb2 = (u64)__double_as_longlong(__longlong_as_double((long long)du0) + er.w);
              const u64 o2 = atomicMin(Dk + ((size_t)(unsigned)er.v << 3), nb2);
              imp = nb2 < o2;
              near = imp && __longlong_as_double((long long)nb2) < T;
            }
            const bool a_near = __any_sync(FULL, near), a_imp = __any_sync(FULL, imp);
            if (!a_imp) continue;
            const unsigned bit = 1u << (er.v & 31);
            unsigned was = 0u;
            if (lane == 0) {
              if (a_near) was = atomicOr(&bits[er.v >> 5], bit);
              else was = (bits[er.v >> 5] & bit) ? bit : atomicOr(&pbits[er.v >> 5], bit);
            }
            was = __shfl_sync(FULL, was, 0);
            if (!(was & bit)) {
              if (a_near) { if (lane == 0) st_u32(&ring[(unsigned)rt & rmask], (unsigned)er.v); ++rt; }
              else { if (lane == 0 && np < p.pend_cap) st_u32(&Pc[np], (unsigned)er.v); ++np; }
            }
          }
        }
        if (np > p.pend_cap || rt - rh > p.ring_cap) { fail = true; break; }
      }
      __syncwarp();
    }
    __syncwarp();
    if (fail) {
      // the pending list outgrew its bound (duplicates): hand the group's trees to the fallback kernel
      const int t = __ldg(&p.tord[grp * 8 + k]);
      if (j == 0 && t >= 0) p.fail_list[atomicAdd(p.fail_count, 1)] = t;
    }
  }
  if (p.stats && lane == 0) {
    atomicAdd((u64*)&p.stats[1], (u64)st_pops);
    atomicAdd((u64*)&p.stats[2], (u64)st_pops * 4ull);
    atomicAdd((u64*)&p.stats[3], (u64)st_trans);
  }
}
