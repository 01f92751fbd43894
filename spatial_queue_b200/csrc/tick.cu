// tick.cu -- the one-second spatial-queue tick on B200 (include/sq_tick.h).
//
// What it replaces (reference file:line, projects/bolinas_civic/): the three Python loops a driver runs every
// simulated second -- Agent.load_vehicle (queue_class.py:587-596), Link.run_link_model (:445-455) and
// Node.run_node_model / non_conflict_vehicles (:226-394) driven by run_traffic_simulation.py:52-69, 111.
//
// Design (not a translation of those loops):
//  * Every link owns a ring of {agent, ready_tick} slots in HBM.  queue_vehicles is ring[head, qtail),
//    run_vehicles is ring[qtail, tail): one entry-ordered sequence with a moving boundary, because the
//    reference only ever appends to the run list and promotes a prefix of it.  `ready_tick` is the first tick
//    at which the reference's f64 predicate `enter_time < t - fft` (:447) holds, computed once at entry with
//    that same f64 expression, so the link model is a boundary advance instead of a list rebuild.
//  * Source (virtual) links have their whole departure sequence written at creation in (departure, agent)
//    order -- the order load_vehicle appends in -- so loading is a tail advance.
//  * Storage is an integer occupancy (length entered - length left) subtracted from the f64 capacity once
//    per tick, exactly the reference's single subtraction (:453); the node model then subtracts lengths one
//    by one in f64 like :348.
//  * Link state is split by writer: {head, out_c, total_left, len_out} belong to the link's end node,
//    {tail, in_c, storage_c, len_in} to its start node.  Within a tick a node only touches the in-side of its
//    in-links and the out-side of its out-links, so nodes are independent: one thread per node, no atomics on
//    simulation state, bit-reproducible.  Atomics appear only on run counters (warp-aggregated).
//  * Turn conflicts: a 64-bit left-turn mask per in-link (bit j = angle to the j-th out-link >= 50 deg, :263)
//    and the node's f64 angle table for the opposite-approach argmin (:277), both built from the reference's
//    own add_connectivity output.
//  * Randomness: counter-based contract of spatial_queue_b200/rng.py.
#include <algorithm>
#include <cfloat>
#include <climits>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <numeric>
#include <utility>
#include <vector>

#include "../../include/sq_sp.h"
#include "../../include/sq_tick.h"
#include "sq_common.cuh"

using sq::SQ_ERR_ARG;
using sq::SQ_ERR_CUDA;
using sq::SQ_ERR_LIMIT;
using sq::SQ_ERR_STATE;
using sq::SQ_OK;

typedef unsigned long long u64;

// =================================================================================================
// Device state
// =================================================================================================
struct __align__(32) AgentRec {
  int4 rec;        // x next_link, y cur_link, z enter_time, w status | vlen << 8
  int destin;      // destination node
  int route_len;   // >= 0: explicit route (link sequence in a pool); < 0: follows the router's next-link tables
  int epoch;       // table epoch slot the route follows (route_len < 0)
  int dslot;       // destination slot of the router
};

// Static words of an in-link's window, one 32-byte sector: ring placement, left-turn mask over the end node's
// out-links, whole lanes, and the local out-index (at the end node) of the reverse link (-1 = none).
struct __align__(32) WinRec {
  long long base;
  u64 left;
  unsigned mask;
  int lanes, rcol, pad;
};

// Static words of a node, one 64-byte record read before the dynamic chain starts: its first eight in-links and, per
// in-link slot i, who the opposite approach is when i is the primary link (:267-293) -- the other slots ordered by
// |turn angle onto the reverse of link i|, so that the choice is "first of the list that has a queue" instead of a walk
// over the angle table.
//   sec[i]: bits 0..2 entries, bits 3..23 the slots (3 bits each); NR_NONE: link i has no reverse link (no opposite
//   approach); NR_SLOW: equal angles / duplicate priority ranks, the order of the walk matters -> pick_secondary.
//   sec[0] & NR_NODE_SLOW: more than eight in-links, the record is not used.
struct __align__(64) NodeRec {
  int in[8];
  unsigned sec[8];
};
constexpr unsigned NR_NONE = 1u << 30, NR_SLOW = 1u << 31, NR_NODE_SLOW = 1u << 29;

struct TickState {
  int N, L, A;
  const NodeRec* n_rec;
  u64 seed;
  // links, static
  const int* l_start; const int* l_end; const int* l_lanes; const int* l_rev; const WinRec* l_win;
  const unsigned char* l_virtual; unsigned char* l_closed;
  const double* l_cap_s; const double* l_storage; const double* l_fft;
  const long long* l_ring_base; const unsigned* l_ring_mask; const int* l_vcount;
  const unsigned char* l_in_local; const unsigned char* l_out_local; const u64* l_left_mask;
  const int4* l_hop;         // {start node, end node, local index among the start node's out-links, 0}: one load per next link
  // links, dynamic
  unsigned* head; unsigned* qtail; unsigned* tail;
  double* out_c; double* in_c; double* storage_c;
  long long* len_in; long long* len_out; int* total_left;
  int2* ring;
  // nodes
  const long long* n_key; const unsigned char* n_type;
  const int* in_ptr; const int* in_links; const unsigned char* in_phase; const int* in_prio;
  const int* out_ptr; const int* out_links;
  const long long* turn_ptr; const double* turn;
  long long* node_move;
  // agents
  AgentRec* ag;  // one 32-byte sector per agent: everything the node phase gathers per queued vehicle
  long long* a_route_off; int* a_route_pos; int* a_arrive;
  const int* route_pool;
  // device router (sq_attach_router): per-destination next-link tables, K epochs of them
  int D, K, freq;
  const int* n_vertex;       // [N] graph vertex of node n
  int* next_tbl;             // [K][D][N] link to take at node n towards destination d, -1 at d / unreachable
  double* dist_tbl;          // [K][D][N] distance to d
  int* cur_epoch;            // slot of the newest tables
  const int* bucket_ptr;     // [freq + 1] agents by (key % freq)
  const int* bucket_agents;
  const int* l_edge;         // [L] edge of the routing graph the link's weight goes to, -1 = none
  const double* l_capacity; const double* l_maxmph;
  double* w_edge;            // [graph edges] staging for sp_set_weights_device
  int* mat_pool;             // routes materialised on the device (agents that could not be rerouted)
  long long* mat_used; long long mat_cap;
  // Butte's rules inside the router (model/dta_meso_butte.py:193-209, 256-263), off unless sq_set_reroute_rules
  double* a_rdist;           // [A] Agent.route_distance: distance of the route last adopted (5e7 before the first)
  double detour_factor;      // > 0: an enroute agent whose new route is longer than factor x the old one shelters (shelter_a2)
  int cfreq;                 // > 0: agents whose next link is closed reroute when t % cfreq == key % cfreq
  const int* cbucket_ptr; const int* cbucket_agents;   // agents by key % cfreq
  const int* a_rslot; const int* a_cslot;              // key % freq, key % cfreq per agent
  int max_cbucket;
  int* l_tomb;               // [L] > 0: vehicles on this link stopped during this tick's agent loop; the link phase strips them
  // agent-loop rules of the drivers
  const int* l_dwell;        // floor(fft) + 2: upper bound of ready_tick - enter_time
  int* l_shelter;            // vehicles that gave up on this link (shelter_a1)
  int shelter_after;         // seconds on one link before a vehicle shelters (run_traffic_simulation.py:103), 0 = off
  int congest_after;         // seconds on one link before a vehicle counts as congested (run_traffic_simulation.py:101), 0 = off
  unsigned* pair_moves;      // [turn table shape] vehicles moved from in-link i to out-link j of a node, or nullptr (off):
                             // the (in_link, out_link) pairs run_node_model returns (model/queue_class.py:377-383)
  u64* n_cand;               // [N] bit i = in-link i of the node has a non-empty queue (written by k_link_phase)
  unsigned* cong_pos;        // [L] ring index of the first vehicle not yet congested (lags behind head when the queue moves)
  long long* l_congested;    // [L] link.congested of the last tick: vehicles on the link for longer than congest_after
  // run control
  int* tick_link;  // tick the next link-phase launch simulates
  int* tick_node;  // tick the next node-phase launch simulates
  u64* counters;   // [0] ticks [1] agent_steps [2] moves [3] active nodes [4] arrivals [5] loaded [6] loaded, frozen
                   // [7] sheltered on a link [8] the same, frozen
  int* err;        // [0] code (0 = ok) [1] agent [2] node [3] link
  // measurement aid (sq_enable_timeline), nullptr = off: %globaltimer at entry and exit of every warp of the two
  // phase kernels for ticks [tl_t0, tl_t0 + tl_ticks); entry [(tick - tl_t0) * 2 + phase][warp]
  ulonglong2* timeline; int tl_t0, tl_ticks, tl_stride;
#ifdef SQ_TICK_PHASES
  u64* phases;     // experiment build only: [tl_ticks][tl_stride][16] checkpoints inside k_node_phase
#endif
  int pdl_early;   // 1: let the next kernel's CTAs be placed as soon as this kernel's dependency is met (see pdl_wait)
};

// Programmatic dependent launch (sm_90+): the three per-tick kernels are launched with
// cudaLaunchAttributeProgrammaticStreamSerialization, so the next kernel's CTAs are placed on the SMs while this one
// still runs.  pdl_wait() returns once the previous kernel has completed and its writes are visible -- nothing written
// by an earlier kernel may be read before it; pdl_release() lets the following kernel start its own prologue.  It is
// issued after the wait, so at most one kernel runs ahead.  Launched without the attribute both are no-ops.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_release() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

__device__ __forceinline__ u64 global_ns() { u64 v; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(v)); return v; }
// Slot of this warp in the timeline, or nullptr (off, outside the window, not lane 0).
__device__ __forceinline__ ulonglong2* timeline_slot(const TickState& s, int t, int phase) {
  if (!s.timeline || (threadIdx.x & 31u)) return nullptr;
  const int rel = t - s.tl_t0;
  const int w = (int)((blockIdx.x * blockDim.x + threadIdx.x) >> 5);
  if (rel < 0 || rel >= s.tl_ticks || w >= s.tl_stride) return nullptr;
  return s.timeline + ((size_t)(rel * 2 + phase) * s.tl_stride + w);
}

__device__ __forceinline__ u64 fmix64(u64 x) {
  x ^= x >> 30; x *= 0xBF58476D1CE4E5B9ull; x ^= x >> 27; x *= 0x94D049BB133111EBull; x ^= x >> 31;
  return x;
}
// spatial_queue_b200/rng.py: u(seed, t, node, k)
__device__ __forceinline__ double rng_uniform(u64 seed, int t, long long node, int k) {
  u64 h = seed ^ 0x243F6A8885A308D3ull;
  h = fmix64(h + (u64)(long long)t * 0x9E3779B97F4A7C15ull);
  h = fmix64(h + (u64)node * 0xBF58476D1CE4E5B9ull);
  h = fmix64(h + (u64)(long long)k * 0x94D049BB133111EBull);
  return (double)(h >> 11) * (1.0 / 9007199254740992.0);
}

__device__ __forceinline__ void raise_error(const TickState& s, int code, int agent, int node, int link) {
  if (atomicCAS(&s.err[0], 0, code) == 0) { s.err[1] = agent; s.err[2] = node; s.err[3] = link; }
}

// First tick t' with (double)t_enter < (double)t' - fft, the promotion test of queue_class.py:447.
__device__ __forceinline__ int ready_tick(int t_enter, double fft) {
  if (!(fft < 1.0e9)) return INT_MAX;
  int r = t_enter + (int)floor(fft) + 1;
  const double te = (double)t_enter;
  while (!(te < (double)r - fft)) ++r;
  while (te < (double)(r - 1) - fft) --r;
  return r;
}

// Next link of `agent` once it stands at the end of link `ol` (find_next_link, queue_class.py:598-608).
// Explicit routes are link sequences in a pool (negative offsets: the device-side pool of k_evacuate);
// table routes (a_route_len < 0) follow the next-link table of the epoch the agent was last routed on.
__device__ __forceinline__ int route_link_at(const TickState& s, int agent, int pos) {
  const long long off = s.a_route_off[agent];
  return off >= 0 ? s.route_pool[off + pos] : s.mat_pool[-(off + 1) + pos];
}
__device__ __forceinline__ size_t tbl_index(const TickState& s, int epoch, int dslot, int node) {
  // the row number (table slot x destination) fits 32 bits; one widening multiply-add for the rest
  return (size_t)((u64)(unsigned)(epoch * s.D + dslot) * (u64)(unsigned)s.N + (u64)(unsigned)node);
}

// Vehicles that stopped on link l since the last tick (status no longer enroute) leave its lists at once, wherever they
// stand (model/dta_meso_butte.py:256-263): the survivors slide towards the tail in order, the head and the queue / run
// boundary move up by what was removed on their side.  Returns the number removed.
__device__ int strip_stopped(const TickState& s, int l) {
  const long long base = s.l_ring_base[l];
  const unsigned mask = s.l_ring_mask[l];
  const unsigned h = s.head[l], qt = s.qtail[l], tl = s.tail[l];
  unsigned j = tl;                                   // next free slot + 1, filling from the tail down
  int removed = 0, removed_run = 0;
  long long len = 0;
  for (unsigned p = tl; p != h; --p) {
    const int2 e = s.ring[base + ((p - 1u) & mask)];
    const int w = s.ag[e.x].rec.w;
    if ((w & 0xff) != SQ_ENROUTE) {
      ++removed; len += (w >> 8) & 0xff;
      if ((int)((p - 1u) - qt) >= 0) ++removed_run;
    } else {
      --j;
      if (j != p - 1u) s.ring[base + (j & mask)] = e;
    }
  }
  if (!removed) return 0;
  s.head[l] = h + (unsigned)removed;
  s.qtail[l] = qt + (unsigned)removed_run;
  s.len_out[l] += len;
  s.l_shelter[l] += removed;
  if (s.cong_pos) s.cong_pos[l] = h + (unsigned)removed;   // the congested prefix is recounted by the next link phase
  return removed;
}

// ---- phase 1: load + link model, one thread per link ---------------------------------------------------
__device__ void reroute_slot(const TickState& s, int i);

// The staggered reroute of this tick (k_reroute's work) rides in the same launch: blocks beyond `link_blocks`
// re-point the agents of this tick's stagger slot.  The two halves touch different words of an agent record
// (reroute: next link, epoch, route kind; load / give-up: enter time, status), both with word-sized stores, and a
// reroute reads nothing the link half writes except the status, where either value gives the same decision.
__global__ void __launch_bounds__(256) k_link_phase(const TickState s, int link_blocks) {
  pdl_wait();
  if (s.pdl_early) pdl_release();
  const int t = *s.tick_link;
  ulonglong2* const tls = timeline_slot(s, t, 0);
  const u64 tl0 = tls ? global_ns() : 0;
  if ((int)blockIdx.x >= link_blocks) {
    reroute_slot(s, ((int)blockIdx.x - link_blocks) * (int)blockDim.x + (int)threadIdx.x);
    if (tls) *tls = make_ulonglong2(tl0, global_ns());
    return;
  }
  const int l = blockIdx.x * blockDim.x + threadIdx.x;
  if (l == 0) {
    *s.tick_node = t;
    // vehicles on links after the previous tick = loaded - arrived up to then; counters[6] is the loaded
    // count frozen by the previous node phase (this launch is adding to counters[5] concurrently)
    // same for the vehicles that sheltered on a link: counters[8] is counters[7] frozen by the node phase
    if (s.counters[0] > 0) s.counters[1] += s.counters[6] - s.counters[4] - s.counters[8];
    s.counters[0] += 1;
  }
  int loaded = 0, sheltered = 0;
  if (l < s.L) {
    if (s.l_tomb && s.l_tomb[l] > 0) {     // shelter_a2 vehicles of this tick's reroutes (k_reroute_rules ran before)
      sheltered += strip_stopped(s, l);
      s.l_tomb[l] = 0;
    }
    // Level 1: every word that does not depend on another load, issued together.  A thread is a chain of dependent
    // L2 round trips (~0.4 us each); read where they are used, these made seven of them.
    const bool virt = s.l_virtual[l] != 0;
    unsigned tl = s.tail[l];
    unsigned qt = s.qtail[l];
    unsigned h = s.head[l];
    const long long base = s.l_ring_base[l];
    const unsigned mask = s.l_ring_mask[l];
    long long li = s.len_in[l];
    long long lout = s.len_out[l];
    const int dwell = s.l_dwell[l];
    const unsigned vcnt = (unsigned)s.l_vcount[l];
    const int in_local = s.l_in_local[l];
    const int end_node = s.l_end[l];
    const double storage = s.l_storage[l];
    const double cap = s.l_cap_s[l];
    const bool closed = s.l_closed[l] != 0;
    // Level 2: the ring entries the three scans below start from (oldest vehicle, first running vehicle, next
    // departure), read together; a scan that goes further loads as it goes.
    const unsigned h0 = h, qt0 = qt, tl0v = tl;
    int2 e_h = make_int2(-1, INT_MAX), e_q = e_h, e_t = e_h;
    const bool want_h = !virt && h != tl && (s.shelter_after > 0);
    const bool want_q = qt != tl;
    const bool want_t = virt && tl < vcnt;
    if (want_h) e_h = s.ring[base + (h & mask)];
    if (want_q) e_q = s.ring[base + (qt & mask)];
    if (want_t) e_t = s.ring[base + tl];
    if (s.congest_after > 0 && !virt) {
      // agent loop of the driver (run_traffic_simulation.py:80-82, 100-102): link.congested = vehicles that have
      // been on this link for more than congest_after seconds, recounted every tick.  Ring order is entry order, so
      // these vehicles are a prefix [head, p); p only moves forward, and the ready tick in the ring entry brackets
      // the enter time (ready - dwell <= enter <= ready - 1), so the agent record is read at the boundary only.
      unsigned p = s.cong_pos[l];
      if ((int)(p - h) < 0) p = h;
      const long long limit = (long long)t - s.congest_after;   // congested <=> enter < limit
      while (p != tl) {
        const int2 e = s.ring[base + (p & mask)];
        bool old;
        if (e.y != INT_MAX && (long long)e.y - 1 < limit) old = true;
        else if (e.y != INT_MAX && (long long)e.y - dwell >= limit) old = false;
        else old = (long long)s.ag[e.x].rec.z < limit;
        if (!old) break;
        ++p;
      }
      s.cong_pos[l] = p;
      s.l_congested[l] = (long long)(p - h);
    }
    if (s.shelter_after > 0 && !virt) {
      // agent loop of the driver (run_traffic_simulation.py:100-104, 113-118): a vehicle that has spent more
      // than shelter_after seconds on this link gives up (status shelter_a1) and is taken off the link before
      // the link model runs.  Ring order is entry order, so these vehicles are a prefix of the ring.
      long long lo = 0;
      int gave_up = 0;
      while (h != tl) {
        const int2 e = h == h0 ? e_h : s.ring[base + (h & mask)];
        if (e.y != INT_MAX && (long long)t - e.y + dwell <= s.shelter_after) break;   // cannot be old enough yet
        const int4 rec = s.ag[e.x].rec;
        if (t - rec.z <= s.shelter_after) break;
        s.ag[e.x].rec.w = (rec.w & ~0xff) | SQ_SHELTER_A1;
        s.a_arrive[e.x] = t;
        lo += (rec.w >> 8) & 0xff;
        ++h; ++gave_up;
      }
      if (gave_up) {
        s.head[l] = h;
        lout += lo;
        s.len_out[l] = lout;
        s.l_shelter[l] += gave_up;
        if ((int)(qt - h) < 0) { qt = h; s.qtail[l] = qt; }   // they were still in the run part
        sheltered += gave_up;
      }
    }
    if (virt) {
      // load_vehicle (:587-596): the next entries of the prefilled departure sequence whose departure == t
      while (tl < vcnt) {
        const int2 e = tl == tl0v ? e_t : s.ring[base + tl];
        if (e.y - 1 > t) break;
        const int w = s.ag[e.x].rec.w;
        s.ag[e.x].rec.z = e.y - 1;                        // current_link_enter_time = departure tick
        s.ag[e.x].rec.w = (w & ~0xff) | SQ_ENROUTE;
        li += (w >> 8) & 0xff;
        ++tl; ++loaded;
      }
      if (loaded) { s.tail[l] = tl; s.len_in[l] = li; }
    }
    // run -> queue promotion (:446-449): advance the boundary over entries whose ready tick has come
    if (qt != tl) {
      const unsigned qs = qt;
      while (qt != tl) {
        const int2 e = (qt == qt0 && want_q) ? e_q : s.ring[base + (qt & mask)];
        if (e.y > t) break;
        ++qt;
      }
      if (qt != qs) s.qtail[l] = qt;
    }
    // candidate bit of the end node: k_node_phase reads one word per node instead of walking its in-links' queue
    // bounds (three dependent loads at the head of every node's chain); fire-and-forget REDs (updating only on change, via a per-link byte, measured slower)
    {
      const u64 bit = 1ull << in_local;
      u64* w = &s.n_cand[end_node];
      if (qt != h) atomicOr(w, bit); else atomicAnd(w, ~bit);
    }
    // capacities (:453-455)
    s.storage_c[l] = storage - (double)(li - lout);
    s.out_c[l] = cap;
    s.in_c[l] = closed ? 0.0 : cap;
  }
  const unsigned any = __ballot_sync(0xffffffffu, (loaded | sheltered) > 0);
  if (any) {
    int tot = loaded, sh = sheltered;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      tot += __shfl_xor_sync(0xffffffffu, tot, o);
      sh += __shfl_xor_sync(0xffffffffu, sh, o);
    }
    if (sq::lane_id() == 0) {
      if (tot) atomicAdd(&s.counters[5], (u64)tot);
      if (sh) atomicAdd(&s.counters[7], (u64)sh);
    }
  }
  if (tls) *tls = make_ulonglong2(tl0, global_ns());
}

// ---- phase 2: node model, one thread per node ---------------------------------------------------------------
struct AdmitCtx {
  int t, node, toss, moves, arrivals;
  long long key;
  long long pair_row;   // index of (this in-link, out-link 0) in the pair-move table, when it is kept
};

// The in-side words of a node's out-links (remaining inflow / storage, run-list tail, length sum) are written by
// this node's thread only.  A vehicle that moves used to pay eight dependent L2 round trips for them; a source
// link can release 27 vehicles a second, one after the other.  They now live in thread-local memory (L1) for
// the duration of the node: filled on the first vehicle that heads for the link, written back at the end.
constexpr int MAXO = 8;     // out-links cached per node (local index < MAXO); further ones take the global path
struct OutCache {
  int link[MAXO];           // -1 = slot not filled
  double in_c[MAXO], storage_c[MAXO];
  unsigned tail[MAXO], room_end[MAXO], mask[MAXO];
  long long len_in[MAXO], base[MAXO];
  int ready[MAXO], end_node[MAXO];
  unsigned dirty;           // bit lc = slot lc must be written back
};

__device__ __forceinline__ void outcache_fill(const TickState& s, OutCache& oc, int lc, int ol, int t) {
  oc.link[lc] = ol;
  oc.in_c[lc] = s.in_c[ol];
  oc.storage_c[lc] = s.storage_c[ol];
  oc.tail[lc] = s.tail[ol];
  oc.mask[lc] = s.l_ring_mask[ol];
  // last tail value the ring has room for, from the head as it is now (the link's end node may be advancing
  // it concurrently, which only adds room)
  oc.room_end[lc] = s.head[ol] + oc.mask[lc];
  oc.len_in[lc] = s.len_in[ol];
  oc.base[lc] = s.l_ring_base[ol];
  oc.ready[lc] = ready_tick(t, s.l_fft[ol]);
  oc.end_node[lc] = s.l_end[ol];
}

__device__ __forceinline__ void outcache_flush(const TickState& s, const OutCache& oc) {
  for (unsigned m = oc.dirty; m; m &= m - 1) {
    const int lc = __ffs((int)m) - 1;
    const int ol = oc.link[lc];
    s.in_c[ol] = oc.in_c[lc];
    s.storage_c[ol] = oc.storage_c[lc];
    s.tail[ol] = oc.tail[lc];
    s.len_in[ol] = oc.len_in[lc];
  }
}

// What the gather phase collected for one queued vehicle.
struct Veh {
  int agent, ol, recw, destin, lc, rlen, nxt;   // nxt: next link after ol for table routes (rlen < 0)
};

// One vehicle of the admit loop (run_node_model :304-386); returns true if it stays in the queue.
__device__ __forceinline__ bool admit_one(const TickState& s, AdmitCtx& c, OutCache& oc, int il, bool secondary, u64 left,
                                          const Veh& v, double& out_c) {
  const int ol = v.ol, agent = v.agent;
  const int vlen = (v.recw >> 8) & 0xff;
  if (secondary && ol >= 0 && ((left >> v.lc) & 1ull)) return true;   // left turn from the opposite approach (:289-292)
  if (c.node == v.destin) {                                           // arrival (:321-328)
    out_c = out_c - 1.0; out_c = out_c > 0.0 ? out_c : 0.0;
    s.ag[agent].rec.w = (v.recw & ~0xff) | SQ_ARRIVE;
    s.a_arrive[agent] = c.t;
    ++c.arrivals;
    return false;
  }
  if (ol < 0) {
    raise_error(s, 1, agent, c.node, il);                             // :312-314
    return true;
  }
  const int lc = v.lc;
  const bool cached = lc < MAXO;
  if (cached && oc.link[lc] != ol) outcache_fill(s, oc, lc, ol, c.t);
  const double storage = cached ? oc.storage_c[lc] : s.storage_c[ol];
  if (storage < (double)vlen) return true;                            // spill-back (:333)
  double in_c = cached ? oc.in_c[lc] : s.in_c[ol];
  bool go;
  if (out_c >= 1.0 && in_c >= 1.0) {                                  // :336
    go = true;
  } else {                                                            // :358-361
    const double cc = out_c < in_c ? out_c : in_c;
    const double w0 = 1.0 - cc;
    const double total = w0 + cc;
    const double u = rng_uniform(s.seed, c.t, c.key, 2 + c.toss);
    ++c.toss;
    go = !(u * total < w0);
    if (!go) {                                                        // :380-386
      if (out_c < in_c) { out_c = out_c - 1.0; out_c = out_c > 0.0 ? out_c : 0.0; }
      else {
        if (!(out_c > in_c)) { out_c = out_c - 1.0; out_c = out_c > 0.0 ? out_c : 0.0; }
        in_c = in_c - 1.0; in_c = in_c > 0.0 ? in_c : 0.0;
        if (cached) { oc.in_c[lc] = in_c; oc.dirty |= 1u << lc; } else s.in_c[ol] = in_c;
      }
      return true;
    }
  }
  // :337-349 / :362-374
  out_c = out_c - 1.0; out_c = out_c > 0.0 ? out_c : 0.0;
  in_c = in_c - 1.0; in_c = in_c > 0.0 ? in_c : 0.0;
  if (cached) {
    oc.in_c[lc] = in_c;
    oc.storage_c[lc] = storage - (double)vlen;
    const unsigned tl = oc.tail[lc];
    if ((int)(tl - oc.room_end[lc]) > 0) raise_error(s, 3, agent, c.node, ol);
    s.ring[oc.base[lc] + (tl & oc.mask[lc])] = make_int2(agent, oc.ready[lc]);   // append to the run list
    oc.tail[lc] = tl + 1u;
    oc.len_in[lc] += vlen;
    oc.dirty |= 1u << lc;
  } else {
    s.in_c[ol] = in_c;
    s.storage_c[ol] = storage - (double)vlen;
    const unsigned tl = s.tail[ol];
    const unsigned omask = s.l_ring_mask[ol];
    if (tl - s.head[ol] > omask) raise_error(s, 3, agent, c.node, ol);
    s.ring[s.l_ring_base[ol] + (tl & omask)] = make_int2(agent, ready_tick(c.t, s.l_fft[ol]));
    s.tail[ol] = tl + 1u;
    s.len_in[ol] += vlen;
  }
  if (s.pair_moves) atomicAdd(&s.pair_moves[c.pair_row + lc], 1u);   // result unused: a RED
  // move_agent_to_next_link + find_next_link (:610-615, :598-608)
  int nxt = v.nxt;
  if (v.rlen >= 0) {
    const int pos = s.a_route_pos[agent] + 1;
    nxt = pos < v.rlen ? route_link_at(s, agent, pos) : -1;
    s.a_route_pos[agent] = pos;
  }
  s.ag[agent].rec = make_int4(nxt, ol, c.t, v.recw);
  return false;
}

#ifndef SQ_GATHER
#define SQ_GATHER 2
#endif
// Vehicles whose records are gathered together (independent loads in flight).  Measured on config 4 (node phase at ticks
// 3000 / 7000, us; first set x later sets): 4x4 12.8 / 22.0, 3x3 13.2 / 21.8, 2x4 13.1 / 22.4, 2x2 11.7 / 21.2, 2x1
// 11.7 / 24.0, 1x1 13.9 / 25.2.  Most windows hold one or two vehicles (whole lanes); the unrolled admit loop runs as
// many iterations as the longest window of the warp has vehicles in the set, and its code size goes with the set.
constexpr int GATHER = SQ_GATHER;

// The window of an in-link the node model works on this tick: its first k queued vehicles.
struct Win {
  long long base, len_out0;
  u64 left;
  double out_c;
  unsigned mask, h;
  int il, k, total_left0, rcol;
};

__device__ __forceinline__ void win_load(const TickState& s, Win& w, int il) {
  w.il = il;
  w.k = 0;
  if (il < 0) return;
  const int4* q = reinterpret_cast<const int4*>(&s.l_win[il]);
  const int4 q0 = q[0], q1 = q[1];
  w.base = ((long long)(unsigned)q0.y << 32) | (unsigned)q0.x;
  w.left = ((u64)(unsigned)q0.w << 32) | (unsigned)q0.z;
  w.mask = (unsigned)q1.x;
  w.rcol = q1.z;
  w.h = s.head[il];
  w.out_c = s.out_c[il];
  w.len_out0 = s.len_out[il];
  w.total_left0 = s.total_left[il];
  w.k = min(q1.y, (int)(s.qtail[il] - w.h));
}

// Everything a vehicle's decision reads from its own records (ring entry -> agent record, destination, route
// kind -> next link's start / local index / end -> next-next link from the router table) is gathered for
// GATHER vehicles at a time, level by level, so the sequential part runs on local data.  The four levels are
// separate functions so that the node kernel can interleave the gathers of two in-links (issue the loads of one
// link's level, then consume the other's).
template <int G>
struct VSetT {
  Veh v[G];
  int st[G], ep[G], dl[G], en[G];
};
typedef VSetT<GATHER> VSet;
#ifndef SQ_GATHER_WIDE
#define SQ_GATHER_WIDE 2
#endif
// Vehicles gathered together beyond the first set of a window (long windows: source links release up to 100 vehicles a
// tick).  8 instead of 4 halves the dependent round trips of a long window but runs slower in the congested hours (node
// phase 23.8 vs 22.0 us at tick 7000): the unrolled admit loop is what a long window costs, not the gather.
constexpr int GATHER_WIDE = SQ_GATHER_WIDE;

template <int G>
__device__ __forceinline__ void lvl_ring(const TickState& s, const Win& w, int i0, VSetT<G>& g) {
  const int cnt = w.k - i0;
#pragma unroll
  for (int j = 0; j < G; ++j) g.v[j].agent = j < cnt ? s.ring[w.base + ((w.h + (unsigned)(i0 + j)) & w.mask)].x : -1;
}

template <int G>
__device__ __forceinline__ void lvl_agent(const TickState& s, VSetT<G>& g) {
#pragma unroll
  for (int j = 0; j < G; ++j) {
    Veh& v = g.v[j];
    v.ol = -1; v.recw = 0; v.destin = -1; v.rlen = 0; g.ep[j] = 0; g.dl[j] = 0;
    if (v.agent >= 0) {
      const int4* q = reinterpret_cast<const int4*>(&s.ag[v.agent]);
      const int4 rec = q[0], aux = q[1];
      v.ol = rec.x; v.recw = rec.w;
      v.destin = aux.x; v.rlen = aux.y;
      g.ep[j] = aux.z; g.dl[j] = aux.w;
    }
  }
}

template <int G>
__device__ __forceinline__ void lvl_hop(const TickState& s, int node, VSetT<G>& g) {
#pragma unroll
  for (int j = 0; j < G; ++j) {
    g.st[j] = node; g.v[j].lc = 0; g.en[j] = 0;
    if (g.v[j].ol >= 0) { const int4 hop = s.l_hop[g.v[j].ol]; g.st[j] = hop.x; g.en[j] = hop.y; g.v[j].lc = hop.z; }
  }
}

template <int G>
__device__ __forceinline__ void lvl_tbl(const TickState& s, VSetT<G>& g) {
#pragma unroll
  for (int j = 0; j < G; ++j) {
    g.v[j].nxt = -1;
    if (g.v[j].ol >= 0 && g.v[j].rlen < 0) g.v[j].nxt = s.next_tbl[tbl_index(s, g.ep[j], g.dl[j], g.en[j])];
  }
}

__device__ __forceinline__ int sel8(const int4& a, const int4& b, int i) {
  int r = a.x;
  r = i == 1 ? a.y : r; r = i == 2 ? a.z : r; r = i == 3 ? a.w : r;
  r = i == 4 ? b.x : r; r = i == 5 ? b.y : r; r = i == 6 ? b.z : r; r = i == 7 ? b.w : r;
  return r;
}

// Experiment build (SQ_NVCC_EXTRA=-DSQ_TICK_PHASES, tools/tick_phases.py): %globaltimer at checkpoints of k_node_phase,
// written by the first active lane of the warp.
#ifdef SQ_TICK_PHASES
#define PHASE(k) phase_mark(s, t, (k))
__device__ __forceinline__ void phase_mark(const TickState& s, int t, int k) {
  if (!s.phases) return;
  const unsigned m = __activemask();
  if ((threadIdx.x & 31u) != (unsigned)(__ffs(m) - 1)) return;
  const int rel = t - s.tl_t0;
  const int w = (int)((blockIdx.x * blockDim.x + threadIdx.x) >> 5);
  if (rel < 0 || rel >= s.tl_ticks || w >= s.tl_stride) return;
  s.phases[((size_t)rel * s.tl_stride + w) * 16 + k] = global_ns();
}
#else
#define PHASE(k)
#endif

// What the admit loop accumulates for one in-link.
struct LinkAcc {
  u64 stay_lo = 0, stay_hi = 0;   // bit i = vehicle i stays in the queue
  long long moved_len = 0;
  int moved = 0;
  bool any_left = false;
};

// Admit loop of run_node_model (:304-386) over the gathered vehicles [i0, i0 + GATHER) of the window; `secondary`
// applies the "no more left turns from the opposite approach" skip of :289-292.  For the primary link the same
// records answer non_conflict_vehicles' question "does any of them turn left" (:248-263): acc.any_left is set
// from the next links the vehicles had before anything moved.
template <int G>
__device__ __forceinline__ void admit_set(const TickState& s, AdmitCtx& c, OutCache& oc, Win& w, bool secondary,
                                          const VSetT<G>& g, int i0, LinkAcc& acc) {
  const int cnt = min(G, w.k - i0);
#pragma unroll
  for (int j = 0; j < G; ++j) {
    if (j >= cnt) break;
    const int i = i0 + j;
    const Veh& v = g.v[j];
    bool stay;
    if (v.ol >= 0 && g.st[j] != c.node) {
      raise_error(s, 2, v.agent, c.node, v.ol);                                    // KeyError of :257-262 / :291
      stay = true;
    } else {
      if (v.ol >= 0 && ((w.left >> v.lc) & 1ull)) acc.any_left = true;
      stay = admit_one(s, c, oc, w.il, secondary, w.left, v, w.out_c);
    }
    if (stay) { if (i < 64) acc.stay_lo |= 1ull << i; else acc.stay_hi |= 1ull << (i - 64); }
    else { ++acc.moved; acc.moved_len += (v.recw >> 8) & 0xff; }
  }
}

// The rest of a window beyond its first (pre-gathered) set, then the queue bookkeeping of the link.
__device__ void admit_finish(const TickState& s, AdmitCtx& c, OutCache& oc, Win& w, bool secondary, LinkAcc& acc) {
  for (int i0 = GATHER; i0 < w.k; i0 += GATHER_WIDE) {
    VSetT<GATHER_WIDE> g;
    lvl_ring(s, w, i0, g);
    lvl_agent(s, g);
    lvl_hop(s, c.node, g);
    lvl_tbl(s, g);
    admit_set(s, c, oc, w, secondary, g, i0, acc);
  }
  if (acc.moved) {
    // close the gaps: stayers slide to the back of the window [h, h+k), preserving order
    const int k = w.k;
    int j = k - 1;
    for (int i = k - 1; i >= 0; --i) {
      const bool st = i < 64 ? ((acc.stay_lo >> i) & 1ull) : ((acc.stay_hi >> (i - 64)) & 1ull);
      if (st) {
        if (j != i) s.ring[w.base + ((w.h + (unsigned)j) & w.mask)] = s.ring[w.base + ((w.h + (unsigned)i) & w.mask)];
        --j;
      }
    }
    s.head[w.il] = w.h + (unsigned)acc.moved;
    s.len_out[w.il] = w.len_out0 + acc.moved_len;
    s.total_left[w.il] = w.total_left0 + acc.moved;
    c.moves += acc.moved;
  }
  s.out_c[w.il] = w.out_c;
}

// Choice of the opposite approach (:267-293) among the remaining candidates `rest`, assuming the primary link
// has no left turner: the in-link whose turn onto the primary's reverse link is the straightest.
__device__ __forceinline__ int pick_secondary(const TickState& s, int n, int ib, int indeg, int type, u64 rest, int first,
                                              int rcol) {
  const int outdeg = s.out_ptr[n + 1] - s.out_ptr[n];
  const double* tb = s.turn + s.turn_ptr[n];
  int secondary = -1;
  double best = 0.0;
  // iteration order of the remaining list: [first (if any), others in original order]; the priority branch
  // iterates in priority order
  if (type == 2) {
    int done_rank = -1;
    for (;;) {
      int br = INT_MAX, bi = -1;
      for (int i = 0; i < indeg; ++i) {
        const int r = s.in_prio[ib + i];
        if (((rest >> i) & 1ull) && r > done_rank && r < br) { br = r; bi = i; }
      }
      if (bi < 0) break;
      done_rank = br;
      const double a = fabs(tb[bi * outdeg + rcol]);
      if (secondary < 0 || a < best) { secondary = bi; best = a; }
    }
  } else {
    if (first >= 0 && ((rest >> first) & 1ull)) { secondary = first; best = fabs(tb[first * outdeg + rcol]); }
    for (int i = 0; i < indeg; ++i) {
      if (!((rest >> i) & 1ull) || i == first) continue;
      const double a = fabs(tb[i * outdeg + rcol]);
      if (secondary < 0 || a < best) { secondary = i; best = a; }
    }
  }
  return secondary;
}

__global__ void __launch_bounds__(32) k_node_phase(const TickState s) {
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  // static words of the node, read while the link phase is still running
  int ib = 0, indeg = 0, type = 0;
  long long key = 0;
  int4 nin0 = make_int4(-1, -1, -1, -1), nin1 = nin0, nsec0 = make_int4(0, 0, 0, 0), nsec1 = nsec0;
  if (n < s.N) {
    ib = s.in_ptr[n]; indeg = s.in_ptr[n + 1] - ib; key = s.n_key[n]; type = s.n_type[n];
    const int4* q = reinterpret_cast<const int4*>(&s.n_rec[n]);
    nin0 = __ldg(q); nin1 = __ldg(q + 1); nsec0 = __ldg(q + 2); nsec1 = __ldg(q + 3);
  }
  pdl_wait();
  if (s.pdl_early) pdl_release();
  const int t = *s.tick_node;
  ulonglong2* const tls = timeline_slot(s, t, 1);
  const u64 tl0 = tls ? global_ns() : 0;
  PHASE(0);
  if (n == 0) { *s.tick_link = t + 1; s.counters[6] = s.counters[5]; s.counters[8] = s.counters[7]; }
  int moves = 0, arrivals = 0, active = 0;
  if (n < s.N) {
    // candidate in-links: non-empty queue, incoming_links order (:230); the link phase keeps the mask
    u64 cand = s.n_cand[n];
    if (cand) {
      PHASE(1);
      active = 1;
      int primary = -1;  // local index of the primary in-link
      // `order` semantics: after the shuffle replacement the list is [c_r, rest in original order].
      int first = -1;    // local index moved to the front by the last shuffle (ties in :277 see it first)
      if (type == 2) {                                                            // priority list (:237-238)
        int best_rank = INT_MAX;
        u64 listed = 0;
        for (int i = 0; i < indeg; ++i) {
          const int r = s.in_prio[ib + i];
          if (r >= 0 && ((cand >> i) & 1ull)) { listed |= 1ull << i; if (r < best_rank) { best_rank = r; primary = i; } }
        }
        cand = listed;
      } else {
        int cnt = __popcll(cand);
        int r = 0;
        if (cnt >= 2) r = (int)(rng_uniform(s.seed, t, key, 0) * (double)cnt);   // shuffle slot 0 (:231)
        {
          u64 m = cand;
          for (int j = 0; j < r; ++j) m &= m - 1;
          first = __ffsll((long long)m) - 1;
        }
        if (type == 1) {                                                          // signal phase (:233-236)
          const int want = (t % 60 < 30) ? 1 : 2;
          u64 f = 0;
          for (int i = 0; i < indeg; ++i)
            if (((cand >> i) & 1ull) && (s.in_phase[ib + i] & want)) f |= 1ull << i;
          // filtered list order: [first (if kept), others in original order]
          const bool first_kept = (f >> first) & 1ull;
          cnt = __popcll(f);
          cand = f;
          if (cnt == 0) { primary = -1; }
          else {
            int r2 = 0;
            if (cnt >= 2) r2 = (int)(rng_uniform(s.seed, t, key, 1) * (double)cnt);   // shuffle slot 1
            // element r2 of [first?, rest...]
            if (first_kept) {
              if (r2 == 0) primary = first;
              else { u64 m = f & ~(1ull << first); for (int j = 0; j < r2 - 1; ++j) m &= m - 1; primary = __ffsll((long long)m) - 1; }
            } else {
              u64 m = f; for (int j = 0; j < r2; ++j) m &= m - 1; primary = __ffsll((long long)m) - 1;
              first = -1;
            }
            // after the second shuffle the list is [primary, (first if kept and != primary), rest]
            if (first == primary) first = -1;
          }
        } else {
          primary = first;
          first = -1;
        }
      }
      if (primary >= 0) {
        PHASE(2);
        AdmitCtx c;
        c.t = t; c.node = n; c.toss = 0; c.moves = 0; c.arrivals = 0; c.key = key;
        OutCache oc;
#pragma unroll
        for (int i = 0; i < MAXO; ++i) oc.link[i] = -1;
        oc.dirty = 0u;
        const bool rec_ok = !((unsigned)nsec0.x & NR_NODE_SLOW);
        const int lp = rec_ok ? sel8(nin0, nin1, primary) : s.in_links[ib + primary];
        // The primary link's vehicles go first (run_node_model walks the go-list in order); the same pass tells
        // whether any of them turns left (:248-263), which decides if the opposite approach moves at all.
        // Nothing the choice of that approach reads (candidate set, angle table, the other links' queues), and
        // nothing its vehicles' decisions are gathered from (their ring entries, agent records, next-link
        // table), is touched by the primary admits.  So the opposite approach is chosen and its vehicles'
        // records are gathered *speculatively*, interleaved level by level with the primary's gather: one
        // thread keeps both chains of dependent loads in flight, and the admits themselves stay in the
        // reference's order.
        Win wp, ws;
        const u64 rest = cand & ~(1ull << primary);
        VSet gp, gs;
        int secondary = -1;
        const unsigned sw = rec_ok ? (unsigned)sel8(nsec0, nsec1, primary) : NR_SLOW;
        if (!(sw & (NR_SLOW | NR_NODE_SLOW))) {
          // the opposite approach from the node's static record: both windows are loaded together
          if (rest && !(sw & NR_NONE)) {
            const int cnt = (int)(sw & 7u);
            for (int a = 0; a < cnt; ++a) {
              const int j = (int)((sw >> (3 + 3 * a)) & 7u);
              if ((rest >> j) & 1ull) { secondary = j; break; }
            }
          }
          win_load(s, wp, lp);
          win_load(s, ws, secondary >= 0 ? sel8(nin0, nin1, secondary) : -1);
          lvl_ring(s, wp, 0, gp);
        } else {
          win_load(s, wp, lp);
          lvl_ring(s, wp, 0, gp);
          if (rest && wp.rcol >= 0) secondary = pick_secondary(s, n, ib, indeg, type, rest, first, wp.rcol);
          win_load(s, ws, secondary >= 0 ? s.in_links[ib + secondary] : -1);
        }
        lvl_ring(s, ws, 0, gs);
        PHASE(3);
        lvl_agent(s, gp);
        lvl_agent(s, gs);
        PHASE(4);
        lvl_hop(s, n, gp);
        lvl_hop(s, n, gs);
        PHASE(5);
        lvl_tbl(s, gp);
        lvl_tbl(s, gs);
        PHASE(6);
        LinkAcc ap;
        long long pair0 = 0;
        int outdeg = 0;
        if (s.pair_moves) { pair0 = s.turn_ptr[n]; outdeg = s.out_ptr[n + 1] - s.out_ptr[n]; }
        c.pair_row = pair0 + (long long)primary * outdeg;
        admit_set(s, c, oc, wp, false, gp, 0, ap);
        PHASE(7);
        admit_finish(s, c, oc, wp, false, ap);
        PHASE(8);
        if (!ap.any_left && ws.il >= 0) {                                         // opposite approach (:267-293)
          LinkAcc as;
          c.pair_row = pair0 + (long long)secondary * outdeg;
          admit_set(s, c, oc, ws, true, gs, 0, as);
          PHASE(9);
          admit_finish(s, c, oc, ws, true, as);
          PHASE(10);
        }
        outcache_flush(s, oc);
        PHASE(11);
        if (c.moves) atomicAdd(reinterpret_cast<u64*>(&s.node_move[n]), (u64)c.moves);   // result unused: a fire-and-forget RED
        moves = c.moves; arrivals = c.arrivals;
      }
    }
  }
  // run counters: one atomic per warp
  const unsigned any = __ballot_sync(0xffffffffu, active != 0);
  if (any) {
    int m = moves, a = arrivals, ac = active;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      m += __shfl_xor_sync(0xffffffffu, m, o);
      a += __shfl_xor_sync(0xffffffffu, a, o);
      ac += __shfl_xor_sync(0xffffffffu, ac, o);
    }
    if (sq::lane_id() == 0) {
      if (m) atomicAdd(&s.counters[2], (u64)m);
      atomicAdd(&s.counters[3], (u64)ac);
      if (a) atomicAdd(&s.counters[4], (u64)a);
    }
  }
  PHASE(12);
  if (tls) *tls = make_ulonglong2(tl0, global_ns());
}

// ---- host-ordered stops (sq_stop_agents): a driver rule took vehicles off the road ----------------------------
// Drivers strip sheltering agents from their link's lists straight away (model/dta_meso_butte.py:256-263,
// run_traffic_simulation.py:113-118).  Pass 1 marks the agents and their links, pass 2 lets one thread per
// marked link close the gaps in its ring: the survivors slide towards the tail, in order, and the head and the
// queue / run boundary move up by what was removed on their side.
__global__ void k_stop_mark(const TickState s, int n, const int* __restrict__ agent_idx, int status, int t, int* tomb) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int a = agent_idx[i];
  const int4 rec = s.ag[a].rec;
  if ((rec.w & 0xff) != SQ_ENROUTE) { raise_error(s, 5, a, -1, rec.y); return; }
  s.ag[a].rec.w = (rec.w & ~0xff) | (status & 0xff);
  s.a_arrive[a] = t;
  atomicAdd(&tomb[rec.y], 1);
}

__global__ void k_stop_compact(const TickState s, int n, const int* __restrict__ agent_idx, int* tomb) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int l = s.ag[agent_idx[i]].rec.y;
  if (atomicExch(&tomb[l], 0) <= 0) return;          // another thread of this launch owns the link (or error above)
  const int removed = strip_stopped(s, l);
  if (removed) atomicAdd(&s.counters[7], (u64)removed);
}

// ---- route replacement (Agent.get_path + find_next_link) -----------------------------------------------
__global__ void k_set_routes(const TickState s, int n, const int* agent_idx, const long long* new_off, const int* new_len) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int a = agent_idx[i];
  int4 rec = s.ag[a].rec;
  const int st = rec.w & 0xff;
  if (st != SQ_UNLOADED && st != SQ_ENROUTE) return;
  s.a_route_off[a] = new_off[i];
  s.ag[a].route_len = new_len[i];
  s.a_route_pos[a] = 0;
  rec.x = new_len[i] > 0 ? s.route_pool[new_off[i]] : -1;   // host-supplied routes always live in route_pool
  s.ag[a].rec = rec;
}

// ---- device router ------------------------------------------------------------------------------------------
// Edge weights for routing (Link.update_travel_time_by_queue_length, queue_class.py:463-466): one thread per
// link, evaluated left to right like the Python expression; closed and source links keep their weight (NaN).
__global__ void __launch_bounds__(256) k_link_weights(const TickState s) {
  const int l = blockIdx.x * 256 + threadIdx.x;
  if (l >= s.L) return;
  const int e = s.l_edge[l];
  if (e < 0) return;
  double w = __longlong_as_double(0x7ff8000000000000ll);
  if (!s.l_closed[l] && !s.l_virtual[l]) {
    const double q = (double)(int)(s.qtail[l] - s.head[l]);
    w = q / (s.l_capacity[l] + 0.1) * 3600.0 * 25.0 / s.l_maxmph[l] + s.l_fft[l];
  }
  s.w_edge[e] = w;
}

// Reverse trees (dist / next vertex towards each destination, indexed by graph vertex) -> tables indexed by
// node: the link to take is the node2link choice for (node, next node) (queue_class.py:149-156, 605).
__global__ void __launch_bounds__(256) k_build_tables(const TickState s, int slot, int V, const double* __restrict__ tdist,
                                                      const int* __restrict__ tnext, const int* __restrict__ pair_ptr,
                                                      const int* __restrict__ pair_endv, const int* __restrict__ pair_link) {
  const int n = blockIdx.x * 256 + threadIdx.x;
  const int d = blockIdx.y;
  if (n >= s.N) return;
  const int v = s.n_vertex[n];
  const double dist = tdist[(size_t)d * V + v];
  const int succ = tnext[(size_t)d * V + v];
  int link = -1;
  if (succ >= 0)
    for (int k = pair_ptr[n]; k < pair_ptr[n + 1]; ++k)
      if (pair_endv[k] == succ) { link = pair_link[k]; break; }
  const size_t i = tbl_index(s, slot, d, n);
  s.dist_tbl[i] = dist;
  s.next_tbl[i] = link;
  if (n == 0 && d == 0) *s.cur_epoch = slot;
}

// Agent.get_path + find_next_link for the agents due this tick (t % freq == agent_id % freq,
// run_traffic_simulation.py:107-109): adopt the newest tables unless the destination is out of reach
// (distance > 1e8 keeps the old route, queue_class.py:511-513).  `all` != 0 routes every agent (t == 0).
__device__ __forceinline__ void reroute_agent(const TickState& s, int a, int t) {
  const int4 rec = s.ag[a].rec;
  const int st = rec.w & 0xff;
  if (st != SQ_UNLOADED && st != SQ_ENROUTE) return;
  const int node = s.l_end[rec.y];
  const int e = *s.cur_epoch;
  const size_t idx = tbl_index(s, e, s.ag[a].dslot, node);
  const double d = s.dist_tbl[idx];
  if (d > 1e8) return;
  s.ag[a].epoch = e;
  s.ag[a].route_len = -1;
  s.ag[a].rec.x = s.next_tbl[idx];
  if (s.a_rdist) {
    // previous_distance = agent.route_distance; get_path; shelter_a2 if the trip more than doubled
    // (model/dta_meso_butte.py:195-198, 201-204)
    const double prev = s.a_rdist[a];
    s.a_rdist[a] = d;
    if (s.detour_factor > 0.0 && st == SQ_ENROUTE && d > prev * s.detour_factor) {
      s.ag[a].rec.w = (rec.w & ~0xff) | SQ_SHELTER_A2;
      s.a_arrive[a] = t;
      atomicAdd(&s.l_tomb[rec.y], 1);      // the link phase of this tick takes the vehicle off the link
    }
  }
}

// i-th agent of this tick's stagger slot (t % freq == key % freq, run_traffic_simulation.py:107)
__device__ void reroute_slot(const TickState& s, int i) {
  if (!s.bucket_ptr) return;
  const int t = *s.tick_link;
  if (t <= 0) return;
  const int b = t % s.freq;
  const int k = s.bucket_ptr[b] + i;
  if (k >= s.bucket_ptr[b + 1]) return;
  reroute_agent(s, s.bucket_agents[k], t);
}

__global__ void __launch_bounds__(128) k_reroute_all(const TickState s) {
  const int a = blockIdx.x * 128 + threadIdx.x;
  if (a < s.A) reroute_agent(s, a, 0);
}

// The agent loop's reroute rules with Butte's additions (model/dta_meso_butte.py:193-209), one launch BEFORE the link
// phase (a sheltering vehicle must be off its link before the link model of the same second, :256-263):
//   regular stagger slot  t % freq  == key % freq   -> get_path (+ the detour test in reroute_agent);
//   closure stagger slot  t % cfreq == key % cfreq  -> get_path again if the agent's next link is closed.
// An agent due in both slots is handled by its closure-slot thread, in the reference's order.
// A thread serves (at most) one agent of the closure slot and one of the regular slot; they are different agents, so
// the words the two decisions start from are read together, level by level (bucket bounds -> agent ids -> records ->
// link words), and reroute_agent_with() gets what it needs in registers: ten dependent round trips became six.
__device__ __forceinline__ void reroute_agent_with(const TickState& s, int a, int t, const int4& rec, int dslot, int node, int e) {
  const int st = rec.w & 0xff;
  if (st != SQ_UNLOADED && st != SQ_ENROUTE) return;
  const size_t idx = tbl_index(s, e, dslot, node);
  const double d = s.dist_tbl[idx];
  const int nl = s.next_tbl[idx];
  double prev = 0.0;
  if (s.a_rdist) prev = s.a_rdist[a];
  if (d > 1e8) return;
  s.ag[a].epoch = e;
  s.ag[a].route_len = -1;
  s.ag[a].rec.x = nl;
  if (s.a_rdist) {
    // previous_distance = agent.route_distance; get_path; shelter_a2 if the trip more than doubled
    // (model/dta_meso_butte.py:195-198, 201-204)
    s.a_rdist[a] = d;
    if (s.detour_factor > 0.0 && st == SQ_ENROUTE && d > prev * s.detour_factor) {
      s.ag[a].rec.w = (rec.w & ~0xff) | SQ_SHELTER_A2;
      s.a_arrive[a] = t;
      atomicAdd(&s.l_tomb[rec.y], 1);      // the link phase of this tick takes the vehicle off the link
    }
  }
}

__global__ void __launch_bounds__(128) k_reroute_rules(const TickState s) {
  const int i = blockIdx.x * 128 + threadIdx.x;
  pdl_wait();
  if (s.pdl_early) pdl_release();
  const int t = *s.tick_link;
  const int e = *s.cur_epoch;
  if (t <= 0) return;
  // level 2: bucket bounds
  const int b = t % s.freq;
  const int b2 = s.cfreq > 0 ? t % s.cfreq : 0;
  const int k = s.bucket_ptr[b] + i;
  const bool have = k < s.bucket_ptr[b + 1];
  int k2 = 0;
  bool have2 = false;
  if (s.cfreq > 0) { k2 = s.cbucket_ptr[b2] + i; have2 = k2 < s.cbucket_ptr[b2 + 1]; }
  // level 3: agent ids
  const int a = have ? s.bucket_agents[k] : 0;
  const int a2 = have2 ? s.cbucket_agents[k2] : 0;
  // level 4: records
  int4 rec = make_int4(-1, 0, 0, SQ_ARRIVE), rec2 = rec;
  int ds = 0, ds2 = 0, cslot = -1, rslot2 = -1;
  if (have) { rec = s.ag[a].rec; ds = s.ag[a].dslot; if (s.cfreq > 0) cslot = s.a_cslot[a]; }
  if (have2) { rec2 = s.ag[a2].rec; ds2 = s.ag[a2].dslot; rslot2 = s.a_rslot[a2]; }
  // level 5: link words
  const int node = have ? s.l_end[rec.y] : 0;
  const int node2 = have2 ? s.l_end[rec2.y] : 0;
  if (have2) {
    if (rslot2 == b) {
      // due in both slots: the regular reroute first, then the closure test on the link it chose
      reroute_agent_with(s, a2, t, rec2, ds2, node2, e);
      rec2 = s.ag[a2].rec;
    }
    const int st2 = rec2.w & 0xff;
    if ((st2 == SQ_UNLOADED || st2 == SQ_ENROUTE) && rec2.x >= 0 && s.l_closed[rec2.x]) reroute_agent_with(s, a2, t, rec2, ds2, node2, e);
  }
  if (have && !(s.cfreq > 0 && cslot == b2)) reroute_agent_with(s, a, t, rec, ds, node, e);
}

// Before table slot `slot` is overwritten: agents that still follow it (their reroutes kept failing) get
// their remaining route written out as a link sequence, so that they keep the old route like the reference.
__global__ void __launch_bounds__(128) k_evacuate(const TickState s, int slot) {
  const int a = blockIdx.x * 128 + threadIdx.x;
  if (a >= s.A) return;
  if (s.ag[a].route_len >= 0 || s.ag[a].epoch != slot) return;
  const int4 rec = s.ag[a].rec;
  const int st = rec.w & 0xff;
  if (st != SQ_UNLOADED && st != SQ_ENROUTE) return;
  const int d = s.ag[a].dslot;
  int len = 0;
  for (int l = rec.x; l >= 0 && len <= s.L; l = s.next_tbl[tbl_index(s, slot, d, s.l_end[l])]) ++len;
  const long long off = (long long)atomicAdd((u64*)s.mat_used, (u64)len);
  if (off + len > s.mat_cap) { raise_error(s, 4, a, -1, -1); return; }
  int k = 0;
  for (int l = rec.x; l >= 0 && k < len; l = s.next_tbl[tbl_index(s, slot, d, s.l_end[l])]) s.mat_pool[off + k++] = l;
  s.a_route_off[a] = -(off + 1);
  s.ag[a].route_len = len;
  s.a_route_pos[a] = 0;
}

// Output cadence (row N4; the t_stats / link_stats reductions of model/dta_meso_butte.py:291-348): one pass over
// agents, links and nodes into 16 counters, so a driver's periodic log line costs a 128-byte read-back instead of
// a full state sync.  out: [0..7] agents by status code, [8] queued vehicles, [9] running vehicles, [10] congested
// vehicles (last tick), [11] closed links, [12] sum of node_move, [13] sum of total_vehicles_left over real links,
// [14] real links with any vehicle on them, [15] vehicles on closed links.
__global__ void __launch_bounds__(256) k_summary(const TickState s, unsigned long long* out) {
  __shared__ unsigned long long acc[16];
  if (threadIdx.x < 16) acc[threadIdx.x] = 0ull;
  __syncthreads();
  const int stride = gridDim.x * blockDim.x;
  for (int a = blockIdx.x * blockDim.x + threadIdx.x; a < s.A; a += stride)
    atomicAdd(&acc[s.ag[a].rec.w & 7], 1ull);
  for (int l = blockIdx.x * blockDim.x + threadIdx.x; l < s.L; l += stride) {
    const unsigned q = s.qtail[l] - s.head[l], r = s.tail[l] - s.qtail[l];
    if (q) atomicAdd(&acc[8], (unsigned long long)q);
    if (r) atomicAdd(&acc[9], (unsigned long long)r);
    if (s.l_congested[l]) atomicAdd(&acc[10], (unsigned long long)s.l_congested[l]);
    if (s.l_closed[l]) { atomicAdd(&acc[11], 1ull); if (q + r) atomicAdd(&acc[15], (unsigned long long)(q + r)); }
    if (!s.l_virtual[l]) {
      if (s.total_left[l]) atomicAdd(&acc[13], (unsigned long long)s.total_left[l]);
      if (q + r) atomicAdd(&acc[14], 1ull);
    }
  }
  for (int n = blockIdx.x * blockDim.x + threadIdx.x; n < s.N; n += stride)
    if (s.node_move[n]) atomicAdd(&acc[12], (unsigned long long)s.node_move[n]);
  __syncthreads();
  if (threadIdx.x < 16 && acc[threadIdx.x]) atomicAdd(&out[threadIdx.x], acc[threadIdx.x]);
}

// Closure schedule (row N3): events [lo, hi) of the device event table take effect.  A closed link gets no
// inflow from the next link phase on (queue_class.py:455) and weighs 1e8 for routing; a reopened one weighs its
// free-flow time again (close_link_to_newcomers / open_link_to_newcomers, queue_class.py:467-474).  w_edge was
// filled with NaN ("keep") by k_fill_nan, so sp_set_weights_device touches only these edges.
__global__ void k_fill_nan(double* w, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) w[i] = __longlong_as_double(0x7ff8000000000000ll);
}
__global__ void k_apply_events(const TickState s, const int* __restrict__ ev_link, const unsigned char* __restrict__ ev_closed,
                               int lo, int hi, int route) {
  const int i = lo + blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= hi) return;
  const int l = ev_link[i];
  const unsigned char c = ev_closed[i];
  s.l_closed[l] = c;
  if (route && s.l_edge[l] >= 0) s.w_edge[s.l_edge[l]] = c ? 1.0e8 : s.l_fft[l];
}

__global__ void k_set_closed(const TickState s, int n, const int* link_idx, const unsigned char* closed) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) s.l_closed[link_idx[i]] = closed[i];
}

// =================================================================================================
// Host side
// =================================================================================================
struct DevBuf {
  void* p = nullptr;
  size_t bytes = 0;
};

struct SqSim {
  TickState st;
  std::vector<void*> allocs;
  int N = 0, L = 0, A = 0;
  long long ring_slots = 0;
  size_t n_turn = 0;   // entries of the turn table (sum over nodes of in-degree x out-degree)
  // host copies needed by getters
  std::vector<long long> h_ring_base;
  std::vector<unsigned> h_ring_mask;
  std::vector<unsigned char> h_virtual, h_closed;
  // closure schedule: host copy (sorted by tick) + device event table
  std::vector<int> ev_tick, ev_link;
  std::vector<unsigned char> ev_closed;
  size_t ev_cursor = 0;
  int* d_ev_link = nullptr;
  unsigned char* d_ev_closed = nullptr;
  unsigned long long* d_summary = nullptr;
  std::vector<double> h_capacity, h_maxmph, h_fft;
  // route pool management
  int* d_pool = nullptr;
  long long pool_cap = 0, pool_used = 0;
  cudaStream_t stream = nullptr;
  cudaStream_t cap_stream = nullptr;
  bool own_stream = true;
  bool pdl = true;       // programmatic dependent launch between the per-tick kernels (SQ_TICK_PDL=0 turns it off)
  int next_t = INT_MIN;  // tick the device counters are armed for
  long long sheltered = 0;
  // device router
  Graph* router_graph = nullptr;
  int wfreq = 0, max_bucket = 0, graph_V = 0, graph_E = 0;
  int last_refresh_t = INT_MIN;
  int host_epoch = 0;
  int* d_dest_vertex = nullptr;     // [D]
  double* d_tdist = nullptr;        // [D][V] reverse-tree scratch
  int* d_tnext = nullptr;
  int* d_pair_ptr = nullptr; int* d_pair_endv = nullptr; int* d_pair_link = nullptr;
  std::vector<long long> agent_key;   // router: agent ids (stagger slots)
  long long launches = 0;
  cudaGraphExec_t graph = nullptr;
  int graph_ticks = 0;
};

template <typename T>
static int dev_upload(SqSim* sim, const T** dst, const std::vector<T>& v) {
  T* d = nullptr;
  const size_t n = std::max<size_t>(v.size(), 1);
  SQ_CUDA_TRY(cudaMalloc(&d, n * sizeof(T)));
  sim->allocs.push_back(d);
  if (!v.empty()) SQ_CUDA_TRY(cudaMemcpy(d, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice));
  *dst = d;
  return SQ_OK;
}
template <typename T>
static int dev_upload_rw(SqSim* sim, T** dst, const std::vector<T>& v) {
  const T* c = nullptr;
  int rc = dev_upload(sim, &c, v);
  *dst = const_cast<T*>(c);
  return rc;
}
template <typename T>
static int dev_upload_ptr(SqSim* sim, const T** dst, const T* src, size_t n) {
  std::vector<T> v(src, src + n);
  return dev_upload(sim, dst, v);
}

// The static node records (NodeRec): host-only, also exported for the CPU test that checks the opposite-approach lists
// against the walk of pick_secondary for every candidate set (sq_build_node_records).
static void build_node_records(const SqNetDesc* d, const std::vector<unsigned char>& out_local, std::vector<NodeRec>& nrec) {
  const int N = d->n_nodes;
  nrec.assign((size_t)N, NodeRec());
  for (int n = 0; n < N; ++n) {
    NodeRec& r = nrec[n];
    const int ib = d->node_in_ptr[n], indeg = d->node_in_ptr[n + 1] - ib;
    const int outdeg = d->node_out_ptr[n + 1] - d->node_out_ptr[n];
    for (int i = 0; i < 8; ++i) { r.in[i] = i < indeg ? d->node_in_links[ib + i] : -1; r.sec[i] = NR_NONE; }
    if (indeg > 8) { r.sec[0] |= NR_NODE_SLOW; continue; }
    bool dup_rank = false;   // pick_secondary's priority walk skips links that share a rank: keep that path
    if (d->node_type[n] == 2)
      for (int a = 0; a < indeg; ++a)
        for (int b = a + 1; b < indeg; ++b)
          if (d->node_in_prio[ib + a] >= 0 && d->node_in_prio[ib + a] == d->node_in_prio[ib + b]) dup_rank = true;
    const double* tb = d->turn_angle + d->node_turn_ptr[n];
    for (int i = 0; i < indeg; ++i) {
      const int rev = d->link_reverse[r.in[i]];
      if (rev < 0) continue;                                    // NR_NONE
      const int rcol = (int)out_local[rev];
      int order[8], cnt = 0;
      bool slow = dup_rank || rcol >= outdeg;
      for (int j = 0; j < indeg; ++j) if (j != i) order[cnt++] = j;
      auto ang = [&](int j) { return std::fabs(tb[(size_t)j * outdeg + rcol]); };
      for (int a = 0; a < cnt && !slow; ++a) {
        if (!(ang(order[a]) == ang(order[a]))) slow = true;     // NaN
        for (int b = a + 1; b < cnt && !slow; ++b) if (ang(order[a]) == ang(order[b])) slow = true;
      }
      if (slow) { r.sec[i] = NR_SLOW; continue; }
      std::sort(order, order + cnt, [&](int a, int b) { return ang(a) < ang(b); });
      unsigned w = (unsigned)cnt;
      for (int a = 0; a < cnt; ++a) w |= (unsigned)order[a] << (3 + 3 * a);
      r.sec[i] = w;
    }
  }
}

static int sim_build(SqSim* sim, const SqNetDesc* d) {
  if (int rc = sq::require_device()) return rc;
  const int N = d->n_nodes, L = d->n_links, A = d->n_agents;
  if (N <= 0 || L <= 0 || A < 0) return sq::set_error(SQ_ERR_ARG, "sq_create: empty network (N=%d L=%d A=%d)", N, L, A);
  sim->N = N; sim->L = L; sim->A = A;
  sim->n_turn = (size_t)d->node_turn_ptr[N];
  TickState& s = sim->st;
  s.N = N; s.L = L; s.A = A; s.seed = d->seed;
  SQ_CUDA_TRY(cudaStreamCreateWithFlags(&sim->stream, cudaStreamNonBlocking));

  for (long long k = 0; k < d->route_ptr[A]; ++k)
    if (d->route_links[k] < 0 || d->route_links[k] >= L)
      return sq::set_error(SQ_ERR_ARG, "route entry %lld is link %d, outside 0..%d", k, d->route_links[k], L - 1);
  // ---- per-link local indices, left-turn masks ----
  std::vector<unsigned char> in_local(L, 0), out_local(L, 0);
  std::vector<u64> left_mask(L, 0);
  for (int n = 0; n < N; ++n) {
    const int ib = d->node_in_ptr[n], ie = d->node_in_ptr[n + 1], ob = d->node_out_ptr[n], oe = d->node_out_ptr[n + 1];
    if (ie - ib > 64 || oe - ob > 64)
      return sq::set_error(SQ_ERR_LIMIT, "node %d has in/out degree %d/%d; the turn masks hold 64", n, ie - ib, oe - ob);
    for (int j = ob; j < oe; ++j) out_local[d->node_out_links[j]] = (unsigned char)(j - ob);
    for (int i = ib; i < ie; ++i) {
      const int l = d->node_in_links[i];
      in_local[l] = (unsigned char)(i - ib);
      u64 m = 0;
      for (int j = ob; j < oe; ++j)
        if (d->turn_angle[d->node_turn_ptr[n] + (long long)(i - ib) * (oe - ob) + (j - ob)] >= 50) m |= 1ull << (j - ob);
      left_mask[l] = m;
    }
  }
  // ---- rings ----
  int min_vlen = 255;
  for (int a = 0; a < A; ++a) {
    if (d->agent_vlen[a] < 1 || d->agent_vlen[a] > 255)
      return sq::set_error(SQ_ERR_LIMIT, "agent %d has vehicle length %d; 1..255 supported", a, d->agent_vlen[a]);
    min_vlen = std::min(min_vlen, d->agent_vlen[a]);
  }
  if (A == 0) min_vlen = 8;
  std::vector<int> virt_of_node(N, -1);
  for (int l = 0; l < L; ++l) {
    if (d->link_lanes_floor[l] > 128)
      return sq::set_error(SQ_ERR_LIMIT, "link %d has %d lanes; 128 supported", l, d->link_lanes_floor[l]);
    if (d->link_is_virtual[l]) virt_of_node[d->link_end_node[l]] = l;
  }
  std::vector<std::vector<int>> departures(L);
  for (int a = 0; a < A; ++a) {
    if (d->agent_status[a] != SQ_UNLOADED) continue;  // shelter_p agents never load (queue_class.py:509, :589)
    const int vl = virt_of_node[d->agent_origin[a]];
    if (vl < 0) return sq::set_error(SQ_ERR_ARG, "agent %d: origin node %d has no source link", a, d->agent_origin[a]);
    departures[vl].push_back(a);
  }
  std::vector<long long> ring_base(L);
  std::vector<unsigned> ring_mask(L);
  std::vector<int> vcount(L, 0);
  long long slots = 0;
  for (int l = 0; l < L; ++l) {
    long long need;
    if (d->link_is_virtual[l]) {
      need = (long long)departures[l].size();
      vcount[l] = (int)need;
    } else {
      need = (long long)std::floor(d->link_storage_cap[l] / (double)min_vlen) + 2;
    }
    long long cap = 1;
    while (cap < need) cap <<= 1;
    if (cap > (1ll << 30)) return sq::set_error(SQ_ERR_LIMIT, "link %d needs a ring of %lld slots", l, cap);
    ring_base[l] = slots;
    ring_mask[l] = (unsigned)(cap - 1);
    slots += cap;
  }
  sim->ring_slots = slots;
  std::vector<int2> ring((size_t)slots, make_int2(-1, INT_MAX));
  for (int l = 0; l < L; ++l) {
    auto& v = departures[l];
    std::stable_sort(v.begin(), v.end(), [&](int a, int b) { return d->agent_dept[a] < d->agent_dept[b]; });
    for (size_t i = 0; i < v.size(); ++i) ring[(size_t)ring_base[l] + i] = make_int2(v[i], d->agent_dept[v[i]] + 1);
  }

  // ---- uploads ----
  int rc;
#define UP(field, ptr, n) if ((rc = dev_upload_ptr(sim, &s.field, ptr, (size_t)(n)))) return rc
  UP(l_start, d->link_start_node, L); UP(l_end, d->link_end_node, L); UP(l_lanes, d->link_lanes_floor, L);
  UP(l_rev, d->link_reverse, L); UP(l_virtual, d->link_is_virtual, L);
  UP(l_cap_s, d->link_cap_per_s, L); UP(l_storage, d->link_storage_cap, L); UP(l_fft, d->link_fft, L);
  UP(n_type, d->node_type, N); UP(in_ptr, d->node_in_ptr, N + 1); UP(in_links, d->node_in_links, d->node_in_ptr[N]);
  UP(in_phase, d->node_in_phase, d->node_in_ptr[N]); UP(in_prio, d->node_in_prio, d->node_in_ptr[N]);
  UP(out_ptr, d->node_out_ptr, N + 1); UP(out_links, d->node_out_links, d->node_out_ptr[N]);
  UP(turn, d->turn_angle, d->node_turn_ptr[N]);
#undef UP
  {
    std::vector<long long> key(d->node_key, d->node_key + N), tp(d->node_turn_ptr, d->node_turn_ptr + N + 1);
    if ((rc = dev_upload(sim, &s.n_key, key))) return rc;
    if ((rc = dev_upload(sim, &s.turn_ptr, tp))) return rc;
  }
  {
    std::vector<unsigned char> closed(d->link_closed, d->link_closed + L);
    if ((rc = dev_upload_rw(sim, &s.l_closed, closed))) return rc;
    sim->h_closed = closed;
  }
  if ((rc = dev_upload(sim, &s.l_ring_base, ring_base))) return rc;
  if ((rc = dev_upload(sim, &s.l_ring_mask, ring_mask))) return rc;
  if ((rc = dev_upload(sim, &s.l_vcount, vcount))) return rc;
  if ((rc = dev_upload(sim, &s.l_in_local, in_local))) return rc;
  if ((rc = dev_upload(sim, &s.l_out_local, out_local))) return rc;
  {
    std::vector<int4> hop((size_t)L);
    for (int l = 0; l < L; ++l) hop[l] = make_int4(d->link_start_node[l], d->link_end_node[l], (int)out_local[l], 0);
    if ((rc = dev_upload(sim, &s.l_hop, hop))) return rc;
  }
  if ((rc = dev_upload(sim, &s.l_left_mask, left_mask))) return rc;
  {
    std::vector<WinRec> win((size_t)L);
    for (int l = 0; l < L; ++l) {
      const int rev = d->link_reverse[l];
      win[l].base = ring_base[l]; win[l].left = left_mask[l]; win[l].mask = ring_mask[l];
      win[l].lanes = d->link_lanes_floor[l]; win[l].rcol = rev >= 0 ? (int)out_local[rev] : -1; win[l].pad = 0;
    }
    if ((rc = dev_upload(sim, &s.l_win, win))) return rc;
  }
  {
    std::vector<NodeRec> nrec;
    build_node_records(d, out_local, nrec);
    if ((rc = dev_upload(sim, &s.n_rec, nrec))) return rc;
  }
  if ((rc = dev_upload_rw(sim, &s.ring, ring))) return rc;
  {
    std::vector<unsigned> z(L, 0u);
    if ((rc = dev_upload_rw(sim, &s.head, z))) return rc;
    if ((rc = dev_upload_rw(sim, &s.qtail, z))) return rc;
    if ((rc = dev_upload_rw(sim, &s.tail, z))) return rc;
    std::vector<double> caps(d->link_cap_per_s, d->link_cap_per_s + L), stor(d->link_storage_cap, d->link_storage_cap + L);
    if ((rc = dev_upload_rw(sim, &s.out_c, caps))) return rc;
    if ((rc = dev_upload_rw(sim, &s.in_c, caps))) return rc;
    if ((rc = dev_upload_rw(sim, &s.storage_c, stor))) return rc;
    std::vector<long long> zl(L, 0);
    if ((rc = dev_upload_rw(sim, &s.len_in, zl))) return rc;
    if ((rc = dev_upload_rw(sim, &s.len_out, zl))) return rc;
    std::vector<int> zi(L, 0);
    if ((rc = dev_upload_rw(sim, &s.total_left, zi))) return rc;
    std::vector<long long> zn(N, 0);
    if ((rc = dev_upload_rw(sim, &s.node_move, zn))) return rc;
  }
  {
    std::vector<int> dwell(L), zi(L, 0);
    for (int l = 0; l < L; ++l) {
      const double f = d->link_fft[l];
      dwell[l] = f < 2.0e9 ? (int)std::floor(f) + 2 : INT_MAX / 2;
    }
    if ((rc = dev_upload(sim, &s.l_dwell, dwell))) return rc;
    if ((rc = dev_upload_rw(sim, &s.l_shelter, zi))) return rc;
    if ((rc = dev_upload_ptr(sim, &s.l_capacity, d->link_capacity, (size_t)L))) return rc;
    if ((rc = dev_upload_ptr(sim, &s.l_maxmph, d->link_maxmph, (size_t)L))) return rc;
    s.shelter_after = 0;
    s.congest_after = 0;
    {
      std::vector<unsigned> zu((size_t)std::max(L, 1), 0u);
      std::vector<long long> zl((size_t)std::max(L, 1), 0ll);
      if ((rc = dev_upload_rw(sim, &s.cong_pos, zu))) return rc;
      std::vector<u64> zc64(N, 0);
      if ((rc = dev_upload_rw(sim, &s.n_cand, zc64))) return rc;
      if ((rc = dev_upload_rw(sim, &s.l_congested, zl))) return rc;
    }
    s.pair_moves = nullptr;
    s.a_rdist = nullptr; s.detour_factor = 0.0; s.cfreq = 0; s.cbucket_ptr = nullptr; s.cbucket_agents = nullptr;
    s.a_rslot = nullptr; s.a_cslot = nullptr; s.max_cbucket = 0; s.l_tomb = nullptr;
    s.D = 0; s.K = 0; s.freq = 1;
    s.n_vertex = nullptr; s.next_tbl = nullptr; s.dist_tbl = nullptr;
    s.cur_epoch = nullptr; s.bucket_ptr = nullptr; s.bucket_agents = nullptr; s.l_edge = nullptr; s.w_edge = nullptr;
    s.mat_pool = nullptr; s.mat_used = nullptr; s.mat_cap = 0;
  }
  // agents
  {
    std::vector<AgentRec> rec((size_t)A);
    std::vector<long long> off((size_t)A);
    std::vector<int> pos((size_t)A, 0), arr((size_t)A, -1);
    for (int a = 0; a < A; ++a) {
      const long long b = d->route_ptr[a], e = d->route_ptr[a + 1];
      off[a] = b;
      const int vl = virt_of_node[d->agent_origin[a]];
      rec[a].rec = make_int4(e > b ? d->route_links[b] : -1, vl, -1, (int)d->agent_status[a] | (d->agent_vlen[a] << 8));
      rec[a].destin = d->agent_destin[a]; rec[a].route_len = (int)(e - b); rec[a].epoch = 0; rec[a].dslot = 0;
      if (d->agent_status[a] == SQ_UNLOADED && e == b && !(d->reserved0 & 1))
        return sq::set_error(SQ_ERR_ARG, "agent %d is to be loaded but has an empty route (queue_class.py:595 raises KeyError)", a);
    }
    if ((rc = dev_upload_rw(sim, &s.ag, rec))) return rc;
    if ((rc = dev_upload_rw(sim, &s.a_route_off, off))) return rc;
    if ((rc = dev_upload_rw(sim, &s.a_route_pos, pos))) return rc;
    if ((rc = dev_upload_rw(sim, &s.a_arrive, arr))) return rc;
    const long long used = d->route_ptr[A];
    sim->pool_cap = used + used / 2 + 4096;
    sim->pool_used = used;
    SQ_CUDA_TRY(cudaMalloc(&sim->d_pool, sizeof(int) * (size_t)sim->pool_cap));
    if (used > 0) SQ_CUDA_TRY(cudaMemcpy(sim->d_pool, d->route_links, sizeof(int) * (size_t)used, cudaMemcpyHostToDevice));
    s.route_pool = sim->d_pool;
  }
  {
    std::vector<int> z2(1, 0);
    if ((rc = dev_upload_rw(sim, &s.tick_link, z2))) return rc;
    if ((rc = dev_upload_rw(sim, &s.tick_node, z2))) return rc;
    std::vector<u64> zc(12, 0);
    if ((rc = dev_upload_rw(sim, &s.counters, zc))) return rc;
    std::vector<int> ze(4, 0);
    if ((rc = dev_upload_rw(sim, &s.err, ze))) return rc;
  }
  sim->h_ring_base = ring_base; sim->h_ring_mask = ring_mask;
  sim->h_virtual.assign(d->link_is_virtual, d->link_is_virtual + L);
  sim->h_capacity.assign(d->link_capacity, d->link_capacity + L);
  sim->h_maxmph.assign(d->link_maxmph, d->link_maxmph + L);
  sim->h_fft.assign(d->link_fft, d->link_fft + L);
  return SQ_OK;
}

// Launch with the programmatic-serialisation attribute (see pdl_wait): in a captured graph it becomes a programmatic
// dependency edge on the previous kernel node.
template <typename... P, typename... A>
static void launch_pdl(bool pdl, void (*kern)(P...), unsigned grid, unsigned block, cudaStream_t st, A&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid); cfg.blockDim = dim3(block); cfg.dynamicSmemBytes = 0; cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at; cfg.numAttrs = pdl ? 1 : 0;
  cudaLaunchKernelEx(&cfg, kern, std::forward<A>(args)...);
}

static int launch_tick(SqSim* sim, cudaStream_t st) {
  const int L = sim->L, N = sim->N;
  const int lb = 128;   // measured: 31.5 / 31.5 / 32.1 us per tick with 64 / 128 / 256 threads
  const int link_blocks = (L + lb - 1) / lb;
  int reroute_blocks = (sim->router_graph && sim->max_bucket > 0) ? (sim->max_bucket + lb - 1) / lb : 0;
  if (sim->st.a_rdist) {
    // Butte's rules: the reroutes run as their own launch before the link phase
    const int work = std::max(sim->max_bucket, sim->st.max_cbucket);
    if (work > 0) launch_pdl(sim->pdl, k_reroute_rules, (unsigned)((work + 127) / 128), 128u, st, sim->st);
    reroute_blocks = 0;
  }
  launch_pdl(sim->pdl, k_link_phase, (unsigned)(link_blocks + reroute_blocks), (unsigned)lb, st, sim->st, link_blocks);
  // one warp per CTA: the warps are independent chains, the finer grain spreads them evenly over the SMs (measured:
  // 31.7 / 32.1 / 32.5 us per tick with 32 / 64 / 128 threads)
  launch_pdl(sim->pdl, k_node_phase, (unsigned)((N + 31) / 32), 32u, st, sim->st);
  return SQ_OK;
}

static const int kGraphTicks = 32;  // ticks per CUDA-graph launch

static int build_graph(SqSim* sim) {
  if (sim->graph) { cudaGraphExecDestroy(sim->graph); sim->graph = nullptr; }
  cudaGraph_t g;
  // capture on a private stream: the run stream may be the caller's (sq_set_stream), e.g. the legacy default
  // stream, which cannot be captured
  if (!sim->cap_stream) SQ_CUDA_TRY(cudaStreamCreateWithFlags(&sim->cap_stream, cudaStreamNonBlocking));
  SQ_CUDA_TRY(cudaStreamBeginCapture(sim->cap_stream, cudaStreamCaptureModeThreadLocal));
  for (int i = 0; i < kGraphTicks; ++i) launch_tick(sim, sim->cap_stream);
  SQ_CUDA_TRY(cudaStreamEndCapture(sim->cap_stream, &g));
  SQ_CUDA_TRY(cudaGraphInstantiate(&sim->graph, g, 0));
  cudaGraphDestroy(g);
  sim->graph_ticks = kGraphTicks;
  return SQ_OK;
}

static int arm_tick(SqSim* sim, int t) {
  if (sim->next_t != t) {
    SQ_CUDA_TRY(cudaMemcpyAsync(sim->st.tick_link, &t, sizeof(int), cudaMemcpyHostToDevice, sim->stream));
    SQ_CUDA_TRY(cudaStreamSynchronize(sim->stream));  // &t is a stack variable
  }
  return SQ_OK;
}

static int check_device_error(SqSim* sim) {
  int e[4];
  SQ_CUDA_TRY(cudaMemcpyAsync(e, sim->st.err, sizeof(e), cudaMemcpyDeviceToHost, sim->stream));
  SQ_CUDA_TRY(cudaStreamSynchronize(sim->stream));
  if (e[0] == 0) return SQ_OK;
  static const char* what[] = {"", "queued vehicle has no next link away from its destination (queue_class.py:312-314)",
                               "vehicle's next link does not leave this node (KeyError at queue_class.py:257-262)",
                               "vehicle ring overflow (link storage bound violated)",
                               "device route pool exhausted while keeping the old route of an unroutable agent",
                               "sq_stop_agents: agent is not on a link (status != enroute)"};
  return sq::set_error(SQ_ERR_STATE, "%s: agent %d, node %d, link %d", what[e[0] < 6 ? e[0] : 0], e[1], e[2], e[3]);
}

// Weight refresh + new routing tables (what the drivers do every `wfreq` ticks before the agent loop,
// run_traffic_simulation.py:81-84, followed by one reverse tree per destination instead of one Dijkstra per agent).
static int router_refresh(SqSim* sim, int t, bool queue_weights = true) {
  if (!sim->router_graph || (queue_weights && sim->last_refresh_t == t)) return SQ_OK;
  TickState& s = sim->st;
  cudaStream_t st = sim->stream;
  const int slot = (sim->host_epoch + 1) % s.K;
  if (sim->last_refresh_t != INT_MIN) {  // the slot is about to be recycled: stragglers keep their old route
    k_evacuate<<<(sim->A + 127) / 128, 128, 0, st>>>(s, slot);
    sq::count_launch(); sim->launches += 1;
  }
  if (queue_weights) {
    k_link_weights<<<(sim->L + 255) / 256, 256, 0, st>>>(s);
    sq::count_launch(); sim->launches += 1;
    SQ_CUDA_TRY(cudaGetLastError());
    if (int rc = sp_set_weights_device(sim->router_graph, s.w_edge, sim->graph_E, (void*)st))
      return sq::set_error(rc, "router refresh: %s", sp_last_error());
  }
  if (int rc = sp_sssp_batch_device_ex(sim->router_graph, sim->d_dest_vertex, s.D, sim->d_tdist, sim->d_tnext, (void*)st, 1))
    return sq::set_error(rc, "router refresh: %s", sp_last_error());
  dim3 grid((unsigned)((sim->N + 255) / 256), (unsigned)s.D);
  k_build_tables<<<grid, 256, 0, st>>>(s, slot, sim->graph_V, sim->d_tdist, sim->d_tnext, sim->d_pair_ptr, sim->d_pair_endv,
                                       sim->d_pair_link);
  sq::count_launch(); sim->launches += 1;
  SQ_CUDA_TRY(cudaGetLastError());
  sim->host_epoch = slot;
  if (queue_weights) sim->last_refresh_t = t;
  return SQ_OK;
}

// Host-side edits of the routing graph (link.close_link_to_newcomers / open_link_to_newcomers -> update_edge,
// queue_class.py:467-474) take effect for every later get_path call in the reference: rebuild the tables before
// the next tick whenever the graph has pending edits.
static int router_catch_up(SqSim* sim, int t) {
  if (!sim->router_graph || !sp_has_pending_edits(sim->router_graph)) return SQ_OK;
  if (sim->wfreq > 0 && t % sim->wfreq == 0 && sim->last_refresh_t != t) return SQ_OK;  // the regular refresh follows
  return router_refresh(sim, t, false);
}

// Events of the closure schedule due at or before tick t, applied before that tick's link phase; with a router
// attached the routing graph and the next-link tables follow at once (the reference's next get_path sees the edit).
static int apply_closure_events(SqSim* sim, int t) {
  if (sim->ev_cursor >= sim->ev_tick.size() || sim->ev_tick[sim->ev_cursor] > t) return SQ_OK;
  const int lo = (int)sim->ev_cursor;
  int hi = lo;
  while (hi < (int)sim->ev_tick.size() && sim->ev_tick[hi] <= t) ++hi;
  sim->ev_cursor = (size_t)hi;
  TickState& s = sim->st;
  cudaStream_t st = sim->stream;
  const int route = sim->router_graph ? 1 : 0;
  if (route) {
    k_fill_nan<<<(sim->graph_E + 255) / 256, 256, 0, st>>>(s.w_edge, sim->graph_E);
    sq::count_launch(); sim->launches += 1;
  }
  // one launch per distinct tick so that a link closed and reopened at two past ticks ends in its last state
  for (int a = lo; a < hi;) {
    int b = a;
    while (b < hi && sim->ev_tick[b] == sim->ev_tick[a]) ++b;
    k_apply_events<<<(b - a + 127) / 128, 128, 0, st>>>(s, sim->d_ev_link, sim->d_ev_closed, a, b, route);
    sq::count_launch(); sim->launches += 1;
    a = b;
  }
  SQ_CUDA_TRY(cudaGetLastError());
  for (int i = lo; i < hi; ++i) sim->h_closed[sim->ev_link[i]] = sim->ev_closed[i];
  if (route) {
    if (int rc = sp_set_weights_device(sim->router_graph, s.w_edge, sim->graph_E, (void*)st))
      return sq::set_error(rc, "closure schedule: %s", sp_last_error());
    if (!(sim->wfreq > 0 && t % sim->wfreq == 0 && sim->last_refresh_t != t))   // else the regular refresh follows
      if (int rc = router_refresh(sim, t, false)) return rc;
  }
  return SQ_OK;
}

static int next_event_tick(const SqSim* sim) {
  return sim->ev_cursor < sim->ev_tick.size() ? sim->ev_tick[sim->ev_cursor] : INT_MAX;
}

static int kernels_per_tick(const SqSim* sim) { return sim->st.a_rdist ? 3 : 2; }

extern "C" {

SqSim* sq_create(const SqNetDesc* desc) {
  if (!desc) { sq::set_error(SQ_ERR_ARG, "sq_create: null description"); return nullptr; }
  SqSim* sim = new SqSim();
  if (const char* e = getenv("SQ_TICK_PDL")) sim->pdl = atoi(e) != 0;
  sim->st.pdl_early = 0;
  if (const char* e = getenv("SQ_TICK_PDL")) sim->st.pdl_early = atoi(e) == 2;
  if (sim_build(sim, desc) != SQ_OK || build_graph(sim) != SQ_OK) { sq_destroy(sim); return nullptr; }
  return sim;
}

void sq_destroy(SqSim* sim) {
  if (!sim) return;
  if (sim->graph) cudaGraphExecDestroy(sim->graph);
  for (void* p : sim->allocs) cudaFree(p);
  cudaFree(sim->d_pool);
  cudaFree(sim->st.timeline);
  cudaFree(sim->d_ev_link); cudaFree(sim->d_ev_closed); cudaFree(sim->d_summary);
  if (sim->stream && sim->own_stream) cudaStreamDestroy(sim->stream);
  if (sim->cap_stream) cudaStreamDestroy(sim->cap_stream);
  delete sim;
}

int sq_step(SqSim* sim, int32_t t) {
  if (!sim) return sq::set_error(SQ_ERR_ARG, "sq_step: null sim");
  if (int rc = arm_tick(sim, t)) return rc;
  if (int rc = router_catch_up(sim, t)) return rc;
  if (int rc = apply_closure_events(sim, t)) return rc;
  if (sim->router_graph && sim->wfreq > 0 && t % sim->wfreq == 0)
    if (int rc = router_refresh(sim, t)) return rc;
  launch_tick(sim, sim->stream);
  SQ_CUDA_TRY(cudaGetLastError());
  sq::count_launch(kernels_per_tick(sim)); sim->launches += kernels_per_tick(sim);
  sim->next_t = t + 1;
  return SQ_OK;
}

int sq_run(SqSim* sim, int32_t t0, int32_t n) {
  if (!sim) return sq::set_error(SQ_ERR_ARG, "sq_run: null sim");
  if (n <= 0) return SQ_OK;
  if (int rc = arm_tick(sim, t0)) return rc;
  if (!sim->graph && n >= kGraphTicks)
    if (int rc = build_graph(sim)) return rc;
  if (int rc = router_catch_up(sim, t0)) return rc;
  // segments between routing refreshes; each is replayed as CUDA graphs of kGraphTicks ticks + single ticks
  int t = t0;
  const int end = t0 + n;
  while (t < end) {
    int seg_end = end;
    if (int rc = apply_closure_events(sim, t)) return rc;
    if (sim->router_graph && sim->wfreq > 0) {
      if (t % sim->wfreq == 0)
        if (int rc = router_refresh(sim, t)) return rc;
      seg_end = std::min(end, (t / sim->wfreq + 1) * sim->wfreq);
    }
    seg_end = std::min(seg_end, next_event_tick(sim));   // > t: everything due by t has just been applied
    int left = seg_end - t;
    while (sim->graph && left >= sim->graph_ticks) {
      SQ_CUDA_TRY(cudaGraphLaunch(sim->graph, sim->stream));
      left -= sim->graph_ticks;
    }
    for (int i = 0; i < left; ++i) launch_tick(sim, sim->stream);
    t = seg_end;
  }
  SQ_CUDA_TRY(cudaGetLastError());
  sq::count_launch((long long)kernels_per_tick(sim) * n); sim->launches += (long long)kernels_per_tick(sim) * n;
  sim->next_t = t0 + n;
  return SQ_OK;
}

int sq_set_stream(SqSim* sim, void* stream) {
  if (!sim) return sq::set_error(SQ_ERR_ARG, "sq_set_stream: null sim");
  SQ_CUDA_TRY(cudaStreamSynchronize(sim->stream));
  if (sim->own_stream) { cudaStreamDestroy(sim->stream); sim->own_stream = false; }
  sim->stream = (cudaStream_t)stream;
  return SQ_OK;
}

int sq_sync(SqSim* sim) {
  if (!sim) return sq::set_error(SQ_ERR_ARG, "sq_sync: null sim");
  SQ_CUDA_TRY(cudaStreamSynchronize(sim->stream));
  return check_device_error(sim);
}

int sq_set_routes(SqSim* sim, int32_t n, const int32_t* agent_idx, const int64_t* ptr, const int32_t* links) {
  if (!sim || (n > 0 && (!agent_idx || !ptr))) return sq::set_error(SQ_ERR_ARG, "sq_set_routes: null argument");
  if (n <= 0) return SQ_OK;
  const long long add = ptr[n] - ptr[0];
  for (int i = 0; i < n; ++i)
    if (agent_idx[i] < 0 || agent_idx[i] >= sim->A) return sq::set_error(SQ_ERR_ARG, "sq_set_routes: agent index %d out of range", agent_idx[i]);
  for (long long k = ptr[0]; k < ptr[n]; ++k)
    if (links[k] < 0 || links[k] >= sim->L) return sq::set_error(SQ_ERR_ARG, "sq_set_routes: link %d out of range", links[k]);
  if (sim->pool_used + add > sim->pool_cap) {
    // grow the pool (old routes stay where they are; superseded ones become garbage until the next rebuild)
    const long long ncap = (sim->pool_used + add) * 2 + 4096;
    int* np_ = nullptr;
    SQ_CUDA_TRY(cudaMalloc(&np_, sizeof(int) * (size_t)ncap));
    SQ_CUDA_TRY(cudaMemcpyAsync(np_, sim->d_pool, sizeof(int) * (size_t)sim->pool_used, cudaMemcpyDeviceToDevice, sim->stream));
    SQ_CUDA_TRY(cudaStreamSynchronize(sim->stream));
    cudaFree(sim->d_pool);
    sim->d_pool = np_; sim->pool_cap = ncap; sim->st.route_pool = np_;
    if (sim->graph) { cudaGraphExecDestroy(sim->graph); sim->graph = nullptr; }  // kernel params changed
  }
  std::vector<long long> off(n);
  std::vector<int> len(n);
  for (int i = 0; i < n; ++i) { off[i] = sim->pool_used + (ptr[i] - ptr[0]); len[i] = (int)(ptr[i + 1] - ptr[i]); }
  int* d_idx = nullptr; long long* d_off = nullptr; int* d_len = nullptr;
  SQ_CUDA_TRY(cudaMalloc(&d_idx, sizeof(int) * (size_t)n));
  SQ_CUDA_TRY(cudaMalloc(&d_off, sizeof(long long) * (size_t)n));
  SQ_CUDA_TRY(cudaMalloc(&d_len, sizeof(int) * (size_t)n));
  if (add > 0)
    SQ_CUDA_TRY(cudaMemcpyAsync(sim->d_pool + sim->pool_used, links + ptr[0], sizeof(int) * (size_t)add, cudaMemcpyHostToDevice, sim->stream));
  SQ_CUDA_TRY(cudaMemcpyAsync(d_idx, agent_idx, sizeof(int) * (size_t)n, cudaMemcpyHostToDevice, sim->stream));
  SQ_CUDA_TRY(cudaMemcpyAsync(d_off, off.data(), sizeof(long long) * (size_t)n, cudaMemcpyHostToDevice, sim->stream));
  SQ_CUDA_TRY(cudaMemcpyAsync(d_len, len.data(), sizeof(int) * (size_t)n, cudaMemcpyHostToDevice, sim->stream));
  k_set_routes<<<(n + 127) / 128, 128, 0, sim->stream>>>(sim->st, n, d_idx, d_off, d_len);
  sq::count_launch(); sim->launches += 1;
  SQ_CUDA_TRY(cudaGetLastError());
  SQ_CUDA_TRY(cudaStreamSynchronize(sim->stream));
  cudaFree(d_idx); cudaFree(d_off); cudaFree(d_len);
  sim->pool_used += add;
  return SQ_OK;
}

int sq_stop_agents(SqSim* sim, int32_t n, const int32_t* agent_idx, int32_t status, int32_t t) {
  if (!sim || (n > 0 && !agent_idx)) return sq::set_error(SQ_ERR_ARG, "sq_stop_agents: null argument");
  if (status < SQ_SHELTER_P || status > 255) return sq::set_error(SQ_ERR_ARG, "sq_stop_agents: status %d is not a stopped status", status);
  if (n <= 0) return SQ_OK;
  for (int i = 0; i < n; ++i)
    if (agent_idx[i] < 0 || agent_idx[i] >= sim->A) return sq::set_error(SQ_ERR_ARG, "sq_stop_agents: agent index %d out of range", agent_idx[i]);
  int* d_idx = nullptr; int* d_tomb = nullptr;
  SQ_CUDA_TRY(cudaMalloc(&d_idx, sizeof(int) * (size_t)n));
  SQ_CUDA_TRY(cudaMalloc(&d_tomb, sizeof(int) * (size_t)sim->L));
  SQ_CUDA_TRY(cudaMemsetAsync(d_tomb, 0, sizeof(int) * (size_t)sim->L, sim->stream));
  SQ_CUDA_TRY(cudaMemcpyAsync(d_idx, agent_idx, sizeof(int) * (size_t)n, cudaMemcpyHostToDevice, sim->stream));
  k_stop_mark<<<(n + 127) / 128, 128, 0, sim->stream>>>(sim->st, n, d_idx, status, t, d_tomb);
  k_stop_compact<<<(n + 127) / 128, 128, 0, sim->stream>>>(sim->st, n, d_idx, d_tomb);
  sq::count_launch(2); sim->launches += 2;
  SQ_CUDA_TRY(cudaGetLastError());
  const int rc = check_device_error(sim);
  cudaFree(d_idx); cudaFree(d_tomb);
  return rc;
}

int sq_set_congest_after(SqSim* sim, int32_t seconds) {
  if (!sim) return sq::set_error(SQ_ERR_ARG, "sq_set_congest_after: null sim");
  SQ_CUDA_TRY(cudaStreamSynchronize(sim->stream));
  sim->st.congest_after = seconds > 0 ? seconds : 0;
  if (sim->graph) { cudaGraphExecDestroy(sim->graph); sim->graph = nullptr; }  // kernel parameters changed
  return SQ_OK;
}

int sq_get_summary(SqSim* sim, int64_t* out16) {
  if (!sim || !out16) return sq::set_error(SQ_ERR_ARG, "sq_get_summary: null argument");
  if (!sim->d_summary) SQ_CUDA_TRY(cudaMalloc(&sim->d_summary, sizeof(unsigned long long) * 16));
  SQ_CUDA_TRY(cudaMemsetAsync(sim->d_summary, 0, sizeof(unsigned long long) * 16, sim->stream));
  const int work = std::max(sim->A, std::max(sim->L, sim->N));
  const int grid = std::max(1, std::min((work + 255) / 256, 148 * 8));
  k_summary<<<grid, 256, 0, sim->stream>>>(sim->st, sim->d_summary);
  sq::count_launch(); sim->launches += 1;
  SQ_CUDA_TRY(cudaGetLastError());
  SQ_CUDA_TRY(cudaMemcpyAsync(out16, sim->d_summary, sizeof(unsigned long long) * 16, cudaMemcpyDeviceToHost, sim->stream));
  return sq_sync(sim);
}

int sq_get_link_congested(SqSim* sim, int64_t* out) {
  if (!sim || !out) return sq::set_error(SQ_ERR_ARG, "sq_get_link_congested: null argument");
  SQ_CUDA_TRY(cudaMemcpyAsync(out, sim->st.l_congested, sizeof(long long) * (size_t)sim->L, cudaMemcpyDeviceToHost, sim->stream));
  return sq_sync(sim);
}

int sq_enable_pair_moves(SqSim* sim) {
  if (!sim) return sq::set_error(SQ_ERR_ARG, "sq_enable_pair_moves: null sim");
  if (sim->st.pair_moves) return SQ_OK;
  SQ_CUDA_TRY(cudaStreamSynchronize(sim->stream));
  const size_t n = std::max<size_t>(sim->n_turn, 1);
  SQ_CUDA_TRY(cudaMalloc(&sim->st.pair_moves, sizeof(unsigned) * n));
  sim->allocs.push_back(sim->st.pair_moves);
  SQ_CUDA_TRY(cudaMemset(sim->st.pair_moves, 0, sizeof(unsigned) * n));
  if (sim->graph) { cudaGraphExecDestroy(sim->graph); sim->graph = nullptr; }  // kernel parameters changed
  return SQ_OK;
}

int sq_get_pair_moves(SqSim* sim, int64_t* out) {
  if (!sim || !out) return sq::set_error(SQ_ERR_ARG, "sq_get_pair_moves: null argument");
  if (!sim->st.pair_moves) return sq::set_error(SQ_ERR_ARG, "sq_get_pair_moves: call sq_enable_pair_moves first");
  std::vector<unsigned> h(sim->n_turn);
  SQ_CUDA_TRY(cudaMemcpyAsync(h.data(), sim->st.pair_moves, sizeof(unsigned) * sim->n_turn, cudaMemcpyDeviceToHost, sim->stream));
  if (int rc = sq_sync(sim)) return rc;
  for (size_t i = 0; i < sim->n_turn; ++i) out[i] = (int64_t)h[i];
  return SQ_OK;
}

int sq_enable_timeline(SqSim* sim, int32_t t0, int32_t n_ticks) {
  if (!sim || n_ticks <= 0) return sq::set_error(SQ_ERR_ARG, "sq_enable_timeline: null sim or empty window");
  SQ_CUDA_TRY(cudaStreamSynchronize(sim->stream));
  TickState& s = sim->st;
  const int stride = std::max((sim->L + 31) / 32 + (sim->A + 31) / 32 / std::max(s.freq, 1) + 64, (sim->N + 31) / 32);
  const size_t bytes = sizeof(ulonglong2) * (size_t)n_ticks * 2 * (size_t)stride;
  if (s.timeline) { cudaFree(s.timeline); s.timeline = nullptr; }
  SQ_CUDA_TRY(cudaMalloc(&s.timeline, bytes));
  SQ_CUDA_TRY(cudaMemset(s.timeline, 0, bytes));
  s.tl_t0 = t0; s.tl_ticks = n_ticks; s.tl_stride = stride;
#ifdef SQ_TICK_PHASES
  if (s.phases) { cudaFree(s.phases); s.phases = nullptr; }
  SQ_CUDA_TRY(cudaMalloc(&s.phases, sizeof(u64) * 16 * (size_t)n_ticks * (size_t)stride));
  SQ_CUDA_TRY(cudaMemset(s.phases, 0, sizeof(u64) * 16 * (size_t)n_ticks * (size_t)stride));
#endif
  if (sim->graph) { cudaGraphExecDestroy(sim->graph); sim->graph = nullptr; }  // kernel parameters changed
  return SQ_OK;
}

int sq_get_timeline(SqSim* sim, int32_t* n_ticks, int32_t* stride, uint64_t* out) {
  if (!sim || !sim->st.timeline) return sq::set_error(SQ_ERR_ARG, "sq_get_timeline: call sq_enable_timeline first");
  if (n_ticks) *n_ticks = sim->st.tl_ticks;
  if (stride) *stride = sim->st.tl_stride;
  if (!out) return SQ_OK;
  const size_t bytes = sizeof(ulonglong2) * (size_t)sim->st.tl_ticks * 2 * (size_t)sim->st.tl_stride;
  SQ_CUDA_TRY(cudaMemcpyAsync(out, sim->st.timeline, bytes, cudaMemcpyDeviceToHost, sim->stream));
  return sq_sync(sim);
}

#ifdef SQ_TICK_PHASES
int sq_get_phases(SqSim* sim, uint64_t* out) {   // experiment build only: [tl_ticks][tl_stride][16]
  if (!sim || !sim->st.phases || !out) return sq::set_error(SQ_ERR_ARG, "sq_get_phases: not enabled");
  SQ_CUDA_TRY(cudaMemcpyAsync(out, sim->st.phases, sizeof(u64) * 16 * (size_t)sim->st.tl_ticks * (size_t)sim->st.tl_stride,
                              cudaMemcpyDeviceToHost, sim->stream));
  return sq_sync(sim);
}
#endif

int sq_build_node_records(const SqNetDesc* d, int32_t* out) {
  if (!d || !out) return sq::set_error(SQ_ERR_ARG, "sq_build_node_records: null argument");
  const int N = d->n_nodes, L = d->n_links;
  std::vector<unsigned char> out_local((size_t)std::max(L, 1), 0);
  for (int n = 0; n < N; ++n)
    for (int j = d->node_out_ptr[n]; j < d->node_out_ptr[n + 1]; ++j) out_local[d->node_out_links[j]] = (unsigned char)(j - d->node_out_ptr[n]);
  std::vector<NodeRec> nrec;
  build_node_records(d, out_local, nrec);
  for (int n = 0; n < N; ++n)
    for (int i = 0; i < 8; ++i) { out[(size_t)n * 16 + i] = nrec[n].in[i]; out[(size_t)n * 16 + 8 + i] = (int32_t)nrec[n].sec[i]; }
  return SQ_OK;
}

int sq_set_shelter_after(SqSim* sim, int32_t seconds) {
  if (!sim) return sq::set_error(SQ_ERR_ARG, "sq_set_shelter_after: null sim");
  SQ_CUDA_TRY(cudaStreamSynchronize(sim->stream));
  sim->st.shelter_after = seconds > 0 ? seconds : 0;
  if (sim->graph) { cudaGraphExecDestroy(sim->graph); sim->graph = nullptr; }  // kernel parameters changed
  return SQ_OK;
}

int sq_attach_router(SqSim* sim, const SqRouterDesc* r) {
  if (!sim || !r || !r->graph) return sq::set_error(SQ_ERR_ARG, "sq_attach_router: null argument");
  if (sim->router_graph) return sq::set_error(SQ_ERR_ARG, "sq_attach_router: a router is already attached");
  const int N = sim->N, L = sim->L, A = sim->A, D = r->n_dest;
  if (D <= 0 || r->reroute_freq <= 0 || r->weight_freq <= 0)
    return sq::set_error(SQ_ERR_ARG, "sq_attach_router: n_dest, reroute_freq and weight_freq must be positive");
  Graph* g = (Graph*)r->graph;
  if (int rc = sp_sync(g)) return sq::set_error(rc, "sq_attach_router: %s", sp_last_error());
  const int V = sp_nvertices(g), E = r->n_graph_edges;
  for (int n = 0; n < N; ++n)
    if (r->node_vertex[n] < 0 || r->node_vertex[n] >= V)
      return sq::set_error(SQ_ERR_ARG, "sq_attach_router: node %d maps to vertex %d outside the graph (%d vertices)", n, r->node_vertex[n], V);
  for (int l = 0; l < L; ++l)
    if (r->link_edge[l] >= E) return sq::set_error(SQ_ERR_ARG, "sq_attach_router: link %d maps to edge %d of %d", l, r->link_edge[l], E);
  for (int a = 0; a < A; ++a)
    if (r->agent_dslot[a] < 0 || r->agent_dslot[a] >= D) return sq::set_error(SQ_ERR_ARG, "sq_attach_router: agent %d has destination slot %d", a, r->agent_dslot[a]);
  SQ_CUDA_TRY(cudaStreamSynchronize(sim->stream));
  TickState& s = sim->st;
  int rc;
  // epochs kept: a route is at most reroute_freq ticks old when it is renewed, so reroute_freq / weight_freq + 1 regular
  // refreshes can be live, + the slot being built, + 8 for refreshes off the weight grid (closure events, host edits of
  // the graph) inside one reroute period.  Beyond that the oldest slot is recycled early and k_evacuate writes its
  // followers' remaining routes out (correct, but it eats the materialisation pool).
  const int K = r->reroute_freq / r->weight_freq + 3 + 8;
  s.D = D; s.K = K; s.freq = r->reroute_freq;
  if ((rc = dev_upload_ptr(sim, &s.n_vertex, r->node_vertex, (size_t)N))) return rc;
  {  // destination slots go into the agents' records (one-time round trip)
    std::vector<AgentRec> rec((size_t)A);
    if (A > 0) SQ_CUDA_TRY(cudaMemcpy(rec.data(), s.ag, sizeof(AgentRec) * (size_t)A, cudaMemcpyDeviceToHost));
    for (int a = 0; a < A; ++a) { rec[a].dslot = r->agent_dslot[a]; rec[a].epoch = 0; }
    if (A > 0) SQ_CUDA_TRY(cudaMemcpy(s.ag, rec.data(), sizeof(AgentRec) * (size_t)A, cudaMemcpyHostToDevice));
  }
  if ((rc = dev_upload_ptr(sim, &s.l_edge, r->link_edge, (size_t)L))) return rc;
  {
    std::vector<int> one(1, 0);
    if ((rc = dev_upload_rw(sim, &s.cur_epoch, one))) return rc;
    std::vector<long long> zl(1, 0);
    if ((rc = dev_upload_rw(sim, &s.mat_used, zl))) return rc;
  }
  const size_t tbl = (size_t)K * D * N;
  SQ_CUDA_TRY(cudaMalloc(&s.next_tbl, sizeof(int) * tbl)); sim->allocs.push_back(s.next_tbl);
  SQ_CUDA_TRY(cudaMalloc(&s.dist_tbl, sizeof(double) * tbl)); sim->allocs.push_back(s.dist_tbl);
  SQ_CUDA_TRY(cudaMalloc(&s.w_edge, sizeof(double) * (size_t)std::max(E, 1))); sim->allocs.push_back(s.w_edge);
  SQ_CUDA_TRY(cudaMemset(s.w_edge, 0xff, sizeof(double) * (size_t)std::max(E, 1)));  // NaN = keep
  s.mat_cap = std::max<long long>(1 << 20, 16ll * A);
  SQ_CUDA_TRY(cudaMalloc(&s.mat_pool, sizeof(int) * (size_t)s.mat_cap)); sim->allocs.push_back(s.mat_pool);
  SQ_CUDA_TRY(cudaMalloc(&sim->d_tdist, sizeof(double) * (size_t)D * V)); sim->allocs.push_back(sim->d_tdist);
  SQ_CUDA_TRY(cudaMalloc(&sim->d_tnext, sizeof(int) * (size_t)D * V)); sim->allocs.push_back(sim->d_tnext);
  {
    std::vector<int> dv((size_t)D);
    for (int d = 0; d < D; ++d) {
      if (r->dest_nodes[d] < 0 || r->dest_nodes[d] >= N) return sq::set_error(SQ_ERR_ARG, "sq_attach_router: destination %d is node %d", d, r->dest_nodes[d]);
      dv[d] = r->node_vertex[r->dest_nodes[d]];
    }
    const int* p = nullptr;
    if ((rc = dev_upload(sim, &p, dv))) return rc;
    sim->d_dest_vertex = const_cast<int*>(p);
  }
  {  // (start node, end node) -> link table as a CSR over start nodes, end nodes as graph vertices
    std::vector<int> ptr((size_t)N + 1, 0), endv((size_t)std::max(r->n_pairs, 1)), link((size_t)std::max(r->n_pairs, 1));
    for (int k = 0; k < r->n_pairs; ++k) {
      if (r->pair_start[k] < 0 || r->pair_start[k] >= N || r->pair_end[k] < 0 || r->pair_end[k] >= N || r->pair_link[k] < 0 || r->pair_link[k] >= L)
        return sq::set_error(SQ_ERR_ARG, "sq_attach_router: pair %d is out of range", k);
      ++ptr[(size_t)r->pair_start[k] + 1];
    }
    for (int n = 0; n < N; ++n) ptr[(size_t)n + 1] += ptr[n];
    std::vector<int> fill(ptr.begin(), ptr.end() - 1);
    for (int k = 0; k < r->n_pairs; ++k) {
      const int pos = fill[r->pair_start[k]]++;
      endv[pos] = r->node_vertex[r->pair_end[k]];
      link[pos] = r->pair_link[k];
    }
    const int* p = nullptr;
    if ((rc = dev_upload(sim, &p, ptr))) return rc; sim->d_pair_ptr = const_cast<int*>(p);
    if ((rc = dev_upload(sim, &p, endv))) return rc; sim->d_pair_endv = const_cast<int*>(p);
    if ((rc = dev_upload(sim, &p, link))) return rc; sim->d_pair_link = const_cast<int*>(p);
  }
  {  // agents by (agent id mod reroute_freq): the reroute kernel only visits the bucket of the tick
    const int F = r->reroute_freq;
    std::vector<int> ptr((size_t)F + 1, 0), ag((size_t)std::max(A, 1));
    auto bucket = [&](int a) { long long m = r->agent_key[a] % F; return (int)(m < 0 ? m + F : m); };
    for (int a = 0; a < A; ++a) ++ptr[(size_t)bucket(a) + 1];
    sim->max_bucket = 0;
    for (int b = 0; b < F; ++b) { sim->max_bucket = std::max(sim->max_bucket, ptr[(size_t)b + 1]); ptr[(size_t)b + 1] += ptr[b]; }
    std::vector<int> fill(ptr.begin(), ptr.end() - 1);
    for (int a = 0; a < A; ++a) ag[fill[bucket(a)]++] = a;
    if ((rc = dev_upload(sim, &s.bucket_ptr, ptr))) return rc;
    if ((rc = dev_upload(sim, &s.bucket_agents, ag))) return rc;
  }
  sim->agent_key.assign(r->agent_key, r->agent_key + A);
  sim->router_graph = g;
  sim->wfreq = r->weight_freq;
  sim->graph_V = V; sim->graph_E = E;
  sim->host_epoch = K - 1;          // the first refresh fills slot 0
  sim->last_refresh_t = INT_MIN;
  if (sim->graph) { cudaGraphExecDestroy(sim->graph); sim->graph = nullptr; }
  // t = 0: weights from the empty network, tables, and every agent routed (run_traffic_simulation.py:107)
  if ((rc = router_refresh(sim, 0))) return rc;
  if (A > 0) {
    k_reroute_all<<<(A + 127) / 128, 128, 0, sim->stream>>>(s);
    sq::count_launch(); sim->launches += 1;
  }
  SQ_CUDA_TRY(cudaGetLastError());
  return sq_sync(sim);
}

int sq_set_reroute_rules(SqSim* sim, double detour_factor, int32_t closure_reroute_freq) {
  if (!sim) return sq::set_error(SQ_ERR_ARG, "sq_set_reroute_rules: null sim");
  if (!sim->router_graph) return sq::set_error(SQ_ERR_ARG, "sq_set_reroute_rules: attach a router first");
  if (detour_factor < 0.0 || closure_reroute_freq < 0) return sq::set_error(SQ_ERR_ARG, "sq_set_reroute_rules: negative argument");
  if (sim->st.a_rdist) return sq::set_error(SQ_ERR_ARG, "sq_set_reroute_rules: rules are already set");
  SQ_CUDA_TRY(cudaStreamSynchronize(sim->stream));
  TickState& s = sim->st;
  const int A = sim->A, L = sim->L, F = s.freq, CF = closure_reroute_freq;
  int rc;
  {
    // route_distance of the t = 0 routes: the distance table entry of each agent's origin
    std::vector<double> rd((size_t)std::max(A, 1), 5e7);
    if ((rc = dev_upload_rw(sim, &s.a_rdist, rd))) return rc;
    std::vector<int> zl((size_t)L, 0);
    if ((rc = dev_upload_rw(sim, &s.l_tomb, zl))) return rc;
  }
  {
    std::vector<int> rs((size_t)std::max(A, 1), 0), cs((size_t)std::max(A, 1), 0);
    auto slot = [&](int a, int m) { long long k = sim->agent_key[a] % m; return (int)(k < 0 ? k + m : k); };
    for (int a = 0; a < A; ++a) { rs[a] = slot(a, F); cs[a] = CF > 0 ? slot(a, CF) : 0; }
    if ((rc = dev_upload(sim, &s.a_rslot, rs))) return rc;
    if ((rc = dev_upload(sim, &s.a_cslot, cs))) return rc;
    if (CF > 0) {
      std::vector<int> ptr((size_t)CF + 1, 0), ag((size_t)std::max(A, 1));
      for (int a = 0; a < A; ++a) ++ptr[(size_t)cs[a] + 1];
      s.max_cbucket = 0;
      for (int b = 0; b < CF; ++b) { s.max_cbucket = std::max(s.max_cbucket, ptr[(size_t)b + 1]); ptr[(size_t)b + 1] += ptr[b]; }
      std::vector<int> fill(ptr.begin(), ptr.end() - 1);
      for (int a = 0; a < A; ++a) ag[fill[cs[a]]++] = a;
      if ((rc = dev_upload(sim, &s.cbucket_ptr, ptr))) return rc;
      if ((rc = dev_upload(sim, &s.cbucket_agents, ag))) return rc;
    }
  }
  s.detour_factor = detour_factor;
  s.cfreq = CF;
  if (sim->graph) { cudaGraphExecDestroy(sim->graph); sim->graph = nullptr; }   // kernel parameters and launches changed
  // the routes given at attach time: route_distance = distance from the origin (never a detour: previous = 5e7)
  if (A > 0) {
    k_reroute_all<<<(A + 127) / 128, 128, 0, sim->stream>>>(s);
    sq::count_launch(); sim->launches += 1;
  }
  SQ_CUDA_TRY(cudaGetLastError());
  return sq_sync(sim);
}

int sq_get_shelter_state(SqSim* sim, int32_t* link_shelter, int32_t* agent_stop_tick) {
  if (!sim) return sq::set_error(SQ_ERR_ARG, "null sim");
  if (link_shelter)
    SQ_CUDA_TRY(cudaMemcpyAsync(link_shelter, sim->st.l_shelter, sizeof(int) * (size_t)sim->L, cudaMemcpyDeviceToHost, sim->stream));
  if (agent_stop_tick)
    SQ_CUDA_TRY(cudaMemcpyAsync(agent_stop_tick, sim->st.a_arrive, sizeof(int) * (size_t)sim->A, cudaMemcpyDeviceToHost, sim->stream));
  return sq_sync(sim);
}

int sq_set_closure_schedule(SqSim* sim, int32_t n, const int32_t* tick, const int32_t* link_idx, const uint8_t* closed) {
  if (!sim || (n > 0 && (!tick || !link_idx || !closed))) return sq::set_error(SQ_ERR_ARG, "sq_set_closure_schedule: null argument");
  for (int i = 0; i < n; ++i)
    if (link_idx[i] < 0 || link_idx[i] >= sim->L) return sq::set_error(SQ_ERR_ARG, "sq_set_closure_schedule: link index %d out of range", link_idx[i]);
  SQ_CUDA_TRY(cudaStreamSynchronize(sim->stream));
  // sort by tick (stable); of several events for one link at one tick the last one given wins
  std::vector<int> order((size_t)std::max(n, 0));
  std::iota(order.begin(), order.end(), 0);
  std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return tick[a] < tick[b]; });
  sim->ev_tick.clear(); sim->ev_link.clear(); sim->ev_closed.clear();
  for (size_t k = 0; k < order.size(); ++k) {
    const int i = order[k];
    bool superseded = false;
    for (size_t m = k + 1; m < order.size() && tick[order[m]] == tick[i]; ++m)
      if (link_idx[order[m]] == link_idx[i]) { superseded = true; break; }
    if (superseded) continue;
    sim->ev_tick.push_back(tick[i]); sim->ev_link.push_back(link_idx[i]); sim->ev_closed.push_back(closed[i] ? 1 : 0);
  }
  sim->ev_cursor = 0;
  cudaFree(sim->d_ev_link); cudaFree(sim->d_ev_closed);
  sim->d_ev_link = nullptr; sim->d_ev_closed = nullptr;
  const size_t m = sim->ev_tick.size();
  if (m > 0) {
    SQ_CUDA_TRY(cudaMalloc(&sim->d_ev_link, sizeof(int) * m));
    SQ_CUDA_TRY(cudaMalloc(&sim->d_ev_closed, m));
    SQ_CUDA_TRY(cudaMemcpy(sim->d_ev_link, sim->ev_link.data(), sizeof(int) * m, cudaMemcpyHostToDevice));
    SQ_CUDA_TRY(cudaMemcpy(sim->d_ev_closed, sim->ev_closed.data(), m, cudaMemcpyHostToDevice));
  }
  return SQ_OK;
}

int sq_set_link_closed(SqSim* sim, int32_t n, const int32_t* link_idx, const uint8_t* closed) {
  if (!sim || (n > 0 && (!link_idx || !closed))) return sq::set_error(SQ_ERR_ARG, "sq_set_link_closed: null argument");
  if (n <= 0) return SQ_OK;
  for (int i = 0; i < n; ++i) {
    if (link_idx[i] < 0 || link_idx[i] >= sim->L) return sq::set_error(SQ_ERR_ARG, "link index %d out of range", link_idx[i]);
    sim->h_closed[link_idx[i]] = closed[i];
  }
  int* d_idx = nullptr; unsigned char* d_c = nullptr;
  SQ_CUDA_TRY(cudaMalloc(&d_idx, sizeof(int) * (size_t)n));
  SQ_CUDA_TRY(cudaMalloc(&d_c, (size_t)n));
  SQ_CUDA_TRY(cudaMemcpyAsync(d_idx, link_idx, sizeof(int) * (size_t)n, cudaMemcpyHostToDevice, sim->stream));
  SQ_CUDA_TRY(cudaMemcpyAsync(d_c, closed, (size_t)n, cudaMemcpyHostToDevice, sim->stream));
  k_set_closed<<<(n + 127) / 128, 128, 0, sim->stream>>>(sim->st, n, d_idx, d_c);
  sq::count_launch(); sim->launches += 1;
  SQ_CUDA_TRY(cudaStreamSynchronize(sim->stream));
  cudaFree(d_idx); cudaFree(d_c);
  return SQ_OK;
}

}  // extern "C"

template <typename T>
static int fetch(SqSim* sim, std::vector<T>& h, const T* d, size_t n) {
  h.resize(n);
  SQ_CUDA_TRY(cudaMemcpyAsync(h.data(), d, n * sizeof(T), cudaMemcpyDeviceToHost, sim->stream));
  return SQ_OK;
}

extern "C" {

int sq_get_link_state(SqSim* sim, int32_t* q, int32_t* r, int32_t* left, double* in_c, double* out_c, double* st) {
  if (!sim) return sq::set_error(SQ_ERR_ARG, "null sim");
  const size_t L = (size_t)sim->L;
  std::vector<unsigned> h, qt, tl;
  int rc;
  if ((rc = fetch(sim, h, sim->st.head, L)) || (rc = fetch(sim, qt, sim->st.qtail, L)) || (rc = fetch(sim, tl, sim->st.tail, L))) return rc;
  if (left) SQ_CUDA_TRY(cudaMemcpyAsync(left, sim->st.total_left, L * sizeof(int), cudaMemcpyDeviceToHost, sim->stream));
  if (in_c) SQ_CUDA_TRY(cudaMemcpyAsync(in_c, sim->st.in_c, L * sizeof(double), cudaMemcpyDeviceToHost, sim->stream));
  if (out_c) SQ_CUDA_TRY(cudaMemcpyAsync(out_c, sim->st.out_c, L * sizeof(double), cudaMemcpyDeviceToHost, sim->stream));
  if (st) SQ_CUDA_TRY(cudaMemcpyAsync(st, sim->st.storage_c, L * sizeof(double), cudaMemcpyDeviceToHost, sim->stream));
  if ((rc = sq_sync(sim))) return rc;
  for (size_t l = 0; l < L; ++l) {
    if (q) q[l] = (int)(qt[l] - h[l]);
    if (r) r[l] = (int)(tl[l] - qt[l]);
  }
  return SQ_OK;
}

int sq_get_link_vehicles(SqSim* sim, int64_t* ptr, int32_t* agents) {
  if (!sim || !ptr || !agents) return sq::set_error(SQ_ERR_ARG, "null argument");
  const size_t L = (size_t)sim->L;
  std::vector<unsigned> h, tl;
  std::vector<int2> ring;
  int rc;
  if ((rc = fetch(sim, h, sim->st.head, L)) || (rc = fetch(sim, tl, sim->st.tail, L)) ||
      (rc = fetch(sim, ring, (const int2*)sim->st.ring, (size_t)sim->ring_slots))) return rc;
  if ((rc = sq_sync(sim))) return rc;
  int64_t n = 0;
  for (size_t l = 0; l < L; ++l) {
    ptr[l] = n;
    for (unsigned i = h[l]; i != tl[l]; ++i) {
      if (n >= sim->A) return sq::set_error(SQ_ERR_STATE, "more vehicles on links than agents");
      agents[n++] = ring[(size_t)sim->h_ring_base[l] + (i & sim->h_ring_mask[l])].x;
    }
  }
  ptr[L] = n;
  return SQ_OK;
}

int sq_get_agent_state(SqSim* sim, uint8_t* status, int32_t* cur, int32_t* next, int32_t* enter, int32_t* arrive) {
  if (!sim) return sq::set_error(SQ_ERR_ARG, "null sim");
  const size_t A = (size_t)sim->A;
  std::vector<AgentRec> rec;
  int rc;
  if ((rc = fetch(sim, rec, (const AgentRec*)sim->st.ag, A))) return rc;
  if (arrive) SQ_CUDA_TRY(cudaMemcpyAsync(arrive, sim->st.a_arrive, A * sizeof(int), cudaMemcpyDeviceToHost, sim->stream));
  if ((rc = sq_sync(sim))) return rc;
  for (size_t a = 0; a < A; ++a) {
    const int4 r = rec[a].rec;
    if (status) status[a] = (uint8_t)(r.w & 0xff);
    if (cur) cur[a] = r.y;
    if (next) next[a] = r.x;
    if (enter) enter[a] = r.z;
    if (arrive && (r.w & 0xff) != SQ_ARRIVE) arrive[a] = -1;   // the slot also holds shelter ticks
  }
  return SQ_OK;
}

int sq_get_node_state(SqSim* sim, int64_t* node_move) {
  if (!sim || !node_move) return sq::set_error(SQ_ERR_ARG, "null argument");
  SQ_CUDA_TRY(cudaMemcpyAsync(node_move, sim->st.node_move, sizeof(long long) * (size_t)sim->N, cudaMemcpyDeviceToHost, sim->stream));
  return sq_sync(sim);
}

int sq_get_link_weights(SqSim* sim, double* w) {
  if (!sim || !w) return sq::set_error(SQ_ERR_ARG, "null argument");
  const size_t L = (size_t)sim->L;
  std::vector<unsigned> h, qt;
  int rc;
  if ((rc = fetch(sim, h, sim->st.head, L)) || (rc = fetch(sim, qt, sim->st.qtail, L))) return rc;
  if ((rc = sq_sync(sim))) return rc;
  for (size_t l = 0; l < L; ++l) {
    if (sim->h_closed[l] || sim->h_virtual[l]) { w[l] = NAN; continue; }
    const double q = (double)(int)(qt[l] - h[l]);
    w[l] = q / (sim->h_capacity[l] + 0.1) * 3600 * 25 / sim->h_maxmph[l] + sim->h_fft[l];
  }
  return SQ_OK;
}

int sq_get_counters(SqSim* sim, int64_t* out6) {
  if (!sim || !out6) return sq::set_error(SQ_ERR_ARG, "null argument");
  u64 c[12];
  SQ_CUDA_TRY(cudaMemcpyAsync(c, sim->st.counters, sizeof(c), cudaMemcpyDeviceToHost, sim->stream));
  if (int rc = sq_sync(sim)) return rc;
  sim->sheltered = (long long)c[7];
  out6[0] = (int64_t)c[0];
  // + vehicles on links after the last tick: loaded - arrived - taken off the road, as frozen by its node phase
  // (c[8]; c[7] may already include sq_stop_agents calls made since)
  out6[1] = (int64_t)(c[1] + (c[0] > 0 ? c[6] - c[4] - c[8] : 0));
  out6[2] = (int64_t)c[2]; out6[3] = (int64_t)c[3]; out6[4] = (int64_t)c[4];
  out6[5] = sim->launches;
  return SQ_OK;
}

const char* sq_last_error(void) { return sq::err_buf(); }
long long sq_kernel_launches(void) { return sq::launch_counter().load(); }

}  // extern "C"
