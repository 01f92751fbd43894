/*
 * tick_oracle.c -- CPU restatement of the reference's one-second tick.  TEST INFRASTRUCTURE ONLY
 * (tests/, __graft_entry__.smoke(), bench.py's cpu_baseline / --impl reference legs).
 *
 * Restates, with plain per-link vehicle lists exactly like the Python lists they stand for:
 *   Agent.load_vehicle            projects/bolinas_civic/queue_class.py:587-596
 *   Link.run_link_model           :445-455
 *   Node.non_conflict_vehicles    :226-296
 *   Node.run_node_model           :298-394
 *   Agent.move_agent_to_next_link :610-615, find_next_link :598-608
 *   driver wrappers link_model / node_model   run_traffic_simulation.py:52-69
 * with random.shuffle / random.choices replaced by the counter-based contract of
 * spatial_queue_b200/rng.py.
 *
 * Parity pin: tests/test_tick_oracle.py replays tests/golden/tick_*.npz -- per-tick digests produced by the
 * UNMODIFIED reference (tests/golden/make_tick_golden.py) -- and requires every tick to match.
 *
 * Nothing here shares data structures with the CUDA implementation (which uses rings, ready ticks and
 * precomputed conflict tables); the two only share the input description (include/sq_tick.h).
 */
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "../include/sq_tick.h"

typedef struct { int* v; int n, cap; } IVec;
static void iv_push(IVec* a, int x) {
  if (a->n == a->cap) { a->cap = a->cap ? a->cap * 2 : 8; a->v = (int*)realloc(a->v, sizeof(int) * (size_t)a->cap); }
  a->v[a->n++] = x;
}
static void iv_remove_value(IVec* a, int x) { /* [v for v in list if v != x] */
  int i, j = 0;
  for (i = 0; i < a->n; ++i) if (a->v[i] != x) a->v[j++] = a->v[i];
  a->n = j;
}

typedef struct {
  int n_nodes, n_links, n_agents;
  uint64_t seed;
  /* copies of the static description */
  int *l_start, *l_end, *l_lanes, *l_rev; uint8_t *l_virtual, *l_closed;
  double *l_cap_s, *l_storage, *l_fft, *l_capacity, *l_maxmph;
  int64_t* n_key; uint8_t* n_type;
  int *in_ptr, *in_links, *in_prio, *out_ptr, *out_links; uint8_t* in_phase;
  int64_t* turn_ptr; double* turn;
  int *a_origin, *a_destin, *a_dept, *a_vlen;
  /* dynamic link state */
  IVec *queue, *run;
  double *in_c, *out_c, *storage_c;
  int* total_left;
  /* dynamic agent state */
  uint8_t* a_status; int *a_cur, *a_next, *a_enter, *a_arrive;
  IVec* a_route; int* a_pos;
  int* virt_of_node;
  int64_t* node_move;
  int64_t* pair_moves;   /* [turn table shape] moves per (in_link, out_link): node_move_link_pairs, model/queue_class.py:291-383 */
  int64_t counters[6];
  int shelter_after; /* run_traffic_simulation.py:103: seconds on one real link before status shelter_a1; 0 = rule off */
  int* l_shelter;
  int congest_after; /* run_traffic_simulation.py:101: seconds on one real link before the vehicle counts as congested; 0 = off */
  long long* l_congested;
  char err[256];
  int failed;
} OSim;

#define DUP(dst, src, n, T) do { (dst) = (T*)malloc(sizeof(T) * (size_t)((n) > 0 ? (n) : 1)); \
                                 if ((n) > 0) memcpy((dst), (src), sizeof(T) * (size_t)(n)); } while (0)

/* ---- RNG contract (spatial_queue_b200/rng.py) ---- */
static uint64_t fmix(uint64_t x) {
  x ^= x >> 30; x *= 0xBF58476D1CE4E5B9ull; x ^= x >> 27; x *= 0x94D049BB133111EBull; x ^= x >> 31; return x;
}
static double uniform(uint64_t seed, int64_t t, int64_t node, int64_t k) {
  uint64_t h = seed ^ 0x243F6A8885A308D3ull;
  h = fmix(h + (uint64_t)t * 0x9E3779B97F4A7C15ull);
  h = fmix(h + (uint64_t)node * 0xBF58476D1CE4E5B9ull);
  h = fmix(h + (uint64_t)k * 0x94D049BB133111EBull);
  return (double)(h >> 11) * (1.0 / 9007199254740992.0);
}
static void shim_shuffle(OSim* s, int t, int node, int slot, int* lst, int n) {
  if (n >= 2) {
    double u = uniform(s->seed, t, s->n_key[node], slot);
    int r = (int)floor(u * (double)n), i, x = lst[r];
    for (i = r; i > 0; --i) lst[i] = lst[i - 1];
    lst[0] = x;
  }
}

OSim* oracle_tick_create(const SqNetDesc* d) {
  OSim* s = (OSim*)calloc(1, sizeof(OSim));
  int L = d->n_links, N = d->n_nodes, A = d->n_agents, i;
  s->n_nodes = N; s->n_links = L; s->n_agents = A; s->seed = d->seed;
  DUP(s->l_start, d->link_start_node, L, int); DUP(s->l_end, d->link_end_node, L, int);
  DUP(s->l_lanes, d->link_lanes_floor, L, int); DUP(s->l_rev, d->link_reverse, L, int);
  DUP(s->l_virtual, d->link_is_virtual, L, uint8_t); DUP(s->l_closed, d->link_closed, L, uint8_t);
  DUP(s->l_cap_s, d->link_cap_per_s, L, double); DUP(s->l_storage, d->link_storage_cap, L, double);
  DUP(s->l_fft, d->link_fft, L, double); DUP(s->l_capacity, d->link_capacity, L, double);
  DUP(s->l_maxmph, d->link_maxmph, L, double);
  DUP(s->n_key, d->node_key, N, int64_t); DUP(s->n_type, d->node_type, N, uint8_t);
  DUP(s->in_ptr, d->node_in_ptr, N + 1, int); DUP(s->out_ptr, d->node_out_ptr, N + 1, int);
  DUP(s->in_links, d->node_in_links, s->in_ptr[N], int); DUP(s->in_phase, d->node_in_phase, s->in_ptr[N], uint8_t);
  DUP(s->in_prio, d->node_in_prio, s->in_ptr[N], int); DUP(s->out_links, d->node_out_links, s->out_ptr[N], int);
  DUP(s->turn_ptr, d->node_turn_ptr, N + 1, int64_t); DUP(s->turn, d->turn_angle, s->turn_ptr[N], double);
  DUP(s->a_origin, d->agent_origin, A, int); DUP(s->a_destin, d->agent_destin, A, int);
  DUP(s->a_dept, d->agent_dept, A, int); DUP(s->a_vlen, d->agent_vlen, A, int);
  DUP(s->a_status, d->agent_status, A, uint8_t);
  s->queue = (IVec*)calloc((size_t)(L ? L : 1), sizeof(IVec)); s->run = (IVec*)calloc((size_t)(L ? L : 1), sizeof(IVec));
  s->in_c = (double*)malloc(sizeof(double) * (size_t)(L ? L : 1));
  s->out_c = (double*)malloc(sizeof(double) * (size_t)(L ? L : 1));
  s->storage_c = (double*)malloc(sizeof(double) * (size_t)(L ? L : 1));
  s->total_left = (int*)calloc((size_t)(L ? L : 1), sizeof(int));
  s->l_shelter = (int*)calloc((size_t)(L ? L : 1), sizeof(int));
  s->l_congested = (long long*)calloc((size_t)(L ? L : 1), sizeof(long long));
  for (i = 0; i < L; ++i) { /* Link.__init__ :417-419 */
    s->in_c[i] = s->out_c[i] = s->l_cap_s[i];
    s->storage_c[i] = s->l_storage[i];
  }
  s->virt_of_node = (int*)malloc(sizeof(int) * (size_t)(N ? N : 1));
  for (i = 0; i < N; ++i) s->virt_of_node[i] = -1;
  for (i = 0; i < L; ++i) if (s->l_virtual[i]) s->virt_of_node[s->l_end[i]] = i;
  s->a_cur = (int*)malloc(sizeof(int) * (size_t)(A ? A : 1)); s->a_next = (int*)malloc(sizeof(int) * (size_t)(A ? A : 1));
  s->a_enter = (int*)malloc(sizeof(int) * (size_t)(A ? A : 1)); s->a_arrive = (int*)malloc(sizeof(int) * (size_t)(A ? A : 1));
  s->a_route = (IVec*)calloc((size_t)(A ? A : 1), sizeof(IVec)); s->a_pos = (int*)calloc((size_t)(A ? A : 1), sizeof(int));
  for (i = 0; i < A; ++i) {
    int64_t k;
    for (k = d->route_ptr[i]; k < d->route_ptr[i + 1]; ++k) iv_push(&s->a_route[i], d->route_links[k]);
    s->a_cur[i] = s->virt_of_node[s->a_origin[i]];           /* find_next_link at t=0: this_link = virtual link */
    s->a_next[i] = s->a_route[i].n > 0 ? s->a_route[i].v[0] : -1;
    s->a_enter[i] = -1; s->a_arrive[i] = -1;
  }
  s->node_move = (int64_t*)calloc((size_t)(N ? N : 1), sizeof(int64_t));
  s->pair_moves = (int64_t*)calloc((size_t)(s->turn_ptr[N] ? s->turn_ptr[N] : 1), sizeof(int64_t));
  return s;
}

void oracle_tick_destroy(OSim* s) {
  int i;
  if (!s) return;
  for (i = 0; i < s->n_links; ++i) { free(s->queue[i].v); free(s->run[i].v); }
  for (i = 0; i < s->n_agents; ++i) free(s->a_route[i].v);
  free(s->l_start); free(s->l_end); free(s->l_lanes); free(s->l_rev); free(s->l_virtual); free(s->l_closed);
  free(s->l_cap_s); free(s->l_storage); free(s->l_fft); free(s->l_capacity); free(s->l_maxmph);
  free(s->n_key); free(s->n_type); free(s->in_ptr); free(s->in_links); free(s->in_prio); free(s->out_ptr);
  free(s->out_links); free(s->in_phase); free(s->turn_ptr); free(s->turn);
  free(s->a_origin); free(s->a_destin); free(s->a_dept); free(s->a_vlen);
  free(s->queue); free(s->run); free(s->in_c); free(s->out_c); free(s->storage_c); free(s->total_left);
  free(s->a_status); free(s->a_cur); free(s->a_next); free(s->a_enter); free(s->a_arrive); free(s->a_route);
  free(s->a_pos); free(s->virt_of_node); free(s->node_move); free(s->pair_moves); free(s->l_shelter); free(s->l_congested); free(s);
}

static int fail(OSim* s, const char* msg, int a, int b, int c) {
  snprintf(s->err, sizeof(s->err), "%s (agent %d, node %d, link %d)", msg, a, b, c);
  s->failed = 1;
  return -4;
}

/* node.incoming_links[in][out]; returns 0 and sets *ok = 0 on a KeyError */
static double turn_angle(const OSim* s, int node, int in_link, int out_link, int* ok) {
  int ib = s->in_ptr[node], ie = s->in_ptr[node + 1], ob = s->out_ptr[node], oe = s->out_ptr[node + 1], i, j;
  for (i = ib; i < ie; ++i) if (s->in_links[i] == in_link) break;
  for (j = ob; j < oe; ++j) if (s->out_links[j] == out_link) break;
  if (i == ie || j == oe) { *ok = 0; return 0.0; }
  *ok = 1;
  return s->turn[s->turn_ptr[node] + (int64_t)(i - ib) * (oe - ob) + (j - ob)];
}

typedef struct { int agent, il, ol; } GoVeh;

/* queue_class.py:226-296 */
static int non_conflict_vehicles(OSim* s, int t, int node, GoVeh* go, int* n_go) {
  int ib = s->in_ptr[node], ie = s->in_ptr[node + 1];
  int cand[256], nc = 0, i, ok, lane;
  int left_turn_phase = 0;
  *n_go = 0;
  for (i = ib; i < ie && nc < 256; ++i) if (s->queue[s->in_links[i]].n > 0) cand[nc++] = s->in_links[i];   /* :230 */
  shim_shuffle(s, t, node, 0, cand, nc);                                                                   /* :231 */
  if (s->n_type[node] == 1) {                                                                              /* :233-236 */
    int want = (t % 60 < 30) ? 1 : 2, m = 0;
    for (i = 0; i < nc; ++i) {
      int k;
      for (k = ib; k < ie; ++k) if (s->in_links[k] == cand[i]) break;
      if (s->in_phase[k] & want) cand[m++] = cand[i];
    }
    nc = m;
    shim_shuffle(s, t, node, 1, cand, nc);
  } else if (s->n_type[node] == 2) {                                                                       /* :237-238 */
    int ordered[256], m = 0, rank, maxrank = -1, k;
    for (k = ib; k < ie; ++k) if (s->in_prio[k] > maxrank) maxrank = s->in_prio[k];
    for (rank = 0; rank <= maxrank; ++rank)
      for (k = ib; k < ie; ++k)
        if (s->in_prio[k] == rank) {
          for (i = 0; i < nc; ++i) if (cand[i] == s->in_links[k]) { ordered[m++] = cand[i]; break; }
        }
    nc = m;
    memcpy(cand, ordered, sizeof(int) * (size_t)nc);
  }
  if (nc == 0) return 0;                                                                                   /* :241 */
  {
    int go_link = cand[0];                                                                                 /* :244 */
    int lanes = s->l_lanes[go_link], vehs = s->queue[go_link].n;
    int take = lanes < vehs ? lanes : vehs;
    for (lane = 0; lane < take; ++lane) {                                                                  /* :248-263 */
      int agent = s->queue[go_link].v[lane], ol = s->a_next[agent];
      go[*n_go].agent = agent; go[*n_go].il = go_link; go[*n_go].ol = ol; ++*n_go;
      if (ol >= 0) {
        double ang = turn_angle(s, node, go_link, ol, &ok);
        if (!ok) return fail(s, "turn not in the node's table (queue_class.py:257-262)", agent, node, go_link);
        if (ang >= 50) left_turn_phase = 1;
      }
    }
    if (left_turn_phase) return 0;                                                                         /* :265 */
    { int m = 0; for (i = 0; i < nc; ++i) if (cand[i] != go_link) cand[m++] = cand[i]; nc = m; }           /* :267 */
    if (nc == 0) return 0;
    {
      int rev = s->l_rev[go_link], best = -1;                                                              /* :272 */
      double best_abs = 0.0;
      if (rev < 0) return 0;                                                                               /* :273-275 */
      for (i = 0; i < nc; ++i) {                                                                           /* :277 */
        double a = fabs(turn_angle(s, node, cand[i], rev, &ok));
        if (!ok) return fail(s, "reverse link missing from the turn table", -1, node, cand[i]);
        if (best < 0 || a < best_abs) { best = cand[i]; best_abs = a; }
      }
      lanes = s->l_lanes[best]; vehs = s->queue[best].n; take = lanes < vehs ? lanes : vehs;
      for (lane = 0; lane < take; ++lane) {                                                                /* :281-293 */
        int agent = s->queue[best].v[lane], ol = s->a_next[agent];
        if (ol >= 0) {
          double ang = turn_angle(s, node, best, ol, &ok);
          if (!ok) return fail(s, "turn not in the node's table (queue_class.py:291)", agent, node, best);
          if (ang >= 50) continue;
        }
        go[*n_go].agent = agent; go[*n_go].il = best; go[*n_go].ol = ol; ++*n_go;
      }
    }
  }
  return 0;
}

static double max0(double x) { return x > 0 ? x : 0; } /* max(0, x) */

/* node_move_link_pairs.append((in_link, out_link)) of model/queue_class.py:327, :352, counted per pair */
static void count_pair(OSim* s, int node, int il, int ol) {
  int ib = s->in_ptr[node], ie = s->in_ptr[node + 1], ob = s->out_ptr[node], oe = s->out_ptr[node + 1], i, j;
  for (i = ib; i < ie && s->in_links[i] != il; ++i) {}
  for (j = ob; j < oe && s->out_links[j] != ol; ++j) {}
  if (i < ie && j < oe) s->pair_moves[s->turn_ptr[node] + (int64_t)(i - ib) * (oe - ob) + (j - ob)] += 1;
}

static void move_agent(OSim* s, int agent, int t) { /* :610-615 + :598-608 */
  s->a_enter[agent] = t;
  s->a_cur[agent] = s->a_next[agent];
  s->a_pos[agent] += 1;
  s->a_next[agent] = s->a_pos[agent] < s->a_route[agent].n ? s->a_route[agent].v[s->a_pos[agent]] : -1;
}

/* queue_class.py:298-394 */
static int run_node_model(OSim* s, int t, int node) {
  static GoVeh go[4096];
  int n_go = 0, g, node_move = 0, toss = 0, rc;
  rc = non_conflict_vehicles(s, t, node, go, &n_go);
  if (rc) return rc;
  for (g = 0; g < n_go; ++g) {
    int agent = go[g].agent, il = go[g].il, ol = go[g].ol;
    if (ol < 0 && node != s->a_destin[agent])
      return fail(s, "queued vehicle has no next link away from its destination (queue_class.py:312-314)", agent, node, il);
    if (node == s->a_destin[agent]) {                                                      /* :321-328 arrival */
      node_move += 1;
      s->total_left[il] += 1;
      iv_remove_value(&s->queue[il], agent);
      s->out_c[il] = max0(s->out_c[il] - 1);
      s->a_status[agent] = SQ_ARRIVE; s->a_arrive[agent] = t;
      s->counters[4] += 1;
    } else if (s->storage_c[ol] < s->a_vlen[agent]) {                                      /* :333 spill-back */
    } else if (s->out_c[il] >= 1 && s->in_c[ol] >= 1) {                                    /* :336-349 */
      node_move += 1;
      s->total_left[il] += 1;
      iv_remove_value(&s->queue[il], agent);
      s->out_c[il] = max0(s->out_c[il] - 1);
      iv_push(&s->run[ol], agent);
      s->in_c[ol] = max0(s->in_c[ol] - 1);
      s->storage_c[ol] -= s->a_vlen[agent];
      count_pair(s, node, il, ol);
      move_agent(s, agent, t);
    } else {                                                                               /* :358-386 */
      double cc = s->out_c[il] < s->in_c[ol] ? s->out_c[il] : s->in_c[ol]; /* min(out, in): first minimal */
      double w0 = 1 - cc, total = w0 + cc;
      double u = uniform(s->seed, t, s->n_key[node], 2 + toss);
      toss += 1;
      if (!(u * total < w0)) {
        node_move += 1;
        s->total_left[il] += 1;
        iv_remove_value(&s->queue[il], agent);
        s->out_c[il] = max0(s->out_c[il] - 1);
        iv_push(&s->run[ol], agent);
        s->in_c[ol] = max0(s->in_c[ol] - 1);
        s->storage_c[ol] -= s->a_vlen[agent];
        count_pair(s, node, il, ol);
        move_agent(s, agent, t);
      } else {
        if (s->out_c[il] < s->in_c[ol]) s->out_c[il] = max0(s->out_c[il] - 1);
        else if (s->out_c[il] > s->in_c[ol]) s->in_c[ol] = max0(s->in_c[ol] - 1);
        else { s->out_c[il] = max0(s->out_c[il] - 1); s->in_c[ol] = max0(s->in_c[ol] - 1); }
      }
    }
  }
  s->node_move[node] += node_move;                                                         /* :388 */
  s->counters[2] += node_move;
  return 0;
}

int oracle_tick_step(OSim* s, int t) {
  int L = s->n_links, A = s->n_agents, N = s->n_nodes, i, k, rc;
  if (s->failed) return -4;
  /* agent loop, give-up rule (run_traffic_simulation.py:100-104, 113-118): an enroute agent that has been on a
   * real link for more than shelter_after seconds becomes shelter_a1 and is stripped from the link's lists */
  /* agent loop, congestion counter (run_traffic_simulation.py:80-82, 100-102): link.congested is reset every tick and
   * then counts the vehicles that have been on the (real) link for more than congest_after seconds, before the
   * give-up test */
  if (s->congest_after > 0) {
    memset(s->l_congested, 0, sizeof(long long) * (size_t)(L ? L : 1));
    for (i = 0; i < A; ++i) {
      int cl;
      if (s->a_status[i] != SQ_ENROUTE) continue;
      cl = s->a_cur[i];
      if (!s->l_virtual[cl] && t - s->a_enter[i] > s->congest_after) s->l_congested[cl] += 1;
    }
  }
  if (s->shelter_after > 0)
    for (i = 0; i < A; ++i) {
      int cl;
      if (s->a_status[i] != SQ_ENROUTE) continue;
      cl = s->a_cur[i];
      if (s->l_virtual[cl] || t - s->a_enter[i] <= s->shelter_after) continue;
      s->a_status[i] = 4; /* shelter_a1 */
      s->a_arrive[i] = t;
      iv_remove_value(&s->queue[cl], i);
      iv_remove_value(&s->run[cl], i);
      s->l_shelter[cl] += 1;
    }
  /* agent loop: load_vehicle for every agent in dict order (run_traffic_simulation.py:111) */
  for (i = 0; i < A; ++i)
    if (s->a_status[i] == SQ_UNLOADED && s->a_dept[i] == t) {
      iv_push(&s->run[s->a_cur[i]], i);
      s->a_status[i] = SQ_ENROUTE;
      s->a_enter[i] = t;
      if (s->a_route[i].n == 0) return fail(s, "loaded agent has an empty route (queue_class.py:595 KeyError)", i, -1, -1);
      s->a_next[i] = s->a_route[i].v[0];
    }
  /* link model (run_traffic_simulation.py:55-56 -> queue_class.py:445-455) */
  for (i = 0; i < L; ++i) {
    IVec* run = &s->run[i];
    IVec* q = &s->queue[i];
    int q0 = q->n, j = 0;
    long sum = 0;
    for (k = 0; k < run->n; ++k)
      if ((double)s->a_enter[run->v[k]] < (double)t - s->l_fft[i]) iv_push(q, run->v[k]);
    for (k = 0; k < run->n; ++k) { /* run = [a for a in run if a not in queue] */
      int a = run->v[k], in_q = 0, m;
      for (m = q0; m < q->n; ++m) if (q->v[m] == a) { in_q = 1; break; }
      if (!in_q) for (m = 0; m < q0; ++m) if (q->v[m] == a) { in_q = 1; break; }
      if (!in_q) run->v[j++] = a;
    }
    run->n = j;
    for (k = 0; k < run->n; ++k) sum += s->a_vlen[run->v[k]];
    for (k = 0; k < q->n; ++k) sum += s->a_vlen[q->v[k]];
    s->storage_c[i] = s->l_storage[i] - (double)sum;
    s->in_c[i] = s->l_cap_s[i]; s->out_c[i] = s->l_cap_s[i];
    if (s->l_closed[i]) s->in_c[i] = 0;
  }
  /* node model for every node with a queued in-link (run_traffic_simulation.py:62-67) */
  for (i = 0; i < N; ++i) {
    int active = 0;
    for (k = s->in_ptr[i]; k < s->in_ptr[i + 1]; ++k) if (s->queue[s->in_links[k]].n > 0) { active = 1; break; }
    if (!active) continue;
    s->counters[3] += 1;
    rc = run_node_model(s, t, i);
    if (rc) return rc;
  }
  s->counters[0] += 1;
  for (i = 0; i < L; ++i) s->counters[1] += s->queue[i].n + s->run[i].n;
  return 0;
}

/* A driver rule stops agents between ticks (model/dta_meso_butte.py:199-263, run_traffic_simulation.py:113-118):
 * status, stop tick, and [v for v in list if v != agent] on both lists of the agent's link. */
int oracle_tick_stop_agents(OSim* s, int n, const int* agent_idx, int status, int t) {
  int k;
  if (status < 3) return fail(s, "stop_agents: not a stopped status", status, -1, -1);
  for (k = 0; k < n; ++k) {
    int i = agent_idx[k], cl;
    if (i < 0 || i >= s->n_agents) return fail(s, "stop_agents: agent index out of range", i, -1, -1);
    if (s->a_status[i] != SQ_ENROUTE) return fail(s, "stop_agents: agent is not on a link", i, -1, -1);
    cl = s->a_cur[i];
    s->a_status[i] = (uint8_t)status;
    s->a_arrive[i] = t;
    iv_remove_value(&s->queue[cl], i);
    iv_remove_value(&s->run[cl], i);
    s->l_shelter[cl] += 1;
  }
  return 0;
}

int oracle_tick_set_congest_after(OSim* s, int seconds) { s->congest_after = seconds > 0 ? seconds : 0; return 0; }
int oracle_tick_get_link_congested(OSim* s, long long* out) {
  memcpy(out, s->l_congested, sizeof(long long) * (size_t)s->n_links);
  return 0;
}
int oracle_tick_set_shelter_after(OSim* s, int seconds) { s->shelter_after = seconds > 0 ? seconds : 0; return 0; }
int oracle_tick_get_shelter_state(OSim* s, int* link_shelter, int* stop_tick) {
  if (link_shelter) memcpy(link_shelter, s->l_shelter, sizeof(int) * (size_t)s->n_links);
  if (stop_tick) memcpy(stop_tick, s->a_arrive, sizeof(int) * (size_t)s->n_agents);
  return 0;
}

int oracle_tick_run(OSim* s, int t0, int n) {
  int t, rc;
  for (t = t0; t < t0 + n; ++t) if ((rc = oracle_tick_step(s, t))) return rc;
  return 0;
}

int oracle_tick_set_routes(OSim* s, int n, const int* agent_idx, const int64_t* ptr, const int* links) {
  int i;
  int64_t k;
  for (i = 0; i < n; ++i) {
    int a = agent_idx[i];
    if (a < 0 || a >= s->n_agents) return -2;
    if (s->a_status[a] != SQ_UNLOADED && s->a_status[a] != SQ_ENROUTE) continue;
    s->a_route[a].n = 0;
    for (k = ptr[i]; k < ptr[i + 1]; ++k) iv_push(&s->a_route[a], links[k]);
    s->a_pos[a] = 0;
    s->a_next[a] = s->a_route[a].n > 0 ? s->a_route[a].v[0] : -1;
  }
  return 0;
}

int oracle_tick_set_link_closed(OSim* s, int n, const int* link_idx, const uint8_t* closed) {
  int i;
  for (i = 0; i < n; ++i) s->l_closed[link_idx[i]] = closed[i];
  return 0;
}

int oracle_tick_get_link_state(OSim* s, int* q, int* r, int* left, double* in_c, double* out_c, double* st) {
  int i;
  for (i = 0; i < s->n_links; ++i) {
    if (q) q[i] = s->queue[i].n;
    if (r) r[i] = s->run[i].n;
    if (left) left[i] = s->total_left[i];
    if (in_c) in_c[i] = s->in_c[i];
    if (out_c) out_c[i] = s->out_c[i];
    if (st) st[i] = s->storage_c[i];
  }
  return 0;
}

int oracle_tick_get_link_vehicles(OSim* s, int64_t* ptr, int* agents) {
  int i, k;
  int64_t n = 0;
  for (i = 0; i < s->n_links; ++i) {
    ptr[i] = n;
    for (k = 0; k < s->queue[i].n; ++k) agents[n++] = s->queue[i].v[k];
    for (k = 0; k < s->run[i].n; ++k) agents[n++] = s->run[i].v[k];
  }
  ptr[s->n_links] = n;
  return 0;
}

int oracle_tick_get_agent_state(OSim* s, uint8_t* status, int* cur, int* next, int* enter, int* arrive) {
  int i;
  for (i = 0; i < s->n_agents; ++i) {
    if (status) status[i] = s->a_status[i];
    if (cur) cur[i] = s->a_cur[i];
    if (next) next[i] = s->a_next[i];
    if (enter) enter[i] = s->a_enter[i];
    if (arrive) arrive[i] = s->a_status[i] == SQ_ARRIVE ? s->a_arrive[i] : -1;
  }
  return 0;
}

int oracle_tick_get_node_state(OSim* s, int64_t* node_move) {
  memcpy(node_move, s->node_move, sizeof(int64_t) * (size_t)s->n_nodes);
  return 0;
}

/* queue_class.py:463-466 */
int oracle_tick_get_link_weights(OSim* s, double* w) {
  int i;
  for (i = 0; i < s->n_links; ++i) {
    if (s->l_closed[i] || s->l_virtual[i]) { w[i] = NAN; continue; }
    w[i] = (double)s->queue[i].n / (s->l_capacity[i] + 0.1) * 3600 * 25 / s->l_maxmph[i] + s->l_fft[i];
  }
  return 0;
}

int oracle_tick_get_pair_moves(OSim* s, int64_t* out) {
  memcpy(out, s->pair_moves, sizeof(int64_t) * (size_t)s->turn_ptr[s->n_nodes]);
  return 0;
}
int oracle_tick_get_counters(OSim* s, int64_t* out6) { memcpy(out6, s->counters, sizeof(s->counters)); return 0; }
const char* oracle_tick_error(OSim* s) { return s->err; }
