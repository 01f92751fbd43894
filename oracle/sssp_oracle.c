/*
 * sssp_oracle.c -- CPU restatement of the reference's shortest-path routine.  TEST INFRASTRUCTURE ONLY:
 * imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg, never by the product path.
 *
 * Follows projects/bolinas_civic/sp/src/graph.cc:153-218 (Graph::dijkstra_priority_queue): binary heap of
 * (distance, vertex) ordered on distance only, strict '>' relaxation (:209), no stale-entry skip, early
 * exit when the destination is popped (:196), DBL_MAX / -1 for unreached vertices (:178-183), empty
 * result for bad ids (:161-163).  Graph construction follows graph.cc:4-28 and :138-150: nvertices =
 * max id + 1, undirected edges are stored (min,max) and only ever relaxed towards .second (:202), all
 * duplicate inserts stay in the adjacency.
 *
 * Parity pin: tests/test_sssp_oracle.py checks this file against the seven known answers of
 * sp/tests/graph_test.cc (:31,:42,:54,:63,:91,:102,:127) and, in this container, bit-for-bit against
 * oracle/_ref/liblsp.so (the reference's own sources compiled by oracle/Makefile) on random graphs.
 *
 * Predecessors: the reference's tie-break depends on heap pop order, so comparisons use the canonical
 * rule parent[v] = min{u : fl(dist[u] + w(u,v)) == dist[v]}; `parent_raw` keeps
 * what this heap produced.
 */
#include <float.h>
#include <stdlib.h>
#include <string.h>

typedef struct {
  int V, E;
  int* row_ptr; /* [V+1] */
  int* head;    /* [E]   */
  double* w;    /* [E]   */
  int* rrow_ptr;
  int* tail;
  double* rw;
} OGraph;

typedef struct {
  double d;
  int v;
} HeapItem;

static void heap_push(HeapItem** h, int* n, int* cap, HeapItem it) {
  if (*n == *cap) {
    *cap = *cap ? *cap * 2 : 1024;
    *h = (HeapItem*)realloc(*h, sizeof(HeapItem) * (size_t)(*cap));
  }
  int i = (*n)++;
  HeapItem* a = *h;
  while (i > 0) {
    int p = (i - 1) / 2;
    if (a[p].d > it.d) { a[i] = a[p]; i = p; } else break;
  }
  a[i] = it;
}

static HeapItem heap_pop(HeapItem* a, int* n) {
  HeapItem top = a[0];
  HeapItem last = a[--(*n)];
  int i = 0;
  for (;;) {
    int l = 2 * i + 1, r = l + 1, m;
    if (l >= *n) break;
    m = (r < *n && a[r].d < a[l].d) ? r : l;
    if (a[m].d < last.d) { a[i] = a[m]; i = m; } else break;
  }
  if (*n > 0) a[i] = last;
  return top;
}

/* Build from a COO list in insertion order.  `directed` = 0 applies the (min,max) storage rule. */
OGraph* oracle_graph_create(const int* starts, const int* ends, const double* wghs, int nedges, int max_vertex_id,
                            int directed) {
  OGraph* g = (OGraph*)calloc(1, sizeof(OGraph));
  int V = max_vertex_id + 1, i;
  g->V = V;
  g->row_ptr = (int*)calloc((size_t)V + 1, sizeof(int));
  g->rrow_ptr = (int*)calloc((size_t)V + 1, sizeof(int));
  int E = 0;
  for (i = 0; i < nedges; ++i) {
    int a = starts[i], b = ends[i];
    if (!directed && a > b) { int t = a; a = b; b = t; }
    if (a < 0 || a >= V || b < 0 || b >= V) continue;
    g->row_ptr[a + 1]++; g->rrow_ptr[b + 1]++; E++;
  }
  for (i = 0; i < V; ++i) { g->row_ptr[i + 1] += g->row_ptr[i]; g->rrow_ptr[i + 1] += g->rrow_ptr[i]; }
  g->E = E;
  g->head = (int*)malloc(sizeof(int) * (size_t)(E ? E : 1));
  g->tail = (int*)malloc(sizeof(int) * (size_t)(E ? E : 1));
  g->w = (double*)malloc(sizeof(double) * (size_t)(E ? E : 1));
  g->rw = (double*)malloc(sizeof(double) * (size_t)(E ? E : 1));
  int* fill = (int*)malloc(sizeof(int) * (size_t)(V ? V : 1));
  int* rfill = (int*)malloc(sizeof(int) * (size_t)(V ? V : 1));
  memcpy(fill, g->row_ptr, sizeof(int) * (size_t)V);
  memcpy(rfill, g->rrow_ptr, sizeof(int) * (size_t)V);
  for (i = 0; i < nedges; ++i) {
    int a = starts[i], b = ends[i];
    if (!directed && a > b) { int t = a; a = b; b = t; }
    if (a < 0 || a >= V || b < 0 || b >= V) continue;
    int s = fill[a]++, r = rfill[b]++;
    g->head[s] = b; g->w[s] = wghs[i];
    g->tail[r] = a; g->rw[r] = wghs[i];
  }
  free(fill); free(rfill);
  return g;
}

void oracle_graph_free(OGraph* g) {
  if (!g) return;
  free(g->row_ptr); free(g->rrow_ptr); free(g->head); free(g->tail); free(g->w); free(g->rw); free(g);
}

int oracle_graph_nvertices(const OGraph* g) { return g->V; }

/* One tree.  dist[V], parent_raw[V] (may be NULL).  Returns 0, or -1 for ids the reference rejects
 * (graph.cc:161-163), in which case dist is filled with DBL_MAX and parents with -1. */
int oracle_dijkstra(const OGraph* g, int source, int destination, double* dist, int* parent_raw) {
  int V = g->V, i;
  for (i = 0; i < V; ++i) dist[i] = DBL_MAX;
  if (parent_raw) for (i = 0; i < V; ++i) parent_raw[i] = -1;
  if (source < 0 || source >= V || ((destination < 0 || destination >= V) && destination != -1)) return -1;
  HeapItem* heap = NULL;
  int n = 0, cap = 0;
  HeapItem it = {0.0, source};
  heap_push(&heap, &n, &cap, it);
  dist[source] = 0.0;
  while (n > 0) {
    HeapItem top = heap_pop(heap, &n);
    int u = top.v, e;
    if (u == destination) break;
    for (e = g->row_ptr[u]; e < g->row_ptr[u + 1]; ++e) {
      int v = g->head[e];
      double du = dist[u] + g->w[e];
      if (dist[v] > du) {
        if (parent_raw) parent_raw[v] = u;
        dist[v] = du;
        HeapItem ni = {du, v};
        heap_push(&heap, &n, &cap, ni);
      }
    }
  }
  free(heap);
  return 0;
}

/* Canonical predecessors from a settled distance array (full trees only). */
void oracle_canonical_parents(const OGraph* g, int source, const double* dist, int* parent) {
  int V = g->V, v, e;
  for (v = 0; v < V; ++v) {
    int best = -1;
    if (v != source && dist[v] != DBL_MAX) {
      for (e = g->rrow_ptr[v]; e < g->rrow_ptr[v + 1]; ++e) {
        int u = g->tail[e];
        if (dist[u] == DBL_MAX) continue;
        volatile double cand = dist[u] + g->rw[e];
        if (cand == dist[v] && (best < 0 || u < best)) best = u;
      }
    }
    parent[v] = best;
  }
}

/* n full trees, outputs [n][V]; the bench's CPU leg times this. */
int oracle_sssp_batch(const OGraph* g, const int* origins, int n, double* dist_out, int* parent_out) {
  int V = g->V, i;
  double* dtmp = dist_out ? NULL : (double*)malloc(sizeof(double) * (size_t)V);
  for (i = 0; i < n; ++i) {
    double* d = dist_out ? dist_out + (size_t)i * V : dtmp;
    oracle_dijkstra(g, origins[i], -1, d, NULL);
    if (parent_out) oracle_canonical_parents(g, origins[i], d, parent_out + (size_t)i * V);
  }
  free(dtmp);
  return 0;
}
