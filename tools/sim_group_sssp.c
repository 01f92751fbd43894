// Design-space simulation (CPU) for the grouped SSSP kernel: G trees of nearby origins share one queue and relax
// together; counts queue pops and edge relaxations per tree against the one-tree-per-queue schedule.
// usage: sim_group_sssp <dir with s.bin e.bin w.bin o.bin> <n_grid> <G> <delta_factor> [max_groups] [order: 0 hilbert,1 random,2 pseudo]
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

static void* rd(const char* dir, const char* name, size_t* n, size_t esz) {
  char p[512]; snprintf(p, sizeof p, "%s/%s", dir, name);
  FILE* f = fopen(p, "rb"); if (!f) { perror(p); exit(1); }
  fseek(f, 0, SEEK_END); long sz = ftell(f); fseek(f, 0, SEEK_SET);
  void* b = malloc(sz); if (fread(b, 1, sz, f) != (size_t)sz) exit(1); fclose(f); *n = sz / esz; return b;
}
static uint32_t hilbert(uint32_t n, uint32_t x, uint32_t y) {
  uint32_t d = 0;
  for (uint32_t s = n / 2; s > 0; s /= 2) {
    uint32_t rx = (x & s) > 0, ry = (y & s) > 0;
    d += s * s * ((3 * rx) ^ ry);
    if (ry == 0) { if (rx == 1) { x = s - 1 - x; y = s - 1 - y; } uint32_t t = x; x = y; y = t; }
  }
  return d;
}
static int n_grid; static uint32_t* hkey;
static int cmp_h(const void* a, const void* b) { uint32_t x = hkey[*(const int*)a], y = hkey[*(const int*)b]; return x < y ? -1 : x > y; }

int main(int argc, char** argv) {
  const char* dir = argv[1]; n_grid = atoi(argv[2]); int G = atoi(argv[3]); double df = atof(argv[4]);
  int max_groups = argc > 5 ? atoi(argv[5]) : 1 << 30; int order = argc > 6 ? atoi(argv[6]) : 0;
  const char* ofile = argc > 7 ? argv[7] : "o.bin";
  size_t E, nO, t;
  int* s = rd(dir, "s.bin", &E, 4); int* e = rd(dir, "e.bin", &t, 4); double* w = rd(dir, "w.bin", &t, 8);
  int* o = rd(dir, ofile, &nO, 4);
  int V = n_grid * n_grid;
  int* rp = calloc(V + 1, 4); for (size_t i = 0; i < E; ++i) rp[s[i] + 1]++;
  for (int v = 0; v < V; ++v) rp[v + 1] += rp[v];
  int* fill = malloc(V * 4); memcpy(fill, rp, V * 4);
  int* av = malloc(E * 4); double* aw = malloc(E * 8); double sum = 0;
  for (size_t i = 0; i < E; ++i) { int p = fill[s[i]]++; av[p] = e[i]; aw[p] = w[i]; sum += w[i]; }
  double delta = df * sum / E;
  hkey = malloc(V * 4);
  if (order == 2) {
    // pseudo-coordinates: hop distances from two landmarks found by BFS sweeps (no geometry needed)
    int* q = malloc(V * 4); int* da = malloc(V * 4); int* dc = malloc(V * 4); int* d0 = malloc(V * 4);
    #define BFS(src, d) do { for (int i = 0; i < V; ++i) d[i] = -1; int h = 0, tl = 0; q[tl++] = src; d[src] = 0; \
      while (h < tl) { int x = q[h++]; for (int i = rp[x]; i < rp[x + 1]; ++i) if (d[av[i]] < 0) { d[av[i]] = d[x] + 1; q[tl++] = av[i]; } } } while (0)
    BFS(0, d0); int a = 0; for (int i = 0; i < V; ++i) if (d0[i] > d0[a]) a = i;
    BFS(a, da); int b = a; for (int i = 0; i < V; ++i) if (da[i] > da[b]) b = i;
    BFS(b, d0);  // d0 = distance from b
    int c = 0; long best = -1; for (int i = 0; i < V; ++i) { long m = da[i] < d0[i] ? da[i] : d0[i]; if (m > best) { best = m; c = i; } }
    BFS(c, dc);
    int pmin = 1 << 30, qmin = 1 << 30, pmax = -(1 << 30), qmax = -(1 << 30);
    for (int i = 0; i < V; ++i) { int p = da[i] + dc[i], qq = da[i] - dc[i]; if (p < pmin) pmin = p; if (p > pmax) pmax = p; if (qq < qmin) qmin = qq; if (qq > qmax) qmax = qq; }
    printf("landmarks a=%d b=%d c=%d  p in [%d,%d] q in [%d,%d]\n", a, b, c, pmin, pmax, qmin, qmax);
    for (int i = 0; i < V; ++i) { int p = (da[i] + dc[i] - pmin) / 2, qq = (da[i] - dc[i] - qmin) / 2; hkey[i] = hilbert(1024, p, qq); }
  } else
  for (int v = 0; v < V; ++v) hkey[v] = order == 1 ? (uint32_t)rand() : hilbert(256, v / n_grid, v % n_grid);
  qsort(o, nO, 4, cmp_h);
  double* dist = malloc((size_t)V * G * 8);
  unsigned char* inring = calloc(V, 1), *inbkt = calloc(V, 1), *dirty = calloc(V, 1);
  int* ring = malloc(V * 4 * 8), *bkt = malloc(V * 4 * 8);
  long long pops = 0, relax_ops = 0, improved = 0, lanes_live = 0, skipped_clean = 0, trans = 0, maxring = 0, store_ops = 0;
  int ngroups = 0;
  for (size_t g0 = 0; g0 + G <= nO && ngroups < max_groups; g0 += G, ++ngroups) {
    for (size_t i = 0; i < (size_t)V * G; ++i) dist[i] = INFINITY;
    int rh = 0, rt = 0, bt = 0;
    for (int k = 0; k < G; ++k) { int sv = o[g0 + k]; dist[(size_t)sv * G + k] = 0; if (!inring[sv]) { inring[sv] = 1; dirty[sv] = 1; ring[rt++] = sv; } }
    double T = delta, T2 = 2 * delta;
    for (;;) {
      if (rh == rt) {
        if (bt == 0) break;
        rh = rt = 0; ++trans;
        for (int i = 0; i < bt; ++i) { int v = bkt[i]; inbkt[v] = 0; if (dirty[v] && !inring[v]) { inring[v] = 1; ring[rt++] = v; } else skipped_clean++; }
        bt = 0; T = T2; T2 = T + delta;
        continue;
      }
      if (rt - rh > maxring) maxring = rt - rh;
      int u = ring[rh++]; inring[u] = 0; dirty[u] = 0; ++pops;
      for (int idx = rp[u]; idx < rp[u + 1]; ++idx) {
        int v = av[idx]; double ww = aw[idx]; int cls = 3; ++relax_ops; int st = 0;
        for (int k = 0; k < G; ++k) {
          double du = dist[(size_t)u * G + k]; if (du == INFINITY) continue; ++lanes_live;
          double nd = du + ww;
          if (nd < dist[(size_t)v * G + k]) { dist[(size_t)v * G + k] = nd; ++improved; st = 1; int c = nd < T ? 0 : nd < T2 ? 1 : 2; if (c < cls) cls = c; }
        }
        store_ops += st;
        if (cls == 3) continue;
        dirty[v] = 1;
        if (cls == 0) { if (!inring[v]) { inring[v] = 1; ring[rt++] = v; } }
        else { if (!inbkt[v] && !inring[v]) { inbkt[v] = 1; bkt[bt++] = v; } }
      }
    }
  }
  double trees = (double)ngroups * G;
  printf("G=%d df=%.1f groups=%d  pops/group=%.0f pops/tree=%.0f  relax_ops/group=%.0f /tree=%.0f  lane-improved/tree=%.0f live-lanes/tree=%.0f stores/group=%.0f skipped_clean/group=%.0f trans/group=%.0f maxring=%lld\n",
         G, df, ngroups, pops / (double)ngroups, pops / trees, relax_ops / (double)ngroups, relax_ops / trees, improved / trees, lanes_live / trees,
         store_ops / (double)ngroups, skipped_clean / (double)ngroups, trans / (double)ngroups, maxring);
  return 0;
}
