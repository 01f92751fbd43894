// CPU model of the grouped SSSP kernel's exact schedule (sssp_group_kernel): G trees share a FIFO ring; dirty marks per
// (vertex, tree) lane; pushes decided per batch of PB pops; one pending list rescanned at bucket transitions.
// usage: sim_group2 <dir> <n_grid> <G> <delta_factor> <max_groups> <PB> [ofile]
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
static uint32_t* hkey;
static int cmp_h(const void* a, const void* b) { uint32_t x = hkey[*(const int*)a], y = hkey[*(const int*)b]; return x < y ? -1 : x > y; }
int main(int argc, char** argv) {
  const char* dir = argv[1]; int n_grid = atoi(argv[2]); int G = atoi(argv[3]); double df = atof(argv[4]);
  int max_groups = atoi(argv[5]); int PB = atoi(argv[6]); const char* ofile = argc > 7 ? argv[7] : "o.bin";
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
  for (int v = 0; v < V; ++v) hkey[v] = hilbert(256, v / n_grid, v % n_grid);
  qsort(o, nO, 4, cmp_h);
  double* dist = malloc((size_t)V * G * 8);
  unsigned char* dirty = calloc((size_t)V * G, 1);
  int cap = 4 * V; int* ring = malloc(cap * 4), *P0 = malloc(cap * 4), *P1 = malloc(cap * 4);
  long long pops = 0, clean_pops = 0, relax_ops = 0, trans = 0, scanned = 0, maxring = 0, maxpend = 0, stores = 0;
  int ngroups = 0;
  // deferred push records of a batch
  int* bv = malloc(PB * 8 * 4); int* bcls = malloc(PB * 8 * 4);
  for (size_t g0 = 0; g0 + G <= nO && ngroups < max_groups; g0 += G, ++ngroups) {
    for (size_t i = 0; i < (size_t)V * G; ++i) { dist[i] = INFINITY; dirty[i] = 0; }
    int rh = 0, rt = 0, np = 0; int* P = P0; int* Pn = P1;
    for (int k = 0; k < G; ++k) { int sv = o[g0 + k]; int was = 0; for (int q = 0; q < G; ++q) was |= dirty[(size_t)sv * G + q];
      dist[(size_t)sv * G + k] = 0; dirty[(size_t)sv * G + k] = 1; if (!was) ring[rt++] = sv; }
    double T = delta;
    for (;;) {
      if (rh == rt) {
        if (np == 0) break;
        ++trans; rh = rt = 0;
        // scan pending: min dirty value per entry
        double T_new = T + delta; int nn = 0; double gmin = INFINITY;
        for (int pass = 0; pass < 2; ++pass) {
          nn = 0;
          for (int i = 0; i < np; ++i) { int v = P[i]; ++scanned; double m = INFINITY;
            for (int k = 0; k < G; ++k) if (dirty[(size_t)v * G + k] && dist[(size_t)v * G + k] < m) m = dist[(size_t)v * G + k];
            if (m == INFINITY) continue;       // cleaned by a pop since it was listed
            if (m < gmin) gmin = m;
            if (m < T_new) ring[rt++] = v; else Pn[nn++] = v; }
          if (rt > 0 || nn == 0) break;
          T_new = gmin + delta; rt = 0;        // nothing within reach: jump
        }
        T = T_new; { int* x = P; P = Pn; Pn = x; } np = nn;
        continue;
      }
      if (rt - rh > maxring) maxring = rt - rh;
      int nb = rt - rh < PB ? rt - rh : PB; int nrec = 0;
      for (int b = 0; b < nb; ++b) {
        int u = ring[rh++]; ++pops; int any = 0;
        for (int k = 0; k < G; ++k) { any |= dirty[(size_t)u * G + k]; dirty[(size_t)u * G + k] = 0; }
        if (!any) ++clean_pops;
        for (int idx = rp[u]; idx < rp[u + 1]; ++idx) {
          int v = av[idx]; double ww = aw[idx]; ++relax_ops;
          int was_dirty = 0, was_near = 0, imp = 0, now_near = 0;
          for (int k = 0; k < G; ++k) { size_t a = (size_t)v * G + k; if (dirty[a]) { was_dirty = 1; if (dist[a] < T) was_near = 1; } }
          for (int k = 0; k < G; ++k) {
            double du = dist[(size_t)u * G + k]; if (du == INFINITY) continue;
            double nd = du + ww; size_t a = (size_t)v * G + k;
            if (nd < dist[a]) { dist[a] = nd; dirty[a] = 1; imp = 1; if (nd < T) now_near = 1; }
          }
          stores += imp;
          if (now_near && !was_near) { bv[nrec] = v; bcls[nrec++] = 0; }
          else if (imp && !now_near && !was_dirty) { bv[nrec] = v; bcls[nrec++] = 1; }
        }
      }
      for (int i = 0; i < nrec; ++i) { if (bcls[i] == 0) ring[rt++] = bv[i]; else P[np++] = bv[i]; }
      if (np > maxpend) maxpend = np;
      if (rt >= cap || np >= cap) { printf("overflow\n"); return 1; }
    }
    if (ngroups < 4) {   // the result must be THE fixed point: d[src] = 0, d[v] = min over in-edges of fl(d[u] + w), no marks left
      long bad = 0;
      double* best = malloc((size_t)V * G * 8);
      for (size_t i = 0; i < (size_t)V * G; ++i) { best[i] = INFINITY; if (dirty[i]) ++bad; }
      for (int u = 0; u < V; ++u) for (int idx = rp[u]; idx < rp[u + 1]; ++idx) for (int k = 0; k < G; ++k) {
        double c = dist[(size_t)u * G + k] + aw[idx]; size_t a = (size_t)av[idx] * G + k; if (c < best[a]) best[a] = c; }
      for (int k = 0; k < G; ++k) best[(size_t)o[g0 + k] * G + k] = 0;
      for (size_t i = 0; i < (size_t)V * G; ++i) if (best[i] != dist[i]) ++bad;
      printf("group %d fixed-point check: %ld bad\n", ngroups, bad);
      free(best);
    }
  }
  double trees = (double)ngroups * G;
  printf("G=%d df=%.1f PB=%d groups=%d pops/group=%.0f (clean %.0f) pops/tree=%.0f relax_ops/group=%.0f stores/group=%.0f trans/group=%.0f scanned/group=%.0f maxring=%lld maxpend=%lld\n",
         G, df, PB, ngroups, pops / (double)ngroups, clean_pops / (double)ngroups, pops / trees, relax_ops / (double)ngroups, stores / (double)ngroups, trans / (double)ngroups,
         scanned / (double)ngroups, maxring, maxpend);
  return 0;
}
