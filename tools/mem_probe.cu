// mem_probe.cu -- what random 8-byte traffic costs on this GPU, in the shape the tree kernels produce it.
//
// Every warp owns a private region of `region_kb` KB (a tree's distance row) and runs STEPS dependent steps; a step is
// `ilp` independent groups per lane of: a random 8-byte load (ld.global.cg), then -- its address derived from the
// loaded value, like dist[u] -> edge -> dist[v] -- a second random operation on the same region:
//   mode 0: load only (one round trip per step)      mode 1: load -> load
//   mode 2: load -> atomicMin.u64 with return        mode 3: load -> RED.MIN (no return)
//   mode 4: atomicMin.u64 with return only
// `spread` lanes share a 64-byte row (spread = 1: every lane its own line, the warp kernel; spread = 8: eight lanes per
// row, the group layout).  Prints operations per second for a sweep of warps per SM and ILP.
// Build and run on the GPU box:  nvcc -O3 -gencode arch=compute_100a,code=sm_100a tools/mem_probe.cu -o /tmp/mem_probe && /tmp/mem_probe
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

typedef unsigned long long u64;

__device__ __forceinline__ u64 ld_cg(const u64* p) {
  u64 v;
  asm volatile("ld.global.cg.u64 %0, [%1];" : "=l"(v) : "l"(p));
  return v;
}
__device__ __forceinline__ unsigned mix(unsigned x) {
  x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
  return x;
}

template <int ILP>
__global__ void probe(u64* mem, size_t region_words, int steps, int mode, int spread, u64* sink) {
  const int lane = threadIdx.x & 31;
  const size_t gwarp = (size_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  u64* R = mem + gwarp * region_words;
  const unsigned rows = (unsigned)(region_words / 8);       // 64-byte rows
  unsigned s[ILP];
#pragma unroll
  for (int j = 0; j < ILP; ++j) s[j] = mix((unsigned)(gwarp * 131 + (lane / spread) * 977 + j * 7919 + 1));
  u64 acc = 0;
  for (int it = 0; it < steps; ++it) {
    u64 a[ILP], b[ILP];
#pragma unroll
    for (int j = 0; j < ILP; ++j) {
      const unsigned row = s[j] % rows;
      const u64* p = R + (size_t)row * 8 + (spread > 1 ? (lane % spread) : 0);
      a[j] = (mode == 4) ? 0ull : ld_cg(p);
    }
#pragma unroll
    for (int j = 0; j < ILP; ++j) {
      s[j] = mix(s[j] + (unsigned)a[j] + 0x9e3779b9u);
      const unsigned row = s[j] % rows;
      u64* q = R + (size_t)row * 8 + (spread > 1 ? (lane % spread) : 0);
      b[j] = 0;
      if (mode == 1) b[j] = ld_cg(q);
      else if (mode == 2 || mode == 4) b[j] = atomicMin(q, ((u64)s[j] << 20) | 0x4000000000000000ull);
      else if (mode == 3) atomicMin(q, ((u64)s[j] << 20) | 0x4000000000000000ull);
    }
#pragma unroll
    for (int j = 0; j < ILP; ++j) {
      s[j] = mix(s[j] ^ (unsigned)b[j]);
      acc += a[j] + b[j];
    }
  }
  if (acc == 0x1234567ull) sink[0] = acc;
}

int main(int argc, char** argv) {
  int dev = 0;
  cudaSetDevice(dev);
  cudaDeviceProp prop;
  cudaGetDeviceProperties(&prop, dev);
  const int n_sm = prop.multiProcessorCount;
  const size_t region_kb = argc > 1 ? (size_t)atoi(argv[1]) : 400;
  const int steps = argc > 2 ? atoi(argv[2]) : 2000;
  const size_t region_words = region_kb * 1024 / 8;
  const int max_wps = 32;
  u64* mem = nullptr;
  u64* sink = nullptr;
  const size_t total = (size_t)n_sm * max_wps * region_words;
  if (cudaMalloc(&mem, total * 8) != cudaSuccess || cudaMalloc(&sink, 8) != cudaSuccess) { printf("alloc failed\n"); return 1; }
  cudaMemset(mem, 0x7f, total * 8);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  const char* names[5] = {"load", "load->load", "load->atomicMin(ret)", "load->RED.MIN", "atomicMin(ret)"};
  printf("{\"gpu\": \"%s\", \"sms\": %d, \"region_kb\": %zu, \"steps\": %d}\n", prop.name, n_sm, region_kb, steps);
  for (int spread = 1; spread <= 8; spread *= 8)
    for (int mode = 0; mode < 5; ++mode)
      for (int wps = 4; wps <= max_wps; wps *= 2)
        for (int ilp = 1; ilp <= 8; ilp *= 2) {
          const int grid = n_sm * wps / 4;
          float best = 1e30f;
          for (int rep = 0; rep < 2; ++rep) {
            cudaEventRecord(e0);
            switch (ilp) {
              case 1: probe<1><<<grid, 128>>>(mem, region_words, steps, mode, spread, sink); break;
              case 2: probe<2><<<grid, 128>>>(mem, region_words, steps, mode, spread, sink); break;
              case 4: probe<4><<<grid, 128>>>(mem, region_words, steps, mode, spread, sink); break;
              default: probe<8><<<grid, 128>>>(mem, region_words, steps, mode, spread, sink); break;
            }
            cudaEventRecord(e1);
            cudaEventSynchronize(e1);
            float ms;
            cudaEventElapsedTime(&ms, e0, e1);
            if (ms < best) best = ms;
          }
          const double lane_ops = (double)grid * 128 * ilp * steps * ((mode == 0 || mode == 4) ? 1 : 2);
          const double cyc_per_step = best * 1e-3 * prop.clockRate * 1e3 / steps;
          printf("{\"spread\": %d, \"mode\": \"%s\", \"warps_per_sm\": %d, \"ilp\": %d, \"ms\": %.3f, \"G_lane_ops_per_s\": %.1f, "
                 "\"cycles_per_step\": %.0f}\n", spread, names[mode], wps, ilp, best, lane_ops / (best * 1e-3) / 1e9, cyc_per_step);
          fflush(stdout);
        }
  return 0;
}
