// sq_common.cuh -- shared helpers for the sm_100a kernels (error channel, launch counter, PTX loads).
#pragma once
#include <cuda_runtime.h>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <atomic>

namespace sq {

// ---- per-thread error text (C ABI functions return codes and leave the text here) -------------
inline char* err_buf() {
  static thread_local char buf[512] = {0};
  return buf;
}
inline int set_error(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(err_buf(), 512, fmt, ap);
  va_end(ap);
  return code;
}
inline void clear_error() { err_buf()[0] = 0; }

// ---- kernel launch counter (bench.py reports it as gpu_launches) ------------------------------
inline std::atomic<long long>& launch_counter() {
  static std::atomic<long long> c{0};
  return c;
}
inline void count_launch(long long n = 1) { launch_counter().fetch_add(n, std::memory_order_relaxed); }

enum : int {
  SQ_OK = 0,
  SQ_ERR_CUDA = -1,      // CUDA runtime failure (text in sp_last_error / sq_last_error)
  SQ_ERR_ARG = -2,       // bad argument
  SQ_ERR_NODEVICE = -3,  // no usable GPU: there is no CPU fallback
  SQ_ERR_STATE = -4,     // simulation state inconsistent (e.g. vehicle with no next link)
  SQ_ERR_LIMIT = -5      // a documented capacity limit was exceeded
};

#define SQ_CUDA_TRY(expr)                                                                      \
  do {                                                                                         \
    cudaError_t _e = (expr);                                                                   \
    if (_e != cudaSuccess)                                                                     \
      return ::sq::set_error(::sq::SQ_ERR_CUDA, "%s failed at %s:%d: %s", #expr, __FILE__,     \
                             __LINE__, cudaGetErrorString(_e));                                \
  } while (0)

inline int require_device() {
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || n <= 0) {
    cudaGetLastError();
    return set_error(SQ_ERR_NODEVICE,
                     "no CUDA device available (%s): this library has no CPU fallback",
                     e == cudaSuccess ? "device count is 0" : cudaGetErrorString(e));
  }
  return SQ_OK;
}

// ---- device helpers ------------------------------------------------------------------------------
#ifdef __CUDACC__
// L2-coherent loads: the tree kernel's distance words are updated with L2 atomics by other
// threads of the CTA, so they must never be served from a stale L1 line.
__device__ __forceinline__ unsigned long long ld_cg_u64(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.global.cg.u64 %0, [%1];" : "=l"(v) : "l"(p));
  return v;
}
__device__ __forceinline__ int ld_cg_s32(const int* p) {
  int v;
  asm volatile("ld.global.cg.s32 %0, [%1];" : "=r"(v) : "l"(p));
  return v;
}
__device__ __forceinline__ unsigned lane_id() { return threadIdx.x & 31u; }
__device__ __forceinline__ unsigned lanemask_lt() {
  unsigned m;
  asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
  return m;
}
#endif

}  // namespace sq
