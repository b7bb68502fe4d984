// Shared device/host helpers for the cxxnet_b200 kernel library (sm_100a only).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>

#define CXN_API extern "C" __attribute__((visibility("default")))

// Number of kernel launches issued by this library in this process (bench.py reports it).
extern "C" unsigned long long cxn_launch_counter_;

// Optional histogram of launches per call site (CXN_LAUNCH_LOG=1: printed to stderr at exit; elementwise.cu).
extern "C" int cxn_launch_log_on_;
extern "C" void cxxnet_note_launch_(const char* file, int line);

// Every launcher returns 0 on success or a cudaError_t / negative framework error.
#define CXN_CHECK_LAUNCH()                                                     \
  do {                                                                         \
    ++cxn_launch_counter_;                                                     \
    if (cxn_launch_log_on_) cxxnet_note_launch_(__FILE__, __LINE__);              \
    cudaError_t e__ = cudaGetLastError();                                      \
    if (e__ != cudaSuccess) {                                                  \
      fprintf(stderr, "[cxxnet_b200] %s:%d launch failed: %s\n", __FILE__,     \
              __LINE__, cudaGetErrorString(e__));                              \
      return (int)e__;                                                         \
    }                                                                          \
  } while (0)

namespace cxn {

typedef __nv_bfloat16 bf16;

__host__ __device__ inline int ceil_div(int a, int b) { return (a + b - 1) / b; }
__host__ __device__ inline long long ceil_div_ll(long long a, long long b) { return (a + b - 1) / b; }

inline int num_sms() {
  static int n = 0;
  if (n == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    if (n <= 0) n = 148;
  }
  return n;
}

// True when [p, p + bytes) is memory of the CURRENT device that this process allocated itself: not a mapping of
// another GPU's memory (CUDA IPC, peer access, cuMemMap), not host memory.  The cluster-multicast TMA routes ask
// this before building a multicast tensor map: TMA multicast loads here read only this GPU's memory, since on
// B200 a TMA multicast load from another GPU's memory hung the GPU.  Anything not provably local answers false.
// (defined in fused_exchange.cu, which also records every range it maps from a peer)
bool memory_is_local(const void* p, size_t bytes);

// grid size for a grid-stride elementwise kernel: enough CTAs to fill the chip
// a few times over, never more than the work needs.
inline int ew_grid(long long n_items, int threads, int waves = 8) {
  long long need = (n_items + threads - 1) / threads;
  long long cap = (long long)num_sms() * waves;
  if (need < 1) need = 1;
  return (int)(need < cap ? need : cap);
}

// ---------------------------------------------------------------- Philox4x32-10
// Counter-based RNG used by dropout / insanity / prelu noise / weight init so
// that a mask can be regenerated from (seed, offset) instead of stored.
struct Philox {
  uint32_t key0, key1;
  __host__ __device__ Philox(uint64_t seed) : key0((uint32_t)seed), key1((uint32_t)(seed >> 32)) {}
  __host__ __device__ static inline void mulhilo(uint32_t a, uint32_t b, uint32_t& hi, uint32_t& lo) {
#ifdef __CUDA_ARCH__
    hi = __umulhi(a, b);
    lo = a * b;
#else
    uint64_t p = (uint64_t)a * b;
    hi = (uint32_t)(p >> 32);
    lo = (uint32_t)p;
#endif
  }
  // 4 x uint32 for 128-bit counter (c0..c3)
  __host__ __device__ inline void operator()(uint64_t ctr_lo, uint64_t ctr_hi, uint32_t out[4]) const {
    uint32_t c0 = (uint32_t)ctr_lo, c1 = (uint32_t)(ctr_lo >> 32);
    uint32_t c2 = (uint32_t)ctr_hi, c3 = (uint32_t)(ctr_hi >> 32);
    uint32_t k0 = key0, k1 = key1;
#pragma unroll
    for (int r = 0; r < 10; ++r) {
      uint32_t hi0, lo0, hi1, lo1;
      mulhilo(0xD2511F53u, c0, hi0, lo0);
      mulhilo(0xCD9E8D57u, c2, hi1, lo1);
      uint32_t n0 = hi1 ^ c1 ^ k0, n1 = lo1, n2 = hi0 ^ c3 ^ k1, n3 = lo0;
      c0 = n0; c1 = n1; c2 = n2; c3 = n3;
      k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
  }
};
// uniform in [0,1)
__host__ __device__ inline float u01(uint32_t x) { return (x >> 8) * (1.0f / 16777216.0f); }

__device__ inline float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ inline float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

}  // namespace cxn
