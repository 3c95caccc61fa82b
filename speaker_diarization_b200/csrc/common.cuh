// Shared helpers for the libeda_b200 kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "eda_b200.h"

namespace eda {

void set_error(const char* fmt, ...);
void count_launch(int n = 1);

#define EDA_REQUIRE(cond, ...)                 \
  do {                                         \
    if (!(cond)) {                             \
      eda::set_error(__VA_ARGS__);             \
      return EDA_EINVAL;                       \
    }                                          \
  } while (0)

#define EDA_CUDA(expr)                                                              \
  do {                                                                              \
    cudaError_t _e = (expr);                                                        \
    if (_e != cudaSuccess) {                                                        \
      eda::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
      return EDA_ECUDA;                                                             \
    }                                                                               \
  } while (0)

#define EDA_LAUNCH_CHECK()                                                          \
  do {                                                                              \
    cudaError_t _e = cudaGetLastError();                                            \
    if (_e != cudaSuccess) {                                                        \
      eda::set_error("kernel launch failed: %s (%s:%d)", cudaGetErrorString(_e), __FILE__, __LINE__); \
      return EDA_ECUDA;                                                             \
    }                                                                               \
    eda::count_launch();                                                            \
  } while (0)

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

// One-time per-DEVICE set-up (cudaFuncSetAttribute is per device, a process may drive several GPUs):
//   static DeviceOnce once;  if (!once.done()) { ...set attributes...; once.mark(); }
struct DeviceOnce {
  unsigned long long mask = 0;
  static unsigned long long bit() { int dev = 0; cudaGetDevice(&dev); return 1ull << (dev & 63); }
  bool done() const { return (__atomic_load_n(&mask, __ATOMIC_ACQUIRE) & bit()) != 0; }
  void mark() { __atomic_fetch_or(&mask, bit(), __ATOMIC_RELEASE); }
};

// Library-owned scratch of the convenience entry points, keyed by (device, stream, tag): calls in flight on
// different streams or GPUs never share a buffer; calls on one stream reuse it in stream order.  Grow-only;
// growing waits for that stream.  Returns nullptr (error set) on failure.
void* stream_scratch(int tag, cudaStream_t st, size_t bytes);

template <typename T>
__host__ __device__ constexpr T ceil_div(T a, T b) { return (a + b - 1) / b; }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// ---- dropout (training step): counter-based Philox4x32-7, the same function in the forward
// epilogues, the backward kernels and eda_dropout_mask (tests), so a mask is never stored.
// Element (row, col) of site `site` keeps its value iff the 16-bit half (col & 1) of word ((col & 7) >> 1) of
// philox(key = seed, counter = (col >> 3, row, site, 0)) >= thresh = round(p * 2^16): one call decides 8 elements.
struct DropoutCfg {
  uint32_t thresh;       // 16-bit threshold; 0 = dropout off
  uint32_t seed_lo, seed_hi, site;
  float scale;           // 1 / (1 - p)
};
__host__ __device__ inline DropoutCfg make_dropout(float p, uint64_t seed, int site) {
  DropoutCfg d;
  d.thresh = p > 0.f ? (uint32_t)((double)p * 65536.0 + 0.5) : 0u;
  d.seed_lo = (uint32_t)seed; d.seed_hi = (uint32_t)(seed >> 32); d.site = (uint32_t)site;
  d.scale = p > 0.f ? 1.0f / (1.0f - p) : 1.0f;
  return d;
}
__device__ __forceinline__ uint4 philox4x32_7(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1) {
#pragma unroll
  for (int r = 0; r < 7; ++r) {
    const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
    const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
    c0 = hi1 ^ c1 ^ k0; c1 = lo1; c2 = hi0 ^ c3 ^ k1; c3 = lo0;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
  return make_uint4(c0, c1, c2, c3);
}
// keep-bits of the 8 elements (row, 8 * colgroup .. + 7): bit j set = element kept
__device__ __forceinline__ uint32_t dropout_keep8(const DropoutCfg& d, uint32_t row, uint32_t colgroup) {
  const uint4 r = philox4x32_7(colgroup, row, d.site, 0u, d.seed_lo, d.seed_hi);
  const uint32_t w[4] = {r.x, r.y, r.z, r.w};
  uint32_t k = 0;
#pragma unroll
  for (int j = 0; j < 4; ++j)
    k |= ((w[j] & 0xffffu) >= d.thresh ? 1u : 0u) << (2 * j) | ((w[j] >> 16) >= d.thresh ? 2u : 0u) << (2 * j);
  return k;
}

// sigmoid the way ATen's CPU/GPU kernels define it: 1 / (1 + exp(-x)), fp32.
__device__ __forceinline__ float sigmoidf_ref(float x) { return 1.0f / (1.0f + expf(-x)); }

}  // namespace eda
