// HBM access-pattern probe for the GEMM operand / result layouts (build: nvcc -arch=sm_100a -O3).
//   w_rows : every thread writes `seg` bytes of its own row (row pitch `pitch`), like a thread-per-row epilogue
//   w_lin  : fully coalesced writes of the same bytes
//   r_rows : a warp reads 128 B segments of 32 different rows (pitch apart), walking along the row (K loop)
//   r_lin  : contiguous reads
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ void st256(void* p, uint32_t v) {
  asm volatile("st.global.v8.b32 [%0], {%1,%1,%1,%1,%1,%1,%1,%1};" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld256(const void* p) {
  uint32_t a, b, c, d, e, f, g, h;
  asm volatile("ld.global.nc.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(a), "=r"(b), "=r"(c), "=r"(d), "=r"(e), "=r"(f), "=r"(g), "=r"(h) : "l"(p));
  return a ^ b ^ c ^ d ^ e ^ f ^ g ^ h;
}

// rows x pitch bytes.  Tile = 128 rows x 512 B (one GEMM output tile of 256 bf16 columns); CTA = 256 threads:
// thread (r = tid & 127, half = tid >> 7) writes 4 x 64 B chunks of row r (columns half*256 B + c*64 B)
__global__ void w_rows(uint8_t* dst, int rows, int pitch) {
  const int tiles_n = pitch / 512, n_tiles = (rows / 128) * tiles_n;
  for (int t = blockIdx.x; t < n_tiles; t += gridDim.x) {
    const int m0 = (t / tiles_n) * 128, n0 = (t % tiles_n) * 512;
    uint8_t* p = dst + (size_t)(m0 + (threadIdx.x & 127)) * pitch + n0 + (threadIdx.x >> 7) * 256;
#pragma unroll
    for (int c = 0; c < 4; ++c) { st256(p + c * 64, t); st256(p + c * 64 + 32, t); }
  }
}
// same bytes, but a warp instruction covers 8 rows x 128 B (4 lanes x 32 B per row): full-line requests
__global__ void w_quad(uint8_t* dst, int rows, int pitch) {
  const int tiles_n = pitch / 512, n_tiles = (rows / 128) * tiles_n;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int t = blockIdx.x; t < n_tiles; t += gridDim.x) {
    const int m0 = (t / tiles_n) * 128, n0 = (t % tiles_n) * 512;
    // warp w: rows 32*(w&3).., column half (w>>2)*256 B; 2 groups of 128 B, 4 row-octets each
    uint8_t* p = dst + (size_t)(m0 + 32 * (warp & 3) + (lane & 7)) * pitch + n0 + (warp >> 2) * 256 + (lane >> 3) * 32;
#pragma unroll
    for (int g = 0; g < 2; ++g)
#pragma unroll
      for (int k = 0; k < 4; ++k) st256(p + (size_t)(8 * k) * pitch + g * 128, t);
  }
}
__global__ void w_lin(uint8_t* dst, size_t bytes) {
  for (size_t i = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * 32; i < bytes; i += (size_t)gridDim.x * blockDim.x * 32)
    st256(dst + i, (uint32_t)i);
}
// K loop: tile of 128 rows, each k step reads 128 B per row (a warp: 32 rows x 128 B = 4 x ld256 per lane)
__global__ void r_rows(const uint8_t* src, int rows, int pitch, uint32_t* sink) {
  uint32_t acc = 0;
  const int n_tiles = rows / 128;
  for (int t = blockIdx.x; t < n_tiles; t += gridDim.x) {
    const uint8_t* p = src + (size_t)(t * 128 + (threadIdx.x & 127)) * pitch + (threadIdx.x >> 7) * 64;
    for (int k = 0; k < pitch; k += 128) { acc ^= ld256(p + k); acc ^= ld256(p + k + 32); }
  }
  if (acc == 0x12345) sink[0] = acc;
}
__global__ void r_lin(const uint8_t* src, size_t bytes, uint32_t* sink) {
  uint32_t acc = 0;
  for (size_t i = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * 32; i < bytes; i += (size_t)gridDim.x * blockDim.x * 32)
    acc ^= ld256(src + i);
  if (acc == 0x12345) sink[0] = acc;
}

template <typename F>
static float timeit(F f, int n = 10) {
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  f(); f();
  cudaEventRecord(a);
  for (int i = 0; i < n; ++i) f();
  cudaEventRecord(b); cudaEventSynchronize(b);
  float ms; cudaEventElapsedTime(&ms, a, b);
  return ms / n;
}

int main() {
  const int rows = 32000;
  uint8_t *buf, *buf2; uint32_t* sink;
  const size_t cap = (size_t)rows * 8192;
  cudaMalloc(&buf, cap); cudaMalloc(&buf2, cap); cudaMalloc(&sink, 4);
  cudaMemset(buf, 1, cap); cudaMemset(buf2, 1, cap);
  for (int pitch : {512, 1024, 4096, 8192}) {
    const size_t bytes = (size_t)rows * pitch;
    uint8_t* b2[2] = {buf, buf2};
    int flip = 0;
    float ms;
    ms = timeit([&] { w_rows<<<148 * 4, 256>>>(b2[flip ^= 1], rows, pitch); });
    printf("pitch %5d  w_rows %7.1f us %6.0f GB/s", pitch, ms * 1e3, bytes / ms / 1e6);
    ms = timeit([&] { w_rows<<<148, 256>>>(b2[flip ^= 1], rows, pitch); });
    printf("   w_rows(1cta/sm) %7.1f us %6.0f GB/s", ms * 1e3, bytes / ms / 1e6);
    ms = timeit([&] { w_quad<<<148, 256>>>(b2[flip ^= 1], rows, pitch); });
    printf("   w_quad(1cta/sm) %7.1f us %6.0f GB/s", ms * 1e3, bytes / ms / 1e6);
    ms = timeit([&] { w_lin<<<148 * 8, 256>>>(b2[flip ^= 1], bytes); });
    printf("   w_lin %7.1f us %6.0f GB/s", ms * 1e3, bytes / ms / 1e6);
    ms = timeit([&] { r_rows<<<148 * 4, 256>>>(b2[flip ^= 1], rows, pitch, sink); });
    printf("   r_rows %7.1f us %6.0f GB/s", ms * 1e3, bytes / ms / 1e6);
    ms = timeit([&] { r_lin<<<148 * 8, 256>>>(b2[flip ^= 1], bytes, sink); });
    printf("   r_lin %7.1f us %6.0f GB/s\n", ms * 1e3, bytes / ms / 1e6);
  }
  {  // L2-resident window (rows = 4096, pitch 4096 = 16 MB), 20 back-to-back passes per launch-equivalent
    const int r2 = 4096, pitch = 4096;
    float ms = timeit([&] { for (int i = 0; i < 8; ++i) w_rows<<<148, 256>>>(buf, r2, pitch); });
    printf("L2 window: w_rows %6.0f GB/s", 8.0 * r2 * pitch / ms / 1e6);
    ms = timeit([&] { for (int i = 0; i < 8; ++i) w_quad<<<148, 256>>>(buf, r2, pitch); });
    printf("   w_quad %6.0f GB/s\n", 8.0 * r2 * pitch / ms / 1e6);
  }
  printf("%s\n", cudaGetErrorString(cudaDeviceSynchronize()));
  return 0;
}
