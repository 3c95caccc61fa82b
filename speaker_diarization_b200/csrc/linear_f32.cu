// nn.Linear with fused epilogues, fp32 CUDA-core path (EDA_PREC_FP32).
//
// This is the reference-accuracy arithmetic (plain FFMA, fp32 accumulation): it is what
// the tcgen05 path is validated against on the device, and the path used when bit-level
// agreement of thresholded labels matters more than speed.
//
// C[m,n] = sum_k A[m,k] * W[n,k]   (A and W both K-contiguous: PyTorch weight layout)
// Tile BM x BN x 16, 256 threads, 8x8 outputs per thread, double-buffered smem with
// register prefetch.  For the LayerNorm epilogues BN == N == 256 so one warp owns whole
// rows and the row statistics are a warp shuffle.
#include "common.cuh"

namespace eda {

constexpr int kBK = 16;
constexpr int kThreads = 256;

template <int BM, int BN, int EPI>
__global__ void __launch_bounds__(kThreads, 2)
linear_f32_kernel(const float* __restrict__ A, int lda, const float* __restrict__ W, int ldw,
                  const float* __restrict__ bias, const float* __restrict__ resid, int ldr,
                  const float* __restrict__ gamma, const float* __restrict__ beta,
                  float* __restrict__ C, int ldc, int M, int N, int K) {
  constexpr int TX = BN / 8;            // threads along N
  constexpr int A_F4 = BM * 4 / kThreads;  // float4 loads per thread for the A tile
  constexpr int B_F4 = BN * 4 / kThreads;
  static_assert(kThreads / TX * 8 == BM, "tile/thread mismatch");
  __shared__ __align__(16) float As[2][kBK][BM + 4];
  __shared__ __align__(16) float Bs[2][kBK][BN + 4];

  const int tid = threadIdx.x;
  const int tx = tid % TX, ty = tid / TX;
  const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;

  float4 pa[A_F4], pb[B_F4];
  auto gload = [&](int k0) {
#pragma unroll
    for (int i = 0; i < A_F4; ++i) {
      int f = tid + i * kThreads, r = f >> 2, k = k0 + (f & 3) * 4;
      pa[i] = (m0 + r < M && k < K)
                  ? __ldg(reinterpret_cast<const float4*>(A + (size_t)(m0 + r) * lda + k))
                  : make_float4(0.f, 0.f, 0.f, 0.f);
    }
#pragma unroll
    for (int i = 0; i < B_F4; ++i) {
      int f = tid + i * kThreads, r = f >> 2, k = k0 + (f & 3) * 4;
      pb[i] = (n0 + r < N && k < K)
                  ? __ldg(reinterpret_cast<const float4*>(W + (size_t)(n0 + r) * ldw + k))
                  : make_float4(0.f, 0.f, 0.f, 0.f);
    }
  };
  auto sstore = [&](int buf) {
#pragma unroll
    for (int i = 0; i < A_F4; ++i) {
      int f = tid + i * kThreads, r = f >> 2, k = (f & 3) * 4;
      As[buf][k + 0][r] = pa[i].x; As[buf][k + 1][r] = pa[i].y;
      As[buf][k + 2][r] = pa[i].z; As[buf][k + 3][r] = pa[i].w;
    }
#pragma unroll
    for (int i = 0; i < B_F4; ++i) {
      int f = tid + i * kThreads, r = f >> 2, k = (f & 3) * 4;
      Bs[buf][k + 0][r] = pb[i].x; Bs[buf][k + 1][r] = pb[i].y;
      Bs[buf][k + 2][r] = pb[i].z; Bs[buf][k + 3][r] = pb[i].w;
    }
  };

  float acc[8][8];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;

  const int nk = (K + kBK - 1) / kBK;
  gload(0);
  sstore(0);
  __syncthreads();
  for (int kt = 0; kt < nk; ++kt) {
    const int buf = kt & 1;
    if (kt + 1 < nk) gload((kt + 1) * kBK);
#pragma unroll
    for (int k = 0; k < kBK; ++k) {
      float4 a0 = *reinterpret_cast<const float4*>(&As[buf][k][ty * 8]);
      float4 a1 = *reinterpret_cast<const float4*>(&As[buf][k][ty * 8 + 4]);
      float4 b0 = *reinterpret_cast<const float4*>(&Bs[buf][k][tx * 4]);
      float4 b1 = *reinterpret_cast<const float4*>(&Bs[buf][k][BN / 2 + tx * 4]);
      float a[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
      float b[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
      for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    if (kt + 1 < nk) {
      sstore(buf ^ 1);
      __syncthreads();
    }
  }

  // ---- epilogue -------------------------------------------------------------------
  const int c0 = n0 + tx * 4, c1 = n0 + BN / 2 + tx * 4;
  float bv[8];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    bv[j] = (bias && c0 + j < N) ? __ldg(bias + c0 + j) : 0.f;
    bv[4 + j] = (bias && c1 + j < N) ? __ldg(bias + c1 + j) : 0.f;
  }
  float gv[8], bt[8];
  if constexpr (EPI == EDA_EPI_BIAS_LN || EPI == EDA_EPI_BIAS_RES_LN) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      gv[j] = __ldg(gamma + c0 + j); gv[4 + j] = __ldg(gamma + c1 + j);
      bt[j] = __ldg(beta + c0 + j);  bt[4 + j] = __ldg(beta + c1 + j);
    }
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int row = m0 + ty * 8 + i;
    const bool row_ok = row < M;                       // warp-uniform for the LN tiles
    float v[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = acc[i][j] + bv[j];
    if constexpr (EPI == EDA_EPI_BIAS_RELU) {
#pragma unroll
      for (int j = 0; j < 8; ++j) v[j] = fmaxf(v[j], 0.f);
    }
    if constexpr (EPI == EDA_EPI_BIAS_RES_LN) {
      if (row_ok) {
        float4 r0 = __ldg(reinterpret_cast<const float4*>(resid + (size_t)row * ldr + c0));
        float4 r1 = __ldg(reinterpret_cast<const float4*>(resid + (size_t)row * ldr + c1));
        v[0] += r0.x; v[1] += r0.y; v[2] += r0.z; v[3] += r0.w;
        v[4] += r1.x; v[5] += r1.y; v[6] += r1.z; v[7] += r1.w;
      }
    }
    if constexpr (EPI == EDA_EPI_BIAS_LN || EPI == EDA_EPI_BIAS_RES_LN) {
      // BN == N == 256 and TX == 32: the 32 lanes of this warp hold the whole row.
      float s = 0.f;
#pragma unroll
      for (int j = 0; j < 8; ++j) s += v[j];
      const float mean = warp_sum(s) * (1.0f / BN);
      float q = 0.f;
#pragma unroll
      for (int j = 0; j < 8; ++j) { float d = v[j] - mean; q = fmaf(d, d, q); }
      const float rstd = 1.0f / sqrtf(warp_sum(q) * (1.0f / BN) + 1e-5f);
#pragma unroll
      for (int j = 0; j < 8; ++j) v[j] = (v[j] - mean) * rstd * gv[j] + bt[j];
    }
    if (row_ok) {
      float* crow = C + (size_t)row * ldc;
      if (c0 + 3 < N) *reinterpret_cast<float4*>(crow + c0) = make_float4(v[0], v[1], v[2], v[3]);
      else
        for (int j = 0; j < 4; ++j) if (c0 + j < N) crow[c0 + j] = v[j];
      if (c1 + 3 < N) *reinterpret_cast<float4*>(crow + c1) = make_float4(v[4], v[5], v[6], v[7]);
      else
        for (int j = 0; j < 4; ++j) if (c1 + j < N) crow[c1 + j] = v[4 + j];
    }
  }
}

int linear_f32(const float* A, int lda, const float* W, int ldw, const float* bias,
               const float* resid, int ldr, const float* gamma, const float* beta, float* C,
               int ldc, int M, int N, int K, int epi, cudaStream_t st) {
  EDA_REQUIRE(lda % 4 == 0 && ldw % 4 == 0 && ldc % 4 == 0 && K % 4 == 0,
              "eda_linear(fp32): lda/ldw/ldc/K must be multiples of 4 (got %d %d %d %d)", lda, ldw, ldc, K);
  EDA_REQUIRE(((uintptr_t)A | (uintptr_t)W | (uintptr_t)C) % 16 == 0, "eda_linear: pointers must be 16B aligned");
  if (epi == EDA_EPI_BIAS_LN || epi == EDA_EPI_BIAS_RES_LN) {
    EDA_REQUIRE(N == 256, "eda_linear: LayerNorm epilogues need N == 256 (got %d)", N);
    EDA_REQUIRE(gamma && beta, "eda_linear: LayerNorm epilogue needs gamma/beta");
    EDA_REQUIRE(epi != EDA_EPI_BIAS_RES_LN || (resid && ldr % 4 == 0), "eda_linear: residual missing/misaligned");
    dim3 grid(1, ceil_div(M, 64));
    if (epi == EDA_EPI_BIAS_LN)
      linear_f32_kernel<64, 256, EDA_EPI_BIAS_LN><<<grid, kThreads, 0, st>>>(A, lda, W, ldw, bias, resid, ldr, gamma, beta, C, ldc, M, N, K);
    else
      linear_f32_kernel<64, 256, EDA_EPI_BIAS_RES_LN><<<grid, kThreads, 0, st>>>(A, lda, W, ldw, bias, resid, ldr, gamma, beta, C, ldc, M, N, K);
  } else {
    dim3 grid(ceil_div(N, 128), ceil_div(M, 128));
    if (epi == EDA_EPI_BIAS)
      linear_f32_kernel<128, 128, EDA_EPI_BIAS><<<grid, kThreads, 0, st>>>(A, lda, W, ldw, bias, resid, ldr, gamma, beta, C, ldc, M, N, K);
    else if (epi == EDA_EPI_BIAS_RELU)
      linear_f32_kernel<128, 128, EDA_EPI_BIAS_RELU><<<grid, kThreads, 0, st>>>(A, lda, W, ldw, bias, resid, ldr, gamma, beta, C, ldc, M, N, K);
    else
      EDA_REQUIRE(false, "eda_linear: unknown epilogue %d", epi);
  }
  EDA_LAUNCH_CHECK();
  return EDA_OK;
}

}  // namespace eda
