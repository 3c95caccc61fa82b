// Bandwidth-bound kernels of the Conformer encoder variant (EendEdaModel(encoder_type="conformer"),
// models.py:498-502, :525-528 -> torchaudio.models.Conformer): stand-alone (pre-) LayerNorm that
// writes the next GEMM's split-bf16 operand, and the convolution module's depthwise conv (k = 31,
// 'same' zero padding over the padded batch) + BatchNorm1d (eval: running statistics) + SiLU.
#include <algorithm>

#include <cuda_bf16.h>

#include "common.cuh"

namespace eda {

typedef __nv_bfloat16 bf16;

__device__ __forceinline__ void cf_split_store2(bf16* hi, bf16* lo, float x0, float x1) {
  const bf16 h0 = __float2bfloat16_rn(x0), h1 = __float2bfloat16_rn(x1);
  *reinterpret_cast<__nv_bfloat162*>(hi) = __halves2bfloat162(h0, h1);
  *reinterpret_cast<__nv_bfloat162*>(lo) = __halves2bfloat162(__float2bfloat16_rn(x0 - __bfloat162float(h0)),
                                                              __float2bfloat16_rn(x1 - __bfloat162float(h1)));
}

// nn.LayerNorm(256), eps 1e-5, two-pass variance like ATen.  Warp per row, 8 columns per lane.
__global__ void __launch_bounds__(256)
layernorm_split_kernel(const float* __restrict__ x, int ld, const float* __restrict__ gamma, const float* __restrict__ beta,
                       int M, float* __restrict__ out, int ld_out, bf16* __restrict__ hi, bf16* __restrict__ lo, int ld_s) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int c0 = lane * 4, c1 = 128 + lane * 4;
  const float4 g0 = __ldg(reinterpret_cast<const float4*>(gamma + c0)), g1 = __ldg(reinterpret_cast<const float4*>(gamma + c1));
  const float4 b0 = __ldg(reinterpret_cast<const float4*>(beta + c0)), b1 = __ldg(reinterpret_cast<const float4*>(beta + c1));
  const float gg[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
  const float bb[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
  for (int row = blockIdx.x * 8 + warp; row < M; row += gridDim.x * 8) {
    const float4 v0 = __ldg(reinterpret_cast<const float4*>(x + (size_t)row * ld + c0));
    const float4 v1 = __ldg(reinterpret_cast<const float4*>(x + (size_t)row * ld + c1));
    float v[8] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
    float s = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) s += v[j];
    const float mean = warp_sum(s) * (1.0f / 256.0f);
    float sq = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) { const float d = v[j] - mean; sq = fmaf(d, d, sq); }
    const float rstd = 1.0f / sqrtf(warp_sum(sq) * (1.0f / 256.0f) + 1e-5f);
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = (v[j] - mean) * rstd * gg[j] + bb[j];
    if (out) {
      *reinterpret_cast<float4*>(out + (size_t)row * ld_out + c0) = make_float4(v[0], v[1], v[2], v[3]);
      *reinterpret_cast<float4*>(out + (size_t)row * ld_out + c1) = make_float4(v[4], v[5], v[6], v[7]);
    }
    if (hi) {
      cf_split_store2(hi + (size_t)row * ld_s + c0, lo + (size_t)row * ld_s + c0, v[0], v[1]);
      cf_split_store2(hi + (size_t)row * ld_s + c0 + 2, lo + (size_t)row * ld_s + c0 + 2, v[2], v[3]);
      cf_split_store2(hi + (size_t)row * ld_s + c1, lo + (size_t)row * ld_s + c1, v[4], v[5]);
      cf_split_store2(hi + (size_t)row * ld_s + c1 + 2, lo + (size_t)row * ld_s + c1 + 2, v[6], v[7]);
    }
  }
}

int layernorm_split(const float* x, int ld, const float* gamma, const float* beta, int M, int D, float* out, int ld_out,
                    void* hi, void* lo, int ld_s, cudaStream_t st) {
  EDA_REQUIRE(D == 256, "eda_layernorm_split: D must be 256 (got %d)", D);
  EDA_REQUIRE(x && gamma && beta && M > 0 && (out || (hi && lo)), "eda_layernorm_split: bad arguments");
  EDA_REQUIRE(ld % 4 == 0 && ((uintptr_t)x | (uintptr_t)gamma | (uintptr_t)beta) % 16 == 0, "eda_layernorm_split: inputs must be 16B aligned");
  EDA_REQUIRE(!out || (ld_out % 4 == 0 && (uintptr_t)out % 16 == 0), "eda_layernorm_split: out misaligned");
  EDA_REQUIRE(!hi || ld_s % 2 == 0, "eda_layernorm_split: split stride must be even");
  layernorm_split_kernel<<<std::min(ceil_div(M, 8), num_sms() * 8), 256, 0, st>>>(x, ld, gamma, beta, M, out, ld_out,
                                                                                 (bf16*)hi, (bf16*)lo, ld_s);
  EDA_LAUNCH_CHECK();
  return EDA_OK;
}

// Depthwise Conv1d(C, C, 31, padding 15, groups C) + BatchNorm1d (eval) + SiLU over (B, T, C = 256)
// rows.  The caller folds conv bias and the running statistics into scale / shift:
//   out[b,t,c] = silu(scale[c] * sum_k w[c,k] x[b, t + k - 15, c] + shift[c]),  x = 0 outside [0, T).
// Thread = channel (coalesced 1 KB rows); a CTA walks 64 consecutive frames of one sequence with the
// 31-frame window held in registers.
constexpr int DW_K = 31, DW_TT = 64;
__global__ void __launch_bounds__(256)
dwconv_bn_silu_kernel(const float* __restrict__ x, int ld, const float* __restrict__ w, const float* __restrict__ scale,
                      const float* __restrict__ shift, int T, bf16* __restrict__ hi, bf16* __restrict__ lo, int ld_s) {
  const int c = threadIdx.x, b = blockIdx.y, t0 = blockIdx.x * DW_TT;
  float wk[DW_K], win[DW_K];
#pragma unroll
  for (int k = 0; k < DW_K; ++k) wk[k] = __ldg(w + c * DW_K + k);
  const float sc = __ldg(scale + c), sh = __ldg(shift + c);
  const float* xb = x + (size_t)b * T * ld + c;
  auto at = [&](int t) { return (t >= 0 && t < T) ? __ldg(xb + (size_t)t * ld) : 0.f; };
#pragma unroll
  for (int k = 0; k < DW_K - 1; ++k) win[k + 1] = at(t0 - 15 + k);     // the loop shifts before it appends
  const int t1 = min(T, t0 + DW_TT);
  for (int t = t0; t < t1; ++t) {
#pragma unroll
    for (int k = 0; k < DW_K - 1; ++k) win[k] = win[k + 1];
    win[DW_K - 1] = at(t + 15);
    float acc = 0.f;
#pragma unroll
    for (int k = 0; k < DW_K; ++k) acc = fmaf(wk[k], win[k], acc);
    const float y = fmaf(sc, acc, sh);
    const float o = y * (1.0f / (1.0f + expf(-y)));
    const bf16 h = __float2bfloat16_rn(o);
    const size_t idx = ((size_t)b * T + t) * ld_s + c;
    hi[idx] = h;
    lo[idx] = __float2bfloat16_rn(o - __bfloat162float(h));
  }
}

int dwconv_bn_silu(const float* x, int ld, const float* w, const float* scale, const float* shift, int B, int T, int C,
                   int K, void* hi, void* lo, int ld_s, cudaStream_t st) {
  EDA_REQUIRE(C == 256 && K == DW_K, "eda_dwconv_bn_silu: C must be 256 and the kernel size 31 (got %d, %d)", C, K);
  EDA_REQUIRE(x && w && scale && shift && hi && lo && B > 0 && B <= 65535 && T > 0, "eda_dwconv_bn_silu: bad arguments");
  dwconv_bn_silu_kernel<<<dim3(ceil_div(T, DW_TT), B), 256, 0, st>>>(x, ld, w, scale, shift, T, (bf16*)hi, (bf16*)lo, ld_s);
  EDA_LAUNCH_CHECK();
  return EDA_OK;
}

}  // namespace eda

extern "C" {

int eda_layernorm_split(const float* x, int ld, const float* gamma, const float* beta, int M, int D, float* out, int ld_out,
                        void* out_hi, void* out_lo, int ld_s, void* stream) {
  return eda::layernorm_split(x, ld, gamma, beta, M, D, out, ld_out, out_hi, out_lo, ld_s, (cudaStream_t)stream);
}

int eda_dwconv_bn_silu(const float* x, int ld, const float* w, const float* scale, const float* shift, int B, int T, int C,
                       int K, void* out_hi, void* out_lo, int ld_s, void* stream) {
  return eda::dwconv_bn_silu(x, ld, w, scale, shift, B, T, C, K, out_hi, out_lo, ld_s, (cudaStream_t)stream);
}

}  // extern "C"
