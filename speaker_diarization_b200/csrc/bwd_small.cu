// Bandwidth-bound kernels of the training step: the backward passes of the
// posterior bmm, the attractor existence head + attractor loss, LayerNorm, bias (column sums),
// the attention row statistic delta = rowsum(dO * O), plus gradient scaling and fused Adam.
// What they differentiate: models.py:254 (bmm), :350-371 (attractor_loss),
// encoder_decoder_attractor.py:50-54 (Linear + sigmoid), nn.LayerNorm inside
// nn.TransformerEncoderLayer / encoder_norm (models.py:186-193), torch.optim.Adam as configured at
// train_eda.py:131.  All reductions over rows go through per-CTA partial sums that a second tiny
// kernel adds in a fixed order (deterministic), except the attractor gradient of the posterior
// (fp32 atomics over a handful of CTAs per sequence).
#include <algorithm>

#include <cuda_bf16.h>

#include "common.cuh"

namespace eda {

typedef __nv_bfloat16 bf16;

__device__ __forceinline__ void split_store2(bf16* hi, bf16* lo, float x0, float x1) {
  const bf16 h0 = __float2bfloat16_rn(x0), h1 = __float2bfloat16_rn(x1);
  *reinterpret_cast<__nv_bfloat162*>(hi) = __halves2bfloat162(h0, h1);
  *reinterpret_cast<__nv_bfloat162*>(lo) = __halves2bfloat162(__float2bfloat16_rn(x0 - __bfloat162float(h0)),
                                                              __float2bfloat16_rn(x1 - __bfloat162float(h1)));
}

// out_k[j] = sum over CTAs of part[cta][k][j] (k-th of n_vec vectors of length n).  A block owns 32
// columns; its 8 warps each add every 8th partial (coalesced 128 B reads), then the 8 sums are
// added in warp order: a fixed summation tree, whatever the launch geometry.
__global__ void __launch_bounds__(256)
partial_finish_kernel(const float* __restrict__ part, int n_cta, int n_vec, int n,
                      float* out0, float* out1, float* out2, int accumulate) {
  __shared__ float red[8][32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int i = blockIdx.x * 32 + lane;
  float acc = 0.f;
  if (i < n_vec * n)
    for (int c = warp; c < n_cta; c += 8) acc += part[(size_t)c * n_vec * n + i];
  red[warp][lane] = acc;
  __syncthreads();
  if (warp == 0 && i < n_vec * n) {
    const int k = i / n, j = i % n;
    float* out = k == 0 ? out0 : k == 1 ? out1 : out2;
    if (out) {
      float t = accumulate ? out[j] : 0.f;
#pragma unroll
      for (int w = 0; w < 8; ++w) t += red[w][lane];
      out[j] = t;
    }
  }
}

// ------------------------------------------------------------------ LayerNorm backward ------
// y = gamma * xhat + beta was the forward output; xhat is recovered as (y - beta) / gamma (gamma
// must be non-zero), rstd comes from the forward epilogue.  Warp per row, 8 columns per lane.
constexpr int LNB_CTAS = 592;
__global__ void __launch_bounds__(256)
ln_bwd_kernel(const float* __restrict__ dy, int ld_dy, const float* __restrict__ y, int ld_y,
              const float* __restrict__ gamma, const float* __restrict__ beta, const float* __restrict__ rstd,
              int M, float* __restrict__ dx, int ld_dx, bf16* __restrict__ dx_hi, bf16* __restrict__ dx_lo, int ld_s,
              float* __restrict__ part, const DropoutCfg drop) {
  __shared__ float red[8][3][256];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int c0 = lane * 4, c1 = 128 + lane * 4;
  const float4 g0 = __ldg(reinterpret_cast<const float4*>(gamma + c0)), g1 = __ldg(reinterpret_cast<const float4*>(gamma + c1));
  const float4 b0 = __ldg(reinterpret_cast<const float4*>(beta + c0)), b1 = __ldg(reinterpret_cast<const float4*>(beta + c1));
  const float gg[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w};
  const float bb[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
  float ig[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) ig[j] = 1.0f / gg[j];
  float a_dg[8] = {}, a_db[8] = {}, a_cs[8] = {};
  for (int row = blockIdx.x * 8 + warp; row < M; row += gridDim.x * 8) {
    const float4 d0 = __ldg(reinterpret_cast<const float4*>(dy + (size_t)row * ld_dy + c0));
    const float4 d1 = __ldg(reinterpret_cast<const float4*>(dy + (size_t)row * ld_dy + c1));
    const float4 y0 = __ldg(reinterpret_cast<const float4*>(y + (size_t)row * ld_y + c0));
    const float4 y1 = __ldg(reinterpret_cast<const float4*>(y + (size_t)row * ld_y + c1));
    const float dd[8] = {d0.x, d0.y, d0.z, d0.w, d1.x, d1.y, d1.z, d1.w};
    const float yy[8] = {y0.x, y0.y, y0.z, y0.w, y1.x, y1.y, y1.z, y1.w};
    float xh[8], a[8], s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      xh[j] = (yy[j] - bb[j]) * ig[j];
      a[j] = dd[j] * gg[j];
      s1 += a[j];
      s2 = fmaf(a[j], xh[j], s2);
      a_dg[j] = fmaf(dd[j], xh[j], a_dg[j]);
      a_db[j] += dd[j];
    }
    s1 = warp_sum(s1) * (1.0f / 256.0f);
    s2 = warp_sum(s2) * (1.0f / 256.0f);
    const float rs = __ldg(rstd + row);
    float o[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) o[j] = rs * (a[j] - s1 - xh[j] * s2);
    if (dx) {        // the residual branch sees the undropped gradient
      *reinterpret_cast<float4*>(dx + (size_t)row * ld_dx + c0) = make_float4(o[0], o[1], o[2], o[3]);
      *reinterpret_cast<float4*>(dx + (size_t)row * ld_dx + c1) = make_float4(o[4], o[5], o[6], o[7]);
    }
    if (drop.thresh) {   // the Linear branch went through dropout in the forward: same mask, same scale
      // this lane's columns 4 lane .. +3 and 128 + 4 lane .. +3 are one half of an 8-column group each
      const uint32_t k0 = dropout_keep8(drop, (uint32_t)row, (uint32_t)(lane >> 1)) >> (4 * (lane & 1));
      const uint32_t k1 = dropout_keep8(drop, (uint32_t)row, 16u + (lane >> 1)) >> (4 * (lane & 1));
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        o[j] = (k0 >> j) & 1u ? o[j] * drop.scale : 0.f;
        o[4 + j] = (k1 >> j) & 1u ? o[4 + j] * drop.scale : 0.f;
      }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) a_cs[j] += o[j];
    if (dx_hi) {
      split_store2(dx_hi + (size_t)row * ld_s + c0, dx_lo + (size_t)row * ld_s + c0, o[0], o[1]);
      split_store2(dx_hi + (size_t)row * ld_s + c0 + 2, dx_lo + (size_t)row * ld_s + c0 + 2, o[2], o[3]);
      split_store2(dx_hi + (size_t)row * ld_s + c1, dx_lo + (size_t)row * ld_s + c1, o[4], o[5]);
      split_store2(dx_hi + (size_t)row * ld_s + c1 + 2, dx_lo + (size_t)row * ld_s + c1 + 2, o[6], o[7]);
    }
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const int c = j < 4 ? c0 + j : c1 + j - 4;
    red[warp][0][c] = a_dg[j]; red[warp][1][c] = a_db[j]; red[warp][2][c] = a_cs[j];
  }
  __syncthreads();
  for (int i = threadIdx.x; i < 3 * 256; i += 256) {
    float s = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) s += (&red[w][0][0])[i];
    part[(size_t)blockIdx.x * 768 + i] = s;
  }
}

int ln_bwd(const float* dy, int ld_dy, const float* y, int ld_y, const float* gamma, const float* beta,
           const float* rstd, int M, int D, float* dx, int ld_dx, void* dx_hi, void* dx_lo, int ld_s,
           float* dgamma, float* dbeta, float* dcolsum, float* work, size_t work_bytes, int accumulate,
           const DropoutCfg& drop, cudaStream_t st) {
  EDA_REQUIRE(D == 256, "eda_ln_bwd: D must be 256 (got %d)", D);
  EDA_REQUIRE(dy && y && gamma && beta && rstd && work && M > 0 && (dx || dx_hi), "eda_ln_bwd: bad arguments");
  EDA_REQUIRE(ld_dy % 4 == 0 && ld_y % 4 == 0 && ((uintptr_t)dy | (uintptr_t)y | (uintptr_t)gamma | (uintptr_t)beta) % 16 == 0, "eda_ln_bwd: inputs must be 16B aligned");
  EDA_REQUIRE(!dx || (ld_dx % 4 == 0 && (uintptr_t)dx % 16 == 0), "eda_ln_bwd: dx misaligned");
  EDA_REQUIRE(!dx_hi || (dx_lo && ld_s % 2 == 0), "eda_ln_bwd: split output needs both halves");
  const int n_cta = std::min(LNB_CTAS, ceil_div(M, 8));
  EDA_REQUIRE(work_bytes >= (size_t)n_cta * 768 * sizeof(float), "eda_ln_bwd: workspace too small (%d CTAs x 3 KB)", n_cta);
  ln_bwd_kernel<<<n_cta, 256, 0, st>>>(dy, ld_dy, y, ld_y, gamma, beta, rstd, M, dx, ld_dx, (bf16*)dx_hi, (bf16*)dx_lo, ld_s, work, drop);
  EDA_LAUNCH_CHECK();
  partial_finish_kernel<<<ceil_div(3 * 256, 32), 256, 0, st>>>(work, n_cta, 3, 256, dgamma, dbeta, dcolsum, accumulate);
  EDA_LAUNCH_CHECK();
  return EDA_OK;
}

// ------------------------------------------------------------------ column sums (bias grads) ---
// out[j] = sum_rows (hi + lo)[row][j]; N % 8 == 0, N <= 2048.  A thread owns 8 columns (16 B of hi,
// 16 B of lo per row); the 256 threads of a CTA cover 2048 / N rows per pass, four passes in flight.
constexpr int CS_CTAS = 592;
__global__ void __launch_bounds__(256)
colsum_split_kernel(const bf16* __restrict__ hi, const bf16* __restrict__ lo, int M, int N, int ld, float* __restrict__ part) {
  __shared__ float red[256][9];
  const int tpr = N / 8;                          // threads per row
  const int rpp = 256 / tpr;                      // rows per pass
  const int sub = threadIdx.x / tpr, c8 = (threadIdx.x % tpr) * 8;
  float acc[8] = {};
  if (sub < rpp) {
    const int stride = gridDim.x * rpp;
    auto add = [&](const uint4& a, const uint4& b) {
      const uint32_t aw[4] = {a.x, a.y, a.z, a.w}, bw[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        acc[2 * j] += __uint_as_float(aw[j] << 16) + __uint_as_float(bw[j] << 16);
        acc[2 * j + 1] += __uint_as_float(aw[j] & 0xffff0000u) + __uint_as_float(bw[j] & 0xffff0000u);
      }
    };
    int r = blockIdx.x * rpp + sub;
    for (; r + 3 * stride < M; r += 4 * stride) {
      uint4 a[4], b[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        a[k] = __ldg(reinterpret_cast<const uint4*>(hi + (size_t)(r + k * stride) * ld + c8));
        b[k] = __ldg(reinterpret_cast<const uint4*>(lo + (size_t)(r + k * stride) * ld + c8));
      }
#pragma unroll
      for (int k = 0; k < 4; ++k) add(a[k], b[k]);
    }
    for (; r < M; r += stride)
      add(__ldg(reinterpret_cast<const uint4*>(hi + (size_t)r * ld + c8)), __ldg(reinterpret_cast<const uint4*>(lo + (size_t)r * ld + c8)));
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) red[threadIdx.x][j] = acc[j];
  __syncthreads();
  if (threadIdx.x < tpr) {                        // add the row groups in order
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float t = 0.f;
      for (int g = 0; g < rpp; ++g) t += red[g * tpr + threadIdx.x][j];
      part[(size_t)blockIdx.x * N + c8 + j] = t;
    }
  }
}

int colsum_split(const void* hi, const void* lo, int M, int N, int ld, float* out, float* work, size_t work_bytes,
                 int accumulate, cudaStream_t st) {
  EDA_REQUIRE(hi && lo && out && work && M > 0 && N > 0 && N % 8 == 0 && N <= 2048 && ld % 8 == 0, "eda_colsum_split: bad arguments (N %% 8 == 0, N <= 2048, ld %% 8 == 0)");
  EDA_REQUIRE(((uintptr_t)hi | (uintptr_t)lo) % 16 == 0, "eda_colsum_split: operands must be 16B aligned");
  const int n_cta = std::min(CS_CTAS, ceil_div(M, std::max(1, 256 / (N / 8))));
  EDA_REQUIRE(work_bytes >= (size_t)n_cta * N * sizeof(float), "eda_colsum_split: workspace too small");
  colsum_split_kernel<<<n_cta, 256, 0, st>>>((const bf16*)hi, (const bf16*)lo, M, N, ld, work);
  EDA_LAUNCH_CHECK();
  partial_finish_kernel<<<ceil_div(N, 32), 256, 0, st>>>(work, n_cta, 1, N, out, nullptr, nullptr, accumulate);
  EDA_LAUNCH_CHECK();
  return EDA_OK;
}

// ------------------------------------------------------------------ posterior backward -------
// pred[b,t,s] = emb[b,t,:] . att[b,s,:]  (models.py:254)
//   demb[b,t,:] = sum_s dpred[b,t,s] att[b,s,:]          (written)
//   datt[b,s,:] += sum_t dpred[b,t,s] emb[b,t,:]          (atomics; caller zeroes datt first)
template <int NATT>
__global__ void __launch_bounds__(256)
posterior_bwd_kernel(const float* __restrict__ dpred, const float* __restrict__ emb, int ld_emb,
                     const float* __restrict__ att, int att_stride_rows, int Tmax, int n_att,
                     float* __restrict__ demb, int ld_demb, float* __restrict__ datt) {
  constexpr int D = 256;
  __shared__ __align__(16) float As[NATT][D];
  __shared__ float red[8][4][D];          // cross-warp reduction, four attractors at a time
  const int b = blockIdx.y;
  for (int i = threadIdx.x; i < n_att * D; i += blockDim.x)
    As[i / D][i % D] = __ldg(att + (size_t)b * att_stride_rows * D + i);
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float acc[NATT][8];
#pragma unroll
  for (int s = 0; s < NATT; ++s)
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[s][j] = 0.f;
  for (int t = blockIdx.x * 8 + warp; t < Tmax; t += gridDim.x * 8) {
    const size_t row = (size_t)b * Tmax + t;
    const float4 e0 = __ldg(reinterpret_cast<const float4*>(emb + row * ld_emb) + lane);
    const float4 e1 = __ldg(reinterpret_cast<const float4*>(emb + row * ld_emb) + 32 + lane);
    const float ee[8] = {e0.x, e0.y, e0.z, e0.w, e1.x, e1.y, e1.z, e1.w};
    const float mine = lane < n_att ? __ldg(dpred + row * n_att + lane) : 0.f;
    float o[8] = {};
#pragma unroll
    for (int s = 0; s < NATT; ++s) {
      if (s < n_att) {
        const float g = __shfl_sync(0xffffffffu, mine, s);
        const float4 a0 = *reinterpret_cast<const float4*>(&As[s][lane * 4]);
        const float4 a1 = *reinterpret_cast<const float4*>(&As[s][128 + lane * 4]);
        o[0] = fmaf(g, a0.x, o[0]); o[1] = fmaf(g, a0.y, o[1]); o[2] = fmaf(g, a0.z, o[2]); o[3] = fmaf(g, a0.w, o[3]);
        o[4] = fmaf(g, a1.x, o[4]); o[5] = fmaf(g, a1.y, o[5]); o[6] = fmaf(g, a1.z, o[6]); o[7] = fmaf(g, a1.w, o[7]);
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[s][j] = fmaf(g, ee[j], acc[s][j]);
      }
    }
    float* dst = demb + row * ld_demb;
    *(reinterpret_cast<float4*>(dst) + lane) = make_float4(o[0], o[1], o[2], o[3]);
    *(reinterpret_cast<float4*>(dst) + 32 + lane) = make_float4(o[4], o[5], o[6], o[7]);
  }
#pragma unroll
  for (int s0 = 0; s0 < NATT; s0 += 4) {
    if (s0 >= n_att) break;
#pragma unroll
    for (int s = 0; s < 4; ++s)
#pragma unroll
      for (int j = 0; j < 8; ++j) red[warp][s][j < 4 ? lane * 4 + j : 128 + lane * 4 + j - 4] = acc[s0 + s][j];
    __syncthreads();
    for (int i = threadIdx.x; i < min(4, n_att - s0) * D; i += 256) {
      float s = 0.f;
#pragma unroll
      for (int w = 0; w < 8; ++w) s += (&red[w][0][0])[i];
      atomicAdd(datt + ((size_t)b * att_stride_rows + s0) * D + i, s);
    }
    __syncthreads();
  }
}

int posterior_bwd(const float* dpred, const float* emb, int ld_emb, const float* att, int att_stride_rows, int B,
                  int Tmax, int D, int n_att, float* demb, int ld_demb, float* datt, cudaStream_t st) {
  EDA_REQUIRE(D == 256, "eda_posterior_bwd: D must be 256 (got %d)", D);
  EDA_REQUIRE(n_att >= 1 && n_att <= 8, "eda_posterior_bwd: n_att must be in [1,8] (got %d)", n_att);
  EDA_REQUIRE(dpred && emb && att && demb && datt && B > 0 && B <= 65535 && Tmax > 0, "eda_posterior_bwd: bad arguments");
  EDA_REQUIRE(ld_emb % 4 == 0 && ld_demb % 4 == 0 && ((uintptr_t)emb | (uintptr_t)demb) % 16 == 0, "eda_posterior_bwd: emb / demb must be 16B aligned");
  // a handful of CTAs per sequence keeps the atomic fan-in (and its rounding-order noise) small
  dim3 grid(std::max(1, std::min(ceil_div(Tmax, 64), ceil_div(num_sms() * 4, B))), B);
  posterior_bwd_kernel<8><<<grid, 256, 0, st>>>(dpred, emb, ld_emb, att, att_stride_rows, Tmax, n_att, demb, ld_demb, datt);
  EDA_LAUNCH_CHECK();
  return EDA_OK;
}

// ------------------------------------------------------------ attractor head backward -------
// loss_att = mean_b mean_n BCEwithlogits(p_bn, y_n), p = sigmoid(att . w + bias)   (models.py:350-371)
// d/dz = (sigmoid(p) - y) * p (1 - p) * scale / (B * n_dec).  One CTA, thread = column d, rows in order.
__global__ void __launch_bounds__(256)
attractor_head_bwd_kernel(const float* __restrict__ att, const float* __restrict__ probs, const float* __restrict__ w,
                          int B, int n_dec, int n_spk_label, float scale, float* __restrict__ datt,
                          float* __restrict__ dw, float* __restrict__ dbias, int accumulate) {
  const int d = threadIdx.x;
  const float wd = __ldg(w + d);
  float a_w = 0.f, a_b = 0.f;
  const float k = scale / (float)(B * n_dec);
  for (int r = 0; r < B * n_dec; ++r) {
    const float p = __ldg(probs + r), yv = (r % n_dec) < n_spk_label ? 1.f : 0.f;
    const float dz = (sigmoidf_ref(p) - yv) * p * (1.f - p) * k;
    datt[(size_t)r * 256 + d] += dz * wd;
    a_w = fmaf(dz, __ldg(att + (size_t)r * 256 + d), a_w);
    a_b += dz;
  }
  dw[d] = (accumulate ? dw[d] : 0.f) + a_w;
  if (d == 0) dbias[0] = (accumulate ? dbias[0] : 0.f) + a_b;
}

int attractor_head_bwd(const float* att, const float* probs, const float* w, int B, int n_dec, int D, int n_spk_label,
                       float scale, float* datt, float* dw, float* dbias, int accumulate, cudaStream_t st) {
  EDA_REQUIRE(D == 256, "eda_attractor_head_bwd: D must be 256 (got %d)", D);
  EDA_REQUIRE(att && probs && w && datt && dw && dbias && B > 0 && n_dec > 0, "eda_attractor_head_bwd: bad arguments");
  attractor_head_bwd_kernel<<<1, 256, 0, st>>>(att, probs, w, B, n_dec, n_spk_label, scale, datt, dw, dbias, accumulate);
  EDA_LAUNCH_CHECK();
  return EDA_OK;
}

// ------------------------------------------------------------ attention row statistic --------
// delta[row][h] = sum_d dO[row][64h + d] * O[row][64h + d]  (both as split-bf16 pairs); warp per row.
__global__ void __launch_bounds__(256)
mha_delta_kernel(const bf16* __restrict__ do_hi, const bf16* __restrict__ do_lo, int ld_do,
                 const bf16* __restrict__ o_hi, const bf16* __restrict__ o_lo, int ld_o, int M, float* __restrict__ delta) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int row = blockIdx.x * 8 + warp; row < M; row += gridDim.x * 8) {
    const uint4 a = __ldg(reinterpret_cast<const uint4*>(do_hi + (size_t)row * ld_do) + lane);
    const uint4 b = __ldg(reinterpret_cast<const uint4*>(do_lo + (size_t)row * ld_do) + lane);
    const uint4 c = __ldg(reinterpret_cast<const uint4*>(o_hi + (size_t)row * ld_o) + lane);
    const uint4 d = __ldg(reinterpret_cast<const uint4*>(o_lo + (size_t)row * ld_o) + lane);
    const uint32_t aw[4] = {a.x, a.y, a.z, a.w}, bw[4] = {b.x, b.y, b.z, b.w}, cw[4] = {c.x, c.y, c.z, c.w}, dw[4] = {d.x, d.y, d.z, d.w};
    float s = 0.f;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float g0 = __uint_as_float(aw[j] << 16) + __uint_as_float(bw[j] << 16);
      const float g1 = __uint_as_float(aw[j] & 0xffff0000u) + __uint_as_float(bw[j] & 0xffff0000u);
      const float o0 = __uint_as_float(cw[j] << 16) + __uint_as_float(dw[j] << 16);
      const float o1 = __uint_as_float(cw[j] & 0xffff0000u) + __uint_as_float(dw[j] & 0xffff0000u);
      s = fmaf(g0, o0, s); s = fmaf(g1, o1, s);
    }
    s += __shfl_xor_sync(0xffffffffu, s, 1); s += __shfl_xor_sync(0xffffffffu, s, 2); s += __shfl_xor_sync(0xffffffffu, s, 4);
    if ((lane & 7) == 0) delta[(size_t)row * 4 + (lane >> 3)] = s;     // 8 lanes x 8 columns = one head
  }
}

int mha_delta(const void* do_hi, const void* do_lo, int ld_do, const void* o_hi, const void* o_lo, int ld_o, int M,
              int D, float* delta, cudaStream_t st) {
  EDA_REQUIRE(D == 256, "eda_mha_delta: D must be 256 (4 heads of 64)");
  EDA_REQUIRE(do_hi && do_lo && o_hi && o_lo && delta && M > 0 && ld_do % 8 == 0 && ld_o % 8 == 0, "eda_mha_delta: bad arguments");
  mha_delta_kernel<<<std::min(ceil_div(M, 8), num_sms() * 8), 256, 0, st>>>((const bf16*)do_hi, (const bf16*)do_lo, ld_do,
                                                                           (const bf16*)o_hi, (const bf16*)o_lo, ld_o, M, delta);
  EDA_LAUNCH_CHECK();
  return EDA_OK;
}

// fp32 (rows, cols) * scale -> split pair written at a column offset of a wider split matrix
__global__ void scale_split_kernel(const float* __restrict__ src, int ld, int rows, int cols, float scale,
                                   bf16* __restrict__ hi, bf16* __restrict__ lo, int ld_s) {
  const int64_t n = (int64_t)rows * (cols / 2);
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int r = (int)(i / (cols / 2)), c = (int)(i % (cols / 2)) * 2;
    const float2 v = *reinterpret_cast<const float2*>(src + (size_t)r * ld + c);
    split_store2(hi + (size_t)r * ld_s + c, lo + (size_t)r * ld_s + c, v.x * scale, v.y * scale);
  }
}

int scale_split(const float* src, int ld, int rows, int cols, float scale, void* hi, void* lo, int ld_s, cudaStream_t st) {
  EDA_REQUIRE(src && hi && lo && rows > 0 && cols > 0 && cols % 2 == 0 && ld % 2 == 0 && ld_s % 2 == 0, "eda_scale_split: bad arguments");
  const int64_t n = (int64_t)rows * (cols / 2);
  scale_split_kernel<<<(int)std::min<int64_t>(ceil_div<int64_t>(n, 256), num_sms() * 16), 256, 0, st>>>(
      src, ld, rows, cols, scale, (bf16*)hi, (bf16*)lo, ld_s);
  EDA_LAUNCH_CHECK();
  return EDA_OK;
}

// keep[row][col] of one dropout site (test / debugging aid: the kernels never materialise a mask)
__global__ void dropout_mask_kernel(const DropoutCfg d, int64_t rows, int cols, uint8_t* __restrict__ keep) {
  const int64_t n = rows * (int64_t)(cols / 8);
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = i / (cols / 8);
    const int g = (int)(i % (cols / 8));
    const uint32_t k = d.thresh ? dropout_keep8(d, (uint32_t)r, (uint32_t)g) : 0xffu;
#pragma unroll
    for (int e = 0; e < 8; ++e) keep[r * cols + 8 * g + e] = (k >> e) & 1u;
  }
}

// The same mask as bits: word w of a row holds columns 32w .. 32w+31 (bit e = column 32w + e kept).  The
// attention kernels read one word per thread and 32 keys instead of evaluating Philox in their softmax warps.
__global__ void dropout_bits_kernel(const DropoutCfg d, int64_t rows, int words, int ld, uint32_t* __restrict__ bits) {
  const int64_t n = rows * (int64_t)words;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = i / words;
    const int w = (int)(i % words);
    uint32_t k = 0;
#pragma unroll
    for (int c = 0; c < 4; ++c) k |= dropout_keep8(d, (uint32_t)r, (uint32_t)(4 * w + c)) << (8 * c);
    bits[r * ld + w] = k;
  }
}
int dropout_bits(float p, uint64_t seed, int site, int64_t rows, int cols, uint32_t* bits, int ld_words, cudaStream_t st) {
  const int words = ceil_div(cols, 32);
  EDA_REQUIRE(bits && rows > 0 && cols > 0 && ld_words >= words && p > 0.f && p < 1.f, "eda_dropout_bits: bad arguments");
  const int64_t n = rows * (int64_t)words;
  dropout_bits_kernel<<<(int)std::min<int64_t>(ceil_div<int64_t>(n, 256), num_sms() * 16), 256, 0, st>>>(make_dropout(p, seed, site), rows, words, ld_words, bits);
  EDA_LAUNCH_CHECK();
  count_launch();
  return EDA_OK;
}

int dropout_mask(float p, uint64_t seed, int site, int64_t rows, int cols, uint8_t* keep, cudaStream_t st) {
  EDA_REQUIRE(keep && rows > 0 && cols > 0 && cols % 8 == 0 && p >= 0.f && p < 1.f, "eda_dropout_mask: bad arguments (cols %% 8 == 0)");
  const int64_t n = rows * (int64_t)(cols / 8);
  dropout_mask_kernel<<<(int)std::min<int64_t>(ceil_div<int64_t>(n, 256), num_sms() * 16), 256, 0, st>>>(make_dropout(p, seed, site), rows, cols, keep);
  EDA_LAUNCH_CHECK();
  return EDA_OK;
}

// ------------------------------------------------------------ gradient scaling, Adam ---------
__global__ void scale_dev_kernel(const float* __restrict__ scale, float* __restrict__ x, int64_t n) {
  const float s = __ldg(scale);
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) x[i] *= s;
}

int scale_dev(const float* scale, float* x, int64_t n, cudaStream_t st) {
  EDA_REQUIRE(scale && x && n > 0, "eda_scale_dev: bad arguments");
  scale_dev_kernel<<<(int)std::min<int64_t>(ceil_div<int64_t>(n, 1024), num_sms() * 8), 256, 0, st>>>(scale, x, n);
  EDA_LAUNCH_CHECK();
  return EDA_OK;
}

// torch.optim.Adam (no weight decay, no amsgrad) on one flat parameter buffer:
//   m = b1 m + (1 - b1) g;  v = b2 v + (1 - b2) g^2
//   p -= step_size * m / (sqrt(v) / bias2_sqrt + eps)     step_size = lr / (1 - b1^t), bias2_sqrt = sqrt(1 - b2^t)
__global__ void adam_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
                            float* __restrict__ v, int64_t n, float b1, float b2, float eps, float step_size,
                            float bias2_sqrt) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const float gi = g[i];
    const float mi = b1 * m[i] + (1.f - b1) * gi;
    const float vi = b2 * v[i] + (1.f - b2) * gi * gi;
    m[i] = mi; v[i] = vi;
    p[i] -= step_size * mi / (sqrtf(vi) / bias2_sqrt + eps);
  }
}

int adam_step(float* p, const float* g, float* m, float* v, int64_t n, float lr, float b1, float b2, float eps,
              int step, cudaStream_t st) {
  EDA_REQUIRE(p && g && m && v && n > 0 && step >= 1, "eda_adam_step: bad arguments");
  const double bc1 = 1.0 - pow((double)b1, step), bc2 = 1.0 - pow((double)b2, step);
  adam_kernel<<<(int)std::min<int64_t>(ceil_div<int64_t>(n, 1024), num_sms() * 8), 256, 0, st>>>(
      p, g, m, v, n, b1, b2, eps, (float)(lr / bc1), (float)sqrt(bc2));
  EDA_LAUNCH_CHECK();
  return EDA_OK;
}

}  // namespace eda

extern "C" {

int eda_ln_bwd(const float* dy, int ld_dy, const float* y, int ld_y, const float* gamma, const float* beta,
               const float* rstd, int M, int D, float* dx, int ld_dx, void* dx_hi, void* dx_lo, int ld_dxs,
               float* dgamma, float* dbeta, float* dcolsum, void* workspace, int64_t workspace_bytes, int accumulate,
               float p_drop, uint64_t seed, int site, void* stream) {
  EDA_REQUIRE(p_drop >= 0.f && p_drop < 1.f, "eda_ln_bwd: bad dropout probability");
  return eda::ln_bwd(dy, ld_dy, y, ld_y, gamma, beta, rstd, M, D, dx, ld_dx, dx_hi, dx_lo, ld_dxs, dgamma, dbeta,
                     dcolsum, (float*)workspace, (size_t)workspace_bytes, accumulate, eda::make_dropout(p_drop, seed, site),
                     (cudaStream_t)stream);
}

int eda_dropout_bits(float p_drop, uint64_t seed, int site, int64_t rows, int cols, uint32_t* bits, int ld_words, void* stream) {
  return eda::dropout_bits(p_drop, seed, site, rows, cols, bits, ld_words, (cudaStream_t)stream);
}
int eda_dropout_mask(float p_drop, uint64_t seed, int site, int64_t rows, int cols, uint8_t* keep, void* stream) {
  return eda::dropout_mask(p_drop, seed, site, rows, cols, keep, (cudaStream_t)stream);
}

int eda_colsum_split(const void* hi, const void* lo, int M, int N, int ld, float* out, void* workspace,
                     int64_t workspace_bytes, int accumulate, void* stream) {
  return eda::colsum_split(hi, lo, M, N, ld, out, (float*)workspace, (size_t)workspace_bytes, accumulate, (cudaStream_t)stream);
}

int eda_posterior_bwd(const float* dpred, const float* emb, int ld_emb, const float* attractors, int att_stride_rows,
                      int B, int Tmax, int D, int n_att, float* demb, int ld_demb, float* dattractors, void* stream) {
  return eda::posterior_bwd(dpred, emb, ld_emb, attractors, att_stride_rows, B, Tmax, D, n_att, demb, ld_demb,
                            dattractors, (cudaStream_t)stream);
}

int eda_attractor_head_bwd(const float* attractors, const float* probs, const float* w, int B, int n_dec, int D,
                           int n_spk_label, float scale, float* dattractors, float* dw, float* dbias, int accumulate,
                           void* stream) {
  return eda::attractor_head_bwd(attractors, probs, w, B, n_dec, D, n_spk_label, scale, dattractors, dw, dbias,
                                 accumulate, (cudaStream_t)stream);
}

int eda_mha_delta(const void* do_hi, const void* do_lo, int ld_do, const void* o_hi, const void* o_lo, int ld_o,
                  int M, int D, float* delta, void* stream) {
  return eda::mha_delta(do_hi, do_lo, ld_do, o_hi, o_lo, ld_o, M, D, delta, (cudaStream_t)stream);
}

int eda_scale_split(const float* src, int ld, int rows, int cols, float scale, void* hi, void* lo, int ld_s, void* stream) {
  return eda::scale_split(src, ld, rows, cols, scale, hi, lo, ld_s, (cudaStream_t)stream);
}

int eda_scale_dev(const float* scale_dev, float* x, int64_t n, void* stream) {
  return eda::scale_dev(scale_dev, x, n, (cudaStream_t)stream);
}

int eda_adam_step(float* params, const float* grads, float* exp_avg, float* exp_avg_sq, int64_t n, float lr,
                  float beta1, float beta2, float eps, int step, void* stream) {
  return eda::adam_step(params, grads, exp_avg, exp_avg_sq, n, lr, beta1, beta2, eps, step, (cudaStream_t)stream);
}

}  // extern "C"
