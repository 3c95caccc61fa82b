// Bandwidth-bound glue kernels of the model path: pad/pack, attractor existence head,
// fused posterior (bmm + sigmoid + threshold).
#include <algorithm>

#include <cuda_bf16.h>

#include "common.cuh"

namespace eda {

// ---------------------------------------------------------------- pad_pack ----------
// models.py:217 pad_sequence(padding_value=-1) into a K-padded staging matrix.
__global__ void pad_pack_kernel(const float* __restrict__ src, const int64_t* __restrict__ seq_off,
                                const int32_t* __restrict__ lens, int Tmax, int n_cols, float pad,
                                float* __restrict__ dst, int ld_dst, int64_t n_rows) {
  const int warps_per_block = blockDim.x >> 5;
  const int lane = threadIdx.x & 31;
  for (int64_t row = (int64_t)blockIdx.x * warps_per_block + (threadIdx.x >> 5); row < n_rows;
       row += (int64_t)gridDim.x * warps_per_block) {
    const int b = (int)(row / Tmax), t = (int)(row % Tmax);
    const bool valid = t < lens[b];
    const float* s = src + (seq_off[b] + t) * (int64_t)n_cols;
    float* d = dst + row * (int64_t)ld_dst;
    for (int c = lane; c < ld_dst; c += 32) d[c] = c < n_cols ? (valid ? __ldg(s + c) : pad) : 0.f;
  }
}

int pad_pack(const float* src, const int64_t* seq_off, const int32_t* lens, int B, int Tmax,
             int n_cols, float pad, float* dst, int ld_dst, cudaStream_t st) {
  EDA_REQUIRE(ld_dst >= n_cols && B > 0 && Tmax > 0, "eda_pad_pack: bad sizes");
  const int64_t n_rows = (int64_t)B * Tmax;
  const int blocks = (int)std::min<int64_t>(ceil_div<int64_t>(n_rows, 8), (int64_t)num_sms() * 8);
  pad_pack_kernel<<<blocks, 256, 0, st>>>(src, seq_off, lens, Tmax, n_cols, pad, dst, ld_dst, n_rows);
  EDA_LAUNCH_CHECK();
  return EDA_OK;
}

// pad_sequence(-1) straight into the split-bf16 operand layout of the tensor-core GEMMs:
// hi = bf16(x), lo = bf16(x - hi); columns [n_cols, ld_dst) are zero.  One bf16x2 per thread.
__global__ void pad_pack_split_kernel(const float* __restrict__ src, const int64_t* __restrict__ seq_off,
                                      const int32_t* __restrict__ lens, int Tmax, int n_cols, int ld_src, float pad,
                                      __nv_bfloat16* __restrict__ hi, __nv_bfloat16* __restrict__ lo, int ld_dst,
                                      int64_t n_rows) {
  const int warps_per_block = blockDim.x >> 5;
  const int lane = threadIdx.x & 31;
  for (int64_t row = (int64_t)blockIdx.x * warps_per_block + (threadIdx.x >> 5); row < n_rows;
       row += (int64_t)gridDim.x * warps_per_block) {
    const int b = (int)(row / Tmax), t = (int)(row % Tmax);
    const bool valid = t < lens[b];
    const float* s = src + (seq_off[b] + t) * (int64_t)ld_src;
    for (int c = lane * 2; c < ld_dst; c += 64) {
      const float x0 = c < n_cols ? (valid ? __ldg(s + c) : pad) : 0.f;
      const float x1 = c + 1 < n_cols ? (valid ? __ldg(s + c + 1) : pad) : 0.f;
      const __nv_bfloat16 h0 = __float2bfloat16_rn(x0), h1 = __float2bfloat16_rn(x1);
      *reinterpret_cast<__nv_bfloat162*>(hi + row * ld_dst + c) = __halves2bfloat162(h0, h1);
      *reinterpret_cast<__nv_bfloat162*>(lo + row * ld_dst + c) = __halves2bfloat162(
          __float2bfloat16_rn(x0 - __bfloat162float(h0)), __float2bfloat16_rn(x1 - __bfloat162float(h1)));
    }
  }
}

int pad_pack_split(const float* src, int ld_src, const int64_t* seq_off, const int32_t* lens, int B, int Tmax,
                   int n_cols, float pad, void* hi, void* lo, int ld_dst, cudaStream_t st) {
  EDA_REQUIRE(ld_dst >= n_cols && ld_dst % 8 == 0 && ld_src >= n_cols && B > 0 && Tmax > 0, "eda_pad_pack_split: bad sizes");
  const int64_t n_rows = (int64_t)B * Tmax;
  const int blocks = (int)std::min<int64_t>(ceil_div<int64_t>(n_rows, 8), (int64_t)num_sms() * 8);
  pad_pack_split_kernel<<<blocks, 256, 0, st>>>(src, seq_off, lens, Tmax, n_cols, ld_src, pad,
                                                (__nv_bfloat16*)hi, (__nv_bfloat16*)lo, ld_dst, n_rows);
  EDA_LAUNCH_CHECK();
  return EDA_OK;
}

// ---------------------------------------------------------- attractor head ----------
// encoder_decoder_attractor.py:50-54 and models.py:350-371.  Warp per (b, n) over the whole
// grid; the attractor loss is a second one-CTA kernel (it needs every probability).
__global__ void __launch_bounds__(256)
attractor_head_kernel(const float* __restrict__ att, const float* __restrict__ w,
                      const float* __restrict__ bias, int total, int D, float* __restrict__ probs) {
  const int lane = threadIdx.x & 31;
  const int i = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (i >= total) return;
  const float* a = att + (size_t)i * D;
  float s = 0.f;
  if (D % 4 == 0 && (((uintptr_t)a | (uintptr_t)w) & 15) == 0) {
    const float4* a4 = reinterpret_cast<const float4*>(a);
    const float4* w4 = reinterpret_cast<const float4*>(w);
    for (int d = lane; d < D / 4; d += 32) {
      const float4 x = __ldg(a4 + d), y = __ldg(w4 + d);
      s = fmaf(x.x, y.x, s); s = fmaf(x.y, y.y, s); s = fmaf(x.z, y.z, s); s = fmaf(x.w, y.w, s);
    }
  } else {
    for (int d = lane; d < D; d += 32) s = fmaf(__ldg(a + d), __ldg(w + d), s);
  }
  s = warp_sum(s);
  if (lane == 0) probs[i] = sigmoidf_ref(s + __ldg(bias));
}

// BCE-with-logits on the PROBABILITIES (reference behaviour).  Row sums in parallel (thread = b,
// n ascending), then ONE thread adds the rows in ascending b: a fixed summation order.
__global__ void __launch_bounds__(1024)
attractor_loss_kernel(const float* __restrict__ probs, int B, int n_dec, int n_spk_label,
                      float* __restrict__ loss_att) {
  __shared__ float rows[1024];
  float total_loss = 0.f;
  for (int b0 = 0; b0 < B; b0 += 1024) {
    const int b = b0 + threadIdx.x;
    if (b < B) {
      float row = 0.f;
      for (int n = 0; n < n_dec; ++n) {
        const float x = probs[b * n_dec + n], y = n < n_spk_label ? 1.f : 0.f;
        row += fmaxf(x, 0.f) - x * y + log1pf(expf(-fabsf(x)));
      }
      rows[threadIdx.x] = row / (float)n_dec;
    }
    __syncthreads();
    if (threadIdx.x == 0)
      for (int j = 0; j < min(1024, B - b0); ++j) total_loss += rows[j];
    __syncthreads();
  }
  if (threadIdx.x == 0) loss_att[0] = total_loss / (float)B;
}

int attractor_head(const float* att, const float* w, const float* bias, int B, int n_dec, int D,
                   float* probs, int n_spk_label, float* loss_att, cudaStream_t st) {
  EDA_REQUIRE(B > 0 && n_dec > 0 && D > 0, "eda_attractor_head: bad sizes");
  const int total = B * n_dec;
  attractor_head_kernel<<<ceil_div(total, 8), 256, 0, st>>>(att, w, bias, total, D, probs);
  EDA_LAUNCH_CHECK();
  if (n_spk_label >= 0 && loss_att != nullptr) {
    attractor_loss_kernel<<<1, 1024, 0, st>>>(probs, B, n_dec, n_spk_label, loss_att);
    EDA_LAUNCH_CHECK();
  }
  return EDA_OK;
}

// --------------------------------------------------------------- posterior ----------
// models.py:329 bmm(emb, attractor[:, :-1]^T) -> :333 sigmoid -> make_rttm.py:29 threshold.
// Warp per frame: the 256-d embedding row is read once (two float4 per lane, coalesced),
// attractors of the sequence sit in shared memory, the n_att dot products are reduced with
// a shuffle butterfly, lane s writes column s.
template <int NATT>
__global__ void __launch_bounds__(256)
posterior_kernel(const float* __restrict__ emb, int ld_emb, const float* __restrict__ att,
                 int att_stride_rows, int Tmax, int n_att, float thr, float* __restrict__ logits,
                 float* __restrict__ post, uint8_t* __restrict__ labels,
                 const int64_t* __restrict__ out_off, const int32_t* __restrict__ lens, uint16_t* __restrict__ mask16) {
  constexpr int D = 256;
  __shared__ __align__(16) float As[NATT][D];
  const int b = blockIdx.y;
  for (int i = threadIdx.x; i < n_att * D; i += blockDim.x)
    As[i / D][i % D] = __ldg(att + (size_t)b * att_stride_rows * D + i);
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  // packed output (eda_posterior_packed): frame t of sequence b goes to row out_off[b] + t, padding is skipped
  const int t_end = out_off ? min(Tmax, __ldg(lens + b)) : Tmax;
  for (int t = blockIdx.x * 8 + warp; t < t_end; t += gridDim.x * 8) {
    const size_t row = (size_t)b * Tmax + t;
    const float4* e = reinterpret_cast<const float4*>(emb + row * ld_emb);
    const float4 e0 = __ldg(e + lane), e1 = __ldg(e + 32 + lane);
    float mine = 0.f;
#pragma unroll
    for (int s = 0; s < NATT; ++s) {
      if (s < n_att) {
        const float4 a0 = *reinterpret_cast<const float4*>(&As[s][lane * 4]);
        const float4 a1 = *reinterpret_cast<const float4*>(&As[s][128 + lane * 4]);
        float v = e0.x * a0.x;
        v = fmaf(e0.y, a0.y, v); v = fmaf(e0.z, a0.z, v); v = fmaf(e0.w, a0.w, v);
        v = fmaf(e1.x, a1.x, v); v = fmaf(e1.y, a1.y, v); v = fmaf(e1.z, a1.z, v); v = fmaf(e1.w, a1.w, v);
        v = warp_sum(v);
        if (lane == s) mine = v;
      }
    }
    if (lane < n_att) {
      const size_t o = (out_off ? (size_t)(__ldg(out_off + b) + t) : row) * n_att + lane;
      const float p = sigmoidf_ref(mine);
      if (logits) logits[o] = mine;
      if (post) post[o] = p;
      if (labels) labels[o] = p > thr ? 1 : 0;
    }
    if (mask16) {      // the frame's decisions as one bit per attractor (bit s = speaker s active): 2 bytes instead of 4 n_att
      const unsigned m = __ballot_sync(0xffffffffu, lane < n_att && sigmoidf_ref(mine) > thr);
      if (lane == 0) mask16[out_off ? (size_t)(__ldg(out_off + b) + t) : row] = (uint16_t)m;
    }
  }
}

int posterior(const float* emb, int ld_emb, const float* att, int att_stride_rows, int B, int Tmax,
              int D, int n_att, float thr, float* logits, float* post, uint8_t* labels,
              cudaStream_t st, const int64_t* out_off, const int32_t* lens, uint16_t* mask16) {
  EDA_REQUIRE(D == 256, "eda_posterior: D must be 256 (got %d)", D);
  EDA_REQUIRE(n_att >= 1 && n_att <= 16, "eda_posterior: n_att must be in [1,16] (got %d)", n_att);
  EDA_REQUIRE(ld_emb % 4 == 0 && (uintptr_t)emb % 16 == 0, "eda_posterior: emb must be 16B aligned");
  EDA_REQUIRE(B <= 65535, "eda_posterior: B too large");
  dim3 grid(max(1, min(ceil_div(Tmax, 8), ceil_div(num_sms() * 8, B))), B);
  EDA_REQUIRE((out_off == nullptr) == (lens == nullptr), "eda_posterior_packed: out_off and lens go together");
  posterior_kernel<16><<<grid, 256, 0, st>>>(emb, ld_emb, att, att_stride_rows, Tmax, n_att, thr, logits, post, labels,
                                             out_off, lens, mask16);
  EDA_LAUNCH_CHECK();
  return EDA_OK;
}

}  // namespace eda
