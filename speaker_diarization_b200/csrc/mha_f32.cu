// Multi-head self-attention forward, fp32 CUDA-core path (EDA_PREC_FP32).
//
// softmax(q k^T / sqrt(64)) v per (batch, head) with NO mask (the reference attends to the
// -1-padded frames too, models.py:223), or with per-sequence key counts (eda_mha_fwd_masked: a batch
// of chunks in which every sequence sees only its own frames, as in the reference's one-chunk-per-call
// loop, infer_eda.py:102-113).  Flash-style: one CTA owns 64 queries of one
// (b, h), streams 64-key tiles, keeps the running max / sum / output in registers.
// 256 threads as 16 x 16: thread (ty, tx) owns score rows ty*4+i, score columns tx+16*j
// (interleaved, so LDS.128 along k is conflict-free with a 68-float row stride) and output
// columns tx*4..tx*4+3.
#include "common.cuh"

namespace eda {

constexpr int kDh = 64;
constexpr int kLd = 68;  // smem row stride in floats

__global__ void __launch_bounds__(256, 2)
mha_f32_kernel(const float* __restrict__ qkv, int ld_qkv, float* __restrict__ out, int ld_out,
               int T, int D, const int32_t* __restrict__ key_lens) {
  extern __shared__ __align__(16) float smem[];
  float(*Qs)[kLd] = reinterpret_cast<float(*)[kLd]>(smem);
  float(*Ks)[kLd] = reinterpret_cast<float(*)[kLd]>(smem + 64 * kLd);
  float(*Vs)[kLd] = reinterpret_cast<float(*)[kLd]>(smem + 2 * 64 * kLd);
  float(*Ps)[kLd] = reinterpret_cast<float(*)[kLd]>(smem + 3 * 64 * kLd);

  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int q0 = blockIdx.x * 64, h = blockIdx.y, b = blockIdx.z;
  const float* base = qkv + (size_t)b * T * ld_qkv + h * kDh;
  const int klen = key_lens ? min(T, max(1, __ldg(key_lens + b))) : T;     // keys >= klen take no part

  // Q tile, pre-scaled by 1/sqrt(64) = 0.125 (exact in fp32)
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    int f = tid + i * 256, r = f >> 4, c = (f & 15) * 4;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (q0 + r < T) v = __ldg(reinterpret_cast<const float4*>(base + (size_t)(q0 + r) * ld_qkv + c));
    v.x *= 0.125f; v.y *= 0.125f; v.z *= 0.125f; v.w *= 0.125f;
    *reinterpret_cast<float4*>(&Qs[r][c]) = v;
  }

  float m[4], l[4], o[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    m[i] = -INFINITY; l[i] = 0.f;
#pragma unroll
    for (int j = 0; j < 4; ++j) o[i][j] = 0.f;
  }

  for (int k0 = 0; k0 < klen; k0 += 64) {
    __syncthreads();  // previous tile fully consumed (also orders the Q stores on the first pass)
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      int f = tid + i * 256, r = f >> 4, c = (f & 15) * 4;
      float4 kv = make_float4(0.f, 0.f, 0.f, 0.f), vv = kv;
      if (k0 + r < klen) {
        const float* p = base + (size_t)(k0 + r) * ld_qkv + c;
        kv = __ldg(reinterpret_cast<const float4*>(p + D));
        vv = __ldg(reinterpret_cast<const float4*>(p + 2 * D));
      }
      *reinterpret_cast<float4*>(&Ks[r][c]) = kv;
      *reinterpret_cast<float4*>(&Vs[r][c]) = vv;
    }
    __syncthreads();

    float s[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) s[i][j] = 0.f;
#pragma unroll 4
    for (int k = 0; k < kDh; k += 4) {
      float4 a[4], bb[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) a[i] = *reinterpret_cast<const float4*>(&Qs[ty * 4 + i][k]);
#pragma unroll
      for (int j = 0; j < 4; ++j) bb[j] = *reinterpret_cast<const float4*>(&Ks[tx + 16 * j][k]);
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          s[i][j] = fmaf(a[i].x, bb[j].x, s[i][j]);
          s[i][j] = fmaf(a[i].y, bb[j].y, s[i][j]);
          s[i][j] = fmaf(a[i].z, bb[j].z, s[i][j]);
          s[i][j] = fmaf(a[i].w, bb[j].w, s[i][j]);
        }
    }
    // mask keys beyond the sequence, online softmax update
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      float mx = -INFINITY;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        if (k0 + tx + 16 * j >= klen) s[i][j] = -INFINITY;
        mx = fmaxf(mx, s[i][j]);
      }
#pragma unroll
      for (int off = 8; off > 0; off >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, off));
      const float mn = fmaxf(m[i], mx);          // finite: key k0 (< klen) is always valid
      const float sc = expf(m[i] - mn);          // exp(-inf) = 0 on the first tile
      float ps = 0.f;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        float p = expf(s[i][j] - mn);
        ps += p;
        Ps[ty * 4 + i][tx + 16 * j] = p;
      }
      l[i] = l[i] * sc + ps;
      m[i] = mn;
#pragma unroll
      for (int j = 0; j < 4; ++j) o[i][j] *= sc;
    }
    __syncthreads();
#pragma unroll 4
    for (int k = 0; k < 64; k += 4) {
      float4 a[4], bb[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) a[i] = *reinterpret_cast<const float4*>(&Ps[ty * 4 + i][k]);
#pragma unroll
      for (int kk = 0; kk < 4; ++kk) bb[kk] = *reinterpret_cast<const float4*>(&Vs[k + kk][tx * 4]);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float av[4] = {a[i].x, a[i].y, a[i].z, a[i].w};
#pragma unroll
        for (int kk = 0; kk < 4; ++kk) {
          o[i][0] = fmaf(av[kk], bb[kk].x, o[i][0]);
          o[i][1] = fmaf(av[kk], bb[kk].y, o[i][1]);
          o[i][2] = fmaf(av[kk], bb[kk].z, o[i][2]);
          o[i][3] = fmaf(av[kk], bb[kk].w, o[i][3]);
        }
      }
    }
  }

#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float lt = l[i];
#pragma unroll
    for (int off = 8; off > 0; off >>= 1) lt += __shfl_xor_sync(0xffffffffu, lt, off);
    const int q = q0 + ty * 4 + i;
    if (q < T) {
      const float inv = 1.0f / lt;
      *reinterpret_cast<float4*>(out + ((size_t)b * T + q) * ld_out + h * kDh + tx * 4) =
          make_float4(o[i][0] * inv, o[i][1] * inv, o[i][2] * inv, o[i][3] * inv);
    }
  }
}

int mha_f32(const float* qkv, int ld_qkv, float* out, int ld_out, int B, int T, int n_heads,
            cudaStream_t st, const int32_t* key_lens) {
  const int D = n_heads * kDh;
  EDA_REQUIRE(ld_qkv % 4 == 0 && ld_out % 4 == 0, "eda_mha_fwd: strides must be multiples of 4");
  EDA_REQUIRE(((uintptr_t)qkv | (uintptr_t)out) % 16 == 0, "eda_mha_fwd: pointers must be 16B aligned");
  EDA_REQUIRE(B <= 65535 && n_heads <= 65535, "eda_mha_fwd: B/n_heads too large for the grid");
  static DeviceOnce attr_set;
  const int smem = 4 * 64 * kLd * (int)sizeof(float);
  if (!attr_set.done()) {
    EDA_CUDA(cudaFuncSetAttribute(mha_f32_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    attr_set.mark();
  }
  dim3 grid(ceil_div(T, 64), n_heads, B);
  mha_f32_kernel<<<grid, 256, smem, st>>>(qkv, ld_qkv, out, ld_out, T, D, key_lens);
  EDA_LAUNCH_CHECK();
  return EDA_OK;
}

}  // namespace eda
