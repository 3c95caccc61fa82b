// nn.Linear with fused epilogues on the 5th-generation tensor cores (tcgen05 / TMEM / TMA).
//
//   C[m,n] = sum_k A[m,k] W[n,k]      fp32 accumulation in TMEM
//
// Operands live in HBM as bf16 "split" pairs: x = hi + lo, hi = bf16(x), lo = bf16(x - hi).
// EDA_PREC_BF16X3 issues three MMAs per k-step (Ahi.Bhi + Ahi.Blo + Alo.Bhi; the dropped
// lo.lo term is ~2^-18 relative), which is what the parity bar needs (one bf16
// pass cannot meet it).  EDA_PREC_BF16 issues only hi.hi.
//
// The producing kernel writes its output already split (and/or as fp32 for the residual
// stream), so no separate conversion pass touches HBM between two GEMMs.
//
// Kernel: persistent, one CTA per SM, 128 x 256 output tiles handed out round-robin.
//   warp 0      TMA producer   (SWIZZLE_128B K-major tiles {A_hi, A_lo, B_hi, B_lo}, 2 stages)
//   warp 1      MMA issuer (one thread) + TMEM owner: 2 accumulator buffers x 256 columns,
//               so the epilogue of tile i overlaps the MMAs of tile i+1
//   warps 2-9   epilogue (two warps per TMEM lane quadrant, 128 columns each): tcgen05.ld ->
//               bias / ReLU / residual + LayerNorm (row held in the registers of two threads,
//               statistics exchanged through smem) -> thread-per-row 32-byte stores (STG.256, one full
//               sector each) of the fp32 and/or split-bf16 output; the QKV variant pre-scales q by 1/sqrt(dh) and
//               stores v TRANSPOSED per (batch, head) for the attention kernel.
#include <algorithm>
#include <mutex>
#include <unordered_map>

#include "common.cuh"
#include "tc_common.cuh"

namespace eda {

using namespace tc;
typedef __nv_bfloat16 bf16;

constexpr int GBM = 128, GBN = 256, GBK = 64;
// Two shared-memory plans.  "Staged" (short K: the epilogue's stores dominate): 2 operand stages and
// 8 KB per epilogue warp from which results leave through TMA stores.  "Direct" (long K): 3
// operand stages, results stored straight from registers.
constexpr int G_STAGES_STAGED = 2, G_STAGES_DIRECT = 3;
constexpr int G_STG_WARP = 2 * 4096;                     // two 32-row x 128 B buffers per epilogue warp
constexpr int G_A_TILE = GBM * GBK * 2;                  // 16 KB: this CTA's 128 rows of A
constexpr int G_B_TILE = (GBN / 2) * GBK * 2;            // 16 KB: this CTA's HALF (128 rows) of the W tile
constexpr int G_STAGE = 2 * G_A_TILE + 2 * G_B_TILE;     // 64 KB (hi + lo of both)
constexpr int G_EPI_WARPS = 8;
constexpr int G_THREADS = 64 + 32 * G_EPI_WARPS;
constexpr int G_LNX_BYTES = 2 * 2 * GBM * 4;             // LayerNorm partial sums [pass][column half][row]
constexpr int G_VEC_MAX = 4096;                          // bias entries staged in smem (zero padded to a tile multiple)
constexpr int G_VEC_BYTES = (G_VEC_MAX + 2 * GBN) * 4;   // bias + LayerNorm gamma, beta
constexpr int G_OPERAND_BYTES = G_STAGES_DIRECT * G_STAGE;    // == 2 stages + 8 staging buffers of 8 KB
static_assert(G_STAGES_STAGED * G_STAGE + G_EPI_WARPS * G_STG_WARP == G_OPERAND_BYTES, "smem plans differ in size");
constexpr int G_SMEM = G_OPERAND_BYTES + G_LNX_BYTES + G_VEC_BYTES + 256 + 1024;

enum { GEPI_QKV = 4 };     // (EDA_EPI_BIAS_RES = 5 ... EDA_EPI_BIAS_GLU = 8 are public, eda_b200.h)

struct GemmParams {
  const float* bias;
  const float* resid; int ldr;
  const float* gamma; const float* beta;
  float* C; int ldc;                       // optional fp32 output
  bf16* Chi; bf16* Clo; int lds;           // optional split output
  int M, N, Kp, passes;
  // QKV variant: rows are (b, t) with t < T; v goes to vt[(b*H + h)*64 + d][t], row stride Tp
  bf16* vt_hi; bf16* vt_lo; int T, Tp, H; float q_scale;
  float out_scale;                         // residual epilogues: C = (A.W^T + bias) * out_scale + resid (0 means 1)
  DropoutCfg drop;                         // training: dropout on relu(A.W^T + bias) (BIAS_RELU) or on A.W^T + bias before
                                           // the residual is added (BIAS_RES_LN); thresh == 0 = off
  float mask_scale;                        // EDA_EPI_RELU_MASK: kept gradients are multiplied by this (0 means 1)
  float* rstd_out;                         // LayerNorm epilogues (training): 1/sqrt(var + eps) per row, optional
  const bf16* mask; int ldm;               // EDA_EPI_RELU_MASK: result is zeroed where mask <= 0
  // weight-gradient form (WG): C[i,j] = sum_r A[r,i] B[r,j] over Kp rows, both operands MN-major;
  // the reduction is cut into `ksplit` slices whose fp32 partial results are `c_slice` elements apart
  int ksplit; long long c_slice;
};

__device__ __forceinline__ void split2(float x0, float x1, __nv_bfloat162& hi, __nv_bfloat162& lo) {
  const bf16 h0 = __float2bfloat16_rn(x0), h1 = __float2bfloat16_rn(x1);
  hi = __halves2bfloat162(h0, h1);
  lo = __halves2bfloat162(__float2bfloat16_rn(x0 - __bfloat162float(h0)),
                          __float2bfloat16_rn(x1 - __bfloat162float(h1)));
}

// dropout on 32 consecutive columns col0.. (col0 % 8 == 0) of one row: 4 Philox calls
__device__ __forceinline__ void apply_dropout32(float (&v)[32], const DropoutCfg& d, int row, int col0) {
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const uint32_t keep = dropout_keep8(d, (uint32_t)row, (uint32_t)(col0 >> 3) + j);
#pragma unroll
    for (int e = 0; e < 8; ++e) v[8 * j + e] = (keep >> e) & 1u ? v[8 * j + e] * d.scale : 0.f;
  }
}

// one 32-byte (full sector) global store
__device__ __forceinline__ void st_global_v8(void* p, uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3,
                                             uint32_t a4, uint32_t a5, uint32_t a6, uint32_t a7) {
  asm volatile("st.global.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};"
               ::"l"(p), "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(a4), "r"(a5), "r"(a6), "r"(a7) : "memory");
}

// two fp32 -> packed bf16 pair (a in the low half), round to nearest even
__device__ __forceinline__ uint32_t pack_bf16x2(float a, float b) {
  uint32_t r;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(b), "f"(a));
  return r;
}
__device__ __forceinline__ void ld_global_nc_v8(const float* p, float4& a, float4& b) {
  asm volatile("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=f"(a.x), "=f"(a.y), "=f"(a.z), "=f"(a.w), "=f"(b.x), "=f"(b.y), "=f"(b.z), "=f"(b.w) : "l"(p));
}

// fp32 (rows x K, ld) -> hi, lo bf16 (rows x ldk), zero padded columns [K, ldk)
__global__ void split_bf16_kernel(const float* __restrict__ src, int ld, int rows, int K, int ldk,
                                  bf16* __restrict__ hi, bf16* __restrict__ lo) {
  const int64_t n = (int64_t)rows * (ldk / 2);
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int r = (int)(i / (ldk / 2)), c = (int)(i % (ldk / 2)) * 2;
    const float x0 = c < K ? src[(size_t)r * ld + c] : 0.f;
    const float x1 = c + 1 < K ? src[(size_t)r * ld + c + 1] : 0.f;
    __nv_bfloat162 h, l;
    split2(x0, x1, h, l);
    *reinterpret_cast<__nv_bfloat162*>(hi + (size_t)r * ldk + c) = h;
    *reinterpret_cast<__nv_bfloat162*>(lo + (size_t)r * ldk + c) = l;
  }
}

// EX: the launch uses the rarely needed epilogue extras (out_scale, dropout); they are compiled out of the
// plain instantiations so the inference kernels keep their register budget.
template <int EPI, bool ST, bool WG, bool EX>
__global__ void __launch_bounds__(G_THREADS, 1)
gemm_split_kernel(const __grid_constant__ CUtensorMap tmA_hi, const __grid_constant__ CUtensorMap tmA_lo,
                  const __grid_constant__ CUtensorMap tmB_hi, const __grid_constant__ CUtensorMap tmB_lo,
                  const __grid_constant__ CUtensorMap tmC, const __grid_constant__ CUtensorMap tmC_hi,
                  const __grid_constant__ CUtensorMap tmC_lo, const GemmParams p) {
  constexpr int GSTAGES = ST ? G_STAGES_STAGED : G_STAGES_DIRECT;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  float (*ln_x)[2][GBM] = reinterpret_cast<float (*)[2][GBM]>(smem + G_OPERAND_BYTES);
  // epilogue row vectors live in shared memory: global loads issued between the epilogue's
  // stores queue up behind them in the LSU
  float* s_bias = reinterpret_cast<float*>(smem + G_OPERAND_BYTES + G_LNX_BYTES);
  float* s_gamma = s_bias + G_VEC_MAX;
  float* s_beta = s_gamma + GBN;
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + G_OPERAND_BYTES + G_LNX_BYTES + G_VEC_BYTES);
  uint64_t* empty = full + GSTAGES;
  uint64_t* tfull = empty + GSTAGES;
  uint64_t* tempty = tfull + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();                     // 0 = leader (issues the MMAs of the pair)
  const int pair = blockIdx.x >> 1, n_pairs = gridDim.x >> 1;
  const int n_tiles_n = (p.N + GBN - 1) / GBN, n_tiles_m = (p.M + 2 * GBM - 1) / (2 * GBM);
  const int n_mn = n_tiles_n * n_tiles_m;                      // 256 x 256 output tiles, one per CTA pair
  const int nkb_all = (p.Kp + GBK - 1) / GBK;
  const int nkb_slice = WG ? (nkb_all + p.ksplit - 1) / p.ksplit : nkb_all;
  const int n_tiles = WG ? n_mn * p.ksplit : n_mn;             // (WG: every reduction slice of a tile is a work item)

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA_hi); tma_prefetch_desc(&tmA_lo); tma_prefetch_desc(&tmB_hi); tma_prefetch_desc(&tmB_lo);
    if (ST) { tma_prefetch_desc(&tmC); tma_prefetch_desc(&tmC_hi); tma_prefetch_desc(&tmC_lo); }
    for (int s = 0; s < GSTAGES; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
    for (int s = 0; s < 2; ++s) { mbar_init(&tfull[s], 1); mbar_init(&tempty[s], 2 * G_EPI_WARPS); }
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc_pair(tmem_slot, 512);
  const bool vec_smem = p.N <= G_VEC_MAX;
  if (vec_smem) {
    const int n_pad = n_tiles_n * GBN;
    for (int i = threadIdx.x; i < n_pad; i += G_THREADS) s_bias[i] = (p.bias && i < p.N) ? __ldg(p.bias + i) : 0.f;
    if (EPI == EDA_EPI_BIAS_LN || EPI == EDA_EPI_BIAS_RES_LN)
      for (int i = threadIdx.x; i < GBN; i += G_THREADS) { s_gamma[i] = __ldg(p.gamma + i); s_beta[i] = __ldg(p.beta + i); }
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();              // both CTAs' barriers and TMEM exist before any cross-CTA signal
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===== TMA producer (both CTAs): own 128 rows of A, own half of the W tile; every byte of
    // the pair is counted on the LEADER's full barrier =====
    if (elect_one()) {
      const uint32_t pair_tx = 2 * (p.passes == 3 ? G_STAGE : (G_A_TILE + G_B_TILE));
      const uint32_t full_leader0 = map_to_rank(&full[0], 0);     // barriers are contiguous 8-byte objects
      uint32_t it = 0;
      for (int tile = pair; tile < n_tiles; tile += n_pairs) {
        const int mn = WG ? tile % n_mn : tile, ks = WG ? tile / n_mn : 0;
        const int m0 = (mn / n_tiles_n) * 2 * GBM + (int)rank * GBM;
        const int n0 = (mn % n_tiles_n) * GBN + (int)rank * (GBN / 2);
        const int kb0 = ks * nkb_slice, kb1 = min(nkb_all, kb0 + nkb_slice);
        for (int kb = kb0; kb < kb1; ++kb, ++it) {
          const int s = it % GSTAGES;
          mbar_wait(&empty[s], ((it / GSTAGES) & 1) ^ 1);      // the pair's MMAs have read this stage (multicast commit)
          uint8_t* st = smem + s * G_STAGE;
          if (rank == 0) mbar_expect_tx(&full[s], pair_tx);
          const uint32_t fb = full_leader0 + 8 * s;
          if constexpr (WG) {
            // MN-major operands: a tile is two boxes of 64 (output index, contiguous in HBM) x 64 reduction rows
#pragma unroll
            for (int blk = 0; blk < 2; ++blk) {
              tma_load_2d_pair(st + blk * 8192, &tmA_hi, fb, m0 + blk * 64, kb * GBK);
              tma_load_2d_pair(st + 2 * G_A_TILE + blk * 8192, &tmB_hi, fb, n0 + blk * 64, kb * GBK);
              if (p.passes == 3) {
                tma_load_2d_pair(st + G_A_TILE + blk * 8192, &tmA_lo, fb, m0 + blk * 64, kb * GBK);
                tma_load_2d_pair(st + 2 * G_A_TILE + G_B_TILE + blk * 8192, &tmB_lo, fb, n0 + blk * 64, kb * GBK);
              }
            }
          } else {
            tma_load_2d_pair(st, &tmA_hi, fb, kb * GBK, m0);
            tma_load_2d_pair(st + 2 * G_A_TILE, &tmB_hi, fb, kb * GBK, n0);
            if (p.passes == 3) {
              tma_load_2d_pair(st + G_A_TILE, &tmA_lo, fb, kb * GBK, m0);
              tma_load_2d_pair(st + 2 * G_A_TILE + G_B_TILE, &tmB_lo, fb, kb * GBK, n0);
            }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer: one thread of the LEADER CTA, cta_group::2, 256 x 256 x 16 per instruction =====
    // (elect.sync, not lane == 0: ptxas then issues the uniform-datapath MMAs back to back
    // instead of wrapping each one in an active-lane loop)
    if (rank == 0 && elect_one()) {
      const uint32_t idesc = idesc_bf16_f32(2 * GBM, GBN) | (WG ? ((1u << 15) | (1u << 16)) : 0u);   // WG: A and B MN-major
      uint32_t it = 0, lt = 0;
      for (int tile = pair; tile < n_tiles; tile += n_pairs, ++lt) {
        const uint32_t ab = lt & 1;
        mbar_wait(&tempty[ab], ((lt >> 1) & 1) ^ 1);       // both CTAs' epilogues have drained this accumulator
        tc_fence_after();
        const uint32_t acc = tmem_base + ab * GBN;
        const int ks = WG ? tile / n_mn : 0;
        const int kb0 = ks * nkb_slice, kb1 = min(nkb_all, kb0 + nkb_slice);
        for (int kb = kb0; kb < kb1; ++kb, ++it) {
          const int s = it % GSTAGES;
          mbar_wait(&full[s], (it / GSTAGES) & 1);
          tc_fence_after();
          const uint32_t sa = smem_u32(smem + s * G_STAGE);
          const uint64_t a_hi = WG ? smem_desc_sw128_mn(sa) : smem_desc_sw128(sa);
          const uint64_t a_lo = WG ? smem_desc_sw128_mn(sa + G_A_TILE) : smem_desc_sw128(sa + G_A_TILE);
          const uint64_t b_hi = WG ? smem_desc_sw128_mn(sa + 2 * G_A_TILE) : smem_desc_sw128(sa + 2 * G_A_TILE);
          const uint64_t b_lo = WG ? smem_desc_sw128_mn(sa + 2 * G_A_TILE + G_B_TILE) : smem_desc_sw128(sa + 2 * G_A_TILE + G_B_TILE);
#pragma unroll
          for (int k = 0; k < GBK / 16; ++k) {
            // K-major: +32 B per 16-element k-step; MN-major: 16 reduction rows of 128 B = +2048 B
            const uint64_t adv = uint64_t((WG ? k * 2048 : k * 32) >> 4);
            umma_bf16_pair(acc, a_hi + adv, b_hi + adv, idesc, ((kb - kb0) | k) != 0);
            if (p.passes == 3) {
              umma_bf16_pair(acc, a_hi + adv, b_lo + adv, idesc, 1);
              umma_bf16_pair(acc, a_lo + adv, b_hi + adv, idesc, 1);
            }
          }
          umma_commit_pair(&empty[s]);   // frees this smem stage in BOTH CTAs once the MMAs above have read it
        }
        umma_commit_pair(&tfull[ab]);    // accumulator halves complete in both CTAs
      }
    }
  } else {
    // ===== epilogue: warps 2..9.  Warp w touches TMEM lanes 32*(w%4)..+31 (rows of the tile);
    // the two warps of a lane quadrant split the 256 columns in halves of 128. =====
    const int q = warp & 3, ch = (warp - 2) >> 2;
    uint32_t lt = 0;
    // staged plan: this warp's two 4 KB buffers (32 rows x 128 B, SWIZZLE_128B like the tensor maps)
    uint8_t* const stg = smem + G_STAGES_STAGED * G_STAGE + (warp - 2) * G_STG_WARP;
    uint32_t n_stores = 0;
    const uint32_t tempty_leader0 = map_to_rank(&tempty[0], 0);
    for (int tile = pair; tile < n_tiles; tile += n_pairs, ++lt) {
      const int mn = WG ? tile % n_mn : tile;
      const int m0 = (mn / n_tiles_n) * 2 * GBM + (int)rank * GBM, n_blk = mn % n_tiles_n, n0 = n_blk * GBN;
      const uint32_t ab = lt & 1;
      const int rbase = m0 + q * 32;                              // first global row of this warp
      // residual chunk c of this thread's row (RES_LN): 8 x 16 B, thread-per-row
      auto load_resid = [&](int c, float4 (&dst)[8]) {
        const int row = rbase + lane;
        const float* src = p.resid + (size_t)row * p.ldr + ch * 128 + c * 32;
        if (row >= p.M) {
#pragma unroll
          for (int j = 0; j < 8; ++j) dst[j] = make_float4(0.f, 0.f, 0.f, 0.f);
        } else if (((uintptr_t)src & 31) == 0) {
#pragma unroll
          for (int j = 0; j < 4; ++j) ld_global_nc_v8(src + 8 * j, dst[2 * j], dst[2 * j + 1]);
        } else {
#pragma unroll
          for (int j = 0; j < 8; ++j) dst[j] = __ldg(reinterpret_cast<const float4*>(src) + j);
        }
      };
      float4 rp[8];
      if constexpr (EPI == EDA_EPI_BIAS_RES_LN || EPI == EDA_EPI_BIAS_RES) load_resid(0, rp);   // in flight while the MMAs finish
      mbar_wait(&tfull[ab], (lt >> 1) & 1);
      tc_fence_after();
      const uint32_t trow = tmem_base + ab * GBN + ch * 128 + (uint32_t(q * 32) << 16);

      // thread-per-row stores straight from registers: 32 consecutive columns of one row are
      // 128 B of fp32 / 64 B of bf16, written as 32-byte (full-sector) STG.256s
      auto store_chunk = [&](const float (&v)[32], int gcol0, bf16* out_hi, bf16* out_lo, int lds) {
        const int grow = rbase + lane;
        if (grow >= p.M || gcol0 >= p.N) return;
        const bool whole = gcol0 + 32 <= p.N;
        if (p.C && EPI != GEPI_QKV) {
          float* dst = (WG ? p.C + (size_t)(tile / n_mn) * p.c_slice : p.C) + (size_t)grow * p.ldc + gcol0;
          if (whole && ((uintptr_t)dst & 31) == 0) {
#pragma unroll
            for (int j = 0; j < 4; ++j)
              st_global_v8(dst + 8 * j, __float_as_uint(v[8 * j]), __float_as_uint(v[8 * j + 1]),
                           __float_as_uint(v[8 * j + 2]), __float_as_uint(v[8 * j + 3]), __float_as_uint(v[8 * j + 4]),
                           __float_as_uint(v[8 * j + 5]), __float_as_uint(v[8 * j + 6]), __float_as_uint(v[8 * j + 7]));
          } else {
#pragma unroll
            for (int j = 0; j < 32; ++j)
              if (gcol0 + j < p.N) dst[j] = v[j];
          }
        }
        if (out_hi) {      // split outputs require N % 4 == 0 (checked on the host)
          uint32_t h[16], l[16];
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            h[j] = pack_bf16x2(v[2 * j], v[2 * j + 1]);
            l[j] = pack_bf16x2(v[2 * j] - __uint_as_float(h[j] << 16), v[2 * j + 1] - __uint_as_float(h[j] & 0xffff0000u));
          }
          bf16* dh = out_hi + (size_t)grow * lds + gcol0;
          bf16* dl = out_lo + (size_t)grow * lds + gcol0;
          if (whole && (((uintptr_t)dh | (uintptr_t)dl) & 31) == 0) {
            st_global_v8(dh, h[0], h[1], h[2], h[3], h[4], h[5], h[6], h[7]);
            st_global_v8(dh + 16, h[8], h[9], h[10], h[11], h[12], h[13], h[14], h[15]);
            st_global_v8(dl, l[0], l[1], l[2], l[3], l[4], l[5], l[6], l[7]);
            st_global_v8(dl + 16, l[8], l[9], l[10], l[11], l[12], l[13], l[14], l[15]);
          } else {
#pragma unroll
            for (int j = 0; j < 16; ++j)
              if (gcol0 + 2 * j < p.N) {
                *reinterpret_cast<uint32_t*>(dh + 2 * j) = h[j];
                *reinterpret_cast<uint32_t*>(dl + 2 * j) = l[j];
              }
          }
        }
      };
      // Staged plan: the lane's 32 words (128 B of its row) go to shared memory with eight
      // conflict-free STS.128 (16 B units XOR-swizzled by row), then ONE TMA store writes the
      // 32 x 128 B block; rows >= M and columns >= N are clipped by the tensor map.  One LSU
      // wavefront per 128 B instead of one per 32 B sector for stores issued from registers, on
      // the data pipe the tensor core's operand reads share.
      auto tma_store_rows = [&](const uint32_t (&w)[32], const CUtensorMap* tm, int col_elem) {
        if (rbase >= p.M) return;                                        // warp-uniform
        // (elect.sync picks the same lane every time for a full mask, so the bulk groups stay with one thread)
        if (elect_one()) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");   // buffer used two stores ago is free
        __syncwarp();
        uint8_t* buf = stg + (n_stores & 1) * 4096;
        const uint32_t row_addr = smem_u32(buf) + lane * 128;
#pragma unroll
        for (int u = 0; u < 8; ++u)
          asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};"
                       ::"r"(row_addr + ((u ^ (lane & 7)) << 4)), "r"(w[4 * u]), "r"(w[4 * u + 1]), "r"(w[4 * u + 2]), "r"(w[4 * u + 3])
                       : "memory");
        fence_proxy_async();
        __syncwarp();
        if (elect_one()) {
          asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];"
                       ::"l"(reinterpret_cast<uint64_t>(tm)), "r"(smem_u32(buf)), "r"(col_elem), "r"(rbase) : "memory");
          asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        }
        ++n_stores;
      };
      // split result: chunk `half` (0/1) of a 64-column pair is packed into words [16 half, 16 half + 16)
      auto pack_split = [&](const float (&v)[32], int half, uint32_t (&hi)[32], uint32_t (&lo)[32]) {
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          const uint32_t h = pack_bf16x2(v[2 * j], v[2 * j + 1]);
          hi[16 * half + j] = h;
          lo[16 * half + j] = pack_bf16x2(v[2 * j] - __uint_as_float(h << 16), v[2 * j + 1] - __uint_as_float(h & 0xffff0000u));
        }
      };
      // one finished 32-column chunk (half 0/1 of a 64-column pair starting at gcol0 - 32 half)
      auto emit_chunk = [&](float (&v)[32], int gcol0, int half, uint32_t (&hi)[32], uint32_t (&lo)[32],
                            bf16* out_hi, bf16* out_lo, int lds) {
        if constexpr (ST) {
          if (out_hi) pack_split(v, half, hi, lo);
          if (p.C && EPI != GEPI_QKV) tma_store_rows(reinterpret_cast<const uint32_t(&)[32]>(v), &tmC, gcol0);
          if (out_hi && half == 1) {
            tma_store_rows(hi, &tmC_hi, gcol0 - 32);
            tma_store_rows(lo, &tmC_lo, gcol0 - 32);
          }
        } else {
          store_chunk(v, gcol0, out_hi, out_lo, lds);
        }
      };
      // v[j] += vec[col0 + j] (bias-like row vector; 16 B loads when aligned, broadcast across lanes)
      auto add_vec = [&](float (&v)[32], const float* vec, int col0) {
        if (vec_smem) {            // (only ever called with vec == p.bias)
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const float4 b = *reinterpret_cast<const float4*>(s_bias + col0 + 4 * j);
            v[4 * j] += b.x; v[4 * j + 1] += b.y; v[4 * j + 2] += b.z; v[4 * j + 3] += b.w;
          }
        } else if (vec == nullptr) {
        } else if (((uintptr_t)(vec + col0) & 15) == 0 && col0 + 32 <= p.N) {
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const float4 b = __ldg(reinterpret_cast<const float4*>(vec + col0) + j);
            v[4 * j] += b.x; v[4 * j + 1] += b.y; v[4 * j + 2] += b.z; v[4 * j + 3] += b.w;
          }
        } else {
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] += col0 + j < p.N ? __ldg(vec + col0 + j) : 0.f;
        }
      };

      if constexpr (EPI == EDA_EPI_BIAS_GLU) {
        // nn.GLU over channel pairs (c, c + N/2): the host interleaves the weight rows in blocks of 32
        // (a-block, gate-block), so both halves of a pair sit in ONE thread's consecutive 32-column
        // chunks; output (M, N/2) fp32, column = 32 * (pair index) + j.
#pragma unroll 1
        for (int cp = 0; cp < 2; ++cp) {
          const int colp = n0 + ch * 128 + cp * 64;
          if (colp >= p.N) break;
          float a[32], g[32];
          tmem_ld32(trow + cp * 64, a);
          tmem_ld32(trow + cp * 64 + 32, g);
          add_vec(a, p.bias, colp);
          add_vec(g, p.bias, colp + 32);
#pragma unroll
          for (int j = 0; j < 32; ++j) a[j] = a[j] * __fdividef(1.f, 1.f + __expf(-g[j]));
          const int grow = rbase + lane;
          if (grow < p.M) {
            float* dst = p.C + (size_t)grow * p.ldc + colp / 2;
#pragma unroll
            for (int j = 0; j < 4; ++j)
              st_global_v8(dst + 8 * j, __float_as_uint(a[8 * j]), __float_as_uint(a[8 * j + 1]), __float_as_uint(a[8 * j + 2]),
                           __float_as_uint(a[8 * j + 3]), __float_as_uint(a[8 * j + 4]), __float_as_uint(a[8 * j + 5]),
                           __float_as_uint(a[8 * j + 6]), __float_as_uint(a[8 * j + 7]));
          }
        }
      } else if constexpr (EPI == EDA_EPI_BIAS || EPI == EDA_EPI_BIAS_RELU || EPI == EDA_EPI_RELU_MASK || EPI == EDA_EPI_BIAS_SILU) {
#pragma unroll 1
        for (int cp = 0; cp < 2; ++cp) {                       // pairs of 32-column chunks
          if (n0 + ch * 128 + cp * 64 >= p.N) break;
          uint32_t hi[32], lo[32];
#pragma unroll
          for (int half = 0; half < 2; ++half) {
            const int col0 = n0 + ch * 128 + cp * 64 + half * 32;
            float v[32];
            tmem_ld32(trow + cp * 64 + half * 32, v);
            add_vec(v, p.bias, col0);
            if (EPI == EDA_EPI_BIAS_RELU) {
#pragma unroll
              for (int j = 0; j < 32; ++j) v[j] = fmaxf(v[j], 0.f);
              if (EX && p.drop.thresh) apply_dropout32(v, p.drop, rbase + lane, col0);
            }
            if (EPI == EDA_EPI_BIAS_SILU) {
#pragma unroll
              for (int j = 0; j < 32; ++j) v[j] = v[j] * __fdividef(1.f, 1.f + __expf(-v[j]));
            }
            if constexpr (EPI == EDA_EPI_RELU_MASK) {
              // ReLU backward: keep the gradient where the forward activation (its bf16 hi part) was > 0
              const int row = rbase + lane;
              if (row < p.M && col0 + 32 <= p.N) {
                const uint4* mp = reinterpret_cast<const uint4*>(p.mask + (size_t)row * p.ldm + col0);
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                  const uint4 mv = __ldg(mp + u);
                  const uint32_t w[4] = {mv.x, mv.y, mv.z, mv.w};
#pragma unroll
                  for (int e = 0; e < 4; ++e) {
                    // positive bf16: sign bit clear and magnitude non-zero
                    if (!((w[e] & 0x7fffu) != 0 && !(w[e] & 0x8000u))) v[8 * u + 2 * e] = 0.f;
                    if (!((w[e] & 0x7fff0000u) != 0 && !(w[e] & 0x80000000u))) v[8 * u + 2 * e + 1] = 0.f;
                  }
                }
              }
              if (p.mask_scale != 0.f) {
#pragma unroll
                for (int j = 0; j < 32; ++j) v[j] *= p.mask_scale;
              }
            }
            emit_chunk(v, col0, half, hi, lo, p.Chi, p.Clo, p.lds);
          }
        }
      } else if constexpr (EPI == GEPI_QKV) {
        // n_blk 0: q (pre-scaled), 1: k -> split (M, 2*GBN); 2: v -> transposed per (b, h)
        const int row = rbase + lane;
        const int b = row / p.T, t = row - b * p.T;
        const float sc = n_blk == 0 ? p.q_scale : 1.0f;
#pragma unroll 1
        for (int cp = 0; cp < 2; ++cp) {
          uint32_t hi[32], lo[32];
#pragma unroll
          for (int half = 0; half < 2; ++half) {
            const int cc = ch * 4 + cp * 2 + half;                   // 32-column chunk within the 256-wide tile
            float v[32];
            tmem_ld32(trow + (cp * 2 + half) * 32, v);
            add_vec(v, p.bias, n0 + cc * 32);
#pragma unroll
            for (int j = 0; j < 32; ++j) v[j] *= sc;
            if (n_blk < 2) {
              emit_chunk(v, n0 + cc * 32, half, hi, lo, p.Chi, p.Clo, p.lds);
            } else if (row < p.M) {
              const int h = cc >> 1, d0 = (cc & 1) * 32;
              const size_t base = ((size_t)(b * p.H + h) * 64 + d0) * p.Tp + t;
#pragma unroll
              for (int j = 0; j < 32; ++j) {
                const bf16 hv = __float2bfloat16_rn(v[j]);
                p.vt_hi[base + (size_t)j * p.Tp] = hv;
                p.vt_lo[base + (size_t)j * p.Tp] = __float2bfloat16_rn(v[j] - __bfloat162float(hv));
              }
            }
          }
        }
      } else {
        // LayerNorm epilogues: N == 256 == GBN.  A row is owned by TWO threads (column halves) and
        // stays in TMEM between the passes; row statistics are exchanged through shared memory
        // with a 64-thread named barrier per lane quadrant.  Two-pass variance like ATen.
        // The residual is read thread-per-row (each 32 B sector is consumed by two back-to-back
        // loads) and prefetched one 32-column chunk ahead so DRAM latency is off the critical path.
        constexpr bool kLN = EPI != EDA_EPI_BIAS_RES;       // BIAS_RES: residual add only, rows leave after one pass
        constexpr bool kRes = EPI == EDA_EPI_BIAS_RES_LN || EPI == EDA_EPI_BIAS_RES;
        float sum = 0.f;
        uint32_t rhi[32], rlo[32];
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          const int col0 = ch * 128 + c * 32;
          float v[32];
          tmem_ld32(trow + c * 32, v);
          // (warp-uniform) Conformer half-step residuals / training dropout act on A.W^T + bias BEFORE the residual
          const bool scaled = EX && (p.out_scale != 0.f || p.drop.thresh != 0);
          if (scaled) {
            add_vec(v, p.bias, col0);
            if (p.out_scale != 0.f) {
#pragma unroll
              for (int j = 0; j < 32; ++j) v[j] *= p.out_scale;
            }
            if (p.drop.thresh) apply_dropout32(v, p.drop, rbase + lane, col0);
          }
          if constexpr (kRes) {
            float4 rn[8];
            if (c < 3) load_resid(c + 1, rn);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              v[4 * j] += rp[j].x; v[4 * j + 1] += rp[j].y; v[4 * j + 2] += rp[j].z; v[4 * j + 3] += rp[j].w;
            }
            if (c < 3) {
#pragma unroll
              for (int j = 0; j < 8; ++j) rp[j] = rn[j];
            }
          }
          if (!scaled) add_vec(v, p.bias, col0);
          if constexpr (!kLN) {
            emit_chunk(v, col0, c & 1, rhi, rlo, p.Chi, p.Clo, p.lds);
          } else {
#pragma unroll
            for (int j = 0; j < 32; ++j) sum += v[j];
            tmem_st32(trow + c * 32, v);
          }
        }
        if constexpr (kLN) {
        tmem_wait_st();
        const int xr = q * 32 + lane;
        ln_x[0][ch][xr] = sum;
        asm volatile("bar.sync %0, 64;" ::"r"(1 + q) : "memory");
        const float mean = (ln_x[0][0][xr] + ln_x[0][1][xr]) * (1.0f / GBN);
        float sq = 0.f;
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          float v[32];
          tmem_ld32(trow + c * 32, v);
#pragma unroll
          for (int j = 0; j < 32; ++j) { const float d = v[j] - mean; sq = fmaf(d, d, sq); }
        }
        ln_x[1][ch][xr] = sq;
        asm volatile("bar.sync %0, 64;" ::"r"(1 + q) : "memory");
        const float rstd = 1.0f / sqrtf((ln_x[1][0][xr] + ln_x[1][1][xr]) * (1.0f / GBN) + 1e-5f);
        if (EX && p.rstd_out && ch == 0 && rbase + lane < p.M) p.rstd_out[rbase + lane] = rstd;
#pragma unroll 1
        for (int cp = 0; cp < 2; ++cp) {
          uint32_t hi[32], lo[32];
#pragma unroll
          for (int half = 0; half < 2; ++half) {
            const int col0 = ch * 128 + cp * 64 + half * 32;
            float v[32];
            tmem_ld32(trow + cp * 64 + half * 32, v);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const float4 g4 = *reinterpret_cast<const float4*>(s_gamma + col0 + 4 * j);
              const float4 b4 = *reinterpret_cast<const float4*>(s_beta + col0 + 4 * j);
              v[4 * j] = (v[4 * j] - mean) * rstd * g4.x + b4.x;
              v[4 * j + 1] = (v[4 * j + 1] - mean) * rstd * g4.y + b4.y;
              v[4 * j + 2] = (v[4 * j + 2] - mean) * rstd * g4.z + b4.z;
              v[4 * j + 3] = (v[4 * j + 3] - mean) * rstd * g4.w + b4.w;
            }
            emit_chunk(v, col0, half, hi, lo, p.Chi, p.Clo, p.lds);
          }
        }
        // (ln_x reuse is safe: a warp rewrites ln_x[k] only after the partner has passed the
        //  barrier that follows its read of ln_x[k])
        }   // kLN
      }
      {

        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster(tempty_leader0 + 8 * ab);
      }
    }
    if (ST && elect_one()) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");   // staging smem outlives its readers
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();              // the leader's MMAs touched the peer's smem and TMEM: leave together
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc_pair(tmem_base, 512);
  }
}

// ------------------------------------------------------------------------- host ---------
typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static PFN_encodeTiled get_encode() {
  static PFN_encodeTiled fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<PFN_encodeTiled>(p);
  }
  return fn;
}

// bf16 tensor maps with a SWIZZLE_128B box whose inner extent is 64 elements (128 B).  Encoding
// costs a few microseconds on the host, so maps are cached on (pointer, geometry).
struct TmapKey {
  const void* ptr; uint64_t d0, d1, d2, s1, s2; uint32_t b1;
  bool operator==(const TmapKey& o) const {
    return ptr == o.ptr && d0 == o.d0 && d1 == o.d1 && d2 == o.d2 && s1 == o.s1 && s2 == o.s2 && b1 == o.b1;
  }
};
struct TmapHash {
  size_t operator()(const TmapKey& k) const {
    size_t h = std::hash<const void*>()(k.ptr);
    for (uint64_t v : {k.d0, k.d1, k.d2, k.s1, k.s2, (uint64_t)k.b1}) h = h * 1000003u ^ std::hash<uint64_t>()(v);
    return h;
  }
};
static std::unordered_map<TmapKey, CUtensorMap, TmapHash> g_tmaps;
static std::mutex g_tmap_mu;

// rank-3 view: dims (d0 inner, d1, d2), byte strides s1 (of d1) and s2 (of d2); box {64, b1, 1}
static int make_tmap_impl(CUtensorMap* m, int rank, const void* ptr, uint64_t d0, uint64_t d1, uint64_t d2,
                          uint64_t s1_bytes, uint64_t s2_bytes, uint32_t box_rows, bool f32 = false) {
  const TmapKey key{ptr, d0, d1, d2, s1_bytes, s2_bytes, box_rows | (uint32_t(rank) << 16) | (uint32_t(f32) << 20)};
  {
    std::lock_guard<std::mutex> g(g_tmap_mu);
    auto it = g_tmaps.find(key);
    if (it != g_tmaps.end()) { *m = it->second; return EDA_OK; }
  }
  PFN_encodeTiled enc = get_encode();
  if (!enc) { set_error("cuTensorMapEncodeTiled entry point not available"); return EDA_ECUDA; }
  cuuint64_t gdim[3] = {d0, d1, d2};
  cuuint64_t gstr[2] = {s1_bytes, s2_bytes};
  cuuint32_t box[3] = {f32 ? 32u : 64u, box_rows, 1};          // 128 B inner extent either way
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = enc(m, f32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, rank, const_cast<void*>(ptr), gdim, gstr, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed (%d) dims=(%llu,%llu,%llu) strides=(%llu,%llu) ptr=%p", (int)r,
              (unsigned long long)d0, (unsigned long long)d1, (unsigned long long)d2,
              (unsigned long long)s1_bytes, (unsigned long long)s2_bytes, ptr);
    return EDA_ECUDA;
  }
  std::lock_guard<std::mutex> g(g_tmap_mu);
  if (g_tmaps.size() > 8192) g_tmaps.clear();
  g_tmaps.emplace(key, *m);
  return EDA_OK;
}

int make_tmap_bf16_3d(CUtensorMap* m, const void* ptr, uint64_t d0, uint64_t d1, uint64_t d2, uint64_t s1_bytes,
                      uint64_t s2_bytes, uint32_t box_rows) {
  return make_tmap_impl(m, 3, ptr, d0, d1, d2, s1_bytes, s2_bytes, box_rows);
}

// 2-D bf16 row-major (rows x cols, row stride ld elements) with a {64, box_rows} box
// fp32 (rows x cols, row stride ld elements), box 32 columns x box_rows
static int make_tmap_f32(CUtensorMap* m, const void* ptr, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_rows) {
  return make_tmap_impl(m, 2, ptr, cols, rows, 1, ld * 4, 0, box_rows, true);
}
int make_tmap_bf16(CUtensorMap* m, const void* ptr, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_rows) {
  return make_tmap_impl(m, 2, ptr, cols, rows, 1, ld * 2, 0, box_rows);
}

int split_bf16(const float* src, int ld, int rows, int K, void* hi, void* lo, int ldk, cudaStream_t st) {
  EDA_REQUIRE(src && hi && lo && rows > 0 && K > 0 && ldk >= K && ldk % 8 == 0, "eda_split_bf16: bad arguments (ldk must be a multiple of 8 and >= K)");
  const int64_t n = (int64_t)rows * (ldk / 2);
  split_bf16_kernel<<<(int)std::min<int64_t>(ceil_div<int64_t>(n, 256), num_sms() * 16), 256, 0, st>>>(
      src, ld, rows, K, ldk, (bf16*)hi, (bf16*)lo);
  EDA_LAUNCH_CHECK();
  return EDA_OK;
}

template <int EPI, bool ST, bool WG = false, bool EX = false>
static int launch_gemm2(const CUtensorMap& a_hi, const CUtensorMap& a_lo, const CUtensorMap& b_hi, const CUtensorMap& b_lo,
                        const CUtensorMap& c, const CUtensorMap& c_hi, const CUtensorMap& c_lo, const GemmParams& p,
                        cudaStream_t st) {
  static DeviceOnce attr;
  if (!attr.done()) {
    EDA_CUDA(cudaFuncSetAttribute(gemm_split_kernel<EPI, ST, WG, EX>, cudaFuncAttributeMaxDynamicSharedMemorySize, G_SMEM));
    attr.mark();
  }
  const int n_tiles = ceil_div(p.N, GBN) * ceil_div(p.M, 2 * GBM) * (WG ? p.ksplit : 1);   // one 256 x 256 tile per CTA pair
  cudaLaunchConfig_t cfg = {};
  cudaLaunchAttribute attrs[1];
  cfg.gridDim = dim3(2 * std::min(n_tiles, num_sms() / 2));
  cfg.blockDim = dim3(G_THREADS);
  cfg.dynamicSmemBytes = G_SMEM;
  cfg.stream = st;
  attrs[0].id = cudaLaunchAttributeClusterDimension;
  attrs[0].val.clusterDim.x = 2; attrs[0].val.clusterDim.y = 1; attrs[0].val.clusterDim.z = 1;
  cfg.attrs = attrs; cfg.numAttrs = 1;
  EDA_CUDA(cudaLaunchKernelEx(&cfg, gemm_split_kernel<EPI, ST, WG, EX>, a_hi, a_lo, b_hi, b_lo, c, c_hi, c_lo, p));
  count_launch();
  return EDA_OK;
}

// Short K (the stores dominate) -> staged plan with TMA stores, if the outputs can be described by
// tensor maps (16 B aligned bases and row pitches); `n_split_cols` = logical width of the split output.
template <int EPI>
static int launch_gemm(const CUtensorMap& a_hi, const CUtensorMap& a_lo, const CUtensorMap& b_hi, const CUtensorMap& b_lo,
                       const GemmParams& p, int n_split_cols, cudaStream_t st) {
  const bool c_out = p.C && EPI != GEPI_QKV;
  bool staged = p.Kp <= 8 * GBK;
  if (c_out && (((uintptr_t)p.C & 15) || (p.ldc & 3))) staged = false;
  if (p.Chi && ((((uintptr_t)p.Chi | (uintptr_t)p.Clo) & 15) || (p.lds & 7))) staged = false;
  // the extras exist for the epilogues that can use them (ReLU dropout, residual scale / dropout)
  constexpr bool kCanEx = EPI == EDA_EPI_BIAS_RELU || EPI == EDA_EPI_BIAS_RES_LN || EPI == EDA_EPI_BIAS_RES || EPI == EDA_EPI_BIAS_LN;
  const bool ex = kCanEx && (p.out_scale != 0.f || p.drop.thresh != 0 || p.rstd_out != nullptr);
  if (!staged) {
    if constexpr (kCanEx) { if (ex) return launch_gemm2<EPI, false, false, true>(a_hi, a_lo, b_hi, b_lo, a_hi, a_hi, a_hi, p, st); }
    return launch_gemm2<EPI, false>(a_hi, a_lo, b_hi, b_lo, a_hi, a_hi, a_hi, p, st);
  }
  CUtensorMap c = a_hi, c_hi = a_hi, c_lo = a_hi;        // placeholders for absent outputs (never dereferenced)
  int rc;
  if (c_out && (rc = make_tmap_f32(&c, p.C, p.M, p.N, p.ldc, 32)) != EDA_OK) return rc;
  if (p.Chi) {
    if ((rc = make_tmap_bf16(&c_hi, p.Chi, p.M, n_split_cols, p.lds, 32)) != EDA_OK) return rc;
    if ((rc = make_tmap_bf16(&c_lo, p.Clo, p.M, n_split_cols, p.lds, 32)) != EDA_OK) return rc;
  }
  if constexpr (kCanEx) { if (ex) return launch_gemm2<EPI, true, false, true>(a_hi, a_lo, b_hi, b_lo, c, c_hi, c_lo, p, st); }
  return launch_gemm2<EPI, true>(a_hi, a_lo, b_hi, b_lo, c, c_hi, c_lo, p, st);
}

// A: split (M x K, row stride lda elements), W: split (N x K, row stride ldw).
int gemm_split(const void* a_hi, const void* a_lo, int lda, const void* w_hi, const void* w_lo, int ldw,
               GemmParams p, int K, int epi, int prec, cudaStream_t st) {
  EDA_REQUIRE(a_hi && a_lo && w_hi && w_lo && p.M > 0 && p.N > 0 && K > 0, "eda_gemm_split: bad arguments");
  EDA_REQUIRE(lda % 8 == 0 && ldw % 8 == 0, "eda_gemm_split: operand row strides must be multiples of 8 elements");
  EDA_REQUIRE(((uintptr_t)a_hi | (uintptr_t)a_lo | (uintptr_t)w_hi | (uintptr_t)w_lo) % 16 == 0, "eda_gemm_split: operands must be 16B aligned");
  EDA_REQUIRE(prec == EDA_PREC_BF16X3 || prec == EDA_PREC_BF16, "eda_gemm_split: precision must be BF16X3 or BF16");
  if (p.C) EDA_REQUIRE(p.ldc % 4 == 0 && (uintptr_t)p.C % 16 == 0, "eda_gemm_split: C must be 16B aligned with ldc %% 4 == 0");
  if (p.Chi) EDA_REQUIRE(p.Clo && p.lds % 4 == 0 && p.N % 4 == 0 && ((uintptr_t)p.Chi | (uintptr_t)p.Clo) % 8 == 0, "eda_gemm_split: split output needs N %% 4 == 0, lds %% 4 == 0, 8B alignment");
  const bool ln = epi == EDA_EPI_BIAS_LN || epi == EDA_EPI_BIAS_RES_LN;
  if (ln) EDA_REQUIRE(p.N == 256 && p.gamma && p.beta && p.bias, "eda_gemm_split: LayerNorm epilogues need N == 256, bias, gamma, beta");
  if (epi == EDA_EPI_BIAS_RES_LN || epi == EDA_EPI_BIAS_RES)
    EDA_REQUIRE(p.N == 256 && p.resid && p.ldr % 4 == 0 && (uintptr_t)p.resid % 16 == 0, "eda_gemm_split: residual epilogues need N == 256 and an aligned residual");
  if (epi == EDA_EPI_RELU_MASK)
    EDA_REQUIRE(p.mask && p.ldm % 8 == 0 && (uintptr_t)p.mask % 16 == 0 && p.N % 32 == 0, "eda_gemm_split: mask must be 16B aligned, ldm %% 8 == 0, N %% 32 == 0");
  p.Kp = K;
  p.passes = prec == EDA_PREC_BF16 ? 1 : 3;
  CUtensorMap ta_hi, ta_lo, tb_hi, tb_lo;
  int rc;
  if ((rc = make_tmap_bf16(&ta_hi, a_hi, p.M, K, lda, GBM)) != EDA_OK) return rc;
  if ((rc = make_tmap_bf16(&ta_lo, a_lo, p.M, K, lda, GBM)) != EDA_OK) return rc;
  if ((rc = make_tmap_bf16(&tb_hi, w_hi, p.N, K, ldw, GBN / 2)) != EDA_OK) return rc;
  if ((rc = make_tmap_bf16(&tb_lo, w_lo, p.N, K, ldw, GBN / 2)) != EDA_OK) return rc;
  switch (epi) {
    case EDA_EPI_BIAS: return launch_gemm<EDA_EPI_BIAS>(ta_hi, ta_lo, tb_hi, tb_lo, p, p.N, st);
    case EDA_EPI_BIAS_RELU: return launch_gemm<EDA_EPI_BIAS_RELU>(ta_hi, ta_lo, tb_hi, tb_lo, p, p.N, st);
    case EDA_EPI_BIAS_LN: return launch_gemm<EDA_EPI_BIAS_LN>(ta_hi, ta_lo, tb_hi, tb_lo, p, p.N, st);
    case EDA_EPI_BIAS_RES_LN: return launch_gemm<EDA_EPI_BIAS_RES_LN>(ta_hi, ta_lo, tb_hi, tb_lo, p, p.N, st);
    case GEPI_QKV: return launch_gemm<GEPI_QKV>(ta_hi, ta_lo, tb_hi, tb_lo, p, 2 * GBN, st);
    case EDA_EPI_BIAS_RES: return launch_gemm<EDA_EPI_BIAS_RES>(ta_hi, ta_lo, tb_hi, tb_lo, p, p.N, st);
    case EDA_EPI_RELU_MASK: return launch_gemm<EDA_EPI_RELU_MASK>(ta_hi, ta_lo, tb_hi, tb_lo, p, p.N, st);
    case EDA_EPI_BIAS_SILU: return launch_gemm<EDA_EPI_BIAS_SILU>(ta_hi, ta_lo, tb_hi, tb_lo, p, p.N, st);
    case EDA_EPI_BIAS_GLU: {
      EDA_REQUIRE(p.C && !p.Chi && p.N % 64 == 0 && p.ldc % 8 == 0 && (uintptr_t)p.C % 32 == 0, "eda_gemm_split_ex: GLU needs an fp32 output (32B aligned, ldc %% 8 == 0) and N %% 64 == 0");
      return launch_gemm2<EDA_EPI_BIAS_GLU, false>(ta_hi, ta_lo, tb_hi, tb_lo, ta_hi, ta_hi, ta_hi, p, st);
    }
  }
  set_error("eda_gemm_split: unknown epilogue %d", epi);
  return EDA_EINVAL;
}

// ---- weight gradient: dW (N_out x K_in) = dY^T X, reduction over the M rows ------------------
// out[i] = (accumulate ? out[i] : 0) + sum_s part[s][i] in slice order (deterministic)
__global__ void splitk_reduce_kernel(const float* __restrict__ part, int ksplit, long long slice, int rows, int cols,
                                     int ld_part, float* __restrict__ out, int ld_out, int accumulate) {
  const long long n = (long long)rows * cols;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const int r = (int)(i / cols), c = (int)(i % cols);
    float acc = accumulate ? out[(size_t)r * ld_out + c] : 0.f;
    for (int s = 0; s < ksplit; ++s) acc += part[(size_t)s * slice + (size_t)r * ld_part + c];
    out[(size_t)r * ld_out + c] = acc;
  }
}

int gemm_wgrad(const void* dy_hi, const void* dy_lo, int ld_dy, const void* x_hi, const void* x_lo, int ld_x,
               int M, int N_out, int K_in, float* dW, int ld_dw, float* work, size_t work_bytes, int accumulate,
               int prec, cudaStream_t st) {
  EDA_REQUIRE(dy_hi && dy_lo && x_hi && x_lo && dW && work && M > 0 && N_out > 0 && K_in > 0, "eda_gemm_wgrad: bad arguments");
  EDA_REQUIRE(ld_dy % 8 == 0 && ld_x % 8 == 0, "eda_gemm_wgrad: operand row strides must be multiples of 8 elements");
  EDA_REQUIRE(((uintptr_t)dy_hi | (uintptr_t)dy_lo | (uintptr_t)x_hi | (uintptr_t)x_lo) % 16 == 0, "eda_gemm_wgrad: operands must be 16B aligned");
  EDA_REQUIRE(prec == EDA_PREC_BF16X3 || prec == EDA_PREC_BF16, "eda_gemm_wgrad: precision must be BF16X3 or BF16");
  const int ldp = (K_in + 3) / 4 * 4;
  const int n_mn = ceil_div(K_in, GBN) * ceil_div(N_out, 2 * GBM);
  const int nkb = ceil_div(M, GBK);
  const size_t slice = (size_t)N_out * ldp;
  int want = std::max(1, (num_sms() / 2) / n_mn);
  want = (int)std::min<size_t>(want, work_bytes / (slice * sizeof(float)));
  EDA_REQUIRE(want >= 1, "eda_gemm_wgrad: workspace smaller than one %d x %d fp32 slice", N_out, ldp);
  const int nkb_slice = ceil_div(nkb, std::min(want, nkb));
  GemmParams p = {};
  p.C = work; p.ldc = ldp; p.M = N_out; p.N = K_in; p.Kp = M;
  p.passes = prec == EDA_PREC_BF16 ? 1 : 3;
  p.ksplit = ceil_div(nkb, nkb_slice);              // every slice owns at least one k-block
  p.c_slice = (long long)slice;
  CUtensorMap ta_hi, ta_lo, tb_hi, tb_lo;
  int rc;
  if ((rc = make_tmap_bf16(&ta_hi, dy_hi, M, N_out, ld_dy, 64)) != EDA_OK) return rc;
  if ((rc = make_tmap_bf16(&ta_lo, dy_lo, M, N_out, ld_dy, 64)) != EDA_OK) return rc;
  if ((rc = make_tmap_bf16(&tb_hi, x_hi, M, K_in, ld_x, 64)) != EDA_OK) return rc;
  if ((rc = make_tmap_bf16(&tb_lo, x_lo, M, K_in, ld_x, 64)) != EDA_OK) return rc;
  if ((rc = launch_gemm2<EDA_EPI_BIAS, false, true>(ta_hi, ta_lo, tb_hi, tb_lo, ta_hi, ta_hi, ta_hi, p, st)) != EDA_OK) return rc;
  const long long n = (long long)N_out * K_in;
  splitk_reduce_kernel<<<(int)std::min<long long>(ceil_div<long long>(n, 256), num_sms() * 8), 256, 0, st>>>(
      work, p.ksplit, (long long)slice, N_out, K_in, ldp, dW, ld_dw, accumulate);
  EDA_LAUNCH_CHECK();
  return EDA_OK;
}

// eda_linear on the tensor cores: fp32 operands are split on the fly, fp32 result.
int linear_tc(const float* A, int lda, const float* W, int ldw, const float* bias, const float* resid, int ldr,
              const float* gamma, const float* beta, float* C, int ldc, int M, int N, int K, int epi, int prec,
              cudaStream_t st) {
  const int Kp = (K + 7) / 8 * 8;
  const size_t a_el = (size_t)M * Kp, w_el = (size_t)N * Kp;
  int rc;
  // on-the-fly operand split into per-(device, stream) scratch (tags 1, 2)
  bf16* sa = (bf16*)stream_scratch(1, st, 2 * a_el * sizeof(bf16));
  bf16* sw = (bf16*)stream_scratch(2, st, 2 * w_el * sizeof(bf16));
  if (!sa || !sw) return EDA_ECUDA;
  bf16 *a_hi = sa, *a_lo = sa + a_el, *w_hi = sw, *w_lo = sw + w_el;
  if ((rc = split_bf16(A, lda, M, K, a_hi, a_lo, Kp, st)) != EDA_OK) return rc;
  if ((rc = split_bf16(W, ldw, N, K, w_hi, w_lo, Kp, st)) != EDA_OK) return rc;
  GemmParams p = {};
  p.bias = bias; p.resid = resid; p.ldr = ldr; p.gamma = gamma; p.beta = beta;
  p.C = C; p.ldc = ldc; p.M = M; p.N = N;
  return gemm_split(a_hi, a_lo, Kp, w_hi, w_lo, Kp, p, Kp, epi, prec, st);
}

}  // namespace eda

extern "C" {

int eda_split_bf16(const float* src, int ld, int rows, int K, void* hi, void* lo, int ldk, void* stream) {
  return eda::split_bf16(src, ld, rows, K, hi, lo, ldk, (cudaStream_t)stream);
}

int eda_gemm_split(const void* a_hi, const void* a_lo, int lda, const void* w_hi, const void* w_lo, int ldw,
                   const float* bias, const float* resid, int ldr, const float* ln_gamma, const float* ln_beta,
                   float* C, int ldc, void* c_hi, void* c_lo, int ldcs, int M, int N, int K, int epilogue,
                   int prec, void* stream) {
  EDA_REQUIRE(epilogue >= EDA_EPI_BIAS && epilogue <= EDA_EPI_BIAS_RES_LN, "eda_gemm_split: unknown epilogue %d", epilogue);
  EDA_REQUIRE(C || c_hi, "eda_gemm_split: no output requested");
  eda::GemmParams p = {};
  p.bias = bias; p.resid = resid; p.ldr = ldr; p.gamma = ln_gamma; p.beta = ln_beta;
  p.C = C; p.ldc = ldc; p.Chi = (eda::bf16*)c_hi; p.Clo = (eda::bf16*)c_lo; p.lds = ldcs; p.M = M; p.N = N;
  return eda::gemm_split(a_hi, a_lo, lda, w_hi, w_lo, ldw, p, K, epilogue, prec, (cudaStream_t)stream);
}

int eda_gemm_split_ex(const void* a_hi, const void* a_lo, int lda, const void* w_hi, const void* w_lo, int ldw,
                      const float* bias, const float* resid, int ldr, const float* ln_gamma, const float* ln_beta,
                      float* C, int ldc, void* c_hi, void* c_lo, int ldcs, float out_scale, int M, int N, int K,
                      int epilogue, int prec, void* stream) {
  EDA_REQUIRE((epilogue >= EDA_EPI_BIAS && epilogue <= EDA_EPI_BIAS_RES_LN) || epilogue == EDA_EPI_BIAS_RES ||
              epilogue == EDA_EPI_BIAS_SILU || epilogue == EDA_EPI_BIAS_GLU, "eda_gemm_split_ex: unknown epilogue %d", epilogue);
  EDA_REQUIRE(C || c_hi, "eda_gemm_split_ex: no output requested");
  eda::GemmParams p = {};
  p.bias = bias; p.resid = resid; p.ldr = ldr; p.gamma = ln_gamma; p.beta = ln_beta;
  p.C = C; p.ldc = ldc; p.Chi = (eda::bf16*)c_hi; p.Clo = (eda::bf16*)c_lo; p.lds = ldcs; p.M = M; p.N = N;
  p.out_scale = (out_scale == 1.0f) ? 0.f : out_scale;
  return eda::gemm_split(a_hi, a_lo, lda, w_hi, w_lo, ldw, p, K, epilogue, prec, (cudaStream_t)stream);
}

int eda_gemm_split_train(const void* a_hi, const void* a_lo, int lda, const void* w_hi, const void* w_lo, int ldw,
                         const float* bias, const float* resid, int ldr, const float* ln_gamma, const float* ln_beta,
                         float* C, int ldc, void* c_hi, void* c_lo, int ldcs, float* rstd_out, float p_drop,
                         uint64_t seed, int site, int M, int N, int K, int epilogue, int prec, void* stream) {
  EDA_REQUIRE(epilogue == EDA_EPI_BIAS_LN || epilogue == EDA_EPI_BIAS_RES_LN || epilogue == EDA_EPI_BIAS_RELU,
              "eda_gemm_split_train: LayerNorm or ReLU epilogues only");
  EDA_REQUIRE(C || c_hi, "eda_gemm_split_train: no output requested");
  EDA_REQUIRE(p_drop >= 0.f && p_drop < 1.f && (p_drop == 0.f || epilogue != EDA_EPI_BIAS_LN), "eda_gemm_split_train: bad dropout");
  eda::GemmParams p = {};
  p.bias = bias; p.resid = resid; p.ldr = ldr; p.gamma = ln_gamma; p.beta = ln_beta;
  p.C = C; p.ldc = ldc; p.Chi = (eda::bf16*)c_hi; p.Clo = (eda::bf16*)c_lo; p.lds = ldcs; p.M = M; p.N = N;
  p.rstd_out = rstd_out;
  p.drop = eda::make_dropout(p_drop, seed, site);
  return eda::gemm_split(a_hi, a_lo, lda, w_hi, w_lo, ldw, p, K, epilogue, prec, (cudaStream_t)stream);
}

int eda_gemm_dgrad(const void* dy_hi, const void* dy_lo, int ld_dy, const void* wt_hi, const void* wt_lo, int ldw,
                   const float* resid, int ldr, const void* relu_mask_hi, int ldm, float mask_scale, float* dX, int ldx,
                   void* dx_hi, void* dx_lo, int ldxs, int M, int K_in, int N_out, int prec, void* stream) {
  EDA_REQUIRE(dX || dx_hi, "eda_gemm_dgrad: no output requested");
  EDA_REQUIRE(!(resid && relu_mask_hi), "eda_gemm_dgrad: residual and ReLU mask are mutually exclusive");
  eda::GemmParams p = {};
  p.resid = resid; p.ldr = ldr; p.mask = (const eda::bf16*)relu_mask_hi; p.ldm = ldm;
  p.mask_scale = mask_scale == 1.0f ? 0.f : mask_scale;
  p.C = dX; p.ldc = ldx; p.Chi = (eda::bf16*)dx_hi; p.Clo = (eda::bf16*)dx_lo; p.lds = ldxs; p.M = M; p.N = K_in;
  const int epi = resid ? EDA_EPI_BIAS_RES : relu_mask_hi ? EDA_EPI_RELU_MASK : EDA_EPI_BIAS;
  return eda::gemm_split(dy_hi, dy_lo, ld_dy, wt_hi, wt_lo, ldw, p, N_out, epi, prec, (cudaStream_t)stream);
}

int eda_gemm_wgrad(const void* dy_hi, const void* dy_lo, int ld_dy, const void* x_hi, const void* x_lo, int ld_x,
                   int M, int N_out, int K_in, float* dW, int ld_dw, void* workspace, int64_t workspace_bytes,
                   int accumulate, int prec, void* stream) {
  return eda::gemm_wgrad(dy_hi, dy_lo, ld_dy, x_hi, x_lo, ld_x, M, N_out, K_in, dW, ld_dw, (float*)workspace,
                         (size_t)workspace_bytes, accumulate, prec, (cudaStream_t)stream);
}

int eda_gemm_qkv(const void* a_hi, const void* a_lo, int lda, const void* w_hi, const void* w_lo, int ldw,
                 const float* bias, void* qk_hi, void* qk_lo, int ld_qk, void* vt_hi, void* vt_lo,
                 int B, int T, int Tp, int D, int n_heads, int prec, void* stream) {
  EDA_REQUIRE(D == eda::GBN && n_heads * 64 == D, "eda_gemm_qkv: D must be 256 with head_dim 64");
  EDA_REQUIRE(bias && qk_hi && qk_lo && vt_hi && vt_lo && B > 0 && T > 0 && Tp >= T && Tp % 8 == 0, "eda_gemm_qkv: bad arguments");
  EDA_REQUIRE(ld_qk >= 2 * D, "eda_gemm_qkv: ld_qk too small");
  eda::GemmParams p = {};
  p.bias = bias; p.Chi = (eda::bf16*)qk_hi; p.Clo = (eda::bf16*)qk_lo; p.lds = ld_qk;
  p.M = B * T; p.N = 3 * D;
  p.vt_hi = (eda::bf16*)vt_hi; p.vt_lo = (eda::bf16*)vt_lo; p.T = T; p.Tp = Tp; p.H = n_heads; p.q_scale = 0.125f;
  return eda::gemm_split(a_hi, a_lo, lda, w_hi, w_lo, ldw, p, D, eda::GEPI_QKV, prec, (cudaStream_t)stream);
}

}  // extern "C"
