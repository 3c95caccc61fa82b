// Multi-head self-attention BACKWARD on tcgen05 / TMEM / TMA (training step).
//
// Differentiates the attention core of nn.TransformerEncoderLayer as the reference runs it
// (models.py:193,223: no masks, padded frames are keys like any other):
//   S = q' k^T (q' = q / 8, already scaled by the QKV epilogue), P = softmax(S), O = P v
// given dO, the saved row log-sum-exp of the forward kernel and delta = rowsum(dO * O):
//   dP = dO v^T,  dS = P * (dP - delta),  dv = P^T dO,  dk = dS^T q',  dq = (dS k) / 8.
//
// One CTA owns 128 KEYS of one (batch, head) and walks the 128-query tiles (key-stationary, like
// FlashAttention-2's backward): dk, dv accumulate in TMEM over the whole walk, dq tiles are added
// to an fp32 matrix in HBM (one float4 reduction per 16 B; 4..16 key tiles add to a row).
// No operand is ever transposed in memory: every product whose contraction index is the row
// index of a stored tile (P^T dO, dS^T q', dS k, dO v^T with v stored as v^T) uses the MN-major
// form of the shared-memory descriptors on the SAME bytes the K-major products read.
//
//   warp 0     TMA producer: K_j, V_j once; Q_i, dO_i per query tile (hi + lo: 32 KB each)
//   warp 1     MMA issuer + TMEM owner: S (128 cols) | dP (128) | dV (64) | dK (64) | dQ (64)
//   warps 2-9  thread = (query row, 64-key half): P = exp2(S log2e - lse), dS = P (dP - delta)
//              -> bf16 into swizzled shared memory (the A operand of the three gradient products);
//              then dQ out of TMEM -> red.global.add.v4.f32; at the end dK / dV -> split bf16.
// Precision: S and dP are bf16x3 products of split operands (EDA_PREC_BF16X3) or single bf16;
// P and dS are rounded to bf16 once (their products take the hi and lo halves of the other
// operand), fp32 accumulation throughout.
#include <algorithm>
#include <type_traits>

#include "common.cuh"
#include "tc_common.cuh"

namespace eda {

using namespace tc;
typedef __nv_bfloat16 bf16;

int make_tmap_bf16_3d(CUtensorMap* m, const void* ptr, uint64_t d0, uint64_t d1, uint64_t d2, uint64_t s1_bytes,
                      uint64_t s2_bytes, uint32_t box_rows);

namespace {

constexpr int BT = 128, BDH = 64;
constexpr int B_TILE = BT * BDH * 2;                 // 16 KB: 128 rows x 64 bf16
constexpr int B_OFF_K = 0;                           // K_hi, K_lo      [key][d]
constexpr int B_OFF_V = 2 * B_TILE;                  // Vt_hi, Vt_lo    two {64 keys, 64 d} boxes each: [d][key]
constexpr int B_OFF_Q = 4 * B_TILE;                  // Q_hi, Q_lo      [q][d]
constexpr int B_OFF_DO = 6 * B_TILE;                 // dO_hi, dO_lo    [q][d]
constexpr int B_OFF_P = 8 * B_TILE;                  // P  (bf16): two 64-key blocks of [q][64 keys]
constexpr int B_OFF_DS = 10 * B_TILE;                // dS (bf16): same layout
constexpr int B_OFF_BAR = 12 * B_TILE;               // 192 KB of operands
constexpr int B_SMEM = B_OFF_BAR + 256 + 1024;
constexpr int B_THREADS = 320;
constexpr uint32_t TM_S = 0, TM_DP = 128, TM_DV = 256, TM_DK = 320, TM_DQ = 384, B_TMEM_COLS = 512;

struct MhaBwdParams {
  int T, H, B, passes;
  const float* lse;        // (B*H, T) log2-domain row log-sum-exp from the forward kernel
  const float* delta;      // (B*T, H)
  float* dq_acc;           // (B*T, H*64) fp32, zeroed by the caller: dq' = dS k summed over key tiles
  bf16* dkv_hi; bf16* dkv_lo; int ld_dkv;     // split (B*T, ld): dk at column dk_col0 + 64h, dv at dv_col0 + 64h
  int dk_col0, dv_col0;
  DropoutCfg drop;         // the forward's attention-weight dropout (thresh == 0 = off)
  const uint32_t* drop_bits; int ld_bits;     // optional: the same keep mask as bits (eda_dropout_bits)
};

__device__ __forceinline__ void tma_load_3d(void* smem_dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1, int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
      ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}
__device__ __forceinline__ uint32_t pack_bf16x2(float a, float b) {
  uint32_t r;
  asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(b), "f"(a));
  return r;
}
__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
constexpr uint32_t kAmn = 1u << 15, kBmn = 1u << 16;      // instruction-descriptor bits: operand is MN-major
// D[tmem] (+)= A[tmem] * B[smem desc]
__device__ __forceinline__ void umma_bf16_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
      ::"r"(tmem_d), "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void tmem_st16u(uint32_t taddr, const uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%16], "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15};"
      ::"r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]),
        "r"(r[8]), "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]),
        "r"(taddr)
      : "memory");
}

#ifdef EDA_MHABWD_TRACE
// probe build only (scripts/mha_bwd_trace.py): clock64 stamps of query tile 3 of every CTA, 16 slots:
// 0 TMA: qfree(2) passed = loads of tile 3 issued | 1 MMA: Q/dO landed | 2 S, dP issued | 3 P/dS arrived | 4 gradient MMAs issued
// 8 softmax warp 2: S/dP committed | 9 first 32 keys done | 10 P/dS stored + arrived | 11 dQ committed | 12 dQ read | 13 reds issued
__device__ long long g_mhabwd_trace[4096 * 16];
#define BWD_STAMP(slot) do { if (i == 3) g_mhabwd_trace[(size_t)blockIdx.x * 16 + (slot)] = clock64(); } while (0)
#else
#define BWD_STAMP(slot) do { } while (0)
#endif

__global__ void __launch_bounds__(B_THREADS, 1)
mha_bwd_tc_kernel(const __grid_constant__ CUtensorMap tmQK_hi, const __grid_constant__ CUtensorMap tmQK_lo,
                  const __grid_constant__ CUtensorMap tmV_hi, const __grid_constant__ CUtensorMap tmV_lo,
                  const __grid_constant__ CUtensorMap tmDO_hi, const __grid_constant__ CUtensorMap tmDO_lo,
                  const MhaBwdParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bar_kv = reinterpret_cast<uint64_t*>(smem + B_OFF_BAR);
  uint64_t* bar_q = bar_kv + 1;        // Q_i landed
  uint64_t* bar_qfree = bar_kv + 2;    // the dK products of tile i have read Q (and dS): Q_{i+1} may land
  uint64_t* bar_s = bar_kv + 3;        // S and dP of tile i are in TMEM
  uint64_t* bar_p = bar_kv + 4;        // P, dS of tile i are in shared memory (8 warps)
  uint64_t* bar_dq = bar_kv + 5;       // dQ of tile i (and every earlier MMA) complete
  uint64_t* bar_dqfree = bar_kv + 6;   // dQ of tile i has been read out of TMEM (8 warps)
  uint64_t* bar_dofree = bar_kv + 7;   // the dV products of tile i have read dO (and P): dO_{i+1} may land
  uint64_t* bar_do = bar_kv + 8;       // dO_i landed (its own barrier: dP = dO V^T is issued while Q_i is still in flight)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bar_kv + 10);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_t = (p.T + BT - 1) / BT;
  const int j = blockIdx.x % n_t, h = (blockIdx.x / n_t) % p.H, b = blockIdx.x / (n_t * p.H);
  const int k0 = j * BT;
  const bool x3 = p.passes == 3;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmQK_hi); tma_prefetch_desc(&tmQK_lo); tma_prefetch_desc(&tmV_hi); tma_prefetch_desc(&tmV_lo);
    tma_prefetch_desc(&tmDO_hi); tma_prefetch_desc(&tmDO_lo);
    mbar_init(bar_kv, 1); mbar_init(bar_q, 1); mbar_init(bar_qfree, 1); mbar_init(bar_s, 1);
    mbar_init(bar_p, 8); mbar_init(bar_dq, 1); mbar_init(bar_dqfree, 8); mbar_init(bar_dofree, 1); mbar_init(bar_do, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, B_TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===== TMA producer =====
    if (elect_one()) {
      mbar_expect_tx(bar_kv, x3 ? 4 * B_TILE : 2 * B_TILE);
      tma_load_3d(smem + B_OFF_K, &tmQK_hi, bar_kv, p.H * BDH + h * BDH, k0, b);
      tma_load_3d(smem + B_OFF_V, &tmV_hi, bar_kv, k0, 0, b * p.H + h);
      tma_load_3d(smem + B_OFF_V + B_TILE / 2, &tmV_hi, bar_kv, k0 + 64, 0, b * p.H + h);
      if (x3) {
        tma_load_3d(smem + B_OFF_K + B_TILE, &tmQK_lo, bar_kv, p.H * BDH + h * BDH, k0, b);
        tma_load_3d(smem + B_OFF_V + B_TILE, &tmV_lo, bar_kv, k0, 0, b * p.H + h);
        tma_load_3d(smem + B_OFF_V + B_TILE + B_TILE / 2, &tmV_lo, bar_kv, k0 + 64, 0, b * p.H + h);
      }
      // the gradient products run dV (reads dO), dK (reads Q), dQ (reads neither): the next tile's dO and Q are
      // requested as soon as their readers have completed, 2/3 and 1/3 of the gradient phase before its end
      // (the 64 KB of loads took ~1900 cycles per tile on the critical path when they waited for all of it)
      for (int i = 0; i < n_t; ++i) {
        if (i > 0) mbar_wait(bar_dofree, (i - 1) & 1);
        BWD_STAMP(0);
        mbar_expect_tx(bar_do, x3 ? 2 * B_TILE : B_TILE);
        tma_load_3d(smem + B_OFF_DO, &tmDO_hi, bar_do, h * BDH, i * BT, b);
        if (x3) tma_load_3d(smem + B_OFF_DO + B_TILE, &tmDO_lo, bar_do, h * BDH, i * BT, b);
        if (i > 0) mbar_wait(bar_qfree, (i - 1) & 1);
        mbar_expect_tx(bar_q, x3 ? 2 * B_TILE : B_TILE);
        tma_load_3d(smem + B_OFF_Q, &tmQK_hi, bar_q, h * BDH, i * BT, b);
        if (x3) tma_load_3d(smem + B_OFF_Q + B_TILE, &tmQK_lo, bar_q, h * BDH, i * BT, b);
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer =====
    if (elect_one()) {
      const uint32_t sb = smem_u32(smem);
      const uint32_t id_s = idesc_bf16_f32(BT, BT);                        // S  = Q K^T        (K-major, K-major)
      const uint32_t id_dp = idesc_bf16_f32(BT, BT) | kBmn;                // dP = dO V^T       (K-major, V^T rows are d: MN-major)
      const uint32_t id_g = idesc_bf16_f32(BT, BDH) | kAmn | kBmn;         // dV = P^T dO, dK = dS^T Q  (both MN-major)
      const uint32_t id_dq = idesc_bf16_f32(BT, BDH) | kBmn;               // dQ = dS K         (K-major, MN-major)
      const uint32_t tS = tmem_base + TM_S, tDP = tmem_base + TM_DP, tDV = tmem_base + TM_DV, tDK = tmem_base + TM_DK,
                     tDQ = tmem_base + TM_DQ;
      mbar_wait(bar_kv, 0);
      for (int i = 0; i < n_t; ++i) {
        mbar_wait(bar_do, i & 1);                            // dO arrives first: dP while Q is still in flight
        tc_fence_after();
#pragma unroll
        for (int k = 0; k < BDH / 16; ++k) {
          const uint32_t kk = k * 32, km = k * 2048;         // V^T tile: 16 d-rows of 128 B per step, key blocks 8 KB apart
          const uint64_t g_hi = smem_desc_sw128(sb + B_OFF_DO + kk), g_lo = smem_desc_sw128(sb + B_OFF_DO + B_TILE + kk);
          const uint64_t v_hi = smem_desc_sw128_mn(sb + B_OFF_V + km, 8192), v_lo = smem_desc_sw128_mn(sb + B_OFF_V + B_TILE + km, 8192);
          umma_bf16(tDP, g_hi, v_hi, id_dp, k != 0);
          if (x3) { umma_bf16(tDP, g_hi, v_lo, id_dp, 1); umma_bf16(tDP, g_lo, v_hi, id_dp, 1); }
        }
        mbar_wait(bar_q, i & 1);
        BWD_STAMP(1);
        tc_fence_after();
#pragma unroll
        for (int k = 0; k < BDH / 16; ++k) {                 // contraction over d: +32 B per step in K-major tiles
          const uint32_t kk = k * 32;
          const uint64_t q_hi = smem_desc_sw128(sb + B_OFF_Q + kk), q_lo = smem_desc_sw128(sb + B_OFF_Q + B_TILE + kk);
          const uint64_t k_hi = smem_desc_sw128(sb + B_OFF_K + kk), k_lo = smem_desc_sw128(sb + B_OFF_K + B_TILE + kk);
          umma_bf16(tS, q_hi, k_hi, id_s, k != 0);
          if (x3) { umma_bf16(tS, q_hi, k_lo, id_s, 1); umma_bf16(tS, q_lo, k_hi, id_s, 1); }
        }
        umma_commit(bar_s);
        BWD_STAMP(2);
        mbar_wait(bar_p, i & 1);
        if (i > 0) mbar_wait(bar_dqfree, (i - 1) & 1);
        BWD_STAMP(3);
        tc_fence_after();
#pragma unroll
        for (int k = 0; k < BT / 16; ++k) {                  // contraction over the 128 queries: 16 rows = 2048 B per step
          const uint32_t km = k * 2048;
          const uint64_t pa = smem_desc_sw128_mn(sb + B_OFF_P + km, 16384);
          const uint64_t g_hi = smem_desc_sw128_mn(sb + B_OFF_DO + km), g_lo = smem_desc_sw128_mn(sb + B_OFF_DO + B_TILE + km);
          umma_bf16(tDV, pa, g_hi, id_g, (i | k) != 0);
          if (x3) umma_bf16(tDV, pa, g_lo, id_g, 1);
        }
        umma_commit(bar_dofree);
#pragma unroll
        for (int k = 0; k < BT / 16; ++k) {
          const uint32_t km = k * 2048;
          const uint64_t da = smem_desc_sw128_mn(sb + B_OFF_DS + km, 16384);
          const uint64_t q_hi = smem_desc_sw128_mn(sb + B_OFF_Q + km), q_lo = smem_desc_sw128_mn(sb + B_OFF_Q + B_TILE + km);
          umma_bf16(tDK, da, q_hi, id_g, (i | k) != 0);
          if (x3) umma_bf16(tDK, da, q_lo, id_g, 1);
        }
        umma_commit(bar_qfree);
#pragma unroll
        for (int k = 0; k < BT / 16; ++k) {                  // contraction over the 128 keys
          const uint32_t da = tS + (k >> 2) * 64 + (k & 3) * 8;          // dS of keys 16k .. 16k+15: 8 packed columns
          const uint64_t k_hi = smem_desc_sw128_mn(sb + B_OFF_K + k * 2048), k_lo = smem_desc_sw128_mn(sb + B_OFF_K + B_TILE + k * 2048);
          umma_bf16_ts(tDQ, da, k_hi, id_dq, k != 0);
          if (x3) umma_bf16_ts(tDQ, da, k_lo, id_dq, 1);
        }
        umma_commit(bar_dq);
        BWD_STAMP(4);
      }
    }
  } else {
    // ===== warps 2..9: thread = (row of the tile, 64-column half) =====
    const int qd = warp & 3, half = (warp - 2) >> 2;
    const int r = qd * 32 + lane;                            // TMEM lane = query row (S, dP, dQ) or key row (dK, dV)
    const uint32_t lane_off = uint32_t(qd * 32) << 16;
    constexpr float kLog2e = 1.4426950408889634f;
    const uint32_t sP = smem_u32(smem + B_OFF_P) + half * 16384 + r * 128;
    const uint32_t sDS = smem_u32(smem + B_OFF_DS) + half * 16384 + r * 128;
    uint32_t keep_n0 = 0xffffffffu, keep_n1 = 0xffffffffu;
    auto load_keep = [&](int tile) {
      const int qn = tile * BT + r;
      if (p.drop_bits != nullptr && qn < p.T) {
        const uint32_t* bw = p.drop_bits + (size_t)((b * p.H + h) * p.T + qn) * p.ld_bits + ((k0 + half * 64) >> 5);
        if (k0 + half * 64 < p.T) keep_n0 = __ldg(bw);
        if (k0 + half * 64 + 32 < p.T) keep_n1 = __ldg(bw + 1);
      }
    };
    load_keep(0);
    for (int i = 0; i < n_t; ++i) {
      const int q = i * BT + r;
      const bool q_ok = q < p.T;
      float lse = 0.f, dl = 0.f;
      if (q_ok) {
        lse = __ldg(p.lse + ((size_t)b * p.H + h) * p.T + q);
        dl = __ldg(p.delta + ((size_t)b * p.T + q) * p.H + h);
      }
      // keep bits of this thread's 2 x 32 keys (precomputed masks), fetched one query tile ahead
      const uint32_t keep_w0 = keep_n0, keep_w1 = keep_n1;
      load_keep(i + 1);
      mbar_wait(bar_s, i & 1);
      if (threadIdx.x == 64) BWD_STAMP(8);
      tc_fence_after();
#pragma unroll 1
      for (int c = 0; c < 2; ++c) {                          // 32 keys at a time
        float sv[32], dv[32];
        tmem_ld32_issue(tmem_base + TM_S + lane_off + half * 64 + c * 32, sv);
        tmem_ld32_issue(tmem_base + TM_DP + lane_off + half * 64 + c * 32, dv);
        tmem_wait_ld();
        const int key0 = k0 + half * 64 + c * 32;
        uint32_t pw[16], dw[16];
        const uint32_t drop_row = (uint32_t)((b * p.H + h) * p.T + q);
        uint32_t kw = c ? keep_w1 : keep_w0;
        if (p.drop.thresh && p.drop_bits == nullptr) {
#pragma unroll
          for (int cc = 0; cc < 4; ++cc)
            kw = (cc == 0 ? 0u : kw) | dropout_keep8(p.drop, drop_row, (uint32_t)(key0 >> 3) + cc) << (8 * cc);
        }
        // P = exp2(S log2e - lse), dS = P (dP - delta); with dropout O = (D P / (1-p)) V: dV sees the dropped weights and
        // dP = D d(dropped) / (1-p).  Compiled four ways: the bounds checks only exist for the ragged last tiles and
        // the keep factors only with dropout (a quarter of this loop's instructions each).
        auto body = [&](auto full_c, auto drop_c) {
          constexpr bool kFull = decltype(full_c)::value, kDrop = decltype(drop_c)::value;
#pragma unroll
          for (int e = 0; e < 32; e += 2) {
            float p0 = ex2_approx(fmaf(sv[e], kLog2e, -lse)), p1 = ex2_approx(fmaf(sv[e + 1], kLog2e, -lse));
            if (!kFull) {
              if (!q_ok || key0 + e >= p.T) p0 = 0.f;
              if (!q_ok || key0 + e + 1 >= p.T) p1 = 0.f;
            }
            float pd0 = p0, pd1 = p1, g0 = dv[e], g1 = dv[e + 1];
            if (kDrop) {
              const float ks0 = (kw >> e) & 1u ? p.drop.scale : 0.f, ks1 = (kw >> (e + 1)) & 1u ? p.drop.scale : 0.f;
              pd0 = p0 * ks0; pd1 = p1 * ks1;
              g0 *= ks0; g1 *= ks1;
            }
            pw[e >> 1] = pack_bf16x2(pd0, pd1);
            dw[e >> 1] = pack_bf16x2(p0 * (g0 - dl), p1 * (g1 - dl));
          }
        };
        const bool full = __all_sync(0xffffffffu, q_ok && key0 + 32 <= p.T);
        using T_ = std::true_type; using F_ = std::false_type;
        if (p.drop.thresh) { if (full) body(T_{}, T_{}); else body(F_{}, T_{}); }
        else               { if (full) body(T_{}, F_{}); else body(F_{}, F_{}); }
#pragma unroll
        for (int u = 0; u < 4; ++u) {                        // 16 B units c*4 + u of the 128 B row, XOR-swizzled by row
          const uint32_t off = uint32_t(((c * 4 + u) ^ (r & 7)) << 4);
          asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(sP + off), "r"(pw[4 * u]), "r"(pw[4 * u + 1]), "r"(pw[4 * u + 2]), "r"(pw[4 * u + 3]) : "memory");
          asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(sDS + off), "r"(dw[4 * u]), "r"(dw[4 * u + 1]), "r"(dw[4 * u + 2]), "r"(dw[4 * u + 3]) : "memory");
        }
        // dS also stays in TMEM, packed two keys per column over the S columns this thread has already consumed
        // ([64 half, 64 half + 32)): it is the A operand of dQ = dS K, which then reads no shared memory for A
        tmem_st16u(tmem_base + TM_S + lane_off + half * 64 + c * 16, dw);
        if (threadIdx.x == 64 && c == 0) BWD_STAMP(9);
      }
      tmem_wait_st();
      fence_proxy_async();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_p);
      if (threadIdx.x == 64) BWD_STAMP(10);
      // dQ tile -> HBM accumulator
      mbar_wait(bar_dq, i & 1);
      if (threadIdx.x == 64) BWD_STAMP(11);
      tc_fence_after();
      float o[32];
      tmem_ld32(tmem_base + TM_DQ + lane_off + half * 32, o);
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_dqfree);
      if (threadIdx.x == 64) BWD_STAMP(12);
      if (q_ok) {
        float* dst = p.dq_acc + ((size_t)b * p.T + q) * (p.H * BDH) + h * BDH + half * 32;
#pragma unroll
        for (int e = 0; e < 32; e += 4)
          asm volatile("red.global.add.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(dst + e), "f"(o[e]), "f"(o[e + 1]), "f"(o[e + 2]), "f"(o[e + 3]) : "memory");
      }
      if (threadIdx.x == 64) BWD_STAMP(13);
    }
    // dK, dV of this key tile (complete: the last bar_dq covers every MMA issued before it)
    const int key = k0 + r;
    float ok[32], ov[32];
    tmem_ld32(tmem_base + TM_DK + lane_off + half * 32, ok);    // warp-collective loads: before the row guard
    tmem_ld32(tmem_base + TM_DV + lane_off + half * 32, ov);
    __syncwarp();
    if (key < p.T) {
      const size_t row = (size_t)b * p.T + key;
      auto put = [&](const float (&o)[32], int col) {
        uint32_t hh[16], ll[16];
#pragma unroll
        for (int e = 0; e < 32; e += 2) {
          hh[e >> 1] = pack_bf16x2(o[e], o[e + 1]);
          ll[e >> 1] = pack_bf16x2(o[e] - __uint_as_float(hh[e >> 1] << 16), o[e + 1] - __uint_as_float(hh[e >> 1] & 0xffff0000u));
        }
        uint4* dh = reinterpret_cast<uint4*>(p.dkv_hi + row * p.ld_dkv + col);
        uint4* dl4 = reinterpret_cast<uint4*>(p.dkv_lo + row * p.ld_dkv + col);
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          dh[u] = make_uint4(hh[4 * u], hh[4 * u + 1], hh[4 * u + 2], hh[4 * u + 3]);
          dl4[u] = make_uint4(ll[4 * u], ll[4 * u + 1], ll[4 * u + 2], ll[4 * u + 3]);
        }
      };
      put(ok, p.dk_col0 + h * BDH + half * 32);
      put(ov, p.dv_col0 + h * BDH + half * 32);
    }
    tc_fence_before();
  }
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, B_TMEM_COLS);
  }
}

}  // namespace

int mha_bwd(const void* qk_hi, const void* qk_lo, int ld_qk, const void* vt_hi, const void* vt_lo, int Tp,
            const void* do_hi, const void* do_lo, int ld_do, const float* lse, const float* delta, int B, int T, int H,
            float* dq_acc, void* dqkv_hi, void* dqkv_lo, int ld_dqkv, DropoutCfg drop, const uint32_t* drop_bits, int ld_bits,
            int prec, cudaStream_t st) {
  EDA_REQUIRE(!drop_bits || (drop.thresh && ld_bits * 32 >= T), "eda_mha_bwd: drop_bits needs p_drop > 0 and ld_bits >= ceil(T / 32)");
  EDA_REQUIRE(qk_hi && qk_lo && vt_hi && vt_lo && do_hi && do_lo && lse && delta && dq_acc && dqkv_hi && dqkv_lo, "eda_mha_bwd: null argument");
  EDA_REQUIRE(B > 0 && T > 0 && H > 0 && ld_qk >= 2 * H * BDH && ld_qk % 8 == 0 && Tp >= T && Tp % 8 == 0 && ld_do % 8 == 0 &&
              ld_dqkv >= 3 * H * BDH && ld_dqkv % 8 == 0, "eda_mha_bwd: bad sizes / strides");
  EDA_REQUIRE(((uintptr_t)dq_acc | (uintptr_t)dqkv_hi | (uintptr_t)dqkv_lo) % 16 == 0, "eda_mha_bwd: outputs must be 16B aligned");
  EDA_REQUIRE(prec == EDA_PREC_BF16X3 || prec == EDA_PREC_BF16, "eda_mha_bwd: precision must be BF16X3 or BF16");
  static DeviceOnce attr;
  if (!attr.done()) {
    EDA_CUDA(cudaFuncSetAttribute(mha_bwd_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, B_SMEM));
    attr.mark();
  }
  EDA_CUDA(cudaMemsetAsync(dq_acc, 0, (size_t)B * T * H * BDH * sizeof(float), st));
  CUtensorMap tq_hi, tq_lo, tv_hi, tv_lo, tg_hi, tg_lo;
  int rc;
  const uint64_t row_b = (uint64_t)ld_qk * 2, row_g = (uint64_t)ld_do * 2;
  if ((rc = make_tmap_bf16_3d(&tq_hi, qk_hi, 2 * H * BDH, T, B, row_b, row_b * T, BT)) != EDA_OK) return rc;
  if ((rc = make_tmap_bf16_3d(&tq_lo, qk_lo, 2 * H * BDH, T, B, row_b, row_b * T, BT)) != EDA_OK) return rc;
  if ((rc = make_tmap_bf16_3d(&tv_hi, vt_hi, T, BDH, (uint64_t)B * H, (uint64_t)Tp * 2, (uint64_t)Tp * 2 * BDH, 64)) != EDA_OK) return rc;
  if ((rc = make_tmap_bf16_3d(&tv_lo, vt_lo, T, BDH, (uint64_t)B * H, (uint64_t)Tp * 2, (uint64_t)Tp * 2 * BDH, 64)) != EDA_OK) return rc;
  if ((rc = make_tmap_bf16_3d(&tg_hi, do_hi, H * BDH, T, B, row_g, row_g * T, BT)) != EDA_OK) return rc;
  if ((rc = make_tmap_bf16_3d(&tg_lo, do_lo, H * BDH, T, B, row_g, row_g * T, BT)) != EDA_OK) return rc;
  MhaBwdParams p = {};
  p.T = T; p.H = H; p.B = B; p.passes = prec == EDA_PREC_BF16 ? 1 : 3;
  p.lse = lse; p.delta = delta; p.dq_acc = dq_acc;
  p.dkv_hi = (bf16*)dqkv_hi; p.dkv_lo = (bf16*)dqkv_lo; p.ld_dkv = ld_dqkv;
  p.dk_col0 = H * BDH; p.dv_col0 = 2 * H * BDH; p.drop = drop;
  p.drop_bits = drop.thresh ? drop_bits : nullptr; p.ld_bits = ld_bits;
  const int64_t n_cta = (int64_t)ceil_div(T, BT) * H * B;
  EDA_REQUIRE(n_cta < (int64_t(1) << 31), "eda_mha_bwd: too many (batch, head, key tile) items");
  mha_bwd_tc_kernel<<<(int)n_cta, B_THREADS, B_SMEM, st>>>(tq_hi, tq_lo, tv_hi, tv_lo, tg_hi, tg_lo, p);
  EDA_LAUNCH_CHECK();
  return EDA_OK;
}

}  // namespace eda

#ifdef EDA_MHABWD_TRACE
extern "C" int eda_mha_bwd_trace_read(long long* host, int n_cta) {
  return cudaMemcpyFromSymbol(host, eda::g_mhabwd_trace, sizeof(long long) * 16 * (size_t)n_cta) == cudaSuccess ? 0 : 1;
}
#endif

extern "C" int eda_mha_bwd(const void* qk_hi, const void* qk_lo, int ld_qk, const void* vt_hi, const void* vt_lo, int Tp,
                           const void* do_hi, const void* do_lo, int ld_do, const float* lse, const float* delta, int B,
                           int T, int n_heads, float* dq_acc, void* dqkv_hi, void* dqkv_lo, int ld_dqkv, float p_drop,
                           uint64_t seed, int site, const uint32_t* drop_bits, int ld_bits, int prec, void* stream) {
  EDA_REQUIRE(p_drop >= 0.f && p_drop < 1.f, "eda_mha_bwd: bad dropout probability");
  return eda::mha_bwd(qk_hi, qk_lo, ld_qk, vt_hi, vt_lo, Tp, do_hi, do_lo, ld_do, lse, delta, B, T, n_heads, dq_acc,
                      dqkv_hi, dqkv_lo, ld_dqkv, eda::make_dropout(p_drop, seed, site), drop_bits, ld_bits, prec,
                      (cudaStream_t)stream);
}
