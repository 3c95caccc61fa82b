// Multi-head self-attention forward on tcgen05 / TMEM / TMA (EDA_PREC_BF16X3 / BF16).
//
// softmax(q k^T / sqrt(64)) v per (batch, head), NO mask: the reference attends to the
// -1-padded frames too (models.py:223).  Flash-style: one CTA owns 128 queries of one (b, h)
// and streams 64-key tiles.
//
// Operands are split-bf16 pairs written by the QKV GEMM epilogue: q (pre-scaled by 1/8) and k
// as K-major rows of a (B*T, 512) matrix, v TRANSPOSED per (b, h) as (64, Tp) so that keys are
// the contraction-major dimension of the second product.  Every product is three MMAs
// (hi.hi + hi.lo + lo.hi), fp32 accumulation in TMEM.
//
//   warp 0     TMA producer: Q once, then K_j / V_j tiles (single stage each; three CTAs share an
//              SM, so one CTA's softmax overlaps the others' MMAs)
//   warp 1     MMA issuer (one thread) + TMEM owner (128 columns: S/P 0..63, O 64..127)
//   warps 2-5  softmax: thread = query row.  S -> running max / sum, P = exp(S - m) (MUFU ex2)
//              split into bf16 hi/lo and written BACK INTO TMEM in place over the 32-column S
//              chunk it was computed from (16 columns of packed hi pairs, 16 of lo pairs; the
//              A operand of P.V comes from TMEM), O rescaled in TMEM when the running max
//              moves; final O / l -> split bf16 (and/or fp32), thread-per-row 32-byte stores.
#include <type_traits>
#include <algorithm>

#include "common.cuh"
#include "tc_common.cuh"

namespace eda {

using namespace tc;
typedef __nv_bfloat16 bf16;

int make_tmap_bf16_3d(CUtensorMap* m, const void* ptr, uint64_t d0, uint64_t d1, uint64_t d2, uint64_t s1_bytes,
                      uint64_t s2_bytes, uint32_t box_rows);

// 64-key tiles with TWO S/P buffers in TMEM: the MMA thread issues S(j+1) = Q.K(j+1)^T before it
// waits for P(j), so the tensor core works on the next scores while the softmax warps turn the
// current ones into P; per key tile the chain is max(softmax, MMA) instead of their sum.  (Three
// CTAs per SM with one buffer each did not help: the per-tile latency chain was the limit.)
constexpr int ABM = 128, ABN = 64, ADH = 64, A_CTAS_PER_SM = 2;
constexpr int A_TILE16 = 128 * 64 * 2;                   // 16 KB: Q, 128 rows x 64 cols bf16
constexpr int A_KTILE = ABN * 64 * 2;                    // 8 KB: K (ABN keys x 64 d) or Vt (64 d x ABN keys)
constexpr int A_OFF_Q = 0;                               // Q_hi, Q_lo
constexpr int A_OFF_K = 2 * A_TILE16;                    // K_hi, K_lo
constexpr int A_KV_STAGES = 2;                           // K / V rings: tile j+1 is in flight while j is used
constexpr int A_KV_STAGE = 2 * A_KTILE;                  // hi + lo
constexpr int A_OFF_V = A_OFF_K + A_KV_STAGES * A_KV_STAGE;          // Vt_hi, Vt_lo
constexpr int A_OFF_BAR = A_OFF_V + A_KV_STAGES * A_KV_STAGE;
constexpr int A_OFF_MX = A_OFF_BAR + 256;                 // row maxima exchanged between the two halves of a row
constexpr int A_OFF_L = A_OFF_MX + 2 * 2 * 128 * 4;       // and their partial row sums, once at the end
constexpr int A_SMEM = A_OFF_L + 2 * 128 * 4 + 1024;
constexpr int A_THREADS = 320;                           // producer warp, MMA warp, 8 softmax warps
constexpr uint32_t TM_S = 0, TM_O = 2 * ABN, A_TMEM_COLS = 256;        // S0 | S1 | O
static_assert(ABN == 64, "the V tile is one 64-key box and P is written in 32-column chunks");

struct MhaParams {
  int T, H, B, passes;
  float* out; int ld_out;                 // optional fp32 (B*T, H*64)
  bf16* out_hi; bf16* out_lo; int ld_os;  // optional split
  float* lse;                             // optional (training): (B*H, T) log2-domain log-sum-exp of each score row
  const int32_t* klens;                   // optional key_padding_mask (Conformer): keys t >= klens[b] are masked out
  DropoutCfg drop;                        // training: dropout on the attention weights (thresh == 0 = off)
  const uint32_t* drop_bits; int ld_bits; // optional: the keep mask of this site as bits (eda_dropout_bits), row (b*H+h)*T + q
};

// D[tmem] (+)= A[tmem] * B[smem desc]
__device__ __forceinline__ void umma_bf16_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc,
                                             uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
      ::"r"(tmem_d), "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%16], "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15};"
      ::"r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]),
        "r"(r[8]), "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]),
        "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tma_load_3d(void* smem_dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1, int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
      ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}
// two fp32 -> packed bf16 pair (a in the low half = element 2c), round to nearest even
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
__device__ __forceinline__ void st_global_v8(void* p, uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3,
                                             uint32_t a4, uint32_t a5, uint32_t a6, uint32_t a7) {
  asm volatile("st.global.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};"
               ::"l"(p), "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(a4), "r"(a5), "r"(a6), "r"(a7) : "memory");
}
__device__ __forceinline__ uint32_t pack_bf16(bf16 a, bf16 b) {   // a -> low half (element 2c), b -> high
  __nv_bfloat162 v = __halves2bfloat162(a, b);
  return *reinterpret_cast<uint32_t*>(&v);
}

#ifdef EDA_MHA_TRACE
// probe build only (scripts/mha_trace.py): clock64 stamps of each CTA's FIRST work item, 96 slots:
//   [0] smid  [1] start  [2] end  [3] items done   MMA thread: [8+4j..] S issued, P ready, V ready, PV issued
//   softmax warp 2: [48+2j] S ready, [49+2j] P arrived;  [80] O ready
__device__ long long g_mha_trace[4096 * 96];
#define MHA_STAMP(slot) do { if (it == 0) g_mha_trace[(size_t)blockIdx.x * 96 + (slot)] = clock64(); } while (0)
#else
#define MHA_STAMP(slot)
#endif

// Persistent: A_CTAS_PER_SM CTAs per SM, each walking the (batch, head, query tile) items
// blockIdx.x, blockIdx.x + gridDim.x, ...  Barriers, the TMEM allocation and the K/V rings live
// across items: the producer prefetches the next item's Q and first K/V tiles while the current
// item's last tiles and epilogue run, so an item costs its key tiles and nothing else (a CTA per
// item spent a quarter of its life in launch latency, setup and waiting for its first loads).
__global__ void __launch_bounds__(A_THREADS, A_CTAS_PER_SM)
mha_tc_kernel(const __grid_constant__ CUtensorMap tmQK_hi, const __grid_constant__ CUtensorMap tmQK_lo,
              const __grid_constant__ CUtensorMap tmK_hi, const __grid_constant__ CUtensorMap tmK_lo,
              const __grid_constant__ CUtensorMap tmV_hi, const __grid_constant__ CUtensorMap tmV_lo,
              const MhaParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bar_q = reinterpret_cast<uint64_t*>(smem + A_OFF_BAR);     // Q of the item has landed
  uint64_t* bar_s = bar_q + 1;        // [2]: S(g) complete in buffer g & 1   (g = key tiles since kernel start)
  uint64_t* bar_p = bar_q + 3;        // [2]: P(g) written over it by the 8 softmax warps
  uint64_t* bar_o = bar_q + 5;        // O of the item is complete
  uint64_t* bar_pv = bar_q + 6;       // P(g).V(g) has been accumulated into O
  uint64_t* q_empty = bar_q + 7;      // the item's S MMAs have all read Q
  uint64_t* o_free = bar_q + 8;       // the item's epilogue has read O out of TMEM
  uint64_t* k_full = bar_q + 9;       // [A_KV_STAGES] each
  uint64_t* k_empty = k_full + A_KV_STAGES;
  uint64_t* v_full = k_empty + A_KV_STAGES;
  uint64_t* v_empty = v_full + A_KV_STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(v_empty + A_KV_STAGES);
  float (*s_mx)[2][128] = reinterpret_cast<float (*)[2][128]>(smem + A_OFF_MX);     // [tile parity][half][row]
  float (*s_l)[128] = reinterpret_cast<float (*)[128]>(smem + A_OFF_L);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_kt_all = (p.T + ABN - 1) / ABN, n_qt = (p.T + ABM - 1) / ABM;
  const int n_items = n_qt * p.H * p.B;
  // keys of batch entry b that take part (all T without a mask) and the 64-key tiles they span
  auto keys_of = [&](int b) { return p.klens ? min(p.T, max(1, __ldg(p.klens + b))) : p.T; };
  auto kt_of = [&](int b) { return p.klens ? (keys_of(b) + ABN - 1) / ABN : n_kt_all; };
#ifdef EDA_MHA_TRACE
  if (threadIdx.x == 0) {
    uint32_t smid;
    asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
    g_mha_trace[(size_t)blockIdx.x * 96] = smid;
    g_mha_trace[(size_t)blockIdx.x * 96 + 1] = clock64();
  }
#endif

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmQK_hi); tma_prefetch_desc(&tmQK_lo); tma_prefetch_desc(&tmK_hi); tma_prefetch_desc(&tmK_lo);
    tma_prefetch_desc(&tmV_hi); tma_prefetch_desc(&tmV_lo);
    mbar_init(bar_q, 1); mbar_init(q_empty, 1); mbar_init(o_free, 8);
    for (int s = 0; s < A_KV_STAGES; ++s) {
      mbar_init(&k_full[s], 1); mbar_init(&k_empty[s], 1); mbar_init(&v_full[s], 1); mbar_init(&v_empty[s], 1);
    }
    mbar_init(&bar_s[0], 1); mbar_init(&bar_s[1], 1); mbar_init(&bar_p[0], 8); mbar_init(&bar_p[1], 8);
    mbar_init(bar_o, 1); mbar_init(bar_pv, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, A_TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const bool x3 = p.passes == 3;

  if (warp == 0) {
    // ===== TMA producer =====
    if (elect_one()) {
      uint32_t g = 0;
      int it = 0;
      for (int item = blockIdx.x; item < n_items; item += gridDim.x, ++it) {
        const int q0 = (item % n_qt) * ABM, h = (item / n_qt) % p.H, b = item / (n_qt * p.H);
        const int n_kt = kt_of(b);
        mbar_wait(q_empty, (it & 1) ^ 1);
        mbar_expect_tx(bar_q, x3 ? 2 * A_TILE16 : A_TILE16);
        tma_load_3d(smem + A_OFF_Q, &tmQK_hi, bar_q, h * ADH, q0, b);
        if (x3) tma_load_3d(smem + A_OFF_Q + A_TILE16, &tmQK_lo, bar_q, h * ADH, q0, b);
        for (int j = 0; j < n_kt; ++j, ++g) {
          const int s = g % A_KV_STAGES;
          const uint32_t ph = (g / A_KV_STAGES) & 1;
          uint8_t* ks = smem + A_OFF_K + s * A_KV_STAGE;
          uint8_t* vs = smem + A_OFF_V + s * A_KV_STAGE;
          mbar_wait(&k_empty[s], ph ^ 1);
          mbar_expect_tx(&k_full[s], x3 ? 2 * A_KTILE : A_KTILE);
          tma_load_3d(ks, &tmK_hi, &k_full[s], p.H * ADH + h * ADH, j * ABN, b);
          if (x3) tma_load_3d(ks + A_KTILE, &tmK_lo, &k_full[s], p.H * ADH + h * ADH, j * ABN, b);
          mbar_wait(&v_empty[s], ph ^ 1);
          mbar_expect_tx(&v_full[s], x3 ? 2 * A_KTILE : A_KTILE);
          tma_load_3d(vs, &tmV_hi, &v_full[s], j * ABN, 0, b * p.H + h);
          if (x3) tma_load_3d(vs + A_KTILE, &tmV_lo, &v_full[s], j * ABN, 0, b * p.H + h);
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer (one thread; elect.sync lets ptxas emit the uniform-datapath MMAs without a
    // per-instruction active-lane loop, which a plain lane == 0 branch costs) =====
    if (elect_one()) {
      const uint32_t idesc_s = idesc_bf16_f32(ABM, ABN), idesc_o = idesc_bf16_f32(ABM, ADH);
      const uint32_t sbase = smem_u32(smem);
      const uint64_t q_hi = smem_desc_sw128(sbase + A_OFF_Q), q_lo = smem_desc_sw128(sbase + A_OFF_Q + A_TILE16);
      const uint32_t tS = tmem_base + TM_S, tO = tmem_base + TM_O;
      uint32_t gs = 0, gp = 0;                     // key tiles whose S / P.V has been issued
      auto issue_s = [&]() {                       // S(gs) = Q K^T into buffer gs & 1
        const int s = gs % A_KV_STAGES;
        mbar_wait(&k_full[s], (gs / A_KV_STAGES) & 1);
        tc_fence_after();
        const uint64_t k_hi = smem_desc_sw128(sbase + A_OFF_K + s * A_KV_STAGE), k_lo = smem_desc_sw128(sbase + A_OFF_K + s * A_KV_STAGE + A_KTILE);
        const uint32_t ts = tS + (gs & 1) * ABN;
#pragma unroll
        for (int k = 0; k < ADH / 16; ++k) {
          const uint64_t adv = uint64_t(k * 32 >> 4);
          umma_bf16(ts, q_hi + adv, k_hi + adv, idesc_s, k != 0);
          if (x3) {
            umma_bf16(ts, q_hi + adv, k_lo + adv, idesc_s, 1);
            umma_bf16(ts, q_lo + adv, k_hi + adv, idesc_s, 1);
          }
        }
        umma_commit(&k_empty[s]);
        umma_commit(&bar_s[gs & 1]);
        ++gs;
      };
      int it = 0;
      for (int item = blockIdx.x; item < n_items; item += gridDim.x, ++it) {
        const int n_kt = kt_of(item / (n_qt * p.H));
        mbar_wait(bar_q, it & 1);
        issue_s();
        for (int j = 0; j < n_kt; ++j, ++gp) {
          // S(g+1) overwrites the buffer P(g-1) lived in: P(g-1).V(g-1) was issued before it, and
          // MMAs of one thread execute in issue order
          if (j + 1 < n_kt) issue_s();
          else umma_commit(q_empty);                     // every S of the item is issued: Q may be replaced once they finish
          MHA_STAMP(8 + 4 * j);
          mbar_wait(&bar_p[gp & 1], (gp >> 1) & 1);     // P(g) is in TMEM (over S(g)), O rescaled if needed
          MHA_STAMP(9 + 4 * j);
          const int s = gp % A_KV_STAGES;
          mbar_wait(&v_full[s], (gp / A_KV_STAGES) & 1);
          if (j == 0) mbar_wait(o_free, (it & 1) ^ 1);   // the previous item's O has been read out
          MHA_STAMP(10 + 4 * j);
          tc_fence_after();
          const uint32_t tPhi = tS + (gp & 1) * ABN;
#pragma unroll
          for (int k = 0; k < ABN / 16; ++k) {
            const uint32_t voff = A_OFF_V + s * A_KV_STAGE + k * 32;                    // 16 keys = 32 B along the key-major rows
            const uint64_t v_hi = smem_desc_sw128(sbase + voff), v_lo = smem_desc_sw128(sbase + voff + A_KTILE);
            const uint32_t pa = tPhi + (k >> 1) * 32 + (k & 1) * 8;    // hi pairs; lo pairs 16 columns on
            umma_bf16_ts(tO, pa, v_hi, idesc_o, (j | k) != 0);
            if (x3) {
              umma_bf16_ts(tO, pa, v_lo, idesc_o, 1);
              umma_bf16_ts(tO, pa + 16, v_hi, idesc_o, 1);
            }
          }
          umma_commit(&v_empty[s]);
          umma_commit(bar_pv);
          if (j == n_kt - 1) umma_commit(bar_o);
          MHA_STAMP(11 + 4 * j);
        }
      }
    }
  } else {
    // ===== softmax + output: warps 2..9.  Two threads per query row: warp w and w + 4 own the same
    // 32 TMEM lanes (lane quarter = warp & 3) and one 32-key half of the tile each, so the row's
    // exp / split chain is half as long and twice as many warps hide its latency.  The halves only
    // share the row maximum (through shared memory, one 64-thread named barrier per tile).
    const int qd = warp & 3, half = (warp - 2) >> 2;
    const int row_in_tile = qd * 32 + lane;
    const uint32_t lane_off = uint32_t(qd * 32) << 16;
    const uint32_t tS0 = tmem_base + TM_S + lane_off + half * 32, tO = tmem_base + TM_O + lane_off + half * 32;
    constexpr float kLog2e = 1.4426950408889634f;
    // kLazy: the running reference maximum only moves when some row of the warp outgrows it by more
    // than 2^kLazy; until then P = exp(S - m_ref) <= 2^kLazy, harmless in the fp32 accumulators, and
    // O / l is unchanged by the common factor.  Saves the O read-modify-write of nearly every tile.
    constexpr float kLazy = 8.0f;
    auto pair_sync = [&]() { asm volatile("bar.sync %0, 64;" ::"r"(1 + qd) : "memory"); };
    uint32_t g = 0;                             // key tiles since kernel start (buffer / phase bookkeeping)
    int it = 0;
    for (int item = blockIdx.x; item < n_items; item += gridDim.x, ++it) {
    const int q0 = (item % n_qt) * ABM, h = (item / n_qt) % p.H, b = item / (n_qt * p.H);
    const int n_kt = kt_of(b), n_keys = keys_of(b);
    float m_ref = -INFINITY, l_sum = 0.f;       // m_ref in log2 units (S * log2 e)
    uint32_t keep_next = 0xffffffffu;
    if (p.drop_bits != nullptr && q0 + row_in_tile < p.T && half * 32 < p.T)
      keep_next = __ldg(p.drop_bits + (size_t)((b * p.H + h) * p.T + q0 + row_in_tile) * p.ld_bits + half);
    for (int j = 0; j < n_kt; ++j, ++g) {
      const uint32_t tS = tS0 + (g & 1) * ABN;
      // keep bits of this thread's 32 keys (precomputed masks: one load instead of 4 Philox evaluations in the
      // softmax warps), fetched one key tile ahead so that the load never sits on the tile's critical path
      uint32_t keep_w = keep_next;
      if (p.drop_bits != nullptr && q0 + row_in_tile < p.T && (j + 1) * ABN + half * 32 < p.T)
        keep_next = __ldg(p.drop_bits + (size_t)((b * p.H + h) * p.T + q0 + row_in_tile) * p.ld_bits + (((j + 1) * ABN + half * 32) >> 5));
      mbar_wait(&bar_s[g & 1], (g >> 1) & 1);
#ifdef EDA_MHA_TRACE
      if (threadIdx.x == 64) MHA_STAMP(48 + 2 * j);
#endif
      tc_fence_after();
      const int n_valid = min(32, n_keys - j * ABN - half * 32);  // keys of this half that exist (may be <= 0)
      const bool whole = n_valid == 32;                           // warp-uniform
      float v[32];                                                // the scores stay in registers: one TMEM read
      tmem_ld32(tS, v);
      float mx4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
      if (whole) {
#pragma unroll
        for (int i = 0; i < 32; ++i) mx4[i & 3] = fmaxf(mx4[i & 3], v[i]);
      } else {
#pragma unroll
        for (int i = 0; i < 32; ++i) mx4[i & 3] = fmaxf(mx4[i & 3], i < n_valid ? v[i] : -INFINITY);
      }
      float mx = fmaxf(fmaxf(mx4[0], mx4[1]), fmaxf(mx4[2], mx4[3])) * kLog2e;
      s_mx[g & 1][half][row_in_tile] = mx;
      pair_sync();
      mx = fmaxf(mx, s_mx[g & 1][half ^ 1][row_in_tile]);         // finite: key 0 of every tile exists
      if (j == 0) {
        m_ref = mx;
      } else if (__any_sync(0xffffffffu, mx > m_ref + kLazy)) {   // same rows, same verdict in both warps of the pair
        // O *= 2^(m_old - m_new) on this half's 32 columns; P(j-1).V must have landed before O is touched
        mbar_wait(bar_pv, (g - 1) & 1);
        tc_fence_after();
        const float m_new = fmaxf(m_ref, mx);
        const float alpha = ex2_approx(m_ref - m_new);
        l_sum *= alpha;
        m_ref = m_new;
        float o[32];
        tmem_ld32(tO, o);
#pragma unroll
        for (int i = 0; i < 32; ++i) o[i] *= alpha;
        tmem_st32(tO, o);
      }
      // P = exp(S - m) -> split bf16, written IN PLACE over the 32 S columns it came from:
      // packed hi pairs in the first 16 columns of the chunk, lo pairs in the last 16
      float l4[4] = {0.f, 0.f, 0.f, 0.f};
      float pk[32];
      const uint32_t drop_row = (uint32_t)((b * p.H + h) * p.T + q0 + row_in_tile);
      if (p.drop.thresh && p.drop_bits == nullptr) {
#pragma unroll
        for (int c = 0; c < 4; ++c)
          keep_w = (c == 0 ? 0u : keep_w) | dropout_keep8(p.drop, drop_row, (uint32_t)((j * ABN + half * 32) >> 3) + c) << (8 * c);
      }
      // compiled four ways (both conditions are warp-uniform): the bounds checks only exist for a ragged last key
      // tile, the keep factors only in train() mode with dropout
      auto body = [&](auto whole_c, auto drop_c) {
        constexpr bool kWhole = decltype(whole_c)::value, kDrop = decltype(drop_c)::value;
#pragma unroll
        for (int i = 0; i < 32; i += 2) {
          float p0 = ex2_approx(fmaf(v[i], kLog2e, -m_ref)), p1 = ex2_approx(fmaf(v[i + 1], kLog2e, -m_ref));
          if (!kWhole) {
            if (i >= n_valid) p0 = 0.f;
            if (i + 1 >= n_valid) p1 = 0.f;
          }
          l4[i & 2] += p0; l4[(i & 2) + 1] += p1;
          if (kDrop) {               // nn.MultiheadAttention drops normalised weights: the row sum keeps every term
            p0 *= (keep_w >> i) & 1u ? p.drop.scale : 0.f;
            p1 *= (keep_w >> (i + 1)) & 1u ? p.drop.scale : 0.f;
          }
          const uint32_t hi = pack_bf16x2(p0, p1);
          pk[i >> 1] = __uint_as_float(hi);
          pk[16 + (i >> 1)] = __uint_as_float(pack_bf16x2(p0 - __uint_as_float(hi << 16),
                                                          p1 - __uint_as_float(hi & 0xffff0000u)));
        }
      };
      using T_ = std::true_type; using F_ = std::false_type;
      if (p.drop.thresh) { if (whole) body(T_{}, T_{}); else body(F_{}, T_{}); }
      else               { if (whole) body(T_{}, F_{}); else body(F_{}, F_{}); }
      tmem_st32(tS, pk);
      l_sum += (l4[0] + l4[1]) + (l4[2] + l4[3]);
      tmem_wait_st();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&bar_p[g & 1]);
#ifdef EDA_MHA_TRACE
      if (threadIdx.x == 64) MHA_STAMP(49 + 2 * j);
#endif
    }
    // ---- epilogue: O / l; thread = (row, 32-column half): 128 B fp32 / 64 B per bf16 half -----
    s_l[half][row_in_tile] = l_sum;
    pair_sync();
    l_sum += s_l[half ^ 1][row_in_tile];
    if (p.lse && half == 0 && q0 + row_in_tile < p.T)      // softmax row = exp2(S log2e - lse): what the backward pass rebuilds P from
      p.lse[((size_t)b * p.H + h) * p.T + q0 + row_in_tile] = m_ref + log2f(l_sum);
    mbar_wait(bar_o, it & 1);
#ifdef EDA_MHA_TRACE
    if (threadIdx.x == 64) MHA_STAMP(80);
#endif
    tc_fence_after();
    const float inv = 1.0f / l_sum;
    const int q = q0 + row_in_tile;
    const size_t row = (size_t)b * p.T + q;
    {
      float o[32];
      tmem_ld32(tO, o);
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(o_free);          // the next item's first P.V may overwrite O
      if (q < p.T) {
#pragma unroll
        for (int i = 0; i < 32; ++i) o[i] *= inv;
        const int col = h * ADH + half * 32;
        if (p.out) {
          float* dst = p.out + row * p.ld_out + col;
          if (((uintptr_t)dst & 31) == 0) {
#pragma unroll
            for (int i = 0; i < 32; i += 8)
              st_global_v8(dst + i, __float_as_uint(o[i]), __float_as_uint(o[i + 1]), __float_as_uint(o[i + 2]),
                           __float_as_uint(o[i + 3]), __float_as_uint(o[i + 4]), __float_as_uint(o[i + 5]),
                           __float_as_uint(o[i + 6]), __float_as_uint(o[i + 7]));
          } else {
#pragma unroll
            for (int i = 0; i < 32; i += 2) *reinterpret_cast<float2*>(dst + i) = make_float2(o[i], o[i + 1]);
          }
        }
        if (p.out_hi) {
          uint32_t hh[16], ll[16];
#pragma unroll
          for (int i = 0; i < 32; i += 2) {
            hh[i >> 1] = pack_bf16x2(o[i], o[i + 1]);
            ll[i >> 1] = pack_bf16x2(o[i] - __uint_as_float(hh[i >> 1] << 16),
                                     o[i + 1] - __uint_as_float(hh[i >> 1] & 0xffff0000u));
          }
          bf16* dh = p.out_hi + row * p.ld_os + col;
          bf16* dl = p.out_lo + row * p.ld_os + col;
          if ((((uintptr_t)dh | (uintptr_t)dl) & 31) == 0) {
            st_global_v8(dh, hh[0], hh[1], hh[2], hh[3], hh[4], hh[5], hh[6], hh[7]);
            st_global_v8(dh + 16, hh[8], hh[9], hh[10], hh[11], hh[12], hh[13], hh[14], hh[15]);
            st_global_v8(dl, ll[0], ll[1], ll[2], ll[3], ll[4], ll[5], ll[6], ll[7]);
            st_global_v8(dl + 16, ll[8], ll[9], ll[10], ll[11], ll[12], ll[13], ll[14], ll[15]);
          } else {
#pragma unroll
            for (int i = 0; i < 16; ++i) {
              *reinterpret_cast<uint32_t*>(dh + 2 * i) = hh[i];
              *reinterpret_cast<uint32_t*>(dl + 2 * i) = ll[i];
            }
          }
        }
      }
    }
    }   // items
    tc_fence_before();
  }
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, A_TMEM_COLS);
  }
#ifdef EDA_MHA_TRACE
  if (threadIdx.x == 0) {
    g_mha_trace[(size_t)blockIdx.x * 96 + 2] = clock64();
    g_mha_trace[(size_t)blockIdx.x * 96 + 3] = (n_items - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;
  }
#endif
}

// ------------------------------------------------------------------------- host ---------
int mha_split(const void* qk_hi, const void* qk_lo, int ld_qk, const void* vt_hi, const void* vt_lo, int Tp,
              int B, int T, int H, float* out, int ld_out, void* out_hi, void* out_lo, int ld_os, int prec,
              cudaStream_t st, float* lse = nullptr, const int32_t* klens = nullptr, DropoutCfg drop = DropoutCfg{},
              const uint32_t* drop_bits = nullptr, int ld_bits = 0) {
  EDA_REQUIRE(qk_hi && qk_lo && vt_hi && vt_lo && B > 0 && T > 0 && H > 0, "eda_mha_split: bad arguments");
  EDA_REQUIRE(ld_qk >= 2 * H * ADH && ld_qk % 8 == 0 && Tp >= T && Tp % 8 == 0, "eda_mha_split: ld_qk / Tp must be multiples of 8 and large enough");
  EDA_REQUIRE(out || (out_hi && out_lo), "eda_mha_split: no output requested");
  EDA_REQUIRE(!out || (ld_out % 2 == 0 && (uintptr_t)out % 8 == 0), "eda_mha_split: fp32 output misaligned");
  EDA_REQUIRE(!out_hi || (ld_os % 2 == 0), "eda_mha_split: split output stride must be even");
  EDA_REQUIRE(prec == EDA_PREC_BF16X3 || prec == EDA_PREC_BF16, "eda_mha_split: precision must be BF16X3 or BF16");
  static DeviceOnce attr;
  if (!attr.done()) {
    EDA_CUDA(cudaFuncSetAttribute(mha_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, A_SMEM));
    attr.mark();
  }
  CUtensorMap tq_hi, tq_lo, tk_hi, tk_lo, tv_hi, tv_lo;
  int rc;
  const uint64_t row_b = (uint64_t)ld_qk * 2;
  if ((rc = make_tmap_bf16_3d(&tq_hi, qk_hi, 2 * H * ADH, T, B, row_b, row_b * T, 128)) != EDA_OK) return rc;
  if ((rc = make_tmap_bf16_3d(&tq_lo, qk_lo, 2 * H * ADH, T, B, row_b, row_b * T, 128)) != EDA_OK) return rc;
  if ((rc = make_tmap_bf16_3d(&tk_hi, qk_hi, 2 * H * ADH, T, B, row_b, row_b * T, ABN)) != EDA_OK) return rc;
  if ((rc = make_tmap_bf16_3d(&tk_lo, qk_lo, 2 * H * ADH, T, B, row_b, row_b * T, ABN)) != EDA_OK) return rc;
  if ((rc = make_tmap_bf16_3d(&tv_hi, vt_hi, T, ADH, (uint64_t)B * H, (uint64_t)Tp * 2, (uint64_t)Tp * 2 * ADH, 64)) != EDA_OK) return rc;
  if ((rc = make_tmap_bf16_3d(&tv_lo, vt_lo, T, ADH, (uint64_t)B * H, (uint64_t)Tp * 2, (uint64_t)Tp * 2 * ADH, 64)) != EDA_OK) return rc;
  MhaParams p = {};
  p.T = T; p.H = H; p.B = B; p.passes = prec == EDA_PREC_BF16 ? 1 : 3;
  p.out = out; p.ld_out = ld_out; p.out_hi = (bf16*)out_hi; p.out_lo = (bf16*)out_lo; p.ld_os = ld_os;
  p.lse = lse; p.klens = klens; p.drop = drop;
  EDA_REQUIRE(!drop_bits || (drop.thresh && ld_bits * 32 >= T), "eda_mha_split_train: drop_bits needs p_drop > 0 and ld_bits >= ceil(T / 32)");
  p.drop_bits = drop.thresh ? drop_bits : nullptr; p.ld_bits = ld_bits;
  const int64_t n_items = (int64_t)ceil_div(T, ABM) * H * B;
  EDA_REQUIRE(n_items < (int64_t(1) << 31), "eda_mha_split: too many (batch, head, query tile) items");
  const int grid = (int)std::min<int64_t>(n_items, (int64_t)A_CTAS_PER_SM * num_sms());
  mha_tc_kernel<<<grid, A_THREADS, A_SMEM, st>>>(tq_hi, tq_lo, tk_hi, tk_lo, tv_hi, tv_lo, p);
  EDA_LAUNCH_CHECK();
  return EDA_OK;
}

// fp32 qkv (B*T, 3*D) -> the split operand layout (eda_mha_fwd's compatibility path)
__global__ void qkv_split_kernel(const float* __restrict__ qkv, int ld_qkv, int B, int T, int H, int Tp,
                                 bf16* __restrict__ qk_hi, bf16* __restrict__ qk_lo,
                                 bf16* __restrict__ vt_hi, bf16* __restrict__ vt_lo) {
  const int D = H * ADH;
  const int64_t n = (int64_t)B * T * 3 * D;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t row = i / (3 * D);
    const int col = (int)(i - row * 3 * D);
    float x = qkv[row * ld_qkv + col];
    if (col < D) x *= 0.125f;
    const bf16 hv = __float2bfloat16_rn(x), lv = __float2bfloat16_rn(x - __bfloat162float(hv));
    if (col < 2 * D) {
      qk_hi[row * 2 * D + col] = hv;
      qk_lo[row * 2 * D + col] = lv;
    } else {
      const int b = (int)(row / T), t = (int)(row - (int64_t)b * T), hd = col - 2 * D;
      const size_t o = ((size_t)b * D + hd) * Tp + t;     // (b*H + h)*64 + d == b*D + hd
      vt_hi[o] = hv;
      vt_lo[o] = lv;
    }
  }
}

int mha_tc(const float* qkv, int ld_qkv, float* out, int ld_out, int B, int T, int H, int prec, cudaStream_t st) {
  const int D = H * ADH, Tp = (T + 7) / 8 * 8;
  const size_t qk_el = (size_t)B * T * 2 * D, vt_el = (size_t)B * D * Tp;
  const size_t need = 2 * qk_el + 2 * vt_el;
  bf16* scratch = (bf16*)stream_scratch(3, st, need * sizeof(bf16));      // per (device, stream)
  if (!scratch) return EDA_ECUDA;
  bf16 *qk_hi = scratch, *qk_lo = qk_hi + qk_el, *vt_hi = qk_lo + qk_el, *vt_lo = vt_hi + vt_el;
  const int64_t n = (int64_t)B * T * 3 * D;
  qkv_split_kernel<<<(int)std::min<int64_t>(ceil_div<int64_t>(n, 256), num_sms() * 16), 256, 0, st>>>(
      qkv, ld_qkv, B, T, H, Tp, qk_hi, qk_lo, vt_hi, vt_lo);
  EDA_LAUNCH_CHECK();
  return mha_split(qk_hi, qk_lo, 2 * D, vt_hi, vt_lo, Tp, B, T, H, out, ld_out, nullptr, nullptr, 0, prec, st);
}

}  // namespace eda

extern "C" int eda_mha_split(const void* qk_hi, const void* qk_lo, int ld_qk, const void* vt_hi, const void* vt_lo,
                             int Tp, int B, int T, int n_heads, float* out, int ld_out, void* out_hi, void* out_lo,
                             int ld_os, int prec, void* stream) {
  return eda::mha_split(qk_hi, qk_lo, ld_qk, vt_hi, vt_lo, Tp, B, T, n_heads, out, ld_out, out_hi, out_lo, ld_os,
                        prec, (cudaStream_t)stream);
}

extern "C" int eda_mha_split_train(const void* qk_hi, const void* qk_lo, int ld_qk, const void* vt_hi, const void* vt_lo,
                                   int Tp, int B, int T, int n_heads, void* out_hi, void* out_lo, int ld_os, float* lse,
                                   float p_drop, uint64_t seed, int site, const uint32_t* drop_bits, int ld_bits,
                                   int prec, void* stream) {
  EDA_REQUIRE(lse && out_hi && out_lo && p_drop >= 0.f && p_drop < 1.f, "eda_mha_split_train: bad argument");
  return eda::mha_split(qk_hi, qk_lo, ld_qk, vt_hi, vt_lo, Tp, B, T, n_heads, nullptr, 0, out_hi, out_lo, ld_os, prec,
                        (cudaStream_t)stream, lse, nullptr, eda::make_dropout(p_drop, seed, site), drop_bits, ld_bits);
}

extern "C" int eda_mha_split_masked(const void* qk_hi, const void* qk_lo, int ld_qk, const void* vt_hi, const void* vt_lo,
                                    int Tp, int B, int T, int n_heads, const int32_t* key_lens, void* out_hi, void* out_lo,
                                    int ld_os, int prec, void* stream) {
  EDA_REQUIRE(key_lens && out_hi && out_lo, "eda_mha_split_masked: null argument");
  return eda::mha_split(qk_hi, qk_lo, ld_qk, vt_hi, vt_lo, Tp, B, T, n_heads, nullptr, 0, out_hi, out_lo, ld_os, prec,
                        (cudaStream_t)stream, nullptr, key_lens);
}

#ifdef EDA_MHA_TRACE
extern "C" int eda_mha_trace_read(long long* host, int n_cta) {
  return (int)cudaMemcpyFromSymbol(host, eda::g_mha_trace, (size_t)n_cta * 96 * sizeof(long long));
}
#endif
