// EDA encoder/decoder LSTM recurrence as a persistent thread-block-cluster kernel.
//
// Reference: encoder_decoder_attractor.py:36-40 (packed encoder LSTM over the shuffled
// frames, decoder LSTM on zero inputs seeded with the encoder's final (h, c)).
//
// One cluster of 8 CTAs advances G = 4 sequences together.  CTA `c` owns hidden units
// [32c, 32c+32), i.e. the 128 gate rows {i,f,g,o} x 32 units of W_hh, RESIDENT IN REGISTERS
// for the whole sequence: thread (warp w, lane = rg*8 + kc) holds the 4 gate rows of unit
// u = 4w + rg restricted to the k-slice {4(kc + 8i) + j : i < 8, j < 4} (128 fp32).
// Per step:
//   wait on the local mbarrier until all 8 CTAs' slices of h_t have landed in shared memory
//   -> 8 conflict-free LDS.128 per sequence -> 256 packed FFMA2 (fma.rn.f32x2)
//   -> warp-shuffle transpose-reduction over the 8 k-slices (16 SHFL: the 16 (gate, sequence)
//      partial sums end up as 4 gates of ONE sequence per lane)
//   -> cell update in registers -> the CTA's 32 x G new h values are staged in shared memory
//      and pushed into the next h buffer of all 8 CTAs with ONE 512-byte cp.async.bulk per
//      destination (distributed shared memory), whose byte count completes the consumer's
//      mbarrier: no cluster-wide barrier in the loop, 8 barrier transactions per step.
// The input projection emb.W_ih^T + b is hoisted out as a tensor-core GEMM and the random
// frame shuffle (models.py:228-232) is applied as an index into it.
#include <cooperative_groups.h>

#include <algorithm>

#include <cuda_bf16.h>

#include "common.cuh"

namespace cg = cooperative_groups;

namespace eda {

constexpr int kD = 256;       // hidden size
constexpr int kCluster = 8;   // CTAs per cluster
constexpr int kUnits = kD / kCluster;  // 32 hidden units per CTA
// Sequences per cluster (template parameter kG): 4 when ceil(B/4) clusters fit the device at
// once (15 on a B200), otherwise 5..8 so that the whole batch still runs as ONE wave.

__device__ __forceinline__ uint32_t s_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint32_t map_rank(uint32_t saddr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(saddr), "r"(rank));
  return r;
}
// local shared memory -> a peer CTA's shared memory, completing the peer's mbarrier
__device__ __forceinline__ void bulk_push(uint32_t raddr, uint32_t laddr, uint32_t bytes, uint32_t rbar) {
  asm volatile("cp.async.bulk.shared::cluster.shared::cta.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(raddr), "r"(laddr), "r"(bytes), "r"(rbar) : "memory");
}
__device__ __forceinline__ void bar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(s_u32(bar)), "r"(count));
}
__device__ __forceinline__ void bar_arm(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(s_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool bar_try(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred P;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 P, [%1], %2;\n\t"   // bulk-copy bytes are visible on completion
      "selp.b32 %0, 1, 0, P;\n\t}"
      : "=r"(ok) : "r"(s_u32(bar)), "r"(parity) : "memory");
  return ok != 0;
}
__device__ __forceinline__ void bar_wait(uint64_t* bar, uint32_t parity) {
  if (bar_try(bar, parity)) return;
  const long long t0 = clock64();
  while (!bar_try(bar, parity)) {
    if (clock64() - t0 > 4000000000LL) { __trap(); }   // protocol bug -> launch error, never a hang
  }
}
// packed fp32x2 FMA (Blackwell FFMA2): d = a * b + c on both halves
__device__ __forceinline__ uint64_t ffma2(uint64_t a, uint64_t b, uint64_t c) {
  uint64_t d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
  return d;
}
__device__ __forceinline__ uint64_t pack2(float lo, float hi) {
  return (uint64_t)__float_as_uint(lo) | ((uint64_t)__float_as_uint(hi) << 32);
}
__device__ __forceinline__ float sum2(uint64_t v) {
  return __uint_as_float((uint32_t)v) + __uint_as_float((uint32_t)(v >> 32));
}

// Gate non-linearities on the MUFU pipe (ex2.approx + rcp.approx, ~3e-7 relative): they sit on
// the per-step dependency chain, where libm's expf/tanhf cost several hundred cycles.
__device__ __forceinline__ float fast_sigmoid(float x) { return __fdividef(1.f, 1.f + __expf(-x)); }
__device__ __forceinline__ float fast_tanh(float x) { return 1.f - __fdividef(2.f, 1.f + __expf(2.f * x)); }

// w2[gate][2*i + {0,1}] = packed (W[row][4(kc+8i)+0..1]), (W[row][4(kc+8i)+2..3])
__device__ __forceinline__ void load_w(uint64_t (&w2)[4][16], const float* __restrict__ W, int cta, int u, int kc) {
#pragma unroll
  for (int gate = 0; gate < 4; ++gate) {
    const float4* src = reinterpret_cast<const float4*>(W + (size_t)(gate * kD + cta * kUnits + u) * kD);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const float4 v = __ldg(src + kc + 8 * i);
      w2[gate][2 * i] = pack2(v.x, v.y);
      w2[gate][2 * i + 1] = pack2(v.z, v.w);
    }
  }
}

// What the training step keeps of the forward recurrence for its backward pass (all optional:
// gates == nullptr means inference).  Encoder rows are indexed by STEP (b * Tmax + s) except hprev,
// which sits at the frame's own row b * Tmax + perm[s] so that dW_hh = dxproj^T . hprev is one GEMM
// with dxproj in frame order; decoder rows are b * n_dec + s.
struct LstmSave {
  float* gates;            // (B*Tmax, 4*D) activated i, f, g, o
  float* c;                // (B*Tmax, D)   cell state after the step
  __nv_bfloat16* hprev_hi; // (B*Tmax, D)   hidden state BEFORE the step, split bf16
  __nv_bfloat16* hprev_lo;
  float* dec_gates;        // (B*n_dec, 4*D)
  float* dec_c;            // (B*n_dec, D)
  __nv_bfloat16* dec_hprev_hi;  // (B*n_dec, D)
  __nv_bfloat16* dec_hprev_lo;
};

#ifdef EDA_LSTM_TRACE
// probe build only (scripts/lstm_trace.py): clock64 stamps of thread 0 of every CTA for steps 64..71,
// 8 slots per step: wake, mat-vec done, reduced, cell done, staged + synced, pushed
__device__ long long g_lstm_trace[1024 * 64];
#define LSTM_STAMP(slot) do { if (tid == 0 && gs >= 64 && gs < 72) g_lstm_trace[(size_t)blockIdx.x * 64 + (gs - 64) * 8 + (slot)] = clock64(); } while (0)
#else
#define LSTM_STAMP(slot) do { } while (0)
#endif

template <int kG>
__global__ void __launch_bounds__(256, 1)
lstm_cluster_kernel(const float* __restrict__ xproj, const int32_t* __restrict__ perm,
                    const int32_t* __restrict__ lens, const int32_t* __restrict__ order,
                    const float* __restrict__ w_enc, const float* __restrict__ w_dec,
                    const float* __restrict__ dec_bias, int B, int Tmax, int n_dec,
                    float* __restrict__ attractors, float* __restrict__ hc_out, const LstmSave sv) {
  cg::cluster_group cluster = cg::this_cluster();
  const int cta = (int)cluster.block_rank();
  const int group = blockIdx.x / kCluster;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int rg = lane >> 3, kc = lane & 7;
  const int u = warp * 4 + rg;                         // local hidden unit of this lane
  // sequence whose gates end up in this lane after the reduction: kG <= 4 leaves every sum in
  // two lanes (kc, kc^1); kG > 4 uses all 8 k-slice lanes for 8 distinct sequences
  const int seq = kG <= 4 ? ((kc >> 2) & 1) * 2 + ((kc >> 1) & 1) : kc;
  const bool owner = kG <= 4 ? ((kc & 1) == 0 && seq < kG) : (kc < kG);   // lane that stores results

  // h_t double buffered, laid out [source CTA][sequence][unit] so that a CTA's slice is 512 B
  __shared__ __align__(128) float hs[2][kCluster][kG][kUnits];
  __shared__ __align__(128) float hstage[2][kG][kUnits];   // this CTA's new slice, double buffered
  __shared__ __align__(8) uint64_t bars[2];
  __shared__ int s_seq[kG], s_len[kG];

  if (tid < kG) {
    const int slot = group * kG + tid;
    int b = -1, n = 0;
    if (slot < B) { b = order ? order[slot] : slot; n = lens[b]; }
    s_seq[tid] = b; s_len[tid] = n;
  }
  for (int i = tid; i < 2 * kG * kD; i += 256) (&hs[0][0][0][0])[i] = 0.f;
  if (tid == 0) {
    bar_init(&bars[0], 1); bar_init(&bars[1], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  int max_len = 0;
#pragma unroll
  for (int g = 0; g < kG; ++g) max_len = max(max_len, s_len[g]);
  const int n_total = max_len + n_dec;
  constexpr uint32_t kStepBytes = kG * kD * 4;
  if (tid == 0) {
    if (n_total > 1) bar_arm(&bars[1], kStepBytes);    // h_1 (written during step 0)
    if (n_total > 2) bar_arm(&bars[0], kStepBytes);    // h_2
  }

  uint64_t w2[4][16];
  load_w(w2, w_enc, cta, u, kc);

  const int my_b = seq < kG ? s_seq[seq] : -1;
  const int my_len = seq < kG ? s_len[seq] : 0;
  float c_state = 0.f, h_state = 0.f;
  float xg[4] = {0.f, 0.f, 0.f, 0.f};
  int p_next = 0, p_cur = 0;
  const size_t col = (size_t)cta * kUnits + u;
  auto load_x = [&](int p, float (&dst)[4]) {
    const float* row = xproj + ((size_t)my_b * Tmax + p) * (4 * kD) + col;
#pragma unroll
    for (int x = 0; x < 4; ++x) dst[x] = __ldg(row + x * kD);
  };
  if (my_len > 0) {   // (my_b >= 0 whenever my_len > 0)
    p_cur = __ldg(perm + (size_t)my_b * Tmax);
    load_x(p_cur, xg);
    if (my_len > 1) p_next = __ldg(perm + (size_t)my_b * Tmax + 1);
  }
  // lane 0 of warp r pushes this CTA's slice to rank r (one bulk copy per warp: 8 copies issued by one
  // warp instruction serialise, ~600 cycles): remote slice / barrier addresses per buffer
  uint32_t r_h0 = 0, r_h1 = 0, r_bar0 = 0, r_bar1 = 0;
  if (lane == 0) {
    r_h0 = map_rank(s_u32(&hs[0][cta][0][0]), warp); r_h1 = map_rank(s_u32(&hs[1][cta][0][0]), warp);
    r_bar0 = map_rank(s_u32(&bars[0]), warp); r_bar1 = map_rank(s_u32(&bars[1]), warp);
  }
  cluster.sync();   // barriers initialised and h buffers zeroed everywhere before any remote write

  for (int gs = 0; gs < n_total; ++gs) {
    const bool dec = gs >= max_len;
    const int s = dec ? gs - max_len : gs;
    const int cur = gs & 1;
    if (gs == max_len) {              // switch to the decoder: new recurrent weights, bias-only input
      load_w(w2, w_dec, cta, u, kc);
#pragma unroll
      for (int x = 0; x < 4; ++x) xg[x] = __ldg(dec_bias + x * kD + col);
      if (hc_out && my_b >= 0 && owner) {
        hc_out[((size_t)my_b * 2 + 0) * kD + col] = h_state;
        hc_out[((size_t)my_b * 2 + 1) * kD + col] = c_state;
      }
    }
    // prefetch next step's input projection while the mat-vec runs
    float xn[4] = {0.f, 0.f, 0.f, 0.f};
    int p_next2 = 0;
    if (!dec && s + 1 < my_len) {
      load_x(p_next, xn);
      if (s + 2 < my_len) p_next2 = __ldg(perm + (size_t)my_b * Tmax + s + 2);
    }
    if (gs > 0) {
      bar_wait(&bars[cur], ((gs - 1) >> 1) & 1);       // all 8 slices of h_gs have landed
      if (tid == 0 && gs + 2 < n_total) bar_arm(&bars[cur], kStepBytes);   // re-arm for h_{gs+2}
    }
    LSTM_STAMP(0);
    uint64_t acc[4][kG];
#pragma unroll
    for (int gate = 0; gate < 4; ++gate)
#pragma unroll
      for (int g = 0; g < kG; ++g) acc[gate][g] = 0ull;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
#pragma unroll
      for (int g = 0; g < kG; ++g) {
        const float4 hv = *reinterpret_cast<const float4*>(&hs[cur][i][g][4 * kc]);   // k = 32 i + 4 kc + j
        const uint64_t h01 = pack2(hv.x, hv.y), h23 = pack2(hv.z, hv.w);
#pragma unroll
        for (int gate = 0; gate < 4; ++gate) {
          acc[gate][g] = ffma2(w2[gate][2 * i], h01, acc[gate][g]);
          acc[gate][g] = ffma2(w2[gate][2 * i + 1], h23, acc[gate][g]);
        }
      }
    }
    // transpose-reduce over the 8 k-slices (lane bits 2,1,0): every round halves the set of
    // sequences a lane keeps and adds the partner's partial sums for them
    float v4[4];
    LSTM_STAMP(1);
    if constexpr (kG <= 4) {
      float v8[4][2];
      {
        const bool up = (kc & 4) != 0;                   // keeps sequences {2,3}
#pragma unroll
        for (int gate = 0; gate < 4; ++gate)
#pragma unroll
          for (int j = 0; j < 2; ++j) {
            const float lo = j < kG ? sum2(acc[gate][j]) : 0.f, hi = 2 + j < kG ? sum2(acc[gate][2 + j]) : 0.f;
            v8[gate][j] = (up ? hi : lo) + __shfl_xor_sync(0xffffffffu, up ? lo : hi, 4);
          }
      }
      {
        const bool up = (kc & 2) != 0;                   // keeps the odd sequence of its pair
#pragma unroll
        for (int gate = 0; gate < 4; ++gate) {
          const float keep = up ? v8[gate][1] : v8[gate][0];
          const float give = up ? v8[gate][0] : v8[gate][1];
          v4[gate] = keep + __shfl_xor_sync(0xffffffffu, give, 2);
        }
      }
#pragma unroll
      for (int gate = 0; gate < 4; ++gate) v4[gate] += __shfl_xor_sync(0xffffffffu, v4[gate], 1);
    } else {
      float va[4][4], vb[4][2];
      {
        const bool up = (kc & 4) != 0;                   // keeps sequences {4..7}
#pragma unroll
        for (int gate = 0; gate < 4; ++gate)
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const float lo = sum2(acc[gate][j]), hi = 4 + j < kG ? sum2(acc[gate][4 + j]) : 0.f;
            va[gate][j] = (up ? hi : lo) + __shfl_xor_sync(0xffffffffu, up ? lo : hi, 4);
          }
      }
      {
        const bool up = (kc & 2) != 0;
#pragma unroll
        for (int gate = 0; gate < 4; ++gate)
#pragma unroll
          for (int j = 0; j < 2; ++j)
            vb[gate][j] = (up ? va[gate][2 + j] : va[gate][j]) +
                          __shfl_xor_sync(0xffffffffu, up ? va[gate][j] : va[gate][2 + j], 2);
      }
      {
        const bool up = (kc & 1) != 0;
#pragma unroll
        for (int gate = 0; gate < 4; ++gate)
          v4[gate] = (up ? vb[gate][1] : vb[gate][0]) + __shfl_xor_sync(0xffffffffu, up ? vb[gate][0] : vb[gate][1], 1);
      }
    }

    LSTM_STAMP(2);
    const bool active = dec ? (my_b >= 0) : (s < my_len);
    if (active) {
      const float ig = fast_sigmoid(v4[0] + xg[0]), fg = fast_sigmoid(v4[1] + xg[1]);
      const float gg = fast_tanh(v4[2] + xg[2]), og = fast_sigmoid(v4[3] + xg[3]);
      if (sv.gates != nullptr && owner) {        // training: keep what the backward recurrence needs
        const size_t row = dec ? (size_t)my_b * n_dec + s : (size_t)my_b * Tmax + s;
        float* gsv = (dec ? sv.dec_gates : sv.gates) + row * (4 * kD) + col;
        gsv[0] = ig; gsv[kD] = fg; gsv[2 * kD] = gg; gsv[3 * kD] = og;
        const size_t hrow = dec ? row : (size_t)my_b * Tmax + p_cur;
        const __nv_bfloat16 hh = __float2bfloat16_rn(h_state);
        (dec ? sv.dec_hprev_hi : sv.hprev_hi)[hrow * kD + col] = hh;
        (dec ? sv.dec_hprev_lo : sv.hprev_lo)[hrow * kD + col] = __float2bfloat16_rn(h_state - __bfloat162float(hh));
      }
      c_state = fg * c_state + ig * gg;
      h_state = og * fast_tanh(c_state);
      if (sv.gates != nullptr && owner) (dec ? sv.dec_c : sv.c)[(dec ? (size_t)my_b * n_dec + s : (size_t)my_b * Tmax + s) * kD + col] = c_state;
      if (dec && owner) attractors[((size_t)my_b * n_dec + s) * kD + col] = h_state;
    }
    LSTM_STAMP(3);
    if (gs + 1 < n_total) {           // publish h (changed or not) into the next buffer of all 8 CTAs
      if (owner) hstage[cur][seq][u] = h_state;
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      __syncthreads();
      LSTM_STAMP(4);
      if (lane == 0)
        bulk_push(cur ? r_h0 : r_h1, s_u32(&hstage[cur][0][0]), kG * kUnits * 4, cur ? r_bar0 : r_bar1);
      LSTM_STAMP(5);
    }
    if (!dec) {
#pragma unroll
      for (int x = 0; x < 4; ++x) xg[x] = xn[x];
      p_cur = p_next;
      p_next = p_next2;
    }
  }
  cluster.sync();    // nobody exits while a peer may still be reading its own buffers' traffic
}

// ------------------------------------------------------------------ tensor-core recurrence -----
// Same cluster protocol, but the per-step mat-vec W_hh[own 128 rows] . h runs on the tensor cores as
// split-bf16 products (W_hi.h_hi + W_lo.h_hi + W_hi.h_lo, fp32 accumulate) with mma.sync m16n8k16:
// the FFMA version above is bound by register-file bandwidth (3 x 64-bit operands per FFMA2, ~1800-2200
// cycles per step measured), the MMA version issues 48 instructions per warp and step (~770 cycles).
// W stays register resident: warp w holds the A fragments of its 16 gate rows (4 units x {i,f,g,o}) for
// all 16 k-steps, hi and lo (128 registers, what the fp32 weights took).  N = 8 = sequences of the cluster.
//   rows of warp w's tile: thread group g = lane / 4 owns tile rows g and g + 8;
//     g even -> unit 4w + g/2, gates i (row g) and f (row g + 8); g odd -> same unit, gates g~ and o
//   one xor-4 shuffle pair swaps the halves so that lane (g, q) ends with the 4 gates of
//     (unit 4w + g/2, sequence 2q + (g & 1)): 4 units x 8 sequences per warp.
//   k ordering is free (it is summed over), so it is chosen for the loads: k-steps 2i, 2i+1 cover source
//     CTA i's 32 units and thread q takes units 32i + 8q + {0..3} / {4..7} -> its B fragments of both
//     steps are ONE 16-byte shared-memory load (hi) + one (lo), conflict-free.
//   h travels as split bf16 by the same bulk copies: slice layout [source CTA][sequence][128 B] with
//     hi | lo in the two 64-byte halves (swapped for odd sequences: keeps the fragment loads conflict-free).
//   (Tried and rejected: per-lane st.async remote stores instead of staging + bulk copy: the exchange took
//    ~1500 cycles instead of ~900.)
__device__ __forceinline__ void mma_bf16(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void split2(float x, float y, uint32_t& hi, uint32_t& lo) {
  const __nv_bfloat16 xh = __float2bfloat16_rn(x), yh = __float2bfloat16_rn(y);
  const __nv_bfloat16 xl = __float2bfloat16_rn(x - __bfloat162float(xh)), yl = __float2bfloat16_rn(y - __bfloat162float(yh));
  hi = (uint32_t)__bfloat16_as_ushort(xh) | ((uint32_t)__bfloat16_as_ushort(yh) << 16);
  lo = (uint32_t)__bfloat16_as_ushort(xl) | ((uint32_t)__bfloat16_as_ushort(yl) << 16);
}
// A fragments of k-step t = 2i + s: a0a1 = (row g, units kb, kb+1), a2a3 = (row g+8, same), a4a5 = (row g, kb+2, kb+3),
// a6a7 = (row g+8, same) with kb = 32i + 8q + 4s
__device__ __forceinline__ void load_w_mma(uint32_t (&wh)[16][4], uint32_t (&wl)[16][4], const float* __restrict__ W,
                                           int cta, int warp, int lane) {
  const int g = lane >> 2, q = lane & 3;
  const int unit = cta * kUnits + warp * 4 + (g >> 1);
  const float* r0 = W + (size_t)(((g & 1) * 2 + 0) * kD + unit) * kD;   // gate i or g~
  const float* r1 = W + (size_t)(((g & 1) * 2 + 1) * kD + unit) * kD;   // gate f or o
#pragma unroll
  for (int t = 0; t < 16; ++t) {
    const int kb = 32 * (t >> 1) + 8 * q + 4 * (t & 1);
    const float4 v0 = __ldg(reinterpret_cast<const float4*>(r0 + kb));
    const float4 v1 = __ldg(reinterpret_cast<const float4*>(r1 + kb));
    split2(v0.x, v0.y, wh[t][0], wl[t][0]);
    split2(v1.x, v1.y, wh[t][1], wl[t][1]);
    split2(v0.z, v0.w, wh[t][2], wl[t][2]);
    split2(v1.z, v1.w, wh[t][3], wl[t][3]);
  }
}

template <int kG>
__global__ void __launch_bounds__(256, 1)
lstm_mma_cluster_kernel(const float* __restrict__ xproj, const int32_t* __restrict__ perm,
                        const int32_t* __restrict__ lens, const int32_t* __restrict__ order,
                        const float* __restrict__ w_enc, const float* __restrict__ w_dec,
                        const float* __restrict__ dec_bias, int B, int Tmax, int n_dec,
                        float* __restrict__ attractors, float* __restrict__ hc_out, const LstmSave sv) {
  cg::cluster_group cluster = cg::this_cluster();
  const int cta = (int)cluster.block_rank();
  const int group = blockIdx.x / kCluster;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int g = lane >> 2, q = lane & 3;
  const int u = warp * 4 + (g >> 1);                   // local hidden unit of this lane
  const int seq = 2 * q + (g & 1);                     // sequence whose gates end up in this lane
  const bool owner = seq < kG;

  // h_t double buffered: [source CTA][sequence 0..7][hi 64 B | lo 64 B] (halves swapped for odd sequences)
  __shared__ __align__(128) uint8_t hs[2][kCluster][8][128];
  __shared__ __align__(128) uint8_t hstage[2][8][128];
  __shared__ __align__(8) uint64_t bars[2];
  __shared__ int s_seq[kG], s_len[kG];

  if (tid < kG) {
    const int slot = group * kG + tid;
    int b = -1, n = 0;
    if (slot < B) { b = order ? order[slot] : slot; n = lens[b]; }
    s_seq[tid] = b; s_len[tid] = n;
  }
  for (int i = tid; i < 2 * kCluster * 8 * 32; i += 256) reinterpret_cast<uint32_t*>(&hs[0][0][0][0])[i] = 0u;
  for (int i = tid; i < 2 * 8 * 32; i += 256) reinterpret_cast<uint32_t*>(&hstage[0][0][0])[i] = 0u;
  if (tid == 0) {
    bar_init(&bars[0], 1); bar_init(&bars[1], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  int max_len = 0;
#pragma unroll
  for (int x = 0; x < kG; ++x) max_len = max(max_len, s_len[x]);
  const int n_total = max_len + n_dec;
  constexpr uint32_t kSliceBytes = kG * 128;
  constexpr uint32_t kStepBytes = kCluster * kSliceBytes;
  if (tid == 0) {
    if (n_total > 1) bar_arm(&bars[1], kStepBytes);
    if (n_total > 2) bar_arm(&bars[0], kStepBytes);
  }

  uint32_t wh[16][4], wl[16][4];
  load_w_mma(wh, wl, w_enc, cta, warp, lane);

  const int my_b = owner ? s_seq[seq] : -1;
  const int my_len = owner ? s_len[seq] : 0;
  float c_state = 0.f, h_state = 0.f;
  float xg[4] = {0.f, 0.f, 0.f, 0.f};
  int p_next = 0, p_cur = 0;
  const size_t col = (size_t)cta * kUnits + u;
  auto load_x = [&](int p, float (&dst)[4]) {
    const float* row = xproj + ((size_t)my_b * Tmax + p) * (4 * kD) + col;
#pragma unroll
    for (int x = 0; x < 4; ++x) dst[x] = __ldg(row + x * kD);
  };
  if (my_len > 0) {
    p_cur = __ldg(perm + (size_t)my_b * Tmax);
    load_x(p_cur, xg);
    if (my_len > 1) p_next = __ldg(perm + (size_t)my_b * Tmax + 1);
  }
  uint32_t r_h0 = 0, r_h1 = 0, r_bar0 = 0, r_bar1 = 0;
  if (lane == 0) {    // lane 0 of warp r pushes this CTA's slice to rank r
    r_h0 = map_rank(s_u32(&hs[0][cta][0][0]), warp); r_h1 = map_rank(s_u32(&hs[1][cta][0][0]), warp);
    r_bar0 = map_rank(s_u32(&bars[0]), warp); r_bar1 = map_rank(s_u32(&bars[1]), warp);
  }
  // this lane's fragment source inside a slice: sequence g, 16-byte chunk q of the hi / lo half
  const int frag_hi = g * 128 + (g & 1) * 64 + q * 16, frag_lo = g * 128 + (1 - (g & 1)) * 64 + q * 16;
  // where this lane's new h goes inside the staged slice
  const int st_hi = seq * 128 + (seq & 1) * 64 + 2 * u, st_lo = seq * 128 + (1 - (seq & 1)) * 64 + 2 * u;
  cluster.sync();

  for (int gs = 0; gs < n_total; ++gs) {
    const bool dec = gs >= max_len;
    const int s = dec ? gs - max_len : gs;
    const int cur = gs & 1;
    if (gs == max_len) {
      load_w_mma(wh, wl, w_dec, cta, warp, lane);
#pragma unroll
      for (int x = 0; x < 4; ++x) xg[x] = __ldg(dec_bias + x * kD + col);
      if (hc_out && my_b >= 0 && owner) {
        hc_out[((size_t)my_b * 2 + 0) * kD + col] = h_state;
        hc_out[((size_t)my_b * 2 + 1) * kD + col] = c_state;
      }
    }
    float xn[4] = {0.f, 0.f, 0.f, 0.f};
    int p_next2 = 0;
    if (!dec && s + 1 < my_len) {
      load_x(p_next, xn);
      if (s + 2 < my_len) p_next2 = __ldg(perm + (size_t)my_b * Tmax + s + 2);
    }
    if (gs > 0) {
      bar_wait(&bars[cur], ((gs - 1) >> 1) & 1);
      if (tid == 0 && gs + 2 < n_total) bar_arm(&bars[cur], kStepBytes);
    }
    LSTM_STAMP(0);
    // one accumulation chain per product (hi.hi, lo.hi, hi.lo): with 2 warps per scheduler that keeps
    // 6 MMAs in flight, and the small terms are summed among themselves first
    float acc[3][4];
#pragma unroll
    for (int a = 0; a < 3; ++a)
#pragma unroll
      for (int x = 0; x < 4; ++x) acc[a][x] = 0.f;
#pragma unroll
    for (int i = 0; i < kCluster; ++i) {
      const uint4 bh = *reinterpret_cast<const uint4*>(&hs[cur][i][0][0] + frag_hi);
      const uint4 bl = *reinterpret_cast<const uint4*>(&hs[cur][i][0][0] + frag_lo);
      mma_bf16(acc[0], wh[2 * i], bh.x, bh.y);
      mma_bf16(acc[1], wl[2 * i], bh.x, bh.y);
      mma_bf16(acc[2], wh[2 * i], bl.x, bl.y);
      mma_bf16(acc[0], wh[2 * i + 1], bh.z, bh.w);
      mma_bf16(acc[1], wl[2 * i + 1], bh.z, bh.w);
      mma_bf16(acc[2], wh[2 * i + 1], bl.z, bl.w);
    }
    LSTM_STAMP(1);
    float c4[4];
#pragma unroll
    for (int x = 0; x < 4; ++x) c4[x] = (acc[1][x] + acc[2][x]) + acc[0][x];
    // c4 = (row g, seq 2q), (row g, seq 2q+1), (row g+8, seq 2q), (row g+8, seq 2q+1): swap halves with lane ^ 4
    const bool odd = (g & 1) != 0;
    const float r0 = __shfl_xor_sync(0xffffffffu, odd ? c4[0] : c4[1], 4);
    const float r1 = __shfl_xor_sync(0xffffffffu, odd ? c4[2] : c4[3], 4);
    float v4[4];
    v4[0] = odd ? r0 : c4[0]; v4[1] = odd ? r1 : c4[2];      // i, f
    v4[2] = odd ? c4[1] : r0; v4[3] = odd ? c4[3] : r1;      // g~, o
    LSTM_STAMP(2);

    const bool active = dec ? (my_b >= 0) : (s < my_len);
    if (active) {
      const float ig = fast_sigmoid(v4[0] + xg[0]), fg = fast_sigmoid(v4[1] + xg[1]);
      const float gg = fast_tanh(v4[2] + xg[2]), og = fast_sigmoid(v4[3] + xg[3]);
      if (sv.gates != nullptr) {
        const size_t row = dec ? (size_t)my_b * n_dec + s : (size_t)my_b * Tmax + s;
        float* gsv = (dec ? sv.dec_gates : sv.gates) + row * (4 * kD) + col;
        gsv[0] = ig; gsv[kD] = fg; gsv[2 * kD] = gg; gsv[3 * kD] = og;
        const size_t hrow = dec ? row : (size_t)my_b * Tmax + p_cur;
        const __nv_bfloat16 hh = __float2bfloat16_rn(h_state);
        (dec ? sv.dec_hprev_hi : sv.hprev_hi)[hrow * kD + col] = hh;
        (dec ? sv.dec_hprev_lo : sv.hprev_lo)[hrow * kD + col] = __float2bfloat16_rn(h_state - __bfloat162float(hh));
      }
      c_state = fg * c_state + ig * gg;
      h_state = og * fast_tanh(c_state);
      if (sv.gates != nullptr) (dec ? sv.dec_c : sv.c)[(dec ? (size_t)my_b * n_dec + s : (size_t)my_b * Tmax + s) * kD + col] = c_state;
      if (dec) attractors[((size_t)my_b * n_dec + s) * kD + col] = h_state;
    }
    LSTM_STAMP(3);
    if (gs + 1 < n_total) {
      if (owner) {
        const __nv_bfloat16 hh = __float2bfloat16_rn(h_state);
        *reinterpret_cast<__nv_bfloat16*>(&hstage[cur][0][0] + st_hi) = hh;
        *reinterpret_cast<__nv_bfloat16*>(&hstage[cur][0][0] + st_lo) = __float2bfloat16_rn(h_state - __bfloat162float(hh));
      }
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      __syncthreads();
      LSTM_STAMP(4);
      if (lane == 0)
        bulk_push(cur ? r_h0 : r_h1, s_u32(&hstage[cur][0][0]), kSliceBytes, cur ? r_bar0 : r_bar1);
      LSTM_STAMP(5);
    }
    if (!dec) {
#pragma unroll
      for (int x = 0; x < 4; ++x) xg[x] = xn[x];
      p_cur = p_next;
      p_next = p_next2;
    }
  }
  cluster.sync();
}

#ifdef EDA_LSTM_TRACE
}  // namespace eda
extern "C" int eda_lstm_trace_read(long long* host, int n_cta) {
  return cudaMemcpyFromSymbol(host, eda::g_lstm_trace, sizeof(long long) * 64 * (size_t)n_cta) == cudaSuccess ? 0 : 1;
}
namespace eda {
#endif

template <int kG>
static int launch_lstm(cudaLaunchConfig_t& cfg, const float* xproj, const int32_t* perm, const int32_t* lens,
                       const int32_t* order, const float* w_enc, const float* w_dec, const float* dec_bias,
                       int B, int Tmax, int n_dec, float* attractors, float* hc_out, const LstmSave& sv, bool tensor) {
  if (tensor)
    EDA_CUDA(cudaLaunchKernelEx(&cfg, lstm_mma_cluster_kernel<kG>, xproj, perm, lens, order, w_enc, w_dec,
                                dec_bias, B, Tmax, n_dec, attractors, hc_out, sv));
  else
    EDA_CUDA(cudaLaunchKernelEx(&cfg, lstm_cluster_kernel<kG>, xproj, perm, lens, order, w_enc, w_dec,
                                dec_bias, B, Tmax, n_dec, attractors, hc_out, sv));
  count_launch();
  return EDA_OK;
}

static void lstm_launch_config(cudaLaunchConfig_t& cfg, cudaLaunchAttribute* attr, int groups, cudaStream_t st) {
  cfg = cudaLaunchConfig_t{};
  cfg.gridDim = dim3(groups * kCluster);
  cfg.blockDim = dim3(256);
  cfg.dynamicSmemBytes = 0;
  cfg.stream = st;
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = kCluster;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
}

// How many 8-CTA clusters of the recurrence kernel the device can keep resident at once
// (each advances kG sequences): batches beyond kG * this run in more than one wave.
int lstm_max_active_clusters() {
  static int cached = -1;
  if (cached < 0) {
    cudaLaunchConfig_t cfg;
    cudaLaunchAttribute attr[1];
    lstm_launch_config(cfg, attr, 64, nullptr);
    int n = 0;
    if (cudaOccupancyMaxActiveClusters(&n, lstm_cluster_kernel<4>, &cfg) != cudaSuccess) { cudaGetLastError(); n = 0; }
    int m = 0;
    if (cudaOccupancyMaxActiveClusters(&m, lstm_mma_cluster_kernel<4>, &cfg) != cudaSuccess) { cudaGetLastError(); m = 0; }
    cached = std::min(n, m);
  }
  return cached;
}

int lstm_attractors(const float* xproj, const int32_t* perm, const int32_t* lens,
                    const int32_t* order, const float* w_enc, const float* w_dec,
                    const float* dec_bias, int B, int Tmax, int D, int n_dec, float* attractors,
                    float* hc_out, const LstmSave& sv, int prec, cudaStream_t st) {
  EDA_REQUIRE(D == kD, "eda_lstm_attractors: D must be 256 (got %d)", D);
  EDA_REQUIRE(prec == EDA_PREC_FP32 || prec == EDA_PREC_BF16X3 || prec == EDA_PREC_BF16, "eda_lstm_attractors: bad precision %d", prec);
  const bool tensor = prec != EDA_PREC_FP32;
  EDA_REQUIRE(B > 0 && Tmax > 0 && n_dec >= 0, "eda_lstm_attractors: bad sizes");
  EDA_REQUIRE(((uintptr_t)w_enc | (uintptr_t)w_dec) % 16 == 0, "eda_lstm_attractors: weights must be 16B aligned");
  // one wave if at all possible: the smallest G in 4..8 whose cluster count fits the device
  if (sv.gates != nullptr)
    EDA_REQUIRE(sv.c && sv.hprev_hi && sv.hprev_lo && sv.dec_gates && sv.dec_c && sv.dec_hprev_hi && sv.dec_hprev_lo && hc_out,
                "eda_lstm_attractors_train: every save buffer (and hc_out) is required");
  const int fit = std::max(1, lstm_max_active_clusters());
  int G = 4;
  while (G < 8 && ceil_div(B, G) > fit) ++G;
  // the MMA tile is 8 sequences wide whatever G is: a step costs about the same for 8 sequences as for 4
  // (+5 % measured), so the inference launch packs 8 per cluster and leaves the other SMs to the batches
  // running beside it on the pipeline's other streams (+4 % end to end).  The training launch (nothing runs
  // beside it, and every owner lane also stores 7 values per step) keeps the smallest G.
  if (tensor && sv.gates == nullptr) G = std::min(8, std::max(4, B));
  cudaLaunchConfig_t cfg;
  cudaLaunchAttribute attr[1];
  lstm_launch_config(cfg, attr, ceil_div(B, G), st);
  switch (G) {
    case 4: return launch_lstm<4>(cfg, xproj, perm, lens, order, w_enc, w_dec, dec_bias, B, Tmax, n_dec, attractors, hc_out, sv, tensor);
    case 5: return launch_lstm<5>(cfg, xproj, perm, lens, order, w_enc, w_dec, dec_bias, B, Tmax, n_dec, attractors, hc_out, sv, tensor);
    case 6: return launch_lstm<6>(cfg, xproj, perm, lens, order, w_enc, w_dec, dec_bias, B, Tmax, n_dec, attractors, hc_out, sv, tensor);
    case 7: return launch_lstm<7>(cfg, xproj, perm, lens, order, w_enc, w_dec, dec_bias, B, Tmax, n_dec, attractors, hc_out, sv, tensor);
    default: return launch_lstm<8>(cfg, xproj, perm, lens, order, w_enc, w_dec, dec_bias, B, Tmax, n_dec, attractors, hc_out, sv, tensor);
  }
}

// ===================================================================== backward recurrence =====
// BPTT through the decoder and encoder LSTMs, same cluster shape as the forward:
// CTA c owns hidden units [32c, 32c+32).  Walking the steps backwards, per step
//   wait until all 8 CTAs' gate gradients of the LATER step have landed (mbarrier + DSMEM bulk
//   pushes of split-bf16 pairs, 4 x the forward's volume: 128 gate rows per CTA and sequence)
//   -> dh[own 32 units] += W_hh[:, own units]^T . dgates on the tensor cores (the 1024 x 32 column slice of
//      W_hh is register resident as split-bf16 MMA fragments, see load_wt_mma; the 8 warps' partial tiles are
//      summed through shared memory)
//   -> lane kc == sequence finishes the cell backward from the saved activated gates / cell states
//   -> its 4 gate gradients go to HBM as the split-bf16 operand of the weight / input gradient
//      GEMMs (frame order, row b*Tmax + perm[s]) and to the peers for the next (earlier) step.
struct LstmBwd {
  const int32_t* perm; const int32_t* lens; const int32_t* order;
  const float* w_enc; const float* w_dec;
  const float* gates; const float* c; const float* dec_gates; const float* dec_c; const float* hc;
  const float* datt;                         // (B, n_dec, D) gradient w.r.t. the attractors
  __nv_bfloat16* dxp_hi; __nv_bfloat16* dxp_lo;     // (B*Tmax, 4D) d loss / d (gate pre-activations), frame order
  __nv_bfloat16* ddec_hi; __nv_bfloat16* ddec_lo;   // (B*n_dec, 4D)
  int B, Tmax, n_dec;
};

// Mat-vec of the backward kernel on the tensor cores (mma.sync m16n8k16, split-bf16 x 3 products like the
// forward): dh[unit (M = 32: two 16-row tiles)][sequence (N = 8)] = sum over the 1024 gate rows (K) of
// W_hh[row][own unit] . dgates[row][sequence].  Warp w takes the K range that arrives from source CTA w
// (8 k-steps) for both M tiles: 2 x 8 x 4 A registers, hi and lo = 128 registers, resident.  The 8 warps'
// partial tiles are summed through shared memory by the lane that owns (unit, sequence).
// k order inside a source slice (free, chosen for the 8-byte stores of the producer and the 16-byte loads of
// the consumer): 16-byte chunk c of a sequence's row holds units 2c, 2c+1 x gates i,f,g,o; thread q reads
// chunks q + 4j: k-step 2j = unit 2(q+4j), k-step 2j+1 = unit 2(q+4j)+1, logical k {2q,2q+1 | 2q+8,2q+9} = gates {0,1 | 2,3}.
__device__ __forceinline__ void load_wt_mma(uint32_t (&wh)[2][8][4], uint32_t (&wl)[2][8][4], const float* __restrict__ W,
                                            int cta, int warp, int lane) {
  const int g = lane >> 2, q = lane & 3;
  constexpr size_t kGate = (size_t)kD * kD;
#pragma unroll
  for (int m = 0; m < 2; ++m)
#pragma unroll
    for (int t = 0; t < 8; ++t) {
      const int v = 2 * (q + 4 * (t >> 1)) + (t & 1);                       // unit inside source CTA `warp`
      const float* base = W + (size_t)(32 * warp + v) * kD + cta * kUnits + m * 16 + g;
      split2(__ldg(base), __ldg(base + kGate), wh[m][t][0], wl[m][t][0]);                       // out unit g,   gates i, f
      split2(__ldg(base + 8), __ldg(base + kGate + 8), wh[m][t][1], wl[m][t][1]);               // out unit g+8, gates i, f
      split2(__ldg(base + 2 * kGate), __ldg(base + 3 * kGate), wh[m][t][2], wl[m][t][2]);       // out unit g,   gates g~, o
      split2(__ldg(base + 2 * kGate + 8), __ldg(base + 3 * kGate + 8), wh[m][t][3], wl[m][t][3]);
    }
}

template <int kG>
__global__ void __launch_bounds__(256, 1)
lstm_bwd_cluster_kernel(const LstmBwd P) {
  cg::cluster_group cluster = cg::this_cluster();
  const int cta = (int)cluster.block_rank();
  const int group = blockIdx.x / kCluster;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int rg = lane >> 3, kc = lane & 7;
  const int u = warp * 4 + rg;                 // (unit, sequence) this lane owns for the cell backward: tid = 8 u + seq
  const int seq = kc;
  const bool owner = kc < kG;
  const int g = lane >> 2, q = lane & 3;       // MMA fragment coordinates

  extern __shared__ __align__(128) uint8_t lsm_raw[];
  // gate gradients of one step as split bf16, [buffer][source CTA][sequence][hi 256 B | lo 256 B]; inside a half,
  // 16-byte chunk c (units 2c, 2c+1 x 4 gates) sits at c ^ (4 * (sequence & 1)): conflict-free fragment loads
  uint8_t* dg = lsm_raw;
  uint8_t* stage = lsm_raw + 2 * kCluster * kG * 512;
  float* red = reinterpret_cast<float*>(lsm_raw + 2 * kCluster * kG * 512 + 2 * kG * 512);   // [warp][unit 32][sequence 8]
  uint64_t* bars = reinterpret_cast<uint64_t*>(red + 8 * 32 * 8);
  int* s_seq = reinterpret_cast<int*>(bars + 2);
  int* s_len = s_seq + 8;

  if (tid < kG) {
    const int slot = group * kG + tid;
    int b = -1, n = 0;
    if (slot < P.B) { b = P.order ? P.order[slot] : slot; n = P.lens[b]; }
    s_seq[tid] = b; s_len[tid] = n;
  }
  if (tid == 0) {
    bar_init(&bars[0], 1); bar_init(&bars[1], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  int max_len = 0;
#pragma unroll
  for (int x = 0; x < kG; ++x) max_len = max(max_len, s_len[x]);
  const int n_dec = P.n_dec, Tmax = P.Tmax;
  const int n_total = max_len + n_dec;
  constexpr uint32_t kStepBytes = kCluster * kG * 512;
  if (tid == 0) {
    if (n_total > 1) bar_arm(&bars[1], kStepBytes);
    if (n_total > 2) bar_arm(&bars[0], kStepBytes);
  }
  uint32_t wh[2][8][4], wl[2][8][4];
  load_wt_mma(wh, wl, n_dec > 0 ? P.w_dec : P.w_enc, cta, warp, lane);

  const int my_b = owner ? s_seq[seq] : -1;
  const int my_len = owner ? s_len[seq] : 0;
  const size_t col = (size_t)cta * kUnits + u;
  float carry_dh = 0.f, carry_dc = 0.f;

  // saved activations of one step for this lane's (sequence, unit)
  struct StepIn { float g[4], c, cp, da; int p; };
  auto load_step = [&](int gs, StepIn& o) {
    o.g[0] = o.g[1] = o.g[2] = o.g[3] = o.c = o.cp = o.da = 0.f; o.p = 0;
    if (my_b < 0 || gs < 0) return;
    const bool dec = gs >= max_len;
    const int s = dec ? gs - max_len : gs;
    if (dec) {
      const size_t row = (size_t)my_b * n_dec + s;
      const float* gp = P.dec_gates + row * (4 * kD) + col;
#pragma unroll
      for (int x = 0; x < 4; ++x) o.g[x] = __ldg(gp + x * kD);
      o.c = __ldg(P.dec_c + row * kD + col);
      o.cp = s > 0 ? __ldg(P.dec_c + (row - 1) * kD + col) : __ldg(P.hc + ((size_t)my_b * 2 + 1) * kD + col);
      o.da = __ldg(P.datt + row * kD + col);
    } else if (s < my_len) {
      const size_t row = (size_t)my_b * Tmax + s;
      const float* gp = P.gates + row * (4 * kD) + col;
#pragma unroll
      for (int x = 0; x < 4; ++x) o.g[x] = __ldg(gp + x * kD);
      o.c = __ldg(P.c + row * kD + col);
      o.cp = s > 0 ? __ldg(P.c + (row - 1) * kD + col) : 0.f;
      o.p = __ldg(P.perm + row);
    }
  };
  StepIn cur_in;
  load_step(n_total - 1, cur_in);

  uint32_t r_d0 = 0, r_d1 = 0, r_bar0 = 0, r_bar1 = 0;
  if (lane == 0) {    // lane 0 of warp r pushes to rank r
    r_d0 = map_rank(s_u32(dg + (size_t)cta * kG * 512), warp);
    r_d1 = map_rank(s_u32(dg + (size_t)(kCluster + cta) * kG * 512), warp);
    r_bar0 = map_rank(s_u32(&bars[0]), warp); r_bar1 = map_rank(s_u32(&bars[1]), warp);
  }
  // fragment source of this lane inside source CTA `warp`'s slice: row of sequence g (clamped: columns >= kG are
  // never used), chunk q + 4j swizzled by the row's parity
  const int g_row = g < kG ? g : kG - 1;
  const int frag = warp * kG * 512 + g_row * 512, frag_sw = 4 * (g_row & 1);   // chunk (q + 4j) ^ frag_sw, + 256 for lo
  // where the owner lane's 4 gate gradients go inside the staged slice
  const int st_off = seq * 512 + (((u >> 1) ^ (4 * (seq & 1))) * 16) + (u & 1) * 8;
  cluster.sync();

  for (int r = 0; r < n_total; ++r) {
    const int gs = n_total - 1 - r;
    const bool dec = gs >= max_len;
    const int s = dec ? gs - max_len : gs;
    const int cur = r & 1;
    StepIn nxt;
    load_step(gs - 1, nxt);                 // in flight under the mat-vec
    if (r > 0) {
      if (n_dec > 0 && gs + 1 == max_len - 1) load_wt_mma(wh, wl, P.w_enc, cta, warp, lane);   // the later step is an encoder step from here on
      bar_wait(&bars[cur], ((r - 1) >> 1) & 1);
      if (tid == 0 && r + 2 < n_total) bar_arm(&bars[cur], kStepBytes);
      float acc[2][3][4];
#pragma unroll
      for (int m = 0; m < 2; ++m)
#pragma unroll
        for (int a = 0; a < 3; ++a)
#pragma unroll
          for (int x = 0; x < 4; ++x) acc[m][a][x] = 0.f;
      const uint8_t* src = dg + (size_t)cur * kCluster * kG * 512 + frag;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const uint4 bh = *reinterpret_cast<const uint4*>(src + ((q + 4 * j) ^ frag_sw) * 16);
        const uint4 bl = *reinterpret_cast<const uint4*>(src + ((q + 4 * j) ^ frag_sw) * 16 + 256);
#pragma unroll
        for (int m = 0; m < 2; ++m) {
          mma_bf16(acc[m][0], wh[m][2 * j], bh.x, bh.y);
          mma_bf16(acc[m][1], wl[m][2 * j], bh.x, bh.y);
          mma_bf16(acc[m][2], wh[m][2 * j], bl.x, bl.y);
          mma_bf16(acc[m][0], wh[m][2 * j + 1], bh.z, bh.w);
          mma_bf16(acc[m][1], wl[m][2 * j + 1], bh.z, bh.w);
          mma_bf16(acc[m][2], wh[m][2 * j + 1], bl.z, bl.w);
        }
      }
      // this warp's partial tile -> shared memory; the owner of (unit, sequence) adds the 8 partials
      float* mine = red + warp * 256;
#pragma unroll
      for (int m = 0; m < 2; ++m) {
        float c4[4];
#pragma unroll
        for (int x = 0; x < 4; ++x) c4[x] = (acc[m][1][x] + acc[m][2][x]) + acc[m][0][x];
        *reinterpret_cast<float2*>(mine + (m * 16 + g) * 8 + 2 * q) = make_float2(c4[0], c4[1]);
        *reinterpret_cast<float2*>(mine + (m * 16 + g + 8) * 8 + 2 * q) = make_float2(c4[2], c4[3]);
      }
      __syncthreads();
      float rsum = 0.f;
#pragma unroll
      for (int w = 0; w < 8; ++w) rsum += red[w * 256 + tid];
      carry_dh += rsum;
    }
    if (owner) {
      const bool active = my_b >= 0 && (dec || s < my_len);
      uint2 hi2 = make_uint2(0u, 0u), lo2 = make_uint2(0u, 0u);
      if (active) {
        const float ig = cur_in.g[0], fg = cur_in.g[1], gg = cur_in.g[2], og = cur_in.g[3];
        const float dh = carry_dh + cur_in.da;
        const float tc = fast_tanh(cur_in.c);
        const float dc = carry_dc + dh * og * (1.f - tc * tc);
        float d4[4];
        d4[0] = dc * gg * ig * (1.f - ig);
        d4[1] = dc * cur_in.cp * fg * (1.f - fg);
        d4[2] = dc * ig * (1.f - gg * gg);
        d4[3] = dh * tc * og * (1.f - og);
        carry_dc = dc * fg;
        carry_dh = 0.f;
        split2(d4[0], d4[1], hi2.x, lo2.x);
        split2(d4[2], d4[3], hi2.y, lo2.y);
        const size_t row = dec ? (size_t)my_b * n_dec + s : (size_t)my_b * Tmax + cur_in.p;
        uint16_t* oh = reinterpret_cast<uint16_t*>(dec ? P.ddec_hi : P.dxp_hi) + row * (4 * kD) + col;
        uint16_t* ol = reinterpret_cast<uint16_t*>(dec ? P.ddec_lo : P.dxp_lo) + row * (4 * kD) + col;
        oh[0] = (uint16_t)hi2.x; oh[kD] = (uint16_t)(hi2.x >> 16); oh[2 * kD] = (uint16_t)hi2.y; oh[3 * kD] = (uint16_t)(hi2.y >> 16);
        ol[0] = (uint16_t)lo2.x; ol[kD] = (uint16_t)(lo2.x >> 16); ol[2 * kD] = (uint16_t)lo2.y; ol[3 * kD] = (uint16_t)(lo2.y >> 16);
      }
      if (r + 1 < n_total) {
        uint8_t* st = stage + cur * kG * 512 + st_off;
        *reinterpret_cast<uint2*>(st) = hi2;
        *reinterpret_cast<uint2*>(st + 256) = lo2;
      }
    }
    if (r + 1 < n_total) {
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      __syncthreads();
      if (lane == 0)
        bulk_push(cur ? r_d0 : r_d1, s_u32(stage + cur * kG * 512), kG * 512, cur ? r_bar0 : r_bar1);
    }
    cur_in = nxt;
  }
  cluster.sync();
}

template <int kG>
static int launch_lstm_bwd(cudaLaunchConfig_t& cfg, const LstmBwd& P) {
  const int smem = 2 * kCluster * kG * 512 + 2 * kG * 512 + 8 * 32 * 8 * 4 + 16 + 64 + 128;
  static DeviceOnce attr;
  if (!attr.done()) {
    EDA_CUDA(cudaFuncSetAttribute(lstm_bwd_cluster_kernel<kG>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    attr.mark();
  }
  cfg.dynamicSmemBytes = smem;
  EDA_CUDA(cudaLaunchKernelEx(&cfg, lstm_bwd_cluster_kernel<kG>, P));
  count_launch();
  return EDA_OK;
}

int lstm_attractors_bwd(const LstmBwd& P, cudaStream_t st) {
  EDA_REQUIRE(P.perm && P.lens && P.w_enc && P.w_dec && P.gates && P.c && P.dec_gates && P.dec_c && P.hc && P.datt &&
              P.dxp_hi && P.dxp_lo && P.ddec_hi && P.ddec_lo, "eda_lstm_attractors_bwd: null argument");
  EDA_REQUIRE(P.B > 0 && P.Tmax > 0 && P.n_dec >= 1, "eda_lstm_attractors_bwd: bad sizes");
  // rows of padded frames (and of sequences that stop early) carry no gradient
  EDA_CUDA(cudaMemsetAsync(P.dxp_hi, 0, (size_t)P.B * P.Tmax * 4 * kD * sizeof(__nv_bfloat16), st));
  EDA_CUDA(cudaMemsetAsync(P.dxp_lo, 0, (size_t)P.B * P.Tmax * 4 * kD * sizeof(__nv_bfloat16), st));
  const int fit = std::max(1, lstm_max_active_clusters());
  int G = 4;
  while (G < 8 && ceil_div(P.B, G) > fit) ++G;
  cudaLaunchConfig_t cfg;
  cudaLaunchAttribute attr[1];
  lstm_launch_config(cfg, attr, ceil_div(P.B, G), st);
  switch (G) {
    case 4: return launch_lstm_bwd<4>(cfg, P);
    case 5: return launch_lstm_bwd<5>(cfg, P);
    case 6: return launch_lstm_bwd<6>(cfg, P);
    case 7: return launch_lstm_bwd<7>(cfg, P);
    default: return launch_lstm_bwd<8>(cfg, P);
  }
}

}  // namespace eda

extern "C" int eda_lstm_max_active_clusters(void) { return eda::lstm_max_active_clusters(); }

extern "C" int eda_lstm_attractors_bwd(const int32_t* perm, const int32_t* lens, const int32_t* order,
                                       const float* w_hh_enc, const float* w_hh_dec, const float* gates, const float* c,
                                       const float* dec_gates, const float* dec_c, const float* hc, const float* dattractors,
                                       int B, int Tmax, int D, int n_dec, void* dxproj_hi, void* dxproj_lo,
                                       void* ddec_hi, void* ddec_lo, void* stream) {
  EDA_REQUIRE(D == eda::kD, "eda_lstm_attractors_bwd: D must be 256 (got %d)", D);
  eda::LstmBwd P = {perm, lens, order, w_hh_enc, w_hh_dec, gates, c, dec_gates, dec_c, hc, dattractors,
                    (__nv_bfloat16*)dxproj_hi, (__nv_bfloat16*)dxproj_lo, (__nv_bfloat16*)ddec_hi, (__nv_bfloat16*)ddec_lo,
                    B, Tmax, n_dec};
  return eda::lstm_attractors_bwd(P, (cudaStream_t)stream);
}
