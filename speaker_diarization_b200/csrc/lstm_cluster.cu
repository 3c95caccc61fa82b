// EDA encoder/decoder LSTM recurrence as a persistent thread-block-cluster kernel.
//
// Reference: encoder_decoder_attractor.py:36-40 (packed encoder LSTM over the shuffled
// frames, decoder LSTM on zero inputs seeded with the encoder's final (h, c)).
//
// One cluster of 8 CTAs advances G sequences together.  CTA `c` of the cluster owns hidden
// units [32c, 32c+32), i.e. the 128 gate rows {i,f,g,o} x 32 units of W_hh.  Those rows stay
// RESIDENT IN REGISTERS for the whole sequence (128 rows x 256 k = 128 fp32 per thread at
// 256 threads), so the per-step critical path is: broadcast-read h from shared memory ->
// 128*G FFMA per thread -> combine the two k-halves through shared memory -> cell update ->
// write the 32 new h values to all 8 CTAs through distributed shared memory -> one cluster
// barrier.  The input projection emb.W_ih^T + b is hoisted out as a GEMM (eda_linear) and
// the random frame shuffle (models.py:228-232) is applied as an index into it.
#include <cooperative_groups.h>

#include "common.cuh"

namespace cg = cooperative_groups;

namespace eda {

constexpr int kD = 256;       // hidden size
constexpr int kCluster = 8;   // CTAs per cluster
constexpr int kUnits = kD / kCluster;  // 32 hidden units per CTA

__device__ __forceinline__ void load_w_rows(float (&w)[128], const float* __restrict__ W, int cta,
                                            int r_local, int kh) {
  // local row r_local = gate*32 + unit  ->  global row gate*256 + 32*cta + unit
  const int grow = (r_local >> 5) * kD + cta * kUnits + (r_local & 31);
  const float4* src = reinterpret_cast<const float4*>(W + (size_t)grow * kD + kh * 128);
#pragma unroll
  for (int i = 0; i < 32; ++i) {
    float4 v = __ldg(src + i);
    w[4 * i + 0] = v.x; w[4 * i + 1] = v.y; w[4 * i + 2] = v.z; w[4 * i + 3] = v.w;
  }
}

template <int G>
__global__ void __launch_bounds__(256, 1)
lstm_cluster_kernel(const float* __restrict__ xproj, const int32_t* __restrict__ perm,
                    const int32_t* __restrict__ lens, const int32_t* __restrict__ order,
                    const float* __restrict__ w_enc, const float* __restrict__ w_dec,
                    const float* __restrict__ dec_bias, int B, int Tmax, int n_dec,
                    float* __restrict__ attractors, float* __restrict__ hc_out) {
  cg::cluster_group cluster = cg::this_cluster();
  const int cta = (int)cluster.block_rank();
  const int group = blockIdx.x / kCluster;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int kh = warp & 1;                    // which half of k this warp reduces
  const int r_local = (warp >> 1) * 32 + lane;  // gate row (0..127) this thread accumulates

  __shared__ __align__(16) float hs[2][G][kD];     // h_{t-1} of the G sequences, double buffered
  __shared__ float part[2][G][128];                // k-half partial sums
  __shared__ int s_seq[G], s_len[G];

  if (tid < G) {
    int slot = group * G + tid;
    int b = -1, n = 0;
    if (slot < B) { b = order ? order[slot] : slot; n = lens[b]; }
    s_seq[tid] = b; s_len[tid] = n;
  }
  for (int i = tid; i < 2 * G * kD; i += 256) (&hs[0][0][0])[i] = 0.f;
  __syncthreads();
  int max_len = 0;
#pragma unroll
  for (int g = 0; g < G; ++g) max_len = max(max_len, s_len[g]);

  float w[128];
  load_w_rows(w, w_enc, cta, r_local, kh);

  // cell-update role: thread (cg_ = tid / 32, unit = lane) for cg_ < G
  const bool cell = warp < G;
  const int my_b = cell ? s_seq[warp] : -1;
  const int my_len = cell ? s_len[warp] : 0;
  float c_state = 0.f, h_state = 0.f;
  float xg[4] = {0.f, 0.f, 0.f, 0.f};
  int p_next = 0;
  const size_t col = (size_t)cta * kUnits + lane;
  auto load_x = [&](int p, float (&dst)[4]) {
    const float* row = xproj + ((size_t)my_b * Tmax + p) * (4 * kD) + col;
#pragma unroll
    for (int x = 0; x < 4; ++x) dst[x] = __ldg(row + x * kD);
  };
  if (cell && my_len > 0) {
    load_x(__ldg(perm + (size_t)my_b * Tmax), xg);
    if (my_len > 1) p_next = __ldg(perm + (size_t)my_b * Tmax + 1);
  }
  cluster.sync();   // every CTA's hs is zeroed before anybody writes into it remotely

  for (int phase = 0; phase < 2; ++phase) {
    const int n_steps = phase == 0 ? max_len : n_dec;
    if (phase == 1) {
      load_w_rows(w, w_dec, cta, r_local, kh);
      if (cell) {
#pragma unroll
        for (int x = 0; x < 4; ++x) xg[x] = __ldg(dec_bias + x * kD + col);
        if (hc_out && my_b >= 0) {
          hc_out[((size_t)my_b * 2 + 0) * kD + col] = h_state;
          hc_out[((size_t)my_b * 2 + 1) * kD + col] = c_state;
        }
      }
    }
    for (int s = 0; s < n_steps; ++s) {
      const int cur = (phase == 0 ? s : max_len + s) & 1;
      // prefetch next step's input projection while the mat-vec runs
      float xn[4] = {0.f, 0.f, 0.f, 0.f};
      int p_next2 = 0;
      if (phase == 0 && cell && s + 1 < my_len) {
        load_x(p_next, xn);
        if (s + 2 < my_len) p_next2 = __ldg(perm + (size_t)my_b * Tmax + s + 2);
      }
      float acc[G];
#pragma unroll
      for (int g = 0; g < G; ++g) acc[g] = 0.f;
#pragma unroll
      for (int i = 0; i < 128; i += 4) {
#pragma unroll
        for (int g = 0; g < G; ++g) {
          const float4 hv = *reinterpret_cast<const float4*>(&hs[cur][g][kh * 128 + i]);
          acc[g] = fmaf(w[i + 0], hv.x, acc[g]);
          acc[g] = fmaf(w[i + 1], hv.y, acc[g]);
          acc[g] = fmaf(w[i + 2], hv.z, acc[g]);
          acc[g] = fmaf(w[i + 3], hv.w, acc[g]);
        }
      }
#pragma unroll
      for (int g = 0; g < G; ++g) part[kh][g][r_local] = acc[g];
      __syncthreads();
      if (cell) {
        const bool active = phase == 1 ? (my_b >= 0) : (s < my_len);
        if (active) {
          float gate[4];
#pragma unroll
          for (int x = 0; x < 4; ++x)
            gate[x] = (part[0][warp][x * 32 + lane] + part[1][warp][x * 32 + lane]) + xg[x];
          const float ig = sigmoidf_ref(gate[0]), fg = sigmoidf_ref(gate[1]);
          const float gg = tanhf(gate[2]), og = sigmoidf_ref(gate[3]);
          c_state = fg * c_state + ig * gg;
          h_state = og * tanhf(c_state);
          if (phase == 1)
            attractors[((size_t)my_b * n_dec + s) * kD + col] = h_state;
        }
        // publish h (changed or not) into the next buffer of all 8 CTAs
        float* dst = &hs[cur ^ 1][warp][cta * kUnits + lane];
#pragma unroll
        for (int r = 0; r < kCluster; ++r) *cluster.map_shared_rank(dst, r) = h_state;
        if (phase == 0) {
#pragma unroll
          for (int x = 0; x < 4; ++x) xg[x] = xn[x];
          p_next = p_next2;
        }
      }
      cluster.sync();
    }
  }
}

template <int G>
static int launch_lstm(const float* xproj, const int32_t* perm, const int32_t* lens,
                       const int32_t* order, const float* w_enc, const float* w_dec,
                       const float* dec_bias, int B, int Tmax, int n_dec, float* attractors,
                       float* hc_out, cudaStream_t st) {
  const int groups = ceil_div(B, G);
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(groups * kCluster);
  cfg.blockDim = dim3(256);
  cfg.dynamicSmemBytes = 0;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = kCluster;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  EDA_CUDA(cudaLaunchKernelEx(&cfg, lstm_cluster_kernel<G>, xproj, perm, lens, order, w_enc, w_dec,
                              dec_bias, B, Tmax, n_dec, attractors, hc_out));
  count_launch();
  return EDA_OK;
}

int lstm_attractors(const float* xproj, const int32_t* perm, const int32_t* lens,
                    const int32_t* order, const float* w_enc, const float* w_dec,
                    const float* dec_bias, int B, int Tmax, int D, int n_dec, float* attractors,
                    float* hc_out, cudaStream_t st) {
  EDA_REQUIRE(D == kD, "eda_lstm_attractors: D must be 256 (got %d)", D);
  EDA_REQUIRE(B > 0 && Tmax > 0 && n_dec >= 0, "eda_lstm_attractors: bad sizes");
  EDA_REQUIRE(((uintptr_t)w_enc | (uintptr_t)w_dec) % 16 == 0, "eda_lstm_attractors: weights must be 16B aligned");
  // 16 clusters of 8 fit a B200 (2 per GPC): prefer the lower-latency G=4 while it fills them.
  if (B <= 64) return launch_lstm<4>(xproj, perm, lens, order, w_enc, w_dec, dec_bias, B, Tmax, n_dec, attractors, hc_out, st);
  return launch_lstm<8>(xproj, perm, lens, order, w_enc, w_dec, dec_bias, B, Tmax, n_dec, attractors, hc_out, st);
}

}  // namespace eda
