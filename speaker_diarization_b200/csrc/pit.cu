// Batched PIT permutation-search loss, its gradient, and the DER counters.
//
// Reference: TransformerEdaModel.pit_loss / pit_loss_single_permute / create_length_mask
// (models.py:373-421) and calc_diarization_error (models.py:423-463).
//
// The reference evaluates S! full masked-BCE passes.  Here one pass builds the S x S cost
// matrix   C[i][j] = sum_{t<len} BCE(pred[t,i], label[t,j])
//                  = sum_t (max(x,0) + log1p(exp(-|x|)))  -  sum_t x * y        (x=pred_i, y=label_j)
// in fp64 with a FIXED summation order (so identical label columns give bit-identical
// costs and exact ties survive), then the S! permutation costs sum_s C[s][p[s]] / S are
// searched in itertools.permutations (= lexicographic) order with lowest-index tie-break.
#include <algorithm>

#include "common.cuh"

namespace eda {

constexpr int kMaxS = 8;

__device__ __forceinline__ void kth_permutation(int k, int S, int* p) {
  // lexicographic rank -> permutation (factoradic decode); == itertools.permutations order
  int avail[kMaxS];
  for (int i = 0; i < S; ++i) avail[i] = i;
  int fact = 1;
  for (int i = 2; i < S; ++i) fact *= i;      // (S-1)!
  int n = S;
  for (int i = 0; i < S; ++i) {
    int q = k / fact;
    k -= q * fact;
    p[i] = avail[q];
    for (int j = q; j < n - 1; ++j) avail[j] = avail[j + 1];
    --n;
    if (n > 1) fact /= n; else fact = 1;
  }
}

template <int S>
__global__ void __launch_bounds__(256)
pit_kernel(const float* __restrict__ pred, const float* __restrict__ label,
           const int32_t* __restrict__ lens, int B, int T, float* __restrict__ loss,
           int64_t* __restrict__ min_idx, float* __restrict__ label_perm,
           int32_t* __restrict__ perm_out, double* __restrict__ best_cost,
           unsigned int* __restrict__ done_counter) {
  const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int len = lens[b];
  const float* P = pred + (size_t)b * T * S;
  const float* L = label + (size_t)b * T * S;

  // ---- pass 1: softplus sums and cross sums, per-thread strided over t, fp64 ----------
  double sp[S], xy[S][S];
#pragma unroll
  for (int i = 0; i < S; ++i) {
    sp[i] = 0.0;
#pragma unroll
    for (int j = 0; j < S; ++j) xy[i][j] = 0.0;
  }
  for (int t = tid; t < len; t += 256) {
    float x[S], y[S];
#pragma unroll
    for (int i = 0; i < S; ++i) { x[i] = __ldg(P + (size_t)t * S + i); y[i] = __ldg(L + (size_t)t * S + i); }
#pragma unroll
    for (int i = 0; i < S; ++i) {
      const double xd = (double)x[i];
      sp[i] += fmax(xd, 0.0) + log1p(exp(-fabs(xd)));
#pragma unroll
      for (int j = 0; j < S; ++j) xy[i][j] += xd * (double)y[j];
    }
  }
  __shared__ double red[8][S * S + S];
  __shared__ double C[S][S];
#pragma unroll
  for (int i = 0; i < S; ++i) {
    double v = warp_sum(sp[i]);
    if (lane == 0) red[warp][S * S + i] = v;
#pragma unroll
    for (int j = 0; j < S; ++j) {
      double u = warp_sum(xy[i][j]);
      if (lane == 0) red[warp][i * S + j] = u;
    }
  }
  __syncthreads();
  if (tid < S * S) {
    const int i = tid / S, j = tid % S;
    double a = 0.0, x = 0.0;
    for (int w = 0; w < 8; ++w) { a += red[w][S * S + i]; x += red[w][i * S + j]; }
    C[i][j] = a - x;
  }
  __syncthreads();

  // ---- pass 2: permutation search ------------------------------------------------------
  int nperm = 1;
#pragma unroll
  for (int i = 2; i <= S; ++i) nperm *= i;
  double best = INFINITY;
  int best_k = 0x7fffffff;
  for (int k = tid; k < nperm; k += 256) {
    int p[kMaxS];
    kth_permutation(k, S, p);
    double c = 0.0;
#pragma unroll
    for (int s = 0; s < S; ++s) c += C[s][p[s]];
    if (c < best) { best = c; best_k = k; }    // k ascending per thread: first minimum kept
  }
  __shared__ double s_best[256];
  __shared__ int s_k[256];
  s_best[tid] = best; s_k[tid] = best_k;
  __syncthreads();
  for (int off = 128; off > 0; off >>= 1) {
    if (tid < off) {
      const double c2 = s_best[tid + off]; const int k2 = s_k[tid + off];
      if (c2 < s_best[tid] || (c2 == s_best[tid] && k2 < s_k[tid])) { s_best[tid] = c2; s_k[tid] = k2; }
    }
    __syncthreads();
  }
  __shared__ int s_perm[kMaxS];
  if (tid == 0) {
    int p[kMaxS];
    kth_permutation(s_k[0], S, p);
    for (int s = 0; s < S; ++s) { s_perm[s] = p[s]; if (perm_out) perm_out[b * S + s] = p[s]; }
    min_idx[b] = s_k[0];
    best_cost[b] = s_best[0] / (double)S;
  }
  __syncthreads();

  // ---- pass 3: label_perm[b,t,s] = label[b,t,perm[s]] over ALL T rows (padding keeps -1)
  if (label_perm) {
    float* O = label_perm + (size_t)b * T * S;
    for (int i = tid; i < T * S; i += 256) {
      const int t = i / S, s = i - t * S;
      O[i] = __ldg(L + (size_t)t * S + s_perm[s]);
    }
  }

  // ---- last CTA: loss = sum_b best / sum_b len, fixed order -----------------------------
  __shared__ bool is_last;
  __threadfence();
  if (tid == 0) is_last = atomicAdd(done_counter, 1u) == (unsigned)(B - 1);
  __syncthreads();
  if (is_last && tid == 0) {
    __threadfence();
    double tot = 0.0; long long nfr = 0;
    for (int i = 0; i < B; ++i) { tot += ((volatile double*)best_cost)[i]; nfr += lens[i]; }
    loss[0] = (float)(tot / (double)nfr);
    *done_counter = 0;
  }
}

template <int S>
static void launch_pit(const float* pred, const float* label, const int32_t* lens, int B, int T,
                       float* loss, int64_t* min_idx, float* label_perm, int32_t* perm_out,
                       double* best, unsigned* counter, cudaStream_t st) {
  pit_kernel<S><<<B, 256, 0, st>>>(pred, label, lens, B, T, loss, min_idx, label_perm, perm_out, best, counter);
}

// scratch (best cost per sequence + completion counter): library-owned, one buffer per (device, stream),
// so calls in flight on different streams or GPUs never share the counter; the counter is zeroed on the
// launch stream before every launch.
int pit_loss(const float* pred, const float* label, const int32_t* lens, int B, int T, int S,
             float* loss, int64_t* min_idx, float* label_perm, int32_t* perm_out, cudaStream_t st) {
  EDA_REQUIRE(S >= 1 && S <= kMaxS, "eda_pit_loss: S must be in [1,%d] (got %d)", kMaxS, S);
  EDA_REQUIRE(B > 0 && T > 0 && loss && min_idx, "eda_pit_loss: bad arguments");
  unsigned char* scratch = (unsigned char*)stream_scratch(4, st, 256 + sizeof(double) * (size_t)std::max(B, 1024));
  if (!scratch) return EDA_ECUDA;
  unsigned* g_counter = reinterpret_cast<unsigned*>(scratch);
  double* g_best = reinterpret_cast<double*>(scratch + 256);
  EDA_CUDA(cudaMemsetAsync(g_counter, 0, sizeof(unsigned), st));
#define EDA_PIT_CASE(s) case s: launch_pit<s>(pred, label, lens, B, T, loss, min_idx, label_perm, perm_out, g_best, g_counter, st); break;
  switch (S) {
    EDA_PIT_CASE(1) EDA_PIT_CASE(2) EDA_PIT_CASE(3) EDA_PIT_CASE(4)
    EDA_PIT_CASE(5) EDA_PIT_CASE(6) EDA_PIT_CASE(7) EDA_PIT_CASE(8)
  }
#undef EDA_PIT_CASE
  EDA_LAUNCH_CHECK();
  return EDA_OK;
}

// ---- eda_pit_loss_static: the same kernel with its scratch in a file-scope device array -------
// For callers that may not allocate (CUDA graph capture, serving loops): no cudaMalloc / cudaFree,
// no growth.  Layout of the array: [0, 8 * kStaticMaxB) best cost per sequence (double),
// then one unsigned completion counter on its own 256-byte line.  Every GPU that loads the
// library holds these bytes until exit; every call on a device shares them, so calls that overlap
// on the GPU (two streams) race - serialise them.
constexpr int kStaticMaxB = EDA_PIT_STATIC_MAX_B;
constexpr size_t kStaticBytes = sizeof(double) * kStaticMaxB + 256;
static __device__ __align__(256) unsigned char g_pit_static[kStaticBytes];

// In host code `g_pit_static` is only a stand-in for the symbol: the device address is looked up once
// per device, with that device current.
constexpr int kMaxDevices = 64;
static void* g_pit_static_addr[kMaxDevices] = {};

int pit_loss_static(const float* pred, const float* label, const int32_t* lens, int B, int T, int S,
                    float* loss, int64_t* min_idx, float* label_perm, int32_t* perm_out, cudaStream_t st) {
  EDA_REQUIRE(S >= 1 && S <= kMaxS, "eda_pit_loss_static: S must be in [1,%d] (got %d)", kMaxS, S);
  EDA_REQUIRE(B > 0 && T > 0 && loss && min_idx, "eda_pit_loss_static: bad arguments");
  EDA_REQUIRE(B <= kStaticMaxB,
              "eda_pit_loss_static: B = %d exceeds the static scratch (EDA_PIT_STATIC_MAX_B = %d sequences); "
              "use eda_pit_loss", B, kStaticMaxB);
  int dev = -1;
  EDA_CUDA(cudaGetDevice(&dev));
  EDA_REQUIRE(dev >= 0 && dev < kMaxDevices, "eda_pit_loss_static: device ordinal %d not supported", dev);
  void* base = __atomic_load_n(&g_pit_static_addr[dev], __ATOMIC_ACQUIRE);
  if (!base) {
    EDA_CUDA(cudaGetSymbolAddress(&base, g_pit_static));
    EDA_REQUIRE(base != nullptr, "eda_pit_loss_static: cudaGetSymbolAddress returned a null address");
    __atomic_store_n(&g_pit_static_addr[dev], base, __ATOMIC_RELEASE);
  }
  double* best = reinterpret_cast<double*>(base);
  unsigned* counter = reinterpret_cast<unsigned*>(static_cast<unsigned char*>(base) + sizeof(double) * kStaticMaxB);
  // the array keeps the previous call's contents: the counter must start at zero (the best costs are
  // all rewritten before they are read)
  EDA_CUDA(cudaMemsetAsync(counter, 0, sizeof(unsigned), st));
#define EDA_PIT_CASE(s) case s: launch_pit<s>(pred, label, lens, B, T, loss, min_idx, label_perm, perm_out, best, counter, st); break;
  switch (S) {
    EDA_PIT_CASE(1) EDA_PIT_CASE(2) EDA_PIT_CASE(3) EDA_PIT_CASE(4)
    EDA_PIT_CASE(5) EDA_PIT_CASE(6) EDA_PIT_CASE(7) EDA_PIT_CASE(8)
  }
#undef EDA_PIT_CASE
  EDA_LAUNCH_CHECK();
  return EDA_OK;
}

// d loss / d pred (models.py:373-413 differentiated): mask*(sigmoid(x) - y)/(S*sum len)
__global__ void pit_bwd_kernel(const float* __restrict__ pred, const float* __restrict__ lab,
                               const int32_t* __restrict__ lens, int B, int T, int S, float scale,
                               float* __restrict__ dpred) {
  long long nfr = 0;
  for (int i = 0; i < B; ++i) nfr += lens[i];
  const float k = scale / ((float)S * (float)nfr);
  const int64_t n = (int64_t)B * T * S;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t bt = i / S;
    const int b = (int)(bt / T), t = (int)(bt % T);
    dpred[i] = t < lens[b] ? (sigmoidf_ref(pred[i]) - lab[i]) * k : 0.f;
  }
}

int pit_loss_bwd(const float* pred, const float* lab, const int32_t* lens, int B, int T, int S,
                 float scale, float* dpred, cudaStream_t st) {
  EDA_REQUIRE(B > 0 && T > 0 && S > 0, "eda_pit_loss_bwd: bad sizes");
  const int64_t n = (int64_t)B * T * S;
  const int blocks = (int)std::min<int64_t>(ceil_div<int64_t>(n, 256), (int64_t)num_sms() * 8);
  pit_bwd_kernel<<<blocks, 256, 0, st>>>(pred, lab, lens, B, T, S, scale, dpred);
  EDA_LAUNCH_CHECK();
  return EDA_OK;
}

// calc_diarization_error (models.py:423-463): integer counters, exact under any order.
__global__ void der_kernel(const float* __restrict__ pred, const float* __restrict__ lab,
                           const int32_t* __restrict__ lens, int B, int T, int S,
                           unsigned long long* __restrict__ counts) {
  unsigned long long c[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
  const int64_t n = (int64_t)B * T;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int b = (int)(i / T), t = (int)(i % T);
    if (t >= lens[b]) continue;
    int n_ref = 0, n_sys = 0, n_map = 0, n_eq = 0;
    for (int s = 0; s < S; ++s) {
      const int l = (int)lab[i * S + s];            // .astype(int): truncation
      const int d = pred[i * S + s] > 0.f ? 1 : 0;
      n_ref += l; n_sys += d; n_map += (l == 1 && d == 1); n_eq += (l == d);
    }
    c[0] += n_eq; c[1] += 1;
    c[2] += n_ref > 0; c[3] += (n_ref > 0 && n_sys == 0); c[4] += (n_ref == 0 && n_sys > 0);
    c[5] += n_ref; c[6] += max(n_ref - n_sys, 0); c[7] += max(n_sys - n_ref, 0);
    c[8] += min(n_ref, n_sys) - n_map;
  }
#pragma unroll
  for (int k = 0; k < 9; ++k) {
    unsigned long long v = c[k];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) == 0 && v) atomicAdd(counts + k, v);
  }
}

int der_counters(const float* pred, const float* lab, const int32_t* lens, int B, int T, int S,
                 int64_t* counts, cudaStream_t st) {
  EDA_REQUIRE(B > 0 && T > 0 && S > 0 && counts, "eda_der_counters: bad arguments");
  EDA_CUDA(cudaMemsetAsync(counts, 0, 9 * sizeof(int64_t), st));
  const int64_t n = (int64_t)B * T;
  const int blocks = (int)std::min<int64_t>(ceil_div<int64_t>(n, 256), (int64_t)num_sms() * 4);
  der_kernel<<<blocks, 256, 0, st>>>(pred, lab, lens, B, T, S, reinterpret_cast<unsigned long long*>(counts));
  EDA_LAUNCH_CHECK();
  return EDA_OK;
}

}  // namespace eda
