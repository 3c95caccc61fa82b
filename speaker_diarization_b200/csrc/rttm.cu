// Post-processing of posteriors into speaker segments (bin/make_rttm.py:29-42):
//   a = T_hat > threshold                               (:29)
//   a = scipy.signal.medfilt(a, (median, 1))            (:30-31, zero padded at the recording's ends)
//   per speaker: pad one zero frame on both sides, segments = pairs of change points   (:32-35)
// One CTA per (recording, speaker) column.  For 0/1 input the median of an odd window of k
// zero-padded samples is 1 iff more than k/2 of them are 1, so the filter is a windowed count.
// Rising / falling edges are numbered with a block-wide scan so that segment i of a column is
// (i-th rising edge, i-th falling edge), in time order - exactly the reference's zip of
// changes[::2] and changes[1::2].
#include "common.cuh"

namespace eda {

constexpr int kRttmThreads = 256;

template <typename SrcT>
__device__ __forceinline__ int active(const SrcT* col, int64_t t, int ld, float thr);
template <>
__device__ __forceinline__ int active<float>(const float* col, int64_t t, int ld, float thr) {
  return __ldg(col + t * ld) > thr ? 1 : 0;
}
template <>
__device__ __forceinline__ int active<uint8_t>(const uint8_t* col, int64_t t, int ld, float) {
  return __ldg(col + t * ld) != 0 ? 1 : 0;
}

template <typename SrcT>
__global__ void __launch_bounds__(kRttmThreads)
rttm_segments_kernel(const SrcT* __restrict__ src, int ld, const int64_t* __restrict__ rec_off,
                     const int32_t* __restrict__ n_spk_rec, int n_spk, float thr, int median, int cap,
                     int32_t* __restrict__ seg, int32_t* __restrict__ counts) {
  const int rec = blockIdx.y, spk = blockIdx.x;
  const int64_t t0 = rec_off[rec];
  const int T = (int)(rec_off[rec + 1] - t0);
  int32_t* out = seg + ((size_t)rec * n_spk + spk) * cap * 2;
  if ((n_spk_rec && spk >= n_spk_rec[rec]) || T <= 0) {
    if (threadIdx.x == 0) counts[rec * n_spk + spk] = 0;
    return;
  }
  const SrcT* col = src + t0 * ld + spk;
  const int h = median > 1 ? median / 2 : 0;
  // filtered decision at frame t; 0 outside [0, T)
  auto filt = [&](int t) -> int {
    if (t < 0 || t >= T) return 0;
    if (h == 0) return active<SrcT>(col, t, ld, thr);
    int c = 0;
    const int lo = max(0, t - h), hi = min(T - 1, t + h);
    for (int u = lo; u <= hi; ++u) c += active<SrcT>(col, u, ld, thr);
    return c > h ? 1 : 0;
  };
  __shared__ int s_warp_r[kRttmThreads / 32], s_warp_f[kRttmThreads / 32];
  __shared__ int s_base_r, s_base_f;
  if (threadIdx.x == 0) { s_base_r = 0; s_base_f = 0; }
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  // change point at index t (0..T) between padded frames t-1 and t
  for (int base = 0; base <= T; base += kRttmThreads) {
    const int t = base + threadIdx.x;
    int rise = 0, fall = 0;
    if (t <= T) {
      const int prev = filt(t - 1), cur = filt(t);
      rise = (cur && !prev); fall = (!cur && prev);
    }
    const unsigned mr = __ballot_sync(0xffffffffu, rise), mf = __ballot_sync(0xffffffffu, fall);
    if (lane == 0) { s_warp_r[warp] = __popc(mr); s_warp_f[warp] = __popc(mf); }
    __syncthreads();
    int off_r = s_base_r, off_f = s_base_f;
    for (int w = 0; w < warp; ++w) { off_r += s_warp_r[w]; off_f += s_warp_f[w]; }
    const unsigned below = (1u << lane) - 1u;
    if (rise) { const int i = off_r + __popc(mr & below); if (i < cap) out[2 * i] = t; }
    if (fall) { const int i = off_f + __popc(mf & below); if (i < cap) out[2 * i + 1] = t; }
    __syncthreads();
    if (threadIdx.x == 0) {
      int r = 0, f = 0;
      for (int w = 0; w < kRttmThreads / 32; ++w) { r += s_warp_r[w]; f += s_warp_f[w]; }
      s_base_r += r; s_base_f += f;
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) counts[rec * n_spk + spk] = s_base_r;      // may exceed cap: caller checks
}

int rttm_segments(const float* post, const uint8_t* labels, int ld, const int64_t* rec_off, int n_rec,
                  const int32_t* n_spk_rec, int n_spk, float thr, int median, int cap, int32_t* seg,
                  int32_t* counts, cudaStream_t st) {
  EDA_REQUIRE((post != nullptr) != (labels != nullptr), "eda_rttm_segments: pass exactly one of post / labels");
  EDA_REQUIRE(rec_off && seg && counts && n_rec > 0 && n_spk > 0 && cap > 0 && ld >= n_spk, "eda_rttm_segments: bad arguments");
  EDA_REQUIRE(median >= 1 && (median & 1), "eda_rttm_segments: median must be odd and >= 1 (got %d)", median);
  EDA_REQUIRE(n_rec <= 65535, "eda_rttm_segments: too many recordings for one launch");
  dim3 grid(n_spk, n_rec);
  if (post)
    rttm_segments_kernel<float><<<grid, kRttmThreads, 0, st>>>(post, ld, rec_off, n_spk_rec, n_spk, thr, median, cap, seg, counts);
  else
    rttm_segments_kernel<uint8_t><<<grid, kRttmThreads, 0, st>>>(labels, ld, rec_off, n_spk_rec, n_spk, thr, median, cap, seg, counts);
  EDA_LAUNCH_CHECK();
  return EDA_OK;
}

}  // namespace eda

extern "C" int eda_rttm_segments(const float* post, const uint8_t* labels, int ld, const int64_t* rec_off, int n_rec,
                                 const int32_t* n_spk_rec, int n_spk, float threshold, int median, int cap,
                                 int32_t* segments, int32_t* counts, void* stream) {
  return eda::rttm_segments(post, labels, ld, rec_off, n_rec, n_spk_rec, n_spk, threshold, median, cap, segments,
                            counts, (cudaStream_t)stream);
}
