// Feature front end: audio -> log-mel (pass 1) -> mean-norm + splice + subsample (pass 2).
//
// Reference: feature.stft (feature.py:155-184), feature.transform 'logmel23_mn'
// (feature.py:64-73), feature.splice (feature.py:130-152), feature.subsample (:122-127).
// The reference runs this in float64 numpy (soundfile hands it float64 audio); to stay
// within ~1e-6 of it the arithmetic here is fp64 as well (the kernel is small next to the
// encoder: ~8 kFLOP per STFT frame).
//
// Pass 1: one warp per STFT frame.  The windowed frame (N = n_fft real samples) is packed
// into N/2 complex points, transformed by a shared-memory radix-2 Stockham FFT, untangled
// into the N/2+1 one-sided bins, squared, and reduced by the 23 Slaney triangles (lane m
// owns filter m and walks its own bin range in ascending order -> deterministic sums).
#include <math.h>

#include <vector>

#include <algorithm>

#include "common.cuh"

struct eda_frontend_plan {
  int sample_rate, frame_size, frame_shift, n_mels, n_fft, n_bins, lpad;
  double* d_window;    // (n_fft) zero-padded periodic Hann
  double2* d_tw;       // (n_fft/2) exp(-2 pi i k / n_fft)
  float* d_melw;       // (n_mels, n_bins) dense Slaney filterbank, float32 like librosa
  int* d_mel_lo;       // (n_mels) first bin with a non-zero weight
  int* d_mel_hi;       // (n_mels) last bin with a non-zero weight
};

namespace eda {

// ---- host: tables (librosa 0.10.2 published algorithm: Slaney scale + area norm) -------
static double hz_to_mel(double f) {
  const double f_sp = 200.0 / 3, min_log_hz = 1000.0, min_log_mel = min_log_hz / f_sp;
  const double logstep = log(6.4) / 27.0;
  return f >= min_log_hz ? min_log_mel + log(f / min_log_hz) / logstep : f / f_sp;
}
static double mel_to_hz(double m) {
  const double f_sp = 200.0 / 3, min_log_hz = 1000.0, min_log_mel = min_log_hz / f_sp;
  const double logstep = log(6.4) / 27.0;
  return m >= min_log_mel ? min_log_hz * exp(logstep * (m - min_log_mel)) : f_sp * m;
}

static void mel_filterbank(int sr, int n_fft, int n_mels, std::vector<float>& w) {
  const int n_bins = n_fft / 2 + 1;
  std::vector<double> mel_f(n_mels + 2);
  const double m0 = hz_to_mel(0.0), m1 = hz_to_mel(sr / 2.0), step = (m1 - m0) / (n_mels + 1);
  for (int i = 0; i < n_mels + 2; ++i) mel_f[i] = mel_to_hz(i == n_mels + 1 ? m1 : m0 + i * step);
  w.assign((size_t)n_mels * n_bins, 0.f);
  for (int i = 0; i < n_mels; ++i) {
    const double enorm = 2.0 / (mel_f[i + 2] - mel_f[i]);
    for (int k = 0; k < n_bins; ++k) {
      const double f = (double)k * (1.0 / (n_fft * (1.0 / sr)));   // np.fft.rfftfreq
      const double lower = -(mel_f[i] - f) / (mel_f[i + 1] - mel_f[i]);
      const double upper = (mel_f[i + 2] - f) / (mel_f[i + 2] - mel_f[i + 1]);
      const float tri = (float)fmax(0.0, fmin(lower, upper));   // float32 storage, then *= enorm
      w[(size_t)i * n_bins + k] = (float)((double)tri * enorm);  // in-place float32 *= float64
    }
  }
}

// ---- pass 1 ---------------------------------------------------------------------------
__device__ __forceinline__ double2 cmul(double2 a, double2 b) {
  return make_double2(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x);
}

template <typename AudioT>
__global__ void __launch_bounds__(256)
logmel_kernel(const AudioT* __restrict__ audio, const int64_t* __restrict__ rec_sample_off,
              const int64_t* __restrict__ rec_frame_off, int n_rec, int64_t total_frames,
              int n_fft, int hop, int n_mels, const double* __restrict__ window,
              const double2* __restrict__ tw, const float* __restrict__ melw,
              const int* __restrict__ mel_lo, const int* __restrict__ mel_hi,
              float* __restrict__ logmel, double* __restrict__ colsum, int frames_per_warp) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int M = n_fft >> 1, n_bins = M + 1;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // layout: [tw: M double2][window: n_fft double][melw: n_mels*n_bins float][per-warp: 2*M double2]
  double2* s_tw = reinterpret_cast<double2*>(smem_raw);
  double* s_win = reinterpret_cast<double*>(s_tw + M);
  float* s_melw = reinterpret_cast<float*>(s_win + n_fft);
  size_t off = (size_t)M * 16 + (size_t)n_fft * 8 + (((size_t)n_mels * n_bins * 4 + 15) / 16) * 16;
  double2* bufA = reinterpret_cast<double2*>(smem_raw + off) + (size_t)warp * 2 * M;
  double2* bufB = bufA + M;
  for (int i = threadIdx.x; i < M; i += blockDim.x) s_tw[i] = tw[i];
  for (int i = threadIdx.x; i < n_fft; i += blockDim.x) s_win[i] = window[i];
  for (int i = threadIdx.x; i < n_mels * n_bins; i += blockDim.x) s_melw[i] = melw[i];
  __syncthreads();
  const int my_lo = lane < n_mels ? mel_lo[lane] : 0, my_hi = lane < n_mels ? mel_hi[lane] : -1;

  const int64_t gwarp = (int64_t)blockIdx.x * (blockDim.x >> 5) + warp;
  const int64_t f_begin = gwarp * frames_per_warp;
  const int64_t f_end = min(total_frames, f_begin + frames_per_warp);
  if (f_begin >= f_end) return;

  // locate the recording of the first frame (largest r with rec_frame_off[r] <= f)
  int rec = 0;
  {
    int lo = 0, hi = n_rec - 1;
    while (lo < hi) {
      int mid = (lo + hi + 1) >> 1;
      if (rec_frame_off[mid] <= f_begin) lo = mid; else hi = mid - 1;
    }
    rec = lo;
  }
  double csum = 0.0;   // lane m: running sum of log-mel bin m over this warp's frames of `rec`
  int64_t r_f0 = rec_frame_off[rec], r_f1 = rec_frame_off[rec + 1];
  int64_t r_s0 = rec_sample_off[rec], r_n = rec_sample_off[rec + 1] - r_s0;

  for (int64_t f = f_begin; f < f_end; ++f) {
    while (f >= r_f1) {   // crossed into the next recording (empty recordings are skipped)
      if (lane < n_mels) atomicAdd(colsum + (size_t)rec * n_mels + lane, csum);
      csum = 0.0;
      ++rec;
      r_f0 = rec_frame_off[rec]; r_f1 = rec_frame_off[rec + 1];
      r_s0 = rec_sample_off[rec]; r_n = rec_sample_off[rec + 1] - r_s0;
    }
    // sample index (within the recording) of padded position 0 of this frame
    const int64_t s_start = (f - r_f0) * hop - M;
    for (int n = lane; n < M; n += 32) {
      const int64_t i0 = s_start + 2 * n, i1 = i0 + 1;
      const double x0 = (i0 >= 0 && i0 < r_n) ? (double)audio[r_s0 + i0] : 0.0;
      const double x1 = (i1 >= 0 && i1 < r_n) ? (double)audio[r_s0 + i1] : 0.0;
      bufA[n] = make_double2(x0 * s_win[2 * n], x1 * s_win[2 * n + 1]);
    }
    __syncwarp();
    double2* in = bufA;
    double2* out = bufB;
    for (int Ns = 1; Ns < M; Ns <<= 1) {
      const int tw_step = n_fft / (2 * Ns);       // exp(-2 pi i k / (2 Ns)) = tw[k * n_fft/(2 Ns)]
      for (int j = lane; j < (M >> 1); j += 32) {
        const int k = j & (Ns - 1);
        const double2 u = in[j];
        const double2 v = cmul(in[j + (M >> 1)], s_tw[k * tw_step]);
        const int j0 = ((j - k) << 1) + k;
        out[j0] = make_double2(u.x + v.x, u.y + v.y);
        out[j0 + Ns] = make_double2(u.x - v.x, u.y - v.y);
      }
      __syncwarp();
      double2* t = in; in = out; out = t;
    }
    // untangle into one-sided bins, power spectrum into `out` (as doubles)
    double* pw = reinterpret_cast<double*>(out);
    for (int k = lane; k <= M; k += 32) {
      double re, im;
      if (k == 0 || k == M) {
        re = k == 0 ? in[0].x + in[0].y : in[0].x - in[0].y;
        im = 0.0;
      } else {
        const double2 zk = in[k], zc = make_double2(in[M - k].x, -in[M - k].y);
        const double2 E = make_double2(0.5 * (zk.x + zc.x), 0.5 * (zk.y + zc.y));
        const double2 d = make_double2(zk.x - zc.x, zk.y - zc.y);
        const double2 O = make_double2(0.5 * d.y, -0.5 * d.x);      // -i/2 * d
        const double2 wo = cmul(s_tw[k], O);
        re = E.x + wo.x; im = E.y + wo.y;
      }
      pw[k] = re * re + im * im;
    }
    __syncwarp();
    if (lane < n_mels) {
      double acc = 0.0;
      const float* wrow = s_melw + (size_t)lane * n_bins;
      for (int k = my_lo; k <= my_hi; ++k) acc += pw[k] * (double)wrow[k];
      const double v = log10(fmax(acc, 1e-10));
      logmel[(size_t)f * n_mels + lane] = (float)v;
      csum += v;
    }
    __syncwarp();
  }
  if (lane < n_mels) atomicAdd(colsum + (size_t)rec * n_mels + lane, csum);
}

// ---- pass 2 ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
splice_kernel(const float* __restrict__ logmel, const double* __restrict__ colsum,
              const int64_t* __restrict__ rec_frame_off, const int64_t* __restrict__ rec_row_off,
              int n_rec, int64_t total_rows, int n_mels, int context, int subsampling,
              float* __restrict__ feats, int ld_feats) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int width = (2 * context + 1) * n_mels;
  for (int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 5) + warp; row < total_rows;
       row += (int64_t)gridDim.x * (blockDim.x >> 5)) {
    int lo = 0, hi = n_rec - 1;
    while (lo < hi) {
      int mid = (lo + hi + 1) >> 1;
      if (rec_row_off[mid] <= row) lo = mid; else hi = mid - 1;
    }
    const int rec = lo;
    const int64_t f0 = rec_frame_off[rec], nfr = rec_frame_off[rec + 1] - f0;
    const int64_t t = (row - rec_row_off[rec]) * subsampling;
    const double inv_n = 1.0 / (double)nfr;
    float* dst = feats + row * (int64_t)ld_feats;
    for (int c = lane; c < ld_feats; c += 32) {
      float v = 0.f;
      if (c < width) {
        const int j = c / n_mels, m = c - j * n_mels;
        const int64_t src = t - context + j;
        if (src >= 0 && src < nfr)
          v = (float)((double)__ldg(logmel + (f0 + src) * n_mels + m) - colsum[(size_t)rec * n_mels + m] * inv_n);
      }
      dst[c] = v;
    }
  }
}

static size_t logmel_smem(int n_fft, int n_mels) {
  const int M = n_fft / 2, n_bins = M + 1;
  return (size_t)M * 16 + (size_t)n_fft * 8 + (((size_t)n_mels * n_bins * 4 + 15) / 16) * 16 +
         (size_t)8 * 2 * M * 16;
}

}  // namespace eda

extern "C" {

int64_t eda_frontend_num_frames(int64_t n_samples, int frame_shift) {
  int64_t n = 1 + n_samples / frame_shift;
  return (n_samples % frame_shift == 0) ? n - 1 : n;
}

int eda_frontend_plan_create(int sample_rate, int frame_size, int frame_shift, int n_mels,
                             eda_frontend_plan** out) {
  using namespace eda;
  EDA_REQUIRE(out && sample_rate > 0 && frame_size > 1 && frame_shift > 0, "frontend_plan_create: bad arguments");
  EDA_REQUIRE(n_mels >= 1 && n_mels <= 32, "frontend_plan_create: n_mels must be <= 32 (one lane per filter)");
  int n_fft = 1;
  while (n_fft < frame_size) n_fft <<= 1;
  EDA_REQUIRE(n_fft >= 64 && n_fft <= 2048, "frontend_plan_create: n_fft %d out of range [64,2048]", n_fft);
  eda_frontend_plan* p = new eda_frontend_plan();
  p->sample_rate = sample_rate; p->frame_size = frame_size; p->frame_shift = frame_shift;
  p->n_mels = n_mels; p->n_fft = n_fft; p->n_bins = n_fft / 2 + 1; p->lpad = (n_fft - frame_size) / 2;
  const double kPi = 3.14159265358979323846;
  std::vector<double> win(n_fft, 0.0);
  for (int i = 0; i < frame_size; ++i) win[p->lpad + i] = 0.5 - 0.5 * cos(2.0 * kPi * i / frame_size);
  std::vector<double2> tw(n_fft / 2);
  for (int k = 0; k < n_fft / 2; ++k) tw[k] = make_double2(cos(2.0 * kPi * k / n_fft), -sin(2.0 * kPi * k / n_fft));
  std::vector<float> melw;
  mel_filterbank(sample_rate, n_fft, n_mels, melw);
  std::vector<int> lo(n_mels), hi(n_mels);
  for (int m = 0; m < n_mels; ++m) {
    lo[m] = p->n_bins; hi[m] = -1;
    for (int k = 0; k < p->n_bins; ++k)
      if (melw[(size_t)m * p->n_bins + k] != 0.f) { lo[m] = std::min(lo[m], k); hi[m] = k; }
    if (hi[m] < 0) lo[m] = 0;
  }
  EDA_CUDA(cudaMalloc(&p->d_window, sizeof(double) * n_fft));
  EDA_CUDA(cudaMalloc(&p->d_tw, sizeof(double2) * (n_fft / 2)));
  EDA_CUDA(cudaMalloc(&p->d_melw, sizeof(float) * melw.size()));
  EDA_CUDA(cudaMalloc(&p->d_mel_lo, sizeof(int) * n_mels));
  EDA_CUDA(cudaMalloc(&p->d_mel_hi, sizeof(int) * n_mels));
  EDA_CUDA(cudaMemcpy(p->d_window, win.data(), sizeof(double) * n_fft, cudaMemcpyHostToDevice));
  EDA_CUDA(cudaMemcpy(p->d_tw, tw.data(), sizeof(double2) * (n_fft / 2), cudaMemcpyHostToDevice));
  EDA_CUDA(cudaMemcpy(p->d_melw, melw.data(), sizeof(float) * melw.size(), cudaMemcpyHostToDevice));
  EDA_CUDA(cudaMemcpy(p->d_mel_lo, lo.data(), sizeof(int) * n_mels, cudaMemcpyHostToDevice));
  EDA_CUDA(cudaMemcpy(p->d_mel_hi, hi.data(), sizeof(int) * n_mels, cudaMemcpyHostToDevice));
  const int smem = (int)logmel_smem(n_fft, n_mels);
  EDA_CUDA(cudaFuncSetAttribute(logmel_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  EDA_CUDA(cudaFuncSetAttribute(logmel_kernel<double>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  *out = p;
  return EDA_OK;
}

int eda_frontend_plan_destroy(eda_frontend_plan* p) {
  if (!p) return EDA_OK;
  cudaFree(p->d_window); cudaFree(p->d_tw); cudaFree(p->d_melw); cudaFree(p->d_mel_lo); cudaFree(p->d_mel_hi);
  delete p;
  return EDA_OK;
}

int eda_frontend_logmel(const eda_frontend_plan* p, const void* audio, int audio_is_f64,
                        const int64_t* rec_sample_off, const int64_t* rec_frame_off, int n_rec,
                        int64_t total_frames, float* logmel, double* colsum, void* stream) {
  using namespace eda;
  EDA_REQUIRE(p && audio && logmel && colsum && n_rec > 0, "eda_frontend_logmel: bad arguments");
  cudaStream_t st = (cudaStream_t)stream;
  EDA_CUDA(cudaMemsetAsync(colsum, 0, sizeof(double) * (size_t)n_rec * p->n_mels, st));
  if (total_frames == 0) return EDA_OK;
  // contiguous runs of frames per warp (amortises the recording lookup and the atomics);
  // aim for >= 4 waves of 8-warp CTAs on the device
  const int64_t target_warps = (int64_t)num_sms() * 8 * 8;
  int fpw = (int)std::max<int64_t>(1, std::min<int64_t>(64, total_frames / target_warps));
  const int64_t n_warps = ceil_div<int64_t>(total_frames, fpw);
  const int blocks = (int)ceil_div<int64_t>(n_warps, 8);
  const size_t smem = logmel_smem(p->n_fft, p->n_mels);
  if (audio_is_f64)
    logmel_kernel<double><<<blocks, 256, smem, st>>>((const double*)audio, rec_sample_off, rec_frame_off, n_rec, total_frames,
        p->n_fft, p->frame_shift, p->n_mels, p->d_window, p->d_tw, p->d_melw, p->d_mel_lo, p->d_mel_hi, logmel, colsum, fpw);
  else
    logmel_kernel<float><<<blocks, 256, smem, st>>>((const float*)audio, rec_sample_off, rec_frame_off, n_rec, total_frames,
        p->n_fft, p->frame_shift, p->n_mels, p->d_window, p->d_tw, p->d_melw, p->d_mel_lo, p->d_mel_hi, logmel, colsum, fpw);
  EDA_LAUNCH_CHECK();
  return EDA_OK;
}

int eda_frontend_splice_subsample(const eda_frontend_plan* p, const float* logmel, const double* colsum,
                                  const int64_t* rec_frame_off, const int64_t* rec_row_off, int n_rec,
                                  int64_t total_rows, int context, int subsampling, float* feats,
                                  int ld_feats, void* stream) {
  using namespace eda;
  EDA_REQUIRE(p && logmel && colsum && feats && n_rec > 0 && subsampling > 0 && context >= 0, "eda_frontend_splice_subsample: bad arguments");
  EDA_REQUIRE(ld_feats >= (2 * context + 1) * p->n_mels, "eda_frontend_splice_subsample: ld_feats too small");
  if (total_rows == 0) return EDA_OK;
  const int blocks = (int)std::min<int64_t>(ceil_div<int64_t>(total_rows, 8), (int64_t)num_sms() * 16);
  splice_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(logmel, colsum, rec_frame_off, rec_row_off, n_rec, total_rows,
                                                           p->n_mels, context, subsampling, feats, ld_feats);
  EDA_LAUNCH_CHECK();
  return EDA_OK;
}

}  // extern "C"
