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
#include <stdlib.h>

#include <type_traits>
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
  double* d_melw64;    // packed non-zero weights (as double), filter m at [mel_off[m], mel_off[m+1])
  int* d_mel_off;      // (n_mels + 1)
  int mel_nnz;
  float* d_melw32;     // the same packed weights as float (what librosa stores), for logmel_t8_kernel
  double2* d_tw_step;  // n_fft 256: (16, 8) W_128^{t k1} at [k1 * 8 + t]; n_fft 512: (16, 16) W_256^{t k1} at [k1 * 16 + t]
                       // (twiddles between the two FFT steps of logmel_t8_kernel / logmel_t16_kernel)
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
__device__ __forceinline__ double audio_sample(const float* a, int64_t i) { return (double)__ldg(a + i); }
__device__ __forceinline__ double audio_sample(const double* a, int64_t i) { return __ldg(a + i); }
// 16-bit PCM, scaled the way soundfile hands it to the reference (x / 32768, exact in fp64)
__device__ __forceinline__ double audio_sample(const int16_t* a, int64_t i) { return (double)__ldg(a + i) * (1.0 / 32768.0); }
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
              float* __restrict__ logmel, double* __restrict__ colsum, int frames_per_warp, int mode,
              const int64_t* __restrict__ sum_range) {
  // mode (EDA_TRANSFORM_*): 0 = log10 mel power (+ column sums), 1 = |Y|, 2 = ln max(|Y|, 1e-10),
  // 3 = log10 max(1e-10, |Y| . mel^T); modes 1 / 2 write n_bins values per frame
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
      if (lane < n_mels && mode == 0) atomicAdd(colsum + (size_t)rec * n_mels + lane, csum);
      csum = 0.0;
      ++rec;
      r_f0 = rec_frame_off[rec]; r_f1 = rec_frame_off[rec + 1];
      r_s0 = rec_sample_off[rec]; r_n = rec_sample_off[rec + 1] - r_s0;
    }
    // sample index (within the recording) of padded position 0 of this frame
    const int64_t s_start = (f - r_f0) * hop - M;
    for (int n = lane; n < M; n += 32) {
      const int64_t i0 = s_start + 2 * n, i1 = i0 + 1;
      const double x0 = (i0 >= 0 && i0 < r_n) ? audio_sample(audio, r_s0 + i0) : 0.0;
      const double x1 = (i1 >= 0 && i1 < r_n) ? audio_sample(audio, r_s0 + i1) : 0.0;
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
      const double p2 = re * re + im * im;
      pw[k] = mode == 0 ? p2 : sqrt(p2);             // np.abs(Y) (feature.py:43) for the magnitude transforms
    }
    __syncwarp();
    if (mode == 1 || mode == 2) {                    // feature.py:44-47: |Y| / log(max(|Y|, 1e-10)), n_bins columns
      for (int k = lane; k <= M; k += 32)
        logmel[(size_t)f * n_bins + k] = (float)(mode == 1 ? pw[k] : log(fmax(pw[k], 1e-10)));
    } else if (lane < n_mels) {
      double acc = 0.0;
      const float* wrow = s_melw + (size_t)lane * n_bins;
      for (int k = my_lo; k <= my_hi; ++k) acc += pw[k] * (double)wrow[k];
      const double v = log10(fmax(acc, 1e-10));       // mode 3 = feature.py:105-116 on magnitudes, no mean removal
      logmel[(size_t)f * n_mels + lane] = (float)v;
      if (!sum_range || (f - r_f0 >= sum_range[2 * rec] && f - r_f0 < sum_range[2 * rec + 1])) csum += v;
    }
    __syncwarp();
  }
  if (lane < n_mels && mode == 0) atomicAdd(colsum + (size_t)rec * n_mels + lane, csum);
}

// ---- pass 1, fast path (n_fft = 256 or 512) ---------------------------------------------
// One warp per frame, the M = n_fft/2 packed complex points held in REGISTERS (R = M/32 per
// lane): a size-R DIF FFT inside each lane, one twiddle, then a 32-point DIF FFT ACROSS the
// lanes with xor-shuffles (no shared-memory round trips, no index arithmetic).  The spectrum
// goes through shared memory once for the real-FFT untangle, which produces the power of bins
// k and M-k together.  fp64 throughout, like the reference's numpy.
struct cd { double x, y; };
__device__ __forceinline__ cd cadd(cd a, cd b) { return {a.x + b.x, a.y + b.y}; }
__device__ __forceinline__ cd csub(cd a, cd b) { return {a.x - b.x, a.y - b.y}; }
__device__ __forceinline__ cd cmulw(cd a, cd w) { return {a.x * w.x - a.y * w.y, a.x * w.y + a.y * w.x}; }
__device__ __forceinline__ cd cmul_mi(cd a) { return {a.y, -a.x}; }          // a * (-i)


template <int R>
__device__ __forceinline__ void fft_local(cd (&a)[R]) {      // DIF, output slot j holds Y[bitrev(j)]
  constexpr double h = 0.70710678118654752440;
  if constexpr (R == 8) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const cd u = a[i], v = a[i + 4];
      a[i] = cadd(u, v);
      const cd d = csub(u, v);
      a[i + 4] = i == 0 ? d : i == 1 ? cd{(d.x + d.y) * h, (d.y - d.x) * h} : i == 2 ? cmul_mi(d)
                                                                              : cd{(d.y - d.x) * h, -(d.x + d.y) * h};
    }
  }
#pragma unroll
  for (int base = 0; base < R; base += 4) {
    const cd u0 = a[base], u1 = a[base + 1], v0 = a[base + 2], v1 = a[base + 3];
    const cd s0 = cadd(u0, v0), s1 = cadd(u1, v1), d0 = csub(u0, v0), d1 = cmul_mi(csub(u1, v1));
    a[base] = cadd(s0, s1); a[base + 1] = csub(s0, s1);
    a[base + 2] = cadd(d0, d1); a[base + 3] = csub(d0, d1);
  }
}

constexpr int kLogmelGroup = 4;      // frames whose mel projection + log10 run together
static size_t logmel_fast_smem(int M, int mel_nnz) {
  return (size_t)M * 16 + (size_t)((mel_nnz + 1) & ~1) * 8 + (size_t)8 * (M * 16 + kLogmelGroup * (M + 2) * 8);
}

template <int R, typename AudioT>
__global__ void __launch_bounds__(256, R == 4 ? 2 : 1)
logmel_fast_kernel(const AudioT* __restrict__ audio, const int64_t* __restrict__ rec_sample_off,
                   const int64_t* __restrict__ rec_frame_off, int n_rec, int64_t total_frames, int hop,
                   int n_mels, const double* __restrict__ window, const double2* __restrict__ tw,
                   const double* __restrict__ melw64, const int* __restrict__ mel_off,
                   const int* __restrict__ mel_lo, int mel_nnz, float* __restrict__ logmel,
                   double* __restrict__ colsum, int frames_per_warp, const int64_t* __restrict__ sum_range) {
  constexpr int M = 32 * R, NFFT = 2 * M, RB = R == 8 ? 3 : 2, FB = kLogmelGroup;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  // layout: [tw: M double2][melw: mel_nnz double (padded to even)][per-warp: M double2 + FB x (M+2) double]
  double2* s_tw = reinterpret_cast<double2*>(smem_raw);
  double* s_melw = reinterpret_cast<double*>(s_tw + M);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  unsigned char* wbase = smem_raw + (size_t)M * 16 + (size_t)((mel_nnz + 1) & ~1) * 8 + (size_t)warp * (M * 16 + FB * (M + 2) * 8);
  double2* zbuf = reinterpret_cast<double2*>(wbase);
  double* pwg = reinterpret_cast<double*>(wbase + M * 16);     // power spectra of the frame group
  for (int i = threadIdx.x; i < M; i += blockDim.x) s_tw[i] = tw[i];
  for (int i = threadIdx.x; i < mel_nnz; i += blockDim.x) s_melw[i] = melw64[i];
  __syncthreads();

  // per-lane constants
  auto wn = [&](int e) -> cd {                 // W_NFFT^e for 0 <= e < NFFT
    const double2 t = s_tw[e & (M - 1)];
    return e >= M ? cd{-t.x, -t.y} : cd{t.x, t.y};
  };
  double win0[R], win1[R];
#pragma unroll
  for (int r = 0; r < R; ++r) { win0[r] = window[2 * (lane + 32 * r)]; win1[r] = window[2 * (lane + 32 * r) + 1]; }
  cd ltw[R];                                   // W_M^{lane * bitrev(j)}
#pragma unroll
  for (int j = 0; j < R; ++j) ltw[j] = wn(2 * lane * (int)(__brev((unsigned)j) >> (32 - RB)));
  cd stw[4];                                   // cross-lane stages d = 16, 8, 4, 2: W_{2d}^{lane % d} on lower lanes
  double sgn[5];
#pragma unroll
  for (int s = 0; s < 5; ++s) {
    const int d = 16 >> s;
    const bool lower = (lane & d) != 0;
    sgn[s] = lower ? -1.0 : 1.0;
    if (s < 4) stw[s] = lower ? wn((lane % d) * (NFFT / (2 * d))) : cd{1.0, 0.0};
  }
  const int k1 = (int)(__brev((unsigned)lane) >> 27);
  const int my_off = lane < n_mels ? mel_off[lane] : 0, my_cnt = lane < n_mels ? mel_off[lane + 1] - my_off : 0;
  const int my_lo = lane < n_mels ? mel_lo[lane] : 0;

  const int64_t gwarp = (int64_t)blockIdx.x * (blockDim.x >> 5) + warp;
  const int64_t f_begin = gwarp * frames_per_warp;
  const int64_t f_end = min(total_frames, f_begin + frames_per_warp);
  if (f_begin >= f_end) return;
  int rec = 0;
  {
    int lo = 0, hi = n_rec - 1;
    while (lo < hi) {
      const int mid = (lo + hi + 1) >> 1;
      if (rec_frame_off[mid] <= f_begin) lo = mid; else hi = mid - 1;
    }
    rec = lo;
  }
  double csum = 0.0;
  int64_t r_f0 = rec_frame_off[rec], r_f1 = rec_frame_off[rec + 1];
  int64_t r_s0 = rec_sample_off[rec], r_n = rec_sample_off[rec + 1] - r_s0;
  // raw samples of one frame, packed as z[n] = x[2n] + i x[2n+1], n = lane + 32 r (center=True zero padding)
  double nx0[R], nx1[R];
  auto load_raw = [&](int64_t ff) {
    const int64_t s_start = (ff - r_f0) * hop - M;
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const int64_t i0 = s_start + 2 * (lane + 32 * r), i1 = i0 + 1;
      nx0[r] = (i0 >= 0 && i0 < r_n) ? audio_sample(audio, r_s0 + i0) : 0.0;
      nx1[r] = (i1 >= 0 && i1 < r_n) ? audio_sample(audio, r_s0 + i1) : 0.0;
    }
  };
  bool have_next = false;                     // (warp-uniform) nx* already hold the next frame
  int64_t f = f_begin;
  while (f < f_end) {
    while (f >= r_f1) {   // crossed into the next recording (empty recordings are skipped)
      if (lane < n_mels) atomicAdd(colsum + (size_t)rec * n_mels + lane, csum);
      csum = 0.0;
      ++rec;
      r_f0 = rec_frame_off[rec]; r_f1 = rec_frame_off[rec + 1];
      r_s0 = rec_sample_off[rec]; r_n = rec_sample_off[rec + 1] - r_s0;
      have_next = false;
    }
    // a group of up to FB frames of ONE recording: FFTs one after the other, mel + log together
    const int64_t f_lim = min(f_end, r_f1);
    const int nfb = (int)min((int64_t)FB, f_lim - f);
    for (int fb = 0; fb < nfb; ++fb) {
      const int64_t ff = f + fb;
      double* pw = pwg + fb * (M + 2);
      if (!have_next) load_raw(ff);
      cd a[R];
#pragma unroll
      for (int r = 0; r < R; ++r) a[r] = {nx0[r] * win0[r], nx1[r] * win1[r]};
      have_next = ff + 1 < f_lim;
      if (have_next) load_raw(ff + 1);          // in flight while this frame's FFT runs
      fft_local<R>(a);
#pragma unroll
      for (int j = 1; j < R; ++j) a[j] = cmulw(a[j], ltw[j]);
      // 32-point DIF across lanes: upper lane keeps own + partner, lower lane (partner - own) * w
#pragma unroll
      for (int s = 0; s < 5; ++s) {
        const int d = 16 >> s;
#pragma unroll
        for (int j = 0; j < R; ++j) {
          const double px = __shfl_xor_sync(0xffffffffu, a[j].x, d), py = __shfl_xor_sync(0xffffffffu, a[j].y, d);
          const cd t = {fma(a[j].x, sgn[s], px), fma(a[j].y, sgn[s], py)};
          a[j] = s < 4 ? cmulw(t, stw[s]) : t;
        }
      }
      // lane holds Z[R * bitrev5(lane) + bitrev(j)] in slot j: natural order through shared memory
      __syncwarp();                               // previous frame's untangle has finished reading zbuf
#pragma unroll
      for (int j = 0; j < R; ++j)
        zbuf[R * k1 + (int)(__brev((unsigned)j) >> (32 - RB))] = make_double2(a[j].x, a[j].y);
      __syncwarp();
      // real-FFT untangle: bins k and M-k from Z[k], Z[M-k]:  X[k] = (E + T)/2, X[M-k] = conj(E - T)/2,
      // E = Z[k] + conj(Z[M-k]),  T = W^k * (-i) (Z[k] - conj(Z[M-k]))
#pragma unroll
      for (int jj = 0; jj < R / 2; ++jj) {
        const int k = 1 + lane + 32 * jj;                // 1 .. M/2
        const double2 zk = zbuf[k], zm = zbuf[M - k];
        const cd E = {zk.x + zm.x, zk.y - zm.y};
        const cd D = {zk.x - zm.x, zk.y + zm.y};
        const double2 w = s_tw[k];
        const cd T = cmulw(cmul_mi(D), cd{w.x, w.y});
        const cd pp = cadd(E, T), q = csub(E, T);
        pw[k] = 0.25 * (pp.x * pp.x + pp.y * pp.y);
        pw[M - k] = 0.25 * (q.x * q.x + q.y * q.y);      // k == M/2 writes the same value twice
      }
      if (lane == 0) {
        const double2 z0 = zbuf[0];
        pw[0] = (z0.x + z0.y) * (z0.x + z0.y);
        pw[M] = (z0.x - z0.y) * (z0.x - z0.y);
      }
    }
    __syncwarp();
    if (lane < n_mels) {
      // sparse mel rows: one weight load feeds the FB frames' independent accumulation chains
      double acc[FB];
#pragma unroll
      for (int fb = 0; fb < FB; ++fb) acc[fb] = 0.0;
      const double* wrow = s_melw + my_off;
      const double* prow = pwg + my_lo;
      for (int i = 0; i < my_cnt; ++i) {
        const double w = wrow[i];
#pragma unroll
        for (int fb = 0; fb < FB; ++fb) acc[fb] = fma(prow[fb * (M + 2) + i], w, acc[fb]);   // rows >= nfb: stale, unused
      }
#pragma unroll
      for (int fb = 0; fb < FB; ++fb) {
        if (fb < nfb) {
          const double v = log10(fmax(acc[fb], 1e-10));
          logmel[(size_t)(f + fb) * n_mels + lane] = (float)v;
          if (!sum_range || (f + fb - r_f0 >= sum_range[2 * rec] && f + fb - r_f0 < sum_range[2 * rec + 1])) csum += v;
        }
      }
    }
    __syncwarp();
    f += nfb;
  }
  if (lane < n_mels) atomicAdd(colsum + (size_t)rec * n_mels + lane, csum);
}


// ---- pass 1, n_fft = 256 (8 kHz configs): eight threads per frame -----------------------------------------
// The previous kernel above (one warp per frame, 4 points per lane, five xor-shuffle stages) is bound by the
// L1 data pipe: every fp64 complex exchange costs four SHFL wavefronts.  Here a frame's M = 128 packed complex
// points live 16 per thread in 8 threads (a warp advances 4 frames at once):
//   step 1  thread t holds z[t + 8 j], j = 0..15  ->  16-point FFT in registers  ->  x W_128^{t k1}
//   transpose through shared memory, ONE pass of 2 KB per frame (STS.128 / LDS.128, xor-swizzled: conflict-free)
//   step 2  thread u holds column k1 = u and column 16 - u (thread 0: columns 0 and 8)  ->  two 8-point FFTs
//           = Z[k1 + 16 k2]
//   real-FFT untangle: the partner of bin u + 16 k2 is bin (16 - u) + 16 (7 - k2) - in the SAME thread, so the
//           bins k and M - k are finished without another exchange; powers leave as fp32
//   mel + log10 in fp32 (sum of <= 29 non-negative terms; the result is stored as fp32 anyway): thread t of the
//           frame's group owns filters t, t + 8, t + 16.
// All FFT / untangle arithmetic is fp64 like the reference's numpy.
struct cdx { double x, y; };
__device__ __forceinline__ cdx operator+(cdx a, cdx b) { return {a.x + b.x, a.y + b.y}; }
__device__ __forceinline__ cdx operator-(cdx a, cdx b) { return {a.x - b.x, a.y - b.y}; }
__device__ __forceinline__ cdx mul_mi(cdx a) { return {a.y, -a.x}; }                    // a * (-i)
__device__ __forceinline__ cdx mul_pi(cdx a) { return {-a.y, a.x}; }                    // a * (+i)
__device__ __forceinline__ cdx cmulx(cdx a, double wx, double wy) { return {a.x * wx - a.y * wy, a.x * wy + a.y * wx}; }

// 4-point DFT (W_4 = -i), in place, natural order
__device__ __forceinline__ void dft4(cdx& a0, cdx& a1, cdx& a2, cdx& a3) {
  const cdx s02 = a0 + a2, d02 = a0 - a2, s13 = a1 + a3, d13 = mul_mi(a1 - a3);
  a0 = s02 + s13; a2 = s02 - s13; a1 = d02 + d13; a3 = d02 - d13;
}

// 16-point DIF FFT in registers: slot 4 r + s ends up holding X[r + 4 s]
__device__ __forceinline__ void fft16(cdx (&a)[16]) {
  constexpr double c1 = 0.92387953251128675613, s1 = 0.38268343236508977173, h = 0.70710678118654752440;
#pragma unroll
  for (int n0 = 0; n0 < 4; ++n0) dft4(a[n0], a[n0 + 4], a[n0 + 8], a[n0 + 12]);        // slot n0 + 4 r = y_r[n0]
  // twiddles W_16^{n0 r}
  a[5] = cmulx(a[5], c1, -s1);                                      // W^1
  a[6] = cdx{(a[6].x + a[6].y) * h, (a[6].y - a[6].x) * h};         // W^2
  a[7] = cmulx(a[7], s1, -c1);                                      // W^3
  a[9] = cdx{(a[9].x + a[9].y) * h, (a[9].y - a[9].x) * h};         // W^2
  a[10] = mul_mi(a[10]);                                            // W^4
  a[11] = cdx{(a[11].y - a[11].x) * h, -(a[11].x + a[11].y) * h};   // W^6
  a[13] = cmulx(a[13], s1, -c1);                                    // W^3
  a[14] = cdx{(a[14].y - a[14].x) * h, -(a[14].x + a[14].y) * h};   // W^6
  a[15] = cmulx(a[15], -c1, s1);                                    // W^9
#pragma unroll
  for (int r = 0; r < 4; ++r) dft4(a[4 * r], a[4 * r + 1], a[4 * r + 2], a[4 * r + 3]);
}

// 8-point DFT, natural order in and out
__device__ __forceinline__ void fft8(cdx (&v)[8]) {
  constexpr double h = 0.70710678118654752440;
  cdx u0 = v[0] + v[4], u1 = v[1] + v[5], u2 = v[2] + v[6], u3 = v[3] + v[7];
  cdx d0 = v[0] - v[4], d1 = v[1] - v[5], d2 = mul_mi(v[2] - v[6]), d3 = v[3] - v[7];
  d1 = cdx{(d1.x + d1.y) * h, (d1.y - d1.x) * h};                   // W_8^1
  d3 = cdx{(d3.y - d3.x) * h, -(d3.x + d3.y) * h};                  // W_8^3
  dft4(u0, u1, u2, u3);
  dft4(d0, d1, d2, d3);
  v[0] = u0; v[2] = u1; v[4] = u2; v[6] = u3; v[1] = d0; v[3] = d1; v[5] = d2; v[7] = d3;
}

// Two neighbouring samples x[i0], x[i0 + 1] exactly as they lie in memory: the loads of the NEXT frame are issued
// while the current one is transformed, so nothing may touch the loaded registers before they are consumed
// (unpacking a 16-bit pair right after its load made every load a full memory round trip: 47 % of all stall samples).
template <typename AudioT, bool WORDS> struct RawPair;
template <> struct RawPair<int16_t, true> {             // every pair is one aligned 32-bit word (all recordings start on even samples, hop even)
  uint32_t v;
  __device__ __forceinline__ void load(const int16_t* rec, int64_t i0, int64_t n) {
    const int16_t* p = rec + i0;
    v = 0;
    if (i0 >= 0 && i0 + 1 < n) v = __ldg(reinterpret_cast<const uint32_t*>(p));
    else if (i0 >= 0 && i0 < n) v = __ldg(reinterpret_cast<const uint16_t*>(p));        // the last sample of an odd-length recording
  }
  // 16-bit PCM scaled the way soundfile hands it to the reference (x / 32768, exact in fp64)
  __device__ __forceinline__ double a() const { return (double)(int)(int16_t)(v & 0xffffu) * (1.0 / 32768.0); }
  __device__ __forceinline__ double b() const { return (double)((int)v >> 16) * (1.0 / 32768.0); }
};
template <> struct RawPair<int16_t, false> {            // general case: two 16-bit loads
  uint16_t x, y;
  __device__ __forceinline__ void load(const int16_t* rec, int64_t i0, int64_t n) {
    x = (i0 >= 0 && i0 < n) ? __ldg(reinterpret_cast<const uint16_t*>(rec + i0)) : (uint16_t)0;
    y = (i0 + 1 >= 0 && i0 + 1 < n) ? __ldg(reinterpret_cast<const uint16_t*>(rec + i0 + 1)) : (uint16_t)0;
  }
  __device__ __forceinline__ double a() const { return (double)(int)(int16_t)x * (1.0 / 32768.0); }
  __device__ __forceinline__ double b() const { return (double)(int)(int16_t)y * (1.0 / 32768.0); }
};
template <bool W> struct RawPair<float, W> {
  float x, y;
  __device__ __forceinline__ void load(const float* rec, int64_t i0, int64_t n) {
    x = (i0 >= 0 && i0 < n) ? __ldg(rec + i0) : 0.f;
    y = (i0 + 1 >= 0 && i0 + 1 < n) ? __ldg(rec + i0 + 1) : 0.f;
  }
  __device__ __forceinline__ double a() const { return (double)x; }
  __device__ __forceinline__ double b() const { return (double)y; }
};
template <bool W> struct RawPair<double, W> {
  double x, y;
  __device__ __forceinline__ void load(const double* rec, int64_t i0, int64_t n) {
    x = (i0 >= 0 && i0 < n) ? __ldg(rec + i0) : 0.0;
    y = (i0 + 1 >= 0 && i0 + 1 < n) ? __ldg(rec + i0 + 1) : 0.0;
  }
  __device__ __forceinline__ double a() const { return x; }
  __device__ __forceinline__ double b() const { return y; }
};

constexpr int kT8Warps = 4;                            // warps per CTA
static size_t logmel_t8_smem(int mel_nnz) {
  // [window 128 double2][step twiddles 128 double2][W_256^k 128 double2][mel weights float][per warp: 4 frames x (2 KB + 136 floats)]
  // (a frame's 2592 bytes = 20 x 128 + 32: the four frames of a warp sit 8 banks apart, so their STS.32 of the powers do not collide)
  return 3 * 2048 + (size_t)((mel_nnz + 31) & ~31) * 4 + (size_t)kT8Warps * 4 * (2048 + 136 * 4);
}

template <typename AudioT>
__global__ void __launch_bounds__(kT8Warps * 32, 3)
logmel_t8_kernel(const AudioT* __restrict__ audio, const int64_t* __restrict__ rec_sample_off,
                 const int64_t* __restrict__ rec_frame_off, int n_rec, int64_t total_frames, int hop, int n_mels,
                 const double* __restrict__ window, const double2* __restrict__ tw_step, const double2* __restrict__ tw256,
                 const float* __restrict__ melw32, const int* __restrict__ mel_off, const int* __restrict__ mel_lo,
                 int mel_nnz, float* __restrict__ logmel, double* __restrict__ colsum, int frames_per_warp,
                 const int64_t* __restrict__ sum_range) {
  constexpr int M = 128;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  double2* s_win = reinterpret_cast<double2*>(smem_raw);                 // (w[2n], w[2n+1])
  double2* s_step = s_win + 128;                                         // W_128^{t k1} at [k1 * 8 + t]
  double2* s_tw = s_step + 128;                                          // W_256^k
  float* s_melw = reinterpret_cast<float*>(s_tw + 128);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, q = lane >> 3, t = lane & 7;
  unsigned char* wbase = smem_raw + 3 * 2048 + (size_t)((mel_nnz + 31) & ~31) * 4 + (size_t)(warp * 4 + q) * (2048 + 136 * 4);
  double2* S = reinterpret_cast<double2*>(wbase);                        // this frame's transpose buffer [k1][t ^ (k1 & 7)]
  float* Pw = reinterpret_cast<float*>(wbase + 2048);                    // this frame's power spectrum (129 floats)
  for (int i = threadIdx.x; i < 128; i += blockDim.x) {
    s_win[i] = make_double2(window[2 * i], window[2 * i + 1]);
    s_step[i] = tw_step[i];
    s_tw[i] = tw256[i];
  }
  for (int i = threadIdx.x; i < mel_nnz; i += blockDim.x) s_melw[i] = melw32[i];
  __syncthreads();
  // are all pairs aligned 32-bit words?  (16-bit audio: every recording starts on an even sample and the hop is even)
  bool words = true;
  if constexpr (sizeof(AudioT) == 2) {
    bool ok = (hop & 1) == 0 && (reinterpret_cast<uintptr_t>(audio) & 3) == 0;
    for (int r = threadIdx.x; r < n_rec; r += blockDim.x) ok = ok && (rec_sample_off[r] & 1) == 0;
    words = __syncthreads_and(ok);
  }

  // this thread's mel filters (t, t + 8, t + 16) and untangle roles
  int m_off[3], m_cnt[3], m_lo[3];
#pragma unroll
  for (int r = 0; r < 3; ++r) {
    const int m = t + 8 * r;
    const bool ok = m < n_mels;
    m_off[r] = ok ? mel_off[m] : 0; m_cnt[r] = ok ? mel_off[m + 1] - m_off[r] : 0; m_lo[r] = ok ? mel_lo[m] : 0;
  }
  const int ka = t, kb = t == 0 ? 8 : 16 - t;                            // the two columns of step 2

  const int64_t gwarp = (int64_t)blockIdx.x * kT8Warps + warp;
  const int64_t f_begin = gwarp * frames_per_warp;
  const int64_t f_end = min(total_frames, f_begin + frames_per_warp);
  if (f_begin >= f_end) return;                                           // warp-uniform
  // recording of this group's first frame
  int64_t f = f_begin + q;
  int rec = 0;
  {
    int lo = 0, hi = n_rec - 1;
    const int64_t f0 = min(f, total_frames - 1);
    while (lo < hi) {
      const int mid = (lo + hi + 1) >> 1;
      if (rec_frame_off[mid] <= f0) lo = mid; else hi = mid - 1;
    }
    rec = lo;
  }
  int64_t r_f0 = rec_frame_off[rec], r_f1 = rec_frame_off[rec + 1];
  int64_t r_s0 = rec_sample_off[rec], r_n = rec_sample_off[rec + 1] - r_s0;
  double csum[3] = {0.0, 0.0, 0.0};
  int rec_acc = rec;                       // the recording csum belongs to

  auto flush = [&]() {
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      if (t + 8 * r < n_mels && csum[r] != 0.0) atomicAdd(colsum + (size_t)rec_acc * n_mels + t + 8 * r, csum[r]);
      csum[r] = 0.0;
    }
  };

  auto run = [&](auto words_tag) {
  constexpr bool WORDS = decltype(words_tag)::value;
  RawPair<AudioT, WORDS> raw[16] = {};
  auto load_raw = [&](int64_t ff) {        // z[n] = x[2n] + i x[2n+1], n = t + 8 j; center=True: padded position 0 = sample -M
    const int64_t s_start = (ff - r_f0) * hop - M + 2 * t;
    const AudioT* rec_audio = audio + r_s0;
#pragma unroll
    for (int j = 0; j < 16; ++j) raw[j].load(rec_audio, s_start + 16 * j, r_n);
  };
  // fetch(ff): advance this group's recording to the one holding frame ff and load its samples
  auto fetch = [&](int64_t ff) -> bool {
    if (ff >= f_end) return false;
    while (ff >= r_f1) {                                   // crossed into the next recording (empty ones are skipped)
      ++rec;
      r_f0 = rec_frame_off[rec]; r_f1 = rec_frame_off[rec + 1];
      r_s0 = rec_sample_off[rec]; r_n = rec_sample_off[rec + 1] - r_s0;
    }
    load_raw(ff);
    return true;
  };

  bool active_next = fetch(f_begin + q);
  for (int64_t fb = f_begin; fb < f_end; fb += 4) {       // warp-uniform loop; f = fb + q is this group's frame
    f = fb + q;
    const bool active = active_next;
    const int rec_cur = rec;                               // recording of the frame whose samples are in `raw`
    const int64_t fl_cur = f - r_f0;                       // its index within that recording
    cdx a[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      const double2 w = s_win[t + 8 * j];
      a[j] = active ? cdx{raw[j].a() * w.x, raw[j].b() * w.y} : cdx{0.0, 0.0};
    }
    active_next = fetch(f + 4);                            // the next frame's samples travel while this one is transformed
    fft16(a);
    // slot 4 r + s = Y_t[k1], k1 = r + 4 s;  x W_128^{t k1};  S[k1][t ^ (k1 & 7)]
    __syncwarp();                                          // the previous iteration's readers of S / Pw are done
#pragma unroll
    for (int r = 0; r < 4; ++r)
#pragma unroll
      for (int s2 = 0; s2 < 4; ++s2) {
        const int k1 = r + 4 * s2;
        cdx y = a[4 * r + s2];
        if (k1 > 0) { const double2 w = s_step[k1 * 8 + t]; y = cmulx(y, w.x, w.y); }
        S[k1 * 8 + (t ^ (k1 & 7))] = make_double2(y.x, y.y);
      }
    __syncwarp();
    cdx za[8], zb[8];
#pragma unroll
    for (int n2 = 0; n2 < 8; ++n2) {
      const double2 va = S[ka * 8 + (n2 ^ (ka & 7))], vb = S[kb * 8 + (n2 ^ (kb & 7))];
      za[n2] = cdx{va.x, va.y}; zb[n2] = cdx{vb.x, vb.y};
    }
    fft8(za);                                              // za[k2] = Z[ka + 16 k2]
    fft8(zb);                                              // zb[k2] = Z[kb + 16 k2]
    // untangle pairs (lower bin k, partner bin M - k).  Threads 1..7: (za[i], zb[7 - i]), k = t + 16 i.
    // Thread 0 (columns 0 and 8): (za[1], za[7]), (za[2], za[6]), (za[3], za[5]), (za[4], za[4]), (zb[i], zb[7 - i]) i = 0..3,
    // plus the purely real bins 0 and M from za[0].
    const bool t0 = t == 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      cdx zk, zm;
      int k;
      if (i < 4) {
        zk = t0 ? za[i + 1] : za[i];
        zm = t0 ? za[7 - i] : zb[7 - i];
        k = t0 ? 16 * (i + 1) : t + 16 * i;
      } else {
        zk = t0 ? zb[i - 4] : za[i];
        zm = t0 ? zb[11 - i] : zb[7 - i];
        k = t0 ? 8 + 16 * (i - 4) : t + 16 * i;
      }
      const cdx E = {zk.x + zm.x, zk.y - zm.y};            // Z[k] + conj(Z[M-k])
      const cdx D = {zk.x - zm.x, zk.y + zm.y};            // Z[k] - conj(Z[M-k])
      const double2 w = s_tw[k];
      const cdx T = cmulx(mul_mi(D), w.x, w.y);
      const cdx pp = E + T, qq = E - T;
      Pw[k] = (float)(0.25 * (pp.x * pp.x + pp.y * pp.y));
      Pw[M - k] = (float)(0.25 * (qq.x * qq.x + qq.y * qq.y));
    }
    if (t0) {
      const double sre = za[0].x + za[0].y, dre = za[0].x - za[0].y;
      Pw[0] = (float)(sre * sre);
      Pw[M] = (float)(dre * dre);
    }
    __syncwarp();
    // mel projection + log10 (fp32): filters t, t + 8, t + 16 of this group's frame
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      float acc = 0.f;
      const float* wrow = s_melw + m_off[r];
      const float* prow = Pw + m_lo[r];
      for (int i = 0; i < m_cnt[r]; ++i) acc = fmaf(prow[i], wrow[i], acc);
      if (active && t + 8 * r < n_mels) {
        if (r == 0 && rec_cur != rec_acc) { flush(); rec_acc = rec_cur; }
        const float v = log10f(fmaxf(acc, 1e-10f));
        logmel[(size_t)f * n_mels + t + 8 * r] = v;
        if (!sum_range || (fl_cur >= sum_range[2 * rec_cur] && fl_cur < sum_range[2 * rec_cur + 1])) csum[r] += (double)v;
      }
    }
  }
  flush();
  };                                                       // run
  if (words) run(std::true_type{}); else run(std::false_type{});
}


// ---- pass 1, n_fft = 512 (16 kHz configs): sixteen threads per frame ------------------------------------------
// Same scheme as logmel_t8_kernel with M = 256 = 16 x 16: thread t holds z[t + 16 j] -> 16-point FFT -> x W_256^{t k1} ->
// transpose through shared memory -> thread u holds column k1 = u -> 16-point FFT = Z[u + 16 k2].  The partner of bin
// u + 16 k2 is (16 - u) + 16 (15 - k2), which lives in ANOTHER thread here, so the spectrum goes through the (now free)
// transpose buffer once more in natural order and thread t untangles the pairs k = 1 + t + 16 i, M - k.  A warp advances
// two frames.  Replaces the one-warp-per-frame kernel for this size (1 CTA per SM at 8 points per lane: latency-bound).
constexpr int kT16Warps = 4;
constexpr int kT16Frame = 4096 + 264 * 4 + 32;         // bytes per frame: spectrum buffer + powers (+ 8 banks of skew)
static size_t logmel_t16_smem(int mel_nnz) {
  return 3 * 4096 + (size_t)((mel_nnz + 31) & ~31) * 4 + (size_t)kT16Warps * 2 * kT16Frame;
}

template <typename AudioT>
__global__ void __launch_bounds__(kT16Warps * 32, 3)
logmel_t16_kernel(const AudioT* __restrict__ audio, const int64_t* __restrict__ rec_sample_off,
                  const int64_t* __restrict__ rec_frame_off, int n_rec, int64_t total_frames, int hop, int n_mels,
                  const double* __restrict__ window, const double2* __restrict__ tw_step, const double2* __restrict__ tw512,
                  const float* __restrict__ melw32, const int* __restrict__ mel_off, const int* __restrict__ mel_lo,
                  int mel_nnz, float* __restrict__ logmel, double* __restrict__ colsum, int frames_per_warp,
                  const int64_t* __restrict__ sum_range) {
  constexpr int M = 256;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  double2* s_win = reinterpret_cast<double2*>(smem_raw);                 // (w[2n], w[2n+1]), n < 256
  double2* s_step = s_win + 256;                                         // W_256^{t k1} at [k1 * 16 + t]
  double2* s_tw = s_step + 256;                                          // W_512^k, k < 256
  float* s_melw = reinterpret_cast<float*>(s_tw + 256);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, q = lane >> 4, t = lane & 15;
  unsigned char* wbase = smem_raw + 3 * 4096 + (size_t)((mel_nnz + 31) & ~31) * 4 + (size_t)(warp * 2 + q) * kT16Frame;
  double2* S = reinterpret_cast<double2*>(wbase);                        // transpose buffer [k1][t ^ k1], then Z[k] in natural order
  float* Pw = reinterpret_cast<float*>(wbase + 4096);                    // power spectrum (257 floats)
  for (int i = threadIdx.x; i < 256; i += blockDim.x) {
    s_win[i] = make_double2(window[2 * i], window[2 * i + 1]);
    s_step[i] = tw_step[i];
    s_tw[i] = tw512[i];
  }
  for (int i = threadIdx.x; i < mel_nnz; i += blockDim.x) s_melw[i] = melw32[i];
  __syncthreads();
  bool words = true;
  if constexpr (sizeof(AudioT) == 2) {
    bool ok = (hop & 1) == 0 && (reinterpret_cast<uintptr_t>(audio) & 3) == 0;
    for (int r = threadIdx.x; r < n_rec; r += blockDim.x) ok = ok && (rec_sample_off[r] & 1) == 0;
    words = __syncthreads_and(ok);
  }
  int m_off[2], m_cnt[2], m_lo[2];                     // this thread's mel filters: t and t + 16
#pragma unroll
  for (int r = 0; r < 2; ++r) {
    const int m = t + 16 * r;
    const bool ok = m < n_mels;
    m_off[r] = ok ? mel_off[m] : 0; m_cnt[r] = ok ? mel_off[m + 1] - m_off[r] : 0; m_lo[r] = ok ? mel_lo[m] : 0;
  }
  const int64_t gwarp = (int64_t)blockIdx.x * kT16Warps + warp;
  const int64_t f_begin = gwarp * frames_per_warp;
  const int64_t f_end = min(total_frames, f_begin + frames_per_warp);
  if (f_begin >= f_end) return;                                           // warp-uniform
  int64_t f = f_begin + q;
  int rec = 0;
  {
    int lo = 0, hi = n_rec - 1;
    const int64_t f0 = min(f, total_frames - 1);
    while (lo < hi) {
      const int mid = (lo + hi + 1) >> 1;
      if (rec_frame_off[mid] <= f0) lo = mid; else hi = mid - 1;
    }
    rec = lo;
  }
  int64_t r_f0 = rec_frame_off[rec], r_f1 = rec_frame_off[rec + 1];
  int64_t r_s0 = rec_sample_off[rec], r_n = rec_sample_off[rec + 1] - r_s0;
  double csum[2] = {0.0, 0.0};
  int rec_acc = rec;
  auto flush = [&]() {
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      if (t + 16 * r < n_mels && csum[r] != 0.0) atomicAdd(colsum + (size_t)rec_acc * n_mels + t + 16 * r, csum[r]);
      csum[r] = 0.0;
    }
  };
  auto run = [&](auto words_tag) {
  constexpr bool WORDS = decltype(words_tag)::value;
  RawPair<AudioT, WORDS> raw[16] = {};
  auto load_raw = [&](int64_t ff) {        // z[n] = x[2n] + i x[2n+1], n = t + 16 j
    const int64_t s_start = (ff - r_f0) * hop - M + 2 * t;
    const AudioT* rec_audio = audio + r_s0;
#pragma unroll
    for (int j = 0; j < 16; ++j) raw[j].load(rec_audio, s_start + 32 * j, r_n);
  };
  auto fetch = [&](int64_t ff) -> bool {
    if (ff >= f_end) return false;
    while (ff >= r_f1) {
      ++rec;
      r_f0 = rec_frame_off[rec]; r_f1 = rec_frame_off[rec + 1];
      r_s0 = rec_sample_off[rec]; r_n = rec_sample_off[rec + 1] - r_s0;
    }
    load_raw(ff);
    return true;
  };
  bool active_next = fetch(f_begin + q);
  for (int64_t fb = f_begin; fb < f_end; fb += 2) {       // warp-uniform loop; f = fb + q is this half-warp's frame
    f = fb + q;
    const bool active = active_next;
    const int rec_cur = rec;
    const int64_t fl_cur = f - r_f0;
    cdx a[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      const double2 w = s_win[t + 16 * j];
      a[j] = active ? cdx{raw[j].a() * w.x, raw[j].b() * w.y} : cdx{0.0, 0.0};
    }
    active_next = fetch(f + 2);
    fft16(a);                                              // slot 4 r + s = Y_t[k1], k1 = r + 4 s
    __syncwarp();                                          // the previous iteration's readers of S / Pw are done
#pragma unroll
    for (int r = 0; r < 4; ++r)
#pragma unroll
      for (int s2 = 0; s2 < 4; ++s2) {
        const int k1 = r + 4 * s2;
        cdx y = a[4 * r + s2];
        if (k1 > 0) { const double2 w = s_step[k1 * 16 + t]; y = cmulx(y, w.x, w.y); }
        S[k1 * 16 + (t ^ k1)] = make_double2(y.x, y.y);
      }
    __syncwarp();
#pragma unroll
    for (int n2 = 0; n2 < 16; ++n2) {
      const double2 v = S[t * 16 + (n2 ^ t)];
      a[n2] = cdx{v.x, v.y};
    }
    fft16(a);                                              // slot 4 r + s = Z[t + 16 k2], k2 = r + 4 s
    __syncwarp();                                          // every lane has read its column: S is free
#pragma unroll
    for (int r = 0; r < 4; ++r)
#pragma unroll
      for (int s2 = 0; s2 < 4; ++s2) S[t + 16 * (r + 4 * s2)] = make_double2(a[4 * r + s2].x, a[4 * r + s2].y);
    __syncwarp();
    // real-FFT untangle of the pairs (k, M - k), k = 1 + t + 16 i (k = M/2 pairs with itself)
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int k = 1 + t + 16 * i;
      const double2 zk = S[k], zm = S[M - k];
      const cdx E = {zk.x + zm.x, zk.y - zm.y};
      const cdx D = {zk.x - zm.x, zk.y + zm.y};
      const double2 w = s_tw[k];
      const cdx T = cmulx(mul_mi(D), w.x, w.y);
      const cdx pp = E + T, qq = E - T;
      Pw[k] = (float)(0.25 * (pp.x * pp.x + pp.y * pp.y));
      Pw[M - k] = (float)(0.25 * (qq.x * qq.x + qq.y * qq.y));
    }
    if (t == 0) {
      const double2 z0 = S[0];
      Pw[0] = (float)((z0.x + z0.y) * (z0.x + z0.y));
      Pw[M] = (float)((z0.x - z0.y) * (z0.x - z0.y));
    }
    __syncwarp();
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      float acc = 0.f;
      const float* wrow = s_melw + m_off[r];
      const float* prow = Pw + m_lo[r];
      for (int i = 0; i < m_cnt[r]; ++i) acc = fmaf(prow[i], wrow[i], acc);
      if (active && t + 16 * r < n_mels) {
        if (r == 0 && rec_cur != rec_acc) { flush(); rec_acc = rec_cur; }
        const float v = log10f(fmaxf(acc, 1e-10f));
        logmel[(size_t)f * n_mels + t + 16 * r] = v;
        if (!sum_range || (fl_cur >= sum_range[2 * rec_cur] && fl_cur < sum_range[2 * rec_cur + 1])) csum[r] += (double)v;
      }
    }
  }
  flush();
  };                                                       // run
  if (words) run(std::true_type{}); else run(std::false_type{});
}

// ---- pass 2 ---------------------------------------------------------------------------
// Row t of the output = frames t*sub - c .. t*sub + c of the (mean-normalised) pass-1 matrix side by side: since that
// matrix is (frames, width) row-major, the (2c+1)*width values are ONE contiguous run of it (zeros where the run leaves
// the recording).  A warp copies a run with coalesced loads and stores; the per-column means of the recording (colsum /
// frames, fp64 like np.mean) sit in shared memory per CTA, which handles consecutive rows (almost always one recording).
// colsum == nullptr: no mean removal (transforms other than logmel23_mn).
__global__ void __launch_bounds__(256)
splice_kernel(const float* __restrict__ src, const double* __restrict__ colsum,
              const int64_t* __restrict__ rec_frame_off, const int64_t* __restrict__ rec_row_off,
              int n_rec, int64_t total_rows, int width, int context, int subsampling,
              float* __restrict__ feats, int ld_feats, int rows_per_cta) {
  extern __shared__ float2 s_mean[];                   // (width) column means of the recording of the CTA's first row, as hi + lo floats
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t row_begin = (int64_t)blockIdx.x * rows_per_cta;
  const int64_t row_end = min(total_rows, row_begin + rows_per_cta);
  if (row_begin >= row_end) return;
  int lo = 0, hi = n_rec - 1;
  while (lo < hi) {
    const int mid = (lo + hi + 1) >> 1;
    if (rec_row_off[mid] <= row_begin) lo = mid; else hi = mid - 1;
  }
  const int rec0 = lo;
  if (colsum) {
    const double inv_n0 = 1.0 / (double)max((int64_t)1, rec_frame_off[rec0 + 1] - rec_frame_off[rec0]);
    for (int m = threadIdx.x; m < width; m += blockDim.x) {
      const double mu = colsum[(size_t)rec0 * width + m] * inv_n0;       // np.mean in fp64
      const float hi_ = (float)mu;
      s_mean[m] = make_float2(hi_, (float)(mu - (double)hi_));
    }
  }
  __syncthreads();
  const int span = (2 * context + 1) * width;
  for (int64_t row = row_begin + warp; row < row_end; row += blockDim.x >> 5) {
    int rec = rec0;
    while (row >= rec_row_off[rec + 1]) ++rec;           // rows ascend: at most a few steps
    const int64_t f0 = rec_frame_off[rec], nfr = rec_frame_off[rec + 1] - f0;
    const int64_t t = (row - rec_row_off[rec]) * subsampling;
    const int64_t e0 = (t - context) * width, e1 = nfr * width;       // run = elements [e0, e0 + span) of this recording
    const float* base = src + f0 * width;
    float* dst = feats + row * (int64_t)ld_feats;
    const bool same = rec == rec0;
    const double inv_n = 1.0 / (double)max((int64_t)1, nfr);
    const int step = 32 % width;
    int m = lane % width;                                 // column of element e0 + c within its frame: (c mod width), e0 = 0 mod width
    // four independent loads in flight per lane before the first store (the kernel is latency-, not bandwidth-bound)
    for (int c0 = lane; c0 < ld_feats; c0 += 128) {
      float x[4];
      bool ok[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int c = c0 + 32 * u;
        const int64_t e = e0 + c;
        ok[u] = c < span && e >= 0 && e < e1;
        x[u] = ok[u] ? __ldg(base + e) : 0.f;
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int c = c0 + 32 * u;
        float v = x[u];
        if (ok[u] && colsum) {
          // x - mean with the mean as two floats: the difference is rounded to fp32 anyway and x already is an fp32
          // rounding of the fp64 log-mel value, so (x - hi) - lo loses nothing against converting to fp64 first
          // (the F2F.F64.F32 conversions were 26 % of this kernel's stall samples)
          if (same) { const float2 mu = s_mean[m]; v = (x[u] - mu.x) - mu.y; }
          else v = (float)((double)x[u] - colsum[(size_t)rec * width + m] * inv_n);
        }
        if (c < ld_feats) dst[c] = v;
        m += step; if (m >= width) m -= width;
      }
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
  std::vector<double> w64;
  std::vector<int> woff(n_mels + 1, 0);
  for (int m = 0; m < n_mels; ++m) {
    woff[m] = (int)w64.size();
    for (int k = lo[m]; k <= hi[m]; ++k) w64.push_back((double)melw[(size_t)m * p->n_bins + k]);
  }
  woff[n_mels] = (int)w64.size();
  p->mel_nnz = (int)w64.size();
  EDA_CUDA(cudaMalloc(&p->d_melw64, sizeof(double) * std::max<size_t>(1, w64.size())));
  EDA_CUDA(cudaMalloc(&p->d_mel_off, sizeof(int) * (n_mels + 1)));
  EDA_CUDA(cudaMemcpy(p->d_melw64, w64.data(), sizeof(double) * w64.size(), cudaMemcpyHostToDevice));
  EDA_CUDA(cudaMemcpy(p->d_mel_off, woff.data(), sizeof(int) * (n_mels + 1), cudaMemcpyHostToDevice));
  {
    std::vector<float> w32(w64.begin(), w64.end());                    // the float32 values librosa stores
    EDA_CUDA(cudaMalloc(&p->d_melw32, sizeof(float) * std::max<size_t>(1, w32.size())));
    EDA_CUDA(cudaMemcpy(p->d_melw32, w32.data(), sizeof(float) * w32.size(), cudaMemcpyHostToDevice));
    const int nt = n_fft == 512 ? 16 : 8, mm = 16 * nt;                 // threads per frame, packed complex points
    std::vector<double2> st(256);
    for (int k1 = 0; k1 < 16; ++k1)
      for (int t = 0; t < nt; ++t) {
        const double a = -2.0 * kPi * (double)((t * k1) % mm) / (double)mm;
        st[k1 * nt + t] = make_double2(cos(a), sin(a));
      }
    EDA_CUDA(cudaMalloc(&p->d_tw_step, sizeof(double2) * 256));
    EDA_CUDA(cudaMemcpy(p->d_tw_step, st.data(), sizeof(double2) * 256, cudaMemcpyHostToDevice));
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
  EDA_CUDA(cudaFuncSetAttribute(logmel_kernel<int16_t>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  if (n_fft == 256) {
    const int tsm = (int)logmel_t8_smem(p->mel_nnz);
    EDA_CUDA(cudaFuncSetAttribute(logmel_t8_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, tsm));
    EDA_CUDA(cudaFuncSetAttribute(logmel_t8_kernel<double>, cudaFuncAttributeMaxDynamicSharedMemorySize, tsm));
    EDA_CUDA(cudaFuncSetAttribute(logmel_t8_kernel<int16_t>, cudaFuncAttributeMaxDynamicSharedMemorySize, tsm));
  }
  if (n_fft == 512) {
    const int tsm = (int)logmel_t16_smem(p->mel_nnz);
    EDA_CUDA(cudaFuncSetAttribute(logmel_t16_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, tsm));
    EDA_CUDA(cudaFuncSetAttribute(logmel_t16_kernel<double>, cudaFuncAttributeMaxDynamicSharedMemorySize, tsm));
    EDA_CUDA(cudaFuncSetAttribute(logmel_t16_kernel<int16_t>, cudaFuncAttributeMaxDynamicSharedMemorySize, tsm));
  }
  if (n_fft == 256 || n_fft == 512) {
    const int M = n_fft / 2;
    const int fsm = (int)logmel_fast_smem(M, p->mel_nnz);
    EDA_CUDA(cudaFuncSetAttribute(logmel_fast_kernel<4, float>, cudaFuncAttributeMaxDynamicSharedMemorySize, fsm));
    EDA_CUDA(cudaFuncSetAttribute(logmel_fast_kernel<4, double>, cudaFuncAttributeMaxDynamicSharedMemorySize, fsm));
    EDA_CUDA(cudaFuncSetAttribute(logmel_fast_kernel<4, int16_t>, cudaFuncAttributeMaxDynamicSharedMemorySize, fsm));
    EDA_CUDA(cudaFuncSetAttribute(logmel_fast_kernel<8, float>, cudaFuncAttributeMaxDynamicSharedMemorySize, fsm));
    EDA_CUDA(cudaFuncSetAttribute(logmel_fast_kernel<8, double>, cudaFuncAttributeMaxDynamicSharedMemorySize, fsm));
    EDA_CUDA(cudaFuncSetAttribute(logmel_fast_kernel<8, int16_t>, cudaFuncAttributeMaxDynamicSharedMemorySize, fsm));
  }
  *out = p;
  return EDA_OK;
}

int eda_frontend_plan_destroy(eda_frontend_plan* p) {
  if (!p) return EDA_OK;
  cudaFree(p->d_window); cudaFree(p->d_tw); cudaFree(p->d_melw); cudaFree(p->d_mel_lo); cudaFree(p->d_mel_hi);
  cudaFree(p->d_melw64); cudaFree(p->d_mel_off); cudaFree(p->d_melw32); cudaFree(p->d_tw_step);
  delete p;
  return EDA_OK;
}

static int frontend_transform(const eda_frontend_plan* p, const void* audio, int audio_dtype,
                              const int64_t* rec_sample_off, const int64_t* rec_frame_off, int n_rec,
                              int64_t total_frames, int transform, float* out, double* colsum,
                              const int64_t* sum_range, void* stream) {
  using namespace eda;
  EDA_REQUIRE(p && audio && out && n_rec > 0, "eda_frontend_transform: bad arguments");
  EDA_REQUIRE(transform >= EDA_TRANSFORM_LOGMEL_MN && transform <= EDA_TRANSFORM_LOGMEL_ESPNET, "eda_frontend_transform: unknown transform %d", transform);
  EDA_REQUIRE(transform != EDA_TRANSFORM_LOGMEL_MN || colsum, "eda_frontend_transform: logmel23_mn needs colsum");
  cudaStream_t st = (cudaStream_t)stream;
  if (transform == EDA_TRANSFORM_LOGMEL_MN) EDA_CUDA(cudaMemsetAsync(colsum, 0, sizeof(double) * (size_t)n_rec * p->n_mels, st));
  if (total_frames == 0) return EDA_OK;
  EDA_REQUIRE(audio_dtype >= EDA_AUDIO_F32 && audio_dtype <= EDA_AUDIO_I16, "eda_frontend_transform: unknown audio dtype %d", audio_dtype);
  const char* force = getenv("EDA_B200_FRONTEND");         // "warp" = the one-warp-per-frame kernel (A/B measurements)
  if (transform == EDA_TRANSFORM_LOGMEL_MN && p->n_fft == 256 && p->n_mels <= 24 && !(force && force[0] == 'w')) {
    // eight threads per frame, four frames per warp step; contiguous runs of frames per warp, ~2 waves of 12 warps per SM
    const int64_t target_warps = (int64_t)num_sms() * 12 * 2;
    int64_t fpw = std::max<int64_t>(4, ceil_div<int64_t>(total_frames, target_warps));
    fpw = (fpw + 3) / 4 * 4;
    const int64_t n_warps = ceil_div<int64_t>(total_frames, fpw);
    const int blocks = (int)ceil_div<int64_t>(n_warps, kT8Warps);
    const size_t smem = logmel_t8_smem(p->mel_nnz);
#define EDA_T8(T)                                                                                                        \
  logmel_t8_kernel<T><<<blocks, kT8Warps * 32, smem, st>>>((const T*)audio, rec_sample_off, rec_frame_off, n_rec,         \
      total_frames, p->frame_shift, p->n_mels, p->d_window, p->d_tw_step, p->d_tw, p->d_melw32, p->d_mel_off, p->d_mel_lo, \
      p->mel_nnz, out, colsum, (int)fpw, sum_range)
    if (audio_dtype == EDA_AUDIO_F64) EDA_T8(double); else if (audio_dtype == EDA_AUDIO_I16) EDA_T8(int16_t); else EDA_T8(float);
#undef EDA_T8
    EDA_LAUNCH_CHECK();
    return EDA_OK;
  }
  if (transform == EDA_TRANSFORM_LOGMEL_MN && p->n_fft == 512 && p->n_mels <= 32 && !(force && force[0] == 'w')) {
    // sixteen threads per frame, two frames per warp step
    const int64_t target_warps = (int64_t)num_sms() * 12 * 2;
    int64_t fpw = std::max<int64_t>(2, ceil_div<int64_t>(total_frames, target_warps));
    fpw = (fpw + 1) / 2 * 2;
    const int64_t n_warps = ceil_div<int64_t>(total_frames, fpw);
    const int blocks = (int)ceil_div<int64_t>(n_warps, kT16Warps);
    const size_t smem = logmel_t16_smem(p->mel_nnz);
#define EDA_T16(T)                                                                                                         \
  logmel_t16_kernel<T><<<blocks, kT16Warps * 32, smem, st>>>((const T*)audio, rec_sample_off, rec_frame_off, n_rec,          \
      total_frames, p->frame_shift, p->n_mels, p->d_window, p->d_tw_step, p->d_tw, p->d_melw32, p->d_mel_off, p->d_mel_lo, \
      p->mel_nnz, out, colsum, (int)fpw, sum_range)
    if (audio_dtype == EDA_AUDIO_F64) EDA_T16(double); else if (audio_dtype == EDA_AUDIO_I16) EDA_T16(int16_t); else EDA_T16(float);
#undef EDA_T16
    EDA_LAUNCH_CHECK();
    return EDA_OK;
  }
  // contiguous runs of frames per warp (amortises the recording lookup and the atomics);
  // aim for >= 4 waves of 8-warp CTAs on the device
  const int64_t target_warps = (int64_t)num_sms() * 8 * 8;
  int fpw = (int)std::max<int64_t>(1, std::min<int64_t>(64, total_frames / target_warps));
  const int64_t n_warps = ceil_div<int64_t>(total_frames, fpw);
  const int blocks = (int)ceil_div<int64_t>(n_warps, 8);
  if (transform == EDA_TRANSFORM_LOGMEL_MN && (p->n_fft == 256 || p->n_fft == 512)) {
    // register-resident FFT, shuffles across lanes
    const int M = p->n_fft / 2;
    const size_t smem = logmel_fast_smem(M, p->mel_nnz);
#define EDA_FAST(R, T)                                                                                      \
  logmel_fast_kernel<R, T><<<blocks, 256, smem, st>>>((const T*)audio, rec_sample_off, rec_frame_off, n_rec, \
      total_frames, p->frame_shift, p->n_mels, p->d_window, p->d_tw, p->d_melw64, p->d_mel_off, p->d_mel_lo, \
      p->mel_nnz, out, colsum, fpw, sum_range)
    if (p->n_fft == 256) {
      if (audio_dtype == EDA_AUDIO_F64) EDA_FAST(4, double); else if (audio_dtype == EDA_AUDIO_I16) EDA_FAST(4, int16_t); else EDA_FAST(4, float);
    } else {
      if (audio_dtype == EDA_AUDIO_F64) EDA_FAST(8, double); else if (audio_dtype == EDA_AUDIO_I16) EDA_FAST(8, int16_t); else EDA_FAST(8, float);
    }
#undef EDA_FAST
    EDA_LAUNCH_CHECK();
    return EDA_OK;
  }
  const size_t smem = logmel_smem(p->n_fft, p->n_mels);
#define EDA_GEN(T)                                                                                                \
  logmel_kernel<T><<<blocks, 256, smem, st>>>((const T*)audio, rec_sample_off, rec_frame_off, n_rec, total_frames, \
      p->n_fft, p->frame_shift, p->n_mels, p->d_window, p->d_tw, p->d_melw, p->d_mel_lo, p->d_mel_hi, out, colsum, fpw, transform, sum_range)
  if (audio_dtype == EDA_AUDIO_F64) EDA_GEN(double); else if (audio_dtype == EDA_AUDIO_I16) EDA_GEN(int16_t); else EDA_GEN(float);
#undef EDA_GEN
  EDA_LAUNCH_CHECK();
  return EDA_OK;
}

int eda_frontend_transform(const eda_frontend_plan* p, const void* audio, int audio_dtype,
                           const int64_t* rec_sample_off, const int64_t* rec_frame_off, int n_rec,
                           int64_t total_frames, int transform, float* out, double* colsum, void* stream) {
  return frontend_transform(p, audio, audio_dtype, rec_sample_off, rec_frame_off, n_rec, total_frames, transform, out,
                            colsum, nullptr, stream);
}

int eda_frontend_logmel_ranges(const eda_frontend_plan* p, const void* audio, int audio_dtype,
                               const int64_t* rec_sample_off, const int64_t* rec_frame_off, const int64_t* sum_range,
                               int n_rec, int64_t total_frames, float* logmel, double* colsum, void* stream) {
  EDA_REQUIRE(colsum && sum_range, "eda_frontend_logmel_ranges: bad arguments");
  return frontend_transform(p, audio, audio_dtype, rec_sample_off, rec_frame_off, n_rec, total_frames,
                            EDA_TRANSFORM_LOGMEL_MN, logmel, colsum, sum_range, stream);
}

int eda_frontend_logmel(const eda_frontend_plan* p, const void* audio, int audio_dtype,
                        const int64_t* rec_sample_off, const int64_t* rec_frame_off, int n_rec,
                        int64_t total_frames, float* logmel, double* colsum, void* stream) {
  EDA_REQUIRE(colsum, "eda_frontend_logmel: bad arguments");
  return eda_frontend_transform(p, audio, audio_dtype, rec_sample_off, rec_frame_off, n_rec, total_frames,
                                EDA_TRANSFORM_LOGMEL_MN, logmel, colsum, stream);
}

int eda_frontend_width(const eda_frontend_plan* p, int transform) {
  if (!p) return 0;
  return (transform == EDA_TRANSFORM_NONE || transform == EDA_TRANSFORM_LOG) ? p->n_bins : p->n_mels;
}

int eda_frontend_splice_subsample_w(const float* src, int width, const double* colsum, const int64_t* rec_frame_off,
                                    const int64_t* rec_row_off, int n_rec, int64_t total_rows, int context,
                                    int subsampling, float* feats, int ld_feats, void* stream) {
  using namespace eda;
  EDA_REQUIRE(src && feats && n_rec > 0 && subsampling > 0 && context >= 0 && width > 0, "eda_frontend_splice_subsample: bad arguments");
  EDA_REQUIRE(ld_feats >= (2 * context + 1) * width, "eda_frontend_splice_subsample: ld_feats too small");
  EDA_REQUIRE(width <= 4096, "eda_frontend_splice_subsample: width too large");
  if (total_rows == 0) return EDA_OK;
  // consecutive rows per CTA (one recording's means in shared memory), ~4 CTAs per SM
  const int64_t want = ceil_div<int64_t>(total_rows, (int64_t)num_sms() * 16);
  const int rows_per_cta = (int)std::max<int64_t>(8, std::min<int64_t>(64, (want + 7) / 8 * 8));
  const int blocks = (int)ceil_div<int64_t>(total_rows, rows_per_cta);
  splice_kernel<<<blocks, 256, sizeof(float2) * width, (cudaStream_t)stream>>>(src, colsum, rec_frame_off, rec_row_off, n_rec,
                                                                              total_rows, width, context, subsampling, feats,
                                                                              ld_feats, rows_per_cta);
  EDA_LAUNCH_CHECK();
  return EDA_OK;
}

int eda_frontend_splice_subsample(const eda_frontend_plan* p, const float* logmel, const double* colsum,
                                  const int64_t* rec_frame_off, const int64_t* rec_row_off, int n_rec,
                                  int64_t total_rows, int context, int subsampling, float* feats,
                                  int ld_feats, void* stream) {
  EDA_REQUIRE(p && colsum, "eda_frontend_splice_subsample: bad arguments");
  return eda_frontend_splice_subsample_w(logmel, p->n_mels, colsum, rec_frame_off, rec_row_off, n_rec, total_rows, context,
                                         subsampling, feats, ld_feats, stream);
}

}  // extern "C"
