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
  double* d_melw64;    // packed non-zero weights (as double), filter m at [mel_off[m], mel_off[m+1])
  int* d_mel_off;      // (n_mels + 1)
  int mel_nnz;
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
                   double* __restrict__ colsum, int frames_per_warp) {
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
          csum += v;
        }
      }
    }
    __syncwarp();
    f += nfb;
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
  cudaFree(p->d_melw64); cudaFree(p->d_mel_off);
  delete p;
  return EDA_OK;
}

int eda_frontend_logmel(const eda_frontend_plan* p, const void* audio, int audio_dtype,
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
  EDA_REQUIRE(audio_dtype >= EDA_AUDIO_F32 && audio_dtype <= EDA_AUDIO_I16, "eda_frontend_logmel: unknown audio dtype %d", audio_dtype);
  if (p->n_fft == 256 || p->n_fft == 512) {
    // fast path: register-resident FFT, shuffles across lanes
    const int M = p->n_fft / 2;
    const size_t smem = logmel_fast_smem(M, p->mel_nnz);
#define EDA_FAST(R, T)                                                                                      \
  logmel_fast_kernel<R, T><<<blocks, 256, smem, st>>>((const T*)audio, rec_sample_off, rec_frame_off, n_rec, \
      total_frames, p->frame_shift, p->n_mels, p->d_window, p->d_tw, p->d_melw64, p->d_mel_off, p->d_mel_lo, \
      p->mel_nnz, logmel, colsum, fpw)
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
      p->n_fft, p->frame_shift, p->n_mels, p->d_window, p->d_tw, p->d_melw, p->d_mel_lo, p->d_mel_hi, logmel, colsum, fpw)
  if (audio_dtype == EDA_AUDIO_F64) EDA_GEN(double); else if (audio_dtype == EDA_AUDIO_I16) EDA_GEN(int16_t); else EDA_GEN(float);
#undef EDA_GEN
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
