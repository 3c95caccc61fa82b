// Frame-level speaker labels from segment lists (feature.py:228-245 in get_frame_labels, the same
// loop at :310-330 in get_labeledSTFT), optionally subsampled like feature.subsample (:122-127).
//
//   start_frame = rint(st * rate / frame_shift), end_frame = rint(et * rate / frame_shift)   (half to even)
//   rel_start = start_frame - start   if start <= start_frame <  end   else None
//   rel_end   = end_frame   - start   if start <  end_frame   <= end   else None
//   if rel_start is not None or rel_end is not None:  T[rel_start:rel_end, spk] = 1
// A segment that covers the whole chunk has BOTH None and is skipped - reference behaviour, kept.
// One warp per segment; overlapping segments write the same 1, so no ordering is needed.
#include "common.cuh"

namespace eda {

template <typename OutT>
__global__ void __launch_bounds__(256)
frame_labels_kernel(const double* __restrict__ seg_st, const double* __restrict__ seg_et,
                    const int32_t* __restrict__ seg_spk, const int32_t* __restrict__ seg_chunk, int64_t n_seg,
                    const int64_t* __restrict__ chunk_start, const int64_t* __restrict__ chunk_end,
                    const int64_t* __restrict__ chunk_frames, const int64_t* __restrict__ chunk_row_off,
                    int n_spk, double rate, double frame_shift, int subsampling, OutT* __restrict__ out, int ld) {
  const int lane = threadIdx.x & 31;
  for (int64_t s = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); s < n_seg;
       s += (int64_t)gridDim.x * (blockDim.x >> 5)) {
    const int spk = seg_spk[s];
    if (spk < 0 || spk >= n_spk) continue;
    const int c = seg_chunk[s];
    const int64_t start = chunk_start[c], end = chunk_end[c], n_frames = chunk_frames[c];
    const int64_t sf = (int64_t)rint(seg_st[s] * rate / frame_shift);
    const int64_t ef = (int64_t)rint(seg_et[s] * rate / frame_shift);
    const bool has_s = start <= sf && sf < end, has_e = start < ef && ef <= end;
    if (!has_s && !has_e) continue;
    // python slice semantics on an array of n_frames rows (indices here are never negative)
    const int64_t lo = has_s ? min(sf - start, n_frames) : 0;
    const int64_t hi = has_e ? min(ef - start, n_frames) : n_frames;
    // kept rows are frames 0, sub, 2 sub, ...: first multiple of sub that is >= lo
    const int64_t r0 = (lo + subsampling - 1) / subsampling;
    OutT* base = out + chunk_row_off[c] * ld + spk;
    for (int64_t r = r0 + lane; r * subsampling < hi; r += 32) base[r * ld] = (OutT)1;
  }
}

int frame_labels(const double* seg_st, const double* seg_et, const int32_t* seg_spk, const int32_t* seg_chunk,
                 int64_t n_seg, const int64_t* chunk_start, const int64_t* chunk_end, const int64_t* chunk_frames,
                 const int64_t* chunk_row_off, int n_chunks, int64_t total_rows, int n_spk, double rate,
                 int frame_shift, int subsampling, void* out, int ld, int out_f32, cudaStream_t st) {
  EDA_REQUIRE(chunk_start && chunk_end && chunk_frames && chunk_row_off && out && n_chunks > 0, "eda_frame_labels: bad arguments");
  EDA_REQUIRE(n_spk > 0 && ld >= n_spk && frame_shift > 0 && subsampling > 0 && rate > 0, "eda_frame_labels: bad sizes");
  EDA_REQUIRE(n_seg == 0 || (seg_st && seg_et && seg_spk && seg_chunk), "eda_frame_labels: segment arrays missing");
  EDA_CUDA(cudaMemsetAsync(out, 0, (size_t)total_rows * ld * 4, st));      // int32 and fp32 are both 4 bytes, 0 is 0
  if (n_seg == 0) return EDA_OK;
  const int blocks = (int)std::min<int64_t>(ceil_div<int64_t>(n_seg, 8), (int64_t)num_sms() * 8);
  if (out_f32)
    frame_labels_kernel<float><<<blocks, 256, 0, st>>>(seg_st, seg_et, seg_spk, seg_chunk, n_seg, chunk_start, chunk_end,
                                                       chunk_frames, chunk_row_off, n_spk, rate, (double)frame_shift,
                                                       subsampling, (float*)out, ld);
  else
    frame_labels_kernel<int32_t><<<blocks, 256, 0, st>>>(seg_st, seg_et, seg_spk, seg_chunk, n_seg, chunk_start, chunk_end,
                                                         chunk_frames, chunk_row_off, n_spk, rate, (double)frame_shift,
                                                         subsampling, (int32_t*)out, ld);
  EDA_LAUNCH_CHECK();
  return EDA_OK;
}

}  // namespace eda

extern "C" int eda_frame_labels(const double* seg_st, const double* seg_et, const int32_t* seg_spk,
                                const int32_t* seg_chunk, int64_t n_seg, const int64_t* chunk_start,
                                const int64_t* chunk_end, const int64_t* chunk_frames, const int64_t* chunk_row_off,
                                int n_chunks, int64_t total_rows, int n_spk, double rate, int frame_shift,
                                int subsampling, void* out, int ld, int out_f32, void* stream) {
  return eda::frame_labels(seg_st, seg_et, seg_spk, seg_chunk, n_seg, chunk_start, chunk_end, chunk_frames,
                           chunk_row_off, n_chunks, total_rows, n_spk, rate, frame_shift, subsampling, out, ld,
                           out_f32, (cudaStream_t)stream);
}
