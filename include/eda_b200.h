/*
 * eda_b200.h — C ABI of the B200-native EEND-EDA hot path (libeda_b200.so).
 *
 * This is the drop-in boundary.  The reference (shanguanma/speaker_diarization) has no
 * FFI layer: its hot path is PyTorch library calls inside
 *   speaker_diarization/feature.py                       (front end)
 *   speaker_diarization/eend_eda/models.py               (TransformerEdaModel)
 *   speaker_diarization/eend_eda/encoder_decoder_attractor.py
 * so every entry point below names the reference function (file:line) it replaces.  All
 * pointers are DEVICE pointers unless the name says `host`; all calls are asynchronous on
 * `stream` (a cudaStream_t passed as void*), allocate nothing unless documented, and return
 * 0 on success or a negative EDA_E* code (text via eda_last_error_string()).  No torch types
 * cross this boundary.  Row-major everywhere; "ld*" arguments are row strides in ELEMENTS.
 */
#ifndef EDA_B200_H_
#define EDA_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define EDA_ABI_VERSION 1

enum {
  EDA_OK = 0,
  EDA_EINVAL = -1,   /* bad argument (shape, alignment, unsupported size) */
  EDA_ECUDA = -2,    /* CUDA runtime / launch error                         */
  EDA_EUNSUPPORTED = -3
};

int eda_abi_version(void);
const char* eda_last_error_string(void);
/* Number of kernels this library has launched since load (bench.py's gpu_launches). */
int64_t eda_launch_count(void);

/* ------------------------------------------------------------------ front end ----
 * Replaces feature.stft (feature.py:155-184; librosa.stft with n_fft = next pow2 of
 * frame_size, periodic Hann, center=True zero padding, last frame dropped when
 * n % shift == 0), the 'logmel23_mn' branch of feature.transform (feature.py:64-73),
 * feature.splice (feature.py:130-152) and the [::subsampling] of feature.subsample
 * (feature.py:122-127) / infer_eda.py:98.
 *
 * A plan owns the device tables (window, twiddles, sparse Slaney mel filterbank).   */
typedef struct eda_frontend_plan eda_frontend_plan;

int eda_frontend_plan_create(int sample_rate, int frame_size, int frame_shift, int n_mels,
                             eda_frontend_plan** plan);            /* allocates tables */
int eda_frontend_plan_destroy(eda_frontend_plan* plan);
/* host helper: number of STFT frames feature.stft returns for n_samples (feature.py:186-193) */
int64_t eda_frontend_num_frames(int64_t n_samples, int frame_shift);

/* Pass 1: audio -> log10(max(|STFT|^2 . mel^T, 1e-10)) per frame, plus per-recording
 * per-bin sums (for the mean).  `audio` holds n_rec recordings back to back;
 * rec_sample_off / rec_frame_off are (n_rec+1) int64 prefix offsets (device).
 * logmel: (total_frames, n_mels) fp32.  colsum: (n_rec, n_mels) fp64, zeroed by the call.
 * audio_dtype: EDA_AUDIO_F32 / F64 (what soundfile hands the reference, kaldi_data.py:82) /
 * I16 (the wav file's own PCM samples; converted as x / 32768 exactly like soundfile does). */
enum { EDA_AUDIO_F32 = 0, EDA_AUDIO_F64 = 1, EDA_AUDIO_I16 = 2 };
int eda_frontend_logmel(const eda_frontend_plan* plan, const void* audio, int audio_dtype,
                        const int64_t* rec_sample_off, const int64_t* rec_frame_off,
                        int n_rec, int64_t total_frames,
                        float* logmel, double* colsum, void* stream);

/* The other feature.transform variants that run under the pinned librosa 0.10.2 (feature.py:43-47, 105-116; the
 * variants that call librosa.filters.mel positionally raise TypeError there and are not reproduced):
 *   EDA_TRANSFORM_LOGMEL_MN      'logmel23_mn'      = eda_frontend_logmel (out (frames, n_mels), colsum required)
 *   EDA_TRANSFORM_NONE           None               |Y|                               out (frames, n_fft/2+1)
 *   EDA_TRANSFORM_LOG            'log'              ln(max(|Y|, 1e-10))               out (frames, n_fft/2+1)
 *   EDA_TRANSFORM_LOGMEL_ESPNET  'logmel23_espnet'  log10(max(1e-10, |Y| . mel^T))    out (frames, n_mels)
 * colsum may be NULL except for LOGMEL_MN (the only variant with a mean).  eda_frontend_width = columns of `out`. */
enum { EDA_TRANSFORM_LOGMEL_MN = 0, EDA_TRANSFORM_NONE = 1, EDA_TRANSFORM_LOG = 2, EDA_TRANSFORM_LOGMEL_ESPNET = 3 };
int eda_frontend_transform(const eda_frontend_plan* plan, const void* audio, int audio_dtype,
                           const int64_t* rec_sample_off, const int64_t* rec_frame_off,
                           int n_rec, int64_t total_frames, int transform,
                           float* out, double* colsum, void* stream);
int eda_frontend_width(const eda_frontend_plan* plan, int transform);

/* eda_frontend_logmel for a SLICE of a recording (chunk-level sharding of one long recording over several GPUs,
 * SURVEY 8e): the log-mel rows of all frames of every "recording" handed in are written, but only frames
 * [sum_range[2r], sum_range[2r+1]) (indices within recording r) enter colsum - the slice's halo frames, which another
 * rank owns, do not.  The ranks then add their colsum and frame counts (one 24-double all-reduce) to obtain the
 * recording-wide mean feature.py:72-73 takes.                                                          */
int eda_frontend_logmel_ranges(const eda_frontend_plan* plan, const void* audio, int audio_dtype,
                               const int64_t* rec_sample_off, const int64_t* rec_frame_off,
                               const int64_t* sum_range, int n_rec, int64_t total_frames,
                               float* logmel, double* colsum, void* stream);

/* Pass 2: subtract the recording mean, splice +-context frames (zeros beyond the edges,
 * of the MEAN-NORMALISED signal, exactly like np.pad after transform) and keep every
 * `subsampling`-th row.  feats: (total_rows, ld_feats) fp32 with (2c+1)*n_mels valid
 * columns; columns beyond that up to ld_feats are written as 0.
 * rec_row_off: (n_rec+1) int64 prefix offsets of output rows (ceil(frames/subsampling)). */
int eda_frontend_splice_subsample(const eda_frontend_plan* plan, const float* logmel,
                                  const double* colsum, const int64_t* rec_frame_off,
                                  const int64_t* rec_row_off, int n_rec, int64_t total_rows,
                                  int context, int subsampling,
                                  float* feats, int ld_feats, void* stream);

/* Pass 2 for any pass-1 width (feature.splice + [::subsampling] after any transform): src (total_frames, width);
 * colsum (n_rec, width) fp64 or NULL = no mean removal.                                          */
int eda_frontend_splice_subsample_w(const float* src, int width, const double* colsum,
                                    const int64_t* rec_frame_off, const int64_t* rec_row_off, int n_rec,
                                    int64_t total_rows, int context, int subsampling,
                                    float* feats, int ld_feats, void* stream);

/* ------------------------------------------------------------------- encoder -----
 * Building blocks of TransformerEdaModel.forward_embedding (models.py:213-234).     */

/* HOST function (no device work): the per-sequence torch.randperm(ilen) draws of models.py:229-232 for a whole batch.
 * left / next / state624 are the fields of torch's serialised CPU generator (32-bit Mersenne Twister); they are advanced
 * exactly as n_seq torch.randperm calls would advance them.  rows[b]: host memory for the permutation of
 * 0..lens[b]-1 of sequence b (sequences are drawn in the order b = 0, 1, ...).                  */
int eda_host_randperm_batch(int32_t* left, uint64_t* next, uint64_t* state624, const int32_t* lens, int n_seq,
                            int32_t* const* rows);

/* pad_sequence(padding_value) of models.py:217 into a K-padded staging matrix:
 * dst (B*Tmax, ld_dst); row (b,t) = src row (seq_off[b]+t) for t < len[b] else `pad`;
 * columns >= n_cols are 0.  src is the concatenation of the sequences, (sum len, n_cols). */
int eda_pad_pack(const float* src, const int64_t* seq_off, const int32_t* lens, int B, int Tmax,
                 int n_cols, float pad, float* dst, int ld_dst, void* stream);

enum {
  EDA_EPI_BIAS = 0,        /* C = A.W^T + bias                                   */
  EDA_EPI_BIAS_RELU = 1,   /* C = relu(A.W^T + bias)          (linear1, models.py:193) */
  EDA_EPI_BIAS_LN = 2,     /* C = LN(A.W^T + bias)            (encoder+encoder_norm, models.py:220-221) */
  EDA_EPI_BIAS_RES_LN = 3, /* C = LN(resid + A.W^T + bias)    (post-LN residual blocks) */
  /* 4 is internal (the QKV scatter of eda_gemm_qkv) */
  EDA_EPI_BIAS_RES = 5,    /* C = resid + A.W^T + bias        (backward: gradient joins of the residual stream) */
  EDA_EPI_RELU_MASK = 6,   /* C = (A.W^T) where mask > 0 else 0   (backward of linear1's ReLU)   */
  EDA_EPI_BIAS_SILU = 7,   /* C = silu(A.W^T + bias)          (Conformer feed-forward, torchaudio conformer.py) */
  EDA_EPI_BIAS_GLU = 8     /* C (M, N/2) = a * sigmoid(g)     (Conformer conv module; weight rows interleaved in
                              32-row blocks a0..31, g0..31, a32..63, ... by the caller)            */
};
enum { EDA_PREC_FP32 = 0, EDA_PREC_BF16X3 = 1, EDA_PREC_BF16 = 2 };

/* nn.Linear (+ fused epilogue).  A (M,K) lda, W (N,K) ldw (PyTorch weight layout),
 * C (M,N) ldc, all fp32 in HBM.  LN epilogues need N == 256; eps = 1e-5.
 * prec selects the arithmetic: FP32 = CUDA-core FFMA; BF16X3 = tcgen05 with error-
 * compensated bf16 split operands (3 MMAs, fp32 TMEM accumulation); BF16 = one MMA.  */
int eda_linear(const float* A, int lda, const float* W, int ldw, const float* bias,
               const float* resid, int ldr, const float* ln_gamma, const float* ln_beta,
               float* C, int ldc, int M, int N, int K, int epilogue, int prec, void* stream);

/* Multi-head self-attention core of nn.TransformerEncoderLayer (no masks: padded frames
 * are attended to, models.py:223).  qkv (B*T, 3*D) = [q | k | v], head h owns columns
 * h*dh..; out (B*T, D) = softmax(q k^T / sqrt(dh)) v per (b, h).  dh must be 64.      */
int eda_mha_fwd(const float* qkv, int ld_qkv, float* out, int ld_out, int B, int T,
                int n_heads, int head_dim, int prec, void* stream);

/* eda_mha_fwd for a batch of INDEPENDENT sequences: keys t >= key_lens[b] take no part (rows of padded
 * queries are computed but meaningless).  With it a padded batch of chunks gives every chunk the result of
 * the reference's own loop, which runs one chunk per model call so that no padding ever exists
 * (infer_eda.py:102-113; models.py:223 passes no mask because B = 1 there).  prec must be EDA_PREC_FP32;
 * the tensor-core pipeline's form is eda_mha_split_masked.                                      */
int eda_mha_fwd_masked(const float* qkv, int ld_qkv, float* out, int ld_out, int B, int T,
                       int n_heads, int head_dim, const int32_t* key_lens, int prec, void* stream);

/* ---------------------------------------------------- split-bf16 tensor-core pipeline --
 * The tensor-core precisions keep activations in HBM as "split" bf16 pairs (x = hi + lo,
 * hi = bf16(x), lo = bf16(x - hi); two bf16 matrices with the same row stride) so that a
 * producing kernel's epilogue feeds the next kernel without a conversion pass.  These entry
 * points are what TransformerEdaModel.forward_embedding (models.py:213-226) runs on for
 * prec = EDA_PREC_BF16X3 / EDA_PREC_BF16; eda_linear / eda_mha_fwd above are the fp32-in,
 * fp32-out forms of the same kernels.  hi/lo pointers are bf16 device pointers (void*).     */

/* fp32 (rows, K) row stride ld -> hi, lo (rows, ldk) with columns [K, ldk) zeroed. ldk % 8 == 0. */
int eda_split_bf16(const float* src, int ld, int rows, int K, void* hi, void* lo, int ldk, void* stream);

/* eda_pad_pack (models.py:217) writing the split layout directly. src rows have stride ld_src. */
int eda_pad_pack_split(const float* src, int ld_src, const int64_t* seq_off, const int32_t* lens, int B,
                       int Tmax, int n_cols, float pad, void* dst_hi, void* dst_lo, int ld_dst, void* stream);

/* nn.Linear + fused epilogue (EDA_EPI_*) on split operands: A (M,K) stride lda, W (N,K) stride
 * ldw.  Outputs: fp32 C (may be NULL) and/or split c_hi/c_lo with stride ldcs (may be NULL). */
int eda_gemm_split(const void* a_hi, const void* a_lo, int lda, const void* w_hi, const void* w_lo, int ldw,
                   const float* bias, const float* resid, int ldr, const float* ln_gamma, const float* ln_beta,
                   float* C, int ldc, void* c_hi, void* c_lo, int ldcs, int M, int N, int K, int epilogue,
                   int prec, void* stream);

/* The packed in_proj of nn.MultiheadAttention (in_proj_weight (3D, D)): rows are (b, t), t < T.
 * q (pre-scaled by 1/sqrt(head_dim)) and k go to split qk (B*T, ld_qk >= 2D) as [q | k];
 * v goes TRANSPOSED per (b, head) to split vt ((B*n_heads*64), Tp): vt[(b*H + h)*64 + d][t],
 * Tp >= T, Tp % 8 == 0.  D must be 256.                                                       */
int eda_gemm_qkv(const void* a_hi, const void* a_lo, int lda, const void* w_hi, const void* w_lo, int ldw,
                 const float* bias, void* qk_hi, void* qk_lo, int ld_qk, void* vt_hi, void* vt_lo,
                 int B, int T, int Tp, int D, int n_heads, int prec, void* stream);

/* eda_mha_fwd on the layouts eda_gemm_qkv writes.  Outputs (B*T, n_heads*64): fp32 `out`
 * (may be NULL) and/or split out_hi/out_lo (may be NULL).                                     */
int eda_mha_split(const void* qk_hi, const void* qk_lo, int ld_qk, const void* vt_hi, const void* vt_lo,
                  int Tp, int B, int T, int n_heads, float* out, int ld_out, void* out_hi, void* out_lo,
                  int ld_os, int prec, void* stream);

/* --------------------------------------------------------------------- EDA -------
 * LstmEncoderDedecoderAttractor.forward (encoder_decoder_attractor.py:19-59) on the
 * frames shuffled by models.py:228-232.
 *
 * xproj (B*Tmax, 4*D) = emb . W_ih^T + b_ih + b_hh (an eda_linear call; the shuffle is
 * applied as an index: step s of sequence b reads row b*Tmax + perm[b*Tmax + s]).
 * The encoder LSTM stops at lens[b] (packed-sequence semantics); its (h, c) seed the
 * decoder LSTM, stepped n_dec times on zero input (gates = W_hh h + dec_bias).
 * w_hh_* are (4*D, D) fp32 in PyTorch gate order i,f,g,o; dec_bias = b_ih + b_hh (4*D).
 * attractors: (B, n_dec, D) fp32.  hc_out (optional, may be NULL): (B, 2, D) final
 * encoder (h, c).  `order` (optional, may be NULL = identity): a permutation of 0..B-1;
 * sequences are co-scheduled in groups in that order (pass them sorted by length so a
 * group finishes together).  D must be 256.
 * prec: EDA_PREC_FP32 = the recurrent mat-vec W_hh.h in fp32 on the CUDA cores; EDA_PREC_BF16X3 /
 * EDA_PREC_BF16 = on the tensor cores as split-bf16 products (W_hi.h_hi + W_lo.h_hi + W_hi.h_lo,
 * fp32 accumulate; both keep the three products: the recurrence is not where bf16 would pay). */
int eda_lstm_attractors(const float* xproj, const int32_t* perm, const int32_t* lens,
                        const int32_t* order, const float* w_hh_enc, const float* w_hh_dec, const float* dec_bias,
                        int B, int Tmax, int D, int n_dec,
                        float* attractors, float* hc_out, int prec, void* stream);

/* Number of 8-CTA clusters of the recurrence kernel the device keeps resident at once; each
 * advances 4 sequences, so batches above 4x this value take more than one wave.            */
int eda_lstm_max_active_clusters(void);

/* Existence head: probs[b,n] = sigmoid(attractors[b,n,:] . w + bias)
 * (encoder_decoder_attractor.py:50-54) and, when n_spk_label >= 0, the attractor loss of
 * models.py:350-371 (BCE-with-logits applied to the probabilities against
 * [1]*n_spk_label + [0]) into loss_att[0].                                              */
int eda_attractor_head(const float* attractors, const float* w, const float* bias,
                       int B, int n_dec, int D, float* probs, int n_spk_label,
                       float* loss_att, void* stream);

/* ---------------------------------------------------------- posterior + threshold --
 * pred = bmm(emb, attractors[:, :n_att]^T) (models.py:254, :329), sigmoid (models.py:333)
 * and the activity decision p > threshold (bin/make_rttm.py:29), fused.
 * emb (B*Tmax, ld_emb); attractors (B, att_stride_rows, D) of which the first n_att rows
 * are used.  Any of logits / post (fp32 (B*Tmax, n_att)) / labels (uint8, same shape) may
 * be NULL.  n_att <= 16.                                                                */
int eda_posterior(const float* emb, int ld_emb, const float* attractors, int att_stride_rows,
                  int B, int Tmax, int D, int n_att, float threshold,
                  float* logits, float* post, uint8_t* labels, void* stream);

/* eda_posterior writing the time-axis concatenation the reference's driver builds with np.vstack(out_chunks)
 * (infer_eda.py:113,121): frame t < lens[b] of sequence b lands in row out_off[b] + t of logits / post / labels
 * ((total rows, n_att) matrices); rows of padded frames are not written.  out_off: (B) int64, lens: (B) int32.
 * mask16 (optional, (total rows) uint16): the frame's thresholded decisions as a bit mask, bit s = attractor s active
 * (bin/make_rttm.py:29 on all columns at once) - 2 bytes per frame, the form in which results are cheapest to collect
 * from the other ranks.                                                                              */
int eda_posterior_packed(const float* emb, int ld_emb, const float* attractors, int att_stride_rows,
                         int B, int Tmax, int D, int n_att, float threshold, const int64_t* out_off,
                         const int32_t* lens, float* logits, float* post, uint8_t* labels, uint16_t* mask16,
                         void* stream);

/* ------------------------------------------------------------------- PIT loss ------
 * TransformerEdaModel.pit_loss (models.py:373-421).  pred, label (B, T, S) fp32; lens (B).
 * Outputs: loss[0] = sum_b min_p cost[b,p] / sum_b lens[b]; min_idx (B) int64 = index
 * into itertools.permutations(range(S)) with the lowest index on ties;
 * label_perm (B, T, S) fp32 = label[b, :, perm[min_idx[b]]]; perm_out (B, S) int32
 * (optional).  S <= 8.  Deterministic (fixed summation order).                           */
int eda_pit_loss(const float* pred, const float* label, const int32_t* lens, int B, int T, int S,
                 float* loss, int64_t* min_idx, float* label_perm, int32_t* perm_out,
                 void* stream);
/* eda_pit_loss for callers that must not allocate (code under CUDA graph capture, serving loops,
 * harnesses that forbid cudaMalloc).  No counterpart in the reference.
 * Same arguments, same kernel and bit-identical results; the per-sequence cost
 * scratch and the completion counter live in a file-scope static device array of
 * 8 * EDA_PIT_STATIC_MAX_B + 256 bytes that every GPU loading the library holds until exit.  The call
 * makes no cudaMalloc / cudaFree (eda_pit_loss grows a scratch buffer on demand) and returns
 * EDA_EINVAL for B > EDA_PIT_STATIC_MAX_B.  All calls on a device share the array: calls that overlap
 * on the GPU (different streams without an ordering between them) race; same-stream calls are safe. */
#define EDA_PIT_STATIC_MAX_B 4096
int eda_pit_loss_static(const float* pred, const float* label, const int32_t* lens, int B, int T, int S,
                        float* loss, int64_t* min_idx, float* label_perm, int32_t* perm_out,
                        void* stream);
/* d loss / d pred = mask * (sigmoid(pred) - label_perm) / (S * sum lens) * grad_scale. */
int eda_pit_loss_bwd(const float* pred, const float* label_perm, const int32_t* lens,
                     int B, int T, int S, float grad_scale, float* dpred, void* stream);

/* TransformerEdaModel.calc_diarization_error (models.py:423-463).  counts (9) int64,
 * zeroed by the call: [n_equal (= correct * S), num_frames, speech_scored, speech_miss,
 * speech_falarm, speaker_scored, speaker_miss, speaker_falarm, speaker_error].          */
int eda_der_counters(const float* pred, const float* label_perm, const int32_t* lens,
                     int B, int T, int S, int64_t* counts, void* stream);

/* Frame-level speaker labels from segment lists: the loop of feature.get_frame_labels
 * (feature.py:228-245; identical at :310-330 in get_labeledSTFT) for a batch of chunks, with
 * the row subsampling of feature.subsample (:122-127) folded in.  Segment s (seconds
 * seg_st/seg_et, speaker column seg_spk, chunk seg_chunk) marks frames
 * rint(st*rate/shift) .. rint(et*rate/shift) of chunk [chunk_start, chunk_end) using the
 * reference's clipping rule (a segment covering the whole chunk is dropped, like there).
 * chunk_frames = rows of the chunk's label matrix before subsampling; its kept rows
 * (frames 0, sub, 2 sub, ...) start at row chunk_row_off[c] of out (total_rows x ld, int32
 * or fp32 by out_f32; zeroed by the call).                                                */
int eda_frame_labels(const double* seg_st, const double* seg_et, const int32_t* seg_spk,
                     const int32_t* seg_chunk, int64_t n_seg, const int64_t* chunk_start,
                     const int64_t* chunk_end, const int64_t* chunk_frames,
                     const int64_t* chunk_row_off, int n_chunks, int64_t total_rows, int n_spk,
                     double rate, int frame_shift, int subsampling, void* out, int ld, int out_f32,
                     void* stream);

/* Post-processing to speaker segments (bin/make_rttm.py:29-35): a = T_hat > threshold (:29),
 * scipy.signal.medfilt(a, (median, 1)) when median > 1 (:30-31; zero padded at the
 * recording's ends, median odd), then per speaker the pairs of change points of the
 * zero-padded column (:32-35).  Recordings are rows [rec_off[r], rec_off[r+1]) of ONE
 * matrix with row stride ld: pass posteriors (post, fp32) or ready 0/1 decisions (labels,
 * uint8), the other NULL.  n_spk_rec (optional, per recording): columns >= it yield nothing.
 * segments (n_rec, n_spk, cap, 2) int32 = [start, end) frames in time order; counts
 * (n_rec, n_spk) int32 = number of segments found (a value > cap means the tail was cut). */
int eda_rttm_segments(const float* post, const uint8_t* labels, int ld, const int64_t* rec_off,
                      int n_rec, const int32_t* n_spk_rec, int n_spk, float threshold, int median,
                      int cap, int32_t* segments, int32_t* counts, void* stream);

/* ============================================================== training step ==========
 * Backward pass + optimiser of the reference's training loop (train_eda.py:182-192:
 * loss, stats = model(y, t); loss.backward(); optimizer.step(); scheduler.step()), SURVEY 8f n1.
 * Forward entry points with a "_train" suffix are the inference ones plus the few saved
 * quantities the backward kernels need.  Gradient GEMMs run on the same tcgen05 kernel as the
 * forward (prec = EDA_PREC_BF16X3 or EDA_PREC_BF16).                                         */

/* Dropout (nn.TransformerEncoderLayer in train() mode: on the attention weights, after out_proj, after the
 * ReLU, after linear2) is a counter-based Philox4x32-7 function of (seed, site, row, column): element
 * (row, col) is kept iff 16-bit half (col & 1) of word ((col & 7) >> 1) of philox(key = seed, counter =
 * (col >> 3, row, site, 0)) >= round(p * 2^16) (one call decides 8 elements; cols % 8 == 0 in eda_dropout_mask),
 * kept values are multiplied by 1 / (1 - p).  The forward epilogues and the backward kernels evaluate the
 * same function, so no mask is stored; eda_dropout_mask writes one out (tests).  For the attention weights
 * row = (b * n_heads + h) * T + query, col = key.                                                          */
int eda_dropout_mask(float p_drop, uint64_t seed, int site, int64_t rows, int cols, uint8_t* keep, void* stream);
/* The same mask as bits: bits[row * ld_words + w] holds columns 32w .. 32w+31 (bit e set = column 32w + e kept;
 * ld_words >= ceil(cols / 32)).  The attention kernels take it as `drop_bits` (may be NULL = evaluate the
 * function in the kernel): one precomputed word per thread and 32 keys replaces 4 Philox calls in the softmax
 * warps of BOTH the forward and the backward kernel (the one place where a mask is stored: 1 bit per weight). */
int eda_dropout_bits(float p_drop, uint64_t seed, int site, int64_t rows, int cols, uint32_t* bits, int ld_words, void* stream);

/* eda_gemm_split for model.train(): LayerNorm epilogues also store rstd (M) = 1/sqrt(var + eps) per row
 * (rstd_out may be NULL); p_drop > 0 applies dropout to relu(A.W^T + bias) (EDA_EPI_BIAS_RELU) or to
 * A.W^T + bias before the residual is added (EDA_EPI_BIAS_RES_LN).                                          */
int eda_gemm_split_train(const void* a_hi, const void* a_lo, int lda, const void* w_hi, const void* w_lo, int ldw,
                         const float* bias, const float* resid, int ldr, const float* ln_gamma, const float* ln_beta,
                         float* C, int ldc, void* c_hi, void* c_lo, int ldcs, float* rstd_out, float p_drop,
                         uint64_t seed, int site, int M, int N, int K, int epilogue, int prec, void* stream);

/* Input gradient of nn.Linear: dX (M, K_in) = dY (M, N_out) . W, with W handed over TRANSPOSED
 * (wt = W^T, (K_in, N_out) row-major split pair, so the contraction index is contiguous).
 * Optional fusions: + resid (fp32 (M, K_in): the other branch of a residual join; K_in == 256)
 * or zeroing where relu_mask_hi (bf16 (M, K_in), the forward activation) is <= 0, kept values times
 * mask_scale (1 / (1 - p) when the activation went through dropout).
 * Outputs: fp32 dX and/or split dx_hi/dx_lo.                                                   */
int eda_gemm_dgrad(const void* dy_hi, const void* dy_lo, int ld_dy, const void* wt_hi, const void* wt_lo, int ldw,
                   const float* resid, int ldr, const void* relu_mask_hi, int ldm, float mask_scale, float* dX, int ldx,
                   void* dx_hi, void* dx_lo, int ldxs, int M, int K_in, int N_out, int prec, void* stream);

/* Weight gradient of nn.Linear: dW (N_out, K_in) (+)= dY^T (N_out, M) . X (M, K_in).  Both operands
 * are read as they lie in HBM (row = sample) through MN-major shared-memory descriptors; the
 * reduction over M is cut into slices whose partial results live in `workspace`
 * (>= N_out * round_up(K_in, 4) * 4 bytes; more lets more CTA pairs work) and are added in
 * slice order.                                                                                 */
int eda_gemm_wgrad(const void* dy_hi, const void* dy_lo, int ld_dy, const void* x_hi, const void* x_lo, int ld_x,
                   int M, int N_out, int K_in, float* dW, int ld_dw, void* workspace, int64_t workspace_bytes,
                   int accumulate, int prec, void* stream);

/* nn.LayerNorm(256) backward from the forward OUTPUT y (xhat = (y - beta) / gamma; gamma != 0)
 * and the saved rstd: dx fp32 and/or split; dgamma, dbeta and (optional) the column sums of dx
 * (= bias gradient of the Linear feeding the norm).  workspace >= 592 * 3 KB.  With p_drop > 0 the
 * split output and the column sums carry the forward's dropout mask of `site` (the Linear branch),
 * the fp32 dx (residual branch) does not.                                                         */
int eda_ln_bwd(const float* dy, int ld_dy, const float* y, int ld_y, const float* gamma, const float* beta,
               const float* rstd, int M, int D, float* dx, int ld_dx, void* dx_hi, void* dx_lo, int ld_dxs,
               float* dgamma, float* dbeta, float* dcolsum, void* workspace, int64_t workspace_bytes, int accumulate,
               float p_drop, uint64_t seed, int site, void* stream);

/* out (N) (+)= column sums of a split matrix (M, N): bias gradients.  N % 8 == 0, N <= 2048;
 * workspace >= 592 * N * 4 B.                                                                    */
int eda_colsum_split(const void* hi, const void* lo, int M, int N, int ld, float* out, void* workspace,
                     int64_t workspace_bytes, int accumulate, void* stream);

/* eda_mha_split that also stores lse (B*n_heads, T): log2-domain log-sum-exp of every score row. */
int eda_mha_split_train(const void* qk_hi, const void* qk_lo, int ld_qk, const void* vt_hi, const void* vt_lo,
                        int Tp, int B, int T, int n_heads, void* out_hi, void* out_lo, int ld_os, float* lse,
                        float p_drop, uint64_t seed, int site, const uint32_t* drop_bits, int ld_bits, int prec, void* stream);
/* delta (B*T, 4) = rowsum over each head's 64 columns of dO * O (both split pairs, D == 256). */
int eda_mha_delta(const void* do_hi, const void* do_lo, int ld_do, const void* o_hi, const void* o_lo, int ld_o,
                  int M, int D, float* delta, void* stream);
/* Attention backward on the layouts eda_gemm_qkv wrote.  dk, dv are written as split pairs into
 * columns [D, 2D) and [2D, 3D) of dqkv (B*T, ld_dqkv); dq' (w.r.t. the pre-scaled q) is
 * accumulated in dq_acc (B*T, D) fp32 (zeroed by the call) - follow with
 * eda_scale_split(dq_acc, 1/8) into columns [0, D) of dqkv.                                      */
int eda_mha_bwd(const void* qk_hi, const void* qk_lo, int ld_qk, const void* vt_hi, const void* vt_lo, int Tp,
                const void* do_hi, const void* do_lo, int ld_do, const float* lse, const float* delta, int B,
                int T, int n_heads, float* dq_acc, void* dqkv_hi, void* dqkv_lo, int ld_dqkv, float p_drop,
                uint64_t seed, int site, const uint32_t* drop_bits, int ld_bits, int prec, void* stream);
/* (hi, lo) = split(src * scale): fp32 (rows, cols) -> split pair with row stride ld_s.           */
int eda_scale_split(const float* src, int ld, int rows, int cols, float scale, void* hi, void* lo, int ld_s, void* stream);

/* eda_lstm_attractors that keeps, per step, the activated gates and cell state and (split bf16)
 * the hidden state BEFORE the step: encoder gates/c rows are b*Tmax + step, hprev rows are
 * b*Tmax + perm[step] (frame order); decoder rows are b*n_dec + step.  hc_out is required.      */
int eda_lstm_attractors_train(const float* xproj, const int32_t* perm, const int32_t* lens, const int32_t* order,
                              const float* w_hh_enc, const float* w_hh_dec, const float* dec_bias, int B, int Tmax,
                              int D, int n_dec, float* attractors, float* hc_out, float* gates, float* c,
                              void* hprev_hi, void* hprev_lo, float* dec_gates, float* dec_c, void* dec_hprev_hi,
                              void* dec_hprev_lo, int prec, void* stream);
/* BPTT through decoder and encoder from dattractors (B, n_dec, D).  Writes the gradient w.r.t.
 * the gate pre-activations as split pairs: dxproj (B*Tmax, 4D) in FRAME order (rows of padded
 * frames are zero), ddec (B*n_dec, 4D).  The caller turns them into weight / input gradients
 * with eda_gemm_wgrad / eda_gemm_dgrad / eda_colsum_split.                                       */
int eda_lstm_attractors_bwd(const int32_t* perm, const int32_t* lens, const int32_t* order,
                            const float* w_hh_enc, const float* w_hh_dec, const float* gates, const float* c,
                            const float* dec_gates, const float* dec_c, const float* hc, const float* dattractors,
                            int B, int Tmax, int D, int n_dec, void* dxproj_hi, void* dxproj_lo,
                            void* ddec_hi, void* ddec_lo, void* stream);

/* Backward of eda_posterior's logits: demb (B*Tmax, ld_demb) written; dattractors
 * (B, att_stride_rows, D) accumulated with atomics (zero or pre-fill it first).  n_att <= 8.     */
int eda_posterior_bwd(const float* dpred, const float* emb, int ld_emb, const float* attractors, int att_stride_rows,
                      int B, int Tmax, int D, int n_att, float* demb, int ld_demb, float* dattractors, void* stream);
/* Backward of eda_attractor_head + the attractor loss scaled by `scale`: dattractors += ...,
 * dw (D) and dbias (1) written (or accumulated).                                                 */
int eda_attractor_head_bwd(const float* attractors, const float* probs, const float* w, int B, int n_dec, int D,
                           int n_spk_label, float scale, float* dattractors, float* dw, float* dbias, int accumulate,
                           void* stream);

/* x *= *scale_dev (the upstream gradient of the loss, kept on the device).                       */
int eda_scale_dev(const float* scale_dev, float* x, int64_t n, void* stream);
/* torch.optim.Adam (train_eda.py:131) on flat buffers; `step` counts from 1.                     */
int eda_adam_step(float* params, const float* grads, float* exp_avg, float* exp_avg_sq, int64_t n, float lr,
                  float beta1, float beta2, float eps, int step, void* stream);

/* ================================================= Conformer encoder variant ===
 * EendEdaModel(encoder_type="conformer") (models.py:498-502, 525-528): torchaudio.models.Conformer
 * layers = half-step SiLU feed-forward, pre-LN self-attention with key_padding_mask, convolution
 * module (pointwise -> GLU -> depthwise k=31 -> BatchNorm1d -> SiLU -> pointwise), half-step
 * feed-forward, final LayerNorm.  GEMMs run through eda_gemm_split_ex; inference (eval) only.   */

/* eda_gemm_split with the extra epilogues EDA_EPI_BIAS_RES / _SILU / _GLU and a scale on the GEMM term
 * of the residual epilogues: C = (A.W^T + bias) * out_scale + resid (then LayerNorm for RES_LN). */
int eda_gemm_split_ex(const void* a_hi, const void* a_lo, int lda, const void* w_hi, const void* w_lo, int ldw,
                      const float* bias, const float* resid, int ldr, const float* ln_gamma, const float* ln_beta,
                      float* C, int ldc, void* c_hi, void* c_lo, int ldcs, float out_scale, int M, int N, int K,
                      int epilogue, int prec, void* stream);
/* nn.LayerNorm(256) on fp32 rows -> fp32 `out` (may be NULL) and/or the split pair of the next GEMM. */
int eda_layernorm_split(const float* x, int ld, const float* gamma, const float* beta, int M, int D, float* out, int ld_out,
                        void* out_hi, void* out_lo, int ld_s, void* stream);
/* Depthwise Conv1d(256, 256, 31, padding 15, groups 256) + BatchNorm1d(eval) + SiLU on (B*T, 256) rows;
 * scale[c] = bn_w / sqrt(running_var + eps), shift[c] = (conv_bias - running_mean) * scale + bn_b. */
int eda_dwconv_bn_silu(const float* x, int ld, const float* w, const float* scale, const float* shift, int B, int T, int C,
                       int K, void* out_hi, void* out_lo, int ld_s, void* stream);
/* eda_mha_split with nn.MultiheadAttention's key_padding_mask: keys t >= key_lens[b] do not take part
 * (Conformer variant, models.py:525-528; and the transformer path's independent-chunk batches, see
 * eda_mha_fwd_masked). */
int eda_mha_split_masked(const void* qk_hi, const void* qk_lo, int ld_qk, const void* vt_hi, const void* vt_lo,
                         int Tp, int B, int T, int n_heads, const int32_t* key_lens, void* out_hi, void* out_lo,
                         int ld_os, int prec, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* EDA_B200_H_ */
