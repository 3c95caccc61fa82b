// extern "C" entry points of the model path: argument checking + dispatch on precision.
#include "common.cuh"

namespace eda {
int linear_f32(const float*, int, const float*, int, const float*, const float*, int, const float*,
               const float*, float*, int, int, int, int, int, cudaStream_t);
int linear_tc(const float*, int, const float*, int, const float*, const float*, int, const float*,
              const float*, float*, int, int, int, int, int, int, cudaStream_t);
int mha_f32(const float*, int, float*, int, int, int, int, cudaStream_t, const int32_t* key_lens = nullptr);
int mha_tc(const float*, int, float*, int, int, int, int, int, cudaStream_t);
struct LstmSave { float* gates; float* c; void* hprev_hi; void* hprev_lo; float* dec_gates; float* dec_c; void* dec_hprev_hi; void* dec_hprev_lo; };
int lstm_attractors(const float*, const int32_t*, const int32_t*, const int32_t*, const float*,
                    const float*, const float*, int, int, int, int, float*, float*, const LstmSave&, int, cudaStream_t);
int pad_pack(const float*, const int64_t*, const int32_t*, int, int, int, float, float*, int, cudaStream_t);
int pad_pack_split(const float*, int, const int64_t*, const int32_t*, int, int, int, float, void*, void*, int, cudaStream_t);
int attractor_head(const float*, const float*, const float*, int, int, int, float*, int, float*, cudaStream_t);
int posterior(const float*, int, const float*, int, int, int, int, int, float, float*, float*, uint8_t*, cudaStream_t,
              const int64_t* out_off = nullptr, const int32_t* lens = nullptr, uint16_t* mask16 = nullptr);
int pit_loss(const float*, const float*, const int32_t*, int, int, int, float*, int64_t*, float*, int32_t*, cudaStream_t);
int pit_loss_static(const float*, const float*, const int32_t*, int, int, int, float*, int64_t*, float*, int32_t*, cudaStream_t);
int pit_loss_bwd(const float*, const float*, const int32_t*, int, int, int, float, float*, cudaStream_t);
int der_counters(const float*, const float*, const int32_t*, int, int, int, int64_t*, cudaStream_t);
}  // namespace eda

extern "C" {

int eda_pad_pack(const float* src, const int64_t* seq_off, const int32_t* lens, int B, int Tmax,
                 int n_cols, float pad, float* dst, int ld_dst, void* stream) {
  return eda::pad_pack(src, seq_off, lens, B, Tmax, n_cols, pad, dst, ld_dst, (cudaStream_t)stream);
}

int eda_pad_pack_split(const float* src, int ld_src, const int64_t* seq_off, const int32_t* lens, int B,
                       int Tmax, int n_cols, float pad, void* dst_hi, void* dst_lo, int ld_dst, void* stream) {
  EDA_REQUIRE(src && seq_off && lens && dst_hi && dst_lo, "eda_pad_pack_split: null argument");
  return eda::pad_pack_split(src, ld_src, seq_off, lens, B, Tmax, n_cols, pad, dst_hi, dst_lo, ld_dst, (cudaStream_t)stream);
}

int eda_linear(const float* A, int lda, const float* W, int ldw, const float* bias,
               const float* resid, int ldr, const float* ln_gamma, const float* ln_beta, float* C,
               int ldc, int M, int N, int K, int epilogue, int prec, void* stream) {
  EDA_REQUIRE(A && W && C && M > 0 && N > 0 && K > 0, "eda_linear: bad arguments");
  if (prec == EDA_PREC_FP32)
    return eda::linear_f32(A, lda, W, ldw, bias, resid, ldr, ln_gamma, ln_beta, C, ldc, M, N, K, epilogue, (cudaStream_t)stream);
  if (prec == EDA_PREC_BF16X3 || prec == EDA_PREC_BF16)
    return eda::linear_tc(A, lda, W, ldw, bias, resid, ldr, ln_gamma, ln_beta, C, ldc, M, N, K, epilogue, prec, (cudaStream_t)stream);
  eda::set_error("eda_linear: unknown precision %d", prec);
  return EDA_EINVAL;
}

int eda_mha_fwd(const float* qkv, int ld_qkv, float* out, int ld_out, int B, int T, int n_heads,
                int head_dim, int prec, void* stream) {
  EDA_REQUIRE(qkv && out && B > 0 && T > 0 && n_heads > 0, "eda_mha_fwd: bad arguments");
  EDA_REQUIRE(head_dim == 64, "eda_mha_fwd: head_dim must be 64 (got %d)", head_dim);
  if (prec == EDA_PREC_FP32) return eda::mha_f32(qkv, ld_qkv, out, ld_out, B, T, n_heads, (cudaStream_t)stream);
  return eda::mha_tc(qkv, ld_qkv, out, ld_out, B, T, n_heads, prec, (cudaStream_t)stream);
}

int eda_mha_fwd_masked(const float* qkv, int ld_qkv, float* out, int ld_out, int B, int T, int n_heads,
                       int head_dim, const int32_t* key_lens, int prec, void* stream) {
  EDA_REQUIRE(qkv && out && key_lens && B > 0 && T > 0 && n_heads > 0, "eda_mha_fwd_masked: bad arguments");
  EDA_REQUIRE(head_dim == 64, "eda_mha_fwd_masked: head_dim must be 64 (got %d)", head_dim);
  EDA_REQUIRE(prec == EDA_PREC_FP32, "eda_mha_fwd_masked: fp32 only (the tensor-core form is eda_mha_split_masked)");
  return eda::mha_f32(qkv, ld_qkv, out, ld_out, B, T, n_heads, (cudaStream_t)stream, key_lens);
}

int eda_lstm_attractors(const float* xproj, const int32_t* perm, const int32_t* lens,
                        const int32_t* order, const float* w_hh_enc, const float* w_hh_dec,
                        const float* dec_bias, int B, int Tmax, int D, int n_dec, float* attractors,
                        float* hc_out, int prec, void* stream) {
  EDA_REQUIRE(xproj && perm && lens && w_hh_enc && w_hh_dec && dec_bias && attractors, "eda_lstm_attractors: null argument");
  eda::LstmSave sv = {};
  return eda::lstm_attractors(xproj, perm, lens, order, w_hh_enc, w_hh_dec, dec_bias, B, Tmax, D, n_dec, attractors, hc_out, sv, prec, (cudaStream_t)stream);
}

int eda_lstm_attractors_train(const float* xproj, const int32_t* perm, const int32_t* lens, const int32_t* order,
                              const float* w_hh_enc, const float* w_hh_dec, const float* dec_bias, int B, int Tmax,
                              int D, int n_dec, float* attractors, float* hc_out, float* gates, float* c,
                              void* hprev_hi, void* hprev_lo, float* dec_gates, float* dec_c, void* dec_hprev_hi,
                              void* dec_hprev_lo, int prec, void* stream) {
  EDA_REQUIRE(xproj && perm && lens && w_hh_enc && w_hh_dec && dec_bias && attractors && gates, "eda_lstm_attractors_train: null argument");
  eda::LstmSave sv = {gates, c, hprev_hi, hprev_lo, dec_gates, dec_c, dec_hprev_hi, dec_hprev_lo};
  return eda::lstm_attractors(xproj, perm, lens, order, w_hh_enc, w_hh_dec, dec_bias, B, Tmax, D, n_dec, attractors, hc_out, sv, prec, (cudaStream_t)stream);
}

int eda_attractor_head(const float* attractors, const float* w, const float* bias, int B, int n_dec,
                       int D, float* probs, int n_spk_label, float* loss_att, void* stream) {
  EDA_REQUIRE(attractors && w && bias && probs, "eda_attractor_head: null argument");
  return eda::attractor_head(attractors, w, bias, B, n_dec, D, probs, n_spk_label, loss_att, (cudaStream_t)stream);
}

int eda_posterior(const float* emb, int ld_emb, const float* attractors, int att_stride_rows, int B,
                  int Tmax, int D, int n_att, float threshold, float* logits, float* post,
                  uint8_t* labels, void* stream) {
  EDA_REQUIRE(emb && attractors && B > 0 && Tmax > 0, "eda_posterior: bad arguments");
  return eda::posterior(emb, ld_emb, attractors, att_stride_rows, B, Tmax, D, n_att, threshold, logits, post, labels, (cudaStream_t)stream);
}

int eda_posterior_packed(const float* emb, int ld_emb, const float* attractors, int att_stride_rows, int B,
                         int Tmax, int D, int n_att, float threshold, const int64_t* out_off, const int32_t* lens,
                         float* logits, float* post, uint8_t* labels, uint16_t* mask16, void* stream) {
  EDA_REQUIRE(emb && attractors && out_off && lens && B > 0 && Tmax > 0, "eda_posterior_packed: bad arguments");
  return eda::posterior(emb, ld_emb, attractors, att_stride_rows, B, Tmax, D, n_att, threshold, logits, post, labels,
                        (cudaStream_t)stream, out_off, lens, mask16);
}

int eda_pit_loss(const float* pred, const float* label, const int32_t* lens, int B, int T, int S,
                 float* loss, int64_t* min_idx, float* label_perm, int32_t* perm_out, void* stream) {
  EDA_REQUIRE(pred && label && lens, "eda_pit_loss: null argument");
  return eda::pit_loss(pred, label, lens, B, T, S, loss, min_idx, label_perm, perm_out, (cudaStream_t)stream);
}

int eda_pit_loss_static(const float* pred, const float* label, const int32_t* lens, int B, int T, int S,
                        float* loss, int64_t* min_idx, float* label_perm, int32_t* perm_out, void* stream) {
  EDA_REQUIRE(pred && label && lens, "eda_pit_loss_static: null argument");
  return eda::pit_loss_static(pred, label, lens, B, T, S, loss, min_idx, label_perm, perm_out, (cudaStream_t)stream);
}

int eda_pit_loss_bwd(const float* pred, const float* label_perm, const int32_t* lens, int B, int T,
                     int S, float grad_scale, float* dpred, void* stream) {
  EDA_REQUIRE(pred && label_perm && lens && dpred, "eda_pit_loss_bwd: null argument");
  return eda::pit_loss_bwd(pred, label_perm, lens, B, T, S, grad_scale, dpred, (cudaStream_t)stream);
}

int eda_der_counters(const float* pred, const float* label_perm, const int32_t* lens, int B, int T,
                     int S, int64_t* counts, void* stream) {
  EDA_REQUIRE(pred && label_perm && lens, "eda_der_counters: null argument");
  return eda::der_counters(pred, label_perm, lens, B, T, S, counts, (cudaStream_t)stream);
}

}  // extern "C"
