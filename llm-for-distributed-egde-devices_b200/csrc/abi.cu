// extern "C" surface of libb200stage.so -- see include/b200stage.h for the contract.
#include "common.cuh"
#include "../../include/b200stage.h"

namespace b2 {
int norm_quant(const void*, const void*, const void*, const float*, int8_t*, float*, int, int, float, cudaStream_t);
int norm_f16(const void*, const void*, const void*, void*, int, int, float, cudaStream_t);
int smooth_quant(const void*, const float*, int8_t*, float*, int, int, cudaStream_t);
int act_mul_quant(const void*, const float*, int8_t*, float*, int, int, int, cudaStream_t);
int gemm_w8a8(const int8_t*, const int8_t*, const float*, const float*, const void*, const void*, void*, int32_t*, int,
              int, int, int, cudaStream_t);
int lm_head(const void*, const void*, float*, int, int, int, cudaStream_t);
bool gemm2_supported(int, int, int);
int gemm2_plain(const int8_t*, const int8_t*, const float*, const float*, const void*, const void*, void*, int32_t*, int, int,
                int, cudaStream_t);
int gemm2_gated(const int8_t*, const int8_t*, const float*, const float*, const void*, const float*, float*, float*, int, int,
                int, int, cudaStream_t);
int gemm2_qkv(const int8_t*, const int8_t*, const float*, const float*, const void*, void*, long long, const float*,
              const float*, const int*, const int*, int, void*, void*, int, int, int, int, int, int, int, cudaStream_t);
int quant_f32(const float*, const float*, int8_t*, float*, int, int, cudaStream_t);
int init_gemm2();
int p2p_send(const void*, void*, long long, unsigned*, unsigned, cudaStream_t);
int p2p_wait(const unsigned*, unsigned, const void*, void*, long long, unsigned*, cudaStream_t);
int init_p2p();
int rope_kv_append(void*, long long, const float*, const float*, const int*, const int*, int, void*, void*, int, int,
                   int, int, int, int, cudaStream_t);
int decode_attn(const void*, long long, const void*, const void*, const int*, const int*, void*, long long, void*,
                size_t, int, int, int, int, int, int, float, int, const void*, long long, const void*, long long,
                cudaStream_t);
int decode_attn_static(const void*, long long, const void*, const void*, const int*, const int*, void*, long long, int,
                       int, int, int, int, int, float, int, cudaStream_t);
int decode_attn_splits(int, int, int, int);
size_t decode_attn_workspace_bytes(int, int, int, int);
int prefill_attn(const void*, long long, const void*, const void*, void*, long long, const int*, int, int, int, int,
                 int, int, int, float, int, cudaStream_t);
int init_gemm();
int act_mul_f16(const void*, void*, int, int, int, cudaStream_t);
int outlier_scan(const void*, long long, int, int, float, float*, int*, int*, unsigned*, cudaStream_t);
int outlier_side(void*, const void*, long long, const int8_t*, const float*, const int*, const int*, const void*, int, int,
                 int, cudaStream_t);
int outlier_side_dense(void*, const void*, long long, const int8_t*, const float*, const float*, const void*, void*, void*,
                       float*, int, int, int, cudaStream_t);
int init_outlier();
int init_attn();
int init_attn_tc();
void set_prefill_attn_impl(int);
int init_rowquant();
int init_misc();
int init_fused();
int init_sample();
int sample(const float*, long long, const long long*, const int*, long long, long long*, int, int, float, int, float, float,
           unsigned long long, const unsigned long long*, unsigned long long, cudaStream_t);
int fused_layer_tail(const b2_fused_desc*, const b2_fused_desc*, const b2_fused_desc*, const b2_fused_desc*, int, unsigned*,
                     cudaStream_t);
int fused_qkv(const void*, long long, int, const void*, const void*, const float*, float, const int8_t*, const float*,
              const void*, void*, const float*, const float*, const int*, const int*, int, void*, void*, int, int, int,
              int, int, int, int, const void*, long long, const void*, long long, cudaStream_t);
int fused_linear(const void*, long long, int, const void*, const void*, const float*, const float*, float, const int8_t*,
                 const float*, const void*, const void*, void*, int, int, const float*, float*, float*, int, int, int,
                 const void*, long long, const void*, long long, cudaStream_t);
int embed(const long long*, const int*, const void*, const void*, void*, int, int, int, cudaStream_t);
int argmax_rows(const float*, long long*, int, int, cudaStream_t);
int confidence(const float*, long long, float*, float*, int, int, cudaStream_t);
int bias_add_f32(float*, const float*, long long, int, int, cudaStream_t);
int argmax_shard(const float*, long long, long long*, float*, long long, const float*, const long long*, int, int, cudaStream_t);
size_t head_argmax_scratch_bytes();
int lm_head_argmax(const void*, const void*, float*, int, int, int, long long, const float*, const long long*, long long*, long long*,
                   float*, void*, const unsigned*, unsigned, unsigned*, unsigned*, unsigned, cudaStream_t);
int p2p_send_multi(const void*, long long, void* const*, unsigned* const*, int, unsigned, cudaStream_t);
int merge_winners(const unsigned*, unsigned, const void*, int, int, long long*, long long*, unsigned*, cudaStream_t);
}  // namespace b2

#define ST(s) reinterpret_cast<cudaStream_t>(s)

extern "C" {

int b2_version(void) { return 100; }

int b2_init(void) {
  int rc = b2::init_gemm();
  if (rc) return rc;
  if ((rc = b2::init_gemm2())) return rc;
  if ((rc = b2::init_p2p())) return rc;
  if ((rc = b2::init_attn())) return rc;
  if ((rc = b2::init_attn_tc())) return rc;
  if ((rc = b2::init_outlier())) return rc;
  if ((rc = b2::init_rowquant())) return rc;
  if ((rc = b2::init_misc())) return rc;
  if ((rc = b2::init_fused())) return rc;
  return b2::init_sample();
}

const char* b2_strerror(int code) {
  switch (code) {
    case B2_OK: return "ok";
    case B2_ERR_BAD_SHAPE: return "bad shape";
    case B2_ERR_BAD_ALIGN: return "pointer not 16-byte aligned";
    case B2_ERR_NULL_PTR: return "null pointer";
    case B2_ERR_ARCH: return "device/attribute not supported (sm_100a required)";
    case B2_ERR_LAUNCH: return "kernel launch failed";
    case B2_ERR_UNSUPPORTED: return "unsupported configuration";
    case B2_ERR_DRIVER: return "cuTensorMapEncodeTiled unavailable or failed";
    case B2_ERR_TOO_LARGE: return "size exceeds the supported maximum";
    default: return "unknown error";
  }
}

int b2_norm_quant(const void* x, const void* gamma, const void* beta, const float* smooth, int8_t* q, float* row_scale,
                  int T, int d, float eps, void* stream) {
  return b2::norm_quant(x, gamma, beta, smooth, q, row_scale, T, d, eps, ST(stream));
}
int b2_norm(const void* x, const void* gamma, const void* beta, void* y, int T, int d, float eps, void* stream) {
  return b2::norm_f16(x, gamma, beta, y, T, d, eps, ST(stream));
}
int b2_smooth_quant(const void* x, const float* smooth, int8_t* q, float* row_scale, int T, int K, void* stream) {
  return b2::smooth_quant(x, smooth, q, row_scale, T, K, ST(stream));
}
int b2_act_mul_quant(const void* gate_up, const float* smooth, int8_t* q, float* row_scale, int T, int F, int act,
                     void* stream) {
  return b2::act_mul_quant(gate_up, smooth, q, row_scale, T, F, act, ST(stream));
}
int b2_gemm_w8a8(const int8_t* A, const int8_t* W, const float* a_scale, const float* w_scale, const void* bias,
                 const void* residual, void* out, int32_t* acc_dump, int M, int N, int K, void* stream) {
  if (b2::gemm2_supported(M, N, K))
    return b2::gemm2_plain(A, W, a_scale, w_scale, bias, residual, out, acc_dump, M, N, K, ST(stream));
  return b2::gemm_w8a8(A, W, a_scale, w_scale, bias, residual, out, acc_dump, M, N, K, 0, ST(stream));
}
int b2_gemm_w8a8_sched(const int8_t* A, const int8_t* W, const float* a_scale, const float* w_scale, const void* bias,
                       const void* residual, void* out, int32_t* acc_dump, int M, int N, int K, int schedule,
                       void* stream) {
  if (schedule == 20) return b2::gemm2_plain(A, W, a_scale, w_scale, bias, residual, out, acc_dump, M, N, K, ST(stream));
  if (schedule == 0 && b2::gemm2_supported(M, N, K))
    return b2::gemm2_plain(A, W, a_scale, w_scale, bias, residual, out, acc_dump, M, N, K, ST(stream));
  return b2::gemm_w8a8(A, W, a_scale, w_scale, bias, residual, out, acc_dump, M, N, K, schedule, ST(stream));
}
int b2_p2p_send(const void* src, void* dst_peer, long long bytes, unsigned int* flag_peer, unsigned int seq, void* stream) {
  return b2::p2p_send(src, dst_peer, bytes, flag_peer, seq, ST(stream));
}
int b2_p2p_wait(const unsigned int* flag, unsigned int seq, const void* src, void* dst, long long bytes, unsigned int* err,
                void* stream) {
  return b2::p2p_wait(flag, seq, src, dst, bytes, err, ST(stream));
}
int b2_argmax_shard(const float* logits, long long ld, long long* out_idx, float* out_val, long long idx_offset,
                    const float* other_val, const long long* other_idx, int T, int V, void* stream) {
  return b2::argmax_shard(logits, ld, out_idx, out_val, idx_offset, other_val, other_idx, T, V, ST(stream));
}
long long b2_lm_head_argmax_scratch_bytes(void) { return (long long)b2::head_argmax_scratch_bytes(); }
int b2_lm_head_argmax(const void* x, const void* W, float* logits, int M, int V, int d, long long idx_offset,
                      const float* other_val, const long long* other_idx, long long* out_idx, long long* out_idx2,
                      float* out_val, void* scratch, const unsigned int* wait_flag, unsigned int wait_seq, unsigned int* err,
                      unsigned int* notify_flag, unsigned int notify_seq, void* stream) {
  return b2::lm_head_argmax(x, W, logits, M, V, d, idx_offset, other_val, other_idx, out_idx, out_idx2, out_val, scratch, wait_flag,
                            wait_seq, err, notify_flag, notify_seq, ST(stream));
}
int b2_p2p_send_multi(const void* src, long long bytes, void* const* dsts, unsigned int* const* flags, int n, unsigned int seq,
                      void* stream) {
  return b2::p2p_send_multi(src, bytes, dsts, flags, n, seq, ST(stream));
}
int b2_merge_winners(const unsigned int* flags, unsigned int seq, const void* records, int W, int B, long long* out_idx,
                     long long* out_idx2, unsigned int* err, void* stream) {
  return b2::merge_winners(flags, seq, records, W, B, out_idx, out_idx2, err, ST(stream));
}
int b2_bias_add_f32(float* logits, const float* bias, long long ld, int T, int V, void* stream) {
  return b2::bias_add_f32(logits, bias, ld, T, V, ST(stream));
}
int b2_confidence(const float* logits, long long ld, float* row_pmax, float* out, int T, int V, void* stream) {
  return b2::confidence(logits, ld, row_pmax, out, T, V, ST(stream));
}
int b2_gemm_gated_act(const int8_t* A, const int8_t* W, const float* a_scale, const float* w_scale, const void* bias,
                      const float* out_smooth, float* y32, float* amax, int M, int F, int K, int act, void* stream) {
  return b2::gemm2_gated(A, W, a_scale, w_scale, bias, out_smooth, y32, amax, M, F, K, act, ST(stream));
}
int b2_quant_f32(const float* y32, const float* amax, int8_t* q, float* row_scale, int T, int F, void* stream) {
  return b2::quant_f32(y32, amax, q, row_scale, T, F, ST(stream));
}
int b2_gemm_qkv_rope(const int8_t* A, const int8_t* W, const float* a_scale, const float* w_scale, const void* bias,
                     void* qkv_out, long long ld, const float* cos_table, const float* sin_table, const int* positions,
                     const int* seq_ids, int seq0, void* kcache, void* vcache, int M, int H, int KV, int hd, int Smax, int K,
                     int use_rope, void* stream) {
  return b2::gemm2_qkv(A, W, a_scale, w_scale, bias, qkv_out, ld, cos_table, sin_table, positions, seq_ids, seq0, kcache,
                       vcache, M, H, KV, hd, Smax, K, use_rope, ST(stream));
}
int b2_fused_linear(const void* x, long long x_ld, int x_kind, const void* gamma, const void* beta, const float* smooth,
                    const float* x_amax, float eps, const int8_t* W, const float* w_scale, const void* bias,
                    const void* residual, void* out, int out_kind, int act, const float* out_smooth, float* out_amax,
                    float* zero_amax, int M, int N, int K, const void* prefetch0, long long prefetch0_bytes,
                    const void* prefetch1, long long prefetch1_bytes, void* stream) {
  return b2::fused_linear(x, x_ld, x_kind, gamma, beta, smooth, x_amax, eps, W, w_scale, bias, residual, out, out_kind,
                          act, out_smooth, out_amax, zero_amax, M, N, K, prefetch0, prefetch0_bytes, prefetch1,
                          prefetch1_bytes, ST(stream));
}
int b2_fused_qkv(const void* x, long long x_ld, int x_kind, const void* gamma, const void* beta, const float* smooth,
                 float eps, const int8_t* W, const float* w_scale, const void* bias, void* qkv_out,
                 const float* cos_table, const float* sin_table, const int* positions, const int* seq_ids, int seq0,
                 void* kcache, void* vcache, int M, int H, int KV, int hd, int Smax, int K, int use_rope,
                 const void* prefetch0, long long prefetch0_bytes, const void* prefetch1, long long prefetch1_bytes,
                 void* stream) {
  return b2::fused_qkv(x, x_ld, x_kind, gamma, beta, smooth, eps, W, w_scale, bias, qkv_out, cos_table, sin_table,
                       positions, seq_ids, seq0, kcache, vcache, M, H, KV, hd, Smax, K, use_rope, prefetch0,
                       prefetch0_bytes, prefetch1, prefetch1_bytes, ST(stream));
}
int b2_fused_layer_tail(const b2_fused_desc* o_proj, const b2_fused_desc* gate_up, const b2_fused_desc* down,
                        const b2_fused_desc* next_qkv, int M, unsigned int* counters, void* stream) {
  return b2::fused_layer_tail(o_proj, gate_up, down, next_qkv, M, counters, ST(stream));
}
int b2_sample(const float* logits, long long ld, const long long* history, const int* hist_len, long long hist_ld,
              long long* out, int B, int V, float temperature, int top_k, float top_p, float repetition_penalty,
              unsigned long long seed, const unsigned long long* step_ptr, unsigned long long step, void* stream) {
  return b2::sample(logits, ld, history, hist_len, hist_ld, out, B, V, temperature, top_k, top_p, repetition_penalty, seed,
                    step_ptr, step, ST(stream));
}
int b2_rope_kv_append(void* qkv, long long ld, const float* cos_table, const float* sin_table, const int* positions,
                      const int* seq_ids, int seq0, void* kcache, void* vcache, int T, int H, int KV, int hd, int Smax,
                      int use_rope, void* stream) {
  return b2::rope_kv_append(qkv, ld, cos_table, sin_table, positions, seq_ids, seq0, kcache, vcache, T, H, KV, hd, Smax,
                            use_rope, ST(stream));
}
int b2_decode_attn(const void* q, long long q_ld, const void* kcache, const void* vcache, const int* seqlens,
                   const int* rows, void* out, long long o_ld, void* workspace, size_t workspace_bytes, int B, int H,
                   int KV, int hd, int Smax, int max_seqlen, float scale, int window, const void* prefetch0,
                   long long prefetch0_bytes, const void* prefetch1, long long prefetch1_bytes, void* stream) {
  return b2::decode_attn(q, q_ld, kcache, vcache, seqlens, rows, out, o_ld, workspace, workspace_bytes, B, H, KV, hd,
                         Smax, max_seqlen, scale, window, prefetch0, prefetch0_bytes, prefetch1, prefetch1_bytes,
                         ST(stream));
}
int b2_decode_attn_static(const void* q, long long q_ld, const void* kcache, const void* vcache, const int* seqlens,
                          const int* rows, void* out, long long o_ld, int B, int H, int KV, int hd, int Smax,
                          int max_seqlen, float scale, int window, void* stream) {
  return b2::decode_attn_static(q, q_ld, kcache, vcache, seqlens, rows, out, o_ld, B, H, KV, hd, Smax, max_seqlen, scale,
                                window, ST(stream));
}
size_t b2_decode_attn_workspace_bytes(int B, int H, int KV, int hd, int max_seqlen, int window) {
  int dev = 0, sms = 148;
  if (cudaGetDevice(&dev) == cudaSuccess) {
    int v = 0;
    if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && v > 0) sms = v;
  }
  const int eff = (window > 0 && max_seqlen > window) ? window : max_seqlen;
  return b2::decode_attn_workspace_bytes(B, H, hd, b2::decode_attn_splits(B, KV, eff, sms));
}
int b2_prefill_attn(const void* q, long long q_ld, const void* kcache, const void* vcache, void* out, long long o_ld,
                    const int* positions, int T, int pos0, int row, int H, int KV, int hd, int Smax, float scale,
                    int window, void* stream) {
  return b2::prefill_attn(q, q_ld, kcache, vcache, out, o_ld, positions, T, pos0, row, H, KV, hd, Smax, scale, window, ST(stream));
}
int b2_act_mul(const void* gate_up, void* y, int T, int F, int act, void* stream) {
  return b2::act_mul_f16(gate_up, y, T, F, act, ST(stream));
}
int b2_outlier_scan(const void* y, long long ld, int T, int K, float threshold, float* mask, int* cols, int* count,
                    unsigned* flags, void* stream) {
  return b2::outlier_scan(y, ld, T, K, threshold, mask, cols, count, flags, ST(stream));
}
int b2_outlier_side(void* out, const void* y, long long ld, const int8_t* W, const float* w_scale, const int* cols,
                    const int* count, const void* residual, int T, int N, int K, void* stream) {
  return b2::outlier_side(out, y, ld, W, w_scale, cols, count, residual, T, N, K, ST(stream));
}
int b2_outlier_side_dense(void* out, const void* y, long long ld, const int8_t* W, const float* w_scale, const float* mask,
                          const void* residual, void* ysub, void* wdeq, float* side, int T, int N, int K, void* stream) {
  return b2::outlier_side_dense(out, y, ld, W, w_scale, mask, residual, ysub, wdeq, side, T, N, K, ST(stream));
}
int b2_set_prefill_attn_impl(int impl) {
  if (impl < 0 || impl > 2) return b2::ERR_UNSUPPORTED;
  b2::set_prefill_attn_impl(impl);
  return b2::OK;
}
int b2_embed(const long long* ids, const int* positions, const void* table, const void* pos_table, void* out, int T,
             int d, int vocab, void* stream) {
  return b2::embed(ids, positions, table, pos_table, out, T, d, vocab, ST(stream));
}
int b2_lm_head(const void* x, const void* W, float* logits, int M, int V, int d, void* stream) {
  return b2::lm_head(x, W, logits, M, V, d, ST(stream));
}
int b2_argmax(const float* logits, long long* out, int T, int V, void* stream) {
  return b2::argmax_rows(logits, out, T, V, ST(stream));
}

}  // extern "C"
