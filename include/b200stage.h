/* b200stage -- C ABI of the B200-native W8A8 decoder-block hot path.
 *
 * This is the drop-in boundary: plain pointers and sizes, no torch types, no C++ exceptions.
 * Every pointer is a DEVICE pointer unless said otherwise; `stream` is a cudaStream_t passed
 * as void* (NULL = legacy default stream).  All calls are asynchronous with respect to `stream`,
 * re-entrant, and allocate nothing: every buffer, including workspaces, belongs to the caller.
 * Return value: 0 (B2_OK) or a negative B2_ERR_* code; b2_strerror() names it.  An unsupported
 * shape is an error, never a silent fallback.
 *
 * The reference (parthabp55/LLM-for-Distributed-Egde-Devices) has no FFI or native code; its
 * decoder-block forward runs inside third-party modules reached from the call sites cited per
 * entry point below (paths relative to the reference root).  "Replaces" therefore names the
 * reference call site plus the third-party routine (transformers 5.5.0 paths under
 * site-packages/transformers/) whose work the entry point takes over.
 */
#ifndef B200STAGE_H
#define B200STAGE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B2_OK 0
#define B2_ERR_BAD_SHAPE (-1)
#define B2_ERR_BAD_ALIGN (-2)   /* pointers must be 16-byte aligned */
#define B2_ERR_NULL_PTR (-3)
#define B2_ERR_ARCH (-4)        /* not an sm_100 device / attribute set failed */
#define B2_ERR_LAUNCH (-5)
#define B2_ERR_UNSUPPORTED (-6)
#define B2_ERR_DRIVER (-7)      /* cuTensorMapEncodeTiled unavailable or failed */
#define B2_ERR_TOO_LARGE (-8)

/* Library version (major*10000 + minor*100 + patch). */
int b2_version(void);
/* Static string for an error code. */
const char* b2_strerror(int code);
/* One-time set-up on the current device: opts every kernel in to its dynamic shared-memory size
 * and resolves cuTensorMapEncodeTiled.  Launch paths also do this lazily, but calling b2_init()
 * first guarantees no launch has first-call side effects -- required before CUDA-graph capture.
 * No counterpart in the reference (library housekeeping). */
int b2_init(void);

/* Fused (RMSNorm | LayerNorm) -> smoothing multiply -> per-token absmax int8 quantise.
 * x [T,d] fp16; gamma [d] fp16; beta [d] fp16 or NULL (NULL selects RMSNorm); smooth [d] fp32 or
 * NULL (the SmoothQuant 1/s multiplier); q [T,d] int8; row_scale [T] fp32 (amax/127).
 * Replaces: `input_layernorm` / `post_attention_layernorm` (models/llama/modeling_llama.py:62-67,
 * :303-334) plus the activation-quantising half of the int8 linear requested at
 * Code/Quantised Models/models_quant_updated.py:30-38 (bitsandbytes Linear8bitLt) and
 * Code/C-DAC Server/try.py:205 (quantize_dynamic). */
int b2_norm_quant(const void* x, const void* gamma, const void* beta, const float* smooth, int8_t* q,
                  float* row_scale, int T, int d, float eps, void* stream);

/* Norm only, fp16 result y [T,d] (the final norm before lm_head).
 * Replaces: `model.norm` (modeling_llama.py:62-67) reached from models_quant_updated.py:66-67, :84-93. */
int b2_norm(const void* x, const void* gamma, const void* beta, void* y, int T, int d, float eps, void* stream);

/* fp16 activation [T,K] -> smoothing multiply -> per-token int8 (input of o_proj).
 * Replaces: the activation quantisation inside Linear8bitLt for `o_proj` (models_quant_updated.py:30-38). */
int b2_smooth_quant(const void* x, const float* smooth, int8_t* q, float* row_scale, int T, int K, void* stream);

/* act(gate)*up -> smoothing -> per-token int8 (input of down_proj).  gate_up [T,2F] fp16, gate in
 * columns [0,F), up in [F,2F).  act: 0 silu*up, 1 gelu_new*up, 2 gelu_new (ungated, input [T,F]), 4 exact-erf gelu (ungated; GPT-NeoX),
 * 3 silu (ungated).
 * Replaces: `act_fn(gate_proj(x)) * up_proj(x)` (modeling_llama.py:182-184) plus the activation
 * quantisation of `down_proj`. */
int b2_act_mul_quant(const void* gate_up, const float* smooth, int8_t* q, float* row_scale, int T, int F, int act,
                     void* stream);

/* W8A8 linear with fused dequantise + bias + residual epilogue.
 * A [M,K] int8 row-major; W [N,K] int8 row-major; a_scale [M], w_scale [N] fp32; bias [N] fp16 or
 * NULL; residual [M,N] fp16 or NULL; out [M,N] fp16; acc_dump [M,N] int32 or NULL (test hook: the
 * raw accumulators).  K % 16 == 0, N % 8 == 0.
 * out = fp16(((float)acc * a_scale[m]) * w_scale[n] + bias[n] + residual[m,n]).
 * Replaces: every nn.Linear of the decoder layer as swapped for bitsandbytes Linear8bitLt by
 * `BitsAndBytesConfig(load_in_8bit=True)` (models_quant_updated.py:30-38; combiner_quant_updated.py:37)
 * or torch.ao dynamic qint8 Linear (try.py:205). */
int b2_gemm_w8a8(const int8_t* A, const int8_t* W, const float* a_scale, const float* w_scale, const void* bias,
                 const void* residual, void* out, int32_t* acc_dump, int M, int N, int K, void* stream);
/* Same, with the schedule forced: 0 = choose (what b2_gemm_w8a8 does), 1 = tcgen05 tensor-core path with
 * the tile heuristic, 2 = dp4a streaming path (M <= 8).  Test hook so both schedules are checked
 * on the same inputs.  10..16 pin the tcgen05 tile / cluster shape (N % 256 == 0 required):
 * 10/11/12 = 128x256 tile, 4 stages, cluster of 1/2/4 CTAs multicasting the weight tile;
 * 14/16/13 = 128x128 tile, 6 stages, cluster of 1/2/4; 15 = 128x64 tile, 8 stages.
 * 17 = 128x32 tile, 8 stages; 18 / 19 = 128x128 tile with only 32 / 64 activation rows staged per k-block
 * (decode batches: M <= 32 / 64, else B2_ERR_UNSUPPORTED) and 10 / 8 stages.
 * The cluster schedules (11, 12, 13, 16) need every byte of W in the current device's own memory (TMA multicast
 * loads here read only this GPU's memory; the whole span is checked, not its first byte) and return
 * B2_ERR_UNSUPPORTED otherwise.  20 = the CTA-pair kernel (tcgen05 cta_group::2, 256x256 tile per two SMs;
 * N % 32 == 0), which schedule 0 takes by itself for M > 128, N >= 256.  Other values: B2_ERR_UNSUPPORTED. */
int b2_gemm_w8a8_sched(const int8_t* A, const int8_t* W, const float* a_scale, const float* w_scale,
                       const void* bias, const void* residual, void* out, int32_t* acc_dump, int M, int N, int K,
                       int schedule, void* stream);

/* Prefill gate/up projection with the activation fused into the GEMM epilogue (CTA-pair tcgen05 kernel).
 * A [M,K] int8, a_scale [M]; W [2F,K] int8 = [gate rows; up rows], w_scale [2F], bias [2F] fp16 or NULL.
 * y32 [M,F] fp32 = act(fp16(gate)) * fp16(up) * out_smooth[f] (out_smooth [F] fp32 or NULL), amax [M] fp32 = the
 * row absmax of y32 (zeroed by this call on `stream`, then raised with atomic max).  act: 0 silu, 1 gelu_new.
 * F % 32 == 0, K % 16 == 0.  Followed by b2_quant_f32 the pair is bit-identical to
 * b2_gemm_w8a8 + b2_act_mul_quant without the [M,2F] fp16 round trip through HBM.
 * Replaces: `act_fn(gate_proj(x)) * up_proj(x)` (modeling_llama.py:182-184) with both projections as the int8
 * linears requested at models_quant_updated.py:30-38, in the prefill forward of `generate` (:84-93). */
int b2_gemm_gated_act(const int8_t* A, const int8_t* W, const float* a_scale, const float* w_scale, const void* bias,
                      const float* out_smooth, float* y32, float* amax, int M, int F, int K, int act, void* stream);
/* y32 [T,F] fp32 with its row absmax amax [T] (as written by b2_gemm_gated_act) -> q [T,F] int8, row_scale [T]
 * (amax/127): the quantise step of b2_act_mul_quant.  F % 16 == 0.
 * Replaces: the activation quantisation of `down_proj` (Linear8bitLt, models_quant_updated.py:30-38). */
int b2_quant_f32(const float* y32, const float* amax, int8_t* q, float* row_scale, int T, int F, void* stream);

/* Prefill q/k/v projection with rotary embedding and the KV-cache append fused into the GEMM epilogue
 * (CTA-pair tcgen05 kernel).  A [M,K] int8, a_scale [M]; W [(H+2KV)*hd, K] int8 = [q; k; v] rows, w_scale, bias
 * likewise.  Rotated q goes to qkv_out [M, ld] (only its q columns [0,H*hd) are written), rotated k and v to the
 * caches [rows][KV][Smax][hd] at (seq_ids[m] or seq0, positions[m]).  hd 64 or 128.  Bit-identical to
 * b2_gemm_w8a8 + b2_rope_kv_append.
 * Replaces: q/k/v_proj + apply_rotary_pos_emb + DynamicCache.update (modeling_llama.py:128-153, :251-289) in the
 * prefill forward of `generate` (models_quant_updated.py:84-93) and of confidence_score (:63-70). */
int b2_gemm_qkv_rope(const int8_t* A, const int8_t* W, const float* a_scale, const float* w_scale, const void* bias,
                     void* qkv_out, long long ld, const float* cos_table, const float* sin_table, const int* positions,
                     const int* seq_ids, int seq0, void* kcache, void* vcache, int M, int H, int KV, int hd, int Smax, int K,
                     int use_rope, void* stream);

/* Decode-time fused linear for M <= 8 tokens: activation prologue -> per-token int8 quantise ->
 * W8A8 weight-streaming GEMV -> epilogue, in ONE launch (programmatic dependent launch: the weight
 * prefetch overlaps the previous kernel's tail).  Bit-identical to the unfused entry points.
 * x_kind: 0 = fp16 x + RMSNorm(gamma) [+smooth]   (== b2_norm_quant, beta NULL)
 *         1 = fp16 x + LayerNorm(gamma, beta) [+smooth]
 *         2 = fp16 x [+smooth]                    (== b2_smooth_quant)
 *         3 = fp32 y [M,K] with row absmax x_amax [M], as written by an out_kind = 1 call
 *             (== the quantise step of b2_act_mul_quant)
 * out_kind: 0 = out fp16 [M,N] = dequantised result + bias + residual   (== b2_gemm_w8a8)
 *           1 = W is the [gate; up] matrix (N = 2F rows): out is fp32 y [M,F] =
 *               act(fp16(gate)) * fp16(up) * out_smooth[f] and out_amax[m] = max(out_amax[m], |y|)
 *               (atomic; the caller -- or an earlier call through zero_amax -- zeroes it first).
 *               act: 0 silu, 1 gelu_new.  residual must be NULL.
 * zero_amax: optional [M] floats this launch sets to 0 (so a chain of calls needs no memset).
 * prefetch0/1 (+_bytes): optional device ranges (NULL/0 = none) -- typically the first ~24 MB of the
 * weight matrices the NEXT launches will stream; the kernel issues L2 bulk prefetches for them at
 * its start so the successor's ramp-up hits L2 and HBM stays busy across the launch boundary.
 * x_ld: row stride of x in elements.  K % 16 == 0.  M > 8 returns B2_ERR_UNSUPPORTED (use the
 * unfused entry points, which run the tensor-core schedule).
 * Replaces: the same reference calls as b2_norm_quant + b2_gemm_w8a8 + b2_act_mul_quant, inside the
 * decode loop of `model.generate` (models_quant_updated.py:84-93). */
int b2_fused_linear(const void* x, long long x_ld, int x_kind, const void* gamma, const void* beta, const float* smooth,
                    const float* x_amax, float eps, const int8_t* W, const float* w_scale, const void* bias,
                    const void* residual, void* out, int out_kind, int act, const float* out_smooth, float* out_amax,
                    float* zero_amax, int M, int N, int K, const void* prefetch0, long long prefetch0_bytes,
                    const void* prefetch1, long long prefetch1_bytes, void* stream);

/* Decode-time (M <= 8) qkv projection with everything around it fused: activation prologue
 * (x_kind 0/1/2 as in b2_fused_linear) -> int8 quantise -> W8A8 GEMV over W = [q;k;v] ->
 * dequantise + bias -> fp16 -> rotary embedding of q and k -> rotated q written to qkv_out
 * [M,(H+2KV)*hd] (only the q columns are written), rotated k and v appended to the caches at
 * (row seq_ids[m] or seq0, positions[m]).  Bit-identical to b2_fused_linear + b2_rope_kv_append.
 * Replaces: q/k/v_proj + apply_rotary_pos_emb + DynamicCache.update (modeling_llama.py:251-289,
 * :128-153) inside the decode loop of `generate` (models_quant_updated.py:84-93). */
int b2_fused_qkv(const void* x, long long x_ld, int x_kind, const void* gamma, const void* beta, const float* smooth,
                 float eps, const int8_t* W, const float* w_scale, const void* bias, void* qkv_out,
                 const float* cos_table, const float* sin_table, const int* positions, const int* seq_ids, int seq0,
                 void* kcache, void* vcache, int M, int H, int KV, int hd, int Smax, int K, int use_rope,
                 const void* prefetch0, long long prefetch0_bytes, const void* prefetch1, long long prefetch1_bytes,
                 void* stream);

/* One fused linear of a decode-time chain: the arguments of b2_fused_linear / b2_fused_qkv as a plain
 * struct (unused fields zero / NULL).  out_kind 2 = qkv projection with rotary + KV append
 * (b2_fused_qkv semantics: N = (H+2KV)*hd, `out` = the qkv buffer). */
typedef struct b2_fused_desc {
  const void* x; long long x_ld; int x_kind;
  const void* gamma; const void* beta; const float* smooth; const float* x_amax; float eps;
  const int8_t* W; const float* w_scale; const void* bias; const void* residual;
  void* out; int out_kind; int act; const float* out_smooth; float* out_amax; float* zero_amax;
  int N, K;
  const float* cos_table; const float* sin_table; const int* positions; const int* seq_ids; int seq0;
  void* kcache; void* vcache; int H, KV, hd, Smax, use_rope;
  const void* prefetch0; long long prefetch0_bytes; const void* prefetch1; long long prefetch1_bytes;
} b2_fused_desc;

/* The dependent tail of a gated decoder layer for M <= 8 tokens in ONE persistent launch:
 *   o_proj (x_kind 2, out_kind 0) -> gate_up (x_kind 0|1, out_kind 1) -> down (x_kind 3, out_kind 0)
 *   [-> next_qkv (x_kind 0|1, out_kind 2), may be NULL]
 * Each phase consumes the previous phase's output; phases are separated by grid-wide barriers and
 * every phase starts streaming its weights before the barrier that precedes it.  Results are
 * bit-identical to the same linears issued one by one.
 * counters: device, 4 x uint32, ALL ZERO on entry (three barrier counters + an error flag; the
 * caller zeroes them again before reusing them; counters[3] != 0 afterwards means a barrier timed
 * out, i.e. the launch was not fully co-resident).  The launch sizes its grid from the occupancy
 * so that all CTAs are resident at once; do not run it concurrently with kernels that would take
 * whole SMs away from it.
 * Replaces: the same reference calls as b2_fused_linear / b2_fused_qkv (decode loop of
 * `model.generate`, models_quant_updated.py:84-93). */
int b2_fused_layer_tail(const b2_fused_desc* o_proj, const b2_fused_desc* gate_up, const b2_fused_desc* down,
                        const b2_fused_desc* next_qkv, int M, unsigned int* counters, void* stream);

/* Rotary embedding of q and k (in fp32 from a caller-supplied cos/sin table [max_pos, hd/2] fp32)
 * and append of rotated k and v to the KV cache.  qkv [T, ld] fp16: q heads at columns
 * [0,H*hd), k at [H*hd,(H+KV)*hd), v after; q is rotated in place.  positions [T] int32;
 * seq_ids [T] int32 cache row per token, or NULL for row `seq0`.  Caches: [rows][KV][Smax][hd] fp16.
 * use_rope = 0 skips the rotation (GPT-2).
 * Replaces: apply_rotary_pos_emb + DynamicCache.update (modeling_llama.py:128-153, :251-289). */
int b2_rope_kv_append(void* qkv, long long ld, const float* cos_table, const float* sin_table, const int* positions,
                      const int* seq_ids, int seq0, void* kcache, void* vcache, int T, int H, int KV, int hd,
                      int Smax, int use_rope, void* stream);

/* Decode attention (one query token per sequence) over the KV cache.
 * q: token b's heads at q + b*q_ld (fp16, H*hd contiguous); seqlens [B] int32 = keys valid per
 * sequence INCLUDING the token just appended; rows [B] int32 cache row per sequence or NULL (= b);
 * out: token b at out + b*o_ld.  window = 0 for full attention, else sliding window size.
 * max_seqlen: host-side upper bound of seqlens (chooses the kv split count without a device read).
 * workspace: scratch for the split-KV partials, at least b2_decode_attn_workspace_bytes(...) bytes;
 * may be NULL/short, in which case fewer splits are used.
 * prefetch0/1 (+ _bytes): optional L2 hints, as in b2_fused_linear: weights that LATER launches will
 * stream.  Decode attention leaves about half of the HBM bandwidth idle, so this is where the
 * o_proj / gate_up weights of the same layer are best pulled into L2.
 * Replaces: eager_attention_forward at q_len = 1 (modeling_llama.py:199-222) inside the decode
 * loop of `model.generate` (models_quant_updated.py:84-93). */
int b2_decode_attn(const void* q, long long q_ld, const void* kcache, const void* vcache, const int* seqlens,
                   const int* rows, void* out, long long o_ld, void* workspace, size_t workspace_bytes, int B, int H,
                   int KV, int hd, int Smax, int max_seqlen, float scale, int window, const void* prefetch0,
                   long long prefetch0_bytes, const void* prefetch1, long long prefetch1_bytes, void* stream);
/* Bytes of workspace b2_decode_attn wants for these sizes (0 if it will not split). */
size_t b2_decode_attn_workspace_bytes(int B, int H, int KV, int hd, int max_seqlen, int window);

/* b2_decode_attn for callers that cannot allocate device memory.  (No counterpart in the
 * reference: it exists for hosts of this library, not to replace any reference routine.)
 * Who needs it: a server that handed all of HBM to its KV-cache pool at start-up and runs with the
 * allocator closed; a latency-critical decode thread in which cudaMalloc -- a device-wide
 * synchronisation with no time bound -- is forbidden; a plugin host that owns device memory and
 * does not let plugins allocate.  Such callers have no workspace to pass, so this variant keeps the
 * split-KV partials in a fixed B2_DECODE_ATTN_STATIC_SCRATCH_BYTES (512 KiB) array inside the library image
 * (every GPU that loads the library holds those bytes until the process exits).  The call itself
 * allocates nothing: no cudaMalloc/cudaFree, no allocating library routine.
 * Arguments and result are those of b2_decode_attn without workspace/workspace_bytes; the output is
 * the same as b2_decode_attn's with a full-size workspace.  It is the slower of the two: it always writes the
 * partials to the array and merges them in a second launch, also for up to 8 splits, where b2_decode_attn merges
 * inside a thread-block cluster without touching global memory.
 * - Sizes whose partials (b2_decode_attn_workspace_bytes) exceed the array are refused with
 *   B2_ERR_TOO_LARGE; nothing is launched and the split count is never silently reduced.
 * - The first call on each device looks up the array's device address (cudaGetSymbolAddress, with
 *   that device current).  Make that first call before starting any stream capture.
 * - All calls share the one array and it keeps the previous call's contents.  Nothing in it needs
 *   to be zero on entry (each call writes every slot it later reads), but two calls whose kernels
 *   can overlap on the GPU -- different streams, or concurrent host threads on one device -- race
 *   on it and corrupt each other.  Serialise them (one stream, or events) or use b2_decode_attn. */
#define B2_DECODE_ATTN_STATIC_SCRATCH_BYTES (1u << 19)
int b2_decode_attn_static(const void* q, long long q_ld, const void* kcache, const void* vcache, const int* seqlens,
                          const int* rows, void* out, long long o_ld, int B, int H, int KV, int hd, int Smax,
                          int max_seqlen, float scale, int window, void* stream);

/* Causal (optionally sliding-window) attention for one contiguous chunk of T query tokens of one
 * sequence at positions [pos0, pos0+T), keys 0..pos0+T-1 read from cache row `row` (the chunk's
 * own k/v must already be appended).  positions: device int32 [T] or NULL; when given, pos0 is
 * read from positions[0] on the device (lets a captured CUDA graph serve every chunk) and the host
 * `pos0` argument is ignored.
 * Replaces: eager_attention_forward at q_len > 1 (modeling_llama.py:199-222): the prefill forward
 * of `generate` and the `model(**inputs).logits` pass of confidence_score
 * (models_quant_updated.py:63-70). */
int b2_prefill_attn(const void* q, long long q_ld, const void* kcache, const void* vcache, void* out, long long o_ld,
                    const int* positions, int T, int pos0, int row, int H, int KV, int hd, int Smax, float scale, int window, void* stream);
/* ---- LLM.int8() mixed decomposition -------------------------------------
 * What `BitsAndBytesConfig(load_in_8bit=True)` (models_quant_updated.py:30-32, combiner_quant_updated.py:37;
 * bitsandbytes Linear8bitLt, threshold 6.0) makes of one nn.Linear, as three calls around b2_gemm_w8a8:
 *   b2_outlier_scan(y)                    which activation columns hold a value of magnitude >= threshold
 *   b2_smooth_quant(y, smooth = mask)     per-row absmax int8 of the other columns
 *   b2_gemm_w8a8(...)                     int8 product, dequantised (+ bias) to fp16
 *   b2_outlier_side(out, y, W, ...)       out += fp16(y[:,cols] . fp16(W[:,cols]*w_scale)^T); out += residual
 * Arithmetic pinned by oracle/w8a8.py llm_int8_linear. */

/* act(gate)*up (act 0/1) or act(x) (act 2/3/4) as fp16 [T,F], not quantised (same act codes as b2_act_mul_quant). */
int b2_act_mul(const void* gate_up, void* y, int T, int F, int act, void* stream);
/* y [T,K] fp16 (row pitch ld): mask [K] fp32 = 0 for columns with any |y| >= threshold else 1; cols [K] int32 =
 * the outlier columns in ascending order; count [1] int32; flags = K/8 words of scratch.  All device memory. */
int b2_outlier_scan(const void* y, long long ld, int T, int K, float threshold, float* mask, int* cols, int* count,
                    unsigned* flags, void* stream);
/* out [T,N] fp16 in place: + the fp16 side product over the `count` columns in `cols` (skipped when 0), then
 * + residual [T,N] fp16 (or NULL).  W [N,K] int8 with row scales w_scale [N]. */
int b2_outlier_side(void* out, const void* y, long long ld, const int8_t* W, const float* w_scale, const int* cols,
                    const int* count, const void* residual, int T, int N, int K, void* stream);

/* b2_outlier_side for long batches, where most columns qualify as outliers (over a 2048-token chunk nearly every
 * column of the MLP intermediate holds a value >= 6): the side product as a tcgen05 fp16 GEMM over y with its
 * NON-outlier columns zeroed (mask from b2_outlier_scan) and W dequantised to fp16 per call.  Same result up
 * to the fp32 summation order.  Caller scratch: ysub [T,K] fp16, wdeq [N,K] fp16, side [T,N] fp32.  K % 16 == 0. */
int b2_outlier_side_dense(void* out, const void* y, long long ld, const int8_t* W, const float* w_scale, const float* mask,
                          const void* residual, void* ysub, void* wdeq, float* side, int T, int N, int K, void* stream);

/* Test hook (no reference counterpart): which kernel b2_prefill_attn uses from now on, process-wide.
 * 0 = choose (the tcgen05/TMEM kernel for head_dim 128 and T >= 64, with or without a sliding window, else the
 * mma.sync kernel), 1 = always the mma.sync kernel, 2 = the tcgen05 kernel or B2_ERR_UNSUPPORTED. */
int b2_set_prefill_attn_impl(int impl);

/* Token (+ learned position) embedding gather.  ids [T] int64; positions [T] int32 (only read
 * when pos_table != NULL); tables fp16 [vocab,d] / [max_pos,d]; out [T,d] fp16.
 * Replaces: `embed_tokens` (and GPT-2 `wpe`) at the head of `model(**inputs)` (models_quant_updated.py:66). */
int b2_embed(const long long* ids, const int* positions, const void* table, const void* pos_table, void* out, int T,
             int d, int vocab, void* stream);

/* Unquantised lm_head: logits [M,V] fp32 = x [M,d] fp16 . W [V,d]^T fp16 (fp32 accumulate).
 * Any V (GPT-2's 50257 included; d % 8 == 0): streaming GEMV for M <= 4, the tcgen05 fp16 tile above.
 * Replaces: `lm_head` (left unquantised by load_in_8bit; models_quant_updated.py:66-67 `.logits`). */
int b2_lm_head(const void* x, const void* W, float* logits, int M, int V, int d, void* stream);

/* logits [T, ld >= V] fp32 += bias [V] fp32, in place.
 * Replaces: the bias of Phi's lm_head (`nn.Linear(hidden, vocab, bias=True)`, transformers models/phi/modeling_phi.py),
 * the first answerer the reference loads (Code/Quantised Models/combiner_quant_updated.py:177). */
int b2_bias_add_f32(float* logits, const float* bias, long long ld, int T, int V, void* stream);

/* Row-wise argmax of fp32 logits [T,V] -> int64 [T]; ties to the lowest index.
 * Replaces: the greedy (do_sample=False) token selection of `generate`. */
int b2_argmax(const float* logits, long long* out, int T, int V, void* stream);

/* Stage-to-stage hand-off of a decode step through NVLink peer memory (one kernel at each end, no host, no NCCL).
 * b2_p2p_send: copy `bytes` (a multiple of 16, 16-byte aligned) from local `src` to `dst_peer` -- a pointer into the
 * receiving GPU's memory mapped into this process (CUDA IPC / symmetric memory; the caller owns the mapping) --
 * then publish `seq` at `flag_peer` (also in the receiver's memory) with release semantics at system scope.
 * b2_p2p_wait: on the receiving GPU, wait until *flag has reached `seq` (wrap-safe), then copy the message from the
 * local inbox `src` to `dst` (dst NULL: no copy, later kernels read the inbox in place).  The wait is bounded (about
 * 30 s); on expiry it sets *err (device, may be NULL) to 1 instead of hanging.  Sequence numbers are the caller's:
 * one counter per (edge, slot), incremented identically at both ends.  A slot must not be re-sent before the
 * receiver has consumed it (in the layer-range pipeline the token's round trip guarantees that).
 * Replaces: the device-to-device hop the reference sketches over gRPC (Code/gRPC/server.py:13-19, client.py:7-11). */
int b2_p2p_send(const void* src, void* dst_peer, long long bytes, unsigned int* flag_peer, unsigned int seq, void* stream);
int b2_p2p_wait(const unsigned int* flag, unsigned int seq, const void* src, void* dst, long long bytes, unsigned int* err,
                void* stream);

/* Greedy selection over one vocabulary shard (the lm_head split between two pipeline stages): logits [T, ld >= V]
 * fp32 hold columns [idx_offset, idx_offset + V) of the full logits; out_idx [T] = idx_offset + row argmax (ties to
 * the lowest index), out_val [T] fp32 = the row maximum (or NULL).  With other_val / other_idx ([T] each, both or
 * neither) the row's winner is merged with another shard's (value, global index) under the same tie rule, so two
 * shards reproduce b2_argmax over the whole row exactly.
 * Replaces: the greedy token selection of `generate`, as b2_argmax. */
int b2_argmax_shard(const float* logits, long long ld, long long* out_idx, float* out_val, long long idx_offset,
                    const float* other_val, const long long* other_idx, int T, int V, void* stream);

/* lm_head and greedy selection in ONE launch, for decode steps of M <= 4 rows (more: B2_ERR_UNSUPPORTED, use
 * b2_lm_head + b2_argmax_shard): the dot products of b2_lm_head (same values) feed a running winner instead of a
 * logits round trip through HBM and a second launch.  x [M,d] fp16 (the normed hidden state), W [V,d] fp16 = rows
 * [idx_offset, idx_offset + V) of the head; out_idx [M] = the winner's global index under b2_argmax_shard's rule
 * (largest value, lowest index among equals), merged with another shard's winner other_val / other_idx ([M] each,
 * both or neither); out_idx2 (or NULL) receives a second copy (the token history beside the next step's input);
 * out_val [M] (or NULL) the winning value; logits [M,V] fp32 (or NULL) the logits themselves.
 * scratch: b2_lm_head_argmax_scratch_bytes() of device memory, zeroed ONCE by the caller; every launch leaves it
 * zeroed.  One scratch per stream of launches (launches sharing it must be stream-ordered).
 * Pipeline hand-off inside the launch (each optional, NULL = off): with wait_flag the kernel first waits (bounded,
 * about 30 s, then *err = 1) until this GPU's word *wait_flag has reached wait_seq -- x is a message another stage
 * stores into this GPU's memory with b2_p2p_send / b2_p2p_send_multi; out_idx / out_val may point into another
 * GPU's memory, and with notify_flag the kernel publishes notify_seq there after storing them (the receiver:
 * b2_merge_winners).  Plain loads and stores only.
 * Replaces: the logits + greedy token selection at the end of every `generate` step
 * (Code/Quantised Models/models_quant_updated.py:84-93). */
long long b2_lm_head_argmax_scratch_bytes(void);
int b2_lm_head_argmax(const void* x, const void* W, float* logits, int M, int V, int d, long long idx_offset,
                      const float* other_val, const long long* other_idx, long long* out_idx, long long* out_idx2,
                      float* out_val, void* scratch, const unsigned int* wait_flag, unsigned int wait_seq,
                      unsigned int* err, unsigned int* notify_flag, unsigned int notify_seq, void* stream);

/* The lm_head shared by ALL stages of a pipeline (each holds V / stages rows): the last stage sends its normed hidden
 * state to every stage at once -- b2_p2p_send_multi: CTA i stores `bytes` (multiple of 16) of src to dsts[i] and
 * publishes seq at flags[i]; dsts / flags are HOST arrays of n <= 16 device pointers (peer-mapped or local) -- every
 * stage runs b2_lm_head_argmax on its slice with the first stage's inbox as output, and the first stage picks the
 * token -- b2_merge_winners: flags [W] and records [W] x 64 bytes (indices [4] int64 at +0, values [4] fp32 at +32)
 * live in this GPU's memory; waits (bounded, then *err = 1) until every flag has reached seq, then out_idx [B <= 4]
 * (and out_idx2, or NULL) = the global index of the largest value per row, lowest index among equals: exactly
 * b2_argmax over the whole row.
 * Replaces: as b2_p2p_send (Code/gRPC/server.py:13-19) and the greedy selection of `generate`. */
int b2_p2p_send_multi(const void* src, long long bytes, void* const* dsts, unsigned int* const* flags, int n,
                      unsigned int seq, void* stream);
int b2_merge_winners(const unsigned int* flags, unsigned int seq, const void* records, int W, int B, long long* out_idx,
                     long long* out_idx2, unsigned int* err, void* stream);

/* Confidence score of a forward pass: out[0] = mean over the T rows of the largest softmax probability,
 * i.e. mean_t 1 / sum_v exp(logits[t,v] - max_v logits[t,v]).  logits [T, ld >= V] fp32; row_pmax [T] fp32 scratch
 * (receives the per-row value); out [1] fp32.  Exponentials in the pinned fp32 sequence, sums in fp64 in a fixed
 * order (oracle/w8a8.py confidence).
 * Replaces: `torch.softmax(logits, dim=-1).max(dim=-1).values.mean()` of confidence_score
 * (Code/Quantised Models/models_quant_updated.py:63-70). */
int b2_confidence(const float* logits, long long ld, float* row_pmax, float* out, int T, int V, void* stream);

/* Sampling with the reference's real generate kwargs (do_sample=True; temperature, top_k, top_p,
 * repetition_penalty: Code/Quantised Models/models_quant_updated.py:84-93, Code/others/config.yaml:3-7).
 * logits [B, ld >= V] fp32; history [B, hist_ld] int64 = every token id seen so far per row (prompt +
 * generated), hist_len [B] int32 valid entries per row (history NULL = no repetition penalty);
 * out [B] int64.  Processor chain and order are HF's (repetition penalty -> temperature -> top-k with
 * ties kept -> top-p on the ascending cumulative softmax, at least one token kept); the draw is an
 * inverse CDF over the kept tokens with a counter-based uniform of (seed, step, row), step read from
 * *step_ptr when step_ptr != NULL (device) else `step`.  top_k <= 0 or > 1024 is capped at 1024
 * candidates.  Replaces: the sampling branch of `model.generate` that the reference always takes. */
int b2_sample(const float* logits, long long ld, const long long* history, const int* hist_len, long long hist_ld,
              long long* out, int B, int V, float temperature, int top_k, float top_p, float repetition_penalty,
              unsigned long long seed, const unsigned long long* step_ptr, unsigned long long step, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200STAGE_H */
