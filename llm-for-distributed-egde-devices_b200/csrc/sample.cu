// On-GPU sampler for the reference's real `generate` kwargs: do_sample=True with temperature, top_k,
// top_p and repetition_penalty (Code/Quantised Models/models_quant_updated.py:84-93; defaults in
// Code/others/config.yaml:3-7; the reference never decodes greedily).
//
// Follows HF's processor chain in HF's order (transformers/generation/utils.py:1089,1214,1219,1223;
// logits_process.py :371-392 repetition penalty, :297-299 temperature, :581-586 top-k with ties kept,
// :521-533 top-p on the ascending cumulative softmax).  torch.multinomial's generator cannot be
// reproduced on a device, so the draw itself is build-defined: inverse CDF over the kept tokens in
// (score desc, id asc) order with a counter-based uniform (splitmix64 of seed, step, row).
// Oracle: oracle/w8a8.py sample_filter / sample_det -- same fp32 operations in the same order, so the
// sampled ids are bit-for-bit reproducible (tests/test_sample_gpu.py).
//
// One CTA per row: history bitmap in shared memory -> 4-pass radix select of the k-th largest
// processed score -> collect the >= threshold candidates -> bitonic sort -> one thread walks the
// (<= 1024) sorted candidates for top-p and the draw.  The vocabulary row (128 KB at V = 32 000) is
// re-read from L2 by the five passes; the kernel is latency-, not bandwidth-bound.
#include "common.cuh"

namespace b2 {

constexpr int SMP_THREADS = 1024;
constexpr int SMP_CAP = 1024;      // max candidates after top-k (k plus ties); also the top_k ceiling

struct SampleParams {
  const float* logits; long long ld;
  const long long* history; const int* hist_len; long long hist_ld;
  long long* out;
  int V, top_k;
  float temperature, top_p, penalty;
  unsigned long long seed;
  const unsigned long long* step_ptr; unsigned long long step;
};

B2_DEVINL uint32_t f2key(float x) {            // order-preserving float -> uint (NaN sorts lowest)
  if (x != x) return 0u;
  const uint32_t u = __float_as_uint(x);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
B2_DEVINL float key2f(uint32_t k) {
  const uint32_t u = (k & 0x80000000u) ? (k & 0x7fffffffu) : ~k;
  return __uint_as_float(u);
}

__global__ void __launch_bounds__(SMP_THREADS) sample_kernel(const SampleParams p) {
  extern __shared__ uint32_t bitmap[];                       // V bits: token occurs in the history
  __shared__ unsigned long long cand[SMP_CAP];               // (key << 32) | (0xffffffff - id)
  __shared__ float ex[SMP_CAP];
  __shared__ int hist[256];
  __shared__ uint32_t s_prefix, s_remaining;
  __shared__ int s_count;
  const int b = blockIdx.x, tid = threadIdx.x;
  const float* row = p.logits + (long long)b * p.ld;
  const int V = p.V;
  const int words = (V + 31) >> 5;
  for (int i = tid; i < words; i += SMP_THREADS) bitmap[i] = 0u;
  for (int i = tid; i < SMP_CAP; i += SMP_THREADS) cand[i] = 0ull;
  if (tid == 0) { s_prefix = 0u; s_remaining = (uint32_t)min(p.top_k, V); s_count = 0; }
  __syncthreads();
  const bool pen_on = p.history != nullptr && p.penalty != 1.0f;
  if (pen_on) {
    const int n = p.hist_len ? p.hist_len[b] : 0;
    const long long* h = p.history + (long long)b * p.hist_ld;
    for (int i = tid; i < n; i += SMP_THREADS) {
      const long long id = h[i];
      if (id >= 0 && id < V) atomicOr(&bitmap[id >> 5], 1u << (id & 31));
    }
  }
  __syncthreads();
  auto score_key = [&](int i) -> uint32_t {
    float x = row[i];
    if (pen_on && ((bitmap[i >> 5] >> (i & 31)) & 1u)) x = (x < 0.0f) ? mul_rn(x, p.penalty) : div_rn(x, p.penalty);
    x = add_rn(div_rn(x, p.temperature), 0.0f);     // -0 -> +0 so equal scores have equal keys
    return f2key(x);
  };
  // ---- radix select: key of the k-th largest processed score ------------------------------------
  uint32_t mask = 0u;
#pragma unroll 1
  for (int shift = 24; shift >= 0; shift -= 8) {
    for (int i = tid; i < 256; i += SMP_THREADS) hist[i] = 0;
    __syncthreads();
    const uint32_t prefix = s_prefix;
    for (int i = tid; i < V; i += SMP_THREADS) {
      const uint32_t k = score_key(i);
      if ((k & mask) == prefix) atomicAdd(&hist[(k >> shift) & 255], 1);
    }
    __syncthreads();
    if (tid == 0) {
      uint32_t rem = s_remaining;
      int d = 255;
      for (; d > 0; --d) {
        const uint32_t c = (uint32_t)hist[d];
        if (c >= rem) break;
        rem -= c;
      }
      s_prefix = prefix | ((uint32_t)d << shift);
      s_remaining = rem;
    }
    mask |= 255u << shift;
    __syncthreads();
  }
  const uint32_t kth = s_prefix;
  // ---- collect every token whose score is >= the k-th largest (HF keeps ties) -------------------
  for (int i = tid; i < V; i += SMP_THREADS) {
    const uint32_t k = score_key(i);
    if (k >= kth) {
      const int slot = atomicAdd(&s_count, 1);
      if (slot < SMP_CAP) cand[slot] = ((unsigned long long)k << 32) | (unsigned long long)(0xffffffffu - (uint32_t)i);
    }
  }
  __syncthreads();
  // ---- bitonic sort, descending: score desc, then id asc ------------------------------------------
  for (int size = 2; size <= SMP_CAP; size <<= 1) {
    for (int stride = size >> 1; stride > 0; stride >>= 1) {
      const int partner = tid ^ stride;
      if (partner > tid) {
        const unsigned long long a = cand[tid], c = cand[partner];
        const bool desc = (tid & size) == 0;
        if (desc ? (a < c) : (a > c)) { cand[tid] = c; cand[partner] = a; }
      }
      __syncthreads();
    }
  }
  // ---- one thread: softmax over the candidates, top-p, draw -------------------------------------
  if (tid == 0) {
    const int n = min(s_count, SMP_CAP);
    const float s0 = key2f((uint32_t)(cand[0] >> 32));
    for (int j = 0; j < n; ++j) ex[j] = exp_det(sub_rn(key2f((uint32_t)(cand[j] >> 32)), s0));
    float z = 0.0f;
    for (int j = n - 1; j >= 0; --j) z = add_rn(z, ex[j]);            // ascending-score order
    const float thr = sub_rn(1.0f, p.top_p);
    float cum = 0.0f;
    int n_keep = n;
    for (int j = n - 1; j >= 1; --j) {
      cum = add_rn(cum, div_rn(ex[j], z));
      if (cum <= thr) n_keep = j; else break;                          // cum is monotone: the rest is kept
    }
    float z2 = 0.0f;
    for (int j = 0; j < n_keep; ++j) z2 = add_rn(z2, ex[j]);
    // counter-based uniform: splitmix64 finaliser of (seed, step, row)
    const unsigned long long step = p.step_ptr ? *p.step_ptr : p.step;
    unsigned long long x = p.seed ^ (step * 0xD1B54A32D192ED03ull) ^ ((unsigned long long)(b + 1) * 0x9E3779B97F4A7C15ull);
    x ^= x >> 30; x *= 0xBF58476D1CE4E5B9ull;
    x ^= x >> 27; x *= 0x94D049BB133111EBull;
    x ^= x >> 31;
    const float u = (float)(x >> 40) * (1.0f / 16777216.0f);
    const float target = mul_rn(u, z2);
    float acc = 0.0f;
    int pick = n_keep - 1;
    for (int j = 0; j < n_keep; ++j) {
      acc = add_rn(acc, ex[j]);
      if (acc > target) { pick = j; break; }
    }
    p.out[b] = (long long)(0xffffffffu - (uint32_t)(cand[pick] & 0xffffffffull));
  }
}

int sample(const float* logits, long long ld, const long long* history, const int* hist_len, long long hist_ld,
           long long* out, int B, int V, float temperature, int top_k, float top_p, float repetition_penalty,
           unsigned long long seed, const unsigned long long* step_ptr, unsigned long long step, cudaStream_t st) {
  if (B == 0) return OK;
  if (!logits || !out) return ERR_NULL_PTR;
  if (B < 0 || V <= 0 || ld < V) return ERR_BAD_SHAPE;
  if (!(temperature > 0.0f) || !(top_p > 0.0f) || top_p > 1.0f || !(repetition_penalty > 0.0f)) return ERR_BAD_SHAPE;
  if (top_k <= 0 || top_k > SMP_CAP) top_k = SMP_CAP;     // top_k = 0 ("off" in HF) is capped at SMP_CAP candidates
  const size_t smem = (size_t)((V + 31) / 32) * 4;
  if (smem > 160 * 1024) return ERR_TOO_LARGE;
  static DeviceOnce attr_once;
  if (!attr_once.get()) {
    if (cudaFuncSetAttribute(sample_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024) != cudaSuccess) return ERR_ARCH;
    attr_once.set();
  }
  SampleParams p;
  p.logits = logits; p.ld = ld; p.history = history; p.hist_len = hist_len; p.hist_ld = hist_ld; p.out = out;
  p.V = V; p.top_k = top_k; p.temperature = temperature; p.top_p = top_p; p.penalty = repetition_penalty;
  p.seed = seed; p.step_ptr = step_ptr; p.step = step;
  sample_kernel<<<B, SMP_THREADS, smem, st>>>(p);
  return launch_status();
}

int init_sample() {
  cudaFuncAttributes fa;
  if (cudaFuncGetAttributes(&fa, sample_kernel) != cudaSuccess) return ERR_ARCH;
  return cudaFuncSetAttribute(sample_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024) == cudaSuccess ? OK : ERR_ARCH;
}

}  // namespace b2
