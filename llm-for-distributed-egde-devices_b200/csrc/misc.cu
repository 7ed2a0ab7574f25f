// Embedding gather (+ learned positions for GPT-2) and greedy argmax over the logits.
// Reference call sites: token embedding inside HF `model(**inputs)` / `generate`
// (Code/Quantised Models/models_quant_updated.py:66-67, :84-93); greedy selection is the
// `do_sample=False` subset of `generate` that the build's Pipeline exposes.
#include "common.cuh"

namespace b2 {

__global__ void __launch_bounds__(256)
embed_kernel(const long long* __restrict__ ids, const int* __restrict__ positions, const __half* __restrict__ table,
             const __half* __restrict__ pos_table, __half* __restrict__ out, int d, int vocab) {
  const int t = blockIdx.x;
  long long id = ids[t];
  id = id < 0 ? 0 : (id >= vocab ? vocab - 1 : id);
  const __half* src = table + id * d;
  const __half* ps = pos_table ? pos_table + (long long)positions[t] * d : nullptr;
  __half* dst = out + (long long)t * d;
  for (int c = threadIdx.x; c < (d >> 3); c += blockDim.x) {
    uint4 u = ld_v4(src + c * 8);
    if (ps) {   // fp32 add, one rounding (oracle: stage_forward)
      const uint4 q = ld_v4(ps + c * 8);
      const uint32_t* a = &u.x;
      const uint32_t* b = &q.x;
      uint32_t o[4];
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const float2 x = h2_to_f2(a[e]), y = h2_to_f2(b[e]);
        o[e] = f2_to_h2(add_rn(x.x, y.x), add_rn(x.y, y.y));
      }
      u = make_uint4(o[0], o[1], o[2], o[3]);
    }
    st_v4(dst + c * 8, u);
  }
}

int embed(const long long* ids, const int* positions, const void* table, const void* pos_table, void* out, int T, int d,
          int vocab, cudaStream_t st) {
  if (T == 0) return OK;
  if (!ids || !table || !out || (pos_table && !positions)) return ERR_NULL_PTR;
  if (T < 0 || d <= 0 || (d & 7) || vocab <= 0) return ERR_BAD_SHAPE;
  embed_kernel<<<T, 256, 0, st>>>(ids, positions, (const __half*)table, (const __half*)pos_table, (__half*)out, d, vocab);
  return launch_status();
}

// one CTA per row; ties resolve to the lowest index (torch.argmax / HF greedy convention)
__global__ void __launch_bounds__(1024)
argmax_kernel(const float* __restrict__ logits, long long ld, long long* __restrict__ out, int V,
              float* __restrict__ out_val, long long idx_offset, const float* __restrict__ other_val,
              const long long* __restrict__ other_idx) {
  __shared__ float sv[32];
  __shared__ int si[32];
  const float* row = logits + (long long)blockIdx.x * ld;
  float best = -INFINITY;
  int bi = 0x7fffffff;
  auto take = [&](float v, int i) { if (v > best || (v == best && i < bi)) { best = v; bi = i; } };
  // 128-bit loads, four in flight per thread (the row is 128 KB: latency-, not bandwidth-bound)
  const bool vec = ((reinterpret_cast<uintptr_t>(row) & 15) == 0);
  const int V4 = vec ? (V >> 2) : 0;
  const float4* row4 = reinterpret_cast<const float4*>(row);
  for (int i0 = threadIdx.x; i0 < V4; i0 += 4 * blockDim.x) {
    float4 q[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int i = i0 + u * blockDim.x;
      q[u] = i < V4 ? __ldg(row4 + i) : make_float4(-INFINITY, -INFINITY, -INFINITY, -INFINITY);
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int i = (i0 + u * blockDim.x) * 4;
      take(q[u].x, i); take(q[u].y, i + 1); take(q[u].z, i + 2); take(q[u].w, i + 3);
    }
  }
  for (int i = V4 * 4 + threadIdx.x; i < V; i += blockDim.x) take(row[i], i);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float ov = __shfl_xor_sync(0xffffffffu, best, o);
    const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
    if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
  }
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  if (l == 0) { sv[w] = best; si[w] = bi; }
  __syncthreads();
  if (w == 0) {
    best = l < (blockDim.x >> 5) ? sv[l] : -INFINITY;
    bi = l < (blockDim.x >> 5) ? si[l] : 0x7fffffff;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float ov = __shfl_xor_sync(0xffffffffu, best, o);
      const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
      if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
    }
    if (l == 0) {
      long long idx = (bi == 0x7fffffff ? 0 : bi) + idx_offset;
      if (other_val != nullptr) {          // merge with the winner of another vocabulary shard: same tie rule
        const float ov = other_val[blockIdx.x];
        const long long oi = other_idx[blockIdx.x];
        if (ov > best || (ov == best && oi < idx)) { best = ov; idx = oi; }
      }
      out[blockIdx.x] = idx;
      if (out_val != nullptr) out_val[blockIdx.x] = best;
    }
  }
}

int argmax_rows(const float* logits, long long* out, int T, int V, cudaStream_t st) {
  if (T == 0) return OK;
  if (!logits || !out) return ERR_NULL_PTR;
  if (T < 0 || V <= 0) return ERR_BAD_SHAPE;
  argmax_kernel<<<T, 1024, 0, st>>>(logits, V, out, V, nullptr, 0, nullptr, nullptr);
  return launch_status();
}

// argmax over one vocabulary shard: out_idx = idx_offset + argmax (ties to the lowest index), out_val = the
// maximum (optional), and optionally merged with another shard's (value, global index) winner per row.
int argmax_shard(const float* logits, long long ld, long long* out_idx, float* out_val, long long idx_offset,
                 const float* other_val, const long long* other_idx, int T, int V, cudaStream_t st) {
  if (T == 0) return OK;
  if (!logits || !out_idx || ((other_val == nullptr) != (other_idx == nullptr))) return ERR_NULL_PTR;
  if (T < 0 || V <= 0 || ld < V) return ERR_BAD_SHAPE;
  argmax_kernel<<<T, 1024, 0, st>>>(logits, ld, out_idx, V, out_val, idx_offset, other_val, other_idx);
  return launch_status();
}

// ---------------------------------------------------------------------------------------------
// confidence score: mean over rows of the largest softmax probability = mean_t 1 / sum_v exp(x[t,v] - max_v x[t,v])
// (the reference's confidence_score: softmax -> max over the vocabulary -> mean over positions,
// Code/Quantised Models/models_quant_updated.py:63-70).  One CTA per row: row maximum (exact), then the
// exponentials in the pinned fp32 sequence (exp_det) summed in fp64; a second one-CTA launch averages the rows in
// fp64.  Both reductions run in a fixed order, so the result is reproducible run to run.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(512)
row_pmax_kernel(const float* __restrict__ logits, long long ld, float* __restrict__ pmax, int V) {
  __shared__ float smax[16];
  __shared__ double ssum[16];
  const float* row = logits + (long long)blockIdx.x * ld;
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  float m = -INFINITY;
  for (int i = threadIdx.x; i < V; i += blockDim.x) m = fmaxf(m, row[i]);
  m = warp_max(m);
  if (l == 0) smax[w] = m;
  __syncthreads();
  m = smax[0];
#pragma unroll
  for (int i = 1; i < 16; ++i) m = fmaxf(m, smax[i]);
  double s = 0.0;
  for (int i = threadIdx.x; i < V; i += blockDim.x) s += (double)exp_det(sub_rn(row[i], m));
  s = warp_sum_d(s);
  if (l == 0) ssum[w] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0;
#pragma unroll
    for (int i = 0; i < 16; ++i) t += ssum[i];
    pmax[blockIdx.x] = (float)(1.0 / t);
  }
}
__global__ void __launch_bounds__(512) mean_rows_kernel(const float* __restrict__ v, float* __restrict__ out, int T) {
  __shared__ double ssum[16];
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  double s = 0.0;
  for (int i = threadIdx.x; i < T; i += blockDim.x) s += (double)v[i];
  s = warp_sum_d(s);
  if (l == 0) ssum[w] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0;
#pragma unroll
    for (int i = 0; i < 16; ++i) t += ssum[i];
    out[0] = (float)(t / (double)T);
  }
}

int confidence(const float* logits, long long ld, float* row_pmax, float* out, int T, int V, cudaStream_t st) {
  if (!logits || !row_pmax || !out) return ERR_NULL_PTR;
  if (T <= 0 || V <= 0 || ld < V) return ERR_BAD_SHAPE;
  row_pmax_kernel<<<T, 512, 0, st>>>(logits, ld, row_pmax, V);
  mean_rows_kernel<<<1, 512, 0, st>>>(row_pmax, out, T);
  return launch_status();
}

// logits[t, v] += bias[v] (fp32): Phi's lm_head carries a bias (modeling_phi.py: lm_head = nn.Linear(..., bias=True))
__global__ void __launch_bounds__(256) bias_add_f32_kernel(float* __restrict__ x, const float* __restrict__ b, long long ld, int V) {
  float* row = x + (long long)blockIdx.y * ld;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < V; i += gridDim.x * blockDim.x) row[i] = add_rn(row[i], b[i]);
}
int bias_add_f32(float* x, const float* bias, long long ld, int T, int V, cudaStream_t st) {
  if (T == 0) return OK;
  if (!x || !bias) return ERR_NULL_PTR;
  if (T < 0 || V <= 0 || ld < V) return ERR_BAD_SHAPE;
  bias_add_f32_kernel<<<dim3((V + 2047) / 2048, T), 256, 0, st>>>(x, bias, ld, V);
  return launch_status();
}

int init_misc() {
  cudaFuncAttributes fa;
  bool ok = cudaFuncGetAttributes(&fa, embed_kernel) == cudaSuccess;
  ok &= cudaFuncGetAttributes(&fa, argmax_kernel) == cudaSuccess;
  ok &= cudaFuncGetAttributes(&fa, row_pmax_kernel) == cudaSuccess;
  ok &= cudaFuncGetAttributes(&fa, bias_add_f32_kernel) == cudaSuccess;
  ok &= cudaFuncGetAttributes(&fa, mean_rows_kernel) == cudaSuccess;
  return ok ? OK : ERR_ARCH;
}

}  // namespace b2
