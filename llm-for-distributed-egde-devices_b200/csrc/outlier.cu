// LLM.int8() mixed decomposition around the W8A8 GEMM.
//
// What the reference's `BitsAndBytesConfig(load_in_8bit=True)` (Code/Quantised Models/
// models_quant_updated.py:30-32, combiner_quant_updated.py:37) makes every nn.Linear do
// (bitsandbytes Linear8bitLt / MatMul8bitLt, threshold 6.0): activation COLUMNS that hold any value
// of magnitude >= threshold are taken out of the int8 product and multiplied in fp16 against the
// dequantised weight columns; the remaining columns are quantised per row (absmax) and go through the
// int8 GEMM.  bitsandbytes is not installable here (no CUDA wheel offline), so the arithmetic is
// pinned by oracle/w8a8.py llm_int8_linear ("parity unpinned").
//
//   b2_outlier_scan : y fp16 [T,K] -> mask fp32 [K] (0 for outlier columns, 1 otherwise: it is passed
//                     as the `smooth` vector of b2_smooth_quant, which then quantises exactly the
//                     sub-matrix), the ascending list of outlier columns and their count (device)
//   b2_outlier_side : out += fp16( y[:,cols] . dequant(W[:,cols])^T ), then out += residual
// Everything is device-resident (the count is read by the kernels), so the chain can be captured in
// a CUDA graph.
#include "common.cuh"

namespace b2 {

constexpr int OS_ROWS = 64;          // rows per CTA in the flag pass

// flags[k/8] bit (k%8) |= any |y[t,k]| >= thr over this CTA's rows
__global__ void __launch_bounds__(128)
outlier_flag_kernel(const __half* __restrict__ y, long long ld, int T, int K, float thr, unsigned* __restrict__ flags) {
  const int ch = blockIdx.x * blockDim.x + threadIdx.x;      // 8-column chunk
  if (ch >= (K >> 3)) return;
  const int t0 = blockIdx.y * OS_ROWS, t1 = min(T, t0 + OS_ROWS);
  unsigned bits = 0u;
  for (int t = t0; t < t1; ++t) {
    const uint4 u = ld_nc_v4(y + (long long)t * ld + ch * 8);
    const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const float2 f = h2_to_f2(w[e]);
      if (fabsf(f.x) >= thr) bits |= 1u << (2 * e);
      if (fabsf(f.y) >= thr) bits |= 1u << (2 * e + 1);
    }
  }
  if (bits) atomicOr(flags + ch, bits);
}

// one CTA: mask[k], ascending cols[], count
__global__ void __launch_bounds__(1024)
outlier_compact_kernel(const unsigned* __restrict__ flags, int K, float* __restrict__ mask, int* __restrict__ cols,
                       int* __restrict__ count) {
  __shared__ int s_warp[32];
  __shared__ int s_base;
  if (threadIdx.x == 0) s_base = 0;
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int k0 = 0; k0 < K; k0 += 1024) {
    const int k = k0 + threadIdx.x;
    const bool out = k < K && ((flags[k >> 3] >> (k & 7)) & 1u);
    if (k < K) mask[k] = out ? 0.0f : 1.0f;
    const unsigned bal = __ballot_sync(0xffffffffu, out);
    if (lane == 0) s_warp[warp] = __popc(bal);
    __syncthreads();
    int before = 0, total = 0;
    for (int w = 0; w < 32; ++w) { const int c = s_warp[w]; if (w < warp) before += c; total += c; }
    if (out) cols[s_base + before + __popc(bal & ((1u << lane) - 1u))] = k;
    __syncthreads();
    if (threadIdx.x == 0) s_base += total;
    __syncthreads();
  }
  if (threadIdx.x == 0) *count = s_base;
}

// out[t,n] = fp16( fp32(out[t,n]) + fp32(fp16( sum_j y[t,cols[j]] * fp16(W[n,cols[j]] * w_scale[n]) )) )  if count > 0
// out[t,n] = fp16( fp32(out[t,n]) + fp32(residual[t,n]) )                                                    if residual
__global__ void __launch_bounds__(256)
outlier_side_kernel(__half* __restrict__ out, const __half* __restrict__ y, long long ld, const int8_t* __restrict__ W,
                    const float* __restrict__ w_scale, const int* __restrict__ cols, const int* __restrict__ count,
                    const __half* __restrict__ residual, int N, int K) {
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  const int t = blockIdx.y;
  if (n >= N) return;
  const int cnt = *count;
  const long long off = (long long)t * N + n;
  float o = __half2float(out[off]);
  if (cnt > 0) {
    const float ws = w_scale[n];
    const __half* yr = y + (long long)t * ld;
    const int8_t* wr = W + (long long)n * K;
    float s = 0.0f;
    for (int j = 0; j < cnt; ++j) {
      const int k = cols[j];
      const float w16 = __half2float(__float2half_rn(mul_rn((float)wr[k], ws)));
      s = add_rn(s, mul_rn(__half2float(yr[k]), w16));
    }
    o = __half2float(__float2half_rn(add_rn(o, __half2float(__float2half_rn(s)))));
  }
  if (residual) o = add_rn(o, __half2float(residual[off]));
  out[off] = __float2half_rn(o);
}

// Decode batches (T <= OS_SMALL_T rows): flags and ordered compaction in ONE launch of one CTA -- a decode step
// runs four of these scans per layer, and each extra launch costs more than the scan itself.
constexpr int OS_SMALL_T = 8;
__global__ void __launch_bounds__(1024)
outlier_scan_small_kernel(const __half* __restrict__ y, long long ld, int T, int K, float thr, float* __restrict__ mask,
                          int* __restrict__ cols, int* __restrict__ count) {
  __shared__ int s_warp[32];
  __shared__ int s_base;
  if (threadIdx.x == 0) s_base = 0;
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int nchunk = K >> 3;                                  // 8 columns (16 bytes) per thread and round
  for (int c0 = 0; c0 < nchunk; c0 += 1024) {
    const int ch = c0 + threadIdx.x;
    unsigned bits = 0u;
    if (ch < nchunk) {
      for (int t = 0; t < T; ++t) {
        const uint4 u = ld_nc_v4(y + (long long)t * ld + ch * 8);
        const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const float2 f = h2_to_f2(w[e]);
          if (fabsf(f.x) >= thr) bits |= 1u << (2 * e);
          if (fabsf(f.y) >= thr) bits |= 1u << (2 * e + 1);
        }
      }
      float m[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) m[e] = ((bits >> e) & 1u) ? 0.0f : 1.0f;
      *reinterpret_cast<float4*>(mask + ch * 8) = make_float4(m[0], m[1], m[2], m[3]);
      *reinterpret_cast<float4*>(mask + ch * 8 + 4) = make_float4(m[4], m[5], m[6], m[7]);
    }
    // ordered compaction: exclusive prefix of the per-thread counts (warp scan, then over the warps)
    const int mine = __popc(bits);
    int incl = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int v = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += v;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    int before = 0, total = 0;
    for (int w = 0; w < 32; ++w) { const int c = s_warp[w]; if (w < warp) before += c; total += c; }
    int pos = s_base + before + incl - mine;
#pragma unroll
    for (int e = 0; e < 8; ++e)
      if ((bits >> e) & 1u) cols[pos++] = ch * 8 + e;
    __syncthreads();
    if (threadIdx.x == 0) s_base += total;
    __syncthreads();
  }
  if (threadIdx.x == 0) *count = s_base;
}

int outlier_scan(const void* y, long long ld, int T, int K, float threshold, float* mask, int* cols, int* count,
                 unsigned* flags, cudaStream_t st) {
  if (K <= 0 || (K & 7) || T < 0 || (ld & 7)) return ERR_BAD_SHAPE;
  if (!mask || !cols || !count || !flags || (T > 0 && !y)) return ERR_NULL_PTR;
  if (reinterpret_cast<uintptr_t>(y) & 15) return ERR_BAD_ALIGN;
  if (T > 0 && T <= OS_SMALL_T) {
    outlier_scan_small_kernel<<<1, 1024, 0, st>>>((const __half*)y, ld, T, K, threshold, mask, cols, count);
    return launch_status();
  }
  if (cudaMemsetAsync(flags, 0, (size_t)(K >> 3) * sizeof(unsigned), st) != cudaSuccess) return ERR_LAUNCH;
  if (T > 0) {
    const dim3 grid(((K >> 3) + 127) / 128, (T + OS_ROWS - 1) / OS_ROWS);
    outlier_flag_kernel<<<grid, 128, 0, st>>>((const __half*)y, ld, T, K, threshold, flags);
  }
  outlier_compact_kernel<<<1, 1024, 0, st>>>(flags, K, mask, cols, count);
  return launch_status();
}

int outlier_side(void* out, const void* y, long long ld, const int8_t* W, const float* w_scale, const int* cols,
                 const int* count, const void* residual, int T, int N, int K, cudaStream_t st) {
  if (T == 0 || N == 0) return OK;
  if (!out || !y || !W || !w_scale || !cols || !count) return ERR_NULL_PTR;
  if (T < 0 || N < 0 || K <= 0) return ERR_BAD_SHAPE;
  outlier_side_kernel<<<dim3((N + 255) / 256, T), 256, 0, st>>>((__half*)out, (const __half*)y, ld, W, w_scale, cols, count,
                                                               (const __half*)residual, N, K);
  return launch_status();
}

// ---------------------------------------------------------------------------------------------
// Dense side product for long batches.  Over a 2048-token prefill chunk most columns of the MLP
// intermediate hold at least one value >= 6, so the "few outlier columns" become thousands and the
// scalar kernel above turns into a naive GEMM.  Then: zero the NON-outlier columns of y, dequantise W
// to fp16 once per call, and run the side product as a tcgen05 fp16 GEMM (the lm_head kernel,
// fp32 accumulators), followed by one combine pass.  Summation order differs from the scalar kernel,
// so this path is compared with the oracle within a tolerance, not bit for bit.
// ---------------------------------------------------------------------------------------------
int lm_head(const void* x, const void* W, float* logits, int M, int V, int d, cudaStream_t st);   // gemm_w8a8.cu

__global__ void __launch_bounds__(256)
outlier_keep_kernel(const __half* __restrict__ y, long long ld, const float* __restrict__ mask, __half* __restrict__ ysub, int K) {
  const int t = blockIdx.y;
  const int ch = blockIdx.x * blockDim.x + threadIdx.x;
  if (ch >= (K >> 3)) return;
  uint4 u = ld_nc_v4(y + (long long)t * ld + ch * 8);
  const float4 m0 = *reinterpret_cast<const float4*>(mask + ch * 8);
  const float4 m1 = *reinterpret_cast<const float4*>(mask + ch * 8 + 4);
  // mask = 1 marks a column that went through the int8 product: drop it here
  if (m0.x != 0.f) u.x &= 0xffff0000u; if (m0.y != 0.f) u.x &= 0x0000ffffu;
  if (m0.z != 0.f) u.y &= 0xffff0000u; if (m0.w != 0.f) u.y &= 0x0000ffffu;
  if (m1.x != 0.f) u.z &= 0xffff0000u; if (m1.y != 0.f) u.z &= 0x0000ffffu;
  if (m1.z != 0.f) u.w &= 0xffff0000u; if (m1.w != 0.f) u.w &= 0x0000ffffu;
  st_v4(ysub + (long long)t * K + ch * 8, u);
}

__global__ void __launch_bounds__(256)
dequant_w_kernel(const int8_t* __restrict__ W, const float* __restrict__ w_scale, __half* __restrict__ wd, int K) {
  const int n = blockIdx.y;
  const int ch = blockIdx.x * blockDim.x + threadIdx.x;      // 16-element chunk
  if (ch >= (K >> 4)) return;
  const uint4 u = ld_nc_v4(W + (long long)n * K + ch * 16);
  const float ws = w_scale[n];
  const uint32_t w4[4] = {u.x, u.y, u.z, u.w};
  uint32_t o[8];
#pragma unroll
  for (int e = 0; e < 4; ++e) {
    const int b0 = (int)(int8_t)(w4[e] & 0xff), b1 = (int)(int8_t)((w4[e] >> 8) & 0xff);
    const int b2_ = (int)(int8_t)((w4[e] >> 16) & 0xff), b3 = (int)(int8_t)((w4[e] >> 24) & 0xff);
    o[2 * e] = f2_to_h2(mul_rn((float)b0, ws), mul_rn((float)b1, ws));
    o[2 * e + 1] = f2_to_h2(mul_rn((float)b2_, ws), mul_rn((float)b3, ws));
  }
  __half* dst = wd + (long long)n * K + ch * 16;
  st_v4(dst, make_uint4(o[0], o[1], o[2], o[3]));
  st_v4(dst + 8, make_uint4(o[4], o[5], o[6], o[7]));
}

__global__ void __launch_bounds__(256)
outlier_combine_kernel(__half* __restrict__ out, const float* __restrict__ side, const __half* __restrict__ residual,
                       long long total) {
  const long long i = ((long long)blockIdx.x * blockDim.x + threadIdx.x) * 4;
  if (i >= total) return;
#pragma unroll
  for (int e = 0; e < 4; ++e) {
    if (i + e < total) {
      float o = __half2float(__float2half_rn(add_rn(__half2float(out[i + e]), __half2float(__float2half_rn(side[i + e])))));
      if (residual) o = add_rn(o, __half2float(residual[i + e]));
      out[i + e] = __float2half_rn(o);
    }
  }
}

// scratch: ysub [T,K] fp16, wdeq [N,K] fp16, side [T,N] fp32 (caller-owned)
int outlier_side_dense(void* out, const void* y, long long ld, const int8_t* W, const float* w_scale, const float* mask,
                       const void* residual, void* ysub, void* wdeq, float* side, int T, int N, int K, cudaStream_t st) {
  if (T == 0 || N == 0) return OK;
  if (!out || !y || !W || !w_scale || !mask || !ysub || !wdeq || !side) return ERR_NULL_PTR;
  if (T < 0 || N < 0 || K <= 0 || (K & 15) || (ld & 7)) return ERR_BAD_SHAPE;
  outlier_keep_kernel<<<dim3(((K >> 3) + 255) / 256, T), 256, 0, st>>>((const __half*)y, ld, mask, (__half*)ysub, K);
  dequant_w_kernel<<<dim3(((K >> 4) + 255) / 256, N), 256, 0, st>>>(W, w_scale, (__half*)wdeq, K);
  const int rc = lm_head(ysub, wdeq, side, T, N, K, st);
  if (rc) return rc;
  const long long total = (long long)T * N;
  outlier_combine_kernel<<<(unsigned)((total / 4 + 255) / 256), 256, 0, st>>>((__half*)out, side, (const __half*)residual, total);
  return launch_status();
}

int init_outlier() {
  cudaFuncAttributes fa;
  bool ok = cudaFuncGetAttributes(&fa, outlier_flag_kernel) == cudaSuccess;
  ok &= cudaFuncGetAttributes(&fa, outlier_compact_kernel) == cudaSuccess;
  ok &= cudaFuncGetAttributes(&fa, outlier_side_kernel) == cudaSuccess;
  ok &= cudaFuncGetAttributes(&fa, outlier_scan_small_kernel) == cudaSuccess;
  ok &= cudaFuncGetAttributes(&fa, outlier_keep_kernel) == cudaSuccess;
  ok &= cudaFuncGetAttributes(&fa, dequant_w_kernel) == cudaSuccess;
  ok &= cudaFuncGetAttributes(&fa, outlier_combine_kernel) == cudaSuccess;
  return ok ? OK : ERR_ARCH;
}

}  // namespace b2
