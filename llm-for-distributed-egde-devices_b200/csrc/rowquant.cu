// Fused row kernels: (RMSNorm | LayerNorm | identity | act(gate)*up | act) -> smoothing multiply
// -> per-token absmax int8 quantise.   HBM-bound: one CTA per token row, 128-bit loads, the row
// stays in registers between the three passes (moments, absmax, quantise), warp-shuffle + one
// shared-memory hop for the block reductions.
//
// Replaces, in the reference's stack, the norm modules and the activation-quantising half of the
// int8 linear that HF/bitsandbytes run under `model.generate` (reference call site
// Code/Quantised Models/models_quant_updated.py:84-93; transformers modeling_llama.py:62-67 RMSNorm,
// :182-184 MLP act).  Operation order is pinned by oracle/w8a8.py (norm_quant, smooth_quant,
// act_mul_quant) and reproduced here exactly: fp64 moment sums, IEEE fp32 elementwise.
#include "common.cuh"

namespace b2 {

enum RowMode : int { ROW_PLAIN = 0, ROW_RMS = 1, ROW_LN = 2, ROW_SILU_MUL = 3, ROW_GELU_MUL = 4, ROW_GELU = 5, ROW_SILU = 6,
                     ROW_GELU_ERF = 7 };

template <int NWARPS>
B2_DEVINL double block_sum_d(double v, double* sh) {
  v = warp_sum_d(v);
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  __syncthreads();
  if (l == 0) sh[w] = v;
  __syncthreads();
  double t = 0.0;
#pragma unroll
  for (int i = 0; i < NWARPS; ++i) t += sh[i];   // same order in every thread
  return t;
}
template <int NWARPS>
B2_DEVINL float block_max_f(float v, float* sh) {
  v = warp_max(v);
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  __syncthreads();
  if (l == 0) sh[w] = v;
  __syncthreads();
  float t = 0.0f;
#pragma unroll
  for (int i = 0; i < NWARPS; ++i) t = fmaxf(t, sh[i]);
  return t;
}

// x: [T, in_stride] fp16.  K = number of quantised columns per row.
// MODE ROW_*_MUL: row holds gate in [0,K) and up in [K,2K).
template <int MODE, int NTHREADS, int MAXC, bool OUT_F16>
__global__ void __launch_bounds__(NTHREADS) rowquant_kernel(
    const __half* __restrict__ x, long long in_stride, const __half* __restrict__ gamma,
    const __half* __restrict__ beta, const float* __restrict__ smooth, int8_t* __restrict__ q,
    float* __restrict__ row_scale, __half* __restrict__ y16, int K, float eps) {
  constexpr int NWARPS = NTHREADS / 32;
  __shared__ double shd[2 * NWARPS];
  __shared__ float shf[NWARPS];
  const int row = blockIdx.x;
  const __half* xr = x + (long long)row * in_stride;
  const int nchunk = K >> 3;   // 8 halves per chunk

  float v[MAXC][8];
  // ---- pass 1: load, activation, moments ----------------------------------------------------
  double s1 = 0.0, s2 = 0.0;
#pragma unroll
  for (int c = 0; c < MAXC; ++c) {
    const int ch = threadIdx.x + c * NTHREADS;
    if (ch < nchunk) {
      uint4 u = ld_nc_v4(xr + ch * 8);
      float2 a0 = h2_to_f2(u.x), a1 = h2_to_f2(u.y), a2 = h2_to_f2(u.z), a3 = h2_to_f2(u.w);
      v[c][0] = a0.x; v[c][1] = a0.y; v[c][2] = a1.x; v[c][3] = a1.y;
      v[c][4] = a2.x; v[c][5] = a2.y; v[c][6] = a3.x; v[c][7] = a3.y;
      if (MODE == ROW_SILU_MUL || MODE == ROW_GELU_MUL) {
        uint4 w = ld_nc_v4(xr + K + ch * 8);
        float2 b0 = h2_to_f2(w.x), b1 = h2_to_f2(w.y), b2_ = h2_to_f2(w.z), b3 = h2_to_f2(w.w);
        const float up[8] = {b0.x, b0.y, b1.x, b1.y, b2_.x, b2_.y, b3.x, b3.y};
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          float a = (MODE == ROW_SILU_MUL) ? silu_det(v[c][j]) : gelu_new_det(v[c][j]);
          v[c][j] = mul_rn(a, up[j]);
        }
      } else if (MODE == ROW_GELU) {
#pragma unroll
        for (int j = 0; j < 8; ++j) v[c][j] = gelu_new_det(v[c][j]);
      } else if (MODE == ROW_SILU) {
#pragma unroll
        for (int j = 0; j < 8; ++j) v[c][j] = silu_det(v[c][j]);
      } else if (MODE == ROW_GELU_ERF) {
#pragma unroll
        for (int j = 0; j < 8; ++j) v[c][j] = gelu_erf_det(v[c][j]);
      } else if (MODE == ROW_RMS || MODE == ROW_LN) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          double d = (double)v[c][j];
          s2 += d * d;
          if (MODE == ROW_LN) s1 += d;
        }
      }
    }
  }
  // ---- normalisation ------------------------------------------------------------------------
  if (MODE == ROW_RMS || MODE == ROW_LN) {
    float mean = 0.0f, var;
    const double ss = block_sum_d<NWARPS>(s2, shd);
    if (MODE == ROW_LN) {
      const double sm = block_sum_d<NWARPS>(s1, shd + NWARPS);
      const double md = sm / (double)K;
      double vd = ss / (double)K - md * md;
      vd = vd > 0.0 ? vd : 0.0;
      mean = (float)md;
      var = (float)vd;
    } else {
      var = (float)(ss / (double)K);
    }
    const float r = div_rn(1.0f, __fsqrt_rn(add_rn(var, eps)));
#pragma unroll
    for (int c = 0; c < MAXC; ++c) {
      const int ch = threadIdx.x + c * NTHREADS;
      if (ch < nchunk) {
        uint4 g = ld_v4(gamma + ch * 8);
        float2 g0 = h2_to_f2(g.x), g1 = h2_to_f2(g.y), g2 = h2_to_f2(g.z), g3 = h2_to_f2(g.w);
        const float gg[8] = {g0.x, g0.y, g1.x, g1.y, g2.x, g2.y, g3.x, g3.y};
        float bb[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        if (MODE == ROW_LN) {
          uint4 b = ld_v4(beta + ch * 8);
          float2 b0 = h2_to_f2(b.x), b1 = h2_to_f2(b.y), b2_ = h2_to_f2(b.z), b3 = h2_to_f2(b.w);
          bb[0] = b0.x; bb[1] = b0.y; bb[2] = b1.x; bb[3] = b1.y;
          bb[4] = b2_.x; bb[5] = b2_.y; bb[6] = b3.x; bb[7] = b3.y;
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          float t = v[c][j];
          if (MODE == ROW_LN) t = sub_rn(t, mean);
          t = mul_rn(mul_rn(t, r), gg[j]);
          if (MODE == ROW_LN) t = add_rn(t, bb[j]);
          v[c][j] = t;
        }
      }
    }
  }
  if (OUT_F16) {   // final norm: fp16 result, no quantisation
#pragma unroll
    for (int c = 0; c < MAXC; ++c) {
      const int ch = threadIdx.x + c * NTHREADS;
      if (ch < nchunk) {
        uint4 o;
        o.x = f2_to_h2(v[c][0], v[c][1]); o.y = f2_to_h2(v[c][2], v[c][3]);
        o.z = f2_to_h2(v[c][4], v[c][5]); o.w = f2_to_h2(v[c][6], v[c][7]);
        st_v4(y16 + (long long)row * K + ch * 8, o);
      }
    }
    return;
  }
  // ---- smoothing + absmax -------------------------------------------------------------------
  float amax = 0.0f;
#pragma unroll
  for (int c = 0; c < MAXC; ++c) {
    const int ch = threadIdx.x + c * NTHREADS;
    if (ch < nchunk) {
      if (smooth != nullptr) {
        const float4 sA = *reinterpret_cast<const float4*>(smooth + ch * 8);
        const float4 sB = *reinterpret_cast<const float4*>(smooth + ch * 8 + 4);
        v[c][0] = mul_rn(v[c][0], sA.x); v[c][1] = mul_rn(v[c][1], sA.y);
        v[c][2] = mul_rn(v[c][2], sA.z); v[c][3] = mul_rn(v[c][3], sA.w);
        v[c][4] = mul_rn(v[c][4], sB.x); v[c][5] = mul_rn(v[c][5], sB.y);
        v[c][6] = mul_rn(v[c][6], sB.z); v[c][7] = mul_rn(v[c][7], sB.w);
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) amax = fmaxf(amax, fabsf(v[c][j]));
    }
  }
  amax = block_max_f<NWARPS>(amax, shf);
  const bool nz = amax > 0.0f;
  const float inv = nz ? div_rn(127.0f, amax) : 0.0f;
  if (threadIdx.x == 0) row_scale[row] = nz ? div_rn(amax, 127.0f) : 0.0f;
  // ---- quantise -----------------------------------------------------------------------------
  int8_t* qr = q + (long long)row * K;
#pragma unroll
  for (int c = 0; c < MAXC; ++c) {
    const int ch = threadIdx.x + c * NTHREADS;
    if (ch < nchunk) {
      uint2 o;
      o.x = pack4(quant1(v[c][0], inv), quant1(v[c][1], inv), quant1(v[c][2], inv), quant1(v[c][3], inv));
      o.y = pack4(quant1(v[c][4], inv), quant1(v[c][5], inv), quant1(v[c][6], inv), quant1(v[c][7], inv));
      *reinterpret_cast<uint2*>(qr + ch * 8) = o;
    }
  }
}

template <int MODE, bool OUT_F16>
static int launch_rowquant(const __half* x, long long in_stride, const __half* gamma, const __half* beta,
                           const float* smooth, int8_t* q, float* row_scale, __half* y16, int T, int K,
                           float eps, cudaStream_t st) {
  if (T == 0) return OK;
  if (T < 0 || K <= 0 || (K & 7)) return ERR_BAD_SHAPE;
  const int nchunk = K >> 3;
  constexpr int MAXC = 4;
  if (nchunk <= 128 * MAXC)
    rowquant_kernel<MODE, 128, MAXC, OUT_F16><<<T, 128, 0, st>>>(x, in_stride, gamma, beta, smooth, q, row_scale, y16, K, eps);
  else if (nchunk <= 256 * MAXC)
    rowquant_kernel<MODE, 256, MAXC, OUT_F16><<<T, 256, 0, st>>>(x, in_stride, gamma, beta, smooth, q, row_scale, y16, K, eps);
  else if (nchunk <= 512 * MAXC)
    rowquant_kernel<MODE, 512, MAXC, OUT_F16><<<T, 512, 0, st>>>(x, in_stride, gamma, beta, smooth, q, row_scale, y16, K, eps);
  else if (nchunk <= 1024 * MAXC)
    rowquant_kernel<MODE, 1024, MAXC, OUT_F16><<<T, 1024, 0, st>>>(x, in_stride, gamma, beta, smooth, q, row_scale, y16, K, eps);
  else
    return ERR_TOO_LARGE;
  return launch_status();
}

static inline bool al16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

int norm_quant(const void* x, const void* gamma, const void* beta, const float* smooth, int8_t* q,
               float* row_scale, int T, int d, float eps, cudaStream_t st) {
  if (T == 0) return OK;
  if (!x || !gamma || !q || !row_scale) return ERR_NULL_PTR;
  if (!al16(x) || !al16(gamma) || !al16(q) || (beta && !al16(beta)) || (smooth && !al16(smooth))) return ERR_BAD_ALIGN;
  if (beta)
    return launch_rowquant<ROW_LN, false>((const __half*)x, d, (const __half*)gamma, (const __half*)beta, smooth, q,
                                          row_scale, nullptr, T, d, eps, st);
  return launch_rowquant<ROW_RMS, false>((const __half*)x, d, (const __half*)gamma, nullptr, smooth, q, row_scale,
                                         nullptr, T, d, eps, st);
}

int norm_f16(const void* x, const void* gamma, const void* beta, void* y, int T, int d, float eps, cudaStream_t st) {
  if (T == 0) return OK;
  if (!x || !gamma || !y) return ERR_NULL_PTR;
  if (!al16(x) || !al16(gamma) || !al16(y) || (beta && !al16(beta))) return ERR_BAD_ALIGN;
  if (beta)
    return launch_rowquant<ROW_LN, true>((const __half*)x, d, (const __half*)gamma, (const __half*)beta, nullptr,
                                         nullptr, nullptr, (__half*)y, T, d, eps, st);
  return launch_rowquant<ROW_RMS, true>((const __half*)x, d, (const __half*)gamma, nullptr, nullptr, nullptr, nullptr,
                                        (__half*)y, T, d, eps, st);
}

int smooth_quant(const void* x, const float* smooth, int8_t* q, float* row_scale, int T, int K, cudaStream_t st) {
  if (T == 0) return OK;
  if (!x || !q || !row_scale) return ERR_NULL_PTR;
  if (!al16(x) || !al16(q) || (smooth && !al16(smooth))) return ERR_BAD_ALIGN;
  return launch_rowquant<ROW_PLAIN, false>((const __half*)x, K, nullptr, nullptr, smooth, q, row_scale, nullptr, T, K,
                                           0.f, st);
}

// act: 0 = silu(gate)*up, 1 = gelu_new(gate)*up, 2 = gelu_new(x) (ungated), 3 = silu(x) (ungated), 4 = exact-erf gelu(x) (ungated)
int act_mul_quant(const void* gate_up, const float* smooth, int8_t* q, float* row_scale, int T, int F, int act,
                  cudaStream_t st) {
  if (T == 0) return OK;
  if (!gate_up || !q || !row_scale) return ERR_NULL_PTR;
  if (!al16(gate_up) || !al16(q) || (smooth && !al16(smooth))) return ERR_BAD_ALIGN;
  const __half* x = (const __half*)gate_up;
  switch (act) {
    case 0: return launch_rowquant<ROW_SILU_MUL, false>(x, 2LL * F, nullptr, nullptr, smooth, q, row_scale, nullptr, T, F, 0.f, st);
    case 1: return launch_rowquant<ROW_GELU_MUL, false>(x, 2LL * F, nullptr, nullptr, smooth, q, row_scale, nullptr, T, F, 0.f, st);
    case 2: return launch_rowquant<ROW_GELU, false>(x, F, nullptr, nullptr, smooth, q, row_scale, nullptr, T, F, 0.f, st);
    case 3: return launch_rowquant<ROW_SILU, false>(x, F, nullptr, nullptr, smooth, q, row_scale, nullptr, T, F, 0.f, st);
    case 4: return launch_rowquant<ROW_GELU_ERF, false>(x, F, nullptr, nullptr, smooth, q, row_scale, nullptr, T, F, 0.f, st);
    default: return ERR_UNSUPPORTED;
  }
}

// act(gate)*up (or act(x)) as fp16, no quantisation: the LLM.int8() path looks for outlier columns in the
// fp16 activation before it quantises (outlier.cu)
int act_mul_f16(const void* gate_up, void* y, int T, int F, int act, cudaStream_t st) {
  if (T == 0) return OK;
  if (!gate_up || !y) return ERR_NULL_PTR;
  if (!al16(gate_up) || !al16(y)) return ERR_BAD_ALIGN;
  const __half* x = (const __half*)gate_up;
  switch (act) {
    case 0: return launch_rowquant<ROW_SILU_MUL, true>(x, 2LL * F, nullptr, nullptr, nullptr, nullptr, nullptr, (__half*)y, T, F, 0.f, st);
    case 1: return launch_rowquant<ROW_GELU_MUL, true>(x, 2LL * F, nullptr, nullptr, nullptr, nullptr, nullptr, (__half*)y, T, F, 0.f, st);
    case 2: return launch_rowquant<ROW_GELU, true>(x, F, nullptr, nullptr, nullptr, nullptr, nullptr, (__half*)y, T, F, 0.f, st);
    case 3: return launch_rowquant<ROW_SILU, true>(x, F, nullptr, nullptr, nullptr, nullptr, nullptr, (__half*)y, T, F, 0.f, st);
    case 4: return launch_rowquant<ROW_GELU_ERF, true>(x, F, nullptr, nullptr, nullptr, nullptr, nullptr, (__half*)y, T, F, 0.f, st);
    default: return ERR_UNSUPPORTED;
  }
}

// Force-load every instantiation (lazy module loading must not happen under graph capture).
template <typename K>
static bool touch(K kern) {
  cudaFuncAttributes fa;
  return cudaFuncGetAttributes(&fa, kern) == cudaSuccess;
}
template <int MODE, bool F16>
static bool touch_mode() {
  return touch(rowquant_kernel<MODE, 128, 4, F16>) && touch(rowquant_kernel<MODE, 256, 4, F16>) &&
         touch(rowquant_kernel<MODE, 512, 4, F16>) && touch(rowquant_kernel<MODE, 1024, 4, F16>);
}
int init_rowquant() {
  bool ok = touch_mode<ROW_PLAIN, false>() && touch_mode<ROW_RMS, false>() && touch_mode<ROW_LN, false>() &&
            touch_mode<ROW_SILU_MUL, false>() && touch_mode<ROW_GELU_MUL, false>() && touch_mode<ROW_GELU, false>() &&
            touch_mode<ROW_SILU, false>() && touch_mode<ROW_RMS, true>() && touch_mode<ROW_LN, true>() &&
            touch_mode<ROW_SILU_MUL, true>() && touch_mode<ROW_GELU_MUL, true>() && touch_mode<ROW_GELU, true>() &&
            touch_mode<ROW_SILU, true>() && touch_mode<ROW_GELU_ERF, false>() && touch_mode<ROW_GELU_ERF, true>();
  return ok ? OK : ERR_ARCH;
}

}  // namespace b2
