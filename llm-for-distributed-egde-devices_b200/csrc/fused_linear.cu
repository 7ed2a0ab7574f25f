// Decode-time fused linear (M <= 8 tokens):   [norm | smoothing | act-requant] -> per-token int8
// quantise -> W8A8 weight-streaming GEMV -> dequantising epilogue [-> act(gate)*up],  ONE kernel.
//
// Why: at M = 1 a 7B decoder layer streams 202 MB of weights in ~31 us at HBM speed, but the unfused
// chain was 11 launches per layer, each with microseconds of launch + dependent-load latency.  Here
// every CTA redoes the (tiny, L2-resident) activation prologue itself, so the quantised activation
// never round-trips through HBM and four launches per layer disappear.  The kernel is launched with
// programmatic dependent launch: it prefetches its first weight tiles BEFORE waiting on the
// predecessor (weights never depend on it), so the launch/ramp latency hides under the previous
// kernel's tail.
//
// Numerics are those of the unfused entry points, operation for operation (oracle/w8a8.py:
// norm_quant / smooth_quant / act_mul_y + quant_rows, gemm_w8a8): results are bit-identical to
// b2_norm_quant + b2_gemm_w8a8 etc.  (tests/test_kernels_gpu.py::test_fused_linear_*).
//
// Reference call sites replaced: the same as b2_norm_quant / b2_gemm_w8a8 / b2_act_mul_quant
// (Code/Quantised Models/models_quant_updated.py:30-38, :84-93 decode loop).
#include <algorithm>
#include "common.cuh"
#include "../../include/b200stage.h"

namespace b2 {

enum : int { X_RMS = 0, X_LN = 1, X_PLAIN = 2, X_Y32 = 3 };
enum : int { O_F16 = 0, O_GATED = 1, O_QKV = 2 };

struct FusedParams {
  const void* x; long long x_ld;                 // fp16 (X_RMS/X_LN/X_PLAIN) or fp32 (X_Y32) [M, x_ld]
  const __half* gamma; const __half* beta;       // norm parameters
  const float* smooth;                           // [K] multiplier or nullptr
  const float* x_amax;                           // X_Y32: row absmax [M] produced by the O_GATED kernel
  float eps;
  const int8_t* W; const float* w_scale; const __half* bias; const __half* residual;
  void* out;                                     // O_F16: fp16 [M,N];  O_GATED: fp32 y [M,N/2]
  const float* out_smooth;                       // O_GATED: [N/2] multiplier of the NEXT linear, or nullptr
  float* out_amax;                               // O_GATED: [M] running absmax (atomicMax; caller zeroes)
  float* zero_amax;                              // housekeeping: [M] floats this launch sets to 0, or nullptr
  int M, N, K, act;
  // L2 prefetch hints: weights a LATER kernel will stream (issued at kernel start, split over the CTAs)
  const uint8_t* pf0; long long pf0_bytes; const uint8_t* pf1; long long pf1_bytes;
  // O_QKV: rotary embedding + KV-cache append in the epilogue (q written rotated to `out`)
  const float* cosT; const float* sinT; const int* positions; const int* seq_ids; int seq0;
  __half* kcache; __half* vcache; int H, KV, hd, Smax, use_rope;
};

constexpr int FL_THREADS = 512, FL_WARPS = 16, FL_ROWS = 2, FL_UNR = 4;
constexpr int FL_SMEM_FULL = 200 * 1024;
#ifndef FL_CTAS_PER_SM
#define FL_CTAS_PER_SM 1
#endif

B2_DEVINL void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
B2_DEVINL void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

B2_DEVINL void fl_cp_async16(void* smem_dst, const void* src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(smem_dst)), "l"(src) : "memory");
}
// this CTA's share of an L2 prefetch of [ptr, ptr+bytes): 16 KB bulk-prefetch requests, dealt out to
// lane 0 of successive warps (a bulk prefetch takes 50-100 ns to issue and they serialise within a
// warp: ten in one thread held that warp's first weight loads back by more than a microsecond,
// tools/timeline.py)
B2_DEVINL void fl_prefetch_slice(const uint8_t* ptr, long long bytes) {
  if (ptr == nullptr || bytes <= 0 || (threadIdx.x & 31)) return;
  constexpr long long PIECE = 16 * 1024;
  const long long npieces = (bytes + PIECE - 1) / PIECE;
  const int warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
  for (long long i = blockIdx.x + (long long)gridDim.x * warp; i < npieces; i += (long long)gridDim.x * nwarps) {
    const long long off = i * PIECE;
    const uint32_t n = (uint32_t)((bytes - off < PIECE ? bytes - off : PIECE) & ~15ll);
    if (n) asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(ptr + off), "r"(n) : "memory");
  }
}
B2_DEVINL double fl_block_sum_d(double v, double* sh) {
  v = warp_sum_d(v);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = v;
  __syncthreads();
  double t = 0.0;
#pragma unroll
  for (int i = 0; i < FL_WARPS; ++i) t += sh[i];
  return t;
}
B2_DEVINL float fl_block_max(float v, float* sh) {
  v = warp_max(v);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = v;
  __syncthreads();
  float t = 0.0f;
#pragma unroll
  for (int i = 0; i < FL_WARPS; ++i) t = fmaxf(t, sh[i]);
  return t;
}

// D[16x8] += A[16x32] * B[32x8], int8 x int8 -> int32 (exact: the order of the k terms does not matter)
B2_DEVINL void fl_mma_s8(int (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.s8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3])
               : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
// Token batches of 2..8 (MT > 1) stream the weights through mma.sync instead of dp4a -- see section 4b.
constexpr int FL_PART_BYTES = 2 * FL_WARPS * 32 * 16;    // two buffers of per-warp partial accumulators
constexpr int FL_XS_PAD = 64;                            // row pitch K + 64: the 8 token rows of a B fragment hit distinct banks
__host__ __device__ constexpr size_t fl_act_bytes(int MT, int K) {
  return MT > 1 ? (size_t)FL_PART_BYTES + (((size_t)MT * (K + FL_XS_PAD) + 127) & ~size_t(127)) : (((size_t)MT * K + 127) & ~size_t(127));
}

// y[0..7] of 8-element chunk `ch` of one row, in the pinned operation order.
template <int XK>
B2_DEVINL void fl_load_y(const FusedParams& p, const void* xrow, int ch, float r, float mean, float (&y)[8]) {
  if (XK == X_Y32) {
    const float4 a = __ldcg(reinterpret_cast<const float4*>(reinterpret_cast<const float*>(xrow) + ch * 8));
    const float4 b = __ldcg(reinterpret_cast<const float4*>(reinterpret_cast<const float*>(xrow) + ch * 8 + 4));
    y[0] = a.x; y[1] = a.y; y[2] = a.z; y[3] = a.w; y[4] = b.x; y[5] = b.y; y[6] = b.z; y[7] = b.w;
    return;                                       // smoothing was applied by the producer
  }
  const uint4 u = ld_cg_v4(reinterpret_cast<const __half*>(xrow) + ch * 8);
  const float2 a0 = h2_to_f2(u.x), a1 = h2_to_f2(u.y), a2 = h2_to_f2(u.z), a3 = h2_to_f2(u.w);
  y[0] = a0.x; y[1] = a0.y; y[2] = a1.x; y[3] = a1.y; y[4] = a2.x; y[5] = a2.y; y[6] = a3.x; y[7] = a3.y;
  if (XK == X_RMS || XK == X_LN) {
    const uint4 g = ld_v4(p.gamma + ch * 8);
    const float2 g0 = h2_to_f2(g.x), g1 = h2_to_f2(g.y), g2 = h2_to_f2(g.z), g3 = h2_to_f2(g.w);
    const float gg[8] = {g0.x, g0.y, g1.x, g1.y, g2.x, g2.y, g3.x, g3.y};
    float bb[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    if (XK == X_LN) {
      const uint4 b = ld_v4(p.beta + ch * 8);
      const float2 b0 = h2_to_f2(b.x), b1 = h2_to_f2(b.y), b2_ = h2_to_f2(b.z), b3 = h2_to_f2(b.w);
      bb[0] = b0.x; bb[1] = b0.y; bb[2] = b1.x; bb[3] = b1.y; bb[4] = b2_.x; bb[5] = b2_.y; bb[6] = b3.x; bb[7] = b3.y;
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float t = y[j];
      if (XK == X_LN) t = sub_rn(t, mean);
      t = mul_rn(mul_rn(t, r), gg[j]);
      if (XK == X_LN) t = add_rn(t, bb[j]);
      y[j] = t;
    }
  }
  if (p.smooth != nullptr) {
    const float4 sA = *reinterpret_cast<const float4*>(p.smooth + ch * 8);
    const float4 sB = *reinterpret_cast<const float4*>(p.smooth + ch * 8 + 4);
    y[0] = mul_rn(y[0], sA.x); y[1] = mul_rn(y[1], sA.y); y[2] = mul_rn(y[2], sA.z); y[3] = mul_rn(y[3], sA.w);
    y[4] = mul_rn(y[4], sB.x); y[5] = mul_rn(y[5], sB.y); y[6] = mul_rn(y[6], sB.z); y[7] = mul_rn(y[7], sB.w);
  }
}

// One fused linear = one "phase".  SYNC is what separates the part that may run ahead of the
// producer of x (cursor set-up, first weight batches, staging of gains/multipliers) from the part that
// consumes x: griddepcontrol.wait when the phase is a kernel of its own, a grid barrier when it is a
// later phase of a multi-phase persistent kernel.  FIRST = false inserts the CTA barrier that
// protects the previous phase's shared memory.  Activations written by other CTAs are read with
// L2-coherent (.cg) loads: L1 is not invalidated between phases of one kernel.
template <int MT, int XK, int OKIND, bool FIRST, class SYNC>
B2_DEVINL void fused_phase(const FusedParams& p, uint8_t* xs, SYNC sync_fn, bool trigger) {
  __shared__ double shd[FL_WARPS];
  __shared__ float shf[FL_WARPS];
  __shared__ float s_scale[MT];
  const int K = p.K, kvec = K >> 4, nchunk = K >> 3;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int F = p.N >> 1;
  constexpr bool TCM = MT > 1;                         // 2..8 tokens: tensor-core stream (section 4b)
  uint8_t* const part = xs;                            // TCM: partial accumulators of the 16 warps, double-buffered
  if (TCM) xs += FL_PART_BYTES;
  const int XS = TCM ? K + FL_XS_PAD : K;              // pitch of a quantised activation row in shared memory
  const int g8 = lane >> 2, t4 = lane & 3;             // mma fragment coordinates of this lane
  constexpr int STEP = FL_UNR * 32;                    // 16-byte vectors per row per batch
  // batches per work item.  TCM: a unit's K is dealt out to the 16 warps in 64-byte pieces
  const int NB = TCM ? (((kvec + 3) >> 2) + FL_WARPS * FL_UNR - 1) / (FL_WARPS * FL_UNR) : (kvec + STEP - 1) / STEP;
  // A work item = the two weight rows one warp reduces together: two adjacent output columns
  // (O_F16), the gate/up rows of one MLP column (O_GATED), or a rotary pair i, i+hd/2 (O_QKV).
  // TCM: a work UNIT = 8 consecutive items = the 16 rows of one mma A tile (rows 0-7: the items' first
  // rows, 8-15: their second rows); the whole CTA works on one unit, each warp on 1/16 of its K.
  const int items_total = (OKIND == O_F16) ? (p.N + 1) / 2 : F;
  const int work_total = TCM ? (items_total + 7) >> 3 : items_total;
  const int stride = TCM ? gridDim.x : gridDim.x * FL_WARPS;
  const int half_hd = (OKIND == O_QKV) ? (p.hd >> 1) : 1;
  struct Cur { int it, j; const int8_t* r0; const int8_t* r1; };
  // Item schedule: warp w of CTA c takes items c*16 + w + k*stride.  Two "fairer" schedules were
  // measured on B200 and are slower: w*grid + c everywhere (518 vs 508 us per 8-layer token: a CTA's
  // 16 warps then touch 32 different 2 MB pages), and dealing only the partial last round out evenly
  // (530 us): when every CTA finishes at the same moment the successor kernel's CTA launch wave and
  // pre-wait loads are fully exposed, whereas CTAs that skip the last round hand their SM over early.
  auto set_rows = [&](Cur& c) {
    int n0, n1;
    const int item = TCM ? min(c.it * 8 + g8, items_total - 1) : c.it;    // TCM: lane (g, t) streams the rows of item 8*unit + g
    if (OKIND == O_GATED) { n0 = item; n1 = F + item; }
    else if (OKIND == O_QKV) { const int hq = item / half_hd; n0 = hq * p.hd + (item - hq * half_hd); n1 = n0 + half_hd; }
    else { n0 = 2 * item; n1 = min(n0 + 1, p.N - 1); }
    c.r0 = p.W + (long long)n0 * K;
    c.r1 = p.W + (long long)n1 * K;
  };
  auto advance = [&](Cur& c) {
    if (++c.j == NB) { c.j = 0; c.it += stride; if (c.it < work_total) set_rows(c); }
  };
  auto load_batch = [&](uint4 (&dst)[FL_UNR][FL_ROWS], const Cur& c) {
    if (c.it < work_total) {
#pragma unroll
      for (int u = 0; u < FL_UNR; ++u) {
        // TCM: 64-byte piece (warp + 16*(u + 4j)) of the row, this lane's 16 bytes of it: one warp-wide
        // load covers 8 rows x 64 contiguous bytes, the 16 warps together 1 KB of each row
        const int idx = TCM ? 4 * (warp + FL_WARPS * (u + FL_UNR * c.j)) + t4 : lane + c.j * STEP + u * 32;
        const bool ok = idx < kvec;
        dst[u][0] = ok ? ld_nc_v4(c.r0 + idx * 16) : make_uint4(0, 0, 0, 0);
        dst[u][1] = ok ? ld_nc_v4(c.r1 + idx * 16) : make_uint4(0, 0, 0, 0);
      }
    }
  };
  // ---- 1. first two weight batches: independent of the predecessor kernel ----------------------
  constexpr unsigned TLK = 16 + OKIND * 4 + XK;     // timeline probe id (tools/timeline.py)
  tl_mark(TLK, 0);
  if (trigger) pdl_launch_dependents();
  tl_mark(TLK, 6);
  uint4 w[FL_UNR][FL_ROWS], w2[FL_UNR][FL_ROWS];
  Cur cc;
  cc.it = TCM ? blockIdx.x : blockIdx.x * FL_WARPS + warp; cc.j = 0; cc.r0 = p.W; cc.r1 = p.W;
  if (cc.it < work_total) set_rows(cc);
  Cur lc = cc;
  load_batch(w, lc); advance(lc);
  load_batch(w2, lc); advance(lc);
  tl_mark(TLK, 7);
  {
    // Pull the head of the next kernels' weights into L2 while this kernel streams its own: the
    // successor's ramp-up then hits L2 and HBM never idles across the kernel boundary.
    fl_prefetch_slice(p.pf0, p.pf0_bytes);
    fl_prefetch_slice(p.pf1, p.pf1_bytes);
  }
  if (!FIRST) __syncthreads();     // every warp has left the previous phase: its shared memory may be reused
  // Norm gains / smoothing multipliers are weights too: stage them in shared memory with cp.async
  // before the dependency wait, so the prologue's critical path is one L2 read of x plus two block
  // reductions (fast path: rows of up to FP_MAXC*256*8 = 6144 elements held in registers).
  constexpr int FP_MAXC = 3;
  const bool fast = !TCM && (XK != X_Y32) && (nchunk <= FP_MAXC * FL_THREADS);
  uint8_t* aux = xs + (((size_t)MT * XS + 127) & ~size_t(127));
  __half* sg = reinterpret_cast<__half*>(aux);                         // gamma [K]
  __half* sb = sg + K;                                                 // beta  [K]
  float* ssm = reinterpret_cast<float*>(aux + (size_t)4 * K);          // smooth [K]
  if (fast) {
    if (XK == X_RMS || XK == X_LN)
      for (int i = threadIdx.x; i < nchunk; i += FL_THREADS) fl_cp_async16(sg + i * 8, p.gamma + i * 8);
    if (XK == X_LN)
      for (int i = threadIdx.x; i < nchunk; i += FL_THREADS) fl_cp_async16(sb + i * 8, p.beta + i * 8);
    if (p.smooth != nullptr)
      for (int i = threadIdx.x; i < 2 * nchunk; i += FL_THREADS) fl_cp_async16(ssm + i * 4, p.smooth + i * 4);
    asm volatile("cp.async.commit_group;" ::: "memory");
  }
  // ---- 2. everything below may read the predecessor's output ---------------------------------
  tl_mark(TLK, 1);
  sync_fn();
  tl_mark(TLK, 2);
  if (p.zero_amax != nullptr && blockIdx.x == 0 && threadIdx.x < p.M) p.zero_amax[threadIdx.x] = 0.0f;
  if (fast) {
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncthreads();
  }
  // ---- 3. activation prologue: quantise the M rows into shared memory ------------------------
  if constexpr (TCM) {
    // Token batches: the rows are quantised SIDE BY SIDE, 16/MT warps per row, with two CTA-wide
    // reductions in total (one row after the other cost 1.5-2.4 us of dependent L2 latency and four
    // barriers per row: 12-19 us per launch at 8 tokens).  Three passes over the (L2-resident) row:
    // moments, absmax, quantise; the fp64 moment sums do not depend on the order of the terms.
    __shared__ double shd1[FL_WARPS];
    constexpr int WPR = FL_WARPS / MT, RT = WPR * 32;
    const int m = warp / WPR, tr = (warp - m * WPR) * 32 + lane;
    const bool rowok = m < p.M;
    uint2* qrow = reinterpret_cast<uint2*>(xs + (size_t)m * XS);
    const void* xrow = (XK == X_Y32)
        ? static_cast<const void*>(reinterpret_cast<const float*>(p.x) + (long long)m * p.x_ld)
        : static_cast<const void*>(reinterpret_cast<const __half*>(p.x) + (long long)m * p.x_ld);
    float r = 0.f, mean = 0.f, amax = 0.f;
    if (XK == X_RMS || XK == X_LN) {
      double s1 = 0.0, s2 = 0.0;
      if (rowok) {
        for (int ch = tr; ch < nchunk; ch += RT) {
          const uint4 u = ld_cg_v4(reinterpret_cast<const __half*>(xrow) + ch * 8);
          const uint32_t* pu = &u.x;
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const float2 f = h2_to_f2(pu[e]);
            const double d0 = (double)f.x, d1 = (double)f.y;
            s2 += d0 * d0; s2 += d1 * d1;
            if (XK == X_LN) { s1 += d0; s1 += d1; }
          }
        }
      }
      s2 = warp_sum_d(s2);
      if (XK == X_LN) s1 = warp_sum_d(s1);
      if (lane == 0) { shd[warp] = s2; shd1[warp] = s1; }
      __syncthreads();
      double ss = 0.0, sm = 0.0;
#pragma unroll
      for (int i = 0; i < WPR; ++i) { ss += shd[m * WPR + i]; sm += shd1[m * WPR + i]; }
      float var;
      if (XK == X_LN) {
        const double md = sm / (double)K;
        double vd = ss / (double)K - md * md;
        vd = vd > 0.0 ? vd : 0.0;
        mean = (float)md; var = (float)vd;
      } else {
        var = (float)(ss / (double)K);
      }
      r = div_rn(1.0f, __fsqrt_rn(add_rn(var, p.eps)));
    }
    if (XK == X_Y32) {
      amax = rowok ? __ldcg(p.x_amax + m) : 0.f;
    } else {
      if (rowok) {
        for (int ch = tr; ch < nchunk; ch += RT) {
          float y[8];
          fl_load_y<XK>(p, xrow, ch, r, mean, y);
#pragma unroll
          for (int j = 0; j < 8; ++j) amax = fmaxf(amax, fabsf(y[j]));
        }
      }
      amax = warp_max(amax);
      if (lane == 0) shf[warp] = amax;
      __syncthreads();
      amax = 0.f;
#pragma unroll
      for (int i = 0; i < WPR; ++i) amax = fmaxf(amax, shf[m * WPR + i]);
    }
    const bool nz = amax > 0.0f;
    const float inv = nz ? div_rn(127.0f, amax) : 0.0f;
    if (tr == 0) s_scale[m] = nz ? div_rn(amax, 127.0f) : 0.0f;
    for (int ch = tr; ch < nchunk; ch += RT) {
      uint2 o = make_uint2(0, 0);
      if (rowok) {
        float y[8];
        fl_load_y<XK>(p, xrow, ch, r, mean, y);
        o.x = pack4(quant1(y[0], inv), quant1(y[1], inv), quant1(y[2], inv), quant1(y[3], inv));
        o.y = pack4(quant1(y[4], inv), quant1(y[5], inv), quant1(y[6], inv), quant1(y[7], inv));
      }
      qrow[ch] = o;
    }
  } else
#pragma unroll 1
  for (int m = 0; m < MT; ++m) {
    uint2* qrow = reinterpret_cast<uint2*>(xs + (size_t)m * XS);
    if (m >= p.M) {
      for (int ch = threadIdx.x; ch < nchunk; ch += FL_THREADS) qrow[ch] = make_uint2(0, 0);
      continue;
    }
    const void* xrow = (XK == X_Y32)
        ? static_cast<const void*>(reinterpret_cast<const float*>(p.x) + (long long)m * p.x_ld)
        : static_cast<const void*>(reinterpret_cast<const __half*>(p.x) + (long long)m * p.x_ld);
    float r = 0.f, mean = 0.f, amax = 0.f;
    if (fast) {
      // single pass: the row lives in registers between moments, absmax and quantise
      float v[FP_MAXC][8];
      double s1 = 0.0, s2 = 0.0;
#pragma unroll
      for (int c = 0; c < FP_MAXC; ++c) {
        const int ch = threadIdx.x + c * FL_THREADS;
        if (ch < nchunk) {
          const uint4 u = ld_cg_v4(reinterpret_cast<const __half*>(xrow) + ch * 8);
          const float2 a0 = h2_to_f2(u.x), a1 = h2_to_f2(u.y), a2 = h2_to_f2(u.z), a3 = h2_to_f2(u.w);
          v[c][0] = a0.x; v[c][1] = a0.y; v[c][2] = a1.x; v[c][3] = a1.y;
          v[c][4] = a2.x; v[c][5] = a2.y; v[c][6] = a3.x; v[c][7] = a3.y;
          if (XK == X_RMS || XK == X_LN) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const double dd = (double)v[c][j];
              s2 += dd * dd;
              if (XK == X_LN) s1 += dd;
            }
          }
        }
      }
      if (XK == X_RMS || XK == X_LN) {
        const double ss = fl_block_sum_d(s2, shd);
        float var;
        if (XK == X_LN) {
          const double sm = fl_block_sum_d(s1, shd);
          const double md = sm / (double)K;
          double vd = ss / (double)K - md * md;
          vd = vd > 0.0 ? vd : 0.0;
          mean = (float)md; var = (float)vd;
        } else {
          var = (float)(ss / (double)K);
        }
        r = div_rn(1.0f, __fsqrt_rn(add_rn(var, p.eps)));
      }
#pragma unroll
      for (int c = 0; c < FP_MAXC; ++c) {
        const int ch = threadIdx.x + c * FL_THREADS;
        if (ch < nchunk) {
          float gg[8], bb[8], sv[8];
          if (XK == X_RMS || XK == X_LN) {
            const uint4 g = *reinterpret_cast<const uint4*>(sg + ch * 8);
            const float2 g0 = h2_to_f2(g.x), g1 = h2_to_f2(g.y), g2 = h2_to_f2(g.z), g3 = h2_to_f2(g.w);
            gg[0] = g0.x; gg[1] = g0.y; gg[2] = g1.x; gg[3] = g1.y; gg[4] = g2.x; gg[5] = g2.y; gg[6] = g3.x; gg[7] = g3.y;
          }
          if (XK == X_LN) {
            const uint4 g = *reinterpret_cast<const uint4*>(sb + ch * 8);
            const float2 g0 = h2_to_f2(g.x), g1 = h2_to_f2(g.y), g2 = h2_to_f2(g.z), g3 = h2_to_f2(g.w);
            bb[0] = g0.x; bb[1] = g0.y; bb[2] = g1.x; bb[3] = g1.y; bb[4] = g2.x; bb[5] = g2.y; bb[6] = g3.x; bb[7] = g3.y;
          }
          if (p.smooth != nullptr) {
            const float4 sA = *reinterpret_cast<const float4*>(ssm + ch * 8);
            const float4 sB = *reinterpret_cast<const float4*>(ssm + ch * 8 + 4);
            sv[0] = sA.x; sv[1] = sA.y; sv[2] = sA.z; sv[3] = sA.w; sv[4] = sB.x; sv[5] = sB.y; sv[6] = sB.z; sv[7] = sB.w;
          }
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            float t = v[c][j];
            if (XK == X_RMS || XK == X_LN) {
              if (XK == X_LN) t = sub_rn(t, mean);
              t = mul_rn(mul_rn(t, r), gg[j]);
              if (XK == X_LN) t = add_rn(t, bb[j]);
            }
            if (p.smooth != nullptr) t = mul_rn(t, sv[j]);
            v[c][j] = t;
            amax = fmaxf(amax, fabsf(t));
          }
        }
      }
      amax = fl_block_max(amax, shf);
      const bool nzf = amax > 0.0f;
      const float invf = nzf ? div_rn(127.0f, amax) : 0.0f;
      if (threadIdx.x == 0) s_scale[m] = nzf ? div_rn(amax, 127.0f) : 0.0f;
#pragma unroll
      for (int c = 0; c < FP_MAXC; ++c) {
        const int ch = threadIdx.x + c * FL_THREADS;
        if (ch < nchunk) {
          uint2 o;
          o.x = pack4(quant1(v[c][0], invf), quant1(v[c][1], invf), quant1(v[c][2], invf), quant1(v[c][3], invf));
          o.y = pack4(quant1(v[c][4], invf), quant1(v[c][5], invf), quant1(v[c][6], invf), quant1(v[c][7], invf));
          qrow[ch] = o;
        }
      }
      continue;
    }
    if (XK == X_RMS || XK == X_LN) {
      double s1 = 0.0, s2 = 0.0;
      for (int ch = threadIdx.x; ch < nchunk; ch += FL_THREADS) {
        const uint4 u = ld_cg_v4(reinterpret_cast<const __half*>(xrow) + ch * 8);
        const uint32_t* pu = &u.x;
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const float2 f = h2_to_f2(pu[e]);
          const double d0 = (double)f.x, d1 = (double)f.y;
          s2 += d0 * d0; s2 += d1 * d1;
          if (XK == X_LN) { s1 += d0; s1 += d1; }
        }
      }
      const double ss = fl_block_sum_d(s2, shd);
      float var;
      if (XK == X_LN) {
        const double sm = fl_block_sum_d(s1, shd);
        const double md = sm / (double)K;
        double vd = ss / (double)K - md * md;
        vd = vd > 0.0 ? vd : 0.0;
        mean = (float)md; var = (float)vd;
      } else {
        var = (float)(ss / (double)K);
      }
      r = div_rn(1.0f, __fsqrt_rn(add_rn(var, p.eps)));
    }
    if (XK == X_Y32) {
      amax = __ldcg(p.x_amax + m);
    } else {
      for (int ch = threadIdx.x; ch < nchunk; ch += FL_THREADS) {
        float y[8];
        fl_load_y<XK>(p, xrow, ch, r, mean, y);
#pragma unroll
        for (int j = 0; j < 8; ++j) amax = fmaxf(amax, fabsf(y[j]));
      }
      amax = fl_block_max(amax, shf);
    }
    const bool nz = amax > 0.0f;
    const float inv = nz ? div_rn(127.0f, amax) : 0.0f;
    if (threadIdx.x == 0) s_scale[m] = nz ? div_rn(amax, 127.0f) : 0.0f;
    for (int ch = threadIdx.x; ch < nchunk; ch += FL_THREADS) {
      float y[8];
      fl_load_y<XK>(p, xrow, ch, r, mean, y);
      uint2 o;
      o.x = pack4(quant1(y[0], inv), quant1(y[1], inv), quant1(y[2], inv), quant1(y[3], inv));
      o.y = pack4(quant1(y[4], inv), quant1(y[5], inv), quant1(y[6], inv), quant1(y[7], inv));
      qrow[ch] = o;
    }
  }
  __syncthreads();
  tl_mark(TLK, 3);
  // ---- epilogue of ONE (item, token): v0 / v1 = the int32 sums of the item's two rows ------------
  auto epi_tok = [&](int it, int m, int v0i, int v1i) {
    if (OKIND == O_F16) {
      __half* out = reinterpret_cast<__half*>(p.out);
#pragma unroll
      for (int r = 0; r < FL_ROWS; ++r) {
        const int n = 2 * it + r;
        if (n < p.N) {
          const float ws = p.w_scale[n];
          const long long off = (long long)m * p.N + n;
          float v = mul_rn(mul_rn((float)(r ? v1i : v0i), s_scale[m]), ws);
          if (p.bias) v = add_rn(v, __half2float(p.bias[n]));
          if (p.residual) v = add_rn(v, __half2float(__ushort_as_half(__ldcg(reinterpret_cast<const unsigned short*>(p.residual + off)))));
          out[off] = __float2half_rn(v);
        }
      }
    } else if (OKIND == O_QKV) {
      __half* out = reinterpret_cast<__half*>(p.out);
      const int q_head = it / half_hd, q_i = it - q_head * half_hd;
      const int n0 = q_head * p.hd + q_i, n1 = n0 + half_hd;
      float v0 = mul_rn(mul_rn((float)v0i, s_scale[m]), p.w_scale[n0]);
      float v1 = mul_rn(mul_rn((float)v1i, s_scale[m]), p.w_scale[n1]);
      if (p.bias) { v0 = add_rn(v0, __half2float(p.bias[n0])); v1 = add_rn(v1, __half2float(p.bias[n1])); }
      __half h0 = __float2half_rn(v0), h1 = __float2half_rn(v1);     // the fp16 qkv of the unfused path
      const int pos = p.positions[m];
      if (q_head < p.H + p.KV && p.use_rope) {
        const float x1 = __half2float(h0), x2 = __half2float(h1);
        const float c = p.cosT[(long long)pos * half_hd + q_i], sn = p.sinT[(long long)pos * half_hd + q_i];
        h0 = __float2half_rn(sub_rn(mul_rn(x1, c), mul_rn(x2, sn)));
        h1 = __float2half_rn(add_rn(mul_rn(x2, c), mul_rn(x1, sn)));
      }
      if (q_head < p.H) {
        out[(long long)m * p.N + n0] = h0;
        out[(long long)m * p.N + n1] = h1;
      } else {
        const bool isk = q_head < p.H + p.KV;
        const int kvh = isk ? q_head - p.H : q_head - p.H - p.KV;
        const int row = p.seq_ids ? p.seq_ids[m] : p.seq0;
        __half* dst = (isk ? p.kcache : p.vcache) + (((long long)row * p.KV + kvh) * p.Smax + pos) * p.hd;
        dst[q_i] = h0;
        dst[q_i + half_hd] = h1;
      }
    } else {
      float* out = reinterpret_cast<float*>(p.out);
      float vg = mul_rn(mul_rn((float)v0i, s_scale[m]), p.w_scale[it]);
      float vu = mul_rn(mul_rn((float)v1i, s_scale[m]), p.w_scale[F + it]);
      if (p.bias) { vg = add_rn(vg, __half2float(p.bias[it])); vu = add_rn(vu, __half2float(p.bias[F + it])); }
      const float g = __half2float(__float2half_rn(vg));      // the fp16 rounding of the unfused path
      const float uu = __half2float(__float2half_rn(vu));
      const float a = (p.act == 0) ? silu_det(g) : gelu_new_det(g);
      float y = mul_rn(a, uu);
      if (p.out_smooth) y = mul_rn(y, p.out_smooth[it]);
      out[(long long)m * F + it] = y;
      atomicMax(reinterpret_cast<int*>(p.out_amax + m), __float_as_int(fabsf(y)));   // |y| >= 0: int order == float order
    }
  };
  if constexpr (!TCM) {
  // ---- 4a. one token: persistent dp4a weight stream, two batches always in flight per lane ------
  int acc[FL_ROWS][MT];
#pragma unroll
  for (int r = 0; r < FL_ROWS; ++r)
#pragma unroll
    for (int m = 0; m < MT; ++m) acc[r][m] = 0;
  auto consume = [&](const uint4 (&src)[FL_UNR][FL_ROWS], int j) {
#pragma unroll
    for (int u = 0; u < FL_UNR; ++u) {
      const int idx = lane + j * STEP + u * 32;
      if (idx < kvec) {
#pragma unroll
        for (int m = 0; m < MT; ++m) {
          const uint4 x = reinterpret_cast<const uint4*>(xs)[m * kvec + idx];
#pragma unroll
          for (int r = 0; r < FL_ROWS; ++r) {
            int a = acc[r][m];
            a = __dp4a((int)src[u][r].x, (int)x.x, a);
            a = __dp4a((int)src[u][r].y, (int)x.y, a);
            a = __dp4a((int)src[u][r].z, (int)x.z, a);
            a = __dp4a((int)src[u][r].w, (int)x.w, a);
            acc[r][m] = a;
          }
        }
      }
    }
  };
  // Deferred, lane-parallel epilogue: a finished item's reduced sums are parked in lane
  // (item count mod 32); when 32 are parked (or at the end) every lane finishes its own item, so the
  // dependent scalar loads, the SiLU and the stores of up to 32 items overlap instead of serialising
  // in one lane while the weight stream waits.
  int res[FL_ROWS][MT];
  int res_it = -1, parked = 0;
  auto flush = [&]() {
    if (res_it >= 0) {
#pragma unroll
      for (int m = 0; m < MT; ++m)
        if (m < p.M) epi_tok(res_it, m, res[0][m], res[1][m]);
    }
    res_it = -1;
    parked = 0;
  };
  auto finish_item = [&](int it) {
#pragma unroll
    for (int r = 0; r < FL_ROWS; ++r)
#pragma unroll
      for (int m = 0; m < MT; ++m) {
        const int v = warp_sum_i(acc[r][m]);
        if (lane == parked) res[r][m] = v;
        acc[r][m] = 0;
      }
    if (lane == parked) res_it = it;
    if (++parked == 32) flush();
  };
  // one step: reduce the batch under the compute cursor, immediately re-arm its registers with the
  // batch two ahead (the load cursor), then close the item if that was its last batch
  auto step = [&](uint4 (&buf)[FL_UNR][FL_ROWS]) {
    consume(buf, cc.j);
    load_batch(buf, lc);
    advance(lc);
    const int it = cc.it;
    const bool last = cc.j == NB - 1;
    advance(cc);
    if (last) finish_item(it);
  };
  while (cc.it < items_total) {
    step(w);
    if (cc.it >= items_total) break;
    step(w2);
  }
  tl_mark(TLK, 4);
  flush();
  } else {
  // ---- 4b. 2..8 tokens: the same register-double-buffered stream feeding mma.sync.m16n8k32 -------
  // (Round 2, measured and dropped: 256-bit loads -- 32 bytes per lane, so one warp-wide request covers 8 rows x a
  //  full 128-byte line instead of 8 x 64 bytes, same register batch -- left every shape where it was: gate_up
  //  28.7 vs 28.3 us, qkv 19.2 vs 18.9, o 10.3 vs 10.0, down 18.5 vs 19.3 at 8 tokens.  The request count is not
  //  what holds the batch stream at half the one-token rate.)
  // With dp4a a batch of 8 tokens costs 32 integer instructions and 8 shared-memory loads per 16 weight
  // bytes per row, and the stream turns compute-bound (gate_up of Llama-2-7B: 45 us against 16 at one
  // token).  Here the 16 rows of a unit are the M dimension of the tile and the tokens its N dimension:
  // lane (g, t) holds 16 contiguous bytes of rows g and g+8, which are two k-slices of its A fragment
  // for two MMAs (the k order inside a tile is free as long as A and B agree, and int32 accumulation
  // is exact in any order: results stay bit-identical to the dp4a path), and the matching 16 bytes of
  // token g's activation row are its B fragments: 1 shared load + 2 MMAs per 2 weight loads.
  // The accumulator fragment is already one value per (row, token): no shuffle reduction.  The 16
  // warps' partial sums (each warp has 1/16 of K) meet in shared memory; 64 threads add them up and
  // run the epilogue of one (item, token) each while the others stream on.
  int c4[4] = {0, 0, 0, 0};
  int ui = 0;
  // Deferred epilogue, as in 4a: the reduced sums of up to 8 units are parked in shared memory, then
  // all 512 threads finish one (item, token) each, so the dependent scale loads, the SiLU and the
  // stores of 8 units overlap instead of holding two warps -- and with them every unit's barrier -- back.
  __shared__ int s_park[8][64][2];
  const int unit0 = blockIdx.x;
  auto flush_units = [&](int first_ui, int count) {       // units first_ui .. first_ui+count-1 (CTA-local numbering)
    __syncthreads();
    const int slot = threadIdx.x >> 6, e = threadIdx.x & 63;
    if (slot < count) {
      const int g = e >> 3, m = e & 7;
      const int it = (unit0 + (first_ui + slot) * stride) * 8 + g;
      if (it < items_total && m < p.M) epi_tok(it, m, s_park[slot][e][0], s_park[slot][e][1]);
    }
  };
  auto consume = [&](const uint4 (&src)[FL_UNR][FL_ROWS], int j) {
#pragma unroll
    for (int u = 0; u < FL_UNR; ++u) {
      const int q4 = 4 * (warp + FL_WARPS * (u + FL_UNR * j));
      if (q4 < kvec) {                                   // warp-uniform
        const int idx = q4 + t4;
        uint4 b = make_uint4(0, 0, 0, 0);
        if (g8 < MT && idx < kvec) b = *reinterpret_cast<const uint4*>(xs + (size_t)g8 * XS + (size_t)idx * 16);
        fl_mma_s8(c4, src[u][0].x, src[u][1].x, src[u][0].y, src[u][1].y, b.x, b.y);
        fl_mma_s8(c4, src[u][0].z, src[u][1].z, src[u][0].w, src[u][1].w, b.z, b.w);
      }
    }
  };
  auto finish_unit = [&](int unit) {
    int4* pb = reinterpret_cast<int4*>(part) + (ui & 1) * (FL_WARPS * 32);
    pb[warp * 32 + lane] = make_int4(c4[0], c4[1], c4[2], c4[3]);
    c4[0] = c4[1] = c4[2] = c4[3] = 0;
    ++ui;
    // One barrier per unit: the buffer written now is read below and written again two units later,
    // after every thread has passed the next unit's barrier.
    __syncthreads();
    if (threadIdx.x < 64) {
      const int g = threadIdx.x >> 3, m = threadIdx.x & 7;          // accumulator (row g | g+8, token m) sits in lane 4g + m/2
      const int* pi = reinterpret_cast<const int*>(pb) + (g * 4 + (m >> 1)) * 4 + (m & 1);
      int v0 = 0, v1 = 0;
#pragma unroll
      for (int w_ = 0; w_ < FL_WARPS; ++w_) { v0 += pi[w_ * 128]; v1 += pi[w_ * 128 + 2]; }
      s_park[(ui - 1) & 7][threadIdx.x][0] = v0;
      s_park[(ui - 1) & 7][threadIdx.x][1] = v1;
    }
    if ((ui & 7) == 0) flush_units(ui - 8, 8);
  };
  auto step = [&](uint4 (&buf)[FL_UNR][FL_ROWS]) {
    consume(buf, cc.j);
    load_batch(buf, lc);
    advance(lc);
    const int unit = cc.it;
    const bool last = cc.j == NB - 1;
    advance(cc);
    if (last) finish_unit(unit);
  };
  while (cc.it < work_total) {
    step(w);
    if (cc.it >= work_total) break;
    step(w2);
  }
  tl_mark(TLK, 4);
  if (ui & 7) flush_units(ui & ~7, ui & 7);
  }
  tl_mark(TLK, 5);
  tl_flush(TLK, 0xFFu);
}

struct SyncPDL {
  B2_DEVINL void operator()() const { pdl_wait(); }
};
// Barrier across all CTAs of a persistent launch (all co-resident: the launcher sizes the grid
// from the occupancy).  `ctr` starts at 0 (zeroed by the caller before the launch) and is used once.
// The spin is bounded so that a broken residency assumption ends as an error flag, not as a hung GPU.
struct SyncGrid {
  unsigned* ctr;
  unsigned* err;
  B2_DEVINL void operator()() const {
    __syncthreads();
    if (threadIdx.x == 0) {
      __threadfence();
      atomicAdd(ctr, 1u);
      unsigned v, spins = 0;
      do {
        asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(ctr) : "memory");
        if (v >= gridDim.x) break;
        // no back-off: the wait is short and latency matters more than the polling traffic
      } while (++spins < (1u << 24));
      if (v < gridDim.x && err) atomicExch(err, 1u);
      __threadfence();
    }
    __syncthreads();
  }
};

template <int MT, int XK, int OKIND>
__global__ void __launch_bounds__(FL_THREADS, FL_CTAS_PER_SM) fused_gemv_kernel(const FusedParams p) {
  extern __shared__ __align__(16) uint8_t xs[];        // int8 activations [MT][K] + staging
  fused_phase<MT, XK, OKIND, true>(p, xs, SyncPDL{}, true);
}

// The dependent tail of a decoder layer in ONE persistent launch:
//   o_proj (+residual) | gate_up (+act*up) | down (+residual) | [next layer's qkv (+rope, KV append)]
// Each phase issues its first weight batches and its L2 hints BEFORE the grid barrier that ends the
// previous phase, so the HBM stream does not stop at the three dependency points.
template <int MT, int NORMK, bool QKV>
__global__ void __launch_bounds__(FL_THREADS, FL_CTAS_PER_SM)
fused_tail_kernel(const FusedParams po, const FusedParams pg, const FusedParams pd, const FusedParams pq, unsigned* ctr) {
  extern __shared__ __align__(16) uint8_t xs[];
  unsigned* err = ctr + 3;
  fused_phase<MT, X_PLAIN, O_F16, true>(po, xs, SyncPDL{}, false);
  fused_phase<MT, NORMK, O_GATED, false>(pg, xs, SyncGrid{ctr + 0, err}, false);
  fused_phase<MT, X_Y32, O_F16, false>(pd, xs, SyncGrid{ctr + 1, err}, !QKV);
  if (QKV) fused_phase<MT, NORMK, O_QKV, false>(pq, xs, SyncGrid{ctr + 2, err}, true);
}

static int fl_sm_count() { return device_sm_count(); }

template <int MT, int XK, int OKIND>
static int launch_fused(const FusedParams& pin, cudaStream_t st) {
  size_t smem = fl_act_bytes(MT, pin.K);
  if (XK != X_Y32 && (pin.K >> 3) <= 3 * FL_THREADS) smem += (size_t)8 * pin.K;   // gamma, beta (fp16) + smooth (fp32) staging
  if (smem > (size_t)FL_SMEM_FULL) return ERR_TOO_LARGE;
  const FusedParams& p = pin;
  static DeviceOnce attr_once;
  if (!attr_once.get()) {
    if (cudaFuncSetAttribute(fused_gemv_kernel<MT, XK, OKIND>, cudaFuncAttributeMaxDynamicSharedMemorySize, FL_SMEM_FULL) != cudaSuccess)
      return ERR_ARCH;
    attr_once.set();
  }
  const int items = (OKIND == O_F16) ? (p.N + 1) / 2 : p.N / 2;
  int grid = MT > 1 ? (items + 7) / 8 : (items + FL_WARPS - 1) / FL_WARPS;    // MT > 1: the whole CTA works on 8 items at a time
  const int cap = fl_sm_count() * FL_CTAS_PER_SM;                   // persistent: one 512-thread CTA per SM (grids of 104..140 CTAs measured slower)
  if (grid > cap) grid = cap;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3(FL_THREADS);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  if (cudaLaunchKernelEx(&cfg, fused_gemv_kernel<MT, XK, OKIND>, p) != cudaSuccess) {
    cudaGetLastError();
    return ERR_LAUNCH;
  }
  return OK;
}

template <int XK, int OKIND>
static int dispatch_m(const FusedParams& p, cudaStream_t st) {
  if (p.M == 1) return launch_fused<1, XK, OKIND>(p, st);
  if (p.M == 2) return launch_fused<2, XK, OKIND>(p, st);
  if (p.M <= 4) return launch_fused<4, XK, OKIND>(p, st);
  return launch_fused<8, XK, OKIND>(p, st);
}
template <int OKIND>
static int dispatch_x(int xk, const FusedParams& p, cudaStream_t st) {
  switch (xk) {
    case X_RMS: return dispatch_m<X_RMS, OKIND>(p, st);
    case X_LN: return dispatch_m<X_LN, OKIND>(p, st);
    case X_PLAIN: return dispatch_m<X_PLAIN, OKIND>(p, st);
    case X_Y32: return dispatch_m<X_Y32, OKIND>(p, st);
    default: return ERR_UNSUPPORTED;
  }
}

int fused_linear(const void* x, long long x_ld, int x_kind, const void* gamma, const void* beta, const float* smooth,
                 const float* x_amax, float eps, const int8_t* W, const float* w_scale, const void* bias,
                 const void* residual, void* out, int out_kind, int act, const float* out_smooth, float* out_amax,
                 float* zero_amax, int M, int N, int K, const void* pf0, long long pf0_bytes, const void* pf1,
                 long long pf1_bytes, cudaStream_t st) {
  if (M == 0 || N == 0) return OK;
  if (!x || !W || !w_scale || !out) return ERR_NULL_PTR;
  if (M < 0 || M > 8) return ERR_UNSUPPORTED;
  if (N < 0 || K <= 0 || (K & 15) || (x_ld & 7)) return ERR_BAD_SHAPE;
  if ((x_kind == X_RMS || x_kind == X_LN) && !gamma) return ERR_NULL_PTR;
  if (x_kind == X_LN && !beta) return ERR_NULL_PTR;
  if (x_kind == X_Y32 && !x_amax) return ERR_NULL_PTR;
  if (out_kind == O_GATED && ((N & 1) || !out_amax || residual)) return ERR_BAD_SHAPE;
  auto mis = [](const void* q) { return q && (reinterpret_cast<uintptr_t>(q) & 15); };
  if (mis(x) || mis(W) || mis(gamma) || mis(beta) || mis(smooth)) return ERR_BAD_ALIGN;
  FusedParams p = {};
  p.x = x; p.x_ld = x_ld; p.gamma = (const __half*)gamma; p.beta = (const __half*)beta; p.smooth = smooth;
  p.x_amax = x_amax; p.eps = eps; p.W = W; p.w_scale = w_scale; p.bias = (const __half*)bias;
  p.residual = (const __half*)residual; p.out = out; p.out_smooth = out_smooth; p.out_amax = out_amax;
  p.zero_amax = zero_amax; p.M = M; p.N = N; p.K = K; p.act = act;
  p.pf0 = (const uint8_t*)pf0; p.pf0_bytes = pf0_bytes; p.pf1 = (const uint8_t*)pf1; p.pf1_bytes = pf1_bytes;
  if (out_kind == O_F16) return dispatch_x<O_F16>(x_kind, p, st);
  if (out_kind == O_GATED) return dispatch_x<O_GATED>(x_kind, p, st);
  return ERR_UNSUPPORTED;
}

// qkv projection with fused rotary embedding + KV append (decode, M <= 8)
int fused_qkv(const void* x, long long x_ld, int x_kind, const void* gamma, const void* beta, const float* smooth,
              float eps, const int8_t* W, const float* w_scale, const void* bias, void* qkv_out, const float* cosT,
              const float* sinT, const int* positions, const int* seq_ids, int seq0, void* kcache, void* vcache, int M,
              int H, int KV, int hd, int Smax, int K, int use_rope, const void* pf0, long long pf0_bytes,
              const void* pf1, long long pf1_bytes, cudaStream_t st) {
  if (M == 0) return OK;
  if (!x || !W || !w_scale || !qkv_out || !positions || !kcache || !vcache) return ERR_NULL_PTR;
  if (use_rope && (!cosT || !sinT)) return ERR_NULL_PTR;
  if (M < 0 || M > 8) return ERR_UNSUPPORTED;
  if (x_kind != X_RMS && x_kind != X_LN && x_kind != X_PLAIN) return ERR_UNSUPPORTED;
  if (H <= 0 || KV <= 0 || hd <= 0 || (hd & 1) || K <= 0 || (K & 15) || (x_ld & 7)) return ERR_BAD_SHAPE;
  if ((x_kind == X_RMS || x_kind == X_LN) && !gamma) return ERR_NULL_PTR;
  if (x_kind == X_LN && !beta) return ERR_NULL_PTR;
  auto mis = [](const void* q) { return q && (reinterpret_cast<uintptr_t>(q) & 15); };
  if (mis(x) || mis(W) || mis(gamma) || mis(beta) || mis(smooth)) return ERR_BAD_ALIGN;
  FusedParams p = {};
  p.x = x; p.x_ld = x_ld; p.gamma = (const __half*)gamma; p.beta = (const __half*)beta; p.smooth = smooth;
  p.eps = eps; p.W = W; p.w_scale = w_scale; p.bias = (const __half*)bias; p.out = qkv_out;
  p.M = M; p.N = (H + 2 * KV) * hd; p.K = K;
  p.cosT = cosT; p.sinT = sinT; p.positions = positions; p.seq_ids = seq_ids; p.seq0 = seq0;
  p.kcache = (__half*)kcache; p.vcache = (__half*)vcache; p.H = H; p.KV = KV; p.hd = hd; p.Smax = Smax;
  p.use_rope = use_rope;
  p.pf0 = (const uint8_t*)pf0; p.pf0_bytes = pf0_bytes; p.pf1 = (const uint8_t*)pf1; p.pf1_bytes = pf1_bytes;
  switch (x_kind) {
    case X_RMS: return dispatch_m<X_RMS, O_QKV>(p, st);
    case X_LN: return dispatch_m<X_LN, O_QKV>(p, st);
    default: return dispatch_m<X_PLAIN, O_QKV>(p, st);
  }
}

// ---------------------------------------------------------------------------------------------
// multi-phase tail
// ---------------------------------------------------------------------------------------------
static int desc_to_params(const b2_fused_desc* d, int M, int want_xk_mask, int want_ok, FusedParams* out) {
  if (!d || !d->x || !d->W || !d->w_scale || !d->out) return ERR_NULL_PTR;
  if (d->out_kind != want_ok || !((1 << d->x_kind) & want_xk_mask)) return ERR_UNSUPPORTED;
  if (d->N <= 0 || d->K <= 0 || (d->K & 15) || (d->x_ld & 7)) return ERR_BAD_SHAPE;
  if ((d->x_kind == X_RMS || d->x_kind == X_LN) && !d->gamma) return ERR_NULL_PTR;
  if (d->x_kind == X_LN && !d->beta) return ERR_NULL_PTR;
  if (d->x_kind == X_Y32 && !d->x_amax) return ERR_NULL_PTR;
  if (d->out_kind == O_GATED && ((d->N & 1) || !d->out_amax || d->residual)) return ERR_BAD_SHAPE;
  if (d->out_kind == O_QKV && (!d->positions || !d->kcache || !d->vcache || (d->use_rope && (!d->cos_table || !d->sin_table))))
    return ERR_NULL_PTR;
  FusedParams p = {};
  p.x = d->x; p.x_ld = d->x_ld; p.gamma = (const __half*)d->gamma; p.beta = (const __half*)d->beta; p.smooth = d->smooth;
  p.x_amax = d->x_amax; p.eps = d->eps; p.W = d->W; p.w_scale = d->w_scale; p.bias = (const __half*)d->bias;
  p.residual = (const __half*)d->residual; p.out = d->out; p.out_smooth = d->out_smooth; p.out_amax = d->out_amax;
  p.zero_amax = d->zero_amax; p.M = M; p.N = d->N; p.K = d->K; p.act = d->act;
  p.cosT = d->cos_table; p.sinT = d->sin_table; p.positions = d->positions; p.seq_ids = d->seq_ids; p.seq0 = d->seq0;
  p.kcache = (__half*)d->kcache; p.vcache = (__half*)d->vcache; p.H = d->H; p.KV = d->KV; p.hd = d->hd; p.Smax = d->Smax;
  p.use_rope = d->use_rope;
  p.pf0 = (const uint8_t*)d->prefetch0; p.pf0_bytes = d->prefetch0_bytes;
  p.pf1 = (const uint8_t*)d->prefetch1; p.pf1_bytes = d->prefetch1_bytes;
  if (d->out_kind == O_QKV && p.N != (p.H + 2 * p.KV) * p.hd) return ERR_BAD_SHAPE;
  *out = p;
  return OK;
}

static size_t phase_smem(int MT, int xk, int K) {
  size_t smem = fl_act_bytes(MT, K);
  if (xk != X_Y32 && (K >> 3) <= 3 * FL_THREADS) smem += (size_t)8 * K;
  return smem;
}

template <int MT, int NORMK, bool QKV>
static int launch_tail(const FusedParams& po, const FusedParams& pg, const FusedParams& pd, const FusedParams& pq,
                       unsigned* ctr, cudaStream_t st) {
  size_t smem = phase_smem(MT, X_PLAIN, po.K);
  smem = std::max(smem, phase_smem(MT, NORMK, pg.K));
  smem = std::max(smem, phase_smem(MT, X_Y32, pd.K));
  if (QKV) smem = std::max(smem, phase_smem(MT, NORMK, pq.K));
  if (smem > (size_t)FL_SMEM_FULL) return ERR_TOO_LARGE;
  auto kern = fused_tail_kernel<MT, NORMK, QKV>;
  static DeviceOnce attr_once;
  if (!attr_once.get()) {
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, FL_SMEM_FULL) != cudaSuccess) return ERR_ARCH;
    attr_once.set();
  }
  // all CTAs must be resident at once (grid barriers): size the grid from the occupancy
  int occ = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, FL_THREADS, smem) != cudaSuccess || occ < 1) return ERR_ARCH;
  if (occ > FL_CTAS_PER_SM) occ = FL_CTAS_PER_SM;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(fl_sm_count() * occ);
  cfg.blockDim = dim3(FL_THREADS);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  if (cudaLaunchKernelEx(&cfg, kern, po, pg, pd, pq, ctr) != cudaSuccess) {
    cudaGetLastError();
    return ERR_LAUNCH;
  }
  return OK;
}

template <int MT>
static int dispatch_tail(int normk, bool qkv, const FusedParams& po, const FusedParams& pg, const FusedParams& pd,
                         const FusedParams& pq, unsigned* ctr, cudaStream_t st) {
  if (normk == X_RMS) return qkv ? launch_tail<MT, X_RMS, true>(po, pg, pd, pq, ctr, st) : launch_tail<MT, X_RMS, false>(po, pg, pd, pq, ctr, st);
  return qkv ? launch_tail<MT, X_LN, true>(po, pg, pd, pq, ctr, st) : launch_tail<MT, X_LN, false>(po, pg, pd, pq, ctr, st);
}

int fused_layer_tail(const b2_fused_desc* o, const b2_fused_desc* gu, const b2_fused_desc* down, const b2_fused_desc* qkv,
                     int M, unsigned* counters, cudaStream_t st) {
  if (M == 0) return OK;
  if (!counters) return ERR_NULL_PTR;
  if (M < 0 || M > 8) return ERR_UNSUPPORTED;
  FusedParams po, pg, pd, pq = {};
  int rc;
  if ((rc = desc_to_params(o, M, 1 << X_PLAIN, O_F16, &po))) return rc;
  if ((rc = desc_to_params(gu, M, (1 << X_RMS) | (1 << X_LN), O_GATED, &pg))) return rc;
  if ((rc = desc_to_params(down, M, 1 << X_Y32, O_F16, &pd))) return rc;
  if (qkv && (rc = desc_to_params(qkv, M, 1 << gu->x_kind, O_QKV, &pq))) return rc;
  const int normk = gu->x_kind;
  if (M == 1) return dispatch_tail<1>(normk, qkv != nullptr, po, pg, pd, pq, counters, st);
  if (M == 2) return dispatch_tail<2>(normk, qkv != nullptr, po, pg, pd, pq, counters, st);
  if (M <= 4) return dispatch_tail<4>(normk, qkv != nullptr, po, pg, pd, pq, counters, st);
  return dispatch_tail<8>(normk, qkv != nullptr, po, pg, pd, pq, counters, st);
}

template <int MT>
static bool fl_touch_m() {
  cudaFuncAttributes fa;
  bool ok = true;
#define B2_T(XK, OKIND) ok &= cudaFuncGetAttributes(&fa, fused_gemv_kernel<MT, XK, OKIND>) == cudaSuccess && \
    cudaFuncSetAttribute(fused_gemv_kernel<MT, XK, OKIND>, cudaFuncAttributeMaxDynamicSharedMemorySize, FL_SMEM_FULL) == cudaSuccess;
  B2_T(X_RMS, O_F16) B2_T(X_LN, O_F16) B2_T(X_PLAIN, O_F16) B2_T(X_Y32, O_F16)
  B2_T(X_RMS, O_GATED) B2_T(X_LN, O_GATED) B2_T(X_PLAIN, O_GATED) B2_T(X_Y32, O_GATED)
  B2_T(X_RMS, O_QKV) B2_T(X_LN, O_QKV) B2_T(X_PLAIN, O_QKV)
#undef B2_T
  return ok;
}
int init_fused() { fl_sm_count(); return (fl_touch_m<1>() && fl_touch_m<2>() && fl_touch_m<4>() && fl_touch_m<8>()) ? OK : ERR_ARCH; }

}  // namespace b2

B2_TL_DEFINE_DUMP(b2_tl_dump_fused)
