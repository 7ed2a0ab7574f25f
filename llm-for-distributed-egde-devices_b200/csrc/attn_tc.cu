// Causal prefill attention on the 5th-generation tensor cores (tcgen05 + TMEM + TMA), head_dim 128.
//
// Same contract as prefill_attn_kernel in attn.cu (which stays for head_dim 64 and sliding
// windows): one chunk of T query tokens of one sequence at positions [pos0, pos0+T), keys
// 0..pos0+T-1 read from the stage-local KV cache.  Replaces eager_attention_forward at q_len > 1
// (transformers modeling_llama.py:199-222) inside the prefill forward of `generate`
// (reference call site Code/Quantised Models/models_quant_updated.py:84-93).
//
// One CTA = 128 query rows of one head, three warpgroups:
//   warp 0      TMA producer: Q once, then K/V tiles of 128 keys through a 2-stage ring
//   warp 1      MMA issuer (one thread) + TMEM owner:  S = Q K^T  (128x128x128, kind::f16, fp32 in TMEM)
//               and PV = P V (P from shared memory, V as an MN-major B operand straight from the
//               cache layout); S, P and PV are double-buffered
//   warps 4-7   softmax group A: the EVEN key tiles;   warps 8-11  softmax group B: the ODD key tiles.
//               Thread = query row = TMEM lane.  Two passes over S with tcgen05.ld (row max, then
//               exp2 / row sum / fp16 P written in the 128B-swizzled K-major layout the MMA reads),
//               then O = O*corr + PV with O kept in registers (128 fp32 per thread) -- no TMEM
//               read-modify-write for the online-softmax rescale.  The two groups keep separate
//               (max, sum, O) states over their halves of the keys, so one group's exponentials
//               run under the other group's PV product; the states are merged once at the end
//               (as for split-KV decode).  384 threads x 168 registers: with 16-column TMEM reads the
//               softmax threads fit without setmaxnreg (which made the TMA/MMA warps spill instead).
// Numerics: fp32 scores and accumulation, P rounded to fp16 before the PV product, exp2f -- the
// same as the mma.sync kernel; compared with the oracle within the attention tolerance.
#include <cuda.h>
#include "common.cuh"

namespace b2 {

constexpr int TC_BM = 128;        // query rows per CTA
constexpr int TC_BN = 128;        // keys per tile
constexpr int TC_HD = 128;
constexpr int TC_THREADS = 384;
constexpr int TC_ATOM = TC_BM * 128;                 // one [128 rows][128 B] swizzle-atom column: 16 KB
constexpr int TC_TILE = 2 * TC_ATOM;                 // a 128 x 128 fp16 tile: 32 KB
constexpr int TC_KV_STAGES = 2;
// shared memory: Q | K[2] | V[2] | P[2] | barriers
constexpr int TC_OFF_Q = 0;
constexpr int TC_OFF_K = TC_OFF_Q + TC_TILE;
constexpr int TC_OFF_V = TC_OFF_K + TC_KV_STAGES * TC_TILE;
constexpr int TC_OFF_P = TC_OFF_V + TC_KV_STAGES * TC_TILE;
constexpr int TC_OFF_BAR = TC_OFF_P + 2 * TC_TILE;
constexpr int TC_SMEM = TC_OFF_BAR + 256 + 1024;     // + alignment slack

struct TcAttnParams {
  __half* out; long long o_ld;
  const int* pos_ptr;
  int T, pos0, row, H, KV, Smax, window;
  float scale_log2;
};

// MN-major operand (rows = K index, 128 contiguous bytes = 64 MN elements, 128B swizzle): 8-row groups
// 1024 B apart (SBO), the next 64 MN elements one 16 KB atom column further (LBO).
B2_DEVINL uint64_t umma_desc_sw128_mn(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr >> 4) & 0x3FFF);
  d |= (uint64_t)(TC_ATOM >> 4) << 16;
  d |= (uint64_t)(1024 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}

B2_DEVINL float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// Packed fp32x2 arithmetic (sm_100): the softmax warps are issue-bound, and one FFMA2 / FADD2 does the work of two
// scalar instructions.  (d0,d1) = (a0,a1) * b + (c0,c1);  (d0,d1) += (a0,a1).
B2_DEVINL void ffma2(float& d0, float& d1, float a0, float a1, float b, float c0, float c1) {
  asm("{\n\t.reg .b64 ra, rb, rc, rd;\n\t"
      "mov.b64 ra, {%2, %3};\n\tmov.b64 rb, {%4, %4};\n\tmov.b64 rc, {%5, %6};\n\t"
      "fma.rn.f32x2 rd, ra, rb, rc;\n\tmov.b64 {%0, %1}, rd;\n\t}"
      : "=f"(d0), "=f"(d1) : "f"(a0), "f"(a1), "f"(b), "f"(c0), "f"(c1));
}
B2_DEVINL void fadd2(float& d0, float& d1, float a0, float a1) {
  asm("{\n\t.reg .b64 ra, rd;\n\t"
      "mov.b64 ra, {%2, %3};\n\tmov.b64 rd, {%0, %1};\n\t"
      "add.rn.f32x2 rd, rd, ra;\n\tmov.b64 {%0, %1}, rd;\n\t}"
      : "+f"(d0), "+f"(d1) : "f"(a0), "f"(a1));
}

// WIN = sliding window (a separate instantiation: the softmax warps are issue-bound, and the two extra compares per
// 16 keys cost the full-attention path 10 % when they were unconditional)
template <bool WIN>
__global__ void __launch_bounds__(TC_THREADS, 1)
prefill_attn_tc_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                       const __grid_constant__ CUtensorMap tmV, const TcAttnParams pin) {
  extern __shared__ uint8_t smem_raw_tc[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw_tc) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + TC_OFF_BAR);
  uint64_t* q_full = bars + 0;
  uint64_t* k_full = bars + 1;      // [2]
  uint64_t* k_empty = bars + 3;     // [2]
  uint64_t* v_full = bars + 5;      // [2]
  uint64_t* v_empty = bars + 7;     // [2]
  uint64_t* s_full = bars + 9;      // [2]  MMA -> softmax
  uint64_t* s_empty = bars + 11;    // [2]  softmax -> MMA (4 warps)
  uint64_t* p_full = bars + 13;     // [2]  softmax -> MMA (4 warps)
  uint64_t* pv_full = bars + 15;    // [2]  MMA -> softmax
  uint64_t* pv_empty = bars + 17;   // [2]  softmax -> MMA (4 warps)
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 19);

  TcAttnParams p = pin;
  if (p.pos_ptr) p.pos0 = p.pos_ptr[0];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int qt = gridDim.x - 1 - blockIdx.x;            // heavy (late) query tiles first
  const int h = blockIdx.y;
  const int kvh = h / (p.H / p.KV);
  const int q0 = qt * TC_BM;
  const int last_q = min(q0 + TC_BM, p.T) - 1;
  const int nt_end = (p.pos0 + last_q) / TC_BN + 1;     // key tiles 0 .. nt_end-1 cover keys 0 .. pos0+last_q
  // sliding window (key > qpos - window): the first key any row of this tile needs is pos0 + q0 - window + 1; whole
  // tiles before it are skipped, the rest of the lower bound is masked per row.  Tiles are numbered from jt0 below.
  int jt0 = 0;
  if (WIN) { const int kf = p.pos0 + q0 - p.window + 1; jt0 = kf > 0 ? kf / TC_BN : 0; }
  const int nt = nt_end - jt0;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmQ); tma_prefetch_desc(&tmK); tma_prefetch_desc(&tmV);
  }
  if (warp == 1 && lane == 0) {
    mbar_init(q_full, 1);
    for (int i = 0; i < 2; ++i) {
      mbar_init(&k_full[i], 1); mbar_init(&k_empty[i], 1);
      mbar_init(&v_full[i], 1); mbar_init(&v_empty[i], 1);
      mbar_init(&s_full[i], 1); mbar_init(&s_empty[i], 4);
      mbar_init(&p_full[i], 4); mbar_init(&pv_full[i], 1); mbar_init(&pv_empty[i], 4);
    }
    mbar_fence_init();
  }
  if (warp == 1) tmem_alloc<512>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t tmem_S = tmem_base, tmem_O = tmem_base + 256;

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      mbar_expect_tx(q_full, TC_TILE);
      tma_load_2d(smem + TC_OFF_Q, &tmQ, q_full, h * TC_HD, q0);
      tma_load_2d(smem + TC_OFF_Q + TC_ATOM, &tmQ, q_full, h * TC_HD + 64, q0);
      const int kv_row0 = (p.row * p.KV + kvh) * p.Smax;
      for (int j = 0; j < nt; ++j) {
        const int s = j & 1;
        const uint32_t ph = (j >> 1) & 1;
        mbar_wait(&k_empty[s], ph ^ 1);
        mbar_expect_tx(&k_full[s], TC_TILE);
        tma_load_2d(smem + TC_OFF_K + s * TC_TILE, &tmK, &k_full[s], 0, kv_row0 + (jt0 + j) * TC_BN);
        tma_load_2d(smem + TC_OFF_K + s * TC_TILE + TC_ATOM, &tmK, &k_full[s], 64, kv_row0 + (jt0 + j) * TC_BN);
        mbar_wait(&v_empty[s], ph ^ 1);
        mbar_expect_tx(&v_full[s], TC_TILE);
        tma_load_2d(smem + TC_OFF_V + s * TC_TILE, &tmV, &v_full[s], 0, kv_row0 + (jt0 + j) * TC_BN);
        tma_load_2d(smem + TC_OFF_V + s * TC_TILE + TC_ATOM, &tmV, &v_full[s], 64, kv_row0 + (jt0 + j) * TC_BN);
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    if (lane == 0) {
      constexpr uint32_t idesc_qk = umma_idesc(IDESC_FMT_F32, IDESC_F16, IDESC_F16, TC_BM, TC_BN);
      constexpr uint32_t idesc_pv = umma_idesc(IDESC_FMT_F32, IDESC_F16, IDESC_F16, TC_BM, TC_HD) | (1u << 16);   // B is MN-major
      const uint32_t sQ = smem_u32(smem + TC_OFF_Q), sP = smem_u32(smem + TC_OFF_P);
      auto issue_qk = [&](int j) {
        const int s = j & 1;
        const uint32_t ph = (j >> 1) & 1;
        mbar_wait(&k_full[s], ph);
        mbar_wait(&s_empty[s], ph ^ 1);
        tc_fence_after();
        const uint32_t sK = smem_u32(smem + TC_OFF_K + s * TC_TILE);
#pragma unroll
        for (int k = 0; k < TC_HD / 16; ++k) {
          // k-step = 16 halves = 32 B inside a 128-B atom row; four steps per atom column
          const uint32_t off = (k >> 2) * TC_ATOM + (k & 3) * 32;
          umma_f16(tmem_S + s * TC_BN, umma_desc_sw128(sQ + off), umma_desc_sw128(sK + off), idesc_qk, k != 0);
        }
        umma_commit(&k_empty[s]);
        umma_commit(&s_full[s]);
      };
      mbar_wait(q_full, 0);
      issue_qk(0);
      if (nt > 1) issue_qk(1);
      for (int j = 0; j < nt; ++j) {
        const int s = j & 1;
        const uint32_t ph = (j >> 1) & 1;
        mbar_wait(&p_full[s], ph);
        mbar_wait(&v_full[s], ph);
        mbar_wait(&pv_empty[s], ph ^ 1);
        tc_fence_after();
        const uint32_t sV = smem_u32(smem + TC_OFF_V + s * TC_TILE);
#pragma unroll
        for (int k = 0; k < TC_BN / 16; ++k) {
          // A = P (K-major, K = keys): as above.  B = V (MN-major): a k-step is 16 key rows = 2048 B.
          const uint32_t aoff = (k >> 2) * TC_ATOM + (k & 3) * 32;
          umma_f16(tmem_O + s * TC_HD, umma_desc_sw128(sP + s * TC_TILE + aoff), umma_desc_sw128_mn(sV + k * 2048), idesc_pv,
                   k != 0);
        }
        umma_commit(&v_empty[s]);
        umma_commit(&pv_full[s]);
        if (j + 2 < nt) issue_qk(j + 2);       // S[s] was released before P[s] was announced
      }
    }
  } else if (warp >= 4) {
    // ===================== softmax / accumulate / merge / store =====================
    const int grp = (warp >> 2) - 1;                    // 0: even key tiles, 1: odd key tiles
    const int quad = warp & 3;                          // TMEM lane quadrant this warp may read
    const int r = quad * 32 + lane;                     // query row within the tile
    const int qpos = p.pos0 + q0 + r;
    const uint32_t lane_addr = (uint32_t)(quad * 32) << 16;
    uint8_t* sP = smem + TC_OFF_P + grp * TC_TILE;
    const uint32_t prow = (uint32_t)(r >> 3) * 1024 + (uint32_t)(r & 7) * 128;
    const uint32_t ts = tmem_S + lane_addr + grp * TC_BN;
    const uint32_t to = tmem_O + lane_addr + grp * TC_HD;
    float o[TC_HD];
#pragma unroll
    for (int i = 0; i < TC_HD; ++i) o[i] = 0.f;
    float m_run = -INFINITY, l_run = 0.f;
    int it = 0;
    for (int j = grp; j < nt; j += 2, ++it) {
      const uint32_t ph = it & 1;
      const int key0 = (jt0 + j) * TC_BN;
      // keys beyond a row's own position are masked; only tiles that reach past the first row's position have any
      const int lim = (key0 + TC_BN - 1 > p.pos0 + q0) ? qpos - key0 : TC_BN;     // this row keeps keys [lo, lim] of the tile
      const int lo = WIN ? qpos - p.window + 1 - key0 : 0;                         // <= 0: no lower bound inside this tile
      mbar_wait(&s_full[grp], ph);
      tc_fence_after();
      // ---- pass 1: row maximum ----
      float mx = -INFINITY;
#pragma unroll 1
      for (int c = 0; c < TC_BN / 16; ++c) {
        uint32_t v[16];
        tmem_ld_32x16(ts + c * 16, v);
        tmem_ld_wait();
        if (c * 16 + 15 <= lim && (!WIN || c * 16 >= lo)) {
#pragma unroll
          for (int i = 0; i < 16; ++i) mx = fmaxf(mx, __uint_as_float(v[i]));
        } else {
#pragma unroll
          for (int i = 0; i < 16; ++i)
            mx = fmaxf(mx, (c * 16 + i <= lim && (!WIN || c * 16 + i >= lo)) ? __uint_as_float(v[i]) : -INFINITY);
        }
      }
      // (a group's first tile may be fully masked for a row: odd tile 1 when pos0 + q0 < 128)
      const float m_new = fmaxf(m_run, mx * p.scale_log2);
      const float m_use = (m_new == -INFINITY) ? 0.f : m_new;
      const float corr = ex2_approx(m_run - m_use);               // m_run = -inf -> 0
      // ---- pass 2: P = exp2(s*scale - m), row sum, fp16 P into the swizzled A-operand layout ----
      float ls = 0.f, ls1 = 0.f;
#pragma unroll 1
      for (int c = 0; c < TC_BN / 16; ++c) {
        uint32_t v[16];
        tmem_ld_32x16(ts + c * 16, v);
        tmem_ld_wait();
        uint32_t pk[8];
        const bool full = c * 16 + 15 <= lim && (!WIN || c * 16 >= lo);
#pragma unroll
        for (int i = 0; i < 16; i += 2) {
          float x0, x1;
          ffma2(x0, x1, __uint_as_float(v[i]), __uint_as_float(v[i + 1]), p.scale_log2, -m_use, -m_use);
          float e0 = ex2_approx(x0);
          float e1 = ex2_approx(x1);
          if (!full) {
            if (c * 16 + i > lim || (WIN && c * 16 + i < lo)) e0 = 0.f;
            if (c * 16 + i + 1 > lim || (WIN && c * 16 + i + 1 < lo)) e1 = 0.f;
          }
          // (row sum in fp32 of the unrounded values, as in the mma.sync kernel: converting the fp16 P back
          //  costs one XU-pipe op per element, and XU -- ex2 + conversions -- is this kernel's busiest pipe)
          fadd2(ls, ls1, e0, e1);
          pk[i >> 1] = f2_to_h2(e0, e1);
        }
        // keys c*16 .. c*16+15 = atom column c/4, 16-byte chunks (c%4)*2, +1 of this row
        uint8_t* base = sP + (c >> 2) * TC_ATOM + prow;
#pragma unroll
        for (int q = 0; q < 2; ++q) {
          const uint32_t chunk = (uint32_t)((c & 3) * 2 + q) ^ (uint32_t)(r & 7);
          *reinterpret_cast<uint4*>(base + chunk * 16) = make_uint4(pk[q * 4], pk[q * 4 + 1], pk[q * 4 + 2], pk[q * 4 + 3]);
        }
      }
      // S[grp] may be overwritten by the QK^T of tile j+2; P[grp] is complete
      tc_fence_before();
      fence_proxy_async();
      __syncwarp();
      if (lane == 0) { mbar_arrive(&s_empty[grp]); mbar_arrive(&p_full[grp]); }
      l_run = l_run * corr + (ls + ls1);
      m_run = m_new;
      // ---- O = O*corr + P V  (the other group's softmax runs while this PV product is in flight) ----
      mbar_wait(&pv_full[grp], ph);
      tc_fence_after();
#pragma unroll
      for (int c = 0; c < TC_HD / 16; ++c) {
        uint32_t v[16];
        tmem_ld_32x16(to + c * 16, v);
        tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 16; i += 2)
          ffma2(o[c * 16 + i], o[c * 16 + i + 1], o[c * 16 + i], o[c * 16 + i + 1], corr, __uint_as_float(v[i]),
                __uint_as_float(v[i + 1]));
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&pv_empty[grp]);
    }
    // ---- merge the two groups' states (B -> shared memory -> A), normalise, store ----
    // Both groups have seen their last S tile, so every QK^T product has completed and the K ring is free.
    asm volatile("bar.sync 1, 256;" ::: "memory");
    float* sO = reinterpret_cast<float*>(smem + TC_OFF_K);       // [128 columns][128 rows]: conflict-free for thread = row
    float* sML = reinterpret_cast<float*>(smem + TC_OFF_V);      // [2][128]
    if (grp == 1) {
#pragma unroll
      for (int i = 0; i < TC_HD; ++i) sO[i * TC_BM + r] = o[i];
      sML[r] = m_run;
      sML[TC_BM + r] = l_run;
    }
    asm volatile("bar.sync 1, 256;" ::: "memory");
    if (grp == 0 && q0 + r < p.T) {
      const float mb = sML[r], lb = sML[TC_BM + r];
      const float m = fmaxf(m_run, mb);                          // finite: tile 0 holds key 0 for every row
      const float wa = ex2_approx(m_run - m), wb = ex2_approx(mb - m);
      const float l = l_run * wa + lb * wb;
      const float inv = l > 0.f ? 1.f / l : 0.f;
      const float ca = wa * inv, cb = wb * inv;
      __half* op = p.out + (long long)(q0 + r) * p.o_ld + h * TC_HD;
#pragma unroll
      for (int i = 0; i < TC_HD; i += 8) {
        float y[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) y[e] = fmaf(o[i + e], ca, sO[(i + e) * TC_BM + r] * cb);
        uint4 u;
        u.x = f2_to_h2(y[0], y[1]); u.y = f2_to_h2(y[2], y[3]); u.z = f2_to_h2(y[4], y[5]); u.w = f2_to_h2(y[6], y[7]);
        st_v4(op + i, u);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc<512>(tmem_base);
}

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------
typedef CUresult (*TcEncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static TcEncodeTiledFn tc_encode_fn() {
  static TcEncodeTiledFn fn = []() -> TcEncodeTiledFn {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qr;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qr) != cudaSuccess) return nullptr;
    if (qr != cudaDriverEntryPointSuccess) return nullptr;
    return reinterpret_cast<TcEncodeTiledFn>(p);
  }();
  return fn;
}
// fp16 matrix [rows][cols] with row pitch `pitch_elems`; box = 64 columns (128 B) x 128 rows, 128B swizzle
static int tc_make_tmap(CUtensorMap* tm, const void* base, long long rows, long long cols, long long pitch_elems) {
  TcEncodeTiledFn enc = tc_encode_fn();
  if (!enc) return ERR_DRIVER;
  cuuint64_t gdim[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t gstride[1] = {(cuuint64_t)pitch_elems * 2};
  cuuint32_t box[2] = {64, 128};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<void*>(base), gdim, gstride, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? OK : ERR_DRIVER;
}

// Shapes this kernel takes; everything else stays on the mma.sync kernel.
bool prefill_attn_tc_supported(int T, int hd, int window, long long q_ld, const void* q, const void* k, const void* v) {
  auto al = [](const void* x) { return (reinterpret_cast<uintptr_t>(x) & 15) == 0; };
  return hd == TC_HD && window >= 0 && T >= 64 && (q_ld % 8) == 0 && al(q) && al(k) && al(v);
}

int prefill_attn_tc(const void* q, long long q_ld, const void* kcache, const void* vcache, void* out, long long o_ld,
                    const int* pos_ptr, int T, int pos0, int row, int H, int KV, int Smax, float scale, int window, cudaStream_t st) {
  CUtensorMap tmQ, tmK, tmV;
  int rc = tc_make_tmap(&tmQ, q, T, (long long)H * TC_HD, q_ld);
  if (rc) return rc;
  // extent = up to the end of this cache row: a key tile that overshoots it is zero-filled, never out of the allocation
  const long long kv_rows = (long long)(row + 1) * KV * Smax;
  if ((rc = tc_make_tmap(&tmK, kcache, kv_rows, TC_HD, TC_HD))) return rc;
  if ((rc = tc_make_tmap(&tmV, vcache, kv_rows, TC_HD, TC_HD))) return rc;
  static DeviceOnce attr_once;
  if (!attr_once.get()) {
    if (cudaFuncSetAttribute(prefill_attn_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, TC_SMEM) != cudaSuccess ||
        cudaFuncSetAttribute(prefill_attn_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, TC_SMEM) != cudaSuccess)
      return ERR_ARCH;
    attr_once.set();
  }
  TcAttnParams p;
  p.out = (__half*)out; p.o_ld = o_ld; p.pos_ptr = pos_ptr; p.T = T; p.pos0 = pos0; p.row = row; p.H = H; p.KV = KV;
  p.Smax = Smax; p.window = window; p.scale_log2 = scale * 1.4426950408889634f;
  const dim3 grid((T + TC_BM - 1) / TC_BM, H);
  if (window > 0) prefill_attn_tc_kernel<true><<<grid, TC_THREADS, TC_SMEM, st>>>(tmQ, tmK, tmV, p);
  else prefill_attn_tc_kernel<false><<<grid, TC_THREADS, TC_SMEM, st>>>(tmQ, tmK, tmV, p);
  return launch_status();
}

int init_attn_tc() {
  cudaFuncAttributes fa;
  if (cudaFuncGetAttributes(&fa, prefill_attn_tc_kernel<false>) != cudaSuccess ||
      cudaFuncGetAttributes(&fa, prefill_attn_tc_kernel<true>) != cudaSuccess) return ERR_ARCH;
  if (cudaFuncSetAttribute(prefill_attn_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, TC_SMEM) != cudaSuccess ||
      cudaFuncSetAttribute(prefill_attn_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, TC_SMEM) != cudaSuccess)
    return ERR_ARCH;
  return tc_encode_fn() ? OK : ERR_DRIVER;
}

}  // namespace b2
