// Prefill W8A8 linear on CTA pairs:  tcgen05.mma.cta_group::2 kind::i8, 256 x 256 x 32 per instruction.
//
// Replaces the int8 linear that the reference reaches through bitsandbytes `Linear8bitLt`
// (Code/Quantised Models/models_quant_updated.py:30-38) / `quantize_dynamic` (Code/C-DAC Server/try.py:205)
// for the prefill forward of `model.generate` (models_quant_updated.py:84-93) and the all-position pass of
// `confidence_score` (:63-70), together with what HF runs right after three of the four linears of a decoder
// layer: rotary embedding + KV-cache append after q/k/v_proj (modeling_llama.py:128-153, :251-289) and
// act(gate)*up after gate/up_proj (:182-184).  Those become GEMM epilogues here.
//
// Why a CTA pair: the single-CTA kernel (gemm_w8a8.cu) keeps the tensor pipe ~49 % busy at M = 2048 -- each SM
// must stage 128 x 128 B of A and 256 x 128 B of W per 512 MMA cycles (96 B/clk in, 96 B/clk out of shared
// memory).  With cta_group::2 the two SMs of a TPC compute one 256 x 256 tile together: each stages its own 128
// rows of A and HALF of the weight tile (64 + 64 B/clk), the tensor cores read the other half from the peer's
// shared memory, and the ring is 6 stages of 32 KB instead of 4 of 48 KB.
//
// One cluster = 2 CTAs (ranks 0 = leader, 1 = peer), EW + 4 warps each (EW = 8 epilogue warps, 16 for the act*up
// epilogue; the roles below are given for EW = 8):
//   warps 0..7  epilogue: TMEM lane quadrant = warp % 4, column half (quarter with 16 warps) = warp / 4; arrive on
//            the leader's `tempty` when a buffer is drained (remote mbarrier arrive from the peer)
//   warp 8   TMA producer (each CTA loads its own A rows and its own half of the weight tile; both signal the
//            LEADER's `full` barrier -- cp.async.bulk.tensor ... .cta_group::2)
//   warp 9   MMA issuer (leader only): waits `full`, issues 4 x tcgen05.mma per 128-byte k-block, commits to the
//            `empty` barrier of BOTH CTAs (tcgen05.commit ... multicast::cluster) and, per tile, to `tfull` of both
//   warp 10  TMEM allocator (512 columns = two 256-column accumulator buffers, allocated by the pair)
// The accumulator of CTA r holds rows [128 r, 128 r + 128) of the pair's tile and all 256 columns; columns
// [0,128) come from the leader's weight rows, [128,256) from the peer's.
//
// Epilogues (all bit-identical to the unfused entry points they replace; oracle/w8a8.py):
//   PLAIN  out = fp16(acc * a_scale[m] * w_scale[n] + bias[n] + residual[m,n])                 (== b2_gemm_w8a8)
//   GATED  W = [gate; up] (2F rows): the leader stages 128 gate rows, the peer the matching 128 up rows, so one
//          thread holds gate[f] and up[f]:  y = act(fp16(gate)) * fp16(up) * out_smooth[f] as fp32 [M,F] plus the
//          row absmax by atomicMax on the float bits (exact: max is order-independent)         (== gemm + the
//          activation half of b2_act_mul_quant; b2_quant_f32 finishes the quantisation)
//   QKV    W = [q; k; v]: fp16 round, rotary embedding in the pinned fp32 order, rotated q to the qkv buffer,
//          rotated k and v straight into the KV cache                              (== gemm + b2_rope_kv_append)
#include <cuda.h>
#include "common.cuh"

namespace b2 {

constexpr int G2_BM = 256;       // rows per CTA pair (128 per CTA)
constexpr int G2_BN = 256;       // columns per tile
constexpr int G2_BK = 128;       // int8 per k-block = one 128-byte swizzle atom
// epilogue warps per CTA: 8 (lane quadrant x column half); 16 for the act*up epilogue (lane quadrant x column
// quarter), whose exp / division chains are latency-bound on two warps per scheduler: with 8 warps the epilogue of a
// tile (~16 us) outlasted its MMAs (13.9 us) and the tensor pipe fell from 72 % to 61 % active (ncu, round 2)
template <int EPI>
constexpr int g2_epi_warps() { return EPI == 1 ? 16 : 8; }
template <int EPI>
constexpr int g2_threads() { return (g2_epi_warps<EPI>() + 4) * 32; }
// Warp roles.  The epilogue warps take the LOW warp ids: when several warps of a scheduler are ready the SM issues
// the highest warp id first, and an ALU-heavy epilogue (SiLU, rotary) is ready nearly every cycle -- with the
// producer and the MMA issuer below it they waited for issue slots and the tensor pipe ran dry (measured: the gated
// epilogue cost 14 % of the GEMM with the control warps at ids 0..2).
// warps 0..EW-1 = epilogue, EW = TMA producer, EW+1 = MMA issuer, EW+2 = TMEM allocator, EW+3 idle
constexpr int G2_TMEM_COLS = 512;
constexpr int G2_STAGE_BYTES = 128 * G2_BK * 2;   // per CTA: A 128 x 128 B + W 128 x 128 B

enum : int { EPI_PLAIN = 0, EPI_GATED = 1, EPI_QKV = 2 };

template <int STAGES>
struct G2Smem {
  static constexpr int BAR_OFF = STAGES * G2_STAGE_BYTES;
  static constexpr int TOTAL = BAR_OFF + 256 + 1024;
};

struct G2Params {
  const float* a_scale;      // [M]
  const float* w_scale;      // [N]
  const __half* bias;        // [N] or null
  const __half* residual;    // PLAIN: [M,N] or null
  __half* out;               // PLAIN: [M,N]; QKV: the qkv buffer (q columns written)
  int32_t* acc_dump;         // PLAIN test hook
  long long ldo;             // row pitch (elements) of out / residual / acc_dump
  int M, N, K;
  // GATED
  const float* out_smooth;   // [F] or null
  float* y32;                // [M,F]
  float* amax;               // [M], zeroed before the launch
  int F, act;                // act 0 = silu, 1 = gelu_new
  // QKV
  const float* cosT; const float* sinT; const int* positions; const int* seq_ids; int seq0;
  __half* kcache; __half* vcache; int H, KV, Smax, use_rope;
};

// ---------------------------------------------------------------------------------------------
// PTX for the pair (cta_group::2) forms
// ---------------------------------------------------------------------------------------------
B2_DEVINL uint32_t mapa_shared(uint32_t addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
  return r;
}
// TMA load into THIS CTA's shared memory whose completion bytes are counted on an mbarrier that may live in
// the other CTA of the pair (bar = a shared::cluster address)
B2_DEVINL void tma_load_2d_pair(void* smem_dst, const void* tmap, uint32_t bar_cluster_addr, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(smem_u32(smem_dst)), "l"(tmap), "r"(bar_cluster_addr), "r"(c0), "r"(c1)
      : "memory");
}
B2_DEVINL void umma_i8_pair(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::i8 [%0], %1, %2, %3, p;\n\t}\n" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// all MMAs issued so far by this thread: arrive on the same-offset mbarrier of every CTA in `mask`
B2_DEVINL void umma_commit_pair(uint64_t* bar, uint16_t mask) {
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
               ::"r"(smem_u32(bar)), "h"(mask)
               : "memory");
}
B2_DEVINL void mbar_arrive_cluster(uint32_t bar_cluster_addr) {
  asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(bar_cluster_addr) : "memory");
}
B2_DEVINL void tmem_alloc_pair(uint32_t* smem_result) {   // one full warp in EACH CTA of the pair
  asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_result)),
               "n"(G2_TMEM_COLS)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
}
B2_DEVINL void tmem_dealloc_pair(uint32_t taddr) {
  asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "n"(G2_TMEM_COLS) : "memory");
}

// mbar_wait that turns a protocol deadlock into a trapped launch instead of a hung GPU: try_wait suspends for up
// to a hardware time limit per attempt, so 2^26 failed attempts are tens of seconds, far beyond any legal wait.
B2_DEVINL void g2_wait(uint64_t* bar, uint32_t parity) {
  uint32_t done = 0;
  for (uint32_t spins = 0; !done; ++spins) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}\n"
        : "=r"(done) : "r"(smem_u32(bar)), "r"(parity) : "memory");
    if (!done && spins > (1u << 26)) __trap();
  }
}

// ---------------------------------------------------------------------------------------------
// epilogue pieces (thread = one accumulator row; r[] = 32 consecutive accumulator columns)
// ---------------------------------------------------------------------------------------------
// v[j] = fp32 dequantised (+ bias) value of column col0 + j, j in [0,32), not yet rounded
B2_DEVINL void dequant32(const uint32_t (&r)[32], float as, const float* __restrict__ w_scale,
                         const __half* __restrict__ bias, int col0, float (&v)[32]) {
#pragma unroll
  for (int j = 0; j < 32; j += 4) {
    const float4 w = *reinterpret_cast<const float4*>(w_scale + col0 + j);
    v[j + 0] = mul_rn(mul_rn((float)(int)r[j + 0], as), w.x);
    v[j + 1] = mul_rn(mul_rn((float)(int)r[j + 1], as), w.y);
    v[j + 2] = mul_rn(mul_rn((float)(int)r[j + 2], as), w.z);
    v[j + 3] = mul_rn(mul_rn((float)(int)r[j + 3], as), w.w);
  }
  if (bias) {
#pragma unroll
    for (int j = 0; j < 32; j += 8) {
      const uint4 b = ld_v4(bias + col0 + j);
      const float2 b0 = h2_to_f2(b.x), b1 = h2_to_f2(b.y), b2_ = h2_to_f2(b.z), b3 = h2_to_f2(b.w);
      v[j + 0] = add_rn(v[j + 0], b0.x); v[j + 1] = add_rn(v[j + 1], b0.y);
      v[j + 2] = add_rn(v[j + 2], b1.x); v[j + 3] = add_rn(v[j + 3], b1.y);
      v[j + 4] = add_rn(v[j + 4], b2_.x); v[j + 5] = add_rn(v[j + 5], b2_.y);
      v[j + 6] = add_rn(v[j + 6], b3.x); v[j + 7] = add_rn(v[j + 7], b3.y);
    }
  }
}
B2_DEVINL float round_h(float x) { return __half2float(__float2half_rn(x)); }

template <int EPI, int HD, int STAGES>
__global__ void __launch_bounds__(g2_threads<EPI>(), 1)
gemm2_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const G2Params p) {
  using S = G2Smem<STAGES>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + S::BAR_OFF);   // leader's copy is the live one
  uint64_t* empty_bar = full_bar + STAGES;                               // one per CTA
  uint64_t* tfull_bar = empty_bar + STAGES;                              // [2] one per CTA
  uint64_t* tempty_bar = tfull_bar + 2;                                  // [2] leader's copy is the live one
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 2);

  constexpr int EW = g2_epi_warps<EPI>();
  constexpr int G2_WARP_TMA = EW, G2_WARP_MMA = EW + 1, G2_WARP_ALLOC = EW + 2;
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const int m_tiles = (p.M + G2_BM - 1) / G2_BM;
  const int n_tiles = (EPI == EPI_GATED) ? (p.F + 127) / 128 : (p.N + G2_BN - 1) / G2_BN;
  const int num_tiles = m_tiles * n_tiles;
  const int t_first = blockIdx.x >> 1, t_step = gridDim.x >> 1;
  const int kblocks = (p.K + G2_BK - 1) / G2_BK;

  if (warp == G2_WARP_TMA && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
  }
  if (warp == G2_WARP_MMA && lane == 0) {
    for (int i = 0; i < STAGES; ++i) { mbar_init(&full_bar[i], 1); mbar_init(&empty_bar[i], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&tfull_bar[i], 1); mbar_init(&tempty_bar[i], 2 * EW); }   // every epilogue warp of both CTAs
    mbar_fence_init();
  }
  if (warp == G2_WARP_ALLOC) tmem_alloc_pair(tmem_slot);
  tc_fence_before();
  __syncthreads();
  cluster_sync();                  // both CTAs' barriers exist before any remote arrive / multicast commit
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == G2_WARP_TMA) {
    // ===================== TMA producer (both CTAs) =====================
    if (lane == 0) {
      int stage = 0; uint32_t phase = 0;
      for (int t = t_first; t < num_tiles; t += t_step) {
        const int mt = t % m_tiles, nt = t / m_tiles;
        const int arow = mt * G2_BM + (int)rank * 128;
        int brow;
        if (EPI == EPI_GATED) brow = (rank == 0 ? 0 : p.F) + nt * 128;      // leader: gate rows, peer: up rows
        else brow = nt * G2_BN + (int)rank * 128;
        for (int kb = 0; kb < kblocks; ++kb) {
          g2_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = smem + stage * G2_STAGE_BYTES;
          uint8_t* sb = sa + 128 * G2_BK;
          if (rank == 0) mbar_expect_tx(&full_bar[stage], 2 * G2_STAGE_BYTES);   // both CTAs' bytes land on this barrier
          const uint32_t bar = mapa_shared(smem_u32(&full_bar[stage]), 0);
          tma_load_2d_pair(sa, &tmA, bar, kb * G2_BK, arow);
          tma_load_2d_pair(sb, &tmB, bar, kb * G2_BK, brow);
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == G2_WARP_MMA) {
    // ===================== MMA issuer (leader only) =====================
    if (rank == 0 && lane == 0) {
      constexpr uint32_t idesc = umma_idesc(IDESC_FMT_S32, IDESC_S8, IDESC_S8, G2_BM, G2_BN);
      int stage = 0; uint32_t phase = 0;
      int it = 0;
      for (int t = t_first; t < num_tiles; t += t_step, ++it) {
        const int buf = it & 1;
        const uint32_t bphase = (it >> 1) & 1;
        g2_wait(&tempty_bar[buf], bphase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + buf * G2_BN;
        for (int kb = 0; kb < kblocks; ++kb) {
          g2_wait(&full_bar[stage], phase);
          tc_fence_after();
          const uint32_t sa = smem_u32(smem + stage * G2_STAGE_BYTES);
          const uint64_t adesc = umma_desc_sw128(sa);
          const uint64_t bdesc = umma_desc_sw128(sa + 128 * G2_BK);
#pragma unroll
          for (int k = 0; k < G2_BK / 32; ++k)
            umma_i8_pair(d_tmem, adesc + 2 * k, bdesc + 2 * k, idesc, (kb | k) != 0);
          umma_commit_pair(&empty_bar[stage], 3);
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
        umma_commit_pair(&tfull_bar[buf], 3);
      }
    }
  } else if (warp < EW) {
    // ===================== epilogue (both CTAs) =====================
    const int quad = warp & 3;
    const int half = warp >> 2;          // column half (PLAIN, QKV) or column quarter (GATED: 16 warps)
    int it = 0;
    for (int t = t_first; t < num_tiles; t += t_step, ++it) {
      const int mt = t % m_tiles, nt = t / m_tiles;
      const int buf = it & 1;
      const uint32_t bphase = (it >> 1) & 1;
      g2_wait(&tfull_bar[buf], bphase);
      tc_fence_after();
      const int row = mt * G2_BM + (int)rank * 128 + quad * 32 + lane;
      const bool row_ok = row < p.M;
      const float as = row_ok ? p.a_scale[row] : 0.0f;
      const uint32_t taddr = tmem_base + ((uint32_t)(quad * 32) << 16) + buf * G2_BN;

      if (EPI == EPI_PLAIN) {
#pragma unroll 1
        for (int c = half * 4; c < half * 4 + 4; ++c) {
          uint32_t r[32];
          __syncwarp();
          tmem_ld_32x32(taddr + c * 32, r);
          tmem_ld_wait();
          const int col0 = nt * G2_BN + c * 32;
          if (row_ok && col0 < p.N) {                         // N % 32 == 0 (checked on the host)
            const long long off = (long long)row * p.ldo + col0;
            if (p.acc_dump) {
#pragma unroll
              for (int j = 0; j < 32; j += 4)
                *reinterpret_cast<uint4*>(p.acc_dump + off + j) = make_uint4(r[j], r[j + 1], r[j + 2], r[j + 3]);
            }
            float v[32];
            dequant32(r, as, p.w_scale, p.bias, col0, v);
#pragma unroll
            for (int j = 0; j < 32; j += 8) {
              if (p.residual) {
                const uint4 b = ld_v4(p.residual + off + j);
                const float2 b0 = h2_to_f2(b.x), b1 = h2_to_f2(b.y), b2_ = h2_to_f2(b.z), b3 = h2_to_f2(b.w);
                v[j + 0] = add_rn(v[j + 0], b0.x); v[j + 1] = add_rn(v[j + 1], b0.y);
                v[j + 2] = add_rn(v[j + 2], b1.x); v[j + 3] = add_rn(v[j + 3], b1.y);
                v[j + 4] = add_rn(v[j + 4], b2_.x); v[j + 5] = add_rn(v[j + 5], b2_.y);
                v[j + 6] = add_rn(v[j + 6], b3.x); v[j + 7] = add_rn(v[j + 7], b3.y);
              }
              uint4 o;
              o.x = f2_to_h2(v[j + 0], v[j + 1]); o.y = f2_to_h2(v[j + 2], v[j + 3]);
              o.z = f2_to_h2(v[j + 4], v[j + 5]); o.w = f2_to_h2(v[j + 6], v[j + 7]);
              st_v4(p.out + off + j, o);
            }
          }
        }
      } else if (EPI == EPI_GATED) {
        // warp (quad, half): 32 rows x the 32 MLP columns [half*32, half*32 + 32) of the tile, 16 at a time (the register
        // budget of a 640-thread CTA is 96 per thread)
        float amax = 0.0f;
#pragma unroll 1
        for (int c16 = 0; c16 < 2; ++c16) {
          const int cc = half * 32 + c16 * 16;            // column inside the 128-wide gate (and up) block
          uint32_t rg[16], ru[16];
          __syncwarp();
          tmem_ld_32x16(taddr + cc, rg);                  // gate: the leader's weight rows
          tmem_ld_32x16(taddr + 128 + cc, ru);            // up: the peer's
          tmem_ld_wait();
          const int f0 = nt * 128 + cc;
          if (row_ok && f0 < p.F) {                           // F % 32 == 0 (checked on the host)
            float* yp = p.y32 + (long long)row * p.F + f0;
#pragma unroll
            for (int j = 0; j < 16; j += 4) {
              const float4 wg = *reinterpret_cast<const float4*>(p.w_scale + f0 + j);
              const float4 wu = *reinterpret_cast<const float4*>(p.w_scale + p.F + f0 + j);
              const float wgs[4] = {wg.x, wg.y, wg.z, wg.w}, wus[4] = {wu.x, wu.y, wu.z, wu.w};
              float bg[4] = {0.f, 0.f, 0.f, 0.f}, bu[4] = {0.f, 0.f, 0.f, 0.f};
              if (p.bias) {
                const uint2 b0 = *reinterpret_cast<const uint2*>(p.bias + f0 + j);
                const uint2 b1 = *reinterpret_cast<const uint2*>(p.bias + p.F + f0 + j);
                const float2 g0 = h2_to_f2(b0.x), g1 = h2_to_f2(b0.y), u0 = h2_to_f2(b1.x), u1 = h2_to_f2(b1.y);
                bg[0] = g0.x; bg[1] = g0.y; bg[2] = g1.x; bg[3] = g1.y;
                bu[0] = u0.x; bu[1] = u0.y; bu[2] = u1.x; bu[3] = u1.y;
              }
              float y[4];
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                float vg = mul_rn(mul_rn((float)(int)rg[j + e], as), wgs[e]);
                float vu = mul_rn(mul_rn((float)(int)ru[j + e], as), wus[e]);
                if (p.bias) { vg = add_rn(vg, bg[e]); vu = add_rn(vu, bu[e]); }
                const float gh = round_h(vg), uh = round_h(vu);
                const float a = (p.act == 0) ? silu_det(gh) : gelu_new_det(gh);
                y[e] = mul_rn(a, uh);
              }
              if (p.out_smooth) {
                const float4 sm = *reinterpret_cast<const float4*>(p.out_smooth + f0 + j);
                y[0] = mul_rn(y[0], sm.x); y[1] = mul_rn(y[1], sm.y); y[2] = mul_rn(y[2], sm.z); y[3] = mul_rn(y[3], sm.w);
              }
              amax = fmaxf(amax, fmaxf(fmaxf(fabsf(y[0]), fabsf(y[1])), fmaxf(fabsf(y[2]), fabsf(y[3]))));
              *reinterpret_cast<float4*>(yp + j) = make_float4(y[0], y[1], y[2], y[3]);
            }
          }
        }
        // non-negative floats order like their bit patterns
        if (row_ok && amax > 0.0f) atomicMax(reinterpret_cast<int*>(p.amax + row), __float_as_int(amax));
      } else {   // EPI_QKV
        const int pos = row_ok ? p.positions[row] : 0;
        const int crow = row_ok ? (p.seq_ids ? p.seq_ids[row] : p.seq0) : 0;
        constexpr int HH = HD / 2;
#pragma unroll 1
        for (int h = 0; h < 128 / HD; ++h) {
          const int hcol = half * 128 + h * HD;              // first accumulator column of this head
          const int n0 = nt * G2_BN + hcol;                  // its global output column
          const int head = n0 / HD;                          // q heads, then k heads, then v heads
#pragma unroll 1
          for (int c = 0; c < HD / 64; ++c) {
            uint32_t ra[32], rb[32];
            tmem_ld_32x32(taddr + hcol + c * 32, ra);        // first half of the head
            tmem_ld_32x32(taddr + hcol + HH + c * 32, rb);   // second half (its rotation partners)
            tmem_ld_wait();
            if (row_ok && n0 < p.N) {
            float a[32], b[32];
            dequant32(ra, as, p.w_scale, p.bias, n0 + c * 32, a);
            dequant32(rb, as, p.w_scale, p.bias, n0 + HH + c * 32, b);
            uint32_t o1[16], o2[16];
            if (head < p.H + p.KV && p.use_rope) {
              const float* cp = p.cosT + (long long)pos * HH + c * 32;
              const float* sp = p.sinT + (long long)pos * HH + c * 32;
#pragma unroll
              for (int j = 0; j < 32; j += 4) {
                const float4 cs = *reinterpret_cast<const float4*>(cp + j);
                const float4 sn = *reinterpret_cast<const float4*>(sp + j);
                const float cc[4] = {cs.x, cs.y, cs.z, cs.w}, ss[4] = {sn.x, sn.y, sn.z, sn.w};
                float x[4], y[4];
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                  const float ah = round_h(a[j + e]), bh = round_h(b[j + e]);
                  x[e] = sub_rn(mul_rn(ah, cc[e]), mul_rn(bh, ss[e]));
                  y[e] = add_rn(mul_rn(bh, cc[e]), mul_rn(ah, ss[e]));
                }
                o1[j / 2] = f2_to_h2(x[0], x[1]); o1[j / 2 + 1] = f2_to_h2(x[2], x[3]);
                o2[j / 2] = f2_to_h2(y[0], y[1]); o2[j / 2 + 1] = f2_to_h2(y[2], y[3]);
              }
            } else {
#pragma unroll
              for (int j = 0; j < 32; j += 2) {
                o1[j / 2] = f2_to_h2(a[j], a[j + 1]);
                o2[j / 2] = f2_to_h2(b[j], b[j + 1]);
              }
            }
            __half* dst;
            if (head < p.H) {
              dst = p.out + (long long)row * p.ldo + head * HD;
            } else {
              const bool isk = head < p.H + p.KV;
              const int kvh = isk ? head - p.H : head - p.H - p.KV;
              dst = (isk ? p.kcache : p.vcache) + (((long long)crow * p.KV + kvh) * p.Smax + pos) * HD;
            }
            dst += c * 32;
#pragma unroll
            for (int j = 0; j < 16; j += 4) {
              st_v4(dst + j * 2, make_uint4(o1[j], o1[j + 1], o1[j + 2], o1[j + 3]));
              st_v4(dst + HH + j * 2, make_uint4(o2[j], o2[j + 1], o2[j + 2], o2[j + 3]));
            }
            }
            __syncwarp();
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(mapa_shared(smem_u32(&tempty_bar[buf]), 0));
    }
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync();                  // the leader's MMAs read the peer's shared memory and write its TMEM
  if (warp == G2_WARP_ALLOC) tmem_dealloc_pair(tmem_base);
}

// ---------------------------------------------------------------------------------------------
// fp32 y [T,F] + row absmax -> int8 + row scale: the quantise half of b2_act_mul_quant after a GATED GEMM
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) quant_f32_kernel(const float* __restrict__ y, const float* __restrict__ amax_in,
                                                        int8_t* __restrict__ q, float* __restrict__ row_scale, int F) {
  const int row = blockIdx.x;
  const float amax = amax_in[row];
  const bool nz = amax > 0.0f;
  const float inv = nz ? div_rn(127.0f, amax) : 0.0f;
  if (threadIdx.x == 0) row_scale[row] = nz ? div_rn(amax, 127.0f) : 0.0f;
  const float* yr = y + (long long)row * F;
  int8_t* qr = q + (long long)row * F;
  const int nchunk = F >> 4;     // 16 values = 64 B in, 16 B out
  for (int ch = threadIdx.x; ch < nchunk; ch += blockDim.x) {
    const float4 a = *reinterpret_cast<const float4*>(yr + ch * 16);
    const float4 b = *reinterpret_cast<const float4*>(yr + ch * 16 + 4);
    const float4 c = *reinterpret_cast<const float4*>(yr + ch * 16 + 8);
    const float4 d = *reinterpret_cast<const float4*>(yr + ch * 16 + 12);
    uint4 o;
    o.x = pack4(quant1(a.x, inv), quant1(a.y, inv), quant1(a.z, inv), quant1(a.w, inv));
    o.y = pack4(quant1(b.x, inv), quant1(b.y, inv), quant1(b.z, inv), quant1(b.w, inv));
    o.z = pack4(quant1(c.x, inv), quant1(c.y, inv), quant1(c.z, inv), quant1(c.w, inv));
    o.w = pack4(quant1(d.x, inv), quant1(d.y, inv), quant1(d.z, inv), quant1(d.w, inv));
    *reinterpret_cast<uint4*>(qr + ch * 16) = o;
  }
}

int quant_f32(const float* y, const float* amax, int8_t* q, float* row_scale, int T, int F, cudaStream_t st) {
  if (T == 0) return OK;
  if (!y || !amax || !q || !row_scale) return ERR_NULL_PTR;
  if (T < 0 || F <= 0 || (F & 15)) return ERR_BAD_SHAPE;
  auto mis = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) != 0; };
  if (mis(y) || mis(q)) return ERR_BAD_ALIGN;
  quant_f32_kernel<<<T, 256, 0, st>>>(y, amax, q, row_scale, F);
  return launch_status();
}

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn2)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                   const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                   CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn2 g2_encode_fn() {
  static EncodeTiledFn2 fn = []() -> EncodeTiledFn2 {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qr;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qr) != cudaSuccess) return nullptr;
    if (qr != cudaDriverEntryPointSuccess) return nullptr;
    return reinterpret_cast<EncodeTiledFn2>(p);
  }();
  return fn;
}
// rows x K int8, K contiguous; box = 128 rows x 128 bytes, 128B swizzle, out-of-bounds reads as zero
static int g2_tmap(CUtensorMap* tm, const void* base, long long rows, int K) {
  EncodeTiledFn2 enc = g2_encode_fn();
  if (!enc) return ERR_DRIVER;
  cuuint64_t gdim[2] = {(cuuint64_t)K, (cuuint64_t)rows};
  cuuint64_t gstride[1] = {(cuuint64_t)K};
  cuuint32_t box[2] = {(cuuint32_t)G2_BK, 128};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<void*>(base), gdim, gstride, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? OK : ERR_DRIVER;
}

constexpr int G2_STAGES = 6;

template <int EPI, int HD>
static int g2_launch(const int8_t* A, const int8_t* W, long long w_rows, const G2Params& p, cudaStream_t st) {
  using S = G2Smem<G2_STAGES>;
  auto kern = gemm2_kernel<EPI, HD, G2_STAGES>;
  static DeviceOnce attr_once;
  if (!attr_once.get()) {
    if (cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, S::TOTAL) != cudaSuccess) return ERR_ARCH;
    attr_once.set();
  }
  CUtensorMap tmA, tmB;
  int rc = g2_tmap(&tmA, A, p.M, p.K);
  if (rc) return rc;
  if ((rc = g2_tmap(&tmB, W, w_rows, p.K))) return rc;
  const int m_tiles = (p.M + G2_BM - 1) / G2_BM;
  const int n_tiles = (EPI == EPI_GATED) ? (p.F + 127) / 128 : (p.N + G2_BN - 1) / G2_BN;
  const int tiles = m_tiles * n_tiles;
  const int max_pairs = device_sm_count() / 2;
  const int pairs = tiles < max_pairs ? tiles : max_pairs;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(pairs * 2);
  cfg.blockDim = dim3(g2_threads<EPI>());
  cfg.dynamicSmemBytes = S::TOTAL;
  cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  if (cudaLaunchKernelEx(&cfg, kern, tmA, tmB, p) != cudaSuccess) {
    cudaGetLastError();
    return ERR_LAUNCH;
  }
  return OK;
}

static inline bool g2_mis(const void* q) { return q && (reinterpret_cast<uintptr_t>(q) & 15); }

// Shapes for which the default schedule takes the pair kernel (the single-CTA kernel otherwise).  Measured on B200
// (tools/gemm_bench.py, Llama-2-7B shapes): with at least ~48 pair tiles the pair kernel wins (M = 256: qkv 21.6 vs
// 28.6 us, gate_up 33.6 vs 38.8; M = 2048: 318 vs 409 us for the four linears), with 16-32 tiles (N = 4096 at
// M = 256 / 512) the 128 x 128 single-CTA tiles fill more SMs (o 17.6 / 18.7 vs 21.6 / 22.1 us).
bool gemm2_supported(int M, int N, int K) {
  if (!(M > 128 && N >= 256 && (N & 31) == 0 && K >= 128 && (K & 15) == 0)) return false;
  const long long tiles = (long long)((M + G2_BM - 1) / G2_BM) * ((N + G2_BN - 1) / G2_BN);
  return tiles >= 40;
}

int gemm2_plain(const int8_t* A, const int8_t* W, const float* a_scale, const float* w_scale, const void* bias,
                const void* residual, void* out, int32_t* acc_dump, int M, int N, int K, cudaStream_t st) {
  if (M == 0 || N == 0) return OK;
  if (!A || !W || !a_scale || !w_scale || !out) return ERR_NULL_PTR;
  if (M < 0 || N < 0 || K <= 0 || (K & 15) || (N & 31)) return ERR_BAD_SHAPE;
  if (g2_mis(A) || g2_mis(W) || g2_mis(w_scale) || g2_mis(bias) || g2_mis(residual) || g2_mis(out) || g2_mis(acc_dump))
    return ERR_BAD_ALIGN;
  G2Params p = {};
  p.a_scale = a_scale; p.w_scale = w_scale; p.bias = (const __half*)bias; p.residual = (const __half*)residual;
  p.out = (__half*)out; p.acc_dump = acc_dump; p.ldo = N; p.M = M; p.N = N; p.K = K;
  return g2_launch<EPI_PLAIN, 128>(A, W, N, p, st);
}

// W [2F,K] = [gate; up]; y32 [M,F] fp32 = act(fp16(gate)) * fp16(up) * out_smooth; amax [M] = row absmax of y32
int gemm2_gated(const int8_t* A, const int8_t* W, const float* a_scale, const float* w_scale, const void* bias,
                const float* out_smooth, float* y32, float* amax, int M, int F, int K, int act, cudaStream_t st) {
  if (M == 0 || F == 0) return OK;
  if (!A || !W || !a_scale || !w_scale || !y32 || !amax) return ERR_NULL_PTR;
  if (M < 0 || F < 0 || K <= 0 || (K & 15) || (F & 31)) return ERR_BAD_SHAPE;
  if (act != 0 && act != 1) return ERR_UNSUPPORTED;
  if (g2_mis(A) || g2_mis(W) || g2_mis(w_scale) || g2_mis(bias) || g2_mis(out_smooth) || g2_mis(y32)) return ERR_BAD_ALIGN;
  if (cudaMemsetAsync(amax, 0, (size_t)M * sizeof(float), st) != cudaSuccess) { cudaGetLastError(); return ERR_LAUNCH; }
  G2Params p = {};
  p.a_scale = a_scale; p.w_scale = w_scale; p.bias = (const __half*)bias; p.M = M; p.N = 2 * F; p.K = K;
  p.out_smooth = out_smooth; p.y32 = y32; p.amax = amax; p.F = F; p.act = act;
  return g2_launch<EPI_GATED, 128>(A, W, 2LL * F, p, st);
}

// W [(H+2KV)*hd, K] = [q; k; v]
int gemm2_qkv(const int8_t* A, const int8_t* W, const float* a_scale, const float* w_scale, const void* bias,
              void* qkv_out, long long ld, const float* cosT, const float* sinT, const int* positions,
              const int* seq_ids, int seq0, void* kcache, void* vcache, int M, int H, int KV, int hd, int Smax, int K,
              int use_rope, cudaStream_t st) {
  if (M == 0) return OK;
  if (!A || !W || !a_scale || !w_scale || !qkv_out || !positions || !kcache || !vcache) return ERR_NULL_PTR;
  if (use_rope && (!cosT || !sinT)) return ERR_NULL_PTR;
  if (M < 0 || H <= 0 || KV <= 0 || K <= 0 || (K & 15) || (ld & 7) || Smax <= 0) return ERR_BAD_SHAPE;
  if (hd != 64 && hd != 128) return ERR_UNSUPPORTED;
  if (g2_mis(A) || g2_mis(W) || g2_mis(w_scale) || g2_mis(bias) || g2_mis(qkv_out) || g2_mis(kcache) || g2_mis(vcache) ||
      g2_mis(cosT) || g2_mis(sinT))
    return ERR_BAD_ALIGN;
  G2Params p = {};
  p.a_scale = a_scale; p.w_scale = w_scale; p.bias = (const __half*)bias; p.out = (__half*)qkv_out; p.ldo = ld;
  p.M = M; p.N = (H + 2 * KV) * hd; p.K = K;
  p.cosT = cosT; p.sinT = sinT; p.positions = positions; p.seq_ids = seq_ids; p.seq0 = seq0;
  p.kcache = (__half*)kcache; p.vcache = (__half*)vcache; p.H = H; p.KV = KV; p.Smax = Smax; p.use_rope = use_rope;
  if (hd == 128) return g2_launch<EPI_QKV, 128>(A, W, p.N, p, st);
  return g2_launch<EPI_QKV, 64>(A, W, p.N, p, st);
}

template <int EPI, int HD>
static int g2_init_one() {
  return cudaFuncSetAttribute(gemm2_kernel<EPI, HD, G2_STAGES>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                              G2Smem<G2_STAGES>::TOTAL) == cudaSuccess ? OK : ERR_ARCH;
}
int init_gemm2() {
  if (!g2_encode_fn()) return ERR_DRIVER;
  int rc;
  if ((rc = g2_init_one<EPI_PLAIN, 128>())) return rc;
  if ((rc = g2_init_one<EPI_GATED, 128>())) return rc;
  if ((rc = g2_init_one<EPI_QKV, 128>())) return rc;
  if ((rc = g2_init_one<EPI_QKV, 64>())) return rc;
  cudaFuncAttributes fa;
  return cudaFuncGetAttributes(&fa, quant_f32_kernel) == cudaSuccess ? OK : ERR_ARCH;
}

}  // namespace b2
