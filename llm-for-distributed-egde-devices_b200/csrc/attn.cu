// Rotary embedding + KV-cache append, decode attention (split-KV, HBM-bound) and causal prefill
// attention over the stage-local KV cache.
//
// Replaces `eager_attention_forward` + `apply_rotary_pos_emb` + `DynamicCache.update` that the
// reference executes inside HF `generate` (call site Code/Quantised Models/models_quant_updated.py:84-93;
// transformers modeling_llama.py:199-222, :128-153, :251-289).  Oracle: oracle/w8a8.py apply_rope
// (bit-exact) and attention (tolerance: softmax sum order / exp2 differ).
//
// KV cache layout (per layer): K and V each [rows][KV][Smax][hd] fp16 -- for one (row, kv head) the
// keys are contiguous, so a CTA streams its key range with fully coalesced 16-byte cp.async.
//
// Decode attention: grid = (kv head, sequence, kv split).  The G = H/KV query heads that share a kv
// head are packed into the 16-row A operand of mma.sync.m16n8k16 so K/V are read once per group.
// Each CTA streams its key range through a double-buffered cp.async ring (128 keys a stage), the 4
// warps take 32 keys each, keep private online-softmax state and are merged through shared memory.
// With splits > 1 the unnormalised partials go to a caller-provided workspace and a second small
// kernel merges them.
#include "common.cuh"

namespace b2 {

// ---------------------------------------------------------------------------------------------
// small PTX helpers
// ---------------------------------------------------------------------------------------------
B2_DEVINL void cp_async16(uint32_t dst, const void* src, bool valid) {
  const int sz = valid ? 16 : 0;   // src-size 0 -> zero fill
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(sz) : "memory");
}
B2_DEVINL void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
B2_DEVINL void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

B2_DEVINL void ldsm_x4(uint32_t (&r)[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
B2_DEVINL void ldsm_x4_t(uint32_t (&r)[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
// D(16x8,f32) += A(16x16,f16,row) * B(16x8,f16,col)
B2_DEVINL void mma16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

// =============================================================================================
// RoPE + KV append
// =============================================================================================
// qkv: [T][ld] fp16 with q heads at [0,H*hd), k at [H*hd,(H+KV)*hd), v after.  q is rotated in
// place, rotated k and v are written to the cache at (row = seq_ids[t] or seq0, positions[t]).
__global__ void __launch_bounds__(256)
rope_kv_append_kernel(__half* __restrict__ qkv, long long ld, const float* __restrict__ cosT,
                      const float* __restrict__ sinT, const int* __restrict__ positions,
                      const int* __restrict__ seq_ids, int seq0, __half* __restrict__ kcache,
                      __half* __restrict__ vcache, int H, int KV, int hd, int Smax, int use_rope) {
  const int t = blockIdx.x;
  const int pos = positions[t];
  const int row = seq_ids ? seq_ids[t] : seq0;
  const int half = hd >> 1;
  const int cph = half >> 3;                  // 8-element chunks per half head
  const int items = (H + 2 * KV) * cph;
  __half* base = qkv + (long long)t * ld;
  for (int it = threadIdx.x; it < items; it += blockDim.x) {
    const int head = it / cph, j = it - head * cph;
    __half* hp = base + head * hd;
    uint4 u1 = ld_v4(hp + j * 8);
    uint4 u2 = ld_v4(hp + half + j * 8);
    if (head < H + KV && use_rope) {
      const float4 cA = *reinterpret_cast<const float4*>(cosT + (long long)pos * half + j * 8);
      const float4 cB = *reinterpret_cast<const float4*>(cosT + (long long)pos * half + j * 8 + 4);
      const float4 sA = *reinterpret_cast<const float4*>(sinT + (long long)pos * half + j * 8);
      const float4 sB = *reinterpret_cast<const float4*>(sinT + (long long)pos * half + j * 8 + 4);
      const float c[8] = {cA.x, cA.y, cA.z, cA.w, cB.x, cB.y, cB.z, cB.w};
      const float s[8] = {sA.x, sA.y, sA.z, sA.w, sB.x, sB.y, sB.z, sB.w};
      const uint32_t* p1 = &u1.x;
      const uint32_t* p2 = &u2.x;
      uint32_t o1[4], o2[4];
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const float2 a = h2_to_f2(p1[e]), b = h2_to_f2(p2[e]);
        const float x0 = sub_rn(mul_rn(a.x, c[2 * e]), mul_rn(b.x, s[2 * e]));
        const float x1 = sub_rn(mul_rn(a.y, c[2 * e + 1]), mul_rn(b.y, s[2 * e + 1]));
        const float y0 = add_rn(mul_rn(b.x, c[2 * e]), mul_rn(a.x, s[2 * e]));
        const float y1 = add_rn(mul_rn(b.y, c[2 * e + 1]), mul_rn(a.y, s[2 * e + 1]));
        o1[e] = f2_to_h2(x0, x1);
        o2[e] = f2_to_h2(y0, y1);
      }
      u1 = make_uint4(o1[0], o1[1], o1[2], o1[3]);
      u2 = make_uint4(o2[0], o2[1], o2[2], o2[3]);
    }
    if (head < H) {
      st_v4(hp + j * 8, u1);
      st_v4(hp + half + j * 8, u2);
    } else {
      const bool isk = head < H + KV;
      const int kvh = isk ? head - H : head - H - KV;
      __half* dst = (isk ? kcache : vcache) + (((long long)row * KV + kvh) * Smax + pos) * hd;
      st_v4(dst + j * 8, u1);
      st_v4(dst + half + j * 8, u2);
    }
  }
}

int rope_kv_append(void* qkv, long long ld, const float* cosT, const float* sinT, const int* positions,
                   const int* seq_ids, int seq0, void* kcache, void* vcache, int T, int H, int KV, int hd, int Smax,
                   int use_rope, cudaStream_t st) {
  if (T == 0) return OK;
  if (!qkv || !positions || !kcache || !vcache || (use_rope && (!cosT || !sinT))) return ERR_NULL_PTR;
  if (T < 0 || hd % 16 || H <= 0 || KV <= 0 || (ld & 7)) return ERR_BAD_SHAPE;
  rope_kv_append_kernel<<<T, 256, 0, st>>>((__half*)qkv, ld, cosT, sinT, positions, seq_ids, seq0, (__half*)kcache,
                                           (__half*)vcache, H, KV, hd, Smax, use_rope);
  return launch_status();
}

// =============================================================================================
// decode attention
// =============================================================================================
constexpr int DEC_TK = 64;        // keys per stage (16 per warp)
// cp.async stages: 3 = 192 keys (K and V) in flight per CTA, two CTAs per SM; 2 for head_dim 256 (GPT-NeoX), whose
// stages are twice as large and must leave room for the cluster landing zone
template <int HD>
constexpr int dec_nst() { return HD > 128 ? 2 : 3; }
template <int HD>
constexpr int dec_smem() { return dec_nst<HD>() * 2 * DEC_TK * (HD + 8) * 2; }
template <int HD>
constexpr int dec_smem_cl() { return dec_smem<HD>() + 8 * 8 * (HD + 2) * 4; }
constexpr int DEC_THREADS = 128;

struct DecParams {
  const __half* q; long long q_ld;     // q[b] at q + b*q_ld, heads contiguous
  const __half* k; const __half* v;    // caches
  const int* seqlens;                  // [B] keys valid per sequence (incl. the new token)
  const int* rows;                     // [B] cache row per sequence (nullable -> b)
  __half* out; long long o_ld;
  float* part_o;                       // [B][H][splits][hd]
  float* part_ml;                      // [B][H][splits][2]
  int H, KV, Smax, splits, window;
  float scale_log2;
  // L2 hints for weights later launches stream (this kernel leaves HBM bandwidth to spare)
  const uint8_t* pf0; long long pf0_bytes; const uint8_t* pf1; long long pf1_bytes;
};
// this CTA's share of an L2 prefetch of [ptr, ptr+bytes): 16 KB bulk-prefetch requests
B2_DEVINL void dec_prefetch_slice(const uint8_t* ptr, long long bytes, int cta, int nctas) {
  if (ptr == nullptr || bytes <= 0) return;
  constexpr long long PIECE = 16 * 1024;
  const long long npieces = (bytes + PIECE - 1) / PIECE;
  if (threadIdx.x & 31) return;                      // lane 0 of each warp: the requests serialise within a warp
  const int warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
  for (long long i = cta + (long long)nctas * warp; i < npieces; i += (long long)nctas * nwarps) {
    const long long off = i * PIECE;
    const uint32_t n = (uint32_t)((bytes - off < PIECE ? bytes - off : PIECE) & ~15ll);
    if (n) asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(ptr + off), "r"(n) : "memory");
  }
}

B2_DEVINL void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
B2_DEVINL void cluster_arrive_release() { asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory"); }
B2_DEVINL void cluster_wait_acquire() { asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory"); }
// write a float to the same shared-memory offset in CTA `rank` of this cluster (DSMEM)
B2_DEVINL void st_dsmem_f32(float* local_ptr, uint32_t rank, float v) {
  uint32_t ra;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(ra) : "r"(smem_u32(local_ptr)), "r"(rank));
  asm volatile("st.shared::cluster.f32 [%0], %1;" ::"r"(ra), "f"(v) : "memory");
}

// CL = true: the `splits` CTAs of one (kv head, sequence) form a thread-block cluster (1,1,splits).
// Every CTA PUSHES its partial into a landing zone behind rank 0's K/V ring (remote shared-memory
// stores), arrives at the cluster barrier and exits; rank 0 waits once and merges from its own
// shared memory -- no workspace, no merge launch, one barrier, no remote loads.
template <int HD, bool CL>
__global__ void __launch_bounds__(DEC_THREADS, 2) decode_attn_kernel(const DecParams p) {
  constexpr int LDS = HD + 8;                       // padded row (halves): ldmatrix conflict-free
  constexpr int DEC_NST = dec_nst<HD>();
  constexpr int TILE_HALVES = DEC_TK * LDS;
  extern __shared__ __align__(16) uint8_t smem_dec[];
  __half* sK = reinterpret_cast<__half*>(smem_dec);   // [NST][TK][LDS]
  __half* sV = sK + DEC_NST * TILE_HALVES;            // [NST][TK][LDS]
  const int kvh = blockIdx.x, b = blockIdx.y, split = blockIdx.z;
  const int G = p.H / p.KV;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = lane >> 2, t4 = lane & 3;
  const int seqlen = p.seqlens[b];
  const int row = p.rows ? p.rows[b] : b;
  const int first = (p.window > 0 && seqlen > p.window) ? seqlen - p.window : 0;
  const int n = seqlen - first;
  const int chunk = (((n + p.splits - 1) / p.splits) + 15) & ~15;
  const int kbeg = first + split * chunk;
  const int kend = min(seqlen, kbeg + chunk);
  const int ntiles = kend > kbeg ? (kend - kbeg + DEC_TK - 1) / DEC_TK : 0;
  const __half* kg = p.k + ((long long)row * p.KV + kvh) * p.Smax * HD;
  const __half* vg = p.v + ((long long)row * p.KV + kvh) * p.Smax * HD;

  auto load_tile = [&](int tile, int stage) {
    const int k0 = kbeg + tile * DEC_TK;
    constexpr int CPR = HD / 8;                       // 16-byte chunks per row
    for (int i = threadIdx.x; i < DEC_TK * CPR; i += DEC_THREADS) {
      const int r = i / CPR, c = i - r * CPR;
      // the newest row (seqlen-1) is being written by the predecessor kernel: never prefetch it here,
      // it is patched into shared memory after the dependency wait
      const bool ok = (k0 + r) < kend && (k0 + r) != seqlen - 1;
      const long long goff = (long long)(ok ? k0 + r : kbeg) * HD + c * 8;
      const int soff = stage * TILE_HALVES + r * LDS + c * 8;
      cp_async16(smem_u32(sK + soff), kg + goff, ok);
      cp_async16(smem_u32(sV + soff), vg + goff, ok);
    }
  };
  // Programmatic dependent launch: everything up to the wait touches only state older than the
  // predecessor (sequence lengths, cache rows of earlier tokens).
  tl_mark(1, 0);
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  // Cluster variant: remote shared-memory stores are only defined once every CTA of the cluster has started.
  // Arrive now, wait just before the first remote store -- the K/V loop in between hides the latency.
  if (CL) asm volatile("barrier.cluster.arrive.relaxed.aligned;" ::: "memory");
  // fill the ring: one commit group per stage (empty groups keep the group count uniform)
#pragma unroll
  for (int sidx = 0; sidx < DEC_NST; ++sidx) {
    if (sidx < ntiles) load_tile(sidx, sidx);
    cp_async_commit();
  }
  {
    const int cta = blockIdx.x + gridDim.x * (blockIdx.y + gridDim.y * blockIdx.z);
    const int nctas = gridDim.x * gridDim.y * gridDim.z;
    dec_prefetch_slice(p.pf0, p.pf0_bytes, cta, nctas);
    dec_prefetch_slice(p.pf1, p.pf1_bytes, cta, nctas);
  }

  tl_mark(1, 1);
  asm volatile("griddepcontrol.wait;" ::: "memory");
  tl_mark(1, 2);
  // the newest key/value row, if it falls in this CTA's range: one 32-bit word of K and of V per thread
  const bool has_new = (seqlen - 1) >= kbeg && (seqlen - 1) < kend;
  uint32_t newk = 0u, newv = 0u;
  if (has_new && threadIdx.x < HD / 2) {
    newk = *reinterpret_cast<const uint32_t*>(kg + (long long)(seqlen - 1) * HD + threadIdx.x * 2);
    newv = *reinterpret_cast<const uint32_t*>(vg + (long long)(seqlen - 1) * HD + threadIdx.x * 2);
  }
  const int new_tile = has_new ? (seqlen - 1 - kbeg) / DEC_TK : -1;
  // Q fragments: rows 0..G-1 = the query heads of this kv group, rows >= G zero
  uint32_t qf[HD / 16][4];
  {
    const __half* qb = p.q + (long long)b * p.q_ld + (long long)kvh * G * HD;
#pragma unroll
    for (int kk = 0; kk < HD / 16; ++kk) {
      qf[kk][0] = (g < G) ? *reinterpret_cast<const uint32_t*>(qb + g * HD + kk * 16 + 2 * t4) : 0u;
      qf[kk][2] = (g < G) ? *reinterpret_cast<const uint32_t*>(qb + g * HD + kk * 16 + 8 + 2 * t4) : 0u;
      qf[kk][1] = 0u;   // rows g+8 >= 8 >= G
      qf[kk][3] = 0u;
    }
  }
  float o[HD / 8][4];
#pragma unroll
  for (int i = 0; i < HD / 8; ++i) { o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f; }
  float m_run = -INFINITY, l_run = 0.f;     // row g (rows g+8 are padding)

  for (int tile = 0; tile < ntiles; ++tile) {
    const int stage = tile % DEC_NST;
    cp_async_wait<DEC_NST - 1>();           // the oldest outstanding group (= this tile) has landed
    if (tile == new_tile && threadIdx.x < HD / 2) {
      const int r = (seqlen - 1 - kbeg) - tile * DEC_TK;
      *reinterpret_cast<uint32_t*>(sK + stage * TILE_HALVES + r * LDS + threadIdx.x * 2) = newk;
      *reinterpret_cast<uint32_t*>(sV + stage * TILE_HALVES + r * LDS + threadIdx.x * 2) = newv;
    }
    __syncthreads();
    const __half* tK = sK + stage * TILE_HALVES + warp * 16 * LDS;
    const __half* tV = sV + stage * TILE_HALVES + warp * 16 * LDS;
    // ---- S = Q K^T for this warp's 16 keys ----
    float s[2][4];
#pragma unroll
    for (int nb = 0; nb < 2; ++nb) { s[nb][0] = s[nb][1] = s[nb][2] = s[nb][3] = 0.f; }
#pragma unroll
    for (int kp = 0; kp < HD / 32; ++kp) {
#pragma unroll
      for (int nb = 0; nb < 2; ++nb) {
        uint32_t bf[4];
        const int r = nb * 8 + (lane & 7);
        const int c = kp * 32 + (lane >> 3) * 8;
        ldsm_x4(bf, smem_u32(tK + r * LDS + c));
        mma16816(s[nb], qf[2 * kp], bf[0], bf[1]);
        mma16816(s[nb], qf[2 * kp + 1], bf[2], bf[3]);
      }
    }
    // ---- mask + online softmax (row g only) ----
    const int key0 = kbeg + tile * DEC_TK + warp * 16;
    float mx = -INFINITY;
#pragma unroll
    for (int nb = 0; nb < 2; ++nb) {
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        const int key = key0 + nb * 8 + 2 * t4 + e;
        float v = s[nb][e] * p.scale_log2;
        v = (key < kend) ? v : -INFINITY;
        s[nb][e] = v;
        mx = fmaxf(mx, v);
      }
    }
    mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
    mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
    const float m_new = fmaxf(m_run, mx);
    const float m_use = (m_new == -INFINITY) ? 0.f : m_new;
    const float corr = exp2f(m_run - m_use);          // m_run = -inf -> 0
    float ls = 0.f;
    uint32_t pf[4];
#pragma unroll
    for (int nb = 0; nb < 2; ++nb) {
      const float p0 = exp2f(s[nb][0] - m_use);
      const float p1 = exp2f(s[nb][1] - m_use);
      ls += p0 + p1;
      // C layout of S -> A layout of P: regs {0,1} = key cols 0-7, {2,3} = cols 8-15
      pf[nb * 2] = f2_to_h2(p0, p1);
      pf[nb * 2 + 1] = 0u;                            // rows g+8: padding
    }
    ls += __shfl_xor_sync(0xffffffffu, ls, 1);
    ls += __shfl_xor_sync(0xffffffffu, ls, 2);
    l_run = l_run * corr + ls;
    m_run = m_new;
#pragma unroll
    for (int i = 0; i < HD / 8; ++i) { o[i][0] *= corr; o[i][1] *= corr; }
    // ---- O += P V (16 keys) ----
#pragma unroll
    for (int np = 0; np < HD / 16; ++np) {
      uint32_t bf[4];
      const int mat = lane >> 3;
      const int r = (mat & 1) * 8 + (lane & 7);
      const int c = np * 16 + (mat >> 1) * 8;
      ldsm_x4_t(bf, smem_u32(tV + r * LDS + c));
      mma16816(o[2 * np], pf, bf[0], bf[1]);
      mma16816(o[2 * np + 1], pf, bf[2], bf[3]);
    }
    __syncthreads();   // all warps done with this stage before it is refilled
    if (tile + DEC_NST < ntiles) load_tile(tile + DEC_NST, stage);
    cp_async_commit();
  }
  cp_async_wait<0>();
  __syncthreads();
  tl_mark(1, 4);
  // ---- merge the 4 warps through shared memory (reuse the K buffers) ----
  float* sO = reinterpret_cast<float*>(smem_dec);          // [4][8][HD]
  float* sM = sO + 4 * 8 * HD;                             // [4][8]
  float* sL = sM + 32;                                     // [4][8]
#pragma unroll
  for (int i = 0; i < HD / 8; ++i) {
    sO[(warp * 8 + g) * HD + i * 8 + 2 * t4] = o[i][0];
    sO[(warp * 8 + g) * HD + i * 8 + 2 * t4 + 1] = o[i][1];
  }
  if (t4 == 0) { sM[warp * 8 + g] = m_run; sL[warp * 8 + g] = l_run; }
  __syncthreads();
  if (CL) asm volatile("barrier.cluster.wait.aligned;" ::: "memory");     // phase 1: every CTA of the cluster is running
  for (int idx = threadIdx.x; idx < G * HD; idx += DEC_THREADS) {
    const int gh = idx / HD, d = idx - gh * HD;
    float m = -INFINITY;
#pragma unroll
    for (int w = 0; w < 4; ++w) m = fmaxf(m, sM[w * 8 + gh]);
    float l = 0.f, acc = 0.f;
    if (m != -INFINITY) {
#pragma unroll
      for (int w = 0; w < 4; ++w) {
        const float f = exp2f(sM[w * 8 + gh] - m);
        l += sL[w * 8 + gh] * f;
        acc += sO[(w * 8 + gh) * HD + d] * f;
      }
    }
    const int h = kvh * G + gh;
    if (CL) {
      float* slot = reinterpret_cast<float*>(smem_dec + DEC_NST * 2 * TILE_HALVES * 2) + split * (G * (HD + 2));
      st_dsmem_f32(slot + idx, 0, acc);
      if (d == 0) { st_dsmem_f32(slot + G * HD + gh, 0, m); st_dsmem_f32(slot + G * HD + G + gh, 0, l); }
    } else if (p.splits == 1) {
      p.out[(long long)b * p.o_ld + h * HD + d] = __float2half_rn(l > 0.f ? acc / l : 0.f);
    } else {
      const long long po = (((long long)b * p.H + h) * p.splits + split);
      p.part_o[po * HD + d] = acc;
      if (d == 0) { p.part_ml[po * 2] = m; p.part_ml[po * 2 + 1] = l; }
    }
  }
  if (CL) {
    cluster_arrive_release();                                // this CTA's partial is on its way to rank 0
    if (split == 0) {
      cluster_wait_acquire();                                // ... and everybody's has landed
      const float* land = reinterpret_cast<const float*>(smem_dec + DEC_NST * 2 * TILE_HALVES * 2);
      const int slot = G * (HD + 2);
      for (int idx = threadIdx.x; idx < G * HD; idx += DEC_THREADS) {
        const int gh = idx / HD, d = idx - gh * HD;
        float m = -INFINITY;
        for (int r = 0; r < p.splits; ++r) m = fmaxf(m, land[r * slot + G * HD + gh]);
        float l = 0.f, acc = 0.f;
        if (m != -INFINITY) {
          for (int r = 0; r < p.splits; ++r) {
            const float f = exp2f(land[r * slot + G * HD + gh] - m);
            l += land[r * slot + G * HD + G + gh] * f;
            acc += land[r * slot + idx] * f;
          }
        }
        p.out[(long long)b * p.o_ld + (kvh * G + gh) * HD + d] = __float2half_rn(l > 0.f ? acc / l : 0.f);
      }
    }
  }
  tl_mark(1, 5);
  tl_flush(1, 0x37u);
}

template <int HD>
__global__ void __launch_bounds__(HD) decode_merge_kernel(const DecParams p) {
  const int h = blockIdx.x, b = blockIdx.y, d = threadIdx.x;
  const long long base = ((long long)b * p.H + h) * p.splits;
  float m = -INFINITY;
  for (int s = 0; s < p.splits; ++s) m = fmaxf(m, p.part_ml[(base + s) * 2]);
  float l = 0.f, acc = 0.f;
  if (m != -INFINITY) {
    for (int s = 0; s < p.splits; ++s) {
      const float f = exp2f(p.part_ml[(base + s) * 2] - m);
      l += p.part_ml[(base + s) * 2 + 1] * f;
      acc += p.part_o[(base + s) * HD + d] * f;
    }
  }
  p.out[(long long)b * p.o_ld + h * HD + d] = __float2half_rn(l > 0.f ? acc / l : 0.f);
}

// How many kv splits the decode kernel will use for these sizes; the caller sizes the workspace
// with decode_attn_workspace_bytes() (0 when splits == 1).
int decode_attn_splits(int B, int KV, int max_seqlen, int sms) {
  if (B <= 0 || KV <= 0) return 1;
  const int ctas = B * KV;
  // About 1.25 CTAs per SM.  Measured on B200 at B*KV = 32, 2.1k keys (decode us/token over 8 layers):
  // 3 splits 586, 4: 562, 5: 545, 6: 537, 7: 536, 8: 571 -- with 8 the 32 clusters of 8 only just fit
  // the GPCs and one or two regularly wait for a whole wave (tools/timeline.py), fewer leave too few
  // bytes in flight.
  int want = (5 * sms / 4 + ctas / 2) / ctas;
  const int max_by_len = (max_seqlen + 255) / 256;   // >= 256 keys per split
  int s = want < max_by_len ? want : max_by_len;
  if (s < 1) s = 1;
  if (s > 8 && s <= 12) s = 8;      // a cluster of 8 merges in-kernel (DSMEM): cheaper than 9..12 + merge launch
  if (s > 64) s = 64;
  return s;
}
size_t decode_attn_workspace_bytes(int B, int H, int hd, int splits) {
  if (splits <= 1) return 0;
  return (size_t)B * H * splits * (hd + 2) * sizeof(float);
}

static int sm_count_attn() { return device_sm_count(); }

template <int HD>
static int launch_decode(const DecParams& p, int B, cudaStream_t st, bool cluster = false) {
  constexpr int SMEM = dec_smem<HD>();
  constexpr int SMEM_CL_MAX = dec_smem_cl<HD>();               // + landing zone: 8 splits x 8 heads of a group
  static DeviceOnce attr_once;
  if (!attr_once.get()) {
    if (cudaFuncSetAttribute(decode_attn_kernel<HD, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM) != cudaSuccess ||
        cudaFuncSetAttribute(decode_attn_kernel<HD, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_CL_MAX) != cudaSuccess)
      return ERR_ARCH;
    attr_once.set();
  }
  if (cluster) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(p.KV, B, p.splits);
    cfg.blockDim = dim3(DEC_THREADS);
    cfg.dynamicSmemBytes = SMEM + p.splits * (p.H / p.KV) * (HD + 2) * 4;
    cfg.stream = st;
    cudaLaunchAttribute at[2];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = 1; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = p.splits;
    at[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = 2;
    if (cudaLaunchKernelEx(&cfg, decode_attn_kernel<HD, true>, p) != cudaSuccess) { cudaGetLastError(); return ERR_LAUNCH; }
    return OK;
  }
  {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(p.KV, B, p.splits);
    cfg.blockDim = dim3(DEC_THREADS);
    cfg.dynamicSmemBytes = SMEM;
    cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    if (cudaLaunchKernelEx(&cfg, decode_attn_kernel<HD, false>, p) != cudaSuccess) { cudaGetLastError(); return ERR_LAUNCH; }
  }
  if (p.splits > 1) decode_merge_kernel<HD><<<dim3(p.H, B), HD, 0, st>>>(p);
  return launch_status();
}

// max_seqlen: upper bound of seqlens[] known to the host (sizes the split count; no device read)
int decode_attn(const void* q, long long q_ld, const void* kcache, const void* vcache, const int* seqlens,
                const int* rows, void* out, long long o_ld, void* workspace, size_t workspace_bytes, int B, int H,
                int KV, int hd, int Smax, int max_seqlen, float scale, int window, const void* pf0, long long pf0_bytes,
                const void* pf1, long long pf1_bytes, cudaStream_t st) {
  if (B == 0) return OK;
  if (!q || !kcache || !vcache || !seqlens || !out) return ERR_NULL_PTR;
  if (B < 0 || H <= 0 || KV <= 0 || H % KV || H / KV > 8 || (q_ld & 1) || max_seqlen <= 0) return ERR_BAD_SHAPE;
  if (hd != 64 && hd != 96 && hd != 128 && hd != 256) return ERR_UNSUPPORTED;
  const int eff = (window > 0 && max_seqlen > window) ? window : max_seqlen;
  int splits = decode_attn_splits(B, KV, eff, sm_count_attn());
  // 2..8 splits: merged inside the kernel by a thread-block cluster over DSMEM (workspace unused).
  const bool cluster = splits > 1 && splits <= 8;
  const size_t need = cluster ? 0 : decode_attn_workspace_bytes(B, H, hd, splits);
  if (need > 0 && (!workspace || workspace_bytes < need)) {
    // shrink to what the workspace allows rather than failing; <= 8 splits need none
    while (splits > 8 && (!workspace || decode_attn_workspace_bytes(B, H, hd, splits) > workspace_bytes)) --splits;
  }
  const bool use_cluster = splits > 1 && splits <= 8;
  DecParams p;
  p.q = (const __half*)q; p.q_ld = q_ld; p.k = (const __half*)kcache; p.v = (const __half*)vcache;
  p.seqlens = seqlens; p.rows = rows; p.out = (__half*)out; p.o_ld = o_ld;
  p.part_o = (float*)workspace;
  p.part_ml = p.part_o ? p.part_o + (size_t)B * H * splits * hd : nullptr;
  p.H = H; p.KV = KV; p.Smax = Smax; p.splits = splits; p.window = window;
  p.scale_log2 = scale * 1.4426950408889634f;
  p.pf0 = (const uint8_t*)pf0; p.pf0_bytes = pf0_bytes; p.pf1 = (const uint8_t*)pf1; p.pf1_bytes = pf1_bytes;
  return hd == 128 ? launch_decode<128>(p, B, st, use_cluster) : hd == 256 ? launch_decode<256>(p, B, st, use_cluster)
       : hd == 96 ? launch_decode<96>(p, B, st, use_cluster) : launch_decode<64>(p, B, st, use_cluster);
}

// =============================================================================================
// decode attention for callers that may not allocate
// =============================================================================================
// Some hosts of this library cannot hand us a workspace because they cannot obtain one: a serving
// process that gave all of HBM to its KV-cache pool at start-up and runs with the allocator closed,
// a latency-critical decode thread where cudaMalloc (which synchronises the device and has no
// time bound) is banned, or a plugin host that owns device memory and forbids plugins to allocate.
// For them the split-KV partials live in the fixed array below, which is part of the library
// image: every GPU that loads libb200stage.so carries these bytes until the process exits.
constexpr size_t DEC_STATIC_SCRATCH_BYTES = 1u << 19;   // 512 KiB, == B2_DECODE_ATTN_STATIC_SCRATCH_BYTES
static __device__ __align__(256) unsigned char g_dec_static_scratch[DEC_STATIC_SCRATCH_BYTES];

// `g_dec_static_scratch` named in HOST code is only the host-side handle of the symbol, not a
// device address; the address a kernel may dereference is looked up per device, once, here.
constexpr int DEC_STATIC_MAX_DEVICES = 64;
static void* g_dec_static_addr[DEC_STATIC_MAX_DEVICES] = {nullptr};

static int dec_static_scratch_ptr(void** out) {
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= DEC_STATIC_MAX_DEVICES) return ERR_ARCH;
  void* p = __atomic_load_n(&g_dec_static_addr[dev], __ATOMIC_ACQUIRE);
  if (!p) {
    // First call on this device (it is the current device): resolve the symbol's device address.
    // Racing first calls resolve the same address, so a plain publish is enough.
    if (cudaGetSymbolAddress(&p, g_dec_static_scratch) != cudaSuccess || !p) {
      cudaGetLastError();
      return ERR_DRIVER;
    }
    __atomic_store_n(&g_dec_static_addr[dev], p, __ATOMIC_RELEASE);
  }
  *out = p;
  return OK;
}

// Same computation and same kernels as decode_attn(); only where the partials live differs.
// Nothing in the array has to start at zero: every (sequence, head, split) slot that the merge
// kernel reads is written by the attention kernel of the same call, including empty splits.
int decode_attn_static(const void* q, long long q_ld, const void* kcache, const void* vcache, const int* seqlens,
                       const int* rows, void* out, long long o_ld, int B, int H, int KV, int hd, int Smax,
                       int max_seqlen, float scale, int window, cudaStream_t st) {
  if (B == 0) return OK;
  if (!q || !kcache || !vcache || !seqlens || !out) return ERR_NULL_PTR;
  if (B < 0 || H <= 0 || KV <= 0 || H % KV || H / KV > 8 || (q_ld & 1) || max_seqlen <= 0) return ERR_BAD_SHAPE;
  if (hd != 64 && hd != 96 && hd != 128 && hd != 256) return ERR_UNSUPPORTED;
  const int eff = (window > 0 && max_seqlen > window) ? window : max_seqlen;
  const int splits = decode_attn_splits(B, KV, eff, sm_count_attn());
  const size_t need = decode_attn_workspace_bytes(B, H, hd, splits);
  if (need > DEC_STATIC_SCRATCH_BYTES) return ERR_TOO_LARGE;   // no silent shrink: the caller must know
  void* scratch = nullptr;
  if (need > 0) {
    const int rc = dec_static_scratch_ptr(&scratch);
    if (rc) return rc;
  }
  DecParams p;
  p.q = (const __half*)q; p.q_ld = q_ld; p.k = (const __half*)kcache; p.v = (const __half*)vcache;
  p.seqlens = seqlens; p.rows = rows; p.out = (__half*)out; p.o_ld = o_ld;
  p.part_o = (float*)scratch;
  p.part_ml = p.part_o ? p.part_o + (size_t)B * H * splits * hd : nullptr;
  p.H = H; p.KV = KV; p.Smax = Smax; p.splits = splits; p.window = window;
  p.scale_log2 = scale * 1.4426950408889634f;
  p.pf0 = nullptr; p.pf0_bytes = 0; p.pf1 = nullptr; p.pf1_bytes = 0;
  return hd == 128 ? launch_decode<128>(p, B, st) : hd == 256 ? launch_decode<256>(p, B, st)
       : hd == 96 ? launch_decode<96>(p, B, st) : launch_decode<64>(p, B, st);
}

// =============================================================================================
// prefill attention (one contiguous chunk of one sequence; causal, optional sliding window)
// =============================================================================================
constexpr int PF_TK = 64;     // keys per stage
constexpr int PF_THREADS = 128;

struct PfParams {
  const __half* q; long long q_ld;     // q[t] at q + t*q_ld
  const __half* k; const __half* v;
  __half* out; long long o_ld;
  const int* pos_ptr;                  // device: positions of the chunk (pos0 = pos_ptr[0]) or nullptr
  int T, pos0, row, H, KV, Smax, window;
  float scale_log2;
};

// MB = 16-row MMA blocks per warp: 2 (128 query rows per CTA) for head_dim <= 128; 1 (64 rows) for head_dim 256
// (GPT-NeoX / Pythia-1B), where two output accumulators of 256 columns would not fit the register file.
template <int HD>
constexpr int pf_mb() { return HD > 128 ? 1 : 2; }
template <int HD>
constexpr int pf_smem_bytes() { return (64 * pf_mb<HD>() + 2 * 2 * PF_TK) * (HD + 8) * 2; }

// FA2-style: every K (and V) fragment fetched with ldmatrix is used by both 16-row blocks of the
// warp, the Q tile lives in shared memory and its fragments are re-read per key tile (keeping them
// in registers next to two O accumulators would not fit).
template <int HD>
__global__ void __launch_bounds__(PF_THREADS) prefill_attn_kernel(const PfParams pin) {
  PfParams p = pin;
  if (p.pos_ptr) p.pos0 = p.pos_ptr[0];
  constexpr int MB = pf_mb<HD>();
  constexpr int BM = 64 * MB;          // query rows per CTA: 4 warps x MB blocks of 16
  constexpr int LDS = HD + 8;
  constexpr int TILE_HALVES = PF_TK * LDS;
  extern __shared__ __align__(16) uint8_t smem_pf[];
  __half* sQ = reinterpret_cast<__half*>(smem_pf);     // [BM][LDS]
  __half* sK = sQ + BM * LDS;                       // [2][TK][LDS]
  __half* sV = sK + 2 * TILE_HALVES;                   // [2][TK][LDS]
  const int qt = gridDim.x - 1 - blockIdx.x;           // heavy (late) query tiles first
  const int h = blockIdx.y;
  const int kvh = h / (p.H / p.KV);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = lane >> 2, t4 = lane & 3;
  const int q0 = qt * BM;
  const int last_q = min(q0 + BM, p.T) - 1;
  const int kend = p.pos0 + last_q + 1;                // exclusive
  int kfirst = 0;
  if (p.window > 0) { kfirst = p.pos0 + q0 - p.window + 1; kfirst = kfirst > 0 ? (kfirst / PF_TK) * PF_TK : 0; }
  const int ntiles = (kend - kfirst + PF_TK - 1) / PF_TK;
  const __half* kg = p.k + ((long long)p.row * p.KV + kvh) * p.Smax * HD;
  const __half* vg = p.v + ((long long)p.row * p.KV + kvh) * p.Smax * HD;
  constexpr int CPR = HD / 8;

  auto load_tile = [&](int tile, int stage) {
    const int k0 = kfirst + tile * PF_TK;
    for (int i = threadIdx.x; i < PF_TK * CPR; i += PF_THREADS) {
      const int r = i / CPR, c = i - r * CPR;
      const bool ok = (k0 + r) < kend;
      const long long goff = (long long)(ok ? k0 + r : kfirst) * HD + c * 8;
      const int soff = stage * TILE_HALVES + r * LDS + c * 8;
      cp_async16(smem_u32(sK + soff), kg + goff, ok);
      cp_async16(smem_u32(sV + soff), vg + goff, ok);
    }
  };
  // Q tile (rows beyond T zero-filled) travels with the first K/V tile
  for (int i = threadIdx.x; i < BM * CPR; i += PF_THREADS) {
    const int r = i / CPR, c = i - r * CPR;
    const bool ok = (q0 + r) < p.T;
    cp_async16(smem_u32(sQ + r * LDS + c * 8), p.q + (long long)(ok ? q0 + r : q0) * p.q_ld + h * HD + c * 8, ok);
  }
  load_tile(0, 0);
  cp_async_commit();

  float o[MB][HD / 8][4];
#pragma unroll
  for (int mb = 0; mb < MB; ++mb)
#pragma unroll
    for (int i = 0; i < HD / 8; ++i) { o[mb][i][0] = o[mb][i][1] = o[mb][i][2] = o[mb][i][3] = 0.f; }
  float m_run[MB][2], l_run[MB][2];
#pragma unroll
  for (int mb = 0; mb < MB; ++mb) { m_run[mb][0] = m_run[mb][1] = -INFINITY; l_run[mb][0] = l_run[mb][1] = 0.f; }
  // absolute positions of this thread's four query rows: [mb][0] = row g, [mb][1] = row g + 8
  int qpos[MB][2];
#pragma unroll
  for (int mb = 0; mb < MB; ++mb) {
    qpos[mb][0] = p.pos0 + q0 + warp * (16 * MB) + mb * 16 + g;
    qpos[mb][1] = qpos[mb][0] + 8;
  }
  const __half* wQ = sQ + (warp * 16 * MB) * LDS;

  for (int tile = 0; tile < ntiles; ++tile) {
    const int stage = tile & 1;
    if (tile + 1 < ntiles) load_tile(tile + 1, stage ^ 1);
    cp_async_commit();
    cp_async_wait<1>();
    __syncthreads();
    const __half* tK = sK + stage * TILE_HALVES;
    const __half* tV = sV + stage * TILE_HALVES;
    const int key0 = kfirst + tile * PF_TK;
    // warp-uniform skip: a key tile entirely above this warp's last query row contributes nothing
    const bool live = key0 <= p.pos0 + q0 + warp * (16 * MB) + 16 * MB - 1;
    if (live) {
      float s[MB][PF_TK / 8][4];
#pragma unroll
      for (int mb = 0; mb < MB; ++mb)
#pragma unroll
        for (int nb = 0; nb < PF_TK / 8; ++nb) { s[mb][nb][0] = s[mb][nb][1] = s[mb][nb][2] = s[mb][nb][3] = 0.f; }
#pragma unroll
      for (int kp = 0; kp < HD / 32; ++kp) {
        uint32_t qa[MB][2][4];            // [mb][k-block within the pair][regs]
#pragma unroll
        for (int mb = 0; mb < MB; ++mb)
#pragma unroll
          for (int kb = 0; kb < 2; ++kb) {
            const int mat = lane >> 3;
            const int r = mb * 16 + (mat & 1) * 8 + (lane & 7);
            const int c = kp * 32 + kb * 16 + (mat >> 1) * 8;
            ldsm_x4(qa[mb][kb], smem_u32(wQ + r * LDS + c));
          }
#pragma unroll
        for (int nb = 0; nb < PF_TK / 8; ++nb) {
          uint32_t bf[4];
          const int r = nb * 8 + (lane & 7);
          const int c = kp * 32 + (lane >> 3) * 8;
          ldsm_x4(bf, smem_u32(tK + r * LDS + c));
#pragma unroll
          for (int mb = 0; mb < MB; ++mb) {
            mma16816(s[mb][nb], qa[mb][0], bf[0], bf[1]);
            mma16816(s[mb][nb], qa[mb][1], bf[2], bf[3]);
          }
        }
      }
      uint32_t pf[MB][PF_TK / 16][4];
#pragma unroll
      for (int mb = 0; mb < MB; ++mb) {
        float mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll
        for (int nb = 0; nb < PF_TK / 8; ++nb) {
#pragma unroll
          for (int e = 0; e < 2; ++e) {
            const int key = key0 + nb * 8 + 2 * t4 + e;
            const bool v0 = key <= qpos[mb][0] && (p.window == 0 || key > qpos[mb][0] - p.window);
            const bool v1 = key <= qpos[mb][1] && (p.window == 0 || key > qpos[mb][1] - p.window);
            const float a = v0 ? s[mb][nb][e] * p.scale_log2 : -INFINITY;
            const float c = v1 ? s[mb][nb][2 + e] * p.scale_log2 : -INFINITY;
            s[mb][nb][e] = a; s[mb][nb][2 + e] = c;
            mx0 = fmaxf(mx0, a); mx1 = fmaxf(mx1, c);
          }
        }
        mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1));
        mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
        mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1));
        mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
        const float mn0 = fmaxf(m_run[mb][0], mx0), mn1 = fmaxf(m_run[mb][1], mx1);
        const float mu0 = (mn0 == -INFINITY) ? 0.f : mn0, mu1 = (mn1 == -INFINITY) ? 0.f : mn1;
        const float c0 = exp2f(m_run[mb][0] - mu0), c1 = exp2f(m_run[mb][1] - mu1);
        float ls0 = 0.f, ls1 = 0.f;
#pragma unroll
        for (int nb = 0; nb < PF_TK / 8; ++nb) {
          const float p00 = exp2f(s[mb][nb][0] - mu0), p01 = exp2f(s[mb][nb][1] - mu0);
          const float p10 = exp2f(s[mb][nb][2] - mu1), p11 = exp2f(s[mb][nb][3] - mu1);
          ls0 += p00 + p01; ls1 += p10 + p11;
          pf[mb][nb >> 1][(nb & 1) * 2] = f2_to_h2(p00, p01);
          pf[mb][nb >> 1][(nb & 1) * 2 + 1] = f2_to_h2(p10, p11);
        }
        ls0 += __shfl_xor_sync(0xffffffffu, ls0, 1); ls0 += __shfl_xor_sync(0xffffffffu, ls0, 2);
        ls1 += __shfl_xor_sync(0xffffffffu, ls1, 1); ls1 += __shfl_xor_sync(0xffffffffu, ls1, 2);
        l_run[mb][0] = l_run[mb][0] * c0 + ls0; l_run[mb][1] = l_run[mb][1] * c1 + ls1;
        m_run[mb][0] = mn0; m_run[mb][1] = mn1;
#pragma unroll
        for (int i = 0; i < HD / 8; ++i) { o[mb][i][0] *= c0; o[mb][i][1] *= c0; o[mb][i][2] *= c1; o[mb][i][3] *= c1; }
      }
#pragma unroll
      for (int kk2 = 0; kk2 < PF_TK / 16; ++kk2) {
#pragma unroll
        for (int np = 0; np < HD / 16; ++np) {
          uint32_t bf[4];
          const int mat = lane >> 3;
          const int r = kk2 * 16 + (mat & 1) * 8 + (lane & 7);
          const int c = np * 16 + (mat >> 1) * 8;
          ldsm_x4_t(bf, smem_u32(tV + r * LDS + c));
#pragma unroll
          for (int mb = 0; mb < MB; ++mb) {
            mma16816(o[mb][2 * np], pf[mb][kk2], bf[0], bf[1]);
            mma16816(o[mb][2 * np + 1], pf[mb][kk2], bf[2], bf[3]);
          }
        }
      }
    }
    __syncthreads();
  }
  cp_async_wait<0>();
#pragma unroll
  for (int mb = 0; mb < MB; ++mb) {
#pragma unroll
    for (int hf = 0; hf < 2; ++hf) {
      const int r = q0 + warp * (16 * MB) + mb * 16 + g + hf * 8;
      if (r < p.T) {
        const float l = l_run[mb][hf];
        const float inv = l > 0.f ? 1.f / l : 0.f;
        __half* op = p.out + (long long)r * p.o_ld + h * HD;
#pragma unroll
        for (int i = 0; i < HD / 8; ++i)
          *reinterpret_cast<uint32_t*>(op + i * 8 + 2 * t4) = f2_to_h2(o[mb][i][2 * hf] * inv, o[mb][i][2 * hf + 1] * inv);
      }
    }
  }
}

template <int HD>
static int launch_prefill(const PfParams& p, cudaStream_t st) {
  constexpr int SMEM = pf_smem_bytes<HD>();
  static DeviceOnce attr_once;
  if (!attr_once.get()) {
    if (cudaFuncSetAttribute(prefill_attn_kernel<HD>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM) != cudaSuccess)
      return ERR_ARCH;
    attr_once.set();
  }
  constexpr int BM = 64 * pf_mb<HD>();
  prefill_attn_kernel<HD><<<dim3((p.T + BM - 1) / BM, p.H), PF_THREADS, SMEM, st>>>(p);
  return launch_status();
}

template <int HD>
static bool init_dec() {
  cudaFuncAttributes fa;
  return cudaFuncGetAttributes(&fa, decode_merge_kernel<HD>) == cudaSuccess &&
         cudaFuncSetAttribute(decode_attn_kernel<HD, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, dec_smem<HD>()) == cudaSuccess &&
         cudaFuncSetAttribute(decode_attn_kernel<HD, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, dec_smem_cl<HD>()) == cudaSuccess;
}

int init_attn() {
  sm_count_attn();
  {
    cudaFuncAttributes fa;
    bool ok = cudaFuncGetAttributes(&fa, rope_kv_append_kernel) == cudaSuccess;
    ok &= cudaFuncGetAttributes(&fa, decode_merge_kernel<128>) == cudaSuccess;
    ok &= cudaFuncGetAttributes(&fa, decode_merge_kernel<64>) == cudaSuccess;
    ok &= cudaFuncGetAttributes(&fa, decode_merge_kernel<96>) == cudaSuccess;
    if (!ok) return ERR_ARCH;
  }
  if (!init_dec<64>() || !init_dec<96>() || !init_dec<128>() || !init_dec<256>()) return ERR_ARCH;
  // head_dim 96: Phi-2's 80-wide heads padded to a multiple of 32 (weights.head_slots); 256: GPT-NeoX / Pythia-1B
  if (cudaFuncSetAttribute(prefill_attn_kernel<96>, cudaFuncAttributeMaxDynamicSharedMemorySize, pf_smem_bytes<96>()) != cudaSuccess) return ERR_ARCH;
  if (cudaFuncSetAttribute(prefill_attn_kernel<256>, cudaFuncAttributeMaxDynamicSharedMemorySize, pf_smem_bytes<256>()) != cudaSuccess) return ERR_ARCH;
  if (cudaFuncSetAttribute(prefill_attn_kernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, pf_smem_bytes<128>()) != cudaSuccess) return ERR_ARCH;
  if (cudaFuncSetAttribute(prefill_attn_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, pf_smem_bytes<64>()) != cudaSuccess) return ERR_ARCH;
  return OK;
}

// attn_tc.cu: the tcgen05 kernel (head_dim 128, causal, optional sliding window)
bool prefill_attn_tc_supported(int T, int hd, int window, long long q_ld, const void* q, const void* k, const void* v);
int prefill_attn_tc(const void* q, long long q_ld, const void* kcache, const void* vcache, void* out, long long o_ld,
                    const int* pos_ptr, int T, int pos0, int row, int H, int KV, int Smax, float scale, int window, cudaStream_t st);
// test hook: 0 = choose (tcgen05 where supported), 1 = always the mma.sync kernel, 2 = tcgen05 or fail
static int g_prefill_impl = 0;
void set_prefill_attn_impl(int impl) { g_prefill_impl = impl; }

int prefill_attn(const void* q, long long q_ld, const void* kcache, const void* vcache, void* out, long long o_ld,
                 const int* positions, int T, int pos0, int row, int H, int KV, int hd, int Smax, float scale, int window,
                 cudaStream_t st) {
  if (T == 0) return OK;
  if (!q || !kcache || !vcache || !out) return ERR_NULL_PTR;
  if (T < 0 || (!positions && (pos0 < 0 || pos0 + T > Smax)) || H <= 0 || KV <= 0 || H % KV || (q_ld & 1) || (o_ld & 1)) return ERR_BAD_SHAPE;
  if (hd != 64 && hd != 96 && hd != 128 && hd != 256) return ERR_UNSUPPORTED;
  const bool tc_ok = prefill_attn_tc_supported(T, hd, window, q_ld, q, kcache, vcache) && !(o_ld & 7) &&
                     !(reinterpret_cast<uintptr_t>(out) & 15);
  if (g_prefill_impl == 2 && !tc_ok) return ERR_UNSUPPORTED;
  if (tc_ok && g_prefill_impl != 1)
    return prefill_attn_tc(q, q_ld, kcache, vcache, out, o_ld, positions, T, pos0, row, H, KV, Smax, scale, window, st);
  PfParams p;
  p.q = (const __half*)q; p.q_ld = q_ld; p.k = (const __half*)kcache; p.v = (const __half*)vcache;
  p.out = (__half*)out; p.o_ld = o_ld; p.pos_ptr = positions; p.T = T; p.pos0 = pos0; p.row = row; p.H = H; p.KV = KV; p.Smax = Smax;
  p.window = window; p.scale_log2 = scale * 1.4426950408889634f;
  return hd == 128 ? launch_prefill<128>(p, st) : hd == 256 ? launch_prefill<256>(p, st)
       : hd == 96 ? launch_prefill<96>(p, st) : launch_prefill<64>(p, st);
}

}  // namespace b2

B2_TL_DEFINE_DUMP(b2_tl_dump_attn)
