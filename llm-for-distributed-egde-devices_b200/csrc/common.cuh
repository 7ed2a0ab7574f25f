// Shared device helpers for the b200stage kernels (sm_100a only).
//
// PTX wrappers for mbarrier / TMA / tcgen05 / TMEM, the deterministic fp32 math pieces whose
// operation order is pinned by oracle/w8a8.py, and small vector load/store helpers.
#pragma once
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <stdint.h>

#define B2_DEVINL __device__ __forceinline__

namespace b2 {

// ---------------------------------------------------------------------------------------------
// error codes (mirrored in include/b200stage.h)
// ---------------------------------------------------------------------------------------------
enum : int {
  OK = 0,
  ERR_BAD_SHAPE = -1,
  ERR_BAD_ALIGN = -2,
  ERR_NULL_PTR = -3,
  ERR_ARCH = -4,
  ERR_LAUNCH = -5,
  ERR_UNSUPPORTED = -6,
  ERR_DRIVER = -7,
  ERR_TOO_LARGE = -8,
};

static inline int launch_status() {
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? OK : ERR_LAUNCH;
}

// ---------------------------------------------------------------------------------------------
// per-device host state: function attributes and the SM count belong to the device that is current
// when a launch path runs, not to the process (one process may drive several GPUs)
// ---------------------------------------------------------------------------------------------
constexpr int MAX_DEVICES = 64;
static inline int current_device() {
  int d = 0;
  if (cudaGetDevice(&d) != cudaSuccess || d < 0 || d >= MAX_DEVICES) return 0;
  return d;
}
static inline int device_sm_count() {
  static int n[MAX_DEVICES] = {};
  const int d = current_device();
  if (n[d] == 0) {
    int v = 0;
    cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, d);
    n[d] = v > 0 ? v : 148;
  }
  return n[d];
}
// `static DeviceOnce once; if (!once.get()) { ...set attributes...; once.set(); }` (idempotent work: a race repeats it)
struct DeviceOnce {
  bool done[MAX_DEVICES] = {};
  bool get() const { return done[current_device()]; }
  void set() { done[current_device()] = true; }
};
template <typename T>
struct DeviceValue {
  T v[MAX_DEVICES];
  bool init[MAX_DEVICES] = {};
  T& get(T dflt) { const int d = current_device(); if (!init[d]) { v[d] = dflt; init[d] = true; } return v[d]; }
};

// ---------------------------------------------------------------------------------------------
// warp helpers
// ---------------------------------------------------------------------------------------------
B2_DEVINL float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
B2_DEVINL float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
B2_DEVINL double warp_sum_d(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
B2_DEVINL int warp_sum_i(int v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

B2_DEVINL bool elect_one() {
  uint32_t pred;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "elect.sync _|p, 0xffffffff;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}\n"
      : "=r"(pred));
  return pred != 0;
}

// ---------------------------------------------------------------------------------------------
// 128-bit global loads/stores
// ---------------------------------------------------------------------------------------------
B2_DEVINL uint4 ld_nc_v4(const void* p) {   // streaming read-only, no L1 allocate
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}
B2_DEVINL uint4 ld_v4(const void* p) { return *reinterpret_cast<const uint4*>(p); }
B2_DEVINL uint4 ld_cg_v4(const void* p) {    // L2-coherent: data another CTA of the same launch may have written
  uint4 r;
  asm volatile("ld.global.cg.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p) : "memory");
  return r;
}
B2_DEVINL void st_v4(void* p, const uint4& v) { *reinterpret_cast<uint4*>(p) = v; }
B2_DEVINL uint2 ld_nc_v2(const void* p) {
  uint2 r;
  asm volatile("ld.global.nc.L1::no_allocate.v2.u32 {%0,%1}, [%2];" : "=r"(r.x), "=r"(r.y) : "l"(p));
  return r;
}

B2_DEVINL float2 h2_to_f2(uint32_t u) {
  __half2 h = *reinterpret_cast<__half2*>(&u);
  return __half22float2(h);
}
B2_DEVINL uint32_t f2_to_h2(float a, float b) {
  __half2 h = __floats2half2_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&h);
}

// ---------------------------------------------------------------------------------------------
// pinned fp32 arithmetic: every step individually rounded (never contracted into FMA),
// IEEE division/sqrt.  See oracle/w8a8.py for the CPU statement of the same sequences.
// ---------------------------------------------------------------------------------------------
B2_DEVINL float mul_rn(float a, float b) { return __fmul_rn(a, b); }
B2_DEVINL float add_rn(float a, float b) { return __fadd_rn(a, b); }
B2_DEVINL float sub_rn(float a, float b) { return __fsub_rn(a, b); }
B2_DEVINL float div_rn(float a, float b) { return __fdiv_rn(a, b); }

// exp(t): Cody-Waite reduction + degree-5 polynomial (oracle: exp_det).
B2_DEVINL float exp_det(float t) {
  t = fminf(fmaxf(t, -80.0f), 80.0f);
  // n = rint(t * log2 e), half to even, by the 1.5 * 2^23 trick (exact for |x| < 2^22; here |x| <= 116): two FADDs
  // and an integer subtract instead of FRND + F2I, which run on the conversion (XU) pipe -- the act*up GEMM epilogue
  // had that pipe 82 % busy (ncu, round 2).  Same values as rintf / (int) for every input.
  const float nm = add_rn(mul_rn(t, 1.4426950408889634f), 12582912.0f);
  float n = sub_rn(nm, 12582912.0f);
  float r = sub_rn(t, mul_rn(n, 0.693359375f));
  r = sub_rn(r, mul_rn(n, -2.12194440e-4f));
  float p = 1.9875691500e-4f;
  p = add_rn(mul_rn(p, r), 1.3981999507e-3f);
  p = add_rn(mul_rn(p, r), 8.3334519073e-3f);
  p = add_rn(mul_rn(p, r), 4.1665795894e-2f);
  p = add_rn(mul_rn(p, r), 1.6666665459e-1f);
  p = add_rn(mul_rn(p, r), 5.0000001201e-1f);
  float r2 = mul_rn(r, r);
  float res = add_rn(mul_rn(p, r2), r);
  res = add_rn(res, 1.0f);
  const int ni = __float_as_int(nm) - 0x4B400000;
  float two_n = __int_as_float((ni + 127) << 23);   // exact 2^n, n in [-116,116]
  return mul_rn(res, two_n);
}
B2_DEVINL float silu_det(float g) {
  float e = exp_det(-g);
  return div_rn(g, add_rn(1.0f, e));
}
B2_DEVINL float gelu_new_det(float u) {
  float u3 = mul_rn(mul_rn(u, u), u);
  float inner = add_rn(u, mul_rn(0.044715f, u3));
  float z = mul_rn(0.7978845608028654f, inner);
  float e = exp_det(mul_rn(2.0f, z));
  float th = sub_rn(1.0f, div_rn(2.0f, add_rn(e, 1.0f)));
  return mul_rn(mul_rn(0.5f, u), add_rn(1.0f, th));
}

// exact-erf GELU (HF "gelu": GPT-NeoX / Pythia): erf by Abramowitz-Stegun 7.1.26 on exp_det (oracle: gelu_erf_det)
B2_DEVINL float gelu_erf_det(float u) {
  const float z = mul_rn(u, 0.7071067811865476f);
  const float a = fabsf(z);
  const float t = div_rn(1.0f, add_rn(1.0f, mul_rn(0.3275911f, a)));
  float p = 1.061405429f;
  p = add_rn(mul_rn(p, t), -1.453152027f);
  p = add_rn(mul_rn(p, t), 1.421413741f);
  p = add_rn(mul_rn(p, t), -0.284496736f);
  p = add_rn(mul_rn(p, t), 0.254829592f);
  p = mul_rn(p, t);
  const float e = exp_det(-mul_rn(a, a));
  float er = sub_rn(1.0f, mul_rn(p, e));
  er = z < 0.0f ? -er : er;
  return mul_rn(mul_rn(0.5f, u), add_rn(1.0f, er));
}

// q = clamp(rint(y*inv), -127, 127)
B2_DEVINL int quant1(float y, float inv) {
  int q = __float2int_rn(mul_rn(y, inv));
  return max(-127, min(127, q));
}
B2_DEVINL uint32_t pack4(int a, int b, int c, int d) {
  return (uint32_t)(a & 0xff) | ((uint32_t)(b & 0xff) << 8) | ((uint32_t)(c & 0xff) << 16) |
         ((uint32_t)(d & 0xff) << 24);
}

// ---------------------------------------------------------------------------------------------
// mbarrier
// ---------------------------------------------------------------------------------------------
B2_DEVINL uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

B2_DEVINL void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
B2_DEVINL void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
B2_DEVINL void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
               : "memory");
}
B2_DEVINL void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
B2_DEVINL void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "WAIT_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra DONE_%=;\n\t"
      "bra WAIT_%=;\n\t"
      "DONE_%=:\n\t}\n" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}

// ---------------------------------------------------------------------------------------------
// TMA (cp.async.bulk.tensor), 2D tiled load global -> shared, completion on an mbarrier
// ---------------------------------------------------------------------------------------------
B2_DEVINL void tma_prefetch_desc(const void* tmap) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(tmap) : "memory");
}
B2_DEVINL void tma_load_2d(void* smem_dst, const void* tmap, uint64_t* bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(smem_u32(smem_dst)), "l"(tmap), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
      : "memory");
}
B2_DEVINL void tma_load_2d_hint(void* smem_dst, const void* tmap, uint64_t* bar, int c0, int c1,
                                uint64_t policy) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint "
      "[%0], [%1, {%3, %4}], [%2], %5;"
      ::"r"(smem_u32(smem_dst)), "l"(tmap), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "l"(policy)
      : "memory");
}
// L2 eviction-priority policies (createpolicy encodings used by CUTLASS' TMA::CacheHintSm90)
constexpr uint64_t L2_EVICT_NORMAL = 0x1000000000000000ull;
constexpr uint64_t L2_EVICT_FIRST = 0x12F0000000000000ull;
constexpr uint64_t L2_EVICT_LAST = 0x14F0000000000000ull;

// Cluster helpers.  A shared::cta address is also a valid shared::cluster address of the executing CTA;
// multicast variants deliver to the same CTA-relative offset in every CTA named by `mask`.
B2_DEVINL uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
B2_DEVINL void cluster_sync() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
B2_DEVINL void tma_load_2d_mcast(void* smem_dst, const void* tmap, uint64_t* bar, int c0, int c1, uint16_t mask) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster "
      "[%0], [%1, {%3, %4}], [%2], %5;"
      ::"r"(smem_u32(smem_dst)), "l"(tmap), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "h"(mask)
      : "memory");
}
// all previously issued MMAs of this thread arrive on the same-offset mbarrier of every CTA in `mask`
B2_DEVINL void umma_commit_mcast(uint64_t* bar, uint16_t mask) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
               ::"r"(smem_u32(bar)), "h"(mask)
               : "memory");
}

// Non-tensor bulk copy (TMA engine, no descriptor): `bytes` contiguous bytes global -> shared,
// completion counted on an mbarrier.  src/dst 16-byte aligned, bytes % 16 == 0.
B2_DEVINL void bulk_g2s(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar, uint64_t policy) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
      ::"r"(smem_u32(smem_dst)), "l"(gsrc), "r"(bytes), "r"(smem_u32(bar)), "l"(policy)
      : "memory");
}
B2_DEVINL void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
B2_DEVINL uint4 lds_v4(const void* p) {
  uint4 r;
  asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "r"(smem_u32(p)));
  return r;
}

// ---------------------------------------------------------------------------------------------
// tcgen05 / TMEM
// ---------------------------------------------------------------------------------------------
template <int COLS>
B2_DEVINL void tmem_alloc(uint32_t* smem_result) {   // one full warp, .sync.aligned
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_result)),
               "n"(COLS)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
template <int COLS>
B2_DEVINL void tmem_dealloc(uint32_t taddr) {          // same warp that allocated
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "n"(COLS) : "memory");
}
B2_DEVINL void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
B2_DEVINL void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// D[tmem] (+)= A[smem desc] * B[smem desc]^T, int8 x int8 -> int32, issued by ONE thread.
B2_DEVINL void umma_i8(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}\n" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// fp16 x fp16 -> fp32
B2_DEVINL void umma_f16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}\n" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// all previously issued MMAs of this thread arrive on the mbarrier when complete
B2_DEVINL void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
               : "memory");
}

// TMEM -> registers: this warp's 32 lanes (its quadrant), 32 consecutive 32-bit columns.
B2_DEVINL void tmem_ld_32x32(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"
      "%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
        "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
        "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
}
B2_DEVINL void tmem_ld_32x16(uint32_t taddr, uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
}
B2_DEVINL void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// Shared-memory matrix descriptor for a K-major tile whose rows are exactly one 128-byte
// swizzle atom wide (tile = rows x 128 B, written by TMA with CU_TENSOR_MAP_SWIZZLE_128B,
// base 1024-byte aligned).  8-row groups are 1024 B apart (SBO); LBO is unused for this
// layout.  Bits: [0,14) addr>>4, [16,30) LBO>>4, [32,46) SBO>>4, [46,48) version=1,
// [61,64) layout (2 = SWIZZLE_128B).
B2_DEVINL uint64_t umma_desc_sw128(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr >> 4) & 0x3FFF);
  d |= (uint64_t)1 << 16;
  d |= (uint64_t)(1024 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}

// Instruction descriptor (upper 32 bits of the CUTLASS "idescE"):
// [4,6) D format (1=f32, 2=s32); [7,10) A format; [10,13) B format; bit15/16 A/B major
// (0 = K-major); [17,23) N>>3; [24,29) M>>4.
constexpr uint32_t umma_idesc(uint32_t dfmt, uint32_t afmt, uint32_t bfmt, uint32_t M, uint32_t N) {
  return (dfmt << 4) | (afmt << 7) | (bfmt << 10) | ((N >> 3) << 17) | ((M >> 4) << 24);
}
constexpr uint32_t IDESC_FMT_S32 = 2, IDESC_FMT_F32 = 1, IDESC_S8 = 1, IDESC_F16 = 0;

// ---------------------------------------------------------------------------------------------
// Timeline probe (tools/timeline.py; `make timeline` builds libb200stage_tl.so with -DB2_TIMELINE).
// Thread 0 of every CTA stamps %globaltimer at named points of a kernel into a per-translation-unit
// device array; the product build compiles tl_mark() to nothing.
// ---------------------------------------------------------------------------------------------
#ifdef B2_TIMELINE
constexpr unsigned TL_CAP = 1u << 20;
struct TlEvent { unsigned long long t; unsigned tag; unsigned cta; };
static __device__ TlEvent g_tl[TL_CAP];
static __device__ unsigned g_tl_n;
// A mark only stores the timer into shared memory; the CTA's marks go to global memory together at
// tl_flush() (one atomic per CTA, after the last point), so the probe adds no memory round trip
// between the points it measures.
constexpr unsigned TL_POINTS = 8;
B2_DEVINL unsigned long long* tl_slots() {
  __shared__ unsigned long long s_tl[TL_POINTS];
  return s_tl;
}
B2_DEVINL void tl_mark(unsigned kernel_id, unsigned point) {
  if (threadIdx.x == 0) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    tl_slots()[point] = t;
  }
}
B2_DEVINL void tl_flush(unsigned kernel_id, unsigned point_mask) {
  if (threadIdx.x == 0) {
    const unsigned n = __popc(point_mask);
    unsigned i = atomicAdd(&g_tl_n, n);
    unsigned smid;
    asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
    const unsigned cta = (smid << 20) | (blockIdx.x + blockIdx.y * gridDim.x + blockIdx.z * gridDim.x * gridDim.y);
    for (unsigned pt = 0; pt < TL_POINTS; ++pt) {
      if (!((point_mask >> pt) & 1u)) continue;
      if (i < TL_CAP) { g_tl[i].t = tl_slots()[pt]; g_tl[i].tag = (kernel_id << 8) | pt; g_tl[i].cta = cta; }
      ++i;
    }
  }
}
// copies the events to `host` (cap events), resets the counter, returns the number recorded
#define B2_TL_DEFINE_DUMP(fn)                                                              \
  extern "C" int fn(void* host, int cap) {                                                 \
    unsigned n = 0;                                                                        \
    cudaDeviceSynchronize();                                                               \
    cudaMemcpyFromSymbol(&n, b2::g_tl_n, sizeof(n));                                       \
    if (n > b2::TL_CAP) n = b2::TL_CAP;                                                    \
    if ((int)n > cap) n = cap;                                                             \
    if (host && n) cudaMemcpyFromSymbol(host, b2::g_tl, (size_t)n * sizeof(b2::TlEvent));  \
    const unsigned z = 0;                                                                  \
    cudaMemcpyToSymbol(b2::g_tl_n, &z, sizeof(z));                                         \
    return (int)n;                                                                         \
  }
#else
#define tl_mark(kernel_id, point) ((void)0)
#define tl_flush(kernel_id, point_mask) ((void)0)
#define B2_TL_DEFINE_DUMP(fn)
#endif

}  // namespace b2
