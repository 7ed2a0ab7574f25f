// Stage-to-stage hand-off of a decode step over NVLink peer memory, without NCCL and without the host:
// the sending stage's kernel STORES the hidden-state tile (or, on the way back, the final hidden state and the
// upper vocabulary shard's winner) straight into the receiving GPU's inbox and then publishes a sequence number
// next to it; the receiving stage's kernel spins on that number in its OWN memory.
//
// Why: with N prompts on N stages the decode ring is latency-bound -- a token visits every stage before the next
// step of its sequence can start -- and a torch.distributed send/recv pair cost ~58 us per hop at 8 stages
// (host-posted receive kernel, side streams, events) against ~230 us of compute per stage.  A peer store + flag
// is one small kernel at each end, enqueued back to back with the stage's graph (a whole hop, store + flag ->
// wait + fetch, measures <= 9 us on the device: tools/hop_probe.py).
//
// Replaces: the gRPC hop between two devices that the reference sketches (Code/gRPC/server.py:13-19,
// client.py:7-11) for the hidden-state hand-off north_star moves onto NVLink.
//
// Plain ld/st only: no TMA here (TMA multicast loads here read only this GPU's memory, since on B200 a TMA
// multicast load from another GPU's memory hung the GPU).
#include "common.cuh"

namespace b2 {

constexpr int P2P_MAX_DST = 16;

// dst / flag live in the PEER's memory (mapped into this process); src is local.  One CTA: every thread stores its
// 16-byte vectors and fences them system-wide, then one thread releases the sequence number.
__global__ void __launch_bounds__(1024) p2p_send_kernel(const uint4* __restrict__ src, uint4* __restrict__ dst, int n16,
                                                        unsigned* flag, unsigned seq) {
  for (int i = threadIdx.x; i < n16; i += blockDim.x) dst[i] = src[i];
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(flag), "r"(seq) : "memory");
}

// flag / src are LOCAL (the inbox of this GPU).  Thread 0 waits until the sequence number has reached `seq`
// (wrap-safe compare), then the CTA copies the message to `dst` (may be null: the consumer reads the inbox in
// place).  The wait is bounded -- about 30 s of polling -- so that a lost peer ends as an error flag the host can
// read, not as a hung GPU.
__global__ void __launch_bounds__(256) p2p_wait_kernel(const unsigned* flag, unsigned seq, const uint4* __restrict__ src,
                                                       uint4* __restrict__ dst, int n16, unsigned* err) {
  __shared__ int ok;
  if (threadIdx.x == 0) {
    unsigned v = 0;
    int good = 0;
    for (unsigned spins = 0; spins < (1u << 28); ++spins) {
      asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(flag) : "memory");
      if ((int)(v - seq) >= 0) { good = 1; break; }
      if (spins > 64) __nanosleep(100);
    }
    if (!good && err) atomicExch(err, 1u);
    ok = good;
  }
  __syncthreads();
  if (dst != nullptr && ok)
    for (int i = threadIdx.x; i < n16; i += blockDim.x) dst[i] = src[i];
}

// The same message to several GPUs at once (the last stage's normed hidden state to every stage that holds a slice
// of the lm_head): CTA b stores to destination b and publishes its flag.
struct MultiDst { uint4* dst[P2P_MAX_DST]; unsigned* flag[P2P_MAX_DST]; };
__global__ void __launch_bounds__(256) p2p_send_multi_kernel(const uint4* __restrict__ src, int n16, MultiDst md, unsigned seq) {
  uint4* dst = md.dst[blockIdx.x];
  for (int i = threadIdx.x; i < n16; i += blockDim.x) dst[i] = src[i];
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(md.flag[blockIdx.x]), "r"(seq) : "memory");
}

// Greedy token from the winners of W vocabulary slices (one record per slice in this GPU's inbox: indices [4] int64 at
// +0, values [4] fp32 at +32, 64 bytes apart; one flag per slice): lane r waits for slice r's sequence number, then the
// warp picks, per row, the largest value -- lowest global index among equals, as b2_argmax does over the whole row.
__global__ void __launch_bounds__(32) merge_winners_kernel(const unsigned* flags, unsigned seq, const uint8_t* records, int W, int B,
                                                           long long* out_idx, long long* out_idx2, unsigned* err) {
  const int r = threadIdx.x;
  bool good = true;
  if (r < W) {
    unsigned v = 0;
    good = false;
    for (unsigned spins = 0; spins < (1u << 28); ++spins) {
      asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(flags + r) : "memory");
      if ((int)(v - seq) >= 0) { good = true; break; }
      if (spins > 64) __nanosleep(100);
    }
    if (!good && err) atomicExch(err, 1u);
  }
  __syncwarp();
  for (int m = 0; m < B; ++m) {
    float best = -INFINITY;
    long long bi = 0x7fffffffffffffffll;
    if (r < W) {
      best = __ldcg(reinterpret_cast<const float*>(records + (size_t)r * 64 + 32) + m);
      bi = __ldcg(reinterpret_cast<const long long*>(records + (size_t)r * 64) + m);
      if (best != best) { best = -INFINITY; bi = 0x7fffffffffffffffll; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float ov = __shfl_xor_sync(0xffffffffu, best, o);
      const long long oi = __shfl_xor_sync(0xffffffffu, bi, o);
      if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
    }
    if (r == 0) {
      out_idx[m] = bi;
      if (out_idx2 != nullptr) out_idx2[m] = bi;
    }
  }
}

int p2p_send_multi(const void* src, long long bytes, void* const* dsts, unsigned* const* flags, int n, unsigned seq, cudaStream_t st) {
  if (!src || !dsts || !flags) return ERR_NULL_PTR;
  if (n <= 0 || n > P2P_MAX_DST || bytes <= 0 || (bytes & 15) || bytes > (1ll << 30)) return ERR_BAD_SHAPE;
  MultiDst md{};
  for (int i = 0; i < n; ++i) {
    if (!dsts[i] || !flags[i]) return ERR_NULL_PTR;
    if (reinterpret_cast<uintptr_t>(dsts[i]) & 15) return ERR_BAD_ALIGN;
    md.dst[i] = reinterpret_cast<uint4*>(dsts[i]);
    md.flag[i] = flags[i];
  }
  if (reinterpret_cast<uintptr_t>(src) & 15) return ERR_BAD_ALIGN;
  p2p_send_multi_kernel<<<n, 256, 0, st>>>((const uint4*)src, (int)(bytes >> 4), md, seq);
  return launch_status();
}

int merge_winners(const unsigned* flags, unsigned seq, const void* records, int W, int B, long long* out_idx, long long* out_idx2,
                  unsigned* err, cudaStream_t st) {
  if (!flags || !records || !out_idx) return ERR_NULL_PTR;
  if (W <= 0 || W > 32 || B <= 0 || B > 4) return ERR_BAD_SHAPE;
  if (reinterpret_cast<uintptr_t>(records) & 15) return ERR_BAD_ALIGN;
  merge_winners_kernel<<<1, 32, 0, st>>>(flags, seq, (const uint8_t*)records, W, B, out_idx, out_idx2, err);
  return launch_status();
}

int p2p_send(const void* src, void* dst_peer, long long bytes, unsigned* flag_peer, unsigned seq, cudaStream_t st) {
  if (!src || !dst_peer || !flag_peer) return ERR_NULL_PTR;
  if (bytes <= 0 || (bytes & 15) || bytes > (1ll << 30)) return ERR_BAD_SHAPE;
  if ((reinterpret_cast<uintptr_t>(src) & 15) || (reinterpret_cast<uintptr_t>(dst_peer) & 15)) return ERR_BAD_ALIGN;
  const int n16 = (int)(bytes >> 4);
  p2p_send_kernel<<<1, n16 >= 1024 ? 1024 : 256, 0, st>>>((const uint4*)src, (uint4*)dst_peer, n16, flag_peer, seq);
  return launch_status();
}

int p2p_wait(const unsigned* flag, unsigned seq, const void* src, void* dst, long long bytes, unsigned* err, cudaStream_t st) {
  if (!flag) return ERR_NULL_PTR;
  if (dst && (!src || bytes <= 0 || (bytes & 15))) return ERR_BAD_SHAPE;
  if (dst && ((reinterpret_cast<uintptr_t>(src) & 15) || (reinterpret_cast<uintptr_t>(dst) & 15))) return ERR_BAD_ALIGN;
  p2p_wait_kernel<<<1, 256, 0, st>>>(flag, seq, (const uint4*)src, (uint4*)dst, dst ? (int)(bytes >> 4) : 0, err);
  return launch_status();
}

int init_p2p() {
  cudaFuncAttributes fa;
  return (cudaFuncGetAttributes(&fa, p2p_send_kernel) == cudaSuccess && cudaFuncGetAttributes(&fa, p2p_wait_kernel) == cudaSuccess &&
          cudaFuncGetAttributes(&fa, p2p_send_multi_kernel) == cudaSuccess && cudaFuncGetAttributes(&fa, merge_winners_kernel) == cudaSuccess)
             ? OK : ERR_ARCH;
}

}  // namespace b2
