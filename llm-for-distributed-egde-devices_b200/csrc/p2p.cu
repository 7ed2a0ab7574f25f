// Stage-to-stage hand-off of a decode step over NVLink peer memory, without NCCL and without the host:
// the sending stage's kernel STORES the hidden-state tile (or, on the way back, the final hidden state and the
// upper vocabulary shard's winner) straight into the receiving GPU's inbox and then publishes a sequence number
// next to it; the receiving stage's kernel spins on that number in its OWN memory.
//
// Why: with N prompts on N stages the decode ring is latency-bound -- a token visits every stage before the next
// step of its sequence can start -- and a torch.distributed send/recv pair cost ~58 us per hop at 8 stages
// (host-posted receive kernel, side streams, events) against ~230 us of compute per stage.  A peer store + flag
// is one ~2 us kernel at each end, enqueued back to back with the stage's graph.
//
// Replaces: the gRPC hop between two devices that the reference sketches (Code/gRPC/server.py:13-19,
// client.py:7-11) for the hidden-state hand-off north_star moves onto NVLink.
//
// Plain ld/st only: no TMA here (TMA multicast loads here read only this GPU's memory, since on B200 a TMA
// multicast load from another GPU's memory hung the GPU).
#include "common.cuh"

namespace b2 {

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
  return (cudaFuncGetAttributes(&fa, p2p_send_kernel) == cudaSuccess && cudaFuncGetAttributes(&fa, p2p_wait_kernel) == cudaSuccess)
             ? OK : ERR_ARCH;
}

}  // namespace b2
