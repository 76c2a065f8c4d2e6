// Stream-ordered barrier between the GPUs of one sequence-parallel group: the two Ulysses exchanges
// of every block are remote stores issued by the producer kernels, so what orders them against the consumer is a barrier
// on the stream, 120 times per denoising step at c3.  One CTA, one thread per peer: thread p publishes this rank's epoch
// into slot [rank] of peer p's flag array (st.release.sys over NVLink) and then spins on slot [p] of its own array until
// peer p's epoch arrives (ld.acquire.sys).  Epochs only grow (the host counts the calls), so no slot is ever reset and
// there is no second round trip.  Kernels enqueued after the barrier start only when every peer's stream has reached its
// own call, i.e. after the peers' earlier kernels — and the remote stores they made — have completed.
// Replaces xfuser's all_to_all synchronisation inside usp_attn_forward (not in the reference tree).
#include "common.cuh"

namespace {

struct FlagPtrs {
  unsigned int* p[8];
};

// kPhase: 3 = barrier (publish, then wait); 1 = publish only; 2 = wait only.  The split form lets the host put independent
// work between "my remote stores are done" and "I need the peers' stores" (the interleaved two-chain schedule of
// transformer._run_blocks_interleaved): a peer that is late by less than that work costs nothing.
template <int kPhase>
__global__ void __launch_bounds__(32, 1) peer_barrier_kernel(FlagPtrs flags, int rank, int n_peers, unsigned int epoch,
                                                             unsigned long long timeout_ns, unsigned int* status) {
  const int p = threadIdx.x;
  if (p >= n_peers) return;
  // an earlier barrier of this group timed out: the step is lost; publish the epoch (peers may still be waiting) but do not
  // spend another timeout in each of the remaining barriers of the forward
  const bool failed = (kPhase & 2) ? *reinterpret_cast<volatile unsigned int*>(status) != 0u : false;
  __threadfence_system();
  if ((kPhase & 1) && p != rank) {
    unsigned int* remote = flags.p[p] + rank;
    asm volatile("st.release.sys.global.u32 [%0], %1;\n" ::"l"(remote), "r"(epoch) : "memory");
  }
  if ((kPhase & 2) && p != rank) {
    const unsigned int* mine = flags.p[rank] + p;
    unsigned long long t0;
    asm volatile("mov.u64 %0, %%globaltimer;\n" : "=l"(t0));
    unsigned int seen;
    int spins = 0;
    while (true) {
      asm volatile("ld.acquire.sys.global.u32 %0, [%1];\n" : "=r"(seen) : "l"(mine) : "memory");
      if ((int)(seen - epoch) >= 0 || failed) break;  // wrap-safe: the peer is at this epoch or beyond
      if ((++spins & 1023) == 0) {
        unsigned long long t1;
        asm volatile("mov.u64 %0, %%globaltimer;\n" : "=l"(t1));
        if (t1 - t0 > timeout_ns) {  // a peer died or the ranks disagree on the call count: report, do not hang the GPU
          atomicExch(status, 0x80000000u | (unsigned int)p);
          break;
        }
      }
    }
  }
  __threadfence_system();
}

}  // namespace

// flags[i]: device address of rank i's flag array (>= n_peers u32, zero-initialised, mapped into this process); epoch: the
// 1-based count of barrier calls on this group (same on every rank); status: one u32 in this GPU's memory that receives
// 0x80000000 | peer on a timeout (timeout_ms; 0 = 5000 ms).
extern "C" int vc_peer_barrier(void* const* flags, int rank, int n_peers, unsigned int epoch, int timeout_ms, void* status,
                               void* stream) {
  VC_CHECK_ARG(flags && status, "vc_peer_barrier: null pointer");
  VC_CHECK_ARG(n_peers >= 1 && n_peers <= 8 && rank >= 0 && rank < n_peers, "vc_peer_barrier: bad rank %d of %d", rank, n_peers);
  FlagPtrs fp;
  for (int i = 0; i < 8; ++i) fp.p[i] = (unsigned int*)(i < n_peers ? flags[i] : nullptr);
  for (int i = 0; i < n_peers; ++i) VC_CHECK_ARG(fp.p[i] && (uintptr_t)fp.p[i] % 4 == 0, "vc_peer_barrier: bad flag pointer %d", i);
  const unsigned long long ns = (unsigned long long)(timeout_ms > 0 ? timeout_ms : 5000) * 1000000ull;
  peer_barrier_kernel<3><<<1, 32, 0, (cudaStream_t)stream>>>(fp, rank, n_peers, epoch, ns, (unsigned int*)status);
  VC_RETURN_LAUNCH_STATUS("vc_peer_barrier");
}

// Split form of vc_peer_barrier.  vc_peer_signal publishes `epoch` to every peer (everything enqueued before it on this
// stream, remote stores included, has completed by then); vc_peer_wait blocks the stream until every peer has published an
// epoch >= `epoch`.  Epochs come from ONE counter per group that every rank advances in the same order, so a flag slot
// only grows and several signals may be outstanding.  vc_peer_barrier(e) == vc_peer_signal(e); vc_peer_wait(e).
extern "C" int vc_peer_signal(void* const* flags, int rank, int n_peers, unsigned int epoch, void* stream) {
  VC_CHECK_ARG(flags, "vc_peer_signal: null pointer");
  VC_CHECK_ARG(n_peers >= 1 && n_peers <= 8 && rank >= 0 && rank < n_peers, "vc_peer_signal: bad rank %d of %d", rank, n_peers);
  FlagPtrs fp;
  for (int i = 0; i < 8; ++i) fp.p[i] = (unsigned int*)(i < n_peers ? flags[i] : nullptr);
  for (int i = 0; i < n_peers; ++i) VC_CHECK_ARG(fp.p[i] && (uintptr_t)fp.p[i] % 4 == 0, "vc_peer_signal: bad flag pointer %d", i);
  peer_barrier_kernel<1><<<1, 32, 0, (cudaStream_t)stream>>>(fp, rank, n_peers, epoch, 0ull, nullptr);
  VC_RETURN_LAUNCH_STATUS("vc_peer_signal");
}

extern "C" int vc_peer_wait(void* const* flags, int rank, int n_peers, unsigned int epoch, int timeout_ms, void* status,
                            void* stream) {
  VC_CHECK_ARG(flags && status, "vc_peer_wait: null pointer");
  VC_CHECK_ARG(n_peers >= 1 && n_peers <= 8 && rank >= 0 && rank < n_peers, "vc_peer_wait: bad rank %d of %d", rank, n_peers);
  FlagPtrs fp;
  for (int i = 0; i < 8; ++i) fp.p[i] = (unsigned int*)(i < n_peers ? flags[i] : nullptr);
  for (int i = 0; i < n_peers; ++i) VC_CHECK_ARG(fp.p[i] && (uintptr_t)fp.p[i] % 4 == 0, "vc_peer_wait: bad flag pointer %d", i);
  const unsigned long long ns = (unsigned long long)(timeout_ms > 0 ? timeout_ms : 5000) * 1000000ull;
  peer_barrier_kernel<2><<<1, 32, 0, (cudaStream_t)stream>>>(fp, rank, n_peers, epoch, ns, (unsigned int*)status);
  VC_RETURN_LAUNCH_STATUS("vc_peer_wait");
}
