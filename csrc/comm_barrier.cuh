// Cross-rank flag barrier over the symmetric arenas' signal pads, shared by the NVLink kernels
// (comm_allreduce.cu, comm_adamw.cu).
//
// CTA b of every rank meets CTA b of every other rank: "put" = CAS 0→1 (release.sys) on the peer's pad,
// "wait" = CAS 1→0 (acquire.sys) on ours; every flag returns to 0, so the same words serve every barrier of
// every launch.  One signal slot holds kSignalBlocks x kMaxRanks 32-bit words; a kernel may use at most
// kSignalBlocks CTAs, and the host side (parallel/symm.py) checks the pad against b200_comm_signal_words().
//
// A spin that never ends takes the GPU with it (and, in a multi-rank job, every other rank's), so both spins are
// bounded: after `timeout_ns` (host-set, default 30 s — it only has to outlast the slowest rank's start-up skew)
// the CTA records what it was waiting for in a per-device status block, marks the communicator dead and goes on;
// every later barrier on that device then returns at once and the kernels drain.  The data is garbage from that
// point — the host reads the status block (b200_comm_status) and raises.  That turns "the job hung until the
// scheduler killed it" into an error message naming rank, CTA, peer and phase.
#pragma once
#include "common.cuh"

namespace b200 {
namespace comm {

constexpr int kMaxRanks = 8;
constexpr int kSignalBlocks = 148;  // the largest grid of any kernel that uses the barrier (one CTA per SM)

// status block: [0] dead flag, [1] rank, [2] blockIdx, [3] peer, [4] phase (1 = put, 2 = wait), [5] slot, [6] flag value
struct CommStatus {
  unsigned int dead;
  int rank, block, peer, phase, slot;
  unsigned int flag;
  int pad_;
};

// One copy per translation unit (no relocatable device code in this build); the host getter of each TU is merged
// by b200_comm_status().
static __device__ CommStatus g_comm_status;
static __device__ unsigned long long g_comm_timeout_ns = 30ull * 1000ull * 1000ull * 1000ull;

__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ unsigned cas_release_sys(unsigned* addr, unsigned cmp, unsigned val) {
  unsigned old;
  asm volatile("atom.global.release.sys.cas.b32 %0, [%1], %2, %3;" : "=r"(old) : "l"(addr), "r"(cmp), "r"(val) : "memory");
  return old;
}
__device__ __forceinline__ unsigned cas_acquire_sys(unsigned* addr, unsigned cmp, unsigned val) {
  unsigned old;
  asm volatile("atom.global.acquire.sys.cas.b32 %0, [%1], %2, %3;" : "=r"(old) : "l"(addr), "r"(cmp), "r"(val) : "memory");
  return old;
}

static __device__ __noinline__ void comm_report_timeout(int rank, int peer, int phase, int slot, unsigned flag) {
  if (atomicCAS(&g_comm_status.dead, 0u, 1u) == 0u) {
    g_comm_status.rank = rank;
    g_comm_status.block = blockIdx.x;
    g_comm_status.peer = peer;
    g_comm_status.phase = phase;
    g_comm_status.slot = slot;
    g_comm_status.flag = flag;
    __threadfence_system();
  }
}

// sig[r] = base of rank r's signal pad (peer-mapped).  Must be called by every thread of the CTA.
__device__ __forceinline__ void cross_rank_barrier(unsigned int* const* sig, int rank, int ws, int slot) {
  __syncthreads();
  if (threadIdx.x < ws) {
    const int peer = threadIdx.x;
    unsigned* send = sig[peer] + slot + blockIdx.x * kMaxRanks + rank;
    unsigned* recv = sig[rank] + slot + blockIdx.x * kMaxRanks + peer;
    if (*reinterpret_cast<volatile unsigned int*>(&g_comm_status.dead) == 0u) {
      unsigned long long t0 = 0;
      unsigned spins = 0;
      unsigned old;
      bool ok = true;
      while ((old = cas_release_sys(send, 0u, 1u)) != 0u) {
        if ((++spins & 1023u) == 0u) {
          const unsigned long long now = globaltimer_ns();
          if (t0 == 0) t0 = now;
          if (now - t0 > g_comm_timeout_ns) { comm_report_timeout(rank, peer, 1, slot, old); ok = false; break; }
        }
      }
      spins = 0;
      t0 = 0;
      while (ok && (old = cas_acquire_sys(recv, 1u, 0u)) != 1u) {
        if ((++spins & 1023u) == 0u) {
          if (*reinterpret_cast<volatile unsigned int*>(&g_comm_status.dead) != 0u) break;
          const unsigned long long now = globaltimer_ns();
          if (t0 == 0) t0 = now;
          if (now - t0 > g_comm_timeout_ns) { comm_report_timeout(rank, peer, 2, slot, old); break; }
        }
      }
    }
  }
  __syncthreads();
}

// host side, per translation unit
static inline int comm_status_read(CommStatus* out) {
  return cudaMemcpyFromSymbol(out, g_comm_status, sizeof(CommStatus)) == cudaSuccess ? 0 : 1;
}
static inline int comm_status_reset_and_set_timeout(unsigned long long timeout_ns) {
  CommStatus z = {};
  if (cudaMemcpyToSymbol(g_comm_status, &z, sizeof(z)) != cudaSuccess) return 1;
  if (timeout_ns && cudaMemcpyToSymbol(g_comm_timeout_ns, &timeout_ns, sizeof(timeout_ns)) != cudaSuccess) return 1;
  return 0;
}

}  // namespace comm
}  // namespace b200
