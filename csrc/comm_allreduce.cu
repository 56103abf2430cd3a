// Gradient all-reduce over NVLink 5 / NVSwitch peer memory, fused with the 1/world_size average.
//
// Replaces dist.all_reduce(...) + grad/world_size of the reference's DDP classes
// (cs336_systems/ddp_bucketed_overlapped_sharded.py:236,247,299,314) and the NCCL ring behind them.
// Every rank passes the peer-mapped base pointers of a symmetric arena (set up
// through torch.distributed._symmetric_memory) — as a small host array that travels to the kernel by value; the kernels load and store peer memory directly.
//
//   one-shot  : every rank loads all ranks' copies of the whole message, reduces in rank order in
//               registers, cross-rank barrier, writes its own copy.  One NVLink traversal; for small
//               messages (latency bound).
//   two-shot  : rank r reduces slice r (peer loads) and stores the averaged slice into every rank's
//               buffer (peer stores).  2·(ws-1)/ws·S bytes per GPU each way.
//   multimem  : slice r is reduced INSIDE the switch with multimem.ld_reduce on the NVLS multicast
//               address and broadcast with multimem.st — ≈S/ws loads + S/ws stores per GPU.
//
// Cross-rank ordering: CTA b of every rank meets CTA b of every other rank through flag words in the
// signal pads (put = CAS 0→1 release.sys on the peer's pad, wait = CAS 1→0 acquire.sys on ours), once
// before the first peer access and once after the last; no host involvement.  The reduction order is
// fixed (rank 0..ws-1, or the switch's), and each element is produced by exactly one rank in the
// two-shot / multimem variants, so all replicas end bit-identical.
#include "common.cuh"
#include "comm_barrier.cuh"

namespace b200 {
namespace comm {

constexpr int kThreads = 512;
constexpr int kMaxBlocks = 64;
static_assert(kMaxBlocks <= kSignalBlocks, "a signal slot holds kSignalBlocks CTAs' flag words");

struct Peers {
  void* ptr[kMaxRanks];
  unsigned int* sig[kMaxRanks];
};

// CTA-to-CTA barrier across ranks (same blockIdx on every rank), bounded spin: see comm_barrier.cuh.
__device__ __forceinline__ void cross_rank_barrier(const Peers& P, int rank, int ws, int slot) {
  cross_rank_barrier(P.sig, rank, ws, slot);
}

__device__ __forceinline__ f32x4 ld_peer(const void* p) {
  f32x4 v;
  asm volatile("ld.global.relaxed.sys.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_peer(void* p, const f32x4& v) {
  asm volatile("st.global.relaxed.sys.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

// 16-byte vector arithmetic for the two element types
template <bool BF16> struct Vec;
template <> struct Vec<false> {
  static __device__ __forceinline__ void add(float (&acc)[8], const f32x4& v) { acc[0] += v.x; acc[1] += v.y; acc[2] += v.z; acc[3] += v.w; }
  static __device__ __forceinline__ f32x4 finish(const float (&acc)[8], float s) { return f32x4{acc[0] * s, acc[1] * s, acc[2] * s, acc[3] * s}; }
};
template <> struct Vec<true> {
  static __device__ __forceinline__ void add(float (&acc)[8], const f32x4& v) {
    const uint32_t* u = reinterpret_cast<const uint32_t*>(&v);
#pragma unroll
    for (int k = 0; k < 4; ++k) { const float2 f = unpack_bf16x2(u[k]); acc[2 * k] += f.x; acc[2 * k + 1] += f.y; }
  }
  static __device__ __forceinline__ f32x4 finish(const float (&acc)[8], float s) {
    f32x4 o;
    uint32_t* u = reinterpret_cast<uint32_t*>(&o);
#pragma unroll
    for (int k = 0; k < 4; ++k) u[k] = pack_bf16x2(acc[2 * k] * s, acc[2 * k + 1] * s);
    return o;
  }
};

// ---- one-shot --------------------------------------------------------------------------------
constexpr int kOneShotVecs = 8;  // 16-byte vectors held per thread across the barrier

template <bool BF16>
__global__ void __launch_bounds__(kThreads) allreduce_oneshot_kernel(Peers P, long long byte_off, long long nvec,
                                                                     float scale, int rank, int ws, int slot) {
  cross_rank_barrier(P, rank, ws, slot);
  f32x4 res[kOneShotVecs];
  const long long stride = (long long)gridDim.x * kThreads;
  const long long i0 = (long long)blockIdx.x * kThreads + threadIdx.x;
#pragma unroll
  for (int k = 0; k < kOneShotVecs; ++k) {
    const long long i = i0 + k * stride;
    if (i < nvec) {
      float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
      for (int r = 0; r < ws; ++r) Vec<BF16>::add(acc, ld_peer(static_cast<const char*>(P.ptr[r]) + byte_off + i * 16));
      res[k] = Vec<BF16>::finish(acc, scale);
    }
  }
  cross_rank_barrier(P, rank, ws, slot);  // everyone has read every copy: safe to overwrite ours
  char* mine = static_cast<char*>(P.ptr[rank]) + byte_off;
#pragma unroll
  for (int k = 0; k < kOneShotVecs; ++k) {
    const long long i = i0 + k * stride;
    if (i < nvec) *reinterpret_cast<f32x4*>(mine + i * 16) = res[k];
  }
}

// ---- 256-bit helpers (bandwidth paths) ---------------------------------------------------------
struct v8f { float f[8]; };
__device__ __forceinline__ v8f ld_peer8(const void* p) {
  v8f r;
  asm volatile("ld.global.relaxed.sys.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=f"(r.f[0]), "=f"(r.f[1]), "=f"(r.f[2]), "=f"(r.f[3]), "=f"(r.f[4]), "=f"(r.f[5]), "=f"(r.f[6]), "=f"(r.f[7])
               : "l"(p) : "memory");
  return r;
}
__device__ __forceinline__ void st_peer8(void* p, const v8f& r) {
  asm volatile("st.global.relaxed.sys.v8.f32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p), "f"(r.f[0]), "f"(r.f[1]),
               "f"(r.f[2]), "f"(r.f[3]), "f"(r.f[4]), "f"(r.f[5]), "f"(r.f[6]), "f"(r.f[7]) : "memory");
}
template <bool BF16> struct Vec8;
template <> struct Vec8<false> {
  static constexpr int N = 8;
  static __device__ __forceinline__ void add(float (&acc)[16], const v8f& v) {
#pragma unroll
    for (int k = 0; k < 8; ++k) acc[k] += v.f[k];
  }
  static __device__ __forceinline__ v8f finish(const float (&acc)[16], float s) {
    v8f o;
#pragma unroll
    for (int k = 0; k < 8; ++k) o.f[k] = acc[k] * s;
    return o;
  }
};
template <> struct Vec8<true> {
  static __device__ __forceinline__ void add(float (&acc)[16], const v8f& v) {
#pragma unroll
    for (int k = 0; k < 8; ++k) { const float2 f = unpack_bf16x2(__float_as_uint(v.f[k])); acc[2 * k] += f.x; acc[2 * k + 1] += f.y; }
  }
  static __device__ __forceinline__ v8f finish(const float (&acc)[16], float s) {
    v8f o;
#pragma unroll
    for (int k = 0; k < 8; ++k) o.f[k] = __uint_as_float(pack_bf16x2(acc[2 * k] * s, acc[2 * k + 1] * s));
    return o;
  }
};

// ---- two-shot --------------------------------------------------------------------------------
// 32-byte vectors, U vectors per thread per trip with all WS*U peer loads issued before the first add:
// NVLink round trips are ~2 us, so bandwidth is bought with bytes in flight (grid x 512 threads x U x 32 B per peer).
template <bool BF16, int WS>
__global__ void __launch_bounds__(kThreads) allreduce_twoshot_kernel(Peers P, long long byte_off, long long nvec32,
                                                                     float scale, int rank, int ws, int slot) {
  constexpr int U = WS <= 2 ? 4 : (WS <= 4 ? 2 : 1);
  cross_rank_barrier(P, rank, ws, slot);
  const long long per = (nvec32 + ws - 1) / ws;
  const long long lo = per * rank, hi = (lo + per < nvec32) ? lo + per : nvec32;
  const long long stride = (long long)gridDim.x * kThreads;
  for (long long i0 = lo + (long long)blockIdx.x * kThreads + threadIdx.x; i0 < hi; i0 += stride * U) {
    v8f v[WS][U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long i = i0 + u * stride;
      if (i < hi) {
#pragma unroll
        for (int r = 0; r < WS; ++r)
          if (r < ws) v[r][u] = ld_peer8(static_cast<const char*>(P.ptr[r]) + byte_off + i * 32);
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long i = i0 + u * stride;
      if (i < hi) {
        float acc[16];
#pragma unroll
        for (int k = 0; k < 16; ++k) acc[k] = 0.f;
#pragma unroll
        for (int r = 0; r < WS; ++r)
          if (r < ws) Vec8<BF16>::add(acc, v[r][u]);
        const v8f out = Vec8<BF16>::finish(acc, scale);
#pragma unroll
        for (int r = 0; r < WS; ++r)
          if (r < ws) st_peer8(static_cast<char*>(P.ptr[r]) + byte_off + i * 32, out);
      }
    }
  }
  cross_rank_barrier(P, rank, ws, slot);  // all slices have landed everywhere
}

// ---- NVLS multimem -----------------------------------------------------------------------------
template <bool BF16>
__global__ void __launch_bounds__(kThreads) allreduce_multimem_kernel(Peers P, char* mc, long long byte_off,
                                                                      long long nvec, float scale, int rank, int ws,
                                                                      int slot) {
  constexpr int U = 4;
  cross_rank_barrier(P, rank, ws, slot);
  const long long per = (nvec + ws - 1) / ws;
  const long long lo = per * rank, hi = (lo + per < nvec) ? lo + per : nvec;
  const long long stride = (long long)gridDim.x * kThreads;
  for (long long i0 = lo + (long long)blockIdx.x * kThreads + threadIdx.x; i0 < hi; i0 += stride * U) {
    uint32_t u[U][4];
#pragma unroll
    for (int q = 0; q < U; ++q) {
      const long long i = i0 + q * stride;
      if (i < hi) {
        char* a = mc + byte_off + i * 16;
        if constexpr (!BF16)
          asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                       : "=r"(u[q][0]), "=r"(u[q][1]), "=r"(u[q][2]), "=r"(u[q][3]) : "l"(a) : "memory");
        else
          asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0,%1,%2,%3}, [%4];"
                       : "=r"(u[q][0]), "=r"(u[q][1]), "=r"(u[q][2]), "=r"(u[q][3]) : "l"(a) : "memory");
      }
    }
#pragma unroll
    for (int q = 0; q < U; ++q) {
      const long long i = i0 + q * stride;
      if (i < hi) {
        char* a = mc + byte_off + i * 16;
        if constexpr (!BF16) {
#pragma unroll
          for (int k = 0; k < 4; ++k) u[q][k] = __float_as_uint(__uint_as_float(u[q][k]) * scale);
          asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(a), "r"(u[q][0]), "r"(u[q][1]), "r"(u[q][2]), "r"(u[q][3]) : "memory");
        } else {
#pragma unroll
          for (int k = 0; k < 4; ++k) { const float2 f = unpack_bf16x2(u[q][k]); u[q][k] = pack_bf16x2(f.x * scale, f.y * scale); }
          asm volatile("multimem.st.relaxed.sys.global.v4.bf16x2 [%0], {%1,%2,%3,%4};" ::"l"(a), "r"(u[q][0]), "r"(u[q][1]), "r"(u[q][2]), "r"(u[q][3]) : "memory");
        }
      }
    }
  }
  cross_rank_barrier(P, rank, ws, slot);
}

__global__ void barrier_kernel(Peers P, int rank, int ws, int slot) { cross_rank_barrier(P, rank, ws, slot); }

// Broadcast by pull: every non-source rank copies [byte_off, byte_off + nvec32*32) of the source's arena into its
// own over NVLink (256-bit peer loads, 4 in flight per thread).  Replaces one NCCL broadcast per parameter at DDP
// construction (ddp_bucketed_overlapped_sharded.py:225-226,260-261) with one kernel over the packed arena.
__global__ void __launch_bounds__(kThreads) bcast_pull_kernel(Peers P, long long byte_off, long long nvec32, int src,
                                                              int rank, int ws, int slot) {
  cross_rank_barrier(P, rank, ws, slot);  // the source's data is in place
  if (rank != src) {
    const char* from = static_cast<const char*>(P.ptr[src]) + byte_off;
    char* to = static_cast<char*>(P.ptr[rank]) + byte_off;
    const long long stride = (long long)gridDim.x * kThreads;
    for (long long i0 = (long long)blockIdx.x * kThreads + threadIdx.x; i0 < nvec32; i0 += 4 * stride) {
      v8f v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u)
        if (i0 + u * stride < nvec32) v[u] = ld_peer8(from + (i0 + u * stride) * 32);
#pragma unroll
      for (int u = 0; u < 4; ++u)
        if (i0 + u * stride < nvec32) st_peer8(to + (i0 + u * stride) * 32, v[u]);
    }
  }
  cross_rank_barrier(P, rank, ws, slot);  // the source may reuse its buffer
}

// Broadcast by multicast store (NVLS): the source rank reads its own copy once and writes it through the switch
// with multimem.st — the switch replicates every store into all ranks' arenas, so the source's NVLink egress is S
// bytes instead of the (ws-1)·S that ws-1 pulling peers cost it.  Replaces one NCCL broadcast per parameter at DDP
// construction (ddp_bucketed_overlapped_sharded.py:225-226,260-261).  Non-source ranks only take part in the barriers.
__global__ void __launch_bounds__(kThreads) bcast_mc_kernel(Peers P, char* mc, long long byte_off, long long nvec16,
                                                            int src, int rank, int ws, int slot) {
  cross_rank_barrier(P, rank, ws, slot);  // nobody is still reading what is about to be overwritten
  if (rank == src) {
    const char* from = static_cast<const char*>(P.ptr[src]) + byte_off;
    char* to = mc + byte_off;
    const long long stride = (long long)gridDim.x * kThreads;
    for (long long i0 = (long long)blockIdx.x * kThreads + threadIdx.x; i0 < nvec16; i0 += 4 * stride) {
      f32x4 v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u)
        if (i0 + u * stride < nvec16) v[u] = *reinterpret_cast<const f32x4*>(from + (i0 + u * stride) * 16);
#pragma unroll
      for (int u = 0; u < 4; ++u)
        if (i0 + u * stride < nvec16)
          asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(to + (i0 + u * stride) * 16),
                       "f"(v[u].x), "f"(v[u].y), "f"(v[u].z), "f"(v[u].w) : "memory");
    }
  }
  cross_rank_barrier(P, rank, ws, slot);  // the stores have landed everywhere
}

static int load_peers(const unsigned long long* peer_ptrs_host, const unsigned long long* signal_ptrs_host, int ws,
                      Peers* out) {
  for (int r = 0; r < kMaxRanks; ++r) {
    out->ptr[r] = (r < ws && peer_ptrs_host) ? reinterpret_cast<void*>(peer_ptrs_host[r]) : nullptr;
    out->sig[r] = (r < ws && signal_ptrs_host) ? reinterpret_cast<unsigned int*>(signal_ptrs_host[r]) : nullptr;
  }
  for (int r = 0; r < ws; ++r)
    if (!out->sig[r]) { set_error("comm: null signal-pad pointer for rank %d", r); return B200_ERR_ARG; }
  return 0;
}

}  // namespace comm
}  // namespace b200

using namespace b200;
using namespace b200::comm;

extern "C" int b200_allreduce(const unsigned long long* peer_ptrs_host, const unsigned long long* signal_ptrs_host,
                              void* multicast_ptr,
                              long long byte_off, long long numel, int dtype, float scale, int rank, int world_size,
                              int algo, int blocks, int signal_slot, void* stream) {
  B200_REQUIRE(world_size >= 1 && world_size <= kMaxRanks, "allreduce: world size %d unsupported (1..8)", world_size);
  B200_REQUIRE(dtype == B200_F32 || dtype == B200_BF16, "allreduce: dtype must be fp32 or bf16");
  B200_REQUIRE(byte_off % 16 == 0, "allreduce: buffer offset must be 16-byte aligned");
  if (numel <= 0) return 0;
  const int es = dtype == B200_F32 ? 4 : 2;
  const long long nbytes = numel * es;
  const long long nvec = (nbytes + 15) / 16;  // buckets are padded to 16 bytes by the allocator
  Peers P;
  if (int rc = load_peers(peer_ptrs_host, signal_ptrs_host, world_size, &P)) return rc;
  if (blocks <= 0) blocks = nbytes >= (8 << 20) ? 48 : 16;
  if (blocks > kMaxBlocks) blocks = kMaxBlocks;
  const long long want = (nvec + kThreads - 1) / kThreads;
  if (want < blocks) blocks = (int)(want < 1 ? 1 : want);
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  if (algo == 0) algo = nbytes <= 64 * 1024 ? 1 : ((multicast_ptr && world_size > 2) ? 3 : 2);
  if (algo == 1 && nvec > (long long)blocks * kThreads * kOneShotVecs) algo = 2;  // does not fit in registers
  if (algo == 3 && !multicast_ptr) algo = 2;
  B200_REQUIRE(algo != 2 || byte_off % 32 == 0, "allreduce: the two-shot path moves 32-byte vectors; offset %lld is not 32-byte aligned", byte_off);
  const bool bf = dtype == B200_BF16;
  switch (algo) {
    case 1:
      if (bf) allreduce_oneshot_kernel<true><<<blocks, kThreads, 0, s>>>(P, byte_off, nvec, scale, rank, world_size, signal_slot);
      else allreduce_oneshot_kernel<false><<<blocks, kThreads, 0, s>>>(P, byte_off, nvec, scale, rank, world_size, signal_slot);
      break;
    case 2: {
      const long long nvec32 = (nbytes + 31) / 32;  // slots are 256-byte padded: whole 32-byte vectors are ours
#define TWOSHOT(BFV, WSV) allreduce_twoshot_kernel<BFV, WSV><<<blocks, kThreads, 0, s>>>(P, byte_off, nvec32, scale, rank, world_size, signal_slot)
      if (world_size <= 2) { if (bf) TWOSHOT(true, 2); else TWOSHOT(false, 2); }
      else if (world_size <= 4) { if (bf) TWOSHOT(true, 4); else TWOSHOT(false, 4); }
      else { if (bf) TWOSHOT(true, 8); else TWOSHOT(false, 8); }
#undef TWOSHOT
      break;
    }
    case 3:
      if (bf) allreduce_multimem_kernel<true><<<blocks, kThreads, 0, s>>>(P, static_cast<char*>(multicast_ptr), byte_off, nvec, scale, rank, world_size, signal_slot);
      else allreduce_multimem_kernel<false><<<blocks, kThreads, 0, s>>>(P, static_cast<char*>(multicast_ptr), byte_off, nvec, scale, rank, world_size, signal_slot);
      break;
    default:
      set_error("allreduce: unknown algorithm %d", algo);
      return B200_ERR_ARG;
  }
  B200_LAUNCH_CHECK();
  return 0;
}

extern "C" int b200_symm_barrier(const unsigned long long* signal_ptrs_host, int rank, int world_size, int signal_slot,
                                 void* stream) {
  B200_REQUIRE(world_size >= 1 && world_size <= kMaxRanks, "symm_barrier: world size %d unsupported", world_size);
  Peers P;
  if (int rc = load_peers(nullptr, signal_ptrs_host, world_size, &P)) return rc;
  barrier_kernel<<<1, 32, 0, reinterpret_cast<cudaStream_t>(stream)>>>(P, rank, world_size, signal_slot);
  B200_LAUNCH_CHECK();
  return 0;
}

extern "C" int b200_bcast_pull(const unsigned long long* peer_ptrs_host, const unsigned long long* signal_ptrs_host,
                               long long byte_off, long long nbytes, int src_rank, int rank, int world_size, int blocks,
                               int signal_slot, void* stream) {
  B200_REQUIRE(world_size >= 1 && world_size <= kMaxRanks, "bcast: world size %d unsupported", world_size);
  B200_REQUIRE(src_rank >= 0 && src_rank < world_size, "bcast: bad source rank %d", src_rank);
  B200_REQUIRE(byte_off % 32 == 0, "bcast: offset must be 32-byte aligned");
  if (nbytes <= 0) return 0;
  Peers P;
  if (int rc = load_peers(peer_ptrs_host, signal_ptrs_host, world_size, &P)) return rc;
  if (blocks <= 0) blocks = 48;
  if (blocks > kMaxBlocks) blocks = kMaxBlocks;
  bcast_pull_kernel<<<blocks, kThreads, 0, reinterpret_cast<cudaStream_t>(stream)>>>(P, byte_off, (nbytes + 31) / 32,
                                                                                      src_rank, rank, world_size,
                                                                                      signal_slot);
  B200_LAUNCH_CHECK();
  return 0;
}

extern "C" int b200_bcast_mc(const unsigned long long* peer_ptrs_host, const unsigned long long* signal_ptrs_host,
                             void* multicast_ptr, long long byte_off, long long nbytes, int src_rank, int rank,
                             int world_size, int blocks, int signal_slot, void* stream) {
  B200_REQUIRE(world_size >= 1 && world_size <= kMaxRanks, "bcast_mc: world size %d unsupported", world_size);
  B200_REQUIRE(src_rank >= 0 && src_rank < world_size, "bcast_mc: bad source rank %d", src_rank);
  B200_REQUIRE(byte_off % 16 == 0, "bcast_mc: offset must be 16-byte aligned");
  B200_REQUIRE(multicast_ptr != nullptr, "bcast_mc: the arena has no NVLS multicast mapping (use b200_bcast_pull)");
  if (nbytes <= 0) return 0;
  Peers P;
  if (int rc = load_peers(peer_ptrs_host, signal_ptrs_host, world_size, &P)) return rc;
  // only the source moves data and nothing else runs at DDP construction: one CTA per SM keeps the switch fed
  if (blocks <= 0 || blocks > kSignalBlocks) blocks = kSignalBlocks;
  bcast_mc_kernel<<<blocks, kThreads, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      P, static_cast<char*>(multicast_ptr), byte_off, (nbytes + 15) / 16, src_rank, rank, world_size, signal_slot);
  B200_LAUNCH_CHECK();
  return 0;
}

// 32-bit words of one signal slot: flag words for the largest grid any barrier-using kernel launches.
extern "C" int b200_comm_signal_words(void) { return kSignalBlocks * kMaxRanks; }

namespace b200 { namespace comm { int adamw_tu_status(CommStatus* out); int adamw_tu_reset(unsigned long long timeout_ns); } }

// Health of the NVLink communicator on the current device: 0 = fine; 1 = a cross-rank barrier timed out (info[0..5] =
// rank, CTA, peer, phase (1 = raising the peer's flag, 2 = waiting for the peer), signal slot, flag value seen) —
// everything produced by comm kernels since then is invalid.  Synchronise the device (or the stream) first.
extern "C" int b200_comm_status(int* info6) {
  CommStatus st[2] = {};
  B200_REQUIRE(comm_status_read(&st[0]) == 0, "comm_status: cannot read the status block");
  B200_REQUIRE(adamw_tu_status(&st[1]) == 0, "comm_status: cannot read the status block");
  for (int i = 0; i < 2; ++i)
    if (st[i].dead) {
      if (info6) { info6[0] = st[i].rank; info6[1] = st[i].block; info6[2] = st[i].peer; info6[3] = st[i].phase; info6[4] = st[i].slot; info6[5] = (int)st[i].flag; }
      return 1;
    }
  return 0;
}

// Clears the status block and sets the barrier timeout (ms; 0 keeps the current value) on the current device.
extern "C" int b200_comm_reset(long long timeout_ms) {
  const unsigned long long ns = timeout_ms > 0 ? (unsigned long long)timeout_ms * 1000000ull : 0ull;
  B200_REQUIRE(comm_status_reset_and_set_timeout(ns) == 0, "comm_reset: cudaMemcpyToSymbol failed");
  B200_REQUIRE(adamw_tu_reset(ns) == 0, "comm_reset: cudaMemcpyToSymbol failed");
  return 0;
}
