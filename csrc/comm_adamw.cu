// Owner-rank AdamW fused with the distribution of the updated parameters over NVLink.
//
// The reference's ShardedStateOptimizer steps the owned parameters and then issues one blocking
// dist.broadcast per parameter tensor for ALL parameters (ddp_bucketed_overlapped_sharded.py:335-340:
// 435 serial NCCL calls, 7.4 GiB per step for the XL model).  Here the parameters of every rank live at
// identical offsets of a symmetric arena, and the multi-tensor AdamW kernel of the owner stores each
// new value straight into every replica (peer stores) together with its bf16 GEMM shadow — the
// broadcast disappears; one flag barrier afterwards orders the stores before the next forward.
#include "common.cuh"

namespace b200 {
namespace comm {

constexpr int kMaxRanksD = 8;
constexpr int kChunkD = 4096;  // must match adamw.cu (b200_adamw_chunk_elems)

struct Bases {
  char* param[kMaxRanksD];
  char* shadow[kMaxRanksD];
};

__device__ __forceinline__ int find_tensor_d(const long long* __restrict__ chunk_start, int n, long long chunk) {
  int lo = 0, hi = n - 1;
  while (lo < hi) {
    const int mid = (lo + hi + 1) >> 1;
    if (chunk_start[mid] <= chunk) lo = mid; else hi = mid - 1;
  }
  return lo;
}

__global__ void __launch_bounds__(256) adamw_distribute_kernel(const b200_adamw_tensor* __restrict__ table,
                                                               const long long* __restrict__ chunk_start,
                                                               int n_tensors, long long total_chunks,
                                                               const float* __restrict__ hyper,
                                                               const int* __restrict__ step, int mode, Bases B,
                                                               char* local_param_base, char* local_shadow_base,
                                                               int ws) {
  const float lr = hyper[0], b1 = hyper[1], b2 = hyper[2], eps = hyper[3], wd = hyper[4], gscale = hyper[5];
  const float t = (float)(*step);
  const float bc1 = 1.f - powf(b1, t), bc2 = 1.f - powf(b2, t);
  const float alpha_t = lr * sqrtf(bc2) / bc1;
  const float rsqrt_bc2 = rsqrtf(bc2);
  for (long long chunk = blockIdx.x; chunk < total_chunks; chunk += gridDim.x) {
    const int ti = find_tensor_d(chunk_start, n_tensors, chunk);
    const b200_adamw_tensor td = table[ti];
    const long long off0 = (chunk - chunk_start[ti]) * kChunkD;
    const long long p_off = reinterpret_cast<char*>(td.p) - local_param_base;  // bytes into the arena
    const long long s_off = td.shadow_bf16 ? reinterpret_cast<char*>(td.shadow_bf16) - local_shadow_base : 0;
#pragma unroll
    for (int it = 0; it < 4; ++it) {
      const long long i = off0 + ((long long)it * 256 + threadIdx.x) * 4;
      if (i >= td.n) break;
      const int cnt = (td.n - i) >= 4 ? 4 : (int)(td.n - i);
      float p[4], g[4], m[4], v[4];
      for (int k = 0; k < 4; ++k) {
        const bool ok = k < cnt;
        p[k] = ok ? td.p[i + k] : 0.f; g[k] = ok ? td.g[i + k] * gscale : 0.f;
        m[k] = ok ? td.m[i + k] : 0.f; v[k] = ok ? td.v[i + k] : 0.f;
      }
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        m[k] = b1 * m[k] + (1.f - b1) * g[k];
        v[k] = b2 * v[k] + (1.f - b2) * g[k] * g[k];
        if (mode == 0) {
          p[k] -= alpha_t * m[k] / (sqrtf(v[k]) + eps);
          p[k] -= lr * wd * p[k];
        } else {
          p[k] *= (1.f - lr * wd);
          p[k] -= (lr / bc1) * m[k] / (sqrtf(v[k]) * rsqrt_bc2 + eps);
        }
      }
      for (int k = 0; k < cnt; ++k) { td.m[i + k] = m[k]; td.v[i + k] = v[k]; }
      const bool vec = cnt == 4 && td.aligned16;
      for (int r = 0; r < ws; ++r) {
        float* dst = reinterpret_cast<float*>(B.param[r] + p_off) + i;
        if (vec) {
          asm volatile("st.global.relaxed.sys.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(dst), "f"(p[0]), "f"(p[1]), "f"(p[2]), "f"(p[3]) : "memory");
        } else {
          for (int k = 0; k < cnt; ++k) dst[k] = p[k];
        }
        if (td.shadow_bf16 && B.shadow[r]) {
          __nv_bfloat16* sd = reinterpret_cast<__nv_bfloat16*>(B.shadow[r] + s_off) + i;
          if (vec) {
            asm volatile("st.global.relaxed.sys.v2.b32 [%0], {%1,%2};" ::"l"(sd), "r"(pack_bf16x2(p[0], p[1])), "r"(pack_bf16x2(p[2], p[3])) : "memory");
          } else {
            for (int k = 0; k < cnt; ++k) sd[k] = __float2bfloat16(p[k]);
          }
        }
      }
    }
  }
}

__global__ void incr_step_kernel_d(int* step) { *step += 1; }

}  // namespace comm
}  // namespace b200

namespace b200 {
namespace comm {

// ------------------------------------------------------------------------------------------------
// Fused gradient reduce-scatter + sharded AdamW + parameter all-gather for one DDP bucket.
//
// Gradient bucket, parameter bucket and bf16 shadow bucket have the same element layout on every rank
// (symmetric arenas).  Rank r owns the r-th slice of the bucket: it averages that slice of the gradients
// straight out of every rank's bucket (NVLink loads), applies AdamW with ITS slice of m and v (optimizer state
// exists once across the job), and stores the new parameters (fp32) and their bf16 GEMM shadows into every
// rank's arenas (NVLink stores).  One kernel replaces: all-reduce + 1/ws scale + full-model AdamW on every
// rank + (for the reference's sharded optimizer) one broadcast per parameter.
// ------------------------------------------------------------------------------------------------
struct v8 { float f[8]; };
__device__ __forceinline__ v8 ldsys8(const void* p) {
  v8 r;
  asm volatile("ld.global.relaxed.sys.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=f"(r.f[0]), "=f"(r.f[1]), "=f"(r.f[2]), "=f"(r.f[3]), "=f"(r.f[4]), "=f"(r.f[5]), "=f"(r.f[6]), "=f"(r.f[7])
               : "l"(p) : "memory");
  return r;
}
__device__ __forceinline__ void stsys8(void* p, const v8& r) {
  asm volatile("st.global.relaxed.sys.v8.f32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p), "f"(r.f[0]), "f"(r.f[1]),
               "f"(r.f[2]), "f"(r.f[3]), "f"(r.f[4]), "f"(r.f[5]), "f"(r.f[6]), "f"(r.f[7]) : "memory");
}
__device__ __forceinline__ void stsys4u(void* p, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("st.global.relaxed.sys.v4.b32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
__device__ __forceinline__ unsigned cas_rel(unsigned* a, unsigned c, unsigned v) {
  unsigned o;
  asm volatile("atom.global.release.sys.cas.b32 %0, [%1], %2, %3;" : "=r"(o) : "l"(a), "r"(c), "r"(v) : "memory");
  return o;
}
__device__ __forceinline__ unsigned cas_acq(unsigned* a, unsigned c, unsigned v) {
  unsigned o;
  asm volatile("atom.global.acquire.sys.cas.b32 %0, [%1], %2, %3;" : "=r"(o) : "l"(a), "r"(c), "r"(v) : "memory");
  return o;
}
struct ZPeers {
  char* grad[kMaxRanksD];
  char* param[kMaxRanksD];
  char* shadow[kMaxRanksD];
  unsigned int* sig[kMaxRanksD];
};
__device__ __forceinline__ void zbarrier(const ZPeers& P, int rank, int ws, int slot) {
  __syncthreads();
  if (threadIdx.x < ws) {
    const int peer = threadIdx.x;
    unsigned* send = P.sig[peer] + slot + blockIdx.x * kMaxRanksD + rank;
    unsigned* recv = P.sig[rank] + slot + blockIdx.x * kMaxRanksD + peer;
    while (cas_rel(send, 0u, 1u) != 0u) {
    }
    while (cas_acq(recv, 1u, 0u) != 1u) {
    }
  }
  __syncthreads();
}

struct ZMulticast {
  char* grad;    // NVLS multicast address of the gradient arena (nullptr: peer-to-peer path)
  char* param;
  char* shadow;
};
__device__ __forceinline__ void mm_ld_reduce8(const char* a, v8& r) {
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(r.f[0]), "=f"(r.f[1]), "=f"(r.f[2]), "=f"(r.f[3]) : "l"(a) : "memory");
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(r.f[4]), "=f"(r.f[5]), "=f"(r.f[6]), "=f"(r.f[7]) : "l"(a + 16) : "memory");
}
__device__ __forceinline__ void mm_st4(char* a, float x, float y, float z, float w) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(a), "f"(x), "f"(y), "f"(z), "f"(w) : "memory");
}

constexpr int kZThreads = 256;
constexpr int kZMaxBlocks = 148;  // one CTA per SM, co-resident with the backward GEMMs (<=112 regs x 256 threads next to the GEMM CTA)

// MMLD: the switch reduces the gradient slice (multimem.ld_reduce) instead of WS peer loads per vector.
// MMST: the switch replicates the parameter / shadow stores (multimem.st) instead of one store per peer.
template <int WS, bool MMLD, bool MMST>
__global__ void __launch_bounds__(kZThreads, 2) zero_adamw_kernel(ZPeers P, ZMulticast MC, long long byte_off,
                                                               long long nvec32, float* __restrict__ m_slice,
                                                               float* __restrict__ v_slice,
                                                               const float* __restrict__ hyper,
                                                               const int* __restrict__ step, int mode, int rank, int ws,
                                                               int slot) {
  constexpr int U = 2;  // vectors per thread per trip: all remote loads of a trip are in flight together
  zbarrier(P, rank, ws, slot);  // every rank's gradients for this bucket are final
  const float lr = hyper[0], b1 = hyper[1], b2 = hyper[2], eps = hyper[3], wd = hyper[4];
  const float t = (float)(*step);
  const float bc1 = 1.f - powf(b1, t), bc2 = 1.f - powf(b2, t);
  const float alpha_t = lr * sqrtf(bc2) / bc1, rsqrt_bc2 = rsqrtf(bc2);
  const float inv_ws = 1.f / (float)ws;
  const long long per = (nvec32 + ws - 1) / ws;
  const long long lo = per * rank, hi = (lo + per < nvec32) ? lo + per : nvec32;
  const long long stride = (long long)gridDim.x * kZThreads;
  for (long long i0 = lo + (long long)blockIdx.x * kZThreads + threadIdx.x; i0 < hi; i0 += U * stride) {
    v8 g[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long i = i0 + u * stride;
      if (i < hi) {
        if constexpr (MMLD) {
          mm_ld_reduce8(MC.grad + byte_off + i * 32, g[u]);
        } else {
          v8 part[WS];
#pragma unroll
          for (int r = 0; r < WS; ++r)
            if (r < ws) part[r] = ldsys8(P.grad[r] + byte_off + i * 32);
#pragma unroll
          for (int k = 0; k < 8; ++k) {
            float a = 0.f;
#pragma unroll
            for (int r = 0; r < WS; ++r)
              if (r < ws) a += part[r].f[k];
            g[u].f[k] = a;
          }
        }
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long i = i0 + u * stride;
      if (i >= hi) continue;
      const long long li = (i - lo) * 8;  // index into this rank's m / v slices
      v8 pv = *reinterpret_cast<const v8*>(P.param[rank] + byte_off + i * 32);
      v8 mv = *reinterpret_cast<const v8*>(m_slice + li);
      v8 vv = *reinterpret_cast<const v8*>(v_slice + li);
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const float gk = g[u].f[k] * inv_ws;
        const float mk = b1 * mv.f[k] + (1.f - b1) * gk;
        const float vk = b2 * vv.f[k] + (1.f - b2) * gk * gk;
        float pk = pv.f[k];
        if (mode == 0) {
          pk -= alpha_t * mk / (sqrtf(vk) + eps);
          pk -= lr * wd * pk;
        } else {
          pk *= (1.f - lr * wd);
          pk -= (lr / bc1) * mk / (sqrtf(vk) * rsqrt_bc2 + eps);
        }
        mv.f[k] = mk; vv.f[k] = vk; pv.f[k] = pk;
      }
      *reinterpret_cast<v8*>(m_slice + li) = mv;
      *reinterpret_cast<v8*>(v_slice + li) = vv;
      const uint32_t s0 = pack_bf16x2(pv.f[0], pv.f[1]), s1 = pack_bf16x2(pv.f[2], pv.f[3]);
      const uint32_t s2 = pack_bf16x2(pv.f[4], pv.f[5]), s3 = pack_bf16x2(pv.f[6], pv.f[7]);
      if constexpr (MMST) {
        char* pa = MC.param + byte_off + i * 32;
        mm_st4(pa, pv.f[0], pv.f[1], pv.f[2], pv.f[3]);
        mm_st4(pa + 16, pv.f[4], pv.f[5], pv.f[6], pv.f[7]);
        if (MC.shadow)
          mm_st4(MC.shadow + (byte_off >> 1) + i * 16, __uint_as_float(s0), __uint_as_float(s1), __uint_as_float(s2),
                 __uint_as_float(s3));
      } else {
#pragma unroll
        for (int r = 0; r < kMaxRanksD; ++r)
          if (r < ws) {
            stsys8(P.param[r] + byte_off + i * 32, pv);
            if (P.shadow[r]) stsys4u(P.shadow[r] + (byte_off >> 1) + i * 16, s0, s1, s2, s3);
          }
      }
    }
  }
  zbarrier(P, rank, ws, slot);  // every owner's parameter stores have landed on every rank
}

// NVLS-reduce variant on 16-byte units: consecutive lanes touch consecutive 16-byte chunks, so every
// multimem.ld_reduce warp instruction covers one contiguous 512-byte span (the 32-byte-per-lane layout of the
// kernel above splits each span over two instructions and ran 1.7x slower through the switch).
struct v4 { float f[4]; };
template <bool MMST>
__global__ void __launch_bounds__(kZThreads) zero_adamw_mm_kernel(ZPeers P, ZMulticast MC, long long byte_off,
                                                                  long long nvec32, float* __restrict__ m_slice,
                                                                  float* __restrict__ v_slice,
                                                                  const float* __restrict__ hyper,
                                                                  const int* __restrict__ step, int mode, int rank,
                                                                  int ws, int slot) {
  constexpr int U = 4;
  zbarrier(P, rank, ws, slot);
  const float lr = hyper[0], b1 = hyper[1], b2 = hyper[2], eps = hyper[3], wd = hyper[4];
  const float t = (float)(*step);
  const float bc1 = 1.f - powf(b1, t), bc2 = 1.f - powf(b2, t);
  const float alpha_t = lr * sqrtf(bc2) / bc1, rsqrt_bc2 = rsqrtf(bc2);
  const float inv_ws = 1.f / (float)ws;
  const long long per = (nvec32 + ws - 1) / ws;
  const long long lo32 = per * rank, hi32 = (lo32 + per < nvec32) ? lo32 + per : nvec32;
  const long long lo = lo32 * 2, hi = hi32 * 2;  // 16-byte units
  const long long stride = (long long)gridDim.x * kZThreads;
  for (long long i0 = lo + (long long)blockIdx.x * kZThreads + threadIdx.x; i0 < hi; i0 += U * stride) {
    v4 g[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long i = i0 + u * stride;
      if (i < hi)
        asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                     : "=f"(g[u].f[0]), "=f"(g[u].f[1]), "=f"(g[u].f[2]), "=f"(g[u].f[3])
                     : "l"(MC.grad + byte_off + i * 16) : "memory");
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long i = i0 + u * stride;
      if (i >= hi) continue;
      const long long li = (i - lo) * 4;
      float4 pv = *reinterpret_cast<const float4*>(P.param[rank] + byte_off + i * 16);
      float4 mv = *reinterpret_cast<const float4*>(m_slice + li);
      float4 vv = *reinterpret_cast<const float4*>(v_slice + li);
      float* pp = reinterpret_cast<float*>(&pv);
      float* mp = reinterpret_cast<float*>(&mv);
      float* vp = reinterpret_cast<float*>(&vv);
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const float gk = g[u].f[k] * inv_ws;
        const float mk = b1 * mp[k] + (1.f - b1) * gk;
        const float vk = b2 * vp[k] + (1.f - b2) * gk * gk;
        float pk = pp[k];
        if (mode == 0) {
          pk -= alpha_t * mk / (sqrtf(vk) + eps);
          pk -= lr * wd * pk;
        } else {
          pk *= (1.f - lr * wd);
          pk -= (lr / bc1) * mk / (sqrtf(vk) * rsqrt_bc2 + eps);
        }
        mp[k] = mk; vp[k] = vk; pp[k] = pk;
      }
      *reinterpret_cast<float4*>(m_slice + li) = mv;
      *reinterpret_cast<float4*>(v_slice + li) = vv;
      const uint32_t s0 = pack_bf16x2(pv.x, pv.y), s1 = pack_bf16x2(pv.z, pv.w);
      if constexpr (MMST) {
        mm_st4(MC.param + byte_off + i * 16, pv.x, pv.y, pv.z, pv.w);
        if (MC.shadow)
          asm volatile("multimem.st.relaxed.sys.global.v2.f32 [%0], {%1,%2};" ::"l"(MC.shadow + (byte_off >> 1) + i * 8),
                       "f"(__uint_as_float(s0)), "f"(__uint_as_float(s1)) : "memory");
      } else {
#pragma unroll
        for (int r = 0; r < kMaxRanksD; ++r)
          if (r < ws) {
            asm volatile("st.global.relaxed.sys.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(P.param[r] + byte_off + i * 16),
                         "f"(pv.x), "f"(pv.y), "f"(pv.z), "f"(pv.w) : "memory");
            if (P.shadow[r])
              asm volatile("st.global.relaxed.sys.v2.b32 [%0], {%1,%2};" ::"l"(P.shadow[r] + (byte_off >> 1) + i * 8),
                           "r"(s0), "r"(s1) : "memory");
          }
      }
    }
  }
  zbarrier(P, rank, ws, slot);
}

__global__ void zero_incr_step_kernel(int* step) { *step += 1; }

}  // namespace comm
}  // namespace b200

using namespace b200;

extern "C" int b200_zero_adamw_bucket(const unsigned long long* grad_peers_host, const unsigned long long* param_peers_host,
                                      const unsigned long long* shadow_peers_host /*nullable*/,
                                      const unsigned long long* signal_ptrs_host, const unsigned long long* multicast3_host,
                                      long long byte_off, long long numel, float* m_slice, float* v_slice,
                                      const float* hyper_dev, const int* step_dev, int mode, int rank, int world_size,
                                      int blocks, int signal_slot, int variant, void* stream) {
  using namespace b200::comm;
  B200_REQUIRE(world_size >= 1 && world_size <= kMaxRanksD, "zero_adamw: world size %d unsupported", world_size);
  B200_REQUIRE(byte_off % 32 == 0, "zero_adamw: bucket offset must be 32-byte aligned");
  if (numel <= 0) return 0;
  ZPeers P;
  for (int r = 0; r < kMaxRanksD; ++r) {
    const bool on = r < world_size;
    P.grad[r] = on ? reinterpret_cast<char*>(grad_peers_host[r]) : nullptr;
    P.param[r] = on ? reinterpret_cast<char*>(param_peers_host[r]) : nullptr;
    P.shadow[r] = (on && shadow_peers_host) ? reinterpret_cast<char*>(shadow_peers_host[r]) : nullptr;
    P.sig[r] = on ? reinterpret_cast<unsigned int*>(signal_ptrs_host[r]) : nullptr;
  }
  // variant bit 0: reduce the gradients in the switch (needs the multicast mapping of the gradient arena);
  // variant bit 1: replicate parameter / shadow stores in the switch (needs those arenas' multicast mappings)
  ZMulticast MC{nullptr, nullptr, nullptr};
  const bool mmld = (variant & 1) != 0, mmst = (variant & 2) != 0;
  if (mmld) {
    B200_REQUIRE(multicast3_host && multicast3_host[0], "zero_adamw: NVLS reduce needs the gradient arena's multicast address");
    MC.grad = reinterpret_cast<char*>(multicast3_host[0]);
  }
  if (mmst) {
    B200_REQUIRE(multicast3_host && multicast3_host[1] && (!shadow_peers_host || multicast3_host[2]),
                 "zero_adamw: NVLS stores need the parameter (and shadow) arena's multicast address");
    MC.param = reinterpret_cast<char*>(multicast3_host[1]);
    MC.shadow = shadow_peers_host ? reinterpret_cast<char*>(multicast3_host[2]) : nullptr;
  }
  const long long nvec32 = (numel * 4 + 31) / 32;
  if (blocks <= 0) blocks = kZMaxBlocks;
  if (blocks > kZMaxBlocks) blocks = kZMaxBlocks;
  const long long want = (nvec32 / world_size + 2 * kZThreads - 1) / (2 * kZThreads);
  if (want < blocks) blocks = (int)(want < 1 ? 1 : want);
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
#define ZLAUNCH(WSV, LD, ST) zero_adamw_kernel<WSV, LD, ST><<<blocks, kZThreads, 0, s>>>(P, MC, byte_off, nvec32, m_slice, v_slice, hyper_dev, step_dev, mode, rank, world_size, signal_slot)
  if (mmld) {
    if (mmst) zero_adamw_mm_kernel<true><<<blocks, kZThreads, 0, s>>>(P, MC, byte_off, nvec32, m_slice, v_slice, hyper_dev, step_dev, mode, rank, world_size, signal_slot);
    else zero_adamw_mm_kernel<false><<<blocks, kZThreads, 0, s>>>(P, MC, byte_off, nvec32, m_slice, v_slice, hyper_dev, step_dev, mode, rank, world_size, signal_slot);
  } else if (world_size <= 2) {
    if (mmst) ZLAUNCH(2, false, true); else ZLAUNCH(2, false, false);
  } else if (world_size <= 4) {
    if (mmst) ZLAUNCH(4, false, true); else ZLAUNCH(4, false, false);
  } else {
    if (mmst) ZLAUNCH(8, false, true); else ZLAUNCH(8, false, false);
  }
#undef ZLAUNCH
  B200_LAUNCH_CHECK();
  return 0;
}

extern "C" int b200_incr_step(int* step_dev, void* stream) {
  comm::zero_incr_step_kernel<<<1, 1, 0, reinterpret_cast<cudaStream_t>(stream)>>>(step_dev);
  B200_LAUNCH_CHECK();
  return 0;
}

extern "C" int b200_adamw_multi_distribute(const b200_adamw_tensor* table_dev, const long long* chunk_start_dev,
                                           int n_tensors, long long total_chunks, const float* hyper_dev,
                                           int* step_dev, int mode, int increment_step,
                                           const unsigned long long* param_peer_bases_host,
                                           const unsigned long long* shadow_peer_bases_host,
                                           unsigned long long local_param_base, unsigned long long local_shadow_base,
                                           int world_size, void* stream) {
  B200_REQUIRE(world_size >= 1 && world_size <= comm::kMaxRanksD, "adamw_distribute: world size %d unsupported", world_size);
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  if (n_tensors > 0 && total_chunks > 0) {
    comm::Bases B;
    for (int r = 0; r < comm::kMaxRanksD; ++r) {
      B.param[r] = r < world_size ? reinterpret_cast<char*>(param_peer_bases_host[r]) : nullptr;
      B.shadow[r] = (r < world_size && shadow_peer_bases_host) ? reinterpret_cast<char*>(shadow_peer_bases_host[r]) : nullptr;
    }
    const long long cap = (long long)num_sms() * 8;
    const int grid = (int)(total_chunks < cap ? total_chunks : cap);
    comm::adamw_distribute_kernel<<<grid, 256, 0, s>>>(table_dev, chunk_start_dev, n_tensors, total_chunks, hyper_dev,
                                                       step_dev, mode, B, reinterpret_cast<char*>(local_param_base),
                                                       reinterpret_cast<char*>(local_shadow_base), world_size);
    B200_LAUNCH_CHECK();
  }
  if (increment_step) {
    comm::incr_step_kernel_d<<<1, 1, 0, s>>>(step_dev);
    B200_LAUNCH_CHECK();
  }
  return 0;
}
