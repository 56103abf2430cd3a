// Owner-rank AdamW fused with the distribution of the updated parameters over NVLink.
//
// The reference's ShardedStateOptimizer steps the owned parameters and then issues one blocking
// dist.broadcast per parameter tensor for ALL parameters (ddp_bucketed_overlapped_sharded.py:335-340:
// 435 serial NCCL calls, 7.4 GiB per step for the XL model).  Here the parameters of every rank live at
// identical offsets of a symmetric arena, and the multi-tensor AdamW kernel of the owner stores each
// new value straight into every replica (peer stores) together with its bf16 GEMM shadow — the
// broadcast disappears; one flag barrier afterwards orders the stores before the next forward.
#include "common.cuh"
#include "comm_barrier.cuh"

namespace b200 {
namespace comm {

constexpr int kMaxRanksD = kMaxRanks;
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
// straight out of every rank's bucket (NVLink loads, or one multimem.ld_reduce through the switch), applies AdamW
// with ITS slice of m and v (optimizer state exists once across the job), and stores the new parameters (fp32) and
// their bf16 GEMM shadows into every rank's arenas (NVLink stores / multimem.st).  One kernel replaces:
// all-reduce + 1/ws scale + full-model AdamW on every rank + (for the reference's sharded optimizer) one
// broadcast per parameter.
//
// Deferred fp32 gather (flag bit 0 clear): only the bf16 shadows — what the next forward's GEMMs read — cross
// NVLink inside this kernel, plus the fp32 values of the few "eager" ranges that forward reads in fp32 (RMSNorm
// gains, the embedding table).  The fp32 replicas of everything else are brought up to date later by
// zero_param_gather_kernel, off the critical path (during the next forward).  That takes 4 of the 6 bytes per
// parameter out of the backward-overlapped window.
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
__device__ __forceinline__ void stsys4f(void* p, float a, float b, float c, float d) {
  asm volatile("st.global.relaxed.sys.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}
__device__ __forceinline__ void stsys4u(void* p, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("st.global.relaxed.sys.v4.b32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
__device__ __forceinline__ void stsys2u(void* p, uint32_t a, uint32_t b) {
  asm volatile("st.global.relaxed.sys.v2.b32 [%0], {%1,%2};" ::"l"(p), "r"(a), "r"(b) : "memory");
}
struct ZPeers {
  char* grad[kMaxRanksD];
  char* param[kMaxRanksD];
  char* shadow[kMaxRanksD];
  unsigned int* sig[kMaxRanksD];
};
__device__ __forceinline__ void zbarrier(const ZPeers& P, int rank, int ws, int slot) {
  cross_rank_barrier(P.sig, rank, ws, slot);
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
__device__ __forceinline__ void mm_st2u(char* a, uint32_t x, uint32_t y) {
  asm volatile("multimem.st.relaxed.sys.global.v2.f32 [%0], {%1,%2};" ::"l"(a), "f"(__uint_as_float(x)), "f"(__uint_as_float(y)) : "memory");
}

constexpr int kZThreads = 256;
constexpr int kZMaxBlocks = 148;  // one CTA per SM, co-resident with the backward GEMMs (<=112 regs x 256 threads next to the GEMM CTA)
static_assert(kZMaxBlocks <= kSignalBlocks, "a signal slot holds kSignalBlocks CTAs' flag words");
constexpr int kZEager = 4;        // fp32-eager ranges per bucket (RMSNorm gains, embedding table)

// Ranges (in 16-byte units from the bucket start, [lo, hi)) whose fp32 values every rank needs right after the
// step even when the fp32 gather is deferred.
struct ZEager {
  int n;
  long long lo[kZEager], hi[kZEager];
};
__device__ __forceinline__ bool in_eager(const ZEager& E, long long u16) {
  bool hit = false;
#pragma unroll
  for (int k = 0; k < kZEager; ++k) hit |= (k < E.n) && u16 >= E.lo[k] && u16 < E.hi[k];
  return hit;
}

struct ZHyper {
  float lr, b1, b2, eps, wd, alpha_t, rsqrt_bc2, bc1, inv_ws;
  int mode;
};
__device__ __forceinline__ ZHyper load_hyper(const float* __restrict__ hyper, const int* __restrict__ step, int mode, int ws) {
  ZHyper h;
  h.lr = hyper[0]; h.b1 = hyper[1]; h.b2 = hyper[2]; h.eps = hyper[3]; h.wd = hyper[4];
  const float t = (float)(*step);
  h.bc1 = 1.f - powf(h.b1, t);
  const float bc2 = 1.f - powf(h.b2, t);
  h.alpha_t = h.lr * sqrtf(bc2) / h.bc1;
  h.rsqrt_bc2 = rsqrtf(bc2);
  h.inv_ws = 1.f / (float)ws;
  h.mode = mode;
  return h;
}
// one element of the course AdamW (mode 0: update then decay, cs336_basics/optimizer.py:74-80) or torch.optim.AdamW (1)
__device__ __forceinline__ void adamw1(const ZHyper& h, float g, float& m, float& v, float& p) {
  const float gk = g * h.inv_ws;
  m = h.b1 * m + (1.f - h.b1) * gk;
  v = h.b2 * v + (1.f - h.b2) * gk * gk;
  if (h.mode == 0) {
    p -= h.alpha_t * m / (sqrtf(v) + h.eps);
    p -= h.lr * h.wd * p;
  } else {
    p *= (1.f - h.lr * h.wd);
    p -= (h.lr / h.bc1) * m / (sqrtf(v) * h.rsqrt_bc2 + h.eps);
  }
}

// MMLD: the switch reduces the gradient slice (multimem.ld_reduce) instead of WS peer loads per vector.
// MMST: the switch replicates the parameter / shadow stores (multimem.st) instead of one store per peer.
// flags bit 0: store the fp32 parameters into every rank's arena here (clear: only the eager ranges; the rest
// waits for zero_param_gather_kernel).
template <int WS, bool MMLD, bool MMST>
__global__ void __launch_bounds__(kZThreads, 2) zero_adamw_kernel(ZPeers P, ZMulticast MC, long long byte_off,
                                                               long long nvec32, float* __restrict__ m_slice,
                                                               float* __restrict__ v_slice,
                                                               const float* __restrict__ hyper,
                                                               const int* __restrict__ step, int mode, int rank, int ws,
                                                               int slot, int flags, ZEager E) {
  constexpr int U = 2;  // vectors per thread per trip: all remote loads of a trip are in flight together
  zbarrier(P, rank, ws, slot);  // every rank's gradients for this bucket are final
  const ZHyper h = load_hyper(hyper, step, mode, ws);
  const bool all_fp32 = (flags & 1) != 0;
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
      for (int k = 0; k < 8; ++k) adamw1(h, g[u].f[k], mv.f[k], vv.f[k], pv.f[k]);
      *reinterpret_cast<v8*>(m_slice + li) = mv;
      *reinterpret_cast<v8*>(v_slice + li) = vv;
      const uint32_t s0 = pack_bf16x2(pv.f[0], pv.f[1]), s1 = pack_bf16x2(pv.f[2], pv.f[3]);
      const uint32_t s2 = pack_bf16x2(pv.f[4], pv.f[5]), s3 = pack_bf16x2(pv.f[6], pv.f[7]);
      const bool wide = all_fp32 || in_eager(E, i * 2);
      if (!wide) *reinterpret_cast<v8*>(P.param[rank] + byte_off + i * 32) = pv;  // own replica only, for now
      if constexpr (MMST) {
        if (wide) {
          char* pa = MC.param + byte_off + i * 32;
          mm_st4(pa, pv.f[0], pv.f[1], pv.f[2], pv.f[3]);
          mm_st4(pa + 16, pv.f[4], pv.f[5], pv.f[6], pv.f[7]);
        }
        if (MC.shadow)
          mm_st4(MC.shadow + (byte_off >> 1) + i * 16, __uint_as_float(s0), __uint_as_float(s1), __uint_as_float(s2),
                 __uint_as_float(s3));
      } else {
#pragma unroll
        for (int r = 0; r < kMaxRanksD; ++r)
          if (r < ws) {
            if (wide) stsys8(P.param[r] + byte_off + i * 32, pv);
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
__global__ void __launch_bounds__(kZThreads, 2) zero_adamw_mm_kernel(ZPeers P, ZMulticast MC, long long byte_off,
                                                                     long long nvec32, float* __restrict__ m_slice,
                                                                     float* __restrict__ v_slice,
                                                                     const float* __restrict__ hyper,
                                                                     const int* __restrict__ step, int mode, int rank,
                                                                     int ws, int slot, int flags, ZEager E) {
  constexpr int U = 4;
  zbarrier(P, rank, ws, slot);
  const ZHyper h = load_hyper(hyper, step, mode, ws);
  const bool all_fp32 = (flags & 1) != 0;
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
      for (int k = 0; k < 4; ++k) adamw1(h, g[u].f[k], mp[k], vp[k], pp[k]);
      *reinterpret_cast<float4*>(m_slice + li) = mv;
      *reinterpret_cast<float4*>(v_slice + li) = vv;
      const uint32_t s0 = pack_bf16x2(pv.x, pv.y), s1 = pack_bf16x2(pv.z, pv.w);
      const bool wide = all_fp32 || in_eager(E, i);
      if (!wide) *reinterpret_cast<float4*>(P.param[rank] + byte_off + i * 16) = pv;
      if constexpr (MMST) {
        if (wide) mm_st4(MC.param + byte_off + i * 16, pv.x, pv.y, pv.z, pv.w);
        if (MC.shadow) mm_st2u(MC.shadow + (byte_off >> 1) + i * 8, s0, s1);
      } else {
#pragma unroll
        for (int r = 0; r < kMaxRanksD; ++r)
          if (r < ws) {
            if (wide) stsys4f(P.param[r] + byte_off + i * 16, pv.x, pv.y, pv.z, pv.w);
            if (P.shadow[r]) stsys2u(P.shadow[r] + (byte_off >> 1) + i * 8, s0, s1);
          }
      }
    }
  }
  zbarrier(P, rank, ws, slot);
}

// Deferred fp32 parameter all-gather: for every bucket of the table, this rank pushes ITS slice of the fp32
// parameters (which it alone updated) into every other rank's arena — NVLink stores, or one multimem.st per vector
// through the switch.  Runs on the communication stream while the next forward computes (forward's GEMMs read the
// bf16 shadows, which the bucket kernels already distributed).  One barrier at the end: when it passes, every
// rank's pushes have landed in this rank's arena.
struct ZBucket { long long byte_off, nvec32; };
template <bool MMST>
__global__ void __launch_bounds__(kZThreads, 2) zero_param_gather_kernel(ZPeers P, ZMulticast MC,
                                                                         const ZBucket* __restrict__ buckets,
                                                                         int n_buckets, int rank, int ws, int slot) {
  constexpr int U = 4;
  const long long stride = (long long)gridDim.x * kZThreads;
  for (int b = 0; b < n_buckets; ++b) {
    const long long byte_off = buckets[b].byte_off, nvec32 = buckets[b].nvec32;
    const long long per = (nvec32 + ws - 1) / ws;
    const long long lo32 = per * rank, hi32 = (lo32 + per < nvec32) ? lo32 + per : nvec32;
    const long long lo = lo32 * 2, hi = hi32 * 2;  // 16-byte units
    const char* mine = P.param[rank] + byte_off;
    for (long long i0 = lo + (long long)blockIdx.x * kZThreads + threadIdx.x; i0 < hi; i0 += U * stride) {
      float4 v[U];
#pragma unroll
      for (int u = 0; u < U; ++u)
        if (i0 + u * stride < hi) v[u] = *reinterpret_cast<const float4*>(mine + (i0 + u * stride) * 16);
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const long long i = i0 + u * stride;
        if (i >= hi) continue;
        if constexpr (MMST) {
          mm_st4(MC.param + byte_off + i * 16, v[u].x, v[u].y, v[u].z, v[u].w);
        } else {
#pragma unroll
          for (int r = 0; r < kMaxRanksD; ++r)
            if (r < ws && r != rank) stsys4f(P.param[r] + byte_off + i * 16, v[u].x, v[u].y, v[u].z, v[u].w);
        }
      }
    }
  }
  zbarrier(P, rank, ws, slot);
}

__global__ void zero_incr_step_kernel(int* step) { *step += 1; }

int adamw_tu_status(CommStatus* out) { return comm_status_read(out); }
int adamw_tu_reset(unsigned long long timeout_ns) { return comm_status_reset_and_set_timeout(timeout_ns); }

static void load_zpeers(ZPeers* P, const unsigned long long* g, const unsigned long long* pa, const unsigned long long* sh,
                        const unsigned long long* sig, int ws) {
  for (int r = 0; r < kMaxRanksD; ++r) {
    const bool on = r < ws;
    P->grad[r] = (on && g) ? reinterpret_cast<char*>(g[r]) : nullptr;
    P->param[r] = (on && pa) ? reinterpret_cast<char*>(pa[r]) : nullptr;
    P->shadow[r] = (on && sh) ? reinterpret_cast<char*>(sh[r]) : nullptr;
    P->sig[r] = (on && sig) ? reinterpret_cast<unsigned int*>(sig[r]) : nullptr;
  }
}

}  // namespace comm
}  // namespace b200

using namespace b200;

extern "C" int b200_zero_adamw_bucket(const unsigned long long* grad_peers_host, const unsigned long long* param_peers_host,
                                      const unsigned long long* shadow_peers_host /*nullable*/,
                                      const unsigned long long* signal_ptrs_host, const unsigned long long* multicast3_host,
                                      long long byte_off, long long numel, float* m_slice, float* v_slice,
                                      const float* hyper_dev, const int* step_dev, int mode, int rank, int world_size,
                                      int blocks, int signal_slot, int variant, int defer_fp32, int n_eager,
                                      const long long* eager_lo_elems, const long long* eager_hi_elems, void* stream) {
  using namespace b200::comm;
  B200_REQUIRE(world_size >= 1 && world_size <= kMaxRanksD, "zero_adamw: world size %d unsupported", world_size);
  B200_REQUIRE(byte_off % 32 == 0, "zero_adamw: bucket offset must be 32-byte aligned");
  B200_REQUIRE(n_eager >= 0 && n_eager <= kZEager, "zero_adamw: at most %d fp32-eager ranges per bucket (got %d)", kZEager, n_eager);
  B200_REQUIRE(!defer_fp32 || shadow_peers_host, "zero_adamw: a deferred fp32 gather needs the bf16 shadows on the wire");
  if (numel <= 0) return 0;
  ZPeers P;
  load_zpeers(&P, grad_peers_host, param_peers_host, shadow_peers_host, signal_ptrs_host, world_size);
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
  ZEager E;
  E.n = n_eager;
  for (int k = 0; k < kZEager; ++k) {
    // element ranges widened to whole 32-byte vectors (the kernels' decision granularity), in 16-byte units
    E.lo[k] = k < n_eager ? (eager_lo_elems[k] / 8) * 2 : 0;
    E.hi[k] = k < n_eager ? ((eager_hi_elems[k] + 7) / 8) * 2 : 0;
  }
  const int flags = defer_fp32 ? 0 : 1;
  const long long nvec32 = (numel * 4 + 31) / 32;
  if (blocks <= 0) blocks = kZMaxBlocks;
  if (blocks > kZMaxBlocks) blocks = kZMaxBlocks;
  const long long want = (nvec32 / world_size + 2 * kZThreads - 1) / (2 * kZThreads);
  if (want < blocks) blocks = (int)(want < 1 ? 1 : want);
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
#define ZARGS P, MC, byte_off, nvec32, m_slice, v_slice, hyper_dev, step_dev, mode, rank, world_size, signal_slot, flags, E
#define ZLAUNCH(WSV, LD, ST) zero_adamw_kernel<WSV, LD, ST><<<blocks, kZThreads, 0, s>>>(ZARGS)
  if (mmld) {
    if (mmst) zero_adamw_mm_kernel<true><<<blocks, kZThreads, 0, s>>>(ZARGS);
    else zero_adamw_mm_kernel<false><<<blocks, kZThreads, 0, s>>>(ZARGS);
  } else if (world_size <= 2) {
    if (mmst) ZLAUNCH(2, false, true); else ZLAUNCH(2, false, false);
  } else if (world_size <= 4) {
    if (mmst) ZLAUNCH(4, false, true); else ZLAUNCH(4, false, false);
  } else {
    if (mmst) ZLAUNCH(8, false, true); else ZLAUNCH(8, false, false);
  }
#undef ZLAUNCH
#undef ZARGS
  B200_LAUNCH_CHECK();
  return 0;
}

// The deferred half of b200_zero_adamw_bucket(defer_fp32 = 1): one launch over all buckets.
// buckets_dev: n_buckets x {byte_off, ceil(numel*4/32)} (int64 pairs, device memory).
extern "C" int b200_zero_param_gather(const unsigned long long* param_peers_host, const unsigned long long* signal_ptrs_host,
                                      unsigned long long param_multicast /*0: peer stores*/, const long long* buckets_dev,
                                      int n_buckets, int rank, int world_size, int blocks, int signal_slot, void* stream) {
  using namespace b200::comm;
  B200_REQUIRE(world_size >= 1 && world_size <= kMaxRanksD, "zero_param_gather: world size %d unsupported", world_size);
  if (n_buckets <= 0) return 0;
  ZPeers P;
  load_zpeers(&P, nullptr, param_peers_host, nullptr, signal_ptrs_host, world_size);
  ZMulticast MC{nullptr, reinterpret_cast<char*>(param_multicast), nullptr};
  if (blocks <= 0 || blocks > kZMaxBlocks) blocks = kZMaxBlocks;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  const ZBucket* bk = reinterpret_cast<const ZBucket*>(buckets_dev);
  if (param_multicast) zero_param_gather_kernel<true><<<blocks, kZThreads, 0, s>>>(P, MC, bk, n_buckets, rank, world_size, signal_slot);
  else zero_param_gather_kernel<false><<<blocks, kZThreads, 0, s>>>(P, MC, bk, n_buckets, rank, world_size, signal_slot);
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
