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

using namespace b200;

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
