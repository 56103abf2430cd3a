// Multi-tensor AdamW and gradient-norm clipping: ONE launch over every parameter of the model.
//
// The reference's optimizer runs ≈12 ATen kernels per parameter tensor from a Python loop
// (cs336_basics/optimizer.py:50-86; 435 tensors for the XL model) and clip_gradient another
// two per tensor (nn_utils.py:20-30).  Here a device-resident table of tensor descriptors drives a
// single grid; step counter and hyper-parameters live in device memory so the launch is
// CUDA-graph replayable and an LR schedule is a 4-byte async copy.  The same pass refreshes the
// bf16 weight shadows read by the tcgen05 GEMMs (no autocast cast kernels in the step).
#include "common.cuh"

namespace b200 {

constexpr int kChunk = 4096;  // elements per block-iteration (256 threads × 4 × 4)

__device__ __forceinline__ int find_tensor(const long long* __restrict__ chunk_start, int n_tensors, long long chunk) {
  int lo = 0, hi = n_tensors - 1;  // last i with chunk_start[i] <= chunk
  while (lo < hi) {
    const int mid = (lo + hi + 1) >> 1;
    if (chunk_start[mid] <= chunk) lo = mid; else hi = mid - 1;
  }
  return lo;
}

// 256-bit streaming accesses with L2 evict-first: optimizer traffic (30 B/param, touched once per step) must not
// push the GEMM operands of a concurrently running backward pass out of the 126 MB L2.
struct f32x8 { float v[8]; };
__device__ __forceinline__ f32x8 ld_stream8(const float* p) {
  f32x8 r;
  asm volatile("ld.global.L1::no_allocate.L2::evict_first.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=f"(r.v[0]), "=f"(r.v[1]), "=f"(r.v[2]), "=f"(r.v[3]), "=f"(r.v[4]), "=f"(r.v[5]), "=f"(r.v[6]), "=f"(r.v[7])
               : "l"(p));
  return r;
}
__device__ __forceinline__ void st_stream8(float* p, const f32x8& r) {
  asm volatile("st.global.L1::no_allocate.L2::evict_first.v8.f32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p),
               "f"(r.v[0]), "f"(r.v[1]), "f"(r.v[2]), "f"(r.v[3]), "f"(r.v[4]), "f"(r.v[5]), "f"(r.v[6]), "f"(r.v[7])
               : "memory");
}

// hyper: [lr, beta1, beta2, eps, weight_decay, grad_scale]; step: int (t of this update, starts at 1)
// mode 0: course AdamW (alpha_t = lr*sqrt(1-b2^t)/(1-b1^t); p -= alpha_t*m/(sqrt(v)+eps); p -= lr*wd*p)
// mode 1: torch.optim.AdamW (p *= 1-lr*wd; p -= lr/bc1 * m/(sqrt(v)/sqrt(bc2)+eps))
__global__ void __launch_bounds__(256) adamw_multi_kernel(const b200_adamw_tensor* __restrict__ table,
                                                          const long long* __restrict__ chunk_start, int n_tensors,
                                                          long long total_chunks, const float* __restrict__ hyper,
                                                          const int* __restrict__ step, int mode) {
  const float lr = hyper[0], b1 = hyper[1], b2 = hyper[2], eps = hyper[3], wd = hyper[4], gscale = hyper[5];
  const float t = (float)(*step);
  const float bc1 = 1.f - powf(b1, t), bc2 = 1.f - powf(b2, t);
  const float alpha_t = lr * sqrtf(bc2) / bc1;
  const float rsqrt_bc2 = rsqrtf(bc2);
  for (long long chunk = blockIdx.x; chunk < total_chunks; chunk += gridDim.x) {
    const int ti = find_tensor(chunk_start, n_tensors, chunk);
    const b200_adamw_tensor td = table[ti];
    const long long off0 = (chunk - chunk_start[ti]) * kChunk;
    const bool vec = td.aligned16 != 0;
    if (td.aligned16 >= 2 && off0 + kChunk <= td.n) {
      // full chunk, every pointer 32-byte aligned: 2 x (256 threads x 8 elements), all loads issued up front
      f32x8 P[2], G[2], M[2], V[2];
#pragma unroll
      for (int it = 0; it < 2; ++it) {
        const long long i = off0 + ((long long)it * 256 + threadIdx.x) * 8;
        P[it] = ld_stream8(td.p + i); G[it] = ld_stream8(td.g + i);
        M[it] = ld_stream8(td.m + i); V[it] = ld_stream8(td.v + i);
      }
#pragma unroll
      for (int it = 0; it < 2; ++it) {
        const long long i = off0 + ((long long)it * 256 + threadIdx.x) * 8;
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const float gk = G[it].v[k] * gscale;
          float mk = b1 * M[it].v[k] + (1.f - b1) * gk;
          float vk = b2 * V[it].v[k] + (1.f - b2) * gk * gk;
          float pk = P[it].v[k];
          if (mode == 0) {
            pk -= alpha_t * mk / (sqrtf(vk) + eps);
            pk -= lr * wd * pk;
          } else {
            pk *= (1.f - lr * wd);
            pk -= (lr / bc1) * mk / (sqrtf(vk) * rsqrt_bc2 + eps);
          }
          M[it].v[k] = mk; V[it].v[k] = vk; P[it].v[k] = pk;
        }
        st_stream8(td.p + i, P[it]); st_stream8(td.m + i, M[it]); st_stream8(td.v + i, V[it]);
        if (td.shadow_bf16) {
          u32x4 o{pack_bf16x2(P[it].v[0], P[it].v[1]), pack_bf16x2(P[it].v[2], P[it].v[3]),
                  pack_bf16x2(P[it].v[4], P[it].v[5]), pack_bf16x2(P[it].v[6], P[it].v[7])};
          *reinterpret_cast<u32x4*>(reinterpret_cast<__nv_bfloat16*>(td.shadow_bf16) + i) = o;
        }
      }
      continue;
    }
#pragma unroll
    for (int it = 0; it < 4; ++it) {
      const long long i = off0 + ((long long)it * 256 + threadIdx.x) * 4;
      if (i >= td.n) break;
      float p[4], g[4], m[4], v[4];
      const int cnt = (td.n - i) >= 4 ? 4 : (int)(td.n - i);
      if (vec && cnt == 4) {
        const f32x4 pv = *reinterpret_cast<const f32x4*>(td.p + i), gv = *reinterpret_cast<const f32x4*>(td.g + i);
        const f32x4 mv = *reinterpret_cast<const f32x4*>(td.m + i), vv = *reinterpret_cast<const f32x4*>(td.v + i);
        p[0] = pv.x; p[1] = pv.y; p[2] = pv.z; p[3] = pv.w;
        g[0] = gv.x; g[1] = gv.y; g[2] = gv.z; g[3] = gv.w;
        m[0] = mv.x; m[1] = mv.y; m[2] = mv.z; m[3] = mv.w;
        v[0] = vv.x; v[1] = vv.y; v[2] = vv.z; v[3] = vv.w;
      } else {
        for (int k = 0; k < 4; ++k) {
          const bool ok = k < cnt;
          p[k] = ok ? td.p[i + k] : 0.f; g[k] = ok ? td.g[i + k] : 0.f;
          m[k] = ok ? td.m[i + k] : 0.f; v[k] = ok ? td.v[i + k] : 0.f;
        }
      }
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const float gk = g[k] * gscale;
        m[k] = b1 * m[k] + (1.f - b1) * gk;
        v[k] = b2 * v[k] + (1.f - b2) * gk * gk;
        if (mode == 0) {
          p[k] -= alpha_t * m[k] / (sqrtf(v[k]) + eps);
          p[k] -= lr * wd * p[k];
        } else {
          p[k] *= (1.f - lr * wd);
          p[k] -= (lr / bc1) * m[k] / (sqrtf(v[k]) * rsqrt_bc2 + eps);
        }
      }
      if (vec && cnt == 4) {
        *reinterpret_cast<f32x4*>(td.p + i) = f32x4{p[0], p[1], p[2], p[3]};
        *reinterpret_cast<f32x4*>(td.m + i) = f32x4{m[0], m[1], m[2], m[3]};
        *reinterpret_cast<f32x4*>(td.v + i) = f32x4{v[0], v[1], v[2], v[3]};
        if (td.shadow_bf16) {
          uint2 o;
          o.x = pack_bf16x2(p[0], p[1]);
          o.y = pack_bf16x2(p[2], p[3]);
          *reinterpret_cast<uint2*>(reinterpret_cast<__nv_bfloat16*>(td.shadow_bf16) + i) = o;
        }
      } else {
        for (int k = 0; k < cnt; ++k) {
          td.p[i + k] = p[k]; td.m[i + k] = m[k]; td.v[i + k] = v[k];
          if (td.shadow_bf16) reinterpret_cast<__nv_bfloat16*>(td.shadow_bf16)[i + k] = __float2bfloat16(p[k]);
        }
      }
    }
  }
}

__global__ void incr_step_kernel(int* step) { *step += 1; }

// ---- gradient norm ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256) sumsq_multi_kernel(const b200_adamw_tensor* __restrict__ table,
                                                          const long long* __restrict__ chunk_start, int n_tensors,
                                                          long long total_chunks, float* __restrict__ partial) {
  __shared__ float red[32];
  float acc = 0.f;
  for (long long chunk = blockIdx.x; chunk < total_chunks; chunk += gridDim.x) {
    const int ti = find_tensor(chunk_start, n_tensors, chunk);
    const b200_adamw_tensor td = table[ti];
    const long long off0 = (chunk - chunk_start[ti]) * kChunk;
#pragma unroll
    for (int it = 0; it < 4; ++it) {
      const long long i = off0 + ((long long)it * 256 + threadIdx.x) * 4;
      if (i >= td.n) break;
      if (td.aligned16 && td.n - i >= 4) {
        const f32x4 gv = *reinterpret_cast<const f32x4*>(td.g + i);
        acc += gv.x * gv.x + gv.y * gv.y + gv.z * gv.z + gv.w * gv.w;
      } else {
        for (long long k = i; k < td.n && k < i + 4; ++k) acc += td.g[k] * td.g[k];
      }
    }
  }
  acc = block_sum(acc, red);
  if (threadIdx.x == 0) partial[blockIdx.x] = acc;
}

// norm_out[0] = sqrt(sum partial); norm_out[1] = clip coefficient min(1, max_norm/(norm+eps))
__global__ void finish_norm_kernel(const float* __restrict__ partial, int n, float max_norm, float eps,
                                   float* __restrict__ norm_out) {
  __shared__ float red[32];
  float s = 0.f;
  for (int i = threadIdx.x; i < n; i += blockDim.x) s += partial[i];
  s = block_sum(s, red);
  if (threadIdx.x == 0) {
    const float norm = sqrtf(s);
    norm_out[0] = norm;
    norm_out[1] = fminf(1.f, max_norm / (norm + eps));
  }
}

__global__ void __launch_bounds__(256) scale_multi_kernel(const b200_adamw_tensor* __restrict__ table,
                                                          const long long* __restrict__ chunk_start, int n_tensors,
                                                          long long total_chunks, const float* __restrict__ coef_ptr) {
  const float coef = *coef_ptr;
  if (coef == 1.f) return;
  for (long long chunk = blockIdx.x; chunk < total_chunks; chunk += gridDim.x) {
    const int ti = find_tensor(chunk_start, n_tensors, chunk);
    const b200_adamw_tensor td = table[ti];
    const long long off0 = (chunk - chunk_start[ti]) * kChunk;
    float* g = const_cast<float*>(td.g);
#pragma unroll
    for (int it = 0; it < 4; ++it) {
      const long long i = off0 + ((long long)it * 256 + threadIdx.x) * 4;
      if (i >= td.n) break;
      for (long long k = i; k < td.n && k < i + 4; ++k) g[k] *= coef;
    }
  }
}

static int grid_for(long long total_chunks) {
  const long long cap = (long long)num_sms() * 8;
  return (int)(total_chunks < cap ? total_chunks : cap);
}

}  // namespace b200

using namespace b200;

extern "C" int b200_adamw_chunk_elems(void) { return kChunk; }

extern "C" int b200_adamw_multi_grid(const b200_adamw_tensor* table_dev, const long long* chunk_start_dev, int n_tensors,
                                     long long total_chunks, const float* hyper_dev, int* step_dev, int mode,
                                     int increment_step, int max_blocks, void* stream) {
  if (n_tensors <= 0 || total_chunks <= 0) return 0;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  int grid = grid_for(total_chunks);
  if (max_blocks > 0 && grid > max_blocks) grid = max_blocks;
  adamw_multi_kernel<<<grid, 256, 0, s>>>(table_dev, chunk_start_dev, n_tensors, total_chunks, hyper_dev, step_dev, mode);
  B200_LAUNCH_CHECK();
  if (increment_step) {
    incr_step_kernel<<<1, 1, 0, s>>>(step_dev);
    B200_LAUNCH_CHECK();
  }
  return 0;
}

extern "C" int b200_adamw_multi(const b200_adamw_tensor* table_dev, const long long* chunk_start_dev, int n_tensors,
                                long long total_chunks, const float* hyper_dev, int* step_dev, int mode,
                                int increment_step, void* stream) {
  return b200_adamw_multi_grid(table_dev, chunk_start_dev, n_tensors, total_chunks, hyper_dev, step_dev, mode,
                               increment_step, 0, stream);
}

extern "C" int b200_gradnorm_partials(long long total_chunks) { return grid_for(total_chunks); }

// partial_dev: b200_gradnorm_partials(total_chunks) floats of scratch; norm_out_dev: 2 floats (norm, coef)
extern "C" int b200_clip_grad_norm(const b200_adamw_tensor* table_dev, const long long* chunk_start_dev, int n_tensors,
                                   long long total_chunks, float max_norm, float eps, float* partial_dev,
                                   float* norm_out_dev, int apply_scale, void* stream) {
  if (n_tensors <= 0 || total_chunks <= 0) return 0;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  const int grid = grid_for(total_chunks);
  sumsq_multi_kernel<<<grid, 256, 0, s>>>(table_dev, chunk_start_dev, n_tensors, total_chunks, partial_dev);
  B200_LAUNCH_CHECK();
  finish_norm_kernel<<<1, 1024, 0, s>>>(partial_dev, grid, max_norm, eps, norm_out_dev);
  B200_LAUNCH_CHECK();
  if (apply_scale) {
    scale_multi_kernel<<<grid, 256, 0, s>>>(table_dev, chunk_start_dev, n_tensors, total_chunks, norm_out_dev + 1);
    B200_LAUNCH_CHECK();
  }
  return 0;
}
