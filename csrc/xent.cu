// Fused softmax cross-entropy: forward (online logsumexp + target gather + mean) and backward
// ((softmax − onehot)·dloss/T).  Replaces the ≈9 ATen kernels of the reference's cross_entropy
// (cs336_basics/nn_utils.py:9-17) over the [T, V] logits (V = 10000 for every course model).
#include "common.cuh"

namespace b200 {

template <typename T> __device__ __forceinline__ float ldf(const T* p);
template <> __device__ __forceinline__ float ldf<float>(const float* p) { return *p; }
template <> __device__ __forceinline__ float ldf<__nv_bfloat16>(const __nv_bfloat16* p) { return __bfloat162float(*p); }

// one CTA per row: running (max, sum) pair per thread, merged across the block
template <typename TL>
__global__ void __launch_bounds__(256) xent_fwd_kernel(const TL* __restrict__ logits, const long long* __restrict__ tgt,
                                                       float* __restrict__ row_loss, float* __restrict__ lse_out,
                                                       int V, int ld) {
  __shared__ float red[32];
  const TL* row = logits + (size_t)blockIdx.x * ld;
  float m = -INFINITY, s = 0.f;
  if (sizeof(TL) == 2 && (ld & 7) == 0) {
    const int nv = V / 8;
    for (int i = threadIdx.x; i < nv; i += blockDim.x) {
      const u32x4 v = *reinterpret_cast<const u32x4*>(reinterpret_cast<const __nv_bfloat16*>(row) + 8 * i);
      const uint32_t r[4] = {v.x, v.y, v.z, v.w};
      float x[8];
#pragma unroll
      for (int k = 0; k < 4; ++k) { const float2 f = unpack_bf16x2(r[k]); x[2 * k] = f.x; x[2 * k + 1] = f.y; }
      float mx = x[0];
#pragma unroll
      for (int k = 1; k < 8; ++k) mx = fmaxf(mx, x[k]);
      const float mn = fmaxf(m, mx);
      float acc = 0.f;
#pragma unroll
      for (int k = 0; k < 8; ++k) acc += __expf(x[k] - mn);
      s = s * __expf(m - mn) + acc;
      m = mn;
    }
    for (int i = nv * 8 + threadIdx.x; i < V; i += blockDim.x) {
      const float x = ldf<TL>(row + i);
      const float mn = fmaxf(m, x);
      s = s * __expf(m - mn) + __expf(x - mn);
      m = mn;
    }
  } else {
    for (int i = threadIdx.x; i < V; i += blockDim.x) {
      const float x = ldf<TL>(row + i);
      const float mn = fmaxf(m, x);
      s = s * __expf(m - mn) + __expf(x - mn);
      m = mn;
    }
  }
  const float gm = block_max(m, red);
  const float contrib = (m == -INFINITY) ? 0.f : s * __expf(m - gm);
  const float gs = block_sum(contrib, red);
  if (threadIdx.x == 0) {
    const float lse = gm + __logf(gs);
    const long long t = tgt[blockIdx.x];
    const float xt = (t >= 0 && t < V) ? ldf<TL>(row + t) : 0.f;
    lse_out[blockIdx.x] = lse;
    row_loss[blockIdx.x] = lse - xt;
  }
}

// deterministic mean of the per-row losses (single block)
__global__ void mean_kernel(const float* __restrict__ v, float* __restrict__ out, int n) {
  __shared__ float red[32];
  float s = 0.f;
  for (int i = threadIdx.x; i < n; i += blockDim.x) s += v[i];
  s = block_sum(s, red);
  if (threadIdx.x == 0) *out = s / (float)n;
}

// dlogits = (exp(logit − lse) − [col == target]) · (*dloss) / T
template <typename TL>
__global__ void __launch_bounds__(256) xent_bwd_kernel(const TL* __restrict__ logits, const long long* __restrict__ tgt,
                                                       const float* __restrict__ lse, const float* __restrict__ dloss,
                                                       TL* __restrict__ dlogits, int V, int ld, int ld_out, float inv_T) {
  const TL* row = logits + (size_t)blockIdx.x * ld;
  TL* out = dlogits + (size_t)blockIdx.x * ld_out;
  const float l = lse[blockIdx.x];
  const float g = (*dloss) * inv_T;
  const int t = (int)tgt[blockIdx.x];
  if (sizeof(TL) == 2 && (ld & 7) == 0 && (ld_out & 7) == 0) {
    const int nv = V / 8;
    for (int i = threadIdx.x; i < nv; i += blockDim.x) {
      const u32x4 v = *reinterpret_cast<const u32x4*>(reinterpret_cast<const __nv_bfloat16*>(row) + 8 * i);
      const uint32_t r[4] = {v.x, v.y, v.z, v.w};
      uint32_t o[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const float2 f = unpack_bf16x2(r[k]);
        const int c = 8 * i + 2 * k;
        o[k] = pack_bf16x2((__expf(f.x - l) - (c == t ? 1.f : 0.f)) * g, (__expf(f.y - l) - (c + 1 == t ? 1.f : 0.f)) * g);
      }
      *reinterpret_cast<u32x4*>(reinterpret_cast<__nv_bfloat16*>(out) + 8 * i) = u32x4{o[0], o[1], o[2], o[3]};
    }
    for (int i = nv * 8 + threadIdx.x; i < V; i += blockDim.x) {
      const float p = __expf(ldf<TL>(row + i) - l) - (i == t ? 1.f : 0.f);
      if constexpr (sizeof(TL) == 2) out[i] = __float2bfloat16(p * g); else out[i] = p * g;
    }
  } else {
    for (int i = threadIdx.x; i < V; i += blockDim.x) {
      const float p = __expf(ldf<TL>(row + i) - l) - (i == t ? 1.f : 0.f);
      if constexpr (sizeof(TL) == 2) out[i] = __float2bfloat16(p * g); else out[i] = p * g;
    }
  }
}

}  // namespace b200

using namespace b200;

// row_loss, lse: [T] fp32 scratch/outputs; loss: 1 fp32
extern "C" int b200_xent_fwd(const void* logits, int dtype, const long long* targets, float* row_loss, float* lse,
                             float* loss, int T, int V, int ld, void* stream) {
  B200_REQUIRE(T > 0 && V > 0, "xent_fwd: empty input");
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  if (dtype == B200_BF16)
    xent_fwd_kernel<__nv_bfloat16><<<T, 256, 0, s>>>(reinterpret_cast<const __nv_bfloat16*>(logits), targets, row_loss, lse, V, ld);
  else if (dtype == B200_F32)
    xent_fwd_kernel<float><<<T, 256, 0, s>>>(reinterpret_cast<const float*>(logits), targets, row_loss, lse, V, ld);
  else { set_error("xent_fwd: bad dtype %d", dtype); return B200_ERR_ARG; }
  B200_LAUNCH_CHECK();
  mean_kernel<<<1, 1024, 0, s>>>(row_loss, loss, T);
  B200_LAUNCH_CHECK();
  return 0;
}

extern "C" int b200_xent_bwd(const void* logits, int dtype, const long long* targets, const float* lse,
                             const float* dloss, void* dlogits, int T, int V, int ld, int ld_out, void* stream) {
  B200_REQUIRE(T > 0 && V > 0, "xent_bwd: empty input");
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  const float inv_T = 1.f / (float)T;
  if (dtype == B200_BF16)
    xent_bwd_kernel<__nv_bfloat16><<<T, 256, 0, s>>>(reinterpret_cast<const __nv_bfloat16*>(logits), targets, lse, dloss,
                                                     reinterpret_cast<__nv_bfloat16*>(dlogits), V, ld, ld_out, inv_T);
  else if (dtype == B200_F32)
    xent_bwd_kernel<float><<<T, 256, 0, s>>>(reinterpret_cast<const float*>(logits), targets, lse, dloss,
                                             reinterpret_cast<float*>(dlogits), V, ld, ld_out, inv_T);
  else { set_error("xent_bwd: bad dtype %d", dtype); return B200_ERR_ARG; }
  B200_LAUNCH_CHECK();
  return 0;
}
