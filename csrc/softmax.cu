// Row softmax (forward and backward) with the scale and the causal mask fused in.
//
// Reference counterpart: softmax() of cs336_basics/nn_utils.py:4-7 (max, sub, exp, sum, div: five ATen kernels) as used
// by the materialised attention of model.py:400-432 (plus the where(mask, s, -inf) in front of it).  The LM itself
// runs FlashAttention; this kernel exists for the materialised-attention comparison arm of the attention benchmarks
// (benchmark_attention.py:113-201) and as the public ops.softmax on CUDA tensors.
//
// One CTA per row.  A thread owns 8-element chunks strided by the CTA width; up to kCache chunks per thread stay in
// registers between the max, the sum and the write pass (rows of up to 256·8·kCache = 16384 columns are read from
// memory exactly once); longer rows re-read (L2 hits).
#include "common.cuh"
#include "ptx.cuh"

namespace b200 {
namespace softmax {

constexpr int kThreads = 256;
constexpr int kCache = 8;

template <bool BF16>
__device__ __forceinline__ void load8(const void* base, size_t idx, int valid, float (&v)[8]) {
  if constexpr (BF16) {
    const __nv_bfloat16* p = reinterpret_cast<const __nv_bfloat16*>(base) + idx;
    if (valid == 8 && (reinterpret_cast<uintptr_t>(p) & 15) == 0) {
      const u32x4 u = *reinterpret_cast<const u32x4*>(p);
      const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
      for (int k = 0; k < 4; ++k) { const float2 f = unpack_bf16x2(w[k]); v[2 * k] = f.x; v[2 * k + 1] = f.y; }
    } else {
#pragma unroll
      for (int k = 0; k < 8; ++k) v[k] = k < valid ? __bfloat162float(p[k]) : 0.f;
    }
  } else {
    const float* p = reinterpret_cast<const float*>(base) + idx;
    if (valid == 8 && (reinterpret_cast<uintptr_t>(p) & 15) == 0) {
      const f32x4 a = *reinterpret_cast<const f32x4*>(p), b = *reinterpret_cast<const f32x4*>(p + 4);
      v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
    } else {
#pragma unroll
      for (int k = 0; k < 8; ++k) v[k] = k < valid ? p[k] : 0.f;
    }
  }
}
template <bool BF16>
__device__ __forceinline__ void store8(void* base, size_t idx, int valid, const float (&v)[8]) {
  if constexpr (BF16) {
    __nv_bfloat16* p = reinterpret_cast<__nv_bfloat16*>(base) + idx;
    if (valid == 8 && (reinterpret_cast<uintptr_t>(p) & 15) == 0) {
      *reinterpret_cast<u32x4*>(p) = u32x4{pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]), pack_bf16x2(v[4], v[5]),
                                           pack_bf16x2(v[6], v[7])};
    } else {
#pragma unroll
      for (int k = 0; k < 8; ++k)
        if (k < valid) p[k] = __float2bfloat16(v[k]);
    }
  } else {
    float* p = reinterpret_cast<float*>(base) + idx;
    if (valid == 8 && (reinterpret_cast<uintptr_t>(p) & 15) == 0) {
      *reinterpret_cast<f32x4*>(p) = f32x4{v[0], v[1], v[2], v[3]};
      *reinterpret_cast<f32x4*>(p + 4) = f32x4{v[4], v[5], v[6], v[7]};
    } else {
#pragma unroll
      for (int k = 0; k < 8; ++k)
        if (k < valid) p[k] = v[k];
    }
  }
}

// out[r, c] = softmax_c(scale * in[r, c]) over the columns c <= limit(r); masked columns get 0.
// causal_period > 0: row r is query position r % causal_period and sees columns 0..position (aligned self-attention).
template <bool IN_BF16, bool OUT_BF16>
__global__ void __launch_bounds__(kThreads) softmax_rows_kernel(const void* __restrict__ in, void* __restrict__ out,
                                                                int cols, long long ld_in, long long ld_out,
                                                                float scale, int causal_period) {
  __shared__ float red[32];
  const long long r = blockIdx.x;
  const int limit = causal_period > 0 ? min(cols, (int)(r % causal_period) + 1) : cols;  // visible columns
  const float sl2 = scale * 1.4426950408889634f;
  float cache[kCache][8];
  float mx = -INFINITY;
  int nchunk = 0;
  for (int c0 = threadIdx.x * 8; c0 < limit; c0 += kThreads * 8, ++nchunk) {
    float v[8];
    load8<IN_BF16>(in, (size_t)(r * ld_in + c0), min(8, cols - c0), v);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      v[k] = (c0 + k < limit) ? v[k] * sl2 : -INFINITY;
      mx = fmaxf(mx, v[k]);
    }
    if (nchunk < kCache) {
#pragma unroll
      for (int k = 0; k < 8; ++k) cache[nchunk][k] = v[k];
    }
  }
  mx = block_max(mx, red);
  float sum = 0.f;
  nchunk = 0;
  for (int c0 = threadIdx.x * 8; c0 < limit; c0 += kThreads * 8, ++nchunk) {
    float v[8];
    if (nchunk < kCache) {
#pragma unroll
      for (int k = 0; k < 8; ++k) v[k] = cache[nchunk][k];
    } else {
      load8<IN_BF16>(in, (size_t)(r * ld_in + c0), min(8, cols - c0), v);
#pragma unroll
      for (int k = 0; k < 8; ++k) v[k] = (c0 + k < limit) ? v[k] * sl2 : -INFINITY;
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      v[k] = exp2f(v[k] - mx);
      sum += v[k];
    }
    if (nchunk < kCache) {
#pragma unroll
      for (int k = 0; k < 8; ++k) cache[nchunk][k] = v[k];
    }
  }
  sum = block_sum(sum, red);
  const float inv = 1.f / sum;
  nchunk = 0;
  for (int c0 = threadIdx.x * 8; c0 < cols; c0 += kThreads * 8, ++nchunk) {
    float v[8];
    if (c0 >= limit) {
#pragma unroll
      for (int k = 0; k < 8; ++k) v[k] = 0.f;
    } else if (nchunk < kCache) {
#pragma unroll
      for (int k = 0; k < 8; ++k) v[k] = cache[nchunk][k] * inv;
    } else {
      load8<IN_BF16>(in, (size_t)(r * ld_in + c0), min(8, cols - c0), v);
#pragma unroll
      for (int k = 0; k < 8; ++k) v[k] = (c0 + k < limit) ? exp2f(v[k] * sl2 - mx) * inv : 0.f;
    }
    store8<OUT_BF16>(out, (size_t)(r * ld_out + c0), min(8, cols - c0), v);
  }
}

// ds[r, c] = scale * p[r, c] * (dp[r, c] - sum_c' dp[r, c'] p[r, c'])
template <bool BF16>
__global__ void __launch_bounds__(kThreads) softmax_bwd_rows_kernel(const void* __restrict__ p, const void* __restrict__ dp,
                                                                    void* __restrict__ ds, int cols, long long ld,
                                                                    float scale) {
  __shared__ float red[32];
  const long long r = blockIdx.x;
  float dot = 0.f;
  for (int c0 = threadIdx.x * 8; c0 < cols; c0 += kThreads * 8) {
    float a[8], b[8];
    load8<BF16>(p, (size_t)(r * ld + c0), min(8, cols - c0), a);
    load8<BF16>(dp, (size_t)(r * ld + c0), min(8, cols - c0), b);
#pragma unroll
    for (int k = 0; k < 8; ++k) dot += a[k] * b[k];
  }
  dot = block_sum(dot, red);
  for (int c0 = threadIdx.x * 8; c0 < cols; c0 += kThreads * 8) {
    float a[8], b[8];
    load8<BF16>(p, (size_t)(r * ld + c0), min(8, cols - c0), a);
    load8<BF16>(dp, (size_t)(r * ld + c0), min(8, cols - c0), b);
#pragma unroll
    for (int k = 0; k < 8; ++k) a[k] = scale * a[k] * (b[k] - dot);
    store8<BF16>(ds, (size_t)(r * ld + c0), min(8, cols - c0), a);
  }
}

}  // namespace softmax
}  // namespace b200

using namespace b200;

extern "C" int b200_softmax_rows(const void* in, int in_dtype, void* out, int out_dtype, long long rows, int cols,
                                 long long ld_in, long long ld_out, float scale, int causal_period, void* stream) {
  using namespace b200::softmax;
  B200_REQUIRE(cols > 0 && rows >= 0, "softmax: empty row");
  B200_REQUIRE((in_dtype == B200_F32 || in_dtype == B200_BF16) && (out_dtype == B200_F32 || out_dtype == B200_BF16),
               "softmax: dtypes must be fp32 or bf16");
  if (rows == 0) return 0;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  const dim3 grid((unsigned)rows);
  if (in_dtype == B200_BF16 && out_dtype == B200_BF16) softmax_rows_kernel<true, true><<<grid, kThreads, 0, s>>>(in, out, cols, ld_in, ld_out, scale, causal_period);
  else if (in_dtype == B200_BF16) softmax_rows_kernel<true, false><<<grid, kThreads, 0, s>>>(in, out, cols, ld_in, ld_out, scale, causal_period);
  else if (out_dtype == B200_BF16) softmax_rows_kernel<false, true><<<grid, kThreads, 0, s>>>(in, out, cols, ld_in, ld_out, scale, causal_period);
  else softmax_rows_kernel<false, false><<<grid, kThreads, 0, s>>>(in, out, cols, ld_in, ld_out, scale, causal_period);
  B200_LAUNCH_CHECK();
  return 0;
}

extern "C" int b200_softmax_bwd_rows(const void* p, const void* dp, void* ds, int dtype, long long rows, int cols,
                                     long long ld, float scale, void* stream) {
  using namespace b200::softmax;
  B200_REQUIRE(cols > 0 && rows >= 0, "softmax_bwd: empty row");
  B200_REQUIRE(dtype == B200_F32 || dtype == B200_BF16, "softmax_bwd: dtype must be fp32 or bf16");
  if (rows == 0) return 0;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  if (dtype == B200_BF16) softmax_bwd_rows_kernel<true><<<(unsigned)rows, kThreads, 0, s>>>(p, dp, ds, cols, ld, scale);
  else softmax_bwd_rows_kernel<false><<<(unsigned)rows, kThreads, 0, s>>>(p, dp, ds, cols, ld, scale);
  B200_LAUNCH_CHECK();
  return 0;
}
