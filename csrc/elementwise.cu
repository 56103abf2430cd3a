// Memory-bound fused ops of the LM: RoPE (in place on the packed QKV buffer), SwiGLU gate,
// embedding gather / scatter-add, fp32→bf16 weight-shadow cast.
//
// Reference counterparts (all chains of ATen kernels): RoPE model.py:135-147, SwiGLU pointwise
// model.py:397,526-527, embedding model.py:57, autocast weight casts (implicit in torch.autocast).
#include "common.cuh"
#include "ptx.cuh"

namespace b200 {

// ---- RoPE ------------------------------------------------------------------------------------
// qkv: [B, S, 3, H, dh] bf16 (the fused QKV GEMM output viewed as heads); rotates the q and k parts
// in place.  Pair (2i, 2i+1) of each head is rotated by angle pos * theta^(-2i/dh); cos/sin come from
// the model's (2, ctx, dh/2) fp32 cache.  `inverse` applies the transposed rotation (backward).
__global__ void rope_qk_kernel(__nv_bfloat16* __restrict__ qkv, const float* __restrict__ cache, int ctx, int S,
                               int H, int dh, int n_rot /*2: q and k*/, float sign, size_t total_pairs) {
  const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= total_pairs) return;
  const int half = dh >> 1;
  const int i = (int)(idx % half);
  size_t t = idx / half;
  const int h = (int)(t % H); t /= H;
  const int which = (int)(t % n_rot); t /= n_rot;   // 0 = q, 1 = k
  const int s = (int)(t % S);
  const size_t b = t / S;
  const size_t off = ((((b * S + s) * 3 + which) * H + h) * (size_t)dh) + 2 * i;
  const float c = cache[(size_t)s * half + i];
  const float sn = sign * cache[(size_t)ctx * half + (size_t)s * half + i];
  const uint32_t raw = *reinterpret_cast<const uint32_t*>(qkv + off);
  const float2 v = unpack_bf16x2(raw);
  *reinterpret_cast<uint32_t*>(qkv + off) = pack_bf16x2(c * v.x - sn * v.y, sn * v.x + c * v.y);
}

// ---- SwiGLU ----------------------------------------------------------------------------------
// ab: [T, 2F] bf16 with a = ab[:, :F] (gate pre-activation), b = ab[:, F:] (up); g = silu(a) * b
__global__ void swiglu_fwd_kernel(const __nv_bfloat16* __restrict__ ab, __nv_bfloat16* __restrict__ g, int F,
                                  size_t total_vec /* T*F/8 */) {
  ptx::griddep_launch();
  ptx::griddep_wait();
  const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= total_vec) return;
  const int fv = F / 8;
  const size_t t = idx / fv;
  const int j = (int)(idx % fv) * 8;
  const u32x4 av = *reinterpret_cast<const u32x4*>(ab + t * 2 * F + j);
  const u32x4 bv = *reinterpret_cast<const u32x4*>(ab + t * 2 * F + F + j);
  const uint32_t ar[4] = {av.x, av.y, av.z, av.w}, br[4] = {bv.x, bv.y, bv.z, bv.w};
  uint32_t o[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const float2 a = unpack_bf16x2(ar[k]), b = unpack_bf16x2(br[k]);
    o[k] = pack_bf16x2(a.x / (1.f + __expf(-a.x)) * b.x, a.y / (1.f + __expf(-a.y)) * b.y);
  }
  *reinterpret_cast<u32x4*>(g + t * F + j) = u32x4{o[0], o[1], o[2], o[3]};
}

// dab[:, :F] = dg * b * (sig + a*sig*(1-sig)),  dab[:, F:] = dg * a * sig
__global__ void swiglu_bwd_kernel(const __nv_bfloat16* __restrict__ ab, const __nv_bfloat16* __restrict__ dg,
                                  __nv_bfloat16* __restrict__ dab, int F, size_t total_vec) {
  ptx::griddep_launch();
  ptx::griddep_wait();
  const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= total_vec) return;
  const int fv = F / 8;
  const size_t t = idx / fv;
  const int j = (int)(idx % fv) * 8;
  const u32x4 av = *reinterpret_cast<const u32x4*>(ab + t * 2 * F + j);
  const u32x4 bv = *reinterpret_cast<const u32x4*>(ab + t * 2 * F + F + j);
  const u32x4 gv = *reinterpret_cast<const u32x4*>(dg + t * F + j);
  const uint32_t ar[4] = {av.x, av.y, av.z, av.w}, br[4] = {bv.x, bv.y, bv.z, bv.w}, gr[4] = {gv.x, gv.y, gv.z, gv.w};
  uint32_t oa[4], ob[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const float2 a = unpack_bf16x2(ar[k]), b = unpack_bf16x2(br[k]), d = unpack_bf16x2(gr[k]);
    const float s0 = 1.f / (1.f + __expf(-a.x)), s1 = 1.f / (1.f + __expf(-a.y));
    oa[k] = pack_bf16x2(d.x * b.x * (s0 + a.x * s0 * (1.f - s0)), d.y * b.y * (s1 + a.y * s1 * (1.f - s1)));
    ob[k] = pack_bf16x2(d.x * a.x * s0, d.y * a.y * s1);
  }
  *reinterpret_cast<u32x4*>(dab + t * 2 * F + j) = u32x4{oa[0], oa[1], oa[2], oa[3]};
  *reinterpret_cast<u32x4*>(dab + t * 2 * F + F + j) = u32x4{ob[0], ob[1], ob[2], ob[3]};
}

// ---- embedding -------------------------------------------------------------------------------
__global__ void embedding_fwd_kernel(const long long* __restrict__ ids, const float* __restrict__ w,
                                     float* __restrict__ out, int d, int vocab) {
  const long long id = ids[blockIdx.x];
  const size_t src = (size_t)(id < 0 ? 0 : (id >= vocab ? vocab - 1 : id)) * d;
  const size_t dst = (size_t)blockIdx.x * d;
  if ((d & 3) == 0) {
    for (int i = threadIdx.x; i < d / 4; i += blockDim.x)
      reinterpret_cast<f32x4*>(out + dst)[i] = reinterpret_cast<const f32x4*>(w + src)[i];
  } else {
    for (int i = threadIdx.x; i < d; i += blockDim.x) out[dst + i] = w[src + i];
  }
}
// dW must be zero (or hold a gradient to accumulate into); rows hit several times add up atomically
__global__ void embedding_bwd_kernel(const long long* __restrict__ ids, const float* __restrict__ dout,
                                     float* __restrict__ dw, int d, int vocab) {
  const long long id = ids[blockIdx.x];
  if (id < 0 || id >= vocab) return;
  const size_t dst = (size_t)id * d, src = (size_t)blockIdx.x * d;
  for (int i = threadIdx.x; i < d; i += blockDim.x) atomicAdd(dw + dst + i, dout[src + i]);
}

// ---- casts -----------------------------------------------------------------------------------
__global__ void cast_f32_bf16_kernel(const float* __restrict__ in, __nv_bfloat16* __restrict__ out, size_t n,
                                     bool aligned) {
  const size_t i8 = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * 8;
  if (i8 >= n) return;
  if (aligned && i8 + 7 < n) {
    const f32x4 a = *reinterpret_cast<const f32x4*>(in + i8), b = *reinterpret_cast<const f32x4*>(in + i8 + 4);
    *reinterpret_cast<u32x4*>(out + i8) =
        u32x4{pack_bf16x2(a.x, a.y), pack_bf16x2(a.z, a.w), pack_bf16x2(b.x, b.y), pack_bf16x2(b.z, b.w)};
  } else {
    for (size_t i = i8; i < n && i < i8 + 8; ++i) out[i] = __float2bfloat16(in[i]);
  }
}

// up to three fp32 -> bf16 casts in one launch (blockIdx.y picks the tensor): the q, k, v of an fp32 attention call
struct Cast3 {
  const float* in[3];
  __nv_bfloat16* out[3];
  size_t n[3];
  bool aligned[3];
};
__global__ void cast3_f32_bf16_kernel(Cast3 c) {
  const int t = blockIdx.y;
  const float* __restrict__ in = c.in[t];
  __nv_bfloat16* __restrict__ out = c.out[t];
  const size_t n = c.n[t];
  for (size_t i8 = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * 8; i8 < n; i8 += (size_t)gridDim.x * blockDim.x * 8) {
    if (c.aligned[t] && i8 + 7 < n) {
      const f32x4 a = *reinterpret_cast<const f32x4*>(in + i8), b = *reinterpret_cast<const f32x4*>(in + i8 + 4);
      *reinterpret_cast<u32x4*>(out + i8) =
          u32x4{pack_bf16x2(a.x, a.y), pack_bf16x2(a.z, a.w), pack_bf16x2(b.x, b.y), pack_bf16x2(b.z, b.w)};
    } else {
      for (size_t i = i8; i < n && i < i8 + 8; ++i) out[i] = __float2bfloat16(in[i]);
    }
  }
}

}  // namespace b200

using namespace b200;

extern "C" int b200_cast3_f32_to_bf16(const float* in0, void* out0, long long n0, const float* in1, void* out1,
                                      long long n1, const float* in2, void* out2, long long n2, void* stream) {
  Cast3 c;
  const float* ins[3] = {in0, in1, in2};
  void* outs[3] = {out0, out1, out2};
  const long long ns[3] = {n0, n1, n2};
  long long nmax = 0;
  int cnt = 0;
  for (int t = 0; t < 3; ++t) {
    if (ns[t] <= 0 || !ins[t]) break;
    c.in[t] = ins[t];
    c.out[t] = reinterpret_cast<__nv_bfloat16*>(outs[t]);
    c.n[t] = (size_t)ns[t];
    c.aligned[t] = ((reinterpret_cast<uintptr_t>(ins[t]) | reinterpret_cast<uintptr_t>(outs[t])) & 15) == 0;
    nmax = ns[t] > nmax ? ns[t] : nmax;
    ++cnt;
  }
  if (cnt == 0) return 0;
  long long blocks = ((nmax + 7) / 8 + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  cast3_f32_bf16_kernel<<<dim3((unsigned)blocks, (unsigned)cnt), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(c);
  B200_LAUNCH_CHECK();
  return 0;
}

extern "C" int b200_rope_qk_inplace(void* qkv_bf16, const float* cache, int ctx, int B, int S, int H, int dh,
                                    int inverse, void* stream) {
  B200_REQUIRE(dh % 2 == 0, "rope: head dim must be even");
  B200_REQUIRE(S <= ctx, "rope: sequence length %d exceeds the cache's context length %d", S, ctx);
  const size_t total = (size_t)B * S * 2 * H * (dh / 2);
  if (total == 0) return 0;
  rope_qk_kernel<<<(unsigned)((total + 255) / 256), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<__nv_bfloat16*>(qkv_bf16), cache, ctx, S, H, dh, 2, inverse ? -1.f : 1.f, total);
  B200_LAUNCH_CHECK();
  return 0;
}

extern "C" int b200_swiglu_fwd(const void* ab, void* g, long long T, int F, void* stream) {
  B200_REQUIRE(F % 8 == 0, "swiglu: d_ff must be a multiple of 8");
  const size_t total = (size_t)T * F / 8;
  if (total == 0) return 0;
  B200_CHECK_CUDA(launch_pdl(swiglu_fwd_kernel, dim3((unsigned)((total + 255) / 256)), dim3(256), 0,
                             reinterpret_cast<cudaStream_t>(stream), reinterpret_cast<const __nv_bfloat16*>(ab),
                             reinterpret_cast<__nv_bfloat16*>(g), F, total));
  return 0;
}

extern "C" int b200_swiglu_bwd(const void* ab, const void* dg, void* dab, long long T, int F, void* stream) {
  B200_REQUIRE(F % 8 == 0, "swiglu: d_ff must be a multiple of 8");
  const size_t total = (size_t)T * F / 8;
  if (total == 0) return 0;
  B200_CHECK_CUDA(launch_pdl(swiglu_bwd_kernel, dim3((unsigned)((total + 255) / 256)), dim3(256), 0,
                             reinterpret_cast<cudaStream_t>(stream), reinterpret_cast<const __nv_bfloat16*>(ab),
                             reinterpret_cast<const __nv_bfloat16*>(dg), reinterpret_cast<__nv_bfloat16*>(dab), F, total));
  return 0;
}

extern "C" int b200_embedding_fwd(const long long* ids, const float* w, float* out, long long T, int d, int vocab,
                                  void* stream) {
  if (T == 0) return 0;
  embedding_fwd_kernel<<<(unsigned)T, 128, 0, reinterpret_cast<cudaStream_t>(stream)>>>(ids, w, out, d, vocab);
  B200_LAUNCH_CHECK();
  return 0;
}

extern "C" int b200_embedding_bwd(const long long* ids, const float* dout, float* dw, long long T, int d, int vocab,
                                  int zero_first, void* stream) {
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  if (zero_first) B200_CHECK_CUDA(cudaMemsetAsync(dw, 0, (size_t)vocab * d * sizeof(float), s));
  if (T == 0) return 0;
  embedding_bwd_kernel<<<(unsigned)T, 128, 0, s>>>(ids, dout, dw, d, vocab);
  B200_LAUNCH_CHECK();
  return 0;
}

extern "C" int b200_cast_f32_to_bf16(const float* in, void* out, long long n, void* stream) {
  if (n <= 0) return 0;
  const long long nvec = (n + 7) / 8;
  cast_f32_bf16_kernel<<<(unsigned)((nvec + 255) / 256), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      in, reinterpret_cast<__nv_bfloat16*>(out), (size_t)n,
      ((reinterpret_cast<uintptr_t>(in) | reinterpret_cast<uintptr_t>(out)) & 15) == 0);
  B200_LAUNCH_CHECK();
  return 0;
}
