// Shared between the 1-CTA and the CTA-pair GEMM kernels: epilogue modes, kernel parameters and the
// per-thread epilogue (one TMEM lane = one output row, 32 accumulator columns per call).
#pragma once
#include "common.cuh"

namespace b200 {
namespace gemm {

constexpr int BM = 128;
constexpr int BK = 64;  // 64 bf16 = one 128-byte swizzle row

enum Epilogue : int {
  EPI_BF16 = 0,      // C bf16 = acc
  EPI_F32 = 1,       // C fp32 = acc
  EPI_F32_ACC = 2,   // C fp32 += acc
  EPI_BF16_SWIGLU = 3,  // C bf16 [M, N/2] = silu(acc[:, 2j]) * acc[:, 2j+1]   (interleaved gate/up columns)
  EPI_F32_RESID = 4,    // C fp32 = aux fp32 + acc   (residual add)
  EPI_F32_T = 6,        // Cᵀ fp32 = acc: element (row, col) goes to C[col*ldc + row] — lane = row, so every
                        //   register is one fully coalesced 128-byte store (used to run a tall wgrad as its wide transpose)
  EPI_BF16_ROPE = 5,    // C bf16 = acc with RoPE applied to columns < rope_cols (fused QKV projection):
                        //   row r is token position r % rope_S, column c is element c % rope_dh of its head,
                        //   aux = fp32 cos/sin cache (2, rope_ctx, rope_dh/2)
};

struct Params {
  void* C;
  const void* aux;
  int M, N, K;
  int ldc, ld_aux;
  int num_m, num_n, num_k;
  int rope_S, rope_dh, rope_cols, rope_ctx;
};

template <int EPI>
__device__ __forceinline__ void store_chunk(const Params& p, int row, int col0, const uint32_t (&r)[32]) {
  // one thread owns one output row; r[] holds 32 consecutive fp32 accumulator columns
  if (row >= p.M || col0 >= p.N) return;
  const bool full = (col0 + 32 <= p.N);
  if constexpr (EPI == EPI_BF16_ROPE) {
    __nv_bfloat16* c = reinterpret_cast<__nv_bfloat16*>(p.C) + (size_t)row * p.ldc + col0;
    float v[32];
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
    if (col0 < p.rope_cols) {
      const float* cache = reinterpret_cast<const float*>(p.aux);
      const int half = p.rope_dh >> 1;
      const int pos = row % p.rope_S;
      const float* cosr = cache + (size_t)pos * half;
      const float* sinr = cache + (size_t)p.rope_ctx * half + (size_t)pos * half;
      if ((p.rope_dh & 31) == 0 && col0 + 32 <= p.rope_cols) {
        // the 32-column chunk lies inside one head: its 16 rotation angles are contiguous -> 8 vector loads up front
        const int i0 = (col0 % p.rope_dh) >> 1;
        f32x4 cs[4], sn[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          cs[q] = *reinterpret_cast<const f32x4*>(cosr + i0 + 4 * q);
          sn[q] = *reinterpret_cast<const f32x4*>(sinr + i0 + 4 * q);
        }
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const float c4[4] = {cs[q].x, cs[q].y, cs[q].z, cs[q].w}, s4[4] = {sn[q].x, sn[q].y, sn[q].z, sn[q].w};
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const int j = 4 * q + e;
            const float a = v[2 * j], b = v[2 * j + 1];
            v[2 * j] = c4[e] * a - s4[e] * b;
            v[2 * j + 1] = s4[e] * a + c4[e] * b;
          }
        }
      } else {
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          const int col = col0 + 2 * j;
          if (col < p.rope_cols) {
            const int i = (col % p.rope_dh) >> 1;
            const float c1 = cosr[i], s1 = sinr[i];
            const float a = v[2 * j], b = v[2 * j + 1];
            v[2 * j] = c1 * a - s1 * b;
            v[2 * j + 1] = s1 * a + c1 * b;
          }
        }
      }
    }
    if (full && (p.ldc & 7) == 0) {
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        u32x4 o;
        o.x = pack_bf16x2(v[8 * q + 0], v[8 * q + 1]);
        o.y = pack_bf16x2(v[8 * q + 2], v[8 * q + 3]);
        o.z = pack_bf16x2(v[8 * q + 4], v[8 * q + 5]);
        o.w = pack_bf16x2(v[8 * q + 6], v[8 * q + 7]);
        *reinterpret_cast<u32x4*>(c + 8 * q) = o;
      }
    } else {
#pragma unroll
      for (int j = 0; j < 32; ++j)
        if (col0 + j < p.N) c[j] = __float2bfloat16(v[j]);
    }
  } else if constexpr (EPI == EPI_BF16) {
    __nv_bfloat16* c = reinterpret_cast<__nv_bfloat16*>(p.C) + (size_t)row * p.ldc + col0;
    if (full && (p.ldc & 7) == 0) {
#pragma unroll
      for (int v = 0; v < 4; ++v) {
        u32x4 o;
        o.x = pack_bf16x2(__uint_as_float(r[8 * v + 0]), __uint_as_float(r[8 * v + 1]));
        o.y = pack_bf16x2(__uint_as_float(r[8 * v + 2]), __uint_as_float(r[8 * v + 3]));
        o.z = pack_bf16x2(__uint_as_float(r[8 * v + 4]), __uint_as_float(r[8 * v + 5]));
        o.w = pack_bf16x2(__uint_as_float(r[8 * v + 6]), __uint_as_float(r[8 * v + 7]));
        *reinterpret_cast<u32x4*>(c + 8 * v) = o;
      }
    } else {
#pragma unroll
      for (int j = 0; j < 32; ++j)
        if (col0 + j < p.N) c[j] = __float2bfloat16(__uint_as_float(r[j]));
    }
  } else if constexpr (EPI == EPI_F32 || EPI == EPI_F32_ACC || EPI == EPI_F32_RESID) {
    float* c = reinterpret_cast<float*>(p.C) + (size_t)row * p.ldc + col0;
    const float* x = nullptr;
    if constexpr (EPI == EPI_F32_ACC) x = c;
    if constexpr (EPI == EPI_F32_RESID) x = reinterpret_cast<const float*>(p.aux) + (size_t)row * p.ld_aux + col0;
    const bool vec = full && (p.ldc & 3) == 0 && (EPI != EPI_F32_RESID || (p.ld_aux & 3) == 0);
    if (vec) {
#pragma unroll
      for (int v = 0; v < 8; ++v) {
        f32x4 o;
        o.x = __uint_as_float(r[4 * v + 0]);
        o.y = __uint_as_float(r[4 * v + 1]);
        o.z = __uint_as_float(r[4 * v + 2]);
        o.w = __uint_as_float(r[4 * v + 3]);
        if constexpr (EPI != EPI_F32) {
          f32x4 old = *reinterpret_cast<const f32x4*>(x + 4 * v);
          o.x += old.x; o.y += old.y; o.z += old.z; o.w += old.w;
        }
        *reinterpret_cast<f32x4*>(c + 4 * v) = o;
      }
    } else {
#pragma unroll
      for (int j = 0; j < 32; ++j)
        if (col0 + j < p.N) {
          float v = __uint_as_float(r[j]);
          if constexpr (EPI != EPI_F32) v += x[j];
          c[j] = v;
        }
    }
  } else if constexpr (EPI == EPI_F32_T) {
    float* c = reinterpret_cast<float*>(p.C) + (size_t)col0 * p.ldc + row;
#pragma unroll
    for (int j = 0; j < 32; ++j)
      if (col0 + j < p.N) c[(size_t)j * p.ldc] = __uint_as_float(r[j]);
  } else if constexpr (EPI == EPI_BF16_SWIGLU) {
    // columns (2j, 2j+1) = (gate, up) of hidden unit (col0/2 + j); output has N/2 columns
    __nv_bfloat16* c = reinterpret_cast<__nv_bfloat16*>(p.C) + (size_t)row * p.ldc + (col0 >> 1);
    float g[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      const float a = __uint_as_float(r[2 * j]), b = __uint_as_float(r[2 * j + 1]);
      g[j] = a / (1.f + __expf(-a)) * b;
    }
    if (full && (p.ldc & 7) == 0) {
#pragma unroll
      for (int v = 0; v < 2; ++v) {
        u32x4 o;
        o.x = pack_bf16x2(g[8 * v + 0], g[8 * v + 1]);
        o.y = pack_bf16x2(g[8 * v + 2], g[8 * v + 3]);
        o.z = pack_bf16x2(g[8 * v + 4], g[8 * v + 5]);
        o.w = pack_bf16x2(g[8 * v + 6], g[8 * v + 7]);
        *reinterpret_cast<u32x4*>(c + 8 * v) = o;
      }
    } else {
#pragma unroll
      for (int j = 0; j < 16; ++j)
        if (col0 + 2 * j + 1 < p.N) c[j] = __float2bfloat16(g[j]);
    }
  }
}


}  // namespace gemm
}  // namespace b200
