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
  EPI_SWIGLU_FWD = 8,   // training forward of the SwiGLU up-projection (CTA-pair kernel, BN = 256): a tile holds the gate
                        //   pre-activations a of 128 hidden units in accumulator columns 0..127 and their up-projections b
                        //   in columns 128..255 (the B tile is fetched from w1 rows and w3 rows of the stacked weight);
                        //   C bf16 [M, F] = silu(a)*b, aux (written!) bf16 [M, 2F] = a | b kept for the backward.  N = 2F.
  EPI_SWIGLU_BWD = 9,   // dgrad GEMM of the down-projection with the gate's backward in its epilogue: acc = dg tile,
                        //   aux bf16 [M, 2F] = a | b, C bf16 [M, 2F] = da | db with da = dg*b*(s + a*s*(1-s)), db = dg*a*s,
                        //   s = sigmoid(a).  N = F (dg never reaches HBM).
  EPI_ADAMW = 7,        // wgrad GEMM whose epilogue IS the optimizer: acc = dW tile; reads p, m, v of the same elements,
                        //   applies AdamW and writes p, m, v and the bf16 GEMM shadow — the gradient never reaches HBM
                        //   (C, when non-null, additionally receives it as fp32, for checks)
};

// Optimizer operands of EPI_ADAMW.  The GEMM's M rows are the rows of up to three parameters stacked in the order of
// their fused forward GEMM (q|k|v, w1|w3): segment i covers rows [row0[i], row0[i+1]) and has its own p / m / v
// (each a contiguous [rows_i, N] fp32 tensor); the bf16 shadow is the stacked [M, N] operand of the forward GEMM.
struct AdamwOperands {
  float* p[3];
  float* m[3];
  float* v[3];
  int row0[4];
  int nseg;
  int mode;               // 0: course AdamW, 1: torch.optim.AdamW (same arithmetic as csrc/adamw.cu)
  __nv_bfloat16* shadow;  // nullable
  const float* hyper;     // device: lr, beta1, beta2, eps, weight_decay, grad_scale
  const int* step;        // device: t of this update
};

struct Params {
  void* C;
  const void* aux;
  int M, N, K;
  int ldc, ld_aux;
  int num_m, num_n, num_k;
  int rope_S, rope_dh, rope_cols, rope_ctx;
  AdamwOperands opt;  // EPI_ADAMW only
};

// per-thread constants of the AdamW epilogue (computed once per kernel by each epilogue thread)
struct AdamwConsts {
  float lr, b1, b2, eps, wd, gscale, alpha_t, bc1, rsqrt_bc2;
};
__device__ __forceinline__ AdamwConsts adamw_consts(const AdamwOperands& o) {
  AdamwConsts c;
  c.lr = o.hyper[0]; c.b1 = o.hyper[1]; c.b2 = o.hyper[2]; c.eps = o.hyper[3]; c.wd = o.hyper[4]; c.gscale = o.hyper[5];
  const float t = (float)(*o.step);
  c.bc1 = 1.f - powf(c.b1, t);
  const float bc2 = 1.f - powf(c.b2, t);
  c.alpha_t = c.lr * sqrtf(bc2) / c.bc1;
  c.rsqrt_bc2 = rsqrtf(bc2);
  return c;
}

// AdamW epilogue, transposed orientation.  The GEMM computes dWᵀ (rows = input features, columns = the stacked
// parameter rows), so a thread (= one TMEM lane = one dWᵀ row) holds 32 consecutive dW ROWS of one dW column: for a
// fixed register the 32 lanes of the warp touch 32 consecutive floats of p / m / v — every load and store is one
// fully coalesced 128-byte request.  (The row-per-thread orientation costs 32 L1 wavefronts per 16-byte vector
// instruction: measured 2x slower than running the optimizer as a separate kernel.)
//   p.M = input features (columns of dW), p.N = stacked parameter rows, parameters are contiguous [rows_i, p.M].
__device__ __forceinline__ void adamw_chunk_t(const Params& p, const AdamwConsts& k, int row, int col0,
                                              const uint32_t (&r)[32]) {
  if (row >= p.M || col0 >= p.N) return;
  const AdamwOperands& o = p.opt;
  const size_t ld = (size_t)p.M;
#pragma unroll
  for (int h = 0; h < 2; ++h) {  // halves of 16 registers: 48 independent loads in flight, then the math
    float P[16], M[16], V[16];
    uint32_t off[16];  // element offset inside one parameter (< 2^32)
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      const int dwrow = col0 + 16 * h + j;
      int s = 0;
      if (o.nseg > 1 && dwrow >= o.row0[1]) s = 1;
      if (o.nseg > 2 && dwrow >= o.row0[2]) s = 2;
      // selects, not dynamic indexing: kernel parameters live in the constant bank
      const int r0 = s == 0 ? o.row0[0] : (s == 1 ? o.row0[1] : o.row0[2]);
      const float* pb = s == 0 ? o.p[0] : (s == 1 ? o.p[1] : o.p[2]);
      const float* mb = s == 0 ? o.m[0] : (s == 1 ? o.m[1] : o.m[2]);
      const float* vb = s == 0 ? o.v[0] : (s == 1 ? o.v[1] : o.v[2]);
      off[j] = (uint32_t)((size_t)(dwrow - r0) * ld + row);
      const bool ok = dwrow < p.N;
      P[j] = ok ? pb[off[j]] : 0.f;
      M[j] = ok ? mb[off[j]] : 0.f;
      V[j] = ok ? vb[off[j]] : 0.f;
    }
#pragma unroll
    for (int j = 0; j < 16; ++j) {
      const int dwrow = col0 + 16 * h + j;
      if (dwrow >= p.N) continue;
      int s = 0;
      if (o.nseg > 1 && dwrow >= o.row0[1]) s = 1;
      if (o.nseg > 2 && dwrow >= o.row0[2]) s = 2;
      float* pb = s == 0 ? o.p[0] : (s == 1 ? o.p[1] : o.p[2]);
      float* mb = s == 0 ? o.m[0] : (s == 1 ? o.m[1] : o.m[2]);
      float* vb = s == 0 ? o.v[0] : (s == 1 ? o.v[1] : o.v[2]);
      const float g = __uint_as_float(r[16 * h + j]);
      const float gk = g * k.gscale;
      const float mk = k.b1 * M[j] + (1.f - k.b1) * gk;
      const float vk = k.b2 * V[j] + (1.f - k.b2) * gk * gk;
      float pk = P[j];
      if (o.mode == 0) {
        pk -= k.alpha_t * mk / (sqrtf(vk) + k.eps);
        pk -= k.lr * k.wd * pk;
      } else {
        pk *= (1.f - k.lr * k.wd);
        pk -= (k.lr / k.bc1) * mk / (sqrtf(vk) * k.rsqrt_bc2 + k.eps);
      }
      pb[off[j]] = pk;
      mb[off[j]] = mk;
      vb[off[j]] = vk;
      if (o.shadow) o.shadow[(size_t)dwrow * ld + row] = __float2bfloat16(pk);
      if (p.C) reinterpret_cast<float*>(p.C)[(size_t)dwrow * p.ldc + row] = g;
    }
  }
}

// ---- SwiGLU gate in the GEMM epilogues (training) ------------------------------------------------------------
__device__ __forceinline__ void st_bf16x32(__nv_bfloat16* dst, const float (&v)[32], bool vec, int valid) {
  if (vec) {
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      u32x4 o;
      o.x = pack_bf16x2(v[8 * q + 0], v[8 * q + 1]);
      o.y = pack_bf16x2(v[8 * q + 2], v[8 * q + 3]);
      o.z = pack_bf16x2(v[8 * q + 4], v[8 * q + 5]);
      o.w = pack_bf16x2(v[8 * q + 6], v[8 * q + 7]);
      *reinterpret_cast<u32x4*>(dst + 8 * q) = o;
    }
  } else {
#pragma unroll
    for (int j = 0; j < 32; ++j)
      if (j < valid) dst[j] = __float2bfloat16(v[j]);
  }
}
// ra / rb: 32 gate / up pre-activations of hidden units hid0..hid0+31 of one row
__device__ __forceinline__ void swiglu_fwd_chunk(const Params& p, int row, int hid0, const uint32_t (&ra)[32],
                                                 const uint32_t (&rb)[32]) {
  const int F = p.N >> 1;
  if (row >= p.M || hid0 >= F) return;
  const int valid = F - hid0 < 32 ? F - hid0 : 32;
  float a[32], b[32], g[32];
#pragma unroll
  for (int j = 0; j < 32; ++j) {
    a[j] = __uint_as_float(ra[j]);
    b[j] = __uint_as_float(rb[j]);
    g[j] = a[j] / (1.f + __expf(-a[j])) * b[j];
  }
  __nv_bfloat16* gp = reinterpret_cast<__nv_bfloat16*>(p.C) + (size_t)row * p.ldc + hid0;
  __nv_bfloat16* abp = reinterpret_cast<__nv_bfloat16*>(const_cast<void*>(p.aux)) + (size_t)row * p.ld_aux + hid0;
  const bool vec = valid == 32 && (p.ldc & 7) == 0 && (p.ld_aux & 7) == 0 && (F & 7) == 0;
  st_bf16x32(gp, g, vec, valid);
  st_bf16x32(abp, a, vec, valid);
  st_bf16x32(abp + F, b, vec, valid);
}
// r: 32 columns col0..col0+31 of the dg tile of one row
__device__ __forceinline__ void swiglu_bwd_chunk(const Params& p, int row, int col0, const uint32_t (&r)[32]) {
  const int F = p.N;
  if (row >= p.M || col0 >= F) return;
  const int valid = F - col0 < 32 ? F - col0 : 32;
  const __nv_bfloat16* abp = reinterpret_cast<const __nv_bfloat16*>(p.aux) + (size_t)row * p.ld_aux + col0;
  __nv_bfloat16* dp = reinterpret_cast<__nv_bfloat16*>(p.C) + (size_t)row * p.ldc + col0;
  const bool vec = valid == 32 && (p.ldc & 7) == 0 && (p.ld_aux & 7) == 0 && (F & 7) == 0;
  float a[32], b[32];
  if (vec) {
    u32x4 av[4], bv[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {  // eight 16-byte loads issued up front
      av[q] = *reinterpret_cast<const u32x4*>(abp + 8 * q);
      bv[q] = *reinterpret_cast<const u32x4*>(abp + F + 8 * q);
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const uint32_t aw[4] = {av[q].x, av[q].y, av[q].z, av[q].w}, bw[4] = {bv[q].x, bv[q].y, bv[q].z, bv[q].w};
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const float2 fa = unpack_bf16x2(aw[e]), fb = unpack_bf16x2(bw[e]);
        a[8 * q + 2 * e] = fa.x; a[8 * q + 2 * e + 1] = fa.y;
        b[8 * q + 2 * e] = fb.x; b[8 * q + 2 * e + 1] = fb.y;
      }
    }
  } else {
#pragma unroll
    for (int j = 0; j < 32; ++j) {
      a[j] = j < valid ? __bfloat162float(abp[j]) : 0.f;
      b[j] = j < valid ? __bfloat162float(abp[F + j]) : 0.f;
    }
  }
  float da[32], db[32];
#pragma unroll
  for (int j = 0; j < 32; ++j) {
    const float dg = __uint_as_float(r[j]);
    const float sg = 1.f / (1.f + __expf(-a[j]));
    da[j] = dg * b[j] * (sg + a[j] * sg * (1.f - sg));
    db[j] = dg * a[j] * sg;
  }
  st_bf16x32(dp, da, vec, valid);
  st_bf16x32(dp + F, db, vec, valid);
}

// epilogue warps per CTA: the optimizer epilogue streams 26 bytes per accumulator element and needs twice the loads
// in flight of a plain store epilogue
// (the SwiGLU-backward epilogue has two dependent 16-byte loads per 8 elements on its critical path: with 4 warps the
// dgrad GEMM ran at 35 % tensor-pipe activity — ncu, profiles/r2/ncu_r2_summary.md — i.e. epilogue-bound)
template <int EPI> struct EpiWarps { static constexpr int N = (EPI == EPI_ADAMW || EPI == EPI_SWIGLU_BWD) ? 8 : 4; };

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

// host entry points of the two kernel families (gemm_sm100.cu: one CTA per tile; gemm2_sm100.cu: CTA pairs), shared by
// the C ABI wrappers; `opt` is only read for EPI_ADAMW
int gemm1_run(const void* A, const void* B, void* C, const void* aux, int M, int N, int K, int lda, int ldb, int ldc,
              int ld_aux, int a_mn_major, int b_mn_major, int epilogue, int force_bn, int rope_S, int rope_dh,
              int rope_cols, int rope_ctx, const AdamwOperands* opt, void* stream);
int gemm2_run(const void* A, const void* B, void* C, const void* aux, int M, int N, int K, int lda, int ldb, int ldc,
              int ld_aux, int a_mn_major, int b_mn_major, int epilogue, int force_bn, int rope_S, int rope_dh,
              int rope_cols, int rope_ctx, const AdamwOperands* opt, void* stream);

}  // namespace gemm
}  // namespace b200
