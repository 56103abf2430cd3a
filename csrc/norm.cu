// Fused RMSNorm (+ residual add) forward and backward.
//
// Replaces the ≈7 ATen kernels per RMSNorm call of the reference LM (model.py:101-107) and the
// separate residual adds (model.py:380-386): one pass reads the fp32 residual stream (+ the bf16
// branch output that still has to be added), writes the updated fp32 stream, the normalised bf16
// activations for the next GEMM and the per-row rstd.  fp32 math throughout.
#include "common.cuh"
#include "ptx.cuh"

#include <algorithm>
#include <cstdlib>
#include <mutex>

namespace b200 {

template <typename T> __device__ __forceinline__ float ld_as_float(const T* p, size_t i);
template <> __device__ __forceinline__ float ld_as_float<float>(const float* p, size_t i) { return p[i]; }
template <> __device__ __forceinline__ float ld_as_float<__nv_bfloat16>(const __nv_bfloat16* p, size_t i) {
  return __bfloat162float(p[i]);
}
template <typename T> __device__ __forceinline__ void st_from_float(T* p, size_t i, float v);
template <> __device__ __forceinline__ void st_from_float<float>(float* p, size_t i, float v) { p[i] = v; }
template <> __device__ __forceinline__ void st_from_float<__nv_bfloat16>(__nv_bfloat16* p, size_t i, float v) {
  p[i] = __float2bfloat16(v);
}

// ---- forward ---------------------------------------------------------------------------------
// one CTA per row; the row (after the optional residual add) is staged in shared memory as fp32
template <typename TX, typename TY>
__global__ void __launch_bounds__(256) rmsnorm_fwd_kernel(const TX* __restrict__ x, const __nv_bfloat16* __restrict__ resid,
                                                          const float* __restrict__ w, TY* __restrict__ y,
                                                          float* __restrict__ xout, float* __restrict__ rstd_out,
                                                          int d, float eps) {
  extern __shared__ float row[];
  __shared__ float red[32];
  ptx::griddep_launch();
  ptx::griddep_wait();
  const size_t base = (size_t)blockIdx.x * d;
  float ss = 0.f;
  if (sizeof(TX) == 4 && (d & 3) == 0) {
    const f32x4* x4 = reinterpret_cast<const f32x4*>(reinterpret_cast<const float*>(x) + base);
    for (int i = threadIdx.x; i < d / 4; i += blockDim.x) {
      f32x4 v = x4[i];
      if (resid) {
        const uint2 rb = *reinterpret_cast<const uint2*>(resid + base + 4 * i);
        const float2 r0 = unpack_bf16x2(rb.x), r1 = unpack_bf16x2(rb.y);
        v.x += r0.x; v.y += r0.y; v.z += r1.x; v.w += r1.y;
      }
      if (xout) *reinterpret_cast<f32x4*>(xout + base + 4 * i) = v;
      *reinterpret_cast<f32x4*>(row + 4 * i) = v;
      ss += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
    }
  } else {
    for (int i = threadIdx.x; i < d; i += blockDim.x) {
      float v = ld_as_float<TX>(x, base + i);
      if (resid) v += __bfloat162float(resid[base + i]);
      if (xout) xout[base + i] = v;
      row[i] = v;
      ss += v * v;
    }
  }
  ss = block_sum(ss, red);
  const float rstd = rsqrtf(ss / (float)d + eps);
  if (threadIdx.x == 0 && rstd_out) rstd_out[blockIdx.x] = rstd;
  if (sizeof(TY) == 2 && (d & 7) == 0) {
    for (int i = threadIdx.x; i < d / 8; i += blockDim.x) {
      const f32x4 a = *reinterpret_cast<const f32x4*>(row + 8 * i), b = *reinterpret_cast<const f32x4*>(row + 8 * i + 4);
      const f32x4 wa = *reinterpret_cast<const f32x4*>(w + 8 * i), wb = *reinterpret_cast<const f32x4*>(w + 8 * i + 4);
      u32x4 o;
      o.x = pack_bf16x2(a.x * rstd * wa.x, a.y * rstd * wa.y);
      o.y = pack_bf16x2(a.z * rstd * wa.z, a.w * rstd * wa.w);
      o.z = pack_bf16x2(b.x * rstd * wb.x, b.y * rstd * wb.y);
      o.w = pack_bf16x2(b.z * rstd * wb.z, b.w * rstd * wb.w);
      *reinterpret_cast<u32x4*>(reinterpret_cast<__nv_bfloat16*>(y) + base + 8 * i) = o;
    }
  } else {
    for (int i = threadIdx.x; i < d; i += blockDim.x) st_from_float<TY>(y, base + i, row[i] * rstd * w[i]);
  }
}

// Warp-per-row forward for the LM's hot case (fp32 residual stream in, bf16 activations out, d % 4 == 0): the row
// (after the optional residual add) lives in registers — NV float4 columns per lane — the sum of squares is a warp
// shuffle reduction, no shared memory and no block barrier; eight rows per CTA.  The gain vector is staged in shared
// memory before the dependency wait (it is a parameter, not an output of the previous kernel).
template <int NV>
__global__ void __launch_bounds__(256) rmsnorm_fwd_rows_kernel(const float* __restrict__ x, const __nv_bfloat16* __restrict__ resid,
                                                               const float* __restrict__ w, __nv_bfloat16* __restrict__ y,
                                                               float* __restrict__ xout, float* __restrict__ rstd_out,
                                                               int T, int d, float eps) {
  extern __shared__ float sw[];
  ptx::griddep_launch();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nvec = d >> 2;
  for (int c = threadIdx.x; c < nvec; c += 256) *reinterpret_cast<f32x4*>(sw + 4 * c) = *reinterpret_cast<const f32x4*>(w + 4 * c);
  __syncthreads();
  ptx::griddep_wait();
  for (int r = blockIdx.x * 8 + warp; r < T; r += gridDim.x * 8) {
    const size_t base = (size_t)r * d;
    f32x4 v[NV];
    uint2 rb[NV];
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int c = lane + 32 * i;
      if (c < nvec) {
        v[i] = *reinterpret_cast<const f32x4*>(x + base + 4 * c);
        if (resid) rb[i] = *reinterpret_cast<const uint2*>(resid + base + 4 * c);
      }
    }
    float ss = 0.f;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int c = lane + 32 * i;
      if (c < nvec) {
        if (resid) {
          const float2 r0 = unpack_bf16x2(rb[i].x), r1 = unpack_bf16x2(rb[i].y);
          v[i].x += r0.x; v[i].y += r0.y; v[i].z += r1.x; v[i].w += r1.y;
        }
        if (xout) *reinterpret_cast<f32x4*>(xout + base + 4 * c) = v[i];
        ss += v[i].x * v[i].x + v[i].y * v[i].y + v[i].z * v[i].z + v[i].w * v[i].w;
      }
    }
    ss = warp_sum(ss);
    const float rstd = rsqrtf(ss / (float)d + eps);
    if (lane == 0 && rstd_out) rstd_out[r] = rstd;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int c = lane + 32 * i;
      if (c < nvec) {
        const f32x4 wv = *reinterpret_cast<const f32x4*>(sw + 4 * c);
        *reinterpret_cast<uint2*>(y + base + 4 * c) =
            uint2{pack_bf16x2(v[i].x * rstd * wv.x, v[i].y * rstd * wv.y), pack_bf16x2(v[i].z * rstd * wv.z, v[i].w * rstd * wv.w)};
      }
    }
  }
}

// ---- backward --------------------------------------------------------------------------------
// dx = dres_in + rstd * (g - xhat * mean(g * xhat)),  g = dy * w,  xhat = x * rstd
// dw[j] = sum_rows dy[r,j] * xhat[r,j]   (per-CTA partials in `dw_partial`, reduced by a 2nd kernel)
template <typename TDY, int VPT>
__global__ void __launch_bounds__(256) rmsnorm_bwd_kernel(const TDY* __restrict__ dy, const float* __restrict__ x,
                                                          const float* __restrict__ w, const float* __restrict__ rstd,
                                                          const float* __restrict__ dres_in, float* __restrict__ dx,
                                                          __nv_bfloat16* __restrict__ dx_bf16,
                                                          float* __restrict__ dw_partial, int T, int d) {
  __shared__ float red[32];
  float dw_acc[VPT];
#pragma unroll
  for (int v = 0; v < VPT; ++v) dw_acc[v] = 0.f;
  for (int r = blockIdx.x; r < T; r += gridDim.x) {
    const size_t base = (size_t)r * d;
    const float rs = rstd[r];
    float gx[VPT], xh[VPT];
    float dot = 0.f;
#pragma unroll
    for (int v = 0; v < VPT; ++v) {
      const int j = threadIdx.x + v * blockDim.x;
      gx[v] = 0.f; xh[v] = 0.f;
      if (j < d) {
        const float dyv = ld_as_float<TDY>(dy, base + j);
        xh[v] = x[base + j] * rs;
        gx[v] = dyv * w[j];
        dw_acc[v] += dyv * xh[v];
        dot += gx[v] * xh[v];
      }
    }
    dot = block_sum(dot, red) / (float)d;
#pragma unroll
    for (int v = 0; v < VPT; ++v) {
      const int j = threadIdx.x + v * blockDim.x;
      if (j < d) {
        float o = rs * (gx[v] - xh[v] * dot);
        if (dres_in) o += dres_in[base + j];
        if (dx) dx[base + j] = o;
        if (dx_bf16) dx_bf16[base + j] = __float2bfloat16(o);
      }
    }
  }
#pragma unroll
  for (int v = 0; v < VPT; ++v) {
    const int j = threadIdx.x + v * blockDim.x;
    if (j < d) dw_partial[(size_t)blockIdx.x * d + j] = dw_acc[v];
  }
}

// Fast path (d % 8 == 0): two bandwidth-shaped kernels instead of one register-starved one.
//  (1) dx: one WARP per row, the row's dy·w and x·rstd held in registers (16-byte loads, all issued up
//      front), warp-shuffle dot product, no block barrier;
//  (2) dw: column-parallel partial sums over chunks of rows (each thread owns 4 columns, coalesced),
//      re-reading dy and x (+20 MB at XL, ≈3 µs) — cheaper than starving (1) of registers.
template <typename TDY>
__device__ __forceinline__ f32x4 ld_dy4(const TDY* dy, size_t off) {
  if constexpr (sizeof(TDY) == 2) {
    const uint2 raw = *reinterpret_cast<const uint2*>(reinterpret_cast<const __nv_bfloat16*>(dy) + off);
    const float2 a = unpack_bf16x2(raw.x), b = unpack_bf16x2(raw.y);
    return f32x4{a.x, a.y, b.x, b.y};
  } else {
    return *reinterpret_cast<const f32x4*>(reinterpret_cast<const float*>(dy) + off);
  }
}

// dx: one 128-thread CTA per row, the row's g = dy*w and x held in registers (<= VPT float4 per thread, all
// loads issued up front), one block reduction.  T rows -> T CTAs: enough bytes in flight to run at HBM speed
// even for T = 2048 (a warp-per-row variant had too few loads outstanding per SM).
template <typename TDY, int VPT>
__global__ void __launch_bounds__(128) rmsnorm_bwd_dx_kernel(const TDY* __restrict__ dy, const float* __restrict__ x,
                                                             const float* __restrict__ w, const float* __restrict__ rstd,
                                                             const float* __restrict__ dres_in, float* __restrict__ dx,
                                                             __nv_bfloat16* __restrict__ dx_bf16, int T, int d) {
  __shared__ float red[32];
  ptx::griddep_launch();
  ptx::griddep_wait();
  const int r = blockIdx.x;
  const int nvec = d >> 2;
  const size_t base = (size_t)r * d;
  const float rs = rstd[r];
  f32x4 g[VPT], xv[VPT], dr[VPT];
#pragma unroll
  for (int v = 0; v < VPT; ++v) {
    const int c = threadIdx.x + 128 * v;
    if (c < nvec) {
      g[v] = ld_dy4<TDY>(dy, base + 4 * c);
      xv[v] = *reinterpret_cast<const f32x4*>(x + base + 4 * c);
      if (dres_in) dr[v] = *reinterpret_cast<const f32x4*>(dres_in + base + 4 * c);
    }
  }
  float dot = 0.f;
#pragma unroll
  for (int v = 0; v < VPT; ++v) {
    const int c = threadIdx.x + 128 * v;
    if (c < nvec) {
      const f32x4 wv = *reinterpret_cast<const f32x4*>(w + 4 * c);
      g[v].x *= wv.x; g[v].y *= wv.y; g[v].z *= wv.z; g[v].w *= wv.w;
      dot += g[v].x * xv[v].x + g[v].y * xv[v].y + g[v].z * xv[v].z + g[v].w * xv[v].w;
    }
  }
  dot = block_sum(dot, red) * rs / (float)d;  // mean(g * xhat)
  const float k2 = rs * rs * dot;             // rs*(g - xhat*dot) = rs*g - rs^2*dot*x
#pragma unroll
  for (int v = 0; v < VPT; ++v) {
    const int c = threadIdx.x + 128 * v;
    if (c < nvec) {
      f32x4 o{rs * g[v].x - k2 * xv[v].x, rs * g[v].y - k2 * xv[v].y, rs * g[v].z - k2 * xv[v].z,
              rs * g[v].w - k2 * xv[v].w};
      if (dres_in) { o.x += dr[v].x; o.y += dr[v].y; o.z += dr[v].z; o.w += dr[v].w; }
      if (dx) *reinterpret_cast<f32x4*>(dx + base + 4 * c) = o;
      if (dx_bf16) *reinterpret_cast<uint2*>(dx_bf16 + base + 4 * c) = uint2{pack_bf16x2(o.x, o.y), pack_bf16x2(o.z, o.w)};
    }
  }
}

// dx AND the dw partial sums in one pass, one WARP per row (the split above re-reads dy and x for dw: +20 MB and one
// more launch per norm; a persistent CTA-per-row loop with block-wide reductions measured 23 us per call, its
// iteration time pinned to the memory latency).  A warp holds its row's dy and x in registers (NV float4 columns per
// lane), reduces dy·w·x with shuffles — no block barrier anywhere in the row loop — and writes dx; eight warps per
// CTA and two CTAs per SM keep ~200 KB of loads in flight per SM.  Each warp accumulates its rows' dy·x·rstd into its
// OWN d-wide slot of shared memory (no atomics); the CTA's eight slots are summed into partial[blockIdx.x][:] at the
// end and colsum_rows_kernel reduces the partial matrix.
// the row's dy stays in its storage format (4 bf16 = 2 registers) and is widened where it is used
template <typename TDY> struct DyPack;
template <> struct DyPack<__nv_bfloat16> {
  uint2 raw;
  __device__ __forceinline__ void load(const __nv_bfloat16* p, size_t off) { raw = *reinterpret_cast<const uint2*>(p + off); }
  __device__ __forceinline__ f32x4 get() const {
    const float2 a = unpack_bf16x2(raw.x), b = unpack_bf16x2(raw.y);
    return f32x4{a.x, a.y, b.x, b.y};
  }
};
template <> struct DyPack<float> {
  f32x4 raw;
  __device__ __forceinline__ void load(const float* p, size_t off) { raw = *reinterpret_cast<const f32x4*>(p + off); }
  __device__ __forceinline__ f32x4 get() const { return raw; }
};

template <typename TDY, int NV>
__global__ void __launch_bounds__(256, (NV * (4 + (int)sizeof(TDY)) <= 13 * 6) ? 2 : 1)
rmsnorm_bwd_rows_kernel(const TDY* __restrict__ dy, const float* __restrict__ x, const float* __restrict__ w,
                        const float* __restrict__ rstd, const float* __restrict__ dres_in, float* __restrict__ dx,
                        __nv_bfloat16* __restrict__ dx_bf16, float* __restrict__ partial, int T, int d) {
  extern __shared__ float sacc[];  // [8 warps][d] dw slots, then [d] gains
  ptx::griddep_launch();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nvec = d >> 2;
  float* mine = sacc + (size_t)warp * d;
  float* sw = sacc + (size_t)8 * d;
  // the gain vector is a parameter (not produced by the previous kernel): stage it before the dependency wait
  for (int c = threadIdx.x; c < nvec; c += 256) *reinterpret_cast<f32x4*>(sw + 4 * c) = *reinterpret_cast<const f32x4*>(w + 4 * c);
  __syncthreads();
  ptx::griddep_wait();
  bool first = true;
  for (int r = blockIdx.x * 8 + warp; r < T; r += gridDim.x * 8) {
    const size_t base = (size_t)r * d;
    const float rs = rstd[r];
    if (dres_in) {  // needed last: pull the row towards L2 now, without tying up registers
#pragma unroll
      for (int i = 0; i < NV; ++i) {
        const int c = lane + 32 * i;
        if (c < nvec && (c & 7) == 0) asm volatile("prefetch.global.L2 [%0];" ::"l"(dres_in + base + 4 * c));
      }
    }
    DyPack<TDY> gy[NV];
    f32x4 xv[NV];
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int c = lane + 32 * i;
      if (c < nvec) {
        gy[i].load(dy, base + 4 * c);
        xv[i] = *reinterpret_cast<const f32x4*>(x + base + 4 * c);
      }
    }
    float dot = 0.f;
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int c = lane + 32 * i;
      if (c < nvec) {
        const f32x4 g = gy[i].get();
        // this row's contribution to dw (raw dy), into the warp's own shared-memory slot
        const f32x4 a{g.x * xv[i].x * rs, g.y * xv[i].y * rs, g.z * xv[i].z * rs, g.w * xv[i].w * rs};
        f32x4* slot = reinterpret_cast<f32x4*>(mine + 4 * c);
        if (first) {
          *slot = a;
        } else {
          const f32x4 o = *slot;
          *slot = f32x4{o.x + a.x, o.y + a.y, o.z + a.z, o.w + a.w};
        }
        const f32x4 wv = *reinterpret_cast<const f32x4*>(sw + 4 * c);
        dot += g.x * wv.x * xv[i].x + g.y * wv.y * xv[i].y + g.z * wv.z * xv[i].z + g.w * wv.w * xv[i].w;
      }
    }
    first = false;
    dot = warp_sum(dot) * rs / (float)d;  // mean(dy*w * xhat)
    const float k2 = rs * rs * dot;       // rs*(g - xhat*dot) = rs*g - rs^2*dot*x
#pragma unroll
    for (int i = 0; i < NV; ++i) {
      const int c = lane + 32 * i;
      if (c < nvec) {
        const f32x4 g = gy[i].get();
        const f32x4 wv = *reinterpret_cast<const f32x4*>(sw + 4 * c);
        f32x4 o{rs * g.x * wv.x - k2 * xv[i].x, rs * g.y * wv.y - k2 * xv[i].y, rs * g.z * wv.z - k2 * xv[i].z,
                rs * g.w * wv.w - k2 * xv[i].w};
        if (dres_in) {
          const f32x4 dr = *reinterpret_cast<const f32x4*>(dres_in + base + 4 * c);
          o.x += dr.x; o.y += dr.y; o.z += dr.z; o.w += dr.w;
        }
        if (dx) *reinterpret_cast<f32x4*>(dx + base + 4 * c) = o;
        if (dx_bf16) *reinterpret_cast<uint2*>(dx_bf16 + base + 4 * c) = uint2{pack_bf16x2(o.x, o.y), pack_bf16x2(o.z, o.w)};
      }
    }
  }
  if (first) {  // this warp had no row: its slot must still read as zero
    for (int c = lane; c < nvec; c += 32) *reinterpret_cast<f32x4*>(mine + 4 * c) = f32x4{0.f, 0.f, 0.f, 0.f};
  }
  __syncthreads();
  for (int c = threadIdx.x; c < nvec; c += 256) {
    f32x4 t = *reinterpret_cast<const f32x4*>(sacc + 4 * c);
#pragma unroll
    for (int ww = 1; ww < 8; ++ww) {
      const f32x4 u = *reinterpret_cast<const f32x4*>(sacc + (size_t)ww * d + 4 * c);
      t.x += u.x; t.y += u.y; t.z += u.z; t.w += u.w;
    }
    *reinterpret_cast<f32x4*>(partial + (size_t)blockIdx.x * d + 4 * c) = t;
  }
}

// column sums of a [rows, d] partial matrix: 32 columns x 8 row groups per CTA (coalesced 128-byte row segments,
// rows/8 independent loads per thread), combined through shared memory.  d/32 CTAs.
__global__ void __launch_bounds__(256) colsum_rows_kernel(const float* __restrict__ partial, float* __restrict__ out,
                                                          int rows, int d, int accumulate) {
  __shared__ float red[8][33];
  ptx::griddep_launch();
  ptx::griddep_wait();
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int j = blockIdx.x * 32 + tx;
  float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
  if (j < d) {
    int r = ty;
    for (; r + 24 < rows; r += 32) {
      s0 += partial[(size_t)r * d + j];
      s1 += partial[(size_t)(r + 8) * d + j];
      s2 += partial[(size_t)(r + 16) * d + j];
      s3 += partial[(size_t)(r + 24) * d + j];
    }
    for (; r < rows; r += 8) s0 += partial[(size_t)r * d + j];
  }
  red[ty][tx] = (s0 + s1) + (s2 + s3);
  __syncthreads();
  if (ty == 0 && j < d) {
    float s = 0.f;
#pragma unroll
    for (int q = 0; q < 8; ++q) s += red[q][tx];
    out[j] = accumulate ? out[j] + s : s;
  }
}

// partial[chunk][j] = sum over the chunk's rows of dy[r,j] * x[r,j] * rstd[r]
template <typename TDY>
__global__ void __launch_bounds__(128) rmsnorm_bwd_dw_kernel(const TDY* __restrict__ dy, const float* __restrict__ x,
                                                             const float* __restrict__ rstd, float* __restrict__ partial,
                                                             int T, int d, int rows_per_chunk) {
  ptx::griddep_launch();
  ptx::griddep_wait();
  const int c = blockIdx.x * 128 + threadIdx.x;  // float4 column
  if (c >= (d >> 2)) return;
  const int r0 = blockIdx.y * rows_per_chunk;
  const int r1 = min(T, r0 + rows_per_chunk);
  f32x4 acc{0.f, 0.f, 0.f, 0.f};
#pragma unroll 8
  for (int r = r0; r < r1; ++r) {
    const size_t off = (size_t)r * d + 4 * c;
    const f32x4 dyv = ld_dy4<TDY>(dy, off);
    const f32x4 xv = *reinterpret_cast<const f32x4*>(x + off);
    const float rs = rstd[r];
    acc.x += dyv.x * xv.x * rs; acc.y += dyv.y * xv.y * rs; acc.z += dyv.z * xv.z * rs; acc.w += dyv.w * xv.w * rs;
  }
  *reinterpret_cast<f32x4*>(partial + (size_t)blockIdx.y * d + 4 * c) = acc;
}

// column sums of a [rows, d] partial matrix: 4 row-groups per column tile, combined through shared memory
__global__ void __launch_bounds__(256) colsum4_kernel(const float* __restrict__ partial, float* __restrict__ out,
                                                      int rows, int d, int accumulate) {
  __shared__ float red[4][64];
  ptx::griddep_launch();
  ptx::griddep_wait();
  const int tx = threadIdx.x & 63, ty = threadIdx.x >> 6;
  const int j = blockIdx.x * 64 + tx;
  float s = 0.f;
  if (j < d)
    for (int r = ty; r < rows; r += 4) s += partial[(size_t)r * d + j];
  red[ty][tx] = s;
  __syncthreads();
  if (ty == 0 && j < d) {
    s = red[0][tx] + red[1][tx] + red[2][tx] + red[3][tx];
    out[j] = accumulate ? out[j] + s : s;
  }
}

__global__ void colsum_kernel(const float* __restrict__ partial, float* __restrict__ out, int rows, int d,
                              int accumulate) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= d) return;
  float s = 0.f;
  for (int r = 0; r < rows; ++r) s += partial[(size_t)r * d + j];
  out[j] = accumulate ? out[j] + s : s;
}

// residual add: out fp32 = x fp32 + y bf16
__global__ void residual_add_kernel(const float* __restrict__ x, const __nv_bfloat16* __restrict__ y,
                                    float* __restrict__ out, size_t n) {
  const size_t i4 = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
  if (i4 + 3 < n) {
    f32x4 v = *reinterpret_cast<const f32x4*>(x + i4);
    const uint2 rb = *reinterpret_cast<const uint2*>(y + i4);
    const float2 a = unpack_bf16x2(rb.x), b = unpack_bf16x2(rb.y);
    v.x += a.x; v.y += a.y; v.z += b.x; v.w += b.y;
    *reinterpret_cast<f32x4*>(out + i4) = v;
  } else {
    for (size_t i = i4; i < n; ++i) out[i] = x[i] + __bfloat162float(y[i]);
  }
}

}  // namespace b200

using namespace b200;

extern "C" int b200_rmsnorm_fwd(const void* x, int x_dtype, const void* resid_bf16, const float* w, void* y,
                                int y_dtype, float* xout, float* rstd, int T, int d, float eps, void* stream) {
  B200_REQUIRE(T > 0 && d > 0, "rmsnorm_fwd: empty input");
  B200_REQUIRE(d * 4 <= 160 * 1024, "rmsnorm_fwd: hidden size %d too large for the smem-staged kernel", d);
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  const int nv = ceil_div(d >> 2, 32);
  const bool al16 = ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(w) | reinterpret_cast<uintptr_t>(xout)) & 15) == 0 &&
                    ((reinterpret_cast<uintptr_t>(resid_bf16) | reinterpret_cast<uintptr_t>(y)) & 7) == 0;
  if (x_dtype == B200_F32 && y_dtype == B200_BF16 && (d & 3) == 0 && nv <= 24 && al16) {
    // hot case: one warp per row (getenv B200SYS_RMSNORM_FWD=cta keeps the CTA-per-row kernel for A/B runs)
    static int rows_mode = -1;
    if (rows_mode < 0) {
      const char* e = getenv("B200SYS_RMSNORM_FWD");
      rows_mode = (e && e[0] == 'c') ? 0 : 1;
    }
    if (rows_mode) {
      const int grid = std::min(ceil_div(T, 8), num_sms() * 4);
      const size_t sm = (size_t)d * 4;
#define LAUNCH_FR(NVV)                                                                                               \
  do {                                                                                                               \
    auto kfn = rmsnorm_fwd_rows_kernel<NVV>;                                                                         \
    if (sm > 48 * 1024) B200_ENSURE_DYN_SMEM(kfn, 4 * 128 * NVV);                                                    \
    B200_CHECK_CUDA(launch_pdl(kfn, dim3(grid), dim3(256), sm, s, reinterpret_cast<const float*>(x),                 \
                               reinterpret_cast<const __nv_bfloat16*>(resid_bf16), w,                                \
                               reinterpret_cast<__nv_bfloat16*>(y), xout, rstd, T, d, eps));                         \
  } while (0)
      if (nv <= 4) LAUNCH_FR(4);
      else if (nv <= 8) LAUNCH_FR(8);
      else if (nv <= 13) LAUNCH_FR(13);
      else if (nv <= 16) LAUNCH_FR(16);
      else if (nv <= 20) LAUNCH_FR(20);
      else LAUNCH_FR(24);
#undef LAUNCH_FR
      B200_CHECK_CUDA(cudaGetLastError());
      return 0;
    }
  }
  const int threads = d >= 2048 ? 256 : 128;
  const size_t smem = (size_t)d * 4;
#define LAUNCH(TX, TY)                                                                                            \
  do {                                                                                                            \
    auto k = rmsnorm_fwd_kernel<TX, TY>;                                                                          \
    if (smem > 48 * 1024) B200_CHECK_CUDA(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
    B200_CHECK_CUDA(launch_pdl(k, dim3(T), dim3(threads), smem, s, reinterpret_cast<const TX*>(x),                   \
                               reinterpret_cast<const __nv_bfloat16*>(resid_bf16), w, reinterpret_cast<TY*>(y), xout,  \
                               rstd, d, eps));                                                                     \
  } while (0)
  if (x_dtype == B200_F32 && y_dtype == B200_BF16) LAUNCH(float, __nv_bfloat16);
  else if (x_dtype == B200_F32 && y_dtype == B200_F32) LAUNCH(float, float);
  else if (x_dtype == B200_BF16 && y_dtype == B200_BF16) LAUNCH(__nv_bfloat16, __nv_bfloat16);
  else if (x_dtype == B200_BF16 && y_dtype == B200_F32) LAUNCH(__nv_bfloat16, float);
  else { set_error("rmsnorm_fwd: bad dtypes %d -> %d", x_dtype, y_dtype); return B200_ERR_ARG; }
#undef LAUNCH
  B200_CHECK_CUDA(cudaGetLastError());  // launch_pdl counted the launch
  return 0;
}

extern "C" int b200_rmsnorm_bwd_workspace_rows(int T) {
  const int cap = num_sms() * 2;
  return T < cap ? T : cap;
}

// scratch: b200_rmsnorm_bwd_workspace_rows(T) * d floats
extern "C" int b200_rmsnorm_bwd(const void* dy, int dy_dtype, const float* x, const float* w, const float* rstd,
                                const float* dres_in, float* dx, void* dx_bf16, float* dw, int dw_accumulate,
                                float* scratch, int T, int d, void* stream) {
  B200_REQUIRE(T > 0 && d > 0, "rmsnorm_bwd: empty input");
  B200_REQUIRE(d <= 256 * 48, "rmsnorm_bwd: hidden size %d too large", d);
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  const int grid = b200_rmsnorm_bwd_workspace_rows(T);
  __nv_bfloat16* dxb = reinterpret_cast<__nv_bfloat16*>(dx_bf16);
  const bool aligned = ((reinterpret_cast<uintptr_t>(dy) | reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(w) |
                         reinterpret_cast<uintptr_t>(dres_in) | reinterpret_cast<uintptr_t>(dx) |
                         reinterpret_cast<uintptr_t>(dx_bf16)) & 15) == 0;
  if ((d % 8) == 0 && d <= 128 * 24 && aligned) {
    static int fused_mode = -1;
    if (fused_mode < 0) {
      const char* e = getenv("B200SYS_RMSNORM_BWD");
      fused_mode = (e && e[0] == 's') ? 0 : 1;  // "split": the two-kernel form (A/B measurements)
    }
    const int nv = ceil_div(d >> 2, 32);
    if (fused_mode && nv <= 24) {
      // one pass: dx + per-CTA dw partials (<= workspace rows CTAs, so the scratch contract holds), then the column sum
      const int fgrid = std::min(grid, ceil_div(T, 8));
      const size_t smem = (size_t)9 * d * sizeof(float);
#define LAUNCH_R(TDY, NVV)                                                                                          \
  do {                                                                                                              \
    auto kfn = rmsnorm_bwd_rows_kernel<TDY, NVV>;                                                                   \
    B200_ENSURE_DYN_SMEM(kfn, 9 * 4 * 128 * NVV);                                                                   \
    B200_CHECK_CUDA(launch_pdl(kfn, dim3(fgrid), dim3(256), smem, s, reinterpret_cast<const TDY*>(dy), x, w, rstd,  \
                               dres_in, dx, dxb, scratch, T, d));                                                   \
  } while (0)
#define BY_DT_R(NVV)                                                   \
  do {                                                                 \
    if (dy_dtype == B200_BF16) LAUNCH_R(__nv_bfloat16, NVV);           \
    else LAUNCH_R(float, NVV);                                         \
  } while (0)
      if (nv <= 4) BY_DT_R(4);
      else if (nv <= 8) BY_DT_R(8);
      else if (nv <= 13) BY_DT_R(13);
      else if (nv <= 16) BY_DT_R(16);
      else if (nv <= 20) BY_DT_R(20);
      else BY_DT_R(24);
#undef BY_DT_R
#undef LAUNCH_R
      B200_CHECK_CUDA(launch_pdl(colsum_rows_kernel, dim3(ceil_div(d, 32)), dim3(256), 0, s, (const float*)scratch, dw,
                                 fgrid, d, dw_accumulate));
      return 0;
    }
    // fast path; the dw partial matrix has `chunks` <= workspace rows, so the scratch contract holds
    const int rows_per_chunk = std::max(16, ceil_div(T, grid));
    const int chunks = ceil_div(T, rows_per_chunk);
    const int fgrid = chunks;
#define LAUNCH_W(TDY, VPL)                                                                                        \
  B200_CHECK_CUDA(launch_pdl(rmsnorm_bwd_dx_kernel<TDY, VPL>, dim3(T), dim3(128), 0, s,                            \
                             reinterpret_cast<const TDY*>(dy), x, w, rstd, dres_in, dx, dxb, T, d))
#define BY_DT_W(VPL)                                                   \
  do {                                                                 \
    if (dy_dtype == B200_BF16) LAUNCH_W(__nv_bfloat16, VPL);           \
    else LAUNCH_W(float, VPL);                                         \
  } while (0)
    if (d <= 512 * 2) BY_DT_W(2);
    else if (d <= 512 * 4) BY_DT_W(4);
    else BY_DT_W(6);
#undef BY_DT_W
#undef LAUNCH_W
    B200_CHECK_CUDA(cudaGetLastError());  // launch_pdl counted the launch
    {
      dim3 g2(ceil_div(d >> 2, 128), chunks);
      if (dy_dtype == B200_BF16)
        B200_CHECK_CUDA(launch_pdl(rmsnorm_bwd_dw_kernel<__nv_bfloat16>, g2, dim3(128), 0, s,
                                   reinterpret_cast<const __nv_bfloat16*>(dy), x, (const float*)rstd, scratch, T, d, rows_per_chunk));
      else
        B200_CHECK_CUDA(launch_pdl(rmsnorm_bwd_dw_kernel<float>, g2, dim3(128), 0, s, reinterpret_cast<const float*>(dy), x,
                                   (const float*)rstd, scratch, T, d, rows_per_chunk));
    }
    B200_CHECK_CUDA(launch_pdl(colsum4_kernel, dim3(ceil_div(d, 64)), dim3(256), 0, s, (const float*)scratch, dw, fgrid, d,
                               dw_accumulate));
    return 0;
  }
#define LAUNCH(TDY, VPT, THREADS)                                                                              \
  rmsnorm_bwd_kernel<TDY, VPT><<<grid, THREADS, 0, s>>>(reinterpret_cast<const TDY*>(dy), x, w, rstd, dres_in, dx, \
                                                        dxb, scratch, T, d)
#define BY_DTYPE(VPT, THREADS)                                              \
  do {                                                                      \
    if (dy_dtype == B200_BF16) LAUNCH(__nv_bfloat16, VPT, THREADS);         \
    else LAUNCH(float, VPT, THREADS);                                       \
  } while (0)
  if (d <= 128 * 4) BY_DTYPE(4, 128);
  else if (d <= 128 * 8) BY_DTYPE(8, 128);
  else if (d <= 256 * 8) BY_DTYPE(8, 256);
  else if (d <= 256 * 16) BY_DTYPE(16, 256);
  else BY_DTYPE(48, 256);
#undef BY_DTYPE
#undef LAUNCH
  B200_LAUNCH_CHECK();
  colsum_kernel<<<ceil_div(d, 128), 128, 0, s>>>(scratch, dw, grid, d, dw_accumulate);
  B200_LAUNCH_CHECK();
  return 0;
}

extern "C" int b200_residual_add(const float* x, const void* y_bf16, float* out, long long n, void* stream) {
  if (n <= 0) return 0;
  const int threads = 256;
  const long long nvec = (n + 3) / 4;
  residual_add_kernel<<<(unsigned)((nvec + threads - 1) / threads), threads, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      x, reinterpret_cast<const __nv_bfloat16*>(y_bf16), out, (size_t)n);
  B200_LAUNCH_CHECK();
  return 0;
}

// ------------------------------------------------------------------------------------------------
// Allocation-free variant of b200_rmsnorm_bwd.
//
// Who needs it: callers that are not allowed to create device allocations while they run — a host
// application whose whole device-memory budget was handed to its own pool at start-up (so a foreign
// library calling cudaMalloc could fail or fragment it), a latency-critical loop that cannot afford
// the implicit device-wide synchronisation of cudaMalloc/cudaFree, or a plug-in ABI that gives a
// library input and output buffers and nothing else.  b200_rmsnorm_bwd asks such a caller for a
// scratch buffer of workspace_rows(T) * d floats; this variant keeps that scratch inside the library
// image instead, in the fixed-size array below.
//
// Cost and limits: the array (B200_RMSNORM_BWD_NOALLOC_MAX_FLOATS floats, 4.75 MiB) is part of the
// module, so every GPU that loads libb200sys holds it until the process exits.  Problems whose
// partial-sum matrix does not fit are rejected with B200_ERR_ARG and a message; nothing is truncated.
//
// Sharing: all calls on a device use the same array.  The kernels overwrite every element they
// later read (each CTA stores a full d-wide row of partial sums before the column reduction runs),
// so nothing has to be cleared between calls and stale contents from an earlier call are harmless.
// Two calls that overlap on the GPU (different streams, or the same entry point from two threads
// without ordering) race on the array — serialise them.
//
// Set-up: in host code the array's name is only a stand-in for the symbol, not a device address.
// The first call on each device, made with that device current, resolves the real address with
// cudaGetSymbolAddress and caches it; make that first call before starting any stream capture.
// The entry point calls no cudaMalloc/cudaFree and no allocating library routine.
// ------------------------------------------------------------------------------------------------
#define B200_RMSNORM_BWD_NOALLOC_MAX_FLOATS (304 * 4096)

static __device__ __align__(256) float g_rmsnorm_bwd_static_scratch[B200_RMSNORM_BWD_NOALLOC_MAX_FLOATS];

namespace {
constexpr int kMaxDevices = 64;
float* g_static_scratch_addr[kMaxDevices] = {nullptr};
std::mutex g_static_scratch_mu;

// Device address of the static scratch on the current device (resolved once per device).
int static_scratch_for_current_device(float** out) {
  int dev = -1;
  B200_CHECK_CUDA(cudaGetDevice(&dev));
  B200_REQUIRE(dev >= 0 && dev < kMaxDevices, "rmsnorm_bwd_noalloc: device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lock(g_static_scratch_mu);
  if (g_static_scratch_addr[dev] == nullptr) {
    float* ptr = nullptr;
    B200_CHECK_CUDA(cudaGetSymbolAddress(reinterpret_cast<void**>(&ptr), g_rmsnorm_bwd_static_scratch));
    B200_REQUIRE(ptr != nullptr, "rmsnorm_bwd_noalloc: cudaGetSymbolAddress returned a null address");
    g_static_scratch_addr[dev] = ptr;
  }
  *out = g_static_scratch_addr[dev];
  return 0;
}
}  // namespace

extern "C" long long b200_rmsnorm_bwd_noalloc_max_floats(void) { return B200_RMSNORM_BWD_NOALLOC_MAX_FLOATS; }

extern "C" int b200_rmsnorm_bwd_noalloc(const void* dy, int dy_dtype, const float* x, const float* w,
                                        const float* rstd, const float* dres_in, float* dx, void* dx_bf16, float* dw,
                                        int dw_accumulate, int T, int d, void* stream) {
  B200_REQUIRE(T > 0 && d > 0, "rmsnorm_bwd_noalloc: empty input");
  const long long need = (long long)b200_rmsnorm_bwd_workspace_rows(T) * d;
  B200_REQUIRE(need <= (long long)B200_RMSNORM_BWD_NOALLOC_MAX_FLOATS,
               "rmsnorm_bwd_noalloc: needs %lld floats of scratch (rows %d x d %d) but the built-in buffer holds %d; "
               "use b200_rmsnorm_bwd with caller-provided scratch",
               need, b200_rmsnorm_bwd_workspace_rows(T), d, (int)B200_RMSNORM_BWD_NOALLOC_MAX_FLOATS);
  float* scratch = nullptr;
  if (int rc = static_scratch_for_current_device(&scratch)) return rc;
  return b200_rmsnorm_bwd(dy, dy_dtype, x, w, rstd, dres_in, dx, dx_bf16, dw, dw_accumulate, scratch, T, d, stream);
}
