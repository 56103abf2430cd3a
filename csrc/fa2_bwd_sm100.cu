// FlashAttention-2 backward on tcgen05 / TMEM / TMA for sm_100a.
//
// The reference backward (cs336_systems/flash_attention.py:270-289) is a torch.compile'd graph
// that materialises the full N×N S, P, dP and dS matrices (O(N²) memory).  This is the tiled
// algorithm (FlashAttention-2, Algorithm 2): P is recomputed from the saved logsumexp tile by
// tile, nothing N×N ever reaches HBM.
//
//   pre-pass   delta = rowsum(O ∘ dO)                                   (fa2_delta_kernel)
//   main       one CTA per 128-key tile j of one (batch, head); loop over query tiles i:
//                S  = Q_i·K_jᵀ          dP = dO_i·V_jᵀ                     (tcgen05, both K-major)
//                P  = exp2(S·c − L_i)    dS = P ∘ (dP − delta_i)·scale      (softmax warpgroup, TMEM→regs→smem)
//                dV_j += Pᵀ·dO_i        dK_j += dSᵀ·Q_i                    (A read MN-major from the same smem tile)
//                dQ_i += dS·K_j                                            (partial in TMEM → fp32 red.add.v4 to HBM)
//              dK_j, dV_j stay in tensor memory for the whole loop.
//   post-pass  dQ fp32 accumulator → bf16 / fp32, RoPE un-rotation         (fa2_dq_convert_kernel)
//
// Two main kernels: fa2_bwd_wgr_kernel (head dim <= 64: S/dP of the next query tile are issued before dV/dK/dQ of the
// current one, P/dS staged in registers, four softmax column groups) and fa2_bwd_wg_kernel (head dim <= 128: two
// column groups, the dQ partial shares S's TMEM columns).  fp32 callers get fp32 dQ/dK/dV straight from the kernels
// (out_f32) and may pass fp32 O / dO to the delta pre-pass.
#include "common.cuh"
#include <cstdlib>
#include "ptx.cuh"
#include "tensormap.h"

namespace b200 {
namespace fa2 {

constexpr int BM = 128;
constexpr int BN = 128;
constexpr float kLog2e = 1.4426950408889634f;

int make_bhnd_map(CUtensorMap* out, const void* base, int B, int H, int N, int D, const long long* strides,
                  int box_rows);

template <int DP>
struct BwdCfg {
  static constexpr int NBLK = DP / 64;
  static constexpr int TILE_BYTES = 128 * DP * 2;
  static constexpr int PS_BYTES = BM * BN * 2;
  static constexpr int STAGES = DP == 64 ? 2 : 1;
  static constexpr bool ALIAS = DP != 64;  // dQ partial shares S's TMEM columns when DP = 128
  static constexpr int SMEM_BYTES = 2 * TILE_BYTES + STAGES * 2 * TILE_BYTES + 2 * PS_BYTES + 1024 + 256;
  static constexpr int COL_S = 0, COL_DP = 128, COL_DV = 256, COL_DK = 256 + DP;
  static constexpr int COL_DQ = ALIAS ? 0 : 256 + 2 * DP;
  static constexpr int TMEM_COLS = 512;
};

struct BwdParams {
  const float* lse;
  const float* delta;
  float* dq_acc;  // [B,H,Nq,D] fp32, zeroed
  void *dk, *dv;  // bf16, or fp32 when out_f32
  int out_f32;
  long long dk_sb, dk_sh, dk_sr, dv_sb, dv_sh, dv_sr;
  int H, Nq, Nk, D;
  int causal;
  float scale, scale_log2;
  int num_q_tiles;
  const float* rope_cache;  // nullable: (2, rope_ctx, D/2) cos/sin (used by the dq/dk post-passes)
  int rope_ctx;
  int rope_epilogue;  // 1: dK is un-rotated in the kernel epilogue (head dim 64); 0: by the post-pass
};

__device__ __forceinline__ uint32_t swz_off_b(int row, int col, int rows) {
  const int blk = col >> 6, c = col & 63;
  return blk * (rows * 128) + row * 128 + ((((c >> 3) ^ (row & 7)) << 4) | ((c & 7) << 1));
}

// ------------------------------------------------------------------------------------------------
// NG column groups of softmax warps (this kernel: head dims 65..128, NG = 2).
//
// The P / dS phase has no row reductions (the row constants L and delta come from the forward and the pre-pass), so
// the 128 columns of S and dP can be split among NG warpgroups without any exchange: warps g·4 … g·4+3 own
// columns [g·128/NG, (g+1)·128/NG) of the same 128 rows (a warp may read TMEM lanes 32·(warp%4)…+31 at any column),
// and likewise DP/NG columns of the dQ partial and of the dK / dV accumulators in the epilogue.  With one CTA per SM
// this raises the softmax warps per scheduler from 1 to NG and shortens the serial work per thread by NG.
// ------------------------------------------------------------------------------------------------

// one chunk of W fp32 accumulator columns of a dK / dV row -> global (bf16 or fp32), columns >= D dropped
template <int W>
__device__ __forceinline__ void store_dkv_chunk(void* row_base, bool f32, int col0, int D, const uint32_t (&t)[W]) {
  if (f32) {
    float* out = reinterpret_cast<float*>(row_base);
    if (((reinterpret_cast<uintptr_t>(out) & 15) == 0) && (D % 4 == 0)) {
#pragma unroll
      for (int v = 0; v < W / 4; ++v)
        if (col0 + v * 4 < D)
          *reinterpret_cast<f32x4*>(out + col0 + v * 4) = f32x4{__uint_as_float(t[4 * v]), __uint_as_float(t[4 * v + 1]),
                                                                __uint_as_float(t[4 * v + 2]), __uint_as_float(t[4 * v + 3])};
    } else {
#pragma unroll
      for (int i = 0; i < W; ++i)
        if (col0 + i < D) out[col0 + i] = __uint_as_float(t[i]);
    }
    return;
  }
  __nv_bfloat16* out = reinterpret_cast<__nv_bfloat16*>(row_base);
  if (((reinterpret_cast<uintptr_t>(out) & 15) == 0) && (D % 8 == 0)) {
#pragma unroll
    for (int v = 0; v < W / 8; ++v) {
      if (col0 + v * 8 < D) {
        u32x4 o;
        o.x = pack_bf16x2(__uint_as_float(t[8 * v + 0]), __uint_as_float(t[8 * v + 1]));
        o.y = pack_bf16x2(__uint_as_float(t[8 * v + 2]), __uint_as_float(t[8 * v + 3]));
        o.z = pack_bf16x2(__uint_as_float(t[8 * v + 4]), __uint_as_float(t[8 * v + 5]));
        o.w = pack_bf16x2(__uint_as_float(t[8 * v + 6]), __uint_as_float(t[8 * v + 7]));
        *reinterpret_cast<u32x4*>(out + col0 + v * 8) = o;
      }
    }
  } else {
#pragma unroll
    for (int i = 0; i < W; ++i)
      if (col0 + i < D) out[col0 + i] = __float2bfloat16(__uint_as_float(t[i]));
  }
}

template <int W>
__device__ __forceinline__ void tmem_ld_cols(uint32_t taddr, uint32_t (&t)[W]) {
  static_assert(W == 16 || W == 32, "chunk width");
  if constexpr (W == 32) ptx::tmem_ld_32x32b_x32(taddr, t);
  else ptx::tmem_ld_32x32b_x16(taddr, t);
}

template <int DP, int NG>
__global__ void __launch_bounds__(NG * 128 + 64, 1)
fa2_bwd_wg_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                  const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmdO,
                  const BwdParams p) {
  using C = BwdCfg<DP>;
  constexpr int TMA_WARP = NG * 4, MMA_WARP = NG * 4 + 1;
  constexpr int SC = BN / NG;              // S / dP columns per group
  constexpr int QC = DP / NG;              // dQ / dK / dV columns per group
  constexpr int W = QC >= 32 ? 32 : 16;    // TMEM chunk width for the dQ / dK / dV columns
  static_assert(SC % 32 == 0 && QC % W == 0, "column split");
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sK = smem;
  uint8_t* sV = sK + C::TILE_BYTES;
  uint8_t* sStage = sV + C::TILE_BYTES;
  uint8_t* sP = sStage + C::STAGES * 2 * C::TILE_BYTES;
  uint8_t* sdS = sP + C::PS_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sdS + C::PS_BYTES);
  uint64_t* kv_full = bars;
  uint64_t* qdo_full = bars + 1;
  uint64_t* qdo_empty = qdo_full + C::STAGES;
  uint64_t* sdp_full = qdo_empty + C::STAGES;
  uint64_t* pds_full = sdp_full + 1;
  uint64_t* dq_full = pds_full + 1;
  uint64_t* dq_empty = dq_full + 1;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(dq_empty + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int j = blockIdx.x, h = blockIdx.y, b = blockIdx.z;
  const int kv0 = j * BN;
  const int i_start = p.causal ? (kv0 / BM) : 0;
  const int n_iter = max(0, p.num_q_tiles - i_start);

  if (warp == TMA_WARP && ptx::elect_one()) {
    ptx::tma_prefetch_desc(&tmQ);
    ptx::tma_prefetch_desc(&tmK);
    ptx::tma_prefetch_desc(&tmV);
    ptx::tma_prefetch_desc(&tmdO);
    ptx::mbar_init(kv_full, 1);
    for (int s = 0; s < C::STAGES; ++s) {
      ptx::mbar_init(&qdo_full[s], 1);
      ptx::mbar_init(&qdo_empty[s], 1);
    }
    ptx::mbar_init(sdp_full, 1);
    ptx::mbar_init(pds_full, 4 * NG);  // one arrive per softmax warp
    ptx::mbar_init(dq_full, 1);
    ptx::mbar_init(dq_empty, 4 * NG);
    ptx::fence_mbar_init();
  }
  if (warp == MMA_WARP) {
    ptx::tmem_alloc(tmem_slot, C::TMEM_COLS);
    ptx::tmem_relinquish();
  }
  ptx::griddep_launch();
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  ptx::griddep_wait();

  if (warp == TMA_WARP) {
    // ===================================== TMA producer =====================================
    if (ptx::elect_one() && n_iter > 0) {
      ptx::mbar_expect_tx(kv_full, 2 * C::TILE_BYTES);
#pragma unroll
      for (int blk = 0; blk < C::NBLK; ++blk) {
        ptx::tma_load_4d(sK + blk * (BN * 128), &tmK, kv_full, blk * 64, kv0, h, b, ptx::kEvictFirst);
        ptx::tma_load_4d(sV + blk * (BN * 128), &tmV, kv_full, blk * 64, kv0, h, b, ptx::kEvictFirst);
      }
      int stage = 0;
      uint32_t phase = 0;
      for (int it = 0; it < n_iter; ++it) {
        const int q0 = (i_start + it) * BM;
        ptx::mbar_wait(&qdo_empty[stage], phase ^ 1);
        ptx::mbar_expect_tx(&qdo_full[stage], 2 * C::TILE_BYTES);
        uint8_t* dq_ = sStage + stage * 2 * C::TILE_BYTES;
        uint8_t* ddo = dq_ + C::TILE_BYTES;
#pragma unroll
        for (int blk = 0; blk < C::NBLK; ++blk) {
          ptx::tma_load_4d(dq_ + blk * (BM * 128), &tmQ, &qdo_full[stage], blk * 64, q0, h, b, ptx::kEvictLast);
          ptx::tma_load_4d(ddo + blk * (BM * 128), &tmdO, &qdo_full[stage], blk * 64, q0, h, b, ptx::kEvictLast);
        }
        if (++stage == C::STAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == MMA_WARP) {
    // ===================================== MMA issuer =====================================
    if (ptx::elect_one() && n_iter > 0) {
      constexpr uint32_t idesc_s = ptx::make_idesc(BM, BN, false, false);
      constexpr uint32_t idesc_dkv = ptx::make_idesc(BN, DP, true, true);
      constexpr uint32_t idesc_dq = ptx::make_idesc(BM, DP, false, true);
      const uint32_t k_addr = ptx::smem_u32(sK), v_addr = ptx::smem_u32(sV);
      const uint32_t p_addr = ptx::smem_u32(sP), ds_addr = ptx::smem_u32(sdS);
      const uint32_t st_addr = ptx::smem_u32(sStage);
      ptx::mbar_wait(kv_full, 0);
      int stage = 0;
      uint32_t phase = 0;
      for (int it = 0; it < n_iter; ++it) {
        const uint32_t q_addr = st_addr + stage * 2 * C::TILE_BYTES;
        const uint32_t do_addr = q_addr + C::TILE_BYTES;
        ptx::mbar_wait(&qdo_full[stage], phase);
        if (C::ALIAS && it > 0) ptx::mbar_wait(dq_empty, (it - 1) & 1);
        ptx::tc_fence_after();
#pragma unroll
        for (int kk = 0; kk < DP / 16; ++kk) {
          const uint32_t off = (kk >> 2) * (128 * 128) + (kk & 3) * 32;
          ptx::mma_f16_ss(tmem_base + C::COL_S, ptx::make_smem_desc(q_addr + off, 16, 1024),
                          ptx::make_smem_desc(k_addr + off, 16, 1024), idesc_s, kk != 0);
        }
#pragma unroll
        for (int kk = 0; kk < DP / 16; ++kk) {
          const uint32_t off = (kk >> 2) * (128 * 128) + (kk & 3) * 32;
          ptx::mma_f16_ss(tmem_base + C::COL_DP, ptx::make_smem_desc(do_addr + off, 16, 1024),
                          ptx::make_smem_desc(v_addr + off, 16, 1024), idesc_s, kk != 0);
        }
        ptx::mma_commit(sdp_full);
        ptx::mbar_wait(pds_full, it & 1);
        ptx::tc_fence_after();
#pragma unroll
        for (int kk = 0; kk < BM / 16; ++kk) {
          const uint32_t acc = (it > 0 || kk > 0) ? 1u : 0u;
          ptx::mma_f16_ss(tmem_base + C::COL_DV, ptx::make_smem_desc(p_addr + kk * 2048, BM * 128, 1024),
                          ptx::make_smem_desc(do_addr + kk * 2048, BM * 128, 1024), idesc_dkv, acc);
        }
#pragma unroll
        for (int kk = 0; kk < BM / 16; ++kk) {
          const uint32_t acc = (it > 0 || kk > 0) ? 1u : 0u;
          ptx::mma_f16_ss(tmem_base + C::COL_DK, ptx::make_smem_desc(ds_addr + kk * 2048, BM * 128, 1024),
                          ptx::make_smem_desc(q_addr + kk * 2048, BM * 128, 1024), idesc_dkv, acc);
        }
        if (!C::ALIAS && it > 0) {
          ptx::mbar_wait(dq_empty, (it - 1) & 1);
          ptx::tc_fence_after();
        }
#pragma unroll
        for (int kk = 0; kk < BN / 16; ++kk) {
          const uint32_t a_off = (kk >> 2) * (BM * 128) + (kk & 3) * 32;
          ptx::mma_f16_ss(tmem_base + C::COL_DQ, ptx::make_smem_desc(ds_addr + a_off, 16, 1024),
                          ptx::make_smem_desc(k_addr + kk * 2048, BN * 128, 1024), idesc_dq, kk != 0);
        }
        ptx::mma_commit(&qdo_empty[stage]);
        ptx::mma_commit(dq_full);
        if (++stage == C::STAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else {
    // ===================================== NG softmax / dS warpgroups =====================================
    const int wq = warp & 3;    // TMEM lane quadrant
    const int grp = warp >> 2;  // column group
    const int r = wq * 32 + lane;
    const uint32_t lane_base = static_cast<uint32_t>(wq * 32) << 16;
    const long long bh = (long long)b * p.H + h;
    float lse_next = 0.f, dl_next = 0.f;
    if (n_iter > 0) {
      const int row0 = i_start * BM + r;
      if (row0 < p.Nq) { lse_next = p.lse[bh * p.Nq + row0]; dl_next = p.delta[bh * p.Nq + row0]; }
    }
    for (int it = 0; it < n_iter; ++it) {
      const int q0 = (i_start + it) * BM;
      const int row = q0 + r;
      const bool row_ok = row < p.Nq;
      const float L2 = row_ok ? lse_next * kLog2e : 0.f;
      const float dl = row_ok ? dl_next : 0.f;
      if (it + 1 < n_iter && row + BM < p.Nq) {
        lse_next = p.lse[bh * p.Nq + row + BM];
        dl_next = p.delta[bh * p.Nq + row + BM];
      }
      const bool need_mask = (kv0 + BN > p.Nk) || (p.causal && kv0 + BN - 1 > q0);
      ptx::mbar_wait(sdp_full, it & 1);
      ptx::tc_fence_after();
      int lim = p.Nk - kv0;
      if (p.causal) lim = min(lim, row - kv0 + 1);
      if (!row_ok) lim = 0;
#pragma unroll
      for (int cc = 0; cc < SC / 32; ++cc) {
        const int c = grp * (SC / 32) + cc;
        uint32_t s[32], dp[32];
        ptx::tmem_ld_32x32b_x32(tmem_base + C::COL_S + lane_base + c * 32, s);
        ptx::tmem_ld_32x32b_x32(tmem_base + C::COL_DP + lane_base + c * 32, dp);
        ptx::tmem_ld_wait();
        const bool whole = (!need_mask && row_ok) || lim >= (c + 1) * 32;
#pragma unroll
        for (int v = 0; v < 4; ++v) {
          float e[8], d[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) e[i] = exp2f(fmaf(__uint_as_float(s[8 * v + i]), p.scale_log2, -L2));
          if (!whole) {
#pragma unroll
            for (int i = 0; i < 8; ++i) e[i] = (c * 32 + 8 * v + i < lim) ? e[i] : 0.f;
          }
#pragma unroll
          for (int i = 0; i < 8; ++i) d[i] = e[i] * (__uint_as_float(dp[8 * v + i]) - dl) * p.scale;
          const uint32_t off = swz_off_b(r, c * 32 + v * 8, BM);
          *reinterpret_cast<u32x4*>(sP + off) = u32x4{pack_bf16x2(e[0], e[1]), pack_bf16x2(e[2], e[3]),
                                                       pack_bf16x2(e[4], e[5]), pack_bf16x2(e[6], e[7])};
          *reinterpret_cast<u32x4*>(sdS + off) = u32x4{pack_bf16x2(d[0], d[1]), pack_bf16x2(d[2], d[3]),
                                                        pack_bf16x2(d[4], d[5]), pack_bf16x2(d[6], d[7])};
        }
      }
      ptx::tc_fence_before();
      ptx::fence_proxy_async();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(pds_full);
      // ---- drain this group's columns of the dQ partial into the fp32 accumulator
      ptx::mbar_wait(dq_full, it & 1);
      ptx::tc_fence_after();
      float* dq_row = p.dq_acc + (bh * p.Nq + row) * (long long)p.D;
#pragma unroll
      for (int cc = 0; cc < QC / W; ++cc) {
        const int col0 = grp * QC + cc * W;
        uint32_t t[W];
        tmem_ld_cols<W>(tmem_base + C::COL_DQ + lane_base + col0, t);
        ptx::tmem_ld_wait();
        if (row_ok) {
          if ((p.D & 3) == 0) {
#pragma unroll
            for (int v = 0; v < W / 4; ++v) {
              const int col = col0 + v * 4;
              if (col < p.D)
                asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dq_row + col),
                             "f"(__uint_as_float(t[4 * v + 0])), "f"(__uint_as_float(t[4 * v + 1])),
                             "f"(__uint_as_float(t[4 * v + 2])), "f"(__uint_as_float(t[4 * v + 3]))
                             : "memory");
            }
          } else {
#pragma unroll
            for (int i = 0; i < W; ++i)
              if (col0 + i < p.D) atomicAdd(dq_row + col0 + i, __uint_as_float(t[i]));
          }
        }
      }
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(dq_empty);
    }

    // ---- epilogue: this group's columns of dK_j and dV_j (row r of the tile = key kv0 + r)
    const int key = kv0 + r;
    if (n_iter > 0) ptx::tc_fence_after();
    const int esz = p.out_f32 ? 4 : 2;
    char* dk_row = reinterpret_cast<char*>(p.dk) + ((long long)b * p.dk_sb + (long long)h * p.dk_sh + (long long)key * p.dk_sr) * esz;
    char* dv_row = reinterpret_cast<char*>(p.dv) + ((long long)b * p.dv_sb + (long long)h * p.dv_sh + (long long)key * p.dv_sr) * esz;
#pragma unroll
    for (int which = 0; which < 2; ++which) {
      char* out = which == 0 ? dv_row : dk_row;
      const uint32_t colbase = which == 0 ? C::COL_DV : C::COL_DK;
#pragma unroll
      for (int cc = 0; cc < QC / W; ++cc) {
        const int col0 = grp * QC + cc * W;
        uint32_t t[W];
        if (n_iter > 0) {
          tmem_ld_cols<W>(tmem_base + colbase + lane_base + col0, t);
          ptx::tmem_ld_wait();
        } else {
#pragma unroll
          for (int i = 0; i < W; ++i) t[i] = 0u;
        }
        if (which == 1 && p.rope_epilogue && key < p.Nk) {
          // rotate the pairs of this chunk back by -angle(position = key index): W/2 contiguous angles
          const int half = p.D >> 1;
          const float* cosr = p.rope_cache + (size_t)key * half + (col0 >> 1);
          const float* sinr = p.rope_cache + (size_t)p.rope_ctx * half + (size_t)key * half + (col0 >> 1);
          f32x4 cs[W / 8], sn[W / 8];
#pragma unroll
          for (int q = 0; q < W / 8; ++q) {
            cs[q] = *reinterpret_cast<const f32x4*>(cosr + 4 * q);
            sn[q] = *reinterpret_cast<const f32x4*>(sinr + 4 * q);
          }
#pragma unroll
          for (int q = 0; q < W / 8; ++q) {
            const float c4[4] = {cs[q].x, cs[q].y, cs[q].z, cs[q].w}, s4[4] = {sn[q].x, sn[q].y, sn[q].z, sn[q].w};
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const int i = 4 * q + e;
              const float a = __uint_as_float(t[2 * i]), bb = __uint_as_float(t[2 * i + 1]);
              t[2 * i] = __float_as_uint(c4[e] * a + s4[e] * bb);
              t[2 * i + 1] = __float_as_uint(c4[e] * bb - s4[e] * a);
            }
          }
        }
        if (key < p.Nk) store_dkv_chunk<W>(out, p.out_f32 != 0, col0, p.D, t);
      }
    }
  }

  ptx::tc_fence_before();
  __syncthreads();
  if (warp == MMA_WARP) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, C::TMEM_COLS);
  }
}

// Same column split, with the tensor-core work re-ordered so that it overlaps the softmax warps (head dim <= 64, the
// default there: NG = 4): after P_it / dS_it are published the MMA warp issues S and dP of tile it+1 BEFORE
// dV, dK, dQ of tile it; the softmax warps turn S/dP(it+1) into packed P/dS registers while those three GEMMs run,
// and only wait for them (dq_full) before overwriting the shared P / dS tiles; the dQ partial of tile it is drained
// after P/dS(it+1) have been published.
template <int DP, int NG>
__global__ void __launch_bounds__(NG * 128 + 64, 1)
fa2_bwd_wgr_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                  const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmdO,
                  const BwdParams p) {
  using C = BwdCfg<DP>;
  constexpr int TMA_WARP = NG * 4, MMA_WARP = NG * 4 + 1;
  constexpr int SC = BN / NG;              // S / dP columns per group
  constexpr int QC = DP / NG;              // dQ / dK / dV columns per group
  constexpr int W = QC >= 32 ? 32 : 16;    // TMEM chunk width for the dQ / dK / dV columns
  static_assert(SC % 32 == 0 && QC % W == 0, "column split");
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sK = smem;
  uint8_t* sV = sK + C::TILE_BYTES;
  uint8_t* sStage = sV + C::TILE_BYTES;
  uint8_t* sP = sStage + C::STAGES * 2 * C::TILE_BYTES;
  uint8_t* sdS = sP + C::PS_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sdS + C::PS_BYTES);
  uint64_t* kv_full = bars;
  uint64_t* qdo_full = bars + 1;
  uint64_t* qdo_empty = qdo_full + C::STAGES;
  uint64_t* sdp_full = qdo_empty + C::STAGES;
  uint64_t* pds_full = sdp_full + 1;
  uint64_t* dq_full = pds_full + 1;
  uint64_t* dq_empty = dq_full + 1;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(dq_empty + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int j = blockIdx.x, h = blockIdx.y, b = blockIdx.z;
  const int kv0 = j * BN;
  const int i_start = p.causal ? (kv0 / BM) : 0;
  const int n_iter = max(0, p.num_q_tiles - i_start);

  if (warp == TMA_WARP && ptx::elect_one()) {
    ptx::tma_prefetch_desc(&tmQ);
    ptx::tma_prefetch_desc(&tmK);
    ptx::tma_prefetch_desc(&tmV);
    ptx::tma_prefetch_desc(&tmdO);
    ptx::mbar_init(kv_full, 1);
    for (int s = 0; s < C::STAGES; ++s) {
      ptx::mbar_init(&qdo_full[s], 1);
      ptx::mbar_init(&qdo_empty[s], 1);
    }
    ptx::mbar_init(sdp_full, 1);
    ptx::mbar_init(pds_full, 4 * NG);  // one arrive per softmax warp
    ptx::mbar_init(dq_full, 1);
    ptx::mbar_init(dq_empty, 4 * NG);
    ptx::fence_mbar_init();
  }
  if (warp == MMA_WARP) {
    ptx::tmem_alloc(tmem_slot, C::TMEM_COLS);
    ptx::tmem_relinquish();
  }
  ptx::griddep_launch();
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  ptx::griddep_wait();

  if (warp == TMA_WARP) {
    // ===================================== TMA producer =====================================
    if (ptx::elect_one() && n_iter > 0) {
      ptx::mbar_expect_tx(kv_full, 2 * C::TILE_BYTES);
#pragma unroll
      for (int blk = 0; blk < C::NBLK; ++blk) {
        ptx::tma_load_4d(sK + blk * (BN * 128), &tmK, kv_full, blk * 64, kv0, h, b, ptx::kEvictFirst);
        ptx::tma_load_4d(sV + blk * (BN * 128), &tmV, kv_full, blk * 64, kv0, h, b, ptx::kEvictFirst);
      }
      int stage = 0;
      uint32_t phase = 0;
      for (int it = 0; it < n_iter; ++it) {
        const int q0 = (i_start + it) * BM;
        ptx::mbar_wait(&qdo_empty[stage], phase ^ 1);
        ptx::mbar_expect_tx(&qdo_full[stage], 2 * C::TILE_BYTES);
        uint8_t* dq_ = sStage + stage * 2 * C::TILE_BYTES;
        uint8_t* ddo = dq_ + C::TILE_BYTES;
#pragma unroll
        for (int blk = 0; blk < C::NBLK; ++blk) {
          ptx::tma_load_4d(dq_ + blk * (BM * 128), &tmQ, &qdo_full[stage], blk * 64, q0, h, b, ptx::kEvictLast);
          ptx::tma_load_4d(ddo + blk * (BM * 128), &tmdO, &qdo_full[stage], blk * 64, q0, h, b, ptx::kEvictLast);
        }
        if (++stage == C::STAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == MMA_WARP) {
    // ===================================== MMA issuer =====================================
    if (ptx::elect_one() && n_iter > 0) {
      constexpr uint32_t idesc_s = ptx::make_idesc(BM, BN, false, false);
      constexpr uint32_t idesc_dkv = ptx::make_idesc(BN, DP, true, true);
      constexpr uint32_t idesc_dq = ptx::make_idesc(BM, DP, false, true);
      const uint32_t k_addr = ptx::smem_u32(sK), v_addr = ptx::smem_u32(sV);
      const uint32_t p_addr = ptx::smem_u32(sP), ds_addr = ptx::smem_u32(sdS);
      const uint32_t st_addr = ptx::smem_u32(sStage);
      ptx::mbar_wait(kv_full, 0);
      // S and dP of tile `t` from stage t % STAGES (the stage's n-th use has parity n & 1)
      auto issue_sdp = [&](int t) {
        const int st = t % C::STAGES;
        const uint32_t q_addr = st_addr + st * 2 * C::TILE_BYTES;
        const uint32_t do_addr = q_addr + C::TILE_BYTES;
        ptx::mbar_wait(&qdo_full[st], (uint32_t)((t / C::STAGES) & 1));
        ptx::tc_fence_after();
#pragma unroll
        for (int kk = 0; kk < DP / 16; ++kk) {
          const uint32_t off = (kk >> 2) * (128 * 128) + (kk & 3) * 32;
          ptx::mma_f16_ss(tmem_base + C::COL_S, ptx::make_smem_desc(q_addr + off, 16, 1024),
                          ptx::make_smem_desc(k_addr + off, 16, 1024), idesc_s, kk != 0);
        }
#pragma unroll
        for (int kk = 0; kk < DP / 16; ++kk) {
          const uint32_t off = (kk >> 2) * (128 * 128) + (kk & 3) * 32;
          ptx::mma_f16_ss(tmem_base + C::COL_DP, ptx::make_smem_desc(do_addr + off, 16, 1024),
                          ptx::make_smem_desc(v_addr + off, 16, 1024), idesc_s, kk != 0);
        }
        ptx::mma_commit(sdp_full);
      };
      issue_sdp(0);
      for (int it = 0; it < n_iter; ++it) {
        const int st = it % C::STAGES;
        const uint32_t q_addr = st_addr + st * 2 * C::TILE_BYTES;
        const uint32_t do_addr = q_addr + C::TILE_BYTES;
        ptx::mbar_wait(pds_full, it & 1);  // P_it, dS_it in smem; S_it / dP_it fully read out of tensor memory
        ptx::tc_fence_after();
        // next tile's S / dP first: the softmax warps work on them while dV, dK, dQ of this tile run
        if (it + 1 < n_iter) issue_sdp(it + 1);
#pragma unroll
        for (int kk = 0; kk < BM / 16; ++kk) {
          const uint32_t acc = (it > 0 || kk > 0) ? 1u : 0u;
          ptx::mma_f16_ss(tmem_base + C::COL_DV, ptx::make_smem_desc(p_addr + kk * 2048, BM * 128, 1024),
                          ptx::make_smem_desc(do_addr + kk * 2048, BM * 128, 1024), idesc_dkv, acc);
        }
#pragma unroll
        for (int kk = 0; kk < BM / 16; ++kk) {
          const uint32_t acc = (it > 0 || kk > 0) ? 1u : 0u;
          ptx::mma_f16_ss(tmem_base + C::COL_DK, ptx::make_smem_desc(ds_addr + kk * 2048, BM * 128, 1024),
                          ptx::make_smem_desc(q_addr + kk * 2048, BM * 128, 1024), idesc_dkv, acc);
        }
        if (it > 0) {
          ptx::mbar_wait(dq_empty, (it - 1) & 1);  // the previous dQ partial has been drained
          ptx::tc_fence_after();
        }
#pragma unroll
        for (int kk = 0; kk < BN / 16; ++kk) {
          const uint32_t a_off = (kk >> 2) * (BM * 128) + (kk & 3) * 32;
          ptx::mma_f16_ss(tmem_base + C::COL_DQ, ptx::make_smem_desc(ds_addr + a_off, 16, 1024),
                          ptx::make_smem_desc(k_addr + kk * 2048, BN * 128, 1024), idesc_dq, kk != 0);
        }
        ptx::mma_commit(&qdo_empty[st]);
        ptx::mma_commit(dq_full);
      }
    }
  } else {
    // ===================================== NG softmax / dS warpgroups =====================================
    const int wq = warp & 3;    // TMEM lane quadrant
    const int grp = warp >> 2;  // column group
    const int r = wq * 32 + lane;
    const uint32_t lane_base = static_cast<uint32_t>(wq * 32) << 16;
    const long long bh = (long long)b * p.H + h;
    float lse_next = 0.f, dl_next = 0.f;
    if (n_iter > 0) {
      const int row0 = i_start * BM + r;
      if (row0 < p.Nq) { lse_next = p.lse[bh * p.Nq + row0]; dl_next = p.delta[bh * p.Nq + row0]; }
    }
    // fp32 red.add of this group's columns of a dQ partial (query row `qrow`) into the accumulator
    auto drain_dq = [&](int qrow, int it_done) {
      (void)it_done;
      const bool ok = qrow < p.Nq;
      float* dq_row = p.dq_acc + (bh * p.Nq + qrow) * (long long)p.D;
#pragma unroll
      for (int cc = 0; cc < QC / W; ++cc) {
        const int col0 = grp * QC + cc * W;
        uint32_t t[W];
        tmem_ld_cols<W>(tmem_base + C::COL_DQ + lane_base + col0, t);
        ptx::tmem_ld_wait();
        if (ok) {
          if ((p.D & 3) == 0) {
#pragma unroll
            for (int v = 0; v < W / 4; ++v) {
              const int col = col0 + v * 4;
              if (col < p.D)
                asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dq_row + col),
                             "f"(__uint_as_float(t[4 * v + 0])), "f"(__uint_as_float(t[4 * v + 1])),
                             "f"(__uint_as_float(t[4 * v + 2])), "f"(__uint_as_float(t[4 * v + 3]))
                             : "memory");
            }
          } else {
#pragma unroll
            for (int i = 0; i < W; ++i)
              if (col0 + i < p.D) atomicAdd(dq_row + col0 + i, __uint_as_float(t[i]));
          }
        }
      }
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(dq_empty);
    };
    for (int it = 0; it < n_iter; ++it) {
      const int q0 = (i_start + it) * BM;
      const int row = q0 + r;
      const bool row_ok = row < p.Nq;
      const float L2 = row_ok ? lse_next * kLog2e : 0.f;
      const float dl = row_ok ? dl_next : 0.f;
      if (it + 1 < n_iter && row + BM < p.Nq) {
        lse_next = p.lse[bh * p.Nq + row + BM];
        dl_next = p.delta[bh * p.Nq + row + BM];
      }
      const bool need_mask = (kv0 + BN > p.Nk) || (p.causal && kv0 + BN - 1 > q0);
      ptx::mbar_wait(sdp_full, it & 1);
      ptx::tc_fence_after();
      int lim = p.Nk - kv0;
      if (p.causal) lim = min(lim, row - kv0 + 1);
      if (!row_ok) lim = 0;
      // ---- P and dS of this tile into registers (packed bf16) while the previous tile's dV / dK / dQ MMAs run
      uint32_t pk[SC / 2], dk_[SC / 2];
#pragma unroll
      for (int cc = 0; cc < SC / 32; ++cc) {
        const int c = grp * (SC / 32) + cc;
        uint32_t s[32], dp[32];
        ptx::tmem_ld_32x32b_x32(tmem_base + C::COL_S + lane_base + c * 32, s);
        ptx::tmem_ld_32x32b_x32(tmem_base + C::COL_DP + lane_base + c * 32, dp);
        ptx::tmem_ld_wait();
        const bool whole = (!need_mask && row_ok) || lim >= (c + 1) * 32;
#pragma unroll
        for (int v = 0; v < 4; ++v) {
          float e[8], d[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) e[i] = exp2f(fmaf(__uint_as_float(s[8 * v + i]), p.scale_log2, -L2));
          if (!whole) {
#pragma unroll
            for (int i = 0; i < 8; ++i) e[i] = (c * 32 + 8 * v + i < lim) ? e[i] : 0.f;
          }
#pragma unroll
          for (int i = 0; i < 8; ++i) d[i] = e[i] * (__uint_as_float(dp[8 * v + i]) - dl) * p.scale;
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            pk[cc * 16 + v * 4 + i] = pack_bf16x2(e[2 * i], e[2 * i + 1]);
            dk_[cc * 16 + v * 4 + i] = pack_bf16x2(d[2 * i], d[2 * i + 1]);
          }
        }
      }
      // ---- the P / dS tiles in shared memory are free once the previous tile's dQ MMA (their last reader) is done
      if (it > 0) {
        ptx::mbar_wait(dq_full, (it - 1) & 1);
        ptx::tc_fence_after();
      }
#pragma unroll
      for (int cc = 0; cc < SC / 32; ++cc) {
        const int c = grp * (SC / 32) + cc;
#pragma unroll
        for (int v = 0; v < 4; ++v) {
          const uint32_t off = swz_off_b(r, c * 32 + v * 8, BM);
          *reinterpret_cast<u32x4*>(sP + off) = u32x4{pk[cc * 16 + v * 4 + 0], pk[cc * 16 + v * 4 + 1],
                                                       pk[cc * 16 + v * 4 + 2], pk[cc * 16 + v * 4 + 3]};
          *reinterpret_cast<u32x4*>(sdS + off) = u32x4{dk_[cc * 16 + v * 4 + 0], dk_[cc * 16 + v * 4 + 1],
                                                        dk_[cc * 16 + v * 4 + 2], dk_[cc * 16 + v * 4 + 3]};
        }
      }
      ptx::tc_fence_before();
      ptx::fence_proxy_async();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(pds_full);
      // ---- drain the PREVIOUS tile's dQ partial (its MMA finished: dq_full(it-1) was observed above)
      if (it > 0) {
        drain_dq(row - BM, it - 1);
      }
    }
    if (n_iter > 0) {  // the last tile's dQ partial
      ptx::mbar_wait(dq_full, (n_iter - 1) & 1);
      ptx::tc_fence_after();
      drain_dq((i_start + n_iter - 1) * BM + r, n_iter - 1);
    }

    // ---- epilogue: this group's columns of dK_j and dV_j (row r of the tile = key kv0 + r)
    const int key = kv0 + r;
    if (n_iter > 0) ptx::tc_fence_after();
    const int esz = p.out_f32 ? 4 : 2;
    char* dk_row = reinterpret_cast<char*>(p.dk) + ((long long)b * p.dk_sb + (long long)h * p.dk_sh + (long long)key * p.dk_sr) * esz;
    char* dv_row = reinterpret_cast<char*>(p.dv) + ((long long)b * p.dv_sb + (long long)h * p.dv_sh + (long long)key * p.dv_sr) * esz;
#pragma unroll
    for (int which = 0; which < 2; ++which) {
      char* out = which == 0 ? dv_row : dk_row;
      const uint32_t colbase = which == 0 ? C::COL_DV : C::COL_DK;
#pragma unroll
      for (int cc = 0; cc < QC / W; ++cc) {
        const int col0 = grp * QC + cc * W;
        uint32_t t[W];
        if (n_iter > 0) {
          tmem_ld_cols<W>(tmem_base + colbase + lane_base + col0, t);
          ptx::tmem_ld_wait();
        } else {
#pragma unroll
          for (int i = 0; i < W; ++i) t[i] = 0u;
        }
        if (which == 1 && p.rope_epilogue && key < p.Nk) {
          // rotate the pairs of this chunk back by -angle(position = key index): W/2 contiguous angles
          const int half = p.D >> 1;
          const float* cosr = p.rope_cache + (size_t)key * half + (col0 >> 1);
          const float* sinr = p.rope_cache + (size_t)p.rope_ctx * half + (size_t)key * half + (col0 >> 1);
          f32x4 cs[W / 8], sn[W / 8];
#pragma unroll
          for (int q = 0; q < W / 8; ++q) {
            cs[q] = *reinterpret_cast<const f32x4*>(cosr + 4 * q);
            sn[q] = *reinterpret_cast<const f32x4*>(sinr + 4 * q);
          }
#pragma unroll
          for (int q = 0; q < W / 8; ++q) {
            const float c4[4] = {cs[q].x, cs[q].y, cs[q].z, cs[q].w}, s4[4] = {sn[q].x, sn[q].y, sn[q].z, sn[q].w};
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const int i = 4 * q + e;
              const float a = __uint_as_float(t[2 * i]), bb = __uint_as_float(t[2 * i + 1]);
              t[2 * i] = __float_as_uint(c4[e] * a + s4[e] * bb);
              t[2 * i + 1] = __float_as_uint(c4[e] * bb - s4[e] * a);
            }
          }
        }
        if (key < p.Nk) store_dkv_chunk<W>(out, p.out_f32 != 0, col0, p.D, t);
      }
    }
  }

  ptx::tc_fence_before();
  __syncthreads();
  if (warp == MMA_WARP) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, C::TMEM_COLS);
  }
}

// delta[b,h,i] = sum_d O[b,h,i,d] * dO[b,h,i,d].  TPR threads per row (a power of two >= D/8), 8 elements = one or two
// 16-byte loads per thread and operand, a TPR-wide shuffle reduction: 32/TPR rows per warp (the first version spent a
// whole warp and 2-element loads on a 64-element row: 17 us for 13 MB).  O / dO are bf16 or fp32 (o_f32 / do_f32); with
// dout_bf16 != nullptr the fp32 dO row is also written out as bf16 in the same pass (contiguous [rows, D]) — the MMA
// operand of the main kernel — so fp32 callers need no separate cast kernel.
template <int TPR>
__global__ void __launch_bounds__(256) fa2_delta_kernel(const void* __restrict__ o, const void* __restrict__ dout,
                                                        float* __restrict__ delta, int H, int Nq, int D, long long o_sb,
                                                        long long o_sh, long long o_sr, long long d_sb, long long d_sh,
                                                        long long d_sr, long long rows, int o_f32, int do_f32,
                                                        __nv_bfloat16* __restrict__ dout_bf16) {
  ptx::griddep_launch();
  ptx::griddep_wait();
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long w = t / TPR;          // row
  const int sub = (int)(t % TPR);       // 8-element group inside the row
  const int d = sub * 8;
  float acc = 0.f;
  if (w < rows && d < D) {
    const unsigned rows_per_b = (unsigned)H * (unsigned)Nq;
    const long long bb = w / rows_per_b;
    const unsigned rem = (unsigned)(w - bb * rows_per_b);
    const unsigned hh = rem / (unsigned)Nq, i = rem - hh * (unsigned)Nq;
    const long long o_off = bb * o_sb + (long long)hh * o_sh + (long long)i * o_sr + d;
    const long long d_off = bb * d_sb + (long long)hh * d_sh + (long long)i * d_sr + d;
    float a[8], g[8];
    if (o_f32) {
      const float* p = reinterpret_cast<const float*>(o) + o_off;
      if ((reinterpret_cast<uintptr_t>(p) & 15) == 0) {
        const f32x4 u = *reinterpret_cast<const f32x4*>(p), v = *reinterpret_cast<const f32x4*>(p + 4);
        a[0] = u.x; a[1] = u.y; a[2] = u.z; a[3] = u.w; a[4] = v.x; a[5] = v.y; a[6] = v.z; a[7] = v.w;
      } else {
#pragma unroll
        for (int e = 0; e < 8; ++e) a[e] = p[e];
      }
    } else {
      const __nv_bfloat16* p = reinterpret_cast<const __nv_bfloat16*>(o) + o_off;
      if ((reinterpret_cast<uintptr_t>(p) & 15) == 0) {
        const u32x4 u = *reinterpret_cast<const u32x4*>(p);
        const uint32_t ww[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
        for (int e = 0; e < 4; ++e) { const float2 f = unpack_bf16x2(ww[e]); a[2 * e] = f.x; a[2 * e + 1] = f.y; }
      } else {
#pragma unroll
        for (int e = 0; e < 8; ++e) a[e] = __bfloat162float(p[e]);
      }
    }
    if (do_f32) {
      const float* p = reinterpret_cast<const float*>(dout) + d_off;
#pragma unroll
      for (int e = 0; e < 8; ++e) g[e] = p[e];
      // delta must see the SAME rounding of dO as the dP = dO·Vᵀ MMA does, or P∘(dP − delta) stops cancelling where it
      // should be exactly zero (a causal row that attends to one key only)
      uint32_t gb[4];
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        gb[e] = pack_bf16x2(g[2 * e], g[2 * e + 1]);
        const float2 f = unpack_bf16x2(gb[e]);
        g[2 * e] = f.x; g[2 * e + 1] = f.y;
      }
      if (dout_bf16) *reinterpret_cast<u32x4*>(dout_bf16 + w * D + d) = u32x4{gb[0], gb[1], gb[2], gb[3]};
    } else {
      const __nv_bfloat16* p = reinterpret_cast<const __nv_bfloat16*>(dout) + d_off;
      if ((reinterpret_cast<uintptr_t>(p) & 15) == 0) {
        const u32x4 u = *reinterpret_cast<const u32x4*>(p);
        const uint32_t ww[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
        for (int e = 0; e < 4; ++e) { const float2 f = unpack_bf16x2(ww[e]); g[2 * e] = f.x; g[2 * e + 1] = f.y; }
      } else {
#pragma unroll
        for (int e = 0; e < 8; ++e) g[e] = __bfloat162float(p[e]);
      }
    }
#pragma unroll
    for (int e = 0; e < 8; ++e) acc += a[e] * g[e];
  }
#pragma unroll
  for (int off = TPR / 2; off > 0; off >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, off);
  if (sub == 0 && w < rows) delta[w] = acc;
}

// 8 consecutive elements of one row per thread (D is a multiple of 8): two 16-byte loads, one 16- or 32-byte store,
// and one row decomposition per thread (the first version did a 64-bit div/mod chain per pair: 42 us for 13 MB).
__global__ void __launch_bounds__(256) fa2_dq_convert_kernel(const float* __restrict__ acc, void* __restrict__ dq, int H,
                                                             int Nq, int D, long long sb, long long sh, long long sr,
                                                             long long total, const float* __restrict__ rope_cache,
                                                             int rope_ctx, int rezero, int out_f32) {
  ptx::griddep_launch();
  ptx::griddep_wait();
  const long long idx = ((long long)blockIdx.x * blockDim.x + threadIdx.x) * 8;
  if (idx >= total) return;
  const int dpr = D >> 3;                                   // 8-element groups per row
  const long long grp = idx >> 3;
  const long long rowi = grp / dpr;
  const int d = (int)(grp - rowi * dpr) << 3;
  const unsigned rows_per_b = (unsigned)H * (unsigned)Nq;
  const long long bb = rowi / rows_per_b;
  const unsigned rem = (unsigned)(rowi - bb * rows_per_b);
  const unsigned hh = rem / (unsigned)Nq, i = rem - hh * (unsigned)Nq;
  const f32x4 v0 = *reinterpret_cast<const f32x4*>(acc + idx), v1 = *reinterpret_cast<const f32x4*>(acc + idx + 4);
  if (rezero) {  // ready for the next call
    *reinterpret_cast<f32x4*>(const_cast<float*>(acc) + idx) = f32x4{0.f, 0.f, 0.f, 0.f};
    *reinterpret_cast<f32x4*>(const_cast<float*>(acc) + idx + 4) = f32x4{0.f, 0.f, 0.f, 0.f};
  }
  float v[8] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
  if (rope_cache != nullptr) {  // un-rotate the pairs (d + 2e, d + 2e + 1) by the angle of position i
    const int half = D >> 1;
    const f32x4 cs = *reinterpret_cast<const f32x4*>(rope_cache + (size_t)i * half + (d >> 1));
    const f32x4 sn = *reinterpret_cast<const f32x4*>(rope_cache + (size_t)rope_ctx * half + (size_t)i * half + (d >> 1));
    const float c4[4] = {cs.x, cs.y, cs.z, cs.w}, s4[4] = {sn.x, sn.y, sn.z, sn.w};
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const float a = v[2 * e], b = v[2 * e + 1];
      v[2 * e] = c4[e] * a + s4[e] * b;
      v[2 * e + 1] = c4[e] * b - s4[e] * a;
    }
  }
  const long long off = bb * sb + (long long)hh * sh + (long long)i * sr + d;
  if (out_f32) {
    float* o = reinterpret_cast<float*>(dq) + off;
    if ((reinterpret_cast<uintptr_t>(o) & 15) == 0) {
      *reinterpret_cast<f32x4*>(o) = f32x4{v[0], v[1], v[2], v[3]};
      *reinterpret_cast<f32x4*>(o + 4) = f32x4{v[4], v[5], v[6], v[7]};
    } else {
#pragma unroll
      for (int e = 0; e < 8; ++e) o[e] = v[e];
    }
  } else {
    __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(dq) + off;
    if ((reinterpret_cast<uintptr_t>(o) & 15) == 0) {
      *reinterpret_cast<u32x4*>(o) = u32x4{pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]), pack_bf16x2(v[4], v[5]),
                                           pack_bf16x2(v[6], v[7])};
    } else {
#pragma unroll
      for (int e = 0; e < 4; ++e) *reinterpret_cast<uint32_t*>(o + 2 * e) = pack_bf16x2(v[2 * e], v[2 * e + 1]);
    }
  }
}

// dK was computed w.r.t. the rotated keys: rotate every (2i, 2i+1) pair back by -angle(position = key index), in place
__global__ void fa2_dk_unrope_kernel(__nv_bfloat16* __restrict__ dk, int H, int Nk, int D, long long sb, long long sh,
                                     long long sr, long long total_pairs, const float* __restrict__ rope_cache,
                                     int rope_ctx) {
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= total_pairs) return;
  const int half = D >> 1;
  const int pi = (int)(idx % half);
  const long long rowi = idx / half;
  const int i = (int)(rowi % Nk);
  const long long bh = rowi / Nk;
  const int hh = (int)(bh % H);
  const long long bb = bh / H;
  uint32_t* ptr = reinterpret_cast<uint32_t*>(dk + bb * sb + hh * sh + (long long)i * sr + 2 * pi);
  const float2 v = unpack_bf16x2(*ptr);
  const float cs = rope_cache[(size_t)i * half + pi];
  const float sn = rope_cache[(size_t)rope_ctx * half + (size_t)i * half + pi];
  *ptr = pack_bf16x2(cs * v.x + sn * v.y, cs * v.y - sn * v.x);
}

template <int DP, int NG>
static int launch_bwd_wg(const CUtensorMap& q, const CUtensorMap& k, const CUtensorMap& v, const CUtensorMap& dO,
                         const BwdParams& p, int B, cudaStream_t s) {
  using C = BwdCfg<DP>;
  auto kern = fa2_bwd_wg_kernel<DP, NG>;
  B200_ENSURE_DYN_SMEM(kern, C::SMEM_BYTES);
  dim3 grid(ceil_div(p.Nk, BN), p.H, B);
  B200_CHECK_CUDA(launch_pdl(kern, grid, dim3(NG * 128 + 64), C::SMEM_BYTES, s, q, k, v, dO, p));
  return 0;
}

template <int NG>
static int launch_bwd_wgr(const CUtensorMap& q, const CUtensorMap& k, const CUtensorMap& v, const CUtensorMap& dO,
                          const BwdParams& p, int B, cudaStream_t s) {
  using C = BwdCfg<64>;
  auto kern = fa2_bwd_wgr_kernel<64, NG>;
  B200_ENSURE_DYN_SMEM(kern, C::SMEM_BYTES);
  dim3 grid(ceil_div(p.Nk, BN), p.H, B);
  B200_CHECK_CUDA(launch_pdl(kern, grid, dim3(NG * 128 + 64), C::SMEM_BYTES, s, q, k, v, dO, p));
  return 0;
}

}  // namespace fa2
}  // namespace b200

// flags: bit 0 = dq_acc is kept zero by the caller between calls (the convert pass re-zeroes it);
//        bit 1 = O is fp32;  bit 2 = dO is fp32 (do_bf16_scratch must then hold B*H*Nq*D bf16: the delta pre-pass
//        fills it and the main kernel reads it);  bit 3 = dQ, dK, dV are written as fp32.
extern "C" int b200_fa2_bwd_ex(const void* q, const void* k, const void* v, const void* o, const float* lse,
                            const void* dout, void* dq, void* dk, void* dv, float* delta, float* dq_acc, int B, int H,
                            int Nq, int Nk, int D, const long long* q_strides, const long long* k_strides,
                            const long long* v_strides, const long long* o_strides, const long long* do_strides,
                            const long long* dq_strides, const long long* dk_strides, const long long* dv_strides,
                            int causal, float scale, const float* rope_cache, int rope_ctx, int flags,
                            void* do_bf16_scratch, void* stream) {
  using namespace b200;
  using namespace b200::fa2;
  B200_REQUIRE(B > 0 && H > 0 && Nq > 0 && Nk > 0, "fa2_bwd: empty problem");
  B200_REQUIRE(rope_cache == nullptr || (Nq <= rope_ctx && Nk <= rope_ctx), "fa2_bwd: sequence exceeds the RoPE cache");
  B200_REQUIRE(D >= 8 && D <= 128 && D % 8 == 0, "fa2_bwd: head dim %d unsupported (multiple of 8, <= 128)", D);
  const int prezeroed = flags & 1;
  const int o_f32 = (flags >> 1) & 1, do_f32 = (flags >> 2) & 1, out_f32 = (flags >> 3) & 1;
  B200_REQUIRE(!do_f32 || do_bf16_scratch != nullptr, "fa2_bwd: fp32 dO needs the bf16 scratch buffer");
  B200_REQUIRE(!(out_f32 && rope_cache), "fa2_bwd: fp32 gradients with fused RoPE un-rotation are not supported");
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  const long long rows = (long long)B * H * Nq;
  {
    const int threads = 256;
    __nv_bfloat16* scratch = reinterpret_cast<__nv_bfloat16*>(do_f32 ? do_bf16_scratch : nullptr);
#define LAUNCH_DELTA(TPRV)                                                                                           \
  B200_CHECK_CUDA(launch_pdl(fa2_delta_kernel<TPRV>, dim3((unsigned)((rows * TPRV + threads - 1) / threads)),         \
                             dim3(threads), 0, s, o, dout, delta, H, Nq, D, o_strides[0], o_strides[1], o_strides[2], \
                             do_strides[0], do_strides[1], do_strides[2], rows, o_f32, do_f32, scratch))
    if (D <= 16) LAUNCH_DELTA(2);
    else if (D <= 32) LAUNCH_DELTA(4);
    else if (D <= 64) LAUNCH_DELTA(8);
    else LAUNCH_DELTA(16);
#undef LAUNCH_DELTA
  }
  if (!prezeroed) B200_CHECK_CUDA(cudaMemsetAsync(dq_acc, 0, (size_t)rows * D * sizeof(float), s));
  CUtensorMap tq, tk, tv, tdo;
  int rc;
  if ((rc = make_bhnd_map(&tq, q, B, H, Nq, D, q_strides, BM))) return rc;
  if ((rc = make_bhnd_map(&tk, k, B, H, Nk, D, k_strides, BN))) return rc;
  if ((rc = make_bhnd_map(&tv, v, B, H, Nk, D, v_strides, BN))) return rc;
  if (do_f32) {
    const long long cs[3] = {(long long)H * Nq * D, (long long)Nq * D, (long long)D};  // contiguous [B,H,Nq,D]
    if ((rc = make_bhnd_map(&tdo, do_bf16_scratch, B, H, Nq, D, cs, BM))) return rc;
  } else {
    if ((rc = make_bhnd_map(&tdo, dout, B, H, Nq, D, do_strides, BM))) return rc;
  }
  BwdParams p;
  p.lse = lse; p.delta = delta; p.dq_acc = dq_acc;
  p.dk = dk; p.dv = dv; p.out_f32 = out_f32;
  p.dk_sb = dk_strides[0]; p.dk_sh = dk_strides[1]; p.dk_sr = dk_strides[2];
  p.dv_sb = dv_strides[0]; p.dv_sh = dv_strides[1]; p.dv_sr = dv_strides[2];
  p.H = H; p.Nq = Nq; p.Nk = Nk; p.D = D; p.causal = causal;
  p.scale = scale; p.scale_log2 = scale * kLog2e;
  p.num_q_tiles = ceil_div(Nq, BM);
  p.rope_cache = rope_cache; p.rope_ctx = rope_ctx;
  p.rope_epilogue = (rope_cache != nullptr && D == 64) ? 1 : 0;
  // measured on B200 (profiles/fa_variants_r2.md): head dim <= 64 — re-ordered pipeline with 4 column groups
  // (16x16384x64 causal: 2.46 ms vs 3.40 for one softmax warpgroup); head dim > 64 — 2 column groups (2.69 vs 2.86)
  rc = D <= 64 ? launch_bwd_wgr<4>(tq, tk, tv, tdo, p, B, s) : launch_bwd_wg<128, 2>(tq, tk, tv, tdo, p, B, s);
  if (rc) return rc;
  {
    const long long total = rows * D;
    const int threads = 256;
    B200_CHECK_CUDA(launch_pdl(fa2_dq_convert_kernel, dim3((unsigned)((total / 8 + threads - 1) / threads)),
                               dim3(threads), 0, s, (const float*)dq_acc, dq, H, Nq, D, dq_strides[0], dq_strides[1],
                               dq_strides[2], total, rope_cache, rope_ctx, prezeroed, out_f32));
    if (rope_cache != nullptr && !p.rope_epilogue) {
      const long long pairs = (long long)B * H * Nk * (D / 2);
      fa2_dk_unrope_kernel<<<(unsigned)((pairs + threads - 1) / threads), threads, 0, s>>>(
          reinterpret_cast<__nv_bfloat16*>(dk), H, Nk, D, dk_strides[0], dk_strides[1], dk_strides[2], pairs,
          rope_cache, rope_ctx);
      B200_LAUNCH_CHECK();
    }
  }
  return 0;
}

extern "C" int b200_fa2_bwd(const void* q, const void* k, const void* v, const void* o, const float* lse,
                            const void* dout, void* dq, void* dk, void* dv, float* delta, float* dq_acc, int B, int H,
                            int Nq, int Nk, int D, const long long* q_strides, const long long* k_strides,
                            const long long* v_strides, const long long* o_strides, const long long* do_strides,
                            const long long* dq_strides, const long long* dk_strides, const long long* dv_strides,
                            int causal, float scale, void* stream) {
  return b200_fa2_bwd_ex(q, k, v, o, lse, dout, dq, dk, dv, delta, dq_acc, B, H, Nq, Nk, D, q_strides, k_strides,
                         v_strides, o_strides, do_strides, dq_strides, dk_strides, dv_strides, causal, scale, nullptr, 0,
                         0, nullptr, stream);
}
