// FlashAttention-2 forward on tcgen05 / TMEM / TMA for sm_100a.
//
// Replaces the reference's Triton kernel (cs336_systems/flash_attention.py:137-266: 16×16 tiles,
// mma.sync, no causal tile skipping) with 128×128 tiles on the 5th-gen tensor cores.
//
// One CTA = one 128-row query tile of one (batch, head); 320 threads:
//   warps 0-7  two softmax warpgroups: warps 0-3 own S columns 0-63, warps 4-7 columns 64-127 of the same 128
//              rows (thread = one TMEM lane = one query row, so row max / sum need no shuffles; the two halves
//              exchange their partial row maximum through shared memory once per tile).  Online softmax in the
//              log2 domain, P (bf16) written to shared memory in the 128B-swizzled K-major UMMA layout.
//   warp 8     TMA producer: Q once, then K_j, V_j, K_j+1, … through a ring of shared-memory slots
//   warp 9     TMEM allocator + MMA issuer: S = Q·K_jᵀ (both K-major), O += P·V_j (V as MN-major B)
// O never leaves tensor memory during the key loop: the PV MMAs accumulate into it, and the exponentials are
// taken against a reference maximum that only moves when the true row maximum exceeds it by 2^8 — the rescale of
// O and l is then a rare warp-uniform TMEM read-modify-write (tcgen05.ld → scale → tcgen05.st) instead of a
// read-back every tile.
// CLUSTER: the CTAs of a 2-CTA thread-block cluster work on neighbouring query tiles of one head and share
// their K/V tiles: each CTA fetches one 64-row half of every common key tile and TMA-multicasts it into both CTAs'
// shared memory (a slot is released by a multicast tcgen05.commit once BOTH CTAs' MMAs have read it), halving the
// L2 → SM traffic for K and V; the extra causal tile of the later query tile is loaded solo.
// Two CTAs are co-resident per SM (head dim ≤ 64), so one CTA's softmax overlaps the other's MMAs.
// Head dims are zero-padded by TMA to DP ∈ {64, 128}; causal tiles beyond the diagonal are skipped.
#include "common.cuh"
#include <cstdlib>
#include "ptx.cuh"
#include "tensormap.h"

namespace b200 {
namespace fa2 {

constexpr int BM = 128;  // query rows per CTA
constexpr int BN = 128;  // keys per iteration
constexpr float kLog2e = 1.4426950408889634f;
constexpr float kLn2 = 0.6931471805599453f;
constexpr int kThreads2 = 320;  // 8 softmax warps, TMA warp (8), MMA warp (9)

struct FwdParams {
  void* o;           // bf16 or fp32 (out_f32)
  float* lse;
  long long o_sb, o_sh, o_sr;
  int H, Nq, Nk, D;
  int causal;
  int out_f32;
  float scale_log2;  // softmax scale * log2(e)
  int num_m_tiles;
};

// byte offset of element (row, col) in a K-major [rows × 128] bf16 tile made of two 64-column blocks,
// each block = rows × 128 B with the 128B swizzle (16-byte chunk index XOR row%8)
__device__ __forceinline__ uint32_t swz_off(int row, int col, int rows) {
  const int blk = col >> 6, c = col & 63;
  return blk * (rows * 128) + row * 128 + ((((c >> 3) ^ (row & 7)) << 4) | ((c & 7) << 1));
}

// Shared-memory plan: Q tile, P tile, a ring of K/V slots, the [buffer][half][row] exchange area for the partial
// row maxima of the two softmax warpgroups, barriers.
template <int DP>
struct Fwd2Cfg {
  static constexpr int NBLK = DP / 64;
  static constexpr int TILE_BYTES = BN * DP * 2;
  static constexpr int P_BYTES = BM * BN * 2;
  static constexpr int SLOTS = DP == 64 ? 3 : 4;
  static constexpr int XCH_BYTES = 2 * 2 * BM * 4;  // [buffer][half][row] fp32
  static constexpr int SMEM_BYTES = TILE_BYTES + P_BYTES + SLOTS * TILE_BYTES + XCH_BYTES + 1024 + 256;
  static constexpr int TMEM_COLS = 256;
  static constexpr int MIN_CTAS = DP == 64 ? 2 : 1;
};

template <int DP, bool CLUSTER>
__global__ void __launch_bounds__(kThreads2, Fwd2Cfg<DP>::MIN_CTAS)
fa2_fwd_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
               const __grid_constant__ CUtensorMap tmV, const __grid_constant__ CUtensorMap tmK64,
               const __grid_constant__ CUtensorMap tmV64, const FwdParams p) {
  using C = Fwd2Cfg<DP>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sQ = smem;
  uint8_t* sP = sQ + C::TILE_BYTES;
  uint8_t* sKV = sP + C::P_BYTES;
  float* xch = reinterpret_cast<float*>(sKV + C::SLOTS * C::TILE_BYTES);
  uint64_t* bars = reinterpret_cast<uint64_t*>(reinterpret_cast<uint8_t*>(xch) + C::XCH_BYTES);
  uint64_t* q_full = bars;
  uint64_t* kv_full = bars + 1;
  uint64_t* kv_empty = kv_full + C::SLOTS;
  uint64_t* s_full = kv_empty + C::SLOTS;
  uint64_t* p_full = s_full + 1;
  uint64_t* o_full = p_full + 1;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(o_full + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int m_tile = p.causal ? (p.num_m_tiles - 1 - (int)blockIdx.x) : (int)blockIdx.x;
  const int h = blockIdx.y, b = blockIdx.z;
  const int m0 = m_tile * BM;

  int n_kv = ceil_div(p.Nk, BN);
  if (p.causal) {
    const int lim = ceil_div(min(m0 + BM, p.Nk), BN);
    n_kv = min(n_kv, lim);
  }
  // CLUSTER: the two CTAs of a cluster work on neighbouring query tiles of the same (batch, head).  Key tiles both
  // need (j < n_shared) are fetched cooperatively: each CTA loads one 64-row half and multicasts it into both CTAs'
  // shared memory; the extra causal tile of the later query tile is loaded by its CTA alone.
  uint32_t cr = 0;
  int n_shared = 0;
  if constexpr (CLUSTER) {
    cr = ptx::cluster_ctarank();
    const int peer_bx = (int)blockIdx.x ^ 1;
    const int peer_m0 = (p.causal ? (p.num_m_tiles - 1 - peer_bx) : peer_bx) * BM;
    int n_kv_peer = ceil_div(p.Nk, BN);
    if (p.causal) n_kv_peer = min(n_kv_peer, ceil_div(min(peer_m0 + BM, p.Nk), BN));
    n_shared = min(n_kv, n_kv_peer);
  }

  if (warp == 8 && ptx::elect_one()) {
    ptx::tma_prefetch_desc(&tmQ);
    ptx::tma_prefetch_desc(&tmK);
    ptx::tma_prefetch_desc(&tmV);
    if constexpr (CLUSTER) {
      ptx::tma_prefetch_desc(&tmK64);
      ptx::tma_prefetch_desc(&tmV64);
    }
    ptx::mbar_init(q_full, 1);
    for (int s = 0; s < C::SLOTS; ++s) {
      ptx::mbar_init(&kv_full[s], 1);
      ptx::mbar_init(&kv_empty[s], CLUSTER ? 2 : 1);  // a shared slot is free when BOTH CTAs' MMAs have read it
    }
    ptx::mbar_init(s_full, 1);
    ptx::mbar_init(p_full, 8);  // one arrive per softmax warp
    ptx::mbar_init(o_full, 1);
    ptx::fence_mbar_init();
  }
  if (warp == 9) {
    ptx::tmem_alloc(tmem_slot, C::TMEM_COLS);
    ptx::tmem_relinquish();
  }
  ptx::griddep_launch();
  ptx::tc_fence_before();
  __syncthreads();
  if constexpr (CLUSTER) ptx::cluster_sync();  // the peer's barriers are initialised before anything of ours can signal them
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  ptx::griddep_wait();
  const uint32_t tmem_S = tmem_base;
  const uint32_t tmem_O = tmem_base + BN;

  if (warp == 8) {
    // ===================================== TMA producer =====================================
    if (ptx::elect_one()) {
      ptx::mbar_expect_tx(q_full, C::TILE_BYTES);
#pragma unroll
      for (int blk = 0; blk < C::NBLK; ++blk)
        ptx::tma_load_4d(sQ + blk * (BM * 128), &tmQ, q_full, blk * 64, m0, h, b, ptx::kEvictFirst);
      int slot = 0;
      uint32_t phase = 0;
      for (int j = 0; j < n_kv; ++j) {
#pragma unroll
        for (int kv = 0; kv < 2; ++kv) {  // K_j then V_j
          ptx::mbar_wait(&kv_empty[slot], phase ^ 1);
          ptx::mbar_expect_tx(&kv_full[slot], C::TILE_BYTES);  // (cluster: our half + the peer's half)
          uint8_t* dst = sKV + slot * C::TILE_BYTES;
          if (CLUSTER && j < n_shared) {
#pragma unroll
            for (int blk = 0; blk < C::NBLK; ++blk)
              ptx::tma_load_4d_mcast(dst + blk * (BN * 128) + cr * (64 * 128), kv == 0 ? &tmK64 : &tmV64, &kv_full[slot],
                                     blk * 64, j * BN + (int)cr * 64, h, b, (uint16_t)0x3, ptx::kEvictLast);
          } else {
#pragma unroll
            for (int blk = 0; blk < C::NBLK; ++blk)
              ptx::tma_load_4d(dst + blk * (BN * 128), kv == 0 ? &tmK : &tmV, &kv_full[slot], blk * 64, j * BN, h, b,
                               ptx::kEvictLast);
          }
          if (++slot == C::SLOTS) { slot = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 9) {
    // ===================================== MMA issuer =====================================
    if (ptx::elect_one() && n_kv > 0) {
      constexpr uint32_t idesc_qk = ptx::make_idesc(BM, BN, false, false);
      constexpr uint32_t idesc_pv = ptx::make_idesc(BM, DP, false, true);
      const uint32_t q_addr = ptx::smem_u32(sQ), p_addr = ptx::smem_u32(sP), kv_addr = ptx::smem_u32(sKV);
      int slot = 0;
      uint32_t phase = 0;
      int jq = 0;  // key tile the next S = Q K^T uses
      auto release = [&](int tile) {  // the slot just read is free in both CTAs (shared tile) or only here (solo)
        if (CLUSTER && tile < n_shared) ptx::mma_commit_mcast(&kv_empty[slot], (uint16_t)0x3);
        else ptx::mma_commit(&kv_empty[slot]);
      };
      auto issue_qk = [&]() {
        ptx::mbar_wait(&kv_full[slot], phase);
        ptx::tc_fence_after();
        const uint32_t k_addr = kv_addr + slot * C::TILE_BYTES;
#pragma unroll
        for (int kk = 0; kk < DP / 16; ++kk) {
          const uint32_t off = (kk >> 2) * (BM * 128) + (kk & 3) * 32;
          ptx::mma_f16_ss(tmem_S, ptx::make_smem_desc(q_addr + off, 16, 1024),
                          ptx::make_smem_desc(k_addr + off, 16, 1024), idesc_qk, kk != 0);
        }
        release(jq++);
        ptx::mma_commit(s_full);
        if (++slot == C::SLOTS) { slot = 0; phase ^= 1; }
      };
      ptx::mbar_wait(q_full, 0);
      issue_qk();
      for (int j = 0; j < n_kv; ++j) {
        ptx::mbar_wait(p_full, j & 1);  // P_j in smem; S_j and the previous PV partial fully consumed
        ptx::tc_fence_after();
        ptx::mbar_wait(&kv_full[slot], phase);
        ptx::tc_fence_after();
        const uint32_t v_addr = kv_addr + slot * C::TILE_BYTES;
#pragma unroll
        for (int kk = 0; kk < BN / 16; ++kk) {
          const uint32_t a_off = (kk >> 2) * (BM * 128) + (kk & 3) * 32;
          ptx::mma_f16_ss(tmem_O, ptx::make_smem_desc(p_addr + a_off, 16, 1024),
                          ptx::make_smem_desc(v_addr + kk * 2048, BN * 128, 1024), idesc_pv, (j > 0 || kk != 0) ? 1u : 0u);
        }
        release(j);
        ptx::mma_commit(o_full);
        if (++slot == C::SLOTS) { slot = 0; phase ^= 1; }
        if (j + 1 < n_kv) issue_qk();
      }
    }
  } else {
    // ===================================== two softmax warpgroups =====================================
    constexpr int OC = DP / 2;           // O columns accumulated by one half
    const int wq = warp & 3;             // TMEM lane quadrant
    const int half = warp >> 2;          // 0: S columns 0-63, 1: S columns 64-127
    const int r = wq * 32 + lane;        // row within the tile == TMEM lane
    const int row = m0 + r;
    const uint32_t lane_base = static_cast<uint32_t>(wq * 32) << 16;
    // O stays in tensor memory for the whole key loop (the PV MMAs accumulate); m_ref is the maximum the exponentials
    // are taken against.  It only moves when the true row maximum exceeds it by more than kTau (log2 units), so the
    // rescaling of O and l — a TMEM read-modify-write — happens in a small fraction of the tiles; P then reaches at
    // most 2^kTau, well inside bf16 / fp32 range, and the result is exact because L = m_ref + log2(l) whatever m_ref.
    constexpr float kTau = 8.f;
    float m_ref = -INFINITY, l_part = 0.f;

    for (int j = 0; j < n_kv; ++j) {
      const int kv0 = j * BN;
      const bool need_mask = (kv0 + BN > p.Nk) || (p.causal && kv0 + BN - 1 > m0);
      ptx::mbar_wait(s_full, j & 1);
      ptx::tc_fence_after();
      int lim = p.Nk - kv0;
      if (p.causal) lim = min(lim, row - kv0 + 1);
      // ---- pass 1: maximum over this half's 64 columns, four independent chains
      float mx4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
      for (int cc = 0; cc < 2; ++cc) {
        const int c = half * 2 + cc;
        uint32_t s[32];
        ptx::tmem_ld_32x32b_x32(tmem_S + lane_base + c * 32, s);
        ptx::tmem_ld_wait();
        if (!need_mask || lim >= (c + 1) * 32) {
#pragma unroll
          for (int i = 0; i < 32; i += 8) {
#pragma unroll
            for (int q = 0; q < 4; ++q)
              mx4[q] = fmaxf(mx4[q], fmaxf(__uint_as_float(s[i + 2 * q]), __uint_as_float(s[i + 2 * q + 1])));
          }
        } else {
#pragma unroll
          for (int i = 0; i < 32; ++i)
            mx4[i & 3] = fmaxf(mx4[i & 3], (c * 32 + i < lim) ? __uint_as_float(s[i]) : -INFINITY);
        }
      }
      float mx = fmaxf(fmaxf(mx4[0], mx4[1]), fmaxf(mx4[2], mx4[3]));
      // ---- exchange the partial maxima of the two halves of this row
      float* slot_x = xch + (j & 1) * (2 * BM);
      slot_x[half * BM + r] = mx;
      ptx::named_bar_sync(1, 256);
      mx = fmaxf(mx, slot_x[(half ^ 1) * BM + r]);
      const float m_tile = mx * p.scale_log2;
      const bool move = m_tile > m_ref + kTau;  // also the first finite maximum (m_ref = -inf)
      const float m_new = move ? m_tile : m_ref;
      const float m_use = (m_new == -INFINITY) ? 0.f : m_new;
      const float alpha = move ? exp2f(m_ref - m_use) : 1.f;  // m_ref = -inf -> 0
      // ---- pass 2: P = exp2(S·scale − m) for this half's columns, partial row sum, bf16 P → swizzled smem
      float psum = 0.f;
#pragma unroll
      for (int cc = 0; cc < 2; ++cc) {
        const int c = half * 2 + cc;
        uint32_t s[32];
        ptx::tmem_ld_32x32b_x32(tmem_S + lane_base + c * 32, s);
        ptx::tmem_ld_wait();
        const bool whole = !need_mask || lim >= (c + 1) * 32;
#pragma unroll
        for (int v = 0; v < 4; ++v) {
          float e[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) e[i] = exp2f(fmaf(__uint_as_float(s[8 * v + i]), p.scale_log2, -m_use));
          if (!whole) {
#pragma unroll
            for (int i = 0; i < 8; ++i) e[i] = (c * 32 + 8 * v + i < lim) ? e[i] : 0.f;
          }
          psum += ((e[0] + e[1]) + (e[2] + e[3])) + ((e[4] + e[5]) + (e[6] + e[7]));
          u32x4 o;
          o.x = pack_bf16x2(e[0], e[1]);
          o.y = pack_bf16x2(e[2], e[3]);
          o.z = pack_bf16x2(e[4], e[5]);
          o.w = pack_bf16x2(e[6], e[7]);
          *reinterpret_cast<u32x4*>(sP + swz_off(r, c * 32 + v * 8, BM)) = o;
        }
      }
      // ---- rescale this half's columns of O in tensor memory when some row of the warp moved its reference
      //      (warp-uniform branch: tcgen05.ld / .st are warp-collective; rows that did not move use alpha = 1)
      if (j > 0 && __any_sync(0xffffffffu, move)) {
        ptx::mbar_wait(o_full, (j - 1) & 1);  // PV_{j-1} has landed; PV_j is not issued before our p_full arrive
        ptx::tc_fence_after();
#pragma unroll
        for (int c = 0; c < OC / 32; ++c) {
          uint32_t t[32];
          ptx::tmem_ld_32x32b_x32(tmem_O + lane_base + half * OC + c * 32, t);
          ptx::tmem_ld_wait();
#pragma unroll
          for (int i = 0; i < 32; ++i) t[i] = __float_as_uint(__uint_as_float(t[i]) * alpha);
          ptx::tmem_st_32x32b_x32(tmem_O + lane_base + half * OC + c * 32, t);
        }
        ptx::tmem_st_wait();
      }
      ptx::tc_fence_before();
      ptx::fence_proxy_async();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(p_full);
      l_part = l_part * alpha + psum;
      m_ref = m_new;
    }
    // ---- the accumulated O row (this half's columns) comes out of tensor memory once
    float o_acc[OC];
    if (n_kv > 0) {
      ptx::mbar_wait(o_full, (n_kv - 1) & 1);
      ptx::tc_fence_after();
#pragma unroll
      for (int c = 0; c < OC / 32; ++c) {
        uint32_t t[32];
        ptx::tmem_ld_32x32b_x32(tmem_O + lane_base + half * OC + c * 32, t);
        ptx::tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; ++i) o_acc[c * 32 + i] = __uint_as_float(t[i]);
      }
    } else {
#pragma unroll
      for (int i = 0; i < OC; ++i) o_acc[i] = 0.f;
    }
    const float m_run = m_ref;

    // ---- epilogue: combine the two partial row sums, O / l, logsumexp
    float* slot_x = xch + (n_kv & 1) * (2 * BM);  // the buffer NOT used by the last tile's max exchange
    slot_x[half * BM + r] = l_part;
    ptx::named_bar_sync(1, 256);
    const float l_run = l_part + slot_x[(half ^ 1) * BM + r];
    if (row < p.Nq) {
      const float inv_l = l_run > 0.f ? 1.f / l_run : 0.f;
      const long long o_off = (long long)b * p.o_sb + (long long)h * p.o_sh + (long long)row * p.o_sr + half * OC;
      if (p.out_f32) {
        // fp32 output (fp32 callers: tests/test_attention.py feeds fp32): no rounding of O to bf16, no cast kernel
        float* orow = reinterpret_cast<float*>(p.o) + o_off;
        if (((reinterpret_cast<uintptr_t>(orow) & 15) == 0) && (p.D % 4 == 0)) {
#pragma unroll
          for (int v = 0; v < OC / 4; ++v)
            if (half * OC + v * 4 < p.D)
              *reinterpret_cast<f32x4*>(orow + 4 * v) = f32x4{o_acc[4 * v] * inv_l, o_acc[4 * v + 1] * inv_l,
                                                               o_acc[4 * v + 2] * inv_l, o_acc[4 * v + 3] * inv_l};
        } else {
#pragma unroll
          for (int i = 0; i < OC; ++i)
            if (half * OC + i < p.D) orow[i] = o_acc[i] * inv_l;
        }
      } else {
        __nv_bfloat16* orow = reinterpret_cast<__nv_bfloat16*>(p.o) + o_off;
        const bool vec = ((reinterpret_cast<uintptr_t>(orow) & 15) == 0) && (p.D % 8 == 0);
        if (vec) {
#pragma unroll
          for (int v = 0; v < OC / 8; ++v) {
            if (half * OC + v * 8 < p.D) {
              u32x4 o;
              o.x = pack_bf16x2(o_acc[8 * v + 0] * inv_l, o_acc[8 * v + 1] * inv_l);
              o.y = pack_bf16x2(o_acc[8 * v + 2] * inv_l, o_acc[8 * v + 3] * inv_l);
              o.z = pack_bf16x2(o_acc[8 * v + 4] * inv_l, o_acc[8 * v + 5] * inv_l);
              o.w = pack_bf16x2(o_acc[8 * v + 6] * inv_l, o_acc[8 * v + 7] * inv_l);
              *reinterpret_cast<u32x4*>(orow + 8 * v) = o;
            }
          }
        } else {
#pragma unroll
          for (int i = 0; i < OC; ++i)
            if (half * OC + i < p.D) orow[i] = __float2bfloat16(o_acc[i] * inv_l);
        }
      }
      if (half == 0)
        p.lse[((long long)b * p.H + h) * p.Nq + row] = l_run > 0.f ? (m_run + log2f(l_run)) * kLn2 : -INFINITY;
    }
  }

  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 9) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, C::TMEM_COLS);
  }
  if constexpr (CLUSTER) ptx::cluster_sync();  // the peer may still multicast into our shared memory / signal our barriers
}

// 4-D tensor map over a (B, H, N, D) bf16 view: dims innermost-first (D, N, H, B), box (64, rows, 1, 1)
int make_bhnd_map(CUtensorMap* out, const void* base, int B, int H, int N, int D, const long long* strides /*b,h,row*/,
                  int box_rows) {
  uint64_t dims[4] = {(uint64_t)D, (uint64_t)N, (uint64_t)H, (uint64_t)B};
  uint64_t str[3] = {(uint64_t)strides[2] * 2, (uint64_t)strides[1] * 2, (uint64_t)strides[0] * 2};
  uint32_t box[4] = {64, (uint32_t)box_rows, 1, 1};
  // a size-1 dimension may carry any stride; TMA still wants 16-byte multiples
  if (H == 1 && str[1] % 16 != 0) str[1] = 16;
  if (B == 1 && str[2] % 16 != 0) str[2] = 16;
  return encode_tensor_map(out, TmDtype::BF16, 4, base, dims, str, box);
}

template <int DP, bool CLUSTER>
static int launch_fwd(const CUtensorMap& q, const CUtensorMap& k, const CUtensorMap& v, const CUtensorMap& k64,
                      const CUtensorMap& v64, const FwdParams& p, int B, cudaStream_t s) {
  using C = Fwd2Cfg<DP>;
  auto kern = fa2_fwd_kernel<DP, CLUSTER>;
  B200_ENSURE_DYN_SMEM(kern, C::SMEM_BYTES);
  dim3 grid(p.num_m_tiles, p.H, B);  // CLUSTER: num_m_tiles is even, clusters of two along x
  B200_CHECK_CUDA(launch_pdl_cluster(kern, grid, dim3(kThreads2), C::SMEM_BYTES, s, CLUSTER ? 2 : 1, q, k, v, k64, v64, p));
  return 0;
}

}  // namespace fa2
}  // namespace b200

// out_f32: O is written as fp32 (same strides, in elements) instead of bf16.
// cluster: 1 = share K/V tiles across 2-CTA clusters when the number of query tiles is even (default), 0 = never.
extern "C" int b200_fa2_fwd_ex(const void* q, const void* k, const void* v, void* o, float* lse, int B, int H, int Nq,
                               int Nk, int D, const long long* q_strides, const long long* k_strides,
                               const long long* v_strides, const long long* o_strides, int causal, float scale,
                               int out_f32, int cluster, void* stream) {
  using namespace b200;
  using namespace b200::fa2;
  B200_REQUIRE(B > 0 && H > 0 && Nq > 0 && Nk > 0, "fa2_fwd: empty problem");
  B200_REQUIRE(D >= 8 && D <= 128 && D % 8 == 0, "fa2_fwd: head dim %d unsupported (multiple of 8, <= 128)", D);
  CUtensorMap tq, tk, tv, tk64, tv64;
  int rc;
  if ((rc = make_bhnd_map(&tq, q, B, H, Nq, D, q_strides, BM))) return rc;
  if ((rc = make_bhnd_map(&tk, k, B, H, Nk, D, k_strides, BN))) return rc;
  if ((rc = make_bhnd_map(&tv, v, B, H, Nk, D, v_strides, BN))) return rc;
  FwdParams p;
  p.o = o;
  p.lse = lse;
  p.o_sb = o_strides[0]; p.o_sh = o_strides[1]; p.o_sr = o_strides[2];
  p.H = H; p.Nq = Nq; p.Nk = Nk; p.D = D;
  p.causal = causal;
  p.out_f32 = out_f32;
  p.scale_log2 = scale * kLog2e;
  p.num_m_tiles = ceil_div(Nq, BM);
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  if (cluster && (p.num_m_tiles % 2) == 0) {
    if ((rc = make_bhnd_map(&tk64, k, B, H, Nk, D, k_strides, 64))) return rc;
    if ((rc = make_bhnd_map(&tv64, v, B, H, Nk, D, v_strides, 64))) return rc;
    return D <= 64 ? launch_fwd<64, true>(tq, tk, tv, tk64, tv64, p, B, s)
                   : launch_fwd<128, true>(tq, tk, tv, tk64, tv64, p, B, s);
  }
  return D <= 64 ? launch_fwd<64, false>(tq, tk, tv, tk, tv, p, B, s) : launch_fwd<128, false>(tq, tk, tv, tk, tv, p, B, s);
}

extern "C" int b200_fa2_fwd(const void* q, const void* k, const void* v, void* o, float* lse, int B, int H, int Nq,
                            int Nk, int D, const long long* q_strides, const long long* k_strides,
                            const long long* v_strides, const long long* o_strides, int causal, float scale,
                            void* stream) {
  return b200_fa2_fwd_ex(q, k, v, o, lse, B, H, Nq, Nk, D, q_strides, k_strides, v_strides, o_strides, causal, scale, 0,
                         1, stream);
}
