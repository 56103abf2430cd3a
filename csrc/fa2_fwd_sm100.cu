// FlashAttention-2 forward on tcgen05 / TMEM / TMA for sm_100a.
//
// Replaces the reference's Triton kernel (cs336_systems/flash_attention.py:137-266: 16×16 tiles,
// mma.sync, no causal tile skipping) with 128×128 tiles on the 5th-gen tensor cores.
//
// One CTA = one 128-row query tile of one (batch, head); 192 threads:
//   warps 0-3  softmax warpgroup: thread t owns query row t.  Reads S from TMEM (one TMEM lane per
//              thread → row max/sum need no shuffles), online softmax in the log2 domain, writes
//              P (bf16) into shared memory in the 128B-swizzled K-major UMMA layout, keeps the
//              running O row in registers (O += PV partial read back from TMEM, rescaled by alpha)
//   warp 4     TMA producer: Q once, then K_j, V_j, K_j+1, … through a ring of shared-memory slots
//   warp 5     TMEM allocator + MMA issuer: S = Q·K_jᵀ (both K-major), PV = P·V_j (V as MN-major B)
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
constexpr int kThreads = 192;
constexpr float kLog2e = 1.4426950408889634f;
constexpr float kLn2 = 0.6931471805599453f;

template <int DP>
struct FwdCfg {
  static constexpr int NBLK = DP / 64;                 // 64-column (128-byte) blocks per row
  static constexpr int TILE_BYTES = BN * DP * 2;       // one K or V tile (also Q)
  static constexpr int P_BYTES = BM * BN * 2;
  static constexpr int SLOTS = DP == 64 ? 3 : 4;
  static constexpr int SMEM_BYTES = TILE_BYTES /*Q*/ + P_BYTES + SLOTS * TILE_BYTES + 1024 + 256;
  static constexpr int TMEM_COLS = 256;                // S: 128 cols, PV partial: DP cols
  static constexpr int MIN_CTAS = DP == 64 ? 2 : 1;
};

struct FwdParams {
  __nv_bfloat16* o;
  float* lse;
  long long o_sb, o_sh, o_sr;
  int H, Nq, Nk, D;
  int causal;
  float scale_log2;  // softmax scale * log2(e)
  int num_m_tiles;
  long long* trace;  // debug: when non-null, thread 0 of CTA (0,0,0) stores clock64() stamps of its softmax loop
};

// byte offset of element (row, col) in a K-major [rows × 128] bf16 tile made of two 64-column blocks,
// each block = rows × 128 B with the 128B swizzle (16-byte chunk index XOR row%8)
__device__ __forceinline__ uint32_t swz_off(int row, int col, int rows) {
  const int blk = col >> 6, c = col & 63;
  return blk * (rows * 128) + row * 128 + ((((c >> 3) ^ (row & 7)) << 4) | ((c & 7) << 1));
}

template <int DP>
__global__ void __launch_bounds__(kThreads, FwdCfg<DP>::MIN_CTAS)
fa2_fwd_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
               const __grid_constant__ CUtensorMap tmV, const FwdParams p) {
  using C = FwdCfg<DP>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sQ = smem;
  uint8_t* sP = sQ + C::TILE_BYTES;
  uint8_t* sKV = sP + C::P_BYTES;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sKV + C::SLOTS * C::TILE_BYTES);
  uint64_t* q_full = bars;
  uint64_t* kv_full = bars + 1;
  uint64_t* kv_empty = kv_full + C::SLOTS;
  uint64_t* s_full = kv_empty + C::SLOTS;
  uint64_t* p_full = s_full + 1;
  uint64_t* o_full = p_full + 1;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(o_full + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // heavy (late) causal tiles first
  const int m_tile = p.causal ? (p.num_m_tiles - 1 - (int)blockIdx.x) : (int)blockIdx.x;
  const int h = blockIdx.y, b = blockIdx.z;
  const int m0 = m_tile * BM;

  int n_kv = ceil_div(p.Nk, BN);
  if (p.causal) {
    const int lim = ceil_div(min(m0 + BM, p.Nk), BN);
    n_kv = min(n_kv, lim);
  }

  if (warp == 4 && ptx::elect_one()) {
    ptx::tma_prefetch_desc(&tmQ);
    ptx::tma_prefetch_desc(&tmK);
    ptx::tma_prefetch_desc(&tmV);
    ptx::mbar_init(q_full, 1);
    for (int s = 0; s < C::SLOTS; ++s) {
      ptx::mbar_init(&kv_full[s], 1);
      ptx::mbar_init(&kv_empty[s], 1);
    }
    ptx::mbar_init(s_full, 1);
    ptx::mbar_init(p_full, 4);  // one arrive per softmax warp
    ptx::mbar_init(o_full, 1);
    ptx::fence_mbar_init();
  }
  if (warp == 5) {
    ptx::tmem_alloc(tmem_slot, C::TMEM_COLS);
    ptx::tmem_relinquish();
  }
  ptx::griddep_launch();
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  ptx::griddep_wait();
  const uint32_t tmem_S = tmem_base;
  const uint32_t tmem_O = tmem_base + BN;

  if (warp == 4) {
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
          ptx::mbar_expect_tx(&kv_full[slot], C::TILE_BYTES);
          uint8_t* dst = sKV + slot * C::TILE_BYTES;
#pragma unroll
          for (int blk = 0; blk < C::NBLK; ++blk)
            ptx::tma_load_4d(dst + blk * (BN * 128), kv == 0 ? &tmK : &tmV, &kv_full[slot], blk * 64, j * BN, h, b,
                             ptx::kEvictLast);
          if (++slot == C::SLOTS) { slot = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 5) {
    // ===================================== MMA issuer =====================================
    if (ptx::elect_one() && n_kv > 0) {
      constexpr uint32_t idesc_qk = ptx::make_idesc(BM, BN, false, false);
      constexpr uint32_t idesc_pv = ptx::make_idesc(BM, DP, false, true);
      const uint32_t q_addr = ptx::smem_u32(sQ), p_addr = ptx::smem_u32(sP), kv_addr = ptx::smem_u32(sKV);
      int slot = 0;
      uint32_t phase = 0;
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
        ptx::mma_commit(&kv_empty[slot]);
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
          const uint32_t a_off = (kk >> 2) * (BM * 128) + (kk & 3) * 32;  // P: K-major, K = keys
          // V: MN-major B (N = head dim contiguous): 16 key rows = 2048 B; 64-wide D blocks BN*128 B apart
          ptx::mma_f16_ss(tmem_O, ptx::make_smem_desc(p_addr + a_off, 16, 1024),
                          ptx::make_smem_desc(v_addr + kk * 2048, BN * 128, 1024), idesc_pv, kk != 0);
        }
        ptx::mma_commit(&kv_empty[slot]);
        ptx::mma_commit(o_full);
        if (++slot == C::SLOTS) { slot = 0; phase ^= 1; }
        if (j + 1 < n_kv) issue_qk();
      }
    }
  } else {
    // ===================================== softmax warpgroup =====================================
    const int r = warp * 32 + lane;  // row within the tile == TMEM lane
    const int row = m0 + r;
    const uint32_t lane_base = static_cast<uint32_t>(warp * 32) << 16;
    float o_acc[DP];
#pragma unroll
    for (int i = 0; i < DP; ++i) o_acc[i] = 0.f;
    float m_run = -INFINITY, l_run = 0.f;
    const bool tr = p.trace != nullptr && threadIdx.x == 0 && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0;
    int tk = 0;
#define FA_STAMP() do { if (tr && tk < 250) p.trace[tk++] = clock64(); } while (0)
    FA_STAMP();

    for (int j = 0; j < n_kv; ++j) {
      const int kv0 = j * BN;
      const bool need_mask = (kv0 + BN > p.Nk) || (p.causal && kv0 + BN - 1 > m0);
      ptx::mbar_wait(s_full, j & 1);
      ptx::tc_fence_after();
      FA_STAMP();  // S ready
      // Per-row count of valid keys in this tile: keys [kv0, kv0 + lim) are visible to this query row
      // (sequence end and, if causal, the diagonal).  lim >= BN: nothing to mask.
      int lim = p.Nk - kv0;
      if (p.causal) lim = min(lim, row - kv0 + 1);
      // ---- pass 1: row max
      float mx = -INFINITY;
#pragma unroll
      for (int c = 0; c < BN / 32; ++c) {
        uint32_t s[32];
        ptx::tmem_ld_32x32b_x32(tmem_S + lane_base + c * 32, s);
        ptx::tmem_ld_wait();
        if (!need_mask || lim >= (c + 1) * 32) {  // whole chunk visible (the common case)
#pragma unroll
          for (int i = 0; i < 32; i += 2) mx = fmaxf(mx, fmaxf(__uint_as_float(s[i]), __uint_as_float(s[i + 1])));
        } else {
#pragma unroll
          for (int i = 0; i < 32; ++i) mx = fmaxf(mx, (c * 32 + i < lim) ? __uint_as_float(s[i]) : -INFINITY);
        }
      }
      FA_STAMP();  // pass 1 done
      const float m_new = fmaxf(m_run, mx * p.scale_log2);
      const float m_use = (m_new == -INFINITY) ? 0.f : m_new;  // fully-masked row so far
      const float alpha = exp2f(m_run - m_use);                // m_run = -inf → 0
      // ---- pass 2: P = exp2(S·scale − m), row sum, bf16 P → swizzled smem (8 elements = one 16-byte store)
      float psum = 0.f;
#pragma unroll
      for (int c = 0; c < BN / 32; ++c) {
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
      ptx::tc_fence_before();      // our TMEM reads of S are done before the MMA warp overwrites S
      ptx::fence_proxy_async();    // P stores visible to the tensor core (async proxy)
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(p_full);
      FA_STAMP();  // P published
      // ---- rescale the running row while the PV MMA runs
      l_run = l_run * alpha + psum;
      m_run = m_new;
#pragma unroll
      for (int i = 0; i < DP; ++i) o_acc[i] *= alpha;
      // ---- accumulate the PV partial
      ptx::mbar_wait(o_full, j & 1);
      ptx::tc_fence_after();
      FA_STAMP();  // PV ready
#pragma unroll
      for (int c = 0; c < DP / 32; ++c) {
        uint32_t t[32];
        ptx::tmem_ld_32x32b_x32(tmem_O + lane_base + c * 32, t);
        ptx::tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; ++i) o_acc[c * 32 + i] += __uint_as_float(t[i]);
      }
      // (the next p_full arrive orders these reads before the next PV overwrite)
      FA_STAMP();  // O accumulated
    }

    // ---- epilogue: O / l, logsumexp
    if (row < p.Nq) {
      const float inv_l = l_run > 0.f ? 1.f / l_run : 0.f;
      __nv_bfloat16* orow = p.o + (long long)b * p.o_sb + (long long)h * p.o_sh + (long long)row * p.o_sr;
      const bool vec = ((reinterpret_cast<uintptr_t>(orow) & 15) == 0) && (p.D % 8 == 0);
      if (vec) {
#pragma unroll
        for (int v = 0; v < DP / 8; ++v) {
          if (v * 8 < p.D) {
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
        for (int i = 0; i < DP; ++i)
          if (i < p.D) orow[i] = __float2bfloat16(o_acc[i] * inv_l);
      }
      p.lse[((long long)b * p.H + h) * p.Nq + row] = l_run > 0.f ? (m_run + log2f(l_run)) * kLn2 : -INFINITY;
    }
  }

  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 5) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, C::TMEM_COLS);
  }
}

// ------------------------------------------------------------------------------------------------
// Variant with TWO softmax warpgroups per CTA (opt-in: B200SYS_FA_FWD_WG2=1).
//
// The softmax of the kernel above is latency-bound: one thread walks a whole 128-column row (two TMEM passes,
// dependent max / sum chains) with two warps per scheduler.  Here warps 0-3 take columns 0-63 and warps 4-7 columns
// 64-127 of the same 128 rows (a warp may read TMEM lanes 32·(warp%4)…+31 at any column), so the serial work per
// thread halves and four softmax warps per scheduler hide each other's TMEM round trips.  The two halves of a row
// exchange their partial row maximum through shared memory once per tile (named barrier, double-buffered slots);
// the partial row sums are only combined in the epilogue because both halves use the same running maximum.  Each
// half accumulates its own DP/2 columns of O.
// ------------------------------------------------------------------------------------------------
constexpr int kThreads2 = 320;  // 8 softmax warps, TMA warp (8), MMA warp (9)

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

template <int DP>
__global__ void __launch_bounds__(kThreads2, Fwd2Cfg<DP>::MIN_CTAS)
fa2_fwd_wg2_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                   const __grid_constant__ CUtensorMap tmV, const FwdParams p) {
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

  if (warp == 8 && ptx::elect_one()) {
    ptx::tma_prefetch_desc(&tmQ);
    ptx::tma_prefetch_desc(&tmK);
    ptx::tma_prefetch_desc(&tmV);
    ptx::mbar_init(q_full, 1);
    for (int s = 0; s < C::SLOTS; ++s) {
      ptx::mbar_init(&kv_full[s], 1);
      ptx::mbar_init(&kv_empty[s], 1);
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
          ptx::mbar_expect_tx(&kv_full[slot], C::TILE_BYTES);
          uint8_t* dst = sKV + slot * C::TILE_BYTES;
#pragma unroll
          for (int blk = 0; blk < C::NBLK; ++blk)
            ptx::tma_load_4d(dst + blk * (BN * 128), kv == 0 ? &tmK : &tmV, &kv_full[slot], blk * 64, j * BN, h, b,
                             ptx::kEvictLast);
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
        ptx::mma_commit(&kv_empty[slot]);
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
                          ptx::make_smem_desc(v_addr + kk * 2048, BN * 128, 1024), idesc_pv, kk != 0);
        }
        ptx::mma_commit(&kv_empty[slot]);
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
    float o_acc[OC];
#pragma unroll
    for (int i = 0; i < OC; ++i) o_acc[i] = 0.f;
    float m_run = -INFINITY, l_part = 0.f;

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
      const float m_new = fmaxf(m_run, mx * p.scale_log2);
      const float m_use = (m_new == -INFINITY) ? 0.f : m_new;
      const float alpha = exp2f(m_run - m_use);
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
      ptx::tc_fence_before();
      ptx::fence_proxy_async();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(p_full);
      l_part = l_part * alpha + psum;
      m_run = m_new;
#pragma unroll
      for (int i = 0; i < OC; ++i) o_acc[i] *= alpha;
      ptx::mbar_wait(o_full, j & 1);
      ptx::tc_fence_after();
#pragma unroll
      for (int c = 0; c < OC / 32; ++c) {
        uint32_t t[32];
        ptx::tmem_ld_32x32b_x32(tmem_O + lane_base + half * OC + c * 32, t);
        ptx::tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; ++i) o_acc[c * 32 + i] += __uint_as_float(t[i]);
      }
    }

    // ---- epilogue: combine the two partial row sums, O / l, logsumexp
    float* slot_x = xch + (n_kv & 1) * (2 * BM);  // the buffer NOT used by the last tile's max exchange
    slot_x[half * BM + r] = l_part;
    ptx::named_bar_sync(1, 256);
    const float l_run = l_part + slot_x[(half ^ 1) * BM + r];
    if (row < p.Nq) {
      const float inv_l = l_run > 0.f ? 1.f / l_run : 0.f;
      __nv_bfloat16* orow = p.o + (long long)b * p.o_sb + (long long)h * p.o_sh + (long long)row * p.o_sr + half * OC;
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
}

// Two-warpgroup kernel with K/V tiles shared across a 2-CTA thread-block cluster by TMA multicast
// (opt-in: B200SYS_FA_FWD_WG2=3; needs an even number of query tiles, else the plain two-warpgroup kernel runs).
template <int DP>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads2, Fwd2Cfg<DP>::MIN_CTAS)
fa2_fwd_wg2c_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
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
  // The two CTAs of a cluster work on neighbouring query tiles of the same (batch, head).  Key tiles both need
  // (j < n_shared) are fetched cooperatively: each CTA loads one 64-row half and multicasts it into both CTAs'
  // shared memory, halving the L2 -> SM traffic for K and V; the extra causal tile of the later query tile is
  // loaded by its CTA alone.
  const uint32_t cr = ptx::cluster_ctarank();
  const int peer_bx = (int)blockIdx.x ^ 1;
  const int peer_m0 = (p.causal ? (p.num_m_tiles - 1 - peer_bx) : peer_bx) * BM;
  int n_kv_peer = ceil_div(p.Nk, BN);
  if (p.causal) n_kv_peer = min(n_kv_peer, ceil_div(min(peer_m0 + BM, p.Nk), BN));
  const int n_shared = min(n_kv, n_kv_peer);

  if (warp == 8 && ptx::elect_one()) {
    ptx::tma_prefetch_desc(&tmQ);
    ptx::tma_prefetch_desc(&tmK);
    ptx::tma_prefetch_desc(&tmV);
    ptx::tma_prefetch_desc(&tmK64);
    ptx::tma_prefetch_desc(&tmV64);
    ptx::mbar_init(q_full, 1);
    for (int s = 0; s < C::SLOTS; ++s) {
      ptx::mbar_init(&kv_full[s], 1);
      ptx::mbar_init(&kv_empty[s], 2);  // a shared slot is free when BOTH CTAs' MMAs have read it
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
  ptx::cluster_sync();  // the peer's barriers are initialised before anything of ours can signal them
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
          ptx::mbar_expect_tx(&kv_full[slot], C::TILE_BYTES);  // our half + the peer's half (or the whole solo tile)
          uint8_t* dst = sKV + slot * C::TILE_BYTES;
          if (j < n_shared) {
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
        if (tile < n_shared) ptx::mma_commit_mcast(&kv_empty[slot], (uint16_t)0x3);
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
                          ptx::make_smem_desc(v_addr + kk * 2048, BN * 128, 1024), idesc_pv, kk != 0);
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
    float o_acc[OC];
#pragma unroll
    for (int i = 0; i < OC; ++i) o_acc[i] = 0.f;
    float m_run = -INFINITY, l_part = 0.f;

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
      const float m_new = fmaxf(m_run, mx * p.scale_log2);
      const float m_use = (m_new == -INFINITY) ? 0.f : m_new;
      const float alpha = exp2f(m_run - m_use);
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
      ptx::tc_fence_before();
      ptx::fence_proxy_async();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(p_full);
      l_part = l_part * alpha + psum;
      m_run = m_new;
#pragma unroll
      for (int i = 0; i < OC; ++i) o_acc[i] *= alpha;
      ptx::mbar_wait(o_full, j & 1);
      ptx::tc_fence_after();
#pragma unroll
      for (int c = 0; c < OC / 32; ++c) {
        uint32_t t[32];
        ptx::tmem_ld_32x32b_x32(tmem_O + lane_base + half * OC + c * 32, t);
        ptx::tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 32; ++i) o_acc[c * 32 + i] += __uint_as_float(t[i]);
      }
    }

    // ---- epilogue: combine the two partial row sums, O / l, logsumexp
    float* slot_x = xch + (n_kv & 1) * (2 * BM);  // the buffer NOT used by the last tile's max exchange
    slot_x[half * BM + r] = l_part;
    ptx::named_bar_sync(1, 256);
    const float l_run = l_part + slot_x[(half ^ 1) * BM + r];
    if (row < p.Nq) {
      const float inv_l = l_run > 0.f ? 1.f / l_run : 0.f;
      __nv_bfloat16* orow = p.o + (long long)b * p.o_sb + (long long)h * p.o_sh + (long long)row * p.o_sr + half * OC;
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
  ptx::cluster_sync();  // the peer may still multicast into our shared memory / signal our barriers
}

// Same two-warpgroup structure, with O resident in tensor memory and lazy rescaling (opt-in: B200SYS_FA_FWD_WG2=2).
template <int DP>
__global__ void __launch_bounds__(kThreads2, Fwd2Cfg<DP>::MIN_CTAS)
fa2_fwd_wg2o_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                   const __grid_constant__ CUtensorMap tmV, const FwdParams p) {
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

  if (warp == 8 && ptx::elect_one()) {
    ptx::tma_prefetch_desc(&tmQ);
    ptx::tma_prefetch_desc(&tmK);
    ptx::tma_prefetch_desc(&tmV);
    ptx::mbar_init(q_full, 1);
    for (int s = 0; s < C::SLOTS; ++s) {
      ptx::mbar_init(&kv_full[s], 1);
      ptx::mbar_init(&kv_empty[s], 1);
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
          ptx::mbar_expect_tx(&kv_full[slot], C::TILE_BYTES);
          uint8_t* dst = sKV + slot * C::TILE_BYTES;
#pragma unroll
          for (int blk = 0; blk < C::NBLK; ++blk)
            ptx::tma_load_4d(dst + blk * (BN * 128), kv == 0 ? &tmK : &tmV, &kv_full[slot], blk * 64, j * BN, h, b,
                             ptx::kEvictLast);
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
        ptx::mma_commit(&kv_empty[slot]);
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
        ptx::mma_commit(&kv_empty[slot]);
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
      __nv_bfloat16* orow = p.o + (long long)b * p.o_sb + (long long)h * p.o_sh + (long long)row * p.o_sr + half * OC;
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

template <int DP>
static int launch_fwd(const CUtensorMap& q, const CUtensorMap& k, const CUtensorMap& v, const FwdParams& p, int B,
                      cudaStream_t s) {
  using C = FwdCfg<DP>;
  auto kern = fa2_fwd_kernel<DP>;
  B200_ENSURE_DYN_SMEM(kern, C::SMEM_BYTES);
  dim3 grid(p.num_m_tiles, p.H, B);
  B200_CHECK_CUDA(launch_pdl(kern, grid, dim3(kThreads), C::SMEM_BYTES, s, q, k, v, p));
  return 0;
}

template <int DP>
static int launch_fwd_wg2(const CUtensorMap& q, const CUtensorMap& k, const CUtensorMap& v, const FwdParams& p, int B,
                          cudaStream_t s) {
  using C = Fwd2Cfg<DP>;
  auto kern = fa2_fwd_wg2_kernel<DP>;
  B200_ENSURE_DYN_SMEM(kern, C::SMEM_BYTES);
  dim3 grid(p.num_m_tiles, p.H, B);
  B200_CHECK_CUDA(launch_pdl(kern, grid, dim3(kThreads2), C::SMEM_BYTES, s, q, k, v, p));
  return 0;
}

template <int DP>
static int launch_fwd_wg2o(const CUtensorMap& q, const CUtensorMap& k, const CUtensorMap& v, const FwdParams& p, int B,
                           cudaStream_t s) {
  using C = Fwd2Cfg<DP>;
  auto kern = fa2_fwd_wg2o_kernel<DP>;
  B200_ENSURE_DYN_SMEM(kern, C::SMEM_BYTES);
  dim3 grid(p.num_m_tiles, p.H, B);
  B200_CHECK_CUDA(launch_pdl(kern, grid, dim3(kThreads2), C::SMEM_BYTES, s, q, k, v, p));
  return 0;
}

template <int DP>
static int launch_fwd_wg2c(const CUtensorMap& q, const CUtensorMap& k, const CUtensorMap& v, const CUtensorMap& k64,
                           const CUtensorMap& v64, const FwdParams& p, int B, cudaStream_t s) {
  using C = Fwd2Cfg<DP>;
  auto kern = fa2_fwd_wg2c_kernel<DP>;
  B200_ENSURE_DYN_SMEM(kern, C::SMEM_BYTES);
  dim3 grid(p.num_m_tiles, p.H, B);  // num_m_tiles is even: clusters of two along x
  B200_CHECK_CUDA(launch_pdl(kern, grid, dim3(kThreads2), C::SMEM_BYTES, s, q, k, v, k64, v64, p));
  return 0;
}

// B200SYS_FA_FWD_WG2: 0 default kernel, 1 two softmax warpgroups, 2 the same with O resident in tensor memory,
// 3 two softmax warpgroups + K/V multicast across a 2-CTA cluster
static int wg2_mode() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("B200SYS_FA_FWD_WG2");
    v = (e && (e[0] == '1' || e[0] == '2' || e[0] == '3')) ? (e[0] - '0') : 0;
  }
  return v;
}

}  // namespace fa2
}  // namespace b200

static long long* g_fa2_fwd_trace = nullptr;
// debug hook: device buffer of >= 256 int64 that receives clock64() stamps of one softmax thread (nullptr = off)
extern "C" void b200_fa2_fwd_set_trace(long long* dev_buf) { g_fa2_fwd_trace = dev_buf; }

extern "C" int b200_fa2_fwd(const void* q, const void* k, const void* v, void* o, float* lse, int B, int H, int Nq,
                            int Nk, int D, const long long* q_strides, const long long* k_strides,
                            const long long* v_strides, const long long* o_strides, int causal, float scale,
                            void* stream) {
  using namespace b200;
  using namespace b200::fa2;
  B200_REQUIRE(B > 0 && H > 0 && Nq > 0 && Nk > 0, "fa2_fwd: empty problem");
  B200_REQUIRE(D >= 8 && D <= 128 && D % 8 == 0, "fa2_fwd: head dim %d unsupported (multiple of 8, <= 128)", D);
  CUtensorMap tq, tk, tv;
  int rc;
  if ((rc = make_bhnd_map(&tq, q, B, H, Nq, D, q_strides, BM))) return rc;
  if ((rc = make_bhnd_map(&tk, k, B, H, Nk, D, k_strides, BN))) return rc;
  if ((rc = make_bhnd_map(&tv, v, B, H, Nk, D, v_strides, BN))) return rc;
  FwdParams p;
  p.o = reinterpret_cast<__nv_bfloat16*>(o);
  p.lse = lse;
  p.o_sb = o_strides[0]; p.o_sh = o_strides[1]; p.o_sr = o_strides[2];
  p.H = H; p.Nq = Nq; p.Nk = Nk; p.D = D;
  p.causal = causal;
  p.scale_log2 = scale * kLog2e;
  p.num_m_tiles = ceil_div(Nq, BM);
  p.trace = g_fa2_fwd_trace;
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  if (wg2_mode() == 3 && (p.num_m_tiles % 2) == 0) {  // ... and K/V multicast across a CTA pair
    CUtensorMap tk64, tv64;
    if ((rc = make_bhnd_map(&tk64, k, B, H, Nk, D, k_strides, 64))) return rc;
    if ((rc = make_bhnd_map(&tv64, v, B, H, Nk, D, v_strides, 64))) return rc;
    return D <= 64 ? launch_fwd_wg2c<64>(tq, tk, tv, tk64, tv64, p, B, s) : launch_fwd_wg2c<128>(tq, tk, tv, tk64, tv64, p, B, s);
  }
  if (wg2_mode() == 1 || wg2_mode() == 3)  // opt-in variant with two softmax warpgroups per CTA
    return D <= 64 ? launch_fwd_wg2<64>(tq, tk, tv, p, B, s) : launch_fwd_wg2<128>(tq, tk, tv, p, B, s);
  if (wg2_mode() == 2)  // ... and O kept in tensor memory with lazy rescaling
    return D <= 64 ? launch_fwd_wg2o<64>(tq, tk, tv, p, B, s) : launch_fwd_wg2o<128>(tq, tk, tv, p, B, s);
  return D <= 64 ? launch_fwd<64>(tq, tk, tv, p, B, s) : launch_fwd<128>(tq, tk, tv, p, B, s);
}
