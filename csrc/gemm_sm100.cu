// Persistent warp-specialised bf16 GEMM on tcgen05 / TMEM / TMA for sm_100a.
//
//   C[M,N] = A · Bᵀ   with fp32 accumulation in tensor memory,
//   A given K-major ([M,K] row-major) or MN-major ([K,M] row-major),
//   B given K-major ([N,K] row-major) or MN-major ([K,N] row-major),
// which covers the three GEMMs of a linear layer without any transposed copy:
//   forward  y  = x·Wᵀ      A = x  (K-major)   B = W  (K-major)
//   dgrad    dx = dy·W      A = dy (K-major)   B = W  (MN-major)
//   wgrad    dW = dyᵀ·x     A = dy (MN-major)  B = x  (MN-major)
// Replaces the cuBLAS calls behind every `Linear` of the reference LM
// (cs336-basics/cs336_basics/model.py:41).
//
// Structure (one CTA per SM, 256 threads):
//   warp 0   TMA producer    — cp.async.bulk.tensor loads of 128×64 A and BN×64 B tiles (128B swizzle)
//                              into a STAGES-deep shared-memory ring, mbarrier complete_tx
//   warp 1   MMA issuer      — one elected thread issues tcgen05.mma (M=128, N=BN, K=16) ×4 per stage,
//                              tcgen05.commit frees the stage / publishes the accumulator
//   warp 2   TMEM allocator  — 512 columns = two BN-wide fp32 accumulators (double buffered)
//   warps 4-7 epilogue       — tcgen05.ld 32 lanes × 32 columns, convert, vectorised global stores;
//                              overlaps the next tile's main loop through the second accumulator
#include "common.cuh"
#include "ptx.cuh"
#include "tensormap.h"
#include "gemm_epilogue.cuh"

namespace b200 {
namespace gemm {

constexpr int kThreads = 256;
constexpr int kTmemCols = 512;

template <int BN>
struct Cfg {
  static constexpr int A_BYTES = BM * BK * 2;
  static constexpr int B_BYTES = BN * BK * 2;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int MAX_STAGES = (227 * 1024 - 2048) / STAGE_BYTES;
  static constexpr int STAGES = MAX_STAGES > 8 ? 8 : MAX_STAGES;
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 /*align slack*/ + 512 /*barriers*/;
};

template <int BN, bool A_MN, bool B_MN, int EPI>
__global__ void __launch_bounds__(128 + 32 * EpiWarps<EPI>::N, 1)
gemm_bf16_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const Params p) {
  using C = Cfg<BN>;
  constexpr int STAGES = C::STAGES;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + STAGES * C::STAGE_BYTES);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tfull_bar = empty_bar + STAGES;
  uint64_t* tempty_bar = tfull_bar + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  if (warp == 0 && ptx::elect_one()) {
    ptx::tma_prefetch_desc(&tmA);
    ptx::tma_prefetch_desc(&tmB);
  }
  if (warp == 1 && ptx::elect_one()) {
    for (int s = 0; s < STAGES; ++s) {
      ptx::mbar_init(&full_bar[s], 1);
      ptx::mbar_init(&empty_bar[s], 1);
    }
    for (int a = 0; a < 2; ++a) {
      ptx::mbar_init(&tfull_bar[a], 1);
      ptx::mbar_init(&tempty_bar[a], EpiWarps<EPI>::N);  // one arrive per epilogue warp
    }
    ptx::fence_mbar_init();
  }
  if (warp == 2) {
    ptx::tmem_alloc(tmem_slot, kTmemCols);
    ptx::tmem_relinquish();
  }
  ptx::griddep_launch();
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  ptx::griddep_wait();  // operands come from the previous kernel; everything above overlapped its tail

  const int total_tiles = p.num_m * p.num_n;

  if (warp == 0) {
    // ================================ TMA producer ================================
    if (ptx::elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      for (int t = blockIdx.x; t < total_tiles; t += gridDim.x) {
        const int m0 = (t % p.num_m) * BM;
        const int n0 = (t / p.num_m) * BN;
        for (int kb = 0; kb < p.num_k; ++kb) {
          ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
          ptx::mbar_expect_tx(&full_bar[stage], C::STAGE_BYTES);
          uint8_t* sa = smem + stage * C::STAGE_BYTES;
          uint8_t* sb = sa + C::A_BYTES;
          if constexpr (!A_MN) {
            ptx::tma_load_2d(sa, &tmA, &full_bar[stage], kb * BK, m0);
          } else {
#pragma unroll
            for (int c = 0; c < BM / 64; ++c)
              ptx::tma_load_2d(sa + c * (64 * BK * 2), &tmA, &full_bar[stage], m0 + c * 64, kb * BK);
          }
          if constexpr (!B_MN) {
            ptx::tma_load_2d(sb, &tmB, &full_bar[stage], kb * BK, n0);
          } else {
#pragma unroll
            for (int c = 0; c < BN / 64; ++c)
              ptx::tma_load_2d(sb + c * (64 * BK * 2), &tmB, &full_bar[stage], n0 + c * 64, kb * BK);
          }
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ================================ MMA issuer ================================
    if (ptx::elect_one()) {
      constexpr uint32_t idesc = ptx::make_idesc(BM, BN, A_MN, B_MN);
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int t = blockIdx.x; t < total_tiles; t += gridDim.x) {
        ptx::mbar_wait(&tempty_bar[acc], acc_phase ^ 1);
        ptx::tc_fence_after();
        const uint32_t d_tmem = tmem_base + acc * BN;
        for (int kb = 0; kb < p.num_k; ++kb) {
          ptx::mbar_wait(&full_bar[stage], phase);
          ptx::tc_fence_after();
          const uint32_t a_addr = ptx::smem_u32(smem + stage * C::STAGE_BYTES);
          const uint32_t b_addr = a_addr + C::A_BYTES;
#pragma unroll
          for (int k = 0; k < BK / 16; ++k) {
            // K-major: advance 16 elements = 32 B inside the swizzled row; MN-major: 16 K-rows = 2048 B
            const uint64_t adesc = A_MN ? ptx::make_smem_desc(a_addr + k * 2048, 64 * BK * 2, 1024)
                                        : ptx::make_smem_desc(a_addr + k * 32, 16, 1024);
            const uint64_t bdesc = B_MN ? ptx::make_smem_desc(b_addr + k * 2048, 64 * BK * 2, 1024)
                                        : ptx::make_smem_desc(b_addr + k * 32, 16, 1024);
            ptx::mma_f16_ss(d_tmem, adesc, bdesc, idesc, (kb | k) != 0 ? 1u : 0u);
          }
          ptx::mma_commit(&empty_bar[stage]);  // stage reusable once these MMAs have read it
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
        ptx::mma_commit(&tfull_bar[acc]);  // accumulator complete
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
      }
    }
  } else if (warp >= 4) {
    // ================================ epilogue ================================
    const int ew = (warp - 4) & 3;  // == warp % 4: TMEM lane quadrant this warp may access
    constexpr int NGRP = EpiWarps<EPI>::N / 4;  // column groups: group g takes the 32-column chunks c = g, g + NGRP, …
    const int grp = (warp - 4) >> 2;
    int acc = 0;
    uint32_t acc_phase = 0;
    AdamwConsts ak;
    if constexpr (EPI == EPI_ADAMW) ak = adamw_consts(p.opt);
    for (int t = blockIdx.x; t < total_tiles; t += gridDim.x) {
      const int m0 = (t % p.num_m) * BM;
      const int n0 = (t / p.num_m) * BN;
      ptx::mbar_wait(&tfull_bar[acc], acc_phase);
      ptx::tc_fence_after();
      const int row = m0 + ew * 32 + lane;
      const uint32_t t_row = tmem_base + (static_cast<uint32_t>(ew * 32) << 16) + acc * BN;
#pragma unroll 1
      for (int c = grp; c < BN / 32; c += NGRP) {
        if (n0 + c * 32 >= p.N) break;  // warp-uniform
        uint32_t r[32];
        ptx::tmem_ld_32x32b_x32(t_row + c * 32, r);
        ptx::tmem_ld_wait();
        if constexpr (EPI == EPI_ADAMW) adamw_chunk_t(p, ak, row, n0 + c * 32, r);
        else if constexpr (EPI == EPI_SWIGLU_BWD) swiglu_bwd_chunk(p, row, n0 + c * 32, r);
        else store_chunk<EPI>(p, row, n0 + c * 32, r);
      }
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(&tempty_bar[acc]);
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
  }

  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, kTmemCols);
  }
}

// ------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------
static int pick_bn(int M, int N, bool need64) {
  // minimise (waves × tile width); wider tiles get a small bonus (less smem traffic per MMA)
  const int cands[3] = {256, 192, 128};
  const double penalty[3] = {1.0, 1.05, 1.38};  // measured in situ: ~1170 / ~1100 / ~845 TFLOP/s per tile width
  int best = 256;
  double best_cost = 1e30;
  const int nm = ceil_div(M, BM);
  for (int i = 0; i < 3; ++i) {
    const int bn = cands[i];
    (void)need64;
    const int tiles = nm * ceil_div(N, bn);
    const int waves = ceil_div(tiles, num_sms());
    const double cost = (double)waves * bn * penalty[i];
    if (cost < best_cost) { best_cost = cost; best = bn; }
  }
  return best;
}

template <int BN, bool A_MN, bool B_MN, int EPI>
static int launch(const CUtensorMap& tmA, const CUtensorMap& tmB, const Params& p, cudaStream_t stream) {
  using C = Cfg<BN>;
  auto kern = gemm_bf16_kernel<BN, A_MN, B_MN, EPI>;
  B200_ENSURE_DYN_SMEM(kern, C::SMEM_BYTES);
  const int tiles = p.num_m * p.num_n;
  const int grid = tiles < num_sms() ? tiles : num_sms();
  B200_CHECK_CUDA(launch_pdl(kern, dim3(grid), dim3(128 + 32 * EpiWarps<EPI>::N), C::SMEM_BYTES, stream, tmA, tmB, p));
  return 0;
}

template <int BN, bool A_MN, bool B_MN>
static int dispatch_epi(int epi, const CUtensorMap& a, const CUtensorMap& b, const Params& p, cudaStream_t s) {
  switch (epi) {
    case EPI_BF16: return launch<BN, A_MN, B_MN, EPI_BF16>(a, b, p, s);
    case EPI_F32: return launch<BN, A_MN, B_MN, EPI_F32>(a, b, p, s);
    case EPI_F32_ACC: return launch<BN, A_MN, B_MN, EPI_F32_ACC>(a, b, p, s);
    case EPI_F32_T:
      if constexpr (A_MN && B_MN) return launch<BN, A_MN, B_MN, EPI_F32_T>(a, b, p, s);
      break;
    case EPI_BF16_SWIGLU:
      if constexpr (!A_MN && !B_MN) return launch<BN, A_MN, B_MN, EPI_BF16_SWIGLU>(a, b, p, s);
      break;
    case EPI_F32_RESID:
      if constexpr (!A_MN && !B_MN) return launch<BN, A_MN, B_MN, EPI_F32_RESID>(a, b, p, s);
      break;
    case EPI_BF16_ROPE:
      if constexpr (!A_MN && !B_MN) return launch<BN, A_MN, B_MN, EPI_BF16_ROPE>(a, b, p, s);
      break;
    case EPI_ADAMW:
      if constexpr (A_MN && B_MN) return launch<BN, A_MN, B_MN, EPI_ADAMW>(a, b, p, s);
      break;
    case EPI_SWIGLU_BWD:
      if constexpr (!A_MN && B_MN) return launch<BN, A_MN, B_MN, EPI_SWIGLU_BWD>(a, b, p, s);
      break;
  }
  set_error("gemm: unsupported epilogue %d for this operand layout", epi);
  return B200_ERR_ARG;
}

template <int BN>
static int dispatch_major(bool a_mn, bool b_mn, int epi, const CUtensorMap& a, const CUtensorMap& b, const Params& p,
                          cudaStream_t s) {
  if (!a_mn && !b_mn) return dispatch_epi<BN, false, false>(epi, a, b, p, s);
  if (!a_mn && b_mn) return dispatch_epi<BN, false, true>(epi, a, b, p, s);
  if (a_mn && b_mn) return dispatch_epi<BN, true, true>(epi, a, b, p, s);
  return dispatch_epi<BN, true, false>(epi, a, b, p, s);
}

}  // namespace gemm
}  // namespace b200

extern "C" int b200_gemm_bf16_ex(const void* A, const void* B, void* C, const void* aux, int M, int N, int K, int lda,
                                 int ldb, int ldc, int ld_aux, int a_mn_major, int b_mn_major, int epilogue,
                                 int force_bn, int rope_S, int rope_dh, int rope_cols, int rope_ctx, void* stream);

extern "C" int b200_gemm_bf16(const void* A, const void* B, void* C, const void* aux, int M, int N, int K, int lda,
                              int ldb, int ldc, int ld_aux, int a_mn_major, int b_mn_major, int epilogue,
                              int force_bn, void* stream) {
  return b200_gemm_bf16_ex(A, B, C, aux, M, N, K, lda, ldb, ldc, ld_aux, a_mn_major, b_mn_major, epilogue, force_bn,
                           0, 0, 0, 0, stream);
}

extern "C" int b200_gemm_bf16_ex(const void* A, const void* B, void* C, const void* aux, int M, int N, int K, int lda,
                                 int ldb, int ldc, int ld_aux, int a_mn_major, int b_mn_major, int epilogue,
                                 int force_bn, int rope_S, int rope_dh, int rope_cols, int rope_ctx, void* stream) {
  return b200::gemm::gemm1_run(A, B, C, aux, M, N, K, lda, ldb, ldc, ld_aux, a_mn_major, b_mn_major, epilogue, force_bn,
                               rope_S, rope_dh, rope_cols, rope_ctx, nullptr, stream);
}

int b200::gemm::gemm1_run(const void* A, const void* B, void* C, const void* aux, int M, int N, int K, int lda, int ldb,
                          int ldc, int ld_aux, int a_mn_major, int b_mn_major, int epilogue, int force_bn, int rope_S,
                          int rope_dh, int rope_cols, int rope_ctx, const AdamwOperands* opt, void* stream) {
  using namespace b200;
  using namespace b200::gemm;
  B200_REQUIRE(M > 0 && N > 0 && K > 0, "gemm: empty problem %dx%dx%d", M, N, K);
  B200_REQUIRE((lda % 8) == 0 && (ldb % 8) == 0, "gemm: lda/ldb must be multiples of 8 elements (TMA 16-byte strides)");
  const bool a_mn = a_mn_major != 0, b_mn = b_mn_major != 0;
  int bn = force_bn > 0 ? force_bn : pick_bn(M, N, a_mn || b_mn);
  B200_REQUIRE(bn == 128 || bn == 192 || bn == 256, "gemm: BN must be 128, 192 or 256");

  CUtensorMap tmA, tmB;
  {
    uint64_t dims[2], strides[1];
    uint32_t box[2];
    if (!a_mn) { dims[0] = (uint64_t)K; dims[1] = (uint64_t)M; box[0] = BK; box[1] = BM; }
    else       { dims[0] = (uint64_t)M; dims[1] = (uint64_t)K; box[0] = 64; box[1] = BK; }
    strides[0] = (uint64_t)lda * 2;
    int rc = encode_tensor_map(&tmA, TmDtype::BF16, 2, A, dims, strides, box);
    if (rc) return rc;
    if (!b_mn) { dims[0] = (uint64_t)K; dims[1] = (uint64_t)N; box[0] = BK; box[1] = (uint32_t)bn; }
    else       { dims[0] = (uint64_t)N; dims[1] = (uint64_t)K; box[0] = 64; box[1] = BK; }
    strides[0] = (uint64_t)ldb * 2;
    rc = encode_tensor_map(&tmB, TmDtype::BF16, 2, B, dims, strides, box);
    if (rc) return rc;
  }
  Params p;
  p.C = C; p.aux = aux; p.M = M; p.N = N; p.K = K; p.ldc = ldc; p.ld_aux = ld_aux;
  p.num_m = ceil_div(M, BM); p.num_n = ceil_div(N, bn); p.num_k = ceil_div(K, BK);
  p.rope_S = rope_S; p.rope_dh = rope_dh; p.rope_cols = rope_cols; p.rope_ctx = rope_ctx;
  p.opt = AdamwOperands{};
  if (opt) p.opt = *opt;
  B200_REQUIRE(epilogue != EPI_ADAMW || opt != nullptr, "gemm: the AdamW epilogue needs its optimizer operands");
  if (epilogue == EPI_BF16_ROPE)
    B200_REQUIRE(aux && rope_S > 0 && rope_dh > 0 && (rope_dh % 2) == 0 && rope_S <= rope_ctx && (rope_cols % 2) == 0,
                 "gemm: RoPE epilogue needs the cos/sin cache, S <= ctx and an even head dim");
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  switch (bn) {
    case 256: return dispatch_major<256>(a_mn, b_mn, epilogue, tmA, tmB, p, s);
    case 192: return dispatch_major<192>(a_mn, b_mn, epilogue, tmA, tmB, p, s);
    default: return dispatch_major<128>(a_mn, b_mn, epilogue, tmA, tmB, p, s);
  }
}
