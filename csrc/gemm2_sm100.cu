// CTA-pair (cta_group::2) variant of the bf16 GEMM: two SMs of one cluster compute one 256×BN tile.
//
// Why: tcgen05.mma is issued by ONE thread.  With cta_group::1 that thread has to keep 64–128-cycle
// instructions fed (measured: 128-wide tiles reach only ~845 TFLOP/s, 256-wide ~1170); with cta_group::2 every
// instruction drives both SMs' tensor cores (M = 256), so the issue cost per FLOP halves, and each CTA stages
// only its half of the B tile — the pair fetches B from L2 once instead of twice.
//
// Per CTA (rank r in the pair): A = its own 128 rows of the 256-row tile, B = rows [r·BN/2, (r+1)·BN/2) of the
// BN-wide tile, accumulator = its 128 rows × BN columns in its own tensor memory.
//   warp 0  TMA producer (both CTAs): loads its A and B halves into its OWN shared memory; the bytes are credited to
//           the LEADER's full barrier (cp.async.bulk.tensor ... cta_group::2)
//   warp 1  MMA issuer (leader CTA only): tcgen05.mma.cta_group::2; tcgen05.commit multicasts the stage-free and
//           accumulator-ready arrivals to the barriers of BOTH CTAs
//   warp 2  TMEM allocator (cta_group::2, both CTAs)
//   warps 4-7 epilogue (both CTAs): own accumulator rows → registers → global; arrive on the LEADER's
//           accumulator-free barrier (remote arrive for the peer)
#include "common.cuh"
#include "ptx.cuh"
#include "tensormap.h"
#include "gemm_epilogue.cuh"

namespace b200 {
namespace gemm {

constexpr int kThreads2 = 256;
constexpr int kTmemCols2 = 512;

template <int BN>
struct Cfg2 {
  static constexpr int A_BYTES = BM * BK * 2;         // this CTA's 128 rows
  static constexpr int B_BYTES = (BN / 2) * BK * 2;   // this CTA's half of the B tile
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int MAX_STAGES = (227 * 1024 - 2048) / STAGE_BYTES;
  static constexpr int STAGES = MAX_STAGES > 8 ? 8 : MAX_STAGES;
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 + 512;
};

template <int BN, bool A_MN, bool B_MN, int EPI>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(128 + 32 * EpiWarps<EPI>::N, 1)
gemm2_bf16_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const Params p) {
  using C = Cfg2<BN>;
  constexpr int STAGES = C::STAGES;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + STAGES * C::STAGE_BYTES);  // used on the leader
  uint64_t* empty_bar = full_bar + STAGES;                                           // own
  uint64_t* tfull_bar = empty_bar + STAGES;                                          // own
  uint64_t* tempty_bar = tfull_bar + 2;                                              // used on the leader
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty_bar + 2);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t cta_rank = ptx::cluster_ctarank();
  const bool leader = cta_rank == 0;

  if (warp == 0 && ptx::elect_one()) {
    ptx::tma_prefetch_desc(&tmA);
    ptx::tma_prefetch_desc(&tmB);
  }
  if (warp == 1 && ptx::elect_one()) {
    for (int s = 0; s < STAGES; ++s) {
      ptx::mbar_init(&full_bar[s], 2);   // one arrive(+expect_tx) per CTA's producer
      ptx::mbar_init(&empty_bar[s], 1);  // multicast tcgen05.commit
    }
    for (int a = 0; a < 2; ++a) {
      ptx::mbar_init(&tfull_bar[a], 1);   // multicast tcgen05.commit
      ptx::mbar_init(&tempty_bar[a], 2 * EpiWarps<EPI>::N);  // the epilogue warps of both CTAs
    }
    ptx::fence_mbar_init();
  }
  if (warp == 2) {
    ptx::tmem_alloc_cg2(tmem_slot, kTmemCols2);
    ptx::tmem_relinquish_cg2();
  }
  ptx::griddep_launch();
  ptx::tc_fence_before();
  __syncthreads();
  ptx::cluster_sync();  // both CTAs' barriers are initialised and both halves of tensor memory are allocated
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  ptx::griddep_wait();

  const int num_m2 = (p.num_m + 1) / 2;  // 256-row tile pairs along M
  const int total_tiles = num_m2 * p.num_n;
  const int cluster_id = blockIdx.x >> 1;
  const int num_clusters = gridDim.x >> 1;

  if (warp == 0) {
    // ================================ TMA producer (both CTAs) ================================
    if (ptx::elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      for (int t = cluster_id; t < total_tiles; t += num_clusters) {
        const int m0 = ((t % num_m2) * 2 + (int)cta_rank) * BM;
        // this CTA's half of the B tile; EPI_SWIGLU_FWD: rank 0 takes 128 rows of w1, rank 1 the same rows of w3
        // (row F + ... of the stacked weight), so a tile's accumulator is [a of 128 hidden units | their b]
        const int n0 = EPI == EPI_SWIGLU_FWD ? (t / num_m2) * (BN / 2) + (int)cta_rank * (p.N >> 1)
                                             : (t / num_m2) * BN + (int)cta_rank * (BN / 2);
        for (int kb = 0; kb < p.num_k; ++kb) {
          ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
          const uint32_t lead_full = ptx::mapa(ptx::smem_u32(&full_bar[stage]), 0);
          ptx::mbar_expect_tx_cluster(lead_full, C::STAGE_BYTES);
          uint8_t* sa = smem + stage * C::STAGE_BYTES;
          uint8_t* sb = sa + C::A_BYTES;
          if constexpr (!A_MN) {
            ptx::tma_load_2d_cg2(sa, &tmA, lead_full, kb * BK, m0, ptx::kEvictNormal);
          } else {
#pragma unroll
            for (int c = 0; c < BM / 64; ++c)
              ptx::tma_load_2d_cg2(sa + c * (64 * BK * 2), &tmA, lead_full, m0 + c * 64, kb * BK, ptx::kEvictNormal);
          }
          if constexpr (!B_MN) {
            ptx::tma_load_2d_cg2(sb, &tmB, lead_full, kb * BK, n0, ptx::kEvictNormal);
          } else {
#pragma unroll
            for (int c = 0; c < BN / 128; ++c)
              ptx::tma_load_2d_cg2(sb + c * (64 * BK * 2), &tmB, lead_full, n0 + c * 64, kb * BK, ptx::kEvictNormal);
          }
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ================================ MMA issuer (leader only) ================================
    if (leader && ptx::elect_one()) {
      constexpr uint32_t idesc = ptx::make_idesc(2 * BM, BN, A_MN, B_MN);
      int stage = 0;
      uint32_t phase = 0;
      int acc = 0;
      uint32_t acc_phase = 0;
      for (int t = cluster_id; t < total_tiles; t += num_clusters) {
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
            const uint64_t adesc = A_MN ? ptx::make_smem_desc(a_addr + k * 2048, 64 * BK * 2, 1024)
                                        : ptx::make_smem_desc(a_addr + k * 32, 16, 1024);
            const uint64_t bdesc = B_MN ? ptx::make_smem_desc(b_addr + k * 2048, 64 * BK * 2, 1024)
                                        : ptx::make_smem_desc(b_addr + k * 32, 16, 1024);
            ptx::mma_f16_ss_cg2(d_tmem, adesc, bdesc, idesc, (kb | k) != 0 ? 1u : 0u);
          }
          ptx::mma_commit_cg2_mcast(&empty_bar[stage], 0x3);  // stage reusable in BOTH CTAs
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
        ptx::mma_commit_cg2_mcast(&tfull_bar[acc], 0x3);  // accumulator halves complete in both CTAs
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
      }
    }
  } else if (warp >= 4) {
    // ================================ epilogue (both CTAs) ================================
    const int ew = (warp - 4) & 3;
    constexpr int NGRP = EpiWarps<EPI>::N / 4;  // column groups: group g takes the 32-column chunks c = g, g + NGRP, …
    const int grp = (warp - 4) >> 2;
    int acc = 0;
    uint32_t acc_phase = 0;
    AdamwConsts ak;
    if constexpr (EPI == EPI_ADAMW) ak = adamw_consts(p.opt);
    for (int t = cluster_id; t < total_tiles; t += num_clusters) {
      const int m0 = ((t % num_m2) * 2 + (int)cta_rank) * BM;
      const int n0 = (t / num_m2) * BN;
      ptx::mbar_wait(&tfull_bar[acc], acc_phase);
      ptx::tc_fence_after();
      const int row = m0 + ew * 32 + lane;
      const uint32_t t_row = tmem_base + (static_cast<uint32_t>(ew * 32) << 16) + acc * BN;
      if constexpr (EPI == EPI_SWIGLU_FWD) {
        const int hid0 = (t / num_m2) * (BN / 2);  // first hidden unit of this tile
#pragma unroll 1
        for (int c = 0; c < BN / 64; ++c) {
          if (hid0 + c * 32 >= (p.N >> 1)) break;
          uint32_t ra[32], rb[32];
          ptx::tmem_ld_32x32b_x32(t_row + c * 32, ra);
          ptx::tmem_ld_32x32b_x32(t_row + BN / 2 + c * 32, rb);
          ptx::tmem_ld_wait();
          swiglu_fwd_chunk(p, row, hid0 + c * 32, ra, rb);
        }
      } else {
#pragma unroll 1
        for (int c = grp; c < BN / 32; c += NGRP) {
          if (n0 + c * 32 >= p.N) break;
          uint32_t r[32];
          ptx::tmem_ld_32x32b_x32(t_row + c * 32, r);
          ptx::tmem_ld_wait();
          if constexpr (EPI == EPI_ADAMW) adamw_chunk_t(p, ak, row, n0 + c * 32, r);
          else if constexpr (EPI == EPI_SWIGLU_BWD) swiglu_bwd_chunk(p, row, n0 + c * 32, r);
          else store_chunk<EPI>(p, row, n0 + c * 32, r);
        }
      }
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive_cluster(ptx::mapa(ptx::smem_u32(&tempty_bar[acc]), 0));
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
  }

  ptx::tc_fence_before();
  __syncthreads();
  ptx::cluster_sync();  // the peer may still be reading our shared memory / signalling our barriers
  if (warp == 2) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc_cg2(tmem_base, kTmemCols2);
  }
}

// ------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------
static int pick_bn2(int M, int N, bool b_mn) {
  const int cands[3] = {256, 192, 128};
  const double penalty[3] = {1.0, 1.02, 1.10};
  const int clusters = num_sms() / 2;
  int best = 256;
  double best_cost = 1e30;
  const int nm2 = ceil_div(ceil_div(M, BM), 2);
  for (int i = 0; i < 3; ++i) {
    const int bn = cands[i];
    if (b_mn && (bn % 128) != 0) continue;  // each CTA's half must be whole 64-wide MN chunks
    const int tiles = nm2 * ceil_div(N, bn);
    const int waves = ceil_div(tiles, clusters);
    const double cost = (double)waves * bn * penalty[i];
    if (cost < best_cost) { best_cost = cost; best = bn; }
  }
  return best;
}

template <int BN, bool A_MN, bool B_MN, int EPI>
static int launch2(const CUtensorMap& tmA, const CUtensorMap& tmB, const Params& p, cudaStream_t stream) {
  using C = Cfg2<BN>;
  auto kern = gemm2_bf16_kernel<BN, A_MN, B_MN, EPI>;
  B200_ENSURE_DYN_SMEM(kern, C::SMEM_BYTES);
  const int tiles = ceil_div(p.num_m, 2) * p.num_n;
  const int clusters = num_sms() / 2;
  const int grid = 2 * (tiles < clusters ? tiles : clusters);
  B200_CHECK_CUDA(launch_pdl(kern, dim3(grid), dim3(128 + 32 * EpiWarps<EPI>::N), C::SMEM_BYTES, stream, tmA, tmB, p));
  return 0;
}

template <int BN, bool A_MN, bool B_MN>
static int dispatch_epi2(int epi, const CUtensorMap& a, const CUtensorMap& b, const Params& p, cudaStream_t s) {
  switch (epi) {
    case EPI_BF16: return launch2<BN, A_MN, B_MN, EPI_BF16>(a, b, p, s);
    case EPI_F32: return launch2<BN, A_MN, B_MN, EPI_F32>(a, b, p, s);
    case EPI_F32_T:
      if constexpr (A_MN && B_MN) return launch2<BN, A_MN, B_MN, EPI_F32_T>(a, b, p, s);
      break;
    case EPI_BF16_ROPE:
      if constexpr (!A_MN && !B_MN) return launch2<BN, A_MN, B_MN, EPI_BF16_ROPE>(a, b, p, s);
      break;
    case EPI_ADAMW:
      if constexpr (A_MN && B_MN) return launch2<BN, A_MN, B_MN, EPI_ADAMW>(a, b, p, s);
      break;
    case EPI_SWIGLU_FWD:
      if constexpr (!A_MN && !B_MN && BN == 256) return launch2<BN, A_MN, B_MN, EPI_SWIGLU_FWD>(a, b, p, s);
      break;
    case EPI_SWIGLU_BWD:
      if constexpr (!A_MN && B_MN) return launch2<BN, A_MN, B_MN, EPI_SWIGLU_BWD>(a, b, p, s);
      break;
  }
  set_error("gemm2: epilogue %d not available in the CTA-pair kernel for this layout", epi);
  return B200_ERR_UNSUPPORTED;
}

template <int BN>
static int dispatch_major2(bool a_mn, bool b_mn, int epi, const CUtensorMap& a, const CUtensorMap& b, const Params& p,
                           cudaStream_t s) {
  if (!a_mn && !b_mn) return dispatch_epi2<BN, false, false>(epi, a, b, p, s);
  if (!a_mn && b_mn) return dispatch_epi2<BN, false, true>(epi, a, b, p, s);
  if (a_mn && b_mn) return dispatch_epi2<BN, true, true>(epi, a, b, p, s);
  set_error("gemm2: A MN-major with B K-major is not instantiated");
  return B200_ERR_UNSUPPORTED;
}

}  // namespace gemm
}  // namespace b200

// Same contract as b200_gemm_bf16_ex, executed by CTA pairs (tcgen05 cta_group::2, 256 x BN tiles).
// Supported epilogues: 0 (bf16), 1 (fp32), 5 (RoPE, NT only), 6 (transposed fp32, TN only).
extern "C" int b200_gemm2_bf16_ex(const void* A, const void* B, void* C, const void* aux, int M, int N, int K, int lda,
                                  int ldb, int ldc, int ld_aux, int a_mn_major, int b_mn_major, int epilogue,
                                  int force_bn, int rope_S, int rope_dh, int rope_cols, int rope_ctx, void* stream) {
  return b200::gemm::gemm2_run(A, B, C, aux, M, N, K, lda, ldb, ldc, ld_aux, a_mn_major, b_mn_major, epilogue, force_bn,
                               rope_S, rope_dh, rope_cols, rope_ctx, nullptr, stream);
}

// Weight-gradient GEMM with the optimizer in its epilogue: dW[M, N] = dyᵀ·x (dy: [T, M], x: [T, N], bf16 row-major) is
// never stored — each accumulator tile goes straight through AdamW against the p / m / v of the same elements.
extern "C" int b200_wgrad_adamw(const void* dy, const void* x, int T, int M, int N, int ld_dy, int ld_x,
                                const b200_wgrad_adamw_args* a, void* grad_out_f32, int pair, int force_bn, void* stream) {
  using namespace b200;
  using namespace b200::gemm;
  B200_REQUIRE(a != nullptr && a->nseg >= 1 && a->nseg <= 3, "wgrad_adamw: 1..3 parameter segments");
  B200_REQUIRE(a->hyper_dev && a->step_dev, "wgrad_adamw: hyper-parameters and step counter must be device pointers");
  AdamwOperands o{};
  int rows = 0;
  for (int i = 0; i < a->nseg; ++i) {
    B200_REQUIRE(a->p[i] && a->m[i] && a->v[i] && a->rows[i] > 0, "wgrad_adamw: segment %d incomplete", i);
    B200_REQUIRE((reinterpret_cast<uintptr_t>(a->p[i]) | reinterpret_cast<uintptr_t>(a->m[i]) |
                  reinterpret_cast<uintptr_t>(a->v[i])) % 16 == 0, "wgrad_adamw: p/m/v must be 16-byte aligned");
    o.p[i] = a->p[i]; o.m[i] = a->m[i]; o.v[i] = a->v[i];
    o.row0[i] = rows;
    rows += a->rows[i];
  }
  for (int i = a->nseg; i < 4; ++i) o.row0[i] = rows;
  B200_REQUIRE(rows == M, "wgrad_adamw: segments cover %d rows, the gradient has %d", rows, M);
  B200_REQUIRE(!a->shadow_bf16 || reinterpret_cast<uintptr_t>(a->shadow_bf16) % 8 == 0, "wgrad_adamw: shadow alignment");
  o.nseg = a->nseg; o.mode = a->mode;
  o.shadow = reinterpret_cast<__nv_bfloat16*>(a->shadow_bf16);
  o.hyper = a->hyper_dev; o.step = a->step_dev;
  // the GEMM runs on the transpose (dWᵀ = xᵀ·dy: rows = input features, columns = stacked parameter rows) so that the
  // epilogue's accesses to p / m / v are coalesced — see adamw_chunk_t
  auto run = (pair && N >= 256) ? gemm2_run : gemm1_run;
  return run(x, dy, grad_out_f32, nullptr, N, M, T, ld_x, ld_dy, N, 0, 1, 1, EPI_ADAMW, force_bn, 0, 0, 0, 0, &o, stream);
}

int b200::gemm::gemm2_run(const void* A, const void* B, void* C, const void* aux, int M, int N, int K, int lda, int ldb,
                          int ldc, int ld_aux, int a_mn_major, int b_mn_major, int epilogue, int force_bn, int rope_S,
                          int rope_dh, int rope_cols, int rope_ctx, const AdamwOperands* opt, void* stream) {
  using namespace b200;
  using namespace b200::gemm;
  B200_REQUIRE(M > 0 && N > 0 && K > 0, "gemm2: empty problem %dx%dx%d", M, N, K);
  B200_REQUIRE((lda % 8) == 0 && (ldb % 8) == 0, "gemm2: lda/ldb must be multiples of 8 elements");
  const bool a_mn = a_mn_major != 0, b_mn = b_mn_major != 0;
  const int bn = force_bn > 0 ? force_bn : pick_bn2(M, N, b_mn);
  B200_REQUIRE(bn == 128 || bn == 256 || (bn == 192 && !b_mn), "gemm2: unsupported tile width %d", bn);
  CUtensorMap tmA, tmB;
  {
    uint64_t dims[2], strides[1];
    uint32_t box[2];
    if (!a_mn) { dims[0] = (uint64_t)K; dims[1] = (uint64_t)M; box[0] = BK; box[1] = BM; }
    else       { dims[0] = (uint64_t)M; dims[1] = (uint64_t)K; box[0] = 64; box[1] = BK; }
    strides[0] = (uint64_t)lda * 2;
    int rc = encode_tensor_map(&tmA, TmDtype::BF16, 2, A, dims, strides, box);
    if (rc) return rc;
    if (!b_mn) { dims[0] = (uint64_t)K; dims[1] = (uint64_t)N; box[0] = BK; box[1] = (uint32_t)(bn / 2); }
    else       { dims[0] = (uint64_t)N; dims[1] = (uint64_t)K; box[0] = 64; box[1] = BK; }
    strides[0] = (uint64_t)ldb * 2;
    rc = encode_tensor_map(&tmB, TmDtype::BF16, 2, B, dims, strides, box);
    if (rc) return rc;
  }
  Params p;
  p.C = C; p.aux = aux; p.M = M; p.N = N; p.K = K; p.ldc = ldc; p.ld_aux = ld_aux;
  p.num_m = ceil_div(M, BM); p.num_n = ceil_div(N, bn); p.num_k = ceil_div(K, BK);
  if (epilogue == EPI_SWIGLU_FWD) {
    B200_REQUIRE(bn == 256 && (N % 256) == 0 && !a_mn && !b_mn && aux != nullptr,
                 "gemm2: the SwiGLU forward epilogue needs 256-wide tiles, N = 2F with F a multiple of 128, K-major operands");
    p.num_n = (N / 2) / 128;  // one tile = 128 hidden units = 128 rows of w1 + the same 128 rows of w3
  }
  p.rope_S = rope_S; p.rope_dh = rope_dh; p.rope_cols = rope_cols; p.rope_ctx = rope_ctx;
  p.opt = AdamwOperands{};
  if (opt) p.opt = *opt;
  B200_REQUIRE(epilogue != EPI_ADAMW || opt != nullptr, "gemm2: the AdamW epilogue needs its optimizer operands");
  cudaStream_t s = reinterpret_cast<cudaStream_t>(stream);
  switch (bn) {
    case 256: return dispatch_major2<256>(a_mn, b_mn, epilogue, tmA, tmB, p, s);
    case 192: return dispatch_major2<192>(a_mn, b_mn, epilogue, tmA, tmB, p, s);
    default: return dispatch_major2<128>(a_mn, b_mn, epilogue, tmA, tmB, p, s);
  }
}
