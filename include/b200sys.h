/* libb200sys — C ABI of the hand-written sm_100a (B200) kernels.
 *
 * Conventions
 *  - every entry point returns 0 on success; on failure a non-zero code and a message retrievable
 *    with b200_last_error() (thread-local);
 *  - pointers are device pointers unless the name ends in _host; `stream` is a cudaStream_t;
 *  - nothing here synchronises the host with the device;
 *  - "ref:" comments name the code of the reference (kkaitlyn111/cs336-a2-systems) whose work the
 *    kernel replaces.
 */
#ifndef B200SYS_H_
#define B200SYS_H_

#ifdef __cplusplus
extern "C" {
#endif

enum { B200_OK = 0, B200_ERR_ARG = 1, B200_ERR_CUDA = 2, B200_ERR_UNSUPPORTED = 3 };
enum { B200_BF16 = 0, B200_F32 = 1 };

const char* b200_last_error(void);
int b200_version(void);
/* kernels launched by this library in this process so far (every entry point counts its own launches) */
long long b200_launch_count(void);
int b200_device_info(int* sm_count, int* cc_major, int* cc_minor);

/* ---- tcgen05 GEMM --------------------------------------------------------------------------
 * C[M,N] = A · Bᵀ, bf16 operands, fp32 accumulation in tensor memory.
 *   a_mn_major = 0: A stored [M,K] row-major (lda = row stride);  1: A stored [K,M] row-major.
 *   b_mn_major = 0: B stored [N,K] row-major (ldb = row stride);  1: B stored [K,N] row-major.
 *   epilogue: 0 C bf16 = acc | 1 C fp32 = acc | 2 C fp32 += acc |
 *             3 C bf16 [M,N/2] = silu(acc[:,2j])*acc[:,2j+1] | 4 C fp32 = aux fp32 + acc |
 *             6 Ct fp32 [N,M] = acc transposed (ldc = row stride of Ct; both operands MN-major)
 *   force_bn: 0 = heuristic, else tile width 128/192/256.
 * ref: every Linear of the LM → cuBLAS (cs336-basics/cs336_basics/model.py:41). */
int b200_gemm_bf16(const void* A, const void* B, void* C, const void* aux, int M, int N, int K, int lda, int ldb,
                   int ldc, int ld_aux, int a_mn_major, int b_mn_major, int epilogue, int force_bn, void* stream);
/* Same, plus epilogue 5: C bf16 = acc with RoPE (interleaved pairs, ref model.py:135-147) applied to columns
 * < rope_cols — the fused QKV projection writes rotated q,k directly.  Row r is position r % rope_S, column c is
 * element c % rope_dh of its head; aux = fp32 cos/sin cache (2, rope_ctx, rope_dh/2). */
int b200_gemm_bf16_ex(const void* A, const void* B, void* C, const void* aux, int M, int N, int K, int lda, int ldb,
                      int ldc, int ld_aux, int a_mn_major, int b_mn_major, int epilogue, int force_bn, int rope_S,
                      int rope_dh, int rope_cols, int rope_ctx, void* stream);

/* Weight-gradient GEMM whose epilogue is the optimizer.  dW[M, N] = dy^T x with dy: [T, M], x: [T, N] (bf16, row
 * strides ld_dy / ld_x); the accumulator tile never becomes a gradient tensor: the epilogue reads p, m, v of the same
 * elements, applies AdamW (mode 0: course arithmetic, 1: torch.optim.AdamW) and writes p, m, v and the bf16 GEMM
 * shadow.  The M rows are the rows of 1..3 parameters stacked in the order of their fused forward GEMM (q|k|v, w1|w3);
 * each segment has its own contiguous [rows_i, N] fp32 p / m / v, the shadow is the stacked [M, N] bf16 operand.
 * hyper_dev: device floats lr, beta1, beta2, eps, weight_decay, grad_scale; step_dev: device int t of this update
 * (advance it once per optimisation step with b200_incr_step).  grad_out_f32 (nullable, [M, N]) also receives dW.
 * ref: the wgrad cuBLAS GEMM of Linear.backward (model.py:41) + the per-parameter AdamW (optimizer.py:50-86). */
typedef struct {
  float* p[3];
  float* m[3];
  float* v[3];
  int rows[3];
  int nseg;
  int mode;
  void* shadow_bf16;
  const float* hyper_dev;
  const int* step_dev;
} b200_wgrad_adamw_args;
int b200_wgrad_adamw(const void* dy, const void* x, int T, int M, int N, int ld_dy, int ld_x,
                     const b200_wgrad_adamw_args* args, void* grad_out_f32, int pair, int force_bn, void* stream);

/* ---- row softmax -------------------------------------------------------------------------------
 * out[r, :] = softmax(scale * in[r, :]) for `rows` rows of `cols` columns (row strides ld_in / ld_out in elements;
 * fp32 or bf16 in and out).  causal_period > 0: row r is query position r % causal_period and only sees columns
 * 0..position; masked outputs are 0.  b200_softmax_bwd_rows: ds = scale * p * (dp - rowsum(dp * p)).
 * ref: softmax (cs336_basics/nn_utils.py:4-7) and the where(mask, s, -inf) + softmax of the materialised
 * attention (model.py:425-432). */
int b200_softmax_rows(const void* in, int in_dtype, void* out, int out_dtype, long long rows, int cols,
                      long long ld_in, long long ld_out, float scale, int causal_period, void* stream);
int b200_softmax_bwd_rows(const void* p, const void* dp, void* ds, int dtype, long long rows, int cols, long long ld,
                          float scale, void* stream);

/* ---- FlashAttention-2 (tcgen05/TMEM/TMA) ------------------------------------------------------
 * q,k,v,o: bf16 with element strides (batch, head, row); last dim contiguous, D in {16,32,64,80,96,128}.
 * lse: fp32 [B,H,Nq] contiguous.  scale = 1/sqrt(D) is applied inside.
 * ref: Triton flash_attention_kernel (cs336_systems/flash_attention.py:137-266). */
int b200_fa2_fwd(const void* q, const void* k, const void* v, void* o, float* lse, int B, int H, int Nq, int Nk,
                 int D, const long long* q_strides, const long long* k_strides, const long long* v_strides,
                 const long long* o_strides, int causal, float scale, void* stream);
/* Extended form.  out_f32 = 1: o is fp32 (same element strides) — fp32 callers get an unrounded O and no cast pass.
 * cluster = 1 (what b200_fa2_fwd uses): when the number of 128-row query tiles is even, pairs of CTAs form a
 * thread-block cluster and share every K/V tile through TMA multicast; 0 = one independent CTA per query tile. */
int b200_fa2_fwd_ex(const void* q, const void* k, const void* v, void* o, float* lse, int B, int H, int Nq, int Nk,
                    int D, const long long* q_strides, const long long* k_strides, const long long* v_strides,
                    const long long* o_strides, int causal, float scale, int out_f32, int cluster, void* stream);
/* delta: fp32 [B,H,Nq] scratch (rowsum(O∘dO)); dq_acc: fp32 [B,H,Nq,D] scratch, zeroed inside.
 * dq,dk,dv: bf16 with their own strides.
 * ref: torch.compile'd backward_pass_recomp materialising N×N matrices (flash_attention.py:270-289). */
int b200_fa2_bwd(const void* q, const void* k, const void* v, const void* o, const float* lse, const void* dout,
                 void* dq, void* dk, void* dv, float* delta, float* dq_acc, int B, int H, int Nq, int Nk, int D,
                 const long long* q_strides, const long long* k_strides, const long long* v_strides,
                 const long long* o_strides, const long long* do_strides, const long long* dq_strides,
                 const long long* dk_strides, const long long* dv_strides, int causal, float scale, void* stream);

/* Extended form: rope_cache (nullable, fp32 (2, rope_ctx, D/2)) = q,k were rotated before attention, so dq and dk are
 * un-rotated on the way out (fused into the fp32->bf16 pass of dq and the dK epilogue).  flags bit 0: dq_acc is
 * already all-zero and must be left all-zero (the convert pass re-zeroes it; saves a memset per call); bit 1: o is
 * fp32; bit 2: dout is fp32 — do_bf16_scratch (B*H*Nq*D bf16) then receives its bf16 copy from the delta pre-pass;
 * bit 3: dq, dk, dv are fp32 (not together with rope_cache).  The fp32 forms serve fp32 callers
 * (reference tests/test_attention.py:29-40) without any separate cast kernel on the way in or out. */
int b200_fa2_bwd_ex(const void* q, const void* k, const void* v, const void* o, const float* lse, const void* dout,
                    void* dq, void* dk, void* dv, float* delta, float* dq_acc, int B, int H, int Nq, int Nk, int D,
                    const long long* q_strides, const long long* k_strides, const long long* v_strides,
                    const long long* o_strides, const long long* do_strides, const long long* dq_strides,
                    const long long* dk_strides, const long long* dv_strides, int causal, float scale,
                    const float* rope_cache, int rope_ctx, int flags, void* do_bf16_scratch, void* stream);

/* ---- fused normalisation / elementwise ---------------------------------------------------------
 * ref: RMSNorm model.py:101-107, residual adds model.py:380-386. */
int b200_rmsnorm_fwd(const void* x, int x_dtype, const void* resid_bf16 /*nullable*/, const float* w, void* y,
                     int y_dtype, float* xout /*nullable: x+resid in fp32*/, float* rstd /*nullable*/, int T, int d,
                     float eps, void* stream);
int b200_rmsnorm_bwd_workspace_rows(int T);
int b200_rmsnorm_bwd(const void* dy, int dy_dtype, const float* x, const float* w, const float* rstd,
                     const float* dres_in /*nullable*/, float* dx /*nullable*/, void* dx_bf16 /*nullable*/, float* dw,
                     int dw_accumulate, float* scratch /* workspace_rows(T)*d floats */, int T, int d, void* stream);
/* Allocation-free variant of b200_rmsnorm_bwd (no counterpart in the reference: it is an addition of this
 * library).  For callers that may not allocate device memory while running — a host whose memory budget is
 * owned by its own pool, a loop that cannot take cudaMalloc's device-wide synchronisation, a plug-in ABI that
 * passes inputs and outputs only.  The partial-sum scratch lives in a fixed array inside the library image
 * (b200_rmsnorm_bwd_noalloc_max_floats() floats = 4.75 MiB, resident on every GPU that loads the library
 * until exit); a problem with workspace_rows(T)*d beyond that is rejected with B200_ERR_ARG.  The first call
 * on each device (with that device current) resolves the array's device address — make it before any stream
 * capture.  All calls share the array: calls that overlap on the GPU race, serialise them.  Never calls
 * cudaMalloc/cudaFree or an allocating library routine. */
long long b200_rmsnorm_bwd_noalloc_max_floats(void);
int b200_rmsnorm_bwd_noalloc(const void* dy, int dy_dtype, const float* x, const float* w, const float* rstd,
                             const float* dres_in /*nullable*/, float* dx /*nullable*/, void* dx_bf16 /*nullable*/,
                             float* dw, int dw_accumulate, int T, int d, void* stream);
int b200_residual_add(const float* x, const void* y_bf16, float* out, long long n, void* stream);
/* ref: RoPE model.py:135-147 (interleaved pairs); qkv is [B,S,3,H,dh] bf16, q and k rotated in place. */
int b200_rope_qk_inplace(void* qkv_bf16, const float* cache /*(2,ctx,dh/2)*/, int ctx, int B, int S, int H, int dh,
                         int inverse, void* stream);
/* ref: silu(w1 x) * w3 x, model.py:397,526-527.  ab = [T, 2F] (gate | up). */
int b200_swiglu_fwd(const void* ab, void* g, long long T, int F, void* stream);
int b200_swiglu_bwd(const void* ab, const void* dg, void* dab, long long T, int F, void* stream);
/* ref: Embedding gather model.py:57 and its scatter-add backward. */
int b200_embedding_fwd(const long long* ids, const float* w, float* out, long long T, int d, int vocab, void* stream);
int b200_embedding_bwd(const long long* ids, const float* dout, float* dw, long long T, int d, int vocab,
                       int zero_first, void* stream);
int b200_cast_f32_to_bf16(const float* in, void* out, long long n, void* stream);
/* Up to three contiguous fp32 -> bf16 casts in ONE launch (stop at the first n <= 0): the q, k, v operands of an fp32
 * attention call.  ref: the implicit per-tensor casts of the Triton kernel's tl.dot operands (flash_attention.py:228). */
int b200_cast3_f32_to_bf16(const float* in0, void* out0, long long n0, const float* in1, void* out1, long long n1,
                           const float* in2, void* out2, long long n2, void* stream);

/* ---- softmax cross-entropy -------------------------------------------------------------------
 * ref: cs336_basics/nn_utils.py:9-17. */
int b200_xent_fwd(const void* logits, int dtype, const long long* targets, float* row_loss, float* lse, float* loss,
                  int T, int V, int ld, void* stream);
int b200_xent_bwd(const void* logits, int dtype, const long long* targets, const float* lse, const float* dloss,
                  void* dlogits, int T, int V, int ld, int ld_out, void* stream);

/* ---- multi-tensor AdamW / gradient clipping ----------------------------------------------------
 * ref: cs336_basics/optimizer.py:50-86 and nn_utils.py:20-30 (per-tensor Python loops). */
typedef struct {
  float* p;
  const float* g;
  float* m;
  float* v;
  void* shadow_bf16; /* nullable: bf16 copy of p refreshed by the update */
  long long n;
  int aligned16; /* 0: unaligned, 1: all pointers 16-byte aligned, 2: all 32-byte aligned (256-bit streaming path) */
  int pad_;
} b200_adamw_tensor;
int b200_adamw_chunk_elems(void);
/* hyper_dev: [lr, beta1, beta2, eps, weight_decay, grad_scale]; mode 0 = course AdamW, 1 = torch.optim.AdamW */
int b200_adamw_multi(const b200_adamw_tensor* table_dev, const long long* chunk_start_dev, int n_tensors,
                     long long total_chunks, const float* hyper_dev, int* step_dev, int mode, int increment_step,
                     void* stream);
/* Same with a cap on the grid (max_blocks > 0): a small footprint (e.g. 2 CTAs per SM) lets the update run
 * on a side stream underneath backward GEMMs without crowding their CTAs out of the SMs. */
int b200_adamw_multi_grid(const b200_adamw_tensor* table_dev, const long long* chunk_start_dev, int n_tensors,
                          long long total_chunks, const float* hyper_dev, int* step_dev, int mode, int increment_step,
                          int max_blocks, void* stream);
int b200_gradnorm_partials(long long total_chunks);
int b200_clip_grad_norm(const b200_adamw_tensor* table_dev, const long long* chunk_start_dev, int n_tensors,
                        long long total_chunks, float max_norm, float eps, float* partial_dev, float* norm_out_dev,
                        int apply_scale, void* stream);

/* ---- NVLink symmetric-memory collectives -------------------------------------------------------
 * peer_ptrs_host / signal_ptrs_host: HOST arrays of world_size device addresses — the peer-mapped base of the
 * symmetric arena and of its signal pad on every rank (from torch.distributed._symmetric_memory); byte_off
 * selects the bucket.  The buffer is averaged in place on every rank (sum over ranks x scale).
 * algo: 0 auto, 1 one-shot, 2 two-shot, 3 multimem (needs multicast_ptr != 0).  Uses
 * b200_comm_signal_words() 32-bit words of the signal pad starting at signal_slot.
 * ref: dist.all_reduce(...) + grad/world_size (ddp_bucketed_overlapped_sharded.py:236,247,299,314). */
int b200_comm_signal_words(void);
int b200_allreduce(const unsigned long long* peer_ptrs_host, const unsigned long long* signal_ptrs_host,
                   void* multicast_ptr, long long byte_off, long long numel, int dtype, float scale, int rank,
                   int world_size, int algo, int blocks, int signal_slot, void* stream);
/* Broadcast by pull over NVLink: every rank != src_rank copies nbytes at byte_off of src_rank's arena into its own.
 * ref: one dist.broadcast(param, src=0) per parameter at DDP construction (...sharded.py:225-226,260-261). */
int b200_bcast_pull(const unsigned long long* peer_ptrs_host, const unsigned long long* signal_ptrs_host,
                    long long byte_off, long long nbytes, int src_rank, int rank, int world_size, int blocks,
                    int signal_slot, void* stream);
/* Broadcast by NVLS multicast store: rank src_rank writes nbytes at byte_off of its arena through the switch into
 * every rank's arena (one multimem.st per 16 bytes; the source's NVLink egress is nbytes, not (ws-1)*nbytes).
 * ref: dist.broadcast(param.data, src=0) per parameter at DDP construction (…sharded.py:225-226,260-261). */
int b200_bcast_mc(const unsigned long long* peer_ptrs_host, const unsigned long long* signal_ptrs_host,
                  void* multicast_ptr, long long byte_off, long long nbytes, int src_rank, int rank, int world_size,
                  int blocks, int signal_slot, void* stream);
/* Health of the NVLink kernels on the current device.  Their cross-rank barriers spin with a time limit
 * (b200_comm_reset sets it, default 30 s); a barrier that runs out records (rank, CTA, peer, phase, slot, flag) and
 * lets every later barrier fall through, so a lost rank shows up as status 1 here instead of a hung GPU.
 * The reference has no counterpart (NCCL's watchdog plays this role for its collectives). */
int b200_comm_status(int* info6 /*nullable*/);
int b200_comm_reset(long long timeout_ms /*0: keep*/);
/* Cross-rank barrier on the signal pads (orders peer stores before what follows on every rank). */
int b200_symm_barrier(const unsigned long long* signal_ptrs_host, int rank, int world_size, int signal_slot,
                      void* stream);
/* Owner-rank multi-tensor AdamW fused with the distribution of the new values: table[i].p points into this
 * rank's parameter arena (base local_param_base); the update is stored at the same offset of EVERY rank's
 * arena (param_peer_bases_host) and, when given, its bf16 copy at the same offset of every rank's shadow arena.
 * Follow with b200_symm_barrier before parameters are read.
 * ref: local optimizer.step + one dist.broadcast per parameter (...sharded.py:335-340). */
int b200_adamw_multi_distribute(const b200_adamw_tensor* table_dev, const long long* chunk_start_dev, int n_tensors,
                                long long total_chunks, const float* hyper_dev, int* step_dev, int mode,
                                int increment_step, const unsigned long long* param_peer_bases_host,
                                const unsigned long long* shadow_peer_bases_host /*nullable*/,
                                unsigned long long local_param_base, unsigned long long local_shadow_base,
                                int world_size, void* stream);

/* One DDP bucket, one kernel: rank r averages slice r of the gradient bucket out of every rank's symmetric grad
 * arena, applies AdamW with its slice of m / v (optimizer state sharded 1/world_size), and stores the new fp32
 * parameters and bf16 shadows into every rank's parameter / shadow arenas (same element layout as the gradient
 * bucket; shadow offset = byte_off / 2).  m_slice / v_slice: ceil(ceil(numel/8)/ws)*8 floats, this rank only.
 * defer_fp32 = 1: only the bf16 shadows cross NVLink here, plus the fp32 values of the n_eager (<= 4) element ranges
 * [eager_lo, eager_hi) of the bucket that the next forward reads in fp32; the other fp32 replicas are brought up to
 * date by b200_zero_param_gather (run it before anything reads them).
 * ref: all_reduce + grad/ws (…sharded.py:299,314) + optimizer.step + per-parameter broadcast (…:335-340). */
int b200_zero_adamw_bucket(const unsigned long long* grad_peers_host, const unsigned long long* param_peers_host,
                           const unsigned long long* shadow_peers_host /*nullable*/,
                           const unsigned long long* signal_ptrs_host,
                           const unsigned long long* multicast3_host /*nullable: NVLS addresses of grad, param, shadow arenas*/,
                           long long byte_off, long long numel, float* m_slice, float* v_slice, const float* hyper_dev,
                           const int* step_dev, int mode, int rank, int world_size, int blocks, int signal_slot,
                           int variant /*bit0: NVLS reduce, bit1: NVLS stores*/, int defer_fp32, int n_eager,
                           const long long* eager_lo_elems, const long long* eager_hi_elems, void* stream);
/* Deferred half of the above: every rank pushes its slice of the fp32 parameters of all n_buckets buckets
 * (buckets_dev: device array of {byte_off, ceil(numel*4/32)} int64 pairs) into every other rank's parameter arena
 * (peer stores, or multimem.st when param_multicast != 0), then one cross-rank barrier.
 * ref: the per-parameter dist.broadcast loop (…sharded.py:337-340). */
int b200_zero_param_gather(const unsigned long long* param_peers_host, const unsigned long long* signal_ptrs_host,
                           unsigned long long param_multicast, const long long* buckets_dev, int n_buckets, int rank,
                           int world_size, int blocks, int signal_slot, void* stream);
int b200_incr_step(int* step_dev, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200SYS_H_ */
