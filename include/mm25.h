/*
 * mm25.h — C-ABI of libmm25.so: B200 (sm_100a) kernels for the contrastive-loss hot path
 * of neskech/Multimodal-2025 (losses/clipLoss.py, cloobLoss.py, vclipLoss.py,
 * alignCLIPLoss.py, losses/distributions/*, metrics/topk.py).
 *
 * Conventions:
 *  - every pointer is a DEVICE pointer owned by the caller; kernels never allocate or free
 *    and retain nothing past return; launches are asynchronous on `stream`
 *    (a cudaStream_t / CUstream passed as void*);
 *  - return 0 on success, a negative MM25_ERR_* otherwise (checked host-side before launch);
 *  - matrices are row-major contiguous; `dtype` is MM25_F32 or MM25_BF16 for embedding
 *    inputs, all statistics / losses / scratch gradients are fp32;
 *  - stateless and re-entrant (autograd calls backward from another host thread).
 *
 * The reference has no FFI: each entry point replaces a block of eager ATen calls, cited
 * per function as `reference file:line`.
 */
#ifndef MM25_H_
#define MM25_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum { MM25_F32 = 0, MM25_BF16 = 1 };

enum {
  MM25_OK = 0,
  MM25_ERR_SHAPE = -1,     /* bad / unsupported sizes */
  MM25_ERR_DTYPE = -2,
  MM25_ERR_ALIGN = -3,     /* pointer or leading dimension not aligned as required */
  MM25_ERR_WORKSPACE = -4, /* workspace too small */
  MM25_ERR_ARCH = -5,      /* device is not sm_100 */
  MM25_ERR_CUDA = -6,      /* a CUDA runtime / driver call failed (see mm25_last_cuda_error) */
  MM25_ERR_ARG = -7
};

/* slots of the per-kernel timing hook (mm25_debug_timing) */
enum {
  MM25_T_SIM_FWD = 0,      /* sim_fwd_tc_kernel */
  MM25_T_GEMM_TE = 1,      /* gemm_te_kernel (kept-E backward GEMMs) */
  MM25_T_SIM_GRAD = 2,     /* sim_grad_tc_kernel */
  MM25_T_GEMM_GROUPED = 3, /* gemm_grouped_kernel / gemm_grouped2_kernel */
  MM25_T_SIM_MOD = 4,      /* sim_mod_tc_kernel (MODULATED fwd / grad, Hopfield backward sweep) */
  MM25_T_SIMT_FWD = 5,     /* sim_fwd_kernel (fp32 FFMA path) */
  MM25_T_SIMT_BWD = 6,     /* sim_bwd_kernel / hopfield_bwd_kernel (fp32 FFMA path) */
  MM25_T_SAMPLER = 7,      /* projnormal_* / sphere_* sampler kernels */
  MM25_T_SIM_RANK = 8,     /* sim_rank_tc_kernel */
  MM25_TIMING_SLOTS = 12
};

/* number of [rows, d] blocks of the concatenated two-term bf16 split operands of the fp32-grade
 * path (mm25_split_bf16x2): hi.hi + hi.lo + lo.hi; the lo x lo product is dropped */
#define MM25_SPLIT_K 3

/* logit modes of the fused similarity kernels */
enum {
  MM25_SIM_PLAIN = 0,    /* z = s * a_i.b_j                       (clipLoss.py:106; loss.py:215) */
  MM25_SIM_MASKDIAG = 1, /* z = s * a_i.b_j, z_ii := -10000 in LSE (cloobLoss.py:26-31)         */
  MM25_SIM_MODULATED = 2 /* z = s * [(a_i.b_j)(1 - u_i.u_j) + delta_ij D_i]  (loss.py:229-242)  */
};

/* how `scale` (a device scalar) is interpreted */
enum {
  MM25_SCALE_MUL = 0, /* multiplicative, used as is (AlignCLIP convention, loss.py:215)          */
  MM25_SCALE_LOG = 1  /* log-space: s = exp(min(l, clamp_max)) (clipLoss.py:35,105)               */
};

const char* mm25_version(void);
/* cudaGetErrorString of the last failing CUDA call seen by this thread (static storage). */
const char* mm25_last_cuda_error(void);
/* number of kernels this library has launched in this process (all threads). */
long long mm25_launch_count(void);
/* measurement hook (bench.py `roofline`): while on, the launchers of the hot kernels bracket each
 * launch with CUDA events on the launch stream (eager launches only; ignored during graph
 * capture).  mm25_debug_timing(on) also clears the record.  Slots: the MM25_T_* enum above. */
void mm25_debug_timing(int on);
/* summed duration (ms) and number of recorded launches of a slot; synchronises on the events. */
int mm25_debug_timing_read(int slot, double* ms_total, long long* launches);
/* test hook: non-zero forces the robust (two-exp) epilogue path of the tcgen05 forward. */
void mm25_debug_force_slow_epilogue(int on);
/* test hook: cap (bytes) of the bf16 gradient panel of the tensor-core backward; 0 = default. */
void mm25_debug_set_g_panel_bytes(size_t bytes);
/* test hook: C[Mg,Ng] fp32 (+)= alpha * op(A) op(B) with the tcgen05 GEMM of the backward;
 * a_mn / b_mn = 1 when the bf16 operand is stored [K, M] / [K, N] instead of [M, K] / [N, K]. */
int mm25_debug_gemm_bf16(const void* A, int a_mn, const void* B, int b_mn, float* C, int64_t Mg,
                         int64_t Ng, int64_t K, float alpha, int accumulate, void* stream);
/* test hook: two GEMMs sharing Ng in one grouped launch, C0[M0,Ng] = op(A0) op(B0) and
 * C1[M1,Ng] = op(A1) op(B1) (the dA / dB pair of the tensor-core backward). */
int mm25_debug_gemm_grouped(const void* A0, int a0_mn, const void* B0, int b0_mn, float* C0,
                            int64_t M0, int64_t K0, const void* A1, int a1_mn, const void* B1,
                            int b1_mn, float* C1, int64_t M1, int64_t K1, int64_t Ng,
                            void* stream);
/* test hook: the kept-E backward GEMMs with the E -> G' transform fused into the operand load
 * (csrc/gemm_te.cu; replaces the autograd of `logits = scale * x @ y.T` + cross_entropy,
 * losses/clipLoss.py:106-119).  E [batch * Me, Ne] bf16; C0 [batch * Me, Ng] = T(E) B0 with
 * B0 [batch * Ne, Ng]; C1 [batch * Ne, Ng] = T(E)^T B1 with B1 [batch * Me, Ng] (bf16 row-major,
 * fp32 out; either C may be NULL).  For entry z, when *e_flag != 0:
 * T(E)_rc = E_rc (rowf_r + colf_c) - consts[2 z], except T(E)_rc = gdiag_r (fp32 [batch * Me],
 * may be NULL) where c == r + diag_offset; identity when *e_flag == 0.
 * Me, Ne, Ng multiples of 256. */
int mm25_debug_gemm_te(const void* E, int64_t Me, int64_t Ne, const float* rowf,
                       const float* colf, const float* consts, const float* gdiag,
                       const int* e_flag, int64_t diag_offset, const void* B0, float* C0,
                       const void* B1, float* C1, int64_t Ng, int64_t batch, void* stream);
/* measurement hook: cycle counters of one A-producer thread of gemm_te_kernel (8 values:
 * load wait, stage wait, stage fill, tiles, ...); enable != 0 allocates, 0 frees; out8 may be NULL. */
int mm25_debug_te_stats(long long* out8, int enable);
/* test hook: the scatter-mode GEMM of the fused reduce-scatter (mm25_sim_bwd_rs) with the owner
 * buffers as a HOST array of `nowners` device pointers to [rows_per_rank, Ng] fp32: rows
 * [q R, (q+1) R) of op(A) op(B) are ADDED into owners[q].  On one GPU the owners are local
 * buffers, so the epilogue variants selected by the environment variable MM25_RS_MODE
 * (0 = per-lane red.global, 1 = 128-byte TMA reductions, 2 = staged row-contiguous red.global)
 * can be checked without NVLink. */
int mm25_debug_gemm_scatter(const void* A, int a_mn, const void* B, int b_mn,
                            float* const* owners, int nowners, int64_t rows_per_rank,
                            int self_rank, int64_t Mg, int64_t Ng, int64_t K, void* stream);
/* host-only test hook (no device work): the static balanced schedule of the grouped GEMM for
 * mb0 blocks of nk0 k-blocks (problem 0, walked round-robin over the groups) and mb1 blocks of
 * nk1 (problem 1, contiguous runs sized by water-filling) on `ngroups` CTA groups.
 * out[5] = {r, nbA, nbB, nsA, nsB}: groups [0, r) take nbA + nsA blocks, the others nbB + nsB. */
int mm25_debug_gemm_schedule(int64_t mb0, int64_t nk0, int64_t mb1, int64_t nk1, int ngroups,
                             int* out);
/* 1 if the current device is compute capability 10.x, else 0. */
int mm25_device_ok(void);

/* ------------------------------------------------------------------------------------------
 * Row L2-normalisation prologue.  Replaces F.normalize(x, dim=-1) (clipLoss.py:28-29,95-96),
 * semantic_features / norm (loss.py:229).
 *   x      [B, M, d] dtype       (B = batch of independent problems, stride M*d)
 *   xhat   [B, M, d] out_dtype   x / max(||x||, eps)   (or a plain cast when normalize == 0)
 *   inv_norm [B, M]  fp32        1 / max(||x||, eps)   (1.0 when normalize == 0)
 *   colsum           fp32        sum_i xhat_i in its first [B, d] floats; the buffer must hold
 *                                mm25_rownorm_colsum_floats(B, M, d) floats (result + partials
 *                                of the deterministic two-stage reduction); may be NULL
 * ------------------------------------------------------------------------------------------ */
int64_t mm25_rownorm_colsum_floats(int64_t B, int64_t M, int64_t d);
int mm25_rownorm_fwd(const void* x, int dtype, void* xhat, int out_dtype, float* inv_norm,
                     float* colsum, int64_t B, int64_t M, int64_t d, int normalize, float eps,
                     void* stream);

/* Allocation-free variant of mm25_rownorm_fwd with column sums.  NO COUNTERPART IN THE REFERENCE:
 * added for callers that cannot allocate (CUDA graph
 * capture, tight serving loops, harnesses that forbid cudaMalloc).  Same kernels and results as
 * mm25_rownorm_fwd, but the partials of the two-stage column sum live in a file-scope device
 * array of the library (MM25_NOALLOC_MAX_FLOATS floats = 2.89 MiB, held by every GPU that loads
 * the library until exit) instead of behind `colsum`: the caller passes inputs and outputs only,
 * `colsum` is exactly [B, d] floats, and the call performs no cudaMalloc / cudaFree.
 *   needs  d % 128 == 0, d <= 2048, M >= 64                       (else MM25_ERR_SHAPE)
 *   and    mm25_rownorm_colsum_noalloc_floats(B, M, d) = B * d * min(ceil(M / 8), 740)
 *          <= MM25_NOALLOC_MAX_FLOATS                             (else MM25_ERR_WORKSPACE;
 *          e.g. B = 1: any M at d <= 1024, M <= 2960 at d = 2048)
 * All calls on a device share the one array: it keeps the last call's partials (nothing needs
 * to start at zero — stage 1 overwrites all it reads back), and two calls that overlap on the GPU
 * (different streams, or graphs replayed concurrently) RACE on it — serialise them. */
#define MM25_NOALLOC_MAX_FLOATS (740 * 1024)
int64_t mm25_rownorm_colsum_noalloc_floats(int64_t B, int64_t M, int64_t d);
int mm25_rownorm_colsum_noalloc(const void* x, int dtype, void* xhat, int out_dtype,
                                float* inv_norm, float* colsum, int64_t B, int64_t M, int64_t d,
                                int normalize, float eps, void* stream);

/* Backward of the prologue: dx = inv_norm * (g - xhat (xhat.g)) (or a cast when !normalize).
 *   g [B,M,d] fp32 (gradient w.r.t. xhat), xhat [B,M,d] xhat_dtype, dx [B,M,d] out_dtype.
 *   gscale: optional device scalar multiplied into dx (upstream grad); NULL = 1.
 *   accumulate: if non-zero dx += (used when a tensor feeds several terms). */
int mm25_rownorm_bwd(const float* g, const void* xhat, int xhat_dtype, const float* inv_norm,
                     void* dx, int out_dtype, int64_t B, int64_t M, int64_t d, int normalize,
                     const float* gscale, int accumulate, void* stream);

/* ------------------------------------------------------------------------------------------
 * Fused similarity + online row/column log-sum-exp.  Never materialises the M x N logits.
 * Replaces clipLoss.py:106-117 (logits + 2x cross_entropy), cloobLoss.py:26-31,
 * AlignCLIP loss.py:209-216,229-269.
 *
 *   a [B,M,d], b [B,N,d]  (dtype)         — already normalised / cast by mm25_rownorm_fwd
 *   u [B,M,d2], v [B,N,d2] fp32 or NULL    — MODULATED only: normalised semantic features
 *   diag_add [B,M] fp32 or NULL            — MODULATED only: D_i added on the diagonal
 *   scale: device scalar, see MM25_SCALE_*
 *   diag_offset: column of the positive of local row 0 (rank * M when rows are sharded,
 *                cf. loss.py:94-95)
 * Outputs (all fp32):
 *   row_lse [B,M]   LSE_j z_ij
 *   col_max, col_sum [B,N]  partial column statistics over the M local rows, such that
 *                   LSE_i z_ij = col_max + log(col_sum); NULL to skip (row-only losses)
 *   diag [B,M]      z_{i, i+diag_offset} (un-masked value)
 * Implementation selection: bf16 inputs with d % 64 == 0, M % 128 == 0, N % 128 == 0 in
 * PLAIN mode run the tcgen05/TMEM/TMA kernel; everything else the fp32 SIMT kernel.
 * `impl`: 0 = auto, 1 = force SIMT, 2 = force tcgen05 (error if not eligible).
 * `dot_bound`: a caller-guaranteed bound on |a_i . b_j| (1.01 for rows normalised by
 * mm25_rownorm_fwd), or <= 0 when unknown.  With a bound and 2 dot_bound |s| log2(e) <= 96 the
 * tcgen05 kernels use a static softmax shift (one exp per logit, no running maximum); results
 * are the same statistics either way.
 * `e_keep`, `e_flag` (both or neither; NULL to skip): a [B*M, N] bf16 buffer and a device int.
 * When the tcgen05 static-shift path runs it also stores E_ij = 2^(z_ij log2e - dot_bound |s|
 * log2e) there and sets *e_flag = 1 (0 on any other path, buffer then untouched).  Handing both
 * to mm25_sim_bwd with the same scale and dot_bound replaces the S-tile recompute of the
 * backward by an elementwise pass over E (2 B read + 2 B written per logit); the buffer is
 * CONSUMED (overwritten by G').  This is the activation autograd would have saved
 * (clipLoss.py:106 keeps the fp32 logits); at B = 32768 it is 2 GiB of the 180 GB.
 * ------------------------------------------------------------------------------------------ */
size_t mm25_sim_fwd_workspace_bytes(int64_t B, int64_t M, int64_t N, int64_t d, int dtype,
                                    int impl);
int mm25_sim_fwd(const void* a, const void* b, int dtype, const float* u, const float* v,
                 const float* diag_add, int64_t B, int64_t M, int64_t N, int64_t d, int64_t d2,
                 const float* scale, int scale_kind, float clamp_max, int mode,
                 int64_t diag_offset, float* row_lse, float* col_max, float* col_sum,
                 float* diag, void* workspace, size_t workspace_bytes, int impl, float dot_bound,
                 void* e_keep, int* e_flag, void* stream);

/* ------------------------------------------------------------------------------------------
 * Fused backward with tile recompute (autograd of the block above).
 *   G_ij = w_row e^{z_ij - row_lse_i} + w_col e^{z_ij - col_lse_j} - w_diag delta_ij - w_all
 *   da  [B,M,d] fp32 = s * sum_j G_ij (dz_ij/ds_ij-factor) b_j        (row-stationary pass)
 *   db  [B,N,d] fp32 = s * sum_i G_ij (...) a_i                        (NULL: skip)
 *   dscale_part: fp32 partial sums of sum_ij G_ij z_ij / s, one per CTA of the row pass,
 *                length mm25_sim_bwd_nparts(); reduce with mm25_sum_parts (NULL: skip)
 *   ddiag [B,M] fp32 = s * G_ii (MODULATED: gradient w.r.t. D_i), NULL to skip
 * col_lse may be NULL iff w_col == 0.  Weights are host scalars; `gout` is an optional device
 * array [B] (upstream gradient per batch entry) folded into all four.
 * `e_keep`, `e_flag`: the pair mm25_sim_fwd filled (NULL, NULL = always recompute).  With them
 * the workspace shrinks to the column factors (`with_e` = 1 in the size query) and, when the
 * flag is up, d loss / d scale is taken as (a . da) / s instead of from the logits — so da or
 * db must be requested whenever dscale_part is (else the call recomputes as if NULL).
 * ------------------------------------------------------------------------------------------ */
size_t mm25_sim_bwd_workspace_bytes(int64_t B, int64_t M, int64_t N, int64_t d, int dtype,
                                    int impl, int with_e);
int64_t mm25_sim_bwd_nparts(int64_t B, int64_t M, int64_t N, int64_t d, int dtype, int impl);
/* MODULATED mode (AlignCLIP semantic term, loss.py:229-242) also has a tcgen05 path: bf16 a/b,
 * B = 1, M, N % 128 == 0, d, d2 % 64 == 0, row statistics only (col_max = NULL, w_col = 0);
 * backward additionally d % 256 == 0, d <= 1024, N % 256 == 0.  Each 128 x 128 tile keeps two
 * TMEM accumulators (a.b over d, u.v over d2).  It is taken when `impl` != 1 and the caller's
 * workspace is at least the size below (0 = not eligible: the SIMT kernels run); dscale_part
 * then needs max(mm25_sim_bwd_nparts, mm25_sim_mod_bwd_nparts) entries. */
/* Kernel family taken by the last mm25_sim_fwd / mm25_sim_bwd call of this process:
 * 0 = fp32 SIMT, 1 = tcgen05 PLAIN, 2 = tcgen05 MODULATED, -1 = none yet (test hook). */
int mm25_debug_last_sim_path(void);
size_t mm25_sim_mod_workspace_bytes(int64_t B, int64_t M, int64_t N, int64_t d, int64_t d2,
                                    int dtype, int bwd);
int64_t mm25_sim_mod_bwd_nparts(int64_t B, int64_t M, int64_t N, int64_t d, int64_t d2,
                                int dtype);
int mm25_sim_bwd(const void* a, const void* b, int dtype, const float* u, const float* v,
                 const float* diag_add, int64_t B, int64_t M, int64_t N, int64_t d, int64_t d2,
                 const float* scale, int scale_kind, float clamp_max, int mode,
                 int64_t diag_offset, const float* row_lse, const float* col_lse, float w_row,
                 float w_col, float w_diag, float w_all, const float* gout, float* da,
                 float* db, float* dscale_part, float* ddiag, void* workspace,
                 size_t workspace_bytes, int impl, float dot_bound, void* e_keep,
                 const int* e_flag, void* stream);

/* Fused backward + reduce-scatter for the row-sharded loss (tcgen05 path, PLAIN mode, B = 1):
 * as mm25_sim_bwd, but instead of writing the partial db [N, d] of this rank, rows
 * [q N/P, (q+1) N/P) of it are ADDED (fp32 red.global over NVLink) into db_peers[q], the
 * peer-mapped [N/P, d] gradient buffer of rank q (symmetric memory; HOST array of P device
 * pointers).  The buffers must be zeroed and a cross-rank barrier passed before the call, and a
 * barrier must follow it before they are read.  Replaces the reduce-scatter that autograd of
 * `gather_features(gather_with_grad=True)` performs (AlignCLIP/align_clip/loss.py:49-51).
 * The summation order across ranks is not fixed (fp32 atomics). */
int mm25_sim_bwd_rs(const void* a, const void* b, int dtype, int64_t M, int64_t N, int64_t d,
                    const float* scale, int scale_kind, float clamp_max, int64_t diag_offset,
                    const float* row_lse, const float* col_lse, float w_row, float w_col,
                    float w_diag, float w_all, const float* gout, float* da,
                    float* const* db_peers, int64_t npeers, float* dscale_part, void* workspace,
                    size_t workspace_bytes, float dot_bound, void* e_keep, const int* e_flag,
                    void* stream);

/* Hopfield-retrieval backward on the tensor cores (cloobLoss.py:43-47; O = softmax(beta Q K^T) K):
 * q [M,d], k [N,d] bf16; d_o [M,d] fp32 = dL/dO; lse [M] = LSE_j beta q_i.k_j (mm25_sim_fwd row
 * LSE); delta [M] = d_o_i . O_i (mm25_rowdot); beta: device scalar.  Writes dq [M,d] and/or
 * dk [N,d] fp32 (NULL to skip).  One sweep with two TMEM accumulators (Q K^T and dO K^T) emits
 * the bf16 panels dZ = beta P (dP - delta) and P; dQ = dZ K, dK = dZ^T Q + P^T dO follow on the
 * grouped GEMM.  Eligible: M % 128 == 0, N % 256 == 0, d % 256 == 0, d <= 1024 (workspace query
 * returns 0 otherwise; use mm25_hopfield_bwd_pass then). */
size_t mm25_hopfield_bwd_tc_workspace_bytes(int64_t M, int64_t N, int64_t d);
int mm25_hopfield_bwd_tc(const void* q, const void* k, const float* d_o, int64_t M, int64_t N,
                         int64_t d, const float* beta, const float* lse, const float* delta,
                         float* dq, float* dk, void* workspace, size_t workspace_bytes,
                         void* stream);

/* Deterministic fixed-order sum of n fp32 values into out[0] (one CTA). out = alpha * sum. */
int mm25_sum_parts(const float* parts, int64_t n, float alpha, float* out, void* stream);

/* Merge P partial column statistics (max,sum)[P,N] in fixed order into col_lse[N]
 * (single GPU: P = 1).  Used after the cross-rank all-gather of column partials. */
int mm25_col_merge(const float* col_max, const float* col_sum, int64_t P, int64_t N,
                   float* col_lse, void* stream);
/* The same with a leading dimension: the partials are rows of a gathered [P, ld] payload. */
int mm25_col_merge_ld(const float* col_max, const float* col_sum, int64_t P, int64_t N,
                      int64_t ld, float* col_lse, void* stream);
/* One-shot all-gather over NVLink peer memory: copies `bytes` from `src` to byte offset
 * `dst_offset_bytes` of each of the `npeers` peer-mapped buffers (HOST array of device pointers,
 * symmetric memory, the caller's own buffer included).  16-byte aligned sizes / offsets take the
 * vector kernel; 4-byte aligned payloads up to 1 MiB (scalars) a word kernel.  A
 * cross-rank barrier must follow before any rank reads its gathered buffer.  Replaces
 * `torch.distributed.nn.all_gather` of the reference's gather_features (loss.py:49-51). */
int mm25_peer_bcast(const void* src, int64_t bytes, void* const* peer_ptrs, int64_t npeers,
                    int64_t dst_offset_bytes, void* stream);
/* The same gather through the NVSwitch multicast mapping of the symmetric buffer: one
 * `multimem.st` per 16 bytes lands in every rank's copy (the sender's included), so each rank
 * sends its shard once instead of `npeers` times.  `mc_ptr` = multicast address of the buffer
 * (torch.distributed._symmetric_memory handle `multicast_ptr`); bytes, offset and pointers
 * 16-byte aligned.  A cross-rank barrier must follow, as for mm25_peer_bcast. */
int mm25_peer_bcast_mc(const void* src, int64_t bytes, void* mc_ptr, int64_t dst_offset_bytes,
                       void* stream);
/* Label-smoothing term of the row-sharded loss (clipLoss.py:111-117 in closed form):
 * loss[0] -= coef * s * (sum_p sx_p) . (sum_p sy_p), sx_p / sy_p = the [d] column sums of
 * xhat / yhat of rank p at column offsets ox / oy of a gathered [P, ld] payload.  One CTA. */
int mm25_smooth_term(const float* payload, int64_t P, int64_t ld, int64_t ox, int64_t oy,
                     int64_t d, const float* scale, int scale_kind, float clamp_max, float coef,
                     float* loss, void* stream);
/* The scalar tail of the row-sharded forward in one launch: loss[0] = sum_p payload[p, orow]
 * (each rank's row part of the loss, rank order) + (0.5 / N) sum_j col_lse[j] (fp64, fixed
 * order: identical bits on every rank) - coef * s * (sum_p sx_p).(sum_p sy_p) (the smoothing term
 * of mm25_smooth_term; d = 0 skips it).  Replaces a scalar all-reduce: after the column
 * statistics are merged every rank holds all N column LSEs (dist.ShardedClipLoss). */
int mm25_sharded_loss(const float* payload, int64_t P, int64_t ld, int64_t orow,
                      const float* col_lse, int64_t N, int64_t ox, int64_t oy, int64_t d,
                      const float* scale, int scale_kind, float clamp_max, float coef,
                      float* loss, void* stream);
/* dL/d(raw scale parameter) from dL/ds (device scalars): MM25_SCALE_LOG: dscale * exp(l) when
 * l <= clamp_max else 0 (clipLoss.py:106 `clamp(max=4.6052).exp()`); MM25_SCALE_MUL: dscale. */
int mm25_scale_param_grad(const float* dscale, const float* scale, int scale_kind,
                          float clamp_max, float* out, void* stream);

/* ------------------------------------------------------------------------------------------
 * CLIP loss value from statistics (clipLoss.py:116-119 in closed form):
 *   L_b = (sum_i r_i + sum_j c_j) / (2 N) - (1-eps)/N sum_i z_ii - eps/N^2 * s * (sx . sy)
 * row_lse,diag [B,M_total]; col_lse [B,N]; sx, sy [B,d] column sums of xhat / yhat (may be
 * NULL when eps == 0, or to leave
 * the smoothing dot term to the caller as the sharded loss does).  loss [B].  `n_norm` = global batch N used in the means.
 * ------------------------------------------------------------------------------------------ */
int mm25_clip_loss_finalize(const float* row_lse, const float* diag, const float* col_lse,
                            const float* sx, const float* sy, int64_t B, int64_t M, int64_t N,
                            int64_t d, int64_t n_norm, const float* scale, int scale_kind,
                            float clamp_max, float eps_ls, float* loss, void* stream);

/* ------------------------------------------------------------------------------------------
 * CLOOB Hopfield retrieval, backward (cloobLoss.py:43-47; forward = mm25_sim_fwd row LSE +
 * the row pass of mm25_sim_bwd with w_row = 1 + mm25_rownorm_fwd).  All fp32.
 *   out[rows, d] = sum_cols G1 bm1[cols] (+ G2 bm2[cols]),
 *   p = exp(beta a.b - lse[q]),  G1 = beta p (a2.b2 - D[q]),  G2 = p,
 *   q = row (stat_by_col = 0) or column (stat_by_col = 1) of the [M, N] tile grid.
 * dQ: a=Q b=K a2=dO b2=K bm1=K bm2=NULL stat_by_col=0;  dK: a=K b=Q a2=K b2=dO bm1=Q bm2=dO, 1.
 * mm25_rowdot: out[i] = x_i . y_i  (D_q = dO_q . O_q).
 * ------------------------------------------------------------------------------------------ */
int mm25_rowdot(const float* x, const float* y, float* out, int64_t rows, int64_t d,
                void* stream);
int mm25_hopfield_bwd_pass(const float* a, const float* b, const float* a2, const float* b2,
                           const float* bm1, const float* bm2, int64_t M, int64_t N, int64_t d,
                           float beta, const float* lse, const float* D, int stat_by_col,
                           float* out, void* stream);

/* ------------------------------------------------------------------------------------------
 * vCLIP sampling path (losses/vclipLoss.py:98-121, losses/distributions/*.py, power_spherical).
 * All tensors fp32.  Sample layout is sample-major [S, B, d] as the reference returns
 * (VonMisesFisher.py:238-239).  Noise: pass the draws the reference would make (parity mode)
 * or NULL to generate them in-kernel with Philox4x32-7 keyed by `seed` (the backward
 * regenerates the same noise from the same seed).
 * ------------------------------------------------------------------------------------------ */
/* ProjectedNormal.rsample (ProjectedNormal.py:41-49): out = (mu + sigma*eps)/||.||;
 * eps_in [S,B,d] or NULL; inv_norm [S,B] (saved for the backward). */
int mm25_sample_projnormal_fwd(const float* mu, const float* sigma, const float* eps_in,
                               uint64_t seed, int64_t S, int64_t B, int64_t d, float* out,
                               float* inv_norm, void* stream);
/* its backward: dmu, dsigma [B,d] (sum over the S samples). */
int mm25_sample_projnormal_bwd(const float* gout, const float* y, const float* inv_norm,
                               const float* eps_in, uint64_t seed, int64_t S, int64_t B,
                               int64_t d, float* dmu, float* dsigma, void* stream);
/* x = [t, sqrt(1-t^2) v/nv], y = Householder(e1 -> loc) x.  t [S,B], loc [B,d], v_in [S,B,d-1]
 * or NULL.  nv = max(||v||, v_clamp) + v_add; u = (e1-loc)/(max(||e1-loc||, h_clamp) + h_add);
 * 1-t^2 clamped below by omt2_min when omt2_min >= 0.
 * vMF (VonMisesFisher.py:181-183,224-236): 1e-10, 0, 1e-10, 0, -1.
 * PowerSpherical (power_spherical _TTransform/_HouseholderRotationTransform): 0, 1e-7, 0, 1e-7, 1e-7. */
int mm25_sphere_sample_fwd(const float* t, const float* loc, const float* v_in, uint64_t seed,
                           int64_t S, int64_t B, int64_t d, float v_clamp, float v_add,
                           float h_clamp, float h_add, float omt2_min, float* out, void* stream);
/* backward: dloc [B,d] (sum over samples) and dt [S,B] from gout, the samples y, t and loc. */
int mm25_sphere_sample_bwd(const float* gout, const float* y, const float* t, const float* loc,
                           int64_t S, int64_t B, int64_t d, float v_clamp, float v_add,
                           float h_clamp, float h_add, float omt2_min, float* dloc, float* dt,
                           void* stream);
/* Wood's rejection sampler of the vMF marginal (VonMisesFisher.py:184-222): w [S,B].
 * prop_eps/prop_u [B,P]: injected Beta((d-1)/2,(d-1)/2) / Uniform proposals scanned in order
 * (count[b] = accepted, missing entries NaN); NULL: Philox. */
int mm25_vmf_sample_w(const float* kappa, int64_t d, int64_t S, int64_t B, uint64_t seed,
                      const float* prop_eps, const float* prop_u, int64_t P, float* w_out,
                      int* count, void* stream);
/* z[s,b] ~ Beta(alpha_b, beta_b) (PowerSpherical marginal, t = 2z-1). */
int mm25_beta_sample(const float* alpha, const float* beta, int64_t S, int64_t B, uint64_t seed,
                     float* z, void* stream);
/* implicit-reparameterisation gradient dz/dalpha of Beta draws (what the backward of
 * torch.distributions.Beta.rsample evaluates: (1-z) * dirichlet_grad_one(z, alpha, alpha+beta)). */
int mm25_beta_grad(const float* z, const float* alpha, const float* beta, int64_t S, int64_t B,
                   float* dz_dalpha, void* stream);
/* log I_nu(kappa) and log I_{nu+1}(kappa) in fp64 (VonMisesFisher.py:19-62 mpmath round trip);
 * either output may be NULL. */
int mm25_log_bessel(const double* kappa, int64_t n, double nu, double* log_i_nu,
                    double* log_i_nu1, void* stream);

/* ------------------------------------------------------------------------------------------
 * Top-k retrieval indices (metrics/topk.py:43-48 `similarities.topk(k).indices`).
 * scores [Q, N] fp32 with row stride ld; out_idx [Q, k] int64 and (optional) out_val [Q, k], sorted
 * by descending score, ties lowest index first, NaN largest (torch order).  k <= 2048, k <= N.
 * ------------------------------------------------------------------------------------------ */
int mm25_topk_rows(const float* scores, int64_t Q, int64_t N, int64_t ld, int64_t k,
                   int64_t* out_idx, float* out_val, void* stream);
/* The same result (bit for bit) through the sliced path: a row is cut into slices that one CTA
 * each loads ONCE into shared memory (radix select, gather and sort never return to HBM), so a
 * small-Q call — the reference's own shape, one query against all texts, metrics/topk.py:43-48
 * — uses every SM; a merge kernel sorts the slices' candidates.  `workspace` holds them:
 * mm25_topk_workspace_bytes(Q, N, k) bytes (0 when one slice per row suffices or when the shape
 * falls back to mm25_topk_rows: k * slices > 8192, Q > 65535). */
size_t mm25_topk_workspace_bytes(int64_t Q, int64_t N, int64_t k);
int mm25_topk_rows_ws(const float* scores, int64_t Q, int64_t N, int64_t ld, int64_t k,
                      int64_t* out_idx, float* out_val, void* workspace, size_t workspace_bytes,
                      void* stream);

/* ------------------------------------------------------------------------------------------
 * Retrieval ranks without the sort (AlignCLIP/main/train.py:338-356 `get_clip_metrics`:
 * `ranking = argsort(logit, descending=True); preds = where(ranking == ground_truth)[1]`, once
 * per direction on the CPU copy of the N x N logits; main/retrieval.py:69-71,88-90 the same
 * argsort for recall@k).
 * scores [Q, N] fp32 slab (row stride ld) holding global rows row0 .. row0+Q-1.  One pass reads
 * each score once and ACCUMULATES (atomic integer adds; the caller zeroes the outputs)
 *   row_rank[q] += #{j : s[q,j] sorts ahead of the row target (value row_tval[q], column
 *                     row_tcol[q])},
 *   col_rank[j] += #{q : s[q,j] sorts ahead of the column target (value col_tval[j], global row
 *                     col_trow[j])},
 * where "ahead" is the stable descending order: larger value, or equal value and smaller
 * position (NaN largest, as torch sorts).  Either direction may be skipped by passing NULL for
 * its rank pointer.  Integer result: bit-exact and independent of the slab split.
 * ------------------------------------------------------------------------------------------ */
int mm25_rank_both(const float* scores, int64_t Q, int64_t N, int64_t ld, int64_t row0,
                   const float* row_tval, const int64_t* row_tcol, const float* col_tval,
                   const int64_t* col_trow, int* row_rank, int* col_rank, void* stream);

/* ------------------------------------------------------------------------------------------
 * fp32-grade similarity kernels on the tensor cores (losses/clipLoss.py:106 computes the logits
 * in fp32: `.float()`; contract rel 1e-5 loss / 1e-4 gradients).  fp32 rows are split into two
 * bf16 terms x = hi + lo (16 significant bits) and laid out so that ONE bf16 sweep over
 * K = 3 d reproduces hi_a.hi_b + hi_a.lo_b + lo_a.hi_b (the lo x lo product, 2^-18 of the dot
 * product, is dropped):
 *   mm25_split_bf16x2(x [rows, d] fp32, role) -> out4 [rows, 3 d] bf16 = [hi hi lo] (role 0,
 *   row operand) or [hi lo hi] (role 1, column operand), plus hi, lo [rows, d] bf16.
 * Forward: mm25_sim_fwd(out4_a, out4_b, MM25_BF16, ..., d = 3 d, MM25_SIM_PLAIN).
 * Backward: mm25_sim_bwd_split — the gradient panel is split the same way (the recompute sweep
 * runs twice, the second pass emits g - bf16(g)) and dA = G_hi Y_hi + G_hi Y_lo + G_lo Y_hi,
 * accumulated in fp32 by three grouped-GEMM launches; likewise dB.  da [B, M, d],
 * db [B, N, d] fp32 (either may be NULL); dscale_part: mm25_sim_bwd_split_nparts(B, M, N)
 * floats (NULL: skip); workspace: mm25_sim_bwd_split_workspace_bytes(M, N); mode = MM25_SIM_PLAIN
 * or MM25_SIM_MASKDIAG (InfoLOOB, cloobLoss.py:26-31).
 * mm25_sim_split_eligible: 1 when the shape is covered (d % 256 == 0, d <= 1024, M % 128 == 0,
 * N % 256 == 0). */
/* MM25_SPLIT_K (defined with the enums above) as a call, for bindings that cannot read macros */
int mm25_split_width(void);
int mm25_split_bf16x2(const float* x, int64_t rows, int64_t d, int role, void* out4, void* hi,
                      void* lo, void* stream);
int mm25_sim_split_eligible(int64_t B, int64_t M, int64_t N, int64_t d);
size_t mm25_sim_bwd_split_workspace_bytes(int64_t M, int64_t N);
int64_t mm25_sim_bwd_split_nparts(int64_t B, int64_t M, int64_t N);
int mm25_sim_bwd_split(const void* a4, const void* b4, const void* a_hi, const void* a_lo,
                       const void* b_hi, const void* b_lo, int64_t B, int64_t M, int64_t N,
                       int64_t d, const float* scale, int scale_kind, float clamp_max,
                       int64_t diag_offset, const float* row_lse, const float* col_lse,
                       float w_row, float w_col, float w_diag, float w_all, const float* gout,
                       float* da, float* db, float* dscale_part, void* workspace,
                       size_t workspace_bytes, float dot_bound, int mode, void* stream);
/* fp32-grade Hopfield backward (autograd of `hopfield`, losses/cloobLoss.py:43-47, at the fp32
 * contract on the tensor cores): the dual-accumulator sweep of mm25_hopfield_bwd_tc on the split
 * operands q4, k4 (mm25_split_bf16x2 roles 0 / 1) with the panels dZ and P split as well; d_o is
 * fp32 [M, d] and is split inside.  dq [M, d], dk [N, d] fp32 (either may be NULL). */
int mm25_hopfield_split_eligible(int64_t M, int64_t N, int64_t d);
size_t mm25_hopfield_bwd_split_workspace_bytes(int64_t M, int64_t N, int64_t d);
int mm25_hopfield_bwd_split(const void* q4, const void* k4, const void* q_hi, const void* q_lo,
                            const void* k_hi, const void* k_lo, const float* d_o, int64_t M,
                            int64_t N, int64_t d, const float* beta, const float* lse,
                            const float* delta, float* dq, float* dk, void* workspace,
                            size_t workspace_bytes, void* stream);

/* Retrieval ranks fused into the similarity GEMM: the scores a_q . b_j (bf16 rows, fp32
 * accumulation on the tensor cores) are counted against the row's target as they leave TMEM and
 * are never written (AlignCLIP/main/train.py:338-356 forms `logits_per_image` [N, N], copies it
 * to the CPU and argsorts it per direction; main/retrieval.py:62-71; metrics/topk.py:64-82).
 * a [Q, d], b [N, d] bf16 row-major (d % 64 == 0, 64 <= d <= 1024; any Q, N); ADDS into
 *   rank[q] += #{ j != tcol[q] : a_q.b_j > tval[q], or == tval[q] and j < tcol[q] }
 * (the stable descending order of mm25_rank_both; the caller zeroes `rank`).  tval[q] is the
 * caller's value of the target score a_q . b_tcol[q] (its own column is excluded by position).
 * The column direction of a paired batch is a second call with a and b swapped.  A positive
 * logit scale does not change the order, so the raw dot products are ranked. */
int mm25_sim_rank(const void* a, const void* b, int dtype, int64_t Q, int64_t N, int64_t d,
                  const float* tval, const int64_t* tcol, int* rank, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* MM25_H_ */
