/* dgraph_b200 -- C ABI of the sm_100a kernel library (dgraph_b200/lib/libdgraph_b200.so).
 *
 * Every launcher takes raw device pointers and the CUDA stream to launch on (cudaStream_t passed as void*), returns 0 on
 * success or a cudaError_t value (see dgb_error_string), never synchronises the device and -- unless stated otherwise --
 * allocates nothing.  Row strides ("ld") are in ELEMENTS unless the name says bytes.  bf16 matrices need 16-byte aligned
 * base pointers and strides that are multiples of 8 elements.
 *
 * Citations name the piece of the reference (LBANN/DGraph, paths relative to its repository root) whose role the entry
 * point takes over.
 */
#ifndef DGRAPH_B200_H_
#define DGRAPH_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- runtime ------------------------------------------------------------------------------------------------------ */
const char* dgb_error_string(int code);
int dgb_version(void);
int dgb_sm_count(void);

/* ---- segment aggregation (csrc/segment_ops.cu) ----------------------------------------------------------------------
 * Reference: scatter_add of GraphConvLayer (experiments/OGB/GCN.py:57-65); K6/K7 masked scatter/gather kernels
 * (DGraph/distributed/csrc/local_data_kernels.cuh:301-406).  CSR/CSC walk, one sub-warp per output row, no atomics. */
/* Pull variant (peer_base != null): idx[k] < 0 names a row of another rank, ~idx[k] = (owner << peer_shift) | row; the kernel reads it
 * from peer_base[owner] + row * ld_peer over NVLink in the same batch of loads as the local rows (reference: the in-kernel remote get
 * of DGraph/distributed/csrc/nvshmem_comm_kernels.cuh:255-299).  out2 (optional): rows >= out_split are written to out2[r - out_split]. */
int dgb_seg_sum(const int* ptr, const int* idx, int64_t n_rows, int H, const void* src, int64_t ld_src, const float* slot_w,
                int mean, void* out, int64_t ld_out, const void* const* peer_base, int world, int peer_shift, int64_t ld_peer,
                void* out2, int64_t ld_out2, int64_t out_split, void* stream);
/* out[r] = sum_k relu(P[r] + Q[idx[k]]); optional per-(row,col) active-edge counts and per-edge ReLU bit masks */
int dgb_seg_add_relu_fwd(const int* ptr, const int* idx, int64_t n_rows, int H, const void* P, int64_t ld_p, const void* Q,
                         int64_t ld_q, void* out, int64_t ld_out, unsigned short* cnt, int64_t ld_cnt, unsigned char* mask,
                         void* stream);

/* Extended form: mask row of edge slot k at mask_slot_map[k] (null = k); ptr_end (null = ptr + 1) restricts every row to
 * slots [ptr[r], ptr_end[r]); accumulate = 1 adds to the out / cnt rows of an earlier launch.  With peer_base the Q rows of remote
 * neighbours (idx < 0, encoded as in dgb_seg_sum) are pulled from their owners inside the kernel: halo exchange + gather + add + ReLU +
 * segment sum of a distributed GraphConv layer in ONE launch (reference op chain P1 + P2: haloExchange.py:137 -> GCN.py:43-65). */
int dgb_seg_add_relu_fwd_ex(const int* ptr, const int* ptr_end, const int* idx, int64_t n_rows, int H, const void* P,
                            int64_t ld_p, const void* Q, int64_t ld_q, void* out, int64_t ld_out, unsigned short* cnt,
                            int64_t ld_cnt, unsigned char* mask, const int* mask_slot_map, int accumulate,
                            const void* const* peer_base, int world, int peer_shift, int64_t ld_peer, void* stream);
int dgb_seg_add_relu_bwd(const int* colptr, const int* row, int64_t n_aug, int H, const void* dOut, int64_t ld_g,
                         const void* P, int64_t ld_p, const void* Q, int64_t ld_q, void* dQ, int64_t ld_dq, void* stream);
int dgb_seg_masked_sum(const int* colptr, const int* colptr_end, const int* row, const int* slot_map, const unsigned char* mask, int64_t n_aug,
                       int H, const void* dOut, int64_t ld_g, void* dQ, int64_t ld_dq, const unsigned short* cnt,
                       int64_t ld_cnt, void* dP, int64_t ld_dp, int64_t dp_rows, void* dQ2, int64_t ld_dq2, int64_t dq_split,
                       int dq2_skip_empty, void* stream);
int dgb_mul_cnt(const void* g, int64_t ld_g, const unsigned short* cnt, int64_t ld_c, void* out, int64_t ld_o, int64_t rows,
                int H, void* stream);

/* ---- distributed BatchNorm (csrc/bn_ops.cu) ----------------------------------------------------------------------------
 * Reference: DistributedBatchNorm1D (experiments/OGB-LSC/distributed_layers.py:22-74).  Statistics come from dgb_colsum_bf16
 * (with_sq) + one all-reduce; these kernels do the rest without an fp32 copy of the activations.  act: 0 none, 1 ReLU (its
 * backward masks dy with the saved output y). */
int dgb_bn_apply_bf16(const void* x, int64_t ld_x, int64_t M, int N, const float* mean, const float* rstd, const float* gamma,
                      const float* beta, int act, void* y, int64_t ld_y, void* stream);
int dgb_bn_finalize_stats(const float* buf, int N, float eps, float* mean, float* var, float* rstd, void* stream);
int64_t dgb_bn_bwd_scratch_bytes(int N);
int dgb_bn_bwd_stats_bf16(const void* dy, int64_t ld_dy, const void* x, int64_t ld_x, const void* y, int64_t ld_y, int64_t M, int N,
                          const float* mean, const float* rstd, float* scratch, float* sums, void* stream);
int dgb_bn_bwd_dx_bf16(const void* dy, int64_t ld_dy, const void* x, int64_t ld_x, const void* y, int64_t ld_y, int64_t M, int N,
                       const float* mean, const float* rstd, const float* gamma, const float* sums, const float* n_rows, void* dx,
                       int64_t ld_dx, void* stream);

/* ---- weight packing (csrc/pack_ops.cu) -------------------------------------------------------------------------------
 * fp32 W = [W_i | W_j] of GraphConvLayer.conv (experiments/OGB/GCN.py:31) -> bf16 wcat [2H, Fp], wcat_t [Fp, 2H] (optional),
 * fp32 bcat [2H] (optional) in one launch per layer and optimizer step. */
int dgb_pack_graphconv_weights(const float* W, const float* bias, int H, int f_in, int Fp, void* wcat, void* wcat_t, float* bcat,
                               void* stream);

/* ---- tcgen05 GEMM / gather-GEMM (csrc/gemm_tcgen05.cu) ---------------------------------------------------------------
 * Reference: nn.Linear -> cuBLAS sgemm on gathered + concatenated operands (experiments/OGB/GCN.py:43-52,
 * experiments/GraphCast/layers.py:52-63, 205-216).  act: 0 none, 1 ReLU, 2 SiLU. */
int dgb_gemm_bf16_tn(const void* A, int64_t lda, const void* B, int64_t ldb, void* C, int64_t ldc, int M, int N, int K,
                     const float* bias, int act, int out_f32, void* stream);
int dgb_gather_gemm_bf16(int n_seg, const void* const* seg_base, const int64_t* seg_ld, const int* const* seg_idx,
                         const int* seg_k, const void* W, int64_t ldw, void* C, int64_t ldc, int M, int N, const float* bias,
                         int act, int out_f32, void* stream);

/* Gather-GEMM whose gathered rows may live on other GPUs: rows j >= seg_nlocal[s] of segment s are read from
 * seg_peer[s][halo_owner[s][j - nlocal]] + halo_row[s][j - nlocal] * seg_ld[s] by the producer warps (NVLink loads inside the
 * tcgen05 GEMM).  Reference: NVSHMEM gather + index + cat + nn.Linear (experiments/GraphCast/layers.py:205-216). */
int dgb_gather_gemm_peer_bf16(int n_seg, const void* const* seg_base, const int64_t* seg_ld, const int* const* seg_idx,
                              const int* seg_k, const void* const* const* seg_peer, const int* const* seg_halo_owner,
                              const int* const* seg_halo_row, const int64_t* seg_nlocal, const void* W, int64_t ldw, void* C,
                              int64_t ldc, int M, int N, const float* bias, int act, int out_f32, void* stream);
/* Split-K weight gradient dW[N1,N2] (fp32) += G[M,N1]^T @ X'[M,N2], X' = X or X[x_idx] (rows gathered inside the kernel);
 * csrc/gemm_wgrad_tcgen05.cu.  Reference: autograd of nn.Linear -> cuBLAS. */
int dgb_gemm_wgrad_bf16(const void* G, int64_t ldg, const void* X, int64_t ldx, float* dW, int64_t ldw, int64_t M, int N1,
                        int N2, const int* x_idx, void* stream);

/* ---- segment softmax + attention-weighted aggregation (csrc/softmax_ops.cu) ------------------------------------------
 * Reference: GAT attention chain, experiments/OGB-LSC/RGAT.py:147-165,191-202 (5 all-to-alls per layer). */
int dgb_seg_softmax_fwd(const int* rowptr, const int* col, int64_t n_dst, int C, float slope, const float* s_i, const float* s_j,
                        const void* h_j, int64_t ld_h, void* out, int64_t ld_out, float* row_max, float* row_den, void* stream);
int dgb_seg_softmax_bwd(const int* rowptr, const int* col, const int* colptr, const int* row, const int* slot_map, int64_t n_dst,
                        int64_t n_src, int C, float slope, const float* s_i, const float* s_j, const void* h_j, int64_t ld_h,
                        const void* out, int64_t ld_out, const float* row_max, const float* row_den, const void* g, int64_t ld_g,
                        float* dscore, float* ds_i, float* ds_j, void* dh, int64_t ld_dh, void* stream);

/* ---- fused LayerNorm (+ residual) and SiLU gradient (csrc/norm_ops.cu) ------------------------------------------------
 * Reference: nn.LayerNorm / nn.SiLU / residual adds of MeshGraphMLP and the mesh blocks (experiments/GraphCast/layers.py:52-63,
 * 142, 215).  Only the first C_valid columns are normalised; columns [C_valid, C) are zero padding. */
int dgb_layernorm_fwd(const void* z, int64_t ld_z, const void* res, int64_t ld_r, const float* gamma, const float* beta, void* y,
                      int64_t ld_y, float* mean, float* rstd, int64_t M, int C, int C_valid, float eps, void* stream);
int64_t dgb_layernorm_bwd_scratch_bytes(int64_t M, int C);
int dgb_layernorm_bwd(const void* gy, int64_t ld_g, const void* z, int64_t ld_z, const float* gamma, const float* mean,
                      const float* rstd, void* dz, int64_t ld_dz, float* scratch, float* dgamma_dbeta, int64_t M, int C,
                      int C_valid, void* stream);
int dgb_silu_bwd(const void* g, const void* pre, void* out, int64_t n, void* stream);

/* ---- rank-local masked row copy / accumulate (csrc/rank_local.cu) -----------------------------------------------------
 * Reference: torch_local.{local_masked_gather, local_masked_scatter, local_masked_scatter_gather,
 * local_masked_scatter_add_gather} (DGraph/distributed/csrc/torch_local_kernels.cu:28-254).  mode 0 = set, 1 = add. */
int dgb_rows_indexed(void* out, int64_t ld_out_bytes, const void* src, int64_t ld_src_bytes, const int64_t* dst_idx,
                     const int64_t* src_idx, const int64_t* placement, int rank, int64_t n, int row_bytes, int elem_bytes,
                     int mode, void* stream);

/* ---- column reductions (csrc/reduce_ops.cu) ---------------------------------------------------------------------------
 * Reference: statistics of DistributedBatchNorm1D (experiments/OGB-LSC/distributed_layers.py:22-74); ATen sum(0) for bias
 * gradients.  Two-phase and deterministic; the caller provides the scratch buffer. */
int64_t dgb_colsum_scratch_bytes(int64_t M, int N, int with_sq);
int dgb_colsum_bf16(const void* x, int64_t ld, int64_t M, int N, int with_sq, float* scratch, float* out, void* stream);
/* Same column sum for callers that cannot create device memory (the GPU is fully owned by their own pool, or a cudaMalloc /
 * cudaFree stall is not acceptable where they run).  This entry point has no counterpart in the reference.  The partial
 * sums go to a 592 x 1024-float (2,424,832-byte) array built into the library and resident on every GPU that loads it, so
 * N * (with_sq ? 2 : 1) must be <= 1024; wider requests return cudaErrorInvalidValue with a message on stderr.  The first
 * call on each device resolves the array's address (do it once before any stream capture).  Calls share the array: do
 * not let two of them overlap on one GPU.  Allocates nothing. */
int dgb_colsum_bf16_noalloc(const void* x, int64_t ld, int64_t M, int N, int with_sq, float* out, void* stream);

/* ---- optimizer (csrc/adam.cu) -------------------------------------------------------------------------------------------
 * Reference: torch.optim.Adam (experiments/OGB/main.py:114).  One kernel over a flat parameter arena. */
int dgb_adam_flat(float* p, const float* g, float* m, float* v, void* p_bf16, int64_t n, float lr, float beta1, float beta2,
                  float eps, float weight_decay, float grad_scale, int step, void* stream);

/* Capturable Adam: step counter (int32) and {lr, 1-beta1^t, 1-beta2^t, -} (4 floats) live in device memory, so the two
 * launches (tick + update) can sit inside a CUDA graph that is replayed every step (reference: torch.optim.Adam only). */
int dgb_adam_flat_capturable(float* p, const float* g, float* m, float* v, void* p_bf16, int64_t n, float beta1,
                                     float beta2, float eps, float weight_decay, float grad_scale, int* step_dev,
                                     float* dev_state, void* stream);

/* ---- peer-memory runtime and kernels (csrc/peer.cu) ---------------------------------------------------------------------
 * Reference: NVSHMEMP2P::{AllocateSymmetricMemory, dist_get, dist_put, barrier} (DGraph/distributed/csrc/
 * torch_nvshmem_p2p.cu:96-376) and NCCLBackendEngine.put (DGraph/distributed/nccl/NCCLBackendEngine.py:281-306).
 * dgb_symm_alloc / dgb_symm_import are the only entry points of the library that allocate or map device memory. */
int dgb_symm_alloc(size_t bytes, void** out);
int dgb_symm_free(void* p);
int dgb_symm_export(void* p, void* handle64);
int dgb_symm_import(const void* handle64, void** out);
int dgb_symm_unmap(void* p);
int dgb_enable_peer_access(int peer);
int dgb_peer_barrier(void* const* pads, int rank, int world, unsigned epoch, int* timeout_flag, long long max_spin,
                     void* stream);
int dgb_halo_pull(void* dst, int64_t ld_dst_bytes, int64_t n_halo, int row_bytes, const void* const* peer_base,
                  int64_t ld_peer_bytes, const int* halo_owner, const int* halo_row, int64_t rot, void* stream);
int dgb_gather_rows_peer(void* out, int64_t ld_out_bytes, int64_t n, int row_bytes, const void* const* peer_base,
                         int64_t ld_peer_bytes, const int* owner, const int* row, void* stream);
/* peer[owner[i]][row[i], :] += src[i, :] (fp32 vector reductions on NVLink-mapped memory): NVSHMEMP2P::dist_put */
int dgb_scatter_add_rows_peer_f32(const float* src, int64_t ld_src, int64_t n, int F, void* const* peer_base, int64_t ld_peer,
                                  const int* owner, const int* row, void* stream);
int dgb_halo_grad_reduce(void* gx, int64_t ld_gx, const void* g_local, int64_t ld_gl, int H, int64_t n_bnd,
                         const int* bnd_row, const int* bnd_ptr, const int* src_rank, const int* src_row,
                         const void* const* peer_base, int64_t ld_peer, void* stream);
int dgb_adam_allreduce(float* p, const void* const* peer_g, int world, float* m, float* v, float* g_out, int64_t n, float lr,
                       float beta1, float beta2, float eps, float weight_decay, float grad_scale, int step, void* stream);
int dgb_allreduce_sum_f32(float* out, const void* const* peer, int world, int64_t n, void* stream);
/* CUDA-graph-capturable forms: the barrier epoch, the optimizer step and the learning rate are read from device memory instead of
 * being kernel arguments (which a graph freezes at capture time).  An engine uses either dgb_peer_barrier or dgb_peer_barrier_dev
 * for its whole life.  No counterpart in the reference (no CUDA graphs there). */
int dgb_peer_barrier_dev(void* const* pads, int rank, int world, unsigned* epoch_dev, int* timeout_flag, long long max_spin,
                         void* stream);
int dgb_adam_allreduce_capturable(float* p, const void* const* peer_g, int world, float* m, float* v, float* g_out, int64_t n,
                                  float beta1, float beta2, float eps, float weight_decay, float grad_scale, int* step_dev,
                                  float* dev_state, void* stream);

/* ---- fused masked cross-entropy (csrc/loss_ops.cu); reference: ATen chain in experiments/OGB/main.py:115,140-148 ---- */
int dgb_masked_ce_fwd(const void* logits, int64_t ld, int is_f32, const int64_t* labels, const int64_t* idx,
                              int64_t n_sel, int C, float* loss_sum, float* lse, void* stream);
int dgb_masked_ce_bwd(const void* logits, int64_t ld, int is_f32, const int64_t* labels, const int64_t* idx,
                              int64_t n_sel, int C, const float* lse, const float* gout, void* dsel, int64_t ld_o,
                              void* stream);

/* The same GEMM with its N tile of B resident in shared memory and 5-8 A stages (tall-skinny, small K): the Python layer takes it
 * for K <= 128 (2.45 M x 256 x 104: 0.47 -> 0.32 ms on B200).  Returns 801 (cudaErrorNotSupported) when the shape does not fit.  No
 * counterpart in the reference. */
int dgb_gemm_bf16_tn_bres(const void* A, int64_t lda, const void* B, int64_t ldb, void* C, int64_t ldc, int M, int N, int K,
                          const float* bias, int act, void* stream);

/* A whole GraphCast MLP block in one kernel (csrc/mlp_fused_tcgen05.cu; the default for the block shapes it handles),
 *   y = LayerNorm(SiLU(cat_s(A_s[idx_s]) @ W1^T + b1) @ W2^T + b2) * gamma + beta (+ residual),   hidden = out = 128;
 * reference: MeshGraphMLP inside MeshEdgeBlock / MeshNodeBlock, experiments/GraphCast/layers.py:24-75,136-143,205-216 (nine ATen
 * kernels there).  W1p as for dgb_gather_gemm_bf16; pre / z / mean / rstd are saved for the backward.  Returns 801 when the
 * shapes or alignments are not handled (the caller then runs the gather-GEMM / GEMM / LayerNorm chain). */
int dgb_mlp2_ln_fused_bf16(int n_seg, const void* const* seg_base, const int64_t* seg_ld, const int* const* seg_idx,
                           const int* seg_k, const void* W1p, int64_t ldw1, const float* b1, const void* W2, int64_t ldw2,
                           const float* b2, const float* gamma, const float* beta, float eps, const void* residual,
                           int64_t ld_res, void* pre, int64_t ld_pre, void* z, int64_t ld_z, void* y, int64_t ld_y, float* mean,
                           float* rstd, int M, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* DGRAPH_B200_H_ */
