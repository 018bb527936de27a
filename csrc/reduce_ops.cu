// Column reductions over tall-skinny activations: bias gradients (sum over rows), BatchNorm statistics (sum and
// sum of squares).  Two-phase: every block reduces a strip of rows into one partial row of a scratch buffer, then a
// second kernel folds the partials in a fixed order (deterministic; no atomics).
//
// Replaces `g.sum(0)` ATen reductions on the hot path and the statistics half of the reference's
// DistributedBatchNorm1D (experiments/OGB-LSC/distributed_layers.py:22-74), whose cross-rank part then is one
// all-reduce of a [2F+1] vector (F5 in SURVEY §8).
#include "common.cuh"

namespace dgb {

constexpr int CS_THREADS = 256;

// partial[b, 0:N]   = sum over the block's rows of x[r, :]
// partial[b, N:2N]  = sum of squares (only when SQ)
template <bool SQ>
__global__ void __launch_bounds__(CS_THREADS) colsum_partial_kernel(const __nv_bfloat16* __restrict__ x, int64_t ld,
                                                                    int64_t M, int N, float* __restrict__ partial) {
  extern __shared__ float sm[];  // [rows_per_iter][N * (SQ ? 2 : 1)]
  const int chunks = N >> 3;                       // 8-column chunks per row
  const int rows_per_iter = CS_THREADS / chunks;   // host guarantees chunks <= CS_THREADS
  const int rl = threadIdx.x / chunks;             // row lane
  const int ch = threadIdx.x % chunks;
  const bool active = rl < rows_per_iter;
  float s[8], q[8];
#pragma unroll
  for (int t = 0; t < 8; ++t) { s[t] = 0.f; q[t] = 0.f; }
  const int64_t rows_per_block = (M + gridDim.x - 1) / gridDim.x;
  const int64_t r0 = (int64_t)blockIdx.x * rows_per_block;
  const int64_t r1 = min(M, r0 + rows_per_block);
  if (active) {
    for (int64_t r = r0 + rl; r < r1; r += rows_per_iter) {
      float f[8];
      unpack8(ldg_nc_v4(x + r * ld + ch * 8), f);
#pragma unroll
      for (int t = 0; t < 8; ++t) {
        s[t] += f[t];
        if (SQ) q[t] = fmaf(f[t], f[t], q[t]);
      }
    }
  }
  const int W = N * (SQ ? 2 : 1);
  if (active) {
#pragma unroll
    for (int t = 0; t < 8; ++t) {
      sm[rl * W + ch * 8 + t] = s[t];
      if (SQ) sm[rl * W + N + ch * 8 + t] = q[t];
    }
  }
  __syncthreads();
  for (int c = threadIdx.x; c < W; c += CS_THREADS) {
    float a = 0.f;
    for (int k = 0; k < rows_per_iter; ++k) a += sm[k * W + c];
    partial[(int64_t)blockIdx.x * W + c] = a;
  }
}

__global__ void colsum_final_kernel(const float* __restrict__ partial, int nblocks, int W, float* __restrict__ out) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= W) return;
  float a = 0.f;
  for (int b = 0; b < nblocks; ++b) a += partial[(int64_t)b * W + c];
  out[c] = a;
}

static int colsum_blocks(int64_t M) {
  int64_t b = (M + 1023) / 1024;
  if (b < 1) b = 1;
  if (b > 592) b = 592;  // 148 SMs x 4
  return (int)b;
}

static int colsum_launch(const void* x, int64_t ld, int64_t M, int N, int with_sq, float* scratch, float* out,
                         cudaStream_t s) {
  const int chunks = N / 8;
  const int W = N * (with_sq ? 2 : 1);
  const int nb = colsum_blocks(M);
  const int rows_per_iter = CS_THREADS / chunks;
  const size_t smem = (size_t)rows_per_iter * W * sizeof(float);
  if (with_sq)
    colsum_partial_kernel<true><<<nb, CS_THREADS, smem, s>>>((const __nv_bfloat16*)x, ld, M, N, scratch);
  else
    colsum_partial_kernel<false><<<nb, CS_THREADS, smem, s>>>((const __nv_bfloat16*)x, ld, M, N, scratch);
  DGB_CHECK_LAUNCH();
  colsum_final_kernel<<<(W + 255) / 256, 256, 0, s>>>(scratch, nb, W, out);
  DGB_CHECK_LAUNCH();
  return 0;
}

}  // namespace dgb

using namespace dgb;

// Bytes of scratch dgb_colsum_bf16 needs for an [M, N] input (with_sq: also sum of squares).
DGB_API int64_t dgb_colsum_scratch_bytes(int64_t M, int N, int with_sq) {
  return (int64_t)colsum_blocks(M) * N * (with_sq ? 2 : 1) * (int64_t)sizeof(float);
}

// out[0:N] = sum_r x[r, :]  (and out[N:2N] = sum_r x[r, :]^2 when with_sq).  x bf16 row-major [M, N] with row stride ld;
// N multiple of 8, 8 <= N <= 2048.  `scratch` must hold dgb_colsum_scratch_bytes(M, N, with_sq) bytes.
DGB_API int dgb_colsum_bf16(const void* x, int64_t ld, int64_t M, int N, int with_sq, float* scratch, float* out,
                            void* stream) {
  if (N % 8 || N < 8 || N > 2048 || ld % 8 || M < 0 || !scratch || !out) return (int)cudaErrorInvalidValue;
  if (M == 0) {
    DGB_CUDA(cudaMemsetAsync(out, 0, sizeof(float) * N * (with_sq ? 2 : 1), (cudaStream_t)stream));
    return 0;
  }
  return colsum_launch(x, ld, M, N, with_sq, scratch, out, (cudaStream_t)stream);
}
