// Distributed BatchNorm over the vertex dimension, kernel side (F5).
//
// Reference: DistributedBatchNorm1D (experiments/OGB-LSC/distributed_layers.py:22-74): mean / variance over all vertices of all
// ranks, y = gamma * (x - mean) / sqrt(var + eps) + beta, and its backward -- there a chain of ATen elementwise / reduce kernels
// around 3 + 4 all-reduces.  Here, per call and direction, ONE statistics kernel pair (colsum_partial / colsum_final of
// reduce_ops.cu), ONE all-reduce of a [2F (+1)] vector, and ONE of the kernels below:
//   bn_apply_kernel      y = act(x_hat * gamma + beta), x_hat = (x - mean) * rstd          bf16 in, bf16 out, fp32 math
//   bn_bwd_stats_kernel  partial column sums of dy' and dy' * x_hat (dy' = dy masked by the fused ReLU), x_hat recomputed from x
//   bn_bwd_dx_kernel     dx = gamma * rstd * (dy' - sum(dy') / n - x_hat * sum(dy' * x_hat) / n)
// No fp32 copy of the [N, F] activations exists in either direction.
#include "common.cuh"

namespace dgb {

constexpr int BN_THREADS = 256;
constexpr int BN_MAX_BLOCKS = 592;

// y[r, c] = act((x[r, c] - mean[c]) * rstd[c] * gamma[c] + beta[c]);   act: 0 none, 1 ReLU
__global__ void __launch_bounds__(BN_THREADS) bn_apply_kernel(const __nv_bfloat16* __restrict__ x, int64_t ld_x, int64_t M, int N,
                                                              const float* __restrict__ mean, const float* __restrict__ rstd,
                                                              const float* __restrict__ gamma, const float* __restrict__ beta,
                                                              int act, __nv_bfloat16* __restrict__ y, int64_t ld_y) {
  const int chunks = N >> 3;
  const int64_t total = M * chunks;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = i / chunks;
    const int c = (int)(i - r * chunks) * 8;
    float f[8];
    unpack8(ldg_nc_v4(x + r * ld_x + c), f);
#pragma unroll
    for (int t = 0; t < 8; ++t) {
      const float sc = rstd[c + t] * (gamma ? gamma[c + t] : 1.f);
      float v = (f[t] - mean[c + t]) * sc + (beta ? beta[c + t] : 0.f);
      if (act == 1) v = fmaxf(v, 0.f);
      f[t] = v;
    }
    stg_v4(y + r * ld_y + c, pack8(f));
  }
}

// partial[b, 0:N] = sum_r dy'[r, :], partial[b, N:2N] = sum_r dy'[r, :] * x_hat[r, :] over the block's rows; dy' = dy where the saved
// output y is > 0 (fused ReLU) or dy itself (y == nullptr)
__global__ void __launch_bounds__(BN_THREADS) bn_bwd_stats_kernel(const __nv_bfloat16* __restrict__ dy, int64_t ld_dy,
                                                                  const __nv_bfloat16* __restrict__ x, int64_t ld_x,
                                                                  const __nv_bfloat16* __restrict__ y, int64_t ld_y, int64_t M, int N,
                                                                  const float* __restrict__ mean, const float* __restrict__ rstd,
                                                                  float* __restrict__ partial) {
  extern __shared__ float sm[];  // [rows_per_iter][2N]
  const int chunks = N >> 3;
  const int rows_per_iter = BN_THREADS / chunks;
  const int rl = threadIdx.x / chunks, ch = threadIdx.x % chunks;
  const bool active = rl < rows_per_iter;
  float s[8], q[8], mu[8], rs[8];
#pragma unroll
  for (int t = 0; t < 8; ++t) { s[t] = 0.f; q[t] = 0.f; mu[t] = mean[ch * 8 + t]; rs[t] = rstd[ch * 8 + t]; }
  const int64_t rows_per_block = (M + gridDim.x - 1) / gridDim.x;
  const int64_t r0 = (int64_t)blockIdx.x * rows_per_block, r1 = min(M, r0 + rows_per_block);
  if (active) {
    for (int64_t r = r0 + rl; r < r1; r += rows_per_iter) {
      float g[8], xv[8];
      unpack8(ldg_nc_v4(dy + r * ld_dy + ch * 8), g);
      unpack8(ldg_nc_v4(x + r * ld_x + ch * 8), xv);
      if (y) {
        float yv[8];
        unpack8(ldg_nc_v4(y + r * ld_y + ch * 8), yv);
#pragma unroll
        for (int t = 0; t < 8; ++t) g[t] = yv[t] > 0.f ? g[t] : 0.f;
      }
#pragma unroll
      for (int t = 0; t < 8; ++t) {
        s[t] += g[t];
        q[t] = fmaf(g[t], (xv[t] - mu[t]) * rs[t], q[t]);
      }
    }
  }
  const int W = 2 * N;
  if (active) {
#pragma unroll
    for (int t = 0; t < 8; ++t) { sm[rl * W + ch * 8 + t] = s[t]; sm[rl * W + N + ch * 8 + t] = q[t]; }
  }
  __syncthreads();
  for (int c = threadIdx.x; c < W; c += BN_THREADS) {
    float a = 0.f;
    for (int k = 0; k < rows_per_iter; ++k) a += sm[k * W + c];
    partial[(int64_t)blockIdx.x * W + c] = a;
  }
}

__global__ void bn_fold_kernel(const float* __restrict__ partial, int nblocks, int W, float* __restrict__ out) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= W) return;
  float a = 0.f;
  for (int b = 0; b < nblocks; ++b) a += partial[(int64_t)b * W + c];
  out[c] = a;
}

// dx = gamma * rstd * (dy' - sums[c] * inv_n - x_hat * sums[N + c] * inv_n);  sums = all-reduced [2N] (sum dy', sum dy' * x_hat);
// inv_n is read from device memory (the global row count arrives with the forward statistics all-reduce)
__global__ void __launch_bounds__(BN_THREADS) bn_bwd_dx_kernel(const __nv_bfloat16* __restrict__ dy, int64_t ld_dy,
                                                               const __nv_bfloat16* __restrict__ x, int64_t ld_x,
                                                               const __nv_bfloat16* __restrict__ y, int64_t ld_y, int64_t M, int N,
                                                               const float* __restrict__ mean, const float* __restrict__ rstd,
                                                               const float* __restrict__ gamma, const float* __restrict__ sums,
                                                               const float* __restrict__ n_rows, __nv_bfloat16* __restrict__ dx,
                                                               int64_t ld_dx) {
  const int chunks = N >> 3;
  const int64_t total = M * chunks;
  const float inv_n = 1.f / fmaxf(*n_rows, 1.f);
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = i / chunks;
    const int c = (int)(i - r * chunks) * 8;
    float g[8], xv[8];
    unpack8(ldg_nc_v4(dy + r * ld_dy + c), g);
    unpack8(ldg_nc_v4(x + r * ld_x + c), xv);
    if (y) {
      float yv[8];
      unpack8(ldg_nc_v4(y + r * ld_y + c), yv);
#pragma unroll
      for (int t = 0; t < 8; ++t) g[t] = yv[t] > 0.f ? g[t] : 0.f;
    }
#pragma unroll
    for (int t = 0; t < 8; ++t) {
      const float xh = (xv[t] - mean[c + t]) * rstd[c + t];
      g[t] = (gamma ? gamma[c + t] : 1.f) * rstd[c + t] * (g[t] - sums[c + t] * inv_n - xh * sums[N + c + t] * inv_n);
    }
    stg_v4(dx + r * ld_dx + c, pack8(g));
  }
}

static int bn_grid(int64_t work) {
  int64_t b = (work + BN_THREADS - 1) / BN_THREADS;
  if (b > 148 * 16) b = 148 * 16;
  return (int)(b < 1 ? 1 : b);
}

}  // namespace dgb

using namespace dgb;

static inline bool bn_ok(const void* p, int64_t ld) { return p && (reinterpret_cast<uintptr_t>(p) & 15) == 0 && ld % 8 == 0; }

DGB_API int dgb_bn_apply_bf16(const void* x, int64_t ld_x, int64_t M, int N, const float* mean, const float* rstd, const float* gamma,
                              const float* beta, int act, void* y, int64_t ld_y, void* stream) {
  if (M == 0) return 0;
  if (N % 8 || N < 8 || !bn_ok(x, ld_x) || !bn_ok(y, ld_y) || !mean || !rstd || act < 0 || act > 1) return (int)cudaErrorInvalidValue;
  bn_apply_kernel<<<bn_grid(M * (N / 8)), BN_THREADS, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)x, ld_x, M, N, mean, rstd, gamma,
                                                                               beta, act, (__nv_bfloat16*)y, ld_y);
  DGB_CHECK_LAUNCH();
  return 0;
}

// Bytes of scratch dgb_bn_bwd_stats_bf16 needs.
DGB_API int64_t dgb_bn_bwd_scratch_bytes(int N) { return (int64_t)BN_MAX_BLOCKS * 2 * N * (int64_t)sizeof(float); }

// sums[0:N] = sum dy', sums[N:2N] = sum dy' * x_hat over this rank's rows (y: saved output of a fused ReLU, or null).  N <= 2048.
DGB_API int dgb_bn_bwd_stats_bf16(const void* dy, int64_t ld_dy, const void* x, int64_t ld_x, const void* y, int64_t ld_y, int64_t M,
                                  int N, const float* mean, const float* rstd, float* scratch, float* sums, void* stream) {
  if (N % 8 || N < 8 || N > 2048 || !bn_ok(dy, ld_dy) || !bn_ok(x, ld_x) || (y && !bn_ok(y, ld_y)) || !scratch || !sums)
    return (int)cudaErrorInvalidValue;
  cudaStream_t s = (cudaStream_t)stream;
  if (M == 0) {
    DGB_CUDA(cudaMemsetAsync(sums, 0, sizeof(float) * 2 * N, s));
    return 0;
  }
  int64_t nb = (M + 1023) / 1024;
  if (nb > BN_MAX_BLOCKS) nb = BN_MAX_BLOCKS;
  const int rows_per_iter = BN_THREADS / (N / 8);
  bn_bwd_stats_kernel<<<(unsigned)nb, BN_THREADS, (size_t)rows_per_iter * 2 * N * sizeof(float), s>>>(
      (const __nv_bfloat16*)dy, ld_dy, (const __nv_bfloat16*)x, ld_x, (const __nv_bfloat16*)y, ld_y, M, N, mean, rstd, scratch);
  DGB_CHECK_LAUNCH();
  bn_fold_kernel<<<(2 * N + 255) / 256, 256, 0, s>>>(scratch, (int)nb, 2 * N, sums);
  DGB_CHECK_LAUNCH();
  return 0;
}

DGB_API int dgb_bn_bwd_dx_bf16(const void* dy, int64_t ld_dy, const void* x, int64_t ld_x, const void* y, int64_t ld_y, int64_t M, int N,
                               const float* mean, const float* rstd, const float* gamma, const float* sums, const float* n_rows,
                               void* dx, int64_t ld_dx, void* stream) {
  if (M == 0) return 0;
  if (N % 8 || N < 8 || !bn_ok(dy, ld_dy) || !bn_ok(x, ld_x) || (y && !bn_ok(y, ld_y)) || !bn_ok(dx, ld_dx) || !sums || !n_rows)
    return (int)cudaErrorInvalidValue;
  bn_bwd_dx_kernel<<<bn_grid(M * (N / 8)), BN_THREADS, 0, (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)dy, ld_dy, (const __nv_bfloat16*)x, ld_x, (const __nv_bfloat16*)y, ld_y, M, N, mean, rstd, gamma, sums,
      n_rows, (__nv_bfloat16*)dx, ld_dx);
  DGB_CHECK_LAUNCH();
  return 0;
}

// mean = s / n, var = max(q / n - mean^2, 0), rstd = 1 / sqrt(var + eps) from the all-reduced buffer [sum (N) | sum of squares (N) | rows (1)];
// one launch instead of seven [N]-sized ATen kernels
namespace dgb {
__global__ void bn_finalize_stats_kernel(const float* __restrict__ buf, int N, float eps, float* __restrict__ mean,
                                         float* __restrict__ var, float* __restrict__ rstd) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= N) return;
  const float n = fmaxf(buf[2 * N], 1.f);
  const float m = buf[c] / n;
  const float v = fmaxf(buf[N + c] / n - m * m, 0.f);
  mean[c] = m;
  var[c] = v;
  rstd[c] = rsqrtf(v + eps);
}
}  // namespace dgb

DGB_API int dgb_bn_finalize_stats(const float* buf, int N, float eps, float* mean, float* var, float* rstd, void* stream) {
  if (N <= 0 || !buf || !mean || !var || !rstd) return (int)cudaErrorInvalidValue;
  dgb::bn_finalize_stats_kernel<<<(N + 255) / 256, 256, 0, (cudaStream_t)stream>>>(buf, N, eps, mean, var, rstd);
  DGB_CHECK_LAUNCH();
  return 0;
}
