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
// Most blocks (= rows of partial sums) any column reduction here launches: 148 SMs x 4.  The allocation-free entry point below sizes
// its built-in scratch array with the same constant, so raising it can never make a launch write past that array.
constexpr int COLSUM_MAX_BLOCKS = 592;

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
  if (b > COLSUM_MAX_BLOCKS) b = COLSUM_MAX_BLOCKS;
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

// =====================================================================================================================
// Allocation-free variant of dgb_colsum_bf16.
//
// Who needs it: some hosts of this library are not allowed to create device memory at the moment they need a column
// sum -- for example a runtime that handed the whole GPU to its own memory pool at start-up (a later cudaMalloc simply
// fails there), or a latency-critical loop that must not call cudaMalloc / cudaFree because both can stall the whole
// device.  Those callers still pass their input and output pointers, but the partial-sum workspace that
// dgb_colsum_bf16 expects from them lives in the library image instead: a fixed array in device global memory.
//
// Limits and rules:
//   * Width limit: N * (with_sq ? 2 : 1) <= DGB_COLSUM_NOALLOC_MAX_WIDTH (1024 fp32 columns of partials), i.e. N <= 1024
//     for plain sums and N <= 512 with sums of squares.  Wider requests are refused (message on stderr, return value
//     cudaErrorInvalidValue); use dgb_colsum_bf16 with your own scratch for those.  M is not limited.
//   * Cost: the array is 592 x 1024 floats = 2,424,832 bytes and is resident on EVERY GPU that loads this library, from
//     load until the process exits, whether or not this entry point is ever called.
//   * The first call on each device looks up the array's device address (cudaGetSymbolAddress, once per device, with
//     that device current) before it launches anything.  Make that first call early, outside of any stream capture.
//   * The function calls no cudaMalloc / cudaFree and no allocating library routine.
//   * All calls on a device share the one array, and it keeps whatever the previous call left in it.  Nothing in it has
//     to start at zero: every block overwrites its complete row of partials and the second phase reads exactly the rows
//     written by this call.  Because the array is shared, two calls that may run at the same time on one GPU (different
//     streams, or different host threads using different streams) race with each other and give wrong sums; issue them
//     on one stream or order them yourself.
// =====================================================================================================================
#include <cstdio>
#include <mutex>

namespace dgb {
constexpr int COLSUM_NOALLOC_MAX_WIDTH = 1024;
constexpr int COLSUM_NOALLOC_MAX_BLOCKS = COLSUM_MAX_BLOCKS;  // the one cap colsum_blocks() uses
constexpr int COLSUM_NOALLOC_MAX_DEVICES = 64;
}  // namespace dgb

// File-scope device array.  NOTE for maintainers: in host code this name is only a handle for the runtime's symbol
// table, not a device address -- never pass it to a kernel; pass the pointer obtained from cudaGetSymbolAddress below.
static __device__ __align__(256) float g_colsum_noalloc_scratch[dgb::COLSUM_NOALLOC_MAX_BLOCKS * dgb::COLSUM_NOALLOC_MAX_WIDTH];

static std::mutex g_colsum_noalloc_mu;
static float* g_colsum_noalloc_ptr[dgb::COLSUM_NOALLOC_MAX_DEVICES] = {nullptr};

// Device address of the array on the current device (looked up once per device).
static int colsum_noalloc_scratch(float** out) {
  int dev = -1;
  DGB_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= dgb::COLSUM_NOALLOC_MAX_DEVICES) return (int)cudaErrorInvalidDevice;
  std::lock_guard<std::mutex> lock(g_colsum_noalloc_mu);
  if (g_colsum_noalloc_ptr[dev] == nullptr) {
    float* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress((void**)&p, g_colsum_noalloc_scratch);
    if (e != cudaSuccess) return (int)e;
    if (p == nullptr) return (int)cudaErrorInvalidSymbol;
    g_colsum_noalloc_ptr[dev] = p;
  }
  *out = g_colsum_noalloc_ptr[dev];
  return 0;
}

// out[0:N] = sum_r x[r, :] (and out[N:2N] = sum of squares when with_sq), exactly like dgb_colsum_bf16, but the partial
// sums go to the library's built-in array (see the block comment above for the width limit, the per-GPU memory cost,
// the first-call set-up and the no-overlap rule).
DGB_API int dgb_colsum_bf16_noalloc(const void* x, int64_t ld, int64_t M, int N, int with_sq, float* out, void* stream) {
  if (N % 8 || N < 8 || ld % 8 || M < 0 || !out || (M > 0 && !x)) return (int)cudaErrorInvalidValue;
  const int width = N * (with_sq ? 2 : 1);
  if (width > dgb::COLSUM_NOALLOC_MAX_WIDTH) {
    fprintf(stderr,
            "dgb_colsum_bf16_noalloc: N=%d%s needs %d columns of partial sums, the built-in scratch holds at most %d; "
            "call dgb_colsum_bf16 with a caller-provided scratch buffer instead\n",
            N, with_sq ? " with sums of squares" : "", width, dgb::COLSUM_NOALLOC_MAX_WIDTH);
    return (int)cudaErrorInvalidValue;
  }
  float* scratch = nullptr;
  const int rc = colsum_noalloc_scratch(&scratch);
  if (rc) return rc;
  if (M == 0) {
    DGB_CUDA(cudaMemsetAsync(out, 0, sizeof(float) * width, (cudaStream_t)stream));
    return 0;
  }
  static_assert(dgb::COLSUM_NOALLOC_MAX_BLOCKS >= dgb::COLSUM_MAX_BLOCKS, "built-in scratch must cover every launch");
  return dgb::colsum_launch(x, ld, M, N, with_sq, scratch, out, (cudaStream_t)stream);
}
