// Fused masked cross-entropy (F9): loss over a subset of rows of a [V, C] logits matrix, one warp per selected row.
//
// Replaces the ATen chain of the reference trainer (experiments/OGB/main.py:115,140-148: `out[mask]` -> CrossEntropyLoss ->
// backward = nll_loss_backward + log_softmax_backward + index_select_backward + zero fill): 8 launches / ~1.2 ms per step on the
// products-shaped graph, dominated by single-block reduction kernels.  Here: one forward kernel (log-sum-exp, loss, mean folded
// in, one atomicAdd per block) and one backward kernel that writes the *compact* gradient rows (softmax - onehot) * scale; the
// consumer (ops/gemm.py::_LinearTC.backward) keeps working on the compact rows, because every other row of dLogits is
// structurally zero.
#include "common.cuh"

namespace dgb {

template <typename T>
__device__ __forceinline__ float ld_logit(const T* p);
template <>
__device__ __forceinline__ float ld_logit<float>(const float* p) { return __ldg(p); }
template <>
__device__ __forceinline__ float ld_logit<__nv_bfloat16>(const __nv_bfloat16* p) {
  return __bfloat162float(*p);
}

__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// loss_sum += (1/n_sel) * sum_w (lse_w - logit[w, y_w]);  lse[w] saved for the backward kernel
template <typename T>
__global__ void __launch_bounds__(256) masked_ce_fwd_kernel(const T* __restrict__ logits, int64_t ld,
                                                            const int64_t* __restrict__ labels,
                                                            const int64_t* __restrict__ idx, int64_t n_sel, int C,
                                                            float inv_n, float* __restrict__ loss_sum,
                                                            float* __restrict__ lse_out) {
  __shared__ float s_part[8];
  const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
  const int64_t w = (int64_t)blockIdx.x * 8 + wib;
  float li = 0.f;
  if (w < n_sel) {
    const int64_t row = idx ? idx[w] : w;
    const T* lr = logits + row * ld;
    float m = -INFINITY;
    for (int c = lane; c < C; c += 32) m = fmaxf(m, ld_logit<T>(lr + c));
    m = warp_max(m);
    float s = 0.f;
    for (int c = lane; c < C; c += 32) s += __expf(ld_logit<T>(lr + c) - m);
    s = warp_sum(s);
    const float lse = m + __logf(s);
    const int64_t y = labels[row];
    if (lane == 0) {
      lse_out[w] = lse;
      li = (y >= 0 && y < C) ? lse - ld_logit<T>(lr + y) : 0.f;
    }
  }
  if (lane == 0) s_part[wib] = li;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) t += s_part[i];
    atomicAdd(loss_sum, t * inv_n);
  }
}

// dsel[w, c] = (exp(logit[w, c] - lse_w) - [c == y_w]) * gout / n_sel   (bf16, columns C..ld_o-1 zero)
template <typename T>
__global__ void __launch_bounds__(256) masked_ce_bwd_kernel(const T* __restrict__ logits, int64_t ld,
                                                            const int64_t* __restrict__ labels,
                                                            const int64_t* __restrict__ idx, int64_t n_sel, int C,
                                                            const float* __restrict__ lse, const float* __restrict__ gout,
                                                            float inv_n, __nv_bfloat16* __restrict__ dsel, int64_t ld_o) {
  const int lane = threadIdx.x & 31;
  const int64_t w = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
  if (w >= n_sel) return;
  const int64_t row = idx ? idx[w] : w;
  const T* lr = logits + row * ld;
  const float l = lse[w];
  const float scale = gout[0] * inv_n;
  const int64_t y = labels[row];
  __nv_bfloat16* o = dsel + w * ld_o;
  for (int c = lane; c < (int)ld_o; c += 32) {
    float d = 0.f;
    if (c < C) d = (__expf(ld_logit<T>(lr + c) - l) - (c == y ? 1.f : 0.f)) * scale;
    o[c] = __float2bfloat16_rn(d);
  }
}

}  // namespace dgb

// logits: [V, C] with row stride ld (elements), bf16 (is_f32 = 0) or fp32; labels: int64 [V]; idx: int64 [n_sel] row ids (null =
// the first n_sel rows); loss_sum: fp32 scalar, must be zero on entry; lse: fp32 [n_sel].
DGB_API int dgb_masked_ce_fwd(const void* logits, int64_t ld, int is_f32, const int64_t* labels, const int64_t* idx,
                              int64_t n_sel, int C, float* loss_sum, float* lse, void* stream) {
  if (n_sel == 0) return 0;
  if (C <= 0 || !loss_sum || !lse) return (int)cudaErrorInvalidValue;
  const unsigned blocks = (unsigned)((n_sel + 7) / 8);
  const float inv_n = 1.f / (float)n_sel;
  if (is_f32)
    dgb::masked_ce_fwd_kernel<float><<<blocks, 256, 0, (cudaStream_t)stream>>>((const float*)logits, ld, labels, idx, n_sel, C,
                                                                               inv_n, loss_sum, lse);
  else
    dgb::masked_ce_fwd_kernel<__nv_bfloat16><<<blocks, 256, 0, (cudaStream_t)stream>>>(
        (const __nv_bfloat16*)logits, ld, labels, idx, n_sel, C, inv_n, loss_sum, lse);
  DGB_CHECK_LAUNCH();
  return 0;
}

// gout: fp32 device scalar (gradient of the mean loss); dsel: bf16 [n_sel, ld_o], ld_o >= C (padding columns are zeroed).
DGB_API int dgb_masked_ce_bwd(const void* logits, int64_t ld, int is_f32, const int64_t* labels, const int64_t* idx,
                              int64_t n_sel, int C, const float* lse, const float* gout, void* dsel, int64_t ld_o,
                              void* stream) {
  if (n_sel == 0) return 0;
  if (C <= 0 || ld_o < C || !lse || !gout || !dsel) return (int)cudaErrorInvalidValue;
  const unsigned blocks = (unsigned)((n_sel + 7) / 8);
  const float inv_n = 1.f / (float)n_sel;
  if (is_f32)
    dgb::masked_ce_bwd_kernel<float><<<blocks, 256, 0, (cudaStream_t)stream>>>(
        (const float*)logits, ld, labels, idx, n_sel, C, lse, gout, inv_n, (__nv_bfloat16*)dsel, ld_o);
  else
    dgb::masked_ce_bwd_kernel<__nv_bfloat16><<<blocks, 256, 0, (cudaStream_t)stream>>>(
        (const __nv_bfloat16*)logits, ld, labels, idx, n_sel, C, lse, gout, inv_n, (__nv_bfloat16*)dsel, ld_o);
  DGB_CHECK_LAUNCH();
  return 0;
}
