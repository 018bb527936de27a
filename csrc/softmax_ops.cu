// Segment softmax + attention-weighted aggregation for GAT / RGAT layers (F3 in SURVEY §8).
//
// Reference chain (experiments/OGB-LSC/RGAT.py:147-165,191-202): cat -> Linear(2C->1) -> LeakyReLU(0.2) -> exp ->
// scatter(sum) -> gather -> divide -> multiply -> scatter(sum): 3 gathers + 2 scatters = 5 all-to-alls per layer, and the
// softmax is not max-stabilised.  Here the attention logit is split like the GraphConv message,
//      score_e = LeakyReLU(s_i[dst] + s_j[src]),   s_i = h_i . a_i + b,   s_j = h_j . a_j,
// edges are walked in CSR order of their destination row, and ONE kernel does the max-stabilised (online) softmax and
// the weighted sum of neighbour rows in a single pass:  out[i] = sum_e softmax_e(score) * h_j[src_e].
// Backward: one CSR walk (per-edge logit gradients + ds_i) and one CSC walk (dh_j, ds_j); deterministic, no atomics.
#include "common.cuh"

namespace dgb {

struct SoftmaxParams {
  const int* ptr;   // CSR over destination rows (fwd, bwd1) or CSC over source rows (bwd2)
  const int* idx;   // fwd/bwd1: source row of each slot; bwd2: destination row of each slot
  int64_t n_rows;
  int C;
  float slope;
  const float* s_i;  // [n_dst]
  const float* s_j;  // [n_src_aug]
  const __nv_bfloat16* h_j;  // [n_src_aug, C]
  int64_t ld_h;
  // forward outputs / saved statistics
  __nv_bfloat16* out;  // [n_dst, C]
  int64_t ld_out;
  float* row_max;  // [n_dst]
  float* row_den;  // [n_dst]
  // backward
  const __nv_bfloat16* g;  // [n_dst, C]
  int64_t ld_g;
  float* dscore;      // [E] per CSR slot: d loss / d (pre-LeakyReLU logit)
  float* ds_i;        // [n_dst]
  float* ds_j;        // [n_src_aug]
  __nv_bfloat16* dh;  // [n_src_aug, C]
  int64_t ld_dh;
  const int* slot_map;  // bwd2: CSR slot of each CSC slot
};

__device__ __forceinline__ float leaky(float x, float slope) { return x > 0.f ? x : slope * x; }

template <int G>
__device__ __forceinline__ float group_sum(float v, unsigned gmask) {
#pragma unroll
  for (int o = G / 2; o > 0; o >>= 1) v += __shfl_xor_sync(gmask, v, o, G);
  return v;
}

// ---------------------------------------------------------------------------------------------- forward
template <int G, int NC>
__global__ void __launch_bounds__(256) seg_softmax_fwd_kernel(const SoftmaxParams p) {
  const int lane = threadIdx.x & 31;
  const int g = lane & (G - 1);
  const int sub = lane / G;
  const int64_t r = ((int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5)) * (32 / G) + sub;
  if (r >= p.n_rows) return;
  const unsigned gmask = (G == 32) ? 0xffffffffu : (((1u << G) - 1u) << (sub * G));
  const int nchunk = p.C >> 3;
  float acc[NC][8];
#pragma unroll
  for (int c = 0; c < NC; ++c)
#pragma unroll
    for (int t = 0; t < 8; ++t) acc[c][t] = 0.f;
  float m = -INFINITY, d = 0.f;
  const float si = p.s_i[r];
  const int beg = p.ptr[r], end = p.ptr[r + 1];
  for (int base = beg; base < end; base += G) {
    const int n = min(G, end - base);
    int my_j = 0;
    float my_sc = 0.f;
    if (g < n) {
      my_j = p.idx[base + g];
      my_sc = leaky(si + p.s_j[my_j], p.slope);
    }
    for (int t = 0; t < n; ++t) {
      const int j = __shfl_sync(gmask, my_j, t, G);
      const float sc = __shfl_sync(gmask, my_sc, t, G);
      const float m_new = fmaxf(m, sc);
      const float scale = __expf(m - m_new);  // exp(-inf) = 0 on the first edge
      const float pe = __expf(sc - m_new);
      d = d * scale + pe;
      m = m_new;
      const __nv_bfloat16* row = p.h_j + (int64_t)j * p.ld_h;
#pragma unroll
      for (int c = 0; c < NC; ++c) {
        const int ch = g + c * G;
        if (ch < nchunk) {
          float f[8];
          unpack8(ldg_nc_v4(row + ch * 8), f);
#pragma unroll
          for (int u = 0; u < 8; ++u) acc[c][u] = fmaf(acc[c][u], scale, pe * f[u]);
        }
      }
    }
  }
  const float inv = d > 0.f ? 1.f / d : 0.f;
#pragma unroll
  for (int c = 0; c < NC; ++c) {
    const int ch = g + c * G;
    if (ch < nchunk) {
#pragma unroll
      for (int u = 0; u < 8; ++u) acc[c][u] *= inv;
      stg_v4(p.out + r * p.ld_out + ch * 8, pack8(acc[c]));
    }
  }
  if (g == 0) {
    p.row_max[r] = m;
    p.row_den[r] = d;
  }
}

// ---------------------------------------------------------------------------------------------- backward, CSR walk
// dscore[e] = alpha_e * (g_i . h_j[e] - g_i . out_i) * LeakyReLU'(pre_e);  ds_i[i] = sum_e dscore[e]
template <int G, int NC>
__global__ void __launch_bounds__(256) seg_softmax_bwd_dst_kernel(const SoftmaxParams p) {
  const int lane = threadIdx.x & 31;
  const int g = lane & (G - 1);
  const int sub = lane / G;
  const int64_t r = ((int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5)) * (32 / G) + sub;
  if (r >= p.n_rows) return;
  const unsigned gmask = (G == 32) ? 0xffffffffu : (((1u << G) - 1u) << (sub * G));
  const int nchunk = p.C >> 3;
  float gi[NC][8];
  float go = 0.f;  // g_i . out_i
#pragma unroll
  for (int c = 0; c < NC; ++c) {
    const int ch = g + c * G;
#pragma unroll
    for (int t = 0; t < 8; ++t) gi[c][t] = 0.f;
    if (ch < nchunk) {
      float o[8];
      unpack8(ldg_nc_v4(p.g + r * p.ld_g + ch * 8), gi[c]);
      unpack8(ldg_nc_v4(p.out + r * p.ld_out + ch * 8), o);
#pragma unroll
      for (int t = 0; t < 8; ++t) go = fmaf(gi[c][t], o[t], go);
    }
  }
  go = group_sum<G>(go, gmask);
  const float si = p.s_i[r], m = p.row_max[r], d = p.row_den[r];
  const float inv = d > 0.f ? 1.f / d : 0.f;
  float dsi = 0.f;
  const int beg = p.ptr[r], end = p.ptr[r + 1];
  for (int e = beg; e < end; ++e) {
    const int j = p.idx[e];
    const __nv_bfloat16* row = p.h_j + (int64_t)j * p.ld_h;
    float dot = 0.f;
#pragma unroll
    for (int c = 0; c < NC; ++c) {
      const int ch = g + c * G;
      if (ch < nchunk) {
        float f[8];
        unpack8(ldg_nc_v4(row + ch * 8), f);
#pragma unroll
        for (int t = 0; t < 8; ++t) dot = fmaf(gi[c][t], f[t], dot);
      }
    }
    dot = group_sum<G>(dot, gmask);
    const float pre = si + p.s_j[j];
    const float alpha = __expf(leaky(pre, p.slope) - m) * inv;
    const float ds = alpha * (dot - go) * (pre > 0.f ? 1.f : p.slope);
    if (g == 0) p.dscore[e] = ds;
    dsi += ds;
  }
  if (g == 0) p.ds_i[r] = dsi;
}

// ---------------------------------------------------------------------------------------------- backward, CSC walk
// dh[j] = sum_{e into j} alpha_e * g[dst_e];  ds_j[j] = sum_e dscore[e]
template <int G, int NC>
__global__ void __launch_bounds__(256) seg_softmax_bwd_src_kernel(const SoftmaxParams p) {
  const int lane = threadIdx.x & 31;
  const int g = lane & (G - 1);
  const int sub = lane / G;
  const int64_t r = ((int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5)) * (32 / G) + sub;
  if (r >= p.n_rows) return;
  const unsigned gmask = (G == 32) ? 0xffffffffu : (((1u << G) - 1u) << (sub * G));
  const int nchunk = p.C >> 3;
  float acc[NC][8];
#pragma unroll
  for (int c = 0; c < NC; ++c)
#pragma unroll
    for (int t = 0; t < 8; ++t) acc[c][t] = 0.f;
  const float sj = p.s_j[r];
  float dsj = 0.f;
  const int beg = p.ptr[r], end = p.ptr[r + 1];
  for (int base = beg; base < end; base += G) {
    const int n = min(G, end - base);
    int my_i = 0;
    float my_a = 0.f;
    if (g < n) {
      my_i = p.idx[base + g];
      const float pre = p.s_i[my_i] + sj;
      const float d = p.row_den[my_i];
      my_a = d > 0.f ? __expf(leaky(pre, p.slope) - p.row_max[my_i]) / d : 0.f;
      dsj += p.dscore[p.slot_map[base + g]];
    }
    for (int t = 0; t < n; ++t) {
      const int i = __shfl_sync(gmask, my_i, t, G);
      const float a = __shfl_sync(gmask, my_a, t, G);
      const __nv_bfloat16* row = p.g + (int64_t)i * p.ld_g;
#pragma unroll
      for (int c = 0; c < NC; ++c) {
        const int ch = g + c * G;
        if (ch < nchunk) {
          float f[8];
          unpack8(ldg_nc_v4(row + ch * 8), f);
#pragma unroll
          for (int u = 0; u < 8; ++u) acc[c][u] = fmaf(a, f[u], acc[c][u]);
        }
      }
    }
  }
  dsj = group_sum<G>(dsj, gmask);
#pragma unroll
  for (int c = 0; c < NC; ++c) {
    const int ch = g + c * G;
    if (ch < nchunk) stg_v4(p.dh + r * p.ld_dh + ch * 8, pack8(acc[c]));
  }
  if (g == 0) p.ds_j[r] = dsj;
}

template <int WHICH>
static int launch_softmax(const SoftmaxParams& p, cudaStream_t s) {
  if (p.n_rows == 0) return 0;
  const int chunks = p.C / 8;
#define DGB_SM(G, NC)                                                                                        \
  do {                                                                                                       \
    const int64_t rows_per_block = 8 * (32 / (G));                                                           \
    const unsigned blocks = (unsigned)((p.n_rows + rows_per_block - 1) / rows_per_block);                    \
    if (WHICH == 0) seg_softmax_fwd_kernel<G, NC><<<blocks, 256, 0, s>>>(p);                                 \
    else if (WHICH == 1) seg_softmax_bwd_dst_kernel<G, NC><<<blocks, 256, 0, s>>>(p);                        \
    else seg_softmax_bwd_src_kernel<G, NC><<<blocks, 256, 0, s>>>(p);                                        \
  } while (0)
  if (chunks <= 4) DGB_SM(4, 1);
  else if (chunks <= 8) DGB_SM(8, 1);
  else if (chunks <= 16) DGB_SM(16, 1);
  else if (chunks <= 32) DGB_SM(32, 1);
  else if (chunks <= 64) DGB_SM(32, 2);
  else if (chunks <= 128) DGB_SM(32, 4);
  else return (int)cudaErrorInvalidValue;
#undef DGB_SM
  DGB_CHECK_LAUNCH();
  return 0;
}

}  // namespace dgb

using namespace dgb;

// out[i,:] = sum_{e: dst_e = i} softmax_e(LeakyReLU(s_i[i] + s_j[src_e])) * h_j[src_e,:]; saves row max / denominator.
DGB_API int dgb_seg_softmax_fwd(const int* rowptr, const int* col, int64_t n_dst, int C, float slope, const float* s_i,
                                const float* s_j, const void* h_j, int64_t ld_h, void* out, int64_t ld_out, float* row_max,
                                float* row_den, void* stream) {
  if (C % 8 || ld_h % 8 || ld_out % 8) return (int)cudaErrorInvalidValue;
  SoftmaxParams p{};
  p.ptr = rowptr; p.idx = col; p.n_rows = n_dst; p.C = C; p.slope = slope; p.s_i = s_i; p.s_j = s_j;
  p.h_j = (const __nv_bfloat16*)h_j; p.ld_h = ld_h; p.out = (__nv_bfloat16*)out; p.ld_out = ld_out;
  p.row_max = row_max; p.row_den = row_den;
  return launch_softmax<0>(p, (cudaStream_t)stream);
}

// Backward of dgb_seg_softmax_fwd.  dscore is an [E] fp32 work array (per CSR slot) written by the first kernel and
// read by the second through slot_map (CSR slot of each CSC slot).
DGB_API int dgb_seg_softmax_bwd(const int* rowptr, const int* col, const int* colptr, const int* row, const int* slot_map,
                                int64_t n_dst, int64_t n_src, int C, float slope, const float* s_i, const float* s_j,
                                const void* h_j, int64_t ld_h, const void* out, int64_t ld_out, const float* row_max,
                                const float* row_den, const void* g, int64_t ld_g, float* dscore, float* ds_i, float* ds_j,
                                void* dh, int64_t ld_dh, void* stream) {
  if (C % 8 || ld_h % 8 || ld_out % 8 || ld_g % 8 || ld_dh % 8) return (int)cudaErrorInvalidValue;
  SoftmaxParams p{};
  p.ptr = rowptr; p.idx = col; p.n_rows = n_dst; p.C = C; p.slope = slope; p.s_i = s_i; p.s_j = s_j;
  p.h_j = (const __nv_bfloat16*)h_j; p.ld_h = ld_h; p.out = (__nv_bfloat16*)const_cast<void*>(out); p.ld_out = ld_out;
  p.row_max = const_cast<float*>(row_max); p.row_den = const_cast<float*>(row_den);
  p.g = (const __nv_bfloat16*)g; p.ld_g = ld_g; p.dscore = dscore; p.ds_i = ds_i;
  int rc = launch_softmax<1>(p, (cudaStream_t)stream);
  if (rc) return rc;
  p.ptr = colptr; p.idx = row; p.n_rows = n_src; p.slot_map = slot_map; p.ds_j = ds_j;
  p.dh = (__nv_bfloat16*)dh; p.ld_dh = ld_dh;
  return launch_softmax<2>(p, (cudaStream_t)stream);
}
