// Fused LayerNorm (+ residual) and SiLU-gradient kernels for the GraphCast MLP blocks (F4 epilogue work in SURVEY §8).
//
// Reference: every MeshGraphMLP ends in nn.LayerNorm and every block adds a residual (experiments/GraphCast/layers.py:52-63,
// 142, 215); through ATen that is a bf16->fp32 copy, the fp32 LayerNorm kernels, a cast back and an add in forward, plus
// layer_norm_grad_input + GammaBetaBackward + casts in backward -- ~35 % of a GraphCast training step on B200 (ncu launch
// list).  Here: ONE bf16-in / bf16-out kernel each way (fp32 statistics in registers), gamma/beta gradients reduced in the
// same backward pass into per-block partials (deterministic two-phase reduction, no atomics).
#include "common.cuh"

namespace dgb {

template <int G>
__device__ __forceinline__ float gsum(float v, unsigned gmask) {
#pragma unroll
  for (int o = G / 2; o > 0; o >>= 1) v += __shfl_xor_sync(gmask, v, o, G);
  return v;
}

// y = LN(z) * gamma + beta (+ res);  mean / rstd saved per row
template <int G, int NC>
__global__ void __launch_bounds__(256) ln_fwd_kernel(const __nv_bfloat16* __restrict__ z, int64_t ld_z,
                                                     const __nv_bfloat16* __restrict__ res, int64_t ld_r,
                                                     const float* __restrict__ gamma, const float* __restrict__ beta,
                                                     __nv_bfloat16* __restrict__ y, int64_t ld_y, float* __restrict__ mean,
                                                     float* __restrict__ rstd, int64_t M, int C, int Cv, float eps) {
  // C = padded width (multiple of 8); Cv <= C = columns that take part in the normalisation (the rest is written as 0)
  const int lane = threadIdx.x & 31, g = lane & (G - 1), sub = lane / G;
  const unsigned gmask = (G == 32) ? 0xffffffffu : (((1u << G) - 1u) << (sub * G));
  const int nchunk = C >> 3;
  const int64_t r = ((int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5)) * (32 / G) + sub;
  if (r >= M) return;
  float x[NC][8];
  float s = 0.f;
#pragma unroll
  for (int c = 0; c < NC; ++c) {
    const int ch = g + c * G;
#pragma unroll
    for (int t = 0; t < 8; ++t) x[c][t] = 0.f;
    if (ch < nchunk) {
      unpack8(ldg_nc_v4(z + r * ld_z + ch * 8), x[c]);
#pragma unroll
      for (int t = 0; t < 8; ++t) {
        if (ch * 8 + t >= Cv) x[c][t] = 0.f;
        s += x[c][t];
      }
    }
  }
  const float mu = gsum<G>(s, gmask) / (float)Cv;
  float v = 0.f;
#pragma unroll
  for (int c = 0; c < NC; ++c) {
    const int ch = g + c * G;
    if (ch < nchunk) {
#pragma unroll
      for (int t = 0; t < 8; ++t) {
        const float d = (ch * 8 + t < Cv) ? x[c][t] - mu : 0.f;
        v = fmaf(d, d, v);
      }
    }
  }
  const float rs = rsqrtf(gsum<G>(v, gmask) / (float)Cv + eps);
#pragma unroll
  for (int c = 0; c < NC; ++c) {
    const int ch = g + c * G;
    if (ch < nchunk) {
      float o[8], rr[8];
      if (res) unpack8(ldg_nc_v4(res + r * ld_r + ch * 8), rr);
#pragma unroll
      for (int t = 0; t < 8; ++t) {
        const int col = ch * 8 + t;
        o[t] = (col < Cv) ? (x[c][t] - mu) * rs * __ldg(gamma + col) + __ldg(beta + col) + (res ? rr[t] : 0.f) : 0.f;
      }
      stg_v4(y + r * ld_y + ch * 8, pack8(o));
    }
  }
  if (g == 0) {
    mean[r] = mu;
    rstd[r] = rs;
  }
}

// dz = rstd * (gy*gamma - mean_C(gy*gamma) - xhat * mean_C(gy*gamma*xhat));  partial[b] = [sum gy*xhat | sum gy] per column
template <int G, int NC>
__global__ void __launch_bounds__(256) ln_bwd_kernel(const __nv_bfloat16* __restrict__ gy, int64_t ld_g,
                                                     const __nv_bfloat16* __restrict__ z, int64_t ld_z,
                                                     const float* __restrict__ gamma, const float* __restrict__ mean,
                                                     const float* __restrict__ rstd, __nv_bfloat16* __restrict__ dz,
                                                     int64_t ld_dz, float* __restrict__ partial, int64_t M, int C, int Cv) {
  extern __shared__ float sm[];  // [groups_per_block][2C]
  const int lane = threadIdx.x & 31, g = lane & (G - 1), sub = lane / G;
  const unsigned gmask = (G == 32) ? 0xffffffffu : (((1u << G) - 1u) << (sub * G));
  const int nchunk = C >> 3;
  const int groups_per_block = (blockDim.x >> 5) * (32 / G);
  const int grp_in_block = (threadIdx.x >> 5) * (32 / G) + sub;
  float dgam[NC][8], dbet[NC][8], gm[NC][8];
#pragma unroll
  for (int c = 0; c < NC; ++c) {
    const int ch = g + c * G;
#pragma unroll
    for (int t = 0; t < 8; ++t) {
      dgam[c][t] = 0.f;
      dbet[c][t] = 0.f;
      gm[c][t] = (ch < nchunk && ch * 8 + t < Cv) ? __ldg(gamma + ch * 8 + t) : 0.f;
    }
  }
  for (int64_t r = (int64_t)blockIdx.x * groups_per_block + grp_in_block; r < M; r += (int64_t)gridDim.x * groups_per_block) {
    const float mu = mean[r], rs = rstd[r];
    float gg[NC][8], xh[NC][8];
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int c = 0; c < NC; ++c) {
      const int ch = g + c * G;
      if (ch < nchunk) {
        float a[8], b[8];
        unpack8(ldg_nc_v4(gy + r * ld_g + ch * 8), a);
        unpack8(ldg_nc_v4(z + r * ld_z + ch * 8), b);
#pragma unroll
        for (int t = 0; t < 8; ++t) {
          const bool valid = ch * 8 + t < Cv;
          if (!valid) a[t] = 0.f;
          xh[c][t] = valid ? (b[t] - mu) * rs : 0.f;
          dgam[c][t] = fmaf(a[t], xh[c][t], dgam[c][t]);
          dbet[c][t] += a[t];
          gg[c][t] = a[t] * gm[c][t];
          s1 += gg[c][t];
          s2 = fmaf(gg[c][t], xh[c][t], s2);
        }
      }
    }
    const float c1 = gsum<G>(s1, gmask) / (float)Cv;
    const float c2 = gsum<G>(s2, gmask) / (float)Cv;
#pragma unroll
    for (int c = 0; c < NC; ++c) {
      const int ch = g + c * G;
      if (ch < nchunk) {
        float o[8];
#pragma unroll
        for (int t = 0; t < 8; ++t) o[t] = (ch * 8 + t < Cv) ? rs * (gg[c][t] - c1 - xh[c][t] * c2) : 0.f;
        stg_v4(dz + r * ld_dz + ch * 8, pack8(o));
      }
    }
  }
  // block reduction of the per-thread column partials
#pragma unroll
  for (int c = 0; c < NC; ++c) {
    const int ch = g + c * G;
    if (ch < nchunk) {
#pragma unroll
      for (int t = 0; t < 8; ++t) {
        sm[grp_in_block * 2 * C + ch * 8 + t] = dgam[c][t];
        sm[grp_in_block * 2 * C + C + ch * 8 + t] = dbet[c][t];
      }
    }
  }
  __syncthreads();
  for (int col = threadIdx.x; col < 2 * C; col += blockDim.x) {
    float a = 0.f;
    for (int k = 0; k < groups_per_block; ++k) a += sm[k * 2 * C + col];
    partial[(int64_t)blockIdx.x * 2 * C + col] = a;
  }
}

__global__ void ln_partial_final_kernel(const float* __restrict__ partial, int nblocks, int W, float* __restrict__ out) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= W) return;
  float a = 0.f;
  for (int b = 0; b < nblocks; ++b) a += partial[(int64_t)b * W + c];
  out[c] = a;
}

// g_pre = g * silu'(pre),  silu'(x) = s (1 + x (1 - s)),  s = sigmoid(x)
__global__ void __launch_bounds__(256) silu_bwd_kernel(const __nv_bfloat16* __restrict__ g, const __nv_bfloat16* __restrict__ pre,
                                                       __nv_bfloat16* __restrict__ out, int64_t n8) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n8) return;
  float a[8], x[8], o[8];
  unpack8(ldg_nc_v4(g + i * 8), a);
  unpack8(ldg_nc_v4(pre + i * 8), x);
#pragma unroll
  for (int t = 0; t < 8; ++t) {
    const float s = 1.f / (1.f + __expf(-x[t]));
    o[t] = a[t] * s * (1.f + x[t] * (1.f - s));
  }
  stg_v4(out + i * 8, pack8(o));
}

static int ln_blocks(int64_t M) {
  int64_t b = (M + 255) / 256;
  return (int)(b < 1 ? 1 : (b > 592 ? 592 : b));
}

}  // namespace dgb

using namespace dgb;

#define DGB_LN_DISPATCH(CALL)                        \
  do {                                               \
    const int chunks = C / 8;                        \
    if (chunks <= 4) { CALL(4, 1); }                 \
    else if (chunks <= 8) { CALL(8, 1); }            \
    else if (chunks <= 16) { CALL(16, 1); }          \
    else if (chunks <= 32) { CALL(32, 1); }          \
    else if (chunks <= 64) { CALL(32, 2); }          \
    else if (chunks <= 128) { CALL(32, 4); }         \
    else return (int)cudaErrorInvalidValue;          \
  } while (0)

// y = LayerNorm(z; gamma, beta, eps) (+ res).  z, res, y bf16 [M, C] (C multiple of 8, <= 1024); gamma, beta fp32 [C];
// mean, rstd fp32 [M] (saved for backward).  res may be null.  Only the first C_valid columns are normalised (statistics
// over C_valid columns); columns [C_valid, C) are zero padding and are written as 0 (e.g. GraphCast's 73-wide output).
DGB_API int dgb_layernorm_fwd(const void* z, int64_t ld_z, const void* res, int64_t ld_r, const float* gamma, const float* beta,
                              void* y, int64_t ld_y, float* mean, float* rstd, int64_t M, int C, int C_valid, float eps,
                              void* stream) {
  if (M == 0) return 0;
  if (C % 8 || C > 1024 || C_valid < 1 || C_valid > C || ld_z % 8 || ld_y % 8 || (res && ld_r % 8))
    return (int)cudaErrorInvalidValue;
  cudaStream_t s = (cudaStream_t)stream;
#define DGB_LNF(G, NC)                                                                                              \
  do {                                                                                                              \
    const int64_t rows_per_block = 8 * (32 / (G));                                                                  \
    ln_fwd_kernel<G, NC><<<(unsigned)((M + rows_per_block - 1) / rows_per_block), 256, 0, s>>>(                      \
        (const __nv_bfloat16*)z, ld_z, (const __nv_bfloat16*)res, ld_r, gamma, beta, (__nv_bfloat16*)y, ld_y, mean, \
        rstd, M, C, C_valid, eps);                                                                                  \
  } while (0)
  DGB_LN_DISPATCH(DGB_LNF);
#undef DGB_LNF
  DGB_CHECK_LAUNCH();
  return 0;
}

DGB_API int64_t dgb_layernorm_bwd_scratch_bytes(int64_t M, int C) { return (int64_t)ln_blocks(M) * 2 * C * (int64_t)sizeof(float); }

// dz (bf16 [M, C]) and dgamma_dbeta (fp32 [2C]: gamma gradient then beta gradient) from gy, z, saved mean / rstd.
// `scratch` holds dgb_layernorm_bwd_scratch_bytes(M, C) bytes.
DGB_API int dgb_layernorm_bwd(const void* gy, int64_t ld_g, const void* z, int64_t ld_z, const float* gamma, const float* mean,
                              const float* rstd, void* dz, int64_t ld_dz, float* scratch, float* dgamma_dbeta, int64_t M,
                              int C, int C_valid, void* stream) {
  if (C_valid < 1 || C_valid > C) return (int)cudaErrorInvalidValue;
  if (C % 8 || C > 512 || ld_g % 8 || ld_z % 8 || ld_dz % 8) return (int)cudaErrorInvalidValue;  // smem: 8 groups x 2C floats
  cudaStream_t s = (cudaStream_t)stream;
  if (M == 0) {
    DGB_CUDA(cudaMemsetAsync(dgamma_dbeta, 0, sizeof(float) * 2 * C, s));
    return 0;
  }
  const int nb = ln_blocks(M);
#define DGB_LNB(G, NC)                                                                                               \
  do {                                                                                                               \
    const size_t smem = (size_t)8 * (32 / (G)) * 2 * C * sizeof(float);                                              \
    ln_bwd_kernel<G, NC><<<nb, 256, smem, s>>>((const __nv_bfloat16*)gy, ld_g, (const __nv_bfloat16*)z, ld_z, gamma,  \
                                              mean, rstd, (__nv_bfloat16*)dz, ld_dz, scratch, M, C, C_valid);        \
  } while (0)
  DGB_LN_DISPATCH(DGB_LNB);
#undef DGB_LNB
  DGB_CHECK_LAUNCH();
  ln_partial_final_kernel<<<(2 * C + 255) / 256, 256, 0, s>>>(scratch, nb, 2 * C, dgamma_dbeta);
  DGB_CHECK_LAUNCH();
  return 0;
}

// out = g * silu'(pre), bf16, n multiple of 8, contiguous
DGB_API int dgb_silu_bwd(const void* g, const void* pre, void* out, int64_t n, void* stream) {
  if (n == 0) return 0;
  if (n % 8) return (int)cudaErrorInvalidValue;
  const int64_t n8 = n / 8;
  silu_bwd_kernel<<<(unsigned)((n8 + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)g, (const __nv_bfloat16*)pre, (__nv_bfloat16*)out, n8);
  DGB_CHECK_LAUNCH();
  return 0;
}
