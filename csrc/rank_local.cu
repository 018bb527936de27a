// Rank-local masked row copy / row accumulate: the sm_100a equivalents of the reference's `torch_local` extension
// (DGraph/distributed/csrc/torch_local_kernels.cu:28-254, kernels K4-K7 in local_data_kernels.cuh:160-406):
//     set:  out[dst_idx[i], :]  = src[src_idx[i], :]        (local_masked_scatter_gather)
//     add:  out[dst_idx[i], :] += src[src_idx[i], :]        (local_masked_scatter_add_gather)
// with an optional per-row ownership mask (rank_placement[i] == rank; local_masked_gather / local_masked_scatter).
// Rows move in 16-byte chunks; the accumulate variant uses vector reductions (red.global.add.v4.f32 for fp32,
// red.global.add.noftz.bf16x2 for bf16) instead of the reference's 4 scalar atomicAdds per float4, and -- unlike the
// reference's masked scatter -- applies no ReLU.  Launched on the caller's stream.
#include "common.cuh"

namespace dgb {

template <int G>
__global__ void __launch_bounds__(256) rows_set_kernel(uint8_t* __restrict__ out, int64_t ld_out_b, const uint8_t* __restrict__ src,
                                                       int64_t ld_src_b, const int64_t* __restrict__ dst_idx,
                                                       const int64_t* __restrict__ src_idx, const int64_t* __restrict__ placement,
                                                       int rank, int64_t n, int row_bytes) {
  const int g = threadIdx.x & (G - 1);
  const int64_t grp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) / G;
  const int64_t ngrp = ((int64_t)gridDim.x * blockDim.x) / G;
  const int chunks = row_bytes >> 4;
  for (int64_t i = grp; i < n; i += ngrp) {
    if (placement && placement[i] != rank) continue;
    const uint8_t* s = src + (src_idx ? src_idx[i] : i) * ld_src_b;
    uint8_t* o = out + (dst_idx ? dst_idx[i] : i) * ld_out_b;
    for (int c = g; c < chunks; c += G) stg_v4(o + (c << 4), ldg_nc_v4(s + (c << 4)));
  }
}

__device__ __forceinline__ void red_add_v4_f32(float* p, const uint4& v) {
  asm volatile("red.global.add.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(__uint_as_float(v.x)), "f"(__uint_as_float(v.y)),
               "f"(__uint_as_float(v.z)), "f"(__uint_as_float(v.w))
               : "memory");
}
__device__ __forceinline__ void red_add_bf16x2(void* p, uint32_t v) {
  asm volatile("red.global.add.noftz.bf16x2 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

template <int G, bool BF16>
__global__ void __launch_bounds__(256) rows_add_kernel(uint8_t* __restrict__ out, int64_t ld_out_b, const uint8_t* __restrict__ src,
                                                       int64_t ld_src_b, const int64_t* __restrict__ dst_idx,
                                                       const int64_t* __restrict__ src_idx, const int64_t* __restrict__ placement,
                                                       int rank, int64_t n, int row_bytes) {
  const int g = threadIdx.x & (G - 1);
  const int64_t grp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) / G;
  const int64_t ngrp = ((int64_t)gridDim.x * blockDim.x) / G;
  const int chunks = row_bytes >> 4;
  for (int64_t i = grp; i < n; i += ngrp) {
    if (placement && placement[i] != rank) continue;
    const uint8_t* s = src + (src_idx ? src_idx[i] : i) * ld_src_b;
    uint8_t* o = out + (dst_idx ? dst_idx[i] : i) * ld_out_b;
    for (int c = g; c < chunks; c += G) {
      const uint4 v = ldg_nc_v4(s + (c << 4));
      if (BF16) {
        red_add_bf16x2(o + (c << 4), v.x);
        red_add_bf16x2(o + (c << 4) + 4, v.y);
        red_add_bf16x2(o + (c << 4) + 8, v.z);
        red_add_bf16x2(o + (c << 4) + 12, v.w);
      } else {
        red_add_v4_f32(reinterpret_cast<float*>(o + (c << 4)), v);
      }
    }
  }
}

}  // namespace dgb

using namespace dgb;

static int pick_g(int chunks) { return chunks <= 4 ? 4 : (chunks <= 8 ? 8 : (chunks <= 16 ? 16 : 32)); }

// mode 0: set, 1: add.  elem_bytes 4 (fp32) or 2 (bf16); row_bytes multiple of 16; index arrays int64 or null (identity);
// placement (int64, may be null) restricts the operation to entries with placement[i] == rank.
DGB_API int dgb_rows_indexed(void* out, int64_t ld_out_bytes, const void* src, int64_t ld_src_bytes, const int64_t* dst_idx,
                             const int64_t* src_idx, const int64_t* placement, int rank, int64_t n, int row_bytes,
                             int elem_bytes, int mode, void* stream) {
  if (n == 0) return 0;
  if ((row_bytes & 15) || (ld_out_bytes & 15) || (ld_src_bytes & 15) || (elem_bytes != 2 && elem_bytes != 4))
    return (int)cudaErrorInvalidValue;
  const int G = pick_g(row_bytes >> 4);
  int64_t blocks = (n + (256 / G) - 1) / (256 / G);
  if (blocks > 148 * 16) blocks = 148 * 16;
  cudaStream_t s = (cudaStream_t)stream;
#define DGB_RL(GG)                                                                                                     \
  do {                                                                                                                 \
    if (mode == 0)                                                                                                     \
      rows_set_kernel<GG><<<(unsigned)blocks, 256, 0, s>>>((uint8_t*)out, ld_out_bytes, (const uint8_t*)src,           \
                                                          ld_src_bytes, dst_idx, src_idx, placement, rank, n, row_bytes); \
    else if (elem_bytes == 2)                                                                                          \
      rows_add_kernel<GG, true><<<(unsigned)blocks, 256, 0, s>>>((uint8_t*)out, ld_out_bytes, (const uint8_t*)src,     \
                                                                ld_src_bytes, dst_idx, src_idx, placement, rank, n,   \
                                                                row_bytes);                                           \
    else                                                                                                               \
      rows_add_kernel<GG, false><<<(unsigned)blocks, 256, 0, s>>>((uint8_t*)out, ld_out_bytes, (const uint8_t*)src,    \
                                                                 ld_src_bytes, dst_idx, src_idx, placement, rank, n,  \
                                                                 row_bytes);                                          \
  } while (0)
  if (G == 4) DGB_RL(4); else if (G == 8) DGB_RL(8); else if (G == 16) DGB_RL(16); else DGB_RL(32);
#undef DGB_RL
  DGB_CHECK_LAUNCH();
  return 0;
}
