// Segment (CSR) aggregation kernels: the bf16 GCN / GraphSAGE / GraphConv message-passing hot loop.
//
// Replaces: P2 `zeros -> scatter_add(0, src.expand, m)` (experiments/OGB/GCN.py:57-65), the
// `x[src]`, `x[dst]`, `cat`, ReLU chain of P1 (GCN.py:43-52) after the algebraic split
//      relu(W [x_i ; x_j] + b) = relu(P[i] + Q[j]),  P = X W_i^T + b,  Q = X W_j^T,
// and the local halves of the reference's K6/K7 scatter/gather kernels (local_data_kernels.cuh:301-406).
//
// Design: one sub-warp group (G lanes) owns one output row and walks its CSR segment; every lane keeps
// 8 bf16 columns (one 16-byte load) in fp32 registers, so the reduction is atomic-free and deterministic.
// Neighbour indices are loaded coalesced and broadcast with shuffles; UNROLL independent 16-byte row loads
// are in flight per lane.  When HALO is set, neighbour rows >= n_local are read *directly from the owner
// GPU's matrix* through NVLink peer pointers (pull), so no halo buffer is materialised.
#include "common.cuh"

namespace dgb {

enum SegMode { SEG_SUM = 0, SEG_ADD_RELU = 1, SEG_ADD_RELU_BWD = 2 };

struct SegParams {
  // CSR (or CSC) structure: segment r covers slots [ptr[r], ptr[r+1]) of idx
  const int* ptr;
  const int* idx;
  int64_t n_rows;
  int H;  // columns (bf16), multiple of 8
  // gathered operand (rows indexed by idx)
  const __nv_bfloat16* src;
  int64_t ld_src;
  // second gathered operand (SEG_ADD_RELU_BWD: P rows indexed by idx)
  const __nv_bfloat16* src2;
  int64_t ld_src2;
  // per-output-row operand (SEG_ADD_RELU: P[r]; SEG_ADD_RELU_BWD: Q[r])
  const __nv_bfloat16* rowop;
  int64_t ld_rowop;
  // optional per-slot weight (fp32) and mean normalisation
  const float* slot_w;
  int mean;
  // outputs
  __nv_bfloat16* out;
  int64_t ld_out;
  unsigned short* cnt;  // SEG_ADD_RELU: number of active (pre-activation > 0) edges per (row, col); may be null
  int64_t ld_cnt;
  // halo pull: rows >= n_local of `src` live on peers
  int64_t n_local;
  const __nv_bfloat16* const* peer_base;  // [W] base pointer of the same matrix on every rank
  const int* halo_owner;                  // [n_halo]
  const int* halo_row;                    // [n_halo] row inside the owner's matrix
  int64_t ld_peer;
};

template <bool HALO>
__device__ __forceinline__ const __nv_bfloat16* src_row(const SegParams& p, int j) {
  if (HALO) {
    if (j >= p.n_local) {
      const int h = j - (int)p.n_local;
      return p.peer_base[p.halo_owner[h]] + (int64_t)p.halo_row[h] * p.ld_peer;
    }
  }
  return p.src + (int64_t)j * p.ld_src;
}

template <int MODE, int G, int NC, bool HALO>
__global__ void __launch_bounds__(256) seg_reduce_kernel(const SegParams p) {
  constexpr int UNROLL = 4;
  const int lane = threadIdx.x & 31;
  const int g = lane & (G - 1);
  const int sub = lane / G;
  const int64_t warp = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int64_t r = warp * (32 / G) + sub;
  if (r >= p.n_rows) return;
  const unsigned gmask = (G == 32) ? 0xffffffffu : (((1u << G) - 1u) << (sub * G));

  float acc[NC][8];
  float rop[NC][8];
  unsigned short cn[NC][8];
#pragma unroll
  for (int c = 0; c < NC; ++c) {
    const int colo = (g + c * G) * 8;
#pragma unroll
    for (int t = 0; t < 8; ++t) { acc[c][t] = 0.f; rop[c][t] = 0.f; cn[c][t] = 0; }
    if (MODE != SEG_SUM && colo < p.H) {
      uint4 v = ldg_nc_v4(p.rowop + r * p.ld_rowop + colo);
      unpack8(v, rop[c]);
    }
  }

  const int beg = p.ptr[r], end = p.ptr[r + 1];
  for (int base = beg; base < end; base += G) {
    const int n = min(G, end - base);
    int my_j = 0;
    float my_w = 1.f;
    if (g < n) {
      my_j = p.idx[base + g];
      if (MODE == SEG_SUM && p.slot_w) my_w = p.slot_w[base + g];
    }
    for (int t0 = 0; t0 < n; t0 += UNROLL) {
      uint4 v[UNROLL][NC];
      uint4 v2[UNROLL][NC];
      float w[UNROLL];
#pragma unroll
      for (int u = 0; u < UNROLL; ++u) {
        const int t = t0 + u;
        const int j = __shfl_sync(gmask, my_j, t < n ? t : 0, G);
        w[u] = __shfl_sync(gmask, my_w, t < n ? t : 0, G);
        if (t < n) {
          const __nv_bfloat16* row = src_row<HALO>(p, j);
#pragma unroll
          for (int c = 0; c < NC; ++c) {
            const int colo = (g + c * G) * 8;
            if (colo < p.H) {
              v[u][c] = HALO ? ldg_v4(row + colo) : ldg_nc_v4(row + colo);
              if (MODE == SEG_ADD_RELU_BWD) v2[u][c] = ldg_nc_v4(p.src2 + (int64_t)j * p.ld_src2 + colo);
            }
          }
        }
      }
#pragma unroll
      for (int u = 0; u < UNROLL; ++u) {
        if (t0 + u < n) {
#pragma unroll
          for (int c = 0; c < NC; ++c) {
            const int colo = (g + c * G) * 8;
            if (colo < p.H) {
              float f[8];
              unpack8(v[u][c], f);
              if (MODE == SEG_SUM) {
#pragma unroll
                for (int t = 0; t < 8; ++t) acc[c][t] = fmaf(w[u], f[t], acc[c][t]);
              } else if (MODE == SEG_ADD_RELU) {
#pragma unroll
                for (int t = 0; t < 8; ++t) {
                  const float m = rop[c][t] + f[t];
                  const bool on = m > 0.f;
                  acc[c][t] += on ? m : 0.f;
                  cn[c][t] += on ? 1 : 0;
                }
              } else {  // SEG_ADD_RELU_BWD: f = dOut[i], f2 = P[i], rop = Q[j]
                float f2[8];
                unpack8(v2[u][c], f2);
#pragma unroll
                for (int t = 0; t < 8; ++t) acc[c][t] += (f2[t] + rop[c][t] > 0.f) ? f[t] : 0.f;
              }
            }
          }
        }
      }
    }
  }

  const float scale = (MODE == SEG_SUM && p.mean) ? 1.f / (float)max(end - beg, 1) : 1.f;
#pragma unroll
  for (int c = 0; c < NC; ++c) {
    const int colo = (g + c * G) * 8;
    if (colo < p.H) {
      if (MODE == SEG_SUM && p.mean) {
#pragma unroll
        for (int t = 0; t < 8; ++t) acc[c][t] *= scale;
      }
      stg_v4(p.out + r * p.ld_out + colo, pack8(acc[c]));
      if (MODE == SEG_ADD_RELU && p.cnt) {
        uint4 cv;
        cv.x = cn[c][0] | ((uint32_t)cn[c][1] << 16);
        cv.y = cn[c][2] | ((uint32_t)cn[c][3] << 16);
        cv.z = cn[c][4] | ((uint32_t)cn[c][5] << 16);
        cv.w = cn[c][6] | ((uint32_t)cn[c][7] << 16);
        stg_v4(p.cnt + r * p.ld_cnt + colo, cv);
      }
    }
  }
}

template <int MODE, bool HALO>
static int launch_seg(const SegParams& p, cudaStream_t s) {
  if (p.n_rows == 0) return 0;
  const int chunks = p.H / 8;
  const int warps_per_block = 8;
#define DGB_LAUNCH(G, NC)                                                                 \
  do {                                                                                    \
    const int64_t rows_per_block = (int64_t)warps_per_block * (32 / (G));                 \
    const int64_t blocks = (p.n_rows + rows_per_block - 1) / rows_per_block;              \
    seg_reduce_kernel<MODE, G, NC, HALO><<<(unsigned)blocks, warps_per_block * 32, 0, s>>>(p); \
  } while (0)
  if (chunks <= 4) DGB_LAUNCH(4, 1);
  else if (chunks <= 8) DGB_LAUNCH(8, 1);
  else if (chunks <= 16) DGB_LAUNCH(16, 1);
  else if (chunks <= 32) DGB_LAUNCH(32, 1);
  else if (chunks <= 64) DGB_LAUNCH(32, 2);
  else if (chunks <= 128) DGB_LAUNCH(32, 4);
  else return (int)cudaErrorInvalidValue;
#undef DGB_LAUNCH
  DGB_CHECK_LAUNCH();
  return 0;
}

// ---------------------------------------------------------------------------------------------------
// elementwise: dP = dOut * cnt  (backward of the row-operand of SEG_ADD_RELU)
// ---------------------------------------------------------------------------------------------------
__global__ void mul_cnt_kernel(const __nv_bfloat16* __restrict__ g, int64_t ld_g, const unsigned short* __restrict__ cnt,
                               int64_t ld_c, __nv_bfloat16* __restrict__ out, int64_t ld_o, int64_t rows, int H8) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= rows * H8) return;
  const int64_t r = i / H8;
  const int c = (int)(i % H8) * 8;
  uint4 gv = ldg_nc_v4(g + r * ld_g + c);
  uint4 cv = ldg_nc_v4(cnt + r * ld_c + c);
  float f[8];
  unpack8(gv, f);
  f[0] *= (float)(cv.x & 0xffff); f[1] *= (float)(cv.x >> 16);
  f[2] *= (float)(cv.y & 0xffff); f[3] *= (float)(cv.y >> 16);
  f[4] *= (float)(cv.z & 0xffff); f[5] *= (float)(cv.z >> 16);
  f[6] *= (float)(cv.w & 0xffff); f[7] *= (float)(cv.w >> 16);
  stg_v4(out + r * ld_o + c, pack8(f));
}

}  // namespace dgb

using namespace dgb;

static inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

// out[r,:] = (mean ? 1/deg : 1) * sum_{k in seg r} w[k] * src[idx[k],:]          (bf16 in/out, fp32 accumulate)
// Halo-pull variant: rows >= n_local are fetched from peer_base[halo_owner[h]] + halo_row[h]*ld_peer.
DGB_API int dgb_seg_sum(const int* ptr, const int* idx, int64_t n_rows, int H, const void* src, int64_t ld_src,
                        const float* slot_w, int mean, void* out, int64_t ld_out, int64_t n_local,
                        const void* const* peer_base, const int* halo_owner, const int* halo_row, int64_t ld_peer,
                        void* stream) {
  if (H % 8 || ld_src % 8 || ld_out % 8 || !aligned16(src) || !aligned16(out)) return (int)cudaErrorInvalidValue;
  SegParams p{};
  p.ptr = ptr; p.idx = idx; p.n_rows = n_rows; p.H = H;
  p.src = (const __nv_bfloat16*)src; p.ld_src = ld_src;
  p.slot_w = slot_w; p.mean = mean;
  p.out = (__nv_bfloat16*)out; p.ld_out = ld_out;
  p.n_local = n_local; p.peer_base = (const __nv_bfloat16* const*)peer_base;
  p.halo_owner = halo_owner; p.halo_row = halo_row; p.ld_peer = ld_peer;
  return peer_base ? launch_seg<SEG_SUM, true>(p, (cudaStream_t)stream) : launch_seg<SEG_SUM, false>(p, (cudaStream_t)stream);
}

// out[r,:] = sum_{k in seg r} relu(P[r,:] + Q[idx[k],:]);   cnt[r,c] = #active edges (optional)
DGB_API int dgb_seg_add_relu_fwd(const int* ptr, const int* idx, int64_t n_rows, int H, const void* P, int64_t ld_p,
                                 const void* Q, int64_t ld_q, void* out, int64_t ld_out, unsigned short* cnt,
                                 int64_t ld_cnt, int64_t n_local, const void* const* peer_base,
                                 const int* halo_owner, const int* halo_row, int64_t ld_peer, void* stream) {
  if (H % 8 || ld_p % 8 || ld_q % 8 || ld_out % 8 || (cnt && ld_cnt % 8) || !aligned16(P) || !aligned16(Q) ||
      !aligned16(out))
    return (int)cudaErrorInvalidValue;
  SegParams p{};
  p.ptr = ptr; p.idx = idx; p.n_rows = n_rows; p.H = H;
  p.src = (const __nv_bfloat16*)Q; p.ld_src = ld_q;
  p.rowop = (const __nv_bfloat16*)P; p.ld_rowop = ld_p;
  p.out = (__nv_bfloat16*)out; p.ld_out = ld_out; p.cnt = cnt; p.ld_cnt = ld_cnt;
  p.n_local = n_local; p.peer_base = (const __nv_bfloat16* const*)peer_base;
  p.halo_owner = halo_owner; p.halo_row = halo_row; p.ld_peer = ld_peer;
  return peer_base ? launch_seg<SEG_ADD_RELU, true>(p, (cudaStream_t)stream)
                   : launch_seg<SEG_ADD_RELU, false>(p, (cudaStream_t)stream);
}

// dQ[j,:] = sum_{k in CSC seg j} (P[row[k],:] + Q[j,:] > 0) ? dOut[row[k],:] : 0       (transposed walk)
DGB_API int dgb_seg_add_relu_bwd(const int* colptr, const int* row, int64_t n_aug, int H, const void* dOut,
                                 int64_t ld_g, const void* P, int64_t ld_p, const void* Q, int64_t ld_q, void* dQ,
                                 int64_t ld_dq, void* stream) {
  if (H % 8 || ld_g % 8 || ld_p % 8 || ld_q % 8 || ld_dq % 8) return (int)cudaErrorInvalidValue;
  SegParams p{};
  p.ptr = colptr; p.idx = row; p.n_rows = n_aug; p.H = H;
  p.src = (const __nv_bfloat16*)dOut; p.ld_src = ld_g;
  p.src2 = (const __nv_bfloat16*)P; p.ld_src2 = ld_p;
  p.rowop = (const __nv_bfloat16*)Q; p.ld_rowop = ld_q;
  p.out = (__nv_bfloat16*)dQ; p.ld_out = ld_dq;
  return launch_seg<SEG_ADD_RELU_BWD, false>(p, (cudaStream_t)stream);
}

// dP = dOut * cnt
DGB_API int dgb_mul_cnt(const void* g, int64_t ld_g, const unsigned short* cnt, int64_t ld_c, void* out, int64_t ld_o,
                        int64_t rows, int H, void* stream) {
  if (H % 8 || ld_g % 8 || ld_c % 8 || ld_o % 8) return (int)cudaErrorInvalidValue;
  const int64_t n = rows * (H / 8);
  if (n == 0) return 0;
  mul_cnt_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
      (const __nv_bfloat16*)g, ld_g, cnt, ld_c, (__nv_bfloat16*)out, ld_o, rows, H / 8);
  DGB_CHECK_LAUNCH();
  return 0;
}
