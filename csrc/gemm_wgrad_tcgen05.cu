// Split-K tcgen05 weight-gradient GEMM:  dW[N1, N2] += G[M, N1]^T @ X[M, N2]   (bf16 in, fp32 out, reduction over M rows)
//
// Both operands are "MN-major" for this product (the reduction index M is the slow, strided dimension of the row-major
// activations), so TMA loads [64 k-rows x 64 columns] boxes with the 128-byte swizzle and the UMMA descriptors use the
// MN-major canonical layout (8-row atoms of 1024 B along K, 64-column panels LBO apart), a_major = b_major = 1 in the
// instruction descriptor.  The grid is (N1 / 128 output tiles) x (K splits): every CTA reduces a contiguous slice of the
// M rows into a 128 x BLOCK_N2 fp32 accumulator in TMEM and adds it to dW with vector reductions (red.global.add.v4.f32).
// Replaces the `dY^T X` library GEMMs of every Linear backward (the reference reaches cuBLAS through autograd).
#include "tc_common.cuh"

namespace dgb {

constexpr int WG_BM = 128;  // rows of dW per CTA (columns of G)
constexpr int WG_BK = 64;   // reduction rows per pipeline stage

template <int BN2>
struct WgSmem {
  static constexpr int A_BYTES = 2 * WG_BK * 128;          // two 64-column panels of G: 16 KB
  static constexpr int B_BYTES = (BN2 / 64) * WG_BK * 128; // BN2/64 panels of X
  static constexpr int STAGE = A_BYTES + B_BYTES;
  static constexpr int STAGES = (BN2 >= 256) ? 4 : 5;
  static constexpr int TOTAL = STAGES * STAGE + 256 + 1024;
};

// GX: the rows of X are gathered (X[x_idx[m], :]) by four extra producer warps that write the MN-major swizzled panels
// themselves -- the weight gradient of a gather-GEMM without materialising the gathered operand.
template <int BN2, bool GX>
__global__ void __launch_bounds__(GX ? 320 : 192, 1)
gemm_wgrad_kernel(const __grid_constant__ CUtensorMap tmap_g, const __grid_constant__ CUtensorMap tmap_x, float* __restrict__ dW,
                  int64_t ldw, int N1, int N2, int64_t M, int kb_per_split, const __nv_bfloat16* __restrict__ Xg, int64_t ldx,
                  const int* __restrict__ x_idx) {
  using L = WgSmem<BN2>;
  constexpr int STAGES = L::STAGES;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + STAGES * L::STAGE);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* acc_full = empty_bar + STAGES;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(acc_full + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int m0 = blockIdx.x * WG_BM;
  const int64_t total_kb = (M + WG_BK - 1) / WG_BK;
  const int64_t kb0 = (int64_t)blockIdx.y * kb_per_split;
  const int64_t kb1 = min(total_kb, kb0 + kb_per_split);
  const int nkb = (int)max((int64_t)0, kb1 - kb0);

  if (warp == 0 && lane == 0) {
    tc::tma_prefetch_desc(&tmap_g);
    tc::tma_prefetch_desc(&tmap_x);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < STAGES; ++s) {
      tc::mbar_init(&full_bar[s], GX ? 129 : 1);
      tc::mbar_init(&empty_bar[s], 1);
    }
    tc::mbar_init(acc_full, 1);
    tc::fence_mbar_init();
  }
  if (warp == 2) {
    tc::tmem_alloc(tmem_ptr, BN2 < 32 ? 32 : BN2);
    tc::tmem_relinquish();
  }
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;

  if (nkb > 0) {
    if (warp == 0) {
      if (lane == 0) {
        int stage = 0;
        uint32_t phase = 0;
        for (int kb = 0; kb < nkb; ++kb) {
          tc::mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = smem + stage * L::STAGE;
          uint8_t* sb = sa + L::A_BYTES;
          tc::mbar_expect_tx(&full_bar[stage], GX ? L::A_BYTES : L::STAGE);
          const int krow = (int)((kb0 + kb) * WG_BK);
#pragma unroll
          for (int pnl = 0; pnl < 2; ++pnl) tc::tma_load_2d(sa + pnl * (WG_BK * 128), &tmap_g, &full_bar[stage], m0 + pnl * 64, krow);
          if (!GX) {
#pragma unroll
            for (int pnl = 0; pnl < BN2 / 64; ++pnl)
              tc::tma_load_2d(sb + pnl * (WG_BK * 128), &tmap_x, &full_bar[stage], pnl * 64, krow);
          }
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      }
    } else if (warp == 1) {
      if (lane == 0) {
        constexpr uint32_t idesc = tc::make_idesc_bf16(WG_BM, BN2, 1, 1);
        int stage = 0;
        uint32_t phase = 0;
        for (int kb = 0; kb < nkb; ++kb) {
          tc::mbar_wait(&full_bar[stage], phase);
          tc::tc_fence_after();
          const uint32_t sa = tc::smem_u32(smem + stage * L::STAGE);
          const uint32_t sb = sa + L::A_BYTES;
          const uint64_t adesc = tc::make_desc_mnmajor_sw128(sa, WG_BK * 128);
          const uint64_t bdesc = tc::make_desc_mnmajor_sw128(sb, WG_BK * 128);
#pragma unroll
          for (int k = 0; k < WG_BK / 16; ++k) {
            // 16 reduction rows = two 8-row atoms = 2048 B further along K: +128 in the (addr >> 4) field
            tc::umma_bf16(tmem_base, adesc + (uint64_t)(k * 128), bdesc + (uint64_t)(k * 128), idesc, (kb | k) ? 1u : 0u);
          }
          tc::umma_commit(&empty_bar[stage]);
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
        tc::umma_commit(acc_full);
      }
    } else if (GX && warp >= 6) {
      // gather producers: 128 threads fill the X panels of every stage (row kk of the stage = X[x_idx[krow + kk], :])
      const int gt = threadIdx.x - 6 * 32;
      const int chunk = gt & 7;  // 16-byte chunk inside a 128-byte panel row
      const int rr = gt >> 3;    // rows rr + 16 * i
      int stage = 0;
      uint32_t phase = 0;
      for (int kb = 0; kb < nkb; ++kb) {
        const int64_t krow = (kb0 + kb) * WG_BK;
        int64_t roff[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const int64_t m = krow + rr + 16 * i;
          roff[i] = (m < M) ? (int64_t)__ldg(x_idx + m) * ldx : -1;
        }
        tc::mbar_wait(&empty_bar[stage], phase ^ 1);
        uint8_t* sb = smem + stage * L::STAGE + L::A_BYTES;
#pragma unroll
        for (int pnl = 0; pnl < BN2 / 64; ++pnl) {
          const int col = pnl * 64 + chunk * 8;
          uint4 v[4];
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            v[i] = make_uint4(0, 0, 0, 0);
            if (roff[i] >= 0 && col < N2) v[i] = ldg_nc_v4(Xg + roff[i] + col);
          }
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const int kk = rr + 16 * i;
            *reinterpret_cast<uint4*>(sb + pnl * (WG_BK * 128) + kk * 128 + ((chunk ^ (kk & 7)) << 4)) = v[i];
          }
        }
        tc::fence_proxy_async_smem();
        tc::mbar_arrive(&full_bar[stage]);
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
    } else if (warp >= 2 && warp < 6) {
      // epilogue: 4 warps, thread = one row of the 128 x BN2 accumulator
      const int q = warp & 3;
      tc::mbar_wait(acc_full, 0);
      tc::tc_fence_after();
      const int row = m0 + q * 32 + lane;
      const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16);
#pragma unroll 1
      for (int c = 0; c < BN2; c += 32) {
        if (c >= N2) break;
        uint32_t r[32];
        tc::tmem_ld_32x32(taddr + c, r);
        tc::tmem_ld_wait();
        if (row < N1) {
          float* dst = dW + (int64_t)row * ldw + c;
          if (c + 32 <= N2 && ((reinterpret_cast<uintptr_t>(dst) & 15) == 0)) {
#pragma unroll
            for (int t = 0; t < 32; t += 4)
              asm volatile("red.global.add.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(dst + t), "f"(__uint_as_float(r[t])),
                           "f"(__uint_as_float(r[t + 1])), "f"(__uint_as_float(r[t + 2])), "f"(__uint_as_float(r[t + 3]))
                           : "memory");
          } else {
#pragma unroll
            for (int t = 0; t < 32; ++t)
              if (c + t < N2) atomicAdd(dst + t, __uint_as_float(r[t]));
          }
        }
      }
    }
  }
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 2) tc::tmem_dealloc(tmem_base, BN2 < 32 ? 32 : BN2);
}

template <int BN2, bool GX>
static int launch_wgrad(const CUtensorMap& tg, const CUtensorMap& tx, float* dW, int64_t ldw, int N1, int N2, int64_t M,
                        const void* X, int64_t ldx, const int* x_idx, cudaStream_t s) {
  using L = WgSmem<BN2>;
  auto kern = gemm_wgrad_kernel<BN2, GX>;
  DGB_OPT_IN_SMEM(kern, L::TOTAL);
  int dev = 0, sms = 0;
  DGB_CUDA(cudaGetDevice(&dev));
  DGB_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  const int m_tiles = (N1 + WG_BM - 1) / WG_BM;
  const int64_t total_kb = (M + WG_BK - 1) / WG_BK;
  int splits = sms / m_tiles;
  if (splits < 1) splits = 1;
  if (splits > total_kb) splits = (int)total_kb;
  const int kb_per_split = (int)((total_kb + splits - 1) / splits);
  splits = (int)((total_kb + kb_per_split - 1) / kb_per_split);
  kern<<<dim3(m_tiles, splits), GX ? 320 : 192, L::TOTAL, s>>>(tg, tx, dW, ldw, N1, N2, M, kb_per_split,
                                                               (const __nv_bfloat16*)X, ldx, x_idx);
  DGB_CHECK_LAUNCH();
  return 0;
}

}  // namespace dgb

using namespace dgb;

// dW[N1, N2] (fp32, row stride ldw) += G[M, N1]^T @ X'[M, N2] with X' = X (x_idx null) or X'[m] = X[x_idx[m]] (gathered rows).
// G, X bf16 row-major (16-byte aligned bases, strides multiples of 8 elements).  N2 <= 256.  dW must be initialised by the
// caller (zero for a plain gradient).  `x_rows` = number of rows of X (only needed for the TMA map when x_idx is null).
DGB_API int dgb_gemm_wgrad_bf16(const void* G, int64_t ldg, const void* X, int64_t ldx, float* dW, int64_t ldw, int64_t M,
                                int N1, int N2, const int* x_idx, void* stream) {
  if (M <= 0 || N1 <= 0 || N2 <= 0) return 0;
  if (N2 > 256 || (ldg % 8) || (ldx % 8) || (reinterpret_cast<uintptr_t>(G) & 15) || (reinterpret_cast<uintptr_t>(X) & 15) ||
      M >= (int64_t(1) << 31))
    return (int)cudaErrorInvalidValue;
  CUtensorMap tg, tx;
  int rc = make_tmap_2d(&tg, G, (uint64_t)M, (uint64_t)N1, (uint64_t)ldg, WG_BK, 64, 2, CU_TENSOR_MAP_SWIZZLE_128B);
  if (rc) return rc;
  cudaStream_t s = (cudaStream_t)stream;
  if (x_idx) {
    if (N2 <= 64) return launch_wgrad<64, true>(tg, tg, dW, ldw, N1, N2, M, X, ldx, x_idx, s);
    if (N2 <= 128) return launch_wgrad<128, true>(tg, tg, dW, ldw, N1, N2, M, X, ldx, x_idx, s);
    return launch_wgrad<256, true>(tg, tg, dW, ldw, N1, N2, M, X, ldx, x_idx, s);
  }
  rc = make_tmap_2d(&tx, X, (uint64_t)M, (uint64_t)N2, (uint64_t)ldx, WG_BK, 64, 2, CU_TENSOR_MAP_SWIZZLE_128B);
  if (rc) return rc;
  if (N2 <= 64) return launch_wgrad<64, false>(tg, tx, dW, ldw, N1, N2, M, X, ldx, nullptr, s);
  if (N2 <= 128) return launch_wgrad<128, false>(tg, tx, dW, ldw, N1, N2, M, X, ldx, nullptr, s);
  return launch_wgrad<256, false>(tg, tx, dW, ldw, N1, N2, M, X, ldx, nullptr, s);
}
