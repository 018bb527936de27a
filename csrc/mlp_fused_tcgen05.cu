// F4: one kernel for a whole GraphCast MLP block (the default GraphCast path; DGRAPH_B200_FUSED_MLP=0 selects the three-kernel
// chain; numerics: tests/test_mlp_fused_gpu.py):
//
//     pre = cat([A_0[idx_0], A_1[idx_1], A_2[idx_2]], 1) @ W1^T + b1          (gather-GEMM, tcgen05, K streamed)
//     h   = SiLU(pre)                                                          (TMEM -> registers -> swizzled smem, never HBM)
//     z   = h @ W2^T + b2                                                      (second tcgen05 GEMM, W2 resident in smem)
//     y   = LayerNorm(z) * gamma + beta (+ residual)                           (row statistics in the epilogue)
//
// for hidden = out = 128, i.e. `MeshGraphMLP` with one hidden layer as used by every edge / node block of the reference's
// GraphCast (experiments/GraphCast/layers.py:24-75 MeshGraphMLP, :78-143 MeshEdgeBlock, :146-216 MeshNodeBlock).  The
// reference runs this as index_select x2, cat, Linear, SiLU, Linear, LayerNorm, add: nine ATen kernels and seven [E, *]
// round trips through HBM.  Here the only HBM traffic is the gathered operand rows, the residual, the output and the two
// tensors the backward needs (`pre` for the SiLU gradient, `z` + mean / rstd for the LayerNorm gradient).
//
// Roles (448 threads): warp 0 = TMA producer (W1 k-blocks; W2 once), warp 1 = MMA issuer, warps 2..9 = epilogue (two warps
// per TMEM lane quarter, 64 columns each), warps 10..13 = gather producers (A rows -> 128-byte-swizzled smem).
// Per 128-row tile the MMA thread issues GEMM1(i+1) *before* GEMM2(i), so the tensor pipe and the producers never wait for
// the SiLU epilogue of tile i; the epilogue runs phase 1 (pre, h) and phase 2 (z, LayerNorm, y) back to back per tile.
// TMEM: two GEMM1 accumulators (2 x 128 columns) + one GEMM2 accumulator (128 columns).
// Shared memory: 3 x (A 16 KB + W1 16 KB) ring, W2 32 KB, H 32 KB, 8 x 4 KB store staging.  The H tile of an epilogue warp
// (32 rows x 64 columns, 4 KB, SWIZZLE_128B) is at the same time that warp's slice of the GEMM2 A operand and -- once GEMM2
// has consumed it -- its staging tile for the TMA store of z.
#include "tc_common.cuh"

namespace dgb {
namespace mlpf {

constexpr int BM = 128;      // rows per tile
constexpr int BN = 128;      // hidden = output width
constexpr int BK = 64;       // k-block: 64 bf16 = one 128-byte swizzle row
constexpr int UK = 16;       // UMMA K
constexpr int STAGES = 3;
constexpr int A_BYTES = BM * BK * 2;          // 16 KB
constexpr int B_BYTES = BN * BK * 2;          // 16 KB
constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
constexpr int W2_BYTES = BN * BN * 2;         // 32 KB = two k-blocks
constexpr int H_BYTES = BM * BN * 2;          // 32 KB = two k-blocks
constexpr int NUM_EPI = 8;
constexpr int EPI0 = 2;
constexpr int GATHER0 = EPI0 + NUM_EPI;       // warps 10..13
constexpr int THREADS = (GATHER0 + 4) * 32;   // 448
constexpr int STORE_BYTES = NUM_EPI * 4096;
constexpr int PARAM_BYTES = 4 * BN * 4 + 2 * BM * 8;  // b1, b2, gamma, beta [128] fp32 + row statistics [2][128] float2
constexpr int BAR_BYTES = 256;
constexpr int SMEM_TOTAL = STAGES * STAGE_BYTES + W2_BYTES + H_BYTES + STORE_BYTES + BAR_BYTES + PARAM_BYTES + 1024;
constexpr int TMEM_COLS = 512;                // 2 x 128 (GEMM1) + 128 (GEMM2) rounded up to a power of two

struct Params {
  int M;
  int n_seg;
  const __nv_bfloat16* seg_base[3];
  const int* seg_idx[3];  // null = identity
  int64_t seg_ld[3];
  int seg_kblocks[3];
  int seg_k[3];
  const float *b1, *b2, *gamma, *beta;  // [128] fp32 (b1 / b2 may be null)
  float eps;
  const __nv_bfloat16* residual;  // [M, 128] or null
  int64_t ld_res;
  float *mean, *rstd;  // [M] fp32
};

__device__ __forceinline__ float silu_f(float v) { return v / (1.f + __expf(-v)); }

__device__ __forceinline__ void epi_bar_sync() { asm volatile("bar.sync 1, %0;" ::"n"(NUM_EPI * 32) : "memory"); }

// 32 rows x 64 bf16 columns of one epilogue warp: lane = row, v = that row's 64 values (already rounded where it matters)
// -> SWIZZLE_128B tile at `st` (the layout TMA stores read and tcgen05 K-major descriptors expect).
__device__ __forceinline__ void stage_rows(uint8_t* st, int lane, const float* v) {
#pragma unroll
  for (int t = 0; t < 8; ++t) *reinterpret_cast<uint4*>(st + lane * 128 + ((t ^ (lane & 7)) << 4)) = pack8(v + t * 8);
}

// wait until every bulk store this warp has issued has finished READING its shared-memory source, then let the warp reuse it
__device__ __forceinline__ void staging_reusable(int lane) {
  if (lane == 0) tc::tma_store_wait_read();
  __syncwarp();
}

__global__ void __launch_bounds__(THREADS, 1)
mlp2_ln_fused_kernel(const __grid_constant__ CUtensorMap tmap_w1, const __grid_constant__ CUtensorMap tmap_w2,
                     const __grid_constant__ CUtensorMap tmap_pre, const __grid_constant__ CUtensorMap tmap_z,
                     const __grid_constant__ CUtensorMap tmap_y, const Params p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* w2_s = smem + STAGES * STAGE_BYTES;  // every region below is a multiple of 1024 bytes: all stay 1024-aligned
  uint8_t* h_s = w2_s + W2_BYTES;
  uint8_t* store_s = h_s + H_BYTES;
  uint8_t* bar_base = store_s + STORE_BYTES;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(bar_base);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* acc1_full = empty_bar + STAGES;  // [2]
  uint64_t* acc1_empty = acc1_full + 2;      // [2]
  uint64_t* h_full = acc1_empty + 2;
  uint64_t* acc2_full = h_full + 1;
  uint64_t* w2_full = acc2_full + 1;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(w2_full + 1);
  float* s_b1 = reinterpret_cast<float*>(bar_base + BAR_BYTES);
  float* s_b2 = s_b1 + BN;
  float* s_gamma = s_b2 + BN;
  float* s_beta = s_gamma + BN;
  float2* s_stat = reinterpret_cast<float2*>(s_beta + BN);  // [2 halves][128 rows] (sum, sum of squares)

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int num_tiles = (p.M + BM - 1) / BM;
  int num_kb = 0;
  for (int s = 0; s < p.n_seg; ++s) num_kb += p.seg_kblocks[s];

  if (warp == 0 && lane == 0) {
    tc::tma_prefetch_desc(&tmap_w1);
    tc::tma_prefetch_desc(&tmap_w2);
    tc::tma_prefetch_desc(&tmap_pre);
    tc::tma_prefetch_desc(&tmap_z);
    tc::tma_prefetch_desc(&tmap_y);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < STAGES; ++s) {
      tc::mbar_init(&full_bar[s], 1 + 128);  // TMA expect_tx arrival + 128 gather threads
      tc::mbar_init(&empty_bar[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      tc::mbar_init(&acc1_full[s], 1);
      tc::mbar_init(&acc1_empty[s], NUM_EPI * 32);
    }
    tc::mbar_init(h_full, NUM_EPI * 32);
    tc::mbar_init(acc2_full, 1);
    tc::mbar_init(w2_full, 1);
    tc::fence_mbar_init();
  }
  if (warp == 2) {
    tc::tmem_alloc(tmem_ptr, TMEM_COLS);
    tc::tmem_relinquish();
  }
  for (int i = threadIdx.x; i < BN; i += blockDim.x) {
    s_b1[i] = p.b1 ? p.b1[i] : 0.f;
    s_b2[i] = p.b2 ? p.b2[i] : 0.f;
    s_gamma[i] = p.gamma[i];
    s_beta[i] = p.beta[i];
  }
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;
  const uint32_t tmem_acc2 = tmem_base + 2 * BN;

  if (warp == 0) {
    // ===================== TMA producer: W2 once, then the W1 k-blocks of every tile =====================
    if (lane == 0) {
      tc::mbar_expect_tx(w2_full, W2_BYTES);
      tc::tma_load_2d(w2_s, &tmap_w2, w2_full, 0, 0);
      tc::tma_load_2d(w2_s + B_BYTES, &tmap_w2, w2_full, BK, 0);
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        for (int kb = 0; kb < num_kb; ++kb) {
          tc::mbar_wait(&empty_bar[stage], phase ^ 1);
          tc::mbar_expect_tx(&full_bar[stage], B_BYTES);
          tc::tma_load_2d(smem + stage * STAGE_BYTES + A_BYTES, &tmap_w1, &full_bar[stage], kb * BK, 0);
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    if (lane == 0) {
      constexpr uint32_t idesc = tc::make_idesc_bf16(BM, BN);
      const uint32_t h_sa = tc::smem_u32(h_s);
      const uint32_t w2_sa = tc::smem_u32(w2_s);
      // GEMM2 of tile number j (per-CTA count): A = H (written by the epilogue), B = W2, D = the third TMEM accumulator
      auto gemm2 = [&](int j) {
        if (j == 0) tc::mbar_wait(w2_full, 0);
        tc::mbar_wait(h_full, (uint32_t)(j & 1));
        tc::tc_fence_after();
#pragma unroll
        for (int kb = 0; kb < BN / BK; ++kb) {
          const uint64_t adesc = tc::make_desc_kmajor_sw128(h_sa + kb * A_BYTES);
          const uint64_t bdesc = tc::make_desc_kmajor_sw128(w2_sa + kb * B_BYTES);
#pragma unroll
          for (int k = 0; k < BK / UK; ++k)
            tc::umma_bf16(tmem_acc2, adesc + (uint64_t)(k * 2), bdesc + (uint64_t)(k * 2), idesc, (kb | k) ? 1u : 0u);
        }
        tc::umma_commit(acc2_full);
      };
      int stage = 0;
      uint32_t phase = 0;
      int it = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
        const int as = it & 1;
        tc::mbar_wait(&acc1_empty[as], ((it >> 1) & 1) ^ 1);
        tc::tc_fence_after();
        const uint32_t tmem_d = tmem_base + as * BN;
        for (int kb = 0; kb < num_kb; ++kb) {
          tc::mbar_wait(&full_bar[stage], phase);
          tc::tc_fence_after();
          const uint32_t sa = tc::smem_u32(smem + stage * STAGE_BYTES);
          const uint64_t adesc = tc::make_desc_kmajor_sw128(sa);
          const uint64_t bdesc = tc::make_desc_kmajor_sw128(sa + A_BYTES);
#pragma unroll
          for (int k = 0; k < BK / UK; ++k)
            tc::umma_bf16(tmem_d, adesc + (uint64_t)(k * 2), bdesc + (uint64_t)(k * 2), idesc, (kb | k) ? 1u : 0u);
          tc::umma_commit(&empty_bar[stage]);
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
        tc::umma_commit(&acc1_full[as]);
        if (it > 0) gemm2(it - 1);  // the previous tile's second GEMM goes behind this tile's first one
      }
      if (it > 0) gemm2(it - 1);
    }
  } else if (warp < GATHER0) {
    // ===================== epilogue =====================
    const int q = warp & 3;                  // TMEM lane quarter this warp may access
    const int half = (warp - EPI0) >> 2;     // which 64 columns (= which k-block of H)
    const int c0 = half * 64;
    uint8_t* my_h = h_s + half * A_BYTES + q * 32 * 128;       // this warp's 4-KB slice of H (rows q*32.., k-block `half`)
    uint8_t* my_st = store_s + (warp - EPI0) * 4096;           // this warp's private staging tile
    int it = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
      const int m0 = tile * BM;
      const int as = it & 1;
      const int row = m0 + q * 32 + lane;
      const bool row_ok = row < p.M;
      float v[64];
      uint32_t* r = reinterpret_cast<uint32_t*>(v);

      // ---------- phase 1: pre = acc1 + b1 (saved for the backward), h = SiLU(pre) -> H ----------
      tc::mbar_wait(&acc1_full[as], (uint32_t)((it >> 1) & 1));
      tc::tc_fence_after();
      {
        const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + as * BN + c0;
        tc::tmem_ld_32x32(taddr, r);
        tc::tmem_ld_32x32(taddr + 32, r + 32);
        tc::tmem_ld_wait();
      }
      tc::tc_fence_before();
      tc::mbar_arrive(&acc1_empty[as]);  // the accumulator is in registers: GEMM1 of tile it + 2 may overwrite it
#pragma unroll
      for (int t = 0; t < 64; ++t) {
        // round to bf16 here: the backward recomputes h = SiLU(pre) from the stored bf16 `pre`
        v[t] = __bfloat162float(__float2bfloat16_rn(__uint_as_float(r[t]) + s_b1[c0 + t]));
      }
      staging_reusable(lane);  // y of the previous tile (staging tile) and z of the previous tile (H slice) have been read out
      stage_rows(my_st, lane, v);
      tc::fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) {
        tc::tma_store_2d(&tmap_pre, my_st, c0, m0 + q * 32);
        tc::tma_store_commit();
      }
#pragma unroll
      for (int t = 0; t < 64; ++t) v[t] = silu_f(v[t]);
      stage_rows(my_h, lane, v);
      tc::fence_proxy_async_smem();  // generic-proxy writes of H -> visible to tcgen05.mma (async proxy)
      tc::mbar_arrive(h_full);

      // ---------- phase 2: z = acc2 + b2 (saved), row statistics, y = LN(z) * gamma + beta (+ residual) ----------
      tc::mbar_wait(acc2_full, (uint32_t)(it & 1));
      tc::tc_fence_after();
      {
        const uint32_t taddr = tmem_acc2 + ((uint32_t)(q * 32) << 16) + c0;
        tc::tmem_ld_32x32(taddr, r);
        tc::tmem_ld_32x32(taddr + 32, r + 32);
        tc::tmem_ld_wait();
      }
      tc::tc_fence_before();  // orders these TMEM reads before the h_full arrival of the next tile (-> next GEMM2)
      float sum = 0.f, sq = 0.f;
#pragma unroll
      for (int t = 0; t < 64; ++t) {
        v[t] = __bfloat162float(__float2bfloat16_rn(__uint_as_float(r[t]) + s_b2[c0 + t]));  // statistics of the stored z
        sum += v[t];
        sq = fmaf(v[t], v[t], sq);
      }
      s_stat[half * BM + q * 32 + lane] = make_float2(sum, sq);
      // GEMM2 of this tile has completed (acc2_full), so nothing reads H any more: the slice doubles as z's staging tile
      stage_rows(my_h, lane, v);
      tc::fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) {
        tc::tma_store_2d(&tmap_z, my_h, c0, m0 + q * 32);
        tc::tma_store_commit();
      }
      epi_bar_sync();  // both halves of every row have published their partial sums
      const float2 other = s_stat[(half ^ 1) * BM + q * 32 + lane];
      const float mean = (sum + other.x) * (1.f / BN);
      const float var = fmaxf((sq + other.y) * (1.f / BN) - mean * mean, 0.f);
      const float rstd = rsqrtf(var + p.eps);
      if (half == 0 && row_ok) {
        p.mean[row] = mean;
        p.rstd[row] = rstd;
      }
#pragma unroll
      for (int t = 0; t < 64; ++t) v[t] = (v[t] - mean) * rstd * s_gamma[c0 + t] + s_beta[c0 + t];
      if (p.residual != nullptr && row_ok) {
        const __nv_bfloat16* rr = p.residual + (int64_t)row * p.ld_res + c0;
#pragma unroll
        for (int t = 0; t < 8; ++t) {
          float f[8];
          unpack8(ldg_nc_v4(rr + t * 8), f);
#pragma unroll
          for (int u = 0; u < 8; ++u) v[t * 8 + u] += f[u];
        }
      }
      staging_reusable(lane);  // the store of `pre` has read the staging tile
      stage_rows(my_st, lane, v);
      tc::fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) {
        tc::tma_store_2d(&tmap_y, my_st, c0, m0 + q * 32);
        tc::tma_store_commit();
      }
    }
    if (lane == 0) tc::tma_store_wait_all();
  } else {
    // ===================== gather producers (4 warps): A rows -> swizzled smem =====================
    const int gt = threadIdx.x - GATHER0 * 32;  // 0..127
    const int chunk = gt & 7;                   // 16-byte chunk inside the 128-byte k-block row
    const int r0 = gt >> 3;                     // rows r0 + 16 * i, i = 0..7
    int stage = 0;
    uint32_t phase = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      const int m0 = tile * BM;
      for (int s = 0; s < p.n_seg; ++s) {
        const __nv_bfloat16* rowptr[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const int m = m0 + r0 + 16 * i;
          rowptr[i] = nullptr;
          if (m < p.M) {
            const int j = p.seg_idx[s] ? __ldg(p.seg_idx[s] + m) : m;
            rowptr[i] = p.seg_base[s] + (int64_t)j * p.seg_ld[s];
          }
        }
        for (int kb = 0; kb < p.seg_kblocks[s]; ++kb) {
          tc::mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = smem + stage * STAGE_BYTES;
          const int k = kb * BK + chunk * 8;
          uint4 v[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            v[i] = make_uint4(0, 0, 0, 0);
            if (rowptr[i] != nullptr && k < p.seg_k[s]) v[i] = ldg_nc_v4(rowptr[i] + k);
          }
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const int rr = r0 + 16 * i;
            *reinterpret_cast<uint4*>(sa + rr * 128 + ((chunk ^ (rr & 7)) << 4)) = v[i];
          }
          tc::fence_proxy_async_smem();
          tc::mbar_arrive(&full_bar[stage]);
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  }

  tc::tc_fence_before();
  __syncthreads();
  if (warp == 2) tc::tmem_dealloc(tmem_base, TMEM_COLS);
}

}  // namespace mlpf
}  // namespace dgb

using namespace dgb;

// y = LayerNorm(SiLU(cat_s(A_s[idx_s]) @ W1^T + b1) @ W2^T + b2) * gamma + beta (+ residual), hidden = out = 128.
//   W1p  bf16 [128, sum_s pad64(K_s)] (segment after segment, each zero-padded to a multiple of 64 columns; the layout of
//        dgb_gather_gemm_bf16), W2 bf16 [128, 128]; b1 / b2 / gamma / beta fp32 [128] (b1, b2 may be null).
//   pre, z, y bf16 [M, 128] (row strides multiples of 8, 16-byte aligned bases); mean, rstd fp32 [M].
// Returns 801 when the shape / alignment is outside what this kernel handles (the caller then runs the unfused chain).
DGB_API int dgb_mlp2_ln_fused_bf16(int n_seg, const void* const* seg_base, const int64_t* seg_ld, const int* const* seg_idx,
                                   const int* seg_k, const void* W1p, int64_t ldw1, const float* b1, const void* W2,
                                   int64_t ldw2, const float* b2, const float* gamma, const float* beta, float eps,
                                   const void* residual, int64_t ld_res, void* pre, int64_t ld_pre, void* z, int64_t ld_z,
                                   void* y, int64_t ld_y, float* mean, float* rstd, int M, void* stream) {
  using namespace dgb::mlpf;
  if (M <= 0) return 0;
  if (n_seg < 1 || n_seg > 3 || !gamma || !beta || !mean || !rstd) return (int)cudaErrorInvalidValue;
  auto misaligned = [](const void* ptr, int64_t ld) { return (reinterpret_cast<uintptr_t>(ptr) & 15) != 0 || (ld % 8) != 0; };
  if (misaligned(W1p, ldw1) || misaligned(W2, ldw2) || misaligned(pre, ld_pre) || misaligned(z, ld_z) || misaligned(y, ld_y))
    return 801;
  if (residual && misaligned(residual, ld_res)) return 801;
  Params p{};
  p.M = M; p.n_seg = n_seg;
  int ktot = 0;
  for (int s = 0; s < n_seg; ++s) {
    if (misaligned(seg_base[s], seg_ld[s]) || (seg_k[s] % 8) || seg_k[s] <= 0) return 801;
    p.seg_base[s] = (const __nv_bfloat16*)seg_base[s];
    p.seg_ld[s] = seg_ld[s];
    p.seg_idx[s] = seg_idx ? seg_idx[s] : nullptr;
    p.seg_k[s] = seg_k[s];
    p.seg_kblocks[s] = (seg_k[s] + BK - 1) / BK;
    ktot += p.seg_kblocks[s] * BK;
  }
  p.b1 = b1; p.b2 = b2; p.gamma = gamma; p.beta = beta; p.eps = eps;
  p.residual = (const __nv_bfloat16*)residual; p.ld_res = ld_res;
  p.mean = mean; p.rstd = rstd;
  CUtensorMap t_w1, t_w2, t_pre, t_z, t_y;
  int rc = make_tmap_bf16_2d(&t_w1, W1p, BN, (uint64_t)ktot, (uint64_t)ldw1, BN, BK);
  if (rc) return rc;
  rc = make_tmap_bf16_2d(&t_w2, W2, BN, BN, (uint64_t)ldw2, BN, BK);
  if (rc) return rc;
  rc = make_tmap_bf16_2d(&t_pre, pre, (uint64_t)M, BN, (uint64_t)ld_pre, 32, 64);
  if (rc) return rc;
  rc = make_tmap_bf16_2d(&t_z, z, (uint64_t)M, BN, (uint64_t)ld_z, 32, 64);
  if (rc) return rc;
  rc = make_tmap_bf16_2d(&t_y, y, (uint64_t)M, BN, (uint64_t)ld_y, 32, 64);
  if (rc) return rc;
  DGB_OPT_IN_SMEM(mlp2_ln_fused_kernel, SMEM_TOTAL);
  int dev = 0, sms = 0;
  DGB_CUDA(cudaGetDevice(&dev));
  DGB_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  const int tiles = (M + BM - 1) / BM;
  const int grid = tiles < sms ? tiles : sms;
  mlp2_ln_fused_kernel<<<grid, THREADS, SMEM_TOTAL, (cudaStream_t)stream>>>(t_w1, t_w2, t_pre, t_z, t_y, p);
  DGB_CHECK_LAUNCH();
  return 0;
}
