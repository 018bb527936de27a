// The Python layer takes this kernel for K <= 128 (DGRAPH_B200_GEMM_BRES=auto; 0 = never, 1 = every supported shape); measured on
// B200 at 2.45 M x 256 x 104: 0.473 -> 0.322 ms (cuBLAS 0.307 ms), numerics in tests/test_gemm_bres_gpu.py.  The streaming kernel
// for larger K is csrc/gemm_tcgen05.cu.
//
// B-resident variant of the persistent tcgen05 GEMM for small K:  C[M,N] = act(A[M,K] @ B[N,K]^T + bias), bf16 in / bf16 out.
//
// Why: for the K <= 128 projections of the GCN / GraphCast models the streaming kernel reaches only 54-59 % of the HBM roofline
// (profiles/README.md section 3).  Per 128-row tile it loads 32 KB of A but re-loads the whole B operand (64 KB at N = 256) from
// L2, and with three pipeline stages only 48 KB of A per SM are in flight -- less than the ~65 KB per SM that 6.4 TB/s x 1.5 us
// of HBM latency asks for.  Here every CTA loads its N-tile of B ONCE into shared memory (all k-blocks, 128-B swizzled), and
// the ring buffer holds A k-blocks only: 5-8 stages of 16 KB.  With several N-tiles a CTA keeps one of them for its whole life
// (tile order: CTA b owns N-tile b % tiles_n and walks M-tiles b / tiles_n, + gridDim / tiles_n, ...).
//
// Everything else -- roles, TMEM double buffering, epilogue through swizzled staging tiles and TMA stores -- is the scheme of
// gemm_tcgen05.cu.  The reference has no tensor-core code.
#include "tc_common.cuh"

namespace dgb {
namespace bres {

constexpr int BLOCK_M = 128;
constexpr int BLOCK_K = 64;
constexpr int UMMA_K = 16;
constexpr int MAX_N = 2048;
constexpr int MAX_STAGES = 8;
constexpr int NUM_EPI_WARPS = 8;
constexpr int EPI_WARP0 = 2;
constexpr int THREADS = (EPI_WARP0 + NUM_EPI_WARPS) * 32;
constexpr int A_BYTES = BLOCK_M * BLOCK_K * 2;        // 16 KB per k-block of A
constexpr int STORE_BYTES = NUM_EPI_WARPS * 2 * 4096;  // two 4-KB staging tiles per epilogue warp
constexpr int BAR_BYTES = 256 + MAX_N * 4;             // mbarriers + tmem pointer, then the staged bias
constexpr int SMEM_LIMIT = 232448;                     // 227 KB opt-in maximum per CTA

enum Act { ACT_NONE = 0, ACT_RELU = 1, ACT_SILU = 2 };

struct Params {
  int M, N, K;
  const float* bias;
  int num_kb;   // k-blocks of 64
  int stages;   // A ring depth chosen by the host (4 .. MAX_STAGES)
};

template <int ACT>
__device__ __forceinline__ float apply_act(float v) {
  if (ACT == ACT_RELU) return fmaxf(v, 0.f);
  if (ACT == ACT_SILU) return v / (1.f + __expf(-v));
  return v;
}

template <int BLOCK_N, int ACT>
__global__ void __launch_bounds__(THREADS, 1)
gemm_bres_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                 const __grid_constant__ CUtensorMap tmap_c, const Params p) {
  constexpr int B_BYTES = BLOCK_N * BLOCK_K * 2;  // one k-block of this CTA's N-tile
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int nst = p.stages;
  uint8_t* a_ring = smem;                                  // nst x 16 KB
  uint8_t* b_res = a_ring + nst * A_BYTES;                 // num_kb x B_BYTES, resident for the CTA's lifetime
  uint8_t* store_base = b_res + p.num_kb * B_BYTES;        // 1024-B aligned: every block above is a multiple of 8 KB
  uint8_t* bar_base = store_base + STORE_BYTES;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(bar_base);
  uint64_t* empty_bar = full_bar + MAX_STAGES;
  uint64_t* tmem_full = empty_bar + MAX_STAGES;
  uint64_t* tmem_empty = tmem_full + 2;
  uint64_t* b_full = tmem_empty + 2;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(b_full + 1);
  float* s_bias = reinterpret_cast<float*>(bar_base + 256);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int tiles_m = (p.M + BLOCK_M - 1) / BLOCK_M;
  const int tiles_n = (p.N + BLOCK_N - 1) / BLOCK_N;
  // this CTA's fixed N-tile and its walk over M-tiles (host guarantees gridDim.x % tiles_n == 0)
  const int nt = blockIdx.x % tiles_n;
  const int mt0 = blockIdx.x / tiles_n;
  const int mstep = gridDim.x / tiles_n;
  const int n0 = nt * BLOCK_N;

  if (warp == 0 && lane == 0) {
    tc::tma_prefetch_desc(&tmap_a);
    tc::tma_prefetch_desc(&tmap_b);
    tc::tma_prefetch_desc(&tmap_c);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < MAX_STAGES; ++s) {
      tc::mbar_init(&full_bar[s], 1);
      tc::mbar_init(&empty_bar[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      tc::mbar_init(&tmem_full[s], 1);
      tc::mbar_init(&tmem_empty[s], NUM_EPI_WARPS * 32);
    }
    tc::mbar_init(b_full, 1);
    tc::fence_mbar_init();
  }
  if (warp == 2) {
    tc::tmem_alloc(tmem_ptr, 2 * BLOCK_N);
    tc::tmem_relinquish();
  }
  for (int i = threadIdx.x; i < BLOCK_N; i += blockDim.x) s_bias[i] = (p.bias && n0 + i < p.N) ? p.bias[n0 + i] : 0.f;
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;

  if (warp == 0) {
    // ===================== TMA producer: B once, then A k-blocks through the ring =====================
    if (lane == 0) {
      tc::mbar_expect_tx(b_full, (uint32_t)(p.num_kb * B_BYTES));
      for (int kb = 0; kb < p.num_kb; ++kb) tc::tma_load_2d(b_res + kb * B_BYTES, &tmap_b, b_full, kb * BLOCK_K, n0);
      int stage = 0;
      uint32_t phase = 0;
      for (int mt = mt0; mt < tiles_m; mt += mstep) {
        const int m0 = mt * BLOCK_M;
        for (int kb = 0; kb < p.num_kb; ++kb) {
          tc::mbar_wait(&empty_bar[stage], phase ^ 1);
          tc::mbar_expect_tx(&full_bar[stage], A_BYTES);
          tc::tma_load_2d(a_ring + stage * A_BYTES, &tmap_a, &full_bar[stage], kb * BLOCK_K, m0);
          if (++stage == nst) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    if (lane == 0) {
      constexpr uint32_t idesc = tc::make_idesc_bf16(BLOCK_M, BLOCK_N);
      tc::mbar_wait(b_full, 0);
      tc::tc_fence_after();
      int stage = 0;
      uint32_t phase = 0;
      int it = 0;
      for (int mt = mt0; mt < tiles_m; mt += mstep, ++it) {
        const int as = it & 1;
        tc::mbar_wait(&tmem_empty[as], ((it >> 1) & 1) ^ 1);
        tc::tc_fence_after();
        const uint32_t tmem_d = tmem_base + as * BLOCK_N;
        for (int kb = 0; kb < p.num_kb; ++kb) {
          tc::mbar_wait(&full_bar[stage], phase);
          tc::tc_fence_after();
          const uint64_t adesc = tc::make_desc_kmajor_sw128(tc::smem_u32(a_ring + stage * A_BYTES));
          const uint64_t bdesc = tc::make_desc_kmajor_sw128(tc::smem_u32(b_res + kb * B_BYTES));
#pragma unroll
          for (int k = 0; k < BLOCK_K / UMMA_K; ++k)
            tc::umma_bf16(tmem_d, adesc + (uint64_t)(k * 2), bdesc + (uint64_t)(k * 2), idesc, (kb | k) ? 1u : 0u);
          tc::umma_commit(&empty_bar[stage]);
          if (++stage == nst) { stage = 0; phase ^= 1; }
        }
        tc::umma_commit(&tmem_full[as]);
      }
    }
  } else {
    // ===================== epilogue: TMEM -> regs -> +bias -> act -> bf16 -> swizzled staging -> TMA store =====================
    const int q = warp & 3;
    const int half = (warp - EPI_WARP0) >> 2;
    constexpr int COLS = BLOCK_N / 2;
    constexpr int CH = 64;  // 64 bf16 = one 128-byte swizzled store row
    int it = 0;
    int sbuf = 0;
    for (int mt = mt0; mt < tiles_m; mt += mstep, ++it) {
      const int m0 = mt * BLOCK_M;
      const int as = it & 1;
      tc::mbar_wait(&tmem_full[as], (it >> 1) & 1);
      tc::tc_fence_after();
      const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + as * BLOCK_N + half * COLS;
      const uint32_t bias_sa = tc::smem_u32(s_bias + half * COLS);
#pragma unroll 1
      for (int c = 0; c < COLS; c += CH) {
        if (p.N - (n0 + half * COLS + c) <= 0) break;  // warp-uniform: no valid column left in this chunk
        uint32_t r[CH];
        tc::tmem_ld_32x32(taddr + c, r);
        tc::tmem_ld_32x32(taddr + c + 32, r + 32);
        tc::tmem_ld_wait();
        float v[CH];
#pragma unroll
        for (int t = 0; t < CH; t += 4) {
          float4 b4;
          asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];"
                       : "=f"(b4.x), "=f"(b4.y), "=f"(b4.z), "=f"(b4.w)
                       : "r"(bias_sa + (uint32_t)((c + t) * 4)));
          v[t] = apply_act<ACT>(__uint_as_float(r[t]) + b4.x);
          v[t + 1] = apply_act<ACT>(__uint_as_float(r[t + 1]) + b4.y);
          v[t + 2] = apply_act<ACT>(__uint_as_float(r[t + 2]) + b4.z);
          v[t + 3] = apply_act<ACT>(__uint_as_float(r[t + 3]) + b4.w);
        }
        uint8_t* st = store_base + ((warp - EPI_WARP0) * 2 + sbuf) * 4096;
        if (lane == 0) tc::tma_store_wait_read1();  // the staging tile issued two chunks ago has been read out
        __syncwarp();
#pragma unroll
        for (int t = 0; t < 8; ++t) {
          const uint4 w = pack8(v + t * 8);
          *reinterpret_cast<uint4*>(st + lane * 128 + ((t ^ (lane & 7)) << 4)) = w;
        }
        tc::fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) {
          tc::tma_store_2d(&tmap_c, st, n0 + half * COLS + c, m0 + q * 32);  // rows >= M / columns >= N are clipped
          tc::tma_store_commit();
        }
        sbuf ^= 1;
      }
      tc::tc_fence_before();
      tc::mbar_arrive(&tmem_empty[as]);
    }
    if (lane == 0) tc::tma_store_wait_all();
  }

  tc::tc_fence_before();
  __syncthreads();
  if (warp == 2) tc::tmem_dealloc(tmem_base, 2 * BLOCK_N);
}

template <int BLOCK_N, int ACT>
static int launch(const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& tcm, const Params& p, int smem_bytes,
                  cudaStream_t stream) {
  auto kern = gemm_bres_kernel<BLOCK_N, ACT>;
  DGB_OPT_IN_SMEM(kern, SMEM_LIMIT);
  int dev = 0, sms = 0;
  DGB_CUDA(cudaGetDevice(&dev));
  DGB_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  const int tiles_m = (p.M + BLOCK_M - 1) / BLOCK_M;
  const int tiles_n = (p.N + BLOCK_N - 1) / BLOCK_N;
  int per_n = sms / tiles_n;                 // CTAs that share one N-tile
  if (per_n > tiles_m) per_n = tiles_m;
  if (per_n < 1) return (int)cudaErrorNotSupported;
  const int grid = per_n * tiles_n;          // a multiple of tiles_n, as the kernel's tile walk requires
  kern<<<grid, THREADS, smem_bytes, stream>>>(ta, tb, tcm, p);
  DGB_CHECK_LAUNCH();
  return 0;
}

template <int BLOCK_N>
static int dispatch(const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& tcm, const Params& p, int act, int smem,
                    cudaStream_t s) {
  if (act == ACT_NONE) return launch<BLOCK_N, ACT_NONE>(ta, tb, tcm, p, smem, s);
  if (act == ACT_RELU) return launch<BLOCK_N, ACT_RELU>(ta, tb, tcm, p, smem, s);
  if (act == ACT_SILU) return launch<BLOCK_N, ACT_SILU>(ta, tb, tcm, p, smem, s);
  return (int)cudaErrorInvalidValue;
}

}  // namespace bres
}  // namespace dgb

// C[M,N] (bf16, 16-byte aligned rows) = act(A[M,K] @ B[N,K]^T + bias) with B resident in shared memory.  Returns
// cudaErrorNotSupported (801) when the shape does not fit this variant (B tile + at least four A stages must fit in 227 KB, N tile
// of 128 or 256, aligned C) -- the caller then uses dgb_gemm_bf16_tn.
DGB_API int dgb_gemm_bf16_tn_bres(const void* A, int64_t lda, const void* B, int64_t ldb, void* C, int64_t ldc, int M, int N,
                                  int K, const float* bias, int act, void* stream) {
  using namespace dgb;
  using namespace dgb::bres;
  if (M <= 0 || N <= 0 || K <= 0) return M == 0 || N == 0 ? 0 : (int)cudaErrorInvalidValue;
  if ((lda % 8) || (ldb % 8) || (ldc % 8) || (reinterpret_cast<uintptr_t>(A) & 15) || (reinterpret_cast<uintptr_t>(B) & 15) ||
      (reinterpret_cast<uintptr_t>(C) & 15))
    return (int)cudaErrorNotSupported;
  if (N <= 64 || N > MAX_N) return (int)cudaErrorNotSupported;
  const int bn = N <= 128 ? 128 : 256;
  const int num_kb = (K + BLOCK_K - 1) / BLOCK_K;
  const int b_bytes = num_kb * bn * BLOCK_K * 2;
  const int fixed = b_bytes + STORE_BYTES + BAR_BYTES + 1024;
  int stages = (SMEM_LIMIT - fixed) / A_BYTES;
  if (stages > MAX_STAGES) stages = MAX_STAGES;
  if (stages < 4) return (int)cudaErrorNotSupported;
  CUtensorMap ta, tb, tcm;
  int rc = make_tmap_bf16_2d(&ta, A, (uint64_t)M, (uint64_t)K, (uint64_t)lda, BLOCK_M, BLOCK_K);
  if (rc) return rc;
  rc = make_tmap_bf16_2d(&tb, B, (uint64_t)N, (uint64_t)K, (uint64_t)ldb, bn, BLOCK_K);
  if (rc) return rc;
  rc = make_tmap_2d(&tcm, C, (uint64_t)M, (uint64_t)N, (uint64_t)ldc, 32, 64, 2, CU_TENSOR_MAP_SWIZZLE_128B);
  if (rc) return rc;
  Params p{};
  p.M = M; p.N = N; p.K = K; p.bias = bias; p.num_kb = num_kb; p.stages = stages;
  const int smem = stages * A_BYTES + fixed;
  if (bn == 128) return dispatch<128>(ta, tb, tcm, p, act, smem, (cudaStream_t)stream);
  return dispatch<256>(ta, tb, tcm, p, act, smem, (cudaStream_t)stream);
}
