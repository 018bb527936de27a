// Persistent warp-specialised tcgen05 GEMM:  C[M,N] = act(A[M,K] @ B[N,K]^T + bias)   (bf16 in, fp32 TMEM accumulate)
//
// This is the node-/edge-level Linear of every model in the framework (GCN P/Q projections, SAGE, RGAT, GraphCast
// MLPs).  The reference reaches cuBLAS sgemm through nn.Linear (experiments/OGB/GCN.py:51-52,
// experiments/GraphCast/layers.py:52-63); it has no tensor-core code at all.
//
// Roles (320 threads): warp 0 = TMA producer (A and B k-blocks, 128-B swizzle), warp 1 = MMA issuer (one elected
// lane, tcgen05.mma cta_group::1, M=128 x N=BLOCK_N x K=16), warps 2..9 = epilogue (tcgen05.ld 32x32b -> smem-staged
// bias -> compile-time activation -> bf16/fp32 -> swizzled 32-row staging tile -> TMA store, double-buffered per warp;
// two warps per TMEM lane quarter; direct register stores only when C is not 16-byte aligned).  Two TMEM accumulator stages let the epilogue of tile i overlap
// the main loop of tile i+1; the smem ring has STAGES k-blocks in flight.  The gathered variant replaces the TMA
// load of A by four producer warps that gather rows (local HBM or NVLink peer memory) into the same swizzled layout.
#include "tc_common.cuh"
#include <type_traits>

namespace dgb {

constexpr int BLOCK_M = 128;
constexpr int BLOCK_K = 64;  // 64 bf16 = 128 B = one swizzle row
constexpr int UMMA_K = 16;
constexpr int MAX_N = 2048;  // bias for all N-tiles is staged in shared memory once per CTA

enum Act { ACT_NONE = 0, ACT_RELU = 1, ACT_SILU = 2 };

struct GemmParams {
  int M, N, K;
  const float* bias;  // [N] fp32 or null
  void* C;
  int64_t ldc;
  int act;
  // gathered-A segments (GATHER variant): A row m, k-range of segment s = seg_base[s][ idx_s[m] * ld_s + k ]
  int n_seg;
  const __nv_bfloat16* seg_base[3];
  const int* seg_idx[3];  // null = identity
  int64_t seg_ld[3];
  int seg_kblocks[3];  // k-blocks (of 64) per segment; weights are laid out segment after segment, each padded to 64
  int seg_k[3];        // true K of the segment (columns beyond are zero-filled)
  // optional peer addressing per segment: gathered row j >= seg_nlocal[s] lives on another GPU, in the same matrix at
  // peer_base[halo_owner[j - nlocal]] + halo_row[j - nlocal] * ld  (NVLink pull inside the GEMM's producer warps)
  const __nv_bfloat16* const* seg_peer[3];
  const int* seg_halo_owner[3];
  const int* seg_halo_row[3];
  int64_t seg_nlocal[3];
};

template <int BLOCK_N>
struct SmemLayout {
  static constexpr int A_BYTES = BLOCK_M * BLOCK_K * 2;  // 16 KB
  static constexpr int B_BYTES = BLOCK_N * BLOCK_K * 2;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int STAGES = (BLOCK_N >= 256) ? 3 : (BLOCK_N >= 128 ? 4 : 6);
  static constexpr int BAR_BYTES = 256 + MAX_N * 4;  // mbarriers + tmem pointer, then the staged bias [MAX_N]
  static constexpr int STORE_BYTES = 8 * 2 * 4096;   // per epilogue warp: two 4-KB staging tiles for TMA stores
  static constexpr int TOTAL = STAGES * STAGE_BYTES + STORE_BYTES + BAR_BYTES + 1024;  // + alignment slack
};

template <int ACT>
__device__ __forceinline__ float apply_act(float v) {
  if (ACT == ACT_RELU) return fmaxf(v, 0.f);
  if (ACT == ACT_SILU) return v / (1.f + __expf(-v));
  return v;
}

constexpr int NUM_EPI_WARPS = 8;  // two warps per TMEM lane quarter, each owns half of the tile's columns
constexpr int EPI_WARP0 = 2;
constexpr int GATHER_WARP0 = EPI_WARP0 + NUM_EPI_WARPS;  // warps 10..13 when GATHER
constexpr int THREADS_DENSE = (EPI_WARP0 + NUM_EPI_WARPS) * 32;
constexpr int THREADS_GATHER = THREADS_DENSE + 128;

template <int BLOCK_N, bool OUT_F32, bool GATHER, int ACT, bool STORE_TMA>
__global__ void __launch_bounds__(GATHER ? THREADS_GATHER : THREADS_DENSE, 1)
gemm_bf16_tn_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                    const __grid_constant__ CUtensorMap tmap_c, const GemmParams p) {
  using L = SmemLayout<BLOCK_N>;
  constexpr int STAGES = L::STAGES;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* store_base = smem + STAGES * L::STAGE_BYTES;  // 1024-B aligned (stage sizes are multiples of 1024)
  uint8_t* bar_base = store_base + L::STORE_BYTES;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(bar_base);
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tmem_full = empty_bar + STAGES;
  uint64_t* tmem_empty = tmem_full + 2;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(tmem_empty + 2);
  float* s_bias = reinterpret_cast<float*>(bar_base + 256);  // [tiles_n * BLOCK_N] fp32, zero beyond N

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int tiles_m = (p.M + BLOCK_M - 1) / BLOCK_M;
  const int tiles_n = (p.N + BLOCK_N - 1) / BLOCK_N;
  const int num_tiles = tiles_m * tiles_n;
  int num_kb = 0;
  if (GATHER) {
    for (int s = 0; s < p.n_seg; ++s) num_kb += p.seg_kblocks[s];
  } else {
    num_kb = (p.K + BLOCK_K - 1) / BLOCK_K;
  }

  if (warp == 0 && lane == 0) {
    if (!GATHER) tc::tma_prefetch_desc(&tmap_a);
    tc::tma_prefetch_desc(&tmap_b);
    if (STORE_TMA) tc::tma_prefetch_desc(&tmap_c);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < STAGES; ++s) {
      tc::mbar_init(&full_bar[s], GATHER ? (1 + 128) : 1);
      tc::mbar_init(&empty_bar[s], 1);
    }
    for (int s = 0; s < 2; ++s) {
      tc::mbar_init(&tmem_full[s], 1);
      tc::mbar_init(&tmem_empty[s], NUM_EPI_WARPS * 32);
    }
    tc::fence_mbar_init();
  }
  if (warp == 2) {
    tc::tmem_alloc(tmem_ptr, 2 * BLOCK_N);
    tc::tmem_relinquish();
  }
  // bias of every N-tile (N <= MAX_N is enforced by the host), zero padded up to the last tile's end
  for (int i = threadIdx.x; i < tiles_n * BLOCK_N; i += blockDim.x) s_bias[i] = (p.bias && i < p.N) ? p.bias[i] : 0.f;
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
        const int m0 = (tile / tiles_n) * BLOCK_M;
        const int n0 = (tile % tiles_n) * BLOCK_N;
        for (int kb = 0; kb < num_kb; ++kb) {
          tc::mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = smem + stage * L::STAGE_BYTES;
          uint8_t* sb = sa + L::A_BYTES;
          if (GATHER) {
            tc::mbar_expect_tx(&full_bar[stage], L::B_BYTES);
          } else {
            tc::mbar_expect_tx(&full_bar[stage], L::STAGE_BYTES);
            tc::tma_load_2d(sa, &tmap_a, &full_bar[stage], kb * BLOCK_K, m0);
          }
          tc::tma_load_2d(sb, &tmap_b, &full_bar[stage], kb * BLOCK_K, n0);
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    if (lane == 0) {
      constexpr uint32_t idesc = tc::make_idesc_bf16(BLOCK_M, BLOCK_N);
      int stage = 0;
      uint32_t phase = 0;
      int it = 0;
      for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
        const int as = it & 1;
        tc::mbar_wait(&tmem_empty[as], ((it >> 1) & 1) ^ 1);
        tc::tc_fence_after();
        const uint32_t tmem_d = tmem_base + as * BLOCK_N;
        for (int kb = 0; kb < num_kb; ++kb) {
          tc::mbar_wait(&full_bar[stage], phase);
          tc::tc_fence_after();
          const uint32_t sa = tc::smem_u32(smem + stage * L::STAGE_BYTES);
          const uint32_t sb = sa + L::A_BYTES;
          const uint64_t adesc = tc::make_desc_kmajor_sw128(sa);
          const uint64_t bdesc = tc::make_desc_kmajor_sw128(sb);
#pragma unroll
          for (int k = 0; k < BLOCK_K / UMMA_K; ++k) {
            // advance 16 bf16 = 32 B inside the 128-B swizzle row: +2 in the (addr >> 4) field
            tc::umma_bf16(tmem_d, adesc + (uint64_t)(k * 2), bdesc + (uint64_t)(k * 2), idesc, (kb | k) ? 1u : 0u);
          }
          tc::umma_commit(&empty_bar[stage]);
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
        tc::umma_commit(&tmem_full[as]);
      }
    }
  } else if (warp < EPI_WARP0 + NUM_EPI_WARPS) {
    // ===================== epilogue: TMEM -> regs -> +bias -> act -> bf16/fp32 -> global =====================
    const int q = warp & 3;                      // TMEM lane quarter this warp may access
    const int half = (warp - EPI_WARP0) >> 2;    // which half of the tile's columns
    constexpr int COLS = BLOCK_N / 2;            // columns per epilogue warp
    constexpr int CH = OUT_F32 ? 32 : ((COLS >= 64) ? 64 : 32);  // columns per TMEM round trip (= one TMA store box)
    using OutT = typename std::conditional<OUT_F32, float, __nv_bfloat16>::type;
    int it = 0;
    int sbuf = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
      const int m0 = (tile / tiles_n) * BLOCK_M;
      const int nt = tile % tiles_n;
      const int n0 = nt * BLOCK_N;
      const int as = it & 1;
      tc::mbar_wait(&tmem_full[as], (it >> 1) & 1);
      tc::tc_fence_after();
      const int row = m0 + q * 32 + lane;
      const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + as * BLOCK_N + half * COLS;
      const uint32_t bias_sa = tc::smem_u32(s_bias + nt * BLOCK_N + half * COLS);
      OutT* crow = reinterpret_cast<OutT*>(p.C) + (int64_t)row * p.ldc + n0 + half * COLS;
      const bool row_ok = row < p.M;
      const bool vec_ok = (reinterpret_cast<uintptr_t>(crow) & 15) == 0;
#pragma unroll 1
      for (int c = 0; c < COLS; c += CH) {
        const int ncol = p.N - (n0 + half * COLS + c);  // valid columns from here (may be <= 0 or > CH); warp-uniform
        if (ncol <= 0) break;
        uint32_t r[CH];
        tc::tmem_ld_32x32(taddr + c, r);
        if (CH == 64) tc::tmem_ld_32x32(taddr + c + 32, r + 32);
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
        if (STORE_TMA) {
          // stage the 32 x CH tile of this warp in shared memory (swizzled, conflict-free) and let TMA write full lines;
          // rows >= M and columns >= N are clipped by the tensor map
          constexpr int ROW_B = CH * (OUT_F32 ? 4 : 2);  // 128 bytes (SWIZZLE_128B box) or 64 bytes (unswizzled box)
          uint8_t* st = store_base + ((warp - EPI_WARP0) * 2 + sbuf) * 4096;
          if (lane == 0) tc::tma_store_wait_read1();  // the tile issued two chunks ago has been read out
          __syncwarp();
#pragma unroll
          for (int t = 0; t < ROW_B / 16; ++t) {
            uint4 w;
            if (OUT_F32) {
              const float* f = v + t * 4;
              w = make_uint4(__float_as_uint(f[0]), __float_as_uint(f[1]), __float_as_uint(f[2]), __float_as_uint(f[3]));
            } else {
              w = pack8(v + t * 8);
            }
            const int off = (ROW_B == 128) ? (lane * 128 + ((t ^ (lane & 7)) << 4)) : (lane * 64 + (t << 4));
            *reinterpret_cast<uint4*>(st + off) = w;
          }
          tc::fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) {
            tc::tma_store_2d(&tmap_c, st, n0 + half * COLS + c, m0 + q * 32);
            tc::tma_store_commit();
          }
          sbuf ^= 1;
        } else if (row_ok) {
          if (ncol >= CH && vec_ok) {
            if (OUT_F32) {
#pragma unroll
              for (int t = 0; t < CH; t += 4)
                *reinterpret_cast<float4*>(reinterpret_cast<float*>(crow) + c + t) =
                    make_float4(v[t], v[t + 1], v[t + 2], v[t + 3]);
            } else {
#pragma unroll
              for (int t = 0; t < CH; t += 8) stg_v4(reinterpret_cast<__nv_bfloat16*>(crow) + c + t, pack8(v + t));
            }
          } else {
#pragma unroll
            for (int t = 0; t < CH; ++t) {
              if (t < ncol) {
                if (OUT_F32) reinterpret_cast<float*>(crow)[c + t] = v[t];
                else reinterpret_cast<__nv_bfloat16*>(crow)[c + t] = __float2bfloat16_rn(v[t]);
              }
            }
          }
        }
      }
      tc::tc_fence_before();
      tc::mbar_arrive(&tmem_empty[as]);
    }
    if (STORE_TMA && lane == 0) tc::tma_store_wait_all();  // all bulk stores of this warp have landed before the CTA exits
  } else if (GATHER) {
    // ===================== gather producers (4 warps): A rows -> swizzled smem =====================
    const int gt = threadIdx.x - GATHER_WARP0 * 32;  // 0..127
    const int chunk = gt & 7;                        // 16-byte chunk inside the 128-byte k-block row
    const int r0 = gt >> 3;                          // rows r0 + 16*i, i = 0..7
    int stage = 0;
    uint32_t phase = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      const int m0 = (tile / tiles_n) * BLOCK_M;
      for (int s = 0; s < p.n_seg; ++s) {
        const __nv_bfloat16* rowptr[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const int m = m0 + r0 + 16 * i;
          rowptr[i] = nullptr;
          if (m < p.M) {
            const int j = p.seg_idx[s] ? __ldg(p.seg_idx[s] + m) : m;
            if (p.seg_peer[s] && j >= p.seg_nlocal[s]) {
              const int h = j - (int)p.seg_nlocal[s];
              rowptr[i] = p.seg_peer[s][__ldg(p.seg_halo_owner[s] + h)] + (int64_t)__ldg(p.seg_halo_row[s] + h) * p.seg_ld[s];
            } else {
              rowptr[i] = p.seg_base[s] + (int64_t)j * p.seg_ld[s];
            }
          }
        }
        const bool remote_ok = p.seg_peer[s] != nullptr;  // peer rows were written by another GPU: coherent loads
        for (int kb = 0; kb < p.seg_kblocks[s]; ++kb) {
          tc::mbar_wait(&empty_bar[stage], phase ^ 1);
          uint8_t* sa = smem + stage * L::STAGE_BYTES;
          const int k = kb * BLOCK_K + chunk * 8;
          uint4 v[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            v[i] = make_uint4(0, 0, 0, 0);
            if (rowptr[i] != nullptr && k < p.seg_k[s]) v[i] = remote_ok ? ldg_v4(rowptr[i] + k) : ldg_nc_v4(rowptr[i] + k);
          }
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const int r = r0 + 16 * i;
            *reinterpret_cast<uint4*>(sa + r * 128 + ((chunk ^ (r & 7)) << 4)) = v[i];
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
  if (warp == 2) tc::tmem_dealloc(tmem_base, 2 * BLOCK_N);
}

template <int BLOCK_N, bool OUT_F32, bool GATHER, int ACT, bool STORE_TMA>
static int launch_gemm(const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& tcm, const GemmParams& p,
                       cudaStream_t stream) {
  using L = SmemLayout<BLOCK_N>;
  auto kern = gemm_bf16_tn_kernel<BLOCK_N, OUT_F32, GATHER, ACT, STORE_TMA>;
  DGB_OPT_IN_SMEM(kern, L::TOTAL);
  int dev = 0, sms = 0;
  DGB_CUDA(cudaGetDevice(&dev));
  DGB_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  const int tiles = ((p.M + BLOCK_M - 1) / BLOCK_M) * ((p.N + BLOCK_N - 1) / BLOCK_N);
  const int grid = tiles < sms ? tiles : sms;
  kern<<<grid, GATHER ? THREADS_GATHER : THREADS_DENSE, L::TOTAL, stream>>>(ta, tb, tcm, p);
  DGB_CHECK_LAUNCH();
  return 0;
}

template <int BLOCK_N, bool OUT_F32, bool GATHER, int ACT>
static int launch_store(const CUtensorMap& ta, const CUtensorMap& tb, const GemmParams& p, cudaStream_t s) {
  // TMA store needs a 16-byte aligned base and row stride; otherwise fall back to direct (clipped) register stores
  const int eb = OUT_F32 ? 4 : 2;
  const bool tma_ok = ((reinterpret_cast<uintptr_t>(p.C) & 15) == 0) && (((int64_t)p.ldc * eb) % 16 == 0);
  if (tma_ok) {
    constexpr int CH = OUT_F32 ? 32 : ((BLOCK_N / 2 >= 64) ? 64 : 32);
    CUtensorMap tcm;
    const int rc = make_tmap_2d(&tcm, p.C, (uint64_t)p.M, (uint64_t)p.N, (uint64_t)p.ldc, 32, CH, eb,
                                (CH * eb == 128) ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE);
    if (rc == 0) return launch_gemm<BLOCK_N, OUT_F32, GATHER, ACT, true>(ta, tb, tcm, p, s);
  }
  return launch_gemm<BLOCK_N, OUT_F32, GATHER, ACT, false>(ta, tb, tb, p, s);
}

template <int BLOCK_N, bool GATHER>
static int dispatch_act(const CUtensorMap& ta, const CUtensorMap& tb, const GemmParams& p, int out_f32, cudaStream_t s) {
  if (out_f32) {
    if (p.act == ACT_NONE) return launch_store<BLOCK_N, true, GATHER, ACT_NONE>(ta, tb, p, s);
    if (p.act == ACT_SILU) return launch_store<BLOCK_N, true, GATHER, ACT_SILU>(ta, tb, p, s);
    return launch_store<BLOCK_N, true, GATHER, ACT_RELU>(ta, tb, p, s);
  }
  if (p.act == ACT_NONE) return launch_store<BLOCK_N, false, GATHER, ACT_NONE>(ta, tb, p, s);
  if (p.act == ACT_SILU) return launch_store<BLOCK_N, false, GATHER, ACT_SILU>(ta, tb, p, s);
  return launch_store<BLOCK_N, false, GATHER, ACT_RELU>(ta, tb, p, s);
}

template <bool GATHER>
static int dispatch_gemm(const CUtensorMap& ta, const CUtensorMap& tb, const GemmParams& p, int out_f32, int block_n,
                         cudaStream_t s) {
  if (p.act < 0 || p.act > 2) return (int)cudaErrorInvalidValue;
  if (block_n == 64) return dispatch_act<64, GATHER>(ta, tb, p, out_f32, s);
  if (block_n == 128) return dispatch_act<128, GATHER>(ta, tb, p, out_f32, s);
  if (block_n == 256) return dispatch_act<256, GATHER>(ta, tb, p, out_f32, s);
  return (int)cudaErrorInvalidValue;
}

static int pick_block_n(int N) { return N <= 64 ? 64 : (N <= 128 ? 128 : 256); }

}  // namespace dgb

using namespace dgb;

// C[M,N] = act(A[M,K] @ B[N,K]^T + bias).  A, B bf16 row-major with 16-byte aligned bases and row strides that are
// multiples of 8 elements; C bf16 (out_f32 = 0) or fp32 (out_f32 = 1) with row stride ldc; bias fp32 [N] or null.
DGB_API int dgb_gemm_bf16_tn(const void* A, int64_t lda, const void* B, int64_t ldb, void* C, int64_t ldc, int M, int N,
                             int K, const float* bias, int act, int out_f32, void* stream) {
  if (M <= 0 || N <= 0 || K <= 0) return M == 0 || N == 0 ? 0 : (int)cudaErrorInvalidValue;
  if (N > MAX_N) return (int)cudaErrorInvalidValue;
  if ((lda % 8) || (ldb % 8) || (reinterpret_cast<uintptr_t>(A) & 15) || (reinterpret_cast<uintptr_t>(B) & 15))
    return (int)cudaErrorInvalidValue;
  const int bn = pick_block_n(N);
  CUtensorMap ta, tb;
  int rc = make_tmap_bf16_2d(&ta, A, (uint64_t)M, (uint64_t)K, (uint64_t)lda, BLOCK_M, BLOCK_K);
  if (rc) return rc;
  rc = make_tmap_bf16_2d(&tb, B, (uint64_t)N, (uint64_t)K, (uint64_t)ldb, bn, BLOCK_K);
  if (rc) return rc;
  GemmParams p{};
  p.M = M; p.N = N; p.K = K; p.bias = bias; p.C = C; p.ldc = ldc; p.act = act;
  return dispatch_gemm<false>(ta, tb, p, out_f32, bn, (cudaStream_t)stream);
}

// Gather-GEMM: C[m,:] = act( sum_s  A_s[idx_s[m], :K_s] @ W[:, koff_s : koff_s + K_s]^T + bias ), i.e. the Linear applied
// to cat([A_0[idx_0], A_1[idx_1], A_2[idx_2]], dim=1) without materialising the gathered rows or the concatenation.
// W is [N, sum_s pad64(K_s)] bf16 row-major (each segment's columns padded to a multiple of 64 with zeros).
DGB_API int dgb_gather_gemm_bf16(int n_seg, const void* const* seg_base, const int64_t* seg_ld, const int* const* seg_idx,
                                 const int* seg_k, const void* W, int64_t ldw, void* C, int64_t ldc, int M, int N,
                                 const float* bias, int act, int out_f32, void* stream) {
  if (M <= 0 || N <= 0) return 0;
  if (n_seg < 1 || n_seg > 3 || (ldw % 8) || N > MAX_N) return (int)cudaErrorInvalidValue;
  GemmParams p{};
  p.M = M; p.N = N; p.bias = bias; p.C = C; p.ldc = ldc; p.act = act; p.n_seg = n_seg;
  int ktot = 0;
  for (int s = 0; s < n_seg; ++s) {
    if ((seg_ld[s] % 8) || (seg_k[s] % 8) || (reinterpret_cast<uintptr_t>(seg_base[s]) & 15)) return (int)cudaErrorInvalidValue;
    p.seg_base[s] = (const __nv_bfloat16*)seg_base[s];
    p.seg_ld[s] = seg_ld[s];
    p.seg_idx[s] = seg_idx[s];
    p.seg_k[s] = seg_k[s];
    p.seg_kblocks[s] = (seg_k[s] + BLOCK_K - 1) / BLOCK_K;
    ktot += p.seg_kblocks[s] * BLOCK_K;
  }
  p.K = ktot;
  const int bn = pick_block_n(N);
  CUtensorMap tb;
  int rc = make_tmap_bf16_2d(&tb, W, (uint64_t)N, (uint64_t)ktot, (uint64_t)ldw, bn, BLOCK_K);
  if (rc) return rc;
  return dispatch_gemm<true>(tb, tb, p, out_f32, bn, (cudaStream_t)stream);
}

// Gather-GEMM whose gathered rows may live on other GPUs (F1: halo gather fused into the first edge/node MLP GEMM).
// Same as dgb_gather_gemm_bf16; for every segment s with seg_peer[s] != null, gathered row j >= seg_nlocal[s] is read from
// seg_peer[s][halo_owner[s][j - nlocal]] + halo_row[s][j - nlocal] * seg_ld[s] -- a plain 16-byte load on an NVLink-mapped
// peer pointer issued by the producer warps that assemble the tcgen05 A tile.  The caller orders the exchange (barrier).
DGB_API int dgb_gather_gemm_peer_bf16(int n_seg, const void* const* seg_base, const int64_t* seg_ld, const int* const* seg_idx,
                                      const int* seg_k, const void* const* const* seg_peer, const int* const* seg_halo_owner,
                                      const int* const* seg_halo_row, const int64_t* seg_nlocal, const void* W, int64_t ldw,
                                      void* C, int64_t ldc, int M, int N, const float* bias, int act, int out_f32,
                                      void* stream) {
  if (M <= 0 || N <= 0) return 0;
  if (n_seg < 1 || n_seg > 3 || (ldw % 8) || N > MAX_N) return (int)cudaErrorInvalidValue;
  GemmParams p{};
  p.M = M; p.N = N; p.bias = bias; p.C = C; p.ldc = ldc; p.act = act; p.n_seg = n_seg;
  int ktot = 0;
  for (int s = 0; s < n_seg; ++s) {
    if ((seg_ld[s] % 8) || (seg_k[s] % 8) || (reinterpret_cast<uintptr_t>(seg_base[s]) & 15)) return (int)cudaErrorInvalidValue;
    p.seg_base[s] = (const __nv_bfloat16*)seg_base[s];
    p.seg_ld[s] = seg_ld[s];
    p.seg_idx[s] = seg_idx[s];
    p.seg_k[s] = seg_k[s];
    p.seg_kblocks[s] = (seg_k[s] + BLOCK_K - 1) / BLOCK_K;
    p.seg_peer[s] = seg_peer ? (const __nv_bfloat16* const*)seg_peer[s] : nullptr;
    p.seg_halo_owner[s] = seg_halo_owner ? seg_halo_owner[s] : nullptr;
    p.seg_halo_row[s] = seg_halo_row ? seg_halo_row[s] : nullptr;
    p.seg_nlocal[s] = seg_nlocal ? seg_nlocal[s] : 0;
    ktot += p.seg_kblocks[s] * BLOCK_K;
  }
  p.K = ktot;
  const int bn = pick_block_n(N);
  CUtensorMap tb;
  int rc = make_tmap_bf16_2d(&tb, W, (uint64_t)N, (uint64_t)ktot, (uint64_t)ldw, bn, BLOCK_K);
  if (rc) return rc;
  return dispatch_gemm<true>(tb, tb, p, out_f32, bn, (cudaStream_t)stream);
}
