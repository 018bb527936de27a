// Peer-memory (NVLink / NVSwitch) runtime and kernels: symmetric arena, signal-pad barrier, halo pull, halo-gradient
// pull-reduce, one-shot all-reduce fused into Adam.
//
// This is the B200-native replacement of the reference's NVSHMEM extension (csrc/torch_nvshmem_p2p.cu:28-376:
// init / AllocateSymmetricMemory / dist_get / dist_put / barrier) and of its NCCL `put` (NCCLBackendEngine.py:281-306):
//   * one persistent arena per rank (cudaMalloc once), exported with CUDA IPC and mapped by every peer of the node -- no
//     per-call nvshmem_malloc, no per-call clone of the input into the symmetric heap (NVSHMEMBackendEngine.py:22-63);
//   * data moves with plain 16-byte loads from peer pointers issued inside the kernels ("pull"): the halo exchange is ONE
//     kernel (no send-side index kernel, no all-to-all, no cat), its backward ONE deterministic pull-reduce kernel
//     (no CAS-loop float atomics, nvshmem_comm_kernels.cuh:17-54);
//   * cross-GPU ordering uses epoch-numbered flags on a symmetric signal pad (st.release.sys / ld.acquire.sys).
#include "common.cuh"
#include <cstring>

namespace dgb {

// ------------------------------------------------------------------------------------------------
// Signal-pad barrier: pad[r] (on every rank) is written by rank r.  One block, W threads.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

__global__ void peer_barrier_kernel(uint32_t* const* pads, int rank, int world, uint32_t epoch, int* timeout_flag,
                                    long long max_spin) {
  const int d = threadIdx.x;
  __threadfence_system();  // publish everything earlier kernels of this stream wrote
  if (d < world) {
    st_release_sys(pads[d] + rank, epoch);
    const uint32_t* mine = pads[rank] + d;
    long long spins = 0;
    // epochs only grow; compare with wrap-around-safe signed difference
    while ((int32_t)(ld_acquire_sys(mine) - epoch) < 0) {
      if (++spins > max_spin) {
        if (timeout_flag) *timeout_flag = 1;
        break;
      }
      __nanosleep(64);
    }
  }
  __syncthreads();
  __threadfence_system();
}

// Same barrier with the epoch kept in device memory: the kernel takes the next epoch from `*epoch_dev` and stores it back, so a
// CUDA graph that contains barriers replays correctly (a kernel argument would be frozen at capture time).  Every rank launches
// the same sequence of barriers, hence all counters advance in lock-step.  One engine uses either this form or the host-epoch
// form for its whole life, never both (the two counters would drift apart).
__global__ void peer_barrier_dev_kernel(uint32_t* const* pads, int rank, int world, uint32_t* epoch_dev, int* timeout_flag,
                                        long long max_spin) {
  __shared__ uint32_t epoch_s;
  const int d = threadIdx.x;
  if (d == 0) epoch_s = *epoch_dev + 1u;
  __syncthreads();
  const uint32_t epoch = epoch_s;
  __threadfence_system();  // publish everything earlier kernels of this stream wrote
  if (d < world) {
    st_release_sys(pads[d] + rank, epoch);
    const uint32_t* mine = pads[rank] + d;
    long long spins = 0;
    while ((int32_t)(ld_acquire_sys(mine) - epoch) < 0) {
      if (++spins > max_spin) {
        if (timeout_flag) *timeout_flag = 1;
        break;
      }
      __nanosleep(64);
    }
  }
  __syncthreads();
  if (d == 0) *epoch_dev = epoch;
  __threadfence_system();
}

// ------------------------------------------------------------------------------------------------
// Halo pull: dst[h, :] = peer[halo_owner[h]][halo_row[h], :]   (bf16 or any 2-byte / 4-byte type: copies 16-B chunks)
// One sub-warp group per row; `row_bytes` multiple of 16.
// ------------------------------------------------------------------------------------------------
template <int G>
__global__ void __launch_bounds__(256) halo_pull_kernel(uint8_t* __restrict__ dst, int64_t ld_dst_bytes, int64_t n_halo,
                                                        int row_bytes, const uint8_t* const* __restrict__ peer_base,
                                                        int64_t ld_peer_bytes, const int* __restrict__ halo_owner,
                                                        const int* __restrict__ halo_row, int64_t rot) {
  // Halo rows are stored owner-major; every rank starts `rot` rows in (the segment of owner rank+1), so at any moment the
  // ranks read from different owners instead of all hitting owner 0 first (NVLink incast).
  const int lane = threadIdx.x & 31;
  const int g = lane & (G - 1);
  const int64_t grp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) / G;
  const int64_t ngrp = ((int64_t)gridDim.x * blockDim.x) / G;
  const int chunks = row_bytes >> 4;
  for (int64_t h0 = grp; h0 < n_halo; h0 += ngrp) {
    int64_t h = h0 + rot;
    if (h >= n_halo) h -= n_halo;
    const uint8_t* src = peer_base[halo_owner[h]] + (int64_t)halo_row[h] * ld_peer_bytes;
    uint8_t* out = dst + h * ld_dst_bytes;
    for (int c = g; c < chunks; c += G) stg_v4(out + (c << 4), ldg_v4(src + (c << 4)));
  }
}

// ------------------------------------------------------------------------------------------------
// Halo-gradient pull-reduce (backward of the halo exchange):
//   gx[bnd_row[b], :] = g_local[bnd_row[b], :] + sum_{k in [bnd_ptr[b], bnd_ptr[b+1])} peer[src_rank[k]][src_row[k], :]
// Rows that are not boundary rows are copied by the caller (or alias g_local).  bf16 in/out, fp32 accumulate,
// fixed summation order (deterministic).
// ------------------------------------------------------------------------------------------------
template <int G, int NC>
__global__ void __launch_bounds__(256)
halo_grad_reduce_kernel(__nv_bfloat16* __restrict__ gx, int64_t ld_gx, const __nv_bfloat16* __restrict__ g_local,
                        int64_t ld_gl, int H, int64_t n_bnd, const int* __restrict__ bnd_row,
                        const int* __restrict__ bnd_ptr, const int* __restrict__ src_rank, const int* __restrict__ src_row,
                        const __nv_bfloat16* const* __restrict__ peer_base, int64_t ld_peer) {
  const int lane = threadIdx.x & 31;
  const int g = lane & (G - 1);
  const int64_t b = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) / G;
  if (b >= n_bnd) return;
  const int r = bnd_row[b];
  float acc[NC][8];
#pragma unroll
  for (int c = 0; c < NC; ++c) {
    const int colo = (g + c * G) * 8;
    if (colo < H) unpack8(ldg_v4(g_local + (int64_t)r * ld_gl + colo), acc[c]);  // coherent load: gx may alias g_local
  }
  for (int k = bnd_ptr[b]; k < bnd_ptr[b + 1]; ++k) {
    const __nv_bfloat16* src = peer_base[src_rank[k]] + (int64_t)src_row[k] * ld_peer;
#pragma unroll
    for (int c = 0; c < NC; ++c) {
      const int colo = (g + c * G) * 8;
      if (colo < H) fadd_bf16x8(ldg_v4(src + colo), acc[c]);
    }
  }
#pragma unroll
  for (int c = 0; c < NC; ++c) {
    const int colo = (g + c * G) * 8;
    if (colo < H) stg_v4(gx + (int64_t)r * ld_gx + colo, pack8(acc[c]));
  }
}

// ------------------------------------------------------------------------------------------------
// One-shot all-reduce(SUM) of fp32 vectors living at the same arena offset on every rank, fused into Adam:
//   g = sum_r peer_g[r][i];  Adam(p, g * grad_scale, m, v)
// Every rank reads all W gradient buffers (W x n x 4 bytes over NVLink; n <= ~1 M for the models here) in the same
// order, so all ranks compute bit-identical updates and the replicated weights never diverge.
// ------------------------------------------------------------------------------------------------
struct AdamArgs {
  float lr, beta1, beta2, eps, weight_decay, grad_scale, bias_c1, bias_c2;
};

__global__ void __launch_bounds__(256) adam_allreduce_kernel(float* __restrict__ p, const float* const* __restrict__ peer_g,
                                                             int world, float* __restrict__ m, float* __restrict__ v,
                                                             float* __restrict__ g_out, int64_t n, AdamArgs a) {
  const int64_t i4 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
  if (i4 >= n) return;  // n is padded to a multiple of 4 by the host
  float4 gs = make_float4(0.f, 0.f, 0.f, 0.f);
  for (int r = 0; r < world; ++r) {
    const uint4 raw = ldg_v4(peer_g[r] + i4);
    gs.x += __uint_as_float(raw.x); gs.y += __uint_as_float(raw.y);
    gs.z += __uint_as_float(raw.z); gs.w += __uint_as_float(raw.w);
  }
  if (g_out) *reinterpret_cast<float4*>(g_out + i4) = gs;
  float4 pv = *reinterpret_cast<float4*>(p + i4);
  float4 mv = *reinterpret_cast<float4*>(m + i4);
  float4 vv = *reinterpret_cast<float4*>(v + i4);
  float pp[4] = {pv.x, pv.y, pv.z, pv.w}, gg[4] = {gs.x, gs.y, gs.z, gs.w};
  float mm[4] = {mv.x, mv.y, mv.z, mv.w}, vw[4] = {vv.x, vv.y, vv.z, vv.w};
  const float inv_c1 = 1.f / a.bias_c1;
  const float inv_sqrt_c2 = rsqrtf(a.bias_c2);
#pragma unroll
  for (int t = 0; t < 4; ++t) {
    const float gr = gg[t] * a.grad_scale + a.weight_decay * pp[t];
    mm[t] = a.beta1 * mm[t] + (1.f - a.beta1) * gr;
    vw[t] = a.beta2 * vw[t] + (1.f - a.beta2) * gr * gr;
    pp[t] -= a.lr * (mm[t] * inv_c1) / (sqrtf(vw[t]) * inv_sqrt_c2 + a.eps);
  }
  *reinterpret_cast<float4*>(p + i4) = make_float4(pp[0], pp[1], pp[2], pp[3]);
  *reinterpret_cast<float4*>(m + i4) = make_float4(mm[0], mm[1], mm[2], mm[3]);
  *reinterpret_cast<float4*>(v + i4) = make_float4(vw[0], vw[1], vw[2], vw[3]);
}

// Capturable form of the fused all-reduce + Adam step: step counter and {lr, 1 - beta1^t, 1 - beta2^t} live in device memory
// (same convention as dgb_adam_flat_capturable in adam.cu), so the launch can sit inside a CUDA graph.
__global__ void adam_tick_peer_kernel(int* __restrict__ step, float* __restrict__ dev_state, float beta1, float beta2) {
  const int t = *step + 1;
  *step = t;
  dev_state[1] = 1.f - powf(beta1, (float)t);
  dev_state[2] = 1.f - powf(beta2, (float)t);
}

__global__ void __launch_bounds__(256) adam_allreduce_dev_kernel(float* __restrict__ p, const float* const* __restrict__ peer_g,
                                                                 int world, float* __restrict__ m, float* __restrict__ v,
                                                                 float* __restrict__ g_out, int64_t n, AdamArgs a,
                                                                 const float* __restrict__ dev_state) {
  const int64_t i4 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
  if (i4 >= n) return;
  a.lr = dev_state[0];
  a.bias_c1 = dev_state[1];
  a.bias_c2 = dev_state[2];
  float4 gs = make_float4(0.f, 0.f, 0.f, 0.f);
  for (int r = 0; r < world; ++r) {
    const uint4 raw = ldg_v4(peer_g[r] + i4);
    gs.x += __uint_as_float(raw.x); gs.y += __uint_as_float(raw.y);
    gs.z += __uint_as_float(raw.z); gs.w += __uint_as_float(raw.w);
  }
  if (g_out) *reinterpret_cast<float4*>(g_out + i4) = gs;
  float4 pv = *reinterpret_cast<float4*>(p + i4);
  float4 mv = *reinterpret_cast<float4*>(m + i4);
  float4 vv = *reinterpret_cast<float4*>(v + i4);
  float pp[4] = {pv.x, pv.y, pv.z, pv.w}, gg[4] = {gs.x, gs.y, gs.z, gs.w};
  float mm[4] = {mv.x, mv.y, mv.z, mv.w}, vw[4] = {vv.x, vv.y, vv.z, vv.w};
  const float inv_c1 = 1.f / a.bias_c1;
  const float inv_sqrt_c2 = rsqrtf(a.bias_c2);
#pragma unroll
  for (int t = 0; t < 4; ++t) {
    const float gr = gg[t] * a.grad_scale + a.weight_decay * pp[t];
    mm[t] = a.beta1 * mm[t] + (1.f - a.beta1) * gr;
    vw[t] = a.beta2 * vw[t] + (1.f - a.beta2) * gr * gr;
    pp[t] -= a.lr * (mm[t] * inv_c1) / (sqrtf(vw[t]) * inv_sqrt_c2 + a.eps);
  }
  *reinterpret_cast<float4*>(p + i4) = make_float4(pp[0], pp[1], pp[2], pp[3]);
  *reinterpret_cast<float4*>(m + i4) = make_float4(mm[0], mm[1], mm[2], mm[3]);
  *reinterpret_cast<float4*>(v + i4) = make_float4(vw[0], vw[1], vw[2], vw[3]);
}

// out[i] = sum_r peer[r][i]  (fp32, n multiple of 4): small all-reduces (BatchNorm statistics, loss scalars)
__global__ void __launch_bounds__(256) allreduce_sum_kernel(float* __restrict__ out, const float* const* __restrict__ peer,
                                                            int world, int64_t n) {
  const int64_t i4 = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) * 4;
  if (i4 >= n) return;
  float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
  for (int r = 0; r < world; ++r) {
    const uint4 raw = ldg_v4(peer[r] + i4);
    s.x += __uint_as_float(raw.x); s.y += __uint_as_float(raw.y);
    s.z += __uint_as_float(raw.z); s.w += __uint_as_float(raw.w);
  }
  *reinterpret_cast<float4*>(out + i4) = s;
}

// Row gather with peer addressing: out[e, :] = (owner[e] == rank ? local : peer[owner[e]])[row[e], :]
template <int G>
__global__ void __launch_bounds__(256) gather_rows_peer_kernel(uint8_t* __restrict__ out, int64_t ld_out_bytes, int64_t n,
                                                               int row_bytes, const uint8_t* const* __restrict__ peer_base,
                                                               int64_t ld_peer_bytes, const int* __restrict__ owner,
                                                               const int* __restrict__ row) {
  const int lane = threadIdx.x & 31;
  const int g = lane & (G - 1);
  const int64_t grp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) / G;
  const int64_t ngrp = ((int64_t)gridDim.x * blockDim.x) / G;
  const int chunks = row_bytes >> 4;
  for (int64_t e = grp; e < n; e += ngrp) {
    const uint8_t* src = peer_base[owner[e]] + (int64_t)row[e] * ld_peer_bytes;
    uint8_t* o = out + e * ld_out_bytes;
    for (int c = g; c < chunks; c += G) stg_v4(o + (c << 4), ldg_v4(src + (c << 4)));
  }
}

// Row scatter-add with peer addressing (fp32): peer[owner[i]][row[i], :] += src[i, :] -- native vector reductions on
// peer-mapped memory instead of the reference's int CAS loop (nvshmem_comm_kernels.cuh:17-54, 91-128).
template <int G>
__global__ void __launch_bounds__(256) scatter_add_rows_peer_kernel(const float* __restrict__ src, int64_t ld_src, int64_t n,
                                                                    int F, float* const* __restrict__ peer_base,
                                                                    int64_t ld_peer, const int* __restrict__ owner,
                                                                    const int* __restrict__ row) {
  const int lane = threadIdx.x & 31;
  const int g = lane & (G - 1);
  const int64_t grp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) / G;
  const int64_t ngrp = ((int64_t)gridDim.x * blockDim.x) / G;
  const int chunks = F >> 2;
  for (int64_t i = grp; i < n; i += ngrp) {
    float* dst = peer_base[owner[i]] + (int64_t)row[i] * ld_peer;
    const float* s = src + i * ld_src;
    for (int c = g; c < chunks; c += G) {
      const float4 v = *reinterpret_cast<const float4*>(s + c * 4);
      asm volatile("red.global.add.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(dst + c * 4), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
                   : "memory");
    }
  }
}

}  // namespace dgb

using namespace dgb;

// ---------------------------------------------------------------- arena / IPC runtime
DGB_API int dgb_symm_alloc(size_t bytes, void** out) {
  DGB_CUDA(cudaMalloc(out, bytes));
  DGB_CUDA(cudaMemset(*out, 0, bytes));
  return 0;
}
DGB_API int dgb_symm_free(void* p) {
  DGB_CUDA(cudaFree(p));
  return 0;
}
// 64-byte CUDA IPC handle of an arena returned by dgb_symm_alloc
DGB_API int dgb_symm_export(void* p, void* handle64) {
  cudaIpcMemHandle_t h;
  DGB_CUDA(cudaIpcGetMemHandle(&h, p));
  memcpy(handle64, &h, sizeof(h));
  return 0;
}
DGB_API int dgb_symm_import(const void* handle64, void** out) {
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, sizeof(h));
  DGB_CUDA(cudaIpcOpenMemHandle(out, h, cudaIpcMemLazyEnablePeerAccess));
  return 0;
}
DGB_API int dgb_symm_unmap(void* p) {
  DGB_CUDA(cudaIpcCloseMemHandle(p));
  return 0;
}
// same-process peers (several GPUs driven by one process, e.g. tests): enable direct access to device `peer`
DGB_API int dgb_enable_peer_access(int peer) {
  cudaError_t e = cudaDeviceEnablePeerAccess(peer, 0);
  if (e == cudaErrorPeerAccessAlreadyEnabled) {
    cudaGetLastError();
    return 0;
  }
  return (int)e;
}

// ---------------------------------------------------------------- kernels
DGB_API int dgb_peer_barrier(void* const* pads, int rank, int world, unsigned epoch, int* timeout_flag, long long max_spin,
                             void* stream) {
  if (world < 1 || world > 64) return (int)cudaErrorInvalidValue;
  peer_barrier_kernel<<<1, 64, 0, (cudaStream_t)stream>>>((uint32_t* const*)pads, rank, world, epoch, timeout_flag, max_spin);
  DGB_CHECK_LAUNCH();
  return 0;
}

static int pick_group(int chunks) { return chunks <= 4 ? 4 : (chunks <= 8 ? 8 : (chunks <= 16 ? 16 : 32)); }

DGB_API int dgb_halo_pull(void* dst, int64_t ld_dst_bytes, int64_t n_halo, int row_bytes, const void* const* peer_base,
                          int64_t ld_peer_bytes, const int* halo_owner, const int* halo_row, int64_t rot, void* stream) {
  if (n_halo == 0) return 0;
  if (rot < 0 || rot >= n_halo) rot = 0;
  if ((row_bytes & 15) || (ld_dst_bytes & 15) || (ld_peer_bytes & 15)) return (int)cudaErrorInvalidValue;
  const int G = pick_group(row_bytes >> 4);
  const int64_t groups_per_block = 256 / G;
  int64_t blocks = (n_halo + groups_per_block - 1) / groups_per_block;
  if (blocks > 148 * 16) blocks = 148 * 16;
  cudaStream_t s = (cudaStream_t)stream;
#define DGB_HP(GG)                                                                                                   \
  halo_pull_kernel<GG><<<(unsigned)blocks, 256, 0, s>>>((uint8_t*)dst, ld_dst_bytes, n_halo, row_bytes,              \
                                                       (const uint8_t* const*)peer_base, ld_peer_bytes, halo_owner, \
                                                       halo_row, rot)
  if (G == 4) DGB_HP(4); else if (G == 8) DGB_HP(8); else if (G == 16) DGB_HP(16); else DGB_HP(32);
#undef DGB_HP
  DGB_CHECK_LAUNCH();
  return 0;
}

DGB_API int dgb_gather_rows_peer(void* out, int64_t ld_out_bytes, int64_t n, int row_bytes, const void* const* peer_base,
                                 int64_t ld_peer_bytes, const int* owner, const int* row, void* stream) {
  if (n == 0) return 0;
  if ((row_bytes & 15) || (ld_out_bytes & 15) || (ld_peer_bytes & 15)) return (int)cudaErrorInvalidValue;
  const int G = pick_group(row_bytes >> 4);
  const int64_t groups_per_block = 256 / G;
  int64_t blocks = (n + groups_per_block - 1) / groups_per_block;
  if (blocks > 148 * 16) blocks = 148 * 16;
  cudaStream_t s = (cudaStream_t)stream;
#define DGB_GR(GG)                                                                                                        \
  gather_rows_peer_kernel<GG><<<(unsigned)blocks, 256, 0, s>>>((uint8_t*)out, ld_out_bytes, n, row_bytes,                  \
                                                              (const uint8_t* const*)peer_base, ld_peer_bytes, owner, row)
  if (G == 4) DGB_GR(4); else if (G == 8) DGB_GR(8); else if (G == 16) DGB_GR(16); else DGB_GR(32);
#undef DGB_GR
  DGB_CHECK_LAUNCH();
  return 0;
}

DGB_API int dgb_scatter_add_rows_peer_f32(const float* src, int64_t ld_src, int64_t n, int F, void* const* peer_base,
                                          int64_t ld_peer, const int* owner, const int* row, void* stream) {
  if (n == 0) return 0;
  if ((F & 3) || (ld_src & 3) || (ld_peer & 3)) return (int)cudaErrorInvalidValue;
  const int G = pick_group(F >> 2);
  int64_t blocks = (n + (256 / G) - 1) / (256 / G);
  if (blocks > 148 * 16) blocks = 148 * 16;
  cudaStream_t s = (cudaStream_t)stream;
#define DGB_SA(GG)                                                                                                  \
  scatter_add_rows_peer_kernel<GG><<<(unsigned)blocks, 256, 0, s>>>(src, ld_src, n, F, (float* const*)peer_base,    \
                                                                   ld_peer, owner, row)
  if (G == 4) DGB_SA(4); else if (G == 8) DGB_SA(8); else if (G == 16) DGB_SA(16); else DGB_SA(32);
#undef DGB_SA
  DGB_CHECK_LAUNCH();
  return 0;
}

DGB_API int dgb_halo_grad_reduce(void* gx, int64_t ld_gx, const void* g_local, int64_t ld_gl, int H, int64_t n_bnd,
                                 const int* bnd_row, const int* bnd_ptr, const int* src_rank, const int* src_row,
                                 const void* const* peer_base, int64_t ld_peer, void* stream) {
  if (n_bnd == 0) return 0;
  if ((H % 8) || (ld_gx % 8) || (ld_gl % 8) || (ld_peer % 8)) return (int)cudaErrorInvalidValue;
  const int chunks = H / 8;
  cudaStream_t s = (cudaStream_t)stream;
#define DGB_HG(G, NC)                                                                                              \
  do {                                                                                                             \
    const int64_t per_block = 256 / (G);                                                                           \
    const int64_t blocks = (n_bnd + per_block - 1) / per_block;                                                    \
    halo_grad_reduce_kernel<G, NC><<<(unsigned)blocks, 256, 0, s>>>(                                               \
        (__nv_bfloat16*)gx, ld_gx, (const __nv_bfloat16*)g_local, ld_gl, H, n_bnd, bnd_row, bnd_ptr, src_rank,     \
        src_row, (const __nv_bfloat16* const*)peer_base, ld_peer);                                                 \
  } while (0)
  if (chunks <= 4) DGB_HG(4, 1);
  else if (chunks <= 8) DGB_HG(8, 1);
  else if (chunks <= 16) DGB_HG(16, 1);
  else if (chunks <= 32) DGB_HG(32, 1);
  else if (chunks <= 64) DGB_HG(32, 2);
  else if (chunks <= 128) DGB_HG(32, 4);
  else return (int)cudaErrorInvalidValue;
#undef DGB_HG
  DGB_CHECK_LAUNCH();
  return 0;
}

DGB_API int dgb_adam_allreduce(float* p, const void* const* peer_g, int world, float* m, float* v, float* g_out, int64_t n,
                               float lr, float beta1, float beta2, float eps, float weight_decay, float grad_scale,
                               int step, void* stream) {
  if (n == 0) return 0;
  if (step < 1 || (n & 3)) return (int)cudaErrorInvalidValue;
  AdamArgs a{lr, beta1, beta2, eps, weight_decay, grad_scale, 1.f - powf(beta1, (float)step), 1.f - powf(beta2, (float)step)};
  const int64_t threads = n / 4;
  adam_allreduce_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
      p, (const float* const*)peer_g, world, m, v, g_out, n, a);
  DGB_CHECK_LAUNCH();
  return 0;
}

// Barrier whose epoch lives in device memory (`epoch_dev`, one uint32 per engine, zero-initialised): capturable in a CUDA graph.
DGB_API int dgb_peer_barrier_dev(void* const* pads, int rank, int world, unsigned* epoch_dev, int* timeout_flag,
                                 long long max_spin, void* stream) {
  if (world < 1 || world > 64 || !epoch_dev) return (int)cudaErrorInvalidValue;
  peer_barrier_dev_kernel<<<1, 64, 0, (cudaStream_t)stream>>>((uint32_t* const*)pads, rank, world, epoch_dev, timeout_flag,
                                                              max_spin);
  DGB_CHECK_LAUNCH();
  return 0;
}

// dgb_adam_allreduce with `step_dev` (int32) and `dev_state` = {lr, 1-beta1^t, 1-beta2^t, unused} in device memory: two launches
// (tick + update), nothing that changes from step to step is a kernel argument.
DGB_API int dgb_adam_allreduce_capturable(float* p, const void* const* peer_g, int world, float* m, float* v, float* g_out,
                                          int64_t n, float beta1, float beta2, float eps, float weight_decay, float grad_scale,
                                          int* step_dev, float* dev_state, void* stream) {
  if (n == 0) return 0;
  if ((n & 3) || !step_dev || !dev_state) return (int)cudaErrorInvalidValue;
  adam_tick_peer_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(step_dev, dev_state, beta1, beta2);
  DGB_CHECK_LAUNCH();
  AdamArgs a{0.f, beta1, beta2, eps, weight_decay, grad_scale, 1.f, 1.f};
  const int64_t threads = n / 4;
  adam_allreduce_dev_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
      p, (const float* const*)peer_g, world, m, v, g_out, n, a, dev_state);
  DGB_CHECK_LAUNCH();
  return 0;
}

DGB_API int dgb_allreduce_sum_f32(float* out, const void* const* peer, int world, int64_t n, void* stream) {
  if (n == 0) return 0;
  if (n & 3) return (int)cudaErrorInvalidValue;
  const int64_t threads = n / 4;
  allreduce_sum_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, (cudaStream_t)stream>>>(out, (const float* const*)peer,
                                                                                           world, n);
  DGB_CHECK_LAUNCH();
  return 0;
}
