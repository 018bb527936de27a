// Shared device helpers for the dgraph_b200 sm_100a kernels.
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>

#define DGB_API extern "C" __attribute__((visibility("default")))

#define DGB_CHECK_LAUNCH()                         \
  do {                                             \
    cudaError_t e__ = cudaGetLastError();          \
    if (e__ != cudaSuccess) return (int)e__;       \
  } while (0)

#define DGB_CUDA(call)                             \
  do {                                             \
    cudaError_t e__ = (call);                      \
    if (e__ != cudaSuccess) return (int)e__;       \
  } while (0)

// The opt-in for more than 48 KB of dynamic shared memory is a per-DEVICE attribute of a kernel: remember it per device, so that a
// process driving several GPUs configures the kernel on each of them (a single per-process flag left the second device unconfigured).
#define DGB_OPT_IN_SMEM(kern, bytes)                                                                     \
  do {                                                                                                   \
    static bool configured__[64] = {false};                                                              \
    int dev__ = 0;                                                                                       \
    DGB_CUDA(cudaGetDevice(&dev__));                                                                     \
    if (dev__ < 0 || dev__ >= 64 || !configured__[dev__]) {                                              \
      DGB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(bytes)));  \
      if (dev__ >= 0 && dev__ < 64) configured__[dev__] = true;                                         \
    }                                                                                                    \
  } while (0)

namespace dgb {

constexpr int kWarp = 32;

__device__ __forceinline__ uint4 ldg_nc_v4(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}
// plain (coherent) 16-byte load: used for peer / symmetric memory that another GPU wrote in this epoch
__device__ __forceinline__ uint4 ldg_v4(const void* p) {
  uint4 r;
  asm volatile("ld.global.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}
// L2 eviction policies (createpolicy + .L2::cache_hint): gathered rows that other warps re-read are kept
// (evict_last); read-once streams are dropped first (evict_first).
__device__ __forceinline__ uint64_t l2_policy_evict_last() {
  uint64_t pol;
  asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
  return pol;
}
__device__ __forceinline__ uint64_t l2_policy_evict_first() {
  uint64_t pol;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
  return pol;
}
__device__ __forceinline__ uint4 ldg_hint_v4(const void* p, uint64_t pol) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v4.u32 {%0,%1,%2,%3}, [%4], %5;"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p), "l"(pol));
  return r;
}
__device__ __forceinline__ uint32_t ldg_stream_u8(const void* p) {  // read-once byte stream: do not keep it in L1
  uint32_t v;
  asm volatile("ld.global.nc.L1::no_allocate.u8 %0, [%1];" : "=r"(v) : "l"(p));
  return v;
}

__device__ __forceinline__ void stg_cs_u8(void* p, uint32_t v) {
  asm volatile("st.global.cs.u8 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void stg_v4(void* p, const uint4& v) {
  asm volatile("st.global.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w)
               : "memory");
}
__device__ __forceinline__ void stg_cs_v4(void* p, const uint4& v) {  // streaming store
  asm volatile("st.global.cs.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w)
               : "memory");
}

// bf16 pair packed in a 32-bit word -> two fp32 (bit tricks: bf16 is the top half of fp32)
__device__ __forceinline__ float bf_lo(uint32_t w) { return __uint_as_float(w << 16); }
__device__ __forceinline__ float bf_hi(uint32_t w) { return __uint_as_float(w & 0xffff0000u); }
__device__ __forceinline__ uint32_t pack_bf2(float lo, float hi) {
  __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&v);
}
// acc_lo += lo half, acc_hi += hi half of a packed bf16 pair, fp32 accumulate: the sm_100 mixed-precision add (`add.f32.bf16`, SASS
// FHADD.BF16 with an .H0 / .H1 operand selector) takes the bf16 operand as it sits in the register -- no shift / mask to widen it first.
__device__ __forceinline__ void fadd_bf16x2(uint32_t w, float& acc_lo, float& acc_hi) {
  unsigned short l, h;
  asm("mov.b32 {%0, %1}, %2;" : "=h"(l), "=h"(h) : "r"(w));
  asm("add.rn.f32.bf16 %0, %1, %0;" : "+f"(acc_lo) : "h"(l));
  asm("add.rn.f32.bf16 %0, %1, %0;" : "+f"(acc_hi) : "h"(h));
}
__device__ __forceinline__ void fadd_bf16x8(const uint4& v, float* acc) {
  fadd_bf16x2(v.x, acc[0], acc[1]); fadd_bf16x2(v.y, acc[2], acc[3]);
  fadd_bf16x2(v.z, acc[4], acc[5]); fadd_bf16x2(v.w, acc[6], acc[7]);
}
__device__ __forceinline__ void unpack8(const uint4& v, float* f) {
  f[0] = bf_lo(v.x); f[1] = bf_hi(v.x); f[2] = bf_lo(v.y); f[3] = bf_hi(v.y);
  f[4] = bf_lo(v.z); f[5] = bf_hi(v.z); f[6] = bf_lo(v.w); f[7] = bf_hi(v.w);
}
__device__ __forceinline__ uint4 pack8(const float* f) {
  uint4 v;
  v.x = pack_bf2(f[0], f[1]); v.y = pack_bf2(f[2], f[3]);
  v.z = pack_bf2(f[4], f[5]); v.w = pack_bf2(f[6], f[7]);
  return v;
}

}  // namespace dgb
