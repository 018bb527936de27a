// Weight packing for the split GraphConv Linear: one launch per layer and optimizer step turns the fp32 master weight
// W = [W_i | W_j] ([H, 2 * f_in], reference layout of `GraphConvLayer.conv`, experiments/OGB/GCN.py:31) into the three bf16 / fp32
// operands the fused layer consumes:
//   wcat   [2H, Fp]  rows 0..H-1 = W_i, rows H..2H-1 = W_j, columns zero-padded from f_in to Fp (16-byte feature rows)
//   wcat_t [Fp, 2H]  its transpose (B operand of the dX GEMM)
//   bcat   [2H]      bias on the P half, zeros on the Q half
// Before, every forward built wcat / bcat with a zero fill and slice copies and every backward transposed wcat again
// (8 ATen launches per layer and step).
#include "common.cuh"

namespace dgb {

__global__ void __launch_bounds__(256) pack_graphconv_kernel(const float* __restrict__ W, const float* __restrict__ bias, int H,
                                                             int f_in, int Fp, __nv_bfloat16* __restrict__ wcat,
                                                             __nv_bfloat16* __restrict__ wcat_t, float* __restrict__ bcat) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int total = 2 * H * Fp;
  if (i < 2 * H && bcat) bcat[i] = (i < H && bias) ? bias[i] : 0.f;
  if (i >= total) return;
  const int o = i / Fp, k = i - o * Fp;
  float v = 0.f;
  if (k < f_in) v = W[(int64_t)(o < H ? o : o - H) * (2 * f_in) + (o < H ? 0 : f_in) + k];
  const __nv_bfloat16 b = __float2bfloat16_rn(v);
  wcat[i] = b;
  if (wcat_t) wcat_t[(int64_t)k * (2 * H) + o] = b;
}

}  // namespace dgb

using namespace dgb;

DGB_API int dgb_pack_graphconv_weights(const float* W, const float* bias, int H, int f_in, int Fp, void* wcat, void* wcat_t,
                                       float* bcat, void* stream) {
  if (H <= 0 || f_in <= 0 || Fp < f_in || !W || !wcat) return (int)cudaErrorInvalidValue;
  const int total = 2 * H * Fp;
  pack_graphconv_kernel<<<(total + 255) / 256, 256, 0, (cudaStream_t)stream>>>(W, bias, H, f_in, Fp, (__nv_bfloat16*)wcat,
                                                                                (__nv_bfloat16*)wcat_t, bcat);
  DGB_CHECK_LAUNCH();
  return 0;
}
