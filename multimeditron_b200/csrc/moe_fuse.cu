// moe_fuse.cu — MoE 'weighted_average' fusion: out[b,p,:] = sum_e w[b, perm[e]] * x[e,b,p,:]
// Reference: src/multimeditron/model/modalities/image_modality_moe.py:171-177 (weights permuted into expert
// order) and image_modality_moe_pep.py:212-216 (no permutation -> perm = identity).  HBM-bound.
// The reference multiplies in the tensor dtype (bf16): stacked * w rounds each product to bf16, then sum(dim=1)
// accumulates in fp32 internally and rounds once; we reproduce those rounding points.
#include "common.cuh"

namespace mm {
constexpr int kMaxExperts = 16;

__global__ void __launch_bounds__(256) weighted_expert_sum_kernel(const __nv_bfloat16* __restrict__ x,
                                                                  const float* __restrict__ w,
                                                                  const int32_t* __restrict__ perm,
                                                                  __nv_bfloat16* __restrict__ out, int E, int64_t B,
                                                                  int64_t P, int64_t D) {
  const int64_t vpr = D >> 3;
  const int64_t rows = B * P;
  const int64_t total = rows * vpr;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = i / vpr, v = i - r * vpr;
    const int64_t b = r / P;
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (int e = 0; e < E; ++e) {
      const int pe = perm ? perm[e] : e;
      // weight is used in bf16 like the reference (weights.to(dtype) broadcast-multiplied)
      const float we = __bfloat162float(__float2bfloat16_rn(w[b * E + pe]));
      int4 xv = __ldg(reinterpret_cast<const int4*>(x + ((int64_t)e * rows + r) * D) + v);
      const uint32_t* u = reinterpret_cast<const uint32_t*>(&xv);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        acc[2 * j] += __bfloat162float(__float2bfloat16_rn(bf16_lo(u[j]) * we));
        acc[2 * j + 1] += __bfloat162float(__float2bfloat16_rn(bf16_hi(u[j]) * we));
      }
    }
    int4 o;
    o.x = (int)pack_bf16x2(acc[0], acc[1]);
    o.y = (int)pack_bf16x2(acc[2], acc[3]);
    o.z = (int)pack_bf16x2(acc[4], acc[5]);
    o.w = (int)pack_bf16x2(acc[6], acc[7]);
    reinterpret_cast<int4*>(out + r * D)[v] = o;
  }
}
}  // namespace mm

extern "C" int mm_weighted_expert_sum(const void* x, const float* w, const int32_t* perm, void* out, int64_t E,
                                      int64_t B, int64_t P, int64_t D, void* stream) {
  using namespace mm;
  if (E <= 0 || E > kMaxExperts || B < 0 || P < 0 || D <= 0 || D % 8) return MM_ERR_BAD_ARG;
  if (B * P == 0) return MM_OK;
  if (!x || !w || !out) return MM_ERR_BAD_ARG;
  const int64_t total = B * P * (D / 8);
  int64_t need = (total + 255) / 256;
  int64_t cap = (int64_t)sm_count() * 8;
  int grid = (int)(need < cap ? need : cap);
  weighted_expert_sum_kernel<<<grid, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const __nv_bfloat16*>(x), w, perm, reinterpret_cast<__nv_bfloat16*>(out), (int)E, B, P, D);
  MM_CUDA_CHECK(cudaGetLastError());
  return MM_OK;
}
