// layernorm.cu — row LayerNorm, bf16 in/out, fp32 statistics (two-pass over registers), one warp per row.
// HBM-bound: 2*M*D*2 bytes.  Replaces nn.LayerNorm at HF:models/clip/modeling_clip.py:371,380,677
// (eps 1e-5) and HF:models/siglip/modeling_siglip.py:346,355,617 (eps 1e-6).
#include "common.cuh"

namespace mm {

// D = 8 * 32 * VPL elements per row handled fully in registers (VPL 16-byte vectors per lane).
template <int VPL>
__global__ void __launch_bounds__(256) layernorm_kernel(const __nv_bfloat16* __restrict__ x,
                                                        __nv_bfloat16* __restrict__ y,
                                                        const __nv_bfloat16* __restrict__ gamma,
                                                        const __nv_bfloat16* __restrict__ beta, int64_t M, int D,
                                                        float eps, int64_t rows_in_group, int64_t rows_out_group,
                                                        int64_t in_row_offset) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int64_t nwarps = (int64_t)gridDim.x * (blockDim.x >> 5);
  const int nvec = D >> 3;  // 16-B vectors per row
  for (int64_t m = warp; m < M; m += nwarps) {
    int64_t mi = m;
    if (rows_in_group > 0) mi = (m / rows_out_group) * rows_in_group + in_row_offset + (m % rows_out_group);
    const int4* xr = reinterpret_cast<const int4*>(x + mi * D);
    int4 v[VPL];
    float sum = 0.f;
#pragma unroll
    for (int i = 0; i < VPL; ++i) {
      const int idx = lane + i * 32;
      if (idx < nvec) {
        v[i] = __ldg(xr + idx);
        const uint32_t* u = reinterpret_cast<const uint32_t*>(&v[i]);
#pragma unroll
        for (int j = 0; j < 4; ++j) sum += bf16_lo(u[j]) + bf16_hi(u[j]);
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    const float mean = sum / (float)D;
    float var = 0.f;
#pragma unroll
    for (int i = 0; i < VPL; ++i) {
      const int idx = lane + i * 32;
      if (idx < nvec) {
        const uint32_t* u = reinterpret_cast<const uint32_t*>(&v[i]);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          float a = bf16_lo(u[j]) - mean, b = bf16_hi(u[j]) - mean;
          var += a * a + b * b;
        }
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) var += __shfl_xor_sync(0xffffffffu, var, o);
    const float rstd = rsqrtf(var / (float)D + eps);
    int4* yr = reinterpret_cast<int4*>(y + m * D);
#pragma unroll
    for (int i = 0; i < VPL; ++i) {
      const int idx = lane + i * 32;
      if (idx < nvec) {
        int4 g = __ldg(reinterpret_cast<const int4*>(gamma) + idx);
        int4 b = __ldg(reinterpret_cast<const int4*>(beta) + idx);
        const uint32_t* u = reinterpret_cast<const uint32_t*>(&v[i]);
        const uint32_t* gu = reinterpret_cast<const uint32_t*>(&g);
        const uint32_t* bu = reinterpret_cast<const uint32_t*>(&b);
        int4 o;
        uint32_t* ou = reinterpret_cast<uint32_t*>(&o);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          float lo = (bf16_lo(u[j]) - mean) * rstd * bf16_lo(gu[j]) + bf16_lo(bu[j]);
          float hi = (bf16_hi(u[j]) - mean) * rstd * bf16_hi(gu[j]) + bf16_hi(bu[j]);
          ou[j] = pack_bf16x2(lo, hi);
        }
        yr[idx] = o;
      }
    }
  }
}

}  // namespace mm

extern "C" int mm_layernorm(const void* x, void* y, const void* gamma, const void* beta, int64_t M, int64_t D,
                            float eps, int64_t rows_in_group, int64_t rows_out_group, int64_t in_row_offset,
                            void* stream) {
  using namespace mm;
  if (M < 0 || D <= 0 || D % 8 != 0 || D > 8 * 32 * 8) return MM_ERR_BAD_ARG;
  if (M == 0) return MM_OK;
  if (!x || !y || !gamma || !beta) return MM_ERR_BAD_ARG;
  if (rows_in_group > 0 && rows_out_group <= 0) return MM_ERR_BAD_ARG;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const int warps_per_block = 8;
  int64_t need = (M + warps_per_block - 1) / warps_per_block;
  int64_t cap = (int64_t)sm_count() * 8;
  int grid = (int)(need < cap ? need : cap);
  const int nvec = (int)(D / 8);
  const __nv_bfloat16* xp = reinterpret_cast<const __nv_bfloat16*>(x);
  __nv_bfloat16* yp = reinterpret_cast<__nv_bfloat16*>(y);
  const __nv_bfloat16* gp = reinterpret_cast<const __nv_bfloat16*>(gamma);
  const __nv_bfloat16* bp = reinterpret_cast<const __nv_bfloat16*>(beta);
#define MM_LN_LAUNCH(VPL)                                                                                   \
  layernorm_kernel<VPL><<<grid, 256, 0, st>>>(xp, yp, gp, bp, M, (int)D, eps, rows_in_group, rows_out_group, \
                                              in_row_offset)
  if (nvec <= 32) MM_LN_LAUNCH(1);
  else if (nvec <= 64) MM_LN_LAUNCH(2);
  else if (nvec <= 128) MM_LN_LAUNCH(4);
  else MM_LN_LAUNCH(8);
#undef MM_LN_LAUNCH
  MM_CUDA_CHECK(cudaGetLastError());
  return MM_OK;
}

// ---- mm_fold_layernorm: weight preparation for the LayerNorm fold of mm_gemm_ln_bias_act (load time) --------------
// One thread per weight row n (rows are independent; the error feedback is serial along k):
//   wg[k]  = W[n,k]*gamma[k] - mean_k(W[n,:]*gamma)          (zero row sum: the LN mean term cancels inside the MMA)
//   t      = wg[k] + err;  q = bf16(t);  err = t - q         (sigma-delta rounding: the bf16 row sum stays < 1/2 ulp)
//   b'[n]  = bias[n] + sum_k W[n,k]*beta[k]
// Row means and the beta dot product are accumulated in double (load-time kernel, accuracy over speed).
namespace mm {
__global__ void __launch_bounds__(128) fold_layernorm_kernel(const __nv_bfloat16* __restrict__ W, int64_t ldw,
                                                             const __nv_bfloat16* __restrict__ bias,
                                                             const __nv_bfloat16* __restrict__ gamma,
                                                             const __nv_bfloat16* __restrict__ beta,
                                                             __nv_bfloat16* __restrict__ Wout, int64_t ldo,
                                                             __nv_bfloat16* __restrict__ bout, int N, int K) {
  extern __shared__ float s_gb[];   // gamma[K] | beta[K]
  for (int k = threadIdx.x; k < K; k += blockDim.x) {
    s_gb[k] = __bfloat162float(gamma[k]);
    s_gb[K + k] = __bfloat162float(beta[k]);
  }
  __syncthreads();
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= N) return;
  const __nv_bfloat16* w = W + (int64_t)n * ldw;
  double sum = 0.0, dot = 0.0;
  for (int k = 0; k < K; ++k) {
    const float x = __bfloat162float(w[k]);
    sum += (double)(x * s_gb[k]);
    dot += (double)x * (double)s_gb[K + k];
  }
  const float mean = (float)(sum / (double)K);
  float err = 0.f;
  __nv_bfloat16* o = Wout + (int64_t)n * ldo;
  for (int k = 0; k < K; ++k) {
    const float t = (__bfloat162float(w[k]) * s_gb[k] - mean) + err;
    const __nv_bfloat16 q = __float2bfloat16_rn(t);
    o[k] = q;
    err = t - __bfloat162float(q);
  }
  bout[n] = __float2bfloat16_rn((float)(dot + (bias ? (double)__bfloat162float(bias[n]) : 0.0)));
}
}  // namespace mm

extern "C" int mm_fold_layernorm(const void* W, int64_t ldw, const void* bias, const void* gamma, const void* beta,
                                 void* W_out, int64_t ldo, void* bias_out, int64_t N, int64_t K, void* stream) {
  using namespace mm;
  if (N < 0 || K <= 0 || K > 12000 || ldw < K || ldo < K || N > INT32_MAX / 2) return MM_ERR_BAD_ARG;
  if (N == 0) return MM_OK;
  if (!W || !gamma || !beta || !W_out || !bias_out) return MM_ERR_BAD_ARG;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const int threads = 128;
  const int grid = (int)((N + threads - 1) / threads);
  const size_t smem = 2 * (size_t)K * sizeof(float);
  if (smem > 48 * 1024)
    MM_CUDA_CHECK(cudaFuncSetAttribute(fold_layernorm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  fold_layernorm_kernel<<<grid, threads, smem, st>>>(
      reinterpret_cast<const __nv_bfloat16*>(W), ldw, reinterpret_cast<const __nv_bfloat16*>(bias),
      reinterpret_cast<const __nv_bfloat16*>(gamma), reinterpret_cast<const __nv_bfloat16*>(beta),
      reinterpret_cast<__nv_bfloat16*>(W_out), ldo, reinterpret_cast<__nv_bfloat16*>(bias_out), (int)N, (int)K);
  MM_CUDA_CHECK(cudaGetLastError());
  return MM_OK;
}
