// train_ops.cu — element-wise / layout kernels of the projector's training step (SURVEY §8(f) rank 2:
// MultiModalModelForCausalLM.freeze_for_alignment, model.py:310-326 -> only MLPProjector trains).
// The GEMMs of the backward (dgrad = dY·W, wgrad = dYᵀ·X) run on the tcgen05 GEMM of gemm_sm100.cu with
// explicitly transposed operands; everything here is HBM-bound byte work:
//   mm_gelu_fwd            a = GELU_erf(z)            (the same apply_act the fused GEMM epilogue uses: bit-identical)
//   mm_gelu_bwd_colsum     dz = da ⊙ GELU'(z), and bias_grad[n] = Σ_m dz[m,n]  (deterministic two-stage column sum)
//   mm_transpose           out[c,r] = in[r,c]         (bf16, 64x64 shared-memory tiles, 16-byte global accesses)
// Replaces (training regime): torch.nn.GELU forward/backward kernels and the implicit transposes inside cuBLAS's
// NT/TN GEMMs that autograd issues for nn.Linear (projectors/mlp.py:33-39).
#include "act.cuh"

namespace mm {

constexpr int kEwThreads = 256;

__global__ void __launch_bounds__(kEwThreads) gelu_fwd_kernel(const uint4* __restrict__ z, uint4* __restrict__ a,
                                                              int64_t nvec) {
  for (int64_t i = (int64_t)blockIdx.x * kEwThreads + threadIdx.x; i < nvec; i += (int64_t)gridDim.x * kEwThreads) {
    uint4 v = __ldg(z + i);
    uint32_t* w = reinterpret_cast<uint32_t*>(&v);
#pragma unroll
    for (int e = 0; e < 4; ++e)
      w[e] = pack_bf16x2(apply_act<MM_ACT_GELU_ERF>(bf16_lo(w[e])), apply_act<MM_ACT_GELU_ERF>(bf16_hi(w[e])));
    a[i] = v;
  }
}

// d/dz [ z Φ(z) ] = Φ(z) + z φ(z)   (torch.nn.GELU(approximate='none') backward, fp32 math, one bf16 rounding)
__device__ __forceinline__ float gelu_grad(float z) {
  const float cdf = 0.5f * (1.0f + erff(z * 0.70710678118654752440f));
  const float pdf = 0.3989422804014327f * __expf(-0.5f * z * z);
  return fmaf(z, pdf, cdf);
}

// grid = (column tiles of 256, row slabs).  Thread (cx = tid % 32, ry = tid / 32) owns 8 columns and walks rows
// ry, ry+8, ... of its slab; partial column sums of the slab go to partial[slab][col] (fp32), summed by
// colsum_finish_kernel in slab order (deterministic).
__global__ void __launch_bounds__(kEwThreads)
    gelu_bwd_colsum_kernel(const __nv_bfloat16* __restrict__ da, const __nv_bfloat16* __restrict__ z,
                           __nv_bfloat16* __restrict__ dz, float* __restrict__ partial, int64_t M, int N,
                           int64_t rows_per_slab) {
  __shared__ float red[8][256 + 8];
  const int cx = threadIdx.x & 31, ry = threadIdx.x >> 5;
  const int col = blockIdx.x * 256 + cx * 8;
  const int64_t r0 = (int64_t)blockIdx.y * rows_per_slab;
  const int64_t r1 = (r0 + rows_per_slab < M) ? r0 + rows_per_slab : M;
  float acc[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) acc[e] = 0.f;
  if (col < N) {
    for (int64_t r = r0 + ry; r < r1; r += 8) {
      uint4 g = __ldg(reinterpret_cast<const uint4*>(da + r * N + col));
      uint32_t* gw = reinterpret_cast<uint32_t*>(&g);
      if (z) {
        const uint4 zz = __ldg(reinterpret_cast<const uint4*>(z + r * N + col));
        const uint32_t* zw = reinterpret_cast<const uint32_t*>(&zz);
#pragma unroll
        for (int e = 0; e < 4; ++e)
          gw[e] = pack_bf16x2(bf16_lo(gw[e]) * gelu_grad(bf16_lo(zw[e])), bf16_hi(gw[e]) * gelu_grad(bf16_hi(zw[e])));
        *reinterpret_cast<uint4*>(dz + r * N + col) = g;
      }
#pragma unroll
      for (int e = 0; e < 4; ++e) {   // the sum runs over the bf16 values as stored (what autograd's reduction sees)
        acc[2 * e] += bf16_lo(gw[e]);
        acc[2 * e + 1] += bf16_hi(gw[e]);
      }
    }
  }
#pragma unroll
  for (int e = 0; e < 8; ++e) red[ry][cx * 8 + e] = acc[e];
  __syncthreads();
  const int c = threadIdx.x;   // 256 threads -> 256 columns of the tile
  if (blockIdx.x * 256 + c < N) {
    float s = 0.f;
#pragma unroll
    for (int y = 0; y < 8; ++y) s += red[y][c];
    partial[(int64_t)blockIdx.y * N + blockIdx.x * 256 + c] = s;
  }
}

__global__ void __launch_bounds__(kEwThreads) colsum_finish_kernel(const float* __restrict__ partial,
                                                                    float* __restrict__ out, int slabs, int N) {
  const int c = blockIdx.x * kEwThreads + threadIdx.x;
  if (c >= N) return;
  float s = 0.f;
  for (int y = 0; y < slabs; ++y) s += partial[(int64_t)y * N + c];
  out[c] = s;
}

// 64 x 64 bf16 tiles: 16-byte loads along the input rows, 16-byte stores along the output rows; the tile is held in
// shared memory as 32-bit words with a +1 padding so that the transposed reads are conflict-free.
__global__ void __launch_bounds__(256) transpose_kernel(const __nv_bfloat16* __restrict__ in,
                                                        __nv_bfloat16* __restrict__ out, int64_t R, int64_t C) {
  __shared__ __nv_bfloat16 tile[64][64 + 2];
  const int64_t r0 = (int64_t)blockIdx.y * 64, c0 = (int64_t)blockIdx.x * 64;
  const int tx = threadIdx.x & 7, ty = threadIdx.x >> 3;   // 8 vectors per row, 32 rows per pass
#pragma unroll
  for (int pass = 0; pass < 2; ++pass) {
    const int r = ty + pass * 32;
    uint4 v = make_uint4(0, 0, 0, 0);
    if (r0 + r < R && c0 + tx * 8 < C) v = __ldg(reinterpret_cast<const uint4*>(in + (r0 + r) * C + c0 + tx * 8));
    const __nv_bfloat16* e = reinterpret_cast<const __nv_bfloat16*>(&v);
#pragma unroll
    for (int k = 0; k < 8; ++k) tile[r][tx * 8 + k] = e[k];
  }
  __syncthreads();
#pragma unroll
  for (int pass = 0; pass < 2; ++pass) {
    const int c = ty + pass * 32;          // output row = input column
    if (c0 + c < C && r0 + tx * 8 < R) {
      uint4 v;
      __nv_bfloat16* e = reinterpret_cast<__nv_bfloat16*>(&v);
#pragma unroll
      for (int k = 0; k < 8; ++k) e[k] = tile[tx * 8 + k][c];
      *reinterpret_cast<uint4*>(out + (c0 + c) * R + r0 + tx * 8) = v;
    }
  }
}

}  // namespace mm

using namespace mm;

static bool al16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

extern "C" int mm_gelu_fwd(const void* z, void* a, int64_t n, void* stream) {
  if (n < 0 || n % 8) return MM_ERR_BAD_ARG;
  if (n == 0) return MM_OK;
  if (!z || !a || !al16(z) || !al16(a)) return MM_ERR_BAD_ARG;
  const int64_t nvec = n / 8;
  int64_t blocks = (nvec + kEwThreads - 1) / kEwThreads;
  const int64_t cap = (int64_t)sm_count() * 8;
  if (blocks > cap) blocks = cap;
  gelu_fwd_kernel<<<(int)blocks, kEwThreads, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const uint4*>(z), reinterpret_cast<uint4*>(a), nvec);
  MM_CUDA_CHECK(cudaGetLastError());
  return MM_OK;
}

extern "C" int64_t mm_colsum_workspace_bytes(int64_t M, int64_t N) {
  if (M < 0 || N <= 0) return -1;
  const int64_t tiles = (N + 255) / 256;
  int64_t slabs = ((int64_t)sm_count() * 4 + tiles - 1) / tiles;
  if (slabs > (M + 7) / 8) slabs = (M + 7) / 8;
  if (slabs < 1) slabs = 1;
  return slabs * N * (int64_t)sizeof(float);
}

extern "C" int mm_gelu_bwd_colsum(const void* da, const void* z, void* dz, float* bias_grad, void* workspace,
                                  int64_t M, int64_t N, void* stream) {
  if (M < 0 || N <= 0 || N % 8 || N > INT32_MAX / 2) return MM_ERR_BAD_ARG;
  if (!bias_grad || !workspace || !da) return MM_ERR_BAD_ARG;
  if ((z != nullptr) != (dz != nullptr)) return MM_ERR_BAD_ARG;
  if (!al16(da) || !al16(z) || !al16(dz)) return MM_ERR_BAD_ARG;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const int64_t tiles = (N + 255) / 256;
  int64_t slabs = mm_colsum_workspace_bytes(M, N) / (N * (int64_t)sizeof(float));
  if (M == 0) {
    MM_CUDA_CHECK(cudaMemsetAsync(bias_grad, 0, N * sizeof(float), st));
    return MM_OK;
  }
  int64_t rows_per_slab = (M + slabs - 1) / slabs;
  rows_per_slab = (rows_per_slab + 7) / 8 * 8;
  slabs = (M + rows_per_slab - 1) / rows_per_slab;
  dim3 grid((unsigned)tiles, (unsigned)slabs);
  gelu_bwd_colsum_kernel<<<grid, kEwThreads, 0, st>>>(
      reinterpret_cast<const __nv_bfloat16*>(da), reinterpret_cast<const __nv_bfloat16*>(z),
      reinterpret_cast<__nv_bfloat16*>(dz), reinterpret_cast<float*>(workspace), M, (int)N, rows_per_slab);
  MM_CUDA_CHECK(cudaGetLastError());
  colsum_finish_kernel<<<(int)((N + kEwThreads - 1) / kEwThreads), kEwThreads, 0, st>>>(
      reinterpret_cast<const float*>(workspace), bias_grad, (int)slabs, (int)N);
  MM_CUDA_CHECK(cudaGetLastError());
  return MM_OK;
}

extern "C" int mm_transpose(const void* in, void* out, int64_t R, int64_t C, void* stream) {
  if (R < 0 || C < 0 || R % 8 || C % 8) return MM_ERR_BAD_ARG;
  if (R == 0 || C == 0) return MM_OK;
  if (!in || !out || !al16(in) || !al16(out)) return MM_ERR_BAD_ARG;
  if ((R + 63) / 64 > 65535) return MM_ERR_BAD_ARG;
  dim3 grid((unsigned)((C + 63) / 64), (unsigned)((R + 63) / 64));
  transpose_kernel<<<grid, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const __nv_bfloat16*>(in), reinterpret_cast<__nv_bfloat16*>(out), R, C);
  MM_CUDA_CHECK(cudaGetLastError());
  return MM_OK;
}
