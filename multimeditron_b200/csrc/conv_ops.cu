// conv_ops.cu — the byte-moving side of the MoE gate's ResNet-50 on the tcgen05 GEMM.
//
// The reference gate (src/multimeditron/model/modalities/moe/gating.py:60-65,85) is torchvision's ResNet-50 on every
// image.  Here every convolution is  rows-matrix x weightᵀ  on mm_gemm_bias_act (eval-mode BatchNorm folded into the
// weight and bias at load time, ReLU / residual-add-then-ReLU in the GEMM epilogue, activations kept NHWC =
// [B*H*W, C] bf16 row-major, which IS the GEMM's A layout): 1x1 stride-1 convolutions need nothing else; the 3x3,
// the strided 1x1 and the 7x7 stem convolutions get their A operand from the row-staging kernels below (an explicit
// im2col matrix, like mm_patch_embed_ws), and the two pooling layers are plain HBM-bound kernels.
//   mm_conv_rows_nhwc   [B,H,W,C] bf16 -> [B*Ho*Wo, Kpad], k = (ky*kw + kx)*C + c, zero padding outside the image
//   mm_conv_rows_stem   [B,3,H,W] fp32|bf16 (the experts' pixel tensor) -> [B*Ho*Wo, Kpad], k = (ky*kw + kx)*3 + c
//   mm_maxpool_nhwc     3x3 / stride 2 / pad 1 (torchvision resnet.py maxpool), bf16
//   mm_avgpool_nhwc     global average over H*W -> [B, C] bf16 (fp32 accumulation, one rounding)
#include "common.cuh"

namespace mm {

// 32-bit index arithmetic throughout (the host wrapper bounds the problem): with 64-bit divisions the kernel was
// instruction-bound at ~2.5 TB/s of real traffic.
__global__ void __launch_bounds__(256) conv_rows_nhwc_kernel(const __nv_bfloat16* __restrict__ x,
                                                             __nv_bfloat16* __restrict__ out, int B, int H, int W, int C,
                                                             int kh, int kw, int stride, int pad, int Ho, int Wo,
                                                             int Kpad) {
  const uint32_t vpr = (uint32_t)Kpad >> 3;        // 16-byte vectors per output row
  const uint32_t cv = (uint32_t)C >> 3;            // vectors per pixel
  const uint32_t kvalid = (uint32_t)(kh * kw) * cv;
  const uint32_t total = (uint32_t)B * Ho * Wo * vpr;
  const int4* xin = reinterpret_cast<const int4*>(x);
  int4* o4 = reinterpret_cast<int4*>(out);
  for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const uint32_t row = i / vpr, v = i - row * vpr;
    int4 val = make_int4(0, 0, 0, 0);
    if (v < kvalid) {
      const uint32_t tap = v / cv, c8 = v - tap * cv;
      const uint32_t ky = tap / (uint32_t)kw, kx = tap - ky * kw;
      const uint32_t t = row / (uint32_t)Wo, wo = row - t * Wo;
      const uint32_t b = t / (uint32_t)Ho, ho = t - b * Ho;
      const int hi = (int)(ho * stride + ky) - pad, wi = (int)(wo * stride + kx) - pad;
      if (hi >= 0 && hi < H && wi >= 0 && wi < W)
        val = __ldg(xin + ((size_t)(b * H + hi) * W + wi) * cv + c8);
    }
    o4[i] = val;     // out is [rows, Kpad] contiguous: vector i of the flat matrix
  }
}

// Stem rows through shared memory: one CTA = kStemPix consecutive output pixels of ONE output row (b, ho).  The
// kh input rows x (kStemPix-1)*stride + kw columns x 3 channels they touch are loaded once with coalesced reads
// (bf16 in smem), then every thread assembles 16-byte pieces of the output rows from smem and stores them coalesced
// (the scalar-gather version spent 1.9 ms on the 1.2 GB stem matrix of 256 x 224^2 images: 0.7 TB/s).
constexpr int kStemPix = 32;
template <typename TIn>
__global__ void __launch_bounds__(256) conv_rows_stem_kernel(const TIn* __restrict__ x, __nv_bfloat16* __restrict__ out,
                                                             int B, int H, int W, int kh, int kw, int stride, int pad,
                                                             int Ho, int Wo, int Kpad) {
  extern __shared__ __nv_bfloat16 s_in[];           // [kh][3][span], span = (kStemPix-1)*stride + kw; then the k table
  const int span = (kStemPix - 1) * stride + kw;
  int16_t* s_tab = reinterpret_cast<int16_t*>(s_in + kh * 3 * span);   // k -> offset of (ky, c, kx) in s_in, -1 = padding
  for (int k = threadIdx.x; k < Kpad; k += blockDim.x) {
    int off = -1;
    if (k < kh * kw * 3) {
      const int tap = k / 3, c = k - tap * 3;
      const int ky = tap / kw, kx = tap - ky * kw;
      off = (ky * 3 + c) * span + kx;
    }
    s_tab[k] = (int16_t)off;
  }
  const int tiles_w = (Wo + kStemPix - 1) / kStemPix;
  const int tile = blockIdx.x % tiles_w;
  const int ho = (blockIdx.x / tiles_w) % Ho, b = blockIdx.x / (tiles_w * Ho);
  const int wo0 = tile * kStemPix;
  const int wi0 = wo0 * stride - pad, hi0 = ho * stride - pad;
  const int64_t plane = (int64_t)H * W;
  for (int i = threadIdx.x; i < kh * 3 * span; i += blockDim.x) {
    const int col = i % span, rc = i / span;
    const int c = rc % 3, ky = rc / 3;
    const int hi = hi0 + ky, wi = wi0 + col;
    float v = 0.f;
    if (hi >= 0 && hi < H && wi >= 0 && wi < W) v = (float)x[((int64_t)b * 3 + c) * plane + (int64_t)hi * W + wi];
    s_in[i] = __float2bfloat16_rn(v);
  }
  __syncthreads();
  const int vpr = Kpad >> 3;
  const int npix = min(kStemPix, Wo - wo0);
  const uint16_t* su = reinterpret_cast<const uint16_t*>(s_in);
  for (int i = threadIdx.x; i < npix * vpr; i += blockDim.x) {
    const int pix = i / vpr, v = i - pix * vpr;
    uint32_t w[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      uint32_t lohi[2];
#pragma unroll
      for (int hh = 0; hh < 2; ++hh) {
        const int off = s_tab[v * 8 + e * 2 + hh];
        lohi[hh] = off >= 0 ? (uint32_t)su[off + pix * stride] : 0u;
      }
      w[e] = lohi[0] | (lohi[1] << 16);
    }
    const int64_t row = ((int64_t)b * Ho + ho) * Wo + wo0 + pix;
    reinterpret_cast<int4*>(out + row * Kpad)[v] = make_int4((int)w[0], (int)w[1], (int)w[2], (int)w[3]);
  }
}

__device__ __forceinline__ uint32_t max_bf16x2(uint32_t a, uint32_t b) {
  uint32_t d;
  asm("max.bf16x2 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b));
  return d;
}

__global__ void __launch_bounds__(256) maxpool_nhwc_kernel(const __nv_bfloat16* __restrict__ x,
                                                           __nv_bfloat16* __restrict__ out, int B, int H, int W, int C,
                                                           int Ho, int Wo) {
  const int cv = C >> 3;
  const int64_t total = (int64_t)B * Ho * Wo * cv;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int c8 = (int)(i % cv);
    const int64_t px = i / cv;
    const int wo = (int)(px % Wo);
    const int64_t t = px / Wo;
    const int ho = (int)(t % Ho), b = (int)(t / Ho);
    const uint32_t ninf = 0xff80ff80u;   // (-inf, -inf) bf16x2: padding never wins (torch pads max-pool with -inf)
    uint32_t m0 = ninf, m1 = ninf, m2 = ninf, m3 = ninf;
#pragma unroll
    for (int ky = 0; ky < 3; ++ky) {
      const int hi = ho * 2 + ky - 1;
      if (hi < 0 || hi >= H) continue;
#pragma unroll
      for (int kx = 0; kx < 3; ++kx) {
        const int wi = wo * 2 + kx - 1;
        if (wi < 0 || wi >= W) continue;
        const int4 v = __ldg(reinterpret_cast<const int4*>(x + (((int64_t)b * H + hi) * W + wi) * C) + c8);
        m0 = max_bf16x2(m0, (uint32_t)v.x), m1 = max_bf16x2(m1, (uint32_t)v.y);
        m2 = max_bf16x2(m2, (uint32_t)v.z), m3 = max_bf16x2(m3, (uint32_t)v.w);
      }
    }
    reinterpret_cast<int4*>(out + px * C)[c8] = make_int4((int)m0, (int)m1, (int)m2, (int)m3);
  }
}

// one warp per (image, 8-channel vector group of 32 lanes): lanes stride over channels, loop over the HW pixels
__global__ void __launch_bounds__(256) avgpool_nhwc_kernel(const __nv_bfloat16* __restrict__ x,
                                                           __nv_bfloat16* __restrict__ out, int B, int HW, int C) {
  const int pairs = C >> 1;
  const int64_t total = (int64_t)B * pairs;
  const float inv = 1.0f / (float)HW;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int b = (int)(i / pairs), c2 = (int)(i - (int64_t)b * pairs);
    const uint32_t* p = reinterpret_cast<const uint32_t*>(x + (int64_t)b * HW * C) + c2;
    float s0 = 0.f, s1 = 0.f;
    for (int q = 0; q < HW; ++q) {
      const uint32_t v = __ldg(p + (int64_t)q * pairs);
      s0 += bf16_lo(v), s1 += bf16_hi(v);
    }
    reinterpret_cast<uint32_t*>(out + (int64_t)b * C)[c2] = pack_bf16x2(s0 * inv, s1 * inv);
  }
}

static int grid_for(int64_t items) {
  const int64_t need = (items + 255) / 256, cap = (int64_t)sm_count() * 16;
  return (int)(need < cap ? (need > 0 ? need : 1) : cap);
}
static bool al16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

}  // namespace mm

using namespace mm;

extern "C" int mm_conv_rows_nhwc(const void* x, void* out, int64_t B, int64_t H, int64_t W, int64_t C, int64_t kh,
                                 int64_t kw, int64_t stride, int64_t pad, int64_t Kpad, void* stream) {
  if (B < 0 || H <= 0 || W <= 0 || C <= 0 || C % 8 || kh <= 0 || kw <= 0 || stride <= 0 || pad < 0) return MM_ERR_BAD_ARG;
  if (Kpad % 8 || Kpad < kh * kw * C) return MM_ERR_BAD_ARG;
  const int64_t Ho = (H + 2 * pad - kh) / stride + 1, Wo = (W + 2 * pad - kw) / stride + 1;
  if (Ho <= 0 || Wo <= 0) return MM_ERR_BAD_ARG;
  if (B == 0) return MM_OK;
  if (!x || !out || !al16(x) || !al16(out)) return MM_ERR_BAD_ARG;
  const int64_t items = B * Ho * Wo * (Kpad / 8);
  if (items >= ((int64_t)1 << 31) || B * H * W * (C / 8) >= ((int64_t)1 << 31)) return MM_ERR_BAD_ARG;   // 32-bit indices
  conv_rows_nhwc_kernel<<<grid_for(items), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const __nv_bfloat16*>(x), reinterpret_cast<__nv_bfloat16*>(out), (int)B, (int)H, (int)W, (int)C,
      (int)kh, (int)kw, (int)stride, (int)pad, (int)Ho, (int)Wo, (int)Kpad);
  MM_CUDA_CHECK(cudaGetLastError());
  return MM_OK;
}

extern "C" int mm_conv_rows_stem(const void* x, int x_dtype, void* out, int64_t B, int64_t H, int64_t W, int64_t kh,
                                 int64_t kw, int64_t stride, int64_t pad, int64_t Kpad, void* stream) {
  if (B < 0 || H <= 0 || W <= 0 || kh <= 0 || kw <= 0 || stride <= 0 || pad < 0) return MM_ERR_BAD_ARG;
  if (Kpad % 8 || Kpad < kh * kw * 3 || (x_dtype != MM_BF16 && x_dtype != MM_F32)) return MM_ERR_BAD_ARG;
  const int64_t Ho = (H + 2 * pad - kh) / stride + 1, Wo = (W + 2 * pad - kw) / stride + 1;
  if (Ho <= 0 || Wo <= 0) return MM_ERR_BAD_ARG;
  if (B == 0) return MM_OK;
  if (!x || !out || !al16(out)) return MM_ERR_BAD_ARG;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const int64_t tiles_w = (Wo + kStemPix - 1) / kStemPix;
  const int64_t blocks = B * Ho * tiles_w;
  if (blocks > 2147483647LL) return MM_ERR_BAD_ARG;
  const size_t smem = (size_t)kh * 3 * ((kStemPix - 1) * stride + kw) * sizeof(__nv_bfloat16) + (size_t)Kpad * sizeof(int16_t);
  if (smem > 48 * 1024) return MM_ERR_BAD_ARG;
  if (x_dtype == MM_F32)
    conv_rows_stem_kernel<float><<<(unsigned)blocks, 256, smem, st>>>(reinterpret_cast<const float*>(x),
                                                                     reinterpret_cast<__nv_bfloat16*>(out), (int)B, (int)H,
                                                                     (int)W, (int)kh, (int)kw, (int)stride, (int)pad,
                                                                     (int)Ho, (int)Wo, (int)Kpad);
  else
    conv_rows_stem_kernel<__nv_bfloat16><<<(unsigned)blocks, 256, smem, st>>>(
        reinterpret_cast<const __nv_bfloat16*>(x), reinterpret_cast<__nv_bfloat16*>(out), (int)B, (int)H, (int)W,
        (int)kh, (int)kw, (int)stride, (int)pad, (int)Ho, (int)Wo, (int)Kpad);
  MM_CUDA_CHECK(cudaGetLastError());
  return MM_OK;
}

extern "C" int mm_maxpool_nhwc(const void* x, void* out, int64_t B, int64_t H, int64_t W, int64_t C, void* stream) {
  if (B < 0 || H <= 0 || W <= 0 || C <= 0 || C % 8) return MM_ERR_BAD_ARG;
  if (B == 0) return MM_OK;
  if (!x || !out || !al16(x) || !al16(out)) return MM_ERR_BAD_ARG;
  const int64_t Ho = (H + 2 - 3) / 2 + 1, Wo = (W + 2 - 3) / 2 + 1;
  maxpool_nhwc_kernel<<<grid_for(B * Ho * Wo * (C / 8)), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const __nv_bfloat16*>(x), reinterpret_cast<__nv_bfloat16*>(out), (int)B, (int)H, (int)W, (int)C,
      (int)Ho, (int)Wo);
  MM_CUDA_CHECK(cudaGetLastError());
  return MM_OK;
}

extern "C" int mm_avgpool_nhwc(const void* x, void* out, int64_t B, int64_t HW, int64_t C, void* stream) {
  if (B < 0 || HW <= 0 || C <= 0 || C % 2) return MM_ERR_BAD_ARG;
  if (B == 0) return MM_OK;
  if (!x || !out) return MM_ERR_BAD_ARG;
  avgpool_nhwc_kernel<<<grid_for(B * (C / 2)), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const __nv_bfloat16*>(x), reinterpret_cast<__nv_bfloat16*>(out), (int)B, (int)HW, (int)C);
  MM_CUDA_CHECK(cudaGetLastError());
  return MM_OK;
}
