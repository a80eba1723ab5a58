// preprocess.cu — device image pre-processing for a ragged batch of raw uint8 HWC images.
// Replaces, bit for bit, what `ImageProcessor.process` (src/multimeditron/model/modalities/image_modality.py:96)
// gets from transformers 5.5.0 `CLIPImageProcessor(TorchvisionBackend)` (HF:image_processing_backends.py:372-420):
//   uint8 antialiased bicubic resize (ATen UpSampleKernel.cpp: `_compute_index_ranges_int16_weights` +
//   the separable int16-weight passes, horizontal first, uint8 intermediate)  -> center crop
//   -> (float(u8) - mean*255) / (std*255)  [-> bf16 RNE, the cast HF:modeling_clip.py:208-209 performs].
// Byte/integer work, HBM-bound; only the crop window is ever computed.  Three launches per batch:
//   resize_weights_kernel   one CTA per (image, axis): double-precision taps -> int16 weights + shift
//   resize_rows_kernel      horizontal pass   src[H,W,3] -> tmp[rows needed, S, 3]   (row staged in smem)
//   resize_cols_norm_kernel vertical pass + normalise -> out[B,3,S,S] fp32 | bf16 (coalesced per plane)
#include "common.cuh"

namespace mm {

// image descriptor, 8 x int64 per image (host-built, see include/mm_b200.h)
enum { D_SRC = 0, D_H = 1, D_W = 2, D_OH = 3, D_OW = 4, D_TOP = 5, D_LEFT = 6, D_TMP = 7, D_N = 8 };

struct PreWs {      // carve-up of the caller's workspace
  int32_t* meta;    // [B][2][4]  {shift, first src index needed, one-past-last src index needed, -}
  int32_t* bounds;  // [B][2][S][2] {xmin, xsize} for the S output indices inside the crop window
  int16_t* w;       // [B][2][S][kmax]
};

__host__ __device__ inline PreWs carve(void* ws, int64_t B, int64_t S, int64_t kmax) {
  PreWs p;
  p.meta = (int32_t*)ws;
  p.bounds = p.meta + B * 2 * 4;
  p.w = (int16_t*)(p.bounds + B * 2 * S * 2);
  return p;
}

// a = -0.5 cubic, evaluated with explicitly rounded operations (no FMA contraction) like the host code
__device__ __forceinline__ double cubic_aa(double x) {
  const double a = -0.5;
  x = fabs(x);
  if (x < 1.0) {
    double t = __dsub_rn(__dmul_rn(a + 2.0, x), a + 3.0);
    return __dadd_rn(__dmul_rn(__dmul_rn(t, x), x), 1.0);
  }
  if (x < 2.0) {
    double t = __dmul_rn(__dsub_rn(x, 5.0), x);
    t = __dmul_rn(__dadd_rn(t, 8.0), x);
    return __dmul_rn(__dsub_rn(t, 4.0), a);
  }
  return 0.0;
}

struct Taps { int lo, n; double tot; };

__device__ __forceinline__ Taps taps_of(int i, double scale, double support, double invscale, int in_size, int ksize) {
  Taps t;
  double center = __dmul_rn(scale, (double)i + 0.5);
  t.lo = max((int)(__dadd_rn(__dsub_rn(center, support), 0.5)), 0);
  int hi = min((int)(__dadd_rn(__dadd_rn(center, support), 0.5)), in_size);
  t.n = max(0, min(hi - t.lo, ksize));
  double tot = 0.0;
  for (int j = 0; j < t.n; ++j)
    tot = __dadd_rn(tot, cubic_aa(__dmul_rn(__dadd_rn(__dsub_rn((double)(j + t.lo), center), 0.5), invscale)));
  t.tot = tot;
  return t;
}

__device__ __forceinline__ double tap_w(const Taps& t, int j, double scale, int i, double invscale) {
  double center = __dmul_rn(scale, (double)i + 0.5);
  double w = cubic_aa(__dmul_rn(__dadd_rn(__dsub_rn((double)(j + t.lo), center), 0.5), invscale));
  return t.tot != 0.0 ? __ddiv_rn(w, t.tot) : w;
}

// grid (B, 2): axis 0 = vertical (H -> out_h, window crop_top..+S), axis 1 = horizontal
__global__ void __launch_bounds__(256) resize_weights_kernel(const int64_t* __restrict__ desc, void* ws, int B, int S,
                                                             int kmax) {
  const int img = blockIdx.x, axis = blockIdx.y;
  const int64_t* d = desc + (int64_t)img * D_N;
  const int in_size = (int)(axis == 0 ? d[D_H] : d[D_W]);
  const int out_size = (int)(axis == 0 ? d[D_OH] : d[D_OW]);
  const int win0 = (int)(axis == 0 ? d[D_TOP] : d[D_LEFT]);
  PreWs p = carve(ws, B, S, kmax);
  const double scale = (double)in_size / (double)out_size;
  const double support = scale >= 1.0 ? 2.0 * scale : 2.0;
  const int ksize = (int)ceilf((float)support) * 2 + 1;
  const double invscale = scale >= 1.0 ? 1.0 / scale : 1.0;

  // pass 1: the largest normalised weight over EVERY output index of the axis fixes the int16 shift
  double wmax = 0.0;
  for (int i = threadIdx.x; i < out_size; i += blockDim.x) {
    Taps t = taps_of(i, scale, support, invscale, in_size, ksize);
    for (int j = 0; j < t.n; ++j) wmax = fmax(wmax, tap_w(t, j, scale, i, invscale));
  }
  __shared__ double red[8];
  __shared__ int s_shift;
  for (int o = 16; o; o >>= 1) wmax = fmax(wmax, __shfl_xor_sync(0xffffffffu, wmax, o));
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = wmax;
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int k = 1; k < 8; ++k) wmax = fmax(wmax, red[k]);
    int shift = 0;
    for (; shift < 22; ++shift)
      if ((int)(0.5 + wmax * (double)(1 << (shift + 1))) >= (1 << 15)) break;
    s_shift = shift;
    Taps a = taps_of(win0, scale, support, invscale, in_size, ksize);
    Taps b = taps_of(win0 + S - 1, scale, support, invscale, in_size, ksize);
    int32_t* m = p.meta + ((int64_t)img * 2 + axis) * 4;
    m[0] = shift; m[1] = a.lo; m[2] = b.lo + b.n; m[3] = ksize;
  }
  __syncthreads();
  const double q = (double)(1 << s_shift);
  // pass 2: int16 taps of the S indices inside the crop window
  for (int r = threadIdx.x; r < S; r += blockDim.x) {
    const int i = win0 + r;
    Taps t = taps_of(i, scale, support, invscale, in_size, ksize);
    int32_t* bd = p.bounds + (((int64_t)img * 2 + axis) * S + r) * 2;
    bd[0] = t.lo; bd[1] = t.n;
    int16_t* w = p.w + (((int64_t)img * 2 + axis) * S + r) * kmax;
    for (int j = 0; j < kmax; ++j) {
      double v = j < t.n ? __dmul_rn(tap_w(t, j, scale, i, invscale), q) : 0.0;
      w[j] = (int16_t)(v < 0.0 ? (v - 0.5) : (v + 0.5));   // C truncation after the +-0.5, as ATen does
    }
  }
}

constexpr int ROWS_PER_CTA = 4;

// Horizontal pass.  grid (ceil(maxH / ROWS_PER_CTA), B); CTA stages the needed byte span of 4 source rows in smem
// (aligned 4-byte loads), then thread <-> (row, output column) runs the int16 taps for the 3 channels.
__global__ void __launch_bounds__(256) resize_rows_kernel(const uint8_t* __restrict__ src, const int64_t* __restrict__ desc,
                                                          const void* ws, uint8_t* __restrict__ tmp, int B, int S, int kmax,
                                                          int row_pitch /* smem bytes per staged row, multiple of 4 */) {
  extern __shared__ __align__(16) uint8_t srow[];
  const int img = blockIdx.y;
  const int64_t* d = desc + (int64_t)img * D_N;
  const int H = (int)d[D_H], W = (int)d[D_W];
  PreWs p = carve(const_cast<void*>(ws), B, S, kmax);
  const int32_t* mv = p.meta + ((int64_t)img * 2 + 0) * 4;   // vertical axis: which source rows are needed at all
  const int32_t* mh = p.meta + ((int64_t)img * 2 + 1) * 4;
  const int y0 = blockIdx.x * ROWS_PER_CTA;
  if (y0 >= H || y0 >= mv[2] || y0 + ROWS_PER_CTA <= mv[1]) return;
  const int shift = mh[0], c_lo = mh[1], c_hi = mh[2];
  const uint8_t* base = src + d[D_SRC];
  for (int r = 0; r < ROWS_PER_CTA; ++r) {
    const int y = y0 + r;
    if (y >= H) break;
    const uint8_t* g0 = base + ((int64_t)y * W + c_lo) * 3;
    const uint8_t* g1 = base + ((int64_t)y * W + c_hi) * 3;
    const uintptr_t a0 = (uintptr_t)g0 & ~(uintptr_t)3, a1 = (uintptr_t)g1 & ~(uintptr_t)3;
    uint32_t* s = (uint32_t*)(srow + r * row_pitch);
    const int nw = (int)((a1 - a0) >> 2);
    for (int k = threadIdx.x; k < nw; k += blockDim.x) s[k] = __ldg((const uint32_t*)a0 + k);
    const int tail = (int)((uintptr_t)g1 - a1);              // < 4 bytes past the last whole word
    if (threadIdx.x < tail) srow[r * row_pitch + nw * 4 + threadIdx.x] = __ldg((const uint8_t*)a1 + threadIdx.x);
  }
  __syncthreads();
  const int32_t* bounds = p.bounds + ((int64_t)img * 2 + 1) * S * 2;
  const int16_t* wts = p.w + ((int64_t)img * 2 + 1) * S * kmax;
  uint8_t* out = tmp + d[D_TMP];
  // thread <-> output column: its taps are fetched once and applied to the CTA's 4 staged rows
  int roff[ROWS_PER_CTA];
#pragma unroll
  for (int r = 0; r < ROWS_PER_CTA; ++r)
    roff[r] = r * row_pitch + (int)(((uintptr_t)(base + ((int64_t)(y0 + r) * W + c_lo) * 3)) & 3) - c_lo * 3;
  const int nrows = min(ROWS_PER_CTA, H - y0);
  const int half = 1 << (shift - 1);
  for (int x = threadIdx.x; x < S; x += blockDim.x) {
    const int xmin = bounds[2 * x], n = bounds[2 * x + 1];
    const int16_t* w = wts + x * kmax;
    int acc[ROWS_PER_CTA][3];
#pragma unroll
    for (int r = 0; r < ROWS_PER_CTA; ++r) acc[r][0] = acc[r][1] = acc[r][2] = half;
    const int s0 = xmin * 3;
    for (int j = 0; j < n; ++j) {
      const int wj = w[j];
#pragma unroll
      for (int r = 0; r < ROWS_PER_CTA; ++r) {
        const uint8_t* s = srow + roff[r] + s0 + 3 * j;   // rows past the image read stale smem; never stored
        acc[r][0] += wj * s[0]; acc[r][1] += wj * s[1]; acc[r][2] += wj * s[2];
      }
    }
#pragma unroll
    for (int r = 0; r < ROWS_PER_CTA; ++r) {
      if (r < nrows) {
        uint8_t* o = out + ((int64_t)(y0 + r) * S + x) * 3;
        o[0] = (uint8_t)min(max(acc[r][0] >> shift, 0), 255);
        o[1] = (uint8_t)min(max(acc[r][1] >> shift, 0), 255);
        o[2] = (uint8_t)min(max(acc[r][2] >> shift, 0), 255);
      }
    }
  }
}

// Vertical pass + crop + normalise.  grid (ceil(S*S/256), B); thread <-> output pixel, three planes written coalesced.
template <typename OutT>
__global__ void __launch_bounds__(256) resize_cols_norm_kernel(const int64_t* __restrict__ desc, const void* ws,
                                                               const uint8_t* __restrict__ tmp, OutT* __restrict__ out, int B,
                                                               int S, int kmax, float m0, float m1, float m2, float s0,
                                                               float s1, float s2) {
  const int img = blockIdx.y;
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= S * S) return;
  const int y = t / S, x = t - y * S;
  const int64_t* d = desc + (int64_t)img * D_N;
  PreWs p = carve(const_cast<void*>(ws), B, S, kmax);
  const int shift = p.meta[((int64_t)img * 2 + 0) * 4];
  const int32_t* bd = p.bounds + (((int64_t)img * 2 + 0) * S + y) * 2;
  const int ymin = bd[0], n = bd[1];
  const int16_t* w = p.w + (((int64_t)img * 2 + 0) * S + y) * kmax;
  const uint8_t* s = tmp + d[D_TMP] + ((int64_t)ymin * S + x) * 3;
  int a0 = 1 << (shift - 1), a1 = a0, a2 = a0;
  for (int j = 0; j < n; ++j) {
    const int wj = w[j];
    const uint8_t* q = s + (int64_t)j * S * 3;
    a0 += wj * q[0]; a1 += wj * q[1]; a2 += wj * q[2];
  }
  const float v0 = (float)min(max(a0 >> shift, 0), 255), v1 = (float)min(max(a1 >> shift, 0), 255),
              v2 = (float)min(max(a2 >> shift, 0), 255);
  OutT* o = out + (int64_t)img * 3 * S * S + t;
  const int64_t plane = (int64_t)S * S;
  // tvF.normalize: sub then IEEE divide in fp32 (HF:image_processing_backends.py:292; torchvision normalize)
  o[0] = (OutT)__fdiv_rn(__fsub_rn(v0, m0), s0);
  o[plane] = (OutT)__fdiv_rn(__fsub_rn(v1, m1), s1);
  o[2 * plane] = (OutT)__fdiv_rn(__fsub_rn(v2, m2), s2);
}

// Vertical pass, 4 horizontally adjacent pixels per thread (S % 4 == 0): 12 contiguous bytes per tap arrive as three
// aligned 32-bit loads, the row's taps are shared by the 4 pixels, each plane gets one 8- or 16-byte store.
template <typename OutT>
__global__ void __launch_bounds__(256) resize_cols_norm4_kernel(const int64_t* __restrict__ desc, const void* ws,
                                                                const uint8_t* __restrict__ tmp, OutT* __restrict__ out, int B,
                                                                int S, int kmax, float m0, float m1, float m2, float s0,
                                                                float s1, float s2) {
  const int img = blockIdx.y;
  const int S4 = S >> 2;
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= S * S4) return;
  const int y = t / S4, x = (t - y * S4) * 4;
  const int64_t* d = desc + (int64_t)img * D_N;
  PreWs p = carve(const_cast<void*>(ws), B, S, kmax);
  const int shift = p.meta[((int64_t)img * 2 + 0) * 4];
  const int32_t* bd = p.bounds + (((int64_t)img * 2 + 0) * S + y) * 2;
  const int ymin = bd[0], n = bd[1];
  const int16_t* w = p.w + (((int64_t)img * 2 + 0) * S + y) * kmax;
  const uint32_t* s = reinterpret_cast<const uint32_t*>(tmp + d[D_TMP] + ((int64_t)ymin * S + x) * 3);
  const int pitch = (S * 3) >> 2;   // 32-bit words per intermediate row
  int acc[12];
#pragma unroll
  for (int i = 0; i < 12; ++i) acc[i] = 1 << (shift - 1);
  for (int j = 0; j < n; ++j) {
    const int wj = w[j];
    const uint32_t u0 = __ldg(s + j * pitch), u1 = __ldg(s + j * pitch + 1), u2 = __ldg(s + j * pitch + 2);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      acc[i] += wj * (int)((u0 >> (8 * i)) & 0xff);
      acc[4 + i] += wj * (int)((u1 >> (8 * i)) & 0xff);
      acc[8 + i] += wj * (int)((u2 >> (8 * i)) & 0xff);
    }
  }
  // acc[3*px + c]: byte order r0 g0 b0 r1 g1 b1 ...
  const float mean[3] = {m0, m1, m2}, stdv[3] = {s0, s1, s2};
  const int64_t plane = (int64_t)S * S;
  OutT* o = out + (int64_t)img * 3 * plane + (int64_t)y * S + x;
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    float v[4];
#pragma unroll
    for (int px = 0; px < 4; ++px)
      v[px] = __fdiv_rn(__fsub_rn((float)min(max(acc[3 * px + c] >> shift, 0), 255), mean[c]), stdv[c]);
    if (sizeof(OutT) == 4) {
      *reinterpret_cast<float4*>(o + c * plane) = make_float4(v[0], v[1], v[2], v[3]);
    } else {
      uint2 q;
      q.x = pack_bf16x2(v[0], v[1]), q.y = pack_bf16x2(v[2], v[3]);
      *reinterpret_cast<uint2*>(o + c * plane) = q;
    }
  }
}

}  // namespace mm

extern "C" int64_t mm_preprocess_workspace_bytes(int64_t B, int64_t S, int64_t kmax) {
  if (B < 0 || S <= 0 || kmax <= 0) return -1;
  int64_t bytes = B * 2 * 4 * 4 + B * 2 * S * 2 * 4 + B * 2 * S * kmax * 2;
  return (bytes + 255) & ~(int64_t)255;
}

extern "C" int mm_preprocess_images(const void* src, const int64_t* desc, int64_t B, int64_t S, int64_t kmax, int64_t max_h,
                                    int64_t max_w, void* ws, void* tmp, const float* mean255, const float* std255, void* out,
                                    int out_dtype, void* stream) {
  using namespace mm;
  if (B == 0) return MM_OK;
  if (!src || !desc || !ws || !tmp || !out || !mean255 || !std255) return MM_ERR_BAD_ARG;
  if (B < 0 || B > 65535 || S <= 0 || S > 4096 || kmax < 5 || kmax > 1024 || max_h <= 0 || max_w <= 0) return MM_ERR_BAD_ARG;
  if (((uintptr_t)src & 3) || (out_dtype != MM_BF16 && out_dtype != MM_F32)) return MM_ERR_BAD_ARG;
  const int row_pitch = (int)((max_w * 3 + 8 + 3) & ~(int64_t)3);
  const size_t smem = (size_t)row_pitch * ROWS_PER_CTA;
  if (smem > 200 * 1024) return MM_ERR_BAD_ARG;
  cudaStream_t st = (cudaStream_t)stream;
  if (smem > 48 * 1024)
    MM_CUDA_CHECK(cudaFuncSetAttribute(resize_rows_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  resize_weights_kernel<<<dim3((unsigned)B, 2), 256, 0, st>>>(desc, ws, (int)B, (int)S, (int)kmax);
  resize_rows_kernel<<<dim3((unsigned)((max_h + ROWS_PER_CTA - 1) / ROWS_PER_CTA), (unsigned)B), 128, smem, st>>>(
      (const uint8_t*)src, desc, ws, (uint8_t*)tmp, (int)B, (int)S, (int)kmax, row_pitch);
  const bool vec4 = (S % 4 == 0) && (((uintptr_t)tmp | (uintptr_t)out) & 15) == 0;   // tmp offsets are multiples of S*3
  if (vec4) {
    dim3 g4((unsigned)((S * (S / 4) + 255) / 256), (unsigned)B);
    if (out_dtype == MM_F32)
      resize_cols_norm4_kernel<float><<<g4, 256, 0, st>>>(desc, ws, (const uint8_t*)tmp, (float*)out, (int)B, (int)S,
                                                          (int)kmax, mean255[0], mean255[1], mean255[2], std255[0],
                                                          std255[1], std255[2]);
    else
      resize_cols_norm4_kernel<__nv_bfloat16><<<g4, 256, 0, st>>>(desc, ws, (const uint8_t*)tmp, (__nv_bfloat16*)out,
                                                                  (int)B, (int)S, (int)kmax, mean255[0], mean255[1],
                                                                  mean255[2], std255[0], std255[1], std255[2]);
    MM_CUDA_CHECK(cudaGetLastError());
    return MM_OK;
  }
  dim3 g((unsigned)((S * S + 255) / 256), (unsigned)B);
  if (out_dtype == MM_F32)
    resize_cols_norm_kernel<float><<<g, 256, 0, st>>>(desc, ws, (const uint8_t*)tmp, (float*)out, (int)B, (int)S, (int)kmax,
                                                      mean255[0], mean255[1], mean255[2], std255[0], std255[1], std255[2]);
  else
    resize_cols_norm_kernel<__nv_bfloat16><<<g, 256, 0, st>>>(desc, ws, (const uint8_t*)tmp, (__nv_bfloat16*)out, (int)B,
                                                              (int)S, (int)kmax, mean255[0], mean255[1], mean255[2],
                                                              std255[0], std255[1], std255[2]);
  MM_CUDA_CHECK(cudaGetLastError());
  return MM_OK;
}
