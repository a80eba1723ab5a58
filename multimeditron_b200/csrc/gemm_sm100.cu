// gemm_sm100.cu — persistent warp-specialised tcgen05 GEMM:  C = act(A · Wᵀ + bias) [+ residual]
//
//   * operands bf16, K-major, staged by TMA (SWIZZLE_128B, 64-element K slabs) into a 4-deep smem ring
//   * tcgen05.mma cta_group::1, UMMA 128x256x16, fp32 accumulators in TMEM, double-buffered
//     (2 x 256 columns) so tile i+1's main loop overlaps tile i's epilogue
//   * epilogue warps: tcgen05.ld -> bias/activation in fp32 -> bf16 -> per-warp swizzled smem
//     transpose -> fully coalesced 16-byte row stores (optionally + residual, optional row remap /
//     scatter map so that drop-CLS, CLS-insert and the splice into inputs_embeds cost no extra pass)
//   * second instantiation (kGatherA): the A operand is gathered straight from NCHW pixels by four
//     producer warps (im2col-free implicit GEMM of the patch convolution).
//
// Warp roles (TMA variant, 384 threads): w0 TMA producer, w1 MMA issuer, w2 TMEM allocator,
// w4..w11 epilogue (TMEM lane quadrant = warp % 4; two warps per quadrant split the 256 columns).
// Gather variant (384 threads): w4..w7 epilogue, w8..w11 = A producers.
//
// Reference call sites replaced: see include/mm_b200.h (mm_gemm_bias_act, mm_patch_embed).
#include "common.cuh"

namespace mm {

constexpr int BM = 128;
constexpr int BN = 256;
constexpr int BK = 64;  // 64 bf16 = 128 B = one SWIZZLE_128B row
constexpr int kUmmaK = 16;
constexpr int kAStageBytes = BM * BK * 2;  // 16 KiB
constexpr int kBStageBytes = BN * BK * 2;  // 32 KiB
constexpr int kEpiWarpBytes = 32 * 128;    // 32 rows x 64 bf16
constexpr int kTmemCols = 512;
constexpr int kKTableMax = 3072 + 64;  // patch-embed k -> pixel offset table (ints)

constexpr int kMaxStages = 4;
struct GemmSmemTail {
  uint64_t full[kMaxStages];
  uint64_t empty[kMaxStages];
  uint64_t tmem_full[2];
  uint64_t tmem_empty[2];
  uint32_t tmem_base;
};

// TMA variant: 4 stages, 8 epilogue warps (2 per TMEM lane quadrant, each owning half of the 256 columns).
// Gather variant (patch embed): 3 stages, 4 epilogue warps, 4 A-producer warps, + 12.5 KiB static k-offset table.
template <bool kGatherA>
struct GemmCfg {
  static constexpr int kStages = kGatherA ? 3 : 4;
  static constexpr int kEpiWarps = kGatherA ? 4 : 8;
  static constexpr int kChunksPerWarp = (BN / 64) / (kEpiWarps / 4);
  static constexpr int kThreads = 128 + kEpiWarps * 32 + (kGatherA ? 128 : 0);
  static constexpr int kEpiBytes = kEpiWarps * kEpiWarpBytes;
  static constexpr int kSmemBytes = 1024 /*align slack*/ + kStages * (kAStageBytes + kBStageBytes) + kEpiBytes + 256;
};
static_assert(sizeof(GemmSmemTail) <= 256, "tail too large");
static_assert(GemmCfg<false>::kSmemBytes <= 232448, "smem budget");
static_assert(GemmCfg<true>::kSmemBytes + kKTableMax * 4 <= 232448, "smem budget (gather)");

struct GemmParams {
  __nv_bfloat16* C;
  const __nv_bfloat16* bias;
  const __nv_bfloat16* residual;
  const int32_t* out_row_map;
  int64_t ldc, ldr;
  int64_t n_out_rows;  // rows of C that may be written through out_row_map (bounds check)
  int M, N, K;
  int res_mod, res_row_off;
  int map_in, map_out, map_off;  // formula row map: out_row = (m / map_in) * map_out + map_off + m % map_in
  int act;
  int tiles_m, tiles_n;
  // gather-A (patch embed) only
  const void* pixels;
  int pixel_is_f32;
  int img, patch, grid;  // grid = img / patch
  int k_real;            // 3*patch*patch
};

__device__ __forceinline__ float fast_ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float fast_rcp(float x) {
  float y;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
// x * sigmoid(k*x) with one ex2 and one rcp (MUFU), k folded with log2(e)
__device__ __forceinline__ float x_sigmoid(float x, float k_log2e) { return x * fast_rcp(1.0f + fast_ex2(-k_log2e * x)); }

template <int ACT>
__device__ __forceinline__ float apply_act(float x) {
  if (ACT == MM_ACT_QUICK_GELU) {  // x * sigmoid(1.702 x)
    return x_sigmoid(x, 1.702f * 1.4426950408889634f);
  } else if (ACT == MM_ACT_GELU_TANH) {  // 0.5x(1+tanh(u)) == x*sigmoid(2u), u = sqrt(2/pi)(x + 0.044715x^3)
    const float u2 = x * (1.5957691216057308f + 0.07135481282636213f * x * x);  // 2u
    return x * fast_rcp(1.0f + fast_ex2(-1.4426950408889634f * u2));
  } else if (ACT == MM_ACT_GELU_ERF) {  // exact GELU; erf by Abramowitz-Stegun 7.1.26 (|err| <= 1.5e-7)
    const float ax = fabsf(x) * 0.70710678118654752440f;
    const float t = fast_rcp(fmaf(0.3275911f, ax, 1.0f));
    float poly = fmaf(t, 1.061405429f, -1.453152027f);
    poly = fmaf(poly, t, 1.421413741f);
    poly = fmaf(poly, t, -0.284496736f);
    poly = fmaf(poly, t, 0.254829592f);
    poly *= t;
    const float erf_abs = 1.0f - poly * fast_ex2(-1.4426950408889634f * ax * ax);
    return 0.5f * x * (1.0f + copysignf(erf_abs, x));
  }
  return x;
}

template <bool kGatherA, int ACT>
__global__ void __launch_bounds__(GemmCfg<kGatherA>::kThreads, 1)
    gemm_bf16_tn_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                        const GemmParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  using Cfg = GemmCfg<kGatherA>;
  constexpr int kStages = Cfg::kStages;
  constexpr int kEpiWarps = Cfg::kEpiWarps;
  uint8_t* sA = smem;
  uint8_t* sB = smem + kStages * kAStageBytes;
  uint8_t* sEpi = sB + kStages * kBStageBytes;
  GemmSmemTail* tail = reinterpret_cast<GemmSmemTail*>(sEpi + Cfg::kEpiBytes);
  __shared__ int s_koff[kGatherA ? kKTableMax : 1];

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int num_kb = (p.K + BK - 1) / BK;
  const int num_tiles = p.tiles_m * p.tiles_n;

  if (warp == 0 && lane == 0) {
    if (!kGatherA) tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < kStages; ++i) {
      mbar_init(&tail->full[i], kGatherA ? 1 + 128 : 1);
      mbar_init(&tail->empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tail->tmem_full[i], 1);
      mbar_init(&tail->tmem_empty[i], kEpiWarps);
    }
    fence_barrier_init();
  }
  if (warp == 2) {
    tmem_alloc(&tail->tmem_base, kTmemCols);
    tmem_relinquish();
  }
  if (kGatherA) {
    // k -> pixel offset table (c*img*img + ky*img + kx); -1 for the zero padding beyond k_real
    const int pp = p.patch * p.patch;
    for (int k = threadIdx.x; k < num_kb * BK; k += blockDim.x) {
      int off = -1;
      if (k < p.k_real) {
        int c = k / pp, r = k - c * pp;
        int ky = r / p.patch, kx = r - ky * p.patch;
        off = c * p.img * p.img + ky * p.img + kx;
      }
      s_koff[k] = off;
    }
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tail->tmem_base;

  if (warp == 0) {
    // =========================== TMA producer ===========================
    if (elect_one_sync()) {
      int stage = 0;
      uint32_t phase = 0;
      for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
        const int m0 = (t / p.tiles_n) * BM, n0 = (t % p.tiles_n) * BN;
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&tail->empty[stage], phase ^ 1);
          if (kGatherA) {
            mbar_arrive_expect_tx(&tail->full[stage], kBStageBytes);
          } else {
            mbar_arrive_expect_tx(&tail->full[stage], kAStageBytes + kBStageBytes);
            tma_load_2d(sA + stage * kAStageBytes, &tmA, &tail->full[stage], kb * BK, m0);
          }
          tma_load_2d(sB + stage * kBStageBytes, &tmB, &tail->full[stage], kb * BK, n0);
          if (++stage == kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    // =========================== MMA issuer (single thread) ===========================
    if (elect_one_sync()) {
      constexpr uint32_t idesc = umma_idesc_bf16(BM, BN, 0, 0);
      int stage = 0;
      uint32_t phase = 0;
      int as = 0;
      uint32_t aphase = 0;
      for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
        mbar_wait(&tail->tmem_empty[as], aphase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)(as * BN);
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&tail->full[stage], phase);
          tc_fence_after();
          const uint32_t a_addr = smem_u32(sA + stage * kAStageBytes);
          const uint32_t b_addr = smem_u32(sB + stage * kBStageBytes);
#pragma unroll
          for (int k = 0; k < BK / kUmmaK; ++k) {
            const uint64_t da = umma_desc_sw128(a_addr + k * kUmmaK * 2, 16, 1024);
            const uint64_t db = umma_desc_sw128(b_addr + k * kUmmaK * 2, 16, 1024);
            tc_mma_f16_ss(d_tmem, da, db, idesc, (kb | k) != 0 ? 1u : 0u);
          }
          tc_commit(&tail->empty[stage]);  // frees the smem slot when these MMAs retire
          if (++stage == kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
        tc_commit(&tail->tmem_full[as]);  // accumulator ready for the epilogue
        if (++as == 2) {
          as = 0;
          aphase ^= 1;
        }
      }
    }
  } else if (warp >= 4 && warp < 4 + kEpiWarps) {
    // =========================== epilogue ===========================
    const int q = warp & 3;             // TMEM lane quadrant (hardware: warp % 4)
    const int half = (warp - 4) >> 2;   // which group of column chunks this warp owns
    uint8_t* stg = sEpi + (warp - 4) * kEpiWarpBytes;
    int as = 0;
    uint32_t aphase = 0;
    for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
      const int m0 = (t / p.tiles_n) * BM, n0 = (t % p.tiles_n) * BN;
      // destination row of this lane's GEMM row
      const int my_m = m0 + q * 32 + lane;
      int my_out = -1;
      if (my_m < p.M) {
        if (p.out_row_map) {
          my_out = p.out_row_map[my_m];
          if ((int64_t)my_out >= p.n_out_rows) my_out = -1;  // never write outside C
        } else if (p.map_in > 0) {
          my_out = (my_m / p.map_in) * p.map_out + p.map_off + (my_m % p.map_in);
        } else {
          my_out = my_m;
        }
      }
      mbar_wait(&tail->tmem_full[as], aphase);
      tc_fence_after();
      const uint32_t t_row = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(as * BN);
#pragma unroll 1
      for (int cc = 0; cc < Cfg::kChunksPerWarp; ++cc) {
        const int c = half * Cfg::kChunksPerWarp + cc;
        const int col0 = n0 + c * 64;
        if (col0 >= p.N) break;  // warp-uniform
        uint32_t acc[64];
        tmem_ld_32x32b_x32(t_row + c * 64, acc);
        tmem_ld_32x32b_x32(t_row + c * 64 + 32, acc + 32);
        // while the TMEM loads are in flight: prefetch this chunk's bias and residual segments
        int4 bv[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          bv[j] = make_int4(0, 0, 0, 0);
          if (p.bias && col0 + j * 8 < p.N) bv[j] = __ldg(reinterpret_cast<const int4*>(p.bias + col0 + j * 8));
        }
        const int ch = lane & 7;
        const int col = col0 + ch * 8;
        int out_rows[8];
        int4 rv[8];
#pragma unroll
        for (int it = 0; it < 8; ++it) {
          const int r = it * 4 + (lane >> 3);
          out_rows[it] = __shfl_sync(0xffffffffu, my_out, r);
          if (col >= p.N) out_rows[it] = -1;
          rv[it] = make_int4(0, 0, 0, 0);
          if (p.residual && out_rows[it] >= 0) {
            const int gm = m0 + q * 32 + r;
            const int64_t rr = (int64_t)((p.res_mod > 0 ? gm % p.res_mod : gm) + p.res_row_off);
            rv[it] = __ldg(reinterpret_cast<const int4*>(p.residual + rr * p.ldr + col));
          }
        }
        tc_wait_ld();
        // bias + activation in fp32, on the bf16-rounded pre-activation (nn.Linear output rounding)
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const uint32_t bw[4] = {(uint32_t)bv[j].x, (uint32_t)bv[j].y, (uint32_t)bv[j].z, (uint32_t)bv[j].w};
          uint4 o;
          uint32_t* ow = reinterpret_cast<uint32_t*>(&o);
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            float x0 = __uint_as_float(acc[j * 8 + 2 * e]) + bf16_lo(bw[e]);
            float x1 = __uint_as_float(acc[j * 8 + 2 * e + 1]) + bf16_hi(bw[e]);
            if (ACT != MM_ACT_NONE) {
              const uint32_t r = pack_bf16x2(x0, x1);
              x0 = apply_act<ACT>(bf16_lo(r));
              x1 = apply_act<ACT>(bf16_hi(r));
            }
            ow[e] = pack_bf16x2(x0, x1);
          }
          *reinterpret_cast<uint4*>(stg + lane * 128 + ((j ^ (lane & 7)) << 4)) = o;
        }
        __syncwarp();
        // coalesced phase: 8 lanes cover one 128-B row segment, 4 rows per instruction
#pragma unroll
        for (int it = 0; it < 8; ++it) {
          const int r = it * 4 + (lane >> 3);
          uint4 v = *reinterpret_cast<const uint4*>(stg + r * 128 + ((ch ^ (r & 7)) << 4));
          if (out_rows[it] >= 0) {
            if (p.residual) {
              v.x = pack_bf16x2(bf16_lo(v.x) + bf16_lo((uint32_t)rv[it].x), bf16_hi(v.x) + bf16_hi((uint32_t)rv[it].x));
              v.y = pack_bf16x2(bf16_lo(v.y) + bf16_lo((uint32_t)rv[it].y), bf16_hi(v.y) + bf16_hi((uint32_t)rv[it].y));
              v.z = pack_bf16x2(bf16_lo(v.z) + bf16_lo((uint32_t)rv[it].z), bf16_hi(v.z) + bf16_hi((uint32_t)rv[it].z));
              v.w = pack_bf16x2(bf16_lo(v.w) + bf16_lo((uint32_t)rv[it].w), bf16_hi(v.w) + bf16_hi((uint32_t)rv[it].w));
            }
            *reinterpret_cast<uint4*>(p.C + (int64_t)out_rows[it] * p.ldc + col) = v;
          }
        }
        __syncwarp();
      }
      // all TMEM reads of this accumulator stage are complete (tc_wait_ld above)
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tail->tmem_empty[as]);
      if (++as == 2) {
        as = 0;
        aphase ^= 1;
      }
    }
  } else if (kGatherA && warp >= 4 + kEpiWarps) {
    // =========================== A producers: gather patches from NCHW pixels ===========================
    const int row = threadIdx.x - (4 + kEpiWarps) * 32;  // 0..127 : one A row (= one patch) per thread
    const int P = p.grid * p.grid;
    const float* px32 = reinterpret_cast<const float*>(p.pixels);
    const __nv_bfloat16* px16 = reinterpret_cast<const __nv_bfloat16*>(p.pixels);
    int stage = 0;
    uint32_t phase = 0;
    for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
      const int m0 = (t / p.tiles_n) * BM;
      const int m = m0 + row;
      int64_t base = -1;
      if (m < p.M) {
        const int b = m / P, pi = m - b * P;
        const int py = pi / p.grid, pxi = pi - py * p.grid;
        base = (int64_t)b * 3 * p.img * p.img + (int64_t)(py * p.patch) * p.img + pxi * p.patch;
      }
      for (int kb = 0; kb < num_kb; ++kb) {
        mbar_wait(&tail->empty[stage], phase ^ 1);
        const uint32_t dst = smem_u32(sA + stage * kAStageBytes) + row * 128;
#pragma unroll 2
        for (int j = 0; j < 8; ++j) {
          float f[8];
#pragma unroll
          for (int e = 0; e < 8; ++e) {
            const int off = s_koff[kb * BK + j * 8 + e];
            float v = 0.f;
            if (base >= 0 && off >= 0) v = p.pixel_is_f32 ? __ldg(px32 + base + off) : __bfloat162float(px16[base + off]);
            f[e] = v;
          }
          asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(dst + ((j ^ (row & 7)) << 4)),
                       "r"(pack_bf16x2(f[0], f[1])), "r"(pack_bf16x2(f[2], f[3])), "r"(pack_bf16x2(f[4], f[5])),
                       "r"(pack_bf16x2(f[6], f[7]))
                       : "memory");
        }
        fence_proxy_async_smem();  // generic-proxy writes -> visible to the tensor-core (async) proxy
        mbar_arrive(&tail->full[stage]);
        if (++stage == kStages) {
          stage = 0;
          phase ^= 1;
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, kTmemCols);
  }
}

// tiny kernel: CLS rows  out[b*T + 0, :] = bf16(cls + pos[0])
__global__ void cls_rows_kernel(const __nv_bfloat16* __restrict__ cls, const __nv_bfloat16* __restrict__ pos,
                                __nv_bfloat16* __restrict__ out, int64_t B, int64_t T, int D) {
  const int64_t n = B * D;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t b = i / D;
    const int d = (int)(i - b * D);
    out[b * T * D + d] = __float2bfloat16_rn(__bfloat162float(cls[d]) + __bfloat162float(pos[d]));
  }
}

// ---- host side ---------------------------------------------------------------------------------------
typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static PFN_encodeTiled get_encode() {
  static PFN_encodeTiled fn = nullptr;
  if (!fn) {
    void* ptr = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<PFN_encodeTiled>(ptr);
    else
      (void)cudaGetLastError();
  }
  return fn;
}

// 2-D bf16 row-major [rows, cols] (ld elements) -> box [box_rows, 64 cols], SWIZZLE_128B
int make_tmap_2d(CUtensorMap* tm, const void* ptr, int64_t rows, int64_t cols, int64_t ld, int box_rows) {
  PFN_encodeTiled enc = get_encode();
  if (!enc) return MM_ERR_CUDA;
  cuuint64_t gdim[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t gstr[1] = {(cuuint64_t)ld * 2};
  cuuint32_t box[2] = {64, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), gdim, gstr, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    g_last_cuda_error = 100000 + (int)r;
    return MM_ERR_CUDA;
  }
  return MM_OK;
}

// 3-D bf16 [d2, d1, d0(cols)] with strides (ld1, ld2 elements) -> box [1, box_rows, 64]
int make_tmap_3d(CUtensorMap* tm, const void* ptr, int64_t d0, int64_t d1, int64_t d2, int64_t ld1, int64_t ld2,
                 int box_rows) {
  PFN_encodeTiled enc = get_encode();
  if (!enc) return MM_ERR_CUDA;
  cuuint64_t gdim[3] = {(cuuint64_t)d0, (cuuint64_t)d1, (cuuint64_t)d2};
  cuuint64_t gstr[2] = {(cuuint64_t)ld1 * 2, (cuuint64_t)ld2 * 2};
  cuuint32_t box[3] = {64, (cuuint32_t)box_rows, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(ptr), gdim, gstr, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    g_last_cuda_error = 100000 + (int)r;
    return MM_ERR_CUDA;
  }
  return MM_OK;
}

// 3-D bf16 map with a [1, box_rows, 16]-element box, SWIZZLE_32B (the 16 extra head columns of head_dim 72)
int make_tmap_3d_sw32(CUtensorMap* tm, const void* ptr, int64_t d0, int64_t d1, int64_t d2, int64_t ld1, int64_t ld2,
                      int box_rows) {
  PFN_encodeTiled enc = get_encode();
  if (!enc) return MM_ERR_CUDA;
  cuuint64_t gdim[3] = {(cuuint64_t)d0, (cuuint64_t)d1, (cuuint64_t)d2};
  cuuint64_t gstr[2] = {(cuuint64_t)ld1 * 2, (cuuint64_t)ld2 * 2};
  cuuint32_t box[3] = {16, (cuuint32_t)box_rows, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(ptr), gdim, gstr, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_32B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    g_last_cuda_error = 100000 + (int)r;
    return MM_ERR_CUDA;
  }
  return MM_OK;
}

template <bool kGatherA, int ACT>
static int launch_gemm_inst(const CUtensorMap& tmA, const CUtensorMap& tmB, GemmParams& p, cudaStream_t st) {
  using Cfg = GemmCfg<kGatherA>;
  static bool attr_set[64] = {false};
  int dev = 0;
  MM_CUDA_CHECK(cudaGetDevice(&dev));
  if (dev >= 0 && dev < 64 && !attr_set[dev]) {
    MM_CUDA_CHECK(cudaFuncSetAttribute(gemm_bf16_tn_kernel<kGatherA, ACT>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       Cfg::kSmemBytes));
    attr_set[dev] = true;
  }
  p.tiles_m = (p.M + BM - 1) / BM;
  p.tiles_n = (p.N + BN - 1) / BN;
  const int tiles = p.tiles_m * p.tiles_n;
  const int sms = sm_count();
  const int grid = tiles < sms ? tiles : sms;
  gemm_bf16_tn_kernel<kGatherA, ACT><<<grid, Cfg::kThreads, Cfg::kSmemBytes, st>>>(tmA, tmB, p);
  MM_CUDA_CHECK(cudaGetLastError());
  return MM_OK;
}

static int launch_gemm(const CUtensorMap& tmA, const CUtensorMap& tmB, GemmParams& p, cudaStream_t st) {
  switch (p.act) {
    case MM_ACT_NONE: return launch_gemm_inst<false, MM_ACT_NONE>(tmA, tmB, p, st);
    case MM_ACT_QUICK_GELU: return launch_gemm_inst<false, MM_ACT_QUICK_GELU>(tmA, tmB, p, st);
    case MM_ACT_GELU_ERF: return launch_gemm_inst<false, MM_ACT_GELU_ERF>(tmA, tmB, p, st);
    case MM_ACT_GELU_TANH: return launch_gemm_inst<false, MM_ACT_GELU_TANH>(tmA, tmB, p, st);
  }
  return MM_ERR_BAD_ARG;
}

}  // namespace mm

using namespace mm;

static bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

extern "C" int mm_gemm_bias_act(const void* A, int64_t lda, const void* W, int64_t ldw, const void* bias, void* C,
                                int64_t ldc, int64_t M, int64_t N, int64_t K, int act, const void* residual,
                                int64_t ldr, int64_t res_mod, int64_t res_row_off, const int32_t* out_row_map,
                                int64_t n_out_rows, void* stream) {
  if (M < 0 || N <= 0 || K <= 0 || M > INT32_MAX / 2 || N > INT32_MAX / 2 || K > INT32_MAX / 2) return MM_ERR_BAD_ARG;
  if (M == 0) return MM_OK;
  if (!A || !W || !C) return MM_ERR_BAD_ARG;
  if (K % 8 || N % 8 || lda % 8 || ldw % 8 || ldc % 8 || (residual && ldr % 8)) return MM_ERR_BAD_ARG;
  if (lda < K || ldw < K || ldc < N) return MM_ERR_BAD_ARG;
  if (!aligned16(A) || !aligned16(W) || !aligned16(C) || !aligned16(bias) || !aligned16(residual))
    return MM_ERR_BAD_ARG;
  if (act < MM_ACT_NONE || act > MM_ACT_GELU_TANH) return MM_ERR_BAD_ARG;
  if (out_row_map && n_out_rows <= 0) return MM_ERR_BAD_ARG;
  CUtensorMap tmA, tmB;
  int rc = make_tmap_2d(&tmA, A, M, K, lda, BM);
  if (rc != MM_OK) return rc;
  rc = make_tmap_2d(&tmB, W, N, K, ldw, BN);
  if (rc != MM_OK) return rc;
  GemmParams p{};
  p.C = reinterpret_cast<__nv_bfloat16*>(C);
  p.bias = reinterpret_cast<const __nv_bfloat16*>(bias);
  p.residual = reinterpret_cast<const __nv_bfloat16*>(residual);
  p.out_row_map = out_row_map;
  p.n_out_rows = out_row_map ? n_out_rows : M;
  p.ldc = ldc;
  p.ldr = ldr;
  p.M = (int)M, p.N = (int)N, p.K = (int)K;
  p.res_mod = (int)res_mod, p.res_row_off = (int)res_row_off;
  p.act = act;
  return launch_gemm(tmA, tmB, p, reinterpret_cast<cudaStream_t>(stream));
}

extern "C" int mm_patch_embed(const void* pixels, int pixel_dtype, int64_t B, int64_t img, int64_t patch,
                              const void* weight, int64_t ldw, const void* conv_bias, const void* pos_emb,
                              const void* cls, void* out, int64_t D, void* stream) {
  if (B < 0 || img <= 0 || patch <= 0 || img < patch || D <= 0 || D % 8) return MM_ERR_BAD_ARG;  // img % patch != 0 is fine: the conv (stride = kernel) drops the remainder, SigLIP 384/14 -> 27x27
  if (B == 0) return MM_OK;
  if (!pixels || !weight || !pos_emb || !out) return MM_ERR_BAD_ARG;
  if (pixel_dtype != MM_F32 && pixel_dtype != MM_BF16) return MM_ERR_BAD_ARG;
  const int64_t G = img / patch, P = G * G, Kreal = 3 * patch * patch;
  const int64_t Kpad = (Kreal + BK - 1) / BK * BK;
  if (Kpad > kKTableMax || ldw < Kpad || ldw % 8) return MM_ERR_BAD_ARG;  // weight must be zero-padded to Kpad
  if (B * P > INT32_MAX / 2 || B * 3 * img * img > (int64_t)INT32_MAX * 4) return MM_ERR_BAD_ARG;
  if (!aligned16(weight) || !aligned16(conv_bias) || !aligned16(pos_emb) || !aligned16(out)) return MM_ERR_BAD_ARG;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const int cls_rows = cls ? 1 : 0;
  const int64_t T = P + cls_rows;
  CUtensorMap tmB;
  int rc = make_tmap_2d(&tmB, weight, D, Kpad, ldw, BN);
  if (rc != MM_OK) return rc;
  GemmParams p{};
  p.C = reinterpret_cast<__nv_bfloat16*>(out);
  p.bias = reinterpret_cast<const __nv_bfloat16*>(conv_bias);
  p.residual = reinterpret_cast<const __nv_bfloat16*>(pos_emb);
  p.ldc = D;
  p.ldr = D;
  p.M = (int)(B * P), p.N = (int)D, p.K = (int)Kpad;
  p.res_mod = (int)P, p.res_row_off = cls_rows;
  p.map_in = (int)P, p.map_out = (int)T, p.map_off = cls_rows;
  p.act = MM_ACT_NONE;
  p.pixels = pixels;
  p.pixel_is_f32 = (pixel_dtype == MM_F32);
  p.img = (int)img, p.patch = (int)patch, p.grid = (int)G;
  p.k_real = (int)Kreal;
  rc = launch_gemm_inst<true, MM_ACT_NONE>(tmB, tmB, p, st);
  if (rc != MM_OK) return rc;
  if (cls) {
    const int64_t n = B * D;
    int grid = (int)((n + 255) / 256);
    if (grid > sm_count() * 8) grid = sm_count() * 8;
    cls_rows_kernel<<<grid, 256, 0, st>>>(reinterpret_cast<const __nv_bfloat16*>(cls),
                                          reinterpret_cast<const __nv_bfloat16*>(pos_emb),
                                          reinterpret_cast<__nv_bfloat16*>(out), B, T, (int)D);
    MM_CUDA_CHECK(cudaGetLastError());
  }
  return MM_OK;
}
