// attention_sm100.cu — non-causal multi-head attention forward for ViT towers (head_dim 64, T <= 640).
//
// One persistent CTA per SM; work unit = (image, head).  K and V of the unit (ceil(T/128) tiles each) are
// TMA-loaded once and stay resident in shared memory.  The CTA walks the 128-row query tiles TWO AT A TIME
// (two "streams", one softmax warpgroup each), so that one stream's MUFU-bound softmax overlaps the other
// stream's tensor-core work:
//   S = Q·Kᵀ      tcgen05.mma 128x128x16, SS form (both operands K-major SWIZZLE_128B smem), fp32 in TMEM
//   P = softmax   one query row per thread straight out of TMEM (tcgen05.ld 32x32b); online max/sum in the
//                 exp2 domain; the running max only moves when it grows by > 8 (log2 units), so the O
//                 accumulator is rescaled rarely; P goes back to TMEM as packed bf16 (tcgen05.st)
//   O += P·V      tcgen05.mma 128x64x16, TS form: A = P read from TMEM, B = V consumed MN-major directly from
//                 its [kv, head_dim] TMA tile (no transpose pass); fp32 in TMEM
// TMEM map (512 columns): S0 0..127 | S1 128..255 | P0 256..319 | P1 320..383 | O0 384..447 | O1 448..511.
//
// Replaces HF:models/clip/modeling_clip.py:318-331 (SDPA; scale = head_dim**-0.5, no mask, dropout 0).
#include "common.cuh"

namespace mm {

int make_tmap_3d(CUtensorMap* tm, const void* ptr, int64_t d0, int64_t d1, int64_t d2, int64_t ld1, int64_t ld2,
                 int box_rows);

constexpr int kHD = 64;
constexpr int kBQ = 128;
constexpr int kBKV = 128;
constexpr int kMaxKV = 5;                  // T <= 640
constexpr int kTileBytes = 128 * kHD * 2;  // 16 KiB : [128 rows x 64 bf16]
constexpr int kAttnThreads = 384;          // w0 TMA, w1 MMA, w2 TMEM alloc, w3 idle, w4-7 stream 0, w8-11 stream 1
constexpr float kRescaleThreshold = 8.0f;  // log2 units: P <= 2^8, harmless in bf16/fp32

struct AttnBars {
  uint64_t kv_full[kMaxKV];
  uint64_t kv_free;  // all MMAs of the unit retired -> K/V smem reusable
  uint64_t q_full[2], q_empty[2];
  uint64_t s_full[2], s_empty[2];
  uint64_t p_full[2], pv_done[2];
  uint64_t o_empty[2];
  uint32_t tmem_base;
};
constexpr int kAttnSmemBytes = 1024 + kTileBytes * (2 + 2 * kMaxKV) + 256;
static_assert(sizeof(AttnBars) <= 256, "bars");
static_assert(kAttnSmemBytes <= 232448, "smem budget");

struct AttnParams {
  __nv_bfloat16* out;  // [B*T, D]
  int B, T, heads, D;
  float scale_log2;  // scale * log2(e)
};

__device__ __forceinline__ float ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// D[tmem] (+)= A[tmem] * B[smem desc]
__device__ __forceinline__ void tc_mma_f16_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc,
                                              uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t"
      "}\n" ::"r"(d_tmem),
      "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}

__global__ void __launch_bounds__(kAttnThreads, 1)
    attention_fwd_kernel(const __grid_constant__ CUtensorMap tmQKV, const AttnParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sQ = smem;                   // 2 tiles (one per stream)
  uint8_t* sK = sQ + 2 * kTileBytes;
  uint8_t* sV = sK + kMaxKV * kTileBytes;
  AttnBars* bars = reinterpret_cast<AttnBars*>(sV + kMaxKV * kTileBytes);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nkv = (p.T + kBKV - 1) / kBKV;
  const int nq = (p.T + kBQ - 1) / kBQ;
  const int npair = (nq + 1) >> 1;
  const int units = p.B * p.heads;

  if (warp == 0 && lane == 0) tma_prefetch_desc(&tmQKV);
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < kMaxKV; ++i) mbar_init(&bars->kv_full[i], 1);
    mbar_init(&bars->kv_free, 1);
    for (int s = 0; s < 2; ++s) {
      mbar_init(&bars->q_full[s], 1);
      mbar_init(&bars->q_empty[s], 1);
      mbar_init(&bars->s_full[s], 1);
      mbar_init(&bars->s_empty[s], 4);
      mbar_init(&bars->p_full[s], 4);
      mbar_init(&bars->pv_done[s], 1);
      mbar_init(&bars->o_empty[s], 4);
    }
    fence_barrier_init();
  }
  if (warp == 2) {
    tmem_alloc(&bars->tmem_base, 512);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = bars->tmem_base;

  if (warp == 0) {
    // =========================== TMA producer ===========================
    if (elect_one_sync()) {
      uint32_t unit_i = 0;
      uint32_t qi[2] = {0, 0};
      auto load_q = [&](int s, int qt, int h, int b) {
        mbar_wait(&bars->q_empty[s], (qi[s] & 1) ^ 1);
        mbar_arrive_expect_tx(&bars->q_full[s], kTileBytes);
        tma_load_3d(sQ + s * kTileBytes, &tmQKV, &bars->q_full[s], h * kHD, qt * kBQ, b);
        ++qi[s];
      };
      for (int u = blockIdx.x; u < units; u += gridDim.x, ++unit_i) {
        const int b = u / p.heads, h = u - b * p.heads;
        mbar_wait(&bars->kv_free, (unit_i & 1) ^ 1);  // previous unit's MMAs done reading K/V
        for (int kt = 0; kt < nkv; ++kt) {
          mbar_arrive_expect_tx(&bars->kv_full[kt], 2 * kTileBytes);
          tma_load_3d(sK + kt * kTileBytes, &tmQKV, &bars->kv_full[kt], p.D + h * kHD, kt * kBKV, b);
          tma_load_3d(sV + kt * kTileBytes, &tmQKV, &bars->kv_full[kt], 2 * p.D + h * kHD, kt * kBKV, b);
          if (kt == 0) {  // first pair of Q tiles right behind the first K/V tile so S(0) can start early
            load_q(0, 0, h, b);
            if (nq > 1) load_q(1, 1, h, b);
          }
        }
        for (int pi = 1; pi < npair; ++pi) {
          load_q(0, 2 * pi, h, b);
          if (2 * pi + 1 < nq) load_q(1, 2 * pi + 1, h, b);
        }
      }
    }
  } else if (warp == 1) {
    // =========================== MMA issuer ===========================
    if (elect_one_sync()) {
      constexpr uint32_t idesc_s = umma_idesc_bf16(kBQ, kBKV, 0, 0);
      constexpr uint32_t idesc_pv = umma_idesc_bf16(kBQ, kHD, 0, 1);  // B (= V) is MN-major
      uint32_t unit_i = 0;
      uint32_t qi[2] = {0, 0}, it_s[2] = {0, 0}, it_pv[2] = {0, 0};
      auto issue_s = [&](int s, int kt) {
        mbar_wait(&bars->s_empty[s], (it_s[s] & 1) ^ 1);
        tc_fence_after();
        const uint32_t q_addr = smem_u32(sQ + s * kTileBytes);
        const uint32_t k_addr = smem_u32(sK + kt * kTileBytes);
#pragma unroll
        for (int k = 0; k < kHD / 16; ++k) {
          tc_mma_f16_ss(tmem_base + s * 128, umma_desc_sw128(q_addr + k * 32, 16, 1024),
                        umma_desc_sw128(k_addr + k * 32, 16, 1024), idesc_s, k != 0);
        }
        tc_commit(&bars->s_full[s]);
        ++it_s[s];
      };
      auto issue_pv = [&](int s, int kt) {
        mbar_wait(&bars->p_full[s], it_pv[s] & 1);
        if (kt == 0) mbar_wait(&bars->o_empty[s], (qi[s] & 1) ^ 1);  // previous tile's epilogue drained O
        tc_fence_after();
        const uint32_t v_addr = smem_u32(sV + kt * kTileBytes);
#pragma unroll
        for (int k = 0; k < kBKV / 16; ++k) {
          // A = P from TMEM (16 bf16 = 8 columns per k-step); B = V MN-major, 16 kv rows = 2 KiB per k-step
          tc_mma_f16_ts(tmem_base + 384 + s * 64, tmem_base + 256 + s * 64 + k * 8,
                        umma_desc_sw128(v_addr + k * 2048, 1024, 1024), idesc_pv, (kt | k) != 0);
        }
        tc_commit(&bars->pv_done[s]);
        ++it_pv[s];
      };
      for (int u = blockIdx.x; u < units; u += gridDim.x, ++unit_i) {
        for (int pi = 0; pi < npair; ++pi) {
          const int ns = (2 * pi + 1 < nq) ? 2 : 1;  // active streams in this pair
          for (int s = 0; s < ns; ++s) mbar_wait(&bars->q_full[s], qi[s] & 1);
          tc_fence_after();
          for (int kt = 0; kt < nkv; ++kt) {
            if (pi == 0) mbar_wait(&bars->kv_full[kt], unit_i & 1);
            for (int s = 0; s < ns; ++s) {
              issue_s(s, kt);
              if (kt == nkv - 1) tc_commit(&bars->q_empty[s]);
            }
            if (kt >= 1)
              for (int s = 0; s < ns; ++s) issue_pv(s, kt - 1);
          }
          for (int s = 0; s < ns; ++s) {
            issue_pv(s, nkv - 1);
            ++qi[s];
          }
        }
        tc_commit(&bars->kv_free);
      }
    }
  } else if (warp >= 4) {
    // =========================== softmax / correction / epilogue (stream = warpgroup) ===========================
    const int s = (warp - 4) >> 2;
    const int q = warp & 3;
    const int row = q * 32 + lane;  // query row within the tile == TMEM lane
    const uint32_t lane_sel = (uint32_t)(q * 32) << 16;
    const uint32_t t_S = tmem_base + s * 128 + lane_sel;
    const uint32_t t_P = tmem_base + 256 + s * 64 + lane_sel;
    const uint32_t t_O = tmem_base + 384 + s * 64 + lane_sel;
    uint32_t it = 0;
    for (int u = blockIdx.x; u < units; u += gridDim.x) {
      const int b = u / p.heads, h = u - b * p.heads;
      for (int qt = s; qt < nq; qt += 2) {
        float m_run = -INFINITY, l_run = 0.f;  // running max (raw score units) and row sum
        for (int kt = 0; kt < nkv; ++kt, ++it) {
          const int valid = p.T - kt * kBKV;  // columns >= valid are padding (only in the last kv tile)
          mbar_wait(&bars->s_full[s], it & 1);
          tc_fence_after();
          // ---- pass 1: row max
          float mx = -INFINITY;
#pragma unroll 1
          for (int c = 0; c < kBKV / 32; ++c) {
            if (c * 32 >= valid) break;
            uint32_t sv[32];
            tmem_ld_32x32b_x32(t_S + c * 32, sv);
            tc_wait_ld();
            if (c * 32 + 32 <= valid) {
#pragma unroll
              for (int j = 0; j < 32; ++j) mx = fmaxf(mx, __uint_as_float(sv[j]));
            } else {
#pragma unroll
              for (int j = 0; j < 32; ++j)
                if (c * 32 + j < valid) mx = fmaxf(mx, __uint_as_float(sv[j]));
            }
          }
          // the reference max only moves when it grows by more than the threshold (keeps O rescales rare)
          float m_new = m_run;
          if ((mx - m_run) * p.scale_log2 > kRescaleThreshold) m_new = mx;  // also true for m_run = -inf
          const float m_scaled = m_new * p.scale_log2;
          // ---- pass 2: p = exp2(s*scale - m*scale) -> packed bf16, row sum
          uint32_t pk[kBKV / 2];
          float sum = 0.f;
#pragma unroll
          for (int c = 0; c < kBKV / 32; ++c) {
            if (c * 32 < valid) {
              uint32_t sv[32];
              tmem_ld_32x32b_x32(t_S + c * 32, sv);
              tc_wait_ld();
              if (c * 32 + 32 <= valid) {
#pragma unroll
                for (int j = 0; j < 32; j += 2) {
                  const float p0 = ex2(fmaf(__uint_as_float(sv[j]), p.scale_log2, -m_scaled));
                  const float p1 = ex2(fmaf(__uint_as_float(sv[j + 1]), p.scale_log2, -m_scaled));
                  sum += p0 + p1;
                  pk[c * 16 + (j >> 1)] = pack_bf16x2(p0, p1);
                }
              } else {
#pragma unroll
                for (int j = 0; j < 32; j += 2) {
                  const float p0 =
                      (c * 32 + j < valid) ? ex2(fmaf(__uint_as_float(sv[j]), p.scale_log2, -m_scaled)) : 0.f;
                  const float p1 =
                      (c * 32 + j + 1 < valid) ? ex2(fmaf(__uint_as_float(sv[j + 1]), p.scale_log2, -m_scaled)) : 0.f;
                  sum += p0 + p1;
                  pk[c * 16 + (j >> 1)] = pack_bf16x2(p0, p1);
                }
              }
            } else {
#pragma unroll
              for (int j = 0; j < 16; ++j) pk[c * 16 + j] = 0u;
            }
          }
          // S fully read -> the issuer may overwrite it with S(kt+1)
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&bars->s_empty[s]);
          // P (TMEM) and the O accumulator are free once PV of the previous iteration retired
          if (it > 0) {
            mbar_wait(&bars->pv_done[s], (it - 1) & 1);
            tc_fence_after();
          }
          if (kt > 0) {
            const float alpha = ex2((m_run - m_new) * p.scale_log2);  // 1 when the max did not move
            l_run = l_run * alpha + sum;
            if (__any_sync(0xffffffffu, m_new != m_run)) {
#pragma unroll 1
              for (int c = 0; c < kHD / 32; ++c) {
                uint32_t o[32];
                tmem_ld_32x32b_x32(t_O + c * 32, o);
                tc_wait_ld();
#pragma unroll
                for (int j = 0; j < 32; ++j) o[j] = __float_as_uint(__uint_as_float(o[j]) * alpha);
                tmem_st_32x32b_x32(t_O + c * 32, o);
              }
            }
          } else {
            l_run = sum;
          }
          m_run = m_new;
          tmem_st_32x32b_x32(t_P, pk);
          tmem_st_32x32b_x32(t_P + 32, pk + 32);
          tc_wait_st();
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&bars->p_full[s]);
        }
        // ---- epilogue of the q tile: O / l -> bf16 -> global ----
        mbar_wait(&bars->pv_done[s], (it - 1) & 1);
        tc_fence_after();
        const float inv_l = 1.0f / l_run;
        const int t = qt * kBQ + row;
        __nv_bfloat16* dst = p.out + ((int64_t)b * p.T + t) * p.D + h * kHD;
#pragma unroll 1
        for (int c = 0; c < kHD / 32; ++c) {
          uint32_t o[32];
          tmem_ld_32x32b_x32(t_O + c * 32, o);
          tc_wait_ld();
          if (t < p.T) {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              int4 v;
              v.x = (int)pack_bf16x2(__uint_as_float(o[j * 8 + 0]) * inv_l, __uint_as_float(o[j * 8 + 1]) * inv_l);
              v.y = (int)pack_bf16x2(__uint_as_float(o[j * 8 + 2]) * inv_l, __uint_as_float(o[j * 8 + 3]) * inv_l);
              v.z = (int)pack_bf16x2(__uint_as_float(o[j * 8 + 4]) * inv_l, __uint_as_float(o[j * 8 + 5]) * inv_l);
              v.w = (int)pack_bf16x2(__uint_as_float(o[j * 8 + 6]) * inv_l, __uint_as_float(o[j * 8 + 7]) * inv_l);
              *reinterpret_cast<int4*>(dst + c * 32 + j * 8) = v;
            }
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&bars->o_empty[s]);
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

}  // namespace mm

using namespace mm;

extern "C" int mm_attention_fwd(const void* qkv, void* out, int64_t B, int64_t T, int64_t heads, int64_t head_dim,
                                float scale, void* stream) {
  if (B < 0 || T <= 0 || heads <= 0) return MM_ERR_BAD_ARG;
  if (head_dim != kHD || T > (int64_t)kMaxKV * kBKV) return MM_ERR_BAD_ARG;
  if (B == 0) return MM_OK;
  if (!qkv || !out || (reinterpret_cast<uintptr_t>(qkv) & 15) || (reinterpret_cast<uintptr_t>(out) & 15))
    return MM_ERR_BAD_ARG;
  if (B * heads > INT32_MAX / 2) return MM_ERR_BAD_ARG;
  if (!(scale > 0.f)) return MM_ERR_BAD_ARG;  // the online max runs in raw-score units (needs scale > 0)
  const int64_t D = heads * head_dim;
  CUtensorMap tm;
  int rc = make_tmap_3d(&tm, qkv, 3 * D, T, B, 3 * D, T * 3 * D, 128);
  if (rc != MM_OK) return rc;
  static bool attr_set[64] = {false};
  int dev = 0;
  MM_CUDA_CHECK(cudaGetDevice(&dev));
  if (dev >= 0 && dev < 64 && !attr_set[dev]) {
    MM_CUDA_CHECK(
        cudaFuncSetAttribute(attention_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kAttnSmemBytes));
    attr_set[dev] = true;
  }
  AttnParams p;
  p.out = reinterpret_cast<__nv_bfloat16*>(out);
  p.B = (int)B, p.T = (int)T, p.heads = (int)heads, p.D = (int)D;
  p.scale_log2 = scale * 1.4426950408889634f;
  const int units = (int)(B * heads);
  const int sms = sm_count();
  const int grid = units < sms ? units : sms;
  attention_fwd_kernel<<<grid, kAttnThreads, kAttnSmemBytes, reinterpret_cast<cudaStream_t>(stream)>>>(tm, p);
  MM_CUDA_CHECK(cudaGetLastError());
  return MM_OK;
}
