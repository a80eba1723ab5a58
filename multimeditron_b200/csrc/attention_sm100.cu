// attention_sm100.cu — non-causal multi-head attention forward for ViT towers (head_dim 64, T <= 640).
//
// One persistent CTA per SM; work unit = (image, head).  K and V of the unit (ceil(T/128) tiles each)
// are TMA-loaded once and stay resident in shared memory; the CTA then walks the 128-row query tiles.
//   S = Q·Kᵀ      tcgen05.mma 128x128x16 (both operands K-major, SWIZZLE_128B), fp32 in TMEM, 2 buffers
//   P = softmax   4 warps, one query row per thread straight out of TMEM (tcgen05.ld 32x32b), online
//                 max/sum in the exp2 domain, P written as bf16 into a K-major SWIZZLE_128B smem tile
//   O += P·V      tcgen05.mma 128x64x16, V consumed MN-major directly from its [kv, head_dim] TMA tile
//                 (no transpose pass), fp32 in TMEM, rescaled in place when the running max moves
// S(kt+1) is issued before softmax(kt) finishes, so the tensor pipe overlaps the MUFU-bound softmax.
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
constexpr int kPBytes = kBQ * kBKV * 2;    // 32 KiB
constexpr int kAttnThreads = 256;

struct AttnBars {
  uint64_t kv_full[kMaxKV];
  uint64_t kv_free;   // all MMAs of the unit retired -> K/V smem reusable
  uint64_t q_full, q_empty;
  uint64_t s_full[2], s_empty[2];
  uint64_t p_full, pv_done;
  uint64_t o_empty;
  uint32_t tmem_base;
};
constexpr int kAttnSmemBytes = 1024 + kTileBytes * (1 + 2 * kMaxKV) + kPBytes + 256;
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

__global__ void __launch_bounds__(kAttnThreads, 1)
    attention_fwd_kernel(const __grid_constant__ CUtensorMap tmQKV, const AttnParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sQ = smem;
  uint8_t* sK = sQ + kTileBytes;
  uint8_t* sV = sK + kMaxKV * kTileBytes;
  uint8_t* sP = sV + kMaxKV * kTileBytes;
  AttnBars* bars = reinterpret_cast<AttnBars*>(sP + kPBytes);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nkv = (p.T + kBKV - 1) / kBKV;
  const int nq = (p.T + kBQ - 1) / kBQ;
  const int units = p.B * p.heads;

  if (warp == 0 && lane == 0) tma_prefetch_desc(&tmQKV);
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < kMaxKV; ++i) mbar_init(&bars->kv_full[i], 1);
    mbar_init(&bars->kv_free, 1);
    mbar_init(&bars->q_full, 1);
    mbar_init(&bars->q_empty, 1);
    for (int i = 0; i < 2; ++i) {
      mbar_init(&bars->s_full[i], 1);
      mbar_init(&bars->s_empty[i], 4);
    }
    mbar_init(&bars->p_full, 4);
    mbar_init(&bars->pv_done, 1);
    mbar_init(&bars->o_empty, 4);
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
  const uint32_t tmem_S = tmem_base;        // 2 x 128 columns
  const uint32_t tmem_O = tmem_base + 256;  // 64 columns

  if (warp == 0) {
    // =========================== TMA producer ===========================
    if (elect_one_sync()) {
      uint32_t unit_i = 0, q_i = 0;
      for (int u = blockIdx.x; u < units; u += gridDim.x, ++unit_i) {
        const int b = u / p.heads, h = u - b * p.heads;
        mbar_wait(&bars->kv_free, (unit_i & 1) ^ 1);  // previous unit's MMAs done reading K/V
        for (int kt = 0; kt < nkv; ++kt) {
          mbar_arrive_expect_tx(&bars->kv_full[kt], 2 * kTileBytes);
          tma_load_3d(sK + kt * kTileBytes, &tmQKV, &bars->kv_full[kt], p.D + h * kHD, kt * kBKV, b);
          tma_load_3d(sV + kt * kTileBytes, &tmQKV, &bars->kv_full[kt], 2 * p.D + h * kHD, kt * kBKV, b);
          if (kt == 0) {
            // first Q tile goes out right after the first K/V tile so S(0) can start early
            mbar_wait(&bars->q_empty, (q_i & 1) ^ 1);
            mbar_arrive_expect_tx(&bars->q_full, kTileBytes);
            tma_load_3d(sQ, &tmQKV, &bars->q_full, h * kHD, 0, b);
            ++q_i;
          }
        }
        for (int qt = 1; qt < nq; ++qt, ++q_i) {
          mbar_wait(&bars->q_empty, (q_i & 1) ^ 1);
          mbar_arrive_expect_tx(&bars->q_full, kTileBytes);
          tma_load_3d(sQ, &tmQKV, &bars->q_full, h * kHD, qt * kBQ, b);
        }
      }
    }
  } else if (warp == 1) {
    // =========================== MMA issuer ===========================
    if (elect_one_sync()) {
      constexpr uint32_t idesc_s = umma_idesc_bf16(kBQ, kBKV, 0, 0);
      constexpr uint32_t idesc_pv = umma_idesc_bf16(kBQ, kHD, 0, 1);  // B (= V) is MN-major
      uint32_t unit_i = 0, q_i = 0, it_s = 0, it_pv = 0;
      const uint32_t q_addr = smem_u32(sQ), p_addr = smem_u32(sP);
      auto issue_s = [&](int kt) {
        const uint32_t buf = it_s & 1;
        mbar_wait(&bars->s_empty[buf], ((it_s >> 1) & 1) ^ 1);
        tc_fence_after();
        const uint32_t k_addr = smem_u32(sK + kt * kTileBytes);
#pragma unroll
        for (int k = 0; k < kHD / 16; ++k) {
          tc_mma_f16_ss(tmem_S + buf * kBKV, umma_desc_sw128(q_addr + k * 32, 16, 1024),
                        umma_desc_sw128(k_addr + k * 32, 16, 1024), idesc_s, k != 0);
        }
        tc_commit(&bars->s_full[buf]);
        ++it_s;
      };
      auto issue_pv = [&](int kt) {
        mbar_wait(&bars->p_full, it_pv & 1);
        if (kt == 0) mbar_wait(&bars->o_empty, (q_i & 1) ^ 1);  // previous q tile's epilogue has drained O
        tc_fence_after();
        const uint32_t v_addr = smem_u32(sV + kt * kTileBytes);
#pragma unroll
        for (int k = 0; k < kBKV / 16; ++k) {
          // A = P: K-major, two 64-wide k-blocks of 16 KiB; B = V: MN-major, 16 kv rows = 2 KiB per step
          const uint64_t da = umma_desc_sw128(p_addr + (k >> 2) * (kBQ * 128) + (k & 3) * 32, 16, 1024);
          const uint64_t db = umma_desc_sw128(v_addr + k * 2048, 1024, 1024);
          tc_mma_f16_ss(tmem_O, da, db, idesc_pv, (kt | k) != 0);
        }
        tc_commit(&bars->pv_done);
        ++it_pv;
      };
      for (int u = blockIdx.x; u < units; u += gridDim.x, ++unit_i) {
        for (int qt = 0; qt < nq; ++qt, ++q_i) {
          mbar_wait(&bars->q_full, q_i & 1);
          tc_fence_after();
          for (int kt = 0; kt < nkv; ++kt) {
            if (qt == 0) mbar_wait(&bars->kv_full[kt], unit_i & 1);
            issue_s(kt);
            if (kt == nkv - 1) tc_commit(&bars->q_empty);
            if (kt >= 1) issue_pv(kt - 1);
          }
          issue_pv(nkv - 1);
        }
        tc_commit(&bars->kv_free);
      }
    }
  } else if (warp >= 4) {
    // =========================== softmax / correction / epilogue ===========================
    const int q = warp & 3;
    const int row = q * 32 + lane;  // query row within the tile == TMEM lane
    const uint32_t lane_sel = (uint32_t)(q * 32) << 16;
    const uint32_t p_row = smem_u32(sP) + row * 128;
    uint32_t it = 0, q_i = 0;
    for (int u = blockIdx.x; u < units; u += gridDim.x) {
      const int b = u / p.heads, h = u - b * p.heads;
      for (int qt = 0; qt < nq; ++qt, ++q_i) {
        float m_run = -INFINITY, l_run = 0.f;
        for (int kt = 0; kt < nkv; ++kt, ++it) {
          const uint32_t buf = it & 1;
          const uint32_t s_addr = tmem_S + buf * kBKV + lane_sel;
          const int valid = p.T - kt * kBKV;  // columns >= valid are padding
          mbar_wait(&bars->s_full[buf], (it >> 1) & 1);
          tc_fence_after();
          // pass 1: row max
          float mx = m_run;
#pragma unroll 1
          for (int c = 0; c < kBKV / 32; ++c) {
            uint32_t s[32];
            tmem_ld_32x32b_x32(s_addr + c * 32, s);
            tc_wait_ld();
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              float v = __uint_as_float(s[j]);
              if (c * 32 + j < valid) mx = fmaxf(mx, v);
            }
          }
          const float m_new = mx;
          const float m_scaled = m_new * p.scale_log2;
          // pass 2: p = exp2(s*scale - m*scale), bf16 P tile, row sum
          uint32_t pk[kBKV / 2];
          float sum = 0.f;
#pragma unroll
          for (int c = 0; c < kBKV / 32; ++c) {
            uint32_t s[32];
            tmem_ld_32x32b_x32(s_addr + c * 32, s);
            tc_wait_ld();
#pragma unroll
            for (int j = 0; j < 32; j += 2) {
              float p0 = (c * 32 + j < valid) ? ex2(fmaf(__uint_as_float(s[j]), p.scale_log2, -m_scaled)) : 0.f;
              float p1 = (c * 32 + j + 1 < valid) ? ex2(fmaf(__uint_as_float(s[j + 1]), p.scale_log2, -m_scaled)) : 0.f;
              sum += p0 + p1;
              pk[c * 16 + (j >> 1)] = pack_bf16x2(p0, p1);
            }
          }
          // S buffer fully read -> the issuer may overwrite it with S(kt+2)
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&bars->s_empty[buf]);
          // P smem / O accumulator are free once PV of the previous iteration retired
          if (it > 0) {
            mbar_wait(&bars->pv_done, (it - 1) & 1);
            tc_fence_after();
          }
          if (kt > 0) {
            const float alpha = ex2((m_run - m_new) * p.scale_log2);
            l_run = l_run * alpha + sum;
            if (__any_sync(0xffffffffu, m_new != m_run)) {
#pragma unroll 1
              for (int c = 0; c < kHD / 32; ++c) {
                uint32_t o[32];
                tmem_ld_32x32b_x32(tmem_O + lane_sel + c * 32, o);
                tc_wait_ld();
#pragma unroll
                for (int j = 0; j < 32; ++j) o[j] = __float_as_uint(__uint_as_float(o[j]) * alpha);
                tmem_st_32x32b_x32(tmem_O + lane_sel + c * 32, o);
              }
              tc_wait_st();
            }
          } else {
            l_run = sum;
          }
          m_run = m_new;
#pragma unroll
          for (int j = 0; j < kBKV / 8; ++j) {  // 16 chunks of 8 bf16
            const uint32_t a = p_row + (j >> 3) * (kBQ * 128) + (((j & 7) ^ (row & 7)) << 4);
            asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(a), "r"(pk[j * 4]), "r"(pk[j * 4 + 1]),
                         "r"(pk[j * 4 + 2]), "r"(pk[j * 4 + 3])
                         : "memory");
          }
          fence_proxy_async_smem();
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&bars->p_full);
        }
        // ---- epilogue of the q tile: O / l -> bf16 -> global ----
        mbar_wait(&bars->pv_done, (it - 1) & 1);
        tc_fence_after();
        const float inv_l = 1.0f / l_run;
        const int t = qt * kBQ + row;
        __nv_bfloat16* dst = p.out + ((int64_t)b * p.T + t) * p.D + h * kHD;
#pragma unroll 1
        for (int c = 0; c < kHD / 32; ++c) {
          uint32_t o[32];
          tmem_ld_32x32b_x32(tmem_O + lane_sel + c * 32, o);
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
        if (lane == 0) mbar_arrive(&bars->o_empty);
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
