// gemm_sm100.cu — persistent warp-specialised tcgen05 GEMM:  C = act(A · Wᵀ + bias) [+ residual]
//
//   * operands bf16, K-major, staged by TMA (SWIZZLE_128B, 64-element K slabs) into a shared-memory ring
//   * CTA-pair variant (kPair, the default for M > 128): the two CTAs of a 2-CTA cluster compute one 256 x 256 tile
//     with tcgen05.mma cta_group::2 (UMMA 256x256x16) issued by the leader CTA; each CTA stages its own 128 rows of
//     A and HALF of the B tile, so shared-memory operand bytes per FLOP halve and the ring is 6 deep.
//     Protocol: every TMA load of both CTAs signals the LEADER's full barrier; the leader's producer posts the
//     bytes of both CTAs; tcgen05.commit multicasts to both CTAs' empty / tmem_full barriers; the peer's epilogue
//     warps release the accumulator stage on the leader's tmem_empty barrier (mapa + remote arrive).
//   * 1-CTA variant (small M, and the gather kernel): tcgen05.mma cta_group::1, UMMA 128x256x16, 4-deep ring
//   * fp32 accumulators in TMEM, double-buffered (2 x 256 columns) so tile i+1's main loop overlaps tile i's epilogue
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
#include "act.cuh"

namespace mm {

constexpr int BM = 128;
constexpr int BN = 256;
constexpr int BK = 64;  // 64 bf16 = 128 B = one SWIZZLE_128B row
constexpr int kUmmaK = 16;
constexpr int kAStageBytes = BM * BK * 2;  // 16 KiB
constexpr int kBStageBytes = BN * BK * 2;  // 32 KiB (CTA pair: each CTA holds half of it)
constexpr int kEpiWarpBytes = 32 * 128;    // 32 rows x 64 bf16
constexpr int kTmemCols = 512;
constexpr int kKTableMax = 3072 + 64;  // patch-embed k -> pixel offset table (ints)

constexpr int kStatSlots = 16;  // row-statistics slots per row: one 128-byte line, one slot per 128 output columns
constexpr int kMaxStages = 6;
struct GemmSmemTail {
  uint64_t full[kMaxStages];
  uint64_t empty[kMaxStages];
  uint64_t tmem_full[2];
  uint64_t tmem_empty[2];
  uint32_t tmem_base;
};

// TMA variant: 4 stages, 8 epilogue warps (2 per TMEM lane quadrant, each owning half of the 256 columns).
// Gather variant (patch embed): 3 stages, 4 epilogue warps, 4 A-producer warps, + 12.5 KiB static k-offset table.
// CTA-pair variant: 6 stages of (16 KiB A + 16 KiB half-B).
template <bool kGatherA, bool kPair = false>
struct GemmCfg {
  static_assert(!(kGatherA && kPair), "the gather kernel is 1-CTA");
  static constexpr int kStages = kGatherA ? 3 : (kPair ? 6 : 4);
  static constexpr int kBStage = kPair ? kBStageBytes / 2 : kBStageBytes;
  static constexpr int kEpiWarps = kGatherA ? 4 : 8;
  static constexpr int kChunksPerWarp = (BN / 64) / (kEpiWarps / 4);
  static constexpr int kThreads = 128 + kEpiWarps * 32 + (kGatherA ? 128 : 0);
  static constexpr int kEpiBytes = kEpiWarps * kEpiWarpBytes;
  static constexpr int kSmemBytes = 1024 /*align slack*/ + kStages * (kAStageBytes + kBStage) + kEpiBytes + 256;
};
static_assert(sizeof(GemmSmemTail) <= 256, "tail too large");
static_assert(GemmCfg<false>::kSmemBytes <= 232448, "smem budget");
static_assert(GemmCfg<false, true>::kSmemBytes <= 232448, "smem budget (pair)");
static_assert(GemmCfg<true>::kSmemBytes + kKTableMax * 4 + kKTableMax / 8 <= 232448, "smem budget (gather)");

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
  int pair;  // 1: CTA-pair kernel (256-row tiles; the B tensor map has a 128-row box)
  // grouped launch (mm_gemm_grouped): `groups` problems of identical shape stacked along the rows of A / C / residual /
  // statistics and along the rows of W (and bias).  Row tile r belongs to group r / group_tiles_m and multiplies the
  // W rows [group * group_n, +N).  0 = one problem.
  int group_tiles_m, group_n;
  // LayerNorm folded into the GEMM (mm_gemm_ln_bias_act):  LN(x)·Wᵀ + b = rstd·(x·W''ᵀ) + b'
  // with W'' = gamma ⊙ W re-centred so that every row sums to zero (the mean term cancels inside the MMA).
  //   producer side: per 64-column chunk (sum, sum of squares) of the rows this GEMM writes
  //   consumer side: row statistics of A from those partials -> rstd; `bias` is b'
  float2* stats_out;        // [rows of C][slots] or NULL
  const float2* ln_stats;   // [M][slots] or NULL
  float ln_eps, ln_inv_k;
  // gather-A (patch embed) only
  const void* pixels;
  int pixel_is_f32;
  int img, patch, grid;  // grid = img / patch
  int k_real;            // 3*patch*patch
};

// 8 values per lane x 8 lanes (ch = lane & 7) -> the total of value index `ch` ends up in this lane.
// Halving butterfly: 7 shuffles, fixed summation order (bit-reproducible).
__device__ __forceinline__ float butterfly8(const float (&a)[8], int ch) {
  const bool b4 = ch & 4, b2 = ch & 2, b1 = ch & 1;
  float u[4], v[2];
#pragma unroll
  for (int i = 0; i < 4; ++i)
    u[i] = (b4 ? a[i + 4] : a[i]) + __shfl_xor_sync(0xffffffffu, b4 ? a[i] : a[i + 4], 4);
#pragma unroll
  for (int i = 0; i < 2; ++i)
    v[i] = (b2 ? u[i + 2] : u[i]) + __shfl_xor_sync(0xffffffffu, b2 ? u[i] : u[i + 2], 2);
  return (b1 ? v[1] : v[0]) + __shfl_xor_sync(0xffffffffu, b1 ? v[0] : v[1], 1);
}

// LN: 0 = plain epilogue, 1 = also emits row statistics (stats_out), 2 = applies the folded LayerNorm (ln_stats),
// 3 = plain with a modular residual row (res_mod > 0).
// Compile-time, not run-time: each variant keeps the lean epilogue it needs (the epilogue is issue/i-cache bound).
template <bool kGatherA, int ACT, int LN, bool kPair>
__global__ void __launch_bounds__(GemmCfg<kGatherA, kPair>::kThreads, 1)
    gemm_bf16_tn_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                        const GemmParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  using Cfg = GemmCfg<kGatherA, kPair>;
  constexpr int kStages = Cfg::kStages;
  constexpr int kEpiWarps = Cfg::kEpiWarps;
  constexpr int kBStage = Cfg::kBStage;
  uint8_t* sA = smem;
  uint8_t* sB = smem + kStages * kAStageBytes;
  uint8_t* sEpi = sB + kStages * kBStage;
  GemmSmemTail* tail = reinterpret_cast<GemmSmemTail*>(sEpi + Cfg::kEpiBytes);
  __shared__ int s_koff[kGatherA ? kKTableMax : 1];
  __shared__ uint8_t s_kcontig[kGatherA ? kKTableMax / 8 : 1];  // 1: the 8 pixels of this k-group are consecutive in memory

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int num_kb = (p.K + BK - 1) / BK;
  const int num_tiles = p.tiles_m * p.tiles_n;
  // CTA pair: a tile is 256 rows (p.tiles_m counts those), this CTA owns rows [rank*128, +128) and the B rows
  // [rank*128, +128) of it; tiles are strided over PAIRS.
  const uint32_t rank = kPair ? cluster_ctarank() : 0u;
  constexpr int kTileRows = kPair ? 2 * BM : BM;
  const int m_off = (int)rank * BM;
  const int tile0 = kPair ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;
  const int tile_step = kPair ? (int)(gridDim.x >> 1) : (int)gridDim.x;

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
      mbar_init(&tail->tmem_empty[i], kPair ? 2 * kEpiWarps : kEpiWarps);  // pair: both CTAs' warps, leader's barrier
    }
    fence_barrier_init();
  }
  if (kPair) cluster_sync_all();  // the peer's barriers exist before anything can signal them
  if (warp == 2) {
    if (kPair) {
      tmem_alloc_pair(&tail->tmem_base, kTmemCols);
    } else {
      tmem_alloc(&tail->tmem_base, kTmemCols);
      tmem_relinquish();
    }
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
    for (int g = threadIdx.x; g < num_kb * (BK / 8); g += blockDim.x) {
      const int k0 = g * 8;   // same channel, same patch row, 8 consecutive kx, all real
      const int c0 = k0 / pp, r0 = k0 - c0 * pp;
      s_kcontig[g] = (k0 + 7 < p.k_real && (k0 + 7) / pp == c0 && (r0 % p.patch) + 7 < p.patch) ? 1 : 0;
    }
  }
  tc_fence_before();
  if (kPair) cluster_sync_all(); else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tail->tmem_base;

  if (warp == 0) {
    // =========================== TMA producer ===========================
    if (elect_one_sync()) {
      int stage = 0;
      uint32_t phase = 0;
      for (int t = tile0; t < num_tiles; t += tile_step) {
        const int grp = p.group_tiles_m ? (t / p.tiles_n) / p.group_tiles_m : 0;
        const int m0 = (t / p.tiles_n) * kTileRows + m_off;
        const int n0 = (t % p.tiles_n) * BN + (kPair ? m_off : 0) + grp * p.group_n;   // row of the (stacked) W
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&tail->empty[stage], phase ^ 1);
          if (kPair) {
            if (rank == 0) mbar_arrive_expect_tx(&tail->full[stage], 2 * (kAStageBytes + kBStage));
            tma_load_2d_pair(sA + stage * kAStageBytes, &tmA, &tail->full[stage], kb * BK, m0);
            tma_load_2d_pair(sB + stage * kBStage, &tmB, &tail->full[stage], kb * BK, n0);
          } else if (kGatherA) {
            mbar_arrive_expect_tx(&tail->full[stage], kBStageBytes);
          } else {
            mbar_arrive_expect_tx(&tail->full[stage], kAStageBytes + kBStageBytes);
            tma_load_2d(sA + stage * kAStageBytes, &tmA, &tail->full[stage], kb * BK, m0);
          }
          if (!kPair) tma_load_2d(sB + stage * kBStageBytes, &tmB, &tail->full[stage], kb * BK, n0);
          if (++stage == kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    // =========================== MMA issuer (single thread) ===========================
    if (rank == 0 && elect_one_sync()) {
      constexpr uint32_t idesc = umma_idesc_bf16(kTileRows, BN, 0, 0);
      int stage = 0;
      uint32_t phase = 0;
      int as = 0;
      uint32_t aphase = 0;
      for (int t = tile0; t < num_tiles; t += tile_step) {
        mbar_wait(&tail->tmem_empty[as], aphase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)(as * BN);
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&tail->full[stage], phase);
          tc_fence_after();
          const uint32_t a_addr = smem_u32(sA + stage * kAStageBytes);
          const uint32_t b_addr = smem_u32(sB + stage * kBStage);
#pragma unroll
          for (int k = 0; k < BK / kUmmaK; ++k) {
            const uint64_t da = umma_desc_sw128(a_addr + k * kUmmaK * 2, 16, 1024);
            const uint64_t db = umma_desc_sw128(b_addr + k * kUmmaK * 2, 16, 1024);
            if (kPair) tc_mma_f16_ss_pair(d_tmem, da, db, idesc, (kb | k) != 0 ? 1u : 0u);
            else tc_mma_f16_ss(d_tmem, da, db, idesc, (kb | k) != 0 ? 1u : 0u);
          }
          // frees the smem slot (pair: in both CTAs) when these MMAs retire
          if (kPair) tc_commit_pair(&tail->empty[stage]); else tc_commit(&tail->empty[stage]);
          if (++stage == kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
        // accumulator ready for the epilogue (pair: both CTAs' halves)
        if (kPair) tc_commit_pair(&tail->tmem_full[as]); else tc_commit(&tail->tmem_full[as]);
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
    // rv: residual segments of the current chunk — or, in LayerNorm-fold mode (no residual there), the NEXT tile's
    // row-statistics lines, requested a whole tile ahead because global-load latency is fully exposed in this
    // 2-warps-per-scheduler epilogue.
    int4 rv[8];
    constexpr bool fold_mode = !kGatherA && LN == 2;
    constexpr int ln_slots = kStatSlots;
    auto load_stats = [&](int tile) {
      const int mb = (tile / p.tiles_n) * kTileRows + m_off + q * 32 + (lane >> 3);
#pragma unroll
      for (int it = 0; it < 8; ++it) {
        const int gm = mb + it * 4;
        rv[it] = make_int4(0, 0, 0, 0);
        if (tile < num_tiles && gm < p.M)
          rv[it] = __ldg(reinterpret_cast<const int4*>(p.ln_stats + (int64_t)gm * ln_slots) + (lane & 7));
      }
    };
    if (fold_mode) load_stats(tile0);
    for (int t = tile0; t < num_tiles; t += tile_step) {
      const int m0 = (t / p.tiles_n) * kTileRows + m_off, n0 = (t % p.tiles_n) * BN;
      const __nv_bfloat16* bias_g = p.bias;
      if (p.bias && p.group_tiles_m) bias_g += (size_t)((t / p.tiles_n) / p.group_tiles_m) * p.group_n;
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
      // folded LayerNorm: this lane's row statistics from the producer GEMM's per-chunk partials (fixed order)
      float ln_rstd = 1.f;
      if (fold_mode) {
        float4 sv[8];   // this tile's statistics lines: requested during the previous tile (first tile: above)
#pragma unroll
        for (int it = 0; it < 8; ++it)
          sv[it] = make_float4(__int_as_float(rv[it].x), __int_as_float(rv[it].y), __int_as_float(rv[it].z),
                               __int_as_float(rv[it].w));
        float ps[8], pq[8];
#pragma unroll
        for (int it = 0; it < 8; ++it) {
          ps[it] = sv[it].x + sv[it].z, pq[it] = sv[it].y + sv[it].w;
        }
        const float ts = butterfly8(ps, lane & 7), tq = butterfly8(pq, lane & 7);  // row (lane&7)*4 + (lane>>3)
        const int src = ((lane & 3) << 3) | (lane >> 2);                            // lane that holds row `lane`
        const float s1 = __shfl_sync(0xffffffffu, ts, src), s2 = __shfl_sync(0xffffffffu, tq, src);
        const float mean = s1 * p.ln_inv_k;
        const float var = fmaxf(s2 * p.ln_inv_k - mean * mean, 0.f);
        ln_rstd = rsqrtf(var + p.ln_eps);
        load_stats(t + tile_step);   // rv is free again: request the next tile's lines now
      }
      mbar_wait(&tail->tmem_full[as], aphase);
      tc_fence_after();
      const uint32_t t_row = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(as * BN);
      float tile_ts = 0.f, tile_tq = 0.f;   // LN == 1: this warp's 128-column group of row statistics
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
          if (p.bias && col0 + j * 8 < p.N) bv[j] = __ldg(reinterpret_cast<const int4*>(bias_g + col0 + j * 8));
        }
        const int ch = lane & 7;
        const int col = col0 + ch * 8;
        int out_rows[8];
#pragma unroll
        for (int it = 0; it < 8; ++it) {
          const int r = it * 4 + (lane >> 3);
          out_rows[it] = __shfl_sync(0xffffffffu, my_out, r);
          if (col >= p.N) out_rows[it] = -1;
          if (!fold_mode) rv[it] = make_int4(0, 0, 0, 0);
          if (!fold_mode && p.residual && out_rows[it] >= 0) {
            const int gm = m0 + q * 32 + r;
            // the modular row (positional embeddings) costs an integer division per row: compiled only where used
            constexpr bool kModRes = kGatherA || LN == 3;
            const int64_t rr = (int64_t)((kModRes && p.res_mod > 0 ? gm % p.res_mod : gm) + p.res_row_off);
            rv[it] = __ldg(reinterpret_cast<const int4*>(p.residual + rr * p.ldr + col));
          }
        }
        tc_wait_ld();
        // bias + activation in fp32, on the bf16-rounded pre-activation (nn.Linear output rounding)
        constexpr bool fold = fold_mode;
        const uint64_t rstd2 = f32x2(ln_rstd, ln_rstd);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const uint32_t bw[4] = {(uint32_t)bv[j].x, (uint32_t)bv[j].y, (uint32_t)bv[j].z, (uint32_t)bv[j].w};
          uint4 o;
          uint32_t* ow = reinterpret_cast<uint32_t*>(&o);
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const uint64_t a2 = f32x2(__uint_as_float(acc[j * 8 + 2 * e]), __uint_as_float(acc[j * 8 + 2 * e + 1]));
            const uint64_t b2 = f32x2(bf16_lo(bw[e]), bf16_hi(bw[e]));
            const uint64_t x2 = fold ? fma_f32x2(rstd2, a2, b2) : add_f32x2(a2, b2);  // fold: rstd*(x·W''ᵀ) + b'
            float x0, x1;
            f32x2_unpack(x2, x0, x1);
            if (ACT != MM_ACT_NONE && ACT != MM_ACT_RELU) {   // (ReLU is applied last, after the residual)
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
        constexpr bool want_stats = !kGatherA && LN == 1;
        float st_s[8], st_q[8];
#pragma unroll
        for (int it = 0; it < 8; ++it) {
          const int r = it * 4 + (lane >> 3);
          uint4 v = *reinterpret_cast<const uint4*>(stg + r * 128 + ((ch ^ (r & 7)) << 4));
          st_s[it] = 0.f, st_q[it] = 0.f;
          if (out_rows[it] >= 0) {
            if (!fold_mode && p.residual) {  // bf16 + bf16, one rounding (HADD2.BF16)
              v.x = add_bf16x2(v.x, (uint32_t)rv[it].x);
              v.y = add_bf16x2(v.y, (uint32_t)rv[it].y);
              v.z = add_bf16x2(v.z, (uint32_t)rv[it].z);
              v.w = add_bf16x2(v.w, (uint32_t)rv[it].w);
            }
            if (ACT == MM_ACT_RELU) {   // relu(bf16(acc + bias) [+ res]): packed bf16 max against +0
              asm("max.bf16x2 %0, %0, %1;" : "+r"(v.x) : "r"(0u));
              asm("max.bf16x2 %0, %0, %1;" : "+r"(v.y) : "r"(0u));
              asm("max.bf16x2 %0, %0, %1;" : "+r"(v.z) : "r"(0u));
              asm("max.bf16x2 %0, %0, %1;" : "+r"(v.w) : "r"(0u));
            }
            *reinterpret_cast<uint4*>(p.C + (int64_t)out_rows[it] * p.ldc + col) = v;
            if (want_stats) {  // statistics of the values as stored (bf16-rounded), fp32, two lanes per instruction
              const uint32_t vw[4] = {v.x, v.y, v.z, v.w};
              uint64_t s2 = 0ull, q2 = 0ull;
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                const uint64_t ab = f32x2(bf16_lo(vw[e]), bf16_hi(vw[e]));
                s2 = add_f32x2(s2, ab);
                q2 = fma_f32x2(ab, ab, q2);
              }
              float a, b;
              f32x2_unpack(s2, a, b);
              st_s[it] = a + b;
              f32x2_unpack(q2, a, b);
              st_q[it] = a + b;
            }
          }
        }
        if (want_stats) {
          // after the butterfly this lane owns row (ch*4 + lane/8) of the warp's 32 rows; chunks add in a fixed order
          tile_ts += butterfly8(st_s, ch);
          tile_tq += butterfly8(st_q, ch);
        }
        __syncwarp();
      }
      if (!kGatherA && LN == 1) {
        const int gcol = n0 + half * (BN / 2);
        const int orow = __shfl_sync(0xffffffffu, my_out, (lane & 7) * 4 + (lane >> 3));
        if (orow >= 0 && gcol < p.N) p.stats_out[(int64_t)orow * kStatSlots + (gcol >> 7)] = make_float2(tile_ts, tile_tq);
      }
      // all TMEM reads of this accumulator stage are complete (tc_wait_ld above)
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if (kPair) mbar_arrive_cta(&tail->tmem_empty[as], 0);   // the leader's MMA thread owns the accumulator ring
        else mbar_arrive(&tail->tmem_empty[as]);
      }
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
    for (int t = tile0; t < num_tiles; t += tile_step) {
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
          uint32_t w4[4] = {0u, 0u, 0u, 0u};   // 8 bf16 of this row's k-group
          const int g = kb * (BK / 8) + j;
          const int64_t a0 = base + s_koff[g * 8];
          if (base >= 0 && s_kcontig[g] && (a0 & 1) == 0) {
            // 8 consecutive pixels: vector loads as wide as the alignment allows (patch 32: always 16 bytes)
            if (p.pixel_is_f32) {
              if ((a0 & 3) == 0) {
                const float4 u = __ldg(reinterpret_cast<const float4*>(px32 + a0));
                const float4 v = __ldg(reinterpret_cast<const float4*>(px32 + a0) + 1);
                w4[0] = pack_bf16x2(u.x, u.y), w4[1] = pack_bf16x2(u.z, u.w);
                w4[2] = pack_bf16x2(v.x, v.y), w4[3] = pack_bf16x2(v.z, v.w);
              } else {
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                  const float2 u = __ldg(reinterpret_cast<const float2*>(px32 + a0) + e);
                  w4[e] = pack_bf16x2(u.x, u.y);
                }
              }
            } else {
              if ((a0 & 7) == 0) {
                const uint4 u = __ldg(reinterpret_cast<const uint4*>(px16 + a0));
                w4[0] = u.x, w4[1] = u.y, w4[2] = u.z, w4[3] = u.w;
              } else {
#pragma unroll
                for (int e = 0; e < 4; ++e) w4[e] = __ldg(reinterpret_cast<const uint32_t*>(px16 + a0) + e);
              }
            }
          } else {
            float f[8];
#pragma unroll
            for (int e = 0; e < 8; ++e) {
              const int off = s_koff[g * 8 + e];
              float v = 0.f;
              if (base >= 0 && off >= 0) v = p.pixel_is_f32 ? __ldg(px32 + base + off) : __bfloat162float(px16[base + off]);
              f[e] = v;
            }
            w4[0] = pack_bf16x2(f[0], f[1]), w4[1] = pack_bf16x2(f[2], f[3]);
            w4[2] = pack_bf16x2(f[4], f[5]), w4[3] = pack_bf16x2(f[6], f[7]);
          }
          asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(dst + ((j ^ (row & 7)) << 4)), "r"(w4[0]),
                       "r"(w4[1]), "r"(w4[2]), "r"(w4[3])
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

  // pair: nobody leaves (or frees TMEM) while the pair's MMAs / multicast commits can still touch this CTA
  tc_fence_before();
  if (kPair) cluster_sync_all(); else __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    if (kPair) tmem_dealloc_pair(tmem_base, kTmemCols); else tmem_dealloc(tmem_base, kTmemCols);
  }
}

// Patch rows for the staged patch embed: A[m, k] = bf16(pixel[b, c, py*patch + ky, px*patch + kx]), m = (b, py, px),
// k = (c, ky, kx), zero for k >= k_real.  One thread per 8 consecutive k (one 16-byte store); the k -> pixel offset
// table lives in shared memory (no integer divisions per element).  The pixels are read once (HBM), the rows feed
// the CTA-pair GEMM through TMA — 8x faster end to end than gathering the same pixels once per 256-column tile from
// inside the GEMM's producer warps.
template <typename PixT>
__global__ void __launch_bounds__(256) patch_rows_kernel(const PixT* __restrict__ px, __nv_bfloat16* __restrict__ A,
                                                         int64_t M, int P, int grid, int img, int patch, int k_real,
                                                         int Kpad) {
  extern __shared__ int s_off[];
  const int pp = patch * patch;
  for (int k = threadIdx.x; k < Kpad; k += blockDim.x) {
    int off = -1;
    if (k < k_real) {
      const int c = k / pp, r = k - c * pp;
      const int ky = r / patch, kx = r - ky * patch;
      off = c * img * img + ky * img + kx;
    }
    s_off[k] = off;
  }
  __syncthreads();
  const int G8 = Kpad >> 3;
  const int64_t items = M * G8;
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < items; idx += (int64_t)gridDim.x * blockDim.x) {
    const int64_t m = idx / G8;
    const int g = (int)(idx - m * G8);
    const int64_t b = m / P;
    const int pi = (int)(m - b * P);
    const int py = pi / grid, pxi = pi - py * grid;
    const PixT* src = px + b * 3 * img * img + (int64_t)(py * patch) * img + pxi * patch;
    float v[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int off = s_off[g * 8 + j];
      v[j] = 0.f;
      if (off >= 0) {
        if constexpr (sizeof(PixT) == 4) v[j] = __ldg(reinterpret_cast<const float*>(src) + off);
        else v[j] = __bfloat162float(src[off]);
      }
    }
    uint4 o;
    o.x = pack_bf16x2(v[0], v[1]), o.y = pack_bf16x2(v[2], v[3]), o.z = pack_bf16x2(v[4], v[5]), o.w = pack_bf16x2(v[6], v[7]);
    *reinterpret_cast<uint4*>(A + m * Kpad + g * 8) = o;
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

template <bool kGatherA, int ACT, int LN, bool kPair>
static int launch_gemm_cfg(const CUtensorMap& tmA, const CUtensorMap& tmB, GemmParams& p, cudaStream_t st) {
  using Cfg = GemmCfg<kGatherA, kPair>;
  static std::atomic<bool> attr_set[64];   // per device, set once (idempotent call; atomics make the race benign)
  int dev = 0;
  MM_CUDA_CHECK(cudaGetDevice(&dev));
  auto* kern = gemm_bf16_tn_kernel<kGatherA, ACT, LN, kPair>;
  if (dev >= 0 && dev < 64 && !attr_set[dev].load(std::memory_order_acquire)) {
    MM_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::kSmemBytes));
    attr_set[dev].store(true, std::memory_order_release);
  }
  constexpr int kTileRows = kPair ? 2 * BM : BM;
  p.tiles_m = (p.M + kTileRows - 1) / kTileRows;
  p.tiles_n = (p.N + BN - 1) / BN;
  const int tiles = p.tiles_m * p.tiles_n;
  const int units = kPair ? sm_count() / 2 : sm_count();   // persistent: one CTA (or CTA pair) per SM (TPC)
  const int grid = (tiles < units ? tiles : units) * (kPair ? 2 : 1);
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(grid), cfg.blockDim = dim3(Cfg::kThreads);
  cfg.dynamicSmemBytes = Cfg::kSmemBytes, cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2, attr[0].val.clusterDim.y = 1, attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr, cfg.numAttrs = kPair ? 1 : 0;
  MM_CUDA_CHECK(cudaLaunchKernelEx(&cfg, kern, tmA, tmB, (const GemmParams)p));
  return MM_OK;
}

// CTA-pair kernel on/off: per call through MM_GEMM_FLAG_1CTA in the `act` argument (the pair-vs-single parity test),
// else the environment read ONCE (MM_B200_GEMM_PAIR=0 forces the 1-CTA kernel everywhere for A/B measurements), else
// on.  No mutable process state: concurrent launches from several threads / streams never see each other's choice.
static bool pair_enabled() {
  static const bool on = [] {
    const char* e = getenv("MM_B200_GEMM_PAIR");
    return !(e && e[0] == '0');
  }();
  return on;
}

template <bool kGatherA, int ACT, int LN = 0>
static int launch_gemm_inst(const CUtensorMap& tmA, const CUtensorMap& tmB, GemmParams& p, cudaStream_t st) {
  if constexpr (!kGatherA) {
    if (p.pair) return launch_gemm_cfg<false, ACT, LN, true>(tmA, tmB, p, st);
  }
  return launch_gemm_cfg<kGatherA, ACT, LN, false>(tmA, tmB, p, st);
}

static int launch_gemm(const CUtensorMap& tmA, const CUtensorMap& tmB, GemmParams& p, cudaStream_t st) {
  if (p.ln_stats) {   // consumer of a folded LayerNorm (QKV: no activation; fc1: QuickGELU / tanh-GELU)
    if (p.stats_out) return MM_ERR_BAD_ARG;
    switch (p.act) {
      case MM_ACT_NONE: return launch_gemm_inst<false, MM_ACT_NONE, 2>(tmA, tmB, p, st);
      case MM_ACT_QUICK_GELU: return launch_gemm_inst<false, MM_ACT_QUICK_GELU, 2>(tmA, tmB, p, st);
      case MM_ACT_GELU_TANH: return launch_gemm_inst<false, MM_ACT_GELU_TANH, 2>(tmA, tmB, p, st);
    }
    return MM_ERR_BAD_ARG;
  }
  if (p.stats_out) {  // producer of row statistics (out-proj / fc2: bias + residual, no activation)
    if (p.act != MM_ACT_NONE) return MM_ERR_BAD_ARG;
    return launch_gemm_inst<false, MM_ACT_NONE, 1>(tmA, tmB, p, st);
  }
  if (p.res_mod > 0) {
    if (p.act != MM_ACT_NONE) return MM_ERR_BAD_ARG;
    return launch_gemm_inst<false, MM_ACT_NONE, 3>(tmA, tmB, p, st);
  }
  switch (p.act) {
    case MM_ACT_NONE: return launch_gemm_inst<false, MM_ACT_NONE>(tmA, tmB, p, st);
    case MM_ACT_QUICK_GELU: return launch_gemm_inst<false, MM_ACT_QUICK_GELU>(tmA, tmB, p, st);
    case MM_ACT_GELU_ERF: return launch_gemm_inst<false, MM_ACT_GELU_ERF>(tmA, tmB, p, st);
    case MM_ACT_GELU_TANH: return launch_gemm_inst<false, MM_ACT_GELU_TANH>(tmA, tmB, p, st);
    case MM_ACT_RELU: return launch_gemm_inst<false, MM_ACT_RELU>(tmA, tmB, p, st);
  }
  return MM_ERR_BAD_ARG;
}

}  // namespace mm

using namespace mm;

static bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

static int gemm_entry(const void* A, int64_t lda, const void* W, int64_t ldw, const void* bias, void* C, int64_t ldc,
                      int64_t M, int64_t N, int64_t K, int act, const void* residual, int64_t ldr, int64_t res_mod,
                      int64_t res_row_off, const int32_t* out_row_map, int64_t n_out_rows, void* stats_out,
                      const void* ln_stats, float ln_eps, void* stream, int64_t groups = 1) {
  if (groups < 1 || groups > 64) return MM_ERR_BAD_ARG;
  const int64_t Mg = M;          // rows per group
  M *= groups;                   // rows of the stacked problem
  if (M < 0 || N <= 0 || K <= 0 || M > INT32_MAX / 2 || N * groups > INT32_MAX / 2 || K > INT32_MAX / 2) return MM_ERR_BAD_ARG;
  if (M == 0) return MM_OK;
  if (!A || !W || !C) return MM_ERR_BAD_ARG;
  if (K % 8 || N % 8 || lda % 8 || ldw % 8 || ldc % 8 || (residual && ldr % 8)) return MM_ERR_BAD_ARG;
  if (lda < K || ldw < K || ldc < N) return MM_ERR_BAD_ARG;
  if (!aligned16(A) || !aligned16(W) || !aligned16(C) || !aligned16(bias) || !aligned16(residual))
    return MM_ERR_BAD_ARG;
  const bool force_1cta = (act & MM_GEMM_FLAG_1CTA) != 0;
  act &= ~MM_GEMM_FLAG_1CTA;
  if (act < MM_ACT_NONE || act > MM_ACT_RELU) return MM_ERR_BAD_ARG;
  if (out_row_map && n_out_rows <= 0) return MM_ERR_BAD_ARG;
  CUtensorMap tmA, tmB;
  int rc = make_tmap_2d(&tmA, A, M, K, lda, BM);
  if (rc != MM_OK) return rc;
  const bool pair = M > BM && !force_1cta && pair_enabled();   // at most one row tile: the peer CTA would only see padding
  if (groups > 1 && (Mg % (pair ? 2 * BM : BM) || res_mod > 0)) return MM_ERR_BAD_ARG;   // a row tile must not straddle two groups
  rc = make_tmap_2d(&tmB, W, N * groups, K, ldw, pair ? BN / 2 : BN);
  if (rc != MM_OK) return rc;
  GemmParams p{};
  p.pair = pair ? 1 : 0;
  p.group_tiles_m = groups > 1 ? (int)(Mg / (pair ? 2 * BM : BM)) : 0;
  p.group_n = (int)N;
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
  if (ln_stats && residual) return MM_ERR_BAD_ARG;  // the epilogue's residual registers carry the statistics
  if ((ln_stats && K > 128 * kStatSlots) || (stats_out && N > 128 * kStatSlots)) return MM_ERR_BAD_ARG;
  if (!aligned16(ln_stats) || !aligned16(stats_out))
    return MM_ERR_BAD_ARG;
  p.stats_out = reinterpret_cast<float2*>(stats_out);
  p.ln_stats = reinterpret_cast<const float2*>(ln_stats);
  p.ln_eps = ln_eps;
  p.ln_inv_k = 1.0f / (float)K;
  return launch_gemm(tmA, tmB, p, reinterpret_cast<cudaStream_t>(stream));
}

extern "C" int mm_gemm_bias_act(const void* A, int64_t lda, const void* W, int64_t ldw, const void* bias, void* C,
                                int64_t ldc, int64_t M, int64_t N, int64_t K, int act, const void* residual,
                                int64_t ldr, int64_t res_mod, int64_t res_row_off, const int32_t* out_row_map,
                                int64_t n_out_rows, void* stream) {
  return gemm_entry(A, lda, W, ldw, bias, C, ldc, M, N, K, act, residual, ldr, res_mod, res_row_off, out_row_map,
                    n_out_rows, nullptr, nullptr, 0.f, stream);
}

extern "C" int mm_gemm_grouped(const void* A, int64_t lda, const void* W, int64_t ldw, const void* bias, void* C, int64_t ldc,
                               int64_t M, int64_t N, int64_t K, int64_t groups, int act, const void* residual, int64_t ldr,
                               const int32_t* out_row_map, int64_t n_out_rows, void* stats_out, const void* ln_stats,
                               float ln_eps, void* stream) {
  return gemm_entry(A, lda, W, ldw, bias, C, ldc, M, N, K, act, residual, ldr, 0, 0, out_row_map, n_out_rows, stats_out,
                    ln_stats, ln_eps, stream, groups);
}

extern "C" int64_t mm_gemm_row_stats_slots(int64_t N) { return (N > 0 && N <= 128 * kStatSlots) ? kStatSlots : -1; }

extern "C" int mm_gemm_ln_bias_act(const void* A, int64_t lda, const void* W, int64_t ldw, const void* bias, void* C,
                                   int64_t ldc, int64_t M, int64_t N, int64_t K, int act, const void* residual,
                                   int64_t ldr, const int32_t* out_row_map, int64_t n_out_rows, void* stats_out,
                                   const void* ln_stats, float ln_eps, void* stream) {
  return gemm_entry(A, lda, W, ldw, bias, C, ldc, M, N, K, act, residual, ldr, 0, 0, out_row_map, n_out_rows, stats_out,
                    ln_stats, ln_eps, stream);
}

extern "C" int64_t mm_patch_embed_workspace_bytes(int64_t B, int64_t img, int64_t patch) {
  if (B < 0 || img <= 0 || patch <= 0 || img < patch) return -1;
  const int64_t G = img / patch, Kpad = (3 * patch * patch + BK - 1) / BK * BK;
  return B * G * G * Kpad * 2;
}

static int patch_embed_impl(const void* pixels, int pixel_dtype, int64_t B, int64_t img, int64_t patch,
                            const void* weight, int64_t ldw, const void* conv_bias, const void* pos_emb,
                            const void* cls, void* out, int64_t D, void* workspace, int64_t workspace_bytes,
                            void* stream) {
  if (B < 0 || img <= 0 || patch <= 0 || img < patch || D <= 0 || D % 8) return MM_ERR_BAD_ARG;  // img % patch != 0 is fine: the conv (stride = kernel) drops the remainder, SigLIP 384/14 -> 27x27
  if (B == 0) return MM_OK;
  if (!pixels || !weight || !pos_emb || !out) return MM_ERR_BAD_ARG;
  if (pixel_dtype != MM_F32 && pixel_dtype != MM_BF16) return MM_ERR_BAD_ARG;
  const int64_t G = img / patch, P = G * G, Kreal = 3 * patch * patch;
  const int64_t Kpad = (Kreal + BK - 1) / BK * BK;
  if (Kpad > kKTableMax || ldw < Kpad || ldw % 8) return MM_ERR_BAD_ARG;  // weight must be zero-padded to Kpad
  if (B * P > INT32_MAX / 2 || B * 3 * img * img > (int64_t)INT32_MAX * 4) return MM_ERR_BAD_ARG;
  if (!aligned16(pixels) || !aligned16(weight) || !aligned16(conv_bias) || !aligned16(pos_emb) || !aligned16(out))
    return MM_ERR_BAD_ARG;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const int cls_rows = cls ? 1 : 0;
  const int64_t T = P + cls_rows;
  const bool staged = workspace != nullptr;   // patch rows staged as a bf16 [B*P, Kpad] matrix, then the TMA GEMM
  if (staged && (workspace_bytes < B * P * Kpad * 2 || !aligned16(workspace))) return MM_ERR_BAD_ARG;
  const bool pair = staged && B * P > BM && pair_enabled();
  CUtensorMap tmA, tmB;
  int rc = make_tmap_2d(&tmB, weight, D, Kpad, ldw, pair ? BN / 2 : BN);
  if (rc != MM_OK) return rc;
  if (staged) {
    rc = make_tmap_2d(&tmA, workspace, B * P, Kpad, Kpad, BM);
    if (rc != MM_OK) return rc;
    const int64_t items = B * P * (Kpad / 8);
    int64_t blocks = (items + 255) / 256;
    if (blocks > (int64_t)sm_count() * 16) blocks = (int64_t)sm_count() * 16;
    __nv_bfloat16* A = reinterpret_cast<__nv_bfloat16*>(workspace);
    if (pixel_dtype == MM_F32)
      patch_rows_kernel<float><<<(int)blocks, 256, Kpad * sizeof(int), st>>>(
          reinterpret_cast<const float*>(pixels), A, B * P, (int)P, (int)G, (int)img, (int)patch, (int)Kreal, (int)Kpad);
    else
      patch_rows_kernel<__nv_bfloat16><<<(int)blocks, 256, Kpad * sizeof(int), st>>>(
          reinterpret_cast<const __nv_bfloat16*>(pixels), A, B * P, (int)P, (int)G, (int)img, (int)patch, (int)Kreal,
          (int)Kpad);
    MM_CUDA_CHECK(cudaGetLastError());
  }
  GemmParams p{};
  p.pair = pair ? 1 : 0;
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
  p.n_out_rows = B * T;
  rc = staged ? launch_gemm(tmA, tmB, p, st)   // modular-residual variant (LN = 3), formula row map
              : launch_gemm_inst<true, MM_ACT_NONE>(tmB, tmB, p, st);
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

extern "C" int mm_patch_embed(const void* pixels, int pixel_dtype, int64_t B, int64_t img, int64_t patch,
                              const void* weight, int64_t ldw, const void* conv_bias, const void* pos_emb,
                              const void* cls, void* out, int64_t D, void* stream) {
  return patch_embed_impl(pixels, pixel_dtype, B, img, patch, weight, ldw, conv_bias, pos_emb, cls, out, D, nullptr, 0,
                          stream);
}

extern "C" int mm_patch_embed_ws(const void* pixels, int pixel_dtype, int64_t B, int64_t img, int64_t patch,
                                 const void* weight, int64_t ldw, const void* conv_bias, const void* pos_emb,
                                 const void* cls, void* out, int64_t D, void* workspace, int64_t workspace_bytes,
                                 void* stream) {
  if (!workspace) return MM_ERR_BAD_ARG;
  return patch_embed_impl(pixels, pixel_dtype, B, img, patch, weight, ldw, conv_bias, pos_emb, cls, out, D, workspace,
                          workspace_bytes, stream);
}
