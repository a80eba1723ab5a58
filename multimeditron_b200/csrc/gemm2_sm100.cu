// gemm2_sm100.cu — EXPERIMENTAL CTA-pair (cta_group::2) variant of the persistent tcgen05 GEMM:  C = A·Wᵀ + bias.
// Two CTAs of one cluster (one TPC) compute a 256 x 256 tile: each CTA holds its own 128 rows of A and HALF of the
// B tile (128 of the 256 weight rows); one `tcgen05.mma.cta_group::2` issued by the leader CTA feeds both SMs' tensor
// cores from both shared memories, so per-SM shared-memory operand traffic per FLOP halves (B is read once for 256
// rows instead of once per 128) and the smaller stages allow a 6-deep ring.
// Protocol (as in CUTLASS's 2-SM pipelines): every TMA load (both CTAs) signals the LEADER's full barrier
// (barrier address with the peer bit cleared); the leader's producer posts the expected bytes of both CTAs;
// tcgen05.commit multicasts to both CTAs' empty / tmem_full barriers; the peer's epilogue warps release the
// accumulator stage on the leader's tmem_empty barrier (mapa + remote arrive).
// Entry point: mm_gemm2_bias (plain bias epilogue only) — used by tools/gemm2_bench.py to decide whether the pair
// variant is worth carrying through all epilogue variants of gemm_sm100.cu.
#include "common.cuh"

namespace mm {

int make_tmap_2d(CUtensorMap* tm, const void* ptr, int64_t rows, int64_t cols, int64_t ld, int box_rows);

namespace pair {

constexpr int BM = 128;         // rows per CTA (256 per pair)
constexpr int BN = 256;
constexpr int BNH = BN / 2;     // B rows held by each CTA
constexpr int BK = 64;
constexpr int kStages = 6;
constexpr int kAStage = BM * BK * 2;    // 16 KiB
constexpr int kBStage = BNH * BK * 2;   // 16 KiB
constexpr int kEpiWarps = 8;
constexpr int kEpiWarpBytes = 32 * 128;
constexpr int kThreads = 128 + kEpiWarps * 32;
constexpr int kSmemBytes = 1024 + kStages * (kAStage + kBStage) + kEpiWarps * kEpiWarpBytes + 256;
constexpr uint32_t kPeerMask = 0xFEFFFFFFu;   // clears the CTA-rank bit of a shared::cluster address (CTA 0 of the pair)

struct Tail {
  uint64_t full[kStages], empty[kStages], tmem_full[2], tmem_empty[2];
  uint32_t tmem_base;
};
static_assert(sizeof(Tail) <= 256, "tail");

struct Params {
  __nv_bfloat16* C;
  const __nv_bfloat16* bias;
  int64_t ldc;
  int M, N, K, tiles_m2, tiles_n;
};

__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void tma_load_2d_pair(void* smem_dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(m)), "r"(smem_u32(bar) & kPeerMask), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tc_commit_pair(uint64_t* bar) {   // arrives on BOTH CTAs' barrier at this offset
  asm volatile(
      "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(
          smem_u32(bar)),
      "h"((uint16_t)3)
      : "memory");
}
__device__ __forceinline__ void tc_mma_f16_ss_pair(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                                   uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t"
      "}\n" ::"r"(d_tmem),
      "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void mbar_arrive_cta(uint64_t* bar, uint32_t cta) {   // arrive on `bar` of CTA `cta` of the cluster
  asm volatile(
      "{\n\t"
      ".reg .b32 ra;\n\t"
      "mapa.shared::cluster.u32 ra, %0, %1;\n\t"
      "mbarrier.arrive.shared::cluster.b64 _, [ra];\n\t"
      "}\n" ::"r"(smem_u32(bar)),
      "r"(cta)
      : "memory");
}

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads, 1)
    gemm_pair_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const Params p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sA = smem;
  uint8_t* sB = smem + kStages * kAStage;
  uint8_t* sEpi = sB + kStages * kBStage;
  Tail* tail = reinterpret_cast<Tail*>(sEpi + kEpiWarps * kEpiWarpBytes);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const bool leader = rank == 0;
  const int pair_id = blockIdx.x >> 1, num_pairs = gridDim.x >> 1;
  const int num_kb = (p.K + BK - 1) / BK;
  const int num_tiles = p.tiles_m2 * p.tiles_n;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < kStages; ++i) {
      mbar_init(&tail->full[i], 1);    // the leader producer's arrive.expect_tx (bytes of both CTAs)
      mbar_init(&tail->empty[i], 1);   // one multicast commit
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&tail->tmem_full[i], 1);
      mbar_init(&tail->tmem_empty[i], 2 * kEpiWarps);   // epilogue warps of BOTH CTAs (used on the leader only)
    }
    fence_barrier_init();
  }
  cluster_sync_all();   // barriers of both CTAs are initialised before anything can signal them
  if (warp == 2) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tail->tmem_base)),
                 "r"(512)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tmem_base = tail->tmem_base;

  if (warp == 0) {
    // =========================== TMA producer (both CTAs) ===========================
    if (elect_one_sync()) {
      int stage = 0;
      uint32_t phase = 0;
      for (int t = pair_id; t < num_tiles; t += num_pairs) {
        const int m0 = (t / p.tiles_n) * (2 * BM) + (int)rank * BM;
        const int n0 = (t % p.tiles_n) * BN + (int)rank * BNH;
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&tail->empty[stage], phase ^ 1);
          if (leader) mbar_arrive_expect_tx(&tail->full[stage], 2 * (kAStage + kBStage));
          tma_load_2d_pair(sA + stage * kAStage, &tmA, &tail->full[stage], kb * BK, m0);
          tma_load_2d_pair(sB + stage * kBStage, &tmB, &tail->full[stage], kb * BK, n0);
          if (++stage == kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    // =========================== MMA issuer (leader CTA, single thread) ===========================
    if (leader && elect_one_sync()) {
      constexpr uint32_t idesc = umma_idesc_bf16(2 * BM, BN, 0, 0);
      int stage = 0, as = 0;
      uint32_t phase = 0, aphase = 0;
      for (int t = pair_id; t < num_tiles; t += num_pairs) {
        mbar_wait(&tail->tmem_empty[as], aphase ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)(as * BN);
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(&tail->full[stage], phase);
          tc_fence_after();
          const uint32_t a_addr = smem_u32(sA + stage * kAStage);
          const uint32_t b_addr = smem_u32(sB + stage * kBStage);
#pragma unroll
          for (int k = 0; k < BK / 16; ++k) {
            const uint64_t da = umma_desc_sw128(a_addr + k * 32, 16, 1024);
            const uint64_t db = umma_desc_sw128(b_addr + k * 32, 16, 1024);
            tc_mma_f16_ss_pair(d_tmem, da, db, idesc, (kb | k) != 0 ? 1u : 0u);
          }
          tc_commit_pair(&tail->empty[stage]);   // frees the slot in BOTH CTAs when these MMAs retire
          if (++stage == kStages) {
            stage = 0;
            phase ^= 1;
          }
        }
        tc_commit_pair(&tail->tmem_full[as]);    // accumulators (both CTAs' TMEM) ready
        if (++as == 2) {
          as = 0;
          aphase ^= 1;
        }
      }
    }
  } else if (warp >= 4) {
    // =========================== epilogue (both CTAs: own 128 rows x 256 columns) ===========================
    const int q = warp & 3;
    const int half = (warp - 4) >> 2;
    uint8_t* stg = sEpi + (warp - 4) * kEpiWarpBytes;
    int as = 0;
    uint32_t aphase = 0;
    for (int t = pair_id; t < num_tiles; t += num_pairs) {
      const int m0 = (t / p.tiles_n) * (2 * BM) + (int)rank * BM, n0 = (t % p.tiles_n) * BN;
      mbar_wait(&tail->tmem_full[as], aphase);
      tc_fence_after();
      const uint32_t t_row = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(as * BN);
#pragma unroll 1
      for (int cc = 0; cc < 2; ++cc) {
        const int c = half * 2 + cc;
        const int col0 = n0 + c * 64;
        if (col0 >= p.N) break;
        uint32_t acc[64];
        tmem_ld_32x32b_x32(t_row + c * 64, acc);
        tmem_ld_32x32b_x32(t_row + c * 64 + 32, acc + 32);
        int4 bv[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          bv[j] = make_int4(0, 0, 0, 0);
          if (p.bias && col0 + j * 8 < p.N) bv[j] = __ldg(reinterpret_cast<const int4*>(p.bias + col0 + j * 8));
        }
        const int ch = lane & 7;
        const int col = col0 + ch * 8;
        tc_wait_ld();
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const uint32_t bw[4] = {(uint32_t)bv[j].x, (uint32_t)bv[j].y, (uint32_t)bv[j].z, (uint32_t)bv[j].w};
          uint4 o;
          uint32_t* ow = reinterpret_cast<uint32_t*>(&o);
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const uint64_t a2 = f32x2(__uint_as_float(acc[j * 8 + 2 * e]), __uint_as_float(acc[j * 8 + 2 * e + 1]));
            const uint64_t b2 = f32x2(bf16_lo(bw[e]), bf16_hi(bw[e]));
            float x0, x1;
            f32x2_unpack(add_f32x2(a2, b2), x0, x1);
            ow[e] = pack_bf16x2(x0, x1);
          }
          *reinterpret_cast<uint4*>(stg + lane * 128 + ((j ^ (lane & 7)) << 4)) = o;
        }
        __syncwarp();
#pragma unroll
        for (int it = 0; it < 8; ++it) {
          const int r = it * 4 + (lane >> 3);
          const uint4 v = *reinterpret_cast<const uint4*>(stg + r * 128 + ((ch ^ (r & 7)) << 4));
          const int gm = m0 + q * 32 + r;
          if (gm < p.M && col < p.N) *reinterpret_cast<uint4*>(p.C + (int64_t)gm * p.ldc + col) = v;
        }
        __syncwarp();
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_cta(&tail->tmem_empty[as], 0);   // the leader's MMA thread owns the accumulator ring
      if (++as == 2) {
        as = 0;
        aphase ^= 1;
      }
    }
  }
  // teardown: nobody leaves (or frees TMEM) while the pair's MMAs / multicast commits can still touch this CTA
  tc_fence_before();
  cluster_sync_all();
  if (warp == 2) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512) : "memory");
  }
}

}  // namespace pair
}  // namespace mm

using namespace mm;

// C[M,N] = A[M,K]·W[N,K]ᵀ + bias   (bf16, fp32 accumulate) on CTA pairs.  Experimental: plain epilogue only.
extern "C" int mm_gemm2_bias(const void* A, int64_t lda, const void* W, int64_t ldw, const void* bias, void* C,
                             int64_t ldc, int64_t M, int64_t N, int64_t K, void* stream) {
  using namespace mm::pair;
  if (M <= 0 || N <= 0 || K <= 0 || !A || !W || !C) return MM_ERR_BAD_ARG;
  if (K % 8 || N % 8 || lda % 8 || ldw % 8 || ldc % 8) return MM_ERR_BAD_ARG;
  CUtensorMap tmA, tmB;
  int rc = make_tmap_2d(&tmA, A, M, K, lda, BM);
  if (rc != MM_OK) return rc;
  rc = make_tmap_2d(&tmB, W, N, K, ldw, BNH);
  if (rc != MM_OK) return rc;
  Params p{};
  p.C = reinterpret_cast<__nv_bfloat16*>(C);
  p.bias = reinterpret_cast<const __nv_bfloat16*>(bias);
  p.ldc = ldc;
  p.M = (int)M, p.N = (int)N, p.K = (int)K;
  p.tiles_m2 = (p.M + 2 * BM - 1) / (2 * BM);
  p.tiles_n = (p.N + BN - 1) / BN;
  static bool attr_set = false;
  if (!attr_set) {
    MM_CUDA_CHECK(cudaFuncSetAttribute(gemm_pair_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes));
    attr_set = true;
  }
  const int tiles = p.tiles_m2 * p.tiles_n;
  const int pairs_max = sm_count() / 2;
  const int pairs = tiles < pairs_max ? tiles : pairs_max;
  gemm_pair_kernel<<<2 * pairs, kThreads, kSmemBytes, reinterpret_cast<cudaStream_t>(stream)>>>(tmA, tmB, p);
  MM_CUDA_CHECK(cudaGetLastError());
  return MM_OK;
}
