// attention_sm100.cu — non-causal multi-head attention forward for ViT towers (CLIP head_dim 64, SigLIP head_dim 72).
//
// One persistent CTA per SM; work unit = (image, head).  The CTA walks the unit's 128-row query tiles TWO AT A
// TIME ("streams", one softmax warpgroup each) so that one stream's MUFU-bound softmax overlaps the other
// stream's tensor-core work.  K/V tiles travel through a shared-memory ring; when the whole unit fits in the
// ring (CLIP: T <= 6*128) it is loaded once per unit and reused by every query pair, otherwise (SigLIP, T=729)
// it is streamed once per query pair (L2 hits).
//   S = Q·Kᵀ      tcgen05.mma 128 x BKV x 16, SS form (K-major SWIZZLE_128B smem operands), fp32 in TMEM
//   P = softmax   one query row per thread; the whole S row is pulled into registers with ONE pass of
//                 tcgen05.ld, which frees the S accumulator for the next QKᵀ before any exp is computed;
//                 online max/sum in the exp2 domain; the running max only moves when it grows by > 8 (log2
//                 units), so the O accumulator is rescaled rarely; P goes back to TMEM as packed bf16
//   O += P·V      tcgen05.mma 128 x HD x 16, TS form: A = P from TMEM, B = V consumed MN-major directly from
//                 its [kv, head_dim] TMA tile (no transpose pass); fp32 in TMEM
// head_dim 72 runs as HD = 80: the fused QKV GEMM writes each head into an 80-column slot whose last 8 columns
// are exact zeros (zero weight rows, packed at load time), the first 64 columns of a head travel as a
// SWIZZLE_128B tile and columns 64..79 as a SWIZZLE_32B tile (one extra k-step for QKᵀ, one extra N=16 MMA
// per k-step for PV); only the 72 real columns of O are stored.  BKV is 96 there so that TMEM holds
// 2 x (S 96 + P 48 + O 80) = 448 columns.
// TMEM map: S0 | S1 | P0 | P1 | O0 | O1 (hd64: 128,128,64,64,64,64 = 512 columns).
//
// Replaces HF:models/clip/modeling_clip.py:318-331 and HF:models/siglip/modeling_siglip.py:285-300
// (SDPA; scale = head_dim**-0.5, no mask, dropout 0).
#include "common.cuh"

namespace mm {

constexpr int kBQ = 128;
constexpr int kAttnThreads = 384;          // w0 TMA, w1 MMA stream 0, w2 TMEM alloc, w3 MMA stream 1, w4-7 / w8-11 softmax of stream 0 / 1
constexpr float kRescaleThreshold = 8.0f;  // log2 units: P <= 2^8, harmless in bf16/fp32
constexpr int kMaxStages = 8;
#ifndef MM_ATTN_POLL_NS
#define MM_ATTN_POLL_NS 32
#endif
constexpr unsigned kPollSleepNs = MM_ATTN_POLL_NS;  // idle back-off of the two polling service threads
// MM_ATTN_TOKEN (default 1 = ON, the build that ships and that every test and bench number comes from): the two
// streams take turns in the exp phase — a named-barrier token passed between the two softmax warps that share a
// scheduler — so one stream's exps run while the other stream's S tile is produced.  -DMM_ATTN_TOKEN=0 lets both
// streams exponentiate whenever they like; kept as an A/B switch only (a micro-benchmark, tools/micro/tmem_probe2,
// had two interleaved warps at 8.7 clk per ex2 against 10.5 for a lone warp, but in the whole kernel the token's
// regular hand-over wins).
#ifndef MM_ATTN_TOKEN
#define MM_ATTN_TOKEN 1
#endif
// Every kv tile decides its reference max from ITS OWN scores before its exps run.  (A "stale max" variant, which
// took the reference from the previous tiles and re-exponentiated a tile whose scores jumped past it, was measured
// in round 2, never shipped, and has been removed.)
// MM_ATTN_PACKED=1: the scale-and-shift before each ex2 and the row-sum adds run as packed fp32x2 instructions
// (FFMA2 / FADD2): 2.5 instead of 3.5 issue slots per score in the exp phase.
#ifndef MM_ATTN_PACKED
#define MM_ATTN_PACKED 1
#endif

template <int HD, int BKV>
struct AttnCfg {
  static_assert(HD == 64 || HD == 80, "head slots of 64 or 80 columns");
  static_assert(BKV % 32 == 0 && BKV <= 128, "kv tile");
  static constexpr int kHDX = HD - 64;                          // extra columns (SWIZZLE_32B tile)
  static constexpr int kQMain = kBQ * 128;                      // bytes: [128 x 64] bf16
  static constexpr int kQX = kHDX ? kBQ * 32 : 0;               //        [128 x 16] bf16
  static constexpr int kKMain = BKV * 128;
  static constexpr int kKX = kHDX ? BKV * 32 : 0;
  static constexpr int kStageBytes = 2 * (kKMain + kKX);        // K + V
  static constexpr int kQBytes = 2 * (kQMain + kQX);            // two streams
  static constexpr int kBarBytes = 512;
  static constexpr int kStagesRaw = (232448 - 1024 - kBarBytes - kQBytes) / kStageBytes;
  static constexpr int kStages = kStagesRaw > kMaxStages ? kMaxStages : kStagesRaw;
  static constexpr int kSmemBytes = 1024 + kQBytes + kStages * kStageBytes + kBarBytes;
  // TMEM columns
  static constexpr int kColS = 0;                 // + s*BKV
  static constexpr int kColP = 2 * BKV;           // + s*BKV/2
  static constexpr int kColO = 3 * BKV;           // + s*HD
  static_assert(3 * BKV + 2 * HD <= 512, "TMEM budget");
  static_assert(kStages >= 2, "ring");
};

struct AttnBars {
  uint64_t kv_full[kMaxStages], kv_empty[kMaxStages];
  uint64_t q_full[2], q_empty[2];
  uint64_t s_full[2], s_empty[2];
  uint64_t p_full[2], pv_done[2];
  uint64_t o_empty[2];
  uint32_t tmem_base;
};
static_assert(sizeof(AttnBars) <= 512, "bars");

struct AttnMaps {
  CUtensorMap q_main, kv_main, q_x, kv_x;
};

struct AttnParams {
  __nv_bfloat16* out;  // [B*T, ld_out]
  int B, T, heads;
  int Dp;              // heads * HD : width of each of the q | k | v sections of a qkv row
  int hd_out;          // real head_dim (64 or 72)
  int ld_out;          // heads * hd_out
  float scale_log2;    // scale * log2(e)
};

struct RingPos {
  int stage;
  uint32_t phase;
  template <int NS>
  __device__ __forceinline__ void advance() {
    if (++stage == NS) {
      stage = 0;
      phase ^= 1u;
    }
  }
};

__device__ __forceinline__ float ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// exp phase of 4 adjacent scores: p = 2^(s*sl2 - m), two packed bf16x2 words out, row sums kept as two fp32x2 pairs
struct ExpAcc {
  uint64_t sa, sb;     // packed partial row sums (MM_ATTN_PACKED)
  float s0, s1, s2, s3;
};
__device__ __forceinline__ void exp4(const uint32_t* sv, uint32_t* pk, float sl2, float m_scaled, ExpAcc& a) {
#if MM_ATTN_PACKED
  const uint64_t k2 = f32x2(sl2, sl2), nm2 = f32x2(-m_scaled, -m_scaled);
  float x0, x1, x2, x3;
  f32x2_unpack(fma_f32x2(f32x2(__uint_as_float(sv[0]), __uint_as_float(sv[1])), k2, nm2), x0, x1);
  f32x2_unpack(fma_f32x2(f32x2(__uint_as_float(sv[2]), __uint_as_float(sv[3])), k2, nm2), x2, x3);
  const float p0 = ex2(x0), p1 = ex2(x1), p2 = ex2(x2), p3 = ex2(x3);
  a.sa = add_f32x2(a.sa, f32x2(p0, p1));
  a.sb = add_f32x2(a.sb, f32x2(p2, p3));
#else
  const float p0 = ex2(fmaf(__uint_as_float(sv[0]), sl2, -m_scaled));
  const float p1 = ex2(fmaf(__uint_as_float(sv[1]), sl2, -m_scaled));
  const float p2 = ex2(fmaf(__uint_as_float(sv[2]), sl2, -m_scaled));
  const float p3 = ex2(fmaf(__uint_as_float(sv[3]), sl2, -m_scaled));
  a.s0 += p0, a.s1 += p1, a.s2 += p2, a.s3 += p3;
#endif
  pk[0] = pack_bf16x2(p0, p1);
  pk[1] = pack_bf16x2(p2, p3);
}
__device__ __forceinline__ void exp_acc_reset(ExpAcc& a) {
  a.sa = a.sb = 0ull;
  a.s0 = a.s1 = a.s2 = a.s3 = 0.f;
}
__device__ __forceinline__ float exp_acc_total(const ExpAcc& a) {
#if MM_ATTN_PACKED
  float u0, u1, u2, u3;
  f32x2_unpack(a.sa, u0, u1);
  f32x2_unpack(a.sb, u2, u3);
  return (u0 + u1) + (u2 + u3);
#else
  return (a.s0 + a.s1) + (a.s2 + a.s3);
#endif
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

// Optional phase trace (tools/attn_trace.py; compiled in only with -DMM_ATTN_TRACE): block 0 records clock64()
// stamps of one softmax warp per stream and of the MMA issuer.
#ifdef MM_ATTN_TRACE
__device__ long long* g_attn_trace = nullptr;  // 3 roles x 1024 (id, clock) pairs; fire-and-forget stores
__device__ __forceinline__ void attn_trace(int role, uint32_t& k, int id) {
  if (g_attn_trace && k < 1024) {
    g_attn_trace[(role * 1024 + k) * 2] = id;
    g_attn_trace[(role * 1024 + k) * 2 + 1] = clock64();
    ++k;
  }
}
#define TRACE_DECL uint32_t trc_k = 0;
#define TRACE_SM(id) do { if (blockIdx.x == 0 && q == 0 && lane == 0) attn_trace(s, trc_k, (id) + 16 * s); } while (0)
#define TRACE_MMA(id) do { if (blockIdx.x == 0) attn_trace(2, trc_k, id); } while (0)
#else
#define TRACE_DECL
#define TRACE_SM(id)
#define TRACE_MMA(id)
#endif

// non-blocking phase test (event loops of the producer and the MMA issuer poll several barriers)
__device__ __forceinline__ bool mbar_test(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.b32 %0, 1, 0, p;\n\t"
      "}\n"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
// softmax-side wait: MM_ATTN_SPIN=1 polls with the non-blocking test_wait (243 clk ping-pong round trip measured)
// instead of try_wait (375 clk: the hardware suspends the thread for a time slice)
#ifndef MM_ATTN_SPIN
#define MM_ATTN_SPIN 0
#endif
__device__ __forceinline__ void sm_wait(uint64_t* bar, uint32_t parity) {
#if MM_ATTN_SPIN
  while (!mbar_test(bar, parity)) {
  }
#else
  mbar_wait(bar, parity);
#endif
}
__device__ __forceinline__ void named_bar_sync(int id, int count) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(count) : "memory");
}
__device__ __forceinline__ void named_bar_arrive(int id, int count) {
  asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(count) : "memory");
}

// Work decomposition shared by every role.  The CTA owns units u_i = blockIdx.x + i*gridDim.x.  A K/V "group" is
// what travels through the ring together: the whole unit when it fits (resident), else one query pair of it.
// Stream s owns, in group g, the query tiles first_qt(s,g) + 2j, j < ntiles(s,g).  Both streams always work inside
// the same group (the exp-phase token below keeps them half a step apart), so ring slots are never awaited across
// groups by a stream that runs ahead.
struct AttnSched {
  int nkv, nq, npair, n_units, ngroups;
  bool resident;
  __device__ __forceinline__ int unit_ord(int g) const { return resident ? g : g / npair; }
  __device__ __forceinline__ int first_qt(int s, int g) const {
    return resident ? s : 2 * (g - unit_ord(g) * npair) + s;
  }
  __device__ __forceinline__ int slots(int /*g*/) const { return resident ? npair : 1; }  // tile slots per stream
  __device__ __forceinline__ int ntiles(int s, int g) const {
    const int f = first_qt(s, g);
    if (f >= nq) return 0;
    return resident ? (nq - f + 1) >> 1 : 1;
  }
};

template <int HD, int BKV>
__global__ void __launch_bounds__(kAttnThreads, 1)
    attention_fwd_kernel(const __grid_constant__ AttnMaps maps, const AttnParams p) {
  using Cfg = AttnCfg<HD, BKV>;
  constexpr int NS = Cfg::kStages;
  constexpr int HDX = Cfg::kHDX;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sQ = smem;                                  // [Q0 main][Q1 main][Q0 x][Q1 x]
  uint8_t* sQX = sQ + 2 * Cfg::kQMain;
  uint8_t* sKV = sQ + Cfg::kQBytes;                    // per stage: [K main][V main][K x][V x]
  AttnBars* bars = reinterpret_cast<AttnBars*>(sKV + NS * Cfg::kStageBytes);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int units = p.B * p.heads;
  AttnSched sc;
  sc.nkv = (p.T + BKV - 1) / BKV;
  sc.nq = (p.T + kBQ - 1) / kBQ;
  sc.npair = (sc.nq + 1) >> 1;
  sc.resident = sc.nkv <= NS;  // whole unit fits in the ring: load K/V once per unit
  sc.n_units = ((int)blockIdx.x < units) ? (units - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;
  sc.ngroups = sc.n_units * (sc.resident ? 1 : sc.npair);
  const int nkv = sc.nkv;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&maps.q_main);
    tma_prefetch_desc(&maps.kv_main);
    if (HDX) {
      tma_prefetch_desc(&maps.q_x);
      tma_prefetch_desc(&maps.kv_x);
    }
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < NS; ++i) {
      mbar_init(&bars->kv_full[i], 1);
      mbar_init(&bars->kv_empty[i], 2);  // one release per stream (real tile: tcgen05.commit; none: plain arrive)
    }
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

  // register re-split (launch: 384 x 168): the service warpgroup keeps 56, each softmax thread gets 224 so that a
  // whole 128-column S row plus its packed P row stay in registers (128*56 + 256*224 = 64512 = 384*168)
  // (the role code is nested inside the two branches so that ptxas allocates each region with its own budget)
  if (warp < 4) {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 56;");
    if (warp == 0) {
      // =========================== TMA producer (event loop: K/V ring, Q of stream 0, Q of stream 1) =============
      if (elect_one_sync()) {
        int kv_g = 0, kv_kt = 0, kv_stage = 0;
        uint32_t kv_phase = 0;
        int q_g[2] = {0, 0}, q_j[2] = {0, 0};
        uint32_t q_n[2] = {0, 0};
        for (int s = 0; s < 2; ++s)
          while (q_g[s] < sc.ngroups && sc.ntiles(s, q_g[s]) == 0) ++q_g[s];
        while (kv_g < sc.ngroups || q_g[0] < sc.ngroups || q_g[1] < sc.ngroups) {
          bool progress = false;
          if (kv_g < sc.ngroups && mbar_test(&bars->kv_empty[kv_stage], kv_phase ^ 1)) {
            progress = true;
            const int u = blockIdx.x + sc.unit_ord(kv_g) * gridDim.x;
            const int b = u / p.heads, h = u - b * p.heads;
            uint64_t* full = &bars->kv_full[kv_stage];
            uint8_t* st = sKV + kv_stage * Cfg::kStageBytes;
            mbar_arrive_expect_tx(full, Cfg::kStageBytes);
            tma_load_3d(st, &maps.kv_main, full, p.Dp + h * HD, kv_kt * BKV, b);
            tma_load_3d(st + Cfg::kKMain, &maps.kv_main, full, 2 * p.Dp + h * HD, kv_kt * BKV, b);
            if (HDX) {
              tma_load_3d(st + 2 * Cfg::kKMain, &maps.kv_x, full, p.Dp + h * HD + 64, kv_kt * BKV, b);
              tma_load_3d(st + 2 * Cfg::kKMain + Cfg::kKX, &maps.kv_x, full, 2 * p.Dp + h * HD + 64, kv_kt * BKV, b);
            }
            if (++kv_stage == NS) kv_stage = 0, kv_phase ^= 1u;
            if (++kv_kt == nkv) kv_kt = 0, ++kv_g;
          }
#pragma unroll
          for (int s = 0; s < 2; ++s) {
            if (q_g[s] < sc.ngroups && mbar_test(&bars->q_empty[s], (q_n[s] & 1) ^ 1)) {
              progress = true;
              const int g = q_g[s];
              const int u = blockIdx.x + sc.unit_ord(g) * gridDim.x;
              const int b = u / p.heads, h = u - b * p.heads;
              const int qt = sc.first_qt(s, g) + 2 * q_j[s];
              mbar_arrive_expect_tx(&bars->q_full[s], Cfg::kQMain + Cfg::kQX);
              tma_load_3d(sQ + s * Cfg::kQMain, &maps.q_main, &bars->q_full[s], h * HD, qt * kBQ, b);
              if (HDX) tma_load_3d(sQX + s * Cfg::kQX, &maps.q_x, &bars->q_full[s], h * HD + 64, qt * kBQ, b);
              ++q_n[s];
              if (++q_j[s] == sc.ntiles(s, g)) {
                q_j[s] = 0;
                ++q_g[s];
                while (q_g[s] < sc.ngroups && sc.ntiles(s, q_g[s]) == 0) ++q_g[s];
              }
            }
          }
          if (!progress) __nanosleep(kPollSleepNs);  // leave the issue slots of this scheduler to the softmax warp
        }
      }
    } else if (warp == 1 || warp == 3) {
      // =========================== MMA issuers: warp 1 serves stream 0, warp 3 stream 1 ===========================
      // (one thread polling the barriers of both streams reacted to each event in ~500-900 clk and was the pacing
      // resource of the whole kernel; per-stream issuers halve the polling loop and never queue behind each other)
      const int s_lo = warp == 1 ? 0 : 1;
      if (elect_one_sync()) {
        constexpr uint32_t idesc_s = umma_idesc_bf16(kBQ, BKV, 0, 0);
        constexpr uint32_t idesc_pv = umma_idesc_bf16(kBQ, 64, 0, 1);   // B (= V) is MN-major
        constexpr uint32_t idesc_pvx = umma_idesc_bf16(kBQ, 16, 0, 1);  // the 16 extra head columns
        TRACE_DECL
        // per-stream cursor: group, tile within group, next S / next PV kv step, and barrier use counts
        const int s = s_lo;
        int cg = 0, cj = 0, ks = 0, kp = 0;
        uint32_t n_tile = 0, n_s = 0, n_pv = 0;
        int cnt = sc.ngroups ? sc.ntiles(s, 0) : 0;
        auto slot_of = [&](int g, int kt, int& stage, uint32_t& phase) {
          const int seq = g * nkv + kt;
          const int r = seq / NS;
          stage = seq - r * NS;
          phase = (uint32_t)r & 1u;
        };
        while (cg < sc.ngroups) {
          bool progress = false;
          {
            int stage;
            uint32_t phase;
            if (cnt == 0) {
              // this stream has no query tile in the group: only hand its share of the K/V slots back
              slot_of(cg, kp, stage, phase);
              if (mbar_test(&bars->kv_full[stage], phase)) {
                mbar_arrive(&bars->kv_empty[stage]);
                progress = true;
                if (++kp == nkv) {
                  kp = 0;
                  if (++cg < sc.ngroups) cnt = sc.ntiles(s, cg);
                }
              }
              continue;
            }
            // ---- S(ks) = Q K^T : needs Q (first step), the K/V slot, and the S accumulator back from the softmax
            if (ks < nkv) {
              slot_of(cg, ks, stage, phase);
              if ((ks != 0 || mbar_test(&bars->q_full[s], n_tile & 1)) &&
                  mbar_test(&bars->kv_full[stage], phase) && mbar_test(&bars->s_empty[s], (n_s & 1) ^ 1)) {
                tc_fence_after();
                progress = true;
                const uint32_t q_addr = smem_u32(sQ + s * Cfg::kQMain);
                const uint32_t k_addr = smem_u32(sKV + stage * Cfg::kStageBytes);
                const uint32_t d = tmem_base + Cfg::kColS + s * BKV;
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                  tc_mma_f16_ss(d, umma_desc_sw128(q_addr + k * 32, 16, 1024),
                                umma_desc_sw128(k_addr + k * 32, 16, 1024), idesc_s, k != 0);
                }
                if (HDX) {
                  const uint32_t qx = smem_u32(sQX + s * Cfg::kQX);
                  const uint32_t kx = smem_u32(sKV + stage * Cfg::kStageBytes + 2 * Cfg::kKMain);
                  tc_mma_f16_ss(d, umma_desc_sw32(qx, 16, 256), umma_desc_sw32(kx, 16, 256), idesc_s, 1);
                }
                tc_commit(&bars->s_full[s]);
                TRACE_MMA(6 + s);
                ++n_s;
                if (++ks == nkv) tc_commit(&bars->q_empty[s]);  // Q tile free once these MMAs retire
              }
            }
            // ---- O += P(kp) V : needs P from the softmax (and, for the first step, the drained O accumulator)
            if (kp < ks) {
              if (mbar_test(&bars->p_full[s], n_pv & 1) &&
                  (kp != 0 || mbar_test(&bars->o_empty[s], (n_tile & 1) ^ 1))) {
                tc_fence_after();
                progress = true;
                slot_of(cg, kp, stage, phase);
                const uint32_t v_addr = smem_u32(sKV + stage * Cfg::kStageBytes + Cfg::kKMain);
                const uint32_t vx_addr = smem_u32(sKV + stage * Cfg::kStageBytes + 2 * Cfg::kKMain + Cfg::kKX);
                const uint32_t d = tmem_base + Cfg::kColO + s * HD;
                const uint32_t a = tmem_base + Cfg::kColP + s * (BKV / 2);
                const uint32_t acc0 = kp != 0;
#pragma unroll
                for (int k = 0; k < BKV / 16; ++k) {
                  // A = P from TMEM (16 bf16 = 8 columns per k-step); B = V MN-major, 16 kv rows per k-step
                  tc_mma_f16_ts(d, a + k * 8, umma_desc_sw128(v_addr + k * 2048, 1024, 1024), idesc_pv, acc0 | (k != 0));
                  if (HDX)
                    tc_mma_f16_ts(d + 64, a + k * 8, umma_desc_sw32(vx_addr + k * 512, 256, 256), idesc_pvx,
                                  acc0 | (k != 0));
                }
                tc_commit(&bars->pv_done[s]);
                TRACE_MMA(8 + s);
                ++n_pv;
                if (cj == cnt - 1) tc_commit(&bars->kv_empty[stage]);  // last tile of this stream on the slot
                if (++kp == nkv) {  // tile finished: next tile / next group
                  ks = kp = 0;
                  ++n_tile;
                  if (++cj == cnt) {
                    cj = 0;
                    if (++cg < sc.ngroups) cnt = sc.ntiles(s, cg);
                  }
                }
              }
            }
          }
          
        }
      }
    }
  } else {
    asm volatile("setmaxnreg.inc.sync.aligned.u32 224;");
    // =========================== softmax / correction / epilogue (stream = warpgroup) ===========================
    // The two streams take turns in the MUFU-bound exp phase (named barriers 1 and 2 pass a token back and forth),
    // which keeps them half a step out of phase: while one stream runs its exps the other does its TMEM loads,
    // row max, P store and barrier traffic.  A stream without a tile in a slot just passes the token on.
    constexpr int NCH = BKV / 32;
    const int s = (warp - 4) >> 2;
    const int q = warp & 3;
    const int row = q * 32 + lane;  // query row within the tile == TMEM lane
    const uint32_t lane_sel = (uint32_t)(q * 32) << 16;
    const uint32_t t_S = tmem_base + Cfg::kColS + s * BKV + lane_sel;
    const uint32_t t_P = tmem_base + Cfg::kColP + s * (BKV / 2) + lane_sel;
    const uint32_t t_O = tmem_base + Cfg::kColO + s * HD + lane_sel;
    const float sl2 = p.scale_log2;
    uint32_t it = 0;
    TRACE_DECL
    // one token per scheduler (warp q of stream 0 and warp q of stream 1 share scheduler q): barriers 1 + 2q + s
    const int tok_mine = 1 + 2 * q + s, tok_other = 1 + 2 * q + (s ^ 1);
#if MM_ATTN_TOKEN
    if (s == 1) named_bar_arrive(tok_other, 64);  // stream 0 goes first
#else
    (void)tok_mine, (void)tok_other;
#endif
    for (int g = 0; g < sc.ngroups; ++g) {
      const int u = blockIdx.x + sc.unit_ord(g) * gridDim.x;
      const int b = u / p.heads, h = u - b * p.heads;
      const int nt = sc.ntiles(s, g);
      for (int j = 0; j < sc.slots(g); ++j) {
        if (j >= nt) {  // no tile in this slot: keep the token moving
#if MM_ATTN_TOKEN
          for (int kt = 0; kt < nkv; ++kt) {
            named_bar_sync(tok_mine, 64);
            named_bar_arrive(tok_other, 64);
          }
#endif
          continue;
        }
        const int qt = sc.first_qt(s, g) + 2 * j;
        const bool warp_live = qt * kBQ + q * 32 < p.T;  // warps whose 32 rows are all padding only keep the protocol
        float m_run = -INFINITY, l_run = 0.f;            // running max (raw score units) and row sum
        for (int kt = 0; kt < nkv; ++kt, ++it) {
          const int valid = p.T - kt * BKV;  // columns >= valid are padding (only in the last kv tile)
          sm_wait(&bars->s_full[s], it & 1);
          TRACE_SM(0);
          tc_fence_after();
          uint32_t sv[BKV];
          if (warp_live) {
#pragma unroll
            for (int c = 0; c < NCH; ++c)
              if (c * 32 < valid) tmem_ld_32x32b_x32(t_S + c * 32, sv + c * 32);
            tc_wait_ld();
            TRACE_SM(1);
          }
          // the S accumulator now lives in registers: hand it back for the next QK^T right away
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&bars->s_empty[s]);
          uint32_t pk[BKV / 2];
          float m_new = m_run, sum = 0.f;
          float m_scaled = 0.f;
          if (warp_live) {
            if (valid < BKV) {
#pragma unroll
              for (int jj = 0; jj < BKV; ++jj)
                if (jj >= valid) sv[jj] = 0xff800000u;  // -inf
            }
            float mx0 = -INFINITY, mx1 = -INFINITY, mx2 = -INFINITY, mx3 = -INFINITY;
#pragma unroll
            for (int jj = 0; jj < BKV; jj += 8) {
              mx0 = fmaxf(mx0, fmaxf(__uint_as_float(sv[jj + 0]), __uint_as_float(sv[jj + 1])));
              mx1 = fmaxf(mx1, fmaxf(__uint_as_float(sv[jj + 2]), __uint_as_float(sv[jj + 3])));
              mx2 = fmaxf(mx2, fmaxf(__uint_as_float(sv[jj + 4]), __uint_as_float(sv[jj + 5])));
              mx3 = fmaxf(mx3, fmaxf(__uint_as_float(sv[jj + 6]), __uint_as_float(sv[jj + 7])));
            }
            const float mx = fmaxf(fmaxf(mx0, mx1), fmaxf(mx2, mx3));
            // the reference max only moves when it grows by more than the threshold (keeps O rescales rare)
            if ((mx - m_run) * sl2 > kRescaleThreshold) m_new = mx;  // also true for m_run = -inf
            m_scaled = m_new * sl2;
          }
          TRACE_SM(14);
#if MM_ATTN_TOKEN
          named_bar_sync(tok_mine, 64);  // ---- exp phase: this warp's turn on its scheduler's MUFU ----
#endif
          TRACE_SM(15);
          if (warp_live) {
            ExpAcc ea;
            exp_acc_reset(ea);
            if (valid >= BKV) {
              // full tile: one straight-line block, so that ptxas interleaves the ex2 of the whole row (a block per
              // 32-column chunk drains the MUFU pipe at every boundary: measured 1900 vs 1250 clk per tile)
#pragma unroll
              for (int jj = 0; jj < BKV; jj += 4) exp4(sv + jj, pk + (jj >> 1), sl2, m_scaled, ea);
            } else {
              // last (partial) kv tile: chunks that are all padding are skipped
#pragma unroll
              for (int c = 0; c < NCH; ++c) {
                if (c * 32 < valid) {
#pragma unroll
                  for (int jj = c * 32; jj < c * 32 + 32; jj += 4) exp4(sv + jj, pk + (jj >> 1), sl2, m_scaled, ea);
                } else {
#pragma unroll
                  for (int jj = 0; jj < 16; ++jj) pk[c * 16 + jj] = 0u;
                }
              }
            }
            sum = exp_acc_total(ea);
          }
          TRACE_SM(2);
#if MM_ATTN_TOKEN
          named_bar_arrive(tok_other, 64);  // ---- token to the other stream's warp on this scheduler ----
#endif
          // P (TMEM) and the O accumulator are free once PV of the previous iteration retired
          if (it > 0) {
            sm_wait(&bars->pv_done[s], (it - 1) & 1);
            tc_fence_after();
          }
          TRACE_SM(3);
          if (warp_live) {
            if (kt > 0) {
              const float alpha = ex2((m_run - m_new) * sl2);  // 1 when the max did not move
              l_run = l_run * alpha + sum;
              if (__any_sync(0xffffffffu, m_new != m_run)) {
#pragma unroll 1
                for (int c = 0; c < HD / 16; ++c) {
                  uint32_t o[16];
                  tmem_ld_32x32b_x16(t_O + c * 16, o);
                  tc_wait_ld();
#pragma unroll
                  for (int jj = 0; jj < 16; ++jj) o[jj] = __float_as_uint(__uint_as_float(o[jj]) * alpha);
                  tmem_st_32x32b_x16(t_O + c * 16, o);
                }
              }
            } else {
              l_run = sum;
            }
            m_run = m_new;
#pragma unroll
            for (int c = 0; c < BKV / 64; ++c) tmem_st_32x32b_x32(t_P + c * 32, pk + c * 32);
            if (BKV % 64) tmem_st_32x32b_x16(t_P + (BKV / 64) * 32, pk + (BKV / 64) * 32);
            tc_wait_st();
            TRACE_SM(4);
          }
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&bars->p_full[s]);
          TRACE_SM(5);
        }
        // ---- epilogue of the q tile: O / l -> bf16 -> global ----
        sm_wait(&bars->pv_done[s], (it - 1) & 1);
        tc_fence_after();
        if (warp_live) {
          const float inv_l = 1.0f / l_run;
          const int t = qt * kBQ + row;
          __nv_bfloat16* dst = p.out + ((int64_t)b * p.T + t) * p.ld_out + h * p.hd_out;
#pragma unroll
          for (int c = 0; c < HD / 16; ++c) {
            uint32_t o[16];
            tmem_ld_32x32b_x16(t_O + c * 16, o);
            tc_wait_ld();
            if (t < p.T) {
#pragma unroll
              for (int jj = 0; jj < 2; ++jj) {
                if (c * 16 + jj * 8 < p.hd_out) {
                  int4 v;
                  v.x = (int)pack_bf16x2(__uint_as_float(o[jj * 8 + 0]) * inv_l, __uint_as_float(o[jj * 8 + 1]) * inv_l);
                  v.y = (int)pack_bf16x2(__uint_as_float(o[jj * 8 + 2]) * inv_l, __uint_as_float(o[jj * 8 + 3]) * inv_l);
                  v.z = (int)pack_bf16x2(__uint_as_float(o[jj * 8 + 4]) * inv_l, __uint_as_float(o[jj * 8 + 5]) * inv_l);
                  v.w = (int)pack_bf16x2(__uint_as_float(o[jj * 8 + 6]) * inv_l, __uint_as_float(o[jj * 8 + 7]) * inv_l);
                  *reinterpret_cast<int4*>(dst + c * 16 + jj * 8) = v;
                }
              }
            }
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&bars->o_empty[s]);
      }
    }
#if MM_ATTN_TOKEN
    if (s == 0) named_bar_sync(tok_mine, 64);  // absorb stream 1's last token pass
#endif
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

}  // namespace mm

namespace mm {
int make_tmap_3d(CUtensorMap* tm, const void* ptr, int64_t d0, int64_t d1, int64_t d2, int64_t ld1, int64_t ld2,
                 int box_rows);
int make_tmap_3d_sw32(CUtensorMap* tm, const void* ptr, int64_t d0, int64_t d1, int64_t d2, int64_t ld1, int64_t ld2,
                      int box_rows);

template <int HD, int BKV>
static int launch_attention(const void* qkv, const AttnParams& p, cudaStream_t st) {
  using Cfg = AttnCfg<HD, BKV>;
  AttnMaps maps;
  const int64_t W = 3 * (int64_t)p.Dp;
  int rc = make_tmap_3d(&maps.q_main, qkv, W, p.T, p.B, W, (int64_t)p.T * W, kBQ);
  if (rc != MM_OK) return rc;
  rc = make_tmap_3d(&maps.kv_main, qkv, W, p.T, p.B, W, (int64_t)p.T * W, BKV);
  if (rc != MM_OK) return rc;
  if (Cfg::kHDX) {
    rc = make_tmap_3d_sw32(&maps.q_x, qkv, W, p.T, p.B, W, (int64_t)p.T * W, kBQ);
    if (rc != MM_OK) return rc;
    rc = make_tmap_3d_sw32(&maps.kv_x, qkv, W, p.T, p.B, W, (int64_t)p.T * W, BKV);
    if (rc != MM_OK) return rc;
  } else {
    maps.q_x = maps.q_main;
    maps.kv_x = maps.kv_main;
  }
  static std::atomic<bool> attr_set[64];
  int dev = 0;
  MM_CUDA_CHECK(cudaGetDevice(&dev));
  if (dev >= 0 && dev < 64 && !attr_set[dev].load(std::memory_order_acquire)) {
    MM_CUDA_CHECK(cudaFuncSetAttribute(attention_fwd_kernel<HD, BKV>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       Cfg::kSmemBytes));
    attr_set[dev].store(true, std::memory_order_release);
  }
  const int units = p.B * p.heads;
  const int sms = sm_count();
  const int grid = units < sms ? units : sms;
  attention_fwd_kernel<HD, BKV><<<grid, kAttnThreads, Cfg::kSmemBytes, st>>>(maps, p);
  MM_CUDA_CHECK(cudaGetLastError());
  return MM_OK;
}
}  // namespace mm

namespace mm {
int launch_attention3(const void* qkv, void* out, int B, int T, int heads, float scale, cudaStream_t st);
}
#ifndef MM_ATTN_V3
#define MM_ATTN_V3 1
#endif
using namespace mm;

#ifdef MM_ATTN_TRACE
namespace mm { int attention3_trace_set(long long* buf); }
extern "C" int mm_attention_trace_set(long long* buf) {
  if (cudaMemcpyToSymbol(g_attn_trace, &buf, sizeof(buf)) != cudaSuccess) return MM_ERR_CUDA;
  return MM_OK;
}
extern "C" int mm_attention3_trace_set(long long* buf) { return mm::attention3_trace_set(buf); }
#endif

extern "C" int64_t mm_attention_head_stride(int64_t head_dim) {
  if (head_dim == 64) return 64;
  if (head_dim == 72) return 80;
  return -1;
}

extern "C" int mm_attention_fwd(const void* qkv, void* out, int64_t B, int64_t T, int64_t heads, int64_t head_dim,
                                float scale, void* stream) {
  if (B < 0 || T <= 0 || heads <= 0) return MM_ERR_BAD_ARG;
  const int64_t hs = mm_attention_head_stride(head_dim);
  if (hs < 0 || T > 1024) return MM_ERR_BAD_ARG;
  if (B == 0) return MM_OK;
  if (!qkv || !out || (reinterpret_cast<uintptr_t>(qkv) & 15) || (reinterpret_cast<uintptr_t>(out) & 15))
    return MM_ERR_BAD_ARG;
  if (B * heads > INT32_MAX / 2 || B * T * 3 * heads * hs > ((int64_t)1 << 40)) return MM_ERR_BAD_ARG;
  if (!(scale > 0.f)) return MM_ERR_BAD_ARG;  // the online max runs in raw-score units (needs scale > 0)
  AttnParams p;
  p.out = reinterpret_cast<__nv_bfloat16*>(out);
  p.B = (int)B, p.T = (int)T, p.heads = (int)heads;
  p.Dp = (int)(heads * hs);
  p.hd_out = (int)head_dim;
  p.ld_out = (int)(heads * head_dim);
  p.scale_log2 = scale * 1.4426950408889634f;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
#if MM_ATTN_V3
  // head_dim 64: three query streams on 64-column kv tiles (attention3_sm100.cu)
  if (hs == 64) return launch_attention3(qkv, out, (int)B, (int)T, (int)heads, scale, st);
#endif
  if (hs == 64) return launch_attention<64, 128>(qkv, p, st);
  return launch_attention<80, 96>(qkv, p, st);
}
