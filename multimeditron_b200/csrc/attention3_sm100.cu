// attention3_sm100.cu — non-causal multi-head attention forward, head_dim 64, NSTR query streams on 64-column kv tiles.
//
// Why this shape (measured on B200, tools/micro/tmem_probe2.cu, profiles/r02d_tmem_probe2.txt):
//   * one softmax step of a stream is a serial chain  S ready -> TMEM load -> row max -> exps -> P store -> PV ; its
//     fixed part (three mbarrier hops of ~170 clk, LDTM + max ~200-370 clk, STTM ~75 clk) does not shrink with the
//     tile, and a lone warp per scheduler only sustains 10.5 clk per ex2 (8.7 with two warps interleaved);
//   * a tcgen05.mma costs >= ~65 clk whatever its N (N=64 and N=128 alike), so tensor time is per INSTRUCTION.
// The two-stream 128-column kernel (attention_sm100.cu) leaves the MUFU pipe idle 59 % of the time: two streams
// cannot cover each other's fixed part, and with T = 577 (five 128-row query tiles) the third tile slot of every
// unit runs ONE stream.  Here three streams (one softmax warpgroup each, 512 threads) walk three query tiles of the
// unit at a time over 64-column kv tiles, free-running (no token): while one warp of a scheduler sits in a hop or
// a TMEM round trip the other two keep the MUFU pipe fed, and five query tiles take two slots (3 + 2) instead of
// three (2 + 2 + 1).
//   S = Q·Kᵀ      tcgen05.mma 128 x 64 x 16 (SS, K-major SWIZZLE_128B operands), fp32 in TMEM
//   P = softmax   one query row per thread: the 64-column S row is pulled into registers (frees S for the next
//                 QKᵀ at once), online max / sum in the exp2 domain with lazy rescale (threshold 2^8), P back to
//                 TMEM as packed bf16
//   O += P·V      tcgen05.mma 128 x 64 x 16 (TS: A = P from TMEM, B = V MN-major straight from its TMA tile)
// TMEM: S0 S1 S2 | P0 P1 P2 | O0 O1 O2 = 3 x (64 + 32 + 64) = 480 columns.
// K/V tiles (64 rows, 16 KB per stage) live in an 11-stage ring: resident for T <= 704 (loaded once per unit, reused
// by every query tile), streamed once per query triple otherwise.
//
// Replaces HF:models/clip/modeling_clip.py:318-331 (SDPA; scale = head_dim**-0.5, no mask, dropout 0).
#include "common.cuh"

namespace mm {
namespace a3 {

constexpr int kBQ = 128;
constexpr float kRescaleThreshold = 8.0f;  // log2 units: P <= 2^8, harmless in bf16/fp32
constexpr int kMaxStages = 12;
#ifndef MM_ATTN3_POLL_NS
#define MM_ATTN3_POLL_NS 32
#endif

// Optional phase trace (tools/attn3_trace.py; only with -DMM_ATTN_TRACE): block 0 records clock64() stamps of warp
// q = 0 of every softmax stream (roles 0..2) and of every MMA issuer (roles 3..5).
#ifdef MM_ATTN_TRACE
__device__ long long* g_trace = nullptr;  // 6 roles x 512 (id, clock) pairs
__device__ __forceinline__ void trace(int role, uint32_t& k, int id) {
  if (g_trace && k < 512) {
    g_trace[(role * 512 + k) * 2] = id;
    g_trace[(role * 512 + k) * 2 + 1] = clock64();
    ++k;
  }
}
#define TR_DECL uint32_t trc_k = 0;
#define TR_SM(id) do { if (blockIdx.x == 0 && q == 0 && lane == 0) trace(s, trc_k, id); } while (0)
#define TR_MMA(id) do { if (blockIdx.x == 0) trace(3 + s, trc_k, id); } while (0)
#else
#define TR_DECL
#define TR_SM(id)
#define TR_MMA(id)
#endif

template <int BKV, int NSTR>
struct Cfg {
  static constexpr int HD = 64;
  static_assert(BKV == 64 || BKV == 128, "kv tile");
  static_assert(NSTR == 2 || NSTR == 3, "streams");
  static constexpr int kThreads = 128 + NSTR * 128;   // w0 TMA (+TMEM alloc), w1..w3 MMA issuer of stream 0..2, then softmax
  static constexpr int kQTile = kBQ * 128;             // bytes: [128 x 64] bf16
  static constexpr int kKTile = BKV * 128;
  static constexpr int kStageBytes = 2 * kKTile;       // K + V
  static constexpr int kQBytes = NSTR * kQTile;
  static constexpr int kBarBytes = 1024;
  static constexpr int kStagesRaw = (232448 - 1024 - kBarBytes - kQBytes) / kStageBytes;
  static constexpr int kStages = kStagesRaw > kMaxStages ? kMaxStages : kStagesRaw;
  static constexpr int kSmemBytes = 1024 + kQBytes + kStages * kStageBytes + kBarBytes;
  static constexpr int kColS = 0;                      // + s*BKV
  static constexpr int kColP = NSTR * BKV;             // + s*BKV/2
  static constexpr int kColO = NSTR * BKV * 3 / 2;     // + s*HD
  static_assert(NSTR * (BKV * 3 / 2 + HD) <= 512, "TMEM budget");
  static constexpr int kRegsService = 56;
  // (threads * launch regs) = 128*56 + NSTR*128*softmax regs
  static constexpr int kRegsLaunch = (65536 / kThreads) / 8 * 8;
  static constexpr int kRegsSoftmax = ((kThreads * kRegsLaunch - 128 * kRegsService) / (NSTR * 128)) / 8 * 8;
};

// MM_ATTN3_PERMITS: how many of a scheduler's NSTR softmax warps may be in the exp phase at once (0 = no limit).
// Free-running streams fall into lockstep (all three load, exponentiate and store at the same time, profiles/
// r02h_attn3_trace.txt), which leaves the MUFU pipe idle during everybody's TMEM round trips; a counting semaphore per
// scheduler keeps one stream out of phase with the other two.
#ifndef MM_ATTN3_PERMITS
#define MM_ATTN3_PERMITS 0
#endif
// MM_ATTN3_ISSUER_SLEEP: nanosleep of an MMA issuer whose poll found nothing to do (0 = spin).  The three issuers sit on
// the schedulers of the softmax warps; ncu (profiles/r02j) shows a third of all issued instructions are their polls.
#ifndef MM_ATTN3_ISSUER_SLEEP
#define MM_ATTN3_ISSUER_SLEEP 0
#endif
// MM_ATTN3_PACKED: scale-and-shift and row-sum adds as fp32x2 instructions (FFMA2 / FADD2)
// MM_ATTN3_POLY: share of the exponentials computed on the FMA pipe instead of the MUFU pipe (FlashAttention-4's trick):
// 2^x = 2^round(x) * p(x - round(x)), round by the magic-number add, p = degree-3 minimax polynomial of 2^f on
// [-0.5, 0.5] (relative error 7.7e-5, far below the bf16 rounding of P), the power of two added into the exponent field.
// 0 = none, 1 = two of every eight scores (25 %), 2 = two of every four (50 %).  The exp phase of the three lock-stepped
// softmax warps is MUFU-pipe bound (8.7 clk per ex2) while the FMA pipe idles at 18 %.
#ifndef MM_ATTN3_POLY
#define MM_ATTN3_POLY 0
#endif
// MM_ATTN3_INDEP: T <= 128 -> every stream works on its own (image, head) unit (see Sched)
#ifndef MM_ATTN3_INDEP
#define MM_ATTN3_INDEP 1
#endif
#ifndef MM_ATTN3_PACKED
#define MM_ATTN3_PACKED 1
#endif

struct Bars {
  int sem[4];
  uint64_t kv_full[kMaxStages], kv_empty[kMaxStages];
  uint64_t q_full[3], q_empty[3];
  uint64_t s_full[3], s_empty[3];
  uint64_t p_full[3], pv_done[3];
  uint64_t o_empty[3];
  uint32_t tmem_base;
};
static_assert(sizeof(Bars) <= 1024, "bars");

struct Maps {
  CUtensorMap q, kv;
};

struct Params {
  __nv_bfloat16* out;  // [B*T, ld_out]
  int B, T, heads;
  int Dp;              // heads * 64 : width of each of the q | k | v sections of a qkv row
  int ld_out;
  float scale_log2;    // scale * log2(e)
};

__device__ __forceinline__ float ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
// two exponentials on the FMA pipe (packed fp32x2): x in [-126, ~16]
__device__ __forceinline__ void ex2_poly2(uint64_t xv, float& pa, float& pb) {
  float xa, xb;
  f32x2_unpack(xv, xa, xb);
  xa = fmaxf(xa, -126.f), xb = fmaxf(xb, -126.f);                       // masked (-inf) and far-below-max scores
  const uint64_t x = f32x2(xa, xb);
  const uint64_t xr = add_f32x2(x, f32x2(12582912.f, 12582912.f));      // 1.5 * 2^23: round(x) lands in the low mantissa bits
  const uint64_t n = add_f32x2(xr, f32x2(-12582912.f, -12582912.f));
  const uint64_t f = fma_f32x2(n, f32x2(-1.f, -1.f), x);                // x - round(x) in [-0.5, 0.5]
  uint64_t pl = fma_f32x2(f, f32x2(0.05508868396282196f, 0.05508868396282196f), f32x2(0.24260404706001282f, 0.24260404706001282f));
  pl = fma_f32x2(pl, f, f32x2(0.6932762265205383f, 0.6932762265205383f));
  pl = fma_f32x2(pl, f, f32x2(0.9999289512634277f, 0.9999289512634277f));
  float ra, rb, qa, qb;
  f32x2_unpack(xr, ra, rb);
  f32x2_unpack(pl, qa, qb);
  pa = __int_as_float(__float_as_int(qa) + (__float_as_int(ra) << 23));
  pb = __int_as_float(__float_as_int(qb) + (__float_as_int(rb) << 23));
}

__device__ __forceinline__ void mma_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t"
      "}\n" ::"r"(d_tmem),
      "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ bool bar_test(uint64_t* bar, uint32_t parity) {
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

// Work decomposition shared by every role.  The CTA owns units u_i = blockIdx.x + i*gridDim.x.  A K/V "group" is
// what travels through the ring together: the whole unit when it fits (resident), else one query NSTR-tuple of it.
// Stream s owns, in group g, the query tiles first_qt(s,g) + NSTR*j, j < ntiles(s,g).
// T <= 128 (one query tile per unit: CLIP ViT-B/32, T = 50): the streams would share nothing but idle time, so each
// stream takes its OWN unit ("indep": group g = units NSTR*g .. NSTR*g + NSTR-1 of this CTA, stream s owns unit
// NSTR*g + s and its private K/V slots) — three units in flight per SM instead of one.
template <int NSTR>
struct Sched {
  int nkv, nq, ntup, n_units, ngroups;
  bool resident, indep;
  __device__ __forceinline__ int unit_ord(int g) const { return resident ? g : g / ntup; }
  // ordinal (in this CTA's unit list) of the unit stream s works on in group g
  __device__ __forceinline__ int unit_of(int s, int g) const { return indep ? NSTR * g + s : unit_ord(g); }
  // ring sequence number of K/V tile kt of the unit stream s works on in group g
  __device__ __forceinline__ int kv_seq(int s, int g, int kt) const { return (indep ? NSTR * g + s : g) * nkv + kt; }
  __device__ __forceinline__ int first_qt(int s, int g) const {
    if (indep) return 0;
    return resident ? s : NSTR * (g - unit_ord(g) * ntup) + s;
  }
  __device__ __forceinline__ int ntiles(int s, int g) const {
    if (indep) return NSTR * g + s < n_units ? 1 : 0;
    const int f = first_qt(s, g);
    if (f >= nq) return 0;
    return resident ? (nq - f + NSTR - 1) / NSTR : 1;
  }
};

template <int BKV, int NSTR>
__global__ void __launch_bounds__(Cfg<BKV, NSTR>::kThreads, 1)
    attention3_fwd_kernel(const __grid_constant__ Maps maps, const Params p) {
  using C = Cfg<BKV, NSTR>;
  constexpr int NS = C::kStages;
  constexpr int HD = 64;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sQ = smem;                                  // [Q0][Q1][Q2]
  uint8_t* sKV = sQ + C::kQBytes;                      // per stage: [K][V]
  Bars* bars = reinterpret_cast<Bars*>(sKV + NS * C::kStageBytes);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int units = p.B * p.heads;
  Sched<NSTR> sc;
  sc.nkv = (p.T + BKV - 1) / BKV;
  sc.nq = (p.T + kBQ - 1) / kBQ;
  sc.ntup = (sc.nq + NSTR - 1) / NSTR;
  sc.resident = sc.nkv <= NS;  // whole unit fits in the ring: load K/V once per unit
  sc.indep = MM_ATTN3_INDEP && sc.nq == 1 && NSTR * sc.nkv <= NS;
  sc.n_units = ((int)blockIdx.x < units) ? (units - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;
  sc.ngroups = sc.indep ? (sc.n_units + NSTR - 1) / NSTR : sc.n_units * (sc.resident ? 1 : sc.ntup);
  const int kv_items = sc.indep ? sc.n_units : sc.ngroups;   // what the K/V producer walks: units (indep) or groups
  const int nkv = sc.nkv;

  if (warp == 1 && lane == 0) {
    tma_prefetch_desc(&maps.q);
    tma_prefetch_desc(&maps.kv);
    for (int i = 0; i < NS; ++i) {
      mbar_init(&bars->kv_full[i], 1);
      // one release per stream that sees the slot (real tile: tcgen05.commit; none: plain arrive); private slots: one
      mbar_init(&bars->kv_empty[i], sc.indep ? 1 : NSTR);
    }
    for (int s = 0; s < NSTR; ++s) {
      mbar_init(&bars->q_full[s], 1);
      mbar_init(&bars->q_empty[s], 1);
      mbar_init(&bars->s_full[s], 1);
      mbar_init(&bars->s_empty[s], 4);
      mbar_init(&bars->p_full[s], 4);
      mbar_init(&bars->pv_done[s], 1);
      mbar_init(&bars->o_empty[s], 4);
    }
    for (int i = 0; i < 4; ++i) bars->sem[i] = MM_ATTN3_PERMITS;
    fence_barrier_init();
  }
  if (warp == 0) {
    tmem_alloc(&bars->tmem_base, 512);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = bars->tmem_base;

  if (warp < 4) {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(C::kRegsService));
    if (warp == 0) {
      // =========================== TMA producer (event loop: K/V ring, Q of every stream) ======================
      if (elect_one_sync()) {
        int kv_g = 0, kv_kt = 0, kv_stage = 0;
        uint32_t kv_phase = 0;
        int q_g[NSTR], q_j[NSTR];
        uint32_t q_n[NSTR];
#pragma unroll
        for (int s = 0; s < NSTR; ++s) {
          q_g[s] = 0, q_j[s] = 0, q_n[s] = 0;
          while (q_g[s] < sc.ngroups && sc.ntiles(s, q_g[s]) == 0) ++q_g[s];
        }
        for (;;) {
          bool pending = kv_g < kv_items;
#pragma unroll
          for (int s = 0; s < NSTR; ++s) pending |= q_g[s] < sc.ngroups;
          if (!pending) break;
          bool progress = false;
          if (kv_g < kv_items && bar_test(&bars->kv_empty[kv_stage], kv_phase ^ 1)) {
            progress = true;
            const int u = blockIdx.x + (sc.indep ? kv_g : sc.unit_ord(kv_g)) * gridDim.x;
            const int b = u / p.heads, h = u - b * p.heads;
            uint64_t* full = &bars->kv_full[kv_stage];
            uint8_t* st = sKV + kv_stage * C::kStageBytes;
            mbar_arrive_expect_tx(full, C::kStageBytes);
            tma_load_3d(st, &maps.kv, full, p.Dp + h * HD, kv_kt * BKV, b);
            tma_load_3d(st + C::kKTile, &maps.kv, full, 2 * p.Dp + h * HD, kv_kt * BKV, b);
            if (++kv_stage == NS) kv_stage = 0, kv_phase ^= 1u;
            if (++kv_kt == nkv) kv_kt = 0, ++kv_g;
          }
#pragma unroll
          for (int s = 0; s < NSTR; ++s) {
            if (q_g[s] < sc.ngroups && bar_test(&bars->q_empty[s], (q_n[s] & 1) ^ 1)) {
              progress = true;
              const int g = q_g[s];
              const int u = blockIdx.x + sc.unit_of(s, g) * gridDim.x;
              const int b = u / p.heads, h = u - b * p.heads;
              const int qt = sc.first_qt(s, g) + NSTR * q_j[s];
              mbar_arrive_expect_tx(&bars->q_full[s], C::kQTile);
              tma_load_3d(sQ + s * C::kQTile, &maps.q, &bars->q_full[s], h * HD, qt * kBQ, b);
              ++q_n[s];
              if (++q_j[s] == sc.ntiles(s, g)) {
                q_j[s] = 0;
                ++q_g[s];
                while (q_g[s] < sc.ngroups && sc.ntiles(s, q_g[s]) == 0) ++q_g[s];
              }
            }
          }
          if (!progress) __nanosleep(MM_ATTN3_POLL_NS);
        }
      }
    } else if (warp - 1 < NSTR) {
      // =========================== MMA issuer of stream (warp - 1) ===============================================
      const int s = warp - 1;
      if (elect_one_sync()) {
        constexpr uint32_t idesc_s = umma_idesc_bf16(kBQ, BKV, 0, 0);
        constexpr uint32_t idesc_pv = umma_idesc_bf16(kBQ, 64, 0, 1);   // B (= V) is MN-major
        int cg = 0, cj = 0, ks = 0, kp = 0;
        uint32_t n_tile = 0, n_s = 0, n_pv = 0;
        TR_DECL
        int cnt = sc.ngroups ? sc.ntiles(s, 0) : 0;
        auto slot_of = [&](int g, int kt, int& stage, uint32_t& phase) {
          const int seq = sc.kv_seq(s, g, kt);
          const int r = seq / NS;
          stage = seq - r * NS;
          phase = (uint32_t)r & 1u;
        };
        const uint32_t q_addr = smem_u32(sQ + s * C::kQTile);
        const uint32_t d_s = tmem_base + C::kColS + s * BKV;
        const uint32_t d_o = tmem_base + C::kColO + s * HD;
        const uint32_t a_p = tmem_base + C::kColP + s * (BKV / 2);
        while (cg < sc.ngroups) {
          int stage;
          uint32_t phase;
          if (cnt == 0 && sc.indep) {   // private slots: nothing to hand back
            if (++cg < sc.ngroups) cnt = sc.ntiles(s, cg);
            continue;
          }
          if (cnt == 0) {
            // this stream has no query tile in the group: only hand its share of the K/V slots back
            slot_of(cg, kp, stage, phase);
            if (bar_test(&bars->kv_full[stage], phase)) {
              mbar_arrive(&bars->kv_empty[stage]);
              if (++kp == nkv) {
                kp = 0;
                if (++cg < sc.ngroups) cnt = sc.ntiles(s, cg);
              }
            } else {
              __nanosleep(MM_ATTN3_POLL_NS);
            }
            continue;
          }
          bool progress = false;
          // ---- S(ks) = Q K^T : needs Q (first step), the K/V slot, and the S accumulator back from the softmax
          if (ks < nkv) {
            slot_of(cg, ks, stage, phase);
            if ((ks != 0 || bar_test(&bars->q_full[s], n_tile & 1)) && bar_test(&bars->kv_full[stage], phase) &&
                bar_test(&bars->s_empty[s], (n_s & 1) ^ 1)) {
              progress = true;
              tc_fence_after();
              const uint32_t k_addr = smem_u32(sKV + stage * C::kStageBytes);
#pragma unroll
              for (int k = 0; k < 4; ++k)
                tc_mma_f16_ss(d_s, umma_desc_sw128(q_addr + k * 32, 16, 1024), umma_desc_sw128(k_addr + k * 32, 16, 1024),
                              idesc_s, k != 0);
              tc_commit(&bars->s_full[s]);
              TR_MMA(6);
              ++n_s;
              if (++ks == nkv) tc_commit(&bars->q_empty[s]);  // Q tile free once these MMAs retire
            }
          }
          // ---- O += P(kp) V : needs P from the softmax (and, for the first step, the drained O accumulator)
          if (kp < ks) {
            if (bar_test(&bars->p_full[s], n_pv & 1) && (kp != 0 || bar_test(&bars->o_empty[s], (n_tile & 1) ^ 1))) {
              progress = true;
              tc_fence_after();
              slot_of(cg, kp, stage, phase);
              const uint32_t v_addr = smem_u32(sKV + stage * C::kStageBytes + C::kKTile);
              const uint32_t acc0 = kp != 0;
#pragma unroll
              for (int k = 0; k < BKV / 16; ++k)   // A = P from TMEM (16 bf16 = 8 columns per k-step); B = V MN-major
                mma_ts(d_o, a_p + k * 8, umma_desc_sw128(v_addr + k * 2048, 1024, 1024), idesc_pv, acc0 | (k != 0));
              tc_commit(&bars->pv_done[s]);
              TR_MMA(8);
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
#if MM_ATTN3_ISSUER_SLEEP
          if (!progress) __nanosleep(MM_ATTN3_ISSUER_SLEEP);
#else
          (void)progress;
#endif
        }
      }
    }
  } else {
    asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(C::kRegsSoftmax));
    // =========================== softmax / correction / epilogue (stream = warpgroup) ===========================
    constexpr int NCH = BKV / 32;
    const int s = (warp - 4) >> 2;
    const int q = warp & 3;
    const int row = q * 32 + lane;  // query row within the tile == TMEM lane
    const uint32_t lane_sel = (uint32_t)(q * 32) << 16;
    const uint32_t t_S = tmem_base + C::kColS + s * BKV + lane_sel;
    const uint32_t t_P = tmem_base + C::kColP + s * (BKV / 2) + lane_sel;
    const uint32_t t_O = tmem_base + C::kColO + s * HD + lane_sel;
    const float sl2 = p.scale_log2;
    uint32_t it = 0;
    TR_DECL
    for (int g = 0; g < sc.ngroups; ++g) {
      const int u = blockIdx.x + sc.unit_of(s, g) * gridDim.x;
      const int b = u / p.heads, h = u - b * p.heads;
      const int nt = sc.ntiles(s, g);
      for (int j = 0; j < nt; ++j) {
        const int qt = sc.first_qt(s, g) + NSTR * j;
        const bool warp_live = qt * kBQ + q * 32 < p.T;  // warps whose 32 rows are all padding only keep the protocol
        float m_run = -INFINITY, l_run = 0.f;            // running max (raw score units) and row sum
        for (int kt = 0; kt < nkv; ++kt, ++it) {
          const int valid = p.T - kt * BKV;  // columns >= valid are padding (only in the last kv tile)
          mbar_wait(&bars->s_full[s], it & 1);
          TR_SM(0);
          tc_fence_after();
          uint32_t sv[BKV];
          if (warp_live) {
#pragma unroll
            for (int c = 0; c < NCH; ++c)
              if (c * 32 < valid) tmem_ld_32x32b_x32(t_S + c * 32, sv + c * 32);
            tc_wait_ld();
          }
          TR_SM(1);
          // the S accumulator now lives in registers: hand it back for the next QK^T right away
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&bars->s_empty[s]);
          uint32_t pk[BKV / 2];
          float m_new = m_run, sum = 0.f;
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
            float m_scaled = m_new * sl2;
            float sum0 = 0.f, sum1 = 0.f, sum2 = 0.f, sum3 = 0.f;
#if MM_ATTN3_PACKED
            uint64_t sa = 0ull, sb = 0ull;
            const uint64_t k2 = f32x2(sl2, sl2), nm2 = f32x2(-m_scaled, -m_scaled);
#define EXP4(jj)                                                                                                  \
  do {                                                                                                            \
    float x0, x1, x2, x3;                                                                                         \
    f32x2_unpack(fma_f32x2(f32x2(__uint_as_float(sv[(jj) + 0]), __uint_as_float(sv[(jj) + 1])), k2, nm2), x0, x1); \
    const uint64_t xq = fma_f32x2(f32x2(__uint_as_float(sv[(jj) + 2]), __uint_as_float(sv[(jj) + 3])), k2, nm2);    \
    float p2, p3;                                                                                                 \
    if (MM_ATTN3_POLY == 2 || (MM_ATTN3_POLY == 1 && (((jj) >> 2) & 1))) {                                         \
      ex2_poly2(xq, p2, p3);                                                                                      \
    } else {                                                                                                      \
      f32x2_unpack(xq, x2, x3);                                                                                   \
      p2 = ex2(x2), p3 = ex2(x3);                                                                                 \
    }                                                                                                             \
    const float p0 = ex2(x0), p1 = ex2(x1);                                                                       \
    sa = add_f32x2(sa, f32x2(p0, p1));                                                                            \
    sb = add_f32x2(sb, f32x2(p2, p3));                                                                            \
    pk[(jj) >> 1] = pack_bf16x2(p0, p1);                                                                          \
    pk[((jj) >> 1) + 1] = pack_bf16x2(p2, p3);                                                                    \
  } while (0)
#else
#define EXP4(jj)                                                                  \
  do {                                                                            \
    const float p0 = ex2(fmaf(__uint_as_float(sv[(jj) + 0]), sl2, -m_scaled));    \
    const float p1 = ex2(fmaf(__uint_as_float(sv[(jj) + 1]), sl2, -m_scaled));    \
    const float p2 = ex2(fmaf(__uint_as_float(sv[(jj) + 2]), sl2, -m_scaled));    \
    const float p3 = ex2(fmaf(__uint_as_float(sv[(jj) + 3]), sl2, -m_scaled));    \
    sum0 += p0, sum1 += p1, sum2 += p2, sum3 += p3;                               \
    pk[(jj) >> 1] = pack_bf16x2(p0, p1);                                          \
    pk[((jj) >> 1) + 1] = pack_bf16x2(p2, p3);                                    \
  } while (0)
#endif
#if MM_ATTN3_PERMITS
            if (lane == 0) {
              while (atomicAdd(&bars->sem[q], -1) <= 0) {
                atomicAdd(&bars->sem[q], 1);
                __nanosleep(32);
              }
            }
            __syncwarp();
            asm volatile("" : "+f"(m_scaled)::"memory");   // every exp below depends on it: none moves above the acquire
#endif
            if (valid >= BKV) {
#pragma unroll
              for (int jj = 0; jj < BKV; jj += 4) EXP4(jj);
            } else {
              // last (partial) kv tile: chunks that are all padding are skipped
#pragma unroll
              for (int c = 0; c < NCH; ++c) {
                if (c * 32 < valid) {
#pragma unroll
                  for (int jj = c * 32; jj < c * 32 + 32; jj += 4) EXP4(jj);
                } else {
#pragma unroll
                  for (int jj = 0; jj < 16; ++jj) pk[c * 16 + jj] = 0u;
                }
              }
            }
#if MM_ATTN3_PACKED
            f32x2_unpack(sa, sum0, sum1);
            f32x2_unpack(sb, sum2, sum3);
#endif
#undef EXP4
            sum = (sum0 + sum1) + (sum2 + sum3);
#if MM_ATTN3_PERMITS
            asm volatile("" ::"f"(sum) : "memory");          // the row sum (= every exp) is complete before the release
            if (lane == 0) atomicAdd(&bars->sem[q], 1);
#endif
          }
          TR_SM(2);
          // P (TMEM) and the O accumulator are free once PV of the previous iteration retired
          if (it > 0) {
            mbar_wait(&bars->pv_done[s], (it - 1) & 1);
            tc_fence_after();
          }
          TR_SM(3);
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
            tc_wait_st();
          }
          TR_SM(4);
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&bars->p_full[s]);
        }
        // ---- epilogue of the q tile: O / l -> bf16 -> global ----
        mbar_wait(&bars->pv_done[s], (it - 1) & 1);
        tc_fence_after();
        if (warp_live) {
          const float inv_l = 1.0f / l_run;
          const int t = qt * kBQ + row;
          __nv_bfloat16* dst = p.out + ((int64_t)b * p.T + t) * p.ld_out + h * HD;
#pragma unroll
          for (int c = 0; c < HD / 16; ++c) {
            uint32_t o[16];
            tmem_ld_32x32b_x16(t_O + c * 16, o);
            tc_wait_ld();
            if (t < p.T) {
#pragma unroll
              for (int jj = 0; jj < 2; ++jj) {
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
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&bars->o_empty[s]);
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

}  // namespace a3

#ifdef MM_ATTN_TRACE
int attention3_trace_set(long long* buf) {
  return cudaMemcpyToSymbol(a3::g_trace, &buf, sizeof(buf)) == cudaSuccess ? MM_OK : MM_ERR_CUDA;
}
#endif

int make_tmap_3d(CUtensorMap* tm, const void* ptr, int64_t d0, int64_t d1, int64_t d2, int64_t ld1, int64_t ld2,
                 int box_rows);

// head_dim 64 entry used by mm_attention_fwd (attention_sm100.cu).  variant: 0 = 3 streams x 64-column kv tiles.
int launch_attention3(const void* qkv, void* out, int B, int T, int heads, float scale, cudaStream_t st) {
  using namespace a3;
  constexpr int BKV = 64, NSTR = 3;
  using C = Cfg<BKV, NSTR>;
  Params p;
  p.out = reinterpret_cast<__nv_bfloat16*>(out);
  p.B = B, p.T = T, p.heads = heads;
  p.Dp = heads * 64;
  p.ld_out = heads * 64;
  p.scale_log2 = scale * 1.4426950408889634f;
  Maps maps;
  const int64_t W = 3 * (int64_t)p.Dp;
  int rc = make_tmap_3d(&maps.q, qkv, W, T, B, W, (int64_t)T * W, kBQ);
  if (rc != MM_OK) return rc;
  rc = make_tmap_3d(&maps.kv, qkv, W, T, B, W, (int64_t)T * W, BKV);
  if (rc != MM_OK) return rc;
  static std::atomic<bool> attr_set[64];
  int dev = 0;
  MM_CUDA_CHECK(cudaGetDevice(&dev));
  if (dev >= 0 && dev < 64 && !attr_set[dev].load(std::memory_order_acquire)) {
    MM_CUDA_CHECK(cudaFuncSetAttribute(attention3_fwd_kernel<BKV, NSTR>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       C::kSmemBytes));
    attr_set[dev].store(true, std::memory_order_release);
  }
  const int units = B * heads;
  const int sms = sm_count();
  const int grid = units < sms ? units : sms;
  attention3_fwd_kernel<BKV, NSTR><<<grid, C::kThreads, C::kSmemBytes, st>>>(maps, p);
  MM_CUDA_CHECK(cudaGetLastError());
  return MM_OK;
}

}  // namespace mm
