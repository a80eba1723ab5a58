// Micro-probe 2 (sm_100a): pure LDTM / STTM cost, mbarrier hand-off flavours, MMA issue patterns (same vs alternating
// accumulators), and the exp phase with a concurrent TMEM-loading neighbour on the same scheduler.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 --expt-relaxed-constexpr -I multimeditron_b200/csrc -I include
//        tools/micro/tmem_probe2.cu -o tools/micro/tmem_probe2 -lcuda
#include <cstdio>
#include <cstdint>
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include "common.cuh"
using namespace mm;

__device__ __forceinline__ void tc_mma_ts(uint32_t d, uint32_t a, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
               "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}\n" ::"r"(d), "r"(a), "l"(bdesc),
               "r"(idesc), "r"(acc) : "memory");
}
__device__ __forceinline__ bool mbar_test(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile("{\n\t.reg .pred p;\n\tmbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.b32 %0, 1, 0, p;\n\t}\n"
               : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
  return ok != 0;
}
__device__ __forceinline__ float ex2(float x) { float y; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }

#define NRES 32
struct Res { long long v[NRES]; };

__global__ void __launch_bounds__(384, 1) probe(Res* res, float* sink, int reps) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + 96 * 1024);
  uint32_t* tbase = reinterpret_cast<uint32_t*>(bars + 16);
  __shared__ long long sres[NRES];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x < NRES) sres[threadIdx.x] = 0;
  if (threadIdx.x == 0) {
    for (int i = 0; i < 8; ++i) mbar_init(&bars[i], 1);
    fence_barrier_init();
  }
  if (warp == 2) { tmem_alloc(tbase, 512); tmem_relinquish(); }
  for (int i = threadIdx.x; i < 96 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tm = *tbase;
  float acc = 0.f;
  const uint32_t tl = tm + ((uint32_t)((warp & 3) * 32) << 16) + (warp >= 8 ? 128 : 0);
  // ---- 0/1: LDTM 4 x .x32 + wait (values consumed by ONE xor tree after the loop), 1 and 2 warps per scheduler
  for (int pass = 0; pass < 2; ++pass) {
    __syncthreads();
    if (warp >= 4 && warp < (pass == 0 ? 8 : 12)) {
      uint32_t sv[128];
      long long t0 = clock64();
      for (int i = 0; i < reps; ++i) {
#pragma unroll
        for (int c = 0; c < 4; ++c) tmem_ld_32x32b_x32(tl + c * 32, sv + c * 32);
        tc_wait_ld();
      }
      long long t1 = clock64();
      uint32_t x = 0;
#pragma unroll
      for (int j = 0; j < 128; ++j) x ^= sv[j];
      acc += __uint_as_float(x & 0xffff);
      if (threadIdx.x == 128) sres[pass] = (t1 - t0) / reps;
    }
  }
  // ---- 2: LDTM 1 x .x32 + wait (latency of one small load)
  __syncthreads();
  if (warp >= 4 && warp < 8) {
    uint32_t sv[32];
    long long t0 = clock64();
    for (int i = 0; i < reps; ++i) { tmem_ld_32x32b_x32(tl, sv); tc_wait_ld(); }
    long long t1 = clock64();
    uint32_t x = 0;
#pragma unroll
    for (int j = 0; j < 32; ++j) x ^= sv[j];
    acc += __uint_as_float(x & 0xffff);
    if (threadIdx.x == 128) sres[2] = (t1 - t0) / reps;
  }
  // ---- 3: STTM 64 columns + wait::st (2 warps per scheduler); 4: STTM 32 columns + wait
  __syncthreads();
  if (warp >= 4) {
    const uint32_t t = tm + ((uint32_t)((warp & 3) * 32) << 16) + 256 + (warp >= 8 ? 64 : 0);
    uint32_t pk[64];
#pragma unroll
    for (int j = 0; j < 64; ++j) pk[j] = 0x3c003c00u + j + lane;
    long long t0 = clock64();
    for (int i = 0; i < reps; ++i) { tmem_st_32x32b_x32(t, pk); tmem_st_32x32b_x32(t + 32, pk + 32); tc_wait_st(); }
    long long t1 = clock64();
    if (threadIdx.x == 128) sres[3] = (t1 - t0) / reps;
    t0 = clock64();
    for (int i = 0; i < reps; ++i) { tmem_st_32x32b_x32(t, pk); tc_wait_st(); }
    t1 = clock64();
    if (threadIdx.x == 128) sres[4] = (t1 - t0) / reps;
  }
  // ---- 5: mbarrier ping-pong, try_wait on both sides; 6: test_wait spin on both sides
  for (int mode = 0; mode < 2; ++mode) {
    __syncthreads();
    uint64_t* b0 = &bars[mode * 2], *b1 = &bars[mode * 2 + 1];
    if (warp == 4 && lane == 0) {
      long long t0 = clock64();
      for (int i = 0; i < reps; ++i) {
        mbar_arrive(b0);
        if (mode == 0) mbar_wait(b1, i & 1); else while (!mbar_test(b1, i & 1)) {}
      }
      long long t1 = clock64();
      sres[5 + mode] = (t1 - t0) / reps;
    } else if (warp == 5 && lane == 0) {
      for (int i = 0; i < reps; ++i) {
        if (mode == 0) mbar_wait(b0, i & 1); else while (!mbar_test(b0, i & 1)) {}
        mbar_arrive(b1);
      }
    }
  }
  // ---- 7: 4-warp arrive (count 4 via 4 separate barriers is not what the kernel does; measure: 4 warps each
  //         fence+syncwarp+lane0 arrive on a count-4 barrier, consumer = warp 1 lane 0 test_wait spin, replies on a
  //         count-1 barrier every softmax warp waits on with try_wait)  -> full softmax<->issuer round trip
  __syncthreads();
  if (threadIdx.x == 0) { mbar_init(&bars[6], 4); mbar_init(&bars[7], 1); fence_barrier_init(); }
  __syncthreads();
  if (warp >= 4 && warp < 8) {
    long long t0 = clock64();
    for (int i = 0; i < reps; ++i) {
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&bars[6]);
      mbar_wait(&bars[7], i & 1);
      tc_fence_after();
    }
    long long t1 = clock64();
    if (threadIdx.x == 128) sres[7] = (t1 - t0) / reps;
  } else if (warp == 1 && lane == 0) {
    for (int i = 0; i < reps; ++i) { while (!mbar_test(&bars[6], i & 1)) {} mbar_arrive(&bars[7]); }
  }
  // ---- MMA patterns (warp 1 lane 0)
  __syncthreads();
  if (threadIdx.x == 0) { mbar_init(&bars[2], 1); fence_barrier_init(); }
  __syncthreads();
  if (warp == 1 && lane == 0) {
    constexpr uint32_t idesc_s = umma_idesc_bf16(128, 128, 0, 0);
    constexpr uint32_t idesc_256 = umma_idesc_bf16(128, 256, 0, 0);
    constexpr uint32_t idesc_pv = umma_idesc_bf16(128, 64, 0, 1);
    const uint32_t a_addr = smem_u32(smem), b_addr = smem_u32(smem + 32768);
    uint32_t ph = 0;
    long long t0, t1;
#define DRAIN() do { tc_commit(&bars[2]); mbar_wait(&bars[2], ph & 1); ++ph; } while (0)
    // 8: 32 SS N128 MMAs on ONE accumulator, one drain
    t0 = clock64();
    for (int i = 0; i < 32; ++i)
      tc_mma_f16_ss(tm, umma_desc_sw128(a_addr + (i & 3) * 32, 16, 1024), umma_desc_sw128(b_addr + (i & 3) * 32, 16, 1024), idesc_s, 1);
    DRAIN();
    t1 = clock64();
    sres[8] = (t1 - t0) / 32;
    // 9: 32 SS N128 MMAs alternating between TWO accumulators
    t0 = clock64();
    for (int i = 0; i < 32; ++i)
      tc_mma_f16_ss(tm + (i & 1) * 128, umma_desc_sw128(a_addr + (i & 3) * 32, 16, 1024), umma_desc_sw128(b_addr + (i & 3) * 32, 16, 1024), idesc_s, 1);
    DRAIN();
    t1 = clock64();
    sres[9] = (t1 - t0) / 32;
    // 10: 32 SS N128 MMAs over FOUR accumulators
    t0 = clock64();
    for (int i = 0; i < 32; ++i)
      tc_mma_f16_ss(tm + (i & 3) * 128, umma_desc_sw128(a_addr + (i & 3) * 32, 16, 1024), umma_desc_sw128(b_addr + (i & 3) * 32, 16, 1024), idesc_s, 1);
    DRAIN();
    t1 = clock64();
    sres[10] = (t1 - t0) / 32;
    // 11: 32 SS N256 MMAs, one accumulator
    t0 = clock64();
    for (int i = 0; i < 32; ++i)
      tc_mma_f16_ss(tm, umma_desc_sw128(a_addr + (i & 3) * 32, 16, 1024), umma_desc_sw128(b_addr + (i & 3) * 32, 16, 1024), idesc_256, 1);
    DRAIN();
    t1 = clock64();
    sres[11] = (t1 - t0) / 32;
    // 12: 32 TS N64 MMAs on one accumulator; 13: alternating two accumulators (and two P sources)
    t0 = clock64();
    for (int i = 0; i < 32; ++i)
      tc_mma_ts(tm + 384, tm + 256 + (i & 7) * 8, umma_desc_sw128(b_addr + (i & 7) * 2048, 1024, 1024), idesc_pv, 1);
    DRAIN();
    t1 = clock64();
    sres[12] = (t1 - t0) / 32;
    t0 = clock64();
    for (int i = 0; i < 32; ++i)
      tc_mma_ts(tm + 384 + (i & 1) * 64, tm + 256 + (i & 1) * 64 + ((i >> 1) & 7) * 8, umma_desc_sw128(b_addr + ((i >> 1) & 7) * 2048, 1024, 1024), idesc_pv, 1);
    DRAIN();
    t1 = clock64();
    sres[13] = (t1 - t0) / 32;
    // 14: 32 SS N128 with the SAME descriptors every time (smem operand reuse pattern irrelevant?) K-advance 0
    t0 = clock64();
    for (int i = 0; i < 32; ++i)
      tc_mma_f16_ss(tm, umma_desc_sw128(a_addr, 16, 1024), umma_desc_sw128(b_addr, 16, 1024), idesc_s, 1);
    DRAIN();
    t1 = clock64();
    sres[14] = (t1 - t0) / 32;
    // 15: issue cost alone: 32 MMAs issued, clock read BEFORE the drain
    t0 = clock64();
    for (int i = 0; i < 32; ++i)
      tc_mma_f16_ss(tm + (i & 3) * 128, umma_desc_sw128(a_addr + (i & 3) * 32, 16, 1024), umma_desc_sw128(b_addr + (i & 3) * 32, 16, 1024), idesc_s, 1);
    t1 = clock64();
    DRAIN();
    sres[15] = (t1 - t0) / 32;
  }
  // ---- 16: exp phase (128 cols: fma+ex2+sum+pack) one warp per scheduler alone; 17: same while the OTHER warp of
  //          the scheduler loops LDTM 128 cols + 43 FMNMX3-like ops; 18: both warps of the scheduler in the exp phase
  for (int mode = 0; mode < 3; ++mode) {
    __syncthreads();
    if (warp >= 4) {
      const bool expw = warp < 8 || mode == 2;
      const bool active = expw || mode == 1;
      float sv[128];
#pragma unroll
      for (int j = 0; j < 128; ++j) sv[j] = (float)(threadIdx.x * 131 + j) * -1e-3f;
      float a = 0.18f, b = -3.f;
      uint32_t pacc = 0;
      long long t0 = clock64();
      if (active) {
        for (int r = 0; r < reps; ++r) {
          if (expw) {
            float s0 = 0, s1 = 0, s2 = 0, s3 = 0;
            uint32_t pk[64];
#pragma unroll
            for (int j = 0; j < 128; j += 4) {
              float p0 = ex2(fmaf(sv[j], a, b)), p1 = ex2(fmaf(sv[j + 1], a, b)), p2 = ex2(fmaf(sv[j + 2], a, b)), p3 = ex2(fmaf(sv[j + 3], a, b));
              s0 += p0; s1 += p1; s2 += p2; s3 += p3;
              pk[j >> 1] = pack_bf16x2(p0, p1); pk[(j >> 1) + 1] = pack_bf16x2(p2, p3);
            }
            acc += (s0 + s1) + (s2 + s3);
#pragma unroll
            for (int j = 0; j < 64; ++j) pacc ^= pk[j];
            b += 1e-7f * acc;
          } else {
            uint32_t lv[128];
#pragma unroll
            for (int c = 0; c < 4; ++c) tmem_ld_32x32b_x32(tl + c * 32, lv + c * 32);
            tc_wait_ld();
            float m0 = -1e30f, m1 = -1e30f;
#pragma unroll
            for (int j = 0; j < 128; j += 4) {
              m0 = fmaxf(m0, fmaxf(__uint_as_float(lv[j]), __uint_as_float(lv[j + 1])));
              m1 = fmaxf(m1, fmaxf(__uint_as_float(lv[j + 2]), __uint_as_float(lv[j + 3])));
            }
            acc += fmaxf(m0, m1) * 1e-9f;
          }
        }
      }
      long long t1 = clock64();
      acc += __uint_as_float(pacc & 0xff);
      if (threadIdx.x == 128) sres[16 + mode] = (t1 - t0) / reps;
      if (threadIdx.x == 256 && mode == 1) sres[19] = (t1 - t0) / reps;   // the LDTM+max loop next to an exp neighbour
    }
  }
  __syncthreads();
  if (threadIdx.x < NRES) res[blockIdx.x].v[threadIdx.x] = sres[threadIdx.x];
  sink[blockIdx.x * blockDim.x + threadIdx.x] = acc;
  tc_fence_before();
  __syncthreads();
  if (warp == 2) { tc_fence_after(); tmem_dealloc(tm, 512); }
}

namespace mm { int sm_count() { return 148; } }

int main() {
  Res* res; float* sink;
  const int grid = 148;
  cudaMalloc(&res, grid * sizeof(Res)); cudaMalloc(&sink, grid * 384 * 4);
  cudaMemset(res, 0, grid * sizeof(Res));
  const int smem = 1024 + 96 * 1024 + 256;
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  for (int it = 0; it < 2; ++it) probe<<<grid, 384, smem>>>(res, sink, 64);
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { printf("cuda error %s\n", cudaGetErrorString(e)); return 1; }
  Res h[148];
  cudaMemcpy(h, res, sizeof(h), cudaMemcpyDeviceToHost);
  const char* names[] = {
      "LDTM 4 x .x32 + wait::ld, 1 warp/scheduler", "LDTM 4 x .x32 + wait::ld, 2 warps/scheduler", "LDTM 1 x .x32 + wait::ld",
      "STTM 64 cols + wait::st (2 warps/sched)", "STTM 32 cols + wait::st (2 warps/sched)",
      "mbarrier ping-pong, try_wait both sides (round trip)", "mbarrier ping-pong, test_wait spin both sides",
      "softmax(4 warps arrive cnt4)<->issuer(test spin) round trip", "SS N128: per MMA, one accumulator",
      "SS N128: per MMA, two accumulators alternating", "SS N128: per MMA, four accumulators", "SS N256: per MMA, one accumulator",
      "TS N64: per MMA, one accumulator", "TS N64: per MMA, two accumulators alternating", "SS N128: per MMA, same smem every time",
      "SS N128: issue cost only (no drain in the window)", "exp phase 128 cols, 1 warp/scheduler alone",
      "exp phase 128 cols, neighbour warp loops LDTM+max", "exp phase 128 cols, both warps of the scheduler", "LDTM+max loop next to an exp neighbour"};
  for (int i = 0; i < 20; ++i) printf("%-60s %6lld clk (SM0)  %6lld clk (SM147)\n", names[i], h[0].v[i], h[147].v[i]);
  return 0;
}
