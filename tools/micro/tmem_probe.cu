// Micro-probe of the primitives that pace the attention softmax chain on one SM (sm_100a):
//   1 LDTM   clk for a warp to pull 128 fp32 columns (4 x 32x32b.x32 + wait::ld) with 1 / 2 warps per scheduler
//   2 STTM   clk for 64 columns (2 x .x32 + wait::st)
//   3 mbarrier ping-pong between two warps (arrive -> try_wait spin -> arrive back): clk per round trip
//   4 named-barrier ping-pong (bar.sync / bar.arrive, 64 threads)
//   5 tcgen05.mma + commit -> mbarrier wait: 4 x SS 128x128x16, 8 x TS 128x64x16, and an EMPTY commit
//   6 ALU rates: 128 fmax (2-input), 3-input max, 64 bf16x2 packs, per warp with 2 warps per scheduler
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I multimeditron_b200/csrc tools/micro/tmem_probe.cu -o tools/micro/tmem_probe
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
__device__ __forceinline__ float max3(float a, float b, float c) {
  float d;
  asm("max.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
  return d;
}

struct Res { long long v[32]; };

__global__ void __launch_bounds__(384, 1) probe(Res* res, float* sink, int reps) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + 96 * 1024);
  uint32_t* tbase = reinterpret_cast<uint32_t*>(bars + 16);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int i = 0; i < 8; ++i) mbar_init(&bars[i], 1);
    fence_barrier_init();
  }
  if (warp == 2) { tmem_alloc(tbase, 512); tmem_relinquish(); }
  // smem operands: finite bf16 garbage
  for (int i = threadIdx.x; i < 96 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tm = *tbase;
  Res r{};
  float acc = 0.f;
  // ---- 1 LDTM: warps 4..7 (one per scheduler) then warps 4..11 (two per scheduler)
  for (int pass = 0; pass < 2; ++pass) {
    __syncthreads();
    const bool active = warp >= 4 && warp < (pass == 0 ? 8 : 12);
    if (active) {
      const uint32_t t = tm + ((uint32_t)((warp & 3) * 32) << 16) + (warp >= 8 ? 128 : 0);
      long long t0 = clock64();
      for (int i = 0; i < reps; ++i) {
        uint32_t sv[128];
#pragma unroll
        for (int c = 0; c < 4; ++c) tmem_ld_32x32b_x32(t + c * 32, sv + c * 32);
        tc_wait_ld();
#pragma unroll
        for (int j = 0; j < 128; ++j) acc += __uint_as_float(sv[j]);
      }
      long long t1 = clock64();
      if (threadIdx.x == 128) r.v[pass] = (t1 - t0) / reps;
    }
  }
  // ---- 2 STTM 64 columns + wait::st, two warps per scheduler
  __syncthreads();
  if (warp >= 4) {
    const uint32_t t = tm + ((uint32_t)((warp & 3) * 32) << 16) + 256 + (warp >= 8 ? 64 : 0);
    uint32_t pk[64];
#pragma unroll
    for (int j = 0; j < 64; ++j) pk[j] = 0x3c003c00u + j;
    long long t0 = clock64();
    for (int i = 0; i < reps; ++i) {
      tmem_st_32x32b_x32(t, pk);
      tmem_st_32x32b_x32(t + 32, pk + 32);
      tc_wait_st();
      pk[i & 63] ^= 1;
    }
    long long t1 = clock64();
    if (threadIdx.x == 128) r.v[2] = (t1 - t0) / reps;
  }
  // ---- 3 mbarrier ping-pong: warp 4 lane 0 <-> warp 5 lane 0
  __syncthreads();
  if (warp == 4 && lane == 0) {
    long long t0 = clock64();
    for (int i = 0; i < reps; ++i) { mbar_arrive(&bars[0]); mbar_wait(&bars[1], i & 1); }
    long long t1 = clock64();
    r.v[3] = (t1 - t0) / reps;
  } else if (warp == 5 && lane == 0) {
    for (int i = 0; i < reps; ++i) { mbar_wait(&bars[0], i & 1); mbar_arrive(&bars[1]); }
  }
  // ---- 4 named-barrier ping-pong: warp 4 <-> warp 8 (same scheduler), ids 1/2, 64 threads
  __syncthreads();
  if (warp == 4) {
    long long t0 = clock64();
    for (int i = 0; i < reps; ++i) {
      asm volatile("bar.arrive 1, 64;" ::: "memory");
      asm volatile("bar.sync 2, 64;" ::: "memory");
    }
    long long t1 = clock64();
    if (lane == 0) r.v[4] = (t1 - t0) / reps;
  } else if (warp == 8) {
    for (int i = 0; i < reps; ++i) {
      asm volatile("bar.sync 1, 64;" ::: "memory");
      asm volatile("bar.arrive 2, 64;" ::: "memory");
    }
  }
  // ---- 5 MMA + commit + wait (single issuing thread = warp 1 lane 0)
  __syncthreads();
  if (warp == 1 && lane == 0) {
    constexpr uint32_t idesc_s = umma_idesc_bf16(128, 128, 0, 0);
    constexpr uint32_t idesc_pv = umma_idesc_bf16(128, 64, 0, 1);
    constexpr uint32_t idesc_s64 = umma_idesc_bf16(128, 64, 0, 0);
    const uint32_t a_addr = smem_u32(smem), b_addr = smem_u32(smem + 32768);
    uint32_t ph = 0;
    long long t0 = clock64();
    for (int i = 0; i < reps; ++i) {  // empty commit
      tc_commit(&bars[2]);
      mbar_wait(&bars[2], ph & 1); ++ph;
    }
    long long t1 = clock64();
    r.v[5] = (t1 - t0) / reps;
    t0 = clock64();
    for (int i = 0; i < reps; ++i) {  // QK^T: 4 x SS 128x128x16
#pragma unroll
      for (int k = 0; k < 4; ++k)
        tc_mma_f16_ss(tm, umma_desc_sw128(a_addr + k * 32, 16, 1024), umma_desc_sw128(b_addr + k * 32, 16, 1024), idesc_s, k != 0);
      tc_commit(&bars[2]);
      mbar_wait(&bars[2], ph & 1); ++ph;
    }
    t1 = clock64();
    r.v[6] = (t1 - t0) / reps;
    t0 = clock64();
    for (int i = 0; i < reps; ++i) {  // PV: 8 x TS 128x64x16 (A = TMEM cols 256.., B MN-major)
#pragma unroll
      for (int k = 0; k < 8; ++k)
        tc_mma_ts(tm + 384, tm + 256 + k * 8, umma_desc_sw128(b_addr + k * 2048, 1024, 1024), idesc_pv, k != 0);
      tc_commit(&bars[2]);
      mbar_wait(&bars[2], ph & 1); ++ph;
    }
    t1 = clock64();
    r.v[7] = (t1 - t0) / reps;
    t0 = clock64();
    for (int i = 0; i < reps; ++i) {  // back-to-back: QK + PV, one commit (throughput of the pair)
#pragma unroll
      for (int k = 0; k < 4; ++k)
        tc_mma_f16_ss(tm, umma_desc_sw128(a_addr + k * 32, 16, 1024), umma_desc_sw128(b_addr + k * 32, 16, 1024), idesc_s, k != 0);
#pragma unroll
      for (int k = 0; k < 8; ++k)
        tc_mma_ts(tm + 384, tm + 256 + k * 8, umma_desc_sw128(b_addr + k * 2048, 1024, 1024), idesc_pv, k != 0);
      if ((i & 7) == 7) { tc_commit(&bars[2]); mbar_wait(&bars[2], ph & 1); ++ph; }
    }
    t1 = clock64();
    r.v[8] = (t1 - t0) / reps;
    t0 = clock64();
    for (int i = 0; i < reps; ++i) {  // QK^T with a 64-column kv tile: 4 x SS 128x64x16, 8 in flight per commit
#pragma unroll
      for (int k = 0; k < 4; ++k)
        tc_mma_f16_ss(tm, umma_desc_sw128(a_addr + k * 32, 16, 1024), umma_desc_sw128(b_addr + k * 32, 16, 1024), idesc_s64, k != 0);
      if ((i & 7) == 7) { tc_commit(&bars[2]); mbar_wait(&bars[2], ph & 1); ++ph; }
    }
    t1 = clock64();
    r.v[9] = (t1 - t0) / reps;
    t0 = clock64();
    for (int i = 0; i < reps; ++i) {  // SS 128x128x16 x4 throughput (8 groups per commit)
#pragma unroll
      for (int k = 0; k < 4; ++k)
        tc_mma_f16_ss(tm, umma_desc_sw128(a_addr + k * 32, 16, 1024), umma_desc_sw128(b_addr + k * 32, 16, 1024), idesc_s, k != 0);
      if ((i & 7) == 7) { tc_commit(&bars[2]); mbar_wait(&bars[2], ph & 1); ++ph; }
    }
    t1 = clock64();
    r.v[10] = (t1 - t0) / reps;
    t0 = clock64();
    for (int i = 0; i < reps; ++i) {  // TS 128x64x16 x8 throughput
#pragma unroll
      for (int k = 0; k < 8; ++k)
        tc_mma_ts(tm + 384, tm + 256 + k * 8, umma_desc_sw128(b_addr + k * 2048, 1024, 1024), idesc_pv, k != 0);
      if ((i & 7) == 7) { tc_commit(&bars[2]); mbar_wait(&bars[2], ph & 1); ++ph; }
    }
    t1 = clock64();
    r.v[11] = (t1 - t0) / reps;
  }
  // ---- 6 ALU rates, warps 4..11 (two per scheduler)
  __syncthreads();
  if (warp >= 4) {
    float sv[128];
#pragma unroll
    for (int j = 0; j < 128; ++j) sv[j] = (float)(threadIdx.x * 131 + j) * 1e-3f + acc * 1e-9f;
    long long t0 = clock64();
    float m = -1e30f;
    for (int i = 0; i < reps; ++i) {
      float m0 = m, m1 = m, m2 = m, m3 = m;
#pragma unroll
      for (int j = 0; j < 128; j += 8) {
        m0 = fmaxf(m0, fmaxf(sv[j], sv[j + 1])); m1 = fmaxf(m1, fmaxf(sv[j + 2], sv[j + 3]));
        m2 = fmaxf(m2, fmaxf(sv[j + 4], sv[j + 5])); m3 = fmaxf(m3, fmaxf(sv[j + 6], sv[j + 7]));
      }
      m = fmaxf(fmaxf(m0, m1), fmaxf(m2, m3)) * 0.999f;
      sv[i & 127] += m * 1e-6f;
    }
    long long t1 = clock64();
    if (threadIdx.x == 128) r.v[12] = (t1 - t0) / reps;
    acc += m;
    t0 = clock64();
    for (int i = 0; i < reps; ++i) {
      float m0 = m, m1 = m;
#pragma unroll
      for (int j = 0; j < 128; j += 4) { m0 = max3(m0, sv[j], sv[j + 1]); m1 = max3(m1, sv[j + 2], sv[j + 3]); }
      m = fmaxf(m0, m1) * 0.999f;
      sv[i & 127] += m * 1e-6f;
    }
    t1 = clock64();
    if (threadIdx.x == 128) r.v[13] = (t1 - t0) / reps;
    acc += m;
    uint32_t x = 0;
    t0 = clock64();
    for (int i = 0; i < reps; ++i) {
#pragma unroll
      for (int j = 0; j < 128; j += 2) x ^= pack_bf16x2(sv[j], sv[j + 1]);
      sv[i & 127] += __uint_as_float(x & 0xff) * 1e-30f;
    }
    t1 = clock64();
    if (threadIdx.x == 128) r.v[14] = (t1 - t0) / reps;
    acc += __uint_as_float(x & 0xffff);
  }
  __syncthreads();
  if (threadIdx.x == 128) {
    for (int i = 0; i < 3; ++i) res[blockIdx.x].v[i] = r.v[i];
    for (int i = 12; i < 15; ++i) res[blockIdx.x].v[i] = r.v[i];
  }
  if (warp == 4 && lane == 0) res[blockIdx.x].v[3] = r.v[3], res[blockIdx.x].v[4] = r.v[4];
  if (warp == 1 && lane == 0) for (int i = 5; i < 12; ++i) res[blockIdx.x].v[i] = r.v[i];
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
  const char* names[] = {"LDTM 128 cols + wait, 1 warp/scheduler", "LDTM 128 cols + wait, 2 warps/scheduler",
                         "STTM 64 cols + wait::st, 2 warps/scheduler", "mbarrier ping-pong round trip",
                         "named-barrier ping-pong round trip", "empty tcgen05.commit -> mbarrier wait",
                         "4 x SS 128x128x16 + commit + wait", "8 x TS 128x64x16 + commit + wait",
                         "QK(4 SS N128) + PV(8 TS N64) throughput per pair", "4 x SS 128x64x16 throughput",
                         "4 x SS 128x128x16 throughput", "8 x TS 128x64x16 throughput",
                         "128 fmax (2-input), 2 warps/scheduler", "128 max via 3-input max.f32", "64 bf16x2 packs"};
  for (int i = 0; i < 15; ++i) printf("%-52s %6lld clk (SM0)  %6lld clk (SM147)\n", names[i], h[0].v[i], h[147].v[i]);
  return 0;
}
