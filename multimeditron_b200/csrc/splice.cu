// splice.cu — attachment-position prefix sum, row scatter (splice), embedding gather and the fused
// embed+splice pass.  HBM-bound byte movement: 16-byte vector loads/stores, streaming cache hints,
// grids sized in multiples of the SM count.  Bit-exact by construction (rows are copied, or cast
// with round-to-nearest-even exactly like Tensor.to()).
//
// Reference statements (relative to /root/reference):
//   src/multimeditron/model/model.py:433        embedded_tokens = embed_tokens(input_ids)
//   src/multimeditron/model/model.py:441-444    embedded_tokens[batch_idx, token_range] = proj.view(-1,H).to(dtype)
//   src/multimeditron/model/data_loader.py:124-141 + prompt_tokenizers.py:255-286  (index construction)
#include "common.cuh"

namespace mm {

constexpr int kScanThreads = 256;
constexpr int kScanItems = 8;                           // ids per thread
constexpr int kScanChunk = kScanThreads * kScanItems;   // ids per block

__device__ __forceinline__ int4 ld_stream(const int4* p) {
  int4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.s32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ void st_stream(int4* p, const int4& v) {
  asm volatile("st.global.L1::no_allocate.v4.s32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z),
               "r"(v.w)
               : "memory");
}

// ---- attach positions ------------------------------------------------------------------------------
__device__ __forceinline__ int load_flags(const int64_t* __restrict__ ids, int64_t n, int64_t base,
                                          int64_t attach, bool (&f)[kScanItems]) {
  int c = 0;
  if (base + kScanItems <= n) {
    const longlong2* p = reinterpret_cast<const longlong2*>(ids + base);  // base % 8 == 0 -> 16-B aligned
#pragma unroll
    for (int i = 0; i < kScanItems / 2; ++i) {
      longlong2 v = __ldg(p + i);
      f[2 * i] = (v.x == attach);
      f[2 * i + 1] = (v.y == attach);
    }
  } else {
#pragma unroll
    for (int i = 0; i < kScanItems; ++i) f[i] = (base + i < n) && (ids[base + i] == attach);
  }
#pragma unroll
  for (int i = 0; i < kScanItems; ++i) c += f[i];
  return c;
}

__device__ __forceinline__ int block_exclusive_scan(int v, int* total, int* smem /*>= 33 ints*/) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int incl = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    int t = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += t;
  }
  if (lane == 31) smem[warp] = incl;
  __syncthreads();
  if (warp == 0) {
    int w = (lane < (int)(blockDim.x >> 5)) ? smem[lane] : 0;
    int wi = w;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      int t = __shfl_up_sync(0xffffffffu, wi, o);
      if (lane >= o) wi += t;
    }
    smem[lane] = wi - w;  // exclusive warp offsets
    if (lane == 31) smem[32] = wi;
  }
  __syncthreads();
  int r = smem[warp] + incl - v;
  if (total) *total = smem[32];
  __syncthreads();
  return r;
}

__global__ void __launch_bounds__(kScanThreads) attach_count_kernel(const int64_t* __restrict__ ids, int64_t n,
                                                                   int64_t attach, int32_t* __restrict__ block_counts) {
  __shared__ int s[33];
  bool f[kScanItems];
  const int64_t base = (int64_t)blockIdx.x * kScanChunk + (int64_t)threadIdx.x * kScanItems;
  int c = load_flags(ids, n, base, attach, f);
  int total;
  block_exclusive_scan(c, &total, s);
  if (threadIdx.x == 0) block_counts[blockIdx.x] = total;
}

// single block: exclusive scan of block_counts in place; total -> count[0]
__global__ void __launch_bounds__(1024) attach_scan_kernel(int32_t* __restrict__ block_counts, int nblk,
                                                            int32_t* __restrict__ count, int64_t expected,
                                                            int32_t* __restrict__ status) {
  __shared__ int s[33];
  int carry = 0;
  for (int base = 0; base < nblk; base += blockDim.x) {
    int i = base + threadIdx.x;
    int v = (i < nblk) ? block_counts[i] : 0;
    int total;
    int ex = block_exclusive_scan(v, &total, s);
    if (i < nblk) block_counts[i] = carry + ex;
    carry += total;
  }
  if (threadIdx.x == 0) {
    count[0] = carry;
    if (expected >= 0 && (int64_t)carry != expected && status) atomicOr(status, MM_STATUS_COUNT_MISMATCH);
  }
}

__global__ void __launch_bounds__(kScanThreads)
    attach_write_kernel(const int64_t* __restrict__ ids, int64_t n, int64_t attach,
                        const int32_t* __restrict__ block_offsets, int32_t* __restrict__ dest_rows, int64_t cap,
                        int32_t* __restrict__ src_of_pos) {
  __shared__ int s[33];
  bool f[kScanItems];
  const int64_t base = (int64_t)blockIdx.x * kScanChunk + (int64_t)threadIdx.x * kScanItems;
  int c = load_flags(ids, n, base, attach, f);
  int rank = block_offsets[blockIdx.x] + block_exclusive_scan(c, nullptr, s);
  int32_t sp[kScanItems];
#pragma unroll
  for (int i = 0; i < kScanItems; ++i) {
    sp[i] = -1;
    if (f[i]) {
      sp[i] = rank;
      if (rank < cap) dest_rows[rank] = (int32_t)(base + i);
      ++rank;
    }
  }
  if (src_of_pos) {
    if (base + kScanItems <= n) {
      int4* o = reinterpret_cast<int4*>(src_of_pos + base);
      o[0] = make_int4(sp[0], sp[1], sp[2], sp[3]);
      o[1] = make_int4(sp[4], sp[5], sp[6], sp[7]);
    } else {
#pragma unroll
      for (int i = 0; i < kScanItems; ++i)
        if (base + i < n) src_of_pos[base + i] = sp[i];
    }
  }
}

// ---- row movers --------------------------------------------------------------------------------------
// One "vec" = 8 elements: 16 B of bf16 or 32 B of fp32.
template <typename T>
struct Vec8;
template <>
struct Vec8<__nv_bfloat16> {
  int4 a;
  __device__ __forceinline__ void load(const __nv_bfloat16* p) { a = ld_stream(reinterpret_cast<const int4*>(p)); }
  __device__ __forceinline__ void store(__nv_bfloat16* p) const { st_stream(reinterpret_cast<int4*>(p), a); }
};
template <>
struct Vec8<float> {
  int4 a, b;
  __device__ __forceinline__ void load(const float* p) {
    a = ld_stream(reinterpret_cast<const int4*>(p));
    b = ld_stream(reinterpret_cast<const int4*>(p) + 1);
  }
  __device__ __forceinline__ void store(float* p) const {
    st_stream(reinterpret_cast<int4*>(p), a);
    st_stream(reinterpret_cast<int4*>(p) + 1, b);
  }
};
__device__ __forceinline__ void convert(const Vec8<__nv_bfloat16>& s, Vec8<__nv_bfloat16>& d) { d.a = s.a; }
__device__ __forceinline__ void convert(const Vec8<float>& s, Vec8<float>& d) {
  d.a = s.a;
  d.b = s.b;
}
__device__ __forceinline__ void convert(const Vec8<__nv_bfloat16>& s, Vec8<float>& d) {  // exact widening
  d.a = make_int4(s.a.x << 16, s.a.x & 0xFFFF0000, s.a.y << 16, s.a.y & 0xFFFF0000);
  d.b = make_int4(s.a.z << 16, s.a.z & 0xFFFF0000, s.a.w << 16, s.a.w & 0xFFFF0000);
}
__device__ __forceinline__ void convert(const Vec8<float>& s, Vec8<__nv_bfloat16>& d) {  // RNE like Tensor.to()
  d.a.x = pack_bf16x2(__int_as_float(s.a.x), __int_as_float(s.a.y));
  d.a.y = pack_bf16x2(__int_as_float(s.a.z), __int_as_float(s.a.w));
  d.a.z = pack_bf16x2(__int_as_float(s.b.x), __int_as_float(s.b.y));
  d.a.w = pack_bf16x2(__int_as_float(s.b.z), __int_as_float(s.b.w));
}
template <typename S, typename D>
__device__ __forceinline__ D cvt1(S v);
template <>
__device__ __forceinline__ float cvt1<float, float>(float v) { return v; }
template <>
__device__ __forceinline__ __nv_bfloat16 cvt1<__nv_bfloat16, __nv_bfloat16>(__nv_bfloat16 v) { return v; }
template <>
__device__ __forceinline__ float cvt1<__nv_bfloat16, float>(__nv_bfloat16 v) { return __bfloat162float(v); }
template <>
__device__ __forceinline__ __nv_bfloat16 cvt1<float, __nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }

constexpr int kMoveThreads = 256;
constexpr int kMoveUnroll = 8;

// mode 0: scatter  dest[rows[k]] = src[k]          (k over n_rows)
// mode 1: scatter with (batch_idx, token_range) int64 pairs
// mode 2: fused    out[r] = sop[r]>=0 ? src[sop[r]] : table[ids[r]]   (r over n_pos)
// mode 3: gather   out[r] = table[ids[r]]   (rows with ids[r] == skip_id are left untouched: the splice overwrites them)
// mode 4: gather   out[k] = src[rows[k]]     (backward of the splice: gradient rows of the projected tokens)
struct MoveArgs {
  void* dest;
  const void* src;
  const void* table;
  const int32_t* rows;
  const int64_t* bidx;
  const int64_t* trange;
  const int64_t* ids;
  const int32_t* sop;
  int32_t* status;
  int64_t n_units;     // rows iterated
  int64_t n_dest_rows; // bound for scatter destinations
  int64_t n_src_rows;  // bound for src rows (fused)
  int64_t batch, seq, vocab, hidden;
  int64_t skip_id;     // mode 3: id whose rows are not written (-1: none)
};

template <int MODE, typename S, typename D>
__device__ __forceinline__ bool resolve(const MoveArgs& a, int64_t u, const S*& sp, D*& dp) {
  const int64_t H = a.hidden;
  if (MODE == 0) {
    int64_t r = a.rows[u];
    if (r < 0 || r >= a.n_dest_rows) {
      if (a.status) atomicOr(a.status, MM_STATUS_INDEX_OOB);
      return false;
    }
    sp = reinterpret_cast<const S*>(a.src) + u * H;
    dp = reinterpret_cast<D*>(a.dest) + r * H;
    return true;
  } else if (MODE == 4) {
    int64_t r = a.rows[u];
    if (r < 0 || r >= a.n_src_rows) {
      if (a.status) atomicOr(a.status, MM_STATUS_INDEX_OOB);
      return false;
    }
    sp = reinterpret_cast<const S*>(a.src) + r * H;
    dp = reinterpret_cast<D*>(a.dest) + u * H;
    return true;
  } else if (MODE == 1) {
    int64_t b = a.bidx[u], t = a.trange[u];
    if (b < 0) b += a.batch;  // Python negative-index wrap (ATen index_put_)
    if (t < 0) t += a.seq;
    if (b < 0 || b >= a.batch || t < 0 || t >= a.seq) {
      if (a.status) atomicOr(a.status, MM_STATUS_INDEX_OOB);
      return false;
    }
    sp = reinterpret_cast<const S*>(a.src) + u * H;
    dp = reinterpret_cast<D*>(a.dest) + (b * a.seq + t) * H;
    return true;
  } else {
    dp = reinterpret_cast<D*>(a.dest) + u * H;
    int32_t k = (MODE == 2) ? a.sop[u] : -1;
    if (k >= 0) {
      if (k >= a.n_src_rows) {
        if (a.status) atomicOr(a.status, MM_STATUS_COUNT_MISMATCH);
        return false;
      }
      sp = reinterpret_cast<const S*>(a.src) + (int64_t)k * H;
      return true;
    }
    return false;  // table rows are handled by the caller (different source dtype = D)
  }
}

// One warp moves one 4-KiB segment (32 lanes x kMoveUnroll 16-B vectors) of one row per iteration: the row lookup
// (index load, bounds check, 64-bit address math) is warp-uniform and amortised over the segment, and every lane
// keeps kMoveUnroll independent 16-B loads in flight before the first store.
template <int MODE, typename S, typename D>
__global__ void __launch_bounds__(kMoveThreads) move_rows_kernel(MoveArgs a) {
  const int lane = threadIdx.x & 31;
  const int vpr = (int)(a.hidden >> 3);  // vecs per row
  constexpr int kSegVecs = 32 * kMoveUnroll;
  const int segs = (vpr + kSegVecs - 1) / kSegVecs;
  const int64_t total = a.n_units * segs;
  const int64_t nwarps = (int64_t)gridDim.x * (kMoveThreads / 32);
  for (int64_t w = (int64_t)blockIdx.x * (kMoveThreads / 32) + (threadIdx.x >> 5); w < total; w += nwarps) {
    const int64_t u = w / segs;
    const int v0 = (int)(w - u * segs) * kSegVecs + lane;
    const S* sp = nullptr;
    D* dp = nullptr;
    int kind = 0;  // 0 skip, 1 from src (S), 2 from table (D)
    const D* tp = nullptr;
    if (resolve<MODE, S, D>(a, u, sp, dp)) {
      kind = 1;
    } else if (MODE == 2 || MODE == 3) {
      const bool is_src = (MODE == 2) && (a.sop[u] >= 0);
      if (!is_src) {
        const int64_t id = a.ids[u];
        if (id < 0 || id >= a.vocab) {
          if (a.status && lane == 0) atomicOr(a.status, MM_STATUS_INDEX_OOB);
        } else if (MODE == 3 && id == a.skip_id) {
          // attachment position: the splice writes this row
        } else {
          kind = 2;
          tp = reinterpret_cast<const D*>(a.table) + id * a.hidden;
        }
      }
    }
    if (kind == 1) {
      Vec8<S> vs[kMoveUnroll];
#pragma unroll
      for (int j = 0; j < kMoveUnroll; ++j)
        if (v0 + j * 32 < vpr) vs[j].load(sp + (int64_t)(v0 + j * 32) * 8);
#pragma unroll
      for (int j = 0; j < kMoveUnroll; ++j)
        if (v0 + j * 32 < vpr) {
          Vec8<D> vd;
          convert(vs[j], vd);
          vd.store(dp + (int64_t)(v0 + j * 32) * 8);
        }
    } else if (kind == 2) {
      Vec8<D> vd[kMoveUnroll];
#pragma unroll
      for (int j = 0; j < kMoveUnroll; ++j)
        if (v0 + j * 32 < vpr) vd[j].load(tp + (int64_t)(v0 + j * 32) * 8);
#pragma unroll
      for (int j = 0; j < kMoveUnroll; ++j)
        if (v0 + j * 32 < vpr) vd[j].store(dp + (int64_t)(v0 + j * 32) * 8);
    }
  }
}

// generic element-wise path for hidden % 8 != 0 (never hit by the BASELINE configs, kept for API totality)
template <int MODE, typename S, typename D>
__global__ void __launch_bounds__(kMoveThreads) move_rows_scalar_kernel(MoveArgs a) {
  const int64_t total = a.n_units * a.hidden;
  const int64_t stride = (int64_t)gridDim.x * kMoveThreads;
  for (int64_t i = (int64_t)blockIdx.x * kMoveThreads + threadIdx.x; i < total; i += stride) {
    const int64_t u = i / a.hidden, c = i - u * a.hidden;
    const S* sp = nullptr;
    D* dp = nullptr;
    if (resolve<MODE, S, D>(a, u, sp, dp)) {
      dp[c] = cvt1<S, D>(sp[c]);
    } else if (MODE == 2 || MODE == 3) {
      bool is_src = (MODE == 2) && (a.sop[u] >= 0);
      int64_t id = a.ids[u];
      if (!is_src) {
        if (id < 0 || id >= a.vocab) {
          if (a.status) atomicOr(a.status, MM_STATUS_INDEX_OOB);
        } else if (MODE == 3 && id == a.skip_id) {
        } else {
          dp[c] = reinterpret_cast<const D*>(a.table)[id * a.hidden + c];
        }
      }
    }
  }
}

template <int MODE, typename S, typename D>
static int launch_move(const MoveArgs& a, cudaStream_t st) {
  if (a.n_units == 0 || a.hidden == 0) return MM_OK;
  const bool vec = (a.hidden % 8 == 0) && ((reinterpret_cast<uintptr_t>(a.dest) & 15) == 0) &&
                   ((reinterpret_cast<uintptr_t>(a.src) & 15) == 0) &&
                   ((reinterpret_cast<uintptr_t>(a.table) & 15) == 0);
  const int sms = sm_count();
  if (vec) {
    const int64_t vpr = a.hidden >> 3;
    const int64_t segs = (vpr + 32 * kMoveUnroll - 1) / (32 * kMoveUnroll);
    const int64_t total = a.n_units * segs;               // warp-iterations
    int64_t need = (total + kMoveThreads / 32 - 1) / (kMoveThreads / 32);
    int64_t cap = (int64_t)sms * 8;  // 8 resident 256-thread CTAs per SM
    int grid = (int)(need < cap ? need : cap);
    move_rows_kernel<MODE, S, D><<<grid, kMoveThreads, 0, st>>>(a);
  } else {
    const int64_t total = a.n_units * a.hidden;
    int64_t need = (total + kMoveThreads - 1) / kMoveThreads;
    int64_t cap = (int64_t)sms * 8;
    int grid = (int)(need < cap ? need : cap);
    move_rows_scalar_kernel<MODE, S, D><<<grid, kMoveThreads, 0, st>>>(a);
  }
  MM_CUDA_CHECK(cudaGetLastError());
  return MM_OK;
}

template <int MODE>
static int dispatch_move(const MoveArgs& a, int src_dtype, int dst_dtype, cudaStream_t st) {
  if (src_dtype == MM_BF16 && dst_dtype == MM_BF16) return launch_move<MODE, __nv_bfloat16, __nv_bfloat16>(a, st);
  if (src_dtype == MM_F32 && dst_dtype == MM_F32) return launch_move<MODE, float, float>(a, st);
  if (src_dtype == MM_BF16 && dst_dtype == MM_F32) return launch_move<MODE, __nv_bfloat16, float>(a, st);
  if (src_dtype == MM_F32 && dst_dtype == MM_BF16) return launch_move<MODE, float, __nv_bfloat16>(a, st);
  return MM_ERR_BAD_ARG;
}

}  // namespace mm

using namespace mm;

extern "C" int64_t mm_attach_positions_workspace_bytes(int64_t n_ids) {
  int64_t nblk = (n_ids + kScanChunk - 1) / kScanChunk;
  if (nblk < 1) nblk = 1;
  return nblk * (int64_t)sizeof(int32_t);
}

extern "C" int mm_attach_positions(const int64_t* input_ids, int64_t n_ids, int64_t attachment_id,
                                   int32_t* dest_rows, int64_t n_rows_cap, int32_t* src_of_pos, int32_t* count,
                                   int64_t expected_rows, int32_t* status, void* workspace, void* stream) {
  if (n_ids < 0 || n_ids >= (int64_t)INT32_MAX || !count || !workspace || (n_ids > 0 && !input_ids))
    return MM_ERR_BAD_ARG;
  if ((reinterpret_cast<uintptr_t>(input_ids) & 15) || (reinterpret_cast<uintptr_t>(src_of_pos) & 15))
    return MM_ERR_BAD_ARG;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  int32_t* blk = reinterpret_cast<int32_t*>(workspace);
  const int nblk = (int)((n_ids + kScanChunk - 1) / kScanChunk);
  // rows beyond the actual count stay -1 (= "skip") so a count mismatch can never scatter through garbage
  if (dest_rows && n_rows_cap > 0) MM_CUDA_CHECK(cudaMemsetAsync(dest_rows, 0xFF, (size_t)n_rows_cap * 4, st));
  if (nblk > 0) {
    attach_count_kernel<<<nblk, kScanThreads, 0, st>>>(input_ids, n_ids, attachment_id, blk);
    MM_CUDA_CHECK(cudaGetLastError());
  }
  attach_scan_kernel<<<1, 1024, 0, st>>>(blk, nblk, count, expected_rows, status);
  MM_CUDA_CHECK(cudaGetLastError());
  if (nblk > 0 && (dest_rows || src_of_pos)) {
    attach_write_kernel<<<nblk, kScanThreads, 0, st>>>(input_ids, n_ids, attachment_id, blk, dest_rows,
                                                       dest_rows ? n_rows_cap : 0, src_of_pos);
    MM_CUDA_CHECK(cudaGetLastError());
  }
  return MM_OK;
}

// ---- mm_attach_positions_static: the same prefix sum for hosts that cannot allocate device memory -------
// Some hosts of this library have no way to get scratch memory on the GPU: a serving runtime that reserved
// all of HBM for its own pool at start-up and forbids cudaMalloc afterwards, a plugin that is handed raw
// device pointers for the inputs and outputs but owns no allocator, a binding from a language where the
// only CUDA the caller sees is this library.  For them the per-block attachment counters live in the array
// below, which is part of the library image: every GPU that loads libmm_b200.so holds these 32 KB until the
// process exits, and no call here allocates or frees anything.
//
// In HOST code `g_attach_static_scratch` is only a handle for the symbol, not an address the GPU can use —
// passing it to a kernel would fault (or quietly hit host memory on systems where the GPU can reach it).  The
// device address is looked up with cudaGetSymbolAddress, once per device, with that device current, and
// remembered in g_attach_static_ptr[device].
constexpr int64_t kAttachStaticMaxBlocks = MM_ATTACH_STATIC_MAX_IDS / kScanChunk;
static_assert(kAttachStaticMaxBlocks * kScanChunk == MM_ATTACH_STATIC_MAX_IDS, "limit must be whole scan chunks");
static __device__ __align__(256) int32_t g_attach_static_scratch[kAttachStaticMaxBlocks];
constexpr int kAttachStaticMaxDevices = 64;
static std::atomic<int32_t*> g_attach_static_ptr[kAttachStaticMaxDevices];

extern "C" int mm_attach_positions_static(const int64_t* input_ids, int64_t n_ids, int64_t attachment_id,
                                          int32_t* dest_rows, int64_t n_rows_cap, int32_t* src_of_pos,
                                          int32_t* count, int64_t expected_rows, int32_t* status, void* stream) {
  // the array has room for MM_ATTACH_STATIC_MAX_IDS ids and cannot grow: say so instead of running past it
  if (n_ids < 0 || n_ids > MM_ATTACH_STATIC_MAX_IDS) return MM_ERR_BAD_ARG;
  int dev = -1;
  MM_CUDA_CHECK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kAttachStaticMaxDevices) return MM_ERR_BAD_ARG;
  int32_t* scratch = g_attach_static_ptr[dev].load(std::memory_order_acquire);
  if (!scratch) {
    // first call on this device: resolve the symbol to this device's copy of the array.  Two threads racing here
    // get the same answer, so the unsynchronised store is harmless.
    void* p = nullptr;
    MM_CUDA_CHECK(cudaGetSymbolAddress(&p, g_attach_static_scratch));
    if (!p) return MM_ERR_CUDA;
    scratch = static_cast<int32_t*>(p);
    g_attach_static_ptr[dev].store(scratch, std::memory_order_release);
  }
  // Nothing has to be cleared between calls: attach_count_kernel writes counter i of every block i < nblk before
  // attach_scan_kernel / attach_write_kernel read it, and counters >= nblk are never read.  What the array keeps
  // from the previous call is therefore invisible — as long as calls do not overlap on the GPU.
  return mm_attach_positions(input_ids, n_ids, attachment_id, dest_rows, n_rows_cap, src_of_pos, count,
                             expected_rows, status, scratch, stream);
}

extern "C" int mm_splice_scatter(void* dest, int64_t n_dest_rows, int64_t hidden, int dst_dtype, const void* src,
                                 int64_t n_rows, int src_dtype, const int32_t* dest_rows, int32_t* status,
                                 void* stream) {
  if (n_rows < 0 || hidden < 0 || n_dest_rows < 0) return MM_ERR_BAD_ARG;
  if (n_rows > 0 && (!dest || !src || !dest_rows)) return MM_ERR_BAD_ARG;
  MoveArgs a{};
  a.dest = dest;
  a.src = src;
  a.rows = dest_rows;
  a.status = status;
  a.n_units = n_rows;
  a.n_dest_rows = n_dest_rows;
  a.hidden = hidden;
  return dispatch_move<0>(a, src_dtype, dst_dtype, reinterpret_cast<cudaStream_t>(stream));
}

extern "C" int mm_splice_scatter_idx(void* dest, int64_t batch, int64_t seq, int64_t hidden, int dst_dtype,
                                     const void* src, int64_t n_rows, int src_dtype, const int64_t* batch_idx,
                                     const int64_t* token_range, int32_t* status, void* stream) {
  if (n_rows < 0 || hidden < 0 || batch < 0 || seq < 0) return MM_ERR_BAD_ARG;
  if (n_rows > 0 && (!dest || !src || !batch_idx || !token_range)) return MM_ERR_BAD_ARG;
  MoveArgs a{};
  a.dest = dest;
  a.src = src;
  a.bidx = batch_idx;
  a.trange = token_range;
  a.status = status;
  a.n_units = n_rows;
  a.batch = batch;
  a.seq = seq;
  a.hidden = hidden;
  return dispatch_move<1>(a, src_dtype, dst_dtype, reinterpret_cast<cudaStream_t>(stream));
}

extern "C" int mm_embed_splice_fused(void* out, int64_t n_pos, int64_t hidden, int dst_dtype, const void* table,
                                     int64_t vocab, const int64_t* input_ids, const int32_t* src_of_pos,
                                     const void* src, int64_t n_rows, int src_dtype, int32_t* status,
                                     void* stream) {
  if (n_pos < 0 || hidden < 0 || vocab < 0 || n_rows < 0) return MM_ERR_BAD_ARG;
  if (n_pos > 0 && (!out || !table || !input_ids || !src_of_pos)) return MM_ERR_BAD_ARG;
  MoveArgs a{};
  a.dest = out;
  a.src = src;
  a.table = table;
  a.ids = input_ids;
  a.sop = src_of_pos;
  a.status = status;
  a.n_units = n_pos;
  a.n_src_rows = n_rows;
  a.vocab = vocab;
  a.hidden = hidden;
  return dispatch_move<2>(a, src_dtype, dst_dtype, reinterpret_cast<cudaStream_t>(stream));
}

static int embed_gather_impl(void* out, int64_t n_pos, int64_t hidden, int dtype, const void* table, int64_t vocab,
                             const int64_t* input_ids, int64_t skip_id, int32_t* status, void* stream) {
  if (n_pos < 0 || hidden < 0 || vocab < 0) return MM_ERR_BAD_ARG;
  if (n_pos > 0 && (!out || !table || !input_ids)) return MM_ERR_BAD_ARG;
  MoveArgs a{};
  a.dest = out;
  a.table = table;
  a.ids = input_ids;
  a.status = status;
  a.n_units = n_pos;
  a.vocab = vocab;
  a.hidden = hidden;
  a.skip_id = skip_id;
  return dispatch_move<3>(a, dtype, dtype, reinterpret_cast<cudaStream_t>(stream));
}

extern "C" int mm_embed_gather(void* out, int64_t n_pos, int64_t hidden, int dtype, const void* table,
                               int64_t vocab, const int64_t* input_ids, int32_t* status, void* stream) {
  return embed_gather_impl(out, n_pos, hidden, dtype, table, vocab, input_ids, -1, status, stream);
}

extern "C" int mm_embed_gather_skip(void* out, int64_t n_pos, int64_t hidden, int dtype, const void* table,
                                    int64_t vocab, const int64_t* input_ids, int64_t skip_id, int32_t* status,
                                    void* stream) {
  return embed_gather_impl(out, n_pos, hidden, dtype, table, vocab, input_ids, skip_id, status, stream);
}

extern "C" int mm_gather_rows(void* out, int64_t n_rows, int64_t hidden, int dst_dtype, const void* src,
                              int64_t n_src_rows, int src_dtype, const int32_t* rows, int32_t* status, void* stream) {
  if (n_rows < 0 || hidden < 0 || n_src_rows < 0) return MM_ERR_BAD_ARG;
  if (n_rows > 0 && (!out || !src || !rows)) return MM_ERR_BAD_ARG;
  MoveArgs a{};
  a.dest = out;
  a.src = src;
  a.rows = rows;
  a.status = status;
  a.n_units = n_rows;
  a.n_src_rows = n_src_rows;
  a.hidden = hidden;
  a.skip_id = -1;
  return dispatch_move<4>(a, src_dtype, dst_dtype, reinterpret_cast<cudaStream_t>(stream));
}

// ---------------------------------------------------------------------------------------------
// position ids: the collator's `cumsum(mask) - 1` and its 2-D image variant
// (src/multimeditron/model/data_loader.py:155-188) built on the device from input_ids + attention_mask.
// One CTA per row: two block scans (mask count, image starts), a short serial pass over the row's images
// (each image shifts everything behind it), one write pass.  Integer work, bit-exact.
// ---------------------------------------------------------------------------------------------
namespace mm {

constexpr int kPosThreads = 512;
constexpr int kPosMaxImages = 256;  // images per row

__device__ __forceinline__ int block_exclusive_scan(int v, int* warp_sums, int& total) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int inc = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int t = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += t;
  }
  __syncthreads();  // warp_sums may still be read from a previous scan
  if (lane == 31) warp_sums[warp] = inc;
  __syncthreads();
  int base = 0, tot = 0;
  for (int w = 0; w < kPosThreads / 32; ++w) {
    const int s = warp_sums[w];
    if (w < warp) base += s;
    tot += s;
  }
  total = tot;
  return base + inc - v;
}

template <typename MaskT>
__global__ void __launch_bounds__(kPosThreads) position_ids_kernel(const int64_t* __restrict__ ids,
                                                                   const MaskT* __restrict__ mask, int64_t* __restrict__ out,
                                                                   int S, int64_t att, int P, int g, int use_2d,
                                                                   int32_t* status) {
  __shared__ int warp_sums[kPosThreads / 32];
  __shared__ int s_start[kPosMaxImages];
  __shared__ long long s_base[kPosMaxImages], s_shift[kPosMaxImages];
  __shared__ int s_nimg;
  const int b = blockIdx.x;
  const int64_t* rid = ids + (int64_t)b * S;
  const MaskT* rm = mask + (int64_t)b * S;
  const int chunk = (S + kPosThreads - 1) / kPosThreads;
  const int c0 = min(threadIdx.x * chunk, S), c1 = min(c0 + chunk, S);
  // An image starts at every P-th attachment token of the row — the reference's own rule (compute_token_range,
  // prompt_tokenizers.py:255-286: "the k-th modality starts at the (sum of the previous lengths)-th attachment token"),
  // so two images that touch without delimiters are still two images.
  int n_mask = 0, n_att = 0;
  for (int t = c0; t < c1; ++t) {
    n_mask += rm[t] != 0;
    if (use_2d) n_att += rid[t] == att;
  }
  int tot_mask, tot_start = 0, tot_att = 0;
  const int pre_mask = block_exclusive_scan(n_mask, warp_sums, tot_mask);
  if (!use_2d) {
    int run = pre_mask;
    for (int t = c0; t < c1; ++t) {
      const bool m = rm[t] != 0;
      run += m;
      out[(int64_t)b * S + t] = m ? run - 1 : 0;
    }
    return;
  }
  __syncthreads();
  const int pre_att = block_exclusive_scan(n_att, warp_sums, tot_att);
  int n_start = 0;
  {
    int a = pre_att;
    for (int t = c0; t < c1; ++t)
      if (rid[t] == att) n_start += (a++ % P) == 0;
  }
  __syncthreads();
  const int pre_start = block_exclusive_scan(n_start, warp_sums, tot_start);
  int64_t* ro = out + (int64_t)b * S * 2;
  {  // pass 1: 1-D ids (kept in component 0 for the serial pass) and the ordered list of image starts
    int run = pre_mask, k = pre_start, a = pre_att;
    for (int t = c0; t < c1; ++t) {
      const bool m = rm[t] != 0;
      run += m;
      ro[2 * (int64_t)t] = m ? run - 1 : 0;
      if (rid[t] == att && (a++ % P) == 0) {
        if (k < kPosMaxImages) s_start[k] = t;
        ++k;
      }
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    int n = tot_start;
    if (n > kPosMaxImages) {
      n = kPosMaxImages;
      if (status) atomicOr(status, MM_STATUS_COUNT_MISMATCH);
    }
    long long shift = 0;
    for (int k = 0; k < n; ++k) {
      const int t0 = s_start[k];
      int t1 = t0 + P;
      if (t1 > S) {  // an image block that does not fit the row: data error, reported like a count mismatch
        t1 = S;
        if (status) atomicOr(status, MM_STATUS_COUNT_MISMATCH);
      }
      const long long base = ro[2 * (int64_t)t0] + shift;                            // position at the first token (:179)
      const long long old_last = t0 > 0 ? ro[2 * (int64_t)(t1 - 1)] + shift : 0;      // :178 (0 when the row starts with it)
      const long long next_last = base + (g - 1);                                     // max of the last grid id (:180)
      shift += next_last - old_last;                                                  // :186
      s_base[k] = base;
      s_shift[k] = shift;
    }
    s_nimg = n;
  }
  __syncthreads();
  {  // pass 2: final ids
    const int n = s_nimg;
    int run = pre_mask, k = min(pre_start, n);
    for (int t = c0; t < c1; ++t) {
      const bool m = rm[t] != 0;
      run += m;
      while (k < n && s_start[k] <= t) ++k;   // images started at or before t
      long long v0, v1;
      if (k > 0 && t < s_start[k - 1] + P) {
        const int p = t - s_start[k - 1];
        v0 = s_base[k - 1] + p / g, v1 = s_base[k - 1] + p % g;
      } else {
        v0 = v1 = (m ? run - 1 : 0) + (k > 0 ? s_shift[k - 1] : 0);
      }
      ro[2 * (int64_t)t] = v0;
      ro[2 * (int64_t)t + 1] = v1;
    }
  }
}

}  // namespace mm

extern "C" int mm_position_ids(const int64_t* input_ids, const void* attention_mask, int mask_is_u8, int64_t batch,
                               int64_t seq, int64_t attachment_id, int64_t tokens_per_image, int64_t grid, int use_2d,
                               int64_t* out, int32_t* status, void* stream) {
  using namespace mm;
  if (batch < 0 || seq < 0 || seq > INT32_MAX / 4) return MM_ERR_BAD_ARG;
  if (batch == 0 || seq == 0) return MM_OK;
  if (!attention_mask || !out || (use_2d && (!input_ids || grid <= 0 || tokens_per_image != grid * grid)))
    return MM_ERR_BAD_ARG;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (mask_is_u8)
    position_ids_kernel<uint8_t><<<(unsigned)batch, kPosThreads, 0, st>>>(
        input_ids, reinterpret_cast<const uint8_t*>(attention_mask), out, (int)seq, attachment_id, (int)tokens_per_image,
        (int)grid, use_2d, status);
  else
    position_ids_kernel<int64_t><<<(unsigned)batch, kPosThreads, 0, st>>>(
        input_ids, reinterpret_cast<const int64_t*>(attention_mask), out, (int)seq, attachment_id, (int)tokens_per_image,
        (int)grid, use_2d, status);
  MM_CUDA_CHECK(cudaGetLastError());
  return MM_OK;
}
