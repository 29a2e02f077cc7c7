// Ordered stream compaction (three-phase: per-block counts -> scan of block sums -> scatter) and
// the token-array merge compaction of K2 (bpe.py:1955-1970 token_pos / bond_to_token update).
#include "common.cuh"
#include <stdarg.h>
#include <mutex>

static thread_local char g_err[512] = "";
void gb_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
extern "C" const char* gb_last_error(void) { return g_err; }
extern "C" int gb_abi_version(void) { return 1; }
extern "C" int gb_sm_count(void) {
  int dev = 0, n = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return -1;
  if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return -1;
  return n;
}

#define SEL_THREADS 256
#define SEL_ITEMS 8
#define SEL_CHUNK (SEL_THREADS * SEL_ITEMS)

extern "C" int64_t gb_select_scratch_elems(int64_t n) { return (n + SEL_CHUNK - 1) / SEL_CHUNK + 2; }

struct PredFlag {
  const u8* f; u32 mask;
  __device__ __forceinline__ bool operator()(i64 i) const { return (f[i] & mask) != 0; }
};
struct PredEq64 {
  const u64* k; u64 key;
  __device__ __forceinline__ bool operator()(i64 i) const { return k[i] == key; }
};
// token i survives a merge step unless pair (i-1, i) was kept (flags bit1 on i-1)
struct PredSurvive {
  const u8* f;
  __device__ __forceinline__ bool operator()(i64 i) const { return !(i > 0 && (f[i - 1] & 2)); }
};

template <class Pred>
__global__ void __launch_bounds__(SEL_THREADS) k_sel_count(Pred pred, i64 n, u32* __restrict__ bsum) {
  __shared__ u32 wsum[SEL_THREADS / 32];
  i64 base = (i64)blockIdx.x * SEL_CHUNK;
  u32 c = 0;
#pragma unroll
  for (int it = 0; it < SEL_ITEMS; ++it) {
    i64 i = base + it * SEL_THREADS + threadIdx.x;
    if (i < n && pred(i)) ++c;
  }
  for (int o = 16; o; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
  if ((threadIdx.x & 31) == 0) wsum[threadIdx.x >> 5] = c;
  __syncthreads();
  if (threadIdx.x == 0) {
    u32 t = 0;
    for (int w = 0; w < SEL_THREADS / 32; ++w) t += wsum[w];
    bsum[blockIdx.x] = t;
  }
}

// in-place exclusive scan of bsum[0..nb), total -> *total (single block)
__global__ void __launch_bounds__(1024) k_sel_scan(u32* __restrict__ bsum, i64 nb, u32* __restrict__ total) {
  __shared__ u32 wtot[32];
  __shared__ u32 carry_s;
  if (threadIdx.x == 0) carry_s = 0;
  __syncthreads();
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (i64 base = 0; base < nb; base += 1024) {
    i64 i = base + threadIdx.x;
    u32 v = (i < nb) ? bsum[i] : 0u;
    u32 x = v;  // inclusive warp scan
    for (int o = 1; o < 32; o <<= 1) {
      u32 y = __shfl_up_sync(0xffffffffu, x, o);
      if (lane >= (unsigned)o) x += y;
    }
    if (lane == 31) wtot[warp] = x;
    __syncthreads();
    u32 woff = 0, tot = 0;
    for (unsigned w = 0; w < 32; ++w) {
      u32 t = wtot[w];
      if (w < warp) woff += t;
      tot += t;
    }
    u32 carry = carry_s;
    if (i < nb) bsum[i] = carry + woff + x - v;
    __syncthreads();
    if (threadIdx.x == 0) carry_s = carry + tot;
    __syncthreads();
  }
  if (threadIdx.x == 0) *total = carry_s;
}

template <class Pred>
__global__ void __launch_bounds__(SEL_THREADS) k_sel_scatter_idx(Pred pred, i64 n, const u32* __restrict__ boff,
                                                                  u32* __restrict__ out_idx) {
  __shared__ u32 wsum[SEL_THREADS / 32];
  i64 base = (i64)blockIdx.x * SEL_CHUNK;
  u32 run = boff[blockIdx.x];
#pragma unroll
  for (int it = 0; it < SEL_ITEMS; ++it) {
    i64 i = base + it * SEL_THREADS + threadIdx.x;
    bool f = (i < n) && pred(i);
    u32 tot;
    u32 ex = gb_block_excl_scan_flag(f, wsum, tot);
    if (f) out_idx[run + ex] = (u32)i;
    run += tot;
  }
}

template <class Pred>
static int select_impl(Pred pred, i64 n, u32* out_idx, u32* total, u32* scratch, cudaStream_t st) {
  i64 nb = (n + SEL_CHUNK - 1) / SEL_CHUNK;
  if (n <= 0) {
    GB_CK(cudaMemsetAsync(total, 0, sizeof(u32), st));
    return 0;
  }
  k_sel_count<<<(unsigned)nb, SEL_THREADS, 0, st>>>(pred, n, scratch);
  GB_LAUNCH_CK();
  k_sel_scan<<<1, 1024, 0, st>>>(scratch, nb, total);
  GB_LAUNCH_CK();
  k_sel_scatter_idx<<<(unsigned)nb, SEL_THREADS, 0, st>>>(pred, n, scratch, out_idx);
  GB_LAUNCH_CK();
  return 0;
}

extern "C" int gb_select_flagged(const uint8_t* flags, int64_t n, uint32_t mask, uint32_t* out_idx,
                                 uint32_t* total, uint32_t* scratch, void* stream) {
  PredFlag p{flags, mask};
  return select_impl(p, n, out_idx, total, scratch, (cudaStream_t)stream);
}

extern "C" int gb_select_eq_u64(const uint64_t* keys, int64_t n, uint64_t key, uint32_t* out_idx,
                                uint32_t* total, uint32_t* scratch, void* stream) {
  PredEq64 p{(const u64*)keys, (u64)key};
  return select_impl(p, n, out_idx, total, scratch, (cudaStream_t)stream);
}

// ---- merge compaction --------------------------------------------------------------------------
struct MergeArgs {
  const u8* flags;
  const u32* cls; const u64* hsh; const u32* glue; const u32* pos; const u32* len; const u32* tid;
  u32* cls_o; u64* hsh_o; u32* glue_o; u32* pos_o; u32* len_o; u32* tid_o;
  const u32* aux; const u32* p2cls; const u32* p2tid;
  u64 key; u32 new_tid;
};

__global__ void __launch_bounds__(SEL_THREADS) k_merge_scatter(MergeArgs a, i64 n, const u32* __restrict__ boff) {
  __shared__ u32 wsum[SEL_THREADS / 32];
  PredSurvive pred{a.flags};
  i64 base = (i64)blockIdx.x * SEL_CHUNK;
  u32 run = boff[blockIdx.x];
#pragma unroll
  for (int it = 0; it < SEL_ITEMS; ++it) {
    i64 i = base + it * SEL_THREADS + threadIdx.x;
    bool f = (i < n) && pred(i);
    u32 tot;
    u32 ex = gb_block_excl_scan_flag(f, wsum, tot);
    if (f) {
      u32 o = run + ex;
      bool merged = (a.flags[i] & 2) != 0;  // token i absorbs token i+1
      a.pos_o[o] = a.pos[i];
      if (merged) {
        a.glue_o[o] = a.glue[i + 1];
        a.len_o[o] = a.len[i] + a.len[i + 1];
        if (a.aux) {
          u32 p = a.aux[i];
          if (a.cls_o) a.cls_o[o] = a.p2cls[p];
          a.tid_o[o] = a.p2tid[p];
        } else {
          if (a.cls_o) a.cls_o[o] = a.cls[i];
          a.tid_o[o] = a.new_tid;
        }
        if (a.hsh_o) a.hsh_o[o] = a.key;
      } else {
        a.glue_o[o] = a.glue[i];
        a.len_o[o] = a.len[i];
        a.tid_o[o] = a.tid[i];
        if (a.cls_o) a.cls_o[o] = a.cls[i];
        if (a.hsh_o) a.hsh_o[o] = a.hsh[i];
      }
    }
    run += tot;
  }
}

extern "C" int gb_merge_compact(const uint8_t* flags, int64_t T, const uint32_t* cls, const uint64_t* hsh,
                                const uint32_t* glue, const uint32_t* pos, const uint32_t* len,
                                const uint32_t* tid, uint32_t* cls_o, uint64_t* hsh_o, uint32_t* glue_o,
                                uint32_t* pos_o, uint32_t* len_o, uint32_t* tid_o, const uint32_t* aux,
                                const uint32_t* p2cls, const uint32_t* p2tid, uint64_t key, uint32_t new_tid,
                                uint32_t* newT, uint32_t* scratch, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  if (T <= 0) {
    GB_CK(cudaMemsetAsync(newT, 0, sizeof(u32), st));
    return 0;
  }
  if ((cls == nullptr) != (cls_o == nullptr) || (hsh == nullptr) != (hsh_o == nullptr)) {
    gb_set_error("gb_merge_compact: cls/hsh in/out must be given together");
    return -2;
  }
  if (aux && (!p2cls || !p2tid)) {
    gb_set_error("gb_merge_compact: aux needs p2cls and p2tid");
    return -2;
  }
  i64 nb = (T + SEL_CHUNK - 1) / SEL_CHUNK;
  PredSurvive pred{flags};
  k_sel_count<<<(unsigned)nb, SEL_THREADS, 0, st>>>(pred, (i64)T, scratch);
  GB_LAUNCH_CK();
  k_sel_scan<<<1, 1024, 0, st>>>(scratch, nb, newT);
  GB_LAUNCH_CK();
  MergeArgs a{flags, cls, (const u64*)hsh, glue, pos, len, tid, cls_o, (u64*)hsh_o, glue_o, pos_o, len_o, tid_o,
              aux, p2cls, p2tid, (u64)key, new_tid};
  k_merge_scatter<<<(unsigned)nb, SEL_THREADS, 0, st>>>(a, (i64)T, scratch);
  GB_LAUNCH_CK();
  return 0;
}

// ---- ordered compaction without any allocation ---------------------------------------------------
// gb_select_flagged with its block-sum scratch in a file-scope device array instead of a caller buffer, for callers that may
// not allocate (CUDA graph capture, serving loops).  No counterpart in the reference.
// The array costs GB_SELECT_STATIC_MAX_N / SEL_CHUNK + 2 words (128 KB) on every GPU that loads the library, until
// exit.  All calls share it: two calls that overlap on the GPU (different streams, no dependency) race.  Nothing in it has to
// start at zero -- the count pass overwrites words [0, nb) before the scan reads them -- so stale contents are harmless.
#define GB_SELECT_STATIC_MAX_N (1ll << 26)
#define SEL_STATIC_WORDS (GB_SELECT_STATIC_MAX_N / SEL_CHUNK + 2)
static __device__ __align__(256) u32 g_sel_static[SEL_STATIC_WORDS];

#define SEL_MAX_DEVICES 64
static u32* g_sel_static_addr[SEL_MAX_DEVICES];   // device address of g_sel_static per device (host code cannot use the name itself)
static std::mutex g_sel_static_mu;

extern "C" int64_t gb_select_flagged_static_max_n(void) { return GB_SELECT_STATIC_MAX_N; }

extern "C" int gb_select_flagged_static(const uint8_t* flags, int64_t n, uint32_t mask, uint32_t* out_idx, uint32_t* total,
                                        void* stream) {
  if (n > GB_SELECT_STATIC_MAX_N) {
    gb_set_error("gb_select_flagged_static: n = %lld exceeds the static scratch (at most %lld entries); use gb_select_flagged "
                 "with a caller buffer", (long long)n, (long long)GB_SELECT_STATIC_MAX_N);
    return -2;
  }
  int dev = 0;
  GB_CK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= SEL_MAX_DEVICES) { gb_set_error("gb_select_flagged_static: device ordinal %d out of range", dev); return -2; }
  u32* scratch;
  {
    std::lock_guard<std::mutex> lk(g_sel_static_mu);
    if (!g_sel_static_addr[dev]) {               // once per device, with that device current
      u32* p = nullptr;
      GB_CK(cudaGetSymbolAddress((void**)&p, g_sel_static));
      g_sel_static_addr[dev] = p;
    }
    scratch = g_sel_static_addr[dev];
  }
  PredFlag p{flags, mask};
  return select_impl(p, n, out_idx, total, scratch, (cudaStream_t)stream);
}
