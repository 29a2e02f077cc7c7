// K1: adjacent motif-pair frequency counting + argmax.
// Replaces BPE.bin (bpe.py:1431-1474), the per-step recount (bpe.py:2077-2138) and the
// SortedDict.peekitem(0) pop (bpe.py:1796) by a full recount every iteration:
//   packed 64-bit pair keys -> warp-aggregated (match.any) insertion into a per-CTA shared-memory
//   hash table -> one global atomic per distinct key per CTA -> table scan for the max count and the
//   tie list.  HBM-bound: cls + glue = 8 B per live token (medoid mode), hsh+glue+len = 16 B (histogram mode),
//   read once with 128-bit coalesced loads.
#include "common.cuh"
#include "k1_keys.cuh"

#define K1_THREADS 512
#define K1_TILE (K1_THREADS * K1_VEC)
#define K1_SLOTS 8192           // direct-mapped shared-memory cache per CTA: u64 key + u32 count = 96 KB

struct GTable { u64* k; u32* c; u64 mask; int32_t* overflow; };

// count-only insertion into the global (L2-resident) table: plain load first, CAS only on an empty slot,
// then one RED on the count
__device__ __forceinline__ void gtable_count(const GTable& g, u64 key, u32 c) {
  u64 s = gb_mix64(key) & g.mask;
#pragma unroll 1
  for (int probe = 0; probe < GB_MAX_PROBES; ++probe) {
    u64 prev = __ldcg(g.k + s);
    if (prev == GB_EMPTY_KEY) prev = atomicCAS(g.k + s, (u64)GB_EMPTY_KEY, key);
    if (prev == GB_EMPTY_KEY || prev == key) { atomicAdd(g.c + s, c); return; }
    s = (s + 1) & g.mask;
  }
  atomicExch(g.overflow, 1);
}

struct SCache { u64* k; u32* c; };

// direct-mapped, first-come: a key that finds its slot taken by another key goes straight to the global table
// (no probing -> no divergent probe loops; hot keys claim their slots early and absorb their repeats)
__device__ __forceinline__ void scache_count(const SCache& s, const GTable& g, u64 key, u32 c) {
  u32 slot = (u32)(gb_mix64(key) >> 24) & (K1_SLOTS - 1);
  u64 prev = s.k[slot];
  if (prev == GB_EMPTY_KEY) prev = atomicCAS(&s.k[slot], (u64)GB_EMPTY_KEY, key);
  if (prev == GB_EMPTY_KEY || prev == key) atomicAdd(&s.c[slot], c);
  else gtable_count(g, key, c);
}

// warp-aggregated count of one key per lane (key == EMPTY -> lane does not contribute)
template <bool USE_SMEM>
__device__ __forceinline__ void warp_agg_count(const SCache& s, const GTable& g, u64 key) {
  unsigned grp = __match_any_sync(0xffffffffu, key);
  if (key != GB_EMPTY_KEY && (unsigned)(__ffs(grp) - 1) == (threadIdx.x & 31)) {
    u32 c = __popc(grp);
    if (USE_SMEM) scache_count(s, g, key, c);
    else gtable_count(g, key, c);
  }
}

struct KeysArr {
  const u64* k;
  __device__ __forceinline__ void keys(i64 j0, i64 T, u64 out[K1_VEC]) const {
#pragma unroll
    for (int e = 0; e < K1_VEC; ++e) out[e] = (j0 + e < T) ? k[j0 + e] : GB_EMPTY_KEY;
  }
};

template <class KF, bool USE_SMEM>
__global__ void __launch_bounds__(K1_THREADS, 2) k_pair_count(KF kf, i64 T, GTable g) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  SCache s;
  s.k = reinterpret_cast<u64*>(smem_raw);
  s.c = reinterpret_cast<u32*>(smem_raw + sizeof(u64) * K1_SLOTS);
  if (USE_SMEM) {
    for (int i = threadIdx.x; i < K1_SLOTS; i += blockDim.x) { s.k[i] = GB_EMPTY_KEY; s.c[i] = 0; }
    __syncthreads();
  }
  for (i64 base = (i64)blockIdx.x * K1_TILE; base < T; base += (i64)gridDim.x * K1_TILE) {
    i64 j0 = base + (i64)threadIdx.x * K1_VEC;
    u64 key[K1_VEC];
    kf.keys(j0, T, key);
#pragma unroll
    for (int e = 0; e < K1_VEC; ++e) warp_agg_count<USE_SMEM>(s, g, key[e]);
  }
  if (USE_SMEM) {
    __syncthreads();
    for (int i = threadIdx.x; i < K1_SLOTS; i += blockDim.x) {
      u64 key = s.k[i];
      if (key != GB_EMPTY_KEY) gtable_count(g, key, s.c[i]);
    }
  }
}

static const size_t K1_SMEM = (sizeof(u64) + sizeof(u32)) * K1_SLOTS;
#define K1_SMEM_MIN_PER_CTA 32768   // pairs per CTA below which the shared cache only adds a flush

template <class KF>
static int launch_pair_count(KF kf, i64 T, i64 min_T, u64* tkeys, u32* tcnt, i64 M, int32_t* overflow, cudaStream_t st) {
  if (M <= 0 || (M & (M - 1))) { gb_set_error("pair_count: table size %lld is not a power of two", (long long)M); return -2; }
  if (T < min_T) return 0;
  static bool attr_set = false;
  if (!attr_set) {
    GB_CK(cudaFuncSetAttribute(k_pair_count<KF, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)K1_SMEM));
    attr_set = true;
  }
  int sms = gb_sm_count();
  if (sms <= 0) sms = 148;
  int blocks = gb_blocks_for(T, K1_TILE, sms * 2);   // persistent: 2 CTAs per SM
  GTable g{tkeys, tcnt, (u64)(M - 1), overflow};
  if (T / blocks >= K1_SMEM_MIN_PER_CTA) k_pair_count<KF, true><<<blocks, K1_THREADS, K1_SMEM, st>>>(kf, T, g);
  else k_pair_count<KF, false><<<blocks, K1_THREADS, 0, st>>>(kf, T, g);
  GB_LAUNCH_CK();
  return 0;
}

extern "C" int gb_pair_count_m(const uint32_t* cls, const uint32_t* glue, int64_t T, uint64_t* tkeys,
                               uint32_t* tcnt, uint32_t* tmin, int64_t M, int32_t* overflow, void* stream) {
  KeysM kf{cls, glue};
  return launch_pair_count(kf, (i64)T, 2, (u64*)tkeys, tcnt, (i64)M, overflow, (cudaStream_t)stream);
}
extern "C" int gb_pair_count_h(const uint64_t* hsh, const uint32_t* glue, const uint32_t* len, int64_t T,
                               const uint64_t* powtab, uint64_t P, uint64_t* tkeys, uint32_t* tcnt,
                               uint32_t* tmin, int64_t M, int32_t* overflow, void* stream) {
  KeysH kf{(const u64*)hsh, glue, len, (const u64*)powtab, (u64)P};
  return launch_pair_count(kf, (i64)T, 2, (u64*)tkeys, tcnt, (i64)M, overflow, (cudaStream_t)stream);
}
__global__ void __launch_bounds__(256) k_key_count_min(const u64* __restrict__ keys, i64 n, u64* tk, u32* tc, u32* tm,
                                                        u64 mask, int32_t* overflow) {
  for (i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (i64)gridDim.x * blockDim.x)
    if (!gb_gtable_add(tk, tc, tm, mask, keys[i], 1u, (u32)i)) atomicExch(overflow, 1);
}
extern "C" int gb_key_count(const uint64_t* keys, int64_t n, uint64_t* tkeys, uint32_t* tcnt, uint32_t* tmin,
                            int64_t M, int32_t* overflow, void* stream) {
  if (M <= 0 || (M & (M - 1))) { gb_set_error("gb_key_count: table size is not a power of two"); return -2; }
  if (n <= 0) return 0;
  k_key_count_min<<<gb_blocks_for(n, 256 * 4, 148 * 8), 256, 0, (cudaStream_t)stream>>>((const u64*)keys, n, (u64*)tkeys,
                                                                                      tcnt, tmin, (u64)(M - 1), overflow);
  GB_LAUNCH_CK();
  return 0;
}

// ---- first position of every key tied at the max (example occurrence for the host's json tie-break) ----
#define FP_LIST 64
template <class KF>
__global__ void __launch_bounds__(K1_THREADS) k_first_pos(KF kf, i64 T, const u64* __restrict__ tk, const u32* __restrict__ tc,
                                                          u32* __restrict__ tm, u64 mask, const u32* __restrict__ out4,
                                                          const u64* __restrict__ tie_keys, u32* __restrict__ tie_min) {
  __shared__ u64 lk[FP_LIST];
  const u32 mx = out4[0], nt = out4[1];
  if (mx == 0 || nt == 0) return;
  const bool small = nt <= FP_LIST;
  if (small && threadIdx.x < nt) lk[threadIdx.x] = tie_keys[threadIdx.x];
  __syncthreads();
  for (i64 base = (i64)blockIdx.x * K1_TILE; base < T; base += (i64)gridDim.x * K1_TILE) {
    i64 j0 = base + (i64)threadIdx.x * K1_VEC;
    u64 key[K1_VEC];
    kf.keys(j0, T, key);
#pragma unroll
    for (int e = 0; e < K1_VEC; ++e) {
      if (key[e] == GB_EMPTY_KEY) continue;
      if (small) {
        for (u32 i = 0; i < nt; ++i)
          if (key[e] == lk[i]) { atomicMin(&tie_min[i], (u32)(j0 + e)); break; }
      } else {
        u64 s = gb_mix64(key[e]) & mask;
        for (int p = 0; p < GB_MAX_PROBES; ++p) {
          u64 k = tk[s];
          if (k == key[e]) { if (tc[s] == mx) atomicMin(&tm[s], (u32)(j0 + e)); break; }
          if (k == GB_EMPTY_KEY) break;
          s = (s + 1) & mask;
        }
      }
    }
  }
}
__global__ void k_ties_fill(const u64* __restrict__ tk, const u32* __restrict__ tm, u64 mask, const u32* __restrict__ out4,
                            const u64* __restrict__ tie_keys, u32* __restrict__ tie_min, int cap) {
  const u32 nt = out4[1];
  if (nt <= FP_LIST) return;
  u32 n = nt < (u32)cap ? nt : (u32)cap;
  for (u32 i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    u64 key = tie_keys[i];
    u64 s = gb_mix64(key) & mask;
    for (int p = 0; p < GB_MAX_PROBES; ++p) {
      u64 k = tk[s];
      if (k == key) { tie_min[i] = tm[s]; break; }
      if (k == GB_EMPTY_KEY) break;
      s = (s + 1) & mask;
    }
  }
}
template <class KF>
static int launch_first_pos(KF kf, i64 T, const u64* tk, const u32* tc, u32* tm, i64 M, const u32* out4, const u64* tie_keys,
                            u32* tie_min, int cap, cudaStream_t st) {
  if (T < 2) return 0;
  int sms = gb_sm_count();
  if (sms <= 0) sms = 148;
  int blocks = gb_blocks_for(T, K1_TILE, sms * 2);
  k_first_pos<KF><<<blocks, K1_THREADS, 0, st>>>(kf, T, tk, tc, tm, (u64)(M - 1), out4, tie_keys, tie_min);
  GB_LAUNCH_CK();
  k_ties_fill<<<64, 256, 0, st>>>(tk, tm, (u64)(M - 1), out4, tie_keys, tie_min, cap);
  GB_LAUNCH_CK();
  return 0;
}
extern "C" int gb_first_positions_m(const uint32_t* cls, const uint32_t* glue, int64_t T, const uint64_t* tkeys,
                                    const uint32_t* tcnt, uint32_t* tmin, int64_t M, const uint32_t* out4,
                                    const uint64_t* tie_keys, uint32_t* tie_minpos, int cap, void* stream) {
  KeysM kf{cls, glue};
  return launch_first_pos(kf, (i64)T, (const u64*)tkeys, tcnt, tmin, (i64)M, out4, (const u64*)tie_keys, tie_minpos, cap,
                          (cudaStream_t)stream);
}
extern "C" int gb_first_positions_h(const uint64_t* hsh, const uint32_t* glue, const uint32_t* len, int64_t T,
                                    const uint64_t* powtab, uint64_t P, const uint64_t* tkeys, const uint32_t* tcnt,
                                    uint32_t* tmin, int64_t M, const uint32_t* out4, const uint64_t* tie_keys,
                                    uint32_t* tie_minpos, int cap, void* stream) {
  KeysH kf{(const u64*)hsh, glue, len, (const u64*)powtab, (u64)P};
  return launch_first_pos(kf, (i64)T, (const u64*)tkeys, tcnt, tmin, (i64)M, out4, (const u64*)tie_keys, tie_minpos, cap,
                          (cudaStream_t)stream);
}

// ---- table maintenance ---------------------------------------------------------------------------
__global__ void k_table_clear(u64* __restrict__ tk, u32* __restrict__ tc, u32* __restrict__ tm, i64 M) {
  for (i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x; i < M; i += (i64)gridDim.x * blockDim.x) {
    tk[i] = GB_EMPTY_KEY; tc[i] = 0u; tm[i] = 0xFFFFFFFFu;
  }
}
extern "C" int gb_table_clear(uint64_t* tkeys, uint32_t* tcnt, uint32_t* tmin, int64_t M, void* stream) {
  if (M <= 0) return 0;
  k_table_clear<<<gb_blocks_for(M, 256 * 4, 148 * 8), 256, 0, (cudaStream_t)stream>>>((u64*)tkeys, tcnt, tmin, M);
  GB_LAUNCH_CK();
  return 0;
}

// out4 = {max count, #ties, #distinct, 0}
__global__ void __launch_bounds__(256) k_table_max(const u32* __restrict__ tc, i64 M, u32* __restrict__ out4) {
  u32 mx = 0, nd = 0;
  for (i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x; i < M; i += (i64)gridDim.x * blockDim.x) {
    u32 c = tc[i];
    mx = c > mx ? c : mx;
    nd += c != 0;
  }
  for (int o = 16; o; o >>= 1) {
    u32 a = __shfl_xor_sync(0xffffffffu, mx, o); mx = a > mx ? a : mx;
    nd += __shfl_xor_sync(0xffffffffu, nd, o);
  }
  if ((threadIdx.x & 31) == 0) { atomicMax(&out4[0], mx); atomicAdd(&out4[2], nd); }
}
__global__ void __launch_bounds__(256) k_table_ties(const u64* __restrict__ tk, const u32* __restrict__ tc,
                                                     const u32* __restrict__ tm, i64 M, u32* __restrict__ out4,
                                                     u64* __restrict__ tie_keys, u32* __restrict__ tie_min, int cap) {
  const u32 mx = out4[0];
  if (mx == 0) return;
  for (i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x; i < M; i += (i64)gridDim.x * blockDim.x) {
    if (tc[i] == mx) {
      u32 k = atomicAdd(&out4[1], 1u);
      if (k < (u32)cap) { tie_keys[k] = tk[i]; tie_min[k] = tm[i]; }   // tm = first position when the insertion tracked it, else 0xFFFFFFFF
    }
  }
}
extern "C" int gb_table_argmax(const uint64_t* tkeys, const uint32_t* tcnt, const uint32_t* tmin, int64_t M,
                               uint32_t* out4, uint64_t* tie_keys, uint32_t* tie_minpos, int cap, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  GB_CK(cudaMemsetAsync(out4, 0, 4 * sizeof(u32), st));
  if (M <= 0) return 0;
  int blocks = gb_blocks_for(M, 256 * 4, 148 * 8);
  k_table_max<<<blocks, 256, 0, st>>>(tcnt, M, out4);
  GB_LAUNCH_CK();
  k_table_ties<<<blocks, 256, 0, st>>>((const u64*)tkeys, tcnt, tmin, M, out4, (u64*)tie_keys, tie_minpos, cap);
  GB_LAUNCH_CK();
  return 0;
}

__global__ void k_table_lookup(const u64* __restrict__ tk, const u32* __restrict__ tc, const u32* __restrict__ tm,
                               u64 mask, const u64* __restrict__ q, int Q, u32* __restrict__ oc, u32* __restrict__ om) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= Q) return;
  u64 key = q[i];
  u64 s = gb_mix64(key) & mask;
  u32 c = 0, m = 0xFFFFFFFFu;
  for (u64 p = 0; p <= mask; ++p) {
    u64 k = tk[s];
    if (k == key) { c = tc[s]; m = tm[s]; break; }
    if (k == GB_EMPTY_KEY) break;
    s = (s + 1) & mask;
  }
  oc[i] = c; om[i] = m;
}
extern "C" int gb_table_lookup(const uint64_t* tkeys, const uint32_t* tcnt, const uint32_t* tmin, int64_t M,
                               const uint64_t* qkeys, int Q, uint32_t* out_cnt, uint32_t* out_min, void* stream) {
  if (Q <= 0) return 0;
  k_table_lookup<<<(Q + 127) / 128, 128, 0, (cudaStream_t)stream>>>((const u64*)tkeys, tcnt, tmin, (u64)(M - 1),
                                                                    (const u64*)qkeys, Q, out_cnt, out_min);
  GB_LAUNCH_CK();
  return 0;
}

__global__ void __launch_bounds__(256) k_table_dump(const u64* __restrict__ tk, const u32* __restrict__ tc,
                                                     const u32* __restrict__ tm, i64 M, u64* __restrict__ ok,
                                                     u32* __restrict__ oc, u32* __restrict__ om, u32* __restrict__ total) {
  for (i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x; i < M; i += (i64)gridDim.x * blockDim.x) {
    u32 c = tc[i];
    if (c) {
      u32 k = atomicAdd(total, 1u);
      ok[k] = tk[i]; oc[k] = c; om[k] = tm[i];
    }
  }
}
extern "C" int gb_table_dump(const uint64_t* tkeys, const uint32_t* tcnt, const uint32_t* tmin, int64_t M,
                             uint64_t* out_keys, uint32_t* out_cnt, uint32_t* out_min, uint32_t* total, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  GB_CK(cudaMemsetAsync(total, 0, sizeof(u32), st));
  if (M <= 0) return 0;
  k_table_dump<<<gb_blocks_for(M, 256 * 4, 148 * 8), 256, 0, st>>>((const u64*)tkeys, tcnt, tmin, M, (u64*)out_keys,
                                                                   out_cnt, out_min, total);
  GB_LAUNCH_CK();
  return 0;
}

// fold the exact local table into key-less direct-mapped counters for the NCCL allreduce
__global__ void __launch_bounds__(256) k_table_to_slots(const u64* __restrict__ tk, const u32* __restrict__ tc, i64 M,
                                                         u32* __restrict__ slots, u64 smask) {
  for (i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x; i < M; i += (i64)gridDim.x * blockDim.x) {
    u32 c = tc[i];
    if (c) atomicAdd(&slots[(gb_mix64(tk[i] ^ 0x9E3779B97F4A7C15ull)) & smask], c);
  }
}
extern "C" int gb_table_to_slots(const uint64_t* tkeys, const uint32_t* tcnt, int64_t M, uint32_t* slots,
                                 int64_t NS, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  if (NS <= 0 || (NS & (NS - 1))) { gb_set_error("gb_table_to_slots: NS must be a power of two"); return -2; }
  GB_CK(cudaMemsetAsync(slots, 0, sizeof(u32) * (size_t)NS, st));
  if (M <= 0) return 0;
  k_table_to_slots<<<gb_blocks_for(M, 256 * 4, 148 * 8), 256, 0, st>>>((const u64*)tkeys, tcnt, M, slots, (u64)(NS - 1));
  GB_LAUNCH_CK();
  return 0;
}

// multi-GPU: this rank's (key, count) entries whose direct-mapped slot is marked (top slots after the allreduce)
__global__ void __launch_bounds__(256) k_table_dump_marked(const u64* __restrict__ tk, const u32* __restrict__ tc, i64 M,
                                                            const u8* __restrict__ mark, u64 smask, u64* __restrict__ ok,
                                                            u32* __restrict__ oc, u32* __restrict__ total, u32 cap) {
  for (i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x; i < M; i += (i64)gridDim.x * blockDim.x) {
    u32 c = tc[i];
    if (c) {
      u64 key = tk[i];
      if (mark[gb_mix64(key ^ 0x9E3779B97F4A7C15ull) & smask]) {
        u32 k = atomicAdd(total, 1u);
        if (k < cap) { ok[k] = key; oc[k] = c; }
      }
    }
  }
}
extern "C" int gb_table_dump_marked(const uint64_t* tkeys, const uint32_t* tcnt, int64_t M, const uint8_t* slotmark,
                                    int64_t NS, uint64_t* out_keys, uint32_t* out_cnt, uint32_t* total, int cap,
                                    void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  if (NS <= 0 || (NS & (NS - 1))) { gb_set_error("gb_table_dump_marked: NS must be a power of two"); return -2; }
  GB_CK(cudaMemsetAsync(total, 0, sizeof(u32), st));
  if (M <= 0) return 0;
  k_table_dump_marked<<<gb_blocks_for(M, 256 * 4, 148 * 8), 256, 0, st>>>((const u64*)tkeys, tcnt, M, slotmark, (u64)(NS - 1),
                                                                          (u64*)out_keys, out_cnt, total, (u32)cap);
  GB_LAUNCH_CK();
  return 0;
}

// one small result record per iteration so the host needs a single D2H: res i64 [5 + 4*nfirst] =
// {max, #ties, #distinct, 0, overflow, then for the first nfirst ties: key, min position j, pos[j], len[j]+len[j+1]}
__global__ void k_pack_result(const u32* __restrict__ out4, const int32_t* __restrict__ overflow,
                              const u64* __restrict__ tie_keys, const u32* __restrict__ tie_min,
                              const u32* __restrict__ pos, const u32* __restrict__ len, int nfirst,
                              const u32* __restrict__ sout, long long* __restrict__ res) {
  int i = threadIdx.x;
  if (i < 4) res[i] = out4[i];
  if (i == 3 && sout) res[3] = (long long)(((u64)sout[1] << 32) | sout[0]);
  if (i == 4) res[4] = *overflow;
  if (i < nfirst) {
    long long* r = res + 5 + 4 * i;
    if ((u32)i < out4[1]) {
      u32 m = tie_min[i];
      r[0] = (long long)tie_keys[i]; r[1] = m;
      bool ok = pos != nullptr && m != 0xFFFFFFFFu;
      r[2] = ok ? pos[m] : 0; r[3] = ok ? (long long)len[m] + len[m + 1] : 0;
    } else { r[0] = r[1] = r[2] = r[3] = 0; }
  }
}
extern "C" int gb_pack_result(const uint32_t* out4, const int32_t* overflow, const uint64_t* tie_keys,
                              const uint32_t* tie_minpos, const uint32_t* pos, const uint32_t* len, int nfirst,
                              const uint32_t* sout, int64_t* res, void* stream) {
  if (nfirst < 0 || nfirst > 64) { gb_set_error("gb_pack_result: nfirst must be in [0, 64]"); return -2; }
  k_pack_result<<<1, 64, 0, (cudaStream_t)stream>>>(out4, overflow, (const u64*)tie_keys, tie_minpos, pos, len, nfirst,
                                                   sout, (long long*)res);
  GB_LAUNCH_CK();
  return 0;
}
