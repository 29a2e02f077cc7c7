// K1s: adjacent motif-pair frequency counting as a key-less slot sketch + exact resolution of the top slots.
// Replaces BPE.bin (bpe.py:1431-1474), the per-step recount (bpe.py:2077-2138) and the SortedDict.peekitem(0)
// pop (bpe.py:1796).
//
//   pass 1  k_slot_count   packed 64-bit pair key -> slot = (key * phi64) >> shift -> one fire-and-forget RED.ADD
//                          into an L2-resident u32 counter array (no key compare, no CAS, no probe loop, no
//                          divergence); optional match.any warp aggregation and a per-CTA shared-memory stage for
//                          hot slots (tags are slot ids; dynamic first-come claim, or a static hot list carried over
//                          from the previous merge iteration).  HBM-bound by design: cls+glue = 8 B per live token
//                          (medoid mode), hsh+glue+len = 16 B (histogram mode), 128-bit coalesced streaming loads.
//   scan    k_slots_max / k_slots_collect   max slot count c1 and the slots holding it (or >= a threshold).
//   pass 2  k_cand_count   second streaming pass: pairs whose slot is a candidate are counted EXACTLY (key, count,
//                          first position) in a small open-addressing table.  A key outside the candidate slots has
//                          count <= its slot count < c1, so if the best exact count m1 equals c1 the exact ties are
//                          all in the table; otherwise the host asks for the slots >= m1 and runs pass 2 once more.
#include "k1_sketch.cuh"

template <class KF>
static int launch_slot_count(KF kf, i64 T, u32* slots, int log2ns, const u32* hot_tags, int variant, int ctas_per_sm,
                             cudaStream_t st) {
  if (log2ns < 10 || log2ns > 30) { gb_set_error("slot_count: log2(NS) = %d outside [10, 30]", log2ns); return -2; }
  if (T < 2) return 0;
  const bool agg = variant & 1, dry = variant & 16;
  const int cache = (variant >> 1) & 3;
  if (cache == 3 || (cache == 2 && hot_tags == nullptr)) { gb_set_error("slot_count: bad cache variant"); return -2; }
  int sms = gb_sm_count();
  if (sms <= 0) sms = 148;
  if (ctas_per_sm < 1) ctas_per_sm = 2;
  int blocks = gb_blocks_for(T, 2 * KS_TILE, sms * ctas_per_sm);
  const size_t smem = cache ? 2 * sizeof(u32) * KS_CACHE : 0;
  const int shift = 64 - log2ns;
#define KS_LAUNCH(AGG, CACHE, DRY)                                                                                 \
  do {                                                                                                             \
    if (smem) GB_CK(cudaFuncSetAttribute(k_slot_count<KF, AGG, CACHE, DRY>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
    k_slot_count<KF, AGG, CACHE, DRY><<<blocks, KS_THREADS, smem, st>>>(kf, T, slots, shift, hot_tags, nullptr);            \
  } while (0)
  if (dry) { if (agg) KS_LAUNCH(true, 0, true); else KS_LAUNCH(false, 0, true); }
  else if (cache == 0) { if (agg) KS_LAUNCH(true, 0, false); else KS_LAUNCH(false, 0, false); }
  else if (cache == 1) { if (agg) KS_LAUNCH(true, 1, false); else KS_LAUNCH(false, 1, false); }
  else { if (agg) KS_LAUNCH(true, 2, false); else KS_LAUNCH(false, 2, false); }
#undef KS_LAUNCH
  GB_LAUNCH_CK();
  return 0;
}

extern "C" int gb_slot_count_m(const uint32_t* cls, const uint32_t* glue, int64_t T, uint32_t* slots, int log2ns,
                               const uint32_t* hot_tags, int variant, int ctas_per_sm, void* stream) {
  if (variant & 8) { KeysMT<true> kf{cls, glue}; return launch_slot_count(kf, (i64)T, slots, log2ns, hot_tags, variant, ctas_per_sm, (cudaStream_t)stream); }
  KeysMT<false> kf{cls, glue};
  return launch_slot_count(kf, (i64)T, slots, log2ns, hot_tags, variant, ctas_per_sm, (cudaStream_t)stream);
}
extern "C" int gb_slot_count_h(const uint64_t* hsh, const uint32_t* glue, const uint32_t* len, int64_t T,
                               const uint64_t* powtab, uint64_t P, uint32_t* slots, int log2ns,
                               const uint32_t* hot_tags, int variant, int ctas_per_sm, void* stream) {
  if (variant & 8) {
    KeysHT<true> kf{(const u64*)hsh, glue, len, (const u64*)powtab, (u64)P};
    return launch_slot_count(kf, (i64)T, slots, log2ns, hot_tags, variant, ctas_per_sm, (cudaStream_t)stream);
  }
  KeysHT<false> kf{(const u64*)hsh, glue, len, (const u64*)powtab, (u64)P};
  return launch_slot_count(kf, (i64)T, slots, log2ns, hot_tags, variant, ctas_per_sm, (cudaStream_t)stream);
}

// ---- scan: sout u32 [4] = {max slot count c1, #candidate slots, #non-zero slots, 0} --------------------------
// One block per chunk of KS_CHUNK slots; the per-chunk maxima (bmax) let the collect pass skip every chunk that
// cannot hold a candidate, so listing the top slots costs one read of the counters, not two.
__global__ void __launch_bounds__(256) k_slots_max(const uint4* __restrict__ slots4, u32* __restrict__ bmax,
                                                    u32* __restrict__ sout) {
  __shared__ u32 s_mx[8], s_nz[8];
  const uint4* p = slots4 + (i64)blockIdx.x * (KS_CHUNK / 4);
  uint4 v[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) v[k] = p[k * 256 + threadIdx.x];
  u32 mx = 0, nz = 0;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    mx = max(max(mx, v[k].x), max(max(v[k].y, v[k].z), v[k].w));
    nz += (v[k].x != 0) + (v[k].y != 0) + (v[k].z != 0) + (v[k].w != 0);
  }
  for (int o = 16; o; o >>= 1) {
    mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    nz += __shfl_xor_sync(0xffffffffu, nz, o);
  }
  if ((threadIdx.x & 31) == 0) { s_mx[threadIdx.x >> 5] = mx; s_nz[threadIdx.x >> 5] = nz; }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < 8; ++w) { mx = max(mx, s_mx[w]); nz += s_nz[w]; }
    bmax[blockIdx.x] = mx;
    if (mx) atomicMax(&sout[0], mx);
    if (nz) atomicAdd(&sout[2], nz);
  }
}
// candidates: slots with count >= thr (thr = *thr_dev when given: the max just found); list is unordered
__global__ void __launch_bounds__(256) k_slots_collect(const uint4* __restrict__ slots4, const u32* __restrict__ bmax,
                                                        const u32* __restrict__ thr_dev, u32 thr_host,
                                                        u32* __restrict__ cand, u32* __restrict__ n_cand, u32 cap) {
  const u32 thr = thr_dev ? *thr_dev : thr_host;
  if (thr == 0 || bmax[blockIdx.x] < thr) return;
  const i64 base4 = (i64)blockIdx.x * (KS_CHUNK / 4);
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    i64 i = base4 + k * 256 + threadIdx.x;
    uint4 v = slots4[i];
    u32 a[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int e = 0; e < 4; ++e)
      if (a[e] >= thr) {
        u32 q = atomicAdd(n_cand, 1u);
        if (q < cap) cand[q] = (u32)(4 * i + e);
      }
  }
}
extern "C" int gb_slots_top(const uint32_t* slots, int log2ns, uint32_t thr_host, uint32_t* sout, uint32_t* cand,
                            int cap, uint32_t* bmax, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  if (log2ns < 12 || log2ns > 30) { gb_set_error("slots_top: log2(NS) = %d outside [12, 30]", log2ns); return -2; }
  const unsigned blocks = (unsigned)(((i64)1 << log2ns) / KS_CHUNK);
  if (thr_host == 0) {                 // find the max, then the slots holding it
    GB_CK(cudaMemsetAsync(sout, 0, 4 * sizeof(u32), st));
    k_slots_max<<<blocks, 256, 0, st>>>((const uint4*)slots, bmax, sout);
    GB_LAUNCH_CK();
    k_slots_collect<<<blocks, 256, 0, st>>>((const uint4*)slots, bmax, sout, 0u, cand, sout + 1, (u32)cap);
  } else {                             // second round: every slot with count >= thr_host (sout[0], sout[2], bmax kept)
    GB_CK(cudaMemsetAsync(sout + 1, 0, sizeof(u32), st));
    k_slots_collect<<<blocks, 256, 0, st>>>((const uint4*)slots, bmax, nullptr, thr_host, cand, sout + 1, (u32)cap);
  }
  GB_LAUNCH_CK();
  return 0;
}

// ---- pass 2: exact (key, count, first position) of the pairs that fall into candidate slots ------------------
template <class KF>
static int launch_cand_count(KF kf, i64 T, int log2ns, const u32* cand, const u32* n_cand_dev, int cap, u64* tk, u32* tc,
                             u32* tm, i64 M, int32_t* overflow, cudaStream_t st) {
  if (M <= 0 || (M & (M - 1))) { gb_set_error("cand_count: table size %lld is not a power of two", (long long)M); return -2; }
  if (cap > KC_SET / 2) { gb_set_error("cand_count: cap %d > %d", cap, KC_SET / 2); return -2; }
  if (T < 2) return 0;
  int sms = gb_sm_count();
  if (sms <= 0) sms = 148;
  int blocks = gb_blocks_for(T, 2 * KS_TILE, sms * 2);
  k_cand_count<KF><<<blocks, KS_THREADS, 0, st>>>(kf, T, 64 - log2ns, cand, n_cand_dev, (u32)cap, tk, tc, tm, (u64)(M - 1), overflow, nullptr, 1);
  GB_LAUNCH_CK();
  return 0;
}
extern "C" int gb_cand_count_m(const uint32_t* cls, const uint32_t* glue, int64_t T, int log2ns, const uint32_t* cand,
                               const uint32_t* n_cand_dev, int cap, uint64_t* tkeys, uint32_t* tcnt, uint32_t* tmin,
                               int64_t M, int32_t* overflow, int streaming, void* stream) {
  if (streaming) { KeysMT<true> kf{cls, glue}; return launch_cand_count(kf, (i64)T, log2ns, cand, n_cand_dev, cap, (u64*)tkeys, tcnt, tmin, (i64)M, overflow, (cudaStream_t)stream); }
  KeysMT<false> kf{cls, glue};
  return launch_cand_count(kf, (i64)T, log2ns, cand, n_cand_dev, cap, (u64*)tkeys, tcnt, tmin, (i64)M, overflow, (cudaStream_t)stream);
}
extern "C" int gb_cand_count_h(const uint64_t* hsh, const uint32_t* glue, const uint32_t* len, int64_t T,
                               const uint64_t* powtab, uint64_t P, int log2ns, const uint32_t* cand,
                               const uint32_t* n_cand_dev, int cap, uint64_t* tkeys, uint32_t* tcnt, uint32_t* tmin,
                               int64_t M, int32_t* overflow, int streaming, void* stream) {
  if (streaming) {
    KeysHT<true> kf{(const u64*)hsh, glue, len, (const u64*)powtab, (u64)P};
    return launch_cand_count(kf, (i64)T, log2ns, cand, n_cand_dev, cap, (u64*)tkeys, tcnt, tmin, (i64)M, overflow, (cudaStream_t)stream);
  }
  KeysHT<false> kf{(const u64*)hsh, glue, len, (const u64*)powtab, (u64)P};
  return launch_cand_count(kf, (i64)T, log2ns, cand, n_cand_dev, cap, (u64*)tkeys, tcnt, tmin, (i64)M, overflow, (cudaStream_t)stream);
}
