// Constants and helpers shared by the slot-sketch kernels (k1_slots.cu) and the device-driven learn loop (k_hloop.cu).
#pragma once
#include "common.cuh"
#include "k1_keys.cuh"

#define KS_THREADS 512
#define KS_TILE (KS_THREADS * K1_VEC)
#define KS_CACHE 8192                 // shared-memory hot stage per CTA: u32 tag + u32 count = 64 KB
#define KS_EMPTY 0xFFFFFFFFu

__device__ __forceinline__ u32 gb_slot_of(u64 key, int shift) {
  return (u32)((key * 0x9E3779B97F4A7C15ull) >> shift);       // Fibonacci hashing: top log2(NS) bits
}

#define KS_CHUNK 4096                 // slots per scan block (per-chunk maxima)
#define KC_SET 8192                   // shared open-addressing set of candidate slot ids (load <= 1/2 at cap 4096)
#define KC_LINEAR 8

// device-resident loop state (u32 words) of the histogram-mode learn loop; hs == nullptr -> host-driven call
enum {
  HS_T = 0, HS_STEP, HS_HALT, HS_EVN, HS_MPREV, HS_THR,
  HS_C1, HS_NCAND, HS_NZ, HS_PAD0,            // = gb_slots_top's sout[4]
  HS_M1, HS_NT, HS_ND, HS_ROUND2, HS_WIN_R0, HS_WIN_NRES, HS_WIN_CNT, HS_WIN_J, HS_MISMATCH, HS_FLAGGED,
  HS_NEWT, HS_NTOK, HS_OVF, HS_R2COUNT, HS_MAXSTEPS, HS_EVCAP, HS_HALT_STEP, HS_NTMAX,
  HS_N /* array extent incl. dead entries */, HS_NCL /* candidate-pair list length */, HS_CLCAP, HS_CMPT /* compaction total */,
  HS_DIST /* 1 = sharded run: counts are global, occurrences local */, HS_WORDS = 64
};

// ---- pass 1 -----------------------------------------------------------------------------------------------------
// CACHE: 0 none, 1 dynamic first-come claim, 2 static hot list (hot_tags u32 [KS_CACHE], KS_EMPTY = unused)
template <class KF, bool AGG, int CACHE, bool DRY>
__global__ void __launch_bounds__(KS_THREADS, 2) k_slot_count(KF kf, i64 T, u32* __restrict__ slots, int shift,
                                                              const u32* __restrict__ hot_tags, const u32* __restrict__ hs) {
  if (hs) { if (hs[HS_HALT]) return; T = (i64)hs[HS_N]; }      // device-driven loop: the array extent lives on the device
  extern __shared__ __align__(16) u32 ks_smem[];
  u32* s_tag = ks_smem;
  u32* s_cnt = ks_smem + KS_CACHE;
  if (CACHE) {
    for (int i = threadIdx.x; i < KS_CACHE; i += KS_THREADS) {
      s_tag[i] = (CACHE == 2) ? hot_tags[i] : KS_EMPTY;
      s_cnt[i] = 0;
    }
    __syncthreads();
  }
  u32 dry = 0;
  // two tiles per iteration: both tiles' 128-bit loads are issued before the first shuffle waits on them
  for (i64 base = (i64)blockIdx.x * (2 * KS_TILE); base < T; base += (i64)gridDim.x * (2 * KS_TILE)) {
    const i64 ja = base + (i64)threadIdx.x * K1_VEC, jb = ja + KS_TILE;
    typename KF::Raw ra, rb;
    kf.ld(ja, T, ra);
    kf.ld(jb, T, rb);
#pragma unroll
    for (int half = 0; half < 2; ++half) {
      u64 key[K1_VEC];
      if (half == 0) kf.fin(ja, T, ra, key); else kf.fin(jb, T, rb, key);
#pragma unroll
      for (int e = 0; e < K1_VEC; ++e) {
        bool valid = key[e] != GB_EMPTY_KEY;
        u32 slot = gb_slot_of(key[e], shift);
        u32 c = 1;
        if (AGG) {
          unsigned grp = __match_any_sync(0xffffffffu, valid ? slot : KS_EMPTY);
          c = __popc(grp);
          valid = valid && ((unsigned)(__ffs(grp) - 1) == (threadIdx.x & 31));
        }
        if (DRY) { dry ^= slot * c; continue; }
        if (!valid) continue;
        if (CACHE) {
          u32 cs = slot & (KS_CACHE - 1);
          u32 tag = s_tag[cs];
          if (CACHE == 1 && tag == KS_EMPTY) {
            tag = atomicCAS(&s_tag[cs], KS_EMPTY, slot);
            if (tag == KS_EMPTY) tag = slot;
          }
          if (tag == slot) atomicAdd(&s_cnt[cs], c);
          else atomicAdd(slots + slot, c);
        } else {
          atomicAdd(slots + slot, c);
        }
      }
    }
  }
  if (DRY) { if (dry == 0x12345u) slots[0] = dry; return; }
  if (CACHE) {
    __syncthreads();
    for (int i = threadIdx.x; i < KS_CACHE; i += KS_THREADS) {
      u32 c = s_cnt[i];
      if (c) atomicAdd(slots + s_tag[i], c);
    }
  }
}


// ---- pass 2: exact (key, count, first position) of the pairs that fall into candidate slots ------------------
template <class KF>
__global__ void __launch_bounds__(KS_THREADS, 2) k_cand_count(KF kf, i64 T, int shift, const u32* __restrict__ cand,
                                                              const u32* __restrict__ n_cand_dev, u32 cap,
                                                              u64* __restrict__ tk, u32* __restrict__ tc, u32* __restrict__ tm,
                                                              u64 mask, int32_t* __restrict__ overflow, const u32* __restrict__ hs, int round) {
  if (hs) {
    if (hs[HS_HALT] || (round == 2 && !hs[HS_ROUND2])) return;
    T = (i64)hs[HS_N];
  }
  __shared__ u32 s_set[KC_SET];
  __shared__ u32 s_lin[KC_LINEAR];
  u32 n = *n_cand_dev;
  if (n == 0) return;
  if (n > cap) { if (blockIdx.x == 0 && threadIdx.x == 0) atomicExch(overflow, 2); return; }
  const bool linear = n <= KC_LINEAR;
  if (linear) {
    if (threadIdx.x < KC_LINEAR) s_lin[threadIdx.x] = threadIdx.x < n ? cand[threadIdx.x] : KS_EMPTY;
  } else {
    for (int i = threadIdx.x; i < KC_SET; i += KS_THREADS) s_set[i] = KS_EMPTY;
    __syncthreads();
    for (u32 i = threadIdx.x; i < n; i += KS_THREADS) {
      u32 v = cand[i], s = (v * 0x9E3779B1u) >> 19;       // 13 bits
      while (true) {
        u32 prev = atomicCAS(&s_set[s], KS_EMPTY, v);
        if (prev == KS_EMPTY || prev == v) break;
        s = (s + 1) & (KC_SET - 1);
      }
    }
  }
  __syncthreads();
  for (i64 base = (i64)blockIdx.x * (2 * KS_TILE); base < T; base += (i64)gridDim.x * (2 * KS_TILE)) {
    const i64 ja = base + (i64)threadIdx.x * K1_VEC, jb = ja + KS_TILE;
    typename KF::Raw ra, rb;
    kf.ld(ja, T, ra);
    kf.ld(jb, T, rb);
#pragma unroll
    for (int half = 0; half < 2; ++half) {
      u64 key[K1_VEC];
      const i64 j0 = half ? jb : ja;
      if (half == 0) kf.fin(ja, T, ra, key); else kf.fin(jb, T, rb, key);
#pragma unroll
      for (int e = 0; e < K1_VEC; ++e) {
        if (key[e] == GB_EMPTY_KEY) continue;
        u32 slot = gb_slot_of(key[e], shift);
        bool hit = false;
        if (linear) {
#pragma unroll
          for (int i = 0; i < KC_LINEAR; ++i) hit |= (s_lin[i] == slot);
        } else {
          u32 s = (slot * 0x9E3779B1u) >> 19;
          while (true) {
            u32 v = s_set[s];
            if (v == slot) { hit = true; break; }
            if (v == KS_EMPTY) break;
            s = (s + 1) & (KC_SET - 1);
          }
        }
        if (hit && !gb_gtable_add(tk, tc, tm, mask, key[e], 1u, (u32)(j0 + e))) atomicExch(overflow, 1);
      }
    }
  }
}
