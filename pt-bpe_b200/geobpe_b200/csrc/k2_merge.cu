// K2: the merge step (bpe.py:1895-2017): flag the occurrences of the popped key, resolve overlapping
// occurrences left-to-right (bpe.py:1909-1914), log merge events / occurrence-set operations.
// The stream compaction of the token arrays itself lives in k_select.cu (gb_merge_compact).
#include "common.cuh"

// ---- flag --------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_flag_m(const u32* __restrict__ cls, const u32* __restrict__ glue, i64 T,
                                                 u64 key, u8* __restrict__ flags) {
  for (i64 j = (i64)blockIdx.x * blockDim.x + threadIdx.x; j < T; j += (i64)gridDim.x * blockDim.x) {
    u8 f = 0;
    if (j + 1 < T) {
      u32 g = glue[j];
      if (!(g & GB_LAST_FLAG)) f = gb_key_m(cls[j], g, cls[j + 1]) == key;
    }
    flags[j] = f;
  }
}
__global__ void __launch_bounds__(256) k_flag_h(const u64* __restrict__ hsh, const u32* __restrict__ glue,
                                                 const u32* __restrict__ len, i64 T, const u64* __restrict__ powtab,
                                                 u64 P, u64 key, u8* __restrict__ flags) {
  for (i64 j = (i64)blockIdx.x * blockDim.x + threadIdx.x; j < T; j += (i64)gridDim.x * blockDim.x) {
    u8 f = 0;
    if (j + 1 < T) {
      u32 g = glue[j];
      if (!(g & GB_LAST_FLAG))
        f = gb_key_h(hsh[j], g, hsh[j + 1], P, __ldg(powtab + (2u * len[j + 1] - 1u))) == key;
    }
    flags[j] = f;
  }
}
extern "C" int gb_flag_pairs_m(const uint32_t* cls, const uint32_t* glue, int64_t T, uint64_t key, uint8_t* flags,
                               void* stream) {
  if (T <= 0) return 0;
  k_flag_m<<<gb_blocks_for(T, 256 * 4, 148 * 8), 256, 0, (cudaStream_t)stream>>>(cls, glue, T, (u64)key, flags);
  GB_LAUNCH_CK();
  return 0;
}
extern "C" int gb_flag_pairs_h(const uint64_t* hsh, const uint32_t* glue, const uint32_t* len, int64_t T,
                               const uint64_t* powtab, uint64_t P, uint64_t key, uint8_t* flags, void* stream) {
  if (T <= 0) return 0;
  k_flag_h<<<gb_blocks_for(T, 256 * 4, 148 * 8), 256, 0, (cudaStream_t)stream>>>((const u64*)hsh, glue, len, T,
                                                                                (const u64*)powtab, (u64)P, (u64)key, flags);
  GB_LAUNCH_CK();
  return 0;
}

// ---- overlap rule --------------------------------------------------------------------------------
// A maximal run of consecutive flagged pairs j, j+1, ... shares tokens pairwise; processing sorted by
// (structure, index) the reference merges the 1st, skips the 2nd (last_i1 + length > i1), merges the 3rd ...
// One thread per run head walks its run (runs never cross a backbone: the last token's pair is not flagged).
__global__ void __launch_bounds__(256) k_resolve_overlaps(u8* __restrict__ flags, i64 T) {
  for (i64 j = (i64)blockIdx.x * blockDim.x + threadIdx.x; j < T; j += (i64)gridDim.x * blockDim.x) {
    if ((flags[j] & 1) && !(j > 0 && (flags[j - 1] & 1))) {
      i64 i = j;
      bool keep = true;
      while (i < T && (flags[i] & 1)) {
        if (keep) flags[i] = 3;
        keep = !keep;
        ++i;
      }
    }
  }
}
extern "C" int gb_resolve_overlaps(uint8_t* flags, int64_t T, void* stream) {
  if (T <= 0) return 0;
  k_resolve_overlaps<<<gb_blocks_for(T, 256 * 4, 148 * 8), 256, 0, (cudaStream_t)stream>>>(flags, T);
  GB_LAUNCH_CK();
  return 0;
}

__global__ void k_scatter_u32(const u32* __restrict__ occ, const u32* __restrict__ val, i64 Q, u32* __restrict__ aux) {
  i64 q = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (q < Q) aux[occ[q]] = val[q];
}
extern "C" int gb_scatter_u32(const uint32_t* occ, const uint32_t* val, int64_t Q, uint32_t* aux, void* stream) {
  if (Q <= 0) return 0;
  k_scatter_u32<<<(unsigned)((Q + 255) / 256), 256, 0, (cudaStream_t)stream>>>(occ, val, Q, aux);
  GB_LAUNCH_CK();
  return 0;
}

__global__ void k_write_events(const u32* __restrict__ kept, i64 Q, const u32* __restrict__ count_dev,
                               const u32* __restrict__ pos, const u32* __restrict__ aux, u32* __restrict__ el,
                               u32* __restrict__ er, u32* __restrict__ ep) {
  i64 q = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (count_dev && q >= (i64)*count_dev) return;
  if (q < Q) {
    u32 j = kept[q];
    el[q] = pos[j];
    er[q] = pos[j + 1];
    ep[q] = aux ? aux[j] : 0u;
  }
}
extern "C" int gb_write_events(const uint32_t* kept, int64_t Q, const uint32_t* count_dev, const uint32_t* pos,
                               const uint32_t* aux, uint32_t* ev_left, uint32_t* ev_right, uint32_t* ev_p, void* stream) {
  if (Q <= 0) return 0;
  k_write_events<<<(unsigned)((Q + 255) / 256), 256, 0, (cudaStream_t)stream>>>(kept, Q, count_dev, pos, aux, ev_left, ev_right, ev_p);
  GB_LAUNCH_CK();
  return 0;
}

// ---- occurrence-set op log (medoid mode) -------------------------------------------------------
// Reference order per merged occurrence (i1 | i2 | i3 = start bonds of left, right, right-right token):
//   remove key@(i,i2); remove left_key@(i,i1); remove right_key@(i,i3); add new_left@(i,i1); add new_right@(i,i3)
// (bpe.py:1924-1950, 1992-2006).  The popped key's own removals are implicit (its set always ends empty).
// "left token" at processing time is the token produced by the previous kept occurrence if that one ended
// exactly at j-1 (flags[j-2] kept), because occurrences are processed left to right.
__global__ void k_oplog_m(const u32* __restrict__ kept, i64 Q, const u8* __restrict__ flags,
                          const u32* __restrict__ cls, const u32* __restrict__ glue, const u32* __restrict__ pos,
                          i64 T, const u32* __restrict__ aux, const u32* __restrict__ p2cls,
                          u64* __restrict__ lk, u32* __restrict__ lp) {
  i64 q = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= Q) return;
  i64 j = kept[q];
  u32 ncls = p2cls[aux[j]];
  u64 k0 = GB_EMPTY_KEY, k1 = GB_EMPTY_KEY, k2 = GB_EMPTY_KEY, k3 = GB_EMPTY_KEY;
  u32 pl = 0, pr = 0;
  if (j > 0 && !(glue[j - 1] & GB_LAST_FLAG)) {
    u32 lc = (j >= 2 && (flags[j - 2] & 2)) ? p2cls[aux[j - 2]] : cls[j - 1];
    u32 g = glue[j - 1];
    k0 = gb_key_m(lc, g, cls[j]);
    k2 = gb_key_m(lc, g, ncls);
    pl = pos[j];
  }
  u32 g1 = glue[j + 1];
  if (!(g1 & GB_LAST_FLAG)) {
    u32 rc = cls[j + 2];
    k1 = gb_key_m(cls[j + 1], g1, rc);
    k3 = gb_key_m(ncls, g1, rc);
    pr = pos[j + 2];
  }
  lk[4 * q + 0] = k0; lp[4 * q + 0] = pl;
  lk[4 * q + 1] = k1; lp[4 * q + 1] = pr;
  lk[4 * q + 2] = k2; lp[4 * q + 2] = pl;
  lk[4 * q + 3] = k3; lp[4 * q + 3] = pr;
}
extern "C" int gb_oplog_m(const uint32_t* kept, int64_t Q, const uint8_t* flags, const uint32_t* cls,
                          const uint32_t* glue, const uint32_t* pos, int64_t T, const uint32_t* aux,
                          const uint32_t* p2cls, uint64_t* log_key, uint32_t* log_pos, void* stream) {
  if (Q <= 0) return 0;
  k_oplog_m<<<(unsigned)((Q + 255) / 256), 256, 0, (cudaStream_t)stream>>>(kept, Q, flags, cls, glue, pos, T, aux, p2cls,
                                                                           (u64*)log_key, log_pos);
  GB_LAUNCH_CK();
  return 0;
}

__global__ void __launch_bounds__(256) k_pair_keys_m(const u32* __restrict__ cls, const u32* __restrict__ glue,
                                                      const u32* __restrict__ pos, i64 T, u64* __restrict__ lk,
                                                      u32* __restrict__ lp) {
  for (i64 j = (i64)blockIdx.x * blockDim.x + threadIdx.x; j < T; j += (i64)gridDim.x * blockDim.x) {
    u64 k = GB_EMPTY_KEY;
    u32 p = 0;
    if (j + 1 < T) {
      u32 g = glue[j];
      if (!(g & GB_LAST_FLAG)) { k = gb_key_m(cls[j], g, cls[j + 1]); p = pos[j + 1]; }
    }
    lk[j] = k; lp[j] = p;
  }
}
extern "C" int gb_pair_keys_m(const uint32_t* cls, const uint32_t* glue, const uint32_t* pos, int64_t T,
                              uint64_t* log_key, uint32_t* log_pos, void* stream) {
  if (T <= 0) return 0;
  k_pair_keys_m<<<gb_blocks_for(T, 256 * 4, 148 * 8), 256, 0, (cudaStream_t)stream>>>(cls, glue, pos, T, (u64*)log_key, log_pos);
  GB_LAUNCH_CK();
  return 0;
}

// ---- histogram-mode exact window check ---------------------------------------------------------
// the pair key is a 61-bit content hash; before merging verify that all flagged windows hold exactly the
// same bins as the reference window (slots 3..8 of every residue; bond slots are constant with std_bonds)
__global__ void k_verify_windows_h(const u32* __restrict__ occ, i64 Q, const u32* __restrict__ pos,
                                   const u32* __restrict__ len, const u16* __restrict__ rbins, i64 R, u32 ref_r0,
                                   u32 nres, u32* __restrict__ mismatches) {
  i64 w = ((i64)blockIdx.x * blockDim.x + threadIdx.x) >> 5;  // warp per occurrence
  unsigned lane = threadIdx.x & 31;
  if (w >= Q) return;
  u32 j = occ[w];
  u32 r0 = pos[j];
  bool bad = (len[j] + len[j + 1]) != nres;
  if (!bad) {
    for (u32 i = lane; i < nres * 6u; i += 32) {
      u32 slot = 3 + i / nres, k = i % nres;
      if (slot >= 6 && k == nres - 1) continue;  // glue of the last residue lies outside the window
      if (rbins[(i64)slot * R + r0 + k] != rbins[(i64)slot * R + ref_r0 + k]) bad = true;
    }
  }
  bad = __any_sync(0xffffffffu, bad);
  if (bad && lane == 0) atomicAdd(mismatches, 1u);
}
extern "C" int gb_verify_windows_h(const uint32_t* occ, int64_t Q, const uint32_t* pos, const uint32_t* len,
                                   const uint16_t* rbins, int64_t R, uint32_t ref_r0, uint32_t nres,
                                   uint32_t* mismatches, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  GB_CK(cudaMemsetAsync(mismatches, 0, sizeof(u32), st));
  if (Q <= 0) return 0;
  i64 threads = Q * 32;
  k_verify_windows_h<<<(unsigned)((threads + 255) / 256), 256, 0, st>>>(occ, Q, pos, len, rbins, R, ref_r0, nres, mismatches);
  GB_LAUNCH_CK();
  return 0;
}

// same check driven by the flags array (no occurrence list, no host round trip): out2 = {mismatches, flagged}
__global__ void __launch_bounds__(256) k_verify_flagged_h(const u8* __restrict__ flags, i64 T, const u32* __restrict__ pos,
                                                           const u32* __restrict__ len, const u16* __restrict__ rbins, i64 R,
                                                           u32 ref_r0, u32 nres, u32* __restrict__ out2) {
  for (i64 j = (i64)blockIdx.x * blockDim.x + threadIdx.x; j < T; j += (i64)gridDim.x * blockDim.x) {
    if (!(flags[j] & 1)) continue;
    atomicAdd(&out2[1], 1u);
    u32 r0 = pos[j];
    bool bad = (len[j] + len[j + 1]) != nres;
    for (u32 slot = 3; slot < 9 && !bad; ++slot) {
      u32 n = (slot >= 6) ? nres - 1 : nres;   // glue of the last residue lies outside the window
      for (u32 k = 0; k < n; ++k)
        if (rbins[(i64)slot * R + r0 + k] != rbins[(i64)slot * R + ref_r0 + k]) { bad = true; break; }
    }
    if (bad) atomicAdd(&out2[0], 1u);
  }
}
extern "C" int gb_verify_flagged_h(const uint8_t* flags, int64_t T, const uint32_t* pos, const uint32_t* len,
                                   const uint16_t* rbins, int64_t R, uint32_t ref_r0, uint32_t nres, uint32_t* out2,
                                   void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  GB_CK(cudaMemsetAsync(out2, 0, 2 * sizeof(u32), st));
  if (T <= 0) return 0;
  k_verify_flagged_h<<<gb_blocks_for(T, 256 * 4, 148 * 8), 256, 0, st>>>(flags, T, pos, len, rbins, R, ref_r0, nres, out2);
  GB_LAUNCH_CK();
  return 0;
}
