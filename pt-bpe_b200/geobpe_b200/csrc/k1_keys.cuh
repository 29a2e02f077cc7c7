// Pair-key functors shared by the K1 kernels (exact table, slot sketch, candidate resolve) and K2 flagging.
// Two-phase API so a kernel can put the loads of several tiles in flight before the first shuffle waits on them:
//   Raw r;  kf.ld(j0, T, r);  ...  kf.fin(j0, T, r, key);        (kf.keys = ld + fin)
#pragma once
#include "common.cuh"

#define K1_VEC 4

// 128-bit loads; CS = cache-streaming (ld.global.cs: evict-first, for token arrays larger than L2 so that the
// slot counters / pair table stay resident)
template <bool CS> __device__ __forceinline__ uint4 gb_ld4(const u32* p) {
  return CS ? __ldcs(reinterpret_cast<const uint4*>(p)) : *reinterpret_cast<const uint4*>(p);
}
template <bool CS> __device__ __forceinline__ ulonglong2 gb_ld2(const u64* p) {
  return CS ? __ldcs(reinterpret_cast<const ulonglong2*>(p)) : *reinterpret_cast<const ulonglong2*>(p);
}

// ---- medoid mode: key = cls[j]<<45 | cls[j+1]<<27 | glue27(j) ---------------------------------------------------
template <bool CS>
struct KeysMT {
  const u32* cls; const u32* glue;
  struct Raw { u32 c[K1_VEC + 1]; u32 g[K1_VEC]; };
  __device__ __forceinline__ void ld(i64 j0, i64 T, Raw& r) const {
    if (j0 + K1_VEC <= T) {
      uint4 cv = gb_ld4<CS>(cls + j0);
      uint4 gv = gb_ld4<CS>(glue + j0);
      r.c[0] = cv.x; r.c[1] = cv.y; r.c[2] = cv.z; r.c[3] = cv.w;
      r.g[0] = gv.x; r.g[1] = gv.y; r.g[2] = gv.z; r.g[3] = gv.w;
    } else {
#pragma unroll
      for (int e = 0; e < K1_VEC; ++e) {
        bool in = j0 + e < T;
        r.c[e] = in ? cls[j0 + e] : 0u;
        r.g[e] = in ? glue[j0 + e] : GB_LAST_FLAG;
      }
    }
    // right neighbour of the 4th token: first token of the next lane (shuffled in fin), one scalar load for lane 31
    r.c[K1_VEC] = 0u;
    if ((threadIdx.x & 31) == 31 && j0 + K1_VEC < T) r.c[K1_VEC] = cls[j0 + K1_VEC];
  }
  __device__ __forceinline__ void fin(i64 j0, i64 T, Raw& r, u64 out[K1_VEC]) const {
    u32 nxt = __shfl_down_sync(0xffffffffu, r.c[0], 1);
    if ((threadIdx.x & 31) != 31) r.c[K1_VEC] = nxt;
#pragma unroll
    for (int e = 0; e < K1_VEC; ++e) {
      bool valid = (j0 + e + 1 < T) && !(r.g[e] & GB_LAST_FLAG);
      out[e] = valid ? gb_key_m(r.c[e], r.g[e], r.c[e + 1]) : GB_EMPTY_KEY;
    }
  }
  __device__ __forceinline__ void keys(i64 j0, i64 T, u64 out[K1_VEC]) const {
    Raw r;
    ld(j0, T, r);
    fin(j0, T, r, out);
  }
};

// ---- histogram mode: key = content hash of hsh[j] . gluesym(j) . hsh[j+1] (gb_key_h) --------------------------------
template <bool CS>
struct KeysHT {
  const u64* hsh; const u32* glue; const u32* len; const u64* powtab; u64 P;
  struct Raw { u64 h[K1_VEC + 1]; u32 g[K1_VEC]; u32 l[K1_VEC + 1]; };
  __device__ __forceinline__ void ld(i64 j0, i64 T, Raw& r) const {
    if (j0 + K1_VEC <= T) {
      ulonglong2 h01 = gb_ld2<CS>(hsh + j0);
      ulonglong2 h23 = gb_ld2<CS>(hsh + j0 + 2);
      uint4 gv = gb_ld4<CS>(glue + j0);
      uint4 lv = gb_ld4<CS>(len + j0);
      r.h[0] = h01.x; r.h[1] = h01.y; r.h[2] = h23.x; r.h[3] = h23.y;
      r.g[0] = gv.x; r.g[1] = gv.y; r.g[2] = gv.z; r.g[3] = gv.w;
      r.l[0] = lv.x; r.l[1] = lv.y; r.l[2] = lv.z; r.l[3] = lv.w;
    } else {
#pragma unroll
      for (int e = 0; e < K1_VEC; ++e) {
        bool in = j0 + e < T;
        r.h[e] = in ? hsh[j0 + e] : 0ull;
        r.g[e] = in ? glue[j0 + e] : GB_LAST_FLAG;
        r.l[e] = in ? len[j0 + e] : 1u;
      }
    }
    r.h[K1_VEC] = 0ull; r.l[K1_VEC] = 1u;
    if ((threadIdx.x & 31) == 31 && j0 + K1_VEC < T) { r.h[K1_VEC] = hsh[j0 + K1_VEC]; r.l[K1_VEC] = len[j0 + K1_VEC]; }
  }
  __device__ __forceinline__ void fin(i64 j0, i64 T, Raw& r, u64 out[K1_VEC]) const {
    u64 nh = __shfl_down_sync(0xffffffffu, r.h[0], 1);
    u32 nl = __shfl_down_sync(0xffffffffu, r.l[0], 1);
    if ((threadIdx.x & 31) != 31) { r.h[K1_VEC] = nh; r.l[K1_VEC] = nl; }
#pragma unroll
    for (int e = 0; e < K1_VEC; ++e) {
      bool valid = (j0 + e + 1 < T) && !(r.g[e] & GB_LAST_FLAG);
      out[e] = valid ? gb_key_h(r.h[e], r.g[e], r.h[e + 1], P, __ldg(powtab + (2u * r.l[e + 1] - 1u))) : GB_EMPTY_KEY;
    }
  }
  __device__ __forceinline__ void keys(i64 j0, i64 T, u64 out[K1_VEC]) const {
    Raw r;
    ld(j0, T, r);
    fin(j0, T, r, out);
  }
};

// ---- histogram mode, lazily compacted arrays (device-driven loop) ---------------------------------------------------
// ls[j] = len(residues) | extra << 16; extra = dead entries between token j and its right neighbour (absorbed by earlier
// merges and not yet compacted away); ls[j] == 0 marks a dead entry.  Right neighbour of j = j + 1 + extra.
template <bool CS>
struct KeysHLT {
  const u64* hsh; const u32* glue; const u32* ls; const u64* powtab; u64 P;
  struct Raw { u64 h[K1_VEC + 1]; u32 g[K1_VEC]; u32 l[K1_VEC + 1]; };
  __device__ __forceinline__ void ld(i64 j0, i64 N, Raw& r) const {
    if (j0 + K1_VEC <= N) {
      ulonglong2 h01 = gb_ld2<CS>(hsh + j0);
      ulonglong2 h23 = gb_ld2<CS>(hsh + j0 + 2);
      uint4 gv = gb_ld4<CS>(glue + j0);
      uint4 lv = gb_ld4<CS>(ls + j0);
      r.h[0] = h01.x; r.h[1] = h01.y; r.h[2] = h23.x; r.h[3] = h23.y;
      r.g[0] = gv.x; r.g[1] = gv.y; r.g[2] = gv.z; r.g[3] = gv.w;
      r.l[0] = lv.x; r.l[1] = lv.y; r.l[2] = lv.z; r.l[3] = lv.w;
    } else {
#pragma unroll
      for (int e = 0; e < K1_VEC; ++e) {
        bool in = j0 + e < N;
        r.h[e] = in ? hsh[j0 + e] : 0ull;
        r.g[e] = in ? glue[j0 + e] : GB_LAST_FLAG;
        r.l[e] = in ? ls[j0 + e] : 0u;
      }
    }
    r.h[K1_VEC] = 0ull; r.l[K1_VEC] = 1u;
    if ((threadIdx.x & 31) == 31 && j0 + K1_VEC < N) { r.h[K1_VEC] = hsh[j0 + K1_VEC]; r.l[K1_VEC] = ls[j0 + K1_VEC]; }
  }
  __device__ __forceinline__ void fin(i64 j0, i64 N, Raw& r, u64 out[K1_VEC]) const {
    u64 nh = __shfl_down_sync(0xffffffffu, r.h[0], 1);
    u32 nl = __shfl_down_sync(0xffffffffu, r.l[0], 1);
    if ((threadIdx.x & 31) != 31) { r.h[K1_VEC] = nh; r.l[K1_VEC] = nl; }
#pragma unroll
    for (int e = 0; e < K1_VEC; ++e) {
      const u32 le = r.l[e];
      out[e] = GB_EMPTY_KEY;
      if (le == 0u || (r.g[e] & GB_LAST_FLAG) || j0 + e >= N) continue;
      const u32 extra = le >> 16;
      u64 hR = r.h[e + 1];
      u32 lR = r.l[e + 1];
      if (extra) { const i64 idx = j0 + e + 1 + extra; hR = hsh[idx]; lR = ls[idx]; }
      out[e] = gb_key_h(r.h[e], r.g[e], hR, P, __ldg(powtab + (2u * (lR & 0xFFFFu) - 1u)));
    }
  }
};

typedef KeysMT<false> KeysM;
typedef KeysHT<false> KeysH;
