// Shared device helpers for the geobpe_b200 kernels (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include "geobpe_b200.h"

typedef unsigned long long u64;
typedef uint32_t u32;
typedef uint16_t u16;
typedef uint8_t u8;
typedef int64_t i64;

#define GB_TWO_PI 6.283185307179586  // == 2*np.pi in float64

void gb_set_error(const char* fmt, ...);

#define GB_CK(expr)                                                                      \
  do {                                                                                   \
    cudaError_t _e = (expr);                                                             \
    if (_e != cudaSuccess) {                                                             \
      gb_set_error("%s:%d %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e)); \
      return (int)_e ? (int)_e : -1;                                                     \
    }                                                                                    \
  } while (0)

#define GB_LAUNCH_CK() GB_CK(cudaGetLastError())

static inline int gb_blocks_for(i64 n, int per_block, int cap) {
  i64 b = (n + per_block - 1) / per_block;
  if (b < 1) b = 1;
  if (b > cap) b = cap;
  return (int)b;
}

// ---- 64-bit mixing / open-addressing table -------------------------------------------------------
__device__ __forceinline__ u64 gb_mix64(u64 x) {
  x ^= x >> 33; x *= 0xff51afd7ed558ccdULL;
  x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ULL;
  x ^= x >> 33;
  return x;
}

#define GB_MAX_PROBES 8192

// add (key, c, mn) to the global table; returns false when no slot was found within GB_MAX_PROBES
__device__ __forceinline__ bool gb_gtable_add(u64* __restrict__ tk, u32* __restrict__ tc,
                                              u32* __restrict__ tm, u64 mask, u64 key, u32 c, u32 mn) {
  u64 s = gb_mix64(key) & mask;
  for (int probe = 0; probe < GB_MAX_PROBES; ++probe) {
    u64 prev = tk[s];
    if (prev == GB_EMPTY_KEY) prev = atomicCAS(&tk[s], (u64)GB_EMPTY_KEY, key);
    if (prev == GB_EMPTY_KEY || prev == key) {
      atomicAdd(&tc[s], c);
      atomicMin(&tm[s], mn);
      return true;
    }
    s = (s + 1) & mask;
  }
  return false;
}

// ---- pair keys ---------------------------------------------------------------------------------
// medoid mode: 18-bit class ids, glue bins compressed 3x10 -> 3x9 bits (bins <= 512)
__device__ __forceinline__ u64 gb_key_m(u32 cL, u32 g, u32 cR) {
  u32 g27 = (g & 0x1FFu) | (((g >> 10) & 0x1FFu) << 9) | (((g >> 20) & 0x1FFu) << 18);
  return ((u64)cL << 45) | ((u64)cR << 27) | (u64)g27;
}

// histogram mode: polynomial content hash over Z/2^64 (odd base P, natural wrap-around) of
//   content(L) . gluesym . content(R);  PpowR = P^(2*lenR-1), the number of symbols in R.
// Concatenation-stable: H(L.g.R) = (H(L)*P + g)*P^|R| + H(R), so [A][BC] and [AB][C] share a key like the reference's
// whole-window bin vector does.  Two 64-bit multiplies per pair (the former mod 2^61-1 variant needed two 128-bit
// products and made K1 ALU-bound).  Every merge still verifies the flagged windows bin by bin (gb_verify_*).
__device__ __forceinline__ u64 gb_key_h(u64 hL, u32 g, u64 hR, u64 P, u64 PpowR) {
  u64 gs = (u64)(g & 0x3FFFFFFFu) + 1ull;
  u64 k = (hL * P + gs) * PpowR + hR;
  return k == GB_EMPTY_KEY ? k - 1ull : k;     // the all-ones value is reserved for "no pair"
}

// ---- binary flags block scan (blockDim.x multiple of 32, <= 1024) -------------------------------
// returns exclusive prefix of `f` within the block and the block total (all threads)
__device__ __forceinline__ u32 gb_block_excl_scan_flag(bool f, u32* warp_tot /*smem[32]*/, u32& total) {
  const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
  unsigned b = __ballot_sync(0xffffffffu, f);
  u32 pre = __popc(b & ((1u << lane) - 1u));
  if (lane == 0) warp_tot[warp] = __popc(b);
  __syncthreads();
  u32 off = 0, tot = 0;
  for (unsigned w = 0; w < nw; ++w) {
    u32 v = warp_tot[w];
    if (w < warp) off += v;
    tot += v;
  }
  __syncthreads();
  total = tot;
  return off + pre;
}
