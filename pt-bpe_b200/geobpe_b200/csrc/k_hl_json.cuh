// JSON key strings of fully histogram-quantised windows + warp-wide string compare: the reference's tie-break order
// (python string order of json.dumps(compute_geo_key dict, sort_keys=True), bpe.py:1149,1469) evaluated on the device.
// Shared by the two device-driven learn loops (k_hloop.cu, k_hloop2.cu).
#pragma once
#include "common.cuh"

// ---- the reference's tie-break on the device -------------------------------------------------------------------
// json.dumps(compute_geo_key's dict, sort_keys=True) of a fully histogram-quantised window (bpe.py:1225-1298, 1149):
//   {"0C:1N": [0 x m], "C:1N:1CA": [..], "CA:C": [0 x nres], "CA:C:1N": [..m], "N:CA": [0 x nres], "omega": [..],
//    "phi": [..], "psi": [..m], "tau": [..nres]}      (omega / C:1N:1CA / phi only when nres > 1; bond bins are 0
// with std_bonds; m = nres - 1 when the window ends its backbone).  Residue-major rbins slots: 3 tau, 4 CA:C:1N,
// 5 psi, 6 omega, 7 C:1N:1CA, 8 phi(r+1).
__device__ __forceinline__ int hl_puts(char* p, int o, const char* s) {
  while (*s) p[o++] = *s++;
  return o;
}
__device__ __forceinline__ int hl_putu(char* p, int o, u32 v) {
  char d[10];
  int n = 0;
  do { d[n++] = (char)('0' + v % 10u); v /= 10u; } while (v);
  while (n) p[o++] = d[--n];
  return o;
}
__device__ __forceinline__ int hl_zero_list(char* p, int o, u32 n) {
  p[o++] = '[';
  for (u32 k = 0; k < n; ++k) { if (k) { p[o++] = ','; p[o++] = ' '; } p[o++] = '0'; }
  p[o++] = ']';
  return o;
}
__device__ __forceinline__ int hl_bin_list(char* p, int o, const u16* __restrict__ row, u32 n) {
  p[o++] = '[';
  for (u32 k = 0; k < n; ++k) { if (k) { p[o++] = ','; p[o++] = ' '; } o = hl_putu(p, o, row[k]); }
  p[o++] = ']';
  return o;
}
static __device__ int hl_json_key(char* p, const u16* __restrict__ rbins, i64 R, u32 r0, u32 nres) {
  const u32 m = (rbins[4 * R + r0 + nres - 1] == GB_NOBIN) ? nres - 1 : nres;
  int o = 0;
  o = hl_puts(p, o, "{\"0C:1N\": ");      o = hl_zero_list(p, o, m);
  if (nres > 1) { o = hl_puts(p, o, ", \"C:1N:1CA\": "); o = hl_bin_list(p, o, rbins + 7 * R + r0, nres - 1); }
  o = hl_puts(p, o, ", \"CA:C\": ");      o = hl_zero_list(p, o, nres);
  o = hl_puts(p, o, ", \"CA:C:1N\": ");   o = hl_bin_list(p, o, rbins + 4 * R + r0, m);
  o = hl_puts(p, o, ", \"N:CA\": ");      o = hl_zero_list(p, o, nres);
  if (nres > 1) {
    o = hl_puts(p, o, ", \"omega\": ");   o = hl_bin_list(p, o, rbins + 6 * R + r0, nres - 1);
    o = hl_puts(p, o, ", \"phi\": ");     o = hl_bin_list(p, o, rbins + 8 * R + r0, nres - 1);
  }
  o = hl_puts(p, o, ", \"psi\": ");       o = hl_bin_list(p, o, rbins + 5 * R + r0, m);
  o = hl_puts(p, o, ", \"tau\": ");       o = hl_bin_list(p, o, rbins + 3 * R + r0, nres);
  p[o++] = '}';
  const int end = (o + 127) & ~127;       // zero padding to the next 128-byte boundary (warp-wide compare)
  for (int k = o; k < end; ++k) p[k] = 0;
  return o;
}
// the same string written by lane 0 of a warp after all 32 lanes have gathered the window's bins into shared memory
// (sb: u16 [6][cap], cap >= nres): the serial formatter then reads shared memory instead of ~6 nres dependent global loads
__device__ __forceinline__ int hl_json_key_warp(char* p, const u16* __restrict__ rbins, i64 R, u32 r0, u32 nres, u16* sb, u32 cap) {
  const unsigned lane = threadIdx.x & 31;
  for (u32 e = lane; e < 6u * nres; e += 32u) {
    const u32 slot = e / nres, k = e - slot * nres;
    sb[slot * cap + k] = __ldg(rbins + (i64)(3u + slot) * R + r0 + k);
  }
  __syncwarp();
  int o = 0;
  if (lane == 0) o = hl_json_key(p, sb - 3 * (i64)cap, (i64)cap, 0u, nres);
  __syncwarp();
  return __shfl_sync(0xffffffffu, o, 0);
}
// warp-cooperative strcmp of two zero-padded, 128-byte aligned strings: true when a < b
__device__ __forceinline__ bool hl_less(const char* a, const char* b, int la, int lb) {
  const unsigned lane = threadIdx.x & 31;
  const int n = (min(la, lb) + 127) & ~127;
  for (int off = 0; off < n; off += 128) {
    u32 wa = reinterpret_cast<const u32*>(a + off)[lane], wb = reinterpret_cast<const u32*>(b + off)[lane];
    unsigned diff = __ballot_sync(0xffffffffu, wa != wb);
    if (diff) {
      int l = __ffs(diff) - 1;
      u32 xa = __byte_perm(__shfl_sync(0xffffffffu, wa, l), 0, 0x0123);   // memory order -> most significant first
      u32 xb = __byte_perm(__shfl_sync(0xffffffffu, wb, l), 0, 0x0123);
      return xa < xb;
    }
  }
  return la < lb;
}
// residues of the pair whose left token is entry j (lazy layout: ls = len | extra << 16, right neighbour j + 1 + extra)
__device__ __forceinline__ u32 hl_pair_res(const u32* __restrict__ ls, u32 j) {
  const u32 l = ls[j];
  return (l & 0xFFFFu) + (ls[j + 1u + (l >> 16)] & 0xFFFFu);
}
