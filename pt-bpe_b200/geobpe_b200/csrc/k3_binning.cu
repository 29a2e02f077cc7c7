// K3: thresholds and value->bin quantisation.
// Replaces bpe.py:820-876 (_init_thresholds), plotting.py:305-337 (save_histogram: np.histogram),
// bpe.py:1165-1189 (get_ind), 1513-1526 (quant_geo), 715-737 (_set_bond_length_worker),
// 381-391 (glue -> bin centre).  HBM-bound: 8 B per value read.
#include "common.cuh"
#include <math.h>

// ---- backbone index of every residue ----------------------------------------------------------
__global__ void k_residue_backbone(const i64* __restrict__ offsets, i64 S, i64 R, u32* __restrict__ bb) {
  i64 r = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  i64 stride = (i64)gridDim.x * blockDim.x;
  for (; r < R; r += stride) {
    i64 lo = 0, hi = S;  // largest s with offsets[s] <= r
    while (hi - lo > 1) {
      i64 mid = (lo + hi) >> 1;
      if (offsets[mid] <= r) lo = mid; else hi = mid;
    }
    bb[r] = (u32)lo;
  }
}

extern "C" int gb_residue_backbone(const int64_t* offsets, int64_t S, int64_t R, uint32_t* bb, void* stream) {
  if (R <= 0) return 0;
  int blocks = gb_blocks_for(R, 256, 148 * 16);
  k_residue_backbone<<<blocks, 256, 0, (cudaStream_t)stream>>>(offsets, S, R, bb);
  GB_LAUNCH_CK();
  return 0;
}

// ---- column min / max / count -------------------------------------------------------------------
__device__ __forceinline__ u64 enc_f64(double d) {
  u64 b = (u64)__double_as_longlong(d);
  return (b >> 63) ? ~b : (b | 0x8000000000000000ull);
}
__device__ __forceinline__ double dec_f64(u64 e) {
  u64 b = (e >> 63) ? (e & 0x7FFFFFFFFFFFFFFFull) : ~e;
  return __longlong_as_double((long long)b);
}
__device__ __forceinline__ double wrap_angle(double v) { return fmod(v + GB_TWO_PI, GB_TWO_PI); }

// scratch u64 [9][3] = {enc min, enc max, count}
__global__ void k_minmax_init(u64* s) {
  int c = threadIdx.x;
  if (c < 9) { s[c * 3 + 0] = ~0ull; s[c * 3 + 1] = 0ull; s[c * 3 + 2] = 0ull; }
}
__global__ void __launch_bounds__(256) k_col_minmax(const double* __restrict__ vals, i64 R, i64 ld, u64* __restrict__ s) {
  const int c = blockIdx.y;
  const bool wrap = c >= 3;
  const double* col = vals + (i64)c * ld;
  u64 mn = ~0ull, mx = 0ull, cnt = 0;
  for (i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x; i < R; i += (i64)gridDim.x * blockDim.x) {
    double v = col[i];
    if (v == v && v != 0.0) {
      if (wrap) v = wrap_angle(v);
      u64 e = enc_f64(v);
      mn = e < mn ? e : mn;
      mx = e > mx ? e : mx;
      ++cnt;
    }
  }
  for (int o = 16; o; o >>= 1) {
    u64 a = __shfl_xor_sync(0xffffffffu, mn, o); mn = a < mn ? a : mn;
    u64 b = __shfl_xor_sync(0xffffffffu, mx, o); mx = b > mx ? b : mx;
    cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
  }
  if ((threadIdx.x & 31) == 0) {
    atomicMin(&s[c * 3 + 0], mn);
    atomicMax(&s[c * 3 + 1], mx);
    atomicAdd(&s[c * 3 + 2], cnt);
  }
}
__global__ void k_minmax_fin(const u64* s, double* minmax, u64* cnt) {
  int c = threadIdx.x;
  if (c < 9) {
    minmax[c * 2 + 0] = dec_f64(s[c * 3 + 0]);
    minmax[c * 2 + 1] = dec_f64(s[c * 3 + 1]);
    cnt[c] = s[c * 3 + 2];
  }
}

static __device__ u64 g_minmax_scratch[27];

extern "C" int gb_col_minmax(const double* vals, int64_t R, int64_t ld, double* minmax, unsigned long long* cnt,
                             void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  u64* s = nullptr;
  GB_CK(cudaGetSymbolAddress((void**)&s, g_minmax_scratch));
  k_minmax_init<<<1, 32, 0, st>>>(s);
  GB_LAUNCH_CK();
  if (R > 0) {
    dim3 grid(gb_blocks_for(R, 256 * 8, 148 * 4), 9);
    k_col_minmax<<<grid, 256, 0, st>>>(vals, R, ld, s);
    GB_LAUNCH_CK();
  }
  k_minmax_fin<<<1, 32, 0, st>>>(s, minmax, (u64*)cnt);
  GB_LAUNCH_CK();
  return 0;
}

// ---- get_ind (bpe.py:1165-1189): bisect_right(left_edges, v) - 1, last bin closed ------------------
// returns bin index, -1 "below the first bin range", -2 "does not fall into any bin"
__device__ __forceinline__ int find_bin(const double* __restrict__ e, int B, double v) {
  int lo = 0, hi = B;  // upper bound over left edges e[0..B-1]
  while (lo < hi) {
    int mid = (lo + hi) >> 1;
    if (e[mid] <= v) lo = mid + 1; else hi = mid;
  }
  int ind = lo - 1;
  if (ind < 0) return (v != v) ? -2 : -1;
  double start = e[ind], end = e[ind + 1];
  if (ind == B - 1 && v == end) return ind;
  if (start <= v && v < end) return ind;
  return -2;
}

struct NB9 { int nb[9]; };

#define HIST_MAXB 1024
// block-local shared histogram per column, flushed with one atomic per non-empty bin per block
__global__ void __launch_bounds__(256) k_col_hist(const double* __restrict__ vals, i64 R, i64 ld,
                                                   const double* __restrict__ edges, int estride, NB9 nb,
                                                   u64* __restrict__ counts) {
  __shared__ double se[HIST_MAXB + 1];
  __shared__ u32 sc[HIST_MAXB];
  const int c = blockIdx.y;
  const int B = nb.nb[c];
  if (B <= 0) return;
  const bool wrap = c >= 3;
  for (int i = threadIdx.x; i <= B; i += blockDim.x) se[i] = edges[(i64)c * estride + i];
  for (int i = threadIdx.x; i < B; i += blockDim.x) sc[i] = 0;
  __syncthreads();
  const double* col = vals + (i64)c * ld;
  for (i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x; i < R; i += (i64)gridDim.x * blockDim.x) {
    double v = col[i];
    if (v == v && v != 0.0) {
      if (wrap) v = wrap_angle(v);
      int b = find_bin(se, B, v);
      if (b >= 0) atomicAdd(&sc[b], 1u);
    }
  }
  __syncthreads();
  for (int i = threadIdx.x; i < B; i += blockDim.x)
    if (sc[i]) atomicAdd(&counts[(i64)c * estride + i], (u64)sc[i]);
}

extern "C" int gb_col_hist(const double* vals, int64_t R, int64_t ld, const double* edges, int estride,
                           const int* nb_host, unsigned long long* counts, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  NB9 nb;
  for (int c = 0; c < 9; ++c) {
    nb.nb[c] = nb_host[c];
    if (nb.nb[c] > HIST_MAXB || nb.nb[c] + 1 > estride) {
      gb_set_error("gb_col_hist: %d bins exceeds limit %d / stride %d", nb.nb[c], HIST_MAXB, estride);
      return -2;
    }
  }
  GB_CK(cudaMemsetAsync(counts, 0, sizeof(u64) * 9 * (size_t)estride, st));
  if (R <= 0) return 0;
  dim3 grid(gb_blocks_for(R, 256 * 8, 148 * 2), 9);
  k_col_hist<<<grid, 256, 0, st>>>(vals, R, ld, edges, estride, nb, (u64*)counts);
  GB_LAUNCH_CK();
  return 0;
}

// ---- residue-major quantisation ------------------------------------------------------------------
// vals rows: 0 0C:1N, 1 N:CA, 2 CA:C, 3 phi, 4 psi, 5 omega, 6 tau, 7 CA:C:1N, 8 C:1N:1CA
// residue-major slots: 0 N:CA(r) 1 CA:C(r) 2 0C:1N(r) 3 tau(r) 4 CA:C:1N(r) 5 psi(r) 6 omega(r) 7 C:1N:1CA(r) 8 phi(r+1)
struct QArgs {
  const double* vals; i64 R; i64 ld; const u32* bb; const i64* offsets;
  const double* edges; int estride; NB9 nb; int bond_mode;
  double init_nca, init_cac, init_tau;
  u16* rbins; u32* glue; u64* sym; int32_t* status;
};

__device__ __forceinline__ u32 quant_one(const double* __restrict__ e, int B, double v, bool wrap, int clamp,
                                         int& st) {
  if (wrap) v = wrap_angle(v);
  int b = find_bin(e, B, v);
  if (b < 0) {
    if (clamp && v == v) {  // strict=False bond lengths, bpe.py:726-729
      if (v < e[0]) return 0u;
      if (v > e[B]) return (u32)(B - 1);
    }
    st |= (b == -1) ? 1 : 2;
    return 0u;
  }
  return (u32)b;
}

__global__ void __launch_bounds__(256) k_quantize_residues(QArgs a) {
  extern __shared__ double se[];  // [9][estride]
  for (int i = threadIdx.x; i < 9 * a.estride; i += blockDim.x) se[i] = a.edges[i];
  __syncthreads();
  const i64 R = a.R, ld = a.ld;
  const double* V = a.vals;
  for (i64 r = (i64)blockIdx.x * blockDim.x + threadIdx.x; r < R; r += (i64)gridDim.x * blockDim.x) {
    u32 s = a.bb[r];
    i64 b0 = a.offsets[s], b1 = a.offsets[s + 1];
    bool first = (r == b0), last = (r == b1 - 1);
    int st = 0;
    u32 bins[9];
    // bonds
    double nca = first ? a.init_nca : V[1 * ld + r - 1];
    double cac = first ? a.init_cac : V[2 * ld + r - 1];
    double tau = first ? a.init_tau : V[6 * ld + r - 1];
    if (a.bond_mode == 0) {
      bins[0] = 0; bins[1] = 0; bins[2] = last ? GB_NOBIN : 0;
    } else {
      int clamp = a.bond_mode == 2;
      bins[0] = quant_one(se + 0 * a.estride, a.nb.nb[0], nca, false, clamp, st);
      bins[1] = quant_one(se + 1 * a.estride, a.nb.nb[1], cac, false, clamp, st);
      bins[2] = last ? GB_NOBIN : quant_one(se + 2 * a.estride, a.nb.nb[2], V[0 * ld + r], false, clamp, st);
    }
    int st2 = 0;   // motif-internal angles (tau, CA:C:1N, psi): reported in bits 2-3 (induce never bins them)
    bins[3] = quant_one(se + 3 * a.estride, a.nb.nb[3], tau, true, 0, st2);
    if (last) {
      bins[4] = bins[5] = bins[6] = bins[7] = bins[8] = GB_NOBIN;
    } else {
      bins[4] = quant_one(se + 4 * a.estride, a.nb.nb[4], V[7 * ld + r], true, 0, st2);
      bins[5] = quant_one(se + 5 * a.estride, a.nb.nb[5], V[4 * ld + r], true, 0, st2);
      bins[6] = quant_one(se + 6 * a.estride, a.nb.nb[6], V[5 * ld + r], true, 0, st);
      bins[7] = quant_one(se + 7 * a.estride, a.nb.nb[7], V[8 * ld + r], true, 0, st);
      bins[8] = quant_one(se + 8 * a.estride, a.nb.nb[8], V[3 * ld + r + 1], true, 0, st);
    }
#pragma unroll
    for (int k = 0; k < 9; ++k) a.rbins[(i64)k * R + r] = (u16)bins[k];
    a.glue[r] = last ? GB_LAST_FLAG : (bins[6] | (bins[7] << 10) | (bins[8] << 20));
    a.sym[r] = 1ull + (u64)(bins[3] | (bins[4] << 10) | (bins[5] << 20));
    st |= st2 << 2;
    if (st) atomicOr(&a.status[s], st);
  }
}

extern "C" int gb_quantize_residues(const double* vals, int64_t R, int64_t ld, const uint32_t* bb,
                                    const int64_t* offsets, int64_t S, const double* edges, int estride,
                                    const int* nb_host, int bond_mode, double init_nca, double init_cac,
                                    double init_tau, uint16_t* rbins, uint32_t* glue, uint64_t* sym,
                                    int32_t* status, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  QArgs a;
  a.vals = vals; a.R = R; a.ld = ld; a.bb = bb; a.offsets = offsets; a.edges = edges; a.estride = estride;
  for (int c = 0; c < 9; ++c) {
    a.nb.nb[c] = nb_host[c];
    if (nb_host[c] >= (int)GB_NOBIN || nb_host[c] + 1 > estride) {
      gb_set_error("gb_quantize_residues: %d bins exceeds 10-bit field / stride", nb_host[c]);
      return -2;
    }
  }
  a.bond_mode = bond_mode;
  a.init_nca = init_nca; a.init_cac = init_cac; a.init_tau = init_tau;
  a.rbins = rbins; a.glue = glue; a.sym = (u64*)sym; a.status = status;
  GB_CK(cudaMemsetAsync(status, 0, sizeof(int32_t) * (size_t)S, st));
  if (R <= 0) return 0;
  size_t smem = sizeof(double) * 9 * (size_t)estride;
  if (smem > 48 * 1024) GB_CK(cudaFuncSetAttribute(k_quantize_residues, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int blocks = gb_blocks_for(R, 256 * 4, 148 * 4);
  k_quantize_residues<<<blocks, 256, smem, st>>>(a);
  GB_LAUNCH_CK();
  return 0;
}
