// f-4 reconstruction metrics for ALL backbones on the device: what bin/encode.py:341-359 computes per reference
// tokenizer -- Tokenizer.compute_coords() of the current (quantised) geometry against the pristine structure, Kabsch backbone
// RMSD (algo.py:8-65) -- as four passes over resident arrays:
//   k_token_of_residue   residue -> live token
//   k_current_slots      the CURRENT value of every residue-major slot: bin centres (histogram mode), the token's medoid geometry
//                        (medoid mode: class table cg f64 [ncls][maxres][9]) with snapped glue bins at token ends
//   k_vals_to_slots      the pristine values in the same residue-major layout
//   k_chain_coords       full-chain NeRF per backbone (nerf.py:84-210, same operation order as k_window_coords)
//   k_chain_rmsd         Kabsch RMSD of two coordinate sets per backbone
// Residue-major slots: 0 N:CA(r) 1 CA:C(r) 2 0C:1N(r) 3 tau(r) 4 CA:C:1N(r) 5 psi(r) 6 omega(r) 7 C:1N:1CA(r) 8 phi(r+1).
#include "common.cuh"
#include "geom.cuh"

__global__ void __launch_bounds__(256) k_token_of_residue(const u32* __restrict__ pos, const u32* __restrict__ len, i64 T, u32* __restrict__ tok_of) {
  for (i64 j = (i64)blockIdx.x * blockDim.x + threadIdx.x; j < T; j += (i64)gridDim.x * blockDim.x) {
    const u32 p = pos[j], n = len[j] & 0xFFFFu;
    for (u32 k = 0; k < n; ++k) tok_of[p + k] = (u32)j;
  }
}

// centres f64 [9][cstride]: bin centre of every residue-major slot (bond slots: the standard lengths at index 0 when std_bonds)
__global__ void __launch_bounds__(256) k_current_slots(const u16* __restrict__ rbins, i64 R, const double* __restrict__ centres, int cstride,
                                                        const u32* __restrict__ tok_of, const u32* __restrict__ pos, const u32* __restrict__ len,
                                                        const u32* __restrict__ cls, const double* __restrict__ cg, int maxres,
                                                        double* __restrict__ cur) {
  const double nan = __longlong_as_double(0x7ff8000000000000ll);
  for (i64 r = (i64)blockIdx.x * blockDim.x + threadIdx.x; r < R; r += (i64)gridDim.x * blockDim.x) {
    u32 k = 0, n = 1, c = 0;
    if (cg) { const u32 j = tok_of[r]; k = (u32)r - pos[j]; n = len[j] & 0xFFFFu; c = cls[j]; }
#pragma unroll
    for (int s = 0; s < 9; ++s) {
      double v;
      if (cg && !(k == n - 1u && s >= 6) && (int)k < maxres) v = cg[((i64)c * maxres + k) * 9 + s];
      else {
        const u16 b = rbins[(i64)s * R + r];
        v = b == GB_NOBIN ? nan : centres[s * cstride + b];
      }
      cur[(i64)s * R + r] = v;
    }
  }
}

// vals f64 [9][ld] type-major (0C:1N, N:CA, CA:C, phi, psi, omega, tau, CA:C:1N, C:1N:1CA; row r = junction r / residue r+1)
__global__ void __launch_bounds__(256) k_vals_to_slots(const double* __restrict__ V, i64 R, i64 ld, const u32* __restrict__ bb,
                                                        const i64* __restrict__ offsets, double i_nca, double i_cac, double i_tau,
                                                        double* __restrict__ out) {
  for (i64 r = (i64)blockIdx.x * blockDim.x + threadIdx.x; r < R; r += (i64)gridDim.x * blockDim.x) {
    const bool first = r == offsets[bb[r]];
    out[0 * R + r] = first ? i_nca : V[1 * ld + r - 1];
    out[1 * R + r] = first ? i_cac : V[2 * ld + r - 1];
    out[3 * R + r] = first ? i_tau : V[6 * ld + r - 1];
    out[2 * R + r] = V[0 * ld + r];
    out[4 * R + r] = V[7 * ld + r];
    out[5 * R + r] = V[4 * ld + r];
    out[6 * R + r] = V[5 * ld + r];
    out[7 * R + r] = V[8 * ld + r];
    out[8 * R + r] = (r + 1 < R) ? V[3 * ld + r + 1] : __longlong_as_double(0x7ff8000000000000ll);
  }
}

// coords f64 [3 R][3]: atoms N, CA, C of every residue, backbone by backbone
__global__ void __launch_bounds__(64) k_chain_coords(const double* __restrict__ sl, i64 R, const i64* __restrict__ offsets, i64 S,
                                                      double* __restrict__ coords) {
  const i64 b = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= S) return;
  const i64 r0 = offsets[b], n = offsets[b + 1] - r0;
  if (n <= 0) return;
  gb_v3 p0, p1, p2;
  gb_first_residue(sl[0 * R + r0], sl[1 * R + r0], sl[3 * R + r0], p0, p1, p2);
  double* out = coords + 9 * r0;
  out[0] = p0.x; out[1] = p0.y; out[2] = p0.z; out[3] = p1.x; out[4] = p1.y; out[5] = p1.z; out[6] = p2.x; out[7] = p2.y; out[8] = p2.z;
  for (i64 i = 0; i + 1 < n; ++i) {
    const i64 r = r0 + i;
    const gb_v3 nn = gb_place_dihedral(p0, p1, p2, sl[4 * R + r], sl[2 * R + r], sl[5 * R + r]);
    const gb_v3 ca = gb_place_dihedral(p1, p2, nn, sl[7 * R + r], sl[0 * R + r + 1], sl[6 * R + r]);
    const gb_v3 cc = gb_place_dihedral(p2, nn, ca, sl[3 * R + r + 1], sl[1 * R + r + 1], sl[8 * R + r]);
    double* o = out + 9 * (i + 1);
    o[0] = nn.x; o[1] = nn.y; o[2] = nn.z; o[3] = ca.x; o[4] = ca.y; o[5] = ca.z; o[6] = cc.x; o[7] = cc.y; o[8] = cc.z;
    p0 = nn; p1 = ca; p2 = cc;
  }
}

__global__ void __launch_bounds__(64) k_chain_rmsd(const double* __restrict__ A, const double* __restrict__ B, const i64* __restrict__ offsets,
                                                    i64 S, double* __restrict__ rmsd) {
  const i64 b = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= S) return;
  const i64 r0 = offsets[b], n = offsets[b + 1] - r0;
  rmsd[b] = n > 0 ? gb_kabsch_rmsd(A + 9 * r0, B + 9 * r0, (int)(3 * n)) : 0.0;
}

extern "C" int gb_reconstruction(const uint16_t* rbins, int64_t R, const double* centres, int cstride, const uint32_t* pos,
                                 const uint32_t* len, const uint32_t* cls, int64_t T, const double* cg, int maxres,
                                 const double* vals, int64_t ld, const uint32_t* bb, const int64_t* offsets, int64_t S,
                                 double i_nca, double i_cac, double i_tau, uint32_t* tok_of, double* slots_cur, double* slots_orig,
                                 double* coords_cur, double* coords_orig, double* rmsd, void* stream) {
  if (R <= 0 || S <= 0) return 0;
  cudaStream_t s = (cudaStream_t)stream;
  int sms = gb_sm_count();
  if (sms <= 0) sms = 148;
  if (cg) k_token_of_residue<<<gb_blocks_for(T, 256, sms * 8), 256, 0, s>>>(pos, len, T, tok_of);
  k_current_slots<<<gb_blocks_for(R, 256, sms * 8), 256, 0, s>>>(rbins, R, centres, cstride, tok_of, pos, len, cls, cg, maxres, slots_cur);
  k_vals_to_slots<<<gb_blocks_for(R, 256, sms * 8), 256, 0, s>>>(vals, R, ld, bb, offsets, i_nca, i_cac, i_tau, slots_orig);
  const unsigned nb = (unsigned)((S + 63) / 64);
  k_chain_coords<<<nb, 64, 0, s>>>(slots_cur, R, offsets, S, coords_cur);
  k_chain_coords<<<nb, 64, 0, s>>>(slots_orig, R, offsets, S, coords_orig);
  k_chain_rmsd<<<nb, 64, 0, s>>>(coords_cur, coords_orig, offsets, S, rmsd);
  GB_LAUNCH_CK();
  return 0;
}

// ---- lDDT-C-alpha of every backbone (bin/encode.py:351-358, plotting.py:436: chain_recon.lddt_ca(orig_chain).mean()) --------------
// The reference calls esm's ProteinChain.lddt_ca (esm is a dependency absent from the reference tree and from this image); its
// published algorithm (esm.utils.structure.metrics.compute_lddt, the OpenFold / AlphaFold lDDT) is restated: float32 coordinates,
// d = sqrt(1e-10 + |xi - xj|^2) for both chains, pairs j != i with d_true < cutoff are scored 1/4 per threshold 0.5 / 1 / 2 / 4 A that
// |d_true - d_pred| stays under, residue score = (1e-10 + sum) / (1e-10 + pairs), chain value = mean over its residues.
// One CTA per backbone; the C-alpha atoms of both chains go through shared memory in tiles.
#define LD_THREADS 128
#define LD_TILE 256
__global__ void __launch_bounds__(LD_THREADS) k_lddt_ca(const double* __restrict__ pred, const double* __restrict__ truth,
                                                        const i64* __restrict__ offsets, i64 S, float cutoff, double* __restrict__ out) {
  __shared__ float tp[LD_TILE][3], tt[LD_TILE][3];
  __shared__ double wsum[LD_THREADS / 32];
  for (i64 b = blockIdx.x; b < S; b += gridDim.x) {
    const i64 r0 = offsets[b];
    const int n = (int)(offsets[b + 1] - r0);
    double acc = 0.0;                              // this thread's share of the sum of residue scores
    for (int i0 = 0; i0 < n; i0 += LD_THREADS) {
      const int i = i0 + threadIdx.x;
      float px = 0, py = 0, pz = 0, qx = 0, qy = 0, qz = 0;
      if (i < n) {                                 // atom 1 of a residue's three (N, CA, C)
        const double* p = pred + 9 * (r0 + i) + 3;
        const double* q = truth + 9 * (r0 + i) + 3;
        px = (float)p[0]; py = (float)p[1]; pz = (float)p[2];
        qx = (float)q[0]; qy = (float)q[1]; qz = (float)q[2];
      }
      float cnt = 0.f, sc = 0.f;
      for (int j0 = 0; j0 < n; j0 += LD_TILE) {
        __syncthreads();
        for (int t = threadIdx.x; t < LD_TILE && j0 + t < n; t += LD_THREADS) {
          const double* p = pred + 9 * (r0 + j0 + t) + 3;
          const double* q = truth + 9 * (r0 + j0 + t) + 3;
          tp[t][0] = (float)p[0]; tp[t][1] = (float)p[1]; tp[t][2] = (float)p[2];
          tt[t][0] = (float)q[0]; tt[t][1] = (float)q[1]; tt[t][2] = (float)q[2];
        }
        __syncthreads();
        if (i < n) {
          const int m = min(LD_TILE, n - j0);
          for (int t = 0; t < m; ++t) {
            if (j0 + t == i) continue;
            // (no FMA contraction: a distance one ulp off could land on the other side of a threshold)
            const float ax = qx - tt[t][0], ay = qy - tt[t][1], az = qz - tt[t][2];
            const float dt = __fsqrt_rn(__fadd_rn(1e-10f, __fadd_rn(__fadd_rn(__fmul_rn(ax, ax), __fmul_rn(ay, ay)), __fmul_rn(az, az))));
            if (!(dt < cutoff)) continue;
            const float bx = px - tp[t][0], by = py - tp[t][1], bz = pz - tp[t][2];
            const float dp = __fsqrt_rn(__fadd_rn(1e-10f, __fadd_rn(__fadd_rn(__fmul_rn(bx, bx), __fmul_rn(by, by)), __fmul_rn(bz, bz))));
            const float l1 = fabsf(dt - dp);
            cnt += 1.f;
            sc += 0.25f * (float)((l1 < 0.5f) + (l1 < 1.0f) + (l1 < 2.0f) + (l1 < 4.0f));
          }
        }
      }
      if (i < n) acc += (double)((1e-10f + sc) * (1.0f / (1e-10f + cnt)));
    }
    for (int o = 16; o; o >>= 1) acc += __shfl_xor_sync(0xFFFFFFFFu, acc, o);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) wsum[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
      double t = 0.0;
      for (int w = 0; w < LD_THREADS / 32; ++w) t += wsum[w];
      out[b] = n > 0 ? t / n : 0.0;
    }
  }
}

extern "C" int gb_lddt_ca(const double* coords_pred, const double* coords_true, const int64_t* offsets, int64_t S, double cutoff,
                          double* lddt, void* stream) {
  if (S <= 0) return 0;
  int sms = gb_sm_count();
  if (sms <= 0) sms = 148;
  k_lddt_ca<<<gb_blocks_for(S, 1, sms * 16), LD_THREADS, 0, (cudaStream_t)stream>>>(coords_pred, coords_true, offsets, S, (float)cutoff, lddt);
  GB_LAUNCH_CK();
  return 0;
}
