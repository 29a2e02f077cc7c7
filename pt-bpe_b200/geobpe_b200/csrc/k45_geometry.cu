// K4 batched NeRF window reconstruction, K5 Kabsch-RMSD medoid assignment, K6 all-pairs RMSD.
// Replaces Tokenizer.compute_coords (tokenizer.py:347-363) -> geo_nerf (317-344) -> NERFBuilder
// (nerf.py:84-128), BPE._compute_assignment(_inner) (bpe.py:646-657) and the distance matrix of
// k_medoids (algo.py:170-181).  float64 CUDA-core math: token ids are np.argmin over float64 RMSDs in the
// reference, so the first build keeps float64 to stay on the same side of every near-tie.
#include "common.cuh"
#include "geom.cuh"

struct WArgs {
  const double* V; i64 R; i64 ld; const u32* bb; const i64* offsets;
  const u32* r0; i64 Q; int nres; int A;
  const u32* f3cls; const double* f3tab; double d_nca, d_cac, d_tau;
  const u16* rbins; const double* bc; int bc_stride;
  double* coords;
};

__global__ void __launch_bounds__(128) k_window_coords(WArgs a) {
  i64 q = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= a.Q) return;
  const i64 ld = a.ld, R = a.R;
  const double* V = a.V;
  i64 r0 = a.r0[q];
  u32 s = a.bb[r0];
  bool first = (r0 == a.offsets[s]);
  double nca, cac, tau;
  if (first) {
    if (a.f3cls && a.f3cls[q] != 0xFFFFFFFFu) {
      const double* f = a.f3tab + 3 * (i64)a.f3cls[q];
      nca = f[0]; cac = f[1]; tau = f[2];
    } else { nca = a.d_nca; cac = a.d_cac; tau = a.d_tau; }
  } else {
    nca = V[1 * ld + r0 - 1]; cac = V[2 * ld + r0 - 1]; tau = V[6 * ld + r0 - 1];
  }
  if (a.bc) {  // quantised bond lengths (induce): centre of the residue's bond bins
    nca = a.bc[0 * a.bc_stride + a.rbins[0 * R + r0]];
    cac = a.bc[1 * a.bc_stride + a.rbins[1 * R + r0]];
  }
  double* out = a.coords + q * (i64)a.A * 3;
  gb_v3 p0, p1, p2;
  gb_first_residue(nca, cac, tau, p0, p1, p2);
  int na = 0;
  if (na < a.A) { out[0] = p0.x; out[1] = p0.y; out[2] = p0.z; ++na; }
  if (na < a.A) { out[3] = p1.x; out[4] = p1.y; out[5] = p1.z; ++na; }
  if (na < a.A) { out[6] = p2.x; out[7] = p2.y; out[8] = p2.z; ++na; }
  for (int i = 0; na < a.A; ++i) {
    i64 r = r0 + i;
    double len, ang, tor;
    // N(i+1): 0C:1N, CA:C:1N, psi of residue r
    len = a.bc ? a.bc[2 * a.bc_stride + a.rbins[2 * R + r]] : V[0 * ld + r];
    ang = V[7 * ld + r]; tor = V[4 * ld + r];
    gb_v3 n = gb_place_dihedral(p0, p1, p2, ang, len, tor);
    out[3 * na + 0] = n.x; out[3 * na + 1] = n.y; out[3 * na + 2] = n.z; ++na;
    if (na >= a.A) break;
    // CA(i+1): N:CA of residue r+1 (row r), C:1N:1CA, omega of residue r
    len = a.bc ? a.bc[0 * a.bc_stride + a.rbins[0 * R + r + 1]] : V[1 * ld + r];
    ang = V[8 * ld + r]; tor = V[5 * ld + r];
    gb_v3 ca = gb_place_dihedral(p1, p2, n, ang, len, tor);
    out[3 * na + 0] = ca.x; out[3 * na + 1] = ca.y; out[3 * na + 2] = ca.z; ++na;
    if (na >= a.A) break;
    // C(i+1): CA:C, tau of residue r+1 (row r), phi of residue r+1 (row r+1)
    len = a.bc ? a.bc[1 * a.bc_stride + a.rbins[1 * R + r + 1]] : V[2 * ld + r];
    ang = V[6 * ld + r]; tor = V[3 * ld + r + 1];
    gb_v3 c = gb_place_dihedral(p2, n, ca, ang, len, tor);
    out[3 * na + 0] = c.x; out[3 * na + 1] = c.y; out[3 * na + 2] = c.z; ++na;
    p0 = n; p1 = ca; p2 = c;
  }
}

extern "C" int gb_window_coords(const double* vals, int64_t R, int64_t ld, const uint32_t* bb, const int64_t* offsets,
                                const uint32_t* r0, int64_t Q, int nres, int A, const uint32_t* f3cls,
                                const double* f3tab, double d_nca, double d_cac, double d_tau, const uint16_t* rbins,
                                const double* bond_centres, int bc_stride, double* coords, void* stream) {
  if (Q <= 0) return 0;
  if (A < 3 || (A != 3 * nres && A != 3 * nres + 1)) {
    gb_set_error("gb_window_coords: A=%d inconsistent with nres=%d", A, nres);
    return -2;
  }
  if (bond_centres && !rbins) { gb_set_error("gb_window_coords: bond_centres needs rbins"); return -2; }
  WArgs a{vals, R, ld, bb, offsets, r0, Q, nres, A, f3cls, f3tab, d_nca, d_cac, d_tau, rbins, bond_centres, bc_stride, coords};
  k_window_coords<<<(unsigned)((Q + 127) / 128), 128, 0, (cudaStream_t)stream>>>(a);
  GB_LAUNCH_CK();
  return 0;
}

// ---- K5: assignment ----------------------------------------------------------------------------
// thread per occurrence; medoid coordinates staged in shared memory (broadcast reads)
__global__ void __launch_bounds__(128) k_kabsch_assign(const double* __restrict__ P, i64 Q, int A,
                                                        const double* __restrict__ Mq, int K, u32* __restrict__ assign,
                                                        double* __restrict__ best) {
  extern __shared__ double sm[];  // [Kc][A][3] chunk of medoids
  const int per = A * 3;
  const int Kc_max = (int)((48 * 1024 / sizeof(double)) / per);
  i64 q = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  const double* p = P + (q < Q ? q : 0) * (i64)per;
  double bv = INFINITY;
  u32 bi = 0;
  for (int k0 = 0; k0 < K; k0 += Kc_max) {
    int kc = min(Kc_max, K - k0);
    __syncthreads();
    for (int i = threadIdx.x; i < kc * per; i += blockDim.x) sm[i] = Mq[(i64)k0 * per + i];
    __syncthreads();
    if (q < Q) {
      for (int m = 0; m < kc; ++m) {
        double d = gb_kabsch_rmsd(p, sm + m * per, A);
        if (d < bv) { bv = d; bi = (u32)(k0 + m); }   // strict <: first index on ties (np.argmin)
      }
    }
  }
  if (q < Q) { assign[q] = bi; if (best) best[q] = bv; }
}

extern "C" int gb_kabsch_assign(const double* P, int64_t Q, int A, const double* Mq, int K, uint32_t* assign,
                                double* best, void* stream) {
  if (Q <= 0) return 0;
  if (K <= 0 || A < 1) { gb_set_error("gb_kabsch_assign: K=%d A=%d", K, A); return -2; }
  int per = A * 3;
  int Kc_max = (int)((48 * 1024 / sizeof(double)) / per);
  if (Kc_max < 1) { gb_set_error("gb_kabsch_assign: window of %d atoms too large for shared staging", A); return -2; }
  int kc = K < Kc_max ? K : Kc_max;
  size_t smem = sizeof(double) * (size_t)kc * per;
  k_kabsch_assign<<<(unsigned)((Q + 127) / 128), 128, smem, (cudaStream_t)stream>>>(P, Q, A, Mq, K, assign, best);
  GB_LAUNCH_CK();
  return 0;
}

// ---- K6: all-pairs RMSD matrix -----------------------------------------------------------------
__global__ void __launch_bounds__(128) k_rmsd_matrix(const double* __restrict__ P, int N, int A, float* __restrict__ D) {
  i64 t = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= (i64)N * N) return;
  int i = (int)(t / N), j = (int)(t % N);
  if (j < i) return;
  // compute_rmsd(strucs[i], strucs[j]); the diagonal is ~1e-16 in the reference (exact 0 after float32 sums),
  // here E0 - 2*sum(sigma) cancels to ~1e-8, so it is pinned to 0
  double d = (i == j) ? 0.0 : gb_kabsch_rmsd(P + (i64)i * A * 3, P + (i64)j * A * 3, A);
  float f = (float)d;
  D[(i64)i * N + j] = f;
  D[(i64)j * N + i] = f;
}
extern "C" int gb_rmsd_matrix(const double* P, int N, int A, float* D, void* stream) {
  if (N <= 0) return 0;
  i64 tot = (i64)N * N;
  k_rmsd_matrix<<<(unsigned)((tot + 127) / 128), 128, 0, (cudaStream_t)stream>>>(P, N, A, D);
  GB_LAUNCH_CK();
  return 0;
}
