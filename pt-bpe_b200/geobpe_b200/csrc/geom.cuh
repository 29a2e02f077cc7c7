// NeRF forward kinematics and Kabsch RMSD in float64, usable from device code and (for the CPU unit
// tests in tests/csrc) from host code.  Formulas follow the reference:
//   angles_and_coords.py:238-317 update_backbone_positions, 232-235 rotate_vector
//   nerf.py:151-210 place_dihedral (numpy branch), nerf.py:84-128 NERFBuilder.cartesian_coords
//   algo.py:8-45 kabsch, 48-65 compute_rmsd
#pragma once
#include <math.h>
#ifdef __CUDACC__
#define GB_HD __host__ __device__ __forceinline__
#else
#define GB_HD inline
#endif

struct gb_v3 { double x, y, z; };
GB_HD gb_v3 v3(double x, double y, double z) { gb_v3 r; r.x = x; r.y = y; r.z = z; return r; }
GB_HD gb_v3 operator+(gb_v3 a, gb_v3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
GB_HD gb_v3 operator-(gb_v3 a, gb_v3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
GB_HD gb_v3 operator*(gb_v3 a, double s) { return v3(a.x * s, a.y * s, a.z * s); }
GB_HD double dot3(gb_v3 a, gb_v3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
GB_HD gb_v3 cross3(gb_v3 a, gb_v3 b) { return v3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
GB_HD double norm3(gb_v3 a) { return sqrt(dot3(a, a)); }

// nerf.py:22-24
#define GB_N_INIT v3(17.047, 14.099, 3.625)
#define GB_CA_INIT v3(16.967, 12.784, 4.338)
#define GB_C_INIT v3(15.685, 12.755, 5.133)

// first residue: move CA along the C->CA line to length L_CA_C, rotate N about (vecN x vecC) so that the
// N-CA-C angle becomes theta, rescale to L_N_CA (angles_and_coords.py:238-317)
GB_HD void gb_first_residue(double L_N_CA, double L_CA_C, double theta, gb_v3& N, gb_v3& CA, gb_v3& C) {
  gb_v3 N0 = GB_N_INIT, CA0 = GB_CA_INIT, C0 = GB_C_INIT;
  gb_v3 v = CA0 - C0;
  v = v * (1.0 / norm3(v));
  gb_v3 CAn = C0 + v * L_CA_C;
  gb_v3 vN = N0 - CAn, vC = C0 - CAn;
  double ct = dot3(vN, vC) / (norm3(vN) * norm3(vC));
  ct = ct < -1.0 ? -1.0 : (ct > 1.0 ? 1.0 : ct);
  double dtheta = theta - acos(ct);
  gb_v3 axis = cross3(vN, vC);
  axis = axis * (1.0 / norm3(axis));
  double ang = -dtheta, ca = cos(ang), sa = sin(ang);
  gb_v3 rot = vN * ca + cross3(axis, vN) * sa + axis * (dot3(axis, vN) * (1.0 - ca));
  rot = rot * (1.0 / norm3(rot)) * L_N_CA;
  N = CAn + rot; CA = CAn; C = C0;
}

// nerf.py:182-210
GB_HD gb_v3 gb_place_dihedral(gb_v3 a, gb_v3 b, gb_v3 c, double bond_angle, double bond_length, double torsion) {
  gb_v3 ab = b - a;
  gb_v3 bc = c - b;
  bc = bc * (1.0 / norm3(bc));
  gb_v3 n = cross3(ab, bc);
  n = n * (1.0 / norm3(n));
  gb_v3 nbc = cross3(n, bc);
  double d0 = -bond_length * cos(bond_angle);
  double d1 = bond_length * cos(torsion) * sin(bond_angle);
  double d2 = bond_length * sin(torsion) * sin(bond_angle);
  return v3(bc.x * d0 + nbc.x * d1 + n.x * d2 + c.x,
            bc.y * d0 + nbc.y * d1 + n.y * d2 + c.y,
            bc.z * d0 + nbc.z * d1 + n.z * d2 + c.z);
}

// singular values of a 3x3 matrix by one-sided (Hestenes) Jacobi: high relative accuracy, no LAPACK
GB_HD void gb_singular_values3(double G[3][3], double s[3]) {
  for (int sweep = 0; sweep < 10; ++sweep) {
    bool rotated = false;
    for (int p = 0; p < 2; ++p)
      for (int q = p + 1; q < 3; ++q) {
        double alpha = G[0][p] * G[0][p] + G[1][p] * G[1][p] + G[2][p] * G[2][p];
        double beta = G[0][q] * G[0][q] + G[1][q] * G[1][q] + G[2][q] * G[2][q];
        double gamma = G[0][p] * G[0][q] + G[1][p] * G[1][q] + G[2][p] * G[2][q];
        if (gamma != 0.0 && fabs(gamma) > 1e-17 * sqrt(alpha * beta)) {
          rotated = true;
          double zeta = (beta - alpha) / (2.0 * gamma);
          double t = (zeta >= 0.0 ? 1.0 : -1.0) / (fabs(zeta) + sqrt(1.0 + zeta * zeta));
          double c = 1.0 / sqrt(1.0 + t * t), sn = c * t;
          for (int i = 0; i < 3; ++i) {
            double gp = G[i][p], gq = G[i][q];
            G[i][p] = c * gp - sn * gq;
            G[i][q] = sn * gp + c * gq;
          }
        }
      }
    if (!rotated) break;
  }
  for (int p = 0; p < 3; ++p) s[p] = sqrt(G[0][p] * G[0][p] + G[1][p] * G[1][p] + G[2][p] * G[2][p]);
}

// RMSD after optimal superposition of Q onto P (A atoms, xyz interleaved, strides in doubles)
GB_HD double gb_kabsch_rmsd(const double* P, const double* Q, int A) {
  double cp[3] = {0, 0, 0}, cq[3] = {0, 0, 0};
  for (int a = 0; a < A; ++a)
    for (int d = 0; d < 3; ++d) { cp[d] += P[3 * a + d]; cq[d] += Q[3 * a + d]; }
  for (int d = 0; d < 3; ++d) { cp[d] /= A; cq[d] /= A; }
  double H[3][3] = {{0, 0, 0}, {0, 0, 0}, {0, 0, 0}};
  double e0 = 0.0;
  for (int a = 0; a < A; ++a) {
    double p[3], q[3];
    for (int d = 0; d < 3; ++d) { p[d] = P[3 * a + d] - cp[d]; q[d] = Q[3 * a + d] - cq[d]; }
    for (int i = 0; i < 3; ++i) {
      e0 += p[i] * p[i] + q[i] * q[i];
      for (int j = 0; j < 3; ++j) H[i][j] += p[i] * q[j];
    }
  }
  double det = H[0][0] * (H[1][1] * H[2][2] - H[1][2] * H[2][1]) - H[0][1] * (H[1][0] * H[2][2] - H[1][2] * H[2][0]) +
               H[0][2] * (H[1][0] * H[2][1] - H[1][1] * H[2][0]);
  double s[3];
  gb_singular_values3(H, s);
  double smin = s[0] < s[1] ? (s[0] < s[2] ? s[0] : s[2]) : (s[1] < s[2] ? s[1] : s[2]);
  double sum = s[0] + s[1] + s[2];
  if (det < 0.0) sum -= 2.0 * smin;
  double msd = (e0 - 2.0 * sum) / A;
  return msd > 0.0 ? sqrt(msd) : 0.0;
}
