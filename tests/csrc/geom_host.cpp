// TEST INFRASTRUCTURE: host build (g++) of the __host__ __device__ geometry functions in
// pt-bpe_b200/geobpe_b200/csrc/geom.cuh so their numerics can be checked against the oracle without a GPU.
// Never loaded by the product path.
#include "geom.cuh"
extern "C" {
void t_first_residue(double nca, double cac, double tau, double* out9) {
  gb_v3 a, b, c;
  gb_first_residue(nca, cac, tau, a, b, c);
  out9[0] = a.x; out9[1] = a.y; out9[2] = a.z; out9[3] = b.x; out9[4] = b.y; out9[5] = b.z;
  out9[6] = c.x; out9[7] = c.y; out9[8] = c.z;
}
void t_place(const double* abc9, double ang, double len, double tor, double* out3) {
  gb_v3 d = gb_place_dihedral(v3(abc9[0], abc9[1], abc9[2]), v3(abc9[3], abc9[4], abc9[5]), v3(abc9[6], abc9[7], abc9[8]),
                              ang, len, tor);
  out3[0] = d.x; out3[1] = d.y; out3[2] = d.z;
}
double t_rmsd(const double* P, const double* Q, int A) { return gb_kabsch_rmsd(P, Q, A); }
}
