// Simplex quadrature rules of order 2p on the reference segment / triangle / tetrahedron.
//
// Stands in for NGSolve's SIMD_IntegrationRule(eltyp, 2*order)
// (reference call sites src/twavetents.cpp:77,739,771,807,828,972,1042). NGSolve is not
// vendored in the reference; the convention restated here is:
//   SEGM: Gauss-Legendre on [0,1], n = order/2+1 points;
//   TRIG/TET: Duffy-collapsed tensor rules, Gauss-Jacobi (alpha = D-1-k in direction k) x Gauss,
//   n = order/2+1 points per direction, first coordinate outermost, sum of weights 1/D!.
// The 125-point tet rule at order 8 is confirmed by the reference's own flop counter
// (test/tents.py:282). Point ORDER beyond that is unpinned.
#pragma once
#include <vector>

namespace ngst {

struct SimplexRule {
  int dim = 0;             // dimension of the reference simplex
  int nip = 0;
  std::vector<double> xi;  // nip * dim reference coordinates
  std::vector<double> w;   // nip weights
};

// n-point Gauss-Jacobi rule for weight (1-x)^alpha on [0,1] (alpha = 0: Gauss-Legendre).
void gauss_jacobi01(int n, int alpha, std::vector<double>& x, std::vector<double>& w);

SimplexRule make_simplex_rule(int dim, int order);

}  // namespace ngst
