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

// The rule the solver samples the state at: an installed override (below) if there is one, else default_simplex_rule.
SimplexRule make_simplex_rule(int dim, int order);
// The restated NGSolve convention described above, never overridden.
SimplexRule default_simplex_rule(int dim, int order);

// Rule tables dumped from a real NGSolve install: replaces the rule of (dim, order) for every solver
// created afterwards, so that wavefront columns correspond point for point to the reference's. The table must have the
// default rule's number of points (the kernels are sized for it), positive weights, and integrate every monomial of
// total degree <= order exactly (checked to 1e-11 relative; throws std::invalid_argument otherwise). nip == 0 removes
// the override.
void set_simplex_rule_override(int dim, int order, int nip, const double* xi, const double* w);

// Points of the compact rule make_compact_rule(dim, degree) returns, or 0 when the table has none (then the Duffy rule of
// that order is used). constexpr: the fast tent kernel sizes its top-face pass with it at compile time.
constexpr int compact_rule_points(int dim, int degree) {
  if (dim == 2) return degree == 4 ? 6 : degree == 6 ? 12 : degree == 8 ? 16 : degree == 10 ? 25 : 0;
  if (dim == 3) return degree == 2 ? 4 : degree == 4 ? 14 : degree == 6 ? 24 : degree == 8 ? 46 : 0;
  return 0;
}

// Fully symmetric interior rule with positive weights that integrates polynomials of total degree <= `degree` exactly with
// (near-)minimal point count (tables generated and verified by scripts/gen_compact_rules.py: 24 points for degree 6 on the
// tetrahedron where the Duffy rule has 64). Not a reference rule: used only where the integrand is a polynomial of known
// degree (top-face Gram of the fast kernel), so the result equals the reference's up to round-off. Falls back to
// make_simplex_rule(dim, degree) when the table has no entry.
SimplexRule make_compact_rule(int dim, int degree);

// Transfer of point values from rule A to rule B through the polynomials of total degree <= deg: the nB x nA matrix
//   M = Psi(B) Psi(A)^T diag(w_A),   Psi = basis of P_deg orthonormal in A's inner product,
// i.e. (M g)(y_r) = (Pi g)(y_r) with Pi the projection onto P_deg that is orthogonal in A's inner product. If A and B are
// exact to degree 2 deg, then for EVERY g (polynomial or not) and every phi in P_deg
//   sum_q wA_q phi(x_q) g_q = sum_r wB_r phi(y_r) (M g)_r,
// and M g = g restricted to B's points when g is itself in P_deg. This is what lets the right-hand side of a tent
// (test functions of degree <= p against the state row, src/twavetents.cpp:230-267) be integrated on the compact rule
// (DESIGN.md section 9). Row-major M[r * nA + q].
std::vector<double> rule_transfer_matrix(const SimplexRule& A, const SimplexRule& B, int deg);

}  // namespace ngst
