// Wave-polynomial (Trefftz) basis table, built seed by seed.
//
// The table is a contract with the reference (rows = TWaveBasis<D+1>::Basis, src/trefftzfespace.cpp:611-682; columns =
// monomials in evaluation order, src/scalarmappedfe.cpp:526-531), not its construction. Here every basis function is
// written down as the Cauchy-Kovalevskaya series of its seed: a wave polynomial (c = 1) is fixed by its traces u(.,0) and
// d_t u(.,0), so the function with seed x^beta t^s, s in {0, 1}, is
//     u = sum_j  C_j(x) t^(s+2j),   C_0 = x^beta,   C_j = Laplace(C_(j-1)) / ((s+2j)(s+2j-1)),
// i.e. one sweep over the time levels s, s+2, ... of that seed only, each level a sparse spatial polynomial held as a list
// of (exponent, coefficient) pairs. The floating-point operation order per coefficient (sum over the space directions in
// ascending order, then one multiplication by 1/(k(k-1))) is the reference's, so the values agree bit for bit with the
// oracle's literal restatement (tests/test_basis_rule.py).
#include "basis.hpp"

#include <array>
#include <stdexcept>

namespace ngst {

int twave_nbasis(int D, int order, int fowave) {
  return bincoeff(D + order, order) + bincoeff(D + order - 1, order - 1) - (fowave ? 1 : 0);
}

namespace {
using Expo = std::array<int, 4>;

// exponent vectors in n variables with total degree <= budget, variable `first` varying slowest and the following ones
// (first + step, ...) faster: lexicographic enumeration by recursion
void enumerate(int n, int budget, int first, int step, Expo& e, int depth, std::vector<Expo>& out) {
  if (depth == n) { out.push_back(e); return; }
  const int var = first + step * depth;
  for (int a = 0; a <= budget; a++) {
    e[var] = a;
    enumerate(n, budget - a, first, step, e, depth + 1, out);
  }
  e[var] = 0;
}

// position of every exponent vector of degree <= order in the evaluation order (dense table over [0, order]^d1)
struct MonomialIndex {
  int d1, order;
  std::vector<int> pos;
  MonomialIndex(int d1_, int order_) : d1(d1_), order(order_) {
    int cells = 1;
    for (int k = 0; k < d1; k++) cells *= order + 1;
    pos.assign(cells, -1);
    std::vector<Expo> all;
    Expo e{};
    enumerate(d1, order, 0, 1, e, 0, all);
    for (size_t i = 0; i < all.size(); i++) pos[cell(all[i])] = (int)i;
  }
  int cell(const Expo& e) const {
    int c = 0;
    for (int k = 0; k < d1; k++) c = c * (order + 1) + e[k];
    return c;
  }
  int operator()(const Expo& e) const { return pos[cell(e)]; }
};
}  // namespace

std::vector<int> monomial_exponents(int d1, int order) {
  std::vector<Expo> all;
  Expo e{};
  enumerate(d1, order, 0, 1, e, 0, all);
  std::vector<int> ex;
  ex.reserve(all.size() * d1);
  for (const Expo& a : all)
    for (int k = 0; k < d1; k++) ex.push_back(a[k]);
  return ex;
}

int monomial_rank(int d1, const int* expo, int order) {
  // closed form of MonomialIndex: exponent vectors that precede `expo` lexicographically
  int rank = 0, left = order;
  for (int d = 0; d < d1; d++) {
    const int rest = d1 - 1 - d;  // variables after this one
    for (int a = 0; a < expo[d]; a++) rank += bincoeff(rest + left - a, rest);
    left -= expo[d];
  }
  return rank;
}

BasisCsr twave_basis(int d1, int order, int fowave) {
  if (d1 < 2 || d1 > 4) throw std::runtime_error("twave_basis: space-time dimension must be 2..4");
  if (order < 1) throw std::runtime_error("twave_basis: order must be >= 1");
  const int D = d1 - 1;
  const MonomialIndex column(d1, order);
  BasisCsr b;
  b.d1 = d1;
  b.order = order;
  b.npoly = bincoeff(d1 + order, order);
  b.expo = monomial_exponents(d1, order);

  struct Term { Expo x; double c; };
  int row = 0;
  for (int s = 0; s <= 1; s++) {
    // seeds x^beta t^s in the reference's numbering: last space coordinate slowest
    std::vector<Expo> seeds;
    Expo e{};
    enumerate(D, order - s, D - 1, -1, e, 0, seeds);
    for (const Expo& beta : seeds) {
      if (row++ < (fowave ? 1 : 0)) continue;  // the first-order system drops the constant (trefftzfespace.cpp:680)
      std::vector<std::pair<int, double>> entries;  // (column, value) of this row
      std::vector<Term> level{{beta, 1.0}};
      for (int k = s;; k += 2) {
        for (const Term& t : level) {
          Expo m = t.x;
          m[D] = k;
          entries.emplace_back(column(m), t.c);
        }
        // next level: Laplace(level) / ((k+2)(k+1)). Its monomials are beta - 2 mu with |mu| = (k + 2 - s) / 2.
        const int kn = k + 2;
        std::vector<Term> next;
        std::vector<Expo> cand;
        Expo g{};
        enumerate(D, order, 0, 1, g, 0, cand);
        const int drop = kn - s;
        for (const Expo& gam : cand) {
          int diff = 0;
          bool ok = true;
          for (int d = 0; d < D; d++) {
            const int r = beta[d] - gam[d];
            ok = ok && r >= 0 && r % 2 == 0;
            diff += r;
          }
          if (!ok || diff != drop) continue;
          double v = 0.0;
          for (int d = 0; d < D; d++) {
            Expo src = gam;
            src[d] += 2;
            double prev = 0.0;
            for (const Term& t : level)
              if (t.x == src) prev = t.c;
            v += (gam[d] + 1) * (gam[d] + 2) * prev;
          }
          v *= 1.0 / (kn * (kn - 1));
          if (v != 0.0) next.push_back({gam, v});
        }
        if (next.empty()) break;
        level.swap(next);
      }
      // CSR rows list their columns in ascending order
      std::vector<std::pair<int, double>> sorted(entries);
      for (size_t i = 1; i < sorted.size(); i++)
        for (size_t j = i; j > 0 && sorted[j].first < sorted[j - 1].first; j--) std::swap(sorted[j], sorted[j - 1]);
      b.rowptr.push_back((int)b.col.size());
      for (const auto& cv : sorted) {
        b.col.push_back(cv.first);
        b.val.push_back(cv.second);
      }
    }
  }
  b.rowptr.push_back((int)b.col.size());
  b.nbasis = (int)b.rowptr.size() - 1;
  if (b.nbasis != twave_nbasis(D, order, fowave)) throw std::logic_error("twave_basis: wrong number of rows");
  return b;
}

}  // namespace ngst
