// Trefftz basis for the wave equation: sparse coefficient matrix monomials -> wave polynomials.
//
// Produces the table of src/trefftzfespace.cpp:611-682 (TWaveBasis<D+1>::Basis, basistype 0) in the CSR form of
// src/scalarmappedfe.cpp:153-175 (MatToCSR) -- same rows, columns and values -- from the Cauchy-Kovalevskaya series of
// each seed monomial (basis.cpp). Monomials xi^alpha, |alpha| <= p, are enumerated with the FIRST coordinate's
// exponent outermost (src/scalarmappedfe.cpp:526-531).
#pragma once
#include <vector>

namespace ngst {

inline int bincoeff(int n, int k) {
  if (k < 0 || k > n) return 0;
  long long r = 1;
  for (int i = 1; i <= k; i++) r = r * (n - k + i) / i;
  return (int)r;
}

struct BasisCsr {
  int d1 = 0, order = 0, nbasis = 0, npoly = 0;
  std::vector<int> rowptr, col;
  std::vector<double> val;
  std::vector<int> expo;  // npoly * d1 monomial exponents in evaluation order
};

// number of Trefftz functions: C(D+p,p) + C(D+p-1,p-1) (twavetents.hpp:160-161), minus 1 for fowave
int twave_nbasis(int D, int order, int fowave);

// exponents of the npoly = C(d1+p,p) monomials, first coordinate outermost
std::vector<int> monomial_exponents(int d1, int order);

// position of the monomial xi^expo in monomial_exponents(d1, order)
int monomial_rank(int d1, const int* expo, int order);

// wave polynomials for c = 1 in d1 = D+1 space-time dimensions; fowave drops the constant (row 0)
BasisCsr twave_basis(int d1, int order, int fowave);

}  // namespace ngst
