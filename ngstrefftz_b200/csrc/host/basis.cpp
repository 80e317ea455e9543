#include "basis.hpp"

#include <stdexcept>

namespace ngst {

int twave_nbasis(int D, int order, int fowave) {
  return bincoeff(D + order, order) + bincoeff(D + order - 1, order - 1) - (fowave ? 1 : 0);
}

int index_map2(int d1, const int* index, int ord) {
  int sum = 0, temp_size = 0;
  for (int d = 0; d < d1; d++) {
    for (int p = 0; p < index[d]; p++) sum += bincoeff(d1 - 1 - d + ord - p - temp_size, ord - p - temp_size);
    temp_size += index[d];
  }
  return sum;
}

namespace {
// visit all exponent vectors with |alpha| <= order; loop nesting given by `outer_first`:
// true  -> first coordinate outermost (evaluation order, scalarmappedfe.cpp:526-531)
// false -> last coordinate (time) outermost (TraversePol, trefftzfespace.cpp:455-488)
template <typename F>
void traverse(int d1, int order, bool outer_first, F f) {
  int e[4] = {0, 0, 0, 0};
  // generic odometer over nested loops l0 (outermost) .. l_{d1-1} (innermost)
  int loopvar[4] = {0, 0, 0, 0};
  int depth = 0;
  loopvar[0] = 0;
  while (depth >= 0) {
    int used = 0;
    for (int k = 0; k < depth; k++) used += loopvar[k];
    if (loopvar[depth] > order - used) {
      depth--;
      if (depth >= 0) loopvar[depth]++;
      continue;
    }
    if (depth == d1 - 1) {
      for (int k = 0; k < d1; k++) {
        int coord = outer_first ? k : d1 - 1 - k;
        e[coord] = loopvar[k];
      }
      f(e);
      loopvar[depth]++;
    } else {
      depth++;
      loopvar[depth] = 0;
    }
  }
}
}  // namespace

std::vector<int> monomial_exponents(int d1, int order) {
  std::vector<int> ex;
  traverse(d1, order, true, [&](const int* e) {
    for (int k = 0; k < d1; k++) ex.push_back(e[k]);
  });
  return ex;
}

BasisCsr twave_basis(int d1, int order, int fowave) {
  if (d1 < 2 || d1 > 4) throw std::runtime_error("twave_basis: space-time dimension must be 2..4");
  if (order < 1) throw std::runtime_error("twave_basis: order must be >= 1");
  const int D = d1 - 1;
  const int ndof = bincoeff(D + order, order) + bincoeff(D + order - 1, order - 1);
  const int npoly = bincoeff(d1 + order, order);
  std::vector<double> tb((size_t)ndof * npoly, 0.0);
  for (int basis = 0; basis < ndof; basis++) {
    int tracker = 0;
    double* row = &tb[(size_t)basis * npoly];
    traverse(d1, order, false, [&](const int* coeff) {
      if (tracker >= 0) tracker++;
      const int indexmap = index_map2(d1, coeff, order);
      const int k = coeff[d1 - 1];
      if (k == 0 || k == 1) {
        if (tracker > basis) {
          row[indexmap] = 1;
          tracker = -1;
        }
      } else {
        for (int m = 0; m < d1 - 1; m++) {
          int get_coeff[4];
          for (int q = 0; q < d1; q++) get_coeff[q] = coeff[q];
          get_coeff[d1 - 1] -= 2;
          get_coeff[m] += 2;
          row[indexmap] += (coeff[m] + 1) * (coeff[m] + 2) * row[index_map2(d1, get_coeff, order)];
        }
        const double wavespeed = 1.0;
        row[indexmap] *= wavespeed * wavespeed / (k * (k - 1));
      }
    });
  }
  BasisCsr b;
  b.d1 = d1;
  b.order = order;
  b.npoly = npoly;
  b.expo = monomial_exponents(d1, order);
  // MatToCSR on rows fowave..ndof-1
  int sp = 0;
  for (int i = fowave ? 1 : 0; i < ndof; i++) {
    b.rowptr.push_back(sp);
    for (int j = 0; j < npoly; j++)
      if (tb[(size_t)i * npoly + j] != 0.0) {
        sp++;
        b.col.push_back(j);
        b.val.push_back(tb[(size_t)i * npoly + j]);
      }
  }
  b.rowptr.push_back(sp);
  b.nbasis = ndof - (fowave ? 1 : 0);
  return b;
}

}  // namespace ngst
