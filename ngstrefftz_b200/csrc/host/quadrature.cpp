#include "quadrature.hpp"

#include <algorithm>
#include <cmath>
#include <map>
#include <mutex>
#include <stdexcept>
#include <string>
#include <utility>

namespace ngst {

namespace {
// Jacobi polynomial P_n^{(a,b)}(t) by the three-term recurrence, in long double.
long double jacobi_p(int n, int a, int b, long double t) {
  if (n == 0) return 1.0L;
  long double p0 = 1.0L;
  long double p1 = 0.5L * (a - b) + 0.5L * (a + b + 2) * t;
  for (int k = 1; k < n; k++) {
    long double c = 2.0L * k + a + b;
    long double a1 = 2.0L * (k + 1) * (k + a + b + 1) * c;
    long double a2 = (c + 1) * (a * a - b * b);
    long double a3 = c * (c + 1) * (c + 2);
    long double a4 = 2.0L * (k + a) * (k + b) * (c + 2);
    long double p2 = ((a2 + a3 * t) * p1 - a4 * p0) / a1;
    p0 = p1;
    p1 = p2;
  }
  return p1;
}
long double jacobi_dp(int n, int a, int b, long double t) {
  if (n == 0) return 0.0L;
  return 0.5L * (n + a + b + 1) * jacobi_p(n - 1, a + 1, b + 1, t);
}

// eigenvalues of a small dense symmetric matrix by cyclic Jacobi rotations
std::vector<long double> sym_eigenvalues(std::vector<long double> A, int n) {
  for (int sweep = 0; sweep < 100; sweep++) {
    long double off = 0;
    for (int i = 0; i < n; i++)
      for (int j = i + 1; j < n; j++) off += A[i * n + j] * A[i * n + j];
    if (off < 1e-40L) break;
    for (int p = 0; p < n; p++)
      for (int q = p + 1; q < n; q++) {
        long double apq = A[p * n + q];
        if (fabsl(apq) < 1e-300L) continue;
        long double theta = (A[q * n + q] - A[p * n + p]) / (2 * apq);
        long double t = (theta >= 0 ? 1.0L : -1.0L) / (fabsl(theta) + sqrtl(theta * theta + 1));
        long double c = 1 / sqrtl(t * t + 1), s = t * c;
        for (int k = 0; k < n; k++) {
          long double akp = A[k * n + p], akq = A[k * n + q];
          A[k * n + p] = c * akp - s * akq;
          A[k * n + q] = s * akp + c * akq;
        }
        for (int k = 0; k < n; k++) {
          long double apk = A[p * n + k], aqk = A[q * n + k];
          A[p * n + k] = c * apk - s * aqk;
          A[q * n + k] = s * apk + c * aqk;
        }
      }
  }
  std::vector<long double> ev(n);
  for (int i = 0; i < n; i++) ev[i] = A[i * n + i];
  std::sort(ev.begin(), ev.end());
  return ev;
}
}  // namespace

void gauss_jacobi01(int n, int alpha, std::vector<double>& x, std::vector<double>& w) {
  if (n < 1) throw std::runtime_error("gauss_jacobi01: n < 1");
  const int a = alpha, b = 0;
  // Golub-Welsch start values from the Jacobi matrix of P^{(a,0)}
  std::vector<long double> J((size_t)n * n, 0.0L);
  for (int k = 0; k < n; k++) {
    long double c = 2.0L * k + a + b;
    J[k * n + k] = (k == 0) ? (long double)(b - a) / (a + b + 2) : (long double)(b * b - a * a) / (c * (c + 2));
    if (k + 1 < n) {
      int k1 = k + 1;
      long double c1 = 2.0L * k1 + a + b;
      long double num = 4.0L * k1 * (k1 + a) * (k1 + b) * (k1 + a + b);
      long double den = c1 * c1 * (c1 + 1) * (c1 - 1);
      long double off = sqrtl(num / den);
      J[k * n + k1] = J[k1 * n + k] = off;
    }
  }
  std::vector<long double> t = sym_eigenvalues(J, n);
  x.resize(n);
  w.resize(n);
  for (int i = 0; i < n; i++) {
    long double ti = t[i];
    for (int it = 0; it < 50; it++) {  // Newton polish
      long double dt = jacobi_p(n, a, b, ti) / jacobi_dp(n, a, b, ti);
      ti -= dt;
      if (fabsl(dt) < 1e-19L) break;
    }
    long double dp = jacobi_dp(n, a, b, ti);
    // weight on [-1,1] for (1-t)^a: 2^{a+1} / ((1-t^2) P'^2); mapping to [0,1] divides by 2^{a+1}
    long double wi = 1.0L / ((1 - ti * ti) * dp * dp);
    x[i] = (double)(0.5L * (1 + ti));
    w[i] = (double)wi;
  }
}

namespace {
std::mutex g_override_mutex;
std::map<std::pair<int, int>, SimplexRule> g_override;
}  // namespace

SimplexRule make_simplex_rule(int dim, int order) {
  {
    std::lock_guard<std::mutex> lock(g_override_mutex);
    auto it = g_override.find({dim, order});
    if (it != g_override.end()) return it->second;
  }
  return default_simplex_rule(dim, order);
}

void set_simplex_rule_override(int dim, int order, int nip, const double* xi, const double* w) {
  const SimplexRule def = default_simplex_rule(dim, order);  // validates dim / order
  std::lock_guard<std::mutex> lock(g_override_mutex);
  if (nip == 0) {
    g_override.erase({dim, order});
    return;
  }
  if (nip != def.nip)
    throw std::invalid_argument("rule override: the table must have " + std::to_string(def.nip) + " points (the default rule's)");
  if (!xi || !w) throw std::invalid_argument("rule override: null table");
  SimplexRule r;
  r.dim = dim;
  r.nip = nip;
  r.xi.assign(xi, xi + (size_t)nip * dim);
  r.w.assign(w, w + nip);
  for (int q = 0; q < nip; q++) {
    if (!(r.w[q] > 0)) throw std::invalid_argument("rule override: weights must be positive");
    double s = 0;
    for (int d = 0; d < dim; d++) {
      const double x = r.xi[(size_t)q * dim + d];
      if (!(x >= 0 && x <= 1)) throw std::invalid_argument("rule override: point outside the reference simplex");
      s += x;
    }
    if (s > 1 + 1e-14) throw std::invalid_argument("rule override: point outside the reference simplex");
  }
  // exactness: integral of x^a over the reference simplex = prod a_i! / (dim + |a|)!
  auto fact = [](int n) { double f = 1; for (int i = 2; i <= n; i++) f *= i; return f; };
  int a[3] = {0, 0, 0};
  for (a[0] = 0; a[0] <= order; a[0]++)
    for (a[1] = 0; a[1] <= (dim > 1 ? order - a[0] : 0); a[1]++)
      for (a[2] = 0; a[2] <= (dim > 2 ? order - a[0] - a[1] : 0); a[2]++) {
        const int deg = a[0] + a[1] + a[2];
        double exact = 1.0 / fact(dim + deg);
        for (int d = 0; d < dim; d++) exact *= fact(a[d]);
        double sum = 0;
        for (int q = 0; q < nip; q++) {
          double v = r.w[q];
          for (int d = 0; d < dim; d++) v *= std::pow(r.xi[(size_t)q * dim + d], a[d]);
          sum += v;
        }
        if (std::fabs(sum - exact) > 1e-11 * exact)
          throw std::invalid_argument("rule override: the table does not integrate degree " + std::to_string(deg) + " exactly");
      }
  g_override[{dim, order}] = std::move(r);
}

SimplexRule default_simplex_rule(int dim, int order) {
  if (dim < 1 || dim > 3) throw std::runtime_error("make_simplex_rule: dim must be 1..3");
  if (order < 0) throw std::runtime_error("make_simplex_rule: negative order");
  const int n = order / 2 + 1;
  SimplexRule r;
  r.dim = dim;
  std::vector<double> x0, w0, x1, w1, x2, w2;
  if (dim == 1) {
    gauss_jacobi01(n, 0, x0, w0);
    r.nip = n;
    r.xi = x0;
    r.w = w0;
  } else if (dim == 2) {
    gauss_jacobi01(n, 1, x0, w0);
    gauss_jacobi01(n, 0, x1, w1);
    r.nip = n * n;
    for (int i = 0; i < n; i++)
      for (int j = 0; j < n; j++) {
        r.xi.push_back(x0[i]);
        r.xi.push_back(x1[j] * (1 - x0[i]));
        r.w.push_back(w0[i] * w1[j]);
      }
  } else {
    gauss_jacobi01(n, 2, x0, w0);
    gauss_jacobi01(n, 1, x1, w1);
    gauss_jacobi01(n, 0, x2, w2);
    r.nip = n * n * n;
    for (int i = 0; i < n; i++)
      for (int j = 0; j < n; j++)
        for (int k = 0; k < n; k++) {
          r.xi.push_back(x0[i]);
          r.xi.push_back(x1[j] * (1 - x0[i]));
          r.xi.push_back(x2[k] * (1 - x0[i]) * (1 - x1[j]));
          r.w.push_back(w0[i] * w1[j] * w2[k]);
        }
  }
  return r;
}

namespace {
struct CompactRuleTable {
  int dim, degree, n;
  std::vector<double> bary;  // n * (dim + 1)
  std::vector<double> w;     // n
};
const std::vector<CompactRuleTable>& compact_tables() {
  static const std::vector<CompactRuleTable> t = {
#include "compact_rules.inc"
  };
  return t;
}
}  // namespace

SimplexRule make_compact_rule(int dim, int degree) {
  for (const CompactRuleTable& t : compact_tables())
    if (t.dim == dim && t.degree == degree) {
      if (t.n != compact_rule_points(dim, degree) || (int)t.w.size() != t.n || (int)t.bary.size() != t.n * (dim + 1))
        throw std::runtime_error("make_compact_rule: table does not match compact_rule_points");
      SimplexRule r;
      r.dim = dim;
      r.nip = t.n;
      r.w = t.w;
      for (int q = 0; q < t.n; q++)
        for (int d = 0; d < dim; d++) r.xi.push_back(t.bary[(size_t)q * (dim + 1) + d]);
      return r;
    }
  return default_simplex_rule(dim, degree);
}

std::vector<double> rule_transfer_matrix(const SimplexRule& A, const SimplexRule& B, int deg) {
  if (A.dim != B.dim || A.dim < 1 || A.dim > 3 || deg < 0) throw std::invalid_argument("rule_transfer_matrix: bad arguments");
  const int D = A.dim, nA = A.nip, nB = B.nip;
  std::vector<int> ex;  // exponents of the monomials of total degree <= deg, first coordinate outermost
  int e[3] = {0, 0, 0};
  for (e[0] = 0; e[0] <= deg; e[0]++)
    for (e[1] = 0; e[1] <= (D > 1 ? deg - e[0] : 0); e[1]++)
      for (e[2] = 0; e[2] <= (D > 2 ? deg - e[0] - e[1] : 0); e[2]++)
        for (int d = 0; d < D; d++) ex.push_back(e[d]);
  const int nd = (int)ex.size() / D;
  auto mono = [&](const SimplexRule& r, std::vector<double>& V) {
    V.assign((size_t)r.nip * nd, 0.0);
    for (int q = 0; q < r.nip; q++)
      for (int k = 0; k < nd; k++) {
        double m = 1.0;
        for (int d = 0; d < D; d++) m *= std::pow(r.xi[(size_t)q * D + d], ex[(size_t)k * D + d]);
        V[(size_t)q * nd + k] = m;
      }
  };
  std::vector<double> VA, VB;
  mono(A, VA);
  mono(B, VB);
  // modified Gram-Schmidt (twice) on the columns of VA in A's inner product; the same column operations on VB evaluate
  // the orthonormal functions at B's points
  auto dot = [&](int a, int b) {
    double s = 0;
    for (int q = 0; q < nA; q++) s += A.w[q] * VA[(size_t)q * nd + a] * VA[(size_t)q * nd + b];
    return s;
  };
  for (int k = 0; k < nd; k++) {
    for (int pass = 0; pass < 2; pass++)
      for (int j = 0; j < k; j++) {
        const double r = dot(j, k);
        for (int q = 0; q < nA; q++) VA[(size_t)q * nd + k] -= r * VA[(size_t)q * nd + j];
        for (int q = 0; q < nB; q++) VB[(size_t)q * nd + k] -= r * VB[(size_t)q * nd + j];
      }
    const double nrm = std::sqrt(dot(k, k));
    if (!(nrm > 0)) throw std::domain_error("rule_transfer_matrix: rule A does not resolve the polynomial space");
    for (int q = 0; q < nA; q++) VA[(size_t)q * nd + k] /= nrm;
    for (int q = 0; q < nB; q++) VB[(size_t)q * nd + k] /= nrm;
  }
  std::vector<double> M((size_t)nB * nA, 0.0);
  for (int r = 0; r < nB; r++)
    for (int q = 0; q < nA; q++) {
      double s = 0;
      for (int k = 0; k < nd; k++) s += VB[(size_t)r * nd + k] * VA[(size_t)q * nd + k];
      M[(size_t)r * nA + q] = s * A.w[q];
    }
  return M;
}

}  // namespace ngst
