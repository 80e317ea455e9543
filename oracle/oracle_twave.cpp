// TEST INFRASTRUCTURE ONLY (see oracle_common.hpp). CPU restatement of the reference's Trefftz tent solver:
//   TWaveBasis<D+1>::Basis            src/trefftzfespace.cpp:611-682  (+ TraversePol 455-488, IndexMap2 hpp:128-142)
//   MatToCSR                          src/scalarmappedfe.cpp:153-175
//   ScalarMappedElement::CalcShape    src/scalarmappedfe.cpp:437-540,  ::CalcDShape 550-674
//   TWaveTents<D>::Propagate          src/twavetents.cpp:72-175
//   ::CalcTentEl 177-333, ::CalcTentBndEl 335-434, ::CalcTentMacroEl 436-539, ::CalcTentElEval 541-590
//   ::TentFaceVerts 594-631, ::TentFaceArea 633-680, ::TentFaceNormal 682-732
//   ::MakeWavefront 734-764 (sampling is done by the caller), ::Error 766-800, ::L2Error 802-822, ::Energy 824-847
//   ::TentAdiam 861-904, ::MakeMacroEl 916-933, ::GetFacetSurfaceElement 935-963, ::Solve 64-70
// NGSolve facilities the reference leans on are restated as in SURVEY.md section 8.3: the quadrature rule is an
// INPUT (set by the caller), affine maps use lambda = (xi, 1 - sum xi) in element-local vertex order, CalcInverse
// follows LAPACK's dgetrf + dgetri (unblocked), dense products are plain loops. No SIMD padding: snip == nip
// (pad points carry zero weight in the reference and contribute nothing to elmat / elvec).
// Tents are run level by level with OpenMP over the tents of a level; the reference's RunParallelDependency
// (paralleldepend.hpp:20-131) yields the same result because concurrently running tents never share elements.
#include <chrono>
#include <algorithm>
#include <cmath>
#include <cstring>
#include <map>
#include <stdexcept>
#include <unordered_map>
#include <vector>

#include "oracle_common.hpp"

namespace oracle {

static int BinCoeff(int n, int k) {
  if (k < 0 || k > n) return 0;
  long long r = 1;
  for (int i = 1; i <= k; i++) r = r * (n - k + i) / i;
  return (int)r;
}

struct Matrix {
  int h = 0, w = 0;
  std::vector<double> d;
  Matrix() {}
  Matrix(int ah, int aw) : h(ah), w(aw), d((size_t)ah * aw, 0.0) {}
  double& operator()(int i, int j) { return d[(size_t)i * w + j]; }
  double operator()(int i, int j) const { return d[(size_t)i * w + j]; }
  double* Row(int i) { return &d[(size_t)i * w]; }
  const double* Row(int i) const { return &d[(size_t)i * w]; }
};

// C(ri.., ci..) += sign * A * B, with A (n x k) and B (k x m), C addressed with leading dimension ldc
static void AddAB(const Matrix& A, const Matrix& B, double* C, int ldc, double sign = 1.0) {
  const int n = A.h, k = A.w, m = B.w;
  for (int i = 0; i < n; i++) {
    double* c = C + (size_t)i * ldc;
    const double* a = A.Row(i);
    for (int kk = 0; kk < k; kk++) {
      const double aik = sign * a[kk];
      const double* b = B.Row(kk);
      for (int j = 0; j < m; j++) c[j] += aik * b[j];
    }
  }
}
static void AddAv(const Matrix& A, const std::vector<double>& v, double* out, double sign) {
  for (int i = 0; i < A.h; i++) {
    const double* a = A.Row(i);
    double s = 0;
    for (int k = 0; k < A.w; k++) s += a[k] * v[k];
    out[i] += sign * s;
  }
}

struct CSR {
  std::vector<int> rowptr, col;
  std::vector<double> val;
};

// exponent tuples in evaluation order (first coordinate outermost, scalarmappedfe.cpp:526-531) -> index
struct PolIndex {
  int d1, ord;
  std::map<std::vector<int>, int> idx;
  std::vector<std::vector<int>> expo;
  PolIndex(int ad1, int aord) : d1(ad1), ord(aord) {
    std::vector<int> e(d1, 0);
    rec(0, ord, e);
  }
  void rec(int d, int left, std::vector<int>& e) {
    if (d == d1) { idx[e] = (int)expo.size(); expo.push_back(e); return; }
    for (int i = 0; i <= left; i++) { e[d] = i; rec(d + 1, left - i, e); }
    e[d] = 0;
  }
  int operator()(const std::vector<int>& e) const { return idx.at(e); }
};

// TraversePol: last coordinate (time) outermost, first coordinate innermost
template <typename F>
static void TraversePol(int d1, int order, F func) {
  std::vector<int> e(d1, 0);
  std::vector<std::vector<int>> all;
  struct R {  // loop level l handles coordinate d1-1-l
    static void go(int level, int d1, int left, std::vector<int>& e, std::vector<std::vector<int>>& out) {
      if (level == d1) { out.push_back(e); return; }
      const int coord = d1 - 1 - level;
      for (int i = 0; i <= left; i++) { e[coord] = i; go(level + 1, d1, left - i, e, out); }
      e[coord] = 0;
    }
  };
  R::go(0, d1, order, e, all);
  for (size_t ii = 0; ii < all.size(); ii++) func((int)ii, all[ii]);
}

static CSR TWaveBasis(int d1, int ord, int fowave, const PolIndex& pix) {
  const int ndof = BinCoeff(d1 - 1 + ord, ord) + BinCoeff(d1 + ord - 2, ord - 1);
  const int npoly = BinCoeff(d1 + ord, ord);
  Matrix trefftzbasis(ndof, npoly);
  for (int basis = 0; basis < ndof; basis++) {
    int tracker = 0;
    TraversePol(d1, ord, [&](int, const std::vector<int>& coeff) {
      if (tracker >= 0) tracker++;
      const int indexmap = pix(coeff);
      const int k = coeff[d1 - 1];
      if (k == 0 || k == 1) {
        if (tracker > basis) { trefftzbasis(basis, indexmap) = 1; tracker = -1; }
      } else if (k > 1) {
        for (int m = 0; m < d1 - 1; m++) {
          std::vector<int> get_coeff = coeff;
          get_coeff[d1 - 1] -= 2;
          get_coeff[m] += 2;
          trefftzbasis(basis, indexmap) += (coeff[m] + 1) * (coeff[m] + 2) * trefftzbasis(basis, pix(get_coeff));
        }
        const double wavespeed = 1.0;
        trefftzbasis(basis, indexmap) *= wavespeed * wavespeed / (k * (k - 1));
      }
    });
  }
  CSR sp;  // MatToCSR on Rows(fowave, ndof)
  int spsize = 0;
  for (int i = fowave; i < ndof; i++) {
    sp.rowptr.push_back(spsize);
    for (int j = 0; j < npoly; j++)
      if (trefftzbasis(i, j)) { spsize++; sp.col.push_back(j); sp.val.push_back(trefftzbasis(i, j)); }
  }
  sp.rowptr.push_back(spsize);
  return sp;
}

// ScalarMappedElement<D+1>
struct MappedElement {
  int d1, ndof, order, npoly;
  const CSR* localmat;
  const PolIndex* pix;
  double shift[4], scale[4];
  void Monomials(const double* point, std::vector<double>& polxt) const {  // polxt[d*(order+1)+k]
    for (int d = 0; d < d1; d++) {
      const double x = (point[d] - shift[d]) * scale[d];
      double* p = &polxt[(size_t)d * (order + 1)];
      p[0] = 1.0;
      for (int k = 1; k <= order; k++) p[k] = p[k - 1] * x;
    }
  }
  // shape(i, imip)
  void CalcShape(const Matrix& points, Matrix& shape) const {
    std::vector<double> polxt((size_t)d1 * (order + 1)), pol(npoly);
    for (int imip = 0; imip < points.h; imip++) {
      Monomials(points.Row(imip), polxt);
      for (int ii = 0; ii < npoly; ii++) {
        double v = 1.0;
        for (int d = 0; d < d1; d++) v *= polxt[(size_t)d * (order + 1) + pix->expo[ii][d]];
        pol[ii] = v;
      }
      for (int i = 0; i < ndof; i++) {
        double s = 0.0;
        for (int j = localmat->rowptr[i]; j < localmat->rowptr[i + 1]; j++) s += localmat->val[j] * pol[localmat->col[j]];
        shape(i, imip) = s;
      }
    }
  }
  // bbmat(i, d*snip + imip) = d-th derivative of shape i  (the reinterpretation at twavetents.cpp:250-251)
  void CalcDShape(const Matrix& points, Matrix& bbmat) const {
    const int snip = points.h;
    std::vector<double> polxt((size_t)d1 * (order + 1)), pol(npoly);
    for (int imip = 0; imip < snip; imip++) {
      Monomials(points.Row(imip), polxt);
      for (int d = 0; d < d1; d++) {
        for (int ii = 0; ii < npoly; ii++) {
          const std::vector<int>& e = pix->expo[ii];
          double v = e[d];
          for (int dd = 0; dd < d1; dd++) {
            const int ex = e[dd] - (dd == d);
            v *= ex >= 0 ? polxt[(size_t)dd * (order + 1) + ex] : 0.0;
          }
          pol[ii] = v;
        }
        for (int i = 0; i < ndof; i++) {
          double s = 0.0;
          for (int j = localmat->rowptr[i]; j < localmat->rowptr[i + 1]; j++)
            s += localmat->val[j] * pol[localmat->col[j]] * scale[d];
          bbmat(i, d * snip + imip) = s;
        }
      }
    }
  }
};

// CalcInverse(a, INV_LAPACK): dgetrf (partial pivoting) followed by dgetri, unblocked.
static bool CalcInverseLapack(double* a, int n) {
  std::vector<int> ipiv(n);
  bool ok = true;
  for (int j = 0; j < n; j++) {  // dgetf2 (row-major storage, row interchanges)
    int p = j;
    double best = std::fabs(a[(size_t)j * n + j]);
    for (int i = j + 1; i < n; i++)
      if (std::fabs(a[(size_t)i * n + j]) > best) { best = std::fabs(a[(size_t)i * n + j]); p = i; }
    ipiv[j] = p;
    if (best == 0.0) { ok = false; continue; }
    if (p != j)
      for (int c = 0; c < n; c++) std::swap(a[(size_t)j * n + c], a[(size_t)p * n + c]);
    const double r = 1.0 / a[(size_t)j * n + j];
    for (int i = j + 1; i < n; i++) a[(size_t)i * n + j] *= r;
    for (int i = j + 1; i < n; i++) {
      const double l = a[(size_t)i * n + j];
      for (int c = j + 1; c < n; c++) a[(size_t)i * n + c] -= l * a[(size_t)j * n + c];
    }
  }
  if (!ok) return false;
  // dtrti2: inverse of the upper triangle in place
  for (int j = 0; j < n; j++) {
    a[(size_t)j * n + j] = 1.0 / a[(size_t)j * n + j];
    const double ajj = -a[(size_t)j * n + j];
    // column j above the diagonal: x = U(0:j,0:j)^-1-so-far * u(0:j, j)
    for (int i = 0; i < j; i++) {
      double s = 0;
      for (int k = i; k < j; k++) s += a[(size_t)i * n + k] * a[(size_t)k * n + j];
      a[(size_t)i * n + j] = s;  // uses already-inverted leading block (upper triangular, row i from k = i)
    }
    for (int i = 0; i < j; i++) a[(size_t)i * n + j] *= ajj;
  }
  // dgetri: solve inv(A) * L = inv(U), columns from the last to the first
  std::vector<double> work(n);
  for (int j = n - 1; j >= 0; j--) {
    for (int i = j + 1; i < n; i++) { work[i] = a[(size_t)i * n + j]; a[(size_t)i * n + j] = 0.0; }
    for (int k = j + 1; k < n; k++) {
      const double wk = work[k];
      if (wk != 0.0)
        for (int i = 0; i < n; i++) a[(size_t)i * n + j] -= a[(size_t)i * n + k] * wk;
    }
  }
  for (int j = n - 2; j >= 0; j--) {  // undo the row interchanges as column interchanges
    const int p = ipiv[j];
    if (p != j)
      for (int i = 0; i < n; i++) std::swap(a[(size_t)i * n + j], a[(size_t)i * n + p]);
  }
  return true;
}

template <typename T>
static int sgn_nozero(T val) { return (T(0) <= val) - (val < T(0)); }

struct TWaveTents {
  int D, order, nbasis, fosystem = 0;
  const OMesh* ma;
  const OSlab* tps;
  std::vector<double> wavespeed, c_vertex;
  int snip = 0;
  Matrix ir_points;            // snip x D reference points
  std::vector<double> ir_w;    // snip
  Matrix wavefront;            // nel x ncomp*snip
  std::vector<double> bddata;  // boundary datum samples in traversal order
  std::vector<long long> bd_offset;  // per tent: offset into bddata of its first boundary facet
  bool have_bd = false;
  double timeshift = 0;
  // EXPERIMENT (off by default; not part of the restatement): bottom-face terms of CalcTentEl integrated on a compact rule
  // B after transferring the state row with M (cnip x snip, ngst_rule_transfer). Proves on the CPU that the fast kernel's
  // planned compact state reproduces the reference's Propagate up to round-off.
  int cnip = 0;
  Matrix c_points;             // cnip x D
  std::vector<double> c_w;     // cnip
  Matrix c_M;                  // cnip x snip
  // Per-phase thread-seconds with the reference's timer names (twavetents.cpp:84-88, 186-188): [0] "tent top calcshape",
  // [1] "tent top AAt", [2] "tent top bilinearform", [3] "tentel" (all of CalcTentEl), [4] solve, [5] "tenteval",
  // [6] "tent" (whole tent). Filled only when phase_timing is set; bench.py reports them beside the CPU baseline.
  bool phase_timing = false;
  mutable double phase_sec[7] = {0, 0, 0, 0, 0, 0, 0};
  static double Now() {
    return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
  }
  void AddPhase(int k, double sec) const {
#pragma omp atomic
    phase_sec[k] += sec;
  }
  PolIndex pix;
  int singular = 0;
  double max_cond = 0;  // largest 1-norm condition number of a tent system since the last reset
  // wavespeed per (element, quadrature point) for Energy / Error (the reference evaluates the CF at every point,
  // twavetents.cpp:785-793, 835-843); empty = the element value
  std::vector<double> c_point;
  double CAt(int elnr, int imip) const { return c_point.empty() ? wavespeed[elnr] : c_point[(size_t)elnr * snip + imip]; }
  virtual ~TWaveTents() {}

  TWaveTents(int aD, int aorder, const OMesh* ama, const OSlab* atps, const double* c_el, const double* c_vert)
      : D(aD), order(aorder), ma(ama), tps(atps), pix(aD + 1, aorder) {
    nbasis = BinCoeff(D + order, order) + BinCoeff(D + order - 1, order - 1);
    wavespeed.assign(c_el, c_el + ma->nel);
    if (c_vert) c_vertex.assign(c_vert, c_vert + ma->nv);
    else {
      c_vertex.resize(ma->nv);
      for (int v = 0; v < ma->nv; v++) c_vertex[v] = wavespeed[ma->v2el[v][0]];
    }
  }

  // ---- geometry -------------------------------------------------------------------------------------
  typedef std::vector<double> Vert;  // (D+1) x (D+1), v(row, col) = vert[row*(D+1)+col], columns = vertices
  double& V(Vert& v, int r, int c) const { return v[(size_t)r * (D + 1) + c]; }
  double V(const Vert& v, int r, int c) const { return v[(size_t)r * (D + 1) + c]; }
  int NbPos(const OTent& tent, int vnr) const {
    for (size_t k = 0; k < tent.nbv.size(); k++)
      if (tent.nbv[k] == vnr) return (int)k;
    throw std::runtime_error("oracle: vertex is not a neighbour of the tent vertex");
  }
  Vert TentFaceVertsBnd(const OTent& tent, const int* sel_verts) const {
    Vert v((size_t)(D + 1) * (D + 1));
    for (int d = 0; d < D; d++) V(v, d, 0) = ma->point(tent.vertex)[d];
    V(v, D, 0) = tent.tbot;
    for (int n = 0; n < D; n++) {
      for (int d = 0; d < D; d++) V(v, d, n + 1) = ma->point(sel_verts[n])[d];
      V(v, D, n + 1) = tent.vertex == sel_verts[n] ? tent.ttop : tent.nbtime[NbPos(tent, sel_verts[n])];
    }
    return v;
  }
  Vert TentFaceVerts(const OTent& tent, int elnr, int top) const {
    Vert v((size_t)(D + 1) * (D + 1));
    const int* vnr = ma->elverts(elnr);
    for (int ivert = 0; ivert <= D; ivert++) {
      for (int d = 0; d < D; d++) V(v, d, ivert) = ma->point(vnr[ivert])[d];
      if (vnr[ivert] == tent.vertex) V(v, D, ivert) = top == 1 ? tent.ttop : tent.tbot;
      else
        for (size_t k = 0; k < tent.nbv.size(); k++)
          if (vnr[ivert] == tent.nbv[k]) V(v, D, ivert) = tent.nbtime[k];
    }
    return v;
  }
  double ColDist(const Vert& ve, int a, int b) const {
    double s = 0;
    for (int r = 0; r <= D; r++) s += (V(ve, r, a) - V(ve, r, b)) * (V(ve, r, a) - V(ve, r, b));
    return std::sqrt(s);
  }
  double TentFaceArea(const Vert& ve) const {
    switch (D) {
      case 1: return ColDist(ve, 0, 1);
      case 2: {
        double ba[3], baa[3];
        for (int r = 0; r < 3; r++) { ba[r] = V(ve, r, 0) - V(ve, r, 1); baa[r] = V(ve, r, 0) - V(ve, r, 2); }
        const double c[3] = {ba[1] * baa[2] - ba[2] * baa[1], ba[2] * baa[0] - ba[0] * baa[2], ba[0] * baa[1] - ba[1] * baa[0]};
        return std::sqrt(c[0] * c[0] + c[1] * c[1] + c[2] * c[2]) / 2.0 * 2.0;
      }
      default: {
        const double U = ColDist(ve, 0, 1), Vv = ColDist(ve, 1, 2), W = ColDist(ve, 2, 0);
        const double u = ColDist(ve, 3, 2), v = ColDist(ve, 3, 0), w = ColDist(ve, 3, 1);
        const double X = (w - U + v) * (U + v + w), x = (U - v + w) * (v - w + U);
        const double Y = (u - Vv + w) * (Vv + w + u), y = (Vv - w + u) * (w - u + Vv);
        const double Z = (v - W + u) * (W + u + v), z = (W - u + v) * (u - v + W);
        const double a = std::sqrt(x * Y * Z), b = std::sqrt(y * Z * X), c = std::sqrt(z * X * Y), d = std::sqrt(x * y * z);
        return std::sqrt((-a + b + c + d) * (a - b + c + d) * (a + b - c + d) * (a + b + c - d)) / (192.0 * u * v * w) * 6.0;
      }
    }
  }
  std::vector<double> TentFaceNormal(Vert v, int top) const {
    std::vector<double> normv(D + 1);
    switch (D) {
      case 1:
        normv[0] = V(v, 1, 1) - V(v, 1, 0);
        normv[1] = V(v, 0, 0) - V(v, 0, 1);
        break;
      case 2: {
        double ba[3], baa[3];
        for (int r = 0; r < 3; r++) { ba[r] = V(v, r, 0) - V(v, r, 1); baa[r] = V(v, r, 0) - V(v, r, 2); }
        normv[0] = ba[1] * baa[2] - ba[2] * baa[1];
        normv[1] = ba[2] * baa[0] - ba[0] * baa[2];
        normv[2] = ba[0] * baa[1] - ba[1] * baa[0];
        break;
      }
      default: {
        for (int d = 1; d < D + 1; d++)
          for (int r = 0; r <= D; r++) V(v, r, d) = V(v, r, 0) - V(v, r, d);
        for (int i = 0; i < D + 1; i++) {
          double pS[3][3];
          for (int k = 0, c = 0; k < D + 1; k++) {
            if (k == i) continue;
            for (int j = 0; j < 3; j++) pS[c][j] = V(v, k, j + 1);
            c++;
          }
          const double det = pS[0][0] * (pS[1][1] * pS[2][2] - pS[1][2] * pS[2][1]) -
                             pS[0][1] * (pS[1][0] * pS[2][2] - pS[1][2] * pS[2][0]) +
                             pS[0][2] * (pS[1][0] * pS[2][1] - pS[1][1] * pS[2][0]);
          normv[i] = (i % 2) == 0 ? det : -det;
        }
      }
    }
    if (top == 1) { const double s = sgn_nozero<double>(normv[D]); for (auto& x : normv) x *= s; }
    else if (top == -1) { const double s = -sgn_nozero<double>(normv[D]); for (auto& x : normv) x *= s; }
    else normv[D] = 0;
    double l = 0;
    for (double x : normv) l += x * x;
    l = std::sqrt(l);
    for (auto& x : normv) x /= l;
    return normv;
  }
  double TentAdiam(const OTent& tent) const {
    const int vnumber = (int)tent.nbv.size();
    double c1 = wavespeed[0];
    double anisotropicdiam = c1 * tent.ttop - c1 * tent.tbot;
    auto L2Norm2 = [&](const double* a, const double* b) {
      double s = 0;
      for (int d = 0; d < D; d++) s += (a[d] - b[d]) * (a[d] - b[d]);
      return s;
    };
    for (int k = 0; k < vnumber; k++) {
      const double* v1 = ma->point(tent.vertex);
      const double* v2 = ma->point(tent.nbv[k]);
      c1 = c_vertex[tent.vertex];
      const double c2 = c_vertex[tent.nbv[k]];
      anisotropicdiam = std::max(anisotropicdiam, std::sqrt(L2Norm2(v1, v2) + std::pow(c1 * tent.ttop - c2 * tent.nbtime[k], 2)));
      anisotropicdiam = std::max(anisotropicdiam, std::sqrt(L2Norm2(v1, v2) + std::pow(c1 * tent.tbot - c2 * tent.nbtime[k], 2)));
      for (int j = 0; j < vnumber; j++) {
        v1 = ma->point(tent.nbv[j]);
        c1 = c_vertex[tent.nbv[j]];
        anisotropicdiam = std::max(anisotropicdiam, std::sqrt(L2Norm2(v1, v2) + std::pow(c1 * tent.nbtime[j] - c2 * tent.nbtime[k], 2)));
      }
    }
    return anisotropicdiam;
  }
  int MakeMacroEl(const std::vector<int>& tentel, std::unordered_map<int, int>& macroel) const {
    int nrmacroel = 0;
    for (size_t i = 0; i < tentel.size(); i++) {
      size_t j = 0;
      while (wavespeed[tentel[i]] != wavespeed[tentel[j]]) j++;
      if (j == i) macroel[tentel[i]] = nrmacroel++;
      else macroel[tentel[i]] = macroel[tentel[j]];
    }
    return nrmacroel;
  }

  // mapped points of the spatial rule on element elnr, time from the linear face function
  void ElementPoints(int elnr, const Vert& vert, Matrix& pts) const { ElementPointsAt(elnr, vert, ir_points, pts); }
  void ElementPointsAt(int elnr, const Vert& vert, const Matrix& ref, Matrix& pts) const {
    const int* ev = ma->elverts(elnr);
    for (int q = 0; q < ref.h; q++) {
      double lam[4], s = 0;
      for (int d = 0; d < D; d++) { lam[d] = ref(q, d); s += lam[d]; }
      lam[D] = 1.0 - s;
      for (int d = 0; d < D; d++) {
        double x = 0;
        for (int k = 0; k <= D; k++) x += lam[k] * ma->point(ev[k])[d];
        pts(q, d) = x;
      }
      double t = 0;
      for (int k = 0; k <= D; k++) t += lam[k] * V(vert, D, k);  // faceint.Evaluate(sir, linbasis, mirtimes)
      pts(q, D) = t;
    }
  }
  void FacePoints(const Vert& vert, Matrix& pts) const {  // twavetents.cpp:355-365
    for (int q = 0; q < snip; q++)
      for (int r = 0; r <= D; r++) {
        double x = 0;
        for (int i = 0; i < D; i++) x += (V(vert, r, i + 1) - V(vert, r, 0)) * ir_points(q, i);
        pts(q, r) = x + V(vert, r, 0);
      }
  }
  void SetWavespeed(MappedElement& tel, double c) const { tel.scale[D] = tel.scale[0] * c; }

  // ---- per-tent pieces ------------------------------------------------------------------------------
  // cpt: wavespeed per quadrature point (the reference's LocalWavespeed(imip) functor, twavetents.cpp:181, 1005-1015);
  // nullptr = the element's constant c
  void CalcTentEl(int elnr, const OTent& tent, MappedElement& tel, double c, double* elmat, int ld, double* elvec,
                  Matrix& topdshapes, const double* cpt = nullptr) const {
    const int d1 = D + 1;
    const double t_el = phase_timing ? Now() : 0.0;
    Matrix pts(snip, d1);
    Vert vert = TentFaceVerts(tent, elnr, -1);
    double area = TentFaceArea(vert);
    std::vector<double> n = TentFaceNormal(vert, -1);
    const int off = !fosystem;
    if (cnip > 0) {
      // experiment: the same bottom-face terms on the compact rule, state row transferred with M
      if (cpt) throw std::runtime_error("oracle: compact right-hand side needs a constant wavespeed per element");
      const int nc = off + d1;
      Matrix g(nc, cnip);
      for (int cc = 0; cc < nc; cc++)
        for (int r = 0; r < cnip; r++) {
          double sacc = 0;
          for (int q = 0; q < snip; q++) sacc += c_M(r, q) * wavefront(elnr, cc * snip + q);
          g(cc, r) = sacc;
        }
      Matrix cpts(cnip, d1);
      ElementPointsAt(elnr, vert, c_points, cpts);
      std::vector<double> cvec((size_t)d1 * cnip, 0.0);
      for (int r = 0; r < cnip; r++) {
        const double weight = c_w[r] * area;
        cvec[D * cnip + r] += n[D] * std::pow(c, -2) * weight * g(off + D, r);
        for (int d = 0; d < D; d++) {
          cvec[d * cnip + r] += n[D] * weight * g(off + d, r);
          cvec[d * cnip + r] -= n[d] * weight * g(off + D, r);
          cvec[D * cnip + r] -= n[d] * weight * g(off + d, r);
        }
      }
      Matrix cbb(nbasis, d1 * cnip);
      tel.CalcDShape(cpts, cbb);
      AddAv(cbb, cvec, elvec, -1.0);
      if (!fosystem) {
        Matrix cs(nbasis, cnip);
        tel.CalcShape(cpts, cs);
        Matrix ws(nbasis, cnip), wst(cnip, nbasis), ws2(nbasis, cnip);
        std::vector<double> wu(cnip);
        for (int r = 0; r < cnip; r++) wu[r] = g(0, r);
        for (int i = 0; i < nbasis; i++)
          for (int r = 0; r < cnip; r++) {
            const double sq = std::sqrt(area * c_w[r]);
            ws(i, r) = cs(i, r) * sq;
            wst(r, i) = ws(i, r);
            ws2(i, r) = ws(i, r) * sq;
          }
        AddAB(ws, wst, elmat, ld);
        AddAv(ws2, wu, elvec, 1.0);
      }
    }
    ElementPoints(elnr, vert, pts);
    std::vector<double> bdbvec((size_t)d1 * snip, 0.0);
    for (int imip = 0; imip < snip && cnip == 0; imip++) {
      const double weight = ir_w[imip] * area;
      bdbvec[D * snip + imip] += n[D] * std::pow(cpt ? cpt[imip] : c, -2) * weight * wavefront(elnr, (off + D) * snip + imip);
      for (int d = 0; d < D; d++) {
        bdbvec[d * snip + imip] += n[D] * weight * wavefront(elnr, (off + d) * snip + imip);
        bdbvec[d * snip + imip] -= n[d] * weight * wavefront(elnr, (off + D) * snip + imip);
        bdbvec[D * snip + imip] -= n[d] * weight * wavefront(elnr, (off + d) * snip + imip);
      }
    }
    Matrix bbmat(nbasis, d1 * snip);
    tel.CalcDShape(pts, bbmat);
    AddAv(bbmat, bdbvec, elvec, -1.0);

    if (!fosystem && cnip == 0) {  // stabilization to recover the second-order solution
      Matrix shapes(nbasis, snip);
      tel.CalcShape(pts, shapes);
      Matrix ws(nbasis, snip), wst(snip, nbasis);
      for (int i = 0; i < nbasis; i++)
        for (int q = 0; q < snip; q++) {
          const double s = std::sqrt(area * ir_w[q]);
          ws(i, q) = shapes(i, q) * s;
          wst(q, i) = ws(i, q);
        }
      AddAB(ws, wst, elmat, ld);  // AddABt(simdshapes, simdshapes, elmat)
      std::vector<double> wu(snip);
      for (int q = 0; q < snip; q++) wu[q] = wavefront(elnr, q);
      Matrix ws2(nbasis, snip);
      for (int i = 0; i < nbasis; i++)
        for (int q = 0; q < snip; q++) ws2(i, q) = ws(i, q) * std::sqrt(area * ir_w[q]);
      AddAv(ws2, wu, elvec, 1.0);
    }

    vert = TentFaceVerts(tent, elnr, 1);
    ElementPoints(elnr, vert, pts);
    const double t1 = phase_timing ? Now() : 0.0;
    tel.CalcDShape(pts, bbmat);
    const double t2 = phase_timing ? Now() : 0.0;
    area = TentFaceArea(vert);
    n = TentFaceNormal(vert, 1);
    Matrix bdbmat(d1 * snip, nbasis);
    for (int imip = 0; imip < snip; imip++) {
      const double weight = ir_w[imip] * area;
      for (int j = 0; j < nbasis; j++) {
        bdbmat(D * snip + imip, j) += n[D] * std::pow(cpt ? cpt[imip] : c, -2) * weight * bbmat(j, D * snip + imip);
        for (int d = 0; d < D; d++) {
          bdbmat(d * snip + imip, j) += n[D] * weight * bbmat(j, d * snip + imip);
          bdbmat(d * snip + imip, j) -= n[d] * weight * bbmat(j, D * snip + imip);
          bdbmat(D * snip + imip, j) -= n[d] * weight * bbmat(j, d * snip + imip);
        }
      }
    }
    const double t3 = phase_timing ? Now() : 0.0;
    AddAB(bbmat, bdbmat, elmat, ld);
    const double t4 = phase_timing ? Now() : 0.0;
    topdshapes = bbmat;
    if (phase_timing) {
      AddPhase(0, t2 - t1);
      AddPhase(1, t3 - t2);
      AddPhase(2, t4 - t3);
      AddPhase(3, Now() - t_el);
    }
  }

  // invc2: 1/c^2 per boundary point (the reference evaluates the wavespeed CF there, twavetents.cpp:409-412); nullptr = 1/cadj^2
  void CalcTentBndEl(int surfel, const OTent& tent, MappedElement& tel, const double* bd, double cadj, double* elmat,
                     int ld, double* elvec, const double* invc2 = nullptr) const {
    const int d1 = D + 1;
    const int* sel_verts = &ma->se2v[(size_t)surfel * D];
    Vert vert = TentFaceVertsBnd(tent, sel_verts);
    std::vector<double> n = TentFaceNormal(vert, 0);
    for (auto& x : n) x = -x;
    if (D == 1) { n[0] = sgn_nozero<int>(tent.vertex - tent.nbv[0]); n[D] = 0; }
    Matrix pts(snip, d1);
    FacePoints(vert, pts);
    Matrix bbmat(nbasis, d1 * snip);
    tel.CalcDShape(pts, bbmat);
    const double area = TentFaceArea(vert);
    Matrix bdbmat(d1 * snip, nbasis);
    std::vector<double> bdbvec((size_t)d1 * snip, 0.0);
    if (ma->se_kind[surfel] == 1) {  // boundary named "neumann"
      const double beta = 0.5;
      for (int imip = 0; imip < snip; imip++)
        for (int r = 0; r < D; r++)  // r == D multiplies by n(D) = 0
          for (int d = 0; d < d1; d++) {
            const double f = (d < D ? -n[d] * beta : 1.0) * (-n[r]) * ir_w[imip] * area;
            for (int j = 0; j < nbasis; j++) bdbmat(r * snip + imip, j) += f * bbmat(j, d * snip + imip);
            bdbvec[d * snip + imip] += (d < D ? -n[d] * beta : -1.0) * (-n[r]) * (bd ? bd[r * snip + imip] : 0.0) * ir_w[imip] * area;
          }
      AddAB(bbmat, bdbmat, elmat, ld);
      AddAv(bbmat, bdbvec, elvec, 1.0);
    } else {  // dirichlet
      const double alpha = 0.5;
      for (int imip = 0; imip < snip; imip++) {
        const double ws = invc2 ? invc2[imip] : 1.0 / (cadj * cadj);
        const double weight = area * ir_w[imip];
        const double g = bd ? bd[imip] : 0.0;
        for (int j = 0; j < nbasis; j++) bdbmat(D * snip + imip, j) += weight * alpha * ws * bbmat(j, D * snip + imip);
        bdbvec[D * snip + imip] -= weight * alpha * ws * g;
        for (int d = 0; d < D; d++) {
          for (int j = 0; j < nbasis; j++) bdbmat(D * snip + imip, j) -= n[d] * weight * bbmat(j, d * snip + imip);
          bdbvec[d * snip + imip] -= n[d] * weight * g;
        }
      }
      AddAB(bbmat, bdbmat, elmat, ld);
      AddAv(bbmat, bdbvec, elvec, -1.0);
    }
  }

  void CalcTentMacroEl(int fnr, const int* elnums, std::unordered_map<int, int>& macroel, const OTent& tent,
                       MappedElement& tel, double* elmat, int ld) const {
    const int d1 = D + 1;
    const int* sel_verts = &ma->facet2v[(size_t)fnr * D];
    Vert vert = TentFaceVertsBnd(tent, sel_verts);
    // outward unit normal of elnums[0] on its facet fnr (ComputeNormalsAndMeasure)
    std::vector<double> n(D, 0.0);
    {
      const int* ev = ma->elverts(elnums[0]);
      int opp = -1;
      for (int k = 0; k <= D; k++) {
        bool on = false;
        for (int j = 0; j < D; j++) on = on || ev[k] == sel_verts[j];
        if (!on) opp = ev[k];
      }
      const double* a = ma->point(sel_verts[0]);
      if (D == 1) n[0] = 1.0;
      else if (D == 2) {
        const double* b = ma->point(sel_verts[1]);
        n[0] = b[1] - a[1]; n[1] = -(b[0] - a[0]);
      } else {
        const double* b = ma->point(sel_verts[1]);
        const double* c = ma->point(sel_verts[2]);
        const double u[3] = {b[0] - a[0], b[1] - a[1], b[2] - a[2]}, w[3] = {c[0] - a[0], c[1] - a[1], c[2] - a[2]};
        n[0] = u[1] * w[2] - u[2] * w[1]; n[1] = u[2] * w[0] - u[0] * w[2]; n[2] = u[0] * w[1] - u[1] * w[0];
      }
      double l = 0, s = 0;
      for (int d = 0; d < D; d++) { l += n[d] * n[d]; s += n[d] * (ma->point(opp)[d] - a[d]); }
      l = std::sqrt(l);
      for (int d = 0; d < D; d++) n[d] = (s > 0 ? -n[d] : n[d]) / l;
    }
    Matrix pts(snip, d1);
    FacePoints(vert, pts);
    Matrix bbmat[2] = {Matrix(nbasis, d1 * snip), Matrix(nbasis, d1 * snip)};
    SetWavespeed(tel, wavespeed[elnums[0]]);
    tel.CalcDShape(pts, bbmat[0]);
    SetWavespeed(tel, wavespeed[elnums[1]]);
    tel.CalcDShape(pts, bbmat[1]);
    const double area = TentFaceArea(vert);
    for (int el = 0; el < 4; el++) {
      Matrix bdbmat(d1 * snip, nbasis);
      for (int imip = 0; imip < snip; imip++) {
        const double weight = ir_w[imip] * area;
        for (int d = 0; d < D; d++)
          for (int j = 0; j < nbasis; j++) {
            bdbmat(d * snip + imip, j) -= std::pow(-1, el / 2) * 0.5 * n[d] * weight * bbmat[el % 2](j, D * snip + imip);
            bdbmat(D * snip + imip, j) -= std::pow(-1, el / 2) * 0.5 * n[d] * weight * bbmat[el % 2](j, d * snip + imip);
          }
      }
      const int in = macroel[elnums[el / 2]];
      const int out = macroel[elnums[el % 2]];
      AddAB(bbmat[el / 2], bdbmat, elmat + (size_t)in * nbasis * ld + out * nbasis, ld);
    }
  }

  void CalcTentElEval(int elnr, const OTent& tent, MappedElement& tel, const double* sol, const Matrix& dshapes) {
    const int d1 = D + 1;
    Matrix pts(snip, d1);
    Vert v = TentFaceVerts(tent, elnr, 1);
    ElementPoints(elnr, v, pts);
    if (!fosystem) {
      Matrix shapes(nbasis, snip);
      tel.CalcShape(pts, shapes);
      for (int q = 0; q < snip; q++) {
        double s = 0;
        for (int i = 0; i < nbasis; i++) s += shapes(i, q) * sol[i];
        wavefront(elnr, q) = s;
      }
    }
    const int base = snip * (!fosystem);
    for (int k = 0; k < d1 * snip; k++) {
      double s = 0;
      for (int i = 0; i < nbasis; i++) s += dshapes(i, k) * sol[i];
      wavefront(elnr, base + k) = s;
    }
  }

  // facets carrying boundary data, in the traversal order of Propagate (tent, internal_facets)
  template <typename F>
  void ForEachBoundaryFacet(F f) const {
    for (size_t t = 0; t < tps->tents.size(); t++) {
      const OTent& tent = tps->tents[t];
      for (int fnr : tent.internal_facets) {
        if (ma->facet2el[2 * fnr + 1] >= 0) continue;
        const int surfel = ma->facet2se[fnr];  // GetFacetSurfaceElement: first match
        if (surfel < 0) continue;
        f((int)t, fnr, surfel);
      }
    }
  }
  void IndexBoundary() {
    bd_offset.assign(tps->tents.size() + 1, 0);
    long long off = 0;
    int last = -1;
    ForEachBoundaryFacet([&](int t, int, int surfel) {
      for (int k = last + 1; k <= t; k++) bd_offset[k] = off;
      last = t;
      off += (long long)(ma->se_kind[surfel] == 1 ? D : 1) * snip;
    });
    for (size_t k = last + 1; k < bd_offset.size(); k++) bd_offset[k] = off;
  }

  void PropagateTent(int tentnr, const CSR& basismat) {
    const OTent& tent = tps->tents[tentnr];
    const int d1 = D + 1;
    const double t_tent = phase_timing ? Now() : 0.0;
    MappedElement tel;
    tel.d1 = d1; tel.ndof = nbasis; tel.order = order; tel.npoly = BinCoeff(d1 + order, order);
    tel.localmat = &basismat; tel.pix = &pix;
    for (int d = 0; d < D; d++) tel.shift[d] = ma->point(tent.vertex)[d];
    tel.shift[D] = (tent.ttop - tent.tbot) / 2 + tent.tbot;
    const double sc = 1.0 / TentAdiam(tent);
    for (int d = 0; d < d1; d++) tel.scale[d] = sc;

    std::unordered_map<int, int> macroel;
    const int ndomains = MakeMacroEl(tent.els, macroel);
    const int n = ndomains * nbasis;
    std::vector<double> elmat((size_t)n * n, 0.0), elvec(n, 0.0);

    long long bdoff = have_bd ? bd_offset[tentnr] : 0;
    for (int fnr : tent.internal_facets) {
      const int e0 = ma->facet2el[2 * fnr], e1 = ma->facet2el[2 * fnr + 1];
      const int nel = (e0 >= 0) + (e1 >= 0);
      const int surfel = nel == 1 ? ma->facet2se[fnr] : -1;
      if (nel == 1 && surfel >= 0) {
        SetWavespeed(tel, wavespeed[e0]);
        const int eli = ndomains > 1 ? macroel[e0] : 0;
        const int nvals = (ma->se_kind[surfel] == 1 ? D : 1) * snip;
        CalcTentBndEl(surfel, tent, tel, have_bd ? &bddata[bdoff] : nullptr, wavespeed[e0],
                      &elmat[(size_t)eli * nbasis * n + eli * nbasis], n, &elvec[eli * nbasis]);
        bdoff += nvals;
      } else if (nel == 2 && ndomains > 1 && macroel[e0] != macroel[e1]) {
        const int elnums[2] = {e0, e1};
        CalcTentMacroEl(fnr, elnums, macroel, tent, tel, elmat.data(), n);
      }
    }

    std::vector<Matrix> topdshapes(tent.els.size());
    for (size_t elnr = 0; elnr < tent.els.size(); elnr++) {
      SetWavespeed(tel, wavespeed[tent.els[elnr]]);
      const int eli = ndomains > 1 ? macroel[tent.els[elnr]] : 0;
      CalcTentEl(tent.els[elnr], tent, tel, wavespeed[tent.els[elnr]], &elmat[(size_t)eli * nbasis * n + eli * nbasis], n,
                 &elvec[eli * nbasis], topdshapes[elnr]);
    }

    // Solve: inverse, then multiply (twavetents.cpp:64-70)
    const double t_solve = phase_timing ? Now() : 0.0;
    if (!InvertTracked(elmat.data(), n)) {
#pragma omp atomic
      singular++;
    }
    std::vector<double> sol(n, 0.0);
    for (int i = 0; i < n; i++) {
      double s = 0;
      for (int j = 0; j < n; j++) s += elmat[(size_t)i * n + j] * elvec[j];
      sol[i] = s;
    }
    const double t_eval = phase_timing ? Now() : 0.0;
    for (size_t elnr = 0; elnr < tent.els.size(); elnr++) {
      SetWavespeed(tel, wavespeed[tent.els[elnr]]);
      const int eli = ndomains > 1 ? macroel[tent.els[elnr]] : 0;
      CalcTentElEval(tent.els[elnr], tent, tel, &sol[eli * nbasis], topdshapes[elnr]);
    }
    if (phase_timing) {
      const double t_end = Now();
      AddPhase(4, t_eval - t_solve);
      AddPhase(5, t_end - t_eval);
      AddPhase(6, t_end - t_tent);
    }
  }

  // CalcInverseLapack plus the condition number of the system after symmetric diagonal scaling (D = sqrt(|diag A|), which
  // takes out the arbitrary scaling of the basis functions), 1-norm: two backward-stable solvers -- the reference's
  // LAPACK inverse, an elimination on the GPU -- may differ by a small multiple of cond * eps in the solution; the
  // parity tests print it where they need more than 1e-10.
  bool InvertTracked(double* a, int n) {
    std::vector<double> dscale(n);
    for (int i = 0; i < n; i++) {
      const double d = std::sqrt(std::fabs(a[(size_t)i * n + i]));
      dscale[i] = d > 0 ? d : 1.0;
    }
    auto norm1 = [&](bool inverse) {  // max absolute column sum of D^-1 A D^-1 resp. D A^-1 D
      double mx = 0;
      for (int j = 0; j < n; j++) {
        double c = 0;
        for (int i = 0; i < n; i++)
          c += std::fabs(a[(size_t)i * n + j]) * (inverse ? dscale[i] * dscale[j] : 1.0 / (dscale[i] * dscale[j]));
        mx = std::max(mx, c);
      }
      return mx;
    };
    const double na = norm1(false);
    const bool ok = CalcInverseLapack(a, n);
    const double cond = na * norm1(true);
#pragma omp critical(oracle_cond)
    max_cond = std::max(max_cond, cond);
    return ok;
  }

  virtual void PropagateTentQT(int) { throw std::runtime_error("oracle: not a quasi-Trefftz solver"); }
  bool qt = false;
  void Propagate(int nthreads, int first_tent, int ntents_limit) {
    CSR basismat = TWaveBasis(D + 1, order, fosystem, pix);
    const int nt = (int)tps->tents.size();
    int nlev = 0;
    for (const OTent& t : tps->tents) nlev = std::max(nlev, t.level + 1);
    std::vector<std::vector<int>> levels(nlev);
    for (int i = 0; i < nt; i++) levels[tps->tents[i].level].push_back(i);
    long long done = 0;
    (void)first_tent;
    for (int l = 0; l < nlev; l++) {
      const std::vector<int>& lt = levels[l];
      int cnt = (int)lt.size();
      if (ntents_limit > 0 && done + cnt > ntents_limit) cnt = (int)std::max<long long>(0, ntents_limit - done);
#pragma omp parallel for schedule(dynamic, 1) num_threads(nthreads)
      for (int k = 0; k < cnt; k++) {
        if (qt) PropagateTentQT(lt[k]);
        else PropagateTent(lt[k], basismat);
      }
      done += cnt;
      if (ntents_limit > 0 && done >= ntents_limit) break;
    }
    if (ntents_limit <= 0) timeshift += tps->dt;
  }

  // Timing aid of bench.py's CPU baseline: the listed tents, in parallel, WITHOUT regard to their dependencies -- the cost of
  // a tent does not depend on the values in its state rows, so a sample spread over all levels of a slab can be timed
  // without running the whole slab. The state afterwards is meaningless; timeshift is not advanced.
  void PropagateSample(int nthreads, const int* tents, int n) {
    CSR basismat = TWaveBasis(D + 1, order, fosystem, pix);
    // tents of one chunk must not share elements (they write rows concurrently): greedy colouring into conflict-free rounds
    std::vector<int> rest(tents, tents + n);
    std::vector<char> used(ma->nel, 0);
    while (!rest.empty()) {
      std::vector<int> round, later;
      for (int t : rest) {
        bool free_ = true;
        for (int el : tps->tents[t].els) free_ = free_ && !used[el];
        if (free_) { round.push_back(t); for (int el : tps->tents[t].els) used[el] = 1; }
        else later.push_back(t);
      }
      for (int t : round) for (int el : tps->tents[t].els) used[el] = 0;
#pragma omp parallel for schedule(dynamic, 1) num_threads(nthreads)
      for (int k = 0; k < (int)round.size(); k++) {
        if (qt) PropagateTentQT(round[k]);
        else PropagateTent(round[k], basismat);
      }
      rest.swap(later);
    }
  }

  double ElWeight(int elnr, int q) const {
    const int* ev = ma->elverts(elnr);
    double J[9];
    for (int i = 0; i < D; i++)
      for (int d = 0; d < D; d++) J[d * D + i] = ma->point(ev[i])[d] - ma->point(ev[D])[d];
    double det;
    if (D == 1) det = J[0];
    else if (D == 2) det = J[0] * J[3] - J[1] * J[2];
    else det = J[0] * (J[4] * J[8] - J[5] * J[7]) - J[1] * (J[3] * J[8] - J[5] * J[6]) + J[2] * (J[3] * J[7] - J[4] * J[6]);
    return std::fabs(det) * ir_w[q];
  }
  double Energy(const double* wf, int ncomp) const {
    double energy = 0;
    const int w = ncomp * snip;
    const int base = fosystem ? 0 : snip;  // the reference hard-codes snip (second-order layout), twavetents.cpp:842
    for (int elnr = 0; elnr < ma->nel; elnr++)
      for (int imip = 0; imip < snip; imip++)
        for (int d = 0; d < D + 1; d++)
          energy += 0.5 * (std::pow(CAt(elnr, imip), -2 * (d == D)) * std::pow(wf[(size_t)elnr * w + base + d * snip + imip], 2) *
                           ElWeight(elnr, imip));
    return energy;
  }
  double Error(const double* a, const double* b, int ncomp) const {
    double error = 0;
    const int w = ncomp * snip, off = !fosystem;
    for (int elnr = 0; elnr < ma->nel; elnr++)
      for (int imip = 0; imip < snip; imip++)
        for (int d = 0; d < D + 1; d++)
          error += std::pow(CAt(elnr, imip), -2 * (d == D)) *
                   std::pow(a[(size_t)elnr * w + (off + d) * snip + imip] - b[(size_t)elnr * w + (off + d) * snip + imip], 2) *
                   ElWeight(elnr, imip);
    return std::sqrt(error);
  }
  double L2Error(const double* a, const double* b, int ncomp) const {
    double l2 = 0;
    const int w = ncomp * snip;
    for (int elnr = 0; elnr < ma->nel; elnr++)
      for (int imip = 0; imip < snip; imip++) {
        const double dlt = a[(size_t)elnr * w + imip] - b[(size_t)elnr * w + imip];
        l2 += dlt * dlt * ElWeight(elnr, imip);
      }
    return std::sqrt(l2);
  }
};

}  // namespace oracle
