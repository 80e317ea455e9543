// TEST INFRASTRUCTURE ONLY (see oracle_common.hpp): CPU restatement of the quasi-Trefftz tent solver
// QTWaveTents<D>::Propagate (src/twavetents.cpp:967-1144), QTWaveBasis<D+1>::Basis (src/trefftzfespace.cpp:1053-1198,
// src/trefftzfespace.hpp:298-332) and ScalarMappedElement<D+1>::CalcDDWaveOperator (src/scalarmappedfe.cpp:919-1008).
// Included by oracle_capi.cpp after oracle_twave.cpp.
//
// NGSolve's CoefficientFunction machinery is not part of the reference tree. Here the wavespeed c(x) is a polynomial in
// the spatial coordinates (coefficients in the evaluation order of PolIndex), evaluated wherever the reference evaluates
// the CF; the Taylor coefficients d^alpha(1/c^2)(x_v)/alpha! and d^alpha(B)(x_v)/alpha! at the mesh vertices, which the
// reference obtains by symbolic differentiation (ComputeDerivs, src/trefftzfespace.hpp:59-126), are inputs.
namespace oracle {

struct PolyCF {
  int D = 0, deg = 0;
  std::vector<double> coef;  // PolIndex(D, deg) order
  std::vector<std::vector<int>> expo;
  void init(int aD, int adeg, const double* c) {
    D = aD; deg = adeg;
    PolIndex pix(D, deg);
    expo = pix.expo;
    coef.assign(c, c + expo.size());
  }
  double operator()(const double* x) const {
    double s = 0;
    for (size_t j = 0; j < coef.size(); j++) {
      double m = coef[j];
      for (int d = 0; d < D; d++)
        for (int k = 0; k < expo[j][d]; k++) m *= x[d];
      s += m;
    }
    return s;
  }
};

// QTWaveBasis<d1>::Basis: AA(nx, ny), BB(nx, ny) are the Taylor coefficients scaled by elsize^(nx+ny)
static CSR QTWaveBasisMat(int d1, int ord, const Matrix& AA, const Matrix& BB, const PolIndex& pix) {
  const int ndof = BinCoeff(d1 - 1 + ord, ord) + BinCoeff(d1 + ord - 2, ord - 1);
  const int npoly = BinCoeff(d1 + ord, ord);
  Matrix q(ndof, npoly);
  auto IDX = [&](int x, int y, int t) {
    std::vector<int> e(d1, 0);
    if (d1 == 3) e[1] = y;
    e[0] = x;
    e[d1 - 1] = t;
    return pix(e);
  };
  int basisn = 0;
  for (int t = 0; t < 2; t++)
    for (int x = 0; x <= ord - t; x++)
      for (int y = 0; y <= (ord - x - t) * (d1 == 3); y++) q(basisn++, IDX(x, y, t)) = 1.0;
  const double A0 = AA(0, 0);
  for (int b = 0; b < ndof; b++)
    for (int ell = 0; ell < ord - 1; ell++)
      for (int t = 0; t <= ell; t++)
        for (int x = (d1 == 2 ? ell - t : 0); x <= ell - t; x++) {
          const int y = ell - t - x;
          double nc = 0;
          for (int betax = 0; betax <= x; betax++)
            for (int betay = (d1 == 3) ? 0 : y; betay <= y; betay++) {
              const double den = (t + 2) * (t + 1) * A0;
              nc += (betax + 2) * (betax + 1) / den * BB(x - betax, y - betay) * q(b, IDX(betax + 2, betay, t)) +
                    (x - betax + 1) * (betax + 1) / den * BB(x - betax + 1, y - betay) * q(b, IDX(betax + 1, betay, t));
              if (d1 == 3)
                nc += (betay + 2) * (betay + 1) / den * BB(x - betax, y - betay) * q(b, IDX(betax, betay + 2, t)) +
                      (y - betay + 1) * (betay + 1) / den * BB(x - betax, y - betay + 1) * q(b, IDX(betax, betay + 1, t));
              if (betax + betay == x + y) continue;
              nc -= AA(x - betax, y - betay) * q(b, IDX(betax, betay, t + 2)) / A0;
            }
          q(b, IDX(x, y, t + 2)) = nc;
        }
  CSR sp;
  int spsize = 0;
  for (int i = 0; i < ndof; i++) {
    sp.rowptr.push_back(spsize);
    for (int j = 0; j < npoly; j++)
      if (q(i, j)) { spsize++; sp.col.push_back(j); sp.val.push_back(q(i, j)); }
  }
  sp.rowptr.push_back(spsize);
  return sp;
}

struct QTWaveTents : TWaveTents {
  PolyCF cpoly;
  std::vector<double> aa_taylor, bb_taylor;  // nv x nA, nv x nB (unscaled Taylor coefficients at the vertices)
  int nA = 0, nB = 0;
  PolIndex pixA, pixB;       // spatial multi-indices of degree <= order-2 / order-1
  Matrix vir_points;         // (D+1)-simplex rule of order 2p
  std::vector<double> vir_w;

  QTWaveTents(int aD, int aorder, const OMesh* ama, const OSlab* atps, const double* c_el, const double* c_vert)
      : TWaveTents(aD, aorder, ama, atps, c_el, c_vert), pixA(aD, std::max(aorder - 2, 0)), pixB(aD, aorder - 1) {
    qt = true;
    nA = (int)pixA.expo.size();
    nB = (int)pixB.expo.size();
  }

  double TentXdiam(const OTent& tent) const {  // twavetents.cpp:1125-1144
    const int vn = (int)tent.nbv.size();
    double diam = 0;
    auto dist = [&](const double* a, const double* b) {
      double s = 0;
      for (int d = 0; d < D; d++) s += (a[d] - b[d]) * (a[d] - b[d]);
      return std::sqrt(s);
    };
    for (int k = 0; k < vn; k++) {
      const double* v1 = ma->point(tent.nbv[k]);
      diam = std::max(diam, dist(v1, ma->point(tent.vertex)));
      for (int j = k; j < vn; j++) diam = std::max(diam, dist(v1, ma->point(tent.nbv[j])));
    }
    return diam;
  }

  CSR TentBasis(const OTent& tent, double elsize) const {  // trefftzfespace.cpp:1076-1089
    const int d1 = D + 1;
    Matrix BB(order, (order - 1) * (d1 == 3) + 1), AA(order - 1, (order - 2) * (d1 == 3) + 1);
    const double* at = &aa_taylor[(size_t)tent.vertex * nA];
    const double* bt = &bb_taylor[(size_t)tent.vertex * nB];
    for (int j = 0; j < nB; j++) {
      const std::vector<int>& e = pixB.expo[j];
      const int nx = e[0], ny = D > 1 ? e[1] : 0;
      BB(nx, ny) = bt[j] * std::pow(elsize, nx + ny);
      if (nx + ny < order - 1) AA(nx, ny) = at[pixA(e)] * std::pow(elsize, nx + ny);
    }
    return QTWaveBasisMat(d1, order, AA, BB, pix);
  }

  // dd(i, imip) = mu * (Laplace phi_i - invc2 * d_tt phi_i), scalarmappedfe.cpp:919-1008 (only the time slot is non-zero)
  void CalcDDWaveOperator(const MappedElement& tel, const Matrix& points, const std::vector<double>& invc2,
                          const std::vector<double>* mu, Matrix& dd) const {
    const int d1 = D + 1, P1 = order + 1;
    std::vector<double> polxt((size_t)d1 * P1), pol(tel.npoly);
    for (int imip = 0; imip < points.h; imip++) {
      tel.Monomials(points.Row(imip), polxt);
      for (int ii = 0; ii < tel.npoly; ii++) {
        const std::vector<int>& e = pix.expo[ii];
        double v = 0;
        for (int d = 0; d < d1; d++) {
          if (e[d] < 2) continue;
          double term = e[d] * (e[d] - 1) * std::pow(tel.scale[d], 2);
          for (int dd2 = 0; dd2 < d1; dd2++) term *= polxt[(size_t)dd2 * P1 + e[dd2] - 2 * (dd2 == d)];
          v += (d == D) ? -term * invc2[imip] : term;
        }
        pol[ii] = v * (mu ? (*mu)[imip] : 1.0);
      }
      for (int i = 0; i < nbasis; i++) {
        double s = 0;
        for (int j = tel.localmat->rowptr[i]; j < tel.localmat->rowptr[i + 1]; j++) s += tel.localmat->val[j] * pol[tel.localmat->col[j]];
        dd(i, imip) = s;
      }
    }
  }

  void PropagateTentQT(int tentnr) override {
    const OTent& tent = tps->tents[tentnr];
    const int d1 = D + 1;
    const double tentsize = TentXdiam(tent);
    const CSR basismat = TentBasis(tent, tentsize);
    MappedElement tel;
    tel.d1 = d1; tel.ndof = nbasis; tel.order = order; tel.npoly = BinCoeff(d1 + order, order);
    tel.localmat = &basismat; tel.pix = &pix;
    for (int d = 0; d < D; d++) tel.shift[d] = ma->point(tent.vertex)[d];
    tel.shift[D] = (tent.ttop - tent.tbot) / 2 + tent.tbot;
    for (int d = 0; d < d1; d++) tel.scale[d] = 1.0 / tentsize;
    const int n = nbasis;
    std::vector<double> elmat((size_t)n * n, 0.0), elvec(n, 0.0);

    // space-like faces with the wavespeed at the element's quadrature points (twavetents.cpp:1001-1016)
    std::vector<Matrix> topdshapes(tent.els.size());
    for (size_t k = 0; k < tent.els.size(); k++) {
      const int elnr = tent.els[k];
      Matrix pts(snip, d1);
      ElementPoints(elnr, TentFaceVerts(tent, elnr, 1), pts);
      std::vector<double> cq(snip);
      for (int q = 0; q < snip; q++) cq[q] = cpoly(pts.Row(q));
      CalcTentEl(elnr, tent, tel, 0.0, elmat.data(), n, elvec.data(), topdshapes[k], cq.data());
    }
    // boundary facets (twavetents.cpp:1018-1030)
    long long bdoff = have_bd ? bd_offset[tentnr] : 0;
    for (int fnr : tent.internal_facets) {
      const int e0 = ma->facet2el[2 * fnr], e1 = ma->facet2el[2 * fnr + 1];
      if (e1 >= 0 || e0 < 0) continue;
      const int surfel = ma->facet2se[fnr];
      if (surfel < 0) continue;
      Matrix pts(snip, d1);
      FacePoints(TentFaceVertsBnd(tent, &ma->se2v[(size_t)surfel * D]), pts);
      std::vector<double> invc2(snip);
      for (int q = 0; q < snip; q++) invc2[q] = 1.0 / std::pow(cpoly(pts.Row(q)), 2);
      CalcTentBndEl(surfel, tent, tel, have_bd ? &bddata[bdoff] : nullptr, 0.0, elmat.data(), n, elvec.data(), invc2.data());
      bdoff += (long long)(ma->se_kind[surfel] == 1 ? D : 1) * snip;
    }
    // volume terms over the bottom and top cone of every patch element (twavetents.cpp:1032-1108)
    const int vnip = (int)vir_w.size();
    for (size_t k = 0; k < tent.els.size(); k++)
      for (int part = -1; part <= 1; part += 2) {
        Vert map = TentFaceVerts(tent, tent.els[k], part);
        for (int i = 0; i < d1; i++)
          for (int r = 0; r < d1; r++) V(map, r, i) -= tel.shift[r];
        double det;
        if (d1 == 2) det = V(map, 0, 0) * V(map, 1, 1) - V(map, 0, 1) * V(map, 1, 0);
        else
          det = V(map, 0, 0) * (V(map, 1, 1) * V(map, 2, 2) - V(map, 1, 2) * V(map, 2, 1)) -
                V(map, 0, 1) * (V(map, 1, 0) * V(map, 2, 2) - V(map, 1, 2) * V(map, 2, 0)) +
                V(map, 0, 2) * (V(map, 1, 0) * V(map, 2, 1) - V(map, 1, 1) * V(map, 2, 0));
        const double vol = std::fabs(det);
        if (vol < 10e-16) continue;
        Matrix pts(vnip, d1);
        for (int q = 0; q < vnip; q++)
          for (int r = 0; r < d1; r++) {
            double x = tel.shift[r];
            for (int i = 0; i < d1; i++) x += V(map, r, i) * vir_points(q, i);
            pts(q, r) = x;
          }
        std::vector<double> invc2(vnip), mu(vnip);
        double cmax = 0;
        for (int q = 0; q < vnip; q++) {
          invc2[q] = 1.0 / std::pow(cpoly(pts.Row(q)), 2);
          cmax = std::max(cmax, std::fabs(invc2[q]));
        }
        for (int q = 0; q < vnip; q++) mu[q] = tentsize / cmax * std::pow(cpoly(pts.Row(q)), 2);
        Matrix dd(nbasis, vnip), ddcor(nbasis, vnip), bb(nbasis, d1 * vnip);
        CalcDDWaveOperator(tel, pts, invc2, nullptr, dd);
        for (int i = 0; i < nbasis; i++)
          for (int q = 0; q < vnip; q++) dd(i, q) *= vol * vir_w[q];
        tel.CalcDShape(pts, bb);
        CalcDDWaveOperator(tel, pts, invc2, &mu, ddcor);
        for (int i = 0; i < nbasis; i++)
          for (int j = 0; j < nbasis; j++) {
            double s = 0;
            for (int q = 0; q < vnip; q++) s += dd(i, q) * bb(j, D * vnip + q) + ddcor(i, q) * dd(j, q);
            elmat[(size_t)i * n + j] += s;
          }
      }
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
    for (size_t k = 0; k < tent.els.size(); k++) CalcTentElEval(tent.els[k], tent, tel, sol.data(), topdshapes[k]);
  }
};

}  // namespace oracle
