// TEST INFRASTRUCTURE ONLY (see oracle_common.hpp): C interface of the CPU oracle, loaded by tests/ and bench.py.
#include <cstring>
#include <memory>
#include <stdexcept>
#include <string>

#include "oracle_twave.cpp"
#include "oracle_qt.cpp"

using namespace oracle;

struct OracleHandle {
  OMesh ma;
  OSlab slab;
  std::unique_ptr<TWaveTents> tw;
};

static thread_local std::string g_err;
template <typename F>
static int guarded(F f) {
  try { f(); return 0; } catch (const std::exception& e) { g_err = e.what(); return -1; }
}

extern "C" {

const char* oracle_last_error() { return g_err.c_str(); }

void* oracle_create(int D, int nv, const double* pts, int nel, const int* el2v, int nse, const int* se2v,
                    const int* se_kind, int nedges, const int* edge2v, const int* el2edge, int nfacets,
                    const int* facet2v, const int* el2facet, const int* facet2el, const int* facet2se) {
  auto* h = new OracleHandle();
  OMesh& m = h->ma;
  m.D = D; m.nv = nv; m.nel = nel; m.nse = nse; m.nedges = nedges; m.nfacets = nfacets;
  m.pts.assign(pts, pts + (size_t)nv * D);
  m.el2v.assign(el2v, el2v + (size_t)nel * (D + 1));
  if (nse) { m.se2v.assign(se2v, se2v + (size_t)nse * D); m.se_kind.assign(se_kind, se_kind + nse); }
  m.edge2v.assign(edge2v, edge2v + (size_t)nedges * 2);
  m.el2edge.assign(el2edge, el2edge + (size_t)nel * m.nepe());
  m.facet2v.assign(facet2v, facet2v + (size_t)nfacets * D);
  m.el2facet.assign(el2facet, el2facet + (size_t)nel * (D + 1));
  m.facet2el.assign(facet2el, facet2el + (size_t)nfacets * 2);
  m.facet2se.assign(facet2se, facet2se + nfacets);
  m.derive();
  return h;
}
void oracle_destroy(void* hp) { delete (OracleHandle*)hp; }

int oracle_pitch(void* hp, int method, const double* cmax_el, double dt, int local_ct, double global_ct) {
  auto* h = (OracleHandle*)hp;
  std::vector<double> c(cmax_el, cmax_el + h->ma.nel);
  int ok = 0;
  int rc = guarded([&] { ok = PitchTents(h->slab, h->ma, method, c, dt, local_ct != 0, global_ct) ? 1 : 0; });
  return rc ? rc : ok;
}
// the same with the partition-independent pick rule of the product's level-synchronous pitcher (rule 1)
int oracle_pitch_rule(void* hp, int method, const double* cmax_el, double dt, int local_ct, double global_ct, int rule) {
  auto* h = (OracleHandle*)hp;
  std::vector<double> c(cmax_el, cmax_el + h->ma.nel);
  int ok = 0;
  int rc = guarded([&] { ok = PitchTents(h->slab, h->ma, method, c, dt, local_ct != 0, global_ct, rule) ? 1 : 0; });
  return rc ? rc : ok;
}
int oracle_ntents(void* hp) { return (int)((OracleHandle*)hp)->slab.tents.size(); }
int oracle_nlayers(void* hp) { return ((OracleHandle*)hp)->slab.GetNLayers(); }
void oracle_sizes(void* hp, long long* sz) {
  auto* h = (OracleHandle*)hp;
  sz[0] = sz[1] = sz[2] = sz[3] = 0;
  for (const OTent& t : h->slab.tents) {
    sz[0] += t.nbv.size(); sz[1] += t.els.size(); sz[2] += t.internal_facets.size(); sz[3] += t.dependent_tents.size();
  }
}
void oracle_export(void* hp, int* vertex, int* level, double* tbot, double* ttop, double* maxslope, int* nb_off,
                   int* nbv, double* nbtime, int* el_off, int* els, int* facet_off, int* facets, int* dag_off, int* dag) {
  auto* h = (OracleHandle*)hp;
  int a = 0, b = 0, c = 0, d = 0, i = 0;
  for (const OTent& t : h->slab.tents) {
    vertex[i] = t.vertex; level[i] = t.level; tbot[i] = t.tbot; ttop[i] = t.ttop; maxslope[i] = t.maxslope;
    nb_off[i] = a; el_off[i] = b; facet_off[i] = c; dag_off[i] = d;
    for (size_t k = 0; k < t.nbv.size(); k++) { nbv[a] = t.nbv[k]; nbtime[a] = t.nbtime[k]; a++; }
    for (int e : t.els) els[b++] = e;
    for (int f : t.internal_facets) facets[c++] = f;
    for (int x : t.dependent_tents) dag[d++] = x;
    i++;
  }
  nb_off[i] = a; el_off[i] = b; facet_off[i] = c; dag_off[i] = d;
}

// import a tent list instead of pitching (e.g. one dumped from a real ngstents TentSlab.GetTent(i))
int oracle_import_tents(void* hp, int ntents, double dt, const int* vertex, const int* level, const double* tbot,
                        const double* ttop, const int* nb_off, const int* nbv, const double* nbtime, const int* el_off,
                        const int* els, const int* facet_off, const int* facets) {
  auto* h = (OracleHandle*)hp;
  return guarded([&] {
    h->slab.ma = &h->ma;
    h->slab.dt = dt;
    h->slab.tents.assign(ntents, OTent());
    int nl = 0;
    for (int i = 0; i < ntents; i++) {
      OTent& t = h->slab.tents[i];
      t.vertex = vertex[i]; t.level = level[i]; t.tbot = tbot[i]; t.ttop = ttop[i];
      t.nbv.assign(nbv + nb_off[i], nbv + nb_off[i + 1]);
      t.nbtime.assign(nbtime + nb_off[i], nbtime + nb_off[i + 1]);
      t.els.assign(els + el_off[i], els + el_off[i + 1]);
      t.internal_facets.assign(facets + facet_off[i], facets + facet_off[i + 1]);
      nl = std::max(nl, t.level);
    }
    h->slab.nlayers = nl;
    h->slab.has_been_pitched = true;
  });
}

int oracle_twave_setup(void* hp, int order, const double* c_el, const double* c_vertex, int nip, const double* xi,
                       const double* w) {
  auto* h = (OracleHandle*)hp;
  return guarded([&] {
    h->tw.reset(new TWaveTents(h->ma.D, order, &h->ma, &h->slab, c_el, c_vertex));
    TWaveTents& t = *h->tw;
    t.snip = nip;
    t.ir_points = Matrix(nip, h->ma.D);
    std::memcpy(t.ir_points.d.data(), xi, sizeof(double) * nip * h->ma.D);
    t.ir_w.assign(w, w + nip);
    t.IndexBoundary();
  });
}
// quasi-Trefftz solver (QTWaveTents): wavespeed polynomial c(x) of degree cdeg (coefficients in first-coordinate-outermost
// monomial order), Taylor tables of 1/c^2 (degree <= order-2) and B (degree <= order-1) at the mesh vertices, and the
// (D+1)-simplex rule of order 2*order for the volume terms
int oracle_qt_setup(void* hp, int order, const double* c_el, const double* c_vertex, int nip, const double* xi,
                    const double* w, int cdeg, const double* ccoef, const double* aa_taylor, const double* bb_taylor,
                    int vnip, const double* vxi, const double* vw) {
  auto* h = (OracleHandle*)hp;
  return guarded([&] {
    const int D = h->ma.D;
    if (D > 2) throw std::runtime_error("oracle: QTWaveTents exists for D = 1, 2 only (twavetents.cpp:1150-1151)");
    auto* q = new QTWaveTents(D, order, &h->ma, &h->slab, c_el, c_vertex);
    h->tw.reset(q);
    q->snip = nip;
    q->ir_points = Matrix(nip, D);
    std::memcpy(q->ir_points.d.data(), xi, sizeof(double) * nip * D);
    q->ir_w.assign(w, w + nip);
    q->cpoly.init(D, cdeg, ccoef);
    q->aa_taylor.assign(aa_taylor, aa_taylor + (size_t)h->ma.nv * q->nA);
    q->bb_taylor.assign(bb_taylor, bb_taylor + (size_t)h->ma.nv * q->nB);
    q->vir_points = Matrix(vnip, D + 1);
    std::memcpy(q->vir_points.d.data(), vxi, sizeof(double) * vnip * (D + 1));
    q->vir_w.assign(vw, vw + vnip);
    q->IndexBoundary();
  });
}
// wavespeed at every (element, quadrature point) for Energy / Error (nel * nip values); NULL restores the element values
void oracle_set_point_wavespeed(void* hp, const double* c) {
  auto* h = (OracleHandle*)hp;
  if (!c) { h->tw->c_point.clear(); return; }
  h->tw->c_point.assign(c, c + (size_t)h->ma.nel * h->tw->snip);
}
// CSR quasi-Trefftz basis of one tent for tests: Taylor tables already scaled by elsize^|alpha|; returns nnz
int oracle_qt_basis(int D, int order, const double* aa_scaled, const double* bb_scaled, int* rowptr, int* col, double* val) {
  const int d1 = D + 1;
  PolIndex pix(d1, order), pixA(D, std::max(order - 2, 0)), pixB(D, order - 1);
  Matrix BB(order, (order - 1) * (d1 == 3) + 1), AA(order - 1, (order - 2) * (d1 == 3) + 1);
  for (size_t j = 0; j < pixB.expo.size(); j++) {
    const std::vector<int>& e = pixB.expo[j];
    const int nx = e[0], ny = D > 1 ? e[1] : 0;
    BB(nx, ny) = bb_scaled[j];
    if (nx + ny < order - 1) AA(nx, ny) = aa_scaled[pixA(e)];
  }
  CSR b = QTWaveBasisMat(d1, order, AA, BB, pix);
  if (rowptr) std::memcpy(rowptr, b.rowptr.data(), sizeof(int) * b.rowptr.size());
  if (col) std::memcpy(col, b.col.data(), sizeof(int) * b.col.size());
  if (val) std::memcpy(val, b.val.data(), sizeof(double) * b.val.size());
  return (int)b.val.size();
}

int oracle_nbasis(void* hp) { return ((OracleHandle*)hp)->tw->nbasis; }

int oracle_set_initial(void* hp, const double* wf, int ncomp) {
  auto* h = (OracleHandle*)hp;
  return guarded([&] {
    TWaveTents& t = *h->tw;
    const int D = t.D;
    if (ncomp != D + 2 && ncomp != D + 1) throw std::runtime_error("oracle: bad ncomp");
    t.wavefront = Matrix(h->ma.nel, ncomp * t.snip);
    std::memcpy(t.wavefront.d.data(), wf, sizeof(double) * t.wavefront.d.size());
    t.nbasis = BinCoeff(D + t.order, t.order) + BinCoeff(D + t.order - 1, t.order - 1);
    t.fosystem = 0;
    if (ncomp == D + 1) { t.fosystem = 1; t.nbasis -= 1; }  // twavetents.hpp:198-203
  });
}
long long oracle_boundary_nvalues(void* hp) { return ((OracleHandle*)hp)->tw->bd_offset.back(); }
long long oracle_boundary_npoints(void* hp) {
  auto* h = (OracleHandle*)hp;
  long long n = 0;
  h->tw->ForEachBoundaryFacet([&](int, int, int) { n += h->tw->snip; });
  return n;
}
// sample points (x, t_local) of the boundary datum, in traversal order
void oracle_boundary_points(void* hp, double* xt, int* kind) {
  auto* h = (OracleHandle*)hp;
  TWaveTents& t = *h->tw;
  const int d1 = t.D + 1;
  long long p = 0;
  t.ForEachBoundaryFacet([&](int tn, int, int surfel) {
    TWaveTents::Vert vert = t.TentFaceVertsBnd(t.tps->tents[tn], &t.ma->se2v[(size_t)surfel * t.D]);
    Matrix pts(t.snip, d1);
    t.FacePoints(vert, pts);
    for (int q = 0; q < t.snip; q++) {
      for (int r = 0; r < d1; r++) xt[(p + q) * d1 + r] = pts(q, r);
      kind[p + q] = t.ma->se_kind[surfel];
    }
    p += t.snip;
  });
}
void oracle_set_boundary_values(void* hp, const double* v, long long n) {
  auto* h = (OracleHandle*)hp;
  if (!v) { h->tw->have_bd = false; return; }
  h->tw->bddata.assign(v, v + n);
  h->tw->have_bd = true;
}
int oracle_propagate(void* hp, int nthreads, int ntents_limit) {
  auto* h = (OracleHandle*)hp;
  int rc = guarded([&] { h->tw->Propagate(nthreads, 0, ntents_limit); });
  if (rc) return rc;
  const int s = h->tw->singular;
  h->tw->singular = 0;
  return s ? -2 : 0;
}
// largest 1-norm condition number of a tent system met since the last call (reset != 0 clears it)
double oracle_max_cond(void* hp, int reset) {
  auto* tw = ((OracleHandle*)hp)->tw.get();
  if (!tw) return 0.0;
  const double c = tw->max_cond;
  if (reset) tw->max_cond = 0.0;
  return c;
}
// bench.py's CPU baseline: the listed tents only, dependencies ignored (timing; see PropagateSample)
int oracle_propagate_sample(void* hp, int nthreads, const int* tents, int n) {
  auto* h = (OracleHandle*)hp;
  int rc = guarded([&] {
    const int nt = (int)h->tw->tps->tents.size();
    for (int i = 0; i < n; i++)
      if (tents[i] < 0 || tents[i] >= nt) throw std::runtime_error("oracle_propagate_sample: tent id out of range");
    h->tw->PropagateSample(nthreads, tents, n);
  });
  h->tw->singular = 0;
  return rc;
}
// Per-phase timers of the plain Trefftz path (names of twavetents.cpp:84-88, 186-188): on != 0 switches them on and
// clears them; out[7] = thread-seconds {calcshape, AAt, bilinearform, tentel, solve, tenteval, tent} since then.
void oracle_set_phase_timing(void* hp, int on) {
  auto* tw = ((OracleHandle*)hp)->tw.get();
  if (!tw) return;
  tw->phase_timing = on != 0;
  for (double& v : tw->phase_sec) v = 0.0;
}
void oracle_get_phase_times(void* hp, double* out) {
  auto* tw = ((OracleHandle*)hp)->tw.get();
  for (int k = 0; k < 7; k++) out[k] = tw ? tw->phase_sec[k] : 0.0;
}

// EXPERIMENT: bottom-face terms on a compact rule (cnip points xi, weights w) with the state row
// transferred by M (cnip x nip, ngst_rule_transfer). cnip == 0 switches back to the restated reference algorithm.
int oracle_set_compact_rhs(void* hp, int cnip, const double* xi, const double* w, const double* M) {
  auto* h = (OracleHandle*)hp;
  return guarded([&] {
    TWaveTents* t = h->tw.get();
    if (!t) throw std::runtime_error("oracle: setup has not been called");
    t->cnip = 0;
    if (cnip <= 0) return;
    t->c_points = Matrix(cnip, h->ma.D);
    std::memcpy(t->c_points.d.data(), xi, sizeof(double) * cnip * h->ma.D);
    t->c_w.assign(w, w + cnip);
    t->c_M = Matrix(cnip, t->snip);
    std::memcpy(t->c_M.d.data(), M, sizeof(double) * (size_t)cnip * t->snip);
    t->cnip = cnip;
  });
}

void oracle_get_wavefront(void* hp, double* out) {
  auto* h = (OracleHandle*)hp;
  std::memcpy(out, h->tw->wavefront.d.data(), sizeof(double) * h->tw->wavefront.d.size());
}
double oracle_timeshift(void* hp) { return ((OracleHandle*)hp)->tw->timeshift; }
double oracle_energy(void* hp, const double* wf, int ncomp) { return ((OracleHandle*)hp)->tw->Energy(wf, ncomp); }
double oracle_error(void* hp, const double* a, const double* b, int ncomp) { return ((OracleHandle*)hp)->tw->Error(a, b, ncomp); }
double oracle_l2error(void* hp, const double* a, const double* b, int ncomp) { return ((OracleHandle*)hp)->tw->L2Error(a, b, ncomp); }
double oracle_max_adiam(void* hp) {
  auto* h = (OracleHandle*)hp;
  double m = 0;
  for (const OTent& t : h->slab.tents) m = std::max(m, h->tw->TentAdiam(t));
  return m;
}
// CSR basis for tests: returns nnz; arrays may be NULL
int oracle_basis(int D, int order, int fowave, int* rowptr, int* col, double* val) {
  PolIndex pix(D + 1, order);
  CSR b = TWaveBasis(D + 1, order, fowave, pix);
  if (rowptr) std::memcpy(rowptr, b.rowptr.data(), sizeof(int) * b.rowptr.size());
  if (col) std::memcpy(col, b.col.data(), sizeof(int) * b.col.size());
  if (val) std::memcpy(val, b.val.data(), sizeof(double) * b.val.size());
  return (int)b.val.size();
}

}  // extern "C"
