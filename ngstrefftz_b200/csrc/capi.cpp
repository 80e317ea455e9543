// extern "C" boundary of libngst_b200.so (declarations and reference citations: include/ngst_b200.h).
#include "../../include/ngst_b200.h"

#include <algorithm>
#include <cstring>
#include <memory>
#include <stdexcept>
#include <string>

#include "cuda/solver.hpp"
#include "host/basis.hpp"
#include "host/mesh.hpp"
#include "host/partition.hpp"
#include "host/pitcher.hpp"
#include "host/quadrature.hpp"
#include "host/stage.hpp"

using namespace ngst;

struct ngst_mesh { SimplexMesh m; };
struct ngst_slab { const ngst_mesh* mesh; TentSlab s; };
struct ngst_twave {
  const ngst_slab* slab;
  int slab_generation = 0;  // generation of *slab the device tables were staged from
  // multi-GPU: this rank's partition plan and its restricted slab (must outlive the solver, which keeps references)
  std::unique_ptr<PartitionPlan> plan;
  std::unique_ptr<TentSlab> sub;
  std::unique_ptr<Solver> solver;
};

namespace {
thread_local std::string g_err;

template <typename F>
int guarded(F f) {
  try {
    f();
    return NGST_OK;
  } catch (const std::invalid_argument& e) {
    g_err = e.what();
    return NGST_ERR_ARG;
  } catch (const std::domain_error& e) {
    g_err = e.what();
    return NGST_ERR_NUMERIC;
  } catch (const std::logic_error& e) {
    g_err = e.what();
    return NGST_ERR_STATE;
  } catch (const std::exception& e) {
    g_err = e.what();
    if (g_err.rfind("COMM:", 0) == 0) return NGST_ERR_COMM;
    return g_err.rfind("CUDA:", 0) == 0 ? NGST_ERR_CUDA : NGST_ERR_ARG;
  } catch (...) {
    g_err = "unknown error";
    return NGST_ERR_ARG;
  }
}
template <typename T>
void copy_out(T* dst, const std::vector<T>& src) {
  if (dst && !src.empty()) std::memcpy(dst, src.data(), src.size() * sizeof(T));
}
void need(const void* p, const char* what) {
  if (!p) throw std::invalid_argument(std::string("null argument: ") + what);
}
// The device tables of a solver are staged once, at creation. The reference re-reads its TentPitchedSlab on every
// Propagate (src/twavetents.cpp:72-90), so re-pitching under a live TWave is legal there; here it is refused instead of
// running the old tables against the new slab.
void need_current(const ngst_twave* t) {
  need(t, "solver");
  if (t->slab->s.generation != t->slab_generation)
    throw std::logic_error("the tent slab was re-pitched after this TWave was created: create a new TWave for the new tents");
}
}  // namespace

extern "C" {

const char* ngst_last_error(void) { return g_err.c_str(); }
const char* ngst_version(void) { return "ngst_b200 0.1 (sm_100a)"; }

int ngst_mesh_create(int D, int nv, const double* pts, int nel, const int* el2v, int nbnd, const int* bnd2v,
                     const int* bnd_kind, ngst_mesh** out) {
  return guarded([&] {
    need(out, "out");
    if (D < 1 || D > 3 || nv < 2 || nel < 1 || nbnd < 0) throw std::invalid_argument("ngst_mesh_create: bad sizes");
    need(pts, "pts");
    need(el2v, "el2v");
    if (nbnd) need(bnd2v, "bnd2v");
    auto h = std::make_unique<ngst_mesh>();
    SimplexMesh& m = h->m;
    m.D = D; m.nv = nv; m.nel = nel; m.nse = nbnd;
    m.pts.assign(pts, pts + (size_t)nv * D);
    m.el2v.assign(el2v, el2v + (size_t)nel * (D + 1));
    if (nbnd) m.se2v.assign(bnd2v, bnd2v + (size_t)nbnd * D);
    if (bnd_kind) m.se_kind.assign(bnd_kind, bnd_kind + nbnd);
    else m.se_kind.assign(nbnd, 0);
    try { m.finalize(); } catch (const std::runtime_error& e) { throw std::invalid_argument(e.what()); }
    *out = h.release();
  });
}

int ngst_mesh_structured(int D, const int* ncells, const double* lo, const double* hi, ngst_mesh** out) {
  return guarded([&] {
    need(out, "out"); need(ncells, "ncells"); need(lo, "lo"); need(hi, "hi");
    auto h = std::make_unique<ngst_mesh>();
    try { h->m = make_structured(D, ncells, lo, hi); } catch (const std::runtime_error& e) { throw std::invalid_argument(e.what()); }
    *out = h.release();
  });
}

int ngst_mesh_set_boundary_kind(ngst_mesh* m, const int* bnd_kind) {
  return guarded([&] {
    need(m, "mesh"); need(bnd_kind, "bnd_kind");
    m->m.se_kind.assign(bnd_kind, bnd_kind + m->m.nse);
  });
}

int ngst_mesh_counts(const ngst_mesh* m, int* counts) {
  return guarded([&] {
    need(m, "mesh"); need(counts, "counts");
    const SimplexMesh& s = m->m;
    int c[8] = {s.D, s.nv, s.nel, s.nse, s.nedges, s.nfacets, 0, 0};
    std::memcpy(counts, c, sizeof(c));
  });
}

int ngst_mesh_tables(const ngst_mesh* m, double* pts, int* el2v, int* bnd2v, int* bnd_kind, int* edge2v,
                     int* facet2v, int* el2facet, int* facet2el, int* facet2bnd, int* el2edge) {
  return guarded([&] {
    need(m, "mesh");
    const SimplexMesh& s = m->m;
    copy_out(pts, s.pts); copy_out(el2v, s.el2v); copy_out(bnd2v, s.se2v); copy_out(bnd_kind, s.se_kind);
    copy_out(edge2v, s.edge2v); copy_out(facet2v, s.facet2v); copy_out(el2facet, s.el2facet);
    copy_out(facet2el, s.facet2el); copy_out(facet2bnd, s.facet2se); copy_out(el2edge, s.el2edge);
  });
}

void ngst_mesh_destroy(ngst_mesh* m) { delete m; }

int ngst_slab_create(const ngst_mesh* m, int method, ngst_slab** out) {
  return guarded([&] {
    need(m, "mesh"); need(out, "out");
    if (method != kEdgeGrad && method != kVolGrad) throw std::invalid_argument("ngst_slab_create: method must be 1 (vol) or 2 (edge)");
    auto h = std::make_unique<ngst_slab>();
    h->mesh = m;
    h->s.mesh = &m->m;
    h->s.method = method;
    *out = h.release();
  });
}

int ngst_slab_set_max_wavespeed(ngst_slab* s, const double* c, int n) {
  return guarded([&] {
    need(s, "slab"); need(c, "c");
    const int nel = s->s.mesh->nel;
    if (n == 1) s->s.cmax_el.assign(nel, c[0]);
    else if (n == nel) s->s.cmax_el.assign(c, c + nel);
    else throw std::invalid_argument("SetMaxWavespeed: need 1 or nel values");
    for (double v : s->s.cmax_el)
      if (!(v > 0)) throw std::invalid_argument("SetMaxWavespeed: wavespeed must be positive");
  });
}

int ngst_slab_pitch(ngst_slab* s, double dt, int local_ct, double global_ct, int* ok) {
  return guarded([&] {
    need(s, "slab");
    if (!(dt > 0)) throw std::invalid_argument("PitchTents: dt must be positive");
    bool r = s->s.PitchTents(dt, local_ct != 0, global_ct);
    if (ok) *ok = r ? 1 : 0;
  });
}

int ngst_slab_pitch_level_sync(ngst_slab* s, double dt, int local_ct, double global_ct, int* ok) {
  return guarded([&] {
    need(s, "slab"); need(ok, "ok");
    *ok = s->s.PitchTentsLevelSync(dt, local_ct != 0, global_ct) ? 1 : 0;
  });
}

int ngst_slab_import_tents(ngst_slab* s, int ntents, double dt, const int* vertex, const int* level, const double* tbot,
                           const double* ttop, const int* nb_off, const int* nbv, const double* nbtime, const int* el_off,
                           const int* els, const int* facet_off, const int* facets) {
  return guarded([&] {
    need(s, "slab");
    s->s.ImportTents(ntents, dt, vertex, level, tbot, ttop, nb_off, nbv, nbtime, el_off, els, facet_off, facets);
  });
}

int ngst_slab_ntents(const ngst_slab* s) { return s ? s->s.ntents : 0; }
int ngst_slab_nlayers(const ngst_slab* s) { return s ? s->s.GetNLayers() : 0; }
double ngst_slab_height(const ngst_slab* s) { return s ? s->s.dt : 0.0; }
double ngst_slab_max_slope(const ngst_slab* s) { return s ? s->s.MaxSlope() : 0.0; }

int ngst_slab_sizes(const ngst_slab* s, long long* sizes) {
  return guarded([&] {
    need(s, "slab"); need(sizes, "sizes");
    sizes[0] = (long long)s->s.nbv.val.size();
    sizes[1] = (long long)s->s.els.val.size();
    sizes[2] = (long long)s->s.facets.val.size();
    sizes[3] = (long long)s->s.dag.val.size();
  });
}

int ngst_slab_export(const ngst_slab* s, int* vertex, int* level, double* tbot, double* ttop, double* maxslope,
                     int* nb_off, int* nbv, double* nbtime, int* el_off, int* els, int* facet_off, int* facets,
                     int* dag_off, int* dag) {
  return guarded([&] {
    need(s, "slab");
    const TentSlab& t = s->s;
    copy_out(vertex, t.vertex); copy_out(level, t.level); copy_out(tbot, t.tbot); copy_out(ttop, t.ttop);
    copy_out(maxslope, t.maxslope);
    copy_out(nb_off, t.nbv.off); copy_out(nbv, t.nbv.val); copy_out(nbtime, t.nbtime);
    copy_out(el_off, t.els.off); copy_out(els, t.els.val);
    copy_out(facet_off, t.facets.off); copy_out(facets, t.facets.val);
    copy_out(dag_off, t.dag.off); copy_out(dag, t.dag.val);
  });
}

int ngst_slab_levels(const ngst_slab* s, int* nlevels, int* level_off, int* level_tents) {
  return guarded([&] {
    need(s, "slab");
    if (nlevels) *nlevels = (int)s->s.level_off.size() - 1;
    copy_out(level_off, s->s.level_off);
    copy_out(level_tents, s->s.level_tents);
  });
}

void ngst_slab_destroy(ngst_slab* s) { delete s; }

int ngst_twave_create(const ngst_slab* s, int order, const double* c_el, const double* c_vertex, int device,
                      ngst_twave** out) {
  return guarded([&] {
    need(s, "slab"); need(out, "out"); need(c_el, "c_el");
    const SimplexMesh& m = *s->s.mesh;
    std::vector<double> ce(c_el, c_el + m.nel), cv;
    if (c_vertex) cv.assign(c_vertex, c_vertex + m.nv);
    auto h = std::make_unique<ngst_twave>();
    h->slab = s;
    h->slab_generation = s->s.generation;
    h->solver = std::make_unique<Solver>(s->s, order, ce, cv, device);
    *out = h.release();
  });
}

int ngst_partition_plan(const ngst_slab* s, const int* vertex_owner, int nranks, int rank, int* counts, int* own_tents,
                        int* loc2glob, int* owned, int* send_off, int* send_el, int* recv_off, int* recv_el) {
  return guarded([&] {
    need(s, "slab"); need(vertex_owner, "vertex_owner"); need(counts, "counts");
    const SimplexMesh& m = *s->s.mesh;
    const std::vector<int> vo(vertex_owner, vertex_owner + m.nv);
    const PartitionPlan p = make_partition_plan(m, s->s, vo, nranks, rank);
    counts[0] = (int)p.own_tents.size();
    counts[1] = (int)p.loc2glob.size();
    counts[2] = (int)p.send_el.size();
    counts[3] = (int)p.recv_el.size();
    counts[4] = p.nlev;
    copy_out(own_tents, p.own_tents);
    copy_out(loc2glob, p.loc2glob);
    if (owned) for (size_t i = 0; i < p.owned.size(); i++) owned[i] = p.owned[i];
    copy_out(send_off, p.send_off);
    copy_out(send_el, p.send_el);
    copy_out(recv_off, p.recv_off);
    copy_out(recv_el, p.recv_el);
  });
}

int ngst_partition_p2p(const ngst_slab* s, const int* vertex_owner, int nranks, int rank, long long* counts, int* indeg,
                       int* lsucc_off, int* lsucc, int* rsucc_off, int* rsucc_peer, int* rsucc_tent, int* push_off,
                       int* push_peer, int* push_el) {
  return guarded([&] {
    need(s, "slab"); need(vertex_owner, "vertex_owner"); need(counts, "counts");
    const SimplexMesh& m = *s->s.mesh;
    const std::vector<int> vo(vertex_owner, vertex_owner + m.nv);
    const PartitionPlan p = make_partition_plan(m, s->s, vo, nranks, rank);
    counts[0] = (long long)p.own_tents.size();
    counts[1] = (long long)p.p2p_lsucc.size();
    counts[2] = (long long)p.p2p_rsucc_peer.size();
    counts[3] = (long long)p.p2p_push_peer.size();
    counts[4] = (long long)p.p2p_push_off.size() - 1;
    copy_out(indeg, p.p2p_indeg);
    copy_out(lsucc_off, p.p2p_lsucc_off); copy_out(lsucc, p.p2p_lsucc);
    copy_out(rsucc_off, p.p2p_rsucc_off); copy_out(rsucc_peer, p.p2p_rsucc_peer); copy_out(rsucc_tent, p.p2p_rsucc_tent);
    copy_out(push_off, p.p2p_push_off); copy_out(push_peer, p.p2p_push_peer); copy_out(push_el, p.p2p_push_el);
  });
}

int ngst_twave_create_partitioned(const ngst_slab* s, int order, const double* c_el, const double* c_vertex, int device,
                                  const int* vertex_owner, int nranks, int rank, ngst_twave** out) {
  return guarded([&] {
    need(s, "slab"); need(out, "out"); need(c_el, "c_el"); need(vertex_owner, "vertex_owner");
    const SimplexMesh& m = *s->s.mesh;
    std::vector<double> ce(c_el, c_el + m.nel), cv;
    if (c_vertex) cv.assign(c_vertex, c_vertex + m.nv);
    const std::vector<int> vo(vertex_owner, vertex_owner + m.nv);
    auto h = std::make_unique<ngst_twave>();
    h->slab = s;
    h->slab_generation = s->s.generation;
    h->plan = std::make_unique<PartitionPlan>(make_partition_plan(m, s->s, vo, nranks, rank));
    h->sub = std::make_unique<TentSlab>(restrict_slab(s->s, h->plan->own_tents));
    h->solver = std::make_unique<Solver>(*h->sub, order, ce, cv, device, h->plan.get());
    *out = h.release();
  });
}

namespace {
QtHostData make_qt_data(const SimplexMesh& m, int order, int cdeg, const double* ccoef, const double* aa_taylor,
                        const double* bb_taylor) {
  need(ccoef, "ccoef"); need(aa_taylor, "aa_taylor"); need(bb_taylor, "bb_taylor");
  if (order < 2 || cdeg < 0 || cdeg > 8) throw std::invalid_argument("ngst_qtwave_create: order >= 2 and 0 <= cdeg <= 8 required");
  QtHostData q;
  q.cdeg = cdeg;
  q.ccoef.assign(ccoef, ccoef + monomial_exponents(m.D, cdeg).size() / m.D);
  const size_t nA = monomial_exponents(m.D, order - 2).size() / m.D, nB = monomial_exponents(m.D, order - 1).size() / m.D;
  q.aa.assign(aa_taylor, aa_taylor + (size_t)m.nv * nA);
  q.bb.assign(bb_taylor, bb_taylor + (size_t)m.nv * nB);
  return q;
}
}  // namespace

int ngst_qtwave_create(const ngst_slab* s, int order, const double* c_el, const double* c_vertex, int device, int cdeg,
                       const double* ccoef, const double* aa_taylor, const double* bb_taylor, ngst_twave** out) {
  return guarded([&] {
    need(s, "slab"); need(out, "out"); need(c_el, "c_el");
    const SimplexMesh& m = *s->s.mesh;
    const QtHostData q = make_qt_data(m, order, cdeg, ccoef, aa_taylor, bb_taylor);
    std::vector<double> ce(c_el, c_el + m.nel), cv;
    if (c_vertex) cv.assign(c_vertex, c_vertex + m.nv);
    auto h = std::make_unique<ngst_twave>();
    h->slab = s;
    h->slab_generation = s->s.generation;
    h->solver = std::make_unique<Solver>(s->s, order, ce, cv, device, nullptr, &q);
    *out = h.release();
  });
}

int ngst_qtwave_create_partitioned(const ngst_slab* s, int order, const double* c_el, const double* c_vertex, int device,
                                   int cdeg, const double* ccoef, const double* aa_taylor, const double* bb_taylor,
                                   const int* vertex_owner, int nranks, int rank, ngst_twave** out) {
  return guarded([&] {
    need(s, "slab"); need(out, "out"); need(c_el, "c_el"); need(vertex_owner, "vertex_owner");
    const SimplexMesh& m = *s->s.mesh;
    const QtHostData q = make_qt_data(m, order, cdeg, ccoef, aa_taylor, bb_taylor);
    std::vector<double> ce(c_el, c_el + m.nel), cv;
    if (c_vertex) cv.assign(c_vertex, c_vertex + m.nv);
    const std::vector<int> vo(vertex_owner, vertex_owner + m.nv);
    auto h = std::make_unique<ngst_twave>();
    h->slab = s;
    h->slab_generation = s->s.generation;
    h->plan = std::make_unique<PartitionPlan>(make_partition_plan(m, s->s, vo, nranks, rank));
    h->sub = std::make_unique<TentSlab>(restrict_slab(s->s, h->plan->own_tents));
    h->solver = std::make_unique<Solver>(*h->sub, order, ce, cv, device, h->plan.get(), &q);
    *out = h.release();
  });
}

int ngst_twave_set_wavespeed_poly(ngst_twave* t, int cdeg, const double* ccoef) {
  return guarded([&] {
    need(t, "solver");
    if (ccoef && (cdeg < 0 || cdeg > 8)) throw std::invalid_argument("ngst_twave_set_wavespeed_poly: 0 <= cdeg <= 8 required");
    if (t->solver->is_qt() && !ccoef) throw std::logic_error("ngst_twave_set_wavespeed_poly: the quasi-Trefftz solver needs its wavespeed polynomial");
    t->solver->SetWavespeedPoly(cdeg, ccoef);
  });
}

int ngst_comm_unique_id(void* out128) {
  const int rc = guarded([&] { need(out128, "out128"); halo_unique_id(out128); });
  return rc == NGST_OK ? rc : NGST_ERR_COMM;
}

int ngst_twave_comm_init(ngst_twave* t, const void* id128) {
  const int rc = guarded([&] { need(t, "solver"); need(id128, "id128"); t->solver->InitComm(id128); });
  return (rc == NGST_OK || rc == NGST_ERR_STATE || rc == NGST_ERR_CUDA) ? rc : NGST_ERR_COMM;
}

int ngst_twave_partition(const ngst_twave* t, int* own_tents, int* loc2glob, int* owned) {
  return guarded([&] {
    need(t, "solver");
    if (!t->plan) throw std::logic_error("ngst_twave_partition: the solver was not created with a partition");
    copy_out(own_tents, t->plan->own_tents);
    copy_out(loc2glob, t->plan->loc2glob);
    if (owned) for (size_t i = 0; i < t->plan->owned.size(); i++) owned[i] = t->plan->owned[i];
  });
}

int ngst_twave_info(const ngst_twave* t, int* info) {
  return guarded([&] {
    need(t, "solver"); need(info, "info");
    const Solver& s = *t->solver;
    int v[8] = {s.D(), s.order(), s.nbasis(), s.nip(), s.ncomp(), s.fosystem(), s.nel(), s.ntents()};
    std::memcpy(info, v, sizeof(v));
  });
}

int ngst_twave_quadrature(const ngst_twave* t, double* x, double* weight) {
  return guarded([&] { need(t, "solver"); t->solver->Quadrature(x, weight); });
}

int ngst_twave_set_initial(ngst_twave* t, const double* wf, int ncomp) {
  return guarded([&] { need(t, "solver"); need(wf, "wf"); t->solver->SetInitial(wf, ncomp); });
}

long long ngst_twave_boundary_npoints(const ngst_twave* t) { return t ? t->solver->staged().nbnd_points : 0; }
long long ngst_twave_boundary_nvalues(const ngst_twave* t) { return t ? t->solver->staged().nbnd_values : 0; }

int ngst_twave_set_initial_streamed(ngst_twave* t, const double* wf, int ncomp) {
  return guarded([&] { need(t, "solver"); need(wf, "wf"); t->solver->SetInitialStreamed(wf, ncomp); });
}

int ngst_twave_prefetch_input(ngst_twave* t, const double* wf, int ncomp, const double* bnd_values, long long nbnd) {
  return guarded([&] { need(t, "solver"); need(wf, "wf"); t->solver->PrefetchInput(wf, ncomp, bnd_values, nbnd); });
}
int ngst_twave_use_prefetched(ngst_twave* t) {
  return guarded([&] { need_current(t); t->solver->UsePrefetched(); });
}
int ngst_twave_get_wavefront_async(ngst_twave* t, double* out) {
  return guarded([&] { need(t, "solver"); need(out, "out"); t->solver->GetWavefrontAsync(out); });
}
int ngst_twave_wait_output(ngst_twave* t) {
  return guarded([&] { need(t, "solver"); t->solver->WaitOutput(); });
}

int ngst_twave_boundary_points(const ngst_twave* t, double* xt, int* kind_per_point) {
  return guarded([&] {
    need(t, "solver");
    const StagedTents& st = t->solver->staged();
    if (xt) boundary_sample_points(st, t->solver->rule(), xt);
    if (kind_per_point) {
      const int Q = t->solver->nip();
      for (int tf = 0; tf < st.ntf; tf++)
        if (st.tf_point[tf] >= 0)
          for (int q = 0; q < Q; q++) kind_per_point[st.tf_point[tf] + q] = st.tf_kind[tf];
    }
  });
}

int ngst_twave_set_boundary_values(ngst_twave* t, const double* values, long long n) {
  return guarded([&] { need(t, "solver"); t->solver->SetBoundaryValues(values, n); });
}

int ngst_twave_set_initial_analytic(ngst_twave* t, int family, const double* params, int nparams, int first_order) {
  return guarded([&] { need(t, "solver"); t->solver->SetInitialAnalytic(family, params, nparams, first_order); });
}

int ngst_twave_make_wavefront_analytic(ngst_twave* t, int family, const double* params, int nparams, double time,
                                       int first_order, double* out) {
  return guarded([&] {
    need(t, "solver");
    t->solver->MakeWavefrontAnalytic(family, params, nparams, time, first_order, out);
  });
}

int ngst_twave_set_boundary_analytic(ngst_twave* t, int family, const double* params, int nparams) {
  return guarded([&] { need(t, "solver"); t->solver->SetBoundaryAnalytic(family, params, nparams); });
}

int ngst_twave_propagate(ngst_twave* t) {
  return guarded([&] { need_current(t); t->solver->Propagate(); });
}

int ngst_twave_propagate_levels(ngst_twave* t, int nlevels) {
  return guarded([&] {
    need_current(t);
    if (nlevels < 0) throw std::invalid_argument("ngst_twave_propagate_levels: nlevels must be >= 0");
    t->solver->Propagate(nlevels);
  });
}

double ngst_twave_timeshift(const ngst_twave* t) { return t ? t->solver->timeshift() : 0.0; }
int ngst_twave_set_timeshift(ngst_twave* t, double timeshift) {
  return guarded([&] { need(t, "solver"); t->solver->set_timeshift(timeshift); });
}
const char* ngst_twave_kernel_name(const ngst_twave* t) { return t ? t->solver->kernel_name() : ""; }
int ngst_twave_set_kernel_mode(ngst_twave* t, int mode) {
  return guarded([&] {
    need(t, "solver");
    if (mode < 0 || mode > 3) throw std::invalid_argument("ngst_twave_set_kernel_mode: mode must be 0, 1, 2 or 3");
    t->solver->set_kernel_mode(mode);
  });
}
int ngst_twave_ntents_fast(const ngst_twave* t) { return t ? t->solver->ntents_fast() : 0; }
double ngst_twave_last_ms(const ngst_twave* t) { return t ? t->solver->last_ms() : 0.0; }
int ngst_twave_last_launches(const ngst_twave* t) { return t ? t->solver->last_launches() : 0; }

int ngst_twave_get_wavefront(ngst_twave* t, double* out) {
  return guarded([&] { need(t, "solver"); need(out, "out"); t->solver->GetWavefront(out); });
}

int ngst_twave_get_wave(ngst_twave* t, int* sizes, double* coef, double* psi_at_points, double* mono_coef, int* expo) {
  return guarded([&] { need(t, "solver"); t->solver->GetWave(sizes, coef, psi_at_points, mono_coef, expo); });
}

int ngst_twave_set_output_mirror(ngst_twave* t, double* host) {
  return guarded([&] { need_current(t); t->solver->SetOutputMirror(host); });
}

int ngst_twave_set_compact_state(ngst_twave* t, int on) {
  return guarded([&] { need(t, "solver"); t->solver->set_compact_state(on != 0); });
}
int ngst_twave_compact_state(const ngst_twave* t) { return t && t->solver->compact_state_active() ? 1 : 0; }

int ngst_twave_set_factor_cache(ngst_twave* t, int on) {
  return guarded([&] {
    need(t, "solver");
    if (on < 0 || on > 6) throw std::invalid_argument("ngst_twave_set_factor_cache: on must be 0..6");
    // generic kernel, stored inverses: 3 = automatic (default: the 1-D warp-per-tent kernel at orders 4..6), 4 = a CTA per
    // tent, 5 = the CTA code executed by a warp per tent (A/B), 6 = the 1-D warp-per-tent kernel at every order it is compiled for
    if (on >= 3) t->solver->set_generic_warp_apply(on == 3 ? 2 : (on == 5 ? 1 : (on == 6 ? 3 : 0)));
    else t->solver->set_factor_cache(on != 0, on == 2);
  });
}
int ngst_twave_factor_cache(const ngst_twave* t, long long* bytes) {
  if (bytes) *bytes = t ? (long long)t->solver->factor_cache_bytes() : 0;
  return t ? t->solver->factor_cache_state() : 0;
}

int ngst_twave_set_peer_exchange(ngst_twave* t, int on) {
  return guarded([&] { need(t, "solver"); t->solver->set_peer_exchange(on != 0, on == 2); });
}
int ngst_twave_peer_exchange(const ngst_twave* t, char* note, int note_len) {
  if (!t) return 0;
  if (note && note_len > 0) {
    std::strncpy(note, t->solver->peer_exchange_note().c_str(), (size_t)note_len - 1);
    note[note_len - 1] = 0;
  }
  return t->solver->peer_exchange_active() ? 1 : 0;
}

int ngst_twave_set_mirror_ctas(ngst_twave* t, int ctas) {
  return guarded([&] {
    need(t, "solver");
    if (ctas < 1 || ctas > 1024) throw std::invalid_argument("ngst_twave_set_mirror_ctas: ctas must be in 1..1024");
    t->solver->set_mirror_ctas(ctas);
  });
}

int ngst_twave_energy(ngst_twave* t, const double* wf, double* out) {
  return guarded([&] { need(t, "solver"); need(out, "out"); *out = t->solver->Energy(wf); });
}

int ngst_twave_energy_resident_noalloc(ngst_twave* t, double* out) {
  return guarded([&] { need(t, "solver"); need(out, "out"); t->solver->EnergyResidentNoAlloc(out); });
}

int ngst_twave_energy_noalloc_max_elements(void) { return ngst::static_reduce_max_elements(); }

int ngst_twave_error(ngst_twave* t, const double* wf, const double* wf_ref, double* out) {
  return guarded([&] {
    need(t, "solver"); need(wf, "wf"); need(wf_ref, "wf_ref"); need(out, "out");
    *out = t->solver->Error(wf, wf_ref, false);
  });
}

int ngst_twave_error_analytic(ngst_twave* t, int family, const double* params, int nparams, double time, int l2only,
                              double* out) {
  return guarded([&] {
    need(t, "solver"); need(out, "out");
    *out = t->solver->ErrorAnalytic(family, params, nparams, time, l2only != 0);
  });
}

int ngst_twave_l2error(ngst_twave* t, const double* wf, const double* wf_ref, double* out) {
  return guarded([&] {
    need(t, "solver"); need(wf, "wf"); need(wf_ref, "wf_ref"); need(out, "out");
    *out = t->solver->Error(wf, wf_ref, true);
  });
}

int ngst_twave_max_adiam(const ngst_twave* t, double* out) {
  return guarded([&] { need(t, "solver"); need(out, "out"); *out = t->solver->MaxAdiam(); });
}

int ngst_twave_basis(const ngst_twave* t, int* sizes, int* rowptr, int* col, double* val) {
  return guarded([&] {
    need(t, "solver");
    const BasisCsr& b = t->solver->basis();
    if (sizes) { sizes[0] = b.nbasis; sizes[1] = b.npoly; sizes[2] = (int)b.val.size(); }
    copy_out(rowptr, b.rowptr); copy_out(col, b.col); copy_out(val, b.val);
  });
}

void ngst_twave_destroy(ngst_twave* t) { delete t; }

int ngst_device_fp64_peak(int device, int which, double* tflops) {
  return guarded([&] {
    need(tflops, "tflops");
    if (which < 0 || which > 2) throw std::invalid_argument("ngst_device_fp64_peak: which must be 0, 1 or 2");
    cudaError_t e = cudaSetDevice(device);
    if (e == cudaSuccess) e = measure_fp64_peak(which, 5, tflops);
    if (e != cudaSuccess) throw std::runtime_error(std::string("CUDA: fp64 peak: ") + cudaGetErrorString(e));
  });
}

int ngst_basis(int D, int order, int fowave, int* sizes, int* rowptr, int* col, double* val, int* expo) {
  return guarded([&] {
    if (D < 1 || D > 3) throw std::invalid_argument("ngst_basis: D must be 1..3");
    BasisCsr b;
    try { b = twave_basis(D + 1, order, fowave); } catch (const std::runtime_error& e) { throw std::invalid_argument(e.what()); }
    if (sizes) { sizes[0] = b.nbasis; sizes[1] = b.npoly; sizes[2] = (int)b.val.size(); }
    copy_out(rowptr, b.rowptr); copy_out(col, b.col); copy_out(val, b.val); copy_out(expo, b.expo);
  });
}

int ngst_rule(int dim, int order, int* nip, double* xi, double* w) {
  return guarded([&] {
    SimplexRule r;
    try { r = make_simplex_rule(dim, order); } catch (const std::runtime_error& e) { throw std::invalid_argument(e.what()); }
    if (nip) *nip = r.nip;
    copy_out(xi, r.xi); copy_out(w, r.w);
  });
}

int ngst_rule_override(int dim, int order, int nip, const double* xi, const double* w) {
  return guarded([&] {
    try { set_simplex_rule_override(dim, order, nip, xi, w); }
    catch (const std::invalid_argument&) { throw; }
    catch (const std::runtime_error& e) { throw std::invalid_argument(e.what()); }
  });
}

int ngst_rule_transfer(int dim, int order_from, int degree_to, int poly_degree, int* sizes, double* M) {
  return guarded([&] {
    try {
      const SimplexRule A = make_simplex_rule(dim, order_from), B = make_compact_rule(dim, degree_to);
      if (sizes) { sizes[0] = B.nip; sizes[1] = A.nip; }
      if (M) copy_out(M, rule_transfer_matrix(A, B, poly_degree));
    } catch (const std::invalid_argument&) { throw; }
    catch (const std::domain_error&) { throw; }
    catch (const std::runtime_error& e) { throw std::invalid_argument(e.what()); }
  });
}

int ngst_compact_rule(int dim, int degree, int* nip, double* xi, double* w) {
  return guarded([&] {
    SimplexRule r;
    try { r = make_compact_rule(dim, degree); } catch (const std::runtime_error& e) { throw std::invalid_argument(e.what()); }
    if (nip) *nip = r.nip;
    copy_out(xi, r.xi); copy_out(w, r.w);
  });
}

}  // extern "C"
