#include "pitcher.hpp"

#include <algorithm>
#include <climits>
#include <cmath>
#include <cstdio>
#include <limits>
#include <stdexcept>

namespace ngst {

namespace {

// Gradients of the barycentric (P1 nodal) functions of element el: g[k*D + d] = d lambda_k / d x_d.
// Reference elements as in NGSolve's ScalarFE<ET,1>: lambda = (xi_0, .., xi_{D-1}, 1 - sum xi),
// x = P_D + sum_i xi_i (P_i - P_D)  (used at tents.cpp:332-352, 666, 873).
void bary_grads(const SimplexMesh& m, int el, double* g, double* absdet) {
  const int D = m.D;
  const int* v = m.elverts(el);
  double J[9], inv[9];
  for (int i = 0; i < D; i++)
    for (int d = 0; d < D; d++) J[d * D + i] = m.point(v[i])[d] - m.point(v[D])[d];
  double det;
  if (D == 1) {
    det = J[0];
    inv[0] = 1.0 / det;
  } else if (D == 2) {
    det = J[0] * J[3] - J[1] * J[2];
    inv[0] = J[3] / det; inv[1] = -J[1] / det;
    inv[2] = -J[2] / det; inv[3] = J[0] / det;
  } else {
    double c00 = J[4] * J[8] - J[5] * J[7], c01 = J[5] * J[6] - J[3] * J[8], c02 = J[3] * J[7] - J[4] * J[6];
    det = J[0] * c00 + J[1] * c01 + J[2] * c02;
    inv[0] = c00 / det; inv[1] = (J[2] * J[7] - J[1] * J[8]) / det; inv[2] = (J[1] * J[5] - J[2] * J[4]) / det;
    inv[3] = c01 / det; inv[4] = (J[0] * J[8] - J[2] * J[6]) / det; inv[5] = (J[2] * J[3] - J[0] * J[5]) / det;
    inv[6] = c02 / det; inv[7] = (J[1] * J[6] - J[0] * J[7]) / det; inv[8] = (J[0] * J[4] - J[1] * J[3]) / det;
  }
  // grad lambda_i = row i of J^{-1}; grad lambda_D = - sum of the others
  for (int d = 0; d < D; d++) g[D * D + d] = 0;
  for (int i = 0; i < D; i++)
    for (int d = 0; d < D; d++) {
      g[i * D + d] = inv[i * D + d];
      g[D * D + d] -= inv[i * D + d];
    }
  if (absdet) *absdet = std::fabs(det);
}

inline double norm(const double* a, int D) {
  double s = 0;
  for (int d = 0; d < D; d++) s += a[d] * a[d];
  return std::sqrt(s);
}

int local_index(const SimplexMesh& m, int el, int v) {
  const int* ev = m.elverts(el);
  for (int k = 0; k <= m.D; k++)
    if (ev[k] == v) return k;
  return -1;
}

// Ready list with the positional semantics of the ngcore Array used at tents.cpp:135,160-168,458-471.
class ReadyList {
 public:
  ReadyList(int nv, const std::vector<int>& level) : level_(level) {
    cap_ = 2;
    while (cap_ < 2 * nv + 2) cap_ <<= 1;
    slot_vertex_.assign(cap_, -1);
    vslot_.assign(nv, -1);
    tmin_.assign(2 * cap_, INT_MAX);
    tcnt_.assign(2 * cap_, 0);
  }
  int size() const { return tcnt_[1]; }
  bool contains(int v) const { return vslot_[v] >= 0; }
  void append(int v) {
    if (end_ == cap_) compact();
    set_slot(end_++, v);
  }
  // first position holding the minimal level (PickNextVertexForPitching, tents.cpp:431-443)
  int pick(int* minlevel) const {
    int node = 1;
    *minlevel = tmin_[1];
    while (node < cap_) node = (tmin_[2 * node] == tmin_[node]) ? 2 * node : 2 * node + 1;
    return node - cap_;
  }
  int vertex_at(int slot) const { return slot_vertex_[slot]; }
  // Array::DeleteElement: the LAST element is moved into the freed position
  void delete_element(int slot) {
    int node = 1;
    while (node < cap_) node = tcnt_[2 * node + 1] > 0 ? 2 * node + 1 : 2 * node;
    int last = node - cap_;
    if (last == slot) { clear_slot(slot); return; }
    int vlast = slot_vertex_[last];
    clear_slot(last);
    clear_slot(slot);
    set_slot(slot, vlast);
  }
  // Array::RemoveElement(Pos(v)): order-preserving erase
  void remove(int v) {
    if (vslot_[v] >= 0) clear_slot(vslot_[v]);
  }
  void level_changed(int v) {
    if (vslot_[v] >= 0) set_key(vslot_[v], level_[v], 1);
  }

 private:
  void set_key(int slot, int key, int cnt) {
    int node = slot + cap_;
    tmin_[node] = key;
    tcnt_[node] = cnt;
    for (node >>= 1; node >= 1; node >>= 1) {
      const int mn = std::min(tmin_[2 * node], tmin_[2 * node + 1]);
      const int cn = tcnt_[2 * node] + tcnt_[2 * node + 1];
      if (mn == tmin_[node] && cn == tcnt_[node]) break;  // nothing changes further up
      tmin_[node] = mn;
      tcnt_[node] = cn;
    }
  }
  void set_slot(int slot, int v) {
    slot_vertex_[slot] = v;
    vslot_[v] = slot;
    set_key(slot, level_[v], 1);
  }
  void clear_slot(int slot) {
    vslot_[slot_vertex_[slot]] = -1;
    slot_vertex_[slot] = -1;
    set_key(slot, INT_MAX, 0);
  }
  void compact() {
    std::vector<int> live;
    for (int s = 0; s < end_; s++)
      if (slot_vertex_[s] >= 0) live.push_back(slot_vertex_[s]);
    std::fill(slot_vertex_.begin(), slot_vertex_.end(), -1);
    std::fill(tmin_.begin(), tmin_.end(), INT_MAX);
    std::fill(tcnt_.begin(), tcnt_.end(), 0);
    end_ = (int)live.size();
    for (int s = 0; s < end_; s++) {
      slot_vertex_[s] = live[s];
      vslot_[live[s]] = s;
      tmin_[cap_ + s] = level_[live[s]];
      tcnt_[cap_ + s] = 1;
    }
    for (int node = cap_ - 1; node >= 1; node--) {
      tmin_[node] = std::min(tmin_[2 * node], tmin_[2 * node + 1]);
      tcnt_[node] = tcnt_[2 * node] + tcnt_[2 * node + 1];
    }
  }
  const std::vector<int>& level_;
  int cap_ = 0, end_ = 0;
  std::vector<int> slot_vertex_, vslot_, tmin_, tcnt_;
};

struct Pitcher {
  const SimplexMesh& m;
  int method;
  double global_ctau = 1.0;
  bool use_local = false;
  std::vector<double> cmax;       // per edge (edge algo) or per element (vol algo)
  std::vector<double> edge_len;
  std::vector<double> vertex_refdt;
  std::vector<double> local_ctau;  // aligned with v2e.val (edge) or v2el.val (vol)
  std::vector<double> edge_term;   // edge algo: global_ctau * local_ct * length / c_max per (vertex, edge), aligned with v2e.val
  std::vector<double> elgrads;     // barycentric gradients of every element, (D+1)*D doubles each (computed once, in parallel)

  const double* grads(int el) const { return &elgrads[(size_t)el * (m.D + 1) * m.D]; }
  void CacheGradients() {
    if (!elgrads.empty()) return;
    const int stride = (m.D + 1) * m.D;
    elgrads.resize((size_t)m.nel * stride);
#pragma omp parallel for schedule(static)
    for (int el = 0; el < m.nel; el++) bary_grads(m, el, &elgrads[(size_t)el * stride], nullptr);
  }

  Pitcher(const SimplexMesh& mesh, int meth) : m(mesh), method(meth) {}

  // tents.cpp:476-563
  void InitializeMeshData(const std::vector<double>& cmax_el, bool calc_local_ct, double global_ct) {
    global_ctau = global_ct;
    edge_len.assign(m.nedges, 0.0);
    if (method == kEdgeGrad) cmax.assign(m.nedges, -1.0);
    else cmax.assign(m.nel, -1.0);
    const int nepe = m.nEdgesPerEl();
    std::vector<char> fine(m.nedges, 0);
    for (int el = 0; el < m.nel; el++) {
      const double wvspd = cmax_el[el];
      if (method == kVolGrad) cmax[el] = wvspd;
      for (int k = 0; k < nepe; k++) {
        int e = m.el2edge[(size_t)el * nepe + k];
        if (!fine[e]) {
          fine[e] = 1;
          const double* p1 = m.point(m.edge2v[2 * e]);
          const double* p2 = m.point(m.edge2v[2 * e + 1]);
          double s = 0;
          for (int d = 0; d < m.D; d++) s += (p1[d] - p2[d]) * (p1[d] - p2[d]);
          edge_len[e] = std::sqrt(s);
        }
        if (method == kEdgeGrad) cmax[e] = std::max(cmax[e], wvspd);
      }
    }
    use_local = calc_local_ct && m.D > 1;
    if (use_local) {
      if (method == kEdgeGrad) CalcLocalCTauEdge();
      else CalcLocalCTauVol();
    }
    if (method == kEdgeGrad) {
      // The tau-independent summand of GetPoleHeight (tents.cpp:806-808), evaluated once per (vertex, edge) in the
      // reference's operation order, so that the pole height below is bit-identical to recomputing it every time.
      edge_term.resize(m.v2e.val.size());
#pragma omp parallel for schedule(static)
      for (int vi = 0; vi < m.nv; vi++)
        for (int i = 0; i < m.v2e.size(vi); i++) {
          const int edge = m.v2e.row(vi)[i];
          const double local_ct = use_local ? local_ctau[m.v2e.off[vi] + i] : 1.0;
          edge_term[m.v2e.off[vi] + i] = global_ctau * local_ct * edge_len[edge] / cmax[edge];
        }
    }
  }

  // EdgeGradientPitcher::CalcLocalCTau, tents.cpp:815-978. The value the reference computes for (vertex, edge) depends on
  // the edge only (both end points enter symmetrically), so it is computed once per edge, in parallel, and copied.
  void CalcLocalCTauEdge() {
    const int D = m.D;
    CacheGradients();
    std::vector<double> edge_val(m.nedges);
#pragma omp parallel for schedule(dynamic, 1024)
    for (int edge = 0; edge < m.nedges; edge++) {
      double val = std::numeric_limits<double>::max();
      const int v1 = m.edge2v[2 * edge], v2 = m.edge2v[2 * edge + 1];
      for (int ii = 0; ii < m.edge2el.size(edge); ii++) {
        const int iel = m.edge2el.row(edge)[ii];
        const double* g = grads(iel);
        const int l1 = local_index(m, iel, v1), l2 = local_index(m, iel, v2);
        double one_over_max_grad;
        if (D == 2) {
          one_over_max_grad = 1.0 / std::max(norm(&g[l1 * D], D), norm(&g[l2 * D], D));
        } else {
          // faces of iel adjacent to the edge: those opposite the two vertices NOT on the edge.
          // The physical face normal is parallel to the gradient of the opposite vertex' hat function.
          double v = 1;
          for (int k = 0; k < 4; k++) {
            if (k == l1 || k == l2) continue;
            double nrm[3];
            const double ln = norm(&g[k * 3], 3);
            for (int d = 0; d < 3; d++) nrm[d] = g[k * 3 + d] / ln;
            auto proj = [&](int lv) {
              const double ng = norm(&g[lv * 3], 3);
              double u[3];
              for (int d = 0; d < 3; d++) u[d] = g[lv * 3 + d] / ng;
              double c[3] = {nrm[1] * u[2] - nrm[2] * u[1], nrm[2] * u[0] - nrm[0] * u[2], nrm[0] * u[1] - nrm[1] * u[0]};
              return norm(c, 3) / ng;
            };
            const double val_face = std::min(proj(l1), proj(l2));
            v = std::min(v, val_face);
          }
          one_over_max_grad = v;
        }
        const double projGrad = one_over_max_grad / edge_len[edge];
        val = std::min(val, projGrad);
      }
      edge_val[edge] = val;
    }
    local_ctau.resize(m.v2e.val.size());
#pragma omp parallel for schedule(static)
    for (size_t k = 0; k < m.v2e.val.size(); k++) local_ctau[k] = edge_val[m.v2e.val[k]];
  }

  // VolumeGradientPitcher::CalcLocalCTau, tents.cpp:719-794
  void CalcLocalCTauVol() {
    const int D = m.D;
    local_ctau.assign(m.v2el.val.size(), 0.0);
    for (int vi = 0; vi < m.nv; vi++)
      for (int ie = 0; ie < m.v2el.size(vi); ie++) {
        const int el = m.v2el.row(vi)[ie];
        const int* ev = m.elverts(el);
        double val = 1.0;
        if (D == 3) {
          for (int k = 0; k < 4; k++) {  // face k is opposite local vertex k
            if (ev[k] == vi) continue;  // face does not contain vi
            int fv[3], nf = 0;
            for (int j = 0; j < 4; j++)
              if (j != k) fv[nf++] = ev[j];
            double opposite = -1, max_edge = 0;
            for (int a = 0; a < 3; a++)
              for (int b = a + 1; b < 3; b++) {
                const double* p1 = m.point(fv[a]);
                const double* p2 = m.point(fv[b]);
                double s = 0;
                for (int d = 0; d < 3; d++) s += (p1[d] - p2[d]) * (p1[d] - p2[d]);
                const double len = std::sqrt(s);
                if (fv[a] != vi && fv[b] != vi) opposite = len;
                else max_edge = std::max(len, max_edge);
              }
            val = std::min(val, opposite / max_edge);
          }
        } else {
          // 2-D: ma->GetElFaces gives the element itself as its only face
          double opposite = -1, max_edge = 0;
          for (int a = 0; a < 3; a++)
            for (int b = a + 1; b < 3; b++) {
              const double* p1 = m.point(ev[a]);
              const double* p2 = m.point(ev[b]);
              const double len = std::sqrt((p1[0] - p2[0]) * (p1[0] - p2[0]) + (p1[1] - p2[1]) * (p1[1] - p2[1]));
              if (ev[a] != vi && ev[b] != vi) opposite = len;
              else max_edge = std::max(len, max_edge);
            }
          val = std::min(val, opposite / max_edge);
        }
        val = std::min(val, 1.0);
        local_ctau[m.v2el.off[vi] + ie] = val;
      }
  }

  double GetPoleHeight(int vi, const std::vector<double>& tau) const {
    return method == kEdgeGrad ? PoleHeightEdge(vi, tau) : PoleHeightVol(vi, tau);
  }

  // EdgeGradientPitcher::GetPoleHeight, tents.cpp:796-813
  double PoleHeightEdge(int vi, const std::vector<double>& tau) const {
    double kt = std::numeric_limits<double>::max();
    const int n = m.v2v.size(vi);
    const int* nbv = m.v2v.row(vi);
    const double* term = &edge_term[m.v2e.off[vi]];  // v2v and v2e rows are aligned (neighbour i is across edge i)
    const double tv = tau[vi];
    for (int i = 0; i < n; i++) {
      const double kt1 = tau[nbv[i]] - tv + term[i];
      kt = std::min(kt, kt1);
    }
    constexpr double num_tol = std::numeric_limits<double>::epsilon();
    if (kt > num_tol) return kt * (1 - num_tol);
    return 0.0;
  }

  // VolumeGradientPitcher::GetPoleHeight, tents.cpp:627-717
  double PoleHeightVol(int vi, const std::vector<double>& tau) const {
    const int D = m.D;
    constexpr double init_pole_height = std::numeric_limits<double>::max();
    constexpr double num_tol = std::numeric_limits<double>::epsilon();
    double pole_height = init_pole_height;
    for (int iel = 0; iel < m.v2el.size(vi); iel++) {
      const int el = m.v2el.row(vi)[iel];
      const double c_max_sq = cmax[el] * cmax[el];
      const int* ev = m.elverts(el);
      const int local_vi = local_index(m, el, vi);
      const double* g = grads(el);
      double coeff[4];
      for (int k = 0; k <= D; k++) coeff[k] = tau[ev[k]];
      coeff[local_vi] = 0;
      double alpha = 0;
      for (int d = 0; d < D; d++) alpha += g[local_vi * D + d] * g[local_vi * D + d];
      double tg[3] = {0, 0, 0};
      for (int d = 0; d < D; d++)
        for (int k = 0; k <= D; k++) tg[d] += g[k * D + d] * coeff[k];
      double ip = 0, tt = 0;
      for (int d = 0; d < D; d++) { ip += tg[d] * g[local_vi * D + d]; tt += tg[d] * tg[d]; }
      const double beta = 2 * ip / alpha;
      const double gamma = (tt - 1.0 / c_max_sq) / alpha;
      const double delta = beta * beta - 4 * gamma;
      double sol;
      if (delta > num_tol * alpha) {
        if (beta <= num_tol * alpha) sol = (std::sqrt(delta) - beta) / 2.0;
        else sol = -(2.0 * gamma) / (beta + std::sqrt(delta));
      } else if (delta > -num_tol * alpha) sol = -beta / 2.0;
      else sol = init_pole_height;
      double kbar = sol - tau[vi];
      const double local_ct = use_local ? local_ctau[m.v2el.off[vi] + iel] : 1.0;
      kbar *= local_ct * global_ctau;
      pole_height = std::min(pole_height, kbar);
    }
    if (pole_height > 0.75 * init_pole_height) return 0.0;
    return pole_height * (1 - num_tol);
  }
};

}  // namespace

double TentSlab::MaxSlope() const {
  double s = 0;
  for (double x : maxslope) s = std::max(s, x);
  return s;
}

namespace {
// What follows the greedy loop (and an import): dependency table, top-front slopes, level sets.
void finalize_slab(TentSlab& S, const Pitcher& P, const std::vector<int>& dep_src, const std::vector<int>& dep_dst) {
  const SimplexMesh& m = *S.mesh;
  const int D = m.D, nv = m.nv;
  // dependency graph, tents.cpp:313-321
  // (stable counting sort by source tent: every row keeps the order in which its dependents were appended)
  S.dag.off.assign(S.ntents + 1, 0);
  for (int s : dep_src) S.dag.off[s + 1]++;
  for (int i = 0; i < S.ntents; i++) S.dag.off[i + 1] += S.dag.off[i];
  S.dag.val.resize(dep_src.size());
  {
    std::vector<int> pos(S.dag.off.begin(), S.dag.off.end() - 1);
    for (size_t k = 0; k < dep_src.size(); k++) S.dag.val[pos[dep_src[k]]++] = dep_dst[k];
  }

  // slope of the top front per tent, tents.cpp:323-356
  S.maxslope.assign(S.ntents, 0.0);
#pragma omp parallel
  {
    // front height of the patch vertices of the current tent (every S.vertex of a patch element is the tent S.vertex or one of S.nbv)
    std::vector<double> vt(nv, 0.0);
#pragma omp for schedule(static)
    for (int i = 0; i < S.ntents; i++) {
      vt[S.vertex[i]] = S.ttop[i];
      for (int q = S.nbv.off[i]; q < S.nbv.off[i + 1]; q++) vt[S.nbv.val[q]] = S.nbtime[q];
      double ms = 0.0;
      for (int j = S.els.off[i]; j < S.els.off[i + 1]; j++) {
        const int el = S.els.val[j];
        const int* ev = m.elverts(el);
        const double* g = P.grads(el);
        double gp[3] = {0, 0, 0};
        for (int d = 0; d < D; d++)
          for (int k = 0; k <= D; k++) gp[d] += g[k * D + d] * vt[ev[k]];
        const double nrm = norm(gp, D);
        if (nrm > ms) ms = nrm;
      }
      S.maxslope[i] = ms;
    }
  }

  // S.level sets (BSP schedule replacing RunParallelDependency, paralleldepend.hpp:20-131)
  int nl = 0;
  for (int i = 0; i < S.ntents; i++) nl = std::max(nl, S.level[i] + 1);
  S.level_off.assign(nl + 1, 0);
  for (int i = 0; i < S.ntents; i++) S.level_off[S.level[i] + 1]++;
  for (int l = 0; l < nl; l++) S.level_off[l + 1] += S.level_off[l];
  S.level_tents.resize(S.ntents);
  {
    std::vector<int> pos(S.level_off.begin(), S.level_off.end() - 1);
    for (int i = 0; i < S.ntents; i++) S.level_tents[pos[S.level[i]]++] = i;
  }
}
}  // namespace

bool TentSlab::PitchTents(double adt, bool calc_local_ct, double global_ct) {
  generation++;
  if (!mesh) throw std::runtime_error("TentSlab: no mesh");
  if (cmax_el.empty()) throw std::logic_error("Wavespeed has not been set!");  // tents.cpp:85-88
  const SimplexMesh& m = *mesh;
  if ((int)cmax_el.size() != m.nel) throw std::runtime_error("TentSlab: wavespeed array size mismatch");
  if (method != kEdgeGrad && method != kVolGrad) return false;
  const int D = m.D, nv = m.nv;
  dt = adt;
  pitched = false;
  nlayers = 0;
  ntents = 0;
  vertex.clear(); level.clear(); tbot.clear(); ttop.clear(); maxslope.clear();
  nbv = Csr(); els = Csr(); facets = Csr(); dag = Csr();
  nbv.off = {0}; els.off = {0}; facets.off = {0};
  nbtime.clear();

  {
    // a first guess of the slab's size (the vectors grow as needed)
    const size_t guess = (size_t)nv * 4;
    vertex.reserve(guess); level.reserve(guess); tbot.reserve(guess); ttop.reserve(guess);
  }
  Pitcher P(m, method);
  P.CacheGradients();
  P.InitializeMeshData(cmax_el, calc_local_ct, global_ct);

  std::vector<double> tau(nv, 0.0);
  // ComputeVerticesReferenceHeight, tents.cpp:420-429
  P.vertex_refdt.assign(nv, std::numeric_limits<double>::max());
  for (int i = 0; i < nv; i++) P.vertex_refdt[i] = P.GetPoleHeight(i, tau);
  std::vector<double> ktilde = P.vertex_refdt;

  std::vector<int> latest_tent(nv, -1), vertices_level(nv, 0);
  std::vector<int> dep_src, dep_dst;  // dependent_tents (tents.cpp:206) as (tent, dependent) pairs in creation order
  const double adv_factor = 0.5;             // reset after every GetReadyVertices (reset_adv_factor = true)
  ReadyList ready(nv, vertices_level);
  std::vector<char> complete(nv, 0);
  bool slab_complete = false;
  const double num_tol = std::numeric_limits<double>::epsilon() * dt;
  // D = 3: the internal-facet list of a tent (tents.cpp:219-247) depends on its vertex only: the facets of the patch
  // elements that contain the vertex, in order of first occurrence. Built once per vertex instead of once per tent.
  Csr v2f;
  if (D == 3) {
    v2f.off.assign(nv + 1, 0);
    std::vector<std::vector<int>> rows(nv);
#pragma omp parallel for schedule(dynamic, 256)
    for (int vi = 0; vi < nv; vi++) {
      std::vector<int>& r = rows[vi];
      for (int i = 0; i < m.v2el.size(vi); i++) {
        const int el = m.v2el.row(vi)[i];
        for (int k = 0; k < 4; k++) {
          const int f = m.el2facet[(size_t)el * 4 + k];
          const int* fv = &m.facet2v[(size_t)f * 3];
          if ((fv[0] == vi || fv[1] == vi || fv[2] == vi) && std::find(r.begin(), r.end(), f) == r.end()) r.push_back(f);
        }
      }
    }
    for (int vi = 0; vi < nv; vi++) v2f.off[vi + 1] = v2f.off[vi] + (int)rows[vi].size();
    v2f.val.resize(v2f.off[nv]);
    for (int vi = 0; vi < nv; vi++) std::copy(rows[vi].begin(), rows[vi].end(), v2f.val.begin() + v2f.off[vi]);
  }

  while (!slab_complete) {
    // GetReadyVertices, tents.cpp:385-418 (the list is empty on entry)
    bool found = false;
    double af = adv_factor;
    for (int ia = 0; ia < 5; ia++) {
      for (int iv = 0; iv < nv; iv++)
        if (!complete[iv] && ktilde[iv] > af * P.vertex_refdt[iv] && !ready.contains(iv)) ready.append(iv);
      if (ready.size()) { found = true; break; }
      af /= 2;
    }
    if (!found) break;

    while (ready.size()) {
      int minlevel;
      const int slot = ready.pick(&minlevel);
      nlayers = std::max(minlevel, nlayers);
      const int vi = ready.vertex_at(slot);
      ready.delete_element(slot);

      const int tid = ntents;
      vertex.push_back(vi);
      tbot.push_back(tau[vi]);
      const double new_ttop = tau[vi] + ktilde[vi];
      double tt;
      if (dt - new_ttop > num_tol) tt = new_ttop;
      else { tt = dt; complete[vi] = 1; }
      ttop.push_back(tt);
      const int lev = vertices_level[vi];
      level.push_back(lev);
      tau[vi] = tt;
      ktilde[vi] = 0;

      for (int i = 0; i < m.v2v.size(vi); i++) {
        const int nb = m.v2v.row(vi)[i];
        nbv.val.push_back(nb);
        nbtime.push_back(tau[nb]);
        if (vertices_level[nb] < lev + 1) {
          vertices_level[nb] = lev + 1;
          ready.level_changed(nb);
        }
        if (latest_tent[nb] != -1) { dep_src.push_back(latest_tent[nb]); dep_dst.push_back(tid); }
      }
      nbv.off.push_back((int)nbv.val.size());
      latest_tent[vi] = tid;
      vertices_level[vi]++;

      // internal facets, tents.cpp:219-247
      if (D == 1) facets.val.push_back(vi);
      else if (D == 2)
        for (int i = 0; i < m.v2e.size(vi); i++) facets.val.push_back(m.v2e.row(vi)[i]);
      else
        facets.val.insert(facets.val.end(), v2f.row(vi), v2f.row(vi) + v2f.size(vi));
      facets.off.push_back((int)facets.val.size());
      for (int i = 0; i < m.v2el.size(vi); i++) els.val.push_back(m.v2el.row(vi)[i]);
      els.off.push_back((int)els.val.size());

      // UpdateNeighbours, tents.cpp:445-474
      for (int i = 0; i < m.v2v.size(vi); i++) {
        const int nb = m.v2v.row(vi)[i];
        if (complete[nb]) continue;
        const double kt = P.GetPoleHeight(nb, tau);
        ktilde[nb] = kt;
        if (kt > adv_factor * P.vertex_refdt[nb]) {
          if (!ready.contains(nb)) ready.append(nb);
        } else ready.remove(nb);
      }
      ntents++;
    }
    slab_complete = true;
    for (int i = 0; i < nv; i++)
      if (!complete[i]) { slab_complete = false; break; }
  }

  if (!slab_complete) std::printf("Error: the algorithm could not pitch the whole slab\n");

  finalize_slab(*this, P, dep_src, dep_dst);
  pitched = slab_complete;
  return pitched;
}

// Level-synchronous pitcher (no reference counterpart). The reference picks "the first vertex of minimal level in the ready
// LIST" (tents.cpp:431-443); the order of that list is a history of the whole run, so its tents can only be reproduced by the
// serial loop above. This pitcher uses a rule that depends on the state alone -- the ready vertex of minimal (level, vertex
// number) -- and is therefore independent of how the work is split: all ready vertices of the minimal level L form one
// phase; pitching them in ascending order, skipping those a smaller-numbered neighbour has just lifted above L, selects the
// lexicographically first maximal independent set of the phase, whose tents do not interact. Selection is a cheap serial
// scan; the tents of a phase, their tables and the pole heights of the neighbourhood are computed in parallel. Other, on
// the tested meshes fewer, tents and levels than the reference's order; bit-identical to the oracle's SERIAL loop with the
// same rule (tests/test_pitcher.py).
bool TentSlab::PitchTentsLevelSync(double adt, bool calc_local_ct, double global_ct) {
  generation++;
  if (!mesh) throw std::runtime_error("TentSlab: no mesh");
  if (cmax_el.empty()) throw std::logic_error("Wavespeed has not been set!");
  const SimplexMesh& m = *mesh;
  if ((int)cmax_el.size() != m.nel) throw std::runtime_error("TentSlab: wavespeed array size mismatch");
  if (method != kEdgeGrad && method != kVolGrad) return false;
  const int D = m.D, nv = m.nv;
  dt = adt;
  pitched = false;
  nlayers = 0;
  ntents = 0;
  vertex.clear(); level.clear(); tbot.clear(); ttop.clear(); maxslope.clear();
  nbv = Csr(); els = Csr(); facets = Csr(); dag = Csr();
  nbv.off = {0}; els.off = {0}; facets.off = {0};
  nbtime.clear();
  Pitcher P(m, method);
  P.CacheGradients();
  P.InitializeMeshData(cmax_el, calc_local_ct, global_ct);
  std::vector<double> tau(nv, 0.0);
  P.vertex_refdt.assign(nv, std::numeric_limits<double>::max());
#pragma omp parallel for schedule(static)
  for (int i = 0; i < nv; i++) P.vertex_refdt[i] = P.GetPoleHeight(i, tau);
  std::vector<double> ktilde = P.vertex_refdt;
  std::vector<int> latest_tent(nv, -1), vlevel(nv, 0), dep_src, dep_dst;
  std::vector<char> complete(nv, 0), blocked(nv, 0), touched(nv, 0);
  const double num_tol = std::numeric_limits<double>::epsilon() * dt;
  // internal facets per vertex (D = 3), as in PitchTents
  Csr v2f;
  if (D == 3) {
    v2f.off.assign(nv + 1, 0);
    std::vector<std::vector<int>> rows(nv);
#pragma omp parallel for schedule(dynamic, 256)
    for (int vi = 0; vi < nv; vi++) {
      std::vector<int>& r = rows[vi];
      for (int i = 0; i < m.v2el.size(vi); i++) {
        const int el = m.v2el.row(vi)[i];
        for (int k = 0; k < 4; k++) {
          const int f = m.el2facet[(size_t)el * 4 + k];
          const int* fv = &m.facet2v[(size_t)f * 3];
          if ((fv[0] == vi || fv[1] == vi || fv[2] == vi) && std::find(r.begin(), r.end(), f) == r.end()) r.push_back(f);
        }
      }
    }
    for (int vi = 0; vi < nv; vi++) v2f.off[vi + 1] = v2f.off[vi] + (int)rows[vi].size();
    v2f.val.resize(v2f.off[nv]);
    for (int vi = 0; vi < nv; vi++) std::copy(rows[vi].begin(), rows[vi].end(), v2f.val.begin() + v2f.off[vi]);
  }
  auto nfacets_of = [&](int vi) { return D == 1 ? 1 : (D == 2 ? m.v2e.size(vi) : v2f.size(vi)); };

  std::vector<int> phase, sel, nbs;
  bool slab_complete = false;
  while (!slab_complete) {
    // ---- the phase: ready vertices of minimal level (ready set from the state; advance factor halved up to 5 times)
    phase.clear();
    int L = 0;
    double af = 0.5;
    for (int ia = 0; ia < 5 && phase.empty(); ia++, af /= 2) {
      int minlevel = std::numeric_limits<int>::max();
#pragma omp parallel for reduction(min : minlevel) schedule(static)
      for (int iv = 0; iv < nv; iv++)
        if (!complete[iv] && ktilde[iv] > af * P.vertex_refdt[iv]) minlevel = std::min(minlevel, vlevel[iv]);
      if (minlevel == std::numeric_limits<int>::max()) continue;
      for (int iv = 0; iv < nv; iv++)
        if (!complete[iv] && ktilde[iv] > af * P.vertex_refdt[iv] && vlevel[iv] == minlevel) phase.push_back(iv);
      L = minlevel;
    }
    if (phase.empty()) break;
    nlayers = std::max(nlayers, L);
    // ---- lexicographically first maximal independent set of the phase
    sel.clear();
    for (int v : phase) {
      if (blocked[v]) continue;
      sel.push_back(v);
      for (int i = 0; i < m.v2v.size(v); i++) blocked[m.v2v.row(v)[i]] = 1;
    }
    for (int v : sel)
      for (int i = 0; i < m.v2v.size(v); i++) blocked[m.v2v.row(v)[i]] = 0;
    // ---- tables of the new tents: offsets first, then the rows in parallel
    const int n0 = ntents, ns = (int)sel.size();
    vertex.resize(n0 + ns); level.resize(n0 + ns); tbot.resize(n0 + ns); ttop.resize(n0 + ns);
    nbv.off.resize(n0 + ns + 1); els.off.resize(n0 + ns + 1); facets.off.resize(n0 + ns + 1);
    for (int i = 0; i < ns; i++) {
      const int v = sel[i];
      nbv.off[n0 + i + 1] = nbv.off[n0 + i] + m.v2v.size(v);
      els.off[n0 + i + 1] = els.off[n0 + i] + m.v2el.size(v);
      facets.off[n0 + i + 1] = facets.off[n0 + i] + nfacets_of(v);
    }
    nbv.val.resize(nbv.off[n0 + ns]); nbtime.resize(nbv.off[n0 + ns]);
    els.val.resize(els.off[n0 + ns]); facets.val.resize(facets.off[n0 + ns]);
#pragma omp parallel for schedule(static)
    for (int i = 0; i < ns; i++) {
      const int v = sel[i], t = n0 + i;
      vertex[t] = v;
      level[t] = L;
      tbot[t] = tau[v];
      const double new_ttop = tau[v] + ktilde[v];
      ttop[t] = (dt - new_ttop > num_tol) ? new_ttop : dt;
      for (int k = 0; k < m.v2v.size(v); k++) {   // neighbours are not in the set: their heights are those before the phase
        nbv.val[nbv.off[t] + k] = m.v2v.row(v)[k];
        nbtime[nbv.off[t] + k] = tau[m.v2v.row(v)[k]];
      }
      for (int k = 0; k < m.v2el.size(v); k++) els.val[els.off[t] + k] = m.v2el.row(v)[k];
      int* fo = &facets.val[facets.off[t]];
      if (D == 1) fo[0] = v;
      else if (D == 2) std::copy(m.v2e.row(v), m.v2e.row(v) + m.v2e.size(v), fo);
      else std::copy(v2f.row(v), v2f.row(v) + v2f.size(v), fo);
    }
    // ---- state: dependencies (in tent and neighbour order, as the serial loop appends them), levels, front
    nbs.clear();
    for (int i = 0; i < ns; i++) {
      const int v = sel[i], t = n0 + i;
      for (int k = 0; k < m.v2v.size(v); k++) {
        const int nb = m.v2v.row(v)[k];
        if (vlevel[nb] < L + 1) vlevel[nb] = L + 1;
        if (latest_tent[nb] != -1) { dep_src.push_back(latest_tent[nb]); dep_dst.push_back(t); }
        if (!touched[nb]) { touched[nb] = 1; nbs.push_back(nb); }
      }
    }
    for (int i = 0; i < ns; i++) {
      const int v = sel[i], t = n0 + i;
      latest_tent[v] = t;
      vlevel[v] = L + 1;
      if (!(dt - (tau[v] + ktilde[v]) > num_tol)) complete[v] = 1;
      tau[v] = ttop[t];
      ktilde[v] = 0;
    }
    const int nn = (int)nbs.size();
#pragma omp parallel for schedule(static)
    for (int i = 0; i < nn; i++) {
      const int nb = nbs[i];
      if (!complete[nb]) ktilde[nb] = P.GetPoleHeight(nb, tau);
    }
    for (int nb : nbs) touched[nb] = 0;
    ntents = n0 + ns;
    slab_complete = true;
    for (int i = 0; i < nv; i++)
      if (!complete[i]) { slab_complete = false; break; }
  }
  if (!slab_complete) std::printf("Error: the algorithm could not pitch the whole slab\n");
  finalize_slab(*this, P, dep_src, dep_dst);
  pitched = slab_complete;
  return pitched;
}

// A slab from tents pitched elsewhere -- the fields python_tents.cpp:119-128 exposes of every Tent of a real
// TentPitchedSlab (vertex, tbot, ttop, nbv, nbtime, els, level, internal_facets), flattened. The tents are taken as they
// are (ids, order inside the lists, times); what the solver needs beyond them is derived here exactly as after pitching:
// dependency table (tents.cpp:206, 313-321), top-front slopes, level sets. The lists are checked for consistency with the
// mesh and for the property the level-parallel schedule rests on: tents that share an element have increasing levels
// in tent order.
void TentSlab::ImportTents(int n, double adt, const int* vertex_, const int* level_, const double* tbot_,
                           const double* ttop_, const int* nb_off, const int* nbv_, const double* nbtime_,
                           const int* el_off, const int* els_, const int* facet_off, const int* facets_) {
  if (!mesh) throw std::runtime_error("TentSlab: no mesh");
  const SimplexMesh& m = *mesh;
  const int D = m.D, nv = m.nv;
  auto bad = [](const std::string& what) { throw std::invalid_argument("import tents: " + what); };
  if (n < 1 || !(adt > 0)) bad("need at least one tent and a positive slab height");
  if (!vertex_ || !tbot_ || !ttop_ || !nb_off || !nbv_ || !nbtime_ || !el_off || !els_) bad("null array");
  if ((facet_off == nullptr) != (facets_ == nullptr)) bad("facet_off and facets go together");
  if (nb_off[0] != 0 || el_off[0] != 0 || (facet_off && facet_off[0] != 0)) bad("offset arrays must start at 0");
  const double tol = 1e-12 * adt;
  std::vector<char> mark(nv, 0);
  for (int t = 0; t < n; t++) {
    const int v = vertex_[t];
    if (v < 0 || v >= nv) bad("tent vertex out of range");
    if (nb_off[t + 1] < nb_off[t] || el_off[t + 1] <= el_off[t] || (facet_off && facet_off[t + 1] < facet_off[t]))
      bad("offset arrays must be non-decreasing (and every tent needs an element)");
    if (!(ttop_[t] > tbot_[t]) || tbot_[t] < -tol || ttop_[t] > adt + tol) bad("tent times outside [0, dt] or ttop <= tbot");
    mark[v] = 1;
    for (int q = nb_off[t]; q < nb_off[t + 1]; q++) {
      if (nbv_[q] < 0 || nbv_[q] >= nv) bad("neighbour vertex out of range");
      if (!(nbtime_[q] >= tbot_[t] - tol) && !(nbtime_[q] <= ttop_[t] + tol)) bad("neighbour time is not a number");
      mark[nbv_[q]] = 1;
    }
    for (int j = el_off[t]; j < el_off[t + 1]; j++) {
      const int el = els_[j];
      if (el < 0 || el >= m.nel) bad("element out of range");
      const int* ev = m.elverts(el);
      bool has = false;
      for (int k = 0; k <= D; k++) {
        has = has || ev[k] == v;
        if (!mark[ev[k]]) bad("an element of a tent has a vertex that is neither the tent vertex nor one of its neighbours");
      }
      if (!has) bad("an element of a tent does not contain the tent vertex");
    }
    if (facet_off)
      for (int j = facet_off[t]; j < facet_off[t + 1]; j++)
        if (facets_[j] < 0 || facets_[j] >= (D == 1 ? nv : m.nfacets)) bad("facet out of range");
    mark[v] = 0;
    for (int q = nb_off[t]; q < nb_off[t + 1]; q++) mark[nbv_[q]] = 0;
  }

  // levels and dependencies as the greedy loop assigns them (tents.cpp:199-217), replayed over the given tent order
  std::vector<int> vertices_level(nv, 0), latest_tent(nv, -1), dep_src, dep_dst, lev_new(n, 0);
  for (int t = 0; t < n; t++) {
    const int v = vertex_[t];
    const int lev = level_ ? level_[t] : vertices_level[v];
    if (lev < 0) bad("negative level");
    lev_new[t] = lev;
    for (int q = nb_off[t]; q < nb_off[t + 1]; q++) {
      const int nb = nbv_[q];
      vertices_level[nb] = std::max(vertices_level[nb], lev + 1);
      if (latest_tent[nb] != -1) { dep_src.push_back(latest_tent[nb]); dep_dst.push_back(t); }
    }
    latest_tent[v] = t;
    vertices_level[v] = std::max(vertices_level[v], lev) + 1;
  }
  {
    std::vector<int> last(m.nel, -1);
    for (int t = 0; t < n; t++)
      for (int j = el_off[t]; j < el_off[t + 1]; j++) {
        int& l = last[els_[j]];
        if (l >= lev_new[t]) bad("two tents that share an element do not have increasing levels in tent order");
        l = lev_new[t];
      }
  }

  // everything checked: replace the slab's tents
  generation++;
  dt = adt;
  pitched = false;
  ntents = n;
  vertex.assign(vertex_, vertex_ + n);
  level.swap(lev_new);
  tbot.assign(tbot_, tbot_ + n);
  ttop.assign(ttop_, ttop_ + n);
  nbv.off.assign(nb_off, nb_off + n + 1);
  nbv.val.assign(nbv_, nbv_ + nb_off[n]);
  nbtime.assign(nbtime_, nbtime_ + nb_off[n]);
  els.off.assign(el_off, el_off + n + 1);
  els.val.assign(els_, els_ + el_off[n]);
  nlayers = 0;
  for (int t = 0; t < n; t++) nlayers = std::max(nlayers, level[t]);

  if (facet_off) {
    facets.off.assign(facet_off, facet_off + n + 1);
    facets.val.assign(facets_, facets_ + facet_off[n]);
  } else {
    // internal facets of the vertex patch as the pitcher lists them (tents.cpp:219-247)
    facets.off.assign(1, 0);
    facets.val.clear();
    for (int t = 0; t < n; t++) {
      const int vi = vertex[t];
      if (D == 1) facets.val.push_back(vi);
      else if (D == 2)
        for (int i = 0; i < m.v2e.size(vi); i++) facets.val.push_back(m.v2e.row(vi)[i]);
      else {
        const size_t b = facets.val.size();
        for (int i = 0; i < m.v2el.size(vi); i++) {
          const int el = m.v2el.row(vi)[i];
          for (int k = 0; k < 4; k++) {
            const int f = m.el2facet[(size_t)el * 4 + k];
            const int* fv = &m.facet2v[(size_t)f * 3];
            if ((fv[0] == vi || fv[1] == vi || fv[2] == vi) && std::find(facets.val.begin() + b, facets.val.end(), f) == facets.val.end())
              facets.val.push_back(f);
          }
        }
      }
      facets.off.push_back((int)facets.val.size());
    }
  }
  Pitcher P(m, method);
  P.CacheGradients();
  finalize_slab(*this, P, dep_src, dep_dst);
  pitched = true;
}

}  // namespace ngst
