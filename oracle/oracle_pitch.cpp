// TEST INFRASTRUCTURE ONLY (see oracle_common.hpp). Literal restatement of ngstents' tent pitching:
//   TentPitchedSlab::PitchTents<DIM>            tents.cpp:78-359
//   TentSlabPitcher::GetReadyVertices           tents.cpp:385-418
//   ::ComputeVerticesReferenceHeight            tents.cpp:420-429
//   ::PickNextVertexForPitching                 tents.cpp:431-443
//   ::UpdateNeighbours                          tents.cpp:445-474
//   ::InitializeMeshData                        tents.cpp:476-563
//   VolumeGradientPitcher::GetPoleHeight        tents.cpp:627-717,  ::CalcLocalCTau 719-794
//   EdgeGradientPitcher::GetPoleHeight          tents.cpp:796-813,  ::CalcLocalCTau 815-978
// (all paths relative to external_dependencies/ngstents/src/). The ready list is an ngcore Array in the
// reference; its semantics are restated from ngcore's array.hpp (not in the reference tree):
// DeleteElement(i) moves the LAST element into slot i, RemoveElement(i) shifts the tail forward,
// Pos() is a first-match linear search. Periodic identifications (vmap) are the identity here: the solver
// path never applies them.
// Compile with -ffp-contract=off: the schedule depends on the bits of tau / ktilde.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <limits>
#include <stdexcept>

#include "oracle_common.hpp"

namespace oracle {

void OMesh::derive() {
  const int d1 = D + 1;
  v2el.assign(nv, {});
  for (int e = 0; e < nel; e++)
    for (int k = 0; k < d1; k++) v2el[elverts(e)[k]].push_back(e);
  v2v.assign(nv, {});
  v2e.assign(nv, {});
  for (int e = 0; e < nedges; e++) {  // tents.cpp:529-539
    const int v1 = edge2v[2 * e], v2 = edge2v[2 * e + 1];
    v2v[v1].push_back(v2); v2e[v1].push_back(e);
    v2v[v2].push_back(v1); v2e[v2].push_back(e);
  }
  edge2el.assign(nedges, {});
  for (int e = 0; e < nel; e++)
    for (int k = 0; k < nepe(); k++) edge2el[el2edge[(size_t)e * nepe() + k]].push_back(e);
}

namespace {

// ngcore Array<int> semantics used by the pitcher
struct NgArray {
  std::vector<int> d;
  size_t Size() const { return d.size(); }
  void Append(int v) { d.push_back(v); }
  void DeleteElement(size_t i) { d[i] = d.back(); d.pop_back(); }
  void RemoveElement(size_t i) { d.erase(d.begin() + i); }
  static constexpr size_t ILLEGAL_POSITION = (size_t)-1;
  size_t Pos(int v) const {
    for (size_t i = 0; i < d.size(); i++)
      if (d[i] == v) return i;
    return ILLEGAL_POSITION;
  }
  int operator[](size_t i) const { return d[i]; }
};

// mapped gradients of ScalarFE<ET,1> (barycentric functions), rows = vertices
void CalcMappedDShape(const OMesh& ma, int el, double* g) {
  const int D = ma.D;
  const int* v = ma.elverts(el);
  double J[9], inv[9];
  for (int i = 0; i < D; i++)
    for (int d = 0; d < D; d++) J[d * D + i] = ma.point(v[i])[d] - ma.point(v[D])[d];
  if (D == 1) inv[0] = 1.0 / J[0];
  else if (D == 2) {
    const double det = J[0] * J[3] - J[1] * J[2];
    inv[0] = J[3] / det; inv[1] = -J[1] / det; inv[2] = -J[2] / det; inv[3] = J[0] / det;
  } else {
    const double c00 = J[4] * J[8] - J[5] * J[7], c01 = J[5] * J[6] - J[3] * J[8], c02 = J[3] * J[7] - J[4] * J[6];
    const double det = J[0] * c00 + J[1] * c01 + J[2] * c02;
    inv[0] = c00 / det; inv[1] = (J[2] * J[7] - J[1] * J[8]) / det; inv[2] = (J[1] * J[5] - J[2] * J[4]) / det;
    inv[3] = c01 / det; inv[4] = (J[0] * J[8] - J[2] * J[6]) / det; inv[5] = (J[2] * J[3] - J[0] * J[5]) / det;
    inv[6] = c02 / det; inv[7] = (J[1] * J[6] - J[0] * J[7]) / det; inv[8] = (J[0] * J[4] - J[1] * J[3]) / det;
  }
  for (int d = 0; d < D; d++) g[D * D + d] = 0;
  for (int i = 0; i < D; i++)
    for (int d = 0; d < D; d++) { g[i * D + d] = inv[i * D + d]; g[D * D + d] -= inv[i * D + d]; }
}
double L2Norm(const double* a, int n) {
  double s = 0;
  for (int i = 0; i < n; i++) s += a[i] * a[i];
  return std::sqrt(s);
}
int PosInEl(const OMesh& ma, int el, int v) {
  for (int k = 0; k <= ma.D; k++)
    if (ma.elverts(el)[k] == v) return k;
  return -1;
}

struct TentSlabPitcher {
  const OMesh& ma;
  int method;  // 1 vol, 2 edge
  std::vector<double> cmax, vertex_refdt, edge_len;
  std::vector<std::vector<double>> local_ctau_table;
  bool have_table = false;
  double global_ctau = 1.0;
  TentSlabPitcher(const OMesh& m, int meth) : ma(m), method(meth), vertex_refdt(m.nv), edge_len(m.nedges) {
    cmax.assign(method == 2 ? ma.nedges : ma.nel, -1.0);
  }
  double local_ctau(int v, int i) const { return have_table ? local_ctau_table[v][i] : 1.0; }

  void InitializeMeshData(const std::vector<double>& wavespeed, bool calc_local_ct, double global_ct) {
    global_ctau = global_ct;
    std::vector<char> fine_edges(ma.nedges, 0);
    for (int el = 0; el < ma.nel; el++) {
      const double wvspd = wavespeed[el];
      if (method == 1) cmax[el] = wvspd;
      for (int k = 0; k < ma.nepe(); k++) {
        const int e = ma.el2edge[(size_t)el * ma.nepe() + k];
        const int v1 = ma.edge2v[2 * e], v2 = ma.edge2v[2 * e + 1];
        if (!fine_edges[e]) {
          fine_edges[e] = 1;
          double s = 0;
          for (int d = 0; d < ma.D; d++) s += (ma.point(v1)[d] - ma.point(v2)[d]) * (ma.point(v1)[d] - ma.point(v2)[d]);
          edge_len[e] = std::sqrt(s);
        }
        if (method == 2) cmax[e] = std::max(cmax[e], wvspd);
      }
    }
    if (calc_local_ct && ma.D > 1) {
      local_ctau_table = method == 2 ? CalcLocalCTauEdge() : CalcLocalCTauVol();
      have_table = true;
    }
  }

  std::vector<std::vector<double>> CalcLocalCTauEdge() const {
    const int D = ma.D;
    std::vector<std::vector<double>> table(ma.nv);
    std::vector<double> gradphi((D + 1) * D);
    for (int vi = 0; vi < ma.nv; vi++)
      for (int edge : ma.v2e[vi]) {
        double val = std::numeric_limits<double>::max();
        const int v1 = ma.edge2v[2 * edge], v2 = ma.edge2v[2 * edge + 1];
        for (int iel : ma.edge2el[edge]) {
          CalcMappedDShape(ma, iel, gradphi.data());
          const int v1_local = PosInEl(ma, iel, v1), v2_local = PosInEl(ma, iel, v2);
          double one_over_max_grad;
          if (D == 2)
            one_over_max_grad = 1.0 / std::max(L2Norm(&gradphi[v1_local * D], D), L2Norm(&gradphi[v2_local * D], D));
          else {
            double v = 1;
            for (int face_local = 0; face_local < 4; face_local++) {
              // faces of the element containing the edge = faces opposite the other two vertices; the mapped
              // reference normal det*J^{-T}*n_ref is parallel to the gradient of the opposite hat function
              if (face_local == v1_local || face_local == v2_local) continue;
              double normal[3];
              const double len_normal = L2Norm(&gradphi[face_local * 3], 3);
              for (int d = 0; d < 3; d++) normal[d] = gradphi[face_local * 3 + d] / len_normal;
              auto calc_grad_proj = [&](int vertex) {
                const double norm_grad = L2Norm(&gradphi[vertex * 3], 3);
                double mg[3];
                for (int d = 0; d < 3; d++) mg[d] = gradphi[vertex * 3 + d] / norm_grad;
                const double cr[3] = {normal[1] * mg[2] - normal[2] * mg[1], normal[2] * mg[0] - normal[0] * mg[2],
                                      normal[0] * mg[1] - normal[1] * mg[0]};
                return L2Norm(cr, 3) / norm_grad;
              };
              const double val_face = std::min(calc_grad_proj(v1_local), calc_grad_proj(v2_local));
              v = std::min(v, val_face);
            }
            one_over_max_grad = v;
          }
          const double projGrad = one_over_max_grad / edge_len[edge];
          val = std::min(val, projGrad);
        }
        table[vi].push_back(val);
      }
    return table;
  }

  std::vector<std::vector<double>> CalcLocalCTauVol() const {
    const int D = ma.D;
    std::vector<std::vector<double>> table(ma.nv);
    auto len = [&](int a, int b) {
      double s = 0;
      for (int d = 0; d < D; d++) s += (ma.point(a)[d] - ma.point(b)[d]) * (ma.point(a)[d] - ma.point(b)[d]);
      return std::sqrt(s);
    };
    for (int vi = 0; vi < ma.nv; vi++)
      for (int el_num : ma.v2el[vi]) {
        const int* ev = ma.elverts(el_num);
        double val = 1.0;
        const int nfaces = D == 3 ? 4 : 1;
        for (int face = 0; face < nfaces; face++) {
          std::vector<int> fv;
          for (int j = 0; j <= D; j++)
            if (D == 2 || j != face) fv.push_back(ev[j]);
          if (std::find(fv.begin(), fv.end(), vi) == fv.end()) continue;
          double opposite_edge = -1, max_edge = 0;
          for (size_t a = 0; a < fv.size(); a++)
            for (size_t b = a + 1; b < fv.size(); b++) {
              if (fv[a] != vi && fv[b] != vi) opposite_edge = len(fv[a], fv[b]);
              else max_edge = std::max(len(fv[a], fv[b]), max_edge);
            }
          val = std::min(val, opposite_edge / max_edge);
        }
        val = std::min(val, 1.0);
        table[vi].push_back(val);
      }
    return table;
  }

  double GetPoleHeight(int vi, const std::vector<double>& tau) const {
    if (method == 2) {  // EdgeGradientPitcher, tents.cpp:796-813
      double kt = std::numeric_limits<double>::max();
      const std::vector<int>& nbv = ma.v2v[vi];
      const std::vector<int>& nbe = ma.v2e[vi];
      for (size_t nb_index = 0; nb_index < nbv.size(); nb_index++) {
        const int nb = nbv[nb_index];
        const int edge = nbe[nb_index];
        const double length = edge_len[edge];
        const double c_max = cmax[edge];
        const double local_ct = local_ctau(vi, (int)nb_index);
        const double kt1 = tau[nb] - tau[vi] + global_ctau * local_ct * length / c_max;
        kt = std::min(kt, kt1);
      }
      constexpr double num_tol = std::numeric_limits<double>::epsilon();
      if (kt > num_tol) return kt * (1 - num_tol);
      return 0.0;
    }
    // VolumeGradientPitcher, tents.cpp:627-717
    const int D = ma.D;
    constexpr double init_pole_height = std::numeric_limits<double>::max();
    constexpr double num_tol = std::numeric_limits<double>::epsilon();
    double pole_height = init_pole_height;
    double gradphi[12];
    const std::vector<int>& els = ma.v2el[vi];
    for (size_t iel = 0; iel < els.size(); iel++) {
      const int el = els[iel];
      const double c_max_sq = cmax[el] * cmax[el];
      const int* v_indices = ma.elverts(el);
      const int local_vi = PosInEl(ma, el, vi);
      CalcMappedDShape(ma, el, gradphi);
      double coeff_vec[4];
      for (int k = 0; k <= D; k++) coeff_vec[k] = tau[v_indices[k]];
      coeff_vec[local_vi] = 0;
      double alpha = 0;
      for (int d = 0; d < D; d++) alpha += gradphi[local_vi * D + d] * gradphi[local_vi * D + d];
      double tau_i_grad_i[3] = {0, 0, 0};
      for (int d = 0; d < D; d++)
        for (int k = 0; k <= D; k++) tau_i_grad_i[d] += gradphi[k * D + d] * coeff_vec[k];
      double ip = 0, tt = 0;
      for (int d = 0; d < D; d++) { ip += tau_i_grad_i[d] * gradphi[local_vi * D + d]; tt += tau_i_grad_i[d] * tau_i_grad_i[d]; }
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
      kbar *= local_ctau(vi, (int)iel) * global_ctau;
      pole_height = std::min(pole_height, kbar);
    }
    if (pole_height > 0.75 * init_pole_height) return 0.0;
    return pole_height * (1 - num_tol);
  }

  bool GetReadyVertices(double& adv_factor, bool reset_adv_factor, const std::vector<double>& ktilde,
                        const std::vector<char>& complete_vertices, std::vector<char>& vertex_ready,
                        NgArray& ready_vertices) const {
    bool found = false;
    constexpr int n_attempts = 5;
    std::fill(vertex_ready.begin(), vertex_ready.end(), 0);
    const double initial_adv_factor = adv_factor;
    for (int ia = 0; ia < n_attempts; ia++) {
      for (int iv = 0; iv < ma.nv; iv++)
        if (!complete_vertices[iv])
          if (ktilde[iv] > adv_factor * vertex_refdt[iv])
            if (!vertex_ready[iv]) { ready_vertices.Append(iv); vertex_ready[iv] = 1; }
      if (ready_vertices.Size()) { found = true; break; }
      adv_factor /= 2;
    }
    if (reset_adv_factor) adv_factor = initial_adv_factor;
    else if (adv_factor < 0.05) return false;
    return found;
  }

  void UpdateNeighbours(int vi, double adv_factor, const std::vector<double>& tau,
                        const std::vector<char>& complete_vertices, std::vector<double>& ktilde,
                        std::vector<char>& vertex_ready, NgArray& ready_vertices) const {
    for (int nb : ma.v2v[vi]) {
      if (complete_vertices[nb]) continue;
      const double kt = GetPoleHeight(nb, tau);
      ktilde[nb] = kt;
      if (kt > adv_factor * vertex_refdt[nb]) {
        if (!vertex_ready[nb]) { ready_vertices.Append(nb); vertex_ready[nb] = 1; }
      } else {
        vertex_ready[nb] = 0;
        const size_t pos_nb = ready_vertices.Pos(nb);
        if (pos_nb != NgArray::ILLEGAL_POSITION) ready_vertices.RemoveElement(pos_nb);
      }
    }
  }
};

}  // namespace

// rule 0: the reference's pick rule (first vertex of minimal level in the ready list, tents.cpp:431-443).
// rule 1: the partition-independent rule of the product's level-synchronous pitcher (no reference counterpart): the ready
// vertex of minimal (level, vertex number); the ready set is recomputed from the state whenever a level is exhausted.
bool PitchTents(OSlab& slab, const OMesh& ma, int method, const std::vector<double>& cmax_el, double dt,
                bool calc_local_ct, double global_ct, int rule) {
  const int DIM = ma.D;
  slab.ma = &ma;
  slab.tents.clear();
  slab.nlayers = 0;
  slab.dt = dt;
  if (method != 1 && method != 2) return false;
  TentSlabPitcher pitcher(ma, method);
  pitcher.InitializeMeshData(cmax_el, calc_local_ct, global_ct);

  std::vector<double> tau(ma.nv, 0.0);
  std::fill(pitcher.vertex_refdt.begin(), pitcher.vertex_refdt.end(), std::numeric_limits<double>::max());
  for (int i = 0; i < ma.nv; i++) pitcher.vertex_refdt[i] = pitcher.GetPoleHeight(i, tau);
  std::vector<double> ktilde = pitcher.vertex_refdt;

  std::vector<int> latest_tent(ma.nv, -1), vertices_level(ma.nv, 0);
  double adv_factor = 0.5;
  constexpr bool reset_adv_factor = true;
  NgArray ready_vertices;
  std::vector<char> vertex_ready(ma.nv, 0), complete_vertices(ma.nv, 0);
  bool slab_complete = false;
  const double num_tol = std::numeric_limits<double>::epsilon() * dt;
  std::vector<OTent>& tents = slab.tents;

  // rule 1: the vertices of the current phase (ready at its start, minimal level), ascending, and the position in it
  std::vector<int> phase;
  size_t phase_pos = 0;
  int phase_level = 0;
  while (!slab_complete) {
    if (rule == 0) {
      const bool found_vertices =
          pitcher.GetReadyVertices(adv_factor, reset_adv_factor, ktilde, complete_vertices, vertex_ready, ready_vertices);
      if (!found_vertices) break;
    } else {
      // ready set from the state: advance larger than adv_factor * reference height, the factor halved up to 5 times
      phase.clear();
      phase_pos = 0;
      double af = 0.5;
      for (int ia = 0; ia < 5 && phase.empty(); ia++, af /= 2) {
        int minlevel = std::numeric_limits<int>::max();
        for (int iv = 0; iv < ma.nv; iv++)
          if (!complete_vertices[iv] && ktilde[iv] > af * pitcher.vertex_refdt[iv]) minlevel = std::min(minlevel, vertices_level[iv]);
        for (int iv = 0; iv < ma.nv; iv++)
          if (!complete_vertices[iv] && ktilde[iv] > af * pitcher.vertex_refdt[iv] && vertices_level[iv] == minlevel) phase.push_back(iv);
        phase_level = minlevel;
      }
      if (phase.empty()) break;
    }
    while (rule == 0 ? ready_vertices.Size() > 0 : phase_pos < phase.size()) {
      int vi;
      if (rule == 0) {
        // PickNextVertexForPitching
        int minlevel = std::numeric_limits<int>::max(), posmin = -1;
        for (size_t i = 0; i < ready_vertices.Size(); i++)
          if (vertices_level[ready_vertices[i]] < minlevel) { minlevel = vertices_level[ready_vertices[i]]; posmin = (int)i; }
        slab.nlayers = std::max(minlevel, slab.nlayers);
        vi = ready_vertices[posmin];
        ready_vertices.DeleteElement(posmin);
        vertex_ready[vi] = 0;
      } else {
        vi = phase[phase_pos++];
        if (vertices_level[vi] != phase_level) continue;  // a neighbour with a smaller number was pitched in this phase
        slab.nlayers = std::max(phase_level, slab.nlayers);
      }

      OTent tent;
      tent.vertex = vi;
      tent.tbot = tau[vi];
      const double new_ttop = tau[vi] + ktilde[vi];
      if (dt - new_ttop > num_tol) tent.ttop = new_ttop;
      else { tent.ttop = dt; complete_vertices[vi] = 1; }
      tent.level = vertices_level[vi];
      tau[vi] = tent.ttop;
      ktilde[vi] = 0;
      for (int nb : ma.v2v[vi]) {
        tent.nbv.push_back(nb);
        tent.nbtime.push_back(tau[nb]);
        if (vertices_level[nb] < tent.level + 1) vertices_level[nb] = tent.level + 1;
        if (latest_tent[nb] != -1) tents[latest_tent[nb]].dependent_tents.push_back((int)tents.size());
      }
      latest_tent[vi] = (int)tents.size();
      vertices_level[vi]++;

      if (DIM == 1) tent.internal_facets.push_back(vi);
      else if (DIM == 2) for (int e : ma.v2e[vi]) tent.internal_facets.push_back(e);
      else
        for (int elnr : ma.v2el[vi])
          for (int k = 0; k < 4; k++) {
            const int f = ma.el2facet[(size_t)elnr * 4 + k];
            for (int j = 0; j < 3; j++)
              if (ma.facet2v[(size_t)f * 3 + j] == vi &&
                  std::find(tent.internal_facets.begin(), tent.internal_facets.end(), f) == tent.internal_facets.end()) {
                tent.internal_facets.push_back(f);
                break;
              }
          }
      tent.els = ma.v2el[vi];
      if (rule == 0) {
        pitcher.UpdateNeighbours(vi, adv_factor, tau, complete_vertices, ktilde, vertex_ready, ready_vertices);
      } else {
        for (int nb : ma.v2v[vi])
          if (!complete_vertices[nb]) ktilde[nb] = pitcher.GetPoleHeight(nb, tau);
      }
      tents.push_back(std::move(tent));
    }
    slab_complete = true;
    for (int i = 0; i < ma.nv; i++)
      if (!complete_vertices[i]) { slab_complete = false; break; }
  }

  slab.tent_dependency.assign(tents.size(), {});
  for (size_t i = 0; i < tents.size(); i++) slab.tent_dependency[i] = tents[i].dependent_tents;

  // slope of the top front, tents.cpp:323-356
  for (OTent& tent : tents) {
    double dshape_nodal[12];
    for (int el : tent.els) {
      const int* vnums = ma.elverts(el);
      double coef_top[4];
      for (int k = 0; k <= DIM; k++) {
        auto it = std::find(tent.nbv.begin(), tent.nbv.end(), vnums[k]);
        coef_top[k] = it != tent.nbv.end() ? tent.nbtime[it - tent.nbv.begin()] : tent.ttop;
      }
      CalcMappedDShape(ma, el, dshape_nodal);
      double gradphi_top[3] = {0, 0, 0};
      for (int d = 0; d < DIM; d++)
        for (int k = 0; k <= DIM; k++) gradphi_top[d] += dshape_nodal[k * DIM + d] * coef_top[k];
      const double norm = L2Norm(gradphi_top, DIM);
      if (norm > tent.maxslope) tent.maxslope = norm;
    }
  }
  slab.has_been_pitched = slab_complete;
  return slab_complete;
}

}  // namespace oracle
