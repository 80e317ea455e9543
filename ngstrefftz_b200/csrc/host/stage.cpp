#include "stage.hpp"

#include <algorithm>
#include <cmath>
#include <stdexcept>

namespace ngst {

namespace {
inline double sgn_nozero(double v) { return (double)((0.0 <= v) - (v < 0.0)); }
inline double dist(const double* a, const double* b, int n) {
  double s = 0;
  for (int i = 0; i < n; i++) s += (a[i] - b[i]) * (a[i] - b[i]);
  return std::sqrt(s);
}
inline double det3(const double m[3][3]) {
  return m[0][0] * (m[1][1] * m[2][2] - m[1][2] * m[2][1]) - m[0][1] * (m[1][0] * m[2][2] - m[1][2] * m[2][0]) +
         m[0][2] * (m[1][0] * m[2][1] - m[1][1] * m[2][0]);
}
}  // namespace

// twavetents.cpp:633-680. Returns D! times the D-dimensional measure of the face.
double tent_face_area(int D, const double* ve) {
  const int d1 = D + 1;
  auto col = [&](int k) { return ve + k * d1; };
  if (D == 1) return dist(col(0), col(1), 2);
  if (D == 2) {
    double ba[3], baa[3];
    for (int i = 0; i < 3; i++) { ba[i] = col(0)[i] - col(1)[i]; baa[i] = col(0)[i] - col(2)[i]; }
    double c[3] = {ba[1] * baa[2] - ba[2] * baa[1], ba[2] * baa[0] - ba[0] * baa[2], ba[0] * baa[1] - ba[1] * baa[0]};
    return std::sqrt(c[0] * c[0] + c[1] * c[1] + c[2] * c[2]) / 2.0 * 2.0;
  }
  const double U = dist(col(0), col(1), 4), V = dist(col(1), col(2), 4), W = dist(col(2), col(0), 4);
  const double u = dist(col(3), col(2), 4), v = dist(col(3), col(0), 4), w = dist(col(3), col(1), 4);
  const double X = (w - U + v) * (U + v + w), x = (U - v + w) * (v - w + U);
  const double Y = (u - V + w) * (V + w + u), y = (V - w + u) * (w - u + V);
  const double Z = (v - W + u) * (W + u + v), z = (W - u + v) * (u - v + W);
  const double a = std::sqrt(x * Y * Z), b = std::sqrt(y * Z * X), c = std::sqrt(z * X * Y), d = std::sqrt(x * y * z);
  return std::sqrt((-a + b + c + d) * (a - b + c + d) * (a + b - c + d) * (a + b + c - d)) / (192.0 * u * v * w) * 6.0;
}

// twavetents.cpp:682-732
void tent_face_normal(int D, const double* ve, int top, double* n) {
  const int d1 = D + 1;
  auto at = [&](int row, int colk) { return ve[colk * d1 + row]; };
  if (D == 1) {
    n[0] = at(1, 1) - at(1, 0);
    n[1] = at(0, 0) - at(0, 1);
  } else if (D == 2) {
    double ba[3], baa[3];
    for (int i = 0; i < 3; i++) { ba[i] = at(i, 0) - at(i, 1); baa[i] = at(i, 0) - at(i, 2); }
    n[0] = ba[1] * baa[2] - ba[2] * baa[1];
    n[1] = ba[2] * baa[0] - ba[0] * baa[2];
    n[2] = ba[0] * baa[1] - ba[1] * baa[0];
  } else {
    double w[4][4];  // w[row][col], col 1..3 = V_0 - V_col
    for (int r = 0; r < 4; r++)
      for (int c = 1; c < 4; c++) w[r][c] = at(r, 0) - at(r, c);
    for (int i = 0; i < 4; i++) {
      double pS[3][3];
      for (int k = 0, c = 0; k < 4; k++) {
        if (k == i) continue;
        for (int j = 0; j < 3; j++) pS[c][j] = w[k][j + 1];
        c++;
      }
      n[i] = (i % 2 == 0) ? det3(pS) : -det3(pS);
    }
  }
  if (top == 1) { double s = sgn_nozero(n[D]); for (int i = 0; i < d1; i++) n[i] *= s; }
  else if (top == -1) { double s = -sgn_nozero(n[D]); for (int i = 0; i < d1; i++) n[i] *= s; }
  else n[D] = 0;
  double l = 0;
  for (int i = 0; i < d1; i++) l += n[i] * n[i];
  l = std::sqrt(l);
  for (int i = 0; i < d1; i++) n[i] /= l;
}

StagedTents stage_tents(const SimplexMesh& m, const TentSlab& slab, const std::vector<double>& c_el,
                        const std::vector<double>& c_vertex, const SimplexRule& rule, bool qt) {
  if ((int)c_el.size() != m.nel) throw std::runtime_error("stage_tents: c_el size mismatch");
  if ((int)c_vertex.size() != m.nv) throw std::runtime_error("stage_tents: c_vertex size mismatch");
  const int D = m.D, d1 = D + 1, nt = slab.ntents, Q = rule.nip;
  StagedTents st;
  st.D = D; st.d1 = d1; st.ntents = nt;
  st.center.resize((size_t)nt * d1);
  st.scale.resize(nt);
  st.ndom.resize(nt);
  st.te_off.assign(nt + 1, 0);
  st.tf_off.assign(nt + 1, 0);

  // ---- pass 1: macro groups and record counts
  std::vector<int> dom_of(slab.els.val.size());
  for (int t = 0; t < nt; t++) {
    const int e0 = slab.els.off[t], ne = slab.els.size(t);
    // MakeMacroEl, twavetents.cpp:916-933 (exact comparison of doubles)
    int nd = 0;
    for (int i = 0; i < ne; i++) {
      int j = 0;
      while (c_el[slab.els.val[e0 + i]] != c_el[slab.els.val[e0 + j]]) j++;
      dom_of[e0 + i] = (j == i) ? nd++ : dom_of[e0 + j];
    }
    if (qt) {
      nd = 1;
      for (int i = 0; i < ne; i++) dom_of[e0 + i] = 0;
    }
    st.ndom[t] = nd;
    st.max_ndom = std::max(st.max_ndom, nd);
    st.max_els = std::max(st.max_els, ne);
    st.te_off[t + 1] = st.te_off[t] + ne;
    int nf = 0;
    for (int q = slab.facets.off[t]; q < slab.facets.off[t + 1]; q++) {
      const int f = slab.facets.val[q];
      const int ea = m.facet2el[2 * f], eb = m.facet2el[2 * f + 1];
      if (eb < 0) { if (m.facet2se[f] >= 0) nf++; }
      else if (nd > 1) {
        int da = -1, db = -1;
        for (int i = 0; i < ne; i++) {
          if (slab.els.val[e0 + i] == ea) da = dom_of[e0 + i];
          if (slab.els.val[e0 + i] == eb) db = dom_of[e0 + i];
        }
        if (da != db) nf++;  // unordered_map::operator[] yields 0 for elements outside the patch; both lie in it
      }
    }
    st.tf_off[t + 1] = st.tf_off[t] + nf;
  }
  st.nte = st.te_off[nt];
  st.ntf = st.tf_off[nt];
  st.te_el.resize(st.nte); st.te_dom.resize(st.nte);
  st.te_tbot.resize((size_t)st.nte * d1); st.te_ttop.resize((size_t)st.nte * d1);
  st.te_geo.resize((size_t)st.nte * 2 * (d1 + 1));
  st.tf_kind.resize(st.ntf); st.tf_dom.resize((size_t)st.ntf * 2); st.tf_c.resize((size_t)st.ntf * 2);
  st.tf_verts.resize((size_t)st.ntf * d1 * d1); st.tf_n.resize((size_t)st.ntf * D); st.tf_area.resize(st.ntf);
  st.tf_data.assign(st.ntf, -1); st.tf_point.assign(st.ntf, -1);

  // ---- pass 2: geometry (independent per tent)
#pragma omp parallel for schedule(dynamic, 256)
  for (int t = 0; t < nt; t++) {
    const int vtx = slab.vertex[t];
    const double tb = slab.tbot[t], tt = slab.ttop[t];
    const int nb0 = slab.nbv.off[t], nnb = slab.nbv.size(t);
    const int* nbv = slab.nbv.val.data() + nb0;
    const double* nbt = slab.nbtime.data() + nb0;
    auto vtime = [&](int v, double own) {
      if (v == vtx) return own;
      for (int k = 0; k < nnb; k++)
        if (nbv[k] == v) return nbt[k];
      throw std::runtime_error("stage_tents: element vertex is not a tent neighbour");
    };
    for (int d = 0; d < D; d++) st.center[(size_t)t * d1 + d] = m.point(vtx)[d];
    st.center[(size_t)t * d1 + D] = (tt - tb) / 2 + tb;

    // TentAdiam, twavetents.cpp:861-904
    {
      double c1 = c_el[0];
      double adiam = c1 * tt - c1 * tb;
      for (int k = 0; k < nnb; k++) {
        const double* v1 = m.point(vtx);
        const double* v2 = m.point(nbv[k]);
        c1 = c_vertex[vtx];
        const double c2 = c_vertex[nbv[k]];
        double l2 = 0;
        for (int d = 0; d < D; d++) l2 += (v1[d] - v2[d]) * (v1[d] - v2[d]);
        adiam = std::max(adiam, std::sqrt(l2 + std::pow(c1 * tt - c2 * nbt[k], 2)));
        adiam = std::max(adiam, std::sqrt(l2 + std::pow(c1 * tb - c2 * nbt[k], 2)));
        for (int j = 0; j < nnb; j++) {
          v1 = m.point(nbv[j]);
          c1 = c_vertex[nbv[j]];
          double l2j = 0;
          for (int d = 0; d < D; d++) l2j += (v1[d] - v2[d]) * (v1[d] - v2[d]);
          adiam = std::max(adiam, std::sqrt(l2j + std::pow(c1 * nbt[j] - c2 * nbt[k], 2)));
        }
      }
      st.scale[t] = 1.0 / adiam;
    }
    if (qt) {  // TentXdiam, twavetents.cpp:1125-1144
      double diam = 0;
      auto dist = [&](const double* p, const double* q) {
        double l2 = 0;
        for (int d = 0; d < D; d++) l2 += (p[d] - q[d]) * (p[d] - q[d]);
        return std::sqrt(l2);
      };
      for (int k = 0; k < nnb; k++) {
        const double* v1 = m.point(nbv[k]);
        diam = std::max(diam, dist(v1, m.point(vtx)));
        for (int j = k; j < nnb; j++) diam = std::max(diam, dist(v1, m.point(nbv[j])));
      }
      st.scale[t] = 1.0 / diam;
    }

    const int e0 = slab.els.off[t], ne = slab.els.size(t);
    const int nd = st.ndom[t];
    auto group_of = [&](int el) {
      for (int i = 0; i < ne; i++)
        if (slab.els.val[e0 + i] == el) return dom_of[e0 + i];
      return 0;
    };

    // space-like faces of every patch element
    for (int i = 0; i < ne; i++) {
      const int te = st.te_off[t] + i;
      const int el = slab.els.val[e0 + i];
      st.te_el[te] = el;
      st.te_dom[te] = nd > 1 ? dom_of[e0 + i] : 0;
      const int* ev = m.elverts(el);
      double vb[16], vt[16];
      for (int k = 0; k < d1; k++) {
        for (int d = 0; d < D; d++) vb[k * d1 + d] = vt[k * d1 + d] = m.point(ev[k])[d];
        vb[k * d1 + D] = st.te_tbot[(size_t)te * d1 + k] = vtime(ev[k], tb);
        vt[k * d1 + D] = st.te_ttop[(size_t)te * d1 + k] = vtime(ev[k], tt);
      }
      double* g = &st.te_geo[(size_t)te * 2 * (d1 + 1)];
      tent_face_normal(D, vb, -1, g);
      g[d1] = tent_face_area(D, vb);
      tent_face_normal(D, vt, 1, g + d1 + 1);
      g[2 * d1 + 1] = tent_face_area(D, vt);
    }

    // time-like faces: boundary facets and material interfaces (twavetents.cpp:110-139)
    int tf = st.tf_off[t];
    for (int q = slab.facets.off[t]; q < slab.facets.off[t + 1]; q++) {
      const int f = slab.facets.val[q];
      const int ea = m.facet2el[2 * f], eb = m.facet2el[2 * f + 1];
      const int* fverts = nullptr;
      int kind;
      if (eb < 0) {
        const int se = m.facet2se[f];
        if (se < 0) continue;
        kind = m.se_kind[se] == 1 ? kNeumann : kDirichlet;
        fverts = &m.se2v[(size_t)se * D];
      } else {
        if (nd <= 1 || group_of(ea) == group_of(eb)) continue;
        kind = kInterface;
        fverts = &m.facet2v[(size_t)f * D];
      }
      double* V = &st.tf_verts[(size_t)tf * d1 * d1];
      for (int d = 0; d < D; d++) V[d] = m.point(vtx)[d];
      V[D] = tb;
      for (int n = 0; n < D; n++) {
        for (int d = 0; d < D; d++) V[(n + 1) * d1 + d] = m.point(fverts[n])[d];
        V[(n + 1) * d1 + D] = vtime(fverts[n], tt);
      }
      st.tf_kind[tf] = kind;
      st.tf_area[tf] = tent_face_area(D, V);
      double* nn = &st.tf_n[(size_t)tf * D];
      // the adjacent element's vertex opposite to the facet fixes "outward"
      const int* ev = m.elverts(ea);
      int opp = -1;
      for (int k = 0; k < d1 && opp < 0; k++) {
        bool on = false;
        for (int n = 0; n < D; n++) on = on || (ev[k] == fverts[n]);
        if (!on) opp = ev[k];
      }
      if (kind != kInterface) {
        double nf[4];
        if (D == 1) nf[0] = sgn_nozero(m.point(vtx)[0] - m.point(nbv[0])[0]);  // twavetents.cpp:350-354
        else {
          tent_face_normal(D, V, 0, nf);  // twavetents.cpp:345-348
          for (int d = 0; d < D; d++) nf[d] = -nf[d];
        }
        double s = 0;
        for (int d = 0; d < D; d++) s += nf[d] * (m.point(opp)[d] - m.point(fverts[0])[d]);
        if (s > 0) for (int d = 0; d < D; d++) nf[d] = -nf[d];  // surface element oriented the other way round
        for (int d = 0; d < D; d++) nn[d] = nf[d];
        const int g = nd > 1 ? group_of(ea) : 0;
        st.tf_dom[2 * tf] = st.tf_dom[2 * tf + 1] = g;
        st.tf_c[2 * tf] = st.tf_c[2 * tf + 1] = c_el[ea];
      } else {
        // outward unit normal of elnums[0] on the facet (twavetents.cpp:469-474)
        const int* eva = m.elverts(ea);
        double x[3] = {0, 0, 0};
        if (D == 1) x[0] = sgn_nozero(m.point(fverts[0])[0] - m.point(opp)[0]);
        else if (D == 2) {
          const double* a = m.point(fverts[0]);
          const double* b = m.point(fverts[1]);
          x[0] = b[1] - a[1]; x[1] = -(b[0] - a[0]);
        } else {
          const double* a = m.point(fverts[0]);
          const double* b = m.point(fverts[1]);
          const double* c = m.point(fverts[2]);
          double u[3] = {b[0] - a[0], b[1] - a[1], b[2] - a[2]}, w[3] = {c[0] - a[0], c[1] - a[1], c[2] - a[2]};
          x[0] = u[1] * w[2] - u[2] * w[1]; x[1] = u[2] * w[0] - u[0] * w[2]; x[2] = u[0] * w[1] - u[1] * w[0];
        }
        (void)eva;
        double l = 0, s = 0;
        for (int d = 0; d < D; d++) { l += x[d] * x[d]; s += x[d] * (m.point(opp)[d] - m.point(fverts[0])[d]); }
        l = std::sqrt(l);
        for (int d = 0; d < D; d++) nn[d] = (s > 0 ? -x[d] : x[d]) / l;
        st.tf_dom[2 * tf] = group_of(ea);
        st.tf_dom[2 * tf + 1] = group_of(eb);
        st.tf_c[2 * tf] = c_el[ea];
        st.tf_c[2 * tf + 1] = c_el[eb];
      }
      tf++;
    }
  }

  // boundary-data offsets in tent order (serial prefix)
  long long off = 0, poff = 0;
  for (int tf = 0; tf < st.ntf; tf++) {
    if (st.tf_kind[tf] == kInterface) continue;
    st.tf_data[tf] = off;
    st.tf_point[tf] = poff;
    off += (long long)(st.tf_kind[tf] == kNeumann ? D : 1) * Q;
    poff += Q;
  }
  st.nbnd_values = off;
  st.nbnd_points = poff;
  return st;
}

void boundary_sample_points(const StagedTents& st, const SimplexRule& rule, double* out) {
  const int D = st.D, d1 = st.d1, Q = rule.nip;
#pragma omp parallel for schedule(static)
  for (int tf = 0; tf < st.ntf; tf++) {
    if (st.tf_point[tf] < 0) continue;
    const double* V = &st.tf_verts[(size_t)tf * d1 * d1];
    double* o = out + st.tf_point[tf] * d1;
    for (int q = 0; q < Q; q++)
      for (int r = 0; r < d1; r++) {
        double x = V[r];
        for (int n = 0; n < D; n++) x += (V[(n + 1) * d1 + r] - V[r]) * rule.xi[(size_t)q * D + n];
        o[q * d1 + r] = x;
      }
  }
}

// Record layouts: cuda/tent_fast.cuh (the offsets are recomputed here from D; solver.cu static_asserts the strides).
FastStage stage_fast(const SimplexMesh& m, const TentSlab& slab, const StagedTents& st, const std::vector<double>& c_el,
                     const SimplexRule& rule, const BasisCsr& basis_so, const BasisCsr& basis_fo, int fe_stride,
                     int ff_stride) {
  const int D = st.D, d1 = st.d1, nt = st.ntents;
  FastStage fs;
  fs.fe_stride = fe_stride;
  fs.ff_stride = ff_stride;
  fs.fe.assign((size_t)st.nte * fe_stride, 0.0);
  fs.ff.assign((size_t)std::max(st.ntf, 1) * ff_stride, 0.0);
  fs.ok.assign(nt, 1);
  fs.sqrtw.resize(rule.nip);
  for (int q = 0; q < rule.nip; q++) {
    if (!(rule.w[q] > 0)) throw std::runtime_error("stage_fast: quadrature weights must be positive");
    fs.sqrtw[q] = std::sqrt(rule.w[q]);
  }
  const int o_xv = 0, o_tb = d1 * D, o_tt = o_tb + d1, o_top = o_tt + d1, o_bot = o_top + D + 2, o_sc = o_bot + D + 3;
  const int f_v0 = 0, f_e = d1, f_n = f_e + D * d1, f_misc = f_n + D;
  if (o_sc + 3 > fe_stride || f_misc + 5 > ff_stride) throw std::runtime_error("stage_fast: record stride too small");
#pragma omp parallel for schedule(dynamic, 256)
  for (int t = 0; t < nt; t++) {
    const double* x0 = &st.center[(size_t)t * d1];
    const double t0 = x0[D], s = st.scale[t];
    bool ok = st.ndom[t] == 1;
    for (int te = st.te_off[t]; te < st.te_off[t + 1]; te++) {
      const int el = st.te_el[te];
      const double c = c_el[el];
      const int* ev = m.elverts(el);
      double* r = &fs.fe[(size_t)te * fe_stride];
      for (int k = 0; k < d1; k++) {
        for (int u = 0; u < D; u++) r[o_xv + k * D + u] = (m.point(ev[k])[u] - x0[u]) * s;
        r[o_tb + k] = (st.te_tbot[(size_t)te * d1 + k] - t0) * s * c;
        r[o_tt + k] = (st.te_ttop[(size_t)te * d1 + k] - t0) * s * c;
      }
      const double* g = &st.te_geo[(size_t)te * 2 * (d1 + 1)];
      const double* nb = g;
      const double area_b = g[d1];
      const double* ntp = g + d1 + 1;
      const double area_t = g[2 * d1 + 1];
      // W_top / omega = [[n_t I, -n_x], [-n_x^T, n_t / c^2]] = L L^T (twavetents.cpp:293-309)
      const double n_t = ntp[D];
      double nx2 = 0;
      for (int u = 0; u < D; u++) nx2 += ntp[u] * ntp[u];
      const double schur = n_t / (c * c) - (n_t > 0 ? nx2 / n_t : 0.0);
      if (!(n_t > 0) || !(schur > 0) || !(area_t >= 0)) { ok = false; continue; }
      const double sa = std::sqrt(area_t), snt = std::sqrt(n_t);
      r[o_top] = sa * snt * s;
      for (int u = 0; u < D; u++) r[o_top + 1 + u] = sa * (ntp[u] / snt) * s * c;
      r[o_top + 1 + D] = sa * std::sqrt(schur) * s * c;
      r[o_bot] = area_b;
      r[o_bot + 1] = std::sqrt(area_b);
      for (int u = 0; u < d1; u++) r[o_bot + 2 + u] = nb[u];
      r[o_sc] = s;
      r[o_sc + 1] = s * c;
      r[o_sc + 2] = 1.0 / (c * c);
    }
    for (int tf = st.tf_off[t]; tf < st.tf_off[t + 1]; tf++) {
      if (st.tf_kind[tf] == kInterface) { ok = false; continue; }
      const double c = st.tf_c[2 * tf];
      const double* V = &st.tf_verts[(size_t)tf * d1 * d1];
      double* r = &fs.ff[(size_t)tf * ff_stride];
      for (int u = 0; u < D; u++) r[f_v0 + u] = (V[u] - x0[u]) * s;
      r[f_v0 + D] = (V[D] - t0) * s * c;
      for (int k = 0; k < D; k++) {
        for (int u = 0; u < D; u++) r[f_e + k * d1 + u] = (V[(k + 1) * d1 + u] - V[u]) * s;
        r[f_e + k * d1 + D] = (V[(k + 1) * d1 + D] - V[D]) * s * c;
      }
      for (int u = 0; u < D; u++) r[f_n + u] = st.tf_n[(size_t)tf * D + u];
      r[f_misc] = st.tf_area[tf];
      r[f_misc + 1] = s;
      r[f_misc + 2] = s * c;
      r[f_misc + 3] = 1.0 / (c * c);
      r[f_misc + 4] = st.tf_kind[tf] == kNeumann ? 1.0 : 0.0;
    }
    fs.ok[t] = ok ? 1 : 0;
  }
  (void)slab;
  // column-compressed basis: monomial j -> (function, coefficient), functions ascending (fixed summation order)
  for (int fo = 0; fo < 2; fo++) {
    const BasisCsr& b = fo ? basis_fo : basis_so;
    std::vector<int>& cp = fs.cptr[fo];
    cp.assign(b.npoly + 1, 0);
    for (int j : b.col) cp[j + 1]++;
    for (int j = 0; j < b.npoly; j++) cp[j + 1] += cp[j];
    fs.crow[fo].resize(b.col.size());
    fs.cval[fo].resize(b.col.size());
    std::vector<int> fill(cp.begin(), cp.end() - 1);
    for (int i = 0; i < b.nbasis; i++)
      for (int p = b.rowptr[i]; p < b.rowptr[i + 1]; p++) {
        const int j = b.col[p];
        fs.crow[fo][fill[j]] = i;
        fs.cval[fo][fill[j]] = b.val[p];
        fill[j]++;
      }
  }
  // d/dxi_u of sum_j cm_j xi^alpha_j = sum_j' (cm[index(alpha' + e_u)] * (alpha'_u + 1)) xi^alpha', |alpha'| <= p-1
  {
    const int order = basis_so.order;
    const std::vector<int> lo = monomial_exponents(d1, order - 1);
    const int npm1 = (int)lo.size() / d1, npoly = basis_so.npoly;
    fs.dmap.resize((size_t)d1 * npm1);
    fs.dfac.resize((size_t)d1 * npm1);
    for (int u = 0; u < d1; u++)
      for (int j1 = 0; j1 < npm1; j1++) {
        int e[4] = {0, 0, 0, 0};
        for (int k = 0; k < d1; k++) e[k] = lo[(size_t)j1 * d1 + k];
        e[u]++;
        int found = -1;
        for (int j = 0; j < npoly && found < 0; j++) {
          bool same = true;
          for (int k = 0; k < d1; k++) same = same && basis_so.expo[(size_t)j * d1 + k] == e[k];
          if (same) found = j;
        }
        if (found < 0) throw std::runtime_error("stage_fast: monomial not found");
        fs.dmap[(size_t)u * npm1 + j1] = found;
        fs.dfac[(size_t)u * npm1 + j1] = (double)e[u];
      }
  }
  return fs;
}

}  // namespace ngst
