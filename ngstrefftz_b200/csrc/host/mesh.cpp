#include "mesh.hpp"

#include <algorithm>
#include <stdexcept>
#include <unordered_map>

namespace ngst {

namespace {
// NGSolve's local numbering (ElementTopology): used only to fix OUR global numbering.
const int kTrigEdges[3][2] = {{2, 0}, {1, 2}, {0, 1}};
const int kTetEdges[6][2] = {{3, 0}, {3, 1}, {3, 2}, {0, 1}, {0, 2}, {1, 2}};
const int kTetFaces[4][3] = {{3, 1, 2}, {3, 2, 0}, {3, 0, 1}, {0, 2, 1}};

inline uint64_t key2(int a, int b) {
  if (a > b) std::swap(a, b);
  return ((uint64_t)a << 32) | (uint32_t)b;
}
inline uint64_t key3(int a, int b, int c) {
  if (a > b) std::swap(a, b);
  if (b > c) std::swap(b, c);
  if (a > b) std::swap(a, b);
  return ((uint64_t)a << 42) | ((uint64_t)b << 21) | (uint64_t)c;
}

Csr build_csr(int nrows, const std::vector<std::pair<int, int>>& entries) {
  // entries are (row, value) in insertion order; order within a row is preserved
  Csr t;
  t.off.assign(nrows + 1, 0);
  for (auto& e : entries) t.off[e.first + 1]++;
  for (int i = 0; i < nrows; i++) t.off[i + 1] += t.off[i];
  t.val.resize(entries.size());
  std::vector<int> pos(t.off.begin(), t.off.end() - 1);
  for (auto& e : entries) t.val[pos[e.first]++] = e.second;
  return t;
}
}  // namespace

void SimplexMesh::finalize() {
  if (D < 1 || D > 3) throw std::runtime_error("mesh: D must be 1, 2 or 3");
  if ((int)pts.size() != nv * D) throw std::runtime_error("mesh: pts size mismatch");
  if ((int)el2v.size() != nel * (D + 1)) throw std::runtime_error("mesh: el2v size mismatch");
  if ((int)se2v.size() != nse * D) throw std::runtime_error("mesh: se2v size mismatch");
  if ((int)se_kind.size() != nse) throw std::runtime_error("mesh: se_kind size mismatch");
  if (D == 3 && nv >= (1 << 21)) throw std::runtime_error("mesh: too many vertices for face keys");
  for (int x : el2v)
    if (x < 0 || x >= nv) throw std::runtime_error("mesh: element vertex out of range");
  for (int x : se2v)
    if (x < 0 || x >= nv) throw std::runtime_error("mesh: surface element vertex out of range");

  const int d1 = D + 1;
  const int nepe = nEdgesPerEl();

  // ---- edges
  edge2v.clear();
  el2edge.assign((size_t)nel * nepe, -1);
  {
    std::unordered_map<uint64_t, int> emap;
    emap.reserve((size_t)nel * 2);
    for (int e = 0; e < nel; e++) {
      const int* v = elverts(e);
      for (int k = 0; k < nepe; k++) {
        int a, b;
        if (D == 1) { a = v[0]; b = v[1]; }
        else if (D == 2) { a = v[kTrigEdges[k][0]]; b = v[kTrigEdges[k][1]]; }
        else { a = v[kTetEdges[k][0]]; b = v[kTetEdges[k][1]]; }
        auto it = emap.find(key2(a, b));
        int id;
        if (it == emap.end()) {
          id = (int)edge2v.size() / 2;
          emap.emplace(key2(a, b), id);
          edge2v.push_back(std::min(a, b));
          edge2v.push_back(std::max(a, b));
        } else id = it->second;
        el2edge[(size_t)e * nepe + k] = id;
      }
    }
    nedges = (int)edge2v.size() / 2;
  }

  // ---- facets
  el2facet.assign((size_t)nel * d1, -1);
  facet2v.clear();
  if (D == 1) {
    nfacets = nv;
    facet2v.resize(nv);
    for (int i = 0; i < nv; i++) facet2v[i] = i;
    for (int e = 0; e < nel; e++) { el2facet[2 * e] = elverts(e)[0]; el2facet[2 * e + 1] = elverts(e)[1]; }
  } else if (D == 2) {
    nfacets = nedges;
    facet2v = edge2v;
    for (int e = 0; e < nel; e++)
      for (int k = 0; k < 3; k++) el2facet[3 * e + k] = el2edge[3 * e + k];
  } else {
    std::unordered_map<uint64_t, int> fmap;
    fmap.reserve((size_t)nel * 3);
    for (int e = 0; e < nel; e++) {
      const int* v = elverts(e);
      for (int k = 0; k < 4; k++) {
        int a = v[kTetFaces[k][0]], b = v[kTetFaces[k][1]], c = v[kTetFaces[k][2]];
        auto it = fmap.find(key3(a, b, c));
        int id;
        if (it == fmap.end()) {
          id = (int)facet2v.size() / 3;
          fmap.emplace(key3(a, b, c), id);
          int s[3] = {a, b, c};
          std::sort(s, s + 3);
          facet2v.insert(facet2v.end(), s, s + 3);
        } else id = it->second;
        el2facet[(size_t)e * 4 + k] = id;
      }
    }
    nfacets = (int)facet2v.size() / 3;
  }

  // ---- facet -> elements
  facet2el.assign((size_t)nfacets * 2, -1);
  for (int e = 0; e < nel; e++)
    for (int k = 0; k < d1; k++) {
      int f = el2facet[(size_t)e * d1 + k];
      if (facet2el[2 * f] < 0) facet2el[2 * f] = e;
      else if (facet2el[2 * f + 1] < 0) facet2el[2 * f + 1] = e;
      else throw std::runtime_error("mesh: facet shared by more than two elements");
    }

  // ---- surface elements -> facets
  facet2se.assign(nfacets, -1);
  se2facet.assign(nse, -1);
  {
    std::unordered_map<uint64_t, int> fkey;
    if (D > 1) {
      fkey.reserve(nfacets);
      for (int f = 0; f < nfacets; f++) {
        const int* v = &facet2v[(size_t)f * D];
        fkey.emplace(D == 2 ? key2(v[0], v[1]) : key3(v[0], v[1], v[2]), f);
      }
    }
    for (int s = 0; s < nse; s++) {
      const int* v = &se2v[(size_t)s * D];
      int f;
      if (D == 1) f = v[0];
      else {
        auto it = fkey.find(D == 2 ? key2(v[0], v[1]) : key3(v[0], v[1], v[2]));
        if (it == fkey.end()) throw std::runtime_error("mesh: surface element is not a mesh facet");
        f = it->second;
      }
      if (facet2el[2 * f + 1] >= 0) throw std::runtime_error("mesh: surface element on an interior facet");
      se2facet[s] = f;
      if (facet2se[f] < 0) facet2se[f] = s;  // first match, as GetFacetSurfaceElement (twavetents.cpp:949-960)
    }
  }

  // ---- vertex -> elements (ascending element index)
  {
    std::vector<std::pair<int, int>> ent;
    ent.reserve((size_t)nel * d1);
    for (int e = 0; e < nel; e++)
      for (int k = 0; k < d1; k++) ent.emplace_back(elverts(e)[k], e);
    v2el = build_csr(nv, ent);
  }
  // ---- vertex -> vertex / edge, ascending edge index (tents.cpp:526-540)
  {
    std::vector<std::pair<int, int>> ev, ee;
    ev.reserve((size_t)nedges * 2);
    ee.reserve((size_t)nedges * 2);
    for (int e = 0; e < nedges; e++) {
      int v1 = edge2v[2 * e], v2 = edge2v[2 * e + 1];
      ev.emplace_back(v1, v2); ee.emplace_back(v1, e);
      ev.emplace_back(v2, v1); ee.emplace_back(v2, e);
    }
    v2v = build_csr(nv, ev);
    v2e = build_csr(nv, ee);
  }
  // ---- edge -> elements
  {
    std::vector<std::pair<int, int>> ent;
    ent.reserve((size_t)nel * nepe);
    for (int e = 0; e < nel; e++)
      for (int k = 0; k < nepe; k++) ent.emplace_back(el2edge[(size_t)e * nepe + k], e);
    edge2el = build_csr(nedges, ent);
  }
}

SimplexMesh make_structured(int D, const int* n, const double* lo, const double* hi) {
  SimplexMesh m;
  m.D = D;
  if (D < 1 || D > 3) throw std::runtime_error("make_structured: D must be 1, 2 or 3");
  for (int d = 0; d < D; d++)
    if (n[d] < 1) throw std::runtime_error("make_structured: need at least one cell per direction");
  int nx = n[0], ny = D > 1 ? n[1] : 0, nz = D > 2 ? n[2] : 0;
  auto X = [&](int d, int i) { return lo[d] + (hi[d] - lo[d]) * (double)i / (double)n[d]; };
  if (D == 1) {
    m.nv = nx + 1;
    for (int i = 0; i <= nx; i++) m.pts.push_back(X(0, i));
    m.nel = nx;
    for (int i = 0; i < nx; i++) { m.el2v.push_back(i); m.el2v.push_back(i + 1); }
    m.nse = 2;
    m.se2v = {0, nx};
    m.se_kind = {0, 0};
  } else if (D == 2) {
    auto vid = [&](int i, int j) { return j * (nx + 1) + i; };
    m.nv = (nx + 1) * (ny + 1);
    for (int j = 0; j <= ny; j++)
      for (int i = 0; i <= nx; i++) { m.pts.push_back(X(0, i)); m.pts.push_back(X(1, j)); }
    for (int j = 0; j < ny; j++)
      for (int i = 0; i < nx; i++) {
        int a = vid(i, j), b = vid(i + 1, j), c = vid(i + 1, j + 1), d = vid(i, j + 1);
        int t[6] = {a, b, c, a, c, d};  // both counter-clockwise, shared diagonal a-c
        m.el2v.insert(m.el2v.end(), t, t + 6);
      }
    m.nel = 2 * nx * ny;
    // boundary segments, counter-clockwise: bottom, right, top, left
    for (int i = 0; i < nx; i++) { m.se2v.push_back(vid(i, 0)); m.se2v.push_back(vid(i + 1, 0)); }
    for (int j = 0; j < ny; j++) { m.se2v.push_back(vid(nx, j)); m.se2v.push_back(vid(nx, j + 1)); }
    for (int i = nx; i > 0; i--) { m.se2v.push_back(vid(i, ny)); m.se2v.push_back(vid(i - 1, ny)); }
    for (int j = ny; j > 0; j--) { m.se2v.push_back(vid(0, j)); m.se2v.push_back(vid(0, j - 1)); }
    m.nse = 2 * (nx + ny);
    m.se_kind.assign(m.nse, 0);
  } else {
    auto vid = [&](int i, int j, int k) { return (k * (ny + 1) + j) * (nx + 1) + i; };
    m.nv = (nx + 1) * (ny + 1) * (nz + 1);
    for (int k = 0; k <= nz; k++)
      for (int j = 0; j <= ny; j++)
        for (int i = 0; i <= nx; i++) { m.pts.push_back(X(0, i)); m.pts.push_back(X(1, j)); m.pts.push_back(X(2, k)); }
    const int perms[6][3] = {{0, 1, 2}, {0, 2, 1}, {1, 0, 2}, {1, 2, 0}, {2, 0, 1}, {2, 1, 0}};
    for (int k = 0; k < nz; k++)
      for (int j = 0; j < ny; j++)
        for (int i = 0; i < nx; i++)
          for (int p = 0; p < 6; p++) {  // Kuhn: walk from (0,0,0) to (1,1,1) along a permutation of axes
            int c[3] = {i, j, k};
            m.el2v.push_back(vid(c[0], c[1], c[2]));
            for (int s = 0; s < 3; s++) { c[perms[p][s]]++; m.el2v.push_back(vid(c[0], c[1], c[2])); }
          }
    m.nel = 6 * nx * ny * nz;
    // boundary triangles: two per cell face, matching the Kuhn split (diagonal from the low to the high corner)
    // Orientation: (b-a) x (c-a) points out of the domain, which is what makes
    // n = -TentFaceNormal(vert, 0) the outward normal at twavetents.cpp:345-348.
    auto quad = [&](int a, int b, int c, int d, bool flip) {  // a low corner, c high corner
      int t[6] = {a, b, c, a, c, d};
      if (flip) { std::swap(t[1], t[2]); std::swap(t[4], t[5]); }
      m.se2v.insert(m.se2v.end(), t, t + 6);
    };
    for (int j = 0; j < ny; j++)
      for (int i = 0; i < nx; i++) {
        quad(vid(i, j, 0), vid(i + 1, j, 0), vid(i + 1, j + 1, 0), vid(i, j + 1, 0), true);
        quad(vid(i, j, nz), vid(i + 1, j, nz), vid(i + 1, j + 1, nz), vid(i, j + 1, nz), false);
      }
    for (int k = 0; k < nz; k++)
      for (int i = 0; i < nx; i++) {
        quad(vid(i, 0, k), vid(i + 1, 0, k), vid(i + 1, 0, k + 1), vid(i, 0, k + 1), false);
        quad(vid(i, ny, k), vid(i + 1, ny, k), vid(i + 1, ny, k + 1), vid(i, ny, k + 1), true);
      }
    for (int k = 0; k < nz; k++)
      for (int j = 0; j < ny; j++) {
        quad(vid(0, j, k), vid(0, j + 1, k), vid(0, j + 1, k + 1), vid(0, j, k + 1), true);
        quad(vid(nx, j, k), vid(nx, j + 1, k), vid(nx, j + 1, k + 1), vid(nx, j, k + 1), false);
      }
    m.nse = (int)m.se2v.size() / 3;
    m.se_kind.assign(m.nse, 0);
  }
  m.finalize();
  return m;
}

}  // namespace ngst
