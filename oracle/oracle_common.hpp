// TEST INFRASTRUCTURE ONLY -- CPU restatement of the reference's tent path (the "oracle").
// Nothing under oracle/ is linked into, imported by, or called from the product (ngstrefftz_b200/);
// only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference leg use it.
//
// PARITY STATUS: the reference cannot be built or imported here (NGSolve 6.2.2606 is an un-vendored
// dependency, pyproject.toml:11-13), it ships no golden vectors for this path, and its doctests pin only
// leading digits of error norms (test/tents.py:46-54). The oracle is pinned against those digits for the
// reproducible 1-D cases and against mesh-independent known answers (tests/); value-level parity with the
// real reference remains UNPINNED.
#pragma once
#include <vector>

namespace oracle {

// Mesh tables are INPUT (they stand for NGSolve's MeshAccess, which is not part of the reference tree).
struct OMesh {
  int D = 0, nv = 0, nel = 0, nse = 0, nedges = 0, nfacets = 0;
  std::vector<double> pts;
  std::vector<int> el2v, se2v, se_kind, edge2v, el2edge, facet2v, el2facet, facet2el, facet2se;
  // derived like MeshAccess::GetVertexElements / tents.cpp:526-540
  std::vector<std::vector<int>> v2el, v2v, v2e, edge2el;
  void derive();
  const double* point(int v) const { return &pts[(size_t)v * D]; }
  const int* elverts(int e) const { return &el2v[(size_t)e * (D + 1)]; }
  int nepe() const { return D == 1 ? 1 : (D == 2 ? 3 : 6); }
};

// struct Tent, tents.hpp:38-75
struct OTent {
  int vertex = -1;
  double tbot = 0, ttop = 0;
  std::vector<int> nbv;
  std::vector<double> nbtime;
  std::vector<int> els, internal_facets, dependent_tents;
  int level = 0;
  double maxslope = 0;
};

struct OSlab {
  const OMesh* ma = nullptr;
  double dt = 0;
  int nlayers = 0;
  bool has_been_pitched = false;
  std::vector<OTent> tents;
  std::vector<std::vector<int>> tent_dependency;
  int GetNLayers() const { return nlayers + 1; }
};

bool PitchTents(OSlab& slab, const OMesh& ma, int method, const std::vector<double>& cmax_el, double dt,
                bool calc_local_ct, double global_ct, int rule = 0);

}  // namespace oracle
