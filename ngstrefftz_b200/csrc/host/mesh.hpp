// Simplex mesh tables for the tent-pitched Trefftz-DG path.
//
// This replaces the slice of NGSolve's MeshAccess the hot path queries
// (reference call sites: src/twavetents.cpp:113,117,380,455,472,601,616,941-952;
// external_dependencies/ngstents/src/tents.cpp:233-236,248,507,532,568).
// NGSolve itself is not vendored in the reference, so global edge / facet numbers
// are OURS: edges and facets are numbered in order of first appearance, looping
// over elements in index order and over NGSolve's local edge / face order.
#pragma once
#include <cstdint>
#include <string>
#include <vector>

namespace ngst {

struct Csr {
  std::vector<int> off;  // size n+1
  std::vector<int> val;
  int rows() const { return (int)off.size() - 1; }
  int size(int r) const { return off[r + 1] - off[r]; }
  const int* row(int r) const { return val.data() + off[r]; }
};

struct SimplexMesh {
  int D = 0;
  int nv = 0, nel = 0, nse = 0;
  std::vector<double> pts;   // nv * D
  std::vector<int> el2v;     // nel * (D+1), element-local vertex order as given
  std::vector<int> se2v;     // nse * D   boundary (surface) elements
  std::vector<int> se_kind;  // nse       0 = Dirichlet, 1 = Neumann (boundary named "neumann")

  // derived topology
  int nedges = 0, nfacets = 0;
  std::vector<int> edge2v;    // nedges * 2, (min,max)
  std::vector<int> el2edge;   // nel * nEdgesPerEl()
  std::vector<int> facet2v;   // nfacets * D (1-D: the vertex itself)
  std::vector<int> el2facet;  // nel * (D+1)
  std::vector<int> facet2el;  // nfacets * 2, -1 where absent
  std::vector<int> facet2se;  // nfacets, -1 if no surface element
  std::vector<int> se2facet;  // nse
  Csr v2el;                   // ascending element index (MeshAccess::GetVertexElements)
  Csr v2v, v2e;               // ascending edge index (tents.cpp:526-540)
  Csr edge2el;                // elements containing an edge (GetEdgeElements)

  int nEdgesPerEl() const { return D == 1 ? 1 : (D == 2 ? 3 : 6); }
  const double* point(int v) const { return pts.data() + (size_t)v * D; }
  const int* elverts(int e) const { return el2v.data() + (size_t)e * (D + 1); }

  // Builds every derived table. Throws std::runtime_error on malformed input.
  void finalize();
};

// Structured generators standing in for netgen's SegMesh / unit_square / unit_cube
// (test/tents.py:24,38,142). n[d] cells per direction on the box [lo,hi]; 2-D cells are
// split into 2 triangles along the same diagonal, 3-D cells into 6 Kuhn tetrahedra.
SimplexMesh make_structured(int D, const int* n, const double* lo, const double* hi);

}  // namespace ngst
