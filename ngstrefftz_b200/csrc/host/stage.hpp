// Flattening of a pitched tent slab into device-ready SoA tables (staged once in HBM and
// reused by every Propagate, as the reference reuses its tents: docs/notebooks/twavetents.ipynb cell 6).
//
// Per tent the reference recomputes, inside the Propagate lambda, the centre and anisotropic diameter
// (src/twavetents.cpp:96-100, 861-904), the macro-element grouping (916-933), face vertices / areas /
// normals (594-732) and the boundary-facet lookup (935-963). None of that depends on the wave state, so it
// is evaluated here once per slab on the host, with the reference's formulas, and the kernel reads it.
#pragma once
#include <vector>

#include "basis.hpp"
#include "mesh.hpp"
#include "pitcher.hpp"
#include "quadrature.hpp"

namespace ngst {

enum FaceKind { kDirichlet = 0, kNeumann = 1, kInterface = 2 };

struct StagedTents {
  int D = 0, d1 = 0, ntents = 0, nte = 0, ntf = 0;
  int max_ndom = 1, max_els = 0;
  // per tent
  std::vector<double> center;   // ntents * d1  (x_vertex, (ttop+tbot)/2)
  std::vector<double> scale;    // ntents       1 / TentAdiam
  std::vector<int> ndom;        // ntents
  std::vector<int> te_off;      // ntents + 1  -> tent-element records
  std::vector<int> tf_off;      // ntents + 1  -> time-like face records
  // per tent-element
  std::vector<int> te_el;       // element id
  std::vector<int> te_dom;      // macro-element group
  std::vector<double> te_tbot;  // nte * d1   vertex times of the bottom face, element-local order
  std::vector<double> te_ttop;  // nte * d1
  std::vector<double> te_geo;   // nte * 2*(d1+1): n_bot[d1], area_bot, n_top[d1], area_top
  // per time-like face (boundary facet or material interface inside a tent)
  std::vector<int> tf_kind;     // FaceKind
  std::vector<int> tf_dom;      // ntf * 2: (in,out) groups; boundary: group of the adjacent element twice
  std::vector<double> tf_c;     // ntf * 2: wavespeed of elnums[0], elnums[1] (boundary: adjacent element twice)
  std::vector<double> tf_verts; // ntf * d1*d1: vertex k at [k*d1 .. k*d1+d1)
  std::vector<double> tf_n;     // ntf * D   spatial unit normal (outward / out of elnums[0])
  std::vector<double> tf_area;  // ntf
  std::vector<long long> tf_data;  // ntf: offset (in doubles) into the boundary-data array, -1 for interfaces
  long long nbnd_values = 0;       // total doubles of boundary data per Propagate
  long long nbnd_points = 0;       // total boundary sample points
  std::vector<long long> tf_point; // ntf: offset (in points) into the boundary sample-point array, -1 for interfaces
};

// c_vertex: wavespeed "evaluated at the vertices" for TentAdiam (twavetents.cpp:871-895).
// qt = true stages for QTWaveTents (src/twavetents.cpp:967-1144): one domain per tent (no macro elements, 992-993), no
// interface faces, isotropic scale 1 / TentXdiam (984-990, 1125-1144).
StagedTents stage_tents(const SimplexMesh& m, const TentSlab& slab, const std::vector<double>& c_el,
                        const std::vector<double>& c_vertex, const SimplexRule& rule, bool qt = false);

// Operands of the fast single-wavespeed kernel (cuda/tent_fast.cu; record layouts in cuda/tent_fast.cuh): everything
// per tent-element / boundary face that does not depend on the wave state, in scaled tent coordinates.
struct FastStage {
  int fe_stride = 0, ff_stride = 0;
  std::vector<double> fe;          // nte * fe_stride
  std::vector<double> ff;          // ntf * ff_stride (interface faces: zeros, never read by the fast kernel)
  std::vector<double> sqrtw;       // nip
  std::vector<unsigned char> ok;   // ntents: 1 = tent can run on the fast kernel (one wavespeed region, causal top faces)
  // column-compressed basis [fosystem]
  std::vector<int> cptr[2], crow[2];
  std::vector<double> cval[2];
  // derivative maps of the monomial basis (degree <= p-1 enumeration, first coordinate outermost)
  std::vector<int> dmap;           // d1 * npm1
  std::vector<double> dfac;        // d1 * npm1
};

FastStage stage_fast(const SimplexMesh& m, const TentSlab& slab, const StagedTents& st, const std::vector<double>& c_el,
                     const SimplexRule& rule, const BasisCsr& basis_so, const BasisCsr& basis_fo, int fe_stride,
                     int ff_stride);

// Geometry helpers restating twavetents.cpp:633-732; vert holds d1 vertices of d1 coordinates each
// (vertex k at vert[k*d1 .. k*d1+d1), i.e. the reference's Mat column k).
double tent_face_area(int D, const double* vert);
void tent_face_normal(int D, const double* vert, int top, double* n);

// Sample points (x_0..x_{D-1}, t) of all boundary faces, nbnd_points * d1 doubles (twavetents.cpp:355-365).
void boundary_sample_points(const StagedTents& st, const SimplexRule& rule, double* out);

}  // namespace ngst
