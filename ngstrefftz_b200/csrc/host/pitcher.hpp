// Tent pitching: serial greedy advancing front, bit-exact with the reference's schedule.
//
// Follows external_dependencies/ngstents/src/tents.cpp:78-359 (PitchTents<DIM>),
// 385-474 (GetReadyVertices, PickNextVertexForPitching, UpdateNeighbours),
// 476-563 (InitializeMeshData), 796-978 (EdgeGradientPitcher), 627-794 (VolumeGradientPitcher).
// The reference keeps the ready list in an ngcore Array and scans it linearly per tent
// (tents.cpp:431-443); here the same list semantics (Append / DeleteElement = move-last-into-slot /
// RemoveElement = order-preserving erase / first-position-of-minimum-level) are kept in a
// tombstoned slot array under a min segment tree, so one tent costs O(log nv) instead of O(nv)
// while producing the identical tent sequence (tests compare against the literal oracle).
#pragma once
#include <vector>

#include "mesh.hpp"

namespace ngst {

enum PitchMethod { kVolGrad = 1, kEdgeGrad = 2 };

struct TentSlab {
  const SimplexMesh* mesh = nullptr;
  int method = kEdgeGrad;
  std::vector<double> cmax_el;  // per-element maximal wavespeed (SetMaxWavespeed)
  double dt = 0;
  bool pitched = false;
  int generation = 0;  // bumped by every PitchTents / import: solvers staged from an older generation refuse to run
  int nlayers = 0;  // as the reference's member: GetNLayers() == nlayers + 1

  // flattened tents (struct Tent, tents.hpp:38-75)
  int ntents = 0;
  std::vector<int> vertex, level;
  std::vector<double> tbot, ttop, maxslope;
  Csr nbv;                     // neighbour vertices
  std::vector<double> nbtime;  // aligned with nbv.val
  Csr els;                     // vertex patch elements
  Csr facets;                  // internal facets
  Csr dag;                     // tent_dependency: successors of each tent

  // level sets: tents of level l are level_tents[level_off[l] .. level_off[l+1]) in ascending tent index
  std::vector<int> level_off, level_tents;

  int GetNLayers() const { return nlayers + 1; }
  double MaxSlope() const;

  // returns false when the front could not reach dt (tents.cpp:265-288)
  bool PitchTents(double dt, bool local_ct, double global_ct);

  // Level-synchronous pitcher with the partition-independent pick rule "ready vertex of minimal (level, vertex number)"
  // (no reference counterpart; other tents than the reference's order). Parallel over the tents of a level.
  bool PitchTentsLevelSync(double dt, bool local_ct, double global_ct);

  // Tents pitched elsewhere (the Tent fields of python_tents.cpp:119-128, flattened; *_off have n + 1 entries). level and
  // facet_off / facets may be null: levels are then replayed as the greedy loop assigns them, internal facets taken from
  // the mesh. Throws std::invalid_argument on lists that are inconsistent with the mesh or not level-schedulable.
  void ImportTents(int n, double dt, const int* vertex, const int* level, const double* tbot, const double* ttop,
                   const int* nb_off, const int* nbv, const double* nbtime, const int* el_off, const int* els,
                   const int* facet_off, const int* facets);
};

}  // namespace ngst
