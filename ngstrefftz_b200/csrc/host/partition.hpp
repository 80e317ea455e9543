// Spatial partition of a pitched tent slab over GPUs (the reference has no distributed Propagate).
//
// Ownership is by vertex: a tent runs on the rank that owns its vertex (struct Tent::vertex, tents.hpp:41). An element is
// LOCAL to every rank that owns one of its vertices (that rank's tents read and overwrite its state row,
// src/twavetents.cpp:235-246, 586-589) and OWNED (counted in Energy / Error) by the owner of its lowest-numbered vertex.
// Tents of one level share no element, so after every level each rank sends the rows it has just updated to the other
// ranks that hold them; both sides derive the lists from the global slab and the owner array without communicating.
#pragma once
#include <vector>

#include "mesh.hpp"
#include "pitcher.hpp"

namespace ngst {

struct PartitionPlan {
  int rank = 0, nranks = 1, nlev = 0;
  std::vector<int> own_tents;        // global tent ids of this rank, ascending
  std::vector<int> loc2glob;         // local element -> global element, ascending
  std::vector<int> glob2loc;         // global element -> local element or -1
  std::vector<unsigned char> owned;  // per local element: counted by this rank in reductions
  std::vector<int> final_level;      // per local element: level of the last tent of the GLOBAL slab that contains it, -1 if none
  // rows exchanged after level l with peer q: [off[l*nranks+q], off[l*nranks+q+1]) into *_el (LOCAL element ids,
  // ordered by global element id on both sides)
  std::vector<int> send_off, send_el;
  std::vector<int> recv_off, recv_el;
  int max_rows_per_level() const;    // max over levels of max(total send rows, total recv rows)

  // ---- dependency-driven execution across ranks (peer-to-peer exchange inside the tent kernel, solver.cu): the hazards of
  // the GLOBAL slab -- tent t waits for the last earlier tent, on whatever rank, that touched each of its elements.
  // Own tents are numbered in own_tents order on every rank.
  std::vector<int> p2p_indeg;                         // per own tent: distinct predecessor tents
  std::vector<int> p2p_lsucc_off, p2p_lsucc;          // successors on this rank (own-tent index)
  std::vector<int> p2p_rsucc_off, p2p_rsucc_peer, p2p_rsucc_tent;  // successors elsewhere: (rank, its own-tent index)
  // per own tent-element, in the order of the restricted slab's element lists: the OTHER ranks that hold the element,
  // with the element's local index there (where this tent's rows are pushed to)
  std::vector<int> p2p_push_off, p2p_push_peer, p2p_push_el;
};

// vowner[v] in [0, nranks). Throws std::invalid_argument on malformed input.
PartitionPlan make_partition_plan(const SimplexMesh& m, const TentSlab& slab, const std::vector<int>& vowner, int nranks,
                                  int rank);

// The tents of `own` (ascending global ids) as a slab of their own: same mesh, same level numbering (so that levels stay
// aligned across ranks), tent-local successor lists.
TentSlab restrict_slab(const TentSlab& slab, const std::vector<int>& own);

}  // namespace ngst
