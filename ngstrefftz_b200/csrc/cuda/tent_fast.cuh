// Host/device contract of the fast single-wavespeed tent kernel (tent_fast.cu).
#pragma once
#include <cuda_runtime.h>

#include "tent_kernel.cuh"

namespace ngst {

// ---- per tent-element record (doubles), staged once per slab by stage_fast():
//   XV  : d1*D  element vertices in scaled tent coordinates, vertex k at [k*D .. k*D+D): (P_k - x_v) * s
//   TB  : d1    bottom-face vertex times, scaled: (tau_k - t0) * s * c
//   TT  : d1    top-face vertex times, scaled
//   TOP : D+2   k0, kd[D], kt: Cholesky factor of the top-face flux matrix (src/twavetents.cpp:293-309) times
//               sqrt(area) and the derivative scalings s / s*c
//   BOT : d1+2  area, sqrt(area), n[d1] of the bottom face
//   SC  : 3     s, s*c, 1/c^2
constexpr int fast_fe_xv(int) { return 0; }
constexpr int fast_fe_tb(int D) { return (D + 1) * D; }
constexpr int fast_fe_tt(int D) { return fast_fe_tb(D) + D + 1; }
constexpr int fast_fe_top(int D) { return fast_fe_tt(D) + D + 1; }
constexpr int fast_fe_bot(int D) { return fast_fe_top(D) + D + 2; }
constexpr int fast_fe_sc(int D) { return fast_fe_bot(D) + D + 3; }
constexpr int fast_fe_stride(int D) { return (fast_fe_sc(D) + 3 + 3) / 4 * 4; }

// ---- per boundary-face record (doubles):
//   V0 : d1     tent vertex at tbot, scaled;  E : D*d1 edge vectors V_{m+1} - V_0, scaled, edge m at [m*d1 ..)
//   N  : D      outward spatial unit normal;  then area, s, s*c, 1/c^2, kind (0 Dirichlet, 1 Neumann)
constexpr int fast_ff_v0(int) { return 0; }
constexpr int fast_ff_e(int D) { return D + 1; }
constexpr int fast_ff_n(int D) { return fast_ff_e(D) + D * (D + 1); }
constexpr int fast_ff_misc(int D) { return fast_ff_n(D) + D; }
constexpr int fast_ff_stride(int D) { return (fast_ff_misc(D) + 5 + 3) / 4 * 4; }

// Peer-to-peer exchange inside the tent kernel (multi-GPU, dependency-driven single launch; device-resident table):
// the state buffers and hazard counters of every rank of the box, mapped into this process (CUDA IPC over NVLink), the
// holders of every tent-element of this rank and the successors of its tents on other ranks (host/partition.hpp).
constexpr int kMaxPeers = 16;
struct P2PArgs {
  double* cstate[kMaxPeers];     // compact state of rank q (this rank's own buffer at its own index)
  double* wavefront[kMaxPeers];  // public rows of rank q
  int* remaining[kMaxPeers];     // hazard counters of rank q's tents
  const int* push_off;           // nte + 1
  const int* push_peer;
  const int* push_el;
  const int* rsucc_off;          // ntents + 1
  const int* rsucc_peer;
  const int* rsucc_tent;
};

struct FastArgs {
  const double* fe;     // nte * fast_fe_stride(D)
  const double* ff;     // ntf * fast_ff_stride(D)
  const double* sqrtw;  // Q: sqrt of the rule weights
  // rule of order 2p-2 for the top-face Gram (exact for its degree-(2p-2) integrand)
  const double* lam_top;    // Qt * (D+1) barycentric coordinates
  const double* sqrtw_top;  // Qt
  int Qt;
  // compact rule exact to degree 2p for the bottom-face stabilisation Gram (second-order form), when the table has one
  const double* lam_stab;   // Qs * (D+1)
  const double* sqrtw_stab; // Qs
  const double* w_stab;     // Qs
  int Qs;
  // compact state (tent_fast.cu, CS = 1): nel * ncomp * Qs, row = [comp][point of the compact rule]; nullptr = the tents
  // read and write the public order-2p rows. te_final[te] != 0: this tent is the last one of the slab (or of the
  // requested levels) to contain the element -> it also writes the public row.
  double* cstate;
  const unsigned char* te_final;
  // factor cache (tent_fast.cu, MODE): 0 = assemble and solve, 1 = assemble, invert, store in ainv, apply, 2 = apply the
  // stored inverse. ainv: per staged tent NP x N doubles, the inverse transposed (column j of A^-1 at [j * NP, j * NP + N)).
  double* ainv;
  int mode;
  const P2PArgs* p2p;   // nullptr: single GPU, or the per-level NCCL exchange
  // ---- spatial sweep (streamed input and output, single GPU, dependency-driven launch; solver.cu SetInitialStreamed):
  // the copy engine delivers the public rows in chunks of consecutive elements while the kernel runs; tents are ticketed
  // in the order in which their dependency cones are complete.
  const int* tent_chunk;          // per tent: the last input chunk it (not its ancestors) reads; nullptr = no streamed input
  const int* chunk_loaded;        // per chunk: non-zero once it has landed (written by a stream memory operation)
  const unsigned char* te_first;  // per tent-element: first tent of the slab to touch the element -> its compact row does not
                                  // exist yet and is projected here from the public row (transfer: Qs x Q, row-major)
  const double* transfer;
  int* chunk_left;                // per chunk: touched elements whose public row is not final yet (output side); nullptr = off
  int chunk_elems;                // elements per chunk
  // column-compressed basis matrix (monomial j -> functions), for sol -> monomial coefficients
  const int* cptr;      // npoly + 1
  const int* crow;      // nnz
  const double* cval;   // nnz
  // derivative maps: coefficient u*npm1 + j' of d/dxi_u in the degree <= p-1 monomial basis = cm[dmap] * dfac
  const int* dmap;      // d1 * npm1
  const double* dfac;   // d1 * npm1
};

bool tent_fast_available(int D, int order);
int tent_fast_compact_points(int D, int order, int fosystem);
size_t tent_fast_inverse_doubles(int D, int order, int fosystem);
bool tent_fast_stages_rows(int D, int order, int fosystem);
size_t tent_fast_smem_bytes(int D, int order, int fosystem);
cudaError_t launch_tent_fast(int D, int order, int fosystem, const TentKernelArgs& a, const FastArgs& f,
                             cudaStream_t stream);

}  // namespace ngst
