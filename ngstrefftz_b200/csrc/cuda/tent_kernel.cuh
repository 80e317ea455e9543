// Device-side argument block shared by the tent kernels and the host launcher.
#pragma once

namespace ngst {

struct TentKernelArgs {
  // ---- spatial mesh (static)
  const double* pts;    // nv * D
  const int* el2v;      // nel * (D+1)
  const double* c_el;   // nel
  // ---- quadrature rule of order 2p on the D-simplex
  const double* lam;    // Q * (D+1) barycentric coordinates (xi_0..xi_{D-1}, 1 - sum)
  const double* w;      // Q
  // ---- Trefftz basis (CSR, c = 1) and monomial exponents
  const int* rowptr;    // N + 1
  const int* col;       // nnz
  const double* val;    // nnz
  const int* expo;      // npoly * (D+1)
  int N, Npad, npoly, order, Q, Qc, ncomp, fosystem;
  // ---- tents (StagedTents, device copies)
  const double* center;
  const double* scale;
  const int* ndom;
  const int* te_off;
  const int* tf_off;
  const int* te_el;
  const int* te_dom;
  const double* te_tbot;
  const double* te_ttop;
  const double* te_geo;
  const int* tf_kind;
  const int* tf_dom;
  const double* tf_c;
  const double* tf_verts;
  const double* tf_n;
  const double* tf_area;
  const long long* tf_data;
  const double* bnd_data;  // boundary datum samples for this slab; nullptr = homogeneous
  // ---- state
  double* wavefront;       // nel * ncomp * Q, row = [comp][q]
  const int* tents;        // tent ids of this launch (one level)
  int ntents;
  int max_ndom;
  int* singular;           // incremented when a zero pivot is met
  int* stuck;              // incremented when a tent of the dependency-driven launch gives up waiting (its writes are skipped)
  // ---- factor cache of the generic kernel (tent_generic.cu): gmode 0 = assemble and solve, 1 = assemble, invert
  // (Gauss-Jordan on [A | I] in shared memory), store the inverse in ginv and apply it, 2 = right-hand side + stored inverse +
  // evaluation. ginv: per staged tent (max_ndom * Npad)^2 doubles, the inverse transposed.
  double* ginv;
  int gmode;
  int gwarp;  // gmode 2 only: 1 = one warp per tent (tent_kernel_generic<D, true>), 2 = tent_kernel_small1d (D = 1), 0 = one CTA
  int gnnz;   // longest row of the CSR basis (tent_kernel_small1d keeps a lane's row in registers)
  // ---- dependency-driven scheduling (tent_fast.cu, single GPU): one launch per slab; CTAs draw tents from `tents` in
  // order through `ticket` and wait until remaining[tent] (unfinished predecessors) is zero; a finished tent decrements
  // the counters of its successors. ticket == nullptr: `tents` is one level, CTA b solves tents[b].
  int* ticket;
  int* remaining;
  const int* succ_off;
  const int* succ;
};

}  // namespace ngst
