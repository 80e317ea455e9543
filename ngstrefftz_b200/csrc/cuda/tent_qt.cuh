// Host/device contract of the quasi-Trefftz tent kernel (tent_qt.cu): QTWaveTents<D>::Propagate, D = 1, 2
// (src/twavetents.cpp:967-1144; QTWaveTents<3> does not exist in the reference, src/twavetents.cpp:1150-1151).
#pragma once
#include <cuda_runtime.h>

#include "tent_kernel.cuh"

namespace ngst {

struct QtArgs {
  // wavespeed polynomial c(x) = sum_j ccoef[j] x^cexpo[j] (ncterm terms, cexpo: ncterm * D)
  const double* ccoef;
  const int* cexpo;
  int ncterm;
  // Taylor tables at the mesh vertices: d^alpha(1/c^2)(x_v)/alpha! (nA per vertex, degree <= p-2) and d^alpha B /alpha!
  // (nB per vertex, degree <= p-1), first-coordinate-outermost order (PolBasis::IndexMap2)
  const double* aa;
  const double* bb;
  int nA, nB;
  const int* idxA;   // (p-1)^2 table: (nx, ny) -> index into the aa row, -1 outside
  const int* idxB;   // p^2 table
  const int* idx3;   // (p+1)^3 table: (x, y, t) -> monomial index in the (D+1)-variate order-p enumeration, -1 outside
  const int* tent_vertex;   // per tent
  const double* tentsize;   // per tent: TentXdiam (src/twavetents.cpp:1125-1144)
  const double* c_point;    // nel * Q wavespeed at the element quadrature points
  // (D+1)-simplex rule of order 2p for the volume terms
  const double* vxi;        // Qv * (D+1)
  const double* vw;         // Qv
  int Qv;
};

// Fast quasi-Trefftz kernel (tent_qt_fast.cu; D = 2): state-independent records computed once per slab on the device
struct QtFastArgs {
  const double* fe;    // nte * qt_fast_fe_stride(): scaled element vertices, face vertex times, normals, areas
  const double* cone;  // nte * 2 * qt_fast_cone_stride(): cone maps, volume, h / max(1/c^2)
  const double* ff;    // ntf * qt_fast_ff_stride(): boundary faces
};
bool tent_qt_fast_available(int D, int order);
int qt_fast_fe_stride();
int qt_fast_cone_stride();
int qt_fast_ff_stride();
// fills the three record arrays and ok[tent] (1 = the tent can run on the fast kernel)
cudaError_t launch_qt_fast_stage(const TentKernelArgs& a, const QtArgs& qa, int ntents, double* fe, double* cone, double* ff,
                                 int* ok, cudaStream_t stream);
cudaError_t launch_tent_qt_fast(int order, const TentKernelArgs& a, const QtArgs& qa, const QtFastArgs& f,
                                cudaStream_t stream);

size_t tent_qt_smem_bytes(int D, const TentKernelArgs& a, const QtArgs& q);
cudaError_t launch_tent_qt(int D, const TentKernelArgs& a, const QtArgs& q, size_t smem, cudaStream_t stream);

}  // namespace ngst
