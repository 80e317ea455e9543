// Per-element partial sums for Energy / Error / L2Error (src/twavetents.cpp:766-847).
// One thread per element, summation order inside an element as in the reference (point outer,
// component inner); the host adds the nel partial sums in element order, so the result is deterministic.
#include <cuda_runtime.h>

namespace ngst {

namespace {
__global__ void element_reduce_kernel(int mode, int D, int nel, int Q, int ncomp, int fo, const double* __restrict__ elw,
                                      const double* __restrict__ c_el, const double* __restrict__ c_point,
                                      const double* __restrict__ a, const double* __restrict__ b,
                                      double* __restrict__ out) {
  const int el = blockIdx.x * blockDim.x + threadIdx.x;
  if (el >= nel) return;
  const double* wa = a + (size_t)el * ncomp * Q;
  const double* wb = b ? b + (size_t)el * ncomp * Q : nullptr;
  const double* w = elw + (size_t)el * Q;
  // wavespeed per point when given (the reference evaluates the CF at every point, twavetents.cpp:785-793, 835-843)
  const double* cp = c_point ? c_point + (size_t)el * Q : nullptr;
  const double cel = c_el[el];
  const int off = fo ? 0 : 1;
  double s = 0.0;
  if (mode == 0) {  // Energy, twavetents.cpp:824-847
    for (int q = 0; q < Q; q++) {
      const double c = cp ? cp[q] : cel;
      const double cm2 = 1.0 / (c * c);
      for (int d = 0; d <= D; d++) {
        const double v = wa[(size_t)(off + d) * Q + q];
        s += 0.5 * ((d == D ? cm2 : 1.0) * (v * v) * w[q]);
      }
    }
  } else if (mode == 1) {  // Error, twavetents.cpp:766-800
    for (int q = 0; q < Q; q++) {
      const double c = cp ? cp[q] : cel;
      const double cm2 = 1.0 / (c * c);
      for (int d = 0; d <= D; d++) {
        const double v = wa[(size_t)(off + d) * Q + q] - wb[(size_t)(off + d) * Q + q];
        s += (d == D ? cm2 : 1.0) * (v * v) * w[q];
      }
    }
  } else {  // L2Error, twavetents.cpp:802-822 (component 0)
    for (int q = 0; q < Q; q++) {
      const double v = wa[q] - wb[q];
      s += v * v * w[q];
    }
  }
  out[el] = s;
}
}  // namespace

cudaError_t launch_element_reduce(int mode, int D, int nel, int Q, int ncomp, int fosystem, const double* elw,
                                  const double* c_el, const double* c_point, const double* a, const double* b,
                                  double* out, cudaStream_t stream) {
  if (nel <= 0) return cudaSuccess;
  const int threads = 128;
  element_reduce_kernel<<<(nel + threads - 1) / threads, threads, 0, stream>>>(mode, D, nel, Q, ncomp, fosystem, elw,
                                                                                c_el, c_point, a, b, out);
  return cudaGetLastError();
}

}  // namespace ngst
