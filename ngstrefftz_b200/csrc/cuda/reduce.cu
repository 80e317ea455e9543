// Per-element partial sums for Energy / Error / L2Error (src/twavetents.cpp:766-847).
// One thread per element, summation order inside an element as in the reference (point outer,
// component inner); the host adds the nel partial sums in element order, so the result is deterministic.
#include <cuda_runtime.h>

namespace ngst {

// largest element count of the allocation-free Energy (2 MB of static device memory per GPU); the bench slab has 131 712
constexpr int kStaticReduceMaxElements = 262144;

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
// ---- Energy without any allocation (ngst_twave_energy_resident_noalloc). The per-element partial sums and the result
// live in a file-scope device array of fixed size that every GPU loading this library holds until exit. All calls share
// it: calls that overlap on one GPU race. Nothing in it has to start at zero -- the first kernel overwrites entries
// [0, nel), the second one reads exactly those and overwrites the result slot -- so contents left by an earlier call
// (entries >= nel included) are never read.
static __device__ __align__(256) double g_static_scratch[kStaticReduceMaxElements + 32];

// the partial sums added up in element order by ONE thread: the order (and hence the bits) of the host loop in
// Solver::ReduceElements. Loads are issued eight at a time, the additions stay sequential.
__global__ void ordered_sum_kernel(const double* __restrict__ part, int n, double* __restrict__ result) {
  if (blockIdx.x != 0 || threadIdx.x != 0) return;
  double s = 0.0;
  int i = 0;
  for (; i + 8 <= n; i += 8) {
    double v[8];
#pragma unroll
    for (int k = 0; k < 8; k++) v[k] = __ldcg(part + i + k);
#pragma unroll
    for (int k = 0; k < 8; k++) s += v[k];
  }
  for (; i < n; i++) s += __ldcg(part + i);
  *result = s;
}

// Device address of g_static_scratch per device. In host code the array's name is only a handle for the runtime, not a
// device address: the address comes from cudaGetSymbolAddress, once per device, with that device current.
constexpr int kMaxDevices = 64;
double* g_static_scratch_ptr[kMaxDevices] = {};
}  // namespace

int static_reduce_max_elements() { return kStaticReduceMaxElements; }

cudaError_t launch_energy_static(int D, int nel, int Q, int ncomp, int fosystem, const double* elw, const double* c_el,
                                 const double* c_point, const double* a, double* host_out, cudaStream_t stream) {
  if (nel < 0 || nel > kStaticReduceMaxElements) return cudaErrorInvalidValue;
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev < 0 || dev >= kMaxDevices) return cudaErrorInvalidDevice;
  double* scratch = g_static_scratch_ptr[dev];
  if (!scratch) {  // set-up, once per device (handles are not thread-safe; two racing first calls would store the same address)
    e = cudaGetSymbolAddress((void**)&scratch, g_static_scratch);
    if (e != cudaSuccess) return e;
    if (!scratch) return cudaErrorInvalidSymbol;
    g_static_scratch_ptr[dev] = scratch;
  }
  double* result = scratch + kStaticReduceMaxElements;  // 256-byte aligned slot behind the partial sums
  const int threads = 128;
  if (nel > 0)
    element_reduce_kernel<<<(nel + threads - 1) / threads, threads, 0, stream>>>(0, D, nel, Q, ncomp, fosystem, elw, c_el,
                                                                                  c_point, a, nullptr, scratch);
  ordered_sum_kernel<<<1, 32, 0, stream>>>(scratch, nel, result);
  e = cudaGetLastError();
  if (e != cudaSuccess) return e;
  e = cudaMemcpyAsync(host_out, result, sizeof(double), cudaMemcpyDeviceToHost, stream);
  if (e != cudaSuccess) return e;
  return cudaStreamSynchronize(stream);
}

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
