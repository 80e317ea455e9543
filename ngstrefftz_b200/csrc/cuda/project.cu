// GetWave (src/__init__.py:7-43): per-element L2 projection of the quadrature-point state onto polynomials.
//
// The reference assembles an L2 mass matrix and a right-hand side from IntegrationPointFunction
// (src/specialcoefficientfunction.cpp:48-172) and calls a sparse solver. On affine simplices the mass matrix is
// |detJ| times the reference one, so in a basis psi_k that is orthonormal on the reference simplex the projection is
//   coef[el][c][k] = sum_q w_q psi_k(xi_q) v[el][c][q],
// one small matrix P (ndof x Q) applied to every state row: a batched GEMM, HBM-bound (one pass over the rows read).
// One CTA per kElems elements; P is staged in shared memory, a thread owns one (row, k-slice).
#include <cuda_runtime.h>


namespace ngst {

namespace {
constexpr int kElems = 8;

// elems != nullptr: the listed elements only (out is then indexed by the element's own number as well)
__global__ void project_rows_kernel(int nrows, int Q, int ndof, int comp0, int ncomp_used, int rowlen,
                                    const double* __restrict__ P, const double* __restrict__ wf, double* __restrict__ out,
                                    const int* __restrict__ elems) {
  extern __shared__ double sm[];
  double* Ps = sm;                       // ndof * Q
  double* vs = sm + (size_t)ndof * Q;    // kElems * ncomp_used * Q
  for (int i = threadIdx.x; i < ndof * Q; i += blockDim.x) Ps[i] = P[i];
  const int el0 = blockIdx.x * kElems;
  const int ne = min(kElems, nrows - el0);
  const int nv = ne * ncomp_used * Q;
  for (int i = threadIdx.x; i < nv; i += blockDim.x) {
    const int e = i / (ncomp_used * Q), r = i % (ncomp_used * Q);
    const size_t el = elems ? (size_t)elems[el0 + e] : (size_t)(el0 + e);
    vs[i] = wf[el * rowlen + (size_t)comp0 * Q + r];  // consecutive threads -> consecutive doubles
  }
  __syncthreads();
  const int nout = ne * ncomp_used * ndof;
  for (int i = threadIdx.x; i < nout; i += blockDim.x) {
    const int k = i % ndof, ec = i / ndof;
    const double* v = vs + (size_t)ec * Q;
    const double* p = Ps + (size_t)k * Q;
    double s = 0.0;
    for (int q = 0; q < Q; q++) s = fma(p[q], v[q], s);
    if (elems) out[(size_t)elems[el0 + ec / ncomp_used] * ncomp_used * ndof + (size_t)(ec % ncomp_used) * ndof + k] = s;
    else out[(size_t)el0 * ncomp_used * ndof + i] = s;
  }
}
}  // namespace

size_t project_smem_bytes(int Q, int ndof, int ncomp_used) {
  return sizeof(double) * ((size_t)ndof * Q + (size_t)kElems * ncomp_used * Q);
}

// out[el][c][k], c < ncomp_used, from the state components comp0 .. comp0 + ncomp_used - 1
cudaError_t launch_project_rows(int nel, int Q, int ndof, int comp0, int ncomp_used, int rowlen, const double* P,
                                const double* wf, double* out, cudaStream_t stream, const int* elems) {
  if (nel <= 0) return cudaSuccess;
  const size_t smem = project_smem_bytes(Q, ndof, ncomp_used);
  cudaError_t e = cudaFuncSetAttribute(project_rows_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  project_rows_kernel<<<(nel + kElems - 1) / kElems, 256, smem, stream>>>(nel, Q, ndof, comp0, ncomp_used, rowlen, P, wf, out, elems);
  return cudaGetLastError();
}

// out[row][n] = sum_q P[n][q] in[row][q] for nrows contiguous rows of Q doubles (ndof <= 64)
}  // namespace ngst
