// GetWave (src/__init__.py:7-43): per-element L2 projection of the quadrature-point state onto polynomials.
//
// The reference assembles an L2 mass matrix and a right-hand side from IntegrationPointFunction
// (src/specialcoefficientfunction.cpp:48-172) and calls a sparse solver. On affine simplices the mass matrix is
// |detJ| times the reference one, so in a basis psi_k that is orthonormal on the reference simplex the projection is
//   coef[el][c][k] = sum_q w_q psi_k(xi_q) v[el][c][q],
// one small matrix P (ndof x Q) applied to every state row: a batched GEMM, HBM-bound (one pass over the rows read).
// One CTA per kElems elements; P is staged in shared memory, a thread owns one (row, k-slice).
#include <cuda_runtime.h>

#include <algorithm>

#include "dmma_tiles.cuh"

namespace ngst {

namespace {
constexpr int kElems = 8;

// ---- out (nrows x ndof) = in (nrows x Q) * P^T, P (ndof x Q): the state transfer of SetInitial (order-2p rule -> compact
// rule, solver.cu ProjectCompact) over ALL rows of the state. A plain GEMM with a tall-skinny left operand, so it runs on
// the FP64 tensor path: mma.sync m8n8k4 (DMMA), one warp per 8 rows, P^T fragments from shared memory (row stride == 4
// mod 16 doubles: conflict-free), the 32 rows of a CTA pass staged through shared memory with coalesced loads.
// NT = n-tiles of 8 columns; K = Q rounded up to 4 (zero padded).
template <int NT>
__global__ void __launch_bounds__(128) transfer_rows_dmma_kernel(long long nrows, int Q, int ndof, const double* __restrict__ P,
                                                                 const double* __restrict__ in, double* __restrict__ out,
                                                                 const int* __restrict__ groups, int per_group) {
  extern __shared__ double sm[];
  const int KP = (Q + 3) / 4 * 4;          // padded inner dimension
  const int LD = KP + ((4 - KP % 16) + 16) % 16;  // row stride == 4 (mod 16) doubles
  double* const Ps = sm;                    // NT * 8 rows (ndof padded) x LD
  double* const Vs = sm + (size_t)NT * 8 * LD;  // 32 rows x LD
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5, g = lane >> 2, k = lane & 3;
  for (int i = tid; i < NT * 8 * LD; i += 128) {
    const int n = i / LD, q = i % LD;
    Ps[i] = (n < ndof && q < Q) ? P[(size_t)n * Q + q] : 0.0;
  }
  for (int i = tid; i < 32 * LD; i += 128) Vs[i] = 0.0;  // the pad columns stay zero
  __syncthreads();
  for (long long r0 = (long long)blockIdx.x * 32; r0 < nrows; r0 += (long long)gridDim.x * 32) {
    const int nr = (int)min((long long)32, nrows - r0);
    if (groups == nullptr) {
      // 32 consecutive rows are one contiguous block of nr * Q doubles
      for (int i = tid; i < nr * Q; i += 128) Vs[(i / Q) * LD + i % Q] = in[r0 * Q + i];
    } else {
      // row r of the pass is row (r % per_group) of the group groups[r / per_group] (an element's component rows)
      for (int i = tid; i < nr * Q; i += 128) {
        const long long r = r0 + i / Q;
        const long long gr = (long long)groups[r / per_group] * per_group + r % per_group;
        Vs[(i / Q) * LD + i % Q] = in[gr * Q + i % Q];
      }
    }
    __syncthreads();
    double acc[NT][2];
#pragma unroll
    for (int t = 0; t < NT; t++) acc[t][0] = acc[t][1] = 0.0;
    const double* __restrict__ va = Vs + (8 * w + g) * LD + k;
    const double* __restrict__ pb = Ps + g * LD + k;
    for (int ks = 0; ks < KP / 4; ks++) {
      const double a = va[4 * ks];
#pragma unroll
      for (int t = 0; t < NT; t++) tiles::dmma_m8n8k4(acc[t][0], acc[t][1], a, pb[(size_t)8 * t * LD + 4 * ks]);
    }
    const int r = 8 * w + g;
    if (r < nr) {
      const long long rr = r0 + r;
      const long long orow = groups ? (long long)groups[rr / per_group] * per_group + rr % per_group : rr;
      double* __restrict__ o = out + orow * ndof;
#pragma unroll
      for (int t = 0; t < NT; t++) {
        const int n = 8 * t + 2 * k;
        if (n < ndof) o[n] = acc[t][0];
        if (n + 1 < ndof) o[n + 1] = acc[t][1];
      }
    }
    __syncthreads();
  }
}

inline int transfer_ld(int Q) {
  const int KP = (Q + 3) / 4 * 4;
  return KP + ((4 - KP % 16) + 16) % 16;
}

__global__ void project_rows_kernel(int nrows, int Q, int ndof, int comp0, int ncomp_used, int rowlen,
                                    const double* __restrict__ P, const double* __restrict__ wf, double* __restrict__ out) {
  extern __shared__ double sm[];
  double* Ps = sm;                       // ndof * Q
  double* vs = sm + (size_t)ndof * Q;    // kElems * ncomp_used * Q
  for (int i = threadIdx.x; i < ndof * Q; i += blockDim.x) Ps[i] = P[i];
  const int el0 = blockIdx.x * kElems;
  const int ne = min(kElems, nrows - el0);
  const int nv = ne * ncomp_used * Q;
  for (int i = threadIdx.x; i < nv; i += blockDim.x) {
    const int e = i / (ncomp_used * Q), r = i % (ncomp_used * Q);
    vs[i] = wf[(size_t)(el0 + e) * rowlen + (size_t)comp0 * Q + r];  // consecutive threads -> consecutive doubles
  }
  __syncthreads();
  const int nout = ne * ncomp_used * ndof;
  for (int i = threadIdx.x; i < nout; i += blockDim.x) {
    const int k = i % ndof, ec = i / ndof;
    const double* v = vs + (size_t)ec * Q;
    const double* p = Ps + (size_t)k * Q;
    double s = 0.0;
    for (int q = 0; q < Q; q++) s = fma(p[q], v[q], s);
    out[(size_t)el0 * ncomp_used * ndof + i] = s;
  }
}
}  // namespace

size_t project_smem_bytes(int Q, int ndof, int ncomp_used) {
  return sizeof(double) * ((size_t)ndof * Q + (size_t)kElems * ncomp_used * Q);
}

// out[el][c][k], c < ncomp_used, from the state components comp0 .. comp0 + ncomp_used - 1
cudaError_t launch_project_rows(int nel, int Q, int ndof, int comp0, int ncomp_used, int rowlen, const double* P,
                                const double* wf, double* out, cudaStream_t stream) {
  if (nel <= 0) return cudaSuccess;
  const size_t smem = project_smem_bytes(Q, ndof, ncomp_used);
  cudaError_t e = cudaFuncSetAttribute(project_rows_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  project_rows_kernel<<<(nel + kElems - 1) / kElems, 256, smem, stream>>>(nel, Q, ndof, comp0, ncomp_used, rowlen, P, wf, out);
  return cudaGetLastError();
}

// out[row][n] = sum_q P[n][q] in[row][q] for nrows contiguous rows of Q doubles (ndof <= 64)
// groups != nullptr: only the rows of the listed groups (ngroups groups of per_group consecutive rows each; in / out are
// indexed by the group's own number) -- the streamed input transfers the rows of the elements it has just fetched
cudaError_t launch_transfer_rows(long long nrows, int Q, int ndof, const double* P, const double* in, double* out,
                                 cudaStream_t stream, const int* groups, int per_group) {
  if (nrows <= 0) return cudaSuccess;
  const int NT = (ndof + 7) / 8;
  if (NT < 1 || NT > 8) return cudaErrorInvalidValue;
  const int LD = transfer_ld(Q);
  const size_t smem = sizeof(double) * ((size_t)NT * 8 * LD + 32 * (size_t)LD);
  if (smem > 200 * 1024) return cudaErrorInvalidValue;
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const long long want = (nrows + 31) / 32;
  const int grid = (int)std::min<long long>(want, (long long)sms * 2);
  cudaError_t e = cudaSuccess;
#define NGST_TR(nt)                                                                                                       \
  case nt:                                                                                                                \
    e = cudaFuncSetAttribute(transfer_rows_dmma_kernel<nt>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);      \
    if (e != cudaSuccess) return e;                                                                                       \
    transfer_rows_dmma_kernel<nt><<<grid, 128, smem, stream>>>(nrows, Q, ndof, P, in, out, groups, per_group);            \
    break;
  switch (NT) {
    NGST_TR(1) NGST_TR(2) NGST_TR(3) NGST_TR(4) NGST_TR(5) NGST_TR(6) NGST_TR(7) NGST_TR(8)
  }
#undef NGST_TR
  return cudaGetLastError();
}

}  // namespace ngst
