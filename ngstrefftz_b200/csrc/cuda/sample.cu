// Device-side sampling of analytic initial / boundary data.
// The reference samples a CoefficientFunction on the host: MakeWavefront (src/twavetents.cpp:734-764) at the mapped
// quadrature points of every element, the boundary datum at the points of every time-like boundary face at the
// physical time t + timeshift (src/twavetents.cpp:355-374, 386, 407-411). For a small library of closed-form data the
// same samples are produced here directly in HBM, which removes the per-slab host evaluation and the H2D copy.
#include <cuda_runtime.h>

#include "sample.hpp"

namespace ngst {

namespace {

constexpr double kPi = 3.14159265358979323846;

// (u, grad u, d_t u) of one data family at (x, t); out[0] = u, out[1..D] = grad u, out[D+1] = d_t u.
template <int D>
__device__ void eval_family(int family, const double* __restrict__ p, const double* x, double t, double* out) {
  if (family == kStandingWave) {
    // u = prod_d sin(pi x_d) sin(pi sqrt(D) c t) / (sqrt(D) pi)   (test/tents.py:83-105)
    const double c = p[0];
    const double sq = sqrt((double)D);
    double s[D], co[D], prod = 1.0;
    for (int d = 0; d < D; d++) {
      sincos(kPi * x[d], &s[d], &co[d]);
      prod *= s[d];
    }
    double st, ct;
    sincos(kPi * t * c * sq, &st, &ct);
    out[0] = prod * st / (sq * kPi);
    for (int d = 0; d < D; d++) {
      double g = 1.0;
      for (int k = 0; k < D; k++) g *= (k == d ? co[k] : s[k]);
      out[1 + d] = g * st / sq;
    }
    out[D + 1] = prod * ct * c;
  } else if (family == kPlaneGauss) {
    // u = exp(-a s^2), s = k.x - c t - s0 with |k| = 1: a travelling pulse, exact for every t
    const double c = p[0], a = p[1], s0 = p[2];
    double s = -c * t - s0;
    for (int d = 0; d < D; d++) s += p[3 + d] * x[d];
    const double u = exp(-a * s * s);
    out[0] = u;
    for (int d = 0; d < D; d++) out[1 + d] = -2.0 * a * s * u * p[3 + d];
    out[D + 1] = 2.0 * a * s * c * u;
  } else {
    // polynomial in (x, t): p = [nterms, (coef, e_0..e_{D-1}, e_t) per term]
    const int nterms = (int)p[0];
    for (int k = 0; k < D + 2; k++) out[k] = 0.0;
    double z[D + 1];
    for (int d = 0; d < D; d++) z[d] = x[d];
    z[D] = t;
    for (int m = 0; m < nterms; m++) {
      const double* term = p + 1 + m * (D + 2);
      const double coef = term[0];
      int e[D + 1];
      double pw[D + 1], dpw[D + 1];  // z^e and e z^(e-1)
      for (int d = 0; d <= D; d++) {
        e[d] = (int)term[1 + d];
        double a = 1.0, b = 0.0;  // a = z^j, b = j z^(j-1)
        for (int j = 0; j < e[d]; j++) {
          b = a * (double)(j + 1);
          a *= z[d];
        }
        pw[d] = a;
        dpw[d] = b;
      }
      double full = coef;
      for (int d = 0; d <= D; d++) full *= pw[d];
      out[0] += full;
      for (int d = 0; d <= D; d++) {
        double g = coef * dpw[d];
        for (int k = 0; k <= D; k++)
          if (k != d) g *= pw[k];
        out[1 + d] += g;
      }
    }
  }
}

// One thread per (element, point). Row layout [comp][q] with comp = (u, grad u, d_t u) or, first order, (grad u, d_t u).
template <int D>
__global__ void sample_wavefront_kernel(int family, const double* __restrict__ params, int nel, int Q, int first_order,
                                        double time, const double* __restrict__ pts, const int* __restrict__ el2v,
                                        const double* __restrict__ lam, double* __restrict__ wf) {
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (long long)nel * Q) return;
  const int el = (int)(idx / Q), q = (int)(idx % Q);
  constexpr int d1 = D + 1;
  double x[D];
  // x = sum_k lambda_k P_k, summed in vertex order without contraction (same bits as Solver::Quadrature)
  for (int d = 0; d < D; d++) {
    double v = 0.0;
    for (int k = 0; k < d1; k++) v = __dadd_rn(v, __dmul_rn(lam[q * d1 + k], pts[(size_t)el2v[(size_t)el * d1 + k] * D + d]));
    x[d] = v;
  }
  double val[D + 2];
  eval_family<D>(family, params, x, time, val);
  const int ncomp = first_order ? D + 1 : D + 2;
  double* row = wf + (size_t)el * ncomp * Q + q;
  for (int cc = 0; cc < ncomp; cc++) row[(size_t)cc * Q] = val[cc + (first_order ? 1 : 0)];
}

// One thread per (tent face, point); faces without datum (material interfaces) have tf_data < 0.
// Dirichlet faces take Q values of d_t u, Neumann faces D*Q values of grad u laid out [component][q].
template <int D>
__global__ void sample_boundary_kernel(int family, const double* __restrict__ params, int ntf, int Q, double timeshift,
                                       const int* __restrict__ tf_kind, const double* __restrict__ tf_verts,
                                       const long long* __restrict__ tf_data, const double* __restrict__ lam,
                                       int neumann_kind, double* __restrict__ bnd) {
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (long long)ntf * Q) return;
  const int tf = (int)(idx / Q), q = (int)(idx % Q);
  const long long off = tf_data[tf];
  if (off < 0) return;
  constexpr int d1 = D + 1;
  const double* V = tf_verts + (size_t)tf * d1 * d1;
  double xt[d1];
  // affine map of the D-simplex rule onto the face (twavetents.cpp:355-365), same operation order as
  // boundary_sample_points (host/stage.cpp)
  for (int r = 0; r < d1; r++) {
    double x = V[r];
    for (int n = 0; n < D; n++) x = __dadd_rn(x, __dmul_rn(__dadd_rn(V[(n + 1) * d1 + r], -V[r]), lam[q * d1 + n]));
    xt[r] = x;
  }
  double val[D + 2];
  eval_family<D>(family, params, xt, __dadd_rn(xt[D], timeshift), val);
  if (tf_kind[tf] == neumann_kind) {
    for (int d = 0; d < D; d++) bnd[off + (long long)d * Q + q] = val[1 + d];
  } else {
    bnd[off + q] = val[D + 1];
  }
}

}  // namespace

cudaError_t launch_sample_wavefront(int D, int family, const double* d_params, int nel, int Q, int first_order,
                                    double time, const double* pts, const int* el2v, const double* lam, double* wf,
                                    cudaStream_t stream) {
  const long long n = (long long)nel * Q;
  if (n <= 0) return cudaSuccess;
  const int threads = 256;
  const unsigned blocks = (unsigned)((n + threads - 1) / threads);
  if (D == 1) sample_wavefront_kernel<1><<<blocks, threads, 0, stream>>>(family, d_params, nel, Q, first_order, time, pts, el2v, lam, wf);
  else if (D == 2) sample_wavefront_kernel<2><<<blocks, threads, 0, stream>>>(family, d_params, nel, Q, first_order, time, pts, el2v, lam, wf);
  else sample_wavefront_kernel<3><<<blocks, threads, 0, stream>>>(family, d_params, nel, Q, first_order, time, pts, el2v, lam, wf);
  return cudaGetLastError();
}

cudaError_t launch_sample_boundary(int D, int family, const double* d_params, int ntf, int Q, double timeshift,
                                   const int* tf_kind, const double* tf_verts, const long long* tf_data,
                                   const double* lam, int neumann_kind, double* bnd, cudaStream_t stream) {
  const long long n = (long long)ntf * Q;
  if (n <= 0) return cudaSuccess;
  const int threads = 256;
  const unsigned blocks = (unsigned)((n + threads - 1) / threads);
  if (D == 1) sample_boundary_kernel<1><<<blocks, threads, 0, stream>>>(family, d_params, ntf, Q, timeshift, tf_kind, tf_verts, tf_data, lam, neumann_kind, bnd);
  else if (D == 2) sample_boundary_kernel<2><<<blocks, threads, 0, stream>>>(family, d_params, ntf, Q, timeshift, tf_kind, tf_verts, tf_data, lam, neumann_kind, bnd);
  else sample_boundary_kernel<3><<<blocks, threads, 0, stream>>>(family, d_params, ntf, Q, timeshift, tf_kind, tf_verts, tf_data, lam, neumann_kind, bnd);
  return cudaGetLastError();
}

}  // namespace ngst
