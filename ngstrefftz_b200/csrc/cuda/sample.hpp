// Analytic data families sampled on the device (cuda/sample.cu).
#pragma once
#include <cuda_runtime.h>

namespace ngst {

// (u, grad u, d_t u) in closed form. Parameters (doubles):
//   kStandingWave: [c]                              u = prod_d sin(pi x_d) sin(pi sqrt(D) c t) / (sqrt(D) pi)  (test/tents.py:83-105)
//   kPlaneGauss:   [c, a, s0, k_0 .. k_{D-1}]       u = exp(-a (k.x - c t - s0)^2), |k| = 1
//   kPolynomial:   [nterms, (coef, e_0 .. e_{D-1}, e_t) x nterms]   u = sum coef x^e t^e_t
enum AnalyticFamily { kStandingWave = 0, kPlaneGauss = 1, kPolynomial = 2 };

cudaError_t launch_sample_wavefront(int D, int family, const double* d_params, int nel, int Q, int first_order,
                                    double time, const double* pts, const int* el2v, const double* lam, double* wf,
                                    cudaStream_t stream);
cudaError_t launch_sample_boundary(int D, int family, const double* d_params, int ntf, int Q, double timeshift,
                                   const int* tf_kind, const double* tf_verts, const long long* tf_data,
                                   const double* lam, int neumann_kind, double* bnd, cudaStream_t stream);

}  // namespace ngst
