// Fast quasi-Trefftz tent kernel (QTWaveTents<2>::Propagate, src/twavetents.cpp:967-1123), sm_100a.
//
// Same mathematics as tent_qt.cu, re-organised so that the dense work runs on the FP64 tensor path:
//   * every quasi-Trefftz function of the tent is phi_i = sum_j M[j][i] m_j(xi) with the monomials m_j of degree <= p in
//     the scaled tent coordinates xi = (x - x_v, t - t_0) / h (QTWaveBasis<3>::Basis, src/trefftzfespace.cpp:1053-1198,
//     builds M per tent). The element matrix and vector are therefore A = M^T At M, b = M^T bt with At, bt assembled in the
//     MONOMIAL basis: the kernel never applies the dense M at a quadrature point (the generic kernel spends most of its
//     time there), monomials and their derivatives are straight-line code with compile-time exponents, and M is applied
//     once per tent;
//   * At collects (i) the top-face flux Gram B W B^T (twavetents.cpp:1001-1016 with the wavespeed at every point) as the
//     symmetric product Z Z^T, Z = B L, W = L L^T; (ii) the bottom-face stabilisation Gram; (iii) the cone volume terms
//     sum_q vol w box(m_i) (d_t m_j + mu box(m_j)) (1032-1108; wave operator of
//     ScalarMappedElement::CalcDDWaveOperator, src/scalarmappedfe.cpp:919-1008) and (iv) the boundary faces
//     (CalcTentBndEl, 335-434) as rank-1 updates l r^T per point -- all with mma.sync m8n8k4 f64 (DMMA), accumulators in
//     registers (tile helpers: dmma_tiles.cuh);
//   * one CTA of 4 warps per tent; lane = quadrature point, warp = one quarter of the monomials;
//   * in-register Gauss-Jordan solve with partial pivoting (tent_solve.cuh); evaluation of the solution on the top faces
//     in the monomial basis (CalcTentElEval, 541-590).
// State-independent per-face / per-cone data (scaled vertices, normals, cone maps, max(1/c^2) per cone, 1089-1094) is
// computed once per slab by qt_fast_stage_kernel. Tents whose top faces are not strictly causal at some quadrature point
// stay on the generic kernel (tent_qt.cu).
#include <cuda_runtime.h>

#include <type_traits>
#include <utility>

#include "dmma_tiles.cuh"
#include "tent_kernel.cuh"
#include "tent_qt.cuh"
#include "tent_solve.cuh"

namespace ngst {

namespace {

using namespace tiles;

constexpr int kFE = 20;  // per tent-element record (doubles): XV[6] TB[3] TT[3] nbot[3] abot ntop[3] atop
constexpr int kCS = 20;  // per cone record: S[9] (xi = S vxi), Mx[6] (x - x_v = Mx vxi), vol, h/cmax, x_v[2], pad
constexpr int kFF = 20;  // per boundary-face record: V0s[3] Es[6] (scaled), x0[2] ex[4] (physical), n[2], area, kind

template <int P>
struct QCfg {
  static constexpr int d1 = 3;
  static constexpr int NM = (P + 1) * (P + 2) * (P + 3) / 6;  // monomials of degree <= P in (x, y, t)
  static constexpr int N = (P + 1) * (P + 1);                 // quasi-Trefftz functions (second-order form)
  static constexpr int NT = (NM + 7) / 8, NP = NT * 8;        // 8x8 tiles / padded monomial count
  static constexpr int NB = (N + 7) / 8 * 8;                  // padded function count
  static constexpr int G = NP / 4;                            // monomial rows per warp (row j -> warp j % 4)
  static constexpr int Q = (P + 1) * (P + 1);                 // points of the order-2P triangle rule
  static constexpr int ZS = kChunk * d1 + 4;                  // row stride of the symmetric operand (== 4 mod 16)
  static constexpr int LDM = NP + 1;
  static constexpr int LDA = N + 1;
  static constexpr int NSYM = NT * (NT + 1) / 2, NACC = (NSYM + 1) / 2;
  // shared memory (doubles): Z (operands; later At | T | A) | M | bt | bv | cm | AAs | BBs
  static constexpr int oZ = 0;
  static constexpr int oAt = 0, oT = NP * LDM, oA = 0;
  static constexpr int oM = NP * ZS;
  static constexpr int oBt = oM + NM * NB;
  static constexpr int oBv = oBt + NP;
  static constexpr int oCm = oBv + NB;
  static constexpr int oAA = oCm + NP;
  static constexpr int oBB = oAA + (P - 1) * (P - 1);
  // per-point setup of the volume pass for one super-chunk of 128 points (thread = point): xi[3], 1/c^2, L and R weights
  static constexpr int oSP = (oBB + P * P + 1) / 2 * 2;
  static constexpr int nsmem = oSP + 6 * 4 * kChunk;
  static_assert(2 * kBS <= ZS, "general operands must fit in the Z buffer");
  static_assert(NP * LDM + NM * NB <= NP * ZS, "At and T must fit in the Z buffer");
  static_assert(N * LDA <= NP * LDM, "A must fit where At was");
  static_assert(NT >= 2 && (NT * NT + 3) / 4 <= NACC, "accumulator budget");
  static_assert(N <= 64, "solver row mask");
};

// exponents of monomial j, first coordinate outermost (src/scalarmappedfe.cpp:526-531, PolBasis::IndexMap2)
template <int P>
struct Mono {
  static constexpr int NM = QCfg<P>::NM;
  static constexpr int expo(int j, int which) {
    int c = 0;
    for (int ex = 0; ex <= P; ex++)
      for (int ey = 0; ey <= P - ex; ey++)
        for (int et = 0; et <= P - ex - ey; et++) {
          if (c == j) return which == 0 ? ex : (which == 1 ? ey : et);
          c++;
        }
    return 0;
  }
};

template <int P, int W, class F, int... Js>
__device__ __forceinline__ void mono_rows_impl(F&& f, std::integer_sequence<int, Js...>) {
  (f(std::integral_constant<int, 4 * Js + W>{}), ...);
}
// f(jc) for every monomial row j = jc() with j % 4 == W and j < NM
template <int P, int W, class F>
__device__ __forceinline__ void mono_rows(F&& f) {
  mono_rows_impl<P, W>(f, std::make_integer_sequence<int, (Mono<P>::NM - W + 3) / 4>{});
}
template <int P, class F>
__device__ __forceinline__ void mono_rows_w(int w, F&& f) {
  switch (w) {
    case 0: mono_rows<P, 0>(f); break;
    case 1: mono_rows<P, 1>(f); break;
    case 2: mono_rows<P, 2>(f); break;
    default: mono_rows<P, 3>(f); break;
  }
}

template <int P>
struct Pw {
  double x[P + 1], y[P + 1], t[P + 1];
  __device__ __forceinline__ void set(const double* xi) {
    x[0] = y[0] = t[0] = 1.0;
#pragma unroll
    for (int e = 1; e <= P; e++) { x[e] = x[e - 1] * xi[0]; y[e] = y[e - 1] * xi[1]; t[e] = t[e - 1] * xi[2]; }
  }
};

// value and first derivatives (w.r.t. xi) of monomial J
template <int P, int J>
__device__ __forceinline__ void mono_d1(const Pw<P>& p, double& v, double& dx, double& dy, double& dt) {
  constexpr int a = Mono<P>::expo(J, 0), b = Mono<P>::expo(J, 1), c = Mono<P>::expo(J, 2);
  const double yt = p.y[b] * p.t[c];
  v = p.x[a] * yt;
  dx = a > 0 ? (double)a * p.x[a > 0 ? a - 1 : 0] * yt : 0.0;
  dy = b > 0 ? (double)b * p.x[a] * p.y[b > 0 ? b - 1 : 0] * p.t[c] : 0.0;
  dt = c > 0 ? (double)c * p.x[a] * p.y[b] * p.t[c > 0 ? c - 1 : 0] : 0.0;
}

// d_t and the wave operator (Laplace - inv d_tt) of monomial J
template <int P, int J>
__device__ __forceinline__ void mono_box(const Pw<P>& p, double inv, double& dt, double& box) {
  constexpr int a = Mono<P>::expo(J, 0), b = Mono<P>::expo(J, 1), c = Mono<P>::expo(J, 2);
  dt = c > 0 ? (double)c * p.x[a] * p.y[b] * p.t[c > 0 ? c - 1 : 0] : 0.0;
  double s = 0.0;
  if (a > 1) s = (double)(a * (a - 1)) * p.x[a > 1 ? a - 2 : 0] * p.y[b] * p.t[c];
  if (b > 1) s = fma((double)(b * (b - 1)) * p.x[a], p.y[b > 1 ? b - 2 : 0] * p.t[c], s);
  if (c > 1) s = fma(-inv * (double)(c * (c - 1)) * p.x[a], p.y[b] * p.t[c > 1 ? c - 2 : 0], s);
  box = s;
}

__device__ __forceinline__ double qt_poly_eval(const QtArgs& q, const double* x) {
  double s = 0.0;
  for (int j = 0; j < q.ncterm; j++) {
    double m = q.ccoef[j];
    for (int d = 0; d < 2; d++)
      for (int k = 0; k < q.cexpo[j * 2 + d]; k++) m *= x[d];
    s += m;
  }
  return s;
}

// ---------------------------------------------------------------------------------------------- per-slab staging
// One thread per tent: records of its tent-elements, cones and boundary faces; ok[tent] = every top face is strictly
// causal at every quadrature point (n_t > 0 and n_t / c^2 - |n_x|^2 / n_t > 0).
__global__ void qt_fast_stage_kernel(TentKernelArgs a, QtArgs qa, int ntents, double* fe, double* cone, double* ff, int* ok) {
  constexpr int D = 2, d1 = 3;
  const int tent = blockIdx.x * blockDim.x + threadIdx.x;
  if (tent >= ntents) return;
  const double h = qa.tentsize[tent], sc = 1.0 / h;
  double ctr[d1];
  for (int d = 0; d < d1; d++) ctr[d] = a.center[(size_t)tent * d1 + d];
  bool good = true;
  for (int te = a.te_off[tent]; te < a.te_off[tent + 1]; te++) {
    const int el = a.te_el[te];
    double Pv[d1][D];
    for (int k = 0; k < d1; k++) {
      const int vk = a.el2v[(size_t)el * d1 + k];
      for (int d = 0; d < D; d++) Pv[k][d] = a.pts[(size_t)vk * D + d];
    }
    double* r = fe + (size_t)te * kFE;
    for (int k = 0; k < d1; k++) {
      for (int d = 0; d < D; d++) r[k * D + d] = (Pv[k][d] - ctr[d]) * sc;
      r[6 + k] = (a.te_tbot[(size_t)te * d1 + k] - ctr[D]) * sc;
      r[9 + k] = (a.te_ttop[(size_t)te * d1 + k] - ctr[D]) * sc;
    }
    const double* geo = a.te_geo + (size_t)te * 2 * (d1 + 1);
    for (int k = 0; k < d1 + 1; k++) { r[12 + k] = geo[k]; r[16 + k] = geo[d1 + 1 + k]; }
    const double* nt = geo + d1 + 1;
    for (int q = 0; q < a.Q; q++) {
      const double c = qa.c_point[(size_t)el * a.Q + q];
      if (!(nt[D] > 0.0) || !(nt[D] / (c * c) - (nt[0] * nt[0] + nt[1] * nt[1]) / nt[D] > 0.0)) good = false;
    }
    for (int part = 0; part < 2; part++) {
      const double* tau = (part == 0 ? a.te_tbot : a.te_ttop) + (size_t)te * d1;
      double M[d1][d1];
      for (int k = 0; k < d1; k++) {
        for (int d = 0; d < D; d++) M[d][k] = Pv[k][d] - ctr[d];
        M[D][k] = tau[k] - ctr[D];
      }
      const double det = M[0][0] * (M[1][1] * M[2][2] - M[1][2] * M[2][1]) - M[0][1] * (M[1][0] * M[2][2] - M[1][2] * M[2][0]) +
                         M[0][2] * (M[1][0] * M[2][1] - M[1][1] * M[2][0]);
      const double vol = fabs(det);
      double* c = cone + ((size_t)te * 2 + part) * kCS;
      for (int rr = 0; rr < d1; rr++)
        for (int k = 0; k < d1; k++) c[rr * d1 + k] = sc * M[rr][k];
      for (int d = 0; d < D; d++)
        for (int k = 0; k < d1; k++) c[9 + d * d1 + k] = M[d][k];
      double mx = 0.0;
      if (!(vol < 10e-16)) {  // twavetents.cpp:1046-1047 skips degenerate cones
        for (int q = 0; q < qa.Qv; q++) {
          double x[D];
          for (int d = 0; d < D; d++) {
            double v = ctr[d];
            for (int k = 0; k < d1; k++) v += M[d][k] * qa.vxi[(size_t)q * d1 + k];
            x[d] = v;
          }
          const double cc = qt_poly_eval(qa, x);
          mx = fmax(mx, fabs(1.0 / (cc * cc)));
        }
      }
      c[15] = (vol < 10e-16) ? 0.0 : vol;
      c[16] = (vol < 10e-16) ? 0.0 : h / mx;
      c[17] = ctr[0];
      c[18] = ctr[1];
      c[19] = 0.0;
    }
  }
  for (int tf = a.tf_off[tent]; tf < a.tf_off[tent + 1]; tf++) {
    const double* V = a.tf_verts + (size_t)tf * d1 * d1;
    double* r = ff + (size_t)tf * kFF;
    for (int d = 0; d < d1; d++) r[d] = (V[d] - ctr[d]) * sc;
    for (int m = 0; m < D; m++)
      for (int d = 0; d < d1; d++) r[3 + m * d1 + d] = (V[(m + 1) * d1 + d] - V[d]) * sc;
    for (int d = 0; d < D; d++) r[9 + d] = V[d];
    for (int m = 0; m < D; m++)
      for (int d = 0; d < D; d++) r[11 + m * D + d] = V[(m + 1) * d1 + d] - V[d];
    for (int d = 0; d < D; d++) r[15 + d] = a.tf_n[(size_t)tf * D + d];
    r[17] = a.tf_area[tf];
    r[18] = (double)a.tf_kind[tf];
    r[19] = 0.0;
    if (a.tf_kind[tf] > 1) good = false;  // no interface faces in quasi-Trefftz mode (992-993)
  }
  ok[tent] = good ? 1 : 0;
}

// ---------------------------------------------------------------------------------------------- the tent kernel
template <int P>
__global__ void __launch_bounds__(kThreads, P >= 5 ? 2 : 3) tent_kernel_qt_fast(TentKernelArgs a, QtArgs qa, QtFastArgs f) {
  using C = QCfg<P>;
  constexpr int D = 2, d1 = 3, NM = C::NM, N = C::N, NT = C::NT, NP = C::NP, NB = C::NB, G = C::G, Q = C::Q, ZS = C::ZS;
  constexpr int LDM = C::LDM, LDA = C::LDA, NACC = C::NACC;
  extern __shared__ double sm[];
  double* const Z = sm + C::oZ;
  double* const Mq = sm + C::oM;   // Mq[j * NB + i]: coefficient of monomial j in function i
  double* const bt = sm + C::oBt;
  double* const bv = sm + C::oBv;
  double* const cm = sm + C::oCm;
  double* const AAs = sm + C::oAA;
  double* const BBs = sm + C::oBB;

  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  const int g = lane >> 2, k = lane & 3;
  const int th = w >> 1, kh = w & 1;
  const int tent = a.tents[blockIdx.x];
  const int te0 = a.te_off[tent], nel = a.te_off[tent + 1] - te0;
  const int tf0 = a.tf_off[tent], nfc = a.tf_off[tent + 1] - tf0;
  const int npts = nel * Q;
  const double h = qa.tentsize[tent], sc = 1.0 / h;

  // ------------------------------------------------------------------ quasi-Trefftz basis of this tent -> Mq
  // (src/trefftzfespace.cpp:1076-1185; Taylor coefficients of 1/c^2 and B at the tent vertex scaled by h^|alpha|)
  {
    const int v = qa.tent_vertex[tent];
    for (int idx = tid; idx < (P - 1) * (P - 1); idx += kThreads) {
      const int nx = idx / (P - 1), ny = idx % (P - 1);
      const int j = qa.idxA[idx];
      AAs[idx] = j >= 0 ? qa.aa[(size_t)v * qa.nA + j] * pow(h, (double)(nx + ny)) : 0.0;
    }
    for (int idx = tid; idx < P * P; idx += kThreads) {
      const int nx = idx / P, ny = idx % P;
      const int j = qa.idxB[idx];
      BBs[idx] = j >= 0 ? qa.bb[(size_t)v * qa.nB + j] * pow(h, (double)(nx + ny)) : 0.0;
    }
    for (int idx = tid; idx < NM * NB; idx += kThreads) Mq[idx] = 0.0;
    for (int i = tid; i < NP * ZS; i += kThreads) Z[i] = 0.0;  // rows of the padding monomials stay zero
    __syncthreads();
    constexpr int P1 = P + 1;
#define IDX3(x, y, t) qa.idx3[((x) * P1 + (y)) * P1 + (t)]
#define QB(j) Mq[(size_t)(j) * NB + b]
#define AAF(x, y) AAs[(x) * (P - 1) + (y)]
#define BBF(x, y) BBs[(x) * P + (y)]
    // seeds: function b <-> the b-th monomial (t-exponent 0 or 1) in TraversePol order (t outermost)
    for (int b = tid; b < N; b += kThreads) {
      int cnt = 0;
      for (int t = 0; t < 2; t++)
        for (int x = 0; x <= P - t; x++)
          for (int y = 0; y <= P - x - t; y++) {
            if (cnt == b) QB(IDX3(x, y, t)) = 1.0;
            cnt++;
          }
    }
    __syncthreads();
    // recursion over the total degree ell of (x, y, t): the coefficient of (x, y, t + 2) needs coefficients of lower ell
    // and, at the same ell, those with t-exponent t (computed from t - 2). Rows t = tp, tp + 1 of one ell are therefore
    // independent of each other: one thread per (function, t, x) item, a barrier per row pair.
    {
      const double A0 = AAF(0, 0);
      for (int ell = 0; ell < P - 1; ell++)
        for (int tp = 0; tp <= ell; tp += 2) {
        const int n0 = ell - tp + 1, n1 = (tp + 1 <= ell) ? ell - tp : 0;
        for (int it = tid; it < N * (n0 + n1); it += kThreads) {
          const int b = it % N;
          const int r = it / N;
          const int t = tp + (r >= n0 ? 1 : 0);
          const int x = r - (r >= n0 ? n0 : 0), y = ell - t - x;
          const double den = (double)((t + 2) * (t + 1)) * A0;
          double nc = 0.0;
          for (int bx = 0; bx <= x; bx++)
            for (int by = 0; by <= y; by++) {
              nc += (double)((bx + 2) * (bx + 1)) / den * BBF(x - bx, y - by) * QB(IDX3(bx + 2, by, t)) +
                    (double)((x - bx + 1) * (bx + 1)) / den * BBF(x - bx + 1, y - by) * QB(IDX3(bx + 1, by, t));
              nc += (double)((by + 2) * (by + 1)) / den * BBF(x - bx, y - by) * QB(IDX3(bx, by + 2, t)) +
                    (double)((y - by + 1) * (by + 1)) / den * BBF(x - bx, y - by + 1) * QB(IDX3(bx, by + 1, t));
              if (bx + by == x + y) continue;
              nc -= AAF(x - bx, y - by) * QB(IDX3(bx, by, t + 2)) / A0;
            }
          QB(IDX3(x, y, t + 2)) = nc;
        }
        __syncthreads();
      }
    }
#undef IDX3
#undef QB
#undef AAF
#undef BBF
  }
  double accs[NACC][2], accf[NACC][2];
#pragma unroll
  for (int i = 0; i < NACC; i++) accs[i][0] = accs[i][1] = accf[i][0] = accf[i][1] = 0.0;
  double bpart[G];
#pragma unroll
  for (int j = 0; j < G; j++) bpart[j] = 0.0;
  __syncthreads();

  // ------------------------------------------------------------------ top faces: At += B W B^T, wavespeed per point
  for (int c0 = 0; c0 < npts; c0 += kChunk) {
    const int idx = c0 + lane;
    const bool valid = idx < npts;
    const int idc = valid ? idx : npts - 1;
    const int te = te0 + idc / Q, q = idc % Q;
    const double* __restrict__ rec = f.fe + (size_t)te * kFE;
    const double* __restrict__ lm = a.lam + q * d1;
    double xi[d1] = {0.0, 0.0, 0.0};
#pragma unroll
    for (int kk = 0; kk < d1; kk++) {
      const double l = lm[kk];
      xi[0] = fma(l, rec[kk * D], xi[0]);
      xi[1] = fma(l, rec[kk * D + 1], xi[1]);
      xi[2] = fma(l, rec[9 + kk], xi[2]);
    }
    const double c = qa.c_point[(size_t)a.te_el[te] * Q + q];
    const double n0 = rec[16], n1 = rec[17], n2 = rec[18];
    const double so = valid ? sc * sqrt(a.w[q] * rec[19]) : 0.0;
    const double sqn = sqrt(n2);
    const double k0 = so * sqn, kd0 = so * n0 / sqn, kd1 = so * n1 / sqn;
    const double kt = so * sqrt(n2 / (c * c) - (n0 * n0 + n1 * n1) / n2);
    Pw<P> pw;
    pw.set(xi);
    double* const zl = Z + lane;
    mono_rows_w<P>(w, [&](auto jc) {
      constexpr int j = decltype(jc)::value;
      double v, dx, dy, dt;
      mono_d1<P, j>(pw, v, dx, dy, dt);
      double* z = zl + j * ZS;
      z[0] = k0 * dx - kd0 * dt;
      z[kChunk] = k0 * dy - kd1 * dt;
      z[2 * kChunk] = kt * dt;
    });
    __syncthreads();
    if (th == 0) dmma_sym<NT, NACC, 0>(accs, Z, ZS, d1 * (kChunk / 4), kh, g, k);
    else dmma_sym<NT, NACC, 1>(accs, Z, ZS, d1 * (kChunk / 4), kh, g, k);
    __syncthreads();
  }

  // ------------------------------------------------------------------ bottom faces: rhs and stabilisation Gram
  for (int c0 = 0; c0 < npts; c0 += kChunk) {
    const int idx = c0 + lane;
    const bool valid = idx < npts;
    const int idc = valid ? idx : npts - 1;
    const int te = te0 + idc / Q, q = idc % Q;
    const int el = a.te_el[te];
    const double* __restrict__ rec = f.fe + (size_t)te * kFE;
    const double* __restrict__ lm = a.lam + q * d1;
    double xi[d1] = {0.0, 0.0, 0.0};
#pragma unroll
    for (int kk = 0; kk < d1; kk++) {
      const double l = lm[kk];
      xi[0] = fma(l, rec[kk * D], xi[0]);
      xi[1] = fma(l, rec[kk * D + 1], xi[1]);
      xi[2] = fma(l, rec[6 + kk], xi[2]);
    }
    const double c = qa.c_point[(size_t)el * Q + q];
    const double n0 = rec[12], n1 = rec[13], n2 = rec[14];
    const double om = valid ? a.w[q] * rec[15] : 0.0;
    const double* __restrict__ wf = a.wavefront + (size_t)el * (4 * Q) + q;
    const double gu = wf[0], gx0 = wf[Q], gx1 = wf[2 * Q], gt = wf[3 * Q];
    const double r0 = om * gu;
    const double r1 = -sc * (n2 * om * gx0 - n0 * om * gt);
    const double r2 = -sc * (n2 * om * gx1 - n1 * om * gt);
    const double r3 = -sc * (n2 / (c * c) * om * gt - n0 * om * gx0 - n1 * om * gx1);
    const double sqo = sqrt(om);
    Pw<P> pw;
    pw.set(xi);
    double* const zl = Z + lane;
    mono_rows_w<P>(w, [&](auto jc) {
      constexpr int j = decltype(jc)::value;
      double v, dx, dy, dt;
      mono_d1<P, j>(pw, v, dx, dy, dt);
      bpart[j >> 2] += fma(v, r0, fma(dx, r1, fma(dy, r2, dt * r3)));
      zl[j * ZS] = sqo * v;
    });
    __syncthreads();
    if (th == 0) dmma_sym<NT, NACC, 0>(accs, Z, ZS, kChunk / 4, kh, g, k);
    else dmma_sym<NT, NACC, 1>(accs, Z, ZS, kChunk / 4, kh, g, k);
    __syncthreads();
  }

  // ------------------------------------------------------------------ cone volume terms: At += sum L R^T  (1032-1108)
  // L = vol w box(m), R = d_t m + mu box(m), mu = h / max_cone(1/c^2) * c^2 at the point
  double* const ZL = Z;
  double* const ZR = Z + NP * kBS;
  for (int i = tid; i < 2 * NP * kBS; i += kThreads) Z[i] = 0.0;
  __syncthreads();
  {
    const int Qv = qa.Qv, nvp = nel * 2 * Qv;
    constexpr int kSuper = 4 * kChunk;
    double* const sp = sm + C::oSP;
    for (int s0 = 0; s0 < nvp; s0 += kSuper) {
      {  // setup, thread = point: the wavespeed polynomial, two divisions and the cone map are evaluated once per point
        const int idx = s0 + tid;
        const bool valid = idx < nvp;
        const int idc = valid ? idx : nvp - 1;
        const int cn = idc / Qv, q = idc % Qv;
        const double* __restrict__ rec = f.cone + ((size_t)te0 * 2 + cn) * kCS;
        const double* __restrict__ vx = qa.vxi + (size_t)q * d1;
        const double v0 = vx[0], v1 = vx[1], v2 = vx[2];
        double x[D];
#pragma unroll
        for (int r = 0; r < d1; r++) sp[r * kSuper + tid] = rec[r * d1] * v0 + rec[r * d1 + 1] * v1 + rec[r * d1 + 2] * v2;
#pragma unroll
        for (int d = 0; d < D; d++) x[d] = rec[17 + d] + rec[9 + d * d1] * v0 + rec[9 + d * d1 + 1] * v1 + rec[9 + d * d1 + 2] * v2;
        const double cc = qt_poly_eval(qa, x);
        const double inv = 1.0 / (cc * cc);
        sp[3 * kSuper + tid] = inv;
        sp[4 * kSuper + tid] = valid ? rec[15] * qa.vw[q] * sc * sc : 0.0;
        sp[5 * kSuper + tid] = rec[16] / inv * sc * sc;
      }
      __syncthreads();
      for (int sub = 0; sub < kSuper / kChunk && s0 + sub * kChunk < nvp; sub++) {
        const double* __restrict__ spl = sp + sub * kChunk + lane;
        const double xi[d1] = {spl[0], spl[kSuper], spl[2 * kSuper]};
        const double inv = spl[3 * kSuper], lw = spl[4 * kSuper], mu = spl[5 * kSuper];
        Pw<P> pw;
        pw.set(xi);
        double* const zl = ZL + lane;
        double* const zr = ZR + lane;
        mono_rows_w<P>(w, [&](auto jc) {
          constexpr int j = decltype(jc)::value;
          double dt, box;
          mono_box<P, j>(pw, inv, dt, box);
          zl[j * kBS] = lw * box;
          zr[j * kBS] = fma(mu, box, sc * dt);
        });
        __syncthreads();
        switch (w) {
          case 0: dmma_full<NT, NACC, 0>(accf, ZL, ZR, g, k); break;
          case 1: dmma_full<NT, NACC, 1>(accf, ZL, ZR, g, k); break;
          case 2: dmma_full<NT, NACC, 2>(accf, ZL, ZR, g, k); break;
          default: dmma_full<NT, NACC, 3>(accf, ZL, ZR, g, k); break;
        }
        __syncthreads();
      }
    }
  }

  // ------------------------------------------------------------------ boundary faces (CalcTentBndEl, 335-434)
  if (nfc > 0) {
    const int nbp = nfc * Q;
    for (int c0 = 0; c0 < nbp; c0 += kChunk) {
      const int idx = c0 + lane;
      const bool valid = idx < nbp;
      const int idc = valid ? idx : nbp - 1;
      const int tf = tf0 + idc / Q, q = idc % Q;
      const double* __restrict__ rec = f.ff + (size_t)tf * kFF;
      const double l0 = a.lam[q * d1], l1 = a.lam[q * d1 + 1];
      double xi[d1], x[D];
#pragma unroll
      for (int r = 0; r < d1; r++) xi[r] = rec[r] + l0 * rec[3 + r] + l1 * rec[6 + r];
#pragma unroll
      for (int d = 0; d < D; d++) x[d] = rec[9 + d] + l0 * rec[11 + d] + l1 * rec[13 + d];
      const double cc = qt_poly_eval(qa, x);
      const double inv = 1.0 / (cc * cc);
      const double n0 = rec[15], n1 = rec[16];
      const bool neumann = rec[18] != 0.0;
      const double om = valid ? a.w[q] * rec[17] : 0.0;
      double gdat = 0.0;
      if (a.bnd_data) {
        const long long off = a.tf_data[tf];
        if (neumann) gdat = n0 * a.bnd_data[off + q] + n1 * a.bnd_data[off + Q + q];
        else gdat = a.bnd_data[off + q];
      }
      const double og = om * gdat;
      Pw<P> pw;
      pw.set(xi);
      double* const zl = ZL + lane;
      double* const zr = ZR + lane;
      mono_rows_w<P>(w, [&](auto jc) {
        constexpr int j = decltype(jc)::value;
        double v, dx, dy, dt;
        mono_d1<P, j>(pw, v, dx, dy, dt);
        const double p = sc * dt;
        const double nd = sc * (n0 * dx + n1 * dy);
        double lv, rv, bc;
        if (neumann) { lv = -nd; rv = om * (p - 0.5 * nd); bc = og * (0.5 * nd + p); }
        else { lv = p; rv = om * (0.5 * inv * p - nd); bc = og * (0.5 * inv * p + nd); }
        zl[j * kBS] = lv;
        zr[j * kBS] = rv;
        bpart[j >> 2] += bc;
      });
      __syncthreads();
      switch (w) {
        case 0: dmma_full<NT, NACC, 0>(accf, ZL, ZR, g, k); break;
        case 1: dmma_full<NT, NACC, 1>(accf, ZL, ZR, g, k); break;
        case 2: dmma_full<NT, NACC, 2>(accf, ZL, ZR, g, k); break;
        default: dmma_full<NT, NACC, 3>(accf, ZL, ZR, g, k); break;
      }
      __syncthreads();
    }
  }

  // ------------------------------------------------------------------ accumulators -> At (fixed order: deterministic)
  double* const At = sm + C::oAt;
  if (kh == 0) {
    if (th == 0) flush_sym<NT, NACC, 0>(accs, At, LDM, false, g, k);
    else flush_sym<NT, NACC, 1>(accs, At, LDM, false, g, k);
  }
  __syncthreads();
  if (kh == 1) {
    if (th == 0) flush_sym<NT, NACC, 0>(accs, At, LDM, true, g, k);
    else flush_sym<NT, NACC, 1>(accs, At, LDM, true, g, k);
  }
  __syncthreads();
  for (int idx = tid; idx < NP * NP; idx += kThreads) {
    const int i = idx / NP, j = idx % NP;
    if (j < i && (j >> 3) != (i >> 3)) At[i * LDM + j] = At[j * LDM + i];
  }
  __syncthreads();
  switch (w) {
    case 0: flush_full<NT, NACC, 0>(accf, At, LDM, g, k); break;
    case 1: flush_full<NT, NACC, 1>(accf, At, LDM, g, k); break;
    case 2: flush_full<NT, NACC, 2>(accf, At, LDM, g, k); break;
    default: flush_full<NT, NACC, 3>(accf, At, LDM, g, k); break;
  }
#pragma unroll
  for (int j = 0; j < G; j++) {
    double v = bpart[j];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane == 0) bt[4 * j + w] = v;
  }
  __syncthreads();

  // ------------------------------------------------------------------ A = M^T At M, b = M^T bt
  double* const T = sm + C::oT;  // T[j * NB + i] = sum_l At[j][l] M[l][i]
  for (int idx = tid; idx < NM * NB; idx += kThreads) {
    const int j = idx / NB, i = idx % NB;
    double v = 0.0;
    if (i < N) {
      for (int l = 0; l < NM; l++) v = fma(At[j * LDM + l], Mq[l * NB + i], v);
    }
    T[idx] = v;
  }
  for (int i = tid; i < N; i += kThreads) {
    double v = 0.0;
    for (int j = 0; j < NM; j++) v = fma(Mq[j * NB + i], bt[j], v);
    bv[i] = v;
  }
  __syncthreads();
  double* const A = sm + C::oA;
  {
    constexpr int NV = (N * N + kThreads - 1) / kThreads;
    double vals[NV];
#pragma unroll
    for (int cnt = 0; cnt < NV; cnt++) {
      const int idx = tid + cnt * kThreads;
      double v = 0.0;
      if (idx < N * N) {
        const int i = idx / N, i2 = idx % N;
        for (int j = 0; j < NM; j++) v = fma(Mq[j * NB + i], T[j * NB + i2], v);
      }
      vals[cnt] = v;
    }
    __syncthreads();  // A overwrites At (T lies behind it and is dead after this point)
#pragma unroll
    for (int cnt = 0; cnt < NV; cnt++) {
      const int idx = tid + cnt * kThreads;
      if (idx < N * N) A[(idx / N) * LDA + idx % N] = vals[cnt];
    }
  }
  __syncthreads();

  // ------------------------------------------------------------------ Solve (64-70): unsymmetric -> partial pivoting
  gauss_jordan_solve<N, NB, kThreads>(A, LDA, bv, false, a.singular);
  __syncthreads();

  // ------------------------------------------------------------------ CalcTentElEval (541-590) in the monomial basis
  for (int j = tid; j < NM; j += kThreads) {
    double v = 0.0;
    for (int i = 0; i < N; i++) v = fma(Mq[j * NB + i], bv[i], v);
    cm[j] = v;
  }
  __syncthreads();
  for (int idx = tid; idx < npts; idx += kThreads) {
    const int te = te0 + idx / Q, q = idx % Q;
    const double* __restrict__ rec = f.fe + (size_t)te * kFE;
    const double* __restrict__ lm = a.lam + q * d1;
    double xi[d1] = {0.0, 0.0, 0.0};
#pragma unroll
    for (int kk = 0; kk < d1; kk++) {
      const double l = lm[kk];
      xi[0] = fma(l, rec[kk * D], xi[0]);
      xi[1] = fma(l, rec[kk * D + 1], xi[1]);
      xi[2] = fma(l, rec[9 + kk], xi[2]);
    }
    Pw<P> pw;
    pw.set(xi);
    double val = 0.0, ux = 0.0, uy = 0.0, ut = 0.0;
    auto row = [&](auto jc) {
      constexpr int j = decltype(jc)::value;
      double v, dx, dy, dt;
      mono_d1<P, j>(pw, v, dx, dy, dt);
      const double cj = cm[j];
      val = fma(cj, v, val);
      ux = fma(cj, dx, ux);
      uy = fma(cj, dy, uy);
      ut = fma(cj, dt, ut);
    };
    mono_rows<P, 0>(row);
    mono_rows<P, 1>(row);
    mono_rows<P, 2>(row);
    mono_rows<P, 3>(row);
    double* __restrict__ wf = a.wavefront + (size_t)a.te_el[te] * (4 * Q) + q;
    wf[0] = val;
    wf[Q] = sc * ux;
    wf[2 * Q] = sc * uy;
    wf[3 * Q] = sc * ut;
  }
}

template <int P>
cudaError_t launch_p(const TentKernelArgs& a, const QtArgs& qa, const QtFastArgs& f, cudaStream_t stream) {
  using C = QCfg<P>;
  if (a.Q != C::Q || a.N != C::N || a.npoly != C::NM || qa.Qv != (P + 1) * (P + 1) * (P + 1)) return cudaErrorInvalidValue;
  const size_t smem = sizeof(double) * C::nsmem;
  // the attribute is per device: one flag per device ordinal, not per process (a process may hold solvers on several GPUs)
  static bool configured[64] = {};
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev < 0 || dev >= 64 || !configured[dev]) {
    e = cudaFuncSetAttribute(tent_kernel_qt_fast<P>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    if (dev >= 0 && dev < 64) configured[dev] = true;
  }
  tent_kernel_qt_fast<P><<<a.ntents, kThreads, smem, stream>>>(a, qa, f);
  return cudaGetLastError();
}

}  // namespace

#define NGST_QT_FAST_ORDERS(X) X(3) X(4) X(5)

bool tent_qt_fast_available(int D, int order) {
  if (D != 2) return false;
#define X(p) if (order == p) return true;
  NGST_QT_FAST_ORDERS(X)
#undef X
  return false;
}

int qt_fast_fe_stride() { return kFE; }
int qt_fast_cone_stride() { return kCS; }
int qt_fast_ff_stride() { return kFF; }

cudaError_t launch_qt_fast_stage(const TentKernelArgs& a, const QtArgs& qa, int ntents, double* fe, double* cone, double* ff,
                                 int* ok, cudaStream_t stream) {
  if (ntents <= 0) return cudaSuccess;
  qt_fast_stage_kernel<<<(ntents + 127) / 128, 128, 0, stream>>>(a, qa, ntents, fe, cone, ff, ok);
  return cudaGetLastError();
}

cudaError_t launch_tent_qt_fast(int order, const TentKernelArgs& a, const QtArgs& qa, const QtFastArgs& f,
                                cudaStream_t stream) {
  if (a.ntents <= 0) return cudaSuccess;
#define X(p) if (order == p) return launch_p<p>(a, qa, f, stream);
  NGST_QT_FAST_ORDERS(X)
#undef X
  return cudaErrorInvalidValue;
}

}  // namespace ngst
