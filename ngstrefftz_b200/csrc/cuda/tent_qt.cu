// Quasi-Trefftz tent kernel: one CTA per tent, D in {1, 2}, smoothly varying wavespeed.
//
// Per-tent body of QTWaveTents<D>::Propagate (src/twavetents.cpp:967-1123):
//   * the quasi-Trefftz basis of the tent is built in shared memory from the Taylor tables of 1/c^2 and B at the tent
//     vertex (QTWaveBasis<D+1>::Basis, src/trefftzfespace.cpp:1053-1198) -- one thread per basis function; the reference
//     serialises this behind a mutex and a string-keyed cache (trefftzfespace.cpp:1056-1061);
//   * space-like faces with the wavespeed at the quadrature points (CalcTentEl with LocalWavespeed(imip),
//     twavetents.cpp:1001-1016), boundary faces (CalcTentBndEl, 335-434, 1/c^2 at the face points);
//   * volume terms over the two cones of every patch element (1032-1108) with the wave operator of
//     ScalarMappedElement::CalcDDWaveOperator (src/scalarmappedfe.cpp:919-1008):
//       A += sum_q vol w_q box(phi_i) (d_t phi_j + mu_q box(phi_j)),  box = Laplace - c^-2 d_tt,  mu = h / max(c^-2) * c^2;
//   * pivoted LU (Solve, 64-70) and top-face evaluation (CalcTentElEval, 541-590).
// Shared-memory operand layout: E[(kind*Qc + q)*Np + i]; kind 0 = value, 1+u = derivative u, D+2 = wave operator.
#include <cuda_runtime.h>

#include "tent_qt.cuh"

namespace ngst {

namespace {

constexpr int kThreads = 128;

struct Smem {
  double *A, *bv, *MqT, *AAs, *BBs, *pt, *pw, *inv, *mono, *E, *Y, *rr, *vinv;
};

template <int D>
__host__ __device__ inline size_t carve(double* base, const TentKernelArgs& a, const QtArgs& q, Smem* s) {
  constexpr int d1 = D + 1, NK = D + 3;
  const int P = a.order, Np = a.Npad, Qc = a.Qc;
  size_t o = 0;
  auto take = [&](size_t n) { double* p = base ? base + o : nullptr; o += (n + 1) / 2 * 2; return p; };
  double* A = take((size_t)a.N * (a.N + 1));
  double* bv = take(Np);
  double* MqT = take((size_t)a.npoly * Np);
  double* AAs = take((size_t)(P > 1 ? (P - 1) * (P - 1) : 1));
  double* BBs = take((size_t)P * P);
  double* pt = take((size_t)Qc * d1);
  double* pw = take((size_t)Qc * d1 * (P + 1));
  double* inv = take(Qc);
  double* mono = take((size_t)Qc * NK * a.npoly);
  double* E = take((size_t)NK * Qc * Np);
  double* Y = take((size_t)NK * Qc * Np);
  double* rr = take((size_t)NK * Qc);
  double* vinv = take(q.Qv);
  if (s) *s = Smem{A, bv, MqT, AAs, BBs, pt, pw, inv, mono, E, Y, rr, vinv};
  return o;
}

__device__ __forceinline__ double poly_eval(const QtArgs& q, int D, const double* x) {
  double s = 0.0;
  for (int j = 0; j < q.ncterm; j++) {
    double m = q.ccoef[j];
    for (int d = 0; d < D; d++)
      for (int k = 0; k < q.cexpo[j * D + d]; k++) m *= x[d];
    s += m;
  }
  return s;
}

// Evaluate the kinds in kmask at the nq points of s.pt (physical space-time coordinates). s.inv[q] = 1/c^2 at the point
// (used by the wave operator only).
template <int D>
__device__ void eval_qt(const TentKernelArgs& a, const Smem& s, int nq, const double* ctr, double sc, unsigned kmask) {
  constexpr int d1 = D + 1, NK = D + 3;
  const int tid = threadIdx.x, nth = blockDim.x;
  const int P1 = a.order + 1, npoly = a.npoly, Np = a.Npad, N = a.N, Qc = a.Qc;
  for (int idx = tid; idx < nq * d1; idx += nth) {
    const int q = idx / d1, r = idx % d1;
    const double xi = (s.pt[q * d1 + r] - ctr[r]) * sc;
    double* p = s.pw + (q * d1 + r) * P1;
    double v = 1.0;
    p[0] = 1.0;
    for (int k = 1; k < P1; k++) { v *= xi; p[k] = v; }
  }
  __syncthreads();
  for (int idx = tid; idx < nq * npoly; idx += nth) {
    const int q = idx / npoly, j = idx % npoly;
    const int* e = a.expo + j * d1;
    const double* p = s.pw + q * d1 * P1;
    double f[d1];
    double full = 1.0;
#pragma unroll
    for (int r = 0; r < d1; r++) { f[r] = p[r * P1 + e[r]]; full *= f[r]; }
    double* mq = s.mono + (size_t)q * NK * npoly;
    mq[j] = full;
    double wop = 0.0;
#pragma unroll
    for (int r = 0; r < d1; r++) {
      double der = 0.0, dd = 0.0;
      if (e[r] > 0) der = (double)e[r] * p[r * P1 + e[r] - 1];
      if (e[r] > 1) dd = (double)(e[r] * (e[r] - 1)) * p[r * P1 + e[r] - 2];
#pragma unroll
      for (int rr = 0; rr < d1; rr++)
        if (rr != r) { der *= f[rr]; dd *= f[rr]; }
      mq[(1 + r) * npoly + j] = der;
      wop += (r == D) ? -s.inv[q] * dd : dd;
    }
    mq[(D + 2) * npoly + j] = wop;
  }
  __syncthreads();
  for (int idx = tid; idx < NK * nq * Np; idx += nth) {
    const int i = idx % Np;
    const int q = (idx / Np) % nq;
    const int k = idx / (Np * nq);
    if (!((kmask >> k) & 1u)) continue;
    double val = 0.0;
    if (i < N) {
      const double* mq = s.mono + ((size_t)q * NK + k) * npoly;
      for (int j = 0; j < npoly; j++) val = fma(s.MqT[(size_t)j * Np + i], mq[j], val);
      val *= (k == 0) ? 1.0 : ((k == D + 2) ? sc * sc : sc);
    }
    s.E[(size_t)(k * Qc + q) * Np + i] = val;
  }
  __syncthreads();
}

// A[i][j] += sum_{k in [k0,k1), q < nq} L[(k*Qc+q)*Np + i] * R[(k*Qc+q)*Np + j]
__device__ void gram(double* A, int lda, int N, const double* L, const double* R, int k0, int k1, int nq, int Qc, int Np) {
  const int nt = Np / 4;
  for (int tile = threadIdx.x; tile < nt * nt; tile += blockDim.x) {
    const int ti = tile / nt, tj = tile % nt;
    double acc[4][4];
#pragma unroll
    for (int x = 0; x < 4; x++)
#pragma unroll
      for (int y = 0; y < 4; y++) acc[x][y] = 0.0;
    for (int k = k0; k < k1; k++)
      for (int q = 0; q < nq; q++) {
        const size_t o = (size_t)(k * Qc + q) * Np;
        double l[4], r[4];
#pragma unroll
        for (int x = 0; x < 4; x++) { l[x] = L[o + 4 * ti + x]; r[x] = R[o + 4 * tj + x]; }
#pragma unroll
        for (int x = 0; x < 4; x++)
#pragma unroll
          for (int y = 0; y < 4; y++) acc[x][y] = fma(l[x], r[y], acc[x][y]);
      }
#pragma unroll
    for (int x = 0; x < 4; x++)
#pragma unroll
      for (int y = 0; y < 4; y++)
        if (4 * ti + x < N && 4 * tj + y < N) A[(size_t)(4 * ti + x) * lda + 4 * tj + y] += acc[x][y];
  }
}

__device__ void gemv(double* b, const double* E, const double* rr, unsigned kmask, int NK, int nq, int Qc, int Np, int N) {
  for (int i = threadIdx.x; i < N; i += blockDim.x) {
    double acc = 0.0;
    for (int k = 0; k < NK; k++) {
      if (!((kmask >> k) & 1u)) continue;
      for (int q = 0; q < nq; q++) acc = fma(E[(size_t)(k * Qc + q) * Np + i], rr[k * Qc + q], acc);
    }
    b[i] += acc;
  }
}

template <int D>
__global__ void __launch_bounds__(kThreads) tent_kernel_qt(TentKernelArgs a, QtArgs qa) {
  constexpr int d1 = D + 1, NK = D + 3, KW = D + 2;  // KW: kind index of the wave operator
  extern __shared__ double smem_raw[];
  Smem s;
  carve<D>(smem_raw, a, qa, &s);
  const int tid = threadIdx.x, nth = blockDim.x;
  const int tent = a.tents[blockIdx.x];
  const int N = a.N, Np = a.Npad, Qc = a.Qc, Q = a.Q, P = a.order, npoly = a.npoly;
  const int LDA = N + 1;
  double ctr[d1];
#pragma unroll
  for (int d = 0; d < d1; d++) ctr[d] = a.center[(size_t)tent * d1 + d];
  const double h = qa.tentsize[tent];
  const double sc = 1.0 / h;
  const unsigned KALLD = ((1u << d1) - 1u) << 1;  // all derivative kinds

  // ------------------------------------------------------------------ quasi-Trefftz basis of this tent
  {
    const int v = qa.tent_vertex[tent];
    for (int idx = tid; idx < (P - 1) * (P - 1); idx += nth) {
      const int nx = idx / (P - 1), ny = idx % (P - 1);
      const int j = qa.idxA[idx];
      s.AAs[idx] = j >= 0 ? qa.aa[(size_t)v * qa.nA + j] * pow(h, (double)(nx + ny)) : 0.0;
    }
    for (int idx = tid; idx < P * P; idx += nth) {
      const int nx = idx / P, ny = idx % P;
      const int j = qa.idxB[idx];
      s.BBs[idx] = j >= 0 ? qa.bb[(size_t)v * qa.nB + j] * pow(h, (double)(nx + ny)) : 0.0;
    }
    for (int idx = tid; idx < npoly * Np; idx += nth) s.MqT[idx] = 0.0;
    for (int idx = tid; idx < N * LDA; idx += nth) s.A[idx] = 0.0;
    for (int idx = tid; idx < Np; idx += nth) s.bv[idx] = 0.0;
    __syncthreads();
    const int P1 = P + 1;
#define IDX3(x, y, t) qa.idx3[((x) * P1 + (y)) * P1 + (t)]
#define QB(j) s.MqT[(size_t)(j) * Np + b]
#define AAF(x, y) s.AAs[(x) * (P - 1) + (y)]
#define BBF(x, y) s.BBs[(x) * P + (y)]
    for (int b = tid; b < N; b += nth) {
      int cnt = 0;
      for (int t = 0; t < 2; t++)
        for (int x = 0; x <= P - t; x++)
          for (int y = 0; y <= (P - x - t) * (d1 == 3); y++) {
            if (cnt == b) QB(IDX3(x, y, t)) = 1.0;
            cnt++;
          }
      const double A0 = AAF(0, 0);
      for (int ell = 0; ell < P - 1; ell++)
        for (int t = 0; t <= ell; t++)
          for (int x = (d1 == 2 ? ell - t : 0); x <= ell - t; x++) {
            const int y = ell - t - x;
            const double den = (double)((t + 2) * (t + 1)) * A0;
            double nc = 0.0;
            for (int bx = 0; bx <= x; bx++)
              for (int by = (d1 == 3) ? 0 : y; by <= y; by++) {
                nc += (double)((bx + 2) * (bx + 1)) / den * BBF(x - bx, y - by) * QB(IDX3(bx + 2, by, t)) +
                      (double)((x - bx + 1) * (bx + 1)) / den * BBF(x - bx + 1, y - by) * QB(IDX3(bx + 1, by, t));
                if (d1 == 3)
                  nc += (double)((by + 2) * (by + 1)) / den * BBF(x - bx, y - by) * QB(IDX3(bx, by + 2, t)) +
                        (double)((y - by + 1) * (by + 1)) / den * BBF(x - bx, y - by + 1) * QB(IDX3(bx, by + 1, t));
                if (bx + by == x + y) continue;
                nc -= AAF(x - bx, y - by) * QB(IDX3(bx, by, t + 2)) / A0;
              }
            QB(IDX3(x, y, t + 2)) = nc;
          }
    }
#undef IDX3
#undef QB
#undef AAF
#undef BBF
    __syncthreads();
  }

  // ------------------------------------------------------------------ space-like faces (CalcTentEl, per-point wavespeed)
  for (int te = a.te_off[tent]; te < a.te_off[tent + 1]; te++) {
    const int el = a.te_el[te];
    const double* geo = a.te_geo + (size_t)te * 2 * (d1 + 1);
    const double* wf = a.wavefront + (size_t)el * a.ncomp * Q;
    const double* cq = qa.c_point + (size_t)el * Q;
    double Pv[d1][D];
#pragma unroll
    for (int k = 0; k < d1; k++) {
      const int vk = a.el2v[(size_t)el * d1 + k];
#pragma unroll
      for (int d = 0; d < D; d++) Pv[k][d] = a.pts[(size_t)vk * D + d];
    }
    for (int q0 = 0; q0 < Q; q0 += Qc) {
      const int nq = min(Qc, Q - q0);
      for (int side = 0; side < 2; side++) {
        const double* tau = (side == 0 ? a.te_tbot : a.te_ttop) + (size_t)te * d1;
        const double* nn = geo + side * (d1 + 1);
        const double area = nn[d1];
        for (int idx = tid; idx < nq * d1; idx += nth) {
          const int q = idx / d1, r = idx % d1;
          const double* lm = a.lam + (size_t)(q0 + q) * d1;
          double x = 0.0;
          if (r < D) { for (int k = 0; k < d1; k++) x += lm[k] * Pv[k][r]; }
          else { for (int k = 0; k < d1; k++) x += lm[k] * tau[k]; }
          s.pt[q * d1 + r] = x;
        }
        __syncthreads();
        if (side == 0) {
          eval_qt<D>(a, s, nq, ctr, sc, KALLD | 1u);
          for (int q = tid; q < nq; q += nth) {
            const double om = a.w[q0 + q] * area;
            const double c = cq[q0 + q];
            const double gt = wf[(size_t)(1 + D) * Q + q0 + q];
            double rt = nn[D] / (c * c) * om * gt;
#pragma unroll
            for (int d = 0; d < D; d++) {
              const double gx = wf[(size_t)(1 + d) * Q + q0 + q];
              s.rr[(1 + d) * Qc + q] = -(nn[D] * om * gx - nn[d] * om * gt);
              rt -= nn[d] * om * gx;
            }
            s.rr[(1 + D) * Qc + q] = -rt;
            s.rr[q] = om * wf[q0 + q];
          }
          for (int idx = tid; idx < nq * Np; idx += nth) {
            const int q = idx / Np;
            s.Y[idx] = a.w[q0 + q] * area * s.E[idx];
          }
          __syncthreads();
          gemv(s.bv, s.E, s.rr, KALLD | 1u, NK, nq, Qc, Np, N);
          gram(s.A, LDA, N, s.E, s.Y, 0, 1, nq, Qc, Np);
          __syncthreads();
        } else {
          eval_qt<D>(a, s, nq, ctr, sc, KALLD);
          for (int idx = tid; idx < nq * Np; idx += nth) {
            const int q = idx / Np, j = idx % Np;
            const double om = a.w[q0 + q] * area;
            const double c = cq[q0 + q];
            const double bt = s.E[(size_t)((1 + D) * Qc + q) * Np + j];
            double yt = nn[D] / (c * c) * om * bt;
#pragma unroll
            for (int d = 0; d < D; d++) {
              const double bd = s.E[(size_t)((1 + d) * Qc + q) * Np + j];
              s.Y[(size_t)((1 + d) * Qc + q) * Np + j] = nn[D] * om * bd - nn[d] * om * bt;
              yt -= nn[d] * om * bd;
            }
            s.Y[(size_t)((1 + D) * Qc + q) * Np + j] = yt;
          }
          __syncthreads();
          gram(s.A, LDA, N, s.E, s.Y, 1, 2 + D, nq, Qc, Np);
          __syncthreads();
        }
      }
    }
  }

  // ------------------------------------------------------------------ boundary faces (CalcTentBndEl, 1/c^2 at the points)
  for (int tf = a.tf_off[tent]; tf < a.tf_off[tent + 1]; tf++) {
    const int kind = a.tf_kind[tf];
    const double* V = a.tf_verts + (size_t)tf * d1 * d1;
    double nrm[D];
#pragma unroll
    for (int d = 0; d < D; d++) nrm[d] = a.tf_n[(size_t)tf * D + d];
    const double area = a.tf_area[tf];
    const long long off = a.tf_data[tf];
    for (int q0 = 0; q0 < Q; q0 += Qc) {
      const int nq = min(Qc, Q - q0);
      for (int idx = tid; idx < nq * d1; idx += nth) {
        const int q = idx / d1, r = idx % d1;
        double x = V[r];
        for (int m = 0; m < D; m++) x += (V[(m + 1) * d1 + r] - V[r]) * a.lam[(size_t)(q0 + q) * d1 + m];
        s.pt[q * d1 + r] = x;
      }
      __syncthreads();
      for (int q = tid; q < nq; q += nth) {
        const double c = poly_eval(qa, D, s.pt + q * d1);
        s.inv[q] = 1.0 / (c * c);
      }
      __syncthreads();
      eval_qt<D>(a, s, nq, ctr, sc, KALLD);
      if (kind == 0) {  // Dirichlet, twavetents.cpp:404-433
        const double alpha = 0.5;
        for (int idx = tid; idx < nq * Np; idx += nth) {
          const int q = idx / Np, j = idx % Np;
          const double om = a.w[q0 + q] * area;
          double y = om * alpha * s.inv[q] * s.E[(size_t)((1 + D) * Qc + q) * Np + j];
#pragma unroll
          for (int d = 0; d < D; d++) y -= nrm[d] * om * s.E[(size_t)((1 + d) * Qc + q) * Np + j];
          s.Y[(size_t)((1 + D) * Qc + q) * Np + j] = y;
        }
        for (int q = tid; q < nq; q += nth) {
          const double om = a.w[q0 + q] * area;
          const double g = a.bnd_data ? a.bnd_data[off + q0 + q] : 0.0;
          s.rr[(1 + D) * Qc + q] = om * alpha * s.inv[q] * g;
#pragma unroll
          for (int d = 0; d < D; d++) s.rr[(1 + d) * Qc + q] = nrm[d] * om * g;
        }
        __syncthreads();
        gram(s.A, LDA, N, s.E, s.Y, 1 + D, 2 + D, nq, Qc, Np);
        gemv(s.bv, s.E, s.rr, KALLD, NK, nq, Qc, Np, N);
        __syncthreads();
      } else {  // Neumann, twavetents.cpp:380-403
        const double beta = 0.5;
        for (int idx = tid; idx < nq * Np; idx += nth) {
          const int q = idx / Np, j = idx % Np;
          const double om = a.w[q0 + q] * area;
          double sum = s.E[(size_t)((1 + D) * Qc + q) * Np + j];
#pragma unroll
          for (int d = 0; d < D; d++) sum += -nrm[d] * beta * s.E[(size_t)((1 + d) * Qc + q) * Np + j];
#pragma unroll
          for (int r = 0; r < D; r++) s.Y[(size_t)((1 + r) * Qc + q) * Np + j] = -nrm[r] * om * sum;
          s.Y[(size_t)((1 + D) * Qc + q) * Np + j] = 0.0;
        }
        for (int q = tid; q < nq; q += nth) {
          const double om = a.w[q0 + q] * area;
          double gn = 0.0;
#pragma unroll
          for (int r = 0; r < D; r++)
            gn += -nrm[r] * (a.bnd_data ? a.bnd_data[off + (long long)r * Q + q0 + q] : 0.0);
#pragma unroll
          for (int d = 0; d < D; d++) s.rr[(1 + d) * Qc + q] = -nrm[d] * beta * gn * om;
          s.rr[(1 + D) * Qc + q] = -gn * om;
        }
        __syncthreads();
        gram(s.A, LDA, N, s.E, s.Y, 1, 2 + D, nq, Qc, Np);
        gemv(s.bv, s.E, s.rr, KALLD, NK, nq, Qc, Np, N);
        __syncthreads();
      }
    }
  }

  // ------------------------------------------------------------------ volume terms over the bottom / top cones (1032-1108)
  __shared__ double red[kThreads];
  for (int te = a.te_off[tent]; te < a.te_off[tent + 1]; te++) {
    const int el = a.te_el[te];
    for (int part = 0; part < 2; part++) {
      const double* tau = (part == 0 ? a.te_tbot : a.te_ttop) + (size_t)te * d1;
      double M[d1][d1];  // columns: face vertices minus the tent centre
#pragma unroll
      for (int k = 0; k < d1; k++) {
        const int vk = a.el2v[(size_t)el * d1 + k];
#pragma unroll
        for (int d = 0; d < D; d++) M[d][k] = a.pts[(size_t)vk * D + d] - ctr[d];
        M[D][k] = tau[k] - ctr[D];
      }
      double det;
      if (d1 == 2) det = M[0][0] * M[1][1] - M[0][1] * M[1][0];
      else
        det = M[0][0] * (M[1][1] * M[2 % d1][2 % d1] - M[1][2 % d1] * M[2 % d1][1]) -
              M[0][1] * (M[1][0] * M[2 % d1][2 % d1] - M[1][2 % d1] * M[2 % d1][0]) +
              M[0][2 % d1] * (M[1][0] * M[2 % d1][1] - M[1][1] * M[2 % d1][0]);
      const double vol = fabs(det);
      if (vol < 10e-16) continue;  // block-uniform
      // 1/c^2 at every cone point and its maximum (cmax, 1089-1094)
      double mx = 0.0;
      for (int q = tid; q < qa.Qv; q += nth) {
        double x[D];
#pragma unroll
        for (int d = 0; d < D; d++) {
          double v = ctr[d];
#pragma unroll
          for (int k = 0; k < d1; k++) v += M[d][k] * qa.vxi[(size_t)q * d1 + k];
          x[d] = v;
        }
        const double c = poly_eval(qa, D, x);
        const double iv = 1.0 / (c * c);
        s.vinv[q] = iv;
        mx = fmax(mx, fabs(iv));
      }
      red[tid] = mx;
      __syncthreads();
      for (int o = kThreads / 2; o > 0; o >>= 1) {
        if (tid < o) red[tid] = fmax(red[tid], red[tid + o]);
        __syncthreads();
      }
      const double cmax = red[0];
      __syncthreads();
      for (int q0 = 0; q0 < qa.Qv; q0 += Qc) {
        const int nq = min(Qc, qa.Qv - q0);
        for (int idx = tid; idx < nq * d1; idx += nth) {
          const int q = idx / d1, r = idx % d1;
          double v = ctr[r];
#pragma unroll
          for (int k = 0; k < d1; k++) v += M[r][k] * qa.vxi[(size_t)(q0 + q) * d1 + k];
          s.pt[q * d1 + r] = v;
        }
        for (int q = tid; q < nq; q += nth) s.inv[q] = s.vinv[q0 + q];
        __syncthreads();
        eval_qt<D>(a, s, nq, ctr, sc, (1u << (1 + D)) | (1u << KW));
        // L = vol w box(phi_i) into Y[kind KW]; R = d_t phi_j + mu box(phi_j) into Y[kind 1+D]
        for (int idx = tid; idx < nq * Np; idx += nth) {
          const int q = idx / Np, j = idx % Np;
          const double bx = s.E[(size_t)(KW * Qc + q) * Np + j];
          const double mu = h / cmax / s.inv[q];
          s.Y[(size_t)(KW * Qc + q) * Np + j] = vol * qa.vw[q0 + q] * bx;
          s.Y[(size_t)((1 + D) * Qc + q) * Np + j] = s.E[(size_t)((1 + D) * Qc + q) * Np + j] + mu * bx;
        }
        __syncthreads();
        gram(s.A, LDA, N, s.Y + (size_t)(KW - (1 + D)) * Qc * Np, s.Y, 1 + D, 2 + D, nq, Qc, Np);
        __syncthreads();
      }
    }
  }

  // ------------------------------------------------------------------ Solve: LU with partial pivoting (implicit permutation)
  const int lane = tid & 31, w = tid >> 5;
  for (int i = tid; i < N; i += nth) s.A[i * LDA + N] = s.bv[i];
  __shared__ int perm[64];
  __syncthreads();
  {
    unsigned long long done = 0ull;
    for (int kk = 0; kk < N; kk++) {
      double best = -1.0;
      int brow = 0x7fffffff;
      for (int r = lane; r < N; r += 32) {
        if ((done >> r) & 1ull) continue;
        const double v = fabs(s.A[r * LDA + kk]);
        if (v > best) { best = v; brow = r; }
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        const double ob = __shfl_xor_sync(0xffffffffu, best, o);
        const int orow = __shfl_xor_sync(0xffffffffu, brow, o);
        if (ob > best || (ob == best && orow < brow)) { best = ob; brow = orow; }
      }
      const int p = brow;
      done |= 1ull << p;
      double pinv = 0.0;
      if (best > 0.0) pinv = 1.0 / s.A[p * LDA + kk];
      if (tid == 0) {
        perm[kk] = p;
        if (!(best > 0.0)) atomicAdd(a.singular, 1);
      }
      const int c0 = kk + 1 + lane, c1 = c0 + 32;
      const double pr0 = (c0 <= N) ? s.A[p * LDA + c0] : 0.0;
      const double pr1 = (c1 <= N) ? s.A[p * LDA + c1] : 0.0;
      for (int r = w; r < N; r += kThreads / 32) {
        if ((done >> r) & 1ull) continue;
        const double m = s.A[r * LDA + kk] * pinv;
        if (c0 <= N) s.A[r * LDA + c0] -= m * pr0;
        if (c1 <= N) s.A[r * LDA + c1] -= m * pr1;
      }
      __syncthreads();
    }
  }
  if (w == 0) {
    double r0 = (lane < N) ? s.A[lane * LDA + N] : 0.0;
    double r1 = (lane + 32 < N) ? s.A[(lane + 32) * LDA + N] : 0.0;
    for (int kk = N - 1; kk >= 0; kk--) {
      const int p = perm[kk];
      const double d = s.A[p * LDA + kk];
      const double rp = __shfl_sync(0xffffffffu, (p < 32) ? r0 : r1, p & 31);
      const double xk = (d != 0.0) ? rp / d : 0.0;
      if (lane == 0) s.bv[kk] = xk;
      if (lane < N) r0 -= s.A[lane * LDA + kk] * xk;
      if (lane + 32 < N) r1 -= s.A[(lane + 32) * LDA + kk] * xk;
    }
  }
  __syncthreads();

  // ------------------------------------------------------------------ CalcTentElEval (541-590)
  const int nw = nth >> 5;
  for (int te = a.te_off[tent]; te < a.te_off[tent + 1]; te++) {
    const int el = a.te_el[te];
    double* wf = a.wavefront + (size_t)el * a.ncomp * Q;
    const double* tau = a.te_ttop + (size_t)te * d1;
    double Pv[d1][D];
#pragma unroll
    for (int k = 0; k < d1; k++) {
      const int vk = a.el2v[(size_t)el * d1 + k];
#pragma unroll
      for (int d = 0; d < D; d++) Pv[k][d] = a.pts[(size_t)vk * D + d];
    }
    for (int q0 = 0; q0 < Q; q0 += Qc) {
      const int nq = min(Qc, Q - q0);
      for (int idx = tid; idx < nq * d1; idx += nth) {
        const int q = idx / d1, r = idx % d1;
        const double* lm = a.lam + (size_t)(q0 + q) * d1;
        double x = 0.0;
        if (r < D) { for (int k = 0; k < d1; k++) x += lm[k] * Pv[k][r]; }
        else { for (int k = 0; k < d1; k++) x += lm[k] * tau[k]; }
        s.pt[q * d1 + r] = x;
      }
      __syncthreads();
      eval_qt<D>(a, s, nq, ctr, sc, KALLD | 1u);
      const int nk = (d1 + 1) * nq;
      for (int kk = w; kk < nk; kk += nw) {
        const int k = kk / nq, q = kk % nq;
        const double* e = s.E + (size_t)(k * Qc + q) * Np;
        double acc = 0.0;
        for (int i = lane; i < N; i += 32) acc = fma(e[i], s.bv[i], acc);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) acc += __shfl_down_sync(0xffffffffu, acc, o);
        if (lane == 0) wf[(size_t)k * Q + q0 + q] = acc;
      }
      __syncthreads();
    }
  }
}

}  // namespace

size_t tent_qt_smem_bytes(int D, const TentKernelArgs& a, const QtArgs& q) {
  return sizeof(double) * (D == 1 ? carve<1>(nullptr, a, q, nullptr) : carve<2>(nullptr, a, q, nullptr));
}

cudaError_t launch_tent_qt(int D, const TentKernelArgs& a, const QtArgs& q, size_t smem, cudaStream_t stream) {
  if (a.ntents <= 0) return cudaSuccess;
  if (a.N > 64) return cudaErrorInvalidValue;
  cudaError_t err;
  if (D == 1) {
    err = cudaFuncSetAttribute(tent_kernel_qt<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (err != cudaSuccess) return err;
    tent_kernel_qt<1><<<a.ntents, kThreads, smem, stream>>>(a, q);
  } else if (D == 2) {
    err = cudaFuncSetAttribute(tent_kernel_qt<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (err != cudaSuccess) return err;
    tent_kernel_qt<2><<<a.ntents, kThreads, smem, stream>>>(a, q);
  } else {
    return cudaErrorInvalidValue;
  }
  return cudaGetLastError();
}

}  // namespace ngst
