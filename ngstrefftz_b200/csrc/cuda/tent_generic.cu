// Generic per-tent Trefftz-DG kernel: one CTA per tent, any D in {1,2,3}, any order, second- or
// first-order input, Dirichlet / Neumann boundary facets and material interfaces (macro elements).
//
// Implements the per-tent body of TWaveTents<D>::Propagate (src/twavetents.cpp:92-171):
//   CalcTentBndEl (335-434), CalcTentMacroEl (436-539), CalcTentEl (177-333), Solve (64-70),
//   CalcTentElEval (541-590), with the basis evaluation of ScalarMappedElement<D+1>::CalcShape /
//   CalcDShape (src/scalarmappedfe.cpp:437-674) done on the fly in shared memory.
// Shared-memory operand layout: E[(v*Qc + q)*Npad + i], v = 0 value, v = 1+d derivative in direction d.
//
// WARP = true (apply mode of the factor cache only, gmode 2): one WARP per tent instead of one CTA -- the same statements
// in the same order, executed by the 32 lanes of a warp with __syncwarp() in place of __syncthreads() and a private slice
// of shared memory without the matrix and Gram operands. Systems of the cache have at most 32 unknowns (1-D tents,
// BASELINE configs[1]): a CTA per tent leaves most of its threads idle and pays a block-wide barrier per step.
#include <cuda_runtime.h>

#include "tent_kernel.cuh"

namespace ngst {

constexpr int kGenericWarpsPerCta = 4;

// doubles of shared memory one warp of the warp-per-tent variant owns: [b | E0 | rr | pw | mono | pt], rounded up to even
__host__ __device__ inline size_t generic_warp_doubles(int D, const TentKernelArgs& a) {
  const size_t d1 = D + 1, nv = D + 2, Qc = a.Qc;
  const size_t n = (size_t)a.max_ndom * a.Npad + nv * Qc * a.Npad + nv * Qc + Qc * d1 * (a.order + 1) + Qc * nv * a.npoly + Qc * d1;
  return (n + 1) / 2 * 2;
}

namespace {

struct Smem {
  double *A, *b, *E0, *E1, *Y, *rr, *pw, *mono, *pt, *Inv;
};

template <bool WARP>
__device__ __forceinline__ void gsync() {
  if constexpr (WARP) __syncwarp();
  else __syncthreads();
}

template <int D, bool WARP>
__device__ __forceinline__ Smem carve(double* base, const TentKernelArgs& a) {
  constexpr int d1 = D + 1, nv = D + 2;
  const int nmax = a.max_ndom * a.Npad;
  const int esz = nv * a.Qc * a.Npad;
  Smem s;
  if constexpr (WARP) {  // apply only: [b | E0 | rr | pw | mono | pt] (generic_warp_doubles)
    s.A = nullptr; s.E1 = nullptr; s.Y = nullptr; s.Inv = nullptr;
    s.b = base;
    s.E0 = s.b + nmax;
    s.rr = s.E0 + esz;
    s.pw = s.rr + nv * a.Qc;
    s.mono = s.pw + a.Qc * d1 * (a.order + 1);
    s.pt = s.mono + a.Qc * nv * a.npoly;
    return s;
  }
  s.A = base;
  s.b = s.A + (size_t)nmax * nmax;
  s.E0 = s.b + nmax;
  s.E1 = s.E0 + esz;
  s.Y = s.E1 + esz;
  s.rr = s.Y + esz;
  s.pw = s.rr + nv * a.Qc;
  s.mono = s.pw + a.Qc * d1 * (a.order + 1);
  s.pt = s.mono + a.Qc * nv * a.npoly;
  s.Inv = s.pt + a.Qc * d1;  // (max_ndom * Npad)^2 doubles, only when the factor cache is being built (gmode 1)
  return s;
}

// Evaluate the scaled Trefftz basis (variants vlo..D+1) at the nq points in s.pt into E.
template <int D, bool WARP>
__device__ void eval_basis(const TentKernelArgs& a, const Smem& s, int nq, const double* ctr, double sc, double c,
                           int vlo, double* E) {
  constexpr int d1 = D + 1, nv = D + 2;
  const int tid = WARP ? (threadIdx.x & 31) : threadIdx.x, nth = WARP ? 32 : blockDim.x;
  const int P1 = a.order + 1, npoly = a.npoly, Np = a.Npad, N = a.N, Qc = a.Qc;
  for (int idx = tid; idx < nq * d1; idx += nth) {
    const int q = idx / d1, d = idx % d1;
    const double scale = (d == D) ? sc * c : sc;
    const double xi = (s.pt[q * d1 + d] - ctr[d]) * scale;
    double* p = s.pw + (q * d1 + d) * P1;
    double v = 1.0;
    p[0] = 1.0;
    for (int k = 1; k < P1; k++) { v *= xi; p[k] = v; }
  }
  gsync<WARP>();
  for (int idx = tid; idx < nq * npoly; idx += nth) {
    const int q = idx / npoly, ii = idx % npoly;
    const int* e = a.expo + ii * d1;
    const double* p = s.pw + q * d1 * P1;
    double f[d1];
    double full = 1.0;
#pragma unroll
    for (int d = 0; d < d1; d++) { f[d] = p[d * P1 + e[d]]; full *= f[d]; }
    double* mq = s.mono + (size_t)q * nv * npoly;
    mq[ii] = full;
#pragma unroll
    for (int d = 0; d < d1; d++) {
      double der = 0.0;
      if (e[d] > 0) {
        der = (double)e[d] * p[d * P1 + e[d] - 1];
#pragma unroll
        for (int dd = 0; dd < d1; dd++)
          if (dd != d) der *= f[dd];
      }
      mq[(1 + d) * npoly + ii] = der;
    }
  }
  gsync<WARP>();
  const int tot = (nv - vlo) * nq * Np;
  for (int idx = tid; idx < tot; idx += nth) {
    const int i = idx % Np;
    const int q = (idx / Np) % nq;
    const int v = vlo + idx / (Np * nq);
    double val = 0.0;
    if (i < N) {
      const double* mq = s.mono + ((size_t)q * nv + v) * npoly;
      for (int j = a.rowptr[i]; j < a.rowptr[i + 1]; j++) val += a.val[j] * mq[a.col[j]];
      if (v > 0) val *= (v - 1 == D) ? sc * c : sc;
    }
    E[(size_t)(v * Qc + q) * Np + i] = val;
  }
  gsync<WARP>();
}

// Asub[i][j] += sum_{v in [v0,v1), q < nq} L[(v*Qc+q)*Np + i] * R[(v*Qc+q)*Np + j], 4x4 register tiles.
__device__ void gram(double* Asub, int lda, const double* L, const double* R, int v0, int v1, int nq, int Qc, int Np) {
  const int nt = Np / 4;
  for (int tile = threadIdx.x; tile < nt * nt; tile += blockDim.x) {
    const int ti = tile / nt, tj = tile % nt;
    double acc[4][4];
#pragma unroll
    for (int x = 0; x < 4; x++)
#pragma unroll
      for (int y = 0; y < 4; y++) acc[x][y] = 0.0;
    for (int v = v0; v < v1; v++)
      for (int q = 0; q < nq; q++) {
        const size_t k = (size_t)(v * Qc + q) * Np;
        const double2 l0 = *reinterpret_cast<const double2*>(L + k + 4 * ti);
        const double2 l1 = *reinterpret_cast<const double2*>(L + k + 4 * ti + 2);
        const double2 r0 = *reinterpret_cast<const double2*>(R + k + 4 * tj);
        const double2 r1 = *reinterpret_cast<const double2*>(R + k + 4 * tj + 2);
        const double l[4] = {l0.x, l0.y, l1.x, l1.y};
        const double r[4] = {r0.x, r0.y, r1.x, r1.y};
#pragma unroll
        for (int x = 0; x < 4; x++)
#pragma unroll
          for (int y = 0; y < 4; y++) acc[x][y] = fma(l[x], r[y], acc[x][y]);
      }
#pragma unroll
    for (int x = 0; x < 4; x++)
#pragma unroll
      for (int y = 0; y < 4; y++) Asub[(size_t)(4 * ti + x) * lda + 4 * tj + y] += acc[x][y];
  }
}

// bsub[i] += sum_{v in [v0,v1), q<nq} E[(v*Qc+q)*Np + i] * rr[v*Qc+q]
template <bool WARP>
__device__ void gemv(double* bsub, const double* E, const double* rr, int v0, int v1, int nq, int Qc, int Np, int N) {
  const int tid = WARP ? (threadIdx.x & 31) : threadIdx.x, nth = WARP ? 32 : blockDim.x;
  for (int i = tid; i < N; i += nth) {
    double acc = 0.0;
    for (int v = v0; v < v1; v++)
      for (int q = 0; q < nq; q++) acc = fma(E[(size_t)(v * Qc + q) * Np + i], rr[v * Qc + q], acc);
    bsub[i] += acc;
  }
}

template <int D, bool WARP>
__global__ void tent_kernel_generic(TentKernelArgs a) {
  constexpr int d1 = D + 1, nv = D + 2;
  extern __shared__ double smem_raw[];
  const int tid = WARP ? (threadIdx.x & 31) : threadIdx.x, nth = WARP ? 32 : blockDim.x;
  // WARP: warp k of CTA b takes entry b * (warps per CTA) + k of the launch's tent list; surplus warps leave at once
  // (no block-wide barrier anywhere on this path)
  const int slot = WARP ? blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5) : blockIdx.x;
  if (WARP && slot >= a.ntents) return;
  const Smem s = carve<D, WARP>(smem_raw + (WARP ? (size_t)(threadIdx.x >> 5) * generic_warp_doubles(D, a) : 0), a);
  const int tent = a.tents[slot];
  const int nd = a.ndom[tent];
  const int N = a.N, Np = a.Npad, Qc = a.Qc, Q = a.Q;
  const int n = nd * Np;
  const int fo = a.fosystem;
  const int vlo_val = fo ? 1 : 0;
  // factor cache: the tent matrix does not depend on the state. apply = only right-hand side, stored inverse, evaluation
  const bool apply = WARP || a.gmode == 2;

  double ctr[d1];
#pragma unroll
  for (int d = 0; d < d1; d++) ctr[d] = a.center[(size_t)tent * d1 + d];
  const double sc = a.scale[tent];

  if (!apply)
    for (int idx = tid; idx < n * n; idx += nth) s.A[idx] = 0.0;
  for (int idx = tid; idx < n; idx += nth) s.b[idx] = 0.0;
  gsync<WARP>();

  // ------------------------------------------------------------------ time-like faces
  for (int tf = a.tf_off[tent]; tf < a.tf_off[tent + 1]; tf++) {
    const int kind = a.tf_kind[tf];
    if (apply && kind == 2) continue;  // material interfaces contribute to the matrix only
    const double* V = a.tf_verts + (size_t)tf * d1 * d1;
    double nrm[D];
#pragma unroll
    for (int d = 0; d < D; d++) nrm[d] = a.tf_n[(size_t)tf * D + d];
    const double area = a.tf_area[tf];
    const int din = a.tf_dom[2 * tf], dout = a.tf_dom[2 * tf + 1];
    const double c0 = a.tf_c[2 * tf], c1 = a.tf_c[2 * tf + 1];
    for (int q0 = 0; q0 < Q; q0 += Qc) {
      const int nq = min(Qc, Q - q0);
      for (int idx = tid; idx < nq * d1; idx += nth) {
        const int q = idx / d1, r = idx % d1;
        double x = V[r];
        for (int m = 0; m < D; m++) x += (V[(m + 1) * d1 + r] - V[r]) * a.lam[(size_t)(q0 + q) * d1 + m];
        s.pt[q * d1 + r] = x;
      }
      gsync<WARP>();
      eval_basis<D, WARP>(a, s, nq, ctr, sc, c0, 1, s.E0);
      if (kind == 2) eval_basis<D, WARP>(a, s, nq, ctr, sc, c1, 1, s.E1);

      if (kind == 0) {  // Dirichlet, twavetents.cpp:404-433
        const double alpha = 0.5, gam = 1.0 / (c0 * c0);
        const long long off = a.tf_data[tf];
        for (int idx = tid; idx < nq * Np; idx += nth) {
          const int q = idx / Np, j = idx % Np;
          const double om = a.w[q0 + q] * area;
          double y = om * alpha * gam * s.E0[(size_t)((1 + D) * Qc + q) * Np + j];
#pragma unroll
          for (int d = 0; d < D; d++) y -= nrm[d] * om * s.E0[(size_t)((1 + d) * Qc + q) * Np + j];
          s.Y[(size_t)((1 + D) * Qc + q) * Np + j] = y;
        }
        for (int q = tid; q < nq; q += nth) {
          const double om = a.w[q0 + q] * area;
          const double g = a.bnd_data ? a.bnd_data[off + q0 + q] : 0.0;
          s.rr[(1 + D) * Qc + q] = om * alpha * gam * g;
#pragma unroll
          for (int d = 0; d < D; d++) s.rr[(1 + d) * Qc + q] = nrm[d] * om * g;
        }
        gsync<WARP>();
        double* Asub = s.A + (size_t)(din * Np) * n + din * Np;
        if (!apply) gram(Asub, n, s.E0, s.Y, 1 + D, 2 + D, nq, Qc, Np);
        gemv<WARP>(s.b + din * Np, s.E0, s.rr, 1, nv, nq, Qc, Np, N);
        gsync<WARP>();
      } else if (kind == 1) {  // Neumann, twavetents.cpp:380-403
        const double beta = 0.5;
        const long long off = a.tf_data[tf];
        for (int idx = tid; idx < nq * Np; idx += nth) {
          const int q = idx / Np, j = idx % Np;
          const double om = a.w[q0 + q] * area;
          double sum = s.E0[(size_t)((1 + D) * Qc + q) * Np + j];
#pragma unroll
          for (int d = 0; d < D; d++) sum += -nrm[d] * beta * s.E0[(size_t)((1 + d) * Qc + q) * Np + j];
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
        gsync<WARP>();
        double* Asub = s.A + (size_t)(din * Np) * n + din * Np;
        if (!apply) gram(Asub, n, s.E0, s.Y, 1, nv, nq, Qc, Np);
        gemv<WARP>(s.b + din * Np, s.E0, s.rr, 1, nv, nq, Qc, Np, N);
        gsync<WARP>();
      } else if (!WARP) {  // material interface, twavetents.cpp:436-539 (matrix only: never reached when applying)
        for (int el = 0; el < 4; el++) {
          const double* Lm = (el / 2 == 0) ? s.E0 : s.E1;
          const double* Rm = (el % 2 == 0) ? s.E0 : s.E1;
          const double sg = (el / 2 == 0) ? 1.0 : -1.0;
          for (int idx = tid; idx < nq * Np; idx += nth) {
            const int q = idx / Np, j = idx % Np;
            const double om = a.w[q0 + q] * area;
            const double rt = Rm[(size_t)((1 + D) * Qc + q) * Np + j];
            double yt = 0.0;
#pragma unroll
            for (int d = 0; d < D; d++) {
              s.Y[(size_t)((1 + d) * Qc + q) * Np + j] = -sg * 0.5 * nrm[d] * om * rt;
              yt -= sg * 0.5 * nrm[d] * om * Rm[(size_t)((1 + d) * Qc + q) * Np + j];
            }
            s.Y[(size_t)((1 + D) * Qc + q) * Np + j] = yt;
          }
          gsync<WARP>();
          const int in = (el / 2 == 0) ? din : dout;
          const int out = (el % 2 == 0) ? din : dout;
          gram(s.A + (size_t)(in * Np) * n + out * Np, n, Lm, s.Y, 1, nv, nq, Qc, Np);
          gsync<WARP>();
        }
      }
    }
  }

  // ------------------------------------------------------------------ space-like faces (CalcTentEl)
  for (int te = a.te_off[tent]; te < a.te_off[tent + 1]; te++) {
    const int el = a.te_el[te];
    const int dom = a.te_dom[te];
    const double c = a.c_el[el];
    const double* geo = a.te_geo + (size_t)te * 2 * (d1 + 1);
    double* Asub = s.A + (size_t)(dom * Np) * n + dom * Np;
    double* bsub = s.b + dom * Np;
    const double* wf = a.wavefront + (size_t)el * a.ncomp * Q;
    double P[d1][D];
#pragma unroll
    for (int k = 0; k < d1; k++) {
      const int vk = a.el2v[(size_t)el * d1 + k];
#pragma unroll
      for (int d = 0; d < D; d++) P[k][d] = a.pts[(size_t)vk * D + d];
    }
    for (int q0 = 0; q0 < Q; q0 += Qc) {
      const int nq = min(Qc, Q - q0);
      for (int side = 0; side < (apply ? 1 : 2); side++) {  // top faces contribute to the matrix only
        const double* tau = (side == 0 ? a.te_tbot : a.te_ttop) + (size_t)te * d1;
        const double* nn = geo + side * (d1 + 1);
        const double area = nn[d1];
        for (int idx = tid; idx < nq * d1; idx += nth) {
          const int q = idx / d1, r = idx % d1;
          const double* lm = a.lam + (size_t)(q0 + q) * d1;
          double x = 0.0;
          if (r < D) { for (int k = 0; k < d1; k++) x += lm[k] * P[k][r]; }
          else { for (int k = 0; k < d1; k++) x += lm[k] * tau[k]; }
          s.pt[q * d1 + r] = x;
        }
        gsync<WARP>();
        if (side == 0) {
          eval_basis<D, WARP>(a, s, nq, ctr, sc, c, vlo_val, s.E0);
          for (int q = tid; q < nq; q += nth) {
            const double om = a.w[q0 + q] * area;
            const double gt = wf[(size_t)((!fo) + D) * Q + q0 + q];
            double rt = nn[D] / (c * c) * om * gt;
#pragma unroll
            for (int d = 0; d < D; d++) {
              const double gx = wf[(size_t)((!fo) + d) * Q + q0 + q];
              s.rr[(1 + d) * Qc + q] = -(nn[D] * om * gx - nn[d] * om * gt);
              rt -= nn[d] * om * gx;
            }
            s.rr[(1 + D) * Qc + q] = -rt;
            if (!fo) s.rr[q] = om * wf[q0 + q];
          }
          if (!fo && !apply)
            for (int idx = tid; idx < nq * Np; idx += nth) {
              const int q = idx / Np;
              s.Y[idx] = a.w[q0 + q] * area * s.E0[idx];
            }
          gsync<WARP>();
          gemv<WARP>(bsub, s.E0, s.rr, vlo_val, nv, nq, Qc, Np, N);
          if (!fo && !apply) gram(Asub, n, s.E0, s.Y, 0, 1, nq, Qc, Np);
          gsync<WARP>();
        } else {
          eval_basis<D, WARP>(a, s, nq, ctr, sc, c, 1, s.E0);
          for (int idx = tid; idx < nq * Np; idx += nth) {
            const int q = idx / Np, j = idx % Np;
            const double om = a.w[q0 + q] * area;
            const double bt = s.E0[(size_t)((1 + D) * Qc + q) * Np + j];
            double yt = nn[D] / (c * c) * om * bt;
#pragma unroll
            for (int d = 0; d < D; d++) {
              const double bd = s.E0[(size_t)((1 + d) * Qc + q) * Np + j];
              s.Y[(size_t)((1 + d) * Qc + q) * Np + j] = nn[D] * om * bd - nn[d] * om * bt;
              yt -= nn[d] * om * bd;
            }
            s.Y[(size_t)((1 + D) * Qc + q) * Np + j] = yt;
          }
          gsync<WARP>();
          gram(Asub, n, s.E0, s.Y, 1, nv, nq, Qc, Np);
          gsync<WARP>();
        }
      }
    }
  }

  // ------------------------------------------------------------------ Solve
  __shared__ int piv_row;
  __shared__ double piv_inv;
  if (!WARP && a.gmode == 0) {
  // LU with partial pivoting
  for (int g = 0; g < nd; g++)
    for (int i = N + tid; i < Np; i += nth) s.A[(size_t)(g * Np + i) * n + g * Np + i] = 1.0;
  gsync<WARP>();
  for (int k = 0; k < n; k++) {
    if (tid < 32) {
      double best = -1.0;
      int brow = k;
      for (int r = k + tid; r < n; r += 32) {
        const double v = fabs(s.A[(size_t)r * n + k]);
        if (v > best) { best = v; brow = r; }
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        const double ob = __shfl_down_sync(0xffffffffu, best, o);
        const int orow = __shfl_down_sync(0xffffffffu, brow, o);
        if (ob > best || (ob == best && orow < brow)) { best = ob; brow = orow; }
      }
      if (tid == 0) {
        piv_row = brow;
        if (best == 0.0) { atomicAdd(a.singular, 1); piv_inv = 0.0; }
        else piv_inv = 1.0 / s.A[(size_t)brow * n + k];
      }
    }
    gsync<WARP>();
    const int p = piv_row;
    if (p != k) {
      for (int cidx = tid; cidx <= n; cidx += nth) {
        if (cidx < n) {
          const double t0 = s.A[(size_t)k * n + cidx];
          s.A[(size_t)k * n + cidx] = s.A[(size_t)p * n + cidx];
          s.A[(size_t)p * n + cidx] = t0;
        } else {
          const double t0 = s.b[k]; s.b[k] = s.b[p]; s.b[p] = t0;
        }
      }
      gsync<WARP>();
    }
    const double pinv = piv_inv;
    const int nr = n - k - 1, nc = n - k;  // columns k+1..n-1 and the right-hand side
    for (int idx = tid; idx < nr * nc; idx += nth) {
      const int r = k + 1 + idx / nc, cc = idx % nc;
      const double m = s.A[(size_t)r * n + k] * pinv;
      if (cc < nc - 1) s.A[(size_t)r * n + k + 1 + cc] -= m * s.A[(size_t)k * n + k + 1 + cc];
      else s.b[r] -= m * s.b[k];
    }
    gsync<WARP>();
  }
  for (int k = n - 1; k >= 0; k--) {
    if (tid == 0) {
      const double d = s.A[(size_t)k * n + k];
      s.b[k] = (d != 0.0) ? s.b[k] / d : 0.0;
    }
    gsync<WARP>();
    const double xk = s.b[k];
    for (int r = tid; r < k; r += nth) s.b[r] -= s.A[(size_t)r * n + k] * xk;
    gsync<WARP>();
  }

  } else {
    // factor cache. The inverse is kept transposed: invT[j * n + i] = (A^-1)[i][j] (consecutive threads read consecutive
    // doubles in the product). Stride n = nd * Npad of THIS tent inside its slot of (max_ndom * Npad)^2 doubles.
    double* const invT = a.ginv + (size_t)tent * ((size_t)a.max_ndom * Np) * ((size_t)a.max_ndom * Np);
    if (!WARP && a.gmode == 1) {
      // Gauss-Jordan with partial pivoting on [A | I] (the reference forms the inverse as well, twavetents.cpp:64-70)
      for (int g = 0; g < nd; g++)
        for (int i = N + tid; i < Np; i += nth) s.A[(size_t)(g * Np + i) * n + g * Np + i] = 1.0;
      for (int idx = tid; idx < n * n; idx += nth) s.Inv[idx] = (idx / n == idx % n) ? 1.0 : 0.0;
      gsync<WARP>();
      for (int k = 0; k < n; k++) {
        if (tid < 32) {
          double best = -1.0;
          int brow = k;
          for (int r = k + tid; r < n; r += 32) {
            const double v = fabs(s.A[(size_t)r * n + k]);
            if (v > best) { best = v; brow = r; }
          }
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) {
            const double ob = __shfl_down_sync(0xffffffffu, best, o);
            const int orow = __shfl_down_sync(0xffffffffu, brow, o);
            if (ob > best || (ob == best && orow < brow)) { best = ob; brow = orow; }
          }
          if (tid == 0) {
            piv_row = brow;
            if (!(best > 0.0) || !isfinite(best)) { atomicAdd(a.singular, 1); piv_inv = 0.0; }
            else piv_inv = 1.0 / s.A[(size_t)brow * n + k];
          }
        }
        gsync<WARP>();
        const int p = piv_row;
        const double pinv = piv_inv;
        // swap rows k and p and scale the pivot row, in A and in the inverse
        for (int cidx = tid; cidx < 2 * n; cidx += nth) {
          double* M = cidx < n ? s.A : s.Inv;
          const int cc = cidx < n ? cidx : cidx - n;
          const double vk = M[(size_t)k * n + cc], vp = M[(size_t)p * n + cc];
          M[(size_t)p * n + cc] = vk;
          M[(size_t)k * n + cc] = vp * pinv;
        }
        gsync<WARP>();
        // eliminate column k from every other row (the multipliers are read before anybody overwrites column k)
        for (int idx = tid; idx < n * 2 * n; idx += nth) {
          const int r = idx / (2 * n), cidx = idx % (2 * n);
          if (r == k) continue;
          const double m = s.A[(size_t)r * n + k];
          if (cidx < n) { if (cidx != k) s.A[(size_t)r * n + cidx] -= m * s.A[(size_t)k * n + cidx]; }
          else s.Inv[(size_t)r * n + cidx - n] -= m * s.Inv[(size_t)k * n + cidx - n];
        }
        gsync<WARP>();
        for (int r = tid; r < n; r += nth)
          if (r != k) s.A[(size_t)r * n + k] = 0.0;
        gsync<WARP>();
      }
      for (int idx = tid; idx < n * n; idx += nth) invT[(size_t)(idx % n) * n + idx / n] = s.Inv[idx];
      gsync<WARP>();  // this CTA's stores are visible to its own threads
    }
    // sol = A^-1 b (identical arithmetic when building and when applying the cache)
    double acc = 0.0;
    if (tid < n)
      for (int j = 0; j < n; j++) acc = fma(__ldcg(invT + (size_t)j * n + tid), s.b[j], acc);
    gsync<WARP>();
    if (tid < n) s.b[tid] = acc;
    gsync<WARP>();
  }

  // ------------------------------------------------------------------ CalcTentElEval
  const int lane = tid & 31, wid = tid >> 5, nw = nth >> 5;
  for (int te = a.te_off[tent]; te < a.te_off[tent + 1]; te++) {
    const int el = a.te_el[te];
    const int dom = a.te_dom[te];
    const double c = a.c_el[el];
    const double* sol = s.b + dom * Np;
    double* wf = a.wavefront + (size_t)el * a.ncomp * Q;
    const double* tau = a.te_ttop + (size_t)te * d1;
    double P[d1][D];
#pragma unroll
    for (int k = 0; k < d1; k++) {
      const int vk = a.el2v[(size_t)el * d1 + k];
#pragma unroll
      for (int d = 0; d < D; d++) P[k][d] = a.pts[(size_t)vk * D + d];
    }
    for (int q0 = 0; q0 < Q; q0 += Qc) {
      const int nq = min(Qc, Q - q0);
      for (int idx = tid; idx < nq * d1; idx += nth) {
        const int q = idx / d1, r = idx % d1;
        const double* lm = a.lam + (size_t)(q0 + q) * d1;
        double x = 0.0;
        if (r < D) { for (int k = 0; k < d1; k++) x += lm[k] * P[k][r]; }
        else { for (int k = 0; k < d1; k++) x += lm[k] * tau[k]; }
        s.pt[q * d1 + r] = x;
      }
      gsync<WARP>();
      eval_basis<D, WARP>(a, s, nq, ctr, sc, c, vlo_val, s.E0);
      const int nk = (nv - vlo_val) * nq;
      for (int kk = wid; kk < nk; kk += nw) {
        const int v = vlo_val + kk / nq, q = kk % nq;
        const double* e = s.E0 + (size_t)(v * Qc + q) * Np;
        double acc = 0.0;
        for (int i = lane; i < N; i += 32) acc = fma(e[i], sol[i], acc);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) acc += __shfl_down_sync(0xffffffffu, acc, o);
        if (lane == 0) wf[(size_t)(v - fo) * Q + q0 + q] = acc;
      }
      gsync<WARP>();
    }
  }
}


// ---------------------------------------------------------------------------------------------------------------------
// gwarp == 2: 1-D tents with stored inverses, one warp per tent, lanes = basis functions (BASELINE configs[1]).
// The variant above runs the CTA code on a warp and is bound by its index arithmetic (a division per operand entry) and
// by the monomial tables it fills for every function. Here lane i of a half-warp owns basis function i of one element
// of the tent (N <= 16; the two elements of a 1-D tent side by side in the two half-warps): it keeps its CSR row in
// registers, evaluates its own function and derivatives at the Q points of a face from the power tables of that face
// (pairs (xi^k, k xi^(k-1)) per point and direction in shared memory, filled by the half-warp's point lanes) and
// accumulates its own entry of the right-hand side; the product with the stored inverse is one row per lane; the
// evaluation reduces the 3 Q outputs of an element in one transposed 16-lane shuffle tree.
// Every floating-point expression and every summation order is the generic kernel's (the same statements on the same
// operands; the shuffle tree of the evaluation is the 32-lane tree with its upper half known to be zero), so the result
// is bitwise the CTA variant's (tests/test_gpu_generic_cache.py; DESIGN.md section 4.2 has the measurements that led
// from the first version, 8.3 ms per slab of BASELINE configs[1], to 3.6 ms).
constexpr int kSmallQ = 8, kSmallP1 = 8;
// Tables of (xi^k, k xi^(k-1)) pairs, laid out [k][q] with an odd row length: the lanes that fill them (one point q each,
// the same k) store consecutive 16-byte words, and the lanes that read them (the same q, different k) hit different banks
// (k * 9 * 16 bytes apart). r05l: with [q][k] rows of 128 bytes the stores were 7- to 14-way bank conflicts and the L1 data
// pipe was the busiest unit of the kernel (73 % of its peak).
constexpr int kSmallQS = 9;
constexpr int kSmallTab = kSmallP1 * kSmallQS;                   // pairs per direction
constexpr int kSmallHalfDoubles = 4 * kSmallTab + 3 * kSmallQ;   // (px, dpx) pairs | (pt, dpt) pairs | rr
constexpr int kSmallWarpDoubles = 32 + 2 * kSmallHalfDoubles;    // b | half 0 | half 1

constexpr int kSmallNnz = 4;   // a 1-D wave polynomial of degree <= 7 has at most four monomials
struct LaneRow {
  double cv[kSmallNnz];   // zero in unused slots (exponents 0): such a slot adds 0.0 * 1.0 to the sums below
  int ea[kSmallNnz], eb[kSmallNnz];   // exponents of x and t times the table row length
};

// value, d/dx, d/dt of the lane's basis function at point q of the half-warp's face: eval_basis for one (q, i) with the
// monomials formed on the fly. tab: px[q][k] = xi_x^k, pt[q][k] = xi_t^k, dpx[q][k] = (double)k * px[q][k-1], dpt likewise
// (the first factor of the generic kernel's derivative, formed once per point instead of once per function).
__device__ __forceinline__ void lane_basis(const LaneRow& r, const double* __restrict__ tab, int q, double sx, double st,
                                           double& e0, double& e1, double& e2) {
  // (power, derivative factor) pairs: one 16-byte shared-memory load per direction and monomial
  const double2* __restrict__ px = reinterpret_cast<const double2*>(tab) + q;
  const double2* __restrict__ pt = px + kSmallTab;
  double v0 = 0.0, v1 = 0.0, v2 = 0.0;
  // No branches: an exponent 0 reads dp[0] = 0.0, so the derivative is (+-0.0) where the generic kernel has 0.0, and an
  // unused slot contributes 0.0 * 1.0. Adding a zero of either sign leaves the sums unchanged (they start at +0.0 and can
  // never be -0.0), so the values are bitwise the generic kernel's.
#pragma unroll
  for (int j = 0; j < kSmallNnz; j++) {
    const double2 xa = px[r.ea[j]], tb = pt[r.eb[j]];
    double f[2];
    double full = 1.0;
    f[0] = xa.x; full *= f[0];
    f[1] = tb.x; full *= f[1];
    double dx = xa.y;
    dx *= f[1];
    double dt = tb.y;
    dt *= f[0];
    v0 += r.cv[j] * full;
    v1 += r.cv[j] * dx;
    v2 += r.cv[j] * dt;
  }
  v1 *= sx;
  v2 *= st;
  e0 = v0; e1 = v1; e2 = v2;
}

template <int P>   // order; Q = P + 1 points per face, P + 1 powers per direction
__global__ void __launch_bounds__(32 * kGenericWarpsPerCta, 6) tent_kernel_small1d(TentKernelArgs a) {
  constexpr int D = 1, d1 = 2, nv = 3;
  __shared__ __align__(16) double S_all[kGenericWarpsPerCta][kSmallWarpDoubles];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int slot = blockIdx.x * kGenericWarpsPerCta + wid;
  if (slot >= a.ntents) return;   // whole warps; no block-wide barrier in this kernel
  const unsigned full_mask = 0xffffffffu;
  const int tent = a.tents[slot];
  const int half = lane >> 4, i = lane & 15;
  double* const b = S_all[wid];
  double* const tab = b + 32 + half * kSmallHalfDoubles;   // (px, dpx)[q][k] | (pt, dpt)[q][k]
  double* const rr = tab + 4 * kSmallTab;                  // [v][q]
  constexpr int Q = P + 1, P1 = P + 1;
  const int N = a.N, Np = a.Npad;
  const int fo = a.fosystem, vlo = fo ? 1 : 0;
  const int nd = a.ndom[tent];
  const int n = nd * Np;
  const double ctr[d1] = {a.center[(size_t)tent * d1], a.center[(size_t)tent * d1 + 1]};
  const double sc = a.scale[tent];
  // the tent's stored inverse towards L2 now (rows of n doubles, 128 bytes per prefetch): it is read after the right-hand side
  const double* const invT = a.ginv + (size_t)tent * ((size_t)a.max_ndom * Np) * ((size_t)a.max_ndom * Np);
  if (lane < n) {
    asm volatile("prefetch.global.L2 [%0];" ::"l"(invT + (size_t)lane * n));
    if (n > 16) asm volatile("prefetch.global.L2 [%0];" ::"l"(invT + (size_t)lane * n + 16));
  }

  LaneRow row;
#pragma unroll
  for (int j = 0; j < kSmallNnz; j++) { row.cv[j] = 0.0; row.ea[j] = 0; row.eb[j] = 0; }
  if (i < N) {
    const int r0 = a.rowptr[i];
    const int nnz = a.rowptr[i + 1] - r0;
#pragma unroll
    for (int j = 0; j < kSmallNnz; j++) {
      if (j < nnz) {
        row.cv[j] = a.val[r0 + j];
        const int c = a.col[r0 + j];
        row.ea[j] = a.expo[c * d1] * kSmallQS;       // table offsets (rows of kSmallQS pairs)
        row.eb[j] = a.expo[c * d1 + 1] * kSmallQS;
      }
    }
  }
  b[lane] = 0.0;
  __syncwarp();

  // tables of point (x, t) for the lane's q: p[k] = xi^k by repeated multiplication (eval_basis, first loop), dp[k] = k p[k-1]
  auto powers = [&](int q, double x, double t, double c) {
#pragma unroll
    for (int d = 0; d < d1; d++) {
      const double scale = (d == D) ? sc * c : sc;
      const double xi = ((d == 0 ? x : t) - ctr[d]) * scale;
      double2* p = reinterpret_cast<double2*>(tab) + d * kSmallTab + q;   // (xi^k, k xi^(k-1)) at [k][q]
      double v = 1.0;
      p[0] = make_double2(1.0, 0.0);
#pragma unroll
      for (int k = 1; k < P1; k++) {
        const double dk = (double)k * v;
        v *= xi;
        p[k * kSmallQS] = make_double2(v, dk);
      }
    }
  };
  // acc = sum_{v in [v0, 3), q < Q} E[v][q] rr[v][q] for the lane's function, v outer, q inner (gemv)
  auto rhs_entry = [&](int v0, double c) {
    double E[nv][Q];
    const double st = sc * c;
#pragma unroll
    for (int q = 0; q < Q; q++) lane_basis(row, tab, q, sc, st, E[0][q], E[1][q], E[2][q]);
    double acc = 0.0;
#pragma unroll
    for (int v = 0; v < nv; v++) {
      if (v < v0) continue;
#pragma unroll
      for (int q = 0; q < Q; q++) acc = fma(E[v][q], rr[v * kSmallQ + q], acc);
    }
    return acc;
  };

  // ------------------------------------------------------------------ boundary faces (right-hand side only), half-warp 0
  for (int tf = a.tf_off[tent]; tf < a.tf_off[tent + 1]; tf++) {
    const int kind = a.tf_kind[tf];
    if (kind == 2) continue;  // material interfaces contribute to the matrix only
    const double* V = a.tf_verts + (size_t)tf * d1 * d1;
    double nrm[D];
    nrm[0] = a.tf_n[(size_t)tf * D];
    const double area = a.tf_area[tf];
    const int din = a.tf_dom[2 * tf];
    const double c0 = a.tf_c[2 * tf];
    if (half == 0 && i < Q) {
      const int q = i;
      double xt[d1];
#pragma unroll
      for (int r = 0; r < d1; r++) {
        double x = V[r];
        for (int m = 0; m < D; m++) x += (V[(m + 1) * d1 + r] - V[r]) * a.lam[(size_t)q * d1 + m];
        xt[r] = x;
      }
      powers(q, xt[0], xt[1], c0);
      const long long off = a.tf_data[tf];
      const double om = a.w[q] * area;
      if (kind == 0) {  // Dirichlet, twavetents.cpp:404-433
        const double alpha = 0.5, gam = 1.0 / (c0 * c0);
        const double g = a.bnd_data ? a.bnd_data[off + q] : 0.0;
        rr[(1 + D) * kSmallQ + q] = om * alpha * gam * g;
#pragma unroll
        for (int d = 0; d < D; d++) rr[(1 + d) * kSmallQ + q] = nrm[d] * om * g;
      } else {  // Neumann, twavetents.cpp:380-403
        const double beta = 0.5;
        double gn = 0.0;
#pragma unroll
        for (int r = 0; r < D; r++) gn += -nrm[r] * (a.bnd_data ? a.bnd_data[off + (long long)r * Q + q] : 0.0);
#pragma unroll
        for (int d = 0; d < D; d++) rr[(1 + d) * kSmallQ + q] = -nrm[d] * beta * gn * om;
        rr[(1 + D) * kSmallQ + q] = -gn * om;
      }
    }
    __syncwarp();
    if (half == 0 && i < N) {
      const double acc = rhs_entry(1, c0);
      b[din * Np + i] += acc;
    }
    __syncwarp();
  }

  // ------------------------------------------------------------------ bottom faces: right-hand side (CalcTentEl, 209-267)
  const int te0 = a.te_off[tent], te1 = a.te_off[tent + 1];
  for (int tp = te0; tp < te1; tp += 2) {
    const int te = tp + half;
    const bool have = te < te1;
    int dom = 0;
    double c = 1.0;
    if (have) {
      const int el = a.te_el[te];
      dom = a.te_dom[te];
      c = a.c_el[el];
      if (i < Q) {
        const int q = i;
        const double* geo = a.te_geo + (size_t)te * 2 * (d1 + 1);
        const double* wf = a.wavefront + (size_t)el * a.ncomp * Q;
        const double* tau = a.te_tbot + (size_t)te * d1;
        const double* nn = geo;
        const double area = nn[d1];
        double P[d1][D];
#pragma unroll
        for (int k = 0; k < d1; k++) {
          const int vk = a.el2v[(size_t)el * d1 + k];
          P[k][0] = a.pts[(size_t)vk * D];
        }
        const double* lm = a.lam + (size_t)q * d1;
        double xt[d1];
#pragma unroll
        for (int r = 0; r < d1; r++) {
          double x = 0.0;
          if (r < D) { for (int k = 0; k < d1; k++) x += lm[k] * P[k][r]; }
          else { for (int k = 0; k < d1; k++) x += lm[k] * tau[k]; }
          xt[r] = x;
        }
        powers(q, xt[0], xt[1], c);
        const double om = a.w[q] * area;
        const double gt = wf[(size_t)((!fo) + D) * Q + q];
        double rt = nn[D] / (c * c) * om * gt;
#pragma unroll
        for (int d = 0; d < D; d++) {
          const double gx = wf[(size_t)((!fo) + d) * Q + q];
          rr[(1 + d) * kSmallQ + q] = -(nn[D] * om * gx - nn[d] * om * gt);
          rt -= nn[d] * om * gx;
        }
        rr[(1 + D) * kSmallQ + q] = -rt;
        if (!fo) rr[q] = om * wf[q];
      }
    }
    __syncwarp();
    double acc = 0.0;
    if (have && i < N) acc = rhs_entry(vlo, c);
    // b += acc in the order of the tent's elements
    if (half == 0 && i < N) b[dom * Np + i] += acc;
    __syncwarp();
    if (half == 1 && have && i < N) b[dom * Np + i] += acc;
    __syncwarp();
  }

  // ------------------------------------------------------------------ sol = A^-1 b, one row per lane
  {
    double acc = 0.0;
#pragma unroll
    for (int j0 = 0; j0 < 32; j0 += 16) {
      if (j0 >= n) break;
      double col[16];   // sixteen independent loads in flight, then the chain in the order of the generic kernel
#pragma unroll
      for (int j = 0; j < 16; j++) col[j] = (lane < n && j0 + j < n) ? __ldcg(invT + (size_t)(j0 + j) * n + lane) : 0.0;
#pragma unroll
      for (int j = 0; j < 16; j++)
        if (j0 + j < n) acc = fma(col[j], b[j0 + j], acc);
    }
    __syncwarp();
    if (lane < n) b[lane] = acc;
    __syncwarp();
  }

  // ------------------------------------------------------------------ CalcTentElEval (541-590) at the top faces
  for (int tp = te0; tp < te1; tp += 2) {
    const int te = tp + half;
    const bool have = te < te1;
    int dom = 0, el = 0;
    double c = 1.0;
    if (have) {
      el = a.te_el[te];
      dom = a.te_dom[te];
      c = a.c_el[el];
      if (i < Q) {
        const int q = i;
        const double* tau = a.te_ttop + (size_t)te * d1;
        double P[d1][D];
#pragma unroll
        for (int k = 0; k < d1; k++) {
          const int vk = a.el2v[(size_t)el * d1 + k];
          P[k][0] = a.pts[(size_t)vk * D];
        }
        const double* lm = a.lam + (size_t)q * d1;
        double xt[d1];
#pragma unroll
        for (int r = 0; r < d1; r++) {
          double x = 0.0;
          if (r < D) { for (int k = 0; k < d1; k++) x += lm[k] * P[k][r]; }
          else { for (int k = 0; k < d1; k++) x += lm[k] * tau[k]; }
          xt[r] = x;
        }
        powers(q, xt[0], xt[1], c);
      }
    }
    __syncwarp();
    const double sol = (have && i < N) ? b[dom * Np + i] : 0.0;
    double* wf = a.wavefront + (size_t)el * a.ncomp * Q;
    // t[v * Q + q] = this lane's term of output (v, q); lanes >= N hold a zero row, halves without an element garbage
    // that never leaves the half-warp and is not written
    double t[24];
#pragma unroll
    for (int k = 3 * Q; k < 24; k++) t[k] = 0.0;
#pragma unroll
    for (int q = 0; q < Q; q++) {
      double e[nv];
      lane_basis(row, tab, q, sc, sc * c, e[0], e[1], e[2]);
#pragma unroll
      for (int v = 0; v < nv; v++) {
        double acc = fma(e[v], sol, 0.0);
        acc += 0.0;   // the step o = 16 of the 32-lane tree of the generic kernel: lanes 16.. hold zero (N <= 16)
        t[v * Q + q] = acc;
      }
    }
    // The 16-lane tree (partners 8, 4, 2, 1 apart) for all outputs at once: at every level a lane keeps one half of its
    // values and hands the other half to its partner, so 12 + 6 + 3 + 2 exchanges replace 4 per output. Each output still
    // is ((t_c + t_c+8) + (t_c+4 + t_c+12)) + ..., i.e. the pairs and the order of the generic kernel's shuffle tree (an
    // addition with swapped operands gives the same bits).
#pragma unroll
    for (int k = 0; k < 12; k++) {
      const bool up = (i & 8) != 0;
      const double recv = __shfl_xor_sync(full_mask, up ? t[k] : t[k + 12], 8, 16);
      t[k] = (up ? t[k + 12] : t[k]) + recv;
    }
#pragma unroll
    for (int k = 0; k < 6; k++) {
      const bool up = (i & 4) != 0;
      const double recv = __shfl_xor_sync(full_mask, up ? t[k] : t[k + 6], 4, 16);
      t[k] = (up ? t[k + 6] : t[k]) + recv;
    }
#pragma unroll
    for (int k = 0; k < 3; k++) {
      const bool up = (i & 2) != 0;
      const double recv = __shfl_xor_sync(full_mask, up ? t[k] : t[k + 3], 2, 16);
      t[k] = (up ? t[k + 3] : t[k]) + recv;
    }
    t[3] = 0.0;
#pragma unroll
    for (int k = 0; k < 2; k++) {
      const bool up = (i & 1) != 0;
      const double recv = __shfl_xor_sync(full_mask, up ? t[k] : t[k + 2], 1, 16);
      t[k] = (up ? t[k + 2] : t[k]) + recv;
    }
    // lane i now holds outputs base + 0 (and base + 1 on even lanes)
    const int base = ((i & 8) ? 12 : 0) + ((i & 4) ? 6 : 0) + ((i & 2) ? 3 : 0) + ((i & 1) ? 2 : 0);
#pragma unroll
    for (int k = 0; k < 2; k++) {
      const int idx = base + k;
      const int v = idx / Q, q = idx % Q;
      if (have && idx < 3 * Q && (k == 0 || !(i & 1)) && v >= vlo) wf[(size_t)(v - fo) * Q + q] = t[k];
    }
    __syncwarp();
  }
}

}  // namespace

// sizes the 1-D lanes-as-functions kernel is compiled for (gnnz: longest CSR row of the basis, set by the solver)
bool tent_small1d_supported(const TentKernelArgs& a) {
  return a.gnnz >= 1 && a.gnnz <= kSmallNnz && a.Npad <= 16 && a.order >= 1 && a.order <= 7 && a.Q == a.order + 1 && a.Qc >= a.Q &&
         a.max_ndom * a.Npad <= 32;
}

size_t tent_generic_warp_smem_bytes(int D, const TentKernelArgs& a) {
  return (size_t)kGenericWarpsPerCta * generic_warp_doubles(D, a) * sizeof(double);
}

size_t tent_generic_smem_bytes(int D, const TentKernelArgs& a) {
  const int d1 = D + 1, nv = D + 2;
  const size_t nmax = (size_t)a.max_ndom * a.Npad;
  const size_t esz = (size_t)nv * a.Qc * a.Npad;
  size_t doubles = nmax * nmax + nmax + 3 * esz + (size_t)nv * a.Qc + (size_t)a.Qc * d1 * (a.order + 1) +
                   (size_t)a.Qc * nv * a.npoly + (size_t)a.Qc * d1;
  if (a.gmode == 1) doubles += nmax * nmax;  // [A | I] while the inverses are formed
  return doubles * sizeof(double);
}

namespace {
template <int D>
cudaError_t launch_warp(const TentKernelArgs& a, cudaStream_t stream) {
  const size_t smem = tent_generic_warp_smem_bytes(D, a);
  cudaError_t err = cudaFuncSetAttribute(tent_kernel_generic<D, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (err != cudaSuccess) return err;
  const int ctas = (a.ntents + kGenericWarpsPerCta - 1) / kGenericWarpsPerCta;
  tent_kernel_generic<D, true><<<ctas, 32 * kGenericWarpsPerCta, smem, stream>>>(a);
  return cudaGetLastError();
}
}  // namespace

cudaError_t launch_tent_generic(int D, const TentKernelArgs& a, int threads, size_t smem, cudaStream_t stream) {
  if (a.ntents <= 0) return cudaSuccess;
  cudaError_t err;
  if (a.gmode == 2 && a.gwarp == 2) {  // stored inverses, 1-D: one warp per tent, lanes = basis functions
    if (D != 1 || !tent_small1d_supported(a)) return cudaErrorInvalidValue;
    const int ctas = (a.ntents + kGenericWarpsPerCta - 1) / kGenericWarpsPerCta;
    switch (a.order) {
      case 1: tent_kernel_small1d<1><<<ctas, 32 * kGenericWarpsPerCta, 0, stream>>>(a); break;
      case 2: tent_kernel_small1d<2><<<ctas, 32 * kGenericWarpsPerCta, 0, stream>>>(a); break;
      case 3: tent_kernel_small1d<3><<<ctas, 32 * kGenericWarpsPerCta, 0, stream>>>(a); break;
      case 4: tent_kernel_small1d<4><<<ctas, 32 * kGenericWarpsPerCta, 0, stream>>>(a); break;
      case 5: tent_kernel_small1d<5><<<ctas, 32 * kGenericWarpsPerCta, 0, stream>>>(a); break;
      case 6: tent_kernel_small1d<6><<<ctas, 32 * kGenericWarpsPerCta, 0, stream>>>(a); break;
      default: tent_kernel_small1d<7><<<ctas, 32 * kGenericWarpsPerCta, 0, stream>>>(a); break;
    }
    return cudaGetLastError();
  }
  if (a.gmode == 2 && a.gwarp) {  // stored inverses, one warp per tent
    switch (D) {
      case 1: return launch_warp<1>(a, stream);
      case 2: return launch_warp<2>(a, stream);
      case 3: return launch_warp<3>(a, stream);
      default: return cudaErrorInvalidValue;
    }
  }
  switch (D) {
    case 1:
      err = cudaFuncSetAttribute(tent_kernel_generic<1, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      if (err != cudaSuccess) return err;
      tent_kernel_generic<1, false><<<a.ntents, threads, smem, stream>>>(a);
      break;
    case 2:
      err = cudaFuncSetAttribute(tent_kernel_generic<2, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      if (err != cudaSuccess) return err;
      tent_kernel_generic<2, false><<<a.ntents, threads, smem, stream>>>(a);
      break;
    case 3:
      err = cudaFuncSetAttribute(tent_kernel_generic<3, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      if (err != cudaSuccess) return err;
      tent_kernel_generic<3, false><<<a.ntents, threads, smem, stream>>>(a);
      break;
    default:
      return cudaErrorInvalidValue;
  }
  return cudaGetLastError();
}

}  // namespace ngst
