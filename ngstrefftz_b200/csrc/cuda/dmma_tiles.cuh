// FP64 tensor-path (mma.sync m8n8k4 f64, SASS DMMA.8x8x4) tile helpers shared by the tent kernels (tent_fast.cu,
// tent_qt_fast.cu): symmetric rank-k updates Z Z^T on the upper-triangular 8x8 tiles, general updates ZL ZR^T on all
// tiles, accumulators in registers, fragments read from shared-memory operand rows Z[row][k].
#pragma once
#include <cuda_runtime.h>

namespace ngst {
namespace tiles {

constexpr int kChunk = 32;   // quadrature points per chunk (one per lane)
constexpr int kThreads = 128;
constexpr int kBS = 36;      // row stride of the boundary-pass operands (== 4 mod 16)

__device__ __forceinline__ void dmma_m8n8k4(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
               : "+d"(c0), "+d"(c1)
               : "d"(a), "d"(b));
}

constexpr int ipow(int b, int e) { return e == 0 ? 1 : b * ipow(b, e - 1); }
constexpr int cbinom(int n, int k) { return (k < 0 || k > n) ? 0 : (k == 0 ? 1 : cbinom(n - 1, k - 1) * n / k); }

// Symmetric rank-k update of this warp's upper-triangular tiles from the operand rows Z[function][k-row]:
// tile L (row-major over ti <= tj) belongs to tile half L & 1; k-steps ks = kh, kh+2, .. < nks.
template <int NT, int NACC, int TH>
__device__ __forceinline__ void mma_sym_step(double (&acc)[NACC][2], const double (&fr)[NT]) {
  int L = 0;
#pragma unroll
  for (int ti = 0; ti < NT; ti++)
#pragma unroll
    for (int tj = ti; tj < NT; tj++) {
      if ((L & 1) == TH) dmma_m8n8k4(acc[L >> 1][0], acc[L >> 1][1], fr[ti], fr[tj]);
      L++;
    }
}

// nks / 2 (the k-steps of one warp) must be even: fragments are double-buffered so that the shared-memory loads of
// the next k-step are in flight while the DMMAs of the current one issue.
template <int NT, int NACC, int TH>
__device__ __forceinline__ void dmma_sym(double (&acc)[NACC][2], const double* __restrict__ Z, int ZS, int nks, int kh,
                                         int g, int k) {
  const double* __restrict__ zp = Z + g * ZS + k + 4 * kh;
  double fa[NT], fb[NT];
#pragma unroll
  for (int t = 0; t < NT; t++) fa[t] = zp[8 * t * ZS];
  for (int ks = kh; ks < nks; ks += 4) {
#pragma unroll
    for (int t = 0; t < NT; t++) fb[t] = zp[8 * t * ZS + 8];
    mma_sym_step<NT, NACC, TH>(acc, fa);
    zp += 16;
    if (ks + 4 < nks) {
#pragma unroll
      for (int t = 0; t < NT; t++) fa[t] = zp[8 * t * ZS];
    }
    mma_sym_step<NT, NACC, TH>(acc, fb);
  }
}

// acc -> A (upper-triangular tiles of this warp's half); add = false stores, true accumulates.
template <int NT, int NACC, int TH>
__device__ __forceinline__ void flush_sym(const double (&acc)[NACC][2], double* A, int LDA, bool add, int g, int k) {
  int L = 0;
#pragma unroll
  for (int ti = 0; ti < NT; ti++)
#pragma unroll
    for (int tj = ti; tj < NT; tj++) {
      if ((L & 1) == TH) {
        double* p = A + (8 * ti + g) * LDA + 8 * tj + 2 * k;
        if (add) { p[0] += acc[L >> 1][0]; p[1] += acc[L >> 1][1]; }
        else { p[0] = acc[L >> 1][0]; p[1] = acc[L >> 1][1]; }
      }
      L++;
    }
}

// General (non-symmetric) update A += ZL ZR^T over 8 k-steps; tile L = ti*NT + tj belongs to warp L & 3 (= W).
template <int NT, int NACC, int W>
__device__ __forceinline__ void dmma_full(double (&acc)[NACC][2], const double* __restrict__ ZL,
                                          const double* __restrict__ ZR, int g, int k) {
  for (int ks = 0; ks < kChunk / 4; ks++) {
    double fl[NT], fr[NT];
#pragma unroll
    for (int t = 0; t < NT; t++) {
      fl[t] = ZL[(8 * t + g) * kBS + 4 * ks + k];
      fr[t] = ZR[(8 * t + g) * kBS + 4 * ks + k];
    }
    int L = 0;
#pragma unroll
    for (int ti = 0; ti < NT; ti++)
#pragma unroll
      for (int tj = 0; tj < NT; tj++) {
        if ((L & 3) == W) dmma_m8n8k4(acc[L >> 2][0], acc[L >> 2][1], fl[ti], fr[tj]);
        L++;
      }
  }
}

template <int NT, int NACC, int W>
__device__ __forceinline__ void flush_full(const double (&acc)[NACC][2], double* A, int LDA, int g, int k) {
  int L = 0;
#pragma unroll
  for (int ti = 0; ti < NT; ti++)
#pragma unroll
    for (int tj = 0; tj < NT; tj++) {
      if ((L & 3) == W) {
        double* p = A + (8 * ti + g) * LDA + 8 * tj + 2 * k;
        p[0] += acc[L >> 2][0];
        p[1] += acc[L >> 2][1];
      }
      L++;
    }
}


}  // namespace tiles
}  // namespace ngst
