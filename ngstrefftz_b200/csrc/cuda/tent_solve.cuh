// In-register Gauss-Jordan solve of the per-tent system (TWaveTents<D>::Solve, src/twavetents.cpp:64-70, forms the
// inverse with LAPACK and multiplies; the solution differs from that by O(cond * eps)).
//
// The whole augmented system [A | b] is held in REGISTERS: thread (ty, tx) of an 8 x 16 grid owns the rows ty + 8 i and
// the columns tx + 16 j (NR x NCB doubles). Per step the owners publish the pivot row and the pivot column through
// (double-buffered) shared memory, then every thread updates its block with NR * NCB FMAs and no further memory traffic;
// eliminating above the diagonal as well costs nothing extra in this layout and removes the serial back substitution.
// spd = true (symmetric positive definite system): natural order, one barrier per step. Otherwise partial pivoting with
// an implicit row permutation (the search is done redundantly by every warp), two barriers per step.
#pragma once
#include <cuda_runtime.h>

namespace ngst {

// A: N x N row-major with leading dimension LDA (shared memory), bv: right-hand side in, solution out (shared memory).
// Must be called by all THREADS (= 128) threads of the CTA; A and bv must be complete (barrier before the call). The
// caller synchronises before reading bv. *singular is incremented for every zero (spd: non-positive) pivot.
template <int N, int NP, int THREADS>
__device__ __forceinline__ void gauss_jordan_solve(const double* A, int LDA, double* bv, bool spd, int* singular) {
  static_assert(THREADS == 128, "thread grid is 8 x 16");
  static_assert(NP % 8 == 0 && NP >= N, "padded size");
  const int tid = threadIdx.x, lane = tid & 31;
  {
    constexpr int NR = NP / 8, NCB = (N + 1 + 15) / 16;
    static_assert(N <= 64, "row mask is 64 bits wide");
    __shared__ double prow[2][16 * NCB];
    __shared__ double pcol[2][NP];
    __shared__ double pinv_s[2];
    __shared__ double pinv_row[NP];
    __shared__ int rowstep[NP];
    const int ty = tid >> 4, tx = tid & 15;
    double Ar[NR][NCB];
#pragma unroll
    for (int i = 0; i < NR; i++)
#pragma unroll
      for (int j = 0; j < NCB; j++) {
        const int r = ty + 8 * i, c = tx + 16 * j;
        Ar[i][j] = (r < N && c < N) ? A[r * LDA + c] : ((r < N && c == N) ? bv[r] : 0.0);
      }
    unsigned long long done = 0ull;
#pragma unroll
    for (int i0 = 0; i0 < NR; i0++) {
      const int j0 = i0 >> 1;  // column block of the steps kk = 8 i0 + k8
      for (int k8 = 0; k8 < 8; k8++) {
        const int kk = 8 * i0 + k8;
        if (kk >= N) break;
        const int buf = kk & 1;
        int p = kk;
        if (spd) {
          if (ty == k8) {
#pragma unroll
            for (int j = 0; j < NCB; j++) prow[buf][tx + 16 * j] = Ar[i0][j];
            if (tx == (kk & 15)) {
              const double piv = Ar[i0][j0];
              const bool ok = piv > 0.0 && isfinite(piv);
              pinv_s[buf] = ok ? 1.0 / piv : 0.0;
              pinv_row[kk] = ok ? 1.0 / piv : 0.0;
              rowstep[kk] = kk;
              if (!ok) atomicAdd(singular, 1);
            }
          }
          if (tx == (kk & 15)) {
#pragma unroll
            for (int i = 0; i < NR; i++) pcol[buf][ty + 8 * i] = Ar[i][j0];
          }
          __syncthreads();
        } else {
          if (tx == (kk & 15)) {
#pragma unroll
            for (int i = 0; i < NR; i++) pcol[buf][ty + 8 * i] = Ar[i][j0];
          }
          __syncthreads();
          double best = -1.0;
          int brow = 0x7fffffff;
          for (int r = lane; r < N; r += 32) {
            if ((done >> r) & 1ull) continue;
            const double v = fabs(pcol[buf][r]);
            if (v > best) { best = v; brow = r; }
          }
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) {
            const double ob = __shfl_xor_sync(0xffffffffu, best, o);
            const int orow = __shfl_xor_sync(0xffffffffu, brow, o);
            if (ob > best || (ob == best && orow < brow)) { best = ob; brow = orow; }
          }
          // every remaining entry NaN (degenerate geometry): no candidate was selected. Take the first row not yet used
          // so that the indices stay in range; the non-finite pivot is counted as singular below.
          if (brow >= N) brow = __ffsll((long long)~done) - 1;
          p = brow;
          done |= 1ull << p;
          if (ty == (p & 7)) {
            const int ip = p >> 3;
            double tmp[NCB];
#pragma unroll
            for (int j = 0; j < NCB; j++) tmp[j] = 0.0;
#pragma unroll
            for (int i = 0; i < NR; i++)
              if (i == ip) {
#pragma unroll
                for (int j = 0; j < NCB; j++) tmp[j] = Ar[i][j];
              }
#pragma unroll
            for (int j = 0; j < NCB; j++) prow[buf][tx + 16 * j] = tmp[j];
            if (tx == (kk & 15)) {
              const double piv = tmp[j0];
              const bool ok = piv != 0.0 && isfinite(piv);
              pinv_s[buf] = ok ? 1.0 / piv : 0.0;
              pinv_row[p] = ok ? 1.0 / piv : 0.0;
              rowstep[p] = kk;
              if (!ok) atomicAdd(singular, 1);
            }
          }
          __syncthreads();
        }
        const double pinv = pinv_s[buf];
        double m[NR], pr[NCB];
#pragma unroll
        for (int i = 0; i < NR; i++) {
          const double v = pcol[buf][ty + 8 * i];
          m[i] = (ty + 8 * i == p) ? 0.0 : -v * pinv;
        }
#pragma unroll
        for (int j = 0; j < NCB; j++) pr[j] = prow[buf][tx + 16 * j];
#pragma unroll
        for (int i = 0; i < NR; i++)
#pragma unroll
          for (int j = 0; j < NCB; j++) Ar[i][j] = fma(m[i], pr[j], Ar[i][j]);
      }
    }
    __syncthreads();
    // row r was the pivot row of step rowstep[r]: x[rowstep[r]] = [A | b][r][N] / pivot
    if (tx == (N & 15)) {
#pragma unroll
      for (int i = 0; i < NR; i++) {
        const int r = ty + 8 * i;
        if (r < N) bv[rowstep[r]] = Ar[i][N >> 4] * pinv_row[r];
      }
    }
  }
}

// The same elimination on [A | I]: the inverse, written transposed to invT (N columns of NP doubles, global memory).
// The reference forms the inverse as well (LAPACK, src/twavetents.cpp:64-70) and multiplies. 2N columns per thread row
// block: 7 x 7 doubles at N = 55 -- used once per slab geometry (factor cache, tent_fast.cu MODE 1).
template <int N, int NP, int THREADS>
__device__ __forceinline__ void gauss_jordan_invert(const double* A, int LDA, bool spd, int* singular, double* __restrict__ invT) {
  static_assert(THREADS == 128, "thread grid is 8 x 16");
  static_assert(NP % 8 == 0 && NP >= N, "padded size");
  const int tid = threadIdx.x, lane = tid & 31;
  {
    constexpr int NR = NP / 8, NCB = (2 * N + 15) / 16;
    static_assert(N <= 64, "row mask is 64 bits wide");
    __shared__ double prow[2][16 * NCB];
    __shared__ double pcol[2][NP];
    __shared__ double pinv_s[2];
    __shared__ double pinv_row[NP];
    __shared__ int rowstep[NP];
    const int ty = tid >> 4, tx = tid & 15;
    double Ar[NR][NCB];
#pragma unroll
    for (int i = 0; i < NR; i++)
#pragma unroll
      for (int j = 0; j < NCB; j++) {
        const int r = ty + 8 * i, c = tx + 16 * j;
        Ar[i][j] = (r < N && c < N) ? A[r * LDA + c] : ((r < N && c == N + r) ? 1.0 : 0.0);  // [A | I]
      }
    unsigned long long done = 0ull;
#pragma unroll
    for (int i0 = 0; i0 < NR; i0++) {
      const int j0 = i0 >> 1;  // column block of the steps kk = 8 i0 + k8
      for (int k8 = 0; k8 < 8; k8++) {
        const int kk = 8 * i0 + k8;
        if (kk >= N) break;
        const int buf = kk & 1;
        int p = kk;
        if (spd) {
          if (ty == k8) {
#pragma unroll
            for (int j = 0; j < NCB; j++) prow[buf][tx + 16 * j] = Ar[i0][j];
            if (tx == (kk & 15)) {
              const double piv = Ar[i0][j0];
              const bool ok = piv > 0.0 && isfinite(piv);
              pinv_s[buf] = ok ? 1.0 / piv : 0.0;
              pinv_row[kk] = ok ? 1.0 / piv : 0.0;
              rowstep[kk] = kk;
              if (!ok) atomicAdd(singular, 1);
            }
          }
          if (tx == (kk & 15)) {
#pragma unroll
            for (int i = 0; i < NR; i++) pcol[buf][ty + 8 * i] = Ar[i][j0];
          }
          __syncthreads();
        } else {
          if (tx == (kk & 15)) {
#pragma unroll
            for (int i = 0; i < NR; i++) pcol[buf][ty + 8 * i] = Ar[i][j0];
          }
          __syncthreads();
          double best = -1.0;
          int brow = 0x7fffffff;
          for (int r = lane; r < N; r += 32) {
            if ((done >> r) & 1ull) continue;
            const double v = fabs(pcol[buf][r]);
            if (v > best) { best = v; brow = r; }
          }
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) {
            const double ob = __shfl_xor_sync(0xffffffffu, best, o);
            const int orow = __shfl_xor_sync(0xffffffffu, brow, o);
            if (ob > best || (ob == best && orow < brow)) { best = ob; brow = orow; }
          }
          // every remaining entry NaN (degenerate geometry): no candidate was selected. Take the first row not yet used
          // so that the indices stay in range; the non-finite pivot is counted as singular below.
          if (brow >= N) brow = __ffsll((long long)~done) - 1;
          p = brow;
          done |= 1ull << p;
          if (ty == (p & 7)) {
            const int ip = p >> 3;
            double tmp[NCB];
#pragma unroll
            for (int j = 0; j < NCB; j++) tmp[j] = 0.0;
#pragma unroll
            for (int i = 0; i < NR; i++)
              if (i == ip) {
#pragma unroll
                for (int j = 0; j < NCB; j++) tmp[j] = Ar[i][j];
              }
#pragma unroll
            for (int j = 0; j < NCB; j++) prow[buf][tx + 16 * j] = tmp[j];
            if (tx == (kk & 15)) {
              const double piv = tmp[j0];
              const bool ok = piv != 0.0 && isfinite(piv);
              pinv_s[buf] = ok ? 1.0 / piv : 0.0;
              pinv_row[p] = ok ? 1.0 / piv : 0.0;
              rowstep[p] = kk;
              if (!ok) atomicAdd(singular, 1);
            }
          }
          __syncthreads();
        }
        const double pinv = pinv_s[buf];
        double m[NR], pr[NCB];
#pragma unroll
        for (int i = 0; i < NR; i++) {
          const double v = pcol[buf][ty + 8 * i];
          m[i] = (ty + 8 * i == p) ? 0.0 : -v * pinv;
        }
#pragma unroll
        for (int j = 0; j < NCB; j++) pr[j] = prow[buf][tx + 16 * j];
#pragma unroll
        for (int i = 0; i < NR; i++)
#pragma unroll
          for (int j = 0; j < NCB; j++) Ar[i][j] = fma(m[i], pr[j], Ar[i][j]);
      }
    }
    __syncthreads();
    // row r was the pivot row of step rowstep[r]: inverse[rowstep[r]][c] = [A | I][r][N + c] / pivot, stored transposed
    // (invT[c * NP + row]: consecutive threads of the matrix-vector product read consecutive doubles)
#pragma unroll
    for (int i = 0; i < NR; i++) {
      const int r = ty + 8 * i;
      if (r < N) {
        const double pr = pinv_row[r];
        const int row = rowstep[r];
#pragma unroll
        for (int j = 0; j < NCB; j++) {
          const int c = tx + 16 * j - N;
          if (c >= 0 && c < N) invT[(size_t)c * NP + row] = Ar[i][j] * pr;
        }
      }
    }
  }
}

}  // namespace ngst
