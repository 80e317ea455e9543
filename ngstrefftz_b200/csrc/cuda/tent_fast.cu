// Fast per-tent Trefftz-DG kernel for single-wavespeed tents (ndomains == 1), sm_100a.
//
// Same mathematics as tent_generic.cu / TWaveTents<D>::Propagate's per-tent body (src/twavetents.cpp:92-171:
// CalcTentEl 177-333, CalcTentBndEl 335-434, Solve 64-70, CalcTentElEval 541-590), re-organised for the B200:
//   * one CTA of 4 warps per tent, three CTAs per SM (168 registers, 61 KB shared memory at D = 3, p = 4);
//   * the flat list of (element, quadrature point) pairs of the tent is processed in chunks of 32 points:
//     lane = point, warp = one quarter of the basis functions. Each warp evaluates its functions with the
//     generated straight-line evaluators (basis_eval_gen.cuh; the reference evaluates the same polynomials through
//     CSR tables, src/scalarmappedfe.cpp:437-674) and writes the weighted operand Z[function][k-row] to shared memory;
//   * the Gram contractions elmat += B W B^T (top faces, src/twavetents.cpp:293-312) and elmat += S diag(w) S^T
//     (bottom faces, 255-262) run on the FP64 tensor path (mma.sync m8n8k4 f64, "DMMA"). W is symmetric positive
//     definite on a causal face, so with W = L L^T and Z = B L the update is the symmetric rank-k product Z Z^T:
//     only the upper-triangular 8x8 tiles are accumulated, in registers, split over the 4 warps as
//     (tile half) x (k-step parity);
//   * right-hand sides are accumulated per lane in registers and reduced once per tent;
//   * boundary faces (Dirichlet / Neumann) are rank-1 per point, A += l r^T, and use the same DMMA machinery on all
//     tiles; pivoted LU in shared memory; the top-face evaluation runs in the monomial basis (one thread per point).
// Tents with material interfaces (ndomains > 1) are routed to the generic kernel by the host.
#include <cuda_runtime.h>

#include "../host/quadrature.hpp"
#include "basis_eval_gen.cuh"
#include "dmma_tiles.cuh"
#include "tent_solve.cuh"
#include "tma_bulk.cuh"
#include "tent_fast.cuh"
#include "tent_kernel.cuh"

namespace ngst {

namespace {

using namespace tiles;
constexpr int kMaxFinal = 64;  // final elements of one tent listed in shared memory (compact state)

template <int D, int P, int FO>
struct Cfg {
  static constexpr int d1 = D + 1;
  static constexpr int N = gen::BasisInfo<D, P, FO>::N;
  static constexpr int npoly = gen::BasisInfo<D, P, FO>::npoly;
  static constexpr int npm1 = cbinom(d1 + P - 1, P - 1);  // monomials of degree <= P-1
  static constexpr int NT = (N + 7) / 8;          // 8x8 tiles per dimension
  static constexpr int NP = NT * 8;               // padded system size
  static constexpr int G = NP / 4;                // functions per evaluation group (function i -> group i % 4)
  static constexpr int Q = (D == 1) ? (P + 1) : ipow(P + 1, D);  // points of the order-2P rule (quadrature.cpp)
  // The top-face Gram integrand (B L)(B L)^T is a polynomial of degree 2P-2 on the (affine) face: any rule exact to that
  // degree gives the order-2P rule's matrix up to round-off. Compact symmetric rule where the table has one (24 points
  // at D = 3, P = 4 instead of the 64 of the Duffy rule of order 2P-2 and the 125 of order 2P), else the Duffy rule.
  static constexpr int QTD = (D == 1) ? P : ipow(P, D);
  static constexpr int QT = compact_rule_points(D, 2 * P - 2) > 0 ? compact_rule_points(D, 2 * P - 2) : QTD;
  // Likewise the bottom-face stabilisation Gram S diag(w) S^T (degree 2P): compact rule of degree 2P in a pass of its
  // own where the table has one (the right-hand side keeps the order-2P rule: it contains the sampled state)
  static constexpr int QS = compact_rule_points(D, 2 * P);
  static constexpr bool splitStab = !FO && QS > 0 && QS < Q;
  static constexpr int ZS = kChunk * d1 + 4;      // row stride of Z: == 4 (mod 16) -> conflict-free fragment loads
  static constexpr int LDA = NP + 1;
  static constexpr int NSYM = NT * (NT + 1) / 2;  // upper-triangular tiles
  static constexpr int NACC = (NSYM + 1) / 2;     // accumulator tiles per warp (>= ceil(NT*NT/4) for NT >= 2)
  static constexpr int FE = fast_fe_stride(D);
  static constexpr int FF = fast_ff_stride(D);
  // shared memory (doubles). The system matrix A is only needed once the Gram passes are over, so where it fits it
  // aliases the tail of the operand buffer Z (behind the two boundary-pass operands): 61 KB per CTA at D = 3, p = 4,
  // three CTAs per SM.
  static constexpr bool aliasA = 2 * kBS + LDA <= ZS;
  static constexpr int oZ = 0;
  static constexpr int oA = aliasA ? oZ + 2 * NP * kBS : oZ + NP * ZS;
  static constexpr int oB = aliasA ? oZ + NP * ZS : oA + NP * LDA;
  static constexpr int oCM = (oB + NP + 1) / 2 * 2;
  static constexpr int oCMD = oCM + npoly;
  // right-hand side of the bottom faces when it is computed in a pass of its own (rhsFirst)
  static constexpr bool rhsFirst = FO || splitStab;
  static constexpr int oRHS = (oCMD + d1 * npm1 + 1) / 2 * 2;
  static constexpr int nsmem = oRHS + (rhsFirst ? NP : 0);
  static_assert(2 * kBS <= ZS, "boundary operands must fit in the Z buffer");
  static_assert(NT >= 2 && (NT * NT + 3) / 4 <= NACC, "accumulator budget");

  // ---- compact state rows staged in shared memory by bulk asynchronous copies (right-hand-side pass)
  static constexpr int NCS = FO ? d1 : d1 + 1;                       // state components
  static constexpr int ROWLEN = NCS * (QS > 0 ? QS : 1);             // doubles per compact row
  static constexpr bool stageRows = QS > 0 && ROWLEN % 2 == 0;       // bulk copies move multiples of 16 bytes
  // MODE 0 / 1: the operand buffer Z is free during the pass: rows behind the partial sums
  static constexpr int oRows01 = 4 * NP;
  static constexpr int capRows01 = (NP * ZS - oRows01) / ROWLEN < 32 ? (NP * ZS - oRows01) / ROWLEN : 32;
  // MODE 2 (apply): no operand buffer, no matrix. [partial sums | stored inverse | rows | b | cm | cmd | rhs], sized
  // so that three CTAs fit in the 227 KB of an SM
  static constexpr int oInv2 = 4 * NP;
  static constexpr int oRows2 = oInv2 + NP * N;
  static constexpr int tail2 = (NP + 2) + (npoly + 2) + (d1 * npm1 + 2) + NP;
  static constexpr int budget2 = 74 * 1024 / 8;
  static constexpr int capRows2raw = (budget2 - oRows2 - tail2) / ROWLEN;
  static constexpr int capRows2 = capRows2raw > 32 ? 32 : (capRows2raw < 1 ? 1 : capRows2raw);
  static constexpr int oB2 = (oRows2 + capRows2 * ROWLEN + 1) / 2 * 2;
  static constexpr int oCM2 = (oB2 + NP + 1) / 2 * 2;
  static constexpr int oCMD2 = oCM2 + npoly;
  static constexpr int oRHS2 = (oCMD2 + d1 * npm1 + 1) / 2 * 2;
  static constexpr int nsmem2 = oRHS2 + NP;
  // MODE 3 (apply, four CTAs per SM): the stored inverse is read from L2 by the product instead of being staged:
  // [partial sums | rows | b | cm | cmd | rhs] within 55 KB
  static constexpr int oRows3 = 4 * NP;
  static constexpr int budget3 = 55 * 1024 / 8;
  static constexpr int capRows3raw = (budget3 - oRows3 - tail2) / ROWLEN;
  static constexpr int capRows3 = capRows3raw > 32 ? 32 : (capRows3raw < 1 ? 1 : capRows3raw);
  static constexpr int oB3 = (oRows3 + capRows3 * ROWLEN + 1) / 2 * 2;
  static constexpr int oCM3 = (oB3 + NP + 1) / 2 * 2;
  static constexpr int oCMD3 = oCM3 + npoly;
  static constexpr int oRHS3 = (oCMD3 + d1 * npm1 + 1) / 2 * 2;
  static constexpr int nsmem3 = oRHS3 + NP;
};

// group dispatch: warp w evaluates the functions i with i % 4 == w
template <int D, int P, int FO, class F>
__device__ __forceinline__ void eval_group(int w, const double* xi, F&& f) {
  switch (w) {
    case 0: gen::BasisGroup<D, P, FO, 0>::eval(xi, f); break;
    case 1: gen::BasisGroup<D, P, FO, 1>::eval(xi, f); break;
    case 2: gen::BasisGroup<D, P, FO, 2>::eval(xi, f); break;
    default: gen::BasisGroup<D, P, FO, 3>::eval(xi, f); break;
  }
}

// all functions (H == 1) or the half of them in groups 2 part, 2 part + 1 (H == 2)
template <int D, int P, int FO, int H, class F>
__device__ __forceinline__ void eval_part(int part, const double* xi, F&& f) {
  if (H == 1) {
    gen::BasisGroup<D, P, FO, 0>::eval(xi, f);
    gen::BasisGroup<D, P, FO, 1>::eval(xi, f);
    gen::BasisGroup<D, P, FO, 2>::eval(xi, f);
    gen::BasisGroup<D, P, FO, 3>::eval(xi, f);
  } else if (part == 0) {
    gen::BasisGroup<D, P, FO, 0>::eval(xi, f);
    gen::BasisGroup<D, P, FO, 1>::eval(xi, f);
  } else {
    gen::BasisGroup<D, P, FO, 2>::eval(xi, f);
    gen::BasisGroup<D, P, FO, 3>::eval(xi, f);
  }
}

// CS = 1 (compact state): the tents read and write a second, internal state that holds every
// row at the QS points of the compact degree-2P rule instead of the Q points of the order-2P rule. A row written by a
// tent is a polynomial of degree <= P on its element, so the compact samples carry the same information, and for any
// row g the bottom-face integrals against degree-<=P test functions agree: sum_q w_q phi(x_q) g_q = sum_r v_r phi(y_r) (M g)_r
// with M the transfer through the degree-P polynomials (rule_transfer_matrix; applied once by SetInitial). The public
// Q-point rows are written only by the LAST tent of the slab that contains the element (f.te_final).
//
// MODE (factor cache): the tent matrix A depends on the geometry of the tent only, not on the
// state, and a slab is propagated many times (the reference re-assembles and re-inverts A on every Propagate,
// src/twavetents.cpp:119-164). MODE 0 assembles and solves. MODE 1 assembles, forms A^-1 (as the reference does) and
// stores it in HBM (f.ainv, transposed), then applies it. MODE 2 skips the assembly: right-hand side, sol = A^-1 b with
// the stored inverse (fetched by one bulk asynchronous copy behind the right-hand-side pass), evaluation.
template <int D, int P, int FO, int CS, int MODE>
__global__ void __launch_bounds__(kThreads, MODE == 1 ? 2 : (MODE == 3 ? 4 : 3)) tent_kernel_fast(TentKernelArgs a, FastArgs f) {
  using C = Cfg<D, P, FO>;
  static_assert(!CS || (C::QS > 0 && C::rhsFirst), "compact state needs the compact degree-2P rule");
  static_assert(MODE == 0 || CS, "the factor cache is built on the compact-state variant");
  constexpr int d1 = C::d1, N = C::N, NT = C::NT, NP = C::NP, G = C::G, Q = C::Q, QT = C::QT, ZS = C::ZS, LDA = C::LDA;
  constexpr int NACC = C::NACC, FE = C::FE, FF = C::FF, npoly = C::npoly, npm1 = C::npm1;
  extern __shared__ __align__(16) double sm[];
  double* const A = sm + C::oA;
  constexpr bool APPLY = MODE >= 2;  // MODE 3: as 2 with 128 registers, 55 KB of shared memory, four CTAs per SM
  double* const bv = sm + (MODE == 2 ? C::oB2 : (MODE == 3 ? C::oB3 : C::oB));
  double* const Z = sm + C::oZ;
  double* const cm = sm + (MODE == 2 ? C::oCM2 : (MODE == 3 ? C::oCM3 : C::oCM));
  double* const cmd = sm + (MODE == 2 ? C::oCMD2 : (MODE == 3 ? C::oCMD3 : C::oCMD));
  double* const rhsS = sm + (MODE == 2 ? C::oRHS2 : (MODE == 3 ? C::oRHS3 : C::oRHS));
  // compact rows of the tent's elements, fetched by bulk asynchronous copies (cp.async.bulk) ahead of the right-hand-side pass
  constexpr bool STG = CS && C::stageRows;
  constexpr int kRowCap = STG ? (MODE == 2 ? C::capRows2 : (MODE == 3 ? C::capRows3 : C::capRows01)) : (1 << 30);
  double* const rowsS = sm + (MODE == 2 ? C::oRows2 : (MODE == 3 ? C::oRows3 : C::oRows01));

  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  const int g = lane >> 2, k = lane & 3;  // DMMA fragment coordinates
  const int th = w >> 1, kh = w & 1;      // tile half / k-step parity of this warp in the symmetric passes
  int tent;
  if (a.ticket) {
    // dependency-driven order (replaces the reference's RunParallelDependency, ngstents paralleldepend.hpp:20-131): tickets
    // are handed out in level order, so every predecessor holds an earlier ticket and is running or done -- no deadlock.
    __shared__ int s_tent, s_stuck;
    if (tid == 0) {
      const int t = a.tents[atomicAdd(a.ticket, 1)];
      volatile int* r = a.remaining + t;
      long long spins = 0;
      int stuck = 0;
      while (*r > 0) {
        __nanosleep(64);
        // never hang the device: a tent that gives up is counted in its own flag (Propagate reports a scheduling failure,
        // not a singular system), skips its work and still releases its successors so that the launch drains quickly
        if (++spins > (1ll << 26)) { atomicAdd(a.stuck, 1); stuck = 1; break; }
      }
      if (CS && f.tent_chunk && !stuck) {
        // streamed input: the public rows this tent projects from have been delivered by the copy engine
        volatile const int* ld = f.chunk_loaded + f.tent_chunk[t];
        while (*ld == 0) {
          __nanosleep(256);
          if (++spins > (1ll << 26)) { atomicAdd(a.stuck, 1); stuck = 1; break; }
        }
      }
      if (f.p2p) __threadfence_system();  // rows and counters may have been written by another GPU
      else __threadfence();
      s_tent = t;
      s_stuck = stuck;
    }
    __syncthreads();
    tent = s_tent;
    if (s_stuck) {
      for (int s = a.succ_off[tent] + tid; s < a.succ_off[tent + 1]; s += kThreads) atomicSub(a.remaining + a.succ[s], 1);
      if (f.p2p)
        for (int s = f.p2p->rsucc_off[tent] + tid; s < f.p2p->rsucc_off[tent + 1]; s += kThreads)
          atomicSub_system(f.p2p->remaining[f.p2p->rsucc_peer[s]] + f.p2p->rsucc_tent[s], 1);
      return;
    }
  } else {
    tent = a.tents[blockIdx.x];
  }
  const int te0 = a.te_off[tent], nel = a.te_off[tent + 1] - te0;
  const int tf0 = a.tf_off[tent], nfc = a.tf_off[tent + 1] - tf0;
  const int npts = nel * Q;
  const int nptt = nel * QT;
  // stored inverse of this tent (MODE 1 writes, MODE 2 reads): N columns of NP doubles, behind the NS * NP partial sums of
  // the right-hand-side pass in the (otherwise unused) operand buffer
  double* const invS = sm + C::oInv2;
  [[maybe_unused]] double* const invG = (MODE != 0) ? f.ainv + (size_t)tent * (NP * N) : nullptr;
  __shared__ __align__(8) unsigned long long s_bar, s_bar_rows;
  if constexpr (APPLY || STG) {
    if (tid == 0) {
      tma::mbar_init(&s_bar, 1);
      tma::mbar_init(&s_bar_rows, 1);
    }
    __syncthreads();
  }
  if constexpr (MODE == 2) {
    if (tid == 0) {
      tma::mbar_expect(&s_bar, NP * N * 8);
      tma::bulk_load(invS, invG, NP * N * 8, &s_bar);
    }
  }
  if constexpr (MODE == 3) {  // the stored inverse towards L2 now; the product reads it from there after the first pass
    for (int i = tid * 16; i < NP * N; i += kThreads * 16) asm volatile("prefetch.global.L2 [%0];" ::"l"(invG + i));
  }

  if (C::rhsFirst) {
    // ---------------------------------------------------------------- bottom faces: right-hand side  (twavetents.cpp:209-267)
    // No Gram is fused into this pass (first-order form, or the stabilisation Gram runs on its compact rule), so it needs
    // no operand buffer and no barrier: warp w takes the chunks w, w + 4, .. and evaluates ALL functions at its 32 points
    // (per-point setup and shared monomials once per point instead of once per warp). It runs first: the Gram
    // accumulators are not live yet, the partial sums of all N functions fit in registers. The state rows are loaded at the
    // top of an iteration and consumed after the coordinate setup (a register prefetch one chunk ahead costs spills).
    constexpr int NC = FO ? d1 : d1 + 1;
    constexpr int QR = CS ? (C::QS > 0 ? C::QS : 1) : Q;             // points per element of the rule the state lives on
    const double* __restrict__ const lamR = CS ? f.lam_stab : a.lam;
    const double* __restrict__ const wR = CS ? f.w_stab : a.w;
    const double* __restrict__ const stR = CS ? f.cstate : a.wavefront;
    // H = 2 (larger bases): a warp evaluates one half of the functions (groups 2h, 2h + 1) on every second chunk, which
    // keeps the partial sums and the live monomials within the 168-register budget of three CTAs per SM
    constexpr int H = (NP > 32) ? 2 : 1;
    constexpr int NS = 4 / H;          // chunk streams
    constexpr int NB = NP / H;         // partial sums per lane
    const int part = w % H, cs = w / H;
    double bp[NB];
#pragma unroll
    for (int i = 0; i < NB; i++) bp[i] = 0.0;
    // STG: the elements of the tent in batches of kRowCap rows (one batch for all but the largest vertex patches): warp 0
    // issues one bulk copy per row, everybody waits on the mbarrier, then the rows are read from shared memory
    for (int b0 = 0; b0 < nel; b0 += kRowCap) {
    const int nb = STG ? min(kRowCap, nel - b0) : nel;
    if constexpr (STG) {
      constexpr unsigned kRowBytes = C::ROWLEN * 8;
      const unsigned char* __restrict__ first = f.te_first;  // streamed input: rows without a compact copy yet
      if (w == 0) {
        int nload = 0;
        for (int e = lane; e < nb; e += 32) nload += !(first && first[te0 + b0 + e]);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) nload += __shfl_xor_sync(0xffffffffu, nload, o);
        if (lane == 0) tma::mbar_expect(&s_bar_rows, nload * kRowBytes);
        __syncwarp();
        for (int e = lane; e < nb; e += 32) {
          if (first && first[te0 + b0 + e]) continue;
          if (a.ticket) tma::fence_proxy_async();  // rows written by other CTAs of this launch (dependency-driven order)
          tma::bulk_load(rowsS + e * C::ROWLEN, f.cstate + (size_t)a.te_el[te0 + b0 + e] * C::ROWLEN, kRowBytes, &s_bar_rows);
        }
      }
      if (first) {
        // compact row = transfer (Qs x Q) times the public row, component by component (what SetInitial's GEMM does for the
        // whole state; here only for the rows this tent is the first to read, straight from the delivered chunk)
        bool any = false;
        for (int e = 0; e < nb; e++) {
          if (!first[te0 + b0 + e]) continue;
          any = true;
          const double* __restrict__ pub = a.wavefront + (size_t)a.te_el[te0 + b0 + e] * (NC * Q);
          for (int idx = tid; idx < C::ROWLEN; idx += kThreads) {
            const int cc = idx / QR, kq = idx % QR;
            const double* __restrict__ m = f.transfer + (size_t)kq * Q;
            const double* __restrict__ g = pub + cc * Q;
            double acc = 0.0;
            for (int q = 0; q < Q; q++) acc = fma(__ldg(m + q), __ldcg(g + q), acc);
            rowsS[e * C::ROWLEN + idx] = acc;
          }
        }
        if (any) __syncthreads();
      }
      tma::mbar_wait(&s_bar_rows, (b0 / kRowCap) & 1);
    }
    const int nptb = nb * QR;
    for (int c0 = cs * kChunk; c0 < nptb; c0 += NS * kChunk) {
      const int idx = c0 + lane;
      const bool valid = idx < nptb;
      const int idc = valid ? idx : nptb - 1;
      const int eb = idc / QR, q = idc % QR;
      const int te = te0 + b0 + eb;
      if (c0 + NS * kChunk < nptb && lane < 6) {
        // geometry records of the (at most two) elements of this warp's NEXT chunk: into L1 now, so that the loads at the top
        // of the next iteration do not wait for L2 (288 B = three 128-byte lines per record; costs no registers)
        const int first_pt = c0 + NS * kChunk + (lane < 3 ? 0 : kChunk - 1);
        const int ten = te0 + b0 + min(first_pt, nptb - 1) / QR;
        asm volatile("prefetch.global.L1 [%0];" ::"l"(f.fe + (size_t)ten * FE + (lane % 3) * 16));
      }
      double gcur[NC];
      if constexpr (STG) {
        const double* __restrict__ r = rowsS + eb * C::ROWLEN + q;
#pragma unroll
        for (int cc = 0; cc < NC; cc++) gcur[cc] = r[cc * QR];
      } else {
        const double* __restrict__ wf = stR + (size_t)a.te_el[te] * (NC * QR) + q;
#pragma unroll
        for (int cc = 0; cc < NC; cc++) gcur[cc] = __ldcg(wf + cc * QR);  // L2: rows written by other SMs
        if (c0 + NS * kChunk < nptb) {  // next chunk of this warp: HBM -> L2 now (costs no registers)
          const int idn = min(idx + NS * kChunk, nptb - 1);
          const double* __restrict__ wn = stR + (size_t)a.te_el[te0 + idn / QR] * (NC * QR) + idn % QR;
#pragma unroll
          for (int cc = 0; cc < NC; cc++) asm volatile("prefetch.global.L2 [%0];" ::"l"(wn + cc * QR));
        }
      }
      const double* __restrict__ rec = f.fe + (size_t)te * FE;
      const double* __restrict__ lm = lamR + q * d1;
      double xi[d1];
#pragma unroll
      for (int u = 0; u < d1; u++) xi[u] = 0.0;
#pragma unroll
      for (int kk = 0; kk < d1; kk++) {
        const double l = lm[kk];
#pragma unroll
        for (int u = 0; u < D; u++) xi[u] = fma(l, rec[fast_fe_xv(D) + kk * D + u], xi[u]);
        xi[D] = fma(l, rec[fast_fe_tb(D) + kk], xi[D]);
      }
      const double area = rec[fast_fe_bot(D)];
      const double* __restrict__ nn = rec + fast_fe_bot(D) + 2;
      const double s = rec[fast_fe_sc(D)], sc = rec[fast_fe_sc(D) + 1], ic2 = rec[fast_fe_sc(D) + 2];
      const double om = valid ? wR[q] * area : 0.0;
      const double gt = gcur[(FO ? 0 : 1) + D];
      double rho[d1];
      double ndg = 0.0;
#pragma unroll
      for (int u = 0; u < D; u++) {
        const double gx = gcur[(FO ? 0 : 1) + u];
        rho[u] = -om * s * (nn[D] * gx - nn[u] * gt);
        ndg = fma(nn[u], gx, ndg);
      }
      rho[D] = -om * sc * (nn[D] * ic2 * gt - ndg);
      const double rho_u = FO ? 0.0 : om * gcur[0];
      auto fn = [&](auto ic, double v, double e0, double e1, double e2, double e3) {
        constexpr int i = decltype(ic)::value;
        const double dv[4] = {e0, e1, e2, e3};
        double r = FO ? 0.0 : v * rho_u;
#pragma unroll
        for (int u = 0; u < d1; u++) r = fma(dv[u], rho[u], r);
        bp[H == 1 ? i : ((i >> 2) * 2 + (i & 1))] += r;   // H == 2: function i = 4 m + 2 part + b -> slot 2 m + b
      };
      eval_part<D, P, FO, H>(part, xi, fn);
    }
    if (STG && b0 + kRowCap < nel) __syncthreads();  // everybody is done with the staged rows before the next batch lands
    }
    // boundary faces, right-hand side only (CalcTentBndEl, 404-433 / 380-403): b += om g (alpha gamma p + nn) (Dirichlet),
    // b += om (n.g) (beta nn + p) (Neumann); the datum is sampled at the points of the order-2P rule (twavetents.cpp:373-374)
    if (nfc > 0) {
      const int nbp = nfc * Q;
      for (int c0 = cs * kChunk; c0 < nbp; c0 += NS * kChunk) {
        const int idx = c0 + lane;
        const bool valid = idx < nbp;
        const int idc = valid ? idx : nbp - 1;
        const int tf = tf0 + idc / Q, q = idc % Q;
        const double* __restrict__ rec = f.ff + (size_t)tf * FF;
        const double* __restrict__ lm = a.lam + q * d1;
        double xi[d1];
#pragma unroll
        for (int u = 0; u < d1; u++) xi[u] = rec[fast_ff_v0(D) + u];
#pragma unroll
        for (int m = 0; m < D; m++) {
          const double l = lm[m];
#pragma unroll
          for (int u = 0; u < d1; u++) xi[u] = fma(l, rec[fast_ff_e(D) + m * d1 + u], xi[u]);
        }
        double nrm[D];
#pragma unroll
        for (int u = 0; u < D; u++) nrm[u] = rec[fast_ff_n(D) + u];
        const double area = rec[fast_ff_misc(D)], s = rec[fast_ff_misc(D) + 1], sc = rec[fast_ff_misc(D) + 2];
        const double ic2 = rec[fast_ff_misc(D) + 3];
        const bool neumann = rec[fast_ff_misc(D) + 4] != 0.0;
        double gdat = 0.0;
        if (a.bnd_data) {
          const long long off = a.tf_data[tf];
          if (neumann) {
#pragma unroll
            for (int r = 0; r < D; r++) gdat = fma(nrm[r], a.bnd_data[off + (long long)r * Q + q], gdat);
          } else {
            gdat = a.bnd_data[off + q];
          }
        }
        const double og = valid ? a.w[q] * area * gdat : 0.0;
        // bc = og (cp p + cn nd): Dirichlet cp = gamma / 2, cn = 1; Neumann cp = 1, cn = 1 / 2
        const double cp = og * sc * (neumann ? 1.0 : 0.5 * ic2), cn = og * s * (neumann ? 0.5 : 1.0);
        auto fnb = [&](auto ic, double, double e0, double e1, double e2, double e3) {
          constexpr int i = decltype(ic)::value;
          const double dv[4] = {e0, e1, e2, e3};
          double nd = 0.0;
#pragma unroll
          for (int u = 0; u < D; u++) nd = fma(nrm[u], dv[u], nd);
          bp[H == 1 ? i : ((i >> 2) * 2 + (i & 1))] += fma(cp, dv[D], cn * nd);
        };
        eval_part<D, P, FO, H>(part, xi, fnb);
      }
    }
#pragma unroll
    for (int sl = 0; sl < NB; sl++) {
      double v = bp[sl];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
      const int i = (H == 1) ? sl : ((sl >> 1) * 4 + 2 * part + (sl & 1));
      if (lane == 0) Z[cs * NP + i] = v;
    }
    __syncthreads();
    for (int i = tid; i < NP; i += kThreads) {
      double v = 0.0;
#pragma unroll
      for (int c = 0; c < NS; c++) v += Z[c * NP + i];
      rhsS[i] = v;
    }
    __syncthreads();
  }
  double bpart[G];
#pragma unroll
  for (int j = 0; j < G; j++) bpart[j] = 0.0;
  if constexpr (!APPLY) {  // ============================ assembly of the tent matrix (skipped when its inverse is stored)
  for (int i = tid; i < NP * ZS; i += kThreads) Z[i] = 0.0;  // rows of the padding functions stay zero
  double acc[NACC][2];
#pragma unroll
  for (int i = 0; i < NACC; i++) acc[i][0] = acc[i][1] = 0.0;
  __syncthreads();

  // ------------------------------------------------------------------ top faces: A += B W B^T  (twavetents.cpp:270-313)
  for (int c0 = 0; c0 < nptt; c0 += kChunk) {
    const int idx = c0 + lane;
    const bool valid = idx < nptt;
    const int idc = valid ? idx : nptt - 1;
    const int te = te0 + idc / QT, q = idc % QT;
    const double* __restrict__ rec = f.fe + (size_t)te * FE;
    const double* __restrict__ lm = f.lam_top + q * d1;
    double xi[d1];
#pragma unroll
    for (int u = 0; u < d1; u++) xi[u] = 0.0;
#pragma unroll
    for (int kk = 0; kk < d1; kk++) {
      const double l = lm[kk];
#pragma unroll
      for (int u = 0; u < D; u++) xi[u] = fma(l, rec[fast_fe_xv(D) + kk * D + u], xi[u]);
      xi[D] = fma(l, rec[fast_fe_tt(D) + kk], xi[D]);
    }
    const double sq = valid ? f.sqrtw_top[q] : 0.0;
    const double k0 = sq * rec[fast_fe_top(D)];
    double kd[D];
#pragma unroll
    for (int u = 0; u < D; u++) kd[u] = sq * rec[fast_fe_top(D) + 1 + u];
    const double kt = sq * rec[fast_fe_top(D) + 1 + D];
    double* const zl = Z + lane;
    eval_group<D, P, FO>(w, xi, [&](auto ic, double, double e0, double e1, double e2, double e3) {
      constexpr int i = decltype(ic)::value;
      const double dv[4] = {e0, e1, e2, e3};
      double* z = zl + i * ZS;
#pragma unroll
      for (int u = 0; u < D; u++) z[u * kChunk] = k0 * dv[u] - kd[u] * dv[D];
      z[D * kChunk] = kt * dv[D];
    });
    __syncthreads();
    if (th == 0) dmma_sym<NT, NACC, 0>(acc, Z, ZS, d1 * (kChunk / 4), kh, g, k);
    else dmma_sym<NT, NACC, 1>(acc, Z, ZS, d1 * (kChunk / 4), kh, g, k);
    __syncthreads();
  }

  // ------------------------------------------------------------------ bottom faces: rhs (+ stabilisation Gram)  (209-267)
  // the state values of the next chunk are loaded one iteration ahead (HBM/L2 latency hidden behind the evaluation)
  constexpr int NC = FO ? d1 : d1 + 1;
  double gnext[NC];
  if (!C::rhsFirst) {
    const int idc = min(lane, npts - 1);
    const double* __restrict__ wf = a.wavefront + (size_t)a.te_el[te0 + idc / Q] * (NC * Q) + idc % Q;
#pragma unroll
    for (int cc = 0; cc < NC; cc++) gnext[cc] = __ldcg(wf + cc * Q);  // L2: rows written by other SMs
  }
  for (int c0 = 0; c0 < npts && !C::rhsFirst; c0 += kChunk) {
    const int idx = c0 + lane;
    const bool valid = idx < npts;
    const int idc = valid ? idx : npts - 1;
    const int te = te0 + idc / Q, q = idc % Q;
    double gcur[NC];
#pragma unroll
    for (int cc = 0; cc < NC; cc++) gcur[cc] = gnext[cc];
    if (c0 + kChunk < npts) {
      const int idn = min(idx + kChunk, npts - 1);
      const double* __restrict__ wf = a.wavefront + (size_t)a.te_el[te0 + idn / Q] * (NC * Q) + idn % Q;
#pragma unroll
      for (int cc = 0; cc < NC; cc++) gnext[cc] = __ldcg(wf + cc * Q);  // L2: rows written by other SMs
    }
    const double* __restrict__ rec = f.fe + (size_t)te * FE;
    const double* __restrict__ lm = a.lam + q * d1;
    double xi[d1];
#pragma unroll
    for (int u = 0; u < d1; u++) xi[u] = 0.0;
#pragma unroll
    for (int kk = 0; kk < d1; kk++) {
      const double l = lm[kk];
#pragma unroll
      for (int u = 0; u < D; u++) xi[u] = fma(l, rec[fast_fe_xv(D) + kk * D + u], xi[u]);
      xi[D] = fma(l, rec[fast_fe_tb(D) + kk], xi[D]);
    }
    const double area = rec[fast_fe_bot(D)];
    const double* __restrict__ nn = rec + fast_fe_bot(D) + 2;
    const double s = rec[fast_fe_sc(D)], sc = rec[fast_fe_sc(D) + 1], ic2 = rec[fast_fe_sc(D) + 2];
    const double om = valid ? a.w[q] * area : 0.0;
    const double gt = gcur[(FO ? 0 : 1) + D];
    double rho[d1];
    double ndg = 0.0;
#pragma unroll
    for (int u = 0; u < D; u++) {
      const double gx = gcur[(FO ? 0 : 1) + u];
      rho[u] = -om * s * (nn[D] * gx - nn[u] * gt);
      ndg = fma(nn[u], gx, ndg);
    }
    rho[D] = -om * sc * (nn[D] * ic2 * gt - ndg);
    const double rho_u = FO ? 0.0 : om * gcur[0];
    const double sqo = valid ? f.sqrtw[q] * rec[fast_fe_bot(D) + 1] : 0.0;
    double* const zl = Z + lane;
    eval_group<D, P, FO>(w, xi, [&](auto ic, double v, double e0, double e1, double e2, double e3) {
      constexpr int i = decltype(ic)::value;
      const double dv[4] = {e0, e1, e2, e3};
      double r = FO ? 0.0 : v * rho_u;
#pragma unroll
      for (int u = 0; u < d1; u++) r = fma(dv[u], rho[u], r);
      bpart[i >> 2] += r;
      if (!FO && !C::splitStab) zl[i * ZS] = sqo * v;
    });
    if (!FO && !C::splitStab) {
      __syncthreads();
      if (th == 0) dmma_sym<NT, NACC, 0>(acc, Z, ZS, kChunk / 4, kh, g, k);
      else dmma_sym<NT, NACC, 1>(acc, Z, ZS, kChunk / 4, kh, g, k);
      __syncthreads();
    }
  }
  if (C::splitStab) {
    // ---------------------------------------------------------------- stabilisation Gram on the compact degree-2P rule
    constexpr int QS = C::QS > 0 ? C::QS : 1;  // (no compact rule: branch not taken)
    const int npss = nel * QS;
    __syncthreads();  // the top-face pass has released Z; the right-hand-side pass above does not touch it
    for (int c0 = 0; c0 < npss; c0 += kChunk) {
      const int idx = c0 + lane;
      const bool valid = idx < npss;
      const int idc = valid ? idx : npss - 1;
      const int te = te0 + idc / QS, q = idc % QS;
      const double* __restrict__ rec = f.fe + (size_t)te * FE;
      const double* __restrict__ lm = f.lam_stab + q * d1;
      double xi[d1];
#pragma unroll
      for (int u = 0; u < d1; u++) xi[u] = 0.0;
#pragma unroll
      for (int kk = 0; kk < d1; kk++) {
        const double l = lm[kk];
#pragma unroll
        for (int u = 0; u < D; u++) xi[u] = fma(l, rec[fast_fe_xv(D) + kk * D + u], xi[u]);
        xi[D] = fma(l, rec[fast_fe_tb(D) + kk], xi[D]);
      }
      const double sqo = valid ? f.sqrtw_stab[q] * rec[fast_fe_bot(D) + 1] : 0.0;
      double* const zl = Z + lane;
      eval_group<D, P, FO>(w, xi, [&](auto ic, double v, double, double, double, double) {
        constexpr int i = decltype(ic)::value;
        zl[i * ZS] = sqo * v;
      });
      __syncthreads();
      if (th == 0) dmma_sym<NT, NACC, 0>(acc, Z, ZS, kChunk / 4, kh, g, k);
      else dmma_sym<NT, NACC, 1>(acc, Z, ZS, kChunk / 4, kh, g, k);
      __syncthreads();
    }
  }

  // ------------------------------------------------------------------ accumulators -> A (fixed order: deterministic)
  if (kh == 0) {
    if (th == 0) flush_sym<NT, NACC, 0>(acc, A, LDA, false, g, k);
    else flush_sym<NT, NACC, 1>(acc, A, LDA, false, g, k);
  }
  __syncthreads();
  if (kh == 1) {
    if (th == 0) flush_sym<NT, NACC, 0>(acc, A, LDA, true, g, k);
    else flush_sym<NT, NACC, 1>(acc, A, LDA, true, g, k);
  }
  __syncthreads();
  for (int idx = tid; idx < NP * NP; idx += kThreads) {
    const int i = idx / NP, j = idx % NP;
    if (j < i && (j >> 3) != (i >> 3)) A[i * LDA + j] = A[j * LDA + i];  // diagonal tiles were computed in full
  }
  __syncthreads();

  // ------------------------------------------------------------------ boundary faces (CalcTentBndEl, 335-434)
  // Dirichlet (404-433): A += sum_q p (om (alpha gamma p - nn))^T, b += om g (alpha gamma p + nn)
  // Neumann   (380-403): A += sum_q (-nn) (om (p - beta nn))^T,    b += om (n.g) (beta nn + p)
  // with p = d_t phi, nn = n . grad_x phi (physical derivatives), alpha = beta = 1/2, gamma = 1/c^2.
  if (nfc > 0) {
#pragma unroll
    for (int i = 0; i < NACC; i++) acc[i][0] = acc[i][1] = 0.0;
    double* const ZL = Z;
    double* const ZR = Z + NP * kBS;
    for (int i = tid; i < 2 * NP * kBS; i += kThreads) Z[i] = 0.0;
    __syncthreads();
    // rhsFirst: the right-hand side was done in the first pass; the matrix integrand is a polynomial of degree 2P-2 on the
    // (affine) time-like face -> compact rule of the top-face Gram (a D-simplex rule as well)
    constexpr int QB = C::rhsFirst ? QT : Q;
    const int nbp = nfc * QB;
    for (int c0 = 0; c0 < nbp; c0 += kChunk) {
      const int idx = c0 + lane;
      const bool valid = idx < nbp;
      const int idc = valid ? idx : nbp - 1;
      const int tf = tf0 + idc / QB, q = idc % QB;
      const double* __restrict__ rec = f.ff + (size_t)tf * FF;
      const double* __restrict__ lm = (C::rhsFirst ? f.lam_top : a.lam) + q * d1;
      double xi[d1];
#pragma unroll
      for (int u = 0; u < d1; u++) xi[u] = rec[fast_ff_v0(D) + u];
#pragma unroll
      for (int m = 0; m < D; m++) {
        const double l = lm[m];
#pragma unroll
        for (int u = 0; u < d1; u++) xi[u] = fma(l, rec[fast_ff_e(D) + m * d1 + u], xi[u]);
      }
      double nrm[D];
#pragma unroll
      for (int u = 0; u < D; u++) nrm[u] = rec[fast_ff_n(D) + u];
      const double area = rec[fast_ff_misc(D)], s = rec[fast_ff_misc(D) + 1], sc = rec[fast_ff_misc(D) + 2];
      const double ic2 = rec[fast_ff_misc(D) + 3];
      const bool neumann = rec[fast_ff_misc(D) + 4] != 0.0;
      const double wq = C::rhsFirst ? f.sqrtw_top[q] * f.sqrtw_top[q] : a.w[q];
      const double om = valid ? wq * area : 0.0;
      double gdat = 0.0;
      if (!C::rhsFirst && a.bnd_data) {
        const long long off = a.tf_data[tf];
        if (neumann) {
#pragma unroll
          for (int r = 0; r < D; r++) gdat = fma(nrm[r], a.bnd_data[off + (long long)r * Q + q], gdat);
        } else {
          gdat = a.bnd_data[off + q];
        }
      }
      const double og = om * gdat;
      double* const zl = ZL + lane;
      double* const zr = ZR + lane;
      eval_group<D, P, FO>(w, xi, [&](auto ic, double, double e0, double e1, double e2, double e3) {
        constexpr int i = decltype(ic)::value;
        const double dv[4] = {e0, e1, e2, e3};
        const double p = sc * dv[D];
        double nd = 0.0;
#pragma unroll
        for (int u = 0; u < D; u++) nd = fma(nrm[u], dv[u], nd);
        nd *= s;
        double lv, rv, bc;
        if (neumann) { lv = -nd; rv = om * (p - 0.5 * nd); bc = og * (0.5 * nd + p); }
        else { lv = p; rv = om * (0.5 * ic2 * p - nd); bc = og * (0.5 * ic2 * p + nd); }
        zl[i * kBS] = lv;
        zr[i * kBS] = rv;
        if (!C::rhsFirst) bpart[i >> 2] += bc;
      });
      __syncthreads();
      switch (w) {
        case 0: dmma_full<NT, NACC, 0>(acc, ZL, ZR, g, k); break;
        case 1: dmma_full<NT, NACC, 1>(acc, ZL, ZR, g, k); break;
        case 2: dmma_full<NT, NACC, 2>(acc, ZL, ZR, g, k); break;
        default: dmma_full<NT, NACC, 3>(acc, ZL, ZR, g, k); break;
      }
      __syncthreads();
    }
    switch (w) {
      case 0: flush_full<NT, NACC, 0>(acc, A, LDA, g, k); break;
      case 1: flush_full<NT, NACC, 1>(acc, A, LDA, g, k); break;
      case 2: flush_full<NT, NACC, 2>(acc, A, LDA, g, k); break;
      default: flush_full<NT, NACC, 3>(acc, A, LDA, g, k); break;
    }
  }
  }  // ======================================================= MODE != 2

  // ------------------------------------------------------------------ right-hand side
  if (C::rhsFirst) {
    for (int i = tid; i < NP; i += kThreads) bv[i] = rhsS[i];
  } else {
#pragma unroll
    for (int j = 0; j < G; j++) {
      double v = bpart[j];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
      if (lane == 0) bv[4 * j + w] = v;
    }
  }
  __syncthreads();
  // ------------------------------------------------------------------ Solve (64-70 forms the inverse with LAPACK)
  // Interior tents (no boundary face) have a symmetric positive definite matrix (Z Z^T + stabilisation): no pivoting.
  if constexpr (MODE == 0) {
    gauss_jordan_solve<N, NP, kThreads>(A, LDA, bv, nfc == 0, a.singular);
    __syncthreads();
  } else {
    const double* inv;
    if constexpr (MODE == 1) {
      gauss_jordan_invert<N, NP, kThreads>(A, LDA, nfc == 0, a.singular, invG);
      __syncthreads();  // the inverse this CTA has just written is visible to all of its threads
      inv = invG;
    } else if constexpr (MODE == 3) {
      inv = invG;
    } else {
      tma::mbar_wait(&s_bar, 0);
      inv = invS;
    }
    // sol = A^-1 b. Thread (h, i) sums the columns j = h, h + 2, .. of row i (identical arithmetic in MODE 1 and 2)
    const int i = tid & 63, h = tid >> 6;
    double part = 0.0;
    if (i < N)
      for (int j = h; j < N; j += 2) part = fma(inv[(size_t)j * NP + i], bv[j], part);
    __syncthreads();
    if (h == 1 && i < N) cm[i] = part;   // cm is free until the evaluation
    __syncthreads();
    if (h == 0 && i < N) bv[i] = part + cm[i];
    __syncthreads();
  }

  // ------------------------------------------------------------------ CalcTentElEval (541-590) in the monomial basis
  // u = sum_i sol_i phi_i = sum_j cm_j xi^alpha_j with cm = M^T sol; derivative coefficients in the degree <= P-1 basis.
  for (int j = tid; j < npoly; j += kThreads) {
    double v = 0.0;
    for (int p = f.cptr[j]; p < f.cptr[j + 1]; p++) v = fma(f.cval[p], bv[f.crow[p]], v);
    cm[j] = v;
  }
  __syncthreads();
  for (int t = tid; t < d1 * npm1; t += kThreads) cmd[t] = cm[f.dmap[t]] * f.dfac[t];
  __syncthreads();
  // value and gradient of the solution at point q (barycentric coordinates lm) of the top face of tent-element te -> row
  // `row` with `QW` points per component
  auto eval_point = [&](int te, const double* __restrict__ lm, double* __restrict__ row, int QW, int q) {
    const double* __restrict__ rec = f.fe + (size_t)te * FE;
    double xi[4] = {0.0, 0.0, 0.0, 0.0};
#pragma unroll
    for (int kk = 0; kk < d1; kk++) {
      const double l = lm[kk];
#pragma unroll
      for (int u = 0; u < D; u++) xi[u] = fma(l, rec[fast_fe_xv(D) + kk * D + u], xi[u]);
      xi[D] = fma(l, rec[fast_fe_tt(D) + kk], xi[D]);
    }
    double pw[4][P + 1];
#pragma unroll
    for (int u = 0; u < 4; u++) {
      pw[u][0] = 1.0;
#pragma unroll
      for (int e = 1; e <= P; e++) pw[u][e] = pw[u][e - 1] * xi[u];
    }
    double val = 0.0, dv[4] = {0.0, 0.0, 0.0, 0.0};
    int j = 0, j1 = 0;
    // monomial order: first coordinate's exponent outermost (src/scalarmappedfe.cpp:526-531); unused coordinates fixed at 0
#pragma unroll
    for (int e0 = 0; e0 <= P; e0++) {
      const double m0 = pw[0][e0];
#pragma unroll
      for (int e1 = 0; e1 <= P - e0; e1++) {
        const double m1 = m0 * pw[1][e1];
#pragma unroll
        for (int e2 = 0; e2 <= ((d1 > 2) ? P - e0 - e1 : 0); e2++) {
          const double m2 = (d1 > 2) ? m1 * pw[2][e2] : m1;
#pragma unroll
          for (int e3 = 0; e3 <= ((d1 > 3) ? P - e0 - e1 - e2 : 0); e3++) {
            const double m = (d1 > 3) ? m2 * pw[3][e3] : m2;
            if (!FO) val = fma(cm[j], m, val);
            if (e0 + e1 + e2 + e3 <= P - 1) {
#pragma unroll
              for (int u = 0; u < d1; u++) dv[u] = fma(cmd[u * npm1 + j1], m, dv[u]);
              j1++;
            }
            j++;
          }
        }
      }
      asm volatile("" ::: "memory");  // keeps the coefficient loads of later e0 slices from being hoisted (register pressure)
    }
    const double s = rec[fast_fe_sc(D)], sc = rec[fast_fe_sc(D) + 1];
    if (!FO) row[q] = val;
#pragma unroll
    for (int u = 0; u < D; u++) row[((FO ? 0 : 1) + u) * QW + q] = s * dv[u];
    row[((FO ? 0 : 1) + D) * QW + q] = sc * dv[D];
  };
  constexpr int NCW = FO ? d1 : d1 + 1;
  if (CS) {
    // compact rows of every element of the tent (what later tents read) ...
    constexpr int QC = C::QS > 0 ? C::QS : 1;
    const int nptc = nel * QC;
    for (int idx = tid; idx < nptc; idx += kThreads) {
      const int te = te0 + idx / QC, q = idx % QC;
      eval_point(te, f.lam_stab + q * d1, f.cstate + (size_t)a.te_el[te] * (NCW * QC), QC, q);
    }
    // ... and the public rows (order-2P points) of the elements this tent is the last one of the slab to contain
    __shared__ int s_fin[kMaxFinal], s_nfin;
    if (tid == 0) s_nfin = 0;
    __syncthreads();
    for (int e = tid; e < nel; e += kThreads)
      if (f.te_final[te0 + e]) s_fin[atomicAdd(&s_nfin, 1) % kMaxFinal] = e;
    __syncthreads();
    const int nfin = s_nfin;
    if (nfin <= kMaxFinal) {
      for (int idx = tid; idx < nfin * Q; idx += kThreads) {
        const int te = te0 + s_fin[idx / Q], q = idx % Q;
        eval_point(te, a.lam + q * d1, a.wavefront + (size_t)a.te_el[te] * (NCW * Q), Q, q);
      }
    } else {  // more final elements than the list holds (never with vertex patches of simplicial meshes): scan the flags
      for (int idx = tid; idx < npts; idx += kThreads) {
        const int te = te0 + idx / Q, q = idx % Q;
        if (f.te_final[te]) eval_point(te, a.lam + q * d1, a.wavefront + (size_t)a.te_el[te] * (NCW * Q), Q, q);
      }
    }
    if (f.chunk_left) {
      // streamed output: one element less to finish in its chunk; the copy engine takes a chunk out as soon as its count
      // is zero (stream wait on the counter), so the rows must be visible system-wide first
      // (one fence for the CTA: the barrier orders every thread's stores before thread 0's fence, which is cumulative --
      // the pattern of a cooperative-groups grid barrier)
      __syncthreads();
      if (w == 0) {
        if (lane == 0) __threadfence_system();
        __syncwarp();
        for (int e = lane; e < nel; e += 32)
          if (f.te_final[te0 + e]) atomicSub(f.chunk_left + a.te_el[te0 + e] / f.chunk_elems, 1);
      }
    }
  } else {
    for (int idx = tid; idx < npts; idx += kThreads) {
      const int te = te0 + idx / Q, q = idx % Q;
      eval_point(te, a.lam + q * d1, a.wavefront + (size_t)a.te_el[te] * (NCW * Q), Q, q);
    }
  }
  if (a.ticket) {
    bool remote = false;
    if (CS && f.p2p) remote = f.p2p->push_off[te0 + nel] > f.p2p->push_off[te0] || f.p2p->rsucc_off[tent + 1] > f.p2p->rsucc_off[tent];
    if (CS && remote) {
      // multi-GPU: the rows this tent has just written also go to the other ranks that hold the element -- stores to
      // their buffers over NVLink -- before its successors, here or there, are released (interior tents, which neither
      // push nor have successors elsewhere, take the cheaper device-scope path below)
      __syncthreads();
      const P2PArgs& pp = *f.p2p;
      constexpr int RL = C::ROWLEN, RP = NCW * Q;
      for (int e = 0; e < nel; e++) {
        const int te = te0 + e;
        const int k0 = pp.push_off[te], k1 = pp.push_off[te + 1];
        if (k0 == k1) continue;
        const size_t el = (size_t)a.te_el[te];
        const bool fin = f.te_final[te] != 0;
        for (int k = k0; k < k1; k++) {
          const int q = pp.push_peer[k];
          const size_t pel = (size_t)pp.push_el[k];
          double* __restrict__ dst = pp.cstate[q] + pel * RL;
          const double* __restrict__ src = f.cstate + el * RL;
          for (int i = tid; i < RL; i += kThreads) dst[i] = __ldcg(src + i);
          if (fin) {
            double* __restrict__ dw = pp.wavefront[q] + pel * RP;
            const double* __restrict__ sw = a.wavefront + el * RP;
            for (int i = tid; i < RP; i += kThreads) dw[i] = __ldcg(sw + i);
          }
        }
      }
      __syncthreads();
      if (w == 0) {  // one system-scope fence for the CTA (cumulative over the barrier), then the releases
        if (lane == 0) __threadfence_system();
        __syncwarp();
        for (int s = a.succ_off[tent] + lane; s < a.succ_off[tent + 1]; s += 32) atomicSub(a.remaining + a.succ[s], 1);
        for (int s = pp.rsucc_off[tent] + lane; s < pp.rsucc_off[tent + 1]; s += 32)
          atomicSub_system(pp.remaining[pp.rsucc_peer[s]] + pp.rsucc_tent[s], 1);
      }
    } else {
      // the state rows of every thread are visible device-wide before the successors are released: barrier, then ONE
      // fence (cumulative), then the releases by the same warp
      __syncthreads();
      if (w == 0) {
        if (lane == 0) __threadfence();
        __syncwarp();
        for (int s = a.succ_off[tent] + lane; s < a.succ_off[tent + 1]; s += 32) atomicSub(a.remaining + a.succ[s], 1);
      }
    }
  }
}

template <int D, int P, int FO, int CS, int MODE>
cudaError_t launch_cfg(const TentKernelArgs& a, const FastArgs& f, cudaStream_t stream) {
  using C = Cfg<D, P, FO>;
  if (a.Q != C::Q || f.Qt != C::QT || a.N != C::N || a.npoly != C::npoly) return cudaErrorInvalidValue;
  if ((C::splitStab || CS) && f.Qs != C::QS) return cudaErrorInvalidValue;
  if (CS && (!f.cstate || !f.te_final || !f.w_stab)) return cudaErrorInvalidValue;
  if (MODE != 0 && !f.ainv) return cudaErrorInvalidValue;
  const size_t smem = sizeof(double) * (MODE == 2 ? C::nsmem2 : (MODE == 3 ? C::nsmem3 : C::nsmem));
  // the attribute is per device: one flag per device ordinal, not per process (a process may hold solvers on several GPUs)
  static bool configured[64] = {};  // (one flag set per template instance)
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev < 0 || dev >= 64 || !configured[dev]) {
    e = cudaFuncSetAttribute(tent_kernel_fast<D, P, FO, CS, MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    if (dev >= 0 && dev < 64) configured[dev] = true;
  }
  tent_kernel_fast<D, P, FO, CS, MODE><<<a.ntents, kThreads, smem, stream>>>(a, f);
  return cudaGetLastError();
}

// compact state only where the compact degree-2P rule exists and is smaller than the order-2P rule
template <int D, int P, int FO>
cudaError_t launch_any(const TentKernelArgs& a, const FastArgs& f, cudaStream_t stream) {
  if constexpr (Cfg<D, P, FO>::QS > 0 && Cfg<D, P, FO>::QS < Cfg<D, P, FO>::Q && Cfg<D, P, FO>::rhsFirst) {
    if (f.cstate) {
      if (f.mode == 1) return launch_cfg<D, P, FO, 1, 1>(a, f, stream);
      if (f.mode == 2) return launch_cfg<D, P, FO, 1, 2>(a, f, stream);
      if (f.mode == 3) return launch_cfg<D, P, FO, 1, 3>(a, f, stream);
      return launch_cfg<D, P, FO, 1, 0>(a, f, stream);
    }
  } else {
    if (f.cstate) return cudaErrorInvalidValue;
  }
  if (f.mode != 0) return cudaErrorInvalidValue;  // the factor cache needs the compact state
  return launch_cfg<D, P, FO, 0, 0>(a, f, stream);
}

template <int D, int P, int FO>
constexpr size_t smem_cfg() { return sizeof(double) * Cfg<D, P, FO>::nsmem; }

}  // namespace

#define NGST_FAST_CONFIGS(X) X(3, 4) X(3, 3) X(3, 2) X(2, 4) X(2, 5) X(2, 3) X(2, 6)

// points of the compact state rule of (D, order, fosystem), 0 where the variant has no compact-state form
int tent_fast_compact_points(int D, int order, int fosystem) {
#define X(d, p)                                                                                              \
  if (D == d && order == p) {                                                                                \
    using C0 = Cfg<d, p, 0>;                                                                                 \
    using C1 = Cfg<d, p, 1>;                                                                                 \
    if (fosystem) return (C1::QS > 0 && C1::QS < C1::Q && C1::rhsFirst) ? C1::QS : 0;                        \
    return (C0::QS > 0 && C0::QS < C0::Q && C0::rhsFirst) ? C0::QS : 0;                                      \
  }
  NGST_FAST_CONFIGS(X)
#undef X
  return 0;
}

// whether the variant stages its compact rows through shared memory with bulk copies (needed by the spatial sweep)
bool tent_fast_stages_rows(int D, int order, int fosystem) {
#define X(d, p) \
  if (D == d && order == p) return fosystem ? (Cfg<d, p, 1>::stageRows && Cfg<d, p, 1>::QS < Cfg<d, p, 1>::Q) : (Cfg<d, p, 0>::stageRows && Cfg<d, p, 0>::QS < Cfg<d, p, 0>::Q);
  NGST_FAST_CONFIGS(X)
#undef X
  return false;
}

// doubles per tent of the stored inverse (NP x N), 0 without a fast variant
size_t tent_fast_inverse_doubles(int D, int order, int fosystem) {
#define X(d, p) \
  if (D == d && order == p) return fosystem ? (size_t)Cfg<d, p, 1>::NP * Cfg<d, p, 1>::N : (size_t)Cfg<d, p, 0>::NP * Cfg<d, p, 0>::N;
  NGST_FAST_CONFIGS(X)
#undef X
  return 0;
}

bool tent_fast_available(int D, int order) {
#define X(d, p) if (D == d && order == p) return true;
  NGST_FAST_CONFIGS(X)
#undef X
  return false;
}

size_t tent_fast_smem_bytes(int D, int order, int fosystem) {
#define X(d, p) if (D == d && order == p) return fosystem ? smem_cfg<d, p, 1>() : smem_cfg<d, p, 0>();
  NGST_FAST_CONFIGS(X)
#undef X
  return 0;
}

cudaError_t launch_tent_fast(int D, int order, int fosystem, const TentKernelArgs& a, const FastArgs& f,
                             cudaStream_t stream) {
  if (a.ntents <= 0) return cudaSuccess;
#define X(d, p) if (D == d && order == p) return fosystem ? launch_any<d, p, 1>(a, f, stream) : launch_any<d, p, 0>(a, f, stream);
  NGST_FAST_CONFIGS(X)
#undef X
  return cudaErrorInvalidValue;
}

}  // namespace ngst
