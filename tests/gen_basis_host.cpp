// Host-side harness for the generated basis evaluators (ngstrefftz_b200/csrc/cuda/basis_eval_gen.cuh): the same
// straight-line code the fast CUDA kernel inlines, compiled with g++ so that it can be checked without a GPU.
#define __device__
#define __forceinline__ inline
#include "../ngstrefftz_b200/csrc/cuda/basis_eval_gen.cuh"

using namespace ngst::gen;

template <int D, int P, int FO>
static void run(const double* x, double* out) {
  auto f = [&](auto ic, double v, double d0, double d1, double d2, double d3) {
    constexpr int i = decltype(ic)::value;
    out[5 * i + 0] = v; out[5 * i + 1] = d0; out[5 * i + 2] = d1; out[5 * i + 3] = d2; out[5 * i + 4] = d3;
  };
  BasisGroup<D, P, FO, 0>::eval(x, f);
  BasisGroup<D, P, FO, 1>::eval(x, f);
  BasisGroup<D, P, FO, 2>::eval(x, f);
  BasisGroup<D, P, FO, 3>::eval(x, f);
}

#define CFGS(X) X(3, 4) X(3, 3) X(3, 2) X(2, 4) X(2, 5) X(2, 3) X(2, 6)

// out: N x 5 (value, d/dxi_0..3); returns N or -1 when the configuration was not generated
extern "C" int gen_basis_eval(int D, int P, int FO, const double* x, double* out) {
#define X(d, p)                                                            \
  if (D == d && P == p) {                                                  \
    if (FO) { run<d, p, 1>(x, out); return BasisInfo<d, p, 1>::N; }        \
    run<d, p, 0>(x, out);                                                  \
    return BasisInfo<d, p, 0>::N;                                          \
  }
  CFGS(X)
#undef X
  return -1;
}
