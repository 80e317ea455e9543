// Output mirror: state rows are written to mapped, page-locked HOST memory as soon as they are final.
//
// GetWavefront (src/twavetents.hpp:193) hands the caller a copy of the state after Propagate; over PCIe that copy
// (659 MB at config 4) is 12 % of an end-to-end step and starts only when the last tent is done. A row is final once
// the last tent that contains its element has run, which for most rows is several levels before the end of the slab,
// so the rows that became final in level l are streamed out by this kernel on a second stream while the tents of the
// levels after l run. The kernel stores straight into the caller's buffer through its device alias (zero copy):
// no staging, rows land at their element's position. A warp writes 256 contiguous bytes per store instruction.
#include <cuda_runtime.h>

namespace ngst {

namespace {
constexpr int kMirrorThreads = 256;
constexpr int kMirrorUnroll = 8;

// rows[r] = element whose row (rowlen doubles) is copied. The (row, entry) pairs are walked as one flat index space, so
// every thread always has work and consecutive threads touch consecutive doubles except at row ends; each CTA keeps
// kMirrorThreads * kMirrorUnroll * 8 B = 16 KB in flight per iteration.
__global__ void __launch_bounds__(kMirrorThreads) mirror_rows_kernel(const int* __restrict__ rows, int nrows, int rowlen,
                                                                     const double* __restrict__ wf,
                                                                     double* __restrict__ host) {
  const long long total = (long long)nrows * rowlen;
  constexpr int kTile = kMirrorThreads * kMirrorUnroll;
  for (long long base = (long long)blockIdx.x * kTile; base < total; base += (long long)gridDim.x * kTile) {
    double v[kMirrorUnroll];
    size_t off[kMirrorUnroll];
#pragma unroll
    for (int u = 0; u < kMirrorUnroll; u++) {
      const long long idx = base + u * kMirrorThreads + threadIdx.x;
      if (idx < total) {
        const int r = (int)(idx / rowlen);
        off[u] = (size_t)rows[r] * rowlen + (size_t)(idx - (long long)r * rowlen);
        v[u] = __ldcg(wf + off[u]);  // written by other SMs in an earlier kernel: read through L2
      }
    }
#pragma unroll
    for (int u = 0; u < kMirrorUnroll; u++) {
      const long long idx = base + u * kMirrorThreads + threadIdx.x;
      if (idx < total) __stcs(host + off[u], v[u]);  // streaming store: nothing on the device reads it again
    }
  }
}

}  // namespace

// ctas: how many CTAs stream rows out concurrently (each takes the place of one tent CTA while it runs)
cudaError_t launch_mirror_rows(const int* rows, int nrows, int rowlen, const double* wf, double* host_alias, int ctas,
                               cudaStream_t stream) {
  if (nrows <= 0) return cudaSuccess;
  const long long tiles = ((long long)nrows * rowlen + kMirrorThreads * kMirrorUnroll - 1) / (kMirrorThreads * kMirrorUnroll);
  const int grid = (int)(tiles < ctas ? tiles : ctas);
  mirror_rows_kernel<<<grid, kMirrorThreads, 0, stream>>>(rows, nrows, rowlen, wf, host_alias);
  return cudaGetLastError();
}

}  // namespace ngst
