// FP64 roofline denominators measured on the device the solver runs on: the DFMA pipe and the
// mma.sync m8n8k4 f64 ("DMMA") path. MEASURED_PEAKS.json holds only HBM and bf16 figures, and this
// path is FP64-bound, so bench.py measures the FP64 peak live with these kernels.
#include <cuda_runtime.h>

namespace ngst {

namespace {

constexpr int kChains = 8;

__global__ void __launch_bounds__(256) dfma_peak_kernel(double* out, int iters, double a, double b) {
  double x[kChains];
#pragma unroll
  for (int k = 0; k < kChains; k++) x[k] = (double)(threadIdx.x + k) * 1e-3;
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int r = 0; r < 8; r++)
#pragma unroll
      for (int k = 0; k < kChains; k++) x[k] = fma(x[k], a, b);
  }
  double s = 0;
#pragma unroll
  for (int k = 0; k < kChains; k++) s += x[k];
  if (s == 123.456) out[blockIdx.x * blockDim.x + threadIdx.x] = s;  // keeps the chains alive
}

__device__ __forceinline__ void dmma_m8n8k4(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
               : "+d"(c0), "+d"(c1)
               : "d"(a), "d"(b));
}

__global__ void __launch_bounds__(256) dmma_peak_kernel(double* out, int iters, double a, double b) {
  double c[kChains][2];
#pragma unroll
  for (int k = 0; k < kChains; k++) { c[k][0] = threadIdx.x * 1e-3; c[k][1] = k * 1e-3; }
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int r = 0; r < 4; r++)
#pragma unroll
      for (int k = 0; k < kChains; k++) dmma_m8n8k4(c[k][0], c[k][1], a, b);
  }
  double s = 0;
#pragma unroll
  for (int k = 0; k < kChains; k++) s += c[k][0] + c[k][1];
  if (s == 123.456) out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

// DMMA and DFMA issued from the same warps, 8 DMMAs (4096 flops/warp) to 16 DFMAs (1024 flops/warp): tells whether
// the two share one FP64 datapath (total stays at the DMMA peak) or can overlap (total exceeds it).
__global__ void __launch_bounds__(256) mixed_peak_kernel(double* out, int iters, double a, double b) {
  double c[kChains][2];
  double x[kChains];
#pragma unroll
  for (int k = 0; k < kChains; k++) { c[k][0] = threadIdx.x * 1e-3; c[k][1] = k * 1e-3; x[k] = (threadIdx.x + k) * 1e-3; }
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int r = 0; r < 4; r++) {
#pragma unroll
      for (int k = 0; k < kChains; k++) {
        dmma_m8n8k4(c[k][0], c[k][1], a, b);
        x[k] = fma(x[k], a, b);
        x[(k + 4) % kChains] = fma(x[(k + 4) % kChains], a, b);
      }
    }
  }
  double s = 0;
#pragma unroll
  for (int k = 0; k < kChains; k++) s += c[k][0] + c[k][1] + x[k];
  if (s == 123.456) out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

}  // namespace

// which: 0 = DFMA, 1 = DMMA m8n8k4, 2 = both interleaved (4:1 flops). Returns TFLOP/s (2 flops per FMA), best of `reps` timed launches.
cudaError_t measure_fp64_peak(int which, int reps, double* tflops) {
  cudaError_t e;
  int dev = 0, nsm = 0;
  if ((e = cudaGetDevice(&dev)) != cudaSuccess) return e;
  if ((e = cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, dev)) != cudaSuccess) return e;
  const int blocks = nsm * 8, threads = 256, iters = 4096;
  double* d_out = nullptr;
  if ((e = cudaMalloc(&d_out, sizeof(double) * blocks * threads)) != cudaSuccess) return e;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  double best = 0;
  for (int r = 0; r < reps + 1; r++) {
    cudaEventRecord(e0);
    if (which == 0) dfma_peak_kernel<<<blocks, threads>>>(d_out, iters, 0.999999, 1e-9);
    else if (which == 1) dmma_peak_kernel<<<blocks, threads>>>(d_out, iters, 0.999999, 1e-9);
    else mixed_peak_kernel<<<blocks, threads>>>(d_out, iters, 0.999999, 1e-9);
    cudaEventRecord(e1);
    if ((e = cudaEventSynchronize(e1)) != cudaSuccess) break;
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    double fma_count;
    if (which == 0) fma_count = (double)blocks * threads * iters * 8.0 * kChains;
    else if (which == 1) fma_count = (double)blocks * (threads / 32) * iters * 4.0 * kChains * 256.0;  // 8x8x4 FMAs per warp MMA
    else fma_count = (double)blocks * (threads / 32) * iters * 4.0 * kChains * (256.0 + 2 * 32.0);
    const double tf = 2.0 * fma_count / (ms * 1e-3) * 1e-12;
    if (r > 0 && tf > best) best = tf;  // first launch is the warm-up
  }
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  cudaFree(d_out);
  if (e == cudaSuccess) e = cudaGetLastError();
  *tflops = best;
  return e;
}

}  // namespace ngst
