// Measured denominators for the compute-bound kernels' rooflines: the FP64 FMA rate
// of the device (SURVEY 8d asks for the achieved FP64 FLOP/s of the limit-curve and
// phase-plane kernels against a MEASURED FP64 FMA peak).  Diagnostics only.
#include <cstdio>

#include <cuda_runtime.h>

#include "../../include/tpb.h"

namespace {

constexpr int kChains = 8;      // independent accumulators per thread
constexpr int kIters = 4096;    // FMAs per accumulator

__global__ void __launch_bounds__(256) k_fp64_fma_peak(double* out, double a, double b) {
  double acc[kChains];
#pragma unroll
  for (int c = 0; c < kChains; ++c) acc[c] = a * (threadIdx.x + c);
#pragma unroll 4
  for (int i = 0; i < kIters; ++i) {
#pragma unroll
    for (int c = 0; c < kChains; ++c) acc[c] = fma(acc[c], a, b);
  }
  double s = 0.0;
#pragma unroll
  for (int c = 0; c < kChains; ++c) s += acc[c];
  if (s == 123.456) out[0] = s;  // keeps the chains alive; never true for the inputs used
}

}  // namespace

extern "C" int32_t tpb_measure_fp64_fma_peak(int32_t device, double* tflops) {
  if (tflops == nullptr) return TPB_INVALID_ARGUMENT;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
    cudaGetLastError();
    return TPB_FAILED_PRECONDITION;
  }
  if (device < 0 || device >= ndev) return TPB_INVALID_ARGUMENT;
  if (cudaSetDevice(device) != cudaSuccess) return TPB_INTERNAL;
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return TPB_INTERNAL;
  double* out = nullptr;
  if (cudaMalloc(&out, sizeof(double)) != cudaSuccess) return TPB_INTERNAL;
  const int blocks = prop.multiProcessorCount * 8, threads = 256;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  k_fp64_fma_peak<<<blocks, threads>>>(out, 0.999999, 1e-9);  // warm-up
  float best = 1e30f;
  for (int rep = 0; rep < 5; ++rep) {
    cudaEventRecord(e0);
    k_fp64_fma_peak<<<blocks, threads>>>(out, 0.999999, 1e-9);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms = 0.f;
    cudaEventElapsedTime(&ms, e0, e1);
    best = ms < best ? ms : best;
  }
  const cudaError_t err = cudaGetLastError();
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  cudaFree(out);
  if (err != cudaSuccess) {
    fprintf(stderr, "tpb_measure_fp64_fma_peak: %s\n", cudaGetErrorString(err));
    return TPB_INTERNAL;
  }
  const double flops = 2.0 * kChains * kIters * static_cast<double>(blocks) * threads;
  *tflops = flops / (static_cast<double>(best) * 1e-3) / 1e12;
  return TPB_OK;
}
