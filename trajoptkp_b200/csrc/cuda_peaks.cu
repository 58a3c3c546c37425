// Live roofline denominators: measured FP64 FMA peak (dependent-chain-free DFMA loop on every SM) and a
// STREAM-style device copy.  MEASURED_PEAKS.json carries no FP64 figure, so bench.py measures it here
// (SURVEY.md 8(d)).
#include <cuda_runtime.h>

#include "traj.h"

namespace tokp {

__global__ void k_dfma_peak(double* out, int iters) {
  double a0 = threadIdx.x * 1e-3, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, a4 = a0 + 4, a5 = a0 + 5, a6 = a0 + 6, a7 = a0 + 7;
  double x = 1.0000001, y = 1e-9;
  for (int i = 0; i < iters; i++) {
    a0 = fma(a0, x, y); a1 = fma(a1, x, y); a2 = fma(a2, x, y); a3 = fma(a3, x, y);
    a4 = fma(a4, x, y); a5 = fma(a5, x, y); a6 = fma(a6, x, y); a7 = fma(a7, x, y);
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7;
}

__global__ void k_copy(const double4* __restrict__ in, double4* __restrict__ out, size_t n) {
  size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  size_t stride = (size_t)gridDim.x * blockDim.x;
  for (; i < n; i += stride) out[i] = in[i];
}

}  // namespace tokp

using namespace tokp;

// out[0] = FP64 FMA TFLOP/s (2 flops per FMA), out[1] = copy GB/s (read + write bytes)
extern "C" int tokp_measure_peaks(double out[2]) {
  if (!out) return TOKP_ERR_ARG;
  int dev = 0, sms = 0;
  TOKP_CUDA(cudaGetDevice(&dev));
  TOKP_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  int blocks = sms * 8, threads = 256, iters = 20000;
  double* d_out;
  TOKP_CUDA(cudaMalloc(&d_out, (size_t)blocks * threads * 8));
  float best = 1e30f;
  for (int rep = 0; rep < 4; rep++) {
    cudaEventRecord(e0);
    k_dfma_peak<<<blocks, threads>>>(d_out, iters);
    cudaEventRecord(e1);
    TOKP_CUDA(cudaEventSynchronize(e1));
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    if (rep > 0 && ms < best) best = ms;
  }
  g_launches += 4;
  out[0] = 2.0 * 8.0 * iters * (double)blocks * threads / (best * 1e-3) / 1e12;
  cudaFree(d_out);
  size_t n4 = (size_t)1 << 25;  // 32 Mi double4 = 1 GiB per buffer
  double4 *a, *b;
  TOKP_CUDA(cudaMalloc(&a, n4 * 32));
  TOKP_CUDA(cudaMalloc(&b, n4 * 32));
  cudaMemset(a, 0, n4 * 32);
  best = 1e30f;
  for (int rep = 0; rep < 4; rep++) {
    cudaEventRecord(e0);
    k_copy<<<sms * 16, 512>>>(a, b, n4);
    cudaEventRecord(e1);
    TOKP_CUDA(cudaEventSynchronize(e1));
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    if (rep > 0 && ms < best) best = ms;
  }
  g_launches += 4;
  out[1] = 2.0 * n4 * 32 / (best * 1e-3) / 1e9;
  cudaFree(a);
  cudaFree(b);
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  return TOKP_OK;
}
