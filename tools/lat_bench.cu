// Dependent-chain latencies of the operations the batched step is made of (one warp, one SM): cycles per operation.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o trajoptkp_b200/_build/lat_bench tools/lat_bench.cu
#include <cstdio>
#include <cuda_runtime.h>

#define N 512
template <int OP>
__global__ void k(double* out, const int* idx, long long* cyc, double a0, double b0) {
  __shared__ double sh[1024];
  __shared__ int shi[1024];
  for (int i = threadIdx.x; i < 1024; i += blockDim.x) { sh[i] = 1.0 + 1e-9 * i; shi[i] = idx[i]; }
  __syncthreads();
  double a = a0 + threadIdx.x * 1e-12, b = b0;
  int j = threadIdx.x & 31;
  long long t0 = clock64();
#pragma unroll 16
  for (int i = 0; i < N; i++) {
    if (OP == 0) a = fma(a, b, b);
    if (OP == 1) a = a * b;
    if (OP == 2) a = a + b;
    if (OP == 3) a = 1.0 / a + b;
    if (OP == 4) a = sqrt(a) + b;
    if (OP == 5) a = sin(a) + b;
    if (OP == 6) a = __shfl_xor_sync(0xffffffffu, a, 1) + b;
    if (OP == 7) { j = shi[j]; }
    if (OP == 8) { j = idx[j]; }
    if (OP == 9) { a = sh[(int)a & 1023]; }
    if (OP == 10) { a = a * b; __syncwarp(); }
    if (OP == 11) { a = a * b; __syncthreads(); }
    if (OP == 12) { float f = (float)a; f = fmaf(f, 1.0001f, 0.5f); a = f; }
    if (OP == 13) { a = __shfl_sync(0xffffffffu, a, 3) * b; }
    if (OP == 14) { sh[threadIdx.x] = a; __syncwarp(); a = sh[threadIdx.x ^ 1] + b; __syncwarp(); }
    if (OP == 15) { double s, c; sincos(a, &s, &c); a = s + c * b; }
    if (OP == 16) { a = __longlong_as_double(__reduce_max_sync(0xffffffffu, (int)__double2hiint(a))) + a; }
    if (OP == 17) { a = rsqrt(a) + b; }
  }
  long long t1 = clock64();
  out[threadIdx.x] = a + j;
  if (threadIdx.x == 0) *cyc = t1 - t0;
}

template <int OP>
void run(const char* name, double* out, int* idx, long long* cyc, int threads = 32) {
  k<OP><<<1, threads>>>(out, idx, cyc, 1.0000001, 0.9999999);
  k<OP><<<1, threads>>>(out, idx, cyc, 1.0000001, 0.9999999);
  cudaDeviceSynchronize();
  long long h;
  cudaMemcpy(&h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
  printf("%-44s %7.1f cycles\n", name, (double)h / N);
}

int main() {
  double* out; int* idx; long long* cyc;
  cudaMalloc(&out, 1024 * 8); cudaMalloc(&idx, 1024 * 4); cudaMalloc(&cyc, 8);
  int h[1024];
  for (int i = 0; i < 1024; i++) h[i] = (i * 7 + 3) & 1023;
  cudaMemcpy(idx, h, sizeof(h), cudaMemcpyHostToDevice);
  run<0>("dfma chain", out, idx, cyc);
  run<1>("dmul chain", out, idx, cyc);
  run<2>("dadd chain", out, idx, cyc);
  run<3>("1/x + b", out, idx, cyc);
  run<4>("sqrt(x) + b", out, idx, cyc);
  run<17>("rsqrt(x) + b", out, idx, cyc);
  run<5>("sin(x) + b", out, idx, cyc);
  run<15>("sincos", out, idx, cyc);
  run<6>("shfl_xor double + dadd", out, idx, cyc);
  run<13>("shfl double + dmul", out, idx, cyc);
  run<7>("shared int pointer chase", out, idx, cyc);
  run<8>("global (L1) int pointer chase", out, idx, cyc);
  run<9>("cvt + shared double load", out, idx, cyc);
  run<10>("dmul + syncwarp", out, idx, cyc);
  run<11>("dmul + syncthreads (1 warp)", out, idx, cyc);
  run<11>("dmul + syncthreads (4 warps)", out, idx, cyc, 128);
  run<11>("dmul + syncthreads (12 warps)", out, idx, cyc, 384);
  run<12>("double->float fma float->double", out, idx, cyc);
  run<14>("shared store, syncwarp, load neighbour, sync", out, idx, cyc);
  run<16>("redux.max.s32 + dadd", out, idx, cyc);
  return 0;
}
