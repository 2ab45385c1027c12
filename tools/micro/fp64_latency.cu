// Dependent-issue latency of fp64 (and a few companion) instructions on one warp: cycles per instruction of a serial
// chain, and of 2 / 4 / 8 independent chains (issue rate).  nvcc -arch=sm_100a -O3 -o fp64_latency fp64_latency.cu
#include <cstdio>
#include <cuda_runtime.h>
#define N 2048
template <int CHAINS, int OP>
__global__ void probe(double* out, long long* cyc, double a, double b) {
  double x[CHAINS];
  for (int c = 0; c < CHAINS; ++c) x[c] = a + c + threadIdx.x * 1e-9;
  long long t0 = clock64();
#pragma unroll 16
  for (int i = 0; i < N; ++i) {
#pragma unroll
    for (int c = 0; c < CHAINS; ++c) {
      if (OP == 0) x[c] = fma(x[c], b, a);
      else if (OP == 1) x[c] = x[c] + b;
      else if (OP == 2) x[c] = x[c] * b;
      else if (OP == 3) x[c] = __shfl_xor_sync(0xffffffffu, x[c], 1) + b;
      else if (OP == 4) x[c] = rint(x[c] * b);
      else if (OP == 5) x[c] = (double)__double2int_rn(x[c]) * b;
    }
  }
  long long t1 = clock64();
  double s = 0;
  for (int c = 0; c < CHAINS; ++c) s += x[c];
  out[threadIdx.x] = s;
  if (threadIdx.x == 0) cyc[0] = t1 - t0;
}
template <int CHAINS, int OP>
void run(const char* name, int ops_per_iter) {
  double* out; long long* cyc;
  cudaMalloc(&out, 32 * 8); cudaMalloc(&cyc, 8);
  probe<CHAINS, OP><<<1, 32>>>(out, cyc, 1.000001, 0.999999);
  probe<CHAINS, OP><<<1, 32>>>(out, cyc, 1.000001, 0.999999);
  long long h; cudaMemcpy(&h, cyc, 8, cudaMemcpyDeviceToHost);
  printf("%-28s chains=%d  %.2f cycles per loop trip (%d dependent op(s) per chain), %.2f cycles per instruction issued\n", name, CHAINS,
         (double)h / N, ops_per_iter, (double)h / N / (CHAINS * ops_per_iter));
  cudaFree(out); cudaFree(cyc);
}
int main() {
  run<1, 0>("DFMA", 1); run<2, 0>("DFMA", 1); run<4, 0>("DFMA", 1); run<8, 0>("DFMA", 1); run<16, 0>("DFMA", 1);
  run<1, 1>("DADD", 1); run<4, 1>("DADD", 1);
  run<1, 2>("DMUL", 1); run<4, 2>("DMUL", 1);
  run<1, 3>("SHFL + DADD", 2); run<4, 3>("SHFL + DADD", 2);
  run<1, 4>("DMUL + rint", 2);
  run<1, 5>("F2I + I2F + DMUL", 3);
  return 0;
}
