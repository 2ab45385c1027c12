// dense_dmma_ab.cu — A/B of the dense-covariance Gaussian mat-vec y = P x (benchmarks/targets.py:94, `x @ cov_inv`),
// thread per chain, D = 20, C = 65,536 chains, as it sits inside a leapfrog step (x <- x - eps * P x, R times):
//
//   scalar : the loop of mcmc_b200/csrc/targets.cuh — P in shared memory (broadcast reads), 400 DFMA per mat-vec
//   dmma   : mma.sync.aligned.m8n8k4.f64 (SASS DMMA).  A warp's 32 chains form a 32 x 20 matrix X; Y = X P is 4 x 3 x 5 = 60
//            m8n8k4 tiles (N padded 20 -> 24).  With one chain per thread the fragments do not sit where the tensor
//            core wants them, so X goes through a shared-memory tile (20 STS), A fragments are 20 LDS, B fragments
//            (P) 15 LDS, the 12 accumulator tiles go back through shared memory (24 STS + 20 LDS).
//
// tcgen05.mma has no f64 kind; the legacy DMMA is the only fp64 tensor path on sm_100a.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -o dense_dmma_ab.bin dense_dmma_ab.cu
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>

constexpr int D = 20, NP = 24, WARPS = 2, BLOCK = 32 * WARPS;
constexpr int XS = 21;   // row stride of the X tile (odd: the column reads of the A fragments are conflict-free)
constexpr int YS = 25;   // row stride of the Y tile

__device__ __forceinline__ void dmma884(double& d0, double& d1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(d0), "+d"(d1) : "d"(a), "d"(b));
}

template <bool DMMA>
__global__ void __launch_bounds__(BLOCK) matvec_kernel(const double* __restrict__ P, const double* __restrict__ x0, double* __restrict__ xout,
                                                       int C, int R, double eps) {
  __shared__ double Ps[D * NP];                    // P, columns padded with zeros to 24
  __shared__ double Xs[WARPS][32 * XS];
  __shared__ double Ys[WARPS][32 * YS];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int i = tid; i < D * NP; i += BLOCK) Ps[i] = (i % NP < D) ? P[(i / NP) * D + i % NP] : 0.0;
  __syncthreads();
  const int chain = blockIdx.x * BLOCK + tid;
  if (chain >= C) return;                          // C is a multiple of 32: whole warps leave together
  double x[D], y[D];
#pragma unroll
  for (int d = 0; d < D; ++d) x[d] = x0[(size_t)chain * D + d];
  for (int r = 0; r < R; ++r) {
    asm volatile("" ::: "memory");   // P is re-read every repetition, as inside a leapfrog body (no hoisting of 400 loads)
    if (!DMMA) {
#pragma unroll
      for (int j = 0; j < D; ++j) {
        double acc = 0.0;
#pragma unroll
        for (int i = 0; i < D; ++i) acc = fma(x[i], Ps[i * NP + j], acc);   // x P (P symmetric here)
        y[j] = acc;
      }
    } else {
      double* xs = Xs[warp];
      double* ys = Ys[warp];
#pragma unroll
      for (int d = 0; d < D; ++d) xs[lane * XS + d] = x[d];
      __syncwarp();
      const int fr = lane >> 2, fc = lane & 3;      // fragment row / column of this lane
#pragma unroll
      for (int ni = 0; ni < 3; ++ni) {
        double bfrag[5];
#pragma unroll
        for (int ki = 0; ki < 5; ++ki) bfrag[ki] = Ps[(4 * ki + fc) * NP + 8 * ni + fr];   // B[k][n], col-major fragment
#pragma unroll
        for (int mi = 0; mi < 4; ++mi) {
          double d0 = 0.0, d1 = 0.0;
#pragma unroll
          for (int ki = 0; ki < 5; ++ki) dmma884(d0, d1, xs[(8 * mi + fr) * XS + 4 * ki + fc], bfrag[ki]);
          ys[(8 * mi + fr) * YS + 8 * ni + 2 * fc] = d0;
          ys[(8 * mi + fr) * YS + 8 * ni + 2 * fc + 1] = d1;
        }
      }
      __syncwarp();
#pragma unroll
      for (int d = 0; d < D; ++d) y[d] = ys[lane * YS + d];
      __syncwarp();
    }
#pragma unroll
    for (int d = 0; d < D; ++d) x[d] = fma(-eps, y[d], x[d]);
  }
#pragma unroll
  for (int d = 0; d < D; ++d) xout[(size_t)chain * D + d] = x[d];
}

int main() {
  const int C = 65536, R = 256;
  const double eps = 1e-3;
  std::vector<double> P(D * D), x(C * (size_t)D);
  srand(1);
  for (int i = 0; i < D; ++i)
    for (int j = 0; j <= i; ++j) P[i * D + j] = P[j * D + i] = (i == j ? 2.0 : 0.0) + 0.1 * ((double)rand() / RAND_MAX - 0.5);
  for (auto& v : x) v = (double)rand() / RAND_MAX - 0.5;
  double *dP, *dx, *ya, *yb;
  cudaMalloc(&dP, P.size() * 8); cudaMalloc(&dx, x.size() * 8); cudaMalloc(&ya, x.size() * 8); cudaMalloc(&yb, x.size() * 8);
  cudaMemcpy(dP, P.data(), P.size() * 8, cudaMemcpyHostToDevice);
  cudaMemcpy(dx, x.data(), x.size() * 8, cudaMemcpyHostToDevice);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  float best[2] = {1e9f, 1e9f};
  for (int rep = 0; rep < 4; ++rep) {
    for (int v = 0; v < 2; ++v) {
      cudaEventRecord(e0);
      if (v == 0) matvec_kernel<false><<<C / BLOCK, BLOCK>>>(dP, dx, ya, C, R, eps);
      else matvec_kernel<true><<<C / BLOCK, BLOCK>>>(dP, dx, yb, C, R, eps);
      cudaEventRecord(e1); cudaEventSynchronize(e1);
      float ms; cudaEventElapsedTime(&ms, e0, e1);
      if (rep > 0 && ms < best[v]) best[v] = ms;
    }
  }
  cudaError_t err = cudaGetLastError();
  std::vector<double> ha(x.size()), hb(x.size());
  cudaMemcpy(ha.data(), ya, x.size() * 8, cudaMemcpyDeviceToHost);
  cudaMemcpy(hb.data(), yb, x.size() * 8, cudaMemcpyDeviceToHost);
  double maxdiff = 0, maxabs = 0;
  for (size_t i = 0; i < x.size(); ++i) { maxdiff = fmax(maxdiff, fabs(ha[i] - hb[i])); maxabs = fmax(maxabs, fabs(ha[i])); }
  const double flop = 2.0 * D * D * (double)C * R;
  printf("cuda status: %s\n", cudaGetErrorString(err));
  printf("dense mat-vec y = x P, D = %d, C = %d chains, %d dependent repetitions (x <- x - eps y)\n", D, C, R);
  printf("scalar (broadcast P from smem, 400 DFMA)   : %8.3f ms  %7.2f TFLOP/s useful\n", best[0], flop / best[0] / 1e9);
  printf("dmma   (m8n8k4 f64, 60 DMMA, smem staging) : %8.3f ms  %7.2f TFLOP/s useful\n", best[1], flop / best[1] / 1e9);
  printf("max |scalar - dmma| = %.3e (max |x| = %.3e)\n", maxdiff, maxabs);
  return 0;
}
