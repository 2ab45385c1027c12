// nuts_inst.cu — instantiates the NUTS kernels and the batched target evaluator of
// ONE target (-DMCMC_TID=k).
#include <cstdlib>

#include "nuts_kernel.cuh"
#if defined(MCMC_WITH_NUTS_QUAD) && !MCMC_IS_FAITHFUL
// quad-of-lanes-per-chain experiment (measured slower: profiles/nuts_quad_ab_r02.txt); compiled only on request:
// python -m mcmc_b200.csrc.build --suffix _quad -DMCMC_WITH_NUTS_QUAD, then taken with MCMC_NUTS_QUAD=1
#include "nuts_quad_kernel.cuh"
#endif

#ifndef MCMC_TID
#error "compile with -DMCMC_TID=<target id>"
#endif

namespace mcmc {

// vmap(value_and_grad(log_prob_fn)) over n points: the *_init functions (HMC.py:73) and
// direct calls of a target's log_prob_fn from host code.
template <int TID, int DS>
__global__ void __launch_bounds__(128) target_eval_kernel(TargetCtx t, int n_rt, const double* __restrict__ q,
                                                          int64_t npts, double* __restrict__ lp,
                                                          double* __restrict__ grad) {
  constexpr int N = DS > 0 ? DS : MCMC_DYN_MAXD;
  const int n = DS > 0 ? DS : n_rt;
  extern __shared__ __align__(16) double smem[];
  target_table_fill<TID>(t, n, smem);
  __syncthreads();
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= npts) return;
  double x[N], g[N];
#pragma unroll
  for (int d = 0; d < n; ++d) x[d] = q[i * n + d];
  const double v = target_eval<TID, DS, true, true>(t, n, x, g, smem);
  if (lp) lp[i] = v;
  if (grad) {
#pragma unroll
    for (int d = 0; d < n; ++d) grad[i * n + d] = g[d];
  }
}

template <>
int launch_target_eval<MCMC_TID>(const McmcParams& p, const double* cov_inv, const double* q, int64_t npts,
                                 double* lp, double* grad, cudaStream_t s) {
  if (npts == 0) return MCMC_OK;
  const TargetCtx t = make_target_ctx(p, cov_inv);
  const int n = p.n_dim;
  const size_t smem = (size_t)target_table_len(MCMC_TID, n) * sizeof(double);
  const unsigned grid = (unsigned)((npts + 127) / 128);
#define MCMC_TE(DS)                                                                                    \
  do {                                                                                                 \
    auto kern = target_eval_kernel<MCMC_TID, DS>;                                                      \
    if (int rc = cuda_fail(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,     \
                                                (int)smem), "cudaFuncSetAttribute(target_eval)"))      \
      return rc;                                                                                       \
    kern<<<grid, 128, smem, s>>>(t, n, q, npts, lp, grad);                                             \
  } while (0)
#ifndef MCMC_USER_GENERIC_ONLY
  if (n == 20) MCMC_TE(20);
  else if (n == 10) MCMC_TE(10);
  else
#endif
    MCMC_TE(0);
#undef MCMC_TE
  return cuda_fail(cudaGetLastError(), "target_eval launch");
}

template <>
int launch_nuts<MCMC_TID>(const Launch& l) {
  switch (l.p.n_dim) {
#ifndef MCMC_USER_GENERIC_ONLY
    case 20:
#if defined(MCMC_WITH_NUTS_QUAD) && !MCMC_IS_FAITHFUL
      if (getenv("MCMC_NUTS_QUAD")) return launch_nuts_quad<MCMC_TID>(l);
#endif
      return launch_nuts_variant<MCMC_TID, 20, 32>(l);
    case 10: return launch_nuts_variant<MCMC_TID, 10, 32>(l);
#endif
    default: return launch_nuts_variant<MCMC_TID, 0, 32>(l);
  }
}

}  // namespace mcmc
