// leapfrog_kernel.cuh — the trajectory on its own: one launch integrates num_steps leapfrog steps for every chain.
//
// Replaces samplers/HMC.py:85-134 (`leapfrog`, the lax.scan over `lf_step`) and, with FRICTION, samplers/GRAHMC.py:
// 215-261 (`_trajectory_with_schedule` over `_conformal_leapfrog_step`, :152-211).  The sampler kernels integrate the
// same steps inline (hmc_kernel.cuh); this entry exposes them the way the reference does — as a public function of
// (position, momentum, log_prob, grad_log_prob) — so the integrator can be checked by itself: against the oracle, against
// the reference's source run on recorded inputs (tests/golden), and by its reversibility at any size.
//
// Thread per chain, state in registers, operations in the order the reference states them in BOTH builds (the fast
// build only lets the compiler contract a*b+c):
//   [p <- p*s_i]  p <- p + (eps/2) g;  q <- q + eps (p w);  (lp, g) <- value_and_grad(q);  p <- p + (eps/2) g;  [p <- p*s_i]
// with s_i = exp(-gamma(i eps) eps/2) from the same device table as hmc_kernel.cuh.  num_steps == 0 returns the
// inputs (a scan of length 0).  Intermediate log-probs are never read, so only the last step evaluates the value.
#pragma once
#include "common.cuh"
#include "hmc_kernel.cuh"   // friction_gamma_dev
#include "rowstore.cuh"
#include "targets.cuh"

namespace mcmc {

struct LeapfrogArgs {
  McmcParams p;
  TargetCtx t;
  const double* inv_mass;                 // [D] or NULL (ones)
  const double *q, *mom, *lp, *grad;      // inputs  [C][D], [C][D], [C], [C][D]
  double *q_out, *mom_out, *lp_out, *grad_out;
};

template <int TID, int DS, bool FRICTION, int BLOCK>
__global__ void __launch_bounds__(BLOCK) leapfrog_kernel(const LeapfrogArgs a) {
  constexpr int N = DS > 0 ? DS : MCMC_DYN_MAXD;
  const int n = DS > 0 ? DS : a.p.n_dim;
  const int L = a.p.num_steps;
  const int64_t C = a.p.n_chains;
  const int tid = threadIdx.x;
  extern __shared__ __align__(16) double smem[];
  double* sm_w = smem;
  double* sm_s = smem + n;
  double* sm_tab = smem + n + (FRICTION ? L : 0);
  const double eps = a.p.step_size;
  const double half_eps = 0.5 * eps;   // HMC.py:119 `half * step_sz`; GRAHMC.py:193

  for (int d = tid; d < n; d += BLOCK) sm_w[d] = a.inv_mass ? a.inv_mass[d] : 1.0;
  if (FRICTION) {
    const double T = eps * L;          // GRAHMC.py:246
    for (int i = tid; i < L; i += BLOCK)
      sm_s[i] = exp(-friction_gamma_dev(a.p.schedule_id, i * eps, T, a.p.gamma, a.p.steepness) * half_eps);  // :194,251
  }
  target_table_fill<TID>(a.t, n, sm_tab);
  __syncthreads();

  const int64_t c = (int64_t)blockIdx.x * BLOCK + tid;
  if (c >= C) return;
  double q[N], p[N], g[N];
  load_row<DS>(q, a.q + c * n, n);
  load_row<DS>(p, a.mom + c * n, n);
  load_row<DS>(g, a.grad + c * n, n);
  double lp = a.lp ? a.lp[c] : 0.0;
  for (int i = 0; i < L; ++i) {
    const double s = FRICTION ? sm_s[i] : 1.0;
#pragma unroll
    for (int d = 0; d < n; ++d) {
      if (FRICTION) p[d] = p[d] * s;          // GRAHMC.py:197
      p[d] = p[d] + half_eps * g[d];          // HMC.py:119
      q[d] = q[d] + eps * (p[d] * sm_w[d]);   // HMC.py:121
    }
    if (i == L - 1) lp = target_eval<TID, DS, true, true>(a.t, n, q, g, sm_tab);   // HMC.py:123
    else target_eval<TID, DS, false, true>(a.t, n, q, g, sm_tab);
#pragma unroll
    for (int d = 0; d < n; ++d) {
      p[d] = p[d] + half_eps * g[d];          // HMC.py:127
      if (FRICTION) p[d] = p[d] * s;          // GRAHMC.py:209
    }
  }
  if (a.q_out) store_row<DS>(a.q_out + c * n, q, n);
  if (a.mom_out) store_row<DS>(a.mom_out + c * n, p, n);
  if (a.grad_out) store_row<DS>(a.grad_out + c * n, g, n);
  if (a.lp_out) a.lp_out[c] = lp;
}

template <int TID, int DS, bool FRICTION>
int launch_leapfrog_variant(const LeapfrogArgs& a, cudaStream_t s) {
  constexpr int BLOCK = 128;
  const int n = a.p.n_dim;
  const size_t smem = ((size_t)n + (FRICTION ? (size_t)a.p.num_steps : 0) + (size_t)target_table_len(TID, n)) * sizeof(double);
  auto kern = leapfrog_kernel<TID, DS, FRICTION, BLOCK>;
  if (int rc = cuda_fail(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem),
                         "cudaFuncSetAttribute(leapfrog)"))
    return rc;
  const unsigned grid = (unsigned)((a.p.n_chains + BLOCK - 1) / BLOCK);
  kern<<<grid, BLOCK, smem, s>>>(a);
  return cuda_fail(cudaGetLastError(), "leapfrog launch");
}

template <int TID>
int launch_leapfrog_impl(const McmcParams& p, const LeapfrogIO& io, bool friction, cudaStream_t s) {
  if (p.n_chains == 0) return MCMC_OK;
  LeapfrogArgs a;
  a.p = p;
  a.t = make_target_ctx(p, io.cov_inv);
  a.inv_mass = io.inv_mass;
  a.q = io.q; a.mom = io.mom; a.lp = io.lp; a.grad = io.grad;
  a.q_out = io.q_out; a.mom_out = io.mom_out; a.lp_out = io.lp_out; a.grad_out = io.grad_out;
  switch (a.p.n_dim) {
#ifndef MCMC_USER_GENERIC_ONLY
    case 20: return friction ? launch_leapfrog_variant<TID, 20, true>(a, s) : launch_leapfrog_variant<TID, 20, false>(a, s);
    case 10: return friction ? launch_leapfrog_variant<TID, 10, true>(a, s) : launch_leapfrog_variant<TID, 10, false>(a, s);
#endif
    default: return friction ? launch_leapfrog_variant<TID, 0, true>(a, s) : launch_leapfrog_variant<TID, 0, false>(a, s);
  }
}

}  // namespace mcmc
