// rwmh_kernel.cuh — rwMH_run as one kernel launch (HBM-write-bound at large C).
//
// Replaces samplers/RWMH.py:148-178 (rwMH_run) and :91-119 (rwMH_step): Gaussian
// proposal, value-only log-density, Metropolis test, stacked (position, log_prob).
// One thread owns one chain; the current position stays in registers, the sample
// rows leave through the shared-memory tile writer (rowstore.cuh).
#pragma once
#include "common.cuh"
#include "philox.cuh"
#include "rowstore.cuh"
#include "targets.cuh"

namespace mcmc {

struct RwmhKArgs {
  McmcParams p;
  McmcBuffers b;
  TargetCtx t;
};

#ifndef MCMC_RWMH_MINBLOCKS
#define MCMC_RWMH_MINBLOCKS 1
#endif
constexpr int kRwmhMinBlocks = MCMC_RWMH_MINBLOCKS;

template <int TID, int DS, int BLOCK>
__global__ void __launch_bounds__(BLOCK, DS > 0 ? kRwmhMinBlocks : 1) rwmh_kernel(const RwmhKArgs a) {
  constexpr int N = DS > 0 ? DS : MCMC_DYN_MAXD;
  const int n = DS > 0 ? DS : a.p.n_dim;
  const int64_t C = a.p.n_chains, S = a.p.num_samples, B = a.p.burn_in;
  const int tid = threadIdx.x;
  const int64_t chain0 = (int64_t)blockIdx.x * BLOCK;
  const bool active = chain0 + tid < C;
  const int64_t chain = active ? chain0 + tid : C - 1;
  const int nvalid = (int)((C - chain0) < BLOCK ? (C - chain0) : BLOCK);
  const bool store = !(a.p.flags & MCMC_FLAG_NO_SAMPLES) && a.b.samples != nullptr;
  const bool injected = a.p.rng_mode == MCMC_RNG_INJECTED;
  const double scale = a.p.step_size;

  extern __shared__ __align__(16) double smem[];
  const int tabn = target_table_len(TID, n);
  double* sm_logtab = smem;
  double* sm_tab = smem + kRngTableDoubles;
  double* sm_stage = sm_tab + ((tabn + 1) / 2) * 2;
  target_table_fill<TID>(a.t, n, sm_tab);
  rng_table_fill(sm_logtab);
  __syncthreads();

  TileWriter<DS, BLOCK> tw;
  if (store) tw.init(sm_stage, n, nvalid, a.b.samples + chain0 * n);

  RngCtx rng;
  {
    const uint64_t gid = (uint64_t)(a.p.chain_offset + chain);
    rng.k0 = a.p.seed[0]; rng.k1 = a.p.seed[1];
    rng.gid_lo = (uint32_t)gid; rng.gid_hi = (uint32_t)(gid >> 32);
    rng.logtab = sm_logtab;
  }

  double q[N], qn[N];
  {
    const double* src = a.b.init_q + chain * n;
#pragma unroll
    for (int d = 0; d < n; ++d) q[d] = src[d];
  }
  double lp = target_eval<TID, DS, true, false>(a.t, n, q, nullptr, sm_tab);  // RWMH.py:67
  int acc = 0;

  for (int64_t it = 0; it < B + S; ++it) {
    if (it == B) acc = 0;
    const int64_t si = it - B;
    // proposal = position + scale * noise (RWMH.py:98-99)
    if (injected) {
      const double* zsrc = a.b.z_inject + ((size_t)it * C + chain) * n;
#pragma unroll
      for (int d = 0; d < n; ++d) qn[d] = q[d] + scale * zsrc[d];
    } else {
#pragma unroll
      for (int j = 0; 2 * j < n; ++j) {
        double z0, z1;
        philox_normals_block(rng, (uint32_t)it, (uint32_t)j, z0, z1);
        qn[2 * j] = q[2 * j] + scale * z0;
        if (2 * j + 1 < n) qn[2 * j + 1] = q[2 * j + 1] + scale * z1;
      }
    }
    const double lpn = target_eval<TID, DS, true, false>(a.t, n, qn, nullptr, sm_tab);  // RWMH.py:102
    const double log_ratio = lpn - lp;
    const double log_u = injected ? log(a.b.u_inject[(size_t)it * C + chain]) : philox_accept_log_uniform(rng, (uint32_t)it);
    const bool accept = log_u < min_nan(0.0, log_ratio);  // RWMH.py:105-107
    if (accept) {
#pragma unroll
      for (int d = 0; d < n; ++d) q[d] = qn[d];
      lp = lpn;
      acc += 1;
    }
    if (si >= 0 && store) {
      tw.put(q, a.b.samples + ((size_t)si * C + chain0) * n, active);
      if (active && a.b.log_probs) a.b.log_probs[(size_t)si * C + chain] = lp;
    }
  }
  if (store) tw.finish();

  if (active) {
    if (a.b.accept_count) a.b.accept_count[chain] = acc;
    if (a.b.accept_rate) a.b.accept_rate[chain] = (float)acc / (float)S;  // RWMH.py:177
    if (a.b.final_q) store_row<DS>(a.b.final_q + chain * n, q, n);
    if (a.b.final_lp) a.b.final_lp[chain] = lp;
  }
  if (a.b.counters) {
    unsigned int v = active ? (unsigned int)acc : 0u;
    v = __reduce_add_sync(0xffffffffu, v);
    if ((tid & 31) == 0 && v) atomicAdd(a.b.counters + 0, (unsigned long long)v);
  }
}

template <int TID, int DS, int BLOCK>
int launch_rwmh_variant(const Launch& l) {
  RwmhKArgs a;
  a.p = l.p;
  a.b = l.b;
  a.t = make_target_ctx(l.p, l.b.cov_inv);
  const int n = l.p.n_dim;
  const bool store = !(l.p.flags & MCMC_FLAG_NO_SAMPLES) && l.b.samples != nullptr;
  const int tabn = target_table_len(TID, n);
  const size_t smem = ((size_t)kRngTableDoubles + (size_t)((tabn + 1) / 2) * 2 + (store ? 2 * (size_t)n * BLOCK : 0)) * sizeof(double);
  auto kern = rwmh_kernel<TID, DS, BLOCK>;
  if (smem > 227 * 1024) {
    set_error("rwmh: %zu bytes of shared memory needed exceed 227 KB", smem);
    return MCMC_ENOTSUP;
  }
  if (int rc = cuda_fail(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem),
                         "cudaFuncSetAttribute(rwmh)"))
    return rc;
  const int64_t grid = (l.p.n_chains + BLOCK - 1) / BLOCK;
  if (grid == 0) return MCMC_OK;
  kern<<<(unsigned)grid, BLOCK, smem, l.stream>>>(a);
  return cuda_fail(cudaGetLastError(), "rwmh_kernel launch");
}

}  // namespace mcmc
