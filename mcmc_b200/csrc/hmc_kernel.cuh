// hmc_kernel.cuh — the whole hmc_run / rahmc_run loop as ONE kernel launch.
//
// One thread owns one chain for all burn_in + num_samples iterations:
// position, momentum and gradient stay in fp64 registers across the L leapfrog
// steps; the pre-trajectory state (for rejection) is parked in shared memory.
//
// Replaces (reference file:line):
//   hmc_run   samplers/HMC.py:259-314    rahmc_run  samplers/GRAHMC.py:398-456
//   hmc_step  samplers/HMC.py:162-219    rahmc_step samplers/GRAHMC.py:298-349
//   leapfrog  samplers/HMC.py:116-134    _trajectory_with_schedule / _conformal_leapfrog_step
//                                        samplers/GRAHMC.py:246-261 / :190-211
//   friction schedules samplers/GRAHMC.py:69-106 (table s_i = exp(-gamma_i*eps/2) built in the prologue)
#pragma once
#include "common.cuh"
#include "philox.cuh"
#include "rowstore.cuh"
#include "targets.cuh"

namespace mcmc {

struct HmcKArgs {
  McmcParams p;
  McmcBuffers b;
  TargetCtx t;
};

// gamma(t) of samplers/GRAHMC.py:69-106 with t = i*eps, T = eps*L (:246,251)
__device__ __forceinline__ double friction_gamma_dev(int sched, double t, double T, double gam, double steep) {
  switch (sched) {
    case MCMC_SCHED_CONSTANT: return (t < T / 2) ? -gam : gam;
    case MCMC_SCHED_TANH: return gam * tanh(steep * (2.0 * t / T - 1.0));
    case MCMC_SCHED_SIGMOID: return gam * (2.0 / (1.0 + exp(-(steep * (t / T - 0.5)))) - 1.0);
    case MCMC_SCHED_LINEAR: return -gam + (2.0 * gam * t / T);
    default: return gam * sin(kPi * (t / T - 0.5));
  }
}

// shared-memory carve-up (doubles)
template <int DS, int BLOCK>
struct HmcSmem {
  int n, L, tabn;
  bool staged;
  __host__ __device__ HmcSmem(int n_, int L_, int tabn_, bool staged_) : n(n_), L(L_), tabn(tabn_), staged(staged_) {}
  __host__ __device__ size_t off_w() const { return 0; }
  __host__ __device__ size_t off_epsw() const { return (size_t)n; }
  __host__ __device__ size_t off_sqw() const { return 2 * (size_t)n; }
  __host__ __device__ size_t off_s() const { return 3 * (size_t)n; }                 // s_i
  __host__ __device__ size_t off_a() const { return off_s() + (size_t)L; }           // merged scale
  __host__ __device__ size_t off_b() const { return off_a() + (size_t)L; }           // merged kick coeff
  __host__ __device__ size_t off_tab() const { return off_b() + (size_t)L; }
  __host__ __device__ size_t off_q() const { return ((off_tab() + (size_t)tabn + 1) / 2) * 2; }
  __host__ __device__ size_t off_g() const { return off_q() + (size_t)n * BLOCK; }
  __host__ __device__ size_t off_stage() const { return off_g() + (size_t)n * BLOCK; }
  __host__ __device__ size_t total_doubles() const { return off_stage() + (staged ? 2 * (size_t)n * BLOCK : 0); }
};

template <int TID, int DS, bool FRICTION, int BLOCK>
__global__ void __launch_bounds__(BLOCK) hmc_kernel(const HmcKArgs a) {
  constexpr int N = DS > 0 ? DS : MCMC_DYN_MAXD;
  const int n = DS > 0 ? DS : a.p.n_dim;
  const int L = a.p.num_steps;
  const int64_t C = a.p.n_chains;
  const int64_t S = a.p.num_samples, B = a.p.burn_in;
  const int tid = threadIdx.x;
  const int64_t chain0 = (int64_t)blockIdx.x * BLOCK;
  const bool active = chain0 + tid < C;
  const int64_t chain = active ? chain0 + tid : C - 1;  // tail threads shadow the last chain, write nothing
  const int nvalid = (int)((C - chain0) < BLOCK ? (C - chain0) : BLOCK);
  const bool store = !(a.p.flags & MCMC_FLAG_NO_SAMPLES) && a.b.samples != nullptr;
  const bool track = (a.p.flags & MCMC_FLAG_TRACK_PROPOSALS) != 0;
  const bool do_w = (a.p.flags & MCMC_FLAG_ACCUM_WELFORD) && a.b.welford_mean && a.b.welford_m2;
  const bool injected = a.p.rng_mode == MCMC_RNG_INJECTED;

  extern __shared__ __align__(16) double smem[];
  const HmcSmem<DS, BLOCK> lay(n, L, target_table_len(TID, n), store);
  double* sm_w = smem + lay.off_w();
  double* sm_epsw = smem + lay.off_epsw();
  double* sm_sqw = smem + lay.off_sqw();
  double* sm_s = smem + lay.off_s();
  double* sm_a = smem + lay.off_a();
  double* sm_b = smem + lay.off_b();
  double* sm_tab = smem + lay.off_tab();
  double* sm_q = smem + lay.off_q();
  double* sm_g = smem + lay.off_g();

  const double eps = a.p.step_size;
  const double half_eps = 0.5 * eps;  // HMC.py:119 `half * step_sz`; GRAHMC.py:193

  // ---- prologue: per-dimension constants, friction table, target table
  for (int d = tid; d < n; d += BLOCK) {
    const double w = a.b.inv_mass ? a.b.inv_mass[d] : 1.0;
    sm_w[d] = w;
    sm_epsw[d] = eps * w;
#if MCMC_IS_FAITHFUL
    sm_sqw[d] = sqrt(w);  // p0 = z / sqrt(w), HMC.py:172
#else
    sm_sqw[d] = 1.0 / sqrt(w);
#endif
  }
  if (FRICTION) {
    const double T = eps * L;  // GRAHMC.py:246
    for (int i = tid; i < L; i += BLOCK) {
      const double gi = friction_gamma_dev(a.p.schedule_id, i * eps, T, a.p.gamma, a.p.steepness);
      sm_s[i] = exp(-gi * half_eps);  // GRAHMC.py:194
    }
  }
  target_table_fill<TID>(a.t, n, sm_tab);
  __syncthreads();
  if (FRICTION && !MCMC_IS_FAITHFUL) {
    // merged coefficients: between steps i and i+1,  p <- p*(s_i s_{i+1}) + g*(h s_i s_{i+1} + h)
    for (int i = tid; i + 1 < L; i += BLOCK) {
      const double ss = sm_s[i] * sm_s[i + 1];
      sm_a[i] = ss;
      sm_b[i] = half_eps * ss + half_eps;
    }
    __syncthreads();
  }

  TileWriter<DS, BLOCK> tw;
  if (store) tw.init(smem + lay.off_stage(), n, nvalid, a.b.samples + chain0 * n);

  RngCtx rng;
  {
    const uint64_t gid = (uint64_t)(a.p.chain_offset + chain);
    rng.k0 = a.p.seed[0]; rng.k1 = a.p.seed[1];
    rng.gid_lo = (uint32_t)gid; rng.gid_hi = (uint32_t)(gid >> 32);
  }

  // ---- hmc_init (HMC.py:71-82): value and gradient at the initial position
  double q[N], p[N], g[N];
  {
    const double* src = a.b.init_q + chain * n;
#pragma unroll
    for (int d = 0; d < n; ++d) q[d] = src[d];
  }
  double lp = target_eval<TID, DS, true, true>(a.t, n, q, g, sm_tab);
#pragma unroll
  for (int d = 0; d < n; ++d) {
    sm_q[d * BLOCK + tid] = q[d];
    sm_g[d * BLOCK + tid] = g[d];
  }
  int acc = 0;
  double wcount = a.b.welford_count;

  for (int64_t it = 0; it < B + S; ++it) {
    if (it == B) acc = 0;  // HMC.py:276-282
    const int64_t si = it - B;

    // ---- momentum p0 = z / sqrt(w) (HMC.py:172), H0 (HMC.py:176-177)
    if (injected) {
      const double* zsrc = a.b.z_inject + ((size_t)it * C + chain) * n;
#pragma unroll
      for (int d = 0; d < n; ++d) p[d] = zsrc[d];
    } else {
#pragma unroll
      for (int j = 0; 2 * j < n; ++j) {
        double z0, z1;
        philox_normals_block(rng, (uint32_t)it, (uint32_t)j, z0, z1);
        p[2 * j] = z0;
        if (2 * j + 1 < n) p[2 * j + 1] = z1;
      }
    }
    double kin = 0.0;
#pragma unroll
    for (int d = 0; d < n; ++d) {
#if MCMC_IS_FAITHFUL
      p[d] = p[d] / sm_sqw[d];
#else
      p[d] = p[d] * sm_sqw[d];
#endif
      kin += (p[d] * p[d]) * sm_w[d];
    }
    const double H0 = -lp + 0.5 * kin;

    // ---- trajectory: registers q, g hold the current state on entry
    double lpn = lp;
#if MCMC_IS_FAITHFUL
    for (int i = 0; i < L; ++i) {
      const double s = FRICTION ? sm_s[i] : 1.0;
#pragma unroll
      for (int d = 0; d < n; ++d) {
        if (FRICTION) p[d] = p[d] * s;        // GRAHMC.py:197
        p[d] = p[d] + half_eps * g[d];        // HMC.py:119
        q[d] = q[d] + eps * (p[d] * sm_w[d]); // HMC.py:121
      }
      if (i == L - 1) lpn = target_eval<TID, DS, true, true>(a.t, n, q, g, sm_tab);
      else target_eval<TID, DS, false, true>(a.t, n, q, g, sm_tab);  // intermediate log-probs are never read
#pragma unroll
      for (int d = 0; d < n; ++d) {
        p[d] = p[d] + half_eps * g[d];        // HMC.py:127
        if (FRICTION) p[d] = p[d] * s;        // GRAHMC.py:209
      }
    }
#else
    if (L > 0) {
      // first half-kick (with the first friction scaling)
      if (FRICTION) {
        const double s0 = sm_s[0];
#pragma unroll
        for (int d = 0; d < n; ++d) p[d] = fma(p[d], s0, half_eps * g[d]);
      } else {
#pragma unroll
        for (int d = 0; d < n; ++d) p[d] = fma(half_eps, g[d], p[d]);
      }
      for (int i = 0; i < L - 1; ++i) {
#pragma unroll
        for (int d = 0; d < n; ++d) q[d] = fma(p[d], sm_epsw[d], q[d]);
        target_eval<TID, DS, false, true>(a.t, n, q, g, sm_tab);
        // second half-kick of step i merged with the first half-kick of step i+1
        if (FRICTION) {
          const double A = sm_a[i], Bc = sm_b[i];
#pragma unroll
          for (int d = 0; d < n; ++d) p[d] = fma(p[d], A, g[d] * Bc);
        } else {
#pragma unroll
          for (int d = 0; d < n; ++d) p[d] = fma(eps, g[d], p[d]);
        }
      }
#pragma unroll
      for (int d = 0; d < n; ++d) q[d] = fma(p[d], sm_epsw[d], q[d]);
      lpn = target_eval<TID, DS, true, true>(a.t, n, q, g, sm_tab);
      if (FRICTION) {
        const double sl = sm_s[L - 1];
#pragma unroll
        for (int d = 0; d < n; ++d) p[d] = fma(half_eps, g[d], p[d]) * sl;
      } else {
#pragma unroll
        for (int d = 0; d < n; ++d) p[d] = fma(half_eps, g[d], p[d]);
      }
    }
#endif

    // ---- p <- -p, H1 with the non-finite guard, MH test (HMC.py:194-207)
    double kin1 = 0.0;
#pragma unroll
    for (int d = 0; d < n; ++d) kin1 += (p[d] * p[d]) * sm_w[d];  // (-p)^2 == p^2
    double H1 = -lpn + 0.5 * kin1;
    if (!isfinite(H1)) H1 = 1e10;
    const double log_alpha = H0 - H1;
    const double u = injected ? a.b.u_inject[(size_t)it * C + chain] : philox_accept_uniform(rng, (uint32_t)it);
    const bool accept = log(u) < min_nan(0.0, log_alpha);

    if (track && si >= 0 && active) {  // HMC.py:287-305
      const size_t o = (size_t)si * C + chain;
      if (a.b.prop_q) store_row<DS>(a.b.prop_q + o * n, q, n);
      if (a.b.prop_lp) a.b.prop_lp[o] = lpn;
      if (a.b.delta_h) a.b.delta_h[o] = H1 - H0;
      if (a.b.pre_lp) a.b.pre_lp[o] = lp;
      if (a.b.pre_q) {
        double* dst = a.b.pre_q + o * n;
#pragma unroll
        for (int d = 0; d < n; ++d) dst[d] = sm_q[d * BLOCK + tid];
      }
    }

    if (accept) {  // HMC.py:209-212
#pragma unroll
      for (int d = 0; d < n; ++d) {
        sm_q[d * BLOCK + tid] = q[d];
        sm_g[d * BLOCK + tid] = g[d];
      }
      lp = lpn;
      acc += 1;
    } else {
#pragma unroll
      for (int d = 0; d < n; ++d) {
        q[d] = sm_q[d * BLOCK + tid];
        g[d] = sm_g[d * BLOCK + tid];
      }
    }

    if (si >= 0) {
      if (store) {
        tw.put(q, a.b.samples + ((size_t)si * C + chain0) * n, active);
        if (active && a.b.log_probs) a.b.log_probs[(size_t)si * C + chain] = lp;
      }
      if (do_w && active) {  // tuning/welford.py:42-59, one fold per stored sample (adaptation.py:229-235)
        wcount += 1.0;
        double* wm = a.b.welford_mean + chain * n;
        double* w2 = a.b.welford_m2 + chain * n;
#pragma unroll
        for (int d = 0; d < n; ++d) {
          const double mean = wm[d];
          const double delta = q[d] - mean;
          const double mean2 = mean + delta / wcount;
          wm[d] = mean2;
          w2[d] = w2[d] + delta * (q[d] - mean2);
        }
      }
    }
  }
  if (store) tw.finish();

  // ---- epilogue: final state, acceptance statistics
  if (active) {
    if (a.b.accept_count) a.b.accept_count[chain] = acc;
    if (a.b.accept_rate) a.b.accept_rate[chain] = (float)acc / (float)S;  // HMC.py:313
    if (a.b.final_q) store_row<DS>(a.b.final_q + chain * n, q, n);
    if (a.b.final_g) store_row<DS>(a.b.final_g + chain * n, g, n);
    if (a.b.final_lp) a.b.final_lp[chain] = lp;
  }
  if (a.b.counters) {
    // fused cross-chain reduction (adaptation.py:218): exact integer sum of accept counts
    unsigned int v = active ? (unsigned int)acc : 0u;
    v = __reduce_add_sync(0xffffffffu, v);
    if ((tid & 31) == 0 && v) atomicAdd(a.b.counters + 0, (unsigned long long)v);
    if (blockIdx.x == 0 && tid == 0)
      atomicAdd(a.b.counters + 1, (unsigned long long)C * (unsigned long long)(B + S) * (unsigned long long)L);
  }
}

// host-side launcher for one (target, D-variant)
template <int TID, int DS, bool FRICTION, int BLOCK>
int launch_hmc_variant(const Launch& l) {
  HmcKArgs a;
  a.p = l.p;
  a.b = l.b;
  a.t = make_target_ctx(l.p, l.b.cov_inv);
  const int n = l.p.n_dim;
  const bool store = !(l.p.flags & MCMC_FLAG_NO_SAMPLES) && l.b.samples != nullptr;
  const HmcSmem<DS, BLOCK> lay(n, l.p.num_steps, target_table_len(TID, n), store);
  const size_t smem = lay.total_doubles() * sizeof(double);
  auto kern = hmc_kernel<TID, DS, FRICTION, BLOCK>;
  if (smem > 227 * 1024) {
    set_error("hmc: %zu bytes of shared memory needed (D=%d, L=%d) exceed 227 KB", smem, n, l.p.num_steps);
    return MCMC_ENOTSUP;
  }
  if (int rc = cuda_fail(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem),
                         "cudaFuncSetAttribute(hmc)"))
    return rc;
  const int64_t grid = (l.p.n_chains + BLOCK - 1) / BLOCK;
  if (grid == 0) return MCMC_OK;
  kern<<<(unsigned)grid, BLOCK, smem, l.stream>>>(a);
  return cuda_fail(cudaGetLastError(), "hmc_kernel launch");
}

}  // namespace mcmc
