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
#include <cstdlib>

#include "common.cuh"
#include "philox.cuh"
#include "rowstore.cuh"
#include "targets.cuh"

// leapfrog steps per trip of the trajectory loop: with several steps in one basic block ptxas overlaps the
// serial tail of one step's gradient with the drift of the next (2 warps per scheduler cannot hide it)
#ifndef MCMC_STEP_UNROLL
#define MCMC_STEP_UNROLL 4
#endif

namespace mcmc {

constexpr int kStepUnroll = MCMC_STEP_UNROLL;
// Long trajectories run an instantiation register-allocated for SIX resident blocks per SM (168 registers, 3 warps per
// scheduler; ptxas spills ~330 B per thread, none of it inside the trajectory loop): the loop then keeps the fp64 pipe
// 97% busy instead of 88%, while the per-iteration part gets ~30% slower (instruction-fetch stalls, spill reloads) —
// a net gain from L ~ 90 upwards (profiles/hmc_occupancy_sweep_r01.txt: L=128 +4.8%, L=64 -2%, L=16 -11%).
#ifndef MCMC_HMC_DENSE_MIN_L
#define MCMC_HMC_DENSE_MIN_L 96
#endif
#ifndef MCMC_HMC_DENSE_BLOCKS
#define MCMC_HMC_DENSE_BLOCKS 6
#endif
// pairs per stage-interleaved draw batch (philox.cuh; 0: the pair-after-pair draw; must divide DS/2, else the whole row is
// one batch).  Measured at the bench shape (profiles/draw_batch_ab_r02.txt): 0 / 2 / 5 / 10 give the same times at every L —
// the per-iteration part of this kernel is not bound by the draw's instruction scheduling — so the default stays 0.
#ifndef MCMC_HMC_DRAW_BATCH
#define MCMC_HMC_DRAW_BATCH 0
#endif
constexpr int kHmcDrawBatch = MCMC_HMC_DRAW_BATCH;
// Two experiment switches, both OFF by default (profiles/hmc_kin_ab_r02.txt: neither pays).  MCMC_HMC_KIN_WAYS=4 keeps
// four interleaved partial sums for the two kinetic energies of an iteration instead of D dependent DFMAs;
// MCMC_HMC_EARLY_LOAD=1 issues the shared-memory loads of the current state before the momentum draw.  With 16 warps
// per SM the other warps already cover those latencies: the sum over the L grid moved by +0.2% / +0.5%, and the early
// load costs the standard normal 10% (registers held across the draw).
#ifndef MCMC_HMC_KIN_WAYS
#define MCMC_HMC_KIN_WAYS 1
#endif
constexpr int kHmcKinWays = MCMC_IS_FAITHFUL ? 1 : MCMC_HMC_KIN_WAYS;
#ifndef MCMC_HMC_EARLY_LOAD
#define MCMC_HMC_EARLY_LOAD 0
#endif
template <int W>
__device__ __forceinline__ double kin_fold(const double (&k)[W]) {
  if (W == 4) return (k[0] + k[1 % W]) + (k[2 % W] + k[3 % W]);
  if (W == 2) return k[0] + k[1 % W];
  double s = k[0];
#pragma unroll
  for (int i = 1; i < W; ++i) s += k[i];
  return s;
}
constexpr int kHmcDenseMinL = MCMC_HMC_DENSE_MIN_L;
constexpr int kHmcDenseBlocks = MCMC_HMC_DENSE_BLOCKS;

struct HmcKArgs {
  McmcParams p;
  McmcBuffers b;
  TargetCtx t;
  // wave-balanced scheduling (see hmc_kernel): work items = (chain group, iteration chunk)
  int n_groups;          // ceil(C / BLOCK)
  int n_chunks;          // 1: one block owns a group for the whole run (no queue, no hand-off)
  int chunk_iters;       // iterations per chunk
  unsigned int* queue;   // [1]  next item
  unsigned int* flags;   // [n_groups] chunks completed per group
  double* ho_q;          // [C*D] hand-off state between chunks
  double* ho_g;          // [C*D]
  double* ho_lp;         // [C]
  int* ho_acc;           // [C]
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
  bool welf;  // per-chain Welford tiles resident (MCMC_FLAG_ACCUM_WELFORD)
  __host__ __device__ HmcSmem(int n_, int L_, int tabn_, bool welf_) : n(n_), L(L_), tabn(tabn_), welf(welf_) {}
  __host__ __device__ size_t off_w() const { return 0; }
  __host__ __device__ size_t off_epsw() const { return (size_t)n; }
  __host__ __device__ size_t off_sqw() const { return 2 * (size_t)n; }
  __host__ __device__ size_t off_s() const { return 3 * (size_t)n; }                 // s_i
  __host__ __device__ size_t off_a() const { return off_s() + (size_t)L; }           // merged scale
  __host__ __device__ size_t off_b() const { return off_a() + (size_t)L; }           // merged kick coeff
  __host__ __device__ size_t off_logtab() const { return off_b() + (size_t)L; }
  __host__ __device__ size_t off_tab() const { return off_logtab() + kRngTableDoubles; }
  __host__ __device__ size_t off_q() const { return ((off_tab() + (size_t)tabn + 1) / 2) * 2; }
  __host__ __device__ size_t off_g() const { return off_q() + (size_t)n * BLOCK; }
  __host__ __device__ size_t off_wm() const { return off_g() + (size_t)n * BLOCK; }  // Welford mean [d][tid]
  __host__ __device__ size_t off_w2() const { return off_wm() + (welf ? (size_t)n * BLOCK : 0); }
  __host__ __device__ size_t total_doubles() const { return off_w2() + (welf ? (size_t)n * BLOCK : 0); }
};

// WELF: the per-chain Welford accumulation (tuning batches) is compiled only into its own instantiation, so that the
// sampling kernel's register allocation does not depend on it
template <int TID, int DS, bool FRICTION, int BLOCK, bool WELF = false, int MINB = 1>
__global__ void __launch_bounds__(BLOCK, MINB) hmc_kernel(const HmcKArgs a) {
  constexpr int N = DS > 0 ? DS : MCMC_DYN_MAXD;
  const int n = DS > 0 ? DS : a.p.n_dim;
  const int L = a.p.num_steps;
  const int64_t C = a.p.n_chains;
  const int64_t S = a.p.num_samples, B = a.p.burn_in;
  const int tid = threadIdx.x;
  const bool store = !(a.p.flags & MCMC_FLAG_NO_SAMPLES) && a.b.samples != nullptr;
  const bool track = (a.p.flags & MCMC_FLAG_TRACK_PROPOSALS) != 0;
  constexpr bool do_w = WELF;
  const bool injected = a.p.rng_mode == MCMC_RNG_INJECTED;

  extern __shared__ __align__(16) double smem[];
  const HmcSmem<DS, BLOCK> lay(n, L, target_table_len(TID, n), do_w);
  double* sm_w = smem + lay.off_w();
  double* sm_epsw = smem + lay.off_epsw();
  double* sm_sqw = smem + lay.off_sqw();
  double* sm_s = smem + lay.off_s();
  double* sm_a = smem + lay.off_a();
  double* sm_b = smem + lay.off_b();
  double* sm_tab = smem + lay.off_tab();
  double* sm_q = smem + lay.off_q();
  double* sm_g = smem + lay.off_g();
  double* sm_wm = smem + lay.off_wm();
  double* sm_w2 = smem + lay.off_w2();

  const double eps = a.b.step_size_dev ? __ldg(a.b.step_size_dev) : a.p.step_size;
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
  rng_table_fill(smem + lay.off_logtab());
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

  // Work distribution.  With C chains and only (SMs x resident blocks) slots, equal-length blocks
  // leave the last wave partly empty (65,536 chains = 1.73 waves of 256 threads/SM: 13% lost).  So a
  // block does not own a group of chains for the whole run: the run is cut into iteration chunks,
  // items (chunk c, group g) are handed out in chunk-major order from an atomic queue, and the state
  // of a group travels between chunks through HBM.  An item waits (flags[g] >= c) only on an item
  // that a resident block has already taken, so the wait always ends.  Philox counters are keyed on
  // (global chain, absolute iteration): the result does not depend on who runs which item.
  __shared__ unsigned int s_item;
  double* my_q_base = sm_q;
  double q[N], p[N], g[N];
  const unsigned int n_items = (unsigned int)a.n_groups * (unsigned int)a.n_chunks;
  for (;;) {
  unsigned int item;
  if (a.n_chunks == 1) {
    item = blockIdx.x;
    if (item >= n_items) break;
  } else {
    if (tid == 0) s_item = atomicAdd(a.queue, 1u);
    __syncthreads();
    item = s_item;
    __syncthreads();
    if (item >= n_items) break;
  }
  const int grp = (int)(item % (unsigned int)a.n_groups);
  const int chunk = (int)(item / (unsigned int)a.n_groups);
  const int64_t it_begin = (int64_t)chunk * a.chunk_iters;
  const int64_t it_end = (it_begin + a.chunk_iters < B + S) ? it_begin + a.chunk_iters : B + S;
  const bool last_chunk = chunk == a.n_chunks - 1;
  const int64_t chain0 = (int64_t)grp * BLOCK;
  const bool active = chain0 + tid < C;
  const int64_t chain = active ? chain0 + tid : C - 1;  // tail threads shadow the last chain, write nothing
  const int nvalid = (int)((C - chain0) < BLOCK ? (C - chain0) : BLOCK);

  // sample rows leave straight from the row-major current-state tile sm_q (see below)
  const bool bulk = store && ((((size_t)nvalid * n * sizeof(double)) & 15) == 0) &&
                    ((((size_t)(a.b.samples + chain0 * n)) & 15) == 0) && ((((size_t)C * n * sizeof(double)) & 15) == 0);

  RngCtx rng;
  {
    const uint64_t gid = (uint64_t)(a.p.chain_offset + chain);
    rng.k0 = a.p.seed[0]; rng.k1 = a.p.seed[1];
    rng.gid_lo = (uint32_t)gid; rng.gid_hi = (uint32_t)(gid >> 32);
    rng.logtab = smem + lay.off_logtab();
  }

  // Current state of the chain: q row-major [tid][n] in shared memory (it doubles as the TMA source
  // of the sample tile), g as [d][tid]; registers hold q/p/g only during a trajectory.
  double* my_q = my_q_base + (size_t)tid * n;
  double lp;
  int acc;
  if (chunk == 0) {
  // ---- hmc_init (HMC.py:71-82): value and gradient at the initial position
  {
    const double* src = a.b.init_q + chain * n;
#pragma unroll
    for (int d = 0; d < n; ++d) q[d] = src[d];
  }
  lp = target_eval<TID, DS, true, true>(a.t, n, q, g, sm_tab);
  acc = 0;
  } else {
    // ---- resume: wait for the previous chunk of this group, then copy its state tiles from HBM (L2) straight
    // into shared memory: the hand-off arrays hold the block's sm_q / sm_g tiles verbatim, so the copy is flat
    // and fully coalesced
    if (tid == 0) {
      while (ld_acquire_u32(a.flags + grp) < (unsigned int)chunk) __nanosleep(200);
    }
    __syncthreads();
    const double* sq = a.ho_q + (size_t)chain0 * n;
    const double* sg = a.ho_g + (size_t)chain0 * n;
    for (int k = tid; k < BLOCK * n; k += BLOCK) {
      sm_q[k] = __ldcg(sq + k);
      sm_g[k] = __ldcg(sg + k);
    }
    lp = __ldcg(a.ho_lp + chain0 + tid);
    acc = __ldcg(a.ho_acc + chain0 + tid);
    __syncthreads();
  }
  if (chunk == 0) {
    store_row<DS>(my_q, q, n);
#pragma unroll
    for (int d = 0; d < n; ++d) sm_g[d * BLOCK + tid] = g[d];
  }
  double wcount = a.b.welford_count + (double)((it_begin > B) ? (it_begin - B) : 0);
  if (do_w) {
    // the group's Welford states stay in shared memory ([d][tid]) for the whole item: one coalesced read here, one
    // coalesced write after the last iteration, instead of 4*D strided L2 accesses per chain and stored sample
    const double* gm = a.b.welford_mean + chain0 * n;
    const double* g2 = a.b.welford_m2 + chain0 * n;
    for (int k = tid; k < nvalid * n; k += BLOCK) {
      const int r = k / n, d = k - r * n;
      sm_wm[d * BLOCK + r] = __ldcg(gm + k);  // L2: another SM may have run the previous chunk
      sm_w2[d * BLOCK + r] = __ldcg(g2 + k);
    }
    __syncthreads();
  }

  for (int64_t it = it_begin; it < it_end; ++it) {
    if (it == B) acc = 0;  // HMC.py:276-282
    const int64_t si = it - B;

#if MCMC_HMC_EARLY_LOAD
    load_row<DS>(q, my_q, n);
#pragma unroll
    for (int d = 0; d < n; ++d) g[d] = sm_g[d * BLOCK + tid];
#endif
    // ---- momentum p0 = z / sqrt(w) (HMC.py:172), H0 (HMC.py:176-177)
    if (injected) {
      const double* zsrc = a.b.z_inject + ((size_t)it * C + chain) * n;
#pragma unroll
      for (int d = 0; d < n; ++d) p[d] = zsrc[d];
    } else if (DS > 0 && DS % 2 == 0 && kHmcDrawBatch > 0) {
      // all DS/2 Philox blocks and Box-Muller pairs advanced stage by stage (philox.cuh): same bits, shorter stalls
      constexpr int kPairs = DS > 0 ? DS / 2 : 1;
      constexpr int kNB = (kHmcDrawBatch > 0 && kPairs % kHmcDrawBatch == 0) ? kHmcDrawBatch : kPairs;
#pragma unroll
      for (int j0 = 0; j0 < kPairs; j0 += kNB) {
        double zb[2 * kNB];
        philox_normals_batch<kNB>(rng, (uint32_t)it, (uint32_t)j0, zb);
#pragma unroll
        for (int k = 0; k < 2 * kNB; ++k) p[2 * j0 + k] = zb[k];
      }
    } else {
#pragma unroll
      for (int j = 0; 2 * j < n; ++j) {
        double z0, z1;
        philox_normals_block(rng, (uint32_t)it, (uint32_t)j, z0, z1);
        p[2 * j] = z0;
        if (2 * j + 1 < n) p[2 * j + 1] = z1;
      }
    }
    double kin0[kHmcKinWays] = {0.0};
#pragma unroll
    for (int d = 0; d < n; ++d) {
#if MCMC_IS_FAITHFUL
      p[d] = p[d] / sm_sqw[d];
#else
      p[d] = p[d] * sm_sqw[d];
#endif
      kin0[d % kHmcKinWays] += (p[d] * p[d]) * sm_w[d];
    }
    const double H0 = -lp + 0.5 * kin_fold(kin0);

    // ---- trajectory from the current state (cached lp and g: no re-evaluation, HMC.py:179-192)
#if !MCMC_HMC_EARLY_LOAD
    load_row<DS>(q, my_q, n);
#pragma unroll
    for (int d = 0; d < n; ++d) g[d] = sm_g[d * BLOCK + tid];
#endif
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
#pragma unroll kStepUnroll
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
    double kin1[kHmcKinWays] = {0.0};
#pragma unroll
    for (int d = 0; d < n; ++d) kin1[d % kHmcKinWays] += (p[d] * p[d]) * sm_w[d];  // (-p)^2 == p^2
    double H1 = -lpn + 0.5 * kin_fold(kin1);
    if (!isfinite(H1)) H1 = 1e10;
    const double log_alpha = H0 - H1;
    const double log_u = injected ? log(a.b.u_inject[(size_t)it * C + chain]) : philox_accept_log_uniform(rng, (uint32_t)it);
    const bool accept = log_u < min_nan(0.0, log_alpha);

    if (track && si >= 0 && active) {  // HMC.py:287-305
      const size_t o = (size_t)si * C + chain;
      if (a.b.prop_q) store_row<DS>(a.b.prop_q + o * n, q, n);
      if (a.b.prop_lp) a.b.prop_lp[o] = lpn;
      if (a.b.delta_h) a.b.delta_h[o] = H1 - H0;
      if (a.b.pre_lp) a.b.pre_lp[o] = lp;
      if (a.b.pre_q) {
        load_row<DS>(p, my_q, n);  // p is dead here
        store_row<DS>(a.b.pre_q + o * n, p, n);
      }
    }

    // the TMA copy of the previous iteration's tile must have drained sm_q before it is overwritten
    if (bulk) {
      if (tid == 0) bulk_wait_read<0>();
      __syncthreads();
    }
    if (accept) {  // HMC.py:209-212
      store_row<DS>(my_q, q, n);
#pragma unroll
      for (int d = 0; d < n; ++d) sm_g[d * BLOCK + tid] = g[d];
      lp = lpn;
      acc += 1;
    }

    if (si >= 0) {
      if (store) {
        double* gtile = a.b.samples + ((size_t)si * C + chain0) * n;
        if (bulk) {
          fence_proxy_async_smem();
          __syncthreads();
          if (tid == 0) bulk_store_s2g(gtile, sm_q, (uint32_t)((size_t)nvalid * n * sizeof(double)));
        } else if (active) {
          if (!accept) load_row<DS>(q, my_q, n);
          store_row<DS>(gtile + (size_t)tid * n, q, n);
        }
        if (active && a.b.log_probs) a.b.log_probs[(size_t)si * C + chain] = lp;
      }
      if (do_w) {  // tuning/welford.py:42-59, one fold per stored sample (adaptation.py:229-235)
        if (!accept) load_row<DS>(q, my_q, n);
        wcount += 1.0;
#if !MCMC_IS_FAITHFUL
        const double rcount = 1.0 / wcount;
#endif
#pragma unroll
        for (int d = 0; d < n; ++d) {
          const double mean = sm_wm[d * BLOCK + tid];
          const double delta = q[d] - mean;
#if MCMC_IS_FAITHFUL
          const double mean2 = mean + delta / wcount;
#else
          const double mean2 = fma(delta, rcount, mean);
#endif
          sm_wm[d * BLOCK + tid] = mean2;
          sm_w2[d * BLOCK + tid] += delta * (q[d] - mean2);
        }
      }
    }
  }
  if (bulk && tid == 0) bulk_wait_all();
  if (do_w) {
    __syncthreads();
    double* gm = a.b.welford_mean + chain0 * n;
    double* g2 = a.b.welford_m2 + chain0 * n;
    for (int k = tid; k < nvalid * n; k += BLOCK) {
      const int r = k / n, d = k - r * n;
      __stcg(gm + k, sm_wm[d * BLOCK + r]);
      __stcg(g2 + k, sm_w2[d * BLOCK + r]);
    }
  }

  if (!last_chunk) {
    // ---- hand the group's state tiles to whoever takes its next chunk (flat coalesced copy of sm_q / sm_g)
    __syncthreads();
    double* dq = a.ho_q + (size_t)chain0 * n;
    double* dg = a.ho_g + (size_t)chain0 * n;
    for (int k = tid; k < BLOCK * n; k += BLOCK) {
      __stcg(dq + k, sm_q[k]);
      __stcg(dg + k, sm_g[k]);
    }
    __stcg(a.ho_lp + chain0 + tid, lp);
    __stcg(a.ho_acc + chain0 + tid, acc);
    __threadfence();
    __syncthreads();
    if (tid == 0) st_release_u32(a.flags + grp, (unsigned int)chunk + 1u);
  } else {
  // ---- epilogue: final state, acceptance statistics
  if (S == 0) acc = 0;  // the counters are reset after burn-in even when nothing is sampled (HMC.py:276-282)
  load_row<DS>(q, my_q, n);
#pragma unroll
  for (int d = 0; d < n; ++d) g[d] = sm_g[d * BLOCK + tid];
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
    if (grp == 0 && tid == 0)
      atomicAdd(a.b.counters + 1, (unsigned long long)C * (unsigned long long)(B + S) * (unsigned long long)L);
  }
  }
  __syncthreads();  // sm_q / sm_g are rewritten by the next item
  if (a.n_chunks == 1) break;
  }  // item loop
}

// host-side launcher for one (target, D-variant)
template <int TID, int DS, bool FRICTION, int BLOCK, bool WELF, int MINB = 1>
int launch_hmc_variant_w(const Launch& l) {
  HmcKArgs a;
  a.p = l.p;
  a.b = l.b;
  a.t = make_target_ctx(l.p, l.b.cov_inv);
  const int n = l.p.n_dim;
  const HmcSmem<DS, BLOCK> lay(n, l.p.num_steps, target_table_len(TID, n), WELF);
  const size_t smem = lay.total_doubles() * sizeof(double);
  auto kern = hmc_kernel<TID, DS, FRICTION, BLOCK, WELF, MINB>;
  if (smem > 227 * 1024) {
    set_error("hmc: %zu bytes of shared memory needed (D=%d, L=%d) exceed 227 KB", smem, n, l.p.num_steps);
    return MCMC_ENOTSUP;
  }
  if (int rc = cuda_fail(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem),
                         "cudaFuncSetAttribute(hmc)"))
    return rc;
  const int64_t groups = (l.p.n_chains + BLOCK - 1) / BLOCK;
  if (groups == 0) return MCMC_OK;
  a.n_groups = (int)groups;
  a.n_chunks = 1;
  a.chunk_iters = (int)(l.p.burn_in + l.p.num_samples);
  a.queue = nullptr; a.flags = nullptr; a.ho_q = a.ho_g = a.ho_lp = nullptr; a.ho_acc = nullptr;
  int64_t grid = groups;
  // more groups than resident slots -> cut the run into chunks and balance them over persistent blocks
  int per_sm = 0;
  if (int rc = cuda_fail(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, BLOCK, smem), "occupancy(hmc)"))
    return rc;
  const int64_t slots = (int64_t)per_sm * device_sm_count();
  const int64_t iters = l.p.burn_in + l.p.num_samples;
  const size_t need = hmc_workspace_bytes(l.p);
  const bool tuning_off = getenv("MCMC_HMC_NO_QUEUE") != nullptr || no_spin();
  const char* force = getenv("MCMC_HMC_FORCE_CHUNKS");  // test hook: chunk even when one wave would do
  if (!tuning_off && slots > 0 && (groups > slots || force) && iters >= 2 && l.b.workspace &&
      l.b.workspace_bytes >= need) {
    // chunk length: an item should cost about kItemCost leapfrog-step equivalents (an iteration costs L steps
    // + ~14 for the momentum draw / MH test): short enough that the last wave of items is a small tail, long
    // enough that the hand-off through L2 stays in the noise; and never fewer than ~6 items per slot
    const char* cost_env = getenv("MCMC_HMC_ITEM_COST");  // tuning knob (experiments only)
    const int64_t item_cost = cost_env ? atoi(cost_env) : 450;
    int64_t ci = item_cost / (l.p.num_steps + 14);
    const int64_t ci_bal = iters * groups / (6 * slots);
    if (ci > ci_bal) ci = ci_bal;
    if (ci < 1) ci = 1;
    int64_t nc = (iters + ci - 1) / ci;
    if (force) { nc = atoi(force); if (nc < 2) nc = 2; if (nc > iters) nc = iters; ci = (iters + nc - 1) / nc; }
    nc = (iters + ci - 1) / ci;
    if (nc > 1) {
      a.n_chunks = (int)nc;
      a.chunk_iters = (int)ci;
      char* w = reinterpret_cast<char*>(l.b.workspace);
      const size_t C = ((size_t)l.p.n_chains + 255) / 256 * 256, D = (size_t)n;   // whole tiles, also for the tail group
      const size_t flag_bytes = (((C + 31) / 32) * 4 + 255) / 256 * 256;
      a.queue = reinterpret_cast<unsigned int*>(w);
      a.flags = reinterpret_cast<unsigned int*>(w + 256);
      a.ho_q = reinterpret_cast<double*>(w + 256 + flag_bytes);
      a.ho_g = a.ho_q + C * D;
      a.ho_lp = a.ho_g + C * D;
      a.ho_acc = reinterpret_cast<int*>(a.ho_lp + C);
      if (int rc = cuda_fail(cudaMemsetAsync(w, 0, 256 + flag_bytes, l.stream), "memset(hmc queue)")) return rc;
      grid = slots < groups ? slots : groups;
      if (force && grid > 1) grid = (grid + 1) / 2;  // fewer blocks than groups: items really migrate
    }
  }
  kern<<<(unsigned)grid, BLOCK, smem, l.stream>>>(a);
  return cuda_fail(cudaGetLastError(), "hmc_kernel launch");
}

template <int TID, int DS, bool FRICTION, int BLOCK>
int launch_hmc_variant(const Launch& l) {
  const bool welf = (l.p.flags & MCMC_FLAG_ACCUM_WELFORD) && l.b.welford_mean && l.b.welford_m2;
  if (DS == 20 && !FRICTION && !welf && l.p.num_steps >= kHmcDenseMinL && !getenv("MCMC_HMC_NO_DENSE"))
    return launch_hmc_variant_w<TID, DS, FRICTION, BLOCK, false, (DS == 20 && !FRICTION) ? kHmcDenseBlocks : 1>(l);
  return welf ? launch_hmc_variant_w<TID, DS, FRICTION, BLOCK, true>(l) : launch_hmc_variant_w<TID, DS, FRICTION, BLOCK, false>(l);
}

}  // namespace mcmc
