// nuts_kernel.cuh — nuts_run as one kernel launch, thread-per-chain, with the
// transition loop flattened into a per-lane state machine.
//
// Replaces samplers/NUTS.py:479-514 (nuts_run), :399-442 (nuts_step) and
// :202-396 (_nuts_step_single_chain) with its helpers :90-198.  This is the
// reference's own variant (whole-subtree validity from the endpoint only,
// endpoint-as-proposal, U-turn on raw momenta), reproduced as written.
//
// Divergent-path handling: under jax.vmap the while_loop runs every chain to
// the deepest tree of the batch.  Here a lane never waits for its neighbours at
// a transition boundary: each loop trip executes ONE leapfrog step for every
// lane that has one pending, whatever transition / doubling that lane is in, and
// the cheap bookkeeping (subtree end, U-turn test, next transition) is the only
// divergent part.  Lane utilisation is then set by the total work per chain
// (which concentrates over many transitions), not by the per-transition maximum.
//
// Tree ends (left/right q,p,g) and the proposal (q,g) are 8 D-vectors per
// chain: shared memory for compile-time D, a global workspace for runtime D.
#pragma once
#include "common.cuh"
#include "philox.cuh"
#include "rowstore.cuh"
#include "targets.cuh"

namespace mcmc {

struct NutsKArgs {
  McmcParams p;
  McmcBuffers b;
  TargetCtx t;
};

enum { SL_QL = 0, SL_PL, SL_GL, SL_QR, SL_PR, SL_GR, SL_QP, SL_GP, SL_COUNT };

template <int TID, int DS, int BLOCK>
__global__ void __launch_bounds__(BLOCK) nuts_kernel(const NutsKArgs a) {
  constexpr int N = DS > 0 ? DS : MCMC_DYN_MAXD;
  constexpr bool kSlotsInSmem = DS > 0;
  const int n = DS > 0 ? DS : a.p.n_dim;
  const int64_t C = a.p.n_chains, S = a.p.num_samples, B = a.p.burn_in;
  const int tid = threadIdx.x;
  const int64_t chain0 = (int64_t)blockIdx.x * BLOCK;
  const bool active = chain0 + tid < C;
  const int64_t chain = active ? chain0 + tid : C - 1;
  const bool store = !(a.p.flags & MCMC_FLAG_NO_SAMPLES) && a.b.samples != nullptr;
  const bool do_w = (a.p.flags & MCMC_FLAG_ACCUM_WELFORD) && a.b.welford_mean && a.b.welford_m2;
  const bool injected = a.p.rng_mode == MCMC_RNG_INJECTED;
  const int max_depth = a.p.max_tree_depth;
  const double eps = a.p.step_size;
  const double delta_max = a.p.delta_max;

  extern __shared__ __align__(16) double smem[];
  double* sm_w = smem;
  double* sm_sqw = smem + n;
  double* sm_logtab = smem + 2 * n;
  double* sm_tab = sm_logtab + kRngTableDoubles;
  const int tabn = target_table_len(TID, n);
  // slot(s, d) = slots[(s*n + d) * stride + lane]
  double* slots;
  size_t stride;
  if (kSlotsInSmem) {
    slots = smem + ((2 * n + kRngTableDoubles + tabn + 1) / 2) * 2 + tid;
    stride = BLOCK;
  } else {
    const size_t cpad = ((size_t)C + 63) / 64 * 64;  // see nuts_workspace_bytes()
    slots = reinterpret_cast<double*>(a.b.workspace) + (chain0 + tid);  // tail lanes get their own column
    stride = cpad;
  }
#define SLOT(s, d) slots[((size_t)(s) * n + (d)) * stride]

  for (int d = tid; d < n; d += BLOCK) {
    const double w = a.b.inv_mass ? a.b.inv_mass[d] : 1.0;
    sm_w[d] = w;
    sm_sqw[d] = sqrt(w);
  }
  target_table_fill<TID>(a.t, n, sm_tab);
  rng_table_fill(sm_logtab);
  __syncthreads();

  RngCtx rng;
  {
    const uint64_t gid = (uint64_t)(a.p.chain_offset + chain);
    rng.k0 = a.p.seed[0]; rng.k1 = a.p.seed[1];
    rng.gid_lo = (uint32_t)gid; rng.gid_hi = (uint32_t)(gid >> 32);
    rng.logtab = sm_logtab;
  }

  // working point of the subtree being integrated
  double qs[N], ps[N], gs[N];
  double lps;

  // nuts_init (NUTS.py:73-86): the current state lives in the proposal slots
  {
    const double* src = a.b.init_q + chain * n;
#pragma unroll
    for (int d = 0; d < n; ++d) qs[d] = src[d];
  }
  double lpp = target_eval<TID, DS, true, true>(a.t, n, qs, gs, sm_tab);
#pragma unroll
  for (int d = 0; d < n; ++d) {
    SLOT(SL_QP, d) = qs[d];
    SLOT(SL_GP, d) = gs[d];
  }

  // per-transition scalars
  double h0 = 0.0, log_u = 0.0, sum_alpha = 0.0, sa = 0.0, h_new = 0.0, u_prop = 0.0;
  double se = 0.0, hse = 0.0;
  int n_valid = 1, n_steps = 0, depth = 0, direction = 1, k = 0, num_steps = 0;
  bool diverged = false, in_subtree = false, in_transition = false, done = (B + S) == 0;
  int64_t it = 0;
  int acc = 0;
  double wcount = a.b.welford_count;
  unsigned long long n_grad = 0, sum_depth = 0, n_div = 0;

  while (!__all_sync(0xffffffffu, done)) {
    // ------------------------------------------------------------ bookkeeping (divergent, cheap)
    while (k == 0 && !done) {
      if (in_subtree) {
        // subtree endpoint: validity, divergence, end replacement, proposal swap (NUTS.py:281-378)
        const bool in_slice = log_u <= -h_new;
        const bool is_div = (h_new - h0) > delta_max;
        const bool is_valid = in_slice && !is_div;
        const int sq = direction == -1 ? SL_QL : SL_QR;
#pragma unroll
        for (int d = 0; d < n; ++d) {
          SLOT(sq, d) = qs[d];
          SLOT(sq + 1, d) = ps[d];
          SLOT(sq + 2, d) = gs[d];
        }
        const int n_valid_new = is_valid ? num_steps : 0;
        const int total_valid = n_valid + n_valid_new;
        float accept_prob = 0.0f;  // float32, NUTS.py:331-335
        if (total_valid > 0 && is_valid)
          accept_prob = (float)((double)n_valid_new / (double)(total_valid > 1 ? total_valid : 1));
        if (u_prop < (double)accept_prob) {  // NUTS.py:336-357
#pragma unroll
          for (int d = 0; d < n; ++d) {
            SLOT(SL_QP, d) = qs[d];
            SLOT(SL_GP, d) = gs[d];
          }
          lpp = lps;
        }
        n_valid = total_valid;
        sum_alpha = sum_alpha + sa;
        n_steps += num_steps;
        n_grad += (unsigned long long)num_steps;
        depth += 1;
        diverged = diverged || is_div;
        in_subtree = false;
      }
      if (!in_transition) {
        // new transition: momentum, energy, slice (NUTS.py:217-241); ps doubles as p0
        if (it == B) acc = 0;  // NUTS.py:494-500
        if (injected) {
          const double* zsrc = a.b.z_inject + ((size_t)it * C + chain) * n;
#pragma unroll
          for (int d = 0; d < n; ++d) ps[d] = zsrc[d];
        } else {
#pragma unroll
          for (int j = 0; 2 * j < n; ++j) {
            double z0, z1;
            philox_normals_block(rng, (uint32_t)it, (uint32_t)j, z0, z1);
            ps[2 * j] = z0;
            if (2 * j + 1 < n) ps[2 * j + 1] = z1;
          }
        }
        double kin = 0.0;
#pragma unroll
        for (int d = 0; d < n; ++d) {
          ps[d] = ps[d] / sm_sqw[d];
          kin += (ps[d] * ps[d]) * sm_w[d];
          const double qd = SLOT(SL_QP, d), gd = SLOT(SL_GP, d);
          SLOT(SL_QL, d) = qd; SLOT(SL_QR, d) = qd;
          SLOT(SL_GL, d) = gd; SLOT(SL_GR, d) = gd;
          SLOT(SL_PL, d) = ps[d]; SLOT(SL_PR, d) = ps[d];
        }
        h0 = -lpp + 0.5 * kin;
        const double lgu = injected ? log(a.b.u_inject[(size_t)it * C + chain]) : philox_accept_log_uniform(rng, (uint32_t)it);
        log_u = lgu - h0;
        n_valid = 1; n_steps = 0; sum_alpha = 0.0; depth = 0; diverged = false;
        in_transition = true;
      }
      // cond_fn (NUTS.py:243-248): NaN dot products compare false -> no U-turn
      double dl = 0.0, dr = 0.0;
#pragma unroll
      for (int d = 0; d < n; ++d) {
        const double dq = SLOT(SL_QR, d) - SLOT(SL_QL, d);
        dl += dq * SLOT(SL_PL, d);
        dr += dq * SLOT(SL_PR, d);
      }
      const bool u_turn = (dl < 0) | (dr < 0);
      if ((depth < max_depth) && !u_turn && !diverged) {
        // start the next subtree of 2^depth steps (NUTS.py:254-279)
        double u_dir;
        if (injected) {
          const size_t idx = ((size_t)it * C + chain) * max_depth + depth;
          u_dir = a.b.nuts_dir_u[idx];
          u_prop = a.b.nuts_prop_u[idx];
        } else {
          philox_nuts_uniforms(rng, (uint32_t)it, (uint32_t)depth, u_dir, u_prop);
        }
        direction = (u_dir < 0.5) ? 1 : -1;  // bernoulli(0.5) = uniform < 0.5
        const int sq = direction == -1 ? SL_QL : SL_QR;
#pragma unroll
        for (int d = 0; d < n; ++d) {
          qs[d] = SLOT(sq, d);
          ps[d] = SLOT(sq + 1, d);
          gs[d] = SLOT(sq + 2, d);
        }
        num_steps = 1 << depth;
        k = num_steps;
        se = direction * eps;  // NUTS.py:155
        hse = 0.5 * se;        // NUTS.py:107
        sa = 0.0;
        lps = NAN;
        in_subtree = true;
      } else {
        // transition finished (NUTS.py:384-396): the proposal becomes the state
        double map = sum_alpha / (double)(n_steps > 1 ? n_steps : 1);
        if (isnan(map) || isinf(map)) map = 0.65;
        acc += 1;
        sum_depth += (unsigned long long)depth;
        n_div += diverged ? 1ull : 0ull;
        const int64_t si = it - B;
        if (si >= 0 && active) {
          const size_t o = (size_t)si * C + chain;
          if (store || do_w) {
#pragma unroll
            for (int d = 0; d < n; ++d) qs[d] = SLOT(SL_QP, d);
          }
          if (store) {
            store_row<DS>(a.b.samples + o * n, qs, n);
            if (a.b.log_probs) a.b.log_probs[o] = lpp;
          }
          if (a.b.tree_depths) a.b.tree_depths[o] = depth;
          if (a.b.mean_accept_probs) a.b.mean_accept_probs[o] = map;
          if (do_w) {
            wcount += 1.0;
            double* wm = a.b.welford_mean + chain * n;
            double* w2 = a.b.welford_m2 + chain * n;
#pragma unroll
            for (int d = 0; d < n; ++d) {
              const double mean = wm[d];
              const double delta = qs[d] - mean;
              const double mean2 = mean + delta / wcount;
              wm[d] = mean2;
              w2[d] = w2[d] + delta * (qs[d] - mean2);
            }
          }
        }
        in_transition = false;
        it += 1;
        if (it == B + S) done = true;
      }
    }

    // ------------------------------------------------------------ one leapfrog step (convergent, hot)
    if (!done) {
#pragma unroll
      for (int d = 0; d < n; ++d) {
        ps[d] = ps[d] + hse * gs[d];              // NUTS.py:107
        qs[d] = qs[d] + se * (ps[d] * sm_w[d]);   // NUTS.py:109
      }
      lps = target_eval<TID, DS, true, true>(a.t, n, qs, gs, sm_tab);  // NUTS.py:111
      double kin = 0.0;
#pragma unroll
      for (int d = 0; d < n; ++d) {
        ps[d] = ps[d] + hse * gs[d];              // NUTS.py:115
        kin += (ps[d] * ps[d]) * sm_w[d];
      }
      h_new = -lps + 0.5 * kin;                   // NUTS.py:178 / :282
      sa = sa + exp(min_nan(0.0, h0 - h_new));    // NUTS.py:179-181
      k -= 1;
    }
  }
#undef SLOT

  if (active) {
    if (a.b.accept_count) a.b.accept_count[chain] = acc;
    if (a.b.accept_rate) a.b.accept_rate[chain] = (float)acc / (float)S;  // NUTS.py:512
    if (a.b.final_lp) a.b.final_lp[chain] = lpp;
    if (a.b.final_q || a.b.final_g) {
#pragma unroll
      for (int d = 0; d < n; ++d) {
        qs[d] = slots[((size_t)SL_QP * n + d) * stride];
        gs[d] = slots[((size_t)SL_GP * n + d) * stride];
      }
      if (a.b.final_q) store_row<DS>(a.b.final_q + chain * n, qs, n);
      if (a.b.final_g) store_row<DS>(a.b.final_g + chain * n, gs, n);
    }
  }
  if (a.b.counters) {
    if (!active) { acc = 0; n_grad = 0; sum_depth = 0; n_div = 0; }
    atomicAdd(a.b.counters + 0, (unsigned long long)acc);
    atomicAdd(a.b.counters + 1, n_grad);
    atomicAdd(a.b.counters + 2, sum_depth);
    atomicAdd(a.b.counters + 3, n_div);
  }
}

template <int TID, int DS, int BLOCK>
int launch_nuts_variant(const Launch& l) {
  NutsKArgs a;
  a.p = l.p;
  a.b = l.b;
  a.t = make_target_ctx(l.p, l.b.cov_inv);
  const int n = l.p.n_dim;
  const int tabn = target_table_len(TID, n);
  size_t smem_d = (size_t)((2 * n + kRngTableDoubles + tabn + 1) / 2) * 2;
  if (DS > 0) smem_d += (size_t)SL_COUNT * n * BLOCK;
  const size_t smem = smem_d * sizeof(double);
  if (DS == 0) {
    const size_t need = nuts_workspace_bytes(l.p);
    if (!l.b.workspace || l.b.workspace_bytes < need) {
      set_error("nuts: workspace of %zu bytes required for runtime D=%d (got %zu)", need, n, l.b.workspace_bytes);
      return MCMC_EWORKSPACE;
    }
  }
  auto kern = nuts_kernel<TID, DS, BLOCK>;
  if (smem > 227 * 1024) {
    set_error("nuts: %zu bytes of shared memory needed exceed 227 KB", smem);
    return MCMC_ENOTSUP;
  }
  if (int rc = cuda_fail(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem),
                         "cudaFuncSetAttribute(nuts)"))
    return rc;
  const int64_t grid = (l.p.n_chains + BLOCK - 1) / BLOCK;
  if (grid == 0) return MCMC_OK;
  kern<<<(unsigned)grid, BLOCK, smem, l.stream>>>(a);
  return cuda_fail(cudaGetLastError(), "nuts_kernel launch");
}

}  // namespace mcmc
