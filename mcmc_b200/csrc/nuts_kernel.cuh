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
// Tree ends (left/right q,p,g) and the proposal (q,g) are 8 D-vectors per chain.  They live in a
// global workspace, one contiguous [slot][d][lane] tile per warp (coalesced, constant offsets,
// L2-resident): in shared memory they capped the SM at 5 warps, in L2 the kernel is limited by
// registers only (8 warps per SM).  The end integrated last stays in registers and is written back
// only when the tree turns round, so a subtree boundary costs one read of the opposite end.
//
// Native random stream: the momenta and the slice uniform of every (chain, iteration) are drawn by a
// separate, fully convergent kernel (nuts_rng_fill_kernel, abi.cu) into the workspace with the SAME
// Philox counters the in-kernel draw would use, and read back here at the start of a transition.
// Inside this kernel the draw was the expensive divergent branch: some lane of a warp starts a
// transition in almost every trip, so all 32 lanes paid ~1700 instructions of Philox + Box-Muller per
// leapfrog step.  Long runs are cut into segments of iterations so the buffer stays bounded; the
// state travels between the launches of a run through the workspace (bit-identical results).
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
  int64_t it_begin, it_end;   // absolute iterations [it_begin, it_end) of this launch (segment)
  double* slots;              // [SL_COUNT][D][cpad]
  size_t cpad;
  double* carry_lp;           // [cpad] log-prob of the current state between segments
  int* carry_acc;             // [cpad]
  const double* z;            // momenta draws: z_inject (+ it*C*D) or the filled segment buffer
  const double* logu;         // native stream: log of the slice uniform per (segment iteration, chain); else NULL
  int64_t z_it0;              // iteration that row 0 of z / logu corresponds to
};

enum { SL_QL = 0, SL_PL, SL_GL, SL_QR, SL_PR, SL_GR, SL_QP, SL_GP, SL_COUNT };

template <int TID, int DS, int BLOCK>
__global__ void __launch_bounds__(BLOCK) nuts_kernel(const NutsKArgs a) {
  constexpr int N = DS > 0 ? DS : MCMC_DYN_MAXD;
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
  double* sm_tab = smem + 2 * n;
  // Slots of one warp are a contiguous [slot][d][lane] tile: slot(s, d) = wbase[(s*n + d)*32], i.e. every
  // access is one coalesced 256-byte request at a compile-time offset from a per-lane base pointer.
  double* wbase = a.slots + (size_t)((chain0 + tid) >> 5) * ((size_t)SL_COUNT * n * 32) + (tid & 31);
#define SLOT(s, d) wbase[((s) * n + (d)) * 32]

  for (int d = tid; d < n; d += BLOCK) {
    const double w = a.b.inv_mass ? a.b.inv_mass[d] : 1.0;
    sm_w[d] = w;
#if MCMC_IS_FAITHFUL
    sm_sqw[d] = sqrt(w);        // p0 = z / sqrt(w), NUTS.py:219
#else
    sm_sqw[d] = 1.0 / sqrt(w);
#endif
  }
  target_table_fill<TID>(a.t, n, sm_tab);
  __syncthreads();

  RngCtx rng;
  {
    const uint64_t gid = (uint64_t)(a.p.chain_offset + chain);
    rng.k0 = a.p.seed[0]; rng.k1 = a.p.seed[1];
    rng.gid_lo = (uint32_t)gid; rng.gid_hi = (uint32_t)(gid >> 32);
    rng.logtab = nullptr;   // only the integer-valued direction / proposal uniforms are drawn in this kernel
  }

  // Working point = the tree end that was integrated last (direction reg_dir).  It stays in registers; its slot
  // is written only when the tree turns round, the opposite end is always current in its slot.
  double qs[N], ps[N], gs[N];
  double lps = 0.0;

  // nuts_init (NUTS.py:73-86): the current state lives in the proposal slots; a later segment of the
  // same run finds it there and picks up the carried log-prob / accept counter
  double lpp;
  int acc = 0;
  if (a.it_begin == 0) {
    const double* src = a.b.init_q + chain * n;
#pragma unroll
    for (int d = 0; d < n; ++d) qs[d] = src[d];
    lpp = target_eval<TID, DS, true, true>(a.t, n, qs, gs, sm_tab);
#pragma unroll
    for (int d = 0; d < n; ++d) {
      SLOT(SL_QP, d) = qs[d];
      SLOT(SL_GP, d) = gs[d];
    }
  } else {
    lpp = a.carry_lp[chain0 + tid];
    acc = a.carry_acc[chain0 + tid];
  }

  // per-transition scalars
  double h0 = 0.0, log_u = 0.0, sum_alpha = 0.0, sa = 0.0, h_new = 0.0, u_prop = 0.0;
  double se = 0.0, hse = 0.0;
  int n_valid = 1, n_steps = 0, depth = 0, direction = 1, reg_dir = 1, k = 0, num_steps = 0;
  bool diverged = false, in_subtree = false, in_transition = false, done = a.it_begin >= a.it_end;
  int64_t it = a.it_begin;
  double wcount = a.b.welford_count + (double)(a.it_begin > B ? a.it_begin - B : 0);
  unsigned long long n_grad = 0, sum_depth = 0, n_div = 0;

  while (!__all_sync(0xffffffffu, done)) {
    // ------------------------------------------------------------ bookkeeping (divergent, kept short)
    while (k == 0 && !done) {
      bool u_turn = false;
      if (in_subtree) {
        // subtree endpoint: validity, divergence, proposal swap (NUTS.py:281-378); the new end stays in registers
        const bool in_slice = log_u <= -h_new;
        const bool is_div = (h_new - h0) > delta_max;
        const bool is_valid = in_slice && !is_div;
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
        // cond_fn (NUTS.py:243-248) on (q_right - q_left) . p_left / p_right; NaN compares false -> no U-turn
        const double* oth = wbase + (reg_dir > 0 ? SL_QL : SL_QR) * n * 32;
        // with e = q_work - q_other:  working end on the right: (dl, dr) = (e.p_other, e.p_work);
        // on the left q_right - q_left = -e: (dl, dr) = (-(e.p_work), -(e.p_other)) — negation commutes with
        // every rounding, so this is the reference's value bit for bit
        double eo = 0.0, ew = 0.0;
#pragma unroll
        for (int d = 0; d < n; ++d) {
          const double e = qs[d] - oth[d * 32];
          eo += e * oth[(n + d) * 32];
          ew += e * ps[d];
        }
        const double dl = reg_dir > 0 ? eo : -ew, dr = reg_dir > 0 ? ew : -eo;
        u_turn = (dl < 0) | (dr < 0);
      }
      if (!in_transition) {
        // new transition: momentum, energy, slice (NUTS.py:217-241).  Both tree ends are the current state
        // (q_prop, p0, g_prop): it is loaded into the working registers; the end opposite to the first
        // direction is written once that direction is known (below).  q_left == q_right -> no U-turn yet.
        if (it == B) acc = 0;  // NUTS.py:494-500
        load_row<DS>(ps, a.z + ((size_t)(it - a.z_it0) * C + chain) * n, n);
        double kin = 0.0;
#pragma unroll
        for (int d = 0; d < n; ++d) {
#if MCMC_IS_FAITHFUL
          ps[d] = ps[d] / sm_sqw[d];
#else
          ps[d] = ps[d] * sm_sqw[d];
#endif
          kin += (ps[d] * ps[d]) * sm_w[d];
          qs[d] = SLOT(SL_QP, d);
          gs[d] = SLOT(SL_GP, d);
        }
        h0 = -lpp + 0.5 * kin;
        const double lgu = a.logu ? a.logu[(size_t)(it - a.z_it0) * C + chain] : log(a.b.u_inject[(size_t)it * C + chain]);
        log_u = lgu - h0;
        n_valid = 1; n_steps = 0; sum_alpha = 0.0; depth = 0; diverged = false;
        in_transition = true;
        reg_dir = 0;   // registers hold BOTH ends
        u_turn = false;
      }
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
        if (reg_dir == 0) {
          // first subtree of the transition: the initial state becomes the end opposite to `direction`
          double* oth = wbase + (direction > 0 ? SL_QL : SL_QR) * n * 32;
#pragma unroll
          for (int d = 0; d < n; ++d) {
            oth[d * 32] = qs[d];
            oth[(n + d) * 32] = ps[d];
            oth[(2 * n + d) * 32] = gs[d];
          }
        } else if (direction != reg_dir) {
          // the tree turns round: park the working end, fetch the opposite one
          double* mine = wbase + (reg_dir > 0 ? SL_QR : SL_QL) * n * 32;
          const double* oth = wbase + (direction > 0 ? SL_QR : SL_QL) * n * 32;
#pragma unroll
          for (int d = 0; d < n; ++d) {
            mine[d * 32] = qs[d];
            mine[(n + d) * 32] = ps[d];
            mine[(2 * n + d) * 32] = gs[d];
            qs[d] = oth[d * 32];
            ps[d] = oth[(n + d) * 32];
            gs[d] = oth[(2 * n + d) * 32];
          }
        }
        reg_dir = direction;
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
        if (it == a.it_end) done = true;
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
      sa = sa + exp_nonpos(min_nan(0.0, h0 - h_new));    // NUTS.py:179-181
      k -= 1;
    }
  }

  if (a.it_end < B + S) {
    a.carry_lp[chain0 + tid] = lpp;
    a.carry_acc[chain0 + tid] = acc;
  } else if (active) {
    if (a.b.accept_count) a.b.accept_count[chain] = acc;
    if (a.b.accept_rate) a.b.accept_rate[chain] = (float)acc / (float)S;  // NUTS.py:512
    if (a.b.final_lp) a.b.final_lp[chain] = lpp;
    if (a.b.final_q || a.b.final_g) {
#pragma unroll
      for (int d = 0; d < n; ++d) {
        qs[d] = SLOT(SL_QP, d);
        gs[d] = SLOT(SL_GP, d);
      }
      if (a.b.final_q) store_row<DS>(a.b.final_q + chain * n, qs, n);
      if (a.b.final_g) store_row<DS>(a.b.final_g + chain * n, gs, n);
    }
  }
#undef SLOT
  if (a.b.counters) {
    if (!active) { acc = 0; n_grad = 0; sum_depth = 0; n_div = 0; }
    if (a.it_end == B + S) atomicAdd(a.b.counters + 0, (unsigned long long)acc);
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
  const size_t smem = (size_t)(2 * n + tabn) * sizeof(double);
  const NutsWs ws = nuts_ws_layout(l.p);
  if (!l.b.workspace || l.b.workspace_bytes < ws.total) {
    set_error("nuts: workspace of %zu bytes required (got %zu)", ws.total, l.b.workspace_bytes);
    return MCMC_EWORKSPACE;
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
  char* w = reinterpret_cast<char*>(l.b.workspace);
  a.slots = reinterpret_cast<double*>(w + ws.slots_off);
  a.cpad = ws.cpad;
  a.carry_lp = reinterpret_cast<double*>(w + ws.lp_off);
  a.carry_acc = reinterpret_cast<int*>(w + ws.acc_off);
  const int64_t T = l.p.burn_in + l.p.num_samples;
  const bool native = l.p.rng_mode == MCMC_RNG_PHILOX;
  // segments of at most ws.tseg iterations (one segment unless the draw buffer would exceed its cap)
  int64_t it0 = 0;
  do {
    const int64_t it1 = (native && it0 + ws.tseg < T) ? it0 + ws.tseg : T;
    a.it_begin = it0;
    a.it_end = it1;
    if (native) {
      double* zb = reinterpret_cast<double*>(w + ws.z_off);
      double* lb = reinterpret_cast<double*>(w + ws.lu_off);
      if (int rc = launch_rng_fill(l.p, it0, it1, zb, lb, l.stream)) return rc;
      a.z = zb;
      a.logu = lb;
      a.z_it0 = it0;
    } else {
      a.z = l.b.z_inject;
      a.logu = nullptr;
      a.z_it0 = 0;
    }
    kern<<<(unsigned)grid, BLOCK, smem, l.stream>>>(a);
    if (int rc = cuda_fail(cudaGetLastError(), "nuts_kernel launch")) return rc;
    it0 = it1;
  } while (it0 < T);
  return MCMC_OK;
}

}  // namespace mcmc
