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
// Memory per chain in flight: the working tree end (q, p, g: the end integrated last) stays in registers; the
// opposite end (q, p, g) and the proposal (q, g) are five D-vectors in a [slot][d][lane] tile private to the LANE — in
// shared memory for the compiled dimensions (25.6 KB per warp at D=20: eight one-warp blocks per SM, the register
// limit), in an L2-resident workspace tile for the runtime-D variant.  A subtree boundary (U-turn test, turn-round,
// proposal swap) therefore never leaves the SM.
//
// Work distribution: a chain is not tied to a lane.  The run is cut into chunks of iterations; items (chunk, chain)
// are claimed chunk-major from an atomic counter by whichever lane runs out of work, the chain's state (q, g, lp,
// accept counter) travels between items through a per-chain home row in the workspace, and an item whose
// predecessor is still running parks its lane in a WAIT state that is polled once per loop trip (the other lanes
// keep stepping).  With a fixed lane->chain map a warp ran until its slowest chain finished: on the 20-D
// rosenbrock 30% and on Neal's funnel 75% of all lane-steps were idle tail (tools/nuts_tail_probe.py).  Results do
// not depend on who runs which item: every random number is keyed on (global chain, absolute iteration).
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
  double* slots;              // runtime-D variant only: [resident warp][SL_COUNT][D][32] lane tiles
  double* home;               // [C][2D]: (q, g) of every chain's current state between items / launches
  double* home_lp;            // [C]
  int* home_acc;              // [C]
  unsigned int* queue;        // item counter of this launch
  unsigned int* flags;        // [C] chunks of this launch completed, per chain
  int chunk_iters, n_chunks;
  const double* z;            // momenta draws: z_inject (+ it*C*D) or the filled segment buffer
  const double* logu;         // native stream: log of the slice uniform per (segment iteration, chain); else NULL
  int64_t z_it0;              // iteration that row 0 of z / logu corresponds to
};

// other tree end (q, p, g) and proposal (q, g)
enum { SL_QO = 0, SL_PO, SL_GO, SL_QP, SL_GP, SL_COUNT };

#ifndef MCMC_NUTS_MINBLOCKS
#define MCMC_NUTS_MINBLOCKS 1
#endif
constexpr int kNutsMinBlocks = MCMC_NUTS_MINBLOCKS;
#ifndef MCMC_NUTS_BK_TRIPS
#define MCMC_NUTS_BK_TRIPS 8
#endif
#ifndef MCMC_NUTS_BK_LANES
#define MCMC_NUTS_BK_LANES 16
#endif
constexpr int kNutsBkTrips = MCMC_NUTS_BK_TRIPS, kNutsBkLanes = MCMC_NUTS_BK_LANES;
// Steps per loop trip.  The loop control of a trip (three votes, the convergence barrier of __all_sync, ~30 kernel-parameter
// reloads) costs a lone lane ~250 cycles — a quarter of its step time (profiles/nuts_experiments_r02.txt section 7).  So a
// trip runs r = min(steps left over all stepping lanes, kNutsInner) steps: every stepping lane has at least r to do, so
// they stay in lockstep and nobody idles inside; lanes waiting for bookkeeping wait r steps instead of one (they already
// wait up to kNutsBkTrips).  Purely a schedule: results are unchanged.
#ifndef MCMC_NUTS_INNER
#define MCMC_NUTS_INNER 8
#endif
constexpr int kNutsInner = MCMC_NUTS_INNER;

template <int TID, int DS, int BLOCK>
__global__ void __launch_bounds__(BLOCK, DS > 0 ? kNutsMinBlocks : 1) nuts_kernel(const NutsKArgs a) {
  static_assert(BLOCK == 32, "one warp per block: the lane tiles are sized per block");
  constexpr int N = DS > 0 ? DS : MCMC_DYN_MAXD;
  constexpr bool kSmemSlots = DS > 0;
  const int n = DS > 0 ? DS : a.p.n_dim;
  const int64_t C = a.p.n_chains, S = a.p.num_samples, B = a.p.burn_in;
  const int tid = threadIdx.x;
  const bool store = !(a.p.flags & MCMC_FLAG_NO_SAMPLES) && a.b.samples != nullptr;
  const bool do_w = (a.p.flags & MCMC_FLAG_ACCUM_WELFORD) && a.b.welford_mean && a.b.welford_m2;
  const bool injected = a.p.rng_mode == MCMC_RNG_INJECTED;
  const int max_depth = a.p.max_tree_depth;
  const double eps = a.b.step_size_dev ? __ldg(a.b.step_size_dev) : a.p.step_size;
  const double delta_max = a.p.delta_max;

  extern __shared__ __align__(16) double smem[];
  double* sm_w = smem;
  double* sm_sqw = smem + n;
  double* sm_epsw = smem + 2 * n;   // eps * w[d]: the drift of the fast build is one DFMA per dimension
  double* sm_tab = smem + 3 * n;
  // lane tile: slot(s, d) = wbase[(s*n + d)*32] — conflict-free in shared memory, one coalesced 256-byte request in L2
  double* wbase;
  if (kSmemSlots) wbase = smem + ((3 * n + target_table_len(TID, n) + 1) / 2) * 2 + tid;
  else wbase = a.slots + (size_t)blockIdx.x * ((size_t)SL_COUNT * n * 32) + tid;
#define SLOT(s, d) wbase[((s) * n + (d)) * 32]

  for (int d = tid; d < n; d += BLOCK) {
    const double w = a.b.inv_mass ? a.b.inv_mass[d] : 1.0;
    sm_w[d] = w;
    sm_epsw[d] = eps * w;
#if MCMC_IS_FAITHFUL
    sm_sqw[d] = sqrt(w);        // p0 = z / sqrt(w), NUTS.py:219
#else
    sm_sqw[d] = 1.0 / sqrt(w);
#endif
  }
  target_table_fill<TID>(a.t, n, sm_tab);
  __syncthreads();

  RngCtx rng;
  rng.k0 = a.p.seed[0]; rng.k1 = a.p.seed[1];
  rng.gid_lo = rng.gid_hi = 0;
  rng.logtab = nullptr;   // only the integer-valued direction / proposal uniforms are drawn in this kernel

  // Working point = the tree end that was integrated last (direction reg_dir).  It stays in registers; the
  // opposite end is always current in the O slots and the two are swapped when the tree turns round.
  // The working momentum is held SIGN-FOLDED: ps = reg_dir * p.  Integrating backwards (step -eps, NUTS.py:155) is
  // integrating -p forwards, and negation commutes with every rounding, so the step below is direction-free — the
  // same kick and drift constants for every lane, bit-identical values.  The opposite end's momentum in the PO slot is
  // folded the same way with ITS outward direction (-reg_dir), so a turn-round is a plain exchange and the U-turn test
  // reads both dot products with a fixed sign.  A lone warp is bound by the issue rate of the fp64 pipe (one warp-instruction per two cycles), so
  // what a step costs is its fp64 instruction count: the fast build's drift is fma(ps, eps*w, qs), one instead of two.
  double qs[N], ps[N], gs[N];
  double lps = 0.0;
  double lpp = 0.0;
  int acc = 0;

  // per-transition scalars
  double h0 = 0.0, log_u = 0.0, sum_alpha = 0.0, sa = 0.0, h_new = 0.0, u_prop = 0.0;
  // The Metropolis weight exp(min(0, h0 - h)) of a step (NUTS.py:179-181) is only ever summed.  Its argument is parked in
  // `pend` and exponentiated at the START of the lane's next step, where the ~25-deep dependent chain hides under the
  // kicks and drifts instead of ending the step (a lone lane waited ~15% of its time there); the subtree-end bookkeeping
  // adds the last one.  Same additions in the same order (exp(-inf) = +0 where nothing is parked): bit-identical.
  double pend = -INFINITY;
  const double heps = 0.5 * eps;   // NUTS.py:107 with the direction folded into ps
  // (steps of the subtree under way = 1 << depth and steps so far = (1 << depth) - 1 are recomputed where needed: every
  // register this loop does not hold is one ptxas can schedule the step with)
  int n_valid = 1, depth = 0, direction = 1, reg_dir = 1, k = 0;
  bool diverged = false, in_subtree = false, in_transition = false;
  // per-item state
  bool done = false, have_item = false, waiting = false;
  int64_t it = 0, it_item_end = 0, chain = 0;
  unsigned int chunk = 0;
  unsigned int n_grad = 0, sum_depth = 0, n_div = 0, acc_total = 0;   // per-lane tallies, flushed before they can wrap
  const unsigned int n_items = (unsigned int)C * (unsigned int)a.n_chunks;

  int since_bk = 0;
  while (!__all_sync(0xffffffffu, done)) {
    // ------------------------------------------------------------ bookkeeping (divergent, latency-bound)
    // A trip that runs it costs several times a leapfrog step in dependent-latency, whether one lane needs it or
    // all 32 do.  So it is batched: lanes whose subtree ended idle until kNutsBkLanes of them are waiting, the
    // oldest has waited kNutsBkTrips trips, or nobody has a step left.  Purely a schedule: results are unchanged.
    const unsigned int needy = __ballot_sync(0xffffffffu, k == 0 && !done);
    const unsigned int busy = __ballot_sync(0xffffffffu, k > 0);
    since_bk += 1;
    const bool run_bk = needy && (busy == 0u || __popc(needy) >= kNutsBkLanes || since_bk >= kNutsBkTrips);
    if (run_bk) since_bk = 0;
    while (run_bk && k == 0 && !done) {
      if (!have_item) {
        if (!waiting) {
          const unsigned int id = atomicAdd(a.queue, 1u);
          if (id >= n_items) { done = true; break; }
          chunk = id / (unsigned int)C;
          chain = (int64_t)(id - chunk * (unsigned int)C);
          it = a.it_begin + (int64_t)chunk * a.chunk_iters;
          it_item_end = (it + a.chunk_iters < a.it_end) ? it + a.chunk_iters : a.it_end;
          waiting = chunk > 0;
        }
        if (waiting) {
          // the chain's previous chunk is still running on another lane: look again next trip
          if (ld_acquire_u32(a.flags + chain) < chunk) break;
          waiting = false;
        }
        const uint64_t gid = (uint64_t)(a.p.chain_offset + chain);
        rng.gid_lo = (uint32_t)gid; rng.gid_hi = (uint32_t)(gid >> 32);
        if (it == 0) {
          // nuts_init (NUTS.py:73-86): value and gradient at the initial position
          const double* src = a.b.init_q + chain * n;
#pragma unroll
          for (int d = 0; d < n; ++d) qs[d] = src[d];
          lpp = target_eval<TID, DS, true, true>(a.t, n, qs, gs, sm_tab);
          acc = 0;
#pragma unroll
          for (int d = 0; d < n; ++d) {
            SLOT(SL_QP, d) = qs[d];
            SLOT(SL_GP, d) = gs[d];
          }
        } else {
          // state left by the previous item of this chain (another lane, another SM or the previous launch)
          const double* hq = a.home + (size_t)chain * 2 * n;
#pragma unroll
          for (int d = 0; d < n; ++d) {
            SLOT(SL_QP, d) = __ldcg(hq + d);
            SLOT(SL_GP, d) = __ldcg(hq + n + d);
          }
          lpp = __ldcg(a.home_lp + chain);
          acc = __ldcg(a.home_acc + chain);
        }
        have_item = true;
        in_transition = false;
        in_subtree = false;
      }
      bool u_turn = false;
      if (in_subtree) {
        // subtree endpoint: validity, divergence, proposal swap (NUTS.py:281-378); the new end stays in registers
        const bool in_slice = log_u <= -h_new;
        const bool is_div = (h_new - h0) > delta_max;
        const bool is_valid = in_slice && !is_div;
        const int num_steps = 1 << depth;   // the subtree that just ended
        const int n_valid_new = is_valid ? num_steps : 0;
        sa = sa + exp_nonpos(pend);   // the weight of the subtree's last step
        pend = -INFINITY;
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
        n_grad += (unsigned int)num_steps;
        if (n_grad > 0x7f000000u || acc_total > 0x7f000000u || sum_depth > 0x7f000000u) {
          if (a.b.counters) {
            atomicAdd(a.b.counters + 0, (unsigned long long)acc_total);
            atomicAdd(a.b.counters + 1, (unsigned long long)n_grad);
            atomicAdd(a.b.counters + 2, (unsigned long long)sum_depth);
          }
          acc_total = 0; n_grad = 0; sum_depth = 0;
        }
        depth += 1;
        diverged = diverged || is_div;
        in_subtree = false;
        // cond_fn (NUTS.py:243-248) on (q_right - q_left) . p_left / p_right; NaN compares false -> no U-turn.
        // With e = q_work - q_other:  working end on the right: (dl, dr) = (e.p_other, e.p_work);
        // on the left q_right - q_left = -e: (dl, dr) = (-(e.p_work), -(e.p_other)) — negation commutes with
        // every rounding, so this is the reference's value bit for bit
        double eop[kSumWays] = {0.0}, ewp[kSumWays] = {0.0};
#pragma unroll
        for (int d = 0; d < n; ++d) {
          const double e = qs[d] - SLOT(SL_QO, d);
          eop[d % kSumWays] += e * SLOT(SL_PO, d);
          ewp[d % kSumWays] += e * ps[d];
        }
        // Both momenta are folded with their end's OUTWARD direction: ps = reg_dir * p_work, PO slot = -reg_dir * p_other.
        // So e.p_work = reg_dir * ew and e.p_other = -reg_dir * eo, and in either orientation the pair (dl, dr) is
        // {-eo, ew}: right end (e.p_other, e.p_work) = (-eo, ew); left end (-(e.p_work), -(e.p_other)) = (ew, -eo)
        const double eo = sum_ways(eop), ew = sum_ways(ewp);
        u_turn = (-eo < 0) | (ew < 0);
      }
      if (!in_transition) {
        // new transition: momentum, energy, slice (NUTS.py:217-241).  Both tree ends are the current state
        // (q_prop, p0, g_prop): it is loaded into the working registers and copied to the O slots once the
        // first direction is known (below).  q_left == q_right -> no U-turn yet.
        if (it == B) acc = 0;  // NUTS.py:494-500
        const double* zrow = a.z + ((size_t)(it - a.z_it0) * C + chain) * n;
        load_row<DS>(ps, zrow, n);
        if (it + 1 < it_item_end) {
          // the next transition's draws (one row further in the segment buffer) start their way from HBM to L2 now
          const char* nx = reinterpret_cast<const char*>(zrow + (size_t)C * n);
          for (int o = 0; o < n * 8; o += 128) asm volatile("prefetch.global.L2 [%0];" ::"l"(nx + o));
        }
        double kinp[kSumWays] = {0.0};
#pragma unroll
        for (int d = 0; d < n; ++d) {
#if MCMC_IS_FAITHFUL
          ps[d] = ps[d] / sm_sqw[d];
#else
          ps[d] = ps[d] * sm_sqw[d];
#endif
          kinp[d % kSumWays] += (ps[d] * ps[d]) * sm_w[d];
          qs[d] = SLOT(SL_QP, d);
          gs[d] = SLOT(SL_GP, d);
        }
        h0 = -lpp + 0.5 * sum_ways(kinp);
        const double lgu = a.logu ? a.logu[(size_t)(it - a.z_it0) * C + chain] : log(a.b.u_inject[(size_t)it * C + chain]);
        log_u = lgu - h0;
        n_valid = 1; sum_alpha = 0.0; depth = 0; diverged = false;
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
          // first subtree of the transition: the initial state (true momentum p0 in ps) is both ends.  Each end keeps
          // its momentum folded with its own outward direction: working copy direction * p0, other end -direction * p0
#pragma unroll
          for (int d = 0; d < n; ++d) {
            const double pw = direction < 0 ? -ps[d] : ps[d];
            SLOT(SL_QO, d) = qs[d];
            SLOT(SL_PO, d) = -pw;
            SLOT(SL_GO, d) = gs[d];
            ps[d] = pw;
          }
        } else if (direction != reg_dir) {
          // the tree turns round: the working end and the opposite end change places — a plain exchange, because each
          // momentum is already folded with the direction its end is integrated in
#pragma unroll
          for (int d = 0; d < n; ++d) {
            const double tq = SLOT(SL_QO, d), tp = SLOT(SL_PO, d), tg = SLOT(SL_GO, d);
            SLOT(SL_QO, d) = qs[d];
            SLOT(SL_PO, d) = ps[d];
            SLOT(SL_GO, d) = gs[d];
            qs[d] = tq; ps[d] = tp; gs[d] = tg;
          }
        }
        reg_dir = direction;
        k = 1 << depth;
        sa = 0.0;
        lps = NAN;
        in_subtree = true;
      } else {
        // transition finished (NUTS.py:384-396): the proposal becomes the state
        const int n_steps = (1 << depth) - 1;   // 1 + 2 + ... over the `depth` subtrees of this transition
        double map = sum_alpha / (double)(n_steps > 1 ? n_steps : 1);
        if (isnan(map) || isinf(map)) map = 0.65;
        acc += 1;
        sum_depth += (unsigned int)depth;
        n_div += diverged ? 1u : 0u;
        const int64_t si = it - B;
        if (si >= 0) {
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
          if (do_w) {  // tuning/welford.py:42-59; L2 accesses: the previous item of the chain may have run on another SM
            const double wcount = a.b.welford_count + (double)(it - B + 1);   // samples folded so far, this one included
            double* wm = a.b.welford_mean + chain * n;
            double* w2 = a.b.welford_m2 + chain * n;
#if !MCMC_IS_FAITHFUL
            const double rcount = 1.0 / wcount;
#endif
#pragma unroll
            for (int d = 0; d < n; ++d) {
              const double mean = __ldcg(wm + d);
              const double delta = qs[d] - mean;
#if MCMC_IS_FAITHFUL
              const double mean2 = mean + delta / wcount;
#else
              const double mean2 = fma(delta, rcount, mean);
#endif
              __stcg(wm + d, mean2);
              __stcg(w2 + d, __ldcg(w2 + d) + delta * (qs[d] - mean2));
            }
          }
        }
        in_transition = false;
        it += 1;
        if (it == it_item_end) {
          // ---- item finished: the chain's state goes home (or, after the last iteration, to the outputs)
          if (it == B + S) {
            if (S == 0) acc = 0;  // reset after burn-in even when nothing is sampled (NUTS.py:494-500)
            acc_total += (unsigned int)acc;
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
          } else {
            double* hq = a.home + (size_t)chain * 2 * n;
#pragma unroll
            for (int d = 0; d < n; ++d) {
              __stcg(hq + d, SLOT(SL_QP, d));
              __stcg(hq + n + d, SLOT(SL_GP, d));
            }
            __stcg(a.home_lp + chain, lpp);
            __stcg(a.home_acc + chain, acc);
            __threadfence();
            st_release_u32(a.flags + chain, chunk + 1u);
          }
          have_item = false;
        }
      }
    }

    // ------------------------------------------------------------ leapfrog steps (convergent, hot)
    int n_inner = 1;
    if (kNutsInner > 1) {
      n_inner = __reduce_min_sync(0xffffffffu, k > 0 ? k : 0x7fffffff);
      n_inner = n_inner < kNutsInner ? n_inner : kNutsInner;
      since_bk += n_inner - 1;
    }
    if (k > 0) {
#pragma unroll 1
     for (int inner = 0; inner < n_inner; ++inner) {
      sa = sa + exp_nonpos(pend);                 // the previous step's weight (NUTS.py:179-181), see `pend`
#pragma unroll
      for (int d = 0; d < n; ++d) {
        ps[d] = ps[d] + heps * gs[d];             // NUTS.py:107 (ps = direction * p, see above)
#if MCMC_IS_FAITHFUL
        qs[d] = qs[d] + eps * (ps[d] * sm_w[d]);  // NUTS.py:109
#else
        qs[d] = fma(ps[d], sm_epsw[d], qs[d]);
#endif
      }
      lps = target_eval<TID, DS, true, true>(a.t, n, qs, gs, sm_tab);  // NUTS.py:111
      double kinp[kSumWays] = {0.0};
#pragma unroll
      for (int d = 0; d < n; ++d) {
        ps[d] = ps[d] + heps * gs[d];             // NUTS.py:115
        kinp[d % kSumWays] += (ps[d] * ps[d]) * sm_w[d];
      }
      h_new = -lps + 0.5 * sum_ways(kinp);        // NUTS.py:178 / :282
      pend = min_nan(0.0, h0 - h_new);
      k -= 1;
     }
    } else if (__activemask() == 0xffffffffu) {
      __nanosleep(100);  // every lane of the warp is waiting for another warp's item (or finished)
    }
  }
#undef SLOT
  if (a.b.counters) {
    if (acc_total) atomicAdd(a.b.counters + 0, (unsigned long long)acc_total);
    if (n_grad) atomicAdd(a.b.counters + 1, (unsigned long long)n_grad);
    if (sum_depth) atomicAdd(a.b.counters + 2, (unsigned long long)sum_depth);
    if (n_div) atomicAdd(a.b.counters + 3, (unsigned long long)n_div);
  }
}

// burn_in + num_samples == 0: nuts_run returns the nuts_init state (NUTS.py:73-86,479-514) — no item ever runs, so the
// outputs are produced here (accept_rate = 0 / 0 in float32, as count / num_samples gives in jax)
static __global__ void nuts_empty_run_kernel(int64_t C, int32_t* __restrict__ accept_count, float* __restrict__ accept_rate) {
  const int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  if (accept_count) accept_count[c] = 0;
  if (accept_rate) accept_rate[c] = __int_as_float(0x7fc00000);   // 0 / 0
}

template <int TID, int DS, int BLOCK>
int launch_nuts_variant(const Launch& l) {
  NutsKArgs a;
  a.p = l.p;
  a.b = l.b;
  a.t = make_target_ctx(l.p, l.b.cov_inv);
  const int n = l.p.n_dim;
  const int tabn = target_table_len(TID, n);
  const size_t tile = (size_t)SL_COUNT * n * 32;   // doubles per warp
  const size_t smem = ((size_t)((3 * n + tabn + 1) / 2) * 2 + (DS > 0 ? tile : 0)) * sizeof(double);
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
  const int64_t C = l.p.n_chains;
  if (C == 0) return MCMC_OK;
  if (C >= ((int64_t)1 << 31)) { set_error("nuts: more than 2^31 chains"); return MCMC_ENOTSUP; }
  // persistent one-warp blocks: as many as fit (at most ws.max_warps, which sized the runtime-D lane tiles)
  int per_sm = 0;
  if (int rc = cuda_fail(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, BLOCK, smem), "occupancy(nuts)"))
    return rc;
  int64_t grid = (int64_t)per_sm * device_sm_count();
  if (grid > ws.max_warps) grid = ws.max_warps;
  if (grid > (C + 31) / 32) grid = (C + 31) / 32;
  if (grid < 1) grid = 1;
  char* w = reinterpret_cast<char*>(l.b.workspace);
  a.slots = reinterpret_cast<double*>(w + ws.slots_off);
  a.home = reinterpret_cast<double*>(w + ws.home_off);
  a.home_lp = reinterpret_cast<double*>(w + ws.lp_off);
  a.home_acc = reinterpret_cast<int*>(w + ws.acc_off);
  a.queue = reinterpret_cast<unsigned int*>(w + ws.queue_off);
  a.flags = a.queue + 64;
  const int64_t T = l.p.burn_in + l.p.num_samples;
  if (T == 0) {
    if (l.b.final_lp || l.b.final_g)
      if (int rc = launch_target_eval<TID>(l.p, l.b.cov_inv, l.b.init_q, C, l.b.final_lp, l.b.final_g, l.stream)) return rc;
    if (l.b.final_q)
      if (int rc = cuda_fail(cudaMemcpyAsync(l.b.final_q, l.b.init_q, (size_t)C * n * sizeof(double),
                                             cudaMemcpyDeviceToDevice, l.stream), "memcpy(nuts final_q)"))
        return rc;
    nuts_empty_run_kernel<<<(unsigned)((C + 255) / 256), 256, 0, l.stream>>>(C, l.b.accept_count, l.b.accept_rate);
    return cuda_fail(cudaGetLastError(), "nuts_empty_run launch");
  }
  const bool native = l.p.rng_mode == MCMC_RNG_PHILOX;
  const char* chunk_env = getenv("MCMC_NUTS_CHUNK");  // test hook / tuning knob: iterations per item
  // segments of at most ws.tseg iterations (one segment unless the draw buffer would exceed its cap)
  int64_t it0 = 0;
  do {
    const int64_t it1 = (native && it0 + ws.tseg < T) ? it0 + ws.tseg : T;
    a.it_begin = it0;
    a.it_end = it1;
    // items: with more chains than resident lanes, ~6 chunks per chain (measured optimum at 65,536 chains x 288
    // iterations: 3..6 within 2%; more chunks add WAIT polling, fewer leave a longer tail); otherwise every chain
    // has a lane of its own and one item per chain is all there is to do
    int64_t ci = it1 - it0;
    if (chunk_env && atoll(chunk_env) > 0) ci = atoll(chunk_env);
    else if (C > grid * 32 && !no_spin()) ci = (it1 - it0 + 5) / 6;
    if (ci < 1) ci = 1;
    int64_t nch = (it1 - it0 + ci - 1) / ci;
    if (nch * C >= ((int64_t)1 << 32) - (grid * 32 + 1)) {  // the queue counter is 32-bit: fall back to longer items
      nch = (((int64_t)1 << 32) - (grid * 32 + 2)) / C;
      if (nch < 1) nch = 1;
      ci = (it1 - it0 + nch - 1) / nch;
      nch = (it1 - it0 + ci - 1) / ci;
    }
    a.chunk_iters = (int)ci;
    a.n_chunks = (int)nch;
    if (int rc = cuda_fail(cudaMemsetAsync(a.queue, 0, (64 + (size_t)C) * sizeof(unsigned int), l.stream), "memset(nuts queue)"))
      return rc;
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
