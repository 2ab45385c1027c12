// nuts_quad_kernel.cuh — nuts_run with a QUAD of lanes per chain (D = 20, fast build).
//
// Same algorithm, same Philox counters, same work distribution as nuts_kernel.cuh (which it replaces for the compiled
// 20-D case; reference: samples/NUTS.py:202-396, :399-442, :479-514) — what changes is how a chain is mapped onto the
// machine.  nuts_kernel.cuh gives a chain ONE lane: 60 doubles of trajectory state per thread -> 255 registers -> 8 warps
// per SM, and measured (profiles/ncu_nuts_r02c.txt, profiles/nuts_tail_probe_r01.txt):
//   * the leapfrog step executes with 15.8 of 32 lanes active: on rosenbrock / the funnel the run is bound by its
//     heaviest chains (a chain that keeps building depth-10 trees has 1023 serial steps per transition; per-warp
//     lane-time mean/max 0.71 / 0.25), and finer work items only add hand-offs (MCMC_NUTS_CHUNK sweep,
//     profiles/nuts_chunk_r02d.txt: 1 iteration per item is 2.5x slower than 16);
//   * on homogeneous targets (standard normal: depth 3 everywhere) the bound is the divergent bookkeeping between
//     subtrees — one visit per 2.3 steps — with two warps per scheduler to hide it.
// Here four lanes share a chain, five dimensions each.  The per-thread state shrinks four-fold (~100 registers ->
// ~20 warps per SM), so bookkeeping latency hides behind other warps; a leapfrog step of the chain on the critical path
// issues 29 DFMA per lane instead of 116; every copy of the bookkeeping (tree turn-round, proposal swap, home-row
// hand-off) is four times shorter; divergence is per 8 chains instead of per 32.  The price: two-level shuffle
// reductions for the energy / log-density / U-turn sums (quad_sum) and control flow computed redundantly by the four
// lanes.  Sums are taken in a different order than in nuts_kernel.cuh -> results agree with it (and with the oracle)
// to rounding, tree depths identical; the strict build keeps nuts_kernel.cuh, whose order is the reference's.
//
// Warp-level convergence: every shuffle names the mask of lanes that execute it.  The bookkeeping is therefore written
// as a warp-uniform loop of predicated stages; before each stage that communicates inside the quad ALL lanes of the
// parent mask evaluate the stage's predicate in a ballot, and the stage's shuffles use that ballot.  Predicates are
// quad-uniform by construction (the four lanes hold identical control state).
#pragma once
#include "nuts_kernel.cuh"
#include "targets_quad.cuh"

namespace mcmc {

#ifndef MCMC_NUTS_QUAD_BLOCK
#define MCMC_NUTS_QUAD_BLOCK 128
#endif
#ifndef MCMC_NUTS_QUAD_MINBLOCKS
#define MCMC_NUTS_QUAD_MINBLOCKS 5
#endif
#ifndef MCMC_NUTS_QUAD_BK_TRIPS
#define MCMC_NUTS_QUAD_BK_TRIPS 4
#endif
#ifndef MCMC_NUTS_QUAD_BK_LANES
#define MCMC_NUTS_QUAD_BK_LANES 12
#endif

template <int TID, int BLOCK>
__global__ void __launch_bounds__(BLOCK, MCMC_NUTS_QUAD_MINBLOCKS) nuts_quad_kernel(const NutsKArgs a) {
  constexpr int n = kQuadDim, E = kQuadE;
  constexpr unsigned kFull = 0xffffffffu;
  const int64_t C = a.p.n_chains, S = a.p.num_samples, B = a.p.burn_in;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int part = lane & 3;            // which five dimensions of the chain
  const int d0 = part * E;
  const bool leader = part == 0;
  const bool store = !(a.p.flags & MCMC_FLAG_NO_SAMPLES) && a.b.samples != nullptr;
  const bool do_w = (a.p.flags & MCMC_FLAG_ACCUM_WELFORD) && a.b.welford_mean && a.b.welford_m2;
  const bool injected = a.p.rng_mode == MCMC_RNG_INJECTED;
  const int max_depth = a.p.max_tree_depth;
  const double eps = a.b.step_size_dev ? __ldg(a.b.step_size_dev) : a.p.step_size;
  const double delta_max = a.p.delta_max;

  extern __shared__ __align__(16) double smem[];
  double* sm_w = smem;
  double* sm_sqw = smem + n;
  double* sm_tab = smem + 2 * n;
  // lane tile: slot(s, i) = wbase[(s*E + i)*32] — conflict-free; 5 slots x 5 values x 32 lanes = 6.4 KB per warp
  double* wbase = smem + ((2 * n + target_table_len(TID, n) + 1) / 2) * 2 + (size_t)warp * (SL_COUNT * E * 32) + lane;
#define QSLOT(s, i) wbase[((s) * E + (i)) * 32]

  for (int d = tid; d < n; d += BLOCK) {
    const double w = a.b.inv_mass ? a.b.inv_mass[d] : 1.0;
    sm_w[d] = w;
    sm_sqw[d] = 1.0 / sqrt(w);      // p0 = z / sqrt(w), NUTS.py:219 (fast build: reciprocal multiply)
  }
  target_table_fill<TID>(a.t, n, sm_tab);
  __syncthreads();                   // the only block barrier: warps are independent from here on
  double wv[E], sqw[E];
#pragma unroll
  for (int i = 0; i < E; ++i) { wv[i] = sm_w[d0 + i]; sqw[i] = sm_sqw[d0 + i]; }

  RngCtx rng;
  rng.k0 = a.p.seed[0]; rng.k1 = a.p.seed[1];
  rng.gid_lo = rng.gid_hi = 0;
  rng.logtab = nullptr;

  // working tree end (the one integrated last) in registers: this lane's five dimensions
  double qs[E], ps[E], gs[E];
  double lps = 0.0, lpp = 0.0;
  int acc = 0;
  double h0 = 0.0, log_u = 0.0, sum_alpha = 0.0, sa = 0.0, h_new = 0.0, u_prop = 0.0, se = 0.0, hse = 0.0;
  int n_valid = 1, n_steps = 0, depth = 0, direction = 1, reg_dir = 1, k = 0, num_steps = 0;
  bool diverged = false, in_subtree = false, in_transition = false;
  bool done = false, have_item = false, waiting = false;
  int64_t it = 0, it_item_end = 0, chain = 0;
  unsigned int chunk = 0;
  double wcount = 0.0;
  unsigned long long n_grad = 0, sum_depth = 0, n_div = 0, acc_total = 0;
  const unsigned int n_items = (unsigned int)C * (unsigned int)a.n_chunks;

  int since_bk = 0;
  for (;;) {
    const unsigned needy = __ballot_sync(kFull, k == 0 && !done);
    const unsigned busy = __ballot_sync(kFull, k > 0);
    if (needy == 0u && busy == 0u) break;                 // every lane is done
    since_bk += 1;
    const bool run_bk = needy && (busy == 0u || __popc(needy) >= MCMC_NUTS_QUAD_BK_LANES || since_bk >= MCMC_NUTS_QUAD_BK_TRIPS);
    if (run_bk) {
      since_bk = 0;
      bool stalled = false;          // this quad's next item waits for its predecessor: look again next visit
      // ---------------------------------------------------------------- bookkeeping: warp-uniform loop of stages
      for (;;) {
        const bool need = k == 0 && !done && !stalled;
        const unsigned m = __ballot_sync(kFull, need);
        if (m == 0u) break;
        if (need) {
          // ---- stage 1: claim an item (leader's atomic, broadcast in the quad)
          const bool want_claim = !have_item && !waiting;
          const unsigned m1 = __ballot_sync(m, want_claim);
          if (want_claim) {
            unsigned int id = leader ? atomicAdd(a.queue, 1u) : 0u;
            id = __shfl_sync(m1, id, 0, 4);
            if (id >= n_items) {
              done = true;
            } else {
              chunk = id / (unsigned int)C;
              chain = (int64_t)(id - chunk * (unsigned int)C);
              it = a.it_begin + (int64_t)chunk * a.chunk_iters;
              it_item_end = (it + a.chunk_iters < a.it_end) ? it + a.chunk_iters : a.it_end;
              waiting = chunk > 0;
            }
          }
          const bool poll = !done && !have_item && waiting;
          const unsigned m2 = __ballot_sync(m, poll);
          if (poll) {
            unsigned int f = leader ? ld_acquire_u32(a.flags + chain) : 0u;
            f = __shfl_sync(m2, f, 0, 4);
            if (f < chunk) {
              stalled = true;
            } else {
              waiting = false;
              __threadfence();       // the three lanes that learnt the flag by shuffle order their home-row loads after it
            }
          }
          // ---- stage 2: load the chain's state (nuts_init at iteration 0, else the home row)
          const bool load = !done && !stalled && !have_item;
          const unsigned m3 = __ballot_sync(m, load && it == 0);
          if (load) {
            const uint64_t gid = (uint64_t)(a.p.chain_offset + chain);
            rng.gid_lo = (uint32_t)gid; rng.gid_hi = (uint32_t)(gid >> 32);
            if (it == 0) {           // NUTS.py:73-86
              const double* src = a.b.init_q + chain * n + d0;
#pragma unroll
              for (int i = 0; i < E; ++i) qs[i] = src[i];
              lpp = target_eval_quad<TID, true, true>(a.t, qs, gs, sm_tab, part, m3);
              acc = 0;
#pragma unroll
              for (int i = 0; i < E; ++i) { QSLOT(SL_QP, i) = qs[i]; QSLOT(SL_GP, i) = gs[i]; }
            } else {
              const double* hq = a.home + (size_t)chain * 2 * n + d0;
#pragma unroll
              for (int i = 0; i < E; ++i) { QSLOT(SL_QP, i) = __ldcg(hq + i); QSLOT(SL_GP, i) = __ldcg(hq + n + i); }
              lpp = __ldcg(a.home_lp + chain);
              acc = __ldcg(a.home_acc + chain);
            }
            wcount = a.b.welford_count + (double)(it > B ? it - B : 0);
            have_item = true;
            in_transition = false;
            in_subtree = false;
          }
          const bool act = have_item && !done && !stalled;
          // ---- stage 3: a subtree ended (NUTS.py:281-378) — validity, proposal swap, U-turn test
          bool u_turn = false;
          const bool st_end = act && in_subtree;
          const unsigned m4 = __ballot_sync(m, st_end);
          if (st_end) {
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
              for (int i = 0; i < E; ++i) { QSLOT(SL_QP, i) = qs[i]; QSLOT(SL_GP, i) = gs[i]; }
              lpp = lps;
            }
            n_valid = total_valid;
            sa = sa + exp_nonpos(min_nan(0.0, h0 - h_new));   // the last step's weight
            sum_alpha = sum_alpha + sa;
            n_steps += num_steps;
            n_grad += leader ? (unsigned long long)num_steps : 0ull;
            depth += 1;
            diverged = diverged || is_div;
            in_subtree = false;
            // cond_fn (NUTS.py:243-248); with e = q_work - q_other (see nuts_kernel.cuh)
            double eo = 0.0, ew = 0.0;
#pragma unroll
            for (int i = 0; i < E; ++i) {
              const double e = qs[i] - QSLOT(SL_QO, i);
              eo += e * QSLOT(SL_PO, i);
              ew += e * ps[i];
            }
            eo = quad_sum(m4, eo);
            ew = quad_sum(m4, ew);
            const double dl = reg_dir > 0 ? eo : -ew, dr = reg_dir > 0 ? ew : -eo;
            u_turn = (dl < 0) | (dr < 0);
          }
          // ---- stage 4: a new transition (NUTS.py:217-241) — momentum, energy, slice
          const bool new_tr = act && !in_transition;
          const unsigned m5 = __ballot_sync(m, new_tr);
          if (new_tr) {
            if (it == B) acc = 0;  // NUTS.py:494-500
            const double* zrow = a.z + ((size_t)(it - a.z_it0) * C + chain) * n;
            if (it + 1 < it_item_end && part < 2)   // next transition's draws on their way to L2 (160 B = two lines)
              asm volatile("prefetch.global.L2 [%0];" ::"l"(reinterpret_cast<const char*>(zrow + (size_t)C * n) + 128 * part));
            double kin = 0.0;
#pragma unroll
            for (int i = 0; i < E; ++i) {
              ps[i] = zrow[d0 + i] * sqw[i];
              kin += (ps[i] * ps[i]) * wv[i];
              qs[i] = QSLOT(SL_QP, i);
              gs[i] = QSLOT(SL_GP, i);
            }
            h0 = -lpp + 0.5 * quad_sum(m5, kin);
            const double lgu = a.logu ? a.logu[(size_t)(it - a.z_it0) * C + chain] : log(a.b.u_inject[(size_t)it * C + chain]);
            log_u = lgu - h0;
            n_valid = 1; n_steps = 0; sum_alpha = 0.0; depth = 0; diverged = false;
            in_transition = true;
            reg_dir = 0;   // registers hold BOTH ends
            u_turn = false;
          }
          // ---- stage 5: next subtree, or the transition ends
          bool item_ends = false;
          if (act) {
            if ((depth < max_depth) && !u_turn && !diverged) {
              // start the next subtree of 2^depth steps (NUTS.py:254-279); the four lanes draw the same uniforms
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
#pragma unroll
                for (int i = 0; i < E; ++i) { QSLOT(SL_QO, i) = qs[i]; QSLOT(SL_PO, i) = ps[i]; QSLOT(SL_GO, i) = gs[i]; }
              } else if (direction != reg_dir) {
                // the tree turns round: the working end and the opposite end change places
#pragma unroll
                for (int i = 0; i < E; ++i) {
                  const double tq = QSLOT(SL_QO, i), tp = QSLOT(SL_PO, i), tg = QSLOT(SL_GO, i);
                  QSLOT(SL_QO, i) = qs[i]; QSLOT(SL_PO, i) = ps[i]; QSLOT(SL_GO, i) = gs[i];
                  qs[i] = tq; ps[i] = tp; gs[i] = tg;
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
              if (leader) {
                sum_depth += (unsigned long long)depth;
                n_div += diverged ? 1ull : 0ull;
              }
              const int64_t si = it - B;
              if (si >= 0) {
                const size_t o = (size_t)si * C + chain;
                if (store || do_w) {
#pragma unroll
                  for (int i = 0; i < E; ++i) qs[i] = QSLOT(SL_QP, i);
                }
                if (store) {
                  double* dst = a.b.samples + o * n + d0;
#pragma unroll
                  for (int i = 0; i < E; ++i) dst[i] = qs[i];
                  if (leader && a.b.log_probs) a.b.log_probs[o] = lpp;
                }
                if (leader && a.b.tree_depths) a.b.tree_depths[o] = depth;
                if (leader && a.b.mean_accept_probs) a.b.mean_accept_probs[o] = map;
                if (do_w) {  // tuning/welford.py:42-59 on this lane's five dimensions
                  wcount += 1.0;
                  double* wm = a.b.welford_mean + chain * n + d0;
                  double* w2 = a.b.welford_m2 + chain * n + d0;
                  const double rcount = 1.0 / wcount;
#pragma unroll
                  for (int i = 0; i < E; ++i) {
                    const double mean = __ldcg(wm + i);
                    const double delta = qs[i] - mean;
                    const double mean2 = fma(delta, rcount, mean);
                    __stcg(wm + i, mean2);
                    __stcg(w2 + i, __ldcg(w2 + i) + delta * (qs[i] - mean2));
                  }
                }
              }
              in_transition = false;
              it += 1;
              item_ends = it == it_item_end;
            }
          }
          // ---- stage 6: the item is finished: state home (or, after the last iteration, to the outputs)
          // (the ballot names only the quads that take the home-row branch: a quad finishing the LAST iteration in the
          // same pass never reaches the __syncwarp below)
          const unsigned m6 = __ballot_sync(m, item_ends && it != B + S);
          if (item_ends) {
            if (it == B + S) {
              if (S == 0) acc = 0;  // reset after burn-in even when nothing is sampled (NUTS.py:494-500)
              if (leader) {
                acc_total += (unsigned long long)acc;
                if (a.b.accept_count) a.b.accept_count[chain] = acc;
                if (a.b.accept_rate) a.b.accept_rate[chain] = (float)acc / (float)S;  // NUTS.py:512
                if (a.b.final_lp) a.b.final_lp[chain] = lpp;
              }
#pragma unroll
              for (int i = 0; i < E; ++i) {
                if (a.b.final_q) a.b.final_q[chain * n + d0 + i] = QSLOT(SL_QP, i);
                if (a.b.final_g) a.b.final_g[chain * n + d0 + i] = QSLOT(SL_GP, i);
              }
            } else {
              double* hq = a.home + (size_t)chain * 2 * n + d0;
#pragma unroll
              for (int i = 0; i < E; ++i) { __stcg(hq + i, QSLOT(SL_QP, i)); __stcg(hq + n + i, QSLOT(SL_GP, i)); }
              if (leader) { __stcg(a.home_lp + chain, lpp); __stcg(a.home_acc + chain, acc); }
              __threadfence();
              __syncwarp(m6);        // the four lanes' rows are out before the leader publishes
              if (leader) st_release_u32(a.flags + chain, chunk + 1u);
            }
            have_item = false;
          }
        }
      }
    }

    // ---------------------------------------------------------------- one leapfrog step (convergent, hot)
    const unsigned ms = __ballot_sync(kFull, k > 0);
    if (k > 0) {
      const double w_prev = (k != num_steps) ? exp_nonpos(min_nan(0.0, h0 - h_new)) : 0.0;   // deferred, as in nuts_kernel.cuh
#pragma unroll
      for (int i = 0; i < E; ++i) {
        ps[i] = ps[i] + hse * gs[i];              // NUTS.py:107
        qs[i] = qs[i] + se * (ps[i] * wv[i]);     // NUTS.py:109
      }
      lps = target_eval_quad<TID, true, true>(a.t, qs, gs, sm_tab, part, ms);  // NUTS.py:111
      double kin = 0.0;
#pragma unroll
      for (int i = 0; i < E; ++i) {
        ps[i] = ps[i] + hse * gs[i];              // NUTS.py:115
        kin += (ps[i] * ps[i]) * wv[i];
      }
      h_new = -lps + 0.5 * quad_sum(ms, kin);     // NUTS.py:178 / :282
      sa = sa + w_prev;
      k -= 1;
    } else if (ms == 0u) {
      __nanosleep(100);  // every quad of the warp waits for another warp's item (or is finished)
    }
  }
#undef QSLOT
  if (a.b.counters && leader) {
    if (acc_total) atomicAdd(a.b.counters + 0, acc_total);
    if (n_grad) atomicAdd(a.b.counters + 1, n_grad);
    if (sum_depth) atomicAdd(a.b.counters + 2, sum_depth);
    if (n_div) atomicAdd(a.b.counters + 3, n_div);
  }
}

// launcher: same workspace layout, segments and item rule as launch_nuts_variant (nuts_kernel.cuh)
template <int TID>
int launch_nuts_quad(const Launch& l) {
  constexpr int BLOCK = MCMC_NUTS_QUAD_BLOCK;
  constexpr int kChainsPerBlock = BLOCK / 4;
  const int64_t T = l.p.burn_in + l.p.num_samples;
  if (T == 0 || l.p.n_chains == 0) return launch_nuts_variant<TID, 20, 32>(l);   // the empty-run path lives there
  NutsKArgs a;
  a.p = l.p;
  a.b = l.b;
  a.t = make_target_ctx(l.p, l.b.cov_inv);
  constexpr int n = kQuadDim;
  const int tabn = target_table_len(TID, n);
  const size_t smem = ((size_t)((2 * n + tabn + 1) / 2) * 2 + (size_t)(BLOCK / 32) * SL_COUNT * kQuadE * 32) * sizeof(double);
  const NutsWs ws = nuts_ws_layout(l.p);
  if (!l.b.workspace || l.b.workspace_bytes < ws.total) {
    set_error("nuts: workspace of %zu bytes required (got %zu)", ws.total, l.b.workspace_bytes);
    return MCMC_EWORKSPACE;
  }
  auto kern = nuts_quad_kernel<TID, BLOCK>;
  if (int rc = cuda_fail(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem),
                         "cudaFuncSetAttribute(nuts_quad)"))
    return rc;
  const int64_t C = l.p.n_chains;
  if (C >= ((int64_t)1 << 31)) { set_error("nuts: more than 2^31 chains"); return MCMC_ENOTSUP; }
  int per_sm = 0;
  if (int rc = cuda_fail(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, BLOCK, smem), "occupancy(nuts_quad)"))
    return rc;
  int64_t grid = (int64_t)per_sm * device_sm_count();
  if (grid > (C + kChainsPerBlock - 1) / kChainsPerBlock) grid = (C + kChainsPerBlock - 1) / kChainsPerBlock;
  if (grid < 1) grid = 1;
  const int64_t slots = grid * kChainsPerBlock;     // chains resident at a time
  char* w = reinterpret_cast<char*>(l.b.workspace);
  a.slots = nullptr;
  a.home = reinterpret_cast<double*>(w + ws.home_off);
  a.home_lp = reinterpret_cast<double*>(w + ws.lp_off);
  a.home_acc = reinterpret_cast<int*>(w + ws.acc_off);
  a.queue = reinterpret_cast<unsigned int*>(w + ws.queue_off);
  a.flags = a.queue + 64;
  const bool native = l.p.rng_mode == MCMC_RNG_PHILOX;
  const char* chunk_env = getenv("MCMC_NUTS_CHUNK");
  int64_t it0 = 0;
  do {
    const int64_t it1 = (native && it0 + ws.tseg < T) ? it0 + ws.tseg : T;
    a.it_begin = it0;
    a.it_end = it1;
    int64_t ci = it1 - it0;
    if (chunk_env && atoll(chunk_env) > 0) ci = atoll(chunk_env);
    else if (C > slots && !no_spin()) ci = (it1 - it0 + 5) / 6;
    if (ci < 1) ci = 1;
    int64_t nch = (it1 - it0 + ci - 1) / ci;
    if (nch * C >= ((int64_t)1 << 32) - (slots + 1)) {  // the queue counter is 32-bit: fall back to longer items
      nch = (((int64_t)1 << 32) - (slots + 2)) / C;
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
      a.z = zb; a.logu = lb; a.z_it0 = it0;
    } else {
      a.z = l.b.z_inject; a.logu = nullptr; a.z_it0 = 0;
    }
    kern<<<(unsigned)grid, BLOCK, smem, l.stream>>>(a);
    if (int rc = cuda_fail(cudaGetLastError(), "nuts_quad_kernel launch")) return rc;
    it0 = it1;
  } while (it0 < T);
  return MCMC_OK;
}

}  // namespace mcmc
