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
  const double scale = a.b.step_size_dev ? __ldg(a.b.step_size_dev) : a.p.step_size;

  extern __shared__ __align__(16) double smem[];
  const int tabn = target_table_len(TID, n);
  double* sm_logtab = smem;
  double* sm_tab = smem + kRngTableDoubles;
  double* sm_stage = sm_tab + ((tabn + 1) / 2) * 2;
  target_table_fill<TID>(a.t, n, sm_tab);
  rng_table_fill(sm_logtab);
  __syncthreads();

  TileWriter<DS, BLOCK> tw;
  if (store) tw.init(sm_stage, n, nvalid, a.b.samples + chain0 * n, (size_t)C * n * sizeof(double));

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

  if (S == 0) acc = 0;  // reset after burn-in even when nothing is sampled (RWMH.py:160-166)
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

// ---------------------------------------------------------------------------------------------------------------
// rwmh_warp_kernel — the compiled-D kernel.  RWMH has no trajectory: an iteration IS the draw (11 Philox blocks + 10
// Box-Muller pairs at D=20, ~1600 warp-instructions) and that makes the kernel issue / latency bound, not HBM bound.
// With the draw fully unrolled into registers (rwmh_kernel above: 213 registers, 2 warps per scheduler) the SM issued
// 2.2 instructions per clock.  Here the operands of the draw live in shared memory — the current position as the warp's
// row-major tile (which is also the TMA source of the sample tile: no separate staging copy), the proposal as a
// conflict-free [d][lane] tile — so a thread needs 128 registers instead of 213 and four 4-warp blocks (16 warps) share
// an SM; the loop over pairs can be rolled (MCMC_RWMH_DRAW_UNROLL) for still more warps, which measured slower.
// Warps are independent: each owns 32 consecutive chains, its own tiles and its own bulk-store group (no block barrier
// in the iteration).  Arithmetic per element is the same expression as before: results are bit-identical.
// Measured at 4 M chains (profiles/rwmh_sweep_r01.txt): unroll 1 / 2 / 5 / 10 at 5 blocks per SM -> 2.10 / 2.13 / 2.32 /
// 2.38 TB/s, unroll 10 at 4 blocks (128 registers) -> 2.40 TB/s; the register-resident kernel above: 2.22 TB/s.
#ifndef MCMC_RWMH_DRAW_UNROLL
#define MCMC_RWMH_DRAW_UNROLL 10
#endif
#ifndef MCMC_RWMH_WARP_MINBLOCKS
#define MCMC_RWMH_WARP_MINBLOCKS 4
#endif
constexpr int kRwmhDrawUnroll = MCMC_RWMH_DRAW_UNROLL;
// pairs per stage-interleaved draw batch (philox.cuh); measured at 4 M chains (profiles/draw_batch_ab_r02.txt):
// 0 -> 2.93 TB/s, 2 -> 3.08, 5 -> 2.95, 10 -> 2.85
#ifndef MCMC_RWMH_DRAW_BATCH
#define MCMC_RWMH_DRAW_BATCH 2
#endif
constexpr int kRwmhDrawBatch = MCMC_RWMH_DRAW_BATCH;

template <int TID, int DS, int BLOCK>
__global__ void __launch_bounds__(BLOCK, MCMC_RWMH_WARP_MINBLOCKS) rwmh_warp_kernel(const RwmhKArgs a) {
  static_assert(DS > 0 && DS % 2 == 0, "compiled even dimension");
  constexpr int n = DS;
  constexpr int kWarps = BLOCK / 32;
  const int64_t C = a.p.n_chains, S = a.p.num_samples, B = a.p.burn_in;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const bool store = !(a.p.flags & MCMC_FLAG_NO_SAMPLES) && a.b.samples != nullptr;
  const bool injected = a.p.rng_mode == MCMC_RNG_INJECTED;
  const double scale = a.b.step_size_dev ? __ldg(a.b.step_size_dev) : a.p.step_size;

  extern __shared__ __align__(16) double smem[];
  const int tabn = target_table_len(TID, n);
  double* sm_logtab = smem;
  double* sm_tab = smem + kRngTableDoubles;
  double* sm_tiles = sm_tab + ((tabn + 1) / 2) * 2;
  double* q_tile = sm_tiles + (size_t)warp * (2 * 32 * n);   // [lane][n] row-major: state + TMA source
  double* p_tile = q_tile + 32 * n;                           // [d][lane]: proposal
  target_table_fill<TID>(a.t, n, sm_tab);
  rng_table_fill(sm_logtab);
  __syncthreads();

  const int64_t chain0 = ((int64_t)blockIdx.x * kWarps + warp) * 32;
  if (chain0 >= C) return;   // no block barrier below this line
  const int nvalid = (int)((C - chain0) < 32 ? (C - chain0) : 32);
  const bool active = lane < nvalid;
  const int64_t chain = active ? chain0 + lane : C - 1;

  RngCtx rng;
  {
    const uint64_t gid = (uint64_t)(a.p.chain_offset + chain);
    rng.k0 = a.p.seed[0]; rng.k1 = a.p.seed[1];
    rng.gid_lo = (uint32_t)gid; rng.gid_hi = (uint32_t)(gid >> 32);
    rng.logtab = sm_logtab;
  }

  double* my_q = q_tile + lane * n;
  double* my_p = p_tile + lane;
  double qn[n];
  load_row<DS>(qn, a.b.init_q + chain * n, n);
  store_row<DS>(my_q, qn, n);
  double lp = target_eval<TID, DS, true, false>(a.t, n, qn, nullptr, sm_tab);  // RWMH.py:67
  int acc = 0;
  bool pending = false;   // a bulk store of q_tile is in flight (lane 0 only)

  for (int64_t it = 0; it < B + S; ++it) {
    if (it == B) acc = 0;
    const int64_t si = it - B;
    // proposal = position + scale * noise (RWMH.py:98-99), pair by pair
    const double* zsrc = injected ? a.b.z_inject + ((size_t)it * C + chain) * n : nullptr;
    if (!injected && kRwmhDrawBatch > 0 && (n / 2) % kRwmhDrawBatch == 0) {
      // stage-interleaved draw of kRwmhDrawBatch pairs at a time (philox.cuh: same bits as the pair-after-pair loop)
#pragma unroll
      for (int j0 = 0; j0 < n / 2; j0 += kRwmhDrawBatch) {
        double zb[2 * (kRwmhDrawBatch > 0 ? kRwmhDrawBatch : 1)];
        philox_normals_batch<(kRwmhDrawBatch > 0 ? kRwmhDrawBatch : 1)>(rng, (uint32_t)it, (uint32_t)j0, zb);
#pragma unroll
        for (int k = 0; k < kRwmhDrawBatch; ++k) {
          const double2 qq = *reinterpret_cast<const double2*>(my_q + 2 * (j0 + k));
          my_p[(2 * (j0 + k)) * 32] = qq.x + scale * zb[2 * k];
          my_p[(2 * (j0 + k) + 1) * 32] = qq.y + scale * zb[2 * k + 1];
        }
      }
    } else {
#pragma unroll kRwmhDrawUnroll
    for (int j = 0; j < n / 2; ++j) {
      double z0, z1;
      if (injected) {
        z0 = zsrc[2 * j];
        z1 = zsrc[2 * j + 1];
      } else {
        philox_normals_block(rng, (uint32_t)it, (uint32_t)j, z0, z1);
      }
      const double2 qq = *reinterpret_cast<const double2*>(my_q + 2 * j);
      my_p[(2 * j) * 32] = qq.x + scale * z0;
      my_p[(2 * j + 1) * 32] = qq.y + scale * z1;
    }
    }
#pragma unroll
    for (int d = 0; d < n; ++d) qn[d] = my_p[d * 32];
    const double lpn = target_eval<TID, DS, true, false>(a.t, n, qn, nullptr, sm_tab);  // RWMH.py:102
    const double log_ratio = lpn - lp;
    const double log_u = injected ? log(a.b.u_inject[(size_t)it * C + chain]) : philox_accept_log_uniform(rng, (uint32_t)it);
    const bool accept = log_u < min_nan(0.0, log_ratio);  // RWMH.py:105-107
    // the TMA engine may still be reading last iteration's tile
    if (pending) { bulk_wait_read<0>(); pending = false; }
    __syncwarp();
    if (accept) {
      store_row<DS>(my_q, qn, n);
      lp = lpn;
      acc += 1;
    }
    if (si >= 0 && store) {
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) {
        bulk_store_s2g(a.b.samples + ((size_t)si * C + chain0) * n, q_tile, (uint32_t)((size_t)nvalid * n * sizeof(double)));
        pending = true;
      }
      if (active && a.b.log_probs) a.b.log_probs[(size_t)si * C + chain] = lp;
    }
  }
  if (lane == 0) bulk_wait_all();
  __syncwarp();

  if (S == 0) acc = 0;  // reset after burn-in even when nothing is sampled (RWMH.py:160-166)
  if (active) {
    if (a.b.accept_count) a.b.accept_count[chain] = acc;
    if (a.b.accept_rate) a.b.accept_rate[chain] = (float)acc / (float)S;  // RWMH.py:177
    if (a.b.final_q) {
      load_row<DS>(qn, my_q, n);
      store_row<DS>(a.b.final_q + chain * n, qn, n);
    }
    if (a.b.final_lp) a.b.final_lp[chain] = lp;
  }
  if (a.b.counters) {
    unsigned int v = active ? (unsigned int)acc : 0u;
    v = __reduce_add_sync(0xffffffffu, v);
    if (lane == 0 && v) atomicAdd(a.b.counters + 0, (unsigned long long)v);
  }
}

template <int TID, int DS, int BLOCK>
int launch_rwmh_warp_variant(const Launch& l) {
  RwmhKArgs a;
  a.p = l.p;
  a.b = l.b;
  a.t = make_target_ctx(l.p, l.b.cov_inv);
  const int n = l.p.n_dim;
  const int tabn = target_table_len(TID, n);
  const size_t smem = ((size_t)kRngTableDoubles + (size_t)((tabn + 1) / 2) * 2 + (size_t)(BLOCK / 32) * 2 * 32 * n) * sizeof(double);
  auto kern = rwmh_warp_kernel<TID, DS, BLOCK>;
  if (int rc = cuda_fail(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem),
                         "cudaFuncSetAttribute(rwmh_warp)"))
    return rc;
  const int64_t grid = (l.p.n_chains + BLOCK - 1) / BLOCK;
  if (grid == 0) return MCMC_OK;
  kern<<<(unsigned)grid, BLOCK, smem, l.stream>>>(a);
  return cuda_fail(cudaGetLastError(), "rwmh_warp_kernel launch");
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
