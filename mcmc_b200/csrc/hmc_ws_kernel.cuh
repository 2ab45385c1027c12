// hmc_ws_kernel.cuh — warp-specialised variant of the HMC/GRAHMC kernel (native Philox stream only).
//
// Same arithmetic, same Philox counters and therefore the same bits as hmc_kernel.cuh; what changes is WHO
// draws the momenta.  In hmc_kernel.cuh a thread draws its 20 normals (Philox4x32-10 + Box-Muller: ~1200
// integer and ~500 fp64-pipe instructions) and then integrates; with 255 registers per thread only two warps
// share a scheduler, so while a warp is in the integer-heavy draw the fp64 pipe idles (46% busy at L=16,
// profiles/).  Here one CTA per SM runs
//     8 consumer warps  (2 warpgroups, setmaxnreg -> 232 registers): trajectory, MH test, sample tile
//     4 producer warps  (1 warpgroup,  setmaxnreg -> 40 registers):  Philox + Box-Muller for the consumers
// A producer lane draws for one chain of a consumer warp and leaves the momenta in a two-stage shared-memory
// ring ([d][lane] tiles, mbarrier full/empty pairs).  The integer work of the draw now issues in the slots
// the consumers' DFMAs leave free (one DFMA per two cycles per scheduler), so the fp64 pipe stays fed.
//
// Work distribution is the chunked queue of hmc_kernel.cuh with 256-chain groups: items (group, iteration
// chunk) come from an atomic counter, the state of a group travels between chunks through L2.
//
// Replaces the same reference code as hmc_kernel.cuh (HMC.py:138-314, GRAHMC.py:152-456).
#pragma once
#include <cstdlib>

#include "common.cuh"
#include "hmc_kernel.cuh"
#include "philox.cuh"
#include "rowstore.cuh"
#include "targets.cuh"

namespace mcmc {

constexpr int kWsConsumerWarps = 8;
constexpr int kWsProducerWarps = 4;
constexpr int kWsThreads = 32 * (kWsConsumerWarps + kWsProducerWarps);
constexpr int kWsGroup = 32 * kWsConsumerWarps;   // chains per CTA item
constexpr int kWsStages = 2;
// register split: 256 consumer threads x 208 + 128 producer threads x 88 = 64,512 = the launch allocation
// (384 x 168); a split that needs more than the CTA was launched with blocks in setmaxnreg.inc forever
#ifndef MCMC_WS_CREGS
#define MCMC_WS_CREGS 208
#define MCMC_WS_PREGS 88
#endif
constexpr int kWsConsumerRegs = MCMC_WS_CREGS, kWsProducerRegs = MCMC_WS_PREGS;

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  do {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
  } while (!ok);
}
template <int REGS>
__device__ __forceinline__ void warpgroup_reg_inc() {
  asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;\n" ::"n"(REGS));
}
template <int REGS>
__device__ __forceinline__ void warpgroup_reg_dec() {
  asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;\n" ::"n"(REGS));
}
// barrier among the consumer warps only (the producers never join it)
__device__ __forceinline__ void consumer_bar() {
  asm volatile("bar.sync 1, %0;" ::"n"(32 * kWsConsumerWarps) : "memory");
}

// shared-memory carve-up (doubles)
struct HmcWsSmem {
  int n, L, tabn;
  __host__ __device__ HmcWsSmem(int n_, int L_, int tabn_) : n(n_), L(L_), tabn(tabn_) {}
  __host__ __device__ size_t off_w() const { return 0; }
  __host__ __device__ size_t off_epsw() const { return (size_t)n; }
  __host__ __device__ size_t off_sqw() const { return 2 * (size_t)n; }
  __host__ __device__ size_t off_s() const { return 3 * (size_t)n; }
  __host__ __device__ size_t off_a() const { return off_s() + (size_t)L; }
  __host__ __device__ size_t off_b() const { return off_a() + (size_t)L; }
  __host__ __device__ size_t off_logtab() const { return off_b() + (size_t)L; }
  __host__ __device__ size_t off_tab() const { return off_logtab() + kRngTableDoubles; }
  __host__ __device__ size_t off_bars() const { return ((off_tab() + (size_t)tabn + 1) / 2) * 2; }   // 2*W*stages u64
  __host__ __device__ size_t off_warp() const { return off_bars() + 2 * kWsConsumerWarps * kWsStages; }
  // per consumer warp: q tile [lane][n], g tile [d][lane], z ring stages x [d][lane], log-u ring stages x [lane]
  __host__ __device__ size_t warp_doubles() const { return (size_t)(2 + kWsStages) * n * 32 + kWsStages * 32; }
  __host__ __device__ size_t total_doubles() const { return off_warp() + kWsConsumerWarps * warp_doubles(); }
};

template <int TID, int DS, bool FRICTION>
__global__ void __launch_bounds__(kWsThreads, 1) hmc_ws_kernel(const HmcKArgs a) {
  constexpr int N = DS;
  constexpr int n = DS;
  const int L = a.p.num_steps;
  const int64_t C = a.p.n_chains;
  const int64_t S = a.p.num_samples, B = a.p.burn_in;
  const int tid = threadIdx.x;
  const int warp = tid >> 5, lane = tid & 31;
  const bool store = !(a.p.flags & MCMC_FLAG_NO_SAMPLES) && a.b.samples != nullptr;
  const bool do_w = (a.p.flags & MCMC_FLAG_ACCUM_WELFORD) && a.b.welford_mean && a.b.welford_m2;

  extern __shared__ __align__(16) double smem[];
  const HmcWsSmem lay(n, L, target_table_len(TID, n));
  double* sm_w = smem + lay.off_w();
  double* sm_epsw = smem + lay.off_epsw();
  double* sm_sqw = smem + lay.off_sqw();
  double* sm_s = smem + lay.off_s();
  double* sm_a = smem + lay.off_a();
  double* sm_b = smem + lay.off_b();
  double* sm_tab = smem + lay.off_tab();
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + lay.off_bars());   // full[w][s] then empty[w][s]
  __shared__ unsigned int s_item;

  const double eps = (a.b.step_size_dev ? __ldg(a.b.step_size_dev) : a.p.step_size);
  const double half_eps = 0.5 * eps;

  // ---- prologue (all 12 warps): constants, friction table, target table, barriers
  for (int d = tid; d < n; d += kWsThreads) {
    const double w = a.b.inv_mass ? a.b.inv_mass[d] : 1.0;
    sm_w[d] = w;
    sm_epsw[d] = eps * w;
    sm_sqw[d] = 1.0 / sqrt(w);
  }
  if (FRICTION) {
    const double T = eps * L;
    for (int i = tid; i < L; i += kWsThreads) {
      const double gi = friction_gamma_dev(a.p.schedule_id, i * eps, T, a.p.gamma, a.p.steepness);
      sm_s[i] = exp(-gi * half_eps);
    }
  }
  target_table_fill<TID>(a.t, n, sm_tab);
  rng_table_fill(smem + lay.off_logtab());
  if (tid < 2 * kWsConsumerWarps * kWsStages) mbar_init(bars + tid, 1);
  __syncthreads();
  if (FRICTION) {
    for (int i = tid; i + 1 < L; i += kWsThreads) {
      const double ss = sm_s[i] * sm_s[i + 1];
      sm_a[i] = ss;
      sm_b[i] = half_eps * ss + half_eps;
    }
  }
  __syncthreads();

  const unsigned int n_items = (unsigned int)a.n_groups * (unsigned int)a.n_chunks;
  const bool is_producer = warp >= kWsConsumerWarps;
  // ring bookkeeping: completed uses of each stage (consumers: their own ring; producers: [consumer half][stage])
  uint32_t uses[2] = {0, 0};
  uint32_t puses[2] = {0, 0};

  // The two roles never share code after this point: ptxas sizes the register allocation of each branch by the
  // setmaxnreg that dominates it (a common loop head would pin both to the launch value of 168).
  if (is_producer) {
    warpgroup_reg_dec<kWsProducerRegs>();
    for (;;) {
      // ---- next item (all 384 threads meet on barrier 0, each role from its own loop)
      unsigned int item;
      if (a.n_chunks == 1) {
        item = blockIdx.x;
      } else {
        __syncthreads();
        if (tid == 0) s_item = atomicAdd(a.queue, 1u);
        __syncthreads();
        item = s_item;
      }
      if (item >= n_items) break;
      const int grp = (int)(item % (unsigned int)a.n_groups);
      const int chunk = (int)(item / (unsigned int)a.n_groups);
      const int64_t it_begin = (int64_t)chunk * a.chunk_iters;
      const int64_t it_end = (it_begin + a.chunk_iters < B + S) ? it_begin + a.chunk_iters : B + S;
      const int64_t grp0 = (int64_t)grp * kWsGroup;
      // ================================================================== producers
      // producer warp pw serves consumer warps pw and pw + 4, iteration by iteration
      const int pw = warp - kWsConsumerWarps;
      RngCtx rng;
      rng.k0 = a.p.seed[0]; rng.k1 = a.p.seed[1];
      rng.logtab = smem + lay.off_logtab();
      // both consumer warps of this producer are served in ONE interleaved pass: four independent Philox /
      // Box-Muller streams per loop trip (2 warps x 2 pairs) give the scheduler the ILP a single dependent
      // stream lacks (a producer shares its fp64 pipe with two saturating consumers)
      RngCtx rngB = rng;
      {
        int64_t cA = grp0 + pw * 32 + lane, cB = grp0 + (pw + kWsProducerWarps) * 32 + lane;
        if (cA >= C) cA = C - 1;                               // tail lanes shadow the last chain (as the consumers do)
        if (cB >= C) cB = C - 1;
        const uint64_t gA = (uint64_t)(a.p.chain_offset + cA), gB = (uint64_t)(a.p.chain_offset + cB);
        rng.gid_lo = (uint32_t)gA; rng.gid_hi = (uint32_t)(gA >> 32);
        rngB.gid_lo = (uint32_t)gB; rngB.gid_hi = (uint32_t)(gB >> 32);
      }
      double* baseA = smem + lay.off_warp() + pw * lay.warp_doubles();
      double* baseB = smem + lay.off_warp() + (pw + kWsProducerWarps) * lay.warp_doubles();
      for (int64_t it = it_begin; it < it_end; ++it) {
        const int st = (int)((it - it_begin) & 1);
        double* zA = baseA + (size_t)(2 + st) * n * 32 + lane;
        double* zB = baseB + (size_t)(2 + st) * n * 32 + lane;
        // the consumers must have drained this stage (first use of a stage passes: parity of a fresh barrier)
        const uint32_t par = (puses[st] & 1u) ^ 1u;
        mbar_wait(bars + (kWsConsumerWarps * kWsStages + pw * kWsStages + st), par);
        mbar_wait(bars + (kWsConsumerWarps * kWsStages + (pw + kWsProducerWarps) * kWsStages + st), par);
#pragma unroll 2
        for (int j = 0; 2 * j < n; ++j) {
          double a0, a1, b0, b1;
          philox_normals_block(rng, (uint32_t)it, (uint32_t)j, a0, a1);
          philox_normals_block(rngB, (uint32_t)it, (uint32_t)j, b0, b1);
          zA[(2 * j) * 32] = a0;
          zB[(2 * j) * 32] = b0;
          if (2 * j + 1 < n) {
            zA[(2 * j + 1) * 32] = a1;
            zB[(2 * j + 1) * 32] = b1;
          }
        }
        baseA[(size_t)(2 + kWsStages) * n * 32 + st * 32 + lane] = philox_accept_log_uniform(rng, (uint32_t)it);
        baseB[(size_t)(2 + kWsStages) * n * 32 + st * 32 + lane] = philox_accept_log_uniform(rngB, (uint32_t)it);
        __syncwarp();
        if (lane == 0) {
          mbar_arrive(bars + (pw * kWsStages + st));
          mbar_arrive(bars + ((pw + kWsProducerWarps) * kWsStages + st));
        }
        puses[st] += 1u;
      }
      if (a.n_chunks == 1) break;
    }
    return;
  }

  warpgroup_reg_inc<kWsConsumerRegs>();
  for (;;) {
    // ---- next item (all 384 threads meet on barrier 0, each role from its own loop)
    unsigned int item;
    if (a.n_chunks == 1) {
      item = blockIdx.x;
    } else {
      __syncthreads();
      if (tid == 0) s_item = atomicAdd(a.queue, 1u);
      __syncthreads();
      item = s_item;
    }
    if (item >= n_items) break;
    const int grp = (int)(item % (unsigned int)a.n_groups);
    const int chunk = (int)(item / (unsigned int)a.n_groups);
    const int64_t it_begin = (int64_t)chunk * a.chunk_iters;
    const int64_t it_end = (it_begin + a.chunk_iters < B + S) ? it_begin + a.chunk_iters : B + S;
    const int64_t grp0 = (int64_t)grp * kWsGroup;
    // ==================================================================== consumers
    const int64_t chain0 = grp0 + warp * 32;
    const bool active = chain0 + lane < C;
    const int64_t chain = active ? chain0 + lane : C - 1;
    int nvalid = (int)(C - chain0);
    nvalid = nvalid < 0 ? 0 : (nvalid > 32 ? 32 : nvalid);
    double* my_base = smem + lay.off_warp() + warp * lay.warp_doubles();
    double* wq = my_base;                   // [lane][n], TMA source of the warp's sample rows
    double* wg = my_base + (size_t)n * 32;  // [d][lane]
    double* my_q = wq + (size_t)lane * n;
    const bool bulk = store && nvalid > 0 && ((((size_t)nvalid * n * sizeof(double)) & 15) == 0) &&
                      ((((size_t)(a.b.samples + chain0 * n)) & 15) == 0) && ((((size_t)C * n * sizeof(double)) & 15) == 0);

    double q[N], p[N], g[N];
    double lp;
    int acc;
    if (chunk == 0) {
      const double* src = a.b.init_q + chain * n;
#pragma unroll
      for (int d = 0; d < n; ++d) q[d] = src[d];
      lp = target_eval<TID, DS, true, true>(a.t, n, q, g, sm_tab);
      acc = 0;
      store_row<DS>(my_q, q, n);
#pragma unroll
      for (int d = 0; d < n; ++d) wg[d * 32 + lane] = g[d];
    } else {
      if (tid == 0) {
        while (ld_acquire_u32(a.flags + grp) < (unsigned int)chunk) __nanosleep(200);
      }
      consumer_bar();
      const double* sq = a.ho_q + (size_t)chain0 * n;
      const double* sg = a.ho_g + (size_t)chain0 * n;
      for (int k = lane; k < 32 * n; k += 32) {
        wq[k] = __ldcg(sq + k);
        wg[k] = __ldcg(sg + k);
      }
      lp = __ldcg(a.ho_lp + chain0 + lane);
      acc = __ldcg(a.ho_acc + chain0 + lane);
    }
    __syncwarp();
    double wcount = a.b.welford_count + (double)((it_begin > B) ? (it_begin - B) : 0);

    for (int64_t it = it_begin; it < it_end; ++it) {
      if (it == B) acc = 0;
      const int64_t si = it - B;
      const int st = (int)((it - it_begin) & 1);
      const double* wz = my_base + (size_t)(2 + st) * n * 32;
      const double* wlu = my_base + (size_t)(2 + kWsStages) * n * 32 + st * 32;
      uint64_t* full = bars + (warp * kWsStages + st);
      uint64_t* empty = bars + (kWsConsumerWarps * kWsStages + warp * kWsStages + st);

      // ---- momentum from the producers' ring, p0 = z / sqrt(w), H0
      mbar_wait(full, uses[st] & 1u);
      double kin = 0.0;
#pragma unroll
      for (int d = 0; d < n; ++d) {
        p[d] = wz[d * 32 + lane] * sm_sqw[d];
        kin += (p[d] * p[d]) * sm_w[d];
      }
      const double log_u = wlu[lane];
      __syncwarp();
      if (lane == 0) mbar_arrive(empty);
      uses[st] += 1u;
      const double H0 = -lp + 0.5 * kin;

      load_row<DS>(q, my_q, n);
#pragma unroll
      for (int d = 0; d < n; ++d) g[d] = wg[d * 32 + lane];
      double lpn = lp;
      if (L > 0) {
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

      double kin1 = 0.0;
#pragma unroll
      for (int d = 0; d < n; ++d) kin1 += (p[d] * p[d]) * sm_w[d];
      double H1 = -lpn + 0.5 * kin1;
      if (!isfinite(H1)) H1 = 1e10;
      const double log_alpha = H0 - H1;
      const bool accept = log_u < min_nan(0.0, log_alpha);

      // the TMA copy of the previous iteration's tile must have drained wq before it is overwritten
      if (bulk) {
        if (lane == 0) bulk_wait_read<0>();
        __syncwarp();
      }
      if (accept) {
        store_row<DS>(my_q, q, n);
#pragma unroll
        for (int d = 0; d < n; ++d) wg[d * 32 + lane] = g[d];
        lp = lpn;
        acc += 1;
      }
      if (si >= 0) {
        if (store) {
          double* gtile = a.b.samples + ((size_t)si * C + chain0) * n;
          if (bulk) {
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0) bulk_store_s2g(gtile, wq, (uint32_t)((size_t)nvalid * n * sizeof(double)));
          } else if (active) {
            if (!accept) load_row<DS>(q, my_q, n);
            store_row<DS>(gtile + (size_t)lane * n, q, n);
          }
          if (active && a.b.log_probs) a.b.log_probs[(size_t)si * C + chain] = lp;
        }
        if (do_w && active) {
          if (!accept) load_row<DS>(q, my_q, n);
          wcount += 1.0;
          double* wm = a.b.welford_mean + chain * n;
          double* w2 = a.b.welford_m2 + chain * n;
          const double rcount = 1.0 / wcount;  // same rounding sequence as hmc_kernel (fast build)
#pragma unroll
          for (int d = 0; d < n; ++d) {
            const double mean = __ldcg(wm + d);
            const double delta = q[d] - mean;
            const double mean2 = MCMC_IS_FAITHFUL ? mean + delta / wcount : fma(delta, rcount, mean);
            __stcg(wm + d, mean2);
            __stcg(w2 + d, __ldcg(w2 + d) + delta * (q[d] - mean2));
          }
        }
      }
    }
    if (bulk && lane == 0) bulk_wait_all();
    __syncwarp();

    if (chunk != a.n_chunks - 1) {
      // ---- hand the warp's state tiles to whoever takes the group's next chunk
      double* dq = a.ho_q + (size_t)chain0 * n;
      double* dg = a.ho_g + (size_t)chain0 * n;
      for (int k = lane; k < 32 * n; k += 32) {
        __stcg(dq + k, wq[k]);
        __stcg(dg + k, wg[k]);
      }
      __stcg(a.ho_lp + chain0 + lane, lp);
      __stcg(a.ho_acc + chain0 + lane, acc);
      __threadfence();
      consumer_bar();
      if (tid == 0) st_release_u32(a.flags + grp, (unsigned int)chunk + 1u);
    } else {
      if (S == 0) acc = 0;  // reset after burn-in even when nothing is sampled (HMC.py:276-282)
      load_row<DS>(q, my_q, n);
#pragma unroll
      for (int d = 0; d < n; ++d) g[d] = wg[d * 32 + lane];
      if (active) {
        if (a.b.accept_count) a.b.accept_count[chain] = acc;
        if (a.b.accept_rate) a.b.accept_rate[chain] = (float)acc / (float)S;
        if (a.b.final_q) store_row<DS>(a.b.final_q + chain * n, q, n);
        if (a.b.final_g) store_row<DS>(a.b.final_g + chain * n, g, n);
        if (a.b.final_lp) a.b.final_lp[chain] = lp;
      }
      if (a.b.counters) {
        unsigned int v = active ? (unsigned int)acc : 0u;
        v = __reduce_add_sync(0xffffffffu, v);
        if (lane == 0 && v) atomicAdd(a.b.counters + 0, (unsigned long long)v);
        if (grp == 0 && tid == 0)
          atomicAdd(a.b.counters + 1, (unsigned long long)C * (unsigned long long)(B + S) * (unsigned long long)L);
      }
    }
    __syncwarp();
    if (a.n_chunks == 1) break;
  }
}


// host-side launcher; returns MCMC_ENOTSUP (without setting an error) when the plain kernel should be used
template <int TID, int DS, bool FRICTION>
int launch_hmc_ws(const Launch& l, bool* used) {
  *used = false;
  const int n = l.p.n_dim;
  if (l.p.rng_mode != MCMC_RNG_PHILOX || (l.p.flags & MCMC_FLAG_TRACK_PROPOSALS)) return MCMC_OK;
  // Opt-in (MCMC_HMC_WS=1).  Measured on B200, rosenbrock D=20, 65,536 chains, 64 iterations (profiles/
  // ncu_hmc_ws_r01.txt): 1.08 / 1.59 / 2.52 / 4.50 ms at L = 16 / 32 / 64 / 128 against 1.03 / 1.45 / 2.38 / 4.26 ms
  // of hmc_kernel: with 208 instead of 254 registers the consumers' trajectory loop loses more ILP than the
  // overlapped draw gains, and at L = 16 the producers (sharing a saturated fp64 pipe) still starve them.
  const char* force = getenv("MCMC_HMC_WS");
  if (!(force && force[0] == '1')) return MCMC_OK;
  const HmcWsSmem lay(n, l.p.num_steps, target_table_len(TID, n));
  const size_t smem = lay.total_doubles() * sizeof(double);
  if (smem > 227 * 1024) return MCMC_OK;
  const int64_t groups = (l.p.n_chains + kWsGroup - 1) / kWsGroup;
  const int64_t slots = device_sm_count();
  const int64_t iters = l.p.burn_in + l.p.num_samples;
  if (groups == 0 || iters == 0) return MCMC_OK;
  HmcKArgs a;
  a.p = l.p;
  a.b = l.b;
  a.t = make_target_ctx(l.p, l.b.cov_inv);
  a.n_groups = (int)groups;
  a.n_chunks = 1;
  a.chunk_iters = (int)iters;
  a.queue = nullptr; a.flags = nullptr; a.ho_q = a.ho_g = a.ho_lp = nullptr; a.ho_acc = nullptr;
  int64_t grid = groups;
  auto kern = hmc_ws_kernel<TID, DS, FRICTION>;
  if (int rc = cuda_fail(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem),
                         "cudaFuncSetAttribute(hmc_ws)"))
    return rc;
  const size_t need = hmc_workspace_bytes(l.p);
  const char* fchunks = getenv("MCMC_HMC_FORCE_CHUNKS");
  if ((groups > slots || fchunks) && iters >= 2 && l.b.workspace && l.b.workspace_bytes >= need) {
    const char* cost_env = getenv("MCMC_HMC_ITEM_COST");
    const int64_t item_cost = cost_env ? atoi(cost_env) : 450;
    int64_t ci = item_cost / (l.p.num_steps + 14);
    const int64_t ci_bal = iters * groups / (6 * slots);
    if (ci > ci_bal) ci = ci_bal;
    if (ci < 1) ci = 1;
    int64_t nc = (iters + ci - 1) / ci;
    if (fchunks) { nc = atoi(fchunks); if (nc < 2) nc = 2; if (nc > iters) nc = iters; ci = (iters + nc - 1) / nc; }
    nc = (iters + ci - 1) / ci;
    if (nc > 1) {
      a.n_chunks = (int)nc;
      a.chunk_iters = (int)ci;
      char* w = reinterpret_cast<char*>(l.b.workspace);
      const size_t C = ((size_t)l.p.n_chains + 255) / 256 * 256, D = (size_t)n;
      const size_t flag_bytes = (((C + 31) / 32) * 4 + 255) / 256 * 256;
      a.queue = reinterpret_cast<unsigned int*>(w);
      a.flags = reinterpret_cast<unsigned int*>(w + 256);
      a.ho_q = reinterpret_cast<double*>(w + 256 + flag_bytes);
      a.ho_g = a.ho_q + C * D;
      a.ho_lp = a.ho_g + C * D;
      a.ho_acc = reinterpret_cast<int*>(a.ho_lp + C);
      if (int rc = cuda_fail(cudaMemsetAsync(w, 0, 256 + flag_bytes, l.stream), "memset(hmc queue)")) return rc;
      grid = slots < groups ? slots : groups;
      if (fchunks && grid > 1) grid = (grid + 1) / 2;
    }
  }
  kern<<<(unsigned)grid, kWsThreads, smem, l.stream>>>(a);
  *used = true;
  return cuda_fail(cudaGetLastError(), "hmc_ws_kernel launch");
}

}  // namespace mcmc
