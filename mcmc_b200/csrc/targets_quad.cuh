// targets_quad.cuh — the benchmark targets evaluated by a QUAD of lanes per chain (D = 20: five dimensions per lane).
//
// Same log-densities and gradients as targets.cuh (benchmarks/targets.py:50,90,130,166,236,300,367,432 through
// vmap(jax.value_and_grad(.)), NUTS.py:111), for the quad-per-chain NUTS kernel (nuts_quad_kernel.cuh): lane `part`
// (0..3) of a quad holds q[5 part .. 5 part + 4]; sums over the 20 dimensions are a 5-term partial per lane followed by a
// two-level butterfly (`quad_sum`: every lane of the quad ends with the same bits), neighbour couplings (rosenbrock)
// and the special first coordinate (funnel, mixture) travel by shuffles inside the quad.  Only the ORDER of the sums
// differs from targets.cuh, i.e. results agree to rounding (the fast build's tolerance, 1e-10 against the oracle).
//
// `mask` is the set of lanes executing the call together (whole quads); every shuffle names it.
#pragma once
#include "common.cuh"
#include "targets.cuh"

namespace mcmc {

constexpr int kQuadDim = 20;   // the compiled dimension
constexpr int kQuadE = 5;      // dimensions per lane

__device__ __forceinline__ double quad_sum(unsigned mask, double v) {
  v += __shfl_xor_sync(mask, v, 1);
  v += __shfl_xor_sync(mask, v, 2);
  return v;
}
// value held by lane `part_src` of the caller's quad
__device__ __forceinline__ double quad_get(unsigned mask, double v, int part_src) {
  return __shfl_sync(mask, v, part_src, 4);
}

template <int TID, bool WANT_LP, bool WANT_G>
__device__ __forceinline__ double target_eval_quad(const TargetCtx& t, const double (&q)[kQuadE], double (&g)[kQuadE],
                                                   const double* __restrict__ tab, const int part, const unsigned mask) {
  constexpr int n = kQuadDim, E = kQuadE;
  const int d0 = part * E;
  double lp = 0.0;
  if (TID == MCMC_TARGET_STANDARD_NORMAL) {
    double s = 0.0;
#pragma unroll
    for (int i = 0; i < E; ++i) {
      if (WANT_LP) s += q[i] * q[i];
      if (WANT_G) g[i] = -q[i];
    }
    if (WANT_LP) lp = -0.5 * (quad_sum(mask, s) + t.c[0]);
  } else if (TID == MCMC_TARGET_CORRELATED_GAUSSIAN) {
    const double a = t.c[0], b = t.c[1];
    double sx = 0.0;
#pragma unroll
    for (int i = 0; i < E; ++i) sx += q[i];
    const double bs = b * quad_sum(mask, sx);
    double maha = 0.0;
#pragma unroll
    for (int i = 0; i < E; ++i) {
      const double y = a * q[i] + bs;
      if (WANT_LP) maha += y * q[i];
      if (WANT_G) g[i] = -y;
    }
    if (WANT_LP) lp = -0.5 * (quad_sum(mask, maha) + t.c[2] + t.c[3]);
  } else if (TID == MCMC_TARGET_ILL_CONDITIONED) {
    double maha = 0.0;
#pragma unroll
    for (int i = 0; i < E; ++i) {
      const double y = q[i] * tab[d0 + i];
      if (WANT_LP) maha += y * q[i];
      if (WANT_G) g[i] = -y;
    }
    if (WANT_LP) lp = -0.5 * (quad_sum(mask, maha) + t.c[0] + t.c[1]);
  } else if (TID == MCMC_TARGET_NEALS_FUNNEL) {
    const double x0 = quad_get(mask, q[0], 0);
    double s = 0.0;
#pragma unroll
    for (int i = 0; i < E; ++i) s += (part == 0 && i == 0) ? 0.0 : q[i] * q[i];
    s = quad_sum(mask, s);
    const double iv = exp(-x0);
    const double sv = s * iv;
    if (WANT_G) {
      const double niv = -iv;
#pragma unroll
      for (int i = 0; i < E; ++i) g[i] = q[i] * niv;
      if (part == 0) g[0] = -x0 / 9.0 + 0.5 * sv - 0.5 * t.c[2];
    }
    if (WANT_LP) {
      const double lp0 = -0.5 * (x0 * x0 / 9.0 + t.c[0]);
      const double lpr = -0.5 * (sv + t.c[2] * x0 + t.c[1]);
      lp = lp0 + lpr;
    }
  } else if (TID == MCMC_TARGET_LOG_GAMMA) {
    const double am1 = t.par[0] - 1.0, beta = t.par[1], lognorm = t.par[2];
    int valid = 1;
    double s = 0.0;
#pragma unroll
    for (int i = 0; i < E; ++i) {
      const double x = q[i];
      valid &= (x > 0.0) ? 1 : 0;
      const double xm = x > 1e-10 ? x : 1e-10;
      if (WANT_LP) s += am1 * log(xm) - beta * x - lognorm;
      if (WANT_G) {
        const double dmax = x > 1e-10 ? 1.0 : (x == 1e-10 ? 0.5 : 0.0);
        g[i] = dmax * (am1 / xm) - beta;
      }
    }
    valid &= __shfl_xor_sync(mask, valid, 1);
    valid &= __shfl_xor_sync(mask, valid, 2);
    if (WANT_G && !valid) {
#pragma unroll
      for (int i = 0; i < E; ++i) g[i] = 0.0;
    }
    if (WANT_LP) {
      s = quad_sum(mask, s);
      lp = valid ? s : -INFINITY;
    }
  } else if (TID == MCMC_TARGET_STUDENT_T) {
    const double df = t.par[0], lognorm = t.par[1], hk = t.c[0];
    const double dfp1 = df + 1.0;
    double s = 0.0;
#pragma unroll
    for (int i = 0; i < E; ++i) {
      const double x = q[i];
      const double xx = x * x;
      if (WANT_LP) s += lognorm + (-hk) * log(1.0 + xx / df);
      if (WANT_G) g[i] = -dfp1 * x / (df + xx);
    }
    if (WANT_LP) lp = quad_sum(mask, s);
  } else if (TID == MCMC_TARGET_ROSENBROCK) {
    // t_d = x_{d+1} - x_d^2 (d < 19);  g_d = x_d (4a t_d - 2) + 2 - 2a t_{d-1}  (targets.cuh): the lane's last term
    // needs the next lane's first coordinate, its first gradient the previous lane's last term
    const double a = t.c[0];
    const double a4 = 4.0 * a, am2 = -2.0 * a;
    const double xnext = __shfl_down_sync(mask, q[0], 1, 4);      // part 3: own value back, unused
    double tt[E];
#pragma unroll
    for (int i = 0; i < E; ++i) tt[i] = fma(-q[i], q[i], (i + 1 < E) ? q[i + 1] : xnext);
    const bool last = part == 3;                                   // d = 19 has no term of its own
    double tprev = __shfl_up_sync(mask, tt[E - 1], 1, 4);          // part 0: own value back, unused
    double U = 0.0;
#pragma unroll
    for (int i = 0; i < E; ++i) {
      const bool has_t = !(last && i == E - 1);
      if (WANT_LP && has_t) {
        const double om = 1.0 - q[i];
        U = fma(om, om, U);
        U = fma(a * tt[i], tt[i], U);
      }
      if (WANT_G) {
        const double hi = has_t ? fma(q[i], fma(a4, tt[i], -2.0), 2.0) : 0.0;
        const bool has_prev = !(part == 0 && i == 0);
        g[i] = has_prev ? fma(am2, (i == 0) ? tprev : tt[i - 1], hi) : hi;
      }
    }
    if (WANT_LP) lp = -quad_sum(mask, U);
  } else if (TID == MCMC_TARGET_GAUSSIAN_MIXTURE) {
    const double hs = t.c[0];
    const double x0 = quad_get(mask, q[0], 0);
    const double dp = x0 + hs, dm = x0 - hs;
    const double m1 = -0.5 * (dp * dp), m2 = -0.5 * (dm * dm);
    const double mx = m1 > m2 ? m1 : m2;
    const double e1 = exp(m1 - mx), e2 = exp(m2 - mx);
    const double es = e1 + e2;
    double s = 0.0;
#pragma unroll
    for (int i = 0; i < E; ++i) {
      if (WANT_LP) s += (part == 0 && i == 0) ? 0.0 : q[i] * q[i];
      if (WANT_G) g[i] = -q[i];
    }
    if (WANT_G && part == 0) g[0] = (e1 / es) * (-dp) + (e2 / es) * (-dm);
    if (WANT_LP) {
      s = quad_sum(mask, s);
      const double lp0 = t.c[1] + mx + log(es) - t.c[2];
      lp = lp0 + -0.5 * (s + t.c[3]);
    }
  } else if (TID == MCMC_TARGET_DENSE_GAUSSIAN) {
    // y_j = sum_i x_i P[i][j] and (P x)_j for the lane's five j: the other lanes' coordinates arrive by shuffles
    double acc[E], acct[E];
#pragma unroll
    for (int j = 0; j < E; ++j) acc[j] = acct[j] = 0.0;
#pragma unroll
    for (int pp = 0; pp < 4; ++pp) {
#pragma unroll
      for (int i = 0; i < E; ++i) {
        const double xi = quad_get(mask, q[i], pp);
        const int gi = pp * E + i;
#pragma unroll
        for (int j = 0; j < E; ++j) {
          acc[j] += xi * tab[gi * n + d0 + j];
          acct[j] += tab[(d0 + j) * n + gi] * xi;
        }
      }
    }
    double maha = 0.0;
#pragma unroll
    for (int j = 0; j < E; ++j) {
      if (WANT_LP) maha += acc[j] * q[j];
      if (WANT_G) g[j] = -0.5 * (acc[j] + acct[j]);
    }
    if (WANT_LP) lp = -0.5 * (quad_sum(mask, maha) + t.c[0] + t.c[1]);
  }
  return lp;
}

}  // namespace mcmc
