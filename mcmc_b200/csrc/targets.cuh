// targets.cuh — analytic log-density and gradient of every benchmark target,
// fused into the sampler kernels (one thread evaluates one chain's D-vector).
//
// Replaces vmap(jax.value_and_grad(log_prob_fn)) (HMC.py:123, GRAHMC.py:203,
// NUTS.py:111, RWMH.py:102) over the closures of benchmarks/targets.py:
//   standard_normal :50-52   correlated_gaussian :79-98   ill_conditioned :125-137
//   neals_funnel :166-185    log_gamma :236-251           student_t :300-311
//   rosenbrock :367-387      gaussian_mixture :432-454
// The expressions are written in the reference's operation order; the strict
// build (-DMCMC_FAITHFUL -fmad=false) evaluates them one rounding per op, the
// fast build lets nvcc contract them to DFMA and replaces divisions by a
// reciprocal multiply where marked.
#pragma once
#include "common.cuh"

// A user target (mcmc_b200/custom.py): the header defines, in the global namespace,
//   __device__ double mcmc_user_log_prob_grad(const double* q, double* g, int n, const double* par);
// returning log p(q) and writing grad log p(q) to g[0..n) (par = McmcParams.target_params).
#define MCMC_TARGET_USER_TID 9
#ifdef MCMC_USER_TARGET_HEADER
#include MCMC_USER_TARGET_HEADER
#endif

namespace mcmc {

// Host-derived scalar constants of one target (kernel parameter, constant bank).
struct TargetCtx {
  double par[8];
  double c[6];
  const double* cov_inv;  // device, DENSE only
};

// filled by the launcher (host); D-dependent constants folded here
inline TargetCtx make_target_ctx(const McmcParams& p, const double* cov_inv) {
  TargetCtx t{};
  for (int i = 0; i < 8; ++i) t.par[i] = p.target_params[i];
  t.cov_inv = cov_inv;
  const int D = p.n_dim;
  switch (p.target_id) {
    case MCMC_TARGET_STANDARD_NORMAL:
      t.c[0] = (p.target_params[0] != 0.0) ? D * kLog2Pi : 0.0;
      break;
    case MCMC_TARGET_CORRELATED_GAUSSIAN: {
      const double rho = p.target_params[0];
      t.c[0] = 1.0 / (1.0 - rho);                                        // a
      t.c[1] = -rho / ((1.0 - rho) * (1.0 + (D - 1) * rho));             // b
      t.c[2] = (D - 1) * log(1.0 - rho) + log(1.0 + (D - 1) * rho);      // logdet
      t.c[3] = D * kLog2Pi;
      break;
    }
    case MCMC_TARGET_ILL_CONDITIONED: {
      const double kappa = p.target_params[0];
      double logdet = 0.0;
      for (int d = 0; d < D; ++d) {
        const double tt = (D > 1) ? (double)d / (double)(D - 1) : 0.0;
        double lam = 1.0 * (1.0 - tt) + kappa * tt;
        if (d == D - 1 && D > 1) lam = kappa;
        logdet += log(lam);
      }
      t.c[0] = logdet;
      t.c[1] = D * kLog2Pi;
      break;
    }
    case MCMC_TARGET_NEALS_FUNNEL:
      t.c[0] = log(2.0 * kPi * 9.0);
      t.c[1] = (D - 1) * kLog2Pi;
      t.c[2] = (double)(D - 1);
      break;
    case MCMC_TARGET_STUDENT_T:
      t.c[0] = (p.target_params[0] + 1.0) / 2.0;
      break;
    case MCMC_TARGET_ROSENBROCK:
      t.c[0] = 1.0 / (p.target_params[0] * p.target_params[0]);  // a
      break;
    case MCMC_TARGET_GAUSSIAN_MIXTURE:
      t.c[0] = p.target_params[0] / 2.0;
      t.c[1] = log(0.5);
      t.c[2] = 0.5 * kLog2Pi;
      t.c[3] = (D - 1) * kLog2Pi;
      break;
    case MCMC_TARGET_DENSE_GAUSSIAN:
      t.c[0] = p.target_params[0];
      t.c[1] = D * kLog2Pi;
      break;
    default: break;
  }
  return t;
}

// doubles of per-block shared-memory table a target needs (ill-conditioned: 1/lambda_d; dense: P)
__host__ __device__ inline int target_table_len(int tid, int D) {
  return tid == MCMC_TARGET_ILL_CONDITIONED ? D : (tid == MCMC_TARGET_DENSE_GAUSSIAN ? D * D : 0);
}

// cooperative fill of the shared table; call before a __syncthreads()
template <int TID>
__device__ __forceinline__ void target_table_fill(const TargetCtx& t, int D, double* tab) {
  if (TID == MCMC_TARGET_ILL_CONDITIONED) {
    for (int d = threadIdx.x; d < D; d += blockDim.x) {
      const double tt = (D > 1) ? (double)d / (double)(D - 1) : 0.0;
      double lam = 1.0 * (1.0 - tt) + t.par[0] * tt;
      if (d == D - 1 && D > 1) lam = t.par[0];
      tab[d] = 1.0 / lam;
    }
  } else if (TID == MCMC_TARGET_DENSE_GAUSSIAN) {
    for (int i = threadIdx.x; i < D * D; i += blockDim.x) tab[i] = t.cov_inv[i];
  }
}

// log p(q) (returned when WANT_LP, else 0) and, when WANT_G, g = grad log p(q).
// DS > 0: compile-time dimension (arrays live in registers, loops fully unrolled);
// DS == 0: runtime dimension n (arrays in local memory).
template <int TID, int DS, bool WANT_LP, bool WANT_G>
__device__ __forceinline__ double target_eval(const TargetCtx& t, const int n_rt, const double* __restrict__ q,
                                              double* __restrict__ g, const double* __restrict__ tab) {
  const int n = DS > 0 ? DS : n_rt;
  double lp = 0.0;
  if (TID == MCMC_TARGET_STANDARD_NORMAL) {
    double sp[kSumWays] = {0.0};
#pragma unroll
    for (int d = 0; d < n; ++d) {
      if (WANT_LP) sp[d % kSumWays] += q[d] * q[d];
      if (WANT_G) g[d] = -q[d];
    }
    if (WANT_LP) lp = -0.5 * (sum_ways(sp) + t.c[0]);
  } else if (TID == MCMC_TARGET_CORRELATED_GAUSSIAN) {
    // P = a I + b J:  (x P)_j = a x_j + b sum(x); the reference forms the dense product
    // (targets.py:94), whose row sums reduce to this in exact arithmetic.
    const double a = t.c[0], b = t.c[1];
    double sx = 0.0;
#pragma unroll
    for (int d = 0; d < n; ++d) sx += q[d];
    const double bs = b * sx;
    double maha = 0.0;
#pragma unroll
    for (int d = 0; d < n; ++d) {
      const double y = a * q[d] + bs;
      if (WANT_LP) maha += y * q[d];
      if (WANT_G) g[d] = -y;
    }
    if (WANT_LP) lp = -0.5 * (maha + t.c[2] + t.c[3]);
  } else if (TID == MCMC_TARGET_ILL_CONDITIONED) {
    double maha = 0.0;
#pragma unroll
    for (int d = 0; d < n; ++d) {
      const double y = q[d] * tab[d];
      if (WANT_LP) maha += y * q[d];
      if (WANT_G) g[d] = -y;
    }
    if (WANT_LP) lp = -0.5 * (maha + t.c[0] + t.c[1]);
  } else if (TID == MCMC_TARGET_NEALS_FUNNEL) {
    const double x0 = q[0];
    double sp[kSumWays] = {0.0};
#pragma unroll
    for (int d = 1; d < n; ++d) sp[d % kSumWays] += q[d] * q[d];
    const double s = sum_ways(sp);
#if MCMC_IS_FAITHFUL
    const double var = exp(x0);
    const double sv = s / var;
    if (WANT_G) {
#pragma unroll
      for (int d = 1; d < n; ++d) g[d] = -(q[d] / var);
    }
#else
    const double iv = exp(-x0);
    const double sv = s * iv;
    if (WANT_G) {
      const double niv = -iv;
#pragma unroll
      for (int d = 1; d < n; ++d) g[d] = q[d] * niv;
    }
#endif
#if MCMC_IS_FAITHFUL
    const double x0v = x0 / 9.0;                 // targets.py:172-175: x[0] / 3**2
    const double x0x0v = x0 * x0 / 9.0;
#else
    // the two divisions by the constant variance were 16 dependent DFMAs per evaluation (reciprocal iteration + correction)
    const double x0v = x0 * (1.0 / 9.0);
    const double x0x0v = x0 * x0v;
#endif
    if (WANT_G) g[0] = -x0v + 0.5 * sv - 0.5 * t.c[2];
    if (WANT_LP) {
      const double lp0 = -0.5 * (x0x0v + t.c[0]);
      const double lpr = -0.5 * (sv + t.c[2] * x0 + t.c[1]);
      lp = lp0 + lpr;
    }
  } else if (TID == MCMC_TARGET_LOG_GAMMA) {
    const double am1 = t.par[0] - 1.0, beta = t.par[1], lognorm = t.par[2];
    bool valid = true;
    double s = 0.0;
#pragma unroll
    for (int d = 0; d < n; ++d) {
      const double x = q[d];
      valid = valid && (x > 0.0);
      const double xm = x > 1e-10 ? x : 1e-10;
      if (WANT_LP) s += am1 * log(xm) - beta * x - lognorm;
      if (WANT_G) {
        const double dmax = x > 1e-10 ? 1.0 : (x == 1e-10 ? 0.5 : 0.0);
        g[d] = dmax * (am1 / xm) - beta;
      }
    }
    if (WANT_G && !valid) {
#pragma unroll
      for (int d = 0; d < n; ++d) g[d] = 0.0;
    }
    if (WANT_LP) lp = valid ? s : -INFINITY;
  } else if (TID == MCMC_TARGET_STUDENT_T) {
    const double df = t.par[0], lognorm = t.par[1], hk = t.c[0];
    const double dfp1 = df + 1.0;
    double s = 0.0;
#pragma unroll
    for (int d = 0; d < n; ++d) {
      const double x = q[d];
      const double xx = x * x;
      if (WANT_LP) s += lognorm + (-hk) * log(1.0 + xx / df);
      if (WANT_G) g[d] = -dfp1 * x / (df + xx);
    }
    if (WANT_LP) lp = s;
  } else if (TID == MCMC_TARGET_ROSENBROCK) {
    const double a = t.c[0];
    const double a4 = 4.0 * a, am2 = -2.0 * a;
    double U = 0.0;
#if MCMC_IS_FAITHFUL
    double carry = 0.0;  // -2a t_{i-1}, the contribution of term i-1 to g_i
#pragma unroll
    for (int i = 0; i < n - 1; ++i) {
      const double om = 1.0 - q[i];
      const double ti = q[i + 1] - q[i] * q[i];
      if (WANT_LP) U += om * om + a * (ti * ti);
      if (WANT_G) {
        const double gi = 2.0 * om + a4 * q[i] * ti;
        g[i] = (i == 0) ? (0.0 + gi) : (carry + gi);
        carry = am2 * ti;
      }
    }
    if (WANT_G) g[n - 1] = (n > 1) ? (0.0 + carry) : 0.0;
#else
    // 4 DFMA per coupling term: with t_i = x_{i+1} - x_i^2,
    //   g_i = 2(1 - x_i) + 4a x_i t_i - 2a t_{i-1} = x_i (4a t_i - 2) + 2 - 2a t_{i-1}
    double tprev = 0.0;
    double Up[4] = {0.0, 0.0, 0.0, 0.0};  // four interleaved partial sums: the value's dependency chain is 4x shorter
#pragma unroll
    for (int i = 0; i < n - 1; ++i) {
      const double ti = fma(-q[i], q[i], q[i + 1]);
      if (WANT_LP) {
        const double om = 1.0 - q[i];
#ifdef MCMC_ROSEN_SERIAL_SUM
        Up[0] = fma(om, om, Up[0]);
        Up[0] = fma(a * ti, ti, Up[0]);
#else
        Up[i & 3] = fma(om, om, Up[i & 3]);
        Up[(i + 2) & 3] = fma(a * ti, ti, Up[(i + 2) & 3]);
#endif
      }
      if (WANT_G) {
        const double hi = fma(q[i], fma(a4, ti, -2.0), 2.0);
        g[i] = (i == 0) ? hi : fma(am2, tprev, hi);
        tprev = ti;
      }
    }
    if (WANT_G) g[n - 1] = (n > 1) ? am2 * tprev : 0.0;
    if (WANT_LP) U = (Up[0] + Up[1]) + (Up[2] + Up[3]);
#endif
    if (WANT_LP) lp = -U;
  } else if (TID == MCMC_TARGET_GAUSSIAN_MIXTURE) {
    const double hs = t.c[0];
    const double x0 = q[0];
    const double dp = x0 + hs, dm = x0 - hs;
    const double m1 = -0.5 * (dp * dp), m2 = -0.5 * (dm * dm);
    const double mx = m1 > m2 ? m1 : m2;
    const double e1 = exp(m1 - mx), e2 = exp(m2 - mx);
    const double es = e1 + e2;
    double s = 0.0;
#pragma unroll
    for (int d = 1; d < n; ++d) {
      if (WANT_LP) s += q[d] * q[d];
      if (WANT_G) g[d] = -q[d];
    }
    if (WANT_G) g[0] = (e1 / es) * (-dp) + (e2 / es) * (-dm);
    if (WANT_LP) {
      const double lp0 = t.c[1] + mx + log(es) - t.c[2];
      lp = lp0 + -0.5 * (s + t.c[3]);
    }
#ifdef MCMC_USER_TARGET_HEADER
  } else if (TID == MCMC_TARGET_USER_TID) {
    if (WANT_G) {
      lp = mcmc_user_log_prob_grad(q, g, n, t.par);
    } else {
      double gl[DS > 0 ? DS : MCMC_DYN_MAXD];   // value-only callers (RWMH): the gradient stores are dead code
      lp = mcmc_user_log_prob_grad(q, gl, n, t.par);
    }
#endif
  } else if (TID == MCMC_TARGET_DENSE_GAUSSIAN) {
    // y = x P and yt = P x from the shared copy of P (broadcast reads); maha = y.x
    double maha = 0.0;
    for (int j = 0; j < n; ++j) {
      double acc = 0.0, acct = 0.0;
#pragma unroll
      for (int i = 0; i < n; ++i) {
        acc += q[i] * tab[i * n + j];
        acct += tab[j * n + i] * q[i];
      }
      if (WANT_LP) maha += acc * q[j];
      if (WANT_G) g[j] = -0.5 * (acc + acct);
    }
    if (WANT_LP) lp = -0.5 * (maha + t.c[0] + t.c[1]);
  }
  return lp;
}

}  // namespace mcmc
