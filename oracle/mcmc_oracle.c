/*
 * mcmc_oracle.c — CPU restatement of the reference's per-chain sampling loop.
 *
 * TEST INFRASTRUCTURE.  Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may load this; the product path
 * (mcmc_b200/) never does.
 *
 * PARITY: pinned against the reference's own SOURCE, not against jax itself.  The
 * reference ships no tests, golden vectors or fixtures and needs
 * jax==0.8.0, which is not installable here (F3).  tests/golden/make_reference_golden.py
 * therefore imports the reference's files unmodified from /root/reference on top of
 * a torch-fp64 stand-in for the jax API (tests/golden/jax_shim: eager IEEE double
 * arithmetic in source order, vmap = loop, value_and_grad = autograd) and records
 * inputs, draws and outputs of 24 hmc/grahmc/rwmh/nuts runs in
 * tests/golden/reference_golden.npz; this restatement reproduces them within 5e-13
 * relative with identical accept counts and tree depths (tests/test_reference_golden_cpu.py).
 * NOT pinned: XLA's own exp/log/tanh roundings and jax's threefry streams
 * ("parity unpinned" in that narrower sense).  Also pinned on (a) known-answer
 * facts derived from the source, (b) an independent NumPy
 * restatement (oracle/oracle_np.py) and (c) torch fp64 autograd of each log-density
 * as coded, for the gradients.
 *
 * It follows the reference operation by operation (compile with
 * -ffp-contract=off so no FMA is formed, as in the reference's jnp expressions):
 *   hmc      samplers/HMC.py:86-134 (leapfrog), :138-219 (hmc_step), :259-314 (hmc_run)
 *   grahmc   samplers/GRAHMC.py:69-106 (schedules), :152-211, :215-261, :265-349, :398-456
 *   rwmh     samplers/RWMH.py:73-119, :148-178
 *   nuts     samplers/NUTS.py:90-198, :202-396, :399-442, :479-514
 *   targets  benchmarks/targets.py:50,90,130,166,236,300,367,432
 *   welford  tuning/welford.py:42-59, :86-100
 * Randomness is either injected (the caller passes z/u arrays) or the native
 * Philox stream of oracle/rng_common.h.
 *
 * One chain per OpenMP loop iteration; chains are independent (the reference
 * vmaps over them, HMC.py:123), so the (iteration, chain) loops commute.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#include "../include/mcmc_b200.h"
#include "rng_common.h"

#define ORACLE_MAXD MCMC_MAX_DIM

/* ------------------------------------------------------------------ targets */

typedef struct {
  int id;
  int D;
  const double* par;     /* params->target_params */
  const double* cov_inv; /* dense Gaussian */
} Target;

static const double LOG_2PI = 1.8378770664093454836; /* log(2*pi) */

/* jnp.minimum propagates NaN (C fmin does not) */
static inline double min_nan(double a, double b) {
  if (a != a) return a;
  if (b != b) return b;
  return a < b ? a : b;
}

/* log p(x) and (if g != NULL) its gradient; formulas: SURVEY.md §2b table */
static double target_eval(const Target* t, const double* x, double* g) {
  const int D = t->D;
  switch (t->id) {
    case MCMC_TARGET_STANDARD_NORMAL: { /* targets.py:50-52 */
      double s = 0.0;
      for (int d = 0; d < D; ++d) s += x[d] * x[d];
      if (g) for (int d = 0; d < D; ++d) g[d] = -x[d];
      if (t->par[0] != 0.0) return -0.5 * (s + D * LOG_2PI);
      return -0.5 * s; /* README.md:51-52 */
    }
    case MCMC_TARGET_CORRELATED_GAUSSIAN: { /* targets.py:79-98: dense x @ cov_inv @ x */
      const double rho = t->par[0];
      const double a = 1.0 / (1.0 - rho);
      const double b = -rho / ((1.0 - rho) * (1.0 + (D - 1) * rho));
      const double logdet = (D - 1) * log(1.0 - rho) + log(1.0 + (D - 1) * rho);
      double y[ORACLE_MAXD];
      double maha = 0.0;
      for (int j = 0; j < D; ++j) { /* y = x @ cov_inv */
        double acc = 0.0;
        for (int i = 0; i < D; ++i) acc += x[i] * (i == j ? a + b : b);
        y[j] = acc;
      }
      for (int j = 0; j < D; ++j) maha += y[j] * x[j];
      if (g) for (int j = 0; j < D; ++j) g[j] = -y[j]; /* P symmetric */
      return -0.5 * (maha + logdet + D * LOG_2PI);
    }
    case MCMC_TARGET_ILL_CONDITIONED: { /* targets.py:125-137 */
      const double kappa = t->par[0];
      double maha = 0.0, logdet = 0.0;
      for (int d = 0; d < D; ++d) {
        /* jnp.linspace(1, kappa, D): start*(1-t) + stop*t, t = d/(D-1), endpoint appended exactly */
        const double tt = (D > 1) ? (double)d / (double)(D - 1) : 0.0;
        double lam = 1.0 * (1.0 - tt) + kappa * tt;
        if (d == D - 1 && D > 1) lam = kappa;
        double inv = 1.0 / lam;
        maha += (x[d] * inv) * x[d];
        logdet += log(lam);
        if (g) g[d] = -(x[d] * inv);
      }
      return -0.5 * (maha + logdet + D * LOG_2PI);
    }
    case MCMC_TARGET_NEALS_FUNNEL: { /* targets.py:166-185 */
      const double x0 = x[0];
      double s = 0.0;
      for (int d = 1; d < D; ++d) s += x[d] * x[d];
      const double var = exp(x0);
      const int Dr = D - 1;
      const double lp0 = -0.5 * (x0 * x0 / 9.0 + log(2.0 * M_PI * 9.0));
      const double lpr = -0.5 * (s / var + Dr * x0 + Dr * LOG_2PI);
      if (g) {
        g[0] = -x0 / 9.0 + 0.5 * (s / var) - 0.5 * Dr;
        for (int d = 1; d < D; ++d) g[d] = -(x[d] / var);
      }
      return lp0 + lpr;
    }
    case MCMC_TARGET_LOG_GAMMA: { /* targets.py:236-251 */
      const double alpha = t->par[0], beta = t->par[1], lognorm = t->par[2];
      int valid = 1;
      double s = 0.0;
      for (int d = 0; d < D; ++d) {
        if (!(x[d] > 0.0)) valid = 0;
        double xm = x[d] > 1e-10 ? x[d] : 1e-10; /* jnp.maximum(x, 1e-10); NaN handled by valid */
        s += (alpha - 1.0) * log(xm) - beta * x[d] - lognorm;
      }
      if (g) {
        for (int d = 0; d < D; ++d) {
          if (!valid) { g[d] = 0.0; continue; } /* where(valid, ., -inf) kills the cotangent */
          double dmax = x[d] > 1e-10 ? 1.0 : (x[d] == 1e-10 ? 0.5 : 0.0);
          double xm = x[d] > 1e-10 ? x[d] : 1e-10;
          g[d] = dmax * ((alpha - 1.0) / xm) - beta;
        }
      }
      return valid ? s : -INFINITY;
    }
    case MCMC_TARGET_STUDENT_T: { /* targets.py:300-311 */
      const double df = t->par[0], lognorm = t->par[1];
      const double hk = (df + 1.0) / 2.0;
      double s = 0.0;
      for (int d = 0; d < D; ++d) {
        s += lognorm + (-hk) * log(1.0 + x[d] * x[d] / df);
        if (g) g[d] = -(df + 1.0) * x[d] / (df + x[d] * x[d]);
      }
      return s;
    }
    case MCMC_TARGET_ROSENBROCK: { /* targets.py:367-387 */
      const double sc = t->par[0];
      const double a = 1.0 / (sc * sc);
      double U = 0.0;
      if (g) for (int d = 0; d < D; ++d) g[d] = 0.0;
      for (int i = 0; i + 1 < D; ++i) {
        const double om = 1.0 - x[i];
        const double ti = x[i + 1] - x[i] * x[i];
        U += om * om + a * (ti * ti);
        if (g) {
          g[i] += 2.0 * om + 4.0 * a * x[i] * ti; /* -dU/dx_i */
          g[i + 1] += -2.0 * a * ti;
        }
      }
      return -U;
    }
    case MCMC_TARGET_GAUSSIAN_MIXTURE: { /* targets.py:432-454 */
      const double sep = t->par[0];
      const double x0 = x[0];
      const double m1 = -0.5 * ((x0 + sep / 2.0) * (x0 + sep / 2.0));
      const double m2 = -0.5 * ((x0 - sep / 2.0) * (x0 - sep / 2.0));
      const double mx = m1 > m2 ? m1 : m2;
      const double e1 = exp(m1 - mx), e2 = exp(m2 - mx);
      const double lp0 = log(0.5) + mx + log(e1 + e2) - 0.5 * LOG_2PI;
      double s = 0.0;
      for (int d = 1; d < D; ++d) s += x[d] * x[d];
      if (g) {
        const double w1 = e1 / (e1 + e2), w2 = e2 / (e1 + e2);
        g[0] = w1 * (-(x0 + sep / 2.0)) + w2 * (-(x0 - sep / 2.0));
        for (int d = 1; d < D; ++d) g[d] = -x[d];
      }
      return lp0 + -0.5 * (s + (D - 1) * LOG_2PI);
    }
    case MCMC_TARGET_DENSE_GAUSSIAN: { /* x @ P @ x with a general P, same shape as targets.py:94 */
      const double* P = t->cov_inv;
      double y[ORACLE_MAXD], yt[ORACLE_MAXD];
      double maha = 0.0;
      for (int j = 0; j < D; ++j) {
        double acc = 0.0, acct = 0.0;
        for (int i = 0; i < D; ++i) { acc += x[i] * P[i * D + j]; acct += P[j * D + i] * x[i]; }
        y[j] = acc; yt[j] = acct;
      }
      for (int j = 0; j < D; ++j) maha += y[j] * x[j];
      if (g) for (int j = 0; j < D; ++j) g[j] = -0.5 * (y[j] + yt[j]);
      return -0.5 * (maha + t->par[0] + D * LOG_2PI);
    }
  }
  return NAN;
}

/* -------------------------------------------------------------- randomness */

typedef struct {
  const McmcParams* p;
  const McmcBuffers* b;
  int64_t c; /* local chain */
} Rng;

static void draw_normals(const Rng* r, int64_t it, double* z) {
  const int D = r->p->n_dim;
  const int64_t C = r->p->n_chains;
  if (r->p->rng_mode == MCMC_RNG_INJECTED) {
    memcpy(z, r->b->z_inject + ((size_t)it * C + r->c) * D, sizeof(double) * D);
    return;
  }
  const uint64_t gid = (uint64_t)(r->p->chain_offset + r->c);
  for (int j = 0; 2 * j < D; ++j) {
    uint32_t ctr[4] = {(uint32_t)gid, (uint32_t)(gid >> 32), (uint32_t)it, (MCMC_TAG_NORMAL << 24) | (uint32_t)j};
    uint32_t o[4];
    philox4x32_10(ctr, r->p->seed, o);
    const double u1 = (double)(mcmc_u53(o[0], o[1]) + 1) * MCMC_TWO_M53;
    const double u2 = (double)mcmc_u53(o[2], o[3]) * MCMC_TWO_M53;
    const double rad = sqrt(-2.0 * log(u1));
    const double ang = 2.0 * M_PI * u2;
    z[2 * j] = rad * cos(ang);
    if (2 * j + 1 < D) z[2 * j + 1] = rad * sin(ang);
  }
}

static double draw_uniform(const Rng* r, int64_t it) {
  if (r->p->rng_mode == MCMC_RNG_INJECTED) return r->b->u_inject[(size_t)it * r->p->n_chains + r->c];
  const uint64_t gid = (uint64_t)(r->p->chain_offset + r->c);
  uint32_t ctr[4] = {(uint32_t)gid, (uint32_t)(gid >> 32), (uint32_t)it, (MCMC_TAG_ACCEPT << 24)};
  uint32_t o[4];
  philox4x32_10(ctr, r->p->seed, o);
  return (double)mcmc_u53(o[0], o[1]) * MCMC_TWO_M53;
}

static void draw_nuts(const Rng* r, int64_t it, int depth, double* u_dir, double* u_prop) {
  if (r->p->rng_mode == MCMC_RNG_INJECTED) {
    const size_t idx = ((size_t)it * r->p->n_chains + r->c) * r->p->max_tree_depth + depth;
    *u_dir = r->b->nuts_dir_u[idx];
    *u_prop = r->b->nuts_prop_u[idx];
    return;
  }
  const uint64_t gid = (uint64_t)(r->p->chain_offset + r->c);
  uint32_t ctr[4] = {(uint32_t)gid, (uint32_t)(gid >> 32), (uint32_t)it, (MCMC_TAG_NUTS << 24) | (uint32_t)depth};
  uint32_t o[4];
  philox4x32_10(ctr, r->p->seed, o);
  *u_dir = (double)mcmc_u53(o[0], o[1]) * MCMC_TWO_M53;
  *u_prop = (double)mcmc_u53(o[2], o[3]) * MCMC_TWO_M53;
}

/* ------------------------------------------------------- friction schedule */

/* samplers/GRAHMC.py:69-106; t = i*eps, T = eps*L (:246,251) */
static double friction_gamma(int sched, double t, double T, double gam, double steep) {
  switch (sched) {
    case MCMC_SCHED_CONSTANT: return (t < T / 2) ? -gam : gam;
    case MCMC_SCHED_TANH: return gam * tanh(steep * (2.0 * t / T - 1.0));
    case MCMC_SCHED_SIGMOID: return gam * (2.0 / (1.0 + exp(-(steep * (t / T - 0.5)))) - 1.0);
    case MCMC_SCHED_LINEAR: return -gam + (2.0 * gam * t / T);
    case MCMC_SCHED_SINE: return gam * sin(M_PI * (t / T - 0.5));
  }
  return NAN;
}

/* s_i = exp(-gamma_i * (0.5*eps))  (GRAHMC.py:193-194) */
void oracle_friction_table(int sched, int L, double eps, double gam, double steep, double* s_out,
                           double* gamma_out) {
  const double T = eps * L;
  const double half_eps = 0.5 * eps;
  for (int i = 0; i < L; ++i) {
    const double gi = friction_gamma(sched, i * eps, T, gam, steep);
    if (gamma_out) gamma_out[i] = gi;
    s_out[i] = exp(-gi * half_eps);
  }
}

static double kinetic(const double* p, const double* w, int D) { /* HMC.py:176 */
  double s = 0.0;
  for (int d = 0; d < D; ++d) s += (p[d] * p[d]) * w[d];
  return 0.5 * s;
}

static void make_target(const McmcParams* p, const McmcBuffers* b, Target* t) {
  t->id = p->target_id;
  t->D = p->n_dim;
  t->par = p->target_params;
  t->cov_inv = b ? b->cov_inv : NULL;
}

/* tuning/welford.py:42-59 */
static void welford_fold(double* mean, double* m2, double count_new, const double* x, int D) {
  for (int d = 0; d < D; ++d) {
    const double delta = x[d] - mean[d];
    mean[d] = mean[d] + delta / count_new;
    const double delta2 = x[d] - mean[d];
    m2[d] = m2[d] + delta * delta2;
  }
}

/* ----------------------------------------------------------- HMC / GRAHMC */

/* One chain through burn-in and sampling.  friction == 0: samplers/HMC.py;
   friction == 1: samplers/GRAHMC.py with the schedule table s[]. */
static void hmc_chain(const McmcParams* p, const McmcBuffers* b, int64_t c, int friction,
                      const double* s_tab, unsigned long long* n_accept_total) {
  const int D = p->n_dim, L = p->num_steps;
  const int64_t C = p->n_chains, S = p->num_samples, B = p->burn_in;
  const double eps = b->step_size_dev ? *b->step_size_dev : p->step_size;
  const double half_eps = 0.5 * eps; /* HMC.py:119 `half*step_sz`, GRAHMC.py:193 */
  Target tg;
  make_target(p, b, &tg);
  Rng rng = {p, b, c};
  double w[ORACLE_MAXD], sqw[ORACLE_MAXD];
  for (int d = 0; d < D; ++d) {
    w[d] = b->inv_mass ? b->inv_mass[d] : 1.0;
    sqw[d] = sqrt(w[d]);
  }
  double q[ORACLE_MAXD], g[ORACLE_MAXD], lp;
  double qn[ORACLE_MAXD], pn[ORACLE_MAXD], gn[ORACLE_MAXD], z[ORACLE_MAXD];
  memcpy(q, b->init_q + (size_t)c * D, sizeof(double) * D);
  lp = target_eval(&tg, q, g); /* hmc_init, HMC.py:73 */
  int32_t acc = 0;
  double wmean[ORACLE_MAXD], wm2[ORACLE_MAXD], wcount = b->welford_count;
  const int do_w = (p->flags & MCMC_FLAG_ACCUM_WELFORD) && b->welford_mean && b->welford_m2;
  if (do_w) {
    memcpy(wmean, b->welford_mean + (size_t)c * D, sizeof(double) * D);
    memcpy(wm2, b->welford_m2 + (size_t)c * D, sizeof(double) * D);
  }
  const int track = (p->flags & MCMC_FLAG_TRACK_PROPOSALS) != 0;

  for (int64_t it = 0; it < B + S; ++it) {
    if (it == B) acc = 0; /* HMC.py:276-282 */
    const int64_t si = it - B;
    draw_normals(&rng, it, z);
    for (int d = 0; d < D; ++d) pn[d] = z[d] / sqw[d]; /* HMC.py:172 */
    const double H0 = -lp + kinetic(pn, w, D);          /* HMC.py:176-177 */
    memcpy(qn, q, sizeof(double) * D);
    memcpy(gn, g, sizeof(double) * D);
    double lpn = lp;
    for (int i = 0; i < L; ++i) {
      if (friction) for (int d = 0; d < D; ++d) pn[d] = pn[d] * s_tab[i];  /* GRAHMC.py:197 */
      for (int d = 0; d < D; ++d) pn[d] = pn[d] + half_eps * gn[d];        /* HMC.py:119 */
      for (int d = 0; d < D; ++d) qn[d] = qn[d] + eps * (pn[d] * w[d]);    /* HMC.py:121 */
      lpn = target_eval(&tg, qn, gn);                                      /* HMC.py:123 */
      for (int d = 0; d < D; ++d) pn[d] = pn[d] + half_eps * gn[d];        /* HMC.py:127 */
      if (friction) for (int d = 0; d < D; ++d) pn[d] = pn[d] * s_tab[i];  /* GRAHMC.py:209 */
    }
    for (int d = 0; d < D; ++d) pn[d] = -pn[d];      /* HMC.py:194 */
    double H1 = -lpn + kinetic(pn, w, D);             /* HMC.py:196-197 */
    if (!isfinite(H1)) H1 = 1e10;                     /* HMC.py:199 */
    const double log_alpha = H0 - H1;
    const double dH = H1 - H0;
    const double u = draw_uniform(&rng, it);
    const int accept = log(u) < min_nan(0.0, log_alpha); /* HMC.py:207 */
    if (track && si >= 0) {
      const size_t o = (size_t)si * C + c;
      if (b->pre_q) memcpy(b->pre_q + o * D, q, sizeof(double) * D);
      if (b->pre_lp) b->pre_lp[o] = lp;
      if (b->prop_q) memcpy(b->prop_q + o * D, qn, sizeof(double) * D);
      if (b->prop_lp) b->prop_lp[o] = lpn;
      if (b->delta_h) b->delta_h[o] = dH;
    }
    if (accept) {
      memcpy(q, qn, sizeof(double) * D);
      memcpy(g, gn, sizeof(double) * D);
      lp = lpn;
      acc += 1;
    }
    if (si >= 0) {
      const size_t o = (size_t)si * C + c;
      if (b->samples && !(p->flags & MCMC_FLAG_NO_SAMPLES)) memcpy(b->samples + o * D, q, sizeof(double) * D);
      if (b->log_probs && !(p->flags & MCMC_FLAG_NO_SAMPLES)) b->log_probs[o] = lp;
      if (do_w) { wcount += 1.0; welford_fold(wmean, wm2, wcount, q, D); }
    }
  }
  if (S == 0) acc = 0; /* counters are reset after burn-in even when nothing is sampled (HMC.py:276-282) */
  if (b->accept_count) b->accept_count[c] = acc;
  if (b->accept_rate) b->accept_rate[c] = (float)acc / (float)S; /* HMC.py:313 */
  if (b->final_q) memcpy(b->final_q + (size_t)c * D, q, sizeof(double) * D);
  if (b->final_g) memcpy(b->final_g + (size_t)c * D, g, sizeof(double) * D);
  if (b->final_lp) b->final_lp[c] = lp;
  if (do_w) {
    memcpy(b->welford_mean + (size_t)c * D, wmean, sizeof(double) * D);
    memcpy(b->welford_m2 + (size_t)c * D, wm2, sizeof(double) * D);
  }
  *n_accept_total += (unsigned long long)acc;
}

/* The trajectory alone.  friction == 0: `leapfrog` (samplers/HMC.py:85-134: lf_step scanned num_steps times);
   friction != 0: `_trajectory_with_schedule` (samplers/GRAHMC.py:215-261) over `_conformal_leapfrog_step` (:152-211).
   Host pointers; lp may be NULL (taken as 0, only returned when num_steps == 0); outputs may be NULL. */
int oracle_leapfrog(const McmcParams* p, int friction, const double* cov_inv, const double* inv_mass, const double* q,
                    const double* mom, const double* lp, const double* grad, double* q_out, double* mom_out,
                    double* lp_out, double* grad_out) {
  if (!p || p->abi_version != MCMC_ABI_VERSION) return MCMC_EINVAL;
  if (p->n_dim < 1 || p->n_dim > ORACLE_MAXD || p->num_steps < 0) return MCMC_EINVAL;
  if (friction && (p->schedule_id < 0 || p->schedule_id >= MCMC_NUM_SCHEDULES)) return MCMC_EINVAL;
  const int D = p->n_dim, L = p->num_steps;
  const double eps = p->step_size, half_eps = 0.5 * eps; /* HMC.py:119 `half*step_sz`, GRAHMC.py:193 */
  double* s_tab = NULL;
  if (friction && L > 0) {
    s_tab = (double*)malloc(sizeof(double) * (size_t)L);
    if (!s_tab) return MCMC_ECUDA;
    oracle_friction_table(p->schedule_id, L, eps, p->gamma, p->steepness, s_tab, NULL);
  }
  McmcBuffers bb;
  memset(&bb, 0, sizeof(bb));
  bb.cov_inv = cov_inv;
  Target tg;
  make_target(p, &bb, &tg);
#pragma omp parallel for schedule(static)
  for (int64_t c = 0; c < p->n_chains; ++c) {
    double w[ORACLE_MAXD], qn[ORACLE_MAXD], pn[ORACLE_MAXD], gn[ORACLE_MAXD];
    for (int d = 0; d < D; ++d) w[d] = inv_mass ? inv_mass[d] : 1.0;
    memcpy(qn, q + (size_t)c * D, sizeof(double) * D);
    memcpy(pn, mom + (size_t)c * D, sizeof(double) * D);
    memcpy(gn, grad + (size_t)c * D, sizeof(double) * D);
    double lpn = lp ? lp[c] : 0.0;
    for (int i = 0; i < L; ++i) {
      if (friction) for (int d = 0; d < D; ++d) pn[d] = pn[d] * s_tab[i];  /* GRAHMC.py:197 */
      for (int d = 0; d < D; ++d) pn[d] = pn[d] + half_eps * gn[d];        /* HMC.py:119 */
      for (int d = 0; d < D; ++d) qn[d] = qn[d] + eps * (pn[d] * w[d]);    /* HMC.py:121 */
      lpn = target_eval(&tg, qn, gn);                                      /* HMC.py:123 */
      for (int d = 0; d < D; ++d) pn[d] = pn[d] + half_eps * gn[d];        /* HMC.py:127 */
      if (friction) for (int d = 0; d < D; ++d) pn[d] = pn[d] * s_tab[i];  /* GRAHMC.py:209 */
    }
    if (q_out) memcpy(q_out + (size_t)c * D, qn, sizeof(double) * D);
    if (mom_out) memcpy(mom_out + (size_t)c * D, pn, sizeof(double) * D);
    if (grad_out) memcpy(grad_out + (size_t)c * D, gn, sizeof(double) * D);
    if (lp_out) lp_out[c] = lpn;
  }
  free(s_tab);
  return MCMC_OK;
}

static int check_common(const McmcParams* p, const McmcBuffers* b) {
  if (!p || !b || p->abi_version != MCMC_ABI_VERSION) return MCMC_EINVAL;
  if (p->n_dim < 1 || p->n_dim > ORACLE_MAXD || p->n_chains < 0 || p->num_samples < 0 || p->burn_in < 0)
    return MCMC_EINVAL;
  if (p->target_id < 0 || p->target_id >= MCMC_NUM_TARGETS) return MCMC_EINVAL;
  if (!b->init_q && p->n_chains > 0) return MCMC_EINVAL;
  return MCMC_OK;
}

static int hmc_like_run(const McmcParams* p, const McmcBuffers* b, int friction) {
  int rc = check_common(p, b);
  if (rc) return rc;
  if (p->num_steps < 0) return MCMC_EINVAL;
  double* s_tab = NULL;
  if (friction) {
    if (p->schedule_id < 0 || p->schedule_id >= MCMC_NUM_SCHEDULES) return MCMC_EINVAL;
    s_tab = (double*)malloc(sizeof(double) * (p->num_steps > 0 ? p->num_steps : 1));
    oracle_friction_table(p->schedule_id, p->num_steps, b->step_size_dev ? *b->step_size_dev : p->step_size, p->gamma, p->steepness, s_tab, NULL);
  }
  unsigned long long tot = 0;
#pragma omp parallel for schedule(static) reduction(+ : tot)
  for (int64_t c = 0; c < p->n_chains; ++c) hmc_chain(p, b, c, friction, s_tab, &tot);
  if (b->counters) {
    b->counters[0] = tot;
    b->counters[1] = (unsigned long long)p->n_chains * (unsigned long long)(p->burn_in + p->num_samples) *
                     (unsigned long long)p->num_steps;
    b->counters[2] = 0;
    b->counters[3] = 0;
  }
  free(s_tab);
  return MCMC_OK;
}

int oracle_hmc_run(const McmcParams* p, const McmcBuffers* b) { return hmc_like_run(p, b, 0); }
int oracle_grahmc_run(const McmcParams* p, const McmcBuffers* b) { return hmc_like_run(p, b, 1); }

/* ------------------------------------------------------------------- RWMH */

static void rwmh_chain(const McmcParams* p, const McmcBuffers* b, int64_t c, unsigned long long* tot) {
  const int D = p->n_dim;
  const int64_t C = p->n_chains, S = p->num_samples, B = p->burn_in;
  const double scale = b->step_size_dev ? *b->step_size_dev : p->step_size;
  Target tg;
  make_target(p, b, &tg);
  Rng rng = {p, b, c};
  double q[ORACLE_MAXD], qn[ORACLE_MAXD], z[ORACLE_MAXD];
  memcpy(q, b->init_q + (size_t)c * D, sizeof(double) * D);
  double lp = target_eval(&tg, q, NULL); /* RWMH.py:67 */
  int32_t acc = 0;
  for (int64_t it = 0; it < B + S; ++it) {
    if (it == B) acc = 0;
    const int64_t si = it - B;
    draw_normals(&rng, it, z);
    for (int d = 0; d < D; ++d) qn[d] = q[d] + scale * z[d]; /* RWMH.py:99 */
    const double lpn = target_eval(&tg, qn, NULL);            /* RWMH.py:102 */
    const double log_ratio = lpn - lp;
    const double u = draw_uniform(&rng, it);
    const int accept = log(u) < min_nan(0.0, log_ratio);      /* RWMH.py:105-107 */
    if (accept) { memcpy(q, qn, sizeof(double) * D); lp = lpn; acc += 1; }
    if (si >= 0 && !(p->flags & MCMC_FLAG_NO_SAMPLES)) {
      const size_t o = (size_t)si * C + c;
      if (b->samples) memcpy(b->samples + o * D, q, sizeof(double) * D);
      if (b->log_probs) b->log_probs[o] = lp;
    }
  }
  if (S == 0) acc = 0; /* counters are reset after burn-in even when nothing is sampled (HMC.py:276-282) */
  if (b->accept_count) b->accept_count[c] = acc;
  if (b->accept_rate) b->accept_rate[c] = (float)acc / (float)S;
  if (b->final_q) memcpy(b->final_q + (size_t)c * D, q, sizeof(double) * D);
  if (b->final_lp) b->final_lp[c] = lp;
  *tot += (unsigned long long)acc;
}

int oracle_rwmh_run(const McmcParams* p, const McmcBuffers* b) {
  int rc = check_common(p, b);
  if (rc) return rc;
  unsigned long long tot = 0;
#pragma omp parallel for schedule(static) reduction(+ : tot)
  for (int64_t c = 0; c < p->n_chains; ++c) rwmh_chain(p, b, c, &tot);
  if (b->counters) { b->counters[0] = tot; b->counters[1] = 0; b->counters[2] = 0; b->counters[3] = 0; }
  return MCMC_OK;
}

/* ------------------------------------------------------------------- NUTS */

static double dotv(const double* a, const double* b, int D) {
  double s = 0.0;
  for (int d = 0; d < D; ++d) s += a[d] * b[d];
  return s;
}

/* samplers/NUTS.py:202-396 for one chain, one transition */
static void nuts_chain(const McmcParams* p, const McmcBuffers* b, int64_t c, unsigned long long cnt[4]) {
  const int D = p->n_dim;
  const int64_t C = p->n_chains, S = p->num_samples, B = p->burn_in;
  const double eps = b->step_size_dev ? *b->step_size_dev : p->step_size;
  const int max_depth = p->max_tree_depth;
  Target tg;
  make_target(p, b, &tg);
  Rng rng = {p, b, c};
  double w[ORACLE_MAXD], sqw[ORACLE_MAXD];
  for (int d = 0; d < D; ++d) { w[d] = b->inv_mass ? b->inv_mass[d] : 1.0; sqw[d] = sqrt(w[d]); }
  double q[ORACLE_MAXD], g[ORACLE_MAXD], lp, z[ORACLE_MAXD];
  memcpy(q, b->init_q + (size_t)c * D, sizeof(double) * D);
  lp = target_eval(&tg, q, g); /* nuts_init, NUTS.py:77 */
  int32_t acc = 0;
  double wmean[ORACLE_MAXD], wm2[ORACLE_MAXD], wcount = b->welford_count;
  const int do_w = (p->flags & MCMC_FLAG_ACCUM_WELFORD) && b->welford_mean && b->welford_m2;
  if (do_w) {
    memcpy(wmean, b->welford_mean + (size_t)c * D, sizeof(double) * D);
    memcpy(wm2, b->welford_m2 + (size_t)c * D, sizeof(double) * D);
  }
  double ql[ORACLE_MAXD], pl[ORACLE_MAXD], gl[ORACLE_MAXD];
  double qr[ORACLE_MAXD], pr[ORACLE_MAXD], gr[ORACLE_MAXD];
  double qp[ORACLE_MAXD], gp[ORACLE_MAXD], lpp;
  double qs[ORACLE_MAXD], ps[ORACLE_MAXD], gs[ORACLE_MAXD], dq[ORACLE_MAXD];

  for (int64_t it = 0; it < B + S; ++it) {
    if (it == B) acc = 0; /* NUTS.py:494-500 */
    const int64_t si = it - B;
    draw_normals(&rng, it, z);
    double p0[ORACLE_MAXD];
    for (int d = 0; d < D; ++d) p0[d] = z[d] / sqw[d];          /* NUTS.py:219 */
    const double h0 = -lp + kinetic(p0, w, D);                   /* NUTS.py:222 */
    const double log_u = log(draw_uniform(&rng, it)) - h0;       /* NUTS.py:224 */
    memcpy(ql, q, sizeof(double) * D); memcpy(pl, p0, sizeof(double) * D); memcpy(gl, g, sizeof(double) * D);
    memcpy(qr, q, sizeof(double) * D); memcpy(pr, p0, sizeof(double) * D); memcpy(gr, g, sizeof(double) * D);
    memcpy(qp, q, sizeof(double) * D); memcpy(gp, g, sizeof(double) * D);
    lpp = lp;
    long n_valid = 1, n_steps = 0;
    double sum_alpha = 0.0;
    int depth = 0, diverged = 0;
    for (;;) {
      /* cond_fn, NUTS.py:243-248 */
      for (int d = 0; d < D; ++d) dq[d] = qr[d] - ql[d];
      const int u_turn = (dotv(dq, pl, D) < 0) | (dotv(dq, pr, D) < 0); /* NaN -> false */
      if (!((depth < max_depth) && !u_turn && !diverged)) break;
      double u_dir, u_prop;
      draw_nuts(&rng, it, depth, &u_dir, &u_prop);
      const int direction = (u_dir < 0.5) ? 1 : -1; /* bernoulli(0.5): uniform < p, NUTS.py:255-256 */
      if (direction == -1) { memcpy(qs, ql, sizeof(double) * D); memcpy(ps, pl, sizeof(double) * D); memcpy(gs, gl, sizeof(double) * D); }
      else { memcpy(qs, qr, sizeof(double) * D); memcpy(ps, pr, sizeof(double) * D); memcpy(gs, gr, sizeof(double) * D); }
      const long num_steps = 1L << depth;
      const double se = direction * eps;       /* NUTS.py:155 */
      const double hse = 0.5 * se;              /* NUTS.py:107 half*eps */
      double lps = NAN, sa = 0.0;
      for (long k = 0; k < num_steps; ++k) {    /* NUTS.py:173-182 */
        for (int d = 0; d < D; ++d) ps[d] = ps[d] + hse * gs[d];
        for (int d = 0; d < D; ++d) qs[d] = qs[d] + se * (ps[d] * w[d]);
        lps = target_eval(&tg, qs, gs);
        for (int d = 0; d < D; ++d) ps[d] = ps[d] + hse * gs[d];
        const double h_new = -lps + kinetic(ps, w, D);
        sa = sa + exp(min_nan(0.0, h0 - h_new));
      }
      cnt[1] += (unsigned long long)num_steps;
      const double h_new = -lps + kinetic(ps, w, D);             /* NUTS.py:282 */
      const int in_slice = log_u <= -h_new;
      const int is_div = (h_new - h0) > p->delta_max;
      const int is_valid = in_slice && !is_div;
      if (direction == -1) { memcpy(ql, qs, sizeof(double) * D); memcpy(pl, ps, sizeof(double) * D); memcpy(gl, gs, sizeof(double) * D); }
      else { memcpy(qr, qs, sizeof(double) * D); memcpy(pr, ps, sizeof(double) * D); memcpy(gr, gs, sizeof(double) * D); }
      const long n_valid_new = is_valid ? num_steps : 0;
      const long total_valid = n_valid + n_valid_new;
      float accept_prob = 0.0f;                                  /* NUTS.py:331-335 (float32) */
      if (total_valid > 0 && is_valid)
        accept_prob = (float)((double)n_valid_new / (double)(total_valid > 1 ? total_valid : 1));
      const int accept_new = u_prop < (double)accept_prob;      /* NUTS.py:336 */
      if (accept_new) { memcpy(qp, qs, sizeof(double) * D); memcpy(gp, gs, sizeof(double) * D); lpp = lps; }
      n_valid = total_valid;
      sum_alpha = sum_alpha + sa;
      n_steps += num_steps;
      depth += 1;
      diverged = diverged | is_div;
    }
    double map = sum_alpha / (double)(n_steps > 1 ? n_steps : 1); /* NUTS.py:386 */
    if (isnan(map) || isinf(map)) map = 0.65;                      /* NUTS.py:389-393 */
    memcpy(q, qp, sizeof(double) * D); memcpy(g, gp, sizeof(double) * D); lp = lpp;
    acc += 1; /* accepted = True, NUTS.py:385,439 */
    cnt[2] += (unsigned long long)depth;
    cnt[3] += (unsigned long long)diverged;
    if (si >= 0) {
      const size_t o = (size_t)si * C + c;
      if (!(p->flags & MCMC_FLAG_NO_SAMPLES)) {
        if (b->samples) memcpy(b->samples + o * D, q, sizeof(double) * D);
        if (b->log_probs) b->log_probs[o] = lp;
      }
      if (b->tree_depths) b->tree_depths[o] = depth;
      if (b->mean_accept_probs) b->mean_accept_probs[o] = map;
      if (do_w) { wcount += 1.0; welford_fold(wmean, wm2, wcount, q, D); }
    }
  }
  if (S == 0) acc = 0; /* counters are reset after burn-in even when nothing is sampled (HMC.py:276-282) */
  if (b->accept_count) b->accept_count[c] = acc;
  if (b->accept_rate) b->accept_rate[c] = (float)acc / (float)S;
  if (b->final_q) memcpy(b->final_q + (size_t)c * D, q, sizeof(double) * D);
  if (b->final_g) memcpy(b->final_g + (size_t)c * D, g, sizeof(double) * D);
  if (b->final_lp) b->final_lp[c] = lp;
  if (do_w) {
    memcpy(b->welford_mean + (size_t)c * D, wmean, sizeof(double) * D);
    memcpy(b->welford_m2 + (size_t)c * D, wm2, sizeof(double) * D);
  }
  cnt[0] += (unsigned long long)acc;
}

int oracle_nuts_run(const McmcParams* p, const McmcBuffers* b) {
  int rc = check_common(p, b);
  if (rc) return rc;
  if (p->max_tree_depth < 0 || p->max_tree_depth > MCMC_MAX_TREE_DEPTH) return MCMC_EINVAL;
  unsigned long long c0 = 0, c1 = 0, c2 = 0, c3 = 0;
#pragma omp parallel for schedule(dynamic, 16) reduction(+ : c0, c1, c2, c3)
  for (int64_t c = 0; c < p->n_chains; ++c) {
    unsigned long long cnt[4] = {0, 0, 0, 0};
    nuts_chain(p, b, c, cnt);
    c0 += cnt[0]; c1 += cnt[1]; c2 += cnt[2]; c3 += cnt[3];
  }
  if (b->counters) { b->counters[0] = c0; b->counters[1] = c1; b->counters[2] = c2; b->counters[3] = c3; }
  return MCMC_OK;
}

/* ------------------------------------------------------------ small extras */

int oracle_target_eval(const McmcParams* p, const double* cov_inv, const double* q, int64_t n, double* lp,
                       double* grad) {
  Target tg;
  make_target(p, NULL, &tg);
  tg.cov_inv = cov_inv;
  const int D = p->n_dim;
#pragma omp parallel for schedule(static)
  for (int64_t i = 0; i < n; ++i) {
    double gtmp[ORACLE_MAXD];
    const double v = target_eval(&tg, q + (size_t)i * D, grad ? gtmp : NULL);
    if (lp) lp[i] = v;
    if (grad) memcpy(grad + (size_t)i * D, gtmp, sizeof(double) * D);
  }
  return MCMC_OK;
}

/* tuning/adaptation.py:240-248: sum over chains of m2/(max(n,2)-1) */
int oracle_welford_finalize(const double* m2, double count, int64_t C, int32_t D, double* out_sum_var) {
  const double n = count > 2.0 ? count : 2.0;
  for (int d = 0; d < D; ++d) {
    double s = 0.0;
    for (int64_t c = 0; c < C; ++c) s += m2[(size_t)c * D + d] / (n - 1.0);
    out_sum_var[d] = s;
  }
  return MCMC_OK;
}

/* native stream, exposed so tests can pin Philox against the Random123 known-answer vectors */
void oracle_philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]) {
  philox4x32_10(ctr, key, out);
}

/* torchrun exports OMP_NUM_THREADS=1 to its workers; the timed CPU arm asks for the host's cores explicitly */
void oracle_set_num_threads(int n) {
#ifdef _OPENMP
  if (n > 0) omp_set_num_threads(n);
#else
  (void)n;
#endif
}

int oracle_num_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}
