"""CUDA sources of user-supplied densities used by the custom-target tests (mcmc_b200/custom.py)."""

# log p(q) = -0.5 * par[0] * sum(q^2): with par[0] = 1 the built-in standard_normal without its constant (README.md:51-52)
SCALED_NORMAL = '''
__device__ double mcmc_user_log_prob_grad(const double* q, double* g, int n, const double* par) {
  double s = 0.0;
  for (int d = 0; d < n; ++d) { s += q[d] * q[d]; g[d] = -par[0] * q[d]; }
  return -0.5 * par[0] * s;
}'''

# the reference's rosenbrock (benchmarks/targets.py:367-387) with a = par[0]: value and hand-derived gradient
ROSENBROCK = '''
__device__ double mcmc_user_log_prob_grad(const double* q, double* g, int n, const double* par) {
  const double a = par[0];
  double U = 0.0;
  for (int d = 0; d < n; ++d) g[d] = 0.0;
  for (int i = 0; i + 1 < n; ++i) {
    const double om = 1.0 - q[i], t = q[i + 1] - q[i] * q[i];
    U += om * om + a * t * t;
    g[i] += 2.0 * om + 4.0 * a * q[i] * t;
    g[i + 1] += -2.0 * a * t;
  }
  return -U;
}'''

BROKEN = '''
__device__ double mcmc_user_log_prob_grad(const double* q, double* g, int n, const double* par) {
  return this_function_does_not_exist(q[0]);
}'''
