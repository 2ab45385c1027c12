// abi.cu — the extern "C" boundary declared in include/mcmc_b200.h: argument
// validation (the reference's ValueError / KeyError cases), per-target dispatch,
// the cross-chain reduction kernels of the warmup, the host-pointer entry and
// the fp64 peak probe.  No exception crosses this file; no CPU compute path exists.
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <vector>

#include "common.cuh"
#include "hmc_kernel.cuh"
#include "philox.cuh"

namespace mcmc {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

NutsWs nuts_ws_layout(const McmcParams& p) {
  NutsWs w{};
  const size_t C = (size_t)p.n_chains, D = (size_t)p.n_dim;
  auto up = [](size_t x) { return (x + 255) / 256 * 256; };
  size_t off = 0;
  // lane tiles in global memory only for the runtime-D kernel (compiled D keeps them in shared memory);
  // at most 8 resident warps per SM there, and never more warps than chains need
  const bool smem_tiles = (p.n_dim == 10 || p.n_dim == 20);
  int sms = device_sm_count();
  if (sms <= 0) sms = 148;
  w.max_warps = smem_tiles ? (int64_t)sms * 32 : (int64_t)sms * 8;
  const int64_t need_warps = ((int64_t)C + 31) / 32;
  const int64_t tile_warps = smem_tiles ? 0 : (need_warps < w.max_warps ? need_warps : w.max_warps);
  w.slots_off = off; off += up((size_t)tile_warps * 5 * D * 32 * sizeof(double));
  w.home_off = off;  off += up(C * 2 * D * sizeof(double));
  w.lp_off = off;    off += up(C * sizeof(double));
  w.acc_off = off;   off += up(C * sizeof(int));
  w.queue_off = off; off += up((64 + C) * sizeof(unsigned int));
  const int64_t T = p.burn_in + p.num_samples;
  w.tseg = T > 0 ? T : 1;
  w.z_off = w.lu_off = off;
  if (p.rng_mode == MCMC_RNG_PHILOX && C > 0 && T > 0) {
    // draw buffer capped (default 4 GiB; MCMC_NUTS_RNG_CAP_MB is a test hook that forces segmentation)
    size_t cap = (size_t)4 << 30;
    if (const char* e = getenv("MCMC_NUTS_RNG_CAP_MB")) cap = (size_t)atoll(e) << 20;
    const size_t per_it = C * (D + 1) * sizeof(double);
    int64_t ts = (int64_t)(cap / per_it);
    if (ts < 1) ts = 1;
    if (ts < T) w.tseg = ts;
    w.z_off = off;  off += up((size_t)w.tseg * C * D * sizeof(double));
    w.lu_off = off; off += up((size_t)w.tseg * C * sizeof(double));
  }
  w.total = off + 256;
  return w;
}

size_t nuts_workspace_bytes(const McmcParams& p) { return nuts_ws_layout(p).total; }

size_t hmc_workspace_bytes(const McmcParams& p) {
  // queue + flags + hand-off state (q, g, lp, acc) of the chunked scheduler in hmc_kernel.cuh
  const size_t C = ((size_t)p.n_chains + 255) / 256 * 256, D = (size_t)p.n_dim;
  const size_t flag_bytes = (((C + 31) / 32) * 4 + 255) / 256 * 256;
  return 256 + flag_bytes + (2 * C * D + C) * sizeof(double) + C * sizeof(int) + 256;
}

bool no_spin() {
  const char* e = getenv("MCMC_NO_SPIN");
  return e && e[0] && e[0] != '0';
}

int device_sm_count() {
  static int sms[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 0;
  if (sms[dev] == 0) cudaDeviceGetAttribute(&sms[dev], cudaDevAttrMultiProcessorCount, dev);
  return sms[dev];
}

#define MCMC_DECL(T)                                                                                   \
  template <> int launch_hmc<T>(const Launch&, bool);                                                  \
  template <> int launch_rwmh<T>(const Launch&);                                                       \
  template <> int launch_nuts<T>(const Launch&);                                                       \
  template <> int launch_target_eval<T>(const McmcParams&, const double*, const double*, int64_t,      \
                                        double*, double*, cudaStream_t);                              \
  template <> int launch_leapfrog<T>(const McmcParams&, const LeapfrogIO&, bool, cudaStream_t);
MCMC_DECL(0) MCMC_DECL(1) MCMC_DECL(2) MCMC_DECL(3) MCMC_DECL(4) MCMC_DECL(5) MCMC_DECL(6) MCMC_DECL(7) MCMC_DECL(8)
#undef MCMC_DECL

typedef int (*HmcFn)(const Launch&, bool);
typedef int (*RunFn)(const Launch&);
typedef int (*EvalFn)(const McmcParams&, const double*, const double*, int64_t, double*, double*, cudaStream_t);
static const HmcFn kHmc[MCMC_NUM_TARGETS] = {launch_hmc<0>, launch_hmc<1>, launch_hmc<2>, launch_hmc<3>, launch_hmc<4>,
                                             launch_hmc<5>, launch_hmc<6>, launch_hmc<7>, launch_hmc<8>};
static const RunFn kRwmh[MCMC_NUM_TARGETS] = {launch_rwmh<0>, launch_rwmh<1>, launch_rwmh<2>, launch_rwmh<3>, launch_rwmh<4>,
                                              launch_rwmh<5>, launch_rwmh<6>, launch_rwmh<7>, launch_rwmh<8>};
static const RunFn kNuts[MCMC_NUM_TARGETS] = {launch_nuts<0>, launch_nuts<1>, launch_nuts<2>, launch_nuts<3>, launch_nuts<4>,
                                              launch_nuts<5>, launch_nuts<6>, launch_nuts<7>, launch_nuts<8>};
static const EvalFn kEval[MCMC_NUM_TARGETS] = {launch_target_eval<0>, launch_target_eval<1>, launch_target_eval<2>,
                                               launch_target_eval<3>, launch_target_eval<4>, launch_target_eval<5>,
                                               launch_target_eval<6>, launch_target_eval<7>, launch_target_eval<8>};

typedef int (*LeapfrogFn)(const McmcParams&, const LeapfrogIO&, bool, cudaStream_t);
static const LeapfrogFn kLeapfrog[MCMC_NUM_TARGETS] = {launch_leapfrog<0>, launch_leapfrog<1>, launch_leapfrog<2>,
                                                       launch_leapfrog<3>, launch_leapfrog<4>, launch_leapfrog<5>,
                                                       launch_leapfrog<6>, launch_leapfrog<7>, launch_leapfrog<8>};

// user targets (ids MCMC_TARGET_USER0 ..): launchers handed over by a plugin library (mcmc_register_user_target)
struct UserTarget { HmcFn hmc; RunFn rwmh; RunFn nuts; EvalFn eval; LeapfrogFn leapfrog; };
static UserTarget g_user[MCMC_MAX_USER_TARGETS] = {};
static int g_user_count = 0;
static std::mutex g_user_mu;

static bool target_known(int tid) {
  return (tid >= 0 && tid < MCMC_NUM_TARGETS) || (tid >= MCMC_TARGET_USER0 && tid < MCMC_TARGET_USER0 + g_user_count);
}
static HmcFn hmc_fn(int tid) { return tid < MCMC_NUM_TARGETS ? kHmc[tid] : g_user[tid - MCMC_TARGET_USER0].hmc; }
static RunFn rwmh_fn(int tid) { return tid < MCMC_NUM_TARGETS ? kRwmh[tid] : g_user[tid - MCMC_TARGET_USER0].rwmh; }
static RunFn nuts_fn(int tid) { return tid < MCMC_NUM_TARGETS ? kNuts[tid] : g_user[tid - MCMC_TARGET_USER0].nuts; }
static EvalFn eval_fn(int tid) { return tid < MCMC_NUM_TARGETS ? kEval[tid] : g_user[tid - MCMC_TARGET_USER0].eval; }
static LeapfrogFn leapfrog_fn(int tid) { return tid < MCMC_NUM_TARGETS ? kLeapfrog[tid] : g_user[tid - MCMC_TARGET_USER0].leapfrog; }

static int have_device() {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); n = 0; }
  return n;
}

static int validate(const McmcParams* p, const McmcBuffers* b, int sampler) {
  if (!p || !b) { set_error("null params/buffers"); return MCMC_EINVAL; }
  if (p->abi_version != MCMC_ABI_VERSION) { set_error("abi_version %u != %d", p->abi_version, MCMC_ABI_VERSION); return MCMC_EINVAL; }
  if (!target_known(p->target_id)) { set_error("Unknown target id %d", p->target_id); return MCMC_EINVAL; }
  if (p->n_dim < 1 || p->n_dim > MCMC_MAX_DIM) { set_error("n_dim %d outside [1, %d]", p->n_dim, MCMC_MAX_DIM); return MCMC_ENOTSUP; }
  if (p->n_chains < 0 || p->num_samples < 0 || p->burn_in < 0) { set_error("negative size"); return MCMC_EINVAL; }
  if (p->n_chains > 0 && !b->init_q) { set_error("init_q is NULL"); return MCMC_EINVAL; }
  if (p->rng_mode != MCMC_RNG_PHILOX && p->rng_mode != MCMC_RNG_INJECTED) { set_error("bad rng_mode %d", p->rng_mode); return MCMC_EINVAL; }
  if (p->rng_mode == MCMC_RNG_INJECTED && (p->burn_in + p->num_samples) > 0 && p->n_chains > 0) {
    if (!b->z_inject || !b->u_inject) { set_error("rng_mode INJECTED needs z_inject and u_inject"); return MCMC_EINVAL; }
    if (sampler == MCMC_SAMPLER_NUTS && p->max_tree_depth > 0 && (!b->nuts_dir_u || !b->nuts_prop_u)) {
      set_error("rng_mode INJECTED needs nuts_dir_u and nuts_prop_u"); return MCMC_EINVAL;
    }
  }
  if (p->target_id == MCMC_TARGET_DENSE_GAUSSIAN) {
    if (!b->cov_inv) { set_error("dense_gaussian needs cov_inv"); return MCMC_EINVAL; }
    if (p->n_dim > 64) { set_error("dense_gaussian supports n_dim <= 64"); return MCMC_ENOTSUP; }
  }
  if ((sampler == MCMC_SAMPLER_HMC || sampler == MCMC_SAMPLER_GRAHMC)) {
    if (p->num_steps < 0 || p->num_steps > MCMC_MAX_STEPS) { set_error("num_steps %d outside [0, %d]", p->num_steps, MCMC_MAX_STEPS); return MCMC_ENOTSUP; }
  }
  if (sampler == MCMC_SAMPLER_GRAHMC && (p->schedule_id < 0 || p->schedule_id >= MCMC_NUM_SCHEDULES)) {
    set_error("Unknown friction schedule id %d", p->schedule_id); return MCMC_EINVAL;  // KeyError, GRAHMC.py:122
  }
  if (sampler == MCMC_SAMPLER_NUTS && (p->max_tree_depth < 0 || p->max_tree_depth > MCMC_MAX_TREE_DEPTH)) {
    set_error("max_tree_depth %d outside [0, %d]", p->max_tree_depth, MCMC_MAX_TREE_DEPTH); return MCMC_EINVAL;
  }
  if (p->burn_in + p->num_samples > 0xffffffffLL) { set_error("more than 2^32 iterations per call"); return MCMC_ENOTSUP; }
  if (have_device() == 0) { set_error("no CUDA device (there is no CPU fallback)"); return MCMC_ECUDA; }
  return MCMC_OK;
}

static int run_device(int sampler, const McmcParams* p, const McmcBuffers* b, cudaStream_t stream) {
  if (int rc = validate(p, b, sampler)) return rc;
  Launch l;
  l.p = *p;
  l.b = *b;
  l.stream = stream;
  if (b->counters)
    if (int rc = cuda_fail(cudaMemsetAsync(b->counters, 0, 4 * sizeof(unsigned long long), stream), "memset counters"))
      return rc;
  switch (sampler) {
    case MCMC_SAMPLER_HMC: return hmc_fn(p->target_id)(l, false);
    case MCMC_SAMPLER_GRAHMC: return hmc_fn(p->target_id)(l, true);
    case MCMC_SAMPLER_RWMH: return rwmh_fn(p->target_id)(l);
    case MCMC_SAMPLER_NUTS: return nuts_fn(p->target_id)(l);
  }
  set_error("Unknown sampler: %d", sampler);  // ValueError, adaptation.py:212
  return MCMC_EINVAL;
}

// ------------------------------------------------------------------ reductions

// out[d] = sum_c m2[c,d] / (max(count,2) - 1): one block per dimension slice, fixed-order
// tree -> deterministic (tuning/welford.py:97-99, adaptation.py:240-248)
__global__ void __launch_bounds__(256) welford_finalize_kernel(const double* __restrict__ m2, double denom, int64_t C,
                                                               int D, double* __restrict__ partial) {
  // grid = (nblk, 1); each block sums a contiguous range of chains for all D, writes partial[blk*D + d]
  __shared__ double sh[256];
  const int64_t per = (C + gridDim.x - 1) / gridDim.x;
  const int64_t c0 = (int64_t)blockIdx.x * per;
  const int64_t c1 = c0 + per < C ? c0 + per : C;
  for (int d = 0; d < D; ++d) {
    double s = 0.0;
    for (int64_t c = c0 + threadIdx.x; c < c1; c += blockDim.x) s += m2[c * D + d] / denom;
    sh[threadIdx.x] = s;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
      if (threadIdx.x < o) sh[threadIdx.x] += sh[threadIdx.x + o];
      __syncthreads();
    }
    if (threadIdx.x == 0) partial[(size_t)blockIdx.x * D + d] = sh[0];
    __syncthreads();
  }
}

__global__ void sum_partials_kernel(const double* __restrict__ partial, int nblk, int D, double* __restrict__ out) {
  const int d = blockIdx.x * blockDim.x + threadIdx.x;
  if (d >= D) return;
  double s = 0.0;
  for (int b = 0; b < nblk; ++b) s += partial[(size_t)b * D + d];
  out[d] = s;
}

__global__ void __launch_bounds__(256) sum_f64_kernel(const double* __restrict__ x, int64_t n, double* __restrict__ partial) {
  __shared__ double sh[256];
  const int64_t per = (n + gridDim.x - 1) / gridDim.x;
  const int64_t i0 = (int64_t)blockIdx.x * per;
  const int64_t i1 = i0 + per < n ? i0 + per : n;
  double s = 0.0;
  for (int64_t i = i0 + threadIdx.x; i < i1; i += blockDim.x) s += x[i];
  sh[threadIdx.x] = s;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if (threadIdx.x < o) sh[threadIdx.x] += sh[threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x == 0) partial[blockIdx.x] = sh[0];
}

// per-chain sequential Welford over the S axis of samples (S, C, D): thread = (chain, dim) pair,
// consecutive threads read consecutive addresses of one sample slab (coalesced).
__global__ void __launch_bounds__(256) chain_moments_kernel(const double* __restrict__ samples, int64_t S, int64_t CD,
                                                            double* __restrict__ mean, double* __restrict__ m2,
                                                            double count_before) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= CD) return;
  double mu = mean[i], mm = m2[i], cnt = count_before;
  for (int64_t s = 0; s < S; ++s) {
    const double x = samples[s * CD + i];
    cnt += 1.0;
    const double delta = x - mu;
    mu = mu + delta / cnt;
    mm = mm + delta * (x - mu);
  }
  mean[i] = mu;
  m2[i] = mm;
}

// tuning/dual_averaging.py:818-887 on the device; one thread
__global__ void da_init_kernel(double* st, double eps0) {
  const double ls = log(eps0);
  st[0] = ls; st[1] = ls; st[2] = 0.0; st[3] = ls; st[4] = 0.0; st[5] = eps0; st[6] = eps0; st[7] = 0.0;
}
__global__ void da_update_kernel(double* st, const double* __restrict__ accept_sum, double n_terms, double target,
                                 double omega, double t0, double kappa) {
  const double stat = accept_sum[0] / (n_terms > 0.0 ? n_terms : accept_sum[1]);
  const double m = st[4] + 1.0;
  const double eta = 1.0 / (m + t0);
  const double hbar = (1.0 - eta) * st[2] + eta * (target - stat);
  const double ls = st[3] - (sqrt(m) / omega) * hbar;
  const double mk = pow(m, -kappa);
  const double lsb = (m == 1.0) ? ls : mk * ls + (1.0 - mk) * st[1];
  st[0] = ls; st[1] = lsb; st[2] = hbar; st[4] = m; st[5] = exp(ls); st[6] = exp(lsb); st[7] = stat;
}
__global__ void da_reset_kernel(double* st) {
  const double cur = st[4] > 0.0 ? st[1] : st[0];
  st[0] = cur; st[1] = cur; st[2] = 0.0; st[3] = cur; st[4] = 0.0; st[5] = exp(cur); st[6] = st[5];
}

__global__ void friction_table_kernel(int sched, int L, double eps, double gam, double steep, double* s_out,
                                      double* gamma_out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= L) return;
  const double gi = friction_gamma_dev(sched, i * eps, eps * L, gam, steep);
  if (gamma_out) gamma_out[i] = gi;
  s_out[i] = exp(-gi * (0.5 * eps));
}

// fp64 roofline denominator: 8 independent FMA chains per thread, registers only
__global__ void __launch_bounds__(256) fp64_probe_kernel(int iters, double* sink) {
  double a0 = threadIdx.x * 1e-9, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, a4 = a0 + 4, a5 = a0 + 5, a6 = a0 + 6, a7 = a0 + 7;
  const double m = 0.999999, c = 1e-6;
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int r = 0; r < 8; ++r) {
      a0 = fma(a0, m, c); a1 = fma(a1, m, c); a2 = fma(a2, m, c); a3 = fma(a3, m, c);
      a4 = fma(a4, m, c); a5 = fma(a5, m, c); a6 = fma(a6, m, c); a7 = fma(a7, m, c);
    }
  }
  const double s = ((a0 + a1) + (a2 + a3)) + ((a4 + a5) + (a6 + a7));
  if (s == 123.456) sink[0] = s;
}

// ------------------------------------------------------------------ native draws for NUTS (see nuts_kernel.cuh)

// one thread per (iteration, chain): 20 normals (10 Philox blocks) + the log of the accept uniform, with the
// counters of philox.cuh -> identical values to an in-kernel draw; rows leave as 16-byte stores
__global__ void __launch_bounds__(128) nuts_rng_fill_kernel(McmcParams p, int64_t it_begin, int64_t n_it,
                                                            double* __restrict__ z, double* __restrict__ logu) {
  __shared__ double sm_logtab[kRngTableDoubles];
  rng_table_fill(sm_logtab);
  __syncthreads();
  const int64_t C = p.n_chains;
  const int n = p.n_dim;
  const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= n_it * C) return;
  const int64_t t = idx / C, c = idx - t * C;
  RngCtx rng;
  const uint64_t gid = (uint64_t)(p.chain_offset + c);
  rng.k0 = p.seed[0]; rng.k1 = p.seed[1];
  rng.gid_lo = (uint32_t)gid; rng.gid_hi = (uint32_t)(gid >> 32);
  rng.logtab = sm_logtab;
  const uint32_t it = (uint32_t)(it_begin + t);
  double* row = z + (size_t)idx * n;
  const bool vec = ((n & 1) == 0) && ((reinterpret_cast<uintptr_t>(z) & 15) == 0);
  for (int j = 0; z && 2 * j < n; ++j) {
    double z0, z1;
    philox_normals_block(rng, it, (uint32_t)j, z0, z1);
    if (vec) {
      *reinterpret_cast<double2*>(row + 2 * j) = make_double2(z0, z1);
    } else {
      row[2 * j] = z0;
      if (2 * j + 1 < n) row[2 * j + 1] = z1;
    }
  }
  if (logu) logu[idx] = philox_accept_log_uniform(rng, it);
}

int launch_rng_fill(const McmcParams& p, int64_t it_begin, int64_t it_end, double* z, double* logu, cudaStream_t s) {
  const int64_t total = (it_end - it_begin) * p.n_chains;
  if (total <= 0) return MCMC_OK;
  const int64_t grid = (total + 127) / 128;
  if (grid > 0x7fffffffLL) { set_error("nuts rng fill: segment too large"); return MCMC_ENOTSUP; }
  nuts_rng_fill_kernel<<<(unsigned)grid, 128, 0, s>>>(p, it_begin, it_end - it_begin, z, logu);
  return cuda_fail(cudaGetLastError(), "nuts_rng_fill launch");
}

// ------------------------------------------------------------------ host arena (mcmc_run_host)

// One arena + stream per device, each behind its own mutex: host callers driving different GPUs do not serialise on
// each other and a device switch does not free / reallocate.
constexpr int kMaxDevices = 64;
struct Arena {
  std::mutex mu;
  char* base = nullptr;
  size_t cap = 0;
  cudaStream_t stream = nullptr;
};
static Arena g_arena[kMaxDevices];

// makes `device` current for the calling thread and restores the previous device on scope exit (a host framework
// or a worker thread may have another device current)
struct DeviceGuard {
  int prev = -1;
  int rc = MCMC_OK;
  explicit DeviceGuard(int device) {
    if (cudaGetDevice(&prev) != cudaSuccess) { cudaGetLastError(); prev = -1; }
    rc = cuda_fail(cudaSetDevice(device), "cudaSetDevice");
  }
  ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

// caller holds a.mu and has made the arena's device current
static int arena_reserve(Arena& a, size_t bytes) {
  if (a.cap < bytes) {
    if (a.base) { cudaFree(a.base); a.base = nullptr; a.cap = 0; }
    const size_t want = bytes + (bytes >> 3) + (1u << 20);
    if (int rc = cuda_fail(cudaMalloc(&a.base, want), "cudaMalloc(host arena)")) return rc;
    a.cap = want;
  }
  if (!a.stream)
    if (int rc = cuda_fail(cudaStreamCreateWithFlags(&a.stream, cudaStreamNonBlocking), "cudaStreamCreate")) return rc;
  return MCMC_OK;
}

}  // namespace mcmc

using namespace mcmc;

extern "C" {

int mcmc_abi_version(void) { return MCMC_ABI_VERSION; }
const char* mcmc_build_info(void) {
  return "mcmc_b200 sm_100a"
#if MCMC_IS_FAITHFUL
         " strict"
#else
         " fast"
#endif
#ifdef MCMC_WITH_WS
         " +ws"
#endif
      ;
}
const char* mcmc_last_error(void) { return g_err; }
int mcmc_device_count(void) { return have_device(); }
int mcmc_set_device(int device) {
  if (device < 0 || device >= have_device()) { set_error("device %d not available", device); return MCMC_ECUDA; }
  return cuda_fail(cudaSetDevice(device), "cudaSetDevice");
}

int mcmc_register_user_target(void* launch_hmc, void* launch_rwmh, void* launch_nuts, void* launch_eval, void* launch_leapfrog) {
  if (!launch_hmc || !launch_rwmh || !launch_nuts || !launch_eval || !launch_leapfrog) { set_error("register_user_target: null launcher"); return MCMC_EINVAL; }
  std::lock_guard<std::mutex> lk(g_user_mu);
  if (g_user_count >= MCMC_MAX_USER_TARGETS) { set_error("at most %d user targets per process", MCMC_MAX_USER_TARGETS); return MCMC_ENOTSUP; }
  g_user[g_user_count] = UserTarget{reinterpret_cast<HmcFn>(launch_hmc), reinterpret_cast<RunFn>(launch_rwmh),
                                    reinterpret_cast<RunFn>(launch_nuts), reinterpret_cast<EvalFn>(launch_eval),
                                    reinterpret_cast<LeapfrogFn>(launch_leapfrog)};
  return MCMC_TARGET_USER0 + g_user_count++;
}

int mcmc_workspace_bytes(int sampler, const McmcParams* p, size_t* bytes) {
  if (!p || !bytes) { set_error("null argument"); return MCMC_EINVAL; }
  *bytes = (sampler == MCMC_SAMPLER_NUTS) ? nuts_workspace_bytes(*p)
           : (sampler == MCMC_SAMPLER_HMC || sampler == MCMC_SAMPLER_GRAHMC) ? hmc_workspace_bytes(*p) : 0;
  return MCMC_OK;
}

int mcmc_hmc_run(const McmcParams* p, const McmcBuffers* b, void* stream) {
  return run_device(MCMC_SAMPLER_HMC, p, b, (cudaStream_t)stream);
}
int mcmc_grahmc_run(const McmcParams* p, const McmcBuffers* b, void* stream) {
  return run_device(MCMC_SAMPLER_GRAHMC, p, b, (cudaStream_t)stream);
}
int mcmc_rwmh_run(const McmcParams* p, const McmcBuffers* b, void* stream) {
  return run_device(MCMC_SAMPLER_RWMH, p, b, (cudaStream_t)stream);
}
int mcmc_nuts_run(const McmcParams* p, const McmcBuffers* b, void* stream) {
  return run_device(MCMC_SAMPLER_NUTS, p, b, (cudaStream_t)stream);
}

int mcmc_target_eval(const McmcParams* p, const double* cov_inv, const double* q, int64_t n, double* lp, double* grad,
                     void* stream) {
  if (!p || (!q && n > 0)) { set_error("null argument"); return MCMC_EINVAL; }
  if (!target_known(p->target_id)) { set_error("Unknown target id %d", p->target_id); return MCMC_EINVAL; }
  if (p->n_dim < 1 || p->n_dim > MCMC_MAX_DIM) { set_error("n_dim %d outside [1, %d]", p->n_dim, MCMC_MAX_DIM); return MCMC_ENOTSUP; }
  if (p->target_id == MCMC_TARGET_DENSE_GAUSSIAN && (!cov_inv || p->n_dim > 64)) { set_error("dense_gaussian needs cov_inv, n_dim <= 64"); return MCMC_EINVAL; }
  if (have_device() == 0) { set_error("no CUDA device (there is no CPU fallback)"); return MCMC_ECUDA; }
  return eval_fn(p->target_id)(*p, cov_inv, q, n, lp, grad, (cudaStream_t)stream);
}

int mcmc_leapfrog(const McmcParams* p, int friction, const double* cov_inv, const double* inv_mass, const double* q,
                  const double* momentum, const double* log_prob, const double* grad, double* q_out, double* momentum_out,
                  double* log_prob_out, double* grad_out, void* stream) {
  if (!p) { set_error("null params"); return MCMC_EINVAL; }
  if (p->abi_version != MCMC_ABI_VERSION) { set_error("abi_version %u != %d", p->abi_version, MCMC_ABI_VERSION); return MCMC_EINVAL; }
  if (!target_known(p->target_id)) { set_error("Unknown target id %d", p->target_id); return MCMC_EINVAL; }
  if (p->n_dim < 1 || p->n_dim > MCMC_MAX_DIM) { set_error("n_dim %d outside [1, %d]", p->n_dim, MCMC_MAX_DIM); return MCMC_ENOTSUP; }
  if (p->n_chains < 0) { set_error("negative size"); return MCMC_EINVAL; }
  if (p->n_chains > 0 && (!q || !momentum || !grad)) { set_error("leapfrog: position, momentum and grad_log_prob are required"); return MCMC_EINVAL; }
  if (p->num_steps < 0 || p->num_steps > MCMC_MAX_STEPS) { set_error("num_steps %d outside [0, %d]", p->num_steps, MCMC_MAX_STEPS); return MCMC_ENOTSUP; }
  if (friction && (p->schedule_id < 0 || p->schedule_id >= MCMC_NUM_SCHEDULES)) {
    set_error("Unknown friction schedule id %d", p->schedule_id); return MCMC_EINVAL;  // KeyError, GRAHMC.py:122
  }
  if (p->target_id == MCMC_TARGET_DENSE_GAUSSIAN && (!cov_inv || p->n_dim > 64)) { set_error("dense_gaussian needs cov_inv, n_dim <= 64"); return MCMC_EINVAL; }
  if (have_device() == 0) { set_error("no CUDA device (there is no CPU fallback)"); return MCMC_ECUDA; }
  const LeapfrogIO io{cov_inv, inv_mass, q, momentum, log_prob, grad, q_out, momentum_out, log_prob_out, grad_out};
  return leapfrog_fn(p->target_id)(*p, io, friction != 0, (cudaStream_t)stream);
}

int mcmc_philox_fill(const McmcParams* p, int64_t it_begin, int64_t n_it, double* z, double* log_u, void* stream) {
  if (!p) { set_error("null params"); return MCMC_EINVAL; }
  if (p->n_dim < 1 || p->n_dim > MCMC_MAX_DIM || p->n_chains < 0 || n_it < 0 || it_begin < 0 ||
      it_begin + n_it > 0xffffffffLL) { set_error("philox_fill: bad sizes"); return MCMC_EINVAL; }
  if (have_device() == 0) { set_error("no CUDA device (there is no CPU fallback)"); return MCMC_ECUDA; }
  return launch_rng_fill(*p, it_begin, it_begin + n_it, z, log_u, (cudaStream_t)stream);
}

int mcmc_friction_table(int schedule_id, int num_steps, double step_size, double gamma, double steepness, double* s_out,
                        double* gamma_out, void* stream) {
  if (schedule_id < 0 || schedule_id >= MCMC_NUM_SCHEDULES) { set_error("Unknown friction schedule id %d", schedule_id); return MCMC_EINVAL; }
  if (num_steps < 0 || !s_out) { set_error("bad argument"); return MCMC_EINVAL; }
  if (have_device() == 0) { set_error("no CUDA device (there is no CPU fallback)"); return MCMC_ECUDA; }
  if (num_steps == 0) return MCMC_OK;
  friction_table_kernel<<<(num_steps + 127) / 128, 128, 0, (cudaStream_t)stream>>>(schedule_id, num_steps, step_size, gamma,
                                                                                   steepness, s_out, gamma_out);
  return cuda_fail(cudaGetLastError(), "friction_table launch");
}

int mcmc_da_init(double* state, double initial_step_size, void* stream) {
  if (!state || !(initial_step_size > 0.0)) { set_error("da_init: bad argument"); return MCMC_EINVAL; }
  if (have_device() == 0) { set_error("no CUDA device (there is no CPU fallback)"); return MCMC_ECUDA; }
  da_init_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(state, initial_step_size);
  return cuda_fail(cudaGetLastError(), "da_init launch");
}
int mcmc_da_update(double* state, const double* accept_sum, double n_terms, double target_accept, double omega, double t0,
                   double kappa, void* stream) {
  if (!state || !accept_sum || !(n_terms >= 0.0) || !(omega > 0.0)) { set_error("da_update: bad argument"); return MCMC_EINVAL; }
  if (have_device() == 0) { set_error("no CUDA device (there is no CPU fallback)"); return MCMC_ECUDA; }
  da_update_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(state, accept_sum, n_terms, target_accept, omega, t0, kappa);
  return cuda_fail(cudaGetLastError(), "da_update launch");
}
int mcmc_da_reset(double* state, void* stream) {
  if (!state) { set_error("da_reset: null state"); return MCMC_EINVAL; }
  if (have_device() == 0) { set_error("no CUDA device (there is no CPU fallback)"); return MCMC_ECUDA; }
  da_reset_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(state);
  return cuda_fail(cudaGetLastError(), "da_reset launch");
}

#define MCMC_REDUCE_BLOCKS 256

int mcmc_welford_finalize(const double* welford_m2, double count, int64_t n_chains, int32_t n_dim, double* out_sum_var,
                          void* workspace, size_t workspace_bytes, void* stream) {
  if (!welford_m2 || !out_sum_var || n_dim < 1) { set_error("bad argument"); return MCMC_EINVAL; }
  if (have_device() == 0) { set_error("no CUDA device (there is no CPU fallback)"); return MCMC_ECUDA; }
  const size_t need = (size_t)MCMC_REDUCE_BLOCKS * n_dim * sizeof(double);
  if (!workspace || workspace_bytes < need) { set_error("welford_finalize needs %zu workspace bytes", need); return MCMC_EWORKSPACE; }
  const double n = count > 2.0 ? count : 2.0;
  welford_finalize_kernel<<<MCMC_REDUCE_BLOCKS, 256, 0, (cudaStream_t)stream>>>(welford_m2, n - 1.0, n_chains, n_dim,
                                                                               (double*)workspace);
  sum_partials_kernel<<<(n_dim + 127) / 128, 128, 0, (cudaStream_t)stream>>>((const double*)workspace, MCMC_REDUCE_BLOCKS,
                                                                             n_dim, out_sum_var);
  return cuda_fail(cudaGetLastError(), "welford_finalize launch");
}

// ---- allocation-free variant of mcmc_welford_finalize (no counterpart in the reference; added for callers that cannot
// allocate: CUDA graph capture, tight serving loops).  The block partials live in a file-scope device array instead of a
// caller workspace.  In HOST code `g_noalloc_scratch` is only a stand-in for the symbol, not a device address: the
// address is resolved once per device with cudaGetSymbolAddress (that device current) and cached.
static __device__ __align__(256) double g_noalloc_scratch[MCMC_REDUCE_BLOCKS * MCMC_NOALLOC_MAX_DIM];   // 256 KiB per GPU

static std::mutex g_noalloc_mu;
static double* g_noalloc_ptr[64] = {nullptr};

// device address of g_noalloc_scratch on the CURRENT device (set-up on the first call per device)
static int noalloc_scratch(double** out) {
  int dev = 0;
  if (int rc = cuda_fail(cudaGetDevice(&dev), "cudaGetDevice")) return rc;
  if (dev < 0 || dev >= 64) { set_error("device ordinal %d outside [0, 64)", dev); return MCMC_ENOTSUP; }
  std::lock_guard<std::mutex> lk(g_noalloc_mu);
  if (!g_noalloc_ptr[dev]) {
    double* p = nullptr;
    if (int rc = cuda_fail(cudaGetSymbolAddress((void**)&p, g_noalloc_scratch), "cudaGetSymbolAddress(noalloc scratch)"))
      return rc;
    if (!p) { set_error("cudaGetSymbolAddress returned a null address"); return MCMC_ECUDA; }
    g_noalloc_ptr[dev] = p;
  }
  *out = g_noalloc_ptr[dev];
  return MCMC_OK;
}

int mcmc_welford_finalize_noalloc(const double* welford_m2, double count, int64_t n_chains, int32_t n_dim,
                                  double* out_sum_var, void* stream) {
  if (!welford_m2 || !out_sum_var || n_dim < 1 || n_chains < 0) { set_error("bad argument"); return MCMC_EINVAL; }
  if (n_dim > MCMC_NOALLOC_MAX_DIM) {
    set_error("welford_finalize_noalloc: n_dim %d exceeds the static scratch (max %d dimensions = %zu bytes); use "
              "mcmc_welford_finalize with a caller workspace", n_dim, MCMC_NOALLOC_MAX_DIM, sizeof(double) * MCMC_REDUCE_BLOCKS * MCMC_NOALLOC_MAX_DIM);
    return MCMC_ENOTSUP;
  }
  if (have_device() == 0) { set_error("no CUDA device (there is no CPU fallback)"); return MCMC_ECUDA; }
  double* scratch = nullptr;
  if (int rc = noalloc_scratch(&scratch)) return rc;
  // nothing in the scratch has to start at zero: every one of the MCMC_REDUCE_BLOCKS blocks overwrites its n_dim partials
  // (a block with no chains writes 0.0) before sum_partials reads exactly those MCMC_REDUCE_BLOCKS * n_dim values, so
  // whatever an earlier call with another n_dim left behind is never read
  const double n = count > 2.0 ? count : 2.0;
  welford_finalize_kernel<<<MCMC_REDUCE_BLOCKS, 256, 0, (cudaStream_t)stream>>>(welford_m2, n - 1.0, n_chains, n_dim, scratch);
  sum_partials_kernel<<<(n_dim + 127) / 128, 128, 0, (cudaStream_t)stream>>>(scratch, MCMC_REDUCE_BLOCKS, n_dim, out_sum_var);
  return cuda_fail(cudaGetLastError(), "welford_finalize_noalloc launch");
}

int mcmc_chain_moments(const double* samples, int64_t num_samples, int64_t n_chains, int32_t n_dim, double* welford_mean,
                       double* welford_m2, double count_before, void* stream) {
  if (!samples || !welford_mean || !welford_m2) { set_error("null argument"); return MCMC_EINVAL; }
  if (have_device() == 0) { set_error("no CUDA device (there is no CPU fallback)"); return MCMC_ECUDA; }
  const int64_t CD = n_chains * n_dim;
  if (CD == 0) return MCMC_OK;
  chain_moments_kernel<<<(unsigned)((CD + 255) / 256), 256, 0, (cudaStream_t)stream>>>(samples, num_samples, CD, welford_mean,
                                                                                      welford_m2, count_before);
  return cuda_fail(cudaGetLastError(), "chain_moments launch");
}

int mcmc_sum_f64(const double* x, int64_t n, double* out, void* workspace, size_t workspace_bytes, void* stream) {
  if (!x || !out) { set_error("null argument"); return MCMC_EINVAL; }
  if (have_device() == 0) { set_error("no CUDA device (there is no CPU fallback)"); return MCMC_ECUDA; }
  const size_t need = (size_t)MCMC_REDUCE_BLOCKS * sizeof(double);
  if (!workspace || workspace_bytes < need) { set_error("sum_f64 needs %zu workspace bytes", need); return MCMC_EWORKSPACE; }
  sum_f64_kernel<<<MCMC_REDUCE_BLOCKS, 256, 0, (cudaStream_t)stream>>>(x, n, (double*)workspace);
  sum_partials_kernel<<<1, 32, 0, (cudaStream_t)stream>>>((const double*)workspace, MCMC_REDUCE_BLOCKS, 1, out);
  return cuda_fail(cudaGetLastError(), "sum_f64 launch");
}

int mcmc_fp64_peak_probe(int iters, double* flops_out, float* ms_out, void* stream) {
  if (have_device() == 0) { set_error("no CUDA device (there is no CPU fallback)"); return MCMC_ECUDA; }
  int dev = 0, sms = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const int blocks = sms * 8, threads = 256;
  double* sink = nullptr;
  if (int rc = cuda_fail(cudaMalloc(&sink, sizeof(double)), "cudaMalloc")) return rc;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  cudaStream_t s = (cudaStream_t)stream;
  fp64_probe_kernel<<<blocks, threads, 0, s>>>(iters / 8 + 1, sink);  // warm-up
  cudaEventRecord(e0, s);
  fp64_probe_kernel<<<blocks, threads, 0, s>>>(iters, sink);
  cudaEventRecord(e1, s);
  cudaError_t e = cudaEventSynchronize(e1);
  float ms = 0.f;
  cudaEventElapsedTime(&ms, e0, e1);
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  cudaFree(sink);
  if (int rc = cuda_fail(e, "fp64 probe")) return rc;
  if (flops_out) *flops_out = 2.0 * 64.0 * (double)iters * (double)blocks * (double)threads;
  if (ms_out) *ms_out = ms;
  return MCMC_OK;
}

int mcmc_host_arena_release(void) {
  for (int dev = 0; dev < kMaxDevices; ++dev) {
    Arena& a = g_arena[dev];
    std::lock_guard<std::mutex> lk(a.mu);
    if (!a.base && !a.stream) continue;
    DeviceGuard dg(dev);
    if (a.base) cudaFree(a.base);
    if (a.stream) cudaStreamDestroy(a.stream);
    a.base = nullptr; a.cap = 0; a.stream = nullptr;
  }
  return MCMC_OK;
}

int mcmc_run_host(int sampler, const McmcParams* p, const McmcBuffers* h, int device) {
  if (int rc = validate(p, h, sampler)) return rc;
  if (device < 0 || device >= have_device() || device >= kMaxDevices) { set_error("device %d not available", device); return MCMC_ECUDA; }
  // the device is made current for THIS call whatever the thread had before (and restored on return): the launchers
  // read the current device's SM count / occupancy and enqueue on this device's arena stream
  DeviceGuard dg(device);
  if (dg.rc) return dg.rc;
  Arena& arena = g_arena[device];
  std::lock_guard<std::mutex> lk(arena.mu);
  const size_t C = (size_t)p->n_chains, D = (size_t)p->n_dim, S = (size_t)p->num_samples;
  const size_t T = (size_t)(p->burn_in + p->num_samples);
  const size_t md = (size_t)(p->max_tree_depth > 0 ? p->max_tree_depth : 1);
  struct Item { const void* hsrc; void* hdst; void** slot; size_t bytes; };
  McmcBuffers d = *h;
  std::vector<Item> items;
  auto in = [&](const double* const& hp, const double*& dp, size_t n) {
    if (hp) items.push_back({hp, nullptr, (void**)&dp, n * sizeof(double)});
  };
  auto out = [&](auto* const& hp, auto*& dp, size_t bytes, bool also_in = false) {
    if (hp) items.push_back({also_in ? (const void*)hp : nullptr, (void*)hp, (void**)&dp, bytes});
  };
  in(h->init_q, d.init_q, C * D);
  in(h->inv_mass, d.inv_mass, D);
  in(h->cov_inv, d.cov_inv, D * D);
  in(h->step_size_dev, d.step_size_dev, 1);
  if (p->rng_mode == MCMC_RNG_INJECTED) {
    in(h->z_inject, d.z_inject, T * C * D);
    in(h->u_inject, d.u_inject, T * C);
    if (sampler == MCMC_SAMPLER_NUTS) { in(h->nuts_dir_u, d.nuts_dir_u, T * C * md); in(h->nuts_prop_u, d.nuts_prop_u, T * C * md); }
  } else {
    d.z_inject = d.u_inject = d.nuts_dir_u = d.nuts_prop_u = nullptr;
  }
  const bool nos = (p->flags & MCMC_FLAG_NO_SAMPLES) != 0;
  if (!nos) { out(h->samples, d.samples, S * C * D * 8); out(h->log_probs, d.log_probs, S * C * 8); }
  else { d.samples = nullptr; d.log_probs = nullptr; }
  out(h->accept_count, d.accept_count, C * 4);
  out(h->accept_rate, d.accept_rate, C * 4);
  out(h->final_q, d.final_q, C * D * 8);
  out(h->final_lp, d.final_lp, C * 8);
  if (sampler != MCMC_SAMPLER_RWMH) out(h->final_g, d.final_g, C * D * 8); else d.final_g = nullptr;
  if ((p->flags & MCMC_FLAG_TRACK_PROPOSALS) && (sampler == MCMC_SAMPLER_HMC || sampler == MCMC_SAMPLER_GRAHMC)) {
    out(h->pre_q, d.pre_q, S * C * D * 8); out(h->pre_lp, d.pre_lp, S * C * 8);
    out(h->prop_q, d.prop_q, S * C * D * 8); out(h->prop_lp, d.prop_lp, S * C * 8);
    out(h->delta_h, d.delta_h, S * C * 8);
  } else { d.pre_q = d.pre_lp = d.prop_q = d.prop_lp = d.delta_h = nullptr; }
  if (sampler == MCMC_SAMPLER_NUTS) { out(h->tree_depths, d.tree_depths, S * C * 4); out(h->mean_accept_probs, d.mean_accept_probs, S * C * 8); }
  else { d.tree_depths = nullptr; d.mean_accept_probs = nullptr; }
  if (p->flags & MCMC_FLAG_ACCUM_WELFORD) {
    out(h->welford_mean, d.welford_mean, C * D * 8, true); out(h->welford_m2, d.welford_m2, C * D * 8, true);
  } else { d.welford_mean = d.welford_m2 = nullptr; }
  out(h->counters, d.counters, 4 * sizeof(unsigned long long));
  size_t ws = (sampler == MCMC_SAMPLER_NUTS) ? nuts_workspace_bytes(*p)
              : (sampler == MCMC_SAMPLER_HMC || sampler == MCMC_SAMPLER_GRAHMC) ? hmc_workspace_bytes(*p) : 0;
  size_t total = ws + 256;
  for (auto& it : items) total += (it.bytes + 255) / 256 * 256;
  if (int rc = arena_reserve(arena, total)) return rc;
  char* cur = arena.base;
  cudaStream_t s = arena.stream;
  for (auto& it : items) {
    *it.slot = cur;
    if (it.hsrc && it.bytes)
      if (int rc = cuda_fail(cudaMemcpyAsync(cur, it.hsrc, it.bytes, cudaMemcpyHostToDevice, s), "H2D")) return rc;
    cur += (it.bytes + 255) / 256 * 256;
  }
  d.workspace = ws ? cur : nullptr;
  d.workspace_bytes = ws;
  if (int rc = run_device(sampler, p, &d, s)) return rc;
  for (auto& it : items)
    if (it.hdst && it.bytes)
      if (int rc = cuda_fail(cudaMemcpyAsync(it.hdst, *it.slot, it.bytes, cudaMemcpyDeviceToHost, s), "D2H")) return rc;
  return cuda_fail(cudaStreamSynchronize(s), "mcmc_run_host sync");
}

}  // extern "C"
