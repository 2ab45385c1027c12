/*
 * mcmc_b200.h — C ABI of the B200-native sampling hot path.
 *
 * The reference (Chuxuan-a/mcmc) is pure Python/JAX and has no FFI of its own;
 * the drop-in boundary is therefore the call signature of its four *_run
 * functions.  Each entry point below replaces one of them and is what a
 * jax.ffi / ctypes / cgo binding for that function would bind:
 *
 *   mcmc_hmc_run     <- samplers/HMC.py:223-314     (hmc_run; hmc_step :138-219; leapfrog :86-134)
 *   mcmc_grahmc_run  <- samplers/GRAHMC.py:353-456  (rahmc_run; rahmc_step :265-349;
 *                                                    _trajectory_with_schedule :215-261;
 *                                                    _conformal_leapfrog_step :152-211;
 *                                                    friction schedules :69-106)
 *   mcmc_rwmh_run    <- samplers/RWMH.py:123-178    (rwMH_run; rwMH_step :73-119)
 *   mcmc_nuts_run    <- samplers/NUTS.py:446-514    (nuts_run; nuts_step :399-442;
 *                                                    _nuts_step_single_chain :202-396)
 *   mcmc_target_eval <- vmap(jax.value_and_grad(log_prob_fn)) call sites
 *                       HMC.py:73,123  GRAHMC.py:140,203  NUTS.py:77,111  RWMH.py:67,102
 *                       over the closures of benchmarks/targets.py:38-482
 *   mcmc_friction_table <- samplers/GRAHMC.py:69-106,193-194,246-252
 *   mcmc_welford_finalize / mcmc_chain_moments <- tuning/adaptation.py:229-265
 *                       (per-chain Welford, mean of per-chain variances) and
 *                       tuning/welford.py:42-100
 *
 * Conventions
 *  - plain C, POD structs, raw pointers and sizes; no torch / C++ types.
 *  - every entry returns 0 on success, a negative MCMC_E* code otherwise;
 *    mcmc_last_error() gives a thread-local message.  No exception crosses
 *    the ABI.  There is NO CPU fallback: without a CUDA device every compute
 *    entry fails with MCMC_ECUDA.
 *  - *_run entries take DEVICE pointers, are asynchronous on `stream`
 *    (a cudaStream_t passed as void*), allocate nothing and never sync the
 *    host.  mcmc_run_host is the same path with HOST pointers (it owns the
 *    H2D/D2H copies and a cached device arena).
 *  - alignment: every pointer must be aligned to its element type; samples / init_q aligned to 16 bytes take
 *    the TMA bulk-store / 16-byte vector paths, pointers that are only 8-byte aligned take plain stores
 *    (same results).
 *  - co-residency: with more chains than resident threads the HMC/GRAHMC and NUTS launchers cut the run into
 *    (chain group, iteration chunk) items claimed from an atomic queue; an item spins (acquire on a device flag)
 *    until the item before it on the same chains has published its state.  That earlier item is always held by a
 *    resident block because the grid is sized from cudaOccupancyMaxActiveBlocksPerMultiprocessor, which assumes
 *    the library has the GPU's SM resources to itself while a sampler kernel runs.  Under MPS, a debugger, or a
 *    concurrent kernel on another stream that occupies shared memory / registers first, set the environment
 *    variable MCMC_NO_SPIN=1: every chain (group) is then ONE item, nothing waits, results are bit-identical,
 *    and the last wave's tail (about 10% at 65,536 chains) is the price.
 *  - all arrays are row-major with the reference's shapes: samples (S,C,D),
 *    log_probs (S,C), state (C,D) / (C,), everything fp64 except
 *    accept_count (int32), accept_rate (float32), tree_depths (int32).
 */
#ifndef MCMC_B200_H_
#define MCMC_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MCMC_ABI_VERSION 2

/* error codes */
#define MCMC_OK 0
#define MCMC_EINVAL (-1)   /* bad argument (the reference's ValueError / KeyError) */
#define MCMC_ECUDA (-2)    /* CUDA runtime failure or no device */
#define MCMC_ENOTSUP (-3)  /* unsupported combination (e.g. D too large) */
#define MCMC_EWORKSPACE (-4) /* workspace missing or too small */

/* samplers */
enum { MCMC_SAMPLER_HMC = 0, MCMC_SAMPLER_GRAHMC = 1, MCMC_SAMPLER_RWMH = 2, MCMC_SAMPLER_NUTS = 3 };

/* targets: benchmarks/targets.py factories, in file order (+ a general dense Gaussian) */
enum {
  MCMC_TARGET_STANDARD_NORMAL = 0,   /* targets.py:50-52;  params[0]: 1 -> with the D*log(2pi) term (default), 0 -> README.md:51 variant */
  MCMC_TARGET_CORRELATED_GAUSSIAN = 1, /* targets.py:79-98; params[0]=rho (closed form of the compound-symmetry precision) */
  MCMC_TARGET_ILL_CONDITIONED = 2,   /* targets.py:125-137; params[0]=kappa */
  MCMC_TARGET_NEALS_FUNNEL = 3,      /* targets.py:166-185 */
  MCMC_TARGET_LOG_GAMMA = 4,         /* targets.py:236-251; params[0]=alpha, [1]=beta, [2]=lgamma(alpha)+alpha*log(beta) */
  MCMC_TARGET_STUDENT_T = 5,         /* targets.py:300-311; params[0]=df, [1]=log normaliser per dim */
  MCMC_TARGET_ROSENBROCK = 6,        /* targets.py:367-387; params[0]=scale (a = 1/scale^2) */
  MCMC_TARGET_GAUSSIAN_MIXTURE = 7,  /* targets.py:432-454; params[0]=separation */
  MCMC_TARGET_DENSE_GAUSSIAN = 8,    /* general x^T P x with P = buffers.cov_inv (D*D); params[0]=logdet(Sigma) */
  MCMC_NUM_TARGETS = 9,              /* built-in targets */
  /* 9 .. 16: user targets registered at run time (mcmc_register_user_target): value + gradient supplied as CUDA source,
     compiled into the same kernels as one more per-target translation unit (mcmc_b200/custom.py) */
  MCMC_TARGET_USER0 = 9,
  MCMC_MAX_USER_TARGETS = 8
};

/* friction schedules: samplers/GRAHMC.py:113-119 registry order */
enum {
  MCMC_SCHED_CONSTANT = 0, MCMC_SCHED_TANH = 1, MCMC_SCHED_SIGMOID = 2,
  MCMC_SCHED_LINEAR = 3, MCMC_SCHED_SINE = 4, MCMC_NUM_SCHEDULES = 5
};

enum { MCMC_RNG_PHILOX = 0, MCMC_RNG_INJECTED = 1 };

/* flags */
#define MCMC_FLAG_TRACK_PROPOSALS 1u  /* HMC/GRAHMC: fill pre_q/pre_lp/prop_q/prop_lp/delta_h (HMC.py:287-305) */
#define MCMC_FLAG_ACCUM_WELFORD 2u    /* update welford_mean/m2 with every stored sample (adaptation.py:229-235) */
#define MCMC_FLAG_NO_SAMPLES 4u       /* do not store samples/log_probs (warmup batches only need the moments) */

#define MCMC_MAX_DIM 128
#define MCMC_MAX_TREE_DEPTH 20

typedef struct McmcParams {
  uint32_t abi_version;     /* MCMC_ABI_VERSION */
  int32_t target_id;
  int32_t schedule_id;      /* GRAHMC only */
  int32_t rng_mode;
  double target_params[8];
  int64_t n_chains;         /* C: chains in THIS shard */
  int64_t num_samples;      /* S */
  int64_t burn_in;
  int64_t chain_offset;     /* global index of this shard's first chain: Philox streams are keyed on
                               chain_offset + local index, so results do not depend on the sharding */
  int32_t n_dim;            /* D */
  int32_t num_steps;        /* L (HMC/GRAHMC) */
  int32_t max_tree_depth;   /* NUTS */
  uint32_t flags;
  double step_size;         /* epsilon (HMC/GRAHMC/NUTS); proposal scale (RWMH) */
  double gamma;             /* GRAHMC friction amplitude */
  double steepness;         /* GRAHMC tanh/sigmoid steepness */
  double delta_max;         /* NUTS divergence threshold */
  uint32_t seed[2];         /* Philox key (the reference's uint32[2] PRNG key) */
  uint32_t reserved[2];
} McmcParams;

typedef struct McmcBuffers {
  /* inputs */
  const double* init_q;     /* [C*D] */
  const double* inv_mass;   /* [D] diagonal of M^-1, NULL -> ones (HMC.py:264-265) */
  const double* cov_inv;    /* [D*D], MCMC_TARGET_DENSE_GAUSSIAN only */
  const double* z_inject;   /* [(burn+S)*C*D] standard normals, rng_mode INJECTED */
  const double* u_inject;   /* [(burn+S)*C]   uniforms [0,1),   rng_mode INJECTED */
  const double* nuts_dir_u; /* [(burn+S)*C*max_tree_depth] uniforms: direction=+1 iff u<0.5 (NUTS.py:255-256) */
  const double* nuts_prop_u;/* [(burn+S)*C*max_tree_depth] uniforms for the proposal swap (NUTS.py:336) */
  /* outputs (any may be NULL unless stated) */
  double* samples;          /* [S*C*D] post-MH positions */
  double* log_probs;        /* [S*C] */
  int32_t* accept_count;    /* [C] */
  float* accept_rate;       /* [C] = float(accept_count)/float(S) */
  double* final_q;          /* [C*D] */
  double* final_lp;         /* [C] */
  double* final_g;          /* [C*D] (not RWMH) */
  double* pre_q;            /* [S*C*D] tracking */
  double* pre_lp;           /* [S*C] */
  double* prop_q;           /* [S*C*D] */
  double* prop_lp;          /* [S*C] */
  double* delta_h;          /* [S*C] */
  int32_t* tree_depths;     /* [S*C] NUTS */
  double* mean_accept_probs;/* [S*C] NUTS */
  /* in/out */
  double* welford_mean;     /* [C*D] per-chain running mean  (MCMC_FLAG_ACCUM_WELFORD) */
  double* welford_m2;       /* [C*D] per-chain running M2 */
  double welford_count;     /* samples already folded into welford_* before this call */
  /* device counters, fused in-kernel reductions (exact integer atomics):
     [0] sum accept_count over chains, [1] sum of gradient evaluations,
     [2] sum of tree depths (NUTS), [3] number of divergent transitions (NUTS) */
  unsigned long long* counters; /* [4], zeroed by the callee at launch */
  void* workspace;          /* mcmc_workspace_bytes() bytes of device scratch */
  size_t workspace_bytes;
  /* optional input (ABI 2): when not NULL the kernels read epsilon / the RWMH scale from this DEVICE double instead
     of McmcParams.step_size — the dual-averaging state lives on the device (mcmc_da_update, element 5) and a whole
     warmup is enqueued without a host round-trip */
  const double* step_size_dev;
} McmcBuffers;

int mcmc_abi_version(void);
/* "mcmc_b200 sm_100a fast|strict [+ws]": which build of the sources this library is */
const char* mcmc_build_info(void);
const char* mcmc_last_error(void);
/* number of CUDA devices usable by this library (0 -> every compute entry fails) */
int mcmc_device_count(void);
/* make `device` current for this thread's calls into the library (the library links its own
   CUDA runtime; a host framework's cudaSetDevice is not guaranteed to be seen by it) */
int mcmc_set_device(int device);

/* device scratch the sampler needs for these params (NUTS tree ends, friction table ...) */
int mcmc_workspace_bytes(int sampler, const McmcParams* p, size_t* bytes);

int mcmc_hmc_run(const McmcParams* p, const McmcBuffers* b, void* stream);
int mcmc_grahmc_run(const McmcParams* p, const McmcBuffers* b, void* stream);
int mcmc_rwmh_run(const McmcParams* p, const McmcBuffers* b, void* stream);
int mcmc_nuts_run(const McmcParams* p, const McmcBuffers* b, void* stream);

/* Same path with HOST pointers in McmcBuffers (workspace ignored): copies the
   inputs to a cached device arena, runs the sampler on the library's own
   stream, copies every non-NULL output back and synchronises.  `device` is the
   CUDA ordinal. */
int mcmc_run_host(int sampler, const McmcParams* p, const McmcBuffers* host, int device);
/* free the cached arena of mcmc_run_host */
int mcmc_host_arena_release(void);

/* Extension point for densities outside benchmarks/targets.py (the reference's *_run accept ANY jax-differentiable
   log_prob_fn: HMC.py:223, README.md:51-52; here the density must exist as device code).  A plugin library — the sampler
   translation units compiled with -DMCMC_TID=9 and the user's `mcmc_user_log_prob_grad` (mcmc_b200/custom.py drives nvcc) —
   hands its five launchers (opaque C++ function pointers of this library's internal types) to this entry and receives the
   target id to put into McmcParams.target_id (MCMC_TARGET_USER0 + slot), or a negative MCMC_E* code. */
int mcmc_register_user_target(void* launch_hmc, void* launch_rwmh, void* launch_nuts, void* launch_eval, void* launch_leapfrog);
/* log p and grad log p of n points (device pointers; lp or grad may be NULL) */
int mcmc_target_eval(const McmcParams* p, const double* cov_inv, const double* q,
                     int64_t n, double* lp, double* grad, void* stream);

/* The trajectory alone: num_steps leapfrog steps from (q, momentum, log_prob, grad) for every chain (device pointers,
   [C][D] rows / [C]; asynchronous on `stream`; no allocation).  friction == 0 replaces `leapfrog` (samplers/HMC.py:
   85-134: half-kick, drift with the diagonal inverse mass, value_and_grad, half-kick), friction != 0 replaces
   `_trajectory_with_schedule` (samplers/GRAHMC.py:215-261) over `_conformal_leapfrog_step` (:152-211) with
   s_i = exp(-gamma(i eps) eps/2) from p->schedule_id / gamma / steepness.  Uses p->target_id, target_params, n_chains,
   n_dim, num_steps, step_size.  inv_mass NULL = ones; log_prob NULL = 0 (only returned unchanged when num_steps == 0);
   any output may be NULL or alias its input.  num_steps == 0 copies the inputs (a scan of length 0). */
int mcmc_leapfrog(const McmcParams* p, int friction, const double* cov_inv, const double* inv_mass, const double* q,
                  const double* momentum, const double* log_prob, const double* grad, double* q_out, double* momentum_out,
                  double* log_prob_out, double* grad_out, void* stream);

/* The native random stream as a callable (the draws the samplers make in-kernel; counter layout in
   mcmc_b200/csrc/philox.cuh and oracle/rng_common.h).  Replaces the reference's random.normal(k_momentum, (C, D)) /
   random.uniform(k_accept, (C,)) call sites (HMC.py:172,205  GRAHMC.py:306,336  RWMH.py:98,106  NUTS.py:219,224):
   for it in [it_begin, it_begin + n_it) and the p->n_chains chains of this shard (p->chain_offset, p->seed, p->n_dim)
   z[((it - it_begin)*C + c)*D + d] = the N(0,1) momentum / proposal draws, log_u[(it - it_begin)*C + c] = log of the
   accept / slice uniform (-inf for u = 0).  Device pointers, either may be NULL. */
int mcmc_philox_fill(const McmcParams* p, int64_t it_begin, int64_t n_it, double* z, double* log_u, void* stream);

/* s_i = exp(-gamma_i * eps/2), i = 0..L-1, written to a DEVICE array of L doubles
   (and gamma_i to gamma_out if not NULL) */
int mcmc_friction_table(int schedule_id, int num_steps, double step_size, double gamma,
                        double steepness, double* s_out, double* gamma_out, void* stream);

/* Cross-chain reductions used by the warmup (tuning/adaptation.py:238-265).
   mcmc_welford_finalize: out[d] = sum_c m2[c,d]/(max(count,2)-1), d < D  (device pointers;
   the caller divides by the global chain count after its all-reduce).
   mcmc_chain_moments: per-chain mean/M2 over the S axis of samples (S,C,D), merged into
   welford_mean/m2 exactly as S sequential welford_update calls would. */
int mcmc_welford_finalize(const double* welford_m2, double count, int64_t n_chains, int32_t n_dim,
                          double* out_sum_var, void* workspace, size_t workspace_bytes, void* stream);
/* mcmc_welford_finalize_noalloc: NO COUNTERPART IN THE REFERENCE — added for
   callers that cannot allocate device memory (code under CUDA graph capture, tight serving loops, harnesses that forbid
   cudaMalloc).  Same result as mcmc_welford_finalize (same kernels, same order of operations: bit-identical), but the
   256 x n_dim block partials live in a file-scope `static __device__` array of the library instead of a caller
   workspace: 256 * MCMC_NOALLOC_MAX_DIM doubles = 256 KiB, held by every GPU that loads the library until exit.
   The entry calls no cudaMalloc / cudaFree / allocating routine; inputs and outputs are the caller's device pointers
   on the CURRENT device.  n_dim > MCMC_NOALLOC_MAX_DIM is rejected with MCMC_ENOTSUP.  The first call on each device
   resolves the array's device address (cudaGetSymbolAddress): make one call per device outside a graph capture.
   All calls on a device share the one array and it keeps the previous call's contents (none of it is read before it is
   overwritten, so nothing needs clearing); two calls that OVERLAP on the GPU (different streams without an ordering
   between them) race on it — serialise them or use mcmc_welford_finalize with separate workspaces. */
#define MCMC_NOALLOC_MAX_DIM 128
int mcmc_welford_finalize_noalloc(const double* welford_m2, double count, int64_t n_chains, int32_t n_dim,
                                  double* out_sum_var, void* stream);
int mcmc_chain_moments(const double* samples, int64_t num_samples, int64_t n_chains, int32_t n_dim,
                       double* welford_mean, double* welford_m2, double count_before, void* stream);
/* Dual averaging on the device (tuning/dual_averaging.py:806-887: da_init / da_update / da_reset).  `state` is a DEVICE
   array of MCMC_DA_STATE_DOUBLES doubles:
     [0] log_step  [1] log_step_bar  [2] H_bar  [3] mu  [4] count m  [5] exp(log_step)  [6] exp(log_step_bar)
     [7] the last accept statistic
   mcmc_da_init: state of da_init(initial_step_size).  mcmc_da_update: accept_stat = accept_sum[0] / n_terms (accept_sum
   a DEVICE double, e.g. the all-reduced in-kernel accept counter; with n_terms == 0 the divisor is accept_sum[1], so
   that the sum and its term count can come out of the same all-reduce), then eta = 1/(m + t0), H_bar <- (1-eta) H_bar +
   eta (target - stat), log_step = mu - sqrt(m)/omega H_bar, log_step_bar <- m^-kappa log_step + (1 - m^-kappa)
   log_step_bar (first update: = log_step).  mcmc_da_reset: restart around the smoothed estimate (mu <- log_step_bar).
   A sampler launched with McmcBuffers.step_size_dev = state + 5 steps with the current epsilon. */
#define MCMC_DA_STATE_DOUBLES 8
int mcmc_da_init(double* state, double initial_step_size, void* stream);
int mcmc_da_update(double* state, const double* accept_sum, double n_terms, double target_accept, double omega, double t0,
                   double kappa, void* stream);
int mcmc_da_reset(double* state, void* stream);
/* deterministic sum of n doubles -> out[0] (device pointers) */
int mcmc_sum_f64(const double* x, int64_t n, double* out, void* workspace, size_t workspace_bytes,
                 void* stream);

/* ---- cross-chain convergence diagnostics (run_benchmarks.py:161-195: az.rhat, az.ess, az.summary; the
   published ArviZ algorithm, see mcmc_b200/csrc/diag.cu).  Device pointers, asynchronous on `stream`.

   mcmc_rank_normalize: ArviZ `_z_scale` of one variable.  `sorted_vals` = its n values in ascending
   order, `perm` = the permutation that sorted them; out_z[perm[i]] = Phi^-1((r_i - 3/8)/(n + 1/4)) with
   r_i the average 1-based rank of sorted_vals[i] over ties.

   mcmc_rank_normalize_fold: both rank-normalisations R-hat needs, from ONE sort done inside (ArviZ `_z_scale(ary)` for
   the bulk and `_z_scale(abs(ary - median(ary)))` for the folded / tail R-hat).  x = the variable's n values in any
   order (not modified); sorted_out = the values ascending (the caller takes the tail quantiles from it); z_bulk[i],
   z_fold[i] = the z-scores of x[i].  Bit-identical to mcmc_rank_normalize applied to a sort of x and to a sort of the
   folded values.  n < 2^31; workspace (256-byte aligned) of mcmc_rank_normalize_fold_workspace(n) bytes.

   mcmc_split_chain_stats: one variable laid out y[s*stride_s + c*stride_c].  Every chain c contributes
   two sequences of half_len draws: rows [0, half_len) and rows [second_half_row, second_half_row +
   half_len).  Outputs: acov_sum[t] = sum over the 2*n_chains sequences of the biased autocovariance at
   lag t (t < n_lags <= half_len); mean_stats[0..2] = (count, mean, M2) of the sequence means.  R-hat and
   ESS follow from these on the host (Geyer truncation) after the sums have been all-reduced over the
   chain shards of a multi-GPU run.  Results are deterministic (fixed-order reductions). */
int mcmc_rank_normalize(const double* sorted_vals, const int64_t* perm, int64_t n, double* out_z, void* stream);
int mcmc_rank_normalize_fold_workspace(int64_t n, size_t* bytes);
int mcmc_rank_normalize_fold(const double* x, int64_t n, double* sorted_out, double* z_bulk, double* z_fold, void* workspace,
                             size_t workspace_bytes, void* stream);
int mcmc_split_chain_stats_workspace(int64_t half_len, int64_t n_chains, int32_t n_lags, size_t* bytes);
int mcmc_split_chain_stats(const double* y, int64_t half_len, int64_t second_half_row, int64_t n_chains,
                           int64_t stride_s, int64_t stride_c, int32_t n_lags, double* acov_sum,
                           double* mean_stats, void* workspace, size_t workspace_bytes, void* stream);

/* register-resident DFMA throughput probe used as the fp64 roofline denominator:
   runs `iters` dependent-chain FMA rounds on every SM and returns flop count and ms */
int mcmc_fp64_peak_probe(int iters, double* flops_out, float* ms_out, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* MCMC_B200_H_ */
