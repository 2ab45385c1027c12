// common.cuh — shared declarations of the CUDA side (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

#include "../../include/mcmc_b200.h"

// Two builds of the same sources:
//   libmcmc_b200.so         fast: FMA contraction on, per-dimension constants folded
//                           (eps*w), consecutive half-kicks merged.
//   libmcmc_b200_strict.so  -DMCMC_FAITHFUL -fmad=false: the reference's operation order,
//                           one rounding per jnp op; bit-comparable with the C oracle
//                           (oracle/mcmc_oracle.c, -ffp-contract=off) up to libm differences.
#ifdef MCMC_FAITHFUL
#define MCMC_IS_FAITHFUL 1
#else
#define MCMC_IS_FAITHFUL 0
#endif

#define MCMC_DYN_MAXD MCMC_MAX_DIM
#define MCMC_TARGET_USER_TID_ 9   // template id the sampler units of a user-target plugin are compiled with
#define MCMC_MAX_STEPS 2048  // friction table lives in shared memory

namespace mcmc {

constexpr double kLog2Pi = 1.8378770664093454836;
constexpr double kPi = 3.14159265358979323846;
constexpr double kTwoM53 = 1.1102230246251565404e-16;

// thread-local error string (abi.cu)
void set_error(const char* fmt, ...);

struct Launch {
  McmcParams p;
  McmcBuffers b;
  cudaStream_t stream;
};

// jnp.minimum propagates NaN; fmin() does not.
// Long sums (kinetic energy, log-density values, U-turn dot products) are serial DFMA chains: D dependent
// latencies per sum, which is what a NUTS trip with few active lanes waits for.  The fast build keeps kSumWays
// interleaved partial sums (d % kSumWays) and adds them pairwise; the strict build keeps the reference's
// left-to-right order (kSumWays == 1).
#if MCMC_IS_FAITHFUL
constexpr int kSumWays = 1;
#else
constexpr int kSumWays = 4;
#endif
__device__ __forceinline__ double sum_ways(const double (&p)[kSumWays]) {
  if (kSumWays == 4) return (p[0] + p[1 % kSumWays]) + (p[2 % kSumWays] + p[3 % kSumWays]);
  double s = p[0];
  for (int i = 1; i < kSumWays; ++i) s += p[i];
  return s;
}

__device__ __forceinline__ double min_nan(double a, double b) {
  return (a != a) ? a : ((b != b) ? b : (a < b ? a : b));
}

__device__ __forceinline__ unsigned int ld_acquire_u32(const unsigned int* p) {
  unsigned int v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_u32(unsigned int* p, unsigned int v) {
  asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// Per-target launchers, one explicit specialisation per translation unit
// (hmc_inst.cu / rwmh_inst.cu / nuts_inst.cu compiled with -DMCMC_TID=k).
template <int TID> int launch_hmc(const Launch& l, bool friction);
template <int TID> int launch_rwmh(const Launch& l);
template <int TID> int launch_nuts(const Launch& l);
template <int TID> int launch_target_eval(const McmcParams& p, const double* cov_inv, const double* q,
                                          int64_t n, double* lp, double* grad, cudaStream_t s);
// the trajectory alone (leapfrog_kernel.cuh): device pointers of mcmc_leapfrog, instantiated in hmc_inst.cu
struct LeapfrogIO {
  const double *cov_inv, *inv_mass, *q, *mom, *lp, *grad;
  double *q_out, *mom_out, *lp_out, *grad_out;
};
template <int TID> int launch_leapfrog(const McmcParams& p, const LeapfrogIO& io, bool friction, cudaStream_t s);

// workspace layouts (bytes), shared by the launchers and mcmc_workspace_bytes
struct NutsWs {
  int64_t max_warps; // resident one-warp blocks the runtime-D lane tiles are sized for
  size_t slots_off, home_off, lp_off, acc_off, queue_off, z_off, lu_off, total;
  int64_t tseg;      // iterations per segment of the native draw buffer
};
NutsWs nuts_ws_layout(const McmcParams& p);
size_t nuts_workspace_bytes(const McmcParams& p);
// z[(it - it_begin)*C*D + c*D + d], logu[(it - it_begin)*C + c] for it in [it_begin, it_end): the momentum normals and
// the log of the slice / accept uniform of the native Philox stream (counter layout of philox.cuh)
int launch_rng_fill(const McmcParams& p, int64_t it_begin, int64_t it_end, double* z, double* logu, cudaStream_t s);
size_t hmc_workspace_bytes(const McmcParams& p);
int device_sm_count();
// MCMC_NO_SPIN=1: no in-kernel wait on another block's progress.  The chunked schedulers of hmc_kernel.cuh and
// nuts_kernel.cuh let an item wait (acquire on a flag) for the item that precedes it on the same chains; that earlier
// item is always held by a RESIDENT block because the grids are sized from the occupancy query — true on a GPU the
// library has to itself, not guaranteed under MPS, a debugger or a concurrent kernel that takes SM resources first.
// With the switch every chain group / chain is one item (n_chunks = 1): nothing ever waits; the price is the tail
// of the last wave (~10% at 65,536 chains).  Results are bit-identical either way.
bool no_spin();

inline int cuda_fail(cudaError_t e, const char* what) {
  if (e == cudaSuccess) return 0;
  set_error("%s: %s", what, cudaGetErrorString(e));
  return MCMC_ECUDA;
}

}  // namespace mcmc
