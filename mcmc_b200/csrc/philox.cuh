// philox.cuh — the native counter-based random stream (device side).
//
// Replaces the reference's jax.random draws (threefry2x32):
//   momenta  random.normal(k_momentum, (C, D))   HMC.py:172  GRAHMC.py:306  NUTS.py:219
//   noise    random.normal(key_noise, (C, D))    RWMH.py:98
//   accept   random.uniform(k_accept, (C,))      HMC.py:205  GRAHMC.py:336  RWMH.py:106  NUTS.py:224
//   NUTS     random.bernoulli / random.uniform   NUTS.py:256,336
// Counter layout (identical to oracle/rng_common.h, which is the specification):
//   key = seed[0..1];  ctr = { gid_lo, gid_hi, iteration, (tag << 24) | block }
#pragma once
#include "common.cuh"
#include "fastmath.cuh"

namespace mcmc {

constexpr uint32_t kPhiloxM0 = 0xD2511F53u, kPhiloxM1 = 0xCD9E8D57u;
constexpr uint32_t kPhiloxW0 = 0x9E3779B9u, kPhiloxW1 = 0xBB67AE85u;
constexpr uint32_t kTagNormal = 0u, kTagAccept = 1u, kTagNuts = 2u;

struct U4 { uint32_t x, y, z, w; };

__device__ __forceinline__ U4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                                            uint32_t k0, uint32_t k1) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint32_t hi0 = __umulhi(kPhiloxM0, c0), lo0 = kPhiloxM0 * c0;
    const uint32_t hi1 = __umulhi(kPhiloxM1, c2), lo1 = kPhiloxM1 * c2;
    const uint32_t n0 = hi1 ^ c1 ^ k0, n2 = hi0 ^ c3 ^ k1;
    c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
    k0 += kPhiloxW0; k1 += kPhiloxW1;
  }
  return U4{c0, c1, c2, c3};
}

__device__ __forceinline__ uint64_t u53(uint32_t hi, uint32_t lo) {
  return ((uint64_t)hi << 21) | (uint64_t)(lo >> 11);
}

// uniform in [0,1)
__device__ __forceinline__ double unif_co(uint32_t hi, uint32_t lo) {
  return (double)(long long)u53(hi, lo) * kTwoM53;
}
// uniform in (0,1]
__device__ __forceinline__ double unif_oc(uint32_t hi, uint32_t lo) {
  return (double)(long long)(u53(hi, lo) + 1ull) * kTwoM53;
}

// two N(0,1) draws from one Philox block (Box-Muller, fp64).  The log / sqrt / sincos are the
// branch-free integer-argument versions of fastmath.cuh (<= 2 ulp from libm's).
__device__ __forceinline__ void normal_pair(const U4& o, const double* __restrict__ logtab, double& z0, double& z1) {
  const double lg = fast_log_u53(u53(o.x, o.y) + 1ull, logtab);   // log of a uniform in (0,1]
  const double rad = fast_sqrt_pos(-2.0 * lg);
  double s, c;
  fast_sincos_tab_u53(u53(o.z, o.w), logtab + MCMC_SCTAB_OFFSET, &s, &c);   // angle 2*pi*u2, u2 in [0,1)
  z0 = rad * c;
  z1 = rad * s;
}

struct RngCtx {
  uint32_t k0, k1;       // key
  uint32_t gid_lo, gid_hi;
  const double* logtab;  // shared-memory copy of kLogTabRaw (log table, then the sin/cos sector table)
};

constexpr int kRngTableDoubles = 768;
// cooperative copy of the tables into shared memory (6 KB); call before __syncthreads()
__device__ __forceinline__ void rng_table_fill(double* sm_logtab) {
  for (int i = threadIdx.x; i < kRngTableDoubles; i += blockDim.x) sm_logtab[i] = kLogTabRaw[i];
}

__device__ __forceinline__ void philox_normals_block(const RngCtx& r, uint32_t it, uint32_t block,
                                                     double& z0, double& z1) {
  normal_pair(philox4x32_10(r.gid_lo, r.gid_hi, it, (kTagNormal << 24) | block, r.k0, r.k1), r.logtab, z0, z1);
}

// log of the accept / slice uniform u in [0,1)  (log 0 = -inf: always accepted, as in the reference)
__device__ __forceinline__ double philox_accept_log_uniform(const RngCtx& r, uint32_t it) {
  const U4 o = philox4x32_10(r.gid_lo, r.gid_hi, it, (kTagAccept << 24), r.k0, r.k1);
  const uint64_t k = u53(o.x, o.y);
  const double lg = fast_log_u53(k | (k == 0), r.logtab);
  return k == 0 ? -INFINITY : lg;
}

__device__ __forceinline__ void philox_nuts_uniforms(const RngCtx& r, uint32_t it, uint32_t depth,
                                                     double& u_dir, double& u_prop) {
  const U4 o = philox4x32_10(r.gid_lo, r.gid_hi, it, (kTagNuts << 24) | depth, r.k0, r.k1);
  u_dir = unif_co(o.x, o.y);
  u_prop = unif_co(o.z, o.w);
}

}  // namespace mcmc
