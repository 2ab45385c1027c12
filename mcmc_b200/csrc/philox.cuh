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

// ---------------------------------------------------------------------------------------------------------------
// Batched ("vertical") draw: NB Philox blocks and NB Box-Muller pairs advanced stage by stage, every stage looped
// over the NB independent chains.  One block is a ~100-cycle dependent chain (10 rounds of IMAD.WIDE -> LOP3), one
// pair a ~350-cycle one (I2F -> table -> 8 DFMA -> rsqrt -> 5 DFMA ...; a DFMA is 8.2 cycles dependent, 2 cycles issue,
// profiles/fp64_latency_r02.txt); written pair after pair, a warp with one or two companions on its scheduler waits out
// most of those latencies.  Same operations per element in the same order as philox4x32_10 / normal_pair above:
// bit-identical values (the full-size slice tests compare the kernels' draws with mcmc_philox_fill, which uses the
// scalar path).
template <int NB>
__device__ __forceinline__ void philox4x32_10_batch(uint32_t gid_lo, uint32_t gid_hi, uint32_t it, uint32_t c3_first,
                                                    uint32_t k0, uint32_t k1, U4 (&o)[NB]) {
  uint32_t c0[NB], c1[NB], c2[NB], c3[NB];
#pragma unroll
  for (int b = 0; b < NB; ++b) { c0[b] = gid_lo; c1[b] = gid_hi; c2[b] = it; c3[b] = c3_first + (uint32_t)b; }
#pragma unroll
  for (int r = 0; r < 10; ++r) {
#pragma unroll
    for (int b = 0; b < NB; ++b) {
      const uint32_t hi0 = __umulhi(kPhiloxM0, c0[b]), lo0 = kPhiloxM0 * c0[b];
      const uint32_t hi1 = __umulhi(kPhiloxM1, c2[b]), lo1 = kPhiloxM1 * c2[b];
      const uint32_t n0 = hi1 ^ c1[b] ^ k0, n2 = hi0 ^ c3[b] ^ k1;
      c0[b] = n0; c1[b] = lo1; c2[b] = n2; c3[b] = lo0;
    }
    k0 += kPhiloxW0; k1 += kPhiloxW1;
  }
#pragma unroll
  for (int b = 0; b < NB; ++b) o[b] = U4{c0[b], c1[b], c2[b], c3[b]};
}

// z[2b], z[2b+1] = normal_pair(o[b]) for b < NB, stage by stage (see fastmath.cuh for the scalar functions restated here)
template <int NB>
__device__ __forceinline__ void normal_pairs_batch(const U4 (&o)[NB], const double* __restrict__ tab, double (&z)[2 * NB]) {
  const double* __restrict__ sctab = tab + MCMC_SCTAB_OFFSET;
  double f[NB], ed[NB], invc[NB], logc[NB], r[NB], pl[NB], lg[NB];
  double t[NB], si[NB], ci[NB];
  // stage 1: integer range reduction of the log argument and of the angle
#pragma unroll
  for (int b = 0; b < NB; ++b) {
    const uint64_t n = u53(o[b].x, o[b].y) + 1ull;                       // uniform in (0,1]
    const uint64_t xb = double_to_bits((double)(long long)n);
    const uint32_t xhi = (uint32_t)(xb >> 32);
    f[b] = bits_to_double((xb & 0x000FFFFFFFFFFFFFull) | 0x3FF0000000000000ull);
    const int j = (int)((xhi >> 13) & 127u);
    ed[b] = (double)(((int)(xhi >> 20) - (1023 + 53)) + (j >> 6));
    invc[b] = tab[2 * j];
    logc[b] = tab[2 * j + 1];
    const uint64_t k = u53(o[b].z, o[b].w);                              // angle 2*pi*k/2^53
    const uint32_t i = (uint32_t)(k >> 45);
    t[b] = bits_to_double(0x4330000000000000ull | (k & ((1ull << 45) - 1))) - kFm[15];
    si[b] = sctab[2 * i];
    ci[b] = sctab[2 * i + 1];
  }
  // stage 2: log1p core (fast_log_u53)
#pragma unroll
  for (int b = 0; b < NB; ++b) r[b] = fma_(f[b], invc[b], -1.0);
#pragma unroll
  for (int b = 0; b < NB; ++b) pl[b] = fma_(r[b], kFm[0], kFm[1]);
#pragma unroll
  for (int c = 2; c <= 5; ++c) {
#pragma unroll
    for (int b = 0; b < NB; ++b) pl[b] = fma_(r[b], pl[b], kFm[c]);
  }
#pragma unroll
  for (int b = 0; b < NB; ++b) pl[b] = fma_(r[b], pl[b], -0.5);
#pragma unroll
  for (int b = 0; b < NB; ++b) {
    const double r2 = r[b] * r[b];
    const double l1p = fma_(r2, pl[b], r[b]);
    lg[b] = fma_(ed[b], kFm[6], logc[b] + fma_(ed[b], kFm[7], l1p));
  }
  // stage 3: radius = sqrt(-2 log u) (fast_sqrt_pos), interleaved with the sin / cos cores (fast_sincos_tab_u53)
  double x[NB], g[NB], h[NB], t2[NB], ps[NB], pc[NB];
#pragma unroll
  for (int b = 0; b < NB; ++b) {
    x[b] = -2.0 * lg[b];
    double y0;
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y0) : "d"(x[b]));
    g[b] = x[b] * y0;
    h[b] = 0.5 * y0;
    t2[b] = t[b] * t[b];
  }
#pragma unroll
  for (int b = 0; b < NB; ++b) {
    const double rr = fma_(-g[b], h[b], 0.5);
    g[b] = fma_(g[b], rr, g[b]);
    ps[b] = fma_(t2[b], kFm[8], kFm[9]);
    pc[b] = fma_(t2[b], kFm[12], kFm[13]);
  }
#pragma unroll
  for (int b = 0; b < NB; ++b) {
    const double d = fma_(-g[b], g[b], x[b]);
    g[b] = fma_(d, h[b], g[b]);
    ps[b] = fma_(t2[b], ps[b], kFm[10]);
    pc[b] = fma_(t2[b], pc[b], kFm[14]);
  }
#pragma unroll
  for (int b = 0; b < NB; ++b) {
    const double rad = x[b] > 0.0 ? g[b] : 0.0;
    ps[b] = fma_(t2[b], ps[b], kFm[11]);
    const double sd = t[b] * ps[b];
    const double cm = t2[b] * pc[b];
    const double sn = fma_(ci[b], sd, fma_(si[b], cm, si[b]));
    const double cs = fma_(-si[b], sd, fma_(ci[b], cm, ci[b]));
    z[2 * b] = rad * cs;
    z[2 * b + 1] = rad * sn;
  }
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

// the normals of blocks [block0, block0 + NB) of iteration `it`: z[2b], z[2b+1] = philox_normals_block(block0 + b)
template <int NB>
__device__ __forceinline__ void philox_normals_batch(const RngCtx& r, uint32_t it, uint32_t block0, double (&z)[2 * NB]) {
  U4 o[NB];
  philox4x32_10_batch<NB>(r.gid_lo, r.gid_hi, it, (kTagNormal << 24) | block0, r.k0, r.k1, o);
  normal_pairs_batch<NB>(o, r.logtab, z);
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
