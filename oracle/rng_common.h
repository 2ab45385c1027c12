/*
 * rng_common.h — definition of the native counter-based random stream, shared
 * by the CPU oracle (oracle/mcmc_oracle.c) and, as a specification, by the CUDA
 * kernels (mcmc_b200/csrc/philox.cuh restates it for the device).
 *
 * TEST INFRASTRUCTURE (oracle side).  The reference draws from JAX's
 * threefry2x32 (samplers/HMC.py:166-172,205 etc.); jax is not installed here,
 * so the native stream is Philox4x32-10 (Salmon et al., SC'11) with this
 * counter layout (north_star: "a counter-based Philox stream draws momenta and
 * uniforms"):
 *
 *   key  = params.seed[0..1]
 *   ctr  = { gid_lo, gid_hi, iteration, (tag << 24) | block }
 *   gid  = chain_offset + local chain index      (shard-invariant)
 *   tag  = 0: N(0,1) draws, block j yields z[2j], z[2j+1]   (Box-Muller)
 *          1: the accept / slice uniform (block 0, words 0-1)
 *          2: NUTS doubling `block`=depth: words 0-1 direction uniform, words 2-3 proposal uniform
 *
 *   53-bit integer k = (w_hi << 21) | (w_lo >> 11)
 *   uniform in [0,1):  k * 2^-53          (accept / direction / proposal draws)
 *   uniform in (0,1]:  (k + 1) * 2^-53    (argument of the log in Box-Muller)
 *   z0 = sqrt(-2 log u1) * cospi(2 u2),  z1 = sqrt(-2 log u1) * sinpi(2 u2)
 */
#ifndef MCMC_RNG_COMMON_H_
#define MCMC_RNG_COMMON_H_
#include <stdint.h>

#define PHILOX_M0 0xD2511F53u
#define PHILOX_M1 0xCD9E8D57u
#define PHILOX_W0 0x9E3779B9u
#define PHILOX_W1 0xBB67AE85u

#define MCMC_TAG_NORMAL 0u
#define MCMC_TAG_ACCEPT 1u
#define MCMC_TAG_NUTS 2u

static inline void philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]) {
  uint32_t c0 = ctr[0], c1 = ctr[1], c2 = ctr[2], c3 = ctr[3];
  uint32_t k0 = key[0], k1 = key[1];
  for (int r = 0; r < 10; ++r) {
    uint64_t p0 = (uint64_t)PHILOX_M0 * c0;
    uint64_t p1 = (uint64_t)PHILOX_M1 * c2;
    uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
    uint32_t n1 = (uint32_t)p1;
    uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
    uint32_t n3 = (uint32_t)p0;
    c0 = n0; c1 = n1; c2 = n2; c3 = n3;
    k0 += PHILOX_W0; k1 += PHILOX_W1;
  }
  out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

static inline uint64_t mcmc_u53(uint32_t hi, uint32_t lo) {
  return ((uint64_t)hi << 21) | (uint64_t)(lo >> 11);
}
#define MCMC_TWO_M53 1.1102230246251565404e-16 /* 2^-53 */

#endif
