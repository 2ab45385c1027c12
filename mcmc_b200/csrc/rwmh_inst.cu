// rwmh_inst.cu — instantiates the RWMH kernels of ONE target (-DMCMC_TID=k).
#include <cstdlib>

#include "rwmh_kernel.cuh"

#ifndef MCMC_TID
#error "compile with -DMCMC_TID=<target id>"
#endif

namespace mcmc {

template <>
int launch_rwmh<MCMC_TID>(const Launch& l) {
  // MCMC_RWMH_UNROLLED=1: the register-resident kernel (A/B experiments; bit-identical results)
  // the warp kernel hands its sample tiles to the TMA engine (cp.async.bulk needs 16-byte aligned global addresses);
  // a samples pointer that is only 8-byte aligned takes the register-resident kernel, whose tile writer checks the
  // alignment per block and falls back to plain stores
  const bool misaligned = (reinterpret_cast<uintptr_t>(l.b.samples) & 15u) != 0;
  const bool unrolled = misaligned || getenv("MCMC_RWMH_UNROLLED") != nullptr;
  switch (l.p.n_dim) {
#ifndef MCMC_USER_GENERIC_ONLY
    case 20: return unrolled ? launch_rwmh_variant<MCMC_TID, 20, 128>(l) : launch_rwmh_warp_variant<MCMC_TID, 20, 128>(l);
    case 10: return unrolled ? launch_rwmh_variant<MCMC_TID, 10, 128>(l) : launch_rwmh_warp_variant<MCMC_TID, 10, 128>(l);
#endif
    default: return launch_rwmh_variant<MCMC_TID, 0, 32>(l);
  }
}

}  // namespace mcmc
