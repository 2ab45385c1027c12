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
  const bool unrolled = getenv("MCMC_RWMH_UNROLLED") != nullptr;
  switch (l.p.n_dim) {
    case 20: return unrolled ? launch_rwmh_variant<MCMC_TID, 20, 128>(l) : launch_rwmh_warp_variant<MCMC_TID, 20, 128>(l);
    case 10: return unrolled ? launch_rwmh_variant<MCMC_TID, 10, 128>(l) : launch_rwmh_warp_variant<MCMC_TID, 10, 128>(l);
    default: return launch_rwmh_variant<MCMC_TID, 0, 32>(l);
  }
}

}  // namespace mcmc
