// rwmh_inst.cu — instantiates the RWMH kernels of ONE target (-DMCMC_TID=k).
#include "rwmh_kernel.cuh"

#ifndef MCMC_TID
#error "compile with -DMCMC_TID=<target id>"
#endif

namespace mcmc {

template <>
int launch_rwmh<MCMC_TID>(const Launch& l) {
  switch (l.p.n_dim) {
    case 20: return launch_rwmh_variant<MCMC_TID, 20, 128>(l);
    case 10: return launch_rwmh_variant<MCMC_TID, 10, 128>(l);
    default: return launch_rwmh_variant<MCMC_TID, 0, 32>(l);
  }
}

}  // namespace mcmc
