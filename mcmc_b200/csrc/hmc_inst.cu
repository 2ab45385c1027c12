// hmc_inst.cu — instantiates the HMC/GRAHMC kernels of ONE target (-DMCMC_TID=k),
// so the nine targets compile in parallel.
#include "hmc_kernel.cuh"

#ifndef MCMC_TID
#error "compile with -DMCMC_TID=<target id>"
#endif

namespace mcmc {

template <>
int launch_hmc<MCMC_TID>(const Launch& l, bool friction) {
  switch (l.p.n_dim) {
    case 20:
      return friction ? launch_hmc_variant<MCMC_TID, 20, true, 64>(l) : launch_hmc_variant<MCMC_TID, 20, false, 64>(l);
    case 10:
      return friction ? launch_hmc_variant<MCMC_TID, 10, true, 64>(l) : launch_hmc_variant<MCMC_TID, 10, false, 64>(l);
    default:
      return friction ? launch_hmc_variant<MCMC_TID, 0, true, 32>(l) : launch_hmc_variant<MCMC_TID, 0, false, 32>(l);
  }
}

}  // namespace mcmc
