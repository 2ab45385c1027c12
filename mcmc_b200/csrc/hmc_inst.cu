// hmc_inst.cu — instantiates the HMC/GRAHMC kernels of ONE target (-DMCMC_TID=k),
// so the nine targets compile in parallel.
#include "hmc_kernel.cuh"
#include "leapfrog_kernel.cuh"
#ifdef MCMC_WITH_WS
// the warp-specialised variant (measured 4-6% slower, profiles/ncu_hmc_ws_r01.txt) is an experiment: compiled only with
// -DMCMC_WITH_WS (python -m mcmc_b200.csrc.build --suffix _ws -DMCMC_WITH_WS) and then taken with MCMC_HMC_WS=1
#include "hmc_ws_kernel.cuh"
#endif

#ifndef MCMC_TID
#error "compile with -DMCMC_TID=<target id>"
#endif

namespace mcmc {

template <>
int launch_hmc<MCMC_TID>(const Launch& l, bool friction) {
#ifdef MCMC_WITH_WS
  if (l.p.n_dim == 20 || l.p.n_dim == 10) {
    bool used = false;
    int rc;
    if (l.p.n_dim == 20) rc = friction ? launch_hmc_ws<MCMC_TID, 20, true>(l, &used) : launch_hmc_ws<MCMC_TID, 20, false>(l, &used);
    else rc = friction ? launch_hmc_ws<MCMC_TID, 10, true>(l, &used) : launch_hmc_ws<MCMC_TID, 10, false>(l, &used);
    if (rc != MCMC_OK || used) return rc;
  }
#endif
  switch (l.p.n_dim) {
#ifndef MCMC_USER_GENERIC_ONLY   // a user plugin may compile only the runtime-D kernels (seconds instead of minutes)
    case 20:
      return friction ? launch_hmc_variant<MCMC_TID, 20, true, 64>(l) : launch_hmc_variant<MCMC_TID, 20, false, 64>(l);
    case 10:
      return friction ? launch_hmc_variant<MCMC_TID, 10, true, 64>(l) : launch_hmc_variant<MCMC_TID, 10, false, 64>(l);
#endif
    default:
      return friction ? launch_hmc_variant<MCMC_TID, 0, true, 32>(l) : launch_hmc_variant<MCMC_TID, 0, false, 32>(l);
  }
}

template <>
int launch_leapfrog<MCMC_TID>(const McmcParams& p, const LeapfrogIO& io, bool friction, cudaStream_t s) {
  return launch_leapfrog_impl<MCMC_TID>(p, io, friction, s);
}

}  // namespace mcmc
