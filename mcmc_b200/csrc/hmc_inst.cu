// hmc_inst.cu — instantiates the HMC/GRAHMC kernels of ONE target (-DMCMC_TID=k),
// so the nine targets compile in parallel.
#include <cstdlib>
#include <cstring>

#include "hmc_kernel.cuh"
#include "hmc_ws_kernel.cuh"

#ifndef MCMC_TID
#error "compile with -DMCMC_TID=<target id>"
#endif

namespace mcmc {

template <>
int launch_hmc<MCMC_TID>(const Launch& l, bool friction) {
#if MCMC_TID == 6
  // tuning knob (experiments only): MCMC_HMC_VARIANT=<block>x<min blocks per SM>
  if (l.p.n_dim == 20 && !friction) {
    const char* v = getenv("MCMC_HMC_VARIANT");
    if (v) {
      if (!strcmp(v, "64x5")) return launch_hmc_variant<MCMC_TID, 20, false, 64, 5>(l);
      if (!strcmp(v, "64x6")) return launch_hmc_variant<MCMC_TID, 20, false, 64, 6>(l);
      if (!strcmp(v, "64x8")) return launch_hmc_variant<MCMC_TID, 20, false, 64, 8>(l);
      if (!strcmp(v, "32x8")) return launch_hmc_variant<MCMC_TID, 20, false, 32, 8>(l);
      if (!strcmp(v, "32x12")) return launch_hmc_variant<MCMC_TID, 20, false, 32, 12>(l);
      if (!strcmp(v, "32x16")) return launch_hmc_variant<MCMC_TID, 20, false, 32, 16>(l);
      if (!strcmp(v, "128x2")) return launch_hmc_variant<MCMC_TID, 20, false, 128, 2>(l);
      if (!strcmp(v, "128x4")) return launch_hmc_variant<MCMC_TID, 20, false, 128, 4>(l);
    }
  }
#endif
  // large batches under the native stream: the warp-specialised kernel (producer warps draw the momenta)
  if (l.p.n_dim == 20 || l.p.n_dim == 10) {
    bool used = false;
    int rc;
    if (l.p.n_dim == 20) rc = friction ? launch_hmc_ws<MCMC_TID, 20, true>(l, &used) : launch_hmc_ws<MCMC_TID, 20, false>(l, &used);
    else rc = friction ? launch_hmc_ws<MCMC_TID, 10, true>(l, &used) : launch_hmc_ws<MCMC_TID, 10, false>(l, &used);
    if (rc != MCMC_OK || used) return rc;
  }
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
