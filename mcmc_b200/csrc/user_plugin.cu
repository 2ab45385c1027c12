// user_plugin.cu — registration stub of a user-target plugin library (mcmc_b200/custom.py).
//
// The plugin is hmc_inst.cu + rwmh_inst.cu + nuts_inst.cu compiled with -DMCMC_TID=9 and
// -DMCMC_USER_TARGET_HEADER="<file defining mcmc_user_log_prob_grad>", plus this file; it links against
// libmcmc_b200.so and hands its five launchers to mcmc_register_user_target, which returns the target id.
#include "common.cuh"

namespace mcmc {
template <> int launch_hmc<MCMC_TARGET_USER_TID_>(const Launch&, bool);
template <> int launch_rwmh<MCMC_TARGET_USER_TID_>(const Launch&);
template <> int launch_nuts<MCMC_TARGET_USER_TID_>(const Launch&);
template <> int launch_target_eval<MCMC_TARGET_USER_TID_>(const McmcParams&, const double*, const double*, int64_t, double*, double*,
                                                         cudaStream_t);
template <> int launch_leapfrog<MCMC_TARGET_USER_TID_>(const McmcParams&, const LeapfrogIO&, bool, cudaStream_t);
}  // namespace mcmc

extern "C" int mcmc_user_plugin_register(void) {
  using namespace mcmc;
  int (*h)(const Launch&, bool) = &launch_hmc<MCMC_TARGET_USER_TID_>;
  int (*r)(const Launch&) = &launch_rwmh<MCMC_TARGET_USER_TID_>;
  int (*n)(const Launch&) = &launch_nuts<MCMC_TARGET_USER_TID_>;
  int (*e)(const McmcParams&, const double*, const double*, int64_t, double*, double*, cudaStream_t) =
      &launch_target_eval<MCMC_TARGET_USER_TID_>;
  int (*l)(const McmcParams&, const LeapfrogIO&, bool, cudaStream_t) = &launch_leapfrog<MCMC_TARGET_USER_TID_>;
  return mcmc_register_user_target(reinterpret_cast<void*>(h), reinterpret_cast<void*>(r), reinterpret_cast<void*>(n),
                                   reinterpret_cast<void*>(e), reinterpret_cast<void*>(l));
}
