// mcmc_xla_ffi.cc — XLA FFI (jax.ffi) custom-call handlers over the C ABI of include/mcmc_b200.h.
//
// The reference is a JAX program: its callers (tuning/adaptation.py:178-210, run_benchmarks.py:591-693,
// README.md:58) invoke hmc_run / rahmc_run / rwMH_run / nuts_run inside jit-compiled code.  These handlers are what
// `jax.ffi.register_ffi_target(name, jax.ffi.pycapsule(lib.<Handler>), platform="CUDA")` binds, one per sampler
// (mcmc_b200/ffi.py does the registration and wraps them behind the reference's signatures):
//
//   McmcHmcRun      <- samplers/HMC.py:223-314      McmcGrahmcRun <- samplers/GRAHMC.py:353-456
//   McmcRwmhRun     <- samplers/RWMH.py:123-178     McmcNutsRun   <- samplers/NUTS.py:446-514
//   McmcHmcRunTracked / McmcGrahmcRunTracked: the track_proposals=True variants (HMC.py:287-305, GRAHMC.py:422-446)
//
// The XLA FFI header ships inside jaxlib (jax.ffi.include_dir()); jaxlib is not installable in this image (no wheel, no
// network), so everything below the guard is compiled only where the header exists (mcmc_b200/csrc/build.py:
// build_xla_ffi()).  Without it the translation unit still builds and exports mcmc_xla_ffi_available() == 0, which is
// what the CPU test checks here.  The adapter adds no kernel code: it fills McmcParams / McmcBuffers from the call's
// attributes and buffers, takes the sampler workspace from XLA's scratch allocator (no cudaMalloc in the call) and
// launches on XLA's stream.
#include <cstdint>

#include "../../include/mcmc_b200.h"

#if defined(__has_include)
#if __has_include("xla/ffi/api/ffi.h")
#define MCMC_HAVE_XLA_FFI 1
#endif
#endif

extern "C" int mcmc_xla_ffi_available(void) {
#ifdef MCMC_HAVE_XLA_FFI
  return 1;
#else
  return 0;
#endif
}

#ifdef MCMC_HAVE_XLA_FFI
#include <cuda_runtime_api.h>

#include "xla/ffi/api/ffi.h"

namespace ffi = xla::ffi;

namespace {

using F64 = ffi::Buffer<ffi::F64>;
using RF64 = ffi::ResultBuffer<ffi::F64>;
using RF32 = ffi::ResultBuffer<ffi::F32>;
using RS32 = ffi::ResultBuffer<ffi::S32>;

// attributes every handler takes (static arguments of the jitted reference functions)
struct Common {
  int32_t target_id, schedule_id, num_steps, max_tree_depth;
  int64_t num_samples, burn_in, chain_offset;
  double step_size, gamma, steepness, delta_max;
  int64_t seed0, seed1;                    // the uint32[2] PRNG key, passed as two integer attributes
  ffi::Span<const double> target_params;   // up to 8 doubles (rho, kappa, df, ...)
};

struct Outputs {
  double *samples, *log_probs;
  float* accept_rate;
  double *final_q, *final_lp, *final_g;
  int32_t* accept_count;
  double *pre_q = nullptr, *pre_lp = nullptr, *prop_q = nullptr, *prop_lp = nullptr, *delta_h = nullptr;
  int32_t* tree_depths = nullptr;
  double* mean_accept_probs = nullptr;
};

ffi::Error Run(int sampler, cudaStream_t stream, ffi::ScratchAllocator& scratch, const F64& init_q, const F64& inv_mass,
               const F64& cov_inv, const Common& a, const Outputs& o, uint32_t flags) {
  const auto dims = init_q.dimensions();
  if (dims.size() != 2)   // HMC.py:57
    return ffi::Error(ffi::ErrorCode::kInvalidArgument, "Input must have shape (n_chains, n_dim).");
  McmcParams p{};
  p.abi_version = MCMC_ABI_VERSION;
  p.target_id = a.target_id;
  p.schedule_id = a.schedule_id;
  p.rng_mode = MCMC_RNG_PHILOX;
  for (size_t i = 0; i < 8 && i < a.target_params.size(); ++i) p.target_params[i] = a.target_params[i];
  p.n_chains = dims[0];
  p.n_dim = static_cast<int32_t>(dims[1]);
  p.num_samples = a.num_samples;
  p.burn_in = a.burn_in;
  p.chain_offset = a.chain_offset;
  p.num_steps = a.num_steps;
  p.max_tree_depth = a.max_tree_depth;
  p.flags = flags;
  p.step_size = a.step_size;
  p.gamma = a.gamma;
  p.steepness = a.steepness;
  p.delta_max = a.delta_max;
  p.seed[0] = static_cast<uint32_t>(a.seed0);
  p.seed[1] = static_cast<uint32_t>(a.seed1);

  McmcBuffers b{};
  b.init_q = init_q.typed_data();
  b.inv_mass = inv_mass.element_count() ? inv_mass.typed_data() : nullptr;   // empty -> ones (HMC.py:264-265)
  b.cov_inv = cov_inv.element_count() ? cov_inv.typed_data() : nullptr;
  b.samples = o.samples;
  b.log_probs = o.log_probs;
  b.accept_rate = o.accept_rate;
  b.accept_count = o.accept_count;
  b.final_q = o.final_q;
  b.final_lp = o.final_lp;
  b.final_g = o.final_g;
  b.pre_q = o.pre_q; b.pre_lp = o.pre_lp; b.prop_q = o.prop_q; b.prop_lp = o.prop_lp; b.delta_h = o.delta_h;
  b.tree_depths = o.tree_depths;
  b.mean_accept_probs = o.mean_accept_probs;

  size_t ws = 0;
  if (mcmc_workspace_bytes(sampler, &p, &ws) != MCMC_OK)
    return ffi::Error(ffi::ErrorCode::kInvalidArgument, mcmc_last_error());
  if (ws) {
    auto mem = scratch.Allocate(ws, 256);
    if (!mem.has_value()) return ffi::Error(ffi::ErrorCode::kResourceExhausted, "mcmc_b200: sampler workspace");
    b.workspace = *mem;
    b.workspace_bytes = ws;
  }
  int rc = MCMC_EINVAL;
  switch (sampler) {
    case MCMC_SAMPLER_HMC: rc = mcmc_hmc_run(&p, &b, stream); break;
    case MCMC_SAMPLER_GRAHMC: rc = mcmc_grahmc_run(&p, &b, stream); break;
    case MCMC_SAMPLER_RWMH: rc = mcmc_rwmh_run(&p, &b, stream); break;
    case MCMC_SAMPLER_NUTS: rc = mcmc_nuts_run(&p, &b, stream); break;
  }
  if (rc == MCMC_OK) return ffi::Error::Success();
  return ffi::Error(rc == MCMC_EINVAL ? ffi::ErrorCode::kInvalidArgument : ffi::ErrorCode::kInternal, mcmc_last_error());
}

#define MCMC_COMMON_PARAMS                                                                                              \
  int32_t target_id, int32_t schedule_id, int32_t num_steps, int32_t max_tree_depth, int64_t num_samples,                \
      int64_t burn_in, int64_t chain_offset, double step_size, double gamma, double steepness, double delta_max,          \
      int64_t seed0, int64_t seed1, ffi::Span<const double> target_params
#define MCMC_COMMON_PACK                                                                                                \
  Common { target_id, schedule_id, num_steps, max_tree_depth, num_samples, burn_in, chain_offset, step_size, gamma,       \
           steepness, delta_max, seed0, seed1, target_params }

// plain run: samples, log_probs, accept_rate, final state
template <int SAMPLER>
ffi::Error RunPlain(cudaStream_t stream, ffi::ScratchAllocator scratch, F64 init_q, F64 inv_mass, F64 cov_inv,
                    MCMC_COMMON_PARAMS, RF64 samples, RF64 log_probs, RF32 accept_rate, RF64 final_q, RF64 final_lp,
                    RF64 final_g, RS32 accept_count) {
  Outputs o{samples->typed_data(), log_probs->typed_data(), accept_rate->typed_data(), final_q->typed_data(),
            final_lp->typed_data(), SAMPLER == MCMC_SAMPLER_RWMH ? nullptr : final_g->typed_data(),
            accept_count->typed_data()};
  return Run(SAMPLER, stream, scratch, init_q, inv_mass, cov_inv, MCMC_COMMON_PACK, o, 0u);
}

// track_proposals=True: + pre_positions, pre_lps, prop_positions, prop_lps, deltas_H
template <int SAMPLER>
ffi::Error RunTracked(cudaStream_t stream, ffi::ScratchAllocator scratch, F64 init_q, F64 inv_mass, F64 cov_inv,
                      MCMC_COMMON_PARAMS, RF64 samples, RF64 log_probs, RF32 accept_rate, RF64 final_q, RF64 final_lp,
                      RF64 final_g, RS32 accept_count, RF64 pre_q, RF64 pre_lp, RF64 prop_q, RF64 prop_lp, RF64 delta_h) {
  Outputs o{samples->typed_data(), log_probs->typed_data(), accept_rate->typed_data(), final_q->typed_data(),
            final_lp->typed_data(), final_g->typed_data(), accept_count->typed_data()};
  o.pre_q = pre_q->typed_data(); o.pre_lp = pre_lp->typed_data(); o.prop_q = prop_q->typed_data();
  o.prop_lp = prop_lp->typed_data(); o.delta_h = delta_h->typed_data();
  return Run(SAMPLER, stream, scratch, init_q, inv_mass, cov_inv, MCMC_COMMON_PACK, o, MCMC_FLAG_TRACK_PROPOSALS);
}

// NUTS: + tree_depths, mean_accept_probs
ffi::Error RunNuts(cudaStream_t stream, ffi::ScratchAllocator scratch, F64 init_q, F64 inv_mass, F64 cov_inv,
                   MCMC_COMMON_PARAMS, RF64 samples, RF64 log_probs, RF32 accept_rate, RF64 final_q, RF64 final_lp,
                   RF64 final_g, RS32 accept_count, RS32 tree_depths, RF64 mean_accept_probs) {
  Outputs o{samples->typed_data(), log_probs->typed_data(), accept_rate->typed_data(), final_q->typed_data(),
            final_lp->typed_data(), final_g->typed_data(), accept_count->typed_data()};
  o.tree_depths = tree_depths->typed_data();
  o.mean_accept_probs = mean_accept_probs->typed_data();
  return Run(MCMC_SAMPLER_NUTS, stream, scratch, init_q, inv_mass, cov_inv, MCMC_COMMON_PACK, o, 0u);
}

auto CommonBinding() {
  return ffi::Ffi::Bind()
      .Ctx<ffi::PlatformStream<cudaStream_t>>()
      .Ctx<ffi::ScratchAllocator>()
      .Arg<F64>()   // init_position (C, D)
      .Arg<F64>()   // inv_mass_matrix (D,) or empty
      .Arg<F64>()   // cov_inv (D, D) for dense_gaussian, else empty
      .Attr<int32_t>("target_id").Attr<int32_t>("schedule_id").Attr<int32_t>("num_steps").Attr<int32_t>("max_tree_depth")
      .Attr<int64_t>("num_samples").Attr<int64_t>("burn_in").Attr<int64_t>("chain_offset")
      .Attr<double>("step_size").Attr<double>("gamma").Attr<double>("steepness").Attr<double>("delta_max")
      .Attr<int64_t>("seed0").Attr<int64_t>("seed1").Attr<ffi::Span<const double>>("target_params")
      .Ret<F64>().Ret<F64>().Ret<ffi::Buffer<ffi::F32>>()                 // samples, log_probs, accept_rate
      .Ret<F64>().Ret<F64>().Ret<F64>().Ret<ffi::Buffer<ffi::S32>>();     // final q, lp, g; accept_count
}
auto TrackedBinding() { return CommonBinding().Ret<F64>().Ret<F64>().Ret<F64>().Ret<F64>().Ret<F64>(); }
auto NutsBinding() { return CommonBinding().Ret<ffi::Buffer<ffi::S32>>().Ret<F64>(); }

}  // namespace

XLA_FFI_DEFINE_HANDLER_SYMBOL(McmcHmcRun, RunPlain<MCMC_SAMPLER_HMC>, CommonBinding());
XLA_FFI_DEFINE_HANDLER_SYMBOL(McmcGrahmcRun, RunPlain<MCMC_SAMPLER_GRAHMC>, CommonBinding());
XLA_FFI_DEFINE_HANDLER_SYMBOL(McmcRwmhRun, RunPlain<MCMC_SAMPLER_RWMH>, CommonBinding());
XLA_FFI_DEFINE_HANDLER_SYMBOL(McmcHmcRunTracked, RunTracked<MCMC_SAMPLER_HMC>, TrackedBinding());
XLA_FFI_DEFINE_HANDLER_SYMBOL(McmcGrahmcRunTracked, RunTracked<MCMC_SAMPLER_GRAHMC>, TrackedBinding());
XLA_FFI_DEFINE_HANDLER_SYMBOL(McmcNutsRun, RunNuts, NutsBinding());

#endif  // MCMC_HAVE_XLA_FFI
