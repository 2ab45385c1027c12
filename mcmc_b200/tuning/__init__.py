"""Host-side tuning API of the reference (tuning/{welford,dual_averaging,adaptation}.py) over the CUDA samplers."""
from .welford import WelfordState, welford_init, welford_update, welford_update_batch, welford_covariance  # noqa: F401
from .dual_averaging import (DualAveragingState, da_init, da_update, da_reset,  # noqa: F401
                             JointDualAveragingState, joint_da_init, joint_da_update,
                             dual_averaging_tune_rwmh, dual_averaging_tune_hmc, dual_averaging_tune_nuts,
                             joint_tune_grahmc)
from .adaptation import build_schedule, run_adaptive_warmup  # noqa: F401
from .sequential_tune_grahmc import sequential_tune_grahmc  # noqa: F401
