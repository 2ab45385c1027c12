"""Reference-signature samplers (samplers/{HMC,GRAHMC,RWMH,NUTS}.py of Chuxuan-a/mcmc)."""
from .HMC import HMCState, hmc_init, hmc_run, hmc_step, leapfrog  # noqa: F401
from .GRAHMC import (RAHMCState, FRICTION_SCHEDULES, get_friction_schedule, rahmc_init, rahmc_run, rahmc_step,  # noqa: F401
                     constant_schedule, tanh_schedule, sigmoid_schedule, linear_schedule, sine_schedule,
                     constant_schedule_default, standard_normal_log_prob)
from .RWMH import RWMState, rwMH_init, rwMH_run, rwMH_step  # noqa: F401
from .NUTS import NUTSState, nuts_init, nuts_run, nuts_step  # noqa: F401
