"""smc-physics on B200: the particle-filter hot path of jackhdewey/smc-physics as sm_100a CUDA kernels behind a
C ABI (include/smcphys.h), with a host-side mirror of the reference's Args / run_smc / run_inference / PhySMC
interfaces.  The CUDA library is the only implementation: nothing here computes on the CPU."""
from ._lib import CUBE, SPHERE, SmcphysError, lib  # noqa: F401
from .api import (MULTINOMIAL, SYSTEMATIC, K_EXTEND, K_GATHER, K_REJUVENATE, K_WEIGHTS, ParticleFilter, fp32_peak, default_config, offspring_range, resample_multinomial, resample_systematic, sim_step, simulate,  # noqa: F401
                  substeps_for, weight_stats)
