"""hydrogym_b200 -- B200-native drop-in for the hydrogym/jax Kolmogorov-flow environment step.

Only the hot path named in BASELINE.json is implemented, plus the "next" rows of SURVEY.md §8(f)
(wrappers, solver surface, diagnostics, the 3-D channel-flow env); the compute runs in hand-written sm_100a CUDA
kernels behind the C ABI of ``include/kolmo_b200.h`` and ``include/channel_b200.h``.
"""
from .envs.kolmogorov import (FlowConfig, KolmogorovFlow, KolmogorovFlowParams,  # noqa: F401
                              KolmogorovFlowState)
from .envs.channel import (ChannelEnvParams, ChannelEnvState, ChannelFlowSpectralEnv,  # noqa: F401
                           ChannelStepper, run_channel_pseudospectral)
from .solver import KolmogorovStepper  # noqa: F401
from .vector_env import KolmogorovVectorEnv  # noqa: F401
from .slab import SlabKolmogorovSolver  # noqa: F401
from . import diagnostics, io, wrappers  # noqa: F401

__all__ = ["FlowConfig", "KolmogorovFlow", "KolmogorovFlowParams", "KolmogorovFlowState",
           "ChannelEnvParams", "ChannelEnvState", "ChannelFlowSpectralEnv", "ChannelStepper",
           "run_channel_pseudospectral", "KolmogorovStepper", "KolmogorovVectorEnv", "SlabKolmogorovSolver", "diagnostics", "io", "wrappers"]
