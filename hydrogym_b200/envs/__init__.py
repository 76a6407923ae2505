from .kolmogorov import FlowConfig, KolmogorovFlow, KolmogorovFlowParams, KolmogorovFlowState  # noqa: F401
from .channel import ChannelEnvParams, ChannelEnvState, ChannelFlowSpectralEnv, ChannelStepper, run_channel_pseudospectral  # noqa: F401
