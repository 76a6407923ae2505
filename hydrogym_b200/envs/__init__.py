from .kolmogorov import FlowConfig, KolmogorovFlow, KolmogorovFlowParams, KolmogorovFlowState  # noqa: F401
