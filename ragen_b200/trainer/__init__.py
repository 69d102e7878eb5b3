from .agent_trainer import AdvantageEstimator, compute_advantage, compute_response_mask  # noqa: F401
