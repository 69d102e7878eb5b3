from .agent_trainer import AdvantageEstimator, apply_kl_penalty, compute_advantage, compute_response_mask, filter_rollout  # noqa: F401
