"""`compute_advantage` with the reference's signature (ragen/trainer/agent_trainer.py:60-137) over the device estimators.

GAE (flat and bi-level) and GRPO are the estimators on this path (a29-a33).  The other members of the
dispatch table (RLOO, REINFORCE++, ReMax) are 8(f) "next" rows: asking for one raises NotImplementedError here rather than
running anywhere else.
"""
from enum import Enum

from . import core_algos


class AdvantageEstimator(str, Enum):
    """verl/verl/trainer/ppo/ray_trainer.py:79-89."""
    GAE = "gae"
    GRPO = "grpo"
    REINFORCE_PLUS_PLUS = "reinforce_plus_plus"
    REINFORCE_PLUS_PLUS_BASELINE = "reinforce_plus_plus_baseline"
    REMAX = "remax"
    RLOO = "rloo"


def compute_response_mask(data):
    """verl/verl/trainer/ppo/ray_trainer.py:178-182: attention mask of the response part."""
    response_length = data.batch["responses"].size(1)
    return data.batch["attention_mask"][:, -response_length:]


def compute_advantage(data, adv_estimator, gamma=1.0, lam=1.0, num_repeat=1, multi_turn=False, norm_adv_by_std_in_grpo=True,
                      bi_level_gae=False, high_level_gamma=1.0, device=None, sharded=False, group=None):
    """`sharded=True`: the rows of `data` are this rank's env groups only; whitening uses the all-reduced moments."""
    if "response_mask" not in data.batch:
        data.batch["response_mask"] = compute_response_mask(data)
    if adv_estimator == AdvantageEstimator.GAE:
        if bi_level_gae:
            advantages, returns = core_algos.compute_bi_level_gae_advantage_return(
                token_level_rewards=data.batch["token_level_rewards"], values=data.batch["values"],
                loss_mask=data.batch["response_mask"], gamma=gamma, lam=lam, high_level_gamma=high_level_gamma,
                device=device, sharded=sharded, group=group)
        else:
            advantages, returns = core_algos.compute_gae_advantage_return(
                token_level_rewards=data.batch["token_level_rewards"], values=data.batch["values"],
                response_mask=data.batch["response_mask"], gamma=gamma, lam=lam, device=device, sharded=sharded, group=group)
    elif adv_estimator == AdvantageEstimator.GRPO:
        grpo_calculation_mask = data.batch["response_mask"]
        if multi_turn:
            response_length = grpo_calculation_mask.size(1)
            grpo_calculation_mask = data.batch["loss_mask"][:, -response_length:]
        advantages, returns = core_algos.compute_grpo_outcome_advantage(
            token_level_rewards=data.batch["token_level_rewards"], response_mask=grpo_calculation_mask,
            index=data.non_tensor_batch["uid"], norm_adv_by_std_in_grpo=norm_adv_by_std_in_grpo, device=device)
    else:
        raise NotImplementedError(f"{adv_estimator}: not on the accelerated path yet (SURVEY.md 8f rank 2)")
    data.batch["advantages"] = advantages
    data.batch["returns"] = returns
    return data
