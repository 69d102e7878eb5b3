"""`compute_advantage` with the reference's signature (ragen/trainer/agent_trainer.py:60-137) over the device estimators.

GAE (flat and bi-level) and GRPO are the estimators north_star names (a29-a33); RLOO, REINFORCE++ (with and
without baseline) and ReMax complete the dispatch table, and `apply_kl_penalty` / `filter_rollout` are the two passes that sit
between the rollout and it in RayAgentTrainer.fit.
"""
from enum import Enum

import numpy as np
import torch

from .. import ops
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
    elif adv_estimator == AdvantageEstimator.REINFORCE_PLUS_PLUS_BASELINE:
        advantages, returns = core_algos.compute_reinforce_plus_plus_baseline_outcome_advantage(
            token_level_rewards=data.batch["token_level_rewards"], response_mask=data.batch["response_mask"],
            index=data.non_tensor_batch["uid"], device=device, sharded=sharded, group=group)
    elif adv_estimator == AdvantageEstimator.REINFORCE_PLUS_PLUS:
        advantages, returns = core_algos.compute_reinforce_plus_plus_outcome_advantage(
            token_level_rewards=data.batch["token_level_rewards"], response_mask=data.batch["response_mask"], gamma=gamma,
            device=device, sharded=sharded, group=group)
    elif adv_estimator == AdvantageEstimator.REMAX:
        advantages, returns = core_algos.compute_remax_outcome_advantage(
            token_level_rewards=data.batch["token_level_rewards"], reward_baselines=data.batch["reward_baselines"],
            response_mask=data.batch["response_mask"], device=device)
    elif adv_estimator == AdvantageEstimator.RLOO:
        advantages, returns = core_algos.compute_rloo_outcome_advantage(
            token_level_rewards=data.batch["token_level_rewards"], response_mask=data.batch["response_mask"],
            index=data.non_tensor_batch["uid"], device=device)
    else:
        raise NotImplementedError
    data.batch["advantages"] = advantages
    data.batch["returns"] = returns
    return data


def apply_kl_penalty(data, kl_ctrl, kl_penalty="kl", multi_turn=False, device=None):
    """verl/verl/trainer/ppo/ray_trainer.py:145-175: token_level_rewards = token_level_scores - beta * kl * mask in one launch;
    the controller is updated with the mean over rows of the per-row masked mean, as in the reference."""
    response_length = data.batch["responses"].size(1)
    token_level_scores = data.batch["token_level_scores"]
    batch_size = token_level_scores.shape[0]
    if multi_turn:
        response_mask = data.batch["loss_mask"][:, -response_length:]
    else:
        response_mask = data.batch["attention_mask"][:, -response_length:]
    home = token_level_scores.device
    dev = home if home.type == "cuda" else torch.device("cuda", torch.cuda.current_device() if device is None else int(device))
    beta = kl_ctrl.value
    rewards, row_kl = ops.kl_rewards(token_level_scores.to(dev), data.batch["old_log_probs"].to(dev), data.batch["ref_log_prob"].to(dev),
                                     response_mask.to(dev), kl_penalty, beta)
    current_kl = torch.mean(row_kl, dim=0).item()
    kl_ctrl.update(current_kl=current_kl, n_steps=batch_size)
    data.batch["token_level_rewards"] = rewards.to(home)
    return data, {"actor/reward_kl_penalty": current_kl, "actor/reward_kl_penalty_coeff": beta}


def filter_rollout(batch, num_groups, group_size, rollout_filter_ratio, rollout_filter_type="std", device=None, sharded=False,
                   group=None):
    """StarPO-S trajectory filter, the `_filter_rollout` closure of RayAgentTrainer.fit (ragen/trainer/agent_trainer.py:461-500):
    keep the top `ratio` of the env groups by in-group std of the summed scores ("std") or by its negative ("std_rev").
    The statistics are one launch; choosing int(ratio * num_groups) of num_groups floats and the boolean row gather of the
    (host-resident) batch are the reference's own torch calls, so ties break exactly as they do there.

    sharded=True: `batch` holds this rank's env groups only (consecutive rows per group, their GLOBAL ids in
    non_tensor_batch["group_ids"]) and `num_groups` is the number of groups of the WHOLE rollout.  Every rank writes the
    std / max / mean of its groups at their global index of one float64 vector, the vector is all-reduced (3 * num_groups
    doubles), every rank takes the reference's top-k on the full vector and keeps its own rows of the chosen
    groups; the metrics are those of the whole rollout on every rank."""
    scores = batch.batch["original_rm_scores"]
    home = scores.device
    dev = home if home.type == "cuda" else torch.device("cuda", torch.cuda.current_device() if device is None else int(device))
    local_groups = scores.shape[0] // group_size if sharded else num_groups
    if local_groups:
        _, in_group_std, in_group_max, in_group_mean = (t.to(home) for t in ops.group_filter_stats(scores.to(dev), local_groups, group_size))
    else:
        in_group_std = in_group_max = in_group_mean = torch.zeros(0, dtype=torch.float32, device=home)
    my_groups = None
    if sharded:
        from ..sharding import all_reduce_group_stats

        my_groups = torch.tensor([int(g) for g in batch.non_tensor_batch["group_ids"][::group_size]], dtype=torch.long)
        in_group_std, in_group_max, in_group_mean = all_reduce_group_stats(
            num_groups, my_groups, (in_group_std, in_group_max, in_group_mean), group=group, device=dev)
        in_group_std, in_group_max, in_group_mean = (t.to(home) for t in (in_group_std, in_group_max, in_group_mean))
    if rollout_filter_ratio == 1:
        return batch, {"rollout/in_group_std": in_group_std.mean(), "rollout/in_group_max": in_group_max.mean(),
                       "rollout/in_group_mean": in_group_mean.mean(), "rollout/chosen_in_group_std": in_group_std.mean(),
                       "rollout/chosen_in_group_max": in_group_max.mean(), "rollout/chosen_in_group_mean": in_group_mean.mean()}
    k = int(rollout_filter_ratio * num_groups)
    if rollout_filter_type == "std_rev":
        top_groups = (-in_group_std).topk(k).indices
    elif rollout_filter_type == "std":
        top_groups = in_group_std.topk(k).indices
    else:
        raise ValueError(f"Invalid rollout filter type: {rollout_filter_type}")
    mask = torch.zeros(num_groups, dtype=torch.bool)
    mask[top_groups.cpu()] = True
    if my_groups is not None:
        mask = mask[my_groups]  # this rank's groups, in the order of its rows
    mask = mask.unsqueeze(1).expand(-1, group_size).flatten()
    if isinstance(batch.batch, dict):  # the stand-in container (no tensordict installed)
        kept = type(batch.batch)({k_: v[mask.to(v.device)] for k_, v in batch.batch.items()})
        kept.batch_size = int(mask.sum())
        batch.batch = kept
    else:
        batch.batch = batch.batch[mask]
    for key, value in batch.non_tensor_batch.items():
        if isinstance(value, np.ndarray):
            batch.non_tensor_batch[key] = value[mask.numpy()]
        else:
            batch.non_tensor_batch[key] = [v for v, m in zip(value, mask) if m]
    metrics = {"rollout/in_group_std": in_group_std.mean(), "rollout/in_group_max": in_group_max.mean(),
               "rollout/in_group_mean": in_group_mean.mean(), "rollout/chosen_in_group_std": in_group_std[top_groups].mean(),
               "rollout/chosen_in_group_max": in_group_max[top_groups].mean(),
               "rollout/chosen_in_group_mean": in_group_mean[top_groups].mean()}
    return batch, metrics
