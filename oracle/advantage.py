"""ORACLE (test infrastructure, CPU) — advantage estimators restated in plain torch fp32.

Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s CPU legs may import this.  Validated in
this container by executing the unmodified reference functions on the same inputs
(tests/test_oracle_vs_reference.py) and pinned by tests/golden/advantage.npz.

Follows
  * verl/verl/trainer/ppo/core_algos.py:67-109    compute_gae_advantage_return
  * verl/verl/trainer/ppo/core_algos.py:113-165   compute_grpo_outcome_advantage
  * verl/verl/utils/torch_functional.py:117-151   masked_mean / masked_var / masked_whiten
  * ragen/trainer/core_algos.py:4-92              compute_bi_level_gae_advantage_return
  * ragen/trainer/agent_trainer.py:461-500        _filter_rollout group statistics + selection
  * verl/verl/trainer/ppo/core_algos.py:168-305   RF++-baseline, RLOO, RF++, ReMax
  * verl/verl/trainer/ppo/core_algos.py:469-501, verl/verl/trainer/ppo/ray_trainer.py:145-175   kl_penalty / apply_kl_penalty
"""
import torch


def masked_whiten(x, mask):
    m = mask.to(x.dtype)
    n = m.sum()
    mean = (x * m).sum() / (n + 1e-8)
    var = (((x - mean) ** 2) * m).sum() / (n + 1e-8)
    if n == 0:
        raise ValueError("At least one element in the mask has to be 1.")
    if n == 1:
        raise ValueError("The sum of the mask is one, which can cause a division by zero.")
    var = var * (n / (n - 1))
    return (x - mean) * torch.rsqrt(var + 1e-8)


def gae_raw(rewards, values, gamma, lam):
    """Un-whitened GAE: A_t = r_t + gamma V_{t+1} - V_t + gamma lam A_{t+1}; returns = A + V."""
    B, T = rewards.shape
    adv = torch.zeros_like(rewards, dtype=torch.float32)
    carry = torch.zeros(B, dtype=torch.float32)
    for t in range(T - 1, -1, -1):
        nxt = values[:, t + 1] if t < T - 1 else 0.0
        carry = rewards[:, t] + gamma * nxt - values[:, t] + gamma * lam * carry
        adv[:, t] = carry
    return adv, adv + values


def gae(rewards, values, mask, gamma, lam):
    adv, ret = gae_raw(rewards.float(), values.float(), gamma, lam)
    return masked_whiten(adv, mask), ret


def bilevel_gae_raw(rewards, values, loss_mask, gamma, lam, high_gamma):
    rewards = rewards.float()
    B, T = rewards.shape
    adv = torch.zeros_like(rewards)
    ret = torch.zeros_like(rewards)
    upd = rewards.clone()
    for b in range(B):
        turn_ends = torch.nonzero(rewards[b] != 0).flatten().tolist()
        carry = 0.0
        for j in range(len(turn_ends) - 1, -1, -1):
            pos = turn_ends[j]
            nxt = values[b, turn_ends[j + 1]] if j < len(turn_ends) - 1 else 0.0
            carry = upd[b, pos] + high_gamma * nxt - values[b, pos] + high_gamma * lam * carry
            adv[b, pos] = carry
        for pos in turn_ends:
            ret[b, pos] = adv[b, pos] + values[b, pos]
            upd[b, pos] = adv[b, pos] + values[b, pos]
        ends = set(turn_ends)
        toks = torch.nonzero(loss_mask[b] != 0).flatten().tolist()
        carry = 0.0
        for j in range(len(toks) - 1, -1, -1):
            pos = toks[j]
            if pos in ends:
                nxt, carry = 0.0, 0.0
            else:
                nxt = values[b, toks[j + 1]]  # reference raises IndexError when j is last (core_algos.py:79)
            carry = upd[b, pos] + gamma * nxt - values[b, pos] + gamma * lam * carry
            adv[b, pos] = carry
            ret[b, pos] = carry + values[b, pos]
    return adv, ret


def bilevel_gae(rewards, values, loss_mask, gamma, lam, high_gamma):
    adv, ret = bilevel_gae_raw(rewards, values.float(), loss_mask, gamma, lam, high_gamma)
    return masked_whiten(adv, loss_mask), ret


def grpo_outcome(rewards, mask, index, epsilon=1e-6, norm_by_std=True):
    scores = rewards.float().sum(dim=-1)
    groups = {}
    for i, key in enumerate(index):
        groups.setdefault(key, []).append(i)
    out = scores.clone()
    for rows in groups.values():
        if len(rows) == 1:
            mean, std = torch.tensor(0.0), torch.tensor(1.0)
        else:
            vals = scores[torch.tensor(rows)]
            mean, std = vals.mean(), vals.std()
        for i in rows:
            out[i] = (scores[i] - mean) / (std + epsilon) if norm_by_std else scores[i] - mean
    adv = out.unsqueeze(-1) * mask
    return adv, adv


def group_filter_stats(original_scores, num_groups, group_size):
    """agent_trainer.py:466-470: per-group std / max / mean of summed scores."""
    s = original_scores.sum(dim=-1).view(num_groups, group_size)
    return s.std(dim=-1), s.max(dim=-1).values, s.mean(dim=-1)


def _group_rows(index):
    groups = {}
    for i, key in enumerate(index):
        groups.setdefault(key, []).append(i)
    return groups


def rloo(rewards, mask, index):
    """compute_rloo_outcome_advantage (core_algos.py:205-242)."""
    scores = rewards.float().sum(dim=-1)
    out = scores.clone()
    for rows in _group_rows(index).values():
        n = len(rows)
        if n > 1:
            mean = torch.mean(scores[torch.tensor(rows)])
            for i in rows:
                out[i] = scores[i] * n / (n - 1) - mean * n / (n - 1)
    return out.unsqueeze(-1) * mask


def rfpp_baseline(rewards, mask, index):
    """compute_reinforce_plus_plus_baseline_outcome_advantage (core_algos.py:168-202)."""
    scores = rewards.float().sum(dim=-1)
    out = scores.clone()
    for rows in _group_rows(index).values():
        mean = torch.mean(scores[torch.tensor(rows)]) if len(rows) > 1 else torch.tensor(0.0)
        for i in rows:
            out[i] = scores[i] - mean
    tiled = out.unsqueeze(-1).tile([1, rewards.shape[-1]]) * mask
    return masked_whiten(tiled, mask)


def rfpp(rewards, mask, gamma):
    """compute_reinforce_plus_plus_outcome_advantage (core_algos.py:245-278) -> (advantages, returns)."""
    returns = torch.zeros_like(rewards)
    run = 0
    for t in reversed(range(rewards.shape[1])):
        run = rewards[:, t] + gamma * run
        returns[:, t] = run
        run = run * mask[:, t]
    return masked_whiten(returns, mask) * mask, returns


def remax(rewards, baselines, mask):
    """compute_remax_outcome_advantage (core_algos.py:281-305) -> (advantages, returns)."""
    returns = (rewards * mask).flip(dims=[-1]).cumsum(dim=-1).flip(dims=[-1])
    return returns - baselines.unsqueeze(-1) * mask, returns


def kl_penalty(logprob, ref_logprob, kind):
    """core_algos.py:469-501."""
    if kind == "kl":
        return logprob - ref_logprob
    if kind == "abs":
        return (logprob - ref_logprob).abs()
    if kind == "mse":
        return 0.5 * (logprob - ref_logprob).square()
    if kind == "low_var_kl":
        kl = ref_logprob - logprob
        return torch.clamp(torch.exp(kl) - kl - 1, min=-10, max=10)
    raise NotImplementedError


def apply_kl_penalty(scores, logprob, ref_logprob, mask, kind, beta):
    """ray_trainer.py:145-175 -> (token_level_rewards, current_kl)."""
    kld = kl_penalty(logprob, ref_logprob, kind) * mask
    per_row = (kld * mask).sum(-1) / (mask.sum(-1) + 1e-8)
    return scores - beta * kld, torch.mean(per_row, dim=0).item()


def filter_rollout(original_rm_scores, num_groups, group_size, ratio, kind="std"):
    """_filter_rollout (agent_trainer.py:461-500; a closure inside fit(), restated): -> (row keep-mask, metrics)."""
    std, mx, mean = group_filter_stats(original_rm_scores, num_groups, group_size)
    if ratio == 1:
        keep = torch.ones(num_groups * group_size, dtype=torch.bool)
        chosen = slice(None)
    else:
        chosen = ((-std) if kind == "std_rev" else std).topk(int(ratio * num_groups)).indices
        gm = torch.zeros(num_groups, dtype=torch.bool)
        gm[chosen] = True
        keep = gm.unsqueeze(1).expand(-1, group_size).flatten()
    metrics = {"rollout/in_group_std": std.mean(), "rollout/in_group_max": mx.mean(), "rollout/in_group_mean": mean.mean(),
               "rollout/chosen_in_group_std": std[chosen].mean(), "rollout/chosen_in_group_max": mx[chosen].mean(),
               "rollout/chosen_in_group_mean": mean[chosen].mean()}
    return keep, metrics
