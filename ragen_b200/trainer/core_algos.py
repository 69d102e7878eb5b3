"""Advantage estimators with the reference's names and argument meaning, on the device.

  compute_gae_advantage_return            verl/verl/trainer/ppo/core_algos.py:67-109
  compute_bi_level_gae_advantage_return   ragen/trainer/core_algos.py:4-92
  compute_grpo_outcome_advantage          verl/verl/trainer/ppo/core_algos.py:113-165
  masked_mean / masked_var / masked_whiten  verl/verl/utils/torch_functional.py:122-151

Each is one or two launches of libragen_b200 (ragen_gae / ragen_bilevel_gae produce the un-whitened advantages, the
returns and the masked moments in one pass; ragen_whiten applies them).  Inputs may live on the host, as in the
reference's driver process: they are moved to `device` and the results come back where the inputs were.  When the rows of
the batch are sharded over ranks (`group` given, or a default process group with world_size > 1 and `sharded=True`), the
whitening moments are all-reduced (24 bytes over NCCL) so every rank whitens with the batch-global statistics.
"""
from collections import OrderedDict

import numpy as np
import torch

from .. import ops


def _dev(t, device):
    if t.is_cuda:
        return t.device
    return torch.device("cuda", torch.cuda.current_device() if device is None else int(device))


def _whiten_from(adv, mom, sharded, group):
    if sharded:
        ops.all_reduce_moments(mom, group)
    return ops.whiten_(adv, mom)


SCAN_MIN_ROWS = 4096  # from here on the pass is bandwidth-bound and the parallel scan is the default


def compute_gae_advantage_return(token_level_rewards, values, response_mask, gamma, lam, device=None, sharded=False, group=None,
                                 exact=None):
    """exact=True: un-whitened advantages and returns carry the reference's bits (its sequential fp32 loop, verl core_algos.py:
    100-104); exact=False: the recurrence as a parallel scan (ragen_gae_scan), equal up to rounding - a few 1e-7, north_star's
    bar is 1e-5 - at HBM speed; None: exact below SCAN_MIN_ROWS rows (there the pass is latency-bound either way)."""
    home, dev = token_level_rewards.device, _dev(token_level_rewards, device)
    if exact is None:
        exact = token_level_rewards.shape[0] < SCAN_MIN_ROWS
    with torch.no_grad():
        adv, ret, mom = ops.gae_raw(token_level_rewards.to(dev), values.to(dev), response_mask.to(dev), gamma, lam, exact=exact)
        adv = _whiten_from(adv, mom, sharded, group)
    return adv.to(home), ret.to(home)


def compute_bi_level_gae_advantage_return(token_level_rewards, values, loss_mask, gamma, lam, high_level_gamma, device=None,
                                          sharded=False, group=None, exact=None):
    """exact as in compute_gae_advantage_return: None takes the block-scan kernel (ragen_bilevel_gae_scan) from SCAN_MIN_ROWS
    rows on and the reference's fp32 operation order (ragen_bilevel_gae) below."""
    home, dev = token_level_rewards.device, _dev(token_level_rewards, device)
    if exact is None:
        exact = token_level_rewards.shape[0] < SCAN_MIN_ROWS
    with torch.no_grad():
        adv, ret, mom = ops.bilevel_gae_raw(token_level_rewards.to(dev), values.to(dev), loss_mask.to(dev), gamma, lam, high_level_gamma,
                                            exact=exact)
        adv = _whiten_from(adv, mom, sharded, group)
    return adv.to(home), ret.to(home)


def dense_index(index):
    """uid array (any hashables) -> (int32 list of dense group ids in first-seen order, number of groups)."""
    seen = OrderedDict()
    idx = [seen.setdefault(u, len(seen)) for u in (index.tolist() if isinstance(index, np.ndarray) else list(index))]
    return idx, len(seen)


def compute_grpo_outcome_advantage(token_level_rewards, response_mask, index, epsilon: float = 1e-6,
                                   norm_adv_by_std_in_grpo: bool = True, device=None):
    """Returns the same tensor twice, as the reference does (:165)."""
    home, dev = token_level_rewards.device, _dev(token_level_rewards, device)
    gi, n_groups = dense_index(index)
    with torch.no_grad():
        adv, _ = ops.grpo_outcome(token_level_rewards.to(dev), response_mask.to(dev),
                                  torch.tensor(gi, dtype=torch.int32, device=dev), n_groups, epsilon, norm_adv_by_std_in_grpo)
    adv = adv.to(home)
    return adv, adv


def compute_rloo_outcome_advantage(token_level_rewards, response_mask, index, epsilon: float = 1e-6, device=None):
    """verl/verl/trainer/ppo/core_algos.py:205-242."""
    home, dev = token_level_rewards.device, _dev(token_level_rewards, device)
    gi, n_groups = dense_index(index)
    with torch.no_grad():
        adv, _, _ = ops.outcome_advantage(token_level_rewards.to(dev), response_mask.to(dev),
                                          torch.tensor(gi, dtype=torch.int32, device=dev), n_groups, "rloo", epsilon)
    adv = adv.to(home)
    return adv, adv


def compute_reinforce_plus_plus_baseline_outcome_advantage(token_level_rewards, response_mask, index, epsilon: float = 1e-6,
                                                           device=None, sharded=False, group=None):
    """verl/verl/trainer/ppo/core_algos.py:168-202 (the whitening is batch-global: moments all-reduced when sharded)."""
    home, dev = token_level_rewards.device, _dev(token_level_rewards, device)
    gi, n_groups = dense_index(index)
    with torch.no_grad():
        adv, _, mom = ops.outcome_advantage(token_level_rewards.to(dev), response_mask.to(dev),
                                            torch.tensor(gi, dtype=torch.int32, device=dev), n_groups, "rfpp_baseline", epsilon,
                                            want_moments=True)
        adv = _whiten_from(adv, mom, sharded, group)
    adv = adv.to(home)
    return adv, adv


def compute_reinforce_plus_plus_outcome_advantage(token_level_rewards, response_mask, gamma, device=None, sharded=False, group=None,
                                                  exact=None):
    """verl/verl/trainer/ppo/core_algos.py:245-278.  exact as in compute_gae_advantage_return."""
    home, dev = token_level_rewards.device, _dev(token_level_rewards, device)
    if exact is None:
        exact = token_level_rewards.shape[0] < SCAN_MIN_ROWS
    with torch.no_grad():
        ret, m, mom = ops.discounted_returns(token_level_rewards.to(dev), response_mask.to(dev), float(gamma), exact=exact)
        if sharded:
            ops.all_reduce_moments(mom, group)
        adv = ops.whiten_masked(ret, m, mom)
    return adv.to(home), ret.to(home)


def compute_remax_outcome_advantage(token_level_rewards, reward_baselines, response_mask, device=None, exact=None):
    """verl/verl/trainer/ppo/core_algos.py:281-305.  exact as in compute_gae_advantage_return."""
    home, dev = token_level_rewards.device, _dev(token_level_rewards, device)
    if exact is None:
        exact = token_level_rewards.shape[0] < SCAN_MIN_ROWS
    with torch.no_grad():
        adv, ret = ops.remax(token_level_rewards.to(dev), response_mask.to(dev), reward_baselines.to(dev), exact=exact)
    return adv.to(home), ret.to(home)


class AdaptiveKLController:
    """verl/verl/trainer/ppo/core_algos.py:29-44."""

    def __init__(self, init_kl_coef, target_kl, horizon):
        self.value = init_kl_coef
        self.target = target_kl
        self.horizon = horizon

    def update(self, current_kl, n_steps):
        proportional_error = np.clip(current_kl / self.target - 1, -0.2, 0.2)
        self.value *= 1 + proportional_error * n_steps / self.horizon


class FixedKLController:
    """verl/verl/trainer/ppo/core_algos.py:47-54."""

    def __init__(self, kl_coef):
        self.value = kl_coef

    def update(self, current_kl, n_steps):
        pass


def get_kl_controller(kl_ctrl):
    """verl/verl/trainer/ppo/core_algos.py:57-64."""
    if kl_ctrl.type == "fixed":
        return FixedKLController(kl_coef=kl_ctrl.kl_coef)
    if kl_ctrl.type == "adaptive":
        assert kl_ctrl.horizon > 0, f"horizon must be larger than 0. Got {kl_ctrl.horizon}"
        return AdaptiveKLController(init_kl_coef=kl_ctrl.kl_coef, target_kl=kl_ctrl.target_kl, horizon=kl_ctrl.horizon)
    raise NotImplementedError


def masked_moments(values, mask, device=None, sharded=False, group=None):
    dev = _dev(values, device)
    mom = ops.masked_moments(values.to(dev), mask.to(dev))
    if sharded:
        ops.all_reduce_moments(mom, group)
    return mom


def masked_mean(values, mask, device=None):
    n, s, _ = masked_moments(values, mask, device).tolist()
    return torch.tensor(s / (n + 1e-8), dtype=torch.float32, device=values.device)


def masked_whiten(values, mask, shift_mean=True, device=None, sharded=False, group=None):
    home, dev = values.device, _dev(values, device)
    x = values.to(dev).float().contiguous().clone()
    mom = masked_moments(x, mask, device, sharded, group)
    out = ops.whiten_(x, mom)
    if not shift_mean:
        n, s, _ = mom.tolist()
        out += s / (n + 1e-8)
    return out.to(home)
