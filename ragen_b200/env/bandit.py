"""Bandit on the GPU: device-resident batch + per-env view with the reference API
(ragen/env/bandit/env.py:14-85, config.py:4-14)."""
from dataclasses import dataclass
from typing import Dict, Optional

import numpy as np
import torch

from .. import _lib
from .base import BaseDiscreteActionEnv

INIT_PROMPT = (  # ragen/env/bandit/env.py:6-12, trailing blanks included
    "You are playing a bandit game. Goal: Maximize your total reward by choosing which arm to pull. \n"
    "Game Rules: \n"
    "1. There are 2 arms, named {name_a} and {name_b}\n"
    "2. Each arm has its own reward distribution, related to their names. \n"
    "3. Analyze the symbolic meaning of each arm's name to guess how their reward distribution might behave.\n"
    "4. Based on the symbolic meaning of their names, which arm do you think is more likely to give higher "
    "rewards on average? Choose between {name_a} and {name_b}, and output like <answer> {name_a} </answer> "
    "or <answer> {name_b} </answer>.\n"
)


@dataclass
class BanditEnvConfig:
    """Field-for-field the reference dataclass (ragen/env/bandit/config.py:4-14)."""
    lo_arm_name: str = "phoenix"
    hi_arm_name: str = "dragon"
    action_space_start: int = 1
    lo_arm_score: float = 0.1
    hi_arm_loscore: float = 0.0
    hi_arm_hiscore: float = 1.0
    hi_arm_hiscore_prob: float = 0.25
    render_mode: str = "text"
    action_lookup: Dict[int, str] = None  # set at reset, as in the reference


class BanditBatch:
    """n two-armed bandits in HBM: lo_first uint8 [n] (1: action id `start` is the lo arm) and one NumPy-exact
    PCG64 generator per env (uint64 [2, n, 2])."""

    def __init__(self, n, config: Optional[BanditEnvConfig] = None, device=0):
        self.ctx = _lib.Context.get(device)
        self.lib = self.ctx.lib
        self.device = dev = torch.device("cuda", int(device))
        self.n = int(n)
        self.config = config or BanditEnvConfig()
        self.rng = torch.zeros((2, self.n, 2), dtype=torch.int64, device=dev)
        self.lo_first = torch.zeros(self.n, dtype=torch.uint8, device=dev)
        self.reward = torch.zeros(self.n, dtype=torch.float64, device=dev)
        self.flags = torch.zeros(self.n, dtype=torch.uint8, device=dev)

    def reset(self, seeds):
        s = torch.as_tensor(np.asarray(seeds, dtype=np.uint64).astype(np.int64)).to(self.device)
        assert s.numel() == self.n
        _lib.check(self.lib.ragen_bandit_reset(self.ctx.handle, self.n, _lib.ptr(s), _lib.ptr(self.rng), _lib.ptr(self.lo_first),
                                               self.ctx.stream()))
        return self.lo_first

    def reset_slot(self, slot, seed):
        """BaseEnv.reset(seed) of ONE env of a shared batch (ragen/env/base.py:16): arm order and generator of `seed`."""
        one = getattr(self, "_one", None)
        if one is None:
            one = self._one = BanditBatch(1, self.config, device=self.device.index)
        one.reset([seed])
        self.lo_first[slot] = one.lo_first[0]
        self.rng[:, slot].copy_(one.rng[:, 0])

    def step(self, actions, active=None):
        """actions uint8 CUDA [n] (action ids); active uint8 [n] or None.  Returns (reward f64 [n], flags u8 [n])."""
        c = self.config
        assert actions.dtype == torch.uint8 and actions.is_cuda and actions.numel() == self.n
        _lib.check(self.lib.ragen_bandit_step(self.ctx.handle, self.n, int(c.action_space_start), float(c.lo_arm_score),
                                              float(c.hi_arm_loscore), float(c.hi_arm_hiscore), float(c.hi_arm_hiscore_prob),
                                              _lib.ptr(actions.contiguous()), _lib.ptr(active), _lib.ptr(self.rng),
                                              _lib.ptr(self.lo_first), _lib.ptr(self.reward), _lib.ptr(self.flags),
                                              self.ctx.stream()))
        return self.reward, self.flags

    def lookup(self, slot, lf=None):
        c, s = self.config, self.config.action_space_start
        lf = bool(self.lo_first[slot]) if lf is None else lf
        return {s: c.lo_arm_name if lf else c.hi_arm_name, s + 1: c.hi_arm_name if lf else c.lo_arm_name}


class BanditEnv(BaseDiscreteActionEnv):
    """One env with the reference interface, backed by a slot of a BanditBatch."""

    def __init__(self, config=None, batch: Optional[BanditBatch] = None, slot: int = 0, **kwargs):
        BaseDiscreteActionEnv.__init__(self)
        self.config = config if config is not None else BanditEnvConfig()
        self.lo_arm_name, self.hi_arm_name = self.config.lo_arm_name, self.config.hi_arm_name
        self.render_cache = None
        self.render_mode = self.config.render_mode
        assert self.render_mode == "text"
        if batch is None:
            batch = BanditBatch(1, self.config, device=kwargs.get("device", 0))
            slot = 0
        self.batch, self.slot = batch, slot
        self._act = self._on = None  # staging tensors of the single-env step path, made on first use

    def adopt_reset(self, lo_first=None):
        """Host bookkeeping after the batch was reset (arm order -> lookup dicts and the prompt)."""
        self.ACTION_LOOKUP = self.batch.lookup(self.slot, lo_first)
        self.config.action_lookup = self.ACTION_LOOKUP
        self.ARM_IDX_TO_NAME = self.ACTION_LOOKUP
        self.NAME_TO_ARM_IDX = {name: idx for idx, name in self.ACTION_LOOKUP.items()}
        s = self.config.action_space_start
        self.render_cache = INIT_PROMPT.format(name_a=self.ARM_IDX_TO_NAME[s], name_b=self.ARM_IDX_TO_NAME[s + 1])
        return self.render_cache

    def reset(self, seed=None, mode=None):
        if seed is None:
            seed = int(np.random.SeedSequence().entropy % (2 ** 63))
        if self.batch.n != 1:
            self.batch.reset_slot(self.slot, seed)  # the neighbours in the batch are not touched
        else:
            self.batch.reset([seed])
        return self.adopt_reset()

    def _staging(self):
        if self._act is None:
            batch = self.batch
            self._act = torch.zeros(batch.n, dtype=torch.uint8, device=batch.device)
            self._on = torch.zeros(batch.n, dtype=torch.uint8, device=batch.device)

    def step(self, action: int):
        assert action in self.ACTION_LOOKUP, f"Invalid action: {action}"
        self._staging()
        self._act.zero_()
        self._on.zero_()
        self._act[self.slot] = int(action)
        self._on[self.slot] = 1
        reward, flags = self.batch.step(self._act, self._on)
        return self.finish_step(action, float(reward[self.slot]), int(flags[self.slot]))

    def finish_step(self, action, reward, flags):
        """Observation text and info dict from the device result (env.py:62-70)."""
        arm_name = self.ARM_IDX_TO_NAME[action]
        # the device returns the pay-out as a double; hand back the config's own object of that value so that an
        # int-typed score prints the way the reference prints it
        c = self.config
        cands = (c.lo_arm_score,) if arm_name == self.lo_arm_name else (c.hi_arm_hiscore, c.hi_arm_loscore)
        reward = next((v for v in cands if float(v) == reward), reward)
        next_obs = f"{arm_name}: {reward} points"
        self.render_cache = next_obs
        info = {"action_is_effective": True, "action_is_valid": True, "success": arm_name == self.hi_arm_name}
        return next_obs, reward, True, info

    def get_all_actions(self):
        s = self.config.action_space_start
        return [s, s + 1]

    def render(self, mode=None):
        return self.render_cache

    def close(self):
        self.render_cache = None
