"""Countdown on the GPU: the reward pass of ragen/env/countdown/env.py:9-21, 69-78 through ragen_countdown_score, behind
the reference's env API (env.py:36-78, config.py:4-10).  Problems come from the parquet file the config names
(needs `datasets`, as in the reference) or from an in-memory list of {'target': int, 'nums': [int, ...]} rows."""
import itertools
from dataclasses import dataclass, field
from typing import List, Optional

import numpy as np
import torch

from .. import _lib
from .base import BaseLanguageBasedEnv


@dataclass
class CountdownEnvConfig:
    """ragen/env/countdown/config.py:4-10 (score / format_score are class attributes there too) + `data`."""
    train_path: str = "data/countdown/train.parquet"
    max_instances: int = 20000
    render_mode: str = "text"
    score = 1
    format_score = 0.1
    data: Optional[List[dict]] = field(default=None, repr=False)  # in-memory problems instead of train_path


def has_solution(nums, target):
    """env.py:23-33: some +/- combination of the (zero-padded to 4) numbers hits the target."""
    nums = list(nums) + [0] * (4 - len(nums))
    return any(sum(s * v for s, v in zip(signs, nums)) == target for signs in itertools.product([1, -1], repeat=4))


def load_problems(config):
    if config.data is not None:
        rows = list(config.data)[: config.max_instances]
    else:
        try:
            import datasets
        except ImportError as e:  # the reference has the same hard dependency (env.py:3)
            raise RuntimeError("CountdownEnvConfig.train_path needs the `datasets` package; pass config.data instead") from e
        rows = list(datasets.load_dataset("parquet", data_files=config.train_path)["train"].select(range(config.max_instances)))
    return [r for r in rows if has_solution(r["nums"], r["target"])]


class CountdownBatch:
    """Scores a batch of answers on the device.  There is no per-env device state (1-step episodes): the batch only
    owns the launch."""

    def __init__(self, n, config: Optional[CountdownEnvConfig] = None, device=0):
        self.ctx = _lib.Context.get(device)
        self.device = torch.device("cuda", int(device))
        self.n = int(n)
        self.config = config or CountdownEnvConfig()

    def score(self, answers, nums_list, targets):
        """answers: list[str]; returns (result u8 numpy {0, 1, 2}, unsupported u8 numpy)."""
        from .. import ops

        enc = [a.encode() for a in answers]
        k = len(enc)
        lens = np.array([len(e) for e in enc], dtype=np.int32)
        start = np.zeros(k, dtype=np.int64)
        np.cumsum(lens[:-1], out=start[1:])
        nums = np.zeros((k, 8), dtype=np.int32)
        n_nums = np.zeros(k, dtype=np.int32)
        for i, v in enumerate(nums_list):
            if len(v) > 8:
                raise ValueError("at most 8 numbers per Countdown problem")
            nums[i, :len(v)] = v
            n_nums[i] = len(v)
        dev = self.device
        text = torch.frombuffer(bytearray(b"".join(enc) or b"\0"), dtype=torch.uint8).to(dev)
        res, uns = ops.countdown_score(text, torch.from_numpy(start).to(dev), torch.from_numpy(lens).to(dev),
                                       torch.from_numpy(nums).to(dev), torch.from_numpy(n_nums).to(dev),
                                       torch.tensor([int(t) for t in targets], dtype=torch.int64, device=dev))
        return res.cpu().numpy(), uns.cpu().numpy()


class CountdownEnv(BaseLanguageBasedEnv):
    def __init__(self, config=None, batch: Optional[CountdownBatch] = None, slot: int = 0, data=None, **kwargs):
        BaseLanguageBasedEnv.__init__(self)
        self.config = config if config is not None else CountdownEnvConfig()
        self.data = data if data is not None else load_problems(self.config)
        self.index = None
        self.render_cache = None
        self.render_mode = self.config.render_mode
        assert self.render_mode == "text"
        self.batch = batch if batch is not None else CountdownBatch(1, self.config, device=kwargs.get("device", 0))
        self.slot = slot

    def reset(self, seed=None, mode=None):
        self.index = seed % len(self.data)
        data = self.data[self.index]
        self.render_cache = f"Target: {data['target']}, nums: {data['nums']}"
        return self.render_cache

    def reward_from_result(self, result):
        """0 / format_score / score with the config's own objects (ints stay ints, as in env.py:72-78)."""
        return 0 if result == 0 else self.config.format_score if result == 1 else self.config.score

    def compute_reward(self, action, ground_truth=None):
        gt = ground_truth if ground_truth is not None else self.data[self.index]
        res, uns = self.batch.score([action], [list(gt["nums"])], [gt["target"]])
        self.note_unsupported(action, bool(uns[0]))
        return self.reward_from_result(int(res[0]))

    def note_unsupported(self, action, flag):
        """The device evaluator's `unsupported` bit, surfaced: the answer left the covered grammar of Python expressions
        (include/ragen_b200.h: ragen_countdown_score) and was scored as if eval() had raised - the reference's eval
        (countdown/env.py:16-21) might have scored it as correct.  No host re-evaluation happens; the caller is told."""
        self.last_unsupported = flag
        if flag:
            import warnings

            self.n_unsupported = getattr(self, "n_unsupported", 0) + 1
            warnings.warn(f"Countdown answer outside the device evaluator's grammar, scored as an eval() exception: {action!r}",
                          RuntimeWarning, stacklevel=2)

    def finish_step(self, reward):
        next_obs = f"Your answer get {reward} points."
        info = {"action_is_effective": reward > 0, "action_is_valid": True, "success": reward == self.config.score}
        self.render_cache = next_obs
        return next_obs, reward, True, info

    def step(self, action):
        return self.finish_step(self.compute_reward(action, self.data[self.index]))

    def render(self, mode=None):
        return self.render_cache

    def close(self):
        pass
