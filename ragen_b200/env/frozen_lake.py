"""FrozenLake on the GPU: device-resident batch + per-env view with the reference API
(ragen/env/frozen_lake/env.py:9-71, config.py:4-22)."""
from dataclasses import dataclass, field
from typing import Dict, Optional

import numpy as np
import torch

from .. import _lib
from .base import BaseDiscreteActionEnv
from .sokoban import StepOut, codes_to_text


@dataclass
class FrozenLakeEnvConfig:
    """Field-for-field the reference dataclass (ragen/env/frozen_lake/config.py:4-22)."""
    size: int = 4
    p: float = 0.8
    is_slippery: bool = True
    map_seed: Optional[int] = None
    render_mode: str = "text"
    action_map: Dict[int, int] = field(default_factory=lambda: {1: 0, 2: 1, 3: 2, 4: 3})
    map_lookup: Dict[bytes, int] = field(default_factory=lambda: {b"P": 0, b"F": 1, b"H": 2, b"G": 3})
    grid_lookup: Dict[int, str] = field(default_factory=lambda: {0: "P", 1: "_", 2: "O", 3: "G", 4: "X", 5: "√"})
    grid_vocab: Dict[str, str] = field(default_factory=lambda: {"P": "player", "_": "empty", "O": "hole", "G": "goal",
                                                                "X": "player in hole", "√": "player on goal"})
    action_lookup: Dict[int, str] = field(default_factory=lambda: {1: "Left", 2: "Down", 3: "Right", 4: "Up"})


def fl_glyph_table(grid_lookup):
    return [grid_lookup.get(i, "?") for i in range(7)] + ["\n"]


class FrozenLakeBatch:
    """n FrozenLake envs in HBM: tiles uint8 [n, size*size] (0 S, 1 F, 2 H, 3 G), pos uint8 [n], one NumPy-exact
    PCG64 generator per env (uint64 [2, n, 2])."""

    def __init__(self, n, size=4, p=0.8, is_slippery=True, device=0):
        self.ctx = _lib.Context.get(device)
        self.lib = self.ctx.lib
        self.device = dev = torch.device("cuda", int(device))
        self.n, self.size, self.p, self.is_slippery = int(n), int(size), float(p), bool(is_slippery)
        C, G = self.size * self.size, self.size * (self.size + 1)
        self.tiles = torch.zeros((self.n, C), dtype=torch.uint8, device=dev)
        self.pos = torch.zeros(self.n, dtype=torch.uint8, device=dev)
        self.rng = torch.zeros((2, self.n, 2), dtype=torch.int64, device=dev)
        self.status = torch.zeros(self.n, dtype=torch.int32, device=dev)
        self.reward = torch.zeros(self.n, dtype=torch.float64, device=dev)
        self.done = torch.zeros(self.n, dtype=torch.uint8, device=dev)
        self.flags = torch.zeros(self.n, dtype=torch.uint8, device=dev)
        self.n_executed = torch.zeros(self.n, dtype=torch.uint8, device=dev)
        self.obs = torch.zeros((self.n, G), dtype=torch.uint8, device=dev)

    def reset(self, seeds, check=True):
        """seeds: int sequence / tensor [n] (non-negative).  Map generation + start state on the device."""
        s = torch.as_tensor(np.asarray(seeds, dtype=np.uint64).astype(np.int64)).to(self.device)
        assert s.numel() == self.n
        _lib.check(self.lib.ragen_frozenlake_reset(self.ctx.handle, self.n, self.size, self.p, _lib.ptr(s), _lib.ptr(self.tiles),
                                                   _lib.ptr(self.pos), _lib.ptr(self.rng), _lib.ptr(self.status), _lib.ptr(self.obs),
                                                   self.ctx.stream()))
        if check and bool(self.status.any()):
            raise RuntimeError("no valid FrozenLake map found for some seeds")
        return self.obs

    def reset_slot(self, slot, seed):
        """BaseEnv.reset(seed) of ONE env of a shared batch (ragen/env/base.py:16, es_manager.py:93-95): the map, the start
        state and the generator of `seed` go into the slot; every other env keeps its episode."""
        one = getattr(self, "_one", None)
        if one is None:
            one = self._one = FrozenLakeBatch(1, self.size, self.p, self.is_slippery, device=self.device.index)
        one.reset([seed])
        self.tiles[slot].copy_(one.tiles[0])
        self.pos[slot] = one.pos[0]
        self.rng[:, slot].copy_(one.rng[:, 0])
        self.obs[slot].copy_(one.obs[0])

    def step(self, actions, n_actions=None, render=True) -> StepOut:
        if actions.dim() == 1:
            actions = actions.unsqueeze(1)
        assert actions.dtype == torch.uint8 and actions.is_cuda and actions.is_contiguous() and actions.shape[0] == self.n
        _lib.check(self.lib.ragen_frozenlake_step(
            self.ctx.handle, self.n, self.size, int(self.is_slippery), _lib.ptr(actions), actions.shape[1], _lib.ptr(n_actions),
            _lib.ptr(self.tiles), _lib.ptr(self.pos), _lib.ptr(self.rng), _lib.ptr(self.reward), _lib.ptr(self.done),
            _lib.ptr(self.flags), _lib.ptr(self.n_executed), _lib.ptr(self.obs) if render else None, self.ctx.stream()))
        return StepOut(self.reward, self.done, self.flags, self.n_executed, self.obs)

    def render(self):
        _lib.check(self.lib.ragen_frozenlake_render(self.ctx.handle, self.n, self.size, _lib.ptr(self.tiles), _lib.ptr(self.pos),
                                                    _lib.ptr(self.obs), self.ctx.stream()))
        return self.obs

    def texts(self, grid_lookup, rows=None):
        tab = fl_glyph_table(grid_lookup)
        codes = (self.obs if rows is None else self.obs[rows]).cpu().numpy()
        return [codes_to_text(r, tab) for r in codes]


class FrozenLakeEnv(BaseDiscreteActionEnv):
    """One env with the reference interface, backed by a slot of a FrozenLakeBatch."""

    def __init__(self, config=None, batch: Optional[FrozenLakeBatch] = None, slot: int = 0, **kwargs):
        BaseDiscreteActionEnv.__init__(self)
        self.config = config or FrozenLakeEnvConfig()
        self.GRID_LOOKUP = self.config.grid_lookup
        self.ACTION_LOOKUP = self.config.action_lookup
        self.render_mode = self.config.render_mode
        self.action_map = self.config.action_map
        self.MAP_LOOKUP = self.config.map_lookup
        if batch is None:
            batch = FrozenLakeBatch(1, self.config.size, self.config.p, self.config.is_slippery, device=kwargs.get("device", 0))
            slot = 0
        self.batch, self.slot = batch, slot
        self._act = self._na = None  # staging tensors of the single-env step path, made on first use
        self.ncol = self.nrow = batch.size

    @property
    def s(self):
        return int(self.batch.pos[self.slot])

    @property
    def desc(self):
        t = self.batch.tiles[self.slot].cpu().numpy().reshape(self.nrow, self.ncol)
        return np.array([b"S", b"F", b"H", b"G"], dtype="S1")[t]

    @property
    def player_pos(self):
        return (self.s // self.ncol, self.s % self.ncol)

    def reset(self, seed=None, mode=None):
        if seed is None:
            seed = int(np.random.SeedSequence().entropy % (2 ** 63))
        self.config.map_seed = seed
        if self.batch.n != 1:
            self.batch.reset_slot(self.slot, seed)  # the neighbours in the batch are not touched
        else:
            self.batch.reset([seed])
        return self.render()

    def _staging(self):
        if self._act is None:
            batch = self.batch
            self._act = torch.zeros((batch.n, 1), dtype=torch.uint8, device=batch.device)
            self._na = torch.zeros(batch.n, dtype=torch.uint8, device=batch.device)

    def step(self, action: int):
        b = self.batch
        self._staging()
        self._act.zero_()
        self._na.zero_()
        self._act[self.slot, 0] = int(self.action_map[action]) + 1
        self._na[self.slot] = 1
        out = b.step(self._act, self._na)
        fl = int(out.flags[self.slot])
        info = {"action_is_effective": bool(fl & 1), "action_is_valid": True, "success": bool(fl & 2)}
        reward = float(out.reward[self.slot])
        return self.render(), (reward if reward else 0), bool(out.done[self.slot]), info

    def render(self, mode=None):
        if (mode or self.render_mode) != "text":
            raise ValueError(f"Invalid mode: {mode or self.render_mode}")
        return self.batch.texts(self.GRID_LOOKUP, rows=[self.slot])[0]

    def get_all_actions(self):
        return list(self.ACTION_LOOKUP.keys())

    def close(self):
        self.render_cache = None
