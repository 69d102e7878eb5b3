"""Shim of the `gymnasium` names used by ragen/env/{frozen_lake,bandit,countdown}/env.py
(test infrastructure).  Env.reset(seed) follows gymnasium.core.Env.reset: a non-None seed
re-creates ``self.np_random`` via ``seeding.np_random(seed)``."""
from . import spaces
from .utils import seeding


class Env:
    metadata = {}
    _np_random = None
    _np_random_seed = None

    @property
    def np_random(self):
        if self._np_random is None:
            self._np_random, self._np_random_seed = seeding.np_random()
        return self._np_random

    @np_random.setter
    def np_random(self, value):
        self._np_random = value
        self._np_random_seed = -1

    def reset(self, *, seed=None, options=None):
        if seed is not None:
            self._np_random, self._np_random_seed = seeding.np_random(seed)

    def close(self):
        pass
