"""Shim of the `gym` names used by ragen/env/sokoban/env.py:1,18 (test infrastructure)."""
from . import spaces


class Env:
    metadata = {}

    def reset(self, *a, **k):
        raise NotImplementedError

    def step(self, action):
        raise NotImplementedError

    def close(self):
        pass
