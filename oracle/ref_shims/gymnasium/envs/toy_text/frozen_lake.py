"""Shim of gymnasium.envs.toy_text.frozen_lake.FrozenLakeEnv (test infrastructure).

Restates the published behaviour of Farama gymnasium's FrozenLake for the members RAGEN
touches (ragen/env/frozen_lake/env.py:15, 21-26, 33, 41; attributes s, ncol, desc, np_random).
Un-vendored and un-pinned (reference requirements.txt:20-21) -> "parity unpinned" at this
boundary; see SURVEY.md section 8c.

Actions: 0 left, 1 down, 2 right, 3 up.  A step from a terminal cell (G/H) stays put with reward
0; slippery motion draws uniformly among (a-1)%4, a, (a+1)%4 via one np_random.random() call.
"""
import numpy as np

from ... import Env

LEFT, DOWN, RIGHT, UP = 0, 1, 2, 3


def _categorical_sample(probs, rng):
    return int(np.argmax(np.cumsum(np.asarray(probs)) > rng.random()))


class FrozenLakeEnv(Env):
    metadata = {"render_modes": ["human", "ansi", "rgb_array"], "render_fps": 4}

    def __init__(self, render_mode=None, desc=None, map_name="4x4", is_slippery=True):
        if desc is None:
            raise NotImplementedError("shim requires an explicit desc (RAGEN always passes one)")
        self.desc = desc = np.asarray(desc, dtype="c")
        self.nrow, self.ncol = nrow, ncol = desc.shape
        self.reward_range = (0, 1)
        self.initial_state_distrib = np.array(desc == b"S").astype("float64").ravel()
        self.initial_state_distrib /= self.initial_state_distrib.sum()

        def shifted(r, c, a):
            if a == LEFT:
                c = max(c - 1, 0)
            elif a == DOWN:
                r = min(r + 1, nrow - 1)
            elif a == RIGHT:
                c = min(c + 1, ncol - 1)
            elif a == UP:
                r = max(r - 1, 0)
            return r, c

        def outcome(r, c, a):
            nr, nc = shifted(r, c, a)
            letter = desc[nr, nc]
            return nr * ncol + nc, float(letter == b"G"), bytes(letter) in b"GH"

        self.P = {}
        for r in range(nrow):
            for c in range(ncol):
                s = r * ncol + c
                self.P[s] = {}
                for a in range(4):
                    if bytes(desc[r, c]) in b"GH":
                        rows = [(1.0, s, 0, True)]
                    elif is_slippery:
                        rows = [(1.0 / 3.0, *outcome(r, c, b)) for b in ((a - 1) % 4, a, (a + 1) % 4)]
                    else:
                        rows = [(1.0, *outcome(r, c, a))]
                    self.P[s][a] = rows
        self.render_mode = render_mode
        self.lastaction = None

    def step(self, a):
        rows = self.P[self.s][a]
        i = _categorical_sample([row[0] for row in rows], self.np_random)
        p, s, r, t = rows[i]
        self.s = s
        self.lastaction = a
        return int(s), r, t, False, {"prob": p}

    def reset(self, *, seed=None, options=None):
        super().reset(seed=seed)
        self.s = _categorical_sample(self.initial_state_distrib, self.np_random)
        self.lastaction = None
        return int(self.s), {"prob": 1}

    def _render_gui(self, mode):
        raise NotImplementedError("rgb_array rendering is out of scope")
