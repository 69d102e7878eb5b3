"""Cross-check of the two third-party shims against the REAL packages, wherever those are installed.

The arithmetic of SokobanEnv.step lives in gym_sokoban and FrozenLake's transition draw in gymnasium; neither is vendored
under the reference nor installable in the build container, so oracle/ref_shims restates their published algorithms
 and parity is "unpinned" at that boundary.  These tests close the gap on any machine that has the packages:
they are skipped (importorskip) where the packages are absent."""
import importlib
import os
import sys

import numpy as np
import pytest

from conftest import ROOT

SHIMS = os.path.join(ROOT, "oracle", "ref_shims")


def _real(name):
    """Import `name` from site-packages, never from oracle/ref_shims."""
    saved = list(sys.path)
    sys.path[:] = [p for p in sys.path if os.path.abspath(p or ".") != SHIMS]
    for k in [k for k in sys.modules if k == name or k.startswith(name + ".")]:
        if SHIMS in (getattr(sys.modules[k], "__file__", "") or ""):
            del sys.modules[k]
    try:
        return pytest.importorskip(name)
    finally:
        sys.path[:] = saved


def test_gym_sokoban_step_equals_the_oracle():
    _real("gym_sokoban")
    real = importlib.import_module("gym_sokoban.envs.sokoban_env")
    if SHIMS in (real.__file__ or ""):
        pytest.skip("only the shim is importable")
    from oracle.sokoban_np import SokobanBatchOracle, synthetic_rooms

    n, H, W, nb = 64, 7, 7, 2
    fixed, state = synthetic_rooms(n, H, W, nb, seed=5)
    rng = np.random.default_rng(6)
    acts = rng.integers(1, 5, (n, 40))
    o = SokobanBatchOracle(fixed, state, H, W, nb, 30)
    envs = []
    for i in range(n):
        e = real.SokobanEnv(dim_room=(H, W), max_steps=30, num_boxes=nb, reset=False)
        e.room_fixed = fixed[i].reshape(H, W).astype(np.int64).copy()
        e.room_state = state[i].reshape(H, W).astype(np.int64).copy()
        e.player_position = np.argwhere(e.room_state == 5)[0]
        e.num_env_steps, e.reward_last, e.boxes_on_target = 0, 0, 0
        envs.append(e)
    done_before = np.zeros(n, dtype=bool)
    for t in range(acts.shape[1]):
        rew, done, eff, won = o.step_one(acts[:, t], ~done_before)
        for i, e in enumerate(envs):
            if done_before[i]:
                continue
            _, r, d, _ = e.step(int(acts[i, t]), observation_mode="tiny_rgb_array")
            assert r == rew[i] and bool(d) == bool(done[i]), (i, t)
            assert np.array_equal(e.room_state.reshape(-1), o.state[i]), (i, t)
        done_before |= done


def test_gymnasium_frozen_lake_equals_the_oracle():
    _real("gymnasium")
    fl = importlib.import_module("gymnasium.envs.toy_text.frozen_lake")
    if SHIMS in (fl.__file__ or ""):
        pytest.skip("only the shim is importable")
    from oracle import toy_envs as TE

    rng = np.random.default_rng(8)
    for seed in range(40):
        o = TE.FrozenLakeOracle()
        o.reset(seed)
        desc = ["".join(ch.decode() if isinstance(ch, bytes) else str(ch) for ch in row) for row in o.desc]
        env = fl.FrozenLakeEnv(desc=desc, is_slippery=True)
        s, _ = env.reset(seed=seed)
        assert int(s) == o.s
        for a in rng.integers(1, 5, 30):
            want = o.step(int(a))
            s, r, term, trunc, _ = env.step({1: 0, 2: 1, 3: 2, 4: 3}[int(a)])
            assert int(s) == o.s and float(r) == float(want[1]) and bool(term) == bool(want[2])
