"""CPU, only where /root/reference exists (this container): the oracle against the unmodified
reference on inputs beyond the committed fixtures."""
import random

import numpy as np
import pytest
import torch

from oracle import advantage as A
from oracle import ref_loader as R
from oracle import sokoban as S
from oracle import toy_envs as T

pytestmark = pytest.mark.skipif(not R.available(), reason="reference not present (GPU box)")


def _ref_room(H, W, nb, depth, seed):
    U = R.sokoban_utils()
    all_seed = R.module("ragen.utils").all_seed
    try:
        with all_seed(seed):
            f, s, _, _ = U.generate_room(dim=(H, W), num_steps=int(1.7 * (H + W)), num_boxes=nb, search_depth=depth)
        return f.astype(np.uint8).ravel(), s.astype(np.uint8).ravel()
    except (RuntimeError, RuntimeWarning):
        return None


@pytest.mark.parametrize("cfg", [(6, 6, 1, 300, 2000, 40), (8, 8, 2, 10, 3000, 40), (6, 6, 2, 100, 4000, 20)])
def test_generator_matches_reference(cfg):
    H, W, nb, depth, s0, n = cfg
    for seed in range(s0, s0 + n):
        want = _ref_room(H, W, nb, depth, seed)
        try:
            got = S.generate_room(H, W, nb, depth, seed)
        except S.RoomGenerationError:
            got = None
        assert (want is None) == (got is None)
        if want is not None:
            assert np.array_equal(want[0], got[0]) and np.array_equal(want[1], got[1])


def test_generator_leaves_global_rng_untouched():
    random.seed(99)
    np.random.seed(99)
    a = (random.random(), np.random.rand())
    random.seed(99)
    np.random.seed(99)
    S.generate_room(6, 6, 1, 300, 5)
    assert (random.random(), np.random.rand()) == a


def test_sokoban_env_matches_reference_with_generated_reset():
    se = R.sokoban_env()
    rng = np.random.default_rng(0)
    for seed in range(700, 712):
        env = se.SokobanEnv(se.SokobanEnvConfig(dim_room=(6, 6), num_boxes=1, max_steps=20))
        orc = S.SokobanOracle((6, 6), 1, 20)
        assert env.reset(seed=seed) == orc.reset(seed)
        for _ in range(25):
            a = int(rng.integers(0, 5))
            x, y = env.step(a), orc.step(a)
            assert x[0] == y[0] and x[1] == y[1] and bool(x[2]) == y[2]
            assert {k: bool(v) for k, v in x[3].items()} == y[3]


def test_solver_matches_reference_length(rooms=None):
    U = R.sokoban_utils()
    for seed in range(5):
        f, s = S.generate_room(6, 6, 1, 300, seed)
        want = U.get_shortest_action_path(f.reshape(6, 6).astype(np.int64), s.reshape(6, 6).astype(np.int64))
        got = S.solve_bfs(f, s, 6, 6)
        assert len(want) == len(got) and len(got) > 0


def test_frozen_lake_and_bandit_match_reference():
    fl, be = R.frozen_lake_env(), R.bandit_env()
    rng = np.random.default_rng(3)
    for seed in range(300, 330):
        e, o = fl.FrozenLakeEnv(fl.FrozenLakeEnvConfig()), T.FrozenLakeOracle()
        assert e.reset(seed=seed) == o.reset(seed)
        for _ in range(20):
            a = int(rng.integers(1, 5))
            x, y = e.step(a), o.step(a)
            assert x[:3] == y[:3] and {k: bool(v) for k, v in x[3].items()} == y[3]
        b, bo = be.BanditEnv(), T.BanditOracle()
        assert b.reset(seed=seed) == bo.reset(seed)
        assert b.step(1 + seed % 2) == bo.step(1 + seed % 2)


def test_advantage_matches_reference():
    ca = R.ragen_core_algos()
    g = torch.Generator().manual_seed(11)
    B, Tn = 5, 50
    mask = (torch.rand(B, Tn, generator=g) > 0.4).float()
    mask[:, -1] = 1
    vals = torch.randn(B, Tn, generator=g) * mask
    rew = torch.zeros(B, Tn)
    for b in range(B):
        idx = torch.nonzero(mask[b]).flatten()
        rew[b, idx[::4]] = 0.3
        rew[b, idx[-1]] = 1.5
    for gamma, lam, hg in ((1.0, 1.0, 0.95), (0.97, 0.9, 0.8)):
        a1, r1 = ca.compute_gae_advantage_return(rew, vals, mask, gamma, lam)
        a2, r2 = A.gae(rew, vals, mask, gamma, lam)
        assert torch.equal(a1, a2) and torch.equal(r1, r2)
        a1, r1 = ca.compute_bi_level_gae_advantage_return(rew, vals, mask, gamma, lam, hg)
        a2, r2 = A.bilevel_gae(rew, vals, mask, gamma, lam, hg)
        assert torch.equal(a1, a2) and torch.equal(r1, r2)
