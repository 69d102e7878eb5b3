"""GPU parity: FrozenLake and Bandit kernels (through the C ABI) against the reference-made golden fixtures and the
oracle.  Bit-exact: maps, start cells, positions, rewards, done, flags, glyph rows."""
import numpy as np
import pytest
import torch

from conftest import golden_json, golden_npz
from oracle import toy_envs as TE

pytestmark = pytest.mark.gpu


def test_frozenlake_golden():
    from ragen_b200.env.frozen_lake import FrozenLakeBatch

    g = golden_npz("frozen_lake.npz")
    texts = golden_json("frozen_lake_texts.json")
    seeds = g["seeds"]
    b = FrozenLakeBatch(len(seeds), 4, 0.8, True)
    b.reset(seeds)
    assert np.array_equal(b.tiles.cpu().numpy(), g["desc"])
    assert np.array_equal(b.pos.cpu().numpy(), g["start"])
    txt = b.texts(TE.FL_GLYPHS)
    for j in range(4):
        assert txt[j] == texts[f"{j}/reset"]
    acts = g["actions"]
    for t in range(acts.shape[1]):
        out = b.step(torch.from_numpy(acts[:, t].copy()).cuda())
        torch.cuda.synchronize()
        assert np.array_equal(b.pos.cpu().numpy(), g["pos"][:, t]), t
        assert np.array_equal(out.reward.cpu().numpy().astype(np.float32), g["reward"][:, t]), t
        assert np.array_equal(out.done.cpu().numpy(), g["done"][:, t]), t
        assert np.array_equal(out.flags.cpu().numpy() & 3, g["flags"][:, t]), t
        if t < 5:
            txt = b.texts(TE.FL_GLYPHS)
            for j in range(4):
                assert txt[j] == texts[f"{j}/{t}"]


@pytest.mark.parametrize("size,p,slip,n", [(4, 0.8, True, 1024), (8, 0.7, True, 300), (5, 0.9, False, 129), (3, 0.5, True, 77)])
def test_frozenlake_turns_against_oracle(size, p, slip, n):
    """C2 shape (1024 envs, env i seeded with i), K-action turns with budgets; stepping continues after termination."""
    from ragen_b200.env.frozen_lake import FrozenLakeBatch

    rng = np.random.default_rng(2)
    b = FrozenLakeBatch(n, size, p, slip)
    b.reset(np.arange(n))
    envs = [TE.FrozenLakeOracle(size=size, p=p, is_slippery=slip) for _ in range(n)]
    obs0 = [e.reset(i) for i, e in enumerate(envs)]
    assert b.texts(TE.FL_GLYPHS) == obs0
    assert np.array_equal(b.tiles.cpu().numpy(), np.stack([e.tiles() for e in envs]))
    for turn in range(8):
        K = int(rng.integers(1, 5))
        acts = rng.integers(1, 5, (n, K)).astype(np.uint8)
        na = rng.integers(0, K + 1, n).astype(np.uint8)
        out = b.step(torch.from_numpy(acts).cuda(), torch.from_numpy(na).cuda())
        torch.cuda.synchronize()
        rew, done, flags, nex, pos = (x.cpu().numpy() for x in (out.reward, out.done, out.flags, out.n_executed, b.pos))
        txt = b.texts(TE.FL_GLYPHS)
        for i, e in enumerate(envs):
            acc, d, fl, k = 0.0, False, 0, 0
            for a in acts[i, :na[i]]:
                o, r, d, info = e.step(int(a))
                acc += r
                k += 1
                fl = int(info["action_is_effective"]) | (int(info["success"]) << 1) | 4
                if d:
                    break
            assert (rew[i], bool(done[i]), flags[i], nex[i], pos[i]) == (acc, d if k else False, fl, k, e.s), (turn, i)
            assert txt[i] == e.render()


def test_frozenlake_env_view():
    from ragen_b200.env.frozen_lake import FrozenLakeEnv, FrozenLakeEnvConfig

    env = FrozenLakeEnv(FrozenLakeEnvConfig())
    ref = TE.FrozenLakeOracle()
    assert env.reset(seed=42) == ref.reset(42)
    assert env.get_all_actions() == [1, 2, 3, 4]
    for a in [1, 2, 2, 3, 4, 2, 2, 3, 3, 1]:
        assert env.step(a) == ref.step(a)
        assert env.s == ref.s


def test_bandit_golden_and_env_view():
    from ragen_b200.env.bandit import BanditBatch, BanditEnv, BanditEnvConfig

    g = golden_npz("bandit.npz")
    cfg = BanditEnvConfig(lo_arm_name="Phoenix", hi_arm_name="Dragon")
    b = BanditBatch(len(g["seeds"]), cfg)
    b.reset(g["seeds"])
    assert np.array_equal(b.lo_first.cpu().numpy(), g["lo_first"])
    rew, fl = b.step(torch.from_numpy(g["actions"].copy()).cuda())
    torch.cuda.synchronize()
    assert np.array_equal(rew.cpu().numpy(), g["reward"])
    assert np.array_equal((fl.cpu().numpy() >> 1) & 1, g["success"])
    texts = golden_json("bandit_texts.json")
    for j in range(3):
        env = BanditEnv(BanditEnvConfig(lo_arm_name="Phoenix", hi_arm_name="Dragon"))
        obs0 = env.reset(seed=int(g["seeds"][j]))
        obs, r, d, info = env.step(int(g["actions"][j]))
        assert [obs0, obs] == texts[str(j)]
        assert r == g["reward"][j] and d and info["success"] == bool(g["success"][j])
    # inactive envs and invalid action ids execute nothing and do not advance the generator
    n = 64
    b = BanditBatch(n, cfg)
    b.reset(np.arange(n))
    state0 = b.rng.clone()
    a = torch.full((n,), 7, dtype=torch.uint8, device="cuda")
    rew, fl = b.step(a)
    assert int(fl.sum()) == 0 and torch.equal(state0, b.rng)
    ors = [TE.BanditOracle("Phoenix", "Dragon") for _ in range(n)]
    for i, o in enumerate(ors):
        o.reset(i)
    a = torch.from_numpy((1 + np.arange(n) % 2).astype(np.uint8)).cuda()
    rew, fl = b.step(a)
    want = [o.step(int(x))[1] for o, x in zip(ors, a.cpu().numpy())]
    assert rew.cpu().numpy().tolist() == want


def test_env_views_reset_their_own_slot_of_a_shared_batch():
    """BaseEnv.reset(seed) is the plugin contract (ragen/env/base.py:16, called per env at es_manager.py:93-95): a view onto
    a manager-owned batch resets ITS slot; the neighbours keep their episode (map, position, generator state)."""
    from ragen_b200.env.bandit import BanditBatch, BanditEnv, BanditEnvConfig
    from ragen_b200.env.frozen_lake import FrozenLakeBatch, FrozenLakeEnv, FrozenLakeEnvConfig

    n = 9
    b = FrozenLakeBatch(n, 4, 0.8, True)
    b.reset(np.arange(n) + 100)
    refs = [TE.FrozenLakeOracle() for _ in range(n)]
    for i, r in enumerate(refs):
        r.reset(100 + i)
    envs = [FrozenLakeEnv(FrozenLakeEnvConfig(), batch=b, slot=i) for i in range(n)]
    for a in (2, 3):
        for i in (0, 4, 8):
            assert envs[i].step(a) == refs[i].step(a)
    assert envs[4].reset(seed=42) == refs[4].reset(42)  # slot 4 starts over with another seed ...
    assert envs[4].config.map_seed == 42
    for a in (2, 2, 3, 1, 4, 2):
        for i in range(n):  # ... and every env, reset or not, follows its own oracle
            assert envs[i].step(a) == refs[i].step(a), (i, a)
            assert envs[i].s == refs[i].s

    cfg = BanditEnvConfig(lo_arm_name="Phoenix", hi_arm_name="Dragon")
    bb = BanditBatch(n, cfg)
    bb.reset(np.arange(n) + 500)
    benvs = [BanditEnv(BanditEnvConfig(lo_arm_name="Phoenix", hi_arm_name="Dragon"), batch=bb, slot=i) for i in range(n)]
    lo = bb.lo_first.cpu().numpy()
    for e, lf in zip(benvs, lo):
        e.adopt_reset(bool(lf))
    ors = [TE.BanditOracle("Phoenix", "Dragon") for _ in range(n)]
    for i, o in enumerate(ors):
        o.reset(500 + i)
    assert benvs[3].reset(seed=7) == ors[3].reset(7)
    for i in range(n):
        assert benvs[i].step(1 + i % 2) == ors[i].step(1 + i % 2), i


def test_frozenlake_large_batch_properties():
    """1 Mi envs: every generated map has exactly one S and one G, G reachable from S (checked on the host with the
    reference's own DFS rule on a sample), positions stay in range, terminal envs never move."""
    from ragen_b200.env.frozen_lake import FrozenLakeBatch

    n = 1 << 20
    b = FrozenLakeBatch(n, 4, 0.8, True)
    b.reset(np.arange(n))
    t = b.tiles
    assert int((t == 0).sum()) == n and int((t == 3).sum()) == n and int(t.max()) == 3
    assert int(b.status.sum()) == 0
    tiles = t.cpu().numpy()
    for i in range(0, n, 4099):
        board = ["".join("SFHG"[c] for c in row) for row in tiles[i].reshape(4, 4)]
        assert TE._reaches_goal(board, 4)
        assert board == TE.frozenlake_map(4, 0.8, i)
    rng = torch.Generator(device="cuda").manual_seed(1)
    for step in range(20):
        before = b.pos.clone()
        a = torch.randint(1, 5, (n,), dtype=torch.uint8, device="cuda", generator=rng)
        out = b.step(a)
        terminal = torch.gather(t, 1, before.long().unsqueeze(1)).squeeze(1) >= 2
        assert bool((b.pos[terminal] == before[terminal]).all())
        assert int(b.pos.max()) < 16
        assert bool((out.reward[terminal] == 0).all())
