"""CPU: the oracle restatement reproduces every golden vector generated from the unmodified
reference (tests/golden/make_golden.py)."""
import numpy as np
import pytest
import torch

from conftest import golden_json, golden_npz
from oracle import advantage as A
from oracle import managers as M
from oracle import sokoban as S
from oracle import toy_envs as T


@pytest.mark.parametrize("name", ["s6", "s6d30", "s8", "s6b3", "s10", "s7x9"])
def test_sokoban_generator_stream_exact(rooms, name):
    H, W, nb, depth = (int(x) for x in rooms[f"{name}_cfg"])
    seeds = rooms[f"{name}_seeds"]
    if name == "s10":
        seeds = seeds[:3]  # ~0.1 s per room in the oracle; keep the CPU suite short
    for j, seed in enumerate(seeds):
        try:
            fixed, state = S.generate_room(H, W, nb, depth, int(seed))
            ok = True
        except S.RoomGenerationError:
            ok = False
        assert ok == bool(rooms[f"{name}_ok"][j]), (name, seed)
        if ok:
            assert np.array_equal(fixed, rooms[f"{name}_fixed"][j]), (name, seed)
            assert np.array_equal(state, rooms[f"{name}_state"][j]), (name, seed)
            assert S.check_room(fixed, state, H, W, nb)


@pytest.mark.parametrize("name", ["s6", "s8", "s10", "s6b3"])
def test_sokoban_steps(rooms, steps, name):
    H, W, nb, depth = (int(x) for x in rooms[f"{name}_cfg"])
    max_steps = int(steps[f"{name}_max_steps"])
    texts = golden_json("sokoban_texts.json")
    for j, i in enumerate(steps[f"{name}_idx"]):
        env = S.SokobanOracle((H, W), nb, max_steps)
        env.load(rooms[f"{name}_fixed"][i], rooms[f"{name}_state"][i])
        for t, a in enumerate(steps[f"{name}_actions"][j]):
            obs, r, d, info = env.step(int(a))
            assert np.array_equal(env.state, steps[f"{name}_states"][j, t])
            assert r == steps[f"{name}_reward"][j, t]  # bit-exact doubles
            assert d == bool(steps[f"{name}_done"][j, t])
            fl = int(info["action_is_effective"]) | (int(info["success"]) << 1)
            assert fl == steps[f"{name}_flags"][j, t]
            key = f"{name}/{j}/{t}"
            if key in texts:
                assert obs == texts[key]


def test_sokoban_solver_agrees_with_generated_rooms(rooms):
    for j in range(12):
        fixed, state = rooms["s6_fixed"][j], rooms["s6_state"][j]
        sol = S.solve_bfs(fixed, state, 6, 6)
        assert sol, "every generated room is solvable"
        env = S.SokobanOracle((6, 6), 1, 100)
        env.load(fixed, state)
        for a in sol:
            _, r, d, info = env.step(a)
        assert d and info["success"]


def test_frozen_lake():
    g = golden_npz("frozen_lake.npz")
    texts = golden_json("frozen_lake_texts.json")
    for j, seed in enumerate(g["seeds"]):
        env = T.FrozenLakeOracle()
        obs0 = env.reset(int(seed))
        assert np.array_equal(env.tiles(), g["desc"][j])
        assert env.s == g["start"][j]
        if f"{j}/reset" in texts:
            assert obs0 == texts[f"{j}/reset"]
        for t, a in enumerate(g["actions"][j]):
            obs, r, d, info = env.step(int(a))
            assert env.s == g["pos"][j, t]
            assert r == g["reward"][j, t] and d == bool(g["done"][j, t])
            assert (int(info["action_is_effective"]) | (int(info["success"]) << 1)) == g["flags"][j, t]
            if f"{j}/{t}" in texts:
                assert obs == texts[f"{j}/{t}"]
    for s, flat in enumerate(g["maps8"]):
        assert "".join(T.frozenlake_map(8, 0.7, s)) == str(flat)


def test_bandit():
    g = golden_npz("bandit.npz")
    texts = golden_json("bandit_texts.json")
    for j, seed in enumerate(g["seeds"]):
        env = T.BanditOracle("Phoenix", "Dragon")
        obs0 = env.reset(int(seed))
        assert env.lo_first == bool(g["lo_first"][j])
        obs, r, d, info = env.step(int(g["actions"][j]))
        assert r == g["reward"][j] and d and info["success"] == bool(g["success"][j])
        if str(j) in texts:
            assert [obs0, obs] == texts[str(j)]


def test_countdown():
    for row in golden_json("countdown.json"):
        assert T.countdown_reward(row["answer"], row["nums"], row["target"]) == row["reward"], row


def test_parse_response():
    for c in golden_json("ctx.json")["parse"]:
        resp, actions = M.parse_response(c["in"], c["enable_think"], "||", c["max_actions"])
        assert resp == c["llm_response"] and actions == c["actions"], c


def test_masks_scores_and_normalize():
    g = golden_json("ctx.json")
    ms = g["masks_scores"]
    ids = torch.tensor(ms["ids"])
    for uts in (0, 1):
        for erm in (0, 1):
            s, lm, rm = M.masks_and_scores(ids, ms["scores"], 151644, 151645, bool(uts), bool(erm))
            assert torch.equal(s, torch.tensor(ms[f"score_{uts}{erm}"], dtype=torch.float32))
            assert torch.equal(lm.long(), torch.tensor(ms[f"loss_{uts}{erm}"]))
            assert torch.equal(rm.long(), torch.tensor(ms[f"resp_{uts}{erm}"]))
    nm = g["normalize"]
    last = torch.tensor(nm["last"], dtype=torch.float32)
    pen = torch.tensor(nm["penalty"], dtype=torch.float32)
    keys = {"state": [i // 4 for i in range(16)], "inductive": ["A" if i < 8 else "B" for i in range(16)], "batch": [1] * 16}
    for grouping, gk in keys.items():
        for method in ("identity", "mean", "mean_std", "asym_clip"):
            got = M.normalize_scores(last, pen, gk, method)
            assert torch.equal(got, torch.tensor(nm[f"{grouping}/{method}"], dtype=torch.float32)), (grouping, method)


def test_advantage():
    g = golden_npz("advantage.npz")
    r, v, m = (torch.tensor(g[k]) for k in ("rewards", "values", "mask"))
    for tag in ("a", "b"):
        gamma, lam, hg = (float(x) for x in g[f"cfg_{tag}"])
        adv, ret = A.gae(r, v, m, gamma, lam)
        assert torch.equal(adv, torch.tensor(g[f"gae_{tag}_adv"])) and torch.equal(ret, torch.tensor(g[f"gae_{tag}_ret"]))
        adv, ret = A.bilevel_gae(r, v, m, gamma, lam, hg)
        assert torch.equal(adv, torch.tensor(g[f"bi_{tag}_adv"])) and torch.equal(ret, torch.tensor(g[f"bi_{tag}_ret"]))
    uid = [str(u) for u in g["uid"]]
    adv, _ = A.grpo_outcome(torch.tensor(g["outcome"]), m, uid, norm_by_std=True)
    assert torch.allclose(adv, torch.tensor(g["grpo_std_adv"]), rtol=1e-6, atol=0)
    adv, _ = A.grpo_outcome(torch.tensor(g["outcome"]), m, uid, norm_by_std=False)
    assert torch.equal(adv, torch.tensor(g["grpo_mean_adv"]))
    # the reference's own worked example (ragen/trainer/core_algos.py:96-102)
    r = torch.tensor([[0, 0, 0, 0, 1, 0, 0, 0, 0, 1]])
    v = torch.tensor([[0.1, 0.2, 0.3, 0.4, 0.5, 0.6, 0.7, 0.8, 0.9, 1.0]])
    adv, ret = A.bilevel_gae(r, v, torch.ones(1, 10), 1, 1, 0.95)
    assert torch.equal(adv, torch.tensor(g["example_bi_adv"])) and torch.equal(ret, torch.tensor(g["example_bi_ret"]))
    assert torch.allclose(ret, torch.tensor([[1.95] * 5 + [1.0] * 5]))


def test_es_manager_script():
    g = golden_json("es_manager.json")
    n = len(g["rooms_fixed"])
    envs = []
    for i in range(n):
        envs.append({"tag": "SimpleSokoban", "env": S.SokobanOracle((6, 6), 1, 100), "max_actions_per_traj": 10,
                     "action_lookup": {1: "Up", 2: "Down", 3: "Left", 4: "Right"}})
    mgr = M.ManagerOracle(envs, g["group_size"])
    mgr.reset(g["seed"], reset_fn=lambda i, e, s: e["env"].load(g["rooms_fixed"][i], g["rooms_state"][i]))
    for inputs in g["script"]:
        mgr.step(inputs)
    final = mgr.rollout_states()
    for got, want in zip(final, g["final"]):
        assert got["penalty"] == want["penalty"]
        assert got["metrics"] == want["metrics"]
        assert len(got["history"]) == len(want["history"])
        for hg, hw in zip(got["history"], want["history"]):
            for k in ("state", "actions_left", "actions", "reward", "info", "llm_response"):
                assert hg.get(k) == hw.get(k), (k, hg.get(k), hw.get(k))
    for st, want in zip(mgr.status, g["status"]):
        assert st == want


def test_es_manager_rooms_are_group_seeded(rooms):
    """reset(seed=123): group g uses seed 123+g for all its envs (es_manager.py:80-82)."""
    g = golden_json("es_manager.json")
    seeds = list(rooms["s6_seeds"])
    for i, (f, s) in enumerate(zip(g["rooms_fixed"], g["rooms_state"])):
        j = seeds.index(123 + i // g["group_size"])
        assert np.array_equal(np.array(f, dtype=np.uint8), rooms["s6_fixed"][j])
        assert np.array_equal(np.array(s, dtype=np.uint8), rooms["s6_state"][j])


@pytest.mark.parametrize("name", ["s6", "s8", "s10", "s6b3"])
def test_batched_numpy_oracle_matches_golden_steps(rooms, steps, name):
    """oracle/sokoban_np.py (the vectorised checker used at large n) against the reference-made fixtures."""
    from oracle.sokoban_np import SokobanBatchOracle

    H, W, nb, depth = (int(x) for x in rooms[f"{name}_cfg"])
    idx = steps[f"{name}_idx"]
    acts = steps[f"{name}_actions"]
    b = SokobanBatchOracle(rooms[f"{name}_fixed"][idx], rooms[f"{name}_state"][idx], H, W, nb, int(steps[f"{name}_max_steps"]))
    done_before = np.zeros(len(idx), dtype=bool)
    for t in range(acts.shape[1]):
        rew, done, eff, won = b.step_one(acts[:, t])
        live = ~done_before  # the fixtures stop recording meaningful data after done? no: they keep stepping
        assert np.array_equal(b.state, steps[f"{name}_states"][:, t])
        assert np.array_equal(rew, steps[f"{name}_reward"][:, t])
        assert np.array_equal(done.astype(np.uint8), steps[f"{name}_done"][:, t])
        assert np.array_equal(eff.astype(np.uint8) | (won.astype(np.uint8) << 1), steps[f"{name}_flags"][:, t])


def test_batched_numpy_oracle_turns_match_scalar_oracle():
    from oracle.sokoban_np import SokobanBatchOracle, synthetic_rooms

    H, W, nb = 7, 6, 2
    fixed, state = synthetic_rooms(200, H, W, nb, seed=11)
    rng = np.random.default_rng(5)
    b = SokobanBatchOracle(fixed, state, H, W, nb, 9)
    envs = []
    for i in range(200):
        e = S.SokobanOracle((H, W), nb, 9)
        e.load(fixed[i], state[i])
        envs.append(e)
    for turn in range(4):
        K = 4
        acts = rng.integers(0, 5, (200, K)).astype(np.uint8)
        na = rng.integers(0, K + 1, 200)
        acc, done, flags, nexec = b.step_turn(acts, na)
        for i, e in enumerate(envs):
            a_acc, a_done, a_fl, a_n = 0.0, 0, 0, 0
            for k in range(int(na[i])):
                _, r, d, info = e.step(int(acts[i, k]))
                a_acc += r
                a_n += 1
                a_fl = int(info["action_is_effective"]) | (int(info["success"]) << 1) | 4
                a_done = int(d)
                if d:
                    break
            assert (a_acc, a_done, a_fl, a_n) == (acc[i], done[i], flags[i], nexec[i]), (turn, i)
            assert np.array_equal(e.state, b.state[i])
            assert np.array_equal(e.glyph_codes().reshape(-1), b.glyphs()[i].reshape(H, W + 1)[:, :W].reshape(-1))
