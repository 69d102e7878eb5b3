"""Extra checks run by __graft_entry__.smoke(): one small invocation of each remaining kernel family on cuda:0,
each compared with the oracle (the oracle is only the checker here)."""
import numpy as np
import torch


def run():
    from oracle import advantage as A
    from oracle import managers as M
    from oracle import toy_envs as TE
    from ragen_b200 import ops
    from ragen_b200.env.bandit import BanditBatch, BanditEnvConfig
    from ragen_b200.env.frozen_lake import FrozenLakeBatch

    notes = []
    # FrozenLake: map generation + 10 slippery steps, 256 envs
    n = 256
    b = FrozenLakeBatch(n, 4, 0.8, True)
    b.reset(np.arange(n))
    envs = [TE.FrozenLakeOracle() for _ in range(n)]
    for i, e in enumerate(envs):
        e.reset(i)
    rng = np.random.default_rng(0)
    for _ in range(10):
        a = rng.integers(1, 5, n).astype(np.uint8)
        out = b.step(torch.from_numpy(a).cuda())
        want = [e.step(int(x)) for e, x in zip(envs, a)]
        assert out.reward.cpu().tolist() == [float(w[1]) for w in want]
        assert b.pos.cpu().tolist() == [e.s for e in envs]
    assert b.texts(TE.FL_GLYPHS) == [e.render() for e in envs]
    notes.append("frozenlake reset+step bit-exact on 256 envs")
    # Bandit
    bb = BanditBatch(64, BanditEnvConfig())
    bb.reset(np.arange(64))
    ors = [TE.BanditOracle() for _ in range(64)]
    for i, o in enumerate(ors):
        o.reset(i)
    rew, _ = bb.step(torch.ones(64, dtype=torch.uint8, device="cuda"))
    assert rew.cpu().tolist() == [o.step(1)[1] for o in ors]
    notes.append("bandit bit-exact on 64 envs")
    # parse + lookup
    texts = ["<think>t</think><answer>Up || jump || LEFT</answer>", "no tags", "<think>a</think>b</think><answer>Down</answer>"]
    res = ops.parse_actions(texts, ops.ActionTables([{1: "Up", 2: "Down", 3: "Left", 4: "Right"}], "cuda")).to_host()
    for t, (resp, actions, codes) in zip(texts, res):
        assert (resp, actions) == M.parse_response(t)
    assert [r[2] for r in res] == [[1, 3], [], [2]]
    notes.append("parse_actions matches the regex oracle")
    # GAE + whiten
    g = torch.Generator().manual_seed(0)
    B, T = 16, 200
    mask = (torch.rand(B, T, generator=g) > 0.5).float()
    r = torch.randn(B, T, generator=g) * (torch.rand(B, T, generator=g) > 0.9)
    v = torch.randn(B, T, generator=g) * mask
    adv, ret, mom = ops.gae_raw(r.cuda(), v.cuda(), mask.cuda(), 1.0, 1.0)
    wa, wr = A.gae_raw(r, v, 1.0, 1.0)
    assert torch.equal(adv.cpu(), wa) and torch.equal(ret.cpu(), wr)
    ops.whiten_(adv, mom)
    assert float((adv.cpu() - A.masked_whiten(wa, mask)).abs().max()) < 1e-5 * max(1.0, float(wa.abs().max()))
    notes.append("gae bit-exact, whiten within 1e-5")
    # generator + solver: 64 rooms 6x6, every valid one solvable, paths equal the oracle BFS
    from oracle import sokoban as S
    from ragen_b200.env.sokoban import SokobanBatch, solve_rooms

    sb = SokobanBatch(1, (6, 6), 1, 100)
    fixed, state, st, _ = sb.generate(list(range(64)))
    ok = st == 0
    status, length, path, _ = solve_rooms(fixed[ok], state[ok], (6, 6), 1)
    assert int((status != 0).sum()) == 0
    f_np, s_np, l_np, p_np = fixed[ok].cpu().numpy(), state[ok].cpu().numpy(), length.cpu().numpy(), path.cpu().numpy()
    for j in range(8):
        assert p_np[j, : l_np[j]].tolist() == list(S.solve_bfs(f_np[j], s_np[j], 6, 6))
    notes.append("generated 6x6 rooms all solvable, solver paths equal the oracle BFS")
    return notes
