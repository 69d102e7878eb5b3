"""GPU parity: ragen_sokoban_step / render / adopt (through the C ABI) against the oracle and the
reference-made golden fixtures.  Bit-exact: states, doubles, done, flags, glyph rows."""
import numpy as np
import pytest
import torch

from conftest import golden_json
from oracle import sokoban as S
from oracle.sokoban_np import SokobanBatchOracle, synthetic_rooms

pytestmark = pytest.mark.gpu


def _batch(n, H, W, nb, max_steps):
    from ragen_b200.env.sokoban import SokobanBatch

    return SokobanBatch(n, (H, W), nb, max_steps)


@pytest.mark.parametrize("name", ["s6", "s8", "s10", "s6b3"])
def test_golden_step_streams(rooms, steps, name):
    H, W, nb, depth = (int(x) for x in rooms[f"{name}_cfg"])
    idx = steps[f"{name}_idx"]
    acts = steps[f"{name}_actions"]
    b = _batch(len(idx), H, W, nb, int(steps[f"{name}_max_steps"]))
    b.load_rooms(rooms[f"{name}_fixed"][idx], rooms[f"{name}_state"][idx])
    texts = golden_json("sokoban_texts.json")
    for t in range(acts.shape[1]):
        out = b.step(torch.from_numpy(acts[:, t].copy()).cuda())
        torch.cuda.synchronize()
        assert np.array_equal(b.state.cpu().numpy(), steps[f"{name}_states"][:, t]), t
        assert np.array_equal(out.reward.cpu().numpy(), steps[f"{name}_reward"][:, t]), t  # bit-exact doubles
        assert np.array_equal(out.done.cpu().numpy(), steps[f"{name}_done"][:, t]), t
        assert np.array_equal(out.flags.cpu().numpy() & 3, steps[f"{name}_flags"][:, t]), t
        assert (out.n_executed.cpu().numpy() == 1).all()
        txt = b.texts(S.DEFAULT_GLYPHS)
        for j in range(len(idx)):
            key = f"{name}/{j}/{t}"
            if key in texts:
                assert txt[j] == texts[key]


@pytest.mark.parametrize("H,W,nb,n", [(6, 6, 1, 128), (6, 6, 1, 1000), (10, 10, 2, 4099), (8, 8, 2, 515), (7, 9, 3, 77), (5, 4, 1, 1)])
def test_turns_against_batched_oracle(H, W, nb, n):
    """K-action turns with per-env action budgets, ragged tile tails (n not a multiple of 128)."""
    fixed, state = synthetic_rooms(n, H, W, nb, seed=n)
    rng = np.random.default_rng(n + 1)
    max_steps = 12
    o = SokobanBatchOracle(fixed, state, H, W, nb, max_steps)
    b = _batch(n, H, W, nb, max_steps)
    b.load_rooms(fixed, state)
    assert np.array_equal(b.player.cpu().numpy(), o.player)
    assert np.array_equal(b.obs.cpu().numpy(), o.glyphs())
    for turn in range(6):
        K = int(rng.integers(1, 6))
        acts = rng.integers(0, 5, (n, K)).astype(np.uint8)
        na = rng.integers(0, K + 1, n).astype(np.uint8)
        use_na = turn % 2 == 0
        acc, done, flags, nexec = o.step_turn(acts, na if use_na else None)
        out = b.step(torch.from_numpy(acts).cuda(), torch.from_numpy(na).cuda() if use_na else None)
        torch.cuda.synchronize()
        assert np.array_equal(b.state.cpu().numpy(), o.state)
        assert np.array_equal(out.reward.cpu().numpy(), acc)
        assert np.array_equal(out.done.cpu().numpy(), done)
        assert np.array_equal(out.flags.cpu().numpy(), flags)
        assert np.array_equal(out.n_executed.cpu().numpy(), nexec)
        assert np.array_equal(b.player.cpu().numpy(), o.player)
        assert np.array_equal(b.env_steps.cpu().numpy().astype(np.int64), o.env_steps)
        assert np.array_equal(b.on_target.cpu().numpy().astype(np.int64), o.on_target)
        assert np.array_equal(out.obs.cpu().numpy(), o.glyphs())


def test_full_size_properties():
    """BASELINE config 4 size on one GPU (1 Mi envs 10x10/2 boxes): size-independent invariants -
    object conservation, glyph rows == render of the state, reward decomposition, and equality with
    the batched oracle on a strided sample of envs."""
    n, H, W, nb = 1 << 20, 10, 10, 2
    fixed, state = synthetic_rooms(n, H, W, nb, seed=4)
    rng = np.random.default_rng(4)
    b = _batch(n, H, W, nb, 100)
    b.load_rooms(fixed, state)
    sample = np.arange(0, n, 257)
    o = SokobanBatchOracle(fixed[sample], state[sample], H, W, nb, 100)
    total = np.zeros(n)
    for t in range(10):
        a = rng.integers(1, 5, n).astype(np.uint8)
        out = b.step(torch.from_numpy(a).cuda())
        st = b.state
        assert int((st == 5).sum()) == n  # exactly one player per env
        assert int(((st == 3) | (st == 4)).sum()) == n * nb  # boxes conserved
        walls_same = ((st == 0) == (b.fixed == 0)).all()
        assert bool(walls_same)
        rew, done, eff, won = o.step_one(a[sample])
        assert np.array_equal(out.reward.cpu().numpy()[sample], rew)
        assert np.array_equal(b.state.cpu().numpy()[sample], o.state)
        assert np.array_equal(out.obs.cpu().numpy()[sample], o.glyphs())
        assert np.array_equal(out.done.cpu().numpy()[sample], done.astype(np.uint8))
        # rewards only take the values the reward rule can produce
        r = out.reward
        vals = torch.tensor([-0.1, -0.1 + 1, -0.1 + -1, -0.1 + 1 + 10, -0.1 + 10], dtype=torch.float64, device="cuda")
        assert bool(torch.isin(r, vals).all())
    # render kernel alone reproduces the fused rows
    fused = b.obs.clone()
    b.obs.zero_()
    b.render()
    assert torch.equal(fused, b.obs)


def test_full_size_compact_kernel_against_oracle_sample():
    """The instance bench.py times - sokoban_tile_kernel<10, 10, 0, 3>: nibble rooms, three-bit rows - at BASELINE config 4's
    size (1 Mi envs 10x10 / 2 boxes): equality with the batched oracle on a strided sample (states, rewards, done, flags,
    three-bit glyph rows), conservation invariants on ALL envs, K = 1 and K = 2 turns."""
    from ragen_b200.env.sokoban import SokobanBatch, unpack3_glyphs

    n, H, W, nb = 1 << 20, 10, 10, 2
    fixed, state = synthetic_rooms(n, H, W, nb, seed=14)
    rng = np.random.default_rng(14)
    b = SokobanBatch(n, (H, W), nb, 100, compact=True)
    b.load_rooms(fixed, state)
    sample = np.arange(3, n, 251)
    o = SokobanBatchOracle(fixed[sample], state[sample], H, W, nb, 100)
    walls = b.fixed == 0
    for t in range(10):
        K = 1 if t % 3 else 2
        a = rng.integers(1, 5, (n, K)).astype(np.uint8)
        out = b.step(torch.from_numpy(a).cuda(), packed=3)
        acc, done, flags, nexec = o.step_turn(a[sample])
        assert np.array_equal(out.reward.cpu().numpy()[sample], acc)
        assert np.array_equal(out.done.cpu().numpy()[sample], done) and np.array_equal(out.flags.cpu().numpy()[sample], flags)
        assert np.array_equal(out.n_executed.cpu().numpy()[sample], nexec)
        assert np.array_equal(unpack3_glyphs(out.obs[torch.from_numpy(sample).cuda()].cpu().numpy(), H, W), o.glyphs())
        s4 = b.state4
        lo, hi = s4 & 15, s4 >> 4
        assert int((lo == 5).sum() + (hi == 5).sum()) == n  # exactly one player per env
        assert int(((lo == 3) | (lo == 4)).sum() + ((hi == 3) | (hi == 4)).sum()) == n * nb  # boxes conserved
        st = b.state
        assert np.array_equal(st.cpu().numpy()[sample], o.state)
        assert bool(((st == 0) == walls).all())
    # the three-bit rows of ALL envs are the render of the byte view (render kernel, byte rooms) in the same format
    fused = out.obs.clone()
    assert torch.equal(fused, b.render(packed=3).clone())


def test_env_view_matches_reference_api(rooms):
    """SokobanEnv (per-env view with the reference's reset/step/render signature)."""
    from ragen_b200.env.sokoban import SokobanEnv, SokobanEnvConfig

    cfg = SokobanEnvConfig(dim_room=(6, 6), num_boxes=1, max_steps=100)
    env = SokobanEnv(cfg)
    ref = S.SokobanOracle((6, 6), 1, 100)
    obs0 = env.load_room(rooms["s6_fixed"][3], rooms["s6_state"][3])
    assert obs0 == ref.load(rooms["s6_fixed"][3], rooms["s6_state"][3])
    assert env.get_all_actions() == [1, 2, 3, 4]
    for a in [1, 4, 4, 2, 3, 3, 1, 2, 4]:
        got = env.step(a)
        want = ref.step(a)
        assert got == want
        assert np.array_equal(env.room_state.reshape(-1), ref.state)
        assert tuple(env.player_position) == divmod(ref.player, 6)


def test_abi_argument_errors():
    from ragen_b200 import _lib

    ctx = _lib.Context.get(0)
    rc = ctx.lib.ragen_sokoban_step(ctx.handle, 4, 6, 6, 1, 100, None, 1, None, None, None, None, None, None, None, None,
                                    None)
    assert rc == -1 and b"NULL" in ctx.lib.ragen_last_error()
    rc = ctx.lib.ragen_sokoban_step(ctx.handle, 4, 1, 6, 1, 100, None, 1, None, None, None, None, None, None, None, None,
                                    None)
    assert rc == -1 and b"bad shape" in ctx.lib.ragen_last_error()


def test_adopt_rejects_rooms_the_generator_cannot_produce():
    """status 1 no player, 2 border not wall, 3 room_fixed not recoverable from room_state (a plain box
    standing on a target cell) - the step kernel relies on that invariant instead of reading room_fixed."""
    fixed, state = synthetic_rooms(4, 6, 6, 1, seed=3)
    fixed, state = fixed.copy(), state.copy()
    state[0][state[0] == 5] = fixed[0][state[0] == 5]          # no player
    fixed[1][0] = 1; state[1][0] = 1                            # hole in the border
    box = int(np.flatnonzero(state[2] == 4)[0]); fixed[2][box] = 2  # box (code 4) on a target cell
    b = _batch(4, 6, 6, 1, 100)
    with pytest.raises(ValueError):
        b.load_rooms(fixed, state)
    assert b.status.cpu().tolist() == [1, 2, 3, 0]


@pytest.mark.parametrize("name", ["s6", "s6d30", "s8", "s6b3", "s7x9", "s10"])
def test_generator_reproduces_reference_rooms(rooms, name):
    """ragen_sokoban_generate: seed for seed the rooms the unmodified reference generated, failures included."""
    H, W, nb, depth = (int(x) for x in rooms[f"{name}_cfg"])
    seeds = rooms[f"{name}_seeds"]
    from ragen_b200.env.sokoban import SokobanBatch

    b = SokobanBatch(1, (H, W), nb, 100, search_depth=depth)
    fixed, state, status, explored = b.generate(seeds)
    torch.cuda.synchronize()
    ok = rooms[f"{name}_ok"]
    st = status.cpu().numpy()
    assert np.array_equal(st == 0, ok)
    assert np.array_equal(fixed.cpu().numpy()[ok], rooms[f"{name}_fixed"][ok])
    assert np.array_equal(state.cpu().numpy()[ok], rooms[f"{name}_state"][ok])
    assert int(explored.max()) <= 300000


def test_batch_reset_group_seeds_and_env_reset(rooms):
    """EnvStateManager-style reset: 8 groups x 16 envs, group g seeded with 123 + g (es_manager.py:80-95)."""
    from ragen_b200.env.sokoban import SokobanBatch, SokobanEnv, SokobanEnvConfig

    b = SokobanBatch(128, (6, 6), 1, 100)
    seeds = [123 + i // 16 for i in range(128)]
    used = b.reset(seeds)
    assert used == seeds
    idx = {int(s): j for j, s in enumerate(rooms["s6_seeds"])}
    want_f = np.stack([rooms["s6_fixed"][idx[s]] for s in seeds])
    want_s = np.stack([rooms["s6_state"][idx[s]] for s in seeds])
    assert np.array_equal(b.fixed.cpu().numpy(), want_f) and np.array_equal(b.state.cpu().numpy(), want_s)
    env = SokobanEnv(SokobanEnvConfig(dim_room=(6, 6), num_boxes=1))
    ref = S.SokobanOracle((6, 6), 1, 100)
    assert env.reset(seed=1010) == ref.reset(seed=1010)
    for a in [1, 2, 3, 4, 4, 1]:
        assert env.step(a) == ref.step(a)


def test_generated_rooms_are_valid_and_solvable():
    """Seeds the fixtures do not cover: every room passes the reference's validity rules and the BFS solver."""
    from ragen_b200.env.sokoban import SokobanBatch

    b = SokobanBatch(1, (6, 6), 1, 100)
    seeds = np.arange(5000, 5000 + 2048)
    fixed, state, status, explored = b.generate(seeds)
    st = status.cpu().numpy()
    assert (st == 0).mean() > 0.9
    f, s = fixed.cpu().numpy(), state.cpu().numpy()
    for j in np.flatnonzero(st == 0)[:300]:
        S.check_room(f[j], s[j], 6, 6, 1)
        assert S.solve_bfs(f[j], s[j], 6, 6), int(seeds[j])  # solvable (get_shortest_action_path contract)
    for j in np.flatnonzero(st == 0)[:64]:
        want_f, want_s = S.generate_room(6, 6, 1, 300, int(seeds[j]))
        assert np.array_equal(f[j], want_f) and np.array_equal(s[j], want_s)


@pytest.mark.parametrize("H,W,nb,n", [(6, 6, 1, 1000), (10, 10, 2, 4099), (8, 8, 2, 515), (5, 5, 1, 130), (10, 10, 2, 1), (6, 4, 1, 300), (6, 3, 1, 9)])
def test_packed_glyph_rows_equal_the_byte_rows(H, W, nb, n):
    """ragen_sokoban_step_packed / ragen_sokoban_render_packed: the same step, glyph rows as two 4-bit codes per byte.
    Unpacked they equal the oracle's rows; translated with the pair table they are the reference's text."""
    from ragen_b200.env.sokoban import glyph_pair_table, packed_to_text, unpack_glyphs

    fixed, state = synthetic_rooms(n, H, W, nb, seed=n + 7)
    rng = np.random.default_rng(n)
    o = SokobanBatchOracle(fixed, state, H, W, nb, 50)
    b = _batch(n, H, W, nb, 50)
    b.load_rooms(fixed, state)
    assert np.array_equal(unpack_glyphs(b.render(packed=True).cpu().numpy()), o.glyphs())
    pairs = glyph_pair_table(S.DEFAULT_GLYPHS)
    for turn in range(8):
        K = int(rng.integers(1, 4))
        acts = rng.integers(0, 5, (n, K)).astype(np.uint8)
        acc, done, flags, nexec = o.step_turn(acts, None)
        out = b.step(torch.from_numpy(acts).cuda(), packed=True)
        assert out.obs.shape == (n, H * (W + 1) // 2)
        assert np.array_equal(b.state.cpu().numpy(), o.state)
        assert np.array_equal(out.reward.cpu().numpy(), acc) and np.array_equal(out.flags.cpu().numpy(), flags)
        packed = out.obs.cpu().numpy()
        assert np.array_equal(unpack_glyphs(packed), o.glyphs())  # includes code 6, player on a target
    b.render()
    want = b.texts(S.DEFAULT_GLYPHS)
    assert [packed_to_text(r, pairs) for r in packed[:64]] == want[:64]
    # host-buffer path
    acts = rng.integers(1, 5, n).astype(np.uint8)
    o.step_turn(acts[:, None], None)
    h_obs, h_rew, h_done, h_flags, h_nex = b.step_host(acts, packed=True)
    assert np.array_equal(unpack_glyphs(h_obs), o.glyphs())


def test_packed_needs_even_row_stream():
    b = _batch(4, 7, 6, 1, 10)  # H*(W+1) = 49
    with pytest.raises(ValueError):
        b.render(packed=True)
    from ragen_b200 import _lib
    buf = torch.zeros(4 * 25, dtype=torch.uint8, device="cuda")
    rc = b.lib.ragen_sokoban_render_packed(b.ctx.handle, 4, 7, 6, _lib.ptr(b.state), _lib.ptr(b.meta), _lib.ptr(buf), b.ctx.stream())
    assert rc != 0 and b"even" in b.lib.ragen_last_error()


def test_solver_returns_the_reference_paths():
    """ragen_sokoban_solve against the paths the unmodified get_shortest_action_path returned (tests/golden/sokoban_paths.json),
    action for action, one launch per room shape; then the generated rooms of a launch are all solvable (the reference's own
    validity criterion) and the lengths equal the oracle BFS on a sample."""
    from ragen_b200 import _lib
    from ragen_b200.env.sokoban import get_shortest_action_path, solve_rooms

    G = golden_json("sokoban_paths.json")
    by_shape = {}
    for e in G["rooms"]:
        by_shape.setdefault((e["H"], e["W"], e["max_depth"]), []).append(e)
    for (H, W, md), es in by_shape.items():
        f = np.array([e["fixed"] for e in es], dtype=np.uint8)
        s = np.array([e["state"] for e in es], dtype=np.uint8)
        status, length, path, explored = solve_rooms(f, s, (H, W), max_depth=md)
        status, length, path = status.cpu().numpy(), length.cpu().numpy(), path.cpu().numpy()
        for j, e in enumerate(es):
            assert path[j, : length[j]].tolist() == e["path"], e["tag"]
            assert status[j] in (0, 1) and (status[j] == 0 or length[j] == 0), e["tag"]
    e = next(x for x in G["rooms"] if len(x["path"]) > 8)
    assert get_shortest_action_path(np.array(e["fixed"]).reshape(e["H"], e["W"]), np.array(e["state"]).reshape(e["H"], e["W"])) == e["path"]

    b = _batch(16, 10, 10, 2, 100)
    fixed, state, st, _ = b.generate(np.arange(7000, 7000 + 512))
    ok = (st == 0)
    status, length, path, explored = solve_rooms(fixed[ok], state[ok], (10, 10), 2)
    assert int((status != 0).sum()) == 0 and int(length.min()) >= 1  # every generated room is solvable, none starts solved
    f_np, s_np, l_np, p_np = fixed[ok].cpu().numpy(), state[ok].cpu().numpy(), length.cpu().numpy(), path.cpu().numpy()
    for j in range(12):
        assert p_np[j, : l_np[j]].tolist() == list(S.solve_bfs(f_np[j], s_np[j], 10, 10))
    # a queue that is too small is reported, not guessed
    ctx = b.ctx
    st2 = torch.empty(1, dtype=torch.int32, device="cuda"); ln2 = torch.empty(1, dtype=torch.int32, device="cuda")
    big = int(explored.argmax())
    _lib.check(ctx.lib.ragen_sokoban_solve(ctx.handle, 1, 10, 10, _lib.ptr(fixed[ok][big:big + 1].contiguous()), _lib.ptr(state[ok][big:big + 1].contiguous()),
                                           100, 64, _lib.ptr(st2), _lib.ptr(ln2), None, None, ctx.stream()))
    assert int(st2[0]) == 2 and int(ln2[0]) == 0


@pytest.mark.parametrize("H,W,nb,n", [(6, 6, 1, 1000), (10, 10, 2, 4099), (8, 8, 2, 515), (10, 10, 2, 1), (6, 6, 1, 129)])
def test_three_bit_glyph_rows_equal_the_byte_rows(H, W, nb, n):
    """ragen_sokoban_step_packed3 / ragen_sokoban_render_packed3: the same step, glyph rows as three bits per cell without
    separators.  Unpacked (separators put back) they equal the oracle's rows, code 6 (player on a target) included; the
    host-buffer path brings rows and scalars back with one copy; other shapes are refused."""
    from ragen_b200 import _lib
    from ragen_b200.env.sokoban import packed3_bytes, unpack3_glyphs

    fixed, state = synthetic_rooms(n, H, W, nb, seed=n + 11)
    rng = np.random.default_rng(n + 1)
    o = SokobanBatchOracle(fixed, state, H, W, nb, 50)
    b = _batch(n, H, W, nb, 50)
    b.load_rooms(fixed, state)
    assert np.array_equal(unpack3_glyphs(b.render(packed=3).cpu().numpy(), H, W), o.glyphs())
    saw_six = False
    for turn in range(10):
        K = int(rng.integers(1, 4))
        acts = rng.integers(0, 5, (n, K)).astype(np.uint8)
        acc, done, flags, nexec = o.step_turn(acts, None)
        out = b.step(torch.from_numpy(acts).cuda(), packed=3)
        assert out.obs.shape == (n, packed3_bytes(H, W))
        assert np.array_equal(b.state.cpu().numpy(), o.state)
        assert np.array_equal(out.reward.cpu().numpy(), acc) and np.array_equal(out.flags.cpu().numpy(), flags)
        got = unpack3_glyphs(out.obs.cpu().numpy(), H, W)
        assert np.array_equal(got, o.glyphs())
        saw_six = saw_six or bool((got == 6).any())
    assert saw_six or n < 100
    t3 = b.texts(S.DEFAULT_GLYPHS)[:32]  # decoded from the three-bit rows
    b.render()
    assert t3 == b.texts(S.DEFAULT_GLYPHS)[:32]  # the byte rows' text
    acts = rng.integers(1, 5, n).astype(np.uint8)
    acc, done, flags, nexec = o.step_turn(acts[:, None], None)
    h_obs, h_rew, h_done, h_flags, h_nex = b.step_host(acts, packed=3)
    assert np.array_equal(unpack3_glyphs(h_obs, H, W), o.glyphs())
    assert np.array_equal(h_rew, acc) and np.array_equal(h_done, done) and np.array_equal(h_flags, flags) and np.array_equal(h_nex, nexec)
    # the nibble path still comes back right through the shared slab
    acc, done, flags, nexec = o.step_turn(acts[:, None], None)
    h_obs, h_rew, h_done, h_flags, h_nex = b.step_host(acts, packed=True)
    from ragen_b200.env.sokoban import unpack_glyphs
    assert np.array_equal(unpack_glyphs(h_obs), o.glyphs()) and np.array_equal(h_rew, acc) and np.array_equal(h_flags, flags)


def test_three_bit_rows_refuse_other_shapes():
    from ragen_b200 import _lib

    b = _batch(4, 7, 6, 1, 50)
    with pytest.raises(ValueError):
        b.render(packed=3)
    buf = torch.zeros(4 * 16, dtype=torch.uint8, device="cuda")
    rc = b.lib.ragen_sokoban_render_packed3(b.ctx.handle, 4, 7, 6, _lib.ptr(b.state), _lib.ptr(b.meta), _lib.ptr(buf), b.ctx.stream())
    assert rc != 0


@pytest.mark.parametrize("H,W,nb,n", [(6, 6, 1, 1000), (10, 10, 2, 4099), (8, 8, 2, 515), (10, 10, 2, 1), (6, 6, 1, 129), (10, 10, 3, 257)])
def test_compact_rooms_step_like_the_byte_rooms(H, W, nb, n):
    """ragen_sokoban_step_compact: the rooms live in HBM as nibbles (two cells per byte); a turn reads, steps and writes
    them back in that form.  Unpacked, room_state equals the oracle's after every turn, and so do the rewards, flags,
    n_executed and the three-bit glyph rows; turns without a render and n_actions < K are covered; the other glyph
    formats go through render() on the unpacked rooms."""
    from ragen_b200.env.sokoban import SokobanBatch, unpack3_glyphs, unpack_glyphs

    fixed, state = synthetic_rooms(n, H, W, nb, seed=n + 23)
    rng = np.random.default_rng(n + 5)
    o = SokobanBatchOracle(fixed, state, H, W, nb, 40)
    b = SokobanBatch(n, (H, W), nb, 40, compact=True)
    b.load_rooms(fixed, state)
    s4 = b.state4.cpu().numpy()
    assert np.array_equal(s4 & 15, state.reshape(n, -1)[:, 0::2]) and np.array_equal(s4 >> 4, state.reshape(n, -1)[:, 1::2])
    for turn in range(12):
        K = int(rng.integers(1, 4))
        acts = rng.integers(0, 5, (n, K)).astype(np.uint8)
        na = rng.integers(0, K + 1, n).astype(np.uint8) if turn % 3 == 2 else None
        acc, done, flags, nexec = o.step_turn(acts, na)
        render = turn % 4 != 1
        out = b.step(torch.from_numpy(acts).cuda(), None if na is None else torch.from_numpy(na).cuda(), render=render, packed=3)
        assert np.array_equal(b.state.cpu().numpy(), o.state)
        assert np.array_equal(out.reward.cpu().numpy(), acc) and np.array_equal(out.done.cpu().numpy(), done)
        assert np.array_equal(out.flags.cpu().numpy(), flags) and np.array_equal(out.n_executed.cpu().numpy(), nexec)
        if render:
            assert np.array_equal(unpack3_glyphs(out.obs.cpu().numpy(), H, W), o.glyphs())
    assert np.array_equal(b.render().cpu().numpy().reshape(n, -1), o.glyphs())
    assert np.array_equal(unpack_glyphs(b.render(packed=True).cpu().numpy()), o.glyphs())
    acts = rng.integers(1, 5, n).astype(np.uint8)
    acc, done, flags, nexec = o.step_turn(acts[:, None], None)
    h_obs, h_rew, h_done, h_flags, h_nex = b.step_host(acts, packed=3)
    assert np.array_equal(unpack3_glyphs(h_obs, H, W), o.glyphs())
    assert np.array_equal(h_rew, acc) and np.array_equal(h_done, done) and np.array_equal(h_flags, flags) and np.array_equal(h_nex, nexec)
    with pytest.raises(ValueError):
        b.step(torch.from_numpy(acts[:, None]).cuda(), packed=True)
    # a partial reset re-packs the rooms it touched and leaves the others alone
    before = b.state.clone()
    meta_before = b.meta.clone()
    slots = [0, n // 2, n - 1] if n >= 3 else [0]
    b.reset([7, 7, 9][: len(slots)], slots=slots)
    after = b.state
    touched = torch.zeros(n, dtype=torch.bool, device="cuda")
    touched[slots] = True
    assert torch.equal(after[~touched], before[~touched])
    assert torch.equal(b.meta[~touched], meta_before[~touched])  # steps / stored box count of envs in mid-episode survive
    assert bool((b.meta[touched][:, 1] == 0).all())
    s4 = b.state4
    assert torch.equal(s4 & 15, after[:, 0::2]) and torch.equal(s4 >> 4, after[:, 1::2])


@pytest.mark.parametrize("compact", [False, True])
def test_slot_reset_leaves_the_other_envs_in_mid_episode(compact):
    """BaseEnv.reset on ONE env of a shared batch (ragen/env/base.py:16): the neighbours keep num_env_steps and the stored
    boxes_on_target, so their next rewards and their max_steps truncation are the oracle's."""
    from ragen_b200.env.sokoban import SokobanBatch, SokobanEnv, SokobanEnvConfig

    n, H, W, nb, max_steps = 300, 6, 6, 1, 9
    fixed, state = synthetic_rooms(n, H, W, nb, seed=91)
    rng = np.random.default_rng(92)
    o = SokobanBatchOracle(fixed, state, H, W, nb, max_steps)
    b = SokobanBatch(n, (H, W), nb, max_steps, search_depth=30, compact=compact)
    b.load_rooms(fixed, state)
    kw = {"packed": 3} if compact else {}
    for turn in range(3):
        acts = rng.integers(1, 5, (n, 2)).astype(np.uint8)
        want = o.step_turn(acts)
        out = b.step(torch.from_numpy(acts).cuda(), **kw)
        assert np.array_equal(out.reward.cpu().numpy(), want[0])
    # reset two slots (through the batch and through an env view); everybody else carries on against the oracle
    slots = [5, 200]
    b.reset([11], slots=slots[:1])
    env = SokobanEnv(SokobanEnvConfig(dim_room=(H, W), num_boxes=nb, max_steps=max_steps, search_depth=30), batch=b, slot=slots[1])
    env.reset(seed=12)
    keep = np.ones(n, dtype=bool)
    keep[slots] = False
    assert np.array_equal(b.env_steps.cpu().numpy().astype(np.int64)[keep], o.env_steps[keep])
    assert np.array_equal(b.on_target.cpu().numpy().astype(np.int64)[keep], o.on_target[keep])
    assert (b.env_steps.cpu().numpy()[slots] == 0).all()
    for turn in range(3):  # 6 + 6 steps > max_steps 9: truncation must arrive on time for the untouched envs
        acts = rng.integers(1, 5, (n, 2)).astype(np.uint8)
        acc, done, flags, nexec = o.step_turn(acts)
        out = b.step(torch.from_numpy(acts).cuda(), **kw)
        assert np.array_equal(out.reward.cpu().numpy()[keep], acc[keep])
        assert np.array_equal(out.done.cpu().numpy()[keep], done[keep]) and np.array_equal(out.n_executed.cpu().numpy()[keep], nexec[keep])
        assert np.array_equal(b.state.cpu().numpy()[keep], o.state[keep])


@pytest.mark.parametrize("H,W,nb,n,K", [(10, 10, 2, 4099, 1), (6, 6, 1, 1000, 3), (8, 8, 2, 515, 2), (10, 10, 3, 129, 5)])
def test_step_deltas_rebuild_rooms_rewards_and_rows_on_the_host(H, W, nb, n, K):
    """ragen_sokoban_step_delta: 4 bytes per action instead of reward + flags + a re-rendered room.  Over 50 random turns (wins,
    truncations, envs that keep being stepped after done, n_actions < K, partial resets in between) the host mirror patched from
    the deltas equals the full three-bit rows of a twin batch, decode_delta equals the result arrays, and both equal the oracle."""
    from ragen_b200.env.sokoban import HostMirror, SokobanBatch, decode_delta, unpack3_glyphs

    fixed, state = synthetic_rooms(n, H, W, nb, seed=n + 7)
    rng = np.random.default_rng(n + 8)
    max_steps = 30
    o = SokobanBatchOracle(fixed, state, H, W, nb, max_steps)
    b = SokobanBatch(n, (H, W), nb, max_steps, search_depth=20, compact=True)
    twin = SokobanBatch(n, (H, W), nb, max_steps, search_depth=20, compact=True)
    one = SokobanBatch(n, (H, W), nb, max_steps, search_depth=20, rooms="glyph3" if K % 2 else "nibbles")  # one-byte records
    b.load_rooms(fixed, state)
    twin.load_rooms(fixed, state)
    one.load_rooms(fixed, state)
    mirror = b.host_mirror()
    mirror1 = one.host_mirror()
    assert np.array_equal(mirror.glyphs(), o.glyphs()) and np.array_equal(mirror1.glyphs(), o.glyphs())
    for turn in range(50):
        acts = rng.integers(0, 5, (n, K)).astype(np.uint8)
        na = rng.integers(0, K + 1, n).astype(np.uint8) if turn % 4 == 3 else None
        d_act = torch.from_numpy(acts).cuda()
        d_na = None if na is None else torch.from_numpy(na).cuda()
        delta = b.step_delta(d_act, d_na).cpu().numpy()
        delta1 = one.step_delta(d_act, d_na, record_bytes=1).cpu().numpy()
        full = twin.step(d_act, d_na, packed=3)
        acc, done, flags, nexec = o.step_turn(acts, na)
        mirror.apply(delta)
        mirror1.apply(delta1, acts)
        assert delta1.shape == (n, K) and np.array_equal(delta1 & 0x7f, delta[:, :, 3])  # the info byte; bit 7: a box moved
        assert np.array_equal(delta1 >> 7, ((delta[:, :, 2] != 0xff) & (delta[:, :, 3] != 0)).astype(np.uint8))
        for got in (decode_delta(delta), decode_delta(delta1)):
            for g, w, f in zip(got, (acc, done, flags, nexec), (full.reward, full.done, full.flags, full.n_executed)):
                assert np.array_equal(g, w) and np.array_equal(g, f.cpu().numpy()), turn
        assert np.array_equal(mirror.state, o.state) and np.array_equal(b.state.cpu().numpy(), o.state)
        assert np.array_equal(mirror1.state, o.state) and np.array_equal(one.state.cpu().numpy(), o.state)
        assert np.array_equal(mirror.glyphs(), unpack3_glyphs(full.obs.cpu().numpy(), H, W)), turn
        assert np.array_equal(mirror1.glyphs(), mirror.glyphs())
        if turn == 20 and H == 6:  # a partial reset: the mirror is refreshed for the slots that were reset
            slots = [1, n // 2]
            b.reset([5, 6], slots=slots)
            twin.reset([5, 6], slots=slots)
            one.reset([5, 6], slots=slots)
            st, fx = b.state.cpu().numpy(), b.fixed.cpu().numpy()
            mirror.state[slots], mirror.fixed[slots] = st[slots], fx[slots]
            mirror1.refresh(slots, fx[slots], st[slots])
            o = SokobanBatchOracle(fx, st, H, W, nb, max_steps)
            o.env_steps = b.env_steps.cpu().numpy().astype(np.int64)
            o.on_target = b.on_target.cpu().numpy().astype(np.int64)


def test_pipelined_host_steps_equal_the_plain_ones():
    """step_host_begin / step_host_end (two tickets in flight, copy stream) against step_host on a twin batch, both formats."""
    from ragen_b200.env.sokoban import SokobanBatch, decode_delta

    n, H, W, nb = 5000, 10, 10, 2
    fixed, state = synthetic_rooms(n, H, W, nb, seed=3)
    rng = np.random.default_rng(4)
    for fmt in ("delta", "delta1", 3):
        a = SokobanBatch(n, (H, W), nb, 100, compact=True)
        t = SokobanBatch(n, (H, W), nb, 100, compact=True)
        a.load_rooms(fixed, state)
        t.load_rooms(fixed, state)
        acts = [torch.from_numpy(rng.integers(1, 5, n).astype(np.uint8)).pin_memory() for _ in range(6)]
        tickets = [a.step_host_begin(acts[0], fmt)]
        for i in range(6):
            if i + 1 < 6:
                tickets.append(a.step_host_begin(acts[i + 1], fmt))  # step i+1 is queued before step i's results are read
            res = a.step_host_end(tickets[i])
            rows, rew, done, flags, nex = t.step_host(acts[i], packed=3)
            if fmt in ("delta", "delta1"):
                acc, d, f, k = decode_delta(res)
                assert np.array_equal(acc, rew) and np.array_equal(d, done) and np.array_equal(f, flags) and np.array_equal(k, nex)
            else:
                for g, w in zip(res, (rows, rew, done, flags, nex)):
                    assert np.array_equal(g, w), i
        assert torch.equal(a.state, t.state)


@pytest.mark.parametrize("compact", [False, True])
def test_text_rows_rendered_on_the_device_are_the_reference_strings(compact):
    """ragen_glyph_utf8: the env's text observation as UTF-8 bytes written by the device, from byte, nibble and three-bit glyph
    rows; multi-byte glyphs ("√", and a custom vocabulary) included."""
    from ragen_b200.env.sokoban import SokobanBatch

    n, H, W, nb = 777, 6, 6, 2
    fixed, state = synthetic_rooms(n, H, W, nb, seed=31)
    rng = np.random.default_rng(32)
    b = SokobanBatch(n, (H, W), nb, 100, compact=compact)
    b.load_rooms(fixed, state)
    o = SokobanBatchOracle(fixed, state, H, W, nb, 100)
    lookups = [S.DEFAULT_GLYPHS, {0: "█", 1: ".", 2: "◎", 3: "√", 4: "□", 5: "@", 6: "é"}]
    for turn in range(8):
        acts = rng.integers(1, 5, (n, 3)).astype(np.uint8)
        o.step_turn(acts)
        fmts = [3] if compact else [False, True, 3]
        fmt = fmts[turn % len(fmts)]
        b.step(torch.from_numpy(acts).cuda(), packed=fmt)
        for lk in lookups:
            rows, lens = b.text_rows(lk)
            host = rows.cpu().numpy()
            got = [bytes(r[:k]).decode("utf-8") for r, k in zip(host, lens.cpu().numpy())]
            tab = [lk.get(i, "?") for i in range(7)] + ["\n"]
            want = ["".join(tab[c] for c in g) for g in o.glyphs()]
            assert got == want, (turn, fmt)
            as_strings = host.view(f"S{host.shape[1]}").reshape(-1)  # NUL padding: a fixed-width bytes array
            assert as_strings[5].decode("utf-8") == want[5]


@pytest.mark.parametrize("H,W,nb,n", [(6, 6, 1, 1000), (10, 10, 2, 4099), (8, 8, 2, 515), (10, 10, 2, 1), (6, 6, 1, 129), (10, 10, 3, 257)])
def test_glyph_rooms_are_their_own_observation(H, W, nb, n):
    """ragen_sokoban_step_glyph3: the rooms rest in HBM as three-bit glyph codes (6 = player on a target), so the room IS the
    rendered observation.  After every turn state3 equals the oracle's glyph rows in the three-bit layout (what the other step
    entry points RENDER), room_state (6 read as 5) equals the oracle's state, and rewards / done / flags / n_executed agree;
    K-action turns, n_actions < K, ragged tiles, partial resets, step_host, deltas and the other render formats included."""
    from ragen_b200.env.sokoban import SokobanBatch, decode_delta, unpack3_glyphs, unpack_glyphs

    fixed, state = synthetic_rooms(n, H, W, nb, seed=n + 41)
    rng = np.random.default_rng(n + 42)
    o = SokobanBatchOracle(fixed, state, H, W, nb, 40)
    b = SokobanBatch(n, (H, W), nb, 40, search_depth=20, rooms="glyph3")
    b.load_rooms(fixed, state)
    assert np.array_equal(unpack3_glyphs(b.state3.cpu().numpy(), H, W), o.glyphs())  # adopt wrote the glyph rooms
    for turn in range(14):
        K = int(rng.integers(1, 4))
        acts = rng.integers(0, 5, (n, K)).astype(np.uint8)
        na = rng.integers(0, K + 1, n).astype(np.uint8) if turn % 3 == 2 else None
        acc, done, flags, nexec = o.step_turn(acts, na)
        d_act = torch.from_numpy(acts).cuda()
        d_na = None if na is None else torch.from_numpy(na).cuda()
        if turn % 4 == 3:  # the same kernel reporting deltas
            delta = b.step_delta(d_act, d_na).cpu().numpy()
            got = decode_delta(delta)
        else:
            out = b.step(d_act, d_na, packed=3)
            assert out.obs.data_ptr() == b.state3.data_ptr()  # the observation is the room
            got = (out.reward.cpu().numpy(), out.done.cpu().numpy(), out.flags.cpu().numpy(), out.n_executed.cpu().numpy())
            delta = None
        for g, w in zip(got, (acc, done, flags, nexec)):
            assert np.array_equal(g, w), turn
        assert np.array_equal(unpack3_glyphs(b.state3.cpu().numpy(), H, W), o.glyphs()), turn
        assert np.array_equal(b.state.cpu().numpy(), o.state), turn
        assert np.array_equal(b.player.cpu().numpy(), o.player) and np.array_equal(b.on_target.cpu().numpy().astype(np.int64), o.on_target)
    assert np.array_equal(b.render().cpu().numpy().reshape(n, -1), o.glyphs())
    assert np.array_equal(unpack_glyphs(b.render(packed=True).cpu().numpy()), o.glyphs())
    assert b.texts(S.DEFAULT_GLYPHS)[0] == "".join(S.DEFAULT_GLYPHS.get(int(c), "?") if c != 7 else "\n" for c in o.glyphs()[0][:-1])
    acts = rng.integers(1, 5, n).astype(np.uint8)
    acc, done, flags, nexec = o.step_turn(acts[:, None], None)
    h_obs, h_rew, h_done, h_flags, h_nex = b.step_host(acts, packed=3)
    assert np.array_equal(unpack3_glyphs(h_obs, H, W), o.glyphs())
    assert np.array_equal(h_rew, acc) and np.array_equal(h_done, done) and np.array_equal(h_flags, flags) and np.array_equal(h_nex, nexec)
    rows, lens = b.text_rows(S.DEFAULT_GLYPHS)
    tab = [S.DEFAULT_GLYPHS.get(i, "?") for i in range(7)] + ["\n"]
    host = rows.cpu().numpy()
    assert [bytes(r[:k]).decode() for r, k in zip(host[:50], lens.cpu().numpy()[:50])] == ["".join(tab[c] for c in g) for g in o.glyphs()[:50]]
    # a partial reset writes the glyph rooms of the slots it touched and leaves the others alone
    before = b.state3.clone()
    slots = [0, n // 2, n - 1] if n >= 3 else [0]
    b.reset([7, 7, 9][: len(slots)], slots=slots)
    touched = torch.zeros(n, dtype=torch.bool, device="cuda")
    touched[slots] = True
    assert torch.equal(b.state3[~touched], before[~touched])
    st = b.state.cpu().numpy()
    fx = b.fixed.cpu().numpy()
    want = np.where((st == 5) & (fx == 2), 6, st).reshape(n, H, W)
    assert np.array_equal(unpack3_glyphs(b.state3.cpu().numpy(), H, W).reshape(n, H, W + 1)[:, :, :W], want)


def test_glyph_rooms_at_full_size():
    """The bench's kernel instance sokoban_tile_kernel<10, 10, 0, 4> at 1 Mi envs: oracle equality on a strided sample, conservation
    on all envs."""
    from ragen_b200.env.sokoban import SokobanBatch, unpack3_glyphs

    n, H, W, nb = 1 << 20, 10, 10, 2
    fixed, state = synthetic_rooms(n, H, W, nb, seed=24)
    rng = np.random.default_rng(24)
    b = SokobanBatch(n, (H, W), nb, 100, rooms="glyph3")
    b.load_rooms(fixed, state)
    sample = np.arange(5, n, 251)
    d_sample = torch.from_numpy(sample).cuda()
    o = SokobanBatchOracle(fixed[sample], state[sample], H, W, nb, 100)
    walls = b.fixed == 0
    for t in range(10):
        K = 1 if t % 3 else 2
        a = rng.integers(1, 5, (n, K)).astype(np.uint8)
        out = b.step(torch.from_numpy(a).cuda(), packed=3)
        acc, done, flags, nexec = o.step_turn(a[sample])
        assert np.array_equal(out.reward.cpu().numpy()[sample], acc)
        assert np.array_equal(out.done.cpu().numpy()[sample], done) and np.array_equal(out.flags.cpu().numpy()[sample], flags)
        assert np.array_equal(unpack3_glyphs(b.state3[d_sample].cpu().numpy(), H, W), o.glyphs())
        st = b.state
        assert int((st == 5).sum()) == n and int(((st == 3) | (st == 4)).sum()) == n * nb
        assert bool(((st == 0) == walls).all())


def test_compact_rooms_refuse_other_shapes():
    from ragen_b200 import _lib
    from ragen_b200.env.sokoban import SokobanBatch

    with pytest.raises(ValueError):
        SokobanBatch(4, (7, 6), 1, 50, compact=True)
    b = _batch(4, 7, 6, 1, 50)
    z = torch.zeros(4 * 21, dtype=torch.uint8, device="cuda")
    a = torch.ones(4, dtype=torch.uint8, device="cuda")
    rc = b.lib.ragen_sokoban_step_compact(b.ctx.handle, 4, 7, 6, 1, 50, _lib.ptr(a), 1, None, _lib.ptr(z), _lib.ptr(b.meta), _lib.ptr(b.reward),
                                          _lib.ptr(b.done), _lib.ptr(b.flags), _lib.ptr(b.n_executed), None, b.ctx.stream())
    assert rc != 0


def test_captured_host_stepper_equals_the_plain_host_steps():
    """SokobanBatch.host_stepper: [actions up | step | records down] replayed as one CUDA graph, two steppers in flight on two
    batches; records, rooms and counters equal step_host's on twin batches, and building a stepper leaves the rooms alone."""
    from ragen_b200.env.sokoban import SokobanBatch, decode_delta

    n, H, W, nb = 5000, 10, 10, 2
    fixed, state = synthetic_rooms(n, H, W, nb, seed=5)
    rng = np.random.default_rng(6)
    for rooms, rb in (("glyph3", 1), ("nibbles", 4)):
        pair = [SokobanBatch(n, (H, W), nb, 100, rooms=rooms) for _ in range(2)]
        twins = [SokobanBatch(n, (H, W), nb, 100, rooms=rooms) for _ in range(2)]
        for b in pair + twins:
            b.load_rooms(fixed, state)
        bufs = [torch.zeros(n, dtype=torch.uint8).pin_memory() for _ in range(2)]
        steppers = [pair[j].host_stepper(bufs[j], record_bytes=rb) for j in range(2)]
        assert all(torch.equal(p_.state, t_.state) for p_, t_ in zip(pair, twins))  # construction executed no action
        acts = [rng.integers(0, 5, n).astype(np.uint8) for _ in range(8)]
        bufs[0].numpy()[:] = acts[0]
        steppers[0].launch()
        for i in range(8):
            if i + 1 < 8:  # step i+1 (the other batch) is queued before step i's records are read
                bufs[(i + 1) & 1].numpy()[:] = acts[i + 1]
                steppers[(i + 1) & 1].launch()
            rec = steppers[i & 1].wait()
            rows, rew, done, flags, nex = twins[i & 1].step_host(acts[i], packed=3)
            acc, d, f, k = decode_delta(rec)
            assert np.array_equal(acc, rew) and np.array_equal(d, done) and np.array_equal(f, flags) and np.array_equal(k, nex), i
        for p_, t_ in zip(pair, twins):
            assert torch.equal(p_.state, t_.state) and torch.equal(p_.env_steps, t_.env_steps)
