"""GPU parity: ragen_sokoban_render_rgb (through the C ABI) against the oracle's restatement of gym_sokoban's room_to_rgb.
Bit-exact bytes.  The sprites are synthetic (every byte random): gym_sokoban's PNGs are not in this image, so the pixel
VALUES of a real run are unpinned; the blit (which sprite lands where, player-on-target substitution) is what is checked."""
import json

import numpy as np
import pytest
import torch

from oracle import cfg as C
from oracle import sokoban as S
from oracle.sokoban_np import SokobanBatchOracle, synthetic_rooms

pytestmark = pytest.mark.gpu


def _stepped(n, H, W, nb, steps=5):
    from ragen_b200.env.sokoban import SokobanBatch

    fixed, state = synthetic_rooms(n, H, W, nb, seed=11 * n + H)
    o = SokobanBatchOracle(fixed, state, H, W, nb, 100)
    b = SokobanBatch(n, (H, W), nb, 100)
    b.load_rooms(fixed, state)
    rng = np.random.default_rng(n)
    for _ in range(steps):
        acts = rng.integers(1, 5, (n, 1)).astype(np.uint8)
        o.step_turn(acts, None)
        b.step(torch.from_numpy(acts).cuda())
    assert np.array_equal(b.state.cpu().numpy(), o.state)
    return b, o


def _want(o, surfaces, rows):
    H, W = o.H, o.W
    return np.stack([S.room_to_rgb(o.state[i].reshape(H, W), o.fixed[i].reshape(H, W), surfaces) for i in rows])


@pytest.mark.parametrize("H,W,nb,n,side", [(6, 6, 1, 300, 16), (10, 10, 2, 1500, 16), (7, 9, 3, 77, 16), (8, 8, 2, 64, 16),
                                           (6, 6, 1, 50, 1), (10, 10, 2, 33, 3), (5, 4, 1, 9, 8), (6, 6, 1, 5, 32)])
def test_blit_matches_oracle(H, W, nb, n, side):
    b, o = _stepped(n, H, W, nb)
    surfaces = S.synthetic_surfaces(side, seed=side)
    b.set_atlas(surfaces)
    got = b.render_rgb().cpu().numpy()
    assert got.shape == (n, H * side, W * side, 3) and got.dtype == np.uint8
    assert np.array_equal(got, _want(o, surfaces, range(n)))
    # the substitution the text render also makes (env.py:55): some env of the sample stands on a target
    if n >= 300:
        assert ((o.state == S.PLAYER) & (o.fixed == S.TARGET)).any()


def test_row_subset_and_errors():
    b, o = _stepped(200, 6, 6, 1)
    surfaces = S.synthetic_surfaces(16, seed=3)
    b.set_atlas(surfaces)
    rows = [199, 0, 17, 17, 42]
    got = b.render_rgb(rows=rows).cpu().numpy()
    assert np.array_equal(got, _want(o, surfaces, rows))
    assert b.render_rgb(rows=[]).shape == (0, 96, 96, 3)
    with pytest.raises(IndexError):
        b.render_rgb(rows=[200])
    with pytest.raises(ValueError):
        b.set_atlas(np.zeros((6, 16, 16, 3), np.uint8))


def test_unaligned_output_takes_byte_path():
    """An output pointer that is not 16-byte aligned must still give the same bytes (C ABI called directly)."""
    from ragen_b200 import _lib

    b, o = _stepped(40, 6, 6, 1)
    surfaces = S.synthetic_surfaces(16, seed=5)
    b.set_atlas(surfaces)
    size = 40 * 96 * 96 * 3
    buf = torch.zeros(size + 16, dtype=torch.uint8, device="cuda")
    out = buf[4:4 + size]
    _lib.check(b.lib.ragen_sokoban_render_rgb(b.ctx.handle, b.n, 6, 6, 16, _lib.ptr(b.state), _lib.ptr(b.meta), _lib.ptr(b.atlas),
                                              None, 0, _lib.ptr(out), b.ctx.stream()))
    assert np.array_equal(out.cpu().numpy().reshape(40, 96, 96, 3), _want(o, surfaces, range(40)))
    assert int(buf[:4].sum()) == 0 and int(buf[4 + size:].sum()) == 0


def test_env_and_manager_rgb_mode():
    """SokobanEnv.render('rgb_array') and the manager's image states (es_manager.py:212-228, 242-250)."""
    import PIL.Image

    from ragen_b200.env.sokoban import SokobanEnv, SokobanEnvConfig
    from ragen_b200.llm_agent.es_manager import EnvStateManager

    surfaces = S.synthetic_surfaces(16, seed=7)
    env = SokobanEnv(SokobanEnvConfig(dim_room=(6, 6), num_boxes=1, render_mode="rgb_array"))
    env.batch.set_atlas(surfaces)
    img = env.reset(seed=3)
    oracle = S.SokobanOracle((6, 6), 1)
    oracle.reset(seed=3)
    assert isinstance(img, np.ndarray) and np.array_equal(img, oracle.render_rgb(surfaces))
    obs, _, _, _ = env.step(1)
    oracle.step(1)
    assert np.array_equal(obs, oracle.render_rgb(surfaces))
    assert env.render(mode="text") == oracle.render()

    cfg = C.base_config(tags=("SimpleSokoban",), n_groups=(2,), group_size=2)
    cfg.custom_envs["SimpleSokoban"].env_config["render_mode"] = "rgb_array"
    mgr = EnvStateManager(cfg, mode="train")
    mgr.blocks[0].batch.set_atlas(surfaces)
    cache = mgr.reset(seed=5)
    oracles = []
    for c in cache:
        h = c["history"][0]
        assert h["state"] == "<images>" and len(h["images"]) == 1 and isinstance(h["images"][0], PIL.Image.Image)
        o = S.SokobanOracle((6, 6), 1)
        o.reset(seed=5 + c["group_id"])
        assert np.array_equal(np.asarray(h["images"][0]), o.render_rgb(surfaces))
        oracles.append(o)
    inputs = [{"env_id": c["env_id"], "llm_raw_response": "", "llm_response": "", "actions": ["Up", "Left"]} for c in cache]
    mgr.step(json.loads(json.dumps(inputs)))
    for c, o in zip(mgr.get_rollout_states(), oracles):
        if not o.step(1)[2]:  # the manager stops a turn at done (es_manager.py:125-126)
            o.step(3)
        assert np.array_equal(np.asarray(c["history"][-1]["images"][0]), o.render_rgb(surfaces))
