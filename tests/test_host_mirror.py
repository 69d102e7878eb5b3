"""Host side of the step records (CPU): decode_delta and HostMirror on one-byte and 4-byte records built from the oracle's own steps.
The GPU tests check that the kernels write these records; here the host logic that consumes them is checked without a GPU."""
import numpy as np
import pytest

from oracle.sokoban_np import SokobanBatchOracle, synthetic_rooms
from ragen_b200.env.sokoban import REWARD_CODES, HostMirror, decode_delta


def records_of_turn(o, acts):
    """Steps the oracle through one turn and writes the records ragen_sokoban_step_delta / _delta1 would: ([n, K, 4], [n, K])."""
    n, K = acts.shape
    d4 = np.zeros((n, K, 4), dtype=np.uint8)
    d1 = np.zeros((n, K), dtype=np.uint8)
    done = np.zeros(n, dtype=bool)
    for k in range(K):
        active = ~done
        before = o.player.copy()
        boxes = (o.state == 3) | (o.state == 4)
        rew, d, eff, won = o.step_one(acts[:, k], active)
        moved_box = ((o.state == 3) | (o.state == 4)) & ~boxes
        pushed = moved_box.any(axis=1)
        box_to = np.where(pushed, moved_box.argmax(axis=1), 0xff)
        code = np.array([REWARD_CODES.index(r) if a else 0 for r, a in zip(np.round(rew, 12).tolist(), active)])
        info = (code | (d.astype(int) << 3) | (eff.astype(int) << 4) | (won.astype(int) << 5) | (1 << 6)) * active
        d4[:, k, 0], d4[:, k, 1], d4[:, k, 2], d4[:, k, 3] = before * active, o.player * active, box_to * active, info
        d1[:, k] = info | ((pushed & active).astype(int) << 7)
        done |= d
    return d4, d1


@pytest.mark.parametrize("H,W,nb,n,K", [(10, 10, 2, 700, 1), (6, 6, 1, 300, 3), (8, 8, 3, 129, 5)])
def test_mirrors_follow_the_oracle_from_both_record_widths(H, W, nb, n, K):
    assert [round(x, 12) for x in REWARD_CODES] == [-0.1, 0.9, -1.1, 10.9, 9.9]
    fixed, state = synthetic_rooms(n, H, W, nb, seed=n)
    rng = np.random.default_rng(n + 1)
    o = SokobanBatchOracle(fixed, state, H, W, nb, 25)
    twin = SokobanBatchOracle(fixed, state, H, W, nb, 25)
    m4, m1 = HostMirror(fixed, state, H, W), HostMirror(fixed, state, H, W)
    for turn in range(40):
        acts = rng.integers(0, 5, (n, K)).astype(np.uint8)
        d4, d1 = records_of_turn(o, acts)
        want = twin.step_turn(acts)
        m4.apply(d4)
        m1.apply(d1, acts)
        for got in (decode_delta(d4), decode_delta(d1)):
            for g, w in zip(got, want):
                assert np.array_equal(g, w), turn
        assert np.array_equal(m4.state, o.state) and np.array_equal(m1.state, o.state), turn
        assert np.array_equal(m1.glyphs(), o.glyphs()) and np.array_equal(m1.players(), o.player)
    # rooms that were reset: the mirror takes their new rows and finds the players again
    fx2, st2 = synthetic_rooms(2, H, W, nb, seed=99)
    m1.refresh([0, n - 1], fx2.reshape(2, -1), st2.reshape(2, -1))
    assert np.array_equal(m1.players()[[0, n - 1]], np.argmax(st2.reshape(2, -1) == 5, axis=1))
