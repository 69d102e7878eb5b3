"""Synthetic INPUTS for benches, smoke runs and tests (host numpy; nothing here is on the compute
path).  Shapes follow SURVEY.md 8(d)."""
import numpy as np

WALL, FLOOR, TARGET, BOX_ON_TARGET, BOX, PLAYER = 0, 1, 2, 3, 4, 5


def synthetic_rooms(n, H, W, num_boxes, seed):
    """Synthetic (NOT reverse-played) rooms for throughput runs: a random-walk floor like
    room_topology_generation (utils.py:281-355) with player / targets / boxes dropped on distinct floor
    cells.  Border is wall.  Vectorised; used only when the device generator is not the thing measured."""
    rng = np.random.default_rng(seed)
    floor = np.zeros((n, H, W), dtype=bool)
    r = rng.integers(1, H - 1, n)
    c = rng.integers(1, W - 1, n)
    dirs = np.array([[1, 0], [0, 1], [-1, 0], [0, -1]])
    d = rng.integers(0, 4, n)
    idx = np.arange(n)
    stamps = [((0, -1), (0, 0), (0, 1)), ((-1, 0), (0, 0), (1, 0)), ((0, -1), (0, 0), (1, 0)),
              ((0, -1), (0, 0), (1, -1), (1, 0)), ((0, 0), (0, 1), (1, 0))]
    for _ in range(int(1.7 * (H + W))):
        ch = rng.random(n) < 0.35
        d = np.where(ch, rng.integers(0, 4, n), d)
        r = np.clip(r + dirs[d, 0], 1, H - 2)
        c = np.clip(c + dirs[d, 1], 1, W - 2)
        which = rng.integers(0, 5, n)
        for k, st in enumerate(stamps):
            sel = idx[which == k]
            for dr, dc in st:
                floor[sel, r[sel] + dr, c[sel] + dc] = True
    floor[:, 0, :] = floor[:, -1, :] = False
    floor[:, :, 0] = floor[:, :, -1] = False
    # a 2x3 block in the middle guarantees enough floor for player + 2*num_boxes objects
    need = 1 + 2 * num_boxes
    flat = floor.reshape(n, -1)
    short = flat.sum(1) < need
    if short.any():
        floor[short, 1:H - 1, 1:W - 1] = True
        flat = floor.reshape(n, -1)
    keys = rng.random((n, H * W))
    keys[~flat] = 2.0
    order = np.argsort(keys, axis=1)[:, :need]  # distinct random floor cells
    fixed = flat.astype(np.uint8)
    state = fixed.copy()
    rows = idx[:, None]
    tg = order[:, 1:1 + num_boxes]
    bx = order[:, 1 + num_boxes:]
    fixed[rows, tg] = TARGET
    state[rows, tg] = TARGET
    state[rows, bx] = BOX
    state[idx, order[:, 0]] = PLAYER
    return fixed, state
