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


_WORDS = ("the box is near wall so i move it push left right up down target then maybe not because plan step think "
          "first second go around corner player reach goal arm dragon phoenix reward higher likely name symbol").split()


def _filler(rng, lo, hi):
    n = int(rng.integers(lo, hi))
    out, size = [], 0
    while size < n:
        w = _WORDS[int(rng.integers(0, len(_WORDS)))]
        out.append(w)
        size += len(w) + 1
    return " ".join(out)


def synthetic_responses(n_bandit=8192, n_countdown=8192, seed=3, lo_name="Phoenix", hi_name="Dragon"):
    """The C3 corpus of SURVEY.md 8(d): LLM-style responses for Bandit and Countdown.

    Returns dict(responses=list[str], kind=np.uint8 [n] (0 bandit, 1 countdown), nums=int32 [n, 8], n_nums=int32 [n],
    target=int64 [n]).  Shapes: 70 % well-formed, 10 % missing </answer>, 10 % nested / duplicated tags, 10 % no tags."""
    rng = np.random.default_rng(seed)
    n = n_bandit + n_countdown
    responses, kind = [], np.zeros(n, np.uint8)
    nums = np.zeros((n, 8), np.int32)
    n_nums = np.zeros(n, np.int32)
    target = np.zeros(n, np.int64)
    for i in range(n):
        is_cd = i >= n_bandit
        kind[i] = int(is_cd)
        if is_cd:
            k = int(rng.integers(3, 5))
            v = rng.integers(1, 100, k)
            signs = rng.choice([1, -1], k)
            nums[i, :k] = v
            n_nums[i] = k
            target[i] = int((v * signs).sum())
            u = rng.random()
            expr = " ".join(("+" if s > 0 else "-") + " " + str(x) for x, s in zip(v, signs)).lstrip("+ ")
            if u < 0.5:
                a = expr
            elif u < 0.75:
                a = " * ".join(str(x) for x in v)
            elif u < 0.9:
                a = expr + " + 1"
            else:
                a = expr + " = " + str(int(target[i]))
        else:
            u = rng.random()
            name = lo_name if u < 0.45 else hi_name if u < 0.9 else "Griffin"
            a = "".join(ch.upper() if rng.random() < 0.5 else ch.lower() for ch in name)
            if rng.random() < 0.3:
                a = " " + a + " "
        w = _filler(rng, 200, 1600)
        shape = rng.random()
        if shape < 0.7:
            text = f"{w}</think><answer>{a}</answer>"
        elif shape < 0.8:
            text = f"{w}</think><answer>{a}"
        elif shape < 0.9:
            text = f"<think>{w}</answer> 123. </think><answer> <answer> {a} </answer></answer>"
        else:
            text = f"{w} {a}"
        responses.append("<think>" + text)  # get_env_inputs puts the opening tag back (ctx_manager.py:340)
    return dict(responses=responses, kind=kind, nums=nums, n_nums=n_nums, target=target)
