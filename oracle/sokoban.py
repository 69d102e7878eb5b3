"""ORACLE (test infrastructure, CPU) — Sokoban reset/step/render restated.

This file is the checker the CUDA path is compared against.  Only `tests/`,
`__graft_entry__.smoke()` and `bench.py`'s cpu_baseline / `--impl reference` legs may import it.
It travels to the GPU box (where `/root/reference` does not exist); in this container it is
itself validated against the unmodified reference (tests/test_oracle_vs_reference.py) and
pinned by the fixtures under tests/golden/.

Follows
  * ragen/env/sokoban/env.py:28-57        reset / step / text render
  * ragen/env/sokoban/utils.py:221-278    generate_room
  * ragen/env/sokoban/utils.py:281-355    room_topology_generation
  * ragen/env/sokoban/utils.py:358-397    place_boxes_and_player
  * ragen/env/sokoban/utils.py:408-560    reverse_playing / depth_first_search / reverse_move /
                                          box_displacement_score
  * ragen/env/sokoban/utils.py:152-213    add_random_player_movement
  * ragen/utils.py:7-18                   all_seed (global RNG state saved and restored)
  * gym_sokoban SokobanEnv.step/_push/_move/_calc_reward (un-vendored third party, un-pinned;
    restated from the published upstream algorithm - "parity unpinned" at that boundary).

Representation differs from the reference on purpose (flat cell indices, box lists instead of
int64 boards) - the results are identical; the random streams are consumed through the very same
CPython `random` / legacy `numpy.random` calls in the same order, so rooms are stream-exact.
"""
import random

import numpy as np

WALL, FLOOR, TARGET, BOX_ON_TARGET, BOX, PLAYER = 0, 1, 2, 3, 4, 5
PLAYER_ON_TARGET = 6  # render-only glyph index (env.py:55)
DEFAULT_GLYPHS = {0: "#", 1: "_", 2: "O", 3: "√", 4: "X", 5: "P", 6: "S"}  # sokoban/config.py:10
DELTAS = ((-1, 0), (1, 0), (0, -1), (0, 1))  # up, down, left, right (utils.py:588-593)

# the five 3x3 stamps of the topology walk, as (dr, dc) offsets from the walk position (utils.py:295-321)
_STAMPS = (
    ((0, -1), (0, 0), (0, 1)),
    ((-1, 0), (0, 0), (1, 0)),
    ((0, -1), (0, 0), (1, 0)),
    ((0, -1), (0, 0), (1, -1), (1, 0)),
    ((0, 0), (0, 1), (1, 0)),
)
_WALK_DIRS = ((1, 0), (0, 1), (-1, 0), (0, -1))  # utils.py:324
MAX_EXPLORED = 300000  # utils.py:457


class RoomGenerationError(Exception):
    """Stands for the RuntimeError (placement) / RuntimeWarning (score 0) of the reference."""


# --------------------------------------------------------------------------------------------
# generator
# --------------------------------------------------------------------------------------------
def _topology(H, W, num_steps, p_change=0.35):
    direction = random.sample(_WALK_DIRS, 1)[0]
    r = random.randint(1, H - 1)
    c = random.randint(1, W - 1)
    floor = [[0] * W for _ in range(H)]
    for _ in range(num_steps):
        if random.random() < p_change:
            direction = random.sample(_WALK_DIRS, 1)[0]
        r = max(min(r + direction[0], H - 2), 1)
        c = max(min(c + direction[1], W - 2), 1)
        for dr, dc in random.sample(_STAMPS, 1)[0]:
            floor[r + dr][c + dc] = 1
    for i in range(H):
        floor[i][0] = floor[i][W - 1] = 0
    for j in range(W):
        floor[0][j] = floor[H - 1][j] = 0
    return floor


def _place(floor, H, W, num_boxes):
    """Returns (player_cell, [target cells]) drawn with legacy np.random.randint."""
    free = [r * W + c for r in range(H) for c in range(W) if floor[r][c] == 1]
    if len(free) <= num_boxes + 1:
        raise RoomGenerationError("not enough free spots")
    player = free[np.random.randint(len(free))]
    taken = {player}
    targets = []
    for _ in range(num_boxes):
        free = [x for x in free if x not in taken]
        t = free[np.random.randint(len(free))]
        targets.append(t)
        taken.add(t)
    return player, targets


def _reverse_play(struct, H, W, player0, targets, search_depth):
    """Pull-only depth-first search from the solved position (utils.py:408-542).

    `targets` must be in row-major order (the key order of the reference's box_mapping).
    Returns (best_player, best_boxes, displacement) where best_boxes[k] is the cell of the box that
    started on targets[k].
    """
    nb = len(targets)
    trow = [t // W for t in targets]
    tcol = [t % W for t in targets]
    step = [d[0] * W + d[1] for d in DELTAS]

    explored = set()
    best_score = -1
    best = None
    # frame: [player, boxes(tuple), pulled(tuple of bool), last_pull, swaps, ttl, next_action]
    stack = [[player0, tuple(targets), (False,) * nb, -1, 0, search_depth, -1]]
    while stack:
        fr = stack[-1]
        p, boxes, pulled, last_pull, swaps, ttl, nxt = fr
        if nxt < 0:  # function entry
            ttl -= 1
            fr[5] = ttl
            if ttl <= 0 or len(explored) >= MAX_EXPLORED:
                stack.pop()
                continue
            key = (p, tuple(sorted(zip(boxes, pulled))))
            if key in explored:
                stack.pop()
                continue
            disp = 0
            for k in range(nb):
                disp += abs(boxes[k] // W - trow[k]) + abs(boxes[k] % W - tcol[k])
            score = swaps * disp
            open_targets = sum(1 for t in targets if t != p and t not in boxes)
            if open_targets != nb:
                score = 0
            if score > best_score:
                best_score = score
                best = (p, boxes, disp)
            explored.add(key)
            fr[6] = 0
            nxt = 0
        if nxt >= 4:
            stack.pop()
            continue
        fr[6] = nxt + 1
        a = nxt
        n = p + step[a]
        cp, cb, cpl, clp = p, boxes, pulled, last_pull
        if struct[n] != WALL and n not in boxes:  # board[n] in (1, 2)
            cp = n
            behind = p - step[a]
            if behind in boxes:
                k = boxes.index(behind)
                cb = boxes[:k] + (p,) + boxes[k + 1:]
                cpl = pulled[:k] + (True,) + pulled[k + 1:]
                clp = k
        stack.append([cp, cb, cpl, clp, swaps + (1 if clp != last_pull else 0), ttl, -1])
    return best


def _jitter(state_cells, struct, H, W, player, move_probability, continue_probability=0.5, max_steps=3):
    """add_random_player_movement (utils.py:152-213) on a flat board; returns the new player cell."""
    if random.random() > move_probability:
        return player
    visited = [player]
    steps = 0
    while steps < max_steps:
        options = []
        for d in DELTAS:
            n = player + d[0] * W + d[1]
            if state_cells[n] in (FLOOR, TARGET) and n not in visited:
                options.append(n)
        if not options:
            break
        n = random.choice(options)
        state_cells[player] = struct[player]
        state_cells[n] = PLAYER
        player = n
        visited.append(n)
        steps += 1
        if steps >= max_steps or random.random() > continue_probability:
            break
    return player


def generate_room_seeded(H, W, num_boxes, search_depth, num_gen_steps, tries=4):
    """generate_room (utils.py:221-278) using the CURRENT global RNG states.

    Returns (room_fixed uint8[H*W], room_state uint8[H*W]).  Raises RoomGenerationError where the
    reference raises RuntimeError / RuntimeWarning.
    """
    for _ in range(tries):
        floor = _topology(H, W, num_gen_steps)
        player, targets = _place(floor, H, W, num_boxes)
        struct = [floor[r][c] for r in range(H) for c in range(W)]
        for t in targets:
            struct[t] = TARGET
        targets = sorted(targets)
        best_player, best_boxes, disp = _reverse_play(struct, H, W, player, targets, search_depth)
        if disp > 0:
            break
    if disp == 0:
        raise RoomGenerationError("generated model with score == 0")
    state = list(struct)
    for b in best_boxes:
        state[b] = BOX
    state[best_player] = PLAYER
    _jitter(state, struct, H, W, best_player, 0.8 if disp == 1 else 0.5)
    return np.array(struct, dtype=np.uint8), np.array(state, dtype=np.uint8)


def generate_room(H, W, num_boxes, search_depth, seed, num_gen_steps=None):
    """One attempt under all_seed(seed) (ragen/utils.py:7-18): global RNG state is restored."""
    if num_gen_steps is None:
        num_gen_steps = int(1.7 * (H + W))  # gym_sokoban ctor default
    py_state = random.getstate()
    np_state = np.random.get_state()
    try:
        random.seed(seed)
        np.random.seed(seed)
        return generate_room_seeded(H, W, num_boxes, search_depth, num_gen_steps)
    finally:
        random.setstate(py_state)
        np.random.set_state(np_state)


def retry_seed(seed):
    """Seed used after a failed attempt (sokoban/env.py:41).  Process-dependent unless
    PYTHONHASHSEED is pinned - same caveat as the reference."""
    return abs(hash(str(seed))) % (2 ** 32)


# --------------------------------------------------------------------------------------------
# env
# --------------------------------------------------------------------------------------------
class SokobanOracle:
    """One environment; mirrors SokobanEnv's reset/step/render contract (sokoban/env.py)."""

    def __init__(self, dim_room=(6, 6), num_boxes=1, max_steps=100, search_depth=300, glyphs=None):
        self.H, self.W = dim_room
        self.num_boxes = num_boxes
        self.max_steps = max_steps
        self.search_depth = search_depth
        self.glyphs = dict(DEFAULT_GLYPHS if glyphs is None else glyphs)
        self.fixed = None
        self.state = None

    # reset -------------------------------------------------------------------------------
    def load(self, fixed, state):
        """Adopt a given room (used to start parity runs from fixture rooms)."""
        self.fixed = np.asarray(fixed, dtype=np.uint8).reshape(-1).copy()
        self.state = np.asarray(state, dtype=np.uint8).reshape(-1).copy()
        self.player = int(np.flatnonzero(self.state == PLAYER)[0])
        f2 = self.fixed.reshape(self.H, self.W)
        if f2[0].any() or f2[-1].any() or f2[:, 0].any() or f2[:, -1].any():
            raise ValueError("rooms must have a wall border (all generated rooms do, utils.py:352-353)")
        self.num_env_steps = 0
        self.boxes_on_target = 0
        self.reward_last = 0
        return self.render()

    def reset(self, seed=None):
        while True:
            try:
                fixed, state = generate_room(self.H, self.W, self.num_boxes, self.search_depth, seed)
                break
            except RoomGenerationError:
                seed = retry_seed(seed) if seed is not None else None
        self.seed_used = seed
        return self.load(fixed, state)

    # step --------------------------------------------------------------------------------
    def _open_targets(self):
        s, f = self.state, self.fixed
        return int(np.count_nonzero((s == TARGET) | ((f == TARGET) & (s == PLAYER))))

    def step(self, action):
        """action 1..4 = up/down/left/right push-or-move; 0 = no-op (still costs a step)."""
        W = self.W
        before = self.player
        self.num_env_steps += 1
        if 1 <= action <= 4:
            dr, dc = DELTAS[action - 1]
            r, c = divmod(self.player, W)
            ar, ac = r + dr, c + dc
            br, bc = ar + dr, ac + dc
            s, f = self.state, self.fixed
            ahead = ar * W + ac  # always in-grid: rooms have a wall border (checked in load)
            moved = False
            if br >= self.H or bc >= W:
                pass  # upstream _push returns before trying a plain move
            else:
                pushable = br >= 0 and bc >= 0 and s[ahead] in (BOX_ON_TARGET, BOX) \
                    and s[br * W + bc] in (FLOOR, TARGET)
                if pushable:
                    beyond = br * W + bc
                    s[beyond] = BOX_ON_TARGET if f[beyond] == TARGET else BOX
                    moved = True
                elif s[ahead] in (FLOOR, TARGET):
                    moved = True
            if moved:
                s[ahead] = PLAYER
                s[self.player] = f[self.player]
                self.player = ahead
        reward = -0.1
        now_on = self.num_boxes - self._open_targets()
        if now_on > self.boxes_on_target:
            reward += 1
        elif now_on < self.boxes_on_target:
            reward += -1
        won = now_on == self.num_boxes
        if won:
            reward += 10
        self.boxes_on_target = now_on
        self.reward_last = reward
        done = won or (self.max_steps == self.num_env_steps)
        info = {
            "action_is_effective": before != self.player,
            "action_is_valid": True,
            "success": self.boxes_on_target == self.num_boxes,
        }
        return self.render(), reward, done, info

    # render ------------------------------------------------------------------------------
    def glyph_codes(self):
        """uint8[H, W] glyph indices 0..6 (6 = player standing on a target)."""
        g = np.where((self.state == PLAYER) & (self.fixed == TARGET), PLAYER_ON_TARGET, self.state)
        return g.reshape(self.H, self.W).astype(np.uint8)

    def render(self):
        lut = self.glyphs
        return "\n".join("".join(lut.get(int(c), "?") for c in row) for row in self.glyph_codes())

    def render_rgb(self, surfaces):
        return room_to_rgb(self.state.reshape(self.H, self.W), self.fixed.reshape(self.H, self.W), surfaces)


# --------------------------------------------------------------------------------------------
# rgb_array observation (env.py:58-59 -> gym_sokoban get_image(mode='rgb_array') -> room_to_rgb)
# --------------------------------------------------------------------------------------------
# gym_sokoban is un-vendored and absent here, and so are its sprite PNGs: the blit below restates the published
# room_to_rgb loop (parity unpinned), and the sprites are an argument.
def room_to_rgb(room_state, room_fixed, surfaces):
    """surfaces: uint8 [7, S, S, 3] in upstream's order (wall, floor, box_target, box_on_target, box, player,
    player_on_target).  Returns uint8 [H*S, W*S, 3]."""
    room = np.array(room_state).astype(np.int64)
    if room_fixed is not None:
        room[(room == PLAYER) & (np.asarray(room_fixed) == TARGET)] = PLAYER_ON_TARGET
    surfaces = np.asarray(surfaces, dtype=np.uint8)
    S = surfaces.shape[1]
    H, W = room.shape
    img = np.zeros((H * S, W * S, 3), dtype=np.uint8)
    for i in range(H):
        for j in range(W):
            img[i * S:(i + 1) * S, j * S:(j + 1) * S, :] = surfaces[room[i, j]]
    return img


def synthetic_surfaces(S=16, seed=0):
    """Random sprites for parity runs: every pixel of every surface differs, so a misplaced byte shows."""
    return np.random.default_rng(seed).integers(0, 256, size=(7, S, S, 3), dtype=np.uint8)


# --------------------------------------------------------------------------------------------
# validity checks for generated rooms
# --------------------------------------------------------------------------------------------
def solve_bfs(fixed, state, H, W, max_depth=100):
    """Shortest push solution, same contract as get_shortest_action_path (utils.py:11-101):
    list of actions 1..4, or [] if none within max_depth.  Compact-state BFS."""
    fixed = np.asarray(fixed).reshape(-1)
    state = np.asarray(state).reshape(-1)
    targets = frozenset(int(x) for x in np.flatnonzero(fixed == TARGET))
    p0 = int(np.flatnonzero(state == PLAYER)[0])
    b0 = frozenset(int(x) for x in np.flatnonzero((state == BOX) | (state == BOX_ON_TARGET)))
    step = [d[0] * W + d[1] for d in DELTAS]
    from collections import deque

    start = (p0, b0)
    seen = {start}
    q = deque([(start, [])])
    while q:
        (p, boxes), path = q.popleft()
        if len(path) > max_depth:
            return []
        # the reference tests `state == 4` (boxes NOT on target); a box counts as solved only when marked 3
        if all(b in targets for b in boxes):
            return path
        for a in range(4):
            n = p + step[a]
            r, c = divmod(p, W)
            nr, nc = r + DELTAS[a][0], c + DELTAS[a][1]
            if not (0 <= nr < H and 0 <= nc < W) or fixed[n] == WALL:
                continue
            nb = boxes
            if n in boxes:
                br, bc = nr + DELTAS[a][0], nc + DELTAS[a][1]
                if not (0 <= br < H and 0 <= bc < W):
                    continue
                m = n + step[a]
                if fixed[m] == WALL or m in boxes:
                    continue
                nb = (boxes - {n}) | {m}
            nxt = (n, nb)
            if nxt in seen:
                continue
            seen.add(nxt)
            q.append((nxt, path + [a + 1]))
    return []


def check_room(fixed, state, H, W, num_boxes):
    """Structural invariants every reference-generated room satisfies."""
    fixed = np.asarray(fixed).reshape(H, W)
    state = np.asarray(state).reshape(H, W)
    assert set(np.unique(fixed)) <= {0, 1, 2}
    assert (fixed[0] == 0).all() and (fixed[-1] == 0).all() and (fixed[:, 0] == 0).all() and (fixed[:, -1] == 0).all()
    assert (fixed == TARGET).sum() == num_boxes
    assert (state == PLAYER).sum() == 1
    assert ((state == BOX) | (state == BOX_ON_TARGET)).sum() == num_boxes
    assert (state == BOX_ON_TARGET).sum() == 0
    # every target is uncovered by boxes (score > 0 rule); the player jitter may park the player on one
    assert ((state == TARGET) | ((state == PLAYER) & (fixed == TARGET))).sum() == num_boxes
    movable = (state >= 3)
    assert ((fixed != WALL) | ~movable).all()
    assert (state[~movable] == fixed[~movable]).all()
    return True
