"""ORACLE (test infrastructure, CPU) — Sokoban step/render restated over a BATCH with numpy.

Same semantics as oracle/sokoban.py:SokobanOracle.step (which follows ragen/env/sokoban/env.py:44-57
and gym_sokoban step/_push/_move/_calc_reward), vectorised over envs so that parity runs at 10^5..10^6
envs finish in seconds.  tests/test_oracle_golden.py checks it against the scalar oracle and the
golden fixtures.  Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may import this.
"""
import numpy as np

WALL, FLOOR, TARGET, BOX_ON_TARGET, BOX, PLAYER = 0, 1, 2, 3, 4, 5
_DR = np.array([0, -1, 1, 0, 0], dtype=np.int64)
_DC = np.array([0, 0, 0, -1, 1], dtype=np.int64)


class SokobanBatchOracle:
    def __init__(self, fixed, state, H, W, num_boxes, max_steps):
        self.H, self.W, self.num_boxes, self.max_steps = H, W, num_boxes, max_steps
        self.fixed = np.ascontiguousarray(fixed, dtype=np.uint8).reshape(-1, H * W).copy()
        self.state = np.ascontiguousarray(state, dtype=np.uint8).reshape(-1, H * W).copy()
        self.n = self.fixed.shape[0]
        self.player = np.argmax(self.state == PLAYER, axis=1).astype(np.int64)
        self.env_steps = np.zeros(self.n, dtype=np.int64)
        self.on_target = np.zeros(self.n, dtype=np.int64)

    def _open_targets(self, rows):
        s, f = self.state[rows], self.fixed[rows]
        return np.count_nonzero((s == TARGET) | ((f == TARGET) & (s == PLAYER)), axis=1)

    def step_one(self, actions, active=None):
        """One action per env.  actions int [n] (0 = no-op step, 1..4 moves); active bool [n] or None.
        Returns (reward f64 [n], done bool [n], effective bool [n], success bool [n]) - entries of
        inactive envs are zero/False."""
        n, H, W = self.n, self.H, self.W
        a = np.asarray(actions, dtype=np.int64)
        act = np.ones(n, dtype=bool) if active is None else np.asarray(active, dtype=bool)
        idx = np.arange(n)
        s, f = self.state, self.fixed
        pos = self.player
        self.env_steps[act] += 1
        mv = act & (a >= 1) & (a <= 4)
        aa = np.where(mv, a, 0)
        r, c = pos // W, pos % W
        ar, ac = r + _DR[aa], c + _DC[aa]
        br, bc = ar + _DR[aa], ac + _DC[aa]
        in_a = (ar >= 0) & (ar < H) & (ac >= 0) & (ac < W)
        ok = mv & in_a & (br < H) & (bc < W)  # upstream _push gives up when the cell beyond leaves the grid
        ahead = np.where(in_a, ar * W + ac, 0)
        in_b = (br >= 0) & (bc >= 0) & (br < H) & (bc < W)
        beyond = np.where(in_b, br * W + bc, 0)
        sa = s[idx, ahead]
        sb = s[idx, beyond]
        push = ok & in_b & ((sa == BOX) | (sa == BOX_ON_TARGET)) & ((sb == FLOOR) | (sb == TARGET))
        move = ok & ~push & ((sa == FLOOR) | (sa == TARGET))
        go = push | move
        pi = idx[push]
        s[pi, beyond[push]] = np.where(f[pi, beyond[push]] == TARGET, BOX_ON_TARGET, BOX)
        gi = idx[go]
        s[gi, ahead[go]] = PLAYER
        s[gi, pos[go]] = f[gi, pos[go]]
        new_pos = np.where(go, ahead, pos)
        self.player = new_pos
        # reward
        now_on = self.on_target.copy()
        ai = idx[act]
        now_on[ai] = self.num_boxes - self._open_targets(ai)
        rew = np.where(act, -0.1, 0.0)
        rew = np.where(act & (now_on > self.on_target), rew + 1, rew)
        rew = np.where(act & (now_on < self.on_target), rew + -1, rew)
        won = act & (now_on == self.num_boxes)
        rew = np.where(won, rew + 10, rew)
        self.on_target = now_on
        done = won | (act & (self.env_steps == self.max_steps))
        return rew, done, act & go, won

    def step_turn(self, actions, n_actions=None):
        """Up to K actions per env, stopping at the first done (es_manager.py:116-128).
        Returns acc_reward f64, done u8, flags u8 (bit0 effective, bit1 success, bit2 executed), n_executed u8."""
        actions = np.asarray(actions)
        if actions.ndim == 1:
            actions = actions[:, None]
        n, K = actions.shape
        na = np.full(n, K, dtype=np.int64) if n_actions is None else np.minimum(np.asarray(n_actions, dtype=np.int64), K)
        acc = np.zeros(n, dtype=np.float64)
        done = np.zeros(n, dtype=bool)
        flags = np.zeros(n, dtype=np.uint8)
        nexec = np.zeros(n, dtype=np.uint8)
        for k in range(K):
            active = (~done) & (k < na)
            if not active.any():
                break
            rew, d, eff, won = self.step_one(actions[:, k], active)
            acc = np.where(active, acc + rew, acc)  # same left-to-right double sum as the Python loop
            done |= d
            flags = np.where(active, eff.astype(np.uint8) | (won.astype(np.uint8) << 1) | 4, flags).astype(np.uint8)
            nexec += active.astype(np.uint8)
        return acc, done.astype(np.uint8), flags, nexec

    def glyphs(self):
        """uint8 [n, H*(W+1)]: codes 0..6 (6 = player on target), 7 = row separator (incl. last byte)."""
        g = np.where((self.state == PLAYER) & (self.fixed == TARGET), 6, self.state).reshape(self.n, self.H, self.W)
        out = np.full((self.n, self.H, self.W + 1), 7, dtype=np.uint8)
        out[:, :, : self.W] = g
        return out.reshape(self.n, -1)


from ragen_b200.synth import synthetic_rooms  # noqa: E402,F401  (input synthesis lives with the product)
