"""ORACLE (test infrastructure, CPU) — FrozenLake, Bandit and Countdown restated.

Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s CPU legs may import this.  Validated in
this container against the unmodified reference (tests/test_oracle_vs_reference.py) and pinned
by tests/golden/.

Follows
  * ragen/env/frozen_lake/env.py:10-57, utils.py:6-47, config.py:4-22
  * gymnasium FrozenLakeEnv.reset/step + utils.seeding.np_random (un-vendored third party,
    un-pinned; restated from the published algorithm - "parity unpinned" at that boundary)
  * ragen/env/bandit/env.py:14-85, config.py:4-14
  * ragen/env/countdown/env.py:9-21, 69-78
"""
import re

import numpy as np

# ---------------------------------------------------------------------------------------------
# FrozenLake
# ---------------------------------------------------------------------------------------------
FL_LEFT, FL_DOWN, FL_RIGHT, FL_UP = 0, 1, 2, 3
FL_GLYPHS = {0: "P", 1: "_", 2: "O", 3: "G", 4: "X", 5: "√"}  # frozen_lake/config.py:20
# tile codes used by the flat uint8 layout shared with the CUDA path
FL_F, FL_H, FL_G, FL_S = 1, 2, 3, 0  # 'S' is stored as 0 so that glyph(F)=1, glyph(H)=2, glyph(G)=3


def _rng(seed):
    return np.random.Generator(np.random.PCG64(np.random.SeedSequence(int(seed))))


def _reaches_goal(board, n):
    """DFS reachability from S to G avoiding H (frozen_lake/utils.py:6-22)."""
    start = [(r, c) for r in range(n) for c in range(n) if board[r][c] == "S"][0]
    todo, seen = [start], set()
    while todo:
        r, c = todo.pop()
        if (r, c) in seen:
            continue
        seen.add((r, c))
        for dr, dc in ((1, 0), (0, 1), (-1, 0), (0, -1)):
            rr, cc = r + dr, c + dc
            if 0 <= rr < n and 0 <= cc < n:
                if board[rr][cc] == "G":
                    return True
                if board[rr][cc] != "H":
                    todo.append((rr, cc))
    return False


def frozenlake_map(size, p, seed):
    """generate_random_map (frozen_lake/utils.py:25-47): list of row strings over S,F,H,G."""
    rng = _rng(seed)
    while True:
        board = rng.choice(["F", "H"], (size, size), p=[p, 1 - p])
        sr, sc = rng.integers(size, size=2)
        gr, gc = rng.integers(size, size=2)
        if (sr, sc) != (gr, gc):
            board[sr][sc], board[gr][gc] = "S", "G"
            if _reaches_goal(board, size):
                return ["".join(row) for row in board]


class FrozenLakeOracle:
    def __init__(self, size=4, p=0.8, is_slippery=True, glyphs=None):
        self.size, self.p, self.is_slippery = size, p, is_slippery
        self.glyphs = dict(FL_GLYPHS if glyphs is None else glyphs)

    def reset(self, seed):
        """FrozenLakeEnv.reset (env.py:28-37): map from seed, then env RNG re-seeded with the same
        seed and ONE uniform drawn for the (deterministic) start-state sample."""
        self.desc = [list(row) for row in frozenlake_map(self.size, self.p, seed)]
        self.rng = _rng(seed)
        n = self.size
        flat = [ch for row in self.desc for ch in row]
        probs = np.array([1.0 if ch == "S" else 0.0 for ch in flat])
        probs /= probs.sum()
        self.s = int(np.argmax(np.cumsum(probs) > self.rng.random()))
        return self.render()

    def _shift(self, r, c, a):
        n = self.size
        if a == FL_LEFT:
            c = max(c - 1, 0)
        elif a == FL_DOWN:
            r = min(r + 1, n - 1)
        elif a == FL_RIGHT:
            c = min(c + 1, n - 1)
        elif a == FL_UP:
            r = max(r - 1, 0)
        return r, c

    def step(self, action):
        """action 1..4 = Left, Down, Right, Up (config.py:15, 22)."""
        a = action - 1
        n = self.size
        prev = self.s
        r, c = divmod(self.s, n)
        u = self.rng.random()  # one draw per step, also from terminal cells
        if self.desc[r][c] in "GH":
            reward, done = 0, True
        else:
            if self.is_slippery:
                cs = np.cumsum([1.0 / 3.0] * 3)
                b = ((a - 1) % 4, a, (a + 1) % 4)[int(np.argmax(cs > u))]
            else:
                b = a
            r, c = self._shift(r, c, b)
            self.s = r * n + c
            reward = float(self.desc[r][c] == "G")
            done = self.desc[r][c] in "GH"
        info = {"action_is_effective": prev != self.s, "action_is_valid": True,
                "success": self.desc[r][c] == "G"}
        return self.render(), reward, done, info

    def glyph_codes(self):
        n = self.size
        lut = {"S": 1, "F": 1, "H": 2, "G": 3}
        g = np.array([[lut[ch] for ch in row] for row in self.desc], dtype=np.uint8)
        r, c = divmod(self.s, n)
        g[r, c] = 4 if self.desc[r][c] == "H" else 5 if self.desc[r][c] == "G" else 0
        return g

    def render(self):
        return "\n".join("".join(self.glyphs.get(int(x), "?") for x in row) for row in self.glyph_codes())

    def tiles(self):
        """uint8[size*size] in the CUDA layout (S=0, F=1, H=2, G=3)."""
        lut = {"S": FL_S, "F": FL_F, "H": FL_H, "G": FL_G}
        return np.array([lut[ch] for row in self.desc for ch in row], dtype=np.uint8)


# ---------------------------------------------------------------------------------------------
# Bandit
# ---------------------------------------------------------------------------------------------
BANDIT_PROMPT = (
    "You are playing a bandit game. Goal: Maximize your total reward by choosing which arm to pull. \n"
    "Game Rules: \n"
    "1. There are 2 arms, named {name_a} and {name_b}\n"
    "2. Each arm has its own reward distribution, related to their names. \n"
    "3. Analyze the symbolic meaning of each arm's name to guess how their reward distribution might behave.\n"
    "4. Based on the symbolic meaning of their names, which arm do you think is more likely to give higher "
    "rewards on average? Choose between {name_a} and {name_b}, and output like <answer> {name_a} </answer> "
    "or <answer> {name_b} </answer>.\n"
)


class BanditOracle:
    def __init__(self, lo_arm_name="phoenix", hi_arm_name="dragon", action_space_start=1, lo_arm_score=0.1,
                 hi_arm_loscore=0.0, hi_arm_hiscore=1.0, hi_arm_hiscore_prob=0.25):
        self.lo, self.hi, self.start = lo_arm_name, hi_arm_name, action_space_start
        self.lo_score, self.hi_lo, self.hi_hi, self.hi_p = lo_arm_score, hi_arm_loscore, hi_arm_hiscore, hi_arm_hiscore_prob

    def reset(self, seed):
        self.rng = _rng(seed)
        lo_first = self.rng.random() < 0.5  # bandit/env.py:27
        self.action_lookup = {self.start: self.lo if lo_first else self.hi,
                              self.start + 1: self.hi if lo_first else self.lo}
        self.lo_first = bool(lo_first)
        self.obs = BANDIT_PROMPT.format(name_a=self.action_lookup[self.start], name_b=self.action_lookup[self.start + 1])
        return self.obs

    def step(self, action):
        assert action in self.action_lookup
        arm = self.action_lookup[action]
        if arm == self.lo:
            reward = self.lo_score
        else:
            reward = self.hi_hi if self.rng.random() < self.hi_p else self.hi_lo
        self.obs = f"{arm}: {reward} points"
        info = {"action_is_effective": True, "action_is_valid": True, "success": arm == self.hi}
        return self.obs, reward, True, info

    def render(self):
        return self.obs


# ---------------------------------------------------------------------------------------------
# Countdown scoring
# ---------------------------------------------------------------------------------------------
def countdown_reward(answer: str, nums, target, score=1, format_score=0.1):
    """compute_reward (countdown/env.py:69-78) with check_format / check_correctness (:9-21)."""
    try:
        used = sorted(int(x) for x in re.findall(r"\d+", answer))
        fmt_ok = used == sorted(nums)
    except Exception:
        fmt_ok = False
    if not fmt_ok:
        return 0
    try:
        value = eval(answer, {"__builtins__": None}, {})
        ok = abs(value - target) < 1e-5
    except Exception:
        ok = False
    return score if ok else format_score


# ---------------------------------------------------------------------------------------------
# text-env answer scoring
# ---------------------------------------------------------------------------------------------
import string as _string


def normalize_text(text):
    """ragen/env/static/utils.py:7-12."""
    text = re.sub(r"\s+", "", text.lower())
    return text.translate(str.maketrans("", "", _string.punctuation))


def score_metamath(user_answer, correct_answer):
    """MetaMathQAEnv._check_answer (ragen/env/metamathqa/env.py:55-66) -> (is_correct, is_valid)."""
    norm = re.sub(r"\s+", "", user_answer.strip().lower())
    ok = bool(correct_answer) and norm == re.sub(r"\s+", "", correct_answer.lower())
    return ok, norm != ""


def score_exact(pred, label):
    """compute_score_exact_match (static/utils.py:69-82)."""
    return normalize_text(pred) == normalize_text(label), len(normalize_text(pred)) > 0


def score_numeric(pred, label):
    """compute_score_numeric (static/utils.py:84-111)."""
    pm, lm = re.search(r"(\d+(?:\.\d+)?)", pred), re.search(r"(\d+(?:\.\d+)?)", label)
    ok = bool(pm and lm and float(pm.group(0)) == float(lm.group(0)))
    return ok or normalize_text(pred) == normalize_text(label), pm is not None


def score_choice(pred, label):
    """compute_score_multiple_choice (static/utils.py:113-132)."""
    pm, lm = re.search(r"([A-D])", pred.upper()), re.search(r"([A-D])", label.upper())
    ok = pm.group(0) == lm.group(0) if (pm and lm) else normalize_text(pred) == normalize_text(label)
    return ok, pm is not None


SCORERS = {"metamathqa": score_metamath, "exact_match": score_exact, "numeric": score_numeric, "multiple_choice": score_choice}
