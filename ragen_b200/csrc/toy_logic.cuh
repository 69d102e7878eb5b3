// Per-env logic of FrozenLake and Bandit as pure host/device functions (the kernels in toy_envs.cu call these;
// tests/test_hostsim.py runs the same code on the host against the oracle).
//
// Tile codes of the flat uint8 map: 0 'S' (start), 1 'F' (frozen), 2 'H' (hole), 3 'G' (goal) - chosen so that
// the text glyph of a tile is the code itself with S drawn as F (ragen/env/frozen_lake/config.py:17-20).
#pragma once
#include "np_random.cuh"

namespace toy {

constexpr int FL_MAX_SIZE = 8;  // size*size <= 64 cells: reachability runs on one 64-bit mask

// A whole map in two registers: 2 bits per cell.
struct Tiles {
  uint64_t lo, hi;
  RAGEN_HD int get(int i) const { return (int)(((i < 32 ? lo : hi) >> ((i & 31) * 2)) & 3ull); }
  RAGEN_HD void set(int i, int v) {
    const uint64_t m = 3ull << ((i & 31) * 2), b = (uint64_t)v << ((i & 31) * 2);
    if (i < 32) lo = (lo & ~m) | b;
    else hi = (hi & ~m) | b;
  }
};

// generate_random_map (ragen/env/frozen_lake/utils.py:25-47) with is_valid (:6-22).
//   per try: size*size Generator.random() draws (choice(['F','H'], p=[p, 1-p]) = cdf search: 'H' iff u >= cdf0),
//   then integers(size, size=2) twice (start, goal); accept iff start != goal and G is reachable from S over
//   non-hole cells (4-neighbourhood).  Returns the start cell, or -1 after max_tries rejected boards.
RAGEN_HD int fl_generate_map(nprand::Pcg64& g, int size, double cdf0, Tiles& tiles, int max_tries) {
  const int n = size * size;
  uint64_t col_first = 0, col_last = 0;
  for (int r = 0; r < size; ++r) {
    col_first |= 1ull << (r * size);
    col_last |= 1ull << (r * size + size - 1);
  }
  const uint64_t full = (n == 64) ? ~0ull : ((1ull << n) - 1ull);
  for (int t = 0; t < max_tries; ++t) {
    uint64_t open = 0;
    tiles.lo = tiles.hi = 0;
    for (int i = 0; i < n; ++i) {
      const bool hole = !(nprand::pcg_random(g) < cdf0);  // searchsorted(cdf, u, side='right') == 0  <=>  u < cdf[0]
      tiles.set(i, hole ? 2 : 1);
      if (!hole) open |= 1ull << i;
    }
    const int sr = (int)nprand::bounded32(g, (uint32_t)size);
    const int sc = (int)nprand::bounded32(g, (uint32_t)size);
    const int gr = (int)nprand::bounded32(g, (uint32_t)size);
    const int gc = (int)nprand::bounded32(g, (uint32_t)size);
    if (sr == gr && sc == gc) continue;
    const int s = sr * size + sc, e = gr * size + gc;
    tiles.set(s, 0);
    tiles.set(e, 3);
    open |= (1ull << s) | (1ull << e);
    uint64_t reach = 1ull << s;
    for (;;) {
      const uint64_t grow = ((reach << 1) & ~col_first) | ((reach >> 1) & ~col_last) | (reach << size) | (reach >> size);
      const uint64_t next = reach | (grow & open & full);
      if (next == reach) break;
      reach = next;
    }
    if (reach & (1ull << e)) return s;
  }
  return -1;
}

// GymFrozenLakeEnv.step through FrozenLakeEnv.step (ragen/env/frozen_lake/env.py:39-45).  action 1..4 = Left, Down,
// Right, Up (config.py:15,22).  One uniform is drawn per step, also from terminal cells and when not slippery.
// Returns reward (0 or 1); sets done, effective, success.
RAGEN_HD int fl_step(const Tiles& tiles, int size, int& pos, nprand::Pcg64& g, int action, bool slippery, bool& done,
                     bool& effective, bool& success) {
  const int a = action - 1;
  const int before = pos;
  const double u = nprand::pcg_random(g);
  const int here = tiles.get(pos);
  int reward = 0;
  if (here == 2 || here == 3) {
    done = true;
  } else {
    int b = a;
    if (slippery) {
      // argmax(cumsum([1/3, 1/3, 1/3]) > u): the partial sums in binary64
      const double c0 = 1.0 / 3.0;
      const double c1 = c0 + 1.0 / 3.0;
      const int i = (c0 > u) ? 0 : (c1 > u) ? 1 : 2;
      b = (a + i + 3) & 3;  // (a-1)%4, a, (a+1)%4
    }
    int r = pos / size, c = pos - r * size;
    if (b == 0) c = (c > 0) ? c - 1 : 0;
    else if (b == 1) r = (r < size - 1) ? r + 1 : size - 1;
    else if (b == 2) c = (c < size - 1) ? c + 1 : size - 1;
    else r = (r > 0) ? r - 1 : 0;
    pos = r * size + c;
    const int now = tiles.get(pos);
    reward = (now == 3);
    done = (now == 2 || now == 3);
  }
  effective = (before != pos);
  success = (tiles.get(pos) == 3);
  return reward;
}

// FrozenLakeEnv.render text branch (env.py:47-57) as glyph codes: tile code with S shown as F (1); the player's
// cell is 0 'P', 4 when standing in a hole, 5 when standing on the goal; 7 = row separator (also the last byte).
RAGEN_HD void fl_render(const Tiles& tiles, int size, int pos, uint8_t* out) {
  int q = 0;
  for (int r = 0; r < size; ++r) {
    for (int c = 0; c < size; ++c) {
      const int i = r * size + c;
      const int raw = tiles.get(i);
      int t = raw ? raw : 1;
      if (i == pos) t = (raw == 2) ? 4 : (raw == 3) ? 5 : 0;
      out[q++] = (uint8_t)t;
    }
    out[q++] = 7;
  }
}

// BanditEnv.reset -> _randomize_arms (ragen/env/bandit/env.py:24-38): one uniform decides the arm order.
RAGEN_HD bool bandit_reset(nprand::Pcg64& g, uint64_t seed) {
  nprand::pcg_seed(g, seed);
  return nprand::pcg_random(g) < 0.5;  // true: first action id is the lo arm
}
// BanditEnv.step/compute_reward (env.py:40-47, 62-77).  arm_index = action - action_space_start (0 or 1).
// kind: 0 lo arm, 1 hi arm paid hi score, 2 hi arm paid lo score.
RAGEN_HD int bandit_step(nprand::Pcg64& g, bool lo_first, int arm_index, double hi_prob) {
  const bool is_lo = (arm_index == 0) == lo_first;
  if (is_lo) return 0;
  return (nprand::pcg_random(g) < hi_prob) ? 1 : 2;
}

}  // namespace toy
