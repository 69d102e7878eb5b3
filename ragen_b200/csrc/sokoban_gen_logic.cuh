// Sokoban room generator (reverse-play search) as pure host/device functions: one call generates one room with the
// reference's exact random streams and search order, so the room for a seed is THE room the reference generates.
//
// Follows ragen/env/sokoban/utils.py: generate_room :221-278, room_topology_generation :281-355,
// place_boxes_and_player :358-397, reverse_playing :408-437, depth_first_search :440-498, reverse_move :501-542,
// box_displacement_score :545-560, add_random_player_movement :152-213; seeded as SokobanEnv.reset does through
// all_seed (ragen/env/sokoban/env.py:28-42, ragen/utils.py:7-18).
//
// Representation (not the reference's int64 boards): cells are flat indices r*W+c; a search node is the player cell
// plus, per box (in row-major order of its home target), its cell and a "was pulled" bit.  The reference's visited
// set keys on the whole board, which is exactly (player, {(box cell, pulled)}) - kept here as a bitmap over that key
// space (or a hash set when the space is too large), so membership is exact.
#pragma once
#include "mt19937.cuh"

namespace sgen {

constexpr int MAX_BOXES = 4;
constexpr int MAX_CELLS = 256;      // H*W
constexpr int MAX_EXPLORED = 300000;  // utils.py:457
enum Status { OK = 0, NO_SPACE = 1, SCORE_ZERO = 2 };

struct Frame {  // one activation of depth_first_search
  uint8_t p;
  uint8_t box[MAX_BOXES];
  uint8_t pulled;      // bit k: box k was pulled at least once (board code 3 instead of 4)
  int8_t last_pull;    // index of the box pulled last, -1 none
  uint8_t next;        // next action to try (0..4)
  uint16_t swaps;
  uint16_t pad;
};

// Visited set: exact membership over keys < 2^key_bits.  mode 0: bitmap (words = 2^key_bits / 32);
// mode 1: open-addressing table of key+1 (capacity = power of two, `mask` = capacity-1).
struct Visited {
  uint32_t* mem;
  uint64_t* mem64;
  uint32_t mask;
  int mode;
};
RAGEN_HD bool visited_test_and_set(Visited& v, uint64_t key) {
  if (v.mode == 0) {
    const uint32_t bit = 1u << (key & 31);
    uint32_t& w = v.mem[key >> 5];
    if (w & bit) return true;
    w |= bit;
    return false;
  }
  uint32_t h = (uint32_t)((key * 0x9E3779B97F4A7C15ull) >> 32) & v.mask;
  for (;;) {
    const uint64_t cur = v.mem64[h];
    if (cur == key + 1) return true;
    if (cur == 0) {
      v.mem64[h] = key + 1;
      return false;
    }
    h = (h + 1) & v.mask;
  }
}
RAGEN_HD bool visited_test(const Visited& v, uint64_t key) {
  if (v.mode == 0) return (v.mem[key >> 5] >> (key & 31)) & 1u;
  uint32_t h = (uint32_t)((key * 0x9E3779B97F4A7C15ull) >> 32) & v.mask;
  for (;;) {
    const uint64_t cur = v.mem64[h];
    if (cur == key + 1) return true;
    if (cur == 0) return false;
    h = (h + 1) & v.mask;
  }
}

struct Geometry {
  int H, W, nb, cell_bits;  // cell_bits: bits of an interior cell index (r-1)*(W-2)+(c-1)
};
RAGEN_HD int interior(const Geometry& g, int cell) {
  const int r = cell / g.W, c = cell - r * g.W;
  return (r - 1) * (g.W - 2) + (c - 1);
}
// key = player | sorted (box<<1 | pulled) fields: the board, canonically
RAGEN_HD uint64_t node_key(const Geometry& g, int p, const uint8_t* box, unsigned pulled) {
  uint32_t f[MAX_BOXES];
  for (int k = 0; k < g.nb; ++k) f[k] = ((uint32_t)interior(g, box[k]) << 1) | ((pulled >> k) & 1u);
  for (int i = 1; i < g.nb; ++i) {  // insertion sort, nb <= 4
    const uint32_t x = f[i];
    int j = i - 1;
    while (j >= 0 && f[j] > x) {
      f[j + 1] = f[j];
      --j;
    }
    f[j + 1] = x;
  }
  uint64_t key = (uint64_t)interior(g, p);
  int shift = g.cell_bits;
  for (int k = 0; k < g.nb; ++k) {
    key |= (uint64_t)f[k] << shift;
    shift += g.cell_bits + 1;
  }
  return key;
}
RAGEN_HD int key_bits(const Geometry& g) { return g.cell_bits + g.nb * (g.cell_bits + 1); }

// ---- room_topology_generation (utils.py:281-355): floor[cell] = 1 floor / 0 wall ------------------------------------
RAGEN_HD void topology(mt::State& py, const Geometry& g, int num_steps, double p_change, uint8_t* floor) {
  const int H = g.H, W = g.W;
  const int DR[4] = {1, 0, -1, 0}, DC[4] = {0, 1, 0, -1};  // directions (:324)
  // the five 3x3 masks (:295-321) as offsets from the walk position; 0xF = unused
  const int8_t SR[5][4] = {{0, 0, 0, 9}, {-1, 0, 1, 9}, {0, 0, 1, 9}, {0, 0, 1, 1}, {0, 0, 1, 9}};
  const int8_t SC[5][4] = {{-1, 0, 1, 9}, {0, 0, 0, 9}, {-1, 0, 0, 9}, {-1, 0, -1, 0}, {0, 1, 0, 9}};
  for (int i = 0; i < H * W; ++i) floor[i] = 0;
  int d = (int)mt::py_randbelow(py, 4);                      // random.sample(directions, 1)[0]
  int r = 1 + (int)mt::py_randbelow(py, (uint32_t)(H - 1));  // random.randint(1, dim_x - 1)
  int c = 1 + (int)mt::py_randbelow(py, (uint32_t)(W - 1));
  for (int s = 0; s < num_steps; ++s) {
    if (mt::py_random(py) < p_change) d = (int)mt::py_randbelow(py, 4);
    r += DR[d];
    c += DC[d];
    r = r > H - 2 ? H - 2 : r;
    r = r < 1 ? 1 : r;
    c = c > W - 2 ? W - 2 : c;
    c = c < 1 ? 1 : c;
    const int m = (int)mt::py_randbelow(py, 5);  // random.sample(masks, 1)[0]
    for (int k = 0; k < 4; ++k)
      if (SR[m][k] != 9) floor[(r + SR[m][k]) * W + (c + SC[m][k])] = 1;
  }
  for (int i = 0; i < H; ++i) floor[i * W] = floor[i * W + W - 1] = 0;
  for (int j = 0; j < W; ++j) floor[j] = floor[(H - 1) * W + j] = 0;
}

// ---- place_boxes_and_player (utils.py:358-397): uniform picks among the remaining floor cells in row-major order.
// structure[cell]: 0 wall, 1 floor, 2 target.  targets[] come back sorted row-major (the key order of box_mapping).
RAGEN_HD bool place(mt::State& np, const Geometry& g, const uint8_t* floor, uint8_t* structure, int& player, uint8_t* targets) {
  const int HW = g.H * g.W;
  int n_free = 0;
  for (int i = 0; i < HW; ++i) {
    structure[i] = floor[i];
    n_free += floor[i];
  }
  if (n_free <= g.nb + 1) return false;  // RuntimeError('Not enough free spots')
  // k-th free cell, skipping cells already taken (marked 5 / 2 in `structure` while placing)
  int ind = (int)mt::np_randint(np, (uint32_t)n_free);
  player = -1;
  for (int i = 0; i < HW; ++i)
    if (structure[i] == 1 && ind-- == 0) {
      player = i;
      break;
    }
  structure[player] = 5;
  for (int b = 0; b < g.nb; ++b) {
    --n_free;
    ind = (int)mt::np_randint(np, (uint32_t)n_free);
    for (int i = 0; i < HW; ++i)
      if (structure[i] == 1 && ind-- == 0) {
        structure[i] = 2;
        break;
      }
  }
  structure[player] = 1;  // room_structure: the player's cell is plain floor (:250-251)
  int k = 0;
  for (int i = 0; i < HW; ++i)
    if (structure[i] == 2) targets[k++] = (uint8_t)i;
  return true;
}

// ---- reverse_playing / depth_first_search (utils.py:408-498) --------------------------------------------------------
struct Best {
  int p, disp, score;
  uint8_t box[MAX_BOXES];
};

// Enters a node the way depth_first_search does: ttl and explored-count cut-offs, visited test, scoring, insertion.
// Returns true when the node was new (its children must then be expanded).  A node at stack depth d runs with
// ttl = search_depth - 1 - d after the decrement at function entry (:455).
RAGEN_HD bool dfs_enter(const Geometry& g, const uint8_t* targets, Visited& vis, int& explored, int search_depth, int depth,
                        int p, const uint8_t* box, unsigned pulled, int last_pull, int swaps, Frame* stack, Best& best) {
  if (search_depth - 1 - depth <= 0 || explored >= MAX_EXPLORED) return false;
  if (visited_test_and_set(vis, node_key(g, p, box, pulled))) return false;
  ++explored;
  int disp = 0, open_targets = 0;
  for (int k = 0; k < g.nb; ++k) {
    const int br = box[k] / g.W, bc = box[k] - br * g.W, tr = targets[k] / g.W, tc = targets[k] - tr * g.W;
    disp += (br > tr ? br - tr : tr - br) + (bc > tc ? bc - tc : tc - bc);
    bool covered = (targets[k] == p);
    for (int j = 0; j < g.nb; ++j) covered |= (box[j] == targets[k]);
    open_targets += covered ? 0 : 1;
  }
  int score = swaps * disp;                 // box_swaps * box_displacement_score (:466)
  if (open_targets != g.nb) score = 0;      // every target must be uncovered (:467-468)
  if (score > best.score) {
    best.score = score;
    best.p = p;
    best.disp = disp;
    for (int k = 0; k < g.nb; ++k) best.box[k] = box[k];
  }
  Frame& f = stack[depth];
  f.p = (uint8_t)p;
  for (int k = 0; k < g.nb; ++k) f.box[k] = box[k];
  f.pulled = (uint8_t)pulled;
  f.last_pull = (int8_t)last_pull;
  f.next = 0;
  f.swaps = (uint16_t)swaps;
  return true;
}

// `stack` holds search_depth frames.  Returns the number of explored states.
RAGEN_HD int reverse_play(const Geometry& g, const uint8_t* structure, const uint8_t* targets, int player, int search_depth,
                          Visited& vis, Frame* stack, Best& best) {
  const int STEP[4] = {-g.W, g.W, -1, 1};  // CHANGE_COORDINATES: up, down, left, right (utils.py:588-593)
  best.score = -1;
  best.p = player;
  best.disp = 0;
  for (int k = 0; k < g.nb; ++k) best.box[k] = targets[k];
  int explored = 0;
  int depth = 0;
  if (!dfs_enter(g, targets, vis, explored, search_depth, 0, player, targets, 0u, -1, 0, stack, best)) return explored;
  while (depth >= 0) {
    Frame& f = stack[depth];
    if (f.next >= 4) {
      --depth;
      continue;
    }
    const int a = f.next++;
    const int p = f.p, n = p + STEP[a];
    bool blocked = (structure[n] == 0);
    for (int k = 0; k < g.nb; ++k) blocked |= (f.box[k] == n);
    if (blocked) continue;  // the child equals its parent, which is explored: the recursive call returns at once
    uint8_t box[MAX_BOXES];
    for (int k = 0; k < g.nb; ++k) box[k] = f.box[k];
    unsigned pulled = f.pulled;
    int last_pull = f.last_pull, swaps = f.swaps;
    const int behind = p - STEP[a];
    for (int k = 0; k < g.nb; ++k)
      if (f.box[k] == behind) {  // pull that box onto the cell the player leaves (reverse_move :525-540)
        box[k] = (uint8_t)p;
        pulled |= 1u << k;
        if (last_pull != k) ++swaps;
        last_pull = k;
      }
    if (dfs_enter(g, targets, vis, explored, search_depth, depth + 1, n, box, pulled, last_pull, swaps, stack, best)) ++depth;
  }
  return explored;
}

// ---- add_random_player_movement (utils.py:152-213) ------------------------------------------------------------------
RAGEN_HD int jitter(mt::State& py, const Geometry& g, const uint8_t* structure, uint8_t* state, int player, double p_move) {
  if (mt::py_random(py) > p_move) return player;
  const int STEP[4] = {-g.W, g.W, -1, 1};
  int seen[4] = {player, -1, -1, -1};
  int steps = 0;
  while (steps < 3) {
    int opt[4], n_opt = 0;
    for (int a = 0; a < 4; ++a) {
      const int n = player + STEP[a];
      bool ok = (state[n] == 1 || state[n] == 2);
      for (int j = 0; j <= steps; ++j) ok &= (seen[j] != n);
      if (ok) opt[n_opt++] = n;
    }
    if (!n_opt) break;
    const int n = opt[mt::py_randbelow(py, (uint32_t)n_opt)];  // random.choice(valid_moves)
    state[player] = structure[player];
    state[n] = 5;
    player = n;
    seen[++steps] = n;
    if (steps >= 3 || mt::py_random(py) > 0.5) break;
  }
  return player;
}

// ---- generate_room under all_seed(seed) (utils.py:221-278, env.py:28-38) --------------------------------------------
// Work memory, all caller-owned: two MT19937 states (624 words each), `stack` (search_depth frames), the visited set
// (cleared here before every search) and `scratch` (2*H*W bytes).  Writes room_fixed / room_state (codes of
// utils.py:563-570) and returns a Status; *explored_out (may be null) = states explored by the last search.
struct Work {
  uint32_t* mt_py;
  uint32_t* mt_np;
  Frame* stack;
  Visited vis;
  uint64_t vis_words;  // 32-bit words (bitmap) or 64-bit slots (hash) to clear
  uint8_t* scratch;
};

RAGEN_HD int generate_room(const Geometry& g, int search_depth, int num_gen_steps, uint64_t seed, Work& w, uint8_t* fixed,
                           uint8_t* state, int* explored_out) {
  const int HW = g.H * g.W;
  mt::State py{w.mt_py, 0}, np{w.mt_np, 0};
  mt::seed_python(py, seed);
  mt::seed_numpy(np, (uint32_t)seed);
  uint8_t* floor = w.scratch;
  uint8_t targets[MAX_BOXES];
  Best best;
  best.disp = 0;
  int player = 0, explored = 0;
  for (int t = 0; t < 4; ++t) {  // tries (:244)
    topology(py, g, num_gen_steps, 0.35, floor);
    if (!place(np, g, floor, fixed, player, targets)) return NO_SPACE;
    if (w.vis.mode == 0)
      for (uint64_t i = 0; i < w.vis_words; ++i) w.vis.mem[i] = 0;
    else
      for (uint64_t i = 0; i < w.vis_words; ++i) w.vis.mem64[i] = 0;
    explored = reverse_play(g, fixed, targets, player, search_depth, w.vis, w.stack, best);
    if (best.disp > 0) break;
  }
  if (explored_out) *explored_out = explored;
  if (best.disp == 0) return SCORE_ZERO;  // RuntimeWarning('Generated Model with score == 0')
  for (int i = 0; i < HW; ++i) state[i] = fixed[i];
  for (int k = 0; k < g.nb; ++k) state[best.box[k]] = 4;  // pulled boxes (3) become plain boxes (:257)
  state[best.p] = 5;
  jitter(py, g, fixed, state, best.p, best.disp == 1 ? 0.8 : 0.5);
  return OK;
}

}  // namespace sgen
