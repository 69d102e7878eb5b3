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

// ---- the same search, shaped for a GPU thread ------------------------------------------------------------------------
// reverse_play above is the plain restatement (frames that copy the node).  reverse_play_fast visits the same nodes in the
// same order and returns the same `explored` and `best`, but keeps the current node in registers (boxes packed one cell per
// byte), pushes a ONE-BYTE undo record per level instead of a 12-byte frame, takes walls from a bitmask and cell -> interior
// index / (row, col) from tables, and loops once per child tried or level popped with no inner loops - so the 32 rooms of a
// warp advance in lock step.  tests/test_hostsim.py checks the two against each other seed for seed.
struct Fast {
  const uint8_t* interior;  // [H*W] interior index of a cell (garbage on the border: never read there)
  const uint16_t* rc;       // [H*W] row << 8 | col
  const uint32_t* wall;     // nibble (cell & 7) of word wall[(cell >> 3) * wall_stride]: bit a = neighbour a is not a wall
  int wall_stride;
  uint8_t* stack;           // undo record of depth d at stack[d * stack_stride]
  int stack_stride;
};
struct BestFast {
  int p, disp, score;
  uint32_t boxes;
};

RAGEN_HD uint32_t eq4(uint32_t a, uint32_t b) {  // 0xff in every byte where a and b agree
#ifdef __CUDA_ARCH__
  return __vcmpeq4(a, b);
#else
  uint32_t m = 0;
  for (int k = 0; k < 4; ++k)
    if (((a >> (8 * k)) & 0xffu) == ((b >> (8 * k)) & 0xffu)) m |= 0xffu << (8 * k);
  return m;
#endif
}
RAGEN_HD int lowest_byte(uint32_t m) {  // index of the lowest non-zero byte, m != 0
#ifdef __CUDA_ARCH__
  return (__ffs((int)m) - 1) >> 3;
#else
  return __builtin_ctz(m) >> 3;
#endif
}
RAGEN_HD int lowest_bit(unsigned m) {
#ifdef __CUDA_ARCH__
  return __ffs((int)m) - 1;
#else
  return __builtin_ctz(m);
#endif
}
RAGEN_HD uint32_t bcast4(int cell) { return (uint32_t)cell * 0x01010101u; }
RAGEN_HD int step_of(int W, int a) { return a == 0 ? -W : a == 1 ? W : a == 2 ? -1 : 1; }  // utils.py:588-593

// bit a set: moving the player in direction a from p is not blocked by a wall or a box.  Walls come as one nibble per cell
// (bit a: the neighbour in direction a is not a wall), boxes by comparing the packed cells with each neighbour.
RAGEN_HD unsigned open_children(const Fast& m, int W, int p, uint32_t boxes, uint32_t nbmask) {
  unsigned open = (m.wall[(p >> 3) * m.wall_stride] >> ((p & 7) << 2)) & 15u;
  if (eq4(boxes, bcast4(p - W)) & nbmask) open &= ~1u;
  if (eq4(boxes, bcast4(p + W)) & nbmask) open &= ~2u;
  if (eq4(boxes, bcast4(p - 1)) & nbmask) open &= ~4u;
  if (eq4(boxes, bcast4(p + 1)) & nbmask) open &= ~8u;
  return open;
}

#define RAGEN_CSWAP(a, b)          \
  {                                \
    const uint32_t lo_ = a < b ? a : b; \
    b = a < b ? b : a;             \
    a = lo_;                       \
  }
// node_key of the packed node: (interior(box) << 1 | pulled) fields sorted ascending (network of <= 5 exchanges)
template <int NB, typename KeyT>
RAGEN_HD KeyT node_key_fast(const Fast& m, int cell_bits, int p, uint32_t boxes, unsigned pulled) {
  uint32_t f0 = ((uint32_t)m.interior[boxes & 0xffu] << 1) | (pulled & 1u), f1 = 0, f2 = 0, f3 = 0;
  if (NB > 1) f1 = ((uint32_t)m.interior[(boxes >> 8) & 0xffu] << 1) | ((pulled >> 1) & 1u);
  if (NB > 2) f2 = ((uint32_t)m.interior[(boxes >> 16) & 0xffu] << 1) | ((pulled >> 2) & 1u);
  if (NB > 3) f3 = ((uint32_t)m.interior[(boxes >> 24) & 0xffu] << 1) | ((pulled >> 3) & 1u);
  if (NB == 2) RAGEN_CSWAP(f0, f1)
  if (NB == 3) {
    RAGEN_CSWAP(f0, f1)
    RAGEN_CSWAP(f1, f2)
    RAGEN_CSWAP(f0, f1)
  }
  if (NB == 4) {
    RAGEN_CSWAP(f0, f1)
    RAGEN_CSWAP(f2, f3)
    RAGEN_CSWAP(f0, f2)
    RAGEN_CSWAP(f1, f3)
    RAGEN_CSWAP(f1, f2)
  }
  const int fb = cell_bits + 1;
  KeyT key = (KeyT)m.interior[p] | ((KeyT)f0 << cell_bits);
  if (NB > 1) key |= (KeyT)f1 << (cell_bits + fb);
  if (NB > 2) key |= (KeyT)f2 << (cell_bits + 2 * fb);
  if (NB > 3) key |= (KeyT)f3 << (cell_bits + 3 * fb);
  return key;
}
#undef RAGEN_CSWAP

// targets as the search wants them: (row, col) and the broadcast cell of each, fixed for the whole search
template <int NB>
struct TargetInfo {
  uint32_t packed;
  uint16_t rc[NB];
};

// box_swaps * box_displacement_score of a node, 0 unless every target is uncovered (utils.py:466-468); *disp_out = the
// displacement term
template <int NB>
RAGEN_HD int node_score(const Fast& m, const TargetInfo<NB>& tg, uint32_t nbmask, int p, uint32_t boxes, int swaps, int* disp_out) {
  int disp = 0;
  bool any_covered = false;
#pragma unroll
  for (int k = 0; k < NB; ++k) {
    const int b = (boxes >> (8 * k)) & 0xffu, t = (tg.packed >> (8 * k)) & 0xffu;
    const int rb = m.rc[b], rt = tg.rc[k];
    const int dr = (rb >> 8) - (rt >> 8), dc = (rb & 0xff) - (rt & 0xff);
    disp += (dr < 0 ? -dr : dr) + (dc < 0 ? -dc : dc);
    any_covered |= (t == p) || (eq4(boxes, bcast4(t)) & nbmask);
  }
  *disp_out = disp;
  return any_covered ? 0 : swaps * disp;
}

// Bitmap probe through a one-entry register cache of the 64-bit word last touched.  Bits 0..5 of a key are the player's
// interior index (cell_bits >= 6) or all of it, so walking the player around one arrangement of boxes stays in one word and
// only a pull (or a return across one) costs a load; every new bit is written through at once.
struct Granule {
  uint32_t idx;
  uint64_t val;
};
RAGEN_HD bool probe_bitmap(Visited& v, Granule& c, uint32_t key) {
  const uint32_t gi = key >> 6;
  const uint64_t bit = 1ull << (key & 63u);
  if (gi != c.idx) {
    c.val = v.mem64[gi];
    c.idx = gi;
  }
  if (c.val & bit) return true;
  c.val |= bit;
  v.mem64[gi] = c.val;
  return false;
}

// The search as a resumable state: search_begin enters the root, search_step is ONE iteration (return one level if the node
// has no child left, then try one child of the node now current), the search is over when depth < 0.  A caller may stop
// stepping a lane and come back to it later (the generator kernel refills idle lanes of a warp while the others are
// mid-search); reverse_play_nb below is begin + step until done.
// BITMAP: the visited set is the bitmap (keys < 2^24, 32-bit arithmetic) instead of the hash table (64-bit keys).
template <int NB>
struct Search {
  int p, last, swaps, depth, explored;
  uint32_t boxes;
  unsigned pulled, open;
  Granule gran;
  TargetInfo<NB> tg;
};

template <int NB, bool BITMAP>
RAGEN_HD void search_begin(Search<NB>& s, const Geometry& g, const Fast& m, uint32_t targets, int player, int search_depth,
                           Visited& vis, BestFast& best, bool active) {
  const uint32_t nbmask = NB >= 4 ? 0xffffffffu : ((1u << (8 * (NB & 3))) - 1u);
  const int W = g.W, cell_bits = g.cell_bits;
  s.p = player;
  s.last = -1;
  s.swaps = 0;
  s.depth = -1;
  s.explored = 0;
  s.boxes = targets | ~nbmask;  // unused bytes 0xff: the far corner of a 16x16 grid at most, a wall
  s.pulled = 0;
  s.open = 0;
  s.gran = Granule{0xffffffffu, 0ull};
  s.tg.packed = targets;
  if (active) {
#pragma unroll
    for (int k = 0; k < NB; ++k) s.tg.rc[k] = m.rc[(targets >> (8 * k)) & 0xffu];
    best.score = -1;
    best.p = player;
    best.disp = 0;
    best.boxes = s.boxes;
  }
  // the root, entered the way depth_first_search enters any node; a node at stack depth d runs with
  // ttl = search_depth - 1 - d and its children are entered only while that stays positive (:455)
  if (active && search_depth - 1 > 0) {
    if (BITMAP)
      probe_bitmap(vis, s.gran, node_key_fast<NB, uint32_t>(m, cell_bits, s.p, s.boxes, s.pulled));
    else
      visited_test_and_set(vis, node_key_fast<NB, uint64_t>(m, cell_bits, s.p, s.boxes, s.pulled));
    s.explored = 1;
    int disp;
    const int score = node_score<NB>(m, s.tg, nbmask, s.p, s.boxes, s.swaps, &disp);
    if (score > best.score) {
      best.score = score;
      best.p = s.p;
      best.disp = disp;
      best.boxes = s.boxes;
    }
    s.depth = 0;
    s.open = (search_depth - 2 > 0) ? open_children(m, W, s.p, s.boxes, nbmask) : 0u;
  }
}

template <int NB, bool BITMAP>
RAGEN_HD void search_step(Search<NB>& s, const Geometry& g, const Fast& m, int search_depth, Visited& vis, BestFast& best) {
  const uint32_t nbmask = NB >= 4 ? 0xffffffffu : ((1u << (8 * (NB & 3))) - 1u);
  const int W = g.W, cell_bits = g.cell_bits;
  // an iteration = return one level if this node has no child left, then try one child of the node now current
  if (s.depth >= 0 && s.open == 0) {
    if (s.depth == 0) {
      s.depth = -1;
    } else {
      const unsigned code = m.stack[(size_t)s.depth * m.stack_stride];
      const int a = code & 3, combo = code >> 3;
      const int k = combo / 5 - 1, prev_last = combo % 5 - 1;
      const int st = step_of(W, a);
      const int parent = s.p - st;
      if (k >= 0) {
        s.boxes = (s.boxes & ~(0xffu << (8 * k))) | ((uint32_t)(parent - st) << (8 * k));
        if (code & 4u) s.pulled &= ~(1u << k);
        if (k != prev_last) --s.swaps;
      }
      s.last = prev_last;
      s.p = parent;
      --s.depth;
      s.open = open_children(m, W, s.p, s.boxes, nbmask) & ~((2u << a) - 1u);
    }
  }
  if (s.depth >= 0 && s.open != 0) {
    const int a = lowest_bit(s.open);
    s.open &= s.open - 1;
    if (s.explored < MAX_EXPLORED) {
      const int st = step_of(W, a);
      const int n = s.p + st, behind = s.p - st;
      const uint32_t mb = eq4(s.boxes, bcast4(behind)) & nbmask;
      uint32_t cboxes = s.boxes;
      unsigned cpulled = s.pulled, newbit = 0;
      int clast = s.last, cswaps = s.swaps, k = -1;
      if (mb) {  // pull that box onto the cell the player leaves (reverse_move :525-540)
        k = lowest_byte(mb);
        cboxes = (s.boxes & ~(0xffu << (8 * k))) | ((uint32_t)s.p << (8 * k));
        newbit = ((s.pulled >> k) & 1u) ^ 1u;
        cpulled |= 1u << k;
        if (s.last != k) ++cswaps;
        clast = k;
      }
      const bool seen = BITMAP ? probe_bitmap(vis, s.gran, node_key_fast<NB, uint32_t>(m, cell_bits, n, cboxes, cpulled))
                               : visited_test_and_set(vis, node_key_fast<NB, uint64_t>(m, cell_bits, n, cboxes, cpulled));
      if (!seen) {
        ++s.explored;
        int disp;
        const int score = node_score<NB>(m, s.tg, nbmask, n, cboxes, cswaps, &disp);
        if (score > best.score) {
          best.score = score;
          best.p = n;
          best.disp = disp;
          best.boxes = cboxes;
        }
        ++s.depth;
        m.stack[(size_t)s.depth * m.stack_stride] =
            (uint8_t)((unsigned)a | (newbit << 2) | ((unsigned)((k + 1) * 5 + (s.last + 1)) << 3));
        s.p = n;
        s.boxes = cboxes;
        s.pulled = cpulled;
        s.last = clast;
        s.swaps = cswaps;
        s.open = (search_depth - 1 - (s.depth + 1) > 0) ? open_children(m, W, s.p, s.boxes, nbmask) : 0u;
      }
    }
  }
}

// `active` false: the caller only takes part in the warp-wide loop (device) and nothing is touched.
template <int NB, bool BITMAP>
RAGEN_HD int reverse_play_nb(const Geometry& g, const Fast& m, uint32_t targets, int player, int search_depth, Visited& vis,
                             BestFast& best, bool active) {
  Search<NB> s;
  search_begin<NB, BITMAP>(s, g, m, targets, player, search_depth, vis, best, active);
#ifdef __CUDA_ARCH__
  // every lane of the warp is here: the vote is also the point where the lanes meet again after each iteration
  while (__any_sync(0xffffffffu, s.depth >= 0)) {
#else
  while (s.depth >= 0) {
#endif
    search_step<NB, BITMAP>(s, g, m, search_depth, vis, best);
  }
  return s.explored;
}

RAGEN_HD int reverse_play_fast(const Geometry& g, const Fast& m, uint32_t targets, int player, int search_depth, Visited& vis,
                               BestFast& best, bool active = true) {
  // every lane of a warp takes the same case (nb and the kind of visited set are launch-wide)
  if (vis.mode == 0) {
    switch (g.nb) {
      case 1: return reverse_play_nb<1, true>(g, m, targets, player, search_depth, vis, best, active);
      case 2: return reverse_play_nb<2, true>(g, m, targets, player, search_depth, vis, best, active);
      case 3: return reverse_play_nb<3, true>(g, m, targets, player, search_depth, vis, best, active);
      default: return reverse_play_nb<4, true>(g, m, targets, player, search_depth, vis, best, active);
    }
  }
  switch (g.nb) {
    case 1: return reverse_play_nb<1, false>(g, m, targets, player, search_depth, vis, best, active);
    case 2: return reverse_play_nb<2, false>(g, m, targets, player, search_depth, vis, best, active);
    case 3: return reverse_play_nb<3, false>(g, m, targets, player, search_depth, vis, best, active);
    default: return reverse_play_nb<4, false>(g, m, targets, player, search_depth, vis, best, active);
  }
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
  mt::State py{w.mt_py, 0, 1}, np{w.mt_np, 0, 1};
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

// ---- pieces of generate_room for callers that run the search with reverse_play_fast ----------------------------------
RAGEN_HD void fill_tables(const Geometry& g, uint8_t* interior_tab, uint16_t* rc_tab) {
  for (int cell = 0; cell < g.H * g.W; ++cell) {
    const int r = cell / g.W, c = cell - r * g.W;
    interior_tab[cell] = (uint8_t)((r - 1) * (g.W - 2) + (c - 1));
    rc_tab[cell] = (uint16_t)((r << 8) | c);
  }
}
RAGEN_HD void fill_walls(const Geometry& g, const uint8_t* structure, uint32_t* wall, int stride) {
  const int HW = g.H * g.W, W = g.W;
  for (int w = 0; w * 8 < HW; ++w) {
    uint32_t bits = 0;
    for (int b = 0; b < 8 && w * 8 + b < HW; ++b) {
      const int cell = w * 8 + b;
      unsigned nib = 0;  // directions up, down, left, right (utils.py:588-593); off-grid neighbours count as walls
      if (cell - W >= 0 && structure[cell - W] != 0) nib |= 1u;
      if (cell + W < HW && structure[cell + W] != 0) nib |= 2u;
      if (cell - 1 >= 0 && structure[cell - 1] != 0) nib |= 4u;
      if (cell + 1 < HW && structure[cell + 1] != 0) nib |= 8u;
      bits |= nib << (4 * b);
    }
    wall[w * stride] = bits;
  }
}
RAGEN_HD uint32_t pack_cells(const uint8_t* cells, int nb) {
  uint32_t v = 0;
  for (int k = 0; k < nb; ++k) v |= (uint32_t)cells[k] << (8 * k);
  return v;
}
// room_state from the best node (:257-259) + add_random_player_movement
RAGEN_HD void finish_room(mt::State& py, const Geometry& g, const uint8_t* fixed, uint8_t* state, const BestFast& best) {
  const int HW = g.H * g.W;
  for (int i = 0; i < HW; ++i) state[i] = fixed[i];
  for (int k = 0; k < g.nb; ++k) state[(best.boxes >> (8 * k)) & 0xffu] = 4;  // pulled boxes (3) become plain boxes
  state[best.p] = 5;
  jitter(py, g, fixed, state, best.p, best.disp == 1 ? 0.8 : 0.5);
}

#ifndef __CUDACC__
// generate_room with the fast search, as the CUDA kernel composes it (host build only: the cross-check of the two searches)
inline int generate_room_fast(const Geometry& g, int search_depth, int num_gen_steps, uint64_t seed, Work& w, uint8_t* fixed,
                              uint8_t* state, int* explored_out) {
  mt::State py{w.mt_py, 0, 1}, np{w.mt_np, 0, 1};
  mt::seed_python(py, seed);
  mt::seed_numpy(np, (uint32_t)seed);
  uint8_t interior_tab[MAX_CELLS];
  uint16_t rc_tab[MAX_CELLS];
  uint32_t wall[MAX_CELLS / 8];
  fill_tables(g, interior_tab, rc_tab);
  Fast m{interior_tab, rc_tab, wall, 1, reinterpret_cast<uint8_t*>(w.stack), 1};
  uint8_t* floor = w.scratch;
  uint8_t targets[MAX_BOXES];
  BestFast best;
  best.disp = 0;
  int player = 0, explored = 0;
  for (int t = 0; t < 4; ++t) {
    topology(py, g, num_gen_steps, 0.35, floor);
    if (!place(np, g, floor, fixed, player, targets)) return NO_SPACE;
    if (w.vis.mode == 0)
      for (uint64_t i = 0; i < w.vis_words; ++i) w.vis.mem[i] = 0;
    else
      for (uint64_t i = 0; i < w.vis_words; ++i) w.vis.mem64[i] = 0;
    fill_walls(g, fixed, wall, 1);
    explored = reverse_play_fast(g, m, pack_cells(targets, g.nb), player, search_depth, w.vis, best);
    if (best.disp > 0) break;
  }
  if (explored_out) *explored_out = explored;
  if (best.disp == 0) return SCORE_ZERO;
  finish_room(py, g, fixed, state, best);
  return OK;
}
#endif

}  // namespace sgen
