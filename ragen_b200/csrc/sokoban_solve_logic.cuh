// Shortest push solution of a Sokoban room: the per-room logic of get_shortest_action_path
// (ragen/env/sokoban/utils.py:11-101) as a pure host/device function (the kernel in sokoban_solve.cu calls it;
// tests/test_hostsim.py runs the same code on the host against paths the unmodified reference function returned).
//
// The reference is a breadth-first search over whole boards: FIFO queue of (board, path), duplicates dropped when
// POPPED (marshal of the board in a set), goal test "no cell equals 4" at pop time, the four moves tried in the order
// up, down, left, right (actions 1..4), `return []` as soon as a popped entry's path is longer than MAX_DEPTH.
// Restated here on compact states (player cell, sorted box cells):
//   * dropping duplicates when they are PUSHED instead of popped visits the same states in the same order (the first
//     queued copy of a state is the first popped), so the path found - not only its length - is the reference's;
//   * a board is determined by (player, set of boxes) given room_fixed, so that pair is the state key;
//   * "path longer than MAX_DEPTH popped -> []" becomes: stop when the head of the queue is deeper than max_depth.
// The search is sequential by nature when the PATH has to match (discovery order decides the parent of a state).
// Out-of-bound moves are refused as the reference does for the player; for the pushed box the reference indexes
// room_fixed before its bound check (numpy wraps negative indices, raises past the end) - rooms with a wall border,
// which is every room adopt accepts, never get there, and here such a push is simply refused.
#pragma once
#include <stdint.h>

#ifndef RAGEN_HD
#ifdef __CUDACC__
#define RAGEN_HD __host__ __device__ __forceinline__
#else
#define RAGEN_HD inline
#endif
#endif

namespace ssolve {

constexpr int MAX_BOXES = 4;
enum { SOLVED = 0, NO_SOLUTION = 1, OVERFLOW = 2, BAD_ROOM = 3 };

// state key: player cell in bits 0-9, box j (ascending cells) in bits 10+10j .. (cells < 1024), at most 4 boxes: 50 bits
RAGEN_HD uint64_t make_key(int player, const int* boxes, int nb) {
  uint64_t k = (uint64_t)player;
  for (int j = 0; j < nb; ++j) k |= (uint64_t)boxes[j] << (10 + 10 * j);
  return k;
}
RAGEN_HD uint32_t hash_key(uint64_t k) {
  k ^= k >> 33;
  k *= 0xff51afd7ed558ccdull;
  k ^= k >> 33;
  k *= 0xc4ceb9fe1a85ec53ull;
  k ^= k >> 33;
  return (uint32_t)k;
}

// Workspace of one room (caller-owned): table[cap_table] (open addressing, 0 = empty, stored key + 1), queue[max_states]
// (state key | action << 50), parent[max_states].  `stride` lets a warp interleave its rooms' arrays (element i of this
// room sits at [i * stride]); 1 on the host.
struct Workspace {
  uint64_t* table;
  uint64_t* queue;
  uint32_t* parent;
  uint32_t cap_table;  // power of two, >= 2 * max_states
  uint32_t max_states;
  uint32_t stride;
};

// fixed: room_fixed (0 wall, 1 floor, 2 target); state: room_state (0..5).  Writes the actions (1..4) to path[0..len)
// when path != nullptr.  Returns a status; *len_out = number of actions (0 when not SOLVED).
template <typename FixedAt>
RAGEN_HD int solve(const FixedAt& fixed_at, const uint8_t* state, int H, int W, int max_depth, const Workspace& ws,
                   uint8_t* path, int path_stride, int* len_out, int* explored_out) {
  const int HW = H * W;
  *len_out = 0;
  if (explored_out) *explored_out = 0;
  int player = -1, nb = 0;
  int boxes[MAX_BOXES];
  for (int i = 0; i < HW; ++i) {
    const int s = state[i];
    if (s == 5 && player < 0) player = i;  // np.argwhere(room_state == 5)[0]
    if (s == 3 || s == 4) {
      if (nb == MAX_BOXES) return BAD_ROOM;
      boxes[nb++] = i;  // ascending by construction
    }
  }
  if (player < 0 || HW > 1024) return BAD_ROOM;
  const uint32_t S = ws.stride, mask = ws.cap_table - 1u;
  for (uint32_t i = 0; i < ws.cap_table; ++i) ws.table[(size_t)i * S] = 0;
  const int dr[4] = {-1, 1, 0, 0}, dc[4] = {0, 0, -1, 1};

  uint32_t head = 0, tail = 0;
  {
    const uint64_t k = make_key(player, boxes, nb);
    ws.table[(size_t)(hash_key(k) & mask) * S] = k + 1;
    ws.queue[0] = k;
    ws.parent[0] = 0xffffffffu;
    tail = 1;
  }
  uint32_t level_end = 1;  // entries [head, level_end) are at depth `depth`
  int depth = 0;
  while (head < tail) {
    if (head == level_end) {
      ++depth;
      level_end = tail;
    }
    if (depth > max_depth) break;  // the reference pops a path longer than MAX_DEPTH and returns []
    const uint64_t cur = ws.queue[(size_t)head * S] & ((1ull << 50) - 1ull);
    const int p = (int)(cur & 1023u);
    int b[MAX_BOXES];
    bool all_on = true;
    for (int j = 0; j < nb; ++j) {
      b[j] = (int)((cur >> (10 + 10 * j)) & 1023u);
      all_on = all_on && (fixed_at(b[j]) == 2);
    }
    if (all_on) {  // `if not boxes_not_on_target: return path`
      int n = 0;
      for (uint32_t at = head; ws.parent[(size_t)at * S] != 0xffffffffu; at = ws.parent[(size_t)at * S]) ++n;
      if (path) {
        int i = n;
        for (uint32_t at = head; ws.parent[(size_t)at * S] != 0xffffffffu; at = ws.parent[(size_t)at * S])
          path[(size_t)(--i) * path_stride] = (uint8_t)(ws.queue[(size_t)at * S] >> 50);
      }
      *len_out = n;
      if (explored_out) *explored_out = (int)head + 1;
      return SOLVED;
    }
    const int pr = p / W, pc = p - pr * W;
    for (int a = 0; a < 4; ++a) {
      const int nr = pr + dr[a], nc = pc + dc[a];
      if (nr < 0 || nr >= H || nc < 0 || nc >= W) continue;
      const int np = nr * W + nc;
      if (fixed_at(np) == 0) continue;
      int nbx[MAX_BOXES];
      int hit = -1;
      for (int j = 0; j < nb; ++j) {
        nbx[j] = b[j];
        if (b[j] == np) hit = j;
      }
      if (hit >= 0) {
        const int br = nr + dr[a], bc = nc + dc[a];
        if (br < 0 || br >= H || bc < 0 || bc >= W) continue;
        const int q = br * W + bc;
        if (fixed_at(q) == 0) continue;
        bool blocked = false;
        for (int j = 0; j < nb; ++j) blocked = blocked || (b[j] == q);
        if (blocked) continue;
        nbx[hit] = q;
        // keep the box cells ascending (one element moved: bubble it into place)
        for (int j = hit; j + 1 < nb && nbx[j] > nbx[j + 1]; ++j) {
          const int t = nbx[j];
          nbx[j] = nbx[j + 1];
          nbx[j + 1] = t;
        }
        for (int j = hit; j > 0 && nbx[j] < nbx[j - 1]; --j) {
          const int t = nbx[j];
          nbx[j] = nbx[j - 1];
          nbx[j - 1] = t;
        }
      }
      const uint64_t k = make_key(np, nbx, nb);
      uint32_t slot = hash_key(k) & mask;
      bool seen = false;
      for (;;) {
        const uint64_t e = ws.table[(size_t)slot * S];
        if (e == 0) break;
        if (e == k + 1) {
          seen = true;
          break;
        }
        slot = (slot + 1) & mask;
      }
      if (seen) continue;
      if (tail >= ws.max_states) {
        if (explored_out) *explored_out = (int)head + 1;
        return OVERFLOW;
      }
      ws.table[(size_t)slot * S] = k + 1;
      ws.queue[(size_t)tail * S] = k | ((uint64_t)(a + 1) << 50);
      ws.parent[(size_t)tail * S] = head;
      ++tail;
    }
    ++head;
  }
  if (explored_out) *explored_out = (int)head;
  return NO_SOLUTION;
}

}  // namespace ssolve
