// TEST INFRASTRUCTURE: compiles the host/device logic headers of ragen_b200/csrc with g++ so that the CPU test
// suite (tests/test_hostsim.py) can run the exact per-element code the kernels run - without a GPU - against the
// oracle and numpy.  Nothing in the product imports or links this; it is not a CPU fallback.
#include <stdint.h>
#include <string.h>

#include "../../ragen_b200/csrc/np_random.cuh"
#include "../../ragen_b200/csrc/parse_logic.cuh"
#include "../../ragen_b200/csrc/score_logic.cuh"
#include "../../ragen_b200/csrc/sokoban_gen_logic.cuh"
#include "../../ragen_b200/csrc/sokoban_solve_logic.cuh"
#include <stdlib.h>
#include "../../ragen_b200/csrc/toy_logic.cuh"
#include "../../ragen_b200/csrc/scan_tiles.cuh"

extern "C" {

// ---- numpy generator ---------------------------------------------------------------------------------------
void hs_seed_state(uint64_t seed, uint64_t* out4) {
  nprand::Pcg64 g;
  nprand::pcg_seed(g, seed);
  out4[0] = g.state_hi;
  out4[1] = g.state_lo;
  out4[2] = g.inc_hi;
  out4[3] = g.inc_lo;
}
void hs_random(uint64_t seed, int k, double* out) {
  nprand::Pcg64 g;
  nprand::pcg_seed(g, seed);
  for (int i = 0; i < k; ++i) out[i] = nprand::pcg_random(g);
}
// `calls` successive Generator.integers(n, size=size) calls
void hs_integers(uint64_t seed, uint32_t n, int size, int calls, uint32_t* out) {
  nprand::Pcg64 g;
  nprand::pcg_seed(g, seed);
  for (int c = 0; c < calls; ++c)
    for (int i = 0; i < size; ++i) out[c * size + i] = nprand::bounded32(g, n);
}

// ---- FrozenLake ----------------------------------------------------------------------------------------------
// reset + a stream of actions for one env; returns the start cell (or -1)
int hs_frozenlake(uint64_t seed, int size, double p, int slippery, const uint8_t* actions, int T, uint8_t* tiles_out,
                  uint8_t* pos_out, double* reward_out, uint8_t* done_out, uint8_t* flags_out, uint8_t* glyph_reset,
                  uint8_t* glyph_steps) {
  const double cdf0 = p / (p + (1.0 - p));
  nprand::Pcg64 g;
  nprand::pcg_seed(g, seed);
  toy::Tiles tiles{0, 0};
  const int start = toy::fl_generate_map(g, size, cdf0, tiles, 100000);
  if (start < 0) return -1;
  for (int i = 0; i < size * size; ++i) tiles_out[i] = (uint8_t)tiles.get(i);
  nprand::pcg_seed(g, seed);
  (void)nprand::pcg_random(g);
  int pos = start;
  const int G = size * (size + 1);
  toy::fl_render(tiles, size, pos, glyph_reset);
  for (int t = 0; t < T; ++t) {
    bool d, e, s;
    const int r = toy::fl_step(tiles, size, pos, g, actions[t], slippery != 0, d, e, s);
    pos_out[t] = (uint8_t)pos;
    reward_out[t] = r;
    done_out[t] = d;
    flags_out[t] = (uint8_t)((e ? 1 : 0) | (s ? 2 : 0));
    toy::fl_render(tiles, size, pos, glyph_steps + (size_t)t * G);
  }
  return start;
}

// ---- Bandit --------------------------------------------------------------------------------------------------
// returns lo_first; kind_out = bandit_step result for `arm_index`
int hs_bandit(uint64_t seed, int arm_index, double hi_prob, int* kind_out) {
  nprand::Pcg64 g;
  const bool lf = toy::bandit_reset(g, seed);
  *kind_out = toy::bandit_step(g, lf, arm_index, hi_prob);
  return lf ? 1 : 0;
}

// ---- parse ---------------------------------------------------------------------------------------------------
// Sequential driver with the same phases as parse_kernel.  out: flags, spans[4], act_spans[2*max], n_parsed,
// n_total, codes[max], n_valid.  `scratch` has at least n bytes.
void hs_parse(const uint8_t* text, int n, int enable_think, const char* sep, int max_actions, const uint8_t* names,
              const int32_t* name_off, const uint8_t* codes, int first, int count, uint8_t* scratch, int32_t* out) {
  parse::Matcher m;
  parse::matcher_init(m, enable_think != 0);
  for (int i = 0; i < n && m.state != 3; ++i)
    if (text[i] == '<') {
      const int tag = parse::classify_tag(text + i, text + n);
      if (tag != parse::T_NONE) parse::matcher_feed(m, text, n, i, tag);
    }
  int fl = 0, ts = 0, te = 0, as = 0, ae = 0;
  parse::ParseTail tail;
  tail.total = tail.kept = tail.nvalid = 0;
  if (m.state == 3) {
    fl = 1;
    ts = m.think_s;
    te = m.think_e;
    as = m.ans_s;
    ae = m.ans_e;
    const uint8_t* abuf = text;
    for (int which = 0; which < 2; ++which) {
      if (which == 0 && !enable_think) continue;
      int s = which ? as : ts, e = which ? ae : te;
      if (parse::has_special(text, s, e)) {
        memcpy(scratch + s, text + s, (size_t)(e - s));
        parse::clean_content(scratch, s, e);
        fl |= which ? 4 : 2;
        if (which) abuf = scratch;
      } else {
        parse::strip(text, s, e);
      }
      if (which) {
        as = s;
        ae = e;
      } else {
        ts = s;
        te = e;
      }
    }
    parse::ActionTable tab{names, name_off, codes, first, count};
    if (count >= 0) fl |= 16;
    const int sep_len = (int)strlen(sep);
    parse::parse_finish(abuf, as, ae, (const uint8_t*)sep, sep_len, max_actions, tab, tail);
    if (tail.total > max_actions) fl |= 8;
  }
  int k = 0;
  out[k++] = fl;
  out[k++] = ts;
  out[k++] = te - ts;
  out[k++] = as;
  out[k++] = ae - as;
  out[k++] = tail.kept;
  out[k++] = tail.total;
  out[k++] = tail.nvalid;
  for (int a = 0; a < max_actions; ++a) {
    out[k++] = a < tail.kept ? tail.spans[2 * a] : 0;
    out[k++] = a < tail.kept ? tail.spans[2 * a + 1] : 0;
  }
  for (int a = 0; a < max_actions; ++a) out[k++] = a < tail.nvalid ? tail.codes[a] : 0;
}

int hs_countdown(const uint8_t* s, int n, const int32_t* nums, int n_nums, long long target, int* unsupported) {
  bool u = false;
  const int r = parse::countdown_score(s, n, nums, n_nums, target, u);
  *unsupported = u ? 1 : 0;
  return r;
}

// ---- MT19937 streams ---------------------------------------------------------------------------------------------
// out_f[0..k): random.random(); out_u[0..k): random._randbelow(n); out_np[0..k): np.random.randint(n) after seed(seed)
void hs_mt(uint64_t seed, uint32_t n, int k, double* out_f, uint32_t* out_u, uint32_t* out_np) {
  uint32_t a[mt::N], b[mt::N];
  mt::State py{a, 0, 1}, np{b, 0, 1};
  mt::seed_python(py, seed);
  mt::seed_numpy(np, (uint32_t)seed);
  for (int i = 0; i < k; ++i) out_f[i] = mt::py_random(py);
  for (int i = 0; i < k; ++i) out_u[i] = mt::py_randbelow(py, n);
  for (int i = 0; i < k; ++i) out_np[i] = mt::np_randint(np, n);
}

// ---- Sokoban generator -------------------------------------------------------------------------------------------
// force_hash bit 0: hash-table visited set instead of the bitmap; bit 1: reverse_play_fast instead of reverse_play
int hs_sokoban_generate(int H, int W, int nb, int search_depth, int num_gen_steps, uint64_t seed, int force_hash,
                        uint8_t* fixed, uint8_t* state, int* explored) {
  const bool fast = force_hash & 2;
  force_hash &= 1;
  sgen::Geometry g{H, W, nb, 0};
  int cells = (H - 2) * (W - 2);
  while ((1 << g.cell_bits) < cells) ++g.cell_bits;
  uint32_t mt_py[mt::N], mt_np[mt::N];
  sgen::Work w;
  w.mt_py = mt_py;
  w.mt_np = mt_np;
  w.stack = (sgen::Frame*)malloc(sizeof(sgen::Frame) * (size_t)(search_depth + 1));
  w.scratch = (uint8_t*)malloc((size_t)2 * H * W);
  const int kb = sgen::key_bits(g);
  if (kb <= 26 && !force_hash) {
    w.vis.mode = 0;
    w.vis_words = ((uint64_t)1 << kb) / 32 + 1;
    w.vis.mem = (uint32_t*)malloc(w.vis_words * 4 + 8);
    w.vis.mem64 = (uint64_t*)w.vis.mem;  // the fast search probes the bitmap 64 bits at a time
  } else {
    w.vis.mode = 1;
    w.vis_words = (uint64_t)1 << 20;
    w.vis.mask = (uint32_t)(w.vis_words - 1);
    w.vis.mem64 = (uint64_t*)malloc(w.vis_words * 8);
    w.vis.mem = nullptr;
  }
  const int st = fast ? sgen::generate_room_fast(g, search_depth, num_gen_steps, seed, w, fixed, state, explored)
                      : sgen::generate_room(g, search_depth, num_gen_steps, seed, w, fixed, state, explored);
  free(w.stack);
  free(w.scratch);
  free(w.vis.mode == 0 ? (void*)w.vis.mem : (void*)w.vis.mem64);
  return st;
}

// ---- text scoring -----------------------------------------------------------------------------------------------
int hs_text_score(int mode, const uint8_t* pred, int np, const uint8_t* label, int nl) { return score::score_pair(mode, pred, np, label, nl); }

// ---- Sokoban solver (get_shortest_action_path) -----------------------------------------------------------------------
// returns the status; path_out needs max_depth + 1 bytes
int hs_sokoban_solve(const uint8_t* fixed, const uint8_t* state, int H, int W, int max_depth, uint32_t max_states, uint8_t* path_out,
                     int* len_out, int* explored_out) {
  uint32_t cap = 2;
  while (cap < 2 * max_states) cap <<= 1;
  uint64_t* table = (uint64_t*)malloc(sizeof(uint64_t) * cap);
  uint64_t* queue = (uint64_t*)malloc(sizeof(uint64_t) * max_states);
  uint32_t* parent = (uint32_t*)malloc(sizeof(uint32_t) * max_states);
  ssolve::Workspace ws{table, queue, parent, cap, max_states, 1u};
  const int st = ssolve::solve([&](int cell) { return (int)fixed[cell]; }, state, H, W, max_depth, ws, path_out, 1, len_out, explored_out);
  free(table);
  free(queue);
  free(parent);
  return st;
}

}  // extern "C"

// Tiles of the bulk-staged advantage scans (scan_tiles.cuh) over a [B, T] tensor with B * T % 16 == 0: cover[w] counts how often
// flat word w is written by a tile's store plan (bulk range or edge words); returns the number of broken invariants.
extern "C" int hs_scan_tiles(int B, int T, int32_t* cover, int32_t* n_tiles) {
  using namespace scan_tiles;
  const int64_t total = (int64_t)B * T;
  int bad = 0, tiles = 0;
  for (int row = 0; row < B; ++row) {
    const int nc = row_chunks(row, T);
    const int64_t lo = (int64_t)row * T, hi = lo + T;
    int64_t expect_next = -1;  // chunks run right to left and are contiguous
    for (int c = nc - 1; c >= 0; --c) {
      const TileGeom g = geom(ScanTile{row, c}, T);
      ++tiles;
      if (g.cw0 % 4 != 0 || g.nw % 4 != 0 || g.nw <= 0 || g.nw > BT_W) ++bad;
      if (g.cw0 < 0 || g.cw0 + g.nw > total) ++bad;                       // staged words inside the tensor
      if (g.cw0 + g.rel_lo != lo || g.cw0 + g.rel_hi != hi) ++bad;         // the row, seen from the tile
      if (g.right_end != (c == nc - 1)) ++bad;
      if (expect_next >= 0 && g.cw0 + g.nw != expect_next) ++bad;
      expect_next = g.cw0;
      if (g.mshift != (int)(g.cw0 % 16) || g.mshift % 4 != 0) ++bad;
      int64_t mb0, mb1;
      mask_window(g, mb0, mb1);
      if (mb0 % 16 != 0 || mb1 % 16 != 0 || mb0 < 0 || mb1 > total || mb0 > g.cw0 || mb1 < g.cw0 + g.nw || mb1 - mb0 > BT_W + 48) ++bad;
      const StorePlan p = store_plan(g);
      if (p.bulk) {
        if ((g.cw0 + p.rs0) % 4 != 0 || (p.rs1 - p.rs0) % 4 != 0) ++bad;   // 16-byte aligned source and destination, whole groups
        if (g.cw0 + p.rs0 < lo || g.cw0 + p.rs1 > hi) ++bad;               // never a word of another row
        for (int w = p.rs0; w < p.rs1; ++w) ++cover[g.cw0 + w];
        if (p.rs0 - p.ra0 > 3 || p.ra1 - p.rs1 > 3) ++bad;                 // at most three edge words on either side
        for (int w = p.rs1; w < p.ra1; ++w) ++cover[g.cw0 + w];
      } else if (p.ra1 - p.ra0 > 8) {
        ++bad;                                                             // eight lanes store the words of a tile without a whole group
      }
      for (int w = p.ra0; w < p.left_end; ++w) ++cover[g.cw0 + w];
    }
    if (expect_next > lo) ++bad;  // the leftmost tile starts at or below the row's first word
  }
  *n_tiles = tiles;
  return bad;
}
