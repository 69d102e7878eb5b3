// Tile geometry of the bulk-staged advantage scans (advantage.cu: gae_scan_bulk_kernel, bilevel_scan_bulk_kernel,
// returns_scan_bulk_kernel), as pure host/device functions: tests/test_hostsim.py runs them on the host and checks that the
// tiles of a [B, T] tensor cover every token exactly once with aligned bulk ranges.
//
// Bulk copies need 16-byte-aligned addresses and a row of an odd T starts anywhere, so a tile is cut in ALIGNED FLAT WORDS of
// the flat [B * T] fp32 tensors, not in tokens: tile (row, c) covers the words [al_lo + c * BT_W, ...), al_lo = the row's first
// word rounded down to a multiple of four.  The mask bytes of those words sit at the same flat index.
#pragma once
#include <stdint.h>

#ifndef RAGEN_HD
#ifdef __CUDACC__
#define RAGEN_HD __host__ __device__ __forceinline__
#else
#define RAGEN_HD inline
#endif
#endif

namespace scan_tiles {

constexpr int BT_T = 288;          // threads (9 warps)
constexpr int BT_K = 8;            // words per thread
constexpr int BT_W = BT_T * BT_K;  // words per tile: a row of T = 2047 fits one tile whatever its alignment
constexpr int BT_NW = BT_T / 32;   // warps

struct ScanTile {  // which (row, chunk) a CTA works on; chunks of a row run right to left
  int row, c;
};

struct TileGeom {
  int64_t cw0;         // first flat word of the tile (a multiple of four)
  int nw;              // staged words (a multiple of four)
  int rel_lo, rel_hi;  // the row, in words of this tile: word w of the tile belongs to the row iff rel_lo <= w < rel_hi
  int mshift;          // the tile's first mask byte sits at stage.m[mshift] (0, 4, 8 or 12)
  bool right_end;      // the tile holds the row's last token
};

// the flat word index of a row's first token is 64-bit; everything relative to the tile fits an int (|.| <= T + BT_W)
RAGEN_HD int row_chunks(int row, int T) {
  const int off = (int)(((int64_t)row * T) & 3);  // words between the 16-byte boundary below the row and its first token
  return (((off + T + 3) & ~3) + BT_W - 1) / BT_W;
}
RAGEN_HD TileGeom geom(ScanTile t, int T) {
  const int64_t lo = (int64_t)t.row * T;
  const int off = (int)(lo & 3);
  const int span = (off + T + 3) & ~3;  // aligned words that cover the row
  const int c0 = t.c * BT_W;            // the tile's first word, counted from the row's aligned start
  TileGeom g;
  g.cw0 = (lo - off) + c0;
  g.right_end = c0 + BT_W >= span;
  g.nw = g.right_end ? span - c0 : BT_W;
  g.rel_lo = off - c0;
  g.rel_hi = g.rel_lo + T;
  g.mshift = (int)(g.cw0 & 15);
  return g;
}
// the mask bytes staged with the tile: [mb0, mb1), both multiples of 16 (B * T % 16 == 0 keeps mb1 inside the tensor)
RAGEN_HD void mask_window(const TileGeom& g, int64_t& mb0, int64_t& mb1) {
  mb0 = g.cw0 & ~(int64_t)15;
  mb1 = (g.cw0 + g.nw + 15) & ~(int64_t)15;
}
// how a tile's results leave: words [rs0, rs1) of the tile by bulk store when bulk (whole 16-byte groups inside the row), the
// row's words left of them [ra0, left_end) and right of them [rs1, ra1) one by one (up to 8 and up to 3)
struct StorePlan {
  int ra0, ra1, rs0, rs1, left_end;
  bool bulk;
};
RAGEN_HD StorePlan store_plan(const TileGeom& g) {
  StorePlan p;
  p.ra0 = g.rel_lo > 0 ? g.rel_lo : 0;
  p.ra1 = g.rel_hi < g.nw ? g.rel_hi : g.nw;
  const int g0 = (g.rel_lo + 3) & ~3, g1 = g.rel_hi & ~3;
  p.rs0 = g0 > 0 ? g0 : 0;
  p.rs1 = g1 < g.nw ? g1 : g.nw;
  p.bulk = p.rs1 > p.rs0;
  p.left_end = p.bulk ? p.rs0 : p.ra1;
  return p;
}

}  // namespace scan_tiles
