// Sokoban step + reward + text render, fused (sm_100a).
//
// Replaces, per env and per action: SokobanEnv.step (ragen/env/sokoban/env.py:44-51) ->
// gym_sokoban SokobanEnv.step/_push/_move/_calc_reward/_check_if_done, followed by the text branch
// of SokobanEnv.render (env.py:53-57); and per turn the action loop of
// EnvStateManager._execute_actions (ragen/llm_agent/es_manager.py:116-128).
//
// Layout: room_state is struct-of-arrays uint8 [n][H*W] in HBM.  A CTA owns a tile of TE consecutive
// envs whose bytes are CONTIGUOUS in HBM, so the tile moves HBM<->shared memory as 1-D TMA bulk
// copies (state in; state out, glyph rows out).  One thread steps one env inside shared memory
// (per-env stride H*W bytes = an odd number of 32-bit words for 6x6 and 10x10: bank-conflict free);
// the glyph rows are rendered cooperatively, one 32-bit word (4 glyphs) per thread per iteration.
//
// room_fixed is NOT read here.  In every room the reference can produce, room_fixed is a function of
// room_state plus one bit: cells 0/1/2 are themselves, 3 (box on target) sits on a 2, 4 (box) sits on
// a 1, and the cell under the player is carried per env in EnvMeta.under.  ragen_sokoban_adopt checks
// that consistency once at reset and rejects rooms that violate it, so dropping the read is exact
// and saves H*W bytes of HBM traffic per env-step.
//
// HBM-bound by design: nothing here is a contraction, tensor cores are not used.
#include <stdlib.h>

#include "common.cuh"

namespace {

constexpr int TE = 128;  // envs per CTA tile == threads per CTA

struct __align__(8) EnvMeta {  // ragen_b200.h: one 8-byte record per env (int16 meta[n][4])
  int16_t player;     // cell index r*W+c
  uint16_t steps;     // num_env_steps
  int8_t on_target;   // boxes_on_target as the reference stores it (0 after reset, env.py:37)
  uint8_t under;      // room_fixed[player] == 2
  int8_t actual;      // num_boxes - #uncovered targets of the current board (what _calc_reward recounts)
  uint8_t pad;
};
static_assert(sizeof(EnvMeta) == 8, "EnvMeta is one 64-bit word");

struct StepParams {
  int64_t n;
  int H, W, num_boxes, max_steps, K;
  const uint8_t* actions;
  const uint8_t* n_actions;
  uint8_t* state;
  EnvMeta* meta;
  double* acc_reward;
  uint8_t* done;
  uint8_t* flags;
  uint8_t* n_executed;
  uint8_t* obs;
  uint8_t* delta;  // [n][K][4] or NULL: per-action delta records instead of the four result arrays (ragen_sokoban_step_delta)
  int delta1;      // the records are ONE byte each, [n][K] (ragen_sokoban_step_delta1): the info byte alone, bit 7 = a box moved
};

__host__ __device__ inline int round16(int x) { return (x + 15) & ~15; }

// cooperative copy for tiles whose byte count is not a multiple of 16 (only the last tile)
__device__ __forceinline__ void copy_bytes(uint8_t* dst, const uint8_t* src, int bytes) {
  for (int i = threadIdx.x; i < bytes; i += blockDim.x) dst[i] = src[i];
}

// The glyph rows of a tile are ONE uniform byte stream: every env is H rows of W cells + 1 separator,
// and consecutive envs have consecutive cells, so output byte Q maps to cell Q - Q/(W+1) unless
// Q % (W+1) == W (separator, code 7).  glyph_word builds the 4 output bytes Q0..Q0+3 from 4 consecutive cells
// (unaligned 32-bit window = two aligned LDS + funnel shift) and splices the separator in.
__device__ __forceinline__ uint32_t glyph_word(const uint32_t* sw, int Q0, int W) {
  const int RW = W + 1;
  const int r0 = Q0 / RW;
  const int c0 = Q0 - r0 * RW;
  const int s0 = Q0 - r0;
  const int a = s0 >> 2;
  const uint32_t x = __funnelshift_r(sw[a], sw[a + 1], (s0 & 3) << 3);
  const int p = W - c0;  // byte of this word that is a separator (if < 4)
  uint32_t out = x;
  if (p < 4) {
    const uint32_t low = (1u << (p << 3)) - 1u;                        // bytes below p
    const uint32_t high = (p == 3) ? 0u : ~((1u << ((p + 1) << 3)) - 1u);  // bytes above p
    out = (x & low) | (7u << (p << 3)) | ((x << 8) & high);
  }
  return out;
}
// 4 glyph codes (< 16 each, one per byte) -> 2 bytes of two 4-bit codes, low nibble first
__device__ __forceinline__ uint32_t pack_nibbles(uint32_t w) {
  const uint32_t t = w | (w >> 4);
  return (t & 0xffu) | ((t >> 8) & 0xff00u);
}
// PACKED: the same stream with two codes per byte (glyph_bytes is then half the number of codes): half the bytes to
// write to HBM and to send to the host, which turns a byte into its two glyphs with the same one-pass translate.
template <int CW, bool PACKED>
__device__ __forceinline__ void render_tile(const uint8_t* s_state, uint8_t* s_obs, int obs_bytes, int rW) {
  const int W = CW ? CW : rW;
  const uint32_t* sw = reinterpret_cast<const uint32_t*>(s_state);
  uint32_t* ow = reinterpret_cast<uint32_t*>(s_obs);
  const int words = (obs_bytes + 3) >> 2;
  if (PACKED && W >= 3) {
    // 8 codes per output word; a row is W + 1 >= 4 codes, so a word holds at most two separators: pack the 8 CELLS that
    // follow the word's first cell to nibbles, then splice the separator nibble(s) in (cells pushed past the word belong
    // to the next one).  (row, column) of the word's first code advance by a constant per iteration, so the division
    // happens once per thread.
    const int RW = W + 1;
    int Q0 = threadIdx.x << 3;
    int r0 = Q0 / RW, c0 = Q0 - r0 * RW;
    const int dr = (TE * 8) / RW, dc = (TE * 8) - dr * RW;
    for (int j = threadIdx.x; j < words; j += TE) {
      const int s0 = Q0 - r0;  // first cell of the window
      const int a = s0 >> 2, sh = (s0 & 3) << 3;
      const uint32_t w0 = sw[a], w1 = sw[a + 1], w2 = sw[a + 2];
      const uint32_t nib = pack_nibbles(__funnelshift_r(w0, w1, sh)) | (pack_nibbles(__funnelshift_r(w1, w2, sh)) << 16);
      uint32_t out = nib;
      for (int q = W - c0; q < 8; q += RW) {  // code(s) of this word that are separators
        const uint32_t low = (1u << (q << 2)) - 1u;
        const uint32_t high = (q == 7) ? 0u : ~((1u << ((q + 1) << 2)) - 1u);
        out = (out & low) | (7u << (q << 2)) | ((out << 4) & high);
      }
      ow[j] = out;
      Q0 += TE * 8;
      r0 += dr;
      c0 += dc;
      if (c0 >= RW) {
        c0 -= RW;
        ++r0;
      }
    }
  } else {
    for (int j = threadIdx.x; j < words; j += TE) {
      if (PACKED) {
        const uint32_t lo = pack_nibbles(glyph_word(sw, j << 3, W)), hi = pack_nibbles(glyph_word(sw, (j << 3) + 4, W));
        ow[j] = lo | (hi << 16);
      } else {
        ow[j] = glyph_word(sw, j << 2, W);
      }
    }
  }
}

// An env's B packed bytes start at e*B in the tile, word-aligned only for some envs: funnel-shift the NOW-word stream by
// (e*B)&3 bytes, store the interior as whole words and the first / last partial words as bytes.
template <int B, int NOW>
__device__ __forceinline__ void store_env_stream(const uint32_t (&out)[NOW], uint8_t* s_obs, int e) {
  constexpr bool ALIGNED = (B % 4 == 0);
  const int a = ALIGNED ? 0 : ((e * B) & 3);
  uint8_t* dst = s_obs + e * B - a;  // word-aligned; this env owns bytes [a, a + B) from here
  uint32_t prev = 0u;
#pragma unroll
  for (int k = 0; k <= NOW; ++k) {
    if (4 * k >= B + 3) break;
    const uint32_t cur = (k < NOW) ? out[k] : 0u;
    const uint32_t w = ALIGNED ? cur : __funnelshift_l(prev, cur, a << 3);
    prev = cur;
    if ((ALIGNED || k >= 1) && 4 * k + 4 <= B) {
      *reinterpret_cast<uint32_t*>(dst + 4 * k) = w;  // whole word inside [a, a + B) for every a
    } else {
#pragma unroll
      for (int b = 0; b < 4; ++b) {
        const int at = 4 * k + b;
        if (at >= a && at < a + B) dst[at] = (uint8_t)(w >> (8 * b));
      }
    }
  }
}

// the room's words -> a stream of cell nibbles, 8 cells per word (two shift-ors and one byte permute per word)
template <int HW>
__device__ __forceinline__ void cell_nibbles(const uint32_t* sw, uint32_t (&cw)[(HW + 7) / 8 + 1]) {
  constexpr int NW = HW / 4, NCW = (HW + 7) / 8;
#pragma unroll
  for (int j = 0; j < NCW; ++j) {
    const uint32_t x0 = sw[2 * j];
    const uint32_t x1 = (2 * j + 1 < NW) ? sw[2 * j + 1] : 0u;
    cw[j] = __byte_perm(x0 | (x0 >> 4), x1 | (x1 >> 4), 0x6420);
  }
  cw[NCW] = 0u;
}

// Thread-per-env packed render for the compile-time room shapes (rooms that are whole 32-bit words: 6x6, 8x8, 10x10).
// The thread that stepped env e turns ITS room into the env's G/2 packed bytes with every index known at compile time:
//   1. the room's words -> a stream of cell nibbles (two shifts-and-ors and one byte permute per 8 cells);
//   2. output word k (8 codes) = for each of the <= 2 rows it touches, a funnel-shifted window of the cell stream under a
//      constant mask, or-ed with a constant holding the separators (code 7) - row r's cells sit r nibbles further along in
//      the output than in the cell stream;
//   3. the env's bytes start at e*G/2 in the tile, which is word-aligned only for every fourth env (55 or 21 bytes per env), so
//      the stream is funnel-shifted by (e*G/2)&3 bytes and the interior words are stored whole, the first and last as bytes.
// ~5x fewer instructions per env than the cooperative word loop below (which recomputes row/column/alignment per word).
template <int CH, int CW>
__device__ __forceinline__ void render_env_packed(const uint8_t* s_state, uint8_t* s_obs, int e) {
  constexpr int HW = CH * CW, RW = CW + 1, G = CH * RW;
  constexpr int B = G / 2;           // packed bytes per env
  constexpr int NW = HW / 4;         // room words per env
  constexpr int NCW = (HW + 7) / 8;  // words of the cell-nibble stream
  constexpr int NOW = (G + 7) / 8;   // words of the output stream before the alignment shift
  static_assert(HW % 4 == 0 && G % 2 == 0 && RW >= 7, "shape not supported by the per-env render");
  uint32_t cw[NCW + 1];
  cell_nibbles<HW>(reinterpret_cast<const uint32_t*>(s_state) + e * NW, cw);
  uint32_t out[NOW];
#pragma unroll
  for (int k = 0; k < NOW; ++k) {
    uint32_t acc = 0u;
#pragma unroll
    for (int q = 0; q < 8; ++q)  // separators of this word (constant-folded)
      if (8 * k + q < G && (8 * k + q) % RW == CW) acc |= 7u << (4 * q);
    const int r0 = (8 * k) / RW;
#pragma unroll
    for (int dr = 0; dr < 3; ++dr) {
      const int r = r0 + dr;
      const int lo = (8 * k > r * RW) ? 8 * k : r * RW;                       // output nibbles of row r's cells ...
      const int hi = (8 * k + 8 < r * RW + CW) ? 8 * k + 8 : r * RW + CW;    // ... that fall into word k
      if (r < CH && lo < hi) {
        const int s = 8 * k - r;  // cell nibble under output nibble 8k
        uint32_t win;
        if (s < 0) win = cw[0] << (4 * (-s));
        else if ((s & 7) == 0) win = cw[s >> 3];
        else win = __funnelshift_r(cw[s >> 3], cw[(s >> 3) + 1], 4 * (s & 7));
        const uint32_t ones = (hi - lo == 8) ? 0xffffffffu : (((1u << (4 * (hi - lo))) - 1u) << (4 * (lo - 8 * k)));
        acc |= win & ones;
      }
    }
    out[k] = acc;
  }
  store_env_stream<B, NOW>(out, s_obs, e);
}

// Three-bit glyph rows (ragen_sokoban_step_packed3): the H*W cell codes (0..6 fit three bits), row-major, NO separators (the
// host knows W), little-endian bit stream padded to an even number of bytes per env - 38 B for 10x10 instead of 55 B of
// nibbles: the format for bulk transport to the host, where the device->host link is what bounds the step.
// (a & m) | (b & ~m) as ONE LOP3 (written in C the two constant masks become two instructions)
__device__ __forceinline__ uint32_t bitsel(uint32_t a, uint32_t b, uint32_t m) {
  uint32_t d;
  asm("lop3.b32 %0, %1, %2, %3, 0xE4;" : "=r"(d) : "r"(a), "r"(b), "r"(m));
  return d;
}

// cw: the env's cell-nibble stream (8 cells per word) -> its three-bit row stream in the glyph tile
template <int CH, int CW>
__device__ __forceinline__ void pack3_store(const uint32_t (&cw)[(CH * CW + 7) / 8 + 1], uint8_t* s_obs, int e) {
  constexpr int HW = CH * CW, NCW = (HW + 7) / 8;
  constexpr int B = ((HW * 3 + 15) / 16) * 2;  // bytes per env
  constexpr int NOW = (B + 3) / 4;
  uint32_t g24[NCW];  // 8 cells -> 24 bits
#pragma unroll
  for (int j = 0; j < NCW; ++j) {
    // Bit selects (one LOP3 each) instead of mask-and-or: bit 3 of every nibble is 0 (codes 0..6), so the bits a select drags
    // along are either zero or sit in positions a later select overwrites.
    const uint32_t x = cw[j];
    const uint32_t pr = bitsel(x, x >> 1, 0x07070707u);  // per byte: cell 2i | cell 2i+1 << 3 in bits 0-5 (bit 7: junk, never picked)
    const uint32_t t1 = bitsel(pr, pr >> 2, 0x0000003fu);  // bits 0-5 byte 0, bits 6.. from byte 1 on
    const uint32_t t2 = bitsel(t1, pr >> 4, 0x00000fffu);  // bits 12.. from byte 2 on
    g24[j] = bitsel(t2, pr >> 6, 0x0003ffffu);             // bits 18-23 byte 3; bits 24-31 = pr bits 30.. = 0
  }
  uint32_t out[NOW];
#pragma unroll
  for (int k = 0; k < NOW; ++k) {
    uint32_t acc = 0u;
#pragma unroll
    for (int j = 0; j < NCW; ++j) {
      const int sh = 24 * j - 32 * k;  // bit position of group j in word k
      if (sh > -24 && sh < 32) acc |= (sh >= 0) ? (g24[j] << sh) : (g24[j] >> (-sh));
    }
    out[k] = acc;
  }
  store_env_stream<B, NOW>(out, s_obs, e);
}

template <int CH, int CW>
__device__ __forceinline__ void render_env_packed3(const uint8_t* s_state, uint8_t* s_obs, int e) {
  constexpr int HW = CH * CW, NCW = (HW + 7) / 8;
  static_assert(HW % 4 == 0, "shape not supported by the per-env render");
  uint32_t cw[NCW + 1];
  cell_nibbles<HW>(reinterpret_cast<const uint32_t*>(s_state) + e * (HW / 4), cw);
  pack3_store<CH, CW>(cw, s_obs, e);
}

// Rooms kept as NIBBLES in HBM (ragen_sokoban_step_compact): a room's bytes ARE the cell-nibble stream the three-bit render
// starts from, so env e's H*W/2 bytes of the staged tile go straight into registers.  Rows of 50 (18) bytes are only 2-byte
// aligned: aligned words + one funnel shift; the cells past H*W in the last word belong to the next env and are masked off.
template <int HW>
__device__ __forceinline__ void load_env_nibbles(const uint8_t* s_tile4, int e, uint32_t (&cw)[(HW + 7) / 8 + 1]) {
  constexpr int B = HW / 2, NCW = (HW + 7) / 8;
  constexpr bool ALIGNED = (B % 4 == 0);
  const int a = ALIGNED ? 0 : ((e * B) & 3);
  const uint32_t* src = reinterpret_cast<const uint32_t*>(s_tile4 + e * B - a);
  uint32_t raw[NCW + 1];
#pragma unroll
  for (int k = 0; k <= NCW; ++k) raw[k] = (ALIGNED && k == NCW) ? 0u : src[k];  // may read one word past the row: the tile is padded
#pragma unroll
  for (int j = 0; j < NCW; ++j) cw[j] = ALIGNED ? raw[j] : __funnelshift_r(raw[j], raw[j + 1], a << 3);
  if (HW % 8 != 0) cw[NCW - 1] &= (1u << (4 * (HW % 8))) - 1u;
  cw[NCW] = 0u;
}

// PACK: glyph-row format - 0 one code per byte, 1 two 4-bit codes per byte, 2 three bits per cell without separators
template <int CH, int CW, bool RENDER_ONLY, int PACK>
__global__ void __launch_bounds__(TE, (PACK >= 3 && !RENDER_ONLY) ? 16 : 1) sokoban_tile_kernel(const StepParams p) {
  // PACK 3 = "compact": three-bit rows out AND rooms kept as nibbles in HBM (p.state is then uint8 [n][H*W/2]); the tile is
  // staged, stepped and rendered in that form
  // PACK 4 = "glyph rooms": the rooms rest in HBM as THREE-BIT GLYPH CODES (0..5 as room_state, 6 = player on a target) - the room
  // IS its own rendered observation, so a step reads and writes 38 B per 10x10 env and there is nothing left to render
  constexpr bool PACKED = (PACK == 1), PACK3 = (PACK == 2 || PACK == 3), STATE4 = (PACK == 3), STATE3 = (PACK == 4);
  const int H = CH ? CH : p.H, W = CW ? CW : p.W;
  const int HW = H * W, G = H * (W + 1);
  const int B3 = ((HW * 3 + 15) / 16) * 2;  // bytes of a three-bit row stream
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ __align__(8) uint64_t bar;
  // the next kernel of the stream may be scheduled as soon as every CTA of this one has started; this one touches global memory
  // only after the kernel before it has completed (launch_tiles)
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  asm volatile("griddepcontrol.wait;" ::: "memory");

  uint8_t* s_state = smem;
  uint8_t* s_obs = s_state + round16(STATE3 ? TE * B3 : STATE4 ? (TE * HW) >> 1 : TE * HW) + 16;  // +16: the render window may read one word past the tile

  const int64_t env0 = (int64_t)blockIdx.x * TE;
  const int cnt = (int)min((int64_t)TE, p.n - env0);
  const int room_bytes = STATE3 ? cnt * B3 : STATE4 ? (cnt * HW) >> 1 : cnt * HW;  // bytes of the tile in HBM
  const int obs_bytes = PACK3 ? cnt * B3 : PACKED ? (cnt * G) >> 1 : cnt * G;  // PACKED needs an even G (checked at launch)
  const bool bulk_rooms = (room_bytes & 15) == 0;
  const bool bulk_obs = (obs_bytes & 15) == 0;
  uint8_t* g_state = p.state + (STATE3 ? env0 * B3 : STATE4 ? (env0 * HW) >> 1 : env0 * HW);

  // ---- stage the tile: HBM -> shared memory -------------------------------------------------
  if (bulk_rooms) {
    if (threadIdx.x == 0) {
      mbar_init(&bar, 1);
      fence_mbar_init();
      mbar_expect_tx(&bar, (uint32_t)room_bytes);
      tma_load_1d(s_state, g_state, (uint32_t)room_bytes, &bar);
    }
  } else {
    copy_bytes(s_state, g_state, room_bytes);
  }

  // per-env scalars travel in registers while the tile is in flight
  const int e = threadIdx.x;
  const int64_t g = env0 + e;
  const bool mine = e < cnt;
  EnvMeta m{};
  uint32_t acts = 0;  // first 4 actions packed (K <= 4 is the common case); longer turns re-read global
  int na = 0;
  if (mine) {
    if (!RENDER_ONLY) {
      // the record, the action count and the action bytes are three INDEPENDENT loads (one HBM round trip, not three in a row):
      // the first min(K, 4) action bytes are fetched whatever n_actions and the record say, and the ones past `na` are never looked at
      const uint8_t* act = p.actions + g * p.K;
      const int nab = p.n_actions ? (int)p.n_actions[g] : p.K;
      // straight-line predicated byte loads (a counted loop here compiles to three peeled gather loops)
      if (p.K > 0) acts = act[0];
      if (p.K > 1) acts |= (uint32_t)act[1] << 8;
      if (p.K > 2) acts |= (uint32_t)act[2] << 16;
      if (p.K > 3) acts |= (uint32_t)act[3] << 24;
      m = p.meta[g];
      na = min(nab, p.K);
      if (m.player < 0) na = 0;  // room rejected by adopt (no player): leave it alone
    } else {
      m = p.meta[g];
    }
  }
  __syncthreads();  // barrier initialised (bulk) / tile copied (fallback) before anyone reads
  if (bulk_rooms) mbar_wait(&bar, 0);

  // ---- one thread steps one env -----------------------------------------------------------------
  int pos = m.player;
  int under = m.under;
  if (!RENDER_ONLY && mine) {
    uint8_t* S = s_state + (STATE3 ? e * B3 : STATE4 ? (e * HW) >> 1 : e * HW);
    // cell i of the own room: a byte, or (STATE4) nibble i&1 of byte i>>1, or (STATE3) bits 3i..3i+2 - rooms are whole bytes,
    // nobody else touches them (a three-bit cell straddles into the next byte only inside its own room)
    auto cell = [&](int i) -> int {
      if (STATE3) {
        const int bit = 3 * i, b = bit >> 3, sh = bit & 7;
        unsigned w = S[b];
        if (sh > 5) w |= (unsigned)S[b + 1] << 8;
        return (w >> sh) & 7;
      }
      return STATE4 ? (S[i >> 1] >> ((i & 1) << 2)) & 15 : S[i];
    };
    auto put = [&](int i, int v) {
      if (STATE3) {
        const int bit = 3 * i, b = bit >> 3, sh = bit & 7;
        if (sh <= 5) {
          S[b] = (uint8_t)((S[b] & ~(7u << sh)) | ((unsigned)v << sh));
        } else {
          unsigned w = (unsigned)S[b] | ((unsigned)S[b + 1] << 8);
          w = (w & ~(7u << sh)) | ((unsigned)v << sh);
          S[b] = (uint8_t)w;
          S[b + 1] = (uint8_t)(w >> 8);
        }
      } else if (STATE4) {
        const int o = S[i >> 1];
        S[i >> 1] = (uint8_t)((i & 1) ? (o & 0x0f) | (v << 4) : (o & 0xf0) | v);
      } else {
        S[i] = (uint8_t)v;
      }
    };
    int steps = m.steps;
    int on = m.on_target;    // the reference's stored count
    int actual = m.actual;   // the true count of the board; differs from `on` only right after a reset
    double acc = 0.0;
    int fl = 0, nexec = 0, dn = 0;
    uint32_t* drec = p.delta ? reinterpret_cast<uint32_t*>(p.delta) + g * p.K : nullptr;
    for (int k = 0; k < na; ++k) {
      const int a = (k < 4) ? (int)((acts >> (8 * k)) & 0xff) : (int)p.actions[g * p.K + k];
      const int before = pos;
      int box_to = 0xff;
      ++steps;
      int delta = 0;  // change of the on-target count caused by this action
      if (a >= 1 && a <= 4) {
        const int mag = (a <= 2) ? W : 1;          // 1 up, 2 down, 3 left, 4 right (config.py:12) without a jump table
        const int d = (a & 1) ? -mag : mag;
        const int ahead = pos + d;  // in-grid: the border is wall (checked by adopt), the player never stands on it
        const int sa = cell(ahead);
        bool go = (sa == 1) | (sa == 2);
        if (sa == 3 || sa == 4) {
          const int beyond = ahead + d;
          if (beyond >= 0 && beyond < HW) {
            const int sb = cell(beyond);
            if (sb == 1 || sb == 2) {
              put(beyond, (sb == 2) ? 3 : 4);
              delta = (int)(sb == 2) - (int)(sa == 3);
              go = true;
              box_to = beyond;
            }
          }
        }
        if (go) {
          const int now_under = (sa == 2) | (sa == 3);
          put(ahead, (STATE3 && now_under) ? 6 : 5);  // glyph rooms hold the player-on-target code themselves
          put(pos, under ? 2 : 1);  // == room_fixed[pos]
          under = now_under;
          pos = ahead;
        }
      }
      // reward (gym_sokoban _calc_reward): the reference recounts the board and compares with the
      // count stored by the previous step; the recount is the running count plus this action's delta
      const int now_on = actual + delta;
      double rew = -0.1;
      const bool now_on_rose = now_on > on, now_on_fell = now_on < on;
      if (now_on_rose) rew += 1.0;
      else if (now_on_fell) rew += -1.0;
      const bool won = (now_on == p.num_boxes);
      if (won) rew += 10.0;
      on = actual = now_on;
      acc += rew;
      ++nexec;
      fl = (before != pos ? 1 : 0) | (won ? 2 : 0) | 4;
      dn = (won || steps == p.max_steps) ? 1 : 0;
      if (drec) {
        // [player before | player after | box destination or 0xff | reward code (0: -0.1, 1: +1, 2: -1, then +10 when won: 3 with
        // the +1, 4 without) | done << 3 | effective << 4 | success << 5 | executed << 6]
        const int code = won ? (now_on_rose ? 3 : 4) : now_on_rose ? 1 : now_on_fell ? 2 : 0;
        const uint32_t info = (uint32_t)(code | (dn << 3) | ((before != pos) << 4) | ((int)won << 5) | (1 << 6));
        if (p.delta1)  // the consumer knows the action it sent and mirrors the player: "moved" and "a box moved" say the rest
          p.delta[g * p.K + k] = (uint8_t)(info | (box_to != 0xff ? 0x80u : 0u));
        else
          drec[k] = (uint32_t)before | ((uint32_t)pos << 8) | ((uint32_t)box_to << 16) | (info << 24);
      }
      if (dn) break;
    }
    if (drec)
      for (int k = nexec; k < p.K; ++k) {  // not executed
        if (p.delta1) p.delta[g * p.K + k] = 0;
        else drec[k] = 0u;
      }
    m.player = (int16_t)pos;
    m.steps = (uint16_t)steps;
    m.on_target = (int8_t)on;
    m.actual = (int8_t)actual;
    m.under = (uint8_t)under;
    p.meta[g] = m;
    if (p.acc_reward) {
      p.acc_reward[g] = acc;
      p.done[g] = (uint8_t)dn;
      p.flags[g] = (uint8_t)fl;
      p.n_executed[g] = (uint8_t)nexec;
    }
  }
  // rooms of the compile-time shapes are rendered by the thread that owns them (render_env_packed), straight after its
  // own step: no CTA-wide barrier between step and render, and the player-on-target patch touches only the owner's bytes
  constexpr bool ENV_RENDER =
      PACK != 0 && CH != 0 && CW != 0 && ((CH * CW) % 4 == 0) && (PACK3 || ((CH * (CW + 1)) % 2 == 0)) && CW >= 6;
  if constexpr (ENV_RENDER && STATE4) {
    if (p.obs && mine) {
      // player standing on a target (env.py:55): code 5 -> 6 (xor 3) in the owner's nibble for the duration of the render -
      // the room's bulk store is issued after the barrier below
      const bool fix = under && pos >= 0;
      uint8_t* pb = s_state + ((e * HW) >> 1) + (pos >> 1);
      const uint8_t fx = (uint8_t)(3u << ((pos & 1) << 2));
      if (fix) *pb ^= fx;
      uint32_t cw[(CH * CW + 7) / 8 + 1];
      load_env_nibbles<CH * CW>(s_state, e, cw);
      if (fix) *pb ^= fx;
      pack3_store<CH, CW>(cw, s_obs, e);
    }
  }
  if (ENV_RENDER && PACK3 && !STATE4 && p.obs && mine) {
    // three-bit rows: the player-on-target code 6 is put into the room for the duration of the render (the room's bulk store
    // is issued after the barrier below, and the cell is the owner's)
    const bool patch = under && pos >= 0;
    if (patch) s_state[e * HW + pos] = 6;
    if constexpr (ENV_RENDER && PACK3 && !STATE4) render_env_packed3<CH, CW>(s_state, s_obs, e);
    if (patch) s_state[e * HW + pos] = 5;
  }
  if (ENV_RENDER && PACKED && p.obs && mine) {
    if constexpr (ENV_RENDER && PACKED) render_env_packed<CH, CW>(s_state, s_obs, e);
    if (under) {  // player standing on a target (env.py:55)
      const int q = e * G + pos + pos / W;
      const uint8_t old = s_obs[q >> 1];
      s_obs[q >> 1] = (q & 1) ? (uint8_t)((old & 0x0f) | 0x60) : (uint8_t)((old & 0xf0) | 0x06);
    }
  }
  if (!RENDER_ONLY) {
    fence_async_smem();  // the stepped rooms are read by the bulk store below
    __syncthreads();
    if (bulk_rooms) {
      if (threadIdx.x == 0) {
        tma_store_1d(g_state, s_state, (uint32_t)room_bytes);
        tma_store_commit();
      }
    } else {
      copy_bytes(g_state, s_state, room_bytes);
    }
  }

  // ---- render glyph rows, then shared memory -> HBM -----------------------------------------------
  if (p.obs) {
    if (!ENV_RENDER) {
      render_tile<CW, PACKED>(s_state, s_obs, obs_bytes, p.W);
      __syncthreads();
      if (mine && under) {  // player standing on a target (env.py:55)
        const int q = e * G + pos + pos / W;
        if (PACKED) {  // G is even, so no other env's thread touches this byte
          const uint8_t old = s_obs[q >> 1];
          s_obs[q >> 1] = (q & 1) ? (uint8_t)((old & 0x0f) | 0x60) : (uint8_t)((old & 0xf0) | 0x06);
        } else {
          s_obs[q] = 6;
        }
      }
    }
    if (!ENV_RENDER || RENDER_ONLY) {  // otherwise the barrier before the room store already covers the glyph tile
      fence_async_smem();
      __syncthreads();
    }
    uint8_t* g_obs = p.obs + (PACK3 ? env0 * B3 : PACKED ? (env0 * G) >> 1 : env0 * G);
    if (bulk_obs) {
      if (threadIdx.x == 0) {
        tma_store_1d(g_obs, s_obs, (uint32_t)obs_bytes);
        tma_store_commit();
      }
    } else {
      copy_bytes(g_obs, s_obs, obs_bytes);
    }
  }
  if (threadIdx.x == 0) tma_store_wait_read();  // shared memory must outlive the bulk engine's reads
}

// Post-reset bookkeeping (SokobanEnv.reset, env.py:37-38) + the consistency check that lets the step
// kernel drop room_fixed.  Same tiling as the step kernel: state and fixed tiles are staged through
// shared memory, one thread scans one room.
struct AdoptParams {
  int64_t n;
  int H, W, num_boxes;
  const uint8_t* state;
  const uint8_t* fixed;
  EnvMeta* meta;
  int32_t* status;
  const uint8_t* only;  // [n] or NULL: adopt env i only when only[i] != 0 (slot-wise reset of a shared batch)
  uint8_t* state4;      // [n][H*W/2] or NULL: the room as nibbles (what ragen_sokoban_step_compact keeps), written for adopted envs
  uint8_t* state3;      // [n][B3] or NULL: the room as three-bit glyph codes (ragen_sokoban_step_glyph3), written for adopted envs
};

__global__ void __launch_bounds__(TE) sokoban_adopt_kernel(const AdoptParams p) {
  const int H = p.H, W = p.W, HW = H * W;
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ __align__(8) uint64_t bar;
  uint8_t* s_state = smem;
  uint8_t* s_fixed = smem + round16(TE * HW);
  const int64_t env0 = (int64_t)blockIdx.x * TE;
  const int cnt = (int)min((int64_t)TE, p.n - env0);
  const int room_bytes = cnt * HW;
  if ((room_bytes & 15) == 0) {
    if (threadIdx.x == 0) {
      mbar_init(&bar, 1);
      fence_mbar_init();
      mbar_expect_tx(&bar, 2u * (uint32_t)room_bytes);
      tma_load_1d(s_state, p.state + env0 * HW, (uint32_t)room_bytes, &bar);
      tma_load_1d(s_fixed, p.fixed + env0 * HW, (uint32_t)room_bytes, &bar);
    }
    __syncthreads();
    mbar_wait(&bar, 0);
  } else {
    copy_bytes(s_state, p.state + env0 * HW, room_bytes);
    copy_bytes(s_fixed, p.fixed + env0 * HW, room_bytes);
    __syncthreads();
  }
  const int e = threadIdx.x;
  if (e >= cnt) return;
  if (p.only && !p.only[env0 + e]) return;  // an env in mid-episode keeps its record (steps, stored box count)
  const uint8_t* S = s_state + e * HW;
  const uint8_t* F = s_fixed + e * HW;
  if (p.state4) {  // H*W is even (checked at launch): rooms are whole bytes
    uint8_t* d = p.state4 + ((env0 + e) * HW >> 1);
    for (int i = 0; i < HW; i += 2) d[i >> 1] = (uint8_t)((S[i] & 15) | (S[i + 1] << 4));
  }
  if (p.state3) {  // glyph codes: room_state with the player on a target as 6, three bits per cell, rows padded to even bytes
    const int b3 = ((HW * 3 + 15) / 16) * 2;
    uint8_t* d = p.state3 + (env0 + e) * b3;
    unsigned acc = 0;
    int have = 0, at = 0;
    for (int i = 0; i < HW; ++i) {
      const unsigned code = (S[i] == 5 && F[i] == 2) ? 6u : (unsigned)(S[i] & 7);
      acc |= code << have;
      have += 3;
      while (have >= 8) {
        d[at++] = (uint8_t)acc;
        acc >>= 8;
        have -= 8;
      }
    }
    while (at < b3) {
      d[at++] = (uint8_t)acc;
      acc >>= 8;
    }
  }
  int pos = -1, st = 0, open_targets = 0, under = 0;
  int r = 0, c = 0;
  for (int i = 0; i < HW; ++i) {
    const int s = S[i], f = F[i];
    if (s == 5 && pos < 0) {  // argwhere(room_state == 5)[0]: first in row-major order
      pos = i;
      under = (f == 2);
    }
    const bool border = (r == 0) | (r == H - 1) | (c == 0) | (c == W - 1);
    if (border && f != 0 && st == 0) st = 2;
    // room_fixed must be recoverable from room_state: 0/1/2 are themselves, a box-on-target sits on a
    // target, a plain box on floor, the player on floor or target (true of every generated room)
    const bool ok = (s <= 2) ? (f == s) : (s == 3) ? (f == 2) : (s == 4) ? (f == 1) : (s == 5) ? (f == 1 || f == 2) : false;
    if (!ok && st == 0) st = 3;
    open_targets += (s == 2) | ((s == 5) & (f == 2));
    if (++c == W) {
      c = 0;
      ++r;
    }
  }
  if (pos < 0) st = 1;
  EnvMeta m{};
  m.player = (int16_t)pos;
  m.steps = 0;
  m.on_target = 0;
  m.under = (uint8_t)under;
  m.actual = (int8_t)max(-128, p.num_boxes - open_targets);
  p.meta[env0 + e] = m;
  if (p.status) p.status[env0 + e] = st;
}

size_t tile_smem(int H, int W, int pack) {
  const int b3 = ((H * W * 3 + 15) / 16) * 2;
  if (pack == 4) return (size_t)round16(TE * b3) + 32;  // glyph rooms: the room tile only
  const int glyph_bytes = pack >= 2 ? TE * b3 : pack == 1 ? (TE * H * (W + 1)) >> 1 : TE * H * (W + 1);
  return (size_t)round16(pack == 3 ? TE * H * W / 2 : TE * H * W) + 16 + round16(glyph_bytes) + 16;
}

// Programmatic dependent launch: the tile kernels begin with griddepcontrol.launch_dependents + griddepcontrol.wait, and are
// launched with programmaticStreamSerializationAllowed - the CTAs of a step are scheduled into the SM slots the previous kernel
// of the stream vacates while it drains and sit at the wait until that kernel has completed and flushed (stream order as
// before: nothing is read or written ahead of the wait), so the launch gap and the prologue of a launch run under the tail of
// the one before.  RAGEN_PDL=0 launches the plain way.
static bool pdl_enabled() {
  static const bool on = [] {
    const char* e = getenv("RAGEN_PDL");
    return !(e && e[0] == '0');
  }();
  return on;
}

template <bool RENDER_ONLY, int PACK>
int launch_tiles(ragen_ctx* ctx, const StepParams& p, cudaStream_t stream) {
  RAGEN_REQUIRE(PACK != 1 || ((p.H * (p.W + 1)) & 1) == 0, "packed glyph rows need an even H*(W+1), got %dx%d", p.H, p.W);
  RAGEN_REQUIRE(PACK < 2 || (p.H == p.W && (p.H == 6 || p.H == 8 || p.H == 10)),
                "three-bit glyph rows and nibble rooms are built for 6x6, 8x8 and 10x10 rooms, got %dx%d", p.H, p.W);
  const size_t smem = tile_smem(p.H, p.W, PACK);
  RAGEN_REQUIRE(smem <= 227 * 1024, "room %dx%d does not fit the shared-memory tile (%zu B)", p.H, p.W, smem);
  const unsigned grid = (unsigned)((p.n + TE - 1) / TE);
#define RAGEN_TILE_LAUNCH(CH, CW)                                                                   \
  do {                                                                                              \
    auto kern = sokoban_tile_kernel<CH, CW, RENDER_ONLY, PACK>;                                            \
    RAGEN_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
    cudaLaunchConfig_t cfg = {};                                                                    \
    cfg.gridDim = dim3(grid);                                                                       \
    cfg.blockDim = dim3(TE);                                                                        \
    cfg.dynamicSmemBytes = smem;                                                                    \
    cfg.stream = stream;                                                                            \
    cudaLaunchAttribute attr[1];                                                                    \
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;                                \
    attr[0].val.programmaticStreamSerializationAllowed = 1;                                         \
    cfg.attrs = attr;                                                                               \
    cfg.numAttrs = pdl_enabled() ? 1 : 0;                                                           \
    RAGEN_CUDA(cudaLaunchKernelEx(&cfg, kern, p));                                                  \
  } while (0)
  if (p.H == 6 && p.W == 6) RAGEN_TILE_LAUNCH(6, 6);
  else if (p.H == 8 && p.W == 8) RAGEN_TILE_LAUNCH(8, 8);
  else if (p.H == 10 && p.W == 10) RAGEN_TILE_LAUNCH(10, 10);
  else if constexpr (PACK < 2) RAGEN_TILE_LAUNCH(0, 0);
#undef RAGEN_TILE_LAUNCH
  RAGEN_LAUNCHED(ctx);
  return RAGEN_OK;
}

bool aligned16(const void* p) { return ((uintptr_t)p & 15) == 0; }
bool aligned8(const void* p) { return ((uintptr_t)p & 7) == 0; }

}  // namespace

#define RAGEN_SOKOBAN_SHAPE(n, H, W) \
  RAGEN_REQUIRE(n >= 0 && H >= 3 && W >= 3 && H * W <= 32767, "bad shape n=%lld H=%d W=%d", (long long)n, H, W)

namespace {
int sokoban_step_entry(int pack, ragen_ctx* ctx, int64_t n, int H, int W, int num_boxes, int max_steps, const uint8_t* actions,
                       int K, const uint8_t* n_actions, uint8_t* state, int16_t* meta, double* acc_reward, uint8_t* done,
                       uint8_t* flags, uint8_t* n_executed, uint8_t* obs, void* stream, uint8_t* delta = nullptr, int delta1 = 0) {
  RAGEN_REQUIRE(ctx, "ctx is NULL");
  RAGEN_SOKOBAN_SHAPE(n, H, W);
  RAGEN_REQUIRE(K >= 1 && K <= 255, "K=%d out of range", K);
  RAGEN_REQUIRE(num_boxes >= 0 && num_boxes <= 127, "num_boxes=%d out of range", num_boxes);
  RAGEN_REQUIRE(actions && state && meta, "NULL pointer argument");
  RAGEN_REQUIRE(delta || (acc_reward && done && flags && n_executed), "NULL pointer argument");
  RAGEN_REQUIRE(aligned16(state) && (!obs || aligned16(obs)) && aligned8(meta) && aligned8(acc_reward) &&
                    (delta1 || ((uintptr_t)delta & 3) == 0),
                "state/obs must be 16-byte aligned, meta/acc_reward 8-byte aligned, delta 4-byte aligned");
  RAGEN_REQUIRE(!delta || delta1 || H * W <= 255, "delta records hold cell indices in one byte: H*W <= 255");
  if (n == 0) return RAGEN_OK;
  DeviceGuard guard(ctx->device);
  StepParams p{n, H, W, num_boxes, max_steps, K, actions, n_actions, state, reinterpret_cast<EnvMeta*>(meta),
               acc_reward, done, flags, n_executed, obs, delta, delta1};
  RAGEN_REQUIRE(pack != 4 || !obs, "glyph rooms are their own observation: obs must be NULL");
  return pack == 4   ? launch_tiles<false, 4>(ctx, p, (cudaStream_t)stream)
         : pack == 3 ? launch_tiles<false, 3>(ctx, p, (cudaStream_t)stream)
         : pack == 2 ? launch_tiles<false, 2>(ctx, p, (cudaStream_t)stream)
         : pack == 1 ? launch_tiles<false, 1>(ctx, p, (cudaStream_t)stream)
                     : launch_tiles<false, 0>(ctx, p, (cudaStream_t)stream);
}

int sokoban_render_entry(int pack, ragen_ctx* ctx, int64_t n, int H, int W, const uint8_t* state, const int16_t* meta,
                         uint8_t* obs, void* stream) {
  RAGEN_REQUIRE(ctx, "ctx is NULL");
  RAGEN_SOKOBAN_SHAPE(n, H, W);
  RAGEN_REQUIRE(state && meta && obs, "NULL pointer argument");
  RAGEN_REQUIRE(aligned16(state) && aligned16(obs) && aligned8(meta), "state/obs must be 16-byte aligned, meta 8-byte aligned");
  if (n == 0) return RAGEN_OK;
  DeviceGuard guard(ctx->device);
  StepParams p{};
  p.n = n;
  p.H = H;
  p.W = W;
  p.state = const_cast<uint8_t*>(state);
  p.meta = reinterpret_cast<EnvMeta*>(const_cast<int16_t*>(meta));
  p.obs = obs;
  return pack == 2   ? launch_tiles<true, 2>(ctx, p, (cudaStream_t)stream)
         : pack == 1 ? launch_tiles<true, 1>(ctx, p, (cudaStream_t)stream)
                     : launch_tiles<true, 0>(ctx, p, (cudaStream_t)stream);
}
}  // namespace

extern "C" int ragen_sokoban_step(ragen_ctx* ctx, int64_t n, int H, int W, int num_boxes, int max_steps,
                                  const uint8_t* actions, int K, const uint8_t* n_actions, uint8_t* state, int16_t* meta,
                                  double* acc_reward, uint8_t* done, uint8_t* flags, uint8_t* n_executed, uint8_t* obs,
                                  void* stream) {
  return sokoban_step_entry(0, ctx, n, H, W, num_boxes, max_steps, actions, K, n_actions, state, meta, acc_reward, done, flags,
                            n_executed, obs, stream);
}

extern "C" int ragen_sokoban_step_packed(ragen_ctx* ctx, int64_t n, int H, int W, int num_boxes, int max_steps,
                                         const uint8_t* actions, int K, const uint8_t* n_actions, uint8_t* state,
                                         int16_t* meta, double* acc_reward, uint8_t* done, uint8_t* flags,
                                         uint8_t* n_executed, uint8_t* obs_packed, void* stream) {
  return sokoban_step_entry(1, ctx, n, H, W, num_boxes, max_steps, actions, K, n_actions, state, meta, acc_reward, done, flags,
                            n_executed, obs_packed, stream);
}

extern "C" int ragen_sokoban_render(ragen_ctx* ctx, int64_t n, int H, int W, const uint8_t* state, const int16_t* meta,
                                    uint8_t* obs, void* stream) {
  return sokoban_render_entry(0, ctx, n, H, W, state, meta, obs, stream);
}

extern "C" int ragen_sokoban_render_packed(ragen_ctx* ctx, int64_t n, int H, int W, const uint8_t* state,
                                           const int16_t* meta, uint8_t* obs_packed, void* stream) {
  return sokoban_render_entry(1, ctx, n, H, W, state, meta, obs_packed, stream);
}

extern "C" int ragen_sokoban_step_packed3(ragen_ctx* ctx, int64_t n, int H, int W, int num_boxes, int max_steps,
                                          const uint8_t* actions, int K, const uint8_t* n_actions, uint8_t* state,
                                          int16_t* meta, double* acc_reward, uint8_t* done, uint8_t* flags,
                                          uint8_t* n_executed, uint8_t* obs_packed3, void* stream) {
  return sokoban_step_entry(2, ctx, n, H, W, num_boxes, max_steps, actions, K, n_actions, state, meta, acc_reward, done, flags,
                            n_executed, obs_packed3, stream);
}

extern "C" int ragen_sokoban_step_compact(ragen_ctx* ctx, int64_t n, int H, int W, int num_boxes, int max_steps,
                                          const uint8_t* actions, int K, const uint8_t* n_actions, uint8_t* state4,
                                          int16_t* meta, double* acc_reward, uint8_t* done, uint8_t* flags,
                                          uint8_t* n_executed, uint8_t* obs_packed3, void* stream) {
  return sokoban_step_entry(3, ctx, n, H, W, num_boxes, max_steps, actions, K, n_actions, state4, meta, acc_reward, done, flags,
                            n_executed, obs_packed3, stream);
}

extern "C" int ragen_sokoban_step_delta(ragen_ctx* ctx, int64_t n, int H, int W, int num_boxes, int max_steps,
                                        const uint8_t* actions, int K, const uint8_t* n_actions, uint8_t* state4,
                                        int16_t* meta, uint8_t* delta, void* stream) {
  RAGEN_REQUIRE(delta, "delta is NULL");
  return sokoban_step_entry(3, ctx, n, H, W, num_boxes, max_steps, actions, K, n_actions, state4, meta, nullptr, nullptr, nullptr,
                            nullptr, nullptr, stream, delta);
}

extern "C" int ragen_sokoban_step_glyph3(ragen_ctx* ctx, int64_t n, int H, int W, int num_boxes, int max_steps,
                                         const uint8_t* actions, int K, const uint8_t* n_actions, uint8_t* state3,
                                         int16_t* meta, double* acc_reward, uint8_t* done, uint8_t* flags,
                                         uint8_t* n_executed, uint8_t* delta, void* stream) {
  return sokoban_step_entry(4, ctx, n, H, W, num_boxes, max_steps, actions, K, n_actions, state3, meta, acc_reward, done, flags,
                            n_executed, nullptr, stream, delta);
}

extern "C" int ragen_sokoban_step_delta1(ragen_ctx* ctx, int64_t n, int H, int W, int num_boxes, int max_steps,
                                         const uint8_t* actions, int K, const uint8_t* n_actions, int rooms_format, uint8_t* rooms,
                                         int16_t* meta, uint8_t* delta1, void* stream) {
  RAGEN_REQUIRE(delta1, "delta1 is NULL");
  RAGEN_REQUIRE(rooms_format == 3 || rooms_format == 4, "rooms_format %d (3: nibble rooms, 4: three-bit glyph rooms)", rooms_format);
  return sokoban_step_entry(rooms_format, ctx, n, H, W, num_boxes, max_steps, actions, K, n_actions, rooms, meta, nullptr, nullptr,
                            nullptr, nullptr, nullptr, stream, delta1, 1);
}

extern "C" int ragen_sokoban_render_packed3(ragen_ctx* ctx, int64_t n, int H, int W, const uint8_t* state,
                                            const int16_t* meta, uint8_t* obs_packed3, void* stream) {
  return sokoban_render_entry(2, ctx, n, H, W, state, meta, obs_packed3, stream);
}

extern "C" int ragen_sokoban_adopt_slots(ragen_ctx* ctx, int64_t n, int H, int W, int num_boxes, const uint8_t* state,
                                         const uint8_t* fixed, int16_t* meta, int32_t* status, const uint8_t* only,
                                         uint8_t* state4, uint8_t* state3, void* stream);

extern "C" int ragen_sokoban_adopt(ragen_ctx* ctx, int64_t n, int H, int W, int num_boxes, const uint8_t* state,
                                   const uint8_t* fixed, int16_t* meta, int32_t* status, void* stream) {
  return ragen_sokoban_adopt_slots(ctx, n, H, W, num_boxes, state, fixed, meta, status, nullptr, nullptr, nullptr, stream);
}

extern "C" int ragen_sokoban_adopt_slots(ragen_ctx* ctx, int64_t n, int H, int W, int num_boxes, const uint8_t* state,
                                         const uint8_t* fixed, int16_t* meta, int32_t* status, const uint8_t* only,
                                         uint8_t* state4, uint8_t* state3, void* stream) {
  RAGEN_REQUIRE(ctx, "ctx is NULL");
  RAGEN_REQUIRE(!state4 || ((H * W) & 1) == 0, "nibble rooms need an even H*W, got %dx%d", H, W);
  RAGEN_SOKOBAN_SHAPE(n, H, W);
  RAGEN_REQUIRE(num_boxes >= 0 && num_boxes <= 127, "num_boxes=%d out of range", num_boxes);
  RAGEN_REQUIRE(state && fixed && meta, "NULL pointer argument");
  RAGEN_REQUIRE(aligned16(state) && aligned16(fixed) && aligned8(meta), "state/fixed must be 16-byte aligned, meta 8-byte aligned");
  if (n == 0) return RAGEN_OK;
  DeviceGuard guard(ctx->device);
  const size_t smem = 2 * (size_t)round16(TE * H * W);
  RAGEN_REQUIRE(smem <= 227 * 1024, "room %dx%d does not fit the shared-memory tile (%zu B)", H, W, smem);
  RAGEN_CUDA(cudaFuncSetAttribute(sokoban_adopt_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  AdoptParams p{n, H, W, num_boxes, state, fixed, reinterpret_cast<EnvMeta*>(meta), status, only, state4, state3};
  sokoban_adopt_kernel<<<(unsigned)((n + TE - 1) / TE), TE, smem, (cudaStream_t)stream>>>(p);
  RAGEN_LAUNCHED(ctx);
  return RAGEN_OK;
}
