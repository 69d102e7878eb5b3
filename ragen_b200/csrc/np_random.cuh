// NumPy's default generator, bit for bit: Generator(PCG64(SeedSequence(seed))) as created by
// gymnasium.utils.seeding.np_random(seed) (called at ragen/env/frozen_lake/utils.py:35, gym.Env.reset(seed=...)
// at ragen/env/frozen_lake/env.py:33 and ragen/env/bandit/env.py:52).
//
// NumPy is a third-party dependency of the reference (requirements.txt pins no version; NumPy 2.3 here).  This
// restates its published algorithms: SeedSequence (O'Neill's seed_seq_fe128 variant, pool of 4 words), PCG64
// (128-bit LCG, XSL-RR 128/64 output), Generator.random (53-bit doubles), Generator.integers for ranges that
// fit 32 bits (pcg64_next32 + Lemire rejection), Generator.choice with probabilities (cdf search).
// tests/test_hostsim.py runs these very functions on the host against numpy itself.
//
// Pure functions, usable from host (tests) and device.
#pragma once
#include <stdint.h>

#ifdef __CUDACC__
#define RAGEN_HD __host__ __device__ __forceinline__
#else
#define RAGEN_HD inline
#endif

namespace nprand {

typedef unsigned __int128 u128;

struct Pcg64 {
  uint64_t state_hi, state_lo;  // 128-bit LCG state
  uint64_t inc_hi, inc_lo;      // 128-bit odd increment
  uint32_t has32, u32;          // numpy's pcg64_next32 keeps the unused high half of a 64-bit draw (never flushed
                                // by 64-bit draws); only integers() with 32-bit ranges consumes it
};

RAGEN_HD u128 make128(uint64_t hi, uint64_t lo) { return ((u128)hi << 64) | (u128)lo; }

RAGEN_HD void pcg_step(Pcg64& g) {
  const u128 mult = make128(0x2360ED051FC65DA4ull, 0x4385DF649FCCF645ull);
  u128 s = make128(g.state_hi, g.state_lo) * mult + make128(g.inc_hi, g.inc_lo);
  g.state_hi = (uint64_t)(s >> 64);
  g.state_lo = (uint64_t)s;
}

RAGEN_HD uint64_t pcg_next64(Pcg64& g) {
  pcg_step(g);
  const uint64_t x = g.state_hi ^ g.state_lo;
  const unsigned r = (unsigned)(g.state_hi >> 58);
  return (x >> r) | (x << ((64u - r) & 63u));
}

// Generator.random(): 53 random bits scaled to [0, 1)
RAGEN_HD double pcg_random(Pcg64& g) { return (double)(pcg_next64(g) >> 11) * (1.0 / 9007199254740992.0); }

// SeedSequence(seed).generate_state(4, uint64) for a non-negative integer seed, then PCG64 seeding
// (pcg_setseq_128_srandom: state = 0; step; state += initstate; step).
RAGEN_HD void pcg_seed(Pcg64& g, uint64_t seed) {
  const uint32_t INIT_A = 0x43b0d7e5u, MULT_A = 0x931e8875u, INIT_B = 0x8b51f9ddu, MULT_B = 0x58f38dedu;
  const uint32_t MIX_L = 0xca01f9ddu, MIX_R = 0x4973f715u;
  uint32_t hc = INIT_A;
  uint32_t pool[4];
  const uint32_t ent[4] = {(uint32_t)seed, (uint32_t)(seed >> 32), 0u, 0u};  // missing entropy words are zero
#define RAGEN_HASHMIX(dst, value)  \
  {                                \
    uint32_t v_ = (value) ^ hc;    \
    hc *= MULT_A;                  \
    v_ *= hc;                      \
    v_ ^= v_ >> 16;                \
    dst = v_;                      \
  }
  for (int i = 0; i < 4; ++i) RAGEN_HASHMIX(pool[i], ent[i]);
  for (int s = 0; s < 4; ++s)
    for (int d = 0; d < 4; ++d)
      if (s != d) {
        uint32_t h;
        RAGEN_HASHMIX(h, pool[s]);
        uint32_t r = MIX_L * pool[d] - MIX_R * h;
        r ^= r >> 16;
        pool[d] = r;
      }
#undef RAGEN_HASHMIX
  uint32_t hb = INIT_B;
  uint32_t w[8];
  for (int i = 0; i < 8; ++i) {
    uint32_t v = pool[i & 3] ^ hb;
    hb *= MULT_B;
    v *= hb;
    v ^= v >> 16;
    w[i] = v;
  }
  const uint64_t v0 = (uint64_t)w[0] | ((uint64_t)w[1] << 32), v1 = (uint64_t)w[2] | ((uint64_t)w[3] << 32);
  const uint64_t v2 = (uint64_t)w[4] | ((uint64_t)w[5] << 32), v3 = (uint64_t)w[6] | ((uint64_t)w[7] << 32);
  const u128 inc = (make128(v2, v3) << 1) | (u128)1;
  g.inc_hi = (uint64_t)(inc >> 64);
  g.inc_lo = (uint64_t)inc;
  g.state_hi = g.state_lo = 0;
  g.has32 = g.u32 = 0;
  pcg_step(g);
  const u128 s = make128(g.state_hi, g.state_lo) + make128(v0, v1);
  g.state_hi = (uint64_t)(s >> 64);
  g.state_lo = (uint64_t)s;
  pcg_step(g);
}

// Generator.integers(n) for 1 <= n <= 2^32 - 1 (one element): numpy's buffered_bounded_lemire_uint32 draws through
// the bit generator's next_uint32 (low half of a fresh 64-bit word first, then the kept high half) and rejects with
// Lemire's threshold.
RAGEN_HD uint32_t pcg_next32(Pcg64& g) {
  if (g.has32) {
    g.has32 = 0;
    return g.u32;
  }
  const uint64_t v = pcg_next64(g);
  g.has32 = 1;
  g.u32 = (uint32_t)(v >> 32);
  return (uint32_t)v;
}
RAGEN_HD uint32_t bounded32(Pcg64& g, uint32_t n) {
  if (n == 1) return 0;  // numpy returns `low` without consuming randomness when the range is empty
  const uint32_t rng = n - 1;
  uint64_t m = (uint64_t)pcg_next32(g) * (uint64_t)n;
  uint32_t left = (uint32_t)m;
  if (left < n) {
    const uint32_t thr = (0xffffffffu - rng) % n;
    while (left < thr) {
      m = (uint64_t)pcg_next32(g) * (uint64_t)n;
      left = (uint32_t)m;
    }
  }
  return (uint32_t)(m >> 32);
}

}  // namespace nprand
