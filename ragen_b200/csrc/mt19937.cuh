// MT19937 as CPython's `random` module and NumPy's legacy global `np.random` use it - the two streams
// ragen/env/sokoban/utils.py draws from (random.* at :167, 196, 210, 325-347; np.random.randint at :379, 393), both
// seeded by all_seed(seed) (ragen/utils.py:7-18: random.seed(seed); np.random.seed(seed)).
//
// CPython and NumPy are third-party to the reference; this restates their published algorithms (Matsumoto & Nishimura's
// mt19937ar.c: init_genrand / init_by_array / genrand_int32; CPython Lib/random.py: random(), getrandbits, _randbelow,
// randint, sample(pop, 1), choice; NumPy legacy RandomState: seed(int) = init_genrand, randint = masked rejection on
// 32-bit draws).  tests/test_hostsim.py checks these functions against the real modules.
#pragma once
#include <stddef.h>
#include <stdint.h>

#ifndef RAGEN_HD
#ifdef __CUDACC__
#define RAGEN_HD __host__ __device__ __forceinline__
#else
#define RAGEN_HD inline
#endif
#endif

namespace mt {

constexpr int N = 624, M = 397;

// word i of the state lives at mt[i * stride]: stride 1 on the host; on the device the 32 lanes of a warp interleave their
// states (stride 32) so that the lock-step seeding and twist loops of a warp touch whole 128-byte lines
struct State {
  uint32_t* mt;  // 624 words, caller-owned
  int idx;
  int stride;
  RAGEN_HD uint32_t& at(int i) const { return mt[(size_t)i * (size_t)stride]; }
};

RAGEN_HD void init_genrand(State& s, uint32_t seed) {
  uint32_t prev = seed;
  s.at(0) = prev;
  for (int i = 1; i < N; ++i) {
    prev = 1812433253u * (prev ^ (prev >> 30)) + (uint32_t)i;
    s.at(i) = prev;
  }
  s.idx = N;
}

// random.seed(a) for a non-negative int a: key = 32-bit digits of a, least significant first (at least one)
RAGEN_HD void seed_python(State& s, uint64_t a) {
  uint32_t key[2] = {(uint32_t)a, (uint32_t)(a >> 32)};
  const int klen = key[1] ? 2 : 1;
  init_genrand(s, 19650218u);
  int i = 1, j = 0;
  uint32_t prev = s.at(0);  // mt[i - 1], carried in a register: the recurrence never waits on its own store
  for (int k = (N > klen ? N : klen); k; --k) {
    prev = (s.at(i) ^ ((prev ^ (prev >> 30)) * 1664525u)) + key[j] + (uint32_t)j;
    s.at(i) = prev;
    ++i;
    ++j;
    if (i >= N) {
      s.at(0) = prev;
      i = 1;
    }
    if (j >= klen) j = 0;
  }
  for (int k = N - 1; k; --k) {
    prev = (s.at(i) ^ ((prev ^ (prev >> 30)) * 1566083941u)) - (uint32_t)i;
    s.at(i) = prev;
    ++i;
    if (i >= N) {
      s.at(0) = prev;
      i = 1;
    }
  }
  s.at(0) = 0x80000000u;
  s.idx = N;
}

// np.random.seed(a) for 0 <= a < 2^32
RAGEN_HD void seed_numpy(State& s, uint32_t a) { init_genrand(s, a); }

RAGEN_HD uint32_t twist_word(uint32_t cur, uint32_t nxt, uint32_t far) {
  const uint32_t y = (cur & 0x80000000u) | (nxt & 0x7fffffffu);
  return far ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
}
RAGEN_HD void twist(State& s) {
  int kk = 0;
  uint32_t cur = s.at(0);
  for (; kk < N - M; ++kk) {
    const uint32_t nxt = s.at(kk + 1);
    s.at(kk) = twist_word(cur, nxt, s.at(kk + M));
    cur = nxt;
  }
  for (; kk < N - 1; ++kk) {
    const uint32_t nxt = s.at(kk + 1);
    s.at(kk) = twist_word(cur, nxt, s.at(kk + (M - N)));
    cur = nxt;
  }
  s.at(N - 1) = twist_word(cur, s.at(0), s.at(M - 1));
  s.idx = 0;
}

RAGEN_HD uint32_t next32(State& s) {
  if (s.idx >= N) twist(s);
  uint32_t y = s.at(s.idx++);
  y ^= y >> 11;
  y ^= (y << 7) & 0x9d2c5680u;
  y ^= (y << 15) & 0xefc60000u;
  y ^= y >> 18;
  return y;
}

// ---- CPython random ----------------------------------------------------------------------------------------
RAGEN_HD double py_random(State& s) {
  const uint32_t a = next32(s) >> 5, b = next32(s) >> 6;
  return ((double)a * 67108864.0 + (double)b) * (1.0 / 9007199254740992.0);
}
// Random._randbelow_with_getrandbits(n), 1 <= n < 2^32: k = n.bit_length(); r = getrandbits(k) until r < n
RAGEN_HD uint32_t py_randbelow(State& s, uint32_t n) {
  int k = 0;
  for (uint32_t v = n; v; v >>= 1) ++k;
  uint32_t r = next32(s) >> (32 - k);
  while (r >= n) r = next32(s) >> (32 - k);
  return r;
}
// random.randint(a, b) = a + _randbelow(b - a + 1); random.sample(seq, 1)[0] and random.choice(seq) = seq[_randbelow(len)]

// ---- NumPy legacy RandomState.randint(n) (low = 0, high = n, default dtype): masked rejection on 32-bit draws ----
RAGEN_HD uint32_t np_randint(State& s, uint32_t n) {
  const uint32_t rng = n - 1;
  if (rng == 0) return 0;
  uint32_t mask = rng;
  mask |= mask >> 1;
  mask |= mask >> 2;
  mask |= mask >> 4;
  mask |= mask >> 8;
  mask |= mask >> 16;
  uint32_t v = next32(s) & mask;
  while (v > rng) v = next32(s) & mask;
  return v;
}

}  // namespace mt
