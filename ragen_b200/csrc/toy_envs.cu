// FrozenLake (map generation, slippery step, text render) and Bandit (reset, step) kernels (sm_100a).
//
// Replaces ragen/env/frozen_lake/env.py:28-57 + utils.py:6-47 (and the gymnasium FrozenLakeEnv.reset/step they call)
// and ragen/env/bandit/env.py:24-77.  The random streams are NumPy's own (np_random.cuh): every env owns a PCG64
// generator in HBM (state plane + increment plane, 16 B each per env), so maps, start states, slip directions and
// bandit pay-outs are the ones the reference produces for the same seed.
//
// One thread per env.  The map of an env (<= 64 cells) is packed 2 bits per cell into two registers; glyph rows are
// staged in shared memory and leave the CTA as coalesced 128-bit stores.  HBM-bound integer work, no tensor cores.
#include "common.cuh"
#include "toy_logic.cuh"

namespace {

constexpr int TB = 128;  // envs per CTA == threads per CTA

// cooperative coalesced copies between a CTA-contiguous global range and shared memory
__device__ __forceinline__ void tile_g2s(uint8_t* s, const uint8_t* g, int bytes) {
  if (((bytes | (int)((uintptr_t)g & 15)) & 15) == 0) {
    const uint4* gv = reinterpret_cast<const uint4*>(g);
    uint4* sv = reinterpret_cast<uint4*>(s);
    for (int i = threadIdx.x; i < (bytes >> 4); i += blockDim.x) sv[i] = ldg_stream(gv + i);
  } else {
    for (int i = threadIdx.x; i < bytes; i += blockDim.x) s[i] = g[i];
  }
}
__device__ __forceinline__ void tile_s2g(uint8_t* g, const uint8_t* s, int bytes) {
  if (((bytes | (int)((uintptr_t)g & 15)) & 15) == 0) {
    uint4* gv = reinterpret_cast<uint4*>(g);
    const uint4* sv = reinterpret_cast<const uint4*>(s);
    for (int i = threadIdx.x; i < (bytes >> 4); i += blockDim.x) stg_stream(gv + i, sv[i]);
  } else {
    for (int i = threadIdx.x; i < bytes; i += blockDim.x) g[i] = s[i];
  }
}

__device__ __forceinline__ nprand::Pcg64 load_rng(const uint64_t* rng, int64_t n, int64_t g) {
  const ulonglong2 st = reinterpret_cast<const ulonglong2*>(rng)[g];
  const ulonglong2 in = reinterpret_cast<const ulonglong2*>(rng)[n + g];
  return nprand::Pcg64{st.x, st.y, in.x, in.y, 0u, 0u};
}
__device__ __forceinline__ void store_rng_state(uint64_t* rng, int64_t g, const nprand::Pcg64& r) {
  reinterpret_cast<ulonglong2*>(rng)[g] = make_ulonglong2(r.state_hi, r.state_lo);
}
__device__ __forceinline__ void store_rng_inc(uint64_t* rng, int64_t n, int64_t g, const nprand::Pcg64& r) {
  reinterpret_cast<ulonglong2*>(rng)[n + g] = make_ulonglong2(r.inc_hi, r.inc_lo);
}

// ---- FrozenLake ---------------------------------------------------------------------------------------------
struct FlParams {
  int64_t n;
  int size, slippery, K, max_tries;
  double cdf0;
  const uint64_t* seeds;
  const uint8_t* actions;
  const uint8_t* n_actions;
  uint8_t* tiles;
  uint8_t* pos;
  uint64_t* rng;
  double* acc_reward;
  uint8_t* done;
  uint8_t* flags;
  uint8_t* n_executed;
  uint8_t* obs;
  int32_t* status;
};

enum { FL_RESET = 0, FL_STEP = 1, FL_RENDER = 2 };

template <int MODE>
__global__ void __launch_bounds__(TB) frozenlake_kernel(const FlParams p) {
  extern __shared__ __align__(16) uint8_t smem[];
  const int S = p.size, C = S * S, G = S * (S + 1);
  uint8_t* s_tiles = smem;                       // [TB][C]
  uint8_t* s_obs = smem + ((TB * C + 15) & ~15);  // [TB][G]
  const int64_t env0 = (int64_t)blockIdx.x * TB;
  const int cnt = (int)min((int64_t)TB, p.n - env0);
  const int e = threadIdx.x;
  const int64_t g = env0 + e;
  const bool mine = e < cnt;

  toy::Tiles tiles{0, 0};
  int pos = 0;
  if (MODE == FL_RESET) {
    if (mine) {
      // FrozenLakeEnv.reset: the map comes from a generator seeded with `seed` (utils.py:35), then the env's own
      // generator is re-seeded with the same seed and draws ONE uniform for the start state (gymnasium reset)
      const uint64_t seed = p.seeds[g];
      nprand::Pcg64 r;
      nprand::pcg_seed(r, seed);
      const int start = toy::fl_generate_map(r, S, p.cdf0, tiles, p.max_tries);
      nprand::pcg_seed(r, seed);
      (void)nprand::pcg_random(r);
      pos = start < 0 ? 0 : start;
      store_rng_state(p.rng, g, r);
      store_rng_inc(p.rng, p.n, g, r);
      p.pos[g] = (uint8_t)pos;
      if (p.status) p.status[g] = start < 0 ? 1 : 0;
      uint8_t* T = s_tiles + e * C;
      for (int i = 0; i < C; ++i) T[i] = (uint8_t)tiles.get(i);
    }
    __syncthreads();
    tile_s2g(p.tiles + env0 * C, s_tiles, cnt * C);
  } else {
    tile_g2s(s_tiles, p.tiles + env0 * C, cnt * C);
    if (mine) pos = p.pos[g];
    __syncthreads();
    if (mine) {
      const uint8_t* T = s_tiles + e * C;
      for (int i = 0; i < C; ++i) tiles.set(i, T[i] & 3);
    }
  }

  if (MODE == FL_STEP && mine) {
    int na = p.K;
    if (p.n_actions) na = min((int)p.n_actions[g], p.K);
    nprand::Pcg64 r = load_rng(p.rng, p.n, g);
    double acc = 0.0;
    int fl = 0, nexec = 0, dn = 0;
    const uint8_t* act = p.actions + g * p.K;
    for (int k = 0; k < na; ++k) {
      const int a = act[k];
      if (a < 1 || a > 4) break;  // action_map has no such key: the reference raises, nothing more is executed
      bool d, eff, suc;
      const int rew = toy::fl_step(tiles, S, pos, r, a, p.slippery != 0, d, eff, suc);
      acc += (double)rew;
      ++nexec;
      fl = (eff ? 1 : 0) | (suc ? 2 : 0) | 4;
      dn = d ? 1 : 0;
      if (dn) break;
    }
    if (nexec) {
      store_rng_state(p.rng, g, r);
      p.pos[g] = (uint8_t)pos;
    }
    p.acc_reward[g] = acc;
    p.done[g] = (uint8_t)dn;
    p.flags[g] = (uint8_t)fl;
    p.n_executed[g] = (uint8_t)nexec;
  }

  if (p.obs) {
    if (mine) toy::fl_render(tiles, S, pos, s_obs + e * G);
    __syncthreads();
    tile_s2g(p.obs + env0 * G, s_obs, cnt * G);
  }
}

// ---- FrozenLake 4x4 (the reference's default size, C2): no tile staging ----------------------------------------
// A 4x4 map is exactly one 16-byte vector per env, so the thread that owns the env loads it with ONE coalesced 128-bit
// load and squeezes it to 2 bits per cell in a register; only the 20-byte glyph rows go through shared memory (five words
// per env, an odd word stride: conflict-free) so that they leave as coalesced 128-bit stores.
constexpr int TB4 = 256;

__device__ __forceinline__ uint32_t squeeze4(uint32_t w) {  // 4 bytes (0..3 each) -> 8 bits, cell 0 in the low 2 bits
  w &= 0x03030303u;
  return (w | (w >> 6) | (w >> 12) | (w >> 18)) & 0xffu;
}

template <int MODE>
__global__ void __launch_bounds__(TB4) frozenlake4_kernel(const FlParams p) {
  __shared__ __align__(16) uint32_t s_obs[TB4 * 5];
  const int64_t env0 = (int64_t)blockIdx.x * TB4;
  const int cnt = (int)min((int64_t)TB4, p.n - env0);
  const int e = threadIdx.x;
  const int64_t g = env0 + e;
  const bool mine = e < cnt;
  toy::Tiles tiles{0, 0};
  int pos = 0;
  if (mine) {
    const uint4 t = ldg_stream(reinterpret_cast<const uint4*>(p.tiles) + g);
    tiles.lo = (uint64_t)(squeeze4(t.x) | (squeeze4(t.y) << 8) | (squeeze4(t.z) << 16) | (squeeze4(t.w) << 24));
    pos = p.pos[g];
  }
  if (MODE == FL_STEP && mine) {
    int na = p.K;
    if (p.n_actions) na = min((int)p.n_actions[g], p.K);
    nprand::Pcg64 r = load_rng(p.rng, p.n, g);
    double acc = 0.0;
    int fl = 0, nexec = 0, dn = 0;
    const uint8_t* act = p.actions + g * p.K;
    for (int k = 0; k < na; ++k) {
      const int a = act[k];
      if (a < 1 || a > 4) break;
      bool d, eff, suc;
      const int rew = toy::fl_step(tiles, 4, pos, r, a, p.slippery != 0, d, eff, suc);
      acc += (double)rew;
      ++nexec;
      fl = (eff ? 1 : 0) | (suc ? 2 : 0) | 4;
      dn = d ? 1 : 0;
      if (dn) break;
    }
    if (nexec) {
      store_rng_state(p.rng, g, r);
      p.pos[g] = (uint8_t)pos;
    }
    p.acc_reward[g] = acc;
    p.done[g] = (uint8_t)dn;
    p.flags[g] = (uint8_t)fl;
    p.n_executed[g] = (uint8_t)nexec;
  }
  if (p.obs) {
    if (mine) {
      // fl_render with everything in registers: 2-bit tile -> glyph code (S drawn as F), the player's cell patched, a
      // separator after every fourth code
      const uint32_t t32 = (uint32_t)tiles.lo;
      const uint32_t raw_here = (t32 >> (2 * pos)) & 3u;
      const uint32_t pcode = (raw_here == 2u) ? 4u : (raw_here == 3u) ? 5u : 0u;
      uint32_t row[4];
#pragma unroll
      for (int rr = 0; rr < 4; ++rr) {
        uint32_t w = 0;
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          const int i = rr * 4 + c;
          uint32_t raw = (t32 >> (2 * i)) & 3u;
          uint32_t code = raw ? raw : 1u;
          code = (i == pos) ? pcode : code;
          w |= code << (8 * c);
        }
        row[rr] = w;
      }
      // 20-byte stream r0 r0 r0 r0 7 r1 r1 r1 r1 7 ... as five little-endian words
      uint32_t* o = s_obs + e * 5;
      o[0] = row[0];
      o[1] = 7u | (row[1] << 8);
      o[2] = (row[1] >> 24) | (7u << 8) | (row[2] << 16);
      o[3] = (row[2] >> 16) | (7u << 16) | (row[3] << 24);
      o[4] = (row[3] >> 8) | (7u << 24);
    }
    __syncthreads();
    tile_s2g(p.obs + env0 * 20, reinterpret_cast<const uint8_t*>(s_obs), cnt * 20);
  }
}

// ---- Bandit --------------------------------------------------------------------------------------------------
__global__ void bandit_reset_kernel(int64_t n, const uint64_t* __restrict__ seeds, uint64_t* rng, uint8_t* lo_first) {
  const int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= n) return;
  nprand::Pcg64 r;
  const bool lf = toy::bandit_reset(r, seeds[g]);
  store_rng_state(rng, g, r);
  store_rng_inc(rng, n, g, r);
  lo_first[g] = lf ? 1 : 0;
}

__global__ void bandit_step_kernel(int64_t n, int start, double lo_score, double hi_lo, double hi_hi, double hi_prob,
                                   const uint8_t* __restrict__ actions, const uint8_t* __restrict__ active, uint64_t* rng,
                                   const uint8_t* __restrict__ lo_first, double* reward, uint8_t* flags) {
  const int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= n) return;
  double rew = 0.0;
  int fl = 0;
  const int arm = (int)actions[g] - start;
  if ((!active || active[g]) && (arm == 0 || arm == 1)) {
    const bool lf = lo_first[g] != 0;
    const bool is_lo = (arm == 0) == lf;
    if (is_lo) {
      rew = lo_score;
    } else {
      nprand::Pcg64 r = load_rng(rng, n, g);
      rew = (toy::bandit_step(r, lf, arm, hi_prob) == 1) ? hi_hi : hi_lo;
      store_rng_state(rng, g, r);
    }
    fl = 1 | (is_lo ? 0 : 2) | 4;  // effective always, success = pulled the hi arm, executed
  }
  reward[g] = rew;
  flags[g] = (uint8_t)fl;
}

size_t fl_smem(int size) { return (size_t)((TB * size * size + 15) & ~15) + (size_t)TB * size * (size + 1) + 16; }

template <int MODE>
int fl_launch(ragen_ctx* ctx, const FlParams& p, cudaStream_t stream) {
  DeviceGuard guard(ctx->device);
  if constexpr (MODE != FL_RESET) {
    if (p.size == 4 && ((uintptr_t)p.tiles & 15) == 0) {
      frozenlake4_kernel<MODE><<<(unsigned)((p.n + TB4 - 1) / TB4), TB4, 0, stream>>>(p);
      RAGEN_LAUNCHED(ctx);
      return RAGEN_OK;
    }
  }
  frozenlake_kernel<MODE><<<(unsigned)((p.n + TB - 1) / TB), TB, fl_smem(p.size), stream>>>(p);
  RAGEN_LAUNCHED(ctx);
  return RAGEN_OK;
}

}  // namespace

#define RAGEN_FL_SHAPE(n, size) \
  RAGEN_REQUIRE(n >= 0 && size >= 2 && size <= toy::FL_MAX_SIZE, "bad shape n=%lld size=%d (2..8)", (long long)n, size)

extern "C" int ragen_frozenlake_reset(ragen_ctx* ctx, int64_t n, int size, double p, const uint64_t* seeds, uint8_t* tiles,
                                      uint8_t* pos, uint64_t* rng, int32_t* status, uint8_t* obs, void* stream) {
  RAGEN_REQUIRE(ctx, "ctx is NULL");
  RAGEN_FL_SHAPE(n, size);
  RAGEN_REQUIRE(seeds && tiles && pos && rng, "NULL pointer argument");
  RAGEN_REQUIRE(((uintptr_t)rng & 15) == 0, "rng must be 16-byte aligned");
  RAGEN_REQUIRE(p > 0.0 && p <= 1.0, "p=%g out of range", p);
  if (n == 0) return RAGEN_OK;
  FlParams q{};
  q.n = n;
  q.size = size;
  q.max_tries = 100000;
  const double c0 = p, c1 = p + (1.0 - p);  // numpy choice(): cdf = cumsum([p, 1-p]); cdf /= cdf[-1]
  q.cdf0 = c0 / c1;
  q.seeds = seeds;
  q.tiles = tiles;
  q.pos = pos;
  q.rng = rng;
  q.status = status;
  q.obs = obs;
  return fl_launch<FL_RESET>(ctx, q, (cudaStream_t)stream);
}

extern "C" int ragen_frozenlake_step(ragen_ctx* ctx, int64_t n, int size, int is_slippery, const uint8_t* actions, int K,
                                     const uint8_t* n_actions, const uint8_t* tiles, uint8_t* pos, uint64_t* rng,
                                     double* acc_reward, uint8_t* done, uint8_t* flags, uint8_t* n_executed, uint8_t* obs,
                                     void* stream) {
  RAGEN_REQUIRE(ctx, "ctx is NULL");
  RAGEN_FL_SHAPE(n, size);
  RAGEN_REQUIRE(K >= 1 && K <= 255, "K=%d out of range", K);
  RAGEN_REQUIRE(actions && tiles && pos && rng && acc_reward && done && flags && n_executed, "NULL pointer argument");
  RAGEN_REQUIRE(((uintptr_t)rng & 15) == 0, "rng must be 16-byte aligned");
  if (n == 0) return RAGEN_OK;
  FlParams q{};
  q.n = n;
  q.size = size;
  q.slippery = is_slippery;
  q.K = K;
  q.actions = actions;
  q.n_actions = n_actions;
  q.tiles = const_cast<uint8_t*>(tiles);
  q.pos = pos;
  q.rng = rng;
  q.acc_reward = acc_reward;
  q.done = done;
  q.flags = flags;
  q.n_executed = n_executed;
  q.obs = obs;
  return fl_launch<FL_STEP>(ctx, q, (cudaStream_t)stream);
}

extern "C" int ragen_frozenlake_render(ragen_ctx* ctx, int64_t n, int size, const uint8_t* tiles, const uint8_t* pos,
                                       uint8_t* obs, void* stream) {
  RAGEN_REQUIRE(ctx, "ctx is NULL");
  RAGEN_FL_SHAPE(n, size);
  RAGEN_REQUIRE(tiles && pos && obs, "NULL pointer argument");
  if (n == 0) return RAGEN_OK;
  FlParams q{};
  q.n = n;
  q.size = size;
  q.tiles = const_cast<uint8_t*>(tiles);
  q.pos = const_cast<uint8_t*>(pos);
  q.obs = obs;
  return fl_launch<FL_RENDER>(ctx, q, (cudaStream_t)stream);
}

extern "C" int ragen_bandit_reset(ragen_ctx* ctx, int64_t n, const uint64_t* seeds, uint64_t* rng, uint8_t* lo_first,
                                  void* stream) {
  RAGEN_REQUIRE(ctx, "ctx is NULL");
  RAGEN_REQUIRE(n >= 0, "bad n=%lld", (long long)n);
  RAGEN_REQUIRE(seeds && rng && lo_first, "NULL pointer argument");
  RAGEN_REQUIRE(((uintptr_t)rng & 15) == 0, "rng must be 16-byte aligned");
  if (n == 0) return RAGEN_OK;
  DeviceGuard guard(ctx->device);
  bandit_reset_kernel<<<(unsigned)((n + 127) / 128), 128, 0, (cudaStream_t)stream>>>(n, seeds, rng, lo_first);
  RAGEN_LAUNCHED(ctx);
  return RAGEN_OK;
}

extern "C" int ragen_bandit_step(ragen_ctx* ctx, int64_t n, int action_space_start, double lo_arm_score,
                                 double hi_arm_loscore, double hi_arm_hiscore, double hi_arm_hiscore_prob,
                                 const uint8_t* actions, const uint8_t* active, uint64_t* rng, const uint8_t* lo_first,
                                 double* reward, uint8_t* flags, void* stream) {
  RAGEN_REQUIRE(ctx, "ctx is NULL");
  RAGEN_REQUIRE(n >= 0, "bad n=%lld", (long long)n);
  RAGEN_REQUIRE(actions && rng && lo_first && reward && flags, "NULL pointer argument");
  RAGEN_REQUIRE(((uintptr_t)rng & 15) == 0, "rng must be 16-byte aligned");
  if (n == 0) return RAGEN_OK;
  DeviceGuard guard(ctx->device);
  bandit_step_kernel<<<(unsigned)((n + 127) / 128), 128, 0, (cudaStream_t)stream>>>(
      n, action_space_start, lo_arm_score, hi_arm_loscore, hi_arm_hiscore, hi_arm_hiscore_prob, actions, active, rng, lo_first,
      reward, flags);
  RAGEN_LAUNCHED(ctx);
  return RAGEN_OK;
}
