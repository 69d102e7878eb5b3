// Sokoban step + reward + text render, fused (sm_100a).
//
// Replaces, per env and per action: SokobanEnv.step (ragen/env/sokoban/env.py:44-51) ->
// gym_sokoban SokobanEnv.step/_push/_move/_calc_reward/_check_if_done, followed by the text branch
// of SokobanEnv.render (env.py:53-57); and per turn the action loop of
// EnvStateManager._execute_actions (ragen/llm_agent/es_manager.py:116-128).
//
// Layout: rooms are struct-of-arrays uint8 [n][H*W] (state, fixed) in HBM.  A CTA owns a tile of
// TE consecutive envs whose bytes are CONTIGUOUS in HBM, so the tile moves HBM<->shared memory as
// three 1-D TMA bulk copies (state in, fixed in; state out, glyph rows out).  One thread steps one
// env inside shared memory (per-env stride H*W bytes = an odd number of 32-bit words for 6x6 and
// 10x10, i.e. bank-conflict free); the glyph rows are then rendered cooperatively by the whole CTA.
// HBM-bound by design: nothing here is a contraction, tensor cores are not used.
#include "common.cuh"

namespace {

constexpr int TE = 128;  // envs per CTA tile == threads per CTA

struct StepParams {
  int64_t n;
  int H, W, num_boxes, max_steps, K;
  const uint8_t* actions;
  const uint8_t* n_actions;
  uint8_t* state;
  const uint8_t* fixed;
  int16_t* player;
  uint16_t* env_steps;
  uint8_t* on_target;
  double* acc_reward;
  uint8_t* done;
  uint8_t* flags;
  uint8_t* n_executed;
  uint8_t* obs;
};

__host__ __device__ inline int round16(int x) { return (x + 15) & ~15; }

// cooperative copy for tiles whose byte count is not a multiple of 16 (only the last tile)
__device__ __forceinline__ void copy_bytes(uint8_t* dst, const uint8_t* src, int bytes) {
  for (int i = threadIdx.x; i < bytes; i += blockDim.x) dst[i] = src[i];
}

// boxes on target as the reference counts them (gym_sokoban _calc_reward):
// num_boxes - #(state == 2 | (fixed == 2 & state == 5))
__device__ __forceinline__ int count_on_target(const uint8_t* S, const uint8_t* F, int HW, int num_boxes) {
  int open = 0;
  for (int i = 0; i < HW; ++i) {
    const int s = S[i];
    open += (s == 2) | ((s == 5) & (F[i] == 2));
  }
  return num_boxes - open;
}

template <int CH, int CW>
__device__ __forceinline__ void render_tile(const uint8_t* s_state, const uint8_t* s_fixed, uint8_t* s_obs, int cnt,
                                            int rH, int rW) {
  const int H = CH ? CH : rH, W = CW ? CW : rW;
  const int HW = H * W, RW = W + 1, G = H * RW;
  const int total = cnt * G;
  for (int j = threadIdx.x; j < total; j += TE) {
    const int e = j / G;
    const int q = j - e * G;
    const int r = q / RW;
    const int c = q - r * RW;
    int g = 7;  // newline (and the unused trailing byte of the env)
    if (c < W) {
      const int src = e * HW + r * W + c;
      g = s_state[src];
      if (g == 5 && s_fixed[src] == 2) g = 6;
    }
    s_obs[j] = (uint8_t)g;
  }
}

template <int CH, int CW, bool RENDER_ONLY>
__global__ void __launch_bounds__(TE) sokoban_tile_kernel(const StepParams p) {
  const int H = CH ? CH : p.H, W = CW ? CW : p.W;
  const int HW = H * W, G = H * (W + 1);
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ __align__(8) uint64_t bar;

  uint8_t* s_state = smem;
  uint8_t* s_fixed = s_state + round16(TE * HW);
  uint8_t* s_obs = s_fixed + round16(TE * HW);

  const int64_t env0 = (int64_t)blockIdx.x * TE;
  const int cnt = (int)min((int64_t)TE, p.n - env0);
  const int room_bytes = cnt * HW;
  const int obs_bytes = cnt * G;
  const bool bulk_rooms = (room_bytes & 15) == 0;
  const bool bulk_obs = (obs_bytes & 15) == 0;
  const uint8_t* g_state_in = p.state + env0 * HW;
  const uint8_t* g_fixed = p.fixed + env0 * HW;

  // ---- stage the tile: HBM -> shared memory -------------------------------------------------
  if (bulk_rooms) {
    if (threadIdx.x == 0) {
      mbar_init(&bar, 1);
      fence_mbar_init();
      mbar_expect_tx(&bar, 2u * (uint32_t)room_bytes);
      tma_load_1d(s_state, g_state_in, (uint32_t)room_bytes, &bar);
      tma_load_1d(s_fixed, g_fixed, (uint32_t)room_bytes, &bar);
    }
    __syncthreads();  // barrier initialised before anyone polls it
    mbar_wait(&bar, 0);
  } else {
    copy_bytes(s_state, g_state_in, room_bytes);
    copy_bytes(s_fixed, g_fixed, room_bytes);
    __syncthreads();
  }

  // ---- one thread steps one env -----------------------------------------------------------------
  if (!RENDER_ONLY) {
    const int e = threadIdx.x;
    if (e < cnt) {
      const int64_t g = env0 + e;
      uint8_t* S = s_state + e * HW;
      const uint8_t* F = s_fixed + e * HW;
      int pos = p.player[g];
      int steps = p.env_steps[g];
      int on = p.on_target[g];
      int na = p.K;
      if (p.n_actions) na = min((int)p.n_actions[g], p.K);
      const uint8_t* act = p.actions + g * p.K;

      double acc = 0.0;
      int fl = 0, nexec = 0, dn = 0;
      for (int k = 0; k < na; ++k) {
        const int a = act[k];
        const int before = pos;
        const bool first = (steps == 0);
        ++steps;
        int delta = 0;  // change of the on-target count caused by this action
        if (a >= 1 && a <= 4) {
          const int dr = (a == 1) ? -1 : (a == 2) ? 1 : 0;
          const int dc = (a == 3) ? -1 : (a == 4) ? 1 : 0;
          const int r = pos / W, c = pos - r * W;
          const int ar = r + dr, ac = c + dc, br = ar + dr, bc = ac + dc;
          // upstream _push gives up (no plain move either) when the cell beyond leaves the grid
          if (ar >= 0 && ar < H && ac >= 0 && ac < W && br < H && bc < W) {
            const int ahead = ar * W + ac;
            const int sa = S[ahead];
            bool go = (sa == 1) | (sa == 2);
            if ((sa == 3 || sa == 4) && br >= 0 && bc >= 0) {
              const int beyond = br * W + bc;
              const int sb = S[beyond];
              if (sb == 1 || sb == 2) {
                const int onto = (F[beyond] == 2);
                S[beyond] = onto ? 3 : 4;
                delta = onto - (int)(F[ahead] == 2);
                go = true;
              }
            }
            if (go) {
              S[ahead] = 5;
              S[pos] = F[pos];
              pos = ahead;
            }
          }
        }
        // reward (gym_sokoban _calc_reward): the count is absolute in the reference; it equals the
        // stored count plus `delta` except on the first step after reset (stored count is 0 there)
        const int now_on = first ? count_on_target(S, F, HW, p.num_boxes) : on + delta;
        double rew = -0.1;
        if (now_on > on) rew += 1.0;
        else if (now_on < on) rew += -1.0;
        const bool won = (now_on == p.num_boxes);
        if (won) rew += 10.0;
        on = now_on;
        acc += rew;
        ++nexec;
        fl = (before != pos ? 1 : 0) | (won ? 2 : 0) | 4;
        dn = (won || steps == p.max_steps) ? 1 : 0;
        if (dn) break;
      }
      p.player[g] = (int16_t)pos;
      p.env_steps[g] = (uint16_t)steps;
      p.on_target[g] = (uint8_t)on;
      p.acc_reward[g] = acc;
      p.done[g] = (uint8_t)dn;
      p.flags[g] = (uint8_t)fl;
      p.n_executed[g] = (uint8_t)nexec;
    }
    __syncthreads();
  }

  // ---- render glyph rows cooperatively, then shared memory -> HBM ---------------------------------
  if (p.obs) render_tile<CH, CW>(s_state, s_fixed, s_obs, cnt, p.H, p.W);
  fence_async_smem();
  __syncthreads();

  uint8_t* g_state_out = p.state + env0 * HW;
  uint8_t* g_obs = p.obs ? p.obs + env0 * G : nullptr;
  const bool store_state = !RENDER_ONLY;
  if (threadIdx.x == 0) {
    bool any = false;
    if (store_state && bulk_rooms) {
      tma_store_1d(g_state_out, s_state, (uint32_t)room_bytes);
      any = true;
    }
    if (g_obs && bulk_obs) {
      tma_store_1d(g_obs, s_obs, (uint32_t)obs_bytes);
      any = true;
    }
    if (any) {
      tma_store_commit();
      tma_store_wait_read();  // shared memory must stay valid until the bulk engine has read it
    }
  }
  if (store_state && !bulk_rooms) copy_bytes(g_state_out, s_state, room_bytes);
  if (g_obs && !bulk_obs) copy_bytes(g_obs, s_obs, obs_bytes);
}

__global__ void sokoban_adopt_kernel(int64_t n, int H, int W, const uint8_t* __restrict__ state,
                                     const uint8_t* __restrict__ fixed, int16_t* player, uint16_t* env_steps,
                                     uint8_t* on_target, int32_t* status) {
  const int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= n) return;
  const int HW = H * W;
  const uint8_t* S = state + g * HW;
  const uint8_t* F = fixed + g * HW;
  int pos = -1, st = 0;
  for (int i = 0; i < HW; ++i) {
    if (S[i] == 5 && pos < 0) pos = i;  // argwhere(...)[0]: first in row-major order
    const int r = i / W, c = i - r * W;
    if ((r == 0 || r == H - 1 || c == 0 || c == W - 1) && F[i] != 0) st = 2;
  }
  if (pos < 0) st = 1;
  player[g] = (int16_t)pos;
  env_steps[g] = 0;
  on_target[g] = 0;
  if (status) status[g] = st;
}

size_t tile_smem(int H, int W) { return (size_t)2 * round16(TE * H * W) + round16(TE * H * (W + 1)); }

template <bool RENDER_ONLY>
int launch_tiles(ragen_ctx* ctx, const StepParams& p, cudaStream_t stream) {
  const size_t smem = tile_smem(p.H, p.W);
  RAGEN_REQUIRE(smem <= 227 * 1024, "room %dx%d does not fit the shared-memory tile (%zu B)", p.H, p.W, smem);
  const unsigned grid = (unsigned)((p.n + TE - 1) / TE);
#define RAGEN_TILE_LAUNCH(CH, CW)                                                                          \
  do {                                                                                                     \
    auto kern = sokoban_tile_kernel<CH, CW, RENDER_ONLY>;                                                  \
    RAGEN_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));        \
    kern<<<grid, TE, smem, stream>>>(p);                                                                   \
  } while (0)
  if (p.H == 6 && p.W == 6) RAGEN_TILE_LAUNCH(6, 6);
  else if (p.H == 8 && p.W == 8) RAGEN_TILE_LAUNCH(8, 8);
  else if (p.H == 10 && p.W == 10) RAGEN_TILE_LAUNCH(10, 10);
  else RAGEN_TILE_LAUNCH(0, 0);
#undef RAGEN_TILE_LAUNCH
  RAGEN_LAUNCHED(ctx);
  return RAGEN_OK;
}

bool aligned16(const void* p) { return ((uintptr_t)p & 15) == 0; }

}  // namespace

extern "C" int ragen_sokoban_step(ragen_ctx* ctx, int64_t n, int H, int W, int num_boxes, int max_steps,
                                  const uint8_t* actions, int K, const uint8_t* n_actions, uint8_t* state,
                                  const uint8_t* fixed, int16_t* player, uint16_t* env_steps, uint8_t* on_target,
                                  double* acc_reward, uint8_t* done, uint8_t* flags, uint8_t* n_executed, uint8_t* obs,
                                  void* stream) {
  RAGEN_REQUIRE(ctx, "ctx is NULL");
  RAGEN_REQUIRE(n >= 0 && H >= 3 && W >= 3 && H * W <= 32767, "bad shape n=%lld H=%d W=%d", (long long)n, H, W);
  RAGEN_REQUIRE(K >= 1 && K <= 255, "K=%d out of range", K);
  RAGEN_REQUIRE(actions && state && fixed && player && env_steps && on_target && acc_reward && done && flags && n_executed,
                "NULL pointer argument");
  RAGEN_REQUIRE(aligned16(state) && aligned16(fixed) && (!obs || aligned16(obs)), "state/fixed/obs must be 16-byte aligned");
  if (n == 0) return RAGEN_OK;
  DeviceGuard guard(ctx->device);
  StepParams p{n, H, W, num_boxes, max_steps, K, actions, n_actions, state, fixed, player, env_steps, on_target,
               acc_reward, done, flags, n_executed, obs};
  return launch_tiles<false>(ctx, p, (cudaStream_t)stream);
}

extern "C" int ragen_sokoban_render(ragen_ctx* ctx, int64_t n, int H, int W, const uint8_t* state, const uint8_t* fixed,
                                    uint8_t* obs, void* stream) {
  RAGEN_REQUIRE(ctx, "ctx is NULL");
  RAGEN_REQUIRE(n >= 0 && H >= 3 && W >= 3 && H * W <= 32767, "bad shape n=%lld H=%d W=%d", (long long)n, H, W);
  RAGEN_REQUIRE(state && fixed && obs, "NULL pointer argument");
  RAGEN_REQUIRE(aligned16(state) && aligned16(fixed) && aligned16(obs), "state/fixed/obs must be 16-byte aligned");
  if (n == 0) return RAGEN_OK;
  DeviceGuard guard(ctx->device);
  StepParams p{};
  p.n = n; p.H = H; p.W = W;
  p.state = const_cast<uint8_t*>(state);
  p.fixed = fixed;
  p.obs = obs;
  return launch_tiles<true>(ctx, p, (cudaStream_t)stream);
}

extern "C" int ragen_sokoban_adopt(ragen_ctx* ctx, int64_t n, int H, int W, const uint8_t* state, const uint8_t* fixed,
                                   int16_t* player, uint16_t* env_steps, uint8_t* on_target, int32_t* status,
                                   void* stream) {
  RAGEN_REQUIRE(ctx, "ctx is NULL");
  RAGEN_REQUIRE(n >= 0 && H >= 3 && W >= 3 && H * W <= 32767, "bad shape n=%lld H=%d W=%d", (long long)n, H, W);
  RAGEN_REQUIRE(state && fixed && player && env_steps && on_target, "NULL pointer argument");
  if (n == 0) return RAGEN_OK;
  DeviceGuard guard(ctx->device);
  const int threads = 128;
  sokoban_adopt_kernel<<<(unsigned)((n + threads - 1) / threads), threads, 0, (cudaStream_t)stream>>>(
      n, H, W, state, fixed, player, env_steps, on_target, status);
  RAGEN_LAUNCHED(ctx);
  return RAGEN_OK;
}
