// Sokoban rgb_array observation as a sprite blit into a headless device buffer (sm_100a).
//
// Replaces SokobanEnv.render(mode='rgb_array') (ragen/env/sokoban/env.py:58-59) -> gym_sokoban
// SokobanEnv.get_image -> render_utils.room_to_rgb(room_state, room_fixed): the player standing on a target becomes
// surface 6, then every cell (i, j) receives the S x S x 3 sprite of its surface id at pixel rows [i*S, (i+1)*S) and
// pixel columns [j*S, (j+1)*S) of a uint8 [H*S][W*S][3] image (S = 16 upstream).
//
// The sprites are DATA here (atlas uint8 [7][S][S][3], surface order wall, floor, box_target, box_on_target, box, player,
// player_on_target = the room_state codes + 6): upstream loads them from PNG files that ship inside the un-vendored
// gym_sokoban package, so the blit is pinned by the restatement in oracle/ and the pixel values are whatever atlas the
// caller loads.
//
// HBM-write-bound: one image is H*W*S*S*3 bytes out (76 800 B for 10x10) against H*W + 8 bytes in.  A CTA keeps the atlas
// in shared memory and walks images grid-stride; for S = 16 a sprite row is 48 B = three 16-byte vectors, so every thread
// stores whole uint4s, consecutive threads consecutive addresses, with streaming stores (nothing re-reads the image on
// the device).
#include "common.cuh"

namespace {

constexpr int RGB_T = 256;

struct __align__(8) EnvMetaRgb {  // the 8-byte per-env record of sokoban_step.cu
  int16_t player;
  uint16_t steps;
  int8_t on_target;
  uint8_t under;
  int8_t actual;
  uint8_t pad;
};

struct RgbParams {
  int64_t m;            // images to render
  int H, W, S;
  const uint8_t* state;
  const EnvMetaRgb* meta;
  const uint8_t* atlas;
  const int64_t* rows;  // env index of image i, or NULL: image i is env i
  uint8_t* out;
};

// surface ids of one room into shared memory (room_to_rgb: room[(room == 5) & (room_structure == 2)] = 6)
__device__ __forceinline__ void load_cells(const RgbParams& p, int64_t env, uint8_t* s_cells, int HW) {
  const uint8_t* st = p.state + env * HW;
  const EnvMetaRgb m = p.meta[env];
  for (int c = threadIdx.x; c < HW; c += RGB_T) {
    int id = st[c];
    if (id > 6) id = 0;  // never produced by the step kernel; keeps the atlas index in range
    if (c == m.player && m.under && id == 5) id = 6;
    s_cells[c] = (uint8_t)id;
  }
}

// S = 16 * VS16: sprite rows are whole 16-byte vectors
template <int S>
__global__ void __launch_bounds__(RGB_T) sokoban_rgb_vec_kernel(const RgbParams p) {
  constexpr int VS = S * 3 / 16;  // vectors per sprite row
  extern __shared__ __align__(16) uint8_t smem[];
  uint4* s_atlas = reinterpret_cast<uint4*>(smem);                  // [7][S][VS]
  uint8_t* s_cells = smem + 7 * S * VS * 16;                        // [H*W]
  const int H = p.H, W = p.W, HW = H * W;
  const uint4* g_atlas = reinterpret_cast<const uint4*>(p.atlas);
  for (int i = threadIdx.x; i < 7 * S * VS; i += RGB_T) s_atlas[i] = g_atlas[i];

  const int RV = W * VS;          // vectors per pixel row
  const int NV = H * S * RV;      // vectors per image
  // (pixel row, vector in row) of this thread's first vector, and the constant step between its vectors
  const int y_first = threadIdx.x / RV, x_first = threadIdx.x - y_first * RV;
  const int dy = RGB_T / RV, dx = RGB_T - dy * RV;

  for (int64_t img = blockIdx.x; img < p.m; img += gridDim.x) {
    const int64_t env = p.rows ? p.rows[img] : img;
    __syncthreads();  // previous image's cells are no longer read (and the atlas is in place)
    load_cells(p, env, s_cells, HW);
    __syncthreads();
    uint4* o = reinterpret_cast<uint4*>(p.out + img * (int64_t)NV * 16);
    int y = y_first, x = x_first;
    for (int v = threadIdx.x; v < NV; v += RGB_T) {
      const int col = x / VS, part = x - col * VS;
      const int r = y / S, sy = y - r * S;
      const int id = s_cells[r * W + col];
      __stcs(o + v, s_atlas[(id * S + sy) * VS + part]);
      x += dx;
      y += dy;
      if (x >= RV) {
        x -= RV;
        ++y;
      }
    }
  }
}

// any S: one byte per thread per iteration
__global__ void __launch_bounds__(RGB_T) sokoban_rgb_byte_kernel(const RgbParams p) {
  extern __shared__ __align__(16) uint8_t smem[];
  const int H = p.H, W = p.W, S = p.S, HW = H * W;
  const int atlas_bytes = 7 * S * S * 3;
  uint8_t* s_atlas = smem;
  uint8_t* s_cells = smem + ((atlas_bytes + 15) & ~15);
  for (int i = threadIdx.x; i < atlas_bytes; i += RGB_T) s_atlas[i] = p.atlas[i];
  const int RB = W * S * 3, NB = H * S * RB, SB = S * 3;
  for (int64_t img = blockIdx.x; img < p.m; img += gridDim.x) {
    const int64_t env = p.rows ? p.rows[img] : img;
    __syncthreads();
    load_cells(p, env, s_cells, HW);
    __syncthreads();
    uint8_t* o = p.out + img * (int64_t)NB;
    for (int b = threadIdx.x; b < NB; b += RGB_T) {
      const int y = b / RB, xb = b - y * RB;
      const int col = xb / SB, within = xb - col * SB;
      const int r = y / S, sy = y - r * S;
      o[b] = s_atlas[(s_cells[r * W + col] * S + sy) * SB + within];
    }
  }
}

}  // namespace

extern "C" int ragen_sokoban_render_rgb(ragen_ctx* ctx, int64_t n, int H, int W, int S, const uint8_t* state,
                                        const int16_t* meta, const uint8_t* atlas, const int64_t* rows, int64_t n_rows,
                                        uint8_t* out, void* stream) {
  RAGEN_REQUIRE(ctx, "ctx is NULL");
  RAGEN_REQUIRE(n >= 0 && H >= 3 && W >= 3 && H * W <= 32767, "bad shape n=%lld H=%d W=%d", (long long)n, H, W);
  RAGEN_REQUIRE(S >= 1 && S <= 64, "sprite side S=%d out of range [1, 64]", S);
  RAGEN_REQUIRE(n_rows >= 0 && (rows || n_rows == 0), "rows is NULL with n_rows=%lld", (long long)n_rows);
  const int64_t m = rows ? n_rows : n;
  if (m == 0) return RAGEN_OK;
  RAGEN_REQUIRE(state && meta && atlas && out, "NULL pointer argument");
  RAGEN_REQUIRE((int64_t)H * S * W * S * 3 <= INT32_MAX, "image of %dx%d cells at S=%d is too large", H, W, S);
  DeviceGuard guard(ctx->device);
  RgbParams p{m, H, W, S, state, reinterpret_cast<const EnvMetaRgb*>(meta), atlas, rows, out};
  const unsigned grid = (unsigned)(m < (int64_t)ctx->sm_count * 8 ? m : (int64_t)ctx->sm_count * 8);
  const bool vec = (S == 16) && ((uintptr_t)atlas % 16 == 0) && ((uintptr_t)out % 16 == 0);
  if (vec) {
    const size_t smem = 7 * 16 * 16 * 3 + (size_t)((H * W + 15) & ~15);
    sokoban_rgb_vec_kernel<16><<<grid, RGB_T, smem, (cudaStream_t)stream>>>(p);
  } else {
    const size_t smem = (size_t)((7 * S * S * 3 + 15) & ~15) + (size_t)((H * W + 15) & ~15);
    RAGEN_REQUIRE(smem <= 227 * 1024, "atlas + room do not fit shared memory (%zu B)", smem);
    RAGEN_CUDA(cudaFuncSetAttribute(sokoban_rgb_byte_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    sokoban_rgb_byte_kernel<<<grid, RGB_T, smem, (cudaStream_t)stream>>>(p);
  }
  RAGEN_LAUNCHED(ctx);
  return RAGEN_OK;
}
