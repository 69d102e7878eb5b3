// Sokoban batched reset: the reverse-play room generator on the GPU (sm_100a).
//
// Replaces SokobanEnv.reset -> generate_room (ragen/env/sokoban/env.py:28-42, ragen/env/sokoban/utils.py:221-278 and
// everything it calls: :152-213, :281-560) for a batch of seeds.  The per-room algorithm lives in
// sokoban_gen_logic.cuh and is stream- and order-exact, so room i is the room the reference generates for seeds[i].
//
// Parallelisation: rooms are independent and each search is inherently sequential (a depth-limited DFS whose global
// visited set makes the result depend on visiting order), so one THREAD owns one room at a time and the grid is a
// pool of persistent threads that pull rooms from an atomic counter (search lengths vary by 1000x between seeds).
// Each pool slot owns a private workspace in HBM: two MT19937 states, the DFS stack and an exact visited bitmap
// over (player, {(box, pulled)}) keys - 128 KB for 10x10 / 2 boxes.  The search is latency-bound pointer chasing
// into that bitmap, hidden by the number of rooms in flight; outputs are 2*H*W bytes per room.
#include "common.cuh"
#include "sokoban_gen_logic.cuh"

namespace {

struct GenParams {
  int64_t n;
  sgen::Geometry g;
  int search_depth, num_gen_steps;
  const uint64_t* seeds;
  uint8_t* fixed;
  uint8_t* state;
  int32_t* status;
  int32_t* explored;
  uint8_t* ws;
  size_t slot_bytes, off_np, off_stack, off_scratch, off_vis;
  uint64_t vis_words;
  uint32_t vis_mask;
  int vis_mode;
  unsigned long long* counter;
};

__global__ void __launch_bounds__(64) sokoban_generate_kernel(const GenParams p) {
  const size_t slot = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  uint8_t* ws = p.ws + slot * p.slot_bytes;
  sgen::Work w;
  w.mt_py = reinterpret_cast<uint32_t*>(ws);
  w.mt_np = reinterpret_cast<uint32_t*>(ws + p.off_np);
  w.stack = reinterpret_cast<sgen::Frame*>(ws + p.off_stack);
  w.scratch = ws + p.off_scratch;
  w.vis.mode = p.vis_mode;
  w.vis.mask = p.vis_mask;
  w.vis.mem = reinterpret_cast<uint32_t*>(ws + p.off_vis);
  w.vis.mem64 = reinterpret_cast<uint64_t*>(ws + p.off_vis);
  w.vis_words = p.vis_words;
  const int HW = p.g.H * p.g.W;
  for (;;) {
    const unsigned long long room = atomicAdd(p.counter, 1ull);
    if (room >= (unsigned long long)p.n) break;
    const uint64_t seed = p.seeds[room];
    int st, explored = 0;
    if (seed >> 32) {
      st = 3;  // np.random.seed rejects seeds >= 2^32 (ValueError in the reference)
    } else {
      st = sgen::generate_room(p.g, p.search_depth, p.num_gen_steps, seed, w, p.fixed + room * HW, p.state + room * HW, &explored);
    }
    p.status[room] = st;
    if (p.explored) p.explored[room] = explored;
  }
}

size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

}  // namespace

extern "C" int ragen_sokoban_generate(ragen_ctx* ctx, int64_t n, int H, int W, int num_boxes, int search_depth,
                                      int num_gen_steps, const uint64_t* seeds, uint8_t* fixed, uint8_t* state,
                                      int32_t* status, int32_t* explored, void* stream) {
  RAGEN_REQUIRE(ctx, "ctx is NULL");
  RAGEN_REQUIRE(n >= 0 && H >= 4 && W >= 4 && H * W <= sgen::MAX_CELLS, "bad shape n=%lld H=%d W=%d (H*W <= %d)", (long long)n, H,
                W, sgen::MAX_CELLS);
  RAGEN_REQUIRE(num_boxes >= 1 && num_boxes <= sgen::MAX_BOXES, "num_boxes=%d out of range (1..%d)", num_boxes, sgen::MAX_BOXES);
  RAGEN_REQUIRE(search_depth >= 1 && search_depth <= 65535 && num_gen_steps >= 0, "bad search_depth / num_gen_steps");
  RAGEN_REQUIRE(seeds && fixed && state && status, "NULL pointer argument");
  if (n == 0) return RAGEN_OK;
  DeviceGuard guard(ctx->device);
  cudaStream_t st = (cudaStream_t)stream;
  GenParams p{};
  p.n = n;
  p.g = sgen::Geometry{H, W, num_boxes, 0};
  const int cells = (H - 2) * (W - 2);
  while ((1 << p.g.cell_bits) < cells) ++p.g.cell_bits;
  p.search_depth = search_depth;
  p.num_gen_steps = num_gen_steps;
  p.seeds = seeds;
  p.fixed = fixed;
  p.state = state;
  p.status = status;
  p.explored = explored;
  const int kb = sgen::key_bits(p.g);
  size_t vis_bytes;
  if (kb <= 24) {
    p.vis_mode = 0;
    p.vis_words = ((uint64_t)1 << kb) / 32 + 1;
    vis_bytes = p.vis_words * 4;
  } else {
    p.vis_mode = 1;
    p.vis_words = (uint64_t)1 << 20;  // > 3x the 300000-state cap of the search
    p.vis_mask = (uint32_t)(p.vis_words - 1);
    vis_bytes = p.vis_words * 8;
  }
  p.off_np = sizeof(uint32_t) * mt::N;
  p.off_stack = 2 * p.off_np;
  p.off_scratch = align_up(p.off_stack + sizeof(sgen::Frame) * (size_t)(search_depth + 1), 16);
  p.off_vis = align_up(p.off_scratch + 2 * (size_t)H * W, 16);
  p.slot_bytes = align_up(p.off_vis + vis_bytes, 128);
  // pool size: enough threads to hide the bitmap latency, bounded by the rooms to do and by 8 GiB of workspace
  const int threads = 64;
  int64_t slots = (int64_t)ctx->sm_count * 512;
  if (slots > n) slots = n;
  const int64_t cap = (int64_t)(((size_t)8 << 30) / p.slot_bytes);
  if (slots > cap) slots = cap;
  slots = (slots + threads - 1) / threads * threads;
  const size_t ws_bytes = 256 + (size_t)slots * p.slot_bytes;
  uint8_t* ws = (uint8_t*)ragen_gen_workspace(ctx, ws_bytes);
  if (!ws) return RAGEN_ENOMEM;
  p.counter = reinterpret_cast<unsigned long long*>(ws);
  p.ws = ws + 256;
  RAGEN_CUDA(cudaMemsetAsync(p.counter, 0, sizeof(unsigned long long), st));
  sokoban_generate_kernel<<<(unsigned)(slots / threads), threads, 0, st>>>(p);
  RAGEN_LAUNCHED(ctx);
  return RAGEN_OK;
}
