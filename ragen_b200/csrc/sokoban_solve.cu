// Sokoban solvability / shortest push solution on the GPU (sm_100a).
//
// Replaces get_shortest_action_path (ragen/env/sokoban/utils.py:11-101) for a batch of rooms - the reference's own validity
// check of a generated room ("get_shortest_action_path finds a solution").  The per-room algorithm is
// ssolve::solve (sokoban_solve_logic.cuh): the reference's breadth-first search on compact states, discovery order kept, so the
// PATH returned is the reference's, not merely its length.
//
// Parallelisation: one thread per room (the search is sequential when the path has to match), 32 rooms of a warp in lock step,
// their hash sets / queues / parent arrays interleaved element by element in HBM (element i of lane l at [i*32 + l]) so that
// the table clear and the queue walk of a warp touch whole 256-byte lines.  Warps are persistent and take 32 consecutive
// rooms at a time from an atomic counter.  Latency-bound pointer chasing in L2/HBM; no roofline claim (output is a few bytes
// per room), rooms/s is what bench.py --rows reports.
#include "common.cuh"
#include "sokoban_solve_logic.cuh"

namespace {

constexpr int SOLVE_T = 64;  // threads per CTA

struct SolveParams {
  int64_t n;
  int H, W, max_depth;
  const uint8_t* fixed;
  const uint8_t* state;
  int32_t* status;
  int32_t* length;
  uint8_t* path;
  int32_t* explored;
  uint8_t* ws;
  size_t warp_bytes, off_queue, off_parent;
  uint32_t cap_table, max_states;
  unsigned long long* counter;
};

__global__ void __launch_bounds__(SOLVE_T) sokoban_solve_kernel(const SolveParams p) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = ((int64_t)blockIdx.x * SOLVE_T + threadIdx.x) >> 5;
  uint8_t* wb = p.ws + (size_t)warp * p.warp_bytes;
  const ssolve::Workspace ws{reinterpret_cast<uint64_t*>(wb) + lane, reinterpret_cast<uint64_t*>(wb + p.off_queue) + lane,
                             reinterpret_cast<uint32_t*>(wb + p.off_parent) + lane, p.cap_table, p.max_states, 32u};
  const int HW = p.H * p.W;
  for (;;) {
    unsigned long long first = 0;
    if (lane == 0) first = atomicAdd(p.counter, 32ull);
    first = __shfl_sync(0xffffffffu, first, 0);
    if ((int64_t)first >= p.n) break;
    const int64_t room = (int64_t)first + lane;
    if (room < p.n) {
      const uint8_t* F = p.fixed + room * HW;
      int len = 0, explored = 0;
      const int st = ssolve::solve([&](int cell) { return (int)__ldg(F + cell); }, p.state + room * HW, p.H, p.W, p.max_depth, ws,
                                   p.path ? p.path + room * p.max_depth : nullptr, 1, &len, &explored);
      p.status[room] = st;
      p.length[room] = len;
      if (p.explored) p.explored[room] = explored;
    }
    __syncwarp();
  }
}

size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

}  // namespace

extern "C" int ragen_sokoban_solve(ragen_ctx* ctx, int64_t n, int H, int W, const uint8_t* fixed, const uint8_t* state, int max_depth,
                                   int64_t max_states, int32_t* status, int32_t* length, uint8_t* path, int32_t* explored,
                                   void* stream) {
  RAGEN_REQUIRE(ctx, "ctx is NULL");
  RAGEN_REQUIRE(n >= 0 && H >= 3 && W >= 3 && H * W <= 1024, "bad shape n=%lld H=%d W=%d (H*W <= 1024)", (long long)n, H, W);
  RAGEN_REQUIRE(max_depth >= 0 && max_depth <= 65535, "max_depth=%d out of range", max_depth);
  RAGEN_REQUIRE(max_states >= 0 && max_states <= ((int64_t)1 << 26), "max_states=%lld out of range (0 = from the room size)", (long long)max_states);
  RAGEN_REQUIRE(fixed && state && status && length, "NULL pointer argument");
  if (n == 0) return RAGEN_OK;
  if (max_states == 0) max_states = (int64_t)1 << 17;  // > 64 * C(63, 2): every state of a 10x10 room with two boxes
  DeviceGuard guard(ctx->device);
  cudaStream_t st = (cudaStream_t)stream;
  SolveParams p{};
  p.n = n;
  p.H = H;
  p.W = W;
  p.max_depth = max_depth;
  p.fixed = fixed;
  p.state = state;
  p.status = status;
  p.length = length;
  p.path = path;
  p.explored = explored;
  p.max_states = (uint32_t)max_states;
  p.cap_table = 2;
  while (p.cap_table < 2 * (uint64_t)max_states) p.cap_table <<= 1;
  p.off_queue = align_up((size_t)p.cap_table * 8 * 32, 512);
  p.off_parent = align_up(p.off_queue + (size_t)max_states * 8 * 32, 512);
  p.warp_bytes = align_up(p.off_parent + (size_t)max_states * 4 * 32, 512);
  int64_t warps = (int64_t)ctx->sm_count * 8 * (SOLVE_T / 32);
  const int64_t want = (n + 31) / 32;
  if (warps > want) warps = want;
  const int64_t cap = (int64_t)(((size_t)8 << 30) / p.warp_bytes);
  if (warps > cap) warps = cap;
  RAGEN_REQUIRE(warps >= 1, "solver workspace for one warp of rooms (%zu B) exceeds 8 GiB; lower max_states", p.warp_bytes);
  const int64_t ctas = (warps + SOLVE_T / 32 - 1) / (SOLVE_T / 32);
  const size_t ws_bytes = 256 + (size_t)ctas * (SOLVE_T / 32) * p.warp_bytes;
  uint8_t* ws = (uint8_t*)ragen_gen_workspace(ctx, ws_bytes);
  if (!ws) return RAGEN_ENOMEM;
  p.counter = reinterpret_cast<unsigned long long*>(ws);
  p.ws = ws + 256;
  RAGEN_CUDA(cudaMemsetAsync(p.counter, 0, sizeof(unsigned long long), st));
  sokoban_solve_kernel<<<(unsigned)ctas, SOLVE_T, 0, st>>>(p);
  RAGEN_LAUNCHED(ctx);
  return RAGEN_OK;
}
