// Sokoban batched reset: the reverse-play room generator on the GPU (sm_100a).
//
// Replaces SokobanEnv.reset -> generate_room (ragen/env/sokoban/env.py:28-42, ragen/env/sokoban/utils.py:221-278 and
// everything it calls: :152-213, :281-560) for a batch of seeds.  The per-room algorithm lives in
// sokoban_gen_logic.cuh and is stream- and order-exact, so room i is the room the reference generates for seeds[i].
//
// Parallelisation: rooms are independent and each search is inherently sequential (a depth-limited DFS whose global
// visited set makes the result depend on visiting order), so one THREAD owns one room - and the code is shaped so that
// the 32 rooms of a warp run in lock step instead of one after the other:
//   * rooms are handed out LANE BY LANE from an atomic counter: a warp alternates between a set-up phase, where the lanes that
//     owe a try walk seed -> topology -> placement -> clear together, and a search phase that ends as soon as a few searches
//     have ended, so a lane with a short search is given its next room while its neighbours' long searches go on (a room's
//     search explores 2 400 states on average and 25 000 at worst: waiting for the longest of 32 left 5 of 32 lanes busy);
//   * the MT19937 states of a warp are interleaved word by word (lane l's word i at [i][l]), so the seeding and twist
//     loops, whose trip counts are the same for every room, touch whole 128-byte lines;
//   * the visited bitmaps (128 KB per room for 10x10 / 2 boxes) are cleared by the whole warp, 512 contiguous bytes per
//     instruction, for the lanes that are about to search;
//   * the search is sgen::reverse_play_fast: the node lives in registers, a level costs a ONE-BYTE undo record in shared
//     memory, walls are a bitmask in shared memory, cell -> (interior index, row, col) are CTA-wide tables, and one loop
//     iteration tries one child or pops one level without inner loops, so divergence inside a warp is a few predicated
//     branches per iteration.  What remains per iteration is one dependent probe of the room's bitmap in L2/HBM, hidden
//     by the other warps of the SM.
// Outputs are 2*H*W bytes per room.  No roofline claim: rooms/s and validity are what is reported.
#include "common.cuh"
#include "sokoban_gen_logic.cuh"

#include <stdlib.h>

namespace {

constexpr int GEN_T = 64;           // threads per CTA (2 warps)
constexpr int STACK_SMEM_MAX = 40 * 1024;  // undo stacks move to the HBM workspace beyond this (search_depth > 639)

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
  // a warp's block of the workspace: [mt_py 624 x 32 words][mt_np 624 x 32 words][32 lane regions of lane_bytes]; a lane
  // region = [scratch 2*H*W][undo stack, only when it does not fit shared memory][visited set]
  size_t warp_bytes, off_lanes, lane_bytes, off_stack, off_vis, vis_bytes;
  uint32_t vis_mask;
  int vis_mode, stack_in_smem;
  int refill_at;  // searches of a warp that must have ended before it stops to hand their lanes new rooms
  unsigned long long* counter;
};


// One warp, NB boxes, bitmap or hash visited set.  A lane is IDLE (wants a room), owes a TRY (topology -> placement -> clear ->
// search_begin; up to four per room, utils.py:244) or is in a SEARCH.  The warp alternates between a set-up phase, in which
// the lanes that owe a try run it in lock step (the MT19937 seeding and the visited-set clears are the expensive, uniform
// parts), and a search phase that runs until `refill_at` searches have ended (any, once the queue is empty) - so a lane whose
// search is short gets its next room while the long searches of its neighbours go on, instead of idling until the longest of
// 32 ends.  Rooms are handed out lane by lane from one atomic counter.
template <int NB, bool BITMAP>
__device__ __forceinline__ void generate_warp(const GenParams& p, const sgen::Geometry& g, sgen::Fast& fm, sgen::Visited& vis,
                                              uint8_t* wws, uint8_t* floor, uint8_t* vis_mem, uint32_t* s_wall_lane, const int lane) {
  const unsigned FULL = 0xffffffffu;
  const int HW = g.H * g.W;
  enum { IDLE = 0, TRY = 1, SEARCH = 2 };
  int phase = IDLE, tries = 0, st = 0, explored = 0, player = 0;
  int64_t room = 0;
  bool fresh = false, queue_empty = false;
  uint8_t targets[sgen::MAX_BOXES] = {0, 0, 0, 0};
  sgen::BestFast best;
  best.disp = 0;
  best.p = 0;
  best.boxes = 0;
  best.score = -1;
  sgen::Search<NB> S;
  S.depth = -1;
  mt::State py{reinterpret_cast<uint32_t*>(wws) + lane, 0, 32};
  mt::State np{reinterpret_cast<uint32_t*>(wws) + mt::N * 32 + lane, 0, 32};

  for (;;) {
    // ---- idle lanes take the next rooms ---------------------------------------------------------------------------
    const unsigned idle = __ballot_sync(FULL, phase == IDLE);
    if (idle && !queue_empty) {
      const int cnt = __popc(idle);
      unsigned long long base = 0;
      if (lane == 0) base = atomicAdd(p.counter, (unsigned long long)cnt);
      base = __shfl_sync(FULL, base, 0);
      if (phase == IDLE) {
        const unsigned long long r = base + (unsigned)__popc(idle & ((1u << lane) - 1u));
        if (r < (unsigned long long)p.n) {
          room = (int64_t)r;
          phase = TRY;
          fresh = true;
          tries = 0;
        }
      }
      if (base + (unsigned long long)cnt >= (unsigned long long)p.n) queue_empty = true;
    }
    // ---- lanes that owe a try: seed (new rooms), topology, placement, clear, enter the root ----------------------------
    if (__any_sync(FULL, phase == TRY)) {
      if (phase == TRY && fresh) {
        const uint64_t seed = p.seeds[room];
        mt::seed_python(py, seed);
        mt::seed_numpy(np, (uint32_t)seed);
        fresh = false;
        st = (int)sgen::OK;
        explored = 0;
        best.disp = 0;
        if ((seed >> 32) != 0) {  // np.random.seed rejects seeds >= 2^32 (ValueError in the reference)
          p.status[room] = 3;
          if (p.explored) p.explored[room] = 0;
          phase = IDLE;
        }
      }
      bool search = false;
      uint8_t* fixed = p.fixed + room * HW;
      if (phase == TRY) {
        sgen::topology(py, g, p.num_gen_steps, 0.35, floor);
        if (sgen::place(np, g, floor, fixed, player, targets)) {
          search = true;
        } else {  // 'Not enough free spots': generate_room raises before reporting the previous try's count
          p.status[room] = (int)sgen::NO_SPACE;
          if (p.explored) p.explored[room] = 0;
          phase = IDLE;
        }
      }
      __syncwarp();
      // clear the visited sets of the lanes that search, all 32 lanes on one lane's set at a time
      unsigned need = __ballot_sync(FULL, search);
      while (need) {
        const int l = __ffs((int)need) - 1;
        need &= need - 1;
        uint4* dst = reinterpret_cast<uint4*>(reinterpret_cast<uint8_t*>(__shfl_sync(FULL, (unsigned long long)vis_mem, l)));
        const size_t vecs = p.vis_bytes >> 4;
        for (size_t i = lane; i < vecs; i += 32) dst[i] = make_uint4(0u, 0u, 0u, 0u);
      }
      __syncwarp();  // orders the other lanes' zeroes before this lane's probes
      if (search) {
        sgen::fill_walls(g, fixed, s_wall_lane, GEN_T);
        sgen::search_begin<NB, BITMAP>(S, g, fm, sgen::pack_cells(targets, g.nb), player, p.search_depth, vis, best, true);
        phase = SEARCH;
        ++tries;
      }
    }
    // ---- search: one child tried or one level popped per iteration, until enough searches have ended --------------------
    const unsigned searching = __ballot_sync(FULL, phase == SEARCH);
    if (searching == 0) {
      if (queue_empty) break;  // nobody searches, nobody owes a try (they would be searching now), nothing left to hand out
      continue;
    }
    const int stop_at = queue_empty ? 1 : (__popc(searching) < p.refill_at ? __popc(searching) : p.refill_at);
    for (;;) {
      const bool running = phase == SEARCH && S.depth >= 0;
      const unsigned run = __ballot_sync(FULL, running);
      if (__popc(searching & ~run) >= stop_at) break;
      if (running) sgen::search_step<NB, BITMAP>(S, g, fm, p.search_depth, vis, best);
    }
    // ---- lanes whose search has ended: done, or another try ----------------------------------------------------------------
    if (phase == SEARCH && S.depth < 0) {
      explored = S.explored;
      if (best.disp > 0 || tries >= 4) {
        if (best.disp == 0) {
          st = (int)sgen::SCORE_ZERO;  // RuntimeWarning('Generated Model with score == 0')
        } else {
          sgen::finish_room(py, g, p.fixed + room * HW, p.state + room * HW, best);
        }
        p.status[room] = st;
        if (p.explored) p.explored[room] = explored;
        phase = IDLE;
      } else {
        phase = TRY;
      }
    }
    __syncwarp();
  }
}

__global__ void __launch_bounds__(GEN_T) sokoban_generate_kernel(const GenParams p) {
  extern __shared__ __align__(16) uint8_t smem[];
  // CTA-wide tables, then per-thread wall nibbles [ceil(H*W/8)][GEN_T] words, then the undo stacks [depth][GEN_T] bytes
  const sgen::Geometry g = p.g;
  const int HW = g.H * g.W;
  uint16_t* s_rc = reinterpret_cast<uint16_t*>(smem);
  uint32_t* s_wall = reinterpret_cast<uint32_t*>(smem + 512);
  uint8_t* s_interior = smem + 512 + ((HW + 7) / 8) * GEN_T * 4;
  uint8_t* s_stack = s_interior + sgen::MAX_CELLS;
  for (int cell = threadIdx.x; cell < HW; cell += GEN_T) {
    const int r = cell / g.W, c = cell - r * g.W;
    s_interior[cell] = (uint8_t)((r - 1) * (g.W - 2) + (c - 1));
    s_rc[cell] = (uint16_t)((r << 8) | c);
  }
  __syncthreads();

  const int lane = threadIdx.x & 31;
  const size_t warp_slot = ((size_t)blockIdx.x * GEN_T + threadIdx.x) >> 5;
  uint8_t* wws = p.ws + warp_slot * p.warp_bytes;
  uint8_t* lws = wws + p.off_lanes + (size_t)lane * p.lane_bytes;
  uint8_t* floor = lws;
  uint8_t* vis_mem = lws + p.off_vis;
  sgen::Visited vis;
  vis.mode = p.vis_mode;
  vis.mask = p.vis_mask;
  vis.mem = reinterpret_cast<uint32_t*>(vis_mem);
  vis.mem64 = reinterpret_cast<uint64_t*>(vis_mem);
  sgen::Fast fm;
  fm.interior = s_interior;
  fm.rc = s_rc;
  fm.wall = s_wall + threadIdx.x;
  fm.wall_stride = GEN_T;
  fm.stack = p.stack_in_smem ? s_stack + threadIdx.x : lws + p.off_stack;
  fm.stack_stride = p.stack_in_smem ? GEN_T : 1;
  uint32_t* wl = s_wall + threadIdx.x;
  // every lane of the launch takes the same case (the number of boxes and the kind of visited set are launch-wide)
  if (p.vis_mode == 0) {
    switch (g.nb) {
      case 1: generate_warp<1, true>(p, g, fm, vis, wws, floor, vis_mem, wl, lane); break;
      case 2: generate_warp<2, true>(p, g, fm, vis, wws, floor, vis_mem, wl, lane); break;
      case 3: generate_warp<3, true>(p, g, fm, vis, wws, floor, vis_mem, wl, lane); break;
      default: generate_warp<4, true>(p, g, fm, vis, wws, floor, vis_mem, wl, lane); break;
    }
  } else {
    switch (g.nb) {
      case 1: generate_warp<1, false>(p, g, fm, vis, wws, floor, vis_mem, wl, lane); break;
      case 2: generate_warp<2, false>(p, g, fm, vis, wws, floor, vis_mem, wl, lane); break;
      case 3: generate_warp<3, false>(p, g, fm, vis, wws, floor, vis_mem, wl, lane); break;
      default: generate_warp<4, false>(p, g, fm, vis, wws, floor, vis_mem, wl, lane); break;
    }
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
  if (kb <= 24) {
    p.vis_mode = 0;
    p.vis_bytes = align_up((((size_t)1 << kb) / 32 + 1) * 4, 512);
  } else {
    p.vis_mode = 1;
    const size_t slots = (size_t)1 << 20;  // > 3x the 300000-state cap of the search
    p.vis_mask = (uint32_t)(slots - 1);
    p.vis_bytes = slots * 8;
  }
  const size_t stack_bytes = (size_t)(search_depth + 1);
  p.stack_in_smem = stack_bytes * GEN_T <= STACK_SMEM_MAX;
  p.off_stack = align_up(2 * (size_t)H * W, 16);
  p.off_vis = align_up(p.off_stack + (p.stack_in_smem ? 0 : stack_bytes), 512);
  p.lane_bytes = p.off_vis + p.vis_bytes;
  p.off_lanes = 2 * sizeof(uint32_t) * mt::N * 32;
  p.warp_bytes = p.off_lanes + 32 * p.lane_bytes;
  const size_t smem = 512 + (size_t)((H * W + 7) / 8) * GEN_T * 4 + sgen::MAX_CELLS + (p.stack_in_smem ? stack_bytes * GEN_T : 0);
  // pool size: warps of 32 rooms; as many as fit an SM by shared memory (at most 16 warps per SM), bounded by the rooms
  // to do and by 8 GiB of workspace
  int64_t ctas_per_sm = (int64_t)(200 * 1024 / smem);
  if (ctas_per_sm > 8) ctas_per_sm = 8;
  if (ctas_per_sm < 1) ctas_per_sm = 1;
  int64_t warps = (int64_t)ctx->sm_count * ctas_per_sm * (GEN_T / 32);
  const int64_t want = (n + 31) / 32;
  if (warps > want) warps = want;
  const int64_t cap = (int64_t)(((size_t)8 << 30) / p.warp_bytes);
  if (warps > cap) warps = cap;
  RAGEN_REQUIRE(warps >= 1, "generator workspace for one warp of rooms (%zu B) exceeds 8 GiB", p.warp_bytes);
  const int64_t ctas = (warps + GEN_T / 32 - 1) / (GEN_T / 32);
  const size_t ws_bytes = 256 + (size_t)ctas * (GEN_T / 32) * p.warp_bytes;
  uint8_t* ws = (uint8_t*)ragen_gen_workspace(ctx, ws_bytes);
  if (!ws) return RAGEN_ENOMEM;
  p.refill_at = 8;
  if (const char* e = getenv("RAGEN_GEN_REFILL")) {  // probing only
    const int v = atoi(e);
    if (v >= 1 && v <= 32) p.refill_at = v;
  }
  p.counter = reinterpret_cast<unsigned long long*>(ws);
  p.ws = ws + 256;
  RAGEN_CUDA(cudaMemsetAsync(p.counter, 0, sizeof(unsigned long long), st));
  RAGEN_CUDA(cudaFuncSetAttribute(sokoban_generate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  sokoban_generate_kernel<<<(unsigned)ctas, GEN_T, smem, st>>>(p);
  RAGEN_LAUNCHED(ctx);
  return RAGEN_OK;
}
