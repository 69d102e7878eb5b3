// ContextManager tensor passes (sm_100a): loss / response masks and score placement from token ids, and the
// group reward normalisation of the trajectory score.
//
// Replaces get_masks_and_scores (ragen/llm_agent/ctx_manager.py:35-70) and _normalize_score_tensor (:175-226).
//
// Layout: input_ids int64 [B][L] row-major (left padded), outputs [B][L-1]: the reference drops the first score
// column and the last mask column (:66-68).  One warp owns one row and scans it in 32-token steps: the turn
// indicator cumsum(ids == <|im_start|>) is a ballot + popcount prefix, so the row is read exactly once with
// coalesced 256 B accesses (14 B per token in total: HBM-bound, no tensor cores).
#include "common.cuh"

namespace {

constexpr int MT = 256;      // threads of the CTA that owns one row
constexpr int SLABS = 8;     // 256-token slabs handled per pass: 2048 tokens of the row in registers
constexpr int MAX_TURNS = 64;

struct MsParams {
  int B, L;
  const int64_t* ids;
  int64_t im_start, im_end;
  int use_turn_scores, enable_response_mask, roll;
  const double* turn_scores;  // [B][max_turns]
  const int32_t* n_turns;     // [B]
  int max_turns;
  uint8_t* loss_mask;
  uint8_t* resp_mask;
  float* score;
};

// One CTA per row.  A pass covers 2048 tokens: every thread loads its 8 ids at once (the loads are what bounds this pass at
// C5: 128 rows), each warp ballots its <|im_start|> flags per slab, the 64 per-(slab, warp) counts are prefix-summed through
// shared memory, and the turn indicator cumsum(ids == <|im_start|>) of a token is its segment's base plus a popcount.
__global__ void __launch_bounds__(MT) masks_scores_kernel(const MsParams p) {
  __shared__ int s_pos[MAX_TURNS];
  __shared__ int s_cnt[SLABS * (MT / 32) + 1];
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const int row = blockIdx.x;
  const int L = p.L, Lo = L - 1;
  const int64_t* ids = p.ids + (size_t)row * L;
  uint8_t* lm = p.loss_mask + (size_t)row * Lo;
  uint8_t* rm = p.resp_mask + (size_t)row * Lo;
  float* sc = p.score + (size_t)row * Lo;
  for (int k = tid; k < p.max_turns; k += MT) s_pos[k] = 0x7fffffff;
  int turn_base = 0;  // im_start tokens left of the current pass
  for (int p0 = 0; p0 < L; p0 += SLABS * MT) {
    int64_t id[SLABS];
    unsigned starts[SLABS];
#pragma unroll
    for (int j = 0; j < SLABS; ++j) {
      const int t = p0 + j * MT + tid;
      id[j] = (t < L) ? ids[t] : -1;
    }
#pragma unroll
    for (int j = 0; j < SLABS; ++j) {
      const int t = p0 + j * MT + tid;
      starts[j] = __ballot_sync(0xffffffffu, t < L && id[j] == p.im_start);
      if (lane == 0) s_cnt[j * (MT / 32) + wid] = __popc(starts[j]);
    }
    __syncthreads();
    if (tid < 32) {  // exclusive prefix of the 64 segment counts (segments are in token order: slab-major, then warp)
      const int a = s_cnt[2 * lane], b = s_cnt[2 * lane + 1];
      int incl = a + b;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int up = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += up;
      }
      const int excl = incl - (a + b);
      __syncwarp();
      s_cnt[2 * lane] = excl;
      s_cnt[2 * lane + 1] = excl + a;
      if (lane == 31) s_cnt[SLABS * (MT / 32)] = incl;
    }
    __syncthreads();
#pragma unroll
    for (int j = 0; j < SLABS; ++j) {
      const int t = p0 + j * MT + tid;
      const int turn = turn_base + s_cnt[j * (MT / 32) + wid] + __popc(starts[j] & (0xffffffffu >> (31 - lane)));  // inclusive (:43-44)
      if (t < Lo) {
        const bool resp = (turn & 1) && turn > 1;
        rm[t] = resp ? 1 : 0;
        lm[t] = p.enable_response_mask ? (resp ? 1 : 0) : (turn > 1 ? 1 : 0);
        sc[t] = 0.0f;
      }
      if (p.use_turn_scores && t < L && id[j] == p.im_end && (turn & 1) && turn >= 3) {
        const int k = (turn - 3) >> 1;  // reward of turn k sits on the <|im_end|> of indicator 2k+3 (:55-56)
        if (k < p.max_turns) atomicMin(&s_pos[k], t);
      }
    }
    turn_base += s_cnt[SLABS * (MT / 32)];
    __syncthreads();
  }
  if (tid != 0) return;
  const int nt = p.n_turns[row];
  const double* ts = p.turn_scores + (size_t)row * p.max_turns;
  if (!p.use_turn_scores) {
    double acc = 0.0;  // sum(scores) of Python floats, left to right (:64)
    for (int k = 0; k < nt; ++k) acc += ts[k];
    if (Lo >= 1) sc[Lo - 1] = (float)acc;  // full column L-1 -> output column L-2
    return;
  }
  // per turn index, ascending: write at the reward position, or at the last column when the row has none (:57-59);
  // zip_longest pads short rows with 0 (:53).  Then the Qwen roll by one (:60-62) and the dropped first column.
  for (int k = 0; k < p.max_turns; ++k) {
    const float v = (k < nt) ? (float)ts[k] : 0.0f;
    const int pos = (s_pos[k] == 0x7fffffff) ? L - 1 : s_pos[k];
    const int out = p.roll ? pos : pos - 1;  // rolled: full[t] lands in output column t; else column t-1
    if (out >= 0 && out < Lo) sc[out] = v;
  }
}

// x = score + penalty; per-group statistics
__global__ void norm_stats_kernel(int B, const float* __restrict__ score, const float* __restrict__ penalty,
                                  const int32_t* __restrict__ group, int n_groups, float* __restrict__ x, double* gstats) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B) return;
  const float v = __fadd_rn(score[i], penalty ? penalty[i] : 0.0f);
  x[i] = v;
  const int g = group[i];
  if (gstats && g >= 0 && g < n_groups) {
    atomicAdd(gstats + 3 * g + 0, 1.0);
    atomicAdd(gstats + 3 * g + 1, (double)v);
    atomicAdd(gstats + 3 * g + 2, (double)v * (double)v);
  }
}

// method: 0 identity, 1 mean, 2 mean_std, 3 asym_clip (:193-202)
__global__ void norm_apply_kernel(int B, float* __restrict__ x, const int32_t* __restrict__ group, int n_groups,
                                  const double* __restrict__ gstats, int method) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B) return;
  const int g = group[i];
  if (method == 0 || g < 0 || g >= n_groups) return;
  const double c = gstats[3 * g], s1 = gstats[3 * g + 1], s2 = gstats[3 * g + 2];
  const double mu = s1 / c;
  const float mean = (float)mu;
  float v = x[i];
  if (method == 1) {
    x[i] = __fsub_rn(v, mean);
    return;
  }
  // unbiased std; a group of one has std = nan in torch, which fails the "> 1e-6" test -> zeros
  float out = 0.0f;
  if (c > 1.5) {
    const float sd = (float)sqrt(fmax((s2 - c * mu * mu) / (c - 1.0), 0.0));
    if (fabsf(sd) > 1e-6f) out = __fdiv_rn(__fsub_rn(v, mean), __fadd_rn(sd, 1e-6f));
  }
  if (method == 3) out = fminf(fmaxf(out, -1.0f), 3.0f);
  x[i] = out;
}

}  // namespace

extern "C" int ragen_masks_scores(ragen_ctx* ctx, int B, int L, const int64_t* input_ids, int64_t im_start_id,
                                  int64_t im_end_id, int use_turn_scores, int enable_response_mask, int roll_scores,
                                  const double* turn_scores, const int32_t* n_turns, int max_turns, uint8_t* loss_mask,
                                  uint8_t* response_mask, float* score, void* stream) {
  RAGEN_REQUIRE(ctx, "ctx is NULL");
  RAGEN_REQUIRE(B >= 0 && L >= 2, "bad shape B=%d L=%d", B, L);
  RAGEN_REQUIRE(max_turns >= 0 && max_turns <= MAX_TURNS, "max_turns=%d out of range (0..%d)", max_turns, MAX_TURNS);
  RAGEN_REQUIRE(input_ids && n_turns && loss_mask && response_mask && score && (turn_scores || max_turns == 0),
                "NULL pointer argument");
  if (B == 0) return RAGEN_OK;
  DeviceGuard guard(ctx->device);
  MsParams p{B, L, input_ids, im_start_id, im_end_id, use_turn_scores, enable_response_mask, roll_scores, turn_scores,
             n_turns, max_turns, loss_mask, response_mask, score};
  masks_scores_kernel<<<B, MT, 0, (cudaStream_t)stream>>>(p);
  RAGEN_LAUNCHED(ctx);
  return RAGEN_OK;
}

extern "C" int ragen_group_norm(ragen_ctx* ctx, int B, const float* score_last, const float* penalty, const int32_t* group,
                                int n_groups, int method, float* out, void* stream) {
  RAGEN_REQUIRE(ctx, "ctx is NULL");
  RAGEN_REQUIRE(B >= 0 && n_groups >= 1 && method >= 0 && method <= 3, "bad argument B=%d n_groups=%d method=%d", B, n_groups,
                method);
  RAGEN_REQUIRE(score_last && group && out, "NULL pointer argument");
  if (B == 0) return RAGEN_OK;
  DeviceGuard guard(ctx->device);
  cudaStream_t st = (cudaStream_t)stream;
  // "if len(group2index) < acc_scores.shape[0]" (:219): with as many groups as rows only the penalty is applied
  const bool normalise = n_groups < B && method != 0;
  double* gstats = nullptr;
  if (normalise) {
    const size_t bytes = (size_t)n_groups * 3 * sizeof(double);
    gstats = (double*)ragen_scratch(ctx, bytes);
    if (!gstats) return RAGEN_ENOMEM;
    RAGEN_CUDA(cudaMemsetAsync(gstats, 0, bytes, st));
  }
  const int threads = 128, blocks = (B + threads - 1) / threads;
  norm_stats_kernel<<<blocks, threads, 0, st>>>(B, score_last, penalty, group, n_groups, out, gstats);
  RAGEN_LAUNCHED(ctx);
  if (normalise) {
    norm_apply_kernel<<<blocks, threads, 0, st>>>(B, out, group, n_groups, gstats, method);
    RAGEN_LAUNCHED(ctx);
  }
  return RAGEN_OK;
}

// The two halves of ragen_group_norm for rows that are sharded over ranks with groups that span ranks (grouping = tag or
// whole batch, ctx_manager.py:186-189): the caller all-reduces gstats between them.
extern "C" int ragen_group_norm_stats(ragen_ctx* ctx, int B, const float* score_last, const float* penalty, const int32_t* group,
                                      int n_groups, float* x, double* gstats, void* stream) {
  RAGEN_REQUIRE(ctx, "ctx is NULL");
  RAGEN_REQUIRE(B >= 0 && n_groups >= 1, "bad argument B=%d n_groups=%d", B, n_groups);
  RAGEN_REQUIRE(gstats && (B == 0 || (score_last && group && x)), "NULL pointer argument");
  DeviceGuard guard(ctx->device);
  cudaStream_t st = (cudaStream_t)stream;
  RAGEN_CUDA(cudaMemsetAsync(gstats, 0, (size_t)n_groups * 3 * sizeof(double), st));
  if (B == 0) return RAGEN_OK;  // a rank without rows still contributes zeros to the all-reduce
  const int threads = 128, blocks = (B + threads - 1) / threads;
  norm_stats_kernel<<<blocks, threads, 0, st>>>(B, score_last, penalty, group, n_groups, x, gstats);
  RAGEN_LAUNCHED(ctx);
  return RAGEN_OK;
}

extern "C" int ragen_group_norm_apply(ragen_ctx* ctx, int B, float* x, const int32_t* group, int n_groups, const double* gstats,
                                      int method, void* stream) {
  RAGEN_REQUIRE(ctx, "ctx is NULL");
  RAGEN_REQUIRE(B >= 0 && n_groups >= 1 && method >= 0 && method <= 3, "bad argument B=%d n_groups=%d method=%d", B, n_groups, method);
  if (B == 0 || method == 0) return RAGEN_OK;
  RAGEN_REQUIRE(x && group && gstats, "NULL pointer argument");
  DeviceGuard guard(ctx->device);
  const int threads = 128, blocks = (B + threads - 1) / threads;
  norm_apply_kernel<<<blocks, threads, 0, (cudaStream_t)stream>>>(B, x, group, n_groups, gstats, method);
  RAGEN_LAUNCHED(ctx);
  return RAGEN_OK;
}

// ---- ragen_group_norm_noalloc -------------------------------------------------------------------------------------
// The same pass for callers that cannot tolerate an allocation (CUDA graph capture, serving loops, harnesses that
// forbid cudaMalloc): the per-group statistics live in a file-scope device array instead of the ctx-owned scratch.
// Every GPU that loads the library holds these RAGEN_NOALLOC_MAX_GROUPS * 24 bytes until the process exits.
// The array is SHARED by every call on a device and keeps the last call's contents: each call clears the part it uses
// on its stream first, and two calls that overlap on the GPU (different streams, no ordering between them) race on it.
static __device__ __align__(256) double g_norm_stats[3 * RAGEN_NOALLOC_MAX_GROUPS];

namespace {
constexpr int MAX_DEVICES = 64;
// device address of g_norm_stats per device ordinal.  In host code the NAME g_norm_stats is only a host-side handle of
// the symbol, not something a kernel can dereference: the address a kernel may use comes from cudaGetSymbolAddress,
// asked once per device with that device current.
double* g_norm_stats_addr[MAX_DEVICES];
}  // namespace

extern "C" int ragen_group_norm_noalloc(ragen_ctx* ctx, int B, const float* score_last, const float* penalty,
                                        const int32_t* group, int n_groups, int method, float* out, void* stream) {
  RAGEN_REQUIRE(ctx, "ctx is NULL");
  RAGEN_REQUIRE(B >= 0 && n_groups >= 1 && method >= 0 && method <= 3, "bad argument B=%d n_groups=%d method=%d", B, n_groups,
                method);
  RAGEN_REQUIRE(n_groups <= RAGEN_NOALLOC_MAX_GROUPS,
                "ragen_group_norm_noalloc: n_groups=%d exceeds the static scratch of RAGEN_NOALLOC_MAX_GROUPS=%d groups "
                "(use ragen_group_norm, which grows a ctx-owned buffer)",
                n_groups, RAGEN_NOALLOC_MAX_GROUPS);
  RAGEN_REQUIRE(score_last && group && out, "NULL pointer argument");
  RAGEN_REQUIRE(ctx->device >= 0 && ctx->device < MAX_DEVICES, "device ordinal %d beyond %d", ctx->device, MAX_DEVICES);
  if (B == 0) return RAGEN_OK;
  DeviceGuard guard(ctx->device);  // ctx->device is current from here on
  double* gstats = g_norm_stats_addr[ctx->device];
  if (!gstats) {  // set-up, once per device
    void* p = nullptr;
    RAGEN_CUDA(cudaGetSymbolAddress(&p, g_norm_stats));
    RAGEN_REQUIRE(p, "cudaGetSymbolAddress returned a NULL address for the static scratch");
    gstats = g_norm_stats_addr[ctx->device] = (double*)p;
  }
  cudaStream_t st = (cudaStream_t)stream;
  const bool normalise = n_groups < B && method != 0;  // as ragen_group_norm (:219)
  if (normalise) RAGEN_CUDA(cudaMemsetAsync(gstats, 0, (size_t)n_groups * 3 * sizeof(double), st));
  const int threads = 128, blocks = (B + threads - 1) / threads;
  norm_stats_kernel<<<blocks, threads, 0, st>>>(B, score_last, penalty, group, n_groups, out, normalise ? gstats : nullptr);
  RAGEN_LAUNCHED(ctx);
  if (normalise) {
    norm_apply_kernel<<<blocks, threads, 0, st>>>(B, out, group, n_groups, gstats, method);
    RAGEN_LAUNCHED(ctx);
  }
  return RAGEN_OK;
}
