// Response parsing, action lookup and Countdown scoring kernels (sm_100a).
//
// Replaces ContextManager._parse_response (ragen/llm_agent/ctx_manager.py:148-173, called per response from
// get_env_inputs :332-352), EnvStateManager._extract_map_valid_actions (ragen/llm_agent/es_manager.py:230-240) and
// CountdownEnv.compute_reward (ragen/env/countdown/env.py:69-78).
//
// Layout: all responses of a batch are one UTF-8 byte blob in HBM with int64 offsets [n+1].  One warp owns one
// response: the 32 lanes read it in 32-byte steps (coalesced), ballot the '<' positions, classify them against the
// six special strings and feed the occurrences in text order to the regex state machine (parse_logic.cuh), which
// every lane advances identically - no divergence, no shared memory.  The short tail (strip, split on the action
// separator, table lookup) runs uniformly on the matched answer span.  Cleaned contents that differ from a
// sub-span of the input (special strings inside the groups: rare) are written to a caller-provided scratch blob of
// the same size.  Byte-scan work: HBM-bound (each response byte is read once), no tensor cores.
#include "common.cuh"
#include "parse_logic.cuh"

namespace {

constexpr int PWARPS = 4;
using parse::MAX_ACTIONS;

struct ParseParams {
  int64_t n;
  const uint8_t* text;
  const int64_t* offsets;
  int enable_think, sep_len, max_actions;
  uint8_t sep[8];
  const uint8_t* names;
  const int32_t* name_off;
  const uint8_t* codes;
  const int32_t* table_first;
  int n_tables;
  const int32_t* table_id;
  uint8_t* scratch;
  uint8_t* flags;
  int32_t* spans;
  int32_t* act_spans;
  uint8_t* n_parsed;
  int32_t* n_total;
  uint8_t* codes_out;
  uint8_t* n_valid;
};

// any special string starting inside text[s, e)?  (warp-cooperative)
__device__ __forceinline__ bool warp_has_special(const uint8_t* text, int s, int e, int lane) {
  bool any = false;
  for (int b = s; b < e; b += 32) {
    const int i = b + lane;
    const bool hit = i < e && text[i] == '<' && parse::classify_tag(text + i, text + e) != parse::T_NONE;
    if (__ballot_sync(0xffffffffu, hit)) {
      any = true;
      break;
    }
  }
  return any;
}

__global__ void __launch_bounds__(PWARPS * 32) parse_kernel(const ParseParams p) {
  const int lane = threadIdx.x & 31;
  const int64_t r = (int64_t)blockIdx.x * PWARPS + (threadIdx.x >> 5);
  if (r >= p.n) return;
  const int64_t off = p.offsets[r];
  const int n = (int)(p.offsets[r + 1] - off);
  const uint8_t* text = p.text + off;

  // ---- regex search over the tag occurrences ----------------------------------------------------------------
  parse::Matcher m;
  parse::matcher_init(m, p.enable_think != 0);
  for (int b = 0; b < n && m.state != 3; b += 32) {
    const int i = b + lane;
    int tag = parse::T_NONE;
    if (i < n && text[i] == '<') tag = parse::classify_tag(text + i, text + n);
    unsigned hits = __ballot_sync(0xffffffffu, tag != parse::T_NONE);
    while (hits) {
      const int l = __ffs(hits) - 1;
      hits &= hits - 1;
      const int t = __shfl_sync(0xffffffffu, tag, l);
      if (parse::matcher_feed(m, text, n, b + l, t)) break;
    }
  }

  uint8_t fl = 0;
  int think_s = 0, think_e = 0, ans_s = 0, ans_e = 0;
  parse::ParseTail tail;
  tail.total = tail.kept = tail.nvalid = 0;
  if (m.state == 3) {
    fl = 1;
    think_s = m.think_s;
    think_e = m.think_e;
    ans_s = m.ans_s;
    ans_e = m.ans_e;
    const uint8_t* abuf = text;
    // ---- strip / remove special strings from both groups (:162-164) ------------------------------------------
    for (int which = 0; which < 2; ++which) {
      int s = which ? ans_s : think_s, e = which ? ans_e : think_e;
      if (which == 0 && !p.enable_think) continue;
      if (warp_has_special(text, s, e, lane)) {
        uint8_t* scr = p.scratch + off;
        for (int i = s + lane; i < e; i += 32) scr[i] = text[i];
        __syncwarp();
        if (lane == 0) parse::clean_content(scr, s, e);
        __syncwarp();
        s = __shfl_sync(0xffffffffu, s, 0);
        e = __shfl_sync(0xffffffffu, e, 0);
        fl |= which ? 4 : 2;
        if (which) abuf = scr;
      } else {
        parse::strip(text, s, e);
      }
      if (which) {
        ans_s = s;
        ans_e = e;
      } else {
        think_s = s;
        think_e = e;
      }
    }
    // ---- split on the action separator, keep max_actions (:166-171), map through the lookup table ------------
    const int tid = p.table_id ? p.table_id[r] : (p.n_tables > 0 ? 0 : -1);
    parse::ActionTable tab{p.names, p.name_off, p.codes, 0, -1};
    if (tid >= 0 && tid < p.n_tables) {
      tab.first = p.table_first[tid];
      tab.count = p.table_first[tid + 1] - tab.first;
      fl |= 16;  // codes / n_valid are meaningful
    }
    parse::parse_finish(abuf, ans_s, ans_e, p.sep, p.sep_len, p.max_actions, tab, tail);
    if (tail.total > p.max_actions) fl |= 8;
  }
  if (lane == 0) {
    p.flags[r] = fl;
    int32_t* sp = p.spans + r * 4;
    sp[0] = think_s;
    sp[1] = think_e - think_s;
    sp[2] = ans_s;
    sp[3] = ans_e - ans_s;
    p.n_parsed[r] = (uint8_t)tail.kept;
    p.n_total[r] = tail.total;
    p.n_valid[r] = (uint8_t)tail.nvalid;
    for (int k = 0; k < p.max_actions; ++k) {
      p.act_spans[(r * p.max_actions + k) * 2] = k < tail.kept ? tail.spans[2 * k] : 0;
      p.act_spans[(r * p.max_actions + k) * 2 + 1] = k < tail.kept ? tail.spans[2 * k + 1] : 0;
      p.codes_out[r * p.max_actions + k] = k < tail.nvalid ? tail.codes[k] : 0;
    }
  }
}

__global__ void countdown_kernel(int64_t n, const uint8_t* __restrict__ text, const int64_t* __restrict__ start,
                                 const int32_t* __restrict__ len, const int32_t* __restrict__ nums,
                                 const int32_t* __restrict__ n_nums, const int64_t* __restrict__ target, uint8_t* result,
                                 uint8_t* unsupported) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  bool unsup = false;
  const int nn = n_nums[i];
  int res = 0;
  if (nn >= 0 && nn <= parse::CD_MAX_NUMS)
    res = parse::countdown_score(text + start[i], len[i], nums + i * parse::CD_MAX_NUMS, nn, (long long)target[i], unsup);
  result[i] = (uint8_t)res;
  if (unsupported) unsupported[i] = unsup ? 1 : 0;
}

}  // namespace

extern "C" int ragen_parse_actions(ragen_ctx* ctx, int64_t n, const uint8_t* text, const int64_t* offsets, int enable_think,
                                   const char* action_sep /*[host]*/, int max_actions, const uint8_t* names,
                                   const int32_t* name_off, const uint8_t* codes, const int32_t* table_first, int n_tables,
                                   const int32_t* table_id, uint8_t* scratch, uint8_t* flags, int32_t* spans,
                                   int32_t* act_spans, uint8_t* n_parsed, int32_t* n_total, uint8_t* codes_out,
                                   uint8_t* n_valid, void* stream) {
  RAGEN_REQUIRE(ctx, "ctx is NULL");
  RAGEN_REQUIRE(n >= 0 && max_actions >= 1 && max_actions <= MAX_ACTIONS, "bad argument n=%lld max_actions=%d (1..%d)",
                (long long)n, max_actions, MAX_ACTIONS);
  RAGEN_REQUIRE(text && offsets && action_sep && scratch && flags && spans && act_spans && n_parsed && n_total && codes_out &&
                    n_valid,
                "NULL pointer argument");
  RAGEN_REQUIRE(n_tables == 0 || (names && name_off && codes && table_first), "lookup tables missing");
  int sep_len = 0;
  while (action_sep[sep_len]) ++sep_len;
  RAGEN_REQUIRE(sep_len >= 1 && sep_len <= 8, "action separator must be 1..8 bytes");
  if (n == 0) return RAGEN_OK;
  DeviceGuard guard(ctx->device);
  ParseParams p{};
  p.n = n;
  p.text = text;
  p.offsets = offsets;
  p.enable_think = enable_think;
  p.sep_len = sep_len;
  p.max_actions = max_actions;
  for (int i = 0; i < sep_len; ++i) p.sep[i] = (uint8_t)action_sep[i];
  p.names = names;
  p.name_off = name_off;
  p.codes = codes;
  p.table_first = table_first;
  p.n_tables = n_tables;
  p.table_id = table_id;
  p.scratch = scratch;
  p.flags = flags;
  p.spans = spans;
  p.act_spans = act_spans;
  p.n_parsed = n_parsed;
  p.n_total = n_total;
  p.codes_out = codes_out;
  p.n_valid = n_valid;
  parse_kernel<<<(unsigned)((n + PWARPS - 1) / PWARPS), PWARPS * 32, 0, (cudaStream_t)stream>>>(p);
  RAGEN_LAUNCHED(ctx);
  return RAGEN_OK;
}

extern "C" int ragen_countdown_score(ragen_ctx* ctx, int64_t n, const uint8_t* text, const int64_t* start, const int32_t* len,
                                     const int32_t* nums, const int32_t* n_nums, const int64_t* target, uint8_t* result,
                                     uint8_t* unsupported, void* stream) {
  RAGEN_REQUIRE(ctx, "ctx is NULL");
  RAGEN_REQUIRE(n >= 0, "bad n=%lld", (long long)n);
  RAGEN_REQUIRE(text && start && len && nums && n_nums && target && result, "NULL pointer argument");
  if (n == 0) return RAGEN_OK;
  DeviceGuard guard(ctx->device);
  countdown_kernel<<<(unsigned)((n + 127) / 128), 128, 0, (cudaStream_t)stream>>>(n, text, start, len, nums, n_nums, target,
                                                                                 result, unsupported);
  RAGEN_LAUNCHED(ctx);
  return RAGEN_OK;
}
