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
#include "score_logic.cuh"

namespace {

constexpr int PWARPS = 4;
// bytes of shared memory per warp for cleaning a span (longer spans are cleaned in the scratch blob).  Kept small on purpose: every
// byte the parser looks at after the first vector scan comes through L1, and 12 CTAs x 4 warps x 4 KiB of stage left L1 ~30 KB for
// 48 warps' responses (C3: 67 -> 59 us with 1 KiB)
constexpr int CLEAN_CAP = 1024;
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

__device__ __forceinline__ unsigned lt_mask4(unsigned word) {  // bit k set iff byte k of word is '<'
  const unsigned y = __vcmpeq4(word, 0x3c3c3c3cu) & 0x01010101u;
  return ((y * 0x01020408u) >> 24) & 0xfu;
}

// first occurrence of `tag` in buf[s, e), buf 16-byte aligned (the shared-memory stage): 512 bytes per warp step, only the
// '<' bytes are looked at.  -1 when there is none.
__device__ __forceinline__ int warp_find_tag_vec(const uint8_t* buf, int s, int e, int tag, int lane) {
  const int v1 = (e + 15) >> 4;
  for (int vb = s >> 4; vb < v1; vb += 32) {
    const int vi = vb + lane;
    int mine = 0x7fffffff;
    if (vi < v1) {
      const uint4 v = *reinterpret_cast<const uint4*>(buf + 16 * vi);
      unsigned m = lt_mask4(v.x) | (lt_mask4(v.y) << 4) | (lt_mask4(v.z) << 8) | (lt_mask4(v.w) << 12);
      const int lo = 16 * vi;
      if (lo < s) m &= 0xffffu << (s - lo);
      if (lo + 16 > e) m &= (1u << (e - lo)) - 1u;
      while (m) {
        const int k = __ffs(m) - 1;
        m &= m - 1;
        if (parse::classify_tag(buf + lo + k, buf + e) == tag) {
          mine = lo + k;
          break;
        }
      }
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) mine = min(mine, __shfl_xor_sync(0xffffffffu, mine, o));
    if (mine != 0x7fffffff) return mine;
  }
  return -1;
}

// One replace pass of ctx_manager.py:162-164 ("content.replace(tag, '')") on scr[s, e), warp-cooperative and in place:
// 32 bytes per iteration, occurrences found by ballot (they cannot overlap: every special string has its only '<' at
// the front), survivors compacted with a popcount prefix.  Reads run ahead of writes (w <= r), and every lane has
// read and classified its byte before any lane writes.  Returns the new end.
__device__ __forceinline__ int warp_remove_tag(uint8_t* scr, int s, int e, int tag, int lane, bool staged) {
  const int L = parse::tag_len(tag);
  // most passes find nothing to remove: look for the first occurrence with a loop that only reads
  int first = staged ? warp_find_tag_vec(scr, s, e, tag, lane) : -1;
  for (int b = s; !staged && b < e && first < 0; b += 32) {
    const int i = b + lane;
    const bool st = i < e && scr[i] == '<' && parse::classify_tag(scr + i, scr + e) == tag;
    const unsigned hit = __ballot_sync(0xffffffffu, st);
    if (hit) first = b + __ffs(hit) - 1;
  }
  if (first < 0) return e;
  s = first;  // nothing moves before it
  int w = s, carry = 0;  // carry: bytes of an occurrence that spill into this chunk
  for (int b = s; b < e; b += 32) {
    const int i = b + lane;
    const uint8_t c = i < e ? scr[i] : 0;
    const bool st = i < e && c == '<' && parse::classify_tag(scr + i, scr + e) == tag;
    unsigned starts = __ballot_sync(0xffffffffu, st);
    const unsigned valid = (e - b >= 32) ? 0xffffffffu : ((1u << (e - b)) - 1u);
    unsigned del = carry >= 32 ? 0xffffffffu : ((1u << carry) - 1u);
    carry = carry > 32 ? carry - 32 : 0;
    while (starts) {
      const int j = __ffs(starts) - 1;
      starts &= starts - 1;
      del |= (L >= 32 - j) ? (0xffffffffu << j) : (((1u << L) - 1u) << j);
      if (j + L > 32) carry = j + L - 32;
    }
    const unsigned keep = valid & ~del;
    __syncwarp();
    if ((keep >> lane) & 1u) scr[w + __popc(keep & ((1u << lane) - 1u))] = c;
    w += __popc(keep);
    __syncwarp();
  }
  return w;
}

// one response, one warp; `stage` is this warp's CLEAN_CAP bytes of shared memory
__device__ __forceinline__ void parse_one(const ParseParams& p, const int64_t r, const int lane, uint8_t* stage) {
  const int64_t off = p.offsets[r];
  const int n = (int)(p.offsets[r + 1] - off);
  const uint8_t* text = p.text + off;

  // ---- regex search over the tag occurrences ----------------------------------------------------------------
  // The response is read once, 16 bytes per lane per step (512 B per warp step) from the 16-byte aligned address
  // below it; bytes outside [0, n) are masked.  '<' bytes are rare: their positions come out of a SIMD byte compare,
  // and only those are classified and fed, in text order, to the state machine that every lane advances identically.
  // The first 32 special-string occurrences are remembered, one per lane, for the has-special test further down.
  parse::Matcher m;
  parse::matcher_init(m, p.enable_think != 0);
  const int a0 = (int)((uintptr_t)text & 15);
  const uint4* vbase = reinterpret_cast<const uint4*>(text - a0);
  const int nvec = (a0 + n + 15) >> 4;
  const uint8_t* blob_end = p.text + p.offsets[p.n];
  int occ_pos = -1, occ_len = 0, n_occ = 0;
  for (int vb = 0; vb < nvec && m.state != 3; vb += 32) {
    const int vi = vb + lane;
    unsigned mask = 0;
    if (vi < nvec) {
      const uint8_t* vp = reinterpret_cast<const uint8_t*>(vbase + vi);
      uint4 v;
      if (vp + 16 <= blob_end) {
        v = __ldg(vbase + vi);
      } else {  // the last 16 bytes of the whole blob may not be there to read as a vector
        unsigned w[4] = {0, 0, 0, 0};
        for (int k = 0; k < 16 && vp + k < blob_end; ++k) w[k >> 2] |= (unsigned)vp[k] << (8 * (k & 3));
        v = make_uint4(w[0], w[1], w[2], w[3]);
      }
      mask = lt_mask4(v.x) | (lt_mask4(v.y) << 4) | (lt_mask4(v.z) << 8) | (lt_mask4(v.w) << 12);
      const int lo = vi * 16 - a0;  // text offset of this lane's byte 0
      if (lo < 0) mask &= 0xffffu << (-lo);
      if (lo + 16 > n) mask &= (n - lo >= 16) ? 0xffffu : ((1u << (n - lo > 0 ? n - lo : 0)) - 1u);
    }
    unsigned lanes = __ballot_sync(0xffffffffu, mask != 0);
    bool matched = false;
    while (lanes && !matched) {
      const int l = __ffs(lanes) - 1;
      lanes &= lanes - 1;
      unsigned bits = __shfl_sync(0xffffffffu, mask, l);
      while (bits) {
        const int k = __ffs(bits) - 1;
        bits &= bits - 1;
        const int pos = (vb + l) * 16 - a0 + k;
        const int tag = parse::classify_tag(text + pos, text + n);
        if (tag == parse::T_NONE) continue;
        if (lane == n_occ) {
          occ_pos = pos;
          occ_len = parse::tag_len(tag);
        }
        ++n_occ;
        if (parse::matcher_feed(m, text, n, pos, tag)) {
          matched = true;
          break;
        }
      }
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
      // any special string lying wholly inside [s, e)?  All of them were seen by the scan (they start before the match end)
      bool special = __ballot_sync(0xffffffffu, occ_pos >= s && occ_pos + occ_len <= e) != 0;
      if (n_occ > 32 && !special) special = parse::has_special(text, s, e);  // more occurrences than remembered: look again
      if (special) {
        uint8_t* scr = p.scratch + off;
        const int len = e - s;
        // First the common case, in ONE pass: special strings start with '<' and hold no second one, so occurrences never overlap,
        // and removing all of them at once is what the six sequential replace passes do unless a removal splices a new special
        // string together ("</ans<think>wer>") - which can only happen across a junction, so the 11 bytes before every junction of
        // the cleaned text are looked at.  (One strip at the end equals a strip after every pass: stripping creates nothing.)
        // The occurrences are the ones the scan remembered, one per lane, in text order.
        bool fused = n_occ <= 32;
        int w = 0;
        if (fused) {
          unsigned inside = __ballot_sync(0xffffffffu, occ_pos >= s && occ_pos + occ_len <= e);
          int from = s;
          int my_junction = -1;
          while (inside) {
            const int l = __ffs(inside) - 1;
            inside &= inside - 1;
            const int q = __shfl_sync(0xffffffffu, occ_pos, l), ql = __shfl_sync(0xffffffffu, occ_len, l);
            if (q < from) continue;  // starts inside an occurrence already removed: cannot happen (no '<' inside one)
#pragma unroll 8
            for (int i = from + lane; i < q; i += 32) scr[s + w + (i - from)] = text[i];  // unrolled: eight loads in flight per lane
            w += q - from;
            from = q + ql;
            if (lane == l) my_junction = w;
          }
#pragma unroll 8
          for (int i = from + lane; i < e; i += 32) scr[s + w + (i - from)] = text[i];
          w += e - from;
          __syncwarp();
          bool splice = false;
          if (my_junction >= 0) {
            for (int k = 1; k <= 11 && my_junction - k >= 0; ++k) {
              const uint8_t* c = scr + s + my_junction - k;
              if (*c == '<') {
                const int tag = parse::classify_tag(c, scr + s + w);
                if (tag != parse::T_NONE && parse::tag_len(tag) > k) splice = true;
              }
            }
          }
          fused = !__any_sync(0xffffffffu, splice);
        }
        if (fused) {
          int cs = s, ce = s + w;
          parse::strip(scr, cs, ce);
          s = cs;
          e = ce;
        } else {
        // the span is staged in shared memory for the six passes (each sweeps it once); longer spans are cleaned in
        // the scratch blob itself
        const bool staged = len <= CLEAN_CAP;
        uint8_t* buf = staged ? stage : scr + s;
#pragma unroll 4
        for (int i = lane; i < len; i += 32) buf[i] = text[s + i];
        __syncwarp();
        int cs = 0, ce = len;
        for (int tag = parse::T_THINK; tag <= parse::T_IM_END; ++tag) {  // six passes, each followed by strip (:162-164)
          ce = warp_remove_tag(buf, cs, ce, tag, lane, staged);
          parse::strip(buf, cs, ce);
        }
        if (staged) {
          for (int i = cs + lane; i < ce; i += 32) scr[s + i] = buf[i];
          __syncwarp();
        }
        e = s + ce;
        s = s + cs;
        }
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

// Persistent warps: response lengths differ by 10x and one in ten needs the cleaning passes, so warps take the next
// response from a global counter instead of owning a fixed one - a finished warp never idles behind a slow neighbour of
// its CTA, and the tail of the launch is one response long.
__global__ void __launch_bounds__(PWARPS * 32) parse_kernel(const ParseParams p, unsigned long long* next) {
  __shared__ __align__(16) uint8_t stage[PWARPS][CLEAN_CAP];
  const int lane = threadIdx.x & 31;
  for (;;) {
    unsigned long long r = 0;
    if (lane == 0) r = atomicAdd(next, 1ULL);
    r = __shfl_sync(0xffffffffu, r, 0);
    if (r >= (unsigned long long)p.n) break;
    parse_one(p, (int64_t)r, lane, stage[threadIdx.x >> 5]);
    __syncwarp();
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

// Answer scoring of the text envs (score_logic.cuh): one thread per (prediction, label) pair.  The strings are a few dozen
// bytes, the batch a few thousand pairs: latency-bound, nothing to tile.
__global__ void text_score_kernel(int64_t n, int mode, const uint8_t* __restrict__ pred, const int64_t* __restrict__ pred_off,
                                  const uint8_t* __restrict__ label, const int64_t* __restrict__ label_off, uint8_t* out) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int64_t a = pred_off[i], b = label_off[i];
  out[i] = (uint8_t)score::score_pair(mode, pred + a, (int)(pred_off[i + 1] - a), label + b, (int)(label_off[i + 1] - b));
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
  unsigned long long* next = (unsigned long long*)ragen_scratch(ctx, sizeof(unsigned long long));
  if (!next) return RAGEN_ENOMEM;
  RAGEN_CUDA(cudaMemsetAsync(next, 0, sizeof(unsigned long long), (cudaStream_t)stream));
  const int64_t want = (n + PWARPS - 1) / PWARPS, cap = (int64_t)ctx->sm_count * 12;  // 12 CTAs of 4 warps fit an SM (registers)
  parse_kernel<<<(unsigned)(want < cap ? want : cap), PWARPS * 32, 0, (cudaStream_t)stream>>>(p, next);
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

extern "C" int ragen_text_score(ragen_ctx* ctx, int64_t n, int mode, const uint8_t* pred, const int64_t* pred_offsets,
                                const uint8_t* label, const int64_t* label_offsets, uint8_t* result, void* stream) {
  RAGEN_REQUIRE(ctx, "ctx is NULL");
  RAGEN_REQUIRE(n >= 0 && mode >= 0 && mode <= 3, "bad argument n=%lld mode=%d", (long long)n, mode);
  RAGEN_REQUIRE(n == 0 || (pred && pred_offsets && label && label_offsets && result), "NULL pointer argument");
  if (n == 0) return RAGEN_OK;
  DeviceGuard guard(ctx->device);
  text_score_kernel<<<(unsigned)((n + 127) / 128), 128, 0, (cudaStream_t)stream>>>(n, mode, pred, pred_offsets, label, label_offsets,
                                                                                  result);
  RAGEN_LAUNCHED(ctx);
  return RAGEN_OK;
}
