// Token-trajectory assembly on the device (sm_100a): ContextManager.get_lm_inputs without host strings.
//
// Replaces, in ragen/llm_agent/ctx_manager.py:228-280, the per-turn path "chat messages -> apply_chat_template -> tokenizer(texts,
// padding=True, padding_side='left') -> attention_mask.cumsum(-1)": every piece of a prompt that can appear (chat-template glue,
// instruction prefix, "Turn k / State", "You have N actions left ...", "Reward: r", the canonical LLM responses) is tokenised ONCE
// into a token pool in HBM; an observation is rendered from the glyph codes the env kernels wrote (sokoban/env.py:53-57,
// frozen_lake/env.py:47-57) straight into token ids through a per-tokenizer glyph table; a prompt is then a list of
// (pool offset, length, destination column) segments and one launch gathers them into the left-padded int64 [B, L] input_ids
// together with attention_mask and position_ids.
//
// HBM-bound gather: 4 B read + 24 B written per token (three int64 outputs, the reference's dtypes); nothing is a contraction.
#include "common.cuh"

namespace {

constexpr int AT = 256;  // threads per row CTA

struct AssembleParams {
  int B, L;
  const int32_t* pool;
  int64_t pool_tokens;
  const int64_t* ent_src;  // [E] first token of the segment in the pool
  const int32_t* ent_dst;  // [E] first column of the segment in its row (left padding included)
  const int32_t* ent_len;  // [E]
  const int32_t* ent_off;  // [B + 1] segments of row b = [ent_off[b], ent_off[b + 1])
  const int32_t* row_len;  // [B] real tokens of the row
  int64_t pad_id;
  int64_t* ids;
  int64_t* attn;
  int64_t* pos;
};

// One CTA per row.  The pad region is filled by all threads; then warp w copies segments w, w + 8, ... of the row: the lanes
// walk a segment with consecutive columns, so every store instruction writes 32 consecutive int64 (256 B) per output tensor.
__global__ void __launch_bounds__(AT) traj_assemble_kernel(const AssembleParams p) {
  const int row = blockIdx.x;
  const int L = p.L;
  const int len = min(max(p.row_len[row], 0), L);
  const int pad = L - len;
  int64_t* ids = p.ids + (size_t)row * L;
  int64_t* attn = p.attn + (size_t)row * L;
  int64_t* pos = p.pos + (size_t)row * L;
  for (int t = threadIdx.x; t < pad; t += AT) {
    ids[t] = p.pad_id;
    attn[t] = 0;
    pos[t] = 0;  // attention_mask.cumsum(-1) (:280): pads stay 0, the first real token is 1
  }
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int e0 = p.ent_off[row], e1 = p.ent_off[row + 1];
  for (int e = e0 + wid; e < e1; e += AT / 32) {
    const int64_t src = p.ent_src[e];
    const int dst = p.ent_dst[e];
    int n = p.ent_len[e];
    if (src < 0 || src + n > p.pool_tokens || dst < pad || n < 0) continue;  // malformed program: leave the columns untouched
    n = min(n, L - dst);
    const int32_t* s = p.pool + src;
    for (int i = lane; i < n; i += 32) {
      const int t = dst + i;
      ids[t] = (int64_t)s[i];
      attn[t] = 1;
      pos[t] = (int64_t)(t - pad + 1);
    }
  }
}

struct GlyphParams {
  int64_t n;
  int G;       // glyph codes per env, H * (W + 1): every row ends with code 7 (row separator), the last one included
  int packed;  // 0: one code per byte; 1: two 4-bit codes per byte, low nibble first (ragen_sokoban_step_packed);
               // 3: three bits per cell, row-major, no separators (ragen_sokoban_step_packed3 / _compact): W says where they go
  int W;
  int64_t pitch;  // bytes between the glyph rows of consecutive envs
  const uint8_t* codes;
  int32_t tok[8][4];  // token ids of glyph code c (table[c][0 .. len[c]))
  int32_t len[8];
  int stride;
  void* out;         // [n][stride] of int32 (token ids) or uint8 (UTF-8 bytes)
  int32_t* out_len;  // [n] or NULL
};

// One warp per env: lane l handles codes l, l + 32, ...; the output offset of a code is the running sum of the token counts
// of the codes before it (warp prefix sum per 32-code chunk), so multi-token glyphs (a UTF-8 glyph under a byte-level
// tokenizer) cost nothing extra.
template <typename OutT>
__global__ void __launch_bounds__(128) glyph_tokens_kernel(const GlyphParams p) {
  __shared__ int32_t s_tok[8][4];
  __shared__ int32_t s_len[8];
  if (threadIdx.x < 32) s_tok[threadIdx.x >> 2][threadIdx.x & 3] = p.tok[threadIdx.x >> 2][threadIdx.x & 3];
  if (threadIdx.x < 8) s_len[threadIdx.x] = p.len[threadIdx.x];
  __syncthreads();
  const int lane = threadIdx.x & 31;
  const int64_t env = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (env >= p.n) return;
  const uint8_t* c = p.codes + env * p.pitch;
  OutT* out = reinterpret_cast<OutT*>(p.out) + env * p.stride;
  int base = 0;
  for (int g0 = 0; g0 < p.G; g0 += 32) {
    const int g = g0 + lane;
    int code = 8, k = 0;
    if (g < p.G) {
      if (p.packed == 3) {
        const int r = g / (p.W + 1), col = g - r * (p.W + 1);
        if (col == p.W) {
          code = 7;
        } else {
          const int bit = 3 * (r * p.W + col);
          unsigned two = (unsigned)c[bit >> 3];
          if ((bit & 7) > 5) two |= (unsigned)c[(bit >> 3) + 1] << 8;  // the three bits straddle a byte (never past the row's end)
          code = (two >> (bit & 7)) & 7u;
        }
      } else {
        code = p.packed ? ((c[g >> 1] >> ((g & 1) << 2)) & 15) : c[g];
      }
      code = code > 7 ? 7 : code;
      k = s_len[code];
    }
    int incl = k;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int up = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += up;
    }
    const int at = base + incl - k;
    for (int j = 0; j < k; ++j)
      if (at + j < p.stride) out[at + j] = (OutT)s_tok[code][j];
    base += __shfl_sync(0xffffffffu, incl, 31);
  }
  if (sizeof(OutT) == 1)  // text rows are fixed-width byte strings: the tail is NUL
    for (int i = base + lane; i < p.stride; i += 32) out[i] = (OutT)0;
  if (lane == 0 && p.out_len) p.out_len[env] = min(base, p.stride);
}

}  // namespace

extern "C" int ragen_traj_assemble(ragen_ctx* ctx, int B, int L, const int32_t* pool, int64_t pool_tokens, const int64_t* ent_src,
                                   const int32_t* ent_dst, const int32_t* ent_len, const int32_t* ent_off, const int32_t* row_len,
                                   int64_t pad_id, int64_t* input_ids, int64_t* attention_mask, int64_t* position_ids,
                                   void* stream) {
  RAGEN_REQUIRE(ctx, "ctx is NULL");
  RAGEN_REQUIRE(B >= 0 && L >= 1, "bad shape B=%d L=%d", B, L);
  RAGEN_REQUIRE(pool && ent_src && ent_dst && ent_len && ent_off && row_len && input_ids && attention_mask && position_ids,
                "NULL pointer argument");
  if (B == 0) return RAGEN_OK;
  DeviceGuard guard(ctx->device);
  AssembleParams p{B, L, pool, pool_tokens, ent_src, ent_dst, ent_len, ent_off, row_len, pad_id, input_ids, attention_mask, position_ids};
  traj_assemble_kernel<<<B, AT, 0, (cudaStream_t)stream>>>(p);
  RAGEN_LAUNCHED(ctx);
  return RAGEN_OK;
}

namespace {
int glyph_entry(bool utf8, ragen_ctx* ctx, int64_t n, int G, int packed, int W, const uint8_t* codes, int64_t pitch, const int32_t* table,
                const int32_t* table_len, int stride, void* out, int32_t* n_out, void* stream) {
  RAGEN_REQUIRE(ctx, "ctx is NULL");
  RAGEN_REQUIRE(n >= 0 && G >= 1 && stride >= 1, "bad shape n=%lld G=%d stride=%d", (long long)n, G, stride);
  RAGEN_REQUIRE(packed == 0 || packed == 1 || packed == 3, "glyph row format %d (0 bytes, 1 nibbles, 3 three-bit rows)", packed);
  RAGEN_REQUIRE(packed != 1 || (G & 1) == 0, "packed glyph rows need an even number of codes, got %d", G);
  RAGEN_REQUIRE(packed != 3 || (W >= 1 && G % (W + 1) == 0), "three-bit rows need the room width: G=%d W=%d", G, W);
  RAGEN_REQUIRE(codes && table && table_len && out, "NULL pointer argument");
  const int64_t row_bytes = packed == 3 ? (((int64_t)3 * (G / (W + 1)) * W + 15) / 16) * 2 : packed == 1 ? G / 2 : G;
  RAGEN_REQUIRE(pitch >= row_bytes, "pitch %lld is smaller than an env's glyph row (%lld B)", (long long)pitch, (long long)row_bytes);
  GlyphParams p{};
  int worst = 0;
  for (int c = 0; c < 8; ++c) {  // table / table_len are HOST arrays (a few dozen ints, part of the launch parameters)
    RAGEN_REQUIRE(table_len[c] >= 0 && table_len[c] <= 4, "glyph %d has %d tokens (0..4 supported)", c, table_len[c]);
    p.len[c] = table_len[c];
    worst = table_len[c] > worst ? table_len[c] : worst;
    for (int j = 0; j < 4; ++j) {
      RAGEN_REQUIRE(!utf8 || (table[4 * c + j] >= 0 && table[4 * c + j] <= 255), "glyph %d byte %d is not a byte", c, j);
      p.tok[c][j] = table[4 * c + j];
    }
  }
  RAGEN_REQUIRE((int64_t)stride >= (int64_t)G * (worst ? 1 : 0), "stride %d cannot hold %d glyphs", stride, G);
  if (n == 0) return RAGEN_OK;
  DeviceGuard guard(ctx->device);
  p.n = n;
  p.G = G;
  p.packed = packed;
  p.W = W;
  p.pitch = pitch;
  p.codes = codes;
  p.stride = stride;
  p.out = out;
  p.out_len = n_out;
  const int warps = 4;
  const unsigned grid = (unsigned)((n + warps - 1) / warps);
  if (utf8)
    glyph_tokens_kernel<uint8_t><<<grid, warps * 32, 0, (cudaStream_t)stream>>>(p);
  else
    glyph_tokens_kernel<int32_t><<<grid, warps * 32, 0, (cudaStream_t)stream>>>(p);
  RAGEN_LAUNCHED(ctx);
  return RAGEN_OK;
}
}  // namespace

extern "C" int ragen_glyph_tokens(ragen_ctx* ctx, int64_t n, int G, int packed, const uint8_t* codes, int64_t pitch, const int32_t* table,
                                  const int32_t* table_len, int stride, int32_t* tokens, int32_t* n_tokens, void* stream) {
  return glyph_entry(false, ctx, n, G, packed, 0, codes, pitch, table, table_len, stride, tokens, n_tokens, stream);
}

extern "C" int ragen_glyph_utf8(ragen_ctx* ctx, int64_t n, int G, int format, int W, const uint8_t* codes, int64_t pitch,
                                const int32_t* table, const int32_t* table_len, int stride, uint8_t* text, int32_t* n_bytes,
                                void* stream) {
  return glyph_entry(true, ctx, n, G, format, W, codes, pitch, table, table_len, stride, text, n_bytes, stream);
}
