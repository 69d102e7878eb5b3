// Advantage pass (sm_100a): GAE, RAGEN's bi-level GAE, GRPO outcome advantage, masked whitening.
//
// Replaces verl/verl/trainer/ppo/core_algos.py:67-109 (compute_gae_advantage_return), :113-165
// (compute_grpo_outcome_advantage), verl/verl/utils/torch_functional.py:122-151 (masked_mean/var/whiten) and
// ragen/trainer/core_algos.py:4-92 (compute_bi_level_gae_advantage_return), as dispatched by compute_advantage
// (ragen/trainer/agent_trainer.py:60-137).
//
// Layout: row-major fp32 [B][T] for rewards / values / advantages / returns, uint8 [B][T] masks.  One warp owns one
// row and walks it backwards in chunks: all 32 lanes load a chunk with coalesced accesses and form the per-token
// deltas, lane 0 runs the first-order recurrence over the chunk in shared memory in the reference's exact fp32
// operation order (so raw advantages and returns are bit-identical to the reference's sequential loop; only the
// whitening moments are reduced in a different order, in fp64), and all lanes write the chunk back coalesced while
// accumulating the masked moments, which are reduced with warp shuffles and one fp64 atomic per warp.
// Rows are independent, so the recurrence latency is hidden across warps; the pass is HBM-bound (17 B per token).
#include <stdlib.h>

#include <type_traits>

#include "common.cuh"
#include "scan_tiles.cuh"

namespace {

constexpr int RT = 256;       // threads of the CTA that owns one row
constexpr int RCHUNK = 2048;  // tokens of the row staged in shared memory per iteration (C5's T = 2047: one pass)

// block-wide sum of three doubles, then ONE set of atomics per CTA
__device__ __forceinline__ void add_moments(double* moments, double s0, double s1, double s2) {
  __shared__ double red[3][RT / 32];
  s0 = warp_sum(s0);
  s1 = warp_sum(s1);
  s2 = warp_sum(s2);
  const int w = threadIdx.x >> 5;
  if ((threadIdx.x & 31) == 0) {
    red[0][w] = s0;
    red[1][w] = s1;
    red[2][w] = s2;
  }
  __syncthreads();
  if (threadIdx.x < 3 && moments) {
    double t = 0.0;
#pragma unroll
    for (int k = 0; k < RT / 32; ++k) t += red[threadIdx.x][k];
    atomicAdd(moments + threadIdx.x, t);
  }
}

__device__ __forceinline__ double block_sum(double v) {  // every thread gets the sum
  __shared__ double red[RT / 32];
  __syncthreads();
  v = warp_sum(v);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  double t = 0.0;
#pragma unroll
  for (int k = 0; k < RT / 32; ++k) t += red[k];
  return t;
}

// The serial part of every estimator here is x <- d[i] + c * x walked right to left by ONE thread over a row staged in
// shared memory.  Written naively, each step waits for its own shared-memory loads (~30 cycles) before the 8-cycle
// multiply-add; these helpers load a batch of 16 operands into registers first, so the loads of a batch overlap and only
// the dependent multiply-adds remain on the critical path.  Same operations, same order: results are unchanged.
constexpr int BATCH = 16;

// d[i] <- x = d[i] + c * x for i = hi-1 .. lo (c constant)
__device__ __forceinline__ float chain_const(float* d, int lo, int hi, float c, float x) {
  int i = hi - 1;
  for (; i - (BATCH - 1) >= lo; i -= BATCH) {
    float v[BATCH];
#pragma unroll
    for (int k = 0; k < BATCH; ++k) v[k] = d[i - k];
#pragma unroll
    for (int k = 0; k < BATCH; ++k) {
      x = __fadd_rn(v[k], __fmul_rn(c, x));
      v[k] = x;
    }
#pragma unroll
    for (int k = 0; k < BATCH; ++k) d[i - k] = v[k];
  }
  for (; i >= lo; --i) {
    x = __fadd_rn(d[i], __fmul_rn(c, x));
    d[i] = x;
  }
  return x;
}

// d[i] <- x = d[i] + c[i] * x for i = hi-1 .. lo (per-token coefficient)
__device__ __forceinline__ float chain_var(float* d, const float* c, int lo, int hi, float x) {
  int i = hi - 1;
  for (; i - (BATCH - 1) >= lo; i -= BATCH) {
    float v[BATCH], w[BATCH];
#pragma unroll
    for (int k = 0; k < BATCH; ++k) {
      v[k] = d[i - k];
      w[k] = c[i - k];
    }
#pragma unroll
    for (int k = 0; k < BATCH; ++k) {
      x = __fadd_rn(v[k], __fmul_rn(w[k], x));
      v[k] = x;
    }
#pragma unroll
    for (int k = 0; k < BATCH; ++k) d[i - k] = v[k];
  }
  for (; i >= lo; --i) {
    x = __fadd_rn(d[i], __fmul_rn(c[i], x));
    d[i] = x;
  }
  return x;
}

// A_t = (r_t + gamma V_{t+1} - V_t) + (gamma lam) A_{t+1}, V_T = 0, over ALL columns (core_algos.py:98-105);
// returns = A + V (:107).  gamma and gamma*lam arrive already rounded to fp32 the way torch rounds a Python scalar
// that multiplies an fp32 tensor.  One CTA per row: 256 threads stage the row and form the deltas (the loads are what
// bounds this pass at C5, so they are spread over many threads), thread 0 runs the first-order recurrence in the
// reference's exact fp32 order, 256 threads write back and reduce the masked moments.
__global__ void __launch_bounds__(RT) gae_kernel(int B, int T, const float* __restrict__ rewards, const float* __restrict__ values,
                                                 const uint8_t* __restrict__ mask, float gamma, float gl, float* __restrict__ adv,
                                                 float* __restrict__ ret, double* moments) {
  __shared__ float s_a[RCHUNK];
  __shared__ float s_v[RCHUNK];
  __shared__ uint8_t s_m[RCHUNK];
  const int tid = threadIdx.x;
  const size_t base = (size_t)blockIdx.x * T;
  double m0 = 0.0, m1 = 0.0, m2 = 0.0;
  float carry = 0.0f;  // thread 0
  for (int t1 = T; t1 > 0; t1 -= RCHUNK) {
    const int t0 = max(0, t1 - RCHUNK), len = t1 - t0;
#pragma unroll 4
    for (int i = tid; i < len; i += RT) {
      const int t = t0 + i;
      const float v = __ldg(values + base + t);
      const float vn = (t + 1 < T) ? __ldg(values + base + t + 1) : 0.0f;
      const float r = __ldg(rewards + base + t);
      s_m[i] = mask[base + t];
      s_a[i] = __fsub_rn(__fadd_rn(r, __fmul_rn(gamma, vn)), v);
      s_v[i] = v;
    }
    __syncthreads();
    if (tid == 0) carry = chain_const(s_a, 0, len, gl, carry);
    __syncthreads();
#pragma unroll 4
    for (int i = tid; i < len; i += RT) {
      const float a = s_a[i];
      adv[base + t0 + i] = a;
      ret[base + t0 + i] = __fadd_rn(a, s_v[i]);
      if (s_m[i]) {
        m0 += 1.0;
        m1 += (double)a;
        m2 += (double)a * (double)a;
      }
    }
    __syncthreads();
  }
  add_moments(moments, m0, m1, m2);
}

// ragen/trainer/core_algos.py:44-88 for one row, as ONE backward sweep: the turn-level estimate at a reward
// position depends only on later reward positions, and the token-level pass at that position only needs the
// turn-level result of the same position, so both passes advance together.  fp32 operation order as the reference.
//
// Per chunk (right to left) the work is split so that the serial part is only the recurrence itself:
//   A (256 threads) "value of the next loss-mask token to the right": per 32-token group by ballot + shuffle, across
//                   groups through a per-group table in shared memory; it turns every non-reward loss-mask token into
//                   its delta = (0 + gamma * next_value) - value;
//   B (thread 0)    lo = d + c * lo over the tokens BETWEEN reward positions as a branch-free, unrolled loop: c = gamma*lam and
//                   d = delta on a loss-mask token, c = 1 and d = 0 (an exact identity in fp32) elsewhere, so the loop's
//                   shared-memory loads run ahead of the dependent multiply-add; reward positions (a handful per row, listed
//                   by phase A) run the turn-level step, reset the token-level state and restart it from the updated
//                   reward (:73-88);
//   C (256 threads) returns = advantage + value, zeros off the loss mask, moments, coalesced stores.
__global__ void __launch_bounds__(RT) bilevel_gae_kernel(int B, int T, const float* __restrict__ rewards,
                                                         const float* __restrict__ values, const uint8_t* __restrict__ mask,
                                                         float gamma, float gl, float hgamma, float hgl, float* __restrict__ adv,
                                                         float* __restrict__ ret, double* moments) {
  constexpr int NG = RCHUNK / 32;
  __shared__ float s_r[RCHUNK];    // rewards in -> deltas -> advantages out
  __shared__ float s_v[RCHUNK];    // values in (returns out at reward positions)
  __shared__ uint8_t s_m[RCHUNK];  // bit0 loss mask, bit1 reward position
  __shared__ float s_c[RCHUNK];    // recurrence coefficient per token
  __shared__ float g_first[NG];    // value of the leftmost loss-mask token of each 32-token group
  __shared__ uint8_t g_has[NG];
  __shared__ unsigned g_eos[NG];   // reward positions of each group, as a bit mask
  __shared__ float s_carry;        // value of the nearest loss-mask token right of the current chunk
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const size_t base = (size_t)blockIdx.x * T;
  double m0 = 0.0, m1 = 0.0, m2 = 0.0;
  bool hi_has_next = false;                            // thread 0: turn-level state
  float hi_next_v = 0.0f, hi_lg = 0.0f, lo_lg = 0.0f;  // thread 0
  if (tid == 0) s_carry = 0.0f;
  for (int t1 = T; t1 > 0; t1 -= RCHUNK) {
    const int t0 = max(0, t1 - RCHUNK), len = t1 - t0;
    const int ngroups = (len + 31) >> 5;
    // ---- A1: stage, per-group leftmost loss-mask value -----------------------------------------------------------
    for (int g = wid; g < ngroups; g += RT / 32) {
      const int i = (g << 5) + lane;
      const bool in = i < len;
      const float r = in ? __ldg(rewards + base + t0 + i) : 0.0f;
      const float v = in ? __ldg(values + base + t0 + i) : 0.0f;
      const bool m = in && mask[base + t0 + i] != 0;
      const bool eos = in && r != 0.0f;  // reward_mask = token_level_rewards.bool()
      const unsigned bits = __ballot_sync(0xffffffffu, m);
      const float first_v = __shfl_sync(0xffffffffu, v, bits ? (__ffs(bits) - 1) : 0);
      const unsigned ebits = __ballot_sync(0xffffffffu, eos);
      if (lane == 0) {
        g_first[g] = first_v;
        g_has[g] = bits ? 1 : 0;
        g_eos[g] = ebits;
      }
      if (in) {
        s_r[i] = r;
        s_v[i] = v;
        s_m[i] = (uint8_t)((m ? 1 : 0) | (eos ? 2 : 0));
      }
    }
    __syncthreads();
    // ---- A2: deltas of the non-reward loss-mask tokens --------------------------------------------------------------
    for (int g = wid; g < ngroups; g += RT / 32) {
      const int i = (g << 5) + lane;
      const bool in = i < len;
      const int f = in ? s_m[i] : 0;
      const float v = in ? s_v[i] : 0.0f;
      const unsigned bits = __ballot_sync(0xffffffffu, (f & 1) != 0);
      const unsigned higher = (lane == 31) ? 0u : (bits & (0xfffffffeu << lane));
      const float from_group = __shfl_sync(0xffffffffu, v, higher ? (__ffs(higher) - 1) : 0);
      float nv = from_group;
      if (!higher) {
        // the reference indexes valid_positions[i + 1] here and raises IndexError when the last loss-mask position
        // carries no reward; that input is treated as "next value 0" (s_carry starts at 0)
        nv = s_carry;
        for (int h = g + 1; h < ngroups; ++h)
          if (g_has[h]) {
            nv = g_first[h];
            break;
          }
      }
      if (in) {
        if (f == 1) {
          s_r[i] = __fsub_rn(__fadd_rn(s_r[i], __fmul_rn(gamma, nv)), v);
          s_c[i] = gl;
        } else if (f == 0) {
          s_r[i] = 0.0f;
          s_c[i] = 1.0f;
        }
      }
    }
    __syncthreads();
    // ---- B -------------------------------------------------------------------------------------------------------------
    if (tid == 0) {
      for (int h = 0; h < ngroups; ++h)
        if (g_has[h]) {
          s_carry = g_first[h];  // for the chunk to the left
          break;
        }
      auto run_segment = [&](int lo_end, int hi_end) {  // tokens hi_end-1 .. lo_end, none of them a reward position
        lo_lg = chain_var(s_r, s_c, lo_end, hi_end, lo_lg);
      };
      int seg_end = len;
      for (int g = ngroups - 1; g >= 0; --g) {
        unsigned eb = g_eos[g];
        while (eb) {  // reward positions of this group, right to left
          const int bit = 31 - __clz(eb);
          eb &= ~(1u << bit);
          const int pos = (g << 5) + bit;
          run_segment(pos + 1, seg_end);
          const int f = s_m[pos];
          const float r = s_r[pos], v = s_v[pos];
          const float nvh = hi_has_next ? hi_next_v : 0.0f;
          hi_lg = __fadd_rn(__fsub_rn(__fadd_rn(r, __fmul_rn(hgamma, nvh)), v), __fmul_rn(hgl, hi_lg));
          hi_next_v = v;
          hi_has_next = true;
          const float upd = __fadd_rn(hi_lg, v);  // returns[pos] = updated_reward[pos] = advantages[pos] + values[pos]
          float a_out = hi_lg, r_out = upd;
          if (f & 1) {  // token-level pass restarts at a reward position: next value 0, lastgaelam 0
            const float delta = __fsub_rn(__fadd_rn(upd, __fmul_rn(gamma, 0.0f)), v);
            lo_lg = __fadd_rn(delta, __fmul_rn(gl, 0.0f));
            a_out = lo_lg;
            r_out = __fadd_rn(lo_lg, v);
          }
          s_r[pos] = a_out;
          s_v[pos] = r_out;
          seg_end = pos;
        }
      }
      run_segment(0, seg_end);
    }
    __syncthreads();
    // ---- C -------------------------------------------------------------------------------------------------------------
#pragma unroll 4
    for (int i = tid; i < len; i += RT) {
      const int f = s_m[i];
      float a = 0.0f, rt = 0.0f;
      if (f == 1) {
        a = s_r[i];
        rt = __fadd_rn(a, s_v[i]);
      } else if (f & 2) {
        a = s_r[i];
        rt = s_v[i];
      }
      adv[base + t0 + i] = a;
      ret[base + t0 + i] = rt;
      if (f & 1) {
        m0 += 1.0;
        m1 += (double)a;
        m2 += (double)a * (double)a;
      }
    }
    __syncthreads();
  }
  add_moments(moments, m0, m1, m2);
}

// masked_whiten (torch_functional.py:139-151) from the three masked moments {sum m, sum x m, sum x^2 m}:
// mean = S1/(S0+1e-8); var = (S2 - 2 mean S1 + mean^2 S0)/(S0+1e-8) * S0/(S0-1); x <- (x-mean) * rsqrt(var+1e-8).
// Applied to every element (the reference does not re-mask).
__global__ void whiten_kernel(int64_t n, float* __restrict__ x, const double* __restrict__ moments) {
  const double S0 = moments[0], S1 = moments[1], S2 = moments[2];
  const double mean = S1 / (S0 + 1e-8);
  double var = (S2 - 2.0 * mean * S1 + mean * mean * S0) / (S0 + 1e-8);
  var = var * (S0 / (S0 - 1.0));
  const float meanf = (float)mean;
  // fewer than two mask tokens: masked_var raises in the reference; here every output is NaN so that the misuse shows
  const float scale = (S0 > 1.0) ? (float)(1.0 / sqrt(var + 1e-8)) : __int_as_float(0x7fc00000);
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if ((reinterpret_cast<uintptr_t>(x) & 15) == 0) {  // 128-bit accesses, two in flight per thread
    float4* x4 = reinterpret_cast<float4*>(x);
    const int64_t n4 = n >> 2;
    int64_t i = tid;
    for (; i + stride < n4; i += 2 * stride) {
      float4 a = x4[i], b = x4[i + stride];
      a.x = (a.x - meanf) * scale; a.y = (a.y - meanf) * scale; a.z = (a.z - meanf) * scale; a.w = (a.w - meanf) * scale;
      b.x = (b.x - meanf) * scale; b.y = (b.y - meanf) * scale; b.z = (b.z - meanf) * scale; b.w = (b.w - meanf) * scale;
      x4[i] = a;
      x4[i + stride] = b;
    }
    if (i < n4) {
      float4 a = x4[i];
      a.x = (a.x - meanf) * scale; a.y = (a.y - meanf) * scale; a.z = (a.z - meanf) * scale; a.w = (a.w - meanf) * scale;
      x4[i] = a;
    }
    for (int64_t j = (n4 << 2) + tid; j < n; j += stride) x[j] = (x[j] - meanf) * scale;
    return;
  }
  for (int64_t i = tid; i < n; i += stride) x[i] = (x[i] - meanf) * scale;
}

// masked moments of an arbitrary tensor (for whitening inputs produced elsewhere)
__global__ void moments_kernel(int64_t n, const float* __restrict__ x, const uint8_t* __restrict__ mask, double* moments) {
  double m0 = 0.0, m1 = 0.0, m2 = 0.0;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride)
    if (mask[i]) {
      const double a = x[i];
      m0 += 1.0;
      m1 += a;
      m2 += a * a;
    }
  add_moments(moments, m0, m1, m2);
}

// Outcome estimators, step 1: scores[b] = sum_t rewards[b][t] (fp64 accumulation, rounded once); per-group
// {count, sum, sum of squares} (core_algos.py:139-149).  One CTA per row.
// TPR threads per row: RT (one CTA per row - few rows, all SMs busy) or 32 (one warp per row, eight rows per CTA - thousands of
// rows: a CTA per 8 KB row spends as long being scheduled and reducing across warps as streaming its row)
template <int TPR>
__global__ void __launch_bounds__(RT) grpo_scores_kernel(int B, int T, const float* __restrict__ rewards,
                                                         const int32_t* __restrict__ group, int n_groups,
                                                         float* __restrict__ scores, double* gstats) {
  const int row = blockIdx.x * (RT / TPR) + threadIdx.x / TPR;
  const int tr = threadIdx.x % TPR;
  if (TPR != RT && row >= B) return;  // whole warps (no CTA barrier on this path)
  const size_t base = (size_t)row * T;
  double s = 0.0;
  // a row of an odd T starts anywhere: up to three tokens one by one until the address is 16-byte aligned, then 128-bit loads
  const float* p = rewards + base;
  const bool vec = (reinterpret_cast<uintptr_t>(rewards) & 15) == 0;
  const int head = vec ? min((int)((4 - (base & 3)) & 3), T) : T;
  for (int t = tr; t < head; t += TPR) s += (double)__ldg(p + t);
  const int nv = (T - head) >> 2;
  const float4* p4 = reinterpret_cast<const float4*>(p + head);
#pragma unroll 4
  for (int i = tr; i < nv; i += TPR) {
    const float4 x = __ldg(p4 + i);
    s += (double)x.x;
    s += (double)x.y;
    s += (double)x.z;
    s += (double)x.w;
  }
  for (int t = head + 4 * nv + tr; t < T; t += TPR) s += (double)__ldg(p + t);
  s = TPR == RT ? block_sum(s) : warp_sum(s);
  if (tr == 0) {
    const float sf = (float)s;
    scores[row] = sf;
    const int g = group ? group[row] : -1;
    if (g >= 0 && g < n_groups) {
      atomicAdd(gstats + 3 * g + 0, 1.0);
      atomicAdd(gstats + 3 * g + 1, (double)sf);
      atomicAdd(gstats + 3 * g + 2, (double)sf * (double)sf);
    }
  }
}

// Step 2.  mode 0: score - mean (GRPO without std); 1: GRPO, (score - mean) / (std + eps) with the unbiased std
// (core_algos.py:150-165); 2: RLOO, score*n/(n-1) - mean*n/(n-1) for groups of n > 1 (:205-242); 3: REINFORCE++-baseline,
// score - mean, to be whitened by the caller from `moments` (:168-202).  Groups of one use mean 0 (and std 1).  The value is
// broadcast over the row and multiplied by the mask.
template <int TPR>
__global__ void __launch_bounds__(RT) outcome_apply_kernel(int B, int T, const float* __restrict__ scores,
                                                           const int32_t* __restrict__ group, int n_groups,
                                                           const double* __restrict__ gstats, const uint8_t* __restrict__ mask,
                                                           float eps, int mode, float* __restrict__ adv, double* moments) {
  const int row_raw = blockIdx.x * (RT / TPR) + threadIdx.x / TPR;
  const int tr = threadIdx.x % TPR;
  const bool live = row_raw < B;  // a warp past the last row writes nothing but stays for the CTA-wide moments
  const int row = live ? row_raw : B - 1;
  const int g = group[row];
  float mean = 0.0f, sd = 1.0f, cnt = 1.0f;
  if (g >= 0 && g < n_groups) {
    const double c = gstats[3 * g], s1 = gstats[3 * g + 1], s2 = gstats[3 * g + 2];
    cnt = (float)c;
    if (c > 1.5) {
      const double mu = s1 / c;
      mean = (float)mu;
      sd = (float)sqrt(fmax((s2 - c * mu * mu) / (c - 1.0), 0.0));
    }
  }
  const float sc = scores[row];
  float val;
  if (mode == 1) {
    val = __fdiv_rn(__fsub_rn(sc, mean), __fadd_rn(sd, eps));
  } else if (mode == 2) {
    val = sc;
    if (cnt > 1.5f) val = __fsub_rn(__fdiv_rn(__fmul_rn(sc, cnt), cnt - 1.0f), __fdiv_rn(__fmul_rn(mean, cnt), cnt - 1.0f));
  } else {
    val = __fsub_rn(sc, mean);
  }
  const size_t base = (size_t)row * T;
  const float off = __fmul_rn(val, 0.0f);
  int on = 0;
  // token index and flat index agree for the mask bytes and the output words: the same up-to-three head tokens bring both to
  // their alignment (4 bytes of mask, 16 bytes of output), then one mask word and one 128-bit store per four tokens
  const bool vec = (reinterpret_cast<uintptr_t>(adv) & 15) == 0 && (reinterpret_cast<uintptr_t>(mask) & 3) == 0;
  const int Tl = live ? T : 0;
  const int head = vec ? min((int)((4 - (base & 3)) & 3), Tl) : Tl;
  for (int t = tr; t < head; t += TPR) {
    const bool m = mask[base + t] != 0;
    adv[base + t] = m ? val : off;
    on += m;
  }
  const int nv = (Tl - head) >> 2;
  const uint32_t* m4 = reinterpret_cast<const uint32_t*>(mask + base + head);
  float4* a4 = reinterpret_cast<float4*>(adv + base + head);
#pragma unroll 4
  for (int i = tr; i < nv; i += TPR) {
    const uint32_t mw = __ldg(m4 + i);
    const bool b0 = (mw & 0xffu) != 0, b1 = (mw & 0xff00u) != 0, b2 = (mw & 0xff0000u) != 0, b3 = (mw & 0xff000000u) != 0;
    a4[i] = make_float4(b0 ? val : off, b1 ? val : off, b2 ? val : off, b3 ? val : off);
    on += (int)b0 + (int)b1 + (int)b2 + (int)b3;
  }
  for (int t = head + 4 * nv + tr; t < Tl; t += TPR) {
    const bool m = mask[base + t] != 0;
    adv[base + t] = m ? val : off;
    on += m;
  }
  if (moments) add_moments(moments, (double)on, (double)on * (double)val, (double)on * (double)val * (double)val);
}

// from this many rows on a warp per row (eight rows per CTA) beats a CTA per row
constexpr int WARP_ROWS_MIN = 1024;
static void launch_row_sums(cudaStream_t st, int B, int T, const float* rewards, const int32_t* group, int n_groups, float* scores,
                            double* gstats) {
  if (B >= WARP_ROWS_MIN)
    grpo_scores_kernel<32><<<(B + RT / 32 - 1) / (RT / 32), RT, 0, st>>>(B, T, rewards, group, n_groups, scores, gstats);
  else
    grpo_scores_kernel<RT><<<B, RT, 0, st>>>(B, T, rewards, group, n_groups, scores, gstats);
}
static void launch_outcome_apply(cudaStream_t st, int B, int T, const float* scores, const int32_t* group, int n_groups,
                                 const double* gstats, const uint8_t* mask, float eps, int mode, float* adv, double* moments) {
  if (B >= WARP_ROWS_MIN)
    outcome_apply_kernel<32><<<(B + RT / 32 - 1) / (RT / 32), RT, 0, st>>>(B, T, scores, group, n_groups, gstats, mask, eps, mode, adv, moments);
  else
    outcome_apply_kernel<RT><<<B, RT, 0, st>>>(B, T, scores, group, n_groups, gstats, mask, eps, mode, adv, moments);
}

// REINFORCE++ (core_algos.py:245-278): running = r_t + gamma * running, returns_t = running, running *= mask_t, right to
// left, in the reference's fp32 order; plus the masked moments of the returns for the whitening that follows.
// REMAX = 1 (core_algos.py:281-305): returns = reverse cumulative sum of rewards*mask (torch's CPU cumsum accumulates
// fp32 data in double and rounds every prefix), adv = returns - baseline * mask.  One CTA per row, as gae_kernel.
template <int REMAX>
__global__ void __launch_bounds__(RT) returns_kernel(int B, int T, const float* __restrict__ rewards,
                                                     const uint8_t* __restrict__ mask, float gamma,
                                                     const float* __restrict__ baselines, float* __restrict__ adv,
                                                     float* __restrict__ ret, double* moments) {
  __shared__ float s_r[RCHUNK];
  __shared__ uint8_t s_m[RCHUNK];
  const int tid = threadIdx.x;
  const int row = blockIdx.x;
  const size_t base = (size_t)row * T;
  const float bl = REMAX ? baselines[row] : 0.0f;
  double m0 = 0.0, m1 = 0.0, m2 = 0.0;
  float run = 0.0f;   // thread 0
  double acc = 0.0;   // thread 0
  for (int t1 = T; t1 > 0; t1 -= RCHUNK) {
    const int t0 = max(0, t1 - RCHUNK), len = t1 - t0;
#pragma unroll 4
    for (int i = tid; i < len; i += RT) {
      const float r = __ldg(rewards + base + t0 + i);
      const uint8_t m = mask[base + t0 + i];
      s_m[i] = m;
      s_r[i] = (REMAX && !m) ? __fmul_rn(r, 0.0f) : r;  // rewards * mask
    }
    __syncthreads();
    if (tid == 0) {
      int i = len - 1;
      for (; i - (BATCH - 1) >= 0; i -= BATCH) {  // operands of a batch are loaded before the dependent chain starts
        float v[BATCH];
        uint8_t mk[BATCH];
#pragma unroll
        for (int k = 0; k < BATCH; ++k) {
          v[k] = s_r[i - k];
          mk[k] = s_m[i - k];
        }
#pragma unroll
        for (int k = 0; k < BATCH; ++k) {
          if (REMAX) {
            acc += (double)v[k];
            v[k] = (float)acc;
          } else {
            run = __fadd_rn(v[k], __fmul_rn(gamma, run));
            v[k] = run;
            run = mk[k] ? run : __fmul_rn(run, 0.0f);
          }
        }
#pragma unroll
        for (int k = 0; k < BATCH; ++k) s_r[i - k] = v[k];
      }
      for (; i >= 0; --i) {
        if (REMAX) {
          acc += (double)s_r[i];
          s_r[i] = (float)acc;
        } else {
          run = __fadd_rn(s_r[i], __fmul_rn(gamma, run));
          s_r[i] = run;
          run = s_m[i] ? run : __fmul_rn(run, 0.0f);
        }
      }
    }
    __syncthreads();
#pragma unroll 4
    for (int i = tid; i < len; i += RT) {
      const float x = s_r[i];
      ret[base + t0 + i] = x;
      if (REMAX) {
        adv[base + t0 + i] = __fsub_rn(x, s_m[i] ? bl : __fmul_rn(bl, 0.0f));
      } else if (s_m[i]) {
        m0 += 1.0;
        m1 += (double)x;
        m2 += (double)x * (double)x;
      }
    }
    __syncthreads();
  }
  if (!REMAX) add_moments(moments, m0, m1, m2);
}

// out = masked_whiten(x) * mask (the last two lines of compute_reinforce_plus_plus_outcome_advantage, :275-276)
__global__ void whiten_masked_kernel(int64_t n, const float* __restrict__ x, const uint8_t* __restrict__ mask,
                                     const double* __restrict__ moments, float* __restrict__ out) {
  const double S0 = moments[0], S1 = moments[1], S2 = moments[2];
  const double mean = S1 / (S0 + 1e-8);
  double var = (S2 - 2.0 * mean * S1 + mean * mean * S0) / (S0 + 1e-8);
  var = var * (S0 / (S0 - 1.0));
  const float meanf = (float)mean;
  // fewer than two mask tokens: masked_var raises in the reference; here every output is NaN so that the misuse shows
  const float scale = (S0 > 1.0) ? (float)(1.0 / sqrt(var + 1e-8)) : __int_as_float(0x7fc00000);
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  auto one = [&](float xv, uint8_t m) {
    const float v = (xv - meanf) * scale;
    return m ? v : __fmul_rn(v, 0.0f);
  };
  int64_t done = 0;
  if (((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(out)) & 15) == 0 && (reinterpret_cast<uintptr_t>(mask) & 3) == 0) {
    const float4* x4 = reinterpret_cast<const float4*>(x);
    const uchar4* m4 = reinterpret_cast<const uchar4*>(mask);
    float4* o4 = reinterpret_cast<float4*>(out);
    const int64_t n4 = n >> 2;
#pragma unroll 2
    for (int64_t i = tid; i < n4; i += stride) {
      const float4 a = x4[i];
      const uchar4 m = m4[i];
      o4[i] = make_float4(one(a.x, m.x), one(a.y, m.y), one(a.z, m.z), one(a.w, m.w));
    }
    done = n4 << 2;
  }
  for (int64_t i = done + tid; i < n; i += stride) out[i] = one(x[i], mask[i]);
}

// apply_kl_penalty (verl/verl/trainer/ppo/ray_trainer.py:145-175) with kl_penalty (core_algos.py:469-501):
// kld = f(logp, ref) * mask; rewards = scores - beta * kld; row_kl[b] = sum(kld) / (sum(mask) + 1e-8).
// mode 0 "kl", 1 "abs", 2 "mse", 3 "low_var_kl".  One CTA per row.
__global__ void __launch_bounds__(RT) kl_rewards_kernel(int B, int T, const float* __restrict__ scores,
                                                        const float* __restrict__ logp, const float* __restrict__ ref,
                                                        const uint8_t* __restrict__ mask, int mode, float beta,
                                                        float* __restrict__ out, float* __restrict__ row_kl) {
  const int row = blockIdx.x;
  const size_t base = (size_t)row * T;
  double ksum = 0.0, msum = 0.0;
  // one token: the penalty term (masked), the token-level reward, the two row sums
  auto token = [&](float a, float b, float sc, bool m) -> float {
    float k;
    if (mode == 0) {
      k = __fsub_rn(a, b);
    } else if (mode == 1) {
      k = fabsf(__fsub_rn(a, b));
    } else if (mode == 2) {
      const float d = __fsub_rn(a, b);
      k = __fmul_rn(0.5f, __fmul_rn(d, d));
    } else {
      const float d = __fsub_rn(b, a);
      k = __fsub_rn(__fsub_rn(expf(d), d), 1.0f);
      k = fminf(fmaxf(k, -10.0f), 10.0f);
    }
    k = m ? k : __fmul_rn(k, 0.0f);
    ksum += (double)k;
    msum += m ? 1.0 : 0.0;
    return __fsub_rn(sc, __fmul_rn(beta, k));
  };
  auto scalar = [&](int t) { out[base + t] = token(__ldg(logp + base + t), __ldg(ref + base + t), __ldg(scores + base + t), mask[base + t] != 0); };
  // up to three head tokens bring all five arrays to their alignment at once (token index = flat index), then 128-bit accesses
  const bool vec = ((reinterpret_cast<uintptr_t>(scores) | reinterpret_cast<uintptr_t>(logp) | reinterpret_cast<uintptr_t>(ref) |
                     reinterpret_cast<uintptr_t>(out)) & 15) == 0 && (reinterpret_cast<uintptr_t>(mask) & 3) == 0;
  const int head = vec ? min((int)((4 - (base & 3)) & 3), T) : T;
  for (int t = threadIdx.x; t < head; t += RT) scalar(t);
  const int nv = (T - head) >> 2;
  const size_t vb = base + head;
  const float4* a4 = reinterpret_cast<const float4*>(logp + vb);
  const float4* b4 = reinterpret_cast<const float4*>(ref + vb);
  const float4* s4 = reinterpret_cast<const float4*>(scores + vb);
  const uint32_t* m4 = reinterpret_cast<const uint32_t*>(mask + vb);
  float4* o4 = reinterpret_cast<float4*>(out + vb);
#pragma unroll 2
  for (int i = threadIdx.x; i < nv; i += RT) {
    const float4 a = __ldg(a4 + i), b = __ldg(b4 + i), sc = __ldg(s4 + i);
    const uint32_t mw = __ldg(m4 + i);
    float4 o;
    o.x = token(a.x, b.x, sc.x, (mw & 0xffu) != 0);
    o.y = token(a.y, b.y, sc.y, (mw & 0xff00u) != 0);
    o.z = token(a.z, b.z, sc.z, (mw & 0xff0000u) != 0);
    o.w = token(a.w, b.w, sc.w, (mw & 0xff000000u) != 0);
    o4[i] = o;
  }
  for (int t = head + 4 * nv + threadIdx.x; t < T; t += RT) scalar(t);
  ksum = block_sum(ksum);
  msum = block_sum(msum);
  if (threadIdx.x == 0 && row_kl) row_kl[row] = (float)ksum / ((float)msum + 1e-8f);
}

// _filter_rollout's statistics (ragen/trainer/agent_trainer.py:466-469): row sums of the score tensor (grpo_scores_kernel
// with no groups), then per group of `group_size` consecutive rows the unbiased std, the max and the mean.
__global__ void group_stats_kernel(int n_groups, int group_size, const float* __restrict__ row_sum, float* __restrict__ gstd,
                                   float* __restrict__ gmax, float* __restrict__ gmean) {
  const int g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= n_groups) return;
  double s1 = 0.0, s2 = 0.0;
  float mx = -INFINITY;
  for (int r = 0; r < group_size; ++r) {
    const float sf = row_sum[(size_t)g * group_size + r];
    s1 += (double)sf;
    s2 += (double)sf * (double)sf;
    mx = fmaxf(mx, sf);
  }
  const double n = (double)group_size, mu = s1 / n;
  gmean[g] = (float)mu;
  gmax[g] = mx;
  gstd[g] = group_size > 1 ? (float)sqrt(fmax((s2 - n * mu * mu) / (n - 1.0), 0.0)) : NAN;  // torch.std of one value is NaN
}


// ---- many rows (B >= 4096): a CTA of two warps per group of 8 rows, ONE LANE PER ROW in the recurrence ---------------------
// The one-CTA-per-row kernels above put a row's serial recurrence on one lane of one warp: ~5 warp-instructions per token with
// 1/32 of the lanes working, which at B = 8192 rows is ~80 M warp-instructions - more issue slots than the HBM time of the
// pass (192 us measured).  Here a row group's "scan" warp runs the recurrences of its 8 rows side by side, one lane per row,
// in the same fp32 (or double, ReMax) order.  Rows are T floats apart in HBM, so the group is read as 32-token tiles, one
// coalesced 128-byte row segment per request, staged by 4-byte cp.async into a 4-deep ring of shared-memory tiles by the
// group's "mover" warp, which also writes the finished tile back (coalesced) and accumulates the masked moments; one CTA
// barrier per tile hands tiles between the two.  Tile rows are 33 floats apart so the lane-per-row column walk is bank-conflict
// free.  Mask rows start at arbitrary byte addresses (T is odd at C5): the 32 bytes of a tile row are fetched as the 9 aligned
// words that cover them.  Tiles run right to left, the carries (recurrence state, V_{t+1}) stay in the lane's registers.
// Results equal the per-row kernels bit for bit (tests/test_gpu_advantage_ctx.py).  Measured at 8192 x 2047: GAE 192 -> 102 us,
// REINFORCE++ 324 -> 89 us; the pass is still bound by instruction latency per tile step (~2.4 instructions per token), not HBM.
constexpr int TS_STAGES = 4;            // ring slots: one being scanned, one being stored, one just freed, TS_LEAD in flight
constexpr int TS_LEAD = TS_STAGES - 3;  // tiles the loader runs ahead of the scan
constexpr int TS_VS = 33;  // floats between tile rows
constexpr int TS_MS = 40;  // bytes between mask tile rows (32 + 3 bytes of misalignment, rounded up to words)
template <int RPW>  // rows per warp
struct TsStage {
  float v[RPW * TS_VS];
  float r[RPW * TS_VS];
  uint8_t m[RPW * TS_MS];
};
enum { TS_GAE = 0, TS_RFPP = 1, TS_REMAX = 2, TS_BILEVEL = 3 };

__device__ __forceinline__ void cp_async4(void* smem_dst, const void* gmem_src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_u32(smem_dst)), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

template <int MODE, int RPW>
__global__ void __launch_bounds__(96) tile_scan_kernel(int B, int T, const float* __restrict__ rewards,
                                                       const float* __restrict__ values, const uint8_t* __restrict__ mask,
                                                       float gamma, float gl, float hgamma, float hgl,
                                                       const float* __restrict__ baselines, float* __restrict__ adv,
                                                       float* __restrict__ ret, double* moments) {
  constexpr bool HAS_V = (MODE == TS_GAE || MODE == TS_BILEVEL);  // the estimators that read values
  extern __shared__ __align__(16) uint8_t ts_smem[];
  using Stage = TsStage<RPW>;
  Stage* st = reinterpret_cast<Stage*>(ts_smem);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int row0 = blockIdx.x * RPW;
  const int nrows = min(RPW, B - row0);
  const int ntiles = (T + 31) >> 5;
  const uintptr_t mask_begin = (uintptr_t)mask, mask_end = (uintptr_t)(mask + (size_t)B * T);

  // Three warps share a row group.  Step i (tile ntiles-1-i, ring slot i & 3):
  //   warp 0 ("scan")   runs the recurrences of tile i, lane = row;
  //   warp 1 ("loader") issues the cp.async of tile i+TS_LEAD+1 into the slot stored during step i-1 and waits for tile i+1
  //                     (a tile has TS_LEAD steps to arrive: with one step of lead the step time WAS the HBM latency);
  //   warp 2 ("storer") writes tile i-1's results to HBM (lane = token, coalesced row segments) with the masked moments;
  // and one CTA barrier per step hands the tiles on.  A warp here is usually alone on
  // its scheduler, so each phase is "a batch of independent loads, the arithmetic, the stores": nothing else hides latency.
  auto issue = [&](int tile, int s) {
    const int t0 = tile << 5;
    const bool tok = t0 + lane < T;
    Stage& S = st[s];
    const size_t first = (size_t)row0 * T + t0;
    const float* vp = values + first + lane;
    const float* rp = rewards + first + lane;
    const uint8_t* mp = mask + first;
#pragma unroll 8
    for (int r = 0; r < nrows; ++r) {
      if (tok) {
        if (HAS_V) cp_async4(&S.v[r * TS_VS + lane], vp);
        cp_async4(&S.r[r * TS_VS + lane], rp);
      }
      if (lane < 9) {
        const uintptr_t w = (((uintptr_t)mp) & ~(uintptr_t)3) + 4u * lane;
        if (w >= mask_begin && w + 4 <= mask_end) {
          cp_async4(&S.m[r * TS_MS + 4 * lane], reinterpret_cast<const void*>(w));
        } else if (w < mask_end && w + 4 > mask_begin) {  // a word that straddles either end of the mask tensor: only its bytes that exist
          for (int b = 0; b < 4; ++b)
            if (w + b >= mask_begin && w + b < mask_end) S.m[r * TS_MS + 4 * lane + b] = *reinterpret_cast<const uint8_t*>(w + b);
        }
      }
      vp += T;
      rp += T;
      mp += T;
    }
  };

  double m1 = 0.0, m2 = 0.0;  // mover
  int n_masked = 0;           // mover: warp-uniform count of mask tokens (popcounts of ballots)
  auto store = [&](int tile, int s) {
    const int t0 = tile << 5;
    const int cnt = min(32, T - t0);
    Stage& S = st[s];
    const bool tok = lane < cnt;
    const size_t first = (size_t)row0 * T + t0;
    float* ap = adv + first + lane;
    float* tp = ret + first + lane;
    const uint8_t* mp = mask + first;
#pragma unroll
    for (int r0 = 0; r0 < RPW; r0 += 8) {
      if (r0 < nrows) {  // warp-uniform
        float a[8], b[8];
        uint8_t mk[8];
        const uint32_t mp_lo = (uint32_t)(uintptr_t)mp;  // only the low two bits of a mask row's address matter
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const int r = r0 + k;  // rows past nrows read stale shared memory and are never stored
          a[k] = S.r[r * TS_VS + lane];
          if (MODE != TS_RFPP) b[k] = S.v[r * TS_VS + lane];
          if (MODE != TS_REMAX) mk[k] = S.m[r * TS_MS + (int)((mp_lo + (uint32_t)(k * T)) & 3u) + lane];
        }
        auto put = [&](int k) {
          if (HAS_V) {
            ap[(size_t)k * T] = a[k];
            tp[(size_t)k * T] = b[k];
          } else if (MODE == TS_RFPP) {
            tp[(size_t)k * T] = a[k];
          } else {
            tp[(size_t)k * T] = a[k];
            ap[(size_t)k * T] = b[k];
          }
        };
        if ((cnt == 32) && (r0 + 8 <= nrows)) {  // warp-uniform: the common case stores without per-row branches
#pragma unroll
          for (int k = 0; k < 8; ++k) put(k);
        } else {
#pragma unroll
          for (int k = 0; k < 8; ++k)
            if (tok && (r0 + k < nrows)) put(k);
        }
        if (MODE != TS_REMAX) {
#pragma unroll
          for (int k = 0; k < 8; ++k) {
            const bool in = tok && (r0 + k < nrows) && mk[k] != 0;
            n_masked += __popc(__ballot_sync(0xffffffffu, in));
            const double am = (double)(in ? a[k] : 0.0f);  // select in fp32, convert unconditionally: no branch per row
            m1 += am;
            m2 = fma(am, am, m2);
          }
        }
        ap += (size_t)8 * T;
        tp += (size_t)8 * T;
        mp += (size_t)8 * T;
      }
    }
  };

  if (warp == 1) {
    // the row's last tile may be partial: its missing tokens are zeros, which leave every recurrence at its initial 0
    // (0 + c*0, and 0 added to the ReMax sum), so the scan needs no per-token bound check
    if (lane >= T - ((ntiles - 1) << 5)) {  // columns the cp.async of that tile never writes
      for (int r = 0; r < RPW; ++r) {
        st[0].v[r * TS_VS + lane] = 0.0f;
        st[0].r[r * TS_VS + lane] = 0.0f;
      }
    }
#pragma unroll
    for (int i = 0; i <= TS_LEAD; ++i) {
      if (i < ntiles) issue(ntiles - 1 - i, i);
      cp_async_commit();
    }
    cp_async_wait<TS_LEAD>();  // tile 0 has landed
  }
  __syncthreads();

  const bool mine = (warp == 0) && lane < nrows;
  const size_t my_base = (size_t)(row0 + (mine ? lane : 0)) * T;
  const float bl = (MODE == TS_REMAX && mine) ? baselines[row0 + lane] : 0.0f;
  const float bl0 = __fmul_rn(bl, 0.0f);
  float x = 0.0f, vn = 0.0f;  // GAE: A_{t+1}, V_{t+1};  REINFORCE++: x = running return
  float hi_lg = 0.0f, hi_nv = 0.0f;  // bi-level: turn-level lastgaelam and the value at the next reward position
  double acc = 0.0;           // ReMax: reverse cumulative sum
  for (int i = 0; i < ntiles; ++i) {
    const int tile = ntiles - 1 - i, s = i & (TS_STAGES - 1);
    if (warp == 0) {
      if (mine) {  // my row, right to left, 16 tokens at a time
        Stage& S = st[s];
        float* V = S.v + lane * TS_VS;
        float* R = S.r + lane * TS_VS;
        const uint8_t* M = S.m + lane * TS_MS + (int)(((uintptr_t)(mask + my_base + (tile << 5))) & 3);
#pragma unroll
        for (int hb = 1; hb >= 0; --hb) {
          float a[16], b[16];
          uint8_t mk[16];
#pragma unroll
          for (int k = 0; k < 16; ++k) {
            const int l = hb * 16 + 15 - k;
            a[k] = R[l];
            if (HAS_V) b[k] = V[l];
            if (MODE != TS_GAE) mk[k] = M[l];
          }
#pragma unroll
          for (int k = 0; k < 16; ++k) {
            if (MODE == TS_GAE) {
              const float v = b[k];
              const float d = __fsub_rn(__fadd_rn(a[k], __fmul_rn(gamma, vn)), v);
              x = __fadd_rn(d, __fmul_rn(gl, x));
              vn = v;
              a[k] = x;
              b[k] = __fadd_rn(x, v);
            } else if (MODE == TS_BILEVEL) {
              // ragen/trainer/core_algos.py:44-88 as ONE right-to-left sweep (see bilevel_gae_kernel): x = token-level lastgaelam,
              // vn = value of the nearest loss-mask token to the right (0 when there is none: the reference would raise), hi_lg /
              // hi_nv = turn-level state.  Reward positions are a handful per row, so the divergent branch is rare.
              const float r = a[k], v = b[k];
              const bool m = mk[k] != 0;
              float ao = 0.0f, ro = 0.0f;
              if (r != 0.0f) {
                hi_lg = __fadd_rn(__fsub_rn(__fadd_rn(r, __fmul_rn(hgamma, hi_nv)), v), __fmul_rn(hgl, hi_lg));
                hi_nv = v;
                const float upd = __fadd_rn(hi_lg, v);
                ao = hi_lg;
                ro = upd;
                if (m) {  // the token-level pass restarts here: next value 0, lastgaelam 0
                  const float d = __fsub_rn(__fadd_rn(upd, __fmul_rn(gamma, 0.0f)), v);
                  x = __fadd_rn(d, __fmul_rn(gl, 0.0f));
                  ao = x;
                  ro = __fadd_rn(x, v);
                  vn = v;
                }
              } else if (m) {
                const float d = __fsub_rn(__fadd_rn(r, __fmul_rn(gamma, vn)), v);
                x = __fadd_rn(d, __fmul_rn(gl, x));
                ao = x;
                ro = __fadd_rn(x, v);
                vn = v;
              }
              a[k] = ao;
              b[k] = ro;
            } else if (MODE == TS_RFPP) {
              x = __fadd_rn(a[k], __fmul_rn(gamma, x));
              a[k] = x;
              x = mk[k] ? x : __fmul_rn(x, 0.0f);
            } else {
              acc += (double)(mk[k] ? a[k] : __fmul_rn(a[k], 0.0f));
              const float c = (float)acc;
              a[k] = c;
              b[k] = __fsub_rn(c, mk[k] ? bl : bl0);
            }
          }
#pragma unroll
          for (int k = 0; k < 16; ++k) {
            const int l = hb * 16 + 15 - k;
            R[l] = a[k];
            if (MODE != TS_RFPP) V[l] = b[k];
          }
        }
      }
    } else if (warp == 1) {
      // into the slot whose tile was stored during step i-1 (i + TS_LEAD + 1 = i - 2 modulo the ring)
      if (i + TS_LEAD + 1 < ntiles) issue(tile - (TS_LEAD + 1), (i + TS_LEAD + 1) & (TS_STAGES - 1));
      cp_async_commit();
      cp_async_wait<TS_LEAD>();  // tile i+1 has landed (the TS_LEAD tiles after it may still be in flight)
    } else {
      if (i >= 1) store(tile + 1, (i - 1) & (TS_STAGES - 1));
    }
    __syncthreads();
  }
  if (warp == 2) {
    store(0, (ntiles - 1) & (TS_STAGES - 1));
    if (MODE != TS_REMAX && moments) {
      m1 = warp_sum(m1);
      m2 = warp_sum(m2);
      if (lane == 0) {
        atomicAdd(moments + 0, (double)n_masked);
        atomicAdd(moments + 1, m1);
        atomicAdd(moments + 2, m2);
      }
    }
  }
}

// ---- GAE as a parallel scan (ragen_gae_scan) ----------------------------------------------------------------------------------
// A_t = delta_t + c * A_{t+1} with delta_t = r_t + gamma * V_{t+1} - V_t and c = gamma * lam is a linear recurrence: its solution
// A_t = sum_{k >= t} c^(k - t) * delta_k can be formed in any order.  A CTA of 256 threads owns a row and takes 2048 tokens of it
// per step (C5's T = 2047: the whole row at once): the tile comes in with full-line loads, thread i runs the recurrence over
// ITS 8 consecutive tokens (from shared memory, 9-word pitch: conflict-free), the 256 per-thread totals are combined by a
// warp-shuffle suffix scan (weights c^8, c^16 ... c^128) and a pass over the 8 warp totals, every thread folds what lies to its
// right into its 8 values, and the tile leaves through shared memory with full-line stores.  Four barriers per 2048 tokens
// instead of a dependent chain per token.  CTAs are persistent over rows and keep TWO tiles in shared memory: the cp.async
// loads of the next tile run under the scan and the stores of the current one.  The masked moments (fp64) cost one set of
// atomics per CTA.  The result differs from the reference's sequential fp32 loop by rounding only (a few ulps of the largest
// partial sum; north_star's bar is 1e-5) - callers that want the reference's bits take ragen_gae.
constexpr int BS_T = 256;              // threads
constexpr int BS_K = 8;                // tokens per thread
constexpr int BS_CH = BS_T * BS_K;     // tokens per step
constexpr int BS_PITCH = BS_K + 1;     // words of shared memory per thread

struct __align__(16) ScanStage {
  float r[BS_T * BS_PITCH + 1];
  float v[BS_T * BS_PITCH + 1];
};

__global__ void __launch_bounds__(BS_T, 4) gae_scan_kernel(int B, int T, const float* __restrict__ rewards,
                                                      const float* __restrict__ values, const uint8_t* __restrict__ mask,
                                                      float gamma, float c, float* __restrict__ adv, float* __restrict__ ret,
                                                      double* moments) {
  // two stages: while tile i is scanned and stored, the 4-byte cp.async of tile i + 1 (the next chunk of the row, or the next
  // row of this CTA) are in flight into the other stage - the loads never wait for the scan
  __shared__ ScanStage stage[2];
  __shared__ float s_w[BS_T / 32];
  __shared__ float s_carry[2];
  __shared__ double s_mom[BS_T / 32][3];
  const unsigned FULL = 0xffffffffu;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  float cp[BS_K + 1];  // c^0 .. c^8
  cp[0] = 1.0f;
#pragma unroll
  for (int k = 1; k <= BS_K; ++k) cp[k] = cp[k - 1] * c;
  const float w1 = cp[BS_K], w2 = w1 * w1, w4 = w2 * w2, w8 = w4 * w4, w16 = w8 * w8, w32 = w16 * w16;  // (c^8)^(1, 2, 4, 8, 16, 32)
  float lanew = 1.0f;  // (c^8)^(32 - lane): what a thread's total weighs in for everything right of its warp
  {
    float base = w1;
    int e = 32 - lane;
    while (e) {
      if (e & 1) lanew *= base;
      base *= base;
      e >>= 1;
    }
  }
  double m0 = 0.0, m1 = 0.0, m2 = 0.0;
  const int n_chunks = (T + BS_CH - 1) / BS_CH;
  const int my_rows = blockIdx.x < B ? (B - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;
  const int n_items = my_rows * n_chunks;  // (row, chunk) tiles of this CTA: rows in grid-stride order, chunks right to left

  // queue the loads of tile `it` into stage `st`.  raw[j] receives the tile's mask bytes, 8 per thread in the COALESCED order
  // (token tid + 256 j): the store phase below handles exactly those tokens, so the mask never passes through shared memory.
  // The eight byte loads are issued FIRST and looked at a barrier later (pack_mask): asm volatile is a compiler fence, loads
  // mixed in between the cp.async would each be waited for in turn
  auto issue = [&](int it, int st, uint32_t (&raw)[BS_K]) {
    const int row = blockIdx.x + (it / n_chunks) * gridDim.x;
    const int t0 = (n_chunks - 1 - it % n_chunks) * BS_CH;
    const float* r = rewards + (size_t)row * T;
    const float* v = values + (size_t)row * T;
    const uint8_t* m = mask + (size_t)row * T;
    ScanStage& S = stage[st];
#pragma unroll
    for (int j = 0; j < BS_K; ++j) {
      const int t = t0 + tid + BS_T * j;
      raw[j] = t < T ? (uint32_t)m[t] : 0u;
    }
#pragma unroll
    for (int j = 0; j < BS_K; ++j) {
      const int g = tid + BS_T * j, t = t0 + g;
      const int idx = (g >> 3) * BS_PITCH + (g & 7);
      if (t < T) {
        cp_async4(&S.r[idx], r + t);
        cp_async4(&S.v[idx], v + t);
      } else {
        S.r[idx] = 0.0f;
        S.v[idx] = 0.0f;
      }
    }
  };
  auto pack_mask = [&](const uint32_t (&raw)[BS_K]) -> uint32_t {
    uint32_t bits = 0;
#pragma unroll
    for (int j = 0; j < BS_K; ++j) bits |= (raw[j] ? 1u : 0u) << j;
    return bits;
  };

  float carry = 0.0f;   // A of the first token right of the tile
  float v_next = 0.0f;  // V of that token (V_T = 0, core_algos.py:100)
  uint32_t mm = 0;  // bit j: mask of token tid + 256 j of the current tile
  uint32_t raw[BS_K] = {};
  if (n_items > 0) issue(0, 0, raw);
  cp_async_commit();
  mm = pack_mask(raw);
  for (int it = 0; it < n_items; ++it) {
    const int st = it & 1;
#pragma unroll
    for (int j = 0; j < BS_K; ++j) raw[j] = 0u;
    if (it + 1 < n_items) issue(it + 1, st ^ 1, raw);  // that stage was last read before the barrier that ended tile it - 1
    cp_async_commit();
    cp_async_wait<1>();  // this thread's copies of tile `it` have landed
    const int ch = n_chunks - 1 - it % n_chunks;
    const int row = blockIdx.x + (it / n_chunks) * gridDim.x;
    const int t0 = ch * BS_CH;
    if (ch == n_chunks - 1) {  // a new row starts at its right end
      carry = 0.0f;
      v_next = 0.0f;
    }
    ScanStage& S = stage[st];
    float* a_out = adv + (size_t)row * T;
    float* r_out = ret + (size_t)row * T;
    if (tid == 0) S.v[BS_T * BS_PITCH] = v_next;
    __syncthreads();  // everybody's copies are visible
    const int base = tid * BS_PITCH;
    float vv[BS_K + 1];
#pragma unroll
    for (int k = 0; k < BS_K; ++k) vv[k] = S.v[base + k];
    vv[BS_K] = S.v[base + BS_PITCH];  // the next thread's first token (the tile's right neighbour for the last thread)
    float x[BS_K];
    float acc = 0.0f;
#pragma unroll
    for (int k = BS_K - 1; k >= 0; --k) {
      const bool in = t0 + BS_K * tid + k < T;
      const float delta = in ? (S.r[base + k] + gamma * vv[k + 1] - vv[k]) : 0.0f;
      acc = delta + c * acc;
      x[k] = acc;
    }
    // suffix scan of the 256 per-thread totals
    float P = x[0], y;
    y = __shfl_down_sync(FULL, P, 1);
    if (lane + 1 < 32) P += w1 * y;
    y = __shfl_down_sync(FULL, P, 2);
    if (lane + 2 < 32) P += w2 * y;
    y = __shfl_down_sync(FULL, P, 4);
    if (lane + 4 < 32) P += w4 * y;
    y = __shfl_down_sync(FULL, P, 8);
    if (lane + 8 < 32) P += w8 * y;
    y = __shfl_down_sync(FULL, P, 16);
    if (lane + 16 < 32) P += w16 * y;
    if (lane == 0) s_w[warp] = P;
    __syncthreads();
    float G = carry;  // A of the first token of the next warp (of the tile to the right for the last warp)
    for (int w = BS_T / 32 - 1; w > warp; --w) G = s_w[w] + w32 * G;
    const float Sx = P + lanew * G;  // A of this thread's first token
    float In = __shfl_down_sync(FULL, Sx, 1);
    if (lane == 31) In = G;
#pragma unroll
    for (int k = 0; k < BS_K; ++k) S.r[base + k] = x[k] + cp[BS_K - k] * In;  // only this thread reads these words before the barrier
    if (tid == 0) {
      s_carry[0] = Sx;
      s_carry[1] = vv[0];
    }
    __syncthreads();
    carry = s_carry[0];
    v_next = s_carry[1];
#pragma unroll
    for (int j = 0; j < BS_K; ++j) {
      const int g = tid + BS_T * j, t = t0 + g;
      if (t < T) {
        const int idx = (g >> 3) * BS_PITCH + (g & 7);
        const float a = S.r[idx];
        a_out[t] = a;
        r_out[t] = a + S.v[idx];  // returns = advantages + values (:107)
        if ((mm >> j) & 1u) {
          m0 += 1.0;
          m1 += (double)a;
          m2 += (double)a * (double)a;
        }
      }
    }
    mm = pack_mask(raw);  // the next tile's mask bytes: in flight since the top of this iteration
    __syncthreads();  // the tile after next is loaded into this stage
  }
  if (moments) {
    m0 = warp_sum(m0);
    m1 = warp_sum(m1);
    m2 = warp_sum(m2);
    if (lane == 0) {
      s_mom[warp][0] = m0;
      s_mom[warp][1] = m1;
      s_mom[warp][2] = m2;
    }
    __syncthreads();
    if (tid < 3) {
      double acc = 0.0;
      for (int w = 0; w < BS_T / 32; ++w) acc += s_mom[w][tid];
      atomicAdd(moments + tid, acc);
    }
  }
}

// ---- the same scan with bulk-copy staging (ragen_gae_scan's fast path) -----------------------------------------------------------
// ncu on the kernel above (profiles/r28b_gae_scan_cpasync_raw.csv): the LSU pipe is the busiest unit (52 %), a run that ONLY
// loads the tiles takes 51 us of the 73 - 4-byte copies cost one LSU slot per token and array.  Here the tiles travel by 1-D bulk
// copies (cp.async.bulk, no LSU slot at all) and the shared memory is read and written with 128-bit accesses.  Bulk copies need
// 16-byte-aligned addresses, a row of an odd T starts anywhere, so a tile is cut in ALIGNED FLAT WORDS, not in tokens: the tile of
// row `row`, chunk c covers the words [al_lo + c * 2304, ...) of the flat [B * T] tensors, where al_lo is the row's first word
// rounded down to a multiple of four; words of the tile that belong to the neighbouring rows are staged but masked out of the
// recurrence (right of the row's end everything is zero, left of its start nothing is kept).  The mask bytes of those words sit
// at the same flat index.  Results leave by bulk stores too - whole 16-byte groups inside the row - and the up to three words at
// either end of a row by ordinary stores.  288 threads x 8 words = 2304 words per tile: a row of T = 2047 fits one tile whatever
// its alignment.  Needs 16-byte-aligned base pointers and B * T % 16 == 0; everything else takes the kernel above.
using scan_tiles::BT_K;
using scan_tiles::BT_NW;
using scan_tiles::BT_T;
using scan_tiles::BT_W;
using scan_tiles::ScanTile;
using scan_tiles::TileGeom;

struct __align__(128) BulkStage {
  float r[BT_W];           // rewards in, advantages out
  float v[BT_W];           // values in, returns out
  uint8_t m[BT_W + 48];    // mask bytes of the tile's words, from the 16-byte boundary at or below the tile's first word
};

// tiles of a CTA: rows in grid-stride order, the chunks of a row right to left (geometry: scan_tiles.cuh)
__device__ __forceinline__ ScanTile bt_first(int B, int T) {
  ScanTile t{(int)blockIdx.x, 0};
  if (t.row < B) t.c = scan_tiles::row_chunks(t.row, T) - 1;
  return t;
}
__device__ __forceinline__ ScanTile bt_advance(ScanTile t, int B, int T) {
  if (t.c > 0) return ScanTile{t.row, t.c - 1};
  t.row += gridDim.x;
  t.c = t.row < B ? scan_tiles::row_chunks(t.row, T) - 1 : 0;
  return t;
}
__device__ __forceinline__ TileGeom bt_geom(ScanTile t, int T) { return scan_tiles::geom(t, T); }
// one thread: arm the stage's barrier and queue the three bulk loads of tile t
__device__ __forceinline__ void bt_issue(ScanTile t, int T, BulkStage& S, uint64_t* bar, const float* rewards, const float* values,
                                         const uint8_t* mask) {
  const TileGeom g = bt_geom(t, T);
  int64_t mb0, mb1;
  scan_tiles::mask_window(g, mb0, mb1);  // B * T % 16 == 0: never past the tensor
  const uint32_t wbytes = (uint32_t)g.nw * 4u, mbytes = (uint32_t)(mb1 - mb0);
  mbar_expect_tx(bar, (values ? 2u : 1u) * wbytes + mbytes);
  tma_load_1d(S.r, rewards + g.cw0, wbytes, bar);
  if (values) tma_load_1d(S.v, values + g.cw0, wbytes, bar);  // the returns-only estimators have no values
  tma_load_1d(S.m, mask + mb0, mbytes, bar);
}
// warp 0, after the barrier that follows the tile's last shared-memory write (and fence_async_smem): whole 16-byte groups of the
// row leave by bulk store (lane 0, one committed group), the words at the row's ends one by one (lanes 1..11)
// (S.r -> adv, S.v -> ret; ret may be NULL)
__device__ __forceinline__ void bt_store(const TileGeom& g, BulkStage& S, float* adv, float* ret, int lane) {
  const scan_tiles::StorePlan sp = scan_tiles::store_plan(g);
  const int ra0 = sp.ra0, ra1 = sp.ra1, rs0 = sp.rs0, rs1 = sp.rs1, left_end = sp.left_end;
  const bool bulk = sp.bulk;
  if (lane == 0 && bulk) {
    tma_store_1d(adv + g.cw0 + rs0, S.r + rs0, (uint32_t)(rs1 - rs0) * 4u);
    if (ret) tma_store_1d(ret + g.cw0 + rs0, S.v + rs0, (uint32_t)(rs1 - rs0) * 4u);
  }
  if (lane == 0) tma_store_commit();
  int w = -1;
  if (lane >= 1 && lane <= 8 && ra0 + lane - 1 < left_end) w = ra0 + lane - 1;
  if (lane >= 9 && lane <= 11 && bulk && rs1 + lane - 9 < ra1) w = rs1 + lane - 9;
  if (w >= 0) {
    adv[g.cw0 + w] = S.r[w];
    if (ret) ret[g.cw0 + w] = S.v[w];
  }
  __syncwarp();
}
// fp64 masked moments of a CTA -> three atomics
__device__ __forceinline__ void bt_add_moments(double* moments, double m0, double m1, double m2, double (*s_mom)[3], int tid) {
  if (!moments) return;
  const int lane = tid & 31, warp = tid >> 5;
  m0 = warp_sum(m0);
  m1 = warp_sum(m1);
  m2 = warp_sum(m2);
  if (lane == 0) {
    s_mom[warp][0] = m0;
    s_mom[warp][1] = m1;
    s_mom[warp][2] = m2;
  }
  __syncthreads();
  if (tid < 3) {
    double acc = 0.0;
    for (int w = 0; w < BT_NW; ++w) acc += s_mom[w][tid];
    atomicAdd(moments + tid, acc);
  }
}

template <int NS>
__global__ void __launch_bounds__(BT_T) gae_scan_bulk_kernel(int B, int T, const float* __restrict__ rewards,
                                                           const float* __restrict__ values, const uint8_t* __restrict__ mask,
                                                           float gamma, float c, float* __restrict__ adv, float* __restrict__ ret,
                                                           double* moments) {
  extern __shared__ __align__(128) uint8_t bulk_smem[];
  BulkStage* stage = reinterpret_cast<BulkStage*>(bulk_smem);
  __shared__ __align__(8) uint64_t full[NS];
  __shared__ float s_w[BT_NW];
  __shared__ float s_carry[2];
  __shared__ double s_mom[BT_NW][3];
  const unsigned FULL = 0xffffffffu;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  float cp[BT_K + 1];  // c^0 .. c^8
  cp[0] = 1.0f;
#pragma unroll
  for (int k = 1; k <= BT_K; ++k) cp[k] = cp[k - 1] * c;
  const float w1 = cp[BT_K], w2 = w1 * w1, w4 = w2 * w2, w8 = w4 * w4, w16 = w8 * w8, w32 = w16 * w16;
  float lanew = 1.0f;  // (c^8)^(32 - lane)
  {
    float base = w1;
    int e = 32 - lane;
    while (e) {
      if (e & 1) lanew *= base;
      base *= base;
      e >>= 1;
    }
  }
  if (tid == 0) {
    for (int s = 0; s < NS; ++s) mbar_init(&full[s], 1);
    fence_mbar_init();
  }
  __syncthreads();
  ScanTile cur = bt_first(B, T);
  ScanTile nxt = cur;  // the next tile to load (tid 0's)
  if (tid == 0) {
    for (int s = 0; s < NS && nxt.row < B; ++s) {
      bt_issue(nxt, T, stage[s], &full[s], rewards, values, mask);
      nxt = bt_advance(nxt, B, T);
    }
  }
  double m0 = 0.0, m1 = 0.0, m2 = 0.0;
  float carry = 0.0f;   // A of the first word right of the tile
  float v_next = 0.0f;  // V of that word
  for (int it = 0; cur.row < B; ++it, cur = bt_advance(cur, B, T)) {
    const int st = it % NS;
    const TileGeom g = bt_geom(cur, T);
    if (g.right_end) {  // V_T = 0 and nothing to carry in (core_algos.py:100)
      carry = 0.0f;
      v_next = 0.0f;
    }
    BulkStage& S = stage[st];
    mbar_wait(&full[st], (uint32_t)((it / NS) & 1));
    const int w0 = BT_K * tid;
    const float4 r0 = reinterpret_cast<const float4*>(S.r)[2 * tid], r1 = reinterpret_cast<const float4*>(S.r)[2 * tid + 1];
    const float4 v0 = reinterpret_cast<const float4*>(S.v)[2 * tid], v1 = reinterpret_cast<const float4*>(S.v)[2 * tid + 1];
    const uint32_t* mwp = reinterpret_cast<const uint32_t*>(S.m + g.mshift + w0);  // mshift is a multiple of four, w0 of eight
    const uint2 mw = make_uint2(mwp[0], mwp[1]);
    float vn = __shfl_down_sync(FULL, v0.x, 1);
    if (lane == 31) vn = (tid == BT_T - 1) ? v_next : S.v[w0 + BT_K];
    const float rr[BT_K] = {r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w};
    const float vv[BT_K + 1] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w, vn};
    float x[BT_K];
    float acc = 0.0f;
#pragma unroll
    for (int k = BT_K - 1; k >= 0; --k) {
      const bool in = (w0 + k >= g.rel_lo) && (w0 + k < g.rel_hi);
      const float vnext = (w0 + k + 1 < g.rel_hi) ? vv[k + 1] : 0.0f;  // selects, never products: words of other rows may hold anything
      const float delta = in ? (rr[k] + gamma * vnext - vv[k]) : 0.0f;
      acc = delta + c * acc;
      x[k] = acc;
    }
    float P = x[0], y;
    y = __shfl_down_sync(FULL, P, 1);
    if (lane + 1 < 32) P += w1 * y;
    y = __shfl_down_sync(FULL, P, 2);
    if (lane + 2 < 32) P += w2 * y;
    y = __shfl_down_sync(FULL, P, 4);
    if (lane + 4 < 32) P += w4 * y;
    y = __shfl_down_sync(FULL, P, 8);
    if (lane + 8 < 32) P += w8 * y;
    y = __shfl_down_sync(FULL, P, 16);
    if (lane + 16 < 32) P += w16 * y;
    if (lane == 0) s_w[warp] = P;
    __syncthreads();
    float G = carry;
    for (int w = BT_NW - 1; w > warp; --w) G = s_w[w] + w32 * G;
    const float Sx = P + lanew * G;  // A of this thread's first word
    float In = __shfl_down_sync(FULL, Sx, 1);
    if (lane == 31) In = G;
    float a[BT_K];
#pragma unroll
    for (int k = 0; k < BT_K; ++k) {
      a[k] = x[k] + cp[BT_K - k] * In;
      const bool in = (w0 + k >= g.rel_lo) && (w0 + k < g.rel_hi);
      const unsigned mb = ((k < 4 ? mw.x : mw.y) >> (8 * (k & 3))) & 0xffu;
      if (in && mb) {
        m0 += 1.0;
        m1 += (double)a[k];
        m2 += (double)a[k] * (double)a[k];
      }
    }
    reinterpret_cast<float4*>(S.r)[2 * tid] = make_float4(a[0], a[1], a[2], a[3]);
    reinterpret_cast<float4*>(S.r)[2 * tid + 1] = make_float4(a[4], a[5], a[6], a[7]);
    reinterpret_cast<float4*>(S.v)[2 * tid] = make_float4(a[0] + vv[0], a[1] + vv[1], a[2] + vv[2], a[3] + vv[3]);  // returns (:107)
    reinterpret_cast<float4*>(S.v)[2 * tid + 1] = make_float4(a[4] + vv[4], a[5] + vv[5], a[6] + vv[6], a[7] + vv[7]);
    if (tid == 0) {
      s_carry[0] = Sx;
      s_carry[1] = vv[0];
    }
    fence_async_smem();  // the bulk stores below read what was just written
    __syncthreads();
    carry = s_carry[0];
    v_next = s_carry[1];
    if (warp == 0) {
      bt_store(g, S, adv, ret, lane);
      if (lane == 0) {
        tma_store_wait_read();  // the stage is free once the bulk stores have read it
        if (nxt.row < B) {
          bt_issue(nxt, T, S, &full[st], rewards, values, mask);
          nxt = bt_advance(nxt, B, T);
        }
      }
    }
  }
  bt_add_moments(moments, m0, m1, m2, s_mom, tid);
}

// ---- bi-level GAE as block scans (ragen_bilevel_gae's many-row path) ---------------------------------------------------------------
// ragen/trainer/core_algos.py:44-88 read as maps on a state that travels right to left, so that every piece is an associative
// composition and a 2304-word tile is done by 288 threads at once instead of by one lane:
//   Z ("value of the next loss-mask token"):   X -> m_t ? V_t : X          E ("value of the next reward position"): X -> e_t ? V_t : X
//   H (turn-level lastgaelam, :62-71):         X -> e_t ? (r_t + hg * E_in - V_t) + hg*lam * X : X
//   L (token-level lastgaelam, :73-88):        X -> m_t ? (e_t ? (H_t + V_t) - V_t : (gamma * Z_in - V_t) + gamma*lam * X) : X
// where Z_in / E_in are the states ENTERING position t from the right and e_t = (r_t != 0).  Each round composes a thread's 8
// maps, suffix-scans the 288 compositions (warp shuffles, then the 9 warp totals), and re-walks the 8 words with the true
// incoming state.  Three rounds (Z and E together, H, L), one CTA barrier each.  Equal to the sequential kernel up to fp32
// rounding (tests: 1e-5 of the tensor's scale); ragen_bilevel_gae keeps the sequential kernel below 4096 rows and for tensors
// the bulk copies cannot take.
template <int NS, int BI_CTAS>
__global__ void __launch_bounds__(BT_T, BI_CTAS) bilevel_scan_bulk_kernel(int B, int T, const float* __restrict__ rewards,
                                                               const float* __restrict__ values, const uint8_t* __restrict__ mask,
                                                               float gamma, float gl, float hgamma, float hgl,
                                                               float* __restrict__ adv, float* __restrict__ ret, double* moments) {
  extern __shared__ __align__(128) uint8_t bulk_smem[];
  BulkStage* stage = reinterpret_cast<BulkStage*>(bulk_smem);
  __shared__ __align__(8) uint64_t full[NS];
  __shared__ float s_zv[BT_NW], s_ev[BT_NW], s_ha[BT_NW], s_hb[BT_NW], s_ta[BT_NW], s_tb[BT_NW];
  __shared__ uint8_t s_zh[BT_NW], s_eh[BT_NW];
  __shared__ float s_carry[4];
  __shared__ double s_mom[BT_NW][3];
  const unsigned FULL = 0xffffffffu;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) {
    for (int s = 0; s < NS; ++s) mbar_init(&full[s], 1);
    fence_mbar_init();
  }
  __syncthreads();
  ScanTile cur = bt_first(B, T);
  ScanTile nxt = cur;
  if (tid == 0) {
    for (int s = 0; s < NS && nxt.row < B; ++s) {
      bt_issue(nxt, T, stage[s], &full[s], rewards, values, mask);
      nxt = bt_advance(nxt, B, T);
    }
  }
  double m0 = 0.0, m1 = 0.0, m2 = 0.0;
  float z_c = 0.0f, e_c = 0.0f, h_c = 0.0f, t_c = 0.0f;  // the four states entering the tile from the right
  for (int it = 0; cur.row < B; ++it, cur = bt_advance(cur, B, T)) {
    const int st = it % NS;
    const TileGeom g = bt_geom(cur, T);
    if (g.right_end) z_c = e_c = h_c = t_c = 0.0f;  // no next value, both lastgaelam 0 (:60, :75; a missing next loss-mask value reads 0)
    BulkStage& S = stage[st];
    mbar_wait(&full[st], (uint32_t)((it / NS) & 1));
    const int w0 = BT_K * tid;
    // every use of a word's reward / value below sits behind a select keyed by the word's bit in mk / ek, and those bits are
    // zero off the row: words of other rows (or stale ones past a short tile) may hold anything, NaN included
    const float4 r0 = reinterpret_cast<const float4*>(S.r)[2 * tid], r1 = reinterpret_cast<const float4*>(S.r)[2 * tid + 1];
    const float4 v0 = reinterpret_cast<const float4*>(S.v)[2 * tid], v1 = reinterpret_cast<const float4*>(S.v)[2 * tid + 1];
    const float vv[BT_K] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
    unsigned mk, ek = 0;  // bit k: word k is a loss-mask token / a reward position of this row
    {
      const int lo_k = min(max(g.rel_lo - w0, 0), BT_K), hi_k = min(max(g.rel_hi - w0, 0), BT_K);
      const unsigned inb = ((1u << hi_k) - 1u) & ~((1u << lo_k) - 1u);  // words of the row among the eight
      const uint32_t* mwp = reinterpret_cast<const uint32_t*>(S.m + g.mshift + w0);
      // one bit per non-zero byte: 0x01 per byte, then the four bits gathered by a multiply (no two partial products meet)
      const uint32_t nlo = __vsetne4(mwp[0], 0u), nhi = __vsetne4(mwp[1], 0u);
      mk = ((((nlo * 0x00204081u) >> 21) & 0xfu) | ((((nhi * 0x00204081u) >> 21) & 0xfu) << 4)) & inb;
      const float r_[BT_K] = {r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w};  // not kept: the few reward positions re-read theirs
#pragma unroll
      for (int k = 0; k < BT_K; ++k) ek |= (r_[k] != 0.0f) ? (1u << k) : 0u;  // reward_mask = token_level_rewards.bool() (:50)
      ek &= inb;
    }
    const unsigned ew = __reduce_or_sync(FULL, ek);  // warp-uniform: which of the eight steps meet a reward position in ANY lane
    // ---- round 1: Z and E --------------------------------------------------------------------------------------------------
    float zv = 0.0f, ev = 0.0f;
#pragma unroll
    for (int k = BT_K - 1; k >= 0; --k) {  // the leftmost flagged word of the eight wins
      zv = ((mk >> k) & 1u) ? vv[k] : zv;
      if ((ew >> k) & 1u) ev = ((ek >> k) & 1u) ? vv[k] : ev;
    }
    const bool zh = mk != 0, eh = ek != 0;
    // a select map needs no scan: the state entering a lane is the value of the nearest flagged lane to its right (ballot + one
    // shuffle), else whatever enters the warp
    const unsigned zbits = __ballot_sync(FULL, zh), ebits = __ballot_sync(FULL, eh);
    const unsigned above = lane == 31 ? 0u : (0xfffffffeu << lane);
    const unsigned zr = zbits & above, er = ebits & above;  // flagged lanes right of this one
    const float z_from = __shfl_sync(FULL, zv, zr ? __ffs(zr) - 1 : 0), e_from = __shfl_sync(FULL, ev, er ? __ffs(er) - 1 : 0);
    const float z_first = __shfl_sync(FULL, zv, zbits ? __ffs(zbits) - 1 : 0), e_first = __shfl_sync(FULL, ev, ebits ? __ffs(ebits) - 1 : 0);
    if (lane == 0) {
      s_zv[warp] = z_first; s_zh[warp] = zbits ? 1 : 0;
      s_ev[warp] = e_first; s_eh[warp] = ebits ? 1 : 0;
    }
    __syncthreads();
    float Gz = z_c, Ge = e_c;  // the states entering this WARP
#pragma unroll
    for (int w = BT_NW - 1; w >= 1; --w) {  // all loads up front, the same work in every warp
      const float wz = s_zv[w], we = s_ev[w];
      const bool hz = s_zh[w] != 0 && w > warp, he = s_eh[w] != 0 && w > warp;
      Gz = hz ? wz : Gz;
      Ge = he ? we : Ge;
    }
    const float In_z = zr ? z_from : Gz, In_e = er ? e_from : Ge;  // the states entering this thread from the right
    const float Sz = zbits ? z_first : Gz, Se = ebits ? e_first : Ge;  // ... and leaving the warp on the left (tid 0: the tile)
    // A warp whose 256 words hold no reward position (three warps in four at a handful of rewards per row) takes the short way
    // through both rounds: its H maps are the identity and its L walks need no reward tests.  any_e is warp-uniform; the CTA
    // barriers stay outside the branches.
    const bool any_e = ew != 0;
    // ---- round 2: H, the turn-level recurrence over the reward positions -----------------------------------------------------------
    float PA = 1.0f, PB = 0.0f;
    if (any_e) {
      float E = In_e;
#pragma unroll
      for (int k = BT_K - 1; k >= 0; --k)
        if ((ew >> k) & 1u) {  // a handful of reward positions per row: most steps are skipped by the whole warp
          const bool e = (ek >> k) & 1u;
          const float nb = (S.r[w0 + k] + hgamma * E - vv[k]) + hgl * PB;
          PB = e ? nb : PB;
          PA = e ? hgl * PA : PA;
          E = e ? vv[k] : E;
        }
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const float a2 = __shfl_down_sync(FULL, PA, o), b2 = __shfl_down_sync(FULL, PB, o);
        const bool on = lane + o < 32;
        PB = on ? PB + PA * b2 : PB;
        PA = on ? PA * a2 : PA;
      }
    }
    if (lane == 0) { s_ha[warp] = PA; s_hb[warp] = PB; }
    __syncthreads();
    float Gh = h_c;
    if (any_e || warp == 0) {  // warp 0 always: its first thread carries the state out of the tile
#pragma unroll
      for (int w = BT_NW - 1; w >= 1; --w) {
        const float nh = s_hb[w] + s_ha[w] * Gh;
        Gh = w > warp ? nh : Gh;
      }
    }
    const float Sh = PB + PA * Gh;
    float hv[BT_K];  // turn-level advantage at the thread's reward positions
#pragma unroll
    for (int k = 0; k < BT_K; ++k) hv[k] = 0.0f;
    if (any_e) {
      float In_h = __shfl_down_sync(FULL, Sh, 1);
      if (lane == 31) In_h = Gh;
      float E = In_e, H = In_h;
#pragma unroll
      for (int k = BT_K - 1; k >= 0; --k)
        if ((ew >> k) & 1u) {
          const bool e = (ek >> k) & 1u;
          const float nh = (S.r[w0 + k] + hgamma * E - vv[k]) + hgl * H;
          H = e ? nh : H;
          hv[k] = H;
          E = e ? vv[k] : E;
        }
    }
    // ---- round 3: L, the token-level recurrence over the loss-mask tokens ----------------------------------------------------------
    // one step of the walk: the map of word k applied to (A, B) (compose) or to X (re-walk)
    auto step_map = [&](auto has_e, int k, float Z, float& ak, float& bk) {
      ak = gl;
      bk = gamma * Z - vv[k];
      if constexpr (decltype(has_e)::value) {
        if ((ew >> k) & 1u) {  // restart from the updated reward: next value 0, lastgaelam 0 (:78-81)
          const bool e = (ek >> k) & 1u;
          ak = e ? 0.0f : ak;
          bk = e ? (hv[k] + vv[k]) - vv[k] : bk;
        }
      }
    };
    auto compose3 = [&](auto has_e, float& At, float& Bt) {
      float Z = In_z;
#pragma unroll
      for (int k = BT_K - 1; k >= 0; --k) {
        const bool m = (mk >> k) & 1u;
        float ak, bk;
        step_map(has_e, k, Z, ak, bk);
        Bt = m ? bk + ak * Bt : Bt;
        At = m ? ak * At : At;
        Z = m ? vv[k] : Z;
      }
    };
    PA = 1.0f;
    PB = 0.0f;
    if (any_e) compose3(std::true_type{}, PA, PB);
    else compose3(std::false_type{}, PA, PB);
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const float a2 = __shfl_down_sync(FULL, PA, o), b2 = __shfl_down_sync(FULL, PB, o);
      const bool on = lane + o < 32;
      PB = on ? PB + PA * b2 : PB;
      PA = on ? PA * a2 : PA;
    }
    if (lane == 0) { s_ta[warp] = PA; s_tb[warp] = PB; }
    __syncthreads();
    float Gt = t_c;
#pragma unroll
    for (int w = BT_NW - 1; w >= 1; --w) {
      const float nt = s_tb[w] + s_ta[w] * Gt;
      Gt = w > warp ? nt : Gt;
    }
    const float St = PB + PA * Gt;
    float In_t = __shfl_down_sync(FULL, St, 1);
    if (lane == 31) In_t = Gt;
    float a[BT_K], rt[BT_K];
    float s1 = 0.0f, s2 = 0.0f;  // the thread's eight masked advantages summed in fp32, then added to the fp64 totals
    auto rewalk3 = [&](auto has_e) {
      float Z = In_z, X = In_t;
#pragma unroll
      for (int k = BT_K - 1; k >= 0; --k) {
        const bool m = (mk >> k) & 1u;
        float ak, bk;
        step_map(has_e, k, Z, ak, bk);
        X = m ? bk + ak * X : X;
        Z = m ? vv[k] : Z;
        const float am = m ? X : 0.0f;
        a[k] = am;
        rt[k] = m ? X + vv[k] : 0.0f;
        s1 += am;
        s2 = fmaf(am, am, s2);
        if constexpr (decltype(has_e)::value) {
          if ((ew >> k) & 1u) {  // a reward position off the loss mask keeps the turn-level pair (:70-71)
            const bool only_e = !m && ((ek >> k) & 1u);
            a[k] = only_e ? hv[k] : a[k];
            rt[k] = only_e ? hv[k] + vv[k] : rt[k];
          }
        }
      }
    };
    if (any_e) rewalk3(std::true_type{});
    else rewalk3(std::false_type{});
    m0 += (double)__popc(mk);
    m1 += (double)s1;
    m2 += (double)s2;
    reinterpret_cast<float4*>(S.r)[2 * tid] = make_float4(a[0], a[1], a[2], a[3]);
    reinterpret_cast<float4*>(S.r)[2 * tid + 1] = make_float4(a[4], a[5], a[6], a[7]);
    reinterpret_cast<float4*>(S.v)[2 * tid] = make_float4(rt[0], rt[1], rt[2], rt[3]);
    reinterpret_cast<float4*>(S.v)[2 * tid + 1] = make_float4(rt[4], rt[5], rt[6], rt[7]);
    if (tid == 0) {
      s_carry[0] = Sz;
      s_carry[1] = Se;
      s_carry[2] = Sh;
      s_carry[3] = St;
    }
    fence_async_smem();
    __syncthreads();
    z_c = s_carry[0];
    e_c = s_carry[1];
    h_c = s_carry[2];
    t_c = s_carry[3];
    if (warp == 0) {
      bt_store(g, S, adv, ret, lane);
      if (lane == 0) {
        tma_store_wait_read();
        if (nxt.row < B) {
          bt_issue(nxt, T, S, &full[st], rewards, values, mask);
          nxt = bt_advance(nxt, B, T);
        }
      }
    }
  }
  bt_add_moments(moments, m0, m1, m2, s_mom, tid);
}

// ---- REINFORCE++ returns and ReMax as block scans ------------------------------------------------------------------------------
// MODE 0, compute_reinforce_plus_plus_outcome_advantage (core_algos.py:262-270): run = r_t + gamma * run; returns_t = run;
// run *= mask_t - the state that travels right to left is X -> mask_t * (r_t + gamma * X), an affine map, and returns_t =
// r_t + gamma * X_in.  MODE 1, compute_remax_outcome_advantage (:295-303): returns = reverse cumsum of r * mask (X -> X + r_t *
// mask_t, summed in double like the sequential kernel), advantages = returns - baseline * mask.  Same tiles, staging and scan
// skeleton as gae_scan_bulk_kernel; S.r carries returns (MODE 0) / advantages (MODE 1) out, S.v the returns of MODE 1.
template <int MODE>
__global__ void __launch_bounds__(BT_T) returns_scan_bulk_kernel(int B, int T, const float* __restrict__ rewards,
                                                               const uint8_t* __restrict__ mask, float gamma,
                                                               const float* __restrict__ baselines, float* __restrict__ out0,
                                                               float* __restrict__ out1, double* moments) {
  constexpr int NS = 2;
  using acc_t = typename std::conditional<MODE == 1, double, float>::type;
  extern __shared__ __align__(128) uint8_t bulk_smem[];
  BulkStage* stage = reinterpret_cast<BulkStage*>(bulk_smem);
  __shared__ __align__(8) uint64_t full[NS];
  __shared__ acc_t s_a[BT_NW], s_b[BT_NW];
  __shared__ acc_t s_carry;
  __shared__ double s_mom[BT_NW][3];
  const unsigned FULL = 0xffffffffu;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) {
    for (int s = 0; s < NS; ++s) mbar_init(&full[s], 1);
    fence_mbar_init();
  }
  __syncthreads();
  ScanTile cur = bt_first(B, T);
  ScanTile nxt = cur;
  if (tid == 0) {
    for (int s = 0; s < NS && nxt.row < B; ++s) {
      bt_issue(nxt, T, stage[s], &full[s], rewards, nullptr, mask);
      nxt = bt_advance(nxt, B, T);
    }
  }
  double m0 = 0.0, m1 = 0.0, m2 = 0.0;
  acc_t carry = 0;
  for (int it = 0; cur.row < B; ++it, cur = bt_advance(cur, B, T)) {
    const int st = it % NS;
    const TileGeom g = bt_geom(cur, T);
    if (g.right_end) carry = 0;
    BulkStage& S = stage[st];
    mbar_wait(&full[st], (uint32_t)((it / NS) & 1));
    const int w0 = BT_K * tid;
    const float4 r0 = reinterpret_cast<const float4*>(S.r)[2 * tid], r1 = reinterpret_cast<const float4*>(S.r)[2 * tid + 1];
    const float r_[BT_K] = {r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, r1.w};
    unsigned mk, inb;
    {
      const int lo_k = min(max(g.rel_lo - w0, 0), BT_K), hi_k = min(max(g.rel_hi - w0, 0), BT_K);
      inb = ((1u << hi_k) - 1u) & ~((1u << lo_k) - 1u);
      const uint32_t* mwp = reinterpret_cast<const uint32_t*>(S.m + g.mshift + w0);
      const uint32_t nlo = __vsetne4(mwp[0], 0u), nhi = __vsetne4(mwp[1], 0u);
      mk = ((((nlo * 0x00204081u) >> 21) & 0xfu) | ((((nhi * 0x00204081u) >> 21) & 0xfu) << 4)) & inb;
    }
    float rr[BT_K];
#pragma unroll
    for (int k = 0; k < BT_K; ++k) rr[k] = ((inb >> k) & 1u) ? r_[k] : 0.0f;  // words of other rows may hold anything
    // the eight maps of this thread composed: X -> Bq + Aq * X
    acc_t Aq = 1, Bq = 0;
#pragma unroll
    for (int k = BT_K - 1; k >= 0; --k) {
      const bool m = (mk >> k) & 1u;
      if (MODE == 0) {
        const float a = m ? gamma : 0.0f, b = m ? rr[k] : 0.0f;
        Bq = b + a * Bq;
        Aq = a * Aq;
      } else {
        Bq += m ? (acc_t)rr[k] : (acc_t)0;  // Aq stays 1
      }
    }
    // (off the row a = b = 0 (MODE 0) / b = 0 (MODE 1): right of the row's end the state is zero anyway, left of its start
    // nothing is used)
    acc_t PA = Aq, PB = Bq;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const acc_t b2 = __shfl_down_sync(FULL, PB, o);
      const bool on = lane + o < 32;
      if (MODE == 0) {
        const acc_t a2 = __shfl_down_sync(FULL, PA, o);
        PB = on ? PB + PA * b2 : PB;
        PA = on ? PA * a2 : PA;
      } else {
        PB = on ? PB + b2 : PB;
      }
    }
    if (lane == 0) {
      s_a[warp] = PA;
      s_b[warp] = PB;
    }
    __syncthreads();
    acc_t G = carry;
#pragma unroll
    for (int w = BT_NW - 1; w >= 1; --w) {
      const acc_t ng = MODE == 0 ? s_b[w] + s_a[w] * G : s_b[w] + G;
      G = w > warp ? ng : G;
    }
    const acc_t Sx = MODE == 0 ? PB + PA * G : PB + G;  // the state leaving this thread's first word
    acc_t In = __shfl_down_sync(FULL, Sx, 1);
    if (lane == 31) In = G;
    float o0[BT_K], o1[BT_K];
    float s1 = 0.0f, s2 = 0.0f;
    {
      acc_t X = In;
#pragma unroll
      for (int k = BT_K - 1; k >= 0; --k) {
        const bool m = (mk >> k) & 1u;
        if (MODE == 0) {
          const float R = rr[k] + gamma * (float)X;  // returns_t
          o0[k] = R;
          o1[k] = 0.0f;
          X = m ? R : 0.0f;  // run *= mask_t
          const float Rm = m ? R : 0.0f;
          s1 += Rm;
          s2 = fmaf(Rm, Rm, s2);
        } else {
          X += m ? (acc_t)rr[k] : (acc_t)0;
          const float R = (float)X;
          o1[k] = R;
          o0[k] = R - (m ? baselines[cur.row] : 0.0f);
        }
      }
    }
    if (MODE == 0) {
      m0 += (double)__popc(mk);
      m1 += (double)s1;
      m2 += (double)s2;
    }
    reinterpret_cast<float4*>(S.r)[2 * tid] = make_float4(o0[0], o0[1], o0[2], o0[3]);
    reinterpret_cast<float4*>(S.r)[2 * tid + 1] = make_float4(o0[4], o0[5], o0[6], o0[7]);
    if (MODE == 1) {
      reinterpret_cast<float4*>(S.v)[2 * tid] = make_float4(o1[0], o1[1], o1[2], o1[3]);
      reinterpret_cast<float4*>(S.v)[2 * tid + 1] = make_float4(o1[4], o1[5], o1[6], o1[7]);
    }
    if (tid == 0) s_carry = Sx;
    fence_async_smem();
    __syncthreads();
    carry = s_carry;
    if (warp == 0) {
      bt_store(g, S, out0, MODE == 1 ? out1 : nullptr, lane);
      if (lane == 0) {
        tma_store_wait_read();
        if (nxt.row < B) {
          bt_issue(nxt, T, S, &full[st], rewards, nullptr, mask);
          nxt = bt_advance(nxt, B, T);
        }
      }
    }
  }
  if (MODE == 0) bt_add_moments(moments, m0, m1, m2, s_mom, tid);
}

// below this many rows the per-row kernels finish in a wave or two and win (RAGEN_TS_MIN_ROWS overrides, for probing)
static int ts_min_rows() {
  static const int v = [] {
    const char* e = getenv("RAGEN_TS_MIN_ROWS");
    return (e && atoi(e) > 0) ? atoi(e) : 4096;
  }();
  return v;
}

// 8 rows per CTA measured best at every B tried, with two and with three warps per group (three warps, 8 Ki x 2047: GAE 96 us vs
// 118 / 172 us with 16 / 32 rows; 64 Ki x 1023: 317 vs 389 / 353 us): a row group's step is a chain of dependent instructions
// issued by lone warps (~500 in the scan warp per 32-token tile), so more and smaller groups overlap better than fewer, fatter
// ones.  RAGEN_TS_RPW = 16 | 32 overrides, for probing.
constexpr int TS_RPW = 8;
template <int MODE>
void launch_tile_scan(const ragen_ctx* ctx, cudaStream_t st, int B, int T, const float* rewards, const float* values, const uint8_t* mask,
                      float gamma, float gl, const float* baselines, float* adv, float* ret, double* moments, float hgamma = 0.0f,
                      float hgl = 0.0f) {
  (void)ctx;
  static const int force = [] {
    const char* e = getenv("RAGEN_TS_RPW");
    return e ? atoi(e) : 0;
  }();
  const int rpw = force ? force : TS_RPW;
  if (rpw == 32)
    tile_scan_kernel<MODE, 32><<<(B + 31) / 32, 96, TS_STAGES * sizeof(TsStage<32>), st>>>(B, T, rewards, values, mask, gamma, gl, hgamma, hgl,
                                                                                          baselines, adv, ret, moments);
  else if (rpw == 16)
    tile_scan_kernel<MODE, 16><<<(B + 15) / 16, 96, TS_STAGES * sizeof(TsStage<16>), st>>>(B, T, rewards, values, mask, gamma, gl, hgamma, hgl,
                                                                                          baselines, adv, ret, moments);
  else
    tile_scan_kernel<MODE, 8><<<(B + 7) / 8, 96, TS_STAGES * sizeof(TsStage<8>), st>>>(B, T, rewards, values, mask, gamma, gl, hgamma, hgl,
                                                                                      baselines, adv, ret, moments);
}

}  // namespace

#define RAGEN_BT(B, T) RAGEN_REQUIRE(B >= 0 && (B == 0 || T >= 1) && (int64_t)B * T < ((int64_t)1 << 40), "bad shape B=%d T=%d", B, T)

extern "C" int ragen_gae(ragen_ctx* ctx, int B, int T, const float* rewards, const float* values, const uint8_t* mask,
                         double gamma, double lam, float* adv, float* ret, double* moments, void* stream) {
  RAGEN_REQUIRE(ctx, "ctx is NULL");
  RAGEN_BT(B, T);
  RAGEN_REQUIRE(B == 0 || (rewards && values && mask && adv && ret), "NULL pointer argument");
  DeviceGuard guard(ctx->device);
  cudaStream_t st = (cudaStream_t)stream;
  if (moments) RAGEN_CUDA(cudaMemsetAsync(moments, 0, 3 * sizeof(double), st));
  if (B == 0) return RAGEN_OK;  // an empty shard still contributes zero moments to the all-reduce
  if (B >= ts_min_rows())
    launch_tile_scan<TS_GAE>(ctx, st, B, T, rewards, values, mask, (float)gamma, (float)(gamma * lam), nullptr, adv, ret, moments);
  else
    gae_kernel<<<B, RT, 0, st>>>(B, T, rewards, values, mask, (float)gamma, (float)(gamma * lam), adv, ret, moments);
  RAGEN_LAUNCHED(ctx);
  return RAGEN_OK;
}

static int scan_knob(const char* name, int dflt) {  // probing knobs of the scan path (environment, read per call: cheap next to a launch)
  const char* e = getenv(name);
  return e ? atoi(e) : dflt;
}

extern "C" int ragen_gae_scan(ragen_ctx* ctx, int B, int T, const float* rewards, const float* values, const uint8_t* mask,
                              double gamma, double lam, float* adv, float* ret, double* moments, void* stream) {
  RAGEN_REQUIRE(ctx, "ctx is NULL");
  RAGEN_BT(B, T);
  RAGEN_REQUIRE(B == 0 || (rewards && values && mask && adv && ret), "NULL pointer argument");
  DeviceGuard guard(ctx->device);
  cudaStream_t st = (cudaStream_t)stream;
  if (moments) RAGEN_CUDA(cudaMemsetAsync(moments, 0, 3 * sizeof(double), st));
  if (B == 0) return RAGEN_OK;
  const float gf = (float)gamma, cf = (float)(gamma * lam);
  const bool aligned = (((uintptr_t)rewards | (uintptr_t)values | (uintptr_t)mask | (uintptr_t)adv | (uintptr_t)ret) & 15) == 0;
  if (aligned && ((int64_t)B * T) % 16 == 0 && scan_knob("RAGEN_SCAN_BULK", 1)) {
    const int ns = scan_knob("RAGEN_SCAN_STAGES", 2), per_sm = scan_knob("RAGEN_SCAN_CTAS", 3);
    const int grid = B < ctx->sm_count * per_sm ? B : ctx->sm_count * per_sm;
    const size_t smem = (size_t)ns * sizeof(BulkStage);
    auto kern = ns == 3 ? gae_scan_bulk_kernel<3> : gae_scan_bulk_kernel<2>;
    RAGEN_REQUIRE(ns == 2 || ns == 3, "RAGEN_SCAN_STAGES is 2 or 3");
    RAGEN_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kern<<<grid, BT_T, smem, st>>>(B, T, rewards, values, mask, gf, cf, adv, ret, moments);
    RAGEN_LAUNCHED(ctx);
    return RAGEN_OK;
  }
  const int cap = ctx->sm_count * 4;  // persistent CTAs: four of 256 threads fit an SM (64 registers per thread, 2 x 18 KB of stages)
  static bool carved[64] = {};
  if (ctx->device >= 0 && ctx->device < 64 && !carved[ctx->device]) {  // 4 x 37 KB per SM: ask for the large shared-memory split once
    RAGEN_CUDA(cudaFuncSetAttribute(gae_scan_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
    carved[ctx->device] = true;
  }
  gae_scan_kernel<<<B < cap ? B : cap, BS_T, 0, st>>>(B, T, rewards, values, mask, (float)gamma, (float)(gamma * lam), adv, ret, moments);
  RAGEN_LAUNCHED(ctx);
  return RAGEN_OK;
}

extern "C" int ragen_bilevel_gae(ragen_ctx* ctx, int B, int T, const float* rewards, const float* values,
                                 const uint8_t* loss_mask, double gamma, double lam, double high_level_gamma, float* adv,
                                 float* ret, double* moments, void* stream) {
  RAGEN_REQUIRE(ctx, "ctx is NULL");
  RAGEN_BT(B, T);
  RAGEN_REQUIRE(B == 0 || (rewards && values && loss_mask && adv && ret), "NULL pointer argument");
  DeviceGuard guard(ctx->device);
  cudaStream_t st = (cudaStream_t)stream;
  if (moments) RAGEN_CUDA(cudaMemsetAsync(moments, 0, 3 * sizeof(double), st));
  if (B == 0) return RAGEN_OK;  // an empty shard still contributes zero moments to the all-reduce
  if (B >= ts_min_rows())
    launch_tile_scan<TS_BILEVEL>(ctx, st, B, T, rewards, values, loss_mask, (float)gamma, (float)(gamma * lam), nullptr, adv, ret, moments,
                                 (float)high_level_gamma, (float)(high_level_gamma * lam));
  else
    bilevel_gae_kernel<<<B, RT, 0, st>>>(B, T, rewards, values, loss_mask, (float)gamma, (float)(gamma * lam), (float)high_level_gamma,
                                         (float)(high_level_gamma * lam), adv, ret, moments);
  RAGEN_LAUNCHED(ctx);
  return RAGEN_OK;
}

// The bi-level estimator as block scans (bilevel_scan_bulk_kernel): same arguments and outputs as ragen_bilevel_gae, equal to it up
// to fp32 rounding (another summation order).  Tensors the bulk copies cannot take (a base pointer off a 16-byte boundary,
// B * T % 16 != 0) go to ragen_bilevel_gae.
extern "C" int ragen_bilevel_gae_scan(ragen_ctx* ctx, int B, int T, const float* rewards, const float* values,
                                      const uint8_t* loss_mask, double gamma, double lam, double high_level_gamma, float* adv,
                                      float* ret, double* moments, void* stream) {
  RAGEN_REQUIRE(ctx, "ctx is NULL");
  RAGEN_BT(B, T);
  RAGEN_REQUIRE(B == 0 || (rewards && values && loss_mask && adv && ret), "NULL pointer argument");
  const bool aligned = (((uintptr_t)rewards | (uintptr_t)values | (uintptr_t)loss_mask | (uintptr_t)adv | (uintptr_t)ret) & 15) == 0;
  if (B == 0 || !aligned || ((int64_t)B * T) % 16 != 0)
    return ragen_bilevel_gae(ctx, B, T, rewards, values, loss_mask, gamma, lam, high_level_gamma, adv, ret, moments, stream);
  DeviceGuard guard(ctx->device);
  cudaStream_t st = (cudaStream_t)stream;
  if (moments) RAGEN_CUDA(cudaMemsetAsync(moments, 0, 3 * sizeof(double), st));
  const int ns = scan_knob("RAGEN_SCAN_STAGES", 2), per_sm = scan_knob("RAGEN_BISCAN_CTAS", 3);
  RAGEN_REQUIRE(ns == 2 || ns == 3, "RAGEN_SCAN_STAGES is 2 or 3");
  const int grid = B < ctx->sm_count * per_sm ? B : ctx->sm_count * per_sm;
  const size_t smem = (size_t)ns * sizeof(BulkStage);
  auto kern = per_sm >= 4 ? bilevel_scan_bulk_kernel<2, 4> : per_sm == 2 ? bilevel_scan_bulk_kernel<2, 2> : ns == 3 ? bilevel_scan_bulk_kernel<3, 3> : bilevel_scan_bulk_kernel<2, 3>;
  RAGEN_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  kern<<<grid, BT_T, smem, st>>>(B, T, rewards, values, loss_mask, (float)gamma, (float)(gamma * lam), (float)high_level_gamma,
                                 (float)(high_level_gamma * lam), adv, ret, moments);
  RAGEN_LAUNCHED(ctx);
  return RAGEN_OK;
}

extern "C" int ragen_masked_moments(ragen_ctx* ctx, int64_t n, const float* x, const uint8_t* mask, double* moments,
                                    void* stream) {
  RAGEN_REQUIRE(ctx, "ctx is NULL");
  RAGEN_REQUIRE(n >= 0 && moments && (n == 0 || (x && mask)), "bad argument");
  DeviceGuard guard(ctx->device);
  cudaStream_t st = (cudaStream_t)stream;
  RAGEN_CUDA(cudaMemsetAsync(moments, 0, 3 * sizeof(double), st));
  if (n == 0) return RAGEN_OK;
  const int threads = 256;
  const int64_t want = (n + threads - 1) / threads;
  const int blocks = (int)(want < (int64_t)ctx->sm_count * 8 ? want : (int64_t)ctx->sm_count * 8);
  moments_kernel<<<blocks, threads, 0, st>>>(n, x, mask, moments);
  RAGEN_LAUNCHED(ctx);
  return RAGEN_OK;
}

// moments live on the device (they may just have come out of an all-reduce), so the two ValueErrors of masked_var (sum of the mask
// 0 or 1, torch_functional.py:128-136) cannot be raised here without a synchronisation: that check is the caller's
// (ops.whiten_(check=True) reads the 24 bytes back and raises them); the kernel itself writes NaN for such moments.
extern "C" int ragen_whiten(ragen_ctx* ctx, int64_t n, float* x, const double* moments, void* stream) {
  RAGEN_REQUIRE(ctx, "ctx is NULL");
  RAGEN_REQUIRE(n >= 0 && moments && (n == 0 || x), "bad argument");
  if (n == 0) return RAGEN_OK;
  DeviceGuard guard(ctx->device);
  const int threads = 256;
  const int64_t want = (n + threads - 1) / threads;
  const int blocks = (int)(want < (int64_t)ctx->sm_count * 16 ? want : (int64_t)ctx->sm_count * 16);
  whiten_kernel<<<blocks, threads, 0, (cudaStream_t)stream>>>(n, x, moments);
  RAGEN_LAUNCHED(ctx);
  return RAGEN_OK;
}

extern "C" int ragen_grpo_outcome(ragen_ctx* ctx, int B, int T, const float* rewards, const uint8_t* mask,
                                  const int32_t* group, int n_groups, double epsilon, int norm_by_std, float* scores,
                                  float* adv, void* stream) {
  RAGEN_REQUIRE(ctx, "ctx is NULL");
  RAGEN_BT(B, T);
  RAGEN_REQUIRE(rewards && mask && group && scores && adv && n_groups >= 1, "bad argument");
  if (B == 0) return RAGEN_OK;
  DeviceGuard guard(ctx->device);
  cudaStream_t st = (cudaStream_t)stream;
  const size_t bytes = (size_t)n_groups * 3 * sizeof(double);
  double* gstats = (double*)ragen_scratch(ctx, bytes);
  if (!gstats) return RAGEN_ENOMEM;
  RAGEN_CUDA(cudaMemsetAsync(gstats, 0, bytes, st));
  launch_row_sums(st, B, T, rewards, group, n_groups, scores, gstats);
  RAGEN_LAUNCHED(ctx);
  launch_outcome_apply(st, B, T, scores, group, n_groups, gstats, mask, (float)epsilon, norm_by_std ? 1 : 0, adv, nullptr);
  RAGEN_LAUNCHED(ctx);
  return RAGEN_OK;
}

extern "C" int ragen_outcome_advantage(ragen_ctx* ctx, int B, int T, const float* rewards, const uint8_t* mask,
                                       const int32_t* group, int n_groups, int mode, double epsilon, float* scores,
                                       float* adv, double* moments, void* stream) {
  RAGEN_REQUIRE(ctx, "ctx is NULL");
  RAGEN_BT(B, T);
  RAGEN_REQUIRE(mode >= 0 && mode <= 3 && (B == 0 || (rewards && mask && group && scores && adv && n_groups >= 1)), "bad argument");
  DeviceGuard guard(ctx->device);
  cudaStream_t st = (cudaStream_t)stream;
  if (moments) RAGEN_CUDA(cudaMemsetAsync(moments, 0, 3 * sizeof(double), st));
  if (B == 0) return RAGEN_OK;
  const size_t bytes = (size_t)n_groups * 3 * sizeof(double);
  double* gstats = (double*)ragen_scratch(ctx, bytes);
  if (!gstats) return RAGEN_ENOMEM;
  RAGEN_CUDA(cudaMemsetAsync(gstats, 0, bytes, st));
  launch_row_sums(st, B, T, rewards, group, n_groups, scores, gstats);
  RAGEN_LAUNCHED(ctx);
  launch_outcome_apply(st, B, T, scores, group, n_groups, gstats, mask, (float)epsilon, mode, adv, moments);
  RAGEN_LAUNCHED(ctx);
  return RAGEN_OK;
}

extern "C" int ragen_discounted_returns(ragen_ctx* ctx, int B, int T, const float* rewards, const uint8_t* mask, double gamma,
                                        float* ret, double* moments, void* stream) {
  RAGEN_REQUIRE(ctx, "ctx is NULL");
  RAGEN_BT(B, T);
  RAGEN_REQUIRE(moments && (B == 0 || (rewards && mask && ret)), "NULL pointer argument");
  DeviceGuard guard(ctx->device);
  cudaStream_t st = (cudaStream_t)stream;
  RAGEN_CUDA(cudaMemsetAsync(moments, 0, 3 * sizeof(double), st));
  if (B == 0) return RAGEN_OK;
  if (B >= ts_min_rows())
    launch_tile_scan<TS_RFPP>(ctx, st, B, T, rewards, nullptr, mask, (float)gamma, 0.0f, nullptr, nullptr, ret, moments);
  else
    returns_kernel<0><<<B, RT, 0, st>>>(B, T, rewards, mask, (float)gamma, nullptr, nullptr, ret, moments);
  RAGEN_LAUNCHED(ctx);
  return RAGEN_OK;
}

extern "C" int ragen_remax(ragen_ctx* ctx, int B, int T, const float* rewards, const uint8_t* mask, const float* baselines,
                           float* adv, float* ret, void* stream) {
  RAGEN_REQUIRE(ctx, "ctx is NULL");
  RAGEN_BT(B, T);
  RAGEN_REQUIRE(rewards && mask && baselines && adv && ret, "NULL pointer argument");
  if (B == 0) return RAGEN_OK;
  DeviceGuard guard(ctx->device);
  if (B >= ts_min_rows())
    launch_tile_scan<TS_REMAX>(ctx, (cudaStream_t)stream, B, T, rewards, nullptr, mask, 0.0f, 0.0f, baselines, adv, ret, nullptr);
  else
    returns_kernel<1><<<B, RT, 0, (cudaStream_t)stream>>>(B, T, rewards, mask, 0.0f, baselines, adv, ret, nullptr);
  RAGEN_LAUNCHED(ctx);
  return RAGEN_OK;
}

// The two estimators above as block scans over bulk-staged tiles (returns_scan_bulk_kernel): same arguments and outputs, equal up
// to fp32 rounding (another summation order; ReMax still sums in double).  Tensors the bulk copies cannot take go to the
// sequential entry points.
extern "C" int ragen_discounted_returns_scan(ragen_ctx* ctx, int B, int T, const float* rewards, const uint8_t* mask, double gamma,
                                             float* ret, double* moments, void* stream) {
  RAGEN_REQUIRE(ctx, "ctx is NULL");
  RAGEN_BT(B, T);
  RAGEN_REQUIRE(moments && (B == 0 || (rewards && mask && ret)), "NULL pointer argument");
  const bool aligned = (((uintptr_t)rewards | (uintptr_t)mask | (uintptr_t)ret) & 15) == 0;
  if (B == 0 || !aligned || ((int64_t)B * T) % 16 != 0) return ragen_discounted_returns(ctx, B, T, rewards, mask, gamma, ret, moments, stream);
  DeviceGuard guard(ctx->device);
  cudaStream_t st = (cudaStream_t)stream;
  RAGEN_CUDA(cudaMemsetAsync(moments, 0, 3 * sizeof(double), st));
  const int per_sm = scan_knob("RAGEN_RSCAN_CTAS", 4);  // 56 registers: four CTAs of 288 threads per SM
  const int grid = B < ctx->sm_count * per_sm ? B : ctx->sm_count * per_sm;
  const size_t smem = 2 * sizeof(BulkStage);
  RAGEN_CUDA(cudaFuncSetAttribute(returns_scan_bulk_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  returns_scan_bulk_kernel<0><<<grid, BT_T, smem, st>>>(B, T, rewards, mask, (float)gamma, nullptr, ret, nullptr, moments);
  RAGEN_LAUNCHED(ctx);
  return RAGEN_OK;
}

extern "C" int ragen_remax_scan(ragen_ctx* ctx, int B, int T, const float* rewards, const uint8_t* mask, const float* baselines,
                                float* adv, float* ret, void* stream) {
  RAGEN_REQUIRE(ctx, "ctx is NULL");
  RAGEN_BT(B, T);
  RAGEN_REQUIRE(rewards && mask && baselines && adv && ret, "NULL pointer argument");
  const bool aligned = (((uintptr_t)rewards | (uintptr_t)mask | (uintptr_t)adv | (uintptr_t)ret) & 15) == 0;
  if (B == 0 || !aligned || ((int64_t)B * T) % 16 != 0) return ragen_remax(ctx, B, T, rewards, mask, baselines, adv, ret, stream);
  DeviceGuard guard(ctx->device);
  const int per_sm = scan_knob("RAGEN_RSCAN_CTAS", 4);  // 56 registers: four CTAs of 288 threads per SM
  const int grid = B < ctx->sm_count * per_sm ? B : ctx->sm_count * per_sm;
  const size_t smem = 2 * sizeof(BulkStage);
  RAGEN_CUDA(cudaFuncSetAttribute(returns_scan_bulk_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  returns_scan_bulk_kernel<1><<<grid, BT_T, smem, (cudaStream_t)stream>>>(B, T, rewards, mask, 0.0f, baselines, adv, ret, nullptr);
  RAGEN_LAUNCHED(ctx);
  return RAGEN_OK;
}

extern "C" int ragen_whiten_masked(ragen_ctx* ctx, int64_t n, const float* x, const uint8_t* mask, const double* moments,
                                   float* out, void* stream) {
  RAGEN_REQUIRE(ctx, "ctx is NULL");
  RAGEN_REQUIRE(n >= 0 && x && mask && moments && out, "bad argument");
  if (n == 0) return RAGEN_OK;
  DeviceGuard guard(ctx->device);
  const int threads = 256;
  const int64_t want = (n + threads - 1) / threads;
  const int blocks = (int)(want < (int64_t)ctx->sm_count * 16 ? want : (int64_t)ctx->sm_count * 16);
  whiten_masked_kernel<<<blocks, threads, 0, (cudaStream_t)stream>>>(n, x, mask, moments, out);
  RAGEN_LAUNCHED(ctx);
  return RAGEN_OK;
}

extern "C" int ragen_kl_rewards(ragen_ctx* ctx, int B, int T, const float* scores, const float* logp, const float* ref_logp,
                                const uint8_t* mask, int mode, double beta, float* rewards_out, float* row_kl, void* stream) {
  RAGEN_REQUIRE(ctx, "ctx is NULL");
  RAGEN_BT(B, T);
  RAGEN_REQUIRE(scores && logp && ref_logp && mask && rewards_out && mode >= 0 && mode <= 3, "bad argument");
  if (B == 0) return RAGEN_OK;
  DeviceGuard guard(ctx->device);
  kl_rewards_kernel<<<B, RT, 0, (cudaStream_t)stream>>>(B, T, scores, logp, ref_logp, mask, mode,
                                                                                 (float)beta, rewards_out, row_kl);
  RAGEN_LAUNCHED(ctx);
  return RAGEN_OK;
}

extern "C" int ragen_group_filter_stats(ragen_ctx* ctx, int n_groups, int group_size, int T, const float* scores,
                                        float* row_sum, float* gstd, float* gmax, float* gmean, void* stream) {
  RAGEN_REQUIRE(ctx, "ctx is NULL");
  RAGEN_REQUIRE(n_groups >= 0 && group_size >= 1 && T >= 1, "bad shape");
  RAGEN_REQUIRE(scores && row_sum && gstd && gmax && gmean, "NULL pointer argument");
  if (n_groups == 0) return RAGEN_OK;
  DeviceGuard guard(ctx->device);
  cudaStream_t st = (cudaStream_t)stream;
  launch_row_sums(st, n_groups * group_size, T, scores, nullptr, 0, row_sum, nullptr);
  RAGEN_LAUNCHED(ctx);
  group_stats_kernel<<<(n_groups + 127) / 128, 128, 0, st>>>(n_groups, group_size, row_sum, gstd, gmax, gmean);
  RAGEN_LAUNCHED(ctx);
  return RAGEN_OK;
}
