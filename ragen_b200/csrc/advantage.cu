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
#include "common.cuh"

namespace {

constexpr int ROWS_PER_CTA = 4;   // warps per CTA
constexpr int CHUNK = 1024;       // tokens staged per warp per iteration

__device__ __forceinline__ void add_moments(double* moments, double s0, double s1, double s2) {
  s0 = warp_sum(s0);
  s1 = warp_sum(s1);
  s2 = warp_sum(s2);
  if ((threadIdx.x & 31) == 0 && moments) {
    atomicAdd(moments + 0, s0);
    atomicAdd(moments + 1, s1);
    atomicAdd(moments + 2, s2);
  }
}

// A_t = (r_t + gamma V_{t+1} - V_t) + (gamma lam) A_{t+1}, V_T = 0, over ALL columns (core_algos.py:98-105);
// returns = A + V (:107).  gamma and gamma*lam arrive already rounded to fp32 the way torch rounds a Python scalar
// that multiplies an fp32 tensor.
__global__ void __launch_bounds__(ROWS_PER_CTA * 32) gae_kernel(int B, int T, const float* __restrict__ rewards,
                                                                 const float* __restrict__ values,
                                                                 const uint8_t* __restrict__ mask, float gamma, float gl,
                                                                 float* __restrict__ adv, float* __restrict__ ret,
                                                                 double* moments) {
  __shared__ float s_a[ROWS_PER_CTA][CHUNK];
  __shared__ float s_v[ROWS_PER_CTA][CHUNK];
  __shared__ uint8_t s_m[ROWS_PER_CTA][CHUNK];
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int row = blockIdx.x * ROWS_PER_CTA + w;
  double m0 = 0.0, m1 = 0.0, m2 = 0.0;
  if (row < B) {
    const size_t base = (size_t)row * T;
    float carry = 0.0f;
    for (int t1 = T; t1 > 0; t1 -= CHUNK) {
      const int t0 = max(0, t1 - CHUNK), len = t1 - t0;
      // the unrolled loop keeps 8 x 4 independent loads in flight per lane: at C5 (128 rows) the pass is bound by
      // memory latency, not bandwidth
#pragma unroll 8
      for (int i = lane; i < len; i += 32) {
        const int t = t0 + i;
        const float v = __ldg(values + base + t);
        const float vn = (t + 1 < T) ? __ldg(values + base + t + 1) : 0.0f;
        const float r = __ldg(rewards + base + t);
        s_m[w][i] = mask[base + t];
        s_a[w][i] = __fsub_rn(__fadd_rn(r, __fmul_rn(gamma, vn)), v);
        s_v[w][i] = v;
      }
      __syncwarp();
      if (lane == 0) {
        float lg = carry;
#pragma unroll 8
        for (int i = len - 1; i >= 0; --i) {
          lg = __fadd_rn(s_a[w][i], __fmul_rn(gl, lg));
          s_a[w][i] = lg;
        }
        carry = lg;
      }
      __syncwarp();
#pragma unroll 8
      for (int i = lane; i < len; i += 32) {
        const int t = t0 + i;
        const float a = s_a[w][i];
        adv[base + t] = a;
        ret[base + t] = __fadd_rn(a, s_v[w][i]);
        if (s_m[w][i]) {
          m0 += 1.0;
          m1 += (double)a;
          m2 += (double)a * (double)a;
        }
      }
      __syncwarp();
    }
  }
  add_moments(moments, m0, m1, m2);
}

// ragen/trainer/core_algos.py:44-88 for one row, as ONE backward sweep: the turn-level estimate at a reward
// position depends only on later reward positions, and the token-level pass at that position only needs the
// turn-level result of the same position, so both passes advance together.  fp32 operation order as the reference.
__global__ void __launch_bounds__(ROWS_PER_CTA * 32) bilevel_gae_kernel(int B, int T, const float* __restrict__ rewards,
                                                                         const float* __restrict__ values,
                                                                         const uint8_t* __restrict__ mask, float gamma,
                                                                         float gl, float hgamma, float hgl,
                                                                         float* __restrict__ adv, float* __restrict__ ret,
                                                                         double* moments) {
  __shared__ float s_r[ROWS_PER_CTA][CHUNK];  // rewards in, advantages out
  __shared__ float s_v[ROWS_PER_CTA][CHUNK];  // values in, returns out
  __shared__ uint8_t s_m[ROWS_PER_CTA][CHUNK];
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int row = blockIdx.x * ROWS_PER_CTA + w;
  double m0 = 0.0, m1 = 0.0, m2 = 0.0;
  if (row < B) {
    const size_t base = (size_t)row * T;
    // sweep state (lane 0 only)
    bool hi_has_next = false, lo_has_next = false;
    float hi_next_v = 0.0f, lo_next_v = 0.0f, hi_lg = 0.0f, lo_lg = 0.0f;
    for (int t1 = T; t1 > 0; t1 -= CHUNK) {
      const int t0 = max(0, t1 - CHUNK), len = t1 - t0;
#pragma unroll 8
      for (int i = lane; i < len; i += 32) {
        s_r[w][i] = __ldg(rewards + base + t0 + i);
        s_v[w][i] = __ldg(values + base + t0 + i);
        s_m[w][i] = mask[base + t0 + i];
      }
      __syncwarp();
      if (lane == 0) {
        for (int i = len - 1; i >= 0; --i) {
          const float r = s_r[w][i], v = s_v[w][i];
          const bool is_eos = (r != 0.0f);  // reward_mask = token_level_rewards.bool()
          float a_out = 0.0f, r_out = 0.0f, upd = r;
          if (is_eos) {
            const float nv = hi_has_next ? hi_next_v : 0.0f;
            const float delta = __fsub_rn(__fadd_rn(r, __fmul_rn(hgamma, nv)), v);
            hi_lg = __fadd_rn(delta, __fmul_rn(hgl, hi_lg));
            hi_next_v = v;
            hi_has_next = true;
            a_out = hi_lg;
            upd = __fadd_rn(hi_lg, v);  // returns[pos] = updated_reward[pos] = advantages[pos] + values[pos]
            r_out = upd;
          }
          if (s_m[w][i]) {
            float nv;
            if (is_eos) {
              nv = 0.0f;
              lo_lg = 0.0f;
            } else {
              // the reference indexes valid_positions[i + 1] here and raises IndexError when the last loss-mask
              // position carries no reward; that input is treated as "next value 0"
              nv = lo_has_next ? lo_next_v : 0.0f;
            }
            const float delta = __fsub_rn(__fadd_rn(upd, __fmul_rn(gamma, nv)), v);
            lo_lg = __fadd_rn(delta, __fmul_rn(gl, lo_lg));
            lo_next_v = v;
            lo_has_next = true;
            a_out = lo_lg;
            r_out = __fadd_rn(lo_lg, v);
          }
          s_r[w][i] = a_out;
          s_v[w][i] = r_out;
        }
      }
      __syncwarp();
#pragma unroll 8
      for (int i = lane; i < len; i += 32) {
        const float a = s_r[w][i];
        adv[base + t0 + i] = a;
        ret[base + t0 + i] = s_v[w][i];
        if (s_m[w][i]) {
          m0 += 1.0;
          m1 += (double)a;
          m2 += (double)a * (double)a;
        }
      }
      __syncwarp();
    }
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
  const float scale = (float)(1.0 / sqrt(var + 1e-8));
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) x[i] = (x[i] - meanf) * scale;
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

// GRPO step 1: scores[b] = sum_t rewards[b][t]; per-group {count, sum, sum of squares} (core_algos.py:139-149)
__global__ void __launch_bounds__(ROWS_PER_CTA * 32) grpo_scores_kernel(int B, int T, const float* __restrict__ rewards,
                                                                         const int32_t* __restrict__ group, int n_groups,
                                                                         float* __restrict__ scores, double* gstats) {
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int row = blockIdx.x * ROWS_PER_CTA + w;
  if (row >= B) return;
  const size_t base = (size_t)row * T;
  double s = 0.0;
  for (int t = lane; t < T; t += 32) s += (double)rewards[base + t];
  s = warp_sum(s);
  if (lane == 0) {
    const float sf = (float)s;
    scores[row] = sf;
    const int g = group[row];
    if (g >= 0 && g < n_groups) {
      atomicAdd(gstats + 3 * g + 0, 1.0);
      atomicAdd(gstats + 3 * g + 1, (double)sf);
      atomicAdd(gstats + 3 * g + 2, (double)sf * (double)sf);
    }
  }
}

// GRPO step 2: (score - mean) / (std + eps) (unbiased std; groups of one: mean 0, std 1), broadcast over the
// row and multiplied by the mask (core_algos.py:150-165).
__global__ void __launch_bounds__(ROWS_PER_CTA * 32) grpo_apply_kernel(int B, int T, const float* __restrict__ scores,
                                                                        const int32_t* __restrict__ group, int n_groups,
                                                                        const double* __restrict__ gstats,
                                                                        const uint8_t* __restrict__ mask, float eps,
                                                                        int norm_by_std, float* __restrict__ adv) {
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int row = blockIdx.x * ROWS_PER_CTA + w;
  if (row >= B) return;
  const int g = group[row];
  float mean = 0.0f, sd = 1.0f;
  if (g >= 0 && g < n_groups) {
    const double c = gstats[3 * g], s1 = gstats[3 * g + 1], s2 = gstats[3 * g + 2];
    if (c > 1.5) {
      const double mu = s1 / c;
      const double var = fmax((s2 - c * mu * mu) / (c - 1.0), 0.0);
      mean = (float)mu;
      sd = (float)sqrt(var);
    }
  }
  const float sc = scores[row];
  const float val = norm_by_std ? __fdiv_rn(__fsub_rn(sc, mean), __fadd_rn(sd, eps)) : __fsub_rn(sc, mean);
  const size_t base = (size_t)row * T;
  for (int t = lane; t < T; t += 32) adv[base + t] = mask[base + t] ? val : __fmul_rn(val, 0.0f);
}

// ---- the remaining members of compute_advantage's dispatch table (agent_trainer.py:106-134) ---------------------------
// Outcome estimators share GRPO's first step (row scores + per-group count / sum / sum of squares); this is their second
// step.  mode 0: score - mean (GRPO without std); 1: GRPO; 2: RLOO, score*n/(n-1) - mean*n/(n-1) for groups of n > 1
// (core_algos.py:205-242); 3: REINFORCE++-baseline, score - mean, to be whitened by the caller from `moments` (:168-202).
// Groups of one use mean 0 (and std 1).  The value is broadcast over the row and multiplied by the mask.
__global__ void __launch_bounds__(ROWS_PER_CTA * 32) outcome_apply_kernel(int B, int T, const float* __restrict__ scores,
                                                                           const int32_t* __restrict__ group, int n_groups,
                                                                           const double* __restrict__ gstats,
                                                                           const uint8_t* __restrict__ mask, float eps, int mode,
                                                                           float* __restrict__ adv, double* moments) {
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int row = blockIdx.x * ROWS_PER_CTA + w;
  double m0 = 0.0, m1 = 0.0, m2 = 0.0;
  if (row < B) {
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
#pragma unroll 4
    for (int t = lane; t < T; t += 32) {
      const bool m = mask[base + t] != 0;
      adv[base + t] = m ? val : off;
      on += m;
    }
    m0 = (double)on;
    m1 = (double)on * (double)val;
    m2 = (double)on * (double)val * (double)val;
  }
  if (moments) add_moments(moments, m0, m1, m2);
}

// REINFORCE++ (core_algos.py:245-278): running = r_t + gamma * running, returns_t = running, running *= mask_t, right to
// left, in the reference's fp32 order; plus the masked moments of the returns for the whitening that follows.
// REMAX = 1 (core_algos.py:281-305): returns = reverse cumulative sum of rewards*mask (torch's CPU cumsum accumulates
// fp32 data in double and rounds every prefix), adv = returns - baseline * mask.
template <int REMAX>
__global__ void __launch_bounds__(ROWS_PER_CTA * 32) returns_kernel(int B, int T, const float* __restrict__ rewards,
                                                                     const uint8_t* __restrict__ mask, float gamma,
                                                                     const float* __restrict__ baselines,
                                                                     float* __restrict__ adv, float* __restrict__ ret,
                                                                     double* moments) {
  __shared__ float s_r[ROWS_PER_CTA][CHUNK];
  __shared__ uint8_t s_m[ROWS_PER_CTA][CHUNK];
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int row = blockIdx.x * ROWS_PER_CTA + w;
  double m0 = 0.0, m1 = 0.0, m2 = 0.0;
  if (row < B) {
    const size_t base = (size_t)row * T;
    const float bl = REMAX ? baselines[row] : 0.0f;
    float run = 0.0f;
    double acc = 0.0;
    for (int t1 = T; t1 > 0; t1 -= CHUNK) {
      const int t0 = max(0, t1 - CHUNK), len = t1 - t0;
#pragma unroll 8
      for (int i = lane; i < len; i += 32) {
        s_r[w][i] = __ldg(rewards + base + t0 + i);
        s_m[w][i] = mask[base + t0 + i];
      }
      __syncwarp();
      if (lane == 0) {
#pragma unroll 8
        for (int i = len - 1; i >= 0; --i) {
          if (REMAX) {
            acc += s_m[w][i] ? (double)s_r[w][i] : (double)__fmul_rn(s_r[w][i], 0.0f);
            s_r[w][i] = (float)acc;
          } else {
            run = __fadd_rn(s_r[w][i], __fmul_rn(gamma, run));
            s_r[w][i] = run;
            run = s_m[w][i] ? run : __fmul_rn(run, 0.0f);
          }
        }
      }
      __syncwarp();
#pragma unroll 8
      for (int i = lane; i < len; i += 32) {
        const float x = s_r[w][i];
        ret[base + t0 + i] = x;
        if (REMAX) {
          adv[base + t0 + i] = __fsub_rn(x, s_m[w][i] ? bl : __fmul_rn(bl, 0.0f));
        } else if (s_m[w][i]) {
          m0 += 1.0;
          m1 += (double)x;
          m2 += (double)x * (double)x;
        }
      }
      __syncwarp();
    }
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
  const float scale = (float)(1.0 / sqrt(var + 1e-8));
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    const float v = (x[i] - meanf) * scale;
    out[i] = mask[i] ? v : __fmul_rn(v, 0.0f);
  }
}

// apply_kl_penalty (verl/verl/trainer/ppo/ray_trainer.py:145-175) with kl_penalty (core_algos.py:469-501):
// kld = f(logp, ref) * mask; rewards = scores - beta * kld; row_kl[b] = sum(kld) / (sum(mask) + 1e-8).
// mode 0 "kl", 1 "abs", 2 "mse", 3 "low_var_kl".
__global__ void __launch_bounds__(ROWS_PER_CTA * 32) kl_rewards_kernel(int B, int T, const float* __restrict__ scores,
                                                                        const float* __restrict__ logp,
                                                                        const float* __restrict__ ref,
                                                                        const uint8_t* __restrict__ mask, int mode, float beta,
                                                                        float* __restrict__ out, float* __restrict__ row_kl) {
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int row = blockIdx.x * ROWS_PER_CTA + w;
  if (row >= B) return;
  const size_t base = (size_t)row * T;
  float ksum = 0.0f, msum = 0.0f;
#pragma unroll 4
  for (int t = lane; t < T; t += 32) {
    const float a = __ldg(logp + base + t), b = __ldg(ref + base + t);
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
    const bool m = mask[base + t] != 0;
    k = m ? k : __fmul_rn(k, 0.0f);
    out[base + t] = __fsub_rn(__ldg(scores + base + t), __fmul_rn(beta, k));
    ksum += k;
    msum += m ? 1.0f : 0.0f;
  }
  ksum = warp_sum(ksum);
  msum = warp_sum(msum);
  if (lane == 0 && row_kl) row_kl[row] = ksum / (msum + 1e-8f);
}

// _filter_rollout's statistics (ragen/trainer/agent_trainer.py:466-469): row sums of the score tensor, then per group
// of `group_size` consecutive rows the unbiased std, the max and the mean.  One warp per group, lane per row.
__global__ void __launch_bounds__(ROWS_PER_CTA * 32) filter_stats_kernel(int n_groups, int group_size, int T,
                                                                          const float* __restrict__ scores,
                                                                          float* __restrict__ row_sum, float* __restrict__ gstd,
                                                                          float* __restrict__ gmax, float* __restrict__ gmean) {
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = blockIdx.x * ROWS_PER_CTA + w;
  if (g >= n_groups) return;
  double s1 = 0.0, s2 = 0.0;
  float mx = -INFINITY;
  for (int r = 0; r < group_size; ++r) {  // the warp sums one row at a time (coalesced); every lane ends with the sum
    const size_t base = ((size_t)g * group_size + r) * T;
    double s = 0.0;
#pragma unroll 4
    for (int t = lane; t < T; t += 32) s += (double)__ldg(scores + base + t);
    const float sf = (float)warp_sum(s);
    if (lane == 0) row_sum[(size_t)g * group_size + r] = sf;
    s1 += (double)sf;
    s2 += (double)sf * (double)sf;
    mx = fmaxf(mx, sf);
  }
  if (lane == 0) {
    const double n = (double)group_size, mu = s1 / n;
    gmean[g] = (float)mu;
    gmax[g] = mx;
    gstd[g] = group_size > 1 ? (float)sqrt(fmax((s2 - n * mu * mu) / (n - 1.0), 0.0)) : NAN;  // torch.std of one value is NaN
  }
}

int grid_rows(int B) { return (B + ROWS_PER_CTA - 1) / ROWS_PER_CTA; }

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
  gae_kernel<<<grid_rows(B), ROWS_PER_CTA * 32, 0, st>>>(B, T, rewards, values, mask, (float)gamma, (float)(gamma * lam), adv,
                                                        ret, moments);
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
  bilevel_gae_kernel<<<grid_rows(B), ROWS_PER_CTA * 32, 0, st>>>(B, T, rewards, values, loss_mask, (float)gamma,
                                                                (float)(gamma * lam), (float)high_level_gamma,
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
  grpo_scores_kernel<<<grid_rows(B), ROWS_PER_CTA * 32, 0, st>>>(B, T, rewards, group, n_groups, scores, gstats);
  RAGEN_LAUNCHED(ctx);
  grpo_apply_kernel<<<grid_rows(B), ROWS_PER_CTA * 32, 0, st>>>(B, T, scores, group, n_groups, gstats, mask, (float)epsilon,
                                                               norm_by_std, adv);
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
  grpo_scores_kernel<<<grid_rows(B), ROWS_PER_CTA * 32, 0, st>>>(B, T, rewards, group, n_groups, scores, gstats);
  RAGEN_LAUNCHED(ctx);
  outcome_apply_kernel<<<grid_rows(B), ROWS_PER_CTA * 32, 0, st>>>(B, T, scores, group, n_groups, gstats, mask, (float)epsilon,
                                                                  mode, adv, moments);
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
  returns_kernel<0><<<grid_rows(B), ROWS_PER_CTA * 32, 0, st>>>(B, T, rewards, mask, (float)gamma, nullptr, nullptr, ret, moments);
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
  returns_kernel<1><<<grid_rows(B), ROWS_PER_CTA * 32, 0, (cudaStream_t)stream>>>(B, T, rewards, mask, 0.0f, baselines, adv, ret,
                                                                                 nullptr);
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
  kl_rewards_kernel<<<grid_rows(B), ROWS_PER_CTA * 32, 0, (cudaStream_t)stream>>>(B, T, scores, logp, ref_logp, mask, mode,
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
  filter_stats_kernel<<<grid_rows(n_groups), ROWS_PER_CTA * 32, 0, (cudaStream_t)stream>>>(n_groups, group_size, T, scores,
                                                                                          row_sum, gstd, gmax, gmean);
  RAGEN_LAUNCHED(ctx);
  return RAGEN_OK;
}
