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
  const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int row = blockIdx.x * ROWS_PER_CTA + w;
  double m0 = 0.0, m1 = 0.0, m2 = 0.0;
  if (row < B) {
    const size_t base = (size_t)row * T;
    float carry = 0.0f;
    for (int t1 = T; t1 > 0; t1 -= CHUNK) {
      const int t0 = max(0, t1 - CHUNK), len = t1 - t0;
      for (int i = lane; i < len; i += 32) {
        const int t = t0 + i;
        const float v = values[base + t];
        const float vn = (t + 1 < T) ? values[base + t + 1] : 0.0f;
        const float r = rewards[base + t];
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
      for (int i = lane; i < len; i += 32) {
        const int t = t0 + i;
        const float a = s_a[w][i];
        adv[base + t] = a;
        ret[base + t] = __fadd_rn(a, s_v[w][i]);
        if (mask[base + t]) {
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
      for (int i = lane; i < len; i += 32) {
        s_r[w][i] = rewards[base + t0 + i];
        s_v[w][i] = values[base + t0 + i];
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

int grid_rows(int B) { return (B + ROWS_PER_CTA - 1) / ROWS_PER_CTA; }

}  // namespace

#define RAGEN_BT(B, T) RAGEN_REQUIRE(B >= 0 && T >= 1 && (int64_t)B * T < ((int64_t)1 << 40), "bad shape B=%d T=%d", B, T)

extern "C" int ragen_gae(ragen_ctx* ctx, int B, int T, const float* rewards, const float* values, const uint8_t* mask,
                         double gamma, double lam, float* adv, float* ret, double* moments, void* stream) {
  RAGEN_REQUIRE(ctx, "ctx is NULL");
  RAGEN_BT(B, T);
  RAGEN_REQUIRE(rewards && values && mask && adv && ret, "NULL pointer argument");
  if (B == 0) return RAGEN_OK;
  DeviceGuard guard(ctx->device);
  cudaStream_t st = (cudaStream_t)stream;
  if (moments) RAGEN_CUDA(cudaMemsetAsync(moments, 0, 3 * sizeof(double), st));
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
  RAGEN_REQUIRE(rewards && values && loss_mask && adv && ret, "NULL pointer argument");
  if (B == 0) return RAGEN_OK;
  DeviceGuard guard(ctx->device);
  cudaStream_t st = (cudaStream_t)stream;
  if (moments) RAGEN_CUDA(cudaMemsetAsync(moments, 0, 3 * sizeof(double), st));
  bilevel_gae_kernel<<<grid_rows(B), ROWS_PER_CTA * 32, 0, st>>>(B, T, rewards, values, loss_mask, (float)gamma,
                                                                (float)(gamma * lam), (float)high_level_gamma,
                                                                (float)(high_level_gamma * lam), adv, ret, moments);
  RAGEN_LAUNCHED(ctx);
  return RAGEN_OK;
}

extern "C" int ragen_masked_moments(ragen_ctx* ctx, int64_t n, const float* x, const uint8_t* mask, double* moments,
                                    void* stream) {
  RAGEN_REQUIRE(ctx, "ctx is NULL");
  RAGEN_REQUIRE(n >= 0 && x && mask && moments, "bad argument");
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
  RAGEN_REQUIRE(n >= 0 && x && moments, "bad argument");
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
