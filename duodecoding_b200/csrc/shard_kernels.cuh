// Vocabulary-sharded verification (V = 151936 over G GPUs).  Every rank holds a
// contiguous slice [v_begin, v_begin + Vs) of the draft and target logits of ALL rows; the step is
//
//   shard_stats_kernel   per row: (slice max, sum exp(x - slice max), scaled logit of the drafted
//                        token if it lies in the slice)                      -> float4 per row
//        -- exchange 1: all-gather of the float4s (host: NCCL) --
//   shard_accept_kernel  per sequence: merge the G partial (max, sum) pairs of every row, p(x_j),
//                        q(x_j), r <= p/q (raw IEEE division), leading run n; publishes the merged
//                        (M, Z) of the one row pair the token is drawn from
//   shard_sample_kernel  per sequence, over the slice: residual d = max(0, p - q) (or p for the
//                        bonus token): local mass, best race key (d / noise, global index), first
//                        argmax(p - q) for the zero-mass fallback              -> 6 words per sequence
//        -- exchange 2: all-gather of the partials --
//   shard_finish_kernel  per sequence: total mass, best key over the ranks -> token
//
// Rows are identified like in the fused path: job = b * (2*gamma+1) + k (draft j, target j, ..., bonus).
// Reference semantics: src/utils.py:686-709 (softmax of logits / T), src/decoding_ops.py:127-147,
// :190-220, src/utils.py:727-748, :787-810.  No top-k / top-p on this path (C5 runs without filter).
#pragma once

#include "common.cuh"

namespace duodec {

struct ShardRows {
  const void* draft;
  const void* target;
  long long d_sb, d_ss, t_sb, t_ss;  // element strides of the local slices
  int B, gamma, Vs, v_begin;
  float t_temp, d_temp;
};

template <class TIn>
__device__ __forceinline__ const TIn* shard_row(const ShardRows& S, int b, int j, int is_target) {
  return is_target ? reinterpret_cast<const TIn*>(S.target) + (long long)b * S.t_sb + (long long)j * S.t_ss
                   : reinterpret_cast<const TIn*>(S.draft) + (long long)b * S.d_sb + (long long)j * S.d_ss;
}

// one block per row: slice max, then sum exp(x - max) (second read comes out of L1/L2)
template <class TIn>
__global__ void __launch_bounds__(kThreads) shard_stats_kernel(const __grid_constant__ ShardRows S,
                                                               const long long* tokens, long long tok_sb,
                                                               float4* stats) {
  __shared__ Xchg xc;
  const int job = blockIdx.x, per_b = 2 * S.gamma + 1;
  const int b = job / per_b, k = job - b * per_b;
  const int is_target = (k == 2 * S.gamma) ? 1 : (k & 1), j = (k == 2 * S.gamma) ? S.gamma : (k >> 1);
  const TIn* row = shard_row<TIn>(S, b, j, is_target);
  const float temp = is_target ? S.t_temp : S.d_temp;
  float mx = -INFINITY;
  for (int i = threadIdx.x; i < S.Vs; i += kThreads) mx = fmaxf(mx, scale_logit(to_f32(row[i]), temp));
  mx = block_allreduce(mx, OpMaxF(), &xc);
  float s = 0.0f;
  if (mx > -INFINITY)
    for (int i = threadIdx.x; i < S.Vs; i += kThreads) s += expf(scale_logit(to_f32(row[i]), temp) - mx);
  s = block_allreduce(s, OpSumF(), &xc);
  if (threadIdx.x == 0) {
    float tokx = 0.0f, has = 0.0f;
    if (j < S.gamma) {
      const long long t = tokens[(long long)b * tok_sb + j] - S.v_begin;
      if (t >= 0 && t < S.Vs) tokx = scale_logit(to_f32(row[t]), temp), has = 1.0f;
    }
    stats[job] = make_float4(mx, s, tokx, has);
  }
}

struct ShardPlan {  // per sequence, written by shard_accept_kernel
  float Mp, Zp, Mq, Zq;
  int n, bonus, pad0, pad1;
};

// merged (M, Z, token logit) of one row from the G gathered partials
__device__ __forceinline__ void shard_merge(const float4* gathered, int G, int rows, int job, float& M, float& Z,
                                            float& tokx, bool& has) {
  M = -INFINITY;
  for (int g = 0; g < G; ++g) M = fmaxf(M, gathered[(long long)g * rows + job].x);
  Z = 0.0f, tokx = 0.0f, has = false;
  for (int g = 0; g < G; ++g) {  // rank order: the same sum on every rank
    const float4 q = gathered[(long long)g * rows + job];
    if (q.x > -INFINITY) Z += q.y * expf(q.x - M);
    if (q.w != 0.0f) tokx = q.z, has = true;
  }
}

// one warp per sequence (gamma <= 32)
__global__ void shard_accept_kernel(const float4* gathered, int G, int B, int gamma, const float* r, long long r_sb,
                                    unsigned char* accept_mask, float* sel_p, float* sel_q, long long* accepted_len,
                                    unsigned char* all_accepted, ShardPlan* plan) {
  const int b = blockIdx.x, lane = threadIdx.x;
  const int per_b = 2 * gamma + 1, rows = B * per_b, base = b * per_b;
  int accept = 1;
  if (lane < gamma) {
    float Mq, Zq, xq, Mp, Zp, xp;
    bool hq, hp;
    shard_merge(gathered, G, rows, base + 2 * lane, Mq, Zq, xq, hq);
    shard_merge(gathered, G, rows, base + 2 * lane + 1, Mp, Zp, xp, hp);
    const float q = hq ? __fdiv_rn(expf(xq - Mq), Zq) : 0.0f;
    const float p = hp ? __fdiv_rn(expf(xp - Mp), Zp) : 0.0f;
    accept = r[(long long)b * r_sb + lane] <= __fdiv_rn(p, q);  // raw IEEE division
    const long long o = (long long)b * gamma + lane;
    accept_mask[o] = (unsigned char)accept;
    sel_p[o] = p;
    sel_q[o] = q;
  }
  const unsigned rejected = __ballot_sync(0xffffffffu, lane < gamma && !accept);
  const int n = rejected ? __ffs(rejected) - 1 : gamma;
  if (lane == 0) {
    ShardPlan pl;
    float x;
    bool h;
    pl.n = n, pl.bonus = (n == gamma), pl.pad0 = pl.pad1 = 0;
    shard_merge(gathered, G, rows, base + (pl.bonus ? 2 * gamma : 2 * n + 1), pl.Mp, pl.Zp, x, h);
    pl.Mq = 0.0f, pl.Zq = 1.0f;
    if (!pl.bonus) shard_merge(gathered, G, rows, base + 2 * n, pl.Mq, pl.Zq, x, h);
    plan[b] = pl;
    accepted_len[b] = n;
    all_accepted[b] = (unsigned char)pl.bonus;
  }
}

struct ShardPartial {  // per sequence and rank: 24 bytes, exchanged as is
  float mass;
  unsigned pad;
  unsigned long long key;  // (bits(d / noise) << 32) | ~global index, d > 0; 0: none
  unsigned long long bd;   // (okey(p - q) << 32) | ~global index
};

struct OpMaxKey2 {
  __device__ __forceinline__ uint4 operator()(uint4 a, uint4 b) const {
    const unsigned long long ak = ((unsigned long long)a.x << 32) | a.y, bk = ((unsigned long long)b.x << 32) | b.y;
    const unsigned long long ad = ((unsigned long long)a.z << 32) | a.w, bd = ((unsigned long long)b.z << 32) | b.w;
    const unsigned long long k = ak > bk ? ak : bk, d = ad > bd ? ad : bd;
    return make_uint4(unsigned(k >> 32), unsigned(k), unsigned(d >> 32), unsigned(d));
  }
};

// one block per sequence over the local slice of the row pair named by the plan
template <class TIn>
__global__ void __launch_bounds__(kThreads) shard_sample_kernel(const __grid_constant__ ShardRows S,
                                                                const ShardPlan* plan, const float* noise,
                                                                long long noise_sb, ShardPartial* out) {
  __shared__ Xchg xc;
  const int b = blockIdx.x;
  const ShardPlan pl = plan[b];
  const TIn* prow = shard_row<TIn>(S, b, pl.n, 1);
  const TIn* qrow = shard_row<TIn>(S, b, pl.bonus ? 0 : pl.n, 0);
  const float* nz = noise + (long long)b * noise_sb;
  float mass = 0.0f;
  unsigned long long key = 0ull, bdk = 0ull;
  for (int i = threadIdx.x; i < S.Vs; i += kThreads) {
    float d = __fdiv_rn(expf(scale_logit(to_f32(prow[i]), S.t_temp) - pl.Mp), pl.Zp);
    if (!pl.bonus) d -= __fdiv_rn(expf(scale_logit(to_f32(qrow[i]), S.d_temp) - pl.Mq), pl.Zq);
    const unsigned gi = 0xffffffffu - unsigned(S.v_begin + i);
    if (d > 0.0f) {
      mass += d;
      const unsigned long long k = ((unsigned long long)__float_as_uint(__fdiv_rn(d, nz[i])) << 32) | gi;
      key = k > key ? k : key;
    }
    const float dz = (d == d && fabsf(d) != INFINITY) ? d : 0.0f;
    const unsigned long long k2 = ((unsigned long long)okey(dz) << 32) | gi;
    bdk = k2 > bdk ? k2 : bdk;
  }
  mass = block_allreduce(mass, OpSumF(), &xc);
  const uint4 kk = block_allreduce(make_uint4(unsigned(key >> 32), unsigned(key), unsigned(bdk >> 32), unsigned(bdk)),
                                   OpMaxKey2(), &xc);
  if (threadIdx.x == 0) {
    ShardPartial p;
    p.mass = mass, p.pad = 0u;
    p.key = ((unsigned long long)kk.x << 32) | kk.y;
    p.bd = ((unsigned long long)kk.z << 32) | kk.w;
    out[b] = p;
  }
}

__global__ void shard_finish_kernel(const ShardPartial* gathered, int G, int B, long long* token) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  float mass = 0.0f;
  unsigned long long key = 0ull, bd = 0ull;
  for (int g = 0; g < G; ++g) {
    const ShardPartial p = gathered[(long long)g * B + b];
    mass += p.mass;
    key = p.key > key ? p.key : key;
    bd = p.bd > bd ? p.bd : bd;
  }
  const unsigned long long win = (mass > 0.0f && key != 0ull) ? key : bd;
  token[b] = win == 0ull ? 0 : (long long)(0xffffffffu - unsigned(win & 0xffffffffull));
}

}  // namespace duodec
