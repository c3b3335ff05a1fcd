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
//                        argmax(p - q) for the zero-mass fallback   -> 24 bytes per sequence and slice part
//        -- exchange 2: all-gather of the partials --
//   shard_finish_kernel  per sequence: total mass, best key over the ranks -> token
//
// Rows are identified like in the fused path: job = b * (2*gamma+1) + k (draft j, target j, ..., bonus).
// Reference semantics: src/utils.py:686-709 (softmax of logits / T), src/decoding_ops.py:127-147,
// :190-220, src/utils.py:727-748, :787-810.  No top-k / top-p on this path (C5 runs without filter).
//
// Fused exchange (ShardXchg): with peer-mapped symmetric buffers (NVLink / NVSwitch P2P) the two exchanges
// happen INSIDE the kernels - shard_stats / shard_sample store their 16- / 24-byte results straight into every
// peer's gather buffer; the block that finishes last publishes the rank's step number to a flag in every
// peer (st.release.sys); shard_accept / shard_finish spin on their own flags (ld.acquire.sys) and read the
// gathered data locally.  No NCCL call, no extra launch; a step is four kernels in one CUDA graph.
#pragma once

#include <type_traits>

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

// eight consecutive entries of a row as floats (two 128-bit loads for fp32, one for bf16)
template <class TIn>
__device__ __forceinline__ void load8(const TIn* p, float (&x)[8]) {
  if constexpr (sizeof(TIn) == 4) {
    const float4 a = __ldg(reinterpret_cast<const float4*>(p)), b = __ldg(reinterpret_cast<const float4*>(p) + 1);
    x[0] = a.x, x[1] = a.y, x[2] = a.z, x[3] = a.w, x[4] = b.x, x[5] = b.y, x[6] = b.z, x[7] = b.w;
  } else {
    const uint4 u = __ldg(reinterpret_cast<const uint4*>(p));
    x[0] = __uint_as_float(u.x << 16), x[1] = __uint_as_float(u.x & 0xffff0000u);
    x[2] = __uint_as_float(u.y << 16), x[3] = __uint_as_float(u.y & 0xffff0000u);
    x[4] = __uint_as_float(u.z << 16), x[5] = __uint_as_float(u.z & 0xffff0000u);
    x[6] = __uint_as_float(u.w << 16), x[7] = __uint_as_float(u.w & 0xffff0000u);
  }
}
template <class TIn>
__device__ __forceinline__ bool row_vec8(const TIn* row, int n) {
  return (n & 7) == 0 && (reinterpret_cast<uintptr_t>(row) & (8 * sizeof(TIn) - 1)) == 0;
}
__device__ __forceinline__ float ex2f(float t) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(t));
  return y;
}
constexpr float kShardLog2e = 1.4426950408889634f;

// ---- fused exchange over peer-mapped memory --------------------------------------------------------
// Every rank owns one symmetric buffer: a 1 KB header (flags of exchange 1 at +0, of exchange 2 at +256, one u32
// per rank; block counters at +512 / +516; the step number at +520), then the gather areas.
constexpr int kXFlags1 = 0, kXFlags2 = 256, kXCnt1 = 512, kXCnt2 = 516, kXStep = 520;
struct ShardXchg {
  void* const* peers;  // device array of `world` base pointers (own base at [rank]); nullptr: no fused exchange
  int rank, world;
  long long off_stats, off_partial;  // byte offsets of gstats [world][rows] float4 / gpartial [world][B * nslice]
  int mode;  // 0: every block stores its result into every peer; 1: the last block copies the rank's slab to the peers;
             // 2 ("pull"): results stay in the rank's OWN buffer, the caller puts a barrier over all ranks between the
             // producing and the consuming kernel (torch symmetric memory: hdl.barrier()), and the consumers read every
             // peer's slab in place over NVLink - no copy, no flag protocol in the kernels
};
constexpr int kXPull = 2;
// slab of rank g inside the gather area at byte offset `off`: in pull mode it lives in rank g's own buffer
__device__ __forceinline__ const unsigned char* xchg_area(const ShardXchg& X, int g, long long off) {
  return static_cast<const unsigned char*>(X.peers[X.mode == kXPull ? g : X.rank]) + off;
}
__device__ __forceinline__ float4 ld_peer_f4(const float4* p) {  // peer memory: never from a stale L1 line
  float4 v;
  asm volatile("ld.volatile.global.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
  return v;
}
__device__ __forceinline__ unsigned ld_acquire_sys(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_sys(unsigned* p, unsigned v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
// Whole block, after its thread 0 stored the block's result into the rank's OWN slab of the gather area: the block
// that finishes last copies the slab (slab_bytes, a multiple of 8) into every peer's area with coalesced stores,
// then publishes the step to a flag in every peer.  One system-scope fence per grid, not per block.
__device__ __forceinline__ void xchg_publish(const ShardXchg& X, int flags_off, int cnt_off, unsigned nblocks,
                                             long long area_off, long long slab_bytes, long long item_off,
                                             int item_bytes) {
  __shared__ int s_last;
  unsigned char* self = static_cast<unsigned char*>(X.peers[X.rank]);
  if (X.mode == 0) {  // every block pushes its own item (at item_off inside the rank's slab) to every peer
    if (threadIdx.x == 0) {
      const unsigned long long* src =
          reinterpret_cast<const unsigned long long*>(self + area_off + X.rank * slab_bytes + item_off);
      for (int g = 0; g < X.world; ++g) {
        if (g == X.rank) continue;
        unsigned long long* dst = reinterpret_cast<unsigned long long*>(static_cast<unsigned char*>(X.peers[g]) +
                                                                        area_off + X.rank * slab_bytes + item_off);
        for (int i = 0; i < item_bytes / 8; ++i) dst[i] = src[i];
      }
      __threadfence_system();
      if (atomicAdd(reinterpret_cast<unsigned*>(self + cnt_off), 1u) == nblocks - 1u) {
        *reinterpret_cast<unsigned*>(self + cnt_off) = 0u;
        __threadfence_system();
        const unsigned epoch = *reinterpret_cast<volatile unsigned*>(self + kXStep) + 1u;
        for (int g = 0; g < X.world; ++g)
          st_release_sys(reinterpret_cast<unsigned*>(static_cast<unsigned char*>(X.peers[g]) + flags_off) + X.rank, epoch);
      }
    }
    return;
  }
  if (threadIdx.x == 0) {
    __threadfence();
    s_last = atomicAdd(reinterpret_cast<unsigned*>(self + cnt_off), 1u) == nblocks - 1u;
  }
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  const unsigned long long* src = reinterpret_cast<const unsigned long long*>(self + area_off + X.rank * slab_bytes);
  const long long n8 = slab_bytes / 8;
  for (int g = 0; g < X.world; ++g) {
    if (g == X.rank) continue;
    unsigned long long* dst =
        reinterpret_cast<unsigned long long*>(static_cast<unsigned char*>(X.peers[g]) + area_off + X.rank * slab_bytes);
    for (long long i = threadIdx.x; i < n8; i += blockDim.x) dst[i] = __ldcg(src + i);
  }
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    *reinterpret_cast<unsigned*>(self + cnt_off) = 0u;  // (every block of this grid has counted: ready for the next step)
    const unsigned epoch = *reinterpret_cast<volatile unsigned*>(self + kXStep) + 1u;
    for (int g = 0; g < X.world; ++g)
      st_release_sys(reinterpret_cast<unsigned*>(static_cast<unsigned char*>(X.peers[g]) + flags_off) + X.rank, epoch);
  }
}
// whole block: until every rank has published this step (traps after ~2 s instead of hanging the device)
__device__ __forceinline__ void xchg_wait(const ShardXchg& X, int flags_off) {
  const unsigned char* self = static_cast<const unsigned char*>(X.peers[X.rank]);
  if (int(threadIdx.x) < X.world) {
    const unsigned epoch = *reinterpret_cast<const volatile unsigned*>(self + kXStep) + 1u;
    const unsigned* f = reinterpret_cast<const unsigned*>(self + flags_off) + threadIdx.x;
    const long long t0 = clock64();
    while (int(ld_acquire_sys(f) - epoch) < 0) {
      if (clock64() - t0 > 4000000000ll) __trap();
    }
  }
  __syncthreads();
}

struct MaxSum {  // running (max, sum of exp(x - max)); merge is associative, fixed tree order
  float m, s;
};
struct OpMaxSum {
  __device__ __forceinline__ MaxSum operator()(MaxSum a, MaxSum b) const {
    const float m = fmaxf(a.m, b.m);
    if (m == -INFINITY) return MaxSum{m, 0.0f};
    return MaxSum{m, a.s * ex2f((a.m - m) * kShardLog2e) + b.s * ex2f((b.m - m) * kShardLog2e)};
  }
};

// one block per row, ONE pass: every thread keeps an online (max, sum), rescaled when its max moves
template <class TIn>
__global__ void __launch_bounds__(kThreads) shard_stats_kernel(const __grid_constant__ ShardRows S,
                                                               const long long* tokens, long long tok_sb,
                                                               float4* stats, const ShardXchg X) {
  __shared__ Xchg xc;
  const int job = blockIdx.x, per_b = 2 * S.gamma + 1;
  const int b = job / per_b, k = job - b * per_b;
  const int is_target = (k == 2 * S.gamma) ? 1 : (k & 1), j = (k == 2 * S.gamma) ? S.gamma : (k >> 1);
  const TIn* row = shard_row<TIn>(S, b, j, is_target);
  const float temp = is_target ? S.t_temp : S.d_temp;
  MaxSum acc{-INFINITY, 0.0f};
  // the temperature test is hoisted out of the entry loops (kUnit: temperature 1, no division): a per-entry
  // `T == 1 ? x : x / T` keeps a branch and an inlined IEEE division per entry in the loop body
  auto body = [&](auto unit_tag) {
  constexpr bool kUnit = decltype(unit_tag)::value;
  auto push8 = [&](float (&x)[8]) {
    float vm = -INFINITY;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if (!kUnit) x[i] = __fdiv_rn(x[i], temp);
      vm = fmaxf(vm, x[i]);
    }
    if (vm > acc.m) {
      acc.s *= ex2f((acc.m - vm) * kShardLog2e);  // (0 * 0 on the first group)
      acc.m = vm;
    }
    if (acc.m > -INFINITY) {
      const float mo = acc.m * kShardLog2e;
#pragma unroll
      for (int i = 0; i < 8; ++i) acc.s += ex2f(fmaf(x[i], kShardLog2e, -mo));
    }
  };
  if (row_vec8(row, S.Vs)) {
    // kStatBatch groups of eight entries per trip: all their loads are issued before the first is consumed (a
    // slice of V / 8 is under five groups per thread: one trip, every load of the row in flight at once)
    constexpr int kStatBatch = 5;
    for (int i0 = threadIdx.x * 8; i0 < S.Vs; i0 += kThreads * 8 * kStatBatch) {
      float x[kStatBatch][8];
#pragma unroll
      for (int k = 0; k < kStatBatch; ++k) {
        const int i = i0 + k * kThreads * 8;
        if (i < S.Vs) {
          load8<TIn>(row + i, x[k]);
        } else {
#pragma unroll
          for (int u = 0; u < 8; ++u) x[k][u] = -INFINITY;  // (contributes nothing)
        }
      }
#pragma unroll
      for (int k = 0; k < kStatBatch; ++k) push8(x[k]);
    }
  } else {
    for (int i = threadIdx.x; i < S.Vs; i += kThreads) {
      float x[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) x[u] = -INFINITY;
      x[0] = to_f32(row[i]);
      push8(x);
    }
  }
  };
  if (temp == 1.0f) body(std::true_type{});
  else body(std::false_type{});
  if (!(acc.s == acc.s)) acc.s = 0.0f;  // -inf - -inf on a thread that saw nothing
  acc = block_allreduce(acc, OpMaxSum(), &xc);
  if (threadIdx.x == 0) {
    float tokx = 0.0f, has = 0.0f;
    if (j < S.gamma) {
      const long long t = tokens[(long long)b * tok_sb + j] - S.v_begin;
      if (t >= 0 && t < S.Vs) tokx = scale_logit(to_f32(row[t]), temp), has = 1.0f;
    }
    const float4 v = make_float4(acc.m, acc.s, tokx, has);
    if (X.peers == nullptr) stats[job] = v;
    else  // the rank's own slab of the gather area, slot [rank][job]
      reinterpret_cast<float4*>(static_cast<unsigned char*>(X.peers[X.rank]) + X.off_stats)[(long long)X.rank * gridDim.x + job] = v;
  }
  if (X.peers != nullptr && X.mode != kXPull)
    xchg_publish(X, kXFlags1, kXCnt1, gridDim.x, X.off_stats, (long long)gridDim.x * 16, (long long)job * 16, 16);
}

// A row's nucleus filter (shard_nucleus.cuh): entry (x, i) is kept iff x > xstar or (x == xstar and i <= istar)
struct __align__(16) ShardCut {
  float xstar;  // boundary value (scaled logit); -inf: everything is kept
  int istar;    // entries equal to xstar are kept up to this GLOBAL index
  float Zk;     // kept mass relative to the row maximum: p_i = exp(x_i - M) / Zk
  int state;    // 0 filtered, 1 everything kept (no crossing), 2 no filter on this row kind
};
__device__ __forceinline__ bool cut_keeps(float x, long long gi, float xstar, int istar) {
  return x > xstar || (x == xstar && gi <= (long long)istar);
}

struct __align__(16) ShardPlan {  // per sequence, written by shard_accept_kernel
  float Mp, Zp, Mq, Zq;
  int n, bonus;
  float xsp;  // the filter of the row pair the token is drawn from (-inf / INT_MAX: none)
  int isp;
  float xsq;
  int isq, pad0, pad1;
};
static_assert(sizeof(ShardPlan) == 48, "ShardPlan is exchanged with the host as 48 opaque bytes");

// merged (M, Z, token logit) of one row from the G gathered partials
__device__ __forceinline__ void shard_merge(const float4* gathered, int G, int rows, int job, float& M, float& Z,
                                            float& tokx, bool& has, const ShardXchg* pull = nullptr) {
  constexpr int kMaxG = 32;
  M = -INFINITY, Z = 0.0f, tokx = 0.0f, has = false;
  if (pull != nullptr && G <= kMaxG) {
    // every rank's record sits in that rank's own buffer: all G loads are issued before the first is used
    float4 q[kMaxG];
#pragma unroll 8
    for (int g = 0; g < G; ++g)
      q[g] = ld_peer_f4(reinterpret_cast<const float4*>(xchg_area(*pull, g, pull->off_stats)) + (long long)g * rows + job);
    for (int g = 0; g < G; ++g) M = fmaxf(M, q[g].x);
    for (int g = 0; g < G; ++g) {
      if (q[g].x > -INFINITY) Z += q[g].y * expf(q[g].x - M);
      if (q[g].w != 0.0f) tokx = q[g].z, has = true;
    }
    return;
  }
  for (int g = 0; g < G; ++g) M = fmaxf(M, gathered[(long long)g * rows + job].x);
  for (int g = 0; g < G; ++g) {  // rank order: the same sum on every rank
    const float4 q = gathered[(long long)g * rows + job];
    if (q.x > -INFINITY) Z += q.y * expf(q.x - M);
    if (q.w != 0.0f) tokx = q.z, has = true;
  }
}

// one warp per sequence (gamma <= 32)
__global__ void shard_accept_kernel(const float4* gathered, int G, int B, int gamma, const float* r, long long r_sb,
                                    unsigned char* accept_mask, float* sel_p, float* sel_q, long long* accepted_len,
                                    unsigned char* all_accepted, ShardPlan* plan, const ShardXchg X,
                                    const long long* tokens, long long tok_sb, const ShardCut* cuts) {
  const ShardXchg* pull = (X.peers != nullptr && X.mode == kXPull) ? &X : nullptr;
  if (X.peers != nullptr && pull == nullptr) {
    xchg_wait(X, kXFlags1);
    gathered = reinterpret_cast<const float4*>(static_cast<const unsigned char*>(X.peers[X.rank]) + X.off_stats);
  }
  const int b = blockIdx.x, lane = threadIdx.x;
  const int per_b = 2 * gamma + 1, rows = B * per_b, base = b * per_b;
  int accept = 1;
  if (lane < gamma) {
    float Mq, Zq, xq, Mp, Zp, xp;
    bool hq, hp;
    shard_merge(gathered, G, rows, base + 2 * lane, Mq, Zq, xq, hq, pull);
    shard_merge(gathered, G, rows, base + 2 * lane + 1, Mp, Zp, xp, hp, pull);
    if (cuts != nullptr) {  // nucleus rows: the kept mass normalises, a filtered-out token has probability 0
      const ShardCut cq = cuts[base + 2 * lane], cp = cuts[base + 2 * lane + 1];
      const long long tk = tokens[(long long)b * tok_sb + lane];
      Zq = cq.Zk, Zp = cp.Zk;
      hq = hq && cut_keeps(xq, tk, cq.xstar, cq.istar);
      hp = hp && cut_keeps(xp, tk, cp.xstar, cp.istar);
    }
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
    const int prow = base + (pl.bonus ? 2 * gamma : 2 * n + 1), qrow = base + 2 * n;
    shard_merge(gathered, G, rows, prow, pl.Mp, pl.Zp, x, h, pull);
    pl.Mq = 0.0f, pl.Zq = 1.0f;
    if (!pl.bonus) shard_merge(gathered, G, rows, qrow, pl.Mq, pl.Zq, x, h, pull);
    pl.xsp = pl.xsq = -INFINITY, pl.isp = pl.isq = 0x7fffffff;
    if (cuts != nullptr) {
      const ShardCut cp = cuts[prow];
      pl.Zp = cp.Zk, pl.xsp = cp.xstar, pl.isp = cp.istar;
      if (!pl.bonus) {
        const ShardCut cq = cuts[qrow];
        pl.Zq = cq.Zk, pl.xsq = cq.xstar, pl.isq = cq.istar;
      }
    }
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

// grid (nslice, B): block (s, b) covers an eighth-aligned part of the local slice of the row pair
// named by the plan and writes partial[b * nslice + s]
template <class TIn, bool kCut>
__global__ void __launch_bounds__(kThreads) shard_sample_kernel(const __grid_constant__ ShardRows S,
                                                                const ShardPlan* plan, const float* noise,
                                                                long long noise_sb, ShardPartial* out,
                                                                const ShardXchg X) {
  __shared__ Xchg xc;
  const int b = blockIdx.y, nslice = gridDim.x;
  const ShardPlan pl = plan[b];
  const TIn* prow = shard_row<TIn>(S, b, pl.n, 1);
  const TIn* qrow = shard_row<TIn>(S, b, pl.bonus ? 0 : pl.n, 0);
  const float* nz = noise + (long long)b * noise_sb;
  int per = (S.Vs + nslice - 1) / nslice;
  per = (per + 7) & ~7;
  const int lo = min(int(blockIdx.x) * per, S.Vs), hi = min(lo + per, S.Vs);
  const float rZp = __frcp_rn(pl.Zp), rZq = __frcp_rn(pl.Zq);
  const float mp = pl.Mp * kShardLog2e, mq = pl.Mq * kShardLog2e;
  float mass = 0.0f, bkey = -1.0f, bdv = -INFINITY;
  int bidx = 0x7fffffff, bdi = 0x7fffffff;
  // branch-free per entry; a thread meets its indices in increasing order, so strict ">" keeps the lowest
  // index of a tie (same arithmetic as the fused path's sample kernel: __fdividef race keys)
  // the temperature tests and the bonus / residual choice are uniform over the block: hoisted out of the entry loop
  auto body = [&](auto unit_tag, auto bonus_tag) {
    constexpr bool kUnit = decltype(unit_tag)::value, kBonus = decltype(bonus_tag)::value;
    auto one = [&](float xp, float xq, float e, int i) {
      const float xps = kUnit ? xp : __fdiv_rn(xp, S.t_temp);
      float d = ex2f(fmaf(xps, kShardLog2e, -mp)) * rZp;
      if (kCut) d = cut_keeps(xps, S.v_begin + i, pl.xsp, pl.isp) ? d : 0.0f;
      if (!kBonus) {
        const float xqs = kUnit ? xq : __fdiv_rn(xq, S.d_temp);
        float dq = ex2f(fmaf(xqs, kShardLog2e, -mq)) * rZq;
        if (kCut) dq = cut_keeps(xqs, S.v_begin + i, pl.xsq, pl.isq) ? dq : 0.0f;
        d -= dq;
      }
      const bool pos = d > 0.0f;
      mass += pos ? d : 0.0f;
      const float k = pos ? __fdividef(d, e) : -1.0f;
      const bool bk = k > bkey;
      bkey = bk ? k : bkey, bidx = bk ? i : bidx;
      const float dz = (fabsf(d) < INFINITY) ? d : 0.0f;  // NaN and +-inf count as 0
      const bool bd = dz > bdv;
      bdv = bd ? dz : bdv, bdi = bd ? i : bdi;
    };
    const bool vec = row_vec8(prow, S.Vs) && row_vec8(qrow, S.Vs) && (reinterpret_cast<uintptr_t>(nz) & 31) == 0;
    if (vec) {
      for (int i = lo + threadIdx.x * 8; i < hi; i += kThreads * 8) {
        float xp[8], xq[8], e[8];
        load8<TIn>(prow + i, xp);
        if (!kBonus) load8<TIn>(qrow + i, xq);
        load8<float>(nz + i, e);
#pragma unroll
        for (int u = 0; u < 8; ++u) one(xp[u], kBonus ? 0.0f : xq[u], e[u], i + u);
      }
    } else {
      for (int i = lo + threadIdx.x; i < hi; i += kThreads)
        one(to_f32(prow[i]), kBonus ? 0.0f : to_f32(qrow[i]), nz[i], i);
    }
  };
  const bool unit = S.t_temp == 1.0f && S.d_temp == 1.0f;  // (mixed temperatures take the dividing variant: x / 1 == x)
  if (unit) {
    if (pl.bonus) body(std::true_type{}, std::true_type{});
    else body(std::true_type{}, std::false_type{});
  } else {
    if (pl.bonus) body(std::false_type{}, std::true_type{});
    else body(std::false_type{}, std::false_type{});
  }
  const unsigned long long key =
      bidx == 0x7fffffff ? 0ull
                         : ((unsigned long long)__float_as_uint(bkey) << 32) | (0xffffffffu - unsigned(S.v_begin + bidx));
  const unsigned long long bdk =
      bdi == 0x7fffffff ? 0ull : ((unsigned long long)okey(bdv) << 32) | (0xffffffffu - unsigned(S.v_begin + bdi));
  mass = block_allreduce(mass, OpSumF(), &xc);
  const uint4 kk = block_allreduce(make_uint4(unsigned(key >> 32), unsigned(key), unsigned(bdk >> 32), unsigned(bdk)),
                                   OpMaxKey2(), &xc);
  if (threadIdx.x == 0) {
    ShardPartial p;
    p.mass = mass, p.pad = 0u;
    p.key = ((unsigned long long)kk.x << 32) | kk.y;
    p.bd = ((unsigned long long)kk.z << 32) | kk.w;
    if (X.peers == nullptr) out[(long long)b * nslice + blockIdx.x] = p;
    else
      reinterpret_cast<ShardPartial*>(static_cast<unsigned char*>(X.peers[X.rank]) + X.off_partial)
          [((long long)X.rank * gridDim.y + b) * nslice + blockIdx.x] = p;
  }
  if (X.peers != nullptr && X.mode != kXPull)
    xchg_publish(X, kXFlags2, kXCnt2, gridDim.x * gridDim.y, X.off_partial, (long long)gridDim.x * gridDim.y * 24,
                 ((long long)b * nslice + blockIdx.x) * 24, 24);
}

__global__ void shard_finish_kernel(const ShardPartial* gathered, int G, int B, int nslice, long long* token,
                                    const ShardXchg X) {
  const bool pull = X.peers != nullptr && X.mode == kXPull;
  if (X.peers != nullptr && !pull) {  // (one block in this mode: it closes the step)
    xchg_wait(X, kXFlags2);
    gathered = reinterpret_cast<const ShardPartial*>(static_cast<const unsigned char*>(X.peers[X.rank]) + X.off_partial);
  }
  for (int b = blockIdx.x * blockDim.x + threadIdx.x; b < B; b += gridDim.x * blockDim.x) {
    float mass = 0.0f;
    unsigned long long key = 0ull, bd = 0ull;
    for (int g = 0; g < G * nslice; ++g) {  // rank-major, then slice: the same order on every rank
      const long long at = ((long long)(g / nslice) * B + b) * nslice + (g % nslice);
      ShardPartial p;
      if (pull) {  // rank g / nslice keeps its partials in its own buffer
        const volatile unsigned long long* src = reinterpret_cast<const volatile unsigned long long*>(
            reinterpret_cast<const ShardPartial*>(xchg_area(X, g / nslice, X.off_partial)) + at);
        const unsigned long long w0 = src[0];
        p.mass = __uint_as_float(unsigned(w0)), p.pad = 0u, p.key = src[1], p.bd = src[2];
      } else {
        p = gathered[at];
      }
      mass += p.mass;
      key = p.key > key ? p.key : key;
      bd = p.bd > bd ? p.bd : bd;
    }
    const unsigned long long win = (mass > 0.0f && key != 0ull) ? key : bd;
    token[b] = win == 0ull ? 0 : (long long)(0xffffffffu - unsigned(win & 0xffffffffull));
  }
  if (X.peers != nullptr && !pull) {
    __syncthreads();
    if (threadIdx.x == 0) {
      unsigned* step = reinterpret_cast<unsigned*>(static_cast<unsigned char*>(X.peers[X.rank]) + kXStep);
      *step = *step + 1u;  // the next step's kernels (stream order) publish and wait for step + 1
    }
  }
}

}  // namespace duodec
