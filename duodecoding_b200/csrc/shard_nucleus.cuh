// Top-p (nucleus) filter under vocabulary sharding ("top-p under sharding needs a distributed
// radix select").  Reference semantics: top_k_top_p_filter, /root/reference/src/utils.py:650-683 - sort the
// row's probabilities in descending order, keep every entry whose EXCLUSIVE prefix mass is <= top_p (the entry
// that crosses top_p is kept), softmax over the kept set; ties in the oracle's canonical order (lowest index
// first).
//
// Every rank holds a slice of every row.  After the first exchange (global row maximum M) the boundary is found
// by a radix select over the order-preserving 32-bit key of the scaled logit, 4 bits per level:
//
//   shard_nuc_level_kernel   per row: 16 bins over the level's key digit among the slice's entries that match
//                            the prefix found so far: (mass, count) per bin          -> 192 B per row
//        -- exchange: all-gather of the bins --
//   shard_nuc_reduce_kernel  per row: bins summed over the ranks, descending scan, the bin in which the mass
//                            crosses the target -> next prefix; at the last level the boundary value x*, the
//                            number t of entries equal to x* that are kept, the kept mass, and how many of THIS
//                            rank's ties are kept (rank order = index order)
//   shard_nuc_tie_kernel     per row: global index of the rank's last kept tie        -> 4 B per row
//        -- exchange: all-gather of the indices --
//   shard_nuc_close_kernel   per row: ShardCut (x*, i*, kept mass): entry (x, i) is kept iff x > x* or
//                            (x == x* and i <= i*)
//
// Masses are 40-bit fixed point relative to the row maximum (an entry's mass is round(exp(x - M) * 2^40)), summed
// as integers: every sum is exact and independent of the order, of the number of ranks and of the slice
// boundaries - a row has ONE answer for every world size, identical on every rank, from run to run.
// Four levels identify a bf16 value (temperature 1), eight an fp32 one.
//
// The level kernel uses NO atomics: every thread owns a private column of the 16 bins in shared memory
// ([bin][thread], conflict-free read-modify-write, mass and count packed in one 64-bit word).  A first version with
// 256 bins per level and shared-memory atomics took 16.8 ms for the first level of the C5 shape on one GPU (the
// populated bins of the top key byte are few: every atomic of a warp lands on the same two or three addresses).
#pragma once

#include "shard_kernels.cuh"

namespace duodec {

constexpr int kNucBits = 4;                                   // key bits per level
constexpr int kNucBins = 1 << kNucBits;
constexpr int kNucHistBytes = kNucBins * 8 + kNucBins * 4;    // u64 mass[16] | u32 count[16]
constexpr float kNucScale = 1099511627776.0f;                 // 2^40
constexpr int kNucCountShift = 52;                            // private columns: mass | count << 52
constexpr int kNucMaxPerThread = 2047;                        // entries of a row slice per thread (2047 * 2^40 < 2^52)
constexpr size_t kNucLevelSmem = size_t(kNucBins) * kThreads * 8;

struct __align__(16) ShardSel {  // per row, identical on every rank
  unsigned long long above;      // fixed-point mass above the current prefix's key range (closed rows: the kept mass)
  unsigned long long target;     // floor(top_p * total mass)
  unsigned prefix;               // key bits fixed so far (after level l: the top 4 * (l + 1) bits)
  int state;                     // 0 selecting, 1 everything is kept (no crossing), 2 no filter on this row kind
  float M;                       // global row maximum of the scaled logits
  int pad;
};

// (ShardCut, the consumers' view of a row's filter, is declared in shard_kernels.cuh)

__device__ __forceinline__ unsigned long long nuc_mass(float x, float mo) {
  return __float2ull_rn(ex2f(fmaf(x, kShardLog2e, -mo)) * kNucScale);
}
__device__ __forceinline__ unsigned nuc_key(float x) { return okey(x + 0.0f); }  // -0 orders as +0
// the full 32-bit key of a prefix of `bits` bits when the remaining bits are known to be constant (bf16 values:
// zeros below a positive value's 16 bits, ones below a negative one's)
__device__ __forceinline__ unsigned nuc_full_key(unsigned prefix, int bits) {
  if (bits >= 32) return prefix;
  const unsigned low = ((prefix >> (bits - 1)) & 1u) ? 0u : ((1u << (32 - bits)) - 1u);
  return (prefix << (32 - bits)) | low;
}

__device__ __forceinline__ void shard_job(const ShardRows& S, int job, int& b, int& j, int& is_target) {
  const int per_b = 2 * S.gamma + 1;
  b = job / per_b;
  const int k = job - b * per_b;
  is_target = (k == 2 * S.gamma) ? 1 : (k & 1);
  j = (k == 2 * S.gamma) ? S.gamma : (k >> 1);
}

// one block per row
template <class TIn>
__global__ void __launch_bounds__(kThreads) shard_nuc_level_kernel(const __grid_constant__ ShardRows S,
                                                                   const float4* gstats, int G, float top_p_t,
                                                                   float top_p_d, ShardSel* sel, int level,
                                                                   unsigned char* hist) {
  extern __shared__ __align__(16) unsigned long long cols[];  // [bin][thread]
  const int job = blockIdx.x, rows = gridDim.x, tid = threadIdx.x;
  int b, j, is_target;
  shard_job(S, job, b, j, is_target);
  const TIn* row = shard_row<TIn>(S, b, j, is_target);
  const float temp = is_target ? S.t_temp : S.d_temp;
  unsigned long long* col = cols + tid;
#pragma unroll
  for (int i = 0; i < kNucBins; ++i) col[i * kThreads] = 0ull;
  ShardSel st;
  if (level == 0) {
    float M, Z, tx;
    bool has;
    shard_merge(gstats, G, rows, job, M, Z, tx, has);
    st.above = 0ull, st.target = 0ull, st.prefix = 0u, st.M = M, st.pad = 0;
    st.state = (is_target ? top_p_t : top_p_d) > 0.0f ? 0 : 2;
  } else {
    st = sel[job];
  }
  if (st.state == 0) {
    const float mo = st.M * kShardLog2e;
    const int shift = 32 - kNucBits * (level + 1);
    auto body = [&](auto unit_tag) {  // (the temperature test hoisted out of the entry loop, as in shard_stats_kernel)
    constexpr bool kUnit = decltype(unit_tag)::value;
    auto one = [&](float raw) {
      const float x = kUnit ? raw : __fdiv_rn(raw, temp);
      const unsigned key = nuc_key(x);
      // branch-free: an entry outside the prefix adds 0 to its column (a divergent read-modify-write per entry
      // measured 20x slower than the uniform one)
      const bool in = level == 0 || (key >> (shift + kNucBits)) == st.prefix;
      unsigned long long* a = col + ((key >> shift) & unsigned(kNucBins - 1)) * kThreads;
      *a += in ? nuc_mass(x, mo) + (1ull << kNucCountShift) : 0ull;
    };
    if (row_vec8(row, S.Vs)) {
      for (int i = tid * 8; i < S.Vs; i += kThreads * 8) {
        float x[8];
        load8<TIn>(row + i, x);
#pragma unroll
        for (int u = 0; u < 8; ++u) one(x[u]);
      }
    } else {
      for (int i = tid; i < S.Vs; i += kThreads) one(to_f32(row[i]));
    }
    };
    if (temp == 1.0f) body(std::true_type{});
    else body(std::false_type{});
  }
  __syncthreads();
  // warp w sums bin w over the threads (kWarps == kNucBins)
  static_assert(kWarps == kNucBins, "one warp per bin");
  const int w = tid >> 5, lane = tid & 31;
  unsigned long long m = 0ull;
  unsigned c = 0u;
#pragma unroll
  for (int k = 0; k < kThreads / 32; ++k) {
    const unsigned long long v = cols[w * kThreads + k * 32 + lane];
    m += v & ((1ull << kNucCountShift) - 1ull);
    c += unsigned(v >> kNucCountShift);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    m += __shfl_xor_sync(0xffffffffu, m, o);
    c += __shfl_xor_sync(0xffffffffu, c, o);
  }
  if (lane == 0) {
    reinterpret_cast<unsigned long long*>(hist + size_t(job) * kNucHistBytes)[w] = m;
    reinterpret_cast<unsigned*>(hist + size_t(job) * kNucHistBytes + kNucBins * 8)[w] = c;
  }
  if (level == 0 && tid == 0) sel[job] = st;
}

// one warp per row; lane l < 16 owns bin 15 - l (descending key order)
__global__ void __launch_bounds__(128) shard_nuc_reduce_kernel(const unsigned char* ghist, int G, int rows, int gamma,
                                                               float top_p_t, float top_p_d, ShardSel* sel, int level,
                                                               int levels, int rank, int* quota) {
  const int job = blockIdx.x * 4 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (job >= rows) return;
  const bool last = level == levels - 1;
  ShardSel st = sel[job];
  if (st.state != 0) {
    if (last && lane == 0) quota[job] = -1;
    return;
  }
  const int bin = kNucBins - 1 - lane;
  unsigned long long m = 0ull, c = 0ull, before = 0ull, mine = 0ull;
  if (lane < kNucBins) {
    for (int g = 0; g < G; ++g) {
      const unsigned char* h = ghist + (size_t(g) * rows + job) * kNucHistBytes;
      m += reinterpret_cast<const unsigned long long*>(h)[bin];
      const unsigned long long cg = reinterpret_cast<const unsigned*>(h + kNucBins * 8)[bin];
      c += cg;
      if (g < rank) before += cg;
      if (g == rank) mine = cg;
    }
  }
  unsigned long long inc = m;
#pragma unroll
  for (int d = 1; d < kNucBins; d <<= 1) {
    const unsigned long long o = __shfl_up_sync(0xffffffffu, inc, d);
    if (lane >= d) inc += o;
  }
  const unsigned long long total = __shfl_sync(0xffffffffu, inc, kNucBins - 1);
  unsigned long long above = st.above, target = st.target;
  if (level == 0) {
    const int k = job % (2 * gamma + 1);
    const float top_p = (k == 2 * gamma || (k & 1)) ? top_p_t : top_p_d;
    above = 0ull;
    target = top_p >= 1.0f ? total : (unsigned long long)(double(top_p) * double(total));  // floor
  }
  const bool cross = lane < kNucBins && above + inc > target && above + inc - m <= target;  // one bin (m > 0) or none
  const unsigned cm = __ballot_sync(0xffffffffu, cross);
  if (cm == 0u) {  // the mass never exceeds the target (top_p >= 1, an empty row): everything is kept
    if (lane == 0) {
      st.state = 1;
      st.above = above + total;
      st.target = target;
      sel[job] = st;
      if (last) quota[job] = -1;
    }
    return;
  }
  const int src = __ffs(cm) - 1;
  const unsigned long long s_above = __shfl_sync(0xffffffffu, above + inc - m, src);
  const unsigned long long n = __shfl_sync(0xffffffffu, c, src);
  const unsigned long long s_before = __shfl_sync(0xffffffffu, before, src);
  const unsigned long long s_mine = __shfl_sync(0xffffffffu, mine, src);
  if (lane != 0) return;
  st.prefix = (st.prefix << kNucBits) | unsigned(kNucBins - 1 - src);
  st.above = s_above;
  st.target = target;
  if (last) {
    // the bin is ONE value x*: n equal entries of mass f each; the first t of them in index order are kept
    const float xs = okey_inv(nuc_full_key(st.prefix, kNucBits * levels));
    const unsigned long long f = nuc_mass(xs, st.M * kShardLog2e);
    unsigned long long t = f > 0ull ? (target - s_above) / f + 1ull : n;
    if (t > n) t = n;
    st.above = s_above + t * f;  // the kept mass
    long long q = (long long)t - (long long)s_before;  // ranks before this one take their ties first
    q = q < 0 ? 0 : (q > (long long)s_mine ? (long long)s_mine : q);
    quota[job] = t == n ? -1 : int(q);  // -1: every tie is kept
  }
  sel[job] = st;
}

// one block per row: global index of the quota-th entry equal to x* in this rank's slice
template <class TIn>
__global__ void __launch_bounds__(kThreads) shard_nuc_tie_kernel(const __grid_constant__ ShardRows S,
                                                                 const ShardSel* sel, const int* quota, int levels,
                                                                 int* cand) {
  __shared__ Xchg xc;
  __shared__ int s_found;
  const int job = blockIdx.x;
  const ShardSel st = sel[job];
  const int q = quota[job];
  if (st.state != 0 || q <= 0) {  // (uniform over the block)
    if (threadIdx.x == 0) cand[job] = (st.state != 0 || q < 0) ? 0x7fffffff : -1;
    return;
  }
  int b, j, is_target;
  shard_job(S, job, b, j, is_target);
  const TIn* row = shard_row<TIn>(S, b, j, is_target);
  const float temp = is_target ? S.t_temp : S.d_temp;
  const unsigned key = nuc_full_key(st.prefix, kNucBits * levels);
  // threads own contiguous chunks (index order = thread order), read with 128-bit loads where the row allows
  const bool vec = row_vec8(row, S.Vs);
  int chunk = (S.Vs + kThreads - 1) / kThreads;
  chunk = (chunk + 7) & ~7;
  const int i0 = min(int(threadIdx.x) * chunk, S.Vs), i1 = min(i0 + chunk, S.Vs);
  auto is_tie = [&](float raw) { return nuc_key(scale_logit(raw, temp)) == key; };
  int cnt = 0;
  if (vec) {
    for (int i = i0; i < i1; i += 8) {
      float x[8];
      load8<TIn>(row + i, x);
#pragma unroll
      for (int u = 0; u < 8; ++u) cnt += is_tie(x[u]);
    }
  } else {
    for (int i = i0; i < i1; ++i) cnt += is_tie(to_f32(row[i]));
  }
  if (threadIdx.x == 0) s_found = -1;
  int seen = block_exscan(cnt, OpSumI(), 0, &xc);
  if (seen < q && q <= seen + cnt) {
    for (int i = i0; i < i1; ++i)
      if (is_tie(to_f32(row[i])) && ++seen == q) s_found = S.v_begin + i;
  }
  __syncthreads();
  if (threadIdx.x == 0) cand[job] = s_found;
}

__global__ void shard_nuc_close_kernel(const ShardSel* sel, const float4* gstats, const int* gcand, int G, int rows,
                                       int levels, ShardCut* cut) {
  const int job = blockIdx.x * blockDim.x + threadIdx.x;
  if (job >= rows) return;
  const ShardSel st = sel[job];
  ShardCut c;
  c.xstar = -INFINITY, c.istar = 0x7fffffff, c.state = st.state;
  if (st.state == 2) {
    float M, Z, tx;
    bool has;
    shard_merge(gstats, G, rows, job, M, Z, tx, has);
    c.Zk = Z;
  } else {
    c.Zk = float(double(st.above) * (1.0 / double(kNucScale)));
    if (st.state == 0) {
      int is = -1;
      for (int g = 0; g < G; ++g) is = max(is, gcand[(long long)g * rows + job]);
      c.xstar = okey_inv(nuc_full_key(st.prefix, kNucBits * levels));
      c.istar = is;
    }
  }
  cut[job] = c;
}

}  // namespace duodec
