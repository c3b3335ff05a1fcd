// Block scans, team histogram reduction and the count-based k-th-largest
// select used by the top-k warp and the top-k compression ops.
#pragma once

#include "common.cuh"

namespace duodec {

template <class T>
__device__ __forceinline__ T shfl_up_any(T v, int delta) {
  static_assert(sizeof(T) % 4 == 0 && sizeof(T) <= 16, "");
  union {
    T t;
    unsigned u[sizeof(T) / 4];
  } a;
  a.t = v;
#pragma unroll
  for (int i = 0; i < int(sizeof(T) / 4); ++i) a.u[i] = __shfl_up_sync(0xffffffffu, a.u[i], delta);
  return a.t;
}

// Exclusive prefix "sum" over the threads of a block in thread order.
// Returns the exclusive prefix of this thread; *total (if given) gets the
// block total.  Deterministic.
template <class T, class Op>
__device__ __forceinline__ T block_exscan(T v, Op op, T identity, Xchg* xc, T* total = nullptr) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  T inc = v;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    T o = shfl_up_any(inc, d);
    if (lane >= d) inc = op(o, inc);
  }
  __syncthreads();
  if (lane == 31) *reinterpret_cast<T*>(&xc->warp[w]) = inc;
  __syncthreads();
  T base = identity;
  T tot = identity;
#pragma unroll
  for (int i = 0; i < kWarps; ++i) {
    T wv = *reinterpret_cast<T*>(&xc->warp[i]);
    if (i < w) base = op(base, wv);
    tot = op(tot, wv);
  }
  if (total) *total = tot;
  T ex = shfl_up_any(inc, 1);
  if (lane == 0) ex = identity;
  return op(base, ex);
}

// In-place all-reduce (sum) of an integer histogram that every block of the
// team filled for its own slice.  Step 1: rank c sums range c over all peers
// into its own copy; step 2: every block fetches the other ranges from their
// owners.  After the call every block holds the full team histogram.
template <int C>
__device__ __forceinline__ void team_hist_allreduce(int* h, int nb) {
  if constexpr (C == 1) {
    __syncthreads();
  } else {
    const unsigned me = Team<C>::rank();
    const int per = nb / C;  // nb is a multiple of 8*C in every use
    Team<C>::sync();         // all local histograms complete
    for (int i = threadIdx.x; i < per; i += kThreads) {
      const int b = int(me) * per + i;
      int acc = 0;
#pragma unroll
      for (int r = 0; r < C; ++r) acc += Team<C>::peer(h, r)[b];
      h[b] = acc;
    }
    Team<C>::sync();
    for (int i = threadIdx.x; i < nb; i += kThreads) {
      const unsigned owner = unsigned(i / per);
      if (owner != me) h[i] = Team<C>::peer(h, owner)[i];
    }
    Team<C>::sync();
  }
}

// Given a histogram h[0..nb) (nb a multiple of kThreads) find the digit d such
// that  sum_{b>d} h[b] < kk <= sum_{b>=d} h[b]  (scanning from the top bin).
// res[0] = d, res[1] = sum_{b>d} h[b]; res must be block-shared memory.
__device__ __forceinline__ void block_find_kth_desc(const int* h, int nb, int kk, int* res, Xchg* xc) {
  const int per = nb / kThreads;
  const int top = nb - 1 - int(threadIdx.x) * per;  // this thread walks top, top-1, ...
  int local = 0;
  for (int i = 0; i < per; ++i) local += h[top - i];
  int above = block_exscan(local, OpSumI(), 0, xc);
  for (int i = 0; i < per; ++i) {
    const int c = h[top - i];
    if (above < kk && kk <= above + c) {
      res[0] = top - i;
      res[1] = above;
    }
    above += c;
  }
  __syncthreads();
}

// k-th largest value (1-based k, 1 <= k <= number of entries) of a team-staged
// row xs (n4 float4 granules per block), by a 3-pass radix select over the
// order-preserving key with integer shared-memory atomics.  hist: >= 2048 ints.
// Also returns how many entries are strictly greater than the result.
struct KthResult {
  float value;
  int greater;
};
template <int C>
__device__ __forceinline__ KthResult team_kth_largest(const float* xs, int n4, int k, int* hist, int* res,
                                                       Xchg* xc) {
  unsigned prefix = 0u, pmask = 0u;
  int kk = k;
  int greater = 0;
  const int shifts[3] = {21, 10, 0};
  const int widths[3] = {11, 11, 10};
#pragma unroll 1
  for (int pass = 0; pass < 3; ++pass) {
    const int shift = shifts[pass];
    const int nb = 1 << widths[pass];
    __syncthreads();
    for (int i = threadIdx.x; i < nb; i += kThreads) hist[i] = 0;
    __syncthreads();
    const float4* x4 = reinterpret_cast<const float4*>(xs);
    for (int i = threadIdx.x; i < n4; i += kThreads) {
      const float4 v = x4[i];
      const float e[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        const unsigned key = okey(e[c]);
        if ((key & pmask) == prefix) atomicAdd(&hist[(key >> shift) & unsigned(nb - 1)], 1);
      }
    }
    __syncthreads();
    team_hist_allreduce<C>(hist, nb);
    block_find_kth_desc(hist, nb, kk, res, xc);
    const int d = res[0];
    greater += res[1];
    kk -= res[1];
    prefix |= unsigned(d) << shift;
    pmask |= unsigned(nb - 1) << shift;
  }
  return KthResult{okey_inv(prefix), greater};
}

}  // namespace duodec
