// Building blocks of the persistent "row stream" verification kernel:
// 1024-thread block reductions / scans and the TMA chunk ring.
#pragma once

#include "rowcore.cuh"

namespace duodec {

constexpr int kST = 1024;          // threads of a persistent block (one block per SM)
constexpr int kSW = kST / 32;      // warps
constexpr int kChunk = 4096;       // row elements per ring slot: one float4 per thread
constexpr int kMaxChunks = 8;      // rows up to 32768 entries live in 32 registers per thread
constexpr int kSlots = 13;         // ring depth (13 x 16 KB)
constexpr int kSlotBytes = kChunk * 4;

struct __align__(16) StreamScratch {
  unsigned long long full[kSlots];  // "chunk landed" mbarriers
  uint4 red[2][kSW];                // ping-pong reduction scratch
  int red_phase;
  int bstar, ncand, jstar, overflow;
  int flag;
  double m_above;
  unsigned long long keystar;
};

// All-reduce over the 1024 threads of the block with ONE __syncthreads: every
// warp reduces, lane 0 publishes, then every warp reduces the 32 partials again.
// The scratch is ping-ponged so back-to-back calls need no extra barrier.
template <class T, class Op>
__device__ __forceinline__ T sblock_allreduce(T v, Op op, StreamScratch* ss, int& phase) {
  static_assert(sizeof(T) <= 16 && sizeof(T) % 4 == 0, "");
  v = warp_allreduce(v, op);
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  if (lane == 0) *reinterpret_cast<T*>(&ss->red[phase][w]) = v;
  __syncthreads();
  T x = *reinterpret_cast<T*>(&ss->red[phase][lane]);
  phase ^= 1;
  return warp_allreduce(x, op);
}

// Exclusive scan over the 1024 threads (thread order); also returns the total.
template <class T, class Op>
__device__ __forceinline__ T sblock_exscan(T v, Op op, T identity, StreamScratch* ss, int& phase, T* total) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  T inc = v;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    T o = shfl_up_any(inc, d);
    if (lane >= d) inc = op(o, inc);
  }
  if (lane == 31) *reinterpret_cast<T*>(&ss->red[phase][w]) = inc;
  __syncthreads();
  // every warp scans the 32 warp totals
  T wt = *reinterpret_cast<T*>(&ss->red[phase][lane]);
  phase ^= 1;
  T winc = wt;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    T o = shfl_up_any(winc, d);
    if (lane >= d) winc = op(o, winc);
  }
  union {
    T t;
    unsigned u[sizeof(T) / 4];
  } a, b;
  a.t = winc;
#pragma unroll
  for (int i = 0; i < int(sizeof(T) / 4); ++i) b.u[i] = __shfl_sync(0xffffffffu, a.u[i], 31);
  if (total) *total = b.t;
  // exclusive prefix of this warp = inclusive prefix of warp w-1
  T wex = shfl_up_any(winc, 1);
#pragma unroll
  for (int i = 0; i < int(sizeof(T) / 4); ++i) {
    a.t = wex;
    b.u[i] = __shfl_sync(0xffffffffu, a.u[i], w);
  }
  T base = (w == 0) ? identity : b.t;
  T ex = shfl_up_any(inc, 1);
  if (lane == 0) ex = identity;
  return op(base, ex);
}

}  // namespace duodec
