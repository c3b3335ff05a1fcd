// Building blocks of the persistent "row stream" verification kernel: the
// consumer-group barrier / reductions and the TMA chunk ring constants.
#pragma once

#include "rowcore.cuh"

namespace duodec {

constexpr int kCT = 512;            // consumer threads of a persistent block (two blocks per SM)
constexpr int kCW = kCT / 32;       // consumer warps
constexpr int kST = kCT + 32;       // + one producer warp that only issues TMA bulk copies
constexpr int kChunk = 2048;        // row elements per ring slot: one float4 per consumer thread
#ifndef DUODEC_SLOTS
#define DUODEC_SLOTS 8
#endif
#ifndef DUODEC_COLCAP
#define DUODEC_COLCAP 8
#endif
constexpr int kSlots = DUODEC_SLOTS; // ring depth (x 8 KB)
constexpr int kSlotBytes = kChunk * 4;
constexpr int kColCap = DUODEC_COLCAP; // window entries a thread keeps in its own shared-memory column
constexpr int kOvfCap = 256;        // shared overflow list for threads that see more
constexpr int kCandMax = 256;       // entries the all-pairs finish can order
constexpr float kCandGoal = 160.0f; // narrowing stops once the bracket provably holds at most this many
constexpr float kLog2e = 1.4426950408889634f;

// named barrier of the 512 consumer threads (the producer warp never joins it)
__device__ __forceinline__ void cbar() { asm volatile("bar.sync 1, 512;" ::: "memory"); }

struct __align__(16) R8 {  // reduction payload: v[0] max-reduced, v[1..7] summed
  float v[8];
};

struct __align__(16) StreamScratch {
  unsigned long long full[kSlots];   // "chunk landed" (TMA complete_tx)
  unsigned long long empty[kSlots];  // "chunk consumed" (one arrive per consumer warp)
  R8 red[2][kCW];                    // ping-pong reduction scratch
  float ovf_x[kOvfCap];
  int ovf_i[kOvfCap];
  float cand_e[kCandMax];
  float cand_x[kCandMax];
  int cand_i[kCandMax];
  int novf[2], ncand, pad0;
  float res_x;
  int res_i;
  float res_Z;
  int res_ok;
  int seq_last;        // 1: this block finished the last row of its sequence
  int seq_acc[64];     // accept bits of the sequence being closed
  float hint_m0, hint_th, hint_dens;  // prediction for the NEXT row job of this block
  int hint_ok;
};

// All-reduce of an R8 over the 512 consumer threads with ONE named barrier; the
// scratch is ping-ponged so back-to-back calls need no extra barrier.
__device__ __forceinline__ R8 creduce(R8 a, StreamScratch* ss, int& phase) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    a.v[0] = fmaxf(a.v[0], __shfl_xor_sync(0xffffffffu, a.v[0], o));
#pragma unroll
    for (int i = 1; i < 8; ++i) a.v[i] += __shfl_xor_sync(0xffffffffu, a.v[i], o);
  }
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  if (lane == 0) ss->red[phase][w] = a;
  cbar();
  R8 x = ss->red[phase][lane & (kCW - 1)];
  phase ^= 1;
#pragma unroll
  for (int o = kCW / 2; o > 0; o >>= 1) {
    x.v[0] = fmaxf(x.v[0], __shfl_xor_sync(0xffffffffu, x.v[0], o));
#pragma unroll
    for (int i = 1; i < 8; ++i) x.v[i] += __shfl_xor_sync(0xffffffffu, x.v[i], o);
  }
  return x;
}

// generic small all-reduce (<= 16 bytes) over the consumer threads, same protocol
template <class T, class Op>
__device__ __forceinline__ T creduce_small(T v, Op op, StreamScratch* ss, int& phase) {
  static_assert(sizeof(T) <= 16 && sizeof(T) % 4 == 0, "");
  v = warp_allreduce(v, op);
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  if (lane == 0) *reinterpret_cast<T*>(&ss->red[phase][w]) = v;
  cbar();
  T x = *reinterpret_cast<T*>(&ss->red[phase][lane & (kCW - 1)]);
  phase ^= 1;
#pragma unroll
  for (int o = kCW / 2; o > 0; o >>= 1) x = op(x, shfl_xor_any(x, o));
  return x;
}

__device__ __forceinline__ void mbar_arrive(unsigned long long* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}

}  // namespace duodec
