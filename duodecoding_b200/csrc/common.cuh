// Shared device utilities for the verification hot path (sm_100a).
//
// Execution model: one TEAM of C thread blocks (a thread-block cluster, C in
// {1,2,4,8}) owns one vocabulary row (or one draft/target row pair).  Every
// block of the team stages a contiguous slice of the row in shared memory and
// all passes after the first run on shared memory only; scalars are combined
// across the team through distributed shared memory (DSMEM).
#pragma once

#include <cooperative_groups.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace cg = cooperative_groups;

namespace duodec {

constexpr int kThreads = 512;  // threads per block (2 blocks/SM at <=64 regs)
constexpr int kWarps = kThreads / 32;
constexpr int kMaxTeam = 8;  // portable cluster limit

// ---------------------------------------------------------------------------
// team (cluster) helpers
// ---------------------------------------------------------------------------
template <int C>
struct Team {
  static __device__ __forceinline__ unsigned rank() {
    if constexpr (C == 1) return 0u;
    else return cg::this_cluster().block_rank();
  }
  static __device__ __forceinline__ unsigned id() {  // index of the team in the grid
    return blockIdx.x / C;
  }
  static __device__ __forceinline__ void sync() {
    if constexpr (C == 1) __syncthreads();
    else cg::this_cluster().sync();
  }
  template <class T>
  static __device__ __forceinline__ T* peer(T* p, unsigned r) {
    if constexpr (C == 1) return p;
    else return cg::this_cluster().map_shared_rank(p, r);
  }
};

// Small per-block exchange area.  slot[] is double buffered so that one team
// sync per reduction is enough (a slot is rewritten two reductions later, by
// which time every peer has passed the sync in between).
struct __align__(16) Xchg {
  uint4 slot[2];
  uint4 gather[kMaxTeam];
  uint4 warp[kWarps];
  int phase;
  int pad[3];
};

template <class T>
__device__ __forceinline__ T shfl_xor_any(T v, int lane_mask) {
  static_assert(sizeof(T) % 4 == 0 && sizeof(T) <= 16, "4..16 byte payloads");
  union {
    T t;
    unsigned u[sizeof(T) / 4];
  } a;
  a.t = v;
#pragma unroll
  for (int i = 0; i < int(sizeof(T) / 4); ++i) a.u[i] = __shfl_xor_sync(0xffffffffu, a.u[i], lane_mask);
  return a.t;
}

template <class T, class Op>
__device__ __forceinline__ T warp_allreduce(T v, Op op) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = op(v, shfl_xor_any(v, o));
  return v;
}

// Block all-reduce; every thread gets the result.  Deterministic: fixed tree.
template <class T, class Op>
__device__ __forceinline__ T block_allreduce(T v, Op op, Xchg* xc) {
  static_assert(sizeof(T) <= 16, "");
  v = warp_allreduce(v, op);
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  __syncthreads();  // protect xc->warp against the previous use
  if (lane == 0) *reinterpret_cast<T*>(&xc->warp[w]) = v;
  __syncthreads();
  T acc = *reinterpret_cast<T*>(&xc->warp[0]);
#pragma unroll
  for (int i = 1; i < kWarps; ++i) acc = op(acc, *reinterpret_cast<T*>(&xc->warp[i]));
  return acc;
}

// Team all-reduce (rank order, identical result in every block of the team).
template <int C, class T, class Op>
__device__ __forceinline__ T team_allreduce(T v, Op op, Xchg* xc) {
  T b = block_allreduce(v, op, xc);
  if constexpr (C == 1) {
    return b;
  } else {
    const int ph = xc->phase;
    __syncthreads();
    if (threadIdx.x == 0) {
      *reinterpret_cast<T*>(&xc->slot[ph]) = b;
      xc->phase = ph ^ 1;
    }
    Team<C>::sync();
    if (threadIdx.x < C)
      *reinterpret_cast<T*>(&xc->gather[threadIdx.x]) =
          *reinterpret_cast<T*>(&Team<C>::peer(xc, threadIdx.x)->slot[ph]);
    __syncthreads();
    T acc = *reinterpret_cast<T*>(&xc->gather[0]);
#pragma unroll
    for (int i = 1; i < C; ++i) acc = op(acc, *reinterpret_cast<T*>(&xc->gather[i]));
    return acc;
  }
}

// Broadcast a value computed by team rank 0 to the whole team.
template <int C, class T>
__device__ __forceinline__ T team_bcast0(T v, Xchg* xc) {
  static_assert(sizeof(T) <= 16, "");
  const int ph = xc->phase;
  __syncthreads();
  if (threadIdx.x == 0) {
    *reinterpret_cast<T*>(&xc->slot[ph]) = v;
    xc->phase = ph ^ 1;
  }
  Team<C>::sync();
  if (threadIdx.x == 0)
    *reinterpret_cast<T*>(&xc->gather[0]) = *reinterpret_cast<T*>(&Team<C>::peer(xc, 0)->slot[ph]);
  __syncthreads();
  return *reinterpret_cast<T*>(&xc->gather[0]);
}

struct OpMaxF {
  __device__ __forceinline__ float operator()(float a, float b) const { return fmaxf(a, b); }
};
struct OpSumF {
  __device__ __forceinline__ float operator()(float a, float b) const { return a + b; }
};
struct OpSumU64 {
  __device__ __forceinline__ unsigned long long operator()(unsigned long long a, unsigned long long b) const { return a + b; }
};
struct OpSumD {
  __device__ __forceinline__ double operator()(double a, double b) const { return a + b; }
};
struct OpSumI {
  __device__ __forceinline__ int operator()(int a, int b) const { return a + b; }
};
struct OpMinI {
  __device__ __forceinline__ int operator()(int a, int b) const { return a < b ? a : b; }
};
struct OpMaxI {
  __device__ __forceinline__ int operator()(int a, int b) const { return a > b ? a : b; }
};
struct OpMaxU64 {
  __device__ __forceinline__ unsigned long long operator()(unsigned long long a, unsigned long long b) const {
    return a > b ? a : b;
  }
};
struct OpMinU64 {
  __device__ __forceinline__ unsigned long long operator()(unsigned long long a, unsigned long long b) const {
    return a < b ? a : b;
  }
};

// (value, index) with "larger value wins, lower index on ties" == torch.argmax.
struct ValIdx {
  float v;
  int i;
};
struct OpArgMax {
  __device__ __forceinline__ ValIdx operator()(ValIdx a, ValIdx b) const {
    if (b.v > a.v || (b.v == a.v && b.i < a.i)) return b;
    return a;
  }
};
struct F2 {
  float a, b;
};
struct OpSumF2 {
  __device__ __forceinline__ F2 operator()(F2 x, F2 y) const { return F2{x.a + y.a, x.b + y.b}; }
};
struct D2 {
  double a, b;
};
struct OpSumD2 {
  __device__ __forceinline__ D2 operator()(D2 x, D2 y) const { return D2{x.a + y.a, x.b + y.b}; }
};

// Order-preserving map float -> uint32 (larger float -> larger key; -0 < +0).
__device__ __forceinline__ unsigned okey(float f) {
  unsigned u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float okey_inv(unsigned k) {
  unsigned u = (k & 0x80000000u) ? (k & 0x7fffffffu) : ~k;
  return __uint_as_float(u);
}

// ---------------------------------------------------------------------------
// input element access
// ---------------------------------------------------------------------------
template <class T>
__device__ __forceinline__ float to_f32(T v);
template <>
__device__ __forceinline__ float to_f32<float>(float v) {
  return v;
}
template <>
__device__ __forceinline__ float to_f32<__nv_bfloat16>(__nv_bfloat16 v) {
  return __bfloat162float(v);
}

template <class T>
__device__ __forceinline__ T from_f32(float v);
template <>
__device__ __forceinline__ float from_f32<float>(float v) {
  return v;
}
template <>
__device__ __forceinline__ __nv_bfloat16 from_f32<__nv_bfloat16>(float v) {
  return __float2bfloat16_rn(v);
}

// logits / temperature exactly as the reference issues it (src/utils.py:706):
// a true IEEE division, skipped when T == 1.
__device__ __forceinline__ float scale_logit(float l, float temperature) {
  return temperature == 1.0f ? l : __fdiv_rn(l, temperature);
}

// ---------------------------------------------------------------------------
// TMA 1-D bulk copy global -> shared (cp.async.bulk, SASS UBLKCP) + mbarrier
// ---------------------------------------------------------------------------
__device__ __forceinline__ unsigned smem_u32(const void* p) {
  return static_cast<unsigned>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(unsigned long long* bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long* bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, unsigned bytes, unsigned long long* bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
          smem_u32(dst)),
      "l"(src), "r"(bytes), "r"(smem_u32(bar))
      : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, unsigned parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}

// Slice of a row owned by one block of a team: elements [g0, g0+n) of a row of
// V entries, split in 8-element granules so 16-byte accesses stay aligned for
// both fp32 and bf16 rows.
struct Slice {
  int g0;
  int n;
};
template <int C>
__device__ __forceinline__ int team_per(int V) {
  if constexpr (C == 1) return V;
  else return (((V + C - 1) / C) + 7) & ~7;
}
// `per` = team_per<C>(widest row handled by the team): with a common `per` the
// draft and the target row of a pair are cut at the same places, so entry i of
// both rows lives in the same block at the same offset.
__device__ __forceinline__ Slice team_slice(int per, int V, unsigned rank) {
  int g0 = per * int(rank);
  if (g0 > V) g0 = V;
  int g1 = g0 + per;
  if (g1 > V) g1 = V;
  return Slice{g0, g1 - g0};
}

// Stage row[g0 .. g0+n) into shared memory as fp32 scaled logits, padded with
// -inf up to a multiple of 4 so every later pass can run on float4.
//   fp32 input, T == 1, 16-byte aligned: one elected thread issues TMA bulk
//   copies that complete on an mbarrier (stage_begin returns immediately and
//   stage_end waits, so two rows can be in flight at once); everything else
//   uses 128-bit (or scalar, when the row base is misaligned) loads.
template <class TIn>
__device__ __forceinline__ bool stage_begin(float* xs, const TIn* __restrict__ row, Slice s, float temperature,
                                            unsigned long long* bar) {
  const TIn* src = row + s.g0;
  const int n = s.n;
  bool async = false;
  const bool aligned = ((reinterpret_cast<uintptr_t>(src) & 15) == 0);
  if (sizeof(TIn) == 4 && aligned && temperature == 1.0f) {
    const int nb = (n & ~3) * 4;  // bytes moved by the bulk engine
    if (nb > 0) {
      async = true;
      if (threadIdx.x == 0) {
        mbar_expect_tx(bar, unsigned(nb));
        constexpr int kChunk = 32768;
        for (int off = 0; off < nb; off += kChunk) {
          int c = nb - off < kChunk ? nb - off : kChunk;
          bulk_g2s(reinterpret_cast<char*>(xs) + off, reinterpret_cast<const char*>(src) + off, unsigned(c), bar);
        }
      }
    }
    for (int i = (n & ~3) + threadIdx.x; i < n; i += kThreads) xs[i] = to_f32(src[i]);
  } else if (sizeof(TIn) == 4 && aligned) {
    const float4* s4 = reinterpret_cast<const float4*>(src);
    float4* d4 = reinterpret_cast<float4*>(xs);
    for (int i = threadIdx.x; i < (n >> 2); i += kThreads) {
      float4 v = __ldg(s4 + i);
      v.x = scale_logit(v.x, temperature);
      v.y = scale_logit(v.y, temperature);
      v.z = scale_logit(v.z, temperature);
      v.w = scale_logit(v.w, temperature);
      d4[i] = v;
    }
    for (int i = (n & ~3) + threadIdx.x; i < n; i += kThreads) xs[i] = scale_logit(to_f32(src[i]), temperature);
  } else if (sizeof(TIn) == 2 && aligned) {
    const uint4* s8 = reinterpret_cast<const uint4*>(src);
    for (int i = threadIdx.x; i < (n >> 3); i += kThreads) {
      uint4 v = __ldg(s8 + i);
      float4 a, b;
      a.x = scale_logit(__uint_as_float(v.x << 16), temperature);
      a.y = scale_logit(__uint_as_float(v.x & 0xffff0000u), temperature);
      a.z = scale_logit(__uint_as_float(v.y << 16), temperature);
      a.w = scale_logit(__uint_as_float(v.y & 0xffff0000u), temperature);
      b.x = scale_logit(__uint_as_float(v.z << 16), temperature);
      b.y = scale_logit(__uint_as_float(v.z & 0xffff0000u), temperature);
      b.z = scale_logit(__uint_as_float(v.w << 16), temperature);
      b.w = scale_logit(__uint_as_float(v.w & 0xffff0000u), temperature);
      reinterpret_cast<float4*>(xs)[2 * i] = a;
      reinterpret_cast<float4*>(xs)[2 * i + 1] = b;
    }
    for (int i = (n & ~7) + threadIdx.x; i < n; i += kThreads) xs[i] = scale_logit(to_f32(src[i]), temperature);
  } else {
    for (int i = threadIdx.x; i < n; i += kThreads) xs[i] = scale_logit(to_f32(src[i]), temperature);
  }
  if (threadIdx.x < 4) {
    const int i = n + threadIdx.x;
    if (i < ((n + 3) & ~3)) xs[i] = -INFINITY;
  }
  return async;
}

// Wait for the bulk copies of stage_begin (if any) and make the slice visible
// to the whole block.
__device__ __forceinline__ void stage_end(bool async, unsigned long long* bar, unsigned& bar_parity) {
  if (async) {
    mbar_wait(bar, bar_parity);
    bar_parity ^= 1u;
  }
  __syncthreads();
}

}  // namespace duodec
