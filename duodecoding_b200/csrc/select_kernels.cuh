// transfer_top_k for rows that fit one SM's shared memory: duodec_topk_rebuild's default kernel.
//
// The reference sorts every row (torch.topk, /root/reference/src/utils.py:751-784 and
// src/communication.py:288-395) and then makes ~8 passes over it.  Here one block of 1024 threads owns
// a row: the row is read ONCE from global memory into shared memory (as fp32), the k-th largest value is
// found exactly by a three-level radix select over the ordered 32-bit keys (11 + 11 + 10 bits, histograms
// in shared-memory atomics, one block scan per level), ties at the k-th value are cut lowest index first
// (the oracle's canonical order), and the output row is written once.  Algorithmic traffic: one read and
// one write of the row.  The team kernel (row_kernels.cuh) remains for rows that do not fit and for the
// forced team sizes of the tests.
#pragma once

#include "row_kernels.cuh"

namespace duodec {
namespace rowsel {

constexpr int kRT = 1024;     // threads per block
constexpr int kRW = kRT / 32;
constexpr int kBins = 2048;

struct __align__(16) SelScratch {
  unsigned hist[kBins];
  unsigned wsum[kRW];
  float wf[kRW];
  int wi[kRW];
  int bin;
  unsigned above;
  int istar;
  int pad;
};

__host__ __device__ constexpr size_t select_smem_bytes(int V) {
  return size_t((V + 3) & ~3) * 4 + sizeof(SelScratch) + 16;
}

__device__ __forceinline__ unsigned sel_key(float v) { return okey(v + 0.0f); }  // -0 orders as +0

// inclusive scan of one unsigned per thread over the block (thread order); total in *total
__device__ __forceinline__ unsigned block_incl_scan(unsigned v, SelScratch* sc, unsigned* total) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  unsigned inc = v;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const unsigned o = __shfl_up_sync(0xffffffffu, inc, d);
    if (lane >= d) inc += o;
  }
  if (lane == 31) sc->wsum[w] = inc;
  __syncthreads();
  unsigned ws = sc->wsum[lane];  // kRW == 32: one warp sum per lane
  unsigned winc = ws;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const unsigned o = __shfl_up_sync(0xffffffffu, winc, d);
    if (lane >= d) winc += o;
  }
  *total = __shfl_sync(0xffffffffu, winc, 31);
  const unsigned before = __shfl_sync(0xffffffffu, winc - ws, w);
  __syncthreads();  // wsum may be rewritten by the next scan
  return inc + before;
}

// The bin (of `nbins` <= 2048, counted from the top) in which the running count reaches `need` (>= 1),
// and the count of the bins above it.  Thread t owns bins nbins-1-2t and nbins-2-2t.
__device__ __forceinline__ void find_bin(SelScratch* sc, int nbins, unsigned need, int* bin, unsigned* above) {
  const int hi = nbins - 1 - 2 * int(threadIdx.x), lo = hi - 1;
  const unsigned ch = hi >= 0 ? sc->hist[hi] : 0u, cl = lo >= 0 ? sc->hist[lo] : 0u;
  unsigned total;
  const unsigned incl = block_incl_scan(ch + cl, sc, &total);
  const unsigned excl = incl - (ch + cl);
  if (excl < need && need <= incl) {
    if (excl + ch >= need) sc->bin = hi, sc->above = excl;
    else sc->bin = lo, sc->above = excl + ch;
  }
  if (threadIdx.x == 0 && total < need) sc->bin = -1;  // fewer than `need` entries: cannot happen for k < V
  __syncthreads();
  *bin = sc->bin;
  *above = sc->above;
  __syncthreads();
}

__device__ __forceinline__ float block_sum_f(float v, SelScratch* sc) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  if ((threadIdx.x & 31) == 0) sc->wf[threadIdx.x >> 5] = v;
  __syncthreads();
  float s = sc->wf[threadIdx.x & 31];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  __syncthreads();
  return s;
}
__device__ __forceinline__ int block_sum_i(int v, SelScratch* sc) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  if ((threadIdx.x & 31) == 0) sc->wi[threadIdx.x >> 5] = v;
  __syncthreads();
  int s = sc->wi[threadIdx.x & 31];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  __syncthreads();
  return s;
}

template <class T>
__global__ void __launch_bounds__(kRT, 1) topk_select_kernel(const __grid_constant__ TopkParams P) {
  extern __shared__ __align__(128) unsigned char smem[];
  float* xs = reinterpret_cast<float*>(smem);
  const int V = P.V, Vp = (V + 3) & ~3;
  SelScratch* sc = reinterpret_cast<SelScratch*>(xs + Vp);
  const int tid = threadIdx.x;
  const long long rowi = blockIdx.x;
  const T* row = reinterpret_cast<const T*>(P.probs) + rowi * P.in_stride;
  T* out = reinterpret_cast<T*>(P.out) + rowi * P.out_stride;
  const bool select = P.mode != 3;

  for (int i = tid; i < kBins; i += kRT) sc->hist[i] = 0u;
  __syncthreads();

  // ---- the row, once: global -> shared (fp32) + the first-level histogram (top 11 key bits)
  constexpr int kPer = 16 / int(sizeof(T));
  const bool vec = (reinterpret_cast<size_t>(row) & 15) == 0;
  if (vec) {
    const int nv = V / kPer;
    for (int g = tid; g < nv; g += kRT) {
      const uint4 raw = __ldg(reinterpret_cast<const uint4*>(row) + g);
      float v[kPer];
      if constexpr (sizeof(T) == 4) {
        v[0] = __uint_as_float(raw.x), v[1] = __uint_as_float(raw.y);
        v[2] = __uint_as_float(raw.z), v[3] = __uint_as_float(raw.w);
      } else {
        const unsigned w[4] = {raw.x, raw.y, raw.z, raw.w};
#pragma unroll
        for (int q = 0; q < 4; ++q) v[2 * q] = __uint_as_float(w[q] << 16), v[2 * q + 1] = __uint_as_float(w[q] & 0xffff0000u);
      }
#pragma unroll
      for (int q = 0; q < kPer; ++q) {
        xs[g * kPer + q] = v[q];
        if (select) atomicAdd(&sc->hist[sel_key(v[q]) >> 21], 1u);
      }
    }
    for (int i = nv * kPer + tid; i < V; i += kRT) {
      const float v = to_f32(row[i]);
      xs[i] = v;
      if (select) atomicAdd(&sc->hist[sel_key(v) >> 21], 1u);
    }
  } else {
    for (int i = tid; i < V; i += kRT) {
      const float v = to_f32(row[i]);
      xs[i] = v;
      if (select) atomicAdd(&sc->hist[sel_key(v) >> 21], 1u);
    }
  }
  __syncthreads();

  unsigned tau = 0u;        // key of the k-th largest entry
  int istar = 0x7fffffff;   // entries equal to it are kept up to this row index
  if (select) {
    int b1, b2, b3;
    unsigned a1, a2, a3;
    find_bin(sc, kBins, unsigned(P.k), &b1, &a1);
    for (int i = tid; i < kBins; i += kRT) sc->hist[i] = 0u;
    __syncthreads();
    for (int i = tid; i < V; i += kRT) {
      const unsigned key = sel_key(xs[i]);
      if (int(key >> 21) == b1) atomicAdd(&sc->hist[(key >> 10) & 2047u], 1u);
    }
    __syncthreads();
    find_bin(sc, kBins, unsigned(P.k) - a1, &b2, &a2);
    for (int i = tid; i < kBins; i += kRT) sc->hist[i] = 0u;
    __syncthreads();
    const unsigned top22 = (unsigned(b1) << 11) | unsigned(b2);
    for (int i = tid; i < V; i += kRT) {
      const unsigned key = sel_key(xs[i]);
      if ((key >> 10) == top22) atomicAdd(&sc->hist[key & 1023u], 1u);
    }
    __syncthreads();
    const unsigned need3 = unsigned(P.k) - a1 - a2;
    find_bin(sc, 1024, need3, &b3, &a3);
    tau = (top22 << 10) | unsigned(b3);
    const unsigned ties = sc->hist[b3], take = need3 - a3;  // 1 <= take <= ties
    if (take < ties) {
      // lowest index first: the take-th entry equal to tau in index order (threads own contiguous chunks)
      const int chunk = (V + kRT - 1) / kRT;
      const int i0 = tid * chunk, i1 = min(i0 + chunk, V);
      unsigned cnt = 0u;
      for (int i = i0; i < i1; ++i) cnt += (sel_key(xs[i]) == tau);
      unsigned total;
      const unsigned incl = block_incl_scan(cnt, sc, &total);
      unsigned seen = incl - cnt;
      if (seen < take && take <= incl) {
        for (int i = i0; i < i1; ++i) {
          if (sel_key(xs[i]) == tau && ++seen == take) sc->istar = i;
        }
      }
      __syncthreads();
      istar = sc->istar;
    }
  }
  auto kept = [&](float v, int i) {
    const unsigned key = sel_key(v);
    return !select || key > tau || (key == tau && i <= istar);
  };

  if (P.mode == 1) {
    for (int i = tid; i < V; i += kRT) {
      const float v = xs[i];
      out[i] = from_f32<T>(kept(v, i) ? v : 0.0f);
    }
    return;
  }
  // kept mass and the number of exact zeros after the scatter (fixed reduction order: deterministic)
  float ksum = 0.0f;
  int holes = 0;
  for (int i = tid; i < V; i += kRT) {
    const float v = xs[i];
    const float kv = kept(v, i) ? v : 0.0f;
    ksum += kv;
    holes += (kv == 0.0f);
  }
  ksum = block_sum_f(ksum, sc);
  holes = block_sum_i(holes, sc);
  float tail = __fsub_rn(1.0f, ksum);
  if (P.mode == 0) tail = fmaxf(tail, 0.0f);
  const float fill = holes > 0 ? __fdiv_rn(tail, float(holes)) : 0.0f;
  float total = 1.0f;
  if (P.mode == 0) {
    float tsum = 0.0f;
    for (int i = tid; i < V; i += kRT) {
      const float v = xs[i];
      const float kv = kept(v, i) ? v : 0.0f;
      tsum += (kv == 0.0f) ? fill : kv;
    }
    total = fmaxf(block_sum_f(tsum, sc), 1e-12f);
  }
  for (int i = tid; i < V; i += kRT) {
    const float v = xs[i];
    const float kv = kept(v, i) ? v : 0.0f;
    float o = (kv == 0.0f) ? fill : kv;
    if (P.mode == 0) o = __fdiv_rn(o, total);
    out[i] = from_f32<T>(o);
  }
}

// ---- rows of at most 32768 entries with 16-byte aligned rows: the row lives in REGISTERS (32 ordered keys per
// thread), shared memory holds only the level histograms.  Every level is a fully unrolled pass over the thread's
// own keys (shift, compare, predicated shared-memory atomic): ~30 instructions per entry for the whole row.
constexpr int kRegRowMax = 32 * kRT;

__device__ __forceinline__ float key_value(unsigned k) { return okey_inv(k); }

template <class T>
__global__ void __launch_bounds__(kRT, 1) topk_reg_kernel(const __grid_constant__ TopkParams P) {
  __shared__ SelScratch scs;
  __shared__ unsigned ucnt[8];
  SelScratch* sc = &scs;
  constexpr int kPer = 16 / int(sizeof(T));  // entries per 128-bit access
  constexpr int kIt = 32 / kPer;             // accesses per thread
  const int tid = threadIdx.x, V = P.V;
  const long long rowi = blockIdx.x;
  const T* row = reinterpret_cast<const T*>(P.probs) + rowi * P.in_stride;
  T* out = reinterpret_cast<T*>(P.out) + rowi * P.out_stride;
  const bool select = P.mode != 3;

  uint4 raw[kIt];
#pragma unroll
  for (int u = 0; u < kIt; ++u) {
    const int i0 = (u * kRT + tid) * kPer;
    raw[u] = i0 < V ? __ldg(reinterpret_cast<const uint4*>(row + i0)) : make_uint4(0u, 0u, 0u, 0u);
  }
  for (int i = tid; i < kBins; i += kRT) sc->hist[i] = 0u;
  if (tid < 8) ucnt[tid] = 0u;
  unsigned key[32];
#pragma unroll
  for (int u = 0; u < kIt; ++u) {
    const unsigned w[4] = {raw[u].x, raw[u].y, raw[u].z, raw[u].w};
    if constexpr (sizeof(T) == 4) {
#pragma unroll
      for (int c = 0; c < 4; ++c) key[u * 4 + c] = sel_key(__uint_as_float(w[c]));
    } else {
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        key[u * 8 + 2 * c] = sel_key(__uint_as_float(w[c] << 16));
        key[u * 8 + 2 * c + 1] = sel_key(__uint_as_float(w[c] & 0xffff0000u));
      }
    }
  }
  __syncthreads();
#define REG_EACH(BODY)                                         \
  _Pragma("unroll") for (int u = 0; u < kIt; ++u) {            \
    const int i0 = (u * kRT + tid) * kPer;                     \
    if (i0 < V) {                                              \
      _Pragma("unroll") for (int c = 0; c < kPer; ++c) {       \
        const unsigned kk = key[u * kPer + c];                 \
        const int idx = i0 + c;                                \
        (void)idx;                                             \
        BODY                                                   \
      }                                                        \
    }                                                          \
  }

  unsigned tau = 0u;
  int istar = 0x7fffffff;
  if (select) {
    REG_EACH(atomicAdd(&sc->hist[kk >> 21], 1u);)
    __syncthreads();
    int b1, b2, b3;
    unsigned a1, a2, a3;
    find_bin(sc, kBins, unsigned(P.k), &b1, &a1);
    for (int i = tid; i < kBins; i += kRT) sc->hist[i] = 0u;
    __syncthreads();
    REG_EACH(if (int(kk >> 21) == b1) atomicAdd(&sc->hist[(kk >> 10) & 2047u], 1u);)
    __syncthreads();
    find_bin(sc, kBins, unsigned(P.k) - a1, &b2, &a2);
    for (int i = tid; i < kBins; i += kRT) sc->hist[i] = 0u;
    __syncthreads();
    const unsigned top22 = (unsigned(b1) << 11) | unsigned(b2);
    REG_EACH(if ((kk >> 10) == top22) atomicAdd(&sc->hist[kk & 1023u], 1u);)
    __syncthreads();
    const unsigned need3 = unsigned(P.k) - a1 - a2;
    find_bin(sc, 1024, need3, &b3, &a3);
    tau = (top22 << 10) | unsigned(b3);
    const unsigned ties = sc->hist[b3], take = need3 - a3;  // 1 <= take <= ties
    if (take < ties) {
      // lowest index first.  Index order = (access u, thread, component): count the ties per access, find the access
      // in which the take-th one lies, then one block scan over that access's per-thread counts.
      unsigned cu[kIt];
#pragma unroll
      for (int u = 0; u < kIt; ++u) {
        cu[u] = 0u;
        const int i0 = (u * kRT + tid) * kPer;
        if (i0 < V) {
#pragma unroll
          for (int c = 0; c < kPer; ++c) cu[u] += (key[u * kPer + c] == tau);
        }
        if (cu[u]) atomicAdd(&ucnt[u], cu[u]);
      }
      __syncthreads();
      int ustar = 0;
      unsigned before = 0u;
#pragma unroll
      for (int u = 0; u < kIt; ++u) {
        const unsigned n = ucnt[u];
        if (before + n < take && u == ustar) before += n, ustar = u + 1;
      }
      unsigned mine = 0u;
#pragma unroll
      for (int u = 0; u < kIt; ++u) mine = (u == ustar) ? cu[u] : mine;
      unsigned total;
      const unsigned incl = block_incl_scan(mine, sc, &total);
      unsigned seen = before + incl - mine;
      if (mine != 0u && seen < take && take <= seen + mine) {
#pragma unroll
        for (int u = 0; u < kIt; ++u) {
          if (u == ustar) {
#pragma unroll
            for (int c = 0; c < kPer; ++c)
              if (key[u * kPer + c] == tau && ++seen == take) sc->istar = (u * kRT + tid) * kPer + c;
          }
        }
      }
      __syncthreads();
      istar = sc->istar;
    }
  }
  const unsigned zkey = okey(0.0f);
  // kept mass and the number of exact zeros after the scatter (fixed reduction order: deterministic)
  float ksum = 0.0f;
  int holes = 0;
  if (P.mode != 1) {
    REG_EACH({
      const bool kp = !select || kk > tau || (kk == tau && idx <= istar);
      const float kv = kp ? key_value(kk) : 0.0f;
      ksum += kv;
      holes += (!kp || kk == zkey);
    })
    ksum = block_sum_f(ksum, sc);
    holes = block_sum_i(holes, sc);
  }
  float tail = __fsub_rn(1.0f, ksum);
  if (P.mode == 0) tail = fmaxf(tail, 0.0f);
  const float fill = (P.mode != 1 && holes > 0) ? __fdiv_rn(tail, float(holes)) : 0.0f;
  float scale = 1.0f;
  if (P.mode == 0) {
    // sum of the rebuilt row = kept mass + holes * fill (the reference sums the row: same value up to fp32 rounding)
    const float total = fmaxf(__fadd_rn(ksum, __fmul_rn(fill, float(holes))), 1e-12f);
    scale = __frcp_rn(total);
  }
#pragma unroll
  for (int u = 0; u < kIt; ++u) {
    const int i0 = (u * kRT + tid) * kPer;
    if (i0 < V) {
      float o[kPer];
#pragma unroll
      for (int c = 0; c < kPer; ++c) {
        const unsigned kk = key[u * kPer + c];
        const bool kp = !select || kk > tau || (kk == tau && i0 + c <= istar);
        const float kv = kp ? key_value(kk) : 0.0f;
        float x = (kv == 0.0f) ? fill : kv;
        if (P.mode == 0) x *= scale;
        o[c] = x;
      }
      uint4 pk;
      if constexpr (sizeof(T) == 4) {
        pk = make_uint4(__float_as_uint(o[0]), __float_as_uint(o[1]), __float_as_uint(o[2]), __float_as_uint(o[3]));
      } else {
        unsigned w[4];
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          const __nv_bfloat162 h = __floats2bfloat162_rn(o[2 * c], o[2 * c + 1]);
          w[c] = *reinterpret_cast<const unsigned*>(&h);
        }
        pk = make_uint4(w[0], w[1], w[2], w[3]);
      }
      *reinterpret_cast<uint4*>(out + i0) = pk;
    }
  }
#undef REG_EACH
}

}  // namespace rowsel
}  // namespace duodec
