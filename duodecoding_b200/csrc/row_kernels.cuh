// Row kernels that materialise a V-sized output: norm_logits and the
// transfer_top_k compression family.  One team per row, row staged in shared
// memory, every logit/probability read from HBM once and written once.
#pragma once

#include "rowcore.cuh"

namespace duodec {

struct NormParams {
  const void* logits;  // [batch, inner, V], vocab stride 1; row = b * inner + s
  long long in_stride, in_bstride;
  void* out;  // [batch, inner, V]
  long long out_stride, out_bstride;
  int rows, inner, V;
  float temperature, top_p;
  int top_k;
  int per, slice_floats;
};

// norm_logits (src/utils.py:686-709): T == 0 -> one-hot of the first argmax;
// else softmax(top_k_top_p_filter(logits / T)).
template <class TIn, class TOut, int C>
__global__ void __launch_bounds__(kThreads, 1) norm_logits_kernel(const __grid_constant__ NormParams P) {
  extern __shared__ __align__(128) unsigned char smem[];
  float* xs = reinterpret_cast<float*>(smem);
  void* work = xs + P.slice_floats;
  RowScratch* sc = reinterpret_cast<RowScratch*>(reinterpret_cast<unsigned char*>(work) + kHistBytes);
  Xchg* xc = &sc->xc;
  const unsigned rank = Team<C>::rank();
  const long long rowi = Team<C>::id();
  if (threadIdx.x == 0) {
    xc->phase = 0;
    mbar_init(&sc->bar, 1);
    mbar_fence_init();
  }
  __syncthreads();
  const long long rb = rowi / P.inner, rs = rowi - rb * P.inner;
  const TIn* row = reinterpret_cast<const TIn*>(P.logits) + rb * P.in_bstride + rs * P.in_stride;
  TOut* out = reinterpret_cast<TOut*>(P.out) + rb * P.out_bstride + rs * P.out_stride;
  const Slice s = team_slice(P.per, P.V, rank);
  const bool greedy = P.temperature == 0.0f;
  const float temp = greedy ? 1.0f : P.temperature;
  unsigned par = 0;
  const bool a = stage_begin<TIn>(xs, row, s, temp, &sc->bar);
  stage_end(a, &sc->bar, par);
  if constexpr (C > 1) Team<C>::sync();
  if (greedy) {
    const int hot = row_argmax<C>(xs, s, xc).i;
    for (int i = threadIdx.x; i < s.n; i += kThreads) out[s.g0 + i] = from_f32<TOut>((s.g0 + i) == hot ? 1.0f : 0.0f);
  } else {
    const RowStat st = row_core<C, TIn>(xs, s, P.per, P.V, row, temp, P.top_k, P.top_p, work, sc);
    const bool vec = (sizeof(TOut) == 4) && ((reinterpret_cast<uintptr_t>(out + s.g0) & 15) == 0);
    if (vec) {
      const float4* x4 = reinterpret_cast<const float4*>(xs);
      float4* o4 = reinterpret_cast<float4*>(out + s.g0);
      for (int i = threadIdx.x; i < (s.n >> 2); i += kThreads) {
        float4 v = x4[i];
        v.x = __fdiv_rn(keep_e(v.x, st.keep_bits), st.Z);
        v.y = __fdiv_rn(keep_e(v.y, st.keep_bits), st.Z);
        v.z = __fdiv_rn(keep_e(v.z, st.keep_bits), st.Z);
        v.w = __fdiv_rn(keep_e(v.w, st.keep_bits), st.Z);
        o4[i] = v;
      }
      for (int i = (s.n & ~3) + threadIdx.x; i < s.n; i += kThreads)
        out[s.g0 + i] = from_f32<TOut>(__fdiv_rn(keep_e(xs[i], st.keep_bits), st.Z));
    } else {
      for (int i = threadIdx.x; i < s.n; i += kThreads)
        out[s.g0 + i] = from_f32<TOut>(__fdiv_rn(keep_e(xs[i], st.keep_bits), st.Z));
    }
  }
  if constexpr (C > 1) Team<C>::sync();
}

// ---------------------------------------------------------------------------
// transfer_top_k compression
//   mode 0: utils.rebuild_topk_probs        (src/utils.py:751-777)  top-k, uniform refill of
//           clamp(1-sum,0) over exact zeros, renormalise by clamp(sum,1e-12)
//   mode 1: CommunicationSimulator._apply_top_k_compression (src/communication.py:288-308)
//           top-k scattered into zeros, nothing else
//   mode 2: CommunicationSimulator.compress_rebuild_probs (src/communication.py:348-395)
//           top-k, uniform refill of (1-sum) over exact zeros, no clamp, no renorm
//   mode 3: CommunicationSimulator.rebuild_full_probs (src/communication.py:310-346)
//           no top-k: refill the exact zeros of an already compressed row
// Ties at the k-th value: lowest index first (canonical order).
// ---------------------------------------------------------------------------
struct TopkParams {
  const void* probs;
  long long in_stride;
  void* out;
  long long out_stride;
  int rows, V, k, mode;
  int per, slice_floats;
};

// index of the t-th (1-based) entry equal to tau in index order, or INT_MAX
// when the row holds no more than t such entries.
template <int C>
__device__ __forceinline__ int team_tie_cut(const float* xs, Slice s, float tau, int t, RowScratch* sc) {
  Xchg* xc = &sc->xc;
  const int chunk = (s.n + kThreads - 1) / kThreads;
  const int i0 = int(threadIdx.x) * chunk;
  int cnt = 0;
  for (int i = i0; i < i0 + chunk && i < s.n; ++i) cnt += (xs[i] == tau);
  int block_total = 0;
  const int before_in_block = block_exscan(cnt, OpSumI(), 0, xc, &block_total);
  // ties held by lower ranks
  int before_block = 0, all = block_total;
  if constexpr (C > 1) {
    struct I4 {
      int v[4];
    };
    struct OpSumI4 {
      __device__ __forceinline__ I4 operator()(I4 a, I4 b) const {
        return I4{{a.v[0] + b.v[0], a.v[1] + b.v[1], a.v[2] + b.v[2], a.v[3] + b.v[3]}};
      }
    };
    const unsigned me = Team<C>::rank();
    I4 lo{{0, 0, 0, 0}}, hi{{0, 0, 0, 0}};
    if (threadIdx.x == 0) {  // one contribution per block: the block reduce sums over threads
      if (me < 4) lo.v[me] = block_total;
      else hi.v[me - 4] = block_total;
    }
    lo = team_allreduce<C>(lo, OpSumI4(), xc);
    if constexpr (C > 4) hi = team_allreduce<C>(hi, OpSumI4(), xc);
    all = 0;
    for (unsigned r = 0; r < unsigned(C); ++r) {
      const int c = r < 4 ? lo.v[r] : hi.v[r - 4];
      if (r < me) before_block += c;
      all += c;
    }
  }
  int found = 0x7fffffff;
  if (all > t) {
    int seen = before_block + before_in_block;
    for (int i = i0; i < i0 + chunk && i < s.n; ++i) {
      if (xs[i] == tau) {
        ++seen;
        if (seen == t) found = s.g0 + i;
      }
    }
  }
  return team_allreduce<C>(found, OpMinI(), xc);
}

template <class T, int C>
__global__ void __launch_bounds__(kThreads, 1) topk_rebuild_kernel(const __grid_constant__ TopkParams P) {
  extern __shared__ __align__(128) unsigned char smem[];
  float* xs = reinterpret_cast<float*>(smem);
  void* work = xs + P.slice_floats;
  RowScratch* sc = reinterpret_cast<RowScratch*>(reinterpret_cast<unsigned char*>(work) + kHistBytes);
  Xchg* xc = &sc->xc;
  const unsigned rank = Team<C>::rank();
  const long long rowi = Team<C>::id();
  if (threadIdx.x == 0) {
    xc->phase = 0;
    mbar_init(&sc->bar, 1);
    mbar_fence_init();
  }
  __syncthreads();
  const T* row = reinterpret_cast<const T*>(P.probs) + rowi * P.in_stride;
  T* out = reinterpret_cast<T*>(P.out) + rowi * P.out_stride;
  const Slice s = team_slice(P.per, P.V, rank);
  unsigned par = 0;
  const bool a = stage_begin<T>(xs, row, s, 1.0f, &sc->bar);
  stage_end(a, &sc->bar, par);
  if constexpr (C > 1) Team<C>::sync();
  const int n4 = (s.n + 3) >> 2;

  float tau = -INFINITY;
  int istar = 0x7fffffff;
  if (P.mode != 3) {
    const KthResult kr = team_kth_largest<C>(xs, n4, P.k, reinterpret_cast<int*>(work), sc->res, xc);
    tau = kr.value;
    istar = team_tie_cut<C>(xs, s, tau, P.k - kr.greater, sc);
  }
  auto kept = [&](float v, int g) { return v > tau || (v == tau && g <= istar); };

  if (P.mode == 1) {
    for (int i = threadIdx.x; i < s.n; i += kThreads) {
      const float v = xs[i];
      out[s.g0 + i] = from_f32<T>(kept(v, s.g0 + i) ? v : 0.0f);
    }
  } else {
    // kept mass and the number of exact zeros after the scatter
    float ksum = 0.0f;
    int holes = 0;
    for (int i = threadIdx.x; i < s.n; i += kThreads) {
      const float v = xs[i];
      const float kv = kept(v, s.g0 + i) ? v : 0.0f;
      ksum += kv;
      holes += (kv == 0.0f);
    }
    ksum = team_allreduce<C>(ksum, OpSumF(), xc);
    holes = team_allreduce<C>(holes, OpSumI(), xc);
    float tail = __fsub_rn(1.0f, ksum);
    if (P.mode == 0) tail = fmaxf(tail, 0.0f);
    const float fill = holes > 0 ? __fdiv_rn(tail, float(holes)) : 0.0f;
    float total = 1.0f;
    if (P.mode == 0) {
      float tsum = 0.0f;
      for (int i = threadIdx.x; i < s.n; i += kThreads) {
        const float v = xs[i];
        const float kv = kept(v, s.g0 + i) ? v : 0.0f;
        tsum += (kv == 0.0f) ? fill : kv;
      }
      total = fmaxf(team_allreduce<C>(tsum, OpSumF(), xc), 1e-12f);
    }
    for (int i = threadIdx.x; i < s.n; i += kThreads) {
      const float v = xs[i];
      const float kv = kept(v, s.g0 + i) ? v : 0.0f;
      float o = (kv == 0.0f) ? fill : kv;
      if (P.mode == 0) o = __fdiv_rn(o, total);
      out[s.g0 + i] = from_f32<T>(o);
    }
  }
  if constexpr (C > 1) Team<C>::sync();
}

}  // namespace duodec
