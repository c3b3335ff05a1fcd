// Probability-space operators: the reference's decoding_ops work on stored
// probability histories; these kernels are their drop-in device versions.
// One block per row, vocabulary streamed with coalesced loads (later passes
// over the same row hit L2).
#pragma once

#include "rowcore.cuh"

namespace duodec {

__device__ __forceinline__ float nan_to_num0(float v) {
  return (v == v && v != INFINITY && v != -INFINITY) ? v : 0.0f;
}

// The reference subtracts in the tensor dtype (src/decoding_ops.py:165) before
// max_fn upcasts (src/utils.py:791): for bf16 inputs the difference is rounded
// to bf16 first.
template <class T>
__device__ __forceinline__ float sub_in_dtype(float a, float b) {
  const float d = __fsub_rn(a, b);
  if constexpr (sizeof(T) == 2) return __bfloat162float(__float2bfloat16_rn(d));
  else return d;
}
template <class T>
__device__ __forceinline__ float div_in_dtype(float a, float b) {
  const float d = __fdiv_rn(a, b);
  if constexpr (sizeof(T) == 2) return __bfloat162float(__float2bfloat16_rn(d));
  else return d;
}

// ---------------------------------------------------------------------------
// sample(probs) with injected Exp(1) noise            src/utils.py:727-748
// ---------------------------------------------------------------------------
template <class T>
__global__ void __launch_bounds__(kThreads) sample_rows_kernel(const T* __restrict__ probs, long long p_stride,
                                                               const float* __restrict__ noise, long long n_stride,
                                                               int V, long long* __restrict__ out) {
  __shared__ Xchg xc;
  const long long row = blockIdx.x;
  const T* p = probs + row * p_stride;
  const float* nz = noise + row * n_stride;
  float mass = 0.0f;
  for (int i = threadIdx.x; i < V; i += kThreads) mass += fmaxf(nan_to_num0(to_f32(p[i])), 0.0f);
  mass = block_allreduce(mass, OpSumF(), &xc);
  long long tok = 0;  // zero-mass row: one-hot at argmax of an all-zero row == index 0
  if (mass > 0.0f) {
    ValIdx best{-1.0f, 0x7fffffff};
    for (int i = threadIdx.x; i < V; i += kThreads) {
      const float v = fmaxf(nan_to_num0(to_f32(p[i])), 0.0f);
      if (v > 0.0f) {
        const float key = __fdiv_rn(__fdiv_rn(v, mass), nz[i]);
        if (key > best.v || (key == best.v && i < best.i)) best = ValIdx{key, i};
      }
    }
    best = block_allreduce(best, OpArgMax(), &xc);
    tok = best.i == 0x7fffffff ? 0 : best.i;
  }
  if (threadIdx.x == 0) out[row] = tok;
}

// ---------------------------------------------------------------------------
// sample(max_fn(p - q)) with injected noise   src/decoding_ops.py:190-206
// ---------------------------------------------------------------------------
template <class T>
__global__ void __launch_bounds__(kThreads) residual_sample_kernel(const T* __restrict__ tp, long long tp_stride,
                                                                   const T* __restrict__ dp, long long dp_stride,
                                                                   const float* __restrict__ noise,
                                                                   long long n_stride, int V,
                                                                   long long* __restrict__ out) {
  __shared__ Xchg xc;
  const long long row = blockIdx.x;
  const T* p = tp + row * tp_stride;
  const T* q = dp + row * dp_stride;
  const float* nz = noise + row * n_stride;
  float mass = 0.0f;
  for (int i = threadIdx.x; i < V; i += kThreads) {
    const float d = nan_to_num0(sub_in_dtype<T>(to_f32(p[i]), to_f32(q[i])));
    if (d > 0.0f) mass += d;
  }
  mass = block_allreduce(mass, OpSumF(), &xc);
  long long tok;
  if (mass > 0.0f) {
    // max_fn normalises once, sample() normalises again
    float mass2 = 0.0f;
    for (int i = threadIdx.x; i < V; i += kThreads) {
      const float d = nan_to_num0(sub_in_dtype<T>(to_f32(p[i]), to_f32(q[i])));
      if (d > 0.0f) mass2 += __fdiv_rn(d, mass);
    }
    mass2 = block_allreduce(mass2, OpSumF(), &xc);
    ValIdx best{-1.0f, 0x7fffffff};
    for (int i = threadIdx.x; i < V; i += kThreads) {
      const float d = nan_to_num0(sub_in_dtype<T>(to_f32(p[i]), to_f32(q[i])));
      if (d > 0.0f) {
        const float key = __fdiv_rn(__fdiv_rn(__fdiv_rn(d, mass), mass2), nz[i]);
        if (key > best.v || (key == best.v && i < best.i)) best = ValIdx{key, i};
      }
    }
    best = block_allreduce(best, OpArgMax(), &xc);
    tok = best.i == 0x7fffffff ? 0 : best.i;
  } else {
    ValIdx bd{-INFINITY, 0x7fffffff};
    for (int i = threadIdx.x; i < V; i += kThreads) {
      const float d = nan_to_num0(sub_in_dtype<T>(to_f32(p[i]), to_f32(q[i])));
      if (d > bd.v || (d == bd.v && i < bd.i)) bd = ValIdx{d, i};
    }
    bd = block_allreduce(bd, OpArgMax(), &xc);
    tok = bd.i == 0x7fffffff ? 0 : bd.i;
  }
  if (threadIdx.x == 0) out[row] = tok;
}

// ---------------------------------------------------------------------------
// max_fn(x): norm(max(x,0)) with the argmax one-hot fallback  src/utils.py:787-810
// ---------------------------------------------------------------------------
template <class T>
__global__ void __launch_bounds__(kThreads) max_fn_kernel(const T* __restrict__ x, long long x_stride, int V,
                                                          float* __restrict__ out, long long o_stride) {
  __shared__ Xchg xc;
  const long long row = blockIdx.x;
  const T* xr = x + row * x_stride;
  float* o = out + row * o_stride;
  float mass = 0.0f;
  for (int i = threadIdx.x; i < V; i += kThreads) {
    const float d = nan_to_num0(to_f32(xr[i]));
    if (d > 0.0f) mass += d;
  }
  mass = block_allreduce(mass, OpSumF(), &xc);
  if (mass > 0.0f) {
    for (int i = threadIdx.x; i < V; i += kThreads) {
      const float d = nan_to_num0(to_f32(xr[i]));
      o[i] = d > 0.0f ? __fdiv_rn(d, mass) : 0.0f;
    }
  } else {
    ValIdx bd{-INFINITY, 0x7fffffff};
    for (int i = threadIdx.x; i < V; i += kThreads) {
      const float d = nan_to_num0(to_f32(xr[i]));
      if (d > bd.v || (d == bd.v && i < bd.i)) bd = ValIdx{d, i};
    }
    bd = block_allreduce(bd, OpArgMax(), &xc);
    const int hot = bd.i == 0x7fffffff ? 0 : bd.i;
    for (int i = threadIdx.x; i < V; i += kThreads) o[i] = (i == hot) ? 1.0f : 0.0f;
  }
}

// ---------------------------------------------------------------------------
// accept test on stored probabilities          src/decoding_ops.py:127-147
// one block per sequence; thread j handles draft position j (gamma <= 1024)
// ---------------------------------------------------------------------------
template <class T>
__global__ void accept_probs_kernel(const T* __restrict__ dprobs, long long d_sb, long long d_ss,
                                    const T* __restrict__ tprobs, long long t_sb, long long t_ss,
                                    const long long* __restrict__ tokens, long long tok_sb,
                                    const float* __restrict__ r, long long r_sb, int gamma, int V,
                                    unsigned char* __restrict__ mask, T* __restrict__ sel_q, T* __restrict__ sel_p,
                                    long long* __restrict__ accepted_len) {
  __shared__ int first_reject;
  const int b = blockIdx.x;
  if (threadIdx.x == 0) first_reject = gamma;
  __syncthreads();
  for (int j = threadIdx.x; j < gamma; j += blockDim.x) {
    const long long tok = tokens[(long long)b * tok_sb + j];
    float q = 0.0f, p = 0.0f;
    if (tok >= 0 && tok < V) {
      q = to_f32(dprobs[(long long)b * d_sb + (long long)j * d_ss + tok]);
      p = to_f32(tprobs[(long long)b * t_sb + (long long)j * t_ss + tok]);
    }
    const bool acc = r[(long long)b * r_sb + j] <= div_in_dtype<T>(p, q);
    mask[(long long)b * gamma + j] = (unsigned char)acc;
    sel_q[(long long)b * gamma + j] = from_f32<T>(q);
    sel_p[(long long)b * gamma + j] = from_f32<T>(p);
    if (!acc) atomicMin(&first_reject, j);
  }
  __syncthreads();
  if (threadIdx.x == 0) accepted_len[b] = first_reject;
}

// ---------------------------------------------------------------------------
// KV rollback: a cache with per-sequence length counters is truncated by
// clamping the counters (src/model_gpu.py:362-383: end_pos clamp), no copy.
// ---------------------------------------------------------------------------
__global__ void kv_truncate_kernel(int* __restrict__ seq_len, const long long* __restrict__ end_pos, int B,
                                   long long end_pos_scalar) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b < B) {
    long long e = end_pos ? end_pos[b] : end_pos_scalar;
    if (e < 0) e = 0;
    const int cur = seq_len[b];
    seq_len[b] = e < cur ? int(e) : cur;
  }
}

}  // namespace duodec

namespace duodec {

// ---------------------------------------------------------------------------
// transfer_top_k under vocabulary sharding (merge of per-shard top-k candidates)
//   compact_nonzero_kernel   one block per row: the non-zero entries of a (locally top-k compressed) slice as
//                            (value, GLOBAL index) pairs in index order; free slots hold (-1, INT_MAX)
//   shard_topk_apply_kernel  one block per row: with the merged boundary (tau, istar) of the global top-k,
//                            rebuild the slice like utils.rebuild_topk_probs (src/utils.py:751-777): kept
//                            entries keep their value, every exact zero takes `fill`, all divided by `total`
// ---------------------------------------------------------------------------
template <class T>
__global__ void __launch_bounds__(256) compact_nonzero_kernel(const T* __restrict__ x, long long stride, int Vs,
                                                              int v_begin, int kcap, float* __restrict__ cand_val,
                                                              int* __restrict__ cand_idx, int* __restrict__ count) {
  __shared__ int warp_tot[8];
  const long long row = blockIdx.x;
  const T* p = x + row * stride;
  const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  const int chunk = (Vs + 255) / 256;  // a thread owns a contiguous chunk, so list order is index order
  const int lo = min(tid * chunk, Vs), hi = min(lo + chunk, Vs);
  int c = 0;
  for (int i = lo; i < hi; ++i) c += to_f32(p[i]) != 0.0f;
  int inc = c;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const int o = __shfl_up_sync(0xffffffffu, inc, d);
    if (lane >= d) inc += o;
  }
  if (lane == 31) warp_tot[w] = inc;
  __syncthreads();
  int base = inc - c, total = 0;
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    if (k < w) base += warp_tot[k];
    total += warp_tot[k];
  }
  float* cv = cand_val + row * kcap;
  int* ci = cand_idx + row * kcap;
  for (int i = lo; i < hi; ++i) {
    const float v = to_f32(p[i]);
    if (v != 0.0f) {
      if (base < kcap) cv[base] = v, ci[base] = v_begin + i;
      ++base;
    }
  }
  for (int s = min(total, kcap) + tid; s < kcap; s += 256) cv[s] = -1.0f, ci[s] = 0x7fffffff;
  if (tid == 0) count[row] = total;
}

template <class T>
__global__ void __launch_bounds__(256) shard_topk_apply_kernel(const T* __restrict__ x, long long in_stride,
                                                               T* __restrict__ out, long long out_stride, int Vs,
                                                               int v_begin, const float* __restrict__ tau,
                                                               const int* __restrict__ istar,
                                                               const float* __restrict__ fill,
                                                               const float* __restrict__ total) {
  const long long row = blockIdx.x;
  const T* p = x + row * in_stride;
  T* o = out + row * out_stride;
  const float t = tau[row], f = fill[row], tot = total[row];
  const int is = istar[row];
  for (int i = threadIdx.x; i < Vs; i += 256) {
    const float v = to_f32(p[i]);
    const bool kept = (v > t || (v == t && v_begin + i <= is)) && v != 0.0f;
    o[i] = from_f32<T>(__fdiv_rn(kept ? v : f, tot));
  }
}

}  // namespace duodec
