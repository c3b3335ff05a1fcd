// CUHLM helpers of the reference's CommunicationSimulator.
//
//   uncertainty_tokens_kernel   calculate_uncertainty (src/communication.py:541-575): M tempered copies of ONE
//                               logits row, softmax(logits / T_m - max), one multinomial draw each
//                               (= argmax p_i / Exp(1) noise, lowest index on ties).  One block per
//                               temperature; the row (<= 600 KB) is read from L2 three times (max, sum, race).
//                               The reference builds an [M, V] tensor and runs five elementwise passes over it.
//   compressed_vocab_kernel     _calculate_compressed_vocab_size (src/communication.py:681-755): the smallest k with
//                               sum_{i >= k} |s_i - u_k| / denominator <= theta over the probabilities sorted
//                               descending, u_k = (1 - sum_{i<k} s_i) / (V - k).  The reference walks k = 1 .. V-1
//                               with an inner Python loop over the tail (O(V^2) `.item()` calls); here every k
//                               is one thread: s is descending, so the tail splits at one crossover j(k) (binary
//                               search) and the sum follows from float64 prefix sums; atomicMin picks the answer.
#pragma once

#include "prob_kernels.cuh"

namespace duodec {

template <class TIn>
__global__ void __launch_bounds__(kThreads) uncertainty_tokens_kernel(const TIn* __restrict__ logits, int V,
                                                                      const float* __restrict__ temps,
                                                                      const float* __restrict__ noise,
                                                                      long long noise_stride,
                                                                      long long* __restrict__ tokens) {
  __shared__ Xchg xc;
  const int m = blockIdx.x;
  const float temp = temps[m];
  const float* nz = noise + (long long)m * noise_stride;
  float mx = -INFINITY;
  for (int i = threadIdx.x; i < V; i += kThreads) mx = fmaxf(mx, __fdiv_rn(to_f32(logits[i]), temp));
  mx = block_allreduce(mx, OpMaxF(), &xc);
  float z = 0.0f;
  for (int i = threadIdx.x; i < V; i += kThreads) z += expf(__fdiv_rn(to_f32(logits[i]), temp) - mx);
  z = block_allreduce(z, OpSumF(), &xc);
  ValIdx best{-1.0f, 0x7fffffff};
  for (int i = threadIdx.x; i < V; i += kThreads) {
    const float p = __fdiv_rn(expf(__fdiv_rn(to_f32(logits[i]), temp) - mx), z);
    if (p > 0.0f) {
      const float key = __fdiv_rn(p, nz[i]);
      if (key > best.v || (key == best.v && i < best.i)) best = ValIdx{key, i};
    }
  }
  best = block_allreduce(best, OpArgMax(), &xc);
  if (threadIdx.x == 0) tokens[m] = best.i == 0x7fffffff ? 0 : best.i;
}

// sorted[V] descending (float32), prefix[V] inclusive float64 sums of sorted; *k_out preset to INT_MAX
__global__ void __launch_bounds__(256) compressed_vocab_kernel(const float* __restrict__ sorted,
                                                               const double* __restrict__ prefix, int V,
                                                               double denominator, double theta, int* k_out) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x + 1;  // k = 1 .. V-1
  if (k >= V) return;
  const double top = prefix[k - 1], total = prefix[V - 1];
  const double residual = 1.0 - top;
  const double u = residual <= 0.0 ? 0.0 : residual / double(V - k);
  // first j in [k, V) with sorted[j] < u (descending order): entries before it lie at or above u
  int lo = k, hi = V;
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (double(sorted[mid]) < u) hi = mid;
    else lo = mid + 1;
  }
  const int j = lo;
  const double above = (j > k ? prefix[j - 1] - top : 0.0) - double(j - k) * u;
  const double below = double(V - j) * u - (total - (j > 0 ? prefix[j - 1] : 0.0));
  if ((above + below) / denominator <= theta) atomicMin(k_out, k);
}

// Rank bookkeeping of the accepted tokens (_record_accepted_token_ranks, /root/reference/src/baselines.py:121-168:
// per token a `(probs > p).sum().item()`, a sorted torch.topk of the row and a `.nonzero()` - three host syncs and
// a sort per accepted token).  One block per accepted position: the number of entries greater than the token's
// probability and the number of EQUAL entries at a lower index (the canonical order of the sorted top-k) - the
// caller reads both counts of every position back in one copy.
template <class TIn>
__global__ void __launch_bounds__(kThreads) token_rank_kernel(const TIn* __restrict__ probs, long long row_stride,
                                                              int V, const long long* __restrict__ tokens,
                                                              long long tok_stride, int* __restrict__ out) {
  __shared__ Xchg xc;
  const int rowi = blockIdx.x;
  const TIn* row = probs + (long long)rowi * row_stride;
  const long long tok = tokens[(long long)rowi * tok_stride];
  int greater = 0, before = 0;
  if (tok >= 0 && tok < V) {
    const float pt = to_f32(row[tok]);
    for (int i = threadIdx.x; i < V; i += kThreads) {
      const float p = to_f32(row[i]);
      greater += (p > pt);
      before += (p == pt && i < tok);
    }
  }
  greater = block_allreduce(greater, OpSumI(), &xc);
  before = block_allreduce(before, OpSumI(), &xc);
  if (threadIdx.x == 0) {
    out[2 * rowi] = (tok >= 0 && tok < V) ? greater : -1;
    out[2 * rowi + 1] = before;
  }
}

}  // namespace duodec
