// Greedy verification (temperature == 0 on both models): fast path of duodec_verify_fused.
//
// At T == 0 the reference's norm_logits returns the fp32 one-hot of the first arg-max
// (src/utils.py:700-704) for draft and target rows alike, so the whole step needs nothing but the
// arg-max of every row - no exponentials, no filter, no noise:
//   accept test   r_j <= p_j(x_j) / q_j(x_j) with p, q in {0, 1} as raw IEEE division
//                 (src/decoding_ops.py:127-147): 1/1 accepts, 0/1 accepts only r == 0, 1/0 = inf accepts,
//                 0/0 = nan rejects;
//   resample      sample(max_fn(p_n - q_n)) (src/decoding_ops.py:190-206): the target's arg-max when it
//                 differs from the draft's, else the zero-mass fallback argmax(0) = index 0
//                 (src/utils.py:787-810);
//   bonus         sample(one-hot) = the target's arg-max of the last row (src/decoding_ops.py:209-220).
//
//   greedy_argmax_kernel  grid (slices of 8192 entries, rows): 128-bit loads, per-thread (value, lowest
//                         index) race, block reduction of the packed key, one atomicMax per block on the
//                         row's 64-bit cell.  One pass over the logits at copy speed.
//   greedy_close_kernel   one thread per sequence: decodes the 2*gamma+1 cells, accept tests, leading run
//                         (its own, or sequence 0's under the batch rule), token, outputs.
// The cells live in their own zero-initialised area of the workspace; the close kernel leaves them zeroed.
#pragma once

#include "split_kernels.cuh"

namespace duodec {
namespace greedy {

constexpr int kGT = 256;         // threads per block
constexpr int kGSlice = 8192;    // row entries per block

struct GreedyParams {
  const void* draft;
  const void* target;
  long long d_sb, d_ss, t_sb, t_ss;
  int B, gamma, V, row0_rule;
  const long long* tokens;
  long long tok_sb;
  const float* r;
  long long r_sb;
  unsigned char* accept_mask;
  float* sel_p;
  float* sel_q;
  long long* accepted_len;
  long long* token;
  unsigned char* all_accepted;
  unsigned long long* cells;  // [B, 2*gamma+1]: (okey(value) << 32) | ~index of the row's arg-max, 0 = empty
};

// NaN never wins (as in row_argmax of the general kernel); -inf rows end with an empty cell -> index 0
__device__ __forceinline__ void greedy_take(float x, int i, float& bv, int& bi) {
  if (x > bv) bv = x, bi = i;  // a thread meets its indices in increasing order: strict > keeps the lowest
}

template <class TIn>
__global__ void __launch_bounds__(kGT) greedy_argmax_kernel(const __grid_constant__ GreedyParams P) {
  __shared__ unsigned long long wk[kGT / 32];
  const int tid = threadIdx.x, job = blockIdx.y;
  const RowJob jb = decode_job(job, P.gamma);
  const TIn* row = jb.is_target
                       ? reinterpret_cast<const TIn*>(P.target) + (long long)jb.b * P.t_sb + (long long)jb.j * P.t_ss
                       : reinterpret_cast<const TIn*>(P.draft) + (long long)jb.b * P.d_sb + (long long)jb.j * P.d_ss;
  const int s0 = blockIdx.x * kGSlice;
  const int s1 = min(s0 + kGSlice, P.V);
  constexpr int kPer = 16 / int(sizeof(TIn));  // entries per 128-bit load
  float bv = -INFINITY;
  int bi = 0x7fffffff;
#pragma unroll 4
  for (int i0 = s0 + tid * kPer; i0 < s1; i0 += kGT * kPer) {
    const uint4 raw = __ldcs(reinterpret_cast<const uint4*>(row + i0));  // streamed once: evict first
    if constexpr (sizeof(TIn) == 4) {
      greedy_take(__uint_as_float(raw.x), i0, bv, bi), greedy_take(__uint_as_float(raw.y), i0 + 1, bv, bi);
      greedy_take(__uint_as_float(raw.z), i0 + 2, bv, bi), greedy_take(__uint_as_float(raw.w), i0 + 3, bv, bi);
    } else {
      const unsigned w[4] = {raw.x, raw.y, raw.z, raw.w};
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        greedy_take(__uint_as_float(w[k] << 16), i0 + 2 * k, bv, bi);
        greedy_take(__uint_as_float(w[k] & 0xffff0000u), i0 + 2 * k + 1, bv, bi);
      }
    }
  }
  unsigned long long key =
      bi == 0x7fffffff ? 0ull : ((unsigned long long)okey(bv) << 32) | (unsigned long long)(0xffffffffu - unsigned(bi));
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const unsigned long long ok = __shfl_xor_sync(0xffffffffu, key, o);
    key = ok > key ? ok : key;
  }
  if ((tid & 31) == 0) wk[tid >> 5] = key;
  __syncthreads();
  if (tid == 0) {
#pragma unroll
    for (int w = 1; w < kGT / 32; ++w) key = wk[w] > key ? wk[w] : key;
    if (key != 0ull) atomicMax(&P.cells[job], key);
  }
}

__device__ __forceinline__ long long greedy_cell_index(unsigned long long cell) {
  return cell == 0ull ? 0 : (long long)(0xffffffffu - unsigned(cell & 0xffffffffull));
}

// accepted run of one sequence from its cells; writes the per-position outputs when `write` is set
__device__ __forceinline__ int greedy_run(const GreedyParams& P, int b, bool write) {
  const unsigned long long* c = P.cells + (long long)b * (2 * P.gamma + 1);
  int n = 0;
  bool open = true;
  for (int j = 0; j < P.gamma; ++j) {
    const long long am_q = greedy_cell_index(__ldcg(c + 2 * j)), am_p = greedy_cell_index(__ldcg(c + 2 * j + 1));
    const long long tok = P.tokens[(long long)b * P.tok_sb + j];
    const float sp = (tok >= 0 && tok < P.V && tok == am_p) ? 1.0f : 0.0f;
    const float sq = (tok >= 0 && tok < P.V && tok == am_q) ? 1.0f : 0.0f;
    const bool acc = P.r[(long long)b * P.r_sb + j] <= __fdiv_rn(sp, sq);  // raw IEEE division
    if (write) {
      const long long o = (long long)b * P.gamma + j;
      P.accept_mask[o] = (unsigned char)acc;
      P.sel_p[o] = sp;
      P.sel_q[o] = sq;
    }
    if (open && acc) ++n;
    else open = false;
  }
  return n;
}

// One block closes every sequence: all reads of the cells (a sequence's own and, under the batch rule, sequence
// 0's) happen before the barrier, the cells are zeroed after it - the workspace is ready for the next launch
// without a memset.  Launched with programmatic stream serialization: it becomes resident while the arg-max
// kernel drains.
__global__ void __launch_bounds__(256) greedy_close_kernel(const __grid_constant__ GreedyParams P) {
  asm volatile("griddepcontrol.wait;" ::: "memory");
  const int per_b = 2 * P.gamma + 1;
  // reference batch rule (src/decoding_ops.py:144-147): sequence 0's count, read before anything is zeroed
  const int n0 = P.row0_rule ? greedy_run(P, 0, false) : 0;
  for (int b0 = 0; b0 < P.B; b0 += 256) {
    const int b = b0 + threadIdx.x;
    int n = 0;
    long long tok = 0;
    if (b < P.B) {
      n = greedy_run(P, b, true);
      if (P.row0_rule) n = n0;
      const unsigned long long* c = P.cells + (long long)b * per_b;
      if (n == P.gamma) {
        tok = greedy_cell_index(__ldcg(c + 2 * P.gamma));
      } else {
        const long long am_q = greedy_cell_index(__ldcg(c + 2 * n)), am_p = greedy_cell_index(__ldcg(c + 2 * n + 1));
        tok = am_p != am_q ? am_p : 0;  // equal one-hots: zero residual mass -> first arg-max of p - q == 0 is index 0
      }
    }
    __syncthreads();
    if (b < P.B) {
      P.accepted_len[b] = n;
      P.token[b] = tok;
      P.all_accepted[b] = (unsigned char)(n == P.gamma);
      for (int k = 0; k < per_b; ++k) P.cells[(long long)b * per_b + k] = 0ull;
    }
  }
}

}  // namespace greedy
}  // namespace duodec
