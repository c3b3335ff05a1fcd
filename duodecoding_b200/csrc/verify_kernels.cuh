// Fused speculative-decoding verification kernels.
//
// verify_pair_kernel: one team per (sequence b, draft position j).  The team
// stages target row j and draft row j (the bonus position j == gamma has only
// the target row), runs the logit warp + softmax statistics on both while
// they sit in shared memory, evaluates the accept test of its position and -
// when its own position rejects (or is the bonus position) - draws the
// replacement token from the rows it already holds.  The last team of a
// sequence to finish scans the gamma accept bits and picks the token of the
// first rejected position (or the bonus token).  HBM traffic: every logit is
// read exactly once; nothing V-sized is written.
//
// Reference semantics restated (file:line under /root/reference):
//   accept test  r <= p/q, raw IEEE division          src/decoding_ops.py:127-145
//   leading run of accepts, n                          src/decoding_ops.py:145-147
//   residual resample  sample(max_fn(p - q))          src/decoding_ops.py:190-206, src/utils.py:787-810
//   bonus sample       sample(p_next)                  src/decoding_ops.py:209-220, src/utils.py:727-748
//   multinomial == argmax(prob / Exp(1) noise)         ATen multinomial, SURVEY fact 4
#pragma once

#include "rowcore.cuh"

namespace duodec {

struct VerifyParams {
  const void* draft;   // [B, >=gamma, Vd] logits
  const void* target;  // [B, >=gamma+1, Vt] logits
  long long d_sb, d_ss, t_sb, t_ss;  // element strides (batch, position); vocab stride is 1
  int B, gamma, Vd, Vt;
  const long long* tokens;  // [B, gamma] drafted token ids
  long long tok_sb;
  const float* r;  // [B, gamma] uniforms of the accept test
  long long r_sb;
  const float* noise;  // [B, >=max(Vd,Vt)] Exp(1) draws of the one sampled row per sequence
  long long noise_sb;
  float t_temp, t_top_p;
  int t_top_k;
  float d_temp, d_top_p;
  int d_top_k;
  // outputs
  unsigned char* accept_mask;  // [B, gamma]
  float* sel_p;                // [B, gamma] target prob of the drafted token
  float* sel_q;                // [B, gamma] draft prob of the drafted token
  long long* accepted_len;     // [B]
  long long* token;            // [B]
  unsigned char* all_accepted;  // [B]
  // workspace (zero-initialised once by the caller; the kernel leaves it zeroed)
  int* counters;          // [B]
  long long* pair_token;  // [B, gamma+1]
  int per;                // slice pitch of a team
  int slice_floats;       // shared-memory floats reserved per row slice
};

// probability of entry `tok` of a staged row (0 when tok is outside the row)
template <int C>
__device__ __forceinline__ float team_pick(const float* xs, Slice s, long long tok, unsigned keep_bits, float Z,
                                           Xchg* xc) {
  float v = 0.0f;
  if (threadIdx.x == 0 && tok >= s.g0 && tok < s.g0 + s.n) v = __fdiv_rn(keep_e(xs[tok - s.g0], keep_bits), Z);
  return team_allreduce<C>(v, OpSumF(), xc);
}

template <class TIn, int C>
__global__ void __launch_bounds__(kThreads, 1) verify_pair_kernel(const __grid_constant__ VerifyParams P) {
  extern __shared__ __align__(128) unsigned char smem[];
  float* xp = reinterpret_cast<float*>(smem);
  float* xq = xp + P.slice_floats;
  void* work = xq + P.slice_floats;
  RowScratch* sc = reinterpret_cast<RowScratch*>(reinterpret_cast<unsigned char*>(work) + kHistBytes);
  Xchg* xc = &sc->xc;

  const unsigned rank = Team<C>::rank();
  const int pair = int(Team<C>::id());
  const int b = pair / (P.gamma + 1);
  const int j = pair - b * (P.gamma + 1);
  const bool bonus = (j == P.gamma);
  const int vlim = P.Vd < P.Vt ? P.Vd : P.Vt;

  if (threadIdx.x == 0) {
    xc->phase = 0;
    mbar_init(&sc->bar, 1);
    mbar_init(&sc->bar2, 1);
    mbar_fence_init();
  }
  __syncthreads();

  const TIn* prow = reinterpret_cast<const TIn*>(P.target) + (long long)b * P.t_sb + (long long)j * P.t_ss;
  const TIn* qrow = reinterpret_cast<const TIn*>(P.draft) + (long long)b * P.d_sb + (long long)j * P.d_ss;
  const Slice sp = team_slice(P.per, P.Vt, rank);
  const Slice sq = bonus ? Slice{0, 0} : team_slice(P.per, P.Vd, rank);
  const bool greedy = (P.t_temp == 0.0f);  // the host rejects mixed greedy / sampled temperatures
  const float t_temp = greedy ? 1.0f : P.t_temp;
  const float d_temp = greedy ? 1.0f : P.d_temp;

  // both rows in flight before either is touched
  {
    unsigned par0 = 0, par1 = 0;
    const bool ap = stage_begin<TIn>(xp, prow, sp, t_temp, &sc->bar);
    const bool aq = bonus ? false : stage_begin<TIn>(xq, qrow, sq, d_temp, &sc->bar2);
    stage_end(ap, &sc->bar, par0);
    stage_end(aq, &sc->bar2, par1);
  }
  if constexpr (C > 1) Team<C>::sync();  // peers' shared memory is live from here on

  const long long tok = bonus ? -1 : P.tokens[(long long)b * P.tok_sb + j];
  const float* nz = P.noise + (long long)b * P.noise_sb;
  int accept = 0;
  long long cand = -1;
  float sel_p = 0.0f, sel_q = 0.0f;

  const long long tok_in = (tok >= 0 && tok < vlim) ? tok : -1;  // out-of-range ids: p = q = 0 -> nan -> reject
  if (greedy) {
    // temperature == 0: one-hot rows (src/utils.py:700-704)
    const int am_p = row_argmax<C>(xp, sp, xc).i;
    if (bonus) {
      cand = am_p;  // sample(one-hot) is its index
    } else {
      const int am_q = row_argmax<C>(xq, sq, xc).i;
      sel_p = (tok_in == am_p) ? 1.0f : 0.0f;
      sel_q = (tok_in == am_q) ? 1.0f : 0.0f;
      accept = P.r[(long long)b * P.r_sb + j] <= __fdiv_rn(sel_p, sel_q);
      if (!accept) {
        // residual of two one-hots over [0, vlim): +1 at am_p when it differs
        // from am_q; otherwise the mass is zero and the fallback is the first
        // index of max(p - q) (0, or 1 when the only negative entry is index 0)
        if (am_p != am_q && am_p < vlim) cand = am_p;
        else if (am_p == am_q) cand = 0;
        else cand = (am_q == 0 && vlim > 1) ? 1 : 0;
      }
    }
  } else {
    const RowStat stp = row_core<C, TIn>(xp, sp, P.per, P.Vt, prow, t_temp, P.t_top_k, P.t_top_p, work, sc);
    if (bonus) {
      // bonus token: argmax_i p_i / noise_i over the whole target vocabulary
      ValIdx best{-1.0f, 0x7fffffff};
      for (int i = threadIdx.x; i < sp.n; i += kThreads) {
        const float p = __fdiv_rn(keep_e(xp[i], stp.keep_bits), stp.Z);
        if (p > 0.0f) {
          const int g = sp.g0 + i;
          const float key = __fdiv_rn(p, nz[g]);
          if (key > best.v || (key == best.v && g < best.i)) best = ValIdx{key, g};
        }
      }
      best = team_allreduce<C>(best, OpArgMax(), xc);
      cand = best.i == 0x7fffffff ? 0 : best.i;
    } else {
      const RowStat stq = row_core<C, TIn>(xq, sq, P.per, P.Vd, qrow, d_temp, P.d_top_k, P.d_top_p, work, sc);
      sel_p = team_pick<C>(xp, sp, tok_in, stp.keep_bits, stp.Z, xc);
      sel_q = team_pick<C>(xq, sq, tok_in, stq.keep_bits, stq.Z, xc);
      accept = P.r[(long long)b * P.r_sb + j] <= __fdiv_rn(sel_p, sel_q);
      if (!accept) {
        // residual resample on this position: sample(max_fn(p - q)) over [0, vlim)
        const int nloc = min(max(vlim - sp.g0, 0), min(sp.n, sq.n));
        float mass = 0.0f;
        ValIdx best{-1.0f, 0x7fffffff};
        for (int i = threadIdx.x; i < nloc; i += kThreads) {
          const float p = __fdiv_rn(keep_e(xp[i], stp.keep_bits), stp.Z);
          const float q = __fdiv_rn(keep_e(xq[i], stq.keep_bits), stq.Z);
          const float d = __fsub_rn(p, q);
          if (d > 0.0f) {
            mass += d;
            const int g = sp.g0 + i;
            const float key = __fdiv_rn(d, nz[g]);
            if (key > best.v || (key == best.v && g < best.i)) best = ValIdx{key, g};
          }
        }
        mass = team_allreduce<C>(mass, OpSumF(), xc);
        if (mass > 0.0f) {
          best = team_allreduce<C>(best, OpArgMax(), xc);
          cand = best.i == 0x7fffffff ? 0 : best.i;
        } else {
          // zero residual mass: one-hot at argmax(p - q), first index on ties
          ValIdx bd{-INFINITY, 0x7fffffff};
          for (int i = threadIdx.x; i < nloc; i += kThreads) {
            const float p = __fdiv_rn(keep_e(xp[i], stp.keep_bits), stp.Z);
            const float q = __fdiv_rn(keep_e(xq[i], stq.keep_bits), stq.Z);
            float d = __fsub_rn(p, q);
            if (!(d == d) || d == INFINITY || d == -INFINITY) d = 0.0f;  // nan_to_num
            const int g = sp.g0 + i;
            if (d > bd.v || (d == bd.v && g < bd.i)) bd = ValIdx{d, g};
          }
          bd = team_allreduce<C>(bd, OpArgMax(), xc);
          cand = bd.i == 0x7fffffff ? 0 : bd.i;
        }
      }
    }
  }

  // publish this position, then let the last team of the sequence finish it
  if (rank == 0 && threadIdx.x == 0) {
    if (!bonus) {
      P.accept_mask[(long long)b * P.gamma + j] = (unsigned char)accept;
      P.sel_p[(long long)b * P.gamma + j] = sel_p;
      P.sel_q[(long long)b * P.gamma + j] = sel_q;
    }
    // accepted positions carry -1; the token slot doubles as the accept flag
    volatile long long* pt = P.pair_token + (long long)b * (P.gamma + 1);
    pt[j] = (bonus || !accept) ? cand : -1;
    __threadfence();
    const int done = atomicAdd(&P.counters[b], 1);
    if (done == P.gamma) {
      __threadfence();
      int a = 0;
      while (a < P.gamma && pt[a] < 0) ++a;
      P.accepted_len[b] = a;
      P.token[b] = pt[a];
      P.all_accepted[b] = (unsigned char)(a == P.gamma);
      P.counters[b] = 0;
    }
  }
  if constexpr (C > 1) Team<C>::sync();  // no block may exit while a peer can still read its memory
}

}  // namespace duodec
