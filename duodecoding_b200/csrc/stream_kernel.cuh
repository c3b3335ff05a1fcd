// Persistent row-stream verification kernel (the fast path of duodec_verify_fused).
//
// Two 544-thread blocks per SM walk a list of ROW jobs (draft row j, target row
// j, ..., bonus target row of every sequence).  In each block one producer warp
// keeps a ring of 8 KB shared-memory chunks full with TMA bulk copies
// (full/empty mbarriers); 512 consumer threads read every chunk exactly once -
// one float4 per thread - and keep NOTHING V-sized: a single pass yields
//
//   max, Z = sum 2^(x*log2e - M0)          (M0: max*log2e predicted from the last row of the
//                                           same model; any value within 2^32 of the true max
//                                           is exact enough, else the row is redone)
//   for the nucleus (top-p): the mass at or above a predicted window [th_lo, th_hi) of scaled
//   logits and the window's entries themselves, appended to per-thread shared-memory columns
//   (no atomics; a small atomic overflow list catches threads that see more than 12).
//
// If the nucleus boundary falls inside the window (the common case once the
// prediction is warm), it is narrowed on the few window entries held in
// registers (three thresholds per pass, interpolated) until <= 160 remain;
// those are ordered exactly like the reference's sort (scaled logit descending,
// lowest index first) by an all-pairs rank, giving the last kept entry
// (xstar, istar) and the kept mass Z'.  Otherwise the window is moved and the row
// is re-read through L2 (cold start, or a row unlike its predecessors).
//
// A row job ends with a 32-byte record (M0, Z', boundary, probability of the
// drafted token) in global memory and ONE atomic on its sequence's counter; the
// block that finishes the last row of a sequence evaluates the gamma accept
// tests from the records, scans them, and draws the one token the sequence needs
// (residual resample at the first rejection, or the bonus sample) by re-reading
// two rows through L2 and the sequence's noise row from HBM.
// HBM traffic: every logit once, plus one noise row per sequence.
//
// Reference semantics restated (file:line under /root/reference):
//   logits / T -> top-p filter -> softmax                src/utils.py:650-709
//   accept test r <= p/q, leading run                   src/decoding_ops.py:127-147
//   residual resample sample(max_fn(p - q))             src/decoding_ops.py:190-206, src/utils.py:787-810
//   bonus sample sample(p_next)                         src/decoding_ops.py:209-220, src/utils.py:727-748
#pragma once

#include "stream_common.cuh"

namespace duodec {

struct RowRecord {  // 32 bytes per row job, written once, read by the block that closes the sequence
  float m0;     // exponent offset: e_i = 2^(x_i*log2e - m0)
  float Z;      // mass of the kept entries (same scale)
  float xstar;  // last kept entry in (scaled logit descending, index ascending) order;
  int istar;    //   (-inf, INT_MAX) keeps everything
  float sel;    // probability of the drafted token of this position
  float mxl;    // row maximum * log2(e)            } statistics the next launch's
  float dens;   // list entries per unit of logit    } prediction is built from
  int pad = 0;  // row-capture kernel: 1 + number of kept entries in the row's workspace list (0: no list)
};

struct StreamHint {  // per row kind, lives in the workspace, survives across launches
  float m0;
  float theta;  // scaled logit of the last nucleus boundary
  float dens;   // entries per unit of scaled logit around it
  int valid;
  float hw;     // half-width of the predicted window (split kernels; 0: derive from dens)
  int pad[3];
};

struct StreamParams {
  const void* draft;
  const void* target;
  long long d_sb, d_ss, t_sb, t_ss;
  int B, gamma, V;
  const long long* tokens;
  long long tok_sb;
  const float* r;
  long long r_sb;
  const float* noise;
  long long noise_sb;
  float t_temp, t_top_p, d_temp, d_top_p;
  unsigned char* accept_mask;
  float* sel_p;
  float* sel_q;
  long long* accepted_len;
  long long* token;
  unsigned char* all_accepted;
  int* counters;       // [B]  zeroed, self-cleaning
  StreamHint* hints;   // [2]  zeroed once (valid = 0)
  RowRecord* records;  // [B, 2*gamma+1]
  int* stats;          // optional [32] counters (tuning hook), may be null
};

struct RowJob {
  int b, j, is_target;
};

__device__ __forceinline__ RowJob decode_job(int job, int gamma) {
  const int per_b = 2 * gamma + 1;
  const int b = job / per_b;
  const int k = job - b * per_b;
  if (k == 2 * gamma) return RowJob{b, gamma, 1};
  return RowJob{b, k >> 1, k & 1};  // draft row first, then its target row
}

template <class TIn>
__device__ __forceinline__ const TIn* job_row(const StreamParams& P, RowJob jb) {
  return jb.is_target
             ? reinterpret_cast<const TIn*>(P.target) + (long long)jb.b * P.t_sb + (long long)jb.j * P.t_ss
             : reinterpret_cast<const TIn*>(P.draft) + (long long)jb.b * P.d_sb + (long long)jb.j * P.d_ss;
}

__device__ __forceinline__ float ex2_approx(float t) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(t));
  return y;
}
// e = 2^(x*log2e - m0): one FFMA + one MUFU.  m0 is a per-row constant, so its rounding is a
// common factor of the whole row and cancels in e / Z.
__device__ __forceinline__ float row_exp(float x, float m0) { return ex2_approx(fmaf(x, kLog2e, -m0)); }

// keep-rule of a finished row: at or before (xstar, istar) in the reference's sort order
__device__ __forceinline__ bool kept(const RowRecord& rc, float x, int idx) {
  return x > rc.xstar || (x == rc.xstar && idx <= rc.istar);
}

__device__ __forceinline__ RowRecord load_record(const RowRecord* p) {
  const uint4* src = reinterpret_cast<const uint4*>(p);
  const uint4 a = __ldcg(src), b4 = __ldcg(src + 1);
  RowRecord r;
  r.m0 = __uint_as_float(a.x);
  r.Z = __uint_as_float(a.y);
  r.xstar = __uint_as_float(a.z);
  r.istar = int(a.w);
  r.sel = __uint_as_float(b4.x);
  r.mxl = __uint_as_float(b4.y);
  r.dens = __uint_as_float(b4.z);
  r.pad = int(b4.w);
  return r;
}
__device__ __forceinline__ void store_record(RowRecord* p, const RowRecord& r) {
  uint4* dst = reinterpret_cast<uint4*>(p);
  dst[0] = make_uint4(__float_as_uint(r.m0), __float_as_uint(r.Z), __float_as_uint(r.xstar), unsigned(r.istar));
  dst[1] = make_uint4(__float_as_uint(r.sel), __float_as_uint(r.mxl), __float_as_uint(r.dens), unsigned(r.pad));
}

struct MassBest {
  float mass;
  float key;
  int idx;
};
struct OpMassBest {
  __device__ __forceinline__ MassBest operator()(MassBest a, MassBest b) const {
    MassBest o;
    o.mass = a.mass + b.mass;
    const bool tb = b.key > a.key || (b.key == a.key && b.idx < a.idx);
    o.key = tb ? b.key : a.key;
    o.idx = tb ? b.idx : a.idx;
    return o;
  }
};

__device__ __forceinline__ float& f4c(float4& v, int c) {
  return c == 0 ? v.x : (c == 1 ? v.y : (c == 2 ? v.z : v.w));
}

template <class TIn>
__device__ __forceinline__ float4 unpack4(const void* p16) {
  float4 v;
  if constexpr (sizeof(TIn) == 4) {
    v = *reinterpret_cast<const float4*>(p16);
  } else {
    const uint2 u = *reinterpret_cast<const uint2*>(p16);
    v.x = __uint_as_float(u.x << 16);
    v.y = __uint_as_float(u.x & 0xffff0000u);
    v.z = __uint_as_float(u.y << 16);
    v.w = __uint_as_float(u.y & 0xffff0000u);
  }
  return v;
}

template <class TIn>
__device__ __forceinline__ float4 load4_scaled(const TIn* __restrict__ p, float temperature) {
  float4 v;
  if constexpr (sizeof(TIn) == 4) {
    v = __ldcg(reinterpret_cast<const float4*>(p));
  } else {
    const uint2 u = __ldcg(reinterpret_cast<const uint2*>(p));
    v.x = __uint_as_float(u.x << 16);
    v.y = __uint_as_float(u.x & 0xffff0000u);
    v.z = __uint_as_float(u.y << 16);
    v.w = __uint_as_float(u.y & 0xffff0000u);
  }
  if (temperature != 1.0f) {
    v.x = __fdiv_rn(v.x, temperature);
    v.y = __fdiv_rn(v.y, temperature);
    v.z = __fdiv_rn(v.z, temperature);
    v.w = __fdiv_rn(v.w, temperature);
  }
  return v;
}

// ---------------------------------------------------------------------------
// one pass over a row: statistics + the window's entries
// ---------------------------------------------------------------------------
struct ScanAcc {
  float mx, z, ghi, gwin;
  int cnt;  // window entries this thread saw (the first kColCap sit in its column, the rest in the overflow list)
};

__device__ __forceinline__ void scan_elem(float x, int idx, float m0, float lo, float hi, ScanAcc& a, float* col_x,
                                          int* col_i, int* novf, StreamScratch* ss) {
  const float e = row_exp(x, m0);
  a.z += e;
#ifdef DUODEC_BRANCHY
  if (x >= hi) {
    a.ghi += e;
  } else if (x >= lo) {
#else
  const bool above = x >= hi;
  if (above) a.ghi += e;
  if (x >= lo && !above) {  // rare: a window entry
#endif
    a.gwin += e;
    if (a.cnt < kColCap) {
      col_x[a.cnt * kCT + threadIdx.x] = x;
      col_i[a.cnt * kCT + threadIdx.x] = idx;
    } else {
      const int pos = atomicAdd(novf, 1);
      if (pos < kOvfCap) {
        ss->ovf_x[pos] = x;
        ss->ovf_i[pos] = idx;
      }
    }
    ++a.cnt;
  }
}

__device__ __forceinline__ void scan4(float4 v, int i0, float m0, float lo, float hi, ScanAcc& a, float* col_x,
                                      int* col_i, int* novf, StreamScratch* ss) {
  a.mx = fmaxf(a.mx, fmaxf(fmaxf(v.x, v.y), fmaxf(v.z, v.w)));
  scan_elem(v.x, i0, m0, lo, hi, a, col_x, col_i, novf, ss);
  scan_elem(v.y, i0 + 1, m0, lo, hi, a, col_x, col_i, novf, ss);
  scan_elem(v.z, i0 + 2, m0, lo, hi, a, col_x, col_i, novf, ss);
  scan_elem(v.w, i0 + 3, m0, lo, hi, a, col_x, col_i, novf, ss);
}

// the same pass with the row re-read through L2 (window moved, or M0 far off)
template <class TIn>
__device__ __noinline__ void scan_global(const TIn* row, int V, float temp, float m0, float lo, float hi,
                                            ScanAcc& a, float* col_x, int* col_i, int* novf, StreamScratch* ss) {
  // 2048 entries per step (one float4 per thread), four steps in flight
  const int nstep = (V + 2047) / 2048;
  for (int c0 = 0; c0 < nstep; c0 += 4) {
    float4 v[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int i0 = (c0 + u) * 2048 + 4 * int(threadIdx.x);
      if (i0 < V) v[u] = load4_scaled<TIn>(row + i0, temp);
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int i0 = (c0 + u) * 2048 + 4 * int(threadIdx.x);
      if (i0 < V) scan4(v[u], i0, m0, lo, hi, a, col_x, col_i, novf, ss);
    }
  }
}

struct PassResult {
  float mx, z, ghi, gwin;
  int nwin;  // entries in the window (all threads)
  int novf;  // of which in the overflow list (> kOvfCap: the list is incomplete)
};

// pass `pp` counts its overflow pushes in ss->novf[pp & 1]; the other counter is cleared here, before
// the barrier, for the pass after this one (nobody touches it between the two barriers)
__device__ __forceinline__ PassResult reduce_pass(const ScanAcc& a, int pp, StreamScratch* ss, int& phase) {
  if (threadIdx.x == 0) {
    ss->novf[(pp + 1) & 1] = 0;
    ss->ncand = 0;
  }
  R8 r;
  r.v[0] = a.mx;
  r.v[1] = a.z;
  r.v[2] = a.ghi;
  r.v[3] = a.gwin;
  r.v[4] = float(a.cnt);
  r.v[5] = float(a.cnt > kColCap ? a.cnt - kColCap : 0);
  r.v[6] = r.v[7] = 0.0f;
  r = creduce(r, ss, phase);
  PassResult p;
  p.mx = r.v[0];
  p.z = r.v[1];
  p.ghi = r.v[2];
  p.gwin = r.v[3];
  p.nwin = int(r.v[4]);
  p.novf = int(r.v[5]);
  return p;
}

struct F4 {
  float a, b, c, d;
};
struct OpMaxF4 {
  __device__ __forceinline__ F4 operator()(F4 x, F4 y) const {
    return F4{fmaxf(x.a, y.a), fmaxf(x.b, y.b), fmaxf(x.c, y.c), fmaxf(x.d, y.d)};
  }
};
struct OpSumF4 {
  __device__ __forceinline__ F4 operator()(F4 x, F4 y) const { return F4{x.a + y.a, x.b + y.b, x.c + y.c, x.d + y.d}; }
};

// ---------------------------------------------------------------------------
// nucleus boundary of one row.  `pr` is the reduced result of the pass that just
// ran over the window [lo, hi) with this thread's counters in `acc`.
// On return rec.{Z, xstar, istar} hold the keep rule; *theta_out / *dens_out feed the hint.
// ---------------------------------------------------------------------------
template <class TIn>
__device__ __noinline__ void resolve_nucleus(const TIn* row, int V, float temp, float top_p, float m0,
                                                PassResult pr, float lo, float hi, float dens, int& pp,
                                                ScanAcc& acc, float* col_x, int* col_i, StreamScratch* ss, int& phase,
                                                RowRecord& rec, float* theta_out, float* dens_out, int* stats) {
  const int tid = threadIdx.x;
  const double target = double(top_p) * double(pr.z);
  *theta_out = pr.mx;
  *dens_out = dens;
  if (!(target < double(pr.z))) return;  // top_p >= 1: nothing is cut

  // boundary entry x* lies in [A, B):  G(x >= A) = GA > target >= GB = G(x >= B)
  float A = -INFINITY, GA = pr.z, B = nextafterf(pr.mx, INFINITY), GB = 0.0f;
  int npass = 0;
  bool list_ok = false;
#pragma unroll 1
  for (;;) {
    if (double(pr.ghi) > target) {
      A = hi, GA = pr.ghi;
    } else if (double(pr.ghi + pr.gwin) > target) {
      if (lo > A) A = lo, GA = pr.ghi + pr.gwin;
      if (hi < B) B = hi, GB = pr.ghi;
      list_ok = pr.novf <= kOvfCap;
    } else if (lo < B) {
      B = lo, GB = pr.ghi + pr.gwin;
    }
    if (pr.nwin > 0 && hi - lo > 0.0f && hi - lo < 1e30f) dens = float(pr.nwin) / (hi - lo);
    if (list_ok) break;
    if (nextafterf(A, INFINITY) >= B || npass >= 48) break;  // one float value left: tie mode below
    // ---- move the window inside [A, B) and re-read the row through L2
    if (A == -INFINITY) {
      float step = 4.0f * float(1 << (npass < 6 ? npass : 6));
      if (dens > 0.0f) step = fminf(step, fmaxf(1200.0f / dens, 1e-3f));
      lo = B - step, hi = B;
    } else {
      const float W = B - A;
      const float est = dens > 0.0f ? dens * W : __fdividef(GA - GB, row_exp(A, m0));
      if (est <= 1200.0f) {
        lo = A, hi = B;
      } else {
        const float c = A + W * __fdividef(GA - float(target), GA - GB);
        const float h = 0.5f * W * __fdividef(1200.0f, est);
        lo = fmaxf(A, c - h), hi = fminf(B, c + h);
        if (!(hi > lo)) lo = A, hi = fminf(B, nextafterf(A + 2.0f * h, INFINITY));
        if (!(hi > lo)) hi = B;
      }
    }
    acc = ScanAcc{-INFINITY, 0.0f, 0.0f, 0.0f, 0};
    ++pp;
    scan_global<TIn>(row, V, temp, m0, lo, hi, acc, col_x, col_i, &ss->novf[pp & 1], ss);
    pr = reduce_pass(acc, pp, ss, phase);
    ++npass;
  }
  *dens_out = dens;

  bool solved = false;
  if (list_ok) {
    // ---- narrow [A, B) on the window entries, now in registers
    float cx[kColCap + 1], ce[kColCap + 1];
#pragma unroll
    for (int j = 0; j < kColCap; ++j) {
      const bool v = j < acc.cnt;
      cx[j] = v ? col_x[j * kCT + tid] : -INFINITY;
      ce[j] = v ? row_exp(cx[j], m0) : 0.0f;
    }
    {
      const bool v = tid < pr.novf;
      cx[kColCap] = v ? ss->ovf_x[tid] : -INFINITY;
      ce[kColCap] = v ? row_exp(cx[kColCap], m0) : 0.0f;
    }
#pragma unroll 1
    for (int it = 0; it < 24; ++it) {
      if ((GA - GB) <= kCandGoal * row_exp(A, m0)) break;
      const float W = B - A;
      const float c = A + W * __fdividef(GA - float(target), GA - GB);
      const float d = 0.125f * W;
      const float a1 = nextafterf(A, INFINITY), b1 = B;
      float t1 = fminf(fmaxf(c - d, a1), b1), t2 = fminf(fmaxf(c, a1), b1), t3 = fminf(fmaxf(c + d, a1), b1);
      if (!(t2 > t1)) t2 = A + 0.5f * W;
      if (!(t3 > t2)) t3 = A + 0.75f * W;
      F4 g{0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int j = 0; j <= kColCap; ++j) {
        if (cx[j] >= t1) g.a += ce[j];
        if (cx[j] >= t2) g.b += ce[j];
        if (cx[j] >= t3) g.c += ce[j];
      }
      g = creduce_small(g, OpSumF4(), ss, phase);
      const float oA = A, oB = B;
      // every list entry is below the window's upper edge, so G(x >= t) = (mass above the window) + list part
      const float tt[3] = {t1, t2, t3}, gg[3] = {pr.ghi + g.a, pr.ghi + g.b, pr.ghi + g.c};
#pragma unroll
      for (int k = 0; k < 3; ++k) {
        if (tt[k] > oA && tt[k] < oB) {
          if (double(gg[k]) > target) {
            if (tt[k] > A) A = tt[k], GA = gg[k];
          } else {
            if (tt[k] < B) B = tt[k], GB = gg[k];
          }
        }
      }
      if (A == oA && B == oB) break;  // float resolution reached
    }
    // ---- the bracket's entries, ordered exactly like the reference's sort
#pragma unroll
    for (int j = 0; j <= kColCap; ++j) {
      if (cx[j] >= A && cx[j] < B) {
        const int pos = atomicAdd(&ss->ncand, 1);
        if (pos < kCandMax) {
          ss->cand_x[pos] = cx[j];
          ss->cand_e[pos] = ce[j];
          ss->cand_i[pos] = j < kColCap ? col_i[j * kCT + tid] : ss->ovf_i[tid];
        }
      }
    }
    if (tid == 0) ss->res_ok = 0;
    cbar();
    const int n = ss->ncand;
    if (n <= kCandMax) {
      int Pn = 16;
      while (Pn < n) Pn <<= 1;
      const int parts = kCT / Pn;  // 2..32 threads share one candidate (same warp)
      const int t = tid / parts, part = tid - t * parts;
      double cum = 0.0;
      float ex = 0.0f, xx = 0.0f;
      int ii = 0;
      if (t < n) {
        ex = ss->cand_e[t], xx = ss->cand_x[t], ii = ss->cand_i[t];
        for (int j = part; j < n; j += parts) {
          const float xj = ss->cand_x[j];
          const int ij = ss->cand_i[j];
          if (xj > xx || (xj == xx && ij <= ii)) cum += double(ss->cand_e[j]);
        }
      }
      for (int o = parts >> 1; o > 0; o >>= 1) cum += __shfl_xor_sync(0xffffffffu, cum, o);
      const double incl = double(GB) + cum;
      if (t < n && part == 0 && incl > target && !(incl - double(ex) > target)) {
        ss->res_x = xx;
        ss->res_i = ii;
        ss->res_Z = float(incl);
        ss->res_ok = 1;
      }
      cbar();
      if (ss->res_ok) {
        solved = true;
        rec.Z = ss->res_Z;
        rec.xstar = ss->res_x;
        rec.istar = ss->res_i;
      }
    }
    if (!solved) {
      // crowded bracket (ties): bisection on K = (okey(x) << 32) | ~index over the list entries
      auto keyof = [](float x, int idx) {
        return ((unsigned long long)okey(x) << 32) | (unsigned long long)(0xffffffffu - unsigned(idx));
      };
      unsigned long long klo = (unsigned long long)okey(A) << 32, khi = (unsigned long long)okey(B) << 32;
      unsigned long long kk[kColCap + 1];
#pragma unroll
      for (int j = 0; j <= kColCap; ++j) {
        const int idx = j < kColCap ? (j < acc.cnt ? col_i[j * kCT + tid] : 0) : (tid < pr.novf ? ss->ovf_i[tid] : 0);
        kk[j] = (cx[j] >= A && cx[j] < B) ? keyof(cx[j], idx) : 0ull;
      }
      auto F = [&](unsigned long long K) {
        double s = 0.0;
#pragma unroll
        for (int j = 0; j <= kColCap; ++j)
          if (kk[j] >= K && kk[j] != 0ull) s += double(ce[j]);
        return double(GB) + creduce_small(s, OpSumD(), ss, phase);
      };
#pragma unroll 1
      while (khi - klo > 1ull) {
        const unsigned long long mid = klo + ((khi - klo) >> 1);
        if (F(mid) > target) klo = mid;
        else khi = mid;
      }
      rec.Z = float(F(klo));
      rec.xstar = okey_inv(unsigned(klo >> 32));
      rec.istar = int(0xffffffffu - unsigned(klo & 0xffffffffull));
      solved = true;
    }
  } else {
    // one float value A holds more entries than the list can: count them in index order
    const float eA = row_exp(A, m0);
    const int nstep = (V + 2047) / 2048;
    int ilo = -1, ihi = V - 1;  // smallest I with GB + eA * #{x == A, idx <= I} > target
    auto C = [&](int I) {
      float cnt = 0.0f;
      for (int c = 0; c < nstep; ++c) {
        const int i0 = c * 2048 + 4 * tid;
        if (i0 < V) {
          float4 v = load4_scaled<TIn>(row + i0, temp);
#pragma unroll
          for (int k = 0; k < 4; ++k)
            if (f4c(v, k) >= A && f4c(v, k) < B && i0 + k <= I) cnt += 1.0f;
        }
      }
      return creduce_small(cnt, OpSumF(), ss, phase);
    };
#pragma unroll 1
    while (ihi - ilo > 1) {
      const int mid = ilo + ((ihi - ilo) >> 1);
      if (double(GB) + double(eA) * double(C(mid)) > target) ihi = mid;
      else ilo = mid;
    }
    rec.Z = float(double(GB) + double(eA) * double(C(ihi)));
    rec.xstar = A;
    rec.istar = ihi;
  }
  *theta_out = rec.xstar;
  if (stats && tid == 0) {
    atomicAdd(&stats[npass < 15 ? npass : 15], 1);  // [0..15] rows by number of extra (L2) passes
    atomicAdd(&stats[16], 1);
    if (!list_ok) atomicAdd(&stats[17], 1);
    atomicAdd(&stats[18], ss->ncand);
  }
  cbar();  // candidate area and result slots are free for the next row
}

// ---------------------------------------------------------------------------
// the one token a sequence needs: residual resample at position n (< gamma) or
// the bonus sample (n == gamma).  Rows are re-read through L2, noise from HBM.
// ---------------------------------------------------------------------------
template <class TIn>
__device__ __forceinline__ long long sample_sequence(const StreamParams& P, int b, int n, StreamScratch* ss,
                                                     int& phase) {
  const int tid = threadIdx.x;
  const int gamma = P.gamma, V = P.V;
  const int nstep = (V + 2047) / 2048;
  const bool bonus = (n == gamma);
  const int base = b * (2 * gamma + 1);
  const RowRecord rp = load_record(&P.records[base + (bonus ? 2 * gamma : 2 * n + 1)]);
  const RowRecord rq = bonus ? rp : load_record(&P.records[base + 2 * n]);
  const TIn* prow = job_row<TIn>(P, RowJob{b, n, 1});
  const TIn* qrow = bonus ? prow : job_row<TIn>(P, RowJob{b, n, 0});
  const float* nz = P.noise + (long long)b * P.noise_sb;
  const float invZp = __frcp_rn(rp.Z), invZq = __frcp_rn(rq.Z);
  MassBest mb{0.0f, -1.0f, 0x7fffffff};
  ValIdx bd{-INFINITY, 0x7fffffff};
#pragma unroll 2
  for (int c = 0; c < nstep; ++c) {  // 2048 entries per step: one float4 per thread and array
    const int i0 = c * 2048 + 4 * tid;
    if (i0 < V) {
      float4 xp = load4_scaled<TIn>(prow + i0, P.t_temp);
      float4 xq = make_float4(0.f, 0.f, 0.f, 0.f);
      if (!bonus) xq = load4_scaled<TIn>(qrow + i0, P.d_temp);
      float4 nv = __ldg(reinterpret_cast<const float4*>(nz + i0));
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const float x = f4c(xp, k);
        float d = kept(rp, x, i0 + k) ? row_exp(x, rp.m0) * invZp : 0.0f;
        if (!bonus) {
          const float y = f4c(xq, k);
          d -= kept(rq, y, i0 + k) ? row_exp(y, rq.m0) * invZq : 0.0f;
        }
        if (d > 0.0f) {
          mb.mass += d;
          const float key = __fdividef(d, f4c(nv, k));
          if (key > mb.key || (key == mb.key && i0 + k < mb.idx)) {
            mb.key = key;
            mb.idx = i0 + k;
          }
        }
        const float dz = (d == d && fabsf(d) != INFINITY) ? d : 0.0f;
        if (dz > bd.v || (dz == bd.v && i0 + k < bd.i)) bd = ValIdx{dz, i0 + k};
      }
    }
  }
  mb = creduce_small(mb, OpMassBest(), ss, phase);
  long long cand;
  if (mb.mass > 0.0f) {
    cand = mb.idx == 0x7fffffff ? 0 : mb.idx;
  } else {
    bd = creduce_small(bd, OpArgMax(), ss, phase);  // zero mass: first argmax of p - q (src/utils.py:799-806)
    cand = bd.i == 0x7fffffff ? 0 : bd.i;
  }
  return cand;
}

// accept tests of every position of sequence b from the published records, leading run, token
template <class TIn>
__device__ __noinline__ void close_sequence(const StreamParams& P, int b, StreamScratch* ss, int& phase) {
  const int tid = threadIdx.x;
  const int gamma = P.gamma;
  if (tid < gamma) {
    const int base = b * (2 * gamma + 1);
    const RowRecord rq = load_record(&P.records[base + 2 * tid]);
    const RowRecord rp = load_record(&P.records[base + 2 * tid + 1]);
    const long long o = (long long)b * gamma + tid;
    const int accept = P.r[(long long)b * P.r_sb + tid] <= __fdiv_rn(rp.sel, rq.sel);  // raw IEEE division
    P.accept_mask[o] = (unsigned char)accept;
    P.sel_p[o] = rp.sel;
    P.sel_q[o] = rq.sel;
    ss->seq_acc[tid] = accept;
  }
  cbar();
  int n = 0;
  while (n < gamma && ss->seq_acc[n] != 0) ++n;
  const long long cand = sample_sequence<TIn>(P, b, n, ss, phase);
  if (tid == 0) {
    P.accepted_len[b] = n;
    P.token[b] = cand;
    P.all_accepted[b] = (unsigned char)(n == gamma);
  }
}

// prediction for a row of `kind` (0 draft, 1 target): thread 0 only, published through the scratch
__device__ __forceinline__ void fetch_hint(const StreamParams& P, int kind, StreamScratch* ss) {
  const uint4 h = __ldcg(reinterpret_cast<const uint4*>(&P.hints[kind]));
  ss->hint_m0 = __uint_as_float(h.x);
  ss->hint_th = __uint_as_float(h.y);
  ss->hint_dens = __uint_as_float(h.z);
  ss->hint_ok = int(h.w);
}

template <class TIn>
__global__ void __launch_bounds__(kST, 2) verify_stream_kernel(const __grid_constant__ StreamParams P) {
  extern __shared__ __align__(128) unsigned char smem[];
  unsigned char* ring = smem;
  float* col_x = reinterpret_cast<float*>(smem + kSlots * kSlotBytes);
  int* col_i = reinterpret_cast<int*>(col_x + kColCap * kCT);
  StreamScratch* ss = reinterpret_cast<StreamScratch*>(col_i + kColCap * kCT);

  const int tid = threadIdx.x;
  const int gamma = P.gamma;
  const int V = P.V;
  const int njobs = P.B * (2 * gamma + 1);
  const int nchunk = (V + kChunk - 1) / kChunk;
  constexpr int kElt = int(sizeof(TIn));

  if (tid == 0) {
    for (int s = 0; s < kSlots; ++s) {
      mbar_init(&ss->full[s], 1);
      mbar_init(&ss->empty[s], kCW);
    }
    mbar_fence_init();
    ss->novf[0] = ss->novf[1] = 0;
    ss->ncand = 0;
    if (int(blockIdx.x) < njobs) fetch_hint(P, decode_job(blockIdx.x, gamma).is_target, ss);
  }
  __syncthreads();

  if (tid >= kCT) {
    // ---- producer warp: one lane streams this block's rows through the ring, in job order
    if (tid == kCT) {
      unsigned seq = 0;
      for (int job = blockIdx.x; job < njobs; job += gridDim.x) {
        const TIn* row = job_row<TIn>(P, decode_job(job, gamma));
        for (int c = 0; c < nchunk; ++c, ++seq) {
          const unsigned slot = seq % unsigned(kSlots);
          if (seq >= unsigned(kSlots)) mbar_wait(&ss->empty[slot], ((seq / unsigned(kSlots)) - 1u) & 1u);
          const int e0 = c * kChunk;
          const int ne = (V - e0) < kChunk ? (V - e0) : kChunk;
          mbar_expect_tx(&ss->full[slot], unsigned(ne * kElt));
          bulk_g2s(ring + slot * kSlotBytes, row + e0, unsigned(ne * kElt), &ss->full[slot]);
        }
      }
    }
    return;
  }

  // ---- consumers
  int phase = 0, pp = 0;
  unsigned seq = 0;
  const int lane = tid & 31;
#pragma unroll 1
  for (int job = blockIdx.x; job < njobs; job += gridDim.x) {
    const RowJob jb = decode_job(job, gamma);
    const TIn* row = job_row<TIn>(P, jb);
    const float temp = jb.is_target ? P.t_temp : P.d_temp;
    const float top_p = jb.is_target ? P.t_top_p : P.d_top_p;
    const bool nucleus = top_p > 0.0f;
    long long tk0 = 0;
    if (P.stats && tid == 0) tk0 = clock64();
    auto tick = [&](int slot) {  // tuning hook: cycles per phase summed over blocks (units of 16 cycles)
      if (P.stats && tid == 0) {
        const long long t = clock64();
        atomicAdd(&P.stats[slot], int((t - tk0) >> 4));
        tk0 = t;
      }
    };

    // the drafted token's logit is needed at the end of the job: fetch it now (thread 0)
    long long tok = -1;
    float xtok = 0.0f;
#ifndef DUODEC_LATE_TOK
    if (tid == 0 && jb.j < gamma) {
      tok = P.tokens[(long long)jb.b * P.tok_sb + jb.j];
      if (tok >= 0 && tok < V) xtok = scale_logit(to_f32(__ldcg(row + tok)), temp);
    }
#endif

    // prediction (published before the last barrier): exponent offset and nucleus window
    const bool warm = ss->hint_ok != 0;
    float m0 = warm ? ss->hint_m0 : 0.0f;
    float dens = warm ? ss->hint_dens : 0.0f;
    float lo = INFINITY, hi = INFINITY;
    if (nucleus && warm && dens > 0.0f) {
      #ifndef DUODEC_WIN
#define DUODEC_WIN 1100.0f
#endif
      const float hw = fminf(fmaxf(__fdividef(DUODEC_WIN, dens), 0.02f), 6.0f);
      lo = ss->hint_th - hw, hi = ss->hint_th + hw;
    }

    // ---- the one pass over the row, chunk by chunk as the TMA copies land
    ScanAcc acc{-INFINITY, 0.0f, 0.0f, 0.0f, 0};
    ++pp;
#pragma unroll 1
    for (int c = 0; c < nchunk; ++c, ++seq) {
      const unsigned slot = seq % unsigned(kSlots);
      mbar_wait(&ss->full[slot], (seq / unsigned(kSlots)) & 1u);
      const int i0 = c * kChunk + 4 * tid;
      if (i0 < V) {
        float4 v = unpack4<TIn>(ring + slot * kSlotBytes + 4 * tid * kElt);
        if (temp != 1.0f) {
          v.x = __fdiv_rn(v.x, temp);
          v.y = __fdiv_rn(v.y, temp);
          v.z = __fdiv_rn(v.z, temp);
          v.w = __fdiv_rn(v.w, temp);
        }
        scan4(v, i0, m0, lo, hi, acc, col_x, col_i, &ss->novf[pp & 1], ss);
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&ss->empty[slot]);
    }
    tick(20);
    PassResult pr = reduce_pass(acc, pp, ss, phase);
    tick(21);
    if (!(fabsf(pr.mx * kLog2e - m0) <= 32.0f) || !(pr.z < INFINITY) || !(pr.z > 0.0f)) {
      // the predicted offset was far from this row's maximum: redo the pass with the exact one
      m0 = pr.mx * kLog2e;
      acc = ScanAcc{-INFINITY, 0.0f, 0.0f, 0.0f, 0};
      ++pp;
      scan_global<TIn>(row, V, temp, m0, lo, hi, acc, col_x, col_i, &ss->novf[pp & 1], ss);
      pr = reduce_pass(acc, pp, ss, phase);
    }

    RowRecord rec;
    rec.m0 = m0;
    rec.Z = pr.z;
    rec.xstar = -INFINITY;
    rec.istar = 0x7fffffff;
    rec.sel = 0.0f;
    rec.mxl = pr.mx * kLog2e;
    rec.dens = 0.0f;
    float theta = pr.mx;
    if (nucleus)
      resolve_nucleus<TIn>(row, V, temp, top_p, m0, pr, lo, hi, dens, pp, acc, col_x, col_i, ss, phase, rec, &theta,
                           &dens, P.stats);

    tick(22);
    // ---- publish the row; the last row of a sequence closes it
    if (tid == 0) {
#ifdef DUODEC_LATE_TOK
      if (jb.j < gamma) {
        tok = P.tokens[(long long)jb.b * P.tok_sb + jb.j];
        if (tok >= 0 && tok < V) xtok = scale_logit(to_f32(__ldcg(row + tok)), temp);
      }
#endif
      if (tok >= 0 && tok < V) {
        const float e = kept(rec, xtok, int(tok)) ? row_exp(xtok, m0) : 0.0f;
        rec.sel = __fdiv_rn(e, rec.Z);
      }
      store_record(&P.records[job], rec);
      // prediction for later rows of the same model (any finished row's values will do)
      *reinterpret_cast<uint4*>(&P.hints[jb.is_target]) =
          make_uint4(__float_as_uint(pr.mx * kLog2e), __float_as_uint(theta), __float_as_uint(dens), 1u);
      __threadfence();
      const int done = atomicAdd(&P.counters[jb.b], 1);
      const int last = (done == 2 * gamma);
      if (last) {
        P.counters[jb.b] = 0;
        __threadfence();
      }
      ss->seq_last = last;
      const int next = job + int(gridDim.x);
      if (next < njobs) fetch_hint(P, decode_job(next, gamma).is_target, ss);
    }
    cbar();
    tick(23);
    if (ss->seq_last) {
      close_sequence<TIn>(P, jb.b, ss, phase);
      tick(24);
    }
  }
}

}  // namespace duodec
