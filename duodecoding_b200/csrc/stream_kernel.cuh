// Persistent row-stream verification kernel (the fast path of duodec_verify_fused).
//
// One 1024-thread block per SM walks a list of ROW jobs (draft row j, target
// row j, ..., bonus target row of every sequence).  Rows arrive through a ring
// of 16 KB shared-memory chunks filled by TMA bulk copies that one elected
// thread keeps issuing ahead of the consumers, and are pulled into registers -
// one float4 per thread per chunk, a whole <=32768-entry row in 32 registers
// per thread.  Everything after that is register arithmetic plus a handful of
// block reductions:
//
//   max -> e_i = 2^(x_i*log2e - max*log2e) -> Z                (softmax statistics)
//   nucleus (top-p): a threshold search on the register-resident e_i - each
//     pass sums the mass at or above three candidate thresholds (2 instructions
//     per entry and threshold, no shared-memory atomics), brackets the boundary
//     by interpolation, and stops when the bracket provably holds <= 128
//     entries; those few are ordered exactly like the reference's sort (scaled
//     logit descending, lowest index first) by an all-pairs rank in shared
//     memory, which yields the last kept entry and the kept mass Z'.
//
// A row job ends with a 32-byte record (max, Z', keep rule, probability of the
// drafted token) in global memory.  The second row of a (sequence, position)
// pair to finish evaluates the accept test; the last row of a sequence to finish
// scans the accept bits and its block draws the ONE token the sequence needs
// (residual resample at the first rejection, or the bonus sample) by re-reading
// the two rows through L2 and the sequence's noise row from HBM.
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

constexpr float kLog2e = 1.4426950408889634f;
constexpr int kCandMax = 256;        // entries the all-pairs finish can order
constexpr float kCandGoal = 128.0f;  // the search stops once the bracket holds at most this many
constexpr int kSearchPasses = 12;

struct RowRecord {  // 32 bytes per row job, written once, read by the partner / the sampler
  float mL;     // max scaled logit * log2(e)
  float Z;      // mass of the kept entries
  float lo;     // e <  lo -> dropped
  float hi;     // e >= hi -> kept;  in between: the boundary rule
  float xstar;  // boundary entry (scaled logit, index): kept iff at or before it in
  int istar;    //   (logit descending, index ascending) order; mode 1: index rule only
  int mode;     // 0: order by scaled logit then index, 1: all in-between entries share one e (bisection path)
  float sel;    // probability of the drafted token of this position
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
  int* counters;          // [B]           zeroed, self-cleaning
  long long* pair_token;  // [B, gamma+1]  accept bit of every position
  int* pair_counters;     // [B, gamma]    zeroed, self-cleaning
  RowRecord* records;     // [B, 2*gamma+1]
  int* stats;             // optional [32] counters (tuning hook), may be null
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
// e = exp(x - max) as 2^(x*log2e - max*log2e): one FFMA + one MUFU.  The rounding of max*log2e is a
// common factor of the whole row and cancels in e / Z.
__device__ __forceinline__ float row_exp(float x, float mL) { return ex2_approx(fmaf(x, kLog2e, -mL)); }

// keep-rule of a finished row applied to a recomputed entry
__device__ __forceinline__ float apply_record(const RowRecord& rc, float x, float e, int idx) {
  if (e >= rc.hi) return e;
  if (!(e >= rc.lo)) return 0.0f;
  if (rc.mode == 0) return (x > rc.xstar || (x == rc.xstar && idx <= rc.istar)) ? e : 0.0f;
  return idx <= rc.istar ? e : 0.0f;
}

__device__ __forceinline__ RowRecord load_record(const RowRecord* p) {
  const uint4* src = reinterpret_cast<const uint4*>(p);
  const uint4 a = __ldcg(src), b4 = __ldcg(src + 1);
  RowRecord r;
  r.mL = __uint_as_float(a.x);
  r.Z = __uint_as_float(a.y);
  r.lo = __uint_as_float(a.z);
  r.hi = __uint_as_float(a.w);
  r.xstar = __uint_as_float(b4.x);
  r.istar = int(b4.y);
  r.mode = int(b4.z);
  r.sel = __uint_as_float(b4.w);
  return r;
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
struct F4 {
  float a, b, c, d;
};
struct OpSumF4 {
  __device__ __forceinline__ F4 operator()(F4 x, F4 y) const { return F4{x.a + y.a, x.b + y.b, x.c + y.c, x.d + y.d}; }
};

// scratch of the persistent block (after the ring)
struct __align__(16) V1Scratch {
  StreamScratch s;  // ring barriers + reduction slots
  float cand_e[kCandMax];
  float cand_x[kCandMax];
  int cand_i[kCandMax];
  float res_x;
  int res_i;
  float res_Z;
  int res_ok;
  int seq_n;        // >= 0: this block draws the token of its sequence at position seq_n
  float warm[2];    // log2(boundary e / Z) of the last draft / target row of this block
  int warm_ok[2];
};

__device__ __forceinline__ float& f4c(float4& v, int c) {
  return c == 0 ? v.x : (c == 1 ? v.y : (c == 2 ? v.z : v.w));
}

// slow path of the nucleus boundary on a register-resident row: bisection on
// K = (bits(e) << 32) | ~index restricted to [klo, khi); `above` is the mass of
// the entries with K >= khi.  Returns K of the last kept entry (0 when nothing
// is cut) and the kept mass through *Zout.
__device__ __forceinline__ unsigned long long stream_bisect(float4 (&r)[kMaxChunks], int nchunk,
                                                        unsigned long long klo, unsigned long long khi,
                                                        double above, double target, StreamScratch* ss,
                                                        int& phase, double* Zout) {
  auto F = [&](unsigned long long t) {  // mass of entries with t <= K < khi
    double acc = 0.0;
#pragma unroll
    for (int c = 0; c < kMaxChunks; ++c) {
      if (c < nchunk) {
        float4 v = r[c];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const float e = f4c(v, k);
          const unsigned idx = unsigned(c * kChunk + 4 * int(threadIdx.x) + k);
          const unsigned long long K = ((unsigned long long)__float_as_uint(e) << 32) | (0xffffffffu - idx);
          if (K >= t && K < khi && e > 0.0f) acc += double(e);
        }
      }
    }
    return sblock_allreduce(acc, OpSumD(), ss, phase);
  };
  const double all = above + F(klo);
  if (!(all > target)) {
    *Zout = all;
    return 0ull;
  }
  unsigned long long lo = klo, hi = khi;
#pragma unroll 1
  while (hi - lo > 1ull) {
    const unsigned long long mid = lo + ((hi - lo) >> 1);
    if (above + F(mid) > target) lo = mid;
    else hi = mid;
  }
  *Zout = above + F(lo);
  return lo;
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

// mass of the register-resident row at or above three thresholds (t1 <= t2 <= t3)
__device__ __forceinline__ F4 mass_above3(const float4 (&r)[kMaxChunks], int nchunk, float t1, float t2, float t3) {
  float a1 = 0.0f, a2 = 0.0f, a3 = 0.0f;
#pragma unroll
  for (int c = 0; c < kMaxChunks; ++c) {
    if (c < nchunk) {
      const float e[4] = {r[c].x, r[c].y, r[c].z, r[c].w};
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        if (e[k] >= t1) a1 += e[k];
        if (e[k] >= t2) a2 += e[k];
        if (e[k] >= t3) a3 += e[k];
      }
    }
  }
  return F4{a1, a2, a3, 0.0f};
}

// ---------------------------------------------------------------------------
// nucleus boundary of a register-resident row (e values), see the file header.
// On return rec.{Z, lo, hi, xstar, istar, mode} describe the keep rule.
// ---------------------------------------------------------------------------
template <class TIn>
__device__ __forceinline__ void nucleus_search(float4 (&r)[kMaxChunks], int nchunk, const TIn* row, float temp,
                                               float top_p, float Zall, int kind, V1Scratch* vs, int& phase,
                                               RowRecord& rec, int* stats) {
  StreamScratch* ss = &vs->s;
  const int tid = threadIdx.x;
  const double target = double(top_p) * double(Zall);
  if (!(target < double(Zall))) return;  // top_p >= 1: nothing is cut (rec keeps lo = hi = 0)

  // bracket in e-space: G(e >= lo) > target >= G(e >= hi); u = log2 of the thresholds
  float lo = 0.0f, hi = 2.0f, Glo = Zall, Ghi = 0.0f;
  float ulo = -160.0f, uhi = 1.0f;
  bool have_lo = false, have_hi = false;
  const float lgZ = __log2f(Zall);
  const bool warm = vs->warm_ok[kind] != 0;
  const float uwarm = vs->warm[kind] + lgZ;
  if (tid == 0) vs->s.ncand = 0;
  int npass = 0;

#pragma unroll 1
  for (int pass = 0; pass < kSearchPasses; ++pass) {
    if (have_lo && (Glo - Ghi) <= kCandGoal * lo) break;
    npass = pass + 1;
    float u1, u2, u3;
    if (pass == 0) {
      if (warm) {
        u1 = uwarm - 0.25f, u2 = uwarm, u3 = uwarm + 0.25f;
      } else {
        u1 = -12.0f, u2 = -8.0f, u3 = -4.0f;
      }
    } else if (!have_lo) {  // everything tried so far was too high: walk down
      const float top = have_hi ? uhi : 0.0f;
      u1 = top - 14.0f, u2 = top - 6.0f, u3 = top - 2.0f;
    } else if (!have_hi) {  // everything tried so far was too low: walk up towards e = 1
      u1 = 0.75f * ulo, u2 = 0.5f * ulo, u3 = 0.25f * ulo;
    } else {
      const float w = uhi - ulo;
      const float c = ulo + w * __fdividef(Glo - float(target), Glo - Ghi);
      const float d = 0.125f * w;
      const float a = ulo + 0.02f * w, b = uhi - 0.02f * w;
      u1 = fminf(fmaxf(c - d, a), b);
      u2 = fminf(fmaxf(c, a), b);
      u3 = fminf(fmaxf(c + d, a), b);
    }
    u1 = fminf(fmaxf(u1, -150.0f), 0.5f);
    u2 = fminf(fmaxf(u2, -150.0f), 0.5f);
    u3 = fminf(fmaxf(u3, -150.0f), 0.5f);
    const float t1 = ex2_approx(u1), t2 = ex2_approx(u2), t3 = ex2_approx(u3);
    const F4 g = sblock_allreduce(mass_above3(r, nchunk, t1, t2, t3), OpSumF4(), ss, phase);
    const float old_lo = lo, old_hi = hi;
    // ascending thresholds: the largest one still above the target raises lo, the smallest one not above lowers hi
    const float tt[3] = {t1, t2, t3}, uu[3] = {u1, u2, u3}, gg[3] = {g.a, g.b, g.c};
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      if (double(gg[k]) > target) {
        if (tt[k] > lo) lo = tt[k], ulo = uu[k], Glo = gg[k], have_lo = true;
      } else {
        if (tt[k] < hi) hi = tt[k], uhi = uu[k], Ghi = gg[k], have_hi = true;
      }
    }
    if (pass > 0 && lo == old_lo && hi == old_hi) break;  // float resolution reached: no progress possible
  }

  bool solved = false;
  if (have_lo && (Glo - Ghi) <= kCandGoal * lo) {
    // ---- collect the bracket [lo, hi) and order it exactly like the reference's sort
#pragma unroll
    for (int c = 0; c < kMaxChunks; ++c) {
      if (c < nchunk) {
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const float e = f4c(r[c], k);
          if (e >= lo && e < hi) {
            const int idx = c * kChunk + 4 * tid + k;
            const int pos = atomicAdd(&vs->s.ncand, 1);
            if (pos < kCandMax) {
              vs->cand_e[pos] = e;
              vs->cand_i[pos] = idx;
              vs->cand_x[pos] = scale_logit(to_f32(row[idx]), temp);
            }
          }
        }
      }
    }
    if (tid == 0) vs->res_ok = 0;
    __syncthreads();
    const int n = vs->s.ncand;
    if (n <= kCandMax) {
      int Pn = 32;
      while (Pn < n) Pn <<= 1;
      const int parts = kST / Pn;  // 4..32 threads share one candidate (same warp)
      const int t = tid / parts, part = tid - t * parts;
      double cum = 0.0;
      float ex = 0.0f, xx = 0.0f;
      int ii = 0;
      if (t < n) {
        ex = vs->cand_e[t], xx = vs->cand_x[t], ii = vs->cand_i[t];
        for (int j = part; j < n; j += parts) {
          const float xj = vs->cand_x[j];
          const int ij = vs->cand_i[j];
          if (xj > xx || (xj == xx && ij <= ii)) cum += double(vs->cand_e[j]);
        }
      }
      for (int o = parts >> 1; o > 0; o >>= 1) cum += __shfl_xor_sync(0xffffffffu, cum, o);
      const double incl = double(Ghi) + cum;
      if (t < n && part == 0 && incl > target && !(incl - double(ex) > target)) {
        vs->res_x = xx;
        vs->res_i = ii;
        vs->res_Z = float(incl);
        vs->res_ok = 1;
      }
      __syncthreads();
      if (vs->res_ok) {
        solved = true;
        rec.Z = vs->res_Z;
        rec.lo = lo;
        rec.hi = hi;
        rec.xstar = vs->res_x;
        rec.istar = vs->res_i;
        rec.mode = 0;
      }
    }
  }
  if (!solved) {
    // slow path: exact bisection on (e, index) keys inside the bracket found so far
    unsigned long long klo = have_lo ? (unsigned long long)__float_as_uint(lo) << 32 : 0ull;
    unsigned long long khi = (unsigned long long)__float_as_uint(hi) << 32;
    double Zk = 0.0;
    unsigned long long kstar = 0ull;
#pragma unroll 1
    for (int attempt = 0; attempt < 2 && kstar == 0ull; ++attempt) {
      // second attempt: the bracket was inconsistent (rounding), redo on the full range
      const unsigned long long a = attempt == 0 ? klo : 0ull;
      const unsigned long long b = attempt == 0 ? khi : (unsigned long long)__float_as_uint(2.0f) << 32;
      kstar = stream_bisect(r, nchunk, a, b, attempt == 0 ? double(Ghi) : 0.0, target, ss, phase, &Zk);
    }
    rec.Z = float(Zk);
    if (kstar != 0ull) {
      const float estar = __uint_as_float(unsigned(kstar >> 32));
      rec.mode = 1;
      rec.lo = estar;
      rec.hi = __uint_as_float(unsigned(kstar >> 32) + 1u);
      rec.istar = int(0xffffffffu - unsigned(kstar & 0xffffffffull));
      lo = estar;
    }
  }
  if (stats && tid == 0) {
    atomicAdd(&stats[npass < 15 ? npass : 15], 1);          // [0..15] search passes per row
    atomicAdd(&stats[solved ? 16 : 17], 1);                  // fast finish / bisection path
    if (solved) atomicAdd(&stats[18], vs->s.ncand);          // candidates ordered
  }
  if (tid == 0 && lo > 0.0f) {
    vs->warm[kind] = __log2f(lo) - lgZ;
    vs->warm_ok[kind] = 1;
  }
  __syncthreads();  // candidate area and result slots are free for the next row
}

// ---------------------------------------------------------------------------
// the one token a sequence needs: residual resample at position n (< gamma) or
// the bonus sample (n == gamma).  Rows are re-read through L2, noise from HBM.
// ---------------------------------------------------------------------------
template <class TIn>
__device__ __forceinline__ long long sample_sequence(const StreamParams& P, int b, int n, V1Scratch* vs, int& phase) {
  StreamScratch* ss = &vs->s;
  const int tid = threadIdx.x;
  const int gamma = P.gamma, V = P.V;
  const int nchunk = (V + kChunk - 1) / kChunk;
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
  for (int c = 0; c < nchunk; ++c) {
    const int i0 = c * kChunk + 4 * tid;
    if (i0 < V) {
      float4 xp = load4_scaled<TIn>(prow + i0, P.t_temp);
      float4 xq = make_float4(0.f, 0.f, 0.f, 0.f);
      if (!bonus) xq = load4_scaled<TIn>(qrow + i0, P.d_temp);
      const float4 nv = __ldg(reinterpret_cast<const float4*>(nz + i0));
      const float nn[4] = {nv.x, nv.y, nv.z, nv.w};
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const float x = f4c(xp, k);
        float d = apply_record(rp, x, row_exp(x, rp.mL), i0 + k) * invZp;
        if (!bonus) {
          const float y = f4c(xq, k);
          d -= apply_record(rq, y, row_exp(y, rq.mL), i0 + k) * invZq;
        }
        if (d > 0.0f) {
          mb.mass += d;
          const float key = __fdividef(d, nn[k]);
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
  mb = sblock_allreduce(mb, OpMassBest(), ss, phase);
  long long cand;
  if (mb.mass > 0.0f) {
    cand = mb.idx == 0x7fffffff ? 0 : mb.idx;
  } else {
    bd = sblock_allreduce(bd, OpArgMax(), ss, phase);  // zero mass: first argmax of p - q (src/utils.py:799-806)
    cand = bd.i == 0x7fffffff ? 0 : bd.i;
  }
  return cand;
}

template <class TIn>
__global__ void __launch_bounds__(kST, 1) verify_stream_kernel(const __grid_constant__ StreamParams P) {
  extern __shared__ __align__(128) unsigned char smem[];
  unsigned char* ring = smem;
  V1Scratch* vs = reinterpret_cast<V1Scratch*>(smem + kSlots * kSlotBytes);
  StreamScratch* ss = &vs->s;

  const int tid = threadIdx.x;
  const int gamma = P.gamma;
  const int V = P.V;
  const int njobs = P.B * (2 * gamma + 1);
  const int nchunk = (V + kChunk - 1) / kChunk;
  constexpr int kElt = int(sizeof(TIn));

  if (tid == 0) {
    for (int s = 0; s < kSlots; ++s) mbar_init(&ss->full[s], 1);
    mbar_fence_init();
    vs->warm_ok[0] = vs->warm_ok[1] = 0;
  }
  __syncthreads();
  int phase = 0;

  // ---- producer: thread 0 keeps up to kSlots chunks in flight, in job order
  int issue_job = blockIdx.x, issue_chunk = 0;
  unsigned issued = 0, consumed = 0;
  auto produce = [&]() {
    while (issue_job < njobs && issued - consumed < unsigned(kSlots)) {
      const TIn* row = job_row<TIn>(P, decode_job(issue_job, gamma));
      const int e0 = issue_chunk * kChunk;
      const int ne = (V - e0) < kChunk ? (V - e0) : kChunk;
      const unsigned slot = issued % unsigned(kSlots);
      mbar_expect_tx(&ss->full[slot], unsigned(ne * kElt));
      bulk_g2s(ring + slot * kSlotBytes, row + e0, unsigned(ne * kElt), &ss->full[slot]);
      ++issued;
      if (++issue_chunk == nchunk) {
        issue_chunk = 0;
        issue_job += gridDim.x;
      }
    }
  };
  if (tid == 0) produce();

#pragma unroll 1
  for (int job = blockIdx.x; job < njobs; job += gridDim.x) {
    const RowJob jb = decode_job(job, gamma);
    const TIn* row = job_row<TIn>(P, jb);
    const float temp = jb.is_target ? P.t_temp : P.d_temp;
    const float top_p = jb.is_target ? P.t_top_p : P.d_top_p;

    // ---- pull the row out of the ring into registers (scaled logits)
    float4 r[kMaxChunks];
    float mx = -INFINITY;
#pragma unroll
    for (int c = 0; c < kMaxChunks; ++c) {
      float4 v = make_float4(-INFINITY, -INFINITY, -INFINITY, -INFINITY);
      if (c < nchunk) {
        const unsigned seq = consumed + unsigned(c);
        const unsigned slot = seq % unsigned(kSlots);
        mbar_wait(&ss->full[slot], (seq / unsigned(kSlots)) & 1u);
        if (c * kChunk + 4 * tid < V) {
          const unsigned char* sp = ring + slot * kSlotBytes;
          if constexpr (kElt == 4) {
            v = reinterpret_cast<const float4*>(sp)[tid];
          } else {
            const uint2 u = reinterpret_cast<const uint2*>(sp)[tid];
            v.x = __uint_as_float(u.x << 16);
            v.y = __uint_as_float(u.x & 0xffff0000u);
            v.z = __uint_as_float(u.y << 16);
            v.w = __uint_as_float(u.y & 0xffff0000u);
          }
          if (temp != 1.0f) {
            v.x = __fdiv_rn(v.x, temp);
            v.y = __fdiv_rn(v.y, temp);
            v.z = __fdiv_rn(v.z, temp);
            v.w = __fdiv_rn(v.w, temp);
          }
        }
        mx = fmaxf(mx, fmaxf(fmaxf(v.x, v.y), fmaxf(v.z, v.w)));
      }
      r[c] = v;
    }
    consumed += unsigned(nchunk);
    const float m = sblock_allreduce(mx, OpMaxF(), ss, phase);  // its barrier also frees the ring slots
    if (tid == 0) produce();

    // ---- e = exp(x - m) in registers, Z
    const float mL = m * kLog2e;
    float z = 0.0f;
#pragma unroll
    for (int c = 0; c < kMaxChunks; ++c) {
      if (c < nchunk) {
        r[c].x = row_exp(r[c].x, mL);
        r[c].y = row_exp(r[c].y, mL);
        r[c].z = row_exp(r[c].z, mL);
        r[c].w = row_exp(r[c].w, mL);
        z += (r[c].x + r[c].y) + (r[c].z + r[c].w);
      }
    }
    const float Zall = sblock_allreduce(z, OpSumF(), ss, phase);

    RowRecord rec;
    rec.mL = mL;
    rec.Z = Zall;
    rec.lo = 0.0f;
    rec.hi = 0.0f;
    rec.xstar = -INFINITY;
    rec.istar = 0x7fffffff;
    rec.mode = 0;
    rec.sel = 0.0f;
    if (top_p > 0.0f) nucleus_search<TIn>(r, nchunk, row, temp, top_p, Zall, jb.is_target, vs, phase, rec, P.stats);

    // ---- publish the row; resolve the position / the sequence
    const long long tok = jb.j < gamma ? P.tokens[(long long)jb.b * P.tok_sb + jb.j] : -1;
    const bool tok_ok = tok >= 0 && tok < V;
    const int owner = tok_ok ? int((tok & (kChunk - 1)) >> 2) : 0;
    if (tid == owner) {
      if (tok_ok) {
        const int tc = int(tok >> 12), tk = int(tok & 3);
        float e = 0.0f;
#pragma unroll
        for (int c = 0; c < kMaxChunks; ++c)
          if (c == tc) e = f4c(r[c], tk);
        if (e < rec.hi && e >= rec.lo)  // inside the boundary bracket: the exact order decides
          e = apply_record(rec, scale_logit(to_f32(row[tok]), temp), e, int(tok));
        else if (!(e >= rec.hi))
          e = 0.0f;
        rec.sel = __fdiv_rn(e, rec.Z);
      }
      P.records[job] = rec;
      __threadfence();
      int seq_n = -1;
      bool arrive = true;
      if (jb.j < gamma) {
        const long long o = (long long)jb.b * gamma + jb.j;
        const int old = atomicAdd(&P.pair_counters[o], 1);
        arrive = (old == 1);
        if (arrive) {  // second row of the pair: accept test  r <= p / q
          __threadfence();
          const RowRecord pr = load_record(&P.records[jb.is_target ? job - 1 : job + 1]);
          const float selp = jb.is_target ? rec.sel : pr.sel;
          const float selq = jb.is_target ? pr.sel : rec.sel;
          const int accept = P.r[(long long)jb.b * P.r_sb + jb.j] <= __fdiv_rn(selp, selq);
          P.accept_mask[o] = (unsigned char)accept;
          P.sel_p[o] = selp;
          P.sel_q[o] = selq;
          P.pair_counters[o] = 0;
          P.pair_token[(long long)jb.b * (gamma + 1) + jb.j] = accept;
        }
      }
      if (arrive) {
        __threadfence();
        const int done = atomicAdd(&P.counters[jb.b], 1);
        if (done == gamma) {  // last position of the sequence: leading run of accepts
          __threadfence();
          const volatile long long* pt = P.pair_token + (long long)jb.b * (gamma + 1);
          int a = 0;
          while (a < gamma && pt[a] != 0) ++a;
          seq_n = a;
          P.counters[jb.b] = 0;
        }
      }
      vs->seq_n = seq_n;
    }
    __syncthreads();
    const int seq_n = vs->seq_n;
    if (seq_n >= 0) {
      const long long cand = sample_sequence<TIn>(P, jb.b, seq_n, vs, phase);
      if (tid == 0) {
        P.accepted_len[jb.b] = seq_n;
        P.token[jb.b] = cand;
        P.all_accepted[jb.b] = (unsigned char)(seq_n == gamma);
      }
    }
  }
}

}  // namespace duodec
