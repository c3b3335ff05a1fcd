// Persistent row-stream verification kernel (the fast path of duodec_verify_fused).
//
// One 1024-thread block per SM walks a list of ROW jobs (draft row j, target
// row j, ..., bonus target row of every sequence).  Rows arrive through an
// 11-slot ring of 16 KB shared-memory chunks filled by TMA bulk copies that one
// elected thread keeps issuing ahead of the consumers (up to 176 KB in flight
// per SM), and are pulled into registers - one float4 per thread per chunk, a
// whole <=32768-entry row in 32 registers per thread - so the logit warp and
// the softmax statistics never touch shared memory again except for the
// integer histogram of the nucleus search (rowcore.cuh).  A row job ends with
// its statistics and the probability of the drafted token in global memory; the
// second row of a (sequence, position) pair to finish evaluates the accept
// test and, on a rejection, draws the replacement token from the residual with
// its own row still in registers and the partner row re-read through L2; the
// last position of a sequence to finish scans the accept bits.
// HBM traffic: every logit once, plus one noise row per rejected / bonus position.
#pragma once

#include "stream_common.cuh"

namespace duodec {

struct RowRecord {  // 32 bytes per row job, written once, read by the partner
  float m, Z;
  unsigned hi_bits;  // e bits >  hi_bits -> kept
  unsigned lo_bits;  // e bits <= lo_bits -> dropped; in between: the boundary rule
  float xstar;       // boundary entry (scaled logit, index): kept iff before-or-at it in
  int istar;         //   (logit descending, index ascending) order; mode 1: index rule only
  int mode;          // 0: order by scaled logit, 1: all in-between entries share one e (bisection path)
  float sel;         // probability of the drafted token of this position
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
  long long* pair_token;  // [B, gamma+1]
  int* pair_counters;     // [B, gamma]    zeroed, self-cleaning
  RowRecord* records;     // [B, 2*gamma+1]
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

// keep-rule of a finished row applied to a recomputed entry
__device__ __forceinline__ float apply_record(const RowRecord& rc, float x, float e, int idx) {
  const unsigned bits = __float_as_uint(e);
  if (bits > rc.hi_bits) return e;
  if (bits <= rc.lo_bits) return 0.0f;
  if (rc.mode == 0) return (x > rc.xstar || (x == rc.xstar && idx <= rc.istar)) ? e : 0.0f;
  return idx <= rc.istar ? e : 0.0f;
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

// ---------------------------------------------------------------------------
// register-resident row helpers
// ---------------------------------------------------------------------------
__device__ __forceinline__ float& f4c(float4& v, int c) {
  return c == 0 ? v.x : (c == 1 ? v.y : (c == 2 ? v.z : v.w));
}

// slow path of the nucleus boundary on a register-resident row: bisection on
// K = (bits(e) << 32) | ~index restricted to [klo, khi); `above` is the exact
// mass of the entries with K >= khi.  Returns K of the last kept entry (0 when
// nothing is cut) and the kept mass through *Zout.
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
  v.x = scale_logit(v.x, temperature);
  v.y = scale_logit(v.y, temperature);
  v.z = scale_logit(v.z, temperature);
  v.w = scale_logit(v.w, temperature);
  return v;
}

template <class TIn>
__global__ void __launch_bounds__(kST, 1) verify_stream_kernel(const __grid_constant__ StreamParams P) {
  extern __shared__ __align__(128) unsigned char smem[];
  unsigned char* ring = smem;
  unsigned* cnt = reinterpret_cast<unsigned*>(smem + kSlots * kSlotBytes);
  unsigned* low = cnt + kHistBins;
  BandWork* bw = reinterpret_cast<BandWork*>(cnt);
  StreamScratch* ss = reinterpret_cast<StreamScratch*>(smem + kSlots * kSlotBytes + kHistBytes);

  const int tid = threadIdx.x;
  const int gamma = P.gamma;
  const int V = P.V;
  const int njobs = P.B * (2 * gamma + 1);
  const int nchunk = (V + kChunk - 1) / kChunk;
  constexpr int kElt = int(sizeof(TIn));

  if (tid == 0) {
    for (int s = 0; s < kSlots; ++s) mbar_init(&ss->full[s], 1);
    mbar_fence_init();
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
    const bool nucleus = top_p > 0.0f;

    // the histogram is cleared while the row is still landing
    if (nucleus)
      for (int i = tid; i < 2 * kHistBins; i += kST) cnt[i] = 0u;
    if (tid == 0) {
      ss->bstar = kHistBins;
      ss->ncand = 0;
      ss->jstar = 0x7fffffff;
      ss->overflow = 0;
    }

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
          v.x = scale_logit(v.x, temp);
          v.y = scale_logit(v.y, temp);
          v.z = scale_logit(v.z, temp);
          v.w = scale_logit(v.w, temp);
        }
        mx = fmaxf(mx, fmaxf(fmaxf(v.x, v.y), fmaxf(v.z, v.w)));
      }
      r[c] = v;
    }
    consumed += unsigned(nchunk);
    const float m = sblock_allreduce(mx, OpMaxF(), ss, phase);  // its barrier also frees the ring slots
    if (tid == 0) produce();

    // ---- pass 2: e = exp(x - m) in registers; Z or the exact-mass histogram
    float z = 0.0f;
#pragma unroll
    for (int c = 0; c < kMaxChunks; ++c) {
      if (c < nchunk) {
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          float& x = f4c(r[c], k);
          const float ex = expf(__fsub_rn(x, m));
          x = ex;
          if (nucleus) {
            const unsigned d = kOneBits - __float_as_uint(ex);
            const unsigned b = d >> kBinShift;
            if (b < unsigned(kHistBins)) {
              atomicAdd(&cnt[b], 1u);
              atomicAdd(&low[b], d & 0xffffu);
            } else {
              z += ex;
            }
          } else {
            z += ex;
          }
        }
      }
    }
    const float zsum = sblock_allreduce(z, OpSumF(), ss, phase);  // barrier: histogram complete

    RowRecord rec;
    rec.m = m;
    rec.Z = zsum;
    rec.hi_bits = 0u;
    rec.lo_bits = 0u;
    rec.xstar = -INFINITY;
    rec.istar = 0x7fffffff;
    rec.mode = 0;
    rec.sel = 0.0f;

    if (nucleus) {
      // exact bin masses -> boundary bin (4 bins per thread, ascending bin = descending e)
      constexpr int kPer = kHistBins / kST;
      const int b0 = tid * kPer;
      double loc = 0.0;
      bool big = false;
#pragma unroll
      for (int i = 0; i < kPer; ++i) {
        const unsigned c = cnt[b0 + i];
        big |= (c > 65535u);
        loc += bin_mass(b0 + i, c, low[b0 + i]);
      }
      double total = 0.0;
      const double run0 = sblock_exscan(loc, OpSumD(), 0.0, ss, phase, &total);
      const double Z1 = total + double(zsum);
      const double target = double(top_p) * Z1;
      if (big) ss->overflow = 1;
      {
        int mine = kHistBins;
        double run = run0;
#pragma unroll
        for (int i = 0; i < kPer; ++i) {
          run += bin_mass(b0 + i, cnt[b0 + i], low[b0 + i]);
          if (run > target && mine == kHistBins) mine = b0 + i;
        }
        if (mine < kHistBins) atomicMin(&ss->bstar, mine);
      }
      __syncthreads();
      const int bstar = ss->bstar;
      const bool have_bin = bstar < kHistBins && !ss->overflow;
      const unsigned nb = have_bin ? cnt[bstar] : 0u;
      if (have_bin && bstar >= b0 && bstar < b0 + kPer) {
        double run = run0;
        for (int i = 0; b0 + i < bstar; ++i) run += bin_mass(b0 + i, cnt[b0 + i], low[b0 + i]);
        ss->m_above = run;
      }
      __syncthreads();  // m_above visible; histogram reads done before the region is reused
      if (have_bin && nb <= unsigned(kCandCap)) {
        const double m_above = ss->m_above;
        // collect the entries of the boundary bin with their exact ordering keys
#pragma unroll
        for (int c = 0; c < kMaxChunks; ++c) {
          if (c < nchunk) {
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              const float e = f4c(r[c], k);
              if (int((kOneBits - __float_as_uint(e)) >> kBinShift) == bstar) {
                const int idx = c * kChunk + 4 * tid + k;
                const int pos = atomicAdd(&ss->ncand, 1);
                const float x = scale_logit(to_f32(row[idx]), temp);
                bw->key[pos] = ((unsigned long long)(~okey(x)) << 32) | (unsigned long long)unsigned(idx);
                bw->mass[pos] = e;
              }
            }
          }
        }
        __syncthreads();
        const int n = ss->ncand;
        int Pn = 32;
        while (Pn < n) Pn <<= 1;
        if (tid >= n && tid < Pn) {
          bw->key[tid] = ~0ull;
          bw->mass[tid] = 0.0f;
        }
        __syncthreads();
        for (int k2 = 2; k2 <= Pn; k2 <<= 1) {
          for (int j2 = k2 >> 1; j2 > 0; j2 >>= 1) {
            const int t = tid, u = tid ^ j2;
            if (t < Pn && u > t) {
              const unsigned long long a = bw->key[t], b = bw->key[u];
              const bool up = ((t & k2) == 0);
              if ((a > b) == up) {
                bw->key[t] = b;
                bw->key[u] = a;
                const float ma = bw->mass[t];
                bw->mass[t] = bw->mass[u];
                bw->mass[u] = ma;
              }
            }
            __syncthreads();
          }
        }
        const double mine_mass = tid < Pn ? double(bw->mass[tid]) : 0.0;
        const double ex = sblock_exscan(mine_mass, OpSumD(), 0.0, ss, phase, static_cast<double*>(nullptr));
        const double cum = m_above + ex + mine_mass;
        if (tid < n && cum > target) atomicMin(&ss->jstar, tid);
        __syncthreads();
        const int jstar = ss->jstar;
        const int nkeep = jstar == 0x7fffffff ? n : jstar + 1;
        if (tid == nkeep - 1) {
          ss->m_above = cum;
          ss->keystar = bw->key[tid];
        }
        __syncthreads();
        const unsigned long long keystar = ss->keystar;
        rec.Z = float(ss->m_above);
        rec.hi_bits = bin_top_bits(bstar);
        rec.lo_bits = bin_top_bits(bstar + 1);
        rec.xstar = okey_inv(~unsigned(keystar >> 32));
        rec.istar = int(unsigned(keystar & 0xffffffffull));
        // drop this thread's boundary-bin entries that come after the boundary
#pragma unroll
        for (int c = 0; c < kMaxChunks; ++c) {
          if (c < nchunk) {
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              float& e = f4c(r[c], k);
              const unsigned bits = __float_as_uint(e);
              if (bits <= rec.lo_bits) {
                e = 0.0f;
              } else if (bits <= rec.hi_bits) {
                const int idx = c * kChunk + 4 * tid + k;
                const float x = scale_logit(to_f32(row[idx]), temp);
                if (!(x > rec.xstar || (x == rec.xstar && idx <= rec.istar))) e = 0.0f;
              }
            }
          }
        }
      } else {
        // slow path: boundary below the histogram range, or a crowded boundary bin
        unsigned long long klo = 0ull, khi = (unsigned long long)(kOneBits + 1u) << 32;
        double above = 0.0;
        if (have_bin) {
          khi = (unsigned long long)(bin_top_bits(bstar) + 1u) << 32;
          klo = (unsigned long long)(bin_top_bits(bstar + 1) + 1u) << 32;
          above = ss->m_above;
        }
        double Zk = 0.0;
        unsigned long long kstar = stream_bisect(r, nchunk, klo, khi, above, target, ss, phase, &Zk);
        if (kstar == 0ull && have_bin) {
          // nothing inside the bin crosses the target: the cut is below it (cannot
          // happen when the bin was located exactly); redo on the full range
          kstar = stream_bisect(r, nchunk, 0ull, (unsigned long long)(kOneBits + 1u) << 32, 0.0, target, ss,
                                phase, &Zk);
        }
        rec.Z = float(Zk);
        if (kstar != 0ull) {
          rec.mode = 1;
          rec.hi_bits = unsigned(kstar >> 32);
          rec.lo_bits = rec.hi_bits - 1u;
          rec.istar = int(0xffffffffu - unsigned(kstar & 0xffffffffull));
#pragma unroll
          for (int c = 0; c < kMaxChunks; ++c) {
            if (c < nchunk) {
#pragma unroll
              for (int k = 0; k < 4; ++k) {
                float& e = f4c(r[c], k);
                const unsigned idx = unsigned(c * kChunk + 4 * tid + k);
                const unsigned long long K = ((unsigned long long)__float_as_uint(e) << 32) | (0xffffffffu - idx);
                if (K < kstar) e = 0.0f;
              }
            }
          }
        }
      }
    }
    __syncthreads();  // the work region and the scratch flags are free for the next use

    // ---- publish the row, resolve the position / the sequence
    const long long tok = jb.j < gamma ? P.tokens[(long long)jb.b * P.tok_sb + jb.j] : -1;
    const bool tok_ok = tok >= 0 && tok < V;
    const int owner = tok_ok ? int((tok & (kChunk - 1)) >> 2) : 0;
    const float* nz = P.noise + (long long)jb.b * P.noise_sb;
    if (tid == owner) {
      if (tok_ok) {
        const int tc = int(tok >> 12), tk = int(tok & 3);
        float e = 0.0f;
#pragma unroll
        for (int c = 0; c < kMaxChunks; ++c)
          if (c == tc) e = f4c(r[c], tk);
        rec.sel = __fdiv_rn(e, rec.Z);
      }
      int old = 0;
      if (jb.j < gamma) {
        P.records[job] = rec;
        __threadfence();
        old = atomicAdd(&P.pair_counters[jb.b * gamma + jb.j], 1);
        __threadfence();
      }
      ss->flag = old;
    }
    __syncthreads();
    const bool second = (jb.j < gamma) && (ss->flag == 1);
    const bool bonus = (jb.j == gamma);
    if (second || bonus) {
      long long cand = -1;
      int accept = 0;
      float sel_mine = 0.0f, sel_other = 0.0f;
      if (bonus) {
        // bonus token: argmax_i p_i / noise_i
        ValIdx best{-1.0f, 0x7fffffff};
#pragma unroll
        for (int c = 0; c < kMaxChunks; ++c) {
          if (c < nchunk && c * kChunk + 4 * tid < V) {
            const float4 nv = __ldg(reinterpret_cast<const float4*>(nz + c * kChunk + 4 * tid));
            const float nn[4] = {nv.x, nv.y, nv.z, nv.w};
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              const float p = __fdiv_rn(f4c(r[c], k), rec.Z);
              if (p > 0.0f) {
                const int idx = c * kChunk + 4 * tid + k;
                const float key = __fdiv_rn(p, nn[k]);
                if (key > best.v || (key == best.v && idx < best.i)) best = ValIdx{key, idx};
              }
            }
          }
        }
        best = sblock_allreduce(best, OpArgMax(), ss, phase);
        cand = best.i == 0x7fffffff ? 0 : best.i;
      } else {
        const int pjob = jb.is_target ? job - 1 : job + 1;
        RowRecord pr;
        {
          const uint4* src = reinterpret_cast<const uint4*>(&P.records[pjob]);
          const uint4 a = __ldcg(src), b4 = __ldcg(src + 1);
          pr.m = __uint_as_float(a.x);
          pr.Z = __uint_as_float(a.y);
          pr.hi_bits = a.z;
          pr.lo_bits = a.w;
          pr.xstar = __uint_as_float(b4.x);
          pr.istar = int(b4.y);
          pr.mode = int(b4.z);
          pr.sel = __uint_as_float(b4.w);
        }
        // every thread needs sel of this row: broadcast it from the owner through the record
        const float my_sel = __uint_as_float(__ldcg(reinterpret_cast<const unsigned*>(&P.records[job].sel)));
        sel_mine = my_sel;
        sel_other = pr.sel;
        const float selp = jb.is_target ? sel_mine : sel_other;
        const float selq = jb.is_target ? sel_other : sel_mine;
        accept = P.r[(long long)jb.b * P.r_sb + jb.j] <= __fdiv_rn(selp, selq);
        if (!accept) {
          // residual resample: sample(max_fn(p - q)); this row from registers, the partner through L2
          const RowJob oj = decode_job(pjob, gamma);
          const TIn* orow = job_row<TIn>(P, oj);
          const float otemp = oj.is_target ? P.t_temp : P.d_temp;
          MassBest mb{0.0f, -1.0f, 0x7fffffff};
          ValIdx bd{-INFINITY, 0x7fffffff};
#pragma unroll
          for (int c = 0; c < kMaxChunks; ++c) {
            if (c < nchunk && c * kChunk + 4 * tid < V) {
              const int i0 = c * kChunk + 4 * tid;
              float4 ov = load4_scaled<TIn>(orow + i0, otemp);
              const float4 nv = __ldg(reinterpret_cast<const float4*>(nz + i0));
              const float nn[4] = {nv.x, nv.y, nv.z, nv.w};
#pragma unroll
              for (int k = 0; k < 4; ++k) {
                const float xo = f4c(ov, k);
                const float eo = apply_record(pr, xo, expf(__fsub_rn(xo, pr.m)), i0 + k);
                const float po = __fdiv_rn(eo, pr.Z);
                const float pm = __fdiv_rn(f4c(r[c], k), rec.Z);
                const float d = jb.is_target ? __fsub_rn(pm, po) : __fsub_rn(po, pm);
                if (d > 0.0f) {
                  mb.mass += d;
                  const float key = __fdiv_rn(d, nn[k]);
                  if (key > mb.key || (key == mb.key && i0 + k < mb.idx)) {
                    mb.key = key;
                    mb.idx = i0 + k;
                  }
                }
                const float dz = (d == d && d != INFINITY && d != -INFINITY) ? d : 0.0f;
                if (dz > bd.v || (dz == bd.v && i0 + k < bd.i)) bd = ValIdx{dz, i0 + k};
              }
            }
          }
          mb = sblock_allreduce(mb, OpMassBest(), ss, phase);
          if (mb.mass > 0.0f) {
            cand = mb.idx == 0x7fffffff ? 0 : mb.idx;
          } else {
            bd = sblock_allreduce(bd, OpArgMax(), ss, phase);  // zero mass: first argmax of p - q
            cand = bd.i == 0x7fffffff ? 0 : bd.i;
          }
        }
      }
      if (tid == 0) {
        const int b = jb.b;
        volatile long long* pt = P.pair_token + (long long)b * (gamma + 1);
        if (!bonus) {
          const long long o = (long long)b * gamma + jb.j;
          P.accept_mask[o] = (unsigned char)accept;
          P.sel_p[o] = jb.is_target ? sel_mine : sel_other;
          P.sel_q[o] = jb.is_target ? sel_other : sel_mine;
          P.pair_counters[o] = 0;
        }
        pt[jb.j] = (bonus || !accept) ? cand : -1;
        __threadfence();
        const int done = atomicAdd(&P.counters[b], 1);
        if (done == gamma) {
          __threadfence();
          int a = 0;
          while (a < gamma && pt[a] < 0) ++a;
          P.accepted_len[b] = a;
          P.token[b] = pt[a];
          P.all_accepted[b] = (unsigned char)(a == gamma);
          P.counters[b] = 0;
        }
      }
    }
  }
}

}  // namespace duodec
