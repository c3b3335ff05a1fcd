// Per-row "logit warp + softmax statistics" on a team-staged row.
//
// Restates, for one row, what the reference computes with
//   logits / T  ->  top_k_top_p_filter  ->  softmax       (src/utils.py:650-709)
// in ONE pass over HBM: the row slice sits in shared memory as fp32 scaled
// logits x_i; the core turns it in place into e_i = exp(x_i - max) with every
// filtered entry either zeroed or below `keep_bits`, and returns the max and the
// normaliser Z of the kept entries, so that  p_i = keep(e_i) / Z.
//
// Top-p (nucleus) without sorting the row:
//   1. integer histogram of the e_i over 4096 log-spaced bins taken straight
//      from the float bit pattern d = bits(1.0f) - bits(e): bin = d >> 16 (128
//      bins per binade), and next to the count of each bin the integer sum of
//      the low 16 bits of d.  Inside a bin e is linear in d, so
//          mass(bin) = count * e_top(bin) - ulp(bin) * sum_low(bin)
//      is EXACT: two native shared-memory integer atomics per entry give the
//      exact mass of every bin, independent of summation order, team size
//      and launch.
//   2. one block scan over the 4096 bin masses (double) finds the bin b* that
//      holds the nucleus boundary and the exact mass above it; bins above b*
//      are kept, bins below are dropped;
//   3. the few entries of b* are ordered exactly like the reference's sort
//      (scaled logit descending, lowest index first on ties), their masses are
//      prefix-summed after the mass above, and the first entry whose inclusive
//      mass exceeds top_p * Z1 is the last one kept
//      (src/utils.py:669-681: drop sorted position j when cum[j-1] > top_p).
//   A boundary below 2^-32 of the row maximum, or a boundary bin with more
//   than kCandCap entries (near-uniform rows), takes a slow bisection on
//   (e, index) keys instead.
#pragma once

#include "scan.cuh"

namespace duodec {

constexpr int kHistBins = 4096;                // 32 binades x 128 bins
constexpr int kBinShift = 16;                  // low 16 bits of d stay inside a bin
constexpr int kHistBytes = 2 * kHistBins * 4;  // counts + low sums = 32 KB, reused for the band work
constexpr int kCandCap = 1024;
constexpr unsigned kOneBits = 0x3F800000u;

// Layout of the 32 KB work region once the histogram is no longer needed.
struct BandWork {
  int cand_local[kCandCap];         // local indices of this block's band entries
  unsigned long long key[kCandCap];  // (~okey(x) << 32) | global index : ascending == reference order
  float mass[kCandCap];              // e of the entry
  double cum[kCandCap];              // inclusive prefix mass in sorted order
};
static_assert(sizeof(BandWork) <= kHistBytes, "band work must fit the histogram region");

struct RowScratch {
  Xchg xc;
  unsigned long long bar;   // mbarriers for TMA bulk staging
  unsigned long long bar2;
  int res[4];
  int bstar, ncand, jstar, overflow;
  double m_above;
};

struct RowStat {
  float m;             // max scaled logit
  float Z;             // normaliser of the kept entries
  unsigned keep_bits;  // entry kept iff float bits of e  > keep_bits
};

__device__ __forceinline__ float keep_e(float e, unsigned keep_bits) {
  return __float_as_uint(e) > keep_bits ? e : 0.0f;
}

// bin b holds the float bit patterns kOneBits - d with d in [b<<16, (b<<16) + 0xffff]
__device__ __forceinline__ unsigned bin_top_bits(int b) { return kOneBits - (unsigned(b) << kBinShift); }
// exact mass of a bin from its count and the sum of the low bits of d
__device__ __forceinline__ double bin_mass(int b, unsigned cnt, unsigned low) {
  const unsigned top = bin_top_bits(b);
  const double e_top = double(__uint_as_float(top));
  const double ulp = e_top - double(__uint_as_float(top - 1u));
  return double(cnt) * e_top - ulp * double(low);
}

// ---------------------------------------------------------------------------
// argmax of a staged row (temperature == 0 path, src/utils.py:700-704)
// ---------------------------------------------------------------------------
template <int C>
__device__ __forceinline__ ValIdx row_argmax(const float* xs, Slice s, Xchg* xc) {
  ValIdx best{-INFINITY, 0x7fffffff};
  for (int i = threadIdx.x; i < s.n; i += kThreads) {
    const float v = xs[i];
    if (v > best.v || (v == best.v && s.g0 + i < best.i)) best = ValIdx{v, s.g0 + i};
  }
  ValIdx r = team_allreduce<C>(best, OpArgMax(), xc);
  if (r.i == 0x7fffffff) r.i = 0;  // all NaN / -inf row: torch.argmax gives some index; pick 0
  return r;
}

// ---------------------------------------------------------------------------
// slow path for the nucleus boundary: bisection on K = (bits(e) << 32) | ~index
// ---------------------------------------------------------------------------
template <int C>
__device__ __noinline__ RowStat nucleus_bisect(float* xs, Slice s, int n4, float m, double target, Xchg* xc) {
  // F(t) = sum of e_i with K_i >= t is non-increasing in t.  Find the largest t
  // with F(t) > target: that entry is the last one kept.
  unsigned long long lo = 0ull, hi = (unsigned long long)(kOneBits + 1u) << 32;
  auto F = [&](unsigned long long t) {
    double acc = 0.0;
    for (int i = threadIdx.x; i < s.n; i += kThreads) {
      const float e = xs[i];
      const unsigned long long K =
          ((unsigned long long)__float_as_uint(e) << 32) | (unsigned long long)(0xffffffffu - unsigned(s.g0 + i));
      if (K >= t && e > 0.0f) acc += double(e);
    }
    return team_allreduce<C>(acc, OpSumD(), xc);
  };
  const double all = F(0ull);
  if (!(all > target)) return RowStat{m, float(all), 0u};  // nothing is cut
#pragma unroll 1
  while (hi - lo > 1ull) {
    const unsigned long long mid = lo + ((hi - lo) >> 1);
    if (F(mid) > target) lo = mid;
    else hi = mid;
  }
  const double Z = F(lo);
  for (int i = threadIdx.x; i < s.n; i += kThreads) {
    const float e = xs[i];
    const unsigned long long K =
        ((unsigned long long)__float_as_uint(e) << 32) | (unsigned long long)(0xffffffffu - unsigned(s.g0 + i));
    if (K < lo) xs[i] = 0.0f;
  }
  Team<C>::sync();
  return RowStat{m, float(Z), 0u};
}

// ---------------------------------------------------------------------------
// the core
// ---------------------------------------------------------------------------
// xs     : this block's slice, fp32 scaled logits padded with -inf to n4*4
// per    : slice pitch of the team (entry g lives in block g / per)
// row    : the original global row (boundary-bin entries are re-read from it so
//          that their order is the order of the scaled logits, not of the exps)
// work   : kHistBytes of shared memory
template <int C, class TIn>
__device__ __forceinline__ RowStat row_core(float* xs, Slice s, int per, int V, const TIn* __restrict__ row,
                                            float temperature, int top_k, float top_p, void* work,
                                            RowScratch* sc) {
  Xchg* xc = &sc->xc;
  const int n4 = (s.n + 3) >> 2;
  float4* x4 = reinterpret_cast<float4*>(xs);
  unsigned* cnt = reinterpret_cast<unsigned*>(work);
  unsigned* low = cnt + kHistBins;

  // pass 1: max
  float mx = -INFINITY;
  for (int i = threadIdx.x; i < n4; i += kThreads) {
    const float4 v = x4[i];
    mx = fmaxf(mx, fmaxf(fmaxf(v.x, v.y), fmaxf(v.z, v.w)));
  }
  const float m = team_allreduce<C>(mx, OpMaxF(), xc);

  // top-k: keep every entry >= the k-th largest (ties survive, src/utils.py:665-666)
  float tau_k = -INFINITY;
  if (top_k > 0 && top_k < V)
    tau_k = team_kth_largest<C>(xs, n4, top_k, reinterpret_cast<int*>(work), sc->res, xc).value;

  const bool nucleus = top_p > 0.0f;
  if (nucleus) {
    __syncthreads();
    for (int i = threadIdx.x; i < 2 * kHistBins; i += kThreads) cnt[i] = 0u;
    __syncthreads();
  }

  // pass 2: e_i = exp(x_i - m) in place; Z1 (plain rows) or the exact-mass histogram (nucleus rows)
  float z = 0.0f;
  for (int i = threadIdx.x; i < n4; i += kThreads) {
    float4 v = x4[i];
    float e[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      const float x = e[c];
      const float ex = (x >= tau_k) ? expf(__fsub_rn(x, m)) : 0.0f;
      e[c] = ex;
      if (nucleus) {
        const unsigned d = kOneBits - __float_as_uint(ex);
        const unsigned b = d >> kBinShift;
        if (b < unsigned(kHistBins)) {
          atomicAdd(&cnt[b], 1u);
          atomicAdd(&low[b], d & 0xffffu);
        } else {
          z += ex;  // below 2^-32 of the maximum: summed in float, far under the rounding of Z1
        }
      } else {
        z += ex;
      }
    }
    x4[i] = make_float4(e[0], e[1], e[2], e[3]);
  }
  const float zsum = team_allreduce<C>(z, OpSumF(), xc);
  if (!nucleus) return RowStat{m, zsum, 0u};

  team_hist_allreduce<C>(reinterpret_cast<int*>(cnt), 2 * kHistBins);
  if (threadIdx.x == 0) {
    sc->bstar = kHistBins;
    sc->ncand = 0;
    sc->jstar = 0x7fffffff;
    sc->overflow = 0;
  }

  // exact bin masses -> boundary bin
  constexpr int kPer = kHistBins / kThreads;  // 8 bins per thread, ascending bin = descending e
  const int b0 = int(threadIdx.x) * kPer;
  double loc = 0.0;
  bool big = false;
  for (int i = 0; i < kPer; ++i) {
    const unsigned c = cnt[b0 + i];
    big |= (c > 65535u);  // the 32-bit low sum may have wrapped
    loc += bin_mass(b0 + i, c, low[b0 + i]);
  }
  double total = 0.0;
  const double run0 = block_exscan(loc, OpSumD(), 0.0, xc, &total);
  const double Z1 = total + double(zsum);
  const double target = double(top_p) * Z1;
  if (big) sc->overflow = 1;
  {
    int mine = kHistBins;
    double run = run0;
    for (int i = 0; i < kPer; ++i) {
      run += bin_mass(b0 + i, cnt[b0 + i], low[b0 + i]);
      if (run > target && mine == kHistBins) mine = b0 + i;
    }
    if (mine < kHistBins) atomicMin(&sc->bstar, mine);
  }
  __syncthreads();
  const int bstar = sc->bstar;
  if (bstar >= kHistBins || sc->overflow || cnt[bstar] > unsigned(kCandCap))
    return nucleus_bisect<C>(xs, s, n4, m, target, xc);
  {
    // the owner of b* publishes the exact mass above it
    if (bstar >= b0 && bstar < b0 + kPer) {
      double run = run0;
      for (int i = 0; b0 + i < bstar; ++i) run += bin_mass(b0 + i, cnt[b0 + i], low[b0 + i]);
      sc->m_above = run;
    }
  }
  __syncthreads();
  const double m_above = sc->m_above;

  // pass 3: collect the entries of the boundary bin
  BandWork* bw = reinterpret_cast<BandWork*>(work);
  __syncthreads();  // histogram reads done before the region is reused
  for (int i = threadIdx.x; i < n4; i += kThreads) {
    const float4 v = x4[i];
    const float e[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      if (int((kOneBits - __float_as_uint(e[c])) >> kBinShift) == bstar) {
        const int pos = atomicAdd(&sc->ncand, 1);
        if (pos < kCandCap) bw->cand_local[pos] = i * 4 + c;
      }
    }
  }
  __syncthreads();
  // exact ordering keys of this block's entries (re-read the scaled logit)
  const int nloc = sc->ncand;
  for (int j = threadIdx.x; j < nloc; j += kThreads) {
    const int li = bw->cand_local[j];
    const int g = s.g0 + li;
    const float x = scale_logit(to_f32(row[g]), temperature);
    bw->key[j] = ((unsigned long long)(~okey(x)) << 32) | (unsigned long long)unsigned(g);
    bw->mass[j] = xs[li];
  }
  Team<C>::sync();

  // team rank 0 orders the boundary bin and finds the boundary
  float Zf = 0.0f;
  if (Team<C>::rank() == 0) {
    int n = nloc;
    if constexpr (C > 1) {
      for (unsigned r = 1; r < unsigned(C); ++r) {
        RowScratch* psc = Team<C>::peer(sc, r);
        BandWork* pbw = Team<C>::peer(bw, r);
        const int nr = psc->ncand;
        for (int j = threadIdx.x; j < nr; j += kThreads) {
          bw->key[n + j] = pbw->key[j];
          bw->mass[n + j] = pbw->mass[j];
        }
        n += nr;
      }
    }
    int P = 32;
    while (P < n) P <<= 1;
    for (int j = n + threadIdx.x; j < P; j += kThreads) {
      bw->key[j] = ~0ull;
      bw->mass[j] = 0.0f;
    }
    __syncthreads();
    // bitonic sort (ascending key) of (key, mass)
    for (int k = 2; k <= P; k <<= 1) {
      for (int j = k >> 1; j > 0; j >>= 1) {
        for (int t = threadIdx.x; t < P; t += kThreads) {
          const int u = t ^ j;
          if (u > t) {
            const unsigned long long a = bw->key[t], b = bw->key[u];
            const bool up = ((t & k) == 0);
            if ((a > b) == up) {
              bw->key[t] = b;
              bw->key[u] = a;
              const float ma = bw->mass[t];
              bw->mass[t] = bw->mass[u];
              bw->mass[u] = ma;
            }
          }
        }
        __syncthreads();
      }
    }
    // inclusive prefix mass in sorted order (P <= 1024: two entries per thread)
    {
      const int t0 = 2 * int(threadIdx.x);
      const double a = t0 < P ? double(bw->mass[t0]) : 0.0;
      const double b = t0 + 1 < P ? double(bw->mass[t0 + 1]) : 0.0;
      const double ex = block_exscan(a + b, OpSumD(), 0.0, xc);
      if (t0 < P) bw->cum[t0] = m_above + ex + a;
      if (t0 + 1 < P) bw->cum[t0 + 1] = m_above + ex + a + b;
    }
    __syncthreads();
    for (int j = threadIdx.x; j < n; j += kThreads)
      if (bw->cum[j] > target) atomicMin(&sc->jstar, j);
    __syncthreads();
    const int jstar = sc->jstar;
    const int nkeep = jstar == 0x7fffffff ? n : jstar + 1;
    const double Z2 = nkeep > 0 ? bw->cum[nkeep - 1] : m_above;
    Zf = float(Z2);
    // zero the entries after the boundary in their owners' slices
    for (int j = nkeep + threadIdx.x; j < n; j += kThreads) {
      const int g = int(unsigned(bw->key[j] & 0xffffffffull));
      const unsigned owner = (C == 1) ? 0u : unsigned(g / per);
      float* pxs = Team<C>::peer(xs, owner);
      pxs[g - int(owner) * per] = 0.0f;
    }
  }
  Zf = team_bcast0<C>(Zf, xc);
  return RowStat{m, Zf, kOneBits - (unsigned(bstar + 1) << kBinShift)};
}

}  // namespace duodec
