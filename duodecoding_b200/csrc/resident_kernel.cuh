// Resident-row scan kernel: the streaming half of duodec_verify_fused, device path 4 (experimental).
//
//   row_resident_kernel  persistent grid, ONE block per SM; every block OWNS whole logit rows
//                    (row = block + k * grid): the hot loop has no global atomics, fences or
//                    cross-block hand-offs.  One producer lane keeps a ring of kStages 8 KB
//                    shared-memory stages full with TMA bulk copies (cp.async.bulk + mbarrier
//                    complete_tx), running ahead across row boundaries; 512 worker threads take
//                    one 128-bit shared load per stage each.
//
//   A row without nucleus filter is one pass:  Z = sum 2^(x*log2e - M0), stages freed as they go.
//   A row with top-p stays RESIDENT in the ring (a 32000-entry fp32 row is 16 of the 22 stages;
//   the other stages prefetch the next row) and is walked several times out of shared memory,
//   so neither HBM nor L2 sees it twice and nothing is predicted:
//     P1  as the stages land: e = 2^(x*log2e - M0) is added to a PER-THREAD histogram of 16
//         buckets of the exponent t = x*log2e - M0 (two units wide, clamped ends) in shared memory
//         (hist[bucket][thread]: bank = thread, no conflicts, no atomics).  The bucket sums give
//         Z and, scanned from the top, the bucket in which the kept mass crosses top_p * Z,
//         together with the mass above it.
//     P2  the entries of that bucket's value range [lo, hi) only: the same histogram, 16 equal
//         slices of [lo, hi) -> the slice with the crossing and the mass above it.
//     P3  the entries of that slice (typically ~100 of 32000) are appended to a shared-memory
//         list; more than 1024 -> P2/P3 again on the slice (ties aside, 16x fewer every time).
//     list_select (split_kernels.cuh) finds the boundary entry in the list in exact fixed-point
//         arithmetic and the reference's order (scaled logit descending, lowest index first).
//   Rows longer than the ring (V = 151936) walk P2/P3 with plain 128-bit loads out of L2.
//   Rows that do not fit the scheme (crossing in a clamped end bucket because M0 is far from the
//   row's maximum, huge tie groups) take the general search of split_kernels.cuh.
//
//   M0 is the mean row maximum (times log2 e) of the previous launch, kept in the workspace; any
//   value within a few units of the row's maximum is exact enough.  A cold workspace makes a
//   resident row find its maximum in an extra walk first.
//
//   The row's 32-byte record (M0, kept mass, boundary, probability of the drafted token) is
//   what seq_sample_kernel (split_kernels.cuh) evaluates the accept tests and draws tokens from.
#pragma once

#include "split_kernels.cuh"

namespace duodec {
namespace resident {

constexpr int kWT = 512;                          // worker threads
constexpr int kWW = kWT / 32;
constexpr int kRT = kWT + 32;                     // + the producer warp
constexpr int kStageElems = 2048;                 // row entries per ring stage: 4 per worker thread
constexpr int kStageBytes = kStageElems * 4;      // sized for fp32 rows; bf16 rows fill half of it
#ifndef DUODEC_RING_STAGES
#define DUODEC_RING_STAGES 22
#endif
constexpr int kStages = DUODEC_RING_STAGES;
constexpr int kNB = 16;                           // histogram buckets
constexpr float kHTmin = -28.0f, kHTmax = 2.0f;   // bucket b holds round(t / 2) == b - 14 (ends clamped)
constexpr float kHMagic = 25165824.0f;            // 1.5 * 2^24: t + magic rounds t to an even integer (ulp 2)
constexpr int kList2 = 1024;                      // entries the shared-memory list holds
constexpr int kMaxLevels = 8;
constexpr float kMaskT = -13.0f;                  // entries with t >= kMaskT are remembered in the per-thread masks
constexpr int kMaskBucket = 8;                    // = the buckets >= 8

struct __align__(16) RingSync {
  unsigned long long full[kStages];   // "stage landed" (TMA complete_tx)
  unsigned long long empty[kStages];  // "stage consumed" (one arrive per worker warp)
  float hsum[kNB];
  float4 red[2][kWW];
  int nlist;
  int pad[3];
};

constexpr size_t ring_smem_bytes() {
  return size_t(kStages) * kStageBytes + size_t(kNB) * kWT * 4 + size_t(kList2) * sizeof(ListEntry) +
         sizeof(Scratch3) + sizeof(RingSync) + 128;
}

__device__ __forceinline__ void barw() { asm volatile("bar.sync 2, 512;" ::: "memory"); }

// coarse bucket of a scaled logit (monotone in x)
__device__ __forceinline__ int coarse_bucket(float x, float m0) {
  const float t = fmaf(x, kLog2e, -m0);
  return __float_as_int(fminf(fmaxf(t, kHTmin), kHTmax) + kHMagic) - __float_as_int(kHTmin + kHMagic);
}

// smallest float x with coarse_bucket(x, m0) >= b (1 <= b <= 15); false if the search does not settle
__device__ __forceinline__ bool bucket_lower_edge(int b, float m0, float* edge) {
  float g = (float(2 * (b - 14) - 1) + m0) * 0.6931471805599453f;
  int it = 0;
  while (coarse_bucket(g, m0) >= b && it < 256) g = nextafterf(g, -INFINITY), ++it;
  while (coarse_bucket(g, m0) < b && it < 512) g = nextafterf(g, INFINITY), ++it;
  *edge = g;
  return it < 512 && coarse_bucket(g, m0) >= b && coarse_bucket(nextafterf(g, -INFINITY), m0) < b;
}

// all-reduce over the 512 workers: .x .y max-reduced, .z .w summed; one barrier, ping-pong scratch
__device__ __forceinline__ float4 reducew(float4 a, RingSync* rs, int& ph) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    a.x = fmaxf(a.x, __shfl_xor_sync(0xffffffffu, a.x, o));
    a.y = fmaxf(a.y, __shfl_xor_sync(0xffffffffu, a.y, o));
    a.z += __shfl_xor_sync(0xffffffffu, a.z, o);
    a.w += __shfl_xor_sync(0xffffffffu, a.w, o);
  }
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  if (lane == 0) rs->red[ph][w] = a;
  barw();
  float4 x = rs->red[ph][lane & (kWW - 1)];
  ph ^= 1;
#pragma unroll
  for (int o = kWW / 2; o > 0; o >>= 1) {
    x.x = fmaxf(x.x, __shfl_xor_sync(0xffffffffu, x.x, o));
    x.y = fmaxf(x.y, __shfl_xor_sync(0xffffffffu, x.y, o));
    x.z += __shfl_xor_sync(0xffffffffu, x.z, o);
    x.w += __shfl_xor_sync(0xffffffffu, x.w, o);
  }
  return x;
}

// Sums the per-thread histogram columns (and clears them), then every warp scans the 16 sums from
// the top: total mass of the histogram, the bucket in which base + (mass from the top) first
// exceeds `target` (-1: none) and the mass above that bucket (base included).  Fixed reduction
// order: every thread of every launch gets the same values.
__device__ __forceinline__ void hist_close(float* hist, RingSync* rs, float base, double target, float& total,
                                           int& bstar, float& above, unsigned& nonempty) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  barw();  // every column is written
  {
    float s = 0.0f;
#pragma unroll
    for (int k = 0; k < kWT / 32; ++k) {
      float* p = hist + warp * kWT + k * 32 + lane;
      s += *p;
      *p = 0.0f;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane == 0) rs->hsum[warp] = s;
  }
  barw();
  const float hb = lane < kNB ? rs->hsum[kNB - 1 - lane] : 0.0f;  // lane l: bucket 15 - l
  float incl = hb;
#pragma unroll
  for (int d = 1; d < kNB; d <<= 1) {
    const float o = __shfl_up_sync(0xffffffffu, incl, d);
    if (lane >= d) incl += o;
  }
  total = __shfl_sync(0xffffffffu, incl, kNB - 1);
  nonempty = __ballot_sync(0xffffffffu, hb > 0.0f) & 0xffffu;
  const unsigned crossed = __ballot_sync(0xffffffffu, double(base + incl) > target) & 0xffffu;
  bstar = -1;
  above = base;
  if (crossed) {
    const int l = __ffs(crossed) - 1;
    bstar = kNB - 1 - l;
    above = base + __shfl_sync(0xffffffffu, incl - hb, l);
  }
}

// ---- walks over the current row: resident stages (shared memory) or plain loads out of L2
#define RES_ROW_SCALE4(v)                                                                         \
  if (temp != 1.0f) {                                                                         \
    v.x = __fdiv_rn(v.x, temp), v.y = __fdiv_rn(v.y, temp);                                   \
    v.z = __fdiv_rn(v.z, temp), v.w = __fdiv_rn(v.w, temp);                                   \
  }
#define RES_ROW_WALK(F)                                                                           \
  if (resident) {                                                                             \
    _Pragma("unroll 2") for (int k = 0; k < nchunk; ++k) {                                    \
      const unsigned slot = (seq0 + unsigned(k)) % unsigned(kStages);                         \
      const int i0 = k * kStageElems + tid * 4;                                               \
      if (i0 < V) {                                                                           \
        float4 v = unpack4<TIn>(ring + size_t(slot) * kStageBytes + size_t(tid * 4) * kElt);  \
        RES_ROW_SCALE4(v)                                                                         \
        F(v.x, i0) F(v.y, i0 + 1) F(v.z, i0 + 2) F(v.w, i0 + 3)                               \
      }                                                                                       \
    }                                                                                         \
  } else {                                                                                    \
    _Pragma("unroll 1") for (int c0 = 0; c0 < V; c0 += 4 * 4 * kWT) {                         \
      float4 v[4];                                                                            \
      _Pragma("unroll") for (int u = 0; u < 4; ++u) {                                         \
        const int i0 = c0 + (u * kWT + tid) * 4;                                              \
        if (i0 < V) v[u] = load4_scaled<TIn>(row + i0, temp);                                 \
      }                                                                                       \
      _Pragma("unroll") for (int u = 0; u < 4; ++u) {                                         \
        const int i0 = c0 + (u * kWT + tid) * 4;                                              \
        if (i0 < V) { F(v[u].x, i0) F(v[u].y, i0 + 1) F(v[u].z, i0 + 2) F(v[u].w, i0 + 3) }   \
      }                                                                                       \
    }                                                                                         \
  }
// the remembered entries of this thread (resident rows): bit b of (mlo, mhi) = stage b / 4, entry b % 4
#define RES_MASK_WALK(F)                                                                          \
  _Pragma("unroll 1") for (int half = 0; half < 2; ++half) {                                  \
    unsigned mm_ = half ? mhi : mlo;                                                          \
    while (mm_) {                                                                             \
      const int b_ = __ffs(mm_) - 1;                                                          \
      mm_ &= mm_ - 1u;                                                                        \
      const int k_ = half * 8 + (b_ >> 2), j_ = b_ & 3;                                       \
      const unsigned slot = (seq0 + unsigned(k_)) % unsigned(kStages);                        \
      float x_ = to_f32(*reinterpret_cast<const TIn*>(ring + size_t(slot) * kStageBytes +     \
                                                      size_t(tid * 4 + j_) * kElt));          \
      if (temp != 1.0f) x_ = __fdiv_rn(x_, temp);                                             \
      F(x_, k_ * kStageElems + tid * 4 + j_)                                                  \
    }                                                                                         \
  }
// the stages of the row as they land; `release`: hand each stage back to the producer at once
#define RES_ROW_STREAM(F, release)                                                                \
  _Pragma("unroll 1") for (int k = 0; k < nchunk; ++k) {                                      \
    const unsigned sq = seq0 + unsigned(k);                                                   \
    const unsigned slot = sq % unsigned(kStages);                                             \
    mbar_wait(&rs->full[slot], (sq / unsigned(kStages)) & 1u);                                 \
    const int i0 = k * kStageElems + tid * 4;                                                 \
    if (i0 < V) {                                                                             \
      float4 v = unpack4<TIn>(ring + size_t(slot) * kStageBytes + size_t(tid * 4) * kElt);    \
      RES_ROW_SCALE4(v)                                                                           \
      F(v.x, i0) F(v.y, i0 + 1) F(v.z, i0 + 2) F(v.w, i0 + 3)                                 \
    }                                                                                         \
    if (release) {                                                                            \
      __syncwarp();                                                                           \
      if (lane == 0) mbar_arrive(&rs->empty[slot]);                                           \
    }                                                                                         \
  }

// General search of one row by the first 256 workers (rare): exact row maximum first, then the
// windowed search of split_kernels.cuh through L2.
template <class TIn>
__device__ __noinline__ RowRecord general_row(const TIn* row, int V, float temp, float top_p, bool nucleus, float m0,
                                              ListEntry* glist, int* count, Scratch3* sc, int* stats) {
  int phase = 0;
  Pass3 pr = rescan_row<TIn, kListCap>(row, V, temp, m0, INFINITY, INFINITY, glist, count, sc, phase);
  m0 = pr.mx * kLog2e;
  pr = rescan_row<TIn, kListCap>(row, V, temp, m0, INFINITY, INFINITY, glist, count, sc, phase);
  RowRecord rec;
  rec.m0 = m0;
  rec.Z = pr.z;
  rec.xstar = -INFINITY;
  rec.istar = 0x7fffffff;
  rec.sel = 0.0f;
  rec.mxl = m0;
  rec.dens = 0.0f;
  float theta = pr.mx, dens = 0.0f;
  if (nucleus)
    resolve_row3<TIn, kListCap>(row, V, temp, top_p, m0, pr, INFINITY, INFINITY, 0.0f, glist, count, sc, phase, rec,
                                &theta, &dens, stats);
  return rec;
}

template <class TIn>
__global__ void __launch_bounds__(kRT, 1) row_resident_kernel(const __grid_constant__ SplitParams P) {
  extern __shared__ __align__(128) unsigned char smem[];
  unsigned char* ring = smem;
  float* hist = reinterpret_cast<float*>(smem + size_t(kStages) * kStageBytes);
  ListEntry* slist = reinterpret_cast<ListEntry*>(hist + kNB * kWT);
  Scratch3* sc = reinterpret_cast<Scratch3*>(slist + kList2);
  RingSync* rs = reinterpret_cast<RingSync*>(sc + 1);

  const int tid = threadIdx.x;
  const int gamma = P.gamma, V = P.V;
  const int njobs = P.B * (2 * gamma + 1);
  const int nchunk = (V + kStageElems - 1) / kStageElems;
  constexpr int kElt = int(sizeof(TIn));

  if (tid == 0) {
    for (int s = 0; s < kStages; ++s) {
      mbar_init(&rs->full[s], 1);
      mbar_init(&rs->empty[s], kWW);
    }
    mbar_fence_init();
    rs->nlist = 0;
  }
  __syncthreads();

  if (tid >= kWT) {
    // ---- producer: one lane streams this block's rows through the ring, in job order
    if (tid == kWT) {
      unsigned seq = 0;
      for (int job = blockIdx.x; job < njobs; job += gridDim.x) {
        const TIn* row = split_row<TIn>(P, decode_job(job, gamma));
        for (int c = 0; c < nchunk; ++c, ++seq) {
          const unsigned slot = seq % unsigned(kStages);
          if (seq >= unsigned(kStages)) mbar_wait(&rs->empty[slot], ((seq / unsigned(kStages)) - 1u) & 1u);
          const int e0 = c * kStageElems;
          const int ne = (V - e0) < kStageElems ? (V - e0) : kStageElems;
          mbar_expect_tx(&rs->full[slot], unsigned(ne * kElt));
          bulk_g2s(ring + size_t(slot) * kStageBytes, row + e0, unsigned(ne * kElt), &rs->full[slot]);
        }
      }
    }
    return;
  }

  // ---- workers
  int ph = 0, phase = 0;
  unsigned seq0 = 0;  // ring sequence number of the current row's first stage
  const int lane = tid & 31;
  float* hp = hist + tid;  // this thread's histogram column
  const int bits0 = __float_as_int(kHTmin + kHMagic);
#pragma unroll
  for (int b = 0; b < kNB; ++b) hp[b * kWT] = 0.0f;

#pragma unroll 1
  for (int job = blockIdx.x; job < njobs; job += gridDim.x, seq0 += unsigned(nchunk)) {
    const RowJob jb = decode_job(job, gamma);
    const TIn* row = split_row<TIn>(P, jb);
    const float temp = jb.is_target ? P.t_temp : P.d_temp;
    const float top_p = jb.is_target ? P.t_top_p : P.d_top_p;
    const bool nucleus = top_p > 0.0f;
    const bool resident = nucleus && nchunk <= kStages - 2;
    ListEntry* glist = P.lists + (long long)blockIdx.x * kListCap;  // general search only

    // the drafted token's logit is needed when the row closes: fetch it now (thread 0)
    long long tok = -1;
    float xtok = 0.0f;
    if (tid == 0 && jb.j < gamma) {
      tok = P.tokens[(long long)jb.b * P.tok_sb + jb.j];
      if (tok >= 0 && tok < V) xtok = scale_logit(to_f32(__ldcg(row + tok)), temp);
    }

    // exponent offset predicted by the previous launch (mean row maximum of this model's rows)
    const uint4 h = __ldg(reinterpret_cast<const uint4*>(&P.hints[jb.is_target]));
    const bool warm = int(h.w) != 0;
    float m0 = warm ? __uint_as_float(h.x) : 0.0f;

#ifdef DUODEC_TICKS
    long long tk0 = 0;
    if (P.stats && tid == 0) tk0 = clock64();
#define res_tick(k)                                            \
  if (P.stats && tid == 0) {                               \
    const long long t_ = clock64();                        \
    atomicAdd(&P.stats[k], int((t_ - tk0) >> 4));          \
    tk0 = t_;                                              \
  }
#else
#define res_tick(k)
#endif

    RowRecord rec;
    rec.xstar = -INFINITY;
    rec.istar = 0x7fffffff;
    rec.sel = 0.0f;
    rec.dens = 0.0f;
    bool slow = false, selected = false;
    float Z = 0.0f, mxl = m0;

    if (!nucleus) {
      float z = 0.0f, mx = -INFINITY;
#define RES_F_PLAIN(x, i) z += row_exp(x, m0), mx = fmaxf(mx, x);
      RES_ROW_STREAM(RES_F_PLAIN, true)
      const float4 r = reducew(make_float4(mx, -INFINITY, z, 0.0f), rs, ph);
      Z = r.z;
      mxl = r.x * kLog2e;
      slow = !(fabsf(mxl - m0) <= 32.0f) || !(Z < INFINITY) || !(Z > 0.0f);
    } else {
      if (!warm && resident) {
        // cold workspace: the row's own maximum first (the stages stay where they are)
        float mx = -INFINITY;
#define RES_F_MAX(x, i) mx = fmaxf(mx, x);
        RES_ROW_STREAM(RES_F_MAX, false)
        const float4 r = reducew(make_float4(mx, -INFINITY, 0.0f, 0.0f), rs, ph);
        m0 = r.x * kLog2e;
        if (!(fabsf(m0) < INFINITY)) m0 = 0.0f;
      }
      // ---- P1: per-thread mass histogram over the exponent, as the stages land
#define RES_F_HIST(x, i)                                                     \
  {                                                                      \
    const float t = fmaf(x, kLog2e, -m0);                                \
    const int bk = __float_as_int(fminf(fmaxf(t, kHTmin), kHTmax) + kHMagic) - bits0; \
    hp[bk * kWT] += ex2_approx(t);                                       \
  }
      // a resident row of <= 16 stages also remembers, one bit per entry of this thread, which
      // entries lie in the upper buckets (t >= -13: buckets 8..15): P2 / P3 visit only those
      const bool masked = resident && nchunk <= 16;
      unsigned mlo = 0u, mhi = 0u;
      if (masked) {
        unsigned m = 0u, bit = 1u;
#define RES_F_HISTM(x, j)                                                    \
  {                                                                      \
    const float t = fmaf(x, kLog2e, -m0);                                \
    const int bk = __float_as_int(fminf(fmaxf(t, kHTmin), kHTmax) + kHMagic) - bits0; \
    hp[bk * kWT] += ex2_approx(t);                                       \
    if (t >= kMaskT) m |= bit << (j);                                    \
  }
#pragma unroll 1
        for (int k = 0; k < nchunk; ++k) {
          const unsigned sq = seq0 + unsigned(k);
          const unsigned slot = sq % unsigned(kStages);
          mbar_wait(&rs->full[slot], (sq / unsigned(kStages)) & 1u);
          if (k * kStageElems + tid * 4 < V) {
            float4 v = unpack4<TIn>(ring + size_t(slot) * kStageBytes + size_t(tid * 4) * kElt);
            RES_ROW_SCALE4(v)
            RES_F_HISTM(v.x, 0) RES_F_HISTM(v.y, 1) RES_F_HISTM(v.z, 2) RES_F_HISTM(v.w, 3)
          }
          bit <<= 4;
          if (k == 7) mlo = m, m = 0u, bit = 1u;
        }
        if (nchunk <= 8) mlo = m; else mhi = m;
      } else if (resident) {
        RES_ROW_STREAM(RES_F_HIST, false)
      } else {
        RES_ROW_STREAM(RES_F_HIST, true)
      }
      res_tick(24);
      int b1;
      float GB;
      unsigned nonempty;
      hist_close(hist, rs, 0.0f, INFINITY, Z, b1, GB, nonempty);  // (target unknown before Z: scan again below)
      const double target = double(top_p) * double(Z);
      {
        // the crossing bucket, from the published sums (same fixed order as hist_close)
        const float hb = lane < kNB ? rs->hsum[kNB - 1 - lane] : 0.0f;
        float incl = hb;
#pragma unroll
        for (int d = 1; d < kNB; d <<= 1) {
          const float o = __shfl_up_sync(0xffffffffu, incl, d);
          if (lane >= d) incl += o;
        }
        const unsigned crossed = __ballot_sync(0xffffffffu, double(incl) > target) & 0xffffu;
        b1 = -1, GB = 0.0f;
        if (crossed) {
          const int l = __ffs(crossed) - 1;
          b1 = kNB - 1 - l;
          GB = __shfl_sync(0xffffffffu, incl - hb, l);
        }
      }
      const int btop = nonempty ? kNB - 1 - (__ffs(nonempty) - 1) : -1;
      mxl = m0 + (btop == kNB - 1 ? fmaxf(kHTmax, log2f(rs->hsum[kNB - 1])) : float(2 * (btop - 14)));
      slow = !(Z < INFINITY) || !(Z > 0.0f);
      float lo = 0.0f, hi = 0.0f;
      if (!slow && target < double(Z)) {
        if (b1 <= 0 || b1 >= kNB - 1) {
          slow = true;  // the end buckets are unbounded
        } else if (!bucket_lower_edge(b1, m0, &lo) || !bucket_lower_edge(b1 + 1, m0, &hi)) {
          slow = true;
        }
      }
      res_tick(25);
      if (!slow && target < double(Z)) {
        // ---- P2 / P3 levels on [lo, hi): GB = mass of the entries >= hi
        const bool use_mask = masked && b1 >= kMaskBucket;
#pragma unroll 1
        for (int level = 0; level < kMaxLevels && !selected && !slow; ++level) {
          const float fsc = float(kNB) / (hi - lo);
#define RES_F_FINE(x, i)                                           \
  if (x >= lo && x < hi) {                                     \
    int sb = int((x - lo) * fsc);                              \
    sb = sb < 0 ? 0 : (sb > kNB - 1 ? kNB - 1 : sb);           \
    hp[sb * kWT] += row_exp(x, m0);                            \
  }
          if (use_mask) { RES_MASK_WALK(RES_F_FINE) } else { RES_ROW_WALK(RES_F_FINE) }
          res_tick(28);
          float tot2, GB2;
          int b2;
          unsigned ne2;
          hist_close(hist, rs, GB, target, tot2, b2, GB2, ne2);
          res_tick(29);
          if (ne2 == 0u) {
            slow = true;  // empty bracket (cannot happen with consistent sums)
            break;
          }
          if (b2 < 0) {
            // rounding only: the whole bracket is kept; its lowest slice holds the boundary entry
            b2 = kNB - 1 - (31 - __clz(ne2));
            GB2 = GB + tot2;  // (list_select keeps everything when the total does not cross)
            // mass above the lowest slice, recomputed from the published sums
            float s = GB;
            for (int b = kNB - 1; b > b2; --b) s += rs->hsum[b];
            GB2 = s;
          }
          float emx = -INFINITY, engmn = -INFINITY;
#define RES_F_TAKE(x, i)                                           \
  if (x >= lo && x < hi) {                                     \
    int sb = int((x - lo) * fsc);                              \
    sb = sb < 0 ? 0 : (sb > kNB - 1 ? kNB - 1 : sb);           \
    if (sb == b2) {                                            \
      const int pos = atomicAdd(&rs->nlist, 1);                \
      if (pos < kList2) slist[pos] = ListEntry{x, i};          \
      emx = fmaxf(emx, x), engmn = fmaxf(engmn, -x);           \
    }                                                          \
  }
          if (use_mask) { RES_MASK_WALK(RES_F_TAKE) } else { RES_ROW_WALK(RES_F_TAKE) }
          res_tick(30);
          const float4 mm = reducew(make_float4(emx, engmn, 0.0f, 0.0f), rs, ph);  // (barrier: list complete)
          res_tick(31);
          const int n = *reinterpret_cast<volatile int*>(&rs->nlist);
          const float nlo = -mm.y, nhi = nextafterf(mm.x, INFINITY);
          GB = GB2;
          if (n <= 0) {
            slow = true;
          } else if (n <= kList2) {
            selected = true;
            lo = nlo, hi = nhi;
          } else if (nlo == mm.x || (nlo == lo && nhi == hi)) {
            slow = true;  // one value, or no progress: a tie group larger than the list
          } else {
            lo = nlo, hi = nhi;
            barw();
            if (tid == 0) rs->nlist = 0;
            barw();
          }
        }
        if (!selected) slow = true;
      }
      res_tick(26);
      if (resident) {
        // the row's stages go back to the producer
        barw();
        if (lane == 0)
          for (int k = 0; k < nchunk; ++k) mbar_arrive(&rs->empty[(seq0 + unsigned(k)) % unsigned(kStages)]);
      }
      if (selected) {
        const int n = *reinterpret_cast<volatile int*>(&rs->nlist);
        if (P.stats && tid == 0) {
          atomicAdd(&P.stats[0], 1);
          atomicAdd(&P.stats[16], 1);
          atomicAdd(&P.stats[21], n);
          atomicMax(&P.stats[22], n);
        }
        if (tid < kT3) list_select_body<kList2>(slist, n, -1, lo, hi, GB, Z, target, m0, sc, phase, rec, P.stats);
      } else {
        rec.Z = Z;  // top_p >= 1 keeps everything (or the general search below decides)
      }
    }
    if (!nucleus) rec.Z = Z;
    rec.m0 = m0;
    rec.mxl = mxl;
    res_tick(27);
    if (slow) {
      // general search (first 256 workers): exact row maximum first, then the windowed search through L2
      barw();
      if (tid < kT3) rec = general_row<TIn>(row, V, temp, top_p, nucleus, m0, glist, &rs->nlist, sc, P.stats);
    }
    if (tid == 0) {
      if (tok >= 0 && tok < V) {
        const float e = kept(rec, xtok, int(tok)) ? row_exp(xtok, rec.m0) : 0.0f;
        rec.sel = __fdiv_rn(e, rec.Z);
      }
      store_record(&P.records[job], rec);
      rs->nlist = 0;
    }
    barw();  // list, counter and scratch are free for the next row
  }
}

}  // namespace resident
}  // namespace duodec
