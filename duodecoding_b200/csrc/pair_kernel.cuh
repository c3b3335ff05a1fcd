// Row-pair scan kernel: the streaming half of duodec_verify_fused, device path 5.
//
// Same algorithm as row_ring_kernel (ring_kernel.cuh: per-thread exponent histogram in pass 1, the
// boundary bucket's entries in pass 2 out of L2, list_select), but a block works on TWO rows at a
// time: its 256 workers are two halves of 128, each half with its own TMA stage ring (fed by its
// own producer warp), histogram and half-barrier, streaming its own row.  With two blocks per SM
// that is four rows in flight per SM - at the C2 shape every row of the step is in flight at once,
// so the latency-bound row tails (bucket scan, second pass, select) overlap instead of queueing
// behind each other.  Each half selects its own row (half_select); the halves join only for rows that need
// the general search.  Pass 2 appends the boundary bucket's entries to per-THREAD list columns with predicated
// stores (round 1: hit masks, a warp prefix sum and one divergent trip per hit into per-warp regions).
#pragma once

#include "ring_kernel.cuh"

namespace duodec {
namespace pair {

constexpr int kHT = 128;                         // worker threads per half
constexpr int kHW = kHT / 32;
constexpr int kPT = kT3 + 64;                    // 256 workers + one producer warp per half
#ifndef DUODEC_PAIR_STAGES
#define DUODEC_PAIR_STAGES 2
#endif
constexpr int kPStages = DUODEC_PAIR_STAGES;     // 8 KB stages per half
constexpr int kPairCap = 2048;                   // kCap of the block's list_select over one row's (compacted) list
constexpr int kColSlots = kHB;                   // list slots of a worker thread: its own column of boundary-bucket entries
                                                 // = its histogram column, dead after pass 1 (the last 4 slots are slack:
                                                 // the column is re-checked once per float4)

struct __align__(16) HalfRow {
  double target;
  float GB, Z, m0, mxl;
  int state;  // 0 no row, 1 plain row (no filter), 2 select from the list, 3 general search, 4 keep everything
  int pad[3];
};

struct __align__(16) PairSync {
  unsigned long long full[2][kPStages];
  unsigned long long empty[2][kPStages];
  float hsum[2][kHB];
  float2 hred[2][kHW];
  HalfRow half[2];
  int nlist;
  int pad[3];
  int colfull[2];  // a thread of the half ran out of list slots: general search
};

// Scratch of one half's own select (half_select below): the level histogram in native 32-bit
// shared-memory atomics - count, and the 38-bit fixed-point mass split into two 19-bit parts.
struct __align__(16) HalfScratch {
  unsigned hcnt[kHT], hlo[kHT], hhi[kHT];
  unsigned long long cand_f[kHT];
  float cand_x[kHT];
  int cand_i[kHT];
  unsigned long long wscan[kHW];
  float2 red[kHW];
  unsigned long long gb_new, lastkey;
  unsigned kmin, kmax;
  int bstar, bcount, ncand, res_ok;
  float res_x;
  int res_i;
  int pad[2];
};

// bit map of the captured positions: 16 bits per thread and stage (the entries' indices are recovered from it)
__host__ __device__ constexpr size_t pair_bitmap_bytes(int nchunk) { return (size_t(nchunk) * kHT * 2 + 15) & ~size_t(15); }

// rings | histograms [2][32][128] float (pass 2: the list columns) | scratch | sync | select scratch | bit maps [2]
__host__ __device__ constexpr size_t pair_smem_bytes(int nchunk) {
  return size_t(2) * kPStages * kStageBytes + size_t(kHB) * kT3 * 4 + sizeof(Scratch3) + sizeof(PairSync) +
         2 * sizeof(HalfScratch) + 2 * pair_bitmap_bytes(nchunk) + 128;
}

__device__ __forceinline__ void barh(int h) {  // named barrier of one half (immediate ids: 2 and 3)
  if (h == 0) asm volatile("bar.sync 2, 128;" ::: "memory");
  else asm volatile("bar.sync 3, 128;" ::: "memory");
}

// one stage = 16 entries per half-thread: four 128-bit shared loads at (u * 128 + ht) * 4
#define PAIR_STAGE_BEGIN(c)                                                                   \
  const unsigned slot = seq % unsigned(kPStages);                                             \
  mbar_wait(&ps->full[h][slot], (seq / unsigned(kPStages)) & 1u);                              \
  const unsigned char* st = ring + size_t(slot) * kStageBytes;                                \
  float4 v4[4];                                                                               \
  _Pragma("unroll") for (int u = 0; u < 4; ++u) v4[u] = unpack4<TIn>(st + size_t((u * kHT + ht) * 4) * kElt); \
  if (temp != 1.0f) {                                                                         \
    _Pragma("unroll") for (int u = 0; u < 4; ++u) {                                           \
      v4[u].x = __fdiv_rn(v4[u].x, temp), v4[u].y = __fdiv_rn(v4[u].y, temp);                 \
      v4[u].z = __fdiv_rn(v4[u].z, temp), v4[u].w = __fdiv_rn(v4[u].w, temp);                 \
    }                                                                                         \
  }                                                                                           \
  const int i00 = (c) * kStageElems + ht * 4;
#define PAIR_STAGE_END()                                                                      \
  __syncwarp();                                                                               \
  if (lane == 0) mbar_arrive(&ps->empty[h][slot]);                                            \
  ++seq;
#ifndef PAIR_GROUP_BRANCHES  // (default since the A/B of round 2: -0.5 us per C2 step)
// one branch per stage: only the row's last stage can be ragged
#define PAIR_EACH(F)                                                                          \
  if (i00 + 3 * kHT * 4 < V) {                                                                \
    _Pragma("unroll") for (int u = 0; u < 4; ++u) {                                           \
      const int iu = i00 + u * kHT * 4;                                                       \
      F(v4[u].x, iu) F(v4[u].y, iu + 1) F(v4[u].z, iu + 2) F(v4[u].w, iu + 3)                 \
    }                                                                                         \
  } else {                                                                                    \
    _Pragma("unroll") for (int u = 0; u < 4; ++u) {                                           \
      const int iu = i00 + u * kHT * 4;                                                       \
      if (iu < V) { F(v4[u].x, iu) F(v4[u].y, iu + 1) F(v4[u].z, iu + 2) F(v4[u].w, iu + 3) } \
    }                                                                                         \
  }
#define PAIR_EACH4(G)                                                                         \
  if (i00 + 3 * kHT * 4 < V) {                                                                \
    _Pragma("unroll") for (int u = 0; u < 4; ++u) { G(v4[u]) }                                \
  } else {                                                                                    \
    _Pragma("unroll") for (int u = 0; u < 4; ++u) {                                           \
      if (i00 + u * kHT * 4 < V) { G(v4[u]) }                                                 \
    }                                                                                         \
  }
#else
#define PAIR_EACH(F)                                                                          \
  _Pragma("unroll") for (int u = 0; u < 4; ++u) {                                             \
    const int iu = i00 + u * kHT * 4;                                                         \
    if (iu < V) { F(v4[u].x, iu) F(v4[u].y, iu + 1) F(v4[u].z, iu + 2) F(v4[u].w, iu + 3) }   \
  }
#define PAIR_EACH4(G)                                                                         \
  _Pragma("unroll") for (int u = 0; u < 4; ++u) {                                             \
    if (i00 + u * kHT * 4 < V) { G(v4[u]) }                                                   \
  }
#endif

// xa <= x < xb: append x to the thread's own list column (predicated store, no atomics, no shuffles), advance the
// column pointer, mark the entry's bit in the stage's mask
__device__ __forceinline__ void col_take(unsigned& lp, unsigned& hm, float x, float xa, float xb, unsigned bit) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ge.f32 p, %2, %3;\n\tsetp.lt.and.f32 p, %2, %4, p;\n\t@p st.shared.f32 [%0], %2;\n\t"
      "@p add.u32 %0, %0, 512;\n\t@p or.b32 %1, %1, %5;\n\t}"
      : "+r"(lp), "+r"(hm)
      : "f"(x), "f"(xa), "f"(xb), "r"(bit)
      : "memory");
}

// row index of the thread's j-th list entry (capture order) from its bit-map column
__device__ __forceinline__ int col_index_of(const unsigned short* bmt, int j, int ht) {
  int stage = 0;
  unsigned hm = bmt[0];
  while (__popc(hm) <= j) {
    j -= __popc(hm);
    ++stage;
    hm = bmt[stage * kHT];
  }
  for (; j > 0; --j) hm &= hm - 1u;  // drop the j lowest set bits (j < 16)
  const int k = __ffs(hm) - 1;
  return stage * kStageElems + ((k >> 2) * kHT + ht) * 4 + (k & 3);
}

// smallest float x whose histogram bucket (hist_bits of x*log2e - m0) is >= bits; false if the search
// does not settle (the caller then takes the general search)
__device__ __forceinline__ bool bucket_edge(int bits, float m0, float* edge) {
  float g = (float(bits - __float_as_int(kHMagic)) - 0.5f + m0) * 0.6931471805599453f;
  int it = 0;
  while (hist_bits(fmaf(g, kLog2e, -m0)) >= bits && it < 128) g = nextafterf(g, -INFINITY), ++it;
  while (hist_bits(fmaf(g, kLog2e, -m0)) < bits && it < 256) g = nextafterf(g, INFINITY), ++it;
  *edge = g;
  return it < 256 && hist_bits(fmaf(g, kLog2e, -m0)) >= bits &&
         hist_bits(fmaf(nextafterf(g, -INFINITY), kLog2e, -m0)) < bits;
}

// Nucleus boundary of one row by ITS OWN half (128 threads, half barrier), so that the two rows of a
// pair are finished side by side.  Same exact arithmetic and the same answer as list_select_body
// (split_kernels.cuh): 38-bit fixed-point masses, a level of 128 value buckets over the entries'
// value range, suffix scan, all-pairs order of the <= 128 entries of the crossing bucket.  The
// thread's entries are col[j * 128], j < nt (its own list column, capture order); their row indices
// come from its bit-map column bmt.  Returns false when the row needs the general search (more equal
// values than the finish can order).
__device__ __forceinline__ bool half_select(const float* col, int nt, const unsigned short* bmt, int h, int ht, float xa,
                                            float xb, float GB, float Zrow, double target, float m0, HalfScratch* hs,
                                            RowRecord& rec, int* ncand_out, int* stats) {
  constexpr unsigned long long kNone = 0xffffffffffffffffull;
  const int lane = ht & 31, hw = ht >> 5;
  const int ez = int(__float_as_uint(Zrow) >> 23) - 126;               // Z = f * 2^ez, f in [0.5, 1)
  const float S = __uint_as_float(unsigned(127 + 38 - ez) << 23);      // 2^(38 - ez)
  const unsigned long long tgt = (unsigned long long)(target * double(S));  // floor
  unsigned long long GBf = (unsigned long long)(double(GB) * double(S));
  if (GBf > tgt) GBf = tgt;  // GB is an fp32 scan difference: it may land a few ulps above the target; the top
                             // non-empty bucket is then the crossing one and its largest entry the boundary
  auto fx = [&](float x) { return __float2ull_rn(row_exp(x, m0) * S); };  // fixed-point mass of an entry
  // the first bracket is the boundary bucket's own interval [xa, xb) (every list entry lies inside it); only
  // the unbounded top bucket needs the entries' maximum
  float A = xa, B = xb;
  if (!(xb < INFINITY)) {
    float emx = -INFINITY;
    for (int j = 0; j < nt; ++j) emx = fmaxf(emx, col[j * kHT]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) emx = fmaxf(emx, __shfl_xor_sync(0xffffffffu, emx, o));
    if (lane == 0) hs->red[hw] = make_float2(emx, 0.0f);
    barh(h);
#pragma unroll
    for (int w = 0; w < kHW; ++w) emx = fmaxf(emx, hs->red[w].x);
    B = nextafterf(emx, INFINITY);
  }
  bool solved = false, tie_mode = false;
  *ncand_out = 0;
#pragma unroll 1
  for (int level = 0; level < 8 && !solved && !tie_mode; ++level) {
    hs->hcnt[ht] = 0u, hs->hlo[ht] = 0u, hs->hhi[ht] = 0u;
    if (ht == 0) {
      hs->bstar = -1;
      hs->kmin = 0xffffffffu;
      hs->kmax = 0u;
      hs->ncand = 0;
      hs->res_ok = 0;
      hs->lastkey = kNone;
    }
    barh(h);
    const float bsc = float(kHT) / (B - A);
    auto bucket = [&](float x) {
      const int bk = int((x - A) * bsc);
      return bk < 0 ? 0 : (bk > kHT - 1 ? kHT - 1 : bk);
    };
    for (int j = 0; j < nt; ++j) {
      const float x = col[j * kHT];
      if (x >= A && x < B) {
        const int bk = bucket(x);
        const unsigned long long f = fx(x);
        atomicAdd(&hs->hcnt[bk], 1u);
        atomicAdd(&hs->hlo[bk], unsigned(f) & 0x7ffffu);
        atomicAdd(&hs->hhi[bk], unsigned(f >> 19));
      }
    }
    barh(h);
    if (ht == 0) STAMP(stats, 25);
    // suffix scan over buckets 127..0 (thread t owns bucket 127 - t)
    const unsigned hc = hs->hcnt[kHT - 1 - ht];
    const unsigned long long hv = ((unsigned long long)hs->hhi[kHT - 1 - ht] << 19) + hs->hlo[kHT - 1 - ht];
    unsigned long long inc = hv;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const unsigned long long o = __shfl_up_sync(0xffffffffu, inc, d);
      if (lane >= d) inc += o;
    }
    if (lane == 31) hs->wscan[hw] = inc;
    barh(h);
#pragma unroll
    for (int w = 0; w < kHW; ++w)
      if (w < hw) inc += hs->wscan[w];
    const unsigned long long m_incl = GBf + inc, m_excl = m_incl - hv;
    if (hc != 0u && m_incl > tgt && m_excl <= tgt) {
      hs->bstar = kHT - 1 - ht;
      hs->gb_new = m_excl;
      hs->bcount = int(hc);
    }
    if (ht == kHT - 1 && m_incl <= tgt) hs->bstar = -2;  // rounding only: the whole bracket is kept
    barh(h);
    if (ht == 0) STAMP(stats, 26);
    const int bstar = hs->bstar;
    if (bstar < 0) {
      // keep everything down to the last entry of the bracket in the reference's order
      unsigned long long kk = kNone;
      for (int j = 0; j < nt; ++j) {
        const float x = col[j * kHT];
        if (x >= A && x < B) {
          const unsigned idx = unsigned(col_index_of(bmt, j, ht));
          const unsigned long long k2 = ((unsigned long long)okey(x) << 32) | (0xffffffffu - idx);
          kk = k2 < kk ? k2 : kk;
        }
      }
      if (kk != kNone) atomicMin(&hs->lastkey, kk);
      barh(h);
      const unsigned long long lk = hs->lastkey;
      if (ht == kHT - 1) hs->gb_new = m_incl;  // the bracket's total is this thread's inclusive sum
      barh(h);
      if (lk != kNone) {
        rec.Z = float(double(hs->gb_new) / double(S));
        rec.xstar = okey_inv(unsigned(lk >> 32));
        rec.istar = int(0xffffffffu - unsigned(lk & 0xffffffffull));
        solved = true;
      } else {
        tie_mode = true;
      }
      break;
    }
    const int bcount = hs->bcount;
    const unsigned long long gb_new = hs->gb_new;
    if (bcount <= kHT) {
      // ---- the bucket's entries, ordered exactly like the reference's sort
      for (int j = 0; j < nt; ++j) {
        const float x = col[j * kHT];
        if (x >= A && x < B && bucket(x) == bstar) {
          const int pos = atomicAdd(&hs->ncand, 1);
          hs->cand_x[pos] = x;
          hs->cand_f[pos] = fx(x);
          hs->cand_i[pos] = (j << 8) | ht;  // (owner thread, its entry number): the row index is recovered below
        }
      }
      barh(h);
      const int n = hs->ncand;
      // one thread per candidate walks the owner's bit-map column (in parallel: not once per trip of the loop above)
      if (ht < n) {
        const int code = hs->cand_i[ht], owner = code & 0xff;
        hs->cand_i[ht] = col_index_of(bmt - ht + owner, code >> 8, owner);
      }
      barh(h);
      if (ht == 0) STAMP(stats, 27);
      *ncand_out = n;
      int Pn = 16;
      while (Pn < n) Pn <<= 1;
      const int parts = kHT / Pn;  // 1..8 threads share one candidate (same warp)
      const int t = ht / parts, part = ht - t * parts;
      unsigned long long cum = 0ull, fme = 0ull;
      float xx = 0.0f;
      int ii = 0;
      if (t < n) {
        fme = hs->cand_f[t], xx = hs->cand_x[t], ii = hs->cand_i[t];
        for (int j = part; j < n; j += parts) {
          const float xj = hs->cand_x[j];
          const int ij = hs->cand_i[j];
          if (xj > xx || (xj == xx && ij <= ii)) cum += hs->cand_f[j];
        }
      }
      for (int o = parts >> 1; o > 0; o >>= 1) cum += __shfl_xor_sync(0xffffffffu, cum, o);
      const unsigned long long incl = gb_new + cum;
      if (t < n && part == 0 && incl > tgt && incl - fme <= tgt) {
        hs->res_x = xx;
        hs->res_i = ii;
        hs->gb_new = incl;
        hs->res_ok = 1;
      }
      barh(h);
      if (ht == 0) STAMP(stats, 28);
      if (hs->res_ok) {
        solved = true;
        rec.Z = float(double(hs->gb_new) / double(S));
        rec.xstar = hs->res_x;
        rec.istar = hs->res_i;
      } else {
        tie_mode = true;  // cannot happen with exact sums; stay safe
      }
      break;
    }
    // crowded bucket: shrink the bracket to the bucket's value range and go one level down
    for (int j = 0; j < nt; ++j) {
      const float x = col[j * kHT];
      if (x >= A && x < B && bucket(x) == bstar) {
        atomicMin(&hs->kmin, okey(x));
        atomicMax(&hs->kmax, okey(x));
      }
    }
    barh(h);
    const unsigned kmin = hs->kmin, kmax = hs->kmax;
    A = okey_inv(kmin);
    B = nextafterf(okey_inv(kmax), INFINITY);
    GBf = gb_new;
    if (kmin == kmax) tie_mode = true;  // more equal values than the finish can order
    barh(h);
  }
  return solved;
}

// General search of one row by the 256 workers (rare): exact row maximum first, then the windowed
// search of split_kernels.cuh through L2.
template <class TIn>
__device__ __noinline__ RowRecord general_row(const TIn* row, int V, float temp, float top_p, bool nucleus, float m0,
                                              ListEntry* glist, int* count, Scratch3* sc, int* stats) {
  int phase = 0;
  Pass3 pr = rescan_row<TIn, kListCap>(row, V, temp, m0, INFINITY, INFINITY, glist, count, sc, phase);
  m0 = pr.mx * kLog2e;
  pr = rescan_row<TIn, kListCap>(row, V, temp, m0, INFINITY, INFINITY, glist, count, sc, phase);
  RowRecord rec;
  rec.m0 = m0, rec.Z = pr.z, rec.xstar = -INFINITY, rec.istar = 0x7fffffff;
  rec.sel = 0.0f, rec.mxl = m0, rec.dens = 0.0f;
  float theta = pr.mx, dens = 0.0f;
  if (nucleus)
    resolve_row3<TIn, kListCap>(row, V, temp, top_p, m0, pr, INFINITY, INFINITY, 0.0f, glist, count, sc, phase, rec,
                                &theta, &dens, stats);
  return rec;
}

template <class TIn>
__global__ void __launch_bounds__(kPT, 2) row_pair_kernel(const __grid_constant__ SplitParams P) {
  extern __shared__ __align__(128) unsigned char smem[];
  float* hist_all = reinterpret_cast<float*>(smem + size_t(2) * kPStages * kStageBytes);
  Scratch3* sc = reinterpret_cast<Scratch3*>(hist_all + kHB * kT3);
  PairSync* ps = reinterpret_cast<PairSync*>(sc + 1);
  HalfScratch* hsc = reinterpret_cast<HalfScratch*>(ps + 1);
  unsigned char* bm_all = reinterpret_cast<unsigned char*>(hsc + 2);  // [2 rows][nchunk][128] u16 (only with a nucleus)

  const int tid = threadIdx.x;
  const int gamma = P.gamma, V = P.V;
  const int njobs = P.B * (2 * gamma + 1);
  const int npairs = (njobs + 1) / 2;
  const int nchunk = (V + kStageElems - 1) / kStageElems;
  constexpr int kElt = int(sizeof(TIn));

  if (tid == 0) {
    for (int hh = 0; hh < 2; ++hh)
      for (int s = 0; s < kPStages; ++s) {
        mbar_init(&ps->full[hh][s], 1);
        mbar_init(&ps->empty[hh][s], kHW);
      }
    mbar_fence_init();
    ps->nlist = 0;
    ps->colfull[0] = ps->colfull[1] = 0;
  }
  __syncthreads();
  // nothing above touches global memory; inputs, hints and the workspace are ordered from here on
  grid_dep_wait();
  grid_dep_launch();
  if (tid == 0) STAMP(P.stats, 8, true);

  if (tid >= kT3) {
    // ---- producers: lane 0 of warp 8 feeds half 0, lane 0 of warp 9 feeds half 1
    const int h = (tid - kT3) >> 5;
    if (((tid - kT3) & 31) == 0) {
      unsigned char* ring = smem + size_t(h) * kPStages * kStageBytes;
      unsigned seq = 0;
      for (int pr = blockIdx.x; pr < npairs; pr += gridDim.x) {
        const int job = 2 * pr + h;
        if (job >= njobs) continue;
        const RowJob jb = decode_job(job, gamma);
        const TIn* row = split_row<TIn>(P, jb);
#ifdef ABL_NOPASS2
        const int passes = 1;
#else
        const int passes = (jb.is_target ? P.t_top_p : P.d_top_p) > 0.0f ? 2 : 1;
#endif
        for (int pass = 0; pass < passes; ++pass)
          for (int c = 0; c < nchunk; ++c, ++seq) {
            const unsigned slot = seq % unsigned(kPStages);
            if (seq >= unsigned(kPStages))
              mbar_wait(&ps->empty[h][slot], ((seq / unsigned(kPStages)) - 1u) & 1u);
            const int e0 = c * kStageElems;
            const int ne = (V - e0) < kStageElems ? (V - e0) : kStageElems;
            mbar_expect_tx(&ps->full[h][slot], unsigned(ne * kElt));
            bulk_g2s(ring + size_t(slot) * kStageBytes, row + e0, unsigned(ne * kElt), &ps->full[h][slot]);
          }
      }
    }
    return;
  }

  // ---- workers: half h = threads [128 h, 128 h + 128)
  const int h = tid >> 7, ht = tid & (kHT - 1), lane = tid & 31, hw = ht >> 5;
  unsigned char* ring = smem + size_t(h) * kPStages * kStageBytes;
  float* hist = hist_all + h * kHB * kHT;  // this half's [bucket][thread] histogram
  float* hp = hist + ht;
  float* col = hp;  // pass 2: the thread's histogram column becomes its list column (same [slot][thread] layout)
  unsigned short* bmt = reinterpret_cast<unsigned short*>(bm_all + size_t(h) * pair_bitmap_bytes(nchunk)) + ht;
  const int bits0 = __float_as_int(kHTmin + kHMagic);
  ListEntry* list = P.lists + (long long)blockIdx.x * kListCap;
  int phase = 0;
  unsigned seq = 0;
#pragma unroll 1
  for (int pr = blockIdx.x; pr < npairs; pr += gridDim.x) {
    const int job = 2 * pr + h;
    const bool have = job < njobs;
    const RowJob jb = decode_job(have ? job : 0, gamma);
    const TIn* row = split_row<TIn>(P, jb);
    const float temp = jb.is_target ? P.t_temp : P.d_temp;
    const float top_p = jb.is_target ? P.t_top_p : P.d_top_p;
    const bool nucleus = top_p > 0.0f;

    // the drafted token's logit is needed when the row closes: fetch it now (first thread of the half)
    long long tok = -1;
    float xtok = 0.0f;
    if (have && ht == 0 && jb.j < gamma) {
      tok = P.tokens[(long long)jb.b * P.tok_sb + jb.j];
      if (tok >= 0 && tok < V) xtok = scale_logit(to_f32(__ldcg(row + tok)), temp);
    }
    const uint4 hint = __ldg(reinterpret_cast<const uint4*>(&P.hints[jb.is_target]));
    const float m0 = int(hint.w) != 0 ? __uint_as_float(hint.x) : 0.0f;

    int state = 0, nt = 0;
    float xa = INFINITY, xb = INFINITY;  // the boundary bucket's interval of scaled logits
    float Z = 0.0f, mxl = m0, GB = 0.0f;
    double target = 0.0;
    if (have && !nucleus) {
      float z = 0.0f, mx = -INFINITY;
#pragma unroll 1
      for (int c = 0; c < nchunk; ++c) {
        PAIR_STAGE_BEGIN(c)
#define FP_PLAIN(x, i) z += row_exp(x, m0), mx = fmaxf(mx, x);
        PAIR_EACH(FP_PLAIN)
        PAIR_STAGE_END()
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        z += __shfl_xor_sync(0xffffffffu, z, o);
        mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
      }
      if (lane == 0) ps->hred[h][hw] = make_float2(mx, z);
      barh(h);
      mx = -INFINITY, z = 0.0f;
#pragma unroll
      for (int w = 0; w < kHW; ++w) mx = fmaxf(mx, ps->hred[h][w].x), z += ps->hred[h][w].y;
      Z = z;
      mxl = mx * kLog2e;
      state = (!(fabsf(mxl - m0) <= 32.0f) || !(Z < INFINITY) || !(Z > 0.0f)) ? 3 : 1;
    } else if (have) {
      // ---- pass 1: per-thread mass histogram over the exponent (the column held the last row's list: clear it)
#pragma unroll
      for (int b = 0; b < kHB; ++b) hp[b * kHT] = 0.0f;
#pragma unroll 1
      for (int c = 0; c < nchunk; ++c) {
        PAIR_STAGE_BEGIN(c)
#define FP_HIST(x, i)                                 \
  {                                                   \
    const float t = fmaf(x, kLog2e, -m0);             \
    float* a = hp + (hist_bits(t) - bits0) * kHT;     \
    *a += ex2_approx(t);                              \
  }
// two entries per round: both bucket cells are read before either is written (the second takes the
// first's sum when the two fall into the same cell), which halves the dependent shared-memory chain
#define FP_HIST2(xa_, xb_)                                          \
  {                                                                 \
    const float t0 = fmaf(xa_, kLog2e, -m0), t1 = fmaf(xb_, kLog2e, -m0); \
    float* a0 = hp + (hist_bits(t0) - bits0) * kHT;                 \
    float* a1 = hp + (hist_bits(t1) - bits0) * kHT;                 \
    const float h0 = *a0, h1 = *a1;                                 \
    const float s0 = h0 + ex2_approx(t0);                           \
    const float s1 = (a1 == a0 ? s0 : h1) + ex2_approx(t1);         \
    *a0 = s0;                                                       \
    *a1 = s1;                                                       \
  }
#define FP_HIST4(v) FP_HIST2(v.x, v.y) FP_HIST2(v.z, v.w)
#ifndef PAIR_RMW1  // (default: two entries per histogram round, -0.2 us per C2 step)
        PAIR_EACH4(FP_HIST4)
#else
        PAIR_EACH(FP_HIST)
#endif
        PAIR_STAGE_END()
      }
      barh(h);
      if (ht == 0) STAMP(P.stats, 9);
      // bucket sums: warp w of the half owns buckets 8w .. 8w+7 (fixed order: deterministic)
#pragma unroll
      for (int q = 0; q < kHB / kHW; ++q) {
        const int b = hw * (kHB / kHW) + q;
        float s = 0.0f;
#pragma unroll
        for (int k = 0; k < kHT / 32; ++k) {
          s += hist[b * kHT + k * 32 + lane];
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        if (lane == 0) ps->hsum[h][b] = s;
      }
      barh(h);
      // every warp scans the 32 bucket sums from the top (lane l holds bucket 31 - l)
      const float hb = ps->hsum[h][kHB - 1 - lane];
      float incl = hb;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const float o = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += o;
      }
      Z = __shfl_sync(0xffffffffu, incl, 31);
      const unsigned nonempty = __ballot_sync(0xffffffffu, hb > 0.0f);
      const int btop = nonempty ? kHB - 1 - (__ffs(nonempty) - 1) : -1;
      mxl = m0 + (btop == kHB - 1 ? fmaxf(kHTmax, log2f(__shfl_sync(0xffffffffu, hb, 0))) : float(btop) + kHTmin);
      bool slow = !(Z < INFINITY) || !(Z > 0.0f);
      target = double(top_p) * double(Z);
      int bstar = -1;
      if (!slow && target < double(Z)) {
        const unsigned crossed = __ballot_sync(0xffffffffu, double(incl) > target);
        if (crossed) {
          const int l = __ffs(crossed) - 1;
          bstar = kHB - 1 - l;
          // mass above the crossing bucket = the previous lane's inclusive sum (not incl - hb, which is rounded twice)
          const float prev = __shfl_up_sync(0xffffffffu, incl, 1);
          GB = l == 0 ? 0.0f : __shfl_sync(0xffffffffu, prev, l);
        }
        if (bstar <= 0) slow = true;  // the lower end bucket is unbounded below
      }
      // ---- pass 2: the boundary bucket's entries (always consumed: the producer streams the row twice).
      // The bucket is an interval [xa, xb) of scaled logits (the bucket function is monotone): two compares.
      xa = INFINITY, xb = INFINITY;
      if (bstar > 0 && !slow) {
        if (!bucket_edge(bits0 + bstar, m0, &xa)) slow = true;
        if (bstar < kHB - 1 && !bucket_edge(bits0 + bstar + 1, m0, &xb)) slow = true;
        if (slow) xa = INFINITY;
      }
      // Per stage: every entry of the interval goes to the thread's OWN list column with a predicated store
      // (5 instructions per entry, no atomics, no shuffles, no divergent trips); the stage's 16-bit mask of
      // captured positions goes to the bit map, from which the select recovers the indices it needs.
      const unsigned lp0 = smem_u32(col), lim = lp0 + unsigned(kColSlots - 4) * unsigned(kHT * 4);
      unsigned lp = lp0;
#ifndef ABL_NOPASS2
#pragma unroll 1
      for (int c = 0; c < nchunk; ++c) {
        PAIR_STAGE_BEGIN(c)
        unsigned hm = 0u;
#define COL_GROUP(u)                                                  \
  {                                                                   \
    const float xe = lp < lim ? xa : INFINITY; /* column full: stop */ \
    col_take(lp, hm, v4[u].x, xe, xb, 1u << (4 * u));                  \
    col_take(lp, hm, v4[u].y, xe, xb, 2u << (4 * u));                  \
    col_take(lp, hm, v4[u].z, xe, xb, 4u << (4 * u));                  \
    col_take(lp, hm, v4[u].w, xe, xb, 8u << (4 * u));                  \
  }
        if (i00 + 3 * kHT * 4 < V) {  // every stage but a row's last one is whole
          COL_GROUP(0) COL_GROUP(1) COL_GROUP(2) COL_GROUP(3)
        } else {
          if (i00 < V) COL_GROUP(0)
          if (i00 + kHT * 4 < V) COL_GROUP(1)
          if (i00 + 2 * kHT * 4 < V) COL_GROUP(2)
          if (i00 + 3 * kHT * 4 < V) COL_GROUP(3)
        }
        bmt[c * kHT] = (unsigned short)hm;
        PAIR_STAGE_END()
      }
#endif
      nt = int((lp - lp0) / unsigned(kHT * 4));
      if (__any_sync(0xffffffffu, lp >= lim) && lane == 0) ps->colfull[h] = 1;
      if (ht == 0) STAMP(P.stats, 10);
      state = slow ? 3 : (bstar > 0 ? 2 : 4);
#if defined(ABL_NOPASS2) || defined(ABL_NOSELECT) || defined(ABL_NOPICK)
      state = 4;
#endif
    }
    // ---- each half closes its own row: the two selects of a pair run side by side
    RowRecord rec;
    rec.m0 = m0, rec.Z = Z, rec.xstar = -INFINITY, rec.istar = 0x7fffffff;
    rec.sel = 0.0f, rec.mxl = mxl, rec.dens = 0.0f;
    if (state == 2) {
      barh(h);  // every column and bit map of the half is complete
      if (ps->colfull[h]) {
        state = 3;  // a thread ran out of list slots: general search below
      } else {
        int ncand = 0;
        const bool ok = half_select(col, nt, bmt, h, ht, xa, xb, GB, Z, target, m0, hsc + h, rec, &ncand, P.stats);
        if (!ok) state = 3;  // more equal values than the finish can order: general search below
        if (P.stats && ht == 0) {
          atomicAdd(&P.stats[0], 1);
          atomicAdd(&P.stats[ok ? 16 : 17], 1);
          atomicAdd(&P.stats[18], ncand);
        }
      }
    }
    if (ht == 0) {
      HalfRow hr;
      hr.target = target, hr.GB = GB, hr.Z = Z, hr.m0 = m0, hr.mxl = mxl, hr.state = state;
      hr.pad[0] = hr.pad[1] = hr.pad[2] = 0;
      ps->half[h] = hr;
      ps->colfull[h] = 0;  // (read above, behind the half barrier; the next row sets it again after the join)
      if (state == 1 || state == 2 || state == 4) {
        if (tok >= 0 && tok < V) {
          const float e = kept(rec, xtok, int(tok)) ? row_exp(xtok, rec.m0) : 0.0f;
          rec.sel = __fdiv_rn(e, rec.Z);
        }
        store_record(&P.records[2 * pr + h], rec);
      }
    }
    if (ht == 0) STAMP(P.stats, 11);
    bar3();  // ---- the halves join: rows that need the whole block (rare) are finished one after the other

    const int st[2] = {ps->half[0].state, ps->half[1].state};
#pragma unroll 1
    for (int hh = 0; hh < 2; ++hh) {
      if (st[hh] != 3) continue;
      // general search by all 256 workers (rare): exact row maximum first, then the windowed search through L2
      const RowJob j2 = decode_job(2 * pr + hh, gamma);
      bar3();
      RowRecord rec = general_row<TIn>(split_row<TIn>(P, j2), V, j2.is_target ? P.t_temp : P.d_temp,
                                       j2.is_target ? P.t_top_p : P.d_top_p,
                                       (j2.is_target ? P.t_top_p : P.d_top_p) > 0.0f, ps->half[hh].m0, list,
                                       &ps->nlist, sc, P.stats);
      if (tid == hh * kHT) {
        if (tok >= 0 && tok < V) {
          const float e = kept(rec, xtok, int(tok)) ? row_exp(xtok, rec.m0) : 0.0f;
          rec.sel = __fdiv_rn(e, rec.Z);
        }
        store_record(&P.records[2 * pr + hh], rec);
      }
    }
    bar3();  // lists, scratch and the half records are free for the next pair
  }
}

}  // namespace pair
}  // namespace duodec
