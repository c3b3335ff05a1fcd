// Row-capture scan kernel: the streaming half of duodec_verify_fused, device path 6 (default).
//
// ONE pass over every logit row.  A block works on two rows at a time (two halves of 128 worker
// threads, each with its own TMA stage ring and producer warp; two blocks per SM = four rows in
// flight per SM), like the row-pair kernel it replaces, but a row is read once and nothing is
// predicted from earlier launches:
//
//   sample   before a row is streamed its producer fetches eight 256-entry pieces spread evenly over
//            the row (one ring stage).  Every thread takes the maximum of its 16 sample entries; a
//            warp-wide bitonic sort of those maxima gives the row's exponent offset M0 (largest sample
//            entry) and a capture cut c: the order statistic of the group maxima that corresponds to
//            the top `f` of the row by COUNT (f = min(1/12, 2667/V); distribution-free).
//   stream   e = 2^(x*log2e - M0) is summed in registers (FFMA + MUFU.EX2 + FADD); entries with
//            x >= c are appended to the thread's own shared-memory list column (predicated store, no
//            atomics, no shuffles) and marked in a V-bit map (16 bits per thread and stage) from which
//            their indices are recovered later.  7 instructions per logit.
//   select   on the ~2700 captured entries only: two or three levels of 16 value buckets with
//            per-thread float histograms locate the bucket in which the kept mass crosses
//            top_p * Z; the mass above it and the <= 128 entries inside it are then handled in exact
//            38-bit fixed-point arithmetic and ordered like the reference's sort (scaled logit
//            descending, lowest index first) - same arithmetic and answer as the row-pair kernel.
//   dump     a target row also leaves its list columns and bit map in the workspace (two TMA bulk
//            stores issued by one thread while the select runs): the sample kernel draws the sequence's
//            token from the kept entries of that dump instead of re-reading three V-sized rows.
//
// A thread whose column is full appends to a small overflow pool of its half (atomic append, rare).
// A row whose nucleus is not inside the captured set (flat rows, a full overflow pool, overflowing
// exponentials) is finished by the general windowed search through L2
// (general_row, pair_kernel.cuh), exactly as before.
//
// Reference semantics restated: src/utils.py:650-709 (logits / T, top-p filter, softmax); the
// record a row leaves behind is the one seq_sample_tma_kernel evaluates (src/decoding_ops.py:127-220).
#pragma once

#include "pair_kernel.cuh"

namespace duodec {
namespace cap {

constexpr int kHT = 128;            // worker threads per half
constexpr int kHW = kHT / 32;
constexpr int kPT = kT3 + 64;       // 256 workers + one producer warp per half
constexpr int kCStages = 2;         // 8 KB stages per half
constexpr int kSlots = 40;          // list slots per worker thread (the last 4 are slack: the column is re-checked per float4)
constexpr int kOvf = 192;           // entries of a half's overflow pool: what does not fit a thread's column (x, index)
// Dump of a closed target row in its 32 KB workspace slot (read by the sample kernel): the list columns
// [kSlots][128] float, the bit map [nchunk][128] u16, the pool (count, entries).
constexpr int kDumpBitmap = kSlots * kHT * 4;          // 20480
constexpr int kDumpPool = 30720;                       // int count, then entries at +16
constexpr int kDumpMaxChunks = (kDumpPool - kDumpBitmap) / (kHT * 2);  // 40 stages: V <= 81920
static_assert(kDumpPool + 16 + kOvf * 8 <= kListCap * 8, "dump must fit the row's workspace slot");
constexpr int kLB = 16;             // value buckets per select level
constexpr int kMaxLevels = 5;
constexpr int kSamplePieces = 8;    // sample stage: 8 pieces of 256 entries
constexpr int kPieceElems = kStageElems / kSamplePieces;
constexpr int kMinV = 2 * kStageElems;  // smaller vocabularies take the general kernel

struct __align__(16) CapSync {
  unsigned long long full[2][kCStages];
  unsigned long long empty[2][kCStages];
  float2 pro[2][kHW];    // prologue: (cut of the warp, largest sample entry of the warp)
  float2 post[2][kHW];   // after the stream: (sum of exponentials, overflow flag)
  float2 mm[2][kHW];     // select: (max, -min) of the captured entries
  float hsum[2][kLB];
  pair::HalfRow half[2];
  int nlist;
  int pad[3];
};

struct __align__(16) CapScratch {  // one per half
  unsigned long long cand_f[kHT];
  float cand_x[kHT];
  int cand_i[kHT];
  unsigned long long wgb[kHW];
  int wcnt[kHW];
  unsigned long long gb_new;
  int ncand, res_ok, res_i, nemit;
  float res_x;
  int novf;  // entries appended to the overflow pool (may exceed kOvf: then the row takes the general search)
  int pad[2];
  ListEntry ovf[kOvf];
};

__host__ __device__ constexpr size_t cap_bitmap_bytes(int nchunk) { return (size_t(nchunk) * kHT * 2 + 15) & ~size_t(15); }

// ring | value lists | level histograms (aliased by the general search's scratch) | bit maps | sync | select scratch
__host__ __device__ constexpr size_t cap_smem_bytes(int nchunk) {
  return size_t(2) * kCStages * kStageBytes + size_t(2) * kSlots * kHT * 4 + size_t(2) * kLB * kHT * 4 +
         2 * cap_bitmap_bytes(nchunk) + sizeof(CapSync) + 2 * sizeof(CapScratch) + 128;
}
static_assert(sizeof(Scratch3) <= size_t(2) * kLB * kHT * 4, "general-search scratch must fit under the level histograms");

__device__ __forceinline__ void barh(int h) { pair::barh(h); }

// one stage = 16 entries per half-thread: four 128-bit shared loads at (u * 128 + ht) * 4
#define CAP_STAGE_BEGIN(c)                                                                    \
  const unsigned slot = seq % unsigned(kCStages);                                             \
  mbar_wait(&cs->full[h][slot], (seq / unsigned(kCStages)) & 1u);                              \
  const unsigned char* st = ring + size_t(slot) * kStageBytes;                                \
  float4 v4[4];                                                                               \
  _Pragma("unroll") for (int u = 0; u < 4; ++u) v4[u] = unpack4<TIn>(st + size_t((u * kHT + ht) * 4) * kElt); \
  if (temp != 1.0f) {                                                                         \
    _Pragma("unroll") for (int u = 0; u < 4; ++u) {                                           \
      v4[u].x = __fdiv_rn(v4[u].x, temp), v4[u].y = __fdiv_rn(v4[u].y, temp);                 \
      v4[u].z = __fdiv_rn(v4[u].z, temp), v4[u].w = __fdiv_rn(v4[u].w, temp);                 \
    }                                                                                         \
  }
#define CAP_STAGE_END()                                                                       \
  __syncwarp();                                                                               \
  if (lane == 0) mbar_arrive(&cs->empty[h][slot]);                                            \
  ++seq;

// TMA bulk store shared -> global (SASS UBLKCP), bulk-group completion
__device__ __forceinline__ void bulk_s2g(void* dst, const void* src, unsigned bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(smem_u32(src)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }

// x >= ce: append x to the thread's list column, advance the column pointer, mark the entry's bit
__device__ __forceinline__ void cap_one(unsigned& lp, unsigned& hm, float x, float ce, unsigned bit) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ge.f32 p, %2, %3;\n\t@p st.shared.f32 [%0], %2;\n\t@p add.u32 %0, %0, 512;\n\t@p or.b32 "
      "%1, %1, %4;\n\t}"
      : "+r"(lp), "+r"(hm)
      : "f"(x), "f"(ce), "r"(bit)
      : "memory");
}

// A thread whose column filled up stopped capturing at a float4-group boundary: every group after the one
// that holds its last captured entry was skipped.  Cold path, after the stream: those groups are read
// again (plain loads, L2) and their entries above the cut go to the half's overflow pool with their index.
template <class TIn>
__device__ __noinline__ void cap_spill_rest(CapScratch* hs, const unsigned short* bmt, const TIn* row, int V,
                                            int nchunk, float temp, float cut, int ht) {
  int c = nchunk - 1;
  unsigned hm = 0u;
  for (; c >= 0; --c) {
    hm = bmt[c * kHT];
    if (hm != 0u) break;
  }
  int g = c < 0 ? 0 : c * 4 + ((31 - __clz(hm)) >> 2) + 1;  // first skipped group (stage * 4 + u)
  for (; g < nchunk * 4; ++g) {
    const int i0 = (g >> 2) * kStageElems + ((g & 3) * kHT + ht) * 4;
    if (i0 >= V) continue;
    const float4 v = load4_scaled<TIn>(row + i0, temp);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float x = k == 0 ? v.x : (k == 1 ? v.y : (k == 2 ? v.z : v.w));
      if (x >= cut) {
        const int pos = atomicAdd(&hs->novf, 1);
        if (pos < kOvf) hs->ovf[pos] = ListEntry{x, i0 + k};
      }
    }
  }
}

__device__ __forceinline__ float warp_sort_asc(float v, int lane) {
#pragma unroll
  for (int k = 2; k <= 32; k <<= 1) {
#pragma unroll
    for (int j = k >> 1; j > 0; j >>= 1) {
      const float o = __shfl_xor_sync(0xffffffffu, v, j);
      const bool up = (lane & k) == 0, lower = (lane & j) == 0;
      v = (lower == up) ? fminf(v, o) : fmaxf(v, o);
    }
  }
  return v;
}

// bucket(x) = clamp(int((x - A) * sc), 0, kLB - 1): monotone in x
__device__ __forceinline__ int lv_bucket(float A, float sc, float x) {
  const int bk = __float2int_rz((x - A) * sc);
  return bk < 0 ? 0 : (bk > kLB - 1 ? kLB - 1 : bk);
}

// index of the thread's j-th captured entry (capture order) from its bit-map column
__device__ __forceinline__ int cap_index_of(const unsigned short* bmt, int j, int ht) {
  int stage = 0;
  unsigned hm = bmt[0];
  while (__popc(hm) <= j) {
    j -= __popc(hm);
    ++stage;
    hm = bmt[stage * kHT];
  }
  const int k = __fns(hm, 0, j + 1);
  return stage * kStageElems + ((k >> 2) * kHT + ht) * 4 + (k & 3);
}

// Closes one row on the entries its half captured: nucleus boundary (rec.Z / xstar / istar) and, for a
// target row, the list of kept entries in the workspace.  The thread's entries are lst[j * 128], j < nt
// (capture order; bit-map column bmt gives their indices) plus its share of the half's overflow pool
// (entries ht, ht + 128, ... of hs->ovf, with explicit indices): local entry numbers 0 .. ne-1 < 64.
//
//   A  one loop over all entries: per-thread FLOAT histogram (private shared-memory column, no atomics,
//      fixed summation order: deterministic) of e = 2^(x*log2e - m0) over 16 value buckets of [cut, max].
//   B  bucket sums, scanned from the top: the bucket b* in which the kept mass crosses top_p * Z and the
//      mass G above it.
//   C  one loop over all entries: a 64-bit mask per thread of the entries in b* (candidates).  From here
//      on only mask bits are visited.
//   D  while more than 128 candidates remain: 16 sub-buckets of the crossing bucket, histogram over the
//      candidates only, same scan.
//   F  the <= 128 candidates left are ordered exactly like the reference's sort (scaled logit descending,
//      lowest index first) by an all-pairs rank in 38-bit fixed point (integer sums: exactly one entry
//      crosses), as in pair::half_select.
// Returns false when the row needs the general search (nucleus not inside the captured set, more equal
// values than the finish can order, float scan and exact finish disagreeing at a bucket edge).
__device__ __noinline__ bool cap_close(const float* lst, int nt, const unsigned short* bmt, int novf, int h, int ht,
                                       float* hist, float cut, float xmax, float Zrow, double target, float m0,
                                       CapSync* cs, CapScratch* hs, RowRecord& rec, int* stats) {
  const int lane = ht & 31, hw = ht >> 5;
  const int ne = nt + (novf - ht + kHT - 1) / kHT;
  float* hp = hist + ht;
  auto getx = [&](int j) { return j < nt ? lst[j * kHT] : hs->ovf[(j - nt) * kHT + ht].x; };
  auto geti = [&](int j) { return j < nt ? cap_index_of(bmt, j, ht) : hs->ovf[(j - nt) * kHT + ht].idx; };
  if (!(cut > -INFINITY) || !(xmax >= cut) || !(xmax < INFINITY)) return false;
  float fA = cut, fsc = float(kLB) / (nextafterf(xmax, INFINITY) - cut);
  if (!(fsc < INFINITY)) return false;
  if (ht == 0) hs->ncand = 0, hs->res_ok = 0;

  // ---- A: level-0 histogram over all entries
#pragma unroll
  for (int b = 0; b < kLB; ++b) hp[b * kHT] = 0.0f;
#pragma unroll 2
  for (int j = 0; j < ne; ++j) {
    const float x = getx(j);
    hp[lv_bucket(fA, fsc, x) * kHT] += row_exp(x, m0);
  }
  unsigned long long candm = 0ull;
  float G = 0.0f;  // mass above the current crossing bucket (float, fixed order)
  int total = 0;
#pragma unroll 1
  for (int level = 0; level < kMaxLevels; ++level) {
    if (level > 0) {
      // ---- D: histogram of the candidates over the sub-buckets of the crossing bucket, and their number
#pragma unroll
      for (int b = 0; b < kLB; ++b) hp[b * kHT] = 0.0f;
      for (unsigned long long m = candm; m; m &= m - 1ull) {
        const float x = getx(__ffsll((long long)m) - 1);
        hp[lv_bucket(fA, fsc, x) * kHT] += row_exp(x, m0);
      }
      int cnt = __popcll(candm);
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
      if (lane == 0) hs->wcnt[hw] = cnt;
    }
    barh(h);  // histogram columns (and the candidate counts) complete
    if (ht == 0) STAMP(stats, level == 0 ? 26 : 28);
    if (level > 0) {
      total = 0;
#pragma unroll
      for (int w = 0; w < kHW; ++w) total += hs->wcnt[w];
      if (total <= kHT) break;  // few enough for the all-pairs finish
    }
    // ---- B: bucket sums (warp w owns buckets 4w .. 4w+3), then every thread scans them from the top
#pragma unroll
    for (int q = 0; q < kLB / kHW; ++q) {
      const int b = hw * (kLB / kHW) + q;
      float s = 0.0f;
#pragma unroll
      for (int k = 0; k < kHT / 32; ++k) s += hist[b * kHT + k * 32 + lane];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
      if (lane == 0) cs->hsum[h][b] = s;
    }
    barh(h);
    float run = G;
    int bstar = -1;
#pragma unroll
    for (int b = kLB - 1; b >= 0; --b) {
      const float s = cs->hsum[h][b];
      if (bstar < 0 && s > 0.0f && double(run + s) > target) bstar = b;
      if (bstar < 0) run += s;
    }
    if (ht == 0) STAMP(stats, 27);
    if (bstar < 0) return false;  // the kept mass does not cross inside the captured entries
    G = run;
    // ---- C: classify - above the crossing bucket: kept for sure; inside it: candidate
    if (level == 0) {
#pragma unroll 2
      for (int j = 0; j < ne; ++j) {
        candm |= (unsigned long long)(lv_bucket(fA, fsc, getx(j)) == bstar) << j;
      }
    } else {
      for (unsigned long long m = candm; m; m &= m - 1ull) {
        const int j = __ffsll((long long)m) - 1;
        if (lv_bucket(fA, fsc, getx(j)) != bstar) candm &= ~(1ull << j);
      }
    }
    fA = fmaf(float(bstar), 1.0f / fsc, fA);
    fsc = fsc * float(kLB);
    // (the next level's barrier orders these reads of hsum / hist before they are overwritten)
  }
  if (total > kHT || total <= 0) return false;  // more (equal) values than the finish can order
  // ---- F: the candidates left, with their indices, in the reference's order
  const int ez = int(__float_as_uint(Zrow) >> 23) - 126;               // Z = f * 2^ez, f in [0.5, 1)
  const float S = __uint_as_float(unsigned(127 + 38 - ez) << 23);      // 2^(38 - ez)
  const unsigned long long tgt = (unsigned long long)(target * double(S));  // floor
  const unsigned long long GBf = (unsigned long long)(double(G) * double(S));
  for (unsigned long long m = candm; m; m &= m - 1ull) {
    const int j = __ffsll((long long)m) - 1;
    const float x = getx(j);
    const int pos = atomicAdd(&hs->ncand, 1);
    if (pos < kHT) {
      hs->cand_x[pos] = x;
      hs->cand_f[pos] = __float2ull_rn(row_exp(x, m0) * S);
      hs->cand_i[pos] = geti(j);
    }
  }
  barh(h);
  if (ht == 0) STAMP(stats, 29);
  const int n = hs->ncand;
  if (n > kHT || n <= 0) return false;
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
  const unsigned long long incl = GBf + cum;
  const bool mine = t < n && part == 0;
  if (mine && incl > tgt && incl - fme <= tgt) {
    hs->res_x = xx;
    hs->res_i = ii;
    hs->gb_new = incl;
    hs->res_ok = 1;
  }
  barh(h);
  if (!hs->res_ok) return false;  // the float scan chose a neighbouring bucket (edge case): general search
  rec.Z = float(double(hs->gb_new) / double(S));
  rec.xstar = hs->res_x;
  rec.istar = hs->res_i;
  return true;
}

template <class TIn>
__global__ void __launch_bounds__(kPT, 2) row_capture_kernel(const __grid_constant__ SplitParams P) {
  extern __shared__ __align__(128) unsigned char smem[];
  const int V = P.V, gamma = P.gamma;
  const int nchunk = (V + kStageElems - 1) / kStageElems;
  float* lists = reinterpret_cast<float*>(smem + size_t(2) * kCStages * kStageBytes);  // [2][kSlots][128]
  float* hist_all = lists + 2 * kSlots * kHT;                                            // [2][kLB][128]
  Scratch3* sc = reinterpret_cast<Scratch3*>(hist_all);  // only the general search (both halves joined) uses it
  unsigned char* bm_all = reinterpret_cast<unsigned char*>(hist_all + 2 * kLB * kHT);
  CapSync* cs = reinterpret_cast<CapSync*>(bm_all + 2 * cap_bitmap_bytes(nchunk));
  CapScratch* hsc = reinterpret_cast<CapScratch*>(cs + 1);

  const int tid = threadIdx.x;
  const int njobs = P.B * (2 * gamma + 1);
  const int npairs = (njobs + 1) / 2;
  constexpr int kElt = int(sizeof(TIn));

  if (tid == 0) {
    for (int hh = 0; hh < 2; ++hh)
      for (int s = 0; s < kCStages; ++s) {
        mbar_init(&cs->full[hh][s], 1);
        mbar_init(&cs->empty[hh][s], kHW);
      }
    mbar_fence_init();
    cs->nlist = 0;
    hsc[0].novf = 0, hsc[1].novf = 0;
  }
  __syncthreads();
  // nothing above touches global memory; inputs and the workspace are ordered from here on
  grid_dep_wait();
  grid_dep_launch();
  if (tid == 0) STAMP(P.stats, 8, true);

  if (tid >= kT3) {
    // ---- producers: lane 0 of warp 8 feeds half 0, lane 0 of warp 9 feeds half 1
    const int h = (tid - kT3) >> 5;
    if (((tid - kT3) & 31) == 0) {
      unsigned char* ring = smem + size_t(h) * kCStages * kStageBytes;
      unsigned seq = 0;
      const int pstep = ((V / kSamplePieces) / kPieceElems) * kPieceElems;  // > 0: V >= kMinV
      for (int pr = blockIdx.x; pr < npairs; pr += gridDim.x) {
        const int job = 2 * pr + h;
        if (job >= njobs) continue;
        const RowJob jb = decode_job(job, gamma);
        const TIn* row = split_row<TIn>(P, jb);
        for (int c = -1; c < nchunk; ++c, ++seq) {
          const unsigned slot = seq % unsigned(kCStages);
          if (seq >= unsigned(kCStages)) mbar_wait(&cs->empty[h][slot], ((seq / unsigned(kCStages)) - 1u) & 1u);
          unsigned char* dst = ring + size_t(slot) * kStageBytes;
          if (c < 0) {  // the sample stage: eight pieces spread over the row
            mbar_expect_tx(&cs->full[h][slot], unsigned(kStageElems * kElt));
            for (int k = 0; k < kSamplePieces; ++k)
              bulk_g2s(dst + size_t(k) * kPieceElems * kElt, row + size_t(k) * pstep, unsigned(kPieceElems * kElt),
                       &cs->full[h][slot]);
          } else {
            const int e0 = c * kStageElems;
            const int ne = (V - e0) < kStageElems ? (V - e0) : kStageElems;
            mbar_expect_tx(&cs->full[h][slot], unsigned(ne * kElt));
            bulk_g2s(dst, row + e0, unsigned(ne * kElt), &cs->full[h][slot]);
          }
        }
      }
    }
    return;
  }

  // ---- workers: half h = threads [128 h, 128 h + 128)
  const int h = tid >> 7, ht = tid & (kHT - 1), lane = tid & 31, hw = ht >> 5;
  unsigned char* ring = smem + size_t(h) * kCStages * kStageBytes;
  float* lst = lists + h * kSlots * kHT + ht;              // this thread's list column
  float* hist = hist_all + h * kLB * kHT;
  unsigned short* bmt = reinterpret_cast<unsigned short*>(bm_all + size_t(h) * cap_bitmap_bytes(nchunk)) + ht;
  const unsigned lp0 = smem_u32(lst), lim = lp0 + unsigned(kSlots - 4) * unsigned(kHT * 4);
  CapScratch* hs = hsc + h;
  ListEntry* glist = P.lists + (long long)(P.B * (gamma + 1) + int(blockIdx.x)) * kListCap;  // the general search's scratch list
  unsigned seq = 0;

#pragma unroll 1
  for (int pr = blockIdx.x; pr < npairs; pr += gridDim.x) {
    const int job = 2 * pr + h;
    const bool have = job < njobs;
    const RowJob jb = decode_job(have ? job : 0, gamma);
    const TIn* row = split_row<TIn>(P, jb);
    const float temp = jb.is_target ? P.t_temp : P.d_temp;
    const float top_p = jb.is_target ? P.t_top_p : P.d_top_p;
    const bool nucleus = top_p > 0.0f && top_p < 1.0f;

    // the drafted token's logit is needed when the row closes: fetch it now (first thread of the half)
    long long tok = -1;
    float xtok = 0.0f;
    if (have && ht == 0 && jb.j < gamma) {
      tok = P.tokens[(long long)jb.b * P.tok_sb + jb.j];
      if (tok >= 0 && tok < V) xtok = scale_logit(to_f32(__ldcg(row + tok)), temp);
    }

    int state = 0, nt = 0, novf = 0;
    bool dumped = false;
    float Z = 0.0f, m0 = 0.0f, cut = INFINITY, xmax = -INFINITY;
    double target = 0.0;
    if (have) {
      // ---- sample stage: exponent offset and capture cut from the row itself
      {
        CAP_STAGE_BEGIN(-1)
        float m16 = -INFINITY;
#pragma unroll
        for (int u = 0; u < 4; ++u) m16 = fmaxf(m16, fmaxf(fmaxf(v4[u].x, v4[u].y), fmaxf(v4[u].z, v4[u].w)));
        CAP_STAGE_END()
        const float sorted = warp_sort_asc(m16, lane);
        const float cw = __shfl_sync(0xffffffffu, sorted, P.cap_rank);
        const float mw = __shfl_sync(0xffffffffu, sorted, 31);
        if (lane == 0) cs->pro[h][hw] = make_float2(cw, mw);
        barh(h);
        float csum = 0.0f, mx = -INFINITY;
#pragma unroll
        for (int w = 0; w < kHW; ++w) csum += cs->pro[h][w].x, mx = fmaxf(mx, cs->pro[h][w].y);
        cut = csum * (1.0f / kHW);
        xmax = mx;
        m0 = mx * kLog2e;
      }
      if (!nucleus) cut = INFINITY;
      // (hs->novf of this half is zero here: set below at the row's end, ordered by the half barriers in between)
      // ---- the stream: sum of exponentials in registers, entries above the cut into the list column
      float z0 = 0.0f, z1 = 0.0f, z2 = 0.0f, z3 = 0.0f;
      unsigned lp = lp0;
#pragma unroll 1
      for (int c = 0; c < nchunk; ++c) {
        CAP_STAGE_BEGIN(c)
        const int i00 = c * kStageElems + ht * 4;
        unsigned hm = 0u;
#define CAP_GROUP(u)                                                       \
  {                                                                        \
    const float ce = lp < lim ? cut : INFINITY;                            \
    cap_one(lp, hm, v4[u].x, ce, 1u << (4 * u));                           \
    cap_one(lp, hm, v4[u].y, ce, 2u << (4 * u));                           \
    cap_one(lp, hm, v4[u].z, ce, 4u << (4 * u));                           \
    cap_one(lp, hm, v4[u].w, ce, 8u << (4 * u));                           \
    z0 += row_exp(v4[u].x, m0), z1 += row_exp(v4[u].y, m0);                \
    z2 += row_exp(v4[u].z, m0), z3 += row_exp(v4[u].w, m0);                \
  }
        if (i00 + 3 * kHT * 4 < V) {  // every stage but a row's last one is whole
          CAP_GROUP(0) CAP_GROUP(1) CAP_GROUP(2) CAP_GROUP(3)
        } else {
          if (i00 < V) CAP_GROUP(0)
          if (i00 + kHT * 4 < V) CAP_GROUP(1)
          if (i00 + 2 * kHT * 4 < V) CAP_GROUP(2)
          if (i00 + 3 * kHT * 4 < V) CAP_GROUP(3)
        }
        bmt[c * kHT] = (unsigned short)hm;
        CAP_STAGE_END()
      }
      float z = (z0 + z1) + (z2 + z3);
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) z += __shfl_xor_sync(0xffffffffu, z, o);
      if (lp >= lim) cap_spill_rest<TIn>(hs, bmt, row, V, nchunk, temp, cut, ht);
      if (lane == 0) cs->post[h][hw] = make_float2(z, 0.0f);
      const bool dump = nucleus && jb.is_target && nchunk <= kDumpMaxChunks;
      if (dump) asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // list / bit-map writes -> the async proxy
      barh(h);  // also: every list column and bit map of the half is complete
      if (dump && ht == 0) {
        // the row's capture state goes to its workspace slot while the select runs (nobody writes it until the row closes)
        unsigned char* slot = reinterpret_cast<unsigned char*>(P.lists + (long long)(jb.b * (gamma + 1) + jb.j) * kListCap);
        bulk_s2g(slot, lists + h * kSlots * kHT, unsigned(kSlots * kHT * 4));
        bulk_s2g(slot + kDumpBitmap, bm_all + size_t(h) * cap_bitmap_bytes(nchunk), unsigned(cap_bitmap_bytes(nchunk)));
        bulk_commit();
      }
      dumped = dump;
      if (ht == 0) STAMP(P.stats, 9);
      z = 0.0f;
#pragma unroll
      for (int w = 0; w < kHW; ++w) z += cs->post[h][w].x;
      Z = z;
      novf = hs->novf;
      nt = int((lp - lp0) / unsigned(kHT * 4));
      const bool bad = !(Z < INFINITY) || !(Z > 0.0f);
      if (!nucleus) state = bad ? 3 : (top_p >= 1.0f ? 4 : 1);
      else state = (bad || novf > kOvf) ? 3 : 2;
      target = double(top_p) * double(Z);
    }
    // ---- each half closes its own row
    RowRecord rec;
    rec.m0 = m0, rec.Z = Z, rec.xstar = -INFINITY, rec.istar = 0x7fffffff;
    rec.sel = 0.0f, rec.mxl = m0, rec.dens = 0.0f, rec.pad = 0;
    if (state == 2) {
      const bool ok = cap_close(lst, nt, bmt, novf, h, ht, hist, cut, xmax, Z, target, m0, cs, hs, rec, P.stats);
      if (!ok) state = 3;
      if (ht == 0) STAMP(P.stats, 10);
      if (ok && dumped) {
        // the overflow pool (rare) follows with plain stores; the record's flag tells the sample kernel to use the dump
        unsigned char* slot = reinterpret_cast<unsigned char*>(P.lists + (long long)(jb.b * (gamma + 1) + jb.j) * kListCap);
        ListEntry* gp = reinterpret_cast<ListEntry*>(slot + kDumpPool + 16);
        for (int o = ht; o < novf; o += kHT) gp[o] = hs->ovf[o];
        rec.pad = 1 + novf;
      }
      if (P.stats && ht == 0) atomicAdd(&P.stats[ok ? 16 : 17], 1);
    }
    if (ht == 0) {
      pair::HalfRow hr;
      hr.target = target, hr.GB = 0.0f, hr.Z = Z, hr.m0 = m0, hr.mxl = m0, hr.state = state;
      hr.pad[0] = hr.pad[1] = hr.pad[2] = 0;
      cs->half[h] = hr;
      if (state == 1 || state == 2 || state == 4) {
        if (tok >= 0 && tok < V) {
          const float e = kept(rec, xtok, int(tok)) ? row_exp(xtok, rec.m0) : 0.0f;
          rec.sel = __fdiv_rn(e, rec.Z);
        }
        store_record(&P.records[job], rec);
      }
    }
    if (ht == 0) {
      if (dumped) bulk_wait_all();  // the bulk stores are complete: the columns may be overwritten, the dump is in memory
      hs->novf = 0;  // every reader of this row's pool is behind a half barrier; the next row's appends are after the join
      STAMP(P.stats, 11);
    }
    bar3();  // ---- the halves join: rows that need the general search (rare) are finished one after the other

    const int st[2] = {cs->half[0].state, cs->half[1].state};
#pragma unroll 1
    for (int hh = 0; hh < 2; ++hh) {
      if (st[hh] != 3) continue;
      // general search by all 256 workers: exact row maximum first, then the windowed search through L2
      const RowJob j2 = decode_job(2 * pr + hh, gamma);
      const float tp2 = j2.is_target ? P.t_top_p : P.d_top_p;
      bar3();
      RowRecord r2 = pair::general_row<TIn>(split_row<TIn>(P, j2), V, j2.is_target ? P.t_temp : P.d_temp, tp2,
                                            tp2 > 0.0f, cs->half[hh].m0, glist, &cs->nlist, sc, P.stats);
      r2.pad = 0;
      if (tid == hh * kHT) {
        if (tok >= 0 && tok < V) {
          const float e = kept(r2, xtok, int(tok)) ? row_exp(xtok, r2.m0) : 0.0f;
          r2.sel = __fdiv_rn(e, r2.Z);
        }
        store_record(&P.records[2 * pr + hh], r2);
      }
      if (P.stats && tid == 0) atomicAdd(&P.stats[19], 1);
    }
    bar3();  // lists, scratch and the half records are free for the next pair
  }
}

}  // namespace cap
}  // namespace duodec
