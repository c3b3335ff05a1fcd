// Row-ring scan kernel: the streaming half of duodec_verify_fused (default path).
//
//   row_ring_kernel  persistent grid, two blocks per SM; every block OWNS whole logit rows
//                    (row = block + k * grid), so the hot loop has no global atomics, fences or
//                    cross-block hand-offs.  One producer lane keeps a ring of kStages (4) 8 KB
//                    shared-memory stages full with TMA bulk copies (cp.async.bulk + mbarrier
//                    complete_tx), running ahead across passes and row boundaries; 256 worker
//                    threads take 8 entries each per stage (two conflict-free 128-bit shared loads).
//
//   A row without nucleus filter is one pass:  Z = sum 2^(x*log2e - M0).
//   A row with top-p is streamed TWICE (the second time it comes out of L2, HBM sees it once):
//     pass 1  e = 2^(x*log2e - M0) is added to a PER-THREAD histogram of 32 buckets, one per
//             unit of the exponent t = x*log2e - M0 (rounded to nearest, clamped), in shared
//             memory (hist[bucket][thread]: bank = thread, no conflicts, no atomics).  The
//             bucket sums give Z and, scanned from the top, the one bucket in which the kept
//             mass crosses top_p * Z, together with the exact mass above it.  Nothing is
//             predicted: M0 only has to be within a few units of the row's maximum.
//     pass 2  the entries of that bucket (a few hundred of 32000) are appended to the block's list;
//             list_select (split_kernels.cuh) finds the boundary entry among them in exact
//             fixed-point arithmetic and the reference's order.
//   Rows that do not fit this scheme (boundary in a clamped end bucket, a crowded bucket, an
//   exponent offset far from the row's maximum) take the general search of split_kernels.cuh,
//   which re-reads the row through L2 with plain loads.
//
//   The row's 32-byte record (M0, kept mass, boundary, probability of the drafted token) is
//   what seq_sample_kernel (split_kernels.cuh) evaluates the accept tests and draws tokens from.
#pragma once

#include "split_kernels.cuh"

namespace duodec {

constexpr int kStageElems = 2048;                 // row entries per ring stage: 8 per worker thread
constexpr int kStageBytes = kStageElems * 4;      // sized for fp32 rows; bf16 rows fill half of it
#ifndef DUODEC_RING_STAGES
#define DUODEC_RING_STAGES 4
#endif
constexpr int kStages = DUODEC_RING_STAGES;
constexpr int kRingPer = kListCap / kT3;           // boundary-bucket entries a worker thread can keep (its list column)
constexpr int kRT = kT3 + 32;                     // 256 workers + the producer warp
constexpr int kHB = 32;                           // histogram buckets
constexpr float kHTmin = -27.0f, kHTmax = 4.0f;   // bucket b holds round(t) == b - 27 (ends clamped)
constexpr float kHMagic = 12582912.0f;            // 1.5 * 2^23: t + magic rounds t to an integer in the low bits

struct __align__(16) RingSync {
  unsigned long long full[kStages];   // "stage landed" (TMA complete_tx)
  unsigned long long empty[kStages];  // "stage consumed" (one arrive per worker warp)
  float hsum[kHB];
  int nlist;
  int pad[3];
};

constexpr size_t ring_smem_bytes() {
  return size_t(kStages) * kStageBytes + size_t(kHB) * kT3 * 4 + sizeof(Scratch3) + sizeof(RingSync) + 128;
}

struct MinMax {
  float mx, ngmn;  // max and max of the negated values
};
struct OpMinMax {
  __device__ __forceinline__ MinMax operator()(MinMax a, MinMax b) const {
    return MinMax{fmaxf(a.mx, b.mx), fmaxf(a.ngmn, b.ngmn)};
  }
};

__device__ __forceinline__ int hist_bits(float t) {
  return __float_as_int(fminf(fmaxf(t, kHTmin), kHTmax) + kHMagic);
}

// one ring stage = 8 entries per worker thread: two 128-bit shared loads (entries 4*tid.. and
// 4*(256+tid)..); the tail of the last stage of a row is stale shared memory and is skipped.
#define RING_STAGE_BEGIN(c)                                                                   \
  const unsigned slot = seq % unsigned(kStages);                                              \
  mbar_wait(&rs->full[slot], (seq / unsigned(kStages)) & 1u);                                  \
  const unsigned char* st = ring + size_t(slot) * kStageBytes;                                \
  float4 va = unpack4<TIn>(st + size_t(tid * 4) * kElt);                                      \
  float4 vb = unpack4<TIn>(st + size_t((kT3 + tid) * 4) * kElt);                              \
  const int ia = (c) * kStageElems + tid * 4, ib = ia + kT3 * 4;                               \
  if (temp != 1.0f) {                                                                         \
    va.x = __fdiv_rn(va.x, temp), va.y = __fdiv_rn(va.y, temp);                               \
    va.z = __fdiv_rn(va.z, temp), va.w = __fdiv_rn(va.w, temp);                               \
    vb.x = __fdiv_rn(vb.x, temp), vb.y = __fdiv_rn(vb.y, temp);                               \
    vb.z = __fdiv_rn(vb.z, temp), vb.w = __fdiv_rn(vb.w, temp);                               \
  }
#define RING_STAGE_END()                                                                      \
  __syncwarp();                                                                               \
  if (lane == 0) mbar_arrive(&rs->empty[slot]);                                               \
  ++seq;
#define RING_EACH(F)                                                                          \
  if (ia < V) { F(va.x, ia) F(va.y, ia + 1) F(va.z, ia + 2) F(va.w, ia + 3) }                 \
  if (ib < V) { F(vb.x, ib) F(vb.y, ib + 1) F(vb.z, ib + 2) F(vb.w, ib + 3) }

template <class TIn>
__global__ void __launch_bounds__(kRT, 2) row_ring_kernel(const __grid_constant__ SplitParams P) {
  extern __shared__ __align__(128) unsigned char smem[];
  unsigned char* ring = smem;
  float* hist = reinterpret_cast<float*>(smem + size_t(kStages) * kStageBytes);
  Scratch3* sc = reinterpret_cast<Scratch3*>(hist + kHB * kT3);
  RingSync* rs = reinterpret_cast<RingSync*>(sc + 1);

  const int tid = threadIdx.x;
  const int gamma = P.gamma, V = P.V;
  const int njobs = P.B * (2 * gamma + 1);
  const int nchunk = (V + kStageElems - 1) / kStageElems;
  constexpr int kElt = int(sizeof(TIn));

  if (tid == 0) {
    for (int s = 0; s < kStages; ++s) {
      mbar_init(&rs->full[s], 1);
      mbar_init(&rs->empty[s], kW3);
    }
    mbar_fence_init();
    rs->nlist = 0;
  }
  __syncthreads();

  if (tid >= kT3) {
    // ---- producer: one lane streams this block's rows through the ring, in job and pass order
    if (tid == kT3) {
      unsigned seq = 0;
      for (int job = blockIdx.x; job < njobs; job += gridDim.x) {
        const RowJob jb = decode_job(job, gamma);
        const TIn* row = split_row<TIn>(P, jb);
        const int passes = (jb.is_target ? P.t_top_p : P.d_top_p) > 0.0f ? 2 : 1;
        for (int ps = 0; ps < passes; ++ps) {
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
    }
    return;
  }

  // ---- workers
  int phase = 0;
  unsigned seq = 0;
  const int lane = tid & 31, warp = tid >> 5;
  float* hp = hist + tid;                           // this thread's histogram column
  const int bits0 = __float_as_int(kHTmin + kHMagic);  // bucket 0
#pragma unroll
  for (int b = 0; b < kHB; ++b) hp[b * kT3] = 0.0f;

#pragma unroll 1
  for (int job = blockIdx.x; job < njobs; job += gridDim.x) {
    const RowJob jb = decode_job(job, gamma);
    const TIn* row = split_row<TIn>(P, jb);
    const float temp = jb.is_target ? P.t_temp : P.d_temp;
    const float top_p = jb.is_target ? P.t_top_p : P.d_top_p;
    const bool nucleus = top_p > 0.0f;
    ListEntry* list = P.lists + (long long)blockIdx.x * kListCap;

    // the drafted token's logit is needed when the row closes: fetch it now (thread 0)
    long long tok = -1;
    float xtok = 0.0f;
    if (tid == 0 && jb.j < gamma) {
      tok = P.tokens[(long long)jb.b * P.tok_sb + jb.j];
      if (tok >= 0 && tok < V) xtok = scale_logit(to_f32(__ldcg(row + tok)), temp);
    }

    // exponent offset predicted by the previous launch (mean row maximum of this model's rows)
    const uint4 h = __ldg(reinterpret_cast<const uint4*>(&P.hints[jb.is_target]));
    float m0 = int(h.w) != 0 ? __uint_as_float(h.x) : 0.0f;

    RowRecord rec;
    rec.m0 = m0;
    rec.xstar = -INFINITY;
    rec.istar = 0x7fffffff;
    rec.sel = 0.0f;
    rec.dens = 0.0f;
    bool slow = false;
#ifdef DUODEC_TICKS
    long long tk0 = 0;
    if (P.stats && tid == 0) tk0 = clock64();
#define tick(k)                                            \
  if (P.stats && tid == 0) {                               \
    const long long t_ = clock64();                        \
    atomicAdd(&P.stats[k], int((t_ - tk0) >> 4));          \
    tk0 = t_;                                              \
  }
#else
#define tick(k)
#endif
    float Z = 0.0f, mxl = m0;

    if (!nucleus) {
      float z = 0.0f, mx = -INFINITY;
#pragma unroll 1
      for (int c = 0; c < nchunk; ++c) {
        RING_STAGE_BEGIN(c)
#define F_PLAIN(x, i) z += row_exp(x, m0), mx = fmaxf(mx, x);
        RING_EACH(F_PLAIN)
        RING_STAGE_END()
      }
      R8 r;
      r.v[0] = mx, r.v[1] = z;
      r.v[2] = r.v[3] = r.v[4] = r.v[5] = r.v[6] = r.v[7] = 0.0f;
      r = reduce3(r, sc, phase);
      Z = r.v[1];
      mxl = r.v[0] * kLog2e;
      slow = !(fabsf(mxl - m0) <= 32.0f) || !(Z < INFINITY) || !(Z > 0.0f);
    } else {
      // ---- pass 1: per-thread mass histogram over the exponent
#pragma unroll 1
      for (int c = 0; c < nchunk; ++c) {
        RING_STAGE_BEGIN(c)
#define F_HIST(x, i)                                  \
  {                                                   \
    const float t = fmaf(x, kLog2e, -m0);             \
    float* a = hp + (hist_bits(t) - bits0) * kT3;     \
    *a += ex2_approx(t);                              \
  }
        RING_EACH(F_HIST)
        RING_STAGE_END()
      }
      tick(24);
      bar3();
      // bucket sums: warp w owns buckets 4w .. 4w+3 (fixed order: deterministic)
#pragma unroll
      for (int q = 0; q < kHB / kW3; ++q) {
        const int b = warp * (kHB / kW3) + q;
        float s = 0.0f;
#pragma unroll
        for (int k = 0; k < kT3 / 32; ++k) {
          s += hist[b * kT3 + k * 32 + lane];
          hist[b * kT3 + k * 32 + lane] = 0.0f;  // ready for the next row
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        if (lane == 0) rs->hsum[b] = s;
      }
      bar3();
      // every warp scans the 32 bucket sums from the top (lane l holds bucket 31 - l): fixed shuffle
      // tree, the same values in every warp
      const float hb = rs->hsum[kHB - 1 - lane];
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
      slow = !(Z < INFINITY) || !(Z > 0.0f);
      const double target = double(top_p) * double(Z);
      int bstar = -1;
      float GB = 0.0f;
      if (!slow && target < double(Z)) {
        const unsigned crossed = __ballot_sync(0xffffffffu, double(incl) > target);
        if (crossed) {
          const int l = __ffs(crossed) - 1;  // first bucket from the top whose inclusive mass exceeds the target
          bstar = kHB - 1 - l;
          GB = __shfl_sync(0xffffffffu, incl - hb, l);
        }
        if (bstar <= 0) slow = true;  // the lower end bucket is unbounded below
      }
      tick(25);
      // ---- pass 2: the boundary bucket's entries (always consumed: the producer streams them)
      const int want = bstar > 0 && !slow ? bits0 + bstar : 0x7fffffff;
      int mycnt = 0;
      ListEntry* col = list + tid;  // this thread's column: entry j at list[j * 256 + tid]
#define F_PICK(x, i)                                                  \
  if (hist_bits(fmaf(x, kLog2e, -m0)) == want) {                      \
    if (mycnt < kRingPer) col[mycnt * kT3] = ListEntry{x, i};         \
    ++mycnt;                                                          \
  }
#pragma unroll 1
      for (int c = 0; c < nchunk; ++c) {  // always consumed: the producer streams the row twice
        RING_STAGE_BEGIN(c)
        RING_EACH(F_PICK)
        RING_STAGE_END()
      }
      tick(26);
      if (bstar > 0 && !slow) {
        const int cmax = reduce3_small(mycnt, OpMaxI(), sc, phase);
        if (cmax > kRingPer) {
          slow = true;
        } else {
          if (P.stats) {
            const int n = reduce3_small(mycnt, OpSumI(), sc, phase);
            if (tid == 0) {
              atomicAdd(&P.stats[0], 1);
              atomicAdd(&P.stats[16], 1);
              atomicAdd(&P.stats[21], n);
              atomicMax(&P.stats[22], n);
            }
          }
          list_select_body<kListCap>(list, 0, mycnt, 0.0f, 0.0f, GB, Z, target, m0, sc, phase, rec, P.stats);
        }
      } else {
        rec.Z = Z;  // top_p >= 1 keeps everything (or slow)
      }
    }
    if (!nucleus) rec.Z = Z;
    tick(27);
    rec.mxl = mxl;
    if (slow) {
      // general search: exact row maximum first, then the windowed search through L2
      bar3();
      Pass3 pr = rescan_row<TIn, kListCap>(row, V, temp, m0, INFINITY, INFINITY, list, &rs->nlist, sc, phase);
      m0 = pr.mx * kLog2e;
      pr = rescan_row<TIn, kListCap>(row, V, temp, m0, INFINITY, INFINITY, list, &rs->nlist, sc, phase);
      rec.m0 = m0;
      rec.Z = pr.z;
      rec.mxl = m0;
      float theta = pr.mx, dens = 0.0f;
      if (nucleus)
        resolve_row3<TIn, kListCap>(row, V, temp, top_p, m0, pr, INFINITY, INFINITY, 0.0f, list, &rs->nlist, sc,
                                    phase, rec, &theta, &dens, P.stats);
    }
    if (tid == 0) {
      if (tok >= 0 && tok < V) {
        const float e = kept(rec, xtok, int(tok)) ? row_exp(xtok, rec.m0) : 0.0f;
        rec.sel = __fdiv_rn(e, rec.Z);
      }
      store_record(&P.records[job], rec);
      rs->nlist = 0;
    }
    bar3();  // list counter and scratch are free for the next row
  }
}

}  // namespace duodec
