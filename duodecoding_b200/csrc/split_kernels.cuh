// Row-level building blocks shared by the fused-verify fast paths: the parameter block, the list
// entry, block reductions, and the GENERAL nucleus search of one row by one 256-thread block through
// L2 (rescan_row -> resolve_row3 -> list_select_body: windowed passes with plain loads, exact 38-bit
// fixed-point level histogram, all-pairs order of the last candidates, tie bisection).  The scan
// kernels call it for rows their fast scheme cannot finish (flat rows, large tie groups, exponent
// offsets far from the row maximum).  Reference semantics: see stream_kernel.cuh.
#pragma once

#include "stream_kernel.cuh"

namespace duodec {

constexpr int kT3 = 256;        // threads per block
constexpr int kW3 = kT3 / 32;
constexpr int kSlice = 8192;    // row entries per row_scan block: 8 float4 per thread
constexpr int kListCap = 4096;  // window entries a row's global list can hold
constexpr int kCol3 = 8;        // window entries a thread stages in its own shared-memory column
constexpr int kOvf3 = 512;      // staged overflow entries per block
constexpr int kSlice3 = 4096;   // row entries per seq_sample block: 4 float4 per thread and array

// barrier of the 256 worker threads (named, so that a kernel may run a producer warp beside them)
__device__ __forceinline__ void bar3() { asm volatile("bar.sync 1, 256;" ::: "memory"); }

struct RowPartial {
  float mx, z, ghi, gwin;
};
// Programmatic dependent launch (stream_launch.cu: launch_pdl).  wait: everything the previous kernel
// in the stream wrote is visible afterwards (no-op for a plain launch); launch: the next kernel in
// the stream may start becoming resident.
__device__ __forceinline__ void grid_dep_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void grid_dep_launch() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

#ifdef DUODEC_TIMELINE
// tuning build: global-timer stamps (ns, low 31 bits) into the stats buffer; slot k keeps the maximum,
// "first" slots keep 0x7fffffff - t so that the maximum is the earliest
__device__ __forceinline__ void stamp(int* stats, int k, bool first = false) {
  if (!stats) return;
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  const int v = int(t & 0x7fffffffull);
  atomicMax(&stats[k], first ? 0x7fffffff - v : v);
}
#define STAMP(stats, k, ...) stamp(stats, k, ##__VA_ARGS__)
// cycle stamps relative to t0 (same SM): slot k keeps the maximum
#define CSTAMP(stats, k, t0) do { if (stats) atomicMax(&(stats)[k], int(clock64() - (t0))); } while (0)
#else
#define STAMP(stats, k, ...)
#define CSTAMP(stats, k, t0)
#endif

struct ListEntry {
  float x;
  int idx;
};

__device__ __forceinline__ float ld_list_x(const ListEntry* e) { return *reinterpret_cast<const volatile float*>(&e->x); }
__device__ __forceinline__ int ld_list_i(const ListEntry* e) { return *reinterpret_cast<const volatile int*>(&e->idx); }

struct SplitParams {
  const void* draft;
  const void* target;
  long long d_sb, d_ss, t_sb, t_ss;
  int B, gamma, V, S;
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
  // workspace (zero-filled once; every counter below is restored to zero by the kernels)
  StreamHint* hints;      // [2] prediction read by this launch
  StreamHint* hints_new;  // [2] prediction written by this launch (copied over by seq_sample)
  RowRecord* records;     // [rows]
  RowPartial* partials;   // [rows, S]
  int* row_count;         // [rows] entries in the row's list
  int* row_arrive;        // [rows] slices finished
  ListEntry* lists;       // [rows, kListCap]
  unsigned long long* seq_key;  // [B] best (d/noise, index)
  unsigned long long* seq_bd;   // [B] best (p - q, index) for the zero-mass fallback
  float* seq_mass;              // [B]
  int* seq_arrive;              // [B]
  int* stats;
  int row0_rule;     // sample kernel: sequence 0's accepted count positions every sequence (reference batch rule)
};

struct __align__(16) Scratch3 {
  R8 red[2][kW3];
  float cand_e[kCandMax];
  float cand_x[kCandMax];
  int cand_i[kCandMax];
  int ncand, novf, base, last;
  float res_x;
  int res_i;
  float res_Z;
  int res_ok;
  float hint_m0, hint_th, hint_dens;
  int hint_ok;
  int seq_acc[64];
  int wsum[kW3];
  unsigned long long hist[256];      // level histogram: (count << 50) | fixed-point mass
  unsigned long long cand_f[kCandMax];
  unsigned long long wscan[kW3];
  unsigned long long gb_new, lastkey;
  unsigned kmin, kmax;
  int bstar, bcount;
};

// all-reduce of an R8 (v[0] max, v[1..7] sum) over the 256 threads; one __syncthreads, ping-pong scratch
__device__ __forceinline__ R8 reduce3(R8 a, Scratch3* sc, int& phase) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    a.v[0] = fmaxf(a.v[0], __shfl_xor_sync(0xffffffffu, a.v[0], o));
#pragma unroll
    for (int i = 1; i < 8; ++i) a.v[i] += __shfl_xor_sync(0xffffffffu, a.v[i], o);
  }
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  if (lane == 0) sc->red[phase][w] = a;
  bar3();
  R8 x = sc->red[phase][lane & (kW3 - 1)];
  phase ^= 1;
#pragma unroll
  for (int o = kW3 / 2; o > 0; o >>= 1) {
    x.v[0] = fmaxf(x.v[0], __shfl_xor_sync(0xffffffffu, x.v[0], o));
#pragma unroll
    for (int i = 1; i < 8; ++i) x.v[i] += __shfl_xor_sync(0xffffffffu, x.v[i], o);
  }
  return x;
}

template <class T, class Op>
__device__ __forceinline__ T reduce3_small(T v, Op op, Scratch3* sc, int& phase) {
  static_assert(sizeof(T) <= 16 && sizeof(T) % 4 == 0, "");
  v = warp_allreduce(v, op);
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  if (lane == 0) *reinterpret_cast<T*>(&sc->red[phase][w]) = v;
  bar3();
  T x = *reinterpret_cast<T*>(&sc->red[phase][lane & (kW3 - 1)]);
  phase ^= 1;
#pragma unroll
  for (int o = kW3 / 2; o > 0; o >>= 1) x = op(x, shfl_xor_any(x, o));
  return x;
}

struct Pass3 {
  float mx, z, ghi, gwin;
  int n;  // entries appended to the row's list (> kListCap: the list is incomplete)
};

// One more pass over a whole row by a single block, through L2: statistics for the window
// [lo, hi) and its entries appended to the (emptied) global list.
template <class TIn, int kCap>
__device__ __noinline__ Pass3 rescan_row(const TIn* row, int V, float temp, float m0, float lo, float hi,
                                         ListEntry* list, int* count, Scratch3* sc, int& phase) {
  const int tid = threadIdx.x;
  if (tid == 0) *count = 0;
  bar3();
  float mx = -INFINITY, z = 0.0f, ghi = 0.0f, gwin = 0.0f;
  for (int c0 = 0; c0 < V; c0 += 4 * 4 * kT3) {
    float4 v[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int i0 = c0 + u * 4 * kT3 + 4 * tid;
      if (i0 < V) v[u] = load4_scaled<TIn>(row + i0, temp);
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int i0 = c0 + u * 4 * kT3 + 4 * tid;
      if (i0 < V) {
        mx = fmaxf(mx, fmaxf(fmaxf(v[u].x, v[u].y), fmaxf(v[u].z, v[u].w)));
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const float x = f4c(v[u], k);
          const float e = row_exp(x, m0);
          z += e;
          if (x >= hi) {
            ghi += e;
          } else if (x >= lo) {
            gwin += e;
            const int pos = atomicAdd(count, 1);
            if (pos < kCap) list[pos] = ListEntry{x, i0 + k};
          }
        }
      }
    }
  }
  R8 r;
  r.v[0] = mx, r.v[1] = z, r.v[2] = ghi, r.v[3] = gwin;
  r.v[4] = r.v[5] = r.v[6] = r.v[7] = 0.0f;
  r = reduce3(r, sc, phase);  // its barrier orders every append before the reads below
  Pass3 p;
  p.mx = r.v[0], p.z = r.v[1], p.ghi = r.v[2], p.gwin = r.v[3];
  p.n = *reinterpret_cast<volatile int*>(count);
  return p;
}

// Nucleus boundary among the n (<= kCap) entries of `list`, all of them inside [A, B), given
// GB = mass of the row's entries >= B and the row's total mass Z:  the LAST entry, in the
// reference's order (scaled logit descending, lowest index first), that is still kept, i.e. the
// first whose inclusive cumulative mass exceeds target = top_p * Z.  Fills rec.{Z, xstar, istar}.
template <int kCap>
__device__ __forceinline__ void list_select_body(const ListEntry* list, int nlist, int mycnt, float A, float B,
                                                 float GB, float Zrow, double target, float m0, Scratch3* sc,
                                                 int& phase, RowRecord& rec, int* stats,
                                                 const ListEntry* ovf = nullptr, int novf = 0) {
  const int tid = threadIdx.x;
  struct { float z; int n; } pr{Zrow, nlist};
  const float lo = A;
  // ---- boundary search on the list entries (16 per thread, in registers), in exact integer
  // arithmetic: every entry's mass as a 38-bit fixed-point fraction of 2^ceil(log2 Z), so that
  // sums do not depend on the order of the shared-memory atomics.  A level spreads the entries
  // of [A, B) over 256 value buckets (mass and count packed in one 64-bit word), a suffix scan
  // finds the bucket in which the kept mass crosses top_p * Z; with <= kCandMax entries in it
  // the all-pairs finish below orders them exactly like the reference's sort.
  constexpr int kCol = kCap / kT3;  // entries of this thread's list column
  constexpr int kPer = kCol + 2;    // + two entries of the overflow list (ovf[k * 256 + tid], k < 2)
  auto entry = [&](int j) { return j < kCol ? &list[j * kT3 + tid] : &ovf[(j - kCol) * kT3 + tid]; };
  constexpr unsigned long long kMassMask = (1ull << 50) - 1ull;
  const int ez = int(__float_as_uint(pr.z) >> 23) - 126;               // Z = f * 2^ez, f in [0.5, 1)
  const float S = __uint_as_float(unsigned(127 + 38 - ez) << 23);      // 2^(38 - ez)
  const unsigned long long tgt = (unsigned long long)(target * double(S));  // floor
  unsigned long long GBf = (unsigned long long)(double(GB) * double(S));
  if (GBf > tgt) GBf = tgt;  // (a float GB a few ulps above the target: the top non-empty bucket is the crossing one)
  // entry j of this thread sits at list[j * 256 + tid]; it exists if j < mycnt (per-thread columns,
  // mycnt >= 0) or j * 256 + tid < nlist (one shared list, mycnt < 0)
  float cx[kPer];
  auto fx = [&](float x) { return __float2ull_rn(row_exp(x, m0) * S); };  // fixed-point mass of an entry
  float emx = -INFINITY, engmn = -INFINITY;
#pragma unroll
  for (int j = 0; j < kPer; ++j) {
    const int p = j * kT3 + tid;
    const bool have = j < kCol ? (mycnt >= 0 ? j < mycnt : p < pr.n) : ((j - kCol) * kT3 + tid < novf);
    float x = -INFINITY;
    if (have) {
      x = ld_list_x(entry(j));
      emx = fmaxf(emx, x), engmn = fmaxf(engmn, -x);
    }
    cx[j] = x;
  }
  if (mycnt >= 0) {  // the bracket is the value range of the entries
    const F4 mm = reduce3_small(F4{emx, engmn, 0.0f, 0.0f}, OpMaxF4(), sc, phase);
    A = -mm.b, B = nextafterf(mm.a, INFINITY);
  }
  if (A < lo) A = lo;  // every list entry is inside the window
  bool solved = false, tie_mode = false;
  int ncand_stat = 0;
#pragma unroll 1
  for (int level = 0; level < 8 && !solved && !tie_mode; ++level) {
    sc->hist[tid] = 0ull;
    if (tid == 0) {
      sc->bstar = -1;
      sc->kmin = 0xffffffffu;
      sc->kmax = 0u;
      sc->ncand = 0;
      sc->res_ok = 0;
      sc->lastkey = 0xffffffffffffffffull;
    }
    bar3();
    const float bsc = 256.0f / (B - A);
#pragma unroll
    for (int j = 0; j < kPer; ++j) {
      if (cx[j] >= A && cx[j] < B) {
        int bk = int((cx[j] - A) * bsc);
        bk = bk < 0 ? 0 : (bk > 255 ? 255 : bk);
        atomicAdd(&sc->hist[bk], fx(cx[j]) | (1ull << 50));
      }
    }
    bar3();
    // suffix scan over buckets 255..0 (thread t owns bucket 255 - t)
    const unsigned long long hv = sc->hist[255 - tid];
    unsigned long long inc = hv;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const unsigned long long o = __shfl_up_sync(0xffffffffu, inc, d);
      if ((tid & 31) >= d) inc += o;
    }
    if ((tid & 31) == 31) sc->wscan[tid >> 5] = inc;
    bar3();
#pragma unroll
    for (int w = 0; w < kW3; ++w)
      if (w < (tid >> 5)) inc += sc->wscan[w];
    const unsigned long long m_incl = GBf + (inc & kMassMask), m_excl = m_incl - (hv & kMassMask);
    if ((hv >> 50) != 0ull && m_incl > tgt && m_excl <= tgt) {
      sc->bstar = 255 - tid;
      sc->gb_new = m_excl;
      sc->bcount = int(hv >> 50);
    }
    if (tid == kT3 - 1 && m_incl <= tgt) sc->bstar = -2;  // rounding only: the whole bracket is kept
    bar3();
    const int bstar = sc->bstar;
    if (bstar == -2 || bstar == -1) {
      // keep everything down to the last entry of the bracket in the reference's order
      unsigned long long kk = 0xffffffffffffffffull;
#pragma unroll
      for (int j = 0; j < kPer; ++j) {
        if (cx[j] >= A && cx[j] < B) {
          const unsigned idx = unsigned(ld_list_i(entry(j)));
          const unsigned long long k2 = ((unsigned long long)okey(cx[j]) << 32) | (0xffffffffu - idx);
          kk = k2 < kk ? k2 : kk;
        }
      }
      if (kk != 0xffffffffffffffffull) atomicMin(&sc->lastkey, kk);
      bar3();
      const unsigned long long lk = sc->lastkey;
      if (tid == kT3 - 1) sc->gb_new = m_incl;  // the bracket's total is this thread's inclusive sum
      bar3();
      if (lk != 0xffffffffffffffffull) {
        rec.Z = float(double(sc->gb_new) / double(S));
        rec.xstar = okey_inv(unsigned(lk >> 32));
        rec.istar = int(0xffffffffu - unsigned(lk & 0xffffffffull));
        solved = true;
      } else {
        tie_mode = true;  // empty bracket: let the generic bisection decide
      }
      break;
    }
    const int bcount = sc->bcount;
    const unsigned long long gb_new = sc->gb_new;
    if (bcount <= kCandMax) {
      // ---- the bucket's entries, ordered exactly like the reference's sort
#pragma unroll
      for (int j = 0; j < kPer; ++j) {
        if (cx[j] >= A && cx[j] < B) {
          int bk = int((cx[j] - A) * bsc);
          bk = bk < 0 ? 0 : (bk > 255 ? 255 : bk);
          if (bk == bstar) {
            const int pos = atomicAdd(&sc->ncand, 1);
            sc->cand_x[pos] = cx[j];
            sc->cand_f[pos] = fx(cx[j]);
            sc->cand_i[pos] = ld_list_i(entry(j));
          }
        }
      }
      bar3();
      const int n = sc->ncand;
      ncand_stat = n;
      int Pn = 16;
      while (Pn < n) Pn <<= 1;
      const int parts = kT3 / Pn;  // 1..16 threads share one candidate (same warp)
      const int t = tid / parts, part = tid - t * parts;
      unsigned long long cum = 0ull, fx = 0ull;
      float xx = 0.0f;
      int ii = 0;
      if (t < n) {
        fx = sc->cand_f[t], xx = sc->cand_x[t], ii = sc->cand_i[t];
        for (int j = part; j < n; j += parts) {
          const float xj = sc->cand_x[j];
          const int ij = sc->cand_i[j];
          if (xj > xx || (xj == xx && ij <= ii)) cum += sc->cand_f[j];
        }
      }
      for (int o = parts >> 1; o > 0; o >>= 1) cum += __shfl_xor_sync(0xffffffffu, cum, o);
      const unsigned long long incl = gb_new + cum;
      if (t < n && part == 0 && incl > tgt && incl - fx <= tgt) {
        sc->res_x = xx;
        sc->res_i = ii;
        sc->gb_new = incl;
        sc->res_ok = 1;
      }
      bar3();
      if (sc->res_ok) {
        solved = true;
        rec.Z = float(double(sc->gb_new) / double(S));
        rec.xstar = sc->res_x;
        rec.istar = sc->res_i;
      } else {
        tie_mode = true;  // cannot happen with exact sums; stay safe
      }
      break;
    }
    // crowded bucket: shrink the bracket to the bucket's value range and go one level down
#pragma unroll
    for (int j = 0; j < kPer; ++j) {
      if (cx[j] >= A && cx[j] < B) {
        int bk = int((cx[j] - A) * bsc);
        bk = bk < 0 ? 0 : (bk > 255 ? 255 : bk);
        if (bk == bstar) {
          atomicMin(&sc->kmin, okey(cx[j]));
          atomicMax(&sc->kmax, okey(cx[j]));
        }
      }
    }
    bar3();
    const unsigned kmin = sc->kmin, kmax = sc->kmax;
    A = okey_inv(kmin);
    B = nextafterf(okey_inv(kmax), INFINITY);
    GBf = gb_new;
    if (kmin == kmax) tie_mode = true;  // more equal values than the finish can order
    bar3();
  }
  if (stats && tid == 0) atomicAdd(&stats[18], ncand_stat);
  if (!solved) {
    // ties / crowded bracket: bisection on K = (okey(x) << 32) | ~index over the bracket's entries
    if (stats && tid == 0) atomicAdd(&stats[17], 1);
    unsigned long long klo = (unsigned long long)okey(A) << 32, khi = (unsigned long long)okey(B) << 32;
    auto F = [&](unsigned long long K) {
      unsigned long long s = 0ull;
#pragma unroll
      for (int j = 0; j < kPer; ++j) {
        if (cx[j] >= A && cx[j] < B) {
          const unsigned idx = unsigned(ld_list_i(entry(j)));
          const unsigned long long kk = ((unsigned long long)okey(cx[j]) << 32) | (0xffffffffu - idx);
          if (kk >= K) s += fx(cx[j]);
        }
      }
      return GBf + reduce3_small(s, OpSumU64(), sc, phase);
    };
#pragma unroll 1
    while (khi - klo > 1ull) {
      const unsigned long long mid = klo + ((khi - klo) >> 1);
      if (F(mid) > tgt) klo = mid;
      else khi = mid;
    }
    rec.Z = float(double(F(klo)) / double(S));
    rec.xstar = okey_inv(unsigned(klo >> 32));
    rec.istar = int(0xffffffffu - unsigned(klo & 0xffffffffull));
  }
}

template <int kCap>
__device__ __noinline__ void list_select(const ListEntry* list, int nlist, int mycnt, float A, float B, float GB,
                                         float Zrow, double target, float m0, Scratch3* sc, int& phase,
                                         RowRecord& rec, int* stats) {
  list_select_body<kCap>(list, nlist, mycnt, A, B, GB, Zrow, target, m0, sc, phase, rec, stats);
}

// Nucleus boundary of one row by the block that finished it; `pr` describes the pass over
// the window [lo, hi) whose entries sit in `list`.  Fills rec.{Z, xstar, istar}.
template <class TIn, int kCap>
__device__ __noinline__ void resolve_row3(const TIn* row, int V, float temp, float top_p, float m0, Pass3 pr,
                                          float lo, float hi, float dens, ListEntry* list, int* count,
                                          Scratch3* sc, int& phase, RowRecord& rec, float* theta_out,
                                          float* dens_out, int* stats) {
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
      list_ok = pr.n <= kCap;
    } else if (lo < B) {
      B = lo, GB = pr.ghi + pr.gwin;
    }
    if (pr.n > 0 && hi - lo > 0.0f && hi - lo < 1e30f) dens = float(pr.n) / (hi - lo);
    if (list_ok) break;
    if (nextafterf(A, INFINITY) >= B || npass >= 48) break;  // one float value left: tie mode below
    if (A == -INFINITY) {
      float step = 4.0f * float(1 << (npass < 6 ? npass : 6));
      if (dens > 0.0f) step = fminf(step, fmaxf((0.5f * float(kCap)) / dens, 1e-3f));
      lo = B - step, hi = B;
    } else {
      const float W = B - A;
      const float est = dens > 0.0f ? dens * W : __fdividef(GA - GB, row_exp(A, m0));
      if (est <= (0.5f * float(kCap))) {
        lo = A, hi = B;
      } else {
        const float c = A + W * __fdividef(GA - float(target), GA - GB);
        const float h = 0.5f * W * __fdividef((0.5f * float(kCap)), est);
        lo = fmaxf(A, c - h), hi = fminf(B, c + h);
        if (!(hi > lo)) lo = A, hi = fminf(B, nextafterf(A + 2.0f * h, INFINITY));
        if (!(hi > lo)) hi = B;
      }
    }
    pr = rescan_row<TIn, kCap>(row, V, temp, m0, lo, hi, list, count, sc, phase);
    ++npass;
  }
  *dens_out = dens;
  if (stats && tid == 0) {
    atomicAdd(&stats[npass < 15 ? npass : 15], 1);  // [0..15] rows by number of extra (L2) passes
    atomicAdd(&stats[16], 1);
    atomicAdd(&stats[21], pr.n);
    atomicMax(&stats[22], pr.n);
  }

  if (list_ok) {
    if (A < lo) A = lo;  // every list entry is inside the window
    list_select<kCap>(list, pr.n, -1, A, B, GB, pr.z, target, m0, sc, phase, rec, stats);
  } else {
    // one float value A holds more entries than the list can: count them in index order
    if (stats && tid == 0) atomicAdd(&stats[19], 1);
    const float eA = row_exp(A, m0);
    int ilo = -1, ihi = V - 1;  // smallest I with GB + eA * #{A <= x < B, idx <= I} > target
    auto C = [&](int I) {
      float cnt = 0.0f;
      for (int i0 = 4 * tid; i0 < V; i0 += 4 * kT3) {
        float4 v = load4_scaled<TIn>(row + i0, temp);
#pragma unroll
        for (int k = 0; k < 4; ++k)
          if (f4c(v, k) >= A && f4c(v, k) < B && i0 + k <= I) cnt += 1.0f;
      }
      return reduce3_small(cnt, OpSumF(), sc, phase);
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
  bar3();
}

template <class TIn>
__device__ __forceinline__ const TIn* split_row(const SplitParams& P, RowJob jb) {
  return jb.is_target
             ? reinterpret_cast<const TIn*>(P.target) + (long long)jb.b * P.t_sb + (long long)jb.j * P.t_ss
             : reinterpret_cast<const TIn*>(P.draft) + (long long)jb.b * P.d_sb + (long long)jb.j * P.d_ss;
}

}  // namespace duodec
