// Sample kernel of the fused verify: a block's three slices - the noise, the target row p_n and
// the draft row q_n - arrive in shared memory through TMA bulk copies: one L2 round trip for the whole
// slice instead of one per group of four entries.  The noise copy is issued before the grid dependency
// resolves (programmatic launch: it overlaps the scan kernel's tail); the row copies as soon as the
// accepted length n is known.  The row slices share their shared memory with the block's scratch and
// records, which are dead by then (one barrier in between).
#pragma once

#include "ring_kernel.cuh"

namespace duodec {

constexpr size_t sample_smem_bytes() { return size_t(3) * kSlice3 * 4; }
static_assert(sizeof(Scratch3) + (2 * 64 + 1) * sizeof(RowRecord) <= size_t(2) * kSlice3 * 2,
              "scratch + records must fit under the row slices (bf16 rows)");

template <class TIn>
__global__ void __launch_bounds__(kT3, 4) seq_sample_tma_kernel(const __grid_constant__ SplitParams P) {
  extern __shared__ __align__(128) unsigned char sm3[];
  __shared__ unsigned long long wkey[kW3], wbd[kW3];
  __shared__ float wmass[kW3];
  __shared__ unsigned long long mb[2];
  __shared__ int acc0[64];
  constexpr int kElt = int(sizeof(TIn));
  unsigned char* s_noise = sm3;                                   // [kSlice3] float
  unsigned char* s_p = sm3 + size_t(kSlice3) * 4;                 // [kSlice3] TIn
  unsigned char* s_q = s_p + size_t(kSlice3) * kElt;              // [kSlice3] TIn
  Scratch3& sc = *reinterpret_cast<Scratch3*>(s_p);               // until the row copies are issued
  RowRecord* srec = reinterpret_cast<RowRecord*>(s_p + sizeof(Scratch3));
  const int tid = threadIdx.x;
  const int slice = blockIdx.x, b = blockIdx.y;
  const int gamma = P.gamma, V = P.V;
  const int base = b * (2 * gamma + 1);
  const int s0 = slice * kSlice3;
  const int ne = V - s0 < kSlice3 ? V - s0 : kSlice3;  // > 0: the grid has ceil(V / kSlice3) slices

  if (tid == 0) {
    mbar_init(&mb[0], 1);
    mbar_init(&mb[1], 1);
    mbar_fence_init();
    if (b < P.B) {
      // the noise does not depend on the scan kernel (the caller wrote it before the scan kernel was launched)
      mbar_expect_tx(&mb[0], unsigned(ne) * 4u);
      bulk_g2s(s_noise, P.noise + (long long)b * P.noise_sb + s0, unsigned(ne) * 4u, &mb[0]);
    }
  }
  if (tid == 0) STAMP(P.stats, 12, true);
  float rj = 0.0f;  // the acceptance draw of position tid: an input as well
  if (b < P.B && tid < gamma) rj = __ldg(&P.r[(long long)b * P.r_sb + tid]);
  grid_dep_wait();
  grid_dep_launch();
  if (tid == 0) STAMP(P.stats, 13);

  if (b == P.B) {  // the extra block row: block 0 of it aggregates the next launch's prediction, off the critical path
    if (slice != 0) return;
    // prediction for the next launch, per row kind: mean exponent offset, mean / spread of the
    // nucleus boundary and mean list density over this launch's rows (fixed order: deterministic)
    int phase = 0;
    const int rows = P.B * (2 * gamma + 1);
    for (int kind = 0; kind < 2; ++kind) {
      R8 a;
      a.v[0] = -INFINITY;
      for (int i = 1; i < 8; ++i) a.v[i] = 0.0f;
      for (int job = tid; job < rows; job += kT3) {
        if (decode_job(job, gamma).is_target != kind) continue;
        const RowRecord rc = load_record(&P.records[job]);
        a.v[1] += 1.0f;
        a.v[2] += rc.mxl;
        if (rc.dens > 0.0f) {
          a.v[3] += 1.0f;
          a.v[4] += rc.xstar;
          a.v[5] += rc.xstar * rc.xstar;
          a.v[6] += rc.dens;
        }
      }
      a = reduce3(a, &sc, phase);
      if (tid == 0 && a.v[1] > 0.0f) {
        StreamHint hn;
        hn.m0 = a.v[2] / a.v[1];
        hn.theta = 0.0f, hn.dens = 0.0f, hn.hw = 0.0f;
        hn.valid = 1;
        hn.pad[0] = hn.pad[1] = hn.pad[2] = 0;
        if (a.v[3] > 0.0f) {
          const float mean = a.v[4] / a.v[3];
          const float var = fmaxf(a.v[5] / a.v[3] - mean * mean, 0.0f);
          hn.theta = mean;
          hn.dens = a.v[6] / a.v[3];
          // four standard deviations, at least ~200 entries, at most ~1800 entries per side
          hn.hw = fminf(fminf(fmaxf(4.0f * sqrtf(var), 200.0f / hn.dens), 1800.0f / hn.dens), 8.0f);
        }
        P.hints[kind] = hn;
      }
    }
    return;
  }
  // every record of the sequence is read ONCE (one L2 round trip) and kept in shared memory
  if (tid <= 2 * gamma) srec[tid] = load_record(&P.records[base + tid]);
  bar3();  // also orders the barrier initialisation before the waits below
  if (tid < gamma) {
    const RowRecord rq = srec[2 * tid], rp = srec[2 * tid + 1];
    const int accept = rj <= __fdiv_rn(rp.sel, rq.sel);  // raw IEEE division
    sc.seq_acc[tid] = accept;
    if (slice == 0) {
      const long long o = (long long)b * gamma + tid;
      P.accept_mask[o] = (unsigned char)accept;
      P.sel_p[o] = rp.sel;
      P.sel_q[o] = rq.sel;
    }
  }
  if (P.row0_rule && b != 0 && tid < gamma) {
    // the reference's batch rule (src/decoding_ops.py:144-147): sequence 0's accepted count positions every sequence
    const RowRecord q0 = load_record(&P.records[2 * tid]), p0 = load_record(&P.records[2 * tid + 1]);
    acc0[tid] = __ldg(&P.r[tid]) <= __fdiv_rn(p0.sel, q0.sel);
  }
  bar3();
  int n = 0;
  if (P.row0_rule && b != 0) {
    while (n < gamma && acc0[n] != 0) ++n;
  } else {
    while (n < gamma && sc.seq_acc[n] != 0) ++n;
  }
  const bool bonus = (n == gamma);
  const RowRecord rp = srec[bonus ? 2 * gamma : 2 * n + 1];
  const RowRecord rq = bonus ? rp : srec[2 * n];
  const float invZp = __frcp_rn(rp.Z), invZq = __frcp_rn(rq.Z);
  // scratch and records are in registers everywhere: their shared memory becomes the row slices
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic accesses above, async writes below
  bar3();
  if (tid == 0) {
    const TIn* prow = split_row<TIn>(P, RowJob{b, n, 1});
    mbar_expect_tx(&mb[1], unsigned(ne) * unsigned(kElt) * (bonus ? 1u : 2u));
    bulk_g2s(s_p, prow + s0, unsigned(ne) * unsigned(kElt), &mb[1]);
    if (!bonus) bulk_g2s(s_q, split_row<TIn>(P, RowJob{b, n, 0}) + s0, unsigned(ne) * unsigned(kElt), &mb[1]);
  }
  if (tid == 0) STAMP(P.stats, 14);
  mbar_wait(&mb[0], 0u);
  mbar_wait(&mb[1], 0u);
  if (tid == 0) STAMP(P.stats, 24);

  float mass = 0.0f, bkey = -1.0f, bdv = -INFINITY;
  int bidx = 0x7fffffff, bdi = 0x7fffffff;
#pragma unroll
  for (int u = 0; u < kSlice3 / (4 * kT3); ++u) {
    const int l0 = u * 4 * kT3 + 4 * tid, i0 = s0 + l0;
    if (i0 < V) {
      float4 xp = unpack4<TIn>(s_p + size_t(l0) * kElt);
      float4 xq = make_float4(0.f, 0.f, 0.f, 0.f);
      if (!bonus) xq = unpack4<TIn>(s_q + size_t(l0) * kElt);
      if (P.t_temp != 1.0f) {
        xp.x = __fdiv_rn(xp.x, P.t_temp), xp.y = __fdiv_rn(xp.y, P.t_temp);
        xp.z = __fdiv_rn(xp.z, P.t_temp), xp.w = __fdiv_rn(xp.w, P.t_temp);
      }
      if (!bonus && P.d_temp != 1.0f) {
        xq.x = __fdiv_rn(xq.x, P.d_temp), xq.y = __fdiv_rn(xq.y, P.d_temp);
        xq.z = __fdiv_rn(xq.z, P.d_temp), xq.w = __fdiv_rn(xq.w, P.d_temp);
      }
      float4 nv = *reinterpret_cast<const float4*>(s_noise + size_t(l0) * 4);
      // branch-free per entry: both exponentials are always taken, the keep rules are predicates.
      // A thread meets its indices in increasing order, so strict ">" keeps the lowest index of a tie.
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const float x = f4c(xp, k);
        const int i = i0 + k;
        const bool kp = (x > rp.xstar) | ((x == rp.xstar) & (i <= rp.istar));
        const float pe = row_exp(x, rp.m0) * invZp;
        float d = kp ? pe : 0.0f;
        if (!bonus) {
          const float y = f4c(xq, k);
          const bool kq = (y > rq.xstar) | ((y == rq.xstar) & (i <= rq.istar));
          const float qe = row_exp(y, rq.m0) * invZq;
          d -= kq ? qe : 0.0f;
        }
        const bool pos = d > 0.0f;
        mass += pos ? d : 0.0f;
        const float key = pos ? __fdividef(d, f4c(nv, k)) : -1.0f;
        const bool bk = key > bkey;
        bkey = bk ? key : bkey, bidx = bk ? i : bidx;
        const float dz = (fabsf(d) < INFINITY) ? d : 0.0f;  // NaN and +-inf count as 0
        const bool bd = dz > bdv;
        bdv = bd ? dz : bdv, bdi = bd ? i : bdi;
      }
    }
  }
  unsigned long long pk = bidx == 0x7fffffff ? 0ull
                                            : ((unsigned long long)__float_as_uint(bkey) << 32) |
                                                  (unsigned long long)(0xffffffffu - unsigned(bidx));
  unsigned long long pb =
      bdi == 0x7fffffff ? 0ull
                        : ((unsigned long long)okey(bdv) << 32) | (unsigned long long)(0xffffffffu - unsigned(bdi));
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    mass += __shfl_xor_sync(0xffffffffu, mass, o);
    const unsigned long long ok = __shfl_xor_sync(0xffffffffu, pk, o);
    const unsigned long long ob = __shfl_xor_sync(0xffffffffu, pb, o);
    pk = ok > pk ? ok : pk;
    pb = ob > pb ? ob : pb;
  }
  if ((tid & 31) == 0) wmass[tid >> 5] = mass, wkey[tid >> 5] = pk, wbd[tid >> 5] = pb;
  bar3();
  if (tid == 0) STAMP(P.stats, 15);
  if (tid == 0) {
    for (int w = 1; w < kW3; ++w) {
      mass += wmass[w];
      pk = wkey[w] > pk ? wkey[w] : pk;
      pb = wbd[w] > pb ? wbd[w] : pb;
    }
    if (mass > 0.0f) atomicAdd(&P.seq_mass[b], mass);
    if (pk != 0ull) atomicMax(&P.seq_key[b], pk);
    if (pb != 0ull) atomicMax(&P.seq_bd[b], pb);
    __threadfence();
    if (atomicAdd(&P.seq_arrive[b], 1) == int(gridDim.x) - 1) {
      __threadfence();
      const float m = atomicAdd(&P.seq_mass[b], 0.0f);
      const unsigned long long k = atomicMax(&P.seq_key[b], 0ull);
      const unsigned long long d = atomicMax(&P.seq_bd[b], 0ull);
      const unsigned long long win = (m > 0.0f && k != 0ull) ? k : d;
      long long tok = win == 0ull ? 0 : (long long)(0xffffffffu - unsigned(win & 0xffffffffull));
      P.accepted_len[b] = n;
      P.token[b] = tok;
      P.all_accepted[b] = (unsigned char)(n == gamma);
      P.seq_mass[b] = 0.0f;
      P.seq_key[b] = 0ull;
      P.seq_bd[b] = 0ull;
      P.seq_arrive[b] = 0;
    }
    STAMP(P.stats, 23);
  }
}

}  // namespace duodec
