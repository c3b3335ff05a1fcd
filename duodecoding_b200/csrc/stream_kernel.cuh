// Shared device primitives of the fused-verify fast paths: the 32-byte row record and its keep rule,
// row-job numbering, the exponential (FFMA + MUFU.EX2), 128-bit row loads.
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

}  // namespace duodec
