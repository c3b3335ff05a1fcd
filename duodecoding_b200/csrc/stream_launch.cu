// Launcher of the persistent row-stream verification kernel (see stream_kernel.cuh).
#include <mutex>
#include <utility>
#include <vector>

#include "../../include/duodec.h"
#include "ring_kernel.cuh"
#include "resident_kernel.cuh"
#include "pair_kernel.cuh"
#include "capture_kernel.cuh"
#include "sample_kernel.cuh"
#include "stream_launch.h"

namespace duodec {

namespace {

size_t stream_smem_bytes() {
  return size_t(kSlots) * kSlotBytes + size_t(kColCap) * kCT * 8 + sizeof(StreamScratch) + 128;
}

// cudaFuncSetAttribute is per (kernel, device): remember the pairs already set up.
cudaError_t ensure_stream_smem(const void* kernel, size_t bytes) {
  static std::mutex mu;
  static std::vector<std::pair<const void*, int>> done;
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  std::lock_guard<std::mutex> lk(mu);
  for (const auto& d : done)
    if (d.first == kernel && d.second == dev) return cudaSuccess;
  e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(bytes));
  if (e != cudaSuccess) return e;
  done.emplace_back(kernel, dev);
  return cudaSuccess;
}

int stream_sm_count() {
  static std::mutex mu;
  static int cached[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  std::lock_guard<std::mutex> lk(mu);
  if (cached[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) return 148;
    cached[dev] = n;
  }
  return cached[dev];
}

// Launch with programmatic stream serialization: the kernel may become resident while the one
// before it in the stream drains; it orders itself with griddepcontrol.wait (grid_dep_wait()).
template <class... KArgs, class... Args>
cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                       Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kernel, std::forward<Args>(args)...);
}

}  // namespace

constexpr int kMaxGrid = 320;   // blocks of a persistent scan kernel (two per SM): one scratch list each
constexpr int kMaxSlices = 32;  // row_scan slices the workspace is sized for (V <= 262144)

size_t stream_workspace_bytes(int B, int gamma) {
  const size_t rows = size_t(B) * size_t(2 * gamma + 1);
  return 4 * sizeof(StreamHint) + rows * sizeof(RowRecord) + rows * kMaxSlices * sizeof(RowPartial) + rows * 8 +
         size_t(B) * 24 + (rows + 2 + kMaxGrid) * size_t(kListCap) * sizeof(ListEntry);
}

bool stream_try_launch(const StreamCall& c, int* rc) {
  const size_t elt = c.dtype == DUODEC_F32 ? 4 : 2;
  auto al16 = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; };
  const bool ok = c.Vd == c.Vt && c.gamma <= 64 && c.temperature > 0.0f && c.draft_temperature > 0.0f &&
                  c.top_k <= 0 && c.draft_top_k <= 0 && (size_t(c.Vt) * elt) % 16 == 0 && al16(c.target) &&
                  (c.gamma == 0 || al16(c.draft)) && (size_t(c.t_sb) * elt) % 16 == 0 &&
                  (size_t(c.t_ss) * elt) % 16 == 0 && (size_t(c.d_sb) * elt) % 16 == 0 &&
                  (size_t(c.d_ss) * elt) % 16 == 0 && al16(c.noise) && (c.noise_sb % 4) == 0;
  if (!ok) return false;
  const size_t rows = size_t(c.B) * size_t(2 * c.gamma + 1);
  const bool resident = c.path == 4 && c.B <= 65535;
  const bool capture = c.path == 6 && c.B <= 65535 && c.Vt >= cap::kMinV &&
                       cap::cap_smem_bytes((c.Vt + kStageElems - 1) / kStageElems) <= 232448;
  const bool pairp = (c.path == 5 || (c.path == 6 && !capture)) && c.B <= 65535;
  const bool ring = (c.path == 3 || (c.path == 1 && c.Vt > kMaxSlices * kSlice)) && c.B <= 65535;
  const bool split = c.path == 1 && c.Vt <= kMaxSlices * kSlice && rows <= 65535 && c.B <= 65535;
  if (ring || split || resident || pairp || capture) {
    SplitParams Q;
    Q.draft = c.draft;
    Q.target = c.target;
    Q.d_sb = c.d_sb;
    Q.d_ss = c.d_ss;
    Q.t_sb = c.t_sb;
    Q.t_ss = c.t_ss;
    Q.B = c.B;
    Q.gamma = c.gamma;
    Q.V = c.Vt;
    Q.S = (c.Vt + kSlice - 1) / kSlice;
    Q.tokens = reinterpret_cast<const long long*>(c.tokens);
    Q.tok_sb = c.tok_sb;
    Q.r = c.r;
    Q.r_sb = c.r_sb;
    Q.noise = c.noise;
    Q.noise_sb = c.noise_sb;
    Q.t_temp = c.temperature;
    Q.t_top_p = c.top_p;
    Q.d_temp = c.draft_temperature;
    Q.d_top_p = c.draft_top_p;
    Q.accept_mask = c.accept_mask;
    Q.sel_p = c.sel_target_p;
    Q.sel_q = c.sel_draft_p;
    Q.accepted_len = reinterpret_cast<long long*>(c.accepted_len);
    Q.token = reinterpret_cast<long long*>(c.token);
    Q.all_accepted = c.all_accepted;
    // behind the general kernel's [B] counters and [B, gamma+1] tokens, rounded up: hints and records
    // are read with 128-bit loads (B * (gamma + 2) odd would leave them 8-byte aligned)
    const size_t head = (size_t(c.B) * 8 + size_t(c.B) * size_t(c.gamma + 1) * 8 + 15) & ~size_t(15);
    char* w = reinterpret_cast<char*>(c.workspace) + head;
    Q.hints = reinterpret_cast<StreamHint*>(w);
    Q.hints_new = Q.hints + 2;
    w += 4 * sizeof(StreamHint);
    Q.records = reinterpret_cast<RowRecord*>(w);
    w += rows * sizeof(RowRecord);
    Q.partials = reinterpret_cast<RowPartial*>(w);
    w += rows * kMaxSlices * sizeof(RowPartial);
    Q.row_count = reinterpret_cast<int*>(w);
    w += rows * 4;
    Q.row_arrive = reinterpret_cast<int*>(w);
    w += rows * 4;
    Q.seq_key = reinterpret_cast<unsigned long long*>(w);
    w += size_t(c.B) * 8;
    Q.seq_bd = reinterpret_cast<unsigned long long*>(w);
    w += size_t(c.B) * 8;
    Q.seq_mass = reinterpret_cast<float*>(w);
    w += size_t(c.B) * 4;
    Q.seq_arrive = reinterpret_cast<int*>(w);
    w += size_t(c.B) * 4;
    Q.lists = reinterpret_cast<ListEntry*>(w);
    Q.stats = c.stats;
    Q.cap_rank = 0;
    Q.compact_hint = 0;
    const dim3 g1(unsigned(Q.S), unsigned(rows)), g2(unsigned((c.Vt + kSlice3 - 1) / kSlice3), unsigned(c.B) + 1u);  // + the block row that aggregates the hints
    if (capture) {
      const size_t smem = cap::cap_smem_bytes((c.Vt + kStageElems - 1) / kStageElems);
      const int sms = stream_sm_count();
      const int per_sm = smem <= 115712 ? 2 : 1;  // two blocks per SM while the bit maps leave room
      const int npairs = int((rows + 1) / 2);
      int grid = npairs < per_sm * sms ? npairs : per_sm * sms;
      if (grid > kMaxGrid) grid = kMaxGrid;
      // capture cut: the top f of a row by count, f = min(1/12, 2667 / V); a thread's 16 sample entries
      // all lie below the cut with probability (1 - f)^16, which is the rank among the 32 sorted maxima
      double f = 2667.0 / double(c.Vt);
      if (f > 1.0 / 12.0) f = 1.0 / 12.0;
      double pbelow = 1.0;
      for (int i = 0; i < 16; ++i) pbelow *= 1.0 - f;
      int rank = int(32.0 * pbelow);
      Q.cap_rank = rank < 0 ? 0 : (rank > 31 ? 31 : rank);
      Q.compact_hint = (c.top_p > 0.0f && c.top_p < 1.0f) ? 1 : 0;
      const dim3 g2c(unsigned((c.Vt + kSlice3 - 1) / kSlice3), unsigned(c.B));  // no hint row: nothing is predicted
      cudaError_t e;
      const size_t smem2 = sample_smem_bytes();
      if (c.dtype == DUODEC_F32) {
        e = ensure_stream_smem(reinterpret_cast<const void*>(cap::row_capture_kernel<float>), smem);
        if (e == cudaSuccess) e = ensure_stream_smem(reinterpret_cast<const void*>(seq_sample_tma_kernel<float>), smem2);
        if (e == cudaSuccess && (c.phases & 1))
          e = launch_pdl(cap::row_capture_kernel<float>, dim3(grid), dim3(cap::kPT), smem, c.stream, Q);
        if (e == cudaSuccess && (c.phases & 2))
          e = launch_pdl(seq_sample_tma_kernel<float>, g2c, dim3(kT3), smem2, c.stream, Q);
      } else {
        e = ensure_stream_smem(reinterpret_cast<const void*>(cap::row_capture_kernel<__nv_bfloat16>), smem);
        if (e == cudaSuccess)
          e = ensure_stream_smem(reinterpret_cast<const void*>(seq_sample_tma_kernel<__nv_bfloat16>), smem2);
        if (e == cudaSuccess && (c.phases & 1))
          e = launch_pdl(cap::row_capture_kernel<__nv_bfloat16>, dim3(grid), dim3(cap::kPT), smem, c.stream, Q);
        if (e == cudaSuccess && (c.phases & 2))
          e = launch_pdl(seq_sample_tma_kernel<__nv_bfloat16>, g2c, dim3(kT3), smem2, c.stream, Q);
      }
      *rc = int(e != cudaSuccess ? e : cudaGetLastError());
      return true;
    }
    if (pairp) {
      const size_t smem = pair::pair_smem_bytes();
      const int sms = stream_sm_count();
      const int npairs = int((rows + 1) / 2);
      const int grid = npairs < 2 * sms ? npairs : 2 * sms;  // two blocks per SM, two rows per block
      cudaError_t e;
      const size_t smem2 = sample_smem_bytes();
      if (c.dtype == DUODEC_F32) {
        e = ensure_stream_smem(reinterpret_cast<const void*>(pair::row_pair_kernel<float>), smem);
        if (e == cudaSuccess) e = ensure_stream_smem(reinterpret_cast<const void*>(seq_sample_tma_kernel<float>), smem2);
        if (e == cudaSuccess && (c.phases & 1))
          e = launch_pdl(pair::row_pair_kernel<float>, dim3(grid), dim3(pair::kPT), smem, c.stream, Q);
        if (e == cudaSuccess && (c.phases & 2)) e = launch_pdl(seq_sample_tma_kernel<float>, g2, dim3(kT3), smem2, c.stream, Q);
      } else {
        e = ensure_stream_smem(reinterpret_cast<const void*>(pair::row_pair_kernel<__nv_bfloat16>), smem);
        if (e == cudaSuccess)
          e = ensure_stream_smem(reinterpret_cast<const void*>(seq_sample_tma_kernel<__nv_bfloat16>), smem2);
        if (e == cudaSuccess && (c.phases & 1))
          e = launch_pdl(pair::row_pair_kernel<__nv_bfloat16>, dim3(grid), dim3(pair::kPT), smem, c.stream, Q);
        if (e == cudaSuccess && (c.phases & 2))
          e = launch_pdl(seq_sample_tma_kernel<__nv_bfloat16>, g2, dim3(kT3), smem2, c.stream, Q);
      }
      *rc = int(e != cudaSuccess ? e : cudaGetLastError());
      return true;
    }
    if (resident) {
      const size_t smem = resident::ring_smem_bytes();
      const int sms = stream_sm_count();
      const int grid = int(rows) < sms ? int(rows) : sms;  // one block per SM
      cudaError_t e;
      if (c.dtype == DUODEC_F32) {
        e = ensure_stream_smem(reinterpret_cast<const void*>(resident::row_resident_kernel<float>), smem);
        if (e == cudaSuccess) {
          if (c.phases & 1) resident::row_resident_kernel<float><<<grid, resident::kRT, smem, c.stream>>>(Q);
          if (c.phases & 2) launch_pdl(seq_sample_kernel<float>, g2, dim3(kT3), 0, c.stream, Q);
        }
      } else {
        e = ensure_stream_smem(reinterpret_cast<const void*>(resident::row_resident_kernel<__nv_bfloat16>), smem);
        if (e == cudaSuccess) {
          if (c.phases & 1) resident::row_resident_kernel<__nv_bfloat16><<<grid, resident::kRT, smem, c.stream>>>(Q);
          if (c.phases & 2) launch_pdl(seq_sample_kernel<__nv_bfloat16>, g2, dim3(kT3), 0, c.stream, Q);
        }
      }
      *rc = int(e != cudaSuccess ? e : cudaGetLastError());
      return true;
    }
    if (ring) {
      const size_t smem = ring_smem_bytes();
      const int sms = stream_sm_count();
      const int grid = int(rows) < 2 * sms ? int(rows) : 2 * sms;  // two blocks per SM
      cudaError_t e;
      if (c.dtype == DUODEC_F32) {
        e = ensure_stream_smem(reinterpret_cast<const void*>(row_ring_kernel<float>), smem);
        if (e == cudaSuccess) {
          if (c.phases & 1) row_ring_kernel<float><<<grid, kRT, smem, c.stream>>>(Q);
          if (c.phases & 2) launch_pdl(seq_sample_kernel<float>, g2, dim3(kT3), 0, c.stream, Q);
        }
      } else {
        e = ensure_stream_smem(reinterpret_cast<const void*>(row_ring_kernel<__nv_bfloat16>), smem);
        if (e == cudaSuccess) {
          if (c.phases & 1) row_ring_kernel<__nv_bfloat16><<<grid, kRT, smem, c.stream>>>(Q);
          if (c.phases & 2) launch_pdl(seq_sample_kernel<__nv_bfloat16>, g2, dim3(kT3), 0, c.stream, Q);
        }
      }
      *rc = int(e != cudaSuccess ? e : cudaGetLastError());
      return true;
    }
    if (c.dtype == DUODEC_F32) {
      if (c.phases & 1) row_scan_kernel<float><<<g1, kT3, 0, c.stream>>>(Q);
      if (c.phases & 2) launch_pdl(seq_sample_kernel<float>, g2, dim3(kT3), 0, c.stream, Q);
    } else {
      if (c.phases & 1) row_scan_kernel<__nv_bfloat16><<<g1, kT3, 0, c.stream>>>(Q);
      if (c.phases & 2) launch_pdl(seq_sample_kernel<__nv_bfloat16>, g2, dim3(kT3), 0, c.stream, Q);
    }
    *rc = int(cudaGetLastError());
    return true;
  }
  StreamParams S;
  S.draft = c.draft;
  S.target = c.target;
  S.d_sb = c.d_sb;
  S.d_ss = c.d_ss;
  S.t_sb = c.t_sb;
  S.t_ss = c.t_ss;
  S.B = c.B;
  S.gamma = c.gamma;
  S.V = c.Vt;
  S.tokens = reinterpret_cast<const long long*>(c.tokens);
  S.tok_sb = c.tok_sb;
  S.r = c.r;
  S.r_sb = c.r_sb;
  S.noise = c.noise;
  S.noise_sb = c.noise_sb;
  S.t_temp = c.temperature;
  S.t_top_p = c.top_p;
  S.d_temp = c.draft_temperature;
  S.d_top_p = c.draft_top_p;
  S.accept_mask = c.accept_mask;
  S.sel_p = c.sel_target_p;
  S.sel_q = c.sel_draft_p;
  S.accepted_len = reinterpret_cast<long long*>(c.accepted_len);
  S.token = reinterpret_cast<long long*>(c.token);
  S.all_accepted = c.all_accepted;
  char* w = reinterpret_cast<char*>(c.workspace);
  S.counters = reinterpret_cast<int*>(w);
  w += size_t(c.B) * 8;
  w += size_t(c.B) * size_t(c.gamma + 1) * 8;  // (the general kernel's position tokens)
  S.hints = reinterpret_cast<StreamHint*>(w);
  w += 4 * sizeof(StreamHint);  // (this launch's and the next launch's predictions of the split kernels)
  S.records = reinterpret_cast<RowRecord*>(w);
  S.stats = c.stats;
  const size_t smem = stream_smem_bytes();
  const int njobs = c.B * (2 * c.gamma + 1);
  const int sms = stream_sm_count();
  const int grid = njobs < 2 * sms ? njobs : 2 * sms;
  cudaError_t e;
  if (c.dtype == DUODEC_F32) {
    e = ensure_stream_smem(reinterpret_cast<const void*>(verify_stream_kernel<float>), smem);
    if (e == cudaSuccess) verify_stream_kernel<float><<<grid, kST, smem, c.stream>>>(S);
  } else {
    e = ensure_stream_smem(reinterpret_cast<const void*>(verify_stream_kernel<__nv_bfloat16>), smem);
    if (e == cudaSuccess) verify_stream_kernel<__nv_bfloat16><<<grid, kST, smem, c.stream>>>(S);
  }
  *rc = int(e != cudaSuccess ? e : cudaGetLastError());
  return true;
}

}  // namespace duodec
