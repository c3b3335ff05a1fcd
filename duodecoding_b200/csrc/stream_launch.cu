// Launcher of the fused-verify fast paths: the row-pair scan kernel followed by the sample kernel under
// programmatic dependent launch (T > 0), or the greedy arg-max kernels (T == 0).
#include <mutex>
#include <utility>
#include <vector>

#include "../../include/duodec.h"
#include "pair_kernel.cuh"
#include "greedy_kernels.cuh"
#include "sample_kernel.cuh"
#include "stream_launch.h"

namespace duodec {

namespace {

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

constexpr size_t kPairSmemMax = 232448;  // the scan kernel's shared memory depends on V (bit maps): opt in to the limit once
constexpr int kMaxGrid = 320;   // blocks of the persistent scan kernel (two per SM): one scratch list each

// hints | row records | per-sequence cells of the sample kernel | one scratch list per scan block (general search) |
// the greedy path's arg-max cells
size_t stream_workspace_bytes(int B, int gamma) {
  const size_t rows = size_t(B) * size_t(2 * gamma + 1);
  return 4 * sizeof(StreamHint) + rows * sizeof(RowRecord) + size_t(B) * 24 + 128 +
         size_t(kMaxGrid) * size_t(kListCap) * sizeof(ListEntry) + rows * 8;  // + the greedy path's arg-max cells
}

bool stream_try_launch(const StreamCall& c, int* rc) {
  const size_t elt = c.dtype == DUODEC_F32 ? 4 : 2;
  auto al16 = [](const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; };
  const size_t nrows = size_t(c.B) * size_t(2 * c.gamma + 1);
  const bool rows_ok = c.Vd == c.Vt && (size_t(c.Vt) * elt) % 16 == 0 && al16(c.target) && (c.gamma == 0 || al16(c.draft)) &&
                       (size_t(c.t_sb) * elt) % 16 == 0 && (size_t(c.t_ss) * elt) % 16 == 0 &&
                       (size_t(c.d_sb) * elt) % 16 == 0 && (size_t(c.d_ss) * elt) % 16 == 0;
  if (c.temperature == 0.0f && c.draft_temperature == 0.0f && rows_ok && nrows <= 65535 && c.path != 1) {
    // greedy verification: one arg-max pass over the rows, then one thread per sequence (greedy_kernels.cuh)
    greedy::GreedyParams G;
    G.draft = c.draft, G.target = c.target;
    G.d_sb = c.d_sb, G.d_ss = c.d_ss, G.t_sb = c.t_sb, G.t_ss = c.t_ss;
    G.B = c.B, G.gamma = c.gamma, G.V = c.Vt, G.row0_rule = c.row0_rule;
    G.tokens = reinterpret_cast<const long long*>(c.tokens), G.tok_sb = c.tok_sb;
    G.r = c.r, G.r_sb = c.r_sb;
    G.accept_mask = c.accept_mask, G.sel_p = c.sel_target_p, G.sel_q = c.sel_draft_p;
    G.accepted_len = reinterpret_cast<long long*>(c.accepted_len), G.token = reinterpret_cast<long long*>(c.token);
    G.all_accepted = c.all_accepted;
    const size_t head = (size_t(c.B) * 8 + size_t(c.B) * size_t(c.gamma + 1) * 8 + 15) & ~size_t(15);
    // the cells sit behind everything the sampled-temperature paths use: zero since the workspace was created
    G.cells = reinterpret_cast<unsigned long long*>(reinterpret_cast<char*>(c.workspace) + head +
                                                    stream_workspace_bytes(c.B, c.gamma) - nrows * 8);
    const dim3 grid(unsigned((c.Vt + greedy::kGSlice - 1) / greedy::kGSlice), unsigned(nrows));
    if (c.dtype == DUODEC_F32) greedy::greedy_argmax_kernel<float><<<grid, greedy::kGT, 0, c.stream>>>(G);
    else greedy::greedy_argmax_kernel<__nv_bfloat16><<<grid, greedy::kGT, 0, c.stream>>>(G);
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) e = launch_pdl(greedy::greedy_close_kernel, dim3(1), dim3(256), 0, c.stream, G);
    *rc = int(e);
    return true;
  }
  const bool ok = c.Vd == c.Vt && c.gamma <= 64 && c.temperature > 0.0f && c.draft_temperature > 0.0f &&
                  c.top_k <= 0 && c.draft_top_k <= 0 && (size_t(c.Vt) * elt) % 16 == 0 && al16(c.target) &&
                  (c.gamma == 0 || al16(c.draft)) && (size_t(c.t_sb) * elt) % 16 == 0 &&
                  (size_t(c.t_ss) * elt) % 16 == 0 && (size_t(c.d_sb) * elt) % 16 == 0 &&
                  (size_t(c.d_ss) * elt) % 16 == 0 && al16(c.noise) && (c.noise_sb % 4) == 0;
  if (!ok) return false;
  const size_t rows = size_t(c.B) * size_t(2 * c.gamma + 1);
  const bool pairp = c.path == 5 && c.B <= 65535;
  if (pairp) {
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
    Q.S = 0;
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
    Q.partials = nullptr;
    Q.row_count = nullptr;
    Q.row_arrive = nullptr;
    Q.seq_key = reinterpret_cast<unsigned long long*>(w);
    w += size_t(c.B) * 8;
    Q.seq_bd = reinterpret_cast<unsigned long long*>(w);
    w += size_t(c.B) * 8;
    Q.seq_mass = reinterpret_cast<float*>(w);
    w += size_t(c.B) * 4;
    Q.seq_arrive = reinterpret_cast<int*>(w);
    w += size_t(c.B) * 4;
    w = reinterpret_cast<char*>((reinterpret_cast<uintptr_t>(w) + 127) & ~uintptr_t(127));  // TMA bulk stores land here
    Q.lists = reinterpret_cast<ListEntry*>(w);
    Q.stats = c.stats;
    Q.row0_rule = c.row0_rule;
    const dim3 g2(unsigned((c.Vt + kSlice3 - 1) / kSlice3), unsigned(c.B) + 1u);  // + the block row that aggregates the hints
    if (pairp) {
      const bool any_nucleus = (c.top_p > 0.0f) || (c.draft_top_p > 0.0f);
      const size_t smem = pair::pair_smem_bytes(any_nucleus ? int((c.Vt + kStageElems - 1) / kStageElems) : 0);
      if (smem > kPairSmemMax) return false;  // (vocabularies beyond ~700 k with a nucleus: general kernel)
      const int sms = stream_sm_count();
      const int per_sm = smem <= 115712 ? 2 : 1;  // two blocks per SM while the bit maps leave room (V <= ~40 k)
      const int npairs = int((rows + 1) / 2);
      const int grid = npairs < per_sm * sms ? npairs : per_sm * sms;  // two rows per block
      cudaError_t e;
      const size_t smem2 = sample_smem_bytes();
      if (c.dtype == DUODEC_F32) {
        e = ensure_stream_smem(reinterpret_cast<const void*>(pair::row_pair_kernel<float>), kPairSmemMax);
        if (e == cudaSuccess) e = ensure_stream_smem(reinterpret_cast<const void*>(seq_sample_tma_kernel<float>), smem2);
        if (e == cudaSuccess && (c.phases & 1))
          e = launch_pdl(pair::row_pair_kernel<float>, dim3(grid), dim3(pair::kPT), smem, c.stream, Q);
        if (e == cudaSuccess && (c.phases & 2)) e = launch_pdl(seq_sample_tma_kernel<float>, g2, dim3(kT3), smem2, c.stream, Q);
      } else {
        e = ensure_stream_smem(reinterpret_cast<const void*>(pair::row_pair_kernel<__nv_bfloat16>), kPairSmemMax);
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
  }
  return false;
}

}  // namespace duodec
