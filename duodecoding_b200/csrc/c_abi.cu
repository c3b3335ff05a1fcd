// C ABI of the verification hot path (see include/duodec.h).  Host side only
// validates arguments, picks the team size and launches; no allocation, no
// synchronisation.
#include <atomic>
#include <cstdint>
#include <mutex>
#include <utility>
#include <vector>

#include "../../include/duodec.h"
#include "prob_kernels.cuh"
#include "cuhlm_kernels.cuh"
#include "row_kernels.cuh"
#include "select_kernels.cuh"
#include "shard_kernels.cuh"
#include "shard_nucleus.cuh"
#include "stream_launch.h"
#include "verify_kernels.cuh"

using namespace duodec;

namespace {

constexpr size_t kMaxDynSmem = 232448;  // 227 KB opt-in limit per block on sm_100
std::atomic<int> g_team_override{0};
std::atomic<int*> g_stats{nullptr};

size_t team_smem_bytes(int V, int rows_resident, int C) {
  const int per = C == 1 ? V : ((((V + C - 1) / C) + 7) & ~7);
  const size_t slice_floats = size_t((per + 7) & ~7) + 8;
  return size_t(rows_resident) * slice_floats * 4 + kHistBytes + sizeof(RowScratch) + 128;
}

int pick_team(int V, int rows_resident) {
  const int forced = g_team_override.load();
  if (forced == 1 || forced == 2 || forced == 4 || forced == 8) {
    if (team_smem_bytes(V, rows_resident, forced) <= kMaxDynSmem) return forced;
  }
  // smallest team that leaves room for two blocks per SM, else smallest that fits
  for (int C = 1; C <= kMaxTeam; C <<= 1)
    if (team_smem_bytes(V, rows_resident, C) <= kMaxDynSmem / 2) return C;
  for (int C = 1; C <= kMaxTeam; C <<= 1)
    if (team_smem_bytes(V, rows_resident, C) <= kMaxDynSmem) return C;
  return 0;
}

// cudaFuncSetAttribute is per (kernel, device): remember the pairs already set up.
cudaError_t ensure_smem(const void* kernel, size_t bytes = kMaxDynSmem) {
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

int sm_count() {
  static std::mutex mu;
  static int cached[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 0;
  std::lock_guard<std::mutex> lk(mu);
  if (cached[dev] == 0) {
    int n = 0;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return 0;
    cached[dev] = n;
  }
  return cached[dev];
}

template <class K, class P>
cudaError_t launch_team(K kernel, const P& params, int teams, int C, size_t smem, cudaStream_t st) {
  cudaError_t e = ensure_smem(reinterpret_cast<const void*>(kernel));
  if (e != cudaSuccess) return e;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(unsigned(teams) * unsigned(C), 1, 1);
  cfg.blockDim = dim3(kThreads, 1, 1);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = unsigned(C);
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = C > 1 ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, params);
}

#define DISPATCH_TEAM(C, ...)                         \
  switch (C) {                                        \
    case 1: { constexpr int kC = 1; __VA_ARGS__; } break; \
    case 2: { constexpr int kC = 2; __VA_ARGS__; } break; \
    case 4: { constexpr int kC = 4; __VA_ARGS__; } break; \
    case 8: { constexpr int kC = 8; __VA_ARGS__; } break; \
    default: return DUODEC_ETOOBIG;                   \
  }

}  // namespace

extern "C" {

const char* duodec_version(void) { return "duodec-b200 0.1.0 sm_100a"; }

int duodec_team_size(int V, int rows_resident) {
  if (V <= 0 || rows_resident < 1 || rows_resident > 2) return 0;
  return pick_team(V, rows_resident);
}

void duodec_set_team_size(int team) { g_team_override.store(team); }

void duodec_set_stats_buffer(int* device_counters) { g_stats.store(device_counters); }

int duodec_norm_logits_batched(const void* logits, int in_dtype, int64_t in_batch_stride, int64_t in_row_stride,
                               void* out, int out_dtype, int64_t out_batch_stride, int64_t out_row_stride, int batch,
                               int inner, int V, float temperature, int top_k, float top_p, void* stream) {
  if (batch < 0 || inner < 0 || V <= 0) return DUODEC_EINVAL;
  if (batch == 0 || inner == 0) return 0;
  if (int64_t(batch) * int64_t(inner) > 0x7fffffff / 8) return DUODEC_EINVAL;
  if (!logits || !out) return DUODEC_EINVAL;
  if ((in_dtype != DUODEC_F32 && in_dtype != DUODEC_BF16) || (out_dtype != DUODEC_F32 && out_dtype != DUODEC_BF16))
    return DUODEC_EINVAL;
  if (!(temperature >= 0.0f)) return DUODEC_EINVAL;
  const int rows = batch * inner;
  const int C = pick_team(V, 1);
  if (C == 0) return DUODEC_ETOOBIG;
  NormParams P;
  P.logits = logits;
  P.in_stride = in_row_stride;
  P.in_bstride = in_batch_stride;
  P.out = out;
  P.out_stride = out_row_stride;
  P.out_bstride = out_batch_stride;
  P.rows = rows;
  P.inner = inner;
  P.V = V;
  P.temperature = temperature;
  P.top_p = top_p;
  P.top_k = top_k;
  P.per = C == 1 ? V : ((((V + C - 1) / C) + 7) & ~7);
  P.slice_floats = ((P.per + 7) & ~7) + 8;
  const size_t smem = team_smem_bytes(V, 1, C);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  cudaError_t e = cudaSuccess;
  DISPATCH_TEAM(C, {
    if (in_dtype == DUODEC_F32 && out_dtype == DUODEC_F32)
      e = launch_team(norm_logits_kernel<float, float, kC>, P, rows, kC, smem, st);
    else if (in_dtype == DUODEC_BF16 && out_dtype == DUODEC_BF16)
      e = launch_team(norm_logits_kernel<__nv_bfloat16, __nv_bfloat16, kC>, P, rows, kC, smem, st);
    else if (in_dtype == DUODEC_BF16 && out_dtype == DUODEC_F32)
      e = launch_team(norm_logits_kernel<__nv_bfloat16, float, kC>, P, rows, kC, smem, st);
    else
      e = launch_team(norm_logits_kernel<float, __nv_bfloat16, kC>, P, rows, kC, smem, st);
  })
  return int(e);
}

int duodec_norm_logits(const void* logits, int in_dtype, int64_t in_row_stride, void* out, int out_dtype,
                       int64_t out_row_stride, int rows, int V, float temperature, int top_k, float top_p,
                       void* stream) {
  return duodec_norm_logits_batched(logits, in_dtype, 0, in_row_stride, out, out_dtype, 0, out_row_stride, 1, rows, V,
                                    temperature, top_k, top_p, stream);
}

size_t duodec_verify_workspace_bytes(int B, int gamma) {
  if (B <= 0 || gamma < 0) return 0;
  // [B] int counters (8 bytes each) + [B, gamma+1] int64 position tokens + 2 row-kind hints
  // + [B, 2*gamma+1] row records
  return ((size_t(B) * 8 + size_t(B) * size_t(gamma + 1) * 8 + 15) & ~size_t(15)) + stream_workspace_bytes(B, gamma);
}

int duodec_verify_fused(const void* draft, const void* target, int dtype, int64_t d_sb, int64_t d_ss, int64_t t_sb,
                        int64_t t_ss, int B, int gamma, int Vd, int Vt, const int64_t* tokens, int64_t tok_sb,
                        const float* r, int64_t r_sb, const float* noise, int64_t noise_sb, float temperature,
                        int top_k, float top_p, float draft_temperature, int draft_top_k, float draft_top_p,
                        int flags, uint8_t* accept_mask, float* sel_target_p, float* sel_draft_p,
                        int64_t* accepted_len, int64_t* token, uint8_t* all_accepted, void* workspace,
                        size_t workspace_bytes, void* stream) {
  if (B < 0 || gamma < 0 || Vd <= 0 || Vt <= 0) return DUODEC_EINVAL;
  if (B == 0) return 0;
  if (!target || !noise || !accepted_len || !token || !all_accepted || !workspace) return DUODEC_EINVAL;
  if (gamma > 0 && (!draft || !tokens || !r || !accept_mask || !sel_target_p || !sel_draft_p)) return DUODEC_EINVAL;
  if (dtype != DUODEC_F32 && dtype != DUODEC_BF16) return DUODEC_EINVAL;
  if (!(temperature >= 0.0f) || !(draft_temperature >= 0.0f)) return DUODEC_EINVAL;
  if ((temperature == 0.0f) != (draft_temperature == 0.0f)) return DUODEC_EINVAL;  // greedy on both or neither
  if (workspace_bytes < duodec_verify_workspace_bytes(B, gamma)) return DUODEC_EWORKSPACE;
  if (reinterpret_cast<uintptr_t>(workspace) & 15u) return DUODEC_EINVAL;  // records are read with 128-bit loads
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int path = (flags >> 4) & 15, phases = (flags >> 8) & 3;
  const bool row0 = (flags & DUODEC_VERIFY_ROW0_RULE) != 0 && B > 1 && gamma > 0;
  if (path != 0 && path != 1 && path != 5) return DUODEC_EINVAL;
  if (path != 1) {
    // fast path: persistent row-stream kernel (stream_launch.cu); falls through when the call does not qualify
    StreamCall c;
    c.draft = draft, c.target = target, c.dtype = dtype;
    c.d_sb = d_sb, c.d_ss = d_ss, c.t_sb = t_sb, c.t_ss = t_ss;
    c.B = B, c.gamma = gamma, c.Vd = Vd, c.Vt = Vt;
    c.tokens = tokens, c.tok_sb = tok_sb, c.r = r, c.r_sb = r_sb, c.noise = noise, c.noise_sb = noise_sb;
    c.temperature = temperature, c.top_k = top_k, c.top_p = top_p;
    c.draft_temperature = draft_temperature, c.draft_top_k = draft_top_k, c.draft_top_p = draft_top_p;
    c.accept_mask = accept_mask, c.sel_target_p = sel_target_p, c.sel_draft_p = sel_draft_p;
    c.accepted_len = accepted_len, c.token = token, c.all_accepted = all_accepted;
    c.workspace = workspace, c.stats = g_stats.load(), c.path = path == 0 ? 5 : path, c.phases = phases == 0 ? 3 : phases;
    c.row0_rule = row0 ? 1 : 0, c.stream = st;
    int rc = 0;
    if (stream_try_launch(c, &rc)) return rc;
  }
  // the general kernel closes every sequence on its own count: the batch rule needs the fast path
  if (row0) return DUODEC_EINVAL;
  const int Vmax = Vd > Vt ? Vd : Vt;
  const int C = pick_team(Vmax, 2);
  if (C == 0) return DUODEC_ETOOBIG;
  VerifyParams P;
  P.draft = draft;
  P.target = target;
  P.d_sb = d_sb;
  P.d_ss = d_ss;
  P.t_sb = t_sb;
  P.t_ss = t_ss;
  P.B = B;
  P.gamma = gamma;
  P.Vd = Vd;
  P.Vt = Vt;
  P.tokens = reinterpret_cast<const long long*>(tokens);
  P.tok_sb = tok_sb;
  P.r = r;
  P.r_sb = r_sb;
  P.noise = noise;
  P.noise_sb = noise_sb;
  P.t_temp = temperature;
  P.t_top_p = top_p;
  P.t_top_k = top_k;
  P.d_temp = draft_temperature;
  P.d_top_p = draft_top_p;
  P.d_top_k = draft_top_k;
  P.accept_mask = accept_mask;
  P.sel_p = sel_target_p;
  P.sel_q = sel_draft_p;
  P.accepted_len = reinterpret_cast<long long*>(accepted_len);
  P.token = reinterpret_cast<long long*>(token);
  P.all_accepted = all_accepted;
  P.counters = reinterpret_cast<int*>(workspace);
  P.pair_token = reinterpret_cast<long long*>(reinterpret_cast<char*>(workspace) + size_t(B) * 8);
  P.per = C == 1 ? Vmax : ((((Vmax + C - 1) / C) + 7) & ~7);
  P.slice_floats = ((P.per + 7) & ~7) + 8;
  const size_t smem = team_smem_bytes(Vmax, 2, C);
  const int teams = B * (gamma + 1);
  cudaError_t e = cudaSuccess;
  DISPATCH_TEAM(C, {
    if (dtype == DUODEC_F32) e = launch_team(verify_pair_kernel<float, kC>, P, teams, kC, smem, st);
    else e = launch_team(verify_pair_kernel<__nv_bfloat16, kC>, P, teams, kC, smem, st);
  })
  return int(e);
}

// ---- no-allocation variant of duodec_verify_fused (include/duodec.h): the workspace is this array.
// In host code the array's name is a host-side stand-in, not a device address: the address is resolved
// below, once per device, with that device current.
}  // extern "C"

static __device__ __align__(256) unsigned char g_static_workspace[DUODEC_VERIFY_STATIC_BYTES];

namespace {
struct StaticWorkspace {
  void* ptr = nullptr;       // device address of g_static_workspace on this device
  long long shape = -1;      // (B, gamma) of the last call: the layout inside the array depends on them
};
StaticWorkspace g_static_ws[64];
std::mutex g_static_mu;
}  // namespace

extern "C" {

size_t duodec_verify_static_max_bytes(void) { return size_t(DUODEC_VERIFY_STATIC_BYTES); }

int duodec_verify_fused_static(const void* draft, const void* target, int dtype, int64_t d_sb, int64_t d_ss,
                               int64_t t_sb, int64_t t_ss, int B, int gamma, int Vd, int Vt, const int64_t* tokens,
                               int64_t tok_sb, const float* r, int64_t r_sb, const float* noise, int64_t noise_sb,
                               float temperature, int top_k, float top_p, float draft_temperature, int draft_top_k,
                               float draft_top_p, int flags, uint8_t* accept_mask, float* sel_target_p,
                               float* sel_draft_p, int64_t* accepted_len, int64_t* token, uint8_t* all_accepted,
                               void* stream) {
  if (B < 0 || gamma < 0) return DUODEC_EINVAL;
  if (B == 0) return 0;
  const size_t need = duodec_verify_workspace_bytes(B, gamma);
  if (need > size_t(DUODEC_VERIFY_STATIC_BYTES)) return DUODEC_ESTATICMAX;
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return int(e);
  if (dev < 0 || dev >= 64) return DUODEC_EINVAL;
  void* ws = nullptr;
  {
    std::lock_guard<std::mutex> lk(g_static_mu);
    StaticWorkspace& w = g_static_ws[dev];
    if (!w.ptr) {
      void* p = nullptr;
      e = cudaGetSymbolAddress(&p, g_static_workspace);
      if (e != cudaSuccess) return int(e);
      w.ptr = p;
    }
    const long long shape = ((long long)B << 32) | (long long)gamma;
    if (w.shape != shape) {
      // the array holds another layout's contents (or nothing yet): what this call uses starts at zero
      e = cudaMemsetAsync(w.ptr, 0, need, static_cast<cudaStream_t>(stream));
      if (e != cudaSuccess) return int(e);
      w.shape = shape;
    }
    ws = w.ptr;
  }
  return duodec_verify_fused(draft, target, dtype, d_sb, d_ss, t_sb, t_ss, B, gamma, Vd, Vt, tokens, tok_sb, r, r_sb,
                             noise, noise_sb, temperature, top_k, top_p, draft_temperature, draft_top_k, draft_top_p,
                             flags, accept_mask, sel_target_p, sel_draft_p, accepted_len, token, all_accepted, ws,
                             size_t(DUODEC_VERIFY_STATIC_BYTES), stream);
}

int duodec_accept_probs(const void* draft_probs, int64_t d_sb, int64_t d_ss, const void* target_probs,
                        int64_t t_sb, int64_t t_ss, int dtype, const int64_t* tokens, int64_t tok_sb,
                        const float* r, int64_t r_sb, int B, int gamma, int V, uint8_t* accept_mask,
                        void* sel_draft_p, void* sel_target_p, int64_t* accepted_len, void* stream) {
  if (B < 0 || gamma < 0 || V <= 0) return DUODEC_EINVAL;
  if (B == 0) return 0;
  if (!accepted_len) return DUODEC_EINVAL;
  if (gamma > 0 && (!draft_probs || !target_probs || !tokens || !r || !accept_mask || !sel_draft_p || !sel_target_p))
    return DUODEC_EINVAL;
  if (dtype != DUODEC_F32 && dtype != DUODEC_BF16) return DUODEC_EINVAL;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int threads = gamma <= 32 ? 32 : (gamma <= 256 ? 256 : 1024);
  if (dtype == DUODEC_F32)
    accept_probs_kernel<float><<<B, threads, 0, st>>>(
        static_cast<const float*>(draft_probs), d_sb, d_ss, static_cast<const float*>(target_probs), t_sb, t_ss,
        reinterpret_cast<const long long*>(tokens), tok_sb, r, r_sb, gamma, V, accept_mask,
        static_cast<float*>(sel_draft_p), static_cast<float*>(sel_target_p),
        reinterpret_cast<long long*>(accepted_len));
  else
    accept_probs_kernel<__nv_bfloat16><<<B, threads, 0, st>>>(
        static_cast<const __nv_bfloat16*>(draft_probs), d_sb, d_ss, static_cast<const __nv_bfloat16*>(target_probs),
        t_sb, t_ss, reinterpret_cast<const long long*>(tokens), tok_sb, r, r_sb, gamma, V, accept_mask,
        static_cast<__nv_bfloat16*>(sel_draft_p), static_cast<__nv_bfloat16*>(sel_target_p),
        reinterpret_cast<long long*>(accepted_len));
  return int(cudaGetLastError());
}

int duodec_sample_rows(const void* probs, int dtype, int64_t row_stride, const float* noise, int64_t noise_stride,
                       int rows, int V, int64_t* out, void* stream) {
  if (rows < 0 || V <= 0) return DUODEC_EINVAL;
  if (rows == 0) return 0;
  if (!probs || !noise || !out) return DUODEC_EINVAL;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (dtype == DUODEC_F32)
    sample_rows_kernel<float><<<rows, kThreads, 0, st>>>(static_cast<const float*>(probs), row_stride, noise,
                                                         noise_stride, V, reinterpret_cast<long long*>(out));
  else if (dtype == DUODEC_BF16)
    sample_rows_kernel<__nv_bfloat16><<<rows, kThreads, 0, st>>>(static_cast<const __nv_bfloat16*>(probs),
                                                                 row_stride, noise, noise_stride, V,
                                                                 reinterpret_cast<long long*>(out));
  else
    return DUODEC_EINVAL;
  return int(cudaGetLastError());
}

int duodec_residual_sample(const void* target_probs, int64_t t_stride, const void* draft_probs, int64_t d_stride,
                           int dtype, const float* noise, int64_t noise_stride, int rows, int V, int64_t* out,
                           void* stream) {
  if (rows < 0 || V <= 0) return DUODEC_EINVAL;
  if (rows == 0) return 0;
  if (!target_probs || !draft_probs || !noise || !out) return DUODEC_EINVAL;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (dtype == DUODEC_F32)
    residual_sample_kernel<float><<<rows, kThreads, 0, st>>>(static_cast<const float*>(target_probs), t_stride,
                                                             static_cast<const float*>(draft_probs), d_stride, noise,
                                                             noise_stride, V, reinterpret_cast<long long*>(out));
  else if (dtype == DUODEC_BF16)
    residual_sample_kernel<__nv_bfloat16><<<rows, kThreads, 0, st>>>(
        static_cast<const __nv_bfloat16*>(target_probs), t_stride, static_cast<const __nv_bfloat16*>(draft_probs),
        d_stride, noise, noise_stride, V, reinterpret_cast<long long*>(out));
  else
    return DUODEC_EINVAL;
  return int(cudaGetLastError());
}

int duodec_max_fn(const void* x, int dtype, int64_t row_stride, int rows, int V, float* out, int64_t out_stride,
                  void* stream) {
  if (rows < 0 || V <= 0) return DUODEC_EINVAL;
  if (rows == 0) return 0;
  if (!x || !out) return DUODEC_EINVAL;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (dtype == DUODEC_F32)
    max_fn_kernel<float><<<rows, kThreads, 0, st>>>(static_cast<const float*>(x), row_stride, V, out, out_stride);
  else if (dtype == DUODEC_BF16)
    max_fn_kernel<__nv_bfloat16><<<rows, kThreads, 0, st>>>(static_cast<const __nv_bfloat16*>(x), row_stride, V, out,
                                                           out_stride);
  else
    return DUODEC_EINVAL;
  return int(cudaGetLastError());
}

int duodec_topk_rebuild(const void* probs, int dtype, int64_t in_stride, void* out, int64_t out_stride, int rows,
                        int V, int k, int mode, void* stream) {
  if (rows < 0 || V <= 0 || mode < 0 || mode > 3) return DUODEC_EINVAL;
  if (rows == 0) return 0;
  if (!probs || !out) return DUODEC_EINVAL;
  if (mode != 3 && (k <= 0 || k >= V)) return DUODEC_EINVAL;
  if (dtype != DUODEC_F32 && dtype != DUODEC_BF16) return DUODEC_EINVAL;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (g_team_override.load() == 0 && rowsel::select_smem_bytes(V) <= kMaxDynSmem) {
    // the row fits one SM: one block per row, one read and one write of the row (select_kernels.cuh)
    TopkParams S;
    S.probs = probs, S.in_stride = in_stride, S.out = out, S.out_stride = out_stride;
    S.rows = rows, S.V = V, S.k = k, S.mode = mode, S.per = V, S.slice_floats = 0;
    const size_t bytes = rowsel::select_smem_bytes(V);
    const size_t elt = dtype == DUODEC_F32 ? 4 : 2;
    const int per = int(16 / elt);
    // rows of <= 32768 entries with 128-bit aligned rows live in registers; the others in shared memory
    const bool regs = V <= rowsel::kRegRowMax && V % per == 0 && (reinterpret_cast<size_t>(probs) & 15) == 0 &&
                      (reinterpret_cast<size_t>(out) & 15) == 0 && (size_t(in_stride) * elt) % 16 == 0 &&
                      (size_t(out_stride) * elt) % 16 == 0;
    if (regs) {
      if (dtype == DUODEC_F32) rowsel::topk_reg_kernel<float><<<unsigned(rows), rowsel::kRT, 0, st>>>(S);
      else rowsel::topk_reg_kernel<__nv_bfloat16><<<unsigned(rows), rowsel::kRT, 0, st>>>(S);
    } else if (dtype == DUODEC_F32) {
      cudaError_t e = ensure_smem(reinterpret_cast<const void*>(rowsel::topk_select_kernel<float>));
      if (e != cudaSuccess) return int(e);
      rowsel::topk_select_kernel<float><<<unsigned(rows), rowsel::kRT, bytes, st>>>(S);
    } else {
      cudaError_t e = ensure_smem(reinterpret_cast<const void*>(rowsel::topk_select_kernel<__nv_bfloat16>));
      if (e != cudaSuccess) return int(e);
      rowsel::topk_select_kernel<__nv_bfloat16><<<unsigned(rows), rowsel::kRT, bytes, st>>>(S);
    }
    return int(cudaGetLastError());
  }
  const int C = pick_team(V, 1);
  if (C == 0) return DUODEC_ETOOBIG;
  TopkParams P;
  P.probs = probs;
  P.in_stride = in_stride;
  P.out = out;
  P.out_stride = out_stride;
  P.rows = rows;
  P.V = V;
  P.k = k;
  P.mode = mode;
  P.per = C == 1 ? V : ((((V + C - 1) / C) + 7) & ~7);
  P.slice_floats = ((P.per + 7) & ~7) + 8;
  const size_t smem = team_smem_bytes(V, 1, C);
  cudaError_t e = cudaSuccess;
  DISPATCH_TEAM(C, {
    if (dtype == DUODEC_F32) e = launch_team(topk_rebuild_kernel<float, kC>, P, rows, kC, smem, st);
    else e = launch_team(topk_rebuild_kernel<__nv_bfloat16, kC>, P, rows, kC, smem, st);
  })
  return int(e);
}

int duodec_kv_truncate(int32_t* seq_len, const int64_t* end_pos_vec, int64_t end_pos, int B, void* stream) {
  if (B < 0) return DUODEC_EINVAL;
  if (B == 0) return 0;
  if (!seq_len) return DUODEC_EINVAL;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  kv_truncate_kernel<<<(B + 255) / 256, 256, 0, st>>>(seq_len, reinterpret_cast<const long long*>(end_pos_vec), B,
                                                      (long long)end_pos);
  return int(cudaGetLastError());
}

// ---- vocabulary-sharded verification (shard_kernels.cuh) ------------------------------------
static bool shard_rows_ok(const void* draft, const void* target, int dtype, int B, int gamma, int Vs) {
  return B >= 0 && gamma >= 0 && gamma <= 32 && Vs > 0 && target && (gamma == 0 || draft) &&
         (dtype == DUODEC_F32 || dtype == DUODEC_BF16);
}

static bool xchg_ok(const duodec_shard_xchg* x) {
  return x == nullptr || (x->peers != nullptr && x->world >= 1 && x->world <= 32 && x->rank >= 0 && x->rank < x->world &&
                          x->off_stats >= 1024 && x->off_partial >= 1024 && x->off_stats % 16 == 0 &&
                          x->off_partial % 8 == 0);
}
static ShardXchg xchg_of(const duodec_shard_xchg* x) {
  if (x == nullptr) return ShardXchg{nullptr, 0, 1, 0, 0, 0};
  return ShardXchg{x->peers, x->rank, x->world, x->off_stats, x->off_partial, x->mode};
}

int duodec_shard_stats(const void* draft, const void* target, int dtype, int64_t d_sb, int64_t d_ss, int64_t t_sb,
                       int64_t t_ss, int B, int gamma, int Vs, int v_begin, const int64_t* tokens, int64_t tok_sb,
                       float temperature, float draft_temperature, float* stats, const duodec_shard_xchg* xchg,
                       void* stream) {
  if (!shard_rows_ok(draft, target, dtype, B, gamma, Vs) || temperature <= 0.0f || draft_temperature <= 0.0f)
    return DUODEC_EINVAL;
  if (B == 0) return 0;
  if (!xchg_ok(xchg) || (!stats && !xchg) || (gamma > 0 && !tokens)) return DUODEC_EINVAL;
  const ShardXchg X = xchg_of(xchg);
  const ShardRows S{draft, target, d_sb, d_ss, t_sb, t_ss, B, gamma, Vs, v_begin, temperature, draft_temperature};
  const int rows = B * (2 * gamma + 1);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (dtype == DUODEC_F32)
    shard_stats_kernel<float><<<rows, kThreads, 0, st>>>(S, reinterpret_cast<const long long*>(tokens), tok_sb,
                                                         reinterpret_cast<float4*>(stats), X);
  else
    shard_stats_kernel<__nv_bfloat16><<<rows, kThreads, 0, st>>>(S, reinterpret_cast<const long long*>(tokens), tok_sb,
                                                                 reinterpret_cast<float4*>(stats), X);
  return int(cudaGetLastError());
}

int duodec_shard_accept(const float* gathered_stats, int world, int B, int gamma, const float* r, int64_t r_sb,
                        uint8_t* accept_mask, float* sel_target_p, float* sel_draft_p, int64_t* accepted_len,
                        uint8_t* all_accepted, void* plan, const duodec_shard_xchg* xchg, const int64_t* tokens,
                        int64_t tok_sb, const void* cuts, void* stream) {
  if (world <= 0 || B < 0 || gamma < 0 || gamma > 32) return DUODEC_EINVAL;
  if (B == 0) return 0;
  if (!xchg_ok(xchg) || (xchg && xchg->world != world)) return DUODEC_EINVAL;
  if (cuts && ((gamma > 0 && !tokens) || xchg)) return DUODEC_EINVAL;  // the filter runs on gathered statistics
  if ((!gathered_stats && !xchg) || !accepted_len || !all_accepted || !plan ||
      (gamma > 0 && (!r || !accept_mask || !sel_target_p || !sel_draft_p)))
    return DUODEC_EINVAL;
  shard_accept_kernel<<<B, 32, 0, static_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const float4*>(gathered_stats), world, B, gamma, r, r_sb, accept_mask, sel_target_p,
      sel_draft_p, reinterpret_cast<long long*>(accepted_len), all_accepted, static_cast<ShardPlan*>(plan),
      xchg_of(xchg), reinterpret_cast<const long long*>(tokens), tok_sb, static_cast<const ShardCut*>(cuts));
  return int(cudaGetLastError());
}

int duodec_shard_sample(const void* draft, const void* target, int dtype, int64_t d_sb, int64_t d_ss, int64_t t_sb,
                        int64_t t_ss, int B, int gamma, int Vs, int v_begin, float temperature,
                        float draft_temperature, const void* plan, const float* noise, int64_t noise_sb,
                        int nslice, void* partial, const duodec_shard_xchg* xchg, int filtered, void* stream) {
  if (!shard_rows_ok(draft, target, dtype, B, gamma, Vs)) return DUODEC_EINVAL;
  if (B == 0) return 0;
  if (!xchg_ok(xchg) || !plan || !noise || (!partial && !xchg) || nslice <= 0 || nslice > 65535 || B > 65535)
    return DUODEC_EINVAL;
  const ShardXchg X = xchg_of(xchg);
  const ShardRows S{draft, target, d_sb, d_ss, t_sb, t_ss, B, gamma, Vs, v_begin, temperature, draft_temperature};
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const dim3 grid{unsigned(nslice), unsigned(B), 1u};
  const ShardPlan* pl = static_cast<const ShardPlan*>(plan);
  ShardPartial* out = static_cast<ShardPartial*>(partial);
  if (dtype == DUODEC_F32) {
    if (filtered) shard_sample_kernel<float, true><<<grid, kThreads, 0, st>>>(S, pl, noise, noise_sb, out, X);
    else shard_sample_kernel<float, false><<<grid, kThreads, 0, st>>>(S, pl, noise, noise_sb, out, X);
  } else {
    if (filtered) shard_sample_kernel<__nv_bfloat16, true><<<grid, kThreads, 0, st>>>(S, pl, noise, noise_sb, out, X);
    else shard_sample_kernel<__nv_bfloat16, false><<<grid, kThreads, 0, st>>>(S, pl, noise, noise_sb, out, X);
  }
  return int(cudaGetLastError());
}

// ---- top-p under vocabulary sharding (shard_nucleus.cuh)
int duodec_shard_nucleus_levels(int dtype, float temperature, float draft_temperature) {
  return (dtype == DUODEC_BF16 && temperature == 1.0f && draft_temperature == 1.0f) ? 4 : 8;
}

int duodec_shard_nucleus_level(const void* draft, const void* target, int dtype, int64_t d_sb, int64_t d_ss,
                               int64_t t_sb, int64_t t_ss, int B, int gamma, int Vs, int v_begin, float temperature,
                               float draft_temperature, float top_p, float draft_top_p, const float* gathered_stats,
                               int world, void* sel, int level, void* hist, void* stream) {
  if (!shard_rows_ok(draft, target, dtype, B, gamma, Vs) || temperature <= 0.0f || draft_temperature <= 0.0f)
    return DUODEC_EINVAL;
  if (B == 0) return 0;
  if (!gathered_stats || world <= 0 || !sel || !hist || level < 0 || level > 7) return DUODEC_EINVAL;
  if (Vs > kNucMaxPerThread * kThreads) return DUODEC_ETOOBIG;  // packed per-thread (mass, count) columns
  const ShardRows S{draft, target, d_sb, d_ss, t_sb, t_ss, B, gamma, Vs, v_begin, temperature, draft_temperature};
  const int rows = B * (2 * gamma + 1);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const float4* gs = reinterpret_cast<const float4*>(gathered_stats);
  if (dtype == DUODEC_F32) {
    cudaError_t e = ensure_smem(reinterpret_cast<const void*>(shard_nuc_level_kernel<float>), kNucLevelSmem);
    if (e != cudaSuccess) return int(e);
    shard_nuc_level_kernel<float><<<rows, kThreads, kNucLevelSmem, st>>>(S, gs, world, top_p, draft_top_p,
                                                                         static_cast<ShardSel*>(sel), level,
                                                                         static_cast<unsigned char*>(hist));
  } else {
    cudaError_t e = ensure_smem(reinterpret_cast<const void*>(shard_nuc_level_kernel<__nv_bfloat16>), kNucLevelSmem);
    if (e != cudaSuccess) return int(e);
    shard_nuc_level_kernel<__nv_bfloat16><<<rows, kThreads, kNucLevelSmem, st>>>(
        S, gs, world, top_p, draft_top_p, static_cast<ShardSel*>(sel), level, static_cast<unsigned char*>(hist));
  }
  return int(cudaGetLastError());
}

int duodec_shard_nucleus_reduce(const void* gathered_hist, int world, int rank, int B, int gamma, float top_p,
                                float draft_top_p, void* sel, int level, int levels, int32_t* quota, void* stream) {
  if (world <= 0 || rank < 0 || rank >= world || B < 0 || gamma < 0) return DUODEC_EINVAL;
  if (B == 0) return 0;
  if (!gathered_hist || !sel || !quota || (levels != 4 && levels != 8) || level < 0 || level >= levels)
    return DUODEC_EINVAL;
  const int rows = B * (2 * gamma + 1);
  shard_nuc_reduce_kernel<<<(rows + 3) / 4, 128, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const unsigned char*>(gathered_hist), world, rows, gamma, top_p, draft_top_p,
      static_cast<ShardSel*>(sel), level, levels, rank, quota);
  return int(cudaGetLastError());
}

int duodec_shard_nucleus_tie(const void* draft, const void* target, int dtype, int64_t d_sb, int64_t d_ss, int64_t t_sb,
                             int64_t t_ss, int B, int gamma, int Vs, int v_begin, float temperature,
                             float draft_temperature, const void* sel, const int32_t* quota, int levels,
                             int32_t* cand, void* stream) {
  if (!shard_rows_ok(draft, target, dtype, B, gamma, Vs) || temperature <= 0.0f || draft_temperature <= 0.0f)
    return DUODEC_EINVAL;
  if (B == 0) return 0;
  if (!sel || !quota || !cand || (levels != 4 && levels != 8)) return DUODEC_EINVAL;
  const ShardRows S{draft, target, d_sb, d_ss, t_sb, t_ss, B, gamma, Vs, v_begin, temperature, draft_temperature};
  const int rows = B * (2 * gamma + 1);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (dtype == DUODEC_F32)
    shard_nuc_tie_kernel<float><<<rows, kThreads, 0, st>>>(S, static_cast<const ShardSel*>(sel), quota, levels, cand);
  else
    shard_nuc_tie_kernel<__nv_bfloat16><<<rows, kThreads, 0, st>>>(S, static_cast<const ShardSel*>(sel), quota, levels,
                                                                   cand);
  return int(cudaGetLastError());
}

int duodec_shard_nucleus_close(const void* sel, const float* gathered_stats, const int32_t* gathered_cand, int world,
                               int B, int gamma, int levels, void* cuts, void* stream) {
  if (world <= 0 || B < 0 || gamma < 0) return DUODEC_EINVAL;
  if (B == 0) return 0;
  if (!sel || !gathered_stats || !gathered_cand || !cuts || (levels != 4 && levels != 8)) return DUODEC_EINVAL;
  const int rows = B * (2 * gamma + 1);
  shard_nuc_close_kernel<<<(rows + 127) / 128, 128, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const ShardSel*>(sel), reinterpret_cast<const float4*>(gathered_stats), gathered_cand, world, rows,
      levels, static_cast<ShardCut*>(cuts));
  return int(cudaGetLastError());
}

int duodec_shard_finish(const void* gathered_partials, int world, int B, int nslice, int64_t* token,
                        const duodec_shard_xchg* xchg, void* stream) {
  if (world <= 0 || B < 0 || nslice <= 0) return DUODEC_EINVAL;
  if (B == 0) return 0;
  if (!xchg_ok(xchg) || (xchg && xchg->world != world) || (!gathered_partials && !xchg) || !token) return DUODEC_EINVAL;
  const bool one = xchg && xchg->mode != 2;  // the flag protocol closes the step in one block
  const int blocks = one ? 1 : (B + 127) / 128;
  shard_finish_kernel<<<blocks, one ? 256 : 128, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<const ShardPartial*>(gathered_partials), world, B, nslice, reinterpret_cast<long long*>(token),
      xchg_of(xchg));
  return int(cudaGetLastError());
}

int duodec_compact_nonzero(const void* x, int dtype, int64_t stride, int rows, int Vs, int v_begin, int kcap,
                           float* cand_val, int32_t* cand_idx, int32_t* count, void* stream) {
  if (rows < 0 || Vs <= 0 || kcap <= 0) return DUODEC_EINVAL;
  if (rows == 0) return 0;
  if (!x || !cand_val || !cand_idx || !count || (dtype != DUODEC_F32 && dtype != DUODEC_BF16)) return DUODEC_EINVAL;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (dtype == DUODEC_F32)
    compact_nonzero_kernel<float><<<rows, 256, 0, st>>>(static_cast<const float*>(x), stride, Vs, v_begin, kcap,
                                                        cand_val, cand_idx, count);
  else
    compact_nonzero_kernel<__nv_bfloat16><<<rows, 256, 0, st>>>(static_cast<const __nv_bfloat16*>(x), stride, Vs,
                                                                v_begin, kcap, cand_val, cand_idx, count);
  return int(cudaGetLastError());
}

int duodec_shard_topk_apply(const void* x, int dtype, int64_t in_stride, void* out, int64_t out_stride, int rows,
                            int Vs, int v_begin, const float* tau, const int32_t* istar, const float* fill,
                            const float* total, void* stream) {
  if (rows < 0 || Vs <= 0) return DUODEC_EINVAL;
  if (rows == 0) return 0;
  if (!x || !out || !tau || !istar || !fill || !total || (dtype != DUODEC_F32 && dtype != DUODEC_BF16))
    return DUODEC_EINVAL;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (dtype == DUODEC_F32)
    shard_topk_apply_kernel<float><<<rows, 256, 0, st>>>(static_cast<const float*>(x), in_stride,
                                                         static_cast<float*>(out), out_stride, Vs, v_begin, tau,
                                                         istar, fill, total);
  else
    shard_topk_apply_kernel<__nv_bfloat16><<<rows, 256, 0, st>>>(
        static_cast<const __nv_bfloat16*>(x), in_stride, static_cast<__nv_bfloat16*>(out), out_stride, Vs, v_begin,
        tau, istar, fill, total);
  return int(cudaGetLastError());
}

int duodec_token_ranks(const void* probs, int dtype, int64_t row_stride, int rows, int V, const int64_t* tokens,
                       int64_t tok_stride, int32_t* out, void* stream) {
  if (rows < 0 || V <= 0) return DUODEC_EINVAL;
  if (rows == 0) return 0;
  if (!probs || !tokens || !out || (dtype != DUODEC_F32 && dtype != DUODEC_BF16)) return DUODEC_EINVAL;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (dtype == DUODEC_F32)
    token_rank_kernel<float><<<rows, kThreads, 0, st>>>(static_cast<const float*>(probs), row_stride, V,
                                                        reinterpret_cast<const long long*>(tokens), tok_stride, out);
  else
    token_rank_kernel<__nv_bfloat16><<<rows, kThreads, 0, st>>>(static_cast<const __nv_bfloat16*>(probs), row_stride,
                                                                V, reinterpret_cast<const long long*>(tokens),
                                                                tok_stride, out);
  return int(cudaGetLastError());
}

int duodec_uncertainty_tokens(const void* logits, int dtype, int V, const float* temps, const float* noise,
                              int64_t noise_stride, int M, int64_t* tokens, void* stream) {
  if (V <= 0 || M < 0) return DUODEC_EINVAL;
  if (M == 0) return 0;
  if (!logits || !temps || !noise || !tokens || (dtype != DUODEC_F32 && dtype != DUODEC_BF16)) return DUODEC_EINVAL;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (dtype == DUODEC_F32)
    uncertainty_tokens_kernel<float><<<M, kThreads, 0, st>>>(static_cast<const float*>(logits), V, temps, noise,
                                                            noise_stride, reinterpret_cast<long long*>(tokens));
  else
    uncertainty_tokens_kernel<__nv_bfloat16><<<M, kThreads, 0, st>>>(static_cast<const __nv_bfloat16*>(logits), V,
                                                                    temps, noise, noise_stride,
                                                                    reinterpret_cast<long long*>(tokens));
  return int(cudaGetLastError());
}

int duodec_compressed_vocab_size(const float* sorted_probs, const double* prefix, int V, double denominator,
                                 double theta, int32_t* k_out, void* stream) {
  if (V <= 1) return DUODEC_EINVAL;
  if (!sorted_probs || !prefix || !k_out || !(denominator > 0.0)) return DUODEC_EINVAL;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  cudaError_t e = cudaMemsetAsync(k_out, 0x7f, sizeof(int32_t), st);  // 0x7f7f7f7f: "no k found"
  if (e != cudaSuccess) return int(e);
  compressed_vocab_kernel<<<(V - 1 + 255) / 256, 256, 0, st>>>(sorted_probs, prefix, V, denominator, theta, k_out);
  return int(cudaGetLastError());
}

}  // extern "C"
