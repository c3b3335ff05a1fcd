// Host-side interface between c_abi.cu and the row-stream kernel's own translation unit
// (stream_launch.cu), so that the two compile independently.
#pragma once

#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

namespace duodec {

struct StreamCall {  // the arguments of duodec_verify_fused, verbatim
  const void* draft;
  const void* target;
  int dtype;
  int64_t d_sb, d_ss, t_sb, t_ss;
  int B, gamma, Vd, Vt;
  const int64_t* tokens;
  int64_t tok_sb;
  const float* r;
  int64_t r_sb;
  const float* noise;
  int64_t noise_sb;
  float temperature;
  int top_k;
  float top_p;
  float draft_temperature;
  int draft_top_k;
  float draft_top_p;
  uint8_t* accept_mask;
  float* sel_target_p;
  float* sel_draft_p;
  int64_t* accepted_len;
  int64_t* token;
  uint8_t* all_accepted;
  void* workspace;
  int* stats;
  int phases;  // bit 0: scan kernel, bit 1: sample kernel (3 = both; measurement hook)
  int path;  // 6: row-capture scan kernel (default), 5: row-pair scan kernel
  int row0_rule;  // sequence 0's accepted count positions every sequence
  cudaStream_t stream;
};

// bytes the row-stream kernel needs after the general kernel's counters and position tokens
size_t stream_workspace_bytes(int B, int gamma);

// Launches the row-stream kernel when the call qualifies (rows that TMA can move: 16-byte aligned
// rows and noise, equal vocabularies, no top-k, sampled temperatures; any vocabulary size) and
// returns true with the C-ABI return code in *rc; returns false when the general kernel must run.
bool stream_try_launch(const StreamCall& c, int* rc);

}  // namespace duodec
