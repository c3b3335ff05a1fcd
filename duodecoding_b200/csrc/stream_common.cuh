// Constants and the reduction payload shared by the fused-verify fast paths.
#pragma once

#include "rowcore.cuh"

namespace duodec {

constexpr int kCandMax = 256;       // entries the all-pairs finish can order
constexpr float kCandGoal = 160.0f; // narrowing stops once the bracket provably holds at most this many
constexpr float kLog2e = 1.4426950408889634f;

struct __align__(16) R8 {  // reduction payload: v[0] max-reduced, v[1..7] summed
  float v[8];
};

__device__ __forceinline__ void mbar_arrive(unsigned long long* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}

}  // namespace duodec
