// Grid-wide phase flags for persistent kernels (all CTAs co-resident: grid <= SM count, 1 CTA / SM).
//
// A phase has one 32-bit counter in global memory, zero when the kernel starts.  A CTA that finished
// its share of the phase arrives once (release); a consumer spins until the counter reaches the
// number of arriving CTAs (acquire).  Counters are never reused inside a launch, so there is no
// sense reversal and no reset race; the host (or the kernel's last phase) zeroes them between
// launches.  Every spin is bounded: a lost arrival traps instead of hanging the GPU.
#pragma once
#include <stdint.h>

namespace b200 {

__device__ __forceinline__ void flag_arrive(unsigned int* ctr) {
  // generic-proxy writes of this thread (and, through the preceding CTA barrier, of the CTA) become
  // visible before the increment; fence.proxy.async orders them against later TMA (async proxy) reads
  asm volatile("fence.proxy.async;" ::: "memory");
  asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(ctr) : "memory");
}

__device__ __forceinline__ unsigned int flag_peek(const unsigned int* ctr) {
  unsigned int v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(ctr) : "memory");
  return v;
}

static __device__ __noinline__ void flag_wait_slow(const unsigned int* ctr, unsigned int target) {
#pragma unroll 1
  for (unsigned int spins = 0; flag_peek(ctr) < target; ++spins) {
    if (spins > (1u << 22)) __trap();          // ~ seconds: a phase never takes that long
  }
}

__device__ __forceinline__ void flag_wait(const unsigned int* ctr, unsigned int target) {
  if (flag_peek(ctr) < target) flag_wait_slow(ctr, target);
  asm volatile("fence.proxy.async;" ::: "memory");
}

}  // namespace b200
