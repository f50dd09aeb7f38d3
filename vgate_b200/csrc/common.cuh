// Shared device helpers for libb200gate (sm_100a only).
#pragma once
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#if defined(__CUDA_ARCH__) && !defined(__CUDA_ARCH_FEAT_SM100_ALL) && !defined(__CUDA_ARCH_FEAT_SM103_ALL)
#error "libb200gate kernels are written for sm_100a (compile with -gencode arch=compute_100a,code=sm_100a)"
#endif

namespace b200 {

constexpr int kWarp = 32;

// ---- warp / block reductions -------------------------------------------------
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// ---- 128-bit streaming loads -------------------------------------------------
// Weights are read exactly once per step: keep them out of L1.
__device__ __forceinline__ uint4 ldg_stream(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}
// Activations are re-read by every warp of the CTA: let L1 keep them.
__device__ __forceinline__ uint4 ldg_cached(const void* p) {
  return __ldg(reinterpret_cast<const uint4*>(p));
}

// ---- Blackwell mixed-precision FMA: d = fp32(a) * fp32(b) + c, one rounding (SASS FHFMA) ----
__device__ __forceinline__ float fhfma_lo(uint32_t a2, uint32_t b2, float c) {
  asm("{.reg .b16 al, ah, bl, bh;\n\t"
      "mov.b32 {al, ah}, %1;\n\t"
      "mov.b32 {bl, bh}, %2;\n\t"
      "fma.rn.f32.f16 %0, al, bl, %0;}" : "+f"(c) : "r"(a2), "r"(b2));
  return c;
}
__device__ __forceinline__ float fhfma_hi(uint32_t a2, uint32_t b2, float c) {
  asm("{.reg .b16 al, ah, bl, bh;\n\t"
      "mov.b32 {al, ah}, %1;\n\t"
      "mov.b32 {bl, bh}, %2;\n\t"
      "fma.rn.f32.f16 %0, ah, bh, %0;}" : "+f"(c) : "r"(a2), "r"(b2));
  return c;
}
// acc += a2.lo*b2.lo + a2.hi*b2.hi   (two exact-product FMAs in k order)
__device__ __forceinline__ float fhfma2(uint32_t a2, uint32_t b2, float c) {
  return fhfma_hi(a2, b2, fhfma_lo(a2, b2, c));
}

// ---- AWQ int4 -> fp16 dequant --------------------------------------------------
// A packed word holds eight 4-bit weights of one output row for eight consecutive k;
// logical k_i sits in nibble kAwqOrder[i] = {0,4,1,5,2,6,3,7} (the AWQ interleave, applied
// along K at load time).  `(w >> 4j) & 0x000f000f | 0x64006400` is then the fp16 pair
// (1024 + q[2j], 1024 + q[2j+1]).  Subtracting (1024 + z) is exact in fp16 and the one
// multiply by the scale rounds exactly like the reference's fp16 `(q - z) * s`.
__device__ __forceinline__ uint32_t lop3_and_or(uint32_t a, uint32_t b, uint32_t c) {   // (a & b) | c in one LOP3
  uint32_t d;
  asm("lop3.b32 %0, %1, %2, %3, 0xEA;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
  return d;
}
// Even nibble pairs (j = 0, 2) are masked in place -> 1024 + q, minus (1024 + z): exact.  Odd pairs
// (j = 1, 3) are masked one nibble up -> 1024 + 16 q, and one exact FMA (x / 16 - (64 + z)) = q - z
// replaces the shift.  Either way the fp16 value q - z is exact and the single multiply by the scale
// rounds exactly like the reference's fp16 (q - z) * s.  13 instructions per 8 weights.
__device__ __forceinline__ void dequant8(uint32_t w, uint32_t z1024_2, uint32_t s2, uint32_t out[4]) {
  const __half2 z = *reinterpret_cast<const __half2*>(&z1024_2);
  const __half2 s = *reinterpret_cast<const __half2*>(&s2);
  const __half2 z64 = __hsub2(z, __float2half2_rn(960.f));           // 64 + zero (exact)
  const __half2 sixteenth = __float2half2_rn(0.0625f);
  const uint32_t w8 = w >> 8;
  uint32_t v0 = lop3_and_or(w, 0x000f000fu, 0x64006400u);             // nibbles 0,4 -> k0,k1
  uint32_t v1 = lop3_and_or(w, 0x00f000f0u, 0x64006400u);             // nibbles 1,5 -> k2,k3 (x16)
  uint32_t v2 = lop3_and_or(w8, 0x000f000fu, 0x64006400u);            // nibbles 2,6 -> k4,k5
  uint32_t v3 = lop3_and_or(w8, 0x00f000f0u, 0x64006400u);            // nibbles 3,7 -> k6,k7 (x16)
  __half2 h0 = __hmul2(__hsub2(*reinterpret_cast<__half2*>(&v0), z), s);
  __half2 h1 = __hmul2(__hfma2(*reinterpret_cast<__half2*>(&v1), sixteenth, __hneg2(z64)), s);
  __half2 h2 = __hmul2(__hsub2(*reinterpret_cast<__half2*>(&v2), z), s);
  __half2 h3 = __hmul2(__hfma2(*reinterpret_cast<__half2*>(&v3), sixteenth, __hneg2(z64)), s);
  out[0] = *reinterpret_cast<uint32_t*>(&h0);
  out[1] = *reinterpret_cast<uint32_t*>(&h1);
  out[2] = *reinterpret_cast<uint32_t*>(&h2);
  out[3] = *reinterpret_cast<uint32_t*>(&h3);
}

__device__ __forceinline__ uint32_t pack_half2(float a, float b) {
  __half2 h = __floats2half2_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&h);
}

// ---- 16-bit activation / weight storage: fp16 (the AWQ path) or bf16 (dense checkpoints) -----------------------
// `bf` is uniform per launch; the arrays are typed __half* either way (16-bit storage).
__device__ __forceinline__ uint32_t pack_bf16x2(float a, float b) {
  __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&h);
}
__device__ __forceinline__ uint32_t pack16x2(float a, float b, bool bf) { return bf ? pack_bf16x2(a, b) : pack_half2(a, b); }
__device__ __forceinline__ uint16_t to16(float a, bool bf) {
  if (bf) { const __nv_bfloat16 h = __float2bfloat16_rn(a); return *reinterpret_cast<const uint16_t*>(&h); }
  const __half h = __float2half_rn(a);
  return *reinterpret_cast<const uint16_t*>(&h);
}
__device__ __forceinline__ float from16(uint16_t v, bool bf) {
  return bf ? __uint_as_float((uint32_t)v << 16) : __half2float(*reinterpret_cast<const __half*>(&v));
}
__device__ __forceinline__ float2 unpack16x2(uint32_t w, bool bf) {
  if (bf) return make_float2(__uint_as_float(w << 16), __uint_as_float(w & 0xffff0000u));
  return __half22float2(*reinterpret_cast<const __half2*>(&w));
}

// ---- shared-memory addresses, mbarrier, bulk async copies (TMA 1-D) ---------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void fence_mbar_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void fence_proxy_async_smem() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
  return ok != 0;
}
// Bounded wait: a pipeline bug must fail loudly (trap -> sticky CUDA error), never hang the GPU.
// The retry loop is kept out of line and un-unrolled: with a dozen call sites per kernel an inlined,
// unrolled spin loop multiplies the code size and costs instruction-cache misses on the hot path.
static __device__ __noinline__ void mbar_wait_slow(uint64_t* bar, uint32_t parity) {
#pragma unroll 1
  for (uint32_t spin = 0; spin < (1u << 24); ++spin)
    if (mbar_try_wait(bar, parity)) return;
  __trap();
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  if (!mbar_try_wait(bar, parity)) mbar_wait_slow(bar, parity);
}
// global -> shared bulk copy (SASS UBLKCP); bytes % 16 == 0, both addresses 16-byte aligned.
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(smem_u32(dst_smem)), "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

// ---- RoPE rotation (neox halves), one definition so that every kernel rounds identically -----------
// (left to the compiler, x1*c - x2*s may or may not be contracted into an FMA, per call site)
__device__ __forceinline__ void rope_rotate(float x1, float x2, float c, float s, float& o1, float& o2) {
  o1 = __fmaf_rn(x1, c, -__fmul_rn(x2, s));
  o2 = __fmaf_rn(x2, c, __fmul_rn(x1, s));
}

// ---- paged KV cache row layout ----------------------------------------------------------------
// A cached row (one token, one kv head) is 128 fp16 = sixteen 16-byte chunks.  Chunk c of the row at
// token offset r inside its block is stored at chunk position c ^ (r & 7): a bulk copy of the block
// then lands in shared memory already in the bank-conflict-free layout ldmatrix wants.
__host__ __device__ __forceinline__ int kv_swz(int r, int dim) { return ((((dim >> 3) ^ (r & 7)) << 3) | (dim & 7)); }

// ---- programmatic dependent launch (PDL) ------------------------------------------------------
// Every kernel of the step starts with pdl_prologue(): it lets the next kernel in the stream begin
// launching (its set-up and weight prefetch overlap our execution) and then blocks until the
// upstream kernel's results are visible.  Both are no-ops for a plain launch.
__device__ __forceinline__ void grid_dep_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void grid_dep_launch() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_prologue() { grid_dep_launch(); grid_dep_wait(); }

// ---- in-situ timeline (profiling aid, off unless b200_debug_timeline() installed a buffer) --------
// Block (0,0,0) thread 0 of every kernel records %globaltimer (32 ns ticks on B200) at entry, after
// the PDL dependency wait, and at exit: 4 x uint64 per launch {kernel id, entry, ready, exit}.
struct Timeline { unsigned long long* buf; unsigned int* ctr; unsigned int cap; };
// __constant__ and tested last: the check sits at the top of every kernel, and as a global-memory load
// executed by every thread it delayed each kernel by an L2 round trip (7 % of the gate_up GEMM's stall samples)
static __constant__ Timeline d_tl;               // one copy per translation unit, set by its tl_set_*()
#define B200_DEFINE_TL_SETTER(name)                                                         \
  void name(unsigned long long* buf, unsigned int* ctr, unsigned int cap) {               \
    Timeline t{buf, ctr, cap};                                                              \
    cudaMemcpyToSymbol(d_tl, &t, sizeof(t));                                                \
  }
__device__ __forceinline__ unsigned long long gtime() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
enum KernelId { kKidEmbed = 1, kKidRmsnorm, kKidRope, kKidAttn, kKidSwiglu, kKidW4Gemv, kKidF16Gemv, kKidUmmaW4,
                kKidUmmaF16, kKidSample, kKidAdvance, kKidTpTick };
// PDL prologue + timeline entry; returns the timeline slot (or ~0u).
// kWait = false: the kernel places griddepcontrol.wait itself (after work that does not depend on upstream).
template <bool kWait = true>
__device__ __forceinline__ unsigned kernel_prologue(int kid) {
  unsigned slot = ~0u;
  const bool rec = threadIdx.x == 0 && (blockIdx.x | blockIdx.y | blockIdx.z) == 0 && d_tl.buf;
  unsigned long long t0 = 0;
  if (rec) t0 = gtime();
  grid_dep_launch();
  if constexpr (kWait) grid_dep_wait();
  if (rec) {
    slot = atomicAdd(d_tl.ctr, 1u);
    if (slot < d_tl.cap) {
      d_tl.buf[4 * slot] = (unsigned long long)kid | ((unsigned long long)gridDim.x << 8) | ((unsigned long long)gridDim.y << 28) | ((unsigned long long)gridDim.z << 48);
      d_tl.buf[4 * slot + 1] = t0;
      d_tl.buf[4 * slot + 2] = gtime();
    } else slot = ~0u;
  }
  return slot;
}
__device__ __forceinline__ void kernel_epilogue(unsigned slot) {
  if (slot != ~0u) d_tl.buf[4 * slot + 3] = gtime();
}

}  // namespace b200
