// tcgen05 / TMEM / TMA building blocks shared by the GEMM kernels (ops_umma.cu) and the persistent
// decode kernel (ops_mega.cu): PTX wrappers, descriptors, the W4 chunk -> TMEM producer and the
// accumulator epilogue.  See ops_umma.cu for the design.
#pragma once
#include <cuda.h>

#include "common.cuh"
#include "internal.h"

namespace b200 {

constexpr int kUmmaM = 128;          // weight rows per CTA tile
constexpr int kUmmaBK = 64;          // k per pipeline stage = one 128-byte swizzle row of fp16
constexpr int kProducerWarps = 16;   // four warpgroups; warpgroup g owns A stage g and k tiles it = g (mod 4)
constexpr int kWarpTmaW = kProducerWarps;        // weight-chunk TMA warp (W4 only)
constexpr int kWarpTmaX = kProducerWarps + 1;    // activation (and fp16 weight) TMA warp
constexpr int kWarpMma = kProducerWarps + 2;     // MMA issuer + TMEM owner
constexpr int kUmmaThreads = (kProducerWarps + 3) * 32;
constexpr int kChunkStride = 512 + 2 * 4096;     // ring slot: scales 256 B + zeros 256 B + up to two 4 KB packed k tiles
constexpr int kMaxStages = 8, kMaxRing = 12;
constexpr int kTileABytes = kUmmaM * kUmmaBK * 2;          // 16 KB

enum UmmaMode : int { kStoreF32 = 0, kAccumF32 = 1, kPartialF32 = 2, kSwigluF16 = 3, kArgmaxF32 = 4 };


// ---- PTX wrappers -------------------------------------------------------------------------------
__device__ __forceinline__ void tma_load_2d(void* dst_smem, const CUtensorMap* map, int x, int y, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
      ::"r"(smem_u32(dst_smem)), "l"(map), "r"(x), "r"(y), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tmem_alloc(uint32_t* dst_smem, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(dst_smem)), "r"(ncols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void umma_f16(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(acc) : "memory");
}
// A operand from tensor memory (row i = TMEM lane i, 16-bit elements packed two per 32-bit column).
__device__ __forceinline__ void umma_f16_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
      ::"r"(d_tmem), "r"(a_tmem), "l"(bdesc), "r"(idesc), "r"(acc) : "memory");
}
// registers -> TMEM: 32 consecutive 32-bit columns of this thread's lane.
__device__ __forceinline__ void tmem_st32(uint32_t taddr, const uint32_t* v) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,"
               "%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31,%32};"
               ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]),
                 "r"(v[8]), "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]),
                 "r"(v[16]), "r"(v[17]), "r"(v[18]), "r"(v[19]), "r"(v[20]), "r"(v[21]), "r"(v[22]), "r"(v[23]),
                 "r"(v[24]), "r"(v[25]), "r"(v[26]), "r"(v[27]), "r"(v[28]), "r"(v[29]), "r"(v[30]), "r"(v[31])
               : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
// TMEM -> registers, 32 lanes x 32 bit, CW consecutive columns per thread (CW in {4, 8, 16, 32}).
template <int CW>
__device__ __forceinline__ void tmem_ld(uint32_t taddr, uint32_t* v) {
  if constexpr (CW == 4) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]) : "r"(taddr) : "memory");
  } else if constexpr (CW == 8) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7])
                 : "r"(taddr) : "memory");
  } else if constexpr (CW == 16) {
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                   "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
                 : "r"(taddr) : "memory");
  } else {
    static_assert(CW == 32, "tmem_ld width");
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"
                 "%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
                   "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
                   "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
                   "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
                 : "r"(taddr) : "memory");
  }
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void prefetch_tensormap(const CUtensorMap* m) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(m) : "memory");
}

__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "elect.sync _|p, 0xffffffff;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(pred));
  return pred != 0;
}

// K-major SWIZZLE_128B shared-memory matrix descriptor (cute::UMMA::SmemDescriptor bit layout):
// start address >> 4 [0,14), LBO >> 4 [16,30) (unused for swizzled K-major), SBO >> 4 [32,46) =
// 1024 B between 8-row groups, version 1 [46,48), layout type 2 = SWIZZLE_128B [61,64).
__device__ __forceinline__ uint64_t make_smem_desc(uint32_t saddr) {
  return (uint64_t)((saddr >> 4) & 0x3FFFu) | ((uint64_t)1 << 16) | ((uint64_t)(1024 >> 4) << 32) | ((uint64_t)1 << 46) |
         ((uint64_t)2 << 61);
}
// kind::f16 instruction descriptor: D fp32 [4,6)=1, A/B fp16 (0), both K-major, N>>3 [17,23), M>>4 [24,29).
__host__ __device__ constexpr uint32_t make_idesc(int M, int N) {
  return (1u << 4) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

struct UmmaSmem {                      // barriers live after the tile rings
  uint64_t full[kMaxStages], empty[kMaxStages];       // one MMA stage = one quantisation group (W4) / k tile (fp16)
  uint64_t pk_full[kMaxRing], pk_empty[kMaxRing];     // packed int4 chunks
  uint64_t tmem_full;
  uint32_t tmem_base;
};

// A pipeline stage holds everything the MMA thread needs for `kTilesPerStage` k tiles, behind ONE
// mbarrier (an mbarrier wait costs ~200 cycles on the single MMA-issuing thread even when the
// phase is already complete, so waits per k tile are what bound a weight-streaming GEMM):
//   W4   : A = fp16 weights in TMEM (64 columns per stage, written by the dequant warpgroups),
//          X = kTilesPerStage activation tiles in smem (TMA).  Stage = one AWQ group (G = 128: 2 k tiles).
//   fp16 : A = weight tile in smem (TMA), X = activation tile in smem (TMA).  Stage = 1 k tile.
// kDeep (W4, token tile <= 64): seven stages instead of four, which is what the 512 TMEM columns hold next
// to the accumulator.  The producer warpgroups dequantise ahead of the dependency wait, so a CTA with up
// to seven groups (down: 6, qkv / o: 3) has its whole A operand in tensor memory before the activations
// exist, and gate_up (12) has seven of twelve; seven activation stages (112 KB) are requested at once.
// Warpgroup w still takes groups w, w + 4, ...; stage = group % 7 is then shared between warpgroups, which
// is safe for the parity waits because a warpgroup's previous group g - 4 already waited for group
// g - 4 - 7 to be consumed, so the barrier of stage g % 7 is never more than one phase behind.
// Dense MT = 2 (token tile 256): a CTA multiplies 256 weight rows by 256 tokens -- 128 flop per byte through L2 instead
// of 85 for a 128 x 256 tile; a large dense GEMM on 148 SMs is bound by L2->SM bandwidth before it is bound by the
// tensor pipe (26 GB of operand traffic for the 7B gate_up GEMM at 8192 tokens with 128-row tiles).  Three stages of
// {A0, A1, B} = 64 KB; two 256-column accumulators fill the tensor memory.
// MT (W4 prefill, token tile 128): weight-row tiles per CTA.  With MT = 2 a CTA multiplies 256 weight rows by
// the activation tile it stages, which halves the activation traffic through L2 -- the bound of the prefill
// GEMM (6.97 GB of L2->SM reads for one gate_up at 12 800 tokens, profiles/r01_kernels_ncu_full.txt).  Tensor
// memory then holds two accumulators and 2 stages x 2 tiles of A; warpgroup w still owns A slot w.
template <int NT, bool kW4, bool kDeep = true, int MT = 1>
struct UmmaCfg {
  static constexpr int kMaxTilesPerStage = kW4 ? 2 : 1;
  // dense, kDeep (token tile <= 64, a grid of at most one CTA per SM -- decode passes of 16-bit models): eight stages of
  // {weight tile, activation tile} = 192 KB in flight per SM.  A weight-streaming CTA is latency-bound: with four
  // stages (96 KB) the 14B shapes pulled 20 GB/s per SM, half of an SM's share of HBM.  Grids of several waves keep
  // four stages: two CTAs per SM then overlap one's epilogue with the other's stream (lm_head: 84 % of HBM).
  static constexpr int kStages = MT == 2 ? (kW4 ? 2 : 3) : (kW4 && kDeep && NT <= 64) ? 7 : (!kW4 && kDeep && NT <= 64) ? 8 : 4;
  // private ring slots per producer warpgroup.  3 (all 12 chunks of a gate_up CTA requested at once) was
  // measured: no change -- the kernel's stall samples sit in the wait for the first chunk (HBM latency) and
  // in the wait for the accumulator (12 groups x 256 tensor cycles), not in ring refills.
  static constexpr int kSlotsPerWg = 2;
  static constexpr int kRing = kW4 ? 4 * kSlotsPerWg : 0;
  static constexpr int kTileBBytes = NT * kUmmaBK * 2;
  static constexpr int kStageBytes = (kW4 ? 0 : MT * kTileABytes) + kMaxTilesPerStage * kTileBBytes;
  static constexpr int kAccCols = MT * (NT < 32 ? 32 : NT);  // TMEM: accumulator(s), then (W4) the A slots, 64 columns each
  static constexpr int kTmemNeed = kAccCols + (kW4 ? kStages * MT * 64 : 0);
  static constexpr int kTmemCols = kTmemNeed <= 32 ? 32 : kTmemNeed <= 64 ? 64 : kTmemNeed <= 128 ? 128 : kTmemNeed <= 256 ? 256 : 512;
  static constexpr int kRingOff = kStages * kStageBytes;
  static constexpr int kBarOff = kRingOff + kRing * kChunkStride;
  static constexpr int kSmemBytes = kBarOff + (int)sizeof(UmmaSmem) + 1024;
  static_assert(MT == 1 || (MT == 2 && kW4 && NT == 128) || (MT == 2 && !kW4 && NT == 256),
                "two row tiles per CTA: W4 with token tile 128, dense with token tile 256");
  static_assert(kStages * MT >= 4 && kStages <= kMaxStages && kRing <= kMaxRing, "ring sizes");
  static_assert(kTmemNeed <= 512, "tensor memory budget");
  static_assert(kSmemBytes <= 227 * 1024, "shared memory budget");
};

struct UmmaParams {
  unsigned long long* trace;
  // W4 source: chunk layout, see umma_tile_w4_kernel
  const unsigned char* chunks;   // [N/128][K/G] chunks of (512 + (G/64) * 4096) bytes
  const __half* bias;            // [N] or null (kStoreF32 only)
  float* y;               // output, row stride ldy; kPartialF32: + split * part_stride
  int64_t part_stride;
  int ldy, T, K, N, G;
  int kt_per_split;       // k tiles handled by one CTA along grid.z (W4: a multiple of G/64)
  int mode;
  int bf16;               // dense path: weights and activations are bf16 (0: fp16).  The W4 path is fp16 only.
  const unsigned char* l2_prefetch;   // weights of the NEXT kernel to pull into L2 (or null), shared out over the CTAs
  int64_t l2_prefetch_bytes;
  // CTA rasterisation.  0: blockIdx.x = weight tile is the fastest index, so a wave of 148 CTAs walks all weight tiles
  // against a few token tiles.  1: token tiles fastest -- a wave multiplies every token tile by a few weight tiles, so
  // the activations stay in L2 and the weights come from HBM once.  Chosen by the launcher for large passes whose
  // activations fit L2 and whose weights do not (7B gate_up at 8192 tokens, ncu: 8.75 GB of DRAM reads for 330 MB of
  // operands with order 0 -- the 271 MB of weights were re-read by each of the 32 waves).
  int raster;
};

__device__ __forceinline__ uint4 lds128(uint32_t saddr) {
  uint4 r;
  asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "r"(saddr));
  return r;
}
__device__ __forceinline__ uint32_t lds_u16(uint32_t saddr) {
  uint16_t r;
  asm volatile("ld.shared.u16 %0, [%1];" : "=h"(r) : "r"(saddr));
  return r;
}

// One producer warpgroup (thread r = weight row of the tile) turns one packed chunk sitting in ring
// slot `ch` into the fp16 A operand of one MMA stage in tensor memory:
//   wait pk_full -> read scales / zeros / packed words -> dequantise in registers -> wait until the
//   stage's previous MMAs retired (empty) -> tcgen05.st -> arrive on full -> release the ring slot.
// tmem_dst: TMEM address of the stage for this thread's lane quarter (64 columns: two k tiles x 32).
__device__ __forceinline__ void w4_chunk_to_tmem(const uint32_t ch, uint64_t* pk_full, const uint32_t pk_full_parity,
                                                 uint64_t* empty, const uint32_t empty_parity, uint64_t* full,
                                                 uint64_t* pk_empty, const uint32_t tmem_dst, const int r,
                                                 const int nt_here) {
  mbar_wait(pk_full, pk_full_parity);
  const uint32_t sh = lds_u16(ch + r * 2), zh = lds_u16(ch + 256 + r * 2);
  const uint32_t su = sh | (sh << 16), zu = zh | (zh << 16);
  for (int t = 0; t < nt_here; ++t) {
    const uint4 w0 = lds128(ch + 512 + (t * 2 + 0) * 2048 + r * 16);
    const uint4 w1 = lds128(ch + 512 + (t * 2 + 1) * 2048 + r * 16);
    uint32_t d[32];
    dequant8(w0.x, zu, su, d + 0);
    dequant8(w0.y, zu, su, d + 4);
    dequant8(w0.z, zu, su, d + 8);
    dequant8(w0.w, zu, su, d + 12);
    dequant8(w1.x, zu, su, d + 16);
    dequant8(w1.y, zu, su, d + 20);
    dequant8(w1.z, zu, su, d + 24);
    dequant8(w1.w, zu, su, d + 28);
    if (t == 0) {
      mbar_wait(empty, empty_parity);
      tc_fence_after();
    }
    tmem_st32(tmem_dst + (uint32_t)(32 * t), d);   // row r: 64 fp16 = 32 columns
  }
  tmem_st_wait();
  tc_fence_before();
  mbar_arrive(full);
  // Release the ring slot only now: the TMEM stores above depend on the values read from it,
  // so every read of the chunk has completed before the TMA warp may overwrite it.
  mbar_arrive(pk_empty);
}

// Epilogue of one [NT tokens x 128 features] accumulator tile, executed by the kProducerWarps * 32
// producer/epilogue threads (threadIdx.x < 512) once the accumulator is complete (tmem_full waited,
// tcgen05.fence::after_thread_sync issued).  `stage`: NT * 512 bytes of shared memory that nothing
// else touches until the function returns.  part_off: element offset of this split's partial tensor.
template <int NT>
__device__ __forceinline__ void umma_epilogue(const int mode, float* y, const int ldy, const size_t part_off,
                                              const __half* bias, const int T, const int tok0, const int nt,
                                              float* stage, const uint32_t tmem, unsigned long long* tr,
                                              const bool bf = false) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // Stage the whole [NT tokens x 128 features] fp32 accumulator tile in shared memory (the
    // activation stages are idle once the accumulator is complete), then write global memory with
    // 16-byte vectors: a TMEM lane holds one feature for many tokens, and storing it directly costs
    // one 4-byte store instruction per (token, feature) -- the LSU issue rate then dominates.
    const int q = warp & 3;                                          // TMEM lane quarter this warp may read
    constexpr int kColsPerWarp = NT / 4;                             // 16 warps: 4 lane quarters x 4 column groups
    constexpr int CW = kColsPerWarp > 32 ? 32 : kColsPerWarp;
    const int col0 = (warp >> 2) * kColsPerWarp;
#pragma unroll 1
    for (int c = 0; c < kColsPerWarp; c += CW) {
      uint32_t v[CW];
      tmem_ld<CW>(tmem + ((uint32_t)(q * 32) << 16) + (uint32_t)(col0 + c), v);
#pragma unroll
      for (int j = 0; j < CW; ++j) stage[(col0 + c + j) * 128 + q * 32 + lane] = __uint_as_float(v[j]);
    }
    if (tr && threadIdx.x == 0) tr[5] = clock64();                   // accumulator staged in shared memory
    asm volatile("bar.sync 1, %0;" ::"n"(kProducerWarps * 32) : "memory");
    const int tid = threadIdx.x;                                     // 0..511
    if (mode == kSwigluF16) {
      // Fused SwiGLU (vLLM qwen2.py:106-111 SiluAndMul): rows 0..63 of the tile are gate features,
      // rows 64..127 the matching up features.  One thread = one token x 8 features = one 16-byte store.
      __half* act = reinterpret_cast<__half*>(y);
      for (int idx = tid; idx < NT * 8; idx += kProducerWarps * 32) {
        const int bcol = idx >> 3, f8 = (idx & 7) * 8;
        const int b = tok0 + bcol;
        if (b < T) {
          const float4 g0 = *reinterpret_cast<const float4*>(stage + bcol * 128 + f8);
          const float4 g1 = *reinterpret_cast<const float4*>(stage + bcol * 128 + f8 + 4);
          const float4 u0 = *reinterpret_cast<const float4*>(stage + bcol * 128 + 64 + f8);
          const float4 u1 = *reinterpret_cast<const float4*>(stage + bcol * 128 + 64 + f8 + 4);
          auto f = [](float g, float u) { return g / (1.f + expf(-g)) * u; };
          uint4 o;
          o.x = pack16x2(f(g0.x, u0.x), f(g0.y, u0.y), bf);
          o.y = pack16x2(f(g0.z, u0.z), f(g0.w, u0.w), bf);
          o.z = pack16x2(f(g1.x, u1.x), f(g1.y, u1.y), bf);
          o.w = pack16x2(f(g1.z, u1.z), f(g1.w, u1.w), bf);
          *reinterpret_cast<uint4*>(act + (size_t)b * ldy + nt * 64 + f8) = o;
        }
      }
    } else if (mode == kArgmaxF32) {
      // Greedy lm_head: per token the maximum of this tile's 128 features and its lowest index, as
      // float2 {value, bits(row index)} at y[tile * ldy + token]; finished by argmax_reduce_kernel.
      // The [T, V] fp32 logits (39 MB at batch 64) are never written.
      float2* part = reinterpret_cast<float2*>(y);
      for (int bcol = warp; bcol < NT && tok0 + bcol < T; bcol += kProducerWarps) {
        const float4 a = *reinterpret_cast<const float4*>(stage + bcol * 128 + lane * 4);
        float m = a.x;
        int mi = lane * 4;
        if (a.y > m) { m = a.y; mi = lane * 4 + 1; }
        if (a.z > m) { m = a.z; mi = lane * 4 + 2; }
        if (a.w > m) { m = a.w; mi = lane * 4 + 3; }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
          const float om = __shfl_xor_sync(0xffffffffu, m, o);
          const int oi = __shfl_xor_sync(0xffffffffu, mi, o);
          if (om > m || (om == m && oi < mi)) { m = om; mi = oi; }
        }
        if (lane == 0) part[(size_t)nt * ldy + tok0 + bcol] = make_float2(m, __int_as_float(nt * kUmmaM + mi));
      }
    } else {
      float* ybase = y + part_off + (size_t)nt * kUmmaM;
      if (mode == kAccumF32) grid_dep_wait();                      // y is read-modify-written
      for (int idx = tid; idx < NT * 32; idx += kProducerWarps * 32) {
        const int bcol = idx >> 5, f4 = (idx & 31) * 4;
        const int b = tok0 + bcol;
        if (b < T) {
          float4 a = *reinterpret_cast<const float4*>(stage + bcol * 128 + f4);
          float4* yp = reinterpret_cast<float4*>(ybase + (size_t)b * ldy + f4);
          if (mode == kAccumF32) {
            const float4 o = *yp;
            a.x += o.x; a.y += o.y; a.z += o.z; a.w += o.w;
          } else if (mode == kStoreF32 && bias) {
            const uint2 bw = *reinterpret_cast<const uint2*>(bias + nt * kUmmaM + f4);
            const float2 b0 = unpack16x2(bw.x, bf);
            const float2 b1 = unpack16x2(bw.y, bf);
            a.x += b0.x; a.y += b0.y; a.z += b1.x; a.w += b1.y;
          }
          *yp = a;
        }
      }
    }
}

}  // namespace b200
