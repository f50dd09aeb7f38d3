// tcgen05 / TMEM GEMM for the token-batched regime (decode batch >= 16 and prefill), swap-AB:
//
//     D[n, b] (TMEM, fp32)  +=  W[n, k] (MMA "A", 128 weight rows)  x  X[b, k] (MMA "B", NT tokens)
//
// The weight matrix is the M side so that a 16..256-token batch fills the MMA N dimension and no
// tensor-core work is wasted on padding; weights stream through exactly once per token tile.
//
//   * AWQ int4 weights never touch shared memory as fp16: the packed words of one (128-row tile, quantisation group) plus their
//     scales and zero points are one contiguous chunk in HBM; a TMA bulk copy (cp.async.bulk) ring
//     keeps ~50 KB of them in flight per SM, so HBM latency never reaches the math warps.  8
//     producer warps read the packed words from shared memory, dequantise in registers (lop3
//     magic-number trick, exact fp16 (q-z)*s) and write the fp16 tile straight into the canonical
//     K-major SWIZZLE_128B layout that tcgen05.mma consumes (fence.proxy.async, mbarrier arrive).
//   * dense fp16 weights (tied lm_head): TMA (cp.async.bulk.tensor.2d, SWIZZLE_128B) feeds the same
//     layout.
//   * activations: always TMA; rows beyond T are zero-filled by the TMA unit.
//   * one elected thread issues tcgen05.mma (M=128, N=NT, K=16) x4 per 64-wide k tile;
//     tcgen05.commit releases the smem stage and, after the last k tile, publishes the accumulator.
//   * epilogue: tcgen05.ld (32 lanes x 32 bit) -> registers -> fused store / bias / residual add /
//     split-K partial; 32 lanes write 32 consecutive output features (coalesced).
//
// Replaces vLLM's awq_gemm / awq_dequantize+matmul (awq.py:262-286) and the cuBLAS lm_head GEMM.
#include <cuda.h>

#include <mutex>

#include "common.cuh"
#include "internal.h"

namespace b200 {

B200_DEFINE_TL_SETTER(tl_set_umma)

constexpr int kUmmaM = 128;          // weight rows per CTA tile
constexpr int kUmmaBK = 64;          // k per pipeline stage = one 128-byte swizzle row of fp16
constexpr int kProducerWarps = 16;   // four warpgroups; warpgroup g owns A stage g and k tiles it = g (mod 4)
constexpr int kWarpTmaW = kProducerWarps;        // weight-chunk TMA warp (W4 only)
constexpr int kWarpTmaX = kProducerWarps + 1;    // activation (and fp16 weight) TMA warp
constexpr int kWarpMma = kProducerWarps + 2;     // MMA issuer + TMEM owner
constexpr int kUmmaThreads = (kProducerWarps + 3) * 32;
constexpr int kChunkStride = 512 + 2 * 4096;     // ring slot: scales 256 B + zeros 256 B + up to two 4 KB packed k tiles
constexpr int kMaxStages = 8, kMaxRing = 8;
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
template <int NT, bool kW4>
struct UmmaCfg {
  static constexpr int kMaxTilesPerStage = kW4 ? 2 : 1;
  static constexpr int kStages = 4;
  static constexpr int kRing = kW4 ? 8 : 0;           // two private slots per producer warpgroup
  static constexpr int kTileBBytes = NT * kUmmaBK * 2;
  static constexpr int kStageBytes = (kW4 ? 0 : kTileABytes) + kMaxTilesPerStage * kTileBBytes;
  static constexpr int kAccCols = NT < 32 ? 32 : NT;  // TMEM: accumulator, then (W4) the A stages, 64 columns each
  static constexpr int kTmemNeed = kAccCols + (kW4 ? kStages * 64 : 0);
  static constexpr int kTmemCols = kTmemNeed <= 32 ? 32 : kTmemNeed <= 64 ? 64 : kTmemNeed <= 128 ? 128 : kTmemNeed <= 256 ? 256 : 512;
  static constexpr int kRingOff = kStages * kStageBytes;
  static constexpr int kBarOff = kRingOff + kRing * kChunkStride;
  static constexpr int kSmemBytes = kBarOff + (int)sizeof(UmmaSmem) + 1024;
  static_assert(kStages == 4 && kStages <= kMaxStages && kRing <= kMaxRing, "ring sizes (stage s is owned by warpgroup s)");
  static_assert(kTmemNeed <= 512, "tensor memory budget");
  static_assert(kSmemBytes <= 227 * 1024, "shared memory budget");
};

// Optional per-CTA cycle trace (debug/profiling aid): 8 x uint64 per CTA, written when non-null.
static unsigned long long* g_umma_trace = nullptr;
static int g_umma_trace_gridx = 0, g_umma_trace_gridz = 0;       // 0 = trace every launch, else only matching grids
void umma_set_trace(unsigned long long* dev_buf, int grid_x, int grid_z) {
  g_umma_trace = dev_buf; g_umma_trace_gridx = grid_x; g_umma_trace_gridz = grid_z;
}

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

template <int NT, bool kW4>
__global__ void __launch_bounds__(kUmmaThreads, 1)
umma_gemm_kernel(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_w,
                 const UmmaParams p) {
  using Cfg = UmmaCfg<NT, kW4>;
  constexpr int kStages = Cfg::kStages, kRing = Cfg::kRing;
  constexpr int kTileBBytes = Cfg::kTileBBytes, kStageBytes = Cfg::kStageBytes;
  constexpr int kXInStage = kW4 ? 0 : kTileABytes;           // byte offset of the first X tile inside a stage
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  // SWIZZLE_128B needs 1024-byte alignment; the dynamic smem base is only guaranteed 16
  unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  unsigned char* ring = smem + Cfg::kRingOff;
  UmmaSmem* sb = reinterpret_cast<UmmaSmem*>(smem + Cfg::kBarOff);
  constexpr uint32_t kTmemCols = Cfg::kTmemCols;

  // PDL: the next kernel may start its own set-up / weight prefetch now (no dependency wait here:
  // the weight stream below does not depend on the upstream kernel)
  unsigned tl = ~0u;
  unsigned long long tl_t0 = 0;
  const bool tl_rec = d_tl.buf && threadIdx.x == 0 && (blockIdx.x | blockIdx.y | blockIdx.z) == 0;
  if (tl_rec) tl_t0 = gtime();
  grid_dep_launch();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nt = blockIdx.x;                 // weight-row tile
  const int tok0 = blockIdx.y * NT;          // token tile
  const int split = blockIdx.z;
  const int KT = p.K / kUmmaBK;
  const int kt0 = split * p.kt_per_split;
  const int nkt = min(p.kt_per_split, KT - kt0);
  const int tps = kW4 ? p.G / kUmmaBK : 1;   // k tiles per stage (W4: per quantisation group, 1 or 2)
  const int nst = (nkt + tps - 1) / tps;     // stages (= groups) this CTA walks through

  unsigned long long* tr = p.trace ? p.trace + 8 * (size_t)(blockIdx.x + gridDim.x * (blockIdx.y + gridDim.y * blockIdx.z)) : nullptr;
  if (tr && threadIdx.x == 0) tr[0] = clock64();
  if (threadIdx.x == 0) {
    for (int s = 0; s < kStages; ++s) { mbar_init(&sb->full[s], kW4 ? 128 + 1 : 1); mbar_init(&sb->empty[s], 1); }
    for (int s = 0; s < kRing; ++s) { mbar_init(&sb->pk_full[s], 1); mbar_init(&sb->pk_empty[s], 128); }
    mbar_init(&sb->tmem_full, 1);
    fence_mbar_init();
  }
  if (warp == kWarpMma) tmem_alloc(&sb->tmem_base, kTmemCols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = sb->tmem_base;
  if (tr && threadIdx.x == 0) tr[1] = clock64();

  if (warp < kProducerWarps) {
    // ===================== weight producers (W4 only): smem ring -> dequant -> TMEM =====================
    if (kW4) {
      // Warpgroup wg dequantises groups gi = wg, wg + 4, ... ; thread = weight row.  The fp16 tile goes
      // registers -> TMEM (tcgen05.st) and is consumed as the MMA A operand from there: no
      // shared-memory write, no proxy fence, and the MMA reads only the activation tile from shared
      // memory (an SS-mode M=128 MMA alone would saturate the 128 B/clk smem port).  Four groups are
      // in flight per SM, which hides the mbarrier and TMEM-store round trips.
      const int r = threadIdx.x & 127, wg = warp >> 2;
      const uint32_t lane_base = (uint32_t)((warp & 3) * 32) << 16;
      // Ring slots and MMA stages are private to a warpgroup (slot = wg + 4 * (use % 2), stage = wg):
      // a thread only ever waits on a barrier whose previous phase it has itself consumed.  (With
      // slots shared between warpgroups a fast warpgroup could wait on a slot two phases ahead, and
      // an mbarrier parity wait cannot tell "two phases ahead" from "done".)
      int use = 0;
      for (int gi = wg; gi < nst; gi += 4, ++use) {
        const int slot = wg + 4 * (use & 1), s = gi % kStages;
        mbar_wait(&sb->pk_full[slot], (uint32_t)(use >> 1) & 1u);
        const uint32_t ch = smem_u32(ring + slot * kChunkStride);
        const uint32_t sh = lds_u16(ch + r * 2), zh = lds_u16(ch + 256 + r * 2);
        const uint32_t su = sh | (sh << 16), zu = zh | (zh << 16);
        const int nt_here = min(tps, nkt - gi * tps);
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
            mbar_wait(&sb->empty[s], ((uint32_t)(gi / kStages) & 1u) ^ 1u);
            tc_fence_after();
          }
          tmem_st32(tmem + lane_base + (uint32_t)(Cfg::kAccCols + 64 * s + 32 * t), d);   // row r: 64 fp16 = 32 columns
        }
        tmem_st_wait();
        tc_fence_before();
        mbar_arrive(&sb->full[s]);
        // Release the ring slot only now: the TMEM stores above depend on the values read from it,
        // so every read of the chunk has completed before the TMA warp may overwrite it.
        mbar_arrive(&sb->pk_empty[slot]);
      }
    }
    // ===================== epilogue: TMEM -> registers -> global =====================
    mbar_wait(&sb->tmem_full, 0);
    tc_fence_after();
    if (tr && threadIdx.x == 0) tr[4] = clock64();
    {
      // Stage the whole [NT tokens x 128 features] fp32 accumulator tile in shared memory (the
      // activation stages are idle once the accumulator is complete), then write global memory with
      // 16-byte vectors: a TMEM lane holds one feature for many tokens, and storing it directly costs
      // one 4-byte store instruction per (token, feature) -- the LSU issue rate then dominates.
      const int q = warp & 3;                                          // TMEM lane quarter this warp may read
      constexpr int kColsPerWarp = NT / 4;                             // 16 warps: 4 lane quarters x 4 column groups
      constexpr int CW = kColsPerWarp > 32 ? 32 : kColsPerWarp;
      const int col0 = (warp >> 2) * kColsPerWarp;
      float* stage = reinterpret_cast<float*>(smem);                   // [NT][128] fp32 = NT * 512 B <= the X ring
      static_assert(NT * 512 <= Cfg::kStages * Cfg::kStageBytes, "epilogue staging fits in the stage ring");
#pragma unroll 1
      for (int c = 0; c < kColsPerWarp; c += CW) {
        uint32_t v[CW];
        tmem_ld<CW>(tmem + ((uint32_t)(q * 32) << 16) + (uint32_t)(col0 + c), v);
#pragma unroll
        for (int j = 0; j < CW; ++j) stage[(col0 + c + j) * 128 + q * 32 + lane] = __uint_as_float(v[j]);
      }
      if (tr && threadIdx.x == 0) tr[2] = clock64();                   // (reuses the "first stage ready" slot)
      asm volatile("bar.sync 1, %0;" ::"n"(kProducerWarps * 32) : "memory");
      const int tid = threadIdx.x;                                     // 0..511
      if (p.mode == kSwigluF16) {
        // Fused SwiGLU (vLLM qwen2.py:106-111 SiluAndMul): rows 0..63 of the tile are gate features,
        // rows 64..127 the matching up features.  One thread = one token x 8 features = one 16-byte store.
        __half* act = reinterpret_cast<__half*>(p.y);
        for (int idx = tid; idx < NT * 8; idx += kProducerWarps * 32) {
          const int bcol = idx >> 3, f8 = (idx & 7) * 8;
          const int b = tok0 + bcol;
          if (b < p.T) {
            const float4 g0 = *reinterpret_cast<const float4*>(stage + bcol * 128 + f8);
            const float4 g1 = *reinterpret_cast<const float4*>(stage + bcol * 128 + f8 + 4);
            const float4 u0 = *reinterpret_cast<const float4*>(stage + bcol * 128 + 64 + f8);
            const float4 u1 = *reinterpret_cast<const float4*>(stage + bcol * 128 + 64 + f8 + 4);
            auto f = [](float g, float u) { return g / (1.f + expf(-g)) * u; };
            uint4 o;
            o.x = pack_half2(f(g0.x, u0.x), f(g0.y, u0.y));
            o.y = pack_half2(f(g0.z, u0.z), f(g0.w, u0.w));
            o.z = pack_half2(f(g1.x, u1.x), f(g1.y, u1.y));
            o.w = pack_half2(f(g1.z, u1.z), f(g1.w, u1.w));
            *reinterpret_cast<uint4*>(act + (size_t)b * p.ldy + nt * 64 + f8) = o;
          }
        }
      } else if (p.mode == kArgmaxF32) {
        // Greedy lm_head: per token the maximum of this tile's 128 features and its lowest index, as
        // float2 {value, bits(row index)} at y[tile * ldy + token]; finished by argmax_reduce_kernel.
        // The [T, V] fp32 logits (39 MB at batch 64) are never written.
        float2* part = reinterpret_cast<float2*>(p.y);
        for (int bcol = warp; bcol < NT && tok0 + bcol < p.T; bcol += kProducerWarps) {
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
          if (lane == 0) part[(size_t)nt * p.ldy + tok0 + bcol] = make_float2(m, __int_as_float(nt * kUmmaM + mi));
        }
      } else {
        float* ybase = p.y + (p.mode == kPartialF32 ? (size_t)split * p.part_stride : 0) + (size_t)nt * kUmmaM;
        if (p.mode == kAccumF32) grid_dep_wait();                      // y is read-modify-written
        for (int idx = tid; idx < NT * 32; idx += kProducerWarps * 32) {
          const int bcol = idx >> 5, f4 = (idx & 31) * 4;
          const int b = tok0 + bcol;
          if (b < p.T) {
            float4 a = *reinterpret_cast<const float4*>(stage + bcol * 128 + f4);
            float4* yp = reinterpret_cast<float4*>(ybase + (size_t)b * p.ldy + f4);
            if (p.mode == kAccumF32) {
              const float4 o = *yp;
              a.x += o.x; a.y += o.y; a.z += o.z; a.w += o.w;
            } else if (p.mode == kStoreF32 && p.bias) {
              const uint2 bw = *reinterpret_cast<const uint2*>(p.bias + nt * kUmmaM + f4);
              const float2 b0 = __half22float2(*reinterpret_cast<const __half2*>(&bw.x));
              const float2 b1 = __half22float2(*reinterpret_cast<const __half2*>(&bw.y));
              a.x += b0.x; a.y += b0.y; a.z += b1.x; a.w += b1.y;
            }
            *yp = a;
          }
        }
      }
    }
  } else if (warp == kWarpTmaW) {
    // ===================== TMA: packed weight chunks (do not depend on upstream kernels) =====================
    if (kW4 && elect_one()) {
      const int KG = p.K / p.G;
      const uint32_t chunk_bytes = 512u + (uint32_t)tps * 4096u;
      const unsigned char* src = p.chunks + ((size_t)nt * KG + kt0 / tps) * chunk_bytes;
      for (int gi = 0; gi < nst; ++gi) {
        const int use = gi >> 2;                                   // same slot walk as the consuming warpgroup
        const int slot = (gi & 3) + 4 * (use & 1);
        mbar_wait(&sb->pk_empty[slot], ((uint32_t)(use >> 1) & 1u) ^ 1u);
        mbar_arrive_expect_tx(&sb->pk_full[slot], chunk_bytes);
        bulk_g2s(ring + slot * kChunkStride, src + (size_t)gi * chunk_bytes, chunk_bytes, &sb->pk_full[slot]);
      }
    }
  } else if (warp == kWarpTmaX) {
    // ===================== TMA: activation tiles (and fp16 weight tiles) =====================
    if (elect_one()) {
      prefetch_tensormap(&map_x);
      if (!kW4) prefetch_tensormap(&map_w);
      // fp16 weights (lm_head): the first ring of weight tiles does not depend on the upstream kernel
      const int pre = kW4 ? 0 : min(nst, kStages);
      for (int gi = 0; gi < pre; ++gi) {
        mbar_arrive_expect_tx(&sb->full[gi], (uint32_t)(kXInStage + kTileBBytes));
        tma_load_2d(smem + gi * kStageBytes, &map_w, (kt0 + gi) * kUmmaBK, nt * kUmmaM, &sb->full[gi]);
      }
      grid_dep_wait();                         // activations come from the upstream kernel
      for (int gi = 0; gi < nst; ++gi) {
        const int s = gi % kStages;
        const int nt_here = min(tps, nkt - gi * tps);
        unsigned char* stage = smem + s * kStageBytes;
        if (gi >= pre) {
          mbar_wait(&sb->empty[s], ((uint32_t)(gi / kStages) & 1u) ^ 1u);
          mbar_arrive_expect_tx(&sb->full[s], (uint32_t)(kXInStage + nt_here * kTileBBytes));
          if (!kW4) tma_load_2d(stage, &map_w, (kt0 + gi) * kUmmaBK, nt * kUmmaM, &sb->full[s]);
        }
        for (int t = 0; t < nt_here; ++t)
          tma_load_2d(stage + kXInStage + t * kTileBBytes, &map_x, (kt0 + gi * tps + t) * kUmmaBK, tok0, &sb->full[s]);
      }
    }
  } else {
    // ===================== MMA issuer =====================
    if (elect_one()) {
      constexpr uint32_t idesc = make_idesc(kUmmaM, NT);
      long long tw_a = 0, t_is = 0;
      for (int gi = 0; gi < nst; ++gi) {
        const int s = gi % kStages;
        const int nt_here = min(tps, nkt - gi * tps);
        const long long c0 = tr ? clock64() : 0;
        mbar_wait(&sb->full[s], (uint32_t)(gi / kStages) & 1u);
        tc_fence_after();
        const long long c1 = tr ? clock64() : 0;
        if (tr && gi == 0) tr[2] = clock64();
        const uint32_t x_addr = smem_u32(smem + s * kStageBytes + kXInStage);
        for (int t = 0; t < nt_here; ++t) {
          const uint32_t b_addr = x_addr + t * kTileBBytes;
          if constexpr (kW4) {
            const uint32_t a_tmem = tmem + (uint32_t)(Cfg::kAccCols + 64 * s + 32 * t);
#pragma unroll
            for (int k = 0; k < kUmmaBK / 16; ++k)     // 16 k = 8 TMEM columns of A, 32 bytes of the B swizzle row
              umma_f16_ts(tmem, a_tmem + k * 8, make_smem_desc(b_addr + k * 32), idesc, (gi | t | k) != 0 ? 1u : 0u);
          } else {
            const uint32_t a_addr = smem_u32(smem + s * kStageBytes);
#pragma unroll
            for (int k = 0; k < kUmmaBK / 16; ++k)     // advance 16 elements (32 bytes) inside the 128-byte swizzle row
              umma_f16(tmem, make_smem_desc(a_addr + k * 32), make_smem_desc(b_addr + k * 32), idesc, (gi | t | k) != 0 ? 1u : 0u);
          }
        }
        umma_commit(&sb->empty[s]);            // the stage is free again when the MMAs above retire
        if (tr) { const long long c2 = clock64(); tw_a += c1 - c0; t_is += c2 - c1; }
      }
      umma_commit(&sb->tmem_full);             // accumulator complete
      if (tr) { tr[3] = clock64(); tr[7] = (unsigned long long)tw_a | ((unsigned long long)t_is << 32); }
    }
  }

  if (tr && threadIdx.x == 0) tr[5] = clock64();
  tc_fence_before();
  __syncthreads();
  if (warp == kWarpMma) tmem_dealloc(tmem, kTmemCols);
  if (tr && threadIdx.x == 0) tr[6] = clock64();
  if (tl_rec) {
    tl = atomicAdd(d_tl.ctr, 1u);
    if (tl < d_tl.cap) {
      d_tl.buf[4 * tl] = (unsigned long long)(kW4 ? kKidUmmaW4 : kKidUmmaF16) | ((unsigned long long)gridDim.x << 8) | ((unsigned long long)gridDim.y << 28) | ((unsigned long long)gridDim.z << 48);
      d_tl.buf[4 * tl + 1] = tl_t0;
      d_tl.buf[4 * tl + 2] = tl_t0;
      d_tl.buf[4 * tl + 3] = gtime();
    }
  }
}

// ---- weight tiling (one-time) ---------------------------------------------------------------------
// kernel layout A (wq [N, K/8], sc/zp [N, K/G]) -> chunk layout streamed by the TMA ring above:
//   chunk(nt, g) = [sc fp16 x128][zp fp16 x128][k tile 0: half 0 uint4 x128, half 1 uint4 x128][k tile 1 ...]
// where row r of the tile is output feature nt*128 + r and a uint4 holds 32 consecutive k of that row.
// pair_rows > 0 (fused SwiGLU, W = [gate; up] with pair_rows rows each): tile nt row r takes gate row
// 64 nt + r for r < 64 and up row 64 nt + r - 64 otherwise, so gate_i and up_i land in one tile.
__global__ void umma_tile_w4_kernel(const uint32_t* __restrict__ wq, const __half* __restrict__ sc,
                                    const __half* __restrict__ zp, int K, int N, int G, int pair_rows,
                                    unsigned char* __restrict__ chunks) {
  const int KG = K / G, tpg = G / 64;
  const size_t chunk_bytes = 512 + (size_t)tpg * 4096;
  const size_t cidx = blockIdx.x;                       // chunk index = nt * KG + g
  const int g = (int)(cidx % KG), nt = (int)(cidx / KG);
  unsigned char* ch = chunks + cidx * chunk_bytes;
  const int r = threadIdx.x;                            // 128 threads
  const size_t row = pair_rows > 0 ? (size_t)(r < 64 ? 0 : pair_rows) + (size_t)nt * 64 + (r & 63) : (size_t)nt * 128 + r;
  reinterpret_cast<__half*>(ch)[r] = sc[row * KG + g];
  reinterpret_cast<__half*>(ch + 256)[r] = zp[row * KG + g];
  for (int t = 0; t < tpg; ++t)
    for (int h = 0; h < 2; ++h) {
      const uint32_t* src = wq + row * (K / 8) + (size_t)(g * tpg + t) * 8 + h * 4;
      reinterpret_cast<uint4*>(ch + 512 + (t * 2 + h) * 2048)[r] = make_uint4(src[0], src[1], src[2], src[3]);
    }
}

size_t umma_w4_chunk_bytes(int K, int N, int G) { return (size_t)(N / 128) * (K / G) * (512 + (size_t)(G / 64) * 4096); }

int launch_umma_tile_w4(const uint32_t* wq, const __half* sc, const __half* zp, int K, int N, int G, int pair_rows,
                        unsigned char* chunks, cudaStream_t s) {
  B200_REQUIRE(N % 128 == 0 && K % G == 0 && (G == 64 || G == 128), "umma_tile_w4: N % 128, K % G, G in {64, 128}");
  B200_REQUIRE(pair_rows == 0 || (pair_rows * 2 == N && pair_rows % 64 == 0), "umma_tile_w4: pair_rows must be N / 2 and a multiple of 64");
  umma_tile_w4_kernel<<<(unsigned)((N / 128) * (K / G)), 128, 0, s>>>(wq, sc, zp, K, N, G, pair_rows, chunks);
  g_launch_count += 1;
  B200_CUDA(cudaGetLastError());
  return B200_OK;
}

// ---- tensor maps ------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* ptr = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(ptr);
  });
  return fn;
}

// fp16 row-major matrix [rows, cols] with row stride ld (elements); box = [box_rows, 64 cols], SWIZZLE_128B.
int make_tensor_map_f16(CUtensorMap* map, const void* base, int64_t rows, int64_t cols, int64_t ld, int box_rows) {
  EncodeTiledFn fn = get_encode_fn();
  if (!fn) { set_last_error("cuTensorMapEncodeTiled not available from the driver"); return B200_ERR_CUDA; }
  B200_REQUIRE(((uintptr_t)base & 15) == 0 && (ld * 2) % 16 == 0 && box_rows >= 1 && box_rows <= 256,
               "tensor map: 16-byte aligned base / row pitch, box rows <= 256");
  const cuuint64_t gdim[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  const cuuint64_t gstride[1] = {(cuuint64_t)ld * 2};
  const cuuint32_t box[2] = {(cuuint32_t)kUmmaBK, (cuuint32_t)box_rows};
  const cuuint32_t estr[2] = {1, 1};
  const CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<void*>(base), gdim, gstride, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { set_last_error("cuTensorMapEncodeTiled failed with CUresult " + std::to_string((int)r)); return B200_ERR_CUDA; }
  return B200_OK;
}

// ---- launch -------------------------------------------------------------------------------------------
template <int NT, bool kW4>
static int launch_nt(const CUtensorMap& mx, const CUtensorMap& mw, const UmmaParams& p, int n_splits, cudaStream_t s) {
  constexpr int smem_bytes = UmmaCfg<NT, kW4>::kSmemBytes;
  auto kern = umma_gemm_kernel<NT, kW4>;
  static bool configured = false;
  if (!configured) {
    B200_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
    configured = true;
  }
  dim3 grid(p.N / kUmmaM, (p.T + NT - 1) / NT, n_splits);
  UmmaParams pp = p;
  if (g_umma_trace_gridx && ((int)grid.x != g_umma_trace_gridx || (int)grid.z != g_umma_trace_gridz)) pp.trace = nullptr;
  launch_k(kern, grid, dim3(kUmmaThreads), (size_t)smem_bytes, s, mx, mw, pp);
  g_launch_count += 1;
  B200_CUDA(cudaGetLastError());
  return B200_OK;
}

template <bool kW4>
static int dispatch_nt(int nt_sel, const CUtensorMap& mx, const CUtensorMap& mw, const UmmaParams& p, int n_splits,
                       cudaStream_t s) {
  switch (nt_sel) {
    case 16: return launch_nt<16, kW4>(mx, mw, p, n_splits, s);
    case 32: return launch_nt<32, kW4>(mx, mw, p, n_splits, s);
    case 64: return launch_nt<64, kW4>(mx, mw, p, n_splits, s);
    case 128: return launch_nt<128, kW4>(mx, mw, p, n_splits, s);
    default:
      if constexpr (kW4) { set_last_error("umma_w4: token tile must be <= 128"); return B200_ERR_INVALID; }
      else return launch_nt<256, kW4>(mx, mw, p, n_splits, s);
  }
}

// Split boundaries fall on quantisation groups; the number of non-empty splits may be smaller
// than requested.
int umma_w4_effective_splits(int K, int G, int n_splits) {
  const int KG = K / G;
  const int gps = (KG + n_splits - 1) / n_splits;
  return (KG + gps - 1) / gps;
}

// Token-tile height.  The W4 kernel tops out at 128 (TMEM: 128 accumulator columns + 4 x 64 columns
// of dequantised weights); the TMA-fed fp16 kernel goes to 256.
int umma_pick_nt(int T, bool w4) { return T <= 16 ? 16 : T <= 32 ? 32 : T <= 64 ? 64 : (T <= 128 || w4) ? 128 : 256; }

// W4: y = x . W^T.  n_splits > 1 requires mode == kPartialF32 (y then holds n_splits partial tensors).
int launch_umma_w4(const UmmaW4& w, const CUtensorMap& map_x, int nt_sel, int T, float* y, int ldy, int mode,
                   int n_splits, int64_t part_stride, cudaStream_t s) {
  B200_REQUIRE(w.N % 128 == 0 && w.K % w.G == 0 && (w.G == 64 || w.G == 128), "umma_w4: N % 128, K % G, G in {64, 128}");
  B200_REQUIRE(n_splits >= 1 && (n_splits == 1 || mode == kPartialF32), "umma_w4: split-K needs partial mode");
  B200_REQUIRE(mode != kSwigluF16 || nt_sel <= 128, "umma_w4: SwiGLU epilogue needs a token tile <= 128");
  if (T == 0) return B200_OK;
  UmmaParams p{};
  p.trace = g_umma_trace;
  p.chunks = w.chunks; p.bias = w.bias; p.y = y; p.part_stride = part_stride;
  p.ldy = ldy; p.T = T; p.K = w.K; p.N = w.N; p.G = w.G; p.mode = mode;
  const int KG = w.K / w.G, tpg = w.G / kUmmaBK;
  const int groups_per_split = (KG + n_splits - 1) / n_splits;     // split boundaries fall on quantisation groups
  p.kt_per_split = groups_per_split * tpg;
  n_splits = (KG + groups_per_split - 1) / groups_per_split;
  return dispatch_nt<true>(nt_sel, map_x, map_x, p, n_splits, s);
}

int launch_umma_f16(const CUtensorMap& map_w, int K, int N, const CUtensorMap& map_x, int nt_sel, int T, float* y,
                    int ldy, int mode, cudaStream_t s) {
  B200_REQUIRE(N % 128 == 0 && K % 64 == 0, "umma_f16: N % 128, K % 64");
  if (T == 0) return B200_OK;
  UmmaParams p{};
  p.trace = g_umma_trace;
  p.y = y; p.ldy = ldy; p.T = T; p.K = K; p.N = N; p.G = 128; p.mode = mode;
  p.kt_per_split = K / kUmmaBK;
  return dispatch_nt<false>(nt_sel, map_x, map_w, p, 1, s);
}

}  // namespace b200
