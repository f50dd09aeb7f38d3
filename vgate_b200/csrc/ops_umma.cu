// tcgen05 / TMEM GEMM for the token-batched regime (decode batch >= 16 and prefill), swap-AB:
//
//     D[n, b] (TMEM, fp32)  +=  W[n, k] (MMA "A", 128 weight rows)  x  X[b, k] (MMA "B", NT tokens)
//
// The weight matrix is the M side so that a 16..256-token batch fills the MMA N dimension and no
// tensor-core work is wasted on padding; weights stream through exactly once per token tile.
//
//   * AWQ int4 weights: the packed words of one (128-row tile, quantisation group) plus their scales and zero
//     points are one contiguous chunk in HBM; one warp keeps a ring of TMA bulk copies (cp.async.bulk) of them in
//     flight.  16 producer warps (four warpgroups, thread = weight row) read the packed words from shared memory,
//     dequantise in registers (lop3 magic-number trick, exact fp16 (q - z) * s) and write the fp16 tile to TENSOR
//     MEMORY with tcgen05.st; the MMA takes its A operand from there (TS mode), so the dequantised weights never
//     pass through shared memory and the MMA reads only the activation tile from it.
//   * dense fp16 weights (tied lm_head): TMA (cp.async.bulk.tensor.2d, SWIZZLE_128B) feeds shared memory; SS mode.
//   * activations: always TMA; rows beyond T are zero-filled by the TMA unit.
//   * one elected thread issues tcgen05.mma (M=128, N=NT, K=16) x4 per 64-wide k tile; one mbarrier per stage
//     (= one quantisation group); tcgen05.commit releases the stage and, after the last one, publishes the accumulator.
//   * epilogue: tcgen05.ld -> registers -> fp32 tile staged in shared memory -> 16-byte global stores with the
//     fused store / bias / residual add / split-K partial / SwiGLU / per-tile argmax.
//
// Replaces vLLM's awq_gemm / awq_dequantize+matmul (awq.py:262-286) and the cuBLAS lm_head GEMM.
#include <mutex>

#include "umma.cuh"

namespace b200 {

B200_DEFINE_TL_SETTER(tl_set_umma)

// Optional per-CTA cycle trace (debug/profiling aid): 8 x uint64 per CTA, written when non-null.
static unsigned long long* g_umma_trace = nullptr;
static int g_umma_trace_gridx = 0, g_umma_trace_gridz = 0;       // 0 = trace every launch, else only matching grids
void umma_set_trace(unsigned long long* dev_buf, int grid_x, int grid_z) {
  g_umma_trace = dev_buf; g_umma_trace_gridx = grid_x; g_umma_trace_gridz = grid_z;
}

template <int NT, bool kW4, int MT = 1, bool kDeep = true>
__global__ void __launch_bounds__(kUmmaThreads, 1)
umma_gemm_kernel(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_w,
                 const UmmaParams p) {
  using Cfg = UmmaCfg<NT, kW4, kDeep, MT>;
  constexpr int kStages = Cfg::kStages, kRing = Cfg::kRing;
  constexpr int kTileBBytes = Cfg::kTileBBytes, kStageBytes = Cfg::kStageBytes;
  constexpr int kXInStage = kW4 ? 0 : MT * kTileABytes;      // byte offset of the first X tile inside a stage
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
  const bool tl_rec = threadIdx.x == 0 && (blockIdx.x | blockIdx.y | blockIdx.z) == 0 && d_tl.buf;
  if (tl_rec) tl_t0 = gtime();
  grid_dep_launch();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  int nt = blockIdx.x;                       // weight-row tile
  int tt = blockIdx.y;                       // token tile
  if (p.raster) {                            // token tiles fastest (see UmmaParams::raster)
    const int lin = blockIdx.x + gridDim.x * blockIdx.y;
    tt = lin % gridDim.y;
    nt = lin / gridDim.y;
  }
  const int tok0 = tt * NT;
  const int split = blockIdx.z;
  const int KT = p.K / kUmmaBK;
  const int kt0 = split * p.kt_per_split;
  const int nkt = min(p.kt_per_split, KT - kt0);
  const int tps = kW4 ? p.G / kUmmaBK : 1;   // k tiles per stage (W4: per quantisation group, 1 or 2)
  const int nst = (nkt + tps - 1) / tps;     // stages (= groups) this CTA walks through

  unsigned long long* tr = p.trace ? p.trace + 8 * (size_t)(blockIdx.x + gridDim.x * (blockIdx.y + gridDim.y * blockIdx.z)) : nullptr;
  if (tr && threadIdx.x == 0) tr[0] = clock64();
  if (threadIdx.x == 0) {
    for (int s = 0; s < kStages; ++s) { mbar_init(&sb->full[s], kW4 ? MT * 128 + 1 : 1); mbar_init(&sb->empty[s], 1); }
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
      // The chunk sequence v = gi * MT + t (group gi, row tile t) is dealt to the warpgroups round-robin; with
      // MT = 2 a stage holds both tiles of a group (A slot = (gi % 2) * 2 + t = v % 4 = wg, still private).
      int use = 0;
      for (int v = wg; v < nst * MT; v += 4, ++use) {
        const int gi = v / MT, t = v % MT;
        const int slot = wg + 4 * (use % Cfg::kSlotsPerWg), s = gi % kStages;
        w4_chunk_to_tmem(smem_u32(ring + slot * kChunkStride), &sb->pk_full[slot], (uint32_t)(use / Cfg::kSlotsPerWg) & 1u,
                         &sb->empty[s], ((uint32_t)(gi / kStages) & 1u) ^ 1u, &sb->full[s], &sb->pk_empty[slot],
                         tmem + lane_base + (uint32_t)(Cfg::kAccCols + 64 * (s * MT + t)), r, min(tps, nkt - gi * tps));
      }
    }
    // ===================== epilogue: TMEM -> registers -> global =====================
    mbar_wait(&sb->tmem_full, 0);
    tc_fence_after();
    if (tr && threadIdx.x == 0) tr[4] = clock64();
#pragma unroll
    for (int t = 0; t < MT; ++t) {
      if (t > 0) asm volatile("bar.sync 1, %0;" ::"n"(kProducerWarps * 32) : "memory");   // the staging buffer is reused
      umma_epilogue<NT>(p.mode, p.y, p.ldy, p.mode == kPartialF32 ? (size_t)split * p.part_stride : 0, p.bias, p.T, tok0,
                        nt * MT + t, reinterpret_cast<float*>(smem), tmem + (uint32_t)(t * (Cfg::kAccCols / MT)), tr, p.bf16 != 0);
    }
  } else if (warp == kWarpTmaW) {
    // ===================== TMA: packed weight chunks (do not depend on upstream kernels) =====================
    if (kW4 && elect_one()) {
      const int KG = p.K / p.G;
      const uint32_t chunk_bytes = 512u + (uint32_t)tps * 4096u;
      for (int v = 0; v < nst * MT; ++v) {
        const int gi = v / MT, t = v % MT;
        const unsigned char* src = p.chunks + ((size_t)(nt * MT + t) * KG + kt0 / tps + gi) * chunk_bytes;
        const int use = v >> 2;                                    // same slot walk as the consuming warpgroup
        const int slot = (v & 3) + 4 * (use % Cfg::kSlotsPerWg);
        mbar_wait(&sb->pk_empty[slot], ((uint32_t)(use / Cfg::kSlotsPerWg) & 1u) ^ 1u);
        mbar_arrive_expect_tx(&sb->pk_full[slot], chunk_bytes);
        bulk_g2s(ring + slot * kChunkStride, src, chunk_bytes, &sb->pk_full[slot]);
      }
      if (p.l2_prefetch) {       // HBM has spare bandwidth here; the next kernel then finds its weights in L2
        const int64_t n_cta = (int64_t)gridDim.x * gridDim.y * gridDim.z;
        const int64_t cta = blockIdx.x + (int64_t)gridDim.x * (blockIdx.y + (int64_t)gridDim.y * blockIdx.z);
        const int64_t per = ((p.l2_prefetch_bytes + n_cta - 1) / n_cta + 4095) & ~(int64_t)4095;
        const int64_t b0 = cta * per, b1 = b0 + per < p.l2_prefetch_bytes ? b0 + per : p.l2_prefetch_bytes;
        for (int64_t b = b0; b < b1; b += 4096) {
          const uint32_t n = (uint32_t)(b1 - b < 4096 ? (b1 - b) & ~(int64_t)15 : 4096);
          if (n) asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(p.l2_prefetch + b), "r"(n) : "memory");
        }
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
#pragma unroll
        for (int m = 0; m < MT; ++m)
          tma_load_2d(smem + gi * kStageBytes + m * kTileABytes, &map_w, (kt0 + gi) * kUmmaBK, (nt * MT + m) * kUmmaM, &sb->full[gi]);
      }
      grid_dep_wait();                         // activations come from the upstream kernel
      if (tr) tr[7] = clock64();
      for (int gi = 0; gi < nst; ++gi) {
        const int s = gi % kStages;
        const int nt_here = min(tps, nkt - gi * tps);
        unsigned char* stage = smem + s * kStageBytes;
        if (gi >= pre) {
          mbar_wait(&sb->empty[s], ((uint32_t)(gi / kStages) & 1u) ^ 1u);
          mbar_arrive_expect_tx(&sb->full[s], (uint32_t)(kXInStage + nt_here * kTileBBytes));
          if constexpr (!kW4) {
#pragma unroll
            for (int m = 0; m < MT; ++m)
              tma_load_2d(stage + m * kTileABytes, &map_w, (kt0 + gi) * kUmmaBK, (nt * MT + m) * kUmmaM, &sb->full[s]);
          }
        }
        for (int t = 0; t < nt_here; ++t)
          tma_load_2d(stage + kXInStage + t * kTileBBytes, &map_x, (kt0 + gi * tps + t) * kUmmaBK, tok0, &sb->full[s]);
      }
    }
  } else {
    // ===================== MMA issuer =====================
    if (elect_one()) {
      // bf16 operands (dense path): a_format [7,10) = b_format [10,13) = 1
      const uint32_t idesc = make_idesc(kUmmaM, NT) | ((!kW4 && p.bf16) ? ((1u << 7) | (1u << 10)) : 0u);
      for (int gi = 0; gi < nst; ++gi) {
        const int s = gi % kStages;
        const int nt_here = min(tps, nkt - gi * tps);
        mbar_wait(&sb->full[s], (uint32_t)(gi / kStages) & 1u);
        tc_fence_after();
        if (tr && gi == 0) tr[2] = clock64();
        const uint32_t x_addr = smem_u32(smem + s * kStageBytes + kXInStage);
        for (int t = 0; t < nt_here; ++t) {
          const uint32_t b_addr = x_addr + t * kTileBBytes;
          if constexpr (kW4) {
#pragma unroll
            for (int m = 0; m < MT; ++m) {             // row tile m: its own accumulator and A slot, the same X tile
              const uint32_t a_tmem = tmem + (uint32_t)(Cfg::kAccCols + 64 * (s * MT + m) + 32 * t);
#pragma unroll
              for (int k = 0; k < kUmmaBK / 16; ++k)   // 16 k = 8 TMEM columns of A, 32 bytes of the B swizzle row
                umma_f16_ts(tmem + (uint32_t)(m * (Cfg::kAccCols / MT)), a_tmem + k * 8, make_smem_desc(b_addr + k * 32), idesc,
                            (gi | t | k) != 0 ? 1u : 0u);
            }
          } else {
#pragma unroll
            for (int m = 0; m < MT; ++m) {               // row tile m: its own accumulator, the same activation tile
              const uint32_t a_addr = smem_u32(smem + s * kStageBytes + m * kTileABytes);
#pragma unroll
              for (int k = 0; k < kUmmaBK / 16; ++k)     // advance 16 elements (32 bytes) inside the 128-byte swizzle row
                umma_f16(tmem + (uint32_t)(m * (Cfg::kAccCols / MT)), make_smem_desc(a_addr + k * 32), make_smem_desc(b_addr + k * 32), idesc,
                         (gi | t | k) != 0 ? 1u : 0u);
            }
          }
        }
        umma_commit(&sb->empty[s]);            // the stage is free again when the MMAs above retire
      }
      umma_commit(&sb->tmem_full);             // accumulator complete
      if (tr) tr[3] = clock64();
    }
  }

  if (tr && threadIdx.x == 0) tr[6] = clock64();
  tc_fence_before();
  __syncthreads();
  if (warp == kWarpMma) tmem_dealloc(tmem, kTmemCols);
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
template <int NT, bool kW4, int MT = 1, bool kDeep = true>
static int launch_nt_cfg(const CUtensorMap& mx, const CUtensorMap& mw, const UmmaParams& p, int n_splits, cudaStream_t s) {
  constexpr int smem_bytes = UmmaCfg<NT, kW4, kDeep, MT>::kSmemBytes;
  auto kern = umma_gemm_kernel<NT, kW4, MT, kDeep>;
  // per device and cheap: set on every launch (a process-wide "configured" flag would leave a second device unset)
  B200_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
  dim3 grid(p.N / (kUmmaM * MT), (p.T + NT - 1) / NT, n_splits);
  UmmaParams pp = p;
  if constexpr (!kW4) {
    // token tiles fastest when the activations fit comfortably in the 126 MB L2, the weights are the larger operand
    // and the grid is several waves; split-K grids keep the default (their z index already separates the waves)
    const int64_t x_bytes = (int64_t)p.T * p.K * 2, w_bytes = (int64_t)p.N * p.K * 2;
    pp.raster = (n_splits == 1 && grid.y > 1 && (int64_t)grid.x * grid.y > 2 * 148 && x_bytes <= (64ll << 20) && w_bytes > x_bytes) ? 1 : 0;
  }
  if (g_umma_trace_gridx && ((int)grid.x != g_umma_trace_gridx || (int)grid.z != g_umma_trace_gridz)) pp.trace = nullptr;
  launch_k(kern, grid, dim3(kUmmaThreads), (size_t)smem_bytes, s, mx, mw, pp);
  g_launch_count += 1;
  B200_CUDA(cudaGetLastError());
  return B200_OK;
}

template <int NT, bool kW4, int MT = 1>
static int launch_nt(const CUtensorMap& mx, const CUtensorMap& mw, const UmmaParams& p, int n_splits, cudaStream_t s) {
  if constexpr (!kW4 && MT == 1 && NT <= 64) {       // dense: the deep ring only while every CTA has an SM to itself
    if ((int64_t)(p.N / kUmmaM) * ((p.T + NT - 1) / NT) * n_splits > 148) return launch_nt_cfg<NT, kW4, MT, false>(mx, mw, p, n_splits, s);
  }
  return launch_nt_cfg<NT, kW4, MT, true>(mx, mw, p, n_splits, s);
}

template <bool kW4>
static int dispatch_nt(int nt_sel, const CUtensorMap& mx, const CUtensorMap& mw, const UmmaParams& p, int n_splits,
                       cudaStream_t s) {
  switch (nt_sel) {
    case 16: return launch_nt<16, kW4>(mx, mw, p, n_splits, s);
    case 32: return launch_nt<32, kW4>(mx, mw, p, n_splits, s);
    case 64: return launch_nt<64, kW4>(mx, mw, p, n_splits, s);
    case 128:
      if constexpr (kW4) {
        // 256 weight rows per CTA (same bits; half the activation traffic through L2, half the CTAs) as soon as
        // 128-row tiles would not fit on the GPU in one wave: large prefill passes, and decode batches of 129..256
        // rows (two token tiles: gate_up 280 CTAs -> 140, down 288 -> 144)
        if (p.N % (2 * kUmmaM) == 0 && (int64_t)(p.N / kUmmaM) * ((p.T + 127) / 128) * n_splits > 148)
          return launch_nt<128, kW4, 2>(mx, mw, p, n_splits, s);
      }
      return launch_nt<128, kW4>(mx, mw, p, n_splits, s);
    default:
      if constexpr (kW4) { set_last_error("umma_w4: token tile must be <= 128"); return B200_ERR_INVALID; }
      else {
        // 256 weight rows per CTA once 128-row tiles would need more than one wave: half the operand traffic through L2
        if (p.N % (2 * kUmmaM) == 0 && n_splits == 1 && (int64_t)(p.N / kUmmaM) * ((p.T + 255) / 256) > 148)
          return launch_nt<256, kW4, 2>(mx, mw, p, n_splits, s);
        return launch_nt<256, kW4>(mx, mw, p, n_splits, s);
      }
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
  p.l2_prefetch = w.next_chunks; p.l2_prefetch_bytes = w.next_bytes;
  p.ldy = ldy; p.T = T; p.K = w.K; p.N = w.N; p.G = w.G; p.mode = mode;
  const int KG = w.K / w.G, tpg = w.G / kUmmaBK;
  const int groups_per_split = (KG + n_splits - 1) / n_splits;     // split boundaries fall on quantisation groups
  p.kt_per_split = groups_per_split * tpg;
  n_splits = (KG + groups_per_split - 1) / groups_per_split;
  return dispatch_nt<true>(nt_sel, map_x, map_x, p, n_splits, s);
}

int umma_dense_effective_splits(int K, int n_splits) {
  const int KT = K / kUmmaBK;
  const int per = (KT + n_splits - 1) / n_splits;
  return (KT + per - 1) / per;
}

// Dense 16-bit weights W[N, K] (fp16 or bf16) by TMA.  n_splits > 1 requires mode == kPartialF32.
int launch_umma_f16(const CUtensorMap& map_w, int K, int N, const CUtensorMap& map_x, int nt_sel, int T, float* y,
                    int ldy, int mode, cudaStream_t s, int n_splits, int64_t part_stride, const __half* bias, bool bf16) {
  B200_REQUIRE(N % 128 == 0 && K % 64 == 0, "umma_f16: N % 128, K % 64");
  B200_REQUIRE(n_splits >= 1 && (n_splits == 1 || mode == kPartialF32), "umma_f16: split-K needs partial mode");
  if (T == 0) return B200_OK;
  UmmaParams p{};
  p.trace = g_umma_trace;
  p.y = y; p.ldy = ldy; p.T = T; p.K = K; p.N = N; p.G = 128; p.mode = mode; p.bias = bias; p.bf16 = bf16 ? 1 : 0;
  p.part_stride = part_stride;
  const int KT = K / kUmmaBK;
  p.kt_per_split = (KT + n_splits - 1) / n_splits;
  n_splits = (KT + p.kt_per_split - 1) / p.kt_per_split;
  return dispatch_nt<false>(nt_sel, map_x, map_w, p, n_splits, s);
}

}  // namespace b200
