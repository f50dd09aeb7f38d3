// tcgen05 / TMEM GEMM for the token-batched regime (decode batch >= 16 and prefill), swap-AB:
//
//     D[n, b] (TMEM, fp32)  +=  W[n, k] (MMA "A", 128 weight rows)  x  X[b, k] (MMA "B", NT tokens)
//
// The weight matrix is the M side so that a 16..256-token batch fills the MMA N dimension and no
// tensor-core work is wasted on padding; weights stream through exactly once per token tile.
//
//   * AWQ int4 weights: 8 producer warps load pre-tiled packed words with coalesced 128-bit
//     no-allocate loads, dequantise in registers (lop3 magic-number trick, exact fp16 (q-z)*s) and
//     write the fp16 tile straight into the canonical K-major SWIZZLE_128B shared-memory layout
//     that tcgen05.mma consumes (fence.proxy.async, then mbarrier arrive).
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

constexpr int kUmmaM = 128;          // weight rows per CTA tile
constexpr int kUmmaBK = 64;          // k per pipeline stage = one 128-byte swizzle row of fp16
constexpr int kStages = 4;
constexpr int kProducerWarps = 8;
constexpr int kUmmaThreads = (kProducerWarps + 2) * 32;   // + TMA warp + MMA warp
constexpr int kTileABytes = kUmmaM * kUmmaBK * 2;          // 16 KB

enum UmmaMode : int { kStoreF32 = 0, kAccumF32 = 1, kPartialF32 = 2 };

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
__device__ __forceinline__ void tmem_ld8(uint32_t taddr, uint32_t v[8]) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7])
               : "r"(taddr) : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
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

struct UmmaSmem {                      // barriers live after the tile ring
  uint64_t full[kStages];
  uint64_t empty[kStages];
  uint64_t tmem_full;
  uint32_t tmem_base;
};

struct UmmaParams {
  // W4 source (tiled layout, see umma_tile_w4_kernel)
  const uint4* wt;        // [N/128][K/64][2][128] uint4
  const __half* sct;      // [N/128][K/G][128]
  const __half* zpt;      // [N/128][K/G][128]  (1024 + zero)
  const __half* bias;     // [N] or null (kStoreF32 only)
  float* y;               // output, row stride ldy; kPartialF32: + split * part_stride
  int64_t part_stride;
  int ldy, T, K, N, G;
  int kt_per_split;       // k tiles handled by one CTA along grid.z
  int mode;
};

template <int NT, bool kW4>
__global__ void __launch_bounds__(kUmmaThreads, 1)
umma_gemm_kernel(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_w,
                 const UmmaParams p) {
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  // 1024-byte alignment is required by SWIZZLE_128B; dynamic smem base is only 16-byte aligned by contract
  unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  constexpr int kTileBBytes = NT * kUmmaBK * 2;
  constexpr int kStageBytes = kTileABytes + kTileBBytes;
  UmmaSmem* sb = reinterpret_cast<UmmaSmem*>(smem + kStages * kStageBytes);
  constexpr uint32_t kTmemCols = NT < 32 ? 32 : NT;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nt = blockIdx.x;                 // weight-row tile
  const int tok0 = blockIdx.y * NT;          // token tile
  const int split = blockIdx.z;
  const int KT = p.K / kUmmaBK;
  const int kt0 = split * p.kt_per_split;
  const int nkt = min(p.kt_per_split, KT - kt0);

  if (threadIdx.x == 0) {
    const uint32_t full_count = kW4 ? kProducerWarps * 32 + 1 : 1;
    for (int s = 0; s < kStages; ++s) { mbar_init(&sb->full[s], full_count); mbar_init(&sb->empty[s], 1); }
    mbar_init(&sb->tmem_full, 1);
    fence_mbar_init();
  }
  if (warp == kProducerWarps + 1) tmem_alloc(&sb->tmem_base, kTmemCols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = sb->tmem_base;

  if (warp < kProducerWarps) {
    // ===================== weight producers (W4 only) =====================
    if (kW4) {
      const int r = threadIdx.x & 127, hsel = threadIdx.x >> 7;       // row in tile, which 32-k half
      const uint4* wsrc = p.wt + ((size_t)nt * KT * 2 + hsel) * 128 + r;   // + kt * 256
      const int KG = p.K / p.G;
      const __half* ssrc = p.sct + (size_t)nt * KG * 128 + r;
      const __half* zsrc = p.zpt + (size_t)nt * KG * 128 + r;
      constexpr int PF = 4;                                           // register prefetch distance
      uint4 buf[PF];
#pragma unroll
      for (int j = 0; j < PF; ++j)
        if (j < nkt) buf[j] = ldg_stream(wsrc + (size_t)(kt0 + j) * 256);
      for (int it0 = 0; it0 < nkt; it0 += PF) {
#pragma unroll
        for (int j = 0; j < PF; ++j) {
          const int it = it0 + j;
          if (it < nkt) {
            const int s = it % kStages;
            const uint32_t ph = (uint32_t)(it / kStages) & 1u;
            const uint4 w = buf[j];
            if (it + PF < nkt) buf[j] = ldg_stream(wsrc + (size_t)(kt0 + it + PF) * 256);
            const int g = ((kt0 + it) * kUmmaBK + hsel * 32) / p.G;
            const __half2 s2 = __half2half2(ssrc[(size_t)g * 128]);
            const __half2 z2 = __half2half2(zsrc[(size_t)g * 128]);
            const uint32_t su = *reinterpret_cast<const uint32_t*>(&s2), zu = *reinterpret_cast<const uint32_t*>(&z2);
            uint32_t d[16];
            dequant8(w.x, zu, su, d + 0);
            dequant8(w.y, zu, su, d + 4);
            dequant8(w.z, zu, su, d + 8);
            dequant8(w.w, zu, su, d + 12);
            mbar_wait(&sb->empty[s], ph ^ 1u);
            unsigned char* rowp = smem + s * kStageBytes + r * 128;
#pragma unroll
            for (int c = 0; c < 4; ++c) {
              const int chunk = (hsel * 4 + c) ^ (r & 7);             // SWIZZLE_128B: 16 B chunk ^= row % 8
              *reinterpret_cast<uint4*>(rowp + chunk * 16) = make_uint4(d[4 * c], d[4 * c + 1], d[4 * c + 2], d[4 * c + 3]);
            }
            fence_proxy_async_smem();                                  // generic-proxy stores -> async proxy (MMA)
            mbar_arrive(&sb->full[s]);
          }
        }
      }
    }
    // ===================== epilogue: TMEM -> registers -> global =====================
    mbar_wait(&sb->tmem_full, 0);
    tc_fence_after();
    const int q = warp & 3;                                            // TMEM lane quarter this warp may read
    const int n = nt * kUmmaM + q * 32 + lane;                         // output feature
    constexpr int kColsPerWarp = NT >= 32 ? NT / 2 : NT;
    const int col0 = NT >= 32 ? (warp >> 2) * kColsPerWarp : 0;
    if (NT >= 32 || warp < 4) {
      const float bn = (p.mode == kStoreF32 && p.bias) ? __half2float(p.bias[n]) : 0.f;
      float* ybase = p.y + (p.mode == kPartialF32 ? (size_t)split * p.part_stride : 0);
#pragma unroll 1
      for (int c = 0; c < kColsPerWarp; c += 8) {
        uint32_t v[8];
        tmem_ld8(tmem + ((uint32_t)(q * 32) << 16) + (uint32_t)(col0 + c), v);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const int b = tok0 + col0 + c + j;
          if (b < p.T) {
            float* yp = ybase + (size_t)b * p.ldy + n;
            const float acc = __uint_as_float(v[j]);
            if (p.mode == kAccumF32) *yp += acc;
            else *yp = acc + bn;
          }
        }
      }
    }
  } else if (warp == kProducerWarps) {
    // ===================== TMA producer =====================
    if (elect_one()) {
      for (int it = 0; it < nkt; ++it) {
        const int s = it % kStages;
        const uint32_t ph = (uint32_t)(it / kStages) & 1u;
        mbar_wait(&sb->empty[s], ph ^ 1u);
        unsigned char* stage = smem + s * kStageBytes;
        mbar_arrive_expect_tx(&sb->full[s], kW4 ? kTileBBytes : kStageBytes);
        if (!kW4) tma_load_2d(stage, &map_w, (kt0 + it) * kUmmaBK, nt * kUmmaM, &sb->full[s]);
        tma_load_2d(stage + kTileABytes, &map_x, (kt0 + it) * kUmmaBK, tok0, &sb->full[s]);
      }
    }
  } else {
    // ===================== MMA issuer =====================
    if (elect_one()) {
      constexpr uint32_t idesc = make_idesc(kUmmaM, NT);
      for (int it = 0; it < nkt; ++it) {
        const int s = it % kStages;
        const uint32_t ph = (uint32_t)(it / kStages) & 1u;
        mbar_wait(&sb->full[s], ph);
        tc_fence_after();
        const uint32_t a_addr = smem_u32(smem + s * kStageBytes);
        const uint32_t b_addr = a_addr + kTileABytes;
#pragma unroll
        for (int k = 0; k < kUmmaBK / 16; ++k) {
          // advance 16 elements (32 bytes) along K inside the 128-byte swizzle row
          umma_f16(tmem, make_smem_desc(a_addr + k * 32), make_smem_desc(b_addr + k * 32), idesc,
                   (it | k) != 0 ? 1u : 0u);
        }
        umma_commit(&sb->empty[s]);            // frees this smem stage when the MMAs above retire
      }
      umma_commit(&sb->tmem_full);             // accumulator complete
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == kProducerWarps + 1) tmem_dealloc(tmem, kTmemCols);
}

// ---- weight tiling (one-time) ---------------------------------------------------------------------
// kernel layout A (wq [N, K/8], sc/zp [N, K/G]) -> tiled layout for the producers above.
__global__ void umma_tile_w4_kernel(const uint32_t* __restrict__ wq, const __half* __restrict__ sc,
                                    const __half* __restrict__ zp, int K, int N, int G, uint4* __restrict__ wt,
                                    __half* __restrict__ sct, __half* __restrict__ zpt) {
  const int KT = K / 64, KG = K / G;
  const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const size_t total = (size_t)(N / 128) * KT * 256;
  if (idx < total) {
    const int r = idx & 127, h = (idx >> 7) & 1;
    const size_t tk = idx >> 8;
    const int kt = (int)(tk % KT), nt = (int)(tk / KT);
    const uint32_t* src = wq + (size_t)(nt * 128 + r) * (K / 8) + kt * 8 + h * 4;
    wt[idx] = make_uint4(src[0], src[1], src[2], src[3]);
  }
  const size_t total_s = (size_t)(N / 128) * KG * 128;
  if (idx < total_s) {
    const int r = idx & 127;
    const size_t tg = idx >> 7;
    const int g = (int)(tg % KG), nt = (int)(tg / KG);
    sct[idx] = sc[(size_t)(nt * 128 + r) * KG + g];
    zpt[idx] = zp[(size_t)(nt * 128 + r) * KG + g];
  }
}

int launch_umma_tile_w4(const uint32_t* wq, const __half* sc, const __half* zp, int K, int N, int G, uint4* wt,
                        __half* sct, __half* zpt, cudaStream_t s) {
  B200_REQUIRE(N % 128 == 0 && K % 64 == 0 && K % G == 0 && G % 32 == 0, "umma_tile_w4: N % 128, K % 64, K % G");
  const size_t total = (size_t)(N / 128) * (K / 64) * 256;
  umma_tile_w4_kernel<<<(unsigned)((total + 255) / 256), 256, 0, s>>>(wq, sc, zp, K, N, G, wt, sct, zpt);
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
  constexpr int smem_bytes = kStages * (kTileABytes + NT * kUmmaBK * 2) + (int)sizeof(UmmaSmem) + 1024;
  auto kern = umma_gemm_kernel<NT, kW4>;
  static bool configured = false;
  if (!configured) {
    B200_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
    configured = true;
  }
  dim3 grid(p.N / kUmmaM, (p.T + NT - 1) / NT, n_splits);
  kern<<<grid, kUmmaThreads, smem_bytes, s>>>(mx, mw, p);
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
    default: return launch_nt<256, kW4>(mx, mw, p, n_splits, s);
  }
}

int umma_pick_nt(int T) { return T <= 16 ? 16 : T <= 32 ? 32 : T <= 64 ? 64 : T <= 128 ? 128 : 256; }

// W4: y = x . W^T.  n_splits > 1 requires mode == kPartialF32 (y then holds n_splits partial tensors).
int launch_umma_w4(const UmmaW4& w, const CUtensorMap& map_x, int nt_sel, int T, float* y, int ldy, int mode,
                   int n_splits, int64_t part_stride, cudaStream_t s) {
  B200_REQUIRE(w.N % 128 == 0 && w.K % 64 == 0, "umma_w4: N % 128, K % 64");
  B200_REQUIRE(n_splits >= 1 && (n_splits == 1 || mode == kPartialF32), "umma_w4: split-K needs partial mode");
  if (T == 0) return B200_OK;
  UmmaParams p{};
  p.wt = w.wt; p.sct = w.sct; p.zpt = w.zpt; p.bias = w.bias; p.y = y; p.part_stride = part_stride;
  p.ldy = ldy; p.T = T; p.K = w.K; p.N = w.N; p.G = w.G; p.mode = mode;
  const int KT = w.K / kUmmaBK;
  p.kt_per_split = (KT + n_splits - 1) / n_splits;
  n_splits = (KT + p.kt_per_split - 1) / p.kt_per_split;
  return dispatch_nt<true>(nt_sel, map_x, map_x, p, n_splits, s);
}

int launch_umma_f16(const CUtensorMap& map_w, int K, int N, const CUtensorMap& map_x, int nt_sel, int T, float* y,
                    int ldy, int mode, cudaStream_t s) {
  B200_REQUIRE(N % 128 == 0 && K % 64 == 0, "umma_f16: N % 128, K % 64");
  if (T == 0) return B200_OK;
  UmmaParams p{};
  p.y = y; p.ldy = ldy; p.T = T; p.K = K; p.N = N; p.G = 128; p.mode = mode;
  p.kt_per_split = K / kUmmaBK;
  return dispatch_nt<false>(nt_sel, map_x, map_w, p, 1, s);
}

}  // namespace b200
