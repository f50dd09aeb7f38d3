// Persistent decode kernel: all transformer layers of one decode step (<= 64 rows) in ONE launch.
//
// The per-kernel decode step is bound by dependent kernel boundaries (7 per layer, ~2.5 us each with
// PDL), not by HBM.  Here one CTA per SM stays resident for the whole step and the 7 phases of a layer
//     rms1 -> qkv GEMM -> attention -> o GEMM -> rms2 -> gate_up GEMM (+SwiGLU) -> down GEMM
// are separated by grid-wide flags in global memory (gridsync.cuh, ~1.4 us), while the weight stream
// never stops: every CTA's share of every GEMM is a static schedule, so its TMA warp keeps the packed
// int4 ring full across phase boundaries and the dequantising warpgroups convert chunks into TMEM
// before the activations they will be multiplied with exist.
//
// Warp roles are those of umma_gemm_kernel<64, W4> (ops_umma.cu): 16 compute warps (4 warpgroups:
// dequant producers, epilogue, and -- between GEMMs -- the RMSNorm and attention phases), one TMA warp
// for weights, one for activations, one MMA issuer.  The arithmetic is the per-kernel path's, through
// the same device functions (umma.cuh, attn.cuh, norm.cuh): both paths produce identical bits, which
// tests/test_engine_gpu.py checks.
//
// Also here: the grid-synchronisation microbenchmark (b200_debug_gridsync_bench).
#include <mutex>

#include "attn.cuh"
#include "gridsync.cuh"
#include "norm.cuh"
#include "umma.cuh"

namespace b200 {

B200_DEFINE_TL_SETTER(tl_set_mega)

constexpr int kMegaNT = 64;                              // token tile = the whole decode batch
constexpr int kMegaPhases = 7;
constexpr int kMegaComputeThreads = kProducerWarps * 32; // 512
using MegaCfg = UmmaCfg<kMegaNT, true, false>;   // four stages: stage s belongs to warpgroup s

// Shared memory: [attention tiles / X stages + epilogue staging (aliased)] [weight ring] [barriers]
struct MegaSmem {
  static constexpr int kOffRegionA = 0;                  // AttnSmem (135712 B) or 4 X stages (64 KB) or staging (32 KB)
  static constexpr int kRegionABytes = (AttnSmem::kBytes + 1023) / 1024 * 1024;
  static constexpr int kOffRing = kRegionABytes;
  static constexpr int kOffBar = kOffRing + MegaCfg::kRing * kChunkStride;
  static constexpr int kOffAccEmpty = kOffBar + (int)sizeof(UmmaSmem);
  static constexpr int kOffRed = kOffAccEmpty + 16;      // float [4][4] RMSNorm partial sums
  static constexpr int kBytes = kOffRed + 64 + 1024;     // + alignment slack
  static_assert(MegaCfg::kStages * MegaCfg::kStageBytes <= AttnSmem::kOffBar, "X stages stay clear of the attention barriers");
  static_assert(kBytes <= 227 * 1024, "shared memory budget");
};

// ---- static work schedule ---------------------------------------------------------------------------
struct GemmShape { int K, N, tiles, nsplit, kt_per_split, KT, tps, mode, ldy; };

__device__ __forceinline__ GemmShape gemm_shape(const MegaParams& P, int g) {
  GemmShape s;
  s.K = g == 0 ? P.H : g == 1 ? P.n_heads * kHeadDim : g == 2 ? P.H : P.I;
  s.N = g == 0 ? P.QKV : g == 1 ? P.H : g == 2 ? 2 * P.I : P.H;
  s.tiles = s.N / kUmmaM;
  s.tps = P.G / kUmmaBK;
  s.KT = s.K / kUmmaBK;
  const int KG = s.K / P.G;
  const int want = g == 2 ? 1 : min(P.splits[g], KG);
  const int gps = (KG + want - 1) / want;                // same arithmetic as launch_umma_w4
  s.kt_per_split = gps * s.tps;
  s.nsplit = (KG + gps - 1) / gps;
  s.mode = g == 2 ? kSwigluF16 : kPartialF32;
  s.ldy = g == 2 ? P.I : s.N;
  return s;
}
// CTA rotation per GEMM so that the light projections (qkv, o) prefetch on different SMs
__device__ __forceinline__ int gemm_first_unit(int g, int cta, int G) {
  const int off = g == 1 ? 64 % G : g == 3 ? 4 % G : 0;
  return (cta + G - off) % G;
}
__device__ __forceinline__ int flag_of_input(int l, int g) {           // phase whose output GEMM g of layer l reads
  return l * kMegaPhases + (g == 0 ? 0 : g == 1 ? 2 : g == 2 ? 4 : 5);
}

// f(l, g, shape, nt, split, kt0, nkt, nst) for every GEMM unit of this CTA, in execution order
template <typename F>
__device__ __forceinline__ void for_each_gemm_unit(const MegaParams& P, int cta, int G, F&& f) {
  for (int l = 0; l < P.n_layers; ++l)
    for (int g = 0; g < 4; ++g) {
      const GemmShape s = gemm_shape(P, g);
      for (int u = gemm_first_unit(g, cta, G); u < s.tiles * s.nsplit; u += G) {
        const int nt = u % s.tiles, split = u / s.tiles;
        const int kt0 = split * s.kt_per_split;
        const int nkt = min(s.kt_per_split, s.KT - kt0);
        f(l, g, s, nt, split, kt0, nkt, (nkt + s.tps - 1) / s.tps);
      }
    }
}

__device__ __forceinline__ void compute_bar() {                       // the 512 compute threads
  asm volatile("bar.sync 2, %0;" ::"n"(kMegaComputeThreads) : "memory");
}

// ---- the kernel ---------------------------------------------------------------------------------------
template <int GQ>
__global__ void __launch_bounds__(kUmmaThreads, 1)
decode_layers_kernel(const __grid_constant__ CUtensorMap map_xn, const __grid_constant__ CUtensorMap map_attn,
                     const __grid_constant__ CUtensorMap map_act, const MegaParams P) {
  using L = MegaSmem;
  constexpr int kStages = MegaCfg::kStages, kStageBytes = MegaCfg::kStageBytes, kTileBBytes = MegaCfg::kTileBBytes;
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  unsigned char* ring = smem + L::kOffRing;
  UmmaSmem* sb = reinterpret_cast<UmmaSmem*>(smem + L::kOffBar);
  uint64_t* acc_empty = reinterpret_cast<uint64_t*>(smem + L::kOffAccEmpty);
  float (*red)[kNormWarps] = reinterpret_cast<float (*)[kNormWarps]>(smem + L::kOffRed);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int cta = blockIdx.x, G = gridDim.x;

  if (threadIdx.x == 0) {
    for (int s = 0; s < kStages; ++s) { mbar_init(&sb->full[s], 128 + 1); mbar_init(&sb->empty[s], 1); }
    for (int s = 0; s < MegaCfg::kRing; ++s) { mbar_init(&sb->pk_full[s], 1); mbar_init(&sb->pk_empty[s], 128); }
    mbar_init(&sb->tmem_full, 1);
    mbar_init(acc_empty, 1);
    for (int w = 0; w < kAttnTileWarps; ++w) mbar_init(reinterpret_cast<uint64_t*>(smem + AttnSmem::kOffBar) + w, 1);
    fence_mbar_init();
  }
  if (warp == kWarpMma) tmem_alloc(&sb->tmem_base, MegaCfg::kTmemCols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = sb->tmem_base;

  if (warp < kProducerWarps) {
    // =============== compute warps: RMSNorm, dequant producers + epilogue, attention ===============
    const int wg = warp >> 2, r = threadIdx.x & 127;
    const uint32_t lane_base = (uint32_t)((warp & 3) * 32) << 16;
    uint32_t n_fill = 0;            // chunks this warpgroup has converted so far (ring slot / stage phase)
    uint32_t n_units = 0;           // GEMM units this CTA has finished (accumulator phase)
    uint32_t attn_uses = 0;         // completed phases of this warp's attention barrier (warps 0..3)
    unsigned int* const flags = P.flags;

    auto wait_phase = [&](int idx) {                     // every compute thread may read what phase idx wrote
      if (threadIdx.x == 0) {
        flag_wait(flags + idx, (unsigned)G);
        if (P.trace) P.trace[((size_t)cta * P.n_layers * kMegaPhases + idx + 1) * 2] = gtime();   // start of phase idx + 1
      }
      compute_bar();
    };
    auto arrive_phase = [&](int idx) {                   // everything the compute threads wrote is published
      asm volatile("fence.proxy.async;" ::: "memory");   // ... also to the TMA (async-proxy) reads of other CTAs
      compute_bar();
      if (threadIdx.x == 0) {
        __threadfence();
        flag_arrive(flags + idx);
        if (P.trace) P.trace[((size_t)cta * P.n_layers * kMegaPhases + idx) * 2 + 1] = gtime();
      }
    };
    auto rms_phase = [&](const float* parts, int n_parts, const __half* gamma) {
      for (int row = (cta + G - 74 % G) % G; row < P.T; row += G) {
        float4 v[kNormMaxPerThread];
        const float ss = rms_quarter_load(P.h, parts, n_parts, (int64_t)P.T * P.H, row, wg, P.H, r, v);
        if (lane == 0) red[wg][warp & 3] = ss;
        compute_bar();
        const float rs = rms_scale(red, P.H, P.rms_eps);
        rms_quarter_store(v, rs, gamma, P.xn, row, wg, P.H, r);
        compute_bar();                                   // red is reused by the next row
      }
    };
    auto gemm_phase = [&](int l, int g) {
      const GemmShape s = gemm_shape(P, g);
      for (int u = gemm_first_unit(g, cta, G); u < s.tiles * s.nsplit; u += G) {
        const int nt = u % s.tiles, split = u / s.tiles;
        const int kt0 = split * s.kt_per_split;
        const int nkt = min(s.kt_per_split, s.KT - kt0);
        const int nst = (nkt + s.tps - 1) / s.tps;
        for (int gi = wg; gi < nst; gi += 4, ++n_fill) {
          const int slot = wg + 4 * (int)(n_fill & 1u);
          w4_chunk_to_tmem(smem_u32(ring + slot * kChunkStride), &sb->pk_full[slot], (n_fill >> 1) & 1u, &sb->empty[wg],
                           (n_fill & 1u) ^ 1u, &sb->full[wg], &sb->pk_empty[slot],
                           tmem + lane_base + (uint32_t)(MegaCfg::kAccCols + 64 * wg), r, min(s.tps, nkt - gi * s.tps));
        }
        mbar_wait(&sb->tmem_full, n_units & 1u);
        tc_fence_after();
        if (P.trace && threadIdx.x == 0) P.trace[((size_t)cta * P.n_layers * kMegaPhases + l * kMegaPhases + (g == 0 ? 1 : g == 1 ? 3 : g + 3)) * 2] = gtime();   // accumulator complete
        float* y = g == 2 ? reinterpret_cast<float*>(P.act) : P.parts;
        umma_epilogue<kMegaNT>(s.mode, y, s.ldy, (size_t)split * ((size_t)P.T * s.N), nullptr, P.T, 0, nt,
                               reinterpret_cast<float*>(smem), tmem, nullptr);
        tc_fence_before();
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // staging (generic) before the next X tiles (TMA)
        compute_bar();                                   // accumulator and staging buffer are free again
        if (threadIdx.x == 0) mbar_arrive(acc_empty);
        ++n_units;
      }
    };

    grid_dep_wait();
    for (int l = 0; l < P.n_layers; ++l) {
      const MegaLayer& ly = P.layers[l];
      const int f0 = l * kMegaPhases;
      const int np_qkv = gemm_shape(P, 0).nsplit, np_o = gemm_shape(P, 1).nsplit, np_down = gemm_shape(P, 3).nsplit;
      // -- phase 0: h += down partials of the previous layer; xn = RMSNorm(h) * ln1
      if (l > 0) wait_phase(f0 - 1);
      rms_phase(P.parts, l > 0 ? np_down : 0, ly.ln1);
      arrive_phase(f0 + 0);
      // -- phase 1: qkv projection (split-K partials)
      gemm_phase(l, 0);
      arrive_phase(f0 + 1);
      // -- phase 2: RoPE + KV write + attention over the paged cache
      wait_phase(f0 + 1);
      {
        AttnParams ap{};
        ap.qkv = P.parts; ap.n_parts = np_qkv; ap.part_stride = (int64_t)P.T * P.QKV; ap.bias = ly.bias_qkv;
        ap.pos = P.pos; ap.slot = P.slot; ap.cos_t = P.cos_t; ap.sin_t = P.sin_t;
        ap.k_cache = ly.k_cache; ap.v_cache = ly.v_cache;
        ap.block_tables = P.block_tables; ap.bt_stride = P.bt_stride; ap.bt_row = P.bt_row; ap.ctx_len = P.ctx_len;
        ap.n_heads = P.n_heads; ap.n_kv = P.n_kv; ap.n_splits = 1; ap.out = P.attn;
        for (int u = cta; u < P.n_kv * P.T; u += G) {
          if (warp < kAttnWarps)
            paged_attn_unit<GQ, true, 3, false>(ap, smem + L::kOffRegionA, u % P.n_kv, u / P.n_kv, 0, nullptr, attn_uses);
          asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic tile accesses before the next TMA writes
          compute_bar();                                 // the tiles are reused by the next unit / the X stages
        }
      }
      arrive_phase(f0 + 2);
      // -- phase 3: o projection (split-K partials)
      gemm_phase(l, 1);
      arrive_phase(f0 + 3);
      // -- phase 4: h += o partials; xn = RMSNorm(h) * ln2
      wait_phase(f0 + 3);
      rms_phase(P.parts, np_o, ly.ln2);
      arrive_phase(f0 + 4);
      // -- phase 5: gate_up projection + SwiGLU -> act
      gemm_phase(l, 2);
      arrive_phase(f0 + 5);
      // -- phase 6: down projection (split-K partials, summed by the next RMSNorm)
      gemm_phase(l, 3);
      arrive_phase(f0 + 6);
    }
  } else if (warp == kWarpTmaW) {
    // =============== weight stream: never waits for activations ===============
    if (elect_one()) {
      uint32_t nf = 0;                                  // four 8-bit chunk counters, one per warpgroup (no local-memory array)
      for_each_gemm_unit(P, cta, G, [&](int l, int g, const GemmShape& s, int nt, int split, int kt0, int nkt, int nst) {
        const uint32_t chunk_bytes = 512u + (uint32_t)s.tps * 4096u;
        const unsigned char* src = P.layers[l].chunks[g] + ((size_t)nt * (s.K / P.G) + kt0 / s.tps) * chunk_bytes;
        for (int gi = 0; gi < nst; ++gi) {
          const int w = gi & 3;
          const uint32_t n = (nf >> (8 * w)) & 0xffu;
          nf = (nf & ~(0xffu << (8 * w))) | (((n + 1) & 0xffu) << (8 * w));   // only n mod 4 is ever used
          const int slot = w + 4 * (int)(n & 1u);
          mbar_wait(&sb->pk_empty[slot], ((n >> 1) & 1u) ^ 1u);
          mbar_arrive_expect_tx(&sb->pk_full[slot], chunk_bytes);
          bulk_g2s(ring + slot * kChunkStride, src + (size_t)gi * chunk_bytes, chunk_bytes, &sb->pk_full[slot]);
        }
      });
    }
  } else if (warp == kWarpTmaX) {
    // =============== activation tiles: wait for the producing phase, then TMA ===============
    if (elect_one()) {
      prefetch_tensormap(&map_xn);
      prefetch_tensormap(&map_attn);
      prefetch_tensormap(&map_act);
      grid_dep_wait();
      uint32_t par = 0;                                 // bit st = parity of the number of uses of stage st
      uint32_t n_units = 0;
      for_each_gemm_unit(P, cta, G, [&](int l, int g, const GemmShape& s, int nt, int split, int kt0, int nkt, int nst) {
        const CUtensorMap* map = g == 1 ? &map_attn : g == 3 ? &map_act : &map_xn;
        flag_wait(P.flags + flag_of_input(l, g), (unsigned)G);
        if (n_units > 0) mbar_wait(acc_empty, (n_units - 1) & 1u);     // epilogue staging aliases the X stages
        for (int gi = 0; gi < nst; ++gi) {
          const int st = gi & 3;
          const int nt_here = min(s.tps, nkt - gi * s.tps);
          mbar_wait(&sb->empty[st], ((par >> st) & 1u) ^ 1u);
          par ^= 1u << st;
          unsigned char* stage = smem + st * kStageBytes;
          mbar_arrive_expect_tx(&sb->full[st], (uint32_t)(nt_here * kTileBBytes));
          for (int t = 0; t < nt_here; ++t)
            tma_load_2d(stage + t * kTileBBytes, map, (kt0 + gi * s.tps + t) * kUmmaBK, 0, &sb->full[st]);
        }
        ++n_units;
      });
    }
  } else {
    // =============== MMA issuer ===============
    if (elect_one()) {
      constexpr uint32_t idesc = make_idesc(kUmmaM, kMegaNT);
      uint32_t par = 0;                                 // bit st = parity of the number of uses of stage st
      uint32_t n_units = 0;
      for_each_gemm_unit(P, cta, G, [&](int l, int g, const GemmShape& s, int nt, int split, int kt0, int nkt, int nst) {
        if (n_units > 0) {                               // the epilogue has drained the previous accumulator
          mbar_wait(acc_empty, (n_units - 1) & 1u);
          tc_fence_after();
        }
        for (int gi = 0; gi < nst; ++gi) {
          const int st = gi & 3;
          const int nt_here = min(s.tps, nkt - gi * s.tps);
          mbar_wait(&sb->full[st], (par >> st) & 1u);
          par ^= 1u << st;
          tc_fence_after();
          const uint32_t x_addr = smem_u32(smem + st * kStageBytes);
          for (int t = 0; t < nt_here; ++t) {
            const uint32_t b_addr = x_addr + t * kTileBBytes;
            const uint32_t a_tmem = tmem + (uint32_t)(MegaCfg::kAccCols + 64 * st + 32 * t);
#pragma unroll
            for (int k = 0; k < kUmmaBK / 16; ++k)
              umma_f16_ts(tmem, a_tmem + k * 8, make_smem_desc(b_addr + k * 32), idesc, (gi | t | k) != 0 ? 1u : 0u);
          }
          umma_commit(&sb->empty[st]);
        }
        umma_commit(&sb->tmem_full);
        ++n_units;
      });
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == kWarpMma) tmem_dealloc(tmem, MegaCfg::kTmemCols);
}

// =========================================================================================================
// Fused MLP for decode (T <= 64 rows): gate_up (+SwiGLU) and down in ONE launch, joined by dataflow.  EXPERIMENT.
//
// As separate kernels, down could not begin before the last gate_up CTA had exited and the grid had
// drained (3.9 us of boundary per layer), and its CTAs could not even be resident -- hence prefetch weights
// -- while gate_up's were.  Here one CTA per SM first computes its gate_up tile (64 gate + 64 up rows ->
// 64 columns of `act`), publishes a flag for exactly those columns, and then computes its down unit, whose
// activation stages wait only for the flags of the two 64-column k tiles they read.  The weight TMA warp
// streams gate_up's and then down's chunks without ever waiting for activations, and the dequantising
// warpgroups fill tensor memory with down's A operand while the flags are still in flight.
// Warp roles and arithmetic are those of umma_gemm_kernel (same device functions, same split-K): the
// output bits equal the two-kernel path's (tests/test_mega_gpu.py).
//
// MEASURED (batch 64, tools/mlp_trace.py): after the dependency resolves, gate_up's accumulator is complete at
// 5.3 us, its columns are published at 7.5 (epilogue 1.3 + fences / store drain 0.9), the consumers see the
// counters at 8.6, down's accumulator is complete at 10.9 and its partials are written at 12.4 -- against
// 14.2 us for the two kernels.  But the grid holds all 148 SMs until its slowest CTA ends, so the next kernels
// lose their head start, and the decode step comes out 2.5 % SLOWER (1.242 vs 1.212 ms).  Producer -> L2 ->
// consumer costs ~3.5 us whether the signal is a flag or a kernel boundary.  Opt-in: VGATE_B200_FUSED_MLP=1.
struct MlpSmem {
  static constexpr int kOffStages = 0;                   // 4 X stages (64 KB); the epilogue staging aliases them
  static constexpr int kOffRing = MegaCfg::kStages * MegaCfg::kStageBytes;
  static constexpr int kOffBar = kOffRing + MegaCfg::kRing * kChunkStride;
  static constexpr int kOffAccEmpty = kOffBar + (int)sizeof(UmmaSmem);
  static constexpr int kBytes = kOffAccEmpty + 16 + 1024;
};

struct MlpUnit { int g, nt, split, kt0, nkt, nst; };     // g: 0 = gate_up, 1 = down

// f(unit) for every unit of this CTA in execution order: its gate_up tiles, then its down units
template <typename F>
__device__ __forceinline__ void for_each_mlp_unit(const MlpParams& P, int cta, int G, F&& f) {
  const int tps = P.G / kUmmaBK;
  {
    const int tiles = (2 * P.I) / kUmmaM, nkt = P.H / kUmmaBK;
    for (int u = cta; u < tiles; u += G) f(MlpUnit{0, u, 0, 0, nkt, (nkt + tps - 1) / tps});
  }
  {
    const int tiles = P.H / kUmmaM, KT = P.I / kUmmaBK, KG = P.I / P.G;
    const int gps = (KG + P.split_down - 1) / P.split_down;          // same arithmetic as launch_umma_w4
    const int kt_per_split = gps * tps, nsplit = (KG + gps - 1) / gps;
    for (int u = (cta + G - 4 % G) % G; u < tiles * nsplit; u += G) {
      const int nt = u % tiles, split = u / tiles, kt0 = split * kt_per_split;
      const int nkt = min(kt_per_split, KT - kt0);
      f(MlpUnit{1, nt, split, kt0, nkt, (nkt + tps - 1) / tps});
    }
  }
}

__global__ void __launch_bounds__(kUmmaThreads, 1)
mlp_fused_kernel(const __grid_constant__ CUtensorMap map_xn, const __grid_constant__ CUtensorMap map_act, const MlpParams P) {
  using L = MlpSmem;
  constexpr int kStages = MegaCfg::kStages, kStageBytes = MegaCfg::kStageBytes, kTileBBytes = MegaCfg::kTileBBytes;
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  unsigned char* ring = smem + L::kOffRing;
  UmmaSmem* sb = reinterpret_cast<UmmaSmem*>(smem + L::kOffBar);
  uint64_t* acc_empty = reinterpret_cast<uint64_t*>(smem + L::kOffAccEmpty);
  const int warp = threadIdx.x >> 5;
  const int cta = blockIdx.x, G = gridDim.x;
  const int tps = P.G / kUmmaBK;
  const int down_kt_per_split = ((P.I / P.G + P.split_down - 1) / P.split_down) * tps;   // k tiles per down split
  unsigned long long* tr = P.trace ? P.trace + 8 * (size_t)cta : nullptr;
  if (tr && threadIdx.x == 0) tr[0] = clock64();
  unsigned tl = ~0u;
  unsigned long long tl_t0 = 0;
  const bool tl_rec = threadIdx.x == 0 && blockIdx.x == 0 && d_tl.buf;
  if (tl_rec) tl_t0 = gtime();
  grid_dep_launch();                 // every CTA of this grid is resident by the time the next kernel may start
  if (threadIdx.x == 0) {
    for (int s = 0; s < kStages; ++s) { mbar_init(&sb->full[s], 128 + 1); mbar_init(&sb->empty[s], 1); }
    for (int s = 0; s < MegaCfg::kRing; ++s) { mbar_init(&sb->pk_full[s], 1); mbar_init(&sb->pk_empty[s], 128); }
    mbar_init(&sb->tmem_full, 1);
    mbar_init(acc_empty, 1);
    fence_mbar_init();
  }
  if (warp == kWarpMma) tmem_alloc(&sb->tmem_base, MegaCfg::kTmemCols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = sb->tmem_base;

  if (warp < kProducerWarps) {
    // =============== dequant producers + epilogues ===============
    const int wg = warp >> 2, r = threadIdx.x & 127;
    const uint32_t lane_base = (uint32_t)((warp & 3) * 32) << 16;
    uint32_t n_fill = 0, n_units = 0;
    for_each_mlp_unit(P, cta, G, [&](const MlpUnit& u) {
      for (int gi = wg; gi < u.nst; gi += 4, ++n_fill) {
        const int slot = wg + 4 * (int)(n_fill & 1u);
        w4_chunk_to_tmem(smem_u32(ring + slot * kChunkStride), &sb->pk_full[slot], (n_fill >> 1) & 1u, &sb->empty[wg],
                         (n_fill & 1u) ^ 1u, &sb->full[wg], &sb->pk_empty[slot],
                         tmem + lane_base + (uint32_t)(MegaCfg::kAccCols + 64 * wg), r, min(tps, u.nkt - gi * tps));
      }
      mbar_wait(&sb->tmem_full, n_units & 1u);
      tc_fence_after();
      if (tr && threadIdx.x == 0) tr[u.g == 0 ? 1 : 3] = clock64();
      if (u.g == 0)
        umma_epilogue<kMegaNT>(kSwigluF16, reinterpret_cast<float*>(P.act), P.I, 0, nullptr, P.T, 0, u.nt,
                               reinterpret_cast<float*>(smem), tmem, nullptr);
      else
        umma_epilogue<kMegaNT>(kPartialF32, P.parts, P.H, (size_t)u.split * ((size_t)P.T * P.H), nullptr, P.T, 0, u.nt,
                               reinterpret_cast<float*>(smem), tmem, nullptr);
      tc_fence_before();
      asm volatile("fence.proxy.async;" ::: "memory");   // act (generic stores) is read by other CTAs' TMA; staging vs next X tiles
      compute_bar();
      if (threadIdx.x == 0) {
        mbar_arrive(acc_empty);
        if (u.g == 0) {
          // columns [64 nt, 64 nt + 64) of act are visible: count them for the down split whose K range holds them.
          // (release: ordered after every compute thread's stores through the barrier above; the threads already
          // issued their generic->async proxy fences)
          asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(P.flags + u.nt / down_kt_per_split) : "memory");
        }
        if (tr) tr[u.g == 0 ? 2 : 4] = clock64();
      }
      ++n_units;
    });
  } else if (warp == kWarpTmaW) {
    // =============== weight stream ===============
    if (elect_one()) {
      uint32_t nf = 0;
      for_each_mlp_unit(P, cta, G, [&](const MlpUnit& u) {
        const int K = u.g == 0 ? P.H : P.I;
        const uint32_t chunk_bytes = 512u + (uint32_t)tps * 4096u;
        const unsigned char* src = (u.g == 0 ? P.chunks_gu : P.chunks_down) + ((size_t)u.nt * (K / P.G) + u.kt0 / tps) * chunk_bytes;
        for (int gi = 0; gi < u.nst; ++gi) {
          const int w = gi & 3;
          const uint32_t n = (nf >> (8 * w)) & 0xffu;
          nf = (nf & ~(0xffu << (8 * w))) | (((n + 1) & 0xffu) << (8 * w));
          const int slot = w + 4 * (int)(n & 1u);
          mbar_wait(&sb->pk_empty[slot], ((n >> 1) & 1u) ^ 1u);
          mbar_arrive_expect_tx(&sb->pk_full[slot], chunk_bytes);
          bulk_g2s(ring + slot * kChunkStride, src + (size_t)gi * chunk_bytes, chunk_bytes, &sb->pk_full[slot]);
        }
      });
    }
  } else if (warp == kWarpTmaX) {
    // =============== activation tiles: xn after the upstream kernel, act after the producing tiles' flags ===============
    if (elect_one()) {
      prefetch_tensormap(&map_xn);
      prefetch_tensormap(&map_act);
      grid_dep_wait();
      if (tr) tr[5] = clock64();
      uint32_t par = 0, n_units = 0;
      for_each_mlp_unit(P, cta, G, [&](const MlpUnit& u) {
        if (n_units > 0) mbar_wait(acc_empty, (n_units - 1) & 1u);     // epilogue staging aliases the X stages
        for (int gi = 0; gi < u.nst; ++gi) {
          const int st = gi & 3;
          const int nt_here = min(tps, u.nkt - gi * tps);
          if (u.g == 1 && gi == 0) {
            // one counter per down split: all nkt gate_up tiles whose columns this unit's K range reads (k tile j
            // == gate_up tile j).  One acquire poll instead of one per tile: an acquire load is an L2 round trip,
            // and twelve of them in sequence cost 6 us.
            flag_wait(P.flags + u.split, (unsigned)u.nkt);
            if (tr) tr[6] = clock64();
          }
          mbar_wait(&sb->empty[st], ((par >> st) & 1u) ^ 1u);
          par ^= 1u << st;
          unsigned char* stage = smem + st * kStageBytes;
          mbar_arrive_expect_tx(&sb->full[st], (uint32_t)(nt_here * kTileBBytes));
          for (int t = 0; t < nt_here; ++t)
            tma_load_2d(stage + t * kTileBBytes, u.g == 0 ? &map_xn : &map_act, (u.kt0 + gi * tps + t) * kUmmaBK, 0, &sb->full[st]);
        }
        ++n_units;
      });
    }
  } else {
    // =============== MMA issuer ===============
    if (elect_one()) {
      constexpr uint32_t idesc = make_idesc(kUmmaM, kMegaNT);
      uint32_t par = 0, n_units = 0;
      for_each_mlp_unit(P, cta, G, [&](const MlpUnit& u) {
        if (n_units > 0) {
          mbar_wait(acc_empty, (n_units - 1) & 1u);
          tc_fence_after();
        }
        for (int gi = 0; gi < u.nst; ++gi) {
          const int st = gi & 3;
          const int nt_here = min(tps, u.nkt - gi * tps);
          mbar_wait(&sb->full[st], (par >> st) & 1u);
          par ^= 1u << st;
          tc_fence_after();
          const uint32_t x_addr = smem_u32(smem + st * kStageBytes);
          for (int t = 0; t < nt_here; ++t) {
            const uint32_t b_addr = x_addr + t * kTileBBytes;
            const uint32_t a_tmem = tmem + (uint32_t)(MegaCfg::kAccCols + 64 * st + 32 * t);
#pragma unroll
            for (int k = 0; k < kUmmaBK / 16; ++k)
              umma_f16_ts(tmem, a_tmem + k * 8, make_smem_desc(b_addr + k * 32), idesc, (gi | t | k) != 0 ? 1u : 0u);
          }
          umma_commit(&sb->empty[st]);
        }
        umma_commit(&sb->tmem_full);
        ++n_units;
      });
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == kWarpMma) tmem_dealloc(tmem, MegaCfg::kTmemCols);
  if (tl_rec) {
    tl = atomicAdd(d_tl.ctr, 1u);
    if (tl < d_tl.cap) {
      d_tl.buf[4 * tl] = (unsigned long long)kKidUmmaW4 | ((unsigned long long)gridDim.x << 8) | (1ull << 28) | (2ull << 48);
      d_tl.buf[4 * tl + 1] = tl_t0;
      d_tl.buf[4 * tl + 2] = tl_t0;
      d_tl.buf[4 * tl + 3] = gtime();
    }
  }
}

// ---- host side ------------------------------------------------------------------------------------------
static unsigned long long* g_mlp_trace = nullptr;
void mlp_fused_set_trace(unsigned long long* dev_buf) { g_mlp_trace = dev_buf; }

bool mlp_fused_supported(int T, int H, int I, int G) {
  return T >= 1 && T <= kMegaNT && (G == 64 || G == 128) && H % kUmmaM == 0 && (2 * I) % kUmmaM == 0 && I % G == 0 && H % G == 0;
}

int launch_mlp_fused(const CUtensorMap& mxn, const CUtensorMap& mact, const MlpParams& P, cudaStream_t s) {
  B200_REQUIRE(mlp_fused_supported(P.T, P.H, P.I, P.G) && P.split_down >= 1, "mlp_fused: unsupported shape");
  static std::mutex mu;
  static int grid_of_device[64] = {0};
  int dev = 0, grid = 0;
  B200_CUDA(cudaGetDevice(&dev));
  {
    std::lock_guard<std::mutex> lk(mu);
    if (dev >= 0 && dev < 64 && grid_of_device[dev] == 0) {
      B200_CUDA(cudaFuncSetAttribute(mlp_fused_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, MlpSmem::kBytes));
      int sms = 0, per_sm = 0;
      B200_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
      B200_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, mlp_fused_kernel, kUmmaThreads, MlpSmem::kBytes));
      B200_REQUIRE(per_sm >= 1 && sms >= 1, "mlp_fused: the kernel does not fit on an SM");
      grid_of_device[dev] = sms;
    }
    grid = dev >= 0 && dev < 64 ? grid_of_device[dev] : 0;
  }
  B200_REQUIRE(grid > 0, "mlp_fused: device index");
  // One CTA per SM: all CTAs are co-resident (nothing else on the stream can hold an SM forever, and the
  // next kernel is only allowed to start once every CTA of this grid runs), which the column flags rely on.
  MlpParams PT = P;
  PT.trace = g_mlp_trace;
  B200_CUDA(launch_k(mlp_fused_kernel, dim3(grid), dim3(kUmmaThreads), (size_t)MlpSmem::kBytes, s, mxn, mact, PT));
  g_launch_count += 1;
  return B200_OK;
}

bool decode_layers_supported(int n_heads, int n_kv, int G, int T) {
  const int gq = n_kv > 0 ? n_heads / n_kv : 0;
  return T >= 1 && T <= kMegaNT && (G == 64 || G == 128) && (gq == 2 || gq == 4 || gq == 6 || gq == 7 || gq == 8);
}
int decode_layers_flag_count(int n_layers) { return n_layers * kMegaPhases; }

static unsigned long long* g_mega_trace = nullptr;
void decode_layers_set_trace(unsigned long long* dev_buf) { g_mega_trace = dev_buf; }

template <int GQ>
static int launch_decode_layers_gq(const CUtensorMap& mxn, const CUtensorMap& mattn, const CUtensorMap& mact,
                                   const MegaParams& P, cudaStream_t s) {
  static std::mutex mu;
  static int grid_of_device[64] = {0};
  auto kern = decode_layers_kernel<GQ>;
  int dev = 0;
  B200_CUDA(cudaGetDevice(&dev));
  int grid = 0;
  {
    std::lock_guard<std::mutex> lk(mu);
    if (dev >= 0 && dev < 64 && grid_of_device[dev] == 0) {
      B200_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, MegaSmem::kBytes));
      int sms = 0, per_sm = 0;
      B200_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
      B200_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, kUmmaThreads, MegaSmem::kBytes));
      B200_REQUIRE(per_sm >= 1 && sms >= 1, "decode_layers: the persistent kernel does not fit on an SM");
      grid_of_device[dev] = sms;
    }
    grid = dev >= 0 && dev < 64 ? grid_of_device[dev] : 0;
  }
  B200_REQUIRE(grid > 0, "decode_layers: device index");
  // cooperative launch: the driver guarantees that all CTAs are co-resident (the phase flags need it)
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(grid); cfg.blockDim = dim3(kUmmaThreads); cfg.dynamicSmemBytes = MegaSmem::kBytes; cfg.stream = s;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeCooperative;
  at[0].val.cooperative = 1;
  cfg.attrs = at; cfg.numAttrs = 1;
  MegaParams PT = P;
  PT.trace = g_mega_trace;
  B200_CUDA(cudaLaunchKernelEx(&cfg, kern, mxn, mattn, mact, PT));
  g_launch_count += 1;
  return B200_OK;
}

int launch_decode_layers(const CUtensorMap& mxn, const CUtensorMap& mattn, const CUtensorMap& mact, const MegaParams& P,
                         cudaStream_t s) {
  B200_REQUIRE(decode_layers_supported(P.n_heads, P.n_kv, P.G, P.T), "decode_layers: unsupported shape");
  B200_REQUIRE(P.H % (4 * kNormCluster) == 0 && P.H % kUmmaM == 0 && P.QKV % kUmmaM == 0 && (2 * P.I) % kUmmaM == 0 &&
                   P.H % P.G == 0 && P.I % P.G == 0 && (P.n_heads * kHeadDim) % P.G == 0,
               "decode_layers: dimensions must be multiples of 128 and of the group size");
  switch (P.n_heads / P.n_kv) {
    case 2: return launch_decode_layers_gq<2>(mxn, mattn, mact, P, s);
    case 4: return launch_decode_layers_gq<4>(mxn, mattn, mact, P, s);
    case 6: return launch_decode_layers_gq<6>(mxn, mattn, mact, P, s);
    case 7: return launch_decode_layers_gq<7>(mxn, mattn, mact, P, s);
    case 8: return launch_decode_layers_gq<8>(mxn, mattn, mact, P, s);
  }
  return B200_ERR_INVALID;
}

// Microbenchmark: n_phases rounds of {write `payload` floats, arrive, wait for every CTA, read the
// neighbour CTA's floats}.  Measures what one dependent phase boundary costs inside a persistent
// kernel (flag round trip through L2 + one L2 read of freshly written data), the number a kernel
// launch boundary (~2.5 us with PDL) has to be compared with.
__global__ void __launch_bounds__(512, 1)
gridsync_bench_kernel(unsigned int* flags, int n_phases, float* buf, int payload, float* sink) {
  const int cta = blockIdx.x, G = gridDim.x, tid = threadIdx.x;
  float acc = 0.f;
  for (int p = 0; p < n_phases; ++p) {
    float* mine = buf + ((size_t)(p & 1) * G + cta) * payload;
    for (int i = tid; i < payload; i += blockDim.x) mine[i] = (float)(p + i) + acc * 1e-30f;
    __syncthreads();
    if (tid == 0) {
      flag_arrive(flags + p);
      flag_wait(flags + p, (unsigned)G);
    }
    __syncthreads();
    const float* theirs = buf + ((size_t)(p & 1) * G + (cta + 1) % G) * payload;
    for (int i = tid; i < payload; i += blockDim.x) acc += __ldcg(theirs + i);
  }
  if (acc == 12345.678f) sink[0] = acc;      // keep the loads alive
}

int launch_gridsync_bench(unsigned int* flags, int n_phases, float* buf, int payload, int n_ctas, float* sink, cudaStream_t s) {
  B200_REQUIRE(n_ctas >= 1 && n_ctas <= 148 && n_phases >= 1 && payload >= 0, "gridsync bench: 1 <= ctas <= 148");
  B200_CUDA(cudaMemsetAsync(flags, 0, (size_t)n_phases * 4, s));
  gridsync_bench_kernel<<<n_ctas, 512, 0, s>>>(flags, n_phases, buf, payload, sink);
  B200_CUDA(cudaGetLastError());
  return B200_OK;
}

}  // namespace b200
