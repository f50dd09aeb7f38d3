// Paged-KV GQA attention, one query row per (t): decode rows, or prefill rows issued as rows with
// ctx_len = position + 1.  Replaces vLLM's attention backend for this path
// (vllm/model_executor/models/qwen2.py:120-236 -> Attention -> FlashInfer/FlashAttention).
//
// grid (n_kv_heads, T, n_splits), 128 threads.  Split s of row t owns the 64-token chunks
// c = s, s + n_splits, ...  of the row's context (flash-decoding): a 256-token context is scored by
// four CTAs at once instead of one CTA walking four chunks.  KV blocks (16 tokens x 128 x fp16 =
// 4 KB, contiguous per (block, kv head)) are staged into shared memory by 1-D bulk TMA copies
// (cp.async.bulk, SASS UBLKCP) completing on an mbarrier.  All GQ = n_heads / n_kv_heads query
// heads of the group share each staged chunk.  Scores / softmax / PV in fp32.
//
// Each split writes (acc, max, sum) to scratch; the last split to finish (per-(t, kv head) counter)
// merges them in split order -- deterministic -- and writes the fp16 output.
//
// Decode mode additionally fuses RoPE and the KV-cache write (vLLM qwen2.py:181 get_rope,
// reshape_and_cache): q / k / v of the new token arrive as fp32 split-K partials of the qkv
// projection; every CTA rotates its own q heads, and the CTA that owns the chunk holding the new
// position rotates k, writes k / v to the paged cache and patches them into its staged tile.
#pragma once
#include "common.cuh"
#include "internal.h"

namespace b200 {

constexpr int kAttnThreads = 128;
constexpr int kAttnWarps = 4;
// Warps that own a K/V tile pair and score chunks (a warp beyond this count only helps with the prologue
// and the merge).  With 3 the CTA drops to ~101 KB of shared memory, two CTAs share an SM, and the CTAs
// of the neighbouring GEMM kernels can all start (and prefetch weights) while attention runs -- with
// programmatic dependent launch a kernel starts only once EVERY CTA of its predecessor has started.
// Measured at batch 64, context 128..384: the downstream kernels did start ~10 us earlier, but their
// post-dependency part (activation tiles through L2, MMA, epilogue) is what takes the time, and the
// attention itself lost a quarter of its warps: 1.249 ms per step against 1.238 ms with 4.
constexpr int kAttnTileWarps = 4;
constexpr int kChunkTokens = 64;
constexpr int kChunkBlocks = kChunkTokens / kBlockTokens;            // 4
constexpr int kKvBlockBytes = kBlockTokens * kHeadDim * 2;           // 4096
constexpr int kTileBytes = kChunkTokens * kHeadDim * 2;              // 16384 per K or V tile
constexpr int kPartFloats = kHeadDim + 2;                            // acc[128], max, sum

// Shared memory: per warp one K tile + one V tile (a warp owns whole 64-token chunks), then the
// query tile [16 rows][128] fp16 (rows >= GQ are zero) and the new token's k / v (fused mode).
template <int NW>                                                    // NW = warps that own a K/V tile pair
struct AttnSmemT {
  static constexpr int kOffTiles = 0;                                // NW x (K 16 KB + V 16 KB)
  static constexpr int kOffQ = NW * 2 * kTileBytes;                  // 4 KB
  static constexpr int kOffNew = kOffQ + 16 * kHeadDim * 2;          // k[128], v[128] fp16
  static constexpr int kOffBar = kOffNew + 2 * kHeadDim * 2;
  static constexpr int kBytes = kOffBar + NW * 8;
};
using AttnSmem = AttnSmemT<4>;



struct AttnParams {
  unsigned long long* trace;
  const __half* q;               // [T, nh * 128] rotated queries (plain mode) or null
  // fused decode mode: qkv fp32 [T, (nh + 2 nkv) * 128] as n_parts partial tensors (+ bias)
  const float* qkv;
  int n_parts;
  int64_t part_stride;
  const __half* bias;
  const int32_t* pos;
  const int32_t* slot;
  const float* cos_t;
  const float* sin_t;
  __half* k_cache;
  __half* v_cache;
  const int32_t* block_tables;
  int bt_stride;
  const int32_t* bt_row;
  const int32_t* ctx_len;
  int n_heads, n_kv, n_splits;
  float* scratch;                // [T, n_kv, n_splits, GQ, 130]
  int32_t* counters;             // [T, n_kv], zero before and after every launch
  __half* out;                   // [T, nh * 128]
  int bf16;                      // 16-bit storage type of q / K / V / out: 0 fp16, 1 bf16
};

__device__ __forceinline__ void ldmatrix_x4(uint32_t saddr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(saddr));
}
__device__ __forceinline__ void ldmatrix_x4_trans(uint32_t saddr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(saddr));
}
// D (fp32, 16x8) += A (16-bit, 16x16, row) * B (16-bit, 16x8, col); bf (uniform per launch) selects bf16 operands
__device__ __forceinline__ void mma_16816(float d[4], const uint32_t a[4], uint32_t b0, uint32_t b1, const bool bf = false) {
  if (bf)
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
  else
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

// One warp scores whole 64-token chunks with tensor-core MMAs (mma.sync m16n8k16; the 6..8 query
// heads of the group are the M rows, padded to 16):
//     S[16 heads, 64 tok] = Q K^T      (A = Q from smem via ldmatrix, B = K rows via ldmatrix)
//     online softmax in registers (fp32), P rounded to fp16 as the next A operand
//     O[16 heads, 128]  += P V          (B = V via ldmatrix.trans)
// Rows of a KV block are stored with their 16-byte chunks XOR-swizzled by (token & 7) -- in the cache
// itself -- so that the bulk copy needs no re-layout and every ldmatrix is bank-conflict free.
// One unit = one (kv head j, row t, split) of the decode attention; 128 threads (4 warps) cooperate.
// Called by paged_attn_kernel (one unit per CTA, kBarId 0 = the CTA barrier) and by the persistent
// decode kernel (an experiment that lives in the history, commit 8d75821: warps 0..3 of the CTA, a named barrier, mbarriers that live across units:
// `bar_uses` counts the completed phases of this warp's barrier; kInitBars = false there).
template <int kBarId>
__device__ __forceinline__ void attn_bar_sync() {
  asm volatile("bar.sync %0, %1;" ::"n"(kBarId), "n"(kAttnThreads) : "memory");
}

// kPreWait (decode kernel only): the K/V tiles of each warp's first chunk are requested BEFORE the
// programmatic-dependency wait, from metadata that may still be the previous step's (the advance kernel
// of this step may not have finished): K/V of old tokens do not depend on the upstream qkv GEMM, and
// they are the bulk of the kernel's HBM traffic.  After the wait the metadata is read again; if the
// block ids or the block count changed (a sequence crossed into a new block, or the batch changed), the
// prefetched tile is dropped (its barrier phase is consumed) and the chunk is loaded again.  Requires
// ctx_len / bt_row / block_tables to hold valid values at all times (the engine zero-fills them).
template <int GQ, bool kFused, int kBarId, bool kInitBars, bool kPreWait = false, int kPhys = kAttnTileWarps>
__device__ __forceinline__ void paged_attn_unit(const AttnParams& p, unsigned char* smem, const int j, const int t,
                                                const int split, unsigned long long* tr, uint32_t& bar_uses) {
  using L = AttnSmemT<kPhys>;
  static_assert(kPhys == 3 || kPhys == 4, "physical tile warps");
  static_assert(GQ <= 8, "one m16 tile, upper half unused");
  static_assert(!kPreWait || kInitBars, "the prefetch owns phase 0 of a fresh barrier");
  const int tid = threadIdx.x;
  const int warp = tid >> 5, lane = tid & 31;
  if (tr && tid == 0) tr[0] = clock64();
  const int n_heads = p.n_heads, n_kv = p.n_kv, S = p.n_splits;
  const bool bf = p.bf16 != 0;
  unsigned char* q_s = smem + L::kOffQ;
  uint16_t* new_s = reinterpret_cast<uint16_t*>(smem + L::kOffNew);
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem + L::kOffBar) + warp;
  unsigned char* Kt = smem + L::kOffTiles + warp * 2 * kTileBytes;
  unsigned char* Vt = Kt + kTileBytes;
  const int c0 = split + S * warp;                  // this warp's first chunk

  int pre_nb = 0;                                   // lane 0: blocks requested before the wait, and their ids
  int pre_ids[kChunkBlocks] = {0, 0, 0, 0};
  const bool tile_warp = warp < kPhys;
  if (lane == 0 && tile_warp) {
    if constexpr (kInitBars) {
      mbar_init(bar, 1);
      fence_mbar_init();
    }
    if constexpr (kPreWait) {
      const int len0 = max(0, __ldcg(p.ctx_len + t));
      const int nblocks0 = min((len0 + kBlockTokens - 1) / kBlockTokens, p.bt_stride);
      if (c0 * kChunkBlocks < nblocks0) {
        const int32_t* bt0 = p.block_tables + (size_t)__ldcg(p.bt_row + t) * p.bt_stride + c0 * kChunkBlocks;
        pre_nb = min(kChunkBlocks, nblocks0 - c0 * kChunkBlocks);
        mbar_arrive_expect_tx(bar, (uint32_t)pre_nb * 2u * kKvBlockBytes);
#pragma unroll
        for (int i = 0; i < kChunkBlocks; ++i)
          if (i < pre_nb) {
            pre_ids[i] = __ldcg(bt0 + i);
            const size_t off = ((size_t)pre_ids[i] * n_kv + j) * (size_t)(kBlockTokens * kHeadDim);
            bulk_g2s(Kt + i * kKvBlockBytes, p.k_cache + off, kKvBlockBytes, bar);
            bulk_g2s(Vt + i * kKvBlockBytes, p.v_cache + off, kKvBlockBytes, bar);
          }
      }
    }
  }
  // Fused mode: what a thread needs besides the qkv partials -- column indices, bias, and the cos / sin
  // row of the new token's position -- is loaded here, before the dependency wait when kPreWait (bias is a
  // weight; the position is ctx_len - 1, possibly still the previous step's: re-checked after the wait).
  constexpr int kPairs = (GQ * 64 + kAttnThreads - 1) / kAttnThreads;        // rotation pairs per thread (3 for GQ = 6)
  constexpr int NE = 2 * kPairs + 3;                                          // + v[tid], k pair (tid & 63)
  int cols[NE];
  float bia[NE], cs[kPairs + 1], sn[kPairs + 1];
  int rope_pos = -1;
  auto load_rope = [&](int ps) {
    rope_pos = ps;
#pragma unroll
    for (int q3 = 0; q3 <= kPairs; ++q3) {
      const int i = q3 < kPairs ? (min(tid + q3 * kAttnThreads, GQ * 64 - 1) & 63) : (tid & 63);
      cs[q3] = p.cos_t[(size_t)ps * 64 + i];
      sn[q3] = p.sin_t[(size_t)ps * 64 + i];
    }
  };
  if constexpr (kFused) {
#pragma unroll
    for (int q3 = 0; q3 < kPairs; ++q3) {
      const int idx = min(tid + q3 * kAttnThreads, GQ * 64 - 1);
      cols[2 * q3] = (j * GQ + (idx >> 6)) * kHeadDim + (idx & 63);
      cols[2 * q3 + 1] = cols[2 * q3] + 64;
    }
    cols[2 * kPairs] = (n_heads + n_kv + j) * kHeadDim + tid;                 // v
    cols[2 * kPairs + 1] = (n_heads + j) * kHeadDim + (tid & 63);            // k (first half of the pair)
    cols[2 * kPairs + 2] = cols[2 * kPairs + 1] + 64;
#pragma unroll
    for (int e = 0; e < NE; ++e) bia[e] = p.bias ? from16(reinterpret_cast<const uint16_t*>(p.bias)[cols[e]], bf) : 0.f;
    if constexpr (kPreWait) {
      const int len0 = __ldcg(p.ctx_len + t);
      if (len0 > 0) load_rope(len0 - 1);
    }
  }
  if constexpr (kPreWait) grid_dep_wait();          // from here on the upstream kernels' results are visible

  const int len = kPreWait ? __ldcg(p.ctx_len + t) : p.ctx_len[t];
  uint16_t* orow = reinterpret_cast<uint16_t*>(p.out) + ((size_t)t * n_heads + (size_t)j * GQ) * kHeadDim;
  const int nchunks = len > 0 ? (len + kChunkTokens - 1) / kChunkTokens : 0;
  const int n_active = min(S, nchunks);
  const int32_t* bt = p.block_tables + (size_t)(kPreWait ? __ldcg(p.bt_row + t) : p.bt_row[t]) * p.bt_stride;
  const int nblocks = len > 0 ? (len + kBlockTokens - 1) / kBlockTokens : 0;

  // chunks of this CTA: c = split + S * i; warp w takes i = w, w + 4, ...
  auto issue_chunk = [&](int c) {   // lane 0 of the owning warp
    const int b0 = c * kChunkBlocks;
    const int nb = min(kChunkBlocks, nblocks - b0);
    mbar_arrive_expect_tx(bar, (uint32_t)nb * 2u * kKvBlockBytes);
    for (int i = 0; i < nb; ++i) {
      const size_t off = ((size_t)bt[b0 + i] * n_kv + j) * (size_t)(kBlockTokens * kHeadDim);
      bulk_g2s(Kt + i * kKvBlockBytes, p.k_cache + off, kKvBlockBytes, bar);
      bulk_g2s(Vt + i * kKvBlockBytes, p.v_cache + off, kKvBlockBytes, bar);
    }
  };
  int dropped = 0;
  if (lane == 0 && tile_warp) {
    const int nb = c0 < nchunks ? min(kChunkBlocks, nblocks - c0 * kChunkBlocks) : 0;
    bool keep = kPreWait && nb == pre_nb;
    if constexpr (kPreWait) {
#pragma unroll
      for (int i = 0; i < kChunkBlocks; ++i)
        if (i < nb && keep) keep = __ldcg(bt + c0 * kChunkBlocks + i) == pre_ids[i];
      if (!keep && pre_nb > 0) {                    // stale prefetch: let it land, consume its phase
        mbar_wait(bar, 0u);
        dropped = 1;
      }
    }
    if (!keep && nb > 0) issue_chunk(c0);
  }
  if constexpr (kPreWait) bar_uses += (uint32_t)__shfl_sync(0xffffffffu, dropped, 0);
  if (len <= 0) {
    if (split == 0) {
#pragma unroll
      for (int h = 0; h < GQ; ++h) orow[h * kHeadDim + tid] = 0;
    }
    return;
  }
  if (split >= nchunks) return;                                     // nothing to score for this split
  if (tr && tid == 0) tr[1] = clock64();

  // ---- queries -> swizzled smem tile (and, in fused mode, the new token's k / v) ----
  // q_s row h (256 B), logical 16-byte chunk cc -> physical chunk cc ^ (h & 7)
  auto q_addr = [&](int h, int dim) { return q_s + h * 256 + ((((dim >> 3) ^ (h & 7)) << 4) | ((dim & 7) << 1)); };
  for (int i = tid; i < 16 * 16; i += kAttnThreads)                  // zero rows GQ..15 (and everything else first)
    if ((i >> 4) >= GQ) reinterpret_cast<uint4*>(q_s)[i] = make_uint4(0, 0, 0, 0);
  const int new_pos = len - 1;                      // fused: position of the token being decoded
  const int new_chunk = new_pos / kChunkTokens;
  if constexpr (kFused) {
    // Everything this thread needs is loaded up front -- split-K partials of its q pairs and of
    // the new k / v, bias, cos / sin -- so the prologue costs two dependent L2 round trips (position,
    // then data), not one per partial.
    const int total = (n_heads + 2 * n_kv) * kHeadDim;
    const float* row = p.qkv + (size_t)t * total;
    constexpr int kMaxParts = 4;                                   // loaded in one batch; more are added after
    const int ps = p.pos[t];
    const int sl = p.slot[t];
    const bool own = (new_chunk % S) == split;                     // this CTA owns the chunk that holds the new token
    float part[NE][kMaxParts];
#pragma unroll
    for (int e = 0; e < NE; ++e) {
#pragma unroll
      for (int pp = 0; pp < kMaxParts; ++pp)
        part[e][pp] = pp < p.n_parts ? __ldcg(row + (size_t)pp * p.part_stride + cols[e]) : 0.f;
    }
    if (ps != rope_pos) load_rope(ps);                             // not prefetched, or prefetched for a stale position
    float val[NE];
#pragma unroll
    for (int e = 0; e < NE; ++e) {
      float v = part[e][0];
#pragma unroll
      for (int pp = 1; pp < kMaxParts; ++pp) v += part[e][pp];                // fixed order: deterministic
      for (int pp = kMaxParts; pp < p.n_parts; ++pp) v += __ldcg(row + (size_t)pp * p.part_stride + cols[e]);
      val[e] = v + bia[e];
    }
#pragma unroll
    for (int q3 = 0; q3 < kPairs; ++q3) {
      const int idx = tid + q3 * kAttnThreads;
      if (idx < GQ * 64) {
        const int h = idx >> 6, i = idx & 63;
        float r1, r2;
        rope_rotate(val[2 * q3], val[2 * q3 + 1], cs[q3], sn[q3], r1, r2);
        *reinterpret_cast<uint16_t*>(q_addr(h, i)) = to16(r1, bf);
        *reinterpret_cast<uint16_t*>(q_addr(h, i + 64)) = to16(r2, bf);
      }
    }
    if (own) {
      const uint16_t vnew = to16(val[2 * kPairs], bf);
      new_s[kHeadDim + tid] = vnew;
      float r1, r2;
      rope_rotate(val[2 * kPairs + 1], val[2 * kPairs + 2], cs[kPairs], sn[kPairs], r1, r2);
      const uint16_t k1 = to16(r1, bf), k2 = to16(r2, bf);
      if (tid < 64) { new_s[tid] = k1; new_s[tid + 64] = k2; }
      if (sl >= 0) {                                // paged cache write (swizzled row)
        const int r = sl % kBlockTokens;
        const size_t rowoff = (((size_t)(sl / kBlockTokens) * n_kv + j) * kBlockTokens + r) * kHeadDim;
        reinterpret_cast<uint16_t*>(p.v_cache)[rowoff + kv_swz(r, tid)] = vnew;
        if (tid < 64) {
          reinterpret_cast<uint16_t*>(p.k_cache)[rowoff + kv_swz(r, tid)] = k1;
          reinterpret_cast<uint16_t*>(p.k_cache)[rowoff + kv_swz(r, tid + 64)] = k2;
        }
      }
    }
  } else {
    const __half* qg = p.q + ((size_t)t * n_heads + (size_t)j * GQ) * kHeadDim;
    for (int i = tid; i < GQ * 16; i += kAttnThreads) {              // 16-byte chunks
      const int h = i >> 4, cc = i & 15;
      *reinterpret_cast<uint4*>(q_s + h * 256 + ((cc ^ (h & 7)) << 4)) = *reinterpret_cast<const uint4*>(qg + h * kHeadDim + cc * 8);
    }
  }
  attn_bar_sync<kBarId>();

  // Q as the MMA A operand: 8 k-steps x {a0a1, a2a3, a4a5, a6a7}
  uint32_t qa[8][4];
  {
    const int mj = lane >> 3, mr = lane & 7;        // ldmatrix.x4: lane -> (matrix, row)
    const int h = (mj & 1) * 8 + mr;
#pragma unroll
    for (int ks = 0; ks < 8; ++ks) {
      const int cc = 2 * ks + (mj >> 1);
      ldmatrix_x4(smem_u32(q_s + h * 256 + ((cc ^ (h & 7)) << 4)), qa[ks][0], qa[ks][1], qa[ks][2], qa[ks][3]);
    }
  }

  const int g = lane >> 2, tq = lane & 3;           // MMA fragment coordinates
  const float scale = 0.08838834764831845f;         // 1 / sqrt(128)
  // The chunks of a (row, kv head, split) are dealt to kAttnTileWarps VIRTUAL warps, c = split + S * (vw + 4 i),
  // each with its own online-softmax state; the merge below combines the virtual warps in order.  With
  // kPhys == 4 a physical warp is a virtual warp.  With kPhys == 3 (three tile pairs: ~101 KB, two CTAs per SM,
  // used when there are more CTAs than SMs) warp 0 also plays virtual warp 3 with a second state -- the
  // arithmetic, hence the bits, do not depend on kPhys.
  auto run_chunks = [&](const int vw, const bool first_issued, float (&o)[16][2], float& m_run, float& l_run) {
  int it = 0;
  for (int c = split + S * vw; c < nchunks; c += S * kAttnTileWarps, ++it) {
    if (it > 0 || !first_issued) {
      __syncwarp();
      if (lane == 0) issue_chunk(c);
    }
    if (tr && tid == 0 && it == 0) tr[2] = clock64();
    mbar_wait(bar, bar_uses & 1u);
    ++bar_uses;
    if (tr && tid == 0 && it == 0) tr[3] = clock64();
    const int nvalid = min(kChunkTokens, len - c * kChunkTokens);
    const int nb = (nvalid + kBlockTokens - 1) / kBlockTokens;
    if constexpr (kFused) {
      if (c == new_chunk) {                         // the cache row of the new token is not in HBM yet: patch the tile
        const int rr = new_pos - c * kChunkTokens, r = rr % kBlockTokens;
        unsigned char* kr = Kt + (rr / kBlockTokens) * kKvBlockBytes + r * 256;
        unsigned char* vr = Vt + (rr / kBlockTokens) * kKvBlockBytes + r * 256;
        if (lane < 16) {
          *reinterpret_cast<uint4*>(kr + ((lane ^ (r & 7)) << 4)) = reinterpret_cast<const uint4*>(new_s)[lane];
          *reinterpret_cast<uint4*>(vr + ((lane ^ (r & 7)) << 4)) = reinterpret_cast<const uint4*>(new_s + kHeadDim)[lane];
        }
        __syncwarp();
      }
    }

    // ---- S = Q K^T : 8 token tiles of 8 ----
    // The eight token tiles are independent accumulation chains; they are issued interleaved (k outer,
    // tile inner) because a single warp per scheduler has nothing else to hide the MMA latency with, and
    // there is no "tile has valid tokens" branch for the same reason.  Rows beyond the context (or in
    // blocks that were not loaded) hold stale data; their scores are replaced by -inf below.
    float sc[8][2];
    {
      float d[8][4];
#pragma unroll
      for (int nt = 0; nt < 8; ++nt) d[nt][0] = d[nt][1] = d[nt][2] = d[nt][3] = 0.f;
      const int mj = lane >> 3, tl7 = lane & 7;
#pragma unroll
      for (int k2 = 0; k2 < 4; ++k2) {              // two k-steps per ldmatrix.x4
        uint32_t kb[8][4];
#pragma unroll
        for (int nt = 0; nt < 8; ++nt) {
          const int tok = nt * 8 + tl7;
          const uint32_t rowaddr = smem_u32(Kt + (tok >> 4) * kKvBlockBytes + (tok & 15) * 256);
          ldmatrix_x4(rowaddr + ((((4 * k2 + mj) ^ (tok & 7))) << 4), kb[nt][0], kb[nt][1], kb[nt][2], kb[nt][3]);
        }
#pragma unroll
        for (int nt = 0; nt < 8; ++nt) mma_16816(d[nt], qa[2 * k2], kb[nt][0], kb[nt][1], bf);
#pragma unroll
        for (int nt = 0; nt < 8; ++nt) mma_16816(d[nt], qa[2 * k2 + 1], kb[nt][2], kb[nt][3], bf);
      }
#pragma unroll
      for (int nt = 0; nt < 8; ++nt) {
        const int t0 = nt * 8 + 2 * tq;
        sc[nt][0] = t0 < nvalid ? __fmul_rn(d[nt][0], scale) : -INFINITY;
        sc[nt][1] = t0 + 1 < nvalid ? __fmul_rn(d[nt][1], scale) : -INFINITY;
      }
    }
    if (tr && tid == 0 && it == 0) tr[4] = clock64();

    // ---- online softmax for row g (the quad holds the row's 64 scores) ----
    float mx = -INFINITY;
#pragma unroll
    for (int nt = 0; nt < 8; ++nt) mx = fmaxf(mx, fmaxf(sc[nt][0], sc[nt][1]));
    mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
    mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
    const float m_new = fmaxf(m_run, mx);
    const float corr = expf(__fsub_rn(m_run, m_new));
    float sum = 0.f;
    uint32_t pa[4][4];                              // P as A operand: 4 k-steps of 16 tokens; rows g + 8 are zero
#pragma unroll
    for (int nt = 0; nt < 8; ++nt) {
      const float p0 = expf(__fsub_rn(sc[nt][0], m_new)), p1 = expf(__fsub_rn(sc[nt][1], m_new));
      sum = __fadd_rn(sum, __fadd_rn(p0, p1));
      pa[nt >> 1][(nt & 1) * 2] = pack16x2(p0, p1, bf);
      pa[nt >> 1][(nt & 1) * 2 + 1] = 0u;
    }
    sum += __shfl_xor_sync(0xffffffffu, sum, 1);
    sum += __shfl_xor_sync(0xffffffffu, sum, 2);
    l_run = __fmaf_rn(l_run, corr, sum);             // explicit: both callers (kernel, persistent kernel) round alike
    m_run = m_new;
#pragma unroll
    for (int dn = 0; dn < 16; ++dn) { o[dn][0] = __fmul_rn(o[dn][0], corr); o[dn][1] = __fmul_rn(o[dn][1], corr); }
    if (tr && tid == 0 && it == 0) tr[5] = clock64();

    // ---- O += P V : k-step kt = KV block kt of the chunk ----
#pragma unroll
    for (int kt = 0; kt < 4; ++kt) {
      if (kt < nb) {
        const int mj = lane >> 3, r = (mj & 1) * 8 + (lane & 7);
        const uint32_t rowaddr = smem_u32(Vt + kt * kKvBlockBytes + r * 256);
#pragma unroll
        uint32_t vb[8][4];                          // all eight loads first, then sixteen independent MMAs
#pragma unroll
        for (int d2 = 0; d2 < 8; ++d2)              // two 8-dim tiles per ldmatrix.x4.trans
          ldmatrix_x4_trans(rowaddr + (((2 * d2 + (mj >> 1)) ^ (r & 7)) << 4), vb[d2][0], vb[d2][1], vb[d2][2], vb[d2][3]);
#pragma unroll
        for (int d2 = 0; d2 < 8; ++d2) {
          float d[4] = {o[2 * d2][0], o[2 * d2][1], 0.f, 0.f};
          mma_16816(d, pa[kt], vb[d2][0], vb[d2][1], bf);
          o[2 * d2][0] = d[0]; o[2 * d2][1] = d[1];
          float e[4] = {o[2 * d2 + 1][0], o[2 * d2 + 1][1], 0.f, 0.f};
          mma_16816(e, pa[kt], vb[d2][2], vb[d2][3], bf);
          o[2 * d2 + 1][0] = e[0]; o[2 * d2 + 1][1] = e[1];
        }
      }
    }
  }
  };
  float oA[16][2], oB[16][2];                       // O[row g][dims 8 dn + 2 tq, +1]  (rows g + 8 are padding)
#pragma unroll
  for (int dn = 0; dn < 16; ++dn) oA[dn][0] = oA[dn][1] = oB[dn][0] = oB[dn][1] = 0.f;
  float mA = -INFINITY, lA = 0.f, mB = -INFINITY, lB = 0.f;   // running max / sum of row g (replicated over the quad)
  const bool second = kPhys < kAttnTileWarps && warp + kPhys < kAttnTileWarps;
  if (tile_warp) run_chunks(warp, true, oA, mA, lA);
  if constexpr (kPhys < kAttnTileWarps) {
    if (second) run_chunks(warp + kPhys, false, oB, mB, lB);
  }
  if (tr && tid == 0) tr[6] = clock64();

  // ---- merge the four virtual warps in shared memory (the tiles are free now) ----
  attn_bar_sync<kBarId>();
  auto put = [&](const int vw, const float (&o)[16][2], const float m_run, const float l_run) {
    float* mrg = reinterpret_cast<float*>(smem + L::kOffTiles) + (size_t)vw * (8 * kPartFloats);   // [8 rows][130]
    if (g < GQ) {
#pragma unroll
      for (int dn = 0; dn < 16; ++dn) {
        mrg[g * kPartFloats + dn * 8 + 2 * tq] = o[dn][0];
        mrg[g * kPartFloats + dn * 8 + 2 * tq + 1] = o[dn][1];
      }
      if (tq == 0) { mrg[g * kPartFloats + kHeadDim] = m_run; mrg[g * kPartFloats + kHeadDim + 1] = l_run; }
    }
  };
  if (tile_warp) put(warp, oA, mA, lA);
  if constexpr (kPhys < kAttnTileWarps) {
    if (second) put(warp + kPhys, oB, mB, lB);
  }
  attn_bar_sync<kBarId>();
  const float* all = reinterpret_cast<const float*>(smem + L::kOffTiles);
  float num[GQ], den[GQ], mxh[GQ];
#pragma unroll
  for (int h = 0; h < GQ; ++h) {
    float M = -INFINITY;
#pragma unroll
    for (int w = 0; w < kAttnTileWarps; ++w) M = fmaxf(M, all[(w * 8 + h) * kPartFloats + kHeadDim]);
    float nu = 0.f, de = 0.f;
#pragma unroll
    for (int w = 0; w < kAttnTileWarps; ++w) {      // fixed warp order: deterministic
      const float* pw = all + (w * 8 + h) * kPartFloats;
      const float wgt = expf(pw[kHeadDim] - M);     // a warp without chunks has max = -inf -> weight 0
      nu = fmaf(pw[tid], wgt, nu);
      de = fmaf(pw[kHeadDim + 1], wgt, de);
    }
    num[h] = nu; den[h] = de; mxh[h] = M;
  }
  if (n_active == 1) {                              // single split: finished
#pragma unroll
    for (int h = 0; h < GQ; ++h) orow[h * kHeadDim + tid] = to16(num[h] / den[h], bf);
    return;
  }

  // ---- several splits: publish this CTA's partial; the last CTA to arrive merges them all ----
  float* base = p.scratch + ((size_t)(t * n_kv + j) * S) * GQ * kPartFloats;
  float* mine = base + (size_t)split * GQ * kPartFloats;
#pragma unroll
  for (int h = 0; h < GQ; ++h) {
    mine[h * kPartFloats + tid] = num[h];
    if (tid == 0) { mine[h * kPartFloats + kHeadDim] = mxh[h]; mine[h * kPartFloats + kHeadDim + 1] = den[h]; }
  }
  __threadfence();
  attn_bar_sync<kBarId>();
  __shared__ int s_last;
  if (tid == 0) {
    const int old = atomicAdd(&p.counters[t * n_kv + j], 1);
    s_last = (old == n_active - 1);
    if (s_last) p.counters[t * n_kv + j] = 0;       // self-cleaning for the next launch
  }
  attn_bar_sync<kBarId>();
  if (!s_last) return;
  __threadfence();
  float pm[GQ][8], pl[GQ][8];                       // S <= 64 in general; merged 8 at a time below
#pragma unroll
  for (int h = 0; h < GQ; ++h) {
    float M = -INFINITY;
    for (int s0 = 0; s0 < n_active; ++s0) M = fmaxf(M, __ldcg(base + ((size_t)s0 * GQ + h) * kPartFloats + kHeadDim));
    float nu = 0.f, de = 0.f;
    for (int s0 = 0; s0 < n_active; s0 += 8) {      // fixed split order: deterministic; 8 loads in flight
      const int nn = min(8, n_active - s0);
#pragma unroll
      for (int k = 0; k < 8; ++k)
        if (k < nn) {
          const float* ps = base + ((size_t)(s0 + k) * GQ + h) * kPartFloats;
          pm[h][k] = __ldcg(ps + tid);
          pl[h][k] = expf(__ldcg(ps + kHeadDim) - M);
        }
#pragma unroll
      for (int k = 0; k < 8; ++k)
        if (k < nn) {
          const float* ps = base + ((size_t)(s0 + k) * GQ + h) * kPartFloats;
          nu = fmaf(pm[h][k], pl[h][k], nu);
          de = fmaf(__ldcg(ps + kHeadDim + 1), pl[h][k], de);
        }
    }
    orow[h * kHeadDim + tid] = to16(nu / de, bf);
  }
}


}  // namespace b200
