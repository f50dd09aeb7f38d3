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
#include <mutex>

#include "common.cuh"
#include "internal.h"

namespace b200 {

constexpr int kAttnThreads = 128;
constexpr int kChunkTokens = 64;
constexpr int kChunkBlocks = kChunkTokens / kBlockTokens;            // 4
constexpr int kKvBlockBytes = kBlockTokens * kHeadDim * 2;           // 4096
constexpr int kStageBytes = kChunkTokens * kHeadDim * 2;             // 16384 per K or V tile
constexpr int kPartFloats = kHeadDim + 2;                            // acc[128], max, sum

template <int GQ>
struct AttnSmem {
  static constexpr int kOffK = 0;
  static constexpr int kOffV = kStageBytes;
  static constexpr int kOffQ = 2 * kStageBytes;                      // [GQ][128] half
  static constexpr int kOffP = kOffQ + GQ * kHeadDim * 2;            // [GQ][64] float
  static constexpr int kOffM = kOffP + GQ * kChunkTokens * 4;        // m[GQ], l[GQ], corr[GQ]
  static constexpr int kOffBar = (kOffM + 3 * GQ * 4 + 15) / 16 * 16;
  static constexpr int kBytes = kOffBar + 16;
};

static unsigned long long* g_attn_trace = nullptr;     // profiling aid: 8 x uint64 clock stamps per CTA
void attn_set_trace(unsigned long long* dev_buf) { g_attn_trace = dev_buf; }

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
};

template <int GQ, bool kFused>
__global__ void __launch_bounds__(kAttnThreads)
paged_attn_kernel(const AttnParams p) {
  extern __shared__ __align__(128) unsigned char smem[];
  using L = AttnSmem<GQ>;
  pdl_prologue();
  const int j = blockIdx.x, t = blockIdx.y, split = blockIdx.z, tid = threadIdx.x;
  unsigned long long* tr = p.trace ? p.trace + 8 * (size_t)(blockIdx.x + gridDim.x * (blockIdx.y + gridDim.y * blockIdx.z)) : nullptr;
  if (tr && tid == 0) tr[0] = clock64();
  const int n_heads = p.n_heads, n_kv = p.n_kv, S = p.n_splits;
  const int len = p.ctx_len[t];
  __half* orow = p.out + ((size_t)t * n_heads + (size_t)j * GQ) * kHeadDim;
  if (len <= 0) {
    if (split == 0) {
#pragma unroll
      for (int h = 0; h < GQ; ++h) orow[h * kHeadDim + tid] = __float2half_rn(0.f);
    }
    return;
  }
  const int nchunks = (len + kChunkTokens - 1) / kChunkTokens;
  if (split >= nchunks) return;                                     // nothing to score for this split
  const int n_active = min(S, nchunks);
  const int32_t* bt = p.block_tables + (size_t)p.bt_row[t] * p.bt_stride;
  const int nblocks = (len + kBlockTokens - 1) / kBlockTokens;

  __half* q_s = reinterpret_cast<__half*>(smem + L::kOffQ);
  float* p_s = reinterpret_cast<float*>(smem + L::kOffP);
  float* m_s = reinterpret_cast<float*>(smem + L::kOffM);
  float* l_s = m_s + GQ;
  float* c_s = l_s + GQ;
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem + L::kOffBar);
  __half* Ks_h = reinterpret_cast<__half*>(smem + L::kOffK);
  __half* Vs_h = reinterpret_cast<__half*>(smem + L::kOffV);

  if (tid == 0) { mbar_init(bar, 1); fence_mbar_init(); }
  if (tid < GQ) { m_s[tid] = -INFINITY; l_s[tid] = 0.f; }

  auto issue_chunk = [&](int c) {   // one thread
    const int b0 = c * kChunkBlocks;
    const int nb = min(kChunkBlocks, nblocks - b0);
    mbar_arrive_expect_tx(bar, (uint32_t)nb * 2u * kKvBlockBytes);
    for (int i = 0; i < nb; ++i) {
      const size_t off = ((size_t)bt[b0 + i] * n_kv + j) * (size_t)(kBlockTokens * kHeadDim);
      bulk_g2s(smem + L::kOffK + i * kKvBlockBytes, p.k_cache + off, kKvBlockBytes, bar);
      bulk_g2s(smem + L::kOffV + i * kKvBlockBytes, p.v_cache + off, kKvBlockBytes, bar);
    }
  };
  __syncthreads();
  if (tid == 0) issue_chunk(split);
  if (tr && tid == 0) tr[1] = clock64();

  // ---- queries (and, in fused mode, the new token's k / v) ----
  const int new_pos = len - 1;                      // fused: position of the token being decoded
  const int new_chunk = new_pos / kChunkTokens;
  float knew1 = 0.f, knew2 = 0.f, vnew = 0.f;       // fused owner: thread tid < 64 holds k pair (tid, tid+64); all hold v[tid]
  bool owner = false;
  if constexpr (kFused) {
    const int total = (n_heads + 2 * n_kv) * kHeadDim;
    const float* row = p.qkv + (size_t)t * total;
    auto load = [&](int col) {
      float v = row[col];
      for (int pp = 1; pp < p.n_parts; ++pp) v += row[pp * p.part_stride + col];
      if (p.bias) v += __half2float(p.bias[col]);
      return v;
    };
    const int ps = p.pos[t];
    for (int idx = tid; idx < GQ * 64; idx += kAttnThreads) {
      const int h = idx >> 6, i = idx & 63;
      const int col = (j * GQ + h) * kHeadDim + i;
      const float x1 = load(col), x2 = load(col + 64);
      const float c = p.cos_t[(size_t)ps * 64 + i], sn = p.sin_t[(size_t)ps * 64 + i];
      q_s[h * kHeadDim + i] = __float2half_rn(x1 * c - x2 * sn);
      q_s[h * kHeadDim + i + 64] = __float2half_rn(x2 * c + x1 * sn);
    }
    owner = (new_chunk % S) == split;
    if (owner) {
      const int kcol = (n_heads + j) * kHeadDim, vcol = (n_heads + n_kv + j) * kHeadDim;
      vnew = load(vcol + tid);
      if (tid < 64) {
        const float x1 = load(kcol + tid), x2 = load(kcol + tid + 64);
        const float c = p.cos_t[(size_t)ps * 64 + tid], sn = p.sin_t[(size_t)ps * 64 + tid];
        knew1 = x1 * c - x2 * sn;
        knew2 = x2 * c + x1 * sn;
      }
      const int sl = p.slot[t];
      if (sl >= 0) {
        const size_t off = (((size_t)(sl / kBlockTokens) * n_kv + j) * kBlockTokens + sl % kBlockTokens) * kHeadDim;
        p.v_cache[off + tid] = __float2half_rn(vnew);
        if (tid < 64) {
          p.k_cache[off + tid] = __float2half_rn(knew1);
          p.k_cache[off + tid + 64] = __float2half_rn(knew2);
        }
      }
    }
  } else {
    const uint4* qg = reinterpret_cast<const uint4*>(p.q + ((size_t)t * n_heads + (size_t)j * GQ) * kHeadDim);
    for (int i = tid; i < GQ * 16; i += kAttnThreads) reinterpret_cast<uint4*>(q_s)[i] = qg[i];
  }

  float acc[GQ];
#pragma unroll
  for (int h = 0; h < GQ; ++h) acc[h] = 0.f;
  constexpr int HH = (GQ + 1) / 2;          // heads scored per thread (two thread halves)
  const int tok = tid & 63, hg = tid >> 6;
  const float scale = 0.08838834764831845f; // 1 / sqrt(128)
  const uint4* q_s4 = reinterpret_cast<const uint4*>(q_s);

  int it = 0;
  for (int c = split; c < nchunks; c += S, ++it) {
    if (it > 0) {
      __syncthreads();                      // everyone is done with the previous tile
      if (tid == 0) issue_chunk(c);
    }
    if (tr && tid == 0 && it == 0) tr[2] = clock64();
    mbar_wait(bar, (uint32_t)it & 1u);
    if (tr && tid == 0 && it == 0) tr[3] = clock64();
    const int nvalid = min(kChunkTokens, len - c * kChunkTokens);
    if constexpr (kFused) {
      if (owner && c == new_chunk) {        // the cache row of the new token is not in HBM yet: patch the tile
        const int rrow = new_pos - c * kChunkTokens;
        Vs_h[rrow * kHeadDim + tid] = __float2half_rn(vnew);
        if (tid < 64) {
          Ks_h[rrow * kHeadDim + tid] = __float2half_rn(knew1);
          Ks_h[rrow * kHeadDim + tid + 64] = __float2half_rn(knew2);
        }
      }
    }
    __syncthreads();                        // q_s (first pass) and the patched rows are visible

    // ---- scores: thread = (token, half of the head group) ----
    {
      const uint4* Ks = reinterpret_cast<const uint4*>(Ks_h);
      float sacc[HH];
#pragma unroll
      for (int i = 0; i < HH; ++i) sacc[i] = 0.f;
      if (tok < nvalid) {
#pragma unroll
        for (int s = 0; s < 16; ++s) {
          const int cc = (s + tok) & 15;                       // rotated start: conflict-free 16 B reads
          const uint4 kv = Ks[tok * 16 + cc];
#pragma unroll
          for (int i = 0; i < HH; ++i) {
            const int h = hg * HH + i;
            if (h < GQ) {
              const uint4 qv = q_s4[h * 16 + cc];
              sacc[i] = fhfma2(kv.x, qv.x, sacc[i]);
              sacc[i] = fhfma2(kv.y, qv.y, sacc[i]);
              sacc[i] = fhfma2(kv.z, qv.z, sacc[i]);
              sacc[i] = fhfma2(kv.w, qv.w, sacc[i]);
            }
          }
        }
      }
#pragma unroll
      for (int i = 0; i < HH; ++i) {
        const int h = hg * HH + i;
        if (h < GQ) p_s[h * kChunkTokens + tok] = tok < nvalid ? sacc[i] * scale : -INFINITY;
      }
    }
    __syncthreads();
    if (tr && tid == 0 && it == 0) tr[4] = clock64();

    // ---- online softmax: warp w owns heads w, w+4 ----
    {
      const int warp = tid >> 5, lane = tid & 31;
      for (int h = warp; h < GQ; h += kAttnThreads / 32) {
        const float s0 = p_s[h * kChunkTokens + lane], s1 = p_s[h * kChunkTokens + lane + 32];
        const float m_old = m_s[h];
        const float m_new = fmaxf(m_old, warp_max(fmaxf(s0, s1)));
        const float p0 = expf(s0 - m_new), p1 = expf(s1 - m_new);
        const float sum = warp_sum(p0 + p1);
        p_s[h * kChunkTokens + lane] = p0;
        p_s[h * kChunkTokens + lane + 32] = p1;
        if (lane == 0) {
          const float corr = expf(m_old - m_new);
          c_s[h] = corr;
          l_s[h] = l_s[h] * corr + sum;
          m_s[h] = m_new;
        }
      }
    }
    __syncthreads();
    if (tr && tid == 0 && it == 0) tr[5] = clock64();

    // ---- PV: thread owns output dim d = tid for all GQ heads ----
    {
#pragma unroll
      for (int h = 0; h < GQ; ++h) acc[h] *= c_s[h];
      const int n4 = nvalid & ~3;
      for (int tk = 0; tk < n4; tk += 4) {
        const float v0 = __half2float(Vs_h[(tk + 0) * kHeadDim + tid]);
        const float v1 = __half2float(Vs_h[(tk + 1) * kHeadDim + tid]);
        const float v2 = __half2float(Vs_h[(tk + 2) * kHeadDim + tid]);
        const float v3 = __half2float(Vs_h[(tk + 3) * kHeadDim + tid]);
#pragma unroll
        for (int h = 0; h < GQ; ++h) {
          const float4 pv = *reinterpret_cast<const float4*>(p_s + h * kChunkTokens + tk);
          acc[h] = fmaf(pv.x, v0, acc[h]);
          acc[h] = fmaf(pv.y, v1, acc[h]);
          acc[h] = fmaf(pv.z, v2, acc[h]);
          acc[h] = fmaf(pv.w, v3, acc[h]);
        }
      }
      for (int tk = n4; tk < nvalid; ++tk) {
        const float v0 = __half2float(Vs_h[tk * kHeadDim + tid]);
#pragma unroll
        for (int h = 0; h < GQ; ++h) acc[h] = fmaf(p_s[h * kChunkTokens + tk], v0, acc[h]);
      }
    }
  }

  if (tr && tid == 0) tr[6] = clock64();
  if (n_active == 1) {                      // single split: finish locally
    __syncthreads();
#pragma unroll
    for (int h = 0; h < GQ; ++h) orow[h * kHeadDim + tid] = __float2half_rn(acc[h] / l_s[h]);
    return;
  }

  // ---- publish this split's partial, then let the last split to arrive merge all of them ----
  float* base = p.scratch + ((size_t)(t * n_kv + j) * S) * GQ * kPartFloats;
  float* mine = base + (size_t)split * GQ * kPartFloats;
  __syncthreads();
#pragma unroll
  for (int h = 0; h < GQ; ++h) mine[h * kPartFloats + tid] = acc[h];
  if (tid < GQ) { mine[tid * kPartFloats + kHeadDim] = m_s[tid]; mine[tid * kPartFloats + kHeadDim + 1] = l_s[tid]; }
  __threadfence();
  __syncthreads();
  __shared__ int s_last;
  if (tid == 0) {
    const int old = atomicAdd(&p.counters[t * n_kv + j], 1);
    s_last = (old == n_active - 1);
    if (s_last) p.counters[t * n_kv + j] = 0;       // self-cleaning for the next launch
  }
  __syncthreads();
  if (!s_last) return;
  __threadfence();
#pragma unroll
  for (int h = 0; h < GQ; ++h) {
    float M = -INFINITY;
    for (int s = 0; s < n_active; ++s) M = fmaxf(M, __ldcg(base + ((size_t)s * GQ + h) * kPartFloats + kHeadDim));
    float num = 0.f, den = 0.f;
    for (int s = 0; s < n_active; ++s) {            // fixed split order: deterministic
      const float* ps = base + ((size_t)s * GQ + h) * kPartFloats;
      const float w = expf(__ldcg(ps + kHeadDim) - M);
      num = fmaf(__ldcg(ps + tid), w, num);
      den = fmaf(__ldcg(ps + kHeadDim + 1), w, den);
    }
    orow[h * kHeadDim + tid] = __float2half_rn(num / den);
  }
}

template <int GQ, bool kFused>
static int launch_gq(const AttnParams& p, int T, cudaStream_t s) {
  static bool configured = false;   // per process; the attribute is idempotent
  auto kern = paged_attn_kernel<GQ, kFused>;
  if (!configured) {
    B200_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, AttnSmem<GQ>::kBytes));
    configured = true;
  }
  dim3 grid(p.n_kv, T, p.n_splits);
  B200_CUDA(launch_k(kern, grid, dim3(kAttnThreads), (size_t)AttnSmem<GQ>::kBytes, s, p));
  g_launch_count += 1;
  B200_CUDA(cudaGetLastError());
  return B200_OK;
}

size_t attn_scratch_floats(int T, int n_heads, int n_kv, int n_splits) {
  return (size_t)T * n_kv * n_splits * (n_heads / n_kv) * kPartFloats;
}

static int dispatch(const AttnParams& p, int T, bool fused, cudaStream_t s) {
  B200_REQUIRE(p.n_kv > 0 && p.n_heads % p.n_kv == 0, "paged_attn: n_heads % n_kv");
  B200_REQUIRE(T <= 65535 && p.n_splits >= 1 && p.n_splits <= 64, "paged_attn: T <= 65535 rows, 1 <= n_splits <= 64");
  B200_REQUIRE(p.n_splits == 1 || (p.scratch && p.counters), "paged_attn: n_splits > 1 needs scratch and counters");
  if (T == 0) return B200_OK;
#define CASE(G) case G: return fused ? launch_gq<G, true>(p, T, s) : launch_gq<G, false>(p, T, s)
  switch (p.n_heads / p.n_kv) {
    CASE(1); CASE(2); CASE(3); CASE(4); CASE(5); CASE(6); CASE(7); CASE(8);
    default:
      set_last_error("paged_attn: n_heads / n_kv_heads must be in 1..8");
      return B200_ERR_INVALID;
  }
#undef CASE
}

int launch_paged_attn(const __half* q, const __half* k_cache, const __half* v_cache, const int32_t* block_tables,
                      int bt_stride, const int32_t* bt_row, const int32_t* ctx_len, int T, int n_heads, int n_kv,
                      int n_splits, float* scratch, int32_t* counters, __half* out, cudaStream_t s) {
  AttnParams p{};
  p.trace = g_attn_trace;
  p.q = q; p.k_cache = const_cast<__half*>(k_cache); p.v_cache = const_cast<__half*>(v_cache);
  p.block_tables = block_tables; p.bt_stride = bt_stride; p.bt_row = bt_row; p.ctx_len = ctx_len;
  p.n_heads = n_heads; p.n_kv = n_kv; p.n_splits = n_splits; p.scratch = scratch; p.counters = counters; p.out = out;
  return dispatch(p, T, false, s);
}

int launch_attn_decode_fused(const float* qkv, int n_parts, int64_t part_stride, const __half* bias, const int32_t* pos,
                             const int32_t* slot, const float* cos_t, const float* sin_t, __half* k_cache,
                             __half* v_cache, const int32_t* block_tables, int bt_stride, const int32_t* bt_row,
                             const int32_t* ctx_len, int T, int n_heads, int n_kv, int n_splits, float* scratch,
                             int32_t* counters, __half* out, cudaStream_t s) {
  B200_REQUIRE(n_parts >= 1, "attn_decode_fused: n_parts >= 1");
  AttnParams p{};
  p.trace = g_attn_trace;
  p.qkv = qkv; p.n_parts = n_parts; p.part_stride = part_stride; p.bias = bias; p.pos = pos; p.slot = slot;
  p.cos_t = cos_t; p.sin_t = sin_t; p.k_cache = k_cache; p.v_cache = v_cache;
  p.block_tables = block_tables; p.bt_stride = bt_stride; p.bt_row = bt_row; p.ctx_len = ctx_len;
  p.n_heads = n_heads; p.n_kv = n_kv; p.n_splits = n_splits; p.scratch = scratch; p.counters = counters; p.out = out;
  return dispatch(p, T, true, s);
}

// ---------------------------------------------------------------------------------------------
// Split-KV attention for callers that must not allocate.
//
// b200_op_paged_attn needs the caller to bring a scratch buffer and arrival counters.  Some callers
// cannot: code running inside a latency-critical serving loop where a cudaMalloc would synchronise
// the whole device, a plugin host that hands out no allocator, or a process whose memory pool is
// already carved up and frozen.  For them the scratch lives in the library image itself, as
// file-scope __device__ arrays of a fixed, modest size.  Every GPU that loads the library carries
// these ~4.3 MB until the process exits.
constexpr int kStaticScratchUnits = 8192;                          // rows * n_heads * n_splits
constexpr int kStaticCounters = 2048;                              // rows * n_kv_heads
static __device__ __align__(256) float g_static_attn_scratch[(size_t)kStaticScratchUnits * kPartFloats];
static __device__ __align__(256) int32_t g_static_attn_counters[kStaticCounters];

constexpr int kMaxDevices = 64;
static float* g_static_scratch_ptr[kMaxDevices];                   // device addresses, resolved once per device
static int32_t* g_static_counters_ptr[kMaxDevices];
static std::mutex g_static_mu;

}  // namespace b200

extern "C" int32_t b200_op_paged_attn_noalloc(const void* q, const void* k_cache, const void* v_cache,
                                              const int32_t* block_tables, int32_t bt_stride, const int32_t* bt_row,
                                              const int32_t* ctx_len, int32_t T, int32_t n_heads, int32_t n_kv,
                                              int32_t n_splits, void* out, void* stream) {
  using namespace b200;
  B200_REQUIRE(T >= 0 && n_kv > 0 && n_heads > 0 && n_heads % n_kv == 0 && n_splits >= 1, "paged_attn_noalloc: shape");
  if ((int64_t)T * n_heads * n_splits > kStaticScratchUnits || (int64_t)T * n_kv > kStaticCounters) {
    set_last_error("paged_attn_noalloc: rows * n_heads * n_splits = " + std::to_string((int64_t)T * n_heads * n_splits) +
                   " exceeds the built-in scratch (" + std::to_string(kStaticScratchUnits) + "), or rows * n_kv_heads = " +
                   std::to_string((int64_t)T * n_kv) + " exceeds the built-in counters (" + std::to_string(kStaticCounters) +
                   "); use b200_op_paged_attn with caller-provided scratch");
    return B200_ERR_CAPACITY;
  }
  int dev = 0;
  B200_CUDA(cudaGetDevice(&dev));
  B200_REQUIRE(dev >= 0 && dev < kMaxDevices, "paged_attn_noalloc: device ordinal out of range");
  float* scratch;
  int32_t* counters;
  {
    // In host code the array names are host-side handles, not device addresses: ask the runtime for
    // the address the current device gave them.  Done once per device, on the first call there.
    std::lock_guard<std::mutex> lk(g_static_mu);
    if (!g_static_scratch_ptr[dev]) {
      void *ps = nullptr, *pc = nullptr;
      B200_CUDA(cudaGetSymbolAddress(&ps, g_static_attn_scratch));
      B200_CUDA(cudaGetSymbolAddress(&pc, g_static_attn_counters));
      g_static_counters_ptr[dev] = static_cast<int32_t*>(pc);
      g_static_scratch_ptr[dev] = static_cast<float*>(ps);
    }
    scratch = g_static_scratch_ptr[dev];
    counters = g_static_counters_ptr[dev];
  }
  if (T == 0) return B200_OK;
  // The arrival counters must read zero when the kernel starts.  A finished call leaves them zero, but
  // an aborted one might not, and all calls share the array: clear the part this call uses.
  B200_CUDA(cudaMemsetAsync(counters, 0, (size_t)T * n_kv * sizeof(int32_t), (cudaStream_t)stream));
  return launch_paged_attn((const __half*)q, (const __half*)k_cache, (const __half*)v_cache, block_tables, bt_stride,
                           bt_row, ctx_len, T, n_heads, n_kv, n_splits, scratch, counters, (__half*)out,
                           (cudaStream_t)stream);
}
