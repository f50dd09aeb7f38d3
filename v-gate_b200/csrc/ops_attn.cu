// Paged-KV GQA attention, one query row per CTA-row (decode; prefill rows are issued as rows with
// ctx_len = position + 1).  Replaces vLLM's attention backend for the decode path
// (vllm/model_executor/models/qwen2.py:120-236 -> Attention -> FlashInfer/FlashAttention).
//
// grid (n_kv_heads, T), 128 threads.  KV blocks (16 tokens x 128 x fp16 = 4 KB, contiguous per
// (block, kv head)) are staged into shared memory by 1-D bulk TMA copies (cp.async.bulk, SASS
// UBLKCP) completing on an mbarrier, double-buffered in chunks of 64 tokens, so the K/V bytes of
// chunk c+1 stream from HBM while chunk c is scored.  All GQ = n_heads / n_kv_heads query heads
// of the group share each staged K/V chunk.  Scores / softmax / PV in fp32; one rounding to fp16.
#include "common.cuh"
#include "internal.h"

namespace b200 {

constexpr int kAttnThreads = 128;
constexpr int kChunkTokens = 64;
constexpr int kChunkBlocks = kChunkTokens / kBlockTokens;            // 4
constexpr int kKvBlockBytes = kBlockTokens * kHeadDim * 2;           // 4096
constexpr int kStageBytes = kChunkTokens * kHeadDim * 2;             // 16384 per K or V stage

template <int GQ>
struct AttnSmem {
  static constexpr int kOffK = 0;
  static constexpr int kOffV = 2 * kStageBytes;
  static constexpr int kOffQ = 4 * kStageBytes;                      // [GQ][128] half
  static constexpr int kOffP = kOffQ + GQ * kHeadDim * 2;            // [GQ][64] float
  static constexpr int kOffM = kOffP + GQ * kChunkTokens * 4;        // m[GQ], l[GQ], corr[GQ]
  static constexpr int kOffBar = (kOffM + 3 * GQ * 4 + 15) / 16 * 16;
  static constexpr int kBytes = kOffBar + 2 * 8;
};

template <int GQ>
__global__ void __launch_bounds__(kAttnThreads)
paged_attn_kernel(const __half* __restrict__ q, const __half* __restrict__ k_cache, const __half* __restrict__ v_cache,
                  const int32_t* __restrict__ block_tables, int bt_stride, const int32_t* __restrict__ bt_row,
                  const int32_t* __restrict__ ctx_len, int n_heads, int n_kv, __half* __restrict__ out) {
  extern __shared__ __align__(128) unsigned char smem[];
  using L = AttnSmem<GQ>;
  pdl_prologue();
  const int j = blockIdx.x, t = blockIdx.y, tid = threadIdx.x;
  const int len = ctx_len[t];
  __half* orow = out + ((size_t)t * n_heads + (size_t)j * GQ) * kHeadDim;
  if (len <= 0) {
#pragma unroll
    for (int h = 0; h < GQ; ++h) orow[h * kHeadDim + tid] = __float2half_rn(0.f);
    return;
  }
  const int32_t* bt = block_tables + (size_t)bt_row[t] * bt_stride;
  const int nblocks = (len + kBlockTokens - 1) / kBlockTokens;
  const int nchunks = (len + kChunkTokens - 1) / kChunkTokens;

  uint4* q_s = reinterpret_cast<uint4*>(smem + L::kOffQ);
  float* p_s = reinterpret_cast<float*>(smem + L::kOffP);
  float* m_s = reinterpret_cast<float*>(smem + L::kOffM);
  float* l_s = m_s + GQ;
  float* c_s = l_s + GQ;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L::kOffBar);

  // q for this group: GQ * 128 halves = GQ * 16 uint4
  const uint4* qg = reinterpret_cast<const uint4*>(q + ((size_t)t * n_heads + (size_t)j * GQ) * kHeadDim);
  for (int i = tid; i < GQ * 16; i += kAttnThreads) q_s[i] = qg[i];
  if (tid < GQ) { m_s[tid] = -INFINITY; l_s[tid] = 0.f; }
  if (tid == 0) {
    mbar_init(&bars[0], 1);
    mbar_init(&bars[1], 1);
    fence_mbar_init();
  }
  __syncthreads();

  auto issue_chunk = [&](int c) {   // one thread
    const int stage = c & 1;
    const int b0 = c * kChunkBlocks;
    const int nb = min(kChunkBlocks, nblocks - b0);
    mbar_arrive_expect_tx(&bars[stage], (uint32_t)nb * 2u * kKvBlockBytes);
    for (int i = 0; i < nb; ++i) {
      const size_t off = ((size_t)bt[b0 + i] * n_kv + j) * (size_t)(kBlockTokens * kHeadDim);
      bulk_g2s(smem + L::kOffK + stage * kStageBytes + i * kKvBlockBytes, k_cache + off, kKvBlockBytes, &bars[stage]);
      bulk_g2s(smem + L::kOffV + stage * kStageBytes + i * kKvBlockBytes, v_cache + off, kKvBlockBytes, &bars[stage]);
    }
  };
  if (tid == 0) issue_chunk(0);

  float acc[GQ];
#pragma unroll
  for (int h = 0; h < GQ; ++h) acc[h] = 0.f;

  constexpr int HH = (GQ + 1) / 2;          // heads scored per thread (two thread halves)
  const int tok = tid & 63, hg = tid >> 6;
  const float scale = 0.08838834764831845f; // 1 / sqrt(128)

  for (int c = 0; c < nchunks; ++c) {
    const int stage = c & 1;
    if (tid == 0 && c + 1 < nchunks) issue_chunk(c + 1);   // stage^1 was released by the last barrier below
    mbar_wait(&bars[stage], (uint32_t)(c >> 1) & 1u);
    const int nvalid = min(kChunkTokens, len - c * kChunkTokens);

    // ---- scores: thread = (token, half of the head group) ----
    {
      const uint4* Ks = reinterpret_cast<const uint4*>(smem + L::kOffK + stage * kStageBytes);
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
              const uint4 qv = q_s[h * 16 + cc];
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

    // ---- PV: thread owns output dim d = tid for all GQ heads ----
    {
      const __half* Vs = reinterpret_cast<const __half*>(smem + L::kOffV + stage * kStageBytes);
#pragma unroll
      for (int h = 0; h < GQ; ++h) acc[h] *= c_s[h];
      const int n4 = nvalid & ~3;
      for (int tk = 0; tk < n4; tk += 4) {
        const float v0 = __half2float(Vs[(tk + 0) * kHeadDim + tid]);
        const float v1 = __half2float(Vs[(tk + 1) * kHeadDim + tid]);
        const float v2 = __half2float(Vs[(tk + 2) * kHeadDim + tid]);
        const float v3 = __half2float(Vs[(tk + 3) * kHeadDim + tid]);
#pragma unroll
        for (int h = 0; h < GQ; ++h) {
          const float4 p = *reinterpret_cast<const float4*>(p_s + h * kChunkTokens + tk);
          acc[h] = fmaf(p.x, v0, acc[h]);
          acc[h] = fmaf(p.y, v1, acc[h]);
          acc[h] = fmaf(p.z, v2, acc[h]);
          acc[h] = fmaf(p.w, v3, acc[h]);
        }
      }
      for (int tk = n4; tk < nvalid; ++tk) {
        const float v0 = __half2float(Vs[tk * kHeadDim + tid]);
#pragma unroll
        for (int h = 0; h < GQ; ++h) acc[h] = fmaf(p_s[h * kChunkTokens + tk], v0, acc[h]);
      }
    }
    __syncthreads();   // p_s and this stage's K/V are free again
  }

#pragma unroll
  for (int h = 0; h < GQ; ++h) orow[h * kHeadDim + tid] = __float2half_rn(acc[h] / l_s[h]);
}

template <int GQ>
static int launch_gq(const __half* q, const __half* k_cache, const __half* v_cache, const int32_t* block_tables,
                     int bt_stride, const int32_t* bt_row, const int32_t* ctx_len, int T, int n_heads, int n_kv,
                     __half* out, cudaStream_t s) {
  static bool configured = false;   // per process; attribute is per device but idempotent
  auto kern = paged_attn_kernel<GQ>;
  if (!configured) {
    B200_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, AttnSmem<GQ>::kBytes));
    configured = true;
  }
  dim3 grid(n_kv, T);
  launch_k(kern, grid, dim3(kAttnThreads), AttnSmem<GQ>::kBytes, s, q, k_cache, v_cache, block_tables, bt_stride, bt_row,
           ctx_len, n_heads, n_kv, out);
  g_launch_count += 1;
  B200_CUDA(cudaGetLastError());
  return B200_OK;
}

int launch_paged_attn(const __half* q, const __half* k_cache, const __half* v_cache, const int32_t* block_tables,
                      int bt_stride, const int32_t* bt_row, const int32_t* ctx_len, int T, int n_heads, int n_kv,
                      __half* out, cudaStream_t s) {
  B200_REQUIRE(n_kv > 0 && n_heads % n_kv == 0, "paged_attn: n_heads % n_kv");
  B200_REQUIRE(T <= 65535, "paged_attn: T <= 65535 rows per launch");
  if (T == 0) return B200_OK;
  const int gq = n_heads / n_kv;
#define CASE(G) case G: return launch_gq<G>(q, k_cache, v_cache, block_tables, bt_stride, bt_row, ctx_len, T, n_heads, n_kv, out, s)
  switch (gq) {
    CASE(1); CASE(2); CASE(3); CASE(4); CASE(5); CASE(6); CASE(7); CASE(8);
    default:
      set_last_error("paged_attn: n_heads / n_kv_heads must be in 1..8");
      return B200_ERR_INVALID;
  }
#undef CASE
}

}  // namespace b200
