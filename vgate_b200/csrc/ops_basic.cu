// Small memory-bound kernels of the decode/prefill path: AWQ repack + dequant, embedding gather,
// fused residual-sum + RMSNorm, RoPE + paged KV write, SwiGLU.
#include "common.cuh"
#include "norm.cuh"
#include "internal.h"

namespace b200 {

B200_DEFINE_TL_SETTER(tl_set_basic)

__constant__ int kAwqOrder[8] = {0, 4, 1, 5, 2, 6, 3, 7};

// ---------------------------------------------------------------------------------------------
// AWQ checkpoint layout -> kernel layout (one-time, at load).
//   in : qweight[k][n/8] nibble kAwqOrder[n%8]              (awq_triton.py:53-70)
//   out: wq[n][k/8]      nibble kAwqOrder[k%8]
__global__ void awq_repack_weight_kernel(const uint32_t* __restrict__ qweight, int K, int N,
                                         uint32_t* __restrict__ wq, int row_offset) {
  const int n = blockIdx.x * blockDim.x + threadIdx.x;   // n fastest: coalesced reads of qweight rows
  const int j = blockIdx.y;                              // output word along K
  if (n >= N) return;
  const int in_shift = 4 * kAwqOrder[n & 7];
  uint32_t out = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const uint32_t word = qweight[(size_t)(8 * j + i) * (N / 8) + (n >> 3)];
    out |= ((word >> in_shift) & 0xFu) << (4 * kAwqOrder[i]);
  }
  wq[(size_t)(n + row_offset) * (K / 8) + j] = out;
}

__global__ void awq_repack_scales_kernel(const uint32_t* __restrict__ qzeros, const __half* __restrict__ scales,
                                         int NG, int N, __half* __restrict__ sc, __half* __restrict__ zp,
                                         int row_offset) {
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  const int g = blockIdx.y;
  if (n >= N) return;
  const uint32_t zw = qzeros[(size_t)g * (N / 8) + (n >> 3)];
  const int z = (zw >> (4 * kAwqOrder[n & 7])) & 0xF;
  sc[(size_t)(n + row_offset) * NG + g] = scales[(size_t)g * N + n];
  zp[(size_t)(n + row_offset) * NG + g] = __int2half_rn(1024 + z);
}

int launch_awq_repack(const int32_t* qweight, const int32_t* qzeros, const __half* scales, int K, int N, int G,
                      uint32_t* wq, __half* sc, __half* zp, int row_offset, cudaStream_t s) {
  B200_REQUIRE(K % G == 0 && G % 32 == 0 && N % 8 == 0 && K % 32 == 0, "awq_repack: K % G, G % 32, N % 8");
  dim3 block(128);
  dim3 grid_w((N + 127) / 128, K / 8);
  awq_repack_weight_kernel<<<grid_w, block, 0, s>>>(reinterpret_cast<const uint32_t*>(qweight), K, N, wq, row_offset);
  dim3 grid_s((N + 127) / 128, K / G);
  awq_repack_scales_kernel<<<grid_s, block, 0, s>>>(reinterpret_cast<const uint32_t*>(qzeros), scales, K / G, N, sc,
                                                    zp, row_offset);
  g_launch_count += 2;
  B200_CUDA(cudaGetLastError());
  return B200_OK;
}

// kernel layout -> dense fp16 [N, K]; uses the same dequant8 as the GEMV/GEMM kernels.
__global__ void awq_dequant_kernel(const uint32_t* __restrict__ wq, const __half* __restrict__ sc,
                                   const __half* __restrict__ zp, int K, int N, int G, __half* __restrict__ w) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;  // word along K
  const int n = blockIdx.y;
  if (j >= K / 8) return;
  const int g = (8 * j) / G;
  const __half2 s2 = __half2half2(sc[(size_t)n * (K / G) + g]);
  const __half2 z2 = __half2half2(zp[(size_t)n * (K / G) + g]);
  uint32_t o[4];
  dequant8(wq[(size_t)n * (K / 8) + j], *reinterpret_cast<const uint32_t*>(&z2),
           *reinterpret_cast<const uint32_t*>(&s2), o);
  *reinterpret_cast<uint4*>(w + (size_t)n * K + 8 * j) = make_uint4(o[0], o[1], o[2], o[3]);
}

int launch_awq_dequant(const uint32_t* wq, const __half* sc, const __half* zp, int K, int N, int G, __half* w,
                       cudaStream_t s) {
  B200_REQUIRE(K % G == 0 && K % 8 == 0, "awq_dequant: K % G");
  dim3 grid((K / 8 + 127) / 128, N);
  awq_dequant_kernel<<<grid, 128, 0, s>>>(wq, sc, zp, K, N, G, w);
  g_launch_count += 1;
  B200_CUDA(cudaGetLastError());
  return B200_OK;
}

// ---------------------------------------------------------------------------------------------
// Embedding gather into the fp32 residual stream.
__global__ void embed_kernel(const __half* __restrict__ table, const int32_t* __restrict__ ids, int H,
                             float* __restrict__ h, int bf) {
  const unsigned tl = kernel_prologue(kKidEmbed);
  const int t = blockIdx.x;
  const __half* row = table + (size_t)ids[t] * H;
  for (int i = threadIdx.x * 8; i < H; i += blockDim.x * 8) {
    const uint4 v = ldg_cached(row + i);
    float4 a, b;
    float2 f;
    f = unpack16x2(v.x, bf); a.x = f.x; a.y = f.y;
    f = unpack16x2(v.y, bf); a.z = f.x; a.w = f.y;
    f = unpack16x2(v.z, bf); b.x = f.x; b.y = f.y;
    f = unpack16x2(v.w, bf); b.z = f.x; b.w = f.y;
    *reinterpret_cast<float4*>(h + (size_t)t * H + i) = a;
    *reinterpret_cast<float4*>(h + (size_t)t * H + i + 4) = b;
  }
  kernel_epilogue(tl);
}

int launch_embed(const __half* table, const int32_t* ids, int T, int H, float* h, cudaStream_t s, bool bf16) {
  B200_REQUIRE(H % 8 == 0, "embed: H % 8");
  if (T == 0) return B200_OK;
  launch_k(embed_kernel, dim3(T), dim3(192), 0, s, table, ids, H, h, bf16 ? 1 : 0);
  g_launch_count += 1;
  B200_CUDA(cudaGetLastError());
  return B200_OK;
}

// ---------------------------------------------------------------------------------------------
// Fused residual-sum + RMSNorm.  One cluster of 4 CTAs per output row, a quarter row per CTA: 4 x n_rows
// CTAs cover the machine at batch 64 (one CTA per row left 84 SMs idle and took 7 us with 12 partials).
//   v = h[row] + sum_p parts[p][row]     (fixed p order: deterministic split-K reduction)
//   h[row] = v (when n_parts > 0);  xn[r] = fp16(v * rsqrt(mean(v^2) + eps) * gamma)
// The sum of squares is exchanged through distributed shared memory: every warp pushes its partial
// into all four CTAs, one cluster barrier, then every CTA adds the 16 values in the same order.
// Follows vLLM qwen2.py:287-312 (RMSNorm with fused residual add), arithmetic in fp32.
template <bool kOneRow>       // kOneRow: exactly one row per cluster (decode, short prefill): no loop carried state
__global__ void __cluster_dims__(kNormCluster, 1, 1) __launch_bounds__(kNormThreads)
rmsnorm_kernel(float* __restrict__ h, const float* __restrict__ parts, int n_parts, int64_t part_stride,
               const __half* __restrict__ gamma, float eps, const int32_t* __restrict__ rows, int n_rows, int H,
               __half* __restrict__ xn, int bf) {
  // red is double-buffered by row parity: a CTA may push the partials of its next row while a slower
  // CTA of the cluster still reads the previous row's (one cluster barrier per row is then enough:
  // nobody can be two rows ahead of a CTA that has not passed the barrier in between)
  __shared__ float red[2][kNormCluster][kNormWarps];
  const unsigned tl = kernel_prologue(kKidRmsnorm);
  const int q = blockIdx.x % kNormCluster;            // == %cluster_ctarank for a 1-D cluster
  const int n_clusters = gridDim.x / kNormCluster;
  int it = 0;
  for (int r = blockIdx.x / kNormCluster; r < n_rows; r += kOneRow ? n_rows : n_clusters, ++it) {   // prefill: many rows per cluster
    const int row = rows ? rows[r] : r;
    float4 v[kNormMaxPerThread];
    const float ss = rms_quarter_load(h, parts, n_parts, part_stride, row, q, H, threadIdx.x, v);
    if ((threadIdx.x & 31) < kNormCluster) {          // lane c pushes this warp's partial into CTA c
      const uint32_t local = smem_u32(&red[it & 1][q][threadIdx.x >> 5]);
      uint32_t remote;
      asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(local), "r"(threadIdx.x & 31));
      asm volatile("st.shared::cluster.f32 [%0], %1;" ::"r"(remote), "f"(ss) : "memory");
    }
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
    const float rs = rms_scale(red[it & 1], H, eps);
    rms_quarter_store(v, rs, gamma, xn, r, q, H, threadIdx.x, bf != 0);
  }
  kernel_epilogue(tl);
}

// Tensor-parallel decode passes: residual-sum + RMSNorm with the all-reduce of the row-parallel projection's
// partials folded in (norm.cuh: tp_own_sum_push, tp_collect).  One row per cluster.
__global__ void __cluster_dims__(kNormCluster, 1, 1) __launch_bounds__(kNormThreads)
rmsnorm_tp_kernel(float* __restrict__ h, const float* __restrict__ parts, int n_parts, int64_t part_stride, const TpPeers tp,
                  const uint32_t* __restrict__ epoch, const __half* __restrict__ gamma, float eps,
                  const int32_t* __restrict__ rows, int n_rows, int H, __half* __restrict__ xn, int bf) {
  __shared__ float red[kNormCluster][kNormWarps];
  const unsigned tl = kernel_prologue(kKidRmsnorm);
  const uint32_t seq = *epoch * tp.n_sites + tp.site + 1u;
  const int q = blockIdx.x % kNormCluster;
  const int r = blockIdx.x / kNormCluster;
  const int row = rows ? rows[r] : r;
  float4 own[kNormMaxPerThread], hv[kNormMaxPerThread], v[kNormMaxPerThread];
  tp_own_sum_push(h, parts, n_parts, part_stride, tp, seq, row, q, H, threadIdx.x, own, hv);
  const unsigned long long t_own = tl != ~0u ? gtime() : 0;
  const float ss = tp_collect(h, tp, seq, row, q, H, threadIdx.x, own, hv, v);
  if (tl != ~0u) {                     // timeline: a second record {13, own partials summed and pushed, -, peer sums added}
    const unsigned slot = atomicAdd(d_tl.ctr, 1u);
    if (slot < d_tl.cap) { d_tl.buf[4 * slot] = 13; d_tl.buf[4 * slot + 1] = t_own; d_tl.buf[4 * slot + 2] = t_own; d_tl.buf[4 * slot + 3] = gtime(); }
  }
  if ((threadIdx.x & 31) < kNormCluster) {
    const uint32_t local = smem_u32(&red[q][threadIdx.x >> 5]);
    uint32_t remote;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(local), "r"(threadIdx.x & 31));
    asm volatile("st.shared::cluster.f32 [%0], %1;" ::"r"(remote), "f"(ss) : "memory");
  }
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
  const float rs = rms_scale(red, H, eps);
  rms_quarter_store(v, rs, gamma, xn, r, q, H, threadIdx.x, bf != 0);
  kernel_epilogue(tl);
}

// The step counter the rendezvous words are compared with: one tick at the start of every pass that uses them.
__global__ void tp_tick_kernel(uint32_t* epoch) {
  const unsigned tl = kernel_prologue(kKidTpTick);
  if (threadIdx.x == 0) *epoch += 1;
  kernel_epilogue(tl);
}

int launch_tp_tick(uint32_t* epoch, cudaStream_t s) {
  launch_k(tp_tick_kernel, dim3(1), dim3(32), 0, s, epoch);
  g_launch_count += 1;
  B200_CUDA(cudaGetLastError());
  return B200_OK;
}

int launch_rmsnorm_tp(float* h, const float* parts, int n_parts, int64_t part_stride, const TpPeers& tp, const uint32_t* epoch,
                      const __half* gamma, float eps, const int32_t* rows, int n_rows, int H, __half* xn, cudaStream_t s, bool bf16) {
  const int bf = bf16 ? 1 : 0;
  B200_REQUIRE(H % (4 * kNormCluster) == 0 && H <= kNormCluster * kNormThreads * 4 * kNormMaxPerThread,
               "rmsnorm_tp: H % 16 == 0 && H <= 8192");
  // every cluster of the grid spins until its peers' words arrive: the whole grid has to be resident at once
  B200_REQUIRE(n_rows >= 1 && n_rows <= kTpNormMaxRows, "rmsnorm_tp: 1 <= rows <= 256");
  B200_REQUIRE(n_parts >= 1 && tp.n_ranks >= 2 && tp.n_ranks <= kMaxTp && n_rows <= tp.rows_cap, "rmsnorm_tp: 2..8 ranks, >= 1 partial, rows within the receive buffers");
  launch_k(rmsnorm_tp_kernel, dim3(n_rows * kNormCluster), dim3(kNormThreads), 0, s, h, parts, n_parts, part_stride, tp, epoch, gamma, eps,
           rows, n_rows, H, xn, bf);
  g_launch_count += 1;
  B200_CUDA(cudaGetLastError());
  return B200_OK;
}

// Prefill variant (no partials, thousands of rows): one WARP per row, every load of the row in flight at
// once (the cluster kernel keeps only ~185 rows in flight on the whole GPU and is latency-bound there:
// 88 us for 8192 rows).  Bit-identical to the cluster kernel: lane l plays thread (w * 32 + l) of quarter q
// for every (q, w), so each of the 16 partial sums goes through the same FMA chain and the same shuffle
// tree, and they are added in the same order.  H <= 2048 (16 float4 per lane kept for the store).
constexpr int kNormWarpRowMaxH = 2048;
__global__ void __launch_bounds__(256)
rmsnorm_rows_kernel(const float* __restrict__ h, const __half* __restrict__ gamma, float eps, const int32_t* __restrict__ rows,
                    int n_rows, int H, __half* __restrict__ xn, int bf) {
  const unsigned tl = kernel_prologue(kKidRmsnorm);
  const int lane = threadIdx.x & 31;
  const int r = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (r < n_rows) {
    const int row = rows ? rows[r] : r;
    const int HQ = H / kNormCluster, nvec = HQ / 4;             // float4 per quarter
    const int wpq = (nvec + 31) / 32;                           // warps the cluster kernel uses per quarter (<= 4)
    constexpr int kMaxVec = kNormWarpRowMaxH / 4 / 32;          // 16
    float4 v[kMaxVec];
    const float4* src = reinterpret_cast<const float4*>(h + (size_t)row * H);
#pragma unroll
    for (int k = 0; k < kMaxVec; ++k) {
      const int q = k / 4, w = k % 4, i = w * 32 + lane;        // (quarter, warp) slot k; i = thread index in the quarter
      v[k] = (q < kNormCluster && w < wpq && i < nvec) ? src[q * nvec + i] : make_float4(0.f, 0.f, 0.f, 0.f);
    }
    float red[kNormCluster][kNormWarps];
#pragma unroll
    for (int k = 0; k < kMaxVec; ++k) {
      float ss = 0.f;
      ss = __fmaf_rn(v[k].x, v[k].x, ss); ss = __fmaf_rn(v[k].y, v[k].y, ss);
      ss = __fmaf_rn(v[k].z, v[k].z, ss); ss = __fmaf_rn(v[k].w, v[k].w, ss);
      red[k / 4][k % 4] = warp_sum(ss);
    }
    const float rs = rms_scale(red, H, eps);
    __half* dst = xn + (size_t)r * H;
#pragma unroll
    for (int k = 0; k < kMaxVec; ++k) {
      const int q = k / 4, w = k % 4, i = w * 32 + lane;
      if (w < wpq && i < nvec) {
        const int e = (q * nvec + i) * 4;
        const uint2 gw = *reinterpret_cast<const uint2*>(gamma + e);
        const float2 g0 = unpack16x2(gw.x, bf);
        const float2 g1 = unpack16x2(gw.y, bf);
        uint2 o;
        o.x = pack16x2(__fmul_rn(__fmul_rn(v[k].x, rs), g0.x), __fmul_rn(__fmul_rn(v[k].y, rs), g0.y), bf);
        o.y = pack16x2(__fmul_rn(__fmul_rn(v[k].z, rs), g1.x), __fmul_rn(__fmul_rn(v[k].w, rs), g1.y), bf);
        *reinterpret_cast<uint2*>(dst + e) = o;
      }
    }
  }
  kernel_epilogue(tl);
}

int launch_rmsnorm(float* h, const float* parts, int n_parts, int64_t part_stride, const __half* gamma, float eps,
                   const int32_t* rows, int n_rows, int H, __half* xn, cudaStream_t s, bool bf16) {
  const int bf = bf16 ? 1 : 0;
  B200_REQUIRE(H % (4 * kNormCluster) == 0 && H <= kNormCluster * kNormThreads * 4 * kNormMaxPerThread,
               "rmsnorm: H % 16 == 0 && H <= 8192");
  if (n_rows == 0) return B200_OK;
  if (n_parts == 0 && n_rows >= 1024 && H <= kNormWarpRowMaxH) {     // large prefill pass: warp per row, same bits
    launch_k(rmsnorm_rows_kernel, dim3((n_rows + 7) / 8), dim3(256), 0, s, (const float*)h, gamma, eps, rows, n_rows, H, xn, bf);
    g_launch_count += 1;
    B200_CUDA(cudaGetLastError());
    return B200_OK;
  }
  const int n_clusters = n_rows < 2048 ? n_rows : 2048;      // beyond that a cluster walks several rows
  if (n_clusters == n_rows)
    launch_k(rmsnorm_kernel<true>, dim3(n_clusters * kNormCluster), dim3(kNormThreads), 0, s, h, parts, n_parts, part_stride, gamma,
             eps, rows, n_rows, H, xn, bf);
  else
    launch_k(rmsnorm_kernel<false>, dim3(n_clusters * kNormCluster), dim3(kNormThreads), 0, s, h, parts, n_parts, part_stride, gamma,
             eps, rows, n_rows, H, xn, bf);
  g_launch_count += 1;
  B200_CUDA(cudaGetLastError());
  return B200_OK;
}

// ---------------------------------------------------------------------------------------------
// RoPE (neox: rotate halves, vLLM qwen2.py:181 get_rope) fused with the paged KV-cache write.
// One CTA per token; one thread per (head, i < 64) pair.  fp32 in, one rounding to fp16 out.
// The qkv row may arrive as n_parts split-K partial tensors (summed here in fixed order, plus bias).
__global__ void rope_kv_kernel(const float* __restrict__ qkv, int n_parts, int64_t part_stride,
                               const __half* __restrict__ bias, const int32_t* __restrict__ pos,
                               const int32_t* __restrict__ slot, const float* __restrict__ cos_t,
                               const float* __restrict__ sin_t, int n_heads, int n_kv, __half* __restrict__ q_out,
                               __half* __restrict__ k_cache, __half* __restrict__ v_cache, int bf) {
  const unsigned tl = kernel_prologue(kKidRope);
  const int t = blockIdx.x;
  const int p = pos[t];
  const int sl = slot[t];
  const int total_heads = n_heads + 2 * n_kv;
  const float* row = qkv + (size_t)t * total_heads * kHeadDim;
  for (int idx = threadIdx.x; idx < total_heads * 64; idx += blockDim.x) {
    const int head = idx >> 6, i = idx & 63;
    float x1 = 0.f, x2 = 0.f;
    for (int p0 = 0; p0 < n_parts; p0 += 4) {          // loads first, then adds in fixed order
      float a[4], b[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const bool ok = p0 + k < n_parts;
        a[k] = ok ? __ldcg(row + (size_t)(p0 + k) * part_stride + head * kHeadDim + i) : 0.f;
        b[k] = ok ? __ldcg(row + (size_t)(p0 + k) * part_stride + head * kHeadDim + i + 64) : 0.f;
      }
#pragma unroll
      for (int k = 0; k < 4; ++k) { x1 += a[k]; x2 += b[k]; }
    }
    if (bias) {
      const uint16_t* b16 = reinterpret_cast<const uint16_t*>(bias);
      x1 += from16(b16[head * kHeadDim + i], bf);
      x2 += from16(b16[head * kHeadDim + i + 64], bf);
    }
    if (head < n_heads) {
      const float c = cos_t[(size_t)p * 64 + i], sn = sin_t[(size_t)p * 64 + i];
      uint16_t* o = reinterpret_cast<uint16_t*>(q_out) + ((size_t)t * n_heads + head) * kHeadDim;
      float r1, r2;
      rope_rotate(x1, x2, c, sn, r1, r2);
      o[i] = to16(r1, bf);
      o[i + 64] = to16(r2, bf);
    } else if (sl >= 0) {
      const int blk = sl / kBlockTokens, off = sl % kBlockTokens;
      if (head < n_heads + n_kv) {
        const int j = head - n_heads;
        const float c = cos_t[(size_t)p * 64 + i], sn = sin_t[(size_t)p * 64 + i];
        uint16_t* o = reinterpret_cast<uint16_t*>(k_cache) + (((size_t)blk * n_kv + j) * kBlockTokens + off) * kHeadDim;
        float r1, r2;
        rope_rotate(x1, x2, c, sn, r1, r2);
        o[kv_swz(off, i)] = to16(r1, bf);                            // swizzled row layout, see common.cuh
        o[kv_swz(off, i + 64)] = to16(r2, bf);
      } else {
        const int j = head - n_heads - n_kv;
        uint16_t* o = reinterpret_cast<uint16_t*>(v_cache) + (((size_t)blk * n_kv + j) * kBlockTokens + off) * kHeadDim;
        o[kv_swz(off, i)] = to16(x1, bf);
        o[kv_swz(off, i + 64)] = to16(x2, bf);
      }
    }
  }
  kernel_epilogue(tl);
}

int launch_rope_kv(const float* qkv, int n_parts, int64_t part_stride, const __half* bias, const int32_t* pos,
                   const int32_t* slot, const float* cos_t, const float* sin_t, int T, int n_heads, int n_kv,
                   __half* q_out, __half* k_cache, __half* v_cache, cudaStream_t s, bool bf16) {
  B200_REQUIRE(n_parts >= 1, "rope_kv: n_parts >= 1");
  if (T == 0) return B200_OK;
  launch_k(rope_kv_kernel, dim3(T), dim3(256), 0, s, qkv, n_parts, part_stride, bias, pos, slot, cos_t, sin_t, n_heads,
           n_kv, q_out, k_cache, v_cache, bf16 ? 1 : 0);
  g_launch_count += 1;
  B200_CUDA(cudaGetLastError());
  return B200_OK;
}

// ---------------------------------------------------------------------------------------------
// SwiGLU: act = fp16(silu(gate) * up), gate|up fp32 (vLLM qwen2.py:106-111 SiluAndMul).
__global__ void swiglu_kernel(const float* __restrict__ gu, int I, __half* __restrict__ act) {
  const unsigned tl = kernel_prologue(kKidSwiglu);
  const int t = blockIdx.y;
  const int i = (blockIdx.x * blockDim.x + threadIdx.x) * 4;
  if (i >= I) return;
  const float4 g = *reinterpret_cast<const float4*>(gu + (size_t)t * 2 * I + i);
  const float4 u = *reinterpret_cast<const float4*>(gu + (size_t)t * 2 * I + I + i);
  auto f = [](float a, float b) { return a / (1.f + expf(-a)) * b; };
  uint2 o;
  o.x = pack_half2(f(g.x, u.x), f(g.y, u.y));
  o.y = pack_half2(f(g.z, u.z), f(g.w, u.w));
  *reinterpret_cast<uint2*>(act + (size_t)t * I + i) = o;
  kernel_epilogue(tl);
}

int launch_swiglu(const float* gu, int T, int I, __half* act, cudaStream_t s) {
  B200_REQUIRE(I % 4 == 0, "swiglu: I % 4");
  if (T == 0) return B200_OK;
  dim3 grid((I / 4 + 255) / 256, T);
  launch_k(swiglu_kernel, grid, dim3(256), 0, s, gu, I, act);
  g_launch_count += 1;
  B200_CUDA(cudaGetLastError());
  return B200_OK;
}

}  // namespace b200
