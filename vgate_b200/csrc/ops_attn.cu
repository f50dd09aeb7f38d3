// Decode / prefill attention kernels and launchers; the decode unit itself lives in attn.cuh (shared
// with the persistent decode kernel).
#include <mutex>

#include "attn.cuh"

namespace b200 {

B200_DEFINE_TL_SETTER(tl_set_attn)

static unsigned long long* g_attn_trace = nullptr;     // profiling aid: 8 x uint64 clock stamps per CTA
void attn_set_trace(unsigned long long* dev_buf) { g_attn_trace = dev_buf; }

template <int GQ, bool kFused, int kPhys>
__global__ void __launch_bounds__(kAttnThreads)
paged_attn_kernel(const AttnParams p) {
  extern __shared__ __align__(128) unsigned char smem[];
  // fused decode mode: the unit prefetches K/V before it waits for the upstream qkv projection
  const unsigned tl = kernel_prologue<!kFused>(kKidAttn);
  unsigned long long* tr = p.trace ? p.trace + 8 * (size_t)(blockIdx.x + gridDim.x * (blockIdx.y + gridDim.y * blockIdx.z)) : nullptr;
  uint32_t bar_uses = 0;
  paged_attn_unit<GQ, kFused, 0, true, kFused, kPhys>(p, smem, blockIdx.x, blockIdx.y, blockIdx.z, tr, bar_uses);
  if (tr && threadIdx.x == 0) tr[7] = clock64();
  kernel_epilogue(tl);
}

// ---------------------------------------------------------------------------------------------
// Prefill attention: one CTA per (kv head, tile of <= 16 consecutive query tokens of one sequence),
// one warp per query head of the group.  The K/V chunks of the sequence are staged once per CTA and
// shared by all heads and all 16 query rows (the per-row kernel above would stage them once per row).
// Same tensor-core scheme: S = Q K^T with the 16 query tokens as the MMA M rows, causal mask from the
// row's own context length, online softmax for two rows per thread, O += P V.
struct QTile { int32_t row0, nrows, bt_row, ctx0; };   // rows [row0, row0 + nrows); row i sees ctx0 + i tokens

struct PrefillAttnParams {
  const __half* q;               // [T, nh * 128] rotated queries
  const __half* k_cache;
  const __half* v_cache;
  const int32_t* block_tables;
  int bt_stride;
  const QTile* tiles;
  int n_heads, n_kv;
  __half* out;                   // [T, nh * 128]
  int bf16;
};

__device__ __forceinline__ float ex2_approx(float x) {       // 2^x, -inf -> 0
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

template <int GQ>
struct PrefillSmem {
  static constexpr int kOffKV = 0;                                   // 2 buffers x (K 16 KB + V 16 KB)
  static constexpr int kOffQ = 4 * kTileBytes;                       // GQ x [16][128] fp16
  static constexpr int kOffBar = kOffQ + GQ * 16 * kHeadDim * 2;
  static constexpr int kBytes = kOffBar + 16;
};

template <int GQ>
__global__ void __launch_bounds__(GQ * 32)
prefill_attn_kernel(const PrefillAttnParams p) {
  extern __shared__ __align__(128) unsigned char smem[];
  using L = PrefillSmem<GQ>;
  const unsigned tl = kernel_prologue(kKidAttn);
  const int j = blockIdx.x, tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const bool bf = p.bf16 != 0;
  const QTile qt = p.tiles[blockIdx.y];
  const int n_heads = p.n_heads, n_kv = p.n_kv;
  const int head = j * GQ + warp;
  const int max_ctx = qt.ctx0 + qt.nrows - 1;
  const int nchunks = (max_ctx + kChunkTokens - 1) / kChunkTokens;
  const int nblocks = (max_ctx + kBlockTokens - 1) / kBlockTokens;
  const int32_t* bt = p.block_tables + (size_t)qt.bt_row * p.bt_stride;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L::kOffBar);

  auto issue_chunk = [&](int c) {   // thread 0
    const int buf = c & 1;
    unsigned char* Kt = smem + L::kOffKV + buf * 2 * kTileBytes;
    unsigned char* Vt = Kt + kTileBytes;
    const int b0 = c * kChunkBlocks;
    const int nb = min(kChunkBlocks, nblocks - b0);
    mbar_arrive_expect_tx(&bars[buf], (uint32_t)nb * 2u * kKvBlockBytes);
    for (int i = 0; i < nb; ++i) {
      const size_t off = ((size_t)bt[b0 + i] * n_kv + j) * (size_t)(kBlockTokens * kHeadDim);
      bulk_g2s(Kt + i * kKvBlockBytes, p.k_cache + off, kKvBlockBytes, &bars[buf]);
      bulk_g2s(Vt + i * kKvBlockBytes, p.v_cache + off, kKvBlockBytes, &bars[buf]);
    }
  };
  if (tid == 0) {
    mbar_init(&bars[0], 1);
    mbar_init(&bars[1], 1);
    fence_mbar_init();
    issue_chunk(0);
    if (nchunks > 1) issue_chunk(1);
  }

  // ---- this warp's Q tile: 16 rows x 128 dims, swizzled; rows >= nrows are zero ----
  unsigned char* q_s = smem + L::kOffQ + warp * (16 * 256);
  for (int i = lane; i < 16 * 16; i += 32) {
    const int r = i >> 4, cc = i & 15;
    uint4 v = make_uint4(0, 0, 0, 0);
    if (r < qt.nrows) v = *reinterpret_cast<const uint4*>(p.q + ((size_t)(qt.row0 + r) * n_heads + head) * kHeadDim + cc * 8);
    *reinterpret_cast<uint4*>(q_s + r * 256 + ((cc ^ (r & 7)) << 4)) = v;
  }
  __syncwarp();
  uint32_t qa[8][4];
  {
    const int mj = lane >> 3, mr = lane & 7;
    const int r = (mj & 1) * 8 + mr;
#pragma unroll
    for (int ks = 0; ks < 8; ++ks) {
      const int cc = 2 * ks + (mj >> 1);
      ldmatrix_x4(smem_u32(q_s + r * 256 + ((cc ^ (r & 7)) << 4)), qa[ks][0], qa[ks][1], qa[ks][2], qa[ks][3]);
    }
  }
  __syncthreads();                                  // barrier init visible to every warp

  const int g = lane >> 2, tq = lane & 3;
  float o[16][4];
#pragma unroll
  for (int dn = 0; dn < 16; ++dn) o[dn][0] = o[dn][1] = o[dn][2] = o[dn][3] = 0.f;
  float m_run[2] = {-INFINITY, -INFINITY}, l_run[2] = {0.f, 0.f};
  const int ctx_r[2] = {qt.ctx0 + g, qt.ctx0 + g + 8};     // keys visible to rows g and g + 8 (causal)
  // scores are kept in the log2 domain (1/sqrt(128) * log2 e folded into one multiply): p = 2^(s - m) is one
  // FADD + one MUFU.EX2 per element where expf() is a dozen instructions -- the softmax, not the MMAs, paced this kernel
  const float scale = 0.08838834764831845f * 1.4426950408889634f;

  for (int c = 0; c < nchunks; ++c) {
    const int buf = c & 1;
    unsigned char* Kt = smem + L::kOffKV + buf * 2 * kTileBytes;
    unsigned char* Vt = Kt + kTileBytes;
    mbar_wait(&bars[buf], (uint32_t)(c >> 1) & 1u);
    const int nvalid = min(kChunkTokens, max_ctx - c * kChunkTokens);
    const int nb = (nvalid + kBlockTokens - 1) / kBlockTokens;
    // every key of the chunk is visible to every row of the tile (all but the last one or two chunks of a prompt):
    // no causal compares
    const bool need_mask = (c + 1) * kChunkTokens > qt.ctx0;

    float sc[8][4];
#pragma unroll
    for (int nt = 0; nt < 8; ++nt) {
      float d[4] = {0.f, 0.f, 0.f, 0.f};
      if (nt * 8 < nvalid) {
        const int mj = lane >> 3, tok = nt * 8 + (lane & 7);
        const uint32_t rowaddr = smem_u32(Kt + (tok >> 4) * kKvBlockBytes + (tok & 15) * 256);
#pragma unroll
        for (int k2 = 0; k2 < 4; ++k2) {
          uint32_t b0, b1, b2, b3;
          ldmatrix_x4(rowaddr + ((((4 * k2 + mj) ^ (tok & 7))) << 4), b0, b1, b2, b3);
          mma_16816(d, qa[2 * k2], b0, b1, bf);
          mma_16816(d, qa[2 * k2 + 1], b2, b3, bf);
        }
      }
      if (need_mask) {
        const int t0 = c * kChunkTokens + nt * 8 + 2 * tq;     // absolute key position
        sc[nt][0] = t0 < ctx_r[0] ? d[0] * scale : -INFINITY;
        sc[nt][1] = t0 + 1 < ctx_r[0] ? d[1] * scale : -INFINITY;
        sc[nt][2] = t0 < ctx_r[1] ? d[2] * scale : -INFINITY;
        sc[nt][3] = t0 + 1 < ctx_r[1] ? d[3] * scale : -INFINITY;
      } else {
        sc[nt][0] = d[0] * scale; sc[nt][1] = d[1] * scale; sc[nt][2] = d[2] * scale; sc[nt][3] = d[3] * scale;
      }
    }
    uint32_t pa[4][4];
#pragma unroll
    for (int hr = 0; hr < 2; ++hr) {                            // row g (hr = 0) and row g + 8 (hr = 1)
      float mx = -INFINITY;
#pragma unroll
      for (int nt = 0; nt < 8; ++nt) mx = fmaxf(mx, fmaxf(sc[nt][2 * hr], sc[nt][2 * hr + 1]));
      mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
      mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
      const float m_new = fmaxf(m_run[hr], mx);
      // a row whose keys are all masked so far (padding rows of a short tile) keeps max = -inf: use 0
      const float m_use = m_new == -INFINITY ? 0.f : m_new;
      const float corr = ex2_approx(m_run[hr] - m_use);
      float sum = 0.f;
#pragma unroll
      for (int nt = 0; nt < 8; ++nt) {
        const float p0 = ex2_approx(sc[nt][2 * hr] - m_use), p1 = ex2_approx(sc[nt][2 * hr + 1] - m_use);
        sum += p0 + p1;
        pa[nt >> 1][(nt & 1) * 2 + hr] = pack16x2(p0, p1, bf);
      }
      sum += __shfl_xor_sync(0xffffffffu, sum, 1);
      sum += __shfl_xor_sync(0xffffffffu, sum, 2);
      l_run[hr] = l_run[hr] * corr + sum;
      m_run[hr] = m_new;
#pragma unroll
      for (int dn = 0; dn < 16; ++dn) { o[dn][2 * hr] *= corr; o[dn][2 * hr + 1] *= corr; }
    }
#pragma unroll
    for (int kt = 0; kt < 4; ++kt) {
      if (kt < nb) {
        const int mj = lane >> 3, r = (mj & 1) * 8 + (lane & 7);
        const uint32_t rowaddr = smem_u32(Vt + kt * kKvBlockBytes + r * 256);
#pragma unroll
        for (int d2 = 0; d2 < 8; ++d2) {
          uint32_t b0, b1, b2, b3;
          ldmatrix_x4_trans(rowaddr + (((2 * d2 + (mj >> 1)) ^ (r & 7)) << 4), b0, b1, b2, b3);
          mma_16816(o[2 * d2], pa[kt], b0, b1, bf);
          mma_16816(o[2 * d2 + 1], pa[kt], b2, b3, bf);
        }
      }
    }
    __syncthreads();                                            // every warp is done with this buffer
    if (tid == 0 && c + 2 < nchunks) issue_chunk(c + 2);
  }

  // ---- write O / l for rows g and g + 8 ----
#pragma unroll
  for (int hr = 0; hr < 2; ++hr) {
    const int r = g + 8 * hr;
    if (r < qt.nrows) {
      const float inv = 1.f / l_run[hr];
      __half* orow = p.out + ((size_t)(qt.row0 + r) * n_heads + head) * kHeadDim;
#pragma unroll
      for (int dn = 0; dn < 16; ++dn)
        *reinterpret_cast<uint32_t*>(orow + dn * 8 + 2 * tq) = pack16x2(o[dn][2 * hr] * inv, o[dn][2 * hr + 1] * inv, bf);
    }
  }
  kernel_epilogue(tl);
}

template <int GQ>
static int launch_prefill_gq(const PrefillAttnParams& p, int n_tiles, cudaStream_t s) {
  auto kern = prefill_attn_kernel<GQ>;
  B200_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, PrefillSmem<GQ>::kBytes));   // per device: every launch
  dim3 grid(p.n_kv, n_tiles);
  B200_CUDA(launch_k(kern, grid, dim3(GQ * 32), (size_t)PrefillSmem<GQ>::kBytes, s, p));
  g_launch_count += 1;
  B200_CUDA(cudaGetLastError());
  return B200_OK;
}

int launch_prefill_attn(const __half* q, const __half* k_cache, const __half* v_cache, const int32_t* block_tables,
                        int bt_stride, const int32_t* tiles, int n_tiles, int n_heads, int n_kv, __half* out,
                        cudaStream_t s, bool bf16) {
  B200_REQUIRE(n_kv > 0 && n_heads % n_kv == 0, "prefill_attn: n_heads % n_kv");
  B200_REQUIRE(n_tiles <= 65535, "prefill_attn: at most 65535 query tiles per launch");
  if (n_tiles == 0) return B200_OK;
  PrefillAttnParams p{};
  p.q = q; p.k_cache = k_cache; p.v_cache = v_cache; p.block_tables = block_tables; p.bt_stride = bt_stride;
  p.tiles = reinterpret_cast<const QTile*>(tiles); p.n_heads = n_heads; p.n_kv = n_kv; p.out = out; p.bf16 = bf16 ? 1 : 0;
#define CASE(G) case G: return launch_prefill_gq<G>(p, n_tiles, s)
  switch (n_heads / n_kv) {
    CASE(1); CASE(2); CASE(3); CASE(4); CASE(5); CASE(6); CASE(7); CASE(8);
    default:
      set_last_error("prefill_attn: n_heads / n_kv_heads must be in 1..8");
      return B200_ERR_INVALID;
  }
#undef CASE
}

template <int GQ, bool kFused, int kPhys>
static int launch_gq_phys(const AttnParams& p, int T, cudaStream_t s) {
  auto kern = paged_attn_kernel<GQ, kFused, kPhys>;
  constexpr int smem_bytes = AttnSmemT<kPhys>::kBytes;
  B200_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));   // idempotent, host-side
  dim3 grid(p.n_kv, T, p.n_splits);
  B200_CUDA(launch_k(kern, grid, dim3(kAttnThreads), (size_t)smem_bytes, s, p));
  g_launch_count += 1;
  B200_CUDA(cudaGetLastError());
  return B200_OK;
}

template <int GQ, bool kFused>
static int launch_gq(const AttnParams& p, int T, cudaStream_t s) {
  // More CTAs than SMs (decode batches above ~64 rows): three tile-owning warps, ~101 KB, two CTAs per SM --
  // half the waves.  Same bits as the four-warp kernel (virtual warps, attn.cuh).
  if constexpr (kFused) {
    if ((int64_t)p.n_kv * T * p.n_splits > 148) return launch_gq_phys<GQ, kFused, 3>(p, T, s);
  }
  return launch_gq_phys<GQ, kFused, 4>(p, T, s);
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
                      int n_splits, float* scratch, int32_t* counters, __half* out, cudaStream_t s, bool bf16) {
  AttnParams p{};
  p.bf16 = bf16 ? 1 : 0;
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
                             int32_t* counters, __half* out, cudaStream_t s, bool bf16) {
  B200_REQUIRE(n_parts >= 1, "attn_decode_fused: n_parts >= 1");
  AttnParams p{};
  p.bf16 = bf16 ? 1 : 0;
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
