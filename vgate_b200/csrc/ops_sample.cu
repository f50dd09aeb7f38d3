// Sampling kernel: greedy argmax, or temperature + top-p + exponential-race draw.
// One CTA per row of logits [B, V] fp32.
//
// Semantics follow vLLM (sampler.py:223-236 temperature / greedy threshold 1e-5;
// topk_topp_sampler.py:349-369 top-p; :395-416 argmax(probs / Exp(1))):
//   keep token i  <=>  (probability mass of tokens strictly more likely than i) < top_p
// which is what the reference's "ascending sort, cumsum, drop cumsum <= 1 - p" computes.
// No sort: the threshold is found by a 5-pass select over fixed-point probability masses
// (uint64 sums are associative, so the result is deterministic), then one pass draws
// argmax((x_i - max) - log(-log u_i)) over the kept set with the counter hash u(seed, step, i).
#include "common.cuh"
#include "internal.h"

namespace b200 {

B200_DEFINE_TL_SETTER(tl_set_sample)

constexpr int kSampleThreads = 1024;

__device__ __forceinline__ uint32_t float_key(float f) {   // order-preserving float -> uint32
  const uint32_t b = __float_as_uint(f);
  return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}

__device__ __forceinline__ float hash_uniform(uint32_t seed, uint32_t step, uint32_t i) {
  uint32_t x = i * 0x9E3779B1u + step * 0x85EBCA77u + seed * 0xC2B2AE3Du;
  x ^= x >> 16; x *= 0x7FEB352Du; x ^= x >> 15; x *= 0x846CA68Bu; x ^= x >> 16;
  // 23 hash bits: (x >> 9) + 0.5 <= 8388607.5 is exact in fp32 and the product stays strictly inside (0, 1).
  // (With 24 bits, 16777215 + 0.5 rounds up to 2^24 and u becomes 1.0: -log(u) = 0 and the token's race
  // score is +inf whatever its probability.)
  return ((float)(x >> 9) + 0.5f) * (1.0f / 8388608.0f);
}

struct ArgMax { float v; int i; };
__device__ __forceinline__ ArgMax better(ArgMax a, ArgMax b) {
  return (b.v > a.v || (b.v == a.v && b.i < a.i)) ? b : a;
}
__device__ ArgMax block_argmax(ArgMax a, ArgMax* red) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    ArgMax b;
    b.v = __shfl_xor_sync(0xffffffffu, a.v, o);
    b.i = __shfl_xor_sync(0xffffffffu, a.i, o);
    a = better(a, b);
  }
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  __syncthreads();
  if (lane == 0) red[warp] = a;
  __syncthreads();
  if (warp == 0) {
    a = lane < kSampleThreads / 32 ? red[lane] : ArgMax{-INFINITY, 0x7fffffff};
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      ArgMax b;
      b.v = __shfl_xor_sync(0xffffffffu, a.v, o);
      b.i = __shfl_xor_sync(0xffffffffu, a.i, o);
      a = better(a, b);
    }
    if (lane == 0) red[0] = a;
  }
  __syncthreads();
  return red[0];
}

// One CLUSTER of 4 CTAs per row; each CTA owns a contiguous quarter of the vocabulary.  (One CTA per row --
// 64 CTAs at batch 64, seven passes over 151 936 logits each, every thread adding 64-bit masses into one
// shared histogram -- took 1.4 ms per step with the reference's default temperature 0.7 / top_p 0.9: more
// than the model.)  Per-warp private histograms take the contention out of the mass sums; the CTAs
// exchange their maxima and histograms through distributed shared memory and every CTA then takes the same
// decisions from the same integer sums, so the result does not depend on how the row is cut.  (Four, not
// eight: only 33 clusters of 8 fit on the GPU at once -- two waves for 64 rows; 1.32 -> 1.28 ms per step.)
constexpr int kSC = 4;                       // CTAs per row
constexpr int kST = 512;                     // threads per CTA
constexpr int kSW = kST / 32;
constexpr int kSampleCand = 4096;            // candidate list per CTA (elements of the threshold bin): x and index, 32 KB

template <typename T>
__device__ __forceinline__ T dsmem_ld(const T* local, int cta) {
  uint32_t remote;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(smem_u32(local)), "r"(cta));
  T v;
  if constexpr (sizeof(T) == 8) {
    unsigned long long x;
    asm volatile("ld.shared::cluster.u64 %0, [%1];" : "=l"(x) : "r"(remote) : "memory");
    v = *reinterpret_cast<T*>(&x);
  } else {
    static_assert(sizeof(T) == 4, "dsmem_ld: 4 or 8 bytes");
    uint32_t x;
    asm volatile("ld.shared::cluster.u32 %0, [%1];" : "=r"(x) : "r"(remote) : "memory");
    v = *reinterpret_cast<T*>(&x);
  }
  return v;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}

// CTA-wide argmax, then cluster-wide: every CTA returns the row's result.  Each call of a kernel uses its own
// exchange slot, so no barrier is needed after the remote reads -- except before the CTA exits (the caller's job).
__device__ ArgMax cluster_argmax(ArgMax a, ArgMax* red, ArgMax* xchg) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    ArgMax b{__shfl_xor_sync(0xffffffffu, a.v, o), __shfl_xor_sync(0xffffffffu, a.i, o)};
    a = better(a, b);
  }
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) red[warp] = a;
  __syncthreads();
  if (threadIdx.x == 0) {
    ArgMax m = red[0];
    for (int w = 1; w < kSW; ++w) m = better(m, red[w]);
    *xchg = m;
  }
  cluster_sync_all();
  ArgMax m{dsmem_ld(&xchg->v, 0), dsmem_ld(&xchg->i, 0)};
  for (int c = 1; c < kSC; ++c) m = better(m, ArgMax{dsmem_ld(&xchg->v, c), dsmem_ld(&xchg->i, c)});
  return m;
}

__global__ void __cluster_dims__(kSC, 1, 1) __launch_bounds__(kST)
sample_kernel(const float* __restrict__ logits, int V, const float* __restrict__ temperature,
              const float* __restrict__ top_p, const uint32_t* __restrict__ seed, const int32_t* __restrict__ step,
              int32_t* __restrict__ out) {
  __shared__ ArgMax red[kSW];
  __shared__ ArgMax xchg[2];
  // per-warp mass histograms, as (low word, high word) pairs: a 64-bit shared-memory atomicAdd compiles to a
  // compare-and-swap loop (it was half of the kernel's stall samples); 32-bit adds are native, and the carry
  // out of the low word is recovered from the value the add returns
  __shared__ unsigned int wlo[kSW][256], whi[kSW][256];
  __shared__ unsigned long long hist[2][256];          // this CTA's histogram (read by the other CTAs), by level parity
  __shared__ unsigned long long ghist[256];            // the row's histogram
  __shared__ int n_cand;
  extern __shared__ __align__(16) unsigned char dyn_smem[];
  float* cand = reinterpret_cast<float*>(dyn_smem);     // x of this CTA's elements in the threshold bin
  int* cand_i = reinterpret_cast<int*>(dyn_smem + kSampleCand * 4);   // ... and their vocabulary indices
  const unsigned tl = kernel_prologue(kKidSample);
  const int r = blockIdx.x / kSC, rank = blockIdx.x % kSC, tid = threadIdx.x, warp = tid >> 5;
  const float* l = logits + (size_t)r * V;
  const int per = (V + kSC - 1) / kSC;
  const int i0 = rank * per, i1 = min(V, i0 + per);
  const float temp = temperature ? temperature[r] : 0.f;

  // f(logit, index) over this CTA's slice, eight coalesced loads in flight per thread (a plain strided loop
  // is one dependent L2 round trip per element, 37 per thread and pass)
  auto for_slice = [&](auto&& f) {
    for (int base = i0 + tid; base < i1; base += 8 * kST) {
      float v[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) v[u] = base + u * kST < i1 ? __ldg(l + base + u * kST) : 0.f;
#pragma unroll
      for (int u = 0; u < 8; ++u)
        if (base + u * kST < i1) f(v[u], base + u * kST);
    }
  };
  // pass 0: argmax (greedy result, and the max for the softmax)
  ArgMax best{-INFINITY, 0x7fffffff};
  for_slice([&](float v, int i) { best = better(best, ArgMax{v, i}); });
  best = cluster_argmax(best, red, &xchg[0]);
  if (temp < 1e-5f) {
    if (rank == 0 && tid == 0) out[r] = best.i;
    cluster_sync_all();                      // no CTA exits while another still reads its shared memory
    kernel_epilogue(tl);
    return;
  }
  // x_i = l_i / T exactly as the reference divides (sampler.py:232)
  const float m = best.v / temp;
  const float p = top_p ? top_p[r] : 1.f;
  uint32_t thr_key = 0;        // keep iff key(x_i) >= thr_key
  const uint32_t sd = seed ? seed[r] : 0u, st = step ? (uint32_t)step[r] : 0u;
  ArgMax draw{-INFINITY, 0x7fffffff};
  auto race = [&](float x, int i) {         // exponential race in log form: argmax of (x - m) + Gumbel(u_i)
    const float u = hash_uniform(sd, st, (uint32_t)i);
    const float score = (x - m) - logf(-logf(u));
    draw = better(draw, ArgMax{score, i});
  };
  bool drawn = false;          // the kept set has already been raced (top-p path with a candidate list)

  if (p < 1.f) {
    constexpr float kBinsPerUnit = 8.f;     // level 0: linear bins over d = m - x in [0, 32)
    auto bin0 = [&](float x) { return min(255, (int)((m - x) * kBinsPerUnit)); };
    auto mass = [&](float x) { return (unsigned long long)(expf(x - m) * 4294967296.f); };
    int n_hist = 0;
    auto add_mass = [&](int b, unsigned long long mv) {
      const unsigned int lo = (unsigned int)mv, hi = (unsigned int)(mv >> 32);
      const unsigned int old = atomicAdd(&wlo[warp][b], lo);
      const unsigned int carry = old + lo < old ? 1u : 0u;
      if (hi + carry) atomicAdd(&whi[warp][b], hi + carry);
    };
    // Sum the masses of this CTA's slice that pass `pred` into bins, combine the eight CTAs: ghist.
    auto histogram = [&](bool from_cand, auto&& bin_of) {
      for (int b = tid; b < kSW * 256; b += kST) { (&wlo[0][0])[b] = 0u; (&whi[0][0])[b] = 0u; }
      __syncthreads();
      if (from_cand) {
        for (int k = tid; k < n_cand; k += kST) {
          const float x = cand[k];
          const int b = bin_of(x);
          if (b >= 0) add_mass(b, mass(x));
        }
      } else {
        for_slice([&](float v, int) {
          const float x = v / temp;
          const int b = bin_of(x);
          if (b >= 0) add_mass(b, mass(x));
        });
      }
      __syncthreads();
      unsigned long long* mine = hist[n_hist & 1];
      if (tid < 256) {
        unsigned long long t = 0;
#pragma unroll
        for (int w = 0; w < kSW; ++w) t += ((unsigned long long)whi[w][tid] << 32) | wlo[w][tid];
        mine[tid] = t;
      }
      cluster_sync_all();
      if (tid < 256) {
        unsigned long long t = 0;
#pragma unroll
        for (int c = 0; c < kSC; ++c) t += dsmem_ld(&mine[tid], c);
        ghist[tid] = t;
      }
      // no second cluster barrier: the next histogram goes to the other buffer, and the one after that is
      // only written once every CTA has passed the barrier in between
      ++n_hist;
      __syncthreads();
    };
    // The bin (visited in order 0..255, or 255..0 when descending) at which the running mass first reaches
    // `target`: returns the bin, and adds the mass of the bins before it to `above`.  Warp-parallel (every warp
    // computes the same): lane j owns eight consecutive bins of the visiting order.
    auto select_bin = [&](bool descending, unsigned long long& above_io, unsigned long long target_v) {
      const int lane = tid & 31;
      unsigned long long h8[8], sum = 0;
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const int o = lane * 8 + k;
        h8[k] = ghist[descending ? 255 - o : o];
        sum += h8[k];
      }
      unsigned long long incl = sum;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const unsigned long long up = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += up;
      }
      const unsigned hit = __ballot_sync(0xffffffffu, above_io + incl >= target_v && sum != 0);
      const int src = hit ? __ffs(hit) - 1 : 31;          // (no lane reaches the target: cannot happen, take the last)
      unsigned long long run = above_io + incl - sum;
      int sel = -1;
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        if (h8[k] == 0 || sel >= 0) continue;
        if (run + h8[k] >= target_v) sel = descending ? 255 - (lane * 8 + k) : lane * 8 + k;
        else run += h8[k];
      }
      above_io = __shfl_sync(0xffffffffu, run, src);
      return __shfl_sync(0xffffffffu, sel, src);
    };
    // ---- level 0 ----
    histogram(false, [&](float x) { return bin0(x); });
    unsigned long long above = 0, target = 0;
    {
      unsigned long long part = 0;
      for (int b = (tid & 31); b < 256; b += 32) part += ghist[b];
#pragma unroll
      for (int d = 16; d > 0; d >>= 1) part += __shfl_xor_sync(0xffffffffu, part, d);
      target = (unsigned long long)((double)p * (double)part);   // part == the row's total mass (integer: order-free)
    }
    const int sel0 = select_bin(false, above, target);           // bin 0 holds the most likely tokens
    // One more pass over the slice does two things.  Elements in bins before sel0 are more likely than anything
    // in the threshold bin, so they are kept whatever the exact threshold turns out to be: they enter the draw
    // right away.  The elements of the threshold bin (a few percent of the slice) are collected; the four
    // radix levels below and the rest of the draw walk that list instead of the slice.
    if (tid == 0) n_cand = 0;
    __syncthreads();
    for_slice([&](float v, int i) {
      const float x = v / temp;
      const int b = bin0(x);
      if (b < sel0) race(x, i);
      else if (b == sel0) {
        const int k = atomicAdd(&n_cand, 1);
        if (k < kSampleCand) { cand[k] = x; cand_i[k] = i; }
      }
    });
    __syncthreads();
    const bool use_cand = n_cand <= kSampleCand;        // else (degenerate rows) the levels rescan the slice
    // ---- levels 1..4: radix select on the float key inside bin sel0 ----
    uint32_t prefix = 0, pmask = 0;
    for (int level = 0; level < 4; ++level) {
      const int shift = 24 - 8 * level;
      histogram(use_cand, [&](float x) {
        if (bin0(x) != sel0) return -1;
        const uint32_t key = float_key(x);
        return (key & pmask) != prefix ? -1 : (int)((key >> shift) & 255u);
      });
      const int sel = select_bin(true, above, target);           // larger key = more likely
      prefix |= (uint32_t)sel << shift;
      pmask |= 255u << shift;
    }
    thr_key = prefix;
    if (use_cand) {              // finish the draw with the kept part of the threshold bin
      for (int k = tid; k < n_cand; k += kST)
        if (float_key(cand[k]) >= thr_key) race(cand[k], cand_i[k]);
      drawn = true;
    } else {
      draw = ArgMax{-INFINITY, 0x7fffffff};                        // degenerate row: start over with the exact threshold
    }
  }

  // ---- draw over the kept set (no top-p, or a threshold bin too large for the candidate list) ----
  if (!drawn)
    for_slice([&](float v, int i) {
      const float x = v / temp;
      if (float_key(x) >= thr_key) race(x, i);
    });
  draw = cluster_argmax(draw, red, &xchg[1]);
  if (rank == 0 && tid == 0) out[r] = draw.i;
  cluster_sync_all();                        // no CTA exits while another still reads its shared memory
  kernel_epilogue(tl);
}

int launch_sample(const float* logits, int B, int V, const float* temperature, const float* top_p,
                  const uint32_t* seed, const int32_t* step, int32_t* out, cudaStream_t s) {
  if (B == 0) return B200_OK;
  // static + dynamic shared memory exceeds the 48 KB default (host-side attribute, per device, cheap to repeat)
  B200_CUDA(cudaFuncSetAttribute(sample_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSampleCand * 8));
  launch_k(sample_kernel, dim3(B * kSC), dim3(kST), (size_t)kSampleCand * 8, s, logits, V, temperature, top_p, seed, step, out);
  g_launch_count += 1;
  B200_CUDA(cudaGetLastError());
  return B200_OK;
}

// Greedy decoding without materialised logits: the lm_head GEMM epilogue (ops_umma.cu, kUmmaArgmaxF32)
// leaves, per 128-row vocabulary tile and token, the tile maximum and its lowest index as
// float2 {value, bits(index)} in part[tile * ld + token]; this kernel finishes the argmax per token with
// the same tie rule as sample_kernel (highest value, then lowest index).
__global__ void __launch_bounds__(kSampleThreads)
argmax_reduce_kernel(const float2* __restrict__ part, int n_tiles, int ld, int32_t* __restrict__ out) {
  __shared__ ArgMax red[kSampleThreads / 32];
  const unsigned tl = kernel_prologue(kKidSample);
  const int r = blockIdx.x;
  ArgMax best{-INFINITY, 0x7fffffff};
  for (int t = threadIdx.x; t < n_tiles; t += kSampleThreads) {
    const float2 v = __ldcg(part + (size_t)t * ld + r);
    best = better(best, ArgMax{v.x, __float_as_int(v.y)});
  }
  best = block_argmax(best, red);
  if (threadIdx.x == 0) out[r] = best.i;
  kernel_epilogue(tl);
}

int launch_argmax_reduce(const float* part, int n_tiles, int ld, int B, int32_t* out, cudaStream_t s) {
  if (B == 0) return B200_OK;
  launch_k(argmax_reduce_kernel, dim3(B), dim3(kSampleThreads), 0, s, reinterpret_cast<const float2*>(part), n_tiles, ld, out);
  g_launch_count += 1;
  B200_CUDA(cudaGetLastError());
  return B200_OK;
}

}  // namespace b200
