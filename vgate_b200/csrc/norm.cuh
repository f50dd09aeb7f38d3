// Residual-sum + RMSNorm arithmetic shared by rmsnorm_kernel (ops_basic.cu: one 4-CTA cluster per row) and
// experiments that ran the same arithmetic inside other kernels (history: 8d75821 persistent kernel, f28aa97 norm tail of the split-K GEMM).  A row is cut in
// four quarters; 128 threads own a quarter; every quarter's warps contribute one partial sum of squares
// each, and the 16 partials are added in one fixed order, so both callers produce the same bits.
#pragma once
#include "common.cuh"
#include "internal.h"

namespace b200 {

constexpr int kNormCluster = 4;        // quarters per row
constexpr int kNormThreads = 128;      // threads per quarter
constexpr int kNormWarps = kNormThreads / 32;
constexpr int kNormMaxPerThread = 4;   // H <= 4 quarters * 128 threads * 4 floats * 4 = 8192

// v = h[row] + sum_p parts[p][row] over this thread's float4s of quarter q (fixed p order), h updated in
// place when n_parts > 0.  Returns the warp's sum of squares (same value in every lane).
// kBatch: partial loads in flight per pass (a register budget; the order of the additions does not depend on it).
template <int kBatch = 16>
__device__ __forceinline__ float rms_quarter_load(float* __restrict__ h, const float* __restrict__ parts, const int n_parts,
                                                  const int64_t part_stride, const int row, const int q, const int H,
                                                  const int tid, float4 (&v)[kNormMaxPerThread]) {
  const int HQ = H / kNormCluster;
  const size_t in_off = (size_t)row * H + (size_t)q * HQ;
  const int nvec = HQ / 4;
  float ss = 0.f;
#pragma unroll
  for (int it = 0; it < kNormMaxPerThread; ++it) {
    const int i = tid + it * kNormThreads;
    if (i < nvec) {
      float4 a = *reinterpret_cast<const float4*>(h + in_off + 4 * i);
      // partials: all loads (up to 16 per batch) are issued before the first add -- one L2 round trip
      // -- then added in the fixed order p = 0, 1, 2, ... so the sum is deterministic
      for (int p0 = 0; p0 < n_parts; p0 += kBatch) {
        float4 b[kBatch];
#pragma unroll
        for (int k = 0; k < kBatch; ++k)
          b[k] = p0 + k < n_parts
                     ? __ldcg(reinterpret_cast<const float4*>(parts + (size_t)(p0 + k) * part_stride + in_off + 4 * i))
                     : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int k = 0; k < kBatch; ++k) { a.x += b[k].x; a.y += b[k].y; a.z += b[k].z; a.w += b[k].w; }
      }
      if (n_parts > 0) *reinterpret_cast<float4*>(h + in_off + 4 * i) = a;
      v[it] = a;
      // explicit FMAs: the two callers must not be contracted differently by the compiler
      ss = __fmaf_rn(a.x, a.x, ss); ss = __fmaf_rn(a.y, a.y, ss); ss = __fmaf_rn(a.z, a.z, ss); ss = __fmaf_rn(a.w, a.w, ss);
    }
  }
  return warp_sum(ss);
}

// ---- tensor-parallel variant: the all-reduce is part of the norm -----------------------------------
// Every rank's o_proj / down_proj leaves its split-K partials of the row-parallel product in its own HBM.  Instead
// of an all-reduce kernel followed by the norm, the norm cluster of row r on every rank
//   1. adds this rank's partials of the row in their fixed order and EXPORTS the sum (local fp32 buffer),
//   2. tells the other ranks' clusters of the same row "exported" (one arrival word per row, site and writer, written
//      into the peers' HBM; the words only grow, nothing is ever reset) and waits for theirs,
//   3. reads the other ranks' exported sums with NVLink peer loads, adds the sums in rank order into the residual
//      stream -- every rank computes the same bits -- and normalises as rmsnorm_kernel does.
// One third of the partial bytes cross NVLink (three split-K partials per rank at the 14B shapes), and a row
// waits for its own row only.
// peer data: relaxed system-scope loads (peer lines may sit in this SM's L1 from an earlier step otherwise)
__device__ __forceinline__ float4 ld_sys_f4(const float* p) {
  float4 v;
  asm volatile("ld.relaxed.sys.global.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
  return v;
}
// step 1: own[it] = sum_p parts[p][row] (fixed p order) for this thread's float4s of quarter q, stored to `exp_out`
// as well; hv[it] = h[row] (loaded here so that its latency overlaps)
__device__ __forceinline__ void tp_own_sum(const float* __restrict__ h, const float* __restrict__ parts, const int n_parts,
                                           const int64_t part_stride, float* __restrict__ exp_out, const int row, const int q,
                                           const int H, const int tid, float4 (&own)[kNormMaxPerThread],
                                           float4 (&hv)[kNormMaxPerThread]) {
  constexpr int kBatch = 4;
  const int HQ = H / kNormCluster;
  const size_t in_off = (size_t)row * H + (size_t)q * HQ;
  const int nvec = HQ / 4;
#pragma unroll
  for (int it = 0; it < kNormMaxPerThread; ++it) {
    const int i = tid + it * kNormThreads;
    if (i < nvec) {
      hv[it] = *reinterpret_cast<const float4*>(h + in_off + 4 * i);
      float4 a = __ldcg(reinterpret_cast<const float4*>(parts + in_off + 4 * i));
      for (int p0 = 1; p0 < n_parts; p0 += kBatch) {
        float4 b[kBatch];
#pragma unroll
        for (int k = 0; k < kBatch; ++k)
          b[k] = p0 + k < n_parts
                     ? __ldcg(reinterpret_cast<const float4*>(parts + (size_t)(p0 + k) * part_stride + in_off + 4 * i))
                     : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int k = 0; k < kBatch; ++k) { a.x += b[k].x; a.y += b[k].y; a.z += b[k].z; a.w += b[k].w; }
      }
      own[it] = a;
      *reinterpret_cast<float4*>(exp_out + in_off + 4 * i) = a;
    }
  }
}
// step 2 for one row: called by all threads of the four CTAs of the row's cluster, after their export stores.
__device__ __forceinline__ void tp_row_rendezvous(const TpPeers& tp, const int row, const int q, const uint32_t epoch) {
  // the cluster barrier (release / acquire) orders the 512 threads' export stores before the signalling thread's
  // system-scope release store; release is cumulative, so the peer that acquires the word sees the exports.  (A
  // __threadfence_system() per thread instead cost several microseconds per site.)
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
  const int t = threadIdx.x;
  if (t < tp.n_ranks && t != tp.rank) {
    if (q == 0)
      asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(tp.flags[t] + (size_t)row * kMaxTp + tp.rank), "r"(epoch) : "memory");
    const uint32_t* f = tp.flags[tp.rank] + (size_t)row * kMaxTp + t;
    const unsigned long long t0 = gtime();
    for (;;) {
      uint32_t v;
      asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(f) : "memory");
      if ((int32_t)(v - epoch) >= 0) break;
      if (gtime() - t0 > 20000000000ull) __trap();         // 20 s: the peer rank is gone; fail the context rather than hang
    }
    asm volatile("fence.acq_rel.sys;" ::: "memory");       // acquire: the peer's exports, stored before its release, are visible now
  }
  __syncthreads();
}
// step 3: v = h + (sum over ranks, in rank order, of the exported sums); h updated; returns the warp's sum of squares
template <int kRanks>
__device__ __forceinline__ float tp_rank_sum(float* __restrict__ h, const TpPeers& tp, const int row, const int q, const int H,
                                             const int tid, const float4 (&own)[kNormMaxPerThread],
                                             const float4 (&hv)[kNormMaxPerThread], float4 (&v)[kNormMaxPerThread]) {
  const int HQ = H / kNormCluster;
  const size_t in_off = (size_t)row * H + (size_t)q * HQ;
  const int nvec = HQ / 4;
  float4 b[kNormMaxPerThread][kRanks];
#pragma unroll
  for (int it = 0; it < kNormMaxPerThread; ++it) {          // every peer load of the thread in flight at once
    const int i = tid + it * kNormThreads;
#pragma unroll
    for (int r = 0; r < kRanks; ++r)
      b[it][r] = (i < nvec && r != tp.rank) ? ld_sys_f4(tp.exports[r] + in_off + 4 * i) : own[it];
  }
  float ss = 0.f;
#pragma unroll
  for (int it = 0; it < kNormMaxPerThread; ++it) {
    const int i = tid + it * kNormThreads;
    if (i < nvec) {
      float4 s = b[it][0];
#pragma unroll
      for (int r = 1; r < kRanks; ++r) { s.x += b[it][r].x; s.y += b[it][r].y; s.z += b[it][r].z; s.w += b[it][r].w; }
      float4 a = hv[it];
      a.x += s.x; a.y += s.y; a.z += s.z; a.w += s.w;
      *reinterpret_cast<float4*>(h + in_off + 4 * i) = a;
      v[it] = a;
      ss = __fmaf_rn(a.x, a.x, ss); ss = __fmaf_rn(a.y, a.y, ss); ss = __fmaf_rn(a.z, a.z, ss); ss = __fmaf_rn(a.w, a.w, ss);
    }
  }
  return warp_sum(ss);
}

// red[quarter][warp] -> 1 / rms of the row
__device__ __forceinline__ float rms_scale(const float (*red)[kNormWarps], const int H, const float eps) {
  float tot = 0.f;
#pragma unroll
  for (int c = 0; c < kNormCluster; ++c)
#pragma unroll
    for (int w = 0; w < kNormWarps; ++w) tot += red[c][w];
  return rsqrtf(__fadd_rn(__fdiv_rn(tot, (float)H), eps));
}

// xn[r] quarter q = fp16(v * rs * gamma)
__device__ __forceinline__ void rms_quarter_store(const float4 (&v)[kNormMaxPerThread], const float rs,
                                                  const __half* __restrict__ gamma, __half* __restrict__ xn, const int r,
                                                  const int q, const int H, const int tid, const bool bf = false) {
  const int HQ = H / kNormCluster;
  const int nvec = HQ / 4;
#pragma unroll
  for (int it = 0; it < kNormMaxPerThread; ++it) {
    const int i = tid + it * kNormThreads;
    if (i < nvec) {
      const uint2 gw = *reinterpret_cast<const uint2*>(gamma + (size_t)q * HQ + 4 * i);
      const float2 g0 = unpack16x2(gw.x, bf);
      const float2 g1 = unpack16x2(gw.y, bf);
      uint2 o;
      o.x = pack16x2(__fmul_rn(__fmul_rn(v[it].x, rs), g0.x), __fmul_rn(__fmul_rn(v[it].y, rs), g0.y), bf);
      o.y = pack16x2(__fmul_rn(__fmul_rn(v[it].z, rs), g1.x), __fmul_rn(__fmul_rn(v[it].w, rs), g1.y), bf);
      *reinterpret_cast<uint2*>(xn + (size_t)r * H + (size_t)q * HQ + 4 * i) = o;
    }
  }
}

}  // namespace b200
