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
//   1. adds this rank's partials of the row in their fixed order and PUSHES the sum into every other rank's receive
//      buffer over NVLink as 8-byte {value, sequence number} words,
//   2. polls its own receive buffer until every word of the row carries this site's sequence number -- the data is
//      its own arrival flag, so there is no fence and no separate flag round trip (a release store at system scope,
//      which a flag protocol needs, took ~10 us under load here; an 8-byte aligned store is single-copy atomic),
//   3. adds the per-rank sums in rank order into the residual stream -- every rank computes the same bits -- and
//      normalises as rmsnorm_kernel does.
// One third of the partial bytes cross NVLink (three split-K partials per rank at the 14B shapes) at half wire
// efficiency, and a row waits for its own row only.  Sequence numbers are unique per (pass, site) and compared for
// equality, so the words never need resetting; the two receive slots alternate by site (a peer two sites ahead is
// impossible: it needs this rank's words of the site in between).
// Word layout of one row quarter in a receive buffer: plane k (component k of every float4) is contiguous, so the 32
// lanes of a warp store / load 256 contiguous bytes per instruction.
__device__ __forceinline__ void st_sys_u2(uint2* p, uint32_t x, uint32_t y) {
  asm volatile("st.relaxed.sys.global.v2.u32 [%0], {%1, %2};" ::"l"(p), "r"(x), "r"(y) : "memory");
}
__device__ __forceinline__ uint2 ld_sys_u2(const uint2* p) {
  uint2 v;
  asm volatile("ld.relaxed.sys.global.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "l"(p) : "memory");
  return v;
}
// step 1: own[it] = sum_p parts[p][row] (fixed p order) for this thread's float4s of quarter q, pushed to the peers;
// hv[it] = h[row] (loaded here so that its latency overlaps)
__device__ __forceinline__ void tp_own_sum_push(const float* __restrict__ h, const float* __restrict__ parts, const int n_parts,
                                                const int64_t part_stride, const TpPeers& tp, const uint32_t seq, const int row,
                                                const int q, const int H, const int tid, float4 (&own)[kNormMaxPerThread],
                                                float4 (&hv)[kNormMaxPerThread]) {
  constexpr int kBatch = 4;
  const int HQ = H / kNormCluster;
  const size_t in_off = (size_t)row * H + (size_t)q * HQ;
  const int nvec = HQ / 4;
  const size_t w_off = ((size_t)tp.rank * tp.rows_cap + row) * H + (size_t)q * HQ;     // my words in a peer's buffer
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
      for (int r = 0; r < tp.n_ranks; ++r) {
        if (r == tp.rank) continue;
        uint2* dst = tp.recv[r] + w_off + i;
        st_sys_u2(dst, __float_as_uint(a.x), seq);
        st_sys_u2(dst + nvec, __float_as_uint(a.y), seq);
        st_sys_u2(dst + 2 * nvec, __float_as_uint(a.z), seq);
        st_sys_u2(dst + 3 * nvec, __float_as_uint(a.w), seq);
      }
    }
  }
}
// steps 2 and 3: v = h + (sum over ranks, in rank order, of the per-rank sums); h updated; returns the warp's sum of squares
__device__ __forceinline__ float tp_collect(float* __restrict__ h, const TpPeers& tp, const uint32_t seq, const int row, const int q,
                                            const int H, const int tid, const float4 (&own)[kNormMaxPerThread],
                                            const float4 (&hv)[kNormMaxPerThread], float4 (&v)[kNormMaxPerThread]) {
  const int HQ = H / kNormCluster;
  const size_t in_off = (size_t)row * H + (size_t)q * HQ;
  const int nvec = HQ / 4;
  float4 s[kNormMaxPerThread];
  for (int r = 0; r < tp.n_ranks; ++r) {
    float4 x[kNormMaxPerThread];
    if (r == tp.rank) {
#pragma unroll
      for (int it = 0; it < kNormMaxPerThread; ++it) x[it] = own[it];
    } else {
      const uint2* src = tp.recv[tp.rank] + ((size_t)r * tp.rows_cap + row) * H + (size_t)q * HQ;
      const unsigned long long t0 = gtime();
      for (;;) {                                            // every word of the thread in flight at once, until all carry seq
        bool ok = true;
#pragma unroll
        for (int it = 0; it < kNormMaxPerThread; ++it) {
          const int i = tid + it * kNormThreads;
          if (i < nvec) {
            const uint2 w0 = ld_sys_u2(src + i), w1 = ld_sys_u2(src + nvec + i), w2 = ld_sys_u2(src + 2 * nvec + i),
                        w3 = ld_sys_u2(src + 3 * nvec + i);
            ok = ok && w0.y == seq && w1.y == seq && w2.y == seq && w3.y == seq;
            x[it] = make_float4(__uint_as_float(w0.x), __uint_as_float(w1.x), __uint_as_float(w2.x), __uint_as_float(w3.x));
          }
        }
        if (ok) break;
        if (gtime() - t0 > 20000000000ull) __trap();        // 20 s: the peer rank is gone; fail the context rather than hang
      }
    }
#pragma unroll
    for (int it = 0; it < kNormMaxPerThread; ++it) {
      if (r == 0) s[it] = x[it];
      else { s[it].x += x[it].x; s[it].y += x[it].y; s[it].z += x[it].z; s[it].w += x[it].w; }
    }
  }
  float ss = 0.f;
#pragma unroll
  for (int it = 0; it < kNormMaxPerThread; ++it) {
    const int i = tid + it * kNormThreads;
    if (i < nvec) {
      float4 a = hv[it];
      a.x += s[it].x; a.y += s[it].y; a.z += s[it].z; a.w += s[it].w;
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
