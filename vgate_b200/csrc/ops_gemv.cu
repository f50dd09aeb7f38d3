// CUDA-core W4A16 / fp16 GEMV for small decode batches: memory-bound weight streaming with
// 128-bit no-allocate loads, in-register AWQ dequant, Blackwell FHFMA (fp16 x fp16 + fp32) for
// exact-product fp32 accumulation, warp-shuffle reduction.  One warp owns one output row.
//
// Replaces, for the decode regime, what vLLM runs as awq_gemm (awq.py:276-283) and the cuBLAS
// lm_head GEMV (qwen2.py:522-534).  Larger batches go to the tcgen05 GEMM (ops_umma.cu).
#include "common.cuh"
#include "internal.h"

namespace b200 {

B200_DEFINE_TL_SETTER(tl_set_gemv)

constexpr int kGemvWarps = 8;

template <int BT>
__global__ void __launch_bounds__(kGemvWarps * 32)
w4_gemv_kernel(const uint32_t* __restrict__ wq, const __half* __restrict__ sc, const __half* __restrict__ zp,
               const __half* __restrict__ bias, const __half* __restrict__ x, int ldx, float* __restrict__ y,
               int ldy, int B, int K, int N, int G, int accumulate) {
  (void)kernel_prologue(kKidW4Gemv);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n = blockIdx.x * kGemvWarps + warp;
  if (n >= N) return;
  const int nchunks = K / 32;                 // one uint4 = 32 weights
  const int chunks_per_group = G / 32;
  const int ngroups = K / G;
  const uint4* wrow = reinterpret_cast<const uint4*>(wq + (size_t)n * (K / 8));
  const __half* srow = sc + (size_t)n * ngroups;
  const __half* zrow = zp + (size_t)n * ngroups;
  const float bn = bias ? __half2float(bias[n]) : 0.f;

  for (int b0 = 0; b0 < B; b0 += BT) {
    float acc[BT];
#pragma unroll
    for (int i = 0; i < BT; ++i) acc[i] = 0.f;
    for (int c = lane; c < nchunks; c += 32) {
      const uint4 w = ldg_stream(wrow + c);
      const int g = c / chunks_per_group;
      const __half2 s2 = __half2half2(srow[g]);
      const __half2 z2 = __half2half2(zrow[g]);
      const uint32_t su = *reinterpret_cast<const uint32_t*>(&s2), zu = *reinterpret_cast<const uint32_t*>(&z2);
      uint32_t wd[16];
      dequant8(w.x, zu, su, wd + 0);
      dequant8(w.y, zu, su, wd + 4);
      dequant8(w.z, zu, su, wd + 8);
      dequant8(w.w, zu, su, wd + 12);
#pragma unroll
      for (int i = 0; i < BT; ++i) {
        const int b = min(b0 + i, B - 1);
        const __half* xp = x + (size_t)b * ldx + c * 32;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const uint4 xv = ldg_cached(xp + 8 * q);
          acc[i] = fhfma2(wd[4 * q + 0], xv.x, acc[i]);
          acc[i] = fhfma2(wd[4 * q + 1], xv.y, acc[i]);
          acc[i] = fhfma2(wd[4 * q + 2], xv.z, acc[i]);
          acc[i] = fhfma2(wd[4 * q + 3], xv.w, acc[i]);
        }
      }
    }
#pragma unroll
    for (int i = 0; i < BT; ++i) {
      const float v = warp_sum(acc[i]);
      if (lane == 0 && b0 + i < B) {
        float* yp = y + (size_t)(b0 + i) * ldy + n;
        *yp = (accumulate ? *yp : 0.f) + v + bn;
      }
    }
  }
}

int launch_w4_gemv(const W4Linear& w, const __half* x, int ldx, float* y, int ldy, int B, bool accumulate,
                   cudaStream_t s) {
  B200_REQUIRE(w.K % w.G == 0 && w.G % 32 == 0 && ldx % 8 == 0, "w4_gemv: K % G, G % 32, ldx % 8");
  if (B == 0) return B200_OK;
  const int grid = (w.N + kGemvWarps - 1) / kGemvWarps;
  const int threads = kGemvWarps * 32;
#define LAUNCH(BT) \
  launch_k(w4_gemv_kernel<BT>, dim3(grid), dim3(threads), 0, s, w.wq, w.sc, w.zp, w.bias, x, ldx, y, ldy, B, w.K, w.N, w.G, accumulate ? 1 : 0)
  if (B == 1) LAUNCH(1);
  else if (B == 2) LAUNCH(2);
  else if (B <= 4) LAUNCH(4);
  else LAUNCH(8);
#undef LAUNCH
  g_launch_count += 1;
  B200_CUDA(cudaGetLastError());
  return B200_OK;
}

// Dense fp16 weights W[N, K] (tied lm_head).
template <int BT>
__global__ void __launch_bounds__(kGemvWarps * 32)
f16_gemv_kernel(const __half* __restrict__ w, const __half* __restrict__ x, int ldx, float* __restrict__ y, int ldy,
                int B, int K, int N, int accumulate) {
  (void)kernel_prologue(kKidF16Gemv);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n = blockIdx.x * kGemvWarps + warp;
  if (n >= N) return;
  const int nchunks = K / 8;
  const uint4* wrow = reinterpret_cast<const uint4*>(w + (size_t)n * K);
  for (int b0 = 0; b0 < B; b0 += BT) {
    float acc[BT];
#pragma unroll
    for (int i = 0; i < BT; ++i) acc[i] = 0.f;
    for (int c = lane; c < nchunks; c += 32) {
      const uint4 wv = ldg_stream(wrow + c);
#pragma unroll
      for (int i = 0; i < BT; ++i) {
        const int b = min(b0 + i, B - 1);
        const uint4 xv = ldg_cached(x + (size_t)b * ldx + c * 8);
        acc[i] = fhfma2(wv.x, xv.x, acc[i]);
        acc[i] = fhfma2(wv.y, xv.y, acc[i]);
        acc[i] = fhfma2(wv.z, xv.z, acc[i]);
        acc[i] = fhfma2(wv.w, xv.w, acc[i]);
      }
    }
#pragma unroll
    for (int i = 0; i < BT; ++i) {
      const float v = warp_sum(acc[i]);
      if (lane == 0 && b0 + i < B) {
        float* yp = y + (size_t)(b0 + i) * ldy + n;
        *yp = (accumulate ? *yp : 0.f) + v;
      }
    }
  }
}

int launch_f16_gemv(const __half* w, const __half* x, int ldx, float* y, int ldy, int B, int K, int N,
                    bool accumulate, cudaStream_t s) {
  B200_REQUIRE(K % 8 == 0 && ldx % 8 == 0, "f16_gemv: K % 8, ldx % 8");
  if (B == 0) return B200_OK;
  const int grid = (N + kGemvWarps - 1) / kGemvWarps;
  const int threads = kGemvWarps * 32;
#define LAUNCH(BT) launch_k(f16_gemv_kernel<BT>, dim3(grid), dim3(threads), 0, s, w, x, ldx, y, ldy, B, K, N, accumulate ? 1 : 0)
  if (B == 1) LAUNCH(1);
  else if (B == 2) LAUNCH(2);
  else if (B <= 4) LAUNCH(4);
  else LAUNCH(8);
#undef LAUNCH
  g_launch_count += 1;
  B200_CUDA(cudaGetLastError());
  return B200_OK;
}

}  // namespace b200
