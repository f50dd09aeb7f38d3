// Internal C++ launcher interface shared by the op wrappers and the engine.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <stdint.h>
#include <string>

#include "../../include/b200gate.h"
#include "../../include/b200gate_debug.h"

namespace b200 {

constexpr int kHeadDim = 128;
constexpr int kBlockTokens = 16;   // KV block size in tokens (vLLM default block_size)

void set_last_error(const std::string& msg);
int cuda_fail(cudaError_t err, const char* what, const char* file, int line);

#define B200_CUDA(call)                                                        \
  do {                                                                         \
    cudaError_t _e = (call);                                                   \
    if (_e != cudaSuccess) return ::b200::cuda_fail(_e, #call, __FILE__, __LINE__); \
  } while (0)

#define B200_REQUIRE(cond, msg)                                                \
  do {                                                                         \
    if (!(cond)) { ::b200::set_last_error(std::string("invalid argument: ") + msg); return B200_ERR_INVALID; } \
  } while (0)

// Every launcher returns B200_OK or an error code and bumps the launch counter.
extern thread_local int64_t g_launch_count;

// Kernel launch with the programmatic-stream-serialization attribute (PDL) unless disabled
// (VGATE_B200_PDL=0): the kernel may start while its predecessor drains; kernels order themselves
// with griddepcontrol.wait (common.cuh pdl_prologue).
bool pdl_enabled();
template <typename... KArgs, typename... Args>
inline cudaError_t launch_k(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s, Args... args) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = s;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
}

// AWQ linear in kernel layout (see b200_op_awq_repack).
struct W4Linear {
  const uint32_t* wq = nullptr;  // [N, K/8]
  const __half* sc = nullptr;    // [N, K/G]
  const __half* zp = nullptr;    // [N, K/G]  (1024 + zero)
  const __half* bias = nullptr;  // [N] or null
  int K = 0, N = 0, G = 128;
};

// AWQ linear pre-tiled for the tcgen05 GEMM producers (ops_umma.cu).
struct UmmaW4 {
  const unsigned char* chunks = nullptr;   // [N/128][K/G] chunks: scales, zeros, packed k tiles (see ops_umma.cu)
  const __half* bias = nullptr;
  int K = 0, N = 0, G = 128;
  // Weights of the linear that runs next and cannot prefetch for itself (down after gate_up: its CTAs need the
  // SMs gate_up's CTAs occupy).  This kernel's TMA warp asks L2 for them once its own loads are issued.
  const unsigned char* next_chunks = nullptr;
  int64_t next_bytes = 0;
};
enum { kUmmaStoreF32 = 0, kUmmaAccumF32 = 1, kUmmaPartialF32 = 2, kUmmaSwigluF16 = 3, kUmmaArgmaxF32 = 4 };

size_t umma_w4_chunk_bytes(int K, int N, int G);
int umma_w4_effective_splits(int K, int G, int n_splits);
int launch_umma_tile_w4(const uint32_t* wq, const __half* sc, const __half* zp, int K, int N, int G, int pair_rows,
                        unsigned char* chunks, cudaStream_t s);
int make_tensor_map_f16(CUtensorMap* map, const void* base, int64_t rows, int64_t cols, int64_t ld, int box_rows);
int umma_pick_nt(int T, bool w4);
void umma_set_trace(unsigned long long* dev_buf, int grid_x, int grid_z);
int launch_umma_w4(const UmmaW4& w, const CUtensorMap& map_x, int nt_sel, int T, float* y, int ldy, int mode,
                   int n_splits, int64_t part_stride, cudaStream_t s);
int launch_umma_f16(const CUtensorMap& map_w, int K, int N, const CUtensorMap& map_x, int nt_sel, int T, float* y,
                    int ldy, int mode, cudaStream_t s, int n_splits = 1, int64_t part_stride = 0, const __half* bias = nullptr,
                    bool bf16 = false);
int umma_dense_effective_splits(int K, int n_splits);

int launch_awq_repack(const int32_t* qweight, const int32_t* qzeros, const __half* scales, int K, int N, int G,
                      uint32_t* wq, __half* sc, __half* zp, int row_offset, cudaStream_t s);
int launch_awq_dequant(const uint32_t* wq, const __half* sc, const __half* zp, int K, int N, int G, __half* w,
                       cudaStream_t s);
int launch_w4_gemv(const W4Linear& w, const __half* x, int ldx, float* y, int ldy, int B, bool accumulate,
                   cudaStream_t s);
int launch_f16_gemv(const __half* w, const __half* x, int ldx, float* y, int ldy, int B, int K, int N,
                    bool accumulate, cudaStream_t s);
int launch_rmsnorm(float* h, const float* parts, int n_parts, int64_t part_stride, const __half* gamma, float eps,
                   const int32_t* rows, int n_rows, int H, __half* xn, cudaStream_t s, bool bf16 = false);
// Tensor-parallel decode passes: the all-reduce after o_proj / down_proj is folded into the RMSNorm that consumes it
// (norm.cuh).  Plain pointers into every rank's HBM (peer access enabled between the ranks' devices).
constexpr int kMaxTp = 8;
constexpr int kTpNormMaxRows = 256;   // the rendezvous needs the whole grid resident: 4 CTAs of 128 threads per row
struct TpPeers {
  uint2* recv[kMaxTp];              // every rank's receive buffer of this site's slot: [writer rank][rows_cap][H] {value, seq} words
  int n_ranks, rank, rows_cap;
  uint32_t site, n_sites;           // sequence number of a word = pass counter * n_sites + site + 1
};
int launch_rmsnorm_tp(float* h, const float* parts, int n_parts, int64_t part_stride, const TpPeers& tp, const uint32_t* epoch,
                      const __half* gamma, float eps, const int32_t* rows, int n_rows, int H, __half* xn, cudaStream_t s, bool bf16);
int launch_tp_tick(uint32_t* epoch, cudaStream_t s);
int launch_rope_kv(const float* qkv, int n_parts, int64_t part_stride, const __half* bias, const int32_t* pos, const int32_t* slot, const float* cos_t,
                   const float* sin_t, int T, int n_heads, int n_kv, __half* q_out, __half* k_cache,
                   __half* v_cache, cudaStream_t s, bool bf16 = false);
int launch_paged_attn(const __half* q, const __half* k_cache, const __half* v_cache, const int32_t* block_tables,
                      int bt_stride, const int32_t* bt_row, const int32_t* ctx_len, int T, int n_heads, int n_kv,
                      int n_splits, float* scratch, int32_t* counters, __half* out, cudaStream_t s, bool bf16 = false);
int launch_attn_decode_fused(const float* qkv, int n_parts, int64_t part_stride, const __half* bias, const int32_t* pos,
                             const int32_t* slot, const float* cos_t, const float* sin_t, __half* k_cache,
                             __half* v_cache, const int32_t* block_tables, int bt_stride, const int32_t* bt_row,
                             const int32_t* ctx_len, int T, int n_heads, int n_kv, int n_splits, float* scratch,
                             int32_t* counters, __half* out, cudaStream_t s, bool bf16 = false);
void attn_set_trace(unsigned long long* dev_buf);
// tiles: int32 [n_tiles, 4] = {row0, nrows <= 16, block-table row, context length of row0}
int launch_prefill_attn(const __half* q, const __half* k_cache, const __half* v_cache, const int32_t* block_tables,
                        int bt_stride, const int32_t* tiles, int n_tiles, int n_heads, int n_kv, __half* out,
                        cudaStream_t s, bool bf16 = false);
size_t attn_scratch_floats(int T, int n_heads, int n_kv, int n_splits);
int launch_swiglu(const float* gu, int T, int I, __half* act, cudaStream_t s);
int launch_embed(const __half* table, const int32_t* ids, int T, int H, float* h, cudaStream_t s, bool bf16 = false);
// part: float2 {max, bits(argmax)} [n_tiles, ld] from the kUmmaArgmaxF32 epilogue
int launch_argmax_reduce(const float* part, int n_tiles, int ld, int B, int32_t* out, cudaStream_t s);
int launch_sample(const float* logits, int B, int V, const float* temperature, const float* top_p,
                  const uint32_t* seed, const int32_t* step, int32_t* out, cudaStream_t s);

}  // namespace b200
