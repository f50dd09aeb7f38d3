// The worker engine behind the C ABI: weight store, block-table KV allocator, continuous-batching
// scheduler thread, forward pass.  Stands where the reference's vLLM AsyncLLMEngine stands
// (vgate/backends/vllm_backend.py:61-132): independent requests are submitted from any thread,
// batched per step by the scheduler, and each caller blocks until its own request is finished.
#include <algorithm>
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <cstring>
#include <deque>
#include <dlfcn.h>
#include <functional>
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <tuple>
#include <unordered_map>
#include <vector>

#include "common.cuh"
#include "internal.h"

namespace b200 {

// ---------------------------------------------------------------------------------------------
thread_local std::string g_last_error;
thread_local int64_t g_launch_count = 0;

void set_last_error(const std::string& msg) { g_last_error = msg; }
bool pdl_enabled() {
  static const bool on = [] { const char* v = getenv("VGATE_B200_PDL"); return !(v && v[0] == '0'); }();
  return on;
}
int cuda_fail(cudaError_t err, const char* what, const char* file, int line) {
  g_last_error = std::string("CUDA error ") + cudaGetErrorName(err) + " (" + cudaGetErrorString(err) + ") at " +
                 file + ":" + std::to_string(line) + " in " + what;
  return B200_ERR_CUDA;
}

B200_DEFINE_TL_SETTER(tl_set_engine)
void tl_set_basic(unsigned long long*, unsigned int*, unsigned int);
void tl_set_gemv(unsigned long long*, unsigned int*, unsigned int);
void tl_set_attn(unsigned long long*, unsigned int*, unsigned int);
void tl_set_sample(unsigned long long*, unsigned int*, unsigned int);
void tl_set_umma(unsigned long long*, unsigned int*, unsigned int);

// ---- NCCL, loaded on demand (tensor-parallel engines only: a single-GPU worker has no NCCL dependency) ----------
struct NcclApi {
  int (*get_unique_id)(void* id128) = nullptr;
  int (*all_reduce)(const void*, void*, size_t, int, int, void*, cudaStream_t) = nullptr;
  int (*comm_destroy)(void*) = nullptr;
  const char* (*error_string)(int) = nullptr;
  void* raw_init = nullptr;
  bool ok = false;
};
struct NcclId { char b[128]; };
static NcclApi& nccl_api() {
  static NcclApi api;
  static std::once_flag once;
  std::call_once(once, [] {
    void* h = nullptr;
    for (const char* name : {"libnccl.so.2", "libnccl.so"}) if ((h = dlopen(name, RTLD_NOW | RTLD_GLOBAL))) break;
    if (!h) {                                  // the copy PyTorch ships (this image: site-packages/nvidia/nccl/lib)
      if (const char* env = getenv("VGATE_B200_NCCL_LIB")) h = dlopen(env, RTLD_NOW | RTLD_GLOBAL);
    }
    if (!h) return;
    api.get_unique_id = reinterpret_cast<int (*)(void*)>(dlsym(h, "ncclGetUniqueId"));
    api.raw_init = dlsym(h, "ncclCommInitRank");
    api.all_reduce = reinterpret_cast<int (*)(const void*, void*, size_t, int, int, void*, cudaStream_t)>(dlsym(h, "ncclAllReduce"));
    api.comm_destroy = reinterpret_cast<int (*)(void*)>(dlsym(h, "ncclCommDestroy"));
    api.error_string = reinterpret_cast<const char* (*)(int)>(dlsym(h, "ncclGetErrorString"));
    api.ok = api.get_unique_id && api.raw_init && api.all_reduce && api.comm_destroy && api.error_string;
  });
  return api;
}
static int nccl_fail(int code, const char* what) {
  NcclApi& a = nccl_api();
  set_last_error(std::string("NCCL error in ") + what + ": " + (a.error_string ? a.error_string(code) : "?"));
  return B200_ERR_CUDA;
}

// The library selects the engine's device for the duration of a call and gives the caller's current device back
// (a host that drives several devices -- a tensor-parallel group, PyTorch next to it -- must not find it changed).
struct DeviceGuard {
  int prev = -1;
  explicit DeviceGuard(int dev) { if (cudaGetDevice(&prev) != cudaSuccess) prev = -1; cudaSetDevice(dev); }
  ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
  DeviceGuard(const DeviceGuard&) = delete;
  DeviceGuard& operator=(const DeviceGuard&) = delete;
};

static double now_s() {
  return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

// ---------------------------------------------------------------------------------------------
// Decode-step bookkeeping on the device: consume the token sampled by the previous forward pass.
//   out_tokens[i, n_gen[i]] = next_ids[i]; ids[i] = next_ids[i]; pos[i] += 1; slot/ctx follow.
__global__ void advance_decode_kernel(int B, const int32_t* __restrict__ next_ids, int32_t* __restrict__ ids,
                                      int32_t* __restrict__ pos, int32_t* __restrict__ slot,
                                      int32_t* __restrict__ ctx, int32_t* __restrict__ n_gen,
                                      const int32_t* __restrict__ bt_row, const int32_t* __restrict__ block_tables,
                                      int bt_stride, int32_t* __restrict__ out_tokens, int out_stride,
                                      int32_t* __restrict__ sample_step) {
  const unsigned tl = kernel_prologue(kKidAdvance);
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B) return;
  const int tok = next_ids[i];
  const int g = n_gen[i];
  if (out_tokens && g < out_stride) out_tokens[(size_t)i * out_stride + g] = tok;
  n_gen[i] = g + 1;
  sample_step[i] += 1;                    // index of the token about to be sampled (RNG counter)
  ids[i] = tok;
  const int p = pos[i] + 1;
  pos[i] = p;
  const int blk = p / kBlockTokens;
  slot[i] = blk < bt_stride ? block_tables[(size_t)bt_row[i] * bt_stride + blk] * kBlockTokens + p % kBlockTokens : -1;
  ctx[i] = p + 1;
  kernel_epilogue(tl);
}

// ---------------------------------------------------------------------------------------------
struct DevMem {
  void* p = nullptr;
  size_t bytes = 0;
  int alloc(size_t n) {
    bytes = n;
    B200_CUDA(cudaMalloc(&p, n ? n : 16));
    return B200_OK;
  }
  void release() { if (p) cudaFree(p); p = nullptr; }
  template <typename T> T* as() const { return reinterpret_cast<T*>(p); }
};

struct LayerWeights {
  DevMem wq[4], sc[4], zp[4];      // 0 qkv, 1 o, 2 gate_up, 3 down
  DevMem bias_qkv, ln1, ln2;
  W4Linear lin[4];
  DevMem chunks[4];                // same linears in the chunk layout the tcgen05 GEMM streams
  UmmaW4 ulin[4];
  DevMem dense[4];                 // dense checkpoints: W[N, K] 16-bit (qkv fused row-wise, gate/up paired per 128-row tile)
};

struct PendingLinear {             // raw checkpoint tensors waiting for their siblings
  DevMem qweight, qzeros, scales;
  bool have_w = false, have_z = false, have_s = false;
};

enum class ReqState { WAITING, RUNNING, FINISHED, ABORTED, FAILED };

struct Request {
  int64_t id = 0;
  std::vector<int32_t> prompt;
  float temperature = 0.f, top_p = 1.f;
  int max_tokens = 0;
  uint32_t seed = 0;
  ReqState state = ReqState::WAITING;
  bool abort_requested = false;
  bool detached = false;                       // aborted by its owner: nobody will wait or poll, the scheduler erases it
  int seq_slot = -1;
  std::vector<int32_t> blocks;
  std::vector<int32_t> out;
  int n_prefilled = 0;                         // prompt tokens whose K/V are in the cache (chunked prefill, shared prefix)
  int n_registered = 0;                        // full prompt blocks offered to the prefix cache so far
  uint64_t h1 = 0xCBF29CE484222325ull, h2 = 0x1234567;   // running chained hash of the registered blocks
  bool bt_uploaded = false;                    // its block-table row is on the device (blocks never change after admission)
  size_t n_polled = 0;
  double t_submit = 0, t_first = 0, t_last = 0;
};

}  // namespace b200

using namespace b200;

struct b200_engine {
  b200_config cfg{};
  cudaStream_t stream = nullptr;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  bool finalized = false;
  std::string fatal;                           // non-empty: engine poisoned

  // ---- model ----
  int H = 0, I = 0, L = 0, NH = 0, NKV = 0, V = 0, G = 0, QKV = 0;
  bool bf16 = false;                           // 16-bit storage type (cfg.act_dtype)
  bool dense = false;                          // dense 16-bit linears instead of AWQ int4 (cfg.dense_weights)
  std::vector<LayerWeights> layers;
  DevMem embed, lm_head_w, final_norm;
  std::map<std::string, PendingLinear> pending;
  int tensors_loaded = 0, tensors_expected = 0;
  DevMem cos_t, sin_t;

  // ---- KV pool + block-table allocator ----
  DevMem k_cache, v_cache;                     // [L][n_blocks][NKV][16][128] fp16
  size_t kv_layer_elems = 0;
  int n_blocks = 0, max_blocks_per_seq = 0;
  std::vector<int32_t> free_blocks;            // LIFO free list
  // ---- prefix cache (vLLM's automatic prefix caching, at KV-block granularity) ----
  // A FULL block of prompt tokens whose K/V are in the cache is registered under the chained hash of all prompt
  // tokens up to its end.  A new request whose prompt starts with the same tokens shares those blocks (reference
  // counted) and prefills only the rest; K/V rows depend on the tokens before them and their positions only, and a
  // row's arithmetic never depends on what else is in its pass, so the shared rows hold the very bits the request
  // would have written itself.  Blocks nobody references stay registered and reclaimable (least recently used first).
  bool prefix_cache_enabled = !(getenv("VGATE_B200_PREFIX_CACHE") && atoi(getenv("VGATE_B200_PREFIX_CACHE")) == 0);
  struct BlockMeta { int32_t ref = 0; uint64_t h1 = 0, h2 = 0; int64_t stamp = 0; };   // h1 == 0: not registered
  std::vector<BlockMeta> block_meta;
  std::unordered_map<uint64_t, int32_t> prefix_map;      // h1 -> block id
  std::map<int64_t, int32_t> evictable;                   // stamp -> registered block with ref == 0
  int64_t lru_clock = 0;
  static void hash_block(const int32_t* tok, uint64_t& h1, uint64_t& h2) {   // chained: (h1, h2) enter as the parent
    for (int i = 0; i < kBlockTokens; ++i) {
      const uint64_t v = (uint64_t)(uint32_t)tok[i] + 0x9E3779B97F4A7C15ull;
      h1 = (h1 ^ v) * 0x100000001B3ull;                                      // FNV-1a style
      h2 = (h2 + v) * 0xD6E8FEB86659FD93ull; h2 ^= h2 >> 32;
    }
    if (h1 == 0) h1 = 1;
  }
  int reclaimable_blocks() const { return (int)free_blocks.size() + (int)evictable.size(); }
  void drop_prefix_cache() {                   // generate_batch addresses blocks directly: nothing cached survives it
    for (auto& kv : evictable) { block_meta[kv.second] = BlockMeta{}; free_blocks.push_back(kv.second); }
    evictable.clear();
    prefix_map.clear();
    for (auto& m : block_meta) { m.h1 = m.h2 = 0; }
  }
  std::vector<int32_t> free_slots;

  // ---- activations / metadata ----
  int Tmax = 0;
  DevMem h, xn, qkv, qh, attn, gu, act, xn_last, logits;
  DevMem amax;                                 // greedy lm_head: float2 {tile max, argmax} [V/128][max_seqs]
  bool greedy_only = false;                    // every row of the pending pass is greedy: skip the logits, fuse argmax
  DevMem d_ids, d_pos, d_slot, d_btrow, d_ctx, d_lastrows, d_bt, d_temp, d_topp, d_seed, d_step, d_next, d_ngen;
  DevMem d_out_tokens;
  DevMem d_tiles;                              // prefill query tiles {row0, nrows, bt_row, ctx0}
  std::vector<int32_t> h_tiles;
  int n_q_tiles = 0;                           // tiles describing the rows of the pending prefill pass
  int32_t* hp_i32 = nullptr;                   // pinned staging
  float* hp_f32 = nullptr;
  size_t hp_cap = 0;

  // ---- decode CUDA graphs (generate_batch path) ----
  struct GraphKey {
    int n; const void* out; int stride; bool greedy;
    bool operator<(const GraphKey& o) const { return std::tie(n, out, stride, greedy) < std::tie(o.n, o.out, o.stride, o.greedy); }
  };
  struct GraphVal { cudaGraphExec_t exec; int64_t launches; };
  std::map<GraphKey, GraphVal> decode_graphs;  // one captured decode step (advance + forward) per configuration
  std::map<int, int> decode_warm;              // per batch size and sampling kind: eager steps run so far
  int poll_waiters = 0;                        // threads blocked in b200_poll (guarded by mu)
  std::vector<int64_t> last_decode_ids;        // scheduler: request ids of the previous decode step, in row order

  // ---- tensor parallelism (config #5: one engine object per rank, heads and intermediate columns divided) ----
  // Every rank runs the same kernels on its shard; o_proj and down_proj are row-parallel, so their outputs are
  // partial sums that an NCCL all-reduce completes before the residual add (2 per layer).  The
  // residual stream, the norms and the LM head are replicated, so all ranks sample the same token from the same
  // bits and the device-resident decode state stays consistent without any host exchange.  Rank 0 (the leader) owns
  // the scheduler; the other ranks execute the device half of every step on a thread of their own (NCCL expects
  // one thread per device, or group calls).
  int tp_size = 1, tp_rank = 0;
  void* nccl_comm = nullptr;
  // Decode passes (<= kTpNormMaxRows token rows) do not call NCCL: the RMSNorm that consumes a row-parallel
  // projection adds this rank's split-K partials, pushes the sum into the other ranks' receive buffers over NVLink
  // as self-validating {value, sequence} words, and collects theirs from its own buffer (norm.cuh).  An NCCL
  // all-reduce of 64 x 5120 x 3 fp32 partials took 45 us of a 240 us 14B layer, twice per layer.  Prefill passes
  // (MBs per all-reduce, bandwidth-bound) stay on NCCL.  The 2L sites of a pass alternate between two receive slots.
  bool tp_peer_ok = false;                     // peer access enabled between all ranks' devices, pointers exchanged
  bool tp_use_peer = !(getenv("VGATE_B200_TP_ALLREDUCE") && !strcmp(getenv("VGATE_B200_TP_ALLREDUCE"), "nccl"));
  DevMem tp_epoch, tp_recv;                    // this rank's pass counter; receive words [2 slots][tp_size writers][max_seqs][H] x 8 bytes
  uint2* tp_recv_of[kMaxTp] = {};              // every rank's receive buffer
  int tp_last_site = -1;                       // site of the pending partials layers_pass left for the final norm (-1: NCCL path)
  TpPeers tp_site(int site) const {
    TpPeers t{};
    t.n_ranks = tp_size; t.rank = tp_rank; t.rows_cap = cfg.max_seqs;
    t.site = (uint32_t)site; t.n_sites = (uint32_t)(2 * L + 2);
    for (int r = 0; r < tp_size; ++r) t.recv[r] = tp_recv_of[r] + (size_t)(site & 1) * tp_size * cfg.max_seqs * H;
    return t;
  }
  std::vector<b200_engine*> tp_followers;      // leader only
  std::thread tp_thread;                       // followers only
  std::mutex tp_mu;
  std::condition_variable tp_cv;
  std::function<int()> tp_cmd;
  bool tp_has_cmd = false, tp_done = true, tp_stop = false;
  int tp_rc = 0;
  std::string tp_err;
  void tp_loop() {
    cudaSetDevice(cfg.device);
    std::unique_lock<std::mutex> lk(tp_mu);
    while (true) {
      tp_cv.wait(lk, [&] { return tp_stop || tp_has_cmd; });
      if (tp_stop) break;
      std::function<int()> cmd = std::move(tp_cmd);
      tp_has_cmd = false;
      lk.unlock();
      const int rc = cmd();
      lk.lock();
      tp_rc = rc;
      tp_err = rc ? g_last_error : std::string();
      tp_done = true;
      tp_cv.notify_all();
    }
  }
  void tp_post(std::function<int()> f) {
    std::lock_guard<std::mutex> lk(tp_mu);
    tp_cmd = std::move(f); tp_has_cmd = true; tp_done = false;
    tp_cv.notify_all();
  }
  int tp_wait() {
    std::unique_lock<std::mutex> lk(tp_mu);
    tp_cv.wait(lk, [&] { return tp_done; });
    if (tp_rc) set_last_error("tensor-parallel rank " + std::to_string(tp_rank) + ": " + tp_err);
    return tp_rc;
  }
  // f(rank engine) on every rank: followers on their own threads, the leader inline; returns when all have ENQUEUED.
  template <class F> int tp_each(F&& f) {
    for (b200_engine* r : tp_followers) r->tp_post([r, &f] { return f(r); });
    int rc = f(this);
    for (b200_engine* r : tp_followers) { const int rr = r->tp_wait(); if (!rc) rc = rr; }
    return rc;
  }
  int tp_allreduce(float* buf, size_t count, cudaStream_t st) {
    if (tp_size <= 1) return B200_OK;
    const int rc = nccl_api().all_reduce(buf, buf, count, /*ncclFloat32*/ 7, /*ncclSum*/ 0, nccl_comm, st);
    if (rc != 0) return nccl_fail(rc, "ncclAllReduce");
    g_launch_count += 1;
    return B200_OK;
  }

  // ---- scheduler ----
  std::mutex mu;
  std::condition_variable cv_work, cv_done;
  std::thread worker;
  bool stop = false;
  bool exclusive = false;                      // generate_batch owns the device
  bool stepping = false;                       // the scheduler is inside run_step with mu released (admitted requests
                                               // are in neither `waiting` nor `running` during a prefill pass)
  int64_t next_id = 1;
  std::deque<std::shared_ptr<Request>> waiting;
  std::vector<std::shared_ptr<Request>> prefilling;   // admitted (slot and KV blocks reserved), prompt not fully in the cache yet; FIFO
  std::vector<std::shared_ptr<Request>> running;      // decoding
  // Prompt tokens per prefill pass while sequences are decoding: a pass of the whole max_batch_tokens budget
  // (8192 tokens, ~27 ms) would stall every running sequence for that long on each arrival; with a chunk the
  // stall is one chunk, and decode steps alternate with the chunks (the policy vLLM's scheduler follows for
  // vllm_backend.py:105-132 callers).  With nothing decoding the full budget is used.
  int prefill_chunk = getenv("VGATE_B200_PREFILL_CHUNK") ? std::max(16, atoi(getenv("VGATE_B200_PREFILL_CHUNK"))) : 2048;
  // ... and prefill passes run back to back until they have held the decoding sequences up for this long (GPU time of
  // the passes since the last decode step); then one decode step.  0 alternates strictly.
  double max_stall_ms = getenv("VGATE_B200_MAX_STALL_MS") ? atof(getenv("VGATE_B200_MAX_STALL_MS")) : 25.0;
  double stall_ms = 0.0;                       // scheduler thread only
  bool last_step_was_prefill = false;
  std::unordered_map<int64_t, std::shared_ptr<Request>> table;
  b200_stats stats{};

  // =========================================================================================
  size_t layer_kv_offset(int l) const { return (size_t)l * kv_layer_elems; }

  // One GEMM kernel family per engine, whatever the token count: a sequence's bits must not depend
  // on how many other sequences share its step (continuous batching).  gemm_path 1 selects the
  // CUDA-core GEMV family (small-batch workers), anything else the tcgen05 GEMM.
  bool use_umma(int T) const { (void)T; return dense || cfg.gemm_path != 1; }

  // TMA descriptors for the (fixed) activation buffers, one per (buffer, token-tile height).
  std::map<std::pair<const void*, int>, CUtensorMap> tmaps;
  int get_map(const void* base, int64_t rows, int64_t cols, int box_rows, const CUtensorMap** out) {
    auto key = std::make_pair(base, box_rows);
    auto it = tmaps.find(key);
    if (it == tmaps.end()) {
      CUtensorMap m;
      int rc = make_tensor_map_f16(&m, base, rows, cols, cols, box_rows);
      if (rc) return rc;
      it = tmaps.emplace(key, m).first;
    }
    *out = &it->second;
    return B200_OK;
  }

  // Split-K factors per linear (qkv, o, gate_up, down).  They depend on the weight shape only,
  // never on the token count, so a sequence's arithmetic is the same in every batch it rides in.
  int splits[4] = {1, 1, 1, 1};
  int debug_gemv_mask = getenv("VGATE_B200_DEBUG_GEMV_MASK") ? atoi(getenv("VGATE_B200_DEBUG_GEMV_MASK")) : 0;
  // Timing ablation only (results become garbage): bit 0 rmsnorm, 1 qkv, 2 rope, 3 attention, 4 o,
  // 5 gate_up, 6 down, 7 lm_head, 8 sample -- the kernel is simply not launched.
  int debug_skip = getenv("VGATE_B200_DEBUG_SKIP") ? atoi(getenv("VGATE_B200_DEBUG_SKIP")) : 0;
  bool skip(int bit) const { return (debug_skip >> bit) & 1; }
  DevMem parts;                                  // fp32 split-K partials [splits][T][N]
  DevMem attn_scratch, attn_counters;            // flash-decoding partials + per-(row, kv head) arrival counters
  // KV splits of the decode attention across CTAs.  1: the four warps of a CTA already take different
  // 64-token chunks, and a split count that followed the batch size would make bits batch-dependent.
  static constexpr int kDecodeSplits = 1;

  // y (+)= x . W^T.  Returns in *n_parts_out how many partial tensors were written to `pbuf`
  // instead (0: the result is already in y).
  // split_k: decode passes split K over more CTAs (few token rows, every SM needed); prefill passes have
  // token tiles enough and write the result once (at 12 800 tokens the partials of the down projection
  // alone were 0.9 GB written and read back per layer).  Which of the two a token goes through is decided
  // by its request's state (prompt token or generated token), never by what else is in the batch.
  int linear(const LayerWeights& lw, int i, const __half* x, int64_t x_rows, float* y, int ldy, int T, bool accumulate,
             int* n_parts_out, cudaStream_t st, float* pbuf, bool split_k) {
    const W4Linear& w = lw.lin[i];
    *n_parts_out = 0;
    if (dense) {
      const int nt = umma_pick_nt(T, false);
      const CUtensorMap *mx, *mw;
      int rc;
      if ((rc = get_map(x, x_rows, w.K, nt, &mx)) || (rc = get_map(lw.dense[i].p, w.N, w.K, 128, &mw))) return rc;
      const int ns = split_k ? umma_dense_effective_splits(w.K, splits[i]) : 1;
      if (ns > 1 || (tp_size > 1 && (i == 1 || i == 3))) {     // a row-parallel shard's output is a partial sum by nature
        *n_parts_out = ns;
        return launch_umma_f16(*mw, w.K, w.N, *mx, nt, T, pbuf, w.N, kUmmaPartialF32, st, ns, (int64_t)T * w.N, nullptr, bf16);
      }
      if (i == 2) return launch_umma_f16(*mw, w.K, w.N, *mx, nt, T, y, ldy, kUmmaSwigluF16, st, 1, 0, nullptr, bf16);
      return launch_umma_f16(*mw, w.K, w.N, *mx, nt, T, y, ldy, accumulate ? kUmmaAccumF32 : kUmmaStoreF32, st, 1, 0,
                             accumulate ? nullptr : w.bias, bf16);
    }
    if (!use_umma(T) || (debug_gemv_mask >> i & 1)) return launch_w4_gemv(w, x, w.K, y, ldy, T, accumulate, st);
    const int nt = umma_pick_nt(T, true);
    const CUtensorMap* mx;
    int rc = get_map(x, x_rows, w.K, nt, &mx);
    if (rc) return rc;
    const int ns = split_k ? umma_w4_effective_splits(w.K, w.G, splits[i]) : 1;
    if (ns > 1 || (tp_size > 1 && (i == 1 || i == 3))) {
      *n_parts_out = ns;
      return launch_umma_w4(lw.ulin[i], *mx, nt, T, pbuf, w.N, kUmmaPartialF32, ns, (int64_t)T * w.N, st);
    }
    if (i == 2)    // gate_up: fused SwiGLU epilogue writes fp16 act [T, I] (y is the act buffer)
      return launch_umma_w4(lw.ulin[i], *mx, nt, T, y, ldy, kUmmaSwigluF16, 1, 0, st);
    return launch_umma_w4(lw.ulin[i], *mx, nt, T, y, ldy, accumulate ? kUmmaAccumF32 : kUmmaStoreF32, 1, 0, st);
  }

  // *fused_argmax: the GEMM epilogue reduced each 128-row vocabulary tile to {max, argmax} (amax) instead
  // of writing logits; the caller finishes with launch_argmax_reduce.
  int lm_head(const __half* head, int n_rows, bool want_argmax, bool* fused_argmax) {
    *fused_argmax = false;
    if (!bf16 && (!use_umma(n_rows) || V % 128 != 0 || (debug_gemv_mask >> 4 & 1)))
      return launch_f16_gemv(head, xn_last.as<__half>(), H, logits.as<float>(), V, n_rows, H, V, false, stream);
    const int nt = umma_pick_nt(n_rows, false);
    const CUtensorMap *mx, *mw;
    int rc;
    if ((rc = get_map(xn_last.as<__half>(), cfg.max_seqs, H, nt, &mx)) || (rc = get_map(head, V, H, 128, &mw))) return rc;
    if (want_argmax) {
      *fused_argmax = true;
      return launch_umma_f16(*mw, H, V, *mx, nt, n_rows, amax.as<float>(), cfg.max_seqs, kUmmaArgmaxF32, stream, 1, 0, nullptr, bf16);
    }
    return launch_umma_f16(*mw, H, V, *mx, nt, n_rows, logits.as<float>(), V, kUmmaStoreF32, stream, 1, 0, nullptr, bf16);
  }

  // Embedding + all decoder layers for token rows [r0, r0 + T) on stream `st`; every buffer is indexed
  // by token row, so a lane is the same pass on offset pointers.  Leaves the last down-projection's
  // split-K partials pending (*pend tensors in pbuf) for the caller's final RMSNorm.
  int layers_pass(int r0, int T, bool decode_rows, cudaStream_t st, float* pbuf, int* pend_out) {
    int rc;
    float* hh = h.as<float>() + (size_t)r0 * H;
    __half* xx = xn.as<__half>() + (size_t)r0 * H;
    float* qq = qkv.as<float>() + (size_t)r0 * QKV;
    __half* qhh = qh.as<__half>() + (size_t)r0 * NH * kHeadDim;
    __half* aa = attn.as<__half>() + (size_t)r0 * NH * kHeadDim;
    __half* cc = act.as<__half>() + (size_t)r0 * I;
    const int32_t *pos = d_pos.as<int32_t>() + r0, *slot = d_slot.as<int32_t>() + r0, *btrow = d_btrow.as<int32_t>() + r0,
                  *ctx = d_ctx.as<int32_t>() + r0;
    const int64_t xrows = Tmax - r0;
    // tensor-parallel decode pass: no NCCL, the norms read all ranks' partials (see tp_peer_ok)
    const bool peer = tp_size > 1 && tp_peer_ok && tp_use_peer && decode_rows && r0 == 0 && T <= kTpNormMaxRows && T <= cfg.max_seqs;
    int site = 0, pend_site = -1;
    if (peer && (rc = launch_tp_tick(tp_epoch.as<uint32_t>(), st))) return rc;
    if ((rc = launch_embed(embed.as<__half>(), d_ids.as<int32_t>() + r0, T, H, hh, st, bf16))) return rc;
    int pend = 0;                 // split-K partials of the previous projection still to be added into h
    const int64_t pend_stride = (int64_t)T * H;
    // residual-sum + norm of the pending partials: from pbuf (after the NCCL all-reduce, or single GPU) or from every rank's site buffer
    auto norm = [&](const __half* gamma) -> int {
      if (skip(0)) return B200_OK;
      if (peer && pend > 0)
        return launch_rmsnorm_tp(hh, pbuf, pend, pend_stride, tp_site(pend_site), tp_epoch.as<uint32_t>(), gamma, cfg.rms_eps, nullptr, T, H, xx, st, bf16);
      return launch_rmsnorm(hh, pbuf, pend, pend_stride, gamma, cfg.rms_eps, nullptr, T, H, xx, st, bf16);
    };
    // a row-parallel projection (o, down): partials into this site's buffer; all-reduced by NCCL unless the norm will do it
    auto row_parallel = [&](LayerWeights& w, int i, const __half* x) -> int {
      if ((rc = linear(w, i, x, xrows, hh, H, T, true, &pend, st, pbuf, decode_rows))) return rc;
      if (peer) { pend_site = site++; return B200_OK; }
      return tp_allreduce(pbuf, (size_t)pend * T * H, st);
    };
    for (int l = 0; l < L; ++l) {
      LayerWeights& w = layers[l];
      __half* kc = k_cache.as<__half>() + layer_kv_offset(l);
      __half* vc = v_cache.as<__half>() + layer_kv_offset(l);
      int np = 0;
      if ((rc = norm(w.ln1.as<__half>()))) return rc;
      pend = 0;
      if (!skip(1) && (rc = linear(w, 0, xx, xrows, qq, QKV, T, false, &np, st, pbuf, decode_rows))) return rc;
      const float* qsrc = np ? pbuf : qq;
      const __half* qbias = np ? w.bias_qkv.as<__half>() : nullptr;
      if (decode_rows && T <= cfg.max_seqs) {
        if (!skip(3) && (rc = launch_attn_decode_fused(qsrc, np ? np : 1, (int64_t)T * QKV, qbias, pos, slot, cos_t.as<float>(),
                                                        sin_t.as<float>(), kc, vc, d_bt.as<int32_t>(), max_blocks_per_seq, btrow,
                                                        ctx, T, NH, NKV, kDecodeSplits, attn_scratch.as<float>(),
                                                        attn_counters.as<int32_t>(), aa, st, bf16))) return rc;
      } else {
        if (!skip(2) && (rc = launch_rope_kv(qsrc, np ? np : 1, (int64_t)T * QKV, qbias, pos, slot, cos_t.as<float>(),
                                             sin_t.as<float>(), T, NH, NKV, qhh, kc, vc, st, bf16))) return rc;
        if (n_q_tiles > 0) {       // tiles of <= 16 consecutive query rows share the staged K/V chunks
          if (!skip(3) && (rc = launch_prefill_attn(qhh, kc, vc, d_bt.as<int32_t>(), max_blocks_per_seq,
                                                    d_tiles.as<int32_t>(), n_q_tiles, NH, NKV, aa, st, bf16)))
            return rc;
        } else if (!skip(3) && (rc = launch_paged_attn(qhh, kc, vc, d_bt.as<int32_t>(), max_blocks_per_seq, btrow, ctx, T, NH,
                                                       NKV, 1, nullptr, nullptr, aa, st, bf16))) return rc;
      }
      if (!skip(4) && (rc = row_parallel(w, 1, aa))) return rc;
      if ((rc = norm(w.ln2.as<__half>()))) return rc;
      pend = 0;
      if (use_umma(T) && !(debug_gemv_mask >> 2 & 1)) {
        if (!skip(5) && (rc = linear(w, 2, xx, xrows, reinterpret_cast<float*>(cc), I, T, false, &np, st, pbuf, decode_rows))) return rc;
      } else {
        if ((rc = linear(w, 2, xx, xrows, gu.as<float>() + (size_t)r0 * 2 * I, 2 * I, T, false, &np, st, pbuf, decode_rows))) return rc;
        if ((rc = launch_swiglu(gu.as<float>() + (size_t)r0 * 2 * I, T, I, cc, st))) return rc;
      }
      if (np) { set_last_error("gate_up projection must not be split"); return B200_ERR_STATE; }
      if (!skip(6) && (rc = row_parallel(w, 3, cc))) return rc;
    }
    *pend_out = pend;
    tp_last_site = peer ? pend_site : -1;
    return B200_OK;
  }

  // One forward pass over T token rows whose metadata already sits in the d_* buffers; samples one
  // token for each of the n_rows "last" rows into d_next.  `rows` == nullptr: row r is token r.
  // decode_rows: every row is the single new token of a distinct sequence (enables the fused
  // RoPE + KV-write + split-KV attention kernel).
  int forward(int T, int n_rows, const int32_t* rows, bool decode_rows, bool want_sample = true) {
    int rc;
    {
      int pend = 0;
      if ((rc = layers_pass(0, T, decode_rows, stream, parts.as<float>(), &pend))) return rc;
      if (tp_last_site >= 0 && pend > 0) {          // tensor-parallel decode pass: the final norm completes the last down_proj's sum
        if ((rc = launch_rmsnorm_tp(h.as<float>(), parts.as<float>(), pend, (int64_t)T * H, tp_site(tp_last_site), tp_epoch.as<uint32_t>(),
                                    final_norm.as<__half>(), cfg.rms_eps, rows, n_rows, H, xn_last.as<__half>(), stream, bf16))) return rc;
      } else if ((rc = launch_rmsnorm(h.as<float>(), parts.as<float>(), pend, (int64_t)T * H, final_norm.as<__half>(), cfg.rms_eps,
                                      rows, n_rows, H, xn_last.as<__half>(), stream, bf16))) return rc;
    }
    const __half* head = cfg.tie_embeddings ? embed.as<__half>() : lm_head_w.as<__half>();
    bool fused = false;
    if (!skip(7) && (rc = lm_head(head, n_rows, want_sample && greedy_only, &fused))) return rc;
    if (want_sample && !skip(8)) {
      if (fused) {
        if ((rc = launch_argmax_reduce(amax.as<float>(), V / 128, cfg.max_seqs, n_rows, d_next.as<int32_t>(), stream))) return rc;
      } else if ((rc = launch_sample(logits.as<float>(), n_rows, V, d_temp.as<float>(), d_topp.as<float>(),
                                     d_seed.as<uint32_t>(), d_step.as<int32_t>(), d_next.as<int32_t>(), stream))) return rc;
    }
    return B200_OK;
  }

  double decode_step_bytes(int B, int64_t ctx_sum) const {
    // algorithmic HBM bytes of one decode step: weights once + KV read + KV write
    double w = 0;
    for (int i = 0; i < 4; ++i) {
      const W4Linear& lw = layers[0].lin[i];
      w += dense ? (double)lw.N * lw.K * 2 : (double)lw.N * lw.K / 2 + 2.0 * lw.N * (lw.K / lw.G) * 2;
    }
    w = w * L + (double)V * H * 2;                 // + the LM head (tied or not, it is read once per step)
    const double kv_tok = 2.0 * L * NKV * kHeadDim * 2;
    return w + kv_tok * (double)ctx_sum + kv_tok * B;
  }

  // ---- block allocator ---------------------------------------------------------------------
  bool alloc_blocks(int n, std::vector<int32_t>& out) {
    if (reclaimable_blocks() < n) return false;
    for (int i = 0; i < n; ++i) {
      int32_t b;
      if (!free_blocks.empty()) { b = free_blocks.back(); free_blocks.pop_back(); }
      else {                                   // evict the least recently used registered block
        auto it = evictable.begin();
        b = it->second;
        evictable.erase(it);
        prefix_map.erase(block_meta[b].h1);
      }
      block_meta[b] = BlockMeta{};
      block_meta[b].ref = 1;
      out.push_back(b);
    }
    return true;
  }
  // Longest run of registered blocks that matches the start of `prompt`; they are pinned (ref + 1) and appended to
  // `out`.  At least one prompt token is always left to compute (its logits sample the first generated token).
  int share_prefix(const std::vector<int32_t>& prompt, std::vector<int32_t>& out) {
    if (!prefix_cache_enabled) return 0;
    const int nb_full = (int)prompt.size() / kBlockTokens;
    uint64_t h1 = 0xCBF29CE484222325ull, h2 = 0x1234567;
    int hits = 0;
    for (int i = 0; i < nb_full && (i + 1) * kBlockTokens < (int)prompt.size(); ++i) {
      hash_block(prompt.data() + (size_t)i * kBlockTokens, h1, h2);
      auto it = prefix_map.find(h1);
      if (it == prefix_map.end() || block_meta[it->second].h2 != h2) break;
      BlockMeta& m = block_meta[it->second];
      if (m.ref == 0) evictable.erase(m.stamp);
      m.ref++;
      out.push_back(it->second);
      ++hits;
    }
    return hits;
  }
  // Register the full prompt blocks of `r` whose K/V are complete (positions < r.n_prefilled) and not yet registered.
  void register_prefix(Request& r) {
    if (!prefix_cache_enabled) return;
    const int nb_done = std::min((int)r.prompt.size(), r.n_prefilled) / kBlockTokens;
    for (; r.n_registered < nb_done; ++r.n_registered) {
      hash_block(r.prompt.data() + (size_t)r.n_registered * kBlockTokens, r.h1, r.h2);
      const int32_t b = r.blocks[r.n_registered];
      if (block_meta[b].h1 != 0) continue;                      // a shared block: registered by its first owner
      if (prefix_map.count(r.h1)) continue;                     // the same prefix was registered meanwhile: ours stays private
      block_meta[b].h1 = r.h1; block_meta[b].h2 = r.h2;
      prefix_map[r.h1] = b;
    }
  }
  void release_request_resources(Request& r) {
    for (int32_t b : r.blocks) {
      BlockMeta& m = block_meta[b];
      if (--m.ref > 0) continue;
      if (m.h1 != 0) { m.stamp = ++lru_clock; evictable[m.stamp] = b; }
      else free_blocks.push_back(b);
    }
    r.blocks.clear();
    if (r.seq_slot >= 0) free_slots.push_back(r.seq_slot);
    r.seq_slot = -1;
  }

  // Group the rows of a prefill pass into query tiles: consecutive rows of one sequence at consecutive
  // positions, at most 16 per tile.
  int upload_q_tiles(const std::vector<int32_t>& btrow, const std::vector<int32_t>& ctx) {
    std::vector<int32_t>& tiles = h_tiles;     // member: must outlive the asynchronous copy
    tiles.clear();
    const int T = (int)btrow.size();
    for (int r = 0; r < T;) {
      int n = 1;
      while (n < 16 && r + n < T && btrow[r + n] == btrow[r] && ctx[r + n] == ctx[r] + n) ++n;
      tiles.insert(tiles.end(), {r, n, btrow[r], ctx[r]});
      r += n;
    }
    n_q_tiles = (int)tiles.size() / 4;
    return upload_i32(d_tiles, tiles);
  }

  int upload_i32(DevMem& dst, const std::vector<int32_t>& src) {
    if (src.empty()) return B200_OK;
    B200_CUDA(cudaMemcpyAsync(dst.p, src.data(), src.size() * 4, cudaMemcpyHostToDevice, stream));
    return B200_OK;
  }
  int upload_f32(DevMem& dst, const std::vector<float>& src) {
    if (src.empty()) return B200_OK;
    B200_CUDA(cudaMemcpyAsync(dst.p, src.data(), src.size() * 4, cudaMemcpyHostToDevice, stream));
    return B200_OK;
  }

  // ---- one scheduler step over `batch`; prefill rows for new requests, one row for running ones.
  // One decode step for the n sequences whose state sits in the d_* buffers: the advance kernel
  // consumes the previously sampled token (d_next), then the forward pass samples the next one.
  // The first step of a batch size runs eagerly (loads every kernel), later ones replay a CUDA graph.
  int decode_step(int n, int32_t* out_tokens, int out_stride) {
    auto eager = [&]() -> int {
      B200_CUDA(launch_k(advance_decode_kernel, dim3((n + 127) / 128), dim3(128), 0, stream, n, d_next.as<int32_t>(),
                         d_ids.as<int32_t>(), d_pos.as<int32_t>(), d_slot.as<int32_t>(), d_ctx.as<int32_t>(),
                         d_ngen.as<int32_t>(), d_btrow.as<int32_t>(), d_bt.as<int32_t>(), max_blocks_per_seq,
                         out_tokens, out_stride, d_step.as<int32_t>()));
      g_launch_count += 1;
      return forward(n, n, nullptr, true);
    };
    // A configuration's first two steps run eagerly, the third is captured: while a batch ramps up or drains, every
    // step has another size, and capturing + instantiating a graph (milliseconds for hundreds of nodes) for a size
    // that is used once costs more than it saves.
    if (!cfg.use_cuda_graphs || decode_warm[2 * n + (greedy_only ? 1 : 0)] < 2) {
      decode_warm[2 * n + (greedy_only ? 1 : 0)]++;
      return eager();
    }
    const GraphKey key{n, out_tokens, out_stride, greedy_only};
    auto it = decode_graphs.find(key);
    if (it == decode_graphs.end()) {
      if (decode_graphs.size() >= 64) {               // bound the cache
        for (auto& g : decode_graphs) cudaGraphExecDestroy(g.second.exec);
        decode_graphs.clear();
      }
      cudaGraph_t g = nullptr;
      B200_CUDA(cudaStreamBeginCapture(stream, cudaStreamCaptureModeThreadLocal));
      const int64_t before = g_launch_count;
      const int rc = eager();
      const int64_t per_step = g_launch_count - before;
      g_launch_count = before;
      const cudaError_t ce = cudaStreamEndCapture(stream, &g);
      if (rc != B200_OK) { if (g) cudaGraphDestroy(g); return rc; }
      B200_CUDA(ce);
      cudaGraphExec_t exec = nullptr;
      B200_CUDA(cudaGraphInstantiate(&exec, g, 0));
      B200_CUDA(cudaGraphDestroy(g));
      it = decode_graphs.emplace(key, GraphVal{exec, per_step}).first;
    }
    B200_CUDA(cudaGraphLaunch(it->second.exec, stream));
    g_launch_count += it->second.launches;
    return B200_OK;
  }

  // ---- scheduler decode: one step of look-ahead -------------------------------------------------
  // A decode step needs nothing from the host when the batch is unchanged (the device holds the state
  // and the sampled tokens), so step s + 1 is launched BEFORE the tokens of step s are read back: the
  // host's per-step work (event wait, copy, bookkeeping, waking waiters) overlaps GPU work instead of
  // separating two steps.  Look-ahead is only taken when no sequence of the batch can reach max_tokens
  // at step s, nothing waits for admission and no abort is pending.  A sequence that stops at step s on
  // EOS has one more token computed at s + 1 (inside its reserved KV blocks); it is dropped at collection.  A step that was launched ahead is always collected (its tokens belong to sequences that
  // did advance) before anything else touches the device state.
  struct Inflight {
    bool active = false;
    std::vector<std::shared_ptr<Request>> batch;
    std::vector<int64_t> ids;
    int slot = 0;
    int64_t ctx_sum = 0, launches = 0;
  };
  Inflight inflight;
  int32_t* hp_next[2] = {nullptr, nullptr};        // pinned, max_seqs tokens each
  int32_t* hp_up[2] = {nullptr, nullptr};          // pinned, 7 x max_seqs words: decode state upload when the batch changes
  cudaEvent_t ev_up[2] = {nullptr, nullptr};
  int up_slot = 0;
  cudaEvent_t ev_a[2] = {nullptr, nullptr}, ev_b[2] = {nullptr, nullptr}, ev_copied[2] = {nullptr, nullptr};
  bool lookahead_enabled = !(getenv("VGATE_B200_LOOKAHEAD") && atoi(getenv("VGATE_B200_LOOKAHEAD")) == 0);

  // The device half of one decode step, executed by every tensor-parallel rank: optional state upload, the step
  // (CUDA-graph replay), and the copy of the sampled tokens into pinned slot `slot`.
  int decode_device(int n, const int32_t* state, bool greedy, int slot) {
    int rc;
    greedy_only = greedy;
    if (state) {
      const int k = (up_slot ^= 1);
      B200_CUDA(cudaEventSynchronize(ev_up[k]));       // the copies that last used this buffer (long done)
      const size_t S = (size_t)cfg.max_seqs;
      int32_t* base = hp_up[k];
      std::memcpy(base, state, 7 * S * 4);
      auto up = [&](DevMem& dst, const void* src) { return cudaMemcpyAsync(dst.p, src, (size_t)n * 4, cudaMemcpyHostToDevice, stream); };
      B200_CUDA(up(d_next, base)); B200_CUDA(up(d_pos, base + S)); B200_CUDA(up(d_btrow, base + 2 * S)); B200_CUDA(up(d_step, base + 3 * S));
      B200_CUDA(up(d_seed, base + 4 * S)); B200_CUDA(up(d_temp, base + 5 * S)); B200_CUDA(up(d_topp, base + 6 * S));
      B200_CUDA(cudaEventRecord(ev_up[k], stream));
    }
    B200_CUDA(cudaEventRecord(ev_a[slot], stream));
    if ((rc = decode_step(n, nullptr, 0))) return rc;
    B200_CUDA(cudaEventRecord(ev_b[slot], stream));
    B200_CUDA(cudaMemcpyAsync(hp_next[slot], d_next.p, (size_t)n * 4, cudaMemcpyDeviceToHost, stream));
    B200_CUDA(cudaEventRecord(ev_copied[slot], stream));
    return B200_OK;
  }

  // Launch one decode step for `batch` (state upload if the batch changed) and queue the copy of its tokens.
  int launch_decode(const std::vector<std::shared_ptr<Request>>& batch, const std::vector<int64_t>& idsig, int64_t extra_ctx) {
    const int n = (int)batch.size();
    int rc;
    int64_t ctx_sum = extra_ctx;
    greedy_only = true;
    for (int r = 0; r < n; ++r) {
      ctx_sum += (int64_t)batch[r]->prompt.size() + (int64_t)batch[r]->out.size();
      greedy_only = greedy_only && batch[r]->temperature < 1e-5f;
    }
    std::vector<int32_t> state;                      // non-empty: the batch changed, the device state is re-uploaded
    if (idsig != last_decode_ids) {
      // state BEFORE the advance kernel: last token in d_next, position one behind.  Seven small per-row arrays
      // go through a pinned double buffer (no stream synchronisation: the decode pipeline keeps running while
      // the host moves on); block-table rows were uploaded when their sequences were admitted.
      const size_t S = (size_t)cfg.max_seqs;
      state.assign(7 * S, 0);
      int32_t* base = state.data();
      int32_t *next = base, *pos = base + S, *btrow = base + 2 * S, *step = base + 3 * S, *seed = base + 4 * S;
      float *temp = reinterpret_cast<float*>(base + 5 * S), *topp = reinterpret_cast<float*>(base + 6 * S);
      for (int r = 0; r < n; ++r) {
        Request& q = *batch[r];
        const int n_ctx = (int)q.prompt.size() + (int)q.out.size();
        next[r] = q.out.back();
        pos[r] = n_ctx - 2;
        btrow[r] = q.seq_slot;
        step[r] = (int32_t)q.out.size() - 1;
        seed[r] = (int32_t)q.seed;
        temp[r] = q.temperature;
        topp[r] = q.top_p;
      }
      last_decode_ids = idsig;
    }
    const int slot = inflight.slot ^ 1;
    const int64_t launches0 = g_launch_count;
    const bool greedy = greedy_only;
    if ((rc = tp_each([&](b200_engine* e) { return e->decode_device(n, state.empty() ? nullptr : state.data(), greedy, slot); }))) return rc;
    inflight.active = true; inflight.batch = batch; inflight.ids = idsig; inflight.slot = slot;
    inflight.ctx_sum = ctx_sum; inflight.launches = g_launch_count - launches0;
    return B200_OK;
  }

  // Wait for the step in flight and hand its tokens to the sequences that are still running.
  int collect_decode() {
    if (!inflight.active) return B200_OK;
    Inflight f = std::move(inflight);
    inflight = Inflight{};
    inflight.slot = f.slot;
    B200_CUDA(cudaEventSynchronize(ev_copied[f.slot]));
    float ms = 0.f;
    B200_CUDA(cudaEventElapsedTime(&ms, ev_a[f.slot], ev_b[f.slot]));
    const int n = (int)f.batch.size();
    const double t = now_s();
    std::lock_guard<std::mutex> lk(mu);
    stats.kernel_launches += f.launches;
    stats.steps_decode++; stats.gpu_seconds_decode += ms * 1e-3; stats.decode_batch_sum += n;
    stats.hbm_bytes_decode += decode_step_bytes(n, f.ctx_sum);
    for (int r = 0; r < n; ++r) {
      Request& q = *f.batch[r];
      if (q.state != ReqState::RUNNING) continue;   // aborted while the step was in flight: its token is dropped
      q.out.push_back(hp_next[f.slot][r]);
      q.t_last = t;
      stats.tokens_generated++;
    }
    return B200_OK;
  }

  int run_decode(const std::vector<std::shared_ptr<Request>>& batch, bool allow_ahead) {
    const int n = (int)batch.size();
    std::vector<int64_t> idsig(n);
    for (int r = 0; r < n; ++r) idsig[r] = batch[r]->id;
    int rc;
    if (inflight.active && inflight.ids != idsig && (rc = collect_decode())) return rc;   // a different batch was ahead
    if (!inflight.active && (rc = launch_decode(batch, idsig, 0))) return rc;
    // look ahead: may step s + 1 start before the tokens of step s are known?
    bool ahead = lookahead_enabled && allow_ahead;
    for (int r = 0; r < n && ahead; ++r) ahead = (int)batch[r]->out.size() + 1 < batch[r]->max_tokens;
    if (ahead) {
      std::lock_guard<std::mutex> lk(mu);
      ahead = waiting.empty() && !stop;
      for (int r = 0; r < n && ahead; ++r) ahead = !batch[r]->abort_requested;
    }
    if (!ahead) return collect_decode();
    // step s is in flight in slot k; launch s + 1 into the other slot, then collect s
    Inflight cur = std::move(inflight);
    inflight = Inflight{};
    inflight.slot = cur.slot;
    if ((rc = launch_decode(batch, idsig, n))) return rc;          // context one token longer for the roofline accounting
    Inflight nxt = std::move(inflight);
    inflight = std::move(cur);
    if ((rc = collect_decode())) return rc;
    inflight = std::move(nxt);
    return B200_OK;
  }

  // One prefill pass: for every chunk the prompt tokens [start, start + count) of its request, K/V written to the
  // paged cache, attention over everything cached so far (earlier chunks included).  Requests whose prompt ends
  // in this pass get their first token sampled.  A row's arithmetic depends on its own position and context
  // only, so a prompt's results do not depend on how it was cut into chunks.
  struct Chunk { std::shared_ptr<Request> req; int start, count; };
  int run_prefill(const std::vector<Chunk>& chunks) {
    { int rc0 = collect_decode(); if (rc0) return rc0; }   // a step launched ahead belongs to the running sequences
    last_decode_ids.clear();                         // a prefill pass overwrites the d_* state
    std::vector<int32_t> ids, pos, slot, btrow, ctx, lastrows, step, seed;
    std::vector<float> temp, topp;
    std::vector<Request*> done;                      // requests whose prompt completes in this pass
    int rc;
    greedy_only = true;
    std::vector<Request*> new_rows;                  // sequences whose block-table row is not on the device(s) yet
    for (const Chunk& c : chunks) {
      Request& q = *c.req;
      if (!q.bt_uploaded) {                          // once per sequence: its blocks are reserved for its whole life
        new_rows.push_back(&q);
        q.bt_uploaded = true;
      }
      for (int p = c.start; p < c.start + c.count; ++p) {
        ids.push_back(q.prompt[p]);
        pos.push_back(p);
        slot.push_back(q.blocks[p / kBlockTokens] * kBlockTokens + p % kBlockTokens);
        btrow.push_back(q.seq_slot);
        ctx.push_back(p + 1);
      }
      if (c.start + c.count == (int)q.prompt.size()) {
        done.push_back(&q);
        lastrows.push_back((int)ids.size() - 1);
        greedy_only = greedy_only && q.temperature < 1e-5f;
        temp.push_back(q.temperature);
        topp.push_back(q.top_p);
        seed.push_back((int32_t)q.seed);
        step.push_back(0);
      }
    }
    const int T = (int)ids.size(), n = (int)done.size();
    const int64_t launches0 = g_launch_count;
    const bool greedy = greedy_only;
    std::vector<int32_t> next(n);
    // the device half, on every tensor-parallel rank (each ends with a stream synchronisation: the host vectors
    // above stay alive until then); the leader also reads the sampled tokens back
    rc = tp_each([&](b200_engine* e) -> int {
      int r;
      e->greedy_only = greedy;
      for (Request* q : new_rows)
        B200_CUDA(cudaMemcpyAsync(e->d_bt.as<int32_t>() + (size_t)q->seq_slot * e->max_blocks_per_seq, q->blocks.data(),
                                  q->blocks.size() * 4, cudaMemcpyHostToDevice, e->stream));
      if ((r = e->upload_i32(e->d_ids, ids)) || (r = e->upload_i32(e->d_pos, pos)) || (r = e->upload_i32(e->d_slot, slot)) ||
          (r = e->upload_i32(e->d_btrow, btrow)) || (r = e->upload_i32(e->d_ctx, ctx)) || (r = e->upload_i32(e->d_lastrows, lastrows)) ||
          (r = e->upload_q_tiles(btrow, ctx)) || (r = e->upload_f32(e->d_temp, temp)) || (r = e->upload_f32(e->d_topp, topp)) ||
          (r = e->upload_i32(e->d_seed, seed)) || (r = e->upload_i32(e->d_step, step)))
        return r;
      B200_CUDA(cudaEventRecord(e->ev0, e->stream));
      if ((r = e->forward(T, n, e->d_lastrows.as<int32_t>(), false, n > 0))) return r;
      B200_CUDA(cudaEventRecord(e->ev1, e->stream));
      if (e == this && n > 0) B200_CUDA(cudaMemcpyAsync(next.data(), e->d_next.p, (size_t)n * 4, cudaMemcpyDeviceToHost, e->stream));
      B200_CUDA(cudaStreamSynchronize(e->stream));
      return B200_OK;
    });
    if (rc) return rc;
    float ms = 0.f;
    B200_CUDA(cudaEventElapsedTime(&ms, ev0, ev1));
    const double t = now_s();
    std::lock_guard<std::mutex> lk(mu);
    stats.kernel_launches += g_launch_count - launches0;
    stats.steps_prefill++; stats.tokens_prefill += T; stats.gpu_seconds_prefill += ms * 1e-3;
    stats.tokens_generated += n;
    stall_ms += ms;
    for (const Chunk& c : chunks) c.req->n_prefilled = c.start + c.count;
    for (int r = 0; r < n; ++r) {
      Request& q = *done[r];
      q.out.push_back(next[r]);
      q.t_first = t;
      q.t_last = t;
    }
    return B200_OK;
  }

  void fail_all(const std::string& why) {       // mu held
    fatal = why;
    for (auto& kv : table) {
      Request& r = *kv.second;
      if (r.state == ReqState::WAITING || r.state == ReqState::RUNNING) r.state = ReqState::FAILED;
    }
    waiting.clear();
    prefilling.clear();
    running.clear();
    cv_done.notify_all();
  }

  void scheduler_loop() {
    cudaSetDevice(cfg.device);
    std::unique_lock<std::mutex> lk(mu);
    while (true) {
      cv_work.wait(lk, [&] { return stop || (!exclusive && (!waiting.empty() || !prefilling.empty() || !running.empty())); });
      if (stop) break;
      // retire aborted requests
      for (auto* list : {&running, &prefilling}) {
        for (auto it = list->begin(); it != list->end();) {
          if ((*it)->abort_requested) {
            (*it)->state = ReqState::ABORTED; release_request_resources(**it); stats.requests_aborted++;
            if ((*it)->detached) table.erase((*it)->id);
            it = list->erase(it);
          } else ++it;
        }
      }
      for (auto it = waiting.begin(); it != waiting.end();) {
        if ((*it)->abort_requested) {
          (*it)->state = ReqState::ABORTED; stats.requests_aborted++;
          if ((*it)->detached) table.erase((*it)->id);
          it = waiting.erase(it);
        } else ++it;
      }
      cv_done.notify_all();
      // admit: FIFO, a sequence slot and KV blocks reserved for the full generation, so that an admitted
      // sequence never has to be preempted.
      while (!waiting.empty() && !free_slots.empty()) {
        auto& r = waiting.front();
        const int need_tokens = (int)r->prompt.size() + r->max_tokens;
        const int need_blocks = (need_tokens + kBlockTokens - 1) / kBlockTokens;
        const int shared = share_prefix(r->prompt, r->blocks);        // pins the blocks it finds
        if (reclaimable_blocks() < need_blocks - shared) {            // does not fit yet: give the pins back, keep waiting
          release_request_resources(*r);
          break;
        }
        alloc_blocks(need_blocks - shared, r->blocks);
        r->n_prefilled = shared * kBlockTokens;
        r->n_registered = shared;
        for (int i = 0; i < shared; ++i) hash_block(r->prompt.data() + (size_t)i * kBlockTokens, r->h1, r->h2);
        stats.prefix_hit_tokens += r->n_prefilled;
        stats.prefix_query_tokens += (int64_t)r->prompt.size();
        r->seq_slot = free_slots.back(); free_slots.pop_back();
        r->state = ReqState::RUNNING;
        prefilling.push_back(r);
        waiting.pop_front();
      }
      // choose the step: prompt chunks and decode steps alternate while both kinds of work exist
      const bool do_prefill = !prefilling.empty() && (running.empty() || !last_step_was_prefill || stall_ms < max_stall_ms);
      std::vector<Chunk> chunks;
      std::vector<std::shared_ptr<Request>> batch;
      if (do_prefill) {
        int budget = running.empty() ? cfg.max_batch_tokens : std::min(cfg.max_batch_tokens, prefill_chunk);
        for (auto& r : prefilling) {
          if (budget <= 0) break;
          const int take = std::min(budget, (int)r->prompt.size() - r->n_prefilled);
          chunks.push_back(Chunk{r, r->n_prefilled, take});
          budget -= take;
        }
      } else {
        batch = running;
      }
      if (chunks.empty() && batch.empty()) continue;
      const bool allow_ahead = prefilling.empty();   // with prompt chunks pending the next step is not a decode step
      stepping = true;
      lk.unlock();
      const int rc = do_prefill ? run_prefill(chunks) : run_decode(batch, allow_ahead);
      lk.lock();
      stepping = false;
      last_step_was_prefill = do_prefill;
      if (!do_prefill) stall_ms = 0.0;
      if (rc != B200_OK) { inflight = Inflight{}; fail_all(g_last_error); continue; }
      if (do_prefill) {                              // sequences whose prompt is in the cache start decoding
        for (auto& c : chunks) register_prefix(*c.req);
        for (auto it = prefilling.begin(); it != prefilling.end();) {
          if ((*it)->n_prefilled == (int)(*it)->prompt.size()) { running.push_back(*it); it = prefilling.erase(it); }
          else ++it;
        }
      }
      // retire finished sequences
      bool any_done = false;
      for (auto it = running.begin(); it != running.end();) {
        Request& r = **it;
        const bool eos = cfg.eos_token_id >= 0 && !r.out.empty() && r.out.back() == cfg.eos_token_id;
        if ((int)r.out.size() >= r.max_tokens || eos) {
          r.state = ReqState::FINISHED; release_request_resources(r); stats.requests_finished++;
          it = running.erase(it);
          any_done = true;
        } else ++it;
      }
      // waking every blocked b200_wait caller on every step costs more than the step's bookkeeping;
      // only finished requests and streaming pollers need to hear about it
      if (any_done || poll_waiters > 0) cv_done.notify_all();
    }
  }
};

// =============================================================================================
// C ABI
// =============================================================================================
extern "C" {

int32_t b200_abi_version(void) { return B200_ABI_VERSION; }
const char* b200_last_error(void) { return g_last_error.c_str(); }

int32_t b200_engine_create(const b200_config* cfg, b200_engine** out) {
  B200_REQUIRE(cfg && out, "null argument");
  B200_REQUIRE(cfg->struct_size == (int32_t)sizeof(b200_config), "b200_config.struct_size mismatch");
  B200_REQUIRE(cfg->head_dim == kHeadDim, "head_dim must be 128");
  B200_REQUIRE(cfg->hidden > 0 && cfg->hidden % 128 == 0 && cfg->intermediate % 128 == 0, "hidden/intermediate % 128");
  B200_REQUIRE(cfg->tp_size >= 0 && cfg->tp_size <= 8 && cfg->tp_rank >= 0 && cfg->tp_rank < std::max(1, cfg->tp_size), "0 <= tp_rank < tp_size <= 8");
  B200_REQUIRE(cfg->dense_weights == 0 || cfg->dense_weights == 1, "dense_weights in {0, 1}");
  B200_REQUIRE(cfg->act_dtype == B200_DTYPE_F16 || cfg->act_dtype == B200_DTYPE_BF16, "act_dtype is B200_DTYPE_F16 or B200_DTYPE_BF16");
  B200_REQUIRE(cfg->dense_weights == 1 || cfg->act_dtype == B200_DTYPE_F16, "AWQ int4 linears run with fp16 activations");
  B200_REQUIRE(cfg->dense_weights == 1 || cfg->group_size == 128 || cfg->group_size == 64 || cfg->group_size == 32, "group_size in {32,64,128}");
  B200_REQUIRE(cfg->n_heads % cfg->n_kv_heads == 0 && cfg->n_heads / cfg->n_kv_heads <= 8, "GQA group <= 8");
  B200_REQUIRE(cfg->max_seqs > 0 && cfg->max_model_len > 0 && cfg->max_batch_tokens > 0, "max_seqs/max_model_len/max_batch_tokens");
  int ndev = 0;
  B200_CUDA(cudaGetDeviceCount(&ndev));
  B200_REQUIRE(cfg->device >= 0 && cfg->device < ndev, "device ordinal out of range");
  DeviceGuard guard(cfg->device);
  cudaDeviceProp prop;
  B200_CUDA(cudaGetDeviceProperties(&prop, cfg->device));
  if (prop.major != 10) {
    set_last_error(std::string("libb200gate needs an sm_100 GPU (B200); found ") + prop.name + " sm_" +
                   std::to_string(prop.major) + std::to_string(prop.minor));
    return B200_ERR_INVALID;
  }
  auto* e = new b200_engine();
  e->cfg = *cfg;
  e->H = cfg->hidden; e->I = cfg->intermediate; e->L = cfg->n_layers; e->NH = cfg->n_heads; e->NKV = cfg->n_kv_heads;
  e->V = cfg->vocab; e->G = cfg->dense_weights ? 128 : cfg->group_size; e->QKV = (cfg->n_heads + 2 * cfg->n_kv_heads) * kHeadDim;
  e->bf16 = cfg->act_dtype == B200_DTYPE_BF16; e->dense = cfg->dense_weights == 1;
  e->tp_size = std::max(1, cfg->tp_size); e->tp_rank = cfg->tp_rank;
  int rc = B200_OK;
  auto fail = [&](int code) { b200_engine_destroy(e); return code; };
  if (cudaStreamCreateWithFlags(&e->stream, cudaStreamNonBlocking) != cudaSuccess ||
      cudaEventCreate(&e->ev0) != cudaSuccess || cudaEventCreate(&e->ev1) != cudaSuccess ||
      cudaEventCreate(&e->ev_a[0]) != cudaSuccess || cudaEventCreate(&e->ev_a[1]) != cudaSuccess ||
      cudaEventCreate(&e->ev_b[0]) != cudaSuccess || cudaEventCreate(&e->ev_b[1]) != cudaSuccess ||
      cudaEventCreateWithFlags(&e->ev_copied[0], cudaEventDisableTiming) != cudaSuccess ||
      cudaEventCreateWithFlags(&e->ev_copied[1], cudaEventDisableTiming) != cudaSuccess ||
      cudaEventCreateWithFlags(&e->ev_up[0], cudaEventDisableTiming) != cudaSuccess ||
      cudaEventCreateWithFlags(&e->ev_up[1], cudaEventDisableTiming) != cudaSuccess ||
      cudaMallocHost(&e->hp_up[0], (size_t)cfg->max_seqs * 7 * 4) != cudaSuccess ||
      cudaMallocHost(&e->hp_up[1], (size_t)cfg->max_seqs * 7 * 4) != cudaSuccess ||
      cudaMallocHost(&e->hp_next[0], (size_t)cfg->max_seqs * 4) != cudaSuccess ||
      cudaMallocHost(&e->hp_next[1], (size_t)cfg->max_seqs * 4) != cudaSuccess) {
    set_last_error("cudaStreamCreate/cudaEventCreate failed");
    return fail(B200_ERR_CUDA);
  }
  e->layers.resize(e->L);
  const int Ks[4] = {e->H, e->NH * kHeadDim, e->H, e->I};
  const int Ns[4] = {e->QKV, e->H, 2 * e->I, e->H};
  for (int l = 0; l < e->L && rc == B200_OK; ++l) {
    LayerWeights& w = e->layers[l];
    for (int i = 0; i < 4 && rc == B200_OK; ++i) {
      const size_t K = Ks[i], N = Ns[i];
      w.lin[i].K = (int)K; w.lin[i].N = (int)N; w.lin[i].G = e->G;
      if (e->dense) {
        if ((rc = w.dense[i].alloc(N * K * 2))) break;
        continue;
      }
      if ((rc = w.wq[i].alloc(N * (K / 8) * 4)) || (rc = w.sc[i].alloc(N * (K / e->G) * 2)) ||
          (rc = w.zp[i].alloc(N * (K / e->G) * 2))) break;
      w.lin[i].wq = w.wq[i].as<uint32_t>(); w.lin[i].sc = w.sc[i].as<__half>(); w.lin[i].zp = w.zp[i].as<__half>();
    }
    if (rc == B200_OK && !((rc = w.bias_qkv.alloc((size_t)e->QKV * 2)) || (rc = w.ln1.alloc((size_t)e->H * 2)) ||
                           (rc = w.ln2.alloc((size_t)e->H * 2))))
      w.lin[0].bias = w.bias_qkv.as<__half>();
  }
  if (rc == B200_OK) rc = e->embed.alloc((size_t)e->V * e->H * 2);
  if (rc == B200_OK && !cfg->tie_embeddings) rc = e->lm_head_w.alloc((size_t)e->V * e->H * 2);
  if (rc == B200_OK) rc = e->final_norm.alloc((size_t)e->H * 2);
  if (rc != B200_OK) return fail(rc);
  // per layer: 7 linears x 3 tensors + 3 biases + 2 norms; + embed + final norm (+ lm_head)
  e->tensors_expected = e->L * (7 * (e->dense ? 1 : 3) + 3 + 2) + 2 + (cfg->tie_embeddings ? 0 : 1);
  if (e->dense) {
    B200_REQUIRE(e->H % 128 == 0 && e->QKV % 128 == 0 && e->I % 64 == 0 && e->V % 128 == 0,
                 "dense linears: hidden, qkv width and vocab % 128 == 0, intermediate % 64 == 0");
  }
  *out = e;
  return B200_OK;
}

static int copy_to(DevMem& dst, size_t dst_off, const void* src, size_t bytes, cudaStream_t s) {
  if (dst_off + bytes > dst.bytes) { set_last_error("tensor larger than its destination"); return B200_ERR_INVALID; }
  B200_CUDA(cudaMemcpyAsync((char*)dst.p + dst_off, src, bytes, cudaMemcpyHostToDevice, s));
  B200_CUDA(cudaStreamSynchronize(s));   // caller may free `src` right after we return
  return B200_OK;
}

int32_t b200_engine_load_tensor(b200_engine* e, const char* name_c, const void* data, int32_t dtype, int32_t ndim,
                                const int64_t* shape) {
  B200_REQUIRE(e && name_c && data && shape && ndim >= 1 && ndim <= 2, "load_tensor: null / ndim");
  if (e->finalized) { set_last_error("load_tensor after finalize"); return B200_ERR_STATE; }
  DeviceGuard guard(e->cfg.device);
  const std::string name(name_c);
  size_t count = 1;
  for (int i = 0; i < ndim; ++i) count *= (size_t)shape[i];
  const size_t esz = (dtype == B200_DTYPE_F16 || dtype == B200_DTYPE_BF16) ? 2 : 4;
  const int32_t dt16 = e->cfg.act_dtype;             // the 16-bit type every non-quantised tensor must have
  auto expect = [&](bool ok, const char* what) {
    if (!ok) set_last_error("load_tensor " + name + ": unexpected " + what);
    return ok;
  };
  int rc;
  if (name == "model.embed_tokens.weight" || name == "lm_head.weight" || name == "model.norm.weight") {
    if (!expect(dtype == dt16, "dtype")) return B200_ERR_INVALID;
    DevMem& dst = name == "model.norm.weight" ? e->final_norm : name == "lm_head.weight" ? e->lm_head_w : e->embed;
    if (name == "lm_head.weight" && e->cfg.tie_embeddings) return B200_OK;   // tied: ignore the duplicate
    if (!expect(count * esz == dst.bytes, "shape")) return B200_ERR_INVALID;
    if ((rc = copy_to(dst, 0, data, count * esz, e->stream))) return rc;
    e->tensors_loaded++;
    return B200_OK;
  }
  if (name == "b200.rope_cos" || name == "b200.rope_sin") {
    // Optional: host-computed RoPE table fp32 [>= max_model_len + 1, 64] (bit-identical to the
    // HF/vLLM fp32 cos/sin cache).  Without it finalize() computes the table itself.
    if (!expect(dtype == B200_DTYPE_F32 && ndim == 2 && shape[1] == 64 && shape[0] >= e->cfg.max_model_len + 1, "rope table"))
      return B200_ERR_INVALID;
    DevMem& dst = name == "b200.rope_cos" ? e->cos_t : e->sin_t;
    dst.release();
    if ((rc = dst.alloc(count * 4)) || (rc = copy_to(dst, 0, data, count * 4, e->stream))) return rc;
    return B200_OK;
  }
  int layer = -1;
  char rest[128] = {0};
  if (sscanf(name_c, "model.layers.%d.%127s", &layer, rest) != 2 || layer < 0 || layer >= e->L) {
    set_last_error("load_tensor: unknown tensor name " + name);
    return B200_ERR_INVALID;
  }
  LayerWeights& w = e->layers[layer];
  const std::string r(rest);
  if (r == "input_layernorm.weight" || r == "post_attention_layernorm.weight") {
    if (!expect(dtype == dt16 && (int)count == e->H, "dtype/shape")) return B200_ERR_INVALID;
    if ((rc = copy_to(r[0] == 'i' ? w.ln1 : w.ln2, 0, data, count * 2, e->stream))) return rc;
    e->tensors_loaded++;
    return B200_OK;
  }
  struct Proj { const char* key; int lin; int row_off; int K; int N; };
  const int hd = kHeadDim;
  const Proj projs[7] = {
      {"self_attn.q_proj", 0, 0, e->H, e->NH * hd},
      {"self_attn.k_proj", 0, e->NH * hd, e->H, e->NKV * hd},
      {"self_attn.v_proj", 0, (e->NH + e->NKV) * hd, e->H, e->NKV * hd},
      {"self_attn.o_proj", 1, 0, e->NH * hd, e->H},
      {"mlp.gate_proj", 2, 0, e->H, e->I},
      {"mlp.up_proj", 2, e->I, e->H, e->I},
      {"mlp.down_proj", 3, 0, e->I, e->H},
  };
  for (const Proj& pj : projs) {
    const std::string key(pj.key);
    if (r.compare(0, key.size() + 1, key + ".") != 0) continue;
    const std::string kind = r.substr(key.size() + 1);
    if (kind == "bias") {
      if (!expect(pj.lin == 0 && dtype == dt16 && (int)count == pj.N, "bias")) return B200_ERR_INVALID;
      if ((rc = copy_to(w.bias_qkv, (size_t)pj.row_off * 2, data, count * 2, e->stream))) return rc;
      e->tensors_loaded++;
      return B200_OK;
    }
    if (kind == "weight") {
      // dense checkpoint: W[N, K] row-major in the activation dtype.  q / k / v are stacked row-wise; gate and up
      // are interleaved in blocks of 64 rows so that every 128-row GEMM tile holds 64 gate rows and the matching
      // 64 up rows (the SwiGLU epilogue pairs them).
      if (!expect(e->dense && dtype == dt16 && ndim == 2 && shape[0] == pj.N && shape[1] == pj.K, "dense weight")) return B200_ERR_INVALID;
      DevMem& dst = w.dense[pj.lin];
      const size_t row_bytes = (size_t)pj.K * 2;
      if (pj.lin == 2) {
        const size_t blk = 64 * row_bytes;
        const size_t off = (pj.row_off ? 64 : 0) * row_bytes;       // up rows take the second half of every tile
        if (off + ((size_t)pj.N / 64 - 1) * 2 * blk + blk > dst.bytes) { set_last_error("tensor larger than its destination"); return B200_ERR_INVALID; }
        B200_CUDA(cudaMemcpy2DAsync((char*)dst.p + off, 2 * blk, data, blk, blk, (size_t)pj.N / 64, cudaMemcpyHostToDevice, e->stream));
        B200_CUDA(cudaStreamSynchronize(e->stream));
      } else if ((rc = copy_to(dst, (size_t)pj.row_off * row_bytes, data, count * 2, e->stream))) {
        return rc;
      }
      e->tensors_loaded++;
      return B200_OK;
    }
    if (e->dense) { set_last_error("load_tensor: " + name + " is an AWQ tensor but the engine was created with dense_weights"); return B200_ERR_INVALID; }
    PendingLinear& pl = e->pending[std::to_string(layer) + "." + key];
    const int K = pj.K, N = pj.N, Gs = e->G;
    if (kind == "qweight") {
      if (!expect(dtype == B200_DTYPE_I32 && ndim == 2 && shape[0] == K && shape[1] == N / 8, "qweight")) return B200_ERR_INVALID;
      if ((rc = pl.qweight.alloc(count * 4)) || (rc = copy_to(pl.qweight, 0, data, count * 4, e->stream))) return rc;
      pl.have_w = true;
    } else if (kind == "qzeros") {
      if (!expect(dtype == B200_DTYPE_I32 && ndim == 2 && shape[0] == K / Gs && shape[1] == N / 8, "qzeros")) return B200_ERR_INVALID;
      if ((rc = pl.qzeros.alloc(count * 4)) || (rc = copy_to(pl.qzeros, 0, data, count * 4, e->stream))) return rc;
      pl.have_z = true;
    } else if (kind == "scales") {
      if (!expect(dtype == B200_DTYPE_F16 && ndim == 2 && shape[0] == K / Gs && shape[1] == N, "scales")) return B200_ERR_INVALID;
      if ((rc = pl.scales.alloc(count * 2)) || (rc = copy_to(pl.scales, 0, data, count * 2, e->stream))) return rc;
      pl.have_s = true;
    } else {
      set_last_error("load_tensor: unknown tensor kind " + name);
      return B200_ERR_INVALID;
    }
    e->tensors_loaded++;
    if (pl.have_w && pl.have_z && pl.have_s) {
      // The fused destination has N_total rows; this projection fills rows [row_off, row_off + N).
      rc = launch_awq_repack(pl.qweight.as<int32_t>(), pl.qzeros.as<int32_t>(), pl.scales.as<__half>(), K, N, Gs,
                             w.wq[pj.lin].as<uint32_t>(), w.sc[pj.lin].as<__half>(), w.zp[pj.lin].as<__half>(),
                             pj.row_off, e->stream);
      if (rc == B200_OK && cudaStreamSynchronize(e->stream) != cudaSuccess) rc = B200_ERR_CUDA;
      pl.qweight.release(); pl.qzeros.release(); pl.scales.release();
      e->pending.erase(std::to_string(layer) + "." + key);
      return rc;
    }
    return B200_OK;
  }
  set_last_error("load_tensor: unknown tensor name " + name);
  return B200_ERR_INVALID;
}

int32_t b200_engine_finalize(b200_engine* e) {
  B200_REQUIRE(e, "null engine");
  if (e->finalized) { set_last_error("finalize called twice"); return B200_ERR_STATE; }
  if (e->tensors_loaded != e->tensors_expected || !e->pending.empty()) {
    set_last_error("finalize: " + std::to_string(e->tensors_loaded) + " of " + std::to_string(e->tensors_expected) +
                   " checkpoint tensors loaded");
    return B200_ERR_STATE;
  }
  DeviceGuard guard(e->cfg.device);
  const b200_config& c = e->cfg;
  int rc;
  // RoPE table: inv_freq = theta^(-2i/d), angle = pos * inv_freq, all fp32 like the HF/vLLM cache
  if (!e->cos_t.p || !e->sin_t.p) {
    const int P = c.max_model_len + 1;
    std::vector<float> cs((size_t)P * 64), sn((size_t)P * 64);
    for (int i = 0; i < 64; ++i) {
      const float inv = 1.0f / powf(c.rope_theta, (float)(2 * i) / (float)kHeadDim);
      for (int p = 0; p < P; ++p) {
        const float a = (float)p * inv;
        cs[(size_t)p * 64 + i] = cosf(a);
        sn[(size_t)p * 64 + i] = sinf(a);
      }
    }
    e->cos_t.release(); e->sin_t.release();
    if ((rc = e->cos_t.alloc(cs.size() * 4)) || (rc = e->sin_t.alloc(sn.size() * 4))) return rc;
    B200_CUDA(cudaMemcpy(e->cos_t.p, cs.data(), cs.size() * 4, cudaMemcpyHostToDevice));
    B200_CUDA(cudaMemcpy(e->sin_t.p, sn.data(), sn.size() * 4, cudaMemcpyHostToDevice));
  }
  // tiled weight copies for the tcgen05 GEMM
  if (e->dense) {
    for (auto& w : e->layers) w.lin[0].bias = w.bias_qkv.as<__half>();
  } else if (c.gemm_path != 1) {
    for (int l = 0; l < e->L; ++l) {
      LayerWeights& w = e->layers[l];
      for (int i = 0; i < 4; ++i) {
        const W4Linear& a = w.lin[i];
        B200_REQUIRE(a.N % 128 == 0 && (a.G == 64 || a.G == 128) && e->I % 64 == 0,
                     "tcgen05 GEMM needs out_features % 128 == 0, intermediate % 64 == 0 and AWQ group size 64 or 128");
        if ((rc = w.chunks[i].alloc(umma_w4_chunk_bytes(a.K, a.N, a.G)))) return rc;
        if ((rc = launch_umma_tile_w4(a.wq, a.sc, a.zp, a.K, a.N, a.G, i == 2 ? e->I : 0,
                                      w.chunks[i].as<unsigned char>(), e->stream)))
          return rc;
        w.ulin[i].chunks = w.chunks[i].as<unsigned char>();
        w.ulin[i].bias = a.bias; w.ulin[i].K = a.K; w.ulin[i].N = a.N; w.ulin[i].G = a.G;
        if (i == 3) { w.ulin[2].next_chunks = w.chunks[3].as<unsigned char>(); w.ulin[2].next_bytes = (int64_t)w.chunks[3].bytes; }
      }
    }
    B200_CUDA(cudaStreamSynchronize(e->stream));
    // The GEMV-layout copy (wq / sc / zp) was only the source of the tiling: the tcgen05 family never reads it.
    // Keeping both layouts resident cost 0.68 GB on the 1.5B model.  (VGATE_B200_DEBUG_GEMV_MASK routes single
    // projections through the GEMV kernels for debugging and therefore keeps it.)
    if (e->debug_gemv_mask == 0) {
      for (auto& w : e->layers)
        for (int i = 0; i < 4; ++i) {
          w.wq[i].release(); w.sc[i].release(); w.zp[i].release();
          w.lin[i].wq = nullptr; w.lin[i].sc = nullptr; w.lin[i].zp = nullptr;
        }
    }
  }
  // KV pool
  e->max_blocks_per_seq = (c.max_model_len + 1 + kBlockTokens - 1) / kBlockTokens + 1;
  const size_t block_bytes_all_layers = (size_t)2 * e->L * e->NKV * kBlockTokens * kHeadDim * 2;
  size_t kv_bytes = c.kv_bytes > 0 ? (size_t)c.kv_bytes : (size_t)c.max_seqs * e->max_blocks_per_seq * block_bytes_all_layers;
  e->n_blocks = (int)(kv_bytes / block_bytes_all_layers);
  B200_REQUIRE(e->n_blocks >= e->max_blocks_per_seq, "kv_bytes too small for one max_model_len sequence");
  e->kv_layer_elems = (size_t)e->n_blocks * e->NKV * kBlockTokens * kHeadDim;
  if ((rc = e->k_cache.alloc(e->kv_layer_elems * e->L * 2)) || (rc = e->v_cache.alloc(e->kv_layer_elems * e->L * 2))) return rc;
  B200_CUDA(cudaMemset(e->k_cache.p, 0, e->k_cache.bytes));
  B200_CUDA(cudaMemset(e->v_cache.p, 0, e->v_cache.bytes));
  for (int b = e->n_blocks - 1; b >= 0; --b) e->free_blocks.push_back(b);
  e->block_meta.assign((size_t)e->n_blocks, b200_engine::BlockMeta{});
  for (int s = c.max_seqs - 1; s >= 0; --s) e->free_slots.push_back(s);
  // activations
  e->Tmax = std::max(std::max(c.max_batch_tokens, c.max_model_len), c.max_seqs);
  const size_t T = e->Tmax, S = c.max_seqs;
  if ((rc = e->h.alloc(T * e->H * 4)) || (rc = e->xn.alloc(T * e->H * 2)) || (rc = e->qkv.alloc(T * e->QKV * 4)) ||
      (rc = e->qh.alloc(T * e->NH * kHeadDim * 2)) || (rc = e->attn.alloc(T * e->NH * kHeadDim * 2)) ||
      (rc = e->gu.alloc(T * 2 * e->I * 4)) || (rc = e->act.alloc(T * e->I * 2)) || (rc = e->xn_last.alloc(S * e->H * 2)) ||
      (rc = e->logits.alloc(S * e->V * 4)) || (rc = e->amax.alloc(((size_t)e->V / 128 + 1) * S * 8)) || (rc = e->d_ids.alloc(T * 4)) || (rc = e->d_pos.alloc(T * 4)) ||
      (rc = e->d_slot.alloc(T * 4)) || (rc = e->d_btrow.alloc(T * 4)) || (rc = e->d_ctx.alloc(T * 4)) ||
      (rc = e->d_lastrows.alloc(S * 4)) || (rc = e->d_tiles.alloc(T * 16)) || (rc = e->d_bt.alloc(S * e->max_blocks_per_seq * 4)) ||
      (rc = e->d_temp.alloc(S * 4)) || (rc = e->d_topp.alloc(S * 4)) || (rc = e->d_seed.alloc(S * 4)) ||
      (rc = e->d_step.alloc(S * 4)) || (rc = e->d_next.alloc(S * 4)) || (rc = e->d_ngen.alloc(S * 4)))
    return rc;
  // metadata is read speculatively by the decode attention before its dependency wait: never garbage
  for (DevMem* m : {&e->d_ids, &e->d_pos, &e->d_slot, &e->d_btrow, &e->d_ctx, &e->d_lastrows, &e->d_tiles, &e->d_bt, &e->d_step,
                    &e->d_next, &e->d_ngen})
    B200_CUDA(cudaMemset(m->p, 0, m->bytes));
  // split-K factors: enough CTAs to engage every SM's dequant throughput on the skinny projections
  // (o, down: 12 row tiles), bounded so partial-sum traffic stays small next to the weight bytes.
  size_t part_elems = 16;
  if (e->dense) {
    // dense decode passes: enough CTAs to pull the weights at HBM speed (tiles x splits ~ the SM count)
    for (int i = 0; i < 4; ++i) {
      const W4Linear& a = e->layers[0].lin[i];
      const int tiles = a.N / 128, KT = a.K / 64;
      int sp = std::max(1, std::min(148 / tiles, KT / 4));    // tiles x splits within one wave of the 148 SMs
      if (i == 2) sp = 1;                               // gate_up feeds SwiGLU directly
      e->splits[i] = sp;
      const int eff = umma_dense_effective_splits(a.K, sp);
      if (eff > 1) part_elems = std::max(part_elems, (size_t)eff * S * a.N);
    }
  } else if (c.gemm_path != 1) {
    const char* ov = getenv("VGATE_B200_SPLITS");      // "qkv,o,gate_up,down" override for tuning
    int o4[4] = {0, 0, 0, 0};
    if (ov) sscanf(ov, "%d,%d,%d,%d", &o4[0], &o4[1], &o4[2], &o4[3]);
    for (int i = 0; i < 4; ++i) {
      const W4Linear& a = e->layers[0].lin[i];
      const int tiles = a.N / 128, KG = a.K / a.G;
      int sp = std::max(1, std::min((148 + tiles / 2) / tiles, KG / 2));
      if (a.K <= 2048) sp = std::min(sp, 4);
      if (i == 2) sp = 1;                               // gate_up feeds SwiGLU directly
      if (o4[i] > 0 && i != 2) sp = o4[i];
      e->splits[i] = sp;
      const int eff = umma_w4_effective_splits(a.K, a.G, sp);
      if (eff > 1) part_elems = std::max(part_elems, (size_t)eff * S * a.N);   // split-K only runs on decode passes (<= max_seqs rows)
    }
  }
  if (e->tp_size > 1) {
    if (!e->nccl_comm) { set_last_error("finalize: tensor-parallel engine without a communicator (call b200_engine_tp_init first)"); return B200_ERR_STATE; }
    part_elems = std::max(part_elems, T * (size_t)e->H);          // o_proj / down_proj outputs of a prefill pass are all-reduced too
    if ((rc = e->tp_epoch.alloc(4)) || (rc = e->tp_recv.alloc((size_t)2 * e->tp_size * S * e->H * 8))) return rc;
    B200_CUDA(cudaMemset(e->tp_recv.p, 0, e->tp_recv.bytes));      // sequence numbers start at 1
    B200_CUDA(cudaMemset(e->tp_epoch.p, 0, 4));
  }
  if ((rc = e->parts.alloc(part_elems * 4))) return rc;
  if ((rc = e->attn_scratch.alloc(attn_scratch_floats(c.max_seqs, e->NH, e->NKV, b200_engine::kDecodeSplits) * 4)) ||
      (rc = e->attn_counters.alloc((size_t)c.max_seqs * e->NKV * 4)))
    return rc;
  B200_CUDA(cudaMemset(e->attn_counters.p, 0, e->attn_counters.bytes));
  B200_CUDA(cudaDeviceSynchronize());
  e->finalized = true;
  if (e->tp_rank == 0) e->worker = std::thread([e] { e->scheduler_loop(); });
  else e->tp_thread = std::thread([e] { e->tp_loop(); });          // a follower rank: executes what the leader posts
  return B200_OK;
}

void b200_engine_destroy(b200_engine* e) {
  if (!e) return;
  {
    std::lock_guard<std::mutex> lk(e->mu);
    e->stop = true;
    for (auto& kv : e->table) {
      Request& r = *kv.second;
      if (r.state == ReqState::WAITING || r.state == ReqState::RUNNING) r.state = ReqState::ABORTED;
    }
    e->cv_work.notify_all();
    e->cv_done.notify_all();
  }
  if (e->worker.joinable()) e->worker.join();
  {
    std::lock_guard<std::mutex> lk(e->tp_mu);
    e->tp_stop = true;
    e->tp_cv.notify_all();
  }
  if (e->tp_thread.joinable()) e->tp_thread.join();
  DeviceGuard guard(e->cfg.device);
  if (e->stream) cudaStreamSynchronize(e->stream);
  for (auto& g : e->decode_graphs) cudaGraphExecDestroy(g.second.exec);
  for (auto& w : e->layers) {
    for (int i = 0; i < 4; ++i) { w.wq[i].release(); w.sc[i].release(); w.zp[i].release(); w.chunks[i].release(); w.dense[i].release(); }
    w.bias_qkv.release(); w.ln1.release(); w.ln2.release();
  }
  for (auto& kv : e->pending) { kv.second.qweight.release(); kv.second.qzeros.release(); kv.second.scales.release(); }
  DevMem* all[] = {&e->embed, &e->lm_head_w, &e->final_norm, &e->cos_t, &e->sin_t, &e->k_cache, &e->v_cache, &e->h,
                   &e->xn, &e->qkv, &e->qh, &e->attn, &e->gu, &e->act, &e->xn_last, &e->logits, &e->amax, &e->d_ids, &e->d_pos,
                   &e->d_slot, &e->d_btrow, &e->d_ctx, &e->d_lastrows, &e->d_bt, &e->d_temp, &e->d_topp, &e->d_seed,
                   &e->d_step, &e->d_next, &e->d_ngen, &e->d_out_tokens, &e->parts, &e->attn_scratch, &e->attn_counters, &e->d_tiles};
  for (DevMem* m : all) m->release();
  for (int i = 0; i < 2; ++i) {
    if (e->ev_a[i]) cudaEventDestroy(e->ev_a[i]);
    if (e->ev_b[i]) cudaEventDestroy(e->ev_b[i]);
    if (e->ev_copied[i]) cudaEventDestroy(e->ev_copied[i]);
    if (e->hp_next[i]) cudaFreeHost(e->hp_next[i]);
    if (e->hp_up[i]) cudaFreeHost(e->hp_up[i]);
    if (e->ev_up[i]) cudaEventDestroy(e->ev_up[i]);
  }
  if (e->nccl_comm) nccl_api().comm_destroy(e->nccl_comm);
  if (e->ev0) cudaEventDestroy(e->ev0);
  if (e->ev1) cudaEventDestroy(e->ev1);
  if (e->stream) cudaStreamDestroy(e->stream);
  delete e;
}

int64_t b200_submit(b200_engine* e, const int32_t* prompt_ids, int32_t n_prompt, float temperature, float top_p,
                    int32_t max_tokens, uint64_t seed) {
  B200_REQUIRE(e && prompt_ids, "null argument");
  B200_REQUIRE(n_prompt > 0 && max_tokens > 0, "n_prompt > 0 && max_tokens > 0");
  B200_REQUIRE(top_p > 0.f && top_p <= 1.f && temperature >= 0.f, "0 < top_p <= 1, temperature >= 0");
  if (!e->finalized) { set_last_error("submit before finalize"); return B200_ERR_STATE; }
  if (e->tp_rank != 0) { set_last_error("submit on a tensor-parallel follower rank: requests go to rank 0"); return B200_ERR_STATE; }
  if (e->tp_size > 1 && (int)e->tp_followers.size() != e->tp_size - 1) { set_last_error("tensor-parallel leader without its followers (b200_engine_tp_attach)"); return B200_ERR_STATE; }
  // vLLM accepts any prompt shorter than max_model_len and stops the generation when the sequence reaches
  // max_model_len (finish reason "length"); only a prompt that leaves no room for a single token is refused.
  if (n_prompt >= e->cfg.max_model_len) {
    set_last_error("prompt (" + std::to_string(n_prompt) + " tokens) does not fit max_model_len " +
                   std::to_string(e->cfg.max_model_len));
    return B200_ERR_CAPACITY;
  }
  max_tokens = std::min(max_tokens, e->cfg.max_model_len - n_prompt);
  for (int i = 0; i < n_prompt; ++i)
    if (prompt_ids[i] < 0 || prompt_ids[i] >= e->V) { set_last_error("prompt token id out of range"); return B200_ERR_INVALID; }
  auto r = std::make_shared<Request>();
  r->prompt.assign(prompt_ids, prompt_ids + n_prompt);
  r->temperature = temperature; r->top_p = top_p; r->max_tokens = max_tokens; r->seed = (uint32_t)(seed ^ (seed >> 32));
  r->t_submit = now_s();
  std::lock_guard<std::mutex> lk(e->mu);
  if (!e->fatal.empty()) { set_last_error("engine failed earlier: " + e->fatal); return B200_ERR_CUDA; }
  if (e->stop) { set_last_error("engine is shutting down"); return B200_ERR_ABORTED; }
  r->id = e->next_id++;
  e->table[r->id] = r;
  e->waiting.push_back(r);
  e->cv_work.notify_all();
  return r->id;
}

static int finish_code(b200_engine* e, const Request& r) {
  switch (r.state) {
    case ReqState::FINISHED: return B200_OK;
    case ReqState::ABORTED: set_last_error("request aborted"); return B200_ERR_ABORTED;
    case ReqState::FAILED: set_last_error("engine failure: " + e->fatal); return B200_ERR_CUDA;
    default: return B200_OK;
  }
}

int32_t b200_wait(b200_engine* e, int64_t rid, int32_t* out_ids, int32_t cap, int32_t* n_out, double* ttft_s,
                  double* gen_time_s) {
  B200_REQUIRE(e, "null engine");
  std::unique_lock<std::mutex> lk(e->mu);
  auto it = e->table.find(rid);
  if (it == e->table.end()) { set_last_error("unknown request id"); return B200_ERR_NOT_FOUND; }
  std::shared_ptr<Request> r = it->second;
  e->cv_done.wait(lk, [&] { return r->state != ReqState::WAITING && r->state != ReqState::RUNNING; });
  const int n = std::min<int>(cap, (int)r->out.size());
  if (out_ids && n > 0) std::memcpy(out_ids, r->out.data(), (size_t)n * 4);
  if (n_out) *n_out = (int)r->out.size();
  // vLLM semantics (vllm_backend.py:26-31): ttft = first token latency, gen_time = last - first token
  if (ttft_s) *ttft_s = r->out.empty() ? 0.0 : r->t_first - r->t_submit;
  if (gen_time_s) *gen_time_s = r->out.empty() ? 0.0 : r->t_last - r->t_first;
  const int code = finish_code(e, *r);
  e->table.erase(rid);
  return code;
}

int32_t b200_poll(b200_engine* e, int64_t rid, int32_t* out_ids, int32_t cap, int32_t* n_new, int32_t* finished,
                  int32_t timeout_ms) {
  B200_REQUIRE(e && n_new && finished, "null argument");
  std::unique_lock<std::mutex> lk(e->mu);
  auto it = e->table.find(rid);
  if (it == e->table.end()) { set_last_error("unknown request id"); return B200_ERR_NOT_FOUND; }
  std::shared_ptr<Request> r = it->second;
  auto ready = [&] { return r->out.size() > r->n_polled || (r->state != ReqState::WAITING && r->state != ReqState::RUNNING); };
  if (timeout_ms > 0) {
    e->poll_waiters++;
    e->cv_done.wait_for(lk, std::chrono::milliseconds(timeout_ms), ready);
    e->poll_waiters--;
  }
  const int avail = (int)(r->out.size() - r->n_polled);
  const int n = std::min(avail, std::max(cap, 0));
  if (out_ids && n > 0) std::memcpy(out_ids, r->out.data() + r->n_polled, (size_t)n * 4);
  r->n_polled += n;
  *n_new = n;
  const bool done = r->state != ReqState::WAITING && r->state != ReqState::RUNNING;
  *finished = (done && r->n_polled == r->out.size()) ? 1 : 0;
  int code = B200_OK;
  if (*finished) { code = finish_code(e, *r); e->table.erase(rid); }
  return code;
}

int32_t b200_abort(b200_engine* e, int64_t rid) {
  B200_REQUIRE(e, "null engine");
  std::lock_guard<std::mutex> lk(e->mu);
  auto it = e->table.find(rid);
  if (it == e->table.end()) { set_last_error("unknown request id"); return B200_ERR_NOT_FOUND; }
  Request& r = *it->second;
  // The caller gives the request up: its table entry (prompt and output vectors) must not outlive it.  A live
  // request is retired and erased by the scheduler; one that already ended is erased here.  A thread blocked in
  // b200_wait / b200_poll on the same id holds its own reference and still gets B200_ERR_ABORTED.
  if (r.state == ReqState::WAITING || r.state == ReqState::RUNNING) {
    r.abort_requested = true;
    r.detached = true;
    e->cv_work.notify_all();
  } else {
    e->table.erase(it);
  }
  return B200_OK;
}

int32_t b200_stats_get(b200_engine* e, b200_stats* out) {
  B200_REQUIRE(e && out, "null argument");
  std::lock_guard<std::mutex> lk(e->mu);
  *out = e->stats;
  out->running = (int)(e->running.size() + e->prefilling.size());
  out->waiting = (int)e->waiting.size();
  out->kv_blocks_total = e->n_blocks;
  out->kv_blocks_free = e->reclaimable_blocks();           // registered prefix blocks nobody references are reclaimable
  return B200_OK;
}

int32_t b200_tp_unique_id(void* id128) {
  B200_REQUIRE(id128, "null argument");
  NcclApi& a = nccl_api();
  if (!a.ok) { set_last_error("libnccl.so.2 not found (set VGATE_B200_NCCL_LIB)"); return B200_ERR_STATE; }
  const int rc = a.get_unique_id(id128);
  return rc ? nccl_fail(rc, "ncclGetUniqueId") : B200_OK;
}
int32_t b200_engine_tp_init(b200_engine* e, const void* id128) {
  B200_REQUIRE(e && id128, "null argument");
  if (e->finalized || e->nccl_comm) { set_last_error("tp_init: call once, before finalize"); return B200_ERR_STATE; }
  B200_REQUIRE(e->tp_size > 1, "tp_init on an engine created with tp_size <= 1");
  NcclApi& a = nccl_api();
  if (!a.ok) { set_last_error("libnccl.so.2 not found (set VGATE_B200_NCCL_LIB)"); return B200_ERR_STATE; }
  DeviceGuard guard(e->cfg.device);
  NcclId id;
  std::memcpy(&id, id128, sizeof(id));
  typedef int (*InitFn)(void**, int, NcclId, int);
  const int rc = reinterpret_cast<InitFn>(a.raw_init)(&e->nccl_comm, e->tp_size, id, e->tp_rank);   // blocks until every rank has called
  return rc ? nccl_fail(rc, "ncclCommInitRank") : B200_OK;
}
int32_t b200_engine_tp_attach(b200_engine* leader, b200_engine* follower) {
  B200_REQUIRE(leader && follower && leader != follower, "null / identical engines");
  B200_REQUIRE(leader->tp_rank == 0 && follower->tp_rank > 0 && leader->tp_size == follower->tp_size, "attach(rank 0, rank r > 0) of one group");
  if (!leader->finalized || !follower->finalized) { set_last_error("tp_attach before finalize"); return B200_ERR_STATE; }
  B200_REQUIRE(leader->n_blocks == follower->n_blocks && leader->cfg.max_seqs == follower->cfg.max_seqs &&
                   leader->Tmax == follower->Tmax && leader->H == follower->H && leader->V == follower->V,
               "tensor-parallel ranks must share max_seqs, token budgets, KV block count, hidden size and vocabulary");
  std::lock_guard<std::mutex> lk(leader->mu);
  leader->tp_followers.push_back(follower);
  if ((int)leader->tp_followers.size() != leader->tp_size - 1) return B200_OK;
  // the group is complete: peer access between every pair of devices, and everybody learns where the others'
  // partial buffers and arrival words live.  Without peer access the decode passes keep using NCCL.
  std::vector<b200_engine*> ranks((size_t)leader->tp_size, nullptr);
  ranks[0] = leader;
  for (b200_engine* f : leader->tp_followers) {
    B200_REQUIRE(!ranks[(size_t)f->tp_rank], "two tensor-parallel followers with the same rank");
    ranks[(size_t)f->tp_rank] = f;
  }
  bool ok = true;
  for (b200_engine* a : ranks)
    for (b200_engine* b : ranks) {
      if (a == b) continue;
      if (a->cfg.device == b->cfg.device) { ok = false; continue; }   // two ranks' kernels waiting for each other on ONE GPU can starve each other of SMs
      int can = 0;
      B200_CUDA(cudaDeviceCanAccessPeer(&can, a->cfg.device, b->cfg.device));
      if (!can) { ok = false; continue; }
      DeviceGuard guard(a->cfg.device);
      const cudaError_t err = cudaDeviceEnablePeerAccess(b->cfg.device, 0);
      if (err == cudaErrorPeerAccessAlreadyEnabled) (void)cudaGetLastError();
      else if (err != cudaSuccess) return cuda_fail(err, "cudaDeviceEnablePeerAccess", __FILE__, __LINE__);
    }
  for (b200_engine* a : ranks) {
    for (b200_engine* b : ranks) {
      a->tp_recv_of[b->tp_rank] = b->tp_recv.as<uint2>();
    }
    a->tp_peer_ok = ok;
  }
  return B200_OK;
}

int32_t b200_engine_status(b200_engine* e) {
  B200_REQUIRE(e, "null engine");
  std::lock_guard<std::mutex> lk(e->mu);
  if (!e->fatal.empty()) { set_last_error("engine failed: " + e->fatal); return B200_ERR_CUDA; }
  if (!e->finalized) { set_last_error("engine not finalized"); return B200_ERR_STATE; }
  return B200_OK;
}

int32_t b200_generate_batch(b200_engine* e, const int32_t* prompt_ids, int32_t n_seqs, int32_t n_prompt,
                            int32_t max_tokens, int32_t* out_ids, float* logits_out, int32_t n_logit_steps,
                            double* gpu_seconds) {
  B200_REQUIRE(e && prompt_ids && out_ids, "null argument");
  B200_REQUIRE(n_seqs > 0 && n_seqs <= e->cfg.max_seqs, "0 < n_seqs <= max_seqs");
  B200_REQUIRE(n_prompt > 0 && max_tokens > 0 && n_prompt + max_tokens <= e->cfg.max_model_len, "prompt + max_tokens <= max_model_len");
  B200_REQUIRE(n_logit_steps >= 0 && n_logit_steps <= max_tokens, "n_logit_steps <= max_tokens");
  if (!e->finalized) { set_last_error("generate_batch before finalize"); return B200_ERR_STATE; }
  if (e->tp_size > 1) { set_last_error("generate_batch runs on single-GPU engines; tensor-parallel engines take submit / wait"); return B200_ERR_STATE; }
  for (int64_t i = 0; i < (int64_t)n_seqs * n_prompt; ++i)
    if (prompt_ids[i] < 0 || prompt_ids[i] >= e->V) { set_last_error("prompt token id out of range"); return B200_ERR_INVALID; }
  {
    std::unique_lock<std::mutex> lk(e->mu);
    if (!e->fatal.empty()) { set_last_error("engine failed earlier: " + e->fatal); return B200_ERR_CUDA; }
    if (e->exclusive || e->stepping || e->inflight.active || !e->running.empty() || !e->prefilling.empty() || !e->waiting.empty()) {
      set_last_error("generate_batch needs an idle engine");
      return B200_ERR_STATE;
    }
    e->exclusive = true;
  }
  struct Release { b200_engine* e; ~Release() { std::lock_guard<std::mutex> lk(e->mu); e->exclusive = false; e->cv_work.notify_all(); } } release{e};
  DeviceGuard guard(e->cfg.device);
  cudaStream_t s = e->stream;
  { const int rc0 = e->collect_decode(); if (rc0) return rc0; }   // a look-ahead step of sequences aborted since
  { std::lock_guard<std::mutex> lk(e->mu); e->drop_prefix_cache(); }
  e->last_decode_ids.clear();
  const int bps = (n_prompt + max_tokens + 1 + kBlockTokens - 1) / kBlockTokens;   // blocks per sequence
  B200_REQUIRE(bps <= e->max_blocks_per_seq, "sequence needs more blocks than max_blocks_per_seq");
  if ((int64_t)bps * n_seqs > e->n_blocks) { set_last_error("KV pool too small for this batch"); return B200_ERR_CAPACITY; }
  int rc;
  // block tables: sequence i owns blocks [i*bps, (i+1)*bps) (pool is idle, so any assignment is free)
  std::vector<int32_t> bt((size_t)e->cfg.max_seqs * e->max_blocks_per_seq, 0);
  for (int i = 0; i < n_seqs; ++i)
    for (int b = 0; b < bps; ++b) bt[(size_t)i * e->max_blocks_per_seq + b] = i * bps + b;
  if ((rc = e->upload_i32(e->d_bt, bt))) return rc;
  std::vector<float> temp(n_seqs, 0.f), topp(n_seqs, 1.f);
  std::vector<int32_t> zeros(n_seqs, 0);
  if ((rc = e->upload_f32(e->d_temp, temp)) || (rc = e->upload_f32(e->d_topp, topp)) || (rc = e->upload_i32(e->d_seed, zeros)) ||
      (rc = e->upload_i32(e->d_step, zeros)) || (rc = e->upload_i32(e->d_ngen, zeros)))
    return rc;
  if (e->d_out_tokens.bytes < (size_t)n_seqs * max_tokens * 4) {
    e->d_out_tokens.release();
    if ((rc = e->d_out_tokens.alloc((size_t)n_seqs * max_tokens * 4))) return rc;
  }
  const int64_t launches0 = g_launch_count;
  // ---- prefill in chunks of whole sequences ----
  const int seqs_per_chunk = std::max(1, std::min(n_seqs, e->Tmax / n_prompt));
  B200_REQUIRE(n_prompt <= e->Tmax, "prompt longer than the token buffers");
  std::vector<int32_t> first_tokens(n_seqs);
  B200_CUDA(cudaEventRecord(e->ev0, s));
  for (int s0 = 0; s0 < n_seqs; s0 += seqs_per_chunk) {
    const int ns = std::min(seqs_per_chunk, n_seqs - s0);
    std::vector<int32_t> ids, pos, slot, btrow, ctx, lastrows;
    for (int i = 0; i < ns; ++i) {
      for (int p = 0; p < n_prompt; ++p) {
        ids.push_back(prompt_ids[(size_t)(s0 + i) * n_prompt + p]);
        pos.push_back(p);
        slot.push_back(((s0 + i) * bps + p / kBlockTokens) * kBlockTokens + p % kBlockTokens);
        btrow.push_back(s0 + i);
        ctx.push_back(p + 1);
      }
      lastrows.push_back((i + 1) * n_prompt - 1);
    }
    if ((rc = e->upload_i32(e->d_ids, ids)) || (rc = e->upload_i32(e->d_pos, pos)) || (rc = e->upload_i32(e->d_slot, slot)) ||
        (rc = e->upload_i32(e->d_btrow, btrow)) || (rc = e->upload_i32(e->d_ctx, ctx)) ||
        (rc = e->upload_i32(e->d_lastrows, lastrows)) || (rc = e->upload_q_tiles(btrow, ctx)))
      return rc;
    e->greedy_only = !(logits_out && n_logit_steps > 0);
    if ((rc = e->forward(ns * n_prompt, ns, e->d_lastrows.as<int32_t>(), false))) return rc;
    B200_CUDA(cudaMemcpyAsync(first_tokens.data() + s0, e->d_next.p, (size_t)ns * 4, cudaMemcpyDeviceToHost, s));
    if (logits_out && n_logit_steps > 0)
      B200_CUDA(cudaMemcpy2DAsync(logits_out + (size_t)s0 * n_logit_steps * e->V, (size_t)n_logit_steps * e->V * 4,
                                  e->logits.p, (size_t)e->V * 4, (size_t)e->V * 4, ns, cudaMemcpyDeviceToHost, s));
    B200_CUDA(cudaStreamSynchronize(s));   // host vectors above go out of scope
  }
  B200_CUDA(cudaEventRecord(e->ev1, s));
  B200_CUDA(cudaStreamSynchronize(s));
  float ms_prefill = 0.f;
  B200_CUDA(cudaEventElapsedTime(&ms_prefill, e->ev0, e->ev1));
  // ---- decode: device-resident loop (advance -> forward), no host interaction per step ----
  {
    std::vector<int32_t> pos(n_seqs, n_prompt - 1), btrow(n_seqs);
    for (int i = 0; i < n_seqs; ++i) btrow[i] = i;
    if ((rc = e->upload_i32(e->d_pos, pos)) || (rc = e->upload_i32(e->d_btrow, btrow)) ||
        (rc = e->upload_i32(e->d_next, first_tokens)))
      return rc;
    B200_CUDA(cudaStreamSynchronize(s));
  }
  e->last_decode_ids.clear();
  B200_CUDA(cudaEventRecord(e->ev0, s));
  for (int t = 1; t < max_tokens; ++t) {
    e->greedy_only = !(logits_out && t < n_logit_steps);      // logits are materialised only when asked for
    if ((rc = e->decode_step(n_seqs, e->d_out_tokens.as<int32_t>(), max_tokens))) return rc;
    if (logits_out && t < n_logit_steps)
      B200_CUDA(cudaMemcpy2DAsync(logits_out + (size_t)t * e->V, (size_t)n_logit_steps * e->V * 4, e->logits.p,
                                  (size_t)e->V * 4, (size_t)e->V * 4, n_seqs, cudaMemcpyDeviceToHost, s));
  }
  // record the last sampled token
  B200_CUDA(launch_k(advance_decode_kernel, dim3((n_seqs + 127) / 128), dim3(128), 0, s, n_seqs, e->d_next.as<int32_t>(),
                     e->d_ids.as<int32_t>(), e->d_pos.as<int32_t>(), e->d_slot.as<int32_t>(), e->d_ctx.as<int32_t>(),
                     e->d_ngen.as<int32_t>(), e->d_btrow.as<int32_t>(), e->d_bt.as<int32_t>(), e->max_blocks_per_seq,
                     e->d_out_tokens.as<int32_t>(), max_tokens, e->d_step.as<int32_t>()));
  g_launch_count += 1;
  B200_CUDA(cudaEventRecord(e->ev1, s));
  B200_CUDA(cudaMemcpyAsync(out_ids, e->d_out_tokens.p, (size_t)n_seqs * max_tokens * 4, cudaMemcpyDeviceToHost, s));
  B200_CUDA(cudaStreamSynchronize(s));
  float ms_decode = 0.f;
  B200_CUDA(cudaEventElapsedTime(&ms_decode, e->ev0, e->ev1));
  if (gpu_seconds) { gpu_seconds[0] = ms_prefill * 1e-3; gpu_seconds[1] = ms_decode * 1e-3; }
  std::lock_guard<std::mutex> lk(e->mu);
  e->stats.kernel_launches += g_launch_count - launches0;
  e->stats.steps_prefill += (n_seqs + seqs_per_chunk - 1) / seqs_per_chunk;
  e->stats.tokens_prefill += (int64_t)n_seqs * n_prompt;
  e->stats.gpu_seconds_prefill += ms_prefill * 1e-3;
  e->stats.steps_decode += max_tokens - 1;
  e->stats.gpu_seconds_decode += ms_decode * 1e-3;
  e->stats.decode_batch_sum += (int64_t)n_seqs * (max_tokens - 1);
  e->stats.tokens_generated += (int64_t)n_seqs * max_tokens;
  for (int t = 1; t < max_tokens; ++t) e->stats.hbm_bytes_decode += e->decode_step_bytes(n_seqs, (int64_t)n_seqs * (n_prompt + t));
  return B200_OK;
}

// ---- kernel entry points ----------------------------------------------------------------------
int32_t b200_op_awq_repack(const int32_t* qweight, const int32_t* qzeros, const void* scales, int32_t K, int32_t N,
                           int32_t G, uint32_t* wq, void* sc, void* zp, void* stream) {
  return launch_awq_repack(qweight, qzeros, (const __half*)scales, K, N, G, wq, (__half*)sc, (__half*)zp, 0,
                           (cudaStream_t)stream);
}
int32_t b200_op_awq_dequant(const uint32_t* wq, const void* sc, const void* zp, int32_t K, int32_t N, int32_t G,
                            void* w_out, void* stream) {
  return launch_awq_dequant(wq, (const __half*)sc, (const __half*)zp, K, N, G, (__half*)w_out, (cudaStream_t)stream);
}
int32_t b200_op_w4_gemv(const uint32_t* wq, const void* sc, const void* zp, const void* x, int32_t ldx, float* y,
                        int32_t ldy, const void* bias, int32_t B, int32_t K, int32_t N, int32_t G, int32_t accumulate,
                        void* stream) {
  W4Linear w;
  w.wq = wq; w.sc = (const __half*)sc; w.zp = (const __half*)zp; w.bias = (const __half*)bias; w.K = K; w.N = N; w.G = G;
  return launch_w4_gemv(w, (const __half*)x, ldx, y, ldy, B, accumulate != 0, (cudaStream_t)stream);
}
int32_t b200_op_w4_tile(const uint32_t* wq, const void* sc, const void* zp, int32_t K, int32_t N, int32_t G,
                        int32_t pair_rows, void* chunks, void* stream) {
  return launch_umma_tile_w4(wq, (const __half*)sc, (const __half*)zp, K, N, G, pair_rows, (unsigned char*)chunks,
                             (cudaStream_t)stream);
}
int64_t b200_op_w4_tile_bytes(int32_t K, int32_t N, int32_t G) {
  if (K <= 0 || N <= 0 || N % 128 != 0 || (G != 64 && G != 128) || K % G != 0) return B200_ERR_INVALID;
  return (int64_t)umma_w4_chunk_bytes(K, N, G);
}
int32_t b200_op_w4_gemm_umma(const void* chunks, const void* x, int32_t ldx, float* y, int32_t ldy, const void* bias,
                             int32_t T, int32_t K, int32_t N, int32_t G, int32_t epilogue, int32_t n_splits,
                             int64_t part_stride, void* stream) {
  const int accumulate = epilogue == 1;
  B200_REQUIRE(T >= 0 && n_splits >= 1 && epilogue >= 0 && epilogue <= 2, "w4_gemm_umma: T >= 0, n_splits >= 1, epilogue in 0..2");
  B200_REQUIRE(n_splits == 1 || (epilogue == 0 && !bias), "w4_gemm_umma: split-K output is raw partials");
  B200_REQUIRE(epilogue != 2 || !bias, "w4_gemm_umma: SwiGLU epilogue takes no bias");
  B200_REQUIRE(n_splits == umma_w4_effective_splits(K, G, n_splits), "w4_gemm_umma: n_splits must divide the groups evenly enough (use b200_op_w4_effective_splits)");
  if (T == 0) return B200_OK;
  UmmaW4 w;
  w.chunks = (const unsigned char*)chunks; w.bias = (const __half*)bias; w.K = K; w.N = N; w.G = G;
  const int nt = umma_pick_nt(T, true);
  CUtensorMap mx;
  const char* xr = getenv("VGATE_B200_DEBUG_XROWS");     // debug: map more rows than T (stale rows, as in the engine)
  int rc = make_tensor_map_f16(&mx, x, xr ? atoi(xr) : T, K, ldx, nt);
  if (rc) return rc;
  const int mode = n_splits > 1 ? kUmmaPartialF32 : epilogue == 2 ? kUmmaSwigluF16 : accumulate ? kUmmaAccumF32 : kUmmaStoreF32;
  return launch_umma_w4(w, mx, nt, T, y, ldy, mode, n_splits, part_stride, (cudaStream_t)stream);
}
void b200_debug_umma_trace(void* dev_buf) { umma_set_trace((unsigned long long*)dev_buf, 0, 0); }
void b200_debug_umma_trace_grid(void* dev_buf, int32_t grid_x, int32_t grid_z) { umma_set_trace((unsigned long long*)dev_buf, grid_x, grid_z); }
void b200_debug_attn_trace(void* dev_buf) { attn_set_trace((unsigned long long*)dev_buf); }
void b200_debug_timeline(void* buf, void* counter, uint32_t capacity) {
  auto* b = (unsigned long long*)buf; auto* c = (unsigned int*)counter;
  tl_set_engine(b, c, capacity); tl_set_basic(b, c, capacity); tl_set_gemv(b, c, capacity); tl_set_attn(b, c, capacity);
  tl_set_sample(b, c, capacity); tl_set_umma(b, c, capacity);
}
int32_t b200_op_w4_effective_splits(int32_t K, int32_t G, int32_t n_splits) {
  if (K <= 0 || G <= 0 || K % G != 0 || n_splits < 1) return B200_ERR_INVALID;
  return umma_w4_effective_splits(K, G, n_splits);
}
int32_t b200_op_f16_gemm_umma(const void* w, const void* x, int32_t ldx, float* y, int32_t ldy, int32_t T, int32_t K,
                              int32_t N, int32_t accumulate, void* stream) {
  if (T == 0) return B200_OK;
  const int nt = umma_pick_nt(T, false);
  CUtensorMap mx, mw;
  int rc = make_tensor_map_f16(&mx, x, T, K, ldx, nt);
  if (rc) return rc;
  if ((rc = make_tensor_map_f16(&mw, w, N, K, K, 128))) return rc;
  return launch_umma_f16(mw, K, N, mx, nt, T, y, ldy, accumulate ? kUmmaAccumF32 : kUmmaStoreF32, (cudaStream_t)stream);
}
int32_t b200_op_dense_gemm_umma(const void* w, const void* x, int32_t ldx, float* y, int32_t ldy, const void* bias, int32_t T,
                                int32_t K, int32_t N, int32_t dtype, int32_t epilogue, int32_t n_splits, int64_t part_stride,
                                void* stream) {
  B200_REQUIRE(dtype == B200_DTYPE_F16 || dtype == B200_DTYPE_BF16, "dense_gemm_umma: dtype is B200_DTYPE_F16 or B200_DTYPE_BF16");
  B200_REQUIRE(T >= 0 && n_splits >= 1 && epilogue >= 0 && epilogue <= 2, "dense_gemm_umma: T >= 0, n_splits >= 1, epilogue in 0..2");
  B200_REQUIRE(n_splits == 1 || (epilogue == 0 && !bias), "dense_gemm_umma: split-K output is raw partials");
  B200_REQUIRE(epilogue != 2 || !bias, "dense_gemm_umma: SwiGLU epilogue takes no bias");
  B200_REQUIRE(K > 0 && K % 64 == 0 && n_splits == umma_dense_effective_splits(K, n_splits),
               "dense_gemm_umma: K % 64 == 0 and n_splits must equal b200_op_dense_effective_splits(K, n_splits)");
  if (T == 0) return B200_OK;
  const int nt = umma_pick_nt(T, false);
  CUtensorMap mx, mw;
  int rc = make_tensor_map_f16(&mx, x, T, K, ldx, nt);
  if (rc) return rc;
  if ((rc = make_tensor_map_f16(&mw, w, N, K, K, 128))) return rc;
  const int mode = n_splits > 1 ? kUmmaPartialF32 : epilogue == 2 ? kUmmaSwigluF16 : epilogue == 1 ? kUmmaAccumF32 : kUmmaStoreF32;
  return launch_umma_f16(mw, K, N, mx, nt, T, y, ldy, mode, (cudaStream_t)stream, n_splits, part_stride, (const __half*)bias,
                         dtype == B200_DTYPE_BF16);
}
int32_t b200_op_dense_effective_splits(int32_t K, int32_t n_splits) {
  if (K <= 0 || K % 64 != 0 || n_splits < 1) return B200_ERR_INVALID;
  return umma_dense_effective_splits(K, n_splits);
}
int32_t b200_op_lm_head_argmax(const void* w, const void* x, int32_t ldx, int32_t T, int32_t K, int32_t N, float* scratch,
                               int32_t* out, void* stream) {
  if (T == 0) return B200_OK;
  B200_REQUIRE(scratch && out, "lm_head_argmax: null scratch / out");
  const int nt = umma_pick_nt(T, false);
  CUtensorMap mx, mw;
  int rc = make_tensor_map_f16(&mx, x, T, K, ldx, nt);
  if (rc) return rc;
  if ((rc = make_tensor_map_f16(&mw, w, N, K, K, 128))) return rc;
  if ((rc = launch_umma_f16(mw, K, N, mx, nt, T, scratch, T, kUmmaArgmaxF32, (cudaStream_t)stream))) return rc;
  return launch_argmax_reduce(scratch, N / 128, T, T, out, (cudaStream_t)stream);
}
int32_t b200_op_f16_gemv(const void* w, const void* x, int32_t ldx, float* y, int32_t ldy, int32_t B, int32_t K,
                         int32_t N, int32_t accumulate, void* stream) {
  return launch_f16_gemv((const __half*)w, (const __half*)x, ldx, y, ldy, B, K, N, accumulate != 0, (cudaStream_t)stream);
}
int32_t b200_op_rmsnorm(float* h, const float* parts, int32_t n_parts, int64_t part_stride, const void* gamma, float eps,
                        const int32_t* rows, int32_t n_rows, int32_t H, void* xn, void* stream) {
  return launch_rmsnorm(h, parts, n_parts, part_stride, (const __half*)gamma, eps, rows, n_rows, H, (__half*)xn,
                        (cudaStream_t)stream);
}
int32_t b200_op_rope_kv(const float* qkv, int32_t n_parts, int64_t part_stride, const void* bias, const int32_t* pos,
                        const int32_t* slot, const float* cos_t, const float* sin_t, int32_t T, int32_t n_heads,
                        int32_t n_kv, void* q_out, void* k_cache, void* v_cache, void* stream) {
  return launch_rope_kv(qkv, n_parts, part_stride, (const __half*)bias, pos, slot, cos_t, sin_t, T, n_heads, n_kv,
                        (__half*)q_out, (__half*)k_cache, (__half*)v_cache, (cudaStream_t)stream);
}
int32_t b200_op_paged_attn(const void* q, const void* k_cache, const void* v_cache, const int32_t* block_tables,
                           int32_t bt_stride, const int32_t* bt_row, const int32_t* ctx_len, int32_t T, int32_t n_heads,
                           int32_t n_kv, int32_t n_splits, float* scratch, int32_t* counters, void* out, void* stream) {
  return launch_paged_attn((const __half*)q, (const __half*)k_cache, (const __half*)v_cache, block_tables, bt_stride,
                           bt_row, ctx_len, T, n_heads, n_kv, n_splits, scratch, counters, (__half*)out,
                           (cudaStream_t)stream);
}
int64_t b200_op_attn_scratch_bytes(int32_t T, int32_t n_heads, int32_t n_kv, int32_t n_splits) {
  if (T < 0 || n_kv <= 0 || n_heads % n_kv != 0 || n_splits < 1) return B200_ERR_INVALID;
  return (int64_t)attn_scratch_floats(T, n_heads, n_kv, n_splits) * 4;
}
int32_t b200_op_attn_decode_fused(const float* qkv, int32_t n_parts, int64_t part_stride, const void* bias,
                                  const int32_t* pos, const int32_t* slot, const float* cos_t, const float* sin_t,
                                  void* k_cache, void* v_cache, const int32_t* block_tables, int32_t bt_stride,
                                  const int32_t* bt_row, const int32_t* ctx_len, int32_t T, int32_t n_heads, int32_t n_kv,
                                  int32_t n_splits, float* scratch, int32_t* counters, void* out, void* stream) {
  return launch_attn_decode_fused(qkv, n_parts, part_stride, (const __half*)bias, pos, slot, cos_t, sin_t, (__half*)k_cache,
                                  (__half*)v_cache, block_tables, bt_stride, bt_row, ctx_len, T, n_heads, n_kv, n_splits,
                                  scratch, counters, (__half*)out, (cudaStream_t)stream);
}
int32_t b200_op_swiglu(const float* gu, int32_t T, int32_t I, void* act, void* stream) {
  return launch_swiglu(gu, T, I, (__half*)act, (cudaStream_t)stream);
}
int32_t b200_op_embed(const void* table, const int32_t* ids, int32_t T, int32_t H, float* h, void* stream) {
  return launch_embed((const __half*)table, ids, T, H, h, (cudaStream_t)stream);
}
int32_t b200_op_sample(const float* logits, int32_t B, int32_t V, const float* temperature, const float* top_p,
                       const uint32_t* seed, const int32_t* step, int32_t* out, void* stream) {
  return launch_sample(logits, B, V, temperature, top_p, seed, step, out, (cudaStream_t)stream);
}

}  // extern "C"
