#!/usr/bin/env python
"""Headline benchmark: generated tokens/s of the b200 worker engine on Qwen2.5-1.5B W4A16
(AWQ-shape, random-init), 64 concurrent sequences x 128-token prompts x 256 new tokens, greedy.

One "step" = one full pass of the hot path over one batch: prefill of all prompts + the
device-resident decode loop until every sequence has `new_tokens` tokens.

  python bench.py [--gpus N --steps K --warmup W]            this repo's CUDA path
  python bench.py --impl reference ...                        the reference arithmetic on host cores
                                                              (oracle/: fp32 PyTorch eager of the same model)
Under torchrun each rank drives one independent replica on its own GPU (weak scaling, no
collective on the data path: the path shards by request).  Rank 0 prints one JSON line.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "generated_tokens_per_second"
UNIT = "tok/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--preset", default="qwen2.5-1.5b-awq")
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--prompt-len", type=int, default=128)
    ap.add_argument("--new-tokens", type=int, default=256)
    ap.add_argument("--gemm-path", type=int, default=0)
    ap.add_argument("--no-graphs", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--tp", type=int, default=1, help="tensor-parallel ranks per replica (config #5: 2); a replica then spans "
                    "--tp consecutive GPUs and is driven by the torchrun rank of its first GPU")
    ap.add_argument("--no-gateway", action="store_true", help="skip the worker-HTTP / gateway levels (tools/bench_gateway.py)")
    ap.add_argument("--profile", default="peaked", choices=["peaked", "flat"],
                    help="synthetic checkpoint profile (vgate_b200/weights.py); timing does not depend on it, the "
                         "self-check against the oracle needs well-defined greedy tokens")
    return ap.parse_args()


def workload_name(a):
    kind = "W4A16" if "awq" in a.preset else "dense 16-bit"
    return (f"{a.preset} {kind} random-init, batch {a.batch} x {a.prompt_len}-token prompts, greedy "
            f"{a.new_tokens} new tokens (prefill + decode)")


def config_of(a, world):
    """The workload both arms are measured on (the reference arm prints the identical dict)."""
    tp = getattr(a, "tp", 1)
    return {"workload": workload_name(a), "replicas": world // tp,
            "parallelism": f"dp{world} (one replica per GPU)" if tp == 1 else f"dp{world // tp} x tp{tp} (NCCL all-reduce inside a replica)",
            "l2": "per-step working set (weights 1.15 GB + KV) exceeds the 126 MB L2",
            "checkpoint_profile": a.profile}


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        p = json.load(open(path))
        return float(p["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks + throttle reasons during the timed region."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index=0):
        self.proc, self.lines, self.gpu = None, [], gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "200", "-i", str(self.gpu)], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        sm, smax, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for l in self.lines:
            f = [x.strip() for x in l.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); smax = float(f[2])
            except ValueError:
                continue
            for name, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": smax, "reasons": sorted(reasons),
                "samples": len(sm)}


def reduce_over_ranks(values, op="max", device="cpu"):
    """Element-wise max / sum over ranks of a small list of floats (identity when not distributed).
    The timing rule of the bench contract: a multi-GPU time is the MAX over ranks."""
    import torch
    import torch.distributed as dist
    t = torch.tensor(list(values), dtype=torch.float64, device=device)
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX if op == "max" else dist.ReduceOp.SUM)
    return t.tolist()


def aggregate_throughput(tokens_per_rank, seconds_this_rank, device="cpu"):
    """Whole-job tokens/s: all ranks' tokens divided by the slowest rank's time (weak scaling, one
    independent replica per rank, no data-path collective)."""
    total = reduce_over_ranks([tokens_per_rank], "sum", device)[0]
    worst = reduce_over_ranks([seconds_this_rank], "max", device)[0]
    return total / worst, worst


def make_prompts(rng, a, vocab):
    return rng.integers(0, vocab, size=(a.batch, a.prompt_len), dtype=np.int32)


# ---------------------------------------------------------------------------------------------------
class CpuReference:
    """The reference arithmetic (oracle/: PyTorch eager forward of the same architecture, the CPU path
    north_star names) on the host cores, timed on the benchmark's own configuration.

    A full pass (batch x prompt_len prefill + new_tokens decode steps of 1.5 B parameters) is minutes of CPU
    work, so each step times a bounded sample of the SAME batch and extrapolates:
      * prefill: `pf_seqs` of the batch's sequences at the full prompt length; linear in sequences (a CPU GEMM
        is at its plateau from ~1k token rows);
      * decode: `dec_steps` steps of the WHOLE batch (every step streams all weights once, as on the GPU),
        taken at a short context: attention is 1-2 % of a CPU step's flops at context 128..384.
    value = batch * new_tokens / (prefill_s * batch / pf_seqs + (new_tokens - 1) * decode_step_s)."""

    def __init__(self, preset, profile, batch, prompt_len, new_tokens, pf_seqs=None, dec_steps=None, modes=None):
        import torch
        from oracle.qwen2 import OracleQwen2, Qwen2Dims
        from vgate_b200.weights import PRESETS, synthetic_checkpoint
        self.torch = torch
        self.cores = os.cpu_count() or 1
        torch.set_num_threads(self.cores)
        self.dims = PRESETS[preset]
        self.batch, self.prompt_len, self.new_tokens = batch, prompt_len, new_tokens
        d = self.dims
        n_params = d.n_layers * (d.hidden * (d.n_heads + 2 * d.n_kv_heads) * d.head_dim + d.hidden * d.n_heads * d.head_dim +
                                 3 * d.hidden * d.intermediate) + d.vocab * d.hidden
        self.big = big = n_params > 3e9         # 7B / 14B: one arithmetic, a smaller sample (the run stays within minutes)
        modes = modes or (("bf16",) if big else ("fp32", "bf16"))
        # about 1k-2k prompt tokens per sample, whatever the prompt length
        pf_seqs = pf_seqs or max(1, min(8, (1024 if big else 1024) // max(1, prompt_len)))
        dec_steps = dec_steps or (2 if big else 4)
        self.pf_seqs, self.dec_steps = min(pf_seqs, batch), dec_steps
        self.profile = profile
        ck = synthetic_checkpoint(self.dims, seed=0, profile=profile)
        self.models = {m: OracleQwen2(Qwen2Dims(**self.dims.to_dict()), ck, mode=m) for m in modes}
        del ck
        self.rng = np.random.default_rng(1)
        self.mode = None                      # fastest arithmetic, chosen by the first step

    def fp32(self):
        if "fp32" not in self.models:           # the self-check's oracle (big presets only time bf16)
            from oracle.qwen2 import OracleQwen2, Qwen2Dims
            from vgate_b200.weights import synthetic_checkpoint
            self.models["fp32"] = OracleQwen2(Qwen2Dims(**self.dims.to_dict()), synthetic_checkpoint(self.dims, seed=0, profile=self.profile), mode="fp32")
        return self.models["fp32"]

    def _step(self, mode):
        m = self.models[mode]
        V = self.dims.vocab
        m.reset()
        t0 = time.perf_counter()
        m.forward(self.rng.integers(0, V, size=(self.pf_seqs, self.prompt_len)), last_only=True)
        t_pf = time.perf_counter() - t0
        m.reset()
        logits = m.forward(self.rng.integers(0, V, size=(self.batch, 8)), last_only=True)[:, -1]     # a short context, untimed
        t0 = time.perf_counter()
        for _ in range(self.dec_steps):
            nxt = logits.argmax(-1)
            logits = m.forward(nxt[:, None].numpy(), last_only=True)[:, -1]
        t_dec = (time.perf_counter() - t0) / self.dec_steps
        m.reset()
        full = t_pf * self.batch / self.pf_seqs + (self.new_tokens - 1) * t_dec
        return {"tok_per_s": self.batch * self.new_tokens / full, "prefill_s": t_pf, "decode_step_s": t_dec,
                "sample_s": t_pf + t_dec * self.dec_steps, "full_pass_s": full}

    def run(self, steps=1, warmup=0):
        per_mode = {}
        if self.mode is None:                 # one probe step per arithmetic, keep the faster (it also warms up)
            for mode in self.models:
                per_mode[mode] = self._step(mode)
            self.mode = max(per_mode, key=lambda k: per_mode[k]["tok_per_s"])
            warmup = max(0, warmup - 1)
        for _ in range(warmup):
            self._step(self.mode)
        recs = [self._step(self.mode) for _ in range(steps)]
        full = float(np.mean([r["full_pass_s"] for r in recs]))
        out = {"value": self.batch * self.new_tokens / full, "unit": UNIT, "cores": self.cores, "kind": "port",
               "arith": self.mode,
               "sample": (f"per step: prefill of {self.pf_seqs} of the {self.batch} sequences x {self.prompt_len} tokens "
                          f"({np.mean([r['prefill_s'] for r in recs]):.2f} s) + {self.dec_steps} decode steps of all {self.batch} "
                          f"sequences ({np.mean([r['decode_step_s'] for r in recs]) * 1e3:.0f} ms each), {self.mode} PyTorch eager, "
                          f"{len(recs)} step(s); extrapolated to the full pass ({self.batch} x {self.prompt_len} prefill + "
                          f"{self.new_tokens - 1} decode steps = {full:.1f} s)"),
               "probe_tok_per_s": {k: v["tok_per_s"] for k, v in per_mode.items()}}
        return out, full * 1e3, float(np.mean([r["sample_s"] for r in recs]))


def run_reference(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    ref = CpuReference(a.preset, a.profile, a.batch, a.prompt_len, a.new_tokens)
    cb, ms, _ = ref.run(steps=max(1, a.steps), warmup=min(a.warmup, 1))
    line = {
        "impl": "reference", "metric": METRIC, "value": cb["value"], "unit": UNIT, "n_gpus": a.gpus, "steps": a.steps,
        "warmup": a.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": cb["arith"], "data": "synthetic", "config": config_of(a, max(1, a.gpus)),
        "cpu_baseline": cb,
        "e2e": {"value": cb["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------------------------
def dense_kernel_roofline(a, dims):
    """Dense checkpoints (config #4): the dominant kernel of a prefill-heavy pass is the gate_up GEMM with the SwiGLU
    epilogue, [tokens x H] x [2I x H]^T on the tensor cores.  Timed alone with CUDA events on its launch stream
    (graph of 6 launches), against the measured cuBLAS bf16 burst throughput of this pool."""
    import ctypes as C
    import torch
    from vgate_b200 import _lib
    lib = _lib.load()
    K, I = dims.hidden, dims.intermediate // max(1, getattr(a, "tp", 1))   # a tensor-parallel rank holds I / tp of the columns
    N, T = 2 * I, min(8192, a.batch * a.prompt_len)
    peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {}
    peak, src = (float(peaks["bf16_tflops"]), "measured burst (MEASURED_PEAKS.json bf16_tflops)") if peaks else (1590.0, "fallback (B200_PROFILING.md)")
    st = torch.cuda.Stream()
    sp = C.c_void_p(st.cuda_stream)
    ptr = lambda t: C.c_void_p(t.data_ptr())
    dt = torch.bfloat16 if dims.dtype == "bf16" else torch.float16
    code = _lib.DTYPE_BF16 if dims.dtype == "bf16" else _lib.DTYPE_F16
    with torch.cuda.stream(st):
        w = (torch.randn((N, K), device="cuda") * 0.02).to(dt)
        x = torch.randn((T, K), device="cuda").to(dt)
        act = torch.empty((T, I), dtype=dt, device="cuda")
        launch = lambda: _lib.check(lib.b200_op_dense_gemm_umma(ptr(w), ptr(x), K, ptr(act), I, None, T, K, N, code, 2, 1, 0, sp))
        launch()
        st.synchronize()
        g = torch.cuda.CUDAGraph()
        reps = 6
        with torch.cuda.graph(g, stream=st):
            for _ in range(reps):
                launch()
        times = []
        for it in range(6):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(st); g.replay(); e1.record(st); st.synchronize()
            if it >= 3:
                times.append(e0.elapsed_time(e1) * 1e-3 / reps)
    t = float(np.mean(times))
    flops = 2.0 * T * K * N
    ach = flops / t / 1e12
    return {"bound": "tensor", "kernel": f"umma_gemm_kernel<256, dense {dims.dtype}> gate_up + SwiGLU [{T}x{K}]x[{N}x{K}]", "achieved": ach,
            "peak": peak, "unit": "TFLOP/s", "frac": ach / peak, "traffic": None, "peak_source": src,
            "algorithmic_flops_per_launch": flops, "avg_launch_us": t * 1e6,
            "timing": f"CUDA events around a graph of {reps} back-to-back launches, mean of 3 replays after 3 warm-up replays"}


def dominant_kernel_roofline(a, dims, hbm_peak, peak_src):
    """Time the dominant decode kernel (the tcgen05 W4A16 GEMM of the fused gate_up projection, SwiGLU
    epilogue) with CUDA events on its launch stream.  20 back-to-back launches are captured in a CUDA
    graph (no host launch latency inside the timed region) and walk over 10 different weight copies
    (138 MB > the 126 MB L2), so every launch streams its weights from HBM like it does inside a step."""
    import ctypes as C
    import torch
    from vgate_b200 import _lib
    lib = _lib.load()
    K, I, G, T = dims.hidden, dims.intermediate, dims.group_size, a.batch
    N = 2 * I
    rng = np.random.default_rng(0)
    ptr = lambda t: C.c_void_p(t.data_ptr())
    st = torch.cuda.Stream()
    sp = C.c_void_p(st.cuda_stream)
    use_umma = a.gemm_path != 1
    n_copies, reps = 10, 20
    sc = torch.full((N, K // G), 0.003, dtype=torch.float16, device="cuda")
    zp = torch.full((N, K // G), 1032.0, dtype=torch.float16, device="cuda")
    weights = []
    with torch.cuda.stream(st):
        for c in range(n_copies):
            wq = torch.from_numpy(rng.integers(0, 2**31, size=(N, K // 8), dtype=np.int32)).cuda()
            if use_umma:
                ch = torch.empty(lib.b200_op_w4_tile_bytes(K, N, G), dtype=torch.uint8, device="cuda")
                _lib.check(lib.b200_op_w4_tile(ptr(wq), ptr(sc), ptr(zp), K, N, G, I, ptr(ch), sp))
                weights.append(ch)
            else:
                weights.append(wq)
        st.synchronize()
    x = torch.from_numpy(rng.standard_normal((T, K)).astype(np.float16)).cuda()
    act = torch.empty((T, I), dtype=torch.float16, device="cuda")
    y = torch.empty((T, N), dtype=torch.float32, device="cuda")

    def launch(i):
        w = weights[i % n_copies]
        if use_umma:
            _lib.check(lib.b200_op_w4_gemm_umma(ptr(w), ptr(x), K, ptr(act), I, None, T, K, N, G, 2, 1, 0, sp))
        else:
            _lib.check(lib.b200_op_w4_gemv(ptr(w), ptr(sc), ptr(zp), ptr(x), K, ptr(y), N, None, T, K, N, G, 0, sp))

    times = []
    with torch.cuda.stream(st):
        launch(0)
        st.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=st):
            for i in range(reps):
                launch(i)
        for it in range(8):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(st)
            g.replay()
            e1.record(st)
            st.synchronize()
            if it >= 3:
                times.append(e0.elapsed_time(e1) * 1e-3 / reps)
    t = float(np.mean(times))
    # algorithmic bytes per launch: packed weights + scales + zeros + activations in + fp16 act out
    alg = N * K / 2 + 2 * N * (K // G) * 2 + T * K * 2 + (T * I * 2 if use_umma else T * N * 4)
    ach = alg / t / 1e9
    return {"bound": "hbm", "kernel": ("umma_gemm_kernel<64, W4> gate_up + SwiGLU" if use_umma else "w4_gemv_kernel gate_up")
            + f" [{T}x{K}]x[{N}x{K}]", "achieved": ach, "peak": hbm_peak, "unit": "GB/s", "frac": ach / hbm_peak,
            # dram__bytes_read.sum + dram__bytes_write.sum of this kernel at T=64 from one `ncu --set full` capture
            # (profiles/r02_kernels_ncu_full.txt, grid 140): 22.18 MB read + 0 written (the output stays in L2).  Of the
            # read, 7.3 MB are the down projection's weights, which this kernel's TMA warp pulls into L2 for its
            # successor (profiles/r01_gateup_ncu_full.txt, captured before that prefetch existed: 14.86 MB)
            "traffic": (22180000 if use_umma and T == 64 else None),
            "traffic_note": "includes 7.3 MB of L2 prefetch for the next kernel (down projection weights); own operands 14.9 MB",
            "traffic_source": "profiles/r02_kernels_ncu_full.txt", "peak_source": peak_src, "algorithmic_bytes_per_launch": alg,
            "avg_launch_us": t * 1e6,
            "l2": f"{n_copies} weight copies ({n_copies * N * K // 2 >> 20} MiB) walked round-robin: weights always come from HBM",
            "timing": f"CUDA events around a graph of {reps} back-to-back launches, mean of 5 replays after 3 warm-up replays"}


def gateway_levels(a, world):
    """tools/bench_gateway.py on this box: `world` b200 workers (CUDA_VISIBLE_DEVICES=i) behind one gateway."""
    import asyncio
    import importlib.util
    spec = importlib.util.spec_from_file_location("bench_gateway", os.path.join(ROOT, "tools", "bench_gateway.py"))
    bg = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(bg)
    ref = bg.find_reference()
    if ref is None:
        return {"unavailable": "no V-Gate checkout on this box (tools/bench_gateway.py --stage copies it to baseline/_ref)"}
    vis = os.environ.get("CUDA_VISIBLE_DEVICES")
    devs = vis.split(",") if vis else [str(i) for i in range(world)]
    # also with max(2, N) gateway server processes (tools/vgate_launch.py): one process is one event loop, which is
    # what binds behind the unmodified single-process gateway -- already at one GPU (140 of the worker's 178 req/s)
    out = asyncio.run(bg.run_b200(ref, os.path.join(ROOT, "gpurun_out", "gateway_logs"), world, a.batch, a.new_tokens, 4,
                                  a.preset, ["worker_http", "gateway"], devices=devs[:world],
                                  extra_gateway_procs=max(2, world)))
    return out


def run_b200(a):
    import torch
    import torch.distributed as dist
    from vgate_b200.engine import Engine
    from vgate_b200.weights import PRESETS, synthetic_checkpoint

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    ctl = None
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        ctl = dist.new_group(backend="gloo")      # host-side rendezvous: an NCCL barrier spins ON the GPU while it waits
    hbm_peak, peak_src = load_peaks()
    dims = PRESETS[a.preset]
    max_len = a.prompt_len + a.new_tokens + 16
    kw = dict(max_seqs=max(a.batch, 8), max_model_len=max_len, max_batch_tokens=max(8192, a.prompt_len),
              use_cuda_graphs=not a.no_graphs, gemm_path=a.gemm_path)
    drives = local % a.tp == 0                        # with --tp K only every K-th rank drives a replica (K GPUs each)
    eng = None
    if drives:
        ckpt = synthetic_checkpoint(dims, seed=0, profile=a.profile)
        if a.tp > 1:
            from vgate_b200.engine import TPEngine
            eng = TPEngine(dims, ckpt, devices=tuple(range(local, local + a.tp)), **kw)
        else:
            eng = Engine(dims, device=local, **kw).load_checkpoint(ckpt)
        del ckpt
    rng = np.random.default_rng(1000 + rank)          # unique prompts per rank and per step

    def one_pass(prompts):
        """tokens [batch, new_tokens], (prefill_s, decode_s) device-timed on the engine's stream"""
        if a.tp == 1:
            toks, _, secs = eng.generate_batch(prompts, a.new_tokens)
            return toks, secs
        s0 = eng.stats()                              # tensor-parallel engines take requests only: the scheduler path
        rids = [eng.submit(p, 0.0, 1.0, a.new_tokens) for p in prompts]
        toks = np.stack([eng.wait(r, a.new_tokens)[0] for r in rids])
        s1 = eng.stats()
        return toks, (s1["gpu_seconds_prefill"] - s0["gpu_seconds_prefill"], s1["gpu_seconds_decode"] - s0["gpu_seconds_decode"])

    def barrier():
        # host-side (gloo) rendezvous: an NCCL barrier is a kernel that SPINS on the waiting rank's GPU, and with --tp K the
        # ranks that drive no replica wait here for the whole timed region while their GPU serves another rank's engine
        # (measured: the 14B TP=2 step ran 2.3x slower under that spinning kernel)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier(group=ctl)
        torch.cuda.synchronize()

    for _ in range(a.warmup if drives else 0):
        one_pass(make_prompts(rng, a, dims.vocab))
    st0 = eng.stats() if drives else None
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    barrier()
    t0 = time.perf_counter()
    gpu_prefill = gpu_decode = 0.0
    checked = None                                   # prompts and tokens of the first timed step, for the self-check
    for i in range(a.steps if drives else 0):
        pr = make_prompts(rng, a, dims.vocab)
        toks, (tp, td) = one_pass(pr)
        if i == 0:
            checked = (pr[:4].copy(), toks[:4, :8].copy())
        gpu_prefill += tp
        gpu_decode += td
    barrier()
    wall = time.perf_counter() - t0
    clk = clocks.stop() if rank == 0 else None
    st1 = eng.stats() if drives else None

    # device-timed whole-step time (CUDA events on the engine's stream), max over ranks
    dev_time = gpu_prefill + gpu_decode
    tokens_per_rank = a.batch * a.new_tokens * a.steps if drives else 0
    cuda_dev = f"cuda:{local}"
    value, _ = aggregate_throughput(tokens_per_rank, dev_time, device=cuda_dev)
    dev_time, wall, gpu_decode_max, gpu_prefill_max = reduce_over_ranks([dev_time, wall, gpu_decode, gpu_prefill], "max", cuda_dev)

    # ---- e2e: the public API, called the way V-Gate calls it ----
    # B200Backend.generate([one prompt], sampling_params) from one thread per in-flight request
    # (vgate/worker_api.py:112, vgate/batcher.py:422): host strings in, host dicts out; tokenisation,
    # H2D of prompt ids, continuous batching, D2H of every step's sampled ids are all inside the timing.
    e2e = None
    if not a.no_e2e and not drives:                   # a rank whose GPU belongs to another rank's replica: collectives only
        barrier(); barrier()
        aggregate_throughput(0, 0.0, device=f"cuda:{local}")
    if not a.no_e2e and drives:
        from concurrent.futures import ThreadPoolExecutor
        from vgate_b200.backend import B200Backend
        from vgate_b200.tokenizer import ByteTokenizer
        be = B200Backend()
        be.engine, be.tokenizer = eng, ByteTokenizer()
        sp = be.create_sampling_params(0.0, 1.0, a.new_tokens)
        pool = ThreadPoolExecutor(a.batch)

        def e2e_round():
            codes = rng.integers(32, 127, size=(a.batch, a.prompt_len), dtype=np.uint8)
            prompts = [bytes(row).decode("ascii") for row in codes]      # 1 byte = 1 token
            t_a = time.perf_counter()
            results = list(pool.map(lambda pr: be.generate([pr], sp)[0], prompts))
            return time.perf_counter() - t_a, results

        e2e_round()                                       # warm
        barrier()
        tot_t, tot_n, tpots = 0.0, 0, []
        for _ in range(max(1, a.steps)):
            dt, results = e2e_round()
            tot_t += dt
            tot_n += sum(r["num_tokens"] for r in results)
            tpots += [r["metrics"]["gen_time"] / r["num_tokens"] for r in results]   # vgate/batcher.py:431
        barrier()
        pool.shutdown()
        e2e_value, _ = aggregate_throughput(tot_n, tot_t, device=f"cuda:{local}")
        e2e = {"value": e2e_value, "unit": UNIT,
               "h2d_bytes_per_step": int(a.batch * a.prompt_len * 4),       # prompt token ids
               "d2h_bytes_per_step": int(a.batch * a.new_tokens * 4),       # sampled ids, read back every decode step
               "tpot_ms_p50": float(np.median(tpots) * 1e3),
               "path": f"B200Backend.generate([prompt], sp) x {a.batch} concurrent threads (one prompt per call, as "
                       "vgate/worker_api.py:112 issues them); continuous batching inside the engine"}

    checks = {}
    if rank == 0 and not a.no_cpu_baseline and world == 1:
        # what was timed is also checked: 4 sequences x 8 tokens of the first timed step against the fp32
        # oracle (the same object the CPU baseline times) -- tokens identical, logits within 2e-2
        ref = CpuReference(a.preset, a.profile, a.batch, a.prompt_len, a.new_tokens)
        pr4, tok4 = checked
        if ref.big:       # an fp32 oracle of a 7B / 14B model is 30-60 GB and minutes to build: parity at these shapes
            checks["self_check"] = {"skipped": "fp32 oracle too large; tests/test_dense_gpu.py checks these layer shapes at reduced depth"}
            checks["cpu_baseline"], _, _ = ref.run(steps=1)
            del ref
            ref = None
    if checks.get("cpu_baseline") is None and rank == 0 and not a.no_cpu_baseline and world == 1:
        again, logits4, _ = eng.generate_batch(pr4, 8, n_logit_steps=8)
        ref_toks, ref_logits = ref.fp32().generate_greedy(pr4, 8, return_logits=True)
        err = float(np.abs(logits4 - np.asarray(ref_logits)).max()) if np.array_equal(again, ref_toks) else float("nan")
        checks["self_check"] = {"sequences": 4, "tokens": 8, "timed_step_tokens_equal_oracle": bool(np.array_equal(tok4, ref_toks)),
                              "batch_invariant": bool(np.array_equal(again, tok4)), "logits_max_abs_err": err}
        assert np.array_equal(tok4, ref_toks), "bench self-check: timed tokens differ from the fp32 oracle"
        assert np.array_equal(again, tok4), "bench self-check: tokens depend on the batch"
        assert err < 2e-2, f"bench self-check: logits differ from the fp32 oracle by {err}"
        checks["cpu_baseline"], _, _ = ref.run(steps=1)
        del ref
    # ---- the same workload behind V-Gate: worker HTTP level and ONE gateway in front of `world` workers ----
    # (north_star: "one worker replica per GPU ... behind one V-Gate gateway").  Unmodified reference processes,
    # spawned by rank 0 once every rank has released its GPU; skipped when no reference checkout travelled here
    # (baseline/_ref, git-ignored) or with --no-gateway.
    levels = None
    if a.tp > 1:
        a.no_gateway = True                           # the gateway levels are measured on the replica-per-GPU topology
    if not a.no_gateway:
        eng.close()
        eng = None
        torch.cuda.synchronize()
        if ctl is not None:                             # the other ranks wait on the host: their GPUs belong to the workers now
            dist.barrier(group=ctl)
        if rank == 0:
            try:
                levels = gateway_levels(a, world)
            except Exception as exc:                    # reported, never fatal for the engine-level line
                levels = {"error": repr(exc)[:300]}
        if ctl is not None:
            dist.barrier(group=ctl)

    if rank == 0:
        d_steps = st1["steps_decode"] - st0["steps_decode"]
        d_bytes = st1["hbm_bytes_decode"] - st0["hbm_bytes_decode"]
        step_ach = d_bytes / gpu_decode_max / 1e9 if gpu_decode_max > 0 else 0.0
        roof = dense_kernel_roofline(a, dims) if dims.quant == "none" else dominant_kernel_roofline(a, dims, hbm_peak, peak_src)
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
            "ms_per_step": dev_time / a.steps * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": ("f16 (int4 weights, fp16 activations, fp32 accumulate)" if dims.quant == "awq" else
                      f"{dims.dtype} (dense {dims.dtype} weights and activations, fp32 accumulate)"), "data": "synthetic",
            "config": config_of(a, world), "cuda_graphs": not a.no_graphs, "gemm_path": a.gemm_path,
            "replicas_x_tp": f"{world // a.tp} x {a.tp}",
            "wall_ms_per_step": wall / a.steps * 1e3,
            "decode": {"tpot_ms_mean": gpu_decode_max / d_steps * 1e3 if d_steps else None,
                       "steps": d_steps, "tok_per_s": (world // a.tp) * a.batch * d_steps / gpu_decode_max if gpu_decode_max else None,
                       "hbm_gbs": step_ach, "hbm_roofline_frac": step_ach / hbm_peak},
            "prefill": {"tok_per_s": (world // a.tp) * a.batch * a.prompt_len * a.steps / gpu_prefill_max if gpu_prefill_max else None},
            "roofline": roof,
            "gpu_launches": int(st1["kernel_launches"] - st0["kernel_launches"]),
            "clocks": clk,
        }
        if e2e is not None:
            line["e2e"] = e2e
        if levels is not None:
            if "error" in levels or "unavailable" in levels:
                line["gateway"] = levels
            else:
                line["worker_http"] = levels["worker_http_multi"]
                line["worker_http_single_prompt_bodies"] = levels["worker_http_single"]
                line["gateway"] = levels["gateway"]
                if "gateway_multi_process" in levels:
                    line["gateway_multi_process"] = levels["gateway_multi_process"]
                lv = {"engine_c_abi": value, "backend_threads": e2e["value"] if e2e else None,
                      "worker_http": levels["worker_http_multi"]["tok_per_s"], "gateway": levels["gateway"].get("tok_per_s")}
                if "gateway_multi_process" in levels:
                    lv[f"gateway_{levels['gateway_multi_process']['processes']}_processes"] = levels["gateway_multi_process"].get("tok_per_s")
                line["levels_tok_per_s"] = lv
                gw_best = max([v for k, v in lv.items() if k.startswith("gateway") and v] or [0.0])
                line["binding_level"] = ("gateway (CPU control plane: one event loop per process, one thread and one of 100 hop "
                                         "connections per in-flight request)"
                                         if gw_best < 0.9 * lv["worker_http"] else
                                         "worker HTTP" if lv["worker_http"] < 0.9 * lv["engine_c_abi"] else "engine")
        line.update(checks)
        # informational: the reference's own GPU path (vLLM) measured once on this pool, same checkpoint and
        # workload, by tools/vllm_b200_baseline.py -- a recorded number, not re-measured by this run
        try:
            with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", "r01_vllm_b200_baseline.json")) as f:
                v = json.load(f)
            if v.get("batch") == a.batch and v.get("new_tokens") == a.new_tokens and a.preset == v.get("preset"):
                line["reference_gpu_recorded"] = {"impl": "vllm " + v["vllm_version"], "value": v["tok_per_s"], "unit": "tok/s",
                                                  "n_gpus": 1, "source": "profiles/r01_vllm_b200_baseline.json"}
        except Exception:   # noqa: BLE001
            pass
        print(json.dumps(line), flush=True)
    if eng is not None:
        eng.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)
