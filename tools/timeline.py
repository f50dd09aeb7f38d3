"""In-situ timeline of one decode step (CUDA-graph replay, PDL on): per-kernel ready->exit durations,
dependency-wait times and gaps, from %globaltimer stamps written by block 0 of every kernel
(b200_debug_timeline).  Usage: python tools/timeline.py [batch] [prompt_len] [new_tokens]"""
import ctypes as C
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from vgate_b200 import _lib
from vgate_b200.engine import Engine
from vgate_b200.weights import PRESETS, synthetic_checkpoint

NAMES = {1: "embed", 2: "rmsnorm", 3: "rope", 4: "attn", 5: "swiglu", 6: "w4_gemv", 7: "f16_gemv", 8: "umma_w4",
         9: "umma_f16", 10: "sample", 11: "advance"}


def main():
    batch = int(sys.argv[1]) if len(sys.argv) > 1 else 64
    plen = int(sys.argv[2]) if len(sys.argv) > 2 else 128
    new = int(sys.argv[3]) if len(sys.argv) > 3 else 130
    preset = os.environ.get("PRESET", "qwen2.5-1.5b-awq")
    dims = PRESETS[preset]
    eng = Engine(dims, max_seqs=max(batch, 8), max_model_len=plen + new + 16).load_checkpoint(synthetic_checkpoint(dims, seed=0))
    lib = _lib.load()
    rng = np.random.default_rng(0)
    prompts = rng.integers(0, dims.vocab, size=(batch, plen), dtype=np.int32)
    eng.generate_batch(prompts, 8)                                   # warm up + capture the graph
    cap = 400 * (new + 4)
    buf = torch.zeros((cap, 4), dtype=torch.int64, device="cuda")
    ctr = torch.zeros(1, dtype=torch.int32, device="cuda")
    lib.b200_debug_timeline(C.c_void_p(buf.data_ptr()), C.c_void_p(ctr.data_ptr()), cap)
    eng.generate_batch(prompts, new)
    torch.cuda.synchronize()
    lib.b200_debug_timeline(None, None, 0)
    n = int(ctr.item())
    rec = buf[:min(n, cap)].cpu().numpy()
    rec = rec[np.argsort(rec[:, 1])]                                 # order by entry time
    kid = rec[:, 0] & 0xFF
    gx, gy, gz = (rec[:, 0] >> 8) & 0xFFFFF, (rec[:, 0] >> 28) & 0xFFFFF, (rec[:, 0] >> 48) & 0xFFFF
    adv = np.where(kid == 11)[0]
    # last complete decode step: between the last two advance kernels
    a, b = adv[-3], adv[-2]
    step = rec[a:b]
    t0 = step[0, 1]
    print(f"# {preset} batch {batch}, decode step at context ~{plen + new - 2}: {b - a} kernels, "
          f"{(rec[b, 1] - t0) / 1e3:.1f} us from advance to advance")
    print("# kernel        grid            entry    ready     exit   (us from step start)   run = exit-ready, wait = ready-entry")
    agg = {}
    prev_exit = t0
    for r, k, x, y, z in zip(step, kid[a:b], gx[a:b], gy[a:b], gz[a:b]):
        run, wait = (r[3] - r[2]) / 1e3, (r[2] - r[1]) / 1e3
        name = NAMES.get(int(k), str(k)) + (f"[{x}x{y}x{z}]" if k in (8, 9) else "")
        e = agg.setdefault(name, [0, 0.0, 0.0, 0.0])
        e[0] += 1; e[1] += run; e[2] += wait; e[3] += max(0.0, (r[2] - prev_exit) / 1e3)
        prev_exit = max(prev_exit, r[3])
    tot = (rec[b, 1] - t0) / 1e3
    print(f"{'kernel':28s} {'n':>4s} {'run us':>9s} {'avg':>7s} {'%step':>6s} {'wait us':>9s} {'idle-before us':>14s}")
    for name, (c, run, wait, idle) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{name:28s} {c:4d} {run:9.1f} {run / c:7.2f} {100 * run / tot:6.1f} {wait:9.1f} {idle:14.1f}")
    print("# first layer in detail (us from step start: entry ready exit):")
    for r, k, x, y, z in list(zip(step, kid[a:b], gx[a:b], gy[a:b], gz[a:b]))[:12]:
        print(f"  {NAMES.get(int(k), k):10s} grid {x}x{y}x{z:<3d} {(r[1] - t0) / 1e3:8.2f} {(r[2] - t0) / 1e3:8.2f} {(r[3] - t0) / 1e3:8.2f}")
    eng.close()


if __name__ == "__main__":
    main()
