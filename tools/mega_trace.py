"""Phase timeline of the persistent decode kernel (b200_debug_mega_trace) and a bitwise comparison
with the per-kernel path.  usage: python tools/mega_trace.py [preset] [batch] [prompt_len]"""
import ctypes as C
import os
import sys

import numpy as np
import torch

sys.path.insert(0, ".")
from vgate_b200 import _lib
from vgate_b200.engine import Engine
from vgate_b200.weights import PRESETS, synthetic_checkpoint

preset = sys.argv[1] if len(sys.argv) > 1 else "qwen2.5-1.5b-awq-l4"
B = int(sys.argv[2]) if len(sys.argv) > 2 else 64
PL = int(sys.argv[3]) if len(sys.argv) > 3 else 200
dims = PRESETS[preset]
ckpt = synthetic_checkpoint(dims, seed=0)
rng = np.random.default_rng(7)
prompts = rng.integers(0, dims.vocab, size=(B, PL), dtype=np.int32)
lib = _lib.load()
NL, NP, G = dims.n_layers, 7, 148
names = ["rms1", "qkv", "attn", "o", "rms2", "gate_up", "down"]


def run(mega, trace=None, n_new=8, n_logits=4):
    os.environ["VGATE_B200_MEGA"] = "1" if mega else "0"
    eng = Engine(dims, max_seqs=64, max_model_len=512, max_batch_tokens=4096, use_cuda_graphs=False).load_checkpoint(ckpt)
    lib.b200_debug_mega_trace(C.c_void_p(trace.data_ptr()) if trace is not None else None)
    try:
        toks, logits, secs = eng.generate_batch(prompts, n_new, n_logit_steps=n_logits)
    finally:
        lib.b200_debug_mega_trace(None)
        eng.close()
    return toks, logits, secs


trace = torch.zeros(G * NL * NP * 2, dtype=torch.int64, device="cuda")
t1, l1, s1 = run(True, trace)
tr = trace.cpu().numpy().reshape(G, NL * NP, 2).astype(np.float64)
t1b, l1b, _ = run(True)
t0, l0, s0 = run(False)
print(f"decode seconds for 7 steps: mega {s1[1]:.6f}  per-kernel {s0[1]:.6f}")
d = l0.view(np.uint32) != l1.view(np.uint32)
print(f"mega vs per-kernel: {int(d.sum())} of {d.size} logits differ, max abs {np.abs(l0 - l1).max():.3e}; "
      f"rows with a difference {np.unique(np.nonzero(d)[0])[:20]}, steps {np.unique(np.nonzero(d)[1])}")
d2 = l1.view(np.uint32) != l1b.view(np.uint32)
print(f"mega run-to-run: {int(d2.sum())} differ; tokens equal to per-kernel: {np.array_equal(t0, t1)}")

# last traced step: arrival times per phase
arr = tr[:, :, 1]
start = tr[:, :, 0]
t_end = arr.max(axis=0)                        # phase complete = last arrival
base = t_end[0]
print("phase        layer0 dur   mean dur (us)   last-arriver CTA (layer 1)    work-start -> arrive of the slowest CTA")
for p in range(NP):
    durs = []
    for l in range(NL):
        i = l * NP + p
        prev = t_end[i - 1] if i > 0 else np.nan
        durs.append((t_end[i] - prev) / 1e3)
    i1 = (1 if NL > 1 else 0) * NP + p
    slow = int(arr[:, i1].argmax())
    ws = start[slow, i1]
    print(f"{names[p]:8s}  {durs[0]:10.2f}   {np.nanmean(durs[1:] if NL > 1 else durs):10.2f}      cta {slow:3d}"
          f"            {(arr[slow, i1] - ws) / 1e3 if ws > 0 else float('nan'):8.2f}")
print(f"layer time (mean): {(t_end[NP * NL - 1] - t_end[NP - 1]) / 1e3 / max(1, NL - 1):.2f} us")
