"""Decode step time with the reference's default sampling parameters (temperature 0.7, top_p 0.9:
vgate/config.py:138-140) against greedy, batch 64.  usage: python tools/bench_sampled.py"""
import sys
import threading

import numpy as np

sys.path.insert(0, ".")
from vgate_b200.engine import Engine
from vgate_b200.weights import PRESETS, synthetic_checkpoint

dims = PRESETS["qwen2.5-1.5b-awq"]
eng = Engine(dims, max_seqs=64, max_model_len=512, max_batch_tokens=8192).load_checkpoint(synthetic_checkpoint(dims, seed=0))
rng = np.random.default_rng(1)
prompts = rng.integers(0, dims.vocab, size=(64, 128), dtype=np.int32)
for name, temp, top_p in (("greedy", 0.0, 1.0), ("temperature 0.7 top_p 0.9", 0.7, 0.9), ("temperature 1.0 top_p 1.0", 1.0, 1.0)):
    for rep in range(2):
        s0 = eng.stats()
        rids = [eng.submit(p, temp, top_p, 128, seed=i) for i, p in enumerate(prompts)]
        for r in rids:
            eng.wait(r, 128)
        s1 = eng.stats()
    steps = s1["steps_decode"] - s0["steps_decode"]
    print(f"{name:28s} decode {1e3 * (s1['gpu_seconds_decode'] - s0['gpu_seconds_decode']) / steps:.4f} ms/step over {steps} steps, "
          f"mean batch {(s1['decode_batch_sum'] - s0['decode_batch_sum']) / steps:.1f}")
eng.close()
