"""Target for `ncu --set full`: one prefill pass (64 x 200 tokens) and two decode steps of a 2-layer
model with the real Qwen2.5-1.5B layer shapes and the full 151 936-entry vocabulary, eager launches
(no CUDA graph), so that every kernel of the path appears once or twice in its production shape."""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from vgate_b200.engine import Engine
from vgate_b200.weights import ModelDims, synthetic_checkpoint

dims = ModelDims(n_layers=2)
eng = Engine(dims, max_seqs=64, max_model_len=512, max_batch_tokens=12800, use_cuda_graphs=False)
eng.load_checkpoint(synthetic_checkpoint(dims, seed=0))
rng = np.random.default_rng(1)
prompts = rng.integers(0, dims.vocab, size=(64, 200), dtype=np.int32)
toks, _, secs = eng.generate_batch(prompts, 3)
print("ok", toks[:2].tolist(), secs)
eng.close()
