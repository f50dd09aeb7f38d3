"""Target for `ncu --set full`: one prefill pass (64 x 200 tokens) and two decode steps of a 2-layer
model with the real Qwen2.5-1.5B layer shapes and the full 151 936-entry vocabulary, eager launches
(no CUDA graph), so that every kernel of the path appears once or twice in its production shape.
PRESET=<name> N_SEQS=.. PROMPT=..: another model (e.g. qwen2.5-7b-bf16-l2, 32 sequences x 256 tokens: the dense path)."""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from vgate_b200.engine import Engine
from vgate_b200.weights import PRESETS, ModelDims, synthetic_checkpoint

dims = PRESETS[os.environ["PRESET"]] if os.environ.get("PRESET") else ModelDims(n_layers=2)
n_seqs, n_prompt = int(os.environ.get("N_SEQS", "64")), int(os.environ.get("PROMPT", "200"))
eng = Engine(dims, max_seqs=n_seqs, max_model_len=n_prompt + 312, max_batch_tokens=12800, use_cuda_graphs=False)
eng.load_checkpoint(synthetic_checkpoint(dims, seed=0))
rng = np.random.default_rng(1)
prompts = rng.integers(0, dims.vocab, size=(n_seqs, n_prompt), dtype=np.int32)
toks, _, secs = eng.generate_batch(prompts, 3)
print("ok", toks[:2].tolist(), secs)
eng.close()
