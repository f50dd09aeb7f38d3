"""Small target for `compute-sanitizer --tool racecheck|synccheck|memcheck` (ONE tool per gpurun call): every
kernel of the path once or twice at small shapes -- a 2-layer model with narrow layers (hidden 256), prefill of
ragged prompts, decode steps through the scheduler (greedy and sampled rows), a batch above 128 rows (the
256-row GEMM CTAs and the three-warp attention) -- eager launches, no CUDA graph.  Checks nothing itself: the
sanitizer's report is the result."""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from vgate_b200.engine import Engine
from vgate_b200.weights import ModelDims, synthetic_checkpoint

dims = ModelDims(hidden=256, intermediate=512, n_layers=2, n_heads=12, n_kv_heads=2, head_dim=128, vocab=2048)
eng = Engine(dims, max_seqs=160, max_model_len=160, max_batch_tokens=1024, use_cuda_graphs=False)
eng.load_checkpoint(synthetic_checkpoint(dims, seed=0))
rng = np.random.default_rng(1)
toks, logits, _ = eng.generate_batch(rng.integers(0, dims.vocab, size=(5, 70), dtype=np.int32), 4, n_logit_steps=2)
big, _, _ = eng.generate_batch(rng.integers(0, dims.vocab, size=(150, 3), dtype=np.int32), 3)
rids = [eng.submit(rng.integers(0, dims.vocab, size=int(n), dtype=np.int32), temperature=t, top_p=0.9, max_tokens=5, seed=i)
        for i, (n, t) in enumerate([(1, 0.0), (17, 0.8), (66, 0.0), (30, 0.7)])]
outs = [eng.wait(r, 5)[0].tolist() for r in rids]
print("ok", toks[0].tolist(), big[0].tolist(), outs[1])
eng.close()
