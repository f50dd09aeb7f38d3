import sys, os; sys.path.insert(0, '/root/repo')
import numpy as np
from vgate_b200.engine import Engine
from vgate_b200.weights import PRESETS, synthetic_checkpoint
graphs = os.environ.get("G", "1") == "1"
dims = PRESETS["qwen2.5-1.5b-awq-l2-v8k"]
ckpt = synthetic_checkpoint(dims, seed=0)
eng = Engine(dims, max_seqs=16, max_model_len=512, max_batch_tokens=256, use_cuda_graphs=graphs, gemm_path=int(os.environ.get("GP","0"))).load_checkpoint(ckpt)
rng = np.random.default_rng(1)
prompts = rng.integers(0, dims.vocab, size=(6, 40), dtype=np.int32)
ref = None; nbad = 0
for rep in range(40):
    toks, logits, _ = eng.generate_batch(prompts, 12, n_logit_steps=12)
    if ref is None: ref = (toks.copy(), logits.copy())
    else:
        dt = (toks != ref[0]).sum(); dl = np.abs(logits - ref[1]).max()
        if dt or dl > 0:
            nbad += 1
            bad_steps = sorted(set(np.where(np.abs(logits-ref[1]).max(axis=(0,2))>0)[0].tolist()))
            print('rep', rep, 'token diffs', dt, 'max logit diff', dl, 'steps', bad_steps[:6])
print('PDL', os.environ.get('VGATE_B200_PDL','1'), 'graphs', graphs, 'nondeterministic reps:', nbad, 'of 39')
eng.close()
