import ctypes as C, os, sys
sys.path.insert(0, '/root/repo')
import numpy as np, torch
from vgate_b200 import _lib
from vgate_b200.engine import Engine
from vgate_b200.weights import PRESETS, synthetic_checkpoint
dims = PRESETS["qwen2.5-1.5b-awq"]
eng = Engine(dims, max_seqs=64, max_model_len=300).load_checkpoint(synthetic_checkpoint(dims, seed=0))
lib = _lib.load()
rng = np.random.default_rng(0)
prompts = rng.integers(0, dims.vocab, size=(64, 128), dtype=np.int32)
eng.generate_batch(prompts, 8)
tr = torch.zeros((2*64, 8), dtype=torch.int64, device="cuda")
lib.b200_debug_attn_trace(C.c_void_p(tr.data_ptr()))
eng.generate_batch(prompts, 130)
torch.cuda.synchronize()
lib.b200_debug_attn_trace(None)
t = tr.cpu().numpy().astype(np.float64)
d = t[:, 1:7] - t[:, 0:1]
names = ['copies issued', 'wait start', 'tile landed', 'scores done', 'softmax done', 'loop end']
print('in-situ attention (last layer, ctx ~257), median cycles since CTA entry: ' + ', '.join(f'{n} {np.median(d[:, i]):.0f}' for i, n in enumerate(names)))
print('  p90: ' + ', '.join(f'{n} {np.percentile(d[:, i], 90):.0f}' for i, n in enumerate(names)))
span = (t[:, 6].max() - t[:, 0].min())
print('  (entry spread %.0f cycles; first entry -> last loop end %.0f cycles; note clocks are per-SM)' % (t[:, 0].max() - t[:, 0].min(), span))
eng.close()
