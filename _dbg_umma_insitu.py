import ctypes as C, os, sys
sys.path.insert(0, '/root/repo')
import numpy as np, torch
from vgate_b200 import _lib
from vgate_b200.engine import Engine
from vgate_b200.weights import PRESETS, synthetic_checkpoint
dims = PRESETS["qwen2.5-1.5b-awq"]
eng = Engine(dims, max_seqs=64, max_model_len=300, use_cuda_graphs=False).load_checkpoint(synthetic_checkpoint(dims, seed=0))
lib = _lib.load()
rng = np.random.default_rng(0)
prompts = rng.integers(0, dims.vocab, size=(64, 128), dtype=np.int32)
eng.generate_batch(prompts, 8)
for name, gx, gz, nk in (("qkv", 16, 4, 6), ("o", 12, 4, 6), ("gate_up", 140, 1, 24), ("down", 12, 12, 12)):
    tr = torch.zeros((gx * gz, 8), dtype=torch.int64, device="cuda")
    lib.b200_debug_umma_trace_grid(C.c_void_p(tr.data_ptr()), gx, gz)
    eng.generate_batch(prompts, 40)
    torch.cuda.synchronize()
    lib.b200_debug_umma_trace_grid(None, 0, 0)
    t = tr.cpu().numpy().astype(np.float64)
    d = t[:, 1:7] - t[:, 0:1]
    v = t[:, 7].astype(np.int64)
    print(f'{name:8s} ctas={gx*gz:4d}: setup {np.median(d[:,0]):.0f}, tmem loaded {np.median(d[:,1]):.0f}, last mma issued {np.median(d[:,2]):.0f}, '
          f'acc ready {np.median(d[:,3]):.0f}, epilogue done {np.median(d[:,4]):.0f}, exit {np.median(d[:,5]):.0f} (max {d[:,5].max():.0f}) cycles; '
          f'mma thread per k-tile: wait {np.median(v & 0xFFFFFFFF)/nk:.0f} issue {np.median(v >> 32)/nk:.0f}')
eng.close()
