"""In-situ stamps of the fused decode MLP kernel (b200_debug_mlp_trace).  usage: python tools/mlp_trace.py [batch]"""
import ctypes as C
import os
import sys

import numpy as np
import torch

sys.path.insert(0, ".")
os.environ["VGATE_B200_FUSED_MLP"] = "1"
from vgate_b200 import _lib
from vgate_b200.engine import Engine
from vgate_b200.weights import PRESETS, synthetic_checkpoint

B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
dims = PRESETS["qwen2.5-1.5b-awq-l4"]
lib = _lib.load()
rng = np.random.default_rng(1)
prompts = rng.integers(0, dims.vocab, size=(B, 250), dtype=np.int32)
for graphs in (False, True):
    eng = Engine(dims, max_seqs=64, max_model_len=512, max_batch_tokens=16384, use_cuda_graphs=graphs)
    eng.load_checkpoint(synthetic_checkpoint(dims, seed=0))
    buf = torch.zeros(148 * 8, dtype=torch.int64, device="cuda")
    lib.b200_debug_mlp_trace(C.c_void_p(buf.data_ptr()))
    eng.generate_batch(prompts, 8)
    lib.b200_debug_mlp_trace(None)
    eng.close()
    t = buf.cpu().numpy().reshape(148, 8).astype(np.float64) / 1.9e3
    gu = t[:140]
    dn = t[4:148]
    print(f"cuda graphs {graphs}: medians in us (SM clocks / 1.9 GHz), last launch")
    print(f"  entry -> dependency wait returned            {np.median(gu[:,5]-gu[:,0]):6.2f}")
    print(f"  wait returned -> gate_up accumulator done    {np.median(gu[:,1]-gu[:,5]):6.2f}")
    print(f"  -> gate_up columns published                 {np.median(gu[:,2]-gu[:,1]):6.2f}")
    print(f"  published -> first down flags seen           {np.median(dn[:,6]-np.where(dn[:,2]>0, dn[:,2], dn[:,5])):6.2f}")
    print(f"  flags seen -> down accumulator done          {np.median(dn[:,3]-dn[:,6]):6.2f}")
    print(f"  -> down epilogue done                        {np.median(dn[:,4]-dn[:,3]):6.2f}")
    t0 = t[:, 0].min()
    print(f"  absolute (from first CTA entry): entry p50/max {np.median(t[:,0]-t0):.2f}/{(t[:,0]-t0).max():.2f}  wait returned p50/max "
          f"{np.median(t[:,5]-t0):.2f}/{(t[:,5]-t0).max():.2f}  gate_up published p50/max {np.median(gu[:,2]-t0):.2f}/{(gu[:,2]-t0).max():.2f}  "
          f"down flags seen p50/max {np.median(dn[:,6]-t0):.2f}/{(dn[:,6]-t0).max():.2f}  end p50/max {np.median(dn[:,4]-t0):.2f}/{(dn[:,4]-t0).max():.2f}")
    print(f"  entry -> end                                 {np.median(dn[:,4]-dn[:,0]):6.2f}   wait returned -> end {np.median(dn[:,4]-dn[:,5]):6.2f}")
