"""In-situ timeline of one decode step of a tensor-parallel engine (rank 0's kernels; NCCL kernels show up as the
gaps between them).  A 4-layer model with the Qwen2.5-14B layer shapes.
usage: python tools/tp_timeline.py [batch] [prompt_len] [tp] [layers]"""
import ctypes as C
import os
import sys
import time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from vgate_b200 import _lib
from vgate_b200.engine import Engine, TPEngine
from vgate_b200.weights import ModelDims, synthetic_checkpoint

NAMES = {1: "embed", 2: "rmsnorm", 3: "rope", 4: "attn", 5: "swiglu", 6: "w4_gemv", 7: "f16_gemv", 8: "umma_w4",
         9: "umma_dense", 10: "sample", 11: "advance", 12: "tp_tick", 13: " tp phases"}


def main():
    batch = int(sys.argv[1]) if len(sys.argv) > 1 else 64
    plen = int(sys.argv[2]) if len(sys.argv) > 2 else 2048
    tp = int(sys.argv[3]) if len(sys.argv) > 3 else 2
    layers = int(sys.argv[4]) if len(sys.argv) > 4 else 4
    dims = ModelDims(hidden=5120, intermediate=13824, n_layers=layers, n_heads=40, n_kv_heads=8, vocab=152064,
                     tie_embeddings=False, quant="none", dtype="bf16")
    ckpt = synthetic_checkpoint(dims, seed=0, profile="peaked")
    kw = dict(max_seqs=batch, max_model_len=plen + 80, max_batch_tokens=8192)
    os.environ.setdefault("VGATE_B200_PREFILL_CHUNK", "8192")
    os.environ.setdefault("VGATE_B200_MAX_STALL_MS", "1000000")
    eng = TPEngine(dims, ckpt, devices=tuple(range(tp)), **kw) if tp > 1 else Engine(dims, **kw).load_checkpoint(ckpt)
    lib = _lib.load()
    rng = np.random.default_rng(0)
    def run(n_new):
        prompts = rng.integers(0, dims.vocab, size=(batch, plen), dtype=np.int32)   # fresh prompts: no prefix-cache hits
        s0 = eng.stats()
        t0 = time.perf_counter()
        rids = [eng.submit(p, 0.0, 1.0, n_new) for p in prompts]
        for r in rids:
            eng.wait(r, n_new)
        s1 = eng.stats()
        return {k: s1[k] - s0[k] for k in ("gpu_seconds_prefill", "gpu_seconds_decode", "steps_decode", "steps_prefill")}, time.perf_counter() - t0

    print("warm", run(8))
    tl_dev = int(os.environ.get("TL_DEV", "0"))              # which rank's kernels to record
    torch.cuda.set_device(tl_dev)
    cap = 40000
    buf = torch.zeros((cap, 4), dtype=torch.int64, device=f"cuda:{tl_dev}")
    ctr = torch.zeros(1, dtype=torch.int32, device=f"cuda:{tl_dev}")
    lib.b200_debug_timeline(C.c_void_p(buf.data_ptr()), C.c_void_p(ctr.data_ptr()), cap)
    st, wall = run(24)
    torch.cuda.synchronize()
    lib.b200_debug_timeline(None, None, 0)
    print(f"tp {tp} layers {layers} batch {batch} prompt {plen}: prefill {st['gpu_seconds_prefill'] * 1e3:.1f} ms in {st['steps_prefill']} passes, "
          f"decode {st['gpu_seconds_decode'] / max(1, st['steps_decode']) * 1e3:.3f} ms per step over {st['steps_decode']} steps, wall {wall:.2f} s")
    n = int(ctr.item())
    rec = buf[:min(n, cap)].cpu().numpy()
    rec = rec[np.argsort(rec[:, 1])]
    kid = rec[:, 0] & 0xFF
    adv = np.where(kid == 11)[0]
    if len(adv) >= 3:
        a, b = adv[-3], adv[-2]
        step = rec[a:b]
        t0 = step[0, 1]
        print(f"# decode step: {b - a} kernels of rank {tl_dev}, {(rec[b, 1] - t0) / 1e3:.1f} us from advance to advance")
        prev_exit = t0
        for r, k in list(zip(step, kid[a:b]))[:30]:
            gx, gy, gz = (r[0] >> 8) & 0xFFFFF, (r[0] >> 28) & 0xFFFFF, (r[0] >> 48) & 0xFFFF
            print(f"  {NAMES.get(int(k), str(k)):10s} grid {gx}x{gy}x{gz:<3d} entry {(r[1] - t0) / 1e3:9.2f} ready {(r[2] - t0) / 1e3:9.2f} exit {(r[3] - t0) / 1e3:9.2f}   gap before ready {(r[2] - prev_exit) / 1e3:8.2f}")
            prev_exit = max(prev_exit, r[3])
    eng.close()


if __name__ == "__main__":
    main()
