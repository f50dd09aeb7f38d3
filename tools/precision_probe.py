"""Where does the engine's logit error come from?  Engine (tcgen05 GEMM family and CUDA-core GEMV family) against
the fp32 oracle and against the oracle's fp16 mode (the same rounding points, emulated on the CPU), plus the
kernel-level error of the W4 GEMM against an fp64 reference.  Diagnostic, not a test.
usage: python tools/precision_probe.py [profile=flat] [n_prompts=4] [prompt_len=128]"""
import ctypes as C
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from oracle.awq import dequantize_awq
from oracle.qwen2 import OracleQwen2, Qwen2Dims
from vgate_b200 import _lib
from vgate_b200.engine import Engine
from vgate_b200.weights import PRESETS, synthetic_checkpoint


def stats(name, a, b):
    d = np.abs(np.asarray(a, np.float64) - np.asarray(b, np.float64))
    print(f"{name:58s} max {d.max():.3e}  rms {np.sqrt((d ** 2).mean()):.3e}", flush=True)


def gemm_probe():
    from tests.gpu_util import lib, dev, ptr, stream, check, repack, random_awq
    for K, N, T in [(1536, 2048, 64), (8960, 1536, 64), (1536, 17920, 64)]:
        rng = np.random.default_rng(K + N)
        qw, qz, sc = random_awq(rng, K, N)
        wq, s, z = repack(qw, qz, sc)
        nbytes = lib().b200_op_w4_tile_bytes(K, N, 128)
        ch = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
        check(lib().b200_op_w4_tile(ptr(wq), ptr(s), ptr(z), K, N, 128, 0, ptr(ch), stream()))
        x = rng.standard_normal((T, K)).astype(np.float16)
        w = dequantize_awq(qw, qz, sc, 128).astype(np.float64)
        ref = x.astype(np.float64) @ w
        y = torch.empty((T, N), dtype=torch.float32, device="cuda")
        check(lib().b200_op_w4_gemm_umma(ptr(ch), ptr(dev(x)), K, ptr(y), N, None, T, K, N, 128, 0, 1, 0, stream()))
        y2 = torch.empty((T, N), dtype=torch.float32, device="cuda")
        check(lib().b200_op_w4_gemv(ptr(wq), ptr(s), ptr(z), ptr(dev(x)), K, ptr(y2), N, None, T, K, N, 128, 0, stream()))
        f32 = (torch.from_numpy(x).float() @ torch.from_numpy(w).float()).numpy()
        print(f"GEMM K={K} N={N}: |y| rms {np.sqrt((ref ** 2).mean()):.3f}")
        stats("  tcgen05 W4 GEMM vs fp64", y.cpu().numpy(), ref)
        stats("  CUDA-core GEMV  vs fp64", y2.cpu().numpy(), ref)
        stats("  CPU fp32 matmul vs fp64", f32, ref)


def main():
    profile = sys.argv[1] if len(sys.argv) > 1 else "flat"
    n = int(sys.argv[2]) if len(sys.argv) > 2 else 4
    plen = int(sys.argv[3]) if len(sys.argv) > 3 else 128
    gemm_probe()
    dims = PRESETS["qwen2.5-1.5b-awq" if profile == "flat" else "qwen2.5-1.5b-awq-untied"]
    ckpt = synthetic_checkpoint(dims, seed=0, profile=profile)
    rng = np.random.default_rng(1)
    prompts = rng.integers(0, dims.vocab, size=(n, plen), dtype=np.int32)
    got = {}
    for path, name in ((0, "tcgen05"), (1, "gemv")):
        eng = Engine(dims, max_seqs=8, max_model_len=plen + 32, max_batch_tokens=4096, gemm_path=path).load_checkpoint(ckpt)
        _, logits, _ = eng.generate_batch(prompts, 2, n_logit_steps=1)
        got[name] = logits[:, 0].copy()
        eng.close()
    d = Qwen2Dims(**dims.to_dict())
    ref32 = OracleQwen2(d, ckpt, mode="fp32").forward(prompts, last_only=True)[:, -1].numpy()
    ref16 = OracleQwen2(d, ckpt, mode="fp16").forward(prompts, last_only=True)[:, -1].numpy()
    print(f"profile {profile}: {n} prompts x {plen} tokens, logits of the first generated position; logit std {ref32.std():.3f}")
    stats("engine (tcgen05 GEMM family) vs fp32 oracle", got["tcgen05"], ref32)
    stats("engine (CUDA-core GEMV family) vs fp32 oracle", got["gemv"], ref32)
    stats("oracle fp16 mode vs fp32 oracle", ref16, ref32)
    stats("engine (tcgen05) vs oracle fp16 mode", got["tcgen05"], ref16)
    stats("engine (gemv) vs oracle fp16 mode", got["gemv"], ref16)
    stats("engine tcgen05 vs engine gemv", got["tcgen05"], got["gemv"])


if __name__ == "__main__":
    main()
