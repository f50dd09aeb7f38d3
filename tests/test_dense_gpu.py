"""Dense 16-bit checkpoints (configs #4 / #5: Qwen2.5-7B / 14B bf16): the TMA-fed tcgen05 GEMM with fp16 and bf16
operands and all its epilogues against fp64, and the engine end to end against the pinned fp32 oracle -- greedy
tokens identical (peaked profile), logits within a tolerance stated per storage type: 2e-2 for fp16 (north_star), 1e-1
for bf16 (8 mantissa bits instead of 11: every rounding point of the path is 8x coarser; the oracle computes in fp32
on the same bf16 weights)."""
import numpy as np
import pytest
import torch

from oracle.qwen2 import OracleQwen2, Qwen2Dims
from vgate_b200 import _lib
from vgate_b200.weights import PRESETS, synthetic_checkpoint, to_bf16_bits, from_bf16_bits

pytestmark = pytest.mark.gpu

from tests.gpu_util import lib, dev, ptr, stream, check  # noqa: E402

TOL = {"fp16": 2e-2, "bf16": 1e-1}


def _round16(a, dtype):
    return from_bf16_bits(to_bf16_bits(a)) if dtype == "bf16" else a.astype(np.float16).astype(np.float32)


def _bits(a, dtype):
    return to_bf16_bits(a) if dtype == "bf16" else a.astype(np.float16)


@pytest.mark.parametrize("dtype", ["fp16", "bf16"])
@pytest.mark.parametrize("T,K,N", [(1, 1024, 512), (33, 1024, 512), (128, 1024, 512), (300, 1024, 512),
                                   (600, 512, 9728),        # 256 weight rows per CTA (76 tiles x 3 token tiles > 148)
                                   (1100, 256, 19456)])     # ... and token-tile-fastest rasterisation (76 x 5 CTAs > two waves, weights > activations)
def test_dense_gemm_epilogues_match_fp64(dtype, T, K, N):
    rng = np.random.default_rng(T)
    code = _lib.DTYPE_BF16 if dtype == "bf16" else _lib.DTYPE_F16
    w = _round16(rng.standard_normal((N, K)).astype(np.float32) * 0.05, dtype)
    x = _round16(rng.standard_normal((T, K)).astype(np.float32), dtype)
    b = _round16(rng.standard_normal(N).astype(np.float32) * 0.1, dtype)
    ref = x.astype(np.float64) @ w.astype(np.float64).T
    wd, xd, bd = dev(_bits(w, dtype)), dev(_bits(x, dtype)), dev(_bits(b, dtype))
    y = torch.full((T, N), 3.0, dtype=torch.float32, device="cuda")
    check(lib().b200_op_dense_gemm_umma(ptr(wd), ptr(xd), K, ptr(y), N, ptr(bd), T, K, N, code, 0, 1, 0, stream()))
    assert np.abs(y.cpu().numpy() - (ref + b)).max() < 5e-4
    y0 = rng.standard_normal((T, N)).astype(np.float32)
    y = dev(y0)
    check(lib().b200_op_dense_gemm_umma(ptr(wd), ptr(xd), K, ptr(y), N, None, T, K, N, code, 1, 1, 0, stream()))
    assert np.abs(y.cpu().numpy() - (y0 + ref)).max() < 5e-4
    # split-K partials sum to the result
    ns = lib().b200_op_dense_effective_splits(K, 5)
    parts = torch.zeros((ns, T, N), dtype=torch.float32, device="cuda")
    check(lib().b200_op_dense_gemm_umma(ptr(wd), ptr(xd), K, ptr(parts), N, None, T, K, N, code, 0, ns, T * N, stream()))
    assert np.abs(parts.cpu().numpy().astype(np.float64).sum(0) - ref).max() < 5e-4
    # SwiGLU epilogue: gate / up interleaved in blocks of 64 rows
    I = N // 2
    g, u = ref.reshape(T, N // 128, 2, 64)[:, :, 0].reshape(T, I), ref.reshape(T, N // 128, 2, 64)[:, :, 1].reshape(T, I)
    want = g / (1 + np.exp(-g)) * u
    act = torch.zeros((T, I), dtype=torch.int16, device="cuda")
    check(lib().b200_op_dense_gemm_umma(ptr(wd), ptr(xd), K, ptr(act), I, None, T, K, N, code, 2, 1, 0, stream()))
    a16 = act.cpu().numpy().view(np.uint16)
    got = from_bf16_bits(a16) if dtype == "bf16" else a16.view(np.float16).astype(np.float32)
    rel = 2.0 ** -8 if dtype == "bf16" else 2.0 ** -11
    assert (np.abs(got - want) <= rel * np.abs(want) + 1e-4).all()


@pytest.mark.parametrize("preset,n,plen", [("tiny-f16-dense", 5, 33), ("tiny-bf16", 5, 33), ("qwen2.5-7b-bf16-l2", 4, 40)])
def test_dense_engine_matches_fp32_oracle(preset, n, plen):
    from vgate_b200.engine import Engine
    dims = PRESETS[preset]
    ckpt = synthetic_checkpoint(dims, seed=3, profile="peaked")
    rng = np.random.default_rng(12)
    prompts = rng.integers(0, dims.vocab, size=(n, plen), dtype=np.int32)
    n_new = 10
    eng = Engine(dims, max_seqs=8, max_model_len=128, max_batch_tokens=512).load_checkpoint(ckpt)
    try:
        toks, logits, _ = eng.generate_batch(prompts, n_new, n_logit_steps=n_new)
        solo, _, _ = eng.generate_batch(prompts[:1], n_new)
        rids = [eng.submit(p, 0.0, 1.0, n_new) for p in prompts]
        sched = np.stack([eng.wait(r, n_new)[0] for r in rids])
        st = eng.stats()
    finally:
        eng.close()
    assert np.array_equal(solo[0], toks[0]) and np.array_equal(sched, toks)     # batch composition / entry point invariant
    assert st["kv_blocks_free"] == st["kv_blocks_total"]
    ref = OracleQwen2(Qwen2Dims(**dims.to_dict()), ckpt, mode="fp32")
    ref_toks, ref_logits = ref.generate_greedy(prompts, n_new, return_logits=True)
    err = float(np.abs(logits - np.asarray(ref_logits)).max())
    print(f"{preset}: max |logits - fp32 oracle| {err:.3e} (tolerance {TOL[dims.dtype]}), logit std {np.asarray(ref_logits).std():.2f}")
    assert np.array_equal(toks, ref_toks)
    assert err < TOL[dims.dtype], err
