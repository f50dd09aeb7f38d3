"""The persistent decode kernel (csrc/ops_mega.cu: every layer of a decode step in one launch) must
produce the same bits as the one-kernel-per-operation path it replaces -- same device functions, same
split-K factors, same summation orders -- and, like it, must not depend on batch composition.
VGATE_B200_MEGA=0 selects the per-kernel path (read once, when an engine is finalized)."""
import numpy as np
import pytest

from vgate_b200.weights import PRESETS, synthetic_checkpoint

pytestmark = pytest.mark.gpu


def _run(monkeypatch, mega, preset, prompts, n_new, n_logit_steps, fused_mlp=False, **kw):
    from vgate_b200.engine import Engine
    monkeypatch.setenv("VGATE_B200_MEGA", "1" if mega else "0")
    monkeypatch.setenv("VGATE_B200_FUSED_MLP", "1" if fused_mlp else "0")
    dims = PRESETS[preset]
    eng = Engine(dims, **kw).load_checkpoint(synthetic_checkpoint(dims, seed=0))
    try:
        toks, logits, _ = eng.generate_batch(prompts, n_new, n_logit_steps=n_logit_steps)
        launches = eng.stats()["kernel_launches"]
    finally:
        eng.close()
    return toks, logits, launches


@pytest.mark.parametrize("preset,n_seqs,prompt_len", [("tiny-awq", 5, 37), ("qwen2.5-1.5b-awq-l2-v8k", 64, 70),
                                                      ("qwen2.5-1.5b-awq-l2-v8k", 3, 129)])
def test_persistent_decode_equals_per_kernel_path_bitwise(monkeypatch, preset, n_seqs, prompt_len):
    dims = PRESETS[preset]
    rng = np.random.default_rng(7)
    prompts = rng.integers(0, dims.vocab, size=(n_seqs, prompt_len), dtype=np.int32)
    kw = dict(max_seqs=64, max_model_len=256, max_batch_tokens=1024)
    n_new, n_logits = 12, 6
    t0, l0, k0 = _run(monkeypatch, False, preset, prompts, n_new, n_logits, **kw)
    t1, l1, k1 = _run(monkeypatch, True, preset, prompts, n_new, n_logits, **kw)
    assert k1 < k0, (k0, k1)                       # the persistent kernel really replaced the per-layer launches
    assert np.array_equal(l0.view(np.uint32), l1.view(np.uint32)), np.abs(l0 - l1).max()
    assert np.array_equal(t0, t1)


def test_persistent_decode_is_batch_invariant(monkeypatch):
    preset = "qwen2.5-1.5b-awq-l2-v8k"
    dims = PRESETS[preset]
    rng = np.random.default_rng(8)
    prompts = rng.integers(0, dims.vocab, size=(33, 50), dtype=np.int32)
    kw = dict(max_seqs=64, max_model_len=256, max_batch_tokens=2048)
    t_all, l_all, _ = _run(monkeypatch, True, preset, prompts, 10, 4, **kw)
    t_one, l_one, _ = _run(monkeypatch, True, preset, prompts[5:6], 10, 4, **kw)
    assert np.array_equal(l_all[5:6].view(np.uint32), l_one.view(np.uint32))
    assert np.array_equal(t_all[5:6], t_one)


@pytest.mark.parametrize("preset,n_seqs,prompt_len", [("tiny-awq", 5, 37), ("qwen2.5-1.5b-awq-l2-v8k", 64, 70),
                                                      ("qwen2.5-1.5b-awq-l2-v8k", 3, 129), ("qwen2.5-1.5b-awq-l4", 33, 40)])
def test_fused_mlp_equals_two_kernel_path_bitwise(monkeypatch, preset, n_seqs, prompt_len):
    """gate_up + SwiGLU + down as one dataflow kernel (csrc/ops_mega.cu mlp_fused_kernel, opt-in) against the
    two separate GEMM launches: identical logits and tokens."""
    dims = PRESETS[preset]
    rng = np.random.default_rng(9)
    prompts = rng.integers(0, dims.vocab, size=(n_seqs, prompt_len), dtype=np.int32)
    kw = dict(max_seqs=64, max_model_len=256, max_batch_tokens=4096)
    n_new, n_logits = 12, 6
    t0, l0, k0 = _run(monkeypatch, False, preset, prompts, n_new, n_logits, fused_mlp=False, **kw)
    t1, l1, k1 = _run(monkeypatch, False, preset, prompts, n_new, n_logits, fused_mlp=True, **kw)
    assert k1 < k0, (k0, k1)
    assert np.array_equal(l0.view(np.uint32), l1.view(np.uint32)), np.abs(l0 - l1).max()
    assert np.array_equal(t0, t1)
