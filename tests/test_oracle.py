"""The oracle against the golden vectors made from the real third-party code
(tests/golden/make_golden.py): vLLM awq_pack, vLLM top-p, transformers Qwen2."""
import os

import numpy as np

from oracle.awq import unpack_awq, pack_awq, dequantize_awq, AWQ_ORDER
from oracle.qwen2 import OracleQwen2, Qwen2Dims
from oracle.sampler import top_p_keep_mask, sample, hash_uniform, uniform_from_bits
from vgate_b200.weights import ModelDims, synthetic_checkpoint


def test_awq_unpack_inverts_vllm_pack(golden_dir):
    g = np.load(os.path.join(golden_dir, "awq_pack.npz"))
    assert np.array_equal(unpack_awq(g["packed"]), g["q"])
    assert np.array_equal(pack_awq(g["q"]), g["packed"])


def test_awq_order_constant():
    # awq_triton.py:53-57
    assert AWQ_ORDER.tolist() == [0, 4, 1, 5, 2, 6, 3, 7]
    w = np.array([[0x76543210]], dtype=np.uint32).view(np.int32)
    assert unpack_awq(w).tolist() == [[0, 4, 1, 5, 2, 6, 3, 7]]


def test_awq_dequant_formula():
    rng = np.random.default_rng(0)
    K, N, G = 256, 16, 128
    q = rng.integers(0, 16, (K, N)).astype(np.uint8)
    z = rng.integers(0, 16, (K // G, N)).astype(np.uint8)
    s = (rng.random((K // G, N)) * 0.01).astype(np.float16)
    w = dequantize_awq(pack_awq(q), pack_awq(z), s, G)
    ref = ((q.astype(np.float32) - np.repeat(z, G, 0).astype(np.float32)).astype(np.float16)
           * np.repeat(s, G, 0)).astype(np.float16)
    assert w.dtype == np.float16 and np.array_equal(w, ref)


def test_top_p_mask_matches_vllm(golden_dir):
    g = np.load(os.path.join(golden_dir, "top_p.npz"))
    for row in range(g["logits"].shape[0]):
        keep = top_p_keep_mask(g["logits"][row], float(g["p"][row]))
        assert np.array_equal(keep, g["keep"][row]), row


def test_sampler_greedy_and_determinism():
    rng = np.random.default_rng(1)
    l = rng.standard_normal(5000).astype(np.float32)
    assert sample(l, 0.0, 0.9, 1, 0) == int(np.argmax(l))
    a = sample(l, 0.7, 0.9, 123, 4)
    assert a == sample(l, 0.7, 0.9, 123, 4)
    keep = top_p_keep_mask(l / np.float32(0.7), 0.9)
    assert keep[a]
    u = hash_uniform(5, 9, 1000)
    assert u.dtype == np.float32 and (u > 0).all() and (u < 1).all()
    assert abs(u.mean() - 0.5) < 0.05


def test_uniform_is_strictly_inside_unit_interval_for_every_hash_value():
    """The draw is -log(u): u == 1 would give 0 and an infinite race score (a garbage token wins whatever
    its probability), u == 0 an infinite draw.  Only the top 23 hash bits are used, so all 2**23 distinct
    inputs (and the two extremes of the discarded low bits) are checked; fp32 arithmetic, as the kernel's."""
    hi = np.arange(1 << 23, dtype=np.uint64) << np.uint64(9)
    for low in (0, 0x1FF):
        u = uniform_from_bits(hi | np.uint64(low))
        assert u.dtype == np.float32
        assert float(u.min()) > 0.0 and float(u.max()) < 1.0
        assert np.isfinite(-np.log(-np.log(u))).all()
    assert uniform_from_bits(np.uint64(0xFFFFFFFF)) == np.float32(1.0) - np.float32(2.0 ** -24)


def _small():
    d = ModelDims(hidden=128, intermediate=256, n_layers=2, n_heads=2, n_kv_heads=1, head_dim=64, vocab=256)
    return d, synthetic_checkpoint(d, seed=3)


def test_qwen2_logits_match_transformers(golden_dir):
    g = np.load(os.path.join(golden_dir, "hf_qwen2.npz"))
    d, ck = _small()
    m = OracleQwen2(Qwen2Dims(**d.to_dict()), ck, mode="fp32")
    logits = m.forward(g["ids"]).numpy()
    assert np.abs(logits - g["logits"]).max() < 2e-4
    # incremental decode through the KV cache equals the one-shot prefill
    m.reset()
    l0 = m.forward(g["ids"][:, :7]).numpy()
    l1 = np.concatenate([m.forward(g["ids"][:, t:t + 1]).numpy() for t in range(7, 12)], axis=1)
    assert np.abs(np.concatenate([l0, l1], 1) - g["logits"]).max() < 2e-4


def test_qwen2_greedy_matches_transformers(golden_dir):
    g = np.load(os.path.join(golden_dir, "hf_qwen2.npz"))
    d, ck = _small()
    m = OracleQwen2(Qwen2Dims(**d.to_dict()), ck, mode="fp32")
    toks = m.generate_greedy(g["ids"], g["greedy"].shape[1])
    assert np.array_equal(toks, g["greedy"])


def test_last_only_logits_equal_full_forward():
    d, ck = _small()
    m = OracleQwen2(Qwen2Dims(**d.to_dict()), ck, mode="fp32")
    ids = np.random.default_rng(4).integers(0, d.vocab, (3, 11))
    full = m.forward(ids).numpy()
    m.reset()
    last = m.forward(ids, last_only=True).numpy()
    assert last.shape == (3, 1, d.vocab) and np.abs(last[:, 0] - full[:, -1]).max() < 1e-5   # BLAS blocking differs with the shape


def test_peaked_profile_has_wide_margins_and_walks_the_permutation():
    """The parity checkpoint: the head row aligned with the residual-carried embedding leads the field by
    several units (a flat random-init checkpoint has margins below 1e-2 every few dozen steps)."""
    from vgate_b200.weights import next_token_permutation
    d = ModelDims(hidden=256, intermediate=512, n_layers=2, n_heads=2, n_kv_heads=1, head_dim=128, vocab=1024,
                  tie_embeddings=False)
    ck = synthetic_checkpoint(d, seed=5, profile="peaked")
    m = OracleQwen2(Qwen2Dims(**d.to_dict()), ck, mode="fp32")
    prompts = np.random.default_rng(6).integers(0, d.vocab, (4, 9))
    toks, logits = m.generate_greedy(prompts, 6, return_logits=True)
    perm = next_token_permutation(d.vocab, 5)
    assert np.array_equal(toks[:, 0], perm[prompts[:, -1]])
    assert np.array_equal(toks[:, 1:], perm[toks[:, :-1]])
    top2 = np.sort(np.partition(logits, -2, axis=-1)[..., -2:], axis=-1)
    assert (top2[..., 1] - top2[..., 0]).min() > 0.5
    # the flat profile of the same shape is what the docstring says it is
    ckf = synthetic_checkpoint(ModelDims(**{**d.to_dict(), "tie_embeddings": True}), seed=5)
    assert "lm_head.weight" not in ckf and ckf["model.embed_tokens.weight"].std() < 0.03


def test_fp16_mode_close_to_fp32():
    d, ck = _small()
    a = OracleQwen2(Qwen2Dims(**d.to_dict()), ck, mode="fp32")
    b = OracleQwen2(Qwen2Dims(**d.to_dict()), ck, mode="fp16")
    ids = np.random.default_rng(2).integers(0, d.vocab, (2, 9))
    assert np.abs(a.forward(ids).numpy() - b.forward(ids).numpy()).max() < 2e-2
