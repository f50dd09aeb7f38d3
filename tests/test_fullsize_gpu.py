"""North-star parity at the real model size (Qwen2.5-1.5B W4A16 shape: 28 layers, hidden 1536, V = 151936)
on the BASELINE shapes (128-token prompts; batch 8 = config #2, batch 64 = config #3), against the pinned
fp32 oracle:

  * logits within 2e-2 max-abs of the reference at the first 8 generated positions;
  * greedy token ids identical over the first 32 generated tokens on >= 99 % of the prompts -- literally,
    no near-tie allowance.

The gate runs on the ``peaked`` synthetic profile (vgate_b200/weights.py): the same decoder layers, with
logit statistics of a trained model (one candidate several units ahead of a field of sigma 0.78).  On the
``flat`` profile (every logit an exchangeable Gaussian) the top-2 margin is below the rounding noise of ANY
fp16 implementation at a few percent of the steps, so token identity is not defined there; the second test
keeps that checkpoint as a stress case with the near-tie analysis spelled out.

The CPU oracle takes a few minutes on the GPU box's host cores.
"""
import numpy as np
import pytest

from oracle.qwen2 import OracleQwen2, Qwen2Dims
from vgate_b200.weights import PRESETS, synthetic_checkpoint

pytestmark = pytest.mark.gpu

LOGIT_TOL = 2e-2
N_PROMPTS, PROMPT_LEN, N_NEW, N_LOGIT = 128, 128, 32, 8


def _logit_err_until_divergence(logits, ref_logits, toks, ref_toks):
    """max |logits - ref| per prompt over the positions whose inputs are still identical (position t is
    comparable while tokens 0..t-1 agree)."""
    errs = np.zeros(len(toks))
    for b in range(len(toks)):
        n_same = 0
        while n_same < toks.shape[1] and toks[b, n_same] == ref_toks[b, n_same]:
            n_same += 1
        upto = min(logits.shape[1], n_same + 1)
        errs[b] = np.abs(logits[b, :upto] - ref_logits[b, :upto]).max()
    return errs


def test_greedy_gate_on_baseline_shapes():
    from vgate_b200.engine import Engine
    dims = PRESETS["qwen2.5-1.5b-awq-untied"]
    ckpt = synthetic_checkpoint(dims, seed=0, profile="peaked")
    rng = np.random.default_rng(1)
    prompts = rng.integers(0, dims.vocab, size=(N_PROMPTS, PROMPT_LEN), dtype=np.int32)
    eng = Engine(dims, max_seqs=64, max_model_len=PROMPT_LEN + N_NEW + 16, max_batch_tokens=8192).load_checkpoint(ckpt)
    try:
        toks = np.empty((N_PROMPTS, N_NEW), np.int32)
        logits = np.empty((N_PROMPTS, N_LOGIT, dims.vocab), np.float32)
        for s0 in range(0, N_PROMPTS, 64):                                  # config #3's batch
            toks[s0:s0 + 64], logits[s0:s0 + 64], _ = eng.generate_batch(prompts[s0:s0 + 64], N_NEW, n_logit_steps=N_LOGIT)
        toks8, logits8, _ = eng.generate_batch(prompts[:8], N_NEW, n_logit_steps=N_LOGIT)     # config #2's batch
        # the same 64 prompts as independent requests through the scheduler (submit / wait): the path V-Gate drives
        rids = [eng.submit(p, 0.0, 1.0, N_NEW) for p in prompts[:64]]
        sched = np.stack([eng.wait(r, N_NEW)[0] for r in rids])
    finally:
        eng.close()
    # batch composition and entry point never change a sequence's tokens
    assert np.array_equal(toks8, toks[:8])
    assert np.array_equal(logits8, logits[:8])
    assert np.array_equal(sched, toks[:64])

    ref = OracleQwen2(Qwen2Dims(**dims.to_dict()), ckpt, mode="fp32")
    ref_toks, ref_logits = ref.generate_greedy(prompts, N_NEW, return_logits=True, n_logit_steps=N_LOGIT)
    del ref
    ref_logits = np.asarray(ref_logits)

    errs = _logit_err_until_divergence(logits, ref_logits, toks, ref_toks)
    same = (toks == ref_toks).all(axis=1)
    top2 = np.sort(np.partition(ref_logits, -2, axis=-1)[..., -2:], axis=-1)
    margin = top2[..., 1] - top2[..., 0]
    print(f"identical over {N_NEW} tokens: {int(same.sum())}/{N_PROMPTS} prompts; max |logits - fp32 oracle| {errs.max():.2e} "
          f"over {N_LOGIT} positions; reference top-2 margin min {margin.min():.2f} median {np.median(margin):.2f}; "
          f"distinct tokens per sequence min {min(len(set(t)) for t in ref_toks.tolist())}")
    assert errs.max() < LOGIT_TOL, errs.max()
    assert same.mean() >= 0.99, f"{int(same.sum())} of {N_PROMPTS} prompts identical over {N_NEW} tokens"
    # the sequences are not degenerate: greedy decoding walks the vocabulary
    assert min(len(set(t)) for t in ref_toks.tolist()) >= N_NEW // 2


def test_flat_checkpoint_stress_logits_and_near_ties():
    """The ``flat`` profile (tied embeddings, the checkpoint bench.py times): logits of 151 936 exchangeable
    Gaussians.  Logits stay within 2e-2 of the fp32 oracle; a trajectory may leave the oracle's only at a step
    where the oracle's own top-2 margin is smaller than twice that tolerance (two results that are both within
    tolerance can order such a pair differently), never elsewhere."""
    from vgate_b200.engine import Engine
    dims = PRESETS["qwen2.5-1.5b-awq"]
    ckpt = synthetic_checkpoint(dims, seed=0)
    rng = np.random.default_rng(1)
    n, plen = 16, 128
    prompts = rng.integers(0, dims.vocab, size=(n, plen), dtype=np.int32)
    eng = Engine(dims, max_seqs=16, max_model_len=256, max_batch_tokens=4096).load_checkpoint(ckpt)
    try:
        toks, logits, _ = eng.generate_batch(prompts, N_NEW, n_logit_steps=N_NEW)
        solo, _, _ = eng.generate_batch(prompts[:1], N_NEW)
    finally:
        eng.close()
    assert np.array_equal(solo[0], toks[0])
    ref = OracleQwen2(Qwen2Dims(**dims.to_dict()), ckpt, mode="fp32")
    ref_toks, ref_logits = ref.generate_greedy(prompts, N_NEW, return_logits=True)
    del ref
    ref_logits = np.asarray(ref_logits)
    errs = _logit_err_until_divergence(logits, ref_logits, toks, ref_toks)
    n_cmp = np.array([int(np.argmax(np.append(toks[b] != ref_toks[b], True))) + 1 for b in range(n)]).clip(max=N_NEW)
    per_pos = [float(max((np.abs(logits[b, t] - ref_logits[b, t]).max() for b in range(n) if t < n_cmp[b]), default=0.0))
               for t in range(N_NEW)]
    print(f"flat profile: max |logits - fp32 oracle| per prompt {np.round(errs, 4).tolist()}; per position {np.round(per_pos, 4).tolist()}")
    # the first position is the north-star gate's; over 32 positions x 16 prompts x 151 936 logits the extreme of
    # the same error distribution is allowed half as much again
    assert per_pos[0] < LOGIT_TOL and errs.max() < 1.5 * LOGIT_TOL, errs.max()
    top2 = np.sort(np.partition(ref_logits, -2, axis=-1)[..., -2:], axis=-1)
    margin = top2[..., 1] - top2[..., 0]
    same = (toks == ref_toks).all(axis=1)
    div = []
    for b in np.where(~same)[0]:
        t = int(np.argmax(toks[b] != ref_toks[b]))
        div.append((int(b), t, float(margin[b, t])))
        # the token the engine chose instead is within twice the tolerance of the oracle's best
        assert ref_logits[b, t].max() - ref_logits[b, t, toks[b, t]] < 2 * LOGIT_TOL, (b, t, margin[b, t])
    print(f"flat profile: identical {int(same.sum())}/{n} prompts over {N_NEW} tokens; max |logits - fp32 oracle| "
          f"{errs.max():.2e}; divergences (prompt, step, oracle margin): {div}; "
          f"steps with oracle margin < 1e-2: {int((margin < 1e-2).sum())} of {margin.size}")


def test_batch_200_matches_oracle():
    """Decode batches above 128 rows take other kernel variants -- two token tiles with 256-row GEMM CTAs
    (ops_umma.cu dispatch_nt) and the three-warp, two-CTAs-per-SM attention (ops_attn.cu launch_gq) -- so they
    get their own engine-vs-oracle check at the real layer shapes (4 layers, full vocabulary)."""
    from vgate_b200.engine import Engine
    dims = PRESETS["qwen2.5-1.5b-awq-l4-untied"]
    ckpt = synthetic_checkpoint(dims, seed=7, profile="peaked")
    rng = np.random.default_rng(8)
    n, plen, n_new, n_logit = 200, 24, 12, 4
    prompts = rng.integers(0, dims.vocab, size=(n, plen), dtype=np.int32)
    eng = Engine(dims, max_seqs=256, max_model_len=64, max_batch_tokens=8192).load_checkpoint(ckpt)
    try:
        toks, logits, _ = eng.generate_batch(prompts, n_new, n_logit_steps=n_logit)
        few, _, _ = eng.generate_batch(prompts[:3], n_new)
    finally:
        eng.close()
    assert np.array_equal(few, toks[:3])                       # the same bits as the small-batch variants
    ref = OracleQwen2(Qwen2Dims(**dims.to_dict()), ckpt, mode="fp32")
    ref_toks, ref_logits = ref.generate_greedy(prompts, n_new, return_logits=True, n_logit_steps=n_logit)
    assert np.array_equal(toks, ref_toks)
    assert np.abs(logits - np.asarray(ref_logits)).max() < LOGIT_TOL
