"""End-to-end parity of the engine (C ABI) against the pinned fp32 CPU oracle on seeded random-init
checkpoints: logits within 2e-2 max-abs, greedy tokens identical (on the ``peaked`` synthetic profile,
where token identity is well defined -- see vgate_b200/weights.py)."""
import threading

import numpy as np
import pytest

from oracle.qwen2 import OracleQwen2, Qwen2Dims
from vgate_b200.weights import PRESETS, synthetic_checkpoint

pytestmark = pytest.mark.gpu

LOGIT_TOL = 2e-2          # north_star: logits within 2e-2 max-abs of the reference


@pytest.fixture(scope="module")
def small():
    from vgate_b200.engine import Engine
    dims = PRESETS["qwen2.5-1.5b-awq-l2-v8k"]
    ckpt = synthetic_checkpoint(dims, seed=0)
    eng = Engine(dims, max_seqs=16, max_model_len=512, max_batch_tokens=256).load_checkpoint(ckpt)
    yield dims, ckpt, eng
    eng.close()


def _oracle(dims, ckpt, mode):
    return OracleQwen2(Qwen2Dims(**dims.to_dict()), ckpt, mode=mode)


def test_logits_match_oracle_on_flat_checkpoint(small):
    dims, ckpt, eng = small
    rng = np.random.default_rng(1)
    prompts = rng.integers(0, dims.vocab, size=(6, 40), dtype=np.int32)
    n_new = 12
    toks, logits, _ = eng.generate_batch(prompts, n_new, n_logit_steps=n_new)
    ref32_toks, ref32_logits = _oracle(dims, ckpt, "fp32").generate_greedy(prompts, n_new, return_logits=True)
    # logits: compare while the prefixes agree (after a divergence the inputs differ)
    for b in range(prompts.shape[0]):
        n_same = 0
        while n_same < n_new and toks[b, n_same] == ref32_toks[b, n_same]:
            n_same += 1
        upto = min(n_new, n_same + 1)
        assert np.abs(logits[b, :upto] - ref32_logits[b, :upto]).max() < LOGIT_TOL, b
        if n_same < n_new:      # a divergence may only pick a token the oracle ranks within twice the tolerance of its best
            assert ref32_logits[b, n_same].max() - ref32_logits[b, n_same, toks[b, n_same]] < 2 * LOGIT_TOL, b


@pytest.mark.parametrize("preset,batch,plen", [("qwen2.5-1.5b-awq-l2-v8k-untied", 6, 40),
                                                 ("qwen2.5-1.5b-awq-l2-v8k", 5, 17)])
def test_greedy_tokens_identical_to_fp32_oracle(preset, batch, plen):
    """Token identity against the pinned fp32 oracle, untied (permuted head: the walk through the vocabulary
    checks embedding gather and argmax indexing) and tied (the head the 1.5B model ships with)."""
    from vgate_b200.engine import Engine
    dims = PRESETS[preset]
    ckpt = synthetic_checkpoint(dims, seed=2, profile="peaked")
    rng = np.random.default_rng(11)
    prompts = rng.integers(0, dims.vocab, size=(batch, plen), dtype=np.int32)
    n_new = 24
    eng = Engine(dims, max_seqs=8, max_model_len=128, max_batch_tokens=512).load_checkpoint(ckpt)
    try:
        toks, logits, _ = eng.generate_batch(prompts, n_new, n_logit_steps=n_new)
        rids = [eng.submit(p, 0.0, 1.0, n_new) for p in prompts]
        sched = np.stack([eng.wait(r, n_new)[0] for r in rids])
    finally:
        eng.close()
    ref_toks, ref_logits = _oracle(dims, ckpt, "fp32").generate_greedy(prompts, n_new, return_logits=True)
    assert np.array_equal(toks, ref_toks)
    assert np.array_equal(sched, ref_toks)
    assert np.abs(logits - np.asarray(ref_logits)).max() < LOGIT_TOL
    if not dims.tie_embeddings:
        assert len(set(ref_toks[0].tolist())) > n_new // 2          # a walk, not a repeated token


def test_submit_wait_equals_generate_batch(small):
    dims, ckpt, eng = small
    rng = np.random.default_rng(2)
    prompts = rng.integers(0, dims.vocab, size=(5, 33), dtype=np.int32)
    toks, _, _ = eng.generate_batch(prompts, 10)
    rids = [eng.submit(p, temperature=0.0, top_p=1.0, max_tokens=10) for p in prompts]
    for b, rid in enumerate(rids):
        out, ttft, gen = eng.wait(rid, 10)
        assert np.array_equal(out, toks[b])
        assert ttft > 0 and gen >= 0


def test_continuous_batching_ragged_and_staggered(small):
    """Requests of different lengths, submitted from several threads while others decode, produce
    exactly what each produces alone (batch composition never changes a sequence's tokens)."""
    dims, ckpt, eng = small
    rng = np.random.default_rng(3)
    lens = [1, 5, 16, 17, 31, 64, 100, 7, 48, 3, 90, 22]
    news = [9, 3, 12, 1, 7, 16, 5, 11, 2, 14, 6, 8]
    prompts = [rng.integers(0, dims.vocab, size=n, dtype=np.int32) for n in lens]
    solo = []
    for p, n in zip(prompts, news):
        solo.append(eng.wait(eng.submit(p, 0.0, 1.0, n), n)[0])
    results = [None] * len(prompts)

    def worker(i):
        results[i] = eng.wait(eng.submit(prompts[i], 0.0, 1.0, news[i]), news[i])[0]

    threads = [threading.Thread(target=worker, args=(i,)) for i in range(len(prompts))]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    for i in range(len(prompts)):
        assert len(results[i]) == news[i]
        assert np.array_equal(results[i], solo[i]), i
    st = eng.stats()
    assert st["running"] == 0 and st["waiting"] == 0
    assert st["kv_blocks_free"] == st["kv_blocks_total"]      # every block returned to the allocator


def test_poll_streams_tokens_and_abort(small):
    dims, ckpt, eng = small
    rng = np.random.default_rng(4)
    p = rng.integers(0, dims.vocab, size=20, dtype=np.int32)
    full = eng.wait(eng.submit(p, 0.0, 1.0, 15), 15)[0]
    rid = eng.submit(p, 0.0, 1.0, 15)
    got, finished = [], False
    while not finished:
        ids, finished = eng.poll(rid, cap=4, timeout_ms=200)
        assert len(ids) <= 4
        got.extend(ids.tolist())
    assert got == full.tolist()
    # abort: the request ends early and its blocks are returned
    rid = eng.submit(p, 0.0, 1.0, 400)
    ids, _ = eng.poll(rid, cap=8, timeout_ms=2000)
    eng.abort(rid)
    from vgate_b200.engine import B200Error
    with pytest.raises(B200Error):
        while True:
            _, fin = eng.poll(rid, cap=64, timeout_ms=200)
            if fin:
                break
    st = eng.stats()
    assert st["kv_blocks_free"] == st["kv_blocks_total"]
    # an aborted request leaves nothing behind: its id is unknown afterwards (abort releases the table entry),
    # also when nobody ever polls it again (client disconnect, main.py:372-376)
    rid = eng.submit(p, 0.0, 1.0, 400)
    eng.poll(rid, cap=8, timeout_ms=2000)
    eng.abort(rid)
    import time
    for _ in range(200):
        if eng.stats()["running"] == 0:
            break
        time.sleep(0.01)
    with pytest.raises(B200Error) as ei:
        eng.poll(rid, cap=8, timeout_ms=0)
    assert ei.value.code == -5                                  # B200_ERR_NOT_FOUND
    n_fin = eng.stats()["requests_finished"]
    done = eng.submit(p, 0.0, 1.0, 3)
    for _ in range(500):
        if eng.stats()["requests_finished"] > n_fin:
            break
        time.sleep(0.01)
    eng.abort(done)                                             # already finished, never waited on: erased as well
    with pytest.raises(B200Error):
        eng.wait(done, 3)


def test_sampling_requests_are_seeded(small):
    dims, ckpt, eng = small
    rng = np.random.default_rng(5)
    p = rng.integers(0, dims.vocab, size=12, dtype=np.int32)
    a = eng.wait(eng.submit(p, 0.8, 0.9, 12, seed=7), 12)[0]
    b = eng.wait(eng.submit(p, 0.8, 0.9, 12, seed=7), 12)[0]
    c = eng.wait(eng.submit(p, 0.8, 0.9, 12, seed=8), 12)[0]
    assert np.array_equal(a, b)
    assert not np.array_equal(a, c)


def test_capacity_and_argument_errors(small):
    dims, ckpt, eng = small
    from vgate_b200.engine import B200Error
    with pytest.raises(B200Error):
        eng.submit(np.zeros(512, np.int32), 0.0, 1.0, 4)            # a prompt of max_model_len tokens leaves no room
    # vLLM semantics: a prompt shorter than max_model_len is accepted and the generation stops when the
    # sequence reaches max_model_len (500 + 12 = 512), whatever max_tokens asked for
    out = eng.wait(eng.submit(np.zeros(500, np.int32), 0.0, 1.0, 100), 100)[0]
    assert len(out) == 12
    st = eng.stats()
    assert st["kv_blocks_free"] == st["kv_blocks_total"]
    with pytest.raises(B200Error):
        eng.submit(np.array([dims.vocab], np.int32), 0.0, 1.0, 4)   # token id out of range
    with pytest.raises(B200Error):
        eng.submit(np.array([1], np.int32), 0.0, 0.0, 4)            # top_p == 0
    with pytest.raises(B200Error):
        eng.wait(123456, 4)


def test_backend_contract_on_gpu(monkeypatch):
    """B200Backend end to end with the reference's contract assertions (tests/test_backends.py:45-94)."""
    import asyncio
    from types import SimpleNamespace
    from vgate_b200 import B200Backend
    from vgate_b200.tokenizer import ByteTokenizer

    monkeypatch.setenv("VGATE_B200_SYNTHETIC", "qwen2.5-1.5b-awq-l2-v8k")
    monkeypatch.setenv("VGATE_B200_MAX_SEQS", "8")
    monkeypatch.setenv("VGATE_B200_MAX_BATCH_TOKENS", "256")
    be = B200Backend()
    be.load_model(SimpleNamespace(model_id="Qwen/Qwen2.5-1.5B-Instruct-AWQ", max_model_len=256))
    try:
        sp = be.create_sampling_params(0.0, 1.0, 6)
        res = be.generate(["User: hello\nAssistant:", "User: bye\nAssistant:"], sp)
        assert len(res) == 2
        for r in res:
            assert set(r) >= {"text", "token_ids", "num_tokens", "metrics"}
            assert r["num_tokens"] == len(r["token_ids"]) == 6
            assert ByteTokenizer.parse(r["text"]) == r["token_ids"]
            assert r["metrics"]["ttft"] > 0 and r["metrics"]["gen_time"] >= 0

        async def collect():
            return [c async for c in be.stream_generate("User: hello\nAssistant:", sp)]

        chunks = asyncio.run(collect())
        assert "".join(c["delta"] for c in chunks) == res[0]["text"]
        counts = [c["num_tokens"] for c in chunks]
        assert counts == sorted(counts) and counts[-1] == 6
    finally:
        be.shutdown()


def test_lookahead_scheduling_changes_nothing(monkeypatch):
    """The scheduler launches decode step s + 1 before it has read the tokens of step s
    (engine.cu run_decode).  Same tokens as without look-ahead for: different max_tokens in one batch,
    a sequence stopping on EOS mid-way (its over-run token is dropped), requests arriving while others
    decode, sampled (seeded) requests."""
    import time
    from vgate_b200.engine import Engine
    dims = PRESETS["qwen2.5-1.5b-awq-l2-v8k"]
    ckpt = synthetic_checkpoint(dims, seed=0)
    rng = np.random.default_rng(11)
    prompts = [rng.integers(0, dims.vocab, size=n, dtype=np.int32) for n in (20, 33, 47, 64, 9, 70)]
    max_toks = [24, 7, 31, 16, 12, 20]

    def run(lookahead, eos):
        monkeypatch.setenv("VGATE_B200_LOOKAHEAD", "1" if lookahead else "0")
        eng = Engine(dims, max_seqs=16, max_model_len=256, max_batch_tokens=256, eos_token_id=eos).load_checkpoint(ckpt)
        try:
            rids = []
            for i, (p, m) in enumerate(zip(prompts, max_toks)):
                temp = 0.8 if i == 4 else 0.0
                rids.append(eng.submit(p, temperature=temp, top_p=0.9 if i == 4 else 1.0, max_tokens=m, seed=5))
                if i == 2:
                    time.sleep(0.02)                 # the rest arrive while the first three decode
            return [eng.wait(rid, m)[0] for rid, m in zip(rids, max_toks)]
        finally:
            eng.close()

    base = run(False, -1)
    assert [len(o) for o in base] == max_toks
    for a, b in zip(base, run(True, -1)):
        assert np.array_equal(a, b)
    eos = int(base[0][10])                            # a token sequence 0 emits mid-way becomes the EOS id
    cut_off = run(False, eos)
    cut_on = run(True, eos)
    first = int(np.argmax(base[0] == eos))
    assert len(cut_off[0]) == first + 1 and cut_off[0][-1] == eos
    for a, b in zip(cut_off, cut_on):
        assert np.array_equal(a, b)


def test_full_context_2048_matches_oracle():
    """The reference's max_model_len (vllm_backend.py:61-86: 2048): a sequence that fills it exactly.
    prompt 2030 + 18 new tokens on a small model; the logits of the last positions (32 KV chunks, 128 KV
    blocks per head) stay within tolerance of the fp32 oracle."""
    from vgate_b200.engine import Engine
    dims = PRESETS["tiny-awq"]
    ckpt = synthetic_checkpoint(dims, seed=3)
    rng = np.random.default_rng(21)
    prompts = rng.integers(0, dims.vocab, size=(2, 2030), dtype=np.int32)
    n_new = 18
    eng = Engine(dims, max_seqs=4, max_model_len=2048, max_batch_tokens=4096).load_checkpoint(ckpt)
    try:
        toks, logits, _ = eng.generate_batch(prompts, n_new, n_logit_steps=n_new)
        # one token past max_model_len is clamped (vLLM stops at max_model_len): the same 18 tokens come back
        clamped = eng.wait(eng.submit(prompts[0], 0.0, 1.0, n_new + 1), n_new + 1)[0]
        assert np.array_equal(clamped, toks[0])
    finally:
        eng.close()
    ref32 = _oracle(dims, ckpt, "fp32")
    ref_toks, ref_logits = ref32.generate_greedy(prompts, n_new, return_logits=True)
    ref_logits = np.asarray(ref_logits)
    for b in range(2):                       # compare while the trajectories agree (teacher forcing is not available)
        n_same = 0
        while n_same < n_new and toks[b, n_same] == ref_toks[b][n_same]:
            n_same += 1
        upto = min(n_new, n_same + 1)
        assert upto >= 1
        assert np.abs(logits[b, :upto] - ref_logits[b, :upto]).max() < LOGIT_TOL, b


def test_more_requests_than_sequences_and_kv_blocks():
    """Admission control: 3x more requests than max_seqs, and a KV pool that holds only a few of them at a
    time; everything completes, FIFO, with the tokens each request produces alone."""
    from vgate_b200.engine import Engine
    dims = PRESETS["tiny-awq"]
    ckpt = synthetic_checkpoint(dims, seed=4)
    rng = np.random.default_rng(22)
    prompts = [rng.integers(0, dims.vocab, size=int(n), dtype=np.int32) for n in rng.integers(1, 120, size=24)]
    news = [int(n) for n in rng.integers(1, 40, size=24)]
    big = Engine(dims, max_seqs=32, max_model_len=256, max_batch_tokens=512).load_checkpoint(ckpt)
    try:
        solo = [big.wait(big.submit(p, 0.0, 1.0, n), n)[0] for p, n in zip(prompts, news)]
    finally:
        big.close()
    kv_tok_bytes = 2 * dims.n_layers * dims.n_kv_heads * dims.head_dim * 2
    small_eng = Engine(dims, max_seqs=8, max_model_len=256, max_batch_tokens=512,
                       kv_bytes=kv_tok_bytes * 16 * 40).load_checkpoint(ckpt)      # 40 blocks: ~4 long sequences at once
    try:
        rids = [small_eng.submit(p, 0.0, 1.0, n) for p, n in zip(prompts, news)]
        outs = [small_eng.wait(r, n)[0] for r, n in zip(rids, news)]
        st = small_eng.stats()
    finally:
        small_eng.close()
    for i, (a, b) in enumerate(zip(outs, solo)):
        assert np.array_equal(a, b), i
    assert st["requests_finished"] == 24 and st["kv_blocks_free"] == st["kv_blocks_total"]


def test_chunked_prefill_interleaves_with_decode_and_changes_nothing(monkeypatch):
    """Prompts that arrive while other sequences decode are prefilled in chunks that alternate with decode steps
    (VGATE_B200_PREFILL_CHUNK, here 48 tokens so that a 300-token prompt takes 7 passes): every sequence still
    produces exactly the tokens it produces alone, prompts cut at arbitrary chunk boundaries included, and the
    running sequences keep receiving tokens between the chunks."""
    import time
    from vgate_b200.engine import Engine
    dims = PRESETS["qwen2.5-1.5b-awq-l2-v8k"]
    ckpt = synthetic_checkpoint(dims, seed=6)
    rng = np.random.default_rng(41)
    lens = [300, 45, 48, 49, 130, 7]
    late = [rng.integers(0, dims.vocab, size=n, dtype=np.int32) for n in lens]
    early = [rng.integers(0, dims.vocab, size=20, dtype=np.int32) for _ in range(4)]
    monkeypatch.setenv("VGATE_B200_PREFILL_CHUNK", "48")
    monkeypatch.setenv("VGATE_B200_MAX_STALL_MS", "0")        # strict alternation: a decode step after every chunk
    monkeypatch.setenv("VGATE_B200_PREFIX_CACHE", "0")      # the solo runs below would otherwise leave the prompts cached
    eng = Engine(dims, max_seqs=16, max_model_len=512, max_batch_tokens=256).load_checkpoint(ckpt)
    try:
        solo_late = [eng.wait(eng.submit(p, 0.0, 1.0, 10), 10)[0] for p in late]
        solo_early = [eng.wait(eng.submit(p, 0.0, 1.0, 150), 150)[0] for p in early]
        st0 = eng.stats()
        rids_e = [eng.submit(p, 0.0, 1.0, 150) for p in early]
        while eng.stats()["tokens_generated"] < st0["tokens_generated"] + 40:      # the early ones are decoding
            time.sleep(0.002)
        rids_l = [eng.submit(p, 0.0, 1.0, 10) for p in late]
        out_l = [eng.wait(r, 10)[0] for r in rids_l]
        out_e = [eng.wait(r, 150)[0] for r in rids_e]
        st1 = eng.stats()
    finally:
        eng.close()
    for a, b in zip(out_l, solo_late):
        assert np.array_equal(a, b)
    for a, b in zip(out_e, solo_early):
        assert np.array_equal(a, b)
    # 579 prompt tokens in chunks of <= 48 while four sequences decode: at least 12 prefill passes
    assert st1["steps_prefill"] - st0["steps_prefill"] >= 12
    assert st1["kv_blocks_free"] == st1["kv_blocks_total"]


def test_prefix_cache_shares_prompt_blocks_and_changes_nothing(monkeypatch):
    """Requests whose prompts start with the same tokens share the K/V blocks of that prefix (block granularity, 16
    tokens) and prefill only the rest: same tokens as with the cache switched off, fewer prompt tokens computed,
    every block reclaimable afterwards, and eviction under a small KV pool keeps results intact."""
    from vgate_b200.engine import Engine
    dims = PRESETS["qwen2.5-1.5b-awq-l2-v8k"]
    ckpt = synthetic_checkpoint(dims, seed=8)
    rng = np.random.default_rng(51)
    common = rng.integers(0, dims.vocab, size=70, dtype=np.int32)                 # 4 full blocks + 6 tokens
    prompts = [np.concatenate([common, rng.integers(0, dims.vocab, size=int(n), dtype=np.int32)]) for n in (1, 9, 30, 58, 3)]
    prompts.append(common[:64].copy())                                            # exactly 4 blocks: the last one is recomputed
    other = [rng.integers(0, dims.vocab, size=90, dtype=np.int32) for _ in range(3)]
    kv_tok_bytes = 2 * dims.n_layers * dims.n_kv_heads * dims.head_dim * 2

    def run(flag, kv_blocks):
        monkeypatch.setenv("VGATE_B200_PREFIX_CACHE", flag)
        eng = Engine(dims, max_seqs=8, max_model_len=256, max_batch_tokens=512, kv_bytes=kv_tok_bytes * 16 * kv_blocks).load_checkpoint(ckpt)
        try:
            first = eng.wait(eng.submit(prompts[0], 0.0, 1.0, 6), 6)[0]           # registers the common prefix
            rids = [eng.submit(p, 0.0, 1.0, 6) for p in prompts[1:] + other]
            outs = [first] + [eng.wait(r, 6)[0] for r in rids]
            again = eng.wait(eng.submit(prompts[2], 0.0, 1.0, 6), 6)[0]           # after other traffic (possibly evicted)
            st = eng.stats()
        finally:
            eng.close()
        return outs, again, st

    base, base_again, st0 = run("0", 200)
    outs, again, st1 = run("1", 200)
    for a, b in zip(outs, base):
        assert np.array_equal(a, b)
    assert np.array_equal(again, base_again) and np.array_equal(again, outs[2])
    assert st0["prefix_hit_tokens"] == 0
    # prompts[1..4] share the 4 blocks of the common prefix, the 64-token prompt 3 of them (one token must be computed),
    # and the repeat of prompts[2] (100 tokens) all 6 full blocks its first run registered
    assert st1["prefix_hit_tokens"] == 4 * 64 + 48 + 96
    assert st1["tokens_prefill"] == st0["tokens_prefill"] - st1["prefix_hit_tokens"]
    assert st1["kv_blocks_free"] == st1["kv_blocks_total"]
    # a pool of 40 blocks (a 90-token request reserves 6-7): registered blocks are evicted to make room, results hold
    outs_small, again_small, st2 = run("1", 40)
    for a, b in zip(outs_small, base):
        assert np.array_equal(a, b)
    assert np.array_equal(again_small, base_again)
    assert st2["kv_blocks_free"] == st2["kv_blocks_total"] == 40
