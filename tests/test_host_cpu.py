"""Host-side logic of the b200 backend on CPU: tokenizer, synthetic checkpoints, and the adapter
contract of vgate/backends/base.py:65-91 exercised against a stub engine (the real engine needs a B200;
tests/test_engine_gpu.py runs the same assertions on it)."""
import asyncio
import os
import sys
import threading
from types import SimpleNamespace

import numpy as np
import pytest

from vgate_b200.tokenizer import ByteTokenizer, IncrementalDetokenizer, load_tokenizer
from vgate_b200.weights import PRESETS, ModelDims, synthetic_checkpoint, linear_shapes

REFERENCE = "/root/reference"
has_reference = os.path.isdir(os.path.join(REFERENCE, "vgate"))


def test_byte_tokenizer_is_reversible():
    tok = ByteTokenizer()
    ids = tok.encode("User: héllo\nAssistant:")
    assert ids == list("User: héllo\nAssistant:".encode("utf-8"))
    text = tok.decode([5, 151935, 0])
    assert ByteTokenizer.parse(text) == [5, 151935, 0]
    assert tok.encode("") == [0]
    det = IncrementalDetokenizer(tok)
    assert det.push([1, 2]) + det.push([3]) == tok.decode([1, 2, 3])
    assert isinstance(load_tokenizer("Qwen/not-on-disk"), ByteTokenizer)


def test_synthetic_checkpoint_shapes_and_determinism():
    d = PRESETS["tiny-awq"]
    a, b = synthetic_checkpoint(d, seed=1), synthetic_checkpoint(d, seed=1)
    c = synthetic_checkpoint(d, seed=2)
    assert a.keys() == b.keys()
    assert all(np.array_equal(a[k], b[k]) for k in a)
    assert any(not np.array_equal(a[k], c[k]) for k in a)
    for name, K, N, has_bias in linear_shapes(d):
        p = "model.layers.0." + name
        assert a[p + ".qweight"].shape == (K, N // 8) and a[p + ".qweight"].dtype == np.int32
        assert a[p + ".qzeros"].shape == (K // d.group_size, N // 8)
        assert a[p + ".scales"].shape == (K // d.group_size, N) and a[p + ".scales"].dtype == np.float16
        assert (p + ".bias" in a) == has_bias
    full = PRESETS["qwen2.5-1.5b-awq"]
    assert (full.hidden, full.intermediate, full.n_layers, full.n_heads, full.n_kv_heads, full.vocab) == \
        (1536, 8960, 28, 12, 2, 151936)


class _StubEngine:
    """Stands in for vgate_b200.engine.Engine: echoes prompt-dependent token ids, honours max_tokens,
    records aborts.  Same method surface, no GPU."""
    instances = []

    def __init__(self, dims, **kw):
        self.kw, self.aborted, self.closed, self._next, self._reqs = kw, [], False, 1, {}
        self.lock = threading.Lock()
        _StubEngine.instances.append(self)

    def load_checkpoint(self, ckpt):
        self.n_tensors = len(ckpt)
        return self

    def submit(self, prompt_ids, temperature=0.0, top_p=1.0, max_tokens=16, seed=0):
        with self.lock:
            rid, self._next = self._next, self._next + 1
            self._reqs[rid] = {"toks": [(int(sum(prompt_ids)) + i) % 1000 for i in range(max_tokens)], "sent": 0,
                               "sp": (temperature, top_p, max_tokens, seed)}
        return rid

    def wait(self, rid, max_tokens):
        r = self._reqs.pop(rid)
        return np.array(r["toks"], dtype=np.int32), 0.01, 0.002 * len(r["toks"])

    def poll(self, rid, cap=64, timeout_ms=0):
        r = self._reqs[rid]
        n = min(2, len(r["toks"]) - r["sent"], cap)
        out = r["toks"][r["sent"]:r["sent"] + n]
        r["sent"] += n
        done = r["sent"] == len(r["toks"])
        if done:
            self._reqs.pop(rid)
        return np.array(out, dtype=np.int32), done

    def abort(self, rid):
        self.aborted.append(rid)

    def close(self):
        self.closed = True


@pytest.fixture
def backend(monkeypatch):
    import vgate_b200.engine as eng_mod
    monkeypatch.setattr(eng_mod, "Engine", _StubEngine)
    monkeypatch.setenv("VGATE_B200_SYNTHETIC", "tiny-awq")
    from vgate_b200 import B200Backend
    be = B200Backend()
    be.load_model(SimpleNamespace(model_id="Qwen/Qwen2.5-1.5B-Instruct-AWQ", max_model_len=512))
    yield be
    be.shutdown()


def test_backend_generate_contract(backend):
    # tests/test_backends.py:45-58 (result keys / types), vllm_backend.py:122-132 (normalisation)
    sp = backend.create_sampling_params(0.7, 0.9, 5)
    assert sp == {"temperature": 0.7, "top_p": 0.9, "max_tokens": 5}          # survives the worker wire (worker_api.py:99-103)
    res = backend.generate(["User: a\nAssistant:", "User: bb\nAssistant:"], sp)
    assert len(res) == 2                                                       # remote_backend.py:186-193
    for r in res:
        assert isinstance(r["text"], str) and isinstance(r["token_ids"], list)
        assert r["num_tokens"] == len(r["token_ids"]) == 5
        assert all(isinstance(t, int) for t in r["token_ids"])
        assert set(r["metrics"]) == {"ttft", "gen_time"} and all(isinstance(v, float) for v in r["metrics"].values())
        assert ByteTokenizer.parse(r["text"]) == r["token_ids"]
    assert res[0]["token_ids"] != res[1]["token_ids"]
    assert backend.supports_concurrent_calls is True
    assert getattr(backend, "supports_streaming", True) is True               # absence => True (main.py:429)
    import json
    json.dumps(res)                                                            # worker_api.py:139 serialises it


def test_generate_releases_every_request_when_one_fails(backend):
    """A failure in the middle of a multi-prompt call (worker_api.py:77 bodies) must not leave the other
    requests of that call decoding or parked in the engine's table: they are aborted (= released)."""
    eng = backend.engine
    sp = backend.create_sampling_params(0.0, 1.0, 4)
    real_wait, calls = eng.wait, []

    def flaky_wait(rid, max_tokens):
        calls.append(rid)
        if len(calls) == 2:
            raise RuntimeError("engine failure")
        return real_wait(rid, max_tokens)

    eng.wait = flaky_wait
    with pytest.raises(RuntimeError):
        backend.generate(["a", "b", "c", "d"], sp)
    # first request was waited on (released by wait); the failing one and the two never waited on are aborted
    assert sorted(eng.aborted) == sorted(calls[1:2] + [r for r in (calls[1] + 1, calls[1] + 2)])
    eng.wait = real_wait
    # a submit that fails half-way aborts what was already submitted
    eng.aborted.clear()
    real_submit, n = eng.submit, []

    def flaky_submit(*a, **k):
        n.append(1)
        if len(n) == 3:
            raise ValueError("prompt token id out of range")
        return real_submit(*a, **k)

    eng.submit = flaky_submit
    with pytest.raises(ValueError):
        backend.generate(["a", "b", "c"], sp)
    assert len(eng.aborted) == 2
    eng.submit = real_submit


def test_backend_accepts_attribute_style_sampling_params(backend):
    sp = SimpleNamespace(temperature=0.0, top_p=1.0, max_tokens=3, seed=11)   # like vllm.SamplingParams
    r = backend.generate(["x"], sp)[0]
    assert r["num_tokens"] == 3


def test_backend_stream_contract_and_abort(backend):
    # tests/test_backends.py:71-94: cumulative counts, deltas concatenate to generate() text
    sp = backend.create_sampling_params(0.0, 1.0, 7)
    full = backend.generate(["User: hi\nAssistant:"], sp)[0]

    async def collect(limit=None):
        out = []
        agen = backend.stream_generate("User: hi\nAssistant:", sp)
        async for c in agen:
            out.append(c)
            if limit and len(out) >= limit:
                await agen.aclose()                                            # client went away (main.py:372-376)
                break
        return out

    chunks = asyncio.run(collect())
    assert "".join(c["delta"] for c in chunks) == full["text"]
    counts = [c["num_tokens"] for c in chunks]
    assert counts == sorted(counts) and counts[-1] == 7
    eng = _StubEngine.instances[-1]
    assert eng.aborted == []
    asyncio.run(collect(limit=1))
    assert len(eng.aborted) == 1                                               # closed mid-stream -> engine.abort


def test_backend_is_reentrant_from_executor_threads(backend):
    # worker_api.py:112 / batcher.py:374 call generate() from many threads at once
    sp = backend.create_sampling_params(0.0, 1.0, 4)
    outs = [None] * 16

    def work(i):
        outs[i] = backend.generate([f"prompt {i}"], sp)[0]

    ts = [threading.Thread(target=work, args=(i,)) for i in range(16)]
    [t.start() for t in ts]
    [t.join() for t in ts]
    assert all(o["num_tokens"] == 4 for o in outs)
    assert len({tuple(o["token_ids"]) for o in outs}) > 1


def test_generate_before_load_model_raises():
    from vgate_b200 import B200Backend
    with pytest.raises(RuntimeError):
        B200Backend().generate(["x"], {"temperature": 0, "top_p": 1, "max_tokens": 1})


@pytest.mark.skipif(not has_reference, reason="reference checkout not present on this machine")
def test_satisfies_reference_protocol_and_factory(monkeypatch):
    """Against the real V-Gate sources: isinstance(InferenceBackend) (tests/test_backends.py:32-34) and
    engine_type 'b200' through the reference's own factory and config validator."""
    monkeypatch.syspath_prepend(REFERENCE)
    monkeypatch.delenv("VGATE_DRY_RUN", raising=False)
    for m in [m for m in sys.modules if m == "vgate" or m.startswith("vgate.")]:
        monkeypatch.delitem(sys.modules, m)
    import vgate_b200.engine as eng_mod
    monkeypatch.setattr(eng_mod, "Engine", _StubEngine)
    monkeypatch.setenv("VGATE_B200_SYNTHETIC", "tiny-awq")
    sys.modules.pop("vgate_b200.register", None)
    import vgate_b200.register  # noqa: F401
    from vgate.backends.base import InferenceBackend
    from vgate.config import ModelConfig, WorkerConfig
    from vgate.engine import VGateEngine
    from vgate_b200 import B200Backend
    assert isinstance(B200Backend(), InferenceBackend)
    with pytest.raises(Exception):
        ModelConfig(engine_type="nope")
    eng = VGateEngine(ModelConfig(engine_type="b200", max_model_len=256), WorkerConfig())
    assert type(eng.backend).__name__ == "B200Backend" and not eng.is_remote
    sp = eng.backend.create_sampling_params(0.0, 1.0, 3)
    assert eng.backend.generate(["User: x\nAssistant:"], sp)[0]["num_tokens"] == 3
    eng.backend.shutdown()


@pytest.mark.skipif(not has_reference, reason="reference checkout not present on this machine")
def test_reference_worker_http_route_serves_b200_backend(monkeypatch):
    """The worker-side caller (vgate/worker_api.py:85-139) over HTTP, unmodified, with the b200 backend bound:
    wire-dict sampling params, executor hop, JSON response; an engine failure becomes HTTP 500."""
    fastapi = pytest.importorskip("fastapi")
    from fastapi.testclient import TestClient
    monkeypatch.syspath_prepend(REFERENCE)
    monkeypatch.delenv("VGATE_DRY_RUN", raising=False)
    for m in [m for m in sys.modules if m == "vgate" or m.startswith("vgate.")]:
        monkeypatch.delitem(sys.modules, m)
    import vgate_b200.engine as eng_mod
    monkeypatch.setattr(eng_mod, "Engine", _StubEngine)
    monkeypatch.setenv("VGATE_B200_SYNTHETIC", "tiny-awq")
    sys.modules.pop("vgate_b200.register", None)
    import vgate_b200.register  # noqa: F401
    from vgate import worker_api
    from vgate.config import ModelConfig, WorkerConfig
    from vgate.engine import VGateEngine
    eng = VGateEngine(ModelConfig(engine_type="b200", max_model_len=256), WorkerConfig())
    app = fastapi.FastAPI()
    app.include_router(worker_api.router)
    worker_api.set_engine(eng)
    assert worker_api._serialize_inference is False                 # supports_concurrent_calls honoured
    try:
        with TestClient(app) as client:
            body = {"prompts": ["User: a\nAssistant:", "User: bb\nAssistant:"],
                    "sampling_params": {"temperature": 0.0, "top_p": 1.0, "max_tokens": 4}}
            r = client.post("/internal/generate", json=body)
            assert r.status_code == 200, r.text
            res = r.json()["results"]
            assert len(res) == 2
            for x in res:
                assert set(x) >= {"text", "token_ids", "num_tokens", "metrics"} and x["num_tokens"] == 4
                assert len(x["token_ids"]) == 4 and set(x["metrics"]) == {"ttft", "gen_time"}
            # engine failure -> HTTP 500 "Inference failed: <Type>" (worker_api.py:115-128)
            def boom(*a, **k):
                raise RuntimeError("device lost")
            monkeypatch.setattr(eng.backend, "generate", boom)
            r = client.post("/internal/generate", json=body)
            assert r.status_code == 500 and "RuntimeError" in r.json()["detail"]
    finally:
        worker_api.set_engine(None)
        eng.backend.shutdown()


def test_qwen_chat_template_rewrites_the_reference_prompt_format():
    """main.py:269-271 flattens a chat into "Role: content" lines + "Assistant:"; VGATE_B200_CHAT_TEMPLATE=qwen turns
    that back into Qwen2.5-Instruct's ChatML prompt (off by default: the reference hands the flat string to vLLM)."""
    from vgate_b200.tokenizer import qwen_chat_prompt
    flat = "System: Be brief.\nUser: first line\nsecond line: with a colon\nAssistant: earlier answer\nUser: again?\nAssistant:"
    want = ("<|im_start|>system\nBe brief.<|im_end|>\n"
            "<|im_start|>user\nfirst line\nsecond line: with a colon<|im_end|>\n"
            "<|im_start|>assistant\nearlier answer<|im_end|>\n"
            "<|im_start|>user\nagain?<|im_end|>\n"
            "<|im_start|>assistant\n")
    assert qwen_chat_prompt(flat) == want
    one = qwen_chat_prompt("User: hello\nAssistant:")
    assert one.startswith("<|im_start|>system\nYou are Qwen") and one.endswith("<|im_start|>user\nhello<|im_end|>\n<|im_start|>assistant\n")
    assert qwen_chat_prompt("a raw completion prompt") == "a raw completion prompt"          # not the chat format: untouched
    assert qwen_chat_prompt("no role here\nAssistant:") == "no role here\nAssistant:"
