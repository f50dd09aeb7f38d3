"""The gateway-side worker hop for b200 workers (vgate_b200/remote.py) and the worker-side additions
(vgate_b200/worker_stream.py), on CPU, against the real V-Gate code when a checkout is present:
coalescing of concurrent single-prompt calls into multi-prompt /internal/generate bodies, the reference's retry
and failure accounting, streaming through the hop, /health reflecting engine liveness."""
import asyncio
import json
import os
import sys
import threading
import time
from types import SimpleNamespace

import pytest

REFERENCE = "/root/reference"
has_reference = os.path.isdir(os.path.join(REFERENCE, "vgate"))
pytestmark = pytest.mark.skipif(not has_reference, reason="reference checkout not present on this machine")


@pytest.fixture
def vgate_env(monkeypatch):
    monkeypatch.syspath_prepend(REFERENCE)
    for m in [m for m in sys.modules if m == "vgate" or m.startswith("vgate.")]:
        monkeypatch.delitem(sys.modules, m)
    sys.modules.pop("vgate_b200.remote", None)
    sys.modules.pop("vgate_b200.register", None)
    yield


def _backend(handler, endpoints=("http://w0", "http://w1"), **env):
    import httpx
    env.setdefault("VGATE_B200_COALESCE_MODE", "body")       # the reference's wire form unless a test asks for "nd"
    from vgate.config import WorkerConfig
    from vgate_b200.remote import CoalescingRemoteBackend
    for k, v in env.items():
        os.environ[k] = str(v)
    try:
        be = CoalescingRemoteBackend(WorkerConfig(endpoints=list(endpoints)))
    finally:
        for k in env:
            os.environ.pop(k, None)
    be._client.close()
    be._client = httpx.Client(transport=httpx.MockTransport(handler))
    return be


def test_concurrent_single_prompt_calls_share_bodies(vgate_env):
    import httpx
    bodies, lock = [], threading.Lock()

    def handler(request):
        body = json.loads(request.content)
        with lock:
            bodies.append((str(request.url), body))
        time.sleep(0.05)                                    # a generation in progress: later calls queue up meanwhile
        res = [{"text": "echo " + p, "token_ids": [1, 2], "num_tokens": 2, "metrics": {"ttft": 0.0, "gen_time": 0.0}}
               for p in body["prompts"]]
        return httpx.Response(200, json={"results": res})

    be = _backend(handler, VGATE_B200_COALESCE_SENDERS=2)
    try:
        sp = be.create_sampling_params(0.0, 1.0, 8)
        assert sp == {"temperature": 0.0, "top_p": 1.0, "max_tokens": 8}          # same wire dict as remote_backend.py:107-112
        out = [None] * 48
        def call(i):
            out[i] = be.generate([f"prompt {i}"], sp if i % 3 else be.create_sampling_params(0.0, 1.0, 16))[0]
        ts = [threading.Thread(target=call, args=(i,)) for i in range(48)]
        for t in ts:
            t.start()
        for t in ts:
            t.join()
        assert all(out[i]["text"] == f"echo prompt {i}" for i in range(48))        # demultiplexed to the right caller
        assert len(bodies) < 24 and sum(len(b["prompts"]) for _, b in bodies) == 48   # far fewer bodies than prompts
        assert be.bodies_sent == len(bodies) and be.prompts_sent == 48
        for _, b in bodies:                                                         # equal sampling params per body only
            want = 16 if all(int(p.split()[1]) % 3 == 0 for p in b["prompts"]) else 8
            assert b["sampling_params"]["max_tokens"] == want
            assert all((int(p.split()[1]) % 3 == 0) == (want == 16) for p in b["prompts"])
        assert {u for u, _ in bodies} == {"http://w0/internal/generate", "http://w1/internal/generate"}   # round-robin
        # a multi-prompt call keeps its order
        res = be.generate(["a", "b", "c"], sp)
        assert [r["text"] for r in res] == ["echo a", "echo b", "echo c"]
    finally:
        be.shutdown()


def test_nd_mode_releases_every_caller_with_its_own_line(vgate_env):
    """/internal/generate_nd: results arrive in completion order and each caller returns with its own line -- a slow
    prompt does not hold the fast ones of its body; mixed sampling parameters share a body; an unmodified worker (404)
    is served through the reference's route."""
    import httpx
    bodies, lock = [], threading.Lock()
    release_slow = threading.Event()

    class Lines(httpx.SyncByteStream):
        def __init__(self, items):
            self.items = items

        def __iter__(self):
            slow = [i for i, it in enumerate(self.items) if it["prompt"].startswith("slow")]
            for i, it in enumerate(self.items):
                if i not in slow:
                    yield (json.dumps({"i": i, "result": {"text": "echo " + it["prompt"], "token_ids": [], "num_tokens": it["sampling_params"]["max_tokens"],
                                                              "metrics": {}}}) + "\n").encode()
            release_slow.wait(5)
            for i in slow:
                yield (json.dumps({"i": i, "error": "RuntimeError"} if self.items[i]["prompt"] == "slow bad" else
                                  {"i": i, "result": {"text": "echo slow", "token_ids": [], "num_tokens": 1, "metrics": {}}}) + "\n").encode()

    def handler(request):
        body = json.loads(request.content)
        with lock:
            bodies.append((request.url.path, body))
        if request.url.host == "w1":                       # an unmodified worker
            if request.url.path == "/internal/generate_nd":
                return httpx.Response(404, json={"detail": "Not Found"})
            return httpx.Response(200, json={"results": [{"text": "ref " + p, "token_ids": [], "num_tokens": 0, "metrics": {}} for p in body["prompts"]]})
        return httpx.Response(200, stream=Lines(body["items"]))

    from vgate.backends.remote_backend import RemoteInferenceError
    be = _backend(handler, endpoints=("http://w0",), VGATE_B200_COALESCE_MODE="nd", VGATE_B200_COALESCE_SENDERS=1)
    try:
        done, errs = {}, {}
        def call(p, mt):
            try:
                done[p] = be.generate([p], be.create_sampling_params(0.0, 1.0, mt))[0]
            except Exception as exc:  # noqa: BLE001
                errs[p] = exc
        first = threading.Thread(target=call, args=("warm", 4))
        first.start(); first.join()
        ts = [threading.Thread(target=call, args=(p, mt)) for p, mt in (("slow one", 8), ("slow bad", 8), ("fast a", 8), ("fast b", 16), ("fast c", 32))]
        for t in ts:
            t.start()
            time.sleep(0.01)
        deadline = time.time() + 5
        while time.time() < deadline and not all(k in done for k in ("fast a", "fast b", "fast c")):
            time.sleep(0.01)
        assert all(k in done for k in ("fast a", "fast b", "fast c")) and "slow one" not in done    # released before the slow ones finish
        assert done["fast b"]["num_tokens"] == 16 and done["fast c"]["num_tokens"] == 32              # per-item sampling parameters
        release_slow.set()
        for t in ts:
            t.join()
        assert done["slow one"]["text"] == "echo slow"
        assert isinstance(errs["slow bad"], RemoteInferenceError)                                   # one item's failure is its caller's only
        assert all(path == "/internal/generate_nd" for path, _ in bodies)
        assert be.prompts_sent == 6 and be.bodies_sent < 6
    finally:
        be.shutdown()
    bodies.clear()
    be2 = _backend(handler, endpoints=("http://w1",), VGATE_B200_COALESCE_MODE="nd")
    try:
        assert be2.generate(["x"], be2.create_sampling_params(0.0, 1.0, 4))[0]["text"] == "ref x"     # fell back to /internal/generate
        assert be2.generate(["y"], be2.create_sampling_params(0.0, 1.0, 4))[0]["text"] == "ref y"
        assert [p for p, _ in bodies] == ["/internal/generate_nd", "/internal/generate", "/internal/generate"]   # asked once
    finally:
        be2.shutdown()


def test_prefix_affinity_routing_keeps_a_prefix_on_one_worker(vgate_env):
    import httpx
    seen = {}

    def handler(request):
        body = json.loads(request.content)
        for p in body["prompts"]:
            seen.setdefault(p[:40], set()).add(request.url.host)
        return httpx.Response(200, json={"results": [{"text": "", "token_ids": [], "num_tokens": 0, "metrics": {}} for _ in body["prompts"]]})

    be = _backend(handler, endpoints=("http://w0", "http://w1", "http://w2"), VGATE_B200_ROUTE="prefix", VGATE_B200_ROUTE_CHARS=40)
    try:
        sp = be.create_sampling_params(0.0, 1.0, 4)
        systems = [f"System prompt number {i:02d}, forty characters.."[:40] for i in range(12)]
        prompts = [s + f" question {j}" for s in systems for j in range(5)]
        ts = [threading.Thread(target=lambda p=p: be.generate([p], sp)) for p in prompts]
        for t in ts:
            t.start()
        for t in ts:
            t.join()
        assert all(len(hosts) == 1 for hosts in seen.values())            # a prefix class never changes worker
        assert len({next(iter(h)) for h in seen.values()}) >= 2           # and the classes spread over the pool
    finally:
        be.shutdown()


def test_retry_and_failure_policy_follow_the_reference(vgate_env):
    import httpx
    from vgate.backends.remote_backend import RemoteInferenceError
    from vgate.worker_registry import NoHealthyWorkersError
    calls = []

    def handler(request):
        calls.append(str(request.url))
        if "w0" in str(request.url):
            raise httpx.ConnectError("refused", request=request)     # never delivered: another worker may take it
        body = json.loads(request.content)
        if body["prompts"][0] == "bad":
            return httpx.Response(500, json={"detail": "Inference failed: RuntimeError"})
        return httpx.Response(200, json={"results": [{"text": "ok", "token_ids": [], "num_tokens": 0, "metrics": {}}
                                                     for _ in body["prompts"]]})

    be = _backend(handler)
    try:
        sp = be.create_sampling_params(0.7, 0.9, 4)
        assert be.generate(["x"], sp)[0]["text"] == "ok"
        assert any("w1" in c for c in calls)
        with pytest.raises(RemoteInferenceError):                   # an error status is final, not retried elsewhere
            be.generate(["bad"], sp)
    finally:
        be.shutdown()
    be2 = _backend(lambda request: (_ for _ in ()).throw(httpx.ConnectError("refused", request=request)))
    try:
        with pytest.raises(NoHealthyWorkersError):
            be2.generate(["x"], be2.create_sampling_params(0.7, 0.9, 4))
    finally:
        be2.shutdown()


def test_register_selects_the_coalescing_hop_for_a_remote_gateway(vgate_env, monkeypatch):
    monkeypatch.setenv("VGATE_B200_REMOTE", "coalesce")
    import vgate_b200.register  # noqa: F401
    from vgate.backends.base import InferenceBackend
    from vgate.config import ModelConfig, WorkerConfig
    from vgate.engine import VGateEngine
    eng = VGateEngine(ModelConfig(), WorkerConfig(endpoints=["http://127.0.0.1:1"]))
    try:
        assert type(eng.backend).__name__ == "CoalescingRemoteBackend" and eng.is_remote
        assert isinstance(eng.backend, InferenceBackend)
        assert getattr(eng.backend, "supports_streaming", True) is True          # the gateway will not answer 501
    finally:
        eng.backend.shutdown()
    monkeypatch.delenv("VGATE_B200_REMOTE")
    eng = VGateEngine(ModelConfig(), WorkerConfig(endpoints=["http://127.0.0.1:1"]))
    assert type(eng.backend).__name__ == "RemoteBackend"                         # untouched default
    eng.backend.shutdown()


def test_worker_health_reflects_engine_liveness_and_stream_route(vgate_env):
    fastapi = pytest.importorskip("fastapi")
    from fastapi.testclient import TestClient
    from vgate import worker_api
    from vgate_b200.worker_stream import install

    class Backend:
        ok = True
        def healthy(self):
            return (True, "") if self.ok else (False, "engine failed: CUDA error")
        def create_sampling_params(self, temperature, top_p, max_tokens):
            return {"max_tokens": max_tokens}
        async def stream_generate(self, prompt, sp):
            for i in range(sp["max_tokens"]):
                yield {"delta": f"{i} ", "num_tokens": i + 1}

    app = fastapi.FastAPI()

    @app.get("/health")
    async def health():                        # stands for main.py:288-295 (unconditional)
        return {"status": "ok"}

    install(app, SimpleNamespace())
    be = Backend()
    worker_api.set_engine(SimpleNamespace(backend=be))
    try:
        with TestClient(app) as client:
            assert client.get("/health").json() == {"status": "ok"}
            be.ok = False
            r = client.get("/health")
            assert r.status_code == 503 and "CUDA" in r.json()["detail"]
            be.ok = True
            with client.stream("POST", "/internal/stream", json={"prompt": "p", "sampling_params": {"max_tokens": 3}}) as r:
                pieces = [json.loads(l) for l in r.iter_lines() if l]
            assert pieces == [{"delta": "0 ", "num_tokens": 1}, {"delta": "1 ", "num_tokens": 2}, {"delta": "2 ", "num_tokens": 3}]
    finally:
        worker_api.set_engine(None)


def test_streaming_and_coalescing_through_real_processes(tmp_path):
    """Two unmodified dry-run workers (with the b200 worker extras mounted) behind one gateway that uses the
    coalescing hop: non-streaming requests succeed, and a streaming request -- 501 with the reference's
    RemoteBackend (main.py:429-436) -- arrives as SSE chunks whose concatenation is the non-streaming text."""
    import aiohttp
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))
    import bench_gateway as bg

    async def main():
        topo = bg.Topology(REFERENCE, 2, lambda i: {"VGATE_DRY_RUN": "true"},
                           {"VGATE_B200_REMOTE": "coalesce", "VGATE_BATCH__MAX_BATCH_SIZE": "64", "B200_EXECUTOR_THREADS": "80"},
                           str(tmp_path), plain_main=False, ready_timeout=60)
        async with topo:
            async with aiohttp.ClientSession() as s:
                async def chat(i, stream=False):
                    body = {"model": "m", "messages": [{"role": "user", "content": f"question {i}"}], "max_tokens": 16, "stream": stream}
                    async with s.post(topo.base_url + "/v1/chat/completions", json=body) as r:
                        if not stream:
                            return r.status, (await r.json())
                        text, done = "", False
                        async for raw in r.content:
                            line = raw.decode().strip()
                            if line == "data: [DONE]":
                                done = True
                            elif line.startswith("data:"):
                                text += json.loads(line[5:])["choices"][0]["delta"].get("content", "") or ""
                        return r.status, text, done
                res = await asyncio.gather(*[chat(i) for i in range(40)])
                assert all(st == 200 for st, _ in res)
                assert all(f"question {i}" in body["choices"][0]["message"]["content"] for i, (_, body) in enumerate(res))
                st, text, done = await chat(7, stream=True)
                assert st == 200 and done
                # (the dry-run backend streams word by word: base.py:121-136 -- whitespace is normalised)
                assert text.split() == res[7][1]["choices"][0]["message"]["content"].split()
    asyncio.run(main())
