"""The reference's own callers of the hot path, unmodified, driving the REAL engine on a B200 (SURVEY.md section 8 rows
a9, a10): the worker route ``POST /internal/generate`` (vgate/worker_api.py:85-139) and the gateway-side
``RequestBatcher.submit -> _sync_inference -> backend.generate([prompt])`` (vgate/batcher.py:179-438).  Token ids are
checked end to end through the reversible detokeniser (the batcher drops ``token_ids``, batcher.py:433-438).
Needs a V-Gate checkout (/root/reference here, baseline/_ref on a GPU box: tools/bench_gateway.py --stage)."""
import asyncio
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REFERENCE = next((p for p in ("/root/reference", os.path.join(ROOT, "baseline", "_ref")) if os.path.isdir(os.path.join(p, "vgate"))), None)
pytestmark = [pytest.mark.gpu, pytest.mark.skipif(REFERENCE is None, reason="no V-Gate checkout on this machine")]

PRESET = "qwen2.5-1.5b-awq-l2-v8k"


@pytest.fixture
def vgate_b200_engine(monkeypatch):
    monkeypatch.syspath_prepend(REFERENCE)
    monkeypatch.delenv("VGATE_DRY_RUN", raising=False)
    monkeypatch.delenv("VGATE_B200_REMOTE", raising=False)
    for m in [m for m in sys.modules if m == "vgate" or m.startswith("vgate.")]:
        monkeypatch.delitem(sys.modules, m)
    monkeypatch.setenv("VGATE_B200_SYNTHETIC", PRESET)
    monkeypatch.setenv("VGATE_B200_MAX_SEQS", "32")
    sys.modules.pop("vgate_b200.register", None)
    import vgate_b200.register  # noqa: F401
    from vgate.config import ModelConfig, WorkerConfig
    from vgate.engine import VGateEngine
    eng = VGateEngine(ModelConfig(engine_type="b200", max_model_len=256), WorkerConfig())
    yield eng
    eng.backend.shutdown()


def _direct_tokens(backend, prompts, max_tokens):
    """The tokens the engine produces for these prompt strings through its own batch entry point."""
    ids = np.array([backend.tokenizer.encode(p) for p in prompts], dtype=np.int32)
    toks, _, _ = backend.engine.generate_batch(ids, max_tokens)
    return toks.tolist()


def test_internal_generate_route_serves_the_real_engine(vgate_b200_engine):
    fastapi = pytest.importorskip("fastapi")
    from fastapi.testclient import TestClient
    from vgate import worker_api
    from vgate_b200.tokenizer import ByteTokenizer
    eng = vgate_b200_engine
    prompts = [f"User: question {i:04d} about distributed systems\nAssistant:" for i in range(5)]     # equal lengths
    want = _direct_tokens(eng.backend, prompts, 12)
    app = fastapi.FastAPI()
    app.include_router(worker_api.router)
    worker_api.set_engine(eng)
    try:
        with TestClient(app) as client:
            r = client.post("/internal/generate", json={"prompts": prompts,
                                                        "sampling_params": {"temperature": 0.0, "top_p": 1.0, "max_tokens": 12}})
            assert r.status_code == 200, r.text
            res = r.json()["results"]
            assert [x["token_ids"] for x in res] == want
            for x in res:
                assert ByteTokenizer.parse(x["text"]) == x["token_ids"] and x["num_tokens"] == 12
                assert x["metrics"]["ttft"] > 0 and x["metrics"]["gen_time"] >= 0
            # the wire defaults (temperature 0.7, top_p 0.9, max_tokens 256; worker_api.py:70-73) reach the sampler
            r = client.post("/internal/generate", json={"prompts": prompts[:1], "sampling_params": {"max_tokens": 7}})
            assert r.status_code == 200 and r.json()["results"][0]["num_tokens"] == 7
    finally:
        worker_api.set_engine(None)


def test_request_batcher_fans_out_to_the_real_engine(vgate_b200_engine):
    from vgate.batcher import RequestBatcher
    from vgate_b200.tokenizer import ByteTokenizer
    eng = vgate_b200_engine
    prompts = [f"User: question {i:04d} about distributed systems\nAssistant:" for i in range(24)]
    want = _direct_tokens(eng.backend, prompts, 10)

    async def main():
        batcher = RequestBatcher(engine=eng, max_batch_size=32)
        assert batcher._serialize_inference is False and batcher.max_concurrent_inferences == 32
        await batcher.start()
        try:
            res = await asyncio.gather(*[batcher.submit(p, max_tokens=10, temperature=0.0, top_p=1.0) for p in prompts])
            again = await batcher.submit(prompts[3], max_tokens=10, temperature=0.0, top_p=1.0)      # served by the cache
            return res, again, batcher.get_metrics()
        finally:
            await batcher.stop()

    res, again, metrics = asyncio.run(main())
    for r, w in zip(res, want):
        assert ByteTokenizer.parse(r["text"]) == w                 # token ids survive as text only (batcher.py:433-438)
        assert r["total_tokens"] == 10 and r["ttft"] > 0 and r["tpot"] >= 0
    assert again["text"] == res[3]["text"]
    assert metrics["total_requests"] >= 24                         # the repeat was answered by the result cache
    st = eng.backend.engine.stats()
    assert st["kv_blocks_free"] == st["kv_blocks_total"]
