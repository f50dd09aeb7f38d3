"""Worker-side additions a b200 deployment mounts on the UNMODIFIED worker app (``main.app``) from outside the
reference tree (``tools/vgate_launch.py`` calls :func:`install`):

* ``POST /internal/stream`` -- the streaming twin of ``/internal/generate`` (``vgate/worker_api.py:85-139``):
  newline-delimited JSON, one ``{"delta", "num_tokens"}`` object per engine step, exactly what
  ``backend.stream_generate`` yields (``vgate/backends/base.py:65-91``).  It is what
  ``vgate_b200.remote.CoalescingRemoteBackend.stream_generate`` consumes on the gateway, which is how SSE
  requests cross the worker hop (the reference answers 501 there, ``main.py:429-436``).  When the gateway drops
  the connection the response task is cancelled, the backend generator is closed and its ``finally`` aborts the
  engine request.
* ``POST /internal/generate_nd`` -- many prompts in, newline-delimited results out IN COMPLETION ORDER: every prompt
  is its own ``backend.generate([prompt], sp)`` call on the executor (the call ``/internal/generate`` makes for a
  single-prompt body), and its ``{"i": index, "result": {...}}`` line is written the moment it finishes.  One HTTP
  exchange carries up to 64 requests without any of them waiting for the slowest (what a multi-prompt
  ``/internal/generate`` body does); items may differ in sampling parameters.  Consumed by
  ``CoalescingRemoteBackend`` (``VGATE_B200_COALESCE_MODE=nd``, the default of that hop).
* ``GET /health`` that reflects engine liveness: the reference's route is unconditional (``main.py:288-295``), so
  a worker whose engine died keeps passing the gateway's probes (``vgate/health_checker.py:307-312``) and k8s'.
  The original handler still produces the body; a backend that reports itself unhealthy turns it into a 503.
"""
import asyncio
import json
from typing import Any, Dict, List

from fastapi import APIRouter, HTTPException, Request
from fastapi.responses import JSONResponse, StreamingResponse
from pydantic import BaseModel, Field


class StreamRequest(BaseModel):
    prompt: str = Field(..., min_length=1)
    sampling_params: Dict[str, Any] = Field(default_factory=dict)


class ManyItem(BaseModel):
    prompt: str = Field(..., min_length=1)
    sampling_params: Dict[str, Any] = Field(default_factory=dict)


class ManyRequest(BaseModel):
    items: List[ManyItem] = Field(..., min_length=1)


def _params(backend, sp: Dict[str, Any]):
    return backend.create_sampling_params(temperature=float(sp.get("temperature", 0.7)), top_p=float(sp.get("top_p", 0.9)),
                                          max_tokens=int(sp.get("max_tokens", 256)))


def install(app, main_module) -> None:
    from vgate import worker_api
    router = APIRouter()

    @router.post("/internal/stream", summary="Internal streaming generate (worker role)")
    async def internal_stream(req: StreamRequest, request: Request):
        engine = worker_api._engine
        if engine is None:
            raise HTTPException(status_code=503, detail="Worker engine not initialized")
        backend = engine.backend
        sp = _params(backend, req.sampling_params)

        async def lines():
            agen = backend.stream_generate(req.prompt, sp)
            try:
                async for piece in agen:
                    yield json.dumps(piece) + "\n"
            except Exception as exc:  # noqa: BLE001 -- the 200 is already on the wire: report in band (main.py:377-396)
                yield json.dumps({"error": type(exc).__name__}) + "\n"
            finally:
                await agen.aclose()             # -> B200Backend.stream_generate's finally -> engine.abort

        return StreamingResponse(lines(), media_type="application/x-ndjson")

    @router.post("/internal/generate_nd", summary="Internal generate, results as they complete (worker role)")
    async def internal_generate_nd(req: ManyRequest):
        engine = worker_api._engine
        if engine is None:
            raise HTTPException(status_code=503, detail="Worker engine not initialized")
        backend = engine.backend
        if not getattr(backend, "supports_concurrent_calls", False):     # worker_api.py:60: such a backend is called under a lock
            raise HTTPException(status_code=501, detail="backend does not take concurrent generate() calls")
        loop = asyncio.get_running_loop()

        async def one(i: int, item: ManyItem) -> str:
            try:
                res = await loop.run_in_executor(None, backend.generate, [item.prompt], _params(backend, item.sampling_params))
                return json.dumps({"i": i, "result": res[0]}) + "\n"
            except Exception as exc:  # noqa: BLE001 -- in band: the 200 is on the wire, the other items go on
                return json.dumps({"i": i, "error": type(exc).__name__}) + "\n"

        async def lines():
            tasks = [asyncio.ensure_future(one(i, it)) for i, it in enumerate(req.items)]
            try:
                for fut in asyncio.as_completed(tasks):
                    yield await fut
            finally:
                for t in tasks:                   # gateway gone: stop waiting (a generate() already on a thread runs to its end)
                    t.cancel()

        return StreamingResponse(lines(), media_type="application/x-ndjson")

    app.include_router(router)

    # /health: keep the reference's body, add the liveness check
    for i, route in enumerate(list(app.router.routes)):
        if getattr(route, "path", None) == "/health" and "GET" in getattr(route, "methods", ()):
            original = route.endpoint
            del app.router.routes[i]

            @app.get("/health", summary="Health Check")
            async def health():
                engine = worker_api._engine
                backend = getattr(engine, "backend", None)
                check = getattr(backend, "healthy", None)
                if callable(check):
                    ok, why = check()
                    if not ok:
                        return JSONResponse(status_code=503, content={"status": "unhealthy", "detail": why})
                return await original()
            break
