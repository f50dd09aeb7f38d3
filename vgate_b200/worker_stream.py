"""Worker-side additions a b200 deployment mounts on the UNMODIFIED worker app (``main.app``) from outside the
reference tree (``tools/vgate_launch.py`` calls :func:`install`):

* ``POST /internal/stream`` -- the streaming twin of ``/internal/generate`` (``vgate/worker_api.py:85-139``):
  newline-delimited JSON, one ``{"delta", "num_tokens"}`` object per engine step, exactly what
  ``backend.stream_generate`` yields (``vgate/backends/base.py:65-91``).  It is what
  ``vgate_b200.remote.CoalescingRemoteBackend.stream_generate`` consumes on the gateway, which is how SSE
  requests cross the worker hop (the reference answers 501 there, ``main.py:429-436``).  When the gateway drops
  the connection the response task is cancelled, the backend generator is closed and its ``finally`` aborts the
  engine request.
* ``GET /health`` that reflects engine liveness: the reference's route is unconditional (``main.py:288-295``), so
  a worker whose engine died keeps passing the gateway's probes (``vgate/health_checker.py:307-312``) and k8s'.
  The original handler still produces the body; a backend that reports itself unhealthy turns it into a 503.
"""
import json
from typing import Any, Dict

from fastapi import APIRouter, HTTPException, Request
from fastapi.responses import JSONResponse, StreamingResponse
from pydantic import BaseModel, Field


class StreamRequest(BaseModel):
    prompt: str = Field(..., min_length=1)
    sampling_params: Dict[str, Any] = Field(default_factory=dict)


def install(app, main_module) -> None:
    from vgate import worker_api
    router = APIRouter()

    @router.post("/internal/stream", summary="Internal streaming generate (worker role)")
    async def internal_stream(req: StreamRequest, request: Request):
        engine = worker_api._engine
        if engine is None:
            raise HTTPException(status_code=503, detail="Worker engine not initialized")
        backend = engine.backend
        sp = backend.create_sampling_params(temperature=float(req.sampling_params.get("temperature", 0.7)),
                                            top_p=float(req.sampling_params.get("top_p", 0.9)),
                                            max_tokens=int(req.sampling_params.get("max_tokens", 256)))

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
