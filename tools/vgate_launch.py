"""Start an UNMODIFIED V-Gate `main.py` (gateway or worker role, chosen by VGATE_ROLE as usual) with the things a b200
deployment needs from outside the reference tree (hard part 1):

  * `engine_type: b200` registered (vgate_b200.register = the two maintainer edits of INTEGRATION.md section 1,
    applied at import time);
  * a default executor large enough for the decode batch.  Every in-flight request parks one thread of the
    asyncio default executor -- `run_in_executor(None, ...)` at vgate/batcher.py:374 on the gateway and
    vgate/worker_api.py:112 on the worker -- and the default pool is min(32, cpu + 4) threads, the ceiling
    the reference measured at benchmarks/results/scaling.md:161-187.  A worker's B200Backend.load_model
    enlarges its own loop's pool; the gateway never loads a backend of ours, so it is done here
    (B200_EXECUTOR_THREADS);
  * on a worker: /internal/stream and a /health that reflects engine liveness (vgate_b200.worker_stream);
  * on a gateway, optionally, several server processes behind the one listening socket (B200_GATEWAY_PROCESSES=K,
    uvicorn's multi-process mode).  One V-Gate gateway process is one asyncio event loop: the reference measured its
    ceiling at 180-290 requests/s (scaling.md:161-181), about what ONE B200 worker finishes.  K processes run K copies
    of the same unmodified app on the same port; admission limit, result cache and in-flight dedup are then per
    process (like K gateway replicas behind a load balancer, the reference's k8s deployment shape).

Run from the reference checkout (cwd = the directory that holds main.py):
    VGATE_ROLE=gateway B200_EXECUTOR_THREADS=600 python /path/to/tools/vgate_launch.py
"""
import asyncio
import os
import sys
from concurrent.futures import ThreadPoolExecutor

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.getcwd(), REPO, os.path.join(REPO, "tools")):
    if p not in sys.path:
        sys.path.insert(0, p)


def _load_main():
    import vgate_b200.register  # noqa: F401  (before main builds its config / engine)
    import main                 # the reference's module, untouched
    if main.IS_WORKER and os.environ.get("VGATE_B200_WORKER_EXTRAS", "1") != "0":
        from vgate_b200.worker_stream import install
        install(main.app, main)                   # /internal/stream and a /health that reflects engine liveness
    return main


def create_app():
    """ASGI app factory for uvicorn's multi-process mode: main.app behind a shim that sizes the serving loop's default
    executor when the process's event loop first runs a request or lifespan message."""
    main = _load_main()
    n = int(os.environ.get("B200_EXECUTOR_THREADS", "0"))
    state = {"done": n <= 0}

    async def app(scope, receive, send):
        if not state["done"]:
            state["done"] = True
            asyncio.get_running_loop().set_default_executor(ThreadPoolExecutor(n, thread_name_prefix="vgate-call"))
        await main.app(scope, receive, send)
    return app


KEEP_ALIVE_S = int(os.environ.get("B200_KEEP_ALIVE_S", "65"))


async def _serve():
    n = int(os.environ.get("B200_EXECUTOR_THREADS", "0"))
    if n > 0:
        asyncio.get_running_loop().set_default_executor(ThreadPoolExecutor(n, thread_name_prefix="vgate-call"))
    main = _load_main()
    import uvicorn
    # keep-alive longer than the gateway hop's: RemoteBackend's httpx pool expires idle connections after 5 s, uvicorn's
    # default closes them after 5 s too, and a request sent into that race dies with "failed mid-request: ReadError"
    # (seen at 1 of 320 requests with dry-run workers).  With 65 s the client always gives a connection up first.
    cfg = uvicorn.Config(main.app, host=main.config.server.host, port=main.config.server.port, log_level="warning",
                         timeout_keep_alive=KEEP_ALIVE_S)
    await uvicorn.Server(cfg).serve()


if __name__ == "__main__":
    procs = int(os.environ.get("B200_GATEWAY_PROCESSES", "1"))
    if procs > 1 and os.environ.get("VGATE_ROLE", "gateway") != "worker":
        import uvicorn
        from vgate.config import get_config
        c = get_config()
        uvicorn.run("vgate_launch:create_app", factory=True, host=c.server.host, port=c.server.port, workers=procs,
                    log_level="warning", timeout_keep_alive=KEEP_ALIVE_S)
    elif os.environ.get("B200_PROFILE"):          # where does a gateway / worker process spend its CPU (event-loop thread only)
        import cProfile
        import pstats
        prof = cProfile.Profile()

        def dump():
            prof.disable()
            with open(os.environ["B200_PROFILE"], "w") as f:
                pstats.Stats(prof, stream=f).sort_stats("tottime").print_stats(45)
            prof.enable()
            asyncio.get_running_loop().call_later(5.0, dump)

        async def profiled():
            asyncio.get_running_loop().call_later(5.0, dump)      # rewritten every 5 s: the process is ended by a signal
            await _serve()
        prof.enable()
        asyncio.run(profiled())
    else:
        asyncio.run(_serve())
