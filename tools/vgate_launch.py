"""Start an UNMODIFIED V-Gate `main.py` (gateway or worker role, chosen by VGATE_ROLE as usual) with the two
things a b200 deployment needs from outside the reference tree (hard part 1):

  * `engine_type: b200` registered (vgate_b200.register = the two maintainer edits of INTEGRATION.md section 1,
    applied at import time);
  * a default executor large enough for the decode batch.  Every in-flight request parks one thread of the
    asyncio default executor -- `run_in_executor(None, ...)` at vgate/batcher.py:374 on the gateway and
    vgate/worker_api.py:112 on the worker -- and the default pool is min(32, cpu + 4) threads, the ceiling
    the reference measured at benchmarks/results/scaling.md:161-187.  A worker's B200Backend.load_model
    enlarges its own loop's pool; the gateway never loads a backend of ours, so it is done here.

Run from the reference checkout (cwd = the directory that holds main.py):
    VGATE_ROLE=gateway B200_EXECUTOR_THREADS=600 python /path/to/tools/vgate_launch.py
"""
import asyncio
import os
import sys
from concurrent.futures import ThreadPoolExecutor

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.getcwd(), REPO):
    if p not in sys.path:
        sys.path.insert(0, p)


async def _serve():
    n = int(os.environ.get("B200_EXECUTOR_THREADS", "0"))
    if n > 0:
        asyncio.get_running_loop().set_default_executor(ThreadPoolExecutor(n, thread_name_prefix="vgate-call"))
    import vgate_b200.register  # noqa: F401  (before main builds its config / engine)
    import main                 # the reference's module, untouched
    if main.IS_WORKER and os.environ.get("VGATE_B200_WORKER_EXTRAS", "1") != "0":
        from vgate_b200.worker_stream import install
        install(main.app, main)                   # /internal/stream and a /health that reflects engine liveness
    import uvicorn
    cfg = uvicorn.Config(main.app, host=main.config.server.host, port=main.config.server.port, log_level="warning")
    await uvicorn.Server(cfg).serve()


if __name__ == "__main__":
    asyncio.run(_serve())
