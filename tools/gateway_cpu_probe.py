"""CPU-only probe of the gateway level: N DRY-RUN workers (the reference's synthetic backend: latency_ms + 2 ms per
token, vgate/backends/base.py:51-62) behind the gateway, driven like tools/bench_gateway.py drives the b200 workers.
What the gateway / worker hop does at a given concurrency can be studied here without a GPU; the absolute rates are
this container's cores, not the GPU box's.
usage: python tools/gateway_cpu_probe.py [--workers 4] [--conc 256] [--rounds 4] [--remote coalesce] [--gateway-procs K]"""
import argparse
import asyncio
import importlib.util
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
spec = importlib.util.spec_from_file_location("bench_gateway", os.path.join(ROOT, "tools", "bench_gateway.py"))
bg = importlib.util.module_from_spec(spec)
spec.loader.exec_module(bg)


async def run(a):
    ref = bg.find_reference()
    wenv = lambda i: {"VGATE_DRY_RUN": "true", "VGATE_DRYRUN_SIMULATED_LATENCY_MS": str(a.latency_ms),
                      "B200_EXECUTOR_THREADS": str(a.conc + 16)}
    genv = {"VGATE_BATCH__MAX_BATCH_SIZE": str(a.conc), "VGATE_WORKER__TIMEOUT_SECONDS": "600",
            "B200_EXECUTOR_THREADS": str(a.conc + 64)}
    if a.remote:
        genv["VGATE_B200_REMOTE"] = a.remote
    if a.gateway_procs > 1:
        genv["B200_GATEWAY_PROCESSES"] = str(a.gateway_procs)
        genv["VGATE_BATCH__MAX_BATCH_SIZE"] = str(max(8, (a.conc + a.gateway_procs - 1) // a.gateway_procs * 2))
    for k in a.env:
        key, val = k.split("=", 1)
        genv[key] = val
    async with bg.Topology(ref, a.workers, wenv, genv, os.path.join(ROOT, "gpurun_out", "gateway_probe_logs")) as topo:
        await bg.gateway_level(ref, topo.base_url, a.conc, a.conc, a.max_tokens, bg.unique_prompts(a.conc, "warm"))
        res, stats = await bg.gateway_level(ref, topo.base_url, a.conc, a.conc * a.rounds, a.max_tokens,
                                            bg.unique_prompts(a.conc * a.rounds, "probe"))
        out = bg.summarise_gateway(res, stats)
    out.update(workers=a.workers, concurrency=a.conc, worker_hop=a.remote or "RemoteBackend (reference)", gateway_processes=a.gateway_procs,
               per_request_s=a.latency_ms / 1e3 + 0.002 * a.max_tokens)
    print(json.dumps(out))


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--workers", type=int, default=4)
    ap.add_argument("--conc", type=int, default=256)
    ap.add_argument("--rounds", type=int, default=4)
    ap.add_argument("--max-tokens", type=int, default=128)
    ap.add_argument("--latency-ms", type=int, default=100)
    ap.add_argument("--remote", default=None)
    ap.add_argument("--gateway-procs", type=int, default=1)
    ap.add_argument("--env", action="append", default=[], help="extra gateway environment KEY=VALUE")
    asyncio.run(run(ap.parse_args()))
