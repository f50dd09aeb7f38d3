#!/usr/bin/env python
"""The hot path measured where north_star puts it: behind V-Gate.  Spawns UNMODIFIED reference processes
(`main.py`, role worker / gateway -- the bring-up of benchmarks/bench_scaling.py:238-272) and drives them with
the reference's own load generator (benchmarks/bench_load.py:160 run_load_test):

  level "worker_http"  POST /internal/generate on every worker (vgate/worker_api.py:85-139), multi-prompt
                       bodies (the wire format takes a list, worker_api.py:77) and single-prompt bodies
                       (what RemoteBackend sends today, remote_backend.py:150-153)
  level "gateway"      POST /v1/chat/completions on ONE gateway in front of the N workers: security
                       middleware, RequestBatcher admission / dedup / cache, WorkerRegistry round-robin,
                       RemoteBackend hop

  python tools/bench_gateway.py --config 1                  config #1: dry-run worker + gateway on CPU,
                                                            40 requests at concurrency 8 (no GPU)
  python tools/bench_gateway.py --gpus N [--batch 64]       config #3: one b200 worker per GPU
                                                            (CUDA_VISIBLE_DEVICES=i) behind one gateway
  python tools/bench_gateway.py --stage                     copy the reference's control plane to
                                                            baseline/_ref/ (git-ignored; it then travels to a GPU box)

The reference tree is looked up at $VGATE_REFERENCE_DIR, /root/reference, baseline/_ref -- nothing of it is
part of this repository.  Prints one JSON document; `summary` says which level binds.
"""
import argparse
import asyncio
import json
import os
import shutil
import signal
import socket
import statistics
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LAUNCH = os.path.join(ROOT, "tools", "vgate_launch.py")
STAGED = os.path.join(ROOT, "baseline", "_ref")
GATEWAY_PORT = 18110
WORKER_PORT_BASE = 18111


def find_reference():
    for cand in (os.environ.get("VGATE_REFERENCE_DIR"), "/root/reference", STAGED):
        if cand and os.path.exists(os.path.join(cand, "main.py")) and os.path.isdir(os.path.join(cand, "vgate")):
            return cand
    return None


def stage_reference():
    """baseline/_ref/ <- the reference's control plane, byte for byte (git-ignored: never part of the history)."""
    src = find_reference()
    if src is None or os.path.abspath(src) == os.path.abspath(STAGED):
        raise SystemExit("no reference checkout to stage from")
    os.makedirs(STAGED, exist_ok=True)
    for name in ("main.py", "config.yaml"):
        if os.path.exists(os.path.join(src, name)):
            shutil.copy2(os.path.join(src, name), os.path.join(STAGED, name))
    for d in ("vgate", "benchmarks"):
        dst = os.path.join(STAGED, d)
        shutil.rmtree(dst, ignore_errors=True)
        shutil.copytree(os.path.join(src, d), dst, ignore=shutil.ignore_patterns("__pycache__", "results", "*.pyc"))
    return STAGED


# ---- process management (same discipline as bench_scaling.py:92-130: own process group, ports checked) ----------
def _spawn(ref_dir, env_over, log_path, plain_main=False):
    env = os.environ.copy()
    env.update({"VGATE_SECURITY__ENABLED": "false", "VGATE_TRACING__ENABLED": "false", "VGATE_LOGGING__LEVEL": "WARNING",
                "VGATE_METRICS__ENABLED": "true", "PYTHONPATH": ROOT + os.pathsep + env.get("PYTHONPATH", "")})
    env.update(env_over)
    cmd = [sys.executable, "main.py"] if plain_main else [sys.executable, LAUNCH]
    return subprocess.Popen(cmd, cwd=ref_dir, env=env, stdout=open(log_path, "w"), stderr=subprocess.STDOUT,
                            start_new_session=True)


def _terminate(proc):
    if proc is None:
        return
    try:
        os.killpg(os.getpgid(proc.pid), signal.SIGTERM)
        proc.wait(timeout=15)
    except subprocess.TimeoutExpired:
        try:
            os.killpg(os.getpgid(proc.pid), signal.SIGKILL)
        except ProcessLookupError:
            pass
        proc.wait(timeout=10)
    except ProcessLookupError:
        pass


def _port_free(port):
    with socket.socket(socket.AF_INET, socket.SOCK_STREAM) as s:
        return s.connect_ex(("127.0.0.1", port)) != 0


async def _wait_healthy(url, proc, timeout_s):
    import aiohttp
    deadline = time.monotonic() + timeout_s
    async with aiohttp.ClientSession() as session:
        while time.monotonic() < deadline:
            if proc.poll() is not None:
                raise RuntimeError(f"{url}: process exited with code {proc.returncode} before becoming healthy")
            try:
                async with session.get(f"{url}/health", timeout=aiohttp.ClientTimeout(total=1)) as resp:
                    if resp.status == 200:
                        return
            except Exception:  # noqa: BLE001
                pass
            await asyncio.sleep(0.25)
    raise RuntimeError(f"{url} did not become healthy within {timeout_s}s")


class Topology:
    """One gateway in front of N workers (bench_scaling.py:191-272), torn down together."""

    def __init__(self, ref_dir, n_workers, worker_env, gateway_env, log_dir, plain_main=False, ready_timeout=300):
        self.ref, self.n, self.wenv, self.genv, self.logs = ref_dir, n_workers, worker_env, gateway_env, log_dir
        self.plain, self.ready_timeout = plain_main, ready_timeout
        self.gateway, self.workers = None, []

    @property
    def base_url(self):
        return f"http://127.0.0.1:{GATEWAY_PORT}"

    def endpoints(self):
        return [f"http://127.0.0.1:{WORKER_PORT_BASE + i}" for i in range(self.n)]

    async def __aenter__(self):
        try:
            busy = [p for p in [GATEWAY_PORT] + [WORKER_PORT_BASE + i for i in range(self.n)] if not _port_free(p)]
            if busy:
                raise RuntimeError(f"ports in use: {busy}")
            os.makedirs(self.logs, exist_ok=True)
            for i in range(self.n):
                env = {"VGATE_ROLE": "worker", "VGATE_SERVER__PORT": str(WORKER_PORT_BASE + i)}
                env.update(self.wenv(i))
                self.workers.append(_spawn(self.ref, env, os.path.join(self.logs, f"worker-{i}.log"), self.plain))
            for i, w in enumerate(self.workers):          # /health answers only after load_model
                await _wait_healthy(self.endpoints()[i], w, self.ready_timeout)
            # The gateway starts once the workers answer: a worker that is still loading its model fails the
            # gateway's health probes (vgate/health_checker.py), is taken out of rotation after failure_threshold
            # probes and needs success_threshold more rounds of 5 s to come back.
            env = {"VGATE_ROLE": "gateway", "VGATE_SERVER__PORT": str(GATEWAY_PORT),
                   "VGATE_WORKER__ENDPOINTS": json.dumps(self.endpoints())}
            env.update(self.genv)
            self.gateway = _spawn(self.ref, env, os.path.join(self.logs, "gateway.log"), self.plain)
            await _wait_healthy(self.base_url, self.gateway, 60)
            await asyncio.sleep(1.0)                       # the gateway's first health probe
        except BaseException:
            await self.__aexit__()
            raise
        return self

    async def restart_gateway(self, gateway_env):
        """The same workers behind a gateway started with other settings."""
        _terminate(self.gateway)
        for _ in range(100):
            if _port_free(GATEWAY_PORT):
                break
            await asyncio.sleep(0.1)
        env = {"VGATE_ROLE": "gateway", "VGATE_SERVER__PORT": str(GATEWAY_PORT),
               "VGATE_WORKER__ENDPOINTS": json.dumps(self.endpoints())}
        env.update(gateway_env)
        self.gateway = _spawn(self.ref, env, os.path.join(self.logs, "gateway-restarted.log"), self.plain)
        await _wait_healthy(self.base_url, self.gateway, 60)
        await asyncio.sleep(1.0)

    async def __aexit__(self, *exc):
        _terminate(self.gateway)
        for w in self.workers:
            _terminate(w)
        self.gateway, self.workers = None, []


def unique_prompts(count, tag, n_chars=111):
    """All-distinct contents of exactly n_chars ASCII characters: "User: " + content + "\\nAssistant:" is then
    6 + 111 + 11 = 128 byte-level tokens, and neither the cache nor in-flight dedup can
    stand in for work (bench_scaling.py:305-309)."""
    stamp = int(time.time() * 1000)
    out = []
    for i in range(count):
        head = f"{tag} {stamp} {i} "
        out.append((head + "Explain distributed systems concept number " + str(i) + " in one sentence. " * 4)[:n_chars].ljust(n_chars, "."))
    return out


def _pct(vals, p):
    if not vals:
        return None
    s = sorted(vals)
    return s[min(int(len(s) * p / 100), len(s) - 1)]


async def worker_http_level(endpoints, per_worker, max_tokens, rounds, multi_prompt):
    """Direct POST /internal/generate on every worker at once.  multi_prompt: one body carrying the worker's whole
    batch; else `per_worker` concurrent single-prompt bodies (what the gateway's RemoteBackend issues)."""
    import aiohttp
    sp = {"temperature": 0.0, "top_p": 1.0, "max_tokens": max_tokens}
    tokens, tpots, failures = 0, [], 0
    timeout = aiohttp.ClientTimeout(total=600)
    conn = aiohttp.TCPConnector(limit=0)
    async with aiohttp.ClientSession(timeout=timeout, connector=conn) as session:
        async def post(url, prompts):
            nonlocal tokens, failures
            try:
                async with session.post(f"{url}/internal/generate", json={"prompts": prompts, "sampling_params": sp}) as r:
                    if r.status != 200:
                        failures += len(prompts)
                        return
                    body = await r.json()
            except Exception:  # noqa: BLE001
                failures += len(prompts)
                return
            for res in body["results"]:
                tokens += res["num_tokens"]
                gen = (res.get("metrics") or {}).get("gen_time")                  # the dry-run backend reports none
                if res["num_tokens"] and gen is not None:
                    tpots.append(gen / res["num_tokens"])                            # vgate/batcher.py:431

        async def one_round(tag):
            jobs = []
            for w, url in enumerate(endpoints):
                prompts = ["User: " + c + "\nAssistant:" for c in unique_prompts(per_worker, f"{tag}-w{w}")]
                if multi_prompt:
                    jobs.append(post(url, prompts))
                else:
                    jobs += [post(url, [p]) for p in prompts]
            await asyncio.gather(*jobs)

        await one_round("warm")
        tokens, tpots, failures = 0, [], 0
        t0 = time.perf_counter()
        for r in range(rounds):
            await one_round(f"r{r}")
        wall = time.perf_counter() - t0
    n_req = len(endpoints) * per_worker * rounds
    return {"tok_per_s": tokens / wall, "req_per_s": n_req / wall, "tpot_ms_p50": (_pct(tpots, 50) or 0) * 1e3,
            "requests": n_req, "failures": failures, "wall_s": wall,
            "bodies": "one multi-prompt body per worker" if multi_prompt else "one prompt per body"}


async def gateway_level(ref_dir, base_url, concurrency, total_requests, max_tokens, prompts, client_pool=100):
    sys.path.insert(0, ref_dir)
    import benchmarks.bench_load as bl                    # the reference's load generator, as is ...
    import aiohttp

    class _Aiohttp:
        """... except, on request, for its connection pool: run_load_test opens aiohttp.ClientSession() with aiohttp's
        default of 100 connections, so a concurrency above 100 queues inside the CLIENT (measured: N = 2 and N = 4
        workers both stopped at ~100 requests in flight, 190-225 req/s, whatever the gateway did).  client_pool = 0:
        same generator, pool sized to the requested concurrency.  The default stays 100 -- which is also the
        connection limit of the reference gateway's worker hop (httpx defaults, remote_backend.py:87); a single gateway
        process driven beyond it lost requests."""
        def __getattr__(self, name):
            return getattr(aiohttp, name)

        @staticmethod
        def ClientSession(*a, **kw):
            kw.setdefault("connector", aiohttp.TCPConnector(limit=client_pool))
            return aiohttp.ClientSession(*a, **kw)

    saved = bl.aiohttp
    bl.aiohttp = _Aiohttp() if client_pool != 100 else saved
    try:
        res = await bl.run_load_test(base_url=base_url, concurrency=concurrency, total_requests=total_requests, prompts=prompts,
                                     max_tokens=max_tokens)
    finally:
        bl.aiohttp = saved
    async with aiohttp.ClientSession() as session:
        async with session.get(f"{base_url}/stats") as r:
            stats = await r.json()
    return res, stats


def summarise_gateway(res, stats):
    out = {"tok_per_s": res["throughput"]["tokens_per_second"], "req_per_s": res["throughput"]["requests_per_second"],
           "latency_p50_s": res["latency"]["p50_s"], "latency_p95_s": res["latency"]["p95_s"], "failures": res["failures"],
           "requests": res["config"]["total_requests"], "wall_s": res["wall_time_s"]}
    b = res.get("batching", {})
    c = res.get("cache", {})
    out["deduplicated"] = b.get("deduplicated")
    out["cache_hits"] = c.get("hits")
    out["avg_tpot_ms"] = 1e3 * stats.get("batcher", {}).get("avg_tpot_s", 0)
    return out


# ---- configurations ---------------------------------------------------------------------------------------
async def run_config1(ref_dir, log_dir, latency_ms=None):
    """Config #1: the reference untouched -- dry-run worker behind the gateway, 64 unique prompts
    (benchmarks/run_report.py:63-65), 40 requests at concurrency 8, max_tokens 64.  latency_ms switches on the
    dry-run backend's synthetic cost (latency_ms + 2 ms x max_tokens per call, vgate/backends/base.py:51-62), the
    setting of the reference's own dry-run baseline (benchmarks/results/baseline.md:9-35)."""
    wenv = {"VGATE_DRY_RUN": "true"}
    if latency_ms is not None:
        wenv["VGATE_DRYRUN_SIMULATED_LATENCY_MS"] = str(latency_ms)
    topo = Topology(ref_dir, 1, lambda i: dict(wenv), {}, log_dir, plain_main=True, ready_timeout=60)
    async with topo:
        prompts = [f"Explain distributed systems concept #{i} in one sentence." for i in range(64)]
        res, stats = await gateway_level(ref_dir, topo.base_url, 8, 40, 64, prompts)
    s = summarise_gateway(res, stats)
    assert s["failures"] == 0, s
    return {"config": "#1 dry-run worker behind the gateway on CPU, 40 requests at concurrency 8, max_tokens 64"
                      + (f", synthetic latency {latency_ms} ms + 2 ms/token" if latency_ms is not None else ""),
            "cores": os.cpu_count(), "gateway": s, "raw": res}


async def run_b200(ref_dir, log_dir, n_gpus, batch, max_tokens, rounds, preset, levels, devices=None, remote=None, gateway_procs=1,
                   extra_gateway_procs=0, client_pool=100):
    threads = n_gpus * batch + 64
    devices = devices or [str(i) for i in range(n_gpus)]

    def wenv(i):
        if os.environ.get("B200_GATEWAY_DRYRUN_WORKERS"):      # harness check on a box without GPUs: the reference's synthetic backend
            return {"VGATE_DRY_RUN": "true", "VGATE_DRYRUN_SIMULATED_LATENCY_MS": os.environ["B200_GATEWAY_DRYRUN_WORKERS"],
                    "B200_EXECUTOR_THREADS": str(batch + 16)}
        return {"CUDA_VISIBLE_DEVICES": devices[i], "VGATE_MODEL__ENGINE_TYPE": "b200", "VGATE_DRY_RUN": "false",
                "VGATE_B200_SYNTHETIC": preset, "VGATE_B200_MAX_SEQS": str(max(batch, 64)),
                "VGATE_MODEL__MAX_MODEL_LEN": str(128 + max_tokens + 16)}

    genv = {"VGATE_BATCH__MAX_BATCH_SIZE": str(n_gpus * batch),        # admission above N x batch (scaling.md:77-85)
            "VGATE_WORKER__TIMEOUT_SECONDS": "600", "B200_EXECUTOR_THREADS": str(threads)}
    if remote:                                     # "coalesce": vgate_b200.remote.CoalescingRemoteBackend as the worker hop
        genv["VGATE_B200_REMOTE"] = remote
    if gateway_procs > 1:                          # K server processes of the unmodified gateway app on the one port
        genv["B200_GATEWAY_PROCESSES"] = str(gateway_procs)
        genv["VGATE_BATCH__MAX_BATCH_SIZE"] = str(max(8, (n_gpus * batch + gateway_procs - 1) // gateway_procs * 2))
    out = {"config": f"#3 {preset}: {n_gpus} b200 worker(s), one per GPU, behind one gateway; {batch} concurrent "
                     f"128-token prompts per GPU, greedy-by-wire max_tokens {max_tokens}",
           "n_gpus": n_gpus, "batch_per_gpu": batch, "cores": os.cpu_count(), "worker_hop": remote or "RemoteBackend (reference)", "gateway_processes": gateway_procs}
    async with Topology(ref_dir, n_gpus, wenv, genv, log_dir) as topo:
        if "worker_http" in levels:
            out["worker_http_multi"] = await worker_http_level(topo.endpoints(), batch, max_tokens, rounds, True)
            out["worker_http_single"] = await worker_http_level(topo.endpoints(), batch, max_tokens, rounds, False)
        if "gateway" in levels:
            conc = n_gpus * batch

            async def measure(tag, client_pool=100):
                # bounded: a gateway that loses requests would otherwise hold the bench for the generator's 300 s timeout
                warm = unique_prompts(conc, tag + "-warm")
                await asyncio.wait_for(gateway_level(ref_dir, topo.base_url, conc, conc, max_tokens, warm, client_pool), 60)
                prompts = unique_prompts(conc * rounds, tag)
                res, stats = await asyncio.wait_for(gateway_level(ref_dir, topo.base_url, conc, conc * rounds, max_tokens, prompts, client_pool), 90)
                g = summarise_gateway(res, stats)
                # the wire carries temperature 0.7 / top_p 0.9 by default (main.py request defaults): bench_load cannot
                # ask for greedy, so the gateway level samples; token counts are what is measured
                g["note"] = "bench_load.py sends the API's default sampling (temperature 0.7, top_p 0.9)"
                g["client_connections"] = ("sized to the concurrency" if client_pool == 0 else
                                           f"bench_load.py default pool ({client_pool}): at most {min(conc, client_pool)} requests in flight")
                return g

            async def guarded(tag, client_pool=100):
                try:
                    return await measure(tag, client_pool)
                except asyncio.TimeoutError:
                    return {"error": "timed out: the gateway stopped answering part of the requests", "concurrency": conc,
                            "client_connections": client_pool or "sized to the concurrency"}
            out["gateway"] = await guarded("gw", client_pool)
            if extra_gateway_procs and extra_gateway_procs > 1:
                # the same workers behind the same gateway app served by K processes on the one port
                g2 = dict(genv)
                g2["B200_GATEWAY_PROCESSES"] = str(extra_gateway_procs)
                g2["VGATE_BATCH__MAX_BATCH_SIZE"] = str(max(8, (conc + extra_gateway_procs - 1) // extra_gateway_procs * 2))
                await topo.restart_gateway(g2)
                out["gateway_multi_process"] = await guarded("gwmp", client_pool=0)
                out["gateway_multi_process"]["processes"] = extra_gateway_procs
                if "note" in out["gateway_multi_process"]:
                    out["gateway_multi_process"]["note"] += ("; /stats counters (dedup, cache hits, TPOT) are per process and are read "
                                                              "from whichever process answers")
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", type=int, default=3, choices=[1, 3])
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--max-tokens", type=int, default=256)
    ap.add_argument("--rounds", type=int, default=2)
    ap.add_argument("--preset", default="qwen2.5-1.5b-awq")
    ap.add_argument("--levels", default="worker_http,gateway")
    ap.add_argument("--log-dir", default=os.path.join(ROOT, "gpurun_out", "gateway_logs"))
    ap.add_argument("--latency-ms", type=int, default=None, help="config 1: dry-run synthetic latency")
    ap.add_argument("--remote", default=None, choices=[None, "coalesce"], help="gateway-side worker hop (default: the reference's RemoteBackend)")
    ap.add_argument("--gateway-procs", type=int, default=1, help="server processes of the gateway app on its one port (uvicorn workers)")
    ap.add_argument("--extra-gateway-procs", type=int, default=0, help="after the main gateway run, restart the gateway with K processes and measure again")
    ap.add_argument("--client-pool", type=int, default=100, help="connections of the load generator for the main gateway run "
                    "(100: bench_load.py's own default; 0: sized to the concurrency)")
    ap.add_argument("--stage", action="store_true")
    a = ap.parse_args()
    if a.stage:
        print(stage_reference())
        return
    ref = find_reference()
    if ref is None:
        print(json.dumps({"unavailable": "no V-Gate checkout found (VGATE_REFERENCE_DIR, /root/reference, baseline/_ref)"}))
        return
    if a.config == 1:
        out = asyncio.run(run_config1(ref, a.log_dir, a.latency_ms))
    else:
        out = asyncio.run(run_b200(ref, a.log_dir, a.gpus, a.batch, a.max_tokens, a.rounds, a.preset, a.levels.split(","), remote=a.remote, gateway_procs=a.gateway_procs,
                                   extra_gateway_procs=a.extra_gateway_procs, client_pool=a.client_pool))
    out["reference_dir"] = ref
    print(json.dumps(out))


if __name__ == "__main__":
    main()
