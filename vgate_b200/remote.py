"""Gateway-side worker hop for b200 workers: a drop-in for ``vgate/backends/remote_backend.py`` (same
``InferenceBackend`` surface, same ``WorkerRegistry`` health accounting, same retry policy) that closes the two
gaps SURVEY.md section 8(f) ranks first:

1. **Coalescing** (ROADMAP.md:924-926, "not done").  The reference sends one HTTP request per prompt
   (``remote_backend.py:150-153``) although the wire format carries a list (``vgate/worker_api.py:70-82``).
   Here concurrent single-prompt ``generate()`` calls that share sampling parameters are merged into ONE
   multi-prompt ``/internal/generate`` body per worker: callers enqueue and block on a future, a few sender
   threads per gateway drain whatever is queued when they become free (no added delay), pick a worker
   round-robin and demultiplex the response.  One body, one JSON document and one connection instead of 64.
   Two wire forms (``VGATE_B200_COALESCE_MODE``):
   ``nd`` (default) posts to ``/internal/generate_nd`` (``vgate_b200.worker_stream``): the worker answers with one
   newline-delimited result per prompt in COMPLETION order and every caller is released the moment its own line
   arrives -- prompts of any sampling parameters share a body, nobody waits for the slowest.
   ``body`` uses the reference's own ``/internal/generate``: a body completes when its slowest prompt does (so only
   prompts with equal parameters share one, and finished sequences wait for their body: measured slower than the
   reference hop on two B200s).  ``nd`` falls back to ``body`` for a worker that answers 404 (an unmodified worker).

2. **Streaming through the hop** (``remote_backend.py:209-224`` raises; the gateway answers 501,
   ``main.py:429-436``).  ``stream_generate`` opens ``POST /internal/stream`` on a worker (a route
   ``vgate_b200.worker_stream`` adds to a worker process) and yields the same ``{"delta", "num_tokens"}``
   dicts the in-process backends yield; closing the generator closes the connection, the worker sees the
   disconnect and aborts the engine request -- cancellation crosses both connections.

3. **Prefix-affinity routing** (optional, ``VGATE_B200_ROUTE=prefix``; ROADMAP.md:664-675).  Round-robin sends
   prompts with a common start (a system prompt, a conversation so far) to different workers, so each worker's
   engine prefix cache sees a fraction of them.  With this switch a body goes to the healthy worker chosen by the
   hash of the first ``VGATE_B200_ROUTE_CHARS`` (256) characters of its first prompt -- and prompts are grouped into
   bodies by that same key -- falling back to round-robin when that worker is out of rotation.

Selected from outside the reference tree: ``vgate_b200.register`` returns this class instead of
``RemoteBackend`` when ``VGATE_B200_REMOTE=coalesce`` is set for a gateway with worker endpoints.
"""
import json
import os
import threading
import time
from concurrent.futures import Future
from typing import Any, AsyncIterator, Dict, List, Optional

import httpx

from vgate.backends.remote_backend import RemoteInferenceError
from vgate.metrics import WORKER_LATENCY, WORKER_REQUESTS, WORKER_RETRIES
from vgate.worker_registry import NoHealthyWorkersError, WorkerRegistry


class CoalescingRemoteBackend:
    supports_concurrent_calls = True          # vgate/batcher.py:116-118: no inference lock around generate()
    supports_streaming = True                 # main.py:429: absence of False => streaming is offered

    def __init__(self, worker_config, registry: Optional[WorkerRegistry] = None):
        discovering = bool(worker_config.discovery.dns_name)
        if not worker_config.endpoints and not discovering:
            raise ValueError("CoalescingRemoteBackend requires worker.endpoints or worker.discovery.dns_name")
        self.config = worker_config
        self.registry = registry or WorkerRegistry(
            worker_config.endpoints, failure_threshold=worker_config.failure_threshold,
            success_threshold=worker_config.success_threshold, allow_empty=discovering)
        self._headers = {"Authorization": f"Bearer {worker_config.api_key}"} if worker_config.api_key else {}
        self._timeout = httpx.Timeout(worker_config.timeout_seconds, connect=worker_config.connect_timeout_seconds)
        self._client = httpx.Client(timeout=self._timeout, headers=self._headers,
                                    limits=httpx.Limits(max_connections=None, max_keepalive_connections=64))
        self.max_body = int(os.environ.get("VGATE_B200_COALESCE_MAX", "64"))       # prompts per body
        self.route_prefix = os.environ.get("VGATE_B200_ROUTE") == "prefix"
        self.route_chars = int(os.environ.get("VGATE_B200_ROUTE_CHARS", "256"))
        self.mode = os.environ.get("VGATE_B200_COALESCE_MODE", "nd")
        self._no_nd: set = set()                # workers that answered 404 to /internal/generate_nd
        n_senders = int(os.environ.get("VGATE_B200_COALESCE_SENDERS", "16"))
        self._queues: Dict[str, Any] = {}       # sampling-parameter class -> (params, [(prompt, future, t_enqueued)])
        self._cv = threading.Condition()
        self._stop = False
        self.bodies_sent = 0                    # observability: HTTP bodies vs prompts they carried
        self.prompts_sent = 0
        self._senders = [threading.Thread(target=self._sender, name=f"b200-hop-{i}", daemon=True) for i in range(n_senders)]
        for t in self._senders:
            t.start()

    # -- protocol ------------------------------------------------------------------------------------
    def load_model(self, model_config) -> None:
        """No-op: the worker process owns model loading (remote_backend.py:104-105)."""

    def create_sampling_params(self, temperature: float, top_p: float, max_tokens: int) -> Any:
        return {"temperature": temperature, "top_p": top_p, "max_tokens": max_tokens}

    def generate(self, prompts: List[str], sampling_params: Any) -> List[Dict[str, Any]]:
        """Blocking; called from executor threads (vgate/batcher.py:374).  Results in prompt order."""
        key = json.dumps(sampling_params, sort_keys=True) if self.mode != "nd" else "*"   # nd: any mix shares a body
        futs = []
        for p in prompts:
            f: Future = Future()
            k = key if not self.route_prefix else f"{key}|{self._affinity(p)}"    # one body = one worker = one prefix class
            self._enqueue(k, sampling_params, p, f)
            futs.append(f)
        return [f.result() for f in futs]       # raises what the sender set (RemoteInferenceError, NoHealthyWorkersError)

    def _affinity(self, prompt: str) -> int:
        import zlib
        n = max(1, len(self.registry.endpoints()))
        return zlib.crc32(prompt[:self.route_chars].encode("utf-8", "replace")) % n

    def _pick(self, prompts, tried):
        """Round-robin (vgate/worker_registry.py:178-201), or the prefix class's own worker while it is healthy."""
        if self.route_prefix and not tried:
            eps = self.registry.endpoints()
            if eps:
                want = eps[self._affinity(prompts[0]) % len(eps)]
                others = set(eps) - {want}
                try:
                    return self.registry.pick(exclude=others)
                except NoHealthyWorkersError:
                    pass
        return self.registry.pick(exclude=tried)

    # -- coalescing senders ----------------------------------------------------------------------------
    def _enqueue(self, key, sp, prompt, fut):
        with self._cv:
            self._queues.setdefault(key, (sp, []))[1].append((prompt, fut, time.monotonic(), sp))
            self._cv.notify()

    def _take_batch(self):
        """The sampling-parameter class whose oldest prompt has waited longest, up to max_body prompts of it."""
        with self._cv:
            while not self._stop and not self._queues:
                self._cv.wait(timeout=0.2)
            if self._stop or not self._queues:
                return None, None
            key = min(self._queues, key=lambda k: self._queues[k][1][0][2])
            sp, items = self._queues[key]
            batch, rest = items[:self.max_body], items[self.max_body:]
            if rest:
                self._queues[key] = (sp, rest)
            else:
                del self._queues[key]
            return sp, batch

    def _sender(self):
        while not self._stop:
            sp, batch = self._take_batch()
            if not batch:
                continue
            try:
                if self.mode == "nd":
                    self._post_nd(batch)
                else:
                    results = self._post_with_retry([it[0] for it in batch], sp)
                    for it, res in zip(batch, results):
                        it[1].set_result(res)
            except BaseException as exc:        # noqa: BLE001 -- every waiting caller must be released
                for it in batch:
                    if not it[1].done():
                        it[1].set_exception(exc)

    def _post_nd(self, batch):
        """One /internal/generate_nd exchange for `batch` = [(prompt, future, t, sampling_params)]: futures are resolved
        line by line.  Same policy as below: another worker only after a connection failure (nothing was sent)."""
        prompts = [it[0] for it in batch]
        tried: set = set()
        last_connect_error = None
        for attempt in range(max(1, len(self.registry.endpoints()))):
            try:
                endpoint = self._pick(prompts, tried)
            except NoHealthyWorkersError:
                break
            if endpoint in self._no_nd:          # an unmodified worker: the reference's wire form, per parameter class
                by_sp: Dict[str, list] = {}
                for it in batch:
                    by_sp.setdefault(json.dumps(it[3], sort_keys=True), []).append(it)
                for group in by_sp.values():
                    for it, res in zip(group, self._post_with_retry([g[0] for g in group], group[0][3])):
                        it[1].set_result(res)
                return
            if attempt > 0:
                WORKER_RETRIES.labels(worker=endpoint).inc()
            started = time.perf_counter()
            try:
                with self._client.stream("POST", f"{endpoint}/internal/generate_nd",
                                         json={"items": [{"prompt": it[0], "sampling_params": it[3]} for it in batch]}) as resp:
                    if resp.status_code in (404, 405, 501):      # no such route / a backend that is called under a lock
                        self._no_nd.add(endpoint)
                        return self._post_nd(batch)
                    if resp.status_code != 200:
                        body = resp.read()[:200]
                        self.registry.record_failure(endpoint)
                        WORKER_REQUESTS.labels(worker=endpoint, outcome="http_error").inc()
                        raise RemoteInferenceError(f"worker at {endpoint} returned {resp.status_code}: {body!r}")
                    seen = 0
                    for line in resp.iter_lines():
                        if not line:
                            continue
                        msg = json.loads(line)
                        fut = batch[int(msg["i"])][1]
                        if "error" in msg:
                            fut.set_exception(RemoteInferenceError(f"worker at {endpoint} returned 500: Inference failed: {msg['error']}"))
                        else:
                            fut.set_result(msg["result"])
                        seen += 1
                    if seen != len(batch):
                        self.registry.record_failure(endpoint)
                        WORKER_REQUESTS.labels(worker=endpoint, outcome="bad_response").inc()
                        raise RemoteInferenceError(f"worker at {endpoint} returned {seen} results for {len(batch)} prompts")
            except httpx.ConnectError as exc:
                self.registry.record_failure(endpoint)
                WORKER_REQUESTS.labels(worker=endpoint, outcome="connect_error").inc()
                tried.add(endpoint)
                last_connect_error = exc
                continue
            except httpx.RequestError as exc:
                self.registry.record_failure(endpoint)
                WORKER_REQUESTS.labels(worker=endpoint, outcome="request_error").inc()
                raise RemoteInferenceError(f"worker at {endpoint} failed mid-request: {type(exc).__name__}") from exc
            finally:
                WORKER_LATENCY.labels(worker=endpoint).observe(time.perf_counter() - started)
            self.registry.record_success(endpoint)
            WORKER_REQUESTS.labels(worker=endpoint, outcome="success").inc()
            self.bodies_sent += 1
            self.prompts_sent += len(batch)
            return
        detail = f"last error: {type(last_connect_error).__name__}" if last_connect_error else "none reachable"
        raise NoHealthyWorkersError(f"no healthy worker served the request after {len(tried)} attempt(s); {detail}")

    def _post_with_retry(self, prompts, sampling_params):
        """remote_backend.py:114-207: retry on connection failures only; a timeout or an error status is final."""
        tried: set = set()
        last_connect_error = None
        for attempt in range(max(1, len(self.registry.endpoints()))):
            try:
                endpoint = self._pick(prompts, tried)
            except NoHealthyWorkersError:
                break
            if attempt > 0:
                WORKER_RETRIES.labels(worker=endpoint).inc()
            started = time.perf_counter()
            try:
                response = self._client.post(f"{endpoint}/internal/generate",
                                             json={"prompts": prompts, "sampling_params": sampling_params})
            except httpx.ConnectError as exc:
                self.registry.record_failure(endpoint)
                WORKER_REQUESTS.labels(worker=endpoint, outcome="connect_error").inc()
                tried.add(endpoint)
                last_connect_error = exc
                continue
            except httpx.RequestError as exc:
                self.registry.record_failure(endpoint)
                WORKER_REQUESTS.labels(worker=endpoint, outcome="request_error").inc()
                raise RemoteInferenceError(f"worker at {endpoint} failed mid-request: {type(exc).__name__}") from exc
            finally:
                WORKER_LATENCY.labels(worker=endpoint).observe(time.perf_counter() - started)
            if response.status_code != 200:
                self.registry.record_failure(endpoint)
                WORKER_REQUESTS.labels(worker=endpoint, outcome="http_error").inc()
                raise RemoteInferenceError(f"worker at {endpoint} returned {response.status_code}: {response.text[:200]}")
            results = response.json().get("results")
            if not isinstance(results, list) or len(results) != len(prompts):
                self.registry.record_failure(endpoint)
                WORKER_REQUESTS.labels(worker=endpoint, outcome="bad_response").inc()
                raise RemoteInferenceError(f"worker at {endpoint} returned "
                                           f"{len(results) if isinstance(results, list) else 'no'} results for {len(prompts)} prompts")
            self.registry.record_success(endpoint)
            WORKER_REQUESTS.labels(worker=endpoint, outcome="success").inc()
            self.bodies_sent += 1
            self.prompts_sent += len(prompts)
            return results
        detail = f"last error: {type(last_connect_error).__name__}" if last_connect_error else "none reachable"
        raise NoHealthyWorkersError(f"no healthy worker served the request after {len(tried)} attempt(s); {detail}")

    # -- streaming through the hop ---------------------------------------------------------------------------
    async def stream_generate(self, prompt: str, sampling_params: Any) -> AsyncIterator[Dict[str, Any]]:
        """Runs on the gateway's event loop (main.py:342).  One worker, no retry once the stream has started."""
        endpoint = self.registry.pick()
        started = time.perf_counter()
        try:
            async with httpx.AsyncClient(timeout=self._timeout, headers=self._headers) as client:
                async with client.stream("POST", f"{endpoint}/internal/stream",
                                         json={"prompt": prompt, "sampling_params": sampling_params}) as resp:
                    if resp.status_code != 200:
                        body = (await resp.aread())[:200]
                        self.registry.record_failure(endpoint)
                        WORKER_REQUESTS.labels(worker=endpoint, outcome="http_error").inc()
                        raise RemoteInferenceError(f"worker at {endpoint} returned {resp.status_code}: {body!r}")
                    async for line in resp.aiter_lines():
                        if not line:
                            continue
                        piece = json.loads(line)
                        if "error" in piece:
                            raise RemoteInferenceError(f"worker at {endpoint} failed mid-stream: {piece['error']}")
                        yield piece             # {"delta": str, "num_tokens": int}; closing here closes the connection
            self.registry.record_success(endpoint)
            WORKER_REQUESTS.labels(worker=endpoint, outcome="success").inc()
        except httpx.RequestError as exc:
            self.registry.record_failure(endpoint)
            WORKER_REQUESTS.labels(worker=endpoint, outcome="request_error").inc()
            raise RemoteInferenceError(f"worker at {endpoint} failed mid-stream: {type(exc).__name__}") from exc
        finally:
            WORKER_LATENCY.labels(worker=endpoint).observe(time.perf_counter() - started)

    def shutdown(self) -> None:
        self._stop = True
        with self._cv:
            self._cv.notify_all()
        self._client.close()
