"""``B200Backend`` -- V-Gate's ``model.engine_type: b200`` inference backend.

Implements the adapter protocol of ``vgate/backends/base.py:65-91`` with the behaviour of
``vgate/backends/vllm_backend.py:35-149`` (the backend it replaces): every prompt is an
independent engine request, concurrent ``generate()`` calls are continuous-batched inside the
engine (``supports_concurrent_calls = True``, vllm_backend.py:55), results are
``{"text", "token_ids", "num_tokens", "metrics": {"ttft", "gen_time"}}`` (vllm_backend.py:122-132),
and ``stream_generate`` yields ``{"delta", "num_tokens"}`` with a cumulative count
(vllm_backend.py:134-146).  All arithmetic happens in libb200gate.so (sm_100a CUDA); this file
is tokenisation and plumbing.  Knobs that ``ModelConfig`` (vgate/config.py:105-113) has no field
for are read from ``VGATE_B200_*`` environment variables, the pattern the dry-run backend uses
(vgate/backends/base.py:26,45).
"""
import asyncio
import os
import secrets
from concurrent.futures import ThreadPoolExecutor
from typing import Any, AsyncIterator, Dict, List

from .tokenizer import load_tokenizer, IncrementalDetokenizer, qwen_chat_prompt
from .weights import PRESETS, synthetic_checkpoint, load_hf_checkpoint


def _env_int(name: str, default: int) -> int:
    return int(os.environ.get(name, default))


def _sp_get(sp: Any, key: str, default):
    if isinstance(sp, dict):
        return sp.get(key, default)
    return getattr(sp, key, default)


class B200Backend:
    """Inference backend running the model on one B200 through libb200gate."""

    # The engine continuous-batches concurrent requests; serialising calls would leave it
    # decoding one sequence at a time (vgate/batcher.py:116-118, vgate/worker_api.py:60,110).
    supports_concurrent_calls = True

    def __init__(self):
        self.engine = None
        self.tokenizer = None
        # "" (default): the flat "User: ...\nAssistant:" prompt goes to the model as the reference sends it to vLLM;
        # "qwen": rewritten into Qwen2.5-Instruct's ChatML template first (tokenizer.qwen_chat_prompt)
        self._chat_template = os.environ.get("VGATE_B200_CHAT_TEMPLATE", "")
        self._loop = None

    # ------------------------------------------------------------------------------------
    def load_model(self, model_config: Any) -> None:
        from .engine import Engine      # imports libb200gate; raises if it is not built

        model_id = getattr(model_config, "model_id", "") or ""
        max_model_len = int(getattr(model_config, "max_model_len", 2048))
        max_seqs = _env_int("VGATE_B200_MAX_SEQS", 256)
        # load_model runs on the serving loop (vgate/engine.py:93 <- main.py:79).  Each in-flight
        # generate() parks one default-executor thread (vgate/worker_api.py:112), so the default
        # min(32, cpu+4) pool would cap the decode batch; enlarge it for this worker process.
        try:
            self._loop = asyncio.get_event_loop()
            n_threads = _env_int("VGATE_B200_EXECUTOR_THREADS", max_seqs + 8)
            if n_threads > 0:
                self._loop.set_default_executor(ThreadPoolExecutor(n_threads, thread_name_prefix="b200-call"))
        except RuntimeError:
            self._loop = None

        if os.path.isdir(model_id) and os.path.exists(os.path.join(model_id, "config.json")):
            print(f"Loading {model_id} (AWQ safetensors) into the b200 engine...")
            dims, ckpt = load_hf_checkpoint(model_id)
        else:
            preset = os.environ.get("VGATE_B200_SYNTHETIC", "qwen2.5-1.5b-awq")
            seed = _env_int("VGATE_B200_SEED", 0)
            print(f"No local checkpoint for {model_id!r}: random-init {preset} (seed {seed}) on the b200 engine...")
            dims = PRESETS[preset]
            ckpt = synthetic_checkpoint(dims, seed=seed)
        self.tokenizer = load_tokenizer(model_id)
        tp = _env_int("VGATE_B200_TP", 1)
        if tp > 1:                       # the larger-model config: one replica over `tp` consecutive GPUs (NCCL all-reduce inside)
            from .engine import TPEngine
            dev = _env_int("VGATE_B200_DEVICE", 0)
            self.engine = TPEngine(
                dims, ckpt, devices=tuple(range(dev, dev + tp)), max_seqs=max_seqs, max_model_len=max_model_len,
                max_batch_tokens=_env_int("VGATE_B200_MAX_BATCH_TOKENS", 8192), kv_bytes=_env_int("VGATE_B200_KV_BYTES", 0),
                eos_token_id=getattr(self.tokenizer, "eos_token_id", -1),
                use_cuda_graphs=bool(_env_int("VGATE_B200_CUDA_GRAPHS", 1)))
            return
        self.engine = Engine(
            dims, device=_env_int("VGATE_B200_DEVICE", 0), max_seqs=max_seqs, max_model_len=max_model_len,
            max_batch_tokens=_env_int("VGATE_B200_MAX_BATCH_TOKENS", 8192),
            kv_bytes=_env_int("VGATE_B200_KV_BYTES", 0),
            eos_token_id=getattr(self.tokenizer, "eos_token_id", -1),
            use_cuda_graphs=bool(_env_int("VGATE_B200_CUDA_GRAPHS", 1)),
            gemm_path=_env_int("VGATE_B200_GEMM_PATH", 0)).load_checkpoint(ckpt)

    def create_sampling_params(self, temperature: float, top_p: float, max_tokens: int) -> Any:
        # plain dict: it must survive being rebuilt from the wire dict on the worker
        # (vgate/worker_api.py:99-103)
        return {"temperature": temperature, "top_p": top_p, "max_tokens": max_tokens}

    # ------------------------------------------------------------------------------------
    def _submit(self, prompt: str, sampling_params: Any):
        if self.engine is None:
            raise RuntimeError("B200Backend.load_model() has not been called")
        if self._chat_template == "qwen":
            prompt = qwen_chat_prompt(prompt)
        ids = self.tokenizer.encode(prompt)
        max_tokens = int(_sp_get(sampling_params, "max_tokens", 256))
        seed = _sp_get(sampling_params, "seed", None)
        rid = self.engine.submit(
            ids, temperature=float(_sp_get(sampling_params, "temperature", 0.7)),
            top_p=float(_sp_get(sampling_params, "top_p", 0.9)), max_tokens=max_tokens,
            seed=secrets.randbits(32) if seed is None else int(seed))
        return rid, max_tokens

    def generate(self, prompts: List[str], sampling_params: Any) -> List[Dict[str, Any]]:
        """Blocking, batch-shaped entry point; called from executor threads, re-entrant.  The GIL
        is released while waiting inside the C library."""
        pending = []                   # request ids this call still owns (submitted, not yet waited on)
        results = []
        try:
            submitted = []
            for p in prompts:
                submitted.append(self._submit(p, sampling_params))
                pending.append(submitted[-1][0])
            for rid, max_tokens in submitted:
                ids, ttft, gen_time = self.engine.wait(rid, max_tokens)     # releases the request
                pending.remove(rid)
                token_ids = [int(t) for t in ids]
                results.append({
                    "text": self.tokenizer.decode(token_ids),
                    "token_ids": token_ids,
                    "num_tokens": len(token_ids),
                    "metrics": {"ttft": float(ttft), "gen_time": float(gen_time)},
                })
        finally:
            # a failed submit or wait: the rest of this call's requests must neither keep decoding nor stay
            # in the engine's table (abort releases them)
            for rid in pending:
                self.engine.abort(rid)
        return results

    async def stream_generate(self, prompt: str, sampling_params: Any) -> AsyncIterator[Dict[str, Any]]:
        """Runs on the caller's event loop (main.py:342) and never blocks it: the bounded
        blocking poll runs on an executor thread."""
        loop = asyncio.get_running_loop()
        rid, _ = self._submit(prompt, sampling_params)
        detok = IncrementalDetokenizer(self.tokenizer)
        n_tokens, finished = 0, False
        try:
            while not finished:
                ids, finished = await loop.run_in_executor(None, self.engine.poll, rid, 64, 50)
                if len(ids) == 0:
                    continue
                n_tokens += len(ids)
                delta = detok.push(ids)
                if delta:
                    yield {"delta": delta, "num_tokens": n_tokens}
        finally:
            if not finished:          # generator closed mid-way (client disconnect, main.py:372-376)
                self.engine.abort(rid)

    def healthy(self):
        """(ok, reason): False once the engine has hit a CUDA error and failed its requests (it accepts no new
        ones); read by the worker's /health (vgate_b200/worker_stream.py) so that probes see a dead engine."""
        if self.engine is None:
            return False, "engine not loaded"
        return self.engine.status()

    def shutdown(self) -> None:
        if self.engine is not None:
            self.engine.close()
        self.engine = None
