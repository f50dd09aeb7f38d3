"""Serving-like traffic through B200Backend.generate (the call V-Gate makes, one prompt per call from an executor
thread: vgate/worker_api.py:112, vgate/batcher.py:422): Poisson arrivals, mixed generation lengths -- the regime
in which finishes and arrivals change the decode batch all the time and prompts must be prefilled while other
sequences decode.  Reports throughput and the reference's TPOT definition (gen_time / num_tokens,
vgate/batcher.py:431) at p50 / p99, beside the all-at-once number bench.py times.

usage: python tools/bench_traffic.py [--rate REQ_PER_S] [--requests N] [--chunks 2048,512,8192]
"""
import argparse
import json
import os
import sys
import threading
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np


def run(eng_kw, preset, rate, n_req, seed, max_inflight, long_frac=0.0, long_len=1536):
    from vgate_b200.backend import B200Backend
    from vgate_b200.engine import Engine
    from vgate_b200.tokenizer import ByteTokenizer
    from vgate_b200.weights import PRESETS, synthetic_checkpoint
    dims = PRESETS[preset]
    eng = Engine(dims, **eng_kw).load_checkpoint(synthetic_checkpoint(dims, seed=0, profile="peaked"))
    be = B200Backend()
    be.engine, be.tokenizer = eng, ByteTokenizer()
    rng = np.random.default_rng(seed)
    lens = np.where(rng.random(n_req) < long_frac, long_len, 128)       # a share of long prompts: their prefill is what chunking cuts up
    prompts = [bytes(rng.integers(32, 127, size=int(n), dtype=np.uint8)).decode("ascii") for n in lens]
    news = rng.integers(64, 257, size=n_req)
    gaps = rng.exponential(1.0 / rate, size=n_req)
    # warm-up: every decode batch size gets its kernels loaded / graph captured lazily; touch a few
    be.generate(prompts[:8], be.create_sampling_params(0.0, 1.0, 8))
    results, lock, sem = [], threading.Lock(), threading.Semaphore(max_inflight)

    def one(i):
        t0 = time.perf_counter()
        r = be.generate([prompts[i]], be.create_sampling_params(0.0, 1.0, int(news[i])))[0]
        dt = time.perf_counter() - t0
        with lock:
            results.append((r["num_tokens"], r["metrics"]["ttft"], r["metrics"]["gen_time"], dt))
        sem.release()

    threads = []
    st0 = eng.stats()
    t_start = time.perf_counter()
    for i in range(n_req):
        time.sleep(gaps[i])
        sem.acquire()
        th = threading.Thread(target=one, args=(i,))
        th.start()
        threads.append(th)
    for th in threads:
        th.join()
    wall = time.perf_counter() - t_start
    st1 = eng.stats()
    eng.close()
    toks = np.array([r[0] for r in results])
    ttft = np.array([r[1] for r in results])
    tpot = np.array([r[2] / r[0] for r in results])
    return {"requests": n_req, "offered_req_per_s": rate, "tok_per_s": float(toks.sum() / wall), "req_per_s": n_req / wall,
            "tpot_ms_p50": float(np.percentile(tpot, 50) * 1e3), "tpot_ms_p99": float(np.percentile(tpot, 99) * 1e3),
            "ttft_ms_p50": float(np.percentile(ttft, 50) * 1e3), "ttft_ms_p99": float(np.percentile(ttft, 99) * 1e3),
            "prefill_passes": st1["steps_prefill"] - st0["steps_prefill"], "decode_steps": st1["steps_decode"] - st0["steps_decode"],
            "mean_decode_batch": (st1["decode_batch_sum"] - st0["decode_batch_sum"]) / max(1, st1["steps_decode"] - st0["steps_decode"])}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--preset", default="qwen2.5-1.5b-awq")
    ap.add_argument("--rate", type=float, default=150.0, help="offered requests per second (Poisson)")
    ap.add_argument("--requests", type=int, default=600)
    ap.add_argument("--max-seqs", type=int, default=64)
    ap.add_argument("--chunks", default="2048,512,8192", help="VGATE_B200_PREFILL_CHUNK values to compare")
    ap.add_argument("--long-frac", type=float, default=0.0, help="share of requests with a --long-len-token prompt")
    ap.add_argument("--long-len", type=int, default=1536)
    a = ap.parse_args()
    out = {"workload": f"{a.preset}: Poisson arrivals at {a.rate:g} req/s, {a.requests} requests, 128-token prompts, greedy, "
                       f"max_tokens uniform in 64..256, at most {a.max_seqs} sequences in the engine"
                       + (f"; {a.long_frac:.0%} of the prompts have {a.long_len} tokens" if a.long_frac > 0 else ""), "runs": {}}
    for chunk in a.chunks.split(","):
        os.environ["VGATE_B200_PREFILL_CHUNK"] = chunk
        kw = dict(max_seqs=a.max_seqs, max_model_len=max(128, a.long_len if a.long_frac > 0 else 0) + 256 + 16, max_batch_tokens=8192)
        out["runs"][f"prefill_chunk_{chunk}"] = run(kw, a.preset, a.rate, a.requests, 7, a.max_seqs * 2, a.long_frac, a.long_len)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
