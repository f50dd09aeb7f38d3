"""The kernel-to-beat: the reference's own GPU path -- vLLM, the library
`vgate/backends/vllm_backend.py:61-132` delegates to -- on the same B200, same workload as bench.py:
the same seeded random-init AWQ checkpoint written as an HF model directory, 64 prompts x 128 random
token ids, greedy, 256 new tokens, EOS ignored.  Prints one JSON line.  Also checks the engine's
greedy tokens against vLLM's on the first prompts (different kernels: near-tie flips are expected).

Not part of the test suite or bench contract: vLLM is third-party, takes minutes to start, and may
not run offline on every box.  usage: python tools/vllm_b200_baseline.py [batch] [new_tokens] [preset] [profile]"""
import json
import os
import sys
import tempfile
import time

import numpy as np

sys.path.insert(0, ".")
os.environ.setdefault("HF_HUB_OFFLINE", "1")
os.environ.setdefault("TRANSFORMERS_OFFLINE", "1")
os.environ.setdefault("VLLM_NO_USAGE_STATS", "1")
os.environ.setdefault("VLLM_LOGGING_LEVEL", "WARNING")

from vgate_b200.weights import PRESETS, synthetic_checkpoint  # noqa: E402


def write_model_dir(dims, ckpt, path):
    from safetensors.numpy import save_file
    os.makedirs(path, exist_ok=True)
    cfg = {
        "architectures": ["Qwen2ForCausalLM"], "model_type": "qwen2", "hidden_size": dims.hidden,
        "intermediate_size": dims.intermediate, "num_hidden_layers": dims.n_layers,
        "num_attention_heads": dims.n_heads, "num_key_value_heads": dims.n_kv_heads, "vocab_size": dims.vocab,
        "rms_norm_eps": dims.rms_eps, "rope_theta": dims.rope_theta, "max_position_embeddings": 32768,
        "tie_word_embeddings": bool(dims.tie_embeddings), "torch_dtype": "float16", "hidden_act": "silu",
        "bos_token_id": 0, "eos_token_id": 1, "use_sliding_window": False, "attention_dropout": 0.0,
        "quantization_config": {"quant_method": "awq", "bits": 4, "group_size": dims.group_size, "zero_point": True,
                                "version": "gemm"},
    }
    with open(os.path.join(path, "config.json"), "w") as f:
        json.dump(cfg, f)
    with open(os.path.join(path, "generation_config.json"), "w") as f:
        json.dump({"bos_token_id": 0, "eos_token_id": 1}, f)
    save_file({k: np.ascontiguousarray(v) for k, v in ckpt.items()}, os.path.join(path, "model.safetensors"))


def main():
    batch = int(sys.argv[1]) if len(sys.argv) > 1 else 64
    new_tokens = int(sys.argv[2]) if len(sys.argv) > 2 else 256
    preset = sys.argv[3] if len(sys.argv) > 3 else "qwen2.5-1.5b-awq"
    profile = sys.argv[4] if len(sys.argv) > 4 else "flat"
    dims = PRESETS[preset]
    ckpt = synthetic_checkpoint(dims, seed=0, profile=profile)
    rng = np.random.default_rng(1)
    prompts = rng.integers(0, dims.vocab, size=(batch, 128), dtype=np.int32)
    out = {"impl": "vllm", "preset": preset, "profile": profile, "batch": batch, "prompt_len": 128, "new_tokens": new_tokens}
    with tempfile.TemporaryDirectory() as d:
        write_model_dir(dims, ckpt, d)
        import vllm
        from vllm import LLM, SamplingParams
        out["vllm_version"] = vllm.__version__
        llm = LLM(model=d, quantization="awq", dtype="float16", skip_tokenizer_init=True, max_model_len=512,
                  gpu_memory_utilization=0.5, max_num_seqs=max(batch, 64), enforce_eager=False, seed=0)
        sp = SamplingParams(temperature=0.0, top_p=1.0, max_tokens=new_tokens, ignore_eos=True, detokenize=False)
        reqs = [{"prompt_token_ids": p.tolist()} for p in prompts]
        llm.generate(reqs[:8], SamplingParams(temperature=0.0, max_tokens=8, ignore_eos=True, detokenize=False), use_tqdm=False)
        times = []
        for _ in range(3):
            t0 = time.perf_counter()
            res = llm.generate(reqs, sp, use_tqdm=False)
            times.append(time.perf_counter() - t0)
        toks = np.array([r.outputs[0].token_ids for r in res], dtype=np.int32)
        best = min(times)
        out.update({"seconds_per_pass": times, "tok_per_s": batch * new_tokens / best,
                    "ms_per_pass": best * 1e3})
        del llm
    # agreement with this repo's engine on the same inputs (informational)
    try:
        import gc
        import torch
        gc.collect()
        torch.cuda.empty_cache()
        from vgate_b200.engine import Engine
        n = min(batch, 64)
        eng = Engine(dims, max_seqs=64, max_model_len=512, max_batch_tokens=4096).load_checkpoint(ckpt)
        ours, _, _ = eng.generate_batch(prompts[:n], min(new_tokens, 32))
        eng.close()
        k = ours.shape[1]
        same = (ours == toks[:n, :k]).all(axis=1)
        first_div = [int(np.argmax(ours[i] != toks[i, :k])) if not same[i] else -1 for i in range(n)]
        out["greedy_agreement_with_b200_engine"] = {"prompts": n, "tokens": k, "identical": int(same.sum()),
                                                    "first_divergence_step": first_div}
    except Exception as e:  # noqa: BLE001
        out["greedy_agreement_with_b200_engine"] = {"error": repr(e)}
    print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
