"""CPU oracle for the b200 worker hot path.  TEST INFRASTRUCTURE ONLY.

Nothing under ``oracle/`` is product code.  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import it, and only as the checker or the timed
CPU baseline -- never as a fallback for the CUDA path (``vgate_b200`` raises
when ``libb200gate.so`` is missing).

What it restates
----------------
The arithmetic of V-Gate's generate path is not in the reference tree: the
reference delegates it to vLLM (``vgate/backends/vllm_backend.py:62-86``,
unpinned; ``requirements.txt:14`` is a commented ``vllm>=0.14.0``).  The oracle
therefore restates the published algorithm of that dependency (vLLM 0.22.0 as
installed in this image; file:line citations are into its site-packages tree):

* ``awq.py``      AWQ int4 unpack + dequant
                  (``vllm/model_executor/layers/quantization/awq_triton.py:11-105``,
                  tensor shapes ``awq.py:214-253``)
* ``qwen2.py``    Qwen2 decoder forward with a KV cache
                  (``vllm/model_executor/models/qwen2.py:83-312,389-421,494-565``;
                  same maths as ``transformers`` ``Qwen2ForCausalLM``)
* ``sampler.py``  temperature / top-p / greedy
                  (``vllm/v1/sample/sampler.py:223-298``,
                  ``vllm/v1/sample/ops/topk_topp_sampler.py:328-369,395-416``)

Pinning
-------
The reference holds no numerical golden vectors for this path.
The oracle is pinned against outputs of the real third-party code run in the
build container: vLLM's own ``awq_pack`` (inverse of the unpack), vLLM's
``apply_top_k_top_p_pytorch`` (top-p mask) and ``transformers``
``Qwen2ForCausalLM`` fp32 logits / greedy tokens.  The vectors and the script
that made them are committed under ``tests/golden/``.
"""
