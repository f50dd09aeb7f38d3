timeout 900 python -m pytest tests/test_engine_gpu.py tests/test_gateway_gpu.py -m gpu -x -q -k "prefix or gateway or batcher or internal or chunked" 2>&1 | tail -15
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
timeout 900 python tools/vllm_b200_baseline.py 64 32 qwen2.5-1.5b-awq-untied peaked > gpurun_out/t6_vllm_peaked.json 2> gpurun_out/t6_vllm_peaked.err; tail -c 1200 gpurun_out/t6_vllm_peaked.json; tail -3 gpurun_out/t6_vllm_peaked.err
