timeout 600 python tools/bench_gateway.py --gpus 2 --rounds 4 --levels gateway --remote coalesce --client-pool 0 > gpurun_out/t18_gateway_n2_coalesce_nd.json 2> gpurun_out/t18.err
python -c "
import json; d=json.loads(open('gpurun_out/t18_gateway_n2_coalesce_nd.json').read().strip().splitlines()[-1]); print(json.dumps(d.get('gateway'))[:700])"
tail -3 gpurun_out/t18.err
