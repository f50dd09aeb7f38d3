"""The N>1 path of bench.py on CPU: two ranks over gloo, one independent replica each (no data-path
collective), whole-job throughput = all ranks' tokens / the slowest rank's time."""
import os
import sys

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import bench
    seconds = 2.0 if rank == 0 else 3.0                 # rank 1 is the slow replica
    value, worst = bench.aggregate_throughput(1000 * (rank + 1), seconds)
    mx = bench.reduce_over_ranks([float(rank), 5.0 - rank], "max")
    dist.barrier()
    q.put((rank, value, worst, mx))
    dist.destroy_process_group()


def test_two_ranks_aggregate_like_the_bench_contract():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    [p.start() for p in procs]
    res = sorted(q.get(timeout=120) for _ in range(2))
    [p.join(timeout=60) for p in procs]
    assert all(p.exitcode == 0 for p in procs)
    for rank, value, worst, mx in res:
        assert worst == 3.0                              # MAX over ranks
        assert abs(value - 3000 / 3.0) < 1e-9            # all tokens / slowest time
        assert mx == [1.0, 5.0]


def test_single_process_is_identity():
    sys.path.insert(0, ROOT)
    import bench
    assert bench.reduce_over_ranks([1.5, 2.5], "max") == [1.5, 2.5]
    v, w = bench.aggregate_throughput(100, 4.0)
    assert v == 25.0 and w == 4.0


def _tp_worker(rank, world, port, q):
    """One tensor-parallel rank on CPU: its shard of a dense decoder layer (vgate_b200.weights.shard_checkpoint), the
    oracle's layer arithmetic on the shard, and a gloo all-reduce where the engine calls ncclAllReduce (after o_proj
    and after down_proj)."""
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import numpy as np
    from oracle.qwen2 import OracleQwen2, Qwen2Dims
    from vgate_b200.weights import ModelDims, synthetic_checkpoint, shard_checkpoint, shard_dims
    d = ModelDims(hidden=256, intermediate=512, n_layers=2, n_heads=4, n_kv_heads=2, head_dim=128, vocab=512,
                  tie_embeddings=False, quant="none", dtype="bf16")
    ck = synthetic_checkpoint(d, seed=9)
    ids = np.random.default_rng(3).integers(0, d.vocab, (2, 11))
    full = OracleQwen2(Qwen2Dims(**d.to_dict()), ck, mode="fp32")
    want = full.forward(ids).numpy()
    # the rank's model: shard dimensions, shard tensors, and the residual branches summed over ranks
    rd = shard_dims(d, world)
    m = OracleQwen2(Qwen2Dims(**rd.to_dict()), shard_checkpoint(d, ck, world, rank), mode="fp32")
    real_mm = m._mm

    def mm(x, w):                                    # o_proj: [nh/tp * 128, H]; down_proj: [I/tp, H] -> partial sums
        y = real_mm(x, w)
        if w.shape == (rd.n_heads * rd.head_dim, rd.hidden) or w.shape == (rd.intermediate, rd.hidden):
            dist.all_reduce(y)
        return y
    m._mm = mm
    got = m.forward(ids).numpy()
    q.put((rank, float(np.abs(got - want).max())))
    dist.barrier()
    dist.destroy_process_group()


def test_tensor_parallel_partition_sums_to_the_whole_model():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 31500 + os.getpid() % 2000
    procs = [ctx.Process(target=_tp_worker, args=(r, 2, port, q)) for r in range(2)]
    [p.start() for p in procs]
    res = sorted(q.get(timeout=180) for _ in range(2))
    [p.join(timeout=60) for p in procs]
    assert all(p.exitcode == 0 for p in procs)
    for rank, err in res:
        assert err < 1e-4, (rank, err)                # every rank ends with the whole model's logits
