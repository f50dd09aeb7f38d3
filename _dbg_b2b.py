import sys, os; sys.path.insert(0, '/root/repo')
import numpy as np, torch, ctypes as C
from oracle.awq import dequantize_awq
from tests.gpu_util import lib, dev, ptr, stream, check, repack, random_awq
from tests.test_umma_gpu import tile
from vgate_b200 import _lib
T = int(os.environ.get("T", "6"))
shapes = [(1536, 2048, 4), (1536, 1536, 4), (1536, 17920, 1), (8960, 1536, 12)]
ws = []
rng = np.random.default_rng(1)
for (K, N, sp) in shapes:
    qw, qz, sc = random_awq(rng, K, N)
    wq, s, z = repack(qw, qz, sc)
    ch = tile(wq, s, z, K, N)
    x = torch.from_numpy(rng.standard_normal((16, K)).astype(np.float16)).cuda()   # rows >= T are "stale"
    sp = lib().b200_op_w4_effective_splits(K, 128, sp)
    ws.append((K, N, sp, ch, x))
R = 60
outs = [[torch.zeros((sp, T, N), dtype=torch.float32, device="cuda") for (K, N, sp, ch, x) in ws] for _ in range(R)]
torch.cuda.synchronize()
st = stream()
for r in range(R):
    for i, (K, N, sp, ch, x) in enumerate(ws):
        _lib.check(lib().b200_op_w4_gemm_umma(ptr(ch), ptr(x), K, ptr(outs[r][i]), N, None, T, K, N, 128, 0, sp, T * N, st))
torch.cuda.synchronize()
for i, (K, N, sp, ch, x) in enumerate(ws):
    ref = outs[0][i]
    bad = [r for r in range(1, R) if not torch.equal(outs[r][i], ref)]
    msg = ''
    if bad:
        d = (outs[bad[0]][i] - ref).abs()
        idx = torch.nonzero(d > 0)
        msg = f' first bad rep {bad[0]} max diff {d.max().item():.4g} n_diff {idx.shape[0]} splits {sorted(set(idx[:,0].tolist()))} rows {sorted(set(idx[:,1].tolist()))} cols {idx[:,2].min().item()}..{idx[:,2].max().item()}'
    print(f'K={K} N={N} splits={sp} T={T}: {len(bad)} of {R-1} back-to-back launches differ from launch 0.' + msg)
