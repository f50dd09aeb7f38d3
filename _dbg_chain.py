import sys, os; sys.path.insert(0, '/root/repo')
import numpy as np, torch, ctypes as C
from tests.gpu_util import lib, dev, ptr, stream, check, repack, random_awq
from tests.test_umma_gpu import tile
from vgate_b200 import _lib
T = 6
K, N = 1536, 17920
rng = np.random.default_rng(1)
qw, qz, sc = random_awq(rng, K, N)
wq, s, z = repack(qw, qz, sc)
ch = tile(wq, s, z, K, N)
mode = os.environ.get("MODE", "plain")
xn = torch.from_numpy(rng.standard_normal((512, K)).astype(np.float16)).cuda()
if mode == "nan":
    xn[6:16] = float('nan')
h = torch.from_numpy(rng.standard_normal((T, K)).astype(np.float32)).cuda()
g = torch.ones(K, dtype=torch.float16, device="cuda")
R = 60
outs = [torch.zeros((T, N), dtype=torch.float32, device="cuda") for _ in range(R)]
torch.cuda.synchronize()
st = stream()
flush = torch.empty(256*1024*1024, dtype=torch.uint8, device="cuda")
for r in range(R):
    if os.environ.get("FLUSH"): flush.fill_(r)
    if mode in ("rms", "nan"):
        _lib.check(lib().b200_op_rmsnorm(ptr(h), None, 0, 0, ptr(g), 1e-6, None, T, K, ptr(xn), st))
    _lib.check(lib().b200_op_w4_gemm_umma(ptr(ch), ptr(xn), K, ptr(outs[r]), N, None, T, K, N, 128, 0, 1, 0, st))
torch.cuda.synchronize()
bad = [r for r in range(1, R) if not torch.equal(outs[r], outs[0])]
msg = ''
if bad:
    d = (outs[bad[0]] - outs[0]).abs(); idx = torch.nonzero(d > 0)
    msg = f' first bad {bad[0]} max diff {d.max().item():.4g} n_diff {idx.shape[0]} rows {sorted(set(idx[:,0].tolist()))} cols {idx[:,1].min().item()}..{idx[:,1].max().item()} ntiles {len(set((idx[:,1]//128).tolist()))}'
print(f'mode={mode} XROWS={os.environ.get("VGATE_B200_DEBUG_XROWS")}: {len(bad)} of {R-1} differ.' + msg, 'nan in out0:', bool(torch.isnan(outs[0]).any()))
