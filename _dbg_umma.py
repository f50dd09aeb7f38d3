import sys; sys.path.insert(0, '/root/repo')
import numpy as np, torch
from oracle.awq import dequantize_awq
from tests.gpu_util import lib, dev, ptr, stream, check, repack, random_awq
from tests.test_umma_gpu import tile
for (K,N,T) in [(8960,1536,256),(8960,1536,700),(8960,1536,64),(1536,17920,64),(1536,2048,1),(8960,1536,16)]:
    rng = np.random.default_rng(1)
    qw, qz, sc = random_awq(rng, K, N)
    wq, s, z = repack(qw, qz, sc)
    ch = tile(wq, s, z, K, N)
    x = rng.standard_normal((T, K)).astype(np.float16)
    ref = x.astype(np.float64) @ dequantize_awq(qw, qz, sc, 128).astype(np.float64)
    xd = dev(x)
    bad = 0
    for rep in range(30):
        y = torch.full((T, N), 7.0, dtype=torch.float32, device="cuda")
        check(lib().b200_op_w4_gemm_umma(ptr(ch), ptr(xd), K, ptr(y), N, None, T, K, N, 128, 0, 1, 0, stream()))
        e = np.abs(y.cpu().numpy() - ref)
        if e.max() > 5e-4:
            bad += 1
            cols = np.unique(np.where(e>5e-4)[1]); rows=np.unique(np.where(e>5e-4)[0])
            print('  bad rep', rep, 'max', e.max(), 'cols', cols.min(), cols.max(), len(cols), 'rows', rows.min(), rows.max(), len(rows))
    print(K,N,T,'bad reps', bad, 'of 30')
