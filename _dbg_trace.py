import sys; sys.path.insert(0, '/root/repo')
import ctypes as C, numpy as np, torch
from vgate_b200 import _lib
lib=_lib.load(); ptr=lambda t: C.c_void_p(t.data_ptr()) if t is not None else None
def run(K,N,T,splits=1,G=128):
    rng=np.random.default_rng(0)
    wq=torch.from_numpy(rng.integers(0,2**31,size=(N,K//8),dtype=np.int32)).cuda()
    sc=torch.full((N,K//G),0.003,dtype=torch.float16,device='cuda'); zp=torch.full((N,K//G),1032.0,dtype=torch.float16,device='cuda')
    ch=torch.empty(lib.b200_op_w4_tile_bytes(K,N,G),dtype=torch.uint8,device='cuda')
    _lib.check(lib.b200_op_w4_tile(ptr(wq),ptr(sc),ptr(zp),K,N,G,0,ptr(ch),None))
    x=torch.randn((T,K),device='cuda').half(); splits=lib.b200_op_w4_effective_splits(K,G,splits)
    y=torch.empty((splits,T,N),dtype=torch.float32,device='cuda')
    nct=N//128*splits
    tr=torch.zeros((nct,8),dtype=torch.int64,device='cuda')
    for rep in range(3):
        lib.b200_debug_umma_trace(ptr(tr))
        _lib.check(lib.b200_op_w4_gemm_umma(ptr(ch),ptr(x),K,ptr(y),N,None,T,K,N,G,0,splits,T*N,None))
        torch.cuda.synchronize()
    lib.b200_debug_umma_trace(None)
    t=tr.cpu().numpy().astype(np.float64)
    d=t[:,1:7]-t[:,0:1]
    v=t[:,7].astype(np.int64); nk=(K//64/splits); print('   mma thread per k-tile: wait %.0f issue+commit %.0f' % (np.median(v & 0xFFFFFFFF)/nk, np.median(v>>32)/nk))
    print(f'K={K} N={N} T={T} splits={splits} ctas={nct}: median cycles since entry: setup {np.median(d[:,0]):.0f}, first tiles {np.median(d[:,1]):.0f}, last mma issued {np.median(d[:,2]):.0f}, acc ready {np.median(d[:,3]):.0f}, epilogue done {np.median(d[:,4]):.0f}, exit {np.median(d[:,5]):.0f}   (max exit {d[:,5].max():.0f})')
run(128,17920,64); run(1536,17920,64); run(1536,2048,64); run(8960,1536,64,12); run(1536,17920,16)
