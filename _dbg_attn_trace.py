import sys; sys.path.insert(0, '/root/repo')
import ctypes as C, numpy as np, torch
from vgate_b200 import _lib
lib=_lib.load(); ptr=lambda t: C.c_void_p(t.data_ptr()) if t is not None else None
def run(T, ctx, n_splits, n_heads=12, n_kv=2):
    nb_seq=(ctx+15)//16+1; n_blocks=T*nb_seq
    kc=torch.randn((n_blocks,n_kv,16,128),device='cuda').half(); vc=torch.randn((n_blocks,n_kv,16,128),device='cuda').half()
    bt=torch.arange(T*nb_seq,dtype=torch.int32,device='cuda').reshape(T,nb_seq).contiguous()
    btrow=torch.arange(T,dtype=torch.int32,device='cuda'); ctxd=torch.full((T,),ctx,dtype=torch.int32,device='cuda')
    q=torch.randn((T,n_heads*128),device='cuda').half(); out=torch.empty_like(q)
    scratch=torch.empty(max(4,lib.b200_op_attn_scratch_bytes(T,n_heads,n_kv,n_splits)//4),dtype=torch.float32,device='cuda')
    counters=torch.zeros(T*n_kv,dtype=torch.int32,device='cuda')
    ncta=n_kv*T*n_splits
    tr=torch.zeros((ncta,8),dtype=torch.int64,device='cuda')
    for rep in range(3):
        tr.zero_(); lib.b200_debug_attn_trace(ptr(tr))
        _lib.check(lib.b200_op_paged_attn(ptr(q),ptr(kc),ptr(vc),ptr(bt),nb_seq,ptr(btrow),ptr(ctxd),T,n_heads,n_kv,n_splits,ptr(scratch),ptr(counters),ptr(out),None))
        torch.cuda.synchronize()
    lib.b200_debug_attn_trace(None)
    t=tr.cpu().numpy().astype(np.float64); t=t[t[:,6]>0]
    d=t[:,1:7]-t[:,0:1]
    names=['copies issued','wait start','tile landed','scores done','softmax done','loop end']
    print(f'T={T} ctx={ctx} splits={n_splits} active ctas={len(t)}: '+', '.join(f'{n} {np.median(d[:,i]):.0f}' for i,n in enumerate(names)))
run(64,256,1); run(64,384,1); run(8,256,1)
