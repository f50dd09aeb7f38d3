import sys; sys.path.insert(0, '/root/repo')
import ctypes as C, numpy as np, torch
from vgate_b200 import _lib
lib=_lib.load(); ptr=lambda t: C.c_void_p(t.data_ptr()) if t is not None else None
st = torch.cuda.Stream(); sp = C.c_void_p(st.cuda_stream)
T,H=64,1536
h=torch.randn((T,H),device='cuda'); g=torch.ones(H,dtype=torch.float16,device='cuda'); xn=torch.empty((T,H),dtype=torch.float16,device='cuda')
for nparts in (0,4,12):
    parts=[torch.randn((max(nparts,1),T,H),device='cuda') for _ in range(8)]
    def call(i):
        _lib.check(lib.b200_op_rmsnorm(ptr(h), ptr(parts[i%8]), nparts, T*H, ptr(g), 1e-6, None, T, H, ptr(xn), sp))
    with torch.cuda.stream(st):
        call(0); st.synchronize()
        gr=torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr, stream=st):
            for i in range(20): call(i)
        ts=[]
        for _ in range(5):
            e0,e1=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
            e0.record(st); gr.replay(); e1.record(st); st.synchronize(); ts.append(e0.elapsed_time(e1)*1e3/20)
    print(f'rmsnorm T={T} H={H} n_parts={nparts}: {np.median(ts):.2f} us per launch (graph replay of 20)')
