"""Phase-boundary cost inside a persistent kernel (b200_debug_gridsync_bench) for several payloads."""
import ctypes as C
import sys

import torch

sys.path.insert(0, ".")
from vgate_b200 import _lib

lib = _lib.load()
ptr = lambda t: C.c_void_p(t.data_ptr())
st = torch.cuda.Stream()
sp = C.c_void_p(st.cuda_stream)
n_phases = 2000
flags = torch.zeros(n_phases, dtype=torch.int32, device="cuda")
sink = torch.zeros(1, dtype=torch.float32, device="cuda")
for ctas in (148, 64):
    for payload in (0, 1024, 8192, 65536):
        buf = torch.zeros(max(1, 2 * ctas * payload), dtype=torch.float32, device="cuda")
        ts = []
        for it in range(4):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            with torch.cuda.stream(st):
                e0.record(st)
                _lib.check(lib.b200_debug_gridsync_bench(ptr(flags), n_phases, ptr(buf), payload, ctas, ptr(sink), sp))
                e1.record(st)
            st.synchronize()
            ts.append(e0.elapsed_time(e1) * 1e3 / n_phases)
        print(f"ctas {ctas:4d} payload {payload * 4 / 1024:7.1f} KB/CTA: {min(ts[1:]):.3f} us / phase", flush=True)
