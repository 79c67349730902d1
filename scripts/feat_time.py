import sys, time, ctypes as C
sys.path.insert(0, '.')
import torch, numpy as np, importlib
import fdb200 as fd
import bench
model, pk, corpus, stats = bench.build_workload(fd, 1000000, 200, torch.device('cuda', 0), 1)
h = model.handle(0)
P = lambda t: C.c_void_p(t.data_ptr())
sp = C.c_void_p(torch.cuda.current_stream().cuda_stream)
n = pk.n_rows
for rep in range(4):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    nnz = C.c_int64(0)
    h.call("fd_featurize_begin", P(pk.data), P(pk.offs), n, 1, C.byref(nnz), sp)
    torch.cuda.synchronize(); t1 = time.perf_counter()
    rowptr = torch.empty(n + 1, dtype=torch.int64, device='cuda'); idx = torch.empty(nnz.value, dtype=torch.int32, device='cuda'); val = torch.empty(nnz.value, dtype=torch.float64, device='cuda')
    torch.cuda.synchronize(); t2 = time.perf_counter()
    h.call("fd_featurize_finish", P(rowptr), P(idx), P(val), nnz.value, sp)
    torch.cuda.synchronize(); t3 = time.perf_counter()
    print(f"begin {1e3*(t1-t0):.2f} ms  alloc {1e3*(t2-t1):.2f}  finish {1e3*(t3-t2):.2f}  nnz {nnz.value}")
