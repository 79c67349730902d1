"""C1 host-path latency (1,600 dialogues through fd_transform_host), printed with percentiles; FDB200_LIB selects the library."""
import os, sys, time, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "scripts"))
import numpy as np, torch
import bench_configs as bc
ctx = bc.Ctx()
torch.cuda.set_device(0)
r = bc.run_c1(ctx, rows_big=100000)
print(r["batch_1600_host_path_ms"], r["ms_per_pass"])
