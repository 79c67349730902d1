import os, sys, time, ctypes as C
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "scripts"))
import numpy as np, torch
import bench_configs as bc
import fdb200 as fd
ctx = bc.Ctx(); torch.cuda.set_device(0)
def c1():
    r = bc.run_c1(ctx, rows_big=50000); return r["batch_1600_host_path_ms"]
print("fresh", c1())
# big pinned allocation + copies, as bench.py's e2e part does
pk, _ = ctx.corpus.sample(1_000_000, tokens=200, device=ctx.dev, seed=1)
hdata = torch.empty(pk.data.numel(), dtype=torch.uint8, pin_memory=True); hdata.copy_(pk.data)
print("after 1 GB pinned alloc", c1())
del hdata
print("after freeing it", c1())
import threading
from oracle import c_oracle
sw = fd.StopWordsRemover(inputCol="words", outputCol="f").getOrDefault("stopWords")
co = c_oracle.COracle(sw, False, 1 << 18, False, None, None)
d = pk.data.cpu().numpy(); o = pk.offs.cpu().numpy()
co.featurize(d, o[:200001], True)
print("after the C oracle ran", c1())
import pynvml as nv
nv.nvmlInit(); h = nv.nvmlDeviceGetHandleByIndex(0); nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)
print("after NVML init", c1())
