"""Fuzz of fd_transform_host's in-place path (pinned buffers) against the device path (fd_transform) on random batches:
random byte strings, row counts, starting rows inside a larger buffer, classifier kinds.  Usage: fuzz_direct_host.py [iterations]"""
import ctypes as C
import importlib
import os
import sys
import tarfile
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import fdb200 as fd  # noqa: E402

P = importlib.import_module("fraud-detection-spark-kafka-llm_b200._lib").ptr
iters = int(sys.argv[1]) if len(sys.argv) > 1 else 200
d = tempfile.mkdtemp()
with tarfile.open(os.path.join(ROOT, "tests", "golden", "shipped_model.tar.gz")) as tf:
    tf.extractall(d, filter="data")
model = fd.PipelineModel.load(os.path.join(d, "dialogue_classification_model"))
h = model.handle(0)
dev = torch.device("cuda", 0)
rng = np.random.default_rng(123)
alphabet = np.frombuffer(b"abcdefghijklmnopqrstuvwxyz      ABCXYZ.,!?0123456789\t\n\xc4\xb0\xe2\x84\xaa\xc3\xa9", dtype=np.uint8)
words = [b"the", b"a", b"officer", b"account", b"social", b"security", b"verify", b"card", b"you", b"payment"]
bad = 0
for it in range(iters):
    n = int(rng.integers(1, 300))
    rows = []
    for _ in range(n):
        kind = rng.integers(0, 10)
        if kind == 0:
            rows.append(b"")
        elif kind < 6:
            rows.append(b" ".join(words[int(i)] for i in rng.integers(0, len(words), size=int(rng.integers(1, 60)))))
        elif kind < 9:
            rows.append(alphabet[rng.integers(0, alphabet.size, size=int(rng.integers(1, 400)))].tobytes())
        else:
            rows.append(alphabet[rng.integers(0, alphabet.size, size=int(rng.integers(900, 9000)))].tobytes())
    total = sum(len(r) for r in rows)
    data = torch.empty(max(total, 1), dtype=torch.uint8).pin_memory()
    offs = torch.zeros(n + 1, dtype=torch.int64).pin_memory()
    buf = np.frombuffer(b"".join(rows), dtype=np.uint8)
    data.numpy()[:total] = buf
    offs.numpy()[1:] = np.cumsum([len(r) for r in rows])
    k = int(rng.integers(0, n))          # rows [k, n)
    m = n - k
    outs = [torch.full(s_, float("nan"), dtype=torch.float64).pin_memory() for s_ in ((m,), (m, 2), (m, 2))]
    h.call("fd_transform_host", P(data), P(offs[k:]), m, 1, P(outs[0]), P(outs[1]), P(outs[2]))
    # device path on the same rows
    pad = (16 - total % 16) % 16 + 16
    ddata = torch.cat([data, torch.zeros(pad, dtype=torch.uint8)]).to(dev)
    doffs = offs[k:].to(dev)
    douts = [torch.empty(s_, dtype=torch.float64, device=dev) for s_ in ((m,), (m, 2), (m, 2))]
    st = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    h.call("fd_transform", P(ddata), P(doffs), m, 1, P(douts[0]), P(douts[1]), P(douts[2]), st)
    h.call("fd_status", st)
    for a_, b_ in zip(outs, douts):
        if not np.array_equal(a_.numpy(), b_.cpu().numpy()):
            bad += 1
            print("MISMATCH at iteration", it, "rows", n, "from", k, flush=True)
            break
print("iterations", iters, "mismatches", bad)
sys.exit(1 if bad else 0)
