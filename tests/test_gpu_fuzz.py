"""Short runs of the fuzzers under scripts/ (the long runs are manual: 400 trainer problems, 1,500 host batches).

* fuzz_trainer.py: random small training problems (sizes, depths, bag weights, subset sizes, label noise) -- the CUDA
  trainer's trees must equal the oracle's, which also exercises the default-link chains of deep, impure trees.
* fuzz_direct_host.py: random batches through fd_transform_host with pinned buffers (read and written in place) against
  fd_transform on a device copy of the same rows."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("script,iterations", [("fuzz_trainer.py", 16), ("fuzz_direct_host.py", 60)])
def test_fuzzers(fd, script, iterations):
    r = subprocess.run([sys.executable, os.path.join(ROOT, "scripts", script), str(iterations)], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "mismatches 0" in r.stdout
