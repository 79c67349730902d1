"""The C-ABI library loads and exports every symbol include/fdb200.h declares (no compute calls here)."""
import ctypes
import os
import re

from conftest import ROOT


def _declared():
    src = open(os.path.join(ROOT, "include", "fdb200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(fd_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol(fd):
    names = _declared()
    assert "fd_transform" in names and "fd_featurize" in names and len(names) >= 20
    lib = ctypes.CDLL(fd.LIB_PATH)
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, f"declared in fdb200.h but not exported: {missing}"


def test_binding_covers_header(fd):
    from importlib import import_module
    sig = import_module("fraud-detection-spark-kafka-llm_b200._lib").SIGNATURES
    assert sorted(sig) == _declared()
    lib = fd.load_library()
    assert lib.fd_version() >= 100


def test_host_side_hash_terms(fd):
    from importlib import import_module
    synth = import_module("fraud-detection-spark-kafka-llm_b200.synth")
    out = synth.hash_terms([b"innocent", b"", b"suspect", b"hello"], 10000)
    assert out.tolist() == [2833, 3372, 9168, 9889]
    assert synth.hash_terms([b"hello"], 262144).tolist() == [(-488910111) % 262144]
