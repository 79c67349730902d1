"""fd_transform_noalloc -- the transform for callers that cannot allocate device memory (a GPU whose memory a
pool allocator already owns; a micro-batch loop that must not hit the driver's allocation lock).  Its scratch is
a fixed array inside the library, shared by every call, so what matters here is that (1) it returns exactly what
fd_transform returns, (2) a second call right after the first -- with the array still holding the first call's
leftovers -- is just as correct, and (3) a batch over the fixed limit is refused with a clear error.
Collected last (test_zz_*) so a run that stops at the first failure has already run everything else; it builds
its own model and handle, so it also passes when run alone in a fresh process."""
import ctypes as C
import importlib

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _model(fd):
    synth = importlib.import_module("fraud-detection-spark-kafka-llm_b200.synth")
    rng = np.random.default_rng(17)
    nf = 1 << 16
    corpus = synth.SynthCorpus(n_words=4000, seed=31)
    feats = synth.hash_terms(corpus.words[corpus.n_stop:corpus.n_stop + 300] + [b""], nf)
    idf = np.log(2001.0 / (rng.integers(0, 2000, size=nf) + 1.0))
    stages = [fd.Tokenizer(inputCol="clean_text", outputCol="words"), fd.StopWordsRemover(inputCol="words", outputCol="filtered_words"),
              fd.HashingTF(inputCol="filtered_words", outputCol="raw_features", numFeatures=nf),
              fd.IDFModel(idf, inputCol="raw_features", outputCol="features"),
              fd.RandomForestClassificationModel([synth.random_tree(rng, 6, feats, idf, full=False) for _ in range(25)], nf)]
    return fd.PipelineModel(stages), corpus


def _both(fd, model, pk, flags=1):
    """-> (fd_transform outputs, fd_transform_noalloc outputs, status) for the same packed dialogues."""
    lib = importlib.import_module("fraud-detection-spark-kafka-llm_b200._lib")
    h = model.handle(0)
    dev = torch.device("cuda", 0)
    pk = pk.to_device(dev)
    n = pk.n_rows
    st = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    ref = [torch.empty(n, dtype=torch.float64, device=dev), torch.empty((n, 2), dtype=torch.float64, device=dev),
           torch.empty((n, 2), dtype=torch.float64, device=dev)]
    out = [torch.full_like(t, float("nan")) for t in ref]
    status = torch.full((1,), 77, dtype=torch.int32, device=dev)
    h.call("fd_transform", lib.ptr(pk.data), lib.ptr(pk.offs), n, flags, *(lib.ptr(t) for t in ref), st)
    h.call("fd_transform_noalloc", lib.ptr(pk.data), lib.ptr(pk.offs), n, flags, *(lib.ptr(t) for t in out), lib.ptr(status), st)
    torch.cuda.synchronize(dev)
    return ref, out, int(status.item())


def test_noalloc_matches_fd_transform_on_two_calls_in_a_row(fd):
    model, corpus = _model(fd)
    rng = np.random.default_rng(4)
    wide = " ".join(bytes(rng.integers(97, 123, size=6, dtype=np.uint8)).decode() for _ in range(1500))  # > 384 distinct: wide table
    a, _ = corpus.sample(3000, tokens=90, device="cpu", seed=1)
    da, oa = a.data.numpy(), a.offs.numpy()
    rows_a = [bytes(da[oa[i]:oa[i + 1]]) for i in range(3000)] + [wide.encode(), b"", b"  ", b"Hello, WORLD!! the end"]
    b, _ = corpus.sample(500, tokens=300, device="cpu", seed=2, skew=True)
    db, ob = b.data.numpy(), b.offs.numpy()
    rows_b = [wide[: len(wide) // 2].encode()] + [bytes(db[ob[i]:ob[i + 1]]) for i in range(500)]
    for rows in (rows_a, rows_b):  # the second call runs on the scratch the first one left behind
        ref, out, status = _both(fd, model, fd.PackedText.from_strings(rows))
        assert status == 0
        for r, o in zip(ref, out):
            assert torch.equal(r, o)  # same kernels, same arithmetic: bit-identical (tolerance 0, as the integer/label bar)
        assert not torch.isnan(out[0]).any() and len(torch.unique(out[2][:, 1])) > 3
    # ... and against the CPU oracle itself (a slice keeps the pure-Python walk short), not only against fd_transform
    import os
    import sys
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    from helpers import oracle_for
    rows = rows_a[:250] + rows_a[-4:]
    _, out, status = _both(fd, model, fd.PackedText.from_strings(rows))
    pred, prob, raw = oracle_for(model).transform(rows)
    assert status == 0 and np.array_equal(out[0].cpu().numpy(), pred) and np.array_equal(out[2].cpu().numpy(), raw)
    np.testing.assert_allclose(out[1].cpu().numpy(), prob, rtol=1e-9, atol=1e-300)


def test_noalloc_largest_batch_and_tokenizer_only_text(fd):
    """The fixed limit itself (65,536 dialogues: the kernel takes its large-batch shape there) and flags = 0, i.e. text that
    was cleaned by the caller: both run other instantiations of the kernels than the test above."""
    model, corpus = _model(fd)
    pk, _ = corpus.sample(65536, tokens=40, device="cuda", seed=8)
    for flags in (1, 0):
        ref, out, status = _both(fd, model, pk, flags)
        assert status == 0
        for r, o in zip(ref, out):
            assert torch.equal(r, o)
        assert len(torch.unique(out[2][:, 1])) > 3


def test_noalloc_reports_an_over_limit_dialogue_through_its_status_word(fd):
    """A dialogue with more distinct hashed terms than the widest count table holds (about 49k) cannot be scored; the call is
    asynchronous, so the condition comes back through out_status_dev.  (Only a classifier that sees most features can
    get there: a forest or a sparse LogisticRegression drops the terms it cannot see before they are counted.)"""
    lib = importlib.import_module("fraud-detection-spark-kafka-llm_b200._lib")
    rng = np.random.default_rng(3)
    nf = 1 << 18
    stages = [fd.Tokenizer(inputCol="clean_text", outputCol="words"),
              fd.HashingTF(inputCol="words", outputCol="features", numFeatures=nf),
              fd.LogisticRegressionModel(rng.normal(0.0, 1.0, size=nf), 0.1)]
    model = fd.PipelineModel(stages)
    letters = rng.integers(97, 123, size=(70000, 7), dtype=np.uint8)
    huge = b" ".join(bytes(r) for r in letters)  # ~70k distinct 7-letter words -> ~61k distinct buckets of 2^18
    ref, out, status = _both(fd, model, fd.PackedText.from_strings([b"fine one", huge, b"fine two"]))
    assert status == lib.FD_ERR_LIMIT
    assert torch.equal(ref[0][[0, 2]], out[0][[0, 2]]) and not torch.isnan(out[0][[0, 2]]).any()
    h = model.handle(0)
    with pytest.raises(fd.FdError) as e:  # the plain call reports the same condition through fd_status
        h.call("fd_status", None)
    assert e.value.code == lib.FD_ERR_LIMIT
    # the next call starts from a clean status word
    _, _, status = _both(fd, model, fd.PackedText.from_strings([b"all good now"]))
    assert status == 0


def test_noalloc_rejects_batches_over_its_fixed_limit(fd):
    lib = importlib.import_module("fraud-detection-spark-kafka-llm_b200._lib")
    model, _ = _model(fd)
    h = model.handle(0)
    n = 65536 + 1
    dev = torch.device("cuda", 0)
    data = torch.zeros(64, dtype=torch.uint8, device=dev)
    offs = torch.zeros(n + 1, dtype=torch.int64, device=dev)
    outs = [torch.empty(n, dtype=torch.float64, device=dev), torch.empty((n, 2), dtype=torch.float64, device=dev),
            torch.empty((n, 2), dtype=torch.float64, device=dev)]
    with pytest.raises(fd.FdError) as e:
        h.call("fd_transform_noalloc", lib.ptr(data), lib.ptr(offs), n, 1, *(lib.ptr(t) for t in outs), None, None)
    assert e.value.code == lib.FD_ERR_LIMIT and "65536" in str(e.value)
    # an empty batch is fine and touches nothing
    h.call("fd_transform_noalloc", lib.ptr(data), lib.ptr(offs), 0, 1, None, None, None, None, None)
