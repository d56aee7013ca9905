"""gt_column_cuts_noalloc: the cut points of one column computed with no device allocation at all.

Some callers of this library cannot allocate when they need cut points: a serving loop that re-bins a
feature while other CUDA streams are busy (cudaMalloc / cudaFree stall the whole device), a process
that runs inside a fixed memory reservation, a hook that is only allowed to enqueue work.  For them the
library keeps one fixed scratch array on each GPU and runs the same sketch kernel on it.  This test
calls that entry point through the C ABI, twice in a row on one stream (the scratch array still holds
the first call's contents when the second starts), and requires the result to be bit-identical to what
the allocating path (gt_engine_prepare -> gt_engine_get_cuts) produced for the same column.

The file name sorts last on purpose: if this addition ever breaks, `pytest -x` has already run every
other test.
"""
import ctypes as C

import numpy as np
import pytest

from tests.synth import make_dataset

pytestmark = pytest.mark.gpu


def _call(lib, values_dev, cuts_dev, ncuts_dev, stream):
    return lib.gt_column_cuts_noalloc(C.c_void_p(values_dev.data_ptr()), values_dev.numel(),
                                      C.c_void_p(cuts_dev.data_ptr()), C.c_void_p(ncuts_dev.data_ptr()),
                                      C.c_void_p(stream))


def _check_two_in_a_row(x, nfold, fold, columns):
    import torch
    from gentun_b200 import engine
    lib = engine.load_library()
    y = np.zeros(x.shape[0], dtype=np.float32)
    eng = engine.Engine(x, y)
    want_cuts, want_n = eng.cuts(nfold, fold)
    train_rows, _ = eng.fold_rows(nfold, fold)
    assert len(train_rows) <= lib.gt_column_cuts_noalloc_max_rows()
    stream = torch.cuda.current_stream().cuda_stream
    outs = []
    for j in columns:       # back to back, no synchronisation in between
        col = torch.from_numpy(np.ascontiguousarray(x[train_rows, j])).cuda()
        cuts = torch.full((256,), float("nan"), dtype=torch.float32, device="cuda")
        ncuts = torch.zeros(1, dtype=torch.int32, device="cuda")
        assert _call(lib, col, cuts, ncuts, stream) == 0, lib.gt_last_error()
        outs.append((j, col, cuts, ncuts))
    torch.cuda.synchronize()
    for j, _, cuts, ncuts in outs:
        n = int(ncuts.item())
        assert n == want_n[j], (j, n, want_n[j])
        got = cuts.cpu().numpy()[:n]
        assert np.array_equal(got.view(np.uint32), want_cuts[j, :n].view(np.uint32)), j   # bit-exact, as in test_gpu_parity
    eng.close()


def test_noalloc_cuts_equal_engine_cuts_small_columns(wine):
    x = np.ascontiguousarray(wine[0].values, dtype=np.float32)
    _check_two_in_a_row(x, nfold=3, fold=1, columns=[3, 7])      # the two columns with > 256 distinct values


def test_noalloc_cuts_equal_engine_cuts_multilevel_sketch():
    x, _ = make_dataset(60000, 12, seed=77)                        # 45000 train rows: the queue flushes several times
    _check_two_in_a_row(x, nfold=4, fold=2, columns=[0, 9])
    _check_two_in_a_row(x, nfold=4, fold=0, columns=[8, 8])      # same column twice: stale scratch must not matter


def test_noalloc_rejects_sizes_beyond_the_fixed_scratch():
    import torch
    from gentun_b200 import engine
    lib = engine.load_library()
    limit = lib.gt_column_cuts_noalloc_max_rows()
    assert limit == 65536
    col = torch.zeros(limit + 1, dtype=torch.float32, device="cuda")
    cuts = torch.zeros(256, dtype=torch.float32, device="cuda")
    ncuts = torch.zeros(1, dtype=torch.int32, device="cuda")
    rc = _call(lib, col, cuts, ncuts, torch.cuda.current_stream().cuda_stream)
    assert rc == -1
    assert b"65536" in lib.gt_last_error()
    torch.cuda.synchronize()
