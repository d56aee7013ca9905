"""The C-ABI library: loads, exports every symbol include/gentun_b200.h declares, and fails loudly
(instead of falling back to a CPU path) when there is no CUDA device.  No GPU compute here."""
import ctypes as C
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_functions():
    text = open(os.path.join(ROOT, "include", "gentun_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(gt_[a-z_0-9]+)\s*\(", text)))


def test_header_functions_are_exported():
    from gentun_b200 import engine
    lib = engine.load_library()
    names = declared_functions()
    assert len(names) >= 13
    for name in names:
        assert hasattr(lib, name), name
    for name in engine.EXPORTS:
        assert name in names


def test_every_declaration_cites_the_reference_or_says_it_has_no_counterpart():
    text = open(os.path.join(ROOT, "include", "gentun_b200.h")).read()
    assert "xgboost_models.py:32" in text and "xgboost_models.py:33-36" in text and "populations.py:55-58" in text


def test_fold_construction_matches_numpy():
    from gentun_b200 import engine
    for n, nfold in [(4898, 3), (1000, 7), (2, 2), (65537, 5)]:
        perm, fb = engine.make_folds(n, nfold)
        np.random.seed(0)
        ref = np.random.permutation(n)
        assert np.array_equal(perm, ref)
        assert np.diff(fb).tolist() == [len(c) for c in np.array_split(ref, nfold)]


def test_struct_layouts_match_the_header():
    from gentun_b200 import engine
    assert C.sizeof(engine.DataOpts) == 32
    assert C.sizeof(engine.FitOpts) == 32
    assert engine.TREE_NODE.itemsize == 16 and engine.NODE_ROWS.itemsize == 16
    assert C.sizeof(engine.Stats) == 72


def _have_cuda():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


@pytest.mark.skipif(_have_cuda(), reason="checks the no-GPU failure mode")
def test_no_device_means_a_loud_error_not_a_cpu_fallback():
    from gentun_b200 import engine
    x = np.zeros((10, 2), dtype=np.float32)
    y = np.zeros(10, dtype=np.float32)
    with pytest.raises(engine.EngineError) as err:
        engine.Engine(x, y)
    assert "no CPU path" in str(err.value)
    import gentun_b200 as gentun
    model = gentun.XgboostModel(x, y, {k: 1.0 for k in engine.GENE_ORDER}, nfold=2)
    with pytest.raises(engine.EngineError):
        model.cross_validate()


def test_missing_library_is_reported(monkeypatch):
    from gentun_b200 import engine
    monkeypatch.setattr(engine, "_lib", None)
    monkeypatch.setattr(engine, "LIB_PATH", os.path.join(ROOT, "gentun_b200", "does_not_exist.so"))
    with pytest.raises(engine.EngineError) as err:
        engine.load_library()
    assert "no CPU fallback" in str(err.value)


def test_product_never_imports_the_oracle():
    for dirpath, _, files in os.walk(os.path.join(ROOT, "gentun_b200")):
        for fn in files:
            if fn.endswith((".py", ".cu", ".cuh", ".h", ".sh")):
                text = open(os.path.join(dirpath, fn)).read()
                assert "oracle" not in text.replace("the oracle by specification", "").replace("with the oracle by", ""), fn
