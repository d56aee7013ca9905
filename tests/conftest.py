import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def wine():
    import pandas as pd
    data = pd.read_csv(os.path.join(ROOT, "tests", "data", "winequality-white.csv"), delimiter=";")
    return data.drop(["quality"], axis=1), data["quality"]


DEFAULT_GENES = {
    'eta': 0.3, 'min_child_weight': 1, 'max_depth': 6, 'gamma': 0.0, 'max_delta_step': 0,
    'subsample': 1.0, 'colsample_bytree': 1.0, 'colsample_bylevel': 1.0, 'lambda': 1.0,
    'alpha': 0.0, 'scale_pos_weight': 1.0,
}


@pytest.fixture(scope="session")
def default_genes():
    return dict(DEFAULT_GENES)
