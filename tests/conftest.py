import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session", autouse=True)
def _built():
    """Build the product library (nvcc cross-compiles without a GPU) and the C oracle once per session."""
    from spatial_queue_b200 import build as b
    b.build()
    from oracle import sssp as osssp
    osssp.build(ref=os.path.exists("/root/reference"))
    yield


def grid_graph(n, seed=1, two_way=True):
    """SURVEY.md 8(d) synthetic grid recipe, as a COO edge list with fft weights (f64)."""
    import numpy as np
    rng = np.random.default_rng(seed)
    ii, jj = np.meshgrid(np.arange(n), np.arange(n), indexing="ij")
    lon = -122.9 + 0.0022 * ii + rng.normal(0, 1e-4, (n, n))
    lat = 37.6 + 0.0018 * jj + rng.normal(0, 1e-4, (n, n))
    nid = ii * n + jj
    s, e = [], []
    for di, dj in ((1, 0), (0, 1)):
        a = nid[: n - di, : n - dj].ravel()
        b = nid[di:, dj:].ravel()
        s += [a, b]
        e += [b, a]
    s = np.concatenate(s).astype(np.int32)
    e = np.concatenate(e).astype(np.int32)
    x = lon.ravel() * 88000.0
    y = lat.ravel() * 111000.0
    length = np.hypot(x[s] - x[e], y[s] - y[e])
    si, sj = s // n, s % n
    ei, ej = e // n, e % n
    arterial = ((si == ei) & (si % 8 == 0)) | ((sj == ej) & (sj % 8 == 0))
    mph = np.where(arterial, 35.0, 25.0)
    w = length / mph * 2.237
    return s, e, w
