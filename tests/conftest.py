import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


_EXT = {"uint8": "u8bin", "int8": "i8bin", "float": "fbin"}


class Dataset:
    """Seeded synthetic base/query sets plus a graph built by the CPU oracle (cached per session)."""

    def __init__(self, data_type, metric, n, d, nq, R, L, r=24, nc=256, kind=None):
        import oracle_lib as O
        self.data_type, self.metric, self.n, self.d, self.nq, self.R, self.L = data_type, metric, n, d, nq, R, L
        kind = kind or data_type
        self.base = O.gen_points(kind, n, d, r=r, nc=nc)
        self.queries = O.gen_points(kind, nq, d, stream=1, r=r, nc=nc)
        self._graph = None

    @property
    def graph(self):
        if self._graph is None:
            import oracle_lib as O
            g, _, _ = O.vamana_build(self.base, self.data_type, self.metric, R=self.R, L=self.L, sort_neighbors=False)
            self._graph = g
        return self._graph


_cache = {}


@pytest.fixture(scope="session")
def dataset():
    def get(data_type="uint8", metric="Euclidian", n=20000, d=128, nq=500, R=32, L=64, **kw):
        key = (data_type, metric, n, d, nq, R, L, tuple(sorted(kw.items())))
        if key not in _cache:
            _cache[key] = Dataset(data_type, metric, n, d, nq, R, L, **kw)
        return _cache[key]
    return get


def valid_graph(g):
    """zero the slots beyond each row's degree so graphs compare as the file format would"""
    g = np.array(g, dtype=np.uint32, copy=True)
    mask = np.arange(g.shape[1] - 1, dtype=np.uint32)[None, :] < g[:, 0][:, None]
    g[:, 1:][~mask] = 0
    return g
