"""Loader for tests/golden/*.npz (outputs of the unmodified reference; made by tests/golden/make_golden.py)."""
import os

import numpy as np

import oracle_lib as O

HERE = os.path.dirname(os.path.abspath(__file__))
GOLDEN = os.path.join(HERE, "golden")

# must match tests/golden/make_golden.py
CASES = {
    "u8_l2": ("uint8", "Euclidian"), "i8_l2": ("int8", "Euclidian"), "u8_mips": ("uint8", "mips"),
    "i8_mips": ("int8", "mips"), "f32_l2": ("float", "Euclidian"), "f32_mips": ("float", "mips"),
    "u8_l2_2pass": ("uint8", "Euclidian"), "u8_l2_big": ("uint8", "Euclidian"),
    "f32_l2_d100": ("float", "Euclidian"), "f32_mips_d100": ("float", "mips"),   # row length not a multiple of 8 elements
}
SEARCHES = [dict(k=10, beam=10), dict(k=10, beam=13), dict(k=10, beam=40), dict(k=5, beam=128, cut=1.1),
            dict(k=10, beam=20, limit=15), dict(k=10, beam=30, degree_limit=5), dict(k=1, beam=200)]
INT_CASES = [c for c, (dt, _) in CASES.items() if dt != "float"]
FLOAT_CASES = [c for c, (dt, _) in CASES.items() if dt == "float"]

_cache = {}


class Golden:
    def __init__(self, name):
        self.name = name
        self.data_type, self.metric = CASES[name]
        self.z = np.load(os.path.join(GOLDEN, name + ".npz"))
        self.n, self.d, self.nq, self.R, self.L, self.passes, self.seed, self.gt_k = (int(x) for x in self.z["params"])
        self.base = O.gen_points(self.data_type, self.n, self.d, seed=self.seed)
        self.queries = O.gen_points(self.data_type, self.nq, self.d, seed=self.seed, stream=1)
        self.graph = self.z["graph_unsorted"]

    def split_lists(self, counts, flat):
        out, pos = [], 0
        for c in counts:
            out.append(flat[pos:pos + int(c)])
            pos += int(c)
        assert pos == len(flat)
        return out


def golden(name):
    if name not in _cache:
        _cache[name] = Golden(name)
    return _cache[name]


def valid_graph(g):
    g = np.array(g, dtype=np.uint32, copy=True)
    mask = np.arange(g.shape[1] - 1, dtype=np.uint32)[None, :] < g[:, 0][:, None]
    g[:, 1:][~mask] = 0
    return g


def rows_equal_as_sets(a, b):
    a, b = valid_graph(a), valid_graph(b)
    if not np.array_equal(a[:, 0], b[:, 0]):
        return False
    return np.array_equal(np.sort(a[:, 1:], axis=1), np.sort(b[:, 1:], axis=1))
