"""Points / graph / ground-truth file formats of the reference, as numpy readers and writers.

Formats (all little-endian, two-u32 headers):
  * points  ``.fbin/.u8bin/.i8bin``: ``u32 n, u32 d`` then ``n*d`` elements row-major
    (reference: algorithms/utils/point_range.h:86-116, python/_files.py:42-90).
  * graph: ``u32 n, u32 maxDeg, u32 deg[n]`` then the concatenated edge lists
    (reference: algorithms/utils/graph.h:147-204 load, :206-232 save).  In memory the graph is the
    reference's flat ``n x (maxDeg+1)`` uint32 array with slot 0 = degree (graph.h:134-141).
  * ground truth ``.ibin``: ``i32 nq, i32 k`` then ``nq*k`` i32 ids then ``nq*k`` f32 distances
    (reference: data_tools/compute_groundtruth.cpp:94-99 write, algorithms/utils/types.h:48-73 read).
"""
import numpy as np

_DTYPES = {"float": np.float32, "uint8": np.uint8, "int8": np.int8}


def np_dtype(data_type):
    if data_type not in _DTYPES:
        raise ValueError(f"data_type must be one of {sorted(_DTYPES)}, got {data_type!r}")
    return _DTYPES[data_type]


def write_bin(path, arr):
    arr = np.ascontiguousarray(arr)
    assert arr.ndim == 2
    with open(path, "wb") as f:
        np.array(arr.shape, dtype=np.uint32).tofile(f)
        arr.tofile(f)


def read_bin(path, data_type):
    dt = np_dtype(data_type)
    with open(path, "rb") as f:
        n, d = np.fromfile(f, dtype=np.uint32, count=2)
        a = np.fromfile(f, dtype=dt, count=int(n) * int(d))
    if a.size != int(n) * int(d):
        raise ValueError(f"{path}: truncated points file")
    return a.reshape(int(n), int(d))


def write_graph(path, graph):
    """graph: [n, maxDeg+1] uint32, slot 0 = degree."""
    graph = np.ascontiguousarray(graph, dtype=np.uint32)
    n, w = graph.shape
    deg = graph[:, 0]
    mask = np.arange(w - 1, dtype=np.uint32)[None, :] < deg[:, None]
    with open(path, "wb") as f:
        np.array([n, w - 1], dtype=np.uint32).tofile(f)
        deg.tofile(f)
        graph[:, 1:][mask].tofile(f)


def read_graph(path):
    with open(path, "rb") as f:
        n, maxdeg = (int(x) for x in np.fromfile(f, dtype=np.uint32, count=2))
        deg = np.fromfile(f, dtype=np.uint32, count=n)
        edges = np.fromfile(f, dtype=np.uint32, count=int(deg.sum()))
    if deg.size != n or edges.size != int(deg.sum()) or (n and deg.max() > maxdeg):
        raise ValueError(f"{path}: malformed graph file")
    g = np.zeros((n, maxdeg + 1), dtype=np.uint32)
    g[:, 0] = deg
    mask = np.arange(maxdeg, dtype=np.uint32)[None, :] < deg[:, None]
    g[:, 1:][mask] = edges
    return g


def write_ibin(path, ids, dists):
    ids = np.ascontiguousarray(ids, dtype=np.int32)
    dists = np.ascontiguousarray(dists, dtype=np.float32)
    assert ids.shape == dists.shape and ids.ndim == 2
    with open(path, "wb") as f:
        np.array(ids.shape, dtype=np.int32).tofile(f)
        ids.tofile(f)
        dists.tofile(f)


def read_ibin(path):
    with open(path, "rb") as f:
        nq, k = (int(x) for x in np.fromfile(f, dtype=np.int32, count=2))
        ids = np.fromfile(f, dtype=np.int32, count=nq * k).reshape(nq, k)
        dists = np.fromfile(f, dtype=np.float32, count=nq * k).reshape(nq, k)
    return ids, dists
