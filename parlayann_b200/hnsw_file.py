"""HNSW model file, format version 3 (reference: algorithms/HNSW/HNSW.hpp:1615-1686 save, :631-744 load).

Layout (little endian): "HNSW", u32 version = 3, u64 code_U, u64 size_node (both ignored by the loader), u32 dim, f32 m_l,
u32 m, u32 ef_construction, f32 alpha, u32 n; n x (u32 level, u32 point id); for every node, for l = 0..level: u64 count,
count x u32 neighbour (node) ids; u64 number of entrances, the entrances (u32 each).

In memory the adjacency is kept in the file's own order as two flat arrays: counts[(node, level)] and edges."""
import numpy as np


class HnswModel:
    def __init__(self, dim, m_l, m, efc, alpha, level, counts, edges, entrance, code_U=0, size_node=0, ids=None):
        self.dim, self.m_l, self.m, self.efc, self.alpha = int(dim), float(np.float32(m_l)), int(m), int(efc), float(np.float32(alpha))
        self.level = np.ascontiguousarray(level, dtype=np.uint32)
        self.counts = np.ascontiguousarray(counts, dtype=np.uint64)
        self.edges = np.ascontiguousarray(edges, dtype=np.uint32)
        self.entrance = np.ascontiguousarray(entrance, dtype=np.uint32)
        self.code_U, self.size_node = int(code_U), int(size_node)
        self.n = len(self.level)
        self.ids = np.arange(self.n, dtype=np.uint32) if ids is None else np.ascontiguousarray(ids, dtype=np.uint32)
        assert len(self.counts) == int(self.level.astype(np.int64).sum()) + self.n and len(self.edges) == int(self.counts.sum())

    def level_graph(self, l):
        """(node ids with level >= l, their neighbour lists at level l as a list of arrays)"""
        first = np.concatenate([[0], np.cumsum(self.level.astype(np.int64) + 1)])[:-1]
        eoff = np.concatenate([[0], np.cumsum(self.counts.astype(np.int64))])
        nodes = np.nonzero(self.level >= l)[0]
        return nodes, [self.edges[eoff[first[v] + l]:eoff[first[v] + l + 1]] for v in nodes]


def read_hnsw(path):
    buf = np.fromfile(path, dtype=np.uint8)
    if bytes(buf[:4]) != b"HNSW":
        raise ValueError("Wrong type of model")
    version = int(buf[4:8].view(np.uint32)[0])
    if version != 3:
        raise ValueError("Unsupported version")
    code_U, size_node = (int(x) for x in buf[8:24].view(np.uint64))
    dim = int(buf[24:28].view(np.uint32)[0]); m_l = float(buf[28:32].view(np.float32)[0])
    m, efc = (int(x) for x in buf[32:40].view(np.uint32)); alpha = float(buf[40:44].view(np.float32)[0])
    n = int(buf[44:48].view(np.uint32)[0])
    pos = 48
    li = buf[pos:pos + 8 * n].view(np.uint32).reshape(n, 2)
    level, ids = li[:, 0].copy(), li[:, 1].copy()
    pos += 8 * n
    nc = int(level.astype(np.int64).sum()) + n
    counts = np.empty(nc, np.uint64)
    chunks = []
    # the adjacency is a stream of (u64 count, count x u32): walked 4 bytes at a time
    w = buf[pos:].view(np.uint32) if (len(buf) - pos) % 4 == 0 else buf[pos:pos + (len(buf) - pos) // 4 * 4].view(np.uint32)
    p = 0
    for i in range(nc):
        c = int(w[p]) | (int(w[p + 1]) << 32)
        counts[i] = c
        chunks.append(w[p + 2:p + 2 + c])
        p += 2 + c
    nent = int(w[p]) | (int(w[p + 1]) << 32)
    entrance = w[p + 2:p + 2 + nent].copy()
    edges = np.concatenate(chunks) if chunks else np.empty(0, np.uint32)
    return HnswModel(dim, m_l, m, efc, alpha, level, counts, edges, entrance, code_U, size_node, ids)


def write_hnsw(path, M):
    nc = len(M.counts)
    out = np.empty(2 * nc + len(M.edges), dtype=np.uint32)
    eoff = np.concatenate([[0], np.cumsum(M.counts.astype(np.int64))])
    cpos = eoff[:-1] + 2 * np.arange(nc)                # word offset of each count
    out[cpos] = (M.counts & np.uint64(0xffffffff)).astype(np.uint32)
    out[cpos + 1] = (M.counts >> np.uint64(32)).astype(np.uint32)
    mask = np.ones(len(out), bool)
    mask[cpos] = False; mask[cpos + 1] = False
    out[mask] = M.edges
    with open(path, "wb") as f:
        f.write(b"HNSW")
        f.write(np.array([3], np.uint32).tobytes())
        f.write(np.array([M.code_U, M.size_node], np.uint64).tobytes())
        f.write(np.array([M.dim], np.uint32).tobytes()); f.write(np.array([M.m_l], np.float32).tobytes())
        f.write(np.array([M.m, M.efc], np.uint32).tobytes()); f.write(np.array([M.alpha], np.float32).tobytes())
        f.write(np.array([M.n], np.uint32).tobytes())
        f.write(np.stack([M.level, M.ids], axis=1).astype(np.uint32).tobytes())
        f.write(out.tobytes())
        f.write(np.array([len(M.entrance)], np.uint64).tobytes())
        f.write(M.entrance.astype(np.uint32).tobytes())
