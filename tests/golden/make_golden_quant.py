"""Golden vectors for the 1-byte quantised search + exact re-rank (the reference's -quantize_mode 1 path:
vamana/neighbors.h:117-127 -> qsearchAll with Q ranges -> beam_search_rerank, beamSearch.h:390-454), produced by the
UNMODIFIED reference through oracle/_ref/ref_driver qsearch.  Run in the container that has /root/reference:
    python tests/golden/make_golden_quant.py
Inputs are regenerated from the counter-based generator (seed 77), so only outputs are stored."""
import os
import re
import sys
import tempfile
import zlib

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, ROOT)
import oracle_lib as O  # noqa: E402
from parlayann_b200 import fileio  # noqa: E402

CASES = [("quant_f32_l2", 8192, 128, 200, 32, 64), ("quant_f32_l2_d100", 4096, 100, 100, 32, 64)]
BEAMS = [10, 16, 40, 100]


def main():
    assert O.ref_driver_path(), "build oracle/_ref first (make -C oracle ref)"
    tmp = tempfile.mkdtemp(prefix="pann_goldenq_")
    for name, n, d, nq, R, L in CASES:
        base = O.gen_points("float", n, d, seed=77)
        qs = O.gen_points("float", nq, d, seed=77, stream=1)
        bp, qp = os.path.join(tmp, name + ".base.fbin"), os.path.join(tmp, name + ".query.fbin")
        fileio.write_bin(bp, base); fileio.write_bin(qp, qs)
        common = ["-data_type", "float", "-dist_func", "Euclidian", "-base", bp]
        out = os.path.join(tmp, name)
        O.ref_run(["build"] + common + ["-R", R, "-L", L, "-alpha", 1.2, "-passes", 1, "-sort", 1, "-out", out + ".g"])
        g = {"graph": fileio.read_graph(out + ".g.graph")}
        for beam in BEAMS:
            txt = O.ref_run(["qsearch"] + common + ["-graph", out + ".g.graph", "-query", qp, "-k", 10, "-beam", beam, "-out", out + f".q{beam}"])
            g[f"ids_{beam}"] = np.fromfile(out + f".q{beam}.ids", dtype=np.uint32).reshape(nq, 10)
            g[f"visited_{beam}"] = np.fromfile(out + f".q{beam}.visited", dtype=np.uint32)
            g[f"cmps_{beam}"] = np.fromfile(out + f".q{beam}.distances", dtype=np.uint32)
        m = re.search(r"REF_QUANT slope (\S+) offset (-?\d+)", txt)
        g["slope"] = np.float32(float(m.group(1))); g["offset"] = np.int32(int(m.group(2)))
        qb = np.fromfile(out + f".q{BEAMS[-1]}.qbase", dtype=np.uint8).reshape(n, d)
        g["qbase_crc"] = np.uint32(zlib.crc32(qb.tobytes()))
        g["qbase_head"] = qb[:64].copy()
        g["qquery"] = np.fromfile(out + f".q{BEAMS[-1]}.qquery", dtype=np.uint8).reshape(nq, d)
        np.savez_compressed(os.path.join(HERE, name + ".npz"), **g)
        print(name, "slope", g["slope"], "offset", g["offset"], {k: v.shape for k, v in g.items() if hasattr(v, "shape") and v.ndim})


MIPS_CASES = [("quant_f32_mips", 8192, 128, 200, 32, 64), ("quant_f32_mips_d200", 4096, 200, 100, 32, 64)]


def main_mips():
    """Float MIPS data: normalised rows, Quantized_Mips_Point<8,true> copy (python/graph_index.cpp:88-95, :166-181), the
    graph builder.cpp:45-54, :73-86 intends (ref_driver qbuild: built over the quantised copy, alpha overridden)."""
    tmp = tempfile.mkdtemp(prefix="pann_goldenqm_")
    for name, n, d, nq, R, L in MIPS_CASES:
        base = O.gen_points("float", n, d, seed=78)
        qs = O.gen_points("float", nq, d, seed=78, stream=1)
        bp, qp = os.path.join(tmp, name + ".base.fbin"), os.path.join(tmp, name + ".query.fbin")
        fileio.write_bin(bp, base); fileio.write_bin(qp, qs)
        common = ["-data_type", "float", "-dist_func", "mips", "-base", bp]
        out = os.path.join(tmp, name)
        txt = O.ref_run(["qbuild"] + common + ["-R", R, "-L", L, "-alpha", 1.2, "-passes", 1, "-out", out + ".g"])
        g = {"graph": fileio.read_graph(out + ".g.graph")}
        g["max_val"] = np.float32(float(re.search(r"REF_QUANT max_val (\S+)", txt).group(1)))
        qb = np.fromfile(out + ".g.qbase", dtype=np.int8).reshape(n, d)
        for beam in BEAMS:
            txt = O.ref_run(["qsearch"] + common + ["-graph", out + ".g.graph", "-query", qp, "-k", 10, "-beam", beam, "-out", out + f".q{beam}"])
            g[f"ids_{beam}"] = np.fromfile(out + f".q{beam}.ids", dtype=np.uint32).reshape(nq, 10)
            g[f"visited_{beam}"] = np.fromfile(out + f".q{beam}.visited", dtype=np.uint32)
            g[f"cmps_{beam}"] = np.fromfile(out + f".q{beam}.distances", dtype=np.uint32)
        assert np.float32(float(re.search(r"REF_QUANT max_val (\S+)", txt).group(1))) == g["max_val"]
        qb2 = np.fromfile(out + f".q{BEAMS[-1]}.qbase", dtype=np.int8).reshape(n, d)
        assert np.array_equal(qb, qb2)
        g["qbase_crc"] = np.uint32(zlib.crc32(qb.tobytes()))
        g["qbase_head"] = qb[:64].copy()
        g["qquery"] = np.fromfile(out + f".q{BEAMS[-1]}.qquery", dtype=np.int8).reshape(nq, d)
        g["nbase_head"] = np.fromfile(out + f".q{BEAMS[-1]}.nbase_head", dtype=np.float32).reshape(64, d)
        g["nquery_crc"] = np.uint32(zlib.crc32(np.fromfile(out + f".q{BEAMS[-1]}.nquery", dtype=np.float32).tobytes()))
        np.savez_compressed(os.path.join(HERE, name + ".npz"), **g)
        print(name, "max_val", g["max_val"], {k: v.shape for k, v in g.items() if hasattr(v, "shape") and v.ndim})


if __name__ == "__main__":
    if "mips" not in sys.argv[1:]:
        main()
    if "l2" not in sys.argv[1:]:
        main_mips()
