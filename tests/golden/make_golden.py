"""Generates tests/golden/*.npz by running the UNMODIFIED reference (oracle/_ref/ref_driver, built from
/root/reference by oracle/Makefile) on small seeded inputs.  Run in the build container only:

    python tests/golden/make_golden.py

Each fixture holds the generator parameters of the inputs (they are re-created bit-identically from
parlayann_b200/csrc/datagen.h) plus the reference's raw outputs: graph files, ids, per-query counters,
frontier/visited lists, ground truth, permutations.  tests/test_oracle_vs_reference.py pins
oracle/pann_oracle.cpp against them; the -m gpu tests pin the CUDA path against them too."""
import os
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import oracle_lib as O  # noqa: E402
from parlayann_b200 import fileio  # noqa: E402

EXT = {"uint8": "u8bin", "int8": "i8bin", "float": "fbin"}

# (name, data_type, metric, n, d, nq, R, L, passes)
CASES = [
    ("u8_l2", "uint8", "Euclidian", 3000, 64, 60, 16, 32, 1),
    ("i8_l2", "int8", "Euclidian", 2000, 48, 40, 12, 24, 1),
    ("u8_mips", "uint8", "mips", 2000, 32, 40, 12, 24, 1),
    ("i8_mips", "int8", "mips", 2000, 40, 40, 12, 24, 1),
    ("f32_l2", "float", "Euclidian", 2000, 32, 40, 12, 24, 1),
    ("f32_mips", "float", "mips", 1500, 24, 30, 12, 24, 1),
    ("u8_l2_2pass", "uint8", "Euclidian", 1500, 32, 30, 8, 16, 2),
    ("u8_l2_big", "uint8", "Euclidian", 9000, 16, 30, 8, 16, 1),   # n >= 8192: the bucketed random_permutation
    # float rows whose length is not a multiple of 8 (or 16) elements: whatever the reference's compiler does with the tail
    # of its distance loop (vector body / scalar remainder, contraction) is in these outputs
    ("f32_l2_d100", "float", "Euclidian", 2000, 100, 40, 12, 24, 1),
    ("f32_mips_d100", "float", "mips", 1500, 100, 30, 12, 24, 1),
]
SEARCHES = [dict(k=10, beam=10), dict(k=10, beam=13), dict(k=10, beam=40), dict(k=5, beam=128, cut=1.1),
            dict(k=10, beam=20, limit=15), dict(k=10, beam=30, degree_limit=5), dict(k=1, beam=200)]


def parse_full(buf, nq):
    a = np.frombuffer(buf, dtype=np.uint32)
    out, pos = [], 0
    for _ in range(nq):
        nf, nv, cm = int(a[pos]), int(a[pos + 1]), int(a[pos + 2])
        pos += 3
        fr = a[pos:pos + 2 * nf].reshape(nf, 2); pos += 2 * nf
        vi = a[pos:pos + 2 * nv].reshape(nv, 2); pos += 2 * nv
        out.append((fr.copy(), vi.copy(), cm))
    assert pos == a.size
    return out


def main():
    assert O.ref_driver_path(), "build oracle/_ref first (make -C oracle ref)"
    tmp = tempfile.mkdtemp(prefix="pann_golden_")
    only = set(sys.argv[1:])          # `make_golden.py name ...` regenerates just those fixtures
    for name, dt, metric, n, d, nq, R, L, passes in CASES:
        if only and name not in only:
            continue
        base = O.gen_points(dt, n, d, seed=77)
        qs = O.gen_points(dt, nq, d, seed=77, stream=1)
        bp, qp = os.path.join(tmp, f"{name}.base.{EXT[dt]}"), os.path.join(tmp, f"{name}.query.{EXT[dt]}")
        fileio.write_bin(bp, base); fileio.write_bin(qp, qs)
        common = ["-data_type", dt, "-dist_func", metric, "-base", bp]
        out = os.path.join(tmp, name)
        g = {}
        # build, unsorted rows (byte-exact target) and sorted rows; 1 thread and all threads must agree
        O.ref_run(["build"] + common + ["-R", R, "-L", L, "-alpha", 1.2, "-passes", passes, "-sort", 0, "-out", out + ".ns"])
        O.ref_run(["build"] + common + ["-R", R, "-L", L, "-alpha", 1.2, "-passes", passes, "-sort", 0, "-out", out + ".ns1"], threads=1)
        assert open(out + ".ns.graph", "rb").read() == open(out + ".ns1.graph", "rb").read(), "reference build is not deterministic?"
        O.ref_run(["build"] + common + ["-R", R, "-L", L, "-alpha", 1.2, "-passes", passes, "-sort", 1, "-out", out + ".s"])
        g["graph_unsorted"] = fileio.read_graph(out + ".ns.graph")
        g["graph_sorted"] = fileio.read_graph(out + ".s.graph")
        g["build_visited"] = np.fromfile(out + ".ns.bvisited", dtype=np.uint32)
        g["build_distances"] = np.fromfile(out + ".ns.bdistances", dtype=np.uint32)
        # searches on the reference-built (unsorted) graph
        for si, s in enumerate(SEARCHES):
            if s["k"] > s["beam"]:
                continue
            args = ["search"] + common + ["-graph", out + ".ns.graph", "-query", qp, "-k", s["k"], "-beam", s["beam"],
                                          "-cut", s.get("cut", 1.35), "-limit", s.get("limit", -1),
                                          "-degree_limit", s.get("degree_limit", -1), "-out", out + f".s{si}"]
            try:
                O.ref_run(args)
            except RuntimeError:
                continue  # the reference aborts when a query ends with fewer than k results
            g[f"s{si}_ids"] = np.fromfile(out + f".s{si}.ids", dtype=np.uint32).reshape(nq, s["k"])
            g[f"s{si}_visited"] = np.fromfile(out + f".s{si}.visited", dtype=np.uint32)
            g[f"s{si}_distances"] = np.fromfile(out + f".s{si}.distances", dtype=np.uint32)
        # full frontier + visited lists, query mode and build mode (self = base row)
        O.ref_run(["full"] + common + ["-graph", out + ".ns.graph", "-query", qp, "-k", 0, "-beam", L, "-cut", 0.0, "-out", out + ".fq"])
        full = parse_full(open(out + ".fq.full", "rb").read(), nq)
        g["full_counts"] = np.array([[len(f), len(v), c] for f, v, c in full], dtype=np.uint32)
        g["full_frontier"] = np.concatenate([f for f, _, _ in full])
        g["full_visited"] = np.concatenate([v for _, v, _ in full])
        # build-mode: the "queries" file is the first nq base rows, -self 1 makes row i alias base row i
        sp = os.path.join(tmp, f"{name}.self.{EXT[dt]}")
        fileio.write_bin(sp, base[:nq])
        O.ref_run(["full"] + common + ["-graph", out + ".ns.graph", "-query", sp, "-k", 0, "-beam", L, "-cut", 0.0, "-self", 1, "-out", out + ".fs"])
        full = parse_full(open(out + ".fs.full", "rb").read(), nq)
        g["self_counts"] = np.array([[len(f), len(v), c] for f, v, c in full], dtype=np.uint32)
        g["self_visited"] = np.concatenate([v for _, v, _ in full])
        # ground truth
        k = 20
        O.ref_run(["gt"] + common + ["-query", qp, "-k", k, "-out", out + ".gt"])
        gi, gd = fileio.read_ibin(out + ".gt.ibin")
        g["gt_ids"], g["gt_dists"] = gi, gd
        g["params"] = np.array([n, d, nq, R, L, passes, 77, k], dtype=np.int64)
        np.savez_compressed(os.path.join(HERE, f"{name}.npz"), **g)
        print(name, "ok", {k_: v.shape for k_, v in g.items() if k_.startswith("graph")})
    if only:
        return
    # permutations and distance known-answers
    perms = {}
    for n in (1, 2, 7, 100, 8191, 8192, 20000):
        O.ref_run(["perm", "-n", n, "-out", os.path.join(tmp, f"perm{n}.bin")])
        perms[f"perm_{n}"] = np.fromfile(os.path.join(tmp, f"perm{n}.bin"), dtype=np.int32)
    np.savez_compressed(os.path.join(HERE, "perms.npz"), **perms)
    print("perms ok")


if __name__ == "__main__":
    main()
