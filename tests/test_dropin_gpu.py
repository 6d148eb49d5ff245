"""The C++ drop-in header (include/parlayann_b200/dropin.h) compiled INSIDE the unmodified reference
(oracle/_ref/dropin_check, built where /root/reference exists): in one process the reference's knn_index::build_index,
qsearchAll and compute_groundtruth run on the CPU and their b200:: twins on the GPU, and the results must be identical
(graph rows, build statistics, ids, per-query counters, ground-truth distance rows)."""
import os
import subprocess

import pytest

import oracle_lib as O
from parlayann_b200 import fileio

pytestmark = pytest.mark.gpu
BIN = os.path.join(O.ORACLE_DIR, "_ref", "dropin_check")
EXT = {"uint8": "u8bin", "int8": "i8bin", "float": "fbin"}


@pytest.mark.parametrize("data_type,metric,n,d,R,L,passes,mode", [
    ("uint8", "Euclidian", 6000, 128, 32, 64, 1, ""),
    ("float", "Euclidian", 4000, 96, 24, 48, 1, ""),
    ("int8", "mips", 3000, 64, 16, 32, 2, ""),
    ("float", "mips", 4000, 100, 24, 48, 1, "qmips"),   # normalised rows + Quantized_Mips_Point<8,true>: build and search + re-rank
    ("float", "mips", 3000, 64, 16, 32, 2, "qmips"),
])
def test_dropin_matches_reference_in_process(tmp_path, data_type, metric, n, d, R, L, passes, mode):
    if not os.path.exists(BIN):
        pytest.skip("oracle/_ref/dropin_check was not built (needs /root/reference at build time)")
    base = O.gen_points(data_type, n, d, seed=21)
    qs = O.gen_points(data_type, 200, d, seed=21, stream=1)
    bp, qp = str(tmp_path / f"b.{EXT[data_type]}"), str(tmp_path / f"q.{EXT[data_type]}")
    fileio.write_bin(bp, base); fileio.write_bin(qp, qs)
    r = subprocess.run([BIN, "-base", bp, "-query", qp, "-data_type", data_type, "-dist_func", metric, "-R", str(R), "-L", str(L),
                        "-passes", str(passes), "-mode", mode], capture_output=True, text=True, timeout=600)
    lines = [l for l in r.stdout.splitlines() if l.startswith("DROPIN")]
    assert r.returncode == 0 and lines and lines[-1] == "DROPIN_OK", "\n".join(lines) + r.stderr[-2000:]
