"""parlayann_b200/csrc/tie_sort.h — the per-row std::sort emulation the final neighbour sort runs on the device for rows
with tied distances (vamana/index.h:180-185 sorts with a comparator that sees the distance only; where equal keys end up is
a property of libstdc++'s algorithm).  The same header is compiled here for the host and compared with std::sort itself."""
import os
import subprocess
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_tie_sort_equals_std_sort():
    with tempfile.TemporaryDirectory() as tmp:
        exe = os.path.join(tmp, "tie_sort_check")
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-I", os.path.join(ROOT, "parlayann_b200", "csrc"),
                               "-o", exe, os.path.join(ROOT, "tests", "tie_sort_check.cpp")])
        out = subprocess.run([exe, "400000"], capture_output=True, text=True)
        assert out.returncode == 0, out.stdout + out.stderr
        assert "differ 0 heap_differ 0 guard 0" in out.stdout, out.stdout
