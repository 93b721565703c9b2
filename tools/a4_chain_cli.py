#!/usr/bin/env python
"""End-to-end time of the drop-in CLI on bench.py's chain workload (64 banded matrices, n=16384, k=4):
`a4 <folder>` reading the reference's text files and writing ./matrix, and `a4 --binary --checksum`
on the same matrices as BSPMAT01 files.  Wall clock of the whole process (CUDA context creation
included) next to a4's own `time taken` line.  Extra arguments (e.g. --ranks 2 --devices 2) are passed
to a4.  Prints one JSON line."""
import json
import os
import re
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "sparse-matrix-multiplication-using-openmp-mpi-and-cuda_b200")
sys.path[:0] = [ROOT, PKG]


def main():
    import bench
    import bspgemm
    ctx = bspgemm.Context(0)
    n, k = bench.CHAIN["n_mats"], bench.CHAIN["k"]
    exe = os.path.join(PKG, "a4")
    out = {}
    with tempfile.TemporaryDirectory(prefix="bsp_a4_") as dd:
        tdir, bdir = os.path.join(dd, "text"), os.path.join(dd, "bin")
        os.mkdir(tdir); os.mkdir(bdir)
        for i in range(n):
            m = ctx.matrix_from_coo(*bench.chain_matrix(i))
            m.write_text(os.path.join(tdir, "matrix%d" % (i + 1)))
            m.write_bin(os.path.join(bdir, "matrix%d.bin" % (i + 1)))
            m.free()
        for d in (tdir, bdir):
            with open(os.path.join(d, "size"), "w") as f:
                f.write("%d %d\n" % (n, k))
        ctx.close()
        for name, folder, flags in (("text", tdir, []), ("binary", bdir, ["--binary", "--checksum"])):
            best = None
            for _ in range(2):
                t0 = time.perf_counter()
                p = subprocess.run([exe, folder] + flags + sys.argv[1:], cwd=dd, capture_output=True, text=True)
                wall = time.perf_counter() - t0
                assert p.returncode == 0, p.stderr
                own = float(re.search(r"time taken ([0-9.eE+-]+) seconds", p.stdout).group(1))
                if best is None or wall < best["wall_s"]:
                    best = {"wall_s": wall, "a4_time_taken_s": own,
                            "input_MB": sum(os.path.getsize(os.path.join(folder, f)) for f in os.listdir(folder)) / 1e6,
                            "output_MB": os.path.getsize(os.path.join(dd, "matrix" if not flags else "matrix.bin")) / 1e6}
                    m = re.search(r"checksum ([0-9a-f]{16})", p.stdout)
                    if m:
                        best["checksum"] = m.group(1)
            out[name] = best
    out["a4_args"] = sys.argv[1:]
    print(json.dumps(out))


if __name__ == "__main__":
    main()
