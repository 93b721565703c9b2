#!/usr/bin/env python
"""SURVEY.md 8(d) "Reference GPU build on B200": the reference's translation unit compiled by nvcc
for sm_100a (oracle/_ref/ref_gpu: its own kernel cu:44-66 and its own host packing, nothing of
ours) timed on the same inputs as this library, on the same GPU.  TEST/MEASUREMENT TOOL -- it
executes oracle/ and is not part of the product.

For each case: write the reference's text folder, run ref_gpu (its phase timers inside helper =
"hot path", its `time taken` line = end to end), run this repo's `a4` on the same folder, check
the two `matrix` files are byte-identical, and time bsp_spgemm on device-resident inputs.
Prints one JSON line per case.   usage: python tools/ref_gpu_compare.py [cfg1] [cfg2_sample]"""
import hashlib
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "sparse-matrix-multiplication-using-openmp-mpi-and-cuda_b200")
sys.path[:0] = [ROOT, PKG, os.path.join(ROOT, "oracle")]


def main():
    import numpy as np
    import bench
    import bsp_gen
    import bspgemm
    import orc
    cases = sys.argv[1:] or ["cfg1", "cfg2_sample"]
    ctx = bspgemm.Context(0)
    for case in cases:
        if case == "cfg1":
            n, k, d = 1024, 4, 0.05
            a = bsp_gen.random_matrix(n, n, k, d, 1000, "u64")
            b = bsp_gen.random_matrix(n, n, k, d, 1001, "u64")
            nprod = int(np.bincount(b[3], minlength=n // k)[a[4]].sum())
            desc = "configs[0] in full: n=1024 k=4 5%"
        else:
            n, k, d, _ = bench.WORKLOADS["cfg2"]
            a, b, nprod, desc = bench.reference_sample(n, k, d, 1000, 1001)
            desc = "configs[1] sample: " + desc
        flop = 2.0 * k ** 3 * nprod
        with tempfile.TemporaryDirectory(prefix="bsp_refgpu_") as folder:
            orc.write_folder(folder, [orc.OMat.from_coo(*a), orc.OMat.from_coo(*b)], k)
            best = None
            for _ in range(2):  # first run pays CUDA context creation inside its end-to-end time
                r = orc.run_ref(folder, ranks=1, variant="ref_gpu", keep_output=True)
                if r["returncode"] != 0:
                    raise SystemExit("ref_gpu failed: " + r["stderr"][-400:])
                hot = sum(r["phases"].values())
                if best is None or hot < best[0]:
                    best = (hot, r["time_taken"][0] if r["time_taken"] else None, r["phases"])
            ref_sha = hashlib.sha256(r["matrix"]).hexdigest()
            ref_bytes = len(r["matrix"])
            del r
            with tempfile.TemporaryDirectory(prefix="bsp_a4_") as cwd:
                t0 = time.perf_counter()
                p = subprocess.run([os.path.join(PKG, "a4"), folder], cwd=cwd, stdout=subprocess.PIPE, stderr=subprocess.PIPE)
                a4_wall = time.perf_counter() - t0
                if p.returncode != 0:
                    raise SystemExit("a4 failed: " + p.stderr.decode()[-400:])
                with open(os.path.join(cwd, "matrix"), "rb") as f:
                    our_sha = hashlib.sha256(f.read()).hexdigest()
        A, B = ctx.matrix_from_coo(*a), ctx.matrix_from_coo(*b)
        ms = []
        for i in range(6):
            st = bspgemm.Stats()
            C = ctx.helper(A, B, stats=st)
            C.free()
            if i >= 3:
                ms.append(st.ms_total)
        dev_ms = sum(ms) / len(ms)
        print(json.dumps({
            "case": case, "sample": desc, "nprod": nprod, "useful_flop": flop, "matrix_file_bytes": ref_bytes,
            "byte_identical_matrix_file": ref_sha == our_sha,
            "reference_gpu_build": {"hot_path_s": best[0], "hot_path_gflops": flop / best[0] / 1e9,
                                    "end_to_end_s": best[1], "phases_s": best[2]},
            "this_library": {"spgemm_device_ms": dev_ms, "spgemm_gflops": flop / (dev_ms * 1e-3) / 1e9,
                             "a4_cli_wall_s": a4_wall}}), flush=True)
        A.free(); B.free()


if __name__ == "__main__":
    main()
