#!/usr/bin/env python
"""Where the time between two products goes on the bench workload: wall clock of helper() and of
free() against the device time the library reports (ms_total)."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "sparse-matrix-multiplication-using-openmp-mpi-and-cuda_b200")]
import bsp_gen  # noqa: E402
import bspgemm  # noqa: E402

n, k, d = 65536, 8, 0.01
ctx = bspgemm.Context(0)
A = ctx.matrix_from_coo(*bsp_gen.random_matrix(n, n, k, d, 1000, "u64"))
B = ctx.matrix_from_coo(*bsp_gen.random_matrix(n, n, k, d, 1001, "u64"))
st = bspgemm.Stats()
for i in range(8):
    t0 = time.perf_counter()
    C = ctx.helper(A, B, stats=st)
    t1 = time.perf_counter()
    C.free()
    t2 = time.perf_counter()
    print("helper wall %.3f ms (device %.3f: sym %.3f num %.3f)  free %.3f ms" %
          ((t1 - t0) * 1e3, st.ms_total, st.ms_symbolic, st.ms_numeric, (t2 - t1) * 1e3))
