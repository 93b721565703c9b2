#!/usr/bin/env python
"""BASELINE.json configs[3]: power-law (R-MAT) block pattern, n=262144, k=8 (G = 32768 = scale 15,
edge factor 16).  One GPU: full product; prints phase times and the row-work imbalance that
row-panel sharding over P GPUs would see with panels balanced by tile products."""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "sparse-matrix-multiplication-using-openmp-mpi-and-cuda_b200"))
import bsp_dist  # noqa: E402
import bsp_gen  # noqa: E402
import bspgemm  # noqa: E402

scale = int(sys.argv[1]) if len(sys.argv) > 1 else 15
k = 8
G = 1 << scale
mats = []
for s in (11, 12):
    br, bc = bsp_gen.rmat_pattern(scale, 16, s)
    v = bsp_gen.values(s, br.size * k * k, "u64").reshape(-1, k, k)
    mats.append((G * k, G * k, k, br, bc, v))
rowlen_b = np.bincount(mats[1][3], minlength=G)
row_work = np.bincount(mats[0][3], weights=rowlen_b[mats[0][4]], minlength=G)
print("nnzb", mats[0][3].size, mats[1][3].size, "nprod", int(row_work.sum()), "max row work", int(row_work.max()),
      "max A row", int(np.bincount(mats[0][3]).max()))
for P in (2, 4, 8):
    b = bsp_dist.balanced_panels(row_work, P)
    loads = [row_work[b[i]:b[i + 1]].sum() for i in range(P)]
    print("P=%d panel work max/mean = %.3f" % (P, max(loads) / (sum(loads) / P)))
with bspgemm.Context(0) as ctx:
    A, B = ctx.matrix_from_coo(*mats[0]), ctx.matrix_from_coo(*mats[1])
    for _ in range(3):
        st = bspgemm.Stats()
        t0 = time.perf_counter()
        C = ctx.helper(A, B, stats=st)
        print(st.as_dict(), "wall %.1f ms" % ((time.perf_counter() - t0) * 1e3))
        C.free()
    # panels of an 8-way sharding, run one after the other on this GPU
    b = bsp_dist.balanced_panels(row_work, 8)
    times = []
    for i in range(8):
        st = bspgemm.Stats()
        ctx.helper(A, B, stats=st, panel=(b[i], b[i + 1])).free()
        times.append(st.ms_total)
    print("8 panels on one GPU, ms each:", [round(x, 2) for x in times], "max/mean %.3f" % (max(times) / (sum(times) / 8)))
