#!/usr/bin/env python
"""Would walking C in column panels (B panel resident in L2) help the float32 tensor-core kernel?  Emulated without
kernel changes: B is cut into P column panels, each product A * B_p runs on its own, numeric times are summed.
usage: python tools/f32_colpanel_probe.py K [P ...]"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "sparse-matrix-multiplication-using-openmp-mpi-and-cuda_b200"))
import bsp_gen  # noqa: E402
import bspgemm  # noqa: E402

k = int(sys.argv[1]) if len(sys.argv) > 1 else 64
panels = [int(x) for x in sys.argv[2:]] or [1, 2, 4, 8]
n, d = 65536, 0.02
G = n // k
with bspgemm.Context(0) as ctx:
    ctx.set_f32_path("tensor")
    coo = []
    for seed in (21, 22):
        br, bc = bsp_gen.bernoulli_pattern(G, G, d, seed)
        u = bsp_gen.splitmix64(seed ^ 0xABCDEF, br.size * k * k).astype(np.float64) / 2.0 ** 64
        coo.append((br, bc, (2.0 * u - 1.0).astype(np.float32).reshape(-1, k, k)))
    A = ctx.matrix_from_coo_f32(n, n, k, *coo[0])
    for P in panels:
        W = (G + P - 1) // P
        Bs = []
        for p in range(P):
            keep = (coo[1][1] >= p * W) & (coo[1][1] < (p + 1) * W)
            Bs.append(ctx.matrix_from_coo_f32(n, n, k, coo[1][0][keep], coo[1][1][keep], coo[1][2][keep]))
        best = None
        for rep in range(4):
            tot, nprod = 0.0, 0
            for Bp in Bs:
                st = bspgemm.Stats()
                ctx.helper(A, Bp, stats=st).free()
                tot += st.ms_numeric
                nprod += st.nprod
            if rep >= 1:
                best = tot if best is None else min(best, tot)
        print(json.dumps({"k": k, "column_panels": P, "numeric_ms_sum": round(best, 3), "nprod": nprod}), flush=True)
        for Bp in Bs:
            Bp.free()
