#!/usr/bin/env python
"""Times the float32 tile product (BASELINE configs[4]: n=65536, 2 % density, k given) on both numeric paths.
usage: python tools/f32_variants.py K [reps]"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "sparse-matrix-multiplication-using-openmp-mpi-and-cuda_b200"))
import bsp_gen  # noqa: E402
import bspgemm  # noqa: E402

k = int(sys.argv[1]) if len(sys.argv) > 1 else 64
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
n, d = 65536, 0.02
G = n // k
with bspgemm.Context(0) as ctx:
    mats = []
    for seed in (21, 22):
        br, bc = bsp_gen.bernoulli_pattern(G, G, d, seed)
        u = bsp_gen.splitmix64(seed ^ 0xABCDEF, br.size * k * k).astype(np.float64) / 2.0 ** 64
        mats.append(ctx.matrix_from_coo_f32(n, n, k, br, bc, (2.0 * u - 1.0).astype(np.float32).reshape(-1, k, k)))
    for path in ("tensor", "simt"):
        ctx.set_f32_path(path)
        st = bspgemm.Stats()
        ms = []
        for i in range(reps):
            ctx.helper(mats[0], mats[1], stats=st).free()
            if i >= 1:
                ms.append(st.ms_numeric)
        ms.sort()
        print(json.dumps({"k": k, "path": path, "numeric_ms": round(ms[len(ms) // 2], 3), "nprod": st.nprod, "nnzb_c": st.nnzb_c,
                          "tflops": round(2.0 * k ** 3 * st.nprod / (ms[len(ms) // 2] * 1e-3) / 1e12, 1)}), flush=True)
