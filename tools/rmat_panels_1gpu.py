#!/usr/bin/env python
"""Balance of the eight R-MAT row panels (BASELINE configs[3]) measured on ONE GPU: the panels bsp_panel_bounds
returns are multiplied one after the other, so max/mean of their times is what eight GPUs would see.
usage: python tools/rmat_panels_1gpu.py [parts=8] [BSP_PANEL_LONG_W values ...]   ("" = the library's own weight)"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "sparse-matrix-multiplication-using-openmp-mpi-and-cuda_b200"))
import bsp_gen  # noqa: E402
import bspgemm  # noqa: E402

scale, k = 15, 8
args = sys.argv[1:]
parts = 8
if args and args[0].startswith("parts="):
    parts = int(args[0][6:]); args = args[1:]
G = 1 << scale
with bspgemm.Context(0) as ctx:
    ms = []
    for seed in (11, 12):
        br, bc = bsp_gen.rmat_pattern(scale, 16, seed)
        ms.append(ctx.matrix_from_coo(G * k, G * k, k, br, bc, bsp_gen.values(seed, br.size * k * k, "u64").reshape(-1, k, k)))
    for w in (args or [""]):
        if w == "":
            os.environ.pop("BSP_PANEL_LONG_W", None)
        else:
            os.environ["BSP_PANEL_LONG_W"] = w
        bounds = ctx.panel_bounds(ms[0], ms[1], parts)
        tot, sym = [], []
        for p in range(parts):
            best = None
            for rep in range(3):
                st = bspgemm.Stats()
                ctx.helper(ms[0], ms[1], stats=st, panel=(bounds[p], bounds[p + 1])).free()
                if best is None or st.ms_total < best[0]:
                    best = (st.ms_total, st.ms_symbolic)
            tot.append(round(best[0], 3)); sym.append(round(best[1], 3))
        print(json.dumps({"parts": parts, "long_row_weight_pct": int(w) if w else "default", "bounds": [int(b) for b in bounds], "panel_ms": tot, "symbolic_ms": sym,
                          "max_over_mean": round(max(tot) / (sum(tot) / parts), 4)}), flush=True)
