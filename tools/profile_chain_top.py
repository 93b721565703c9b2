#!/usr/bin/env python
"""Builds the two halves of bench.py's chain workload and multiplies them a few times:
the top-of-tree product that dominates the chain (for ncu: numeric_rows launches 63..)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "sparse-matrix-multiplication-using-openmp-mpi-and-cuda_b200"))
import bench  # noqa: E402
import bspgemm  # noqa: E402

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 2
with bspgemm.Context(0) as ctx:
    mats = [ctx.matrix_from_coo(*bench.chain_matrix(i)) for i in range(64)]
    left, right = ctx.helper2(mats[:32]), ctx.helper2(mats[32:])
    for _ in range(reps):
        st = bspgemm.Stats()
        c = ctx.helper(left, right, stats=st)
        print({k: v for k, v in st.as_dict().items()})
        c.free()
