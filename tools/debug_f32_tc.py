import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "sparse-matrix-multiplication-using-openmp-mpi-and-cuda_b200"))
import bspgemm
k = int(sys.argv[1]) if len(sys.argv) > 1 else 32
np.set_printoptions(linewidth=250, precision=0, suppress=True)
with bspgemm.Context(0) as ctx:
    ctx.set_f32_path("tensor")
    z = np.zeros(1, np.int32)
    I = np.eye(k, dtype=np.float32)
    R = (np.arange(k)[:, None] * 100 + np.arange(k)[None, :]).astype(np.float32)  # R[m][n] = 100 m + n
    for name, a, b in (("I*R", I, R), ("R*I", R, I)):
        A = ctx.matrix_from_coo_f32(k, k, k, z, z, a.reshape(1, k, k))
        B = ctx.matrix_from_coo_f32(k, k, k, z, z, b.reshape(1, k, k))
        C = ctx.helper(A, B).to_coo()[2][0]
        print(name, "max abs err", np.abs(C - R).max())
        print(C[:4, :12]); print(C[16:19, :12])
