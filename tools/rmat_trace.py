import os, sys
sys.path.insert(0, '/root/repo/sparse-matrix-multiplication-using-openmp-mpi-and-cuda_b200')
os.environ['BSP_TRACE']='1'
import bsp_gen, bspgemm
scale,k=15,8; G=1<<scale
with bspgemm.Context(0) as ctx:
    ms=[]
    for seed in (11,12):
        br,bc=bsp_gen.rmat_pattern(scale,16,seed)
        ms.append(ctx.matrix_from_coo(G*k,G*k,k,br,bc,bsp_gen.values(seed,br.size*k*k,'u64').reshape(-1,k,k)))
    for i in range(3):
        st=bspgemm.Stats(); ctx.helper(ms[0],ms[1],stats=st).free()
    print(st.as_dict())
