#!/usr/bin/env python
"""Times the numeric pass of configs[1] (n=65536, k=8, 1 %) under the kernel variants selected by
BSP_NUM_KERNEL / BSP_NUM_PANEL_MB / BSP_NUM_Q, checks every result against the oracle's checksum, and
prints one JSON line per variant.  usage: python tools/numeric_variants.py [spec ...]
where spec = comma-separated NAME=VALUE environment settings, e.g. BSP_NUM_KERNEL=2,BSP_NUM_PANEL_MB=64"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "sparse-matrix-multiplication-using-openmp-mpi-and-cuda_b200"))
import bsp_gen  # noqa: E402
import bspgemm  # noqa: E402

KEYS = ("BSP_NUM_KERNEL", "BSP_NUM_PANEL_MB", "BSP_NUM_Q", "BSP_NUM_WARPS", "BSP_NUM_MINB", "BSP_NUM_MACV", "BSP_NUM_SORT", "BSP_HV_MIN_P", "BSP_SYM_NO_DENSE", "BSP_SYM_DENSE_RATIO", "BSP_SYM_BIG")


def main():
    args = sys.argv[1:]
    workload = "cfg2"
    if args and args[0] in ("cfg2", "chaintop", "k16", "k32", "rmat"):
        workload, args = args[0], args[1:]
    specs = args or ["", "BSP_NUM_MINB=4"]
    golden = json.load(open(os.path.join(ROOT, "tests", "golden", "full_size.json")))
    with bspgemm.Context(0) as ctx:
        if workload == "cfg2":
            n, k, d = 65536, 8, 0.01
            want = golden["cfg2"]["checksum"]
            A = ctx.matrix_from_coo(*bsp_gen.random_matrix(n, n, k, d, 1000, "u64"))
            B = ctx.matrix_from_coo(*bsp_gen.random_matrix(n, n, k, d, 1001, "u64"))
        elif workload == "rmat":  # BASELINE configs[3]: R-MAT scale 15, k=8
            scale, k = 15, 8
            G = 1 << scale
            want = golden["rmat"]["checksum"]
            ms_ = []
            for seed in (11, 12):
                br, bc = bsp_gen.rmat_pattern(scale, 16, seed)
                ms_.append(ctx.matrix_from_coo(G * k, G * k, k, br, bc, bsp_gen.values(seed, br.size * k * k, "u64").reshape(-1, k, k)))
            A, B = ms_
        elif workload in ("k16", "k32"):  # bench.py's large-tile u64 workloads; no golden: the variants must agree with each other
            sys.path.insert(0, ROOT)
            import bench
            n, k, d, _ = bench.WORKLOADS[workload]
            want = None
            A = ctx.matrix_from_coo(*bsp_gen.random_matrix(n, n, k, d, 1000, "u64"))
            B = ctx.matrix_from_coo(*bsp_gen.random_matrix(n, n, k, d, 1001, "u64"))
        else:  # the top product of bench.py's chain (k=4, 1.76e8 tile products, 105 pairs per output tile)
            sys.path.insert(0, ROOT)
            import bench
            k = bench.CHAIN["k"]
            want = golden["chain"]["ranks"]["1"]["checksum"]
            mats = [ctx.matrix_from_coo(*bench.chain_matrix(i)) for i in range(64)]
            A, B = ctx.helper2(mats[:32]), ctx.helper2(mats[32:])
            del mats
        results = {spec: [] for spec in specs}
        sym = {}
        info = {}
        for rnd in range(3):  # interleaved rounds: clock / power drift hits every variant alike
            for spec in specs:
                for key in KEYS:
                    os.environ.pop(key, None)
                for kv in filter(None, spec.split(",")):
                    name, val = kv.split("=")
                    os.environ[name] = val
                st = bspgemm.Stats()
                for i in range(5):
                    C = ctx.helper(A, B, stats=st)
                    if i >= 1:
                        results[spec].append(st.ms_numeric)
                        sym.setdefault(spec, []).append(st.ms_symbolic)
                    if i < 4:
                        C.free()
                if rnd == 0:
                    cs = "%016x" % C.checksum()
                    if want is None:
                        want = cs
                    info[spec] = (cs == want, st.nprod, st.ms_symbolic)
                C.free()
        for spec in specs:
            ms = sorted(results[spec])
            ok, nprod, _ = info[spec]
            print(json.dumps({"workload": workload, "variant": spec or "default", "numeric_ms_median": round(ms[len(ms) // 2], 3),
                              "numeric_ms_min": round(ms[0], 3), "numeric_ms_max": round(ms[-1], 3),
                              "u64_mac_per_s": round(k ** 3 * nprod / (ms[len(ms) // 2] * 1e-3) / 1e12, 3), "symbolic_ms": round(sorted(sym[spec])[len(sym[spec]) // 2], 3),
                              "checksum_ok": ok}), flush=True)


if __name__ == "__main__":
    main()
