#!/usr/bin/env python
"""HBM-bound phases around the product (north_star item 3, "compaction and output on device") at
BASELINE.json configs[1] size: zero-tile prune (flag pass alone, and flag + compaction with half the
tiles zero), panel concatenation, content checksum.  Wall clock around synchronous C-ABI calls, best
of 3; bytes = algorithmic (what the phase must read and write).  Prints one JSON line."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "sparse-matrix-multiplication-using-openmp-mpi-and-cuda_b200")]


def best(fn, reps=3):
    ts = []
    for _ in range(reps):
        t0 = time.perf_counter()
        fn()
        ts.append(time.perf_counter() - t0)
    return min(ts)


def main():
    import torch
    import bsp_gen
    import bspgemm
    n, k, d = 65536, 8, 0.01
    G = n // k
    ctx = bspgemm.Context(0)
    A = ctx.matrix_from_coo(*bsp_gen.random_matrix(n, n, k, d, 1000, "u64"))
    B = ctx.matrix_from_coo(*bsp_gen.random_matrix(n, n, k, d, 1001, "u64"))
    C = ctx.helper(A, B)
    nnzb = C.nnzb
    tile = k * k * 8
    out = {"nnzb_c": nnzb, "peak_GBps": json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))).get("hbm_gbs")
           if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else None}

    def rate(nbytes, sec):
        return {"ms": sec * 1e3, "GBps": nbytes / sec / 1e9}
    C.checksum()
    out["checksum"] = rate(nnzb * (tile + 8), best(C.checksum))
    # prune, nothing to drop: one read of every tile
    C.prune_zero()
    out["prune_flags_only"] = rate(nnzb * tile + 8 * nnzb, best(C.prune_zero))
    # concat of 8 panels: every byte read once and written once
    bounds = [G * i // 8 for i in range(9)]
    parts = [ctx.helper(A, B, panel=(bounds[i], bounds[i + 1])) for i in range(8)]
    C.free()

    def cat():
        ctx.concat_panels(parts).free()
        ctx.sync()
    cat()
    out["concat_8_panels"] = rate(2 * nnzb * (tile + 4), best(cat))
    for p in parts:
        p.free()
    # prune with every other tile zero: read all, write half
    C = ctx.helper(A, B)
    nbr, nbc, _, _, _ = C.device_view()
    rowptr = torch.empty(nbr + 1, dtype=torch.int32, device="cuda:0")
    colidx = torch.empty(nnzb, dtype=torch.int32, device="cuda:0")
    vals = torch.empty(nnzb * k * k, dtype=torch.int64, device="cuda:0")
    C.export_device(rowptr.data_ptr(), colidx.data_ptr(), vals.data_ptr())
    C.free()
    vals.view(-1, k * k)[::2] = 0
    torch.cuda.synchronize()
    ts = []
    for _ in range(3):
        Z = ctx.matrix_from_device_bcsr(n, n, k, nbr, nbc, nnzb, rowptr.data_ptr(), colidx.data_ptr(), vals.data_ptr())
        ctx.sync()
        t0 = time.perf_counter()
        Z.prune_zero()
        ctx.sync()  # the compaction kernel is launched after the call's only host wait
        ts.append(time.perf_counter() - t0)
        kept = Z.nnzb
        Z.free()
    out["prune_half_zero"] = dict(rate(nnzb * tile + kept * tile * 2 + 12 * nnzb, min(ts)), kept=kept)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
