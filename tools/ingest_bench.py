#!/usr/bin/env python
"""Ingestion either side of the path: the 64 chain matrices of
bench.py's chain workload and the two `mid` matrices, as text files and as binary files.
Times (wall clock, best of 3): one-by-one text reads, the pipelined multi-file reader, binary
writes and reads.  Prints one JSON line.  usage: python tools/ingest_bench.py"""
import json
import os
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "sparse-matrix-multiplication-using-openmp-mpi-and-cuda_b200")
sys.path[:0] = [ROOT, PKG]


def best(fn, reps=3):
    ts = []
    for _ in range(reps):
        t0 = time.perf_counter()
        fn()
        ts.append(time.perf_counter() - t0)
    return min(ts)


def main():
    import bench
    import bsp_gen
    import bspgemm
    ctx = bspgemm.Context(0)
    out = {"host_threads": os.cpu_count()}
    with tempfile.TemporaryDirectory(prefix="bsp_ingest_") as dd:
        for name, mats, k in (("chain64", [bench.chain_matrix(i) for i in range(bench.CHAIN["n_mats"])], bench.CHAIN["k"]),
                              ("mid2", [bsp_gen.random_matrix(16384, 16384, 8, 0.01, 1000 + i, "u64") for i in range(2)], 8)):
            dev = [ctx.matrix_from_coo(*m) for m in mats]
            tp = [os.path.join(dd, "%s_%d" % (name, i)) for i in range(len(mats))]
            bp = [p + ".bin" for p in tp]
            t_wt = best(lambda: [m.write_text(p) for m, p in zip(dev, tp)], 1)
            t_wb = best(lambda: [m.write_bin(p) for m, p in zip(dev, bp)])
            text_bytes = sum(os.path.getsize(p) for p in tp)
            bin_bytes = sum(os.path.getsize(p) for p in bp)

            def one_by_one():
                ms = [ctx.read_text(p, k) for p in tp]
                ctx.sync()
                return ms

            def many():
                ms = ctx.read_text_many(tp, k)
                ctx.sync()
                return ms

            def binary():
                ms = [ctx.read_bin(p) for p in bp]
                ctx.sync()
                return ms
            sums = [m.checksum() for m in dev]
            for fn in (one_by_one, many, binary):
                assert [m.checksum() for m in fn()] == sums
            r = {"files": len(mats), "text_MB": text_bytes / 1e6, "bin_MB": bin_bytes / 1e6,
                 "write_text_s": t_wt, "write_bin_s": t_wb, "read_text_one_by_one_s": best(one_by_one),
                 "read_text_many_s": best(many), "read_bin_s": best(binary)}
            r["read_text_many_MBps"] = r["text_MB"] / r["read_text_many_s"]
            r["read_bin_MBps"] = r["bin_MB"] / r["read_bin_s"]
            out[name] = r
    print(json.dumps(out))


if __name__ == "__main__":
    main()
