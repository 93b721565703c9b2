#!/usr/bin/env python
"""Generates tests/golden/full_size.json: content checksums of the products bench.py times, at
BASELINE.json's full sizes, computed by the CPU oracle (oracle/liborc.so, pinned to the
reference's own code by tests/test_oracle_vs_ref.py) -- never by the CUDA path.

    python tests/golden/make_full_size.py [cfg2] [chain] [rmat]      (default: all three)

  cfg2   configs[1]: A*B, n=65536, k=8, 1 % density, seeds 1000 / 1001 (bench.py's rank-0 inputs)
  chain  configs[2]: bench.py's banded 64-chain (n=16384, k=4), reduced with the rank trees of
         P = 1, 2, 4, 8 (orc.chain(ranks=P) = cu:438-456 + cu:569-571), zero tiles kept
  rmat   configs[3]: R-MAT scale 15, k=8, seeds 11 / 12

The checksum is the one csrc/bsp_checksum.cu defines (tests/util.py restates it in numpy); here
it is evaluated in slices so that the 19 GB product of cfg2 is never copied.  Needs ~45 GB of
host memory and a few minutes on 8 cores for cfg2.
"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "sparse-matrix-multiplication-using-openmp-mpi-and-cuda_b200"))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
sys.path.insert(0, os.path.join(ROOT, "tests"))

import bsp_gen  # noqa: E402
import orc  # noqa: E402
import util  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden", "full_size.json")
_M64 = (1 << 64) - 1


def checksum_omat(m, chunk=1 << 18):
    """Content checksum of an oracle matrix without copying its values."""
    p = m.ptr.contents
    n, k = p.nnzb, p.k
    header = util._splitmix64(p.rows & 0xFFFFFFFF)
    for x in (p.cols & 0xFFFFFFFF, k, 8, n):
        header = util._splitmix64(header ^ x)
    body = 0
    if n:
        W = k * k * 2
        br = np.ctypeslib.as_array(p.brow, shape=(n,))
        bc = np.ctypeslib.as_array(p.bcol, shape=(n,))
        v = np.ctypeslib.as_array(p.vals, shape=(n * k * k,))
        i = np.arange(W, dtype=np.uint32)
        with np.errstate(over="ignore"):
            for t0 in range(0, n, chunk):
                t1 = min(n, t0 + chunk)
                words = v[t0 * k * k:t1 * k * k].view(np.uint32).reshape(t1 - t0, W)
                s1 = util._fmix32(br[t0:t1].astype(np.uint32) * np.uint32(0x9E3779B1)
                                  + bc[t0:t1].astype(np.uint32) * np.uint32(0x85EBCA77) + np.uint32(0x27D4EB2F))
                s2 = util._fmix32(s1 ^ np.uint32(0x165667B1))
                mm = (s1[:, None] + i[None, :] * np.uint32(0x9E3779B1)) | np.uint32(1)
                tile = (s2.astype(np.uint64) << np.uint64(32)) | s1.astype(np.uint64)
                tile = tile + (words.astype(np.uint64) * mm.astype(np.uint64)).sum(axis=1, dtype=np.uint64)
                body = (body + int(tile.sum(dtype=np.uint64))) & _M64
    return util._splitmix64(body ^ header), body


def cfg2():
    n, k, d = 65536, 8, 0.01
    t0 = time.time()
    A = orc.OMat.from_coo(*bsp_gen.random_matrix(n, n, k, d, 1000, "u64"))
    B = orc.OMat.from_coo(*bsp_gen.random_matrix(n, n, k, d, 1001, "u64"))
    C, nprod = orc.multiply(A, B, threads=os.cpu_count() or 1)
    cs, body = checksum_omat(C)
    return {"workload": "A*B, n=%d, k=%d, %g%% density, u64, seeds 1000/1001" % (n, k, d * 100),
            "nnzb_a": A.nnzb, "nnzb_b": B.nnzb, "nnzb_c": C.nnzb, "nprod": nprod, "rows": C.rows, "cols": C.cols, "k": k,
            "checksum": "%016x" % cs, "body": "%016x" % body, "oracle_seconds": round(time.time() - t0, 1)}


def chain():
    sys.path.insert(0, ROOT)
    import bench
    t0 = time.time()
    mats = [orc.OMat.from_coo(*bench.chain_matrix(i)) for i in range(bench.CHAIN["n_mats"])]
    out = {"workload": "banded chain of bench.py: %r" % (bench.CHAIN,), "ranks": {}}
    for P in (1, 2, 4, 8):
        C, nprod = orc.chain(mats, ranks=P)
        cs, body = checksum_omat(C)
        out["ranks"][str(P)] = {"nnzb": C.nnzb, "nprod_total": nprod, "checksum": "%016x" % cs, "body": "%016x" % body}
        out["rows"], out["cols"], out["k"] = C.rows, C.cols, C.k
    out["oracle_seconds"] = round(time.time() - t0, 1)
    return out


def rmat():
    scale, k = 15, 8
    G = 1 << scale
    t0 = time.time()
    mats = []
    for seed in (11, 12):
        br, bc = bsp_gen.rmat_pattern(scale, 16, seed)
        v = bsp_gen.values(seed, br.size * k * k, "u64").reshape(-1, k, k)
        mats.append(orc.OMat.from_coo(G * k, G * k, k, br, bc, v))
    C, nprod = orc.multiply(mats[0], mats[1], threads=os.cpu_count() or 1)
    cs, body = checksum_omat(C)
    return {"workload": "R-MAT scale %d (edge factor 16, .57/.19/.19/.05), k=%d, u64, seeds 11/12" % (scale, k),
            "nnzb_a": mats[0].nnzb, "nnzb_b": mats[1].nnzb, "nnzb_c": C.nnzb, "nprod": nprod, "rows": C.rows, "cols": C.cols,
            "k": k, "checksum": "%016x" % cs, "body": "%016x" % body, "oracle_seconds": round(time.time() - t0, 1)}


if __name__ == "__main__":
    which = sys.argv[1:] or ["chain", "rmat", "cfg2"]
    data = json.load(open(OUT)) if os.path.exists(OUT) else {}
    for w in which:
        data[w] = {"cfg2": cfg2, "chain": chain, "rmat": rmat}[w]()
        print(w, json.dumps(data[w])[:400], flush=True)
        with open(OUT, "w") as f:
            json.dump(data, f, indent=1, sort_keys=True)
            f.write("\n")
