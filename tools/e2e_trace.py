#!/usr/bin/env python
"""Host-side timeline of bsp_multiply_host on BASELINE configs[1] (bench.py's e2e leg) next to the raw
PCIe rates of the same box, so that the end-to-end figure can be split into upload, ingest, product
and download.  One GPU.

    BSP_TRACE_HOST=1 python tools/e2e_trace.py [calls]

Prints one JSON line per measurement; the library's own timeline goes to stderr.
"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "sparse-matrix-multiplication-using-openmp-mpi-and-cuda_b200"))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

import bsp_gen  # noqa: E402
import bspgemm  # noqa: E402


def pcie_rates(nbytes):
    """Pinned-buffer copy rates in GB/s: (h2d, d2h, d2h on two streams)."""
    host = torch.empty(nbytes, dtype=torch.uint8, pin_memory=True)
    dev = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    out = {}
    for name, fn in (("h2d", lambda: dev.copy_(host, non_blocking=True)), ("d2h", lambda: host.copy_(dev, non_blocking=True))):
        fn(); torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        out[name] = 3 * nbytes / (time.perf_counter() - t0) / 1e9
    half = nbytes // 2
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(3):
        with torch.cuda.stream(s1):
            host[:half].copy_(dev[:half], non_blocking=True)
        with torch.cuda.stream(s2):
            host[half:].copy_(dev[half:], non_blocking=True)
    torch.cuda.synchronize()
    out["d2h_two_streams"] = 3 * nbytes / (time.perf_counter() - t0) / 1e9
    return out


def main():
    calls = int(sys.argv[1]) if len(sys.argv) > 1 else 3
    n, k, d = 65536, 8, 0.01
    torch.cuda.set_device(0)
    print(json.dumps({"pcie_gbs": pcie_rates(4 << 30)}), flush=True)
    a = bsp_gen.random_matrix(n, n, k, d, 1000, "u64")
    b = bsp_gen.random_matrix(n, n, k, d, 1001, "u64")
    ctx = bspgemm.Context(0)
    pin = lambda shape, dt: torch.empty(shape, dtype=dt, pin_memory=True)  # noqa: E731

    def host_side(m):
        h = [pin(m[3].shape, torch.int32), pin(m[4].shape, torch.int32), pin(m[5].size, torch.int64)]
        h[0].numpy()[:] = m[3]; h[1].numpy()[:] = m[4]; h[2].numpy().view(np.uint64)[:] = m[5].reshape(-1)
        return h, (n, n, h[0].numpy(), h[1].numpy(), h[2].numpy().view(np.uint64))

    ha, av = host_side(a)
    hb, bv = host_side(b)
    st = bspgemm.Stats()
    # size of the result through the device path, then preallocated pinned output
    A = ctx.matrix_from_coo(*a); B = ctx.matrix_from_coo(*b)
    Cm = ctx.helper(A, B, stats=st)
    nc = int(st.nnzb_c)
    Cm.free(); A.free(); B.free()
    ho = [pin((nc,), torch.int32), pin((nc,), torch.int32), pin((nc * k * k,), torch.int64)]
    out = (ho[0].numpy(), ho[1].numpy(), ho[2].numpy().view(np.uint64).reshape(nc, k, k))
    ctx.multiply_host(k, av, bv, out=out)  # warm-up
    for i in range(calls):
        t0 = time.perf_counter()
        ctx.multiply_host(k, av, bv, out=out)
        torch.cuda.synchronize()
        ms = (time.perf_counter() - t0) * 1e3
        d2h = sum(x.numel() * x.element_size() for x in ho)
        print(json.dumps({"call": i, "ms": round(ms, 1), "d2h_gb": round(d2h / 1e9, 2),
                          "d2h_gbs_if_alone": round(d2h / ms / 1e6, 1)}), flush=True)


if __name__ == "__main__":
    main()
