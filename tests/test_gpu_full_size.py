"""Full-size parity: the products bench.py times, at BASELINE.json's sizes, against checksums the CPU
oracle computed for the same inputs (tests/golden/full_size.json, generator
tests/golden/make_full_size.py).  Bar: equal (the checksum is a function of every tile's coordinates
and every 32-bit word of its values; tests/util.py restates it, tests/test_gpu_io.py pins the device
implementation against that restatement).  Reference path covered: cu:97-286 (one product),
cu:287-327 + cu:438-456 + cu:569-571 (chain under the P-rank trees)."""
import json
import os

import numpy as np
import pytest

import bsp_dist
import bsp_gen
import bspgemm
import util

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def golden():
    with open(os.path.join(util.GOLDEN, "full_size.json")) as f:
        return json.load(f)


@pytest.fixture()
def ctx():
    with bspgemm.Context(0) as c:
        yield c


def test_golden_file_is_consistent_with_the_numpy_restatement(golden):
    """body -> checksum through the header mix, for every entry (runs without touching the GPU)."""
    for key in ("cfg2", "rmat"):
        g = golden[key]
        assert "%016x" % bspgemm.checksum_finish(g["rows"], g["cols"], g["k"], 8, g["nnzb_c"], int(g["body"], 16)) == g["checksum"]


def test_cfg2_full_product_equals_oracle(ctx, golden):
    """configs[1]: n=65536, k=8, 1 %: all 3.75e7 output tiles, whole and as 4 row panels."""
    g = golden["cfg2"]
    n, k, d = 65536, 8, 0.01
    A = ctx.matrix_from_coo(*bsp_gen.random_matrix(n, n, k, d, 1000, "u64"))
    B = ctx.matrix_from_coo(*bsp_gen.random_matrix(n, n, k, d, 1001, "u64"))
    st = bspgemm.Stats()
    C = ctx.helper(A, B, stats=st)
    assert (st.nprod, C.nnzb) == (g["nprod"], g["nnzb_c"])
    assert "%016x" % C.checksum() == g["checksum"]
    C.free()
    bounds = ctx.panel_bounds(A, B, 4)
    body = nn = 0
    for i in range(4):
        P = ctx.helper(A, B, panel=(bounds[i], bounds[i + 1]))
        body = (body + P.checksum_body()) & 0xFFFFFFFFFFFFFFFF
        nn += P.nnzb
        P.free()
    assert "%016x" % bspgemm.checksum_finish(n, n, k, 8, nn, body) == g["checksum"]


@pytest.mark.parametrize("P", [1, 2, 4, 8])
def test_chain64_equals_oracle_rank_tree(ctx, golden, P):
    """configs[2]: the banded 64-chain of bench.py under the reference's P-rank association tree."""
    import bench
    g = golden["chain"]["ranks"][str(P)]
    mats = [ctx.matrix_from_coo(*bench.chain_matrix(i)) for i in range(bench.CHAIN["n_mats"])]
    res = bsp_dist.chain_ranks_on_one_context(ctx, mats, P)
    assert res.nnzb == g["nnzb"]
    assert "%016x" % res.checksum() == g["checksum"]


def test_rmat_full_product_equals_oracle(ctx, golden):
    """configs[3]: R-MAT scale 15, k=8: whole product and 8 balanced panels."""
    g = golden["rmat"]
    scale, k = 15, 8
    G = 1 << scale
    mats = []
    for seed in (11, 12):
        br, bc = bsp_gen.rmat_pattern(scale, 16, seed)
        mats.append(ctx.matrix_from_coo(G * k, G * k, k, br, bc, bsp_gen.values(seed, br.size * k * k, "u64").reshape(-1, k, k)))
    A, B = mats
    st = bspgemm.Stats()
    C = ctx.helper(A, B, stats=st)
    assert (st.nprod, C.nnzb) == (g["nprod"], g["nnzb_c"])
    assert "%016x" % C.checksum() == g["checksum"]
    C.free()
    bounds = ctx.panel_bounds(A, B, 8)
    body = nn = 0
    for i in range(8):
        P = ctx.helper(A, B, panel=(bounds[i], bounds[i + 1]))
        body = (body + P.checksum_body()) & 0xFFFFFFFFFFFFFFFF
        nn += P.nnzb
        P.free()
    assert "%016x" % bspgemm.checksum_finish(G * k, G * k, k, 8, nn, body) == g["checksum"]
