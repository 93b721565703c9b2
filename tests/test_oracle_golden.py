"""The CPU oracle (oracle/bsp_oracle.c) against the golden fixtures produced by the
reference's own code (tests/golden/make_golden.py).  CPU only."""
import hashlib
import os
import tempfile

import numpy as np
import pytest

import orc
import util

KATS = util.load_kats()


def run_oracle(case):
    k = case["k"]
    mats = [orc.OMat.from_coo(r, c, k, br, bc, v) for r, c, br, bc, v in util.case_mats(case)]
    res = orc.chain(mats, ranks=case["ranks"])[0]
    res.prune_zero()
    return res


@pytest.mark.parametrize("case", KATS, ids=[c["id"] for c in KATS])
def test_oracle_matches_reference_output(case):
    res = run_oracle(case)
    with tempfile.TemporaryDirectory() as d:
        p = os.path.join(d, "matrix")
        res.write(p)
        assert open(p).read() == case["matrix_text"]
    br, bc, v = res.coo()
    assert util.matrix_text(res.rows, res.cols, case["k"], br, bc, v) == case["matrix_text"]


@pytest.mark.parametrize("case", [c for c in KATS if c["multiplying"] is not None], ids=lambda c: c["id"])
def test_oracle_tree_log(case):
    k = case["k"]
    mats = [orc.OMat.from_coo(r, c, k, br, bc, v) for r, c, br, bc, v in util.case_mats(case)]
    _, _, log = orc.chain(mats, want_log=True)
    got = [list(map(int, ln.split()[1:])) for ln in log.splitlines()]
    assert got == case["multiplying"]


@pytest.mark.parametrize("cfg", util.load_cfg1(), ids=lambda c: c["id"])
def test_oracle_cfg1_hash(cfg):
    import bsp_gen
    mats = []
    for s in cfg["seeds"]:
        rows, cols, k, br, bc, v = bsp_gen.random_matrix(cfg["n"], cfg["n"], cfg["k"], cfg["density"], s, cfg["kind"])
        mats.append(orc.OMat.from_coo(rows, cols, k, br, bc, v))
    assert [m.nnzb for m in mats] == cfg["nnzb_in"]
    res = orc.chain(mats)[0].prune_zero()
    assert res.nnzb == cfg["out_nnzb"]
    with tempfile.TemporaryDirectory() as d:
        p = os.path.join(d, "matrix")
        res.write(p)
        assert hashlib.sha256(open(p, "rb").read()).hexdigest() == cfg["sha256"]


def test_scalar_edges():
    L = orc.lib()
    MAX = (1 << 64) - 1
    assert L.orc_mulmod(MAX, 1) == 0                       # product equal to 2^64-1 becomes 0
    assert L.orc_mulmod(1 << 32, 1 << 32) == 0             # wraps to 0 (not 1, which true mod 2^64-1 would give)
    assert L.orc_mulmod((1 << 32) + 1, (1 << 32) - 1) == 0  # 2^64-1 exactly
    assert L.orc_addmod(MAX - 1, 2) == 0                   # wraps
    assert L.orc_addmod(MAX - 1, 1) == 0                   # hits 2^64-1 -> 0
    assert L.orc_addmod(MAX, 0) == 0
    assert L.orc_addmod(5, 6) == 11


def test_multithreaded_fold_is_identical():
    import bsp_gen
    a = orc.OMat.from_coo(*bsp_gen.random_matrix(256, 256, 4, 0.2, 11, "edgy"))
    b = orc.OMat.from_coo(*bsp_gen.random_matrix(256, 256, 4, 0.2, 12, "edgy"))
    c1, n1 = orc.multiply(a, b, threads=1)
    c4, n4 = orc.multiply(a, b, threads=4)
    assert n1 == n4
    for x, y in zip(c1.coo(), c4.coo()):
        assert np.array_equal(x, y)
