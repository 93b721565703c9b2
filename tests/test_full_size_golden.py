"""CPU checks of tests/golden/full_size.json (the oracle's checksums of the full-size bench workloads, made by
tests/golden/make_full_size.py): the stored tile sum and the stored checksum agree through the numpy restatement of
the header mix, the C library's bsp_checksum_finish gives the same value without a GPU, and the chain entries
cover the rank counts bench.py runs."""
import json
import os

import util


def _load():
    with open(os.path.join(util.GOLDEN, "full_size.json")) as f:
        return json.load(f)


def _finish(rows, cols, k, nnzb, body):
    header = util._splitmix64(rows & 0xFFFFFFFF)
    for x in (cols & 0xFFFFFFFF, k, 8, nnzb):
        header = util._splitmix64(header ^ x)
    return util._splitmix64(body ^ header)


def test_bodies_and_checksums_agree():
    g = _load()
    for key in ("cfg2", "rmat"):
        e = g[key]
        assert "%016x" % _finish(e["rows"], e["cols"], e["k"], e["nnzb_c"], int(e["body"], 16)) == e["checksum"]
    c = g["chain"]
    assert sorted(c["ranks"]) == ["1", "2", "4", "8"]
    for e in c["ranks"].values():
        assert "%016x" % _finish(c["rows"], c["cols"], c["k"], e["nnzb"], int(e["body"], 16)) == e["checksum"]


def test_library_finish_matches_numpy_restatement():
    import bspgemm
    g = _load()["cfg2"]
    got = bspgemm.checksum_finish(g["rows"], g["cols"], g["k"], 8, g["nnzb_c"], int(g["body"], 16))
    assert "%016x" % got == g["checksum"]


def test_sizes_are_the_baseline_configs():
    g = _load()
    assert g["cfg2"]["rows"] == 65536 and g["cfg2"]["k"] == 8 and g["cfg2"]["nprod"] > 5 * 10 ** 7
    assert g["rmat"]["rows"] == 262144 and g["rmat"]["k"] == 8
    assert g["chain"]["rows"] == 16384 and g["chain"]["k"] == 4
