#!/usr/bin/env python
"""Regenerate tests/golden/*.json by running the reference's own code (oracle/_ref, the
host build made by oracle/build_ref.sh from /root/reference) on hand-made known-answer
inputs and on small seeded random inputs.  Needs /root/reference (run in the build
container, not on the GPU box).  The fixtures it writes are committed.

    python tests/golden/make_golden.py
"""
import hashlib
import json
import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
sys.path.insert(0, os.path.join(ROOT, "sparse-matrix-multiplication-using-openmp-mpi-and-cuda_b200"))
import orc  # noqa: E402
import bsp_gen  # noqa: E402

MAX = (1 << 64) - 1


def M(rows, cols, *blocks):
    return dict(rows=rows, cols=cols, blocks=[[r, c, list(v)] for r, c, v in blocks])


# inputs are restated here, outputs come from the reference.
KATS = [
    dict(id="t1_chain3_rect", k=2, mats=[
        M(4, 4, (0, 0, [1, 2, 3, 4]), (1, 1, [5, 6, 7, 8])),
        M(4, 6, (0, 0, [1, 0, 0, 1]), (0, 2, [2, 2, 2, 2]), (1, 1, [MAX, 1, 0, 0])),
        M(6, 2, (0, 0, [1, 1, 1, 1]), (2, 0, [3, 0, 0, 3]))]),
    dict(id="t2_max_edges_and_zero_tiles", k=2, mats=[
        M(8, 8, (0, 0, [MAX, 0, 0, 0]), (1, 1, [1, 0, 0, 1]), (2, 2, [0, 0, 0, 0]),
          (3, 3, [4294967296, 1, 18446744073709551614, 2])),
        M(8, 8, (0, 0, [1, 0, 0, 0]), (1, 1, [7, 7, 7, 7]), (2, 2, [9, 9, 9, 9]), (3, 3, [4294967296, 1, 1, 1]))]),
    dict(id="t3_max_edges_no_zero_tiles", k=2, mats=[
        M(8, 8, (1, 1, [1, 0, 0, 1]), (3, 3, [4294967296, 1, 18446744073709551614, 2])),
        M(8, 8, (0, 0, [1, 0, 0, 0]), (1, 1, [7, 7, 7, 7]), (2, 2, [9, 9, 9, 9]), (3, 3, [4294967296, 1, 1, 1]))]),
    dict(id="t4_all_zero_product", k=2, mats=[
        M(4, 4, (0, 0, [0, 0, 0, 0])), M(4, 4, (0, 0, [5, 5, 5, 5]))]),
    dict(id="t5_chain5_tree", k=2, mats=[
        M(4, 4, (0, 0, [i, 1, 0, i]), (1, 1, [1, i, 1, 1])) for i in range(1, 6)]),
    dict(id="t6_single_passthrough_sorted", k=2, mats=[
        M(4, 4, (1, 0, [1, 2, 3, 4]), (0, 1, [9, 8, 7, 6]))]),
    dict(id="t7_sum_hits_max_then_plus5", k=2, mats=[
        M(2, 4, (0, 0, [MAX - 1, 1, 0, 0]), (0, 1, [5, 0, 0, 0])),
        M(4, 2, (0, 0, [1, 0, 1, 0]), (1, 0, [1, 0, 0, 0]))]),
    dict(id="t8_same_terms_other_order", k=2, mats=[
        M(2, 4, (0, 1, [MAX - 1, 1, 0, 0]), (0, 0, [5, 0, 0, 0])),
        M(4, 2, (1, 0, [1, 0, 1, 0]), (0, 0, [1, 0, 0, 0]))]),
    dict(id="t9_sum_exactly_max_pruned", k=2, mats=[
        M(2, 2, (0, 0, [MAX - 1, 1, 0, 0])), M(2, 2, (0, 0, [1, 0, 1, 0]))]),
    dict(id="t10_duplicate_key_last_wins", k=2, mats=[
        M(4, 4, (0, 0, [1, 1, 1, 1]), (0, 0, [2, 0, 0, 2])), M(4, 4, (0, 0, [3, 4, 5, 6]))]),
    dict(id="t11_k1", k=1, mats=[
        M(3, 3, (0, 0, [MAX]), (0, 1, [2]), (2, 2, [MAX - 1])),
        M(3, 3, (0, 0, [1]), (1, 0, [3]), (2, 2, [2]))]),
    dict(id="t12_k3_chain4", k=3, mats=[
        M(6, 6, (0, 0, list(range(1, 10))), (1, 0, [MAX, 1, 2, 3, MAX - 1, 5, 1 << 32, 7, 8])) for _ in range(4)]),
    dict(id="t13_offset_coordinates", k=2, mats=[
        M(100, 100, (40, 96, [1, 2, 3, 4]), (0, 2, [1, 1, 1, 1])),
        M(100, 100, (96, 0, [1, 0, 0, 1]), (2, 98, [2, 0, 0, 2]))]),
]


def to_omat(m, k):
    b = m["blocks"]
    return orc.OMat.from_coo(m["rows"], m["cols"], k, [x[0] for x in b], [x[1] for x in b],
                             np.array([x[2] for x in b], dtype=np.uint64).reshape(-1) if b else np.zeros(0, np.uint64))


def write_raw_folder(folder, mats, k):
    """Writes the blocks in the GIVEN order (duplicates and unsorted input are part of the test)."""
    os.makedirs(folder, exist_ok=True)
    with open(os.path.join(folder, "size"), "w") as f:
        f.write("%d %d\n" % (len(mats), k))
    for i, m in enumerate(mats):
        with open(os.path.join(folder, "matrix%d" % (i + 1)), "w") as f:
            f.write("%d %d\n%d\n" % (m["rows"], m["cols"], len(m["blocks"])))
            for r, c, v in m["blocks"]:
                f.write("%d %d\n" % (r, c))
                for rr in range(k):
                    f.write(" ".join(str(int(x)) for x in v[rr * k:(rr + 1) * k]) + "\n")


def random_case(cid, seed, n_mats, G, k, density, kind, ranks):
    mats = []
    for i in range(n_mats):
        _, _, _, br, bc, v = bsp_gen.random_matrix(G * k, G * k, k, density, seed + i, kind)
        mats.append(dict(rows=G * k, cols=G * k,
                         blocks=[[int(r), int(c), [int(x) for x in t.reshape(-1)]] for r, c, t in zip(br, bc, v)]))
    return dict(id=cid, k=k, mats=mats, ranks=ranks,
                gen=dict(seed=seed, n_mats=n_mats, G=G, density=density, kind=kind))


def main():
    if not orc.have_ref("ref_host_fixed"):
        sys.exit("oracle/_ref missing: run oracle/build_ref.sh where /root/reference exists")
    cases = [dict(c, ranks=1) for c in KATS]
    cases.append(dict(KATS[4], id="t5_chain5_ranks2", ranks=2))
    cases.append(dict(KATS[4], id="t5_chain5_ranks3", ranks=3))
    cases.append(dict(KATS[4], id="t5_chain5_ranks8_fewer_mats_than_ranks", ranks=8))
    cases.append(random_case("r1_edgy_k2_chain7", 101, 7, 6, 2, 0.45, "edgy", 1))
    cases.append(random_case("r2_edgy_k2_chain7_ranks3", 101, 7, 6, 2, 0.45, "edgy", 3))
    cases.append(random_case("r3_edgy_k4_pair", 202, 2, 12, 4, 0.3, "edgy", 1))
    cases.append(random_case("r4_u64_k8_pair", 303, 2, 10, 8, 0.3, "u64", 1))
    cases.append(random_case("r5_u64_k4_chain6_ranks2", 404, 6, 8, 4, 0.3, "u64", 2))
    cases.append(random_case("r6_small_k5_chain3", 505, 3, 7, 5, 0.4, "small", 1))
    cases.append(random_case("r7_edgy_k16_pair", 606, 2, 4, 16, 0.6, "edgy", 1))
    cases.append(random_case("r8_edgy_k32_pair", 707, 2, 3, 32, 0.7, "edgy", 1))
    out = []
    for c in cases:
        with tempfile.TemporaryDirectory() as d:
            write_raw_folder(d, c["mats"], c["k"])
            r = orc.run_ref(d, ranks=c["ranks"], variant="ref_host_fixed")
            assert r["returncode"] == 0 and r["matrix"] is not None, (c["id"], r["stderr"])
            shipped = orc.run_ref(d, ranks=c["ranks"], variant="ref_host")
        rec = dict(id=c["id"], k=c["k"], ranks=c["ranks"], matrix_text=r["matrix"].decode(),
                   multiplying=[list(map(int, x)) for x in r["multiplying"]] if c["ranks"] == 1 else None,
                   shipped_build_returncode=shipped["returncode"],
                   shipped_build_same_output=(shipped["matrix"] == r["matrix"]))
        if "gen" in c:
            rec["gen"] = c["gen"]
        else:
            rec["mats"] = c["mats"]
        out.append(rec)
        print("%-45s ranks=%d out=%4d B  shipped rc=%s same=%s" % (
            c["id"], c["ranks"], len(r["matrix"]), shipped["returncode"], rec["shipped_build_same_output"]))
    with open(os.path.join(ROOT, "tests", "golden", "kats.json"), "w") as f:
        json.dump(out, f, indent=0, separators=(",", ":"))

    # config-1 sized case (BASELINE.json configs[0]): hash of the reference's output file only
    big = []
    for cid, kind, seed in (("cfg1_u64", "u64", 1), ("cfg1_small", "small", 2)):
        mats = [orc.OMat.from_coo(*_rm(1024, 4, 0.05, seed * 1000 + i, kind)) for i in range(2)]
        with tempfile.TemporaryDirectory() as d:
            orc.write_folder(d, mats, 4)
            r = orc.run_ref(d, variant="ref_host_fixed")
        txt = r["matrix"]
        big.append(dict(id=cid, n=1024, k=4, density=0.05, kind=kind, seeds=[seed * 1000, seed * 1000 + 1],
                        nnzb_in=[m.nnzb for m in mats], out_bytes=len(txt), out_nnzb=int(txt.split(b"\n")[1]),
                        sha256=hashlib.sha256(txt).hexdigest()))
        print(big[-1])
    with open(os.path.join(ROOT, "tests", "golden", "cfg1_hash.json"), "w") as f:
        json.dump(big, f, indent=1)


def _rm(n, k, d, seed, kind):
    rows, cols, k, br, bc, v = bsp_gen.random_matrix(n, n, k, d, seed, kind)
    return rows, cols, k, br, bc, v


if __name__ == "__main__":
    main()
