"""Shared helpers for the tests: golden-fixture loading and conversions."""
import json
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")


def case_mats(case):
    """list of (rows, cols, brow[int32], bcol[int32], vals[uint64, nnzb*k*k]) in FILE order
    (duplicates and unsorted keys preserved)."""
    k = case["k"]
    if "mats" in case:
        out = []
        for m in case["mats"]:
            b = m["blocks"]
            out.append((m["rows"], m["cols"], np.array([x[0] for x in b], np.int32), np.array([x[1] for x in b], np.int32),
                        np.array([x[2] for x in b], dtype=np.uint64).reshape(-1) if b else np.zeros(0, np.uint64)))
        return out
    import bsp_gen
    g = case["gen"]
    out = []
    for i in range(g["n_mats"]):
        rows, cols, _, br, bc, v = bsp_gen.random_matrix(g["G"] * k, g["G"] * k, k, g["density"], g["seed"] + i, g["kind"])
        out.append((rows, cols, br, bc, v.reshape(-1)))
    return out


def load_kats():
    with open(os.path.join(GOLDEN, "kats.json")) as f:
        return json.load(f)


def load_cfg1():
    with open(os.path.join(GOLDEN, "cfg1_hash.json")) as f:
        return json.load(f)


def matrix_text(rows, cols, k, brow, bcol, vals):
    """The reference's output layout (cu:595-608) from a sorted coordinate list."""
    vals = np.asarray(vals, dtype=np.uint64).reshape(-1, k, k)
    lines = ["%d %d" % (rows, cols), "%d" % len(brow)]
    for r, c, t in zip(brow, bcol, vals):
        lines.append("%d %d" % (r, c))
        for row in t:
            lines.append(" ".join(str(int(x)) for x in row))
    return "\n".join(lines) + "\n"
