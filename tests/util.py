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


# ---- numpy restatement of the content checksum (csrc/bsp_checksum.cu states the definition) ----
_M64 = (1 << 64) - 1


def _splitmix64(x):
    x = (x + 0x9E3779B97F4A7C15) & _M64
    x = ((x ^ (x >> 30)) * 0xBF58476D1CE4E5B9) & _M64
    x = ((x ^ (x >> 27)) * 0x94D049BB133111EB) & _M64
    return x ^ (x >> 31)


def _fmix32(h):
    h = h.astype(np.uint32).copy()
    h ^= h >> np.uint32(16)
    h *= np.uint32(0x85EBCA6B)
    h ^= h >> np.uint32(13)
    h *= np.uint32(0xC2B2AE35)
    h ^= h >> np.uint32(16)
    return h


def checksum_numpy(rows, cols, k, brow, bcol, vals):
    """vals: uint64 or float32 array of nnzb*k*k entries, tiles sorted or not (the sum is
    order-independent); duplicates must already be resolved."""
    vals = np.ascontiguousarray(vals)
    esize = vals.dtype.itemsize
    nnzb = len(brow)
    header = _splitmix64(rows & 0xFFFFFFFF)
    for x in (cols & 0xFFFFFFFF, k, esize, nnzb):
        header = _splitmix64(header ^ x)
    body = 0
    if nnzb:
        W = k * k * esize // 4
        words = vals.reshape(-1).view(np.uint32).reshape(nnzb, W)
        with np.errstate(over="ignore"):
            s1 = _fmix32(np.asarray(brow).astype(np.uint32) * np.uint32(0x9E3779B1)
                         + np.asarray(bcol).astype(np.uint32) * np.uint32(0x85EBCA77) + np.uint32(0x27D4EB2F))
            s2 = _fmix32(s1 ^ np.uint32(0x165667B1))
            i = np.arange(W, dtype=np.uint32)
            m = (s1[:, None] + i[None, :] * np.uint32(0x9E3779B1)) | np.uint32(1)
            tile = (s2.astype(np.uint64) << np.uint64(32)) | s1.astype(np.uint64)
            tile = tile + (words.astype(np.uint64) * m.astype(np.uint64)).sum(axis=1, dtype=np.uint64)
            body = int(tile.sum(dtype=np.uint64))
    return _splitmix64(body ^ header)
