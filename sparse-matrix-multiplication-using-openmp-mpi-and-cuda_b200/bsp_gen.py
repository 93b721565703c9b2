"""Seeded synthetic block-sparse matrices for the BASELINE.json configurations.

One documented PRNG: splitmix64 (Steele, Lea, Flood 2014), evaluated counter-style
(value i of stream s = mix(s + (i+1)*GOLDEN)), so any slice can be regenerated
independently and the numpy code below is vectorised.  Seed per matrix = base + index.
"""
import numpy as np

GOLDEN = np.uint64(0x9E3779B97F4A7C15)
U64_MAX = np.uint64(0xFFFFFFFFFFFFFFFF)


def splitmix64(seed, n, offset=0):
    """n consecutive outputs of splitmix64 started at `seed`, skipping `offset` outputs."""
    with np.errstate(over="ignore"):
        i = np.arange(offset + 1, offset + n + 1, dtype=np.uint64)
        z = np.uint64(seed & 0xFFFFFFFFFFFFFFFF) + i * GOLDEN
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        return z ^ (z >> np.uint64(31))


def values(seed, n, kind="u64"):
    """Tile entries. kind: 'u64' uniform in [1, 2^64-2]; 'small' uniform in [1, 1000];
    'edgy' drawn from a set that makes the 2^64-1 edge of the reference arithmetic fire
    often (0, 1, 2, 2^32, 2^64-2, 2^64-1, ...)."""
    r = splitmix64(seed ^ 0x5bd1e995, n)
    if kind == "u64":
        return r % (U64_MAX - np.uint64(1)) + np.uint64(1)
    if kind == "small":
        return r % np.uint64(1000) + np.uint64(1)
    if kind == "edgy":
        table = np.array([0, 1, 1, 2, 3, 1 << 32, (1 << 32) + 1, (1 << 63), (1 << 64) - 2, (1 << 64) - 1,
                          (1 << 64) - 3, (1 << 32) - 1, 0xFFFFFFFF00000001, 5, 7, 0x5555555555555555], dtype=np.uint64)
        return table[(r >> np.uint64(7)) % np.uint64(table.size)]
    raise ValueError(kind)


def bernoulli_pattern(G_rows, G_cols, density, seed):
    """Each of the G_rows*G_cols block slots is kept with probability `density`.
    Returns (brow, bcol) int32 sorted by (brow, bcol)."""
    thr = np.uint64(min(int(density * 2.0 ** 64), (1 << 64) - 1))
    rows, cols = [], []
    step = max(1, (1 << 24) // max(G_cols, 1))
    for r0 in range(0, G_rows, step):
        r1 = min(G_rows, r0 + step)
        u = splitmix64(seed, (r1 - r0) * G_cols, offset=r0 * G_cols)
        idx = np.flatnonzero(u < thr)
        rows.append((idx // G_cols + r0).astype(np.int32))
        cols.append((idx % G_cols).astype(np.int32))
    return np.concatenate(rows), np.concatenate(cols)


def rmat_pattern(scale, edge_factor, seed, abcd=(0.57, 0.19, 0.19, 0.05)):
    """R-MAT (Chakrabarti, Zhan, Faloutsos 2004) block pattern on a 2^scale square grid of
    blocks, edge_factor * 2^scale edges before de-duplication. Sorted, unique."""
    n_edges = edge_factor << scale
    a, b, c, _ = abcd
    row = np.zeros(n_edges, dtype=np.int64)
    col = np.zeros(n_edges, dtype=np.int64)
    for lvl in range(scale):
        u = splitmix64(seed + 7919 * (lvl + 1), n_edges).astype(np.float64) / 2.0 ** 64
        rbit = u >= (a + b)
        cbit = ((u >= a) & (u < a + b)) | (u >= a + b + c)
        row = (row << 1) | rbit
        col = (col << 1) | cbit
    key = np.unique((row << 32) | col)
    return (key >> 32).astype(np.int32), (key & 0xFFFFFFFF).astype(np.int32)


def banded_pattern(G, half_band, density, seed):
    """Blocks only within |brow - bcol| <= half_band, each kept with probability `density`
    (a bounded-fill pattern for long chains: the band of a product of m factors is m*half_band)."""
    w = 2 * half_band + 1
    thr = np.uint64(min(int(density * 2.0 ** 64), (1 << 64) - 1))
    u = splitmix64(seed, G * w)
    idx = np.flatnonzero(u < thr)
    r = idx // w
    c = r + (idx % w) - half_band
    ok = (c >= 0) & (c < G)
    return r[ok].astype(np.int32), c[ok].astype(np.int32)


def random_matrix(n_rows, n_cols, k, density, seed, kind="u64", pattern="bernoulli"):
    """(rows, cols, k, brow, bcol, vals[nnzb,k,k]) for an n_rows x n_cols matrix of k x k tiles."""
    Gr, Gc = n_rows // k, n_cols // k
    if pattern == "bernoulli":
        br, bc = bernoulli_pattern(Gr, Gc, density, seed)
    else:
        raise ValueError(pattern)
    v = values(seed, br.size * k * k, kind).reshape(-1, k, k)
    return n_rows, n_cols, k, br, bc, v
