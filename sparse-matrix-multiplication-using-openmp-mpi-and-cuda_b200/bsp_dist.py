"""Host-side partitioning logic that replaces the reference's MPI decomposition
(sparse_matrix_mult.cu, "cu:N"): contiguous sub-chains per rank (cu:438-456), the rank-0
reduction of the P partial products (cu:569-571) re-expressed as a step-doubling tree over
GPUs, and block-row panels of a single product.  Pure Python/numpy; the data movement that
uses it lives in bench.py / the tests (torch.distributed send/recv of the CSR-of-blocks arrays).
"""
import numpy as np


def split_chain(n, P):
    """[(start, end)] inclusive index ranges, one per working rank, exactly as cu:438-456:
    per = n // P; rank r < P-1 owns [r*per, (r+1)*per - 1], the last rank owns the rest.
    If n < P only rank 0 works, on the whole chain (cu:612-616)."""
    per = n // P
    if per == 0:
        return [(0, n - 1)]
    return [(r * per, (n - 1) if r == P - 1 else (r + 1) * per - 1) for r in range(P)]


def tree_levels(n):
    """helper2's schedule (cu:287-327) on n operands as a list of levels; each level is a list
    of (left, right) positions multiplied, positions refer to the level's own input list; an
    odd tail is carried."""
    levels = []
    while n > 1:
        levels.append([(i, i + 1) for i in range(0, n - 1, 2)])
        n = n // 2 + (n & 1)
    return levels


def merge_schedule(P):
    """Step-doubling tree over P ranks with the same association as helper2 over P partials:
    list of steps; each step is a list of (receiver, sender): at step s rank g with
    g % (2s) == 0 and g+s < P receives from g+s and multiplies mine * received."""
    steps = []
    s = 1
    while s < P:
        steps.append([(g, g + s) for g in range(0, P, 2 * s) if g + s < P])
        s *= 2
    return steps


def association(n, P):
    """Fully parenthesised string of the product the P-rank run computes (for tests)."""
    def tree(items):
        while len(items) > 1:
            nxt = ["(%s%s)" % (items[i], items[i + 1]) for i in range(0, len(items) - 1, 2)]
            if len(items) & 1:
                nxt.append(items[-1])
            items = nxt
        return items[0]
    parts = [tree(["M%d" % i for i in range(s, e + 1)]) for s, e in split_chain(n, P)]
    return tree(parts)


def association_by_schedule(n, P):
    """Same, computed by running merge_schedule instead of helper2 over the partials."""
    def tree(items):
        while len(items) > 1:
            nxt = ["(%s%s)" % (items[i], items[i + 1]) for i in range(0, len(items) - 1, 2)]
            if len(items) & 1:
                nxt.append(items[-1])
            items = nxt
        return items[0]
    parts = [tree(["M%d" % i for i in range(s, e + 1)]) for s, e in split_chain(n, P)]
    held = dict(enumerate(parts))
    for step in merge_schedule(len(parts)):
        for recv, send in step:
            held[recv] = "(%s%s)" % (held[recv], held.pop(send))
    return held[0]


def balanced_panels(row_work, P):
    """Cut block rows [0, len(row_work)) into P contiguous panels of near-equal total work
    (row_work = tile products per block row, from the symbolic count). Returns P+1 bounds."""
    w = np.asarray(row_work, dtype=np.float64)
    cum = np.concatenate([[0.0], np.cumsum(w)])
    total = cum[-1]
    bounds = [0]
    for p in range(1, P):
        bounds.append(int(np.searchsorted(cum, total * p / P, side="left")))
    bounds.append(len(w))
    for i in range(1, len(bounds)):
        bounds[i] = max(bounds[i], bounds[i - 1])
    return bounds


def chain_ranks_on_one_context(ctx, mats, P, stats_list=None):
    """The P-rank association tree executed on a single context (used for parity tests of the
    tree itself and by single-GPU runs that must reproduce `mpirun -np P` bit for bit)."""
    parts = []
    for s, e in split_chain(len(mats), P):
        parts.append(ctx.helper2(mats[s:e + 1]))
    return ctx.helper2(parts) if len(parts) > 1 else parts[0]
