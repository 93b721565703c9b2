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


# ---- partial-product exchange over torch.distributed (NCCL on GPUs, gloo in the CPU tests) ----
# Wire format of one matrix (replaces the reference's header / keys / values messages,
# cu:471-506): int64 header [rows, cols, k, nbrows, nbcols, nnzb], then rowptr int32[nbrows+1],
# colidx int32[nnzb], vals int64[nnzb*k*k] (uint64 bit patterns).

HEADER_LEN = 6


class Payload:
    """A matrix as four tensors on one device (CPU for gloo, CUDA for NCCL)."""

    def __init__(self, header, rowptr, colidx, vals):
        self.header, self.rowptr, self.colidx, self.vals = header, rowptr, colidx, vals

    @property
    def nbytes(self):
        return sum(t.numel() * t.element_size() for t in (self.header, self.rowptr, self.colidx, self.vals))


def send_payload(dist, pl, dst):
    dist.send(pl.header, dst)
    dist.send(pl.rowptr, dst)
    if pl.colidx.numel():
        dist.send(pl.colidx, dst)
        dist.send(pl.vals, dst)


def recv_payload(dist, torch, src, device):
    header = torch.empty(HEADER_LEN, dtype=torch.int64, device=device)
    dist.recv(header, src)
    rows, cols, k, nbr, nbc, nnzb = [int(x) for x in header.cpu().tolist()]
    rowptr = torch.empty(nbr + 1, dtype=torch.int32, device=device)
    dist.recv(rowptr, src)
    colidx = torch.empty(nnzb, dtype=torch.int32, device=device)
    vals = torch.empty(nnzb * k * k, dtype=torch.int64, device=device)
    if nnzb:
        dist.recv(colidx, src)
        dist.recv(vals, src)
    return Payload(header, rowptr, colidx, vals)


def matrix_to_payload(torch, mat):
    """Device matrix (bspgemm.Matrix) -> Payload of CUDA tensors (device-to-device copies)."""
    rows, cols, k, nnzb = mat.info()
    nbr, nbc, _, _, _ = mat.device_view()
    header = torch.tensor([rows, cols, k, nbr, nbc, nnzb], dtype=torch.int64, device="cuda")
    rowptr = torch.empty(nbr + 1, dtype=torch.int32, device="cuda")
    colidx = torch.empty(max(nnzb, 1), dtype=torch.int32, device="cuda")
    vals = torch.empty(max(nnzb * k * k, 1), dtype=torch.int64, device="cuda")
    torch.cuda.synchronize()
    mat.export_device(rowptr.data_ptr(), colidx.data_ptr(), vals.data_ptr())
    return Payload(header, rowptr, colidx[:nnzb], vals[:nnzb * k * k])


def payload_to_matrix(torch, ctx, pl):
    rows, cols, k, nbr, nbc, nnzb = [int(x) for x in pl.header.cpu().tolist()]
    torch.cuda.synchronize()
    return ctx.matrix_from_device_bcsr(rows, cols, k, nbr, nbc, nnzb, pl.rowptr.data_ptr(),
                                       pl.colidx.data_ptr() if nnzb else 0, pl.vals.data_ptr() if nnzb else 0)


def tree_merge(rank, nparts, partial, multiply, send, recv):
    """Step-doubling merge of `nparts` partial products held by ranks 0..nparts-1 (one each).
    multiply(left, right) -> product; send(obj, dst); recv(src) -> obj.  Returns the final
    product on rank 0 and None elsewhere.  Same association as helper2 over the partials
    (cu:569-571), see association_by_schedule."""
    held = partial
    for step in merge_schedule(nparts):
        for receiver, sender in step:
            if rank == sender:
                send(held, receiver)
                held = None
            elif rank == receiver:
                other = recv(sender)
                held = multiply(held, other)
    return held if rank == 0 else None


def merge_groups(nparts):
    """All rank groups that tree_merge_panels uses, in schedule order: [(step, [ranks...]), ...]."""
    out = []
    s = 1
    while s < nparts:
        for g in range(0, nparts, 2 * s):
            if g + s < nparts:
                out.append((s, list(range(g, min(g + 2 * s, nparts)))))
        s *= 2
    return out


def _phase_tracer(rank):
    """BSP_DIST_TRACE=1: rank 0 prints the wall-clock time of every phase of the merge (with a device
    synchronisation before each reading, so the numbers add up but the run is slower)."""
    import os
    import sys
    import time
    if os.environ.get("BSP_DIST_TRACE") != "1" or rank != 0:
        return lambda label: None
    import torch
    last = [time.perf_counter()]

    def mark(label):
        torch.cuda.synchronize()
        now = time.perf_counter()
        sys.stderr.write("[bsp dist] %-22s %8.3f ms\n" % (label, (now - last[0]) * 1e3))
        last[0] = now
    return mark


def tree_merge_panels(rank, nparts, partial, send, recv, send_rows, bcast, row_weights, multiply, concat):
    """Step-doubling merge in which every product is ROW-PANEL SHARDED over the ranks of its
    subtree instead of being computed by the receiving rank alone (the reference's rank 0
    multiplies the P partials sequentially, cu:569-571, which leaves P-1 ranks idle exactly
    where the chain is most expensive).  The association tree is unchanged and every output
    element is still folded by one thread in reference order, so the result is bit-identical.

    For the merge of (g, g+s) the group is ranks g..g+2s-1: rank g scatters block-row panels of
    left (balanced by row_weights(left)), rank g+s broadcasts right, every rank multiplies its
    panel by right, the panel products are gathered on g and concatenated.
      send(obj, dst) / recv(src) -> obj            move a whole matrix
      send_rows(obj, lo, hi, dst)                  move block rows [lo, hi) of a matrix (same row space)
      bcast(obj_or_None, src, group) -> obj        broadcast a matrix inside `group` (list of ranks)
      row_weights(left) -> 1-D array               estimated work per block row
      multiply(left_rows, right)                   product (left_rows holds only this rank's panel)
      concat(list_of_panels)                       one matrix
    Returns the final product on rank 0, None elsewhere."""
    held = partial
    trace = _phase_tracer(rank)
    for s, group in merge_groups(nparts):
        if rank not in group:
            continue
        g = group[0]
        trace("level s=%d start" % s)
        # 1. left: block-row panels from g;  right: from g+s to everybody in the group
        if rank == g:
            bounds = balanced_panels(row_weights(held), len(group))
            trace("row weights")
            for i, r in enumerate(group):
                if r != g:
                    send_rows(held, bounds[i], bounds[i + 1], r)
            left_rows = (held, bounds[0], bounds[1])
        else:
            left_rows = (recv(g), None, None)
        trace("left panels moved")
        right = bcast(held if rank == g + s else None, g + s, group)
        trace("right broadcast")
        # 2. this rank's panel
        if rank == g:
            panel = multiply(left_rows[0], right, (left_rows[1], left_rows[2]))
        else:
            panel = multiply(left_rows[0], right, None)
        trace("panel product")
        # 3. gather on g
        if rank == g:
            panels = [panel] + [recv(r) for r in group[1:]]
            trace("panels gathered")
            held = concat(panels)
            trace("concat")
        else:
            send(panel, g)
            held = None
    return held if rank == 0 else None


def panel_orders(nparts):
    """For tree_merge_distributed: which ranks hold the block-row panels of every intermediate
    product, in ascending row order.  Returns (levels, final): levels[i] = (s, group, left_order,
    right_order) for merge_groups(nparts)[i], final = order of the last product's panels."""
    order = {r: [r] for r in range(nparts)}
    levels = []
    for s, group in merge_groups(nparts):
        g = group[0]
        left, right = order[g], order[g + s]
        new = []
        for p in left:
            new.append(p)
            if p + s <= group[-1]:
                new.append(p + s)  # p hands the lower half of its panel (by work) to p+s
        levels.append((s, group, list(left), list(right)))
        order[g] = new
        del order[g + s]
    return levels, order[0]


def tree_merge_distributed(rank, nparts, partial, send, recv, send_rows, bcast, row_weights, multiply, concat,
                           gather=None):
    """Step-doubling merge in which every product is row-panel sharded over the ranks of its subtree
    AND STAYS DISTRIBUTED: the panels of an intermediate product are not gathered on the group
    leader and scattered again one level up (tree_merge_panels), they are split in place.  For the
    merge of the groups led by g (left operand) and g+s (right operand):
      1. every rank that holds a panel of the right operand broadcasts it inside the group; all
         ranks concatenate the panels (ascending rows) into the full right operand;
      2. every rank p that holds a panel of the left operand keeps the upper half of it (by tiles)
         and sends the lower half to rank p+s, which has just given its own panel away;
      3. every rank multiplies its (half) panel by the right operand: a panel of the product.
    Only the final product is gathered (on rank 0).  Per level: one broadcast per right-hand panel
    and one point-to-point message per pair, all pairs at once.  Same association tree, same fold
    order per output element: bit-identical to tree_merge and to the reference's rank-0 loop
    (cu:569-571).  Callables as in tree_merge_panels; send_rows(obj, lo, hi, dst) moves block rows
    [lo, hi), multiply(left, right, (lo, hi) or None); gather(held, order) (optional) -> list of the
    panels in `order` on rank 0, None elsewhere."""
    levels, final = panel_orders(nparts)
    held = partial
    trace = _phase_tracer(rank)
    for s, group, left, right in levels:
        if rank not in group:
            continue
        panels = [bcast(held if rank == q else None, q, group) for q in right]
        right_full = panels[0] if len(panels) == 1 else concat(panels)
        trace("s=%d right operand" % s)
        if rank in left:
            partner = rank + s
            if partner <= group[-1]:
                w = np.asarray(row_weights(held), dtype=np.int64)
                c = np.cumsum(w)
                mid = int(np.searchsorted(c, (c[-1] + 1) // 2, side="left")) + 1 if c.size and c[-1] > 0 else 0
                send_rows(held, mid, w.size, partner)
                trace("s=%d left half sent" % s)
                held = multiply(held, right_full, (0, mid))
            else:
                held = multiply(held, right_full, None)
        else:
            mine = recv(rank - s)
            trace("s=%d left half received" % s)
            held = multiply(mine, right_full, None)
        trace("s=%d panel product" % s)
    if rank >= nparts:
        return None
    if gather is not None:  # all panels at once (GpuTransport.gather: one batch of receives instead of a loop)
        parts = gather(held, final)
        if rank != 0:
            return None
    else:
        if rank != 0:
            send(held, 0)
            return None
        parts = [held if r == 0 else recv(r) for r in final]
    if len(final) == 1:
        return held
    trace("final gather")
    out = concat(parts)
    trace("final concat")
    return out


def device_row_weights(torch, mat):
    """Tiles per block row of a device matrix (the balance proxy for row panels)."""
    nbr = mat.device_view()[0]
    rp = torch.empty(nbr + 1, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    mat.export_device(rp.data_ptr(), 0, 0)
    return np.diff(rp.cpu().numpy())


def slice_rows_payload(torch, pl, lo, hi):
    """Payload holding only block rows [lo, hi) of `pl` (same row space; CSR slices are contiguous)."""
    nbr = pl.rowptr.numel() - 1
    lo, hi = max(0, min(lo, nbr)), max(0, min(hi, nbr))
    s0, s1 = int(pl.rowptr[lo].item()), int(pl.rowptr[hi].item())
    rows, cols, k = [int(x) for x in pl.header[:3].tolist()]
    kk = k * k
    rowptr = (torch.clamp(pl.rowptr, s0, s1) - s0).to(torch.int32)
    header = pl.header.clone()
    header[5] = s1 - s0
    return Payload(header, rowptr, pl.colidx[s0:s1].contiguous(), pl.vals[s0 * kk:s1 * kk].contiguous())


class GroupBroadcaster:
    """Broadcast of a Payload inside the rank groups of tree_merge_panels, over pre-created
    torch.distributed sub-groups (NCCL: one tree/ring broadcast instead of len(group)-1 sends)."""

    def __init__(self, dist, torch, nparts, device):
        self.dist, self.torch, self.device = dist, torch, device
        self.groups = {}
        for _, ranks in merge_groups(nparts):
            key = tuple(ranks)
            if key not in self.groups:  # new_group is collective over the whole world: same order everywhere
                self.groups[key] = dist.new_group(ranks=ranks)

    def bcast(self, pl, src, ranks):
        dist, torch = self.dist, self.torch
        grp = self.groups[tuple(ranks)]
        header = pl.header if pl is not None else torch.empty(HEADER_LEN, dtype=torch.int64, device=self.device)
        dist.broadcast(header, src, group=grp)
        rows, cols, k, nbr, nbc, nnzb = [int(x) for x in header.cpu().tolist()]
        if pl is None:
            pl = Payload(header, torch.empty(nbr + 1, dtype=torch.int32, device=self.device),
                         torch.empty(nnzb, dtype=torch.int32, device=self.device),
                         torch.empty(nnzb * k * k, dtype=torch.int64, device=self.device))
        dist.broadcast(pl.rowptr, src, group=grp)
        if nnzb:
            dist.broadcast(pl.colidx, src, group=grp)
            dist.broadcast(pl.vals, src, group=grp)
        return pl


class GpuTransport:
    """Moves bspgemm matrices between ranks over NCCL with as few launches and host waits as the
    sizes allow: one 6-word header message, one packed body (rowptr | colidx | vals in ONE buffer,
    8-byte aligned sections), one host wait per received message (the receiver has to read the
    header to size the body).  The context must run on torch's current stream: the device-to-device
    copies of the library, torch's slicing kernels and the NCCL calls are then ordered by the stream
    and nothing else synchronises.  Provides the callables of tree_merge_distributed /
    tree_merge_panels.  `sent` counts the bytes this rank put on the wire."""

    def __init__(self, torch, dist, ctx, nparts):
        self.torch, self.dist, self.ctx = torch, dist, ctx
        self.sent = 0
        self.groups = {}
        self._rowptr_host = {}
        for _, ranks in merge_groups(nparts):
            key = tuple(ranks)
            if key not in self.groups:  # new_group is collective over the whole world: same order everywhere
                self.groups[key] = dist.new_group(ranks=ranks)

    # ---- packing ---------------------------------------------------------------------------
    @staticmethod
    def _layout(nbr, nnzb, kk):
        o_col = (4 * (nbr + 1) + 7) // 8 * 8
        o_val = o_col + (4 * nnzb + 7) // 8 * 8
        return o_col, o_val, o_val + 8 * nnzb * kk

    class _Raw:
        """Raw device memory as an object torch.as_tensor can wrap without copying."""

        def __init__(self, ptr, nbytes):
            self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2}

    def _views(self, mat):
        """(header list, uint8 views of the matrix' rowptr / colidx / vals) -- no copy, no library call."""
        torch = self.torch
        rows, cols, k, nnzb = mat.info()
        nbr, nbc, p_row, p_col, p_val = mat.device_view()
        as_u8 = lambda ptr, n: torch.as_tensor(self._Raw(ptr, n), device="cuda") if n else torch.empty(0, dtype=torch.uint8, device="cuda")  # noqa: E731
        return ([rows, cols, k, nbr, nbc, nnzb], as_u8(p_row, 4 * (nbr + 1)), as_u8(p_col, 4 * nnzb),
                as_u8(p_val, 8 * nnzb * k * k))

    def _export(self, mat):
        torch = self.torch
        hdr, v_row, v_col, v_val = self._views(mat)
        nbr, nnzb, kk = hdr[3], hdr[5], hdr[2] * hdr[2]
        o_col, o_val, total = self._layout(nbr, nnzb, kk)
        body = torch.empty(max(total, 8), dtype=torch.uint8, device="cuda")
        body[:v_row.numel()].copy_(v_row)
        if nnzb:
            body[o_col:o_col + v_col.numel()].copy_(v_col)
            body[o_val:o_val + v_val.numel()].copy_(v_val)
        return hdr, body

    def host_rowptr(self, mat):
        """rowptr of a device matrix on the host (one small download and one wait), cached per matrix."""
        key = id(mat)
        hit = self._rowptr_host.get(key)
        if hit is not None and hit[0] is mat:
            return hit[1]
        _, v_row, _, _ = self._views(mat)
        host = v_row.view(self.torch.int32).cpu().numpy()
        self._rowptr_host = {key: (mat, host)}  # only the latest: the merge asks about one matrix at a time
        return host

    def row_weights(self, mat):
        return np.diff(self.host_rowptr(mat))

    def _pack_rows(self, mat, lo, hi):
        torch = self.torch
        hdr, v_row, v_col, v_val = self._views(mat)
        nbr, kk = hdr[3], hdr[2] * hdr[2]
        rp = self.host_rowptr(mat)
        lo, hi = max(0, min(lo, nbr)), max(0, min(hi, nbr))
        s0, s1 = int(rp[lo]), int(rp[max(hi, lo)])
        n = s1 - s0
        o_col, o_val, total = self._layout(nbr, n, kk)
        body = torch.empty(max(total, 8), dtype=torch.uint8, device="cuda")
        body[:4 * (nbr + 1)].view(torch.int32).copy_(torch.clamp(v_row.view(torch.int32), s0, s1) - s0)
        if n:
            body[o_col:o_col + 4 * n].copy_(v_col[4 * s0:4 * s1])
            body[o_val:o_val + 8 * n * kk].copy_(v_val[8 * s0 * kk:8 * s1 * kk])
        return hdr[:5] + [n], body

    def _to_matrix(self, hdr, body):
        rows, cols, k, nbr, nbc, nnzb = hdr
        o_col, o_val, _ = self._layout(nbr, nnzb, k * k)
        base = body.data_ptr()
        return self.ctx.matrix_from_device_bcsr(rows, cols, k, nbr, nbc, nnzb, base, base + o_col if nnzb else 0,
                                                base + o_val if nnzb else 0)

    # ---- point to point --------------------------------------------------------------------
    def _send(self, hdr, body, dst):
        torch, dist = self.torch, self.dist
        dist.send(torch.tensor(hdr, dtype=torch.int64, device="cuda"), dst)
        dist.send(body, dst)
        self.sent += body.numel() + 8 * HEADER_LEN

    def send(self, mat, dst):
        self._send(*self._export(mat), dst)

    def send_rows(self, mat, lo, hi, dst):
        self._send(*self._pack_rows(mat, lo, hi), dst)

    def recv(self, src):
        torch, dist = self.torch, self.dist
        header = torch.empty(HEADER_LEN, dtype=torch.int64, device="cuda")
        dist.recv(header, src)
        hdr = [int(x) for x in header.cpu().tolist()]
        total = self._layout(hdr[3], hdr[5], hdr[2] * hdr[2])[2]
        body = torch.empty(max(total, 8), dtype=torch.uint8, device="cuda")
        dist.recv(body, src)
        return self._to_matrix(hdr, body)

    # ---- broadcast inside a merge group ----------------------------------------------------------
    def bcast(self, mat, src, ranks):
        torch, dist = self.torch, self.dist
        grp = self.groups[tuple(ranks)]
        if mat is not None:
            hdr, body = self._export(mat)
            header = torch.tensor(hdr, dtype=torch.int64, device="cuda")
            self.sent += body.numel() + 8 * HEADER_LEN
        else:
            header = torch.empty(HEADER_LEN, dtype=torch.int64, device="cuda")
        dist.broadcast(header, src, group=grp)
        if mat is not None:
            dist.broadcast(body, src, group=grp)
            return mat  # the owner keeps using its own copy
        hdr = [int(x) for x in header.cpu().tolist()]
        total = self._layout(hdr[3], hdr[5], hdr[2] * hdr[2])[2]
        body = torch.empty(max(total, 8), dtype=torch.uint8, device="cuda")
        dist.broadcast(body, src, group=grp)
        return self._to_matrix(hdr, body)

    # ---- all panels of the final product to rank 0 -------------------------------------------------
    def gather(self, mat, order):
        """Rank 0 posts the header receives of all peers as one batch, waits once, then posts all body
        receives as one batch (the transfers overlap); the other ranks send header and body."""
        torch, dist = self.torch, self.dist
        rank = dist.get_rank()
        if rank != 0:
            if rank in order:
                hdr, body = self._export(mat)
                header = torch.tensor(hdr, dtype=torch.int64, device="cuda")
                for w in dist.batch_isend_irecv([dist.P2POp(dist.isend, header, 0), dist.P2POp(dist.isend, body, 0)]):
                    w.wait()
                self.sent += body.numel() + 8 * HEADER_LEN
            return None
        peers = [r for r in order if r != 0]
        if not peers:
            return [mat]
        headers = torch.empty((len(peers), HEADER_LEN), dtype=torch.int64, device="cuda")
        for w in dist.batch_isend_irecv([dist.P2POp(dist.irecv, headers[i], r) for i, r in enumerate(peers)]):
            w.wait()
        hdrs = [[int(x) for x in row] for row in headers.cpu().tolist()]
        bodies = [torch.empty(max(self._layout(h[3], h[5], h[2] * h[2])[2], 8), dtype=torch.uint8, device="cuda") for h in hdrs]
        for w in dist.batch_isend_irecv([dist.P2POp(dist.irecv, bodies[i], r) for i, r in enumerate(peers)]):
            w.wait()
        got = {r: self._to_matrix(hdrs[i], bodies[i]) for i, r in enumerate(peers)}
        return [mat if r == 0 else got[r] for r in order]
