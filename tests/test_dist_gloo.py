"""world_size-2 (and 3) runs of the partial-product exchange over gloo on the CPU: the wire
format, the step-doubling schedule and the sub-chain split, with the CPU oracle standing in
for the multiply (this is a test of the host-side logic, not of the product's compute path)."""
import os
import socket
import tempfile

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import bsp_dist
import bsp_gen
import orc


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _omat_to_payload(m):
    br, bc, v = m.coo()
    nbr = m.rows // m.k  # the generators use coordinates < rows/k; keeps the row space of slices identical
    nbc = m.cols // m.k
    rowptr = np.zeros(nbr + 1, np.int32)
    np.add.at(rowptr, br + 1, 1)
    rowptr = np.cumsum(rowptr).astype(np.int32)
    header = torch.tensor([m.rows, m.cols, m.k, nbr, nbc, br.size], dtype=torch.int64)
    return bsp_dist.Payload(header, torch.from_numpy(rowptr), torch.from_numpy(bc.copy()),
                            torch.from_numpy(v.reshape(-1).view(np.int64).copy()))


def _payload_to_omat(pl):
    rows, cols, k, nbr, nbc, nnzb = [int(x) for x in pl.header.tolist()]
    rowptr = pl.rowptr.numpy()
    br = np.repeat(np.arange(nbr, dtype=np.int32), np.diff(rowptr))
    return orc.OMat.from_coo(rows, cols, k, br, pl.colidx.numpy(), pl.vals.numpy().view(np.uint64))


def _row_slice(m, lo, hi):
    br, bc, v = m.coo()
    keep = (br >= lo) & (br < hi)
    return orc.OMat.from_coo(m.rows, m.cols, m.k, br[keep], bc[keep], v[keep])


def _concat(panels):
    parts = [p.coo() for p in panels]
    return orc.OMat.from_coo(panels[0].rows, panels[0].cols, panels[0].k, np.concatenate([p[0] for p in parts]),
                             np.concatenate([p[1] for p in parts]), np.concatenate([p[2] for p in parts]))


def _weights(m):
    return np.bincount(m.coo()[0], minlength=m.rows // m.k)


def _worker(rank, world, port, n_mats, k, G, outdir, panels=False):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        mats = [orc.OMat.from_coo(*bsp_gen.random_matrix(G * k, G * k, k, 0.3, 3000 + i, "edgy")) for i in range(n_mats)]
        parts = bsp_dist.split_chain(n_mats, world)
        partial = None
        if rank < len(parts):
            s, e = parts[rank]
            partial = orc.chain(mats[s:e + 1])[0]
        send = lambda m, dst: bsp_dist.send_payload(dist, _omat_to_payload(m), dst)  # noqa: E731
        recv = lambda src: _payload_to_omat(bsp_dist.recv_payload(dist, torch, src, "cpu"))  # noqa: E731
        if panels:
            gb = bsp_dist.GroupBroadcaster(dist, torch, len(parts), "cpu")
            merge = bsp_dist.tree_merge_distributed if panels == "dist" else bsp_dist.tree_merge_panels
            res = merge(
                rank, len(parts), partial, send, recv,
                send_rows=lambda m, lo, hi, dst: bsp_dist.send_payload(
                    dist, bsp_dist.slice_rows_payload(torch, _omat_to_payload(m), lo, hi), dst),
                bcast=lambda m, src, grp: _payload_to_omat(gb.bcast(_omat_to_payload(m) if m is not None else None, src, grp)),
                row_weights=_weights,
                multiply=lambda a, b, rng: orc.multiply(a if rng is None else _row_slice(a, rng[0], rng[1]), b)[0],
                concat=_concat)
        else:
            res = bsp_dist.tree_merge(rank, len(parts), partial, multiply=lambda a, b: orc.multiply(a, b)[0],
                                      send=send, recv=recv)
        if rank == 0:
            res.prune_zero().write(os.path.join(outdir, "matrix"))
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,n_mats,panels", [(2, 5, False), (2, 1, False), (3, 7, False), (2, 5, True), (4, 9, True),
                                                 (2, 5, "dist"), (3, 7, "dist"), (4, 9, "dist"), (2, 1, "dist")])
def test_tree_merge_over_gloo_matches_rank_oracle(world, n_mats, panels):
    k, G = 2, 8
    with tempfile.TemporaryDirectory() as d:
        mp.spawn(_worker, args=(world, _free_port(), n_mats, k, G, d, panels), nprocs=world, join=True)
        mats = [orc.OMat.from_coo(*bsp_gen.random_matrix(G * k, G * k, k, 0.3, 3000 + i, "edgy")) for i in range(n_mats)]
        want = orc.chain(mats, ranks=world)[0].prune_zero()
        p = os.path.join(d, "want")
        want.write(p)
        assert open(os.path.join(d, "matrix"), "rb").read() == open(p, "rb").read()
