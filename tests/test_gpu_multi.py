"""Multi-GPU paths (skipped on a single-GPU box): a4 with sub-chains on two devices and the
torch.distributed (NCCL) tree merge, both against the P-rank oracle."""
import os
import subprocess
import sys
import tempfile

import numpy as np
import pytest

import bsp_gen
import bspgemm
import orc
import util

pytestmark = pytest.mark.gpu


def _ngpu():
    import torch
    return torch.cuda.device_count()


@pytest.mark.skipif("_ngpu() < 2")
def test_a4_two_devices_matches_two_rank_reference_tree():
    k, G, n = 4, 40, 9
    mats = [orc.OMat.from_coo(*bsp_gen.random_matrix(G * k, G * k, k, 0.08, 8100 + i, "edgy")) for i in range(n)]
    exe = os.path.join(os.path.dirname(bspgemm.__file__), "a4")
    with tempfile.TemporaryDirectory() as d:
        orc.write_folder(os.path.join(d, "in"), mats, k)
        want = orc.chain(mats, ranks=2)[0].prune_zero()
        want.write(os.path.join(d, "want"))
        logs = []
        # second run: every product of the merge sharded by block rows over the two devices
        for env in (os.environ, dict(os.environ, BSP_A4_SHARD_MIN="1")):
            p = subprocess.run([exe, os.path.join(d, "in"), "--ranks", "2", "--devices", "2"], cwd=d, env=env,
                               stdout=subprocess.PIPE, stderr=subprocess.PIPE, timeout=300)
            assert p.returncode == 0, p.stderr.decode()
            assert open(os.path.join(d, "matrix"), "rb").read() == open(os.path.join(d, "want"), "rb").read()
            logs.append([ln for ln in p.stdout.decode().splitlines() if ln.startswith("multiplying")])
            os.remove(os.path.join(d, "matrix"))
        assert logs[0] == logs[1] and logs[0][-1] == "multiplying 0 1"
        # four ranks on two devices: two merge levels, the partial of rank 3 is a longer sub-chain
        p = subprocess.run([exe, os.path.join(d, "in"), "--ranks", "4", "--devices", "2"], cwd=d,
                           env=dict(os.environ, BSP_A4_SHARD_MIN="1"), stdout=subprocess.PIPE, stderr=subprocess.PIPE, timeout=300)
        assert p.returncode == 0, p.stderr.decode()
        orc.chain(mats, ranks=4)[0].prune_zero().write(os.path.join(d, "want4"))
        assert open(os.path.join(d, "matrix"), "rb").read() == open(os.path.join(d, "want4"), "rb").read()


NCCL_WORKER = r'''
import os, sys
sys.path.insert(0, sys.argv[1]); sys.path.insert(0, sys.argv[2])
import numpy as np, torch, torch.distributed as dist
import bsp_dist, bsp_gen, bspgemm
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(rank)
dist.init_process_group("nccl", device_id=torch.device("cuda", rank))
k, G, n = 4, 40, 9
stream = torch.cuda.Stream()
torch.cuda.set_stream(stream)   # GpuTransport needs the context on torch's current stream
ctx = bspgemm.Context(rank, stream=stream.cuda_stream)
parts = bsp_dist.split_chain(n, world)
s, e = parts[rank]
mats = [ctx.matrix_from_coo(*bsp_gen.random_matrix(G * k, G * k, k, 0.08, 8100 + i, "edgy")) for i in range(s, e + 1)]
partial = ctx.helper2(mats)
send = lambda m, dst: bsp_dist.send_payload(dist, bsp_dist.matrix_to_payload(torch, m), dst)
recv = lambda src: bsp_dist.payload_to_matrix(torch, ctx, bsp_dist.recv_payload(dist, torch, src, "cuda"))
if sys.argv[4] == "panels":
    gb = bsp_dist.GroupBroadcaster(dist, torch, len(parts), "cuda")
    res = bsp_dist.tree_merge_panels(
        rank, len(parts), partial, send, recv,
        lambda m, lo, hi, dst: bsp_dist.send_payload(dist, bsp_dist.slice_rows_payload(torch, bsp_dist.matrix_to_payload(torch, m), lo, hi), dst),
        lambda m, src, grp: bsp_dist.payload_to_matrix(torch, ctx, gb.bcast(bsp_dist.matrix_to_payload(torch, m) if m is not None else None, src, grp)),
        lambda m: bsp_dist.device_row_weights(torch, m),
        lambda a, b, rng: ctx.helper(a, b, panel=rng), ctx.concat_panels)
elif sys.argv[4] == "distributed":
    tp = bsp_dist.GpuTransport(torch, dist, ctx, len(parts))
    res = bsp_dist.tree_merge_distributed(rank, len(parts), partial, tp.send, tp.recv, tp.send_rows, tp.bcast,
                                          tp.row_weights, lambda a, b, rng: ctx.helper(a, b, panel=rng), ctx.concat_panels,
                                          gather=tp.gather if len(sys.argv) > 5 else None)
    assert tp.sent > 0
else:
    res = bsp_dist.tree_merge(rank, len(parts), partial, lambda a, b: ctx.helper(a, b), send, recv)
if rank == 0:
    res.prune_zero().write_text(sys.argv[3])
dist.barrier()
dist.destroy_process_group()
'''


@pytest.mark.skipif("_ngpu() < 2")
@pytest.mark.parametrize("mode", ["rank0", "panels", "distributed", "distributed+gather"])
def test_nccl_tree_merge_two_ranks(mode):
    k, G, n = 4, 40, 9
    pkg = os.path.dirname(bspgemm.__file__)
    with tempfile.TemporaryDirectory() as d:
        script = os.path.join(d, "w.py")
        open(script, "w").write(NCCL_WORKER)
        out = os.path.join(d, "matrix")
        p = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                            "--master-addr", "127.0.0.1", "--master-port", "29611", script, pkg,
                            os.path.join(util.ROOT, "oracle"), out] + mode.split("+"), stdout=subprocess.PIPE, stderr=subprocess.PIPE, timeout=600)
        assert p.returncode == 0, p.stderr.decode()[-3000:]
        mats = [orc.OMat.from_coo(*bsp_gen.random_matrix(G * k, G * k, k, 0.08, 8100 + i, "edgy")) for i in range(n)]
        want = orc.chain(mats, ranks=2)[0].prune_zero()
        want.write(os.path.join(d, "want"))
        assert open(out, "rb").read() == open(os.path.join(d, "want"), "rb").read()


# ---- the C-ABI transport and the distributed chain (csrc/bsp_dist.cu) ------------------------------
CAPI_WORKER = r'''
import os, sys
sys.path.insert(0, sys.argv[1]); sys.path.insert(0, sys.argv[2])
import numpy as np, torch, torch.distributed as dist
import bsp_dist, bsp_gen, bspgemm
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(rank)
dist.init_process_group("nccl", device_id=torch.device("cuda", rank))
ctx = bspgemm.Context(rank)   # its own stream: nothing here depends on torch's
tok = torch.zeros(bspgemm.COMM_ID_BYTES, dtype=torch.uint8, device="cuda")
if rank == 0:
    tok.copy_(torch.frombuffer(bytearray(bspgemm.comm_unique_id()), dtype=torch.uint8))
dist.broadcast(tok, 0)
torch.cuda.synchronize()
comm = bspgemm.Comm(ctx, tok.cpu().numpy().tobytes(), world, rank)
mode, out = sys.argv[4], sys.argv[3]
if mode == "chain":
    k, G, n = 4, 40, int(sys.argv[5])
    parts = bsp_dist.split_chain(n, world)
    s, e = parts[rank] if rank < len(parts) else (0, -1)
    mats = [ctx.matrix_from_coo(*bsp_gen.random_matrix(G * k, G * k, k, 0.08, 8100 + i, "edgy")) for i in range(s, e + 1)]
    for rep in range(2):   # twice: the second run reuses the pool and the NCCL channels
        res, log = comm.chain(mats, want_log=True, log_base=s)
        if rank == 0:
            assert res is not None
            res.prune_zero().write_text(out)
        else:
            assert res is None
    assert comm.bytes_sent > 0 or world == 1 or n < world
elif mode == "transport":
    k, G = 8, 24
    m = bsp_gen.random_matrix(G * k, G * k, k, 0.1, 4242, "edgy")
    f = (G * k, G * k, k, m[3], m[4], (bsp_gen.splitmix64(7, m[3].size * k * k).astype(np.float64) / 2.0 ** 64).astype(np.float32))
    if rank == 0:
        A = ctx.matrix_from_coo(*m)
        F = ctx.matrix_from_coo_f32(*f)
        comm.send(A, 1)
        comm.send(F, 1)          # float32 tiles travel through the same calls (SURVEY 8e row 3)
        back = comm.recv(1)
        assert back.checksum() == A.checksum()
        assert comm.bcast(A, 0) is A
        Fb = comm.bcast(None, 1)
        assert Fb.is_f32 and Fb.checksum() == F.checksum()
    elif rank == 1:
        A = comm.recv(0)
        F = comm.recv(0)
        assert F.is_f32 and not A.is_f32
        br, bc, v = A.to_coo()
        assert np.array_equal(br, m[3]) and np.array_equal(bc, m[4]) and np.array_equal(v.reshape(-1), m[5].reshape(-1))
        _, _, fv = F.to_coo()
        assert np.array_equal(fv.reshape(-1), f[5].reshape(-1))
        comm.send(A, 0)
        A2 = comm.bcast(None, 0)
        assert A2.checksum() == A.checksum()
        comm.bcast(F, 1)
    else:
        A2 = comm.bcast(None, 0)
        F2 = comm.bcast(None, 1)
        assert F2.is_f32
    # row panels of one product, float32 and uint64, gathered on rank 0 = the whole product
    for mk in ("u64", "f32"):
        X = ctx.matrix_from_coo(*m) if mk == "u64" else ctx.matrix_from_coo_f32(*f)
        bounds = ctx.panel_bounds(X, X, world)
        assert bounds[0] == 0 and bounds[-1] == G and all(bounds[i] <= bounds[i + 1] for i in range(world))
        P = ctx.helper(X, X, panel=(bounds[rank], bounds[rank + 1]))
        full = comm.gather_panels(P, 0)
        if rank == 0:
            whole = ctx.helper(X, X)
            assert full.checksum() == whole.checksum(), mk
            a, b, c = full.to_coo(); a2, b2, c2 = whole.to_coo()
            assert np.array_equal(a, a2) and np.array_equal(b, b2) and np.array_equal(c, c2)
        else:
            assert full is None
    if rank == 0:
        open(out, "w").write("ok")
dist.barrier()
comm.close()
dist.destroy_process_group()
'''


def _run_capi(world, args, port):
    pkg = os.path.dirname(bspgemm.__file__)
    with tempfile.TemporaryDirectory() as d:
        script = os.path.join(d, "w.py")
        open(script, "w").write(CAPI_WORKER)
        out = os.path.join(d, "matrix")
        p = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=%d" % world,
                            "--master-addr", "127.0.0.1", "--master-port", str(port), script, pkg,
                            os.path.join(util.ROOT, "oracle"), out] + args, stdout=subprocess.PIPE, stderr=subprocess.PIPE, timeout=600)
        assert p.returncode == 0, p.stderr.decode()[-3000:]
        return open(out, "rb").read()


@pytest.mark.parametrize("world,n", [(2, 9), (2, 1), (3, 7), (4, 9), (8, 19), (8, 5)])
def test_capi_chain_distributed_matches_rank_tree(world, n):
    """bsp_chain_distributed over NCCL against the oracle's P-rank tree (cu:438-456 + cu:569-571), byte for
    byte through the writer; n < world exercises the 'only rank 0 works' rule (cu:612-616)."""
    if _ngpu() < world:
        pytest.skip("needs %d GPUs" % world)
    k, G = 4, 40
    got = _run_capi(world, ["chain", str(n)], 29621 + world)
    mats = [orc.OMat.from_coo(*bsp_gen.random_matrix(G * k, G * k, k, 0.08, 8100 + i, "edgy")) for i in range(n)]
    with tempfile.TemporaryDirectory() as d:
        orc.chain(mats, ranks=world)[0].prune_zero().write(os.path.join(d, "want"))
        assert got == open(os.path.join(d, "want"), "rb").read()


@pytest.mark.parametrize("world", [2, 3])
def test_capi_transport_send_recv_bcast_gather(world):
    if _ngpu() < world:
        pytest.skip("needs %d GPUs" % world)
    assert _run_capi(world, ["transport"], 29641 + world) == b"ok"


def test_capi_comm_single_rank():
    """One rank: the communicator degenerates to the local chain (runs on the 1-GPU test box)."""
    import ctypes
    k, G, n = 4, 30, 5
    with bspgemm.Context(0) as ctx:
        comm = bspgemm.Comm(ctx, bspgemm.comm_unique_id(), 1, 0)
        host = [bsp_gen.random_matrix(G * k, G * k, k, 0.1, 9100 + i, "edgy") for i in range(n)]
        mats = [ctx.matrix_from_coo(*m) for m in host]
        res = comm.chain(mats)
        want = ctx.helper2(mats)
        assert res.checksum() == want.checksum()
        A = mats[0]
        assert ctx.panel_bounds(A, mats[1], 1) == [0, A.device_view()[0]]
        b4 = ctx.panel_bounds(A, mats[1], 4)
        assert b4[0] == 0 and b4[4] == A.device_view()[0] and b4 == sorted(b4)
        # panels by those bounds add up to the whole product, also by checksum parts
        whole = ctx.helper(A, mats[1])
        body, nn = 0, 0
        for i in range(4):
            P = ctx.helper(A, mats[1], panel=(b4[i], b4[i + 1]))
            body = (body + P.checksum_body()) & 0xFFFFFFFFFFFFFFFF
            nn += P.nnzb
        r, c, kk, _ = whole.info()
        assert bspgemm.checksum_finish(r, c, kk, 8, nn, body) == whole.checksum()
        assert comm.gather_panels(whole, 0).checksum() == whole.checksum()
        comm.close()
