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
