"""bsp_prune_zero_noalloc -- the zero-tile compaction for hosts that must not allocate device
memory (they pre-assigned all of it to their own arena, or cannot take the device-wide stall
of cudaMalloc/cudaFree mid-loop).  The buffers below are created by the test with torch, as
such a host would create them from its own arena; the entry point itself gets only addresses.

Checked against the existing, allocating entry point bsp_prune_zero, bit for bit (integer
data: the tolerance of every parity test here is zero), on two calls in a row that share the
library's built-in scratch: the second call uses a smaller, different matrix, so anything the
first call left behind in the scratch would show up as a wrong result.

Collected last on purpose (file name): a failure here cannot hide another test's result.
"""
import numpy as np
import pytest

import bsp_gen
import bspgemm

pytestmark = pytest.mark.gpu


def matrix_with_zero_tiles(G, k, density, seed, zero_every):
    rows, cols, k, br, bc, v = bsp_gen.random_matrix(G * k, G * k, k, density, seed, "u64")
    v = v.copy()
    v[::zero_every] = 0          # whole tiles of zeros, spread over all rows
    v[1::7, 0, 0] = 0            # tiles with a single zero entry must survive
    return rows, cols, k, br, bc, v


def expected_by_existing_entry_point(ctx, host):
    m = ctx.matrix_from_coo(*host)
    m.prune_zero()
    return m.to_coo()


def run_noalloc(ctx, torch, host):
    k = host[2]
    m = ctx.matrix_from_coo(*host)                      # input buffers (owned by the library here)
    nbr, _, p_rowptr, p_colidx, p_vals = m.device_view()
    n = m.nnzb
    out_rowptr = torch.full((nbr + 1,), -1, dtype=torch.int32, device="cuda")   # caller-owned outputs
    out_colidx = torch.full((max(n, 1),), -1, dtype=torch.int32, device="cuda")
    out_vals = torch.full((max(n, 1) * k * k,), -1, dtype=torch.int64, device="cuda")
    torch.cuda.synchronize()
    kept = ctx.prune_zero_noalloc(k, nbr, n, p_rowptr, p_colidx, p_vals,
                                  out_rowptr.data_ptr(), out_colidx.data_ptr(), out_vals.data_ptr())
    rowptr = out_rowptr.cpu().numpy()
    assert rowptr[0] == 0 and rowptr[-1] == kept and np.all(np.diff(rowptr) >= 0)
    brow = np.repeat(np.arange(nbr, dtype=np.int32), np.diff(rowptr))
    bcol = out_colidx.cpu().numpy()[:kept]
    vals = out_vals.cpu().numpy().view(np.uint64)[:kept * k * k].reshape(kept, k, k)
    return brow, bcol, vals


def test_noalloc_prune_matches_allocating_prune_twice_in_a_row():
    import torch
    with bspgemm.Context(0) as ctx:
        first = matrix_with_zero_tiles(G=160, k=8, density=0.12, seed=9101, zero_every=3)
        second = matrix_with_zero_tiles(G=90, k=4, density=0.2, seed=9102, zero_every=2)
        for host in (first, second):
            want = expected_by_existing_entry_point(ctx, host)
            got = run_noalloc(ctx, torch, host)
            assert got[0].size == want[0].size and got[0].size < host[3].size
            for g, w in zip(got, want):
                assert np.array_equal(g, w)


def test_noalloc_prune_rejects_oversize_and_handles_empty():
    import torch
    with bspgemm.Context(0) as ctx:
        dummy = torch.zeros(8, dtype=torch.int64, device="cuda")
        with pytest.raises(bspgemm.BspError) as e:
            ctx.prune_zero_noalloc(4, 1, (1 << 20) + 1, dummy.data_ptr(), dummy.data_ptr(), dummy.data_ptr(),
                                   dummy.data_ptr(), dummy.data_ptr(), dummy.data_ptr())
        assert e.value.code == 4 and "BSP_NOALLOC_MAX_TILES" in str(e.value)
        rowptr = torch.zeros(6, dtype=torch.int32, device="cuda")
        out = torch.full((6,), -1, dtype=torch.int32, device="cuda")
        kept = ctx.prune_zero_noalloc(4, 5, 0, rowptr.data_ptr(), 0, 0, out.data_ptr(), 0, 0)
        assert kept == 0 and out.cpu().tolist() == [0] * 6
