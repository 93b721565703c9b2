"""Live cross-check of the CPU restatement against the reference's own host build
(oracle/_ref/ref_host_fixed) on seeded random folders.  Skipped where oracle/_ref was not
built (it needs /root/reference at build time; the built binary travels to the GPU box)."""
import os
import tempfile

import pytest

import bsp_gen
import orc

pytestmark = pytest.mark.skipif(not orc.have_ref(), reason="oracle/_ref not built")


@pytest.mark.parametrize("k,G,dens,kind,n,ranks", [
    (2, 9, 0.4, "edgy", 5, 1), (2, 9, 0.4, "edgy", 9, 4), (4, 16, 0.2, "u64", 3, 2),
    (8, 12, 0.25, "edgy", 2, 1), (3, 10, 0.3, "small", 4, 1), (4, 20, 0.15, "edgy", 8, 8)])
def test_restatement_equals_reference_program(k, G, dens, kind, n, ranks):
    mats = [orc.OMat.from_coo(*bsp_gen.random_matrix(G * k, G * k, k, dens, 9000 + 17 * i + k, kind)) for i in range(n)]
    with tempfile.TemporaryDirectory() as d:
        orc.write_folder(d, mats, k)
        ref = orc.run_ref(d, ranks=ranks)
        assert ref["returncode"] == 0
        res = orc.chain(mats, ranks=ranks)[0].prune_zero()
        p = os.path.join(d, "mine")
        res.write(p)
        assert open(p, "rb").read() == ref["matrix"]
        # the reader of the restatement sees what the reference's reader saw
        back = orc.OMat.read(os.path.join(d, "matrix1"), k)
        assert back.nnzb == mats[0].nnzb
