"""Parity of the CUDA path (called through the C ABI) against the CPU oracle and the golden
fixtures made by the reference's own code.  Integer work: the bar is bit-exact."""
import hashlib
import os
import subprocess
import tempfile

import numpy as np
import pytest

import bsp_dist
import bsp_gen
import bspgemm
import orc
import util

pytestmark = pytest.mark.gpu
KATS = util.load_kats()
MAXV = (1 << 64) - 1


@pytest.fixture(scope="module")
def ctx():
    c = bspgemm.Context(0)
    yield c
    c.close()


def dev(ctx, m):
    return ctx.matrix_from_coo(*m)


def assert_same(gpu_mat, oracle_mat):
    br, bc, v = gpu_mat.to_coo()
    rbr, rbc, rv = oracle_mat.coo()
    assert gpu_mat.nnzb == oracle_mat.nnzb
    assert np.array_equal(br, rbr) and np.array_equal(bc, rbc)
    assert np.array_equal(v, rv)
    assert (gpu_mat.rows, gpu_mat.cols) == (oracle_mat.rows, oracle_mat.cols)


@pytest.mark.parametrize("case", KATS, ids=[c["id"] for c in KATS])
def test_golden_cases(ctx, case):
    """Known answers from the reference program, incl. its P-rank association trees."""
    k = case["k"]
    mats = [ctx.matrix_from_coo(r, c, k, br, bc, v) for r, c, br, bc, v in util.case_mats(case)]
    res = bsp_dist.chain_ranks_on_one_context(ctx, mats, case["ranks"])
    if res in mats:  # single matrix: helper2 returns a copy
        res = ctx.helper2([res])
    res.prune_zero()
    br, bc, v = res.to_coo()
    assert util.matrix_text(res.rows, res.cols, k, br, bc, v) == case["matrix_text"]
    with tempfile.TemporaryDirectory() as d:
        p = os.path.join(d, "matrix")
        res.write_text(p)
        assert open(p).read() == case["matrix_text"]


@pytest.mark.parametrize("case", [c for c in KATS if c["multiplying"] is not None], ids=lambda c: c["id"])
def test_golden_tree_log(ctx, case):
    k = case["k"]
    mats = [ctx.matrix_from_coo(r, c, k, br, bc, v) for r, c, br, bc, v in util.case_mats(case)]
    _, log = ctx.helper2(mats, want_log=True)
    assert [list(map(int, ln.split()[1:])) for ln in log.splitlines()] == case["multiplying"]


@pytest.mark.parametrize("k", [1, 2, 3, 4, 5, 7, 8, 12, 16, 32])
@pytest.mark.parametrize("kind", ["edgy", "u64"])
def test_pair_vs_oracle_all_tile_sizes(ctx, k, kind):
    G = 24 if k <= 8 else 8
    a = bsp_gen.random_matrix(G * k, (G + 5) * k, k, 0.2, 40 + k, kind)
    b = bsp_gen.random_matrix((G + 5) * k, (G - 3) * k, k, 0.25, 90 + k, kind)
    st = bspgemm.Stats()
    c = ctx.helper(dev(ctx, a), dev(ctx, b), stats=st)
    ref, nprod = orc.multiply(orc.OMat.from_coo(*a), orc.OMat.from_coo(*b))
    assert st.nprod == nprod and st.nnzb_c == ref.nnzb
    assert_same(c, ref)


@pytest.mark.parametrize("k,G,n,dens", [(4, 40, 6, 0.06), (8, 24, 5, 0.1), (2, 30, 9, 0.1), (4, 64, 16, 0.03)])
def test_chain_vs_oracle(ctx, k, G, n, dens):
    host = [bsp_gen.random_matrix(G * k, G * k, k, dens, 500 + 13 * i, "edgy" if i % 2 else "u64") for i in range(n)]
    res = ctx.helper2([dev(ctx, m) for m in host]).prune_zero()
    ref = orc.chain([orc.OMat.from_coo(*m) for m in host])[0].prune_zero()
    assert_same(res, ref)


@pytest.mark.parametrize("P", [2, 3, 4, 8])
def test_rank_trees_vs_oracle(ctx, P):
    k, G, n = 2, 10, 11
    host = [bsp_gen.random_matrix(G * k, G * k, k, 0.35, 700 + i, "edgy") for i in range(n)]
    res = bsp_dist.chain_ranks_on_one_context(ctx, [dev(ctx, m) for m in host], P).prune_zero()
    ref = orc.chain([orc.OMat.from_coo(*m) for m in host], ranks=P)[0].prune_zero()
    assert_same(res, ref)


@pytest.mark.parametrize("cfg", util.load_cfg1(), ids=lambda c: c["id"])
def test_cfg1_output_file_hash(ctx, cfg):
    """BASELINE.json configs[0]: n=1024, k=4, 5 % -- byte-identical output file."""
    mats = [dev(ctx, bsp_gen.random_matrix(cfg["n"], cfg["n"], cfg["k"], cfg["density"], s, cfg["kind"]))
            for s in cfg["seeds"]]
    res = ctx.helper2(mats).prune_zero()
    assert res.nnzb == cfg["out_nnzb"]
    with tempfile.TemporaryDirectory() as d:
        p = os.path.join(d, "matrix")
        res.write_text(p)
        assert hashlib.sha256(open(p, "rb").read()).hexdigest() == cfg["sha256"]


def test_edge_cases(ctx):
    k = 4
    a = bsp_gen.random_matrix(64, 64, k, 0.3, 1, "u64")
    empty = (64, 64, k, np.zeros(0, np.int32), np.zeros(0, np.int32), np.zeros(0, np.uint64))
    A, E = dev(ctx, a), dev(ctx, empty)
    for x, y in ((A, E), (E, A), (E, E)):
        c = ctx.helper(x, y)
        assert c.nnzb == 0 and c.to_coo()[0].size == 0
        assert c.prune_zero().nnzb == 0
    # inner indices that never match, and A columns beyond B's last block row
    a2 = (64, 64, k, np.array([0, 1], np.int32), np.array([50, 60], np.int32), np.ones(2 * 16, np.uint64))
    b2 = (64, 64, k, np.array([2, 3], np.int32), np.array([0, 1], np.int32), np.ones(2 * 16, np.uint64))
    assert ctx.helper(dev(ctx, a2), dev(ctx, b2)).nnzb == 0
    # ragged: one very long row, many empty rows, max values
    br = np.zeros(300, np.int32)
    bc = np.arange(300, dtype=np.int32) * 3
    v = np.full(300 * 16, MAXV, np.uint64)
    long_a = (1200, 1200, k, br, bc, v)
    b3 = bsp_gen.random_matrix(1200 * 3, 1200, k, 0.05, 5, "edgy")
    c = ctx.helper(dev(ctx, long_a), dev(ctx, b3))
    assert_same(c, orc.multiply(orc.OMat.from_coo(*long_a), orc.OMat.from_coo(*b3))[0])
    # negative coordinates are rejected with a message, not a crash
    with pytest.raises(bspgemm.BspError) as e:
        ctx.matrix_from_coo(8, 8, k, [-1], [0], np.ones(16, np.uint64))
    assert e.value.code == 1
    # tile-edge mismatch
    with pytest.raises(bspgemm.BspError):
        ctx.helper(A, dev(ctx, bsp_gen.random_matrix(64, 64, 8, 0.3, 1, "u64")))


def test_duplicates_and_unsorted_ingest(ctx):
    k = 2
    br = np.array([3, 0, 3, 1, 0, 3], np.int32)
    bc = np.array([1, 2, 1, 1, 2, 0], np.int32)
    v = np.arange(6 * 4, dtype=np.uint64)
    m = ctx.matrix_from_coo(8, 8, k, br, bc, v)
    o = orc.OMat.from_coo(8, 8, k, br, bc, v)
    assert_same(m, o)
    assert m.nnzb == 4


def test_panels_concatenate_to_full_product(ctx):
    k, G = 8, 96
    a = bsp_gen.random_matrix(G * k, G * k, k, 0.08, 21, "u64")
    b = bsp_gen.random_matrix(G * k, G * k, k, 0.08, 22, "u64")
    A, B = dev(ctx, a), dev(ctx, b)
    full = ctx.helper(A, B).to_coo()
    bounds = [0, 7, 40, 41, G]
    parts = [ctx.helper(A, B, panel=(bounds[i], bounds[i + 1])).to_coo() for i in range(4)]
    for j in range(3):
        assert np.array_equal(np.concatenate([p[j] for p in parts]), full[j])


def test_multiply_host_single_call(ctx):
    k = 4
    a = bsp_gen.random_matrix(256, 256, k, 0.1, 31, "edgy")
    b = bsp_gen.random_matrix(256, 256, k, 0.1, 32, "edgy")
    st = bspgemm.Stats()
    br, bc, v = ctx.multiply_host(k, (a[0], a[1], a[3], a[4], a[5]), (b[0], b[1], b[3], b[4], b[5]), stats=st)
    ref = orc.multiply(orc.OMat.from_coo(*a), orc.OMat.from_coo(*b))[0]
    rbr, rbc, rv = ref.coo()
    assert np.array_equal(br, rbr) and np.array_equal(bc, rbc) and np.array_equal(v, rv)
    assert st.launches > 0


def test_a4_cli_matches_reference_files(ctx):
    exe = os.path.join(os.path.dirname(bspgemm.__file__), "a4")
    import make_golden_inputs
    for case in KATS:
        with tempfile.TemporaryDirectory() as d:
            make_golden_inputs.write_raw_folder(os.path.join(d, "in"), util.case_mats(case), case["k"])
            p = subprocess.run([exe, os.path.join(d, "in"), "--ranks", str(case["ranks"]), "--devices", "1"], cwd=d,
                               stdout=subprocess.PIPE, stderr=subprocess.PIPE, timeout=120)
            assert p.returncode == 0, p.stderr.decode()
            assert open(os.path.join(d, "matrix")).read() == case["matrix_text"], case["id"]
            out = p.stdout.decode()
            assert "time taken" in out
            if case["multiplying"] is not None:
                got = [list(map(int, ln.split()[1:])) for ln in out.splitlines() if ln.startswith("multiplying")]
                assert got == case["multiplying"]
    with tempfile.TemporaryDirectory() as d:
        p = subprocess.run([exe, os.path.join(d, "missing")], cwd=d, stdout=subprocess.PIPE, stderr=subprocess.PIPE)
        assert p.returncode == 1 and b"Cannot open size file!" in p.stderr


# ---- size-independent properties at BASELINE.json's full sizes -------------------------------

def block_identity(G, k):
    br = np.arange(G, dtype=np.int32)
    v = np.tile(np.eye(k, dtype=np.uint64).reshape(-1), G)
    return G * k, G * k, k, br, br.copy(), v


def test_full_size_identity_and_sampled_tiles(ctx):
    """configs[1] shape: n=65536, k=8, 1 % density. A*I == A == I*A bit for bit (entries are in
    [1, 2^64-2], so the 2^64-1 edge never fires), and randomly sampled tiles of A*B equal a
    numpy evaluation of the reference arithmetic."""
    n, k, d = 65536, 8, 0.01
    G = n // k
    a = bsp_gen.random_matrix(n, n, k, d, 1234, "u64")
    A = dev(ctx, a)
    I = dev(ctx, block_identity(G, k))
    for prod in (ctx.helper(A, I), ctx.helper(I, A)):
        br, bc, v = prod.to_coo()
        assert np.array_equal(br, a[3]) and np.array_equal(bc, a[4]) and np.array_equal(v, a[5])
    b = bsp_gen.random_matrix(n, n, k, d, 1235, "u64")
    st = bspgemm.Stats()
    Cm = ctx.helper(A, dev(ctx, b), stats=st)
    br, bc, v = Cm.to_coo()
    # pattern: sortedness + counts
    key = br.astype(np.int64) * G + bc
    assert np.all(np.diff(key) > 0)
    rowlen_b = np.bincount(b[3], minlength=G)
    assert st.nprod == int(rowlen_b[a[4]].sum())
    # sampled tiles, evaluated with python ints in reference order
    rng = np.random.default_rng(7)
    a_key = a[3].astype(np.int64) * G + a[4]
    b_key = b[3].astype(np.int64) * G + b[4]
    for t in rng.integers(0, br.size, 40):
        i, c = int(br[t]), int(bc[t])
        js = a[4][np.searchsorted(a[3], i):np.searchsorted(a[3], i + 1)]
        acc = [[0] * k for _ in range(k)]
        hit = 0
        for j in js:
            q = np.searchsorted(b_key, int(j) * G + c)
            if q < b_key.size and b_key[q] == int(j) * G + c:
                hit += 1
                At = a[5][np.searchsorted(a_key, i * G + int(j))]
                Bt = b[5][q]
                for r in range(k):
                    for cc in range(k):
                        s = acc[r][cc]
                        for x in range(k):
                            p = (int(At[r, x]) * int(Bt[x, cc])) & MAXV
                            p = 0 if p == MAXV else p
                            s = (s + p) & MAXV
                            s = 0 if s == MAXV else s
                        acc[r][cc] = s
        assert hit >= 1
        assert np.array_equal(np.array(acc, dtype=np.uint64), v[t])


def test_full_size_chain_is_associative_without_edge_hits(ctx):
    """n=16384, k=4 (configs[2] shape), short chain: with entries that never produce 2^64-1 the
    arithmetic is the ring Z/2^64, so every association tree gives the same bits."""
    n, k = 16384, 4
    G = n // k
    mats = [dev(ctx, bsp_gen.random_matrix(n, n, k, 2.0 / G, 4000 + i, "u64")) for i in range(8)]
    r1 = ctx.helper2(mats).to_coo()
    r2 = bsp_dist.chain_ranks_on_one_context(ctx, mats, 4).to_coo()
    left = mats[0]
    for m in mats[1:]:
        left = ctx.helper(left, m)
    r3 = left.to_coo()
    for x, y, z in zip(r1, r2, r3):
        assert np.array_equal(x, y) and np.array_equal(x, z)


@pytest.mark.parametrize("k,kind", [(4, "edgy"), (8, "edgy"), (4, "u64"), (8, "u64")])
def test_dense_patterns_many_pairs_per_tile(ctx, k, kind):
    """Every output tile folds dozens of pairs (a dense chain level); with 'edgy' values the
    2^64-1 detector of the fast path fires on most rows and the exact refold takes over."""
    G = 37
    a = bsp_gen.random_matrix(G * k, G * k, k, 0.9, 61, kind)
    b = bsp_gen.random_matrix(G * k, G * k, k, 0.85, 62, kind)
    st = bspgemm.Stats()
    c = ctx.helper(dev(ctx, a), dev(ctx, b), stats=st)
    assert_same(c, orc.multiply(orc.OMat.from_coo(*a), orc.OMat.from_coo(*b))[0])
    assert st.numeric_path == k
    if kind == "edgy":
        assert st.exact_refolds > 0


def test_skewed_pairs_per_tile(ctx):
    """One block row of A hits every row of B (thousands of pairs for its tiles), the others hit
    one or two: the pair-space work split must still give every tile to exactly one group."""
    k, G = 8, 200
    rng = np.random.default_rng(3)
    br = np.concatenate([np.zeros(G, np.int32), rng.integers(1, G, 300).astype(np.int32)])
    bc = np.concatenate([np.arange(G, dtype=np.int32), rng.integers(0, G, 300).astype(np.int32)])
    va = bsp_gen.values(71, br.size * k * k, "u64")
    a = (G * k, G * k, k, br, bc, va)
    b = bsp_gen.random_matrix(G * k, G * k, k, 0.3, 72, "u64")
    c = ctx.helper(dev(ctx, a), dev(ctx, b))
    assert_same(c, orc.multiply(orc.OMat.from_coo(*a), orc.OMat.from_coo(*b))[0])


def test_symbolic_row_classes_and_generic_numeric(ctx):
    """Force rows through every capacity class of the symbolic fill (shared-memory classes and
    the global-memory variant) and through the any-k exact numeric kernel; all must agree."""
    import os
    k = 4
    a = bsp_gen.random_matrix(160 * k, 160 * k, k, 0.15, 81, "edgy")
    b = bsp_gen.random_matrix(160 * k, 160 * k, k, 0.2, 82, "edgy")
    ref = orc.multiply(orc.OMat.from_coo(*a), orc.OMat.from_coo(*b))[0]
    try:
        for caps in ("16,64,256", "4,8,16", "100000,100000,100000"):
            os.environ["BSP_SYM_CAPS"] = caps
            assert_same(ctx.helper(dev(ctx, a), dev(ctx, b)), ref)
            # rows beyond the classes: the per-warp-range kernel (default) and the global-memory kernel
            os.environ["BSP_SYM_BIG"] = "0"
            assert_same(ctx.helper(dev(ctx, a), dev(ctx, b)), ref)
            os.environ.pop("BSP_SYM_BIG", None)
        os.environ["BSP_FORCE_GENERIC"] = "1"
        with bspgemm.Context(0) as c2:
            st = bspgemm.Stats()
            assert_same(c2.helper(c2.matrix_from_coo(*a), c2.matrix_from_coo(*b), stats=st), ref)
            assert st.numeric_path == 0
    finally:
        os.environ.pop("BSP_SYM_CAPS", None)
        os.environ.pop("BSP_SYM_BIG", None)
        os.environ.pop("BSP_FORCE_GENERIC", None)


@pytest.mark.parametrize("k,G,da,db,kind", [(4, 160, 0.15, 0.2, "edgy"), (8, 700, 0.05, 0.05, "u64"), (4, 300, 0.4, 0.5, "edgy"),
                                              (2, 2100, 0.02, 0.03, "edgy")])
def test_symbolic_medium_rows_against_oracle(ctx, k, G, da, db, kind):
    """Rows of a few hundred to a few thousand tile products (the shared-memory fill classes at the sizes
    of a single large product): B rows longer than a warp, many pairs per tile, and the 2^64-1 edge
    values whose fold order would expose a mis-ordered pair list; with and without the bit-matrix fill."""
    import os
    a = bsp_gen.random_matrix(G * k, G * k, k, da, 181, kind)
    b = bsp_gen.random_matrix(G * k, G * k, k, db, 182, kind)
    ref = orc.multiply(orc.OMat.from_coo(*a), orc.OMat.from_coo(*b))[0]
    try:
        got = ctx.helper(dev(ctx, a), dev(ctx, b))
        assert_same(got, ref)
        os.environ["BSP_SYM_NO_DENSE"] = "1"
        assert ctx.helper(dev(ctx, a), dev(ctx, b)).checksum() == got.checksum()
    finally:
        os.environ.pop("BSP_SYM_NO_DENSE", None)


def test_concat_panels(ctx):
    k, G = 4, 120
    a = bsp_gen.random_matrix(G * k, G * k, k, 0.06, 91, "edgy")
    b = bsp_gen.random_matrix(G * k, G * k, k, 0.06, 92, "edgy")
    A, B = dev(ctx, a), dev(ctx, b)
    full = ctx.helper(A, B).to_coo()
    bounds = [0, 1, 50, 50, 99, G]
    parts = [ctx.helper(A, B, panel=(bounds[i], bounds[i + 1])) for i in range(5)]
    got = ctx.concat_panels(parts).to_coo()
    for x, y in zip(got, full):
        assert np.array_equal(x, y)


def test_text_io_roundtrip_large(ctx):
    """Files above 1 MB are parsed on several host threads; the writer formats in parallel."""
    k = 8
    a = bsp_gen.random_matrix(400 * k, 400 * k, k, 0.05, 123, "u64")   # ~8000 tiles, ~10 MB of text
    A = dev(ctx, a)
    with tempfile.TemporaryDirectory() as d:
        p = os.path.join(d, "matrix1")
        A.write_text(p)
        assert os.path.getsize(p) > (1 << 20)
        ref = orc.OMat.from_coo(*a)
        q = os.path.join(d, "ref")
        ref.write(q)
        assert open(p, "rb").read() == open(q, "rb").read()
        back = ctx.read_text(p, k)
        assert_same(back, ref)
        # truncated file -> error code, not a crash
        with open(p, "rb") as f:
            data = f.read()
        with open(p, "wb") as f:
            f.write(data[: len(data) // 2])
        with pytest.raises(bspgemm.BspError) as e:
            ctx.read_text(p, k)
        assert e.value.code == 5


def test_chain_levels_run_as_stacked_products(ctx):
    """A tree level with several small products is executed as one stacked product
    (BSP_NO_BATCH=1 turns that off): rectangular factors of different sizes, inner indices
    without partner rows, empty factors -- all must equal the oracle and the unbatched run."""
    import os
    k = 4
    dims = [10, 14, 9, 20, 6, 11, 11, 3, 17, 8, 12]   # factor i is dims[i] x dims[i+1] blocks
    host = []
    for i in range(len(dims) - 1):
        dens = 0.0 if i == 6 else 0.45                  # factor 6 is empty
        m = bsp_gen.random_matrix(dims[i] * k, dims[i + 1] * k, k, dens, 600 + i, "edgy")
        host.append(m)
    # an A column index beyond its partner's last block row
    r, c, kk, br, bc, v = host[2]
    host[2] = (r, c, kk, np.append(br, 0).astype(np.int32), np.append(bc, 500).astype(np.int32),
               np.concatenate([v, np.ones((1, k, k), np.uint64)]))
    ref = orc.chain([orc.OMat.from_coo(*m) for m in host])[0]
    st = bspgemm.Stats()
    res = ctx.helper2([dev(ctx, m) for m in host], stats=st)
    assert_same(res, ref)
    try:
        os.environ["BSP_NO_BATCH"] = "1"
        with bspgemm.Context(0) as c2:
            st2 = bspgemm.Stats()
            res2 = c2.helper2([c2.matrix_from_coo(*m) for m in host], stats=st2)
            assert_same(res2, ref)
            assert st2.nprod == st.nprod and st2.launches > st.launches
    finally:
        os.environ.pop("BSP_NO_BATCH", None)


def test_multiply_host_pipelined_download(ctx):
    """Above 65536 tiles in A the host-to-host call computes block-row panels and downloads one
    while the next is computed; the concatenated result must be the sorted product."""
    k, G = 4, 2048
    a = bsp_gen.random_matrix(G * k, G * k, k, 0.02, 41, "u64")
    b = bsp_gen.random_matrix(G * k, G * k, k, 0.004, 42, "u64")
    assert a[3].size >= 65536
    st = bspgemm.Stats()
    br, bc, v = ctx.multiply_host(k, (a[0], a[1], a[3], a[4], a[5]), (b[0], b[1], b[3], b[4], b[5]), stats=st)
    ref, nprod = orc.multiply(orc.OMat.from_coo(*a), orc.OMat.from_coo(*b), threads=8)
    rbr, rbc, rv = ref.coo()
    assert st.nprod == nprod and st.nnzb_c == ref.nnzb
    assert np.array_equal(br, rbr) and np.array_equal(bc, rbc) and np.array_equal(v, rv)


def test_products_above_the_pair_budget_run_in_row_panels(monkeypatch):
    """A product whose pair lists (8 B per tile product) exceed the context's budget -- dense x
    dense chain levels hold G^3 of them -- is computed in row panels and concatenated; bit-exact
    against the oracle and the unsplit run, same statistics.  BSP_PAIR_BUDGET makes the split
    happen at test sizes; a single row above the budget still gets a panel of its own."""
    k, G = 4, 48
    dense = [bsp_gen.random_matrix(G * k, G * k, k, 1.0, 60 + i, "edgy" if i else "u64") for i in range(2)]
    sparse = [bsp_gen.random_matrix(300 * k, 300 * k, k, 0.05, 70 + i, "u64") for i in range(2)]
    hub = list(bsp_gen.random_matrix(200 * k, 200 * k, k, 0.02, 80, "u64"))
    # block row 7 becomes full (duplicates: the last one wins): one row far above the budget
    hub[3] = np.concatenate([hub[3], np.full(200, 7, np.int32)])
    hub[4] = np.concatenate([hub[4], np.arange(200, dtype=np.int32)])
    hub[5] = np.concatenate([hub[5].reshape(-1, k, k), bsp_gen.values(81, 200 * k * k, "u64").reshape(-1, k, k)])
    cases = [(dense[0], dense[1]), (sparse[0], sparse[1]), (tuple(hub), tuple(hub))]
    plain = bspgemm.Context(0)
    monkeypatch.setenv("BSP_PAIR_BUDGET", "3000")
    split = bspgemm.Context(0)
    monkeypatch.delenv("BSP_PAIR_BUDGET")
    try:
        for a, b in cases:
            want, nprod = orc.multiply(orc.OMat.from_coo(*a), orc.OMat.from_coo(*b))
            s1, s2 = bspgemm.Stats(), bspgemm.Stats()
            c1 = plain.helper(plain.matrix_from_coo(*a), plain.matrix_from_coo(*b), stats=s1)
            c2 = split.helper(split.matrix_from_coo(*a), split.matrix_from_coo(*b), stats=s2)
            assert_same(c1, want)
            assert_same(c2, want)
            assert s1.nprod == s2.nprod == nprod and s1.nnzb_c == s2.nnzb_c == want.nnzb
            assert s2.launches > s1.launches        # it really ran as several panels
            assert c1.checksum() == c2.checksum()
        # a chain through the split context (levels of a densifying chain) equals the oracle's tree
        mats = [bsp_gen.random_matrix(40 * k, 40 * k, k, 0.15, 90 + i, "u64") for i in range(6)]
        want, _ = orc.chain([orc.OMat.from_coo(*m) for m in mats])
        assert_same(split.helper2([split.matrix_from_coo(*m) for m in mats]), want)
    finally:
        plain.close(); split.close()


def test_very_long_rows_multi_launch_fill(ctx, monkeypatch):
    """Rows above hv_min tile products (BSP_HV_MIN_P lowers it to test sizes) go through the
    multi-launch path: global bitmaps, atomic counts and cursors, then a sort of every tile's pairs
    by inner index.  Covers tile lists of 1..32 pairs (warp sort), hundreds (shared-memory sort) and
    > 8192 (global scratch sort); 'edgy' values make any pair out of order visible (the fold is order
    dependent).  The one-CTA kernel and the bit-matrix kernel (default for these rows) must give the same bits."""
    k, G = 4, 600
    rows, cols, _, br, bc, v = bsp_gen.random_matrix(G * k, G * k, k, 0.01, 500, "edgy")
    hub_r = np.repeat(np.array([3, 77, 599], np.int32), G)
    hub_c = np.tile(np.arange(G, dtype=np.int32), 3)
    a = (rows, cols, k, np.concatenate([br, hub_r]), np.concatenate([bc, hub_c]),
         np.concatenate([v.reshape(-1, k, k), bsp_gen.values(501, 3 * G * k * k, "edgy").reshape(-1, k, k)]))
    b = bsp_gen.random_matrix(G * k, G * k, k, 0.3, 502, "edgy")
    want, _ = orc.multiply(orc.OMat.from_coo(*a), orc.OMat.from_coo(*b), threads=8)
    s0, s1 = bspgemm.Stats(), bspgemm.Stats()
    assert_same(ctx.helper(dev(ctx, a), dev(ctx, b)), want)      # default: the hub rows fit the bit-matrix fill
    # one output tile with 9000 pairs: a full row of A (k = 1) times two full columns of B
    n = 9000
    a1 = (1, n, 1, np.zeros(n, np.int32), np.arange(n, dtype=np.int32), bsp_gen.values(510, n, "edgy"))
    b1 = (n, 2, 1, np.repeat(np.arange(n, dtype=np.int32), 2), np.tile(np.array([0, 1], np.int32), n),
          bsp_gen.values(511, 2 * n, "edgy"))
    want1, nprod = orc.multiply(orc.OMat.from_coo(*a1), orc.OMat.from_coo(*b1))
    assert nprod == 2 * n
    assert_same(ctx.helper(dev(ctx, a1), dev(ctx, b1)), want1)
    monkeypatch.setenv("BSP_SYM_NO_DENSE", "1")
    one_cta = ctx.helper(dev(ctx, a), dev(ctx, b), stats=s0)
    assert_same(one_cta, want)
    assert_same(ctx.helper(dev(ctx, a1), dev(ctx, b1)), want1)
    monkeypatch.setenv("BSP_HV_MIN_P", "10000")
    multi = ctx.helper(dev(ctx, a), dev(ctx, b), stats=s1)
    assert_same(multi, want)
    assert s1.launches > s0.launches                 # the three hub rows (108 000 products each) took the other path
    assert_same(ctx.helper(dev(ctx, a1), dev(ctx, b1)), want1)


def test_banded_chain_bit_matrix_fill(monkeypatch):
    """The chain workload of bench.py at test size: banded factors, every output tile of the upper
    levels folds tens of pairs, so the rows go through the bit-matrix fill (both sizes of it: the
    last levels need more than 24 KB).  Against the oracle's tree, and against the same run with the
    bit-matrix classes switched off (BSP_SYM_NO_DENSE=1)."""
    k, G, n = 4, 512, 32
    mats = []
    for i in range(n):
        br, bc = bsp_gen.banded_pattern(G, 4, 0.5, 7000 + i)
        mats.append((G * k, G * k, k, br, bc, bsp_gen.values(7000 + i, br.size * k * k, "edgy" if i % 5 == 0 else "u64")))
    want, nprod = orc.chain([orc.OMat.from_coo(*m) for m in mats])
    with bspgemm.Context(0) as c1:
        st = bspgemm.Stats()
        got = c1.helper2([c1.matrix_from_coo(*m) for m in mats], stats=st)
        assert_same(got, want)
        assert st.nprod == nprod
        sums = got.checksum()
        # the top product alone, whole and as panels
        left, right = c1.helper2([c1.matrix_from_coo(*m) for m in mats[:16]]), c1.helper2([c1.matrix_from_coo(*m) for m in mats[16:]])
        top = c1.helper(left, right)
        assert top.checksum() == sums
        parts = [c1.helper(left, right, panel=(lo, hi)) for lo, hi in ((0, 100), (100, 101), (101, G))]
        assert c1.concat_panels(parts).checksum() == sums
    monkeypatch.setenv("BSP_SYM_NO_DENSE", "1")
    with bspgemm.Context(0) as c2:
        assert c2.helper2([c2.matrix_from_coo(*m) for m in mats]).checksum() == sums


def _rmat(scale, k, seed, kind):
    br, bc = bsp_gen.rmat_pattern(scale, 16, seed)
    G = 1 << scale
    return G * k, G * k, k, br, bc, bsp_gen.values(seed, br.size * k * k, kind).reshape(-1, k, k)


def test_rmat_small_against_oracle_and_panels(ctx):
    """BASELINE.json configs[3] at test size (R-MAT scale 9, edge factor 16: hub rows with thousands
    of tile products next to rows with a handful): whole product against the oracle, and 8 block-row
    panels balanced by tile products, concatenated, against the whole product."""
    k, scale = 8, 9
    G = 1 << scale
    a, b = _rmat(scale, k, 11, "edgy"), _rmat(scale, k, 12, "u64")
    want, nprod = orc.multiply(orc.OMat.from_coo(*a), orc.OMat.from_coo(*b), threads=8)
    A, B = dev(ctx, a), dev(ctx, b)
    st = bspgemm.Stats()
    full = ctx.helper(A, B, stats=st)
    assert_same(full, want)
    assert st.nprod == nprod
    rowlen_b = np.bincount(b[3], minlength=G)
    work = np.bincount(a[3], weights=rowlen_b[a[4]], minlength=G)
    bounds = bsp_dist.balanced_panels(work, 8)
    parts = [ctx.helper(A, B, panel=(bounds[i], bounds[i + 1])) for i in range(8)]
    assert ctx.concat_panels(parts).checksum() == full.checksum()


def test_rmat_full_size_panels_and_sampled_tiles(ctx):
    """configs[3] in full (scale 15, n=262144, k=8, 1.47e8 tile products, 5.8e7 output tiles): the
    product computed whole equals the 8 balanced panels concatenated (content checksum), the hub rows
    take the multi-launch fill, and sampled output tiles -- the longest pair lists among them -- equal
    a Python-int evaluation of the reference arithmetic."""
    k, scale = 8, 15
    G = 1 << scale
    a, b = _rmat(scale, k, 11, "u64"), _rmat(scale, k, 12, "u64")
    A, B = dev(ctx, a), dev(ctx, b)
    st = bspgemm.Stats()
    full = ctx.helper(A, B, stats=st)
    rowlen_b = np.bincount(b[3], minlength=G)
    work = np.bincount(a[3], weights=rowlen_b[a[4]], minlength=G)
    assert st.nprod == int(work.sum()) and work.max() > (1 << 18)   # at least one row above hv_min
    whole = full.checksum()
    bounds = bsp_dist.balanced_panels(work, 8)
    parts = [ctx.helper(A, B, panel=(bounds[i], bounds[i + 1])) for i in range(8)]
    cat = ctx.concat_panels(parts)
    assert cat.nnzb == full.nnzb and cat.checksum() == whole
    del parts, cat
    # tiles of the heaviest row and of random rows; duplicates of the generator were resolved on ingest
    # (last wins), so take the operands back from the device
    abr, abc, av = A.to_coo()
    bbr, bbc, bv = B.to_coo()
    a_key = abr.astype(np.int64) * G + abc
    b_key = bbr.astype(np.int64) * G + bbc
    nbr, nbc, rp_ptr, ci_ptr, _ = full.device_view()
    import torch
    rowptr = torch.empty(nbr + 1, dtype=torch.int32, device="cuda:0")
    colidx = torch.empty(full.nnzb, dtype=torch.int32, device="cuda:0")
    vals = torch.empty(full.nnzb * k * k, dtype=torch.int64, device="cuda:0")
    full.export_device(rowptr.data_ptr(), colidx.data_ptr(), vals.data_ptr())
    rp = rowptr.cpu().numpy()
    hub = int(np.argmax(work))
    rng = np.random.default_rng(3)
    picks = [(hub, int(x)) for x in rng.integers(rp[hub], rp[hub + 1], 6)]
    picks += [(int(np.searchsorted(rp, t, side="right") - 1), int(t)) for t in rng.integers(0, full.nnzb, 10)]
    for i, t in picks:
        c = int(colidx[t].item())
        got = vals[t * k * k:(t + 1) * k * k].cpu().numpy().view(np.uint64).reshape(k, k)
        js = abc[np.searchsorted(abr, i):np.searchsorted(abr, i + 1)]
        acc = [[0] * k for _ in range(k)]
        for j in js:
            q = np.searchsorted(b_key, int(j) * G + c)
            if q < b_key.size and b_key[q] == int(j) * G + c:
                At, Bt = av[np.searchsorted(a_key, i * G + int(j))], bv[q]
                for r in range(k):
                    for cc in range(k):
                        s = acc[r][cc]
                        for x in range(k):
                            p = (int(At[r, x]) * int(Bt[x, cc])) & MAXV
                            p = 0 if p == MAXV else p
                            s = (s + p) & MAXV
                            s = 0 if s == MAXV else s
                        acc[r][cc] = s
        assert np.array_equal(np.array(acc, dtype=np.uint64), got), (i, c)
    del rowptr, colidx, vals
    torch.cuda.empty_cache()
