"""Data formats either side of the path: the binary container, the
device-side content checksum and the multi-file reader that overlaps parsing with upload.  None
of them has a counterpart in the reference; the checks are against a numpy restatement of the
checksum definition, against the text reader/writer (itself byte-exact against the reference's
writer, test_gpu_parity.py) and against the oracle."""
import os
import struct
import tempfile

import numpy as np
import pytest

import bsp_gen
import bspgemm
import orc
import util

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    c = bspgemm.Context(0)
    yield c
    c.close()


def coo_equal(a, b):
    return all(np.array_equal(x, y) for x, y in zip(a.to_coo(), b.to_coo())) and a.info() == b.info()


@pytest.mark.parametrize("k,G,d,kind", [(8, 300, 0.05, "u64"), (4, 200, 0.1, "edgy"), (3, 50, 0.3, "u64"),
                                        (1, 64, 0.5, "small"), (16, 40, 0.2, "u64"), (5, 33, 0.2, "u64")])
def test_checksum_matches_numpy_restatement(ctx, k, G, d, kind):
    rows, cols, _, br, bc, v = bsp_gen.random_matrix(G * k, G * k + k, k, d, 77 + k, kind)
    M = ctx.matrix_from_coo(rows, cols, k, br, bc, v)
    assert M.checksum() == util.checksum_numpy(rows, cols, k, br, bc, v)


def test_checksum_float_and_empty(ctx):
    rng = np.random.default_rng(5)
    k, G = 32, 20
    br, bc = bsp_gen.bernoulli_pattern(G, G, 0.2, 9)
    v = rng.uniform(-1, 1, (br.size, k, k)).astype(np.float32)
    M = ctx.matrix_from_coo_f32(G * k, G * k, k, br, bc, v)
    assert M.checksum() == util.checksum_numpy(G * k, G * k, k, br, bc, v)
    E = ctx.matrix_from_coo(16, 16, 4, np.zeros(0, np.int32), np.zeros(0, np.int32), np.zeros(0, np.uint64))
    assert E.checksum() == util.checksum_numpy(16, 16, 4, [], [], np.zeros(0, np.uint64))
    E2 = ctx.matrix_from_coo(16, 20, 4, np.zeros(0, np.int32), np.zeros(0, np.int32), np.zeros(0, np.uint64))
    assert E.checksum() != E2.checksum()  # the header is part of the content


def test_checksum_sees_what_changes_and_ignores_what_does_not(ctx):
    k, G = 8, 120
    rows, cols, _, br, bc, v = bsp_gen.random_matrix(G * k, G * k, k, 0.1, 4242, "u64")
    base = ctx.matrix_from_coo(rows, cols, k, br, bc, v).checksum()
    # input order is not content
    perm = np.random.default_rng(1).permutation(br.size)
    assert ctx.matrix_from_coo(rows, cols, k, br[perm], bc[perm], v.reshape(-1, k, k)[perm]).checksum() == base
    # one bit of one value, the top bit included
    for bit in (0, 31, 32, 63):
        w = v.copy().reshape(-1)
        w[12345 % w.size] ^= np.uint64(1 << bit)
        assert ctx.matrix_from_coo(rows, cols, k, br, bc, w).checksum() != base
    # two values swapped inside a tile; a tile moved to a free position
    w = v.copy().reshape(-1, k, k)
    w[3, 0, 0], w[3, 0, 1] = v.reshape(-1, k, k)[3, 0, 1], v.reshape(-1, k, k)[3, 0, 0]
    assert ctx.matrix_from_coo(rows, cols, k, br, bc, w).checksum() != base
    bc2 = bc.copy()
    used = set(zip(br.tolist(), bc.tolist()))
    c = next(c for c in range(G) if (int(br[0]), c) not in used)
    bc2[0] = c
    assert ctx.matrix_from_coo(rows, cols, k, br, bc2, v).checksum() != base
    # a product computed whole and by row panels has the same checksum; an all-zero tile counts until pruned
    A = ctx.matrix_from_coo(rows, cols, k, br, bc, v)
    full = ctx.helper(A, A)
    parts = [ctx.helper(A, A, panel=(lo, hi)) for lo, hi in ((0, 17), (17, 80), (80, G))]
    assert ctx.concat_panels(parts).checksum() == full.checksum()
    z = v.copy().reshape(-1, k, k)
    z[7] = 0
    Z = ctx.matrix_from_coo(rows, cols, k, br, bc, z)
    before = Z.checksum()
    Z.prune_zero()
    assert Z.nnzb == br.size - 1 and Z.checksum() != before
    keep = np.arange(br.size) != 7
    assert Z.checksum() == util.checksum_numpy(rows, cols, k, br[keep], bc[keep], z[keep])


@pytest.mark.parametrize("k,G,d", [(8, 700, 0.05), (4, 100, 0.1), (3, 20, 0.5)])
def test_binary_roundtrip(ctx, k, G, d):
    """Small files (k=8: 12.5 MB, k=3: one piece); the multi-piece path is covered further down."""
    m = bsp_gen.random_matrix(G * k, G * k, k, d, 31 + k, "u64")
    M = ctx.matrix_from_coo(*m)
    with tempfile.TemporaryDirectory() as dd:
        p = os.path.join(dd, "m.bin")
        M.write_bin(p)
        size = os.path.getsize(p)
        assert size == 64 + M.nnzb * (8 + 8 * k * k)
        with open(p, "rb") as f:
            h = f.read(64)
        magic, rows, cols, kk, esize, nnzb, csum = struct.unpack("<8siiiiqQ", h[:40])
        assert (magic, rows, cols, kk, esize, nnzb) == (b"BSPMAT01", m[0], m[1], k, 8, M.nnzb)
        assert csum == M.checksum() == util.checksum_numpy(m[0], m[1], k, m[3], m[4], m[5])
        back = ctx.read_bin(p)
        assert coo_equal(back, M)
        # the text writer agrees (so the binary file holds what the reference's `matrix` file would)
        t1, t2 = os.path.join(dd, "a.txt"), os.path.join(dd, "b.txt")
        M.write_text(t1); back.write_text(t2)
        assert open(t1, "rb").read() == open(t2, "rb").read()
        # a flipped payload byte is caught by the checksum; a cut file and a foreign file are refused
        with open(p, "r+b") as f:
            f.seek(size - 5)
            b = f.read(1)
            f.seek(size - 5)
            f.write(bytes([b[0] ^ 0x40]))
        with pytest.raises(bspgemm.BspError) as e:
            ctx.read_bin(p)
        assert e.value.code == 5 and "checksum" in str(e.value)
        with open(p, "r+b") as f:
            f.truncate(size // 2)
        with pytest.raises(bspgemm.BspError) as e:
            ctx.read_bin(p)
        assert e.value.code == 5
        with pytest.raises(bspgemm.BspError) as e:
            ctx.read_bin(t1)
        assert e.value.code == 5
        with pytest.raises(bspgemm.BspError) as e:
            ctx.read_bin(os.path.join(dd, "missing"))
        assert e.value.code == 5


def test_binary_roundtrip_multi_piece_and_float(ctx):
    """far more than one 16 MiB staging piece: the double-buffered staging path wraps around several times."""
    k, G = 16, 260
    m = bsp_gen.random_matrix(G * k, G * k, k, 0.7, 99, "u64")     # ~47 000 tiles * 2 KiB = 97 MB
    M = ctx.matrix_from_coo(*m)
    assert M.nnzb * k * k * 8 > 80 << 20
    rng = np.random.default_rng(3)
    fr, fc = bsp_gen.bernoulli_pattern(30, 30, 0.3, 2)
    F = ctx.matrix_from_coo_f32(30 * 32, 30 * 32, 32, fr, fc, rng.uniform(-1, 1, (fr.size, 32, 32)).astype(np.float32))
    with tempfile.TemporaryDirectory() as dd:
        p, q = os.path.join(dd, "m.bin"), os.path.join(dd, "f.bin")
        M.write_bin(p)
        assert coo_equal(ctx.read_bin(p), M)
        F.write_bin(q)
        back = ctx.read_bin(q)
        assert back.is_f32 and coo_equal(back, F)


def test_read_text_many_equals_one_by_one(ctx):
    """More files than staging slots, different sizes (one above 1 MB so that it is parsed in
    pieces when threads are left over), an empty matrix, duplicates (last wins, cu:383)."""
    k = 4
    mats = []
    for i in range(40):
        G = 30 + 7 * i
        mats.append(bsp_gen.random_matrix(G * k, (G + 3) * k, k, 0.08, 900 + i, "u64" if i % 3 else "edgy"))
    mats.append((16, 16, k, np.zeros(0, np.int32), np.zeros(0, np.int32), np.zeros((0, k, k), np.uint64)))
    dup = bsp_gen.random_matrix(64, 64, k, 0.5, 5, "small")
    mats.append((64, 64, k, np.concatenate([dup[3], dup[3][:5]]), np.concatenate([dup[4], dup[4][:5]]),
                 np.concatenate([dup[5].reshape(-1, k, k), dup[5].reshape(-1, k, k)[:5] + np.uint64(1)])))
    with tempfile.TemporaryDirectory() as dd:
        paths = []
        for i, m in enumerate(mats):
            p = os.path.join(dd, "matrix%d" % (i + 1))
            with open(p, "w") as f:
                f.write(util.matrix_text(m[0], m[1], k, m[3], m[4], m[5]))   # FILE order, duplicates kept
            paths.append(p)
        many = ctx.read_text_many(paths, k)
        assert len(many) == len(mats)
        for p, got, m in zip(paths, many, mats):
            assert coo_equal(got, ctx.read_text(p, k))
            ref = orc.OMat.read(p, k)
            assert np.array_equal(got.to_coo()[2], ref.coo()[2]) and got.nnzb == ref.nnzb
        # two big files: the threads left over split each file
        big = [bsp_gen.random_matrix(300 * 8, 300 * 8, 8, 0.06, 70 + i, "u64") for i in range(2)]
        bp = []
        for i, m in enumerate(big):
            p = os.path.join(dd, "big%d" % i)
            ctx.matrix_from_coo(*m).write_text(p)
            assert os.path.getsize(p) > (1 << 20)
            bp.append(p)
        for p, got in zip(bp, ctx.read_text_many(bp, 8)):
            assert coo_equal(got, ctx.read_text(p, 8))
        assert ctx.read_text_many([], k) == []
        # one bad file among many: an error, every handle released, the context still usable
        with open(paths[17], "r+b") as f:
            f.truncate(os.path.getsize(paths[17]) // 2)
        with pytest.raises(bspgemm.BspError) as e:
            ctx.read_text_many(paths, k)
        assert e.value.code == 5 and "matrix18" in str(e.value)
        with pytest.raises(bspgemm.BspError):
            ctx.read_text_many(paths[:3] + [os.path.join(dd, "absent")], k)
        assert coo_equal(ctx.read_text_many(paths[:2], k)[1], ctx.read_text(paths[1], k))


def test_a4_binary_and_checksum_flags(ctx):
    """`a4 --binary --checksum`: reads matrixI.bin where present (text otherwise), writes
    ./matrix.bin, prints the checksum; the product equals the default text run."""
    import re
    import subprocess
    exe = os.path.join(os.path.dirname(bspgemm.__file__), "a4")
    k = 4
    mats = [bsp_gen.random_matrix(60 * k, 60 * k, k, 0.08, 300 + i, "edgy" if i == 2 else "u64") for i in range(5)]
    want, _ = orc.chain([orc.OMat.from_coo(*m) for m in mats], ranks=2)
    want.prune_zero()
    with tempfile.TemporaryDirectory() as dd:
        folder = os.path.join(dd, "in")
        os.mkdir(folder)
        orc.write_folder(folder, [orc.OMat.from_coo(*m) for m in mats], k)
        for i in (1, 3):  # two of the five inputs also exist as binary files; their text twins are removed
            ctx.matrix_from_coo(*mats[i]).write_bin(os.path.join(folder, "matrix%d.bin" % (i + 1)))
            os.remove(os.path.join(folder, "matrix%d" % (i + 1)))
        p = subprocess.run([exe, folder, "--ranks", "2", "--binary", "--checksum"], cwd=dd, capture_output=True, text=True)
        assert p.returncode == 0, p.stderr
        assert not os.path.exists(os.path.join(dd, "matrix"))
        got = ctx.read_bin(os.path.join(dd, "matrix.bin"))
        br, bc, v = got.to_coo()
        wbr, wbc, wv = want.coo()
        assert np.array_equal(br, wbr) and np.array_equal(bc, wbc) and np.array_equal(v, wv)
        m = re.search(r"^checksum ([0-9a-f]{16})$", p.stdout, re.M)
        assert m and int(m.group(1), 16) == got.checksum() == util.checksum_numpy(want.rows, want.cols, k, wbr, wbc, wv)
        # without --binary the missing text files are an error (exit code 3, as for any unreadable input)
        q = subprocess.run([exe, folder, "--ranks", "2"], cwd=dd, capture_output=True, text=True)
        assert q.returncode == 3 and "matrix2" in q.stderr
        # --no-output leaves no file behind
        os.remove(os.path.join(dd, "matrix.bin"))
        r = subprocess.run([exe, folder, "--binary", "--checksum", "--no-output"], cwd=dd, capture_output=True, text=True)
        assert r.returncode == 0 and not os.path.exists(os.path.join(dd, "matrix.bin"))
        assert re.search(r"^checksum [0-9a-f]{16}$", r.stdout, re.M)
        u = subprocess.run([exe, folder, "--frobnicate"], cwd=dd, capture_output=True, text=True)
        assert u.returncode == 2


def test_full_size_product_checksum_whole_vs_panels(ctx, capsys):
    """BASELINE.json configs[1] at full size (n=65536, k=8, 1 %): the 19.2 GB product computed in
    one piece and as 8 row panels concatenated on the device has the same content checksum -- the
    comparison the reference can only make by diffing 35-50 GB text files.  Also reports the
    checksum kernel's bandwidth (one streaming pass over the tiles)."""
    import time
    n, k, d = 65536, 8, 0.01
    G = n // k
    A = ctx.matrix_from_coo(*bsp_gen.random_matrix(n, n, k, d, 1000, "u64"))
    B = ctx.matrix_from_coo(*bsp_gen.random_matrix(n, n, k, d, 1001, "u64"))
    full = ctx.helper(A, B)
    full.checksum()
    t0 = time.perf_counter()
    whole = full.checksum()
    dt = time.perf_counter() - t0
    nbytes = full.nnzb * (k * k * 8 + 8)
    with capsys.disabled():
        print("\n[checksum] %.2f GB in %.2f ms = %.0f GB/s" % (nbytes / 1e9, dt * 1e3, nbytes / dt / 1e9))
    full.free()
    bounds = [G * i // 8 for i in range(9)]
    parts = [ctx.helper(A, B, panel=(bounds[i], bounds[i + 1])) for i in range(8)]
    cat = ctx.concat_panels(parts)
    for p in parts:
        p.free()
    assert cat.checksum() == whole
    # one flipped bit near the end of the 19.2 GB changes it (the kernel indexes with 64 bits)
    import torch
    nbr, nbc, _, _, _ = cat.device_view()
    rowptr = torch.empty(nbr + 1, dtype=torch.int32, device="cuda:0")
    colidx = torch.empty(cat.nnzb, dtype=torch.int32, device="cuda:0")
    vals = torch.empty(cat.nnzb * k * k, dtype=torch.int64, device="cuda:0")
    cat.export_device(rowptr.data_ptr(), colidx.data_ptr(), vals.data_ptr())
    ctx.sync()
    nnzb = cat.nnzb
    cat.free()
    vals[-5] ^= 1 << 62
    torch.cuda.synchronize()
    flipped = ctx.matrix_from_device_bcsr(n, n, k, nbr, nbc, nnzb, rowptr.data_ptr(), colidx.data_ptr(), vals.data_ptr())
    assert flipped.checksum() != whole
    vals[-5] ^= 1 << 62
    torch.cuda.synchronize()
    again = ctx.matrix_from_device_bcsr(n, n, k, nbr, nbc, nnzb, rowptr.data_ptr(), colidx.data_ptr(), vals.data_ptr())
    assert again.checksum() == whole
    # hand the 19 GB scratch tensors back before the other test modules run in this process
    flipped.free(); again.free(); A.free(); B.free()
    ctx.sync()
    del rowptr, colidx, vals
    torch.cuda.empty_cache()
