"""float32 tiles (BASELINE.json configs[4]).  No reference path exists for this variant (the
reference parses and multiplies uint64 only), so the checker is a float64 evaluation of the same
block products (numpy), and the tolerances are stated here:
  SIMT fp32 kernels : |err| <= 2e-6 * sum_j |a_j||b_j| per element (fp32 FMA rounding, K*pairs terms)
  tensor cores (TF32 inputs, fp32 accumulate): |err| <= 2e-3 * sum_j |a_j||b_j| per element
"""
import numpy as np
import pytest

import bsp_gen
import bspgemm

pytestmark = pytest.mark.gpu


def float_matrix(G_r, G_c, k, dens, seed):
    br, bc = bsp_gen.bernoulli_pattern(G_r, G_c, dens, seed)
    u = bsp_gen.splitmix64(seed ^ 0xABCDEF, br.size * k * k).astype(np.float64) / 2.0 ** 64
    v = (2.0 * u - 1.0).astype(np.float32).reshape(-1, k, k)   # U(-1, 1)
    return G_r * k, G_c * k, k, br, bc, v


def dense_of(m, G_r, G_c, absval=False):
    rows, cols, k, br, bc, v = m
    D = np.zeros((G_r * k, G_c * k), np.float64)
    for r, c, t in zip(br, bc, v):
        D[r * k:(r + 1) * k, c * k:(c + 1) * k] = np.abs(t) if absval else t
    return D


def check(ctx, k, G, dens, path, tol, seeds=(1, 2)):
    a = float_matrix(G, G + 2, k, dens, seeds[0])
    b = float_matrix(G + 2, G - 1, k, dens, seeds[1])
    ctx.set_f32_path(path)
    st = bspgemm.Stats()
    C = ctx.helper(ctx.matrix_from_coo_f32(*a), ctx.matrix_from_coo_f32(*b), stats=st)
    br, bc, v = C.to_coo()
    assert v.dtype == np.float32
    ref = dense_of(a, G, G + 2) @ dense_of(b, G + 2, G - 1)
    bound = dense_of(a, G, G + 2, True) @ dense_of(b, G + 2, G - 1, True)
    # pattern: every tile of the block pattern product is present (zeros are kept), nothing else
    pa = np.zeros((G, G + 2), bool); pa[a[3], a[4]] = True
    pb = np.zeros((G + 2, G - 1), bool); pb[b[3], b[4]] = True
    pc = (pa.astype(np.int32) @ pb.astype(np.int32)) > 0
    got = np.zeros_like(pc); got[br, bc] = True
    assert np.array_equal(got, pc)
    worst = 0.0
    for r, c, t in zip(br, bc, v):
        blk = (slice(r * k, (r + 1) * k), slice(c * k, (c + 1) * k))
        err = np.abs(t.astype(np.float64) - ref[blk])
        worst = max(worst, float((err / np.maximum(bound[blk], 1e-30)).max()))
    assert worst <= tol, (worst, tol, st.numeric_path)
    return st


@pytest.fixture(scope="module")
def ctx():
    c = bspgemm.Context(0)
    yield c
    c.close()


@pytest.mark.parametrize("k", [4, 16, 32, 64])
def test_f32_simt_matches_float64(ctx, k):
    st = check(ctx, k, 10 if k >= 32 else 20, 0.3, "simt", 2e-6)
    assert st.numeric_path == 100


def test_f32_chain_and_prune(ctx):
    k, G = 32, 8
    ctx.set_f32_path("simt")
    ms = [float_matrix(G, G, k, 0.4, 50 + i) for i in range(4)]
    res = ctx.helper2([ctx.matrix_from_coo_f32(*m) for m in ms]).prune_zero()
    D = [dense_of(m, G, G) for m in ms]
    ref = (D[0] @ D[1]) @ (D[2] @ D[3])
    br, bc, v = res.to_coo()
    for r, c, t in zip(br, bc, v):
        assert np.allclose(t, ref[r * k:(r + 1) * k, c * k:(c + 1) * k], rtol=1e-3, atol=1e-3)


def test_mixed_types_rejected(ctx):
    a = float_matrix(4, 4, 4, 0.5, 3)
    A = ctx.matrix_from_coo_f32(*a)
    U = ctx.matrix_from_coo(16, 16, 4, a[3], a[4], np.ones(a[3].size * 16, np.uint64))
    with pytest.raises(bspgemm.BspError):
        ctx.helper(A, U)


@pytest.mark.parametrize("k", [32, 64])
def test_f32_tensor_cores_match_float64_within_tf32(ctx, k):
    st = check(ctx, k, 9, 0.35, "tensor", 2e-3, seeds=(7, 8))
    assert st.numeric_path == 100 + k
    # auto picks the tensor-core kernel where it exists (k = 32 and 64); smaller tiles stay on the fp32 kernels
    st = check(ctx, k, 6, 0.5, "auto", 2e-3, seeds=(9, 10))
    assert st.numeric_path == 100 + k
    assert check(ctx, 16, 6, 0.5, "auto", 2e-6, seeds=(9, 10)).numeric_path == 100


@pytest.mark.parametrize("k", [32, 64])
@pytest.mark.parametrize("tch", [1, 3, 8])
def test_f32_tensor_core_tile_chunks(ctx, k, tch, monkeypatch):
    """The kernel walks output tiles in chunks of BSP_F32_TCH consecutive tiles (a comparison hook; the
    library picks 1..8 itself): many pairs per tile, then mostly single-pair tiles with a ragged last chunk."""
    monkeypatch.setenv("BSP_F32_TCH", str(tch))
    assert check(ctx, k, 7, 0.6, "tensor", 2e-3, seeds=(11, 12)).numeric_path == 100 + k
    assert check(ctx, k, 12, 0.08, "tensor", 2e-3, seeds=(13, 14)).numeric_path == 100 + k


@pytest.mark.parametrize("k", [32, 64])
@pytest.mark.parametrize("tch", [0, 1])
def test_f32_tensor_core_many_tiles(ctx, k, tch, monkeypatch):
    """More chunks than resident CTAs: every CTA moves through several chunks (0 = the library's own chunk size)."""
    if tch:
        monkeypatch.setenv("BSP_F32_TCH", str(tch))
    else:
        monkeypatch.delenv("BSP_F32_TCH", raising=False)
    check(ctx, k, 40, 0.3, "tensor", 2e-3, seeds=(15, 16))
