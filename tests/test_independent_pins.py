"""A second, independently written derivation of the arithmetic on the path, used to pin the C oracle (the reference
has no test vectors and cannot run here: SURVEY.md §8c).  Everything below is written from the Pascal / asm TEXT in
matrix or vectorised form - a different code path from oracle/*.c, which follows the reference's loops:
  * RGBToYUV (main.pas:2656-2679) as array arithmetic;
  * WaveletGS (main.pas:2805-2840, depth 2) as three orthonormal block products W_n B W_n^T on the n x n top-left
    block, n = 8, 4, 2 (the pyramid decomposition the recursion performs);
  * the DCT branch (main.pas:3075-3174, LUT main.pas:615-623) as (Cv X Cu^T) o cUVRatio;
  * the x86-64 dissimilarity (kmodes.pas:357-412) emulated register by register (psubb / pabsb / psadbw / paddw /
    pcmpeqb / pmovmskb / popcnt / pextrw) on numpy integer arrays, and the generic form (kmodes.pas:239-250);
  * exact k-NN as a float64 distance matrix (torch.cdist).
100 000 random tiles / row pairs each."""
import numpy as np
import pytest

N = 100_000


def yuv_planes(rgb, gamma=-1):
    """rgb int32 [n][64] $00BBGGRR -> float64 [n][3][8][8]"""
    r = (rgb & 0xff).astype(np.float64)
    g = ((rgb >> 8) & 0xff).astype(np.float64)
    b = ((rgb >> 16) & 0xff).astype(np.float64)
    if gamma >= 0:
        e = (2.0, 0.6)[gamma]                      # gGamma main.pas:586
        fr, fg, fb = (r / 255.0) ** e, (g / 255.0) ** e, (b / 255.0) ** e
    else:
        fr, fg, fb = r / 255.0, g / 255.0, b / 255.0
    y = (2126 * fr + 7152 * fg + 722 * fb) / 10000.0
    u = (fb - y) * (0.5 / (1.0 - 722 / 10000.0))
    v = (fr - y) * (0.5 / (1.0 - 2126 / 10000.0))
    return np.stack([y, u, v], 1).reshape(-1, 3, 8, 8)


def haar_level(n):
    """one WaveletGS level on n samples: pair sums to the first half, pair differences to the second, both / sqrt 2"""
    w = np.zeros((n, n))
    for i in range(n // 2):
        w[i, 2 * i] = w[i, 2 * i + 1] = 1 / np.sqrt(2.0)
        w[n // 2 + i, 2 * i], w[n // 2 + i, 2 * i + 1] = 1 / np.sqrt(2.0), -1 / np.sqrt(2.0)
    return w


def haar_pyramid(p):
    """WaveletGS(…, 8, 8, depth 2): rows then columns of the n x n top-left block, for n = 8, 4, 2 - the recursion
    re-enters on the low-low quadrant only (main.pas:2838-2839), i.e. the non-standard (pyramid) decomposition, NOT
    the separable H X H^T: the low-pass halves of high-pass rows are left alone."""
    f = p.copy()
    for n in (8, 4, 2):
        w = haar_level(n)
        f[..., :n, :n] = w @ f[..., :n, :n] @ w.T
    return f


def dct_matrices():
    x = np.arange(8)
    c = np.cos((x[None, :] + 0.5) * x[:, None] * np.pi / 16.0)      # c[u][x], main.pas:615-623
    k = np.where(x == 0, np.sqrt(0.5), 1.0)
    ratio = np.outer(k, k)                                            # cUVRatio main.pas:3000-3009 (0.5 at [0][0])
    return c, ratio


def descriptors(rgb, wavelets=True, gamma=-1):
    p = yuv_planes(rgb, gamma)
    if wavelets:
        f = haar_pyramid(p)                                           # rows of the result index the vertical frequency
    else:
        c, ratio = dct_matrices()
        f = (c @ p @ c.T) * ratio
    return f.reshape(len(rgb), 192)


def test_haar_levels_are_orthonormal_so_the_pyramid_is_an_isometry():
    for n in (2, 4, 8):
        w = haar_level(n)
        assert np.allclose(w @ w.T, np.eye(n), atol=1e-15)
    x = np.random.default_rng(0).standard_normal((50, 3, 8, 8))
    assert np.allclose((haar_pyramid(x) ** 2).sum((1, 2, 3)), (x ** 2).sum((1, 2, 3)), rtol=1e-13)


@pytest.mark.parametrize("wavelets,gamma", [(True, -1), (False, -1), (True, 0), (False, 1)])
def test_descriptors_matrix_form_vs_oracle(oracle, wavelets, gamma):
    rng = np.random.default_rng(7 + gamma + 2 * wavelets)
    rgb = rng.integers(0, 1 << 24, (N, 64)).astype(np.int32)
    rgb[0] = 0
    rgb[1] = 0xFFFFFF
    want = descriptors(rgb, wavelets, gamma)
    got32 = oracle.psyv_queries_f32(rgb, use_wavelets=wavelets, gamma=gamma)
    # float32 rows (what crosses the FFI, main.pas:4024-4025): the two derivations differ by double rounding-order
    # noise only, i.e. they round to the same float32 except when the double sits on a rounding boundary
    w32 = want.astype(np.float32)
    ulp = np.spacing(np.maximum(np.abs(w32), np.float32(1e-30)))
    assert np.all(np.abs(got32.astype(np.float64) - want) <= 0.5 * ulp + 1e-12 + 2e-15 * np.abs(want).max())
    assert (got32 == w32).mean() > 0.999
    # a few tiles in full double precision against the oracle's 192 doubles
    for i in range(16):
        d = oracle.psyv(rgb=rgb[i], use_wavelets=wavelets, gamma=gamma)
        assert np.allclose(d, want[i], rtol=0, atol=2e-13)


def test_palette_rendered_rows_and_mirrors_matrix_form(oracle):
    """FromPal = True with H/V mirror by index reversal (main.pas:3041-3046) == flipping the pixel array."""
    rng = np.random.default_rng(3)
    pal = rng.integers(0, 1 << 24, 16).astype(np.int32)
    tiles = rng.integers(0, 16, (2000, 64)).astype(np.uint8)
    rows = oracle.psyv_tileset_f32(tiles, pal)                        # [n*4][192], order (vmir, hmir) = 00, 01, 10, 11
    px = pal[tiles].reshape(-1, 8, 8)
    for a, (vm, hm) in enumerate([(0, 0), (0, 1), (1, 0), (1, 1)]):
        q = px[:, ::-1, :] if vm else px
        q = q[:, :, ::-1] if hm else q
        want = descriptors(np.ascontiguousarray(q).reshape(-1, 64)).astype(np.float32)
        # 16 colours only: many coefficients cancel to exactly 0 in one summation order and to ~1e-17 in the other
        assert (rows[a::4] == want).mean() > 0.99
        assert np.allclose(rows[a::4], want, rtol=0, atol=1e-6)


def _asm_dissim(a, b):
    """kmodes.pas:357-412 on numpy arrays [n][80] uint8: xmm6..10 = item, xmm0..4 = row"""
    a16, b16 = a.astype(np.int16), b.astype(np.int16)
    d0 = ((a16[:, :16] - b16[:, :16]) & 0xff).astype(np.uint8).view(np.int8)          # psubb (wraps)
    ab = np.abs(d0.astype(np.int16)).astype(np.uint8)                                   # pabsb: -128 -> 0x80
    words = ab.view(np.uint16).astype(np.uint32)                                        # xmm11 as eight words (little endian)
    w0, w4 = words[:, 0].copy(), words[:, 4].copy()
    for blk in range(1, 5):                                                              # psadbw on the other four registers
        s = np.abs(a16[:, 16 * blk:16 * blk + 16] - b16[:, 16 * blk:16 * blk + 16])
        w0 = (w0 + s[:, :8].sum(1)) & 0xffff                                            # paddw: 16-bit lanes
        w4 = (w4 + s[:, 8:].sum(1)) & 0xffff
    mism = (a != b).sum(1).astype(np.uint64)                                            # pcmpeqb / pmovmskb / not / popcnt
    return (mism << np.uint64(11)) + w0.astype(np.uint64) + w4.astype(np.uint64)       # pextrw 0 and 4


def _generic_dissim(a, b):
    d = np.abs(a.astype(np.int64) - b.astype(np.int64))
    return ((a != b).sum(1).astype(np.uint64) << np.uint64(11)) + d.sum(1).astype(np.uint64)


@pytest.mark.parametrize("wide", [False, True])
def test_dissimilarity_register_emulation_vs_oracle(oracle, wide):
    rng = np.random.default_rng(11 + wide)
    hi = 256 if wide else 16
    a = rng.integers(0, hi, (N, 80)).astype(np.uint8)
    b = np.where(rng.random((N, 80)) < 0.7, a, rng.integers(0, hi, (N, 80))).astype(np.uint8)
    wa, wg = _asm_dissim(a, b), _generic_dissim(a, b)
    # the oracle, pair by pair through its arg-min over ONE row (clust 0, dis = the dissimilarity)
    for sem, want in ((0, wa), (1, wg)):
        for i in range(0, N, 20000):  # spot rows through the scalar entry point
            assert oracle.dissim(a[i], b[i], sem) == int(want[i])
    # all pairs: assignment of point a[i] against the single centroid b[i] is too slow pair by pair in Python; use the
    # batch entry point on blocks where every point sees 64 centroids and check the reported minimum and arg-min
    for sem, fn in ((0, _asm_dissim), (1, _generic_dissim)):
        for blk in range(0, 2048, 64):
            pts, cen = a[blk:blk + 64], b[blk:blk + 64]
            full = np.stack([fn(pts, np.repeat(cen[j:j + 1], 64, 0)) for j in range(64)], 1)   # [point][centroid]
            clust, dis = oracle.kmodes_assign(pts, cen, sem)
            assert np.array_equal(dis, full.min(1))
            last = 63 - np.argmin(full[:, ::-1], 1)                                            # last index among equals
            assert np.array_equal(clust, last)


def test_knn_float64_distance_matrix_vs_oracle(oracle):
    import torch
    rng = np.random.default_rng(5)
    ds = rng.standard_normal((5000, 192)).astype(np.float32)
    q = (ds[rng.integers(0, 5000, 2000)] + 0.3 * rng.standard_normal((2000, 192))).astype(np.float32)
    d = torch.cdist(torch.from_numpy(q).double(), torch.from_numpy(ds).double()) ** 2
    wd, wi = torch.topk(d, 9, dim=1, largest=False)
    oi, od, nxt = oracle.knn_f64(ds, q, 8)
    assert np.allclose(od, wd[:, :8].numpy(), rtol=1e-9, atol=1e-9)
    gap = np.diff(wd.numpy(), axis=1).min(1) > 1e-9            # no near-ties among the first nine
    assert gap.mean() > 0.99
    assert np.array_equal(oi[gap], wi[:, :8].numpy()[gap])
    assert np.allclose(nxt, wd[:, 8].numpy(), rtol=1e-9, atol=1e-9)
