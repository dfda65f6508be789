"""GPU parity of the TMA-fed voxel-streaming collide kernels (gvx_collide_tma.cu, GVX_OPT_COLLIDE_TMA = 1..5) against
the CPU oracle and against the plain-load kernels: counts and the eBVM_COLLISION write-back, bit-exact, for every
map-type pair with a BitVectorVoxel side, grids with and without a partial last tile, every tile/stage shape."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

BIT, PROB = 0, 4


@pytest.fixture(scope="module")
def gv():
    import gpu_voxels_b200 as g
    g._lib.init()
    yield g
    g.set_option(g.GVX_OPT_COLLIDE_TMA, -1)
    g.set_option(g.GVX_OPT_USE_SUMMARY, 1)


def _random_map(gv, t, dims, seed, density):
    rng = np.random.Generator(np.random.PCG64(seed))
    n = dims[0] * dims[1] * dims[2]
    if t == PROB:
        data = np.full(n, -128, np.int8)
        k = rng.uniform(size=n) < density
        data[k] = rng.integers(-127, 128, size=int(k.sum()))
        m = gv.ProbVoxelMap(dims, 1.0)
    else:
        data = np.zeros((n, 32), np.uint8)
        idx = np.nonzero(rng.uniform(size=n) < density)[0]
        bits = rng.integers(0, 256, size=len(idx))
        data[idx, bits >> 3] |= (1 << (bits & 7)).astype(np.uint8)
        data[rng.uniform(size=n) < density, 0] |= 1   # eBVM_FREE alone is not occupancy
        data[rng.uniform(size=n) < density / 4, 0] |= 4  # already marked
        m = gv.BitVectorVoxelMap(dims, 1.0)
    m.upload(data.reshape(-1).view(np.uint8))
    return m, data.reshape(-1).view(np.uint8).copy()


@pytest.mark.parametrize("dims", [(89, 123, 74), (64, 64, 32), (7, 5, 3), (40, 16, 1)])
@pytest.mark.parametrize("pair", [(BIT, PROB), (PROB, BIT), (BIT, BIT)])
@pytest.mark.parametrize("shape", [1, 2, 3, 4, 5])
def test_tma_collide_equals_oracle_and_plain_kernel(gv, oracle, dims, pair, shape):
    ta, tb = pair
    n = dims[0] * dims[1] * dims[2]
    gv.set_option(gv.GVX_OPT_USE_SUMMARY, 0)
    for thr, mark in ((0.5, True), (1.0, False), (0.1, True)):
        a, a0 = _random_map(gv, ta, dims, 11 + shape, 0.3)
        b, b0 = _random_map(gv, tb, dims, 23, 0.4)
        want_a = a0.copy()
        want = oracle.collide_map_map(want_a, ta, b0, tb, n, thr, 0, mark)
        gv.set_option(gv.GVX_OPT_COLLIDE_TMA, shape)
        got = a.collideWith(b, thr, mark=mark)
        gv.set_option(gv.GVX_OPT_COLLIDE_TMA, 0)
        assert got == want
        assert np.array_equal(a.download().reshape(-1).view(np.uint8), want_a)
        assert np.array_equal(b.download().reshape(-1).view(np.uint8), b0)
        # the plain-load kernel on the (now marked) map gives the same count again
        assert a.collideWith(b, thr, mark=False) == want
    gv.set_option(gv.GVX_OPT_USE_SUMMARY, 1)


def test_tma_option_leaves_unsupported_calls_on_the_plain_kernels(gv, oracle):
    """offsets, collideWithTypes and Prob x Prob are not TMA-fed: results must not change with the option on"""
    dims = (32, 24, 16)
    n = dims[0] * dims[1] * dims[2]
    gv.set_option(gv.GVX_OPT_USE_SUMMARY, 0)
    gv.set_option(gv.GVX_OPT_COLLIDE_TMA, 1)
    a, a0 = _random_map(gv, BIT, dims, 5, 0.3)
    b, b0 = _random_map(gv, PROB, dims, 6, 0.4)
    off = (1, -2, 1)
    lin = oracle.voxel_index_signed(dims, off)
    wa = a0.copy()
    assert a.collideWith(b, 0.5, off) == oracle.collide_map_map(wa, BIT, b0, PROB, n, 0.5, lin, True)
    types = gv.BitVectorVoxel()
    wa2 = a.download().reshape(-1).view(np.uint8).copy()
    mask = np.zeros(32, np.uint8)
    assert a.collideWithTypes(b, types, 0.5) == oracle.collide_bit_prob_types(wa2, b0, n, 0.5, True, mask)[0]
    p1, q0 = _random_map(gv, PROB, dims, 7, 0.5)
    w1 = q0.copy()
    assert p1.collideWith(b, 0.5) == oracle.collide_map_map(w1, PROB, b0, PROB, n, 0.5, 0, True)
    assert gv._lib.load().gvx_set_option(gv.GVX_OPT_COLLIDE_TMA, 9) != 0
    gv.set_option(gv.GVX_OPT_COLLIDE_TMA, -1)
    gv.set_option(gv.GVX_OPT_USE_SUMMARY, 1)


def test_auto_mode_feeds_large_bit_x_prob_through_tma_with_the_same_result(gv, oracle):
    """default (-1): Bit x Prob on >= 4 M voxels takes the TMA kernel; count and marks as the oracle's"""
    dims = (256, 192, 96)  # 4.7 M voxels, 151 MB of BitVectorVoxels
    n = dims[0] * dims[1] * dims[2]
    gv.set_option(gv.GVX_OPT_USE_SUMMARY, 0)
    gv.set_option(gv.GVX_OPT_COLLIDE_TMA, -1)
    a, a0 = _random_map(gv, BIT, dims, 41, 0.2)
    b, b0 = _random_map(gv, PROB, dims, 42, 0.3)
    want_a = a0.copy()
    want = oracle.collide_map_map(want_a, BIT, b0, PROB, n, 0.5, 0, True)
    assert a.collideWith(b, 0.5) == want
    assert np.array_equal(a.download().reshape(-1).view(np.uint8), want_a)
    gv.set_option(gv.GVX_OPT_COLLIDE_TMA, 0)
    assert a.collideWith(b, 0.5, mark=False) == want
    gv.set_option(gv.GVX_OPT_COLLIDE_TMA, -1)
    gv.set_option(gv.GVX_OPT_USE_SUMMARY, 1)
