"""The coarse mask of the summary collide (GVX_OPT_USE_COARSE): 1 bit per 4096 voxels, built lazily on the
second collide of an unmodified map, dropped by every modifying call. Bit-exact against the oracle after EVERY kind of
modification, with the mask's validity and contents checked (bit c == any summary bit of voxels [4096 c, 4096 c + 4096))."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

BIT, PROB = 0, 4
POSE = np.array([[1, 0, 0, 0], [0, 1, 0, 0], [0, 0, 1, 0.001]], np.float32)


@pytest.fixture(scope="module")
def gv():
    import gpu_voxels_b200 as g
    g._lib.init()
    return g


def _cloud(seed, n, lo, hi):
    rng = np.random.Generator(np.random.PCG64(seed))
    return (rng.uniform(0, 1, size=(n, 3)) * (np.asarray(hi) - np.asarray(lo)) + np.asarray(lo)).astype(np.float32)


def _check_coarse(m, expect_valid):
    bits, valid = m.coarse()
    assert valid == expect_valid
    if valid:
        summ, sv = m.summary()
        assert sv
        pad = (-len(summ)) % 4096
        want = np.concatenate([summ, np.zeros(pad, bool)]).reshape(-1, 4096).any(1)
        assert np.array_equal(bits, want)
        assert not want.all() and want.any()   # the test maps are sparse: the mask really skips something


@pytest.mark.parametrize("dims", [(67, 45, 39), (160, 128, 96)])
@pytest.mark.parametrize("pair", [(PROB, BIT), (BIT, PROB), (BIT, BIT), (PROB, PROB)])
def test_coarse_mask_follows_every_modification(gv, oracle, pair, dims):
    side = 0.01
    n = dims[0] * dims[1] * dims[2]
    ext = np.asarray(dims) * side
    mk = lambda t: (gv.ProbVoxelMap if t == PROB else gv.BitVectorVoxelMap)(dims, side)  # noqa: E731
    a, b = mk(pair[0]), mk(pair[1])
    ra, rb = oracle.new_map(pair[0], dims), oracle.new_map(pair[1], dims)

    def insert(m, r, t, pts, meaning=1):
        m.insertPointCloud(pts, meaning)
        oracle.insert_points(r, t, dims, side, pts, meaning)

    def collide3(mark=False, thr=1.0):
        for k in range(3):
            want = oracle.collide_map_map(ra, pair[0], rb, pair[1], n, thr, 0, mark)
            assert a.collideWith(b, thr, mark=mark) == want, (k, mark)
        assert np.array_equal(a.download().reshape(-1).view(np.uint8), ra.view(np.uint8))
        return want

    # sparse, partly overlapping content in two corners
    pa = _cloud(1, 6000, ext * 0.05, ext * 0.30)
    pb = np.concatenate([pa[:2000], _cloud(2, 5000, ext * 0.20, ext * 0.45)])
    insert(a, ra, pair[0], pa)
    insert(b, rb, pair[1], pb)
    _check_coarse(a, False)
    assert collide3() > 0
    _check_coarse(a, True); _check_coarse(b, True)
    # marks do not change occupancy / known-ness: the masks stay valid, the marked voxels are still found
    collide3(mark=True)
    _check_coarse(a, True); _check_coarse(b, True)
    # new voxels far away from everything the masks cover
    far = _cloud(3, 3000, ext * 0.70, ext * 0.95)
    insert(b, rb, pair[1], far)
    _check_coarse(b, False)
    insert(a, ra, pair[0], far[:1500])
    w = collide3()
    assert w >= 1000
    _check_coarse(a, True); _check_coarse(b, True)
    if pair[1] == PROB:      # log-odds frames reach new voxels too
        fr = _cloud(4, 4000, ext * 0.50, ext * 0.65)
        b.insertSensorData(fr, POSE)
        b.insertSensorData(fr, POSE)
        oracle.update_logodds(rb, dims, side, fr, POSE)
        oracle.update_logodds(rb, dims, side, fr, POSE)
        insert(a, ra, pair[0], fr[:2000])
        collide3(thr=0.5)
    if pair[0] == BIT:       # removing a meaning empties voxels
        a.insertPointCloud(far[1500:], 9); oracle.insert_points(ra, BIT, dims, side, far[1500:], 9)
        collide3()
        mask = np.zeros(32, np.uint8); mask[0] = 0x06   # eBVM_OCCUPIED and eBVM_COLLISION
        a.clearBits(mask); oracle.clear_bits(ra, n, mask)
        _check_coarse(a, False)
        collide3()
    # upload of foreign content
    up = oracle.new_map(pair[1], dims)
    oracle.insert_points(up, pair[1], dims, side, _cloud(5, 4000, ext * 0.02, ext * 0.25), 1)
    b.upload(up.reshape(-1, 32) if pair[1] == BIT else up)
    rb[:] = up
    _check_coarse(b, False)
    collide3()
    assert a.collisionCheck(b, 1.0) == (oracle.collide_map_map(ra.copy(), pair[0], rb, pair[1], n, 1.0, 0, False) > 0)
    # clear: nothing collides any more, and a re-filled map is seen again
    b.clearMap(); rb[:] = oracle.new_map(pair[1], dims)
    assert collide3() == 0
    insert(b, rb, pair[1], pa[:3000])
    collide3()
    # the option switches the mask off without changing any result
    gv.set_option(gv.GVX_OPT_USE_COARSE, 0)
    try:
        insert(b, rb, pair[1], far[:10])
        collide3()
        _check_coarse(b, False)
    finally:
        gv.set_option(gv.GVX_OPT_USE_COARSE, 1)
    a.destroy(); b.destroy()
