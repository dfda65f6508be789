"""GPU parity of the swept-volume bit-window check and the swept-volume ID shift: the reference's two
Boost tests through the API, then randomised maps against the CPU oracle (which is pinned against the reference's
BitVector.h compiled for the host, tests/test_oracle_vs_ref_cpu.py) -- with and without summaries."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

DIMS = (89, 123, 74)
N = DIMS[0] * DIMS[1] * DIMS[2]
BIT, PROB = 0, 4
SV_START = 4


@pytest.fixture(scope="module")
def gv():
    import gpu_voxels_b200 as g
    g._lib.init()
    return g


def test_ref_collision_bitshift(gv):
    """GV/test/testing_voxelmap.cu:173-229"""
    pts = np.array([[x, y, z] for x in (2, 3) for y in (2, 3) for z in (2, 3)], np.float32)
    type_1, window = SV_START + 40, 20
    map_1, map_2 = gv.BitVectorVoxelMap(DIMS, 1.0), gv.BitVectorVoxelMap(DIMS, 1.0)
    map_1.insertPointCloud(pts, type_1)
    types_voxel = gv.BitVectorVoxel()
    for i in range(type_1 - window - 9, type_1 + window + 10):
        map_2.clearMap()
        map_2.insertPointCloud(pts, i)
        n = map_1.collisionCheckBitvector(map_2, window, types_voxel)
        if n > 0:
            types_voxel.bytes[type_1 >> 3] &= ~np.uint8(1 << (type_1 & 7))
            assert not types_voxel.bytes.any()
        assert n == (0 if abs(i - type_1) > window else len(pts)), i


def test_ref_bitvoxelmap_bitshift(gv):
    """GV/test/testing_voxelmap.cu:238-304"""
    pts = np.array([[0.3] * 3, [1.3] * 3, [2.3] * 3, [3.3] * 3], np.float32)
    shift_start = 50
    type_int = SV_START + shift_start
    map_1, map_2 = gv.BitVectorVoxelMap(DIMS, 1.0), gv.BitVectorVoxelMap(DIMS, 1.0)
    for shift in range(0, shift_start + SV_START - 5):
        map_1.clearMap(); map_2.clearMap()
        map_2.insertPointCloud(pts, type_int)
        map_1.insertPointCloud(pts, type_int - shift)
        map_2.shiftLeftSweptVolumeIDs(shift)
        n = map_1.collisionCheckBitvector(map_2, 1, gv.BitVectorVoxel())
        assert n == (len(pts) if shift <= shift_start else 0), shift
    with pytest.raises(gv.GvxError):
        map_2.shiftLeftSweptVolumeIDs(57)


def _random_bit_map(seed, n, density, max_bits=4, low_bits=True):
    rng = np.random.Generator(np.random.PCG64(seed))
    m = np.zeros((n, 32), np.uint8)
    idx = np.flatnonzero(rng.uniform(size=n) < density)
    for _ in range(max_bits):
        sel = idx[rng.uniform(size=len(idx)) < 0.6]
        b = rng.integers(0 if low_bits else 4, 256, size=len(sel))
        m[sel, b >> 3] |= (1 << (b & 7)).astype(np.uint8)
    return m.reshape(-1)


@pytest.mark.parametrize("use_summary", [1, 0])
@pytest.mark.parametrize("margin,sv_offset", [(0, 0), (1, 0), (5, 3), (3, 8), (20, 0), (31, 17), (2, 100), (0, 7), (10, 255)])
def test_bitcheck_random_maps_match_oracle(gv, oracle, margin, sv_offset, use_summary):
    dims = (64, 48, 33)
    n = dims[0] * dims[1] * dims[2]
    a_h, b_h = _random_bit_map(81, n, 0.3), _random_bit_map(82, n, 0.3)
    a, b = gv.BitVectorVoxelMap(dims, 0.01), gv.BitVectorVoxelMap(dims, 0.01)
    a.upload(a_h.reshape(n, 32)); b.upload(b_h.reshape(n, 32))
    gv.set_option(gv.GVX_OPT_USE_SUMMARY, use_summary)
    try:
        for mark in (False, True):
            r_a = a_h.copy()
            want, want_types = oracle.collide_bitcheck(r_a, b_h, n, margin, sv_offset, mark)
            tv = gv.BitVectorVoxel()
            got = a.collisionCheckBitvector(b, margin, tv, sv_offset, mark=mark)
            assert got == want and (want > 0 or sv_offset >= 64)  # a large offset pushes every result bit out of its byte
            assert np.array_equal(tv.bytes, want_types)
            assert np.array_equal(a.download().reshape(-1), r_a), "eBVM_COLLISION marks differ"
        # the summary of `a` must have followed the marks: a plain collideWith afterwards agrees with the oracle
        assert a.collideWith(b, 1.0, mark=False) == oracle.collide_map_map(r_a.copy(), BIT, b_h, BIT, n, 1.0, 0, False)
    finally:
        gv.set_option(gv.GVX_OPT_USE_SUMMARY, 1)


def test_free_only_voxels_reach_swept_volume_bits_through_the_window(gv, oracle):
    """`this` voxels holding only eBVM_FREE are not in the occupancy summary, yet with margin + offset/8 >= 4 their
    bit 0 lies within the window of swept-volume bit 4..: the a-side summary must not be used as a filter then."""
    dims = (32, 32, 8)
    n = dims[0] * dims[1] * dims[2]
    a_h = np.zeros((n, 32), np.uint8); b_h = np.zeros((n, 32), np.uint8)
    a_h[::3, 0] = 1          # eBVM_FREE only
    b_h[::2, 0] = 0x30       # swept-volume bits 4 and 5
    a, b = gv.BitVectorVoxelMap(dims, 0.01), gv.BitVectorVoxelMap(dims, 0.01)
    a.upload(a_h); b.upload(b_h)
    for margin in (3, 4, 6):
        want, wt = oracle.collide_bitcheck(a_h.reshape(-1).copy(), b_h.reshape(-1), n, margin, 0, False)
        tv = gv.BitVectorVoxel()
        assert a.collisionCheckBitvector(b, margin, tv, 0, mark=False) == want
        assert np.array_equal(tv.bytes, wt)
        assert (want > 0) == (margin >= 4)


@pytest.mark.parametrize("shift", [0, 1, 7, 8, 31, 32, 33, 56])
def test_shift_matches_oracle_and_keeps_the_summary_exact(gv, oracle, shift):
    dims = (40, 40, 21)
    n = dims[0] * dims[1] * dims[2]
    h = _random_bit_map(90 + shift, n, 0.4, max_bits=6)
    m = gv.BitVectorVoxelMap(dims, 0.01)
    m.upload(h.reshape(n, 32))
    m.shiftLeftSweptVolumeIDs(shift)
    want = h.copy()
    assert oracle.shift_map(want, n, shift)
    got = m.download().reshape(n, 32)
    assert np.array_equal(got.reshape(-1), want)
    occ, valid = m.summary()
    assert valid
    w = want.reshape(n, 32).copy(); w[:, 0] &= 0xFE
    assert np.array_equal(occ, w.any(axis=1))


def test_collide_with_types_bit_x_bit_is_window_zero(gv, oracle):
    """BitVoxelMap::collideWithTypes(BitVectorVoxelMap*), GV/voxelmap/BitVoxelMap.hpp:205-210: SVCollider(thr), window 0"""
    dims = (50, 30, 20)
    n = dims[0] * dims[1] * dims[2]
    a_h, b_h = _random_bit_map(95, n, 0.35), _random_bit_map(96, n, 0.35)
    a, b = gv.BitVectorVoxelMap(dims, 0.01), gv.BitVectorVoxelMap(dims, 0.01)
    a.upload(a_h.reshape(n, 32)); b.upload(b_h.reshape(n, 32))
    r_a = a_h.copy()
    want, wt = oracle.collide_bitcheck(r_a, b_h, n, 0, 0, True)
    tv = gv.BitVectorVoxel()
    assert a.collideWithTypes(b, tv) == want
    assert np.array_equal(tv.bytes, wt)
    # window 0: a bit collides iff the same swept-volume bit is set on both sides
    both = (a_h.reshape(n, 32) & b_h.reshape(n, 32)); both[:, 0] &= 0xF0
    assert want == int(both.any(axis=1).sum())


def test_bitcheck_argument_errors(gv):
    a, b = gv.BitVectorVoxelMap((8, 8, 8), 0.01), gv.BitVectorVoxelMap((8, 8, 8), 0.01)
    p = gv.ProbVoxelMap((8, 8, 8), 0.01)
    tv = gv.BitVectorVoxel()
    for args in ((b, 32, tv, 0), (b, 0, tv, 8 * 32), (b, 0, tv, 70000), (p, 0, tv, 0)):
        with pytest.raises(gv.GvxError):
            a.collisionCheckBitvector(*args)
